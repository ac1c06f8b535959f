#!/usr/bin/env python3
"""Limb-level model of Field::sqr (csrc/field.cuh): the 28 cross products in even/odd aligned accumulators, doubling, the 8
squares, and a separated Montgomery reduction in the shifting even/odd frame.  Checked against big-int arithmetic."""
import random

M32 = (1 << 32) - 1
Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617


def limbs(x, n=8):
    return [(x >> (32 * i)) & M32 for i in range(n)]


def val(l):
    return sum(v << (32 * i) for i, v in enumerate(l))


def chain(acc, k, xs, y):
    """acc[k .. k+2n-1] += {xs} * y on aligned pairs, carry added into acc[k+2n] (must not overflow)"""
    carry = 0
    for t, x in enumerate(xs):
        p = x * y
        lo = acc[k + 2 * t] + (p & M32) + carry
        acc[k + 2 * t] = lo & M32
        hi = acc[k + 2 * t + 1] + (p >> 32) + (lo >> 32)
        acc[k + 2 * t + 1] = hi & M32
        carry = hi >> 32
    acc[k + 2 * len(xs)] += carry
    assert acc[k + 2 * len(xs)] <= M32


def sqr(a_int, p_int):
    a = limbs(a_int)
    p = limbs(p_int)
    inv = (-pow(p_int, -1, 1 << 32)) & M32
    E, O = [0] * 16, [0] * 16
    for i in range(7):
        ja = list(range(i + 1, 8, 2))      # odd positions -> O, index i + j - 1
        jb = list(range(i + 2, 8, 2))      # even positions -> E, index i + j
        if ja:
            chain(O, i + ja[0] - 1, [a[j] for j in ja], a[i])
        if jb:
            chain(E, i + jb[0], [a[j] for j in jb], a[i])
    C = [0] * 16
    carry = 0
    for k in range(16):
        s = E[k] + (O[k - 1] if k else 0) + carry
        C[k] = s & M32
        carry = s >> 32
    assert carry == 0 and O[15] == 0
    T = [((C[k] << 1) | ((C[k - 1] >> 31) if k else 0)) & M32 for k in range(16)]
    assert C[15] >> 31 == 0
    carry = 0
    for i in range(8):
        sq = a[i] * a[i]
        lo = T[2 * i] + (sq & M32) + carry
        T[2 * i] = lo & M32
        hi = T[2 * i + 1] + (sq >> 32) + (lo >> 32)
        T[2 * i + 1] = hi & M32
        carry = hi >> 32
    assert carry == 0 and val(T) == a_int * a_int
    # separated reduction in the shifting frame
    Ef = T[:8]
    hi = 0
    for i in range(8):
        m = (Ef[0] * inv) & M32
        Of = [0] * 8
        for t in range(4):
            pr = p[2 * t + 1] * m
            Of[2 * t], Of[2 * t + 1] = pr & M32, pr >> 32
        c = 0
        for t in range(4):      # E += p_even * m
            pr = p[2 * t] * m
            lo = Ef[2 * t] + (pr & M32) + c
            Ef[2 * t] = lo & M32
            h = Ef[2 * t + 1] + (pr >> 32) + (lo >> 32)
            Ef[2 * t + 1] = h & M32
            c = h >> 32
        assert Ef[0] == 0
        t_in = T[8 + i] + c + hi
        new = [0] * 8
        carry = 0
        for j in range(7):
            s = Ef[j + 1] + Of[j] + carry
            new[j] = s & M32
            carry = s >> 32
        s = Of[7] + (t_in & M32) + carry
        new[7] = s & M32
        hi = (s >> 32) + (t_in >> 32)
        Ef = new
    r = val(Ef) + (hi << 256)
    if r >= p_int:
        r -= p_int
    assert r < p_int
    return r


if __name__ == "__main__":
    rng = random.Random(1)
    for p in (Q, R):
        rinv = pow(1 << 256, -1, p)
        for a in [0, 1, p - 1, p - 2, (1 << 254) % p, (1 << 256) % p] + [rng.randrange(p) for _ in range(3000)]:
            assert sqr(a, p) == a * a * rinv % p, hex(a)
    print("sqr model ok")
