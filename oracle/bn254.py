"""BN254 ("bn128") fields, curve groups and the optimal-ate pairing in Python ints.

Oracle code (test infrastructure, see oracle/__init__.py).  Restates row a17 of
SURVEY.md section 8(a): Fr, Fq, Fq2 = Fq[u]/(u^2+1), G1: y^2 = x^3 + 3 over Fq,
G2: y^2 = x^3 + 3/(9+u) over Fq2.  In-tree anchors for the constants:
q as 6x43-bit limbs in tests/1_sigs_1_batches_5_height/logs/layer_two_setup.log:9,
generators in tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json:5-51.

The pairing follows the construction the reference itself calls through
py_ecc (scripts/sanitize_groth16_proof.py:16-18, :63): Fq12 = Fq[w]/(w^12 - 18 w^6 + 82),
twist (x, y) -> (x' w^2, y' w^3), Miller loop over 6t+2, two Frobenius line
steps, final exponentiation (q^12-1)/r.
"""

Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
MONT_BITS = 256
MONT_R_Q = (1 << MONT_BITS) % Q
MONT_R_R = (1 << MONT_BITS) % R

G1_GEN = (1, 2)
G2_GEN = (
    (10857046999023057135944570762232829481370756359578518086990519993285655852781,
     11559732032986387107991004021392285783925812861821192530917403151452391805634),
    (8495653923123431417604973247489272438418190587263600148770280649306958101930,
     4082367875863433681332203403145435568316851327593401208105741076214120093531),
)

ATE_LOOP_COUNT = 29793968203157093288  # 6t+2, t = 4965661367192848881
LOG_ATE_LOOP_COUNT = 63


def inv_mod(a, m):
    return pow(a % m, -1, m)


# ----------------------------------------------------------------------------
# Fq2 as tuples (c0, c1), u^2 = -1
# ----------------------------------------------------------------------------
def f2_add(a, b):
    return ((a[0] + b[0]) % Q, (a[1] + b[1]) % Q)


def f2_sub(a, b):
    return ((a[0] - b[0]) % Q, (a[1] - b[1]) % Q)


def f2_neg(a):
    return ((-a[0]) % Q, (-a[1]) % Q)


def f2_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % Q, (a[0] * b[1] + a[1] * b[0]) % Q)


def f2_sqr(a):
    return f2_mul(a, a)


def f2_inv(a):
    d = inv_mod(a[0] * a[0] + a[1] * a[1], Q)
    return (a[0] * d % Q, (-a[1]) * d % Q)


def f2_muls(a, s):
    return (a[0] * s % Q, a[1] * s % Q)


F2_ZERO = (0, 0)
F2_ONE = (1, 0)
B1 = 3
# twist coefficient b' = 3 / (9 + u)
B2 = f2_mul((3, 0), f2_inv((9, 1)))


class _FqOps:
    zero = 0
    one = 1
    b = B1

    @staticmethod
    def add(a, b):
        return (a + b) % Q

    @staticmethod
    def sub(a, b):
        return (a - b) % Q

    @staticmethod
    def mul(a, b):
        return a * b % Q

    @staticmethod
    def neg(a):
        return (-a) % Q

    @staticmethod
    def inv(a):
        return inv_mod(a, Q)

    @staticmethod
    def is_zero(a):
        return a % Q == 0


class _Fq2Ops:
    zero = F2_ZERO
    one = F2_ONE
    b = B2
    add = staticmethod(f2_add)
    sub = staticmethod(f2_sub)
    mul = staticmethod(f2_mul)
    neg = staticmethod(f2_neg)
    inv = staticmethod(f2_inv)

    @staticmethod
    def is_zero(a):
        return a[0] % Q == 0 and a[1] % Q == 0


class Curve:
    """Short-Weierstrass a=0 group over a field given by `ops`.

    Affine points are (x, y) tuples or None for infinity; Jacobian points are
    (X, Y, Z) with Z == zero for infinity.
    """

    def __init__(self, ops, gen):
        self.F = ops
        self.gen = gen

    def is_on_curve(self, p):
        if p is None:
            return True
        F = self.F
        x, y = p
        return F.sub(F.mul(y, y), F.add(F.mul(F.mul(x, x), x), F.b)) == F.zero

    def to_jac(self, p):
        F = self.F
        if p is None:
            return (F.one, F.one, F.zero)
        return (p[0], p[1], F.one)

    def to_affine(self, P):
        F = self.F
        X, Y, Z = P
        if F.is_zero(Z):
            return None
        zi = F.inv(Z)
        zi2 = F.mul(zi, zi)
        return (F.mul(X, zi2), F.mul(Y, F.mul(zi2, zi)))

    def jdouble(self, P):
        F = self.F
        X, Y, Z = P
        if F.is_zero(Z) or F.is_zero(Y):
            return (F.one, F.one, F.zero)
        A = F.mul(X, X)
        B = F.mul(Y, Y)
        C = F.mul(B, B)
        t = F.add(X, B)
        D = F.sub(F.sub(F.mul(t, t), A), C)
        D = F.add(D, D)
        E = F.add(F.add(A, A), A)
        Fv = F.mul(E, E)
        X3 = F.sub(Fv, F.add(D, D))
        C8 = F.add(C, C)
        C8 = F.add(C8, C8)
        C8 = F.add(C8, C8)
        Y3 = F.sub(F.mul(E, F.sub(D, X3)), C8)
        Z3 = F.mul(Y, Z)
        Z3 = F.add(Z3, Z3)
        return (X3, Y3, Z3)

    def jadd(self, P, Qp):
        F = self.F
        X1, Y1, Z1 = P
        X2, Y2, Z2 = Qp
        if F.is_zero(Z1):
            return Qp
        if F.is_zero(Z2):
            return P
        Z1Z1 = F.mul(Z1, Z1)
        Z2Z2 = F.mul(Z2, Z2)
        U1 = F.mul(X1, Z2Z2)
        U2 = F.mul(X2, Z1Z1)
        S1 = F.mul(Y1, F.mul(Z2, Z2Z2))
        S2 = F.mul(Y2, F.mul(Z1, Z1Z1))
        H = F.sub(U2, U1)
        Rr = F.sub(S2, S1)
        if F.is_zero(H):
            if F.is_zero(Rr):
                return self.jdouble(P)
            return (F.one, F.one, F.zero)
        HH = F.mul(H, H)
        HHH = F.mul(H, HH)
        V = F.mul(U1, HH)
        X3 = F.sub(F.sub(F.mul(Rr, Rr), HHH), F.add(V, V))
        Y3 = F.sub(F.mul(Rr, F.sub(V, X3)), F.mul(S1, HHH))
        Z3 = F.mul(F.mul(Z1, Z2), H)
        return (X3, Y3, Z3)

    def jneg(self, P):
        return (P[0], self.F.neg(P[1]), P[2])

    def jmul(self, P, k):
        F = self.F
        k = int(k)
        if k < 0:
            return self.jmul(self.jneg(P), -k)
        acc = (F.one, F.one, F.zero)
        for bit in bin(k)[2:] if k else "":
            acc = self.jdouble(acc)
            if bit == "1":
                acc = self.jadd(acc, P)
        return acc

    # affine conveniences
    def add(self, p, q):
        return self.to_affine(self.jadd(self.to_jac(p), self.to_jac(q)))

    def neg(self, p):
        if p is None:
            return None
        return (p[0], self.F.neg(p[1]))

    def mul(self, p, k):
        return self.to_affine(self.jmul(self.to_jac(p), k % R))

    def msm(self, points, scalars):
        """Naive sum of scalar multiples (ground truth for tiny sizes)."""
        acc = self.to_jac(None)
        for p, s in zip(points, scalars):
            if p is None or s % R == 0:
                continue
            acc = self.jadd(acc, self.jmul(self.to_jac(p), s % R))
        return self.to_affine(acc)


G1 = Curve(_FqOps, G1_GEN)
G2 = Curve(_Fq2Ops, G2_GEN)


class FixedBase:
    """Windowed fixed-base multiplier k -> k*gen (speeds up the fixture factory)."""

    def __init__(self, curve, base=None, wbits=8):
        self.c = curve
        self.w = wbits
        base = curve.gen if base is None else base
        self.tables = []
        nwin = (254 + wbits - 1) // wbits
        cur = curve.to_jac(base)
        for _ in range(nwin):
            row = [None]
            acc = curve.to_jac(None)
            for _j in range(1, 1 << wbits):
                acc = curve.jadd(acc, cur)
                row.append(acc)
            self.tables.append(row)
            for _j in range(wbits):
                cur = curve.jdouble(cur)

    def mul(self, k):
        k %= R
        c = self.c
        acc = c.to_jac(None)
        i = 0
        mask = (1 << self.w) - 1
        while k:
            d = k & mask
            if d:
                acc = c.jadd(acc, self.tables[i][d])
            k >>= self.w
            i += 1
        return c.to_affine(acc)


# ----------------------------------------------------------------------------
# Fq12 = Fq[w] / (w^12 - 18 w^6 + 82), dense coefficient lists of length 12
# ----------------------------------------------------------------------------
def f12_one():
    return [1] + [0] * 11


def f12_mul(a, b):
    t = [0] * 23
    for i, ai in enumerate(a):
        if ai:
            for j, bj in enumerate(b):
                t[i + j] += ai * bj
    # w^12 = 18 w^6 - 82
    for k in range(22, 11, -1):
        c = t[k]
        if c:
            t[k - 6] += 18 * c
            t[k - 12] -= 82 * c
    return [x % Q for x in t[:12]]


def f12_sub(a, b):
    return [(x - y) % Q for x, y in zip(a, b)]


def f12_add(a, b):
    return [(x + y) % Q for x, y in zip(a, b)]


def f12_muls(a, s):
    return [x * s % Q for x in a]


def f12_pow(a, e):
    out = f12_one()
    base = a
    while e:
        if e & 1:
            out = f12_mul(out, base)
        base = f12_mul(base, base)
        e >>= 1
    return out


def _poly_deg(p):
    d = len(p) - 1
    while d and p[d] == 0:
        d -= 1
    return d


def f12_inv(a):
    """Extended Euclid over Fq[w] modulo the degree-12 modulus."""
    mod = [82, 0, 0, 0, 0, 0, (-18) % Q, 0, 0, 0, 0, 0, 1]
    lm, hm = [1] + [0] * 12, [0] * 13
    low, high = list(a) + [0], mod
    while _poly_deg(low):
        # r = high // low
        dl = _poly_deg(low)
        temp = list(high)
        r = [0] * 13
        inv_lead = inv_mod(low[dl], Q)
        for i in range(_poly_deg(high) - dl, -1, -1):
            c = temp[dl + i] * inv_lead % Q
            r[i] = c
            if c:
                for j in range(dl + 1):
                    temp[i + j] = (temp[i + j] - c * low[j]) % Q
        nm, new = list(hm), list(high)
        for i in range(13):
            if lm[i] == 0 and low[i] == 0:
                continue
            for j in range(13 - i):
                if r[j]:
                    nm[i + j] = (nm[i + j] - lm[i] * r[j]) % Q
                    new[i + j] = (new[i + j] - low[i] * r[j]) % Q
        lm, low, hm, high = nm, new, lm, low
    s = inv_mod(low[0], Q)
    return [x * s % Q for x in lm[:12]]


def _twist(p):
    """G2 affine point over Fq2 -> point over Fq12 (as (x12, y12))."""
    (x0, x1), (y0, y1) = p
    xc = [(x0 - 9 * x1) % Q, x1]
    yc = [(y0 - 9 * y1) % Q, y1]
    nx = [0] * 12
    ny = [0] * 12
    # x * w^2, y * w^3 with u = w^6 - 9
    nx[2], nx[8] = xc[0], xc[1]
    ny[3], ny[9] = yc[0], yc[1]
    return (nx, ny)


def _cast_g1(p):
    return ([p[0]] + [0] * 11, [p[1]] + [0] * 11)


def _f12_is_zero(a):
    return all(x % Q == 0 for x in a)


def _ec12_double(p):
    x, y = p
    m = f12_mul(f12_muls(f12_mul(x, x), 3), f12_inv(f12_muls(y, 2)))
    nx = f12_sub(f12_mul(m, m), f12_muls(x, 2))
    ny = f12_sub(f12_mul(m, f12_sub(x, nx)), y)
    return (nx, ny)


def _ec12_add(p, q):
    if p is None:
        return q
    if q is None:
        return p
    x1, y1 = p
    x2, y2 = q
    if x1 == x2:
        if y1 == y2:
            return _ec12_double(p)
        return None
    m = f12_mul(f12_sub(y2, y1), f12_inv(f12_sub(x2, x1)))
    nx = f12_sub(f12_sub(f12_mul(m, m), x1), x2)
    ny = f12_sub(f12_mul(m, f12_sub(x1, nx)), y1)
    return (nx, ny)


def _linefunc(p1, p2, t):
    x1, y1 = p1
    x2, y2 = p2
    xt, yt = t
    if x1 != x2:
        m = f12_mul(f12_sub(y2, y1), f12_inv(f12_sub(x2, x1)))
        return f12_sub(f12_mul(m, f12_sub(xt, x1)), f12_sub(yt, y1))
    if y1 == y2:
        m = f12_mul(f12_muls(f12_mul(x1, x1), 3), f12_inv(f12_muls(y1, 2)))
        return f12_sub(f12_mul(m, f12_sub(xt, x1)), f12_sub(yt, y1))
    return f12_sub(xt, x1)


def miller_loop(q2, p1):
    """Unreduced pairing value of (G2 point q2, G1 point p1); 1 for infinity."""
    if q2 is None or p1 is None:
        return f12_one()
    Qt = _twist(q2)
    Pt = _cast_g1(p1)
    Rt = Qt
    f = f12_one()
    for i in range(LOG_ATE_LOOP_COUNT, -1, -1):
        f = f12_mul(f12_mul(f, f), _linefunc(Rt, Rt, Pt))
        Rt = _ec12_double(Rt)
        if ATE_LOOP_COUNT & (1 << i):
            f = f12_mul(f, _linefunc(Rt, Qt, Pt))
            Rt = _ec12_add(Rt, Qt)
    Q1 = (f12_pow(Qt[0], Q), f12_pow(Qt[1], Q))
    nQ2 = (f12_pow(Q1[0], Q), [(-c) % Q for c in f12_pow(Q1[1], Q)])
    f = f12_mul(f, _linefunc(Rt, Q1, Pt))
    Rt = _ec12_add(Rt, Q1)
    f = f12_mul(f, _linefunc(Rt, nQ2, Pt))
    return f


FINAL_EXP = (Q ** 12 - 1) // R


def final_exponentiate(f):
    return f12_pow(f, FINAL_EXP)


def pairing(q2, p1):
    """e(p1, q2) with the argument order py_ecc uses: pairing(Q in G2, P in G1)."""
    return final_exponentiate(miller_loop(q2, p1))


def f12_to_fp2_basis(x):
    """Fq12 element -> six Fq2 coefficients of w^0..w^5 (u = w^6 - 9).

    Same map as scripts/lib/field_helper.py:58-64 / scripts/sanitize_groth16_proof.py:31-37.
    """
    return [((x[i] + 9 * x[i + 6]) % Q, x[i + 6] % Q) for i in range(6)]


def f12_from_tower(t):
    """snarkjs/ffjavascript tower [[c00,c01,c02],[c10,c11,c12]] (Fq12 = Fq6[w]/(w^2-v),
    Fq6 = Fq2[v]/(v^3-(9+u))) -> dense w-polynomial.  v = w^2."""
    out = [0] * 12
    for j in range(2):
        for i in range(3):
            c0, c1 = int(t[j][i][0]), int(t[j][i][1])
            k = 2 * i + j
            out[k] = (out[k] + c0 - 9 * c1) % Q
            out[k + 6] = (out[k + 6] + c1) % Q
    return out


_BN_X = 4965661367192848881
# ffjavascript/wasmcurves compute the "hard part" of the final exponentiation by the
# Fuentes-Castaneda addition chain, which yields the pairing raised to the fixed
# multiple 2x(6x^2+3x+1); pinned by vk_alphabeta_12 in the reference's vkey files
# (tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json:52-81).
SNARKJS_GT_MULTIPLE = 2 * _BN_X * (6 * _BN_X * _BN_X + 3 * _BN_X + 1)


def pairing_snarkjs(q2, p1):
    """The GT element snarkjs prints as vk_alphabeta_12 (dense w-polynomial form)."""
    return f12_pow(pairing(q2, p1), SNARKJS_GT_MULTIPLE % R)
