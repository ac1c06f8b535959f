# Simulation of the even/odd wide-MAD Montgomery multiplication (design check for csrc/field.cuh)
import random
Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
M32 = (1<<32)-1
def limbs(x): return [(x>>(32*i))&M32 for i in range(8)]
def val(l): return sum(v<<(32*i) for i,v in enumerate(l))

def chain(acc, xs, y):
    """acc[0..7] += xs[0..3]*y in pairs; returns carry out (mad.lo.cc / madc.hi.cc ... addc)"""
    cf = 0
    for k in range(4):
        prod = xs[k]*y
        lo, hi = prod & M32, prod >> 32
        t = acc[2*k] + lo + cf; acc[2*k] = t & M32; cf = t >> 32
        t = acc[2*k+1] + hi + cf; acc[2*k+1] = t & M32; cf = t >> 32
    return cf

def chain_shift(src, xs, y, cf):
    """out[j] = prod parts + src[j+2] + carry; src has 8 limbs (src[8],src[9] = 0)"""
    out = [0]*8
    s = src + [0, 0]
    for k in range(4):
        prod = xs[k]*y
        lo, hi = prod & M32, prod >> 32
        t = s[2*k+2] + lo + cf; out[2*k] = t & M32; cf = t >> 32
        t = s[2*k+3] + hi + cf; out[2*k+1] = t & M32; cf = t >> 32
    assert cf == 0
    return out

def montmul(a, b, p):
    inv = (-pow(p, -1, 1<<32)) & M32
    A, B, P = limbs(a), limbs(b), limbs(p)
    Ae, Ao = A[0::2], A[1::2]
    Pe, Po = P[0::2], P[1::2]
    E = [0]*8; O = [0]*8
    for i in range(8):
        if i == 0:
            chain(E, Ae, B[0]); chain(O, Ao, B[0])
        else:
            En = O
            t = En[0] + E[1]; En[0] = t & M32; cf = t >> 32
            On = chain_shift(E, Ao, B[i], cf)
            c = chain(En, Ae, B[i])
            On[7] += c; assert On[7] <= M32
            E, O = En, On
        m = (E[0]*inv) & M32
        co = chain(O, Po, m); assert co == 0
        c = chain(E, Pe, m)
        O[7] += c; assert O[7] <= M32
        assert E[0] == 0
    r = []; cf = 0
    for k in range(8):
        t = O[k] + (E[k+1] if k < 7 else 0) + cf; r.append(t & M32); cf = t >> 32
    assert cf == 0
    v = val(r)
    assert v < 2*p
    if v >= p: v -= p
    return v

if __name__ == "__main__":
    for p in (Q, R):
        Rm = (1<<256) % p
        Rinv = pow(Rm, -1, p)
        edge = [0, 1, p-1, p-2, Rm, (Rm*Rm)%p, 2, (1<<253), p>>1]
        cases = [(x,y) for x in edge for y in edge] + [(random.randrange(p), random.randrange(p)) for _ in range(20000)]
        for a,b in cases:
            assert montmul(a,b,p) == a*b*Rinv % p, (a,b)
    print("ok")
