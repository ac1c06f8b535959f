# index-math prototype of csrc/ntt.cu (pass/tile structure), checked against the oracle NTT
import sys, random
sys.path.insert(0, '/root/repo')
from oracle import groth16 as g
R = g.R
TILE = 64   # small tile for the prototype (CUDA uses 1024)

def plan(logn, maxk):
    np_ = max(1, -(-logn // maxk))
    base, rem = divmod(logn, np_)
    ks = [base + (1 if i < rem else 0) for i in range(np_)]
    return ks

def run_pass(a, logn, s_lo, k, tw, dif, inverse, scale=None):
    n = 1 << logn
    tile = max(TILE, 1 << k)
    S = tile >> k
    nq = n >> k
    for q0 in range(0, nq, S):
        Sq = min(S, nq - q0)
        sh = [0] * tile
        def gidx(e):
            if s_lo > 0:
                qi, t = e % S, e // S
            else:
                t, qi = e % (1 << k), e >> k
            q = q0 + qi
            hi, lo = q >> s_lo, q & ((1 << s_lo) - 1)
            return qi, t, (hi << (s_lo + k)) + (t << s_lo) + lo, lo
        for e in range(tile):
            qi, t, idx, lo = gidx(e)
            if qi < Sq: sh[e] = a[idx]
        rs = range(k - 1, -1, -1) if dif else range(k)
        for r in rs:
            b = s_lo + r
            for u in range(tile // 2):
                if s_lo > 0:
                    qi, tp = u % S, u // S
                else:
                    tp, qi = u % (1 << (k - 1)), u >> (k - 1)
                if qi >= Sq: continue
                tl, th = tp & ((1 << r) - 1), tp >> r
                t0 = (th << (r + 1)) | tl
                t1 = t0 + (1 << r)
                q = q0 + qi
                lo = q & ((1 << s_lo) - 1)
                j = (tl << s_lo) + lo
                ex = j << (logn - 1 - b)
                if inverse:
                    w = 1 if ex == 0 else (R - tw[(n >> 1) - ex]) % R
                else:
                    w = tw[ex]
                e0 = (t0 * S + qi) if s_lo > 0 else (qi << k) + t0
                e1 = (t1 * S + qi) if s_lo > 0 else (qi << k) + t1
                x, y = sh[e0], sh[e1]
                if dif:
                    sh[e0] = (x + y) % R; sh[e1] = (x - y) * w % R
                else:
                    y = y * w % R
                    sh[e0] = (x + y) % R; sh[e1] = (x - y) % R
        for e in range(tile):
            qi, t, idx, lo = gidx(e)
            if qi < Sq:
                v = sh[e]
                if scale is not None: v = v * scale[idx] % R
                a[idx] = v

def brev(x, bits):
    return int(bin(x)[2:].zfill(bits)[::-1], 2) if bits else 0

def coset_pipeline(X, logn, maxk):
    n = 1 << logn
    w = g.root_of_unity(logn)
    tw = [pow(w, e, R) for e in range(max(1, n // 2))]
    inc = g.root_of_unity(logn + 1)
    ninv = pow(n, -1, R)
    scale = [pow(inc, brev(p, logn), R) * ninv % R for p in range(n)]
    a = list(X)
    ks = plan(logn, maxk)
    # DIF inverse: top stages first
    s = logn
    for i, k in enumerate(ks):
        s -= k
        run_pass(a, logn, s, k, tw, True, True, scale if s == 0 else None)
    # DIT forward: low stages first
    s = 0
    for k in reversed(ks):
        run_pass(a, logn, s, k, tw, False, False)
        s += k
    return a

for logn, maxk in ((1,3),(2,3),(3,3),(5,3),(7,3),(8,3),(9,4),(6,6),(10,5)):
    n = 1 << logn
    X = [random.randrange(R) for _ in range(n)]
    w = g.root_of_unity(logn); inc = g.root_of_unity(logn + 1)
    coef = g.intt(X, w); t = 1
    for i in range(n):
        coef[i] = coef[i] * t % R; t = t * inc % R
    want = g.ntt(coef, w)
    got = coset_pipeline(X, logn, maxk)
    assert got == want, (logn, maxk)
print("ntt prototype ok")
