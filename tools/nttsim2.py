# index-math prototype of the split-twiddle pass (csrc/ntt.cuh, ntt_pass_kernel): a pass = pure 2^k-point
# sub-transforms with a 2^(k-1)-entry twiddle table + ONE inter-pass twiddle per element, checked against the oracle
import sys, random
sys.path.insert(0, '/root/repo')
from oracle import groth16 as g
R = g.R
TILE_LOG = 6   # small tile for the prototype (CUDA uses 10)

def plan(logn, maxk):
    np_ = max(1, -(-logn // maxk))
    base, rem = divmod(logn, np_)
    return [base + (1 if i < rem else 0) for i in range(np_)]

def brev(x, bits):
    return int(bin(x)[2:].zfill(bits)[::-1], 2) if bits else 0

def run_pass(a, logn, s_lo, k, tw, dif, inverse, scale=None):
    n = 1 << logn
    tile_log = max(TILE_LOG, k)
    tile = 1 << tile_log
    S = tile >> k
    logS = tile_log - k
    nq = n >> k
    half = n >> 1
    sh_tw = logn - s_lo - k
    # small table: P[j] = w_{2^k}^(+-j), j < 2^(k-1)
    P = []
    for j in range(max(1, 1 << (k - 1))):
        x = j << (logn - k)
        if inverse:
            P.append(1 if x == 0 else (R - tw[half - x]) % R)
        else:
            P.append(tw[x])
    def inter(lo, t):
        E = (lo * brev(t, k)) << sh_tw
        if E == 0: return None
        if inverse:
            return (R - tw[half - E]) % R if E <= half else tw[n - E]
        return tw[E] if E < half else (R - tw[E - half]) % R
    for q0 in range(0, nq, S):
        Sq = min(S, nq - q0)
        sh = [0] * tile
        for qi in range(Sq):
            q = q0 + qi
            hi, lo = q >> s_lo, q & ((1 << s_lo) - 1)
            for t in range(1 << k):
                idx = (hi << (s_lo + k)) + (t << s_lo) + lo
                v = a[idx]
                if not dif and s_lo > 0:
                    w = inter(lo, t)
                    if w is not None: v = v * w % R
                sh[t * S + qi] = v
        def bf(e0, e1, w, unit):
            x, y = sh[e0], sh[e1]
            if dif:
                d = (x - y) % R
                sh[e0] = (x + y) % R
                sh[e1] = d if unit else d * w % R
            else:
                if not unit: y = y * w % R
                sh[e0] = (x + y) % R; sh[e1] = (x - y) % R
        def single(r):
            for u in range(tile // 2):
                qi = u & (S - 1)
                v = u >> logS
                th = v & ((1 << (k - 1 - r)) - 1)
                tl = v >> (k - 1 - r)
                if qi >= Sq: continue
                t0 = (th << (r + 1)) | tl
                bf(t0 * S + qi, (t0 + (1 << r)) * S + qi, P[tl << (k - 1 - r)], tl == 0)
        def pair(r):   # levels r and r+1 on 4 elements held by one thread
            for gidx in range(tile // 4):
                qi = gidx & (S - 1)
                v = gidx >> logS
                th = v & ((1 << (k - 2 - r)) - 1)
                tl = v >> (k - 2 - r)
                if qi >= Sq: continue
                t00 = (th << (r + 2)) | tl
                a = [(t00 + d) * S + qi for d in (0, 1 << r, 2 << r, 3 << r)]
                w = P[tl << (k - 1 - r)]
                w1 = P[tl << (k - 2 - r)]
                w2 = P[(tl + (1 << r)) << (k - 2 - r)]
                if dif:
                    bf(a[0], a[2], w1, tl == 0); bf(a[1], a[3], w2, False)
                    bf(a[0], a[1], w, tl == 0); bf(a[2], a[3], w, tl == 0)
                else:
                    bf(a[0], a[1], w, tl == 0); bf(a[2], a[3], w, tl == 0)
                    bf(a[0], a[2], w1, tl == 0); bf(a[1], a[3], w2, False)
        if dif:
            r = k - 1
            if k & 1:
                single(r); r -= 1
            while r >= 1:
                pair(r - 1); r -= 2
        else:
            r = 0
            while r + 1 < k:
                pair(r); r += 2
            if k & 1: single(k - 1)
        for qi in range(Sq):
            q = q0 + qi
            hi, lo = q >> s_lo, q & ((1 << s_lo) - 1)
            for t in range(1 << k):
                idx = (hi << (s_lo + k)) + (t << s_lo) + lo
                v = sh[t * S + qi]
                if dif and s_lo > 0:
                    w = inter(lo, t)
                    if w is not None: v = v * w % R
                if scale is not None: v = v * scale[idx] % R
                a[idx] = v

def coset_pipeline(X, logn, maxk):
    n = 1 << logn
    w = g.root_of_unity(logn)
    tw = [pow(w, e, R) for e in range(max(1, n // 2))]
    inc = g.root_of_unity(logn + 1)
    ninv = pow(n, -1, R)
    scale = [pow(inc, brev(p, logn), R) * ninv % R for p in range(n)]
    a = list(X)
    ks = plan(logn, maxk)
    s = logn
    for i, k in enumerate(ks):
        s -= k
        run_pass(a, logn, s, k, tw, True, True, scale if s == 0 else None)
    s = 0
    for k in reversed(ks):
        run_pass(a, logn, s, k, tw, False, False)
        s += k
    return a

for logn, maxk in ((1,3),(2,3),(3,3),(5,3),(7,3),(8,3),(9,4),(6,6),(10,5),(4,6),(9,3)):
    n = 1 << logn
    X = [random.randrange(R) for _ in range(n)]
    w = g.root_of_unity(logn); inc = g.root_of_unity(logn + 1)
    coef = g.intt(X, w); t = 1
    for i in range(n):
        coef[i] = coef[i] * t % R; t = t * inc % R
    want = g.ntt(coef, w)
    got = coset_pipeline(X, logn, maxk)
    assert got == want, (logn, maxk)
print("split-twiddle prototype ok")
