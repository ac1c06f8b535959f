"""Groth16 prove, restated (oracle; test infrastructure).

Follows the data flow of SURVEY.md section 3.3 / rows a4-a16 (upstream homes: snarkjs 0.7.2
src/groth16_prove.js, ffjavascript 0.2.62 src/engine_{fft,multiexp,applykey}.js; call site
scripts/g16_prove.sh:246-260).  r and s are injectable (neither upstream prover offers that;
it is a test hook, row a14).
"""
import json

from . import bn254 as bn
from .binfile import ZKey  # noqa: F401  (re-export)

R = bn.R
TWO_ADICITY = 28
NQR = 5  # smallest quadratic non-residue of Fr (row a5)


def root_of_unity(log_n):
    """w[log_n]: w[28] = 5^((r-1)/2^28), w[k] = w[k+1]^2 (ffjavascript F1Field)."""
    w = pow(NQR, (R - 1) >> TWO_ADICITY, R)
    for _ in range(TWO_ADICITY - log_n):
        w = w * w % R
    return w


def ntt(a, w):
    """out[i] = sum_j a[j] w^(ij), natural order in and out (iterative radix 2)."""
    n = len(a)
    if n == 1:
        return list(a)
    logn = n.bit_length() - 1
    assert 1 << logn == n
    a = list(a)
    # bit reversal
    j = 0
    for i in range(1, n):
        bit = n >> 1
        while j & bit:
            j ^= bit
            bit >>= 1
        j |= bit
        if i < j:
            a[i], a[j] = a[j], a[i]
    length = 2
    while length <= n:
        wl = pow(w, n // length, R)
        half = length >> 1
        for start in range(0, n, length):
            t = 1
            for k in range(start, start + half):
                u = a[k]
                v = a[k + half] * t % R
                a[k] = (u + v) % R
                a[k + half] = (u - v) % R
                t = t * wl % R
        length <<= 1
    return a


def intt(a, w):
    n = len(a)
    ninv = bn.inv_mod(n, R)
    return [x * ninv % R for x in ntt(a, bn.inv_mod(w, R))]


def build_abc(zk, w):
    """Row a4 (buildABC1): sparse A,B evaluation over the domain; C = A*B pointwise."""
    n = zk.domain_size
    A = [0] * n
    B = [0] * n
    for m, c, s, coef in zk.coefs:
        if m == 0:
            A[c] = (A[c] + coef * w[s]) % R
        elif m == 1:
            B[c] = (B[c] + coef * w[s]) % R
        else:
            raise ValueError("bad matrix index in zkey coefficient")
    C = [x * y % R for x, y in zip(A, B)]
    return A, B, C


def h_scalars(zk, w):
    """Rows a5-a8: P[i] = A'(x_i) B'(x_i) - C'(x_i) on the odd coset x_i = w_2n^(2i+1)."""
    n = zk.domain_size
    log_n = n.bit_length() - 1
    A, B, C = build_abc(zk, w)
    wn = root_of_unity(log_n)
    inc = root_of_unity(log_n + 1) if log_n < TWO_ADICITY else NQR * NQR % R
    out = []
    for X in (A, B, C):
        coef = intt(X, wn)
        t = 1
        for i in range(n):
            coef[i] = coef[i] * t % R
            t = t * inc % R
        out.append(ntt(coef, wn))
    return [(a * b - c) % R for a, b, c in zip(*out)]


def prove_points(zk, w, r, s):
    """Rows a9-a15.  Returns affine (pi_a in G1, pi_b in G2, pi_c in G1)."""
    if len(w) != zk.n_vars:
        raise ValueError("Invalid witness length. Circuit: %d, witness: %d" % (zk.n_vars, len(w)))
    G1, G2 = bn.G1, bn.G2
    P = h_scalars(zk, w)
    a = G1.msm(zk.A, w)
    b1 = G1.msm(zk.B1, w)
    b2 = G2.msm(zk.B2, w)
    c = G1.msm(zk.C, w[zk.n_public + 1:])
    h = G1.msm(zk.H, P)
    pi_a = G1.add(G1.add(a, zk.alpha1), G1.mul(zk.delta1, r))
    pi_b = G2.add(G2.add(b2, zk.beta2), G2.mul(zk.delta2, s))
    pb1 = G1.add(G1.add(b1, zk.beta1), G1.mul(zk.delta1, s))
    pi_c = G1.add(c, h)
    pi_c = G1.add(pi_c, G1.mul(pi_a, s))
    pi_c = G1.add(pi_c, G1.mul(pb1, r))
    pi_c = G1.add(pi_c, G1.neg(G1.mul(zk.delta1, r * s % R)))
    return pi_a, pi_b, pi_c


# ---- output encoding (row a16) ---------------------------------------------
def proof_object(pi_a, pi_b, pi_c, style="rapidsnark"):
    def g1(p):
        return ["0", "1", "0"] if p is None else [str(p[0]), str(p[1]), "1"]

    def g2(p):
        if p is None:
            return [["0", "0"], ["1", "0"], ["0", "0"]]
        return [[str(p[0][0]), str(p[0][1])], [str(p[1][0]), str(p[1][1])], ["1", "0"]]

    obj = {"pi_a": g1(pi_a), "pi_b": g2(pi_b), "pi_c": g1(pi_c), "protocol": "groth16"}
    if style == "snarkjs":
        obj["curve"] = "bn128"
    return obj


def dumps(obj, style="rapidsnark"):
    """Byte layout of the two provers: rapidsnark = compact, snarkjs = JSON.stringify(x, null, 1);
    neither ends with a newline (see tests/golden/reference_vectors.json `formats`)."""
    if style == "snarkjs":
        return json.dumps(obj, indent=1)
    return json.dumps(obj, separators=(",", ":"))


def prove(zk, w, r, s, style="rapidsnark"):
    """-> (proof.json text, public.json text)."""
    pi_a, pi_b, pi_c = prove_points(zk, w, r, s)
    proof = proof_object(pi_a, pi_b, pi_c, style)
    public = [str(x) for x in w[1:zk.n_public + 1]]
    return dumps(proof, style), dumps(public, style)
