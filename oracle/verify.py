"""Groth16 verifier + proof/vkey decoding (oracle; test infrastructure).

Equation: e(-A,B) * e(alpha,beta) * e(IC0 + sum pub_i IC_{i+1}, gamma) * e(C,delta) = 1,
the same check as scripts/g16_verify.sh:212-216 (`snarkjs groth16 verify`),
git_modules/circom-pairing/circuits/bn254/groth16.circom:79-125 and
git_modules/ed25519-circom/contracts/test-verifier.sol:305-323.
"""
from . import bn254 as bn


def g1_from_json(p):
    x, y, z = (int(v) for v in p)
    if z == 0:
        return None
    assert z == 1
    return (x % bn.Q, y % bn.Q)


def g2_from_json(p):
    x, y, z = ((int(v[0]), int(v[1])) for v in p)
    if z == (0, 0):
        return None
    assert z == (1, 0)
    return (x, y)


def verify(vkey, public, proof):
    """True iff the proof verifies.  All inputs are the parsed JSON objects."""
    A = g1_from_json(proof["pi_a"])
    B = g2_from_json(proof["pi_b"])
    C = g1_from_json(proof["pi_c"])
    alpha = g1_from_json(vkey["vk_alpha_1"])
    beta = g2_from_json(vkey["vk_beta_2"])
    gamma = g2_from_json(vkey["vk_gamma_2"])
    delta = g2_from_json(vkey["vk_delta_2"])
    ic = [g1_from_json(p) for p in vkey["IC"]]
    if len(public) + 1 != len(ic):
        return False
    for pt in (A, C, alpha, *ic):
        if not bn.G1.is_on_curve(pt):
            return False
    for pt in (B, beta, gamma, delta):
        if not bn.G2.is_on_curve(pt):
            return False
    acc = bn.G1.to_jac(ic[0])
    for s, p in zip(public, ic[1:]):
        s = int(s)
        if not 0 <= s < bn.R:
            return False
        acc = bn.G1.jadd(acc, bn.G1.jmul(bn.G1.to_jac(p), s))
    vkx = bn.G1.to_affine(acc)
    f = bn.miller_loop(B, bn.G1.neg(A))
    f = bn.f12_mul(f, bn.miller_loop(beta, alpha))
    f = bn.f12_mul(f, bn.miller_loop(gamma, vkx))
    f = bn.f12_mul(f, bn.miller_loop(delta, C))
    return bn.final_exponentiate(f) == bn.f12_one()


def limbs(x, n=43, k=6):
    """scripts/lib/field_helper.py:17-24 numberToArray."""
    out = []
    for _ in range(k):
        out.append(x % (1 << n))
        x >>= n
    return out


def from_limbs(l, n=43):
    return sum(int(v) << (n * i) for i, v in enumerate(l))


def sanitize(vkey, public, proof, n=43, k=6):
    """Restatement of scripts/sanitize_groth16_proof.py:39-123 (proof + vkey + public ->
    the 43-bit x 6 limb input of the recursive verifier circuit)."""
    A = g1_from_json(proof["pi_a"])
    B = g2_from_json(proof["pi_b"])
    C = g1_from_json(proof["pi_c"])
    alpha = g1_from_json(vkey["vk_alpha_1"])
    beta = g2_from_json(vkey["vk_beta_2"])
    gamma = g2_from_json(vkey["vk_gamma_2"])
    delta = g2_from_json(vkey["vk_delta_2"])
    f1 = lambda v: limbs(v, n, k)
    f2 = lambda v: [limbs(v[0], n, k), limbs(v[1], n, k)]
    nab = bn.pairing(beta, bn.G1.neg(alpha))
    return {
        "gamma2": [f2(gamma[0]), f2(gamma[1])],
        "delta2": [f2(delta[0]), f2(delta[1])],
        "negalfa1xbeta2": [f2(c) for c in bn.f12_to_fp2_basis(nab)],
        "IC": [[f1(p[0]), f1(p[1])] for p in (g1_from_json(q) for q in vkey["IC"])],
        "negpa": [f1(A[0]), f1((-A[1]) % bn.Q)],
        "pb": [f2(B[0]), f2(B[1])],
        "pc": [f1(C[0]), f1(C[1])],
        "pubInput": [int(v) for v in public],
    }


def vkey_from_sanitized(san, alpha_beta_only=False):
    """Rebuild what can be rebuilt of a vkey from a sanitized_proof.json (the 1-sig
    test commits no vkey files; gamma2, delta2, IC and e(-alpha,beta) are embedded)."""
    g2 = lambda v: ((from_limbs(v[0][0]), from_limbs(v[0][1])), (from_limbs(v[1][0]), from_limbs(v[1][1])))
    g1 = lambda v: (from_limbs(v[0]), from_limbs(v[1]))
    nab6 = [(from_limbs(c[0]), from_limbs(c[1])) for c in san["negalfa1xbeta2"]]
    nab = [0] * 12
    for i, (c0, c1) in enumerate(nab6):
        nab[i] = (c0 - 9 * c1) % bn.Q
        nab[i + 6] = c1 % bn.Q
    return {"gamma": g2(san["gamma2"]), "delta": g2(san["delta2"]),
            "IC": [g1(p) for p in san["IC"]], "neg_alpha_beta": nab}


def verify_sanitized(san):
    """Verify the proof embedded in a sanitized_proof.json against its embedded vkey:
    e(negpa, pb) * e(-alpha,beta)^-1 ... arranged as
    e(-A,B) * e(vkx,gamma) * e(C,delta) == e(-alpha,beta)   (sign as the circuit uses it:
    git_modules/circom-pairing/circuits/bn254/groth16.circom:79-125)."""
    vk = vkey_from_sanitized(san)
    g1 = lambda v: (from_limbs(v[0]), from_limbs(v[1]))
    negA = g1(san["negpa"])
    B = ((from_limbs(san["pb"][0][0]), from_limbs(san["pb"][0][1])),
         (from_limbs(san["pb"][1][0]), from_limbs(san["pb"][1][1])))
    C = g1(san["pc"])
    acc = bn.G1.to_jac(vk["IC"][0])
    for s, p in zip(san["pubInput"], vk["IC"][1:]):
        acc = bn.G1.jadd(acc, bn.G1.jmul(bn.G1.to_jac(p), int(s)))
    vkx = bn.G1.to_affine(acc)
    f = bn.miller_loop(B, negA)
    f = bn.f12_mul(f, bn.miller_loop(vk["gamma"], vkx))
    f = bn.f12_mul(f, bn.miller_loop(vk["delta"], C))
    lhs = bn.final_exponentiate(f)
    # e(-A,B) e(vkx,gamma) e(C,delta) * e(alpha,beta) = 1  <=>  lhs = e(alpha,beta)^-1 = e(-alpha,beta)
    return lhs == vk["neg_alpha_beta"]
