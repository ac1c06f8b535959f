// Native Groth16 verifier over BN254, host C++ only, no CUDA.
//
// Replaces `snarkjs groth16 verify <vkey.json> <public.json> <proof.json>` (scripts/g16_verify.sh:212-216) -- the
// consumer of this prover's output, 0.3 s per proof in node -- so that the same binary that proves can check what it
// wrote.  Same equation as git_modules/circom-pairing/circuits/bn254/groth16.circom:79-125 and
// git_modules/ed25519-circom/contracts/test-verifier.sol:305-323:
//     e(-A, B) * e(alpha, beta) * e(IC0 + sum pub_i IC_{i+1}, gamma) * e(C, delta) == 1.
// The pairing is the optimal ate pairing written the plain way: Fq12 = Fq[w]/(w^12 - 18 w^6 + 82), G2 points moved to
// Fq12 by the twist (x w^2, y w^3) with u = w^6 - 9, affine line functions, final exponentiation by (q^12 - 1)/r.
// Clarity over speed: ~0.2 s per proof on one core.  Checked against every (vkey, public, proof) triple the reference
// commits (tests/golden/reference_vectors.json) and against tampered copies (tests/test_verify_native.py).
#pragma once
#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include "hostfield.hpp"

namespace g16v {
using namespace g16h;

typedef HFq::el fq;
typedef std::array<fq, 12> f12;

inline fq fq_small(uint64_t x) { return HFq::from_u64(x); }

inline f12 f12_zero() {
    f12 r;
    for (auto& c : r) c = HFq::zero();
    return r;
}
inline f12 f12_one() {
    f12 r = f12_zero();
    r[0] = HFq::one();
    return r;
}
inline bool f12_eq(const f12& a, const f12& b) {
    for (int i = 0; i < 12; i++)
        if (!HFq::eq(a[i], b[i])) return false;
    return true;
}
inline bool f12_is_zero(const f12& a) { return f12_eq(a, f12_zero()); }
inline f12 f12_add(const f12& a, const f12& b) {
    f12 r;
    for (int i = 0; i < 12; i++) r[i] = HFq::add(a[i], b[i]);
    return r;
}
inline f12 f12_sub(const f12& a, const f12& b) {
    f12 r;
    for (int i = 0; i < 12; i++) r[i] = HFq::sub(a[i], b[i]);
    return r;
}
inline f12 f12_neg(const f12& a) { return f12_sub(f12_zero(), a); }
inline f12 f12_muls(const f12& a, const fq& s) {
    f12 r;
    for (int i = 0; i < 12; i++) r[i] = HFq::mul(a[i], s);
    return r;
}

// schoolbook product, then w^12 = 18 w^6 - 82
inline f12 f12_mul(const f12& a, const f12& b) {
    static const fq c18 = fq_small(18), c82 = fq_small(82);
    fq t[23];
    for (auto& x : t) x = HFq::zero();
    for (int i = 0; i < 12; i++) {
        if (HFq::is_zero(a[i])) continue;
        for (int j = 0; j < 12; j++) t[i + j] = HFq::add(t[i + j], HFq::mul(a[i], b[j]));
    }
    for (int k = 22; k >= 12; k--) {
        if (HFq::is_zero(t[k])) continue;
        t[k - 6] = HFq::add(t[k - 6], HFq::mul(c18, t[k]));
        t[k - 12] = HFq::sub(t[k - 12], HFq::mul(c82, t[k]));
    }
    f12 r;
    for (int i = 0; i < 12; i++) r[i] = t[i];
    return r;
}

// exponent as little-endian 64-bit words
inline f12 f12_pow(const f12& a, const std::vector<uint64_t>& e) {
    f12 out = f12_one(), base = a;
    size_t top = e.size();
    while (top && e[top - 1] == 0) top--;
    for (size_t i = 0; i < top * 64; i++) {
        if ((e[i >> 6] >> (i & 63)) & 1) out = f12_mul(out, base);
        base = f12_mul(base, base);
    }
    return out;
}

// inverse by linear algebra: column j of M is a * w^j; solve M x = 1 (Gauss-Jordan over Fq, 12 x 12)
inline f12 f12_inv(const f12& a) {
    fq m[12][13];
    f12 col = a, wj = f12_zero();
    wj[1] = HFq::one();  // w
    for (int j = 0; j < 12; j++) {
        for (int i = 0; i < 12; i++) m[i][j] = col[i];
        col = f12_mul(col, wj);
    }
    for (int i = 0; i < 12; i++) m[i][12] = i == 0 ? HFq::one() : HFq::zero();
    for (int c = 0; c < 12; c++) {
        int p = c;
        while (p < 12 && HFq::is_zero(m[p][c])) p++;
        if (p == 12) return f12_zero();  // not invertible (a == 0)
        if (p != c)
            for (int k = 0; k < 13; k++) std::swap(m[p][k], m[c][k]);
        fq inv = HFq::inv(m[c][c]);
        for (int k = c; k < 13; k++) m[c][k] = HFq::mul(m[c][k], inv);
        for (int r = 0; r < 12; r++) {
            if (r == c || HFq::is_zero(m[r][c])) continue;
            fq f = m[r][c];
            for (int k = c; k < 13; k++) m[r][k] = HFq::sub(m[r][k], HFq::mul(f, m[c][k]));
        }
    }
    f12 x;
    for (int i = 0; i < 12; i++) x[i] = m[i][12];
    return x;
}

}  // namespace g16v
