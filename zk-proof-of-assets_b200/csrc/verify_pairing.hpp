// optimal ate pairing on BN254 (see verify.hpp for the Fq12 model: Fq12 = Fq[w]/(w^12 - 18 w^6 + 82), u = w^6 - 9)
//
// Round 2: the Miller loop runs on the twist.  A G2 point stays an affine pair over Fq2 (its image in E(Fq12) is
// (x w^2, y w^3)), the slope is an Fq2 element, and the line through T evaluated at P = (xP, yP) in G1 is the sparse
// element   yP + (-lambda xP) w + (lambda xT - yT) w^3   (five non-zero Fq coefficients), multiplied into f with the
// zero-skipping product.  Several pairs share ONE accumulator (one squaring per bit for the whole product of pairings)
// and one Fq2 inversion per step (Montgomery's trick across the pairs).  The final exponentiation is split as
// (q^6 - 1)(q^2 + 1) -- a conjugation, one inversion and a coefficient-wise Frobenius -- times (q^4 - q^2 + 1)/r by a
// windowed power with dedicated squarings.  8 ms per Groth16 verification on one core where the first version
// (points and lines as dense Fq12 elements, one 2800-bit power) took 300 ms.  Lines differ from the textbook ones by
// factors in Fq, which the final exponentiation removes.
#pragma once
#include <utility>

#include "verify.hpp"

namespace g16v {

typedef HFq2El fq2;

// an Fq2 element c0 + c1 u times w^k, added into an Fq12 element: (c0 - 9 c1) w^k + c1 w^(k+6), k + 6 < 12
inline void f12_add_fq2(f12& r, int k, const fq2& c) {
    static const fq nine = fq_small(9);
    r[k] = HFq::add(r[k], HFq::sub(c.c0, HFq::mul(nine, c.c1)));
    r[k + 6] = HFq::add(r[k + 6], c.c1);
}

inline fq2 fq2_pow(const fq2& a, const std::vector<uint64_t>& e) {
    fq2 out = HFq2::one(), base = a;
    for (size_t i = 0; i < e.size() * 64; i++) {
        if ((e[i >> 6] >> (i & 63)) & 1) out = HFq2::mul(out, base);
        base = HFq2::sqr(base);
    }
    return out;
}
inline fq2 fq2_conj(const fq2& a) { return fq2{a.c0, HFq::neg(a.c1)}; }

inline std::vector<uint64_t> words_of(const U256& a) { return {a.v[0], a.v[1], a.v[2], a.v[3]}; }

// little-endian multi-word helpers for the few exponents computed once
inline std::vector<uint64_t> big_mul(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
    std::vector<uint64_t> r(a.size() + b.size(), 0);
    for (size_t i = 0; i < a.size(); i++) {
        unsigned __int128 c = 0;
        for (size_t j = 0; j < b.size(); j++) {
            c += (unsigned __int128)a[i] * b[j] + r[i + j];
            r[i + j] = (uint64_t)c;
            c >>= 64;
        }
        r[i + b.size()] += (uint64_t)c;
    }
    return r;
}
inline std::vector<uint64_t> big_sub(std::vector<uint64_t> a, const std::vector<uint64_t>& b) {   // a >= b
    unsigned __int128 bw = 0;
    for (size_t k = 0; k < a.size(); k++) {
        unsigned __int128 d = (unsigned __int128)a[k] - (k < b.size() ? b[k] : 0) - (uint64_t)bw;
        a[k] = (uint64_t)d;
        bw = (d >> 64) & 1;
    }
    return a;
}
inline std::vector<uint64_t> big_add_small(std::vector<uint64_t> a, uint64_t x) {
    a.push_back(0);
    for (size_t k = 0; k < a.size() && x; k++) {
        unsigned __int128 s = (unsigned __int128)a[k] + x;
        a[k] = (uint64_t)s;
        x = (uint64_t)(s >> 64);
    }
    return a;
}
// floor(p / d), bit by bit (done once per exponent); d small (r, 2, 3, 6 as words)
inline std::vector<uint64_t> big_div(const std::vector<uint64_t>& p, const std::vector<uint64_t>& d) {
    std::vector<uint64_t> quo(p.size(), 0), rem(d.size() + 1, 0);
    for (int i = int(p.size()) * 64 - 1; i >= 0; i--) {
        uint64_t carry = (p[i >> 6] >> (i & 63)) & 1;
        for (auto& w : rem) {
            uint64_t nc = w >> 63;
            w = (w << 1) | carry;
            carry = nc;
        }
        bool ge = rem[d.size()] != 0;
        if (!ge) {
            ge = true;
            for (int k = int(d.size()) - 1; k >= 0; k--)
                if (rem[k] != d[k]) {
                    ge = rem[k] > d[k];
                    break;
                }
        }
        if (ge) {
            rem = big_sub(rem, d);
            quo[i >> 6] |= uint64_t(1) << (i & 63);
        }
    }
    return quo;
}

// constants of the Frobenius maps, computed once: xi = 9 + u = w^6
struct FrobConsts {
    fq2 g2, g3;        // xi^((q-1)/3), xi^((q-1)/2): pi(x, y) = (conj(x) g2, conj(y) g3) on the twist
    fq zeta[12];       // zeta^k, zeta = xi^((q^2-1)/6) in Fq: f^(q^2) scales the coefficient of w^k by zeta^k
};
inline const FrobConsts& frob() {
    static const FrobConsts C = [] {
        FrobConsts c;
        fq2 xi{fq_small(9), HFq::one()};
        std::vector<uint64_t> q = words_of(HFq::mod()), qm1 = q;
        qm1[0] -= 1;
        c.g2 = fq2_pow(xi, big_div(qm1, {3}));
        c.g3 = fq2_pow(xi, big_div(qm1, {2}));
        std::vector<uint64_t> q2m1 = big_mul(q, q);
        q2m1[0] -= 1;
        fq2 z = fq2_pow(xi, big_div(q2m1, {6}));   // a sixth root of unity: lies in Fq because 6 | q - 1
        c.zeta[0] = HFq::one();
        for (int k = 1; k < 12; k++) c.zeta[k] = HFq::mul(c.zeta[k - 1], z.c0);
        return c;
    }();
    return C;
}

// one (Q in G2, P in G1) pair of a product of pairings
struct MillerPair {
    fq2 qx, qy;     // Q on the twist
    fq px, py;      // P
    fq2 tx, ty;     // the running point T
    bool t_inf = false;
};

// f *= line with slope lam through T, evaluated at P
inline void mul_by_line(f12& f, const MillerPair& m, const fq2& lam) {
    f12 l = f12_zero();
    l[0] = m.py;
    f12_add_fq2(l, 1, fq2{HFq::neg(HFq::mul(lam.c0, m.px)), HFq::neg(HFq::mul(lam.c1, m.px))});
    f12_add_fq2(l, 3, HFq2::sub(HFq2::mul(lam, m.tx), m.ty));
    f = f12_mul(l, f);   // the product skips the zero coefficients of its first operand
}
// f *= the vertical line x - xT at P
inline void mul_by_vertical(f12& f, const MillerPair& m) {
    f12 l = f12_zero();
    l[0] = m.px;
    f12_add_fq2(l, 2, HFq2::neg(m.tx));
    f = f12_mul(l, f);
}

// denominators of this step's slopes, inverted together (Montgomery's trick); zero entries are skipped
inline void batch_inverse(std::vector<fq2>& d) {
    std::vector<fq2> pre(d.size());
    fq2 acc = HFq2::one();
    for (size_t i = 0; i < d.size(); i++) {
        pre[i] = acc;
        if (!HFq2::is_zero(d[i])) acc = HFq2::mul(acc, d[i]);
    }
    fq2 inv = HFq2::inv(acc);
    for (size_t i = d.size(); i-- > 0;) {
        if (HFq2::is_zero(d[i])) continue;
        fq2 di = HFq2::mul(inv, pre[i]);
        inv = HFq2::mul(inv, d[i]);
        d[i] = di;
    }
}

// T <- 2T with the tangent line, for every pair
inline void step_double(f12& f, std::vector<MillerPair>& ps) {
    std::vector<fq2> den(ps.size());
    for (size_t i = 0; i < ps.size(); i++) den[i] = ps[i].t_inf ? HFq2::zero() : HFq2::dbl(ps[i].ty);
    batch_inverse(den);
    for (size_t i = 0; i < ps.size(); i++) {
        MillerPair& m = ps[i];
        if (m.t_inf) continue;
        if (HFq2::is_zero(m.ty)) {   // a point of order two: vertical tangent
            mul_by_vertical(f, m);
            m.t_inf = true;
            continue;
        }
        fq2 xx = HFq2::sqr(m.tx);
        fq2 lam = HFq2::mul(HFq2::add(HFq2::dbl(xx), xx), den[i]);
        mul_by_line(f, m, lam);
        fq2 x3 = HFq2::sub(HFq2::sqr(lam), HFq2::dbl(m.tx));
        m.ty = HFq2::sub(HFq2::mul(lam, HFq2::sub(m.tx, x3)), m.ty);
        m.tx = x3;
    }
}

// T <- T + S_i with the chord, for every pair (S given per pair)
inline void step_add(f12& f, std::vector<MillerPair>& ps, const std::vector<std::pair<fq2, fq2>>& S) {
    std::vector<fq2> den(ps.size());
    for (size_t i = 0; i < ps.size(); i++) {
        den[i] = HFq2::zero();
        if (ps[i].t_inf) continue;
        fq2 dx = HFq2::sub(S[i].first, ps[i].tx);
        den[i] = HFq2::is_zero(dx) ? (HFq2::eq(S[i].second, ps[i].ty) ? HFq2::dbl(ps[i].ty) : HFq2::zero()) : dx;
    }
    batch_inverse(den);
    for (size_t i = 0; i < ps.size(); i++) {
        MillerPair& m = ps[i];
        if (m.t_inf) {   // infinity + S = S, the line is 1
            m.tx = S[i].first;
            m.ty = S[i].second;
            m.t_inf = false;
            continue;
        }
        fq2 dx = HFq2::sub(S[i].first, m.tx), lam;
        if (HFq2::is_zero(dx)) {
            if (!HFq2::eq(S[i].second, m.ty) || HFq2::is_zero(m.ty)) {   // S = -T: vertical line, T + S = infinity
                mul_by_vertical(f, m);
                m.t_inf = true;
                continue;
            }
            fq2 xx = HFq2::sqr(m.tx);
            lam = HFq2::mul(HFq2::add(HFq2::dbl(xx), xx), den[i]);
        } else {
            lam = HFq2::mul(HFq2::sub(S[i].second, m.ty), den[i]);
        }
        mul_by_line(f, m, lam);
        fq2 x3 = HFq2::sub(HFq2::sub(HFq2::sqr(lam), m.tx), S[i].first);
        m.ty = HFq2::sub(HFq2::mul(lam, HFq2::sub(m.tx, x3)), m.ty);
        m.tx = x3;
    }
}

// dedicated squaring: the 66 cross products once, doubled, plus the 12 squares (78 products instead of 144)
inline f12 f12_sqr(const f12& a) {
    static const fq c18 = fq_small(18), c82 = fq_small(82);
    fq t[23];
    for (auto& x : t) x = HFq::zero();
    for (int i = 0; i < 12; i++)
        for (int j = i + 1; j < 12; j++) t[i + j] = HFq::add(t[i + j], HFq::mul(a[i], a[j]));
    for (auto& x : t) x = HFq::dbl(x);
    for (int i = 0; i < 12; i++) t[2 * i] = HFq::add(t[2 * i], HFq::sqr(a[i]));
    for (int k = 22; k >= 12; k--) {
        t[k - 6] = HFq::add(t[k - 6], HFq::mul(c18, t[k]));
        t[k - 12] = HFq::sub(t[k - 12], HFq::mul(c82, t[k]));
    }
    f12 r;
    for (int i = 0; i < 12; i++) r[i] = t[i];
    return r;
}

// unreduced value of prod_i e(P_i, Q_i); pairs with a point at infinity contribute 1
struct PairingInput {
    HAffine<HFq2> q;
    HAffine<HFq> p;
};
inline f12 miller_loop_multi(const std::vector<PairingInput>& in) {
    std::vector<MillerPair> ps;
    for (const PairingInput& x : in) {
        if (x.q.inf || x.p.inf) continue;
        MillerPair m;
        m.qx = x.q.x; m.qy = x.q.y; m.px = x.p.x; m.py = x.p.y;
        m.tx = m.qx; m.ty = m.qy;
        ps.push_back(m);
    }
    f12 f = f12_one();
    if (ps.empty()) return f;
    // 6t + 2, t = 4965661367192848881: 0x1_9d797039be763ba8, bit 64 is the leading one
    const uint64_t ATE_LO = 0x9d797039be763ba8ULL;
    std::vector<std::pair<fq2, fq2>> Q(ps.size());
    for (size_t i = 0; i < ps.size(); i++) Q[i] = {ps[i].qx, ps[i].qy};
    for (int i = 63; i >= 0; i--) {
        f = f12_sqr(f);
        step_double(f, ps);
        if ((ATE_LO >> i) & 1) step_add(f, ps, Q);
    }
    // the two Frobenius corrections: T += pi(Q), then T += -pi^2(Q)
    const FrobConsts& C = frob();
    std::vector<std::pair<fq2, fq2>> Q1(ps.size()), nQ2(ps.size());
    for (size_t i = 0; i < ps.size(); i++) {
        Q1[i] = {HFq2::mul(fq2_conj(ps[i].qx), C.g2), HFq2::mul(fq2_conj(ps[i].qy), C.g3)};
        fq2 x2 = HFq2::mul(fq2_conj(Q1[i].first), C.g2), y2 = HFq2::mul(fq2_conj(Q1[i].second), C.g3);
        nQ2[i] = {x2, HFq2::neg(y2)};
    }
    step_add(f, ps, Q1);
    step_add(f, ps, nQ2);
    return f;
}
inline f12 miller_loop(const HAffine<HFq2>& q2, bool q_inf, const HAffine<HFq>& p1, bool p_inf) {
    if (q_inf || p_inf) return f12_one();
    return miller_loop_multi({PairingInput{q2, p1}});
}

// (q^4 - q^2 + 1) / r, the hard part of the final exponent
inline const std::vector<uint64_t>& hard_exponent() {
    static const std::vector<uint64_t> e = [] {
        std::vector<uint64_t> q = words_of(HFq::mod()), q2 = big_mul(q, q), q4 = big_mul(q2, q2);
        std::vector<uint64_t> num = big_add_small(big_sub(q4, q2), 1);
        std::vector<uint64_t> out = big_div(num, words_of(HFr::mod()));
        while (out.size() > 1 && out.back() == 0) out.pop_back();
        return out;
    }();
    return e;
}

// a^e with dedicated squarings and a 4-bit fixed window
inline f12 f12_pow_window(const f12& a, const std::vector<uint64_t>& e) {
    f12 tab[16];
    tab[0] = f12_one();
    for (int i = 1; i < 16; i++) tab[i] = f12_mul(tab[i - 1], a);
    int top = int(e.size()) * 64;
    while (top > 0 && !((e[(top - 1) >> 6] >> ((top - 1) & 63)) & 1)) top--;
    int nibbles = (top + 3) / 4;
    f12 out = f12_one();
    for (int n = nibbles - 1; n >= 0; n--) {
        for (int k = 0; k < 4; k++) out = f12_sqr(out);
        unsigned d = unsigned((e[(4 * n) >> 6] >> ((4 * n) & 63)) & 15);
        if (d) out = f12_mul(tab[d], out);
    }
    return out;
}

// f^((q^12 - 1) / r)
inline f12 final_exponentiate(const f12& f) {
    // q^6 - 1: w^(q^6) = -w, so f^(q^6) negates the odd coefficients
    f12 conj = f;
    for (int k = 1; k < 12; k += 2) conj[k] = HFq::neg(conj[k]);
    f12 a = f12_mul(conj, f12_inv(f));
    // q^2 + 1
    const FrobConsts& C = frob();
    f12 b;
    for (int k = 0; k < 12; k++) b[k] = HFq::mul(a[k], C.zeta[k]);
    a = f12_mul(b, a);
    return f12_pow_window(a, hard_exponent());
}

}  // namespace g16v
