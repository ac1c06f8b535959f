// optimal ate pairing on BN254 in the plain Fq12 model (see verify.hpp)
#pragma once
#include "verify.hpp"

namespace g16v {

struct P12 {  // affine point over Fq12; inf = point at infinity
    f12 x, y;
    bool inf = false;
};

// G2 affine (x, y) over Fq2 -> Fq12: (x w^2, y w^3), with an Fq2 element c0 + c1 u written as (c0 - 9 c1) + c1 w^6
inline P12 twist(const HAffine<HFq2>& q) {
    static const fq nine = fq_small(9);
    P12 r;
    r.x = f12_zero();
    r.y = f12_zero();
    r.x[2] = HFq::sub(q.x.c0, HFq::mul(nine, q.x.c1));
    r.x[8] = q.x.c1;
    r.y[3] = HFq::sub(q.y.c0, HFq::mul(nine, q.y.c1));
    r.y[9] = q.y.c1;
    return r;
}

inline P12 cast_g1(const HAffine<HFq>& p) {
    P12 r;
    r.x = f12_zero();
    r.y = f12_zero();
    r.x[0] = p.x;
    r.y[0] = p.y;
    return r;
}

inline P12 ec12_double(const P12& p) {
    if (p.inf) return p;
    static const fq two = fq_small(2), three = fq_small(3);
    f12 m = f12_mul(f12_muls(f12_mul(p.x, p.x), three), f12_inv(f12_muls(p.y, two)));
    P12 r;
    r.x = f12_sub(f12_mul(m, m), f12_muls(p.x, two));
    r.y = f12_sub(f12_mul(m, f12_sub(p.x, r.x)), p.y);
    return r;
}

inline P12 ec12_add(const P12& p, const P12& q) {
    if (p.inf) return q;
    if (q.inf) return p;
    if (f12_eq(p.x, q.x)) {
        if (f12_eq(p.y, q.y)) return ec12_double(p);
        P12 r;
        r.inf = true;
        return r;
    }
    f12 m = f12_mul(f12_sub(q.y, p.y), f12_inv(f12_sub(q.x, p.x)));
    P12 r;
    r.x = f12_sub(f12_sub(f12_mul(m, m), p.x), q.x);
    r.y = f12_sub(f12_mul(m, f12_sub(p.x, r.x)), p.y);
    return r;
}

// the line through p1 and p2 (tangent if equal) evaluated at t
inline f12 linefunc(const P12& p1, const P12& p2, const P12& t) {
    static const fq two = fq_small(2), three = fq_small(3);
    if (!f12_eq(p1.x, p2.x)) {
        f12 m = f12_mul(f12_sub(p2.y, p1.y), f12_inv(f12_sub(p2.x, p1.x)));
        return f12_sub(f12_mul(m, f12_sub(t.x, p1.x)), f12_sub(t.y, p1.y));
    }
    if (f12_eq(p1.y, p2.y)) {
        f12 m = f12_mul(f12_muls(f12_mul(p1.x, p1.x), three), f12_inv(f12_muls(p1.y, two)));
        return f12_sub(f12_mul(m, f12_sub(t.x, p1.x)), f12_sub(t.y, p1.y));
    }
    return f12_sub(t.x, p1.x);
}

inline std::vector<uint64_t> words_of(const U256& a) { return {a.v[0], a.v[1], a.v[2], a.v[3]}; }

// unreduced pairing value of (Q in G2, P in G1); 1 if either is infinity
inline f12 miller_loop(const HAffine<HFq2>& q2, bool q_inf, const HAffine<HFq>& p1, bool p_inf) {
    if (q_inf || p_inf) return f12_one();
    // 6t + 2, t = 4965661367192848881
    const uint64_t ATE_LO = 0x9d797039be763ba8ULL;  // 29793968203157093288 = 0x1_9d797039be763ba8: bit 64 is the leading one
    P12 Qt = twist(q2), Pt = cast_g1(p1), Rt = Qt;
    f12 f = f12_one();
    auto bit = [&](int i) -> bool { return ((ATE_LO >> i) & 1) != 0; };
    for (int i = 63; i >= 0; i--) {
        f = f12_mul(f12_mul(f, f), linefunc(Rt, Rt, Pt));
        Rt = ec12_double(Rt);
        if (bit(i)) {
            f = f12_mul(f, linefunc(Rt, Qt, Pt));
            Rt = ec12_add(Rt, Qt);
        }
    }
    std::vector<uint64_t> qw = words_of(HFq::mod());
    P12 Q1, nQ2;
    Q1.x = f12_pow(Qt.x, qw);
    Q1.y = f12_pow(Qt.y, qw);
    nQ2.x = f12_pow(Q1.x, qw);
    nQ2.y = f12_neg(f12_pow(Q1.y, qw));
    f = f12_mul(f, linefunc(Rt, Q1, Pt));
    Rt = ec12_add(Rt, Q1);
    f = f12_mul(f, linefunc(Rt, nQ2, Pt));
    return f;
}

// (q^12 - 1) / r as little-endian 64-bit words, computed once with schoolbook big integers
inline const std::vector<uint64_t>& final_exponent() {
    static const std::vector<uint64_t> e = [] {
        auto mulw = [](const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
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
        };
        std::vector<uint64_t> q = words_of(HFq::mod()), r = words_of(HFr::mod()), p = {1};
        for (int i = 0; i < 12; i++) p = mulw(p, q);
        // p -= 1 (q^12 is odd, so no borrow beyond word 0)
        p[0] -= 1;
        // long division of p by r, bit by bit (done once)
        std::vector<uint64_t> quo(p.size(), 0), rem(r.size() + 1, 0);
        for (int i = int(p.size()) * 64 - 1; i >= 0; i--) {
            // rem = rem * 2 + bit
            uint64_t carry = (p[i >> 6] >> (i & 63)) & 1;
            for (auto& w : rem) {
                uint64_t nc = w >> 63;
                w = (w << 1) | carry;
                carry = nc;
            }
            // if rem >= r: rem -= r, set quotient bit
            bool ge = rem[r.size()] != 0;
            if (!ge) {
                ge = true;
                for (int k = int(r.size()) - 1; k >= 0; k--) {
                    if (rem[k] != r[k]) {
                        ge = rem[k] > r[k];
                        break;
                    }
                }
            }
            if (ge) {
                unsigned __int128 b = 0;
                for (size_t k = 0; k < rem.size(); k++) {
                    unsigned __int128 d = (unsigned __int128)rem[k] - (k < r.size() ? r[k] : 0) - (uint64_t)b;
                    rem[k] = (uint64_t)d;
                    b = (d >> 64) & 1;
                }
                quo[i >> 6] |= uint64_t(1) << (i & 63);
            }
        }
        return quo;
    }();
    return e;
}

inline f12 final_exponentiate(const f12& f) { return f12_pow(f, final_exponent()); }

}  // namespace g16v
