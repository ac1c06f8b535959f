// Host-side BN254 arithmetic (4 x 64-bit Montgomery, unsigned __int128) for the O(1) parts of the prover:
// the final combination of the five MSM results with alpha/beta/delta and the blinding factors r, s
// (SURVEY.md row a15), XYZZ -> affine, decimal output (row a16) and the NTT root tables' seeds.
// Not a CPU fallback for the hot path: nothing here touches a vector of more than a few elements.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>

namespace g16h {

typedef unsigned __int128 u128;

struct U256 {
    uint64_t v[4];
};

inline bool u256_is_zero(const U256& a) { return (a.v[0] | a.v[1] | a.v[2] | a.v[3]) == 0; }
inline bool u256_eq(const U256& a, const U256& b) {
    return a.v[0] == b.v[0] && a.v[1] == b.v[1] && a.v[2] == b.v[2] && a.v[3] == b.v[3];
}
inline int u256_cmp(const U256& a, const U256& b) {
    for (int i = 3; i >= 0; i--) {
        if (a.v[i] < b.v[i]) return -1;
        if (a.v[i] > b.v[i]) return 1;
    }
    return 0;
}
inline U256 u256_from_bytes(const uint8_t* p) {
    U256 r;
    std::memcpy(r.v, p, 32);
    return r;
}
inline void u256_to_bytes(const U256& a, uint8_t* p) { std::memcpy(p, a.v, 32); }

// decimal string of a 256-bit integer
inline std::string u256_to_dec(U256 a) {
    if (u256_is_zero(a)) return "0";
    char buf[80];
    int pos = 80;
    while (!u256_is_zero(a)) {
        // divide by 10^18
        uint64_t rem = 0;
        const uint64_t D = 1000000000000000000ull;
        for (int i = 3; i >= 0; i--) {
            u128 cur = ((u128)rem << 64) | a.v[i];
            a.v[i] = (uint64_t)(cur / D);
            rem = (uint64_t)(cur % D);
        }
        bool last = u256_is_zero(a);
        for (int d = 0; d < 18; d++) {
            if (last && rem == 0) break;
            buf[--pos] = char('0' + rem % 10);
            rem /= 10;
        }
    }
    return std::string(buf + pos, 80 - pos);
}

struct FqP {
    static constexpr uint64_t MOD[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull,
                                        0x30644e72e131a029ull};
    static constexpr uint64_t R1[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull,
                                       0x0e0a77c19a07df2full};
    static constexpr uint64_t R2[4] = {0xf32cfc5b538afa89ull, 0xb5e71911d44501fbull, 0x47ab1eff0a417ff6ull,
                                       0x06d89f71cab8351full};
    static constexpr uint64_t INV = 0x87d20782e4866389ull;
};
struct FrP {
    static constexpr uint64_t MOD[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull,
                                        0x30644e72e131a029ull};
    static constexpr uint64_t R1[4] = {0xac96341c4ffffffbull, 0x36fc76959f60cd29ull, 0x666ea36f7879462eull,
                                       0x0e0a77c19a07df2full};
    static constexpr uint64_t R2[4] = {0x1bb8e645ae216da7ull, 0x53fe3ab1e35c59e3ull, 0x8c49833d53bb8085ull,
                                       0x0216d0b17f4e44a5ull};
    static constexpr uint64_t INV = 0xc2e1f593efffffffull;
};

template <class P>
struct HField {
    typedef U256 el;
    static el mod() { return el{{P::MOD[0], P::MOD[1], P::MOD[2], P::MOD[3]}}; }
    static el zero() { return el{{0, 0, 0, 0}}; }
    static el one() { return el{{P::R1[0], P::R1[1], P::R1[2], P::R1[3]}}; }
    static el r2() { return el{{P::R2[0], P::R2[1], P::R2[2], P::R2[3]}}; }
    static bool is_zero(const el& a) { return u256_is_zero(a); }
    static bool eq(const el& a, const el& b) { return u256_eq(a, b); }

    static el add(const el& a, const el& b) {
        el r;
        u128 c = 0;
        for (int i = 0; i < 4; i++) {
            c += (u128)a.v[i] + b.v[i];
            r.v[i] = (uint64_t)c;
            c >>= 64;
        }
        el m = mod();
        if (u256_cmp(r, m) >= 0) r = raw_sub(r, m);
        return r;
    }
    static el raw_sub(const el& a, const el& b) {
        el r;
        uint64_t borrow = 0;
        for (int i = 0; i < 4; i++) {
            u128 d = (u128)a.v[i] - b.v[i] - borrow;
            r.v[i] = (uint64_t)d;
            borrow = (uint64_t)(d >> 64) & 1;
        }
        return r;
    }
    static el sub(const el& a, const el& b) {
        if (u256_cmp(a, b) >= 0) return raw_sub(a, b);
        return raw_sub(add_nored(a, mod()), b);
    }
    static el add_nored(const el& a, const el& b) {
        el r;
        u128 c = 0;
        for (int i = 0; i < 4; i++) {
            c += (u128)a.v[i] + b.v[i];
            r.v[i] = (uint64_t)c;
            c >>= 64;
        }
        return r;
    }
    static el neg(const el& a) { return is_zero(a) ? a : raw_sub(mod(), a); }
    static el dbl(const el& a) { return add(a, a); }

    // CIOS Montgomery product; b may be any 256-bit value
    static el mul(const el& a, const el& b) {
        uint64_t t[6] = {0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 4; i++) {
            u128 c = 0;
            for (int j = 0; j < 4; j++) {
                c += (u128)a.v[j] * b.v[i] + t[j];
                t[j] = (uint64_t)c;
                c >>= 64;
            }
            c += t[4];
            t[4] = (uint64_t)c;
            t[5] = (uint64_t)(c >> 64);
            uint64_t m = t[0] * P::INV;
            c = (u128)m * P::MOD[0] + t[0];
            c >>= 64;
            for (int j = 1; j < 4; j++) {
                c += (u128)m * P::MOD[j] + t[j];
                t[j - 1] = (uint64_t)c;
                c >>= 64;
            }
            c += t[4];
            t[3] = (uint64_t)c;
            t[4] = t[5] + (uint64_t)(c >> 64);
        }
        el r{{t[0], t[1], t[2], t[3]}};
        if (t[4] || u256_cmp(r, mod()) >= 0) r = raw_sub(r, mod());
        return r;
    }
    static el sqr(const el& a) { return mul(a, a); }
    static el to_mont(const el& a) { return mul(r2(), a); }
    static el from_mont(const el& a) { return mul(a, el{{1, 0, 0, 0}}); }
    static el pow(const el& a, const el& e) {
        el r = one();
        for (int i = 255; i >= 0; i--) {
            r = sqr(r);
            if ((e.v[i >> 6] >> (i & 63)) & 1) r = mul(r, a);
        }
        return r;
    }
    static el inv(const el& a) {
        el e = mod();
        e.v[0] -= 2;
        return pow(a, e);
    }
    static el from_u64(uint64_t x) { return to_mont(el{{x, 0, 0, 0}}); }
};

typedef HField<FqP> HFq;
typedef HField<FrP> HFr;

struct HFq2El {
    U256 c0, c1;
};
struct HFq2 {
    typedef HFq2El el;
    static el zero() { return el{HFq::zero(), HFq::zero()}; }
    static el one() { return el{HFq::one(), HFq::zero()}; }
    static bool is_zero(const el& a) { return HFq::is_zero(a.c0) && HFq::is_zero(a.c1); }
    static bool eq(const el& a, const el& b) { return HFq::eq(a.c0, b.c0) && HFq::eq(a.c1, b.c1); }
    static el add(const el& a, const el& b) { return el{HFq::add(a.c0, b.c0), HFq::add(a.c1, b.c1)}; }
    static el sub(const el& a, const el& b) { return el{HFq::sub(a.c0, b.c0), HFq::sub(a.c1, b.c1)}; }
    static el dbl(const el& a) { return add(a, a); }
    static el neg(const el& a) { return el{HFq::neg(a.c0), HFq::neg(a.c1)}; }
    static el mul(const el& a, const el& b) {
        U256 t0 = HFq::mul(a.c0, b.c0), t1 = HFq::mul(a.c1, b.c1);
        U256 s = HFq::mul(HFq::add(a.c0, a.c1), HFq::add(b.c0, b.c1));
        return el{HFq::sub(t0, t1), HFq::sub(HFq::sub(s, t0), t1)};
    }
    static el sqr(const el& a) { return mul(a, a); }
    static el inv(const el& a) {
        U256 d = HFq::inv(HFq::add(HFq::sqr(a.c0), HFq::sqr(a.c1)));
        return el{HFq::mul(a.c0, d), HFq::neg(HFq::mul(a.c1, d))};
    }
};

// XYZZ points on the host; same formulas as csrc/curve.cuh
template <class F>
struct HPoint {
    typename F::el x, y, zz, zzz;
};
template <class F>
struct HAffine {
    typename F::el x, y;
    bool inf;
};

template <class F>
struct HCurve {
    typedef typename F::el fe;
    typedef HPoint<F> pt;
    typedef HAffine<F> aff;
    static pt identity() { return pt{F::zero(), F::zero(), F::zero(), F::zero()}; }
    static bool is_identity(const pt& p) { return F::is_zero(p.zz); }
    static pt from_affine(const aff& a) {
        if (a.inf) return identity();
        return pt{a.x, a.y, F::one(), F::one()};
    }
    static pt dbl(const pt& p) {
        if (is_identity(p)) return p;
        if (F::is_zero(p.y)) return identity();
        fe u = F::dbl(p.y), v = F::sqr(u), w = F::mul(u, v), s = F::mul(p.x, v);
        fe xx = F::sqr(p.x), m = F::add(F::dbl(xx), xx);
        fe x3 = F::sub(F::sqr(m), F::dbl(s));
        fe y3 = F::sub(F::mul(m, F::sub(s, x3)), F::mul(w, p.y));
        return pt{x3, y3, F::mul(v, p.zz), F::mul(w, p.zzz)};
    }
    static pt add(const pt& a, const pt& b) {
        if (is_identity(a)) return b;
        if (is_identity(b)) return a;
        fe u1 = F::mul(a.x, b.zz), u2 = F::mul(b.x, a.zz);
        fe s1 = F::mul(a.y, b.zzz), s2 = F::mul(b.y, a.zzz);
        fe p = F::sub(u2, u1), r = F::sub(s2, s1);
        if (F::is_zero(p)) return F::is_zero(r) ? dbl(a) : identity();
        fe pp = F::sqr(p), ppp = F::mul(p, pp), q = F::mul(u1, pp);
        fe x3 = F::sub(F::sub(F::sqr(r), ppp), F::dbl(q));
        fe y3 = F::sub(F::mul(r, F::sub(q, x3)), F::mul(s1, ppp));
        return pt{x3, y3, F::mul(F::mul(a.zz, b.zz), pp), F::mul(F::mul(a.zzz, b.zzz), ppp)};
    }
    static pt neg(const pt& p) { return pt{p.x, F::neg(p.y), p.zz, p.zzz}; }
    // k * p, k a standard-form 256-bit scalar
    static pt mul(const pt& p, const U256& k) {
        pt r = identity();
        for (int i = 255; i >= 0; i--) {
            r = dbl(r);
            if ((k.v[i >> 6] >> (i & 63)) & 1) r = add(r, p);
        }
        return r;
    }
    static aff to_affine(const pt& p) {
        aff a;
        if (is_identity(p)) {
            a.x = F::zero();
            a.y = F::zero();
            a.inf = true;
            return a;
        }
        a.inf = false;
        a.x = F::mul(p.x, F::inv(p.zz));
        a.y = F::mul(p.y, F::inv(p.zzz));
        return a;
    }
};

typedef HCurve<HFq> HG1;
typedef HCurve<HFq2> HG2;

}  // namespace g16h
