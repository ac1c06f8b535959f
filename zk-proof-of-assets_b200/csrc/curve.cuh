// BN254 G1 (over Fq) and G2 (over Fq2) group law in XYZZ coordinates (x = X/ZZ, y = Y/ZZZ, a = 0).
//
// Replaces: wasmcurves 0.2.2 build_curve_jacobian_a0 / rapidsnark curve.hpp.
// XYZZ instead of Jacobian because the bucket accumulation is one long run of mixed additions
// (8M + 2S each) and never needs the Z coordinate on its own.  Identity: ZZ == 0.
#pragma once
#include "field.cuh"

namespace g16 {

template <class F>
struct Affine {
    typename F::el x, y;  // (0, 0) encodes infinity (zkey convention)
};

template <class F>
struct XYZZ {
    typename F::el x, y, zz, zzz;
};

template <class F>
struct Curve {
    typedef F field;
    typedef typename F::el fe;
    typedef Affine<F> aff;
    typedef XYZZ<F> pt;

    static __device__ __forceinline__ pt identity() { return pt{F::zero(), F::zero(), F::zero(), F::zero()}; }
    static __device__ __forceinline__ bool is_identity(const pt& p) { return F::is_zero(p.zz); }
    static __device__ __forceinline__ bool is_inf(const aff& a) { return F::is_zero(a.x) && F::is_zero(a.y); }
    static __device__ __forceinline__ pt from_affine(const aff& a) {
        if (is_inf(a)) return identity();
        return pt{a.x, a.y, F::one(), F::one()};
    }

    // 2*(x, y)  (dbl-2008-s-1 with ZZ = ZZZ = 1)
    static __device__ __forceinline__ pt dbl_affine(const aff& a) {
        fe u = F::dbl(a.y);
        fe v = F::sqr(u);
        fe w = F::mul(u, v);
        fe s = F::mul(a.x, v);
        fe xx = F::sqr(a.x);
        fe m = F::add(F::dbl(xx), xx);
        fe x3 = F::sub(F::sqr(m), F::dbl(s));
        fe y3 = F::sub(F::mul(m, F::sub(s, x3)), F::mul(w, a.y));
        return pt{x3, y3, v, w};
    }

    static __device__ __forceinline__ pt dbl(const pt& p) {
        if (is_identity(p)) return p;
        fe u = F::dbl(p.y);
        fe v = F::sqr(u);
        fe w = F::mul(u, v);
        fe s = F::mul(p.x, v);
        fe xx = F::sqr(p.x);
        fe m = F::add(F::dbl(xx), xx);
        fe x3 = F::sub(F::sqr(m), F::dbl(s));
        fe y3 = F::sub(F::mul(m, F::sub(s, x3)), F::mul(w, p.y));
        return pt{x3, y3, F::mul(v, p.zz), F::mul(w, p.zzz)};
    }

    // acc += (x2, +-y2)  (madd-2008-s); the affine operand must not be infinity
    static __device__ __forceinline__ void madd(pt& acc, const aff& q) {
        if (is_identity(acc)) {
            acc = pt{q.x, q.y, F::one(), F::one()};
            return;
        }
        fe u2 = F::mul(q.x, acc.zz);
        fe s2 = F::mul(q.y, acc.zzz);
        fe p = F::sub(u2, acc.x);
        fe r = F::sub(s2, acc.y);
        if (F::is_zero(p)) {
            if (F::is_zero(r)) acc = dbl_affine(q);
            else acc = identity();
            return;
        }
        fe pp = F::sqr(p);
        fe ppp = F::mul(p, pp);
        fe qq = F::mul(acc.x, pp);
        fe x3 = F::sub(F::sub(F::sqr(r), ppp), F::dbl(qq));
        fe y3 = F::sub(F::mul(r, F::sub(qq, x3)), F::mul(acc.y, ppp));
        acc.x = x3;
        acc.y = y3;
        acc.zz = F::mul(acc.zz, pp);
        acc.zzz = F::mul(acc.zzz, ppp);
    }

    // acc += b (add-2008-s), all cases
    static __device__ __forceinline__ void add(pt& acc, const pt& b) {
        if (is_identity(b)) return;
        if (is_identity(acc)) {
            acc = b;
            return;
        }
        fe u1 = F::mul(acc.x, b.zz);
        fe u2 = F::mul(b.x, acc.zz);
        fe s1 = F::mul(acc.y, b.zzz);
        fe s2 = F::mul(b.y, acc.zzz);
        fe p = F::sub(u2, u1);
        fe r = F::sub(s2, s1);
        if (F::is_zero(p)) {
            if (F::is_zero(r)) acc = dbl(acc);
            else acc = identity();
            return;
        }
        fe pp = F::sqr(p);
        fe ppp = F::mul(p, pp);
        fe qq = F::mul(u1, pp);
        fe x3 = F::sub(F::sub(F::sqr(r), ppp), F::dbl(qq));
        fe y3 = F::sub(F::mul(r, F::sub(qq, x3)), F::mul(s1, ppp));
        acc.x = x3;
        acc.y = y3;
        acc.zz = F::mul(F::mul(acc.zz, b.zz), pp);
        acc.zzz = F::mul(F::mul(acc.zzz, b.zzz), ppp);
    }

    static __device__ __forceinline__ pt neg(const pt& p) { return pt{p.x, F::neg(p.y), p.zz, p.zzz}; }

    // k * p for a small unsigned k (bucket-chunk weights), double-and-add from the top bit
    static __device__ pt mul_small(const pt& p, uint32_t k) {
        pt r = identity();
        for (int i = 31 - __clz(k | 1); i >= 0; i--) {
            r = dbl(r);
            if ((k >> i) & 1) add(r, p);
        }
        if (k == 0) return identity();
        return r;
    }
};

typedef Curve<Fq> G1;
typedef Curve<Fq2> G2;
typedef Affine<Fq> g1_aff;   // 64 bytes, zkey section 5/6/8/9 record
typedef Affine<Fq2> g2_aff;  // 128 bytes, zkey section 7 record (x.c0 x.c1 y.c0 y.c1)
typedef XYZZ<Fq> g1_pt;      // 128 bytes
typedef XYZZ<Fq2> g2_pt;     // 256 bytes

}  // namespace g16
