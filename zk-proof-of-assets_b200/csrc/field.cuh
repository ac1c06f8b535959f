// BN254 prime fields on the sm_100a integer pipe: 8 x 32-bit limbs, Montgomery form (R = 2^256).
//
// Replaces: wasmcurves 0.2.2 build_f1m / rapidsnark fr.asm+fq.asm (4x64 Montgomery).
// Call sites in the reference: scripts/g16_prove.sh:246-260.  Byte layout of an element is the
// zkey/wtns one: 32 bytes little-endian, so device buffers are the file bytes.
//
// mul(): operand-scanning Montgomery product with the reduction interleaved.  Partial products of the
// even limbs of `a` and of the odd limbs of `a` are kept in two accumulators one limb apart, so every
// 32x32->64 product lands on an aligned register pair and a row is four carry chains of
// mad.lo.cc/madc.hi.cc (ptxas pairs them into IMAD.WIDE.U32 + carry).  128 wide MADs + 8 single
// MULs per product = the 136-op work model of SURVEY.md section 8(d).  The design was checked limb for
// limb by tools/montsim.py before it was written here.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace g16 {

struct alignas(16) fe_t {
    uint32_t v[8];
};

struct FqParams {
    static __host__ __device__ __forceinline__ constexpr uint32_t mod(int i) {
        const uint32_t t[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return t[i];
    }
    static __host__ __device__ __forceinline__ constexpr uint32_t r1(int i) {
        const uint32_t t[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return t[i];
    }
    static __host__ __device__ __forceinline__ constexpr uint32_t r2(int i) {
        const uint32_t t[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u, 0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return t[i];
    }
    static constexpr uint32_t INV = 0xe4866389u;  // -MOD^-1 mod 2^32
};

struct FrParams {
    static __host__ __device__ __forceinline__ constexpr uint32_t mod(int i) {
        const uint32_t t[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return t[i];
    }
    static __host__ __device__ __forceinline__ constexpr uint32_t r1(int i) {
        const uint32_t t[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u, 0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return t[i];
    }
    static __host__ __device__ __forceinline__ constexpr uint32_t r2(int i) {
        const uint32_t t[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u, 0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return t[i];
    }
    static constexpr uint32_t INV = 0xefffffffu;
};

// ---- carry-chain building blocks (one asm statement per chain: the carry flag never crosses statements)

// acc[0..7] += {x0,x1,x2,x3} * y, product k on limbs (2k, 2k+1); returns the carry out of limb 7
__device__ __forceinline__ uint32_t mad_row_c(uint32_t (&acc)[8], uint32_t x0, uint32_t x1, uint32_t x2,
                                              uint32_t x3, uint32_t y) {
    uint32_t c;
    asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32 %8, 0, 0;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]),
          "+r"(acc[7]), "=r"(c)
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
    return c;
}

// same, when the bound on the running value proves there is no carry out
__device__ __forceinline__ void mad_row(uint32_t (&acc)[8], uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3,
                                        uint32_t y) {
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
        "madc.hi.u32 %7, %11, %12, %7;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]),
          "+r"(acc[7])
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
}

// out[0..7] = {x0..x3} * y (aligned pairs, no carries)
__device__ __forceinline__ void mul_row(uint32_t (&out)[8], uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3,
                                        uint32_t y) {
    asm("mul.lo.u32 %0, %8, %12;\n\t"
        "mul.hi.u32 %1, %8, %12;\n\t"
        "mul.lo.u32 %2, %9, %12;\n\t"
        "mul.hi.u32 %3, %9, %12;\n\t"
        "mul.lo.u32 %4, %10, %12;\n\t"
        "mul.hi.u32 %5, %10, %12;\n\t"
        "mul.lo.u32 %6, %11, %12;\n\t"
        "mul.hi.u32 %7, %11, %12;"
        : "=r"(out[0]), "=r"(out[1]), "=r"(out[2]), "=r"(out[3]), "=r"(out[4]), "=r"(out[5]), "=r"(out[6]),
          "=r"(out[7])
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
}

// The one-limb shift between rows, fused with the odd-limb products of the next row:
//   lo0 += sh[1]                                  (carry goes on into limb 1 = out[0])
//   out[j] = {x0..x3}*y on pairs + sh[j+2] + carry   (sh[8] = sh[9] = 0)
__device__ __forceinline__ void mad_row_shift(uint32_t (&out)[8], uint32_t& lo0, const uint32_t (&sh)[8], uint32_t x0,
                                              uint32_t x1, uint32_t x2, uint32_t x3, uint32_t y) {
    asm("add.cc.u32 %8, %8, %9;\n\t"
        "madc.lo.cc.u32 %0, %16, %20, %10;\n\t"
        "madc.hi.cc.u32 %1, %16, %20, %11;\n\t"
        "madc.lo.cc.u32 %2, %17, %20, %12;\n\t"
        "madc.hi.cc.u32 %3, %17, %20, %13;\n\t"
        "madc.lo.cc.u32 %4, %18, %20, %14;\n\t"
        "madc.hi.cc.u32 %5, %18, %20, %15;\n\t"
        "madc.lo.cc.u32 %6, %19, %20, 0;\n\t"
        "madc.hi.u32 %7, %19, %20, 0;"
        : "=&r"(out[0]), "=&r"(out[1]), "=&r"(out[2]), "=&r"(out[3]), "=&r"(out[4]), "=&r"(out[5]), "=&r"(out[6]),
          "=&r"(out[7]), "+r"(lo0)
        : "r"(sh[1]), "r"(sh[2]), "r"(sh[3]), "r"(sh[4]), "r"(sh[5]), "r"(sh[6]), "r"(sh[7]), "r"(x0), "r"(x1),
          "r"(x2), "r"(x3), "r"(y));
}


// ---- building blocks of the dedicated squaring ------------------------------------------------------------------
// acc pairs += {x...} * y on aligned register pairs, carry out returned; 1, 2 and 3 products (4 = mad_row_c)
__device__ __forceinline__ uint32_t mad_pairs1_c(uint32_t& a0, uint32_t& a1, uint32_t x0, uint32_t y) {
    uint32_t c;
    asm("mad.lo.cc.u32 %0, %3, %4, %0;\n\t"
        "madc.hi.cc.u32 %1, %3, %4, %1;\n\t"
        "addc.u32 %2, 0, 0;"
        : "+r"(a0), "+r"(a1), "=r"(c)
        : "r"(x0), "r"(y));
    return c;
}
__device__ __forceinline__ uint32_t mad_pairs2_c(uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3, uint32_t x0, uint32_t x1,
                                                 uint32_t y) {
    uint32_t c;
    asm("mad.lo.cc.u32 %0, %5, %7, %0;\n\t"
        "madc.hi.cc.u32 %1, %5, %7, %1;\n\t"
        "madc.lo.cc.u32 %2, %6, %7, %2;\n\t"
        "madc.hi.cc.u32 %3, %6, %7, %3;\n\t"
        "addc.u32 %4, 0, 0;"
        : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3), "=r"(c)
        : "r"(x0), "r"(x1), "r"(y));
    return c;
}
__device__ __forceinline__ uint32_t mad_pairs3_c(uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3, uint32_t& a4, uint32_t& a5,
                                                 uint32_t x0, uint32_t x1, uint32_t x2, uint32_t y) {
    uint32_t c;
    asm("mad.lo.cc.u32 %0, %7, %10, %0;\n\t"
        "madc.hi.cc.u32 %1, %7, %10, %1;\n\t"
        "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
        "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
        "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
        "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
        "addc.u32 %6, 0, 0;"
        : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3), "+r"(a4), "+r"(a5), "=r"(c)
        : "r"(x0), "r"(x1), "r"(x2), "r"(y));
    return c;
}
__device__ __forceinline__ uint32_t mad_pairs4_c(uint32_t& a0, uint32_t& a1, uint32_t& a2, uint32_t& a3, uint32_t& a4, uint32_t& a5,
                                                 uint32_t& a6, uint32_t& a7, uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3,
                                                 uint32_t y) {
    uint32_t c;
    asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32 %8, 0, 0;"
        : "+r"(a0), "+r"(a1), "+r"(a2), "+r"(a3), "+r"(a4), "+r"(a5), "+r"(a6), "+r"(a7), "=r"(c)
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
    return c;
}

template <class P>
struct Field {
    typedef fe_t el;

    static __device__ __forceinline__ el zero() {
        el r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = 0;
        return r;
    }
    static __device__ __forceinline__ el one() {  // Montgomery form of 1
        el r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = P::r1(i);
        return r;
    }
    static __device__ __forceinline__ el r2() {
        el r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = P::r2(i);
        return r;
    }
    static __device__ __forceinline__ bool is_zero(const el& a) {
        uint32_t o = a.v[0] | a.v[1] | a.v[2] | a.v[3] | a.v[4] | a.v[5] | a.v[6] | a.v[7];
        return o == 0;
    }
    static __device__ __forceinline__ bool eq(const el& a, const el& b) {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o |= a.v[i] ^ b.v[i];
        return o == 0;
    }

    // r = a - MOD if a >= MOD (a < 2*MOD)
    static __device__ __forceinline__ void final_sub(uint32_t (&r)[8]) {
        uint32_t t[8], borrow;
        asm("sub.cc.u32 %0, %9, %17;\n\t"
            "subc.cc.u32 %1, %10, %18;\n\t"
            "subc.cc.u32 %2, %11, %19;\n\t"
            "subc.cc.u32 %3, %12, %20;\n\t"
            "subc.cc.u32 %4, %13, %21;\n\t"
            "subc.cc.u32 %5, %14, %22;\n\t"
            "subc.cc.u32 %6, %15, %23;\n\t"
            "subc.cc.u32 %7, %16, %24;\n\t"
            "subc.u32 %8, 0, 0;"
            : "=r"(t[0]), "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7]),
              "=r"(borrow)
            : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
              "r"(P::mod(0)), "r"(P::mod(1)), "r"(P::mod(2)), "r"(P::mod(3)), "r"(P::mod(4)), "r"(P::mod(5)),
              "r"(P::mod(6)), "r"(P::mod(7)));
        if (borrow == 0) {
#pragma unroll
            for (int i = 0; i < 8; i++) r[i] = t[i];
        }
    }

    static __device__ __forceinline__ el add(const el& a, const el& b) {
        uint32_t r[8];
        asm("add.cc.u32 %0, %8, %16;\n\t"
            "addc.cc.u32 %1, %9, %17;\n\t"
            "addc.cc.u32 %2, %10, %18;\n\t"
            "addc.cc.u32 %3, %11, %19;\n\t"
            "addc.cc.u32 %4, %12, %20;\n\t"
            "addc.cc.u32 %5, %13, %21;\n\t"
            "addc.cc.u32 %6, %14, %22;\n\t"
            "addc.u32 %7, %15, %23;"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
            : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
              "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
        final_sub(r);  // a + b < 2*MOD < 2^255: no carry out of limb 7
        el o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = r[i];
        return o;
    }

    static __device__ __forceinline__ el dbl(const el& a) { return add(a, a); }

    static __device__ __forceinline__ el sub(const el& a, const el& b) {
        uint32_t r[8], borrow;
        asm("sub.cc.u32 %0, %9, %17;\n\t"
            "subc.cc.u32 %1, %10, %18;\n\t"
            "subc.cc.u32 %2, %11, %19;\n\t"
            "subc.cc.u32 %3, %12, %20;\n\t"
            "subc.cc.u32 %4, %13, %21;\n\t"
            "subc.cc.u32 %5, %14, %22;\n\t"
            "subc.cc.u32 %6, %15, %23;\n\t"
            "subc.cc.u32 %7, %16, %24;\n\t"
            "subc.u32 %8, 0, 0;"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(borrow)
            : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
              "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
        // borrow is 0 or 0xffffffff: add MOD & borrow
        asm("add.cc.u32 %0, %0, %8;\n\t"
            "addc.cc.u32 %1, %1, %9;\n\t"
            "addc.cc.u32 %2, %2, %10;\n\t"
            "addc.cc.u32 %3, %3, %11;\n\t"
            "addc.cc.u32 %4, %4, %12;\n\t"
            "addc.cc.u32 %5, %5, %13;\n\t"
            "addc.cc.u32 %6, %6, %14;\n\t"
            "addc.u32 %7, %7, %15;"
            : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
            : "r"(P::mod(0) & borrow), "r"(P::mod(1) & borrow), "r"(P::mod(2) & borrow), "r"(P::mod(3) & borrow),
              "r"(P::mod(4) & borrow), "r"(P::mod(5) & borrow), "r"(P::mod(6) & borrow), "r"(P::mod(7) & borrow));
        el o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = r[i];
        return o;
    }

    static __device__ __forceinline__ el neg(const el& a) {
        if (is_zero(a)) return a;
        el m;
#pragma unroll
        for (int i = 0; i < 8; i++) m.v[i] = P::mod(i);
        return sub(m, a);
    }

    // Montgomery product a*b/R mod MOD; a < MOD, b < 2^256 (so raw witness bytes are a legal `b`)
    static __device__ __forceinline__ el mul(const el& a, const el& b) {
        uint32_t E[8], O[8], T[8];
        // row 0
        mul_row(E, a.v[0], a.v[2], a.v[4], a.v[6], b.v[0]);
        mul_row(O, a.v[1], a.v[3], a.v[5], a.v[7], b.v[0]);
        uint32_t m = E[0] * P::INV;
        mad_row(O, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
        O[7] += mad_row_c(E, P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
#pragma unroll
        for (int i = 1; i < 8; i++) {
            // frame moves up one limb: new even accumulator = old odd one; old even one is shifted into T
            mad_row_shift(T, O[0], E, a.v[1], a.v[3], a.v[5], a.v[7], b.v[i]);
            T[7] += mad_row_c(O, a.v[0], a.v[2], a.v[4], a.v[6], b.v[i]);
            m = O[0] * P::INV;
            mad_row(T, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            T[7] += mad_row_c(O, P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                E[k] = O[k];
                O[k] = T[k];
            }
        }
        // value = O + (E >> 32)   (E[0] == 0)
        uint32_t r[8];
        asm("add.cc.u32 %0, %8, %16;\n\t"
            "addc.cc.u32 %1, %9, %17;\n\t"
            "addc.cc.u32 %2, %10, %18;\n\t"
            "addc.cc.u32 %3, %11, %19;\n\t"
            "addc.cc.u32 %4, %12, %20;\n\t"
            "addc.cc.u32 %5, %13, %21;\n\t"
            "addc.cc.u32 %6, %14, %22;\n\t"
            "addc.u32 %7, %15, 0;"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
            : "r"(O[0]), "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]), "r"(E[1]),
              "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]));
        final_sub(r);
        el o;
#pragma unroll
        for (int k = 0; k < 8; k++) o.v[k] = r[k];
        return o;
    }

    // The squaring every kernel uses is the general product: the dedicated form below has 28 fewer wide multiply-adds
    // but was measured SLOWER inside the kernels that matter (H accumulation 4.61 -> 4.75 ms at 2^21, 33.7 -> 34.7 ms at
    // 2^24; Poseidon tree unchanged): its separated reduction and the doubling add ~130 carry-chain ALU instructions of
    // dependent latency and 40 more bytes of spills at 128 registers, and those kernels are limited by dependency
    // stalls as much as by the pipe (ncu: stalled_wait above math_pipe_throttle).
    static __device__ __forceinline__ el sqr(const el& a) { return mul(a, a); }

    // Dedicated Montgomery squaring (kept, tested through g16_test_field_op op 6): the 28 cross products a_i a_j (i < j)
    // once, doubled, plus the 8 squares = 36 wide multiply-adds instead of 64, then a separated reduction (8 rows of
    // m * MOD): 108 of the integer pipe's slots against 136 for mul(a, a).  Cross products of odd / even limb distance go
    // to two accumulators one limb apart (O holds limb k+1 at index k), so that every product lands on an aligned
    // register pair like in mul(); rows are ordered so that a chain's carry always lands at or above everything written
    // so far.  Limb-level model: tools/sqrsim.py.
    static __device__ __forceinline__ el sqr_dedicated(const el& a) {
        const uint32_t a0 = a.v[0], a1 = a.v[1], a2 = a.v[2], a3 = a.v[3], a4 = a.v[4], a5 = a.v[5], a6 = a.v[6], a7 = a.v[7];
        uint32_t E[16], O[16];
#pragma unroll
        for (int k = 0; k < 16; k++) E[k] = O[k] = 0;
        // row 0: a0 * {a1,a3,a5,a7} -> O[0..7], a0 * {a2,a4,a6} -> E[2..7]
        O[8] += mad_pairs4_c(O[0], O[1], O[2], O[3], O[4], O[5], O[6], O[7], a1, a3, a5, a7, a0);
        E[8] += mad_pairs3_c(E[2], E[3], E[4], E[5], E[6], E[7], a2, a4, a6, a0);
        // row 1: a1 * {a2,a4,a6} -> O[2..7], a1 * {a3,a5,a7} -> E[4..9]
        O[8] += mad_pairs3_c(O[2], O[3], O[4], O[5], O[6], O[7], a2, a4, a6, a1);
        E[10] += mad_pairs3_c(E[4], E[5], E[6], E[7], E[8], E[9], a3, a5, a7, a1);
        // row 2: a2 * {a3,a5,a7} -> O[4..9], a2 * {a4,a6} -> E[6..9]
        O[10] += mad_pairs3_c(O[4], O[5], O[6], O[7], O[8], O[9], a3, a5, a7, a2);
        E[10] += mad_pairs2_c(E[6], E[7], E[8], E[9], a4, a6, a2);
        // row 3: a3 * {a4,a6} -> O[6..9], a3 * {a5,a7} -> E[8..11]
        O[10] += mad_pairs2_c(O[6], O[7], O[8], O[9], a4, a6, a3);
        E[12] += mad_pairs2_c(E[8], E[9], E[10], E[11], a5, a7, a3);
        // row 4: a4 * {a5,a7} -> O[8..11], a4 * a6 -> E[10..11]
        O[12] += mad_pairs2_c(O[8], O[9], O[10], O[11], a5, a7, a4);
        E[12] += mad_pairs1_c(E[10], E[11], a6, a4);
        // row 5: a5 * a6 -> O[10..11], a5 * a7 -> E[12..13]
        O[12] += mad_pairs1_c(O[10], O[11], a6, a5);
        E[14] += mad_pairs1_c(E[12], E[13], a7, a5);
        // row 6: a6 * a7 -> O[12..13]
        O[14] += mad_pairs1_c(O[12], O[13], a7, a6);
        // C = E + (O << 32): 15 limbs of carry chain (E[0] = E[1] = 0: no cross product sits at positions 0, 1 of E)
        uint32_t T[16];
        T[0] = 0;
        asm("add.cc.u32 %0, %15, %30;\n\t"     // T1 = E1 + O0
            "addc.cc.u32 %1, %16, %31;\n\t"
            "addc.cc.u32 %2, %17, %32;\n\t"
            "addc.cc.u32 %3, %18, %33;\n\t"
            "addc.cc.u32 %4, %19, %34;\n\t"
            "addc.cc.u32 %5, %20, %35;\n\t"
            "addc.cc.u32 %6, %21, %36;\n\t"
            "addc.cc.u32 %7, %22, %37;\n\t"
            "addc.cc.u32 %8, %23, %38;\n\t"
            "addc.cc.u32 %9, %24, %39;\n\t"
            "addc.cc.u32 %10, %25, %40;\n\t"
            "addc.cc.u32 %11, %26, %41;\n\t"
            "addc.cc.u32 %12, %27, %42;\n\t"
            "addc.cc.u32 %13, %28, %43;\n\t"
            "addc.u32 %14, %29, %44;"
            : "=r"(T[1]), "=r"(T[2]), "=r"(T[3]), "=r"(T[4]), "=r"(T[5]), "=r"(T[6]), "=r"(T[7]), "=r"(T[8]), "=r"(T[9]),
              "=r"(T[10]), "=r"(T[11]), "=r"(T[12]), "=r"(T[13]), "=r"(T[14]), "=r"(T[15])
            : "r"(E[1]), "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]), "r"(E[8]), "r"(E[9]), "r"(E[10]),
              "r"(E[11]), "r"(E[12]), "r"(E[13]), "r"(E[14]), "r"(E[15]), "r"(O[0]), "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]),
              "r"(O[5]), "r"(O[6]), "r"(O[7]), "r"(O[8]), "r"(O[9]), "r"(O[10]), "r"(O[11]), "r"(O[12]), "r"(O[13]), "r"(O[14]));
        // double (the cross sum is below 2^511: no bit falls off the top)
#pragma unroll
        for (int k = 15; k > 0; k--) T[k] = __funnelshift_l(T[k - 1], T[k], 1);
        // + a_i^2 at limbs (2i, 2i+1): one carry chain through all sixteen limbs
        asm("mad.lo.cc.u32 %0, %16, %16, %0;\n\t"
            "madc.hi.cc.u32 %1, %16, %16, %1;\n\t"
            "madc.lo.cc.u32 %2, %17, %17, %2;\n\t"
            "madc.hi.cc.u32 %3, %17, %17, %3;\n\t"
            "madc.lo.cc.u32 %4, %18, %18, %4;\n\t"
            "madc.hi.cc.u32 %5, %18, %18, %5;\n\t"
            "madc.lo.cc.u32 %6, %19, %19, %6;\n\t"
            "madc.hi.cc.u32 %7, %19, %19, %7;\n\t"
            "madc.lo.cc.u32 %8, %20, %20, %8;\n\t"
            "madc.hi.cc.u32 %9, %20, %20, %9;\n\t"
            "madc.lo.cc.u32 %10, %21, %21, %10;\n\t"
            "madc.hi.cc.u32 %11, %21, %21, %11;\n\t"
            "madc.lo.cc.u32 %12, %22, %22, %12;\n\t"
            "madc.hi.cc.u32 %13, %22, %22, %13;\n\t"
            "madc.lo.cc.u32 %14, %23, %23, %14;\n\t"
            "madc.hi.u32 %15, %23, %23, %15;"
            : "+r"(T[0]), "+r"(T[1]), "+r"(T[2]), "+r"(T[3]), "+r"(T[4]), "+r"(T[5]), "+r"(T[6]), "+r"(T[7]), "+r"(T[8]),
              "+r"(T[9]), "+r"(T[10]), "+r"(T[11]), "+r"(T[12]), "+r"(T[13]), "+r"(T[14]), "+r"(T[15])
            : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7));
        // separated Montgomery reduction in the shifting even/odd frame: F = limbs f..f+7 of the running value
        uint32_t F[8], G[8], hi = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) F[k] = T[k];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t m = F[0] * P::INV;
            mul_row(G, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);                    // limbs f+1 .. f+8
            uint32_t c = mad_row_c(F, P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);     // F[0] becomes 0, carry at limb f+8
            uint32_t t, h1, N[8];
            asm("add.cc.u32 %0, %2, %3;\n\t"
                "addc.u32 %1, 0, 0;"
                : "=r"(t), "=r"(h1)
                : "r"(T[8 + i]), "r"(c + hi));
            asm("add.cc.u32 %0, %9, %16;\n\t"
                "addc.cc.u32 %1, %10, %17;\n\t"
                "addc.cc.u32 %2, %11, %18;\n\t"
                "addc.cc.u32 %3, %12, %19;\n\t"
                "addc.cc.u32 %4, %13, %20;\n\t"
                "addc.cc.u32 %5, %14, %21;\n\t"
                "addc.cc.u32 %6, %15, %22;\n\t"
                "addc.cc.u32 %7, %24, %23;\n\t"
                "addc.u32 %8, %25, 0;"
                : "=r"(N[0]), "=r"(N[1]), "=r"(N[2]), "=r"(N[3]), "=r"(N[4]), "=r"(N[5]), "=r"(N[6]), "=r"(N[7]), "=r"(hi)
                : "r"(F[1]), "r"(F[2]), "r"(F[3]), "r"(F[4]), "r"(F[5]), "r"(F[6]), "r"(F[7]), "r"(G[0]), "r"(G[1]), "r"(G[2]),
                  "r"(G[3]), "r"(G[4]), "r"(G[5]), "r"(G[6]), "r"(G[7]), "r"(t), "r"(h1));
#pragma unroll
            for (int k = 0; k < 8; k++) F[k] = N[k];
        }
        final_sub(F);   // a < MOD: the reduced value is below 2 MOD < 2^255, so hi is 0 here
        el o;
#pragma unroll
        for (int k = 0; k < 8; k++) o.v[k] = F[k];
        return o;
    }

    static __device__ __forceinline__ el to_mont(const el& a) { return mul(r2(), a); }
    static __device__ __forceinline__ el from_mont(const el& a) {
        el o = zero();
        o.v[0] = 1;
        return mul(a, o);
    }

    // a^e, e given as 8 little-endian limbs (host constant); used for inversions off the hot path
    static __device__ el pow(const el& a, const uint32_t (&e)[8]) {
        el r = one();
        for (int i = 255; i >= 0; i--) {
            r = sqr(r);
            if ((e[i >> 5] >> (i & 31)) & 1) r = mul(r, a);
        }
        return r;
    }
    static __device__ el inv(const el& a) {  // Fermat; inv(0) = 0
        uint32_t e[8];
#pragma unroll
        for (int i = 0; i < 8; i++) e[i] = P::mod(i);
        e[0] -= 2;  // MOD[0] >= 2 for both primes, no borrow
        return pow(a, e);
    }
};

typedef Field<FqParams> Fq;
typedef Field<FrParams> Fr;

// ---- Fq2 = Fq[u]/(u^2+1) (wasmcurves build_f2m, non-residue -1) ----------------------------------------
struct alignas(16) fe2_t {
    fe_t c0, c1;
};

struct Fq2 {
    typedef fe2_t el;
    static __device__ __forceinline__ el zero() { return el{Fq::zero(), Fq::zero()}; }
    static __device__ __forceinline__ el one() { return el{Fq::one(), Fq::zero()}; }
    static __device__ __forceinline__ bool is_zero(const el& a) { return Fq::is_zero(a.c0) && Fq::is_zero(a.c1); }
    static __device__ __forceinline__ bool eq(const el& a, const el& b) { return Fq::eq(a.c0, b.c0) && Fq::eq(a.c1, b.c1); }
    static __device__ __forceinline__ el add(const el& a, const el& b) { return el{Fq::add(a.c0, b.c0), Fq::add(a.c1, b.c1)}; }
    static __device__ __forceinline__ el sub(const el& a, const el& b) { return el{Fq::sub(a.c0, b.c0), Fq::sub(a.c1, b.c1)}; }
    static __device__ __forceinline__ el dbl(const el& a) { return el{Fq::dbl(a.c0), Fq::dbl(a.c1)}; }
    static __device__ __forceinline__ el neg(const el& a) { return el{Fq::neg(a.c0), Fq::neg(a.c1)}; }
    // Karatsuba: 3 base-field products
    static __device__ __forceinline__ el mul(const el& a, const el& b) {
        fe_t t0 = Fq::mul(a.c0, b.c0);
        fe_t t1 = Fq::mul(a.c1, b.c1);
        fe_t s = Fq::mul(Fq::add(a.c0, a.c1), Fq::add(b.c0, b.c1));
        return el{Fq::sub(t0, t1), Fq::sub(Fq::sub(s, t0), t1)};
    }
    static __device__ el inv(const el& a) {  // conj(a) / (a0^2 + a1^2)
        fe_t d = Fq::inv(Fq::add(Fq::sqr(a.c0), Fq::sqr(a.c1)));
        return el{Fq::mul(a.c0, d), Fq::neg(Fq::mul(a.c1, d))};
    }
    // (a0+a1)(a0-a1) + 2 a0 a1 u : 2 base-field products
    static __device__ __forceinline__ el sqr(const el& a) {
        fe_t t = Fq::mul(a.c0, a.c1);
        fe_t c0 = Fq::mul(Fq::add(a.c0, a.c1), Fq::sub(a.c0, a.c1));
        return el{c0, Fq::dbl(t)};
    }
};

// ---- Fq2 split over a pair of lanes ---------------------------------------------------------------------------
// Lanes (2k, 2k+1) of a warp hold ONE Fq2 element: the even lane its c0, the odd lane its c1.  Every operation is
// called by both lanes of the pair together (control flow must be uniform within a pair); halves are exchanged with
// shuffles under the pair's own two-lane mask, so different pairs of a warp may sit in different branches.
// Why: a G2 XYZZ accumulator is 64 registers per thread in the one-thread form (255 registers + stack, 2 CTAs per SM,
// 58 % pipe utilisation); split, a thread holds 32 and the kernel fits the G1 budget (128 registers, 4 CTAs per SM).
// Cost: the product is the 4-multiplication schoolbook form, two per lane (Karatsuba's three do not split evenly),
// the squaring stays one per lane.
struct Fq2Half {
    typedef fe_t el;
    static __device__ __forceinline__ uint32_t half() { return threadIdx.x & 1u; }
    static __device__ __forceinline__ uint32_t pair_mask() { return 3u << (threadIdx.x & 30u); }
    static __device__ __forceinline__ el other(const el& a) {
        el o;
        const uint32_t m = pair_mask();
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = __shfl_xor_sync(m, a.v[i], 1);
        return o;
    }
    static __device__ __forceinline__ el zero() { return Fq::zero(); }
    static __device__ __forceinline__ el one() { return half() ? Fq::zero() : Fq::one(); }
    static __device__ __forceinline__ bool is_zero(const el& a) {
        bool z = Fq::is_zero(a);
        bool oz = __shfl_xor_sync(pair_mask(), (int)z, 1) != 0;
        return z && oz;
    }
    static __device__ __forceinline__ bool eq(const el& a, const el& b) {
        bool z = Fq::eq(a, b);
        bool oz = __shfl_xor_sync(pair_mask(), (int)z, 1) != 0;
        return z && oz;
    }
    static __device__ __forceinline__ el add(const el& a, const el& b) { return Fq::add(a, b); }
    static __device__ __forceinline__ el sub(const el& a, const el& b) { return Fq::sub(a, b); }
    static __device__ __forceinline__ el dbl(const el& a) { return Fq::dbl(a); }
    static __device__ __forceinline__ el neg(const el& a) { return Fq::neg(a); }
    // (a0 + a1 u)(b0 + b1 u) = (a0 b0 - a1 b1) + (a0 b1 + a1 b0) u
    static __device__ __forceinline__ el mul(const el& a, const el& b) {
        const uint32_t h = half();
        el ob = other(b);
        el same = Fq::mul(a, b);      // lane 0: a0 b0, lane 1: a1 b1
        el cross = Fq::mul(a, ob);    // lane 0: a0 b1, lane 1: a1 b0
        el send;
#pragma unroll
        for (int i = 0; i < 8; i++) send.v[i] = h ? same.v[i] : cross.v[i];
        el recv = other(send);        // lane 0 gets a1 b1, lane 1 gets a0 b1
        return h ? Fq::add(cross, recv) : Fq::sub(same, recv);
    }
    // c0 = (a0 + a1)(a0 - a1), c1 = 2 a0 a1: one product per lane
    static __device__ __forceinline__ el sqr(const el& a) {
        const uint32_t h = half();
        el oa = other(a);
        el u = h ? a : Fq::add(a, oa);
        el v = h ? oa : Fq::sub(a, oa);
        el r = Fq::mul(u, v);
        return h ? Fq::dbl(r) : r;
    }
};

// 32-byte element load/store as two 16-byte vector accesses
__device__ __forceinline__ fe_t ld_fe(const fe_t* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    fe_t r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ fe_t ldg_fe(const fe_t* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    fe_t r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void st_fe(fe_t* p, const fe_t& r) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
    q[1] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}

}  // namespace g16
