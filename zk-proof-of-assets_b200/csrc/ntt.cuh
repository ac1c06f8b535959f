// Radix-2 NTT passes over Fr and the H-polynomial pipeline  P = (A.B - C) on the odd coset.
//
// Replaces: ffjavascript 0.2.62 engine_fft.js / engine_applykey.js (Fr.ifft, Fr.batchApplyKey,
// Fr.fft), snarkjs 0.7.2 groth16_prove.js joinABC + batchFromMontgomery, rapidsnark FFT<Fr>::ifft/fft; reference call
// site scripts/g16_prove.sh:246-260.  Roots follow ffjavascript: w[28] = 5^((r-1)/2^28), w[k] = w[k+1]^2,
// inc = w[log2 n + 1].
//
// Order trick: the inverse transform runs decimation-in-frequency (natural in, bit-reversed out), the coset shift
// inc^i / n is applied in bit-reversed positions from a table as that transform's last pass stores, and the forward
// transform runs decimation-in-time (bit-reversed in, natural out) -- no bit-reversal pass exists.
//
// A transform of 2^logn points is ceil(logn/10) passes; a pass stages up to 10 butterfly levels of a 1024-element
// tile in shared memory and makes one coalesced round trip to HBM.  Index math prototyped in tools/nttsim.py
// (fused twiddles) and tools/nttsim2.py (split twiddles, the version that runs) against the oracle.
#pragma once
#include <cstdlib>
#include <vector>
#include "common.cuh"
#include "field.cuh"
#include "hostfield.hpp"

namespace g16 {

constexpr int NTT_TILE_LOG = 10;
constexpr int NTT_TILE = 1 << NTT_TILE_LOG;
constexpr int NTT_THREADS = 256;

// ---- first version of the pass (kept selectable with G16_NTT_FUSED=1 for A/B timing): radix-2 butterflies whose
// twiddle w^(j << ..) carries the column offset, i.e. one 32-byte table load per butterfly
// a: data (Montgomery form).  tw[e] = w^e, e < n/2.  scale: optional per-position factor applied when storing.
template <bool DIF, bool INV>
__global__ void __launch_bounds__(NTT_THREADS)
ntt_pass_fused_kernel(fe_t* __restrict__ a, int logn, int s_lo, int k, const fe_t* __restrict__ tw,
                const fe_t* __restrict__ scale) {
    __shared__ fe_t sh[NTT_TILE];
    const uint32_t n = 1u << logn;
    const uint32_t nq = n >> k;                 // sub-transforms of 2^k points, stride 2^s_lo
    const uint32_t S = NTT_TILE >> k;           // sub-transforms per tile
    const uint32_t q0 = blockIdx.x * S;
    const uint32_t Sq = min(S, nq - q0);
    const uint32_t lomask = (1u << s_lo) - 1;

    // load: shared layout follows the (coalesced) global order
    for (uint32_t e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        uint32_t qi, t;
        if (s_lo > 0) { qi = e & (S - 1); t = e >> (NTT_TILE_LOG - k); }
        else { t = e & ((1u << k) - 1); qi = e >> k; }
        if (qi < Sq) {
            uint32_t q = q0 + qi;
            size_t idx = (size_t(q >> s_lo) << (s_lo + k)) + (size_t(t) << s_lo) + (q & lomask);
            sh[e] = ld_fe(a + idx);
        }
    }
    __syncthreads();

    for (int step = 0; step < k; step++) {
        const int r = DIF ? (k - 1 - step) : step;
        const int b = s_lo + r;
        for (uint32_t u = threadIdx.x; u < NTT_TILE / 2; u += NTT_THREADS) {
            uint32_t qi, tp;
            if (s_lo > 0) { qi = u & (S - 1); tp = u >> (NTT_TILE_LOG - k); }
            else { tp = u & ((1u << (k - 1)) - 1); qi = u >> (k - 1); }
            if (qi >= Sq) continue;
            uint32_t tl = tp & ((1u << r) - 1), th = tp >> r;
            uint32_t t0 = (th << (r + 1)) | tl;
            uint32_t t1 = t0 + (1u << r);
            uint32_t lo = (q0 + qi) & lomask;
            uint32_t j = (tl << s_lo) + lo;
            uint32_t ex = j << (logn - 1 - b);
            uint32_t e0 = s_lo > 0 ? (t0 << (NTT_TILE_LOG - k)) + qi : (qi << k) + t0;
            uint32_t e1 = s_lo > 0 ? (t1 << (NTT_TILE_LOG - k)) + qi : (qi << k) + t1;
            fe_t x = sh[e0], y = sh[e1];
            fe_t w;
            bool unit = (ex == 0);
            if (!unit) {
                if (INV) w = Fr::neg(ldg_fe(tw + ((n >> 1) - ex)));
                else w = ldg_fe(tw + ex);
            }
            if (DIF) {
                fe_t d = Fr::sub(x, y);
                sh[e0] = Fr::add(x, y);
                sh[e1] = unit ? d : Fr::mul(d, w);
            } else {
                if (!unit) y = Fr::mul(y, w);
                sh[e0] = Fr::add(x, y);
                sh[e1] = Fr::sub(x, y);
            }
        }
        __syncthreads();
    }

    for (uint32_t e = threadIdx.x; e < NTT_TILE; e += NTT_THREADS) {
        uint32_t qi, t;
        if (s_lo > 0) { qi = e & (S - 1); t = e >> (NTT_TILE_LOG - k); }
        else { t = e & ((1u << k) - 1); qi = e >> k; }
        if (qi < Sq) {
            uint32_t q = q0 + qi;
            size_t idx = (size_t(q >> s_lo) << (s_lo + k)) + (size_t(t) << s_lo) + (q & lomask);
            fe_t v = sh[e];
            if (scale) v = Fr::mul(v, ldg_fe(scale + idx));
            st_fe(a + idx, v);
        }
    }
}

// ---- the pass: split twiddles ------------------------------------------------------------------------------
// A level-b butterfly of the big transform needs w_n^(((tl << s_lo) + lo) << (logn-1-b)) = w_{2^(r+1)}^tl * w_M'^lo.
// The second factor is the same for both inputs of every later (DIF) / earlier (DIT) butterfly of the pass, so it
// moves out of the butterflies: a pass is S pure 2^k-point sub-transforms whose twiddles come from a 2^(k-1)-entry
// table in shared memory, plus ONE inter-pass twiddle per element, w_M^(lo * bitrev_k(t)) with M = 2^(s_lo + k),
// applied while storing (DIF) or loading (DIT); passes with s_lo == 0 have none.  Against the fused version: the
// per-butterfly table load from L2 becomes a shared-memory read, a butterfly whose sub-transform twiddle is 1
// (tl == 0; whole warps, see the thread mapping) skips its product -- 28 % of a 2^7 pass -- and the inverse
// pass negates its table once instead of per butterfly.  Index math prototyped in tools/nttsim2.py.
// Shared layout: element (t, qi) of the tile at slot t * S + qi, two planes of 16-byte halves (a 32-byte element
// per lane would make every access a 2-way bank conflict), one slot of padding per 8.
constexpr int NTT_PLANE = NTT_TILE + NTT_TILE / 8;
static inline size_t ntt_pass_smem(int k) { return size_t(2 * NTT_PLANE) * 16 + (size_t(32) << (k > 1 ? k - 1 : 0)); }

__device__ __forceinline__ uint32_t ntt_slot(uint32_t e) { return e + (e >> 3); }
__device__ __forceinline__ fe_t ntt_lds(const uint4* pl, uint32_t slot) {
    uint4 a = pl[slot], b = pl[NTT_PLANE + slot];
    fe_t r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void ntt_sts(uint4* pl, uint32_t slot, const fe_t& r) {
    pl[slot] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
    pl[NTT_PLANE + slot] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}
// w^(+-E) for 0 < E < n from the half-size table
template <bool INV>
__device__ __forceinline__ fe_t ntt_inter_twiddle(const fe_t* __restrict__ tw, uint32_t E, uint32_t n) {
    const uint32_t half = n >> 1;
    uint32_t at;
    bool negate;
    if (INV) { negate = E <= half; at = negate ? half - E : n - E; }
    else { negate = E >= half; at = negate ? E - half : E; }
    fe_t w = ldg_fe(tw + at);
    return negate ? Fr::neg(w) : w;
}

template <bool DIF, bool INV>
__global__ void __launch_bounds__(NTT_THREADS)
ntt_pass_kernel(fe_t* __restrict__ a, int logn, int s_lo, int k, const fe_t* __restrict__ tw,
                const fe_t* __restrict__ scale) {
    extern __shared__ uint4 ntt_smem[];
    uint4* pl = ntt_smem;                      // two planes of NTT_PLANE slots
    uint4* P = ntt_smem + 2 * NTT_PLANE;       // sub-transform twiddles: entry j at P[2j], P[2j+1]
    const uint32_t n = 1u << logn;
    const uint32_t nq = n >> k;                 // sub-transforms of 2^k points, stride 2^s_lo
    const int logS = NTT_TILE_LOG - k;
    const uint32_t S = 1u << logS;              // sub-transforms per tile
    const uint32_t q0 = blockIdx.x * S;
    const uint32_t Sq = min(S, nq - q0);
    const uint32_t lomask = (1u << s_lo) - 1;
    const uint32_t kmask = (1u << k) - 1;
    const int sh_tw = logn - s_lo - k;

    for (uint32_t j = threadIdx.x; j < (1u << (k - 1)); j += NTT_THREADS) {
        uint32_t x = j << (logn - k);
        fe_t w;
        if (INV) w = x ? Fr::neg(ldg_fe(tw + ((n >> 1) - x))) : ldg_fe(tw);
        else w = ldg_fe(tw + x);
        P[2 * j] = make_uint4(w.v[0], w.v[1], w.v[2], w.v[3]);
        P[2 * j + 1] = make_uint4(w.v[4], w.v[5], w.v[6], w.v[7]);
    }

    // load (global order is coalesced either way: along qi when s_lo > 0, along t when s_lo == 0)
#pragma unroll
    for (uint32_t c = 0; c < NTT_TILE / NTT_THREADS; c++) {
        uint32_t i = threadIdx.x + c * NTT_THREADS;
        uint32_t qi, t;
        if (s_lo > 0) { qi = i & (S - 1); t = i >> logS; }
        else { t = i & kmask; qi = i >> k; }
        if (qi < Sq) {
            uint32_t q = q0 + qi, lo = q & lomask;
            size_t idx = (size_t(q >> s_lo) << (s_lo + k)) + (size_t(t) << s_lo) + lo;
            fe_t v = ld_fe(a + idx);
            if (!DIF && s_lo > 0) {
                uint32_t E = (lo * (__brev(t) >> (32 - k))) << sh_tw;
                if (E) v = Fr::mul(v, ntt_inter_twiddle<INV>(tw, E, n));
            }
            ntt_sts(pl, ntt_slot((t << logS) + qi), v);
        }
    }
    __syncthreads();

    for (int step = 0; step < k; step++) {
        const int r = DIF ? (k - 1 - step) : step;
        const int hb = k - 1 - r;               // bits of th
        // butterfly u: qi in the low bits (one quarter-warp = 8 neighbouring columns of one row pair), then th, tl on
        // top, so that tl is uniform across a warp whenever it can be and the tl == 0 warps skip the product
        for (uint32_t u = threadIdx.x; u < NTT_TILE / 2; u += NTT_THREADS) {
            uint32_t qi = u & (S - 1), v = u >> logS;
            if (qi >= Sq) continue;
            uint32_t th = v & ((1u << hb) - 1), tl = v >> hb;
            uint32_t t0 = (th << (r + 1)) | tl;
            uint32_t s0 = ntt_slot((t0 << logS) + qi), s1 = ntt_slot(((t0 + (1u << r)) << logS) + qi);
            fe_t x = ntt_lds(pl, s0), y = ntt_lds(pl, s1);
            if (DIF) {
                fe_t d = Fr::sub(x, y);
                ntt_sts(pl, s0, Fr::add(x, y));
                if (tl) {
                    uint32_t j = tl << hb;
                    uint4 wa = P[2 * j], wb = P[2 * j + 1];
                    fe_t w;
                    w.v[0] = wa.x; w.v[1] = wa.y; w.v[2] = wa.z; w.v[3] = wa.w;
                    w.v[4] = wb.x; w.v[5] = wb.y; w.v[6] = wb.z; w.v[7] = wb.w;
                    d = Fr::mul(d, w);
                }
                ntt_sts(pl, s1, d);
            } else {
                if (tl) {
                    uint32_t j = tl << hb;
                    uint4 wa = P[2 * j], wb = P[2 * j + 1];
                    fe_t w;
                    w.v[0] = wa.x; w.v[1] = wa.y; w.v[2] = wa.z; w.v[3] = wa.w;
                    w.v[4] = wb.x; w.v[5] = wb.y; w.v[6] = wb.z; w.v[7] = wb.w;
                    y = Fr::mul(y, w);
                }
                ntt_sts(pl, s0, Fr::add(x, y));
                ntt_sts(pl, s1, Fr::sub(x, y));
            }
        }
        __syncthreads();
    }

#pragma unroll
    for (uint32_t c = 0; c < NTT_TILE / NTT_THREADS; c++) {
        uint32_t i = threadIdx.x + c * NTT_THREADS;
        uint32_t qi, t;
        if (s_lo > 0) { qi = i & (S - 1); t = i >> logS; }
        else { t = i & kmask; qi = i >> k; }
        if (qi < Sq) {
            uint32_t q = q0 + qi, lo = q & lomask;
            size_t idx = (size_t(q >> s_lo) << (s_lo + k)) + (size_t(t) << s_lo) + lo;
            fe_t v = ntt_lds(pl, ntt_slot((t << logS) + qi));
            if (DIF && s_lo > 0) {
                uint32_t E = (lo * (__brev(t) >> (32 - k))) << sh_tw;
                if (E) v = Fr::mul(v, ntt_inter_twiddle<INV>(tw, E, n));
            }
            if (scale) v = Fr::mul(v, ldg_fe(scale + idx));
            st_fe(a + idx, v);
        }
    }
}

// out[i] = prod over set bits b of i (or of bitrev(i)) of pw[b], times `mult`.  pw[b] = base^(2^b) (Montgomery).
static __global__ void ntt_powers_kernel(fe_t* __restrict__ out, uint32_t count, int bits, bool bitrev,
                                  const fe_t* __restrict__ pw, fe_t mult) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint32_t e = bitrev ? (bits ? __brev(i) >> (32 - bits) : 0) : i;
    fe_t r = mult;
    for (int b = 0; b < bits; b++)
        if ((e >> b) & 1) r = Fr::mul(r, pw[b]);
    st_fe(out + i, r);
}

// P[i] = from_montgomery(A[i]*B[i] - C[i])  (rows a8: joinABC + batchFromMontgomery)
static __global__ void ntt_join_kernel(const fe_t* __restrict__ A, const fe_t* __restrict__ B, const fe_t* __restrict__ C,
                                fe_t* __restrict__ P, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fe_t v = Fr::sub(Fr::mul(ld_fe(A + i), ld_fe(B + i)), ld_fe(C + i));
    st_fe(P + i, Fr::from_mont(v));
}

// ---- a transform split over two devices (sharded proofs on >= 6 GPUs) ------------------------------------------------
// With the coefficients split by parity, c[2m+t], the coset evaluations are
//     X[k] = Y0[k mod n/2] + inc w^k Y1[k mod n/2],   Yt = odd-coset transform, on the HALF domain, of the polynomial c[2m+t]
// and c[2m+t] = 1/2 iNTT_half(b_t)[m] with b_t[i] = (a[i] + (-1)^t a[i + n/2]) w^(-i t): the first decimation-in-frequency
// level.  The half domain's coset factor is w^m (its `inc` is the full domain's w), so the full domain's scale table
// serves both halves: scale[p] holds w^brev(p) / n for p < n/2 (the 1/2 included) and inc w^brev(p - n/2) / n above
// (the join's factor inc included).  tw[e] = w^e, e < n/2, of the FULL domain in both kernels.
static __global__ void ntt_split_kernel(fe_t* __restrict__ a, uint32_t half, int parity, const fe_t* __restrict__ tw) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= half) return;
    fe_t lo = ld_fe(a + i), hi = ld_fe(a + i + half);
    if (!parity) {
        st_fe(a + i, Fr::add(lo, hi));
        return;
    }
    fe_t d = Fr::sub(lo, hi);
    if (i) d = Fr::neg(Fr::mul(d, ldg_fe(tw + (half - i))));   // w^(-i) = -w^(n/2 - i)
    st_fe(a + i, d);
}

// P[lo + i] = from_montgomery(XA XB - XC), X = E[i] + w^(lo + i) O[i]; E/O: this range's slices of the two half transforms
static __global__ void ntt_join_halves_kernel(const fe_t* __restrict__ eA, const fe_t* __restrict__ oA,
                                              const fe_t* __restrict__ eB, const fe_t* __restrict__ oB,
                                              const fe_t* __restrict__ eC, const fe_t* __restrict__ oC,
                                              fe_t* __restrict__ P, uint32_t count, uint32_t lo, uint32_t half,
                                              const fe_t* __restrict__ tw) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint32_t k = lo + i;
    fe_t w = ldg_fe(tw + (k >= half ? k - half : k));
    if (k >= half) w = Fr::neg(w);
    fe_t xa = Fr::add(ld_fe(eA + i), Fr::mul(w, ld_fe(oA + i)));
    fe_t xb = Fr::add(ld_fe(eB + i), Fr::mul(w, ld_fe(oB + i)));
    fe_t xc = Fr::add(ld_fe(eC + i), Fr::mul(w, ld_fe(oC + i)));
    st_fe(P + i, Fr::from_mont(Fr::sub(Fr::mul(xa, xb), xc)));
}

inline std::vector<int> ntt_plan(int logn) {
    int np = (logn + NTT_TILE_LOG - 1) / NTT_TILE_LOG;
    if (np < 1) np = 1;
    std::vector<int> ks(np);
    for (int i = 0; i < np; i++) ks[i] = logn / np + (i < logn % np ? 1 : 0);
    return ks;
}

// Per-domain tables, built once (zkey open) and kept resident.
struct NttDomain {
    int logn = 0;
    DevBuf tw;     // w^e, e < max(1, n/2)
    DevBuf scale;  // inc^bitrev(p) / n

    static g16h::U256 root_of_unity(int log) {  // ffjavascript F1Field: w[28] = 5^((r-1)/2^28), squared down
        using namespace g16h;
        U256 e = HFr::mod();
        e.v[0] -= 1;
        // (r-1) >> 28
        for (int i = 0; i < 4; i++) e.v[i] = (e.v[i] >> 28) | (i < 3 ? e.v[i + 1] << 36 : 0);
        U256 w = HFr::pow(HFr::from_u64(5), e);
        for (int i = 28; i > log; i--) w = HFr::sqr(w);
        return w;  // Montgomery form
    }

    void init(int logn_, cudaStream_t s, bool with_scale = true) {   // with_scale false: tw only (the half domain)
        using namespace g16h;
        logn = logn_;
        if (logn < 1 || logn > 28) throw std::runtime_error("NTT domain must be 2^1 .. 2^28");
        // a 2^10-point pass needs 36 KB of planes + 16 KB of sub-transform twiddles: above the 48 KB default
        G16_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)ntt_pass_smem(NTT_TILE_LOG)));
        G16_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)ntt_pass_smem(NTT_TILE_LOG)));
        size_t n = size_t(1) << logn;
        tw.alloc((n / 2 ? n / 2 : 1) * sizeof(fe_t));
        if (with_scale) scale.alloc(n * sizeof(fe_t));
        U256 w = root_of_unity(logn);
        // row a6: inc = w[logn + 1]; for logn == 28 ffjavascript uses Fr.shift = nqr^2 = 25
        U256 inc = logn < 28 ? root_of_unity(logn + 1) : HFr::from_u64(25);
        U256 ninv = HFr::inv(HFr::from_u64(uint64_t(n)));
        fe_t pw[64];
        U256 a = w, b = inc;
        for (int i = 0; i < 32; i++) {
            std::memcpy(&pw[i], a.v, 32);
            std::memcpy(&pw[32 + i], b.v, 32);
            a = HFr::sqr(a);
            b = HFr::sqr(b);
        }
        DevBuf dpw;
        dpw.alloc(sizeof(pw));
        G16_CUDA(cudaMemcpyAsync(dpw.p, pw, sizeof(pw), cudaMemcpyHostToDevice, s));
        fe_t one_m, ninv_m;
        U256 o = HFr::one();
        std::memcpy(&one_m, o.v, 32);
        std::memcpy(&ninv_m, ninv.v, 32);
        uint32_t half = (uint32_t)(n / 2 ? n / 2 : 1);
        ntt_powers_kernel<<<(half + 255) / 256, 256, 0, s>>>(tw.as<fe_t>(), half, logn - 1, false, dpw.as<fe_t>(), one_m);
        if (with_scale)
            ntt_powers_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(scale.as<fe_t>(), (uint32_t)n, logn, true,
                                                                          dpw.as<fe_t>() + 32, ninv_m);
        G16_CUDA(cudaGetLastError());
        G16_CUDA(cudaStreamSynchronize(s));
    }

    static bool fused() {   // G16_NTT_FUSED=1: the first version of the pass, for A/B timing
        static bool v = [] { const char* e = std::getenv("G16_NTT_FUSED"); return e && std::atoi(e) != 0; }();
        return v;
    }
    template <bool DIF, bool INV>
    void pass(fe_t* a, int sl, int k, const fe_t* sc, cudaStream_t s) const {
        unsigned grid = (unsigned)(((size_t(1) << logn) + NTT_TILE - 1) / NTT_TILE);
        if (fused()) ntt_pass_fused_kernel<DIF, INV><<<grid, NTT_THREADS, 0, s>>>(a, logn, sl, k, tw.as<fe_t>(), sc);
        else ntt_pass_kernel<DIF, INV><<<grid, NTT_THREADS, ntt_pass_smem(k), s>>>(a, logn, sl, k, tw.as<fe_t>(), sc);
    }

    // evaluations on the domain (natural order) -> evaluations on the odd coset (natural order), in place
    void to_odd_coset(fe_t* a, cudaStream_t s) const {
        inverse_dif(a, true, s);
        forward_dit(a, s);
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(2 * ntt_plan(logn).size());
    }

    // plain transforms (test hooks / standalone operator): natural order in and out is NOT provided by a single call;
    // forward_dit expects bit-reversed input, inverse_dif leaves bit-reversed output (unscaled).
    void inverse_dif(fe_t* a, bool with_scale, cudaStream_t s) const {
        std::vector<int> ks = ntt_plan(logn);
        int sl = logn;
        for (size_t i = 0; i < ks.size(); i++) {
            sl -= ks[i];
            pass<true, true>(a, sl, ks[i], (with_scale && sl == 0) ? scale.as<fe_t>() : nullptr, s);
        }
    }
    // inverse transform whose last pass multiplies by a caller's table (bit-reversed positions)
    void inverse_dif_with(fe_t* a, const fe_t* sc, cudaStream_t s) const {
        std::vector<int> ks = ntt_plan(logn);
        int sl = logn;
        for (size_t i = 0; i < ks.size(); i++) {
            sl -= ks[i];
            pass<true, true>(a, sl, ks[i], sl == 0 ? sc : nullptr, s);
        }
    }
    void forward_dit(fe_t* a, cudaStream_t s) const {
        std::vector<int> ks = ntt_plan(logn);
        int sl = 0;
        for (size_t i = ks.size(); i-- > 0;) {
            pass<false, false>(a, sl, ks[i], nullptr, s);
            sl += ks[i];
        }
    }
};

}  // namespace g16
