// Pippenger multi-scalar multiplication over BN254 G1/G2 for sm_100a.
//
// Replaces: ffjavascript 0.2.62 engine_multiexp.js -> wasmcurves g1m/g2m_multiexpAffine_chunk,
// rapidsnark multiexp.cpp (G1.multiMulByScalar / G2.multiMulByScalar); reference call site scripts/g16_prove.sh:246-260.
// Inputs are the zkey's own records: affine Montgomery bases (64 B G1, 128 B G2; all-zero = infinity) and 32-byte
// little-endian standard-form scalars (wtns section 2, or the H scalars from the NTT pipeline).
//
// Pipeline of one MSM.  Every step is enqueued on CUDA streams without the host ever reading a counter back: sizes
// that depend on the scalars (entries, tasks, cut buckets) live in device memory, kernels that depend on them run
// grid-stride loops over fixed grids, and buffers are sized from bounds known at reserve() time.
//   count    : signed c-bit windows per scalar; every non-zero digit adds one to its bucket's counter
//              (key = window * 2^(c-1) + |digit| - 1; warp-aggregated atomics: a quarter of a circom witness is 0/1)
//   scan     : exclusive sum of the counters = first entry of every bucket (cub::DeviceScan; size nb is static)
//   scatter  : the digits again; entry (point index | sign) goes to start[key] + cursor[key]++.  This is a counting
//              sort by bucket: the order of entries INSIDE a bucket is arbitrary, which group addition does not care
//              about (the affine result is the same point), so no radix sort and no key array exist.
//   tasks    : a bucket with more than T entries is cut into tasks of <= T entries (T chosen on the device from the
//              entry count); tasks are ordered longest-first by a counting sort on their length
//   accumulate (the hot kernel, IMAD bound): one thread per task, XYZZ mixed additions, next point prefetched
//   combine  : buckets that were cut are summed by one CTA each
//   reduce   : per chunk of L buckets a running-sum gives sum (b+1) B_b; chunks are tree-summed per window
//   windows  : the W window sums go to the host, which runs the Horner chain (c doublings per window)
#pragma once
#include <cub/cub.cuh>
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include "common.cuh"
#include "curve.cuh"

namespace g16 {

constexpr uint32_t MSM_INDEX_MASK = 0x7fffffffu;   // entry value: bit 31 = negate the base, bits 0-30 = base index

static inline int msm_num_windows(int c) { return (255 + c - 1) / c; }

// heuristic window width: bucket reduction (2 full adds per bucket per window) stays a few % of accumulation
static inline int msm_pick_c(size_t n) {
    int lg = 0;
    while ((size_t(1) << (lg + 1)) <= n) lg++;
    int c = lg - 5;
    if (c < 6) c = 6;
    if (c > 20) c = 20;
    return c;
}

// block-wide tree sum of XYZZ points held one per thread; result in every thread.
// The staging array is STATIC shared memory (16 KB G1 / 32 KB G2 at 128 threads): with an `extern __shared__`
// byte array cast to pt*, nvcc 12.9 drops the first 16-byte store of `sh[threadIdx.x] = acc` (seen in the PTX).
// Ends with a barrier, so it can be called again in a loop.
template <class C, int THREADS>
__device__ __forceinline__ void block_sum_points(typename C::pt& acc) {
    __shared__ typename C::pt sh[THREADS];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) {
            typename C::pt a = sh[threadIdx.x];
            C::add(a, sh[threadIdx.x + s]);
            sh[threadIdx.x] = a;
        }
        __syncthreads();
    }
    acc = sh[0];
    __syncthreads();
}

// counters of one MSM run, in device memory (MsmParams) and copied to the host with the window sums
struct MsmParams {
    uint32_t m;        // entries = non-zero digits = mixed additions
    uint32_t T;        // max entries per task
    uint32_t nt;       // tasks
    uint32_t np;       // partial sums of cut buckets
    uint32_t nblk;     // first-level combine blocks
    uint32_t ncut;     // cut buckets
    uint32_t pad[2];
};

#ifdef G16_MSM_INSTANTIATE   // the kernels are compiled in msm.cu only (one copy in the library, one module to load)
// ---------------------------------------------------------------------------------------------------------
// digits: count and scatter
// ---------------------------------------------------------------------------------------------------------
// scalars: n x 8 u32 (standard form, < 2^254).  inf_mask: optional bitmask, bit i set = base i is infinity.
// fold_n != 0 (fixed-base window table, see msm_window_table_kernel): every window shares ONE bucket set and the
// entry points at row w of the table, base index w * fold_n + i, which holds 2^(c w) * base_i.
// SCATTER = false: count[key]++.  SCATTER = true: vals[start[key] + cursor[key]++] = entry.
// AGG: lanes of a warp that hit the same bucket in the same window share one atomic (witness-like scalars: the
// digit-1 bucket of window 0 holds a quarter of all entries; without this its counter serialises the kernel).
// Dense uniform scalars (the H MSM) have no such bucket and skip the match.
template <bool SCATTER, bool AGG>
static __global__ void __launch_bounds__(256)
msm_digits_kernel(const uint32_t* __restrict__ scalars, uint32_t n, const uint32_t* __restrict__ inf_mask, int c, int W,
                  uint32_t fold_n, uint32_t* __restrict__ count, const uint32_t* __restrict__ start,
                  uint32_t* __restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t half = 1u << (c - 1);
    const uint32_t mask = (1u << c) - 1;
    const uint32_t lane = threadIdx.x & 31;
    uint32_t s[8];
    bool live = i < n;
    if (live && inf_mask && ((inf_mask[i >> 5] >> (i & 31)) & 1)) live = false;
    if (live) {
        const uint4* p = reinterpret_cast<const uint4*>(scalars + size_t(i) * 8);
        uint4 a = p[0], b = p[1];
        s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
        if ((s[0] | s[1] | s[2] | s[3] | s[4] | s[5] | s[6] | s[7]) == 0) live = false;
    }
    if (!AGG && !live) return;
    if (!live) {
#pragma unroll
        for (int k = 0; k < 8; k++) s[k] = 0;
    }
    int nbits = 0;   // position of the scalar's highest set bit + 1
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (s[k]) nbits = 32 * k + 32 - __clz(s[k]);
    uint32_t carry = 0;
    for (int w = 0; w < W; w++) {
        int bit = w * c, limb = bit >> 5, sh = bit & 31;
        // nothing left above this window in any lane of the warp: done (bit and small wires end early)
        if (AGG && !__any_sync(0xffffffffu, bit < nbits || carry)) break;
        unsigned long long v = limb < 8 ? s[limb] : 0;
        if (limb + 1 < 8) v |= (unsigned long long)s[limb + 1] << 32;
        uint32_t d = (uint32_t)((v >> sh) & mask) + carry;
        carry = d > half;
        bool nz = d != 0 && d != (1u << c);                // d == 2^c becomes digit 0 with a carry
        uint32_t neg = carry;
        uint32_t mag = neg ? (1u << c) - d : d;            // 1 .. half when nz
        uint32_t key = fold_n ? mag - 1 : (uint32_t)w * half + (mag - 1);
        uint32_t val = (fold_n ? (uint32_t)w * fold_n + i : i) | (neg << 31);
        if (AGG) {
            uint32_t grp = __match_any_sync(0xffffffffu, nz ? key : 0xffffffffu);
            if (!nz) continue;
            uint32_t leader = __ffs(grp) - 1, cnt = __popc(grp), rank = __popc(grp & ((1u << lane) - 1));
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(&count[key], cnt);
            if (SCATTER) {
                base = __shfl_sync(grp, base, leader);
                vals[start[key] + base + rank] = val;
            }
        } else {
            if (!nz) continue;
            uint32_t base = atomicAdd(&count[key], 1u);
            if (SCATTER) vals[start[key] + base] = val;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// fixed-base window table (built once per key, for the H section whose scalars are dense)
// ---------------------------------------------------------------------------------------------------------
// table[w * n + i] = 2^(c w) * base_i in affine form, w = 0 .. Wd-1.  With it sum_i s_i B_i = sum_{i,w} d_{i,w} table[w][i]
// is ONE bucket problem of 2^(c-1) buckets instead of Wd of them: c can be 4 bits wider for the same bucket count,
// which removes ~20 % of the mixed additions, and the 255-doubling Horner chain at the end disappears.
// One thread per base: Wd-1 runs of c doublings in XYZZ, then one shared inversion (prefix products of the ZZZ) to
// bring the Wd-1 points back to affine: x = X (ZZ / ZZZ)^2, y = Y / ZZZ.
constexpr int MSM_TABLE_MAX_W = 43;   // c >= 6
template <class C>
static __global__ void __launch_bounds__(128)
msm_window_table_kernel(const typename C::aff* __restrict__ bases, uint32_t n, int c, int Wd, typename C::aff* __restrict__ table) {
    typedef typename C::field F;
    typedef typename C::fe fe;
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    aff a = bases[i];
    table[i] = a;
    if (C::is_inf(a)) {
        for (int w = 1; w < Wd; w++) table[size_t(w) * n + i] = a;
        return;
    }
    fe zz[MSM_TABLE_MAX_W], zzz[MSM_TABLE_MAX_W], pre[MSM_TABLE_MAX_W];
    pt q = C::from_affine(a);
    fe prod = F::one();
    for (int w = 1; w < Wd; w++) {
        for (int k = 0; k < c; k++) q = C::dbl(q);
        table[size_t(w) * n + i] = aff{q.x, q.y};   // numerators; divided below
        zz[w] = q.zz;
        zzz[w] = q.zzz;
        pre[w] = prod;
        prod = F::mul(prod, q.zzz);
    }
    fe inv = F::inv(prod);
    for (int w = Wd - 1; w >= 1; w--) {
        fe u = F::mul(inv, pre[w]);        // 1 / ZZZ_w
        inv = F::mul(inv, zzz[w]);
        fe zi = F::mul(u, zz[w]);          // 1 / Z_w
        aff t = table[size_t(w) * n + i];
        t.x = F::mul(t.x, F::sqr(zi));
        t.y = F::mul(t.y, u);
        table[size_t(w) * n + i] = t;
    }
}

// Dense uniform scalars over a window table (the H MSM): nearly every window of every scalar is a live digit, so the
// entry count is known up front to within 2^-c: digit w of scalar i goes to the FIXED slot w * n + i (coalesced stores,
// no atomics), dead digits carry the sentinel key nb, and a radix sort of all n * W slots (a count the host knows, so
// still no read-back) brings the buckets together with the sentinels at the end.  The counting sort above needs one
// returning atomic and one scattered 4-byte store per entry, which at 2^24 scalars (2 x 10^8 entries) is 3 x slower
// than the radix sort's streaming passes.
static __global__ void __launch_bounds__(256)
msm_digits_dense_kernel(const uint32_t* __restrict__ scalars, uint32_t n, const uint32_t* __restrict__ inf_mask, int c, int W,
                        uint32_t nb, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t half = 1u << (c - 1);
    const uint32_t mask = (1u << c) - 1;
    uint32_t s[8];
    bool live = !(inf_mask && ((inf_mask[i >> 5] >> (i & 31)) & 1));
    {
        const uint4* p = reinterpret_cast<const uint4*>(scalars + size_t(i) * 8);
        uint4 a = p[0], b = p[1];
        s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
    }
    uint32_t carry = 0;
    for (int w = 0; w < W; w++) {
        int bit = w * c, limb = bit >> 5, sh = bit & 31;
        unsigned long long v = limb < 8 ? s[limb] : 0;
        if (limb + 1 < 8) v |= (unsigned long long)s[limb + 1] << 32;
        uint32_t d = (uint32_t)((v >> sh) & mask) + carry;
        carry = d > half;
        bool nz = live && d != 0 && d != (1u << c);
        uint32_t neg = carry;
        uint32_t mag = neg ? (1u << c) - d : d;
        size_t slot = size_t(w) * n + i;
        keys[slot] = nz ? mag - 1 : nb;
        vals[slot] = ((uint32_t)w * n + i) | (neg << 31);
    }
}

// start[k] = first position of key k in the sorted keys (lower bound), k = 0 .. nb; start[nb] = live entries
static __global__ void msm_lower_bound_kernel(const uint32_t* __restrict__ keys, uint32_t m, uint32_t nb, uint32_t* __restrict__ start) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > nb) return;
    uint32_t lo = 0, hi = m;
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (keys[mid] < k) lo = mid + 1;
        else hi = mid;
    }
    start[k] = lo;
}

// ---------------------------------------------------------------------------------------------------------
// tasks
// ---------------------------------------------------------------------------------------------------------
constexpr uint32_t MSM_T_MAX = 512;
constexpr uint32_t COMBINE_GROUP = 1024;
constexpr uint32_t COMBINE_SMALL = 16;  // cut buckets with at most this many partial sums are summed by one thread

// one thread: entry count -> task size.  Enough tasks to fill the machine several times over: a task is one thread's
// dependency chain, so its length bounds the accumulation's tail; a cut bucket costs one extra full addition per task.
static __global__ void msm_plan_kernel(const uint32_t* __restrict__ start, uint32_t nb, uint32_t slots, uint32_t t_min,
                                       uint32_t t_fixed, MsmParams* __restrict__ prm) {
    uint32_t m = start[nb];
    uint32_t T = t_fixed;
    if (T == 0) {
        uint32_t t = m / slots;
        T = t < t_min ? t_min : t > MSM_T_MAX ? MSM_T_MAX : t;
    }
    if (T < t_min) T = t_min;
    prm->m = m;
    prm->T = T;
    prm->ncut = 0;
}

// per bucket: number of tasks, number of partial-sum slots (0 unless the bucket is cut), number of first-level
// combine blocks (COMBINE_GROUP partial sums each), list of cut buckets
static __global__ void msm_count_tasks_kernel(const uint32_t* __restrict__ start, uint32_t nb, MsmParams* __restrict__ prm,
                                              uint32_t* __restrict__ ntasks, uint32_t* __restrict__ nparts,
                                              uint32_t* __restrict__ nblk, uint32_t* __restrict__ cut_list) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > nb) return;
    const uint32_t T = prm->T;
    uint32_t t = 0;
    if (k < nb) t = (start[k + 1] - start[k] + T - 1) / T;
    ntasks[k] = t;
    nparts[k] = t > 1 ? t : 0;
    nblk[k] = t > COMBINE_SMALL ? (t + COMBINE_GROUP - 1) / COMBINE_GROUP : 0;
    if (t > 1) cut_list[atomicAdd(&prm->ncut, 1u)] = k;
}

// last k with base[k] <= j (base has nb+1 non-decreasing entries, base[nb] = total)
static __device__ __forceinline__ uint32_t msm_owner(const uint32_t* __restrict__ base, uint32_t nb, uint32_t j) {
    uint32_t lo = 0, hi = nb;
    while (lo < hi) {
        uint32_t mid = (lo + hi + 1) >> 1;
        if (base[mid] <= j) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

// task j -> (first entry, length, output slot); slot bit 31 set = partial-sum array, else bucket array.
// Also the histogram of task lengths (shared-memory histogram per CTA, one global atomic per non-empty bin).
static __global__ void __launch_bounds__(256)
msm_make_tasks_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ task_base,
                      const uint32_t* __restrict__ part_base, const uint32_t* __restrict__ blk_base, uint32_t nb,
                      MsmParams* __restrict__ prm, uint32_t* __restrict__ t_begin, uint32_t* __restrict__ t_len,
                      uint32_t* __restrict__ t_out, uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[MSM_T_MAX + 1];
    for (uint32_t b = threadIdx.x; b <= MSM_T_MAX; b += blockDim.x) sh[b] = 0;
    __syncthreads();
    const uint32_t nt = task_base[nb], T = prm->T;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        prm->nt = nt;
        prm->np = part_base[nb];
        prm->nblk = blk_base[nb];
    }
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < nt; j += gridDim.x * blockDim.x) {
        uint32_t k = msm_owner(task_base, nb, j);
        uint32_t local = j - task_base[k];
        uint32_t b = start[k] + local * T;
        uint32_t len = min(T, start[k + 1] - b);
        uint32_t nt_k = task_base[k + 1] - task_base[k];
        t_begin[j] = b;
        t_len[j] = len;
        t_out[j] = nt_k > 1 ? (0x80000000u | (part_base[k] + local)) : k;
        atomicAdd(&sh[len], 1u);
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b <= MSM_T_MAX; b += blockDim.x)
        if (sh[b]) atomicAdd(&hist[b], sh[b]);
}

// one CTA: cursor[len] = number of tasks longer than len (longest tasks first)
static __global__ void __launch_bounds__(1024) msm_order_base_kernel(const uint32_t* __restrict__ hist, uint32_t* __restrict__ cursor) {
    __shared__ uint32_t sh[1024];
    uint32_t t = threadIdx.x;                       // bin of length MSM_T_MAX - t
    uint32_t v = t <= MSM_T_MAX ? hist[MSM_T_MAX - t] : 0;
    sh[t] = v;
    __syncthreads();
    for (uint32_t o = 1; o < 1024; o <<= 1) {
        uint32_t a = t >= o ? sh[t - o] : 0;
        __syncthreads();
        sh[t] += a;
        __syncthreads();
    }
    if (t <= MSM_T_MAX) cursor[MSM_T_MAX - t] = sh[t] - v;
}

// order[position] = task, positions by decreasing length: per tile of 1024 tasks a CTA ranks its tasks per length in
// shared memory, reserves one range per length with a single global atomic, and writes
static __global__ void __launch_bounds__(256)
msm_order_tasks_kernel(const uint32_t* __restrict__ t_len, const MsmParams* __restrict__ prm, uint32_t* __restrict__ cursor,
                       uint32_t* __restrict__ order) {
    __shared__ uint32_t cnt[MSM_T_MAX + 1], base[MSM_T_MAX + 1];
    const uint32_t nt = prm->nt;
    for (uint32_t tile = blockIdx.x * 1024; tile < nt; tile += gridDim.x * 1024) {
        for (uint32_t b = threadIdx.x; b <= MSM_T_MAX; b += blockDim.x) cnt[b] = 0;
        __syncthreads();
        uint32_t len[4], rank[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t j = tile + q * 256 + threadIdx.x;
            len[q] = j < nt ? t_len[j] : 0;
            rank[q] = len[q] ? atomicAdd(&cnt[len[q]], 1u) : 0;
        }
        __syncthreads();
        for (uint32_t b = threadIdx.x; b <= MSM_T_MAX; b += blockDim.x)
            if (cnt[b]) base[b] = atomicAdd(&cursor[b], cnt[b]);
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (len[q]) order[base[len[q]] + rank[q]] = tile + q * 256 + threadIdx.x;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------
// accumulate: the hot kernel
// ---------------------------------------------------------------------------------------------------------
// MINB = minimum resident CTAs per SM the register allocation must allow (G1: 144 registers give 3; 4 caps them at 128)
template <class C, int MINB>
__global__ void __launch_bounds__(128, MINB)
msm_accumulate_kernel(const typename C::aff* __restrict__ bases, const uint32_t* __restrict__ vals,
                      const uint32_t* __restrict__ order, const uint32_t* __restrict__ t_begin,
                      const uint32_t* __restrict__ t_len, const uint32_t* __restrict__ t_out,
                      const MsmParams* __restrict__ prm, typename C::pt* __restrict__ buckets,
                      typename C::pt* __restrict__ parts) {
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= prm->nt) return;
    uint32_t j = order[tid];
    uint32_t e = t_begin[j], len = t_len[j], out = t_out[j];
    pt acc = C::identity();
    uint32_t v = vals[e];
    aff q = bases[v & MSM_INDEX_MASK];
    for (uint32_t k = 0; k < len; k++) {
        uint32_t vn = 0;
        aff qn = q;
        if (k + 1 < len) {  // prefetch the next point while this one is added
            vn = vals[e + k + 1];
            qn = bases[vn & MSM_INDEX_MASK];
        }
        if (v >> 31) q.y = C::field::neg(q.y);
        C::madd(acc, q);
        v = vn;
        q = qn;
    }
    if (out >> 31) parts[out & 0x7fffffffu] = acc;
    else buckets[out] = acc;
}

// The G2 accumulation over lane pairs (Fq2Half, field.cuh): lanes (2k, 2k+1) run ONE task, the even lane on the c0
// halves of every coordinate, the odd lane on the c1 halves.  Same registers per thread as the G1 kernel (32 for the
// accumulator, 16 + 16 for the point and the prefetched one), so 4 CTAs per SM instead of 2 and no stack.
__global__ void __launch_bounds__(128, 4)
msm_accumulate_g2_pair_kernel(const g2_aff* __restrict__ bases, const uint32_t* __restrict__ vals,
                              const uint32_t* __restrict__ order, const uint32_t* __restrict__ t_begin,
                              const uint32_t* __restrict__ t_len, const uint32_t* __restrict__ t_out,
                              const MsmParams* __restrict__ prm, g2_pt* __restrict__ buckets, g2_pt* __restrict__ parts) {
    typedef Curve<Fq2Half> CH;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t task = tid >> 1;
    const uint32_t h = tid & 1u;
    if (task >= prm->nt) return;   // uniform within a pair
    uint32_t j = order[task];
    uint32_t e = t_begin[j], len = t_len[j], out = t_out[j];
    auto fetch = [&](uint32_t v) {   // record = x.c0 x.c1 y.c0 y.c1
        const fe_t* rec = reinterpret_cast<const fe_t*>(bases + (v & MSM_INDEX_MASK));
        CH::aff a;
        a.x = ldg_fe(rec + h);
        a.y = ldg_fe(rec + 2 + h);
        return a;
    };
    CH::pt acc = CH::identity();
    uint32_t v = vals[e];
    CH::aff q = fetch(v);
    for (uint32_t k = 0; k < len; k++) {
        uint32_t vn = 0;
        CH::aff qn = q;
        if (k + 1 < len) {
            vn = vals[e + k + 1];
            qn = fetch(vn);
        }
        if (v >> 31) q.y = Fq::neg(q.y);
        CH::madd(acc, q);
        v = vn;
        q = qn;
    }
    fe_t* o = reinterpret_cast<fe_t*>((out >> 31) ? parts + (out & 0x7fffffffu) : buckets + out);
    st_fe(o + h, acc.x);
    st_fe(o + 2 + h, acc.y);
    st_fe(o + 4 + h, acc.zz);
    st_fe(o + 6 + h, acc.zzz);
}

// Most cut buckets hold a handful of partial sums (a bucket a few times over the task size): one thread each.
template <class C>
__global__ void __launch_bounds__(128)
msm_combine_small_kernel(const uint32_t* __restrict__ cut_list, const MsmParams* __restrict__ prm,
                         const uint32_t* __restrict__ part_base, const typename C::pt* __restrict__ parts,
                         typename C::pt* __restrict__ buckets) {
    const uint32_t ncut = prm->ncut;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ncut; i += gridDim.x * blockDim.x) {
        uint32_t k = cut_list[i];
        uint32_t b = part_base[k], e = part_base[k + 1];
        if (e - b > COMBINE_SMALL) continue;
        typename C::pt acc = parts[b];
        for (uint32_t j = b + 1; j < e; j++) C::add(acc, parts[j]);
        buckets[k] = acc;
    }
}

// The heavy cut buckets are summed in two levels so that a bucket holding a quarter of all points (the digit-1 bucket of a
// circom witness) is still a short dependency chain: level 1 = one CTA per COMBINE_GROUP partial sums,
// level 2 = one CTA per cut bucket over its level-1 results.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_combine1_kernel(const uint32_t* __restrict__ blk_base, const uint32_t* __restrict__ part_base, uint32_t nb,
                    const MsmParams* __restrict__ prm, const typename C::pt* __restrict__ parts,
                    typename C::pt* __restrict__ parts2) {
    const uint32_t nblk = prm->nblk;
    for (uint32_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
        uint32_t k = msm_owner(blk_base, nb, blk);
        uint32_t g = blk - blk_base[k];
        uint32_t b = part_base[k] + g * COMBINE_GROUP;
        uint32_t e = min(part_base[k + 1], b + COMBINE_GROUP);
        typename C::pt acc = C::identity();
        for (uint32_t i = b + threadIdx.x; i < e; i += THREADS) C::add(acc, parts[i]);
        block_sum_points<C, THREADS>(acc);
        if (threadIdx.x == 0) parts2[blk] = acc;
    }
}

template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_combine2_kernel(const uint32_t* __restrict__ cut_list, const uint32_t* __restrict__ blk_base,
                    const MsmParams* __restrict__ prm, const typename C::pt* __restrict__ parts2,
                    typename C::pt* __restrict__ buckets) {
    const uint32_t ncut = prm->ncut;
    for (uint32_t i = blockIdx.x; i < ncut; i += gridDim.x) {
        uint32_t k = cut_list[i];
        uint32_t b = blk_base[k], e = blk_base[k + 1];
        if (e == b) continue;  // a small cut bucket: msm_combine_small_kernel wrote it (uniform over the CTA)
        typename C::pt acc = C::identity();
        for (uint32_t j = b + threadIdx.x; j < e; j += THREADS) C::add(acc, parts2[j]);
        block_sum_points<C, THREADS>(acc);
        if (threadIdx.x == 0) buckets[k] = acc;
    }
}

// Bucket reduction, sum_b (b+1) B_b per window, in two levels so that no thread multiplies by its chunk offset at
// the bucket level (that double-and-add was a third of the kernel):  with chunks of L,
//   sum_b (b+1) B_b = sum_ch A0_ch + L * sum_ch ch * S0_ch,   A0_ch = sum_j (j+1) B_{L ch + j},  S0_ch = sum_j B_{L ch + j}
// and the second sum is the same problem over the cpw = bpw/L chunk sums (zero-based weights), where the offset
// multiplication is paid once per L*L buckets.
// Thread per chunk of L consecutive points of one window:
//   A_out[t] = sum_j (j + ONE_BASED) X[lo + j]  (+ lo * sum_j X[lo + j]  if MUL);   S_out[t] = sum_j X[lo + j] (if given)
template <class C, bool ONE_BASED, bool MUL>
__global__ void __launch_bounds__(64)
msm_chunk_reduce_kernel(const typename C::pt* __restrict__ in, uint32_t per_window, uint32_t L,
                        uint32_t chunks_per_window, uint32_t nchunks, typename C::pt* __restrict__ A_out,
                        typename C::pt* __restrict__ S_out) {
    typedef typename C::pt pt;
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nchunks) return;
    uint32_t w = t / chunks_per_window, ch = t % chunks_per_window;
    uint32_t lo = ch * L;
    uint32_t hi = min(lo + L, per_window);
    const pt* X = in + size_t(w) * per_window;
    pt run = C::identity(), acc = C::identity();
    for (uint32_t b = hi; b-- > lo;) {
        if (ONE_BASED) {
            C::add(run, X[b]);
            C::add(acc, run);
        } else {
            C::add(acc, run);
            C::add(run, X[b]);
        }
    }
    if (MUL && lo) {
        pt s = C::mul_small(run, lo);
        C::add(acc, s);
    }
    A_out[t] = acc;
    if (S_out) S_out[t] = run;
}

// CTA (w, p) sums slice p of window w's `per_window` points: out[w * slices + p]
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_sum_slices_kernel(const typename C::pt* __restrict__ in, uint32_t per_window, uint32_t slices,
                      typename C::pt* __restrict__ out) {
    uint32_t w = blockIdx.x / slices, p = blockIdx.x % slices;
    uint32_t len = (per_window + slices - 1) / slices;
    uint32_t lo = p * len, hi = min(lo + len, per_window);
    const typename C::pt* X = in + size_t(w) * per_window;
    typename C::pt acc = C::identity();
    for (uint32_t i = lo + threadIdx.x; i < hi; i += THREADS) C::add(acc, X[i]);
    block_sum_points<C, THREADS>(acc);
    if (threadIdx.x == 0) out[blockIdx.x] = acc;
}

// Level 1 of the bucket reduction and the slice sums of the level-0 results in ONE launch (both read only level-0
// output, both are latency-bound on few threads, so they overlap instead of queueing): CTAs [0, W * slices) sum
// slices of a0, the rest run the level-1 chunks (zero-based weights, chunk offset by double-and-add).
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_level1_kernel(const typename C::pt* __restrict__ a0, uint32_t per_window, uint32_t slices, uint32_t slice_ctas,
                  typename C::pt* __restrict__ slice_out, const typename C::pt* __restrict__ s0, uint32_t L1,
                  uint32_t chunks_per_window, uint32_t nchunks, typename C::pt* __restrict__ a1_out) {
    typedef typename C::pt pt;
    if (blockIdx.x < slice_ctas) {
        uint32_t w = blockIdx.x / slices, p = blockIdx.x % slices;
        uint32_t len = (per_window + slices - 1) / slices;
        uint32_t lo = p * len, hi = min(lo + len, per_window);
        const pt* X = a0 + size_t(w) * per_window;
        pt acc = C::identity();
        for (uint32_t i = lo + threadIdx.x; i < hi; i += THREADS) C::add(acc, X[i]);
        block_sum_points<C, THREADS>(acc);
        if (threadIdx.x == 0) slice_out[blockIdx.x] = acc;
        return;
    }
    uint32_t t = (blockIdx.x - slice_ctas) * THREADS + threadIdx.x;
    if (t >= nchunks) return;
    uint32_t w = t / chunks_per_window, ch = t % chunks_per_window;
    uint32_t lo = ch * L1;
    uint32_t hi = min(lo + L1, per_window);
    const pt* X = s0 + size_t(w) * per_window;
    pt run = C::identity(), acc = C::identity();
    for (uint32_t b = hi; b-- > lo;) {
        C::add(acc, run);
        C::add(run, X[b]);
    }
    if (lo) {
        pt s = C::mul_small(run, lo);
        C::add(acc, s);
    }
    a1_out[t] = acc;
}

// one CTA per window: window sum = sum of a0[n0] + L * sum of a1[n1]
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_window_sum_kernel(const typename C::pt* __restrict__ a0, uint32_t n0, const typename C::pt* __restrict__ a1,
                      uint32_t n1, uint32_t L, typename C::pt* __restrict__ window_out) {
    const typename C::pt* p = a0 + size_t(blockIdx.x) * n0;
    const typename C::pt* q = a1 + size_t(blockIdx.x) * n1;
    typename C::pt acc = C::identity(), hi = C::identity();
    for (uint32_t i = threadIdx.x; i < n0; i += THREADS) C::add(acc, p[i]);
    for (uint32_t i = threadIdx.x; i < n1; i += THREADS) C::add(hi, q[i]);
    // L * (sum of hi) = sum of L * hi: the weight is applied per thread so that ONE block tree finishes the window
    if (!C::is_identity(hi)) {
        hi = C::mul_small(hi, L);
        C::add(acc, hi);
    }
    block_sum_points<C, THREADS>(acc);
    if (threadIdx.x == 0) window_out[blockIdx.x] = acc;
}

#endif  // G16_MSM_INSTANTIATE

// bit i of mask = base i is the all-zero record (infinity)
template <class C>
__global__ void msm_inf_mask_kernel(const typename C::aff* __restrict__ bases, uint32_t n, uint32_t* __restrict__ mask) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool inf = i < n ? C::is_inf(bases[i]) : true;
    uint32_t bits = __ballot_sync(0xffffffffu, inf);
    if ((threadIdx.x & 31) == 0 && i < n) mask[i >> 5] = bits;
}


// ---------------------------------------------------------------------------------------------------------
// host side: workspace + enqueue
// ---------------------------------------------------------------------------------------------------------
// Witness-scalar MSMs (A, B1, B2, C): most scalars are 0/1 or below 2^64, so there is about one non-zero digit per
// point instead of W; narrower windows keep the bucket reduction (2 full additions per bucket per window) in
// proportion.  Measured on c2-layer-one: c = 15 -> 12 shrinks 278 K buckets to 45 K for 25 % more additions.
static inline int msm_pick_c_sparse(size_t n) {
    int c = msm_pick_c(n) - 3;
    return c < 6 ? 6 : c;
}

// window width with a fixed-base table: 4 bits wider (one bucket set instead of W), then the narrowest width that
// keeps the same number of table rows
static inline int msm_pick_c_table(size_t n) {
    int c = msm_pick_c(n) + 4;
    if (c > 24) c = 24;
    int W = msm_num_windows(c);
    while (c > 6 && msm_num_windows(c - 1) == W) c--;
    return c;
}
// the same for witness-like scalars: 2 bits narrower than the dense choice (a third of the entries of a dense MSM,
// so the one bucket set must be smaller for its reduction to stay in proportion)
static inline int msm_pick_c_table_sparse(size_t n) {
    int c = msm_pick_c(n) + 2;
    if (c > 22) c = 22;
    return c;
}

struct MsmConfig {
    bool sparse = false;  // scalars are witness-like (see msm_pick_c_sparse)
    int c = 0;          // window bits (0 = pick from n)
    uint32_t T = 0;     // max entries per task (0 = from the entry count: about 1024 tasks' worth of work per SM)
    uint32_t L = 0;     // buckets per reduction chunk (0 = from the bucket count)
    bool table = false; // fixed-base window table (build_table() before the first run)
};

// The sizes reserve() derives from (n, config), and the limits it enforces -- host arithmetic only, no CUDA call, so that
// the limits can be tested without a device (g16_test_msm_shape).  Entry positions and counters are 32-bit, base indices
// 31-bit (the top bit carries the sign of the digit), and cub's radix sort (the dense path) takes an int item count.
struct MsmShape {
    int c = 0, Wd = 0, W = 0;
    uint32_t bpw = 0, nb = 0;
    size_t m_max = 0;      // digits an MSM of this shape can produce = entries to sort
    bool dense = false;    // fixed-slot digits + radix sort (see Msm::dense)
};
inline size_t msm_dense_min() {
    static size_t v = [] { const char* e = std::getenv("G16_DENSE_MIN"); return e ? size_t(std::atoll(e)) : size_t(80) << 20; }();
    return v;
}
inline MsmShape msm_shape(size_t n, const MsmConfig& cfg) {
    MsmShape sh;
    int c = cfg.c ? cfg.c
          : cfg.table ? (cfg.sparse ? msm_pick_c_table_sparse(n) : msm_pick_c_table(n))
          : cfg.sparse ? msm_pick_c_sparse(n) : msm_pick_c(n);
    if (c < 2) c = 2;
    if (c > 24) c = 24;
    if (cfg.table && c < 6) c = 6;
    sh.c = c;
    sh.Wd = msm_num_windows(c);
    sh.W = cfg.table ? 1 : sh.Wd;
    if (size_t(sh.W) > 256) throw std::runtime_error("MSM: too many windows");
    sh.bpw = 1u << (c - 1);
    sh.nb = uint32_t(sh.W) * sh.bpw;
    if (n > size_t(MSM_INDEX_MASK)) throw std::runtime_error("MSM too large for 31-bit base indices; shard it");
    sh.m_max = n * size_t(sh.Wd);
    if (sh.m_max >= (size_t(1) << 32)) throw std::runtime_error("MSM too large for 32-bit entry indices; shard it");
    if (cfg.table && sh.m_max > size_t(MSM_INDEX_MASK))
        throw std::runtime_error("MSM window table too large for 31-bit base indices; shard it");
    sh.dense = cfg.table && !cfg.sparse && sh.m_max > msm_dense_min();
    if (sh.dense && sh.m_max > size_t(INT_MAX))
        throw std::runtime_error("MSM too large for the radix sort's int item count; shard it");
    // cub's scans take int item counts as well: nb + 1 <= 11 * 2^23 + 1 (c <= 24), far below INT_MAX
    return sh;
}

// Counters a finished MSM leaves behind (for bench.py's work model and tests)
struct MsmStats {
    unsigned long long entries = 0;  // non-zero digits = mixed additions issued
    uint32_t tasks = 0, cut_buckets = 0, buckets = 0, T = 0;
    int c = 0, W = 0;
};

template <class C>
class Msm {
public:
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    typedef typename C::fe fe;
    static constexpr uint32_t T_MIN = sizeof(pt) > 128 ? 16 : 32;
    static constexpr uint32_t T_MAX = 512;

    Msm() {}
    ~Msm() {
        if (h_params) cudaFreeHost(h_params);
        if (h_windows) cudaFreeHost(h_windows);
        if (ev_acc0) cudaEventDestroy(ev_acc0);
        if (ev_acc1) cudaEventDestroy(ev_acc1);
        for (auto& e : ev_t)
            if (e) cudaEventDestroy(e);
    }

    // capacity for n points
    void reserve(size_t n, const MsmConfig& cfg_in = MsmConfig());

    // cfg.table: build the window table of these bases (device pointer, n <= capacity) once; every later run must
    // pass the same bases.  Returns the table's size in bytes.
    size_t build_table(const aff* bases_, size_t n_, cudaStream_t s);
    static size_t table_bytes(size_t n, int c) { return size_t(msm_num_windows(c)) * n * sizeof(aff); }

    // timeline of the last run relative to `base` (an event recorded earlier), milliseconds:
    // [0] start, [1] digits counted, [2] entries scattered + tasks ordered, [3] accumulate starts, [4] accumulate done,
    // [5] window sums on the host
    void trace(cudaEvent_t base, float out[6]) const {
        cudaEvent_t e[6] = {ev_t[0], ev_t[1], ev_t[2], ev_acc0, ev_acc1, ev_t[3]};
        for (int i = 0; i < 6; i++) {
            out[i] = -1;
            if (!enqueued) continue;
            float ms = 0;
            if (cudaEventElapsedTime(&ms, base, e[i]) == cudaSuccess) out[i] = ms;
        }
    }

    // duration of the last accumulate kernel (CUDA events on its own stream); valid once wait() returned
    float accumulate_ms() const {
        float ms = 0;
        if (!enqueued || cudaEventElapsedTime(&ms, ev_acc0, ev_acc1) != cudaSuccess) return 0;
        return ms;
    }

    // Enqueue the whole MSM; nothing here waits for the device.  Counting sort of the digits and the task list go on
    // `s`; the accumulation and the reductions on `s_acc` (nullptr = the same stream): the prover keeps the H MSM's
    // preparation on the high-priority transform stream and its accumulation on the lowest-priority one, so that it
    // fills the machine under everything else.  scalars/bases/mask are device pointers; n <= reserved capacity.
    // `gate` (optional): an event the accumulation waits for (scheduling: see prove_device).
    // `s_red` (optional): a third stream for what follows the accumulation kernel (combine, bucket reduction, copies to
    // the host): latency-bound work on few threads, which the prover puts on high-priority side streams while the
    // accumulation kernels of all five MSMs run back to back on one stream.
    void enqueue(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s,
                 cudaStream_t s_acc = nullptr, cudaEvent_t gate = nullptr, cudaStream_t s_red = nullptr);

    // block until the window sums and counters of the last enqueue() are in host memory
    void wait() {
        if (!enqueued) return;
        G16_CUDA(cudaEventSynchronize(ev_t[3]));
        stats = MsmStats();
        stats.entries = h_params->m; stats.c = c; stats.W = Wd; stats.buckets = nb; stats.T = h_params->T;
        stats.tasks = h_params->nt;
        stats.cut_buckets = h_params->ncut;
    }

    // whole MSM on one stream; returns after the window sums are in host memory
    void run(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s) {
        enqueue(bases_, scalars, inf_mask, n_, s);
        wait();
    }

    // sum_w 2^(c w) S_w on the host (255 dependent doublings: a latency chain, an order of magnitude quicker on a
    // CPU core than on one GPU thread).  HC = g16h::HG1 / HG2, whose XYZZ layout equals the device one.
    // Valid after wait().
    template <class HC>
    typename HC::pt finish() const {
        static_assert(sizeof(typename HC::pt) == sizeof(pt), "host/device point layouts differ");
        typename HC::pt acc = HC::identity(), sw;
        for (int w = W - 1; w >= 0; w--) {
            for (int i = 0; i < c; i++) acc = HC::dbl(acc);
            std::memcpy(&sw, h_windows + w, sizeof(pt));
            acc = HC::add(acc, sw);
        }
        return acc;
    }
    static void prepare() {}

    MsmStats stats;
    int c = 0, W = 0;    // W = windows of the bucket problem (1 with a table)
    int Wd = 0;          // digit windows = rows of the table
    uint32_t L = 0, L1 = 4;

private:
    static int acc_minb() {   // 4 = the 128-register build of the G1 accumulation (4 CTAs/SM; measured 1-2 % faster than
                              // the 148-register one); G16_ACC_MINB=1 selects the other for comparison
        static int v = [] { const char* e = std::getenv("G16_ACC_MINB"); return e ? std::atoi(e) : 4; }();
        return v;
    }
    static constexpr int RED_THREADS = 128;
    // slices a window's chunk results are summed in before the one-CTA window sum: none below 1024 points, else about
    // 4 points per thread of a 128-thread CTA, 32..256 slices
    static uint32_t num_slices(uint32_t per_window) {
        if (per_window < 1024) return 0;
        uint32_t k = per_window / 512;
        return k < 32 ? 32 : k > 256 ? 256 : k;
    }
    MsmConfig cfg;
    size_t cap_n = 0, n = 0, slots = 148 * 1024;
    int sms = 148;
    uint32_t bpw = 0, nb = 0, cpw = 0, cpw1 = 0, ns0 = 0, ns1 = 0;
    uint32_t nt_max = 0, np_max = 0, nblk_max = 0, ncut_max = 0;
    const aff* bases = nullptr;
    DevBuf vals, keys_a, keys_b, vals_b, count, start, ntasks, nparts, nblk, task_base, part_base, blk_base, cut_list,
        t_begin, t_len, t_out, order, hist, buckets, parts, parts2, chunk_out, chunk_sum, chunk2_out, slice_out,
        slice1_out, window_out, params, cub_tmp, table;
    size_t table_n = 0;
    MsmParams* h_params = nullptr;
    pt* h_windows = nullptr;
    cudaEvent_t ev_acc0 = nullptr, ev_acc1 = nullptr, ev_t[4] = {nullptr, nullptr, nullptr, nullptr};
    bool enqueued = false;
    // fixed-slot digits + radix sort instead of the counting sort: dense scalars over a table, and enough entries that the
    // counting sort's scattered stores no longer stay in L2 (measured: 27 M entries 0.48 ms counting / 0.87 ms radix,
    // 201 M entries 8.0 ms / 6.0 ms)
    bool dense() const { return cfg.table && !cfg.sparse && cap_n * size_t(Wd) > dense_min(); }
    static size_t dense_min() { return msm_dense_min(); }
};

#ifdef G16_MSM_INSTANTIATE
// ---- out-of-class definitions: compiled in msm.cu only, so that the kernels exist once in the library
template <class C>
void Msm<C>::reserve(size_t n, const MsmConfig& cfg_in) {
    cfg = cfg_in;
    cap_n = n;
    const MsmShape sh = msm_shape(n, cfg);   // throws beyond the index limits ("shard it")
    c = sh.c;
    Wd = sh.Wd;
    W = sh.W;
    bpw = sh.bpw;
    nb = sh.nb;
    size_t m_max = sh.m_max;
    if (cfg.T > T_MAX) cfg.T = T_MAX;
    // level-0 chunks: 16 buckets when there are enough buckets to fill the machine with chunk threads, else 8
    // (half the dependency chain); level 1 works on bpw/L points only, so its chunks are short (latency, not work)
    L = cfg.L ? cfg.L : (nb >= (2u << 20) ? 16 : 8);
    L1 = 4;
    cpw = (bpw + L - 1) / L;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    slots = size_t(sms) * (sizeof(pt) > 128 ? 512 : 1024);
    // Bounds that hold for every scalar vector (msm_plan_kernel picks T = clamp(m / slots, T_MIN, T_MAX) on the
    // device): q >= m / T, tasks <= non-empty buckets + q, partial sums <= 2 q (a cut bucket has more than T
    // entries), combine blocks < partial sums / 16 (a bucket with blocks has more than COMBINE_SMALL partial sums)
    size_t q = cfg.T ? m_max / (cfg.T < T_MIN ? T_MIN : cfg.T) + 1 : std::max(2 * slots + 2, m_max / T_MAX + 1);
    if (q > m_max / T_MIN + 1) q = m_max / T_MIN + 1;
    nt_max = uint32_t(std::min<size_t>(size_t(nb) + q, m_max ? m_max : 1));
    np_max = uint32_t(std::min<size_t>(2 * q, m_max ? m_max : 1));
    nblk_max = np_max / COMBINE_SMALL + 1;
    ncut_max = uint32_t(std::min<size_t>(nb, q));
    size_t e = m_max ? m_max : 1;
    vals.alloc(e * 4);
    size_t t1 = 0;
    if (dense()) {
        keys_a.alloc(e * 4); keys_b.alloc(e * 4); vals_b.alloc(e * 4);
        cub::DoubleBuffer<uint32_t> dk(keys_a.as<uint32_t>(), keys_b.as<uint32_t>());
        cub::DoubleBuffer<uint32_t> dv(vals.as<uint32_t>(), vals_b.as<uint32_t>());
        int bits = 1;
        while ((size_t(1) << bits) <= nb) bits++;      // keys 0 .. nb (nb = sentinel)
        cub::DeviceRadixSort::SortPairs(nullptr, t1, dk, dv, (int)e, 0, bits);
    }
    count.alloc(size_t(nb + 1) * 4); start.alloc(size_t(nb + 1) * 4);
    ntasks.alloc(size_t(nb + 1) * 4); nparts.alloc(size_t(nb + 1) * 4); nblk.alloc(size_t(nb + 1) * 4);
    task_base.alloc(size_t(nb + 1) * 4); part_base.alloc(size_t(nb + 1) * 4); blk_base.alloc(size_t(nb + 1) * 4);
    cut_list.alloc((size_t(ncut_max) + 1) * 4);
    t_begin.alloc(size_t(nt_max) * 4); t_len.alloc(size_t(nt_max) * 4); t_out.alloc(size_t(nt_max) * 4);
    order.alloc(size_t(nt_max) * 4);
    hist.alloc(2 * (MSM_T_MAX + 1) * 4);   // length histogram, then the cursors
    parts.alloc(size_t(np_max) * sizeof(pt));
    parts2.alloc(size_t(nblk_max) * sizeof(pt));
    buckets.alloc(size_t(nb) * sizeof(pt));
    cpw1 = (cpw + L1 - 1) / L1;
    chunk_out.alloc(size_t(W) * cpw * sizeof(pt));
    chunk_sum.alloc(size_t(W) * cpw * sizeof(pt));
    chunk2_out.alloc(size_t(W) * cpw1 * sizeof(pt));
    ns0 = num_slices(cpw);
    ns1 = num_slices(cpw1);
    slice_out.alloc(size_t(W) * (ns0 ? ns0 : 1) * sizeof(pt));
    slice1_out.alloc(size_t(W) * (ns1 ? ns1 : 1) * sizeof(pt));
    window_out.alloc(size_t(W) * sizeof(pt));
    params.alloc(sizeof(MsmParams));
    size_t t3 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, t3, count.as<uint32_t>(), start.as<uint32_t>(), (int)(nb + 1));
    cub_tmp.alloc((t1 > t3 ? t1 : t3) + 256);
    if (!h_params) G16_CUDA(cudaMallocHost(&h_params, sizeof(MsmParams)));
    if (!h_windows) G16_CUDA(cudaMallocHost(&h_windows, 256 * sizeof(pt)));
    if (!ev_acc0) G16_CUDA(cudaEventCreate(&ev_acc0));
    if (!ev_acc1) G16_CUDA(cudaEventCreate(&ev_acc1));
    for (auto& e : ev_t)
        if (!e) G16_CUDA(cudaEventCreate(&e));
    enqueued = false;
}

template <class C>
size_t Msm<C>::build_table(const aff* bases_, size_t n_, cudaStream_t s) {
    if (!cfg.table) throw std::runtime_error("MSM: build_table without cfg.table");
    if (n_ > cap_n) throw std::runtime_error("MSM: n exceeds reserved capacity");
    table_n = n_;
    if (!n_) return 0;
    table.alloc(size_t(Wd) * n_ * sizeof(aff));
    msm_window_table_kernel<C><<<(unsigned)((n_ + 127) / 128), 128, 0, s>>>(bases_, (uint32_t)n_, c, Wd, table.template as<aff>());
    G16_CUDA(cudaGetLastError());
    return size_t(Wd) * n_ * sizeof(aff);
}

template <class C>
void Msm<C>::enqueue(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s,
                     cudaStream_t s_acc, cudaEvent_t gate, cudaStream_t s_red) {
    if (n_ > cap_n) throw std::runtime_error("MSM: n exceeds reserved capacity");
    if (cfg.table && n_ != table_n) throw std::runtime_error("MSM: window table was built for a different base count");
    bases = cfg.table && n_ ? table.template as<aff>() : bases_;
    n = n_;
    uint32_t* cnt = count.as<uint32_t>();
    uint32_t* st = start.as<uint32_t>();
    MsmParams* prm = params.as<MsmParams>();
    const uint32_t fold = cfg.table ? (uint32_t)n : 0u;
    const unsigned dgrid = (unsigned)((n + 255) / 256);
    // warp-aggregated counters unless the scalars are known to be dense and uniform (the H MSM over its window table)
    const bool agg = !(cfg.table && !cfg.sparse);
    G16_CUDA(cudaEventRecord(ev_t[0], s));
    const uint32_t* entry_vals = vals.as<uint32_t>();
    size_t tb = cub_tmp.bytes;
    if (dense()) {
        // ---- fixed-slot digits + radix sort (see msm_digits_dense_kernel)
        size_t slots_n = n * size_t(Wd);
        uint32_t m_live = 0;
        if (n) {
            msm_digits_dense_kernel<<<dgrid, 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd, nb, keys_a.as<uint32_t>(), vals.as<uint32_t>());
            G16_LAUNCHED(1);
            cub::DoubleBuffer<uint32_t> dk(keys_a.as<uint32_t>(), keys_b.as<uint32_t>());
            cub::DoubleBuffer<uint32_t> dv(vals.as<uint32_t>(), vals_b.as<uint32_t>());
            int bits = 1;
            while ((size_t(1) << bits) <= nb) bits++;
            G16_CUDA(cudaEventRecord(ev_t[1], s));
            G16_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, dk, dv, (int)slots_n, 0, bits, s));
            entry_vals = dv.Current();
            m_live = (uint32_t)slots_n;
            msm_lower_bound_kernel<<<(nb + 1 + 255) / 256, 256, 0, s>>>(dk.Current(), m_live, nb, st);
            G16_LAUNCHED(1);
        } else {
            G16_CUDA(cudaMemsetAsync(st, 0, size_t(nb + 1) * 4, s));
            G16_CUDA(cudaEventRecord(ev_t[1], s));
        }
        msm_plan_kernel<<<1, 1, 0, s>>>(st, nb, (uint32_t)slots, T_MIN, cfg.T, prm);
    } else {
    // ---- counting sort of the digits by bucket
    G16_CUDA(cudaMemsetAsync(cnt, 0, size_t(nb + 1) * 4, s));
    if (n) {
        if (agg) msm_digits_kernel<false, true><<<dgrid, 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd, fold, cnt, nullptr, nullptr);
        else msm_digits_kernel<false, false><<<dgrid, 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd, fold, cnt, nullptr, nullptr);
        G16_LAUNCHED(1);
    }
    tb = cub_tmp.bytes;
    G16_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, cnt, st, (int)(nb + 1), s));
    msm_plan_kernel<<<1, 1, 0, s>>>(st, nb, (uint32_t)slots, T_MIN, cfg.T, prm);
    G16_CUDA(cudaEventRecord(ev_t[1], s));
    G16_CUDA(cudaMemsetAsync(cnt, 0, size_t(nb + 1) * 4, s));
    if (n) {
        if (agg) msm_digits_kernel<true, true><<<dgrid, 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd, fold, cnt, st, vals.as<uint32_t>());
        else msm_digits_kernel<true, false><<<dgrid, 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd, fold, cnt, st, vals.as<uint32_t>());
        G16_LAUNCHED(1);
    }
    }   // counting sort
    // ---- tasks, longest first
    msm_count_tasks_kernel<<<(nb + 1 + 255) / 256, 256, 0, s>>>(st, nb, prm, ntasks.as<uint32_t>(), nparts.as<uint32_t>(),
                                                                nblk.as<uint32_t>(), cut_list.as<uint32_t>());
    uint32_t* src[3] = {ntasks.as<uint32_t>(), nparts.as<uint32_t>(), nblk.as<uint32_t>()};
    uint32_t* dst[3] = {task_base.as<uint32_t>(), part_base.as<uint32_t>(), blk_base.as<uint32_t>()};
    for (int i = 0; i < 3; i++) {
        tb = cub_tmp.bytes;
        G16_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, src[i], dst[i], (int)(nb + 1), s));
    }
    uint32_t* hs = hist.as<uint32_t>();
    uint32_t* cursor = hs + (MSM_T_MAX + 1);
    G16_CUDA(cudaMemsetAsync(hs, 0, (MSM_T_MAX + 1) * 4, s));
    const unsigned tgrid = (unsigned)std::min<size_t>((size_t(nt_max) + 255) / 256, size_t(sms) * 8);
    msm_make_tasks_kernel<<<tgrid, 256, 0, s>>>(st, task_base.as<uint32_t>(), part_base.as<uint32_t>(), blk_base.as<uint32_t>(),
                                                nb, prm, t_begin.as<uint32_t>(), t_len.as<uint32_t>(), t_out.as<uint32_t>(), hs);
    msm_order_base_kernel<<<1, 1024, 0, s>>>(hs, cursor);
    const unsigned ogrid = (unsigned)std::min<size_t>((size_t(nt_max) + 1023) / 1024, size_t(sms) * 8);
    msm_order_tasks_kernel<<<ogrid, 256, 0, s>>>(t_len.as<uint32_t>(), prm, cursor, order.as<uint32_t>());
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(5);
    G16_CUDA(cudaEventRecord(ev_t[2], s));
    // ---- accumulation + reductions, possibly on another stream
    if (s_acc && s_acc != s) {
        G16_CUDA(cudaStreamWaitEvent(s_acc, ev_t[2], 0));
        s = s_acc;
    }
    if (gate) G16_CUDA(cudaStreamWaitEvent(s, gate, 0));
    G16_CUDA(cudaMemsetAsync(buckets.p, 0, size_t(nb) * sizeof(pt), s));
    G16_CUDA(cudaEventRecord(ev_acc0, s));
    const unsigned agrid = (nt_max + 127) / 128;
    bool launched = false;
    if constexpr (sizeof(pt) > 128) {
        // G2: G16_G2_PAIR=1 selects the lane-pair kernel.  Measured alone it is slower than the one-thread kernel (1.45 against
        // 1.33 ms at 2^21, 4.92 against 4.25 ms at 2^24: 36 products per addition against Karatsuba's 28 outweigh the
        // occupancy gained), so it stays an option
        static const bool pair = [] { const char* e = std::getenv("G16_G2_PAIR"); return e && std::atoi(e) != 0; }();
        if (pair) {
            msm_accumulate_g2_pair_kernel<<<(2 * nt_max + 127) / 128, 128, 0, s>>>(
                bases, entry_vals, order.as<uint32_t>(), t_begin.as<uint32_t>(), t_len.as<uint32_t>(), t_out.as<uint32_t>(), prm,
                buckets.as<pt>(), parts.as<pt>());
            launched = true;
        }
    }
    if constexpr (sizeof(pt) <= 128) {
        // G16_ACC_MINB=5 / 6: 96 / 80 registers (5 / 6 CTAs per SM, 124 / 232 bytes of spills) for comparison
        if (acc_minb() == 5 || acc_minb() == 6) {
            if (acc_minb() == 5)
                msm_accumulate_kernel<C, 5><<<agrid, 128, 0, s>>>(bases, entry_vals, order.as<uint32_t>(), t_begin.as<uint32_t>(),
                                                                  t_len.as<uint32_t>(), t_out.as<uint32_t>(), prm, buckets.as<pt>(), parts.as<pt>());
            else
                msm_accumulate_kernel<C, 6><<<agrid, 128, 0, s>>>(bases, entry_vals, order.as<uint32_t>(), t_begin.as<uint32_t>(),
                                                                  t_len.as<uint32_t>(), t_out.as<uint32_t>(), prm, buckets.as<pt>(), parts.as<pt>());
            launched = true;
        }
        if (!launched && acc_minb() == 4) {
            msm_accumulate_kernel<C, 4><<<agrid, 128, 0, s>>>(bases, entry_vals, order.as<uint32_t>(), t_begin.as<uint32_t>(),
                                                              t_len.as<uint32_t>(), t_out.as<uint32_t>(), prm, buckets.as<pt>(), parts.as<pt>());
            launched = true;
        }
    }
    if (!launched)
        msm_accumulate_kernel<C, 1><<<agrid, 128, 0, s>>>(bases, entry_vals, order.as<uint32_t>(), t_begin.as<uint32_t>(),
                                                          t_len.as<uint32_t>(), t_out.as<uint32_t>(), prm, buckets.as<pt>(), parts.as<pt>());
    G16_CUDA(cudaEventRecord(ev_acc1, s));
    if (s_red && s_red != s) {
        G16_CUDA(cudaStreamWaitEvent(s_red, ev_acc1, 0));
        s = s_red;
    }
    const unsigned cgrid = (unsigned)std::min<size_t>((size_t(ncut_max) + 127) / 128, size_t(sms) * 4);
    msm_combine_small_kernel<C><<<cgrid ? cgrid : 1, 128, 0, s>>>(cut_list.as<uint32_t>(), prm, part_base.as<uint32_t>(), parts.as<pt>(),
                                                                  buckets.as<pt>());
    msm_combine1_kernel<C, RED_THREADS><<<(unsigned)std::min<size_t>(nblk_max, size_t(sms) * 4), RED_THREADS, 0, s>>>(
        blk_base.as<uint32_t>(), part_base.as<uint32_t>(), nb, prm, parts.as<pt>(), parts2.as<pt>());
    msm_combine2_kernel<C, RED_THREADS><<<(unsigned)std::min<size_t>(ncut_max ? ncut_max : 1, size_t(sms) * 4), RED_THREADS, 0, s>>>(
        cut_list.as<uint32_t>(), blk_base.as<uint32_t>(), prm, parts2.as<pt>(), buckets.as<pt>());
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(4);
    uint32_t nchunks = uint32_t(W) * cpw, nchunks1 = uint32_t(W) * cpw1;
    msm_chunk_reduce_kernel<C, true, false><<<(nchunks + 63) / 64, 64, 0, s>>>(
        buckets.as<pt>(), bpw, L, cpw, nchunks, chunk_out.as<pt>(), chunk_sum.as<pt>());
    const pt *a0 = chunk_out.as<pt>(), *a1 = chunk2_out.as<pt>();
    uint32_t n0 = cpw, n1 = cpw1;
    if (ns0) {
        // long windows: level 1 and the slice sums of the level-0 results share a launch (msm_level1_kernel); slices
        // keep every thread of the one-CTA-per-window sum below off a long dependent chain of additions
        uint32_t slice_ctas = uint32_t(W) * ns0;
        msm_level1_kernel<C, RED_THREADS><<<slice_ctas + (nchunks1 + RED_THREADS - 1) / RED_THREADS, RED_THREADS, 0, s>>>(
            chunk_out.as<pt>(), cpw, ns0, slice_ctas, slice_out.as<pt>(), chunk_sum.as<pt>(), L1, cpw1, nchunks1,
            chunk2_out.as<pt>());
        a0 = slice_out.as<pt>();
        n0 = ns0;
    } else {
        msm_chunk_reduce_kernel<C, false, true><<<(nchunks1 + 63) / 64, 64, 0, s>>>(
            chunk_sum.as<pt>(), cpw, L1, cpw1, nchunks1, chunk2_out.as<pt>(), nullptr);
    }
    if (ns1) {
        msm_sum_slices_kernel<C, RED_THREADS><<<uint32_t(W) * ns1, RED_THREADS, 0, s>>>(chunk2_out.as<pt>(), cpw1, ns1, slice1_out.as<pt>());
        a1 = slice1_out.as<pt>();
        n1 = ns1;
        G16_LAUNCHED(1);
    }
    msm_window_sum_kernel<C, RED_THREADS><<<W, RED_THREADS, 0, s>>>(a0, n0, a1, n1, L, window_out.as<pt>());
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(3);
    G16_CUDA(cudaMemcpyAsync(h_windows, window_out.p, size_t(W) * sizeof(pt), cudaMemcpyDeviceToHost, s));
    G16_CUDA(cudaMemcpyAsync(h_params, prm, sizeof(MsmParams), cudaMemcpyDeviceToHost, s));
    G16_CUDA(cudaEventRecord(ev_t[3], s));
    enqueued = true;
}
#endif  // G16_MSM_INSTANTIATE

#ifndef G16_MSM_INSTANTIATE
extern template class Msm<G1>;
extern template class Msm<G2>;
#endif

}  // namespace g16
