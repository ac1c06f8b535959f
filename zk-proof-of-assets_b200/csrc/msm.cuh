// Pippenger multi-scalar multiplication over BN254 G1/G2 for sm_100a.
//
// Replaces: ffjavascript 0.2.62 engine_multiexp.js -> wasmcurves g1m/g2m_multiexpAffine_chunk,
// rapidsnark multiexp.cpp (G1.multiMulByScalar / G2.multiMulByScalar); reference call site scripts/g16_prove.sh:246-260.
// Inputs are the zkey's own records: affine Montgomery bases (64 B G1, 128 B G2; all-zero = infinity) and 32-byte
// little-endian standard-form scalars (wtns section 2, or the H scalars from the NTT pipeline).
//
// Pipeline of one MSM (all on one stream; two tiny device->host reads of counters):
//   digits   : signed c-bit windows per scalar; only non-zero digits become (bucket key, point index|sign) entries
//   sort     : cub radix sort of the entries by key = window * 2^(c-1) + (|digit| - 1)
//   bounds   : run start/end per bucket from the sorted keys
//   tasks    : a bucket with more than T entries is cut into tasks of <= T entries (load balance for
//              witness-like scalars where one bucket can hold a quarter of all points); tasks sorted by length
//   accumulate (the hot kernel, IMAD bound): one thread per task, XYZZ mixed additions, next point prefetched
//   combine  : buckets that were cut are summed by one CTA each
//   reduce   : per chunk of L buckets a running-sum gives sum (b+1) B_b; chunks are tree-summed per window
//   windows  : the W window sums go to the host, which runs the Horner chain (c doublings per window)
#pragma once
#include <cub/cub.cuh>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include "common.cuh"
#include "curve.cuh"

namespace g16 {

// entry values: bit 31 = negate the base, bit 30 = the index points into the merged-pair array (msm_pair_merge_kernel)
constexpr uint32_t MSM_MERGED_FLAG = 0x40000000u;
constexpr uint32_t MSM_INDEX_MASK = 0x3fffffffu;

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
}

#ifdef G16_MSM_INSTANTIATE   // the kernels are compiled in msm.cu only (one copy in the library, one module to load)
// ---------------------------------------------------------------------------------------------------------
// digits
// ---------------------------------------------------------------------------------------------------------
// scalars: n x 8 u32 (standard form, < 2^254).  inf_mask: optional bitmask, bit i set = base i is infinity.
static __global__ void msm_digits_kernel(const uint32_t* __restrict__ scalars, uint32_t n, const uint32_t* __restrict__ inf_mask,
                                  int c, int W, uint32_t fold_n, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                  unsigned long long* __restrict__ counter) {
    // fold_n != 0 (fixed-base window table, see msm_window_table_kernel): every window shares ONE bucket set and the
    // entry points at row w of the table, base index w * fold_n + i, which holds 2^(c w) * base_i
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t half = 1u << (c - 1);
    const uint32_t mask = (1u << c) - 1;
    uint32_t cnt = 0;
    uint32_t s[8];
    bool live = i < n;
    if (live && inf_mask && ((inf_mask[i >> 5] >> (i & 31)) & 1)) live = false;
    if (live) {
        const uint4* p = reinterpret_cast<const uint4*>(scalars + size_t(i) * 8);
        uint4 a = p[0], b = p[1];
        s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
        if ((s[0] | s[1] | s[2] | s[3] | s[4] | s[5] | s[6] | s[7]) == 0) live = false;
    }
    // pass 1: count non-zero signed digits
    if (live) {
        uint32_t carry = 0;
        for (int w = 0; w < W; w++) {
            int bit = w * c, limb = bit >> 5, sh = bit & 31;
            unsigned long long v = limb < 8 ? s[limb] : 0;
            if (limb + 1 < 8) v |= (unsigned long long)s[limb + 1] << 32;
            uint32_t d = (uint32_t)((v >> sh) & mask) + carry;
            carry = d > half;
            if (d != 0 && d != (1u << c)) cnt++;  // d == 2^c becomes digit 0 with a carry
        }
    }
    // warp-aggregated reservation of output slots
    uint32_t lane = threadIdx.x & 31;
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long base = 0;
    if (lane == 31 && total) base = atomicAdd(counter, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 31);
    if (!live || cnt == 0) return;
    unsigned long long pos = base + (incl - cnt);
    uint32_t carry = 0;
    for (int w = 0; w < W; w++) {
        int bit = w * c, limb = bit >> 5, sh = bit & 31;
        unsigned long long v = limb < 8 ? s[limb] : 0;
        if (limb + 1 < 8) v |= (unsigned long long)s[limb + 1] << 32;
        uint32_t d = (uint32_t)((v >> sh) & mask) + carry;
        carry = d > half;
        if (d == 0 || d == (1u << c)) continue;
        uint32_t neg = carry;
        uint32_t mag = neg ? (1u << c) - d : d;  // 1 .. half
        keys[pos] = fold_n ? mag - 1 : (uint32_t)w * half + (mag - 1);
        vals[pos] = (fold_n ? (uint32_t)w * fold_n + i : i) | (neg << 31);
        pos++;
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

// ---------------------------------------------------------------------------------------------------------
// bucket bounds and tasks
// ---------------------------------------------------------------------------------------------------------
static __global__ void msm_bounds_kernel(const uint32_t* __restrict__ keys, uint32_t m, uint32_t* __restrict__ start,
                                  uint32_t* __restrict__ end) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    uint32_t k = keys[i];
    if (i == 0 || keys[i - 1] != k) start[k] = i;
    if (i == m - 1 || keys[i + 1] != k) end[k] = i + 1;
}

// per bucket: number of tasks, number of partial-sum slots (0 unless the bucket is cut), number of first-level
// combine blocks (COMBINE_GROUP partial sums each), list of cut buckets
constexpr uint32_t COMBINE_GROUP = 1024;
constexpr uint32_t COMBINE_SMALL = 16;  // cut buckets with at most this many partial sums are summed by one thread
static __global__ void msm_count_tasks_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ end, uint32_t nb,
                                       uint32_t T, uint32_t* __restrict__ ntasks, uint32_t* __restrict__ nparts,
                                       uint32_t* __restrict__ nblk, uint32_t* __restrict__ cut_list,
                                       uint32_t* __restrict__ cut_count) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > nb) return;
    uint32_t t = 0;
    if (k < nb) t = (end[k] - start[k] + T - 1) / T;
    ntasks[k] = t;
    nparts[k] = t > 1 ? t : 0;
    nblk[k] = t > COMBINE_SMALL ? (t + COMBINE_GROUP - 1) / COMBINE_GROUP : 0;
    if (t > 1) cut_list[atomicAdd(cut_count, 1u)] = k;
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

// task j -> (first entry, length, output slot); slot bit 31 set = partial-sum array, else bucket array
static __global__ void msm_make_tasks_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ end,
                                      const uint32_t* __restrict__ task_base, const uint32_t* __restrict__ part_base,
                                      uint32_t nb, uint32_t nt, uint32_t T, uint32_t* __restrict__ t_begin,
                                      uint32_t* __restrict__ t_len, uint32_t* __restrict__ t_out,
                                      uint32_t* __restrict__ sort_key, uint32_t* __restrict__ sort_val) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nt) return;
    uint32_t k = msm_owner(task_base, nb, j);
    uint32_t local = j - task_base[k];
    uint32_t b = start[k] + local * T;
    uint32_t len = min(T, end[k] - b);
    uint32_t nt_k = task_base[k + 1] - task_base[k];
    t_begin[j] = b;
    t_len[j] = len;
    t_out[j] = nt_k > 1 ? (0x80000000u | (part_base[k] + local)) : k;
    sort_key[j] = T - len;  // longest tasks first
    sort_val[j] = j;
}

// ---------------------------------------------------------------------------------------------------------
// accumulate: the hot kernel
// ---------------------------------------------------------------------------------------------------------
// MINB = minimum resident CTAs per SM the register allocation must allow (G1: 144 registers give 3; 4 caps them at 128)
template <class C, int MINB>
__global__ void __launch_bounds__(128, MINB)
msm_accumulate_kernel(const typename C::aff* __restrict__ bases, const typename C::aff* __restrict__ merged,
                      const uint32_t* __restrict__ vals, const uint32_t* __restrict__ order,
                      const uint32_t* __restrict__ t_begin, const uint32_t* __restrict__ t_len,
                      const uint32_t* __restrict__ t_out, uint32_t nt, typename C::pt* __restrict__ buckets,
                      typename C::pt* __restrict__ parts) {
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nt) return;
    uint32_t j = order[tid];
    uint32_t e = t_begin[j], len = t_len[j], out = t_out[j];
    pt acc = C::identity();
    // an entry is a base index (bit 31 = negate) or, after the pair merge, MSM_MERGED_FLAG | index of a merged point
    auto fetch = [&](uint32_t v) -> aff {
        return (v & MSM_MERGED_FLAG) ? merged[v & MSM_INDEX_MASK] : bases[v & MSM_INDEX_MASK];
    };
    uint32_t v = vals[e];
    aff q = fetch(v);
    for (uint32_t k = 0; k < len; k++) {
        uint32_t vn = 0;
        aff qn = q;
        if (k + 1 < len) {  // prefetch the next point while this one is added
            vn = vals[e + k + 1];
            qn = fetch(vn);
        }
        if (v >> 31) q.y = C::field::neg(q.y);
        if (!((v & MSM_MERGED_FLAG) && C::is_inf(q))) C::madd(acc, q);   // a merged pair of opposite points is infinity
        v = vn;
        q = qn;
    }
    if (out >> 31) parts[out & 0x7fffffffu] = acc;
    else buckets[out] = acc;
}

// ---------------------------------------------------------------------------------------------------------
// pair merge: one round of batched-affine addition in front of the accumulation
// ---------------------------------------------------------------------------------------------------------
// Sorted entries that share a bucket and sit on absolute positions (2g, 2g+1) are added in AFFINE coordinates:
// (x3, y3) = (L^2 - x1 - x2, L (x1 - x3) - y1), L = (y2 - y1) / (x2 - x1), with the inversions batched by Montgomery's
// trick over the P <= 256 pair slots of a thread (one Fermat inversion per thread).  That is 1 (prefix product) + 384/P
// (inversion) + 2 (unwinding) + 3 (L, L^2, y3) = about 7.5 field products per addition against 10 for the XYZZ mixed
// addition, on half of all additions.  The merged points go to `merged`, the entry list is rewritten compacted per
// bucket (an entry is a base index | sign, or MERGED_FLAG | merged index), and the accumulation runs on the shorter
// lists.  A bucket that starts on an odd position keeps its first entry single; so does one that ends on an even one.
// output position of input position p of a bucket that starts at s
static __device__ __forceinline__ uint32_t msm_merged_pos(uint32_t s, uint32_t p) { return s + ((p - s + (s & 1u)) >> 1); }

static __global__ void msm_merged_end_kernel(const uint32_t* __restrict__ start, uint32_t* __restrict__ end, uint32_t nb) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nb) return;
    uint32_t s = start[k], e = end[k];
    if (e > s) end[k] = msm_merged_pos(s, e - 1) + 1;
}

// kind of a pair: 0 = generic (d = x2 - x1), 1 = doubling (d = 2 y1), 2 = opposite points (sum = infinity, d = 1)
template <class C>
static __device__ __forceinline__ int msm_pair_den(const typename C::aff& a, const typename C::aff& b, typename C::fe& d) {
    typedef typename C::field F;
    d = F::sub(b.x, a.x);
    if (!F::is_zero(d)) return 0;
    if (F::eq(a.y, b.y)) {
        d = F::dbl(a.y);
        return 1;
    }
    d = F::one();
    return 2;
}

// one pair slot as loaded from memory (the loads of slot j+1 are issued before the arithmetic of slot j)
template <class C>
struct MsmPairSlot {
    typename C::aff a, b;
    uint32_t k0, v0, v1;
    bool live, pair;   // live: g < nslots; pair: both positions exist and share a bucket
};

template <class C>
static __device__ __forceinline__ MsmPairSlot<C> msm_pair_fetch(const typename C::aff* __restrict__ bases,
                                                                const uint32_t* __restrict__ keys,
                                                                const uint32_t* __restrict__ vals, uint32_t m,
                                                                uint32_t nslots, uint32_t g) {
    MsmPairSlot<C> s;
    s.live = g < nslots;
    s.pair = false;
    if (!s.live) return s;
    uint32_t p0 = 2 * g, p1 = p0 + 1;
    s.k0 = keys[p0];
    s.v0 = vals[p0];
    s.v1 = 0;
    if (p1 < m) {
        s.v1 = vals[p1];
        s.pair = keys[p1] == s.k0;
    }
    if (s.pair) {
        s.a = bases[s.v0 & MSM_INDEX_MASK];
        s.b = bases[s.v1 & MSM_INDEX_MASK];
        if (s.v0 >> 31) s.a.y = C::field::neg(s.a.y);
        if (s.v1 >> 31) s.b.y = C::field::neg(s.b.y);
    }
    return s;
}

// Thread t owns the slots t, t + T, t + 2T, ... (coalesced across a warp); the batch of one inversion is a thread's
// per_thread slots.  It has to be large: the Fermat chain is 384 sequential products PER LANE, so it is amortised
// only over the slots of that lane (a warp-wide batch with one shared inversion was measured 3x slower: the
// inversion's instruction stream costs a warp the same whether one lane or thirty-two use its result).
template <class C>
__global__ void __launch_bounds__(128)
msm_pair_merge_kernel(const typename C::aff* __restrict__ bases, const uint32_t* __restrict__ keys,
                      const uint32_t* __restrict__ vals, const uint32_t* __restrict__ start, uint32_t m, uint32_t nslots,
                      uint32_t per_thread, typename C::fe* __restrict__ prefix, typename C::aff* __restrict__ merged,
                      uint32_t* __restrict__ vals_out) {
    typedef typename C::field F;
    typedef typename C::fe fe;
    typedef typename C::aff aff;
    typedef MsmPairSlot<C> slot_t;
    const uint32_t T = gridDim.x * blockDim.x;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    // forward: prefix products of the denominators of this thread's slots
    fe prod = F::one();
    {
        slot_t cur = msm_pair_fetch<C>(bases, keys, vals, m, nslots, t);
        for (uint32_t j = 0; j < per_thread; j++) {
            uint32_t g = t + j * T;
            slot_t nxt = msm_pair_fetch<C>(bases, keys, vals, m, nslots, j + 1 < per_thread ? g + T : nslots);
            if (cur.pair) {
                fe d;
                msm_pair_den<C>(cur.a, cur.b, d);
                prefix[g] = prod;
                prod = F::mul(prod, d);
            }
            cur = nxt;
        }
    }
    fe inv = F::inv(prod);
    // backward: unwind the inverses, finish the additions, rewrite the entries
    uint32_t jl = per_thread - 1;
    slot_t cur = msm_pair_fetch<C>(bases, keys, vals, m, nslots, t + jl * T);
    fe pre = cur.pair ? prefix[t + jl * T] : F::one();
    for (uint32_t j = per_thread; j-- > 0;) {
        uint32_t g = t + j * T;
        slot_t nxt = msm_pair_fetch<C>(bases, keys, vals, m, nslots, j > 0 ? g - T : nslots);
        fe npre = nxt.pair ? prefix[g - T] : F::one();
        if (cur.live) {
            uint32_t p0 = 2 * g, p1 = p0 + 1;
            if (!cur.pair) {
                // two singles: the last entry of one bucket, the first of the next (or the odd entry that ends the list)
                vals_out[msm_merged_pos(start[cur.k0], p0)] = cur.v0;
                if (p1 < m) vals_out[p1] = cur.v1;   // start of its bucket: position unchanged
            } else {
                fe d;
                int kind = msm_pair_den<C>(cur.a, cur.b, d);
                fe dinv = F::mul(inv, pre);
                inv = F::mul(inv, d);
                aff r;
                if (kind == 2) {
                    r.x = F::zero();
                    r.y = F::zero();   // infinity: the accumulation skips it
                } else {
                    fe num;
                    if (kind == 0) num = F::sub(cur.b.y, cur.a.y);
                    else {
                        fe xx = F::sqr(cur.a.x);
                        num = F::add(F::dbl(xx), xx);
                    }
                    fe lam = F::mul(num, dinv);
                    r.x = F::sub(F::sub(F::sqr(lam), cur.a.x), cur.b.x);
                    r.y = F::sub(F::mul(lam, F::sub(cur.a.x, r.x)), cur.a.y);
                }
                merged[g] = r;
                vals_out[msm_merged_pos(start[cur.k0], p0)] = MSM_MERGED_FLAG | g;
            }
        }
        cur = nxt;
        pre = npre;
    }
}

// Most cut buckets hold a handful of partial sums (a bucket a few times over the task size): one thread each.
template <class C>
__global__ void __launch_bounds__(128)
msm_combine_small_kernel(const uint32_t* __restrict__ cut_list, uint32_t ncut, const uint32_t* __restrict__ part_base,
                         const typename C::pt* __restrict__ parts, typename C::pt* __restrict__ buckets) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncut) return;
    uint32_t k = cut_list[i];
    uint32_t b = part_base[k], e = part_base[k + 1];
    if (e - b > COMBINE_SMALL) return;
    typename C::pt acc = parts[b];
    for (uint32_t j = b + 1; j < e; j++) C::add(acc, parts[j]);
    buckets[k] = acc;
}

// The heavy cut buckets are summed in two levels so that a bucket holding a quarter of all points (the digit-1 bucket of a
// circom witness) is still a short dependency chain: level 1 = one CTA per COMBINE_GROUP partial sums,
// level 2 = one CTA per cut bucket over its level-1 results.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_combine1_kernel(const uint32_t* __restrict__ blk_base, const uint32_t* __restrict__ part_base, uint32_t nb,
                    const typename C::pt* __restrict__ parts, typename C::pt* __restrict__ parts2) {
    uint32_t k = msm_owner(blk_base, nb, blockIdx.x);
    uint32_t g = blockIdx.x - blk_base[k];
    uint32_t b = part_base[k] + g * COMBINE_GROUP;
    uint32_t e = min(part_base[k + 1], b + COMBINE_GROUP);
    typename C::pt acc = C::identity();
    for (uint32_t i = b + threadIdx.x; i < e; i += THREADS) C::add(acc, parts[i]);
    block_sum_points<C, THREADS>(acc);
    if (threadIdx.x == 0) parts2[blockIdx.x] = acc;
}

template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
msm_combine2_kernel(const uint32_t* __restrict__ cut_list, const uint32_t* __restrict__ blk_base,
                    const typename C::pt* __restrict__ parts2, typename C::pt* __restrict__ buckets) {
    uint32_t k = cut_list[blockIdx.x];
    uint32_t b = blk_base[k], e = blk_base[k + 1];
    typename C::pt acc = C::identity();
    if (e == b) return;  // a small cut bucket: msm_combine_small_kernel wrote it
    for (uint32_t i = b + threadIdx.x; i < e; i += THREADS) C::add(acc, parts2[i]);
    block_sum_points<C, THREADS>(acc);
    if (threadIdx.x == 0) buckets[k] = acc;
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
// host side: workspace + staged launch
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

struct MsmConfig {
    bool sparse = false;  // scalars are witness-like (see msm_pick_c_sparse)
    int c = 0;          // window bits (0 = pick from n)
    uint32_t T = 0;     // max entries per task (0 = from the entry count: about 1024 tasks' worth of work per SM)
    uint32_t L = 0;     // buckets per reduction chunk (0 = from the bucket count)
    bool table = false; // fixed-base window table (build_table() before the first run); dense scalars only
};

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
        if (h_counters) cudaFreeHost(h_counters);
        if (h_windows) cudaFreeHost(h_windows);
        if (ev_acc0) cudaEventDestroy(ev_acc0);
        if (ev_acc1) cudaEventDestroy(ev_acc1);
        if (ev_mrg0) cudaEventDestroy(ev_mrg0);
        if (ev_mrg1) cudaEventDestroy(ev_mrg1);
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
    // [0] start, [1] digits done, [2] sort + task counts done, [3] accumulate starts, [4] accumulate done, [5] window sums on the host
    void trace(cudaEvent_t base, float out[6]) const {
        cudaEvent_t e[6] = {ev_t[0], ev_t[1], ev_t[2], ev_acc0, ev_acc1, ev_t[3]};
        for (int i = 0; i < 6; i++) {
            out[i] = -1;
            if ((i == 3 || i == 4) && !acc_timed) continue;
            float ms = 0;
            if (cudaEventElapsedTime(&ms, base, e[i]) == cudaSuccess) out[i] = ms;
        }
    }

    // duration of the last accumulate kernel (CUDA events on its own stream); valid once the stream is synchronised
    float accumulate_ms() const {
        float ms = 0, mg = 0;
        if (!acc_timed || cudaEventElapsedTime(&ms, ev_acc0, ev_acc1) != cudaSuccess) return 0;
        if (merged_slots && cudaEventElapsedTime(&mg, ev_mrg0, ev_mrg1) == cudaSuccess) ms += mg;  // the pair merge is part of it
        return ms;
    }

    // stage 1: digits (async).  scalars/bases/mask are device pointers; n <= reserved capacity.
    void stage1(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s);

    // stage 2: sort + task counts (reads the entry count -> one stream sync)
    void stage2();

    // stage 3: accumulate + reductions (reads the task count -> one stream sync); window sums go to the host.
    // `s3` (optional) = a different stream for this stage: the prover keeps the H MSM's digits + sort on the
    // high-priority transform stream (so the sort is not starved by the witness MSMs' accumulations) and its
    // accumulation on the lowest-priority stream (so it fills the machine under everything else).
    void stage3(cudaStream_t s3 = nullptr);

    // whole MSM on one stream; returns after the window sums are in host memory
    void run(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s) {
        stage1(bases_, scalars, inf_mask, n_, s);
        stage2();
        stage3();
        G16_CUDA(cudaStreamSynchronize(s));
    }

    // sum_w 2^(c w) S_w on the host (255 dependent doublings: a latency chain, an order of magnitude quicker on a
    // CPU core than on one GPU thread).  HC = g16h::HG1 / HG2, whose XYZZ layout equals the device one.
    // Valid after the stream is synchronised.
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
    uint32_t T = 0, L = 0, L1 = 4;

private:
    static int acc_minb() {   // 4 = the 128-register build of the G1 accumulation (4 CTAs/SM; measured 1-2 % faster than
                              // the 148-register one); G16_ACC_MINB=1 selects the other for comparison
        static int v = [] { const char* e = std::getenv("G16_ACC_MINB"); return e ? std::atoi(e) : 4; }();
        return v;
    }
    static bool merge_enabled() {   // G16_PAIR_MERGE=1: batched-affine pair round in front of the accumulation
        static bool v = [] { const char* e = std::getenv("G16_PAIR_MERGE"); return e && std::atoi(e) != 0; }();
        return v;
    }
    static size_t merge_min() {     // fewest entries for which the pair round pays (G16_PAIR_MERGE_MIN overrides; tests use 0)
        static size_t v = [] { const char* e = std::getenv("G16_PAIR_MERGE_MIN"); return e ? size_t(std::atoll(e)) : size_t(1) << 20; }();
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
    static void grow(DevBuf& b, size_t bytes) {
        if (bytes > b.bytes) b.alloc(bytes + bytes / 4 + 4096);
    }
    int key_bits() const {
        int b = 1;
        while ((size_t(1) << b) < nb) b++;
        return b;
    }
    MsmConfig cfg;
    size_t cap_n = 0, n = 0, m = 0, slots = 148 * 1024;
    uint32_t bpw = 0, nb = 0, cpw = 0, cpw1 = 0, ns0 = 0, ns1 = 0;
    const aff* bases = nullptr;
    cudaStream_t stream = 0;
    uint32_t *sorted_keys = nullptr, *sorted_vals = nullptr;
    DevBuf keys_a, keys_b, vals_a, vals_b, start, end, ntasks, nparts, nblk, task_base, part_base, blk_base, cut_list,
        t_begin, t_len, t_out, tk_a, tk_b, tv_a, tv_b, buckets, parts, parts2, merged, prefix, chunk_out, chunk_sum, chunk2_out, slice_out, slice1_out,
        window_out, counters, cub_tmp,
        task_tmp, table;
    size_t table_n = 0;
    unsigned long long* h_counters = nullptr;
    pt* h_windows = nullptr;
    cudaEvent_t ev_mrg0 = nullptr, ev_mrg1 = nullptr;
    uint32_t merged_slots = 0;
    cudaEvent_t ev_acc0 = nullptr, ev_acc1 = nullptr, ev_t[4] = {nullptr, nullptr, nullptr, nullptr};
    bool acc_timed = false;
};

#ifdef G16_MSM_INSTANTIATE
// ---- out-of-class definitions: compiled in msm.cu only, so that the kernels exist once in the library
template <class C>
void Msm<C>::reserve(size_t n, const MsmConfig& cfg_in) {
    cfg = cfg_in;
    cap_n = n;
    c = cfg.c ? cfg.c : cfg.table ? msm_pick_c_table(n) : cfg.sparse ? msm_pick_c_sparse(n) : msm_pick_c(n);
    if (c < 2) c = 2;
    if (c > 24) c = 24;
    if (cfg.table && c < 6) c = 6;
    Wd = msm_num_windows(c);
    W = cfg.table ? 1 : Wd;
    bpw = 1u << (c - 1);
    nb = uint32_t(W) * bpw;
    size_t m_max = n * size_t(Wd);
    if (m_max >= (size_t(1) << 32)) throw std::runtime_error("MSM too large for 32-bit entry indices; shard it");
    if (cfg.table && m_max > size_t(MSM_INDEX_MASK)) throw std::runtime_error("MSM window table too large for 30-bit base indices; shard it");
    // level-0 chunks: 16 buckets when there are enough buckets to fill the machine with chunk threads, else 8
    // (half the dependency chain); level 1 works on bpw/L points only, so its chunks are short (latency, not work)
    L = cfg.L ? cfg.L : (nb >= (2u << 20) ? 16 : 8);
    L1 = 4;
    cpw = (bpw + L - 1) / L;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    slots = size_t(sms) * (sizeof(pt) > 128 ? 512 : 1024);
    size_t e = m_max ? m_max : 1;
    keys_a.alloc(e * 4); keys_b.alloc(e * 4); vals_a.alloc(e * 4); vals_b.alloc(e * 4);
    start.alloc(size_t(nb) * 4); end.alloc(size_t(nb) * 4);
    ntasks.alloc(size_t(nb + 1) * 4); nparts.alloc(size_t(nb + 1) * 4); nblk.alloc(size_t(nb + 1) * 4);
    task_base.alloc(size_t(nb + 1) * 4); part_base.alloc(size_t(nb + 1) * 4); blk_base.alloc(size_t(nb + 1) * 4);
    cut_list.alloc((m_max / T_MIN + 2) * 4);  // a cut bucket holds more than T >= T_MIN entries
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
    counters.alloc(64);
    size_t t1 = 0, t3 = 0;
    cub::DoubleBuffer<uint32_t> dk(keys_a.as<uint32_t>(), keys_b.as<uint32_t>());
    cub::DoubleBuffer<uint32_t> dv(vals_a.as<uint32_t>(), vals_b.as<uint32_t>());
    cub::DeviceRadixSort::SortPairs(nullptr, t1, dk, dv, (int)e, 0, key_bits());
    cub::DeviceScan::ExclusiveSum(nullptr, t3, ntasks.as<uint32_t>(), task_base.as<uint32_t>(), (int)(nb + 1));
    cub_tmp.alloc((t1 > t3 ? t1 : t3) + 256);
    if (!h_counters) G16_CUDA(cudaMallocHost(&h_counters, 64));
    if (!h_windows) G16_CUDA(cudaMallocHost(&h_windows, 32 * sizeof(pt) * 8));
    if (size_t(W) > 256) throw std::runtime_error("MSM: too many windows");
    if (!ev_acc0) G16_CUDA(cudaEventCreate(&ev_acc0));
    if (!ev_acc1) G16_CUDA(cudaEventCreate(&ev_acc1));
    if (!ev_mrg0) G16_CUDA(cudaEventCreate(&ev_mrg0));
    if (!ev_mrg1) G16_CUDA(cudaEventCreate(&ev_mrg1));
    for (auto& e : ev_t)
        if (!e) G16_CUDA(cudaEventCreate(&e));
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
void Msm<C>::stage1(const aff* bases_, const uint32_t* scalars, const uint32_t* inf_mask, size_t n_, cudaStream_t s) {
    if (n_ > cap_n) throw std::runtime_error("MSM: n exceeds reserved capacity");
    if (cfg.table && n_ != table_n) throw std::runtime_error("MSM: window table was built for a different base count");
    bases = cfg.table && n_ ? table.template as<aff>() : bases_;
    n = n_;
    stream = s;
    G16_CUDA(cudaEventRecord(ev_t[0], s));
    G16_CUDA(cudaMemsetAsync(counters.p, 0, 64, s));
    if (n) {
        msm_digits_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(scalars, (uint32_t)n, inf_mask, c, Wd,
                                                                      cfg.table ? (uint32_t)n : 0u,
                                                                      keys_a.as<uint32_t>(), vals_a.as<uint32_t>(),
                                                                      counters.as<unsigned long long>());
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(1);
    }
    acc_timed = false;
    G16_CUDA(cudaEventRecord(ev_t[1], s));
    G16_CUDA(cudaMemcpyAsync(h_counters, counters.p, 16, cudaMemcpyDeviceToHost, s));
}

template <class C>
void Msm<C>::stage2() {
    cudaStream_t s = stream;
    G16_CUDA(cudaStreamSynchronize(s));
    m = (size_t)h_counters[0];
    T = cfg.T;
    if (T == 0) {
        // enough tasks to fill the machine several times over: a task is one thread's dependency chain, so its
        // length bounds the kernel's tail; a cut bucket costs one extra full addition per task
        size_t t = (merge_enabled() ? m / 2 : m) / slots;
        T = uint32_t(t < T_MIN ? T_MIN : t > T_MAX ? T_MAX : t);
    }
    if (T < T_MIN) T = T_MIN;
    stats = MsmStats();
    stats.entries = m; stats.c = c; stats.W = Wd; stats.buckets = nb; stats.T = T;
    G16_CUDA(cudaMemsetAsync(buckets.p, 0, size_t(nb) * sizeof(pt), s));
    if (m == 0) {
        G16_CUDA(cudaEventRecord(ev_t[2], s));
        return;
    }
    cub::DoubleBuffer<uint32_t> dk(keys_a.as<uint32_t>(), keys_b.as<uint32_t>());
    cub::DoubleBuffer<uint32_t> dv(vals_a.as<uint32_t>(), vals_b.as<uint32_t>());
    size_t tb = cub_tmp.bytes;
    G16_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, dk, dv, (int)m, 0, key_bits(), s));
    sorted_keys = dk.Current();
    sorted_vals = dv.Current();
    G16_CUDA(cudaMemsetAsync(start.p, 0, size_t(nb) * 4, s));
    G16_CUDA(cudaMemsetAsync(end.p, 0, size_t(nb) * 4, s));
    msm_bounds_kernel<<<(unsigned)((m + 255) / 256), 256, 0, s>>>(sorted_keys, (uint32_t)m, start.as<uint32_t>(),
                                                                  end.as<uint32_t>());
    merged_slots = 0;
    if (merge_enabled() && m >= merge_min() && ((m + 1) / 2) * (sizeof(aff) + sizeof(fe)) <= (size_t(24) << 30)) {
        // one round of batched-affine pair additions in front of the accumulation (see msm_pair_merge_kernel)
        uint32_t nslots = uint32_t((m + 1) / 2);
        grow(merged, size_t(nslots) * sizeof(aff));
        grow(prefix, size_t(nslots) * sizeof(fe));
        uint32_t want = uint32_t(slots / 2);                 // threads that fill the machine
        uint32_t per = (nslots + want - 1) / want;
        per = per < 32 ? 32 : per > 256 ? 256 : per;         // slots per thread = the inversion's batch
        uint32_t blocks = ((nslots + per - 1) / per + 127) / 128;
        per = (nslots + blocks * 128 - 1) / (blocks * 128);
        G16_CUDA(cudaEventRecord(ev_mrg0, s));
        msm_pair_merge_kernel<C><<<blocks, 128, 0, s>>>(bases, sorted_keys, sorted_vals, start.as<uint32_t>(), (uint32_t)m,
                                                        nslots, per, prefix.as<fe>(), merged.as<aff>(), dv.Alternate());
        G16_CUDA(cudaEventRecord(ev_mrg1, s));
        msm_merged_end_kernel<<<(nb + 255) / 256, 256, 0, s>>>(start.as<uint32_t>(), end.as<uint32_t>(), nb);
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(2);
        sorted_vals = dv.Alternate();
        merged_slots = nslots;
    }
    uint32_t* cut_count = counters.as<uint32_t>() + 4;  // byte offset 16
    msm_count_tasks_kernel<<<(nb + 1 + 255) / 256, 256, 0, s>>>(start.as<uint32_t>(), end.as<uint32_t>(), nb, T,
                                                                ntasks.as<uint32_t>(), nparts.as<uint32_t>(),
                                                                nblk.as<uint32_t>(), cut_list.as<uint32_t>(), cut_count);
    G16_LAUNCHED(2);
    uint32_t* src[3] = {ntasks.as<uint32_t>(), nparts.as<uint32_t>(), nblk.as<uint32_t>()};
    uint32_t* dst[3] = {task_base.as<uint32_t>(), part_base.as<uint32_t>(), blk_base.as<uint32_t>()};
    uint32_t* h32 = reinterpret_cast<uint32_t*>(h_counters);
    for (int i = 0; i < 3; i++) {
        tb = cub_tmp.bytes;
        G16_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, src[i], dst[i], (int)(nb + 1), s));
        // h32[4] = tasks, [5] = partial sums, [6] = combine blocks, [7] = cut buckets
        G16_CUDA(cudaMemcpyAsync(h32 + 4 + i, dst[i] + nb, 4, cudaMemcpyDeviceToHost, s));
    }
    G16_CUDA(cudaMemcpyAsync(h32 + 7, cut_count, 4, cudaMemcpyDeviceToHost, s));
    G16_CUDA(cudaEventRecord(ev_t[2], s));
}

template <class C>
void Msm<C>::stage3(cudaStream_t s3) {
    cudaStream_t s = s3 ? s3 : stream;
    if (s != stream) G16_CUDA(cudaStreamWaitEvent(s, ev_t[2], 0));
    if (m != 0) {
        G16_CUDA(cudaStreamSynchronize(stream));
        const uint32_t* h32 = reinterpret_cast<const uint32_t*>(h_counters);
        uint32_t nt = h32[4], np = h32[5], nblocks = h32[6], ncut = h32[7];
        stats.tasks = nt;
        stats.cut_buckets = ncut;
        // the stream is idle here, so growing these is safe
        grow(t_begin, size_t(nt) * 4); grow(t_len, size_t(nt) * 4); grow(t_out, size_t(nt) * 4);
        grow(tk_a, size_t(nt) * 4); grow(tk_b, size_t(nt) * 4); grow(tv_a, size_t(nt) * 4); grow(tv_b, size_t(nt) * 4);
        grow(parts, size_t(np) * sizeof(pt));
        grow(parts2, size_t(nblocks) * sizeof(pt));
        cub::DoubleBuffer<uint32_t> tk(tk_a.as<uint32_t>(), tk_b.as<uint32_t>());
        cub::DoubleBuffer<uint32_t> tv(tv_a.as<uint32_t>(), tv_b.as<uint32_t>());
        int bits = 1;
        while ((1u << bits) <= T) bits++;
        size_t need = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, need, tk, tv, (int)nt, 0, bits);
        grow(task_tmp, need + 256);
        msm_make_tasks_kernel<<<(nt + 255) / 256, 256, 0, s>>>(
            start.as<uint32_t>(), end.as<uint32_t>(), task_base.as<uint32_t>(), part_base.as<uint32_t>(), nb, nt, T,
            t_begin.as<uint32_t>(), t_len.as<uint32_t>(), t_out.as<uint32_t>(), tk_a.as<uint32_t>(),
            tv_a.as<uint32_t>());
        size_t tb = task_tmp.bytes;
        G16_CUDA(cub::DeviceRadixSort::SortPairs(task_tmp.p, tb, tk, tv, (int)nt, 0, bits, s));
        G16_CUDA(cudaEventRecord(ev_acc0, s));
        bool launched = false;
        if constexpr (sizeof(pt) <= 128) {
            if (acc_minb() == 4) {
                msm_accumulate_kernel<C, 4><<<(nt + 127) / 128, 128, 0, s>>>(
                    bases, merged.as<aff>(), sorted_vals, tv.Current(), t_begin.as<uint32_t>(), t_len.as<uint32_t>(), t_out.as<uint32_t>(),
                    nt, buckets.as<pt>(), parts.as<pt>());
                launched = true;
            }
        }
        if (!launched)
            msm_accumulate_kernel<C, 1><<<(nt + 127) / 128, 128, 0, s>>>(
                bases, merged.as<aff>(), sorted_vals, tv.Current(), t_begin.as<uint32_t>(), t_len.as<uint32_t>(), t_out.as<uint32_t>(), nt,
                buckets.as<pt>(), parts.as<pt>());
        G16_CUDA(cudaEventRecord(ev_acc1, s));
        acc_timed = true;
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(2);
        if (ncut) {
            msm_combine_small_kernel<C><<<(ncut + 127) / 128, 128, 0, s>>>(
                cut_list.as<uint32_t>(), ncut, part_base.as<uint32_t>(), parts.as<pt>(), buckets.as<pt>());
            G16_LAUNCHED(1);
            if (nblocks) {
                msm_combine1_kernel<C, RED_THREADS><<<nblocks, RED_THREADS, 0, s>>>(
                    blk_base.as<uint32_t>(), part_base.as<uint32_t>(), nb, parts.as<pt>(), parts2.as<pt>());
                msm_combine2_kernel<C, RED_THREADS><<<ncut, RED_THREADS, 0, s>>>(
                    cut_list.as<uint32_t>(), blk_base.as<uint32_t>(), parts2.as<pt>(), buckets.as<pt>());
                G16_LAUNCHED(2);
            }
            G16_CUDA(cudaGetLastError());
        }
    }
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
    G16_CUDA(cudaEventRecord(ev_t[3], s));
    stream = s;
}
#endif  // G16_MSM_INSTANTIATE

#ifndef G16_MSM_INSTANTIATE
extern template class Msm<G1>;
extern template class Msm<G2>;
#endif

}  // namespace g16
