// Stand-alone operators of the C ABI (MSM, odd-coset NTT), the field/curve micro-operators the parity tests
// drive, and the integer-pipe microbenchmark that gives the MSM roofline its denominator.
#include <cstring>
#include <memory>
#include <vector>

#include "../../include/g16_b200.h"
#include "internal.cuh"

using namespace g16;
using namespace g16h;

struct g16_msm {
    int is_g2 = 0;
    int device = 0;
    size_t max_n = 0;
    Msm<G1> m1;
    Msm<G2> m2;
    cudaStream_t s = nullptr;
    DevBuf mask;  // infinity bitmask of the last base array seen (cached by pointer and length)
    const void* mask_for = nullptr;
    size_t mask_n = 0;
    ~g16_msm() {
        if (s) cudaStreamDestroy(s);
    }
};

namespace {

template <class F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const std::exception& e) {
        std::string m = e.what();
        return fail(m.rfind("CUDA error", 0) == 0 ? G16_ERR_CUDA : G16_ERR_GENERIC, m);
    }
}

void use_device(const g16_opts* o) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw std::runtime_error("CUDA error: no CUDA device: libg16b200 has no CPU fallback");
    if (o && o->device >= 0) G16_CUDA(cudaSetDevice(o->device));
}

void g1_out(const HPoint<HFq>& p, uint8_t* out) {
    HAffine<HFq> a = HG1::to_affine(p);
    std::memset(out, 0, 64);
    if (!a.inf) {
        u256_to_bytes(a.x, out);
        u256_to_bytes(a.y, out + 32);
    }
}
void g2_out(const HPoint<HFq2>& p, uint8_t* out) {
    HAffine<HFq2> a = HG2::to_affine(p);
    std::memset(out, 0, 128);
    if (!a.inf) {
        u256_to_bytes(a.x.c0, out);
        u256_to_bytes(a.x.c1, out + 32);
        u256_to_bytes(a.y.c0, out + 64);
        u256_to_bytes(a.y.c1, out + 96);
    }
}
void g1_out(const g1_pt* r, uint8_t* out) {
    HPoint<HFq> p;
    std::memcpy(&p, r, sizeof(p));
    g1_out(p, out);
}
void g2_out(const g2_pt* r, uint8_t* out) {
    HPoint<HFq2> p;
    std::memcpy(&p, r, sizeof(p));
    g2_out(p, out);
}

// ---- micro-operator kernels ---------------------------------------------------------------------------------
template <class F>
__global__ void field_op_kernel(int op, const fe_t* a, const fe_t* b, fe_t* o, size_t n) {
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    fe_t x = ld_fe(a + i), y = b ? ld_fe(b + i) : F::zero(), r;
    switch (op) {
        case 0: r = F::mul(x, y); break;
        case 1: r = F::add(x, y); break;
        case 2: r = F::sub(x, y); break;
        case 3: r = F::to_mont(x); break;
        case 4: r = F::from_mont(x); break;
        case 6: r = F::sqr_dedicated(x); break;
        default: r = F::inv(x); break;
    }
    st_fe(o + i, r);
}

// op 0: one thread walks all points with mixed additions
template <class C>
__global__ void curve_chain_kernel(const typename C::aff* pts, size_t n, typename C::pt* out) {
    if (blockIdx.x || threadIdx.x) return;
    typename C::pt acc = C::identity();
    for (size_t i = 0; i < n; i++) {
        typename C::aff q = pts[i];
        if (C::is_inf(q)) continue;
        C::madd(acc, q);
    }
    *out = acc;
}
// op 1: block tree of full additions (128 threads)
template <class C>
__global__ void curve_tree_kernel(const typename C::aff* pts, size_t n, typename C::pt* out) {
    typename C::pt acc = C::identity();
    for (size_t i = threadIdx.x; i < n; i += 128) {
        typename C::pt p = C::from_affine(pts[i]);
        C::add(acc, p);
    }
    block_sum_points<C, 128>(acc);
    if (threadIdx.x == 0) *out = acc;
}
// op 2: doubling, one thread per point, through both dbl paths
template <class C>
__global__ void curve_dbl_kernel(const typename C::aff* pts, size_t n, typename C::pt* out) {
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    typename C::aff q = pts[i];
    if (C::is_inf(q)) {
        out[i] = C::identity();
        return;
    }
    typename C::pt a = C::dbl_affine(q);        // 2P
    typename C::pt b = C::dbl(C::from_affine(q));  // 2P again through the XYZZ doubling
    typename C::pt nb = C::neg(b);
    C::add(nb, a);                               // must cancel to the identity
    if (!C::is_identity(nb)) a = C::identity();  // poison the result if the two paths disagree
    out[i] = a;
}

template <class C>
int curve_op(int op, const uint8_t* points, size_t n, uint8_t* out, void (*conv)(const typename C::pt*, uint8_t*),
             size_t affsz) {
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    DevBuf dp, dout;
    dp.alloc(n * sizeof(aff));
    size_t nout = op == 2 ? n : 1;
    dout.alloc(nout * sizeof(pt));
    G16_CUDA(cudaMemcpy(dp.p, points, n * sizeof(aff), cudaMemcpyHostToDevice));
    if (op == 0) curve_chain_kernel<C><<<1, 32>>>(dp.as<aff>(), n, dout.as<pt>());
    else if (op == 1) {
        curve_tree_kernel<C><<<1, 128>>>(dp.as<aff>(), n, dout.as<pt>());
    } else curve_dbl_kernel<C><<<(unsigned)((n + 63) / 64), 64>>>(dp.as<aff>(), n, dout.as<pt>());
    G16_CUDA(cudaGetLastError());
    std::vector<pt> h(nout);
    G16_CUDA(cudaMemcpy(h.data(), dout.p, nout * sizeof(pt), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < nout; i++) conv(&h[i], out + i * affsz);
    return G16_OK;
}

// ---- batch apply key (operator level) ---------------------------------------------------
// out[i] = (first * inc^i) * P_i -- ffjavascript 0.2.62 engine_applykey.js `G.batchApplyKey(buff, first, inc)`, the
// kernel of `snarkjs zkey contribute` / `zkey beacon` (sections 8 and 9 times 1/delta', scripts/g16_setup.sh:240-283)
// and of the powers-of-tau steps.  One thread per point: its scalar from the binary powers of inc (the table trick of
// ntt_powers_kernel), 254 doublings + a mixed addition per set bit in XYZZ, one Fermat inversion back to affine.
template <class C>
__global__ void __launch_bounds__(128)
apply_key_kernel(const typename C::aff* __restrict__ pts, size_t n, fe_t first_m, const fe_t* __restrict__ pw, int bits,
                 typename C::aff* __restrict__ out) {
    typedef typename C::field F;
    typedef typename C::fe fe;
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    aff q = pts[i];
    if (C::is_inf(q)) {
        out[i] = q;
        return;
    }
    fe_t k = first_m;
    for (int b = 0; b < bits; b++)
        if ((i >> b) & 1) k = Fr::mul(k, pw[b]);
    k = Fr::from_mont(k);
    pt acc = C::identity();
    for (int bit = 253; bit >= 0; bit--) {
        acc = C::dbl(acc);
        if ((k.v[bit >> 5] >> (bit & 31)) & 1) C::madd(acc, q);
    }
    aff r;
    if (C::is_identity(acc)) {
        r.x = F::zero();
        r.y = F::zero();
    } else {
        fe u = F::inv(acc.zzz);           // 1 / Z^3
        fe zi = F::mul(u, acc.zz);        // 1 / Z
        r.x = F::mul(acc.x, F::sqr(zi));
        r.y = F::mul(acc.y, u);
    }
    out[i] = r;
}

static bool scalar_from_bytes(const uint8_t* b, U256& out) {
    out = u256_from_bytes(b);
    U256 m = HFr::mod();
    for (int i = 3; i >= 0; i--) {
        if (out.v[i] != m.v[i]) return out.v[i] < m.v[i];
    }
    return false;  // == r
}

static int apply_key_dev(int is_g2, const void* d_points, size_t n, const uint8_t* first, const uint8_t* inc, void* d_out,
                         cudaStream_t s) {
    U256 f, c;
    if (!scalar_from_bytes(first, f) || !scalar_from_bytes(inc, c)) return fail(G16_ERR_GENERIC, "apply key: scalar not below r");
    if (!n) return (int)G16_OK;
    fe_t pw[64], first_m;
    U256 fm = HFr::to_mont(f), a = HFr::to_mont(c);
    std::memcpy(&first_m, fm.v, 32);
    int bits = 0;
    while (bits < 64 && (n - 1) >> bits) bits++;
    for (int b = 0; b < 64; b++) {
        std::memcpy(&pw[b], a.v, 32);
        a = HFr::sqr(a);
    }
    DevBuf dpw;
    dpw.alloc(sizeof(pw));
    G16_CUDA(cudaMemcpyAsync(dpw.p, pw, sizeof(pw), cudaMemcpyHostToDevice, s));
    unsigned grid = (unsigned)((n + 127) / 128);
    if (is_g2)
        apply_key_kernel<G2><<<grid, 128, 0, s>>>((const G2::aff*)d_points, n, first_m, dpw.as<fe_t>(), bits, (G2::aff*)d_out);
    else
        apply_key_kernel<G1><<<grid, 128, 0, s>>>((const G1::aff*)d_points, n, first_m, dpw.as<fe_t>(), bits, (G1::aff*)d_out);
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(1);
    G16_CUDA(cudaStreamSynchronize(s));
    return (int)G16_OK;
}

// ---- integer-pipe microbenchmark ------------------------------------------------------------------------------
// 8 independent accumulators per thread, long unrolled loops; results folded into one store so nothing is elided.
__global__ void int_pipe_kernel(int op, int iters, uint32_t seed, uint32_t* out) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (op == 0) {
        uint32_t a[8];
        for (int k = 0; k < 8; k++) a[k] = seed + t * 9 + k;
        uint32_t m = seed | 1, c = t;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 8; u++) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = a[k] * m + c;  // IMAD
            }
        }
        uint32_t r = 0;
        for (int k = 0; k < 8; k++) r ^= a[k];
        out[t] = r;
    } else if (op == 1) {
        uint32_t acc[8];
        for (int k = 0; k < 8; k++) acc[k] = seed + t + k;
        uint32_t x0 = seed * 3 + t, x1 = x0 ^ 0x9e3779b9u, x2 = x1 + 77, x3 = x2 * 5, y = seed + 12345;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 16; u++) {
                uint32_t c = mad_row_c(acc, x0, x1, x2, x3, y);  // 4 IMAD.WIDE.U32(.X) + carry capture
                y += c + 1;
            }
        }
        uint32_t r = 0;
        for (int k = 0; k < 8; k++) r ^= acc[k];
        out[t] = r;
    } else if (op == 3) {
        // four INDEPENDENT carry-chain accumulators per thread (op 1 has one: each row waits for the carry of the one
        // before), i.e. the instruction-level parallelism a Montgomery product offers the scheduler
        uint32_t a0[8], a1[8], a2[8], a3[8];
        for (int k = 0; k < 8; k++) {
            a0[k] = seed + t + k;
            a1[k] = seed * 3 + t + k;
            a2[k] = (seed ^ 0x55aa55aau) + t * 7 + k;
            a3[k] = seed + t * 13 + k;
        }
        uint32_t x0 = seed * 3 + t, x1 = x0 ^ 0x9e3779b9u, x2 = x1 + 77, x3 = x2 * 5;
        uint32_t y0 = seed + 12345, y1 = seed + 777, y2 = seed * 9 + 1, y3 = seed ^ 0xdeadbeefu;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                uint32_t c0 = mad_row_c(a0, x0, x1, x2, x3, y0);
                uint32_t c1 = mad_row_c(a1, x1, x2, x3, x0, y1);
                uint32_t c2 = mad_row_c(a2, x2, x3, x0, x1, y2);
                uint32_t c3 = mad_row_c(a3, x3, x0, x1, x2, y3);
                y0 += c0 + 1;
                y1 += c1 + 1;
                y2 += c2 + 1;
                y3 += c3 + 1;
            }
        }
        uint32_t r = 0;
        for (int k = 0; k < 8; k++) r ^= a0[k] ^ a1[k] ^ a2[k] ^ a3[k];
        out[t] = r;
    } else {
        fe_t a = Fq::one(), b = Fq::r2();
        a.v[0] ^= t;
        b.v[1] ^= seed;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                a = Fq::mul(a, b);
                b = Fq::mul(b, a);
            }
        }
        out[t] = a.v[0] ^ b.v[7];
    }
}

}  // namespace

extern "C" {

int g16_msm_create(int is_g2, size_t max_n, const g16_opts* opts, g16_msm** out) {
    return guarded([&]() {
        if (!out) return fail(G16_ERR_GENERIC, "null argument");
        use_device(opts);
        std::unique_ptr<g16_msm> m(new g16_msm());
        m->is_g2 = is_g2;
        m->max_n = max_n;
        G16_CUDA(cudaGetDevice(&m->device));
        G16_CUDA(cudaStreamCreateWithFlags(&m->s, cudaStreamNonBlocking));
        MsmConfig cfg;
        if (opts) cfg.c = opts->msm_c;
        if (is_g2) {
            Msm<G2>::prepare();
            m->m2.reserve(max_n ? max_n : 1, cfg);
        } else {
            Msm<G1>::prepare();
            m->m1.reserve(max_n ? max_n : 1, cfg);
        }
        *out = m.release();
        return (int)G16_OK;
    });
}

void g16_msm_destroy(g16_msm* m) { delete m; }

// test hook: the sizes and limits Msm::reserve derives from (n, window bits, table, sparse) -- host arithmetic only, works
// without a device.  G16_ERR_GENERIC + message ("... shard it") beyond the 32-bit entry / 31-bit base index / int item limits.
int g16_test_msm_shape(uint64_t n, int c, int table, int sparse, uint64_t* entries, int* c_out, int* windows, int* dense) {
    return guarded([&]() {
        MsmConfig cfg;
        cfg.c = c;
        cfg.table = table != 0;
        cfg.sparse = sparse != 0;
        MsmShape sh = msm_shape(size_t(n), cfg);
        if (entries) *entries = sh.m_max;
        if (c_out) *c_out = sh.c;
        if (windows) *windows = sh.Wd;
        if (dense) *dense = sh.dense ? 1 : 0;
        return (int)G16_OK;
    });
}

int g16_msm_run_dev(g16_msm* m, const void* d_bases, const void* d_scalars, size_t n, uint8_t* out_affine) {
    return guarded([&]() {
        if (!m || !out_affine) return fail(G16_ERR_GENERIC, "null argument");
        G16_CUDA(cudaSetDevice(m->device));
        if (n > m->max_n) return fail(G16_ERR_GENERIC, "n exceeds the workspace capacity");
        bool fresh = m->mask_for != d_bases || m->mask_n != n;
        if (fresh) {
            m->mask.alloc(((n + 31) / 32 + 1) * 4);
            m->mask_for = d_bases;
            m->mask_n = n;
        }
        unsigned grid = (unsigned)((n + 255) / 256);
        if (m->is_g2) {
            if (fresh && n) msm_inf_mask_kernel<G2><<<grid, 256, 0, m->s>>>((const g2_aff*)d_bases, (uint32_t)n, m->mask.as<uint32_t>());
            m->m2.run((const g2_aff*)d_bases, (const uint32_t*)d_scalars, m->mask.as<uint32_t>(), n, m->s);
            g2_out(m->m2.finish<HG2>(), out_affine);
        } else {
            if (fresh && n) msm_inf_mask_kernel<G1><<<grid, 256, 0, m->s>>>((const g1_aff*)d_bases, (uint32_t)n, m->mask.as<uint32_t>());
            m->m1.run((const g1_aff*)d_bases, (const uint32_t*)d_scalars, m->mask.as<uint32_t>(), n, m->s);
            g1_out(m->m1.finish<HG1>(), out_affine);
        }
        return (int)G16_OK;
    });
}

int g16_msm_stats(const g16_msm* m, uint64_t out[6]) {
    if (!m || !out) return fail(G16_ERR_GENERIC, "null argument");
    const MsmStats& s = m->is_g2 ? m->m2.stats : m->m1.stats;
    out[0] = s.entries; out[1] = s.tasks; out[2] = s.cut_buckets; out[3] = s.buckets; out[4] = s.c; out[5] = s.W;
    return G16_OK;
}

static int msm_host(int is_g2, const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t* out,
                    const g16_opts* opts) {
    return guarded([&]() {
        if (!out || (n && (!bases || !scalars))) return fail(G16_ERR_GENERIC, "null argument");
        use_device(opts);
        size_t rec = is_g2 ? 128 : 64;
        DevBuf db, ds, mask;
        db.alloc(n * rec);
        ds.alloc(n * 32);
        mask.alloc(((n + 31) / 32 + 1) * 4);
        cudaStream_t s;
        G16_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        MsmConfig cfg;
        if (opts) cfg.c = opts->msm_c;
        if (n) {
            G16_CUDA(cudaMemcpyAsync(db.p, bases, n * rec, cudaMemcpyHostToDevice, s));
            G16_CUDA(cudaMemcpyAsync(ds.p, scalars, n * 32, cudaMemcpyHostToDevice, s));
        }
        int rc = G16_OK;
        try {
            if (is_g2) {
                Msm<G2>::prepare();
                Msm<G2> m;
                m.reserve(n ? n : 1, cfg);
                if (n) msm_inf_mask_kernel<G2><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(db.as<g2_aff>(), (uint32_t)n, mask.as<uint32_t>());
                m.run(db.as<g2_aff>(), ds.as<uint32_t>(), mask.as<uint32_t>(), n, s);
                g2_out(m.finish<HG2>(), out);
            } else {
                Msm<G1>::prepare();
                Msm<G1> m;
                m.reserve(n ? n : 1, cfg);
                if (n) msm_inf_mask_kernel<G1><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(db.as<g1_aff>(), (uint32_t)n, mask.as<uint32_t>());
                m.run(db.as<g1_aff>(), ds.as<uint32_t>(), mask.as<uint32_t>(), n, s);
                g1_out(m.finish<HG1>(), out);
            }
        } catch (...) {
            cudaStreamDestroy(s);
            throw;
        }
        cudaStreamDestroy(s);
        return rc;
    });
}

int g16_msm_g1_host(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[64], const g16_opts* opts) {
    return msm_host(0, bases, scalars, n, out, opts);
}
int g16_msm_g2_host(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[128], const g16_opts* opts) {
    return msm_host(1, bases, scalars, n, out, opts);
}

int g16_coset_ntt_dev(void* d_data, int logn, int count) {
    return guarded([&]() {
        if (!d_data) return fail(G16_ERR_GENERIC, "null argument");
        static thread_local std::unique_ptr<NttDomain> dom;  // tables cached per thread for repeated calls
        if (!dom || dom->logn != logn) {
            dom.reset(new NttDomain());
            dom->init(logn, 0);
        }
        for (int i = 0; i < count; i++) dom->to_odd_coset((fe_t*)d_data + (size_t(i) << logn), 0);
        G16_CUDA(cudaStreamSynchronize(0));
        return (int)G16_OK;
    });
}

int g16_coset_ntt_host(uint8_t* data, int logn) {
    return guarded([&]() {
        if (!data) return fail(G16_ERR_GENERIC, "null argument");
        use_device(nullptr);
        size_t bytes = (size_t(1) << logn) * 32;
        DevBuf d;
        d.alloc(bytes);
        G16_CUDA(cudaMemcpy(d.p, data, bytes, cudaMemcpyHostToDevice));
        NttDomain dom;
        dom.init(logn, 0);
        dom.to_odd_coset(d.as<fe_t>(), 0);
        G16_CUDA(cudaMemcpy(data, d.p, bytes, cudaMemcpyDeviceToHost));
        return (int)G16_OK;
    });
}

int g16_batch_apply_key_dev(int is_g2, const void* d_points, size_t n, const uint8_t first[32], const uint8_t inc[32],
                            void* d_out) {
    return guarded([&]() {
        if ((!d_points || !d_out) && n) return fail(G16_ERR_GENERIC, "null argument");
        if (!first || !inc) return fail(G16_ERR_GENERIC, "null argument");
        return apply_key_dev(is_g2, d_points, n, first, inc, d_out, 0);
    });
}

int g16_batch_apply_key(int is_g2, const uint8_t* points, size_t n, const uint8_t first[32], const uint8_t inc[32],
                        uint8_t* out, const g16_opts* opts) {
    return guarded([&]() {
        if ((!points || !out) && n) return fail(G16_ERR_GENERIC, "null argument");
        if (!first || !inc) return fail(G16_ERR_GENERIC, "null argument");
        use_device(opts);
        size_t rec = is_g2 ? 128 : 64;
        DevBuf d;
        d.alloc(n * rec);
        G16_CUDA(cudaMemcpy(d.p, points, n * rec, cudaMemcpyHostToDevice));
        int rc = apply_key_dev(is_g2, d.p, n, first, inc, d.p, 0);   // in place: a thread reads its record before it writes it
        if (rc != G16_OK) return rc;
        G16_CUDA(cudaMemcpy(out, d.p, n * rec, cudaMemcpyDeviceToHost));
        return (int)G16_OK;
    });
}

int g16_test_field_op(int field, int op, const uint8_t* a, const uint8_t* b, uint8_t* out, size_t n) {
    return guarded([&]() {
        use_device(nullptr);
        DevBuf da, db, dout;
        da.alloc(n * 32);
        db.alloc(n * 32);
        dout.alloc(n * 32);
        G16_CUDA(cudaMemcpy(da.p, a, n * 32, cudaMemcpyHostToDevice));
        if (b) G16_CUDA(cudaMemcpy(db.p, b, n * 32, cudaMemcpyHostToDevice));
        unsigned grid = (unsigned)((n + 127) / 128);
        if (field == 0) field_op_kernel<Fq><<<grid, 128>>>(op, da.as<fe_t>(), b ? db.as<fe_t>() : nullptr, dout.as<fe_t>(), n);
        else field_op_kernel<Fr><<<grid, 128>>>(op, da.as<fe_t>(), b ? db.as<fe_t>() : nullptr, dout.as<fe_t>(), n);
        G16_CUDA(cudaGetLastError());
        G16_CUDA(cudaMemcpy(out, dout.p, n * 32, cudaMemcpyDeviceToHost));
        return (int)G16_OK;
    });
}

int g16_test_curve_op(int is_g2, int op, const uint8_t* points, size_t n, uint8_t* out) {
    return guarded([&]() {
        use_device(nullptr);
        if (is_g2) return curve_op<G2>(op, points, n, out, g2_out, 128);
        return curve_op<G1>(op, points, n, out, g1_out, 64);
    });
}

int g16_bench_int_pipe(int op, double* ops_per_s) {
    return guarded([&]() {
        use_device(nullptr);
        cudaDeviceProp prop;
        int dev;
        G16_CUDA(cudaGetDevice(&dev));
        G16_CUDA(cudaGetDeviceProperties(&prop, dev));
        int blocks = prop.multiProcessorCount * 8, threads = 256;
        int iters = op == 2 ? 64 : 512;
        DevBuf out;
        out.alloc(size_t(blocks) * threads * 4);
        cudaEvent_t e0, e1;
        G16_CUDA(cudaEventCreate(&e0));
        G16_CUDA(cudaEventCreate(&e1));
        for (int w = 0; w < 2; w++) int_pipe_kernel<<<blocks, threads>>>(op, iters, 1234567u + w, out.as<uint32_t>());
        G16_CUDA(cudaEventRecord(e0));
        const int reps = 5;
        for (int w = 0; w < reps; w++) int_pipe_kernel<<<blocks, threads>>>(op, iters, 99u + w, out.as<uint32_t>());
        G16_CUDA(cudaEventRecord(e1));
        G16_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        // ops per thread: 0 = IMAD, 1 and 3 = wide multiply-adds (16 rows of 4 per iteration), 2 = Montgomery products
        double per_thread = op == 2 ? 8.0 * iters : 64.0 * iters;
        *ops_per_s = per_thread * blocks * threads * reps / (ms * 1e-3);
        return (int)G16_OK;
    });
}

}  // extern "C"
