// Groth16 prover: resident proving key, sparse R1CS evaluation, orchestration of the NTT pipeline and the five
// MSMs, final combination and the proof.json / public.json writers.  C ABI in include/g16_b200.h.
//
// Replaces: snarkjs 0.7.2 src/groth16_prove.js + zkey_utils.js + wtns_utils.js,
// rapidsnark Groth16::Prover::prove; reference call site scripts/g16_prove.sh:246-260.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/random.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <future>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "../../include/g16_b200.h"
#include "internal.cuh"

using namespace g16;
using namespace g16h;

namespace g16 {

// A proof is spread over a dozen streams (transforms, five MSMs x preparation / accumulation).  CUDA maps streams onto
// CUDA_DEVICE_MAX_CONNECTIONS hardware queues (default 8); streams that share a queue pick up false dependencies on each
// other (measured: the H preparation waited 5 ms behind another stream's accumulations, 14.6 -> 13.9 ms per proof at
// 2^21 once every stream has its own queue).  The variable is read when the CUDA context is created, so it is set when
// the library is loaded, without overriding the user's choice.  The extra queues cost context-creation time (measured:
// 0.8 -> 1.4 s), which a process that makes ONE proof never earns back: G16_CONNECTIONS_DEFAULTED tells such a caller (the
// one-shot command line, csrc/prover_main.cpp) that the value is this library's default and may be lowered again.
static const int g_max_connections_set = [] {
    if (std::getenv("CUDA_DEVICE_MAX_CONNECTIONS")) return 0;
    ::setenv("G16_CONNECTIONS_DEFAULTED", "1", 1);
    return ::setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 1);
}();

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

// ---------------------------------------------------------------------------------------------------------
// iden3 binfile container (row a1): magic[4] | u32 version | u32 nSections | { u32 type | u64 size | payload }*
// ---------------------------------------------------------------------------------------------------------
struct Section {
    const uint8_t* p = nullptr;
    uint64_t size = 0;
};

struct FormatError : std::runtime_error {
    explicit FormatError(const std::string& m) : std::runtime_error(m) {}
};
struct StatusError : std::runtime_error {
    int code;
    StatusError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

static inline uint32_t rd32(const uint8_t* p) {
    uint32_t v;
    std::memcpy(&v, p, 4);
    return v;
}
static inline uint64_t rd64(const uint8_t* p) {
    uint64_t v;
    std::memcpy(&v, p, 8);
    return v;
}

static std::map<uint32_t, Section> parse_binfile(const uint8_t* data, size_t len, const char magic[4], uint32_t version) {
    if (len < 12 || std::memcmp(data, magic, 4) != 0)
        throw FormatError(std::string("Invalid File format: expected magic \"") + std::string(magic, 4) + "\"");
    if (rd32(data + 4) > version) throw FormatError("Version not supported");
    uint32_t nsec = rd32(data + 8);
    std::map<uint32_t, Section> secs;
    size_t pos = 12;
    for (uint32_t i = 0; i < nsec; i++) {
        if (pos + 12 > len) throw FormatError("binfile: truncated section header");
        uint32_t typ = rd32(data + pos);
        uint64_t size = rd64(data + pos + 4);
        pos += 12;
        if (size > len - pos) throw FormatError("binfile: section runs past the end of the file");
        if (!secs.count(typ)) secs[typ] = Section{data + pos, size};
        pos += size;
    }
    return secs;
}

static const Section& need(const std::map<uint32_t, Section>& secs, uint32_t typ, const char* what) {
    auto it = secs.find(typ);
    if (it == secs.end()) throw FormatError(std::string(what) + ": missing section " + std::to_string(typ));
    return it->second;
}

// ---------------------------------------------------------------------------------------------------------
// sparse R1CS evaluation (row a4, buildABC1): rows of A and B in CSR form, built once at key load
// ---------------------------------------------------------------------------------------------------------
struct __align__(4) CoefRec {  // zkey section 4 record, 44 bytes, unaligned for 16-byte loads
    uint32_t m, c, s;
    uint32_t value[8];
};

// key = m * n + c
__global__ void r1cs_count_kernel(const uint8_t* __restrict__ recs, uint64_t ncoef, uint32_t n, uint32_t n_vars,
                                  uint32_t* __restrict__ counts, uint32_t* __restrict__ bad) {
    uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
    if (i >= ncoef) return;
    const uint32_t* r = reinterpret_cast<const uint32_t*>(recs + i * 44);
    uint32_t m = r[0], c = r[1], s = r[2];
    if (m > 1 || c >= n || s >= n_vars) {
        atomicAdd(bad, 1u);
        return;
    }
    atomicAdd(&counts[m * n + c], 1u);
}

__global__ void r1cs_fill_kernel(const uint8_t* __restrict__ recs, uint64_t ncoef, uint32_t n,
                                 const uint32_t* __restrict__ row_ptr, uint32_t* __restrict__ cursor,
                                 uint32_t* __restrict__ col, fe_t* __restrict__ val) {
    uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
    if (i >= ncoef) return;
    const uint32_t* r = reinterpret_cast<const uint32_t*>(recs + i * 44);
    uint32_t key = r[0] * n + r[1];
    uint32_t pos = row_ptr[key] + atomicAdd(&cursor[key], 1u);
    col[pos] = r[2];
    fe_t v;
#pragma unroll
    for (int k = 0; k < 8; k++) v.v[k] = r[3 + k];
    st_fe(val + pos, v);
}

// thread per constraint row c: A[c] = sum coef (x) w[s], B[c] likewise, C[c] = A[c]*B[c].
// coef is stored as coef*R^2 and w is the raw standard-form witness, so one Montgomery product gives coef*w in
// Montgomery form (snarkjs buildABC1 does the same with frm_mul on the raw bytes).
__global__ void __launch_bounds__(128)
r1cs_eval_kernel(const uint32_t* __restrict__ row_ptr, const uint32_t* __restrict__ col, const fe_t* __restrict__ val,
                 const fe_t* __restrict__ w, uint32_t n, fe_t* __restrict__ A, fe_t* __restrict__ B,
                 fe_t* __restrict__ C) {
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    fe_t a = Fr::zero(), b = Fr::zero();
    for (uint32_t p = row_ptr[c], e = row_ptr[c + 1]; p < e; p++)
        a = Fr::add(a, Fr::mul(ld_fe(val + p), ldg_fe(w + col[p])));
    for (uint32_t p = row_ptr[n + c], e = row_ptr[n + c + 1]; p < e; p++)
        b = Fr::add(b, Fr::mul(ld_fe(val + p), ldg_fe(w + col[p])));
    st_fe(A + c, a);
    st_fe(B + c, b);
    st_fe(C + c, Fr::mul(a, b));
}

}  // namespace g16

// ---------------------------------------------------------------------------------------------------------
// the resident proving key
// ---------------------------------------------------------------------------------------------------------
template <class C>
struct BaseSection {
    DevBuf bases;    // this shard's slice of the section
    DevBuf mask;     // infinity bitmask
    size_t lo = 0;   // first index of the slice within the section
    size_t n = 0;    // points in the slice
    Msm<C> msm;
};

struct g16_zkey {
    int device = 0;
    uint32_t n_vars = 0, n_public = 0, domain_size = 0;
    int logn = 0;
    uint64_t n_coefs = 0;
    int shard_rank = 0, shard_count = 1;
    HAffine<HFq> alpha1, beta1, delta1;
    HAffine<HFq2> beta2, delta2;
    BaseSection<G1> A, B1, C, H;
    BaseSection<G2> B2;
    DevBuf row_ptr, col, val;  // CSR of the A rows then the B rows (2n+1 pointers)
    DevBuf w;                  // witness, n_vars x 32 B standard form
    DevBuf w_next;             // the witness of the NEXT proof, filled by g16_witness_prefetch while this one is proved
    cudaStream_t s_copy = nullptr;
    cudaStream_t s_hpre = nullptr;     // H preparation in the chained schedule
    cudaEvent_t ev_copy = nullptr;
    bool next_ready = false;
    std::vector<uint8_t> next_public;
    DevBuf pa, pb, pc;         // A, B, C evaluations; pa ends up holding the H scalars
    NttDomain dom;
    NttDomain dom_half;                // 2^(logn-1) points, twiddles only: g16_h_transform_half (built on first use)
    cudaStream_t s_main = nullptr, s_msm[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    HPoint<HFq> fin_a, fin_b1, fin_c;   // finished witness MSMs of the last prove_device
    HPoint<HFq2> fin_b2;
    float t_transform_ms[2] = {0, 0};  // g16_h_transform: r1cs eval, transforms
    cudaStream_t s_pre[4] = {nullptr, nullptr, nullptr, nullptr};  // witness MSMs: digits + sort (high priority)
    cudaEvent_t ev[16];
    bool ev_ok = false;
    bool witness_loaded = false;
    std::vector<uint8_t> public_bytes;  // w[1..n_public], 32 B each
    double t_ms[10] = {0};
    double t_horner_ms = 0;
    bool partial_done = false;
    bool witness_enqueued = false;   // g16_prove_begin ran: the witness MSMs of this proof are already in flight

    ~g16_zkey() {
        cudaSetDevice(device);
        if (s_main) cudaStreamDestroy(s_main);
        if (s_copy) cudaStreamDestroy(s_copy);
        if (s_hpre) cudaStreamDestroy(s_hpre);
        if (ev_copy) cudaEventDestroy(ev_copy);
        for (auto& s : s_pre)
            if (s) cudaStreamDestroy(s);
        for (auto& s : s_msm)
            if (s) cudaStreamDestroy(s);
        if (ev_ok)
            for (auto& e : ev) cudaEventDestroy(e);
    }
};

namespace g16 {

static HAffine<HFq> dec_g1(const uint8_t* p) {
    HAffine<HFq> a;
    a.x = u256_from_bytes(p);
    a.y = u256_from_bytes(p + 32);
    a.inf = u256_is_zero(a.x) && u256_is_zero(a.y);
    return a;
}
static HAffine<HFq2> dec_g2(const uint8_t* p) {
    HAffine<HFq2> a;
    a.x.c0 = u256_from_bytes(p);
    a.x.c1 = u256_from_bytes(p + 32);
    a.y.c0 = u256_from_bytes(p + 64);
    a.y.c1 = u256_from_bytes(p + 96);
    a.inf = HFq2::is_zero(a.x) && HFq2::is_zero(a.y);
    return a;
}

static void set_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        throw StatusError(G16_ERR_CUDA, "no CUDA device: libg16b200 has no CPU fallback");
    if (device >= 0) {
        if (device >= count) throw StatusError(G16_ERR_CUDA, "CUDA device ordinal out of range");
        G16_CUDA(cudaSetDevice(device));
    }
}

template <class C>
static void load_section(BaseSection<C>& sec, const Section& s, size_t rec, size_t expect, int rank, int count,
                         int msm_c, bool sparse, cudaStream_t st, const char* name, int table = 0) {
    if (s.size != expect * rec) throw FormatError(std::string("zkey: section ") + name + " has the wrong size");
    sec.lo = expect * size_t(rank) / size_t(count);
    size_t hi = expect * size_t(rank + 1) / size_t(count);
    sec.n = hi - sec.lo;
    sec.bases.alloc(sec.n * rec);
    sec.mask.alloc(((sec.n + 31) / 32 + 1) * 4);
    if (sec.n) {
        G16_CUDA(cudaMemcpyAsync(sec.bases.p, s.p + sec.lo * rec, sec.n * rec, cudaMemcpyHostToDevice, st));
        msm_inf_mask_kernel<C><<<(unsigned)((sec.n + 255) / 256), 256, 0, st>>>(sec.bases.template as<typename C::aff>(),
                                                                                (uint32_t)sec.n, sec.mask.template as<uint32_t>());
        G16_CUDA(cudaGetLastError());
    }
    MsmConfig cfg;
    cfg.c = msm_c;
    cfg.sparse = sparse;
    if (table && sec.n) {
        // fixed-base window table (msm_window_table_kernel): on when asked for, or when it fits in a quarter of the
        // memory that is free right now
        int c = msm_c ? (msm_c < 6 ? 6 : msm_c > 24 ? 24 : msm_c) : sparse ? msm_pick_c_table_sparse(sec.n) : msm_pick_c_table(sec.n);
        size_t need = Msm<C>::table_bytes(sec.n, c), free_b = 0, total_b = 0;
        G16_CUDA(cudaMemGetInfo(&free_b, &total_b));
        bool fits = size_t(msm_num_windows(c)) * sec.n <= size_t(MSM_INDEX_MASK);
        cfg.table = fits && (table > 1 || need <= free_b / 4);
    }
    sec.msm.reserve(sec.n ? sec.n : 1, cfg);
    if (cfg.table) sec.msm.build_table(sec.bases.template as<typename C::aff>(), sec.n, st);
}

#ifndef G16_H_TABLE_DEFAULT
#define G16_H_TABLE_DEFAULT 1   // what opts.h_table == 0 means: 0 = off, 1 = on when it fits
#endif

static g16_zkey* zkey_build(const uint8_t* data, size_t len, const g16_opts* opts) {
    g16_opts o;
    if (opts) o = *opts;
    else g16_opts_default(&o);
    set_device(o.device);
    std::unique_ptr<g16_zkey> zk(new g16_zkey());
    G16_CUDA(cudaGetDevice(&zk->device));
    zk->shard_count = o.shard_count > 1 ? o.shard_count : 1;
    zk->shard_rank = o.shard_count > 1 ? o.shard_rank : 0;
    if (zk->shard_rank < 0 || zk->shard_rank >= zk->shard_count) throw std::runtime_error("bad shard rank");

    auto secs = parse_binfile(data, len, "zkey", 1);
    const Section& s1 = need(secs, 1, "zkey");
    if (s1.size < 4 || rd32(s1.p) != 1) throw FormatError("zkey file is not groth16");
    const Section& s2 = need(secs, 2, "zkey");
    const uint8_t* p = s2.p;
    if (s2.size < 4 + 32 + 4 + 32 + 12 + 64 * 3 + 128 * 3) throw FormatError("zkey: header section too short");
    uint32_t n8q = rd32(p);
    if (n8q != 32) throw StatusError(G16_ERR_CURVE, "zkey: curve not supported (n8q != 32)");
    U256 q = u256_from_bytes(p + 4);
    uint32_t n8r = rd32(p + 36);
    if (n8r != 32) throw StatusError(G16_ERR_CURVE, "zkey: curve not supported (n8r != 32)");
    U256 r = u256_from_bytes(p + 40);
    if (!u256_eq(q, HFq::mod()) || !u256_eq(r, HFr::mod())) throw StatusError(G16_ERR_CURVE, "zkey: curve is not bn128");
    p += 72;
    zk->n_vars = rd32(p);
    zk->n_public = rd32(p + 4);
    zk->domain_size = rd32(p + 8);
    p += 12;
    uint32_t n = zk->domain_size;
    if (n < 2 || (n & (n - 1))) throw FormatError("zkey: domain size is not a power of two >= 2");
    while ((1u << zk->logn) < n) zk->logn++;
    if (zk->n_public + 1 > zk->n_vars) throw FormatError("zkey: nPublic exceeds nVars");
    zk->alpha1 = dec_g1(p);
    zk->beta1 = dec_g1(p + 64);
    zk->beta2 = dec_g2(p + 128);
    zk->delta1 = dec_g1(p + 128 + 128 + 128);
    zk->delta2 = dec_g2(p + 128 + 128 + 128 + 64);

    // Priorities order the throughput-bound work so that the latency-bound tails overlap it: transforms first (they
    // gate the H MSM), then the witness-scalar MSMs, whose bucket reductions (short dependency chains on few
    // threads) then run while the H accumulation, lowest, fills the machine.
    int prio_lo = 0, prio_hi = 0;
    G16_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    int prio_mid = prio_hi < prio_lo - 1 ? prio_hi + 1 : prio_hi;
    static const bool chain_main = [] { const char* e = std::getenv("G16_CHAIN"); return e && std::atoi(e) != 0; }();
    G16_CUDA(cudaStreamCreateWithPriority(&zk->s_main, cudaStreamNonBlocking, chain_main ? prio_hi + 1 : prio_hi));
    // G16_HPRIO (scheduling experiment): priority of the H accumulation stream, 0 = lowest (default), 1 = middle, 2 = highest
    static const int hprio = [] { const char* e = std::getenv("G16_HPRIO"); return e ? std::atoi(e) : 0; }();
    int prio_h = hprio == 2 ? prio_hi : hprio == 1 ? prio_mid : prio_lo;
    // with the H accumulation above them, the witness accumulations drop to the lowest priority (they become the filler)
    int prio_w = hprio == 2 ? prio_lo : prio_mid;
    static const bool chain = [] { const char* e = std::getenv("G16_CHAIN"); return e && std::atoi(e) != 0; }();
    if (chain) prio_h = prio_w = prio_hi;   // the side streams only carry latency-bound work then
    for (int i = 0; i < 5; i++)
        G16_CUDA(cudaStreamCreateWithPriority(&zk->s_msm[i], cudaStreamNonBlocking, i == 4 ? prio_h : prio_w));
    G16_CUDA(cudaStreamCreateWithPriority(&zk->s_hpre, cudaStreamNonBlocking, prio_hi));
    // digits + sort of the witness MSMs are HBM-bound and short; at the transforms' priority they interleave with the
    // (integer-bound) transforms instead of being starved by them, so the accumulations are ready when the transforms end
    for (int i = 0; i < 4; i++) G16_CUDA(cudaStreamCreateWithPriority(&zk->s_pre[i], cudaStreamNonBlocking, prio_hi));
    for (auto& e : zk->ev) G16_CUDA(cudaEventCreate(&e));
    zk->ev_ok = true;
    Msm<G1>::prepare();
    Msm<G2>::prepare();
    cudaStream_t st = zk->s_main;

    // sections 5-9: base points (this shard's slice)
    size_t nv = zk->n_vars;
    // G16_W_TABLE=1: window tables for the witness-scalar sections too (one bucket set of wider windows: fewer digits per
    // 64-bit and full-width wire); same memory rule as the H table
    int w_table = 0;
    if (const char* e = std::getenv("G16_W_TABLE")) w_table = std::atoi(e) > 0 ? 1 : 0;
    load_section<G1>(zk->A, need(secs, 5, "zkey"), 64, nv, zk->shard_rank, zk->shard_count, o.msm_c, true, st, "5 (A)", w_table);
    load_section<G1>(zk->B1, need(secs, 6, "zkey"), 64, nv, zk->shard_rank, zk->shard_count, o.msm_c, true, st, "6 (B1)", w_table);
    load_section<G2>(zk->B2, need(secs, 7, "zkey"), 128, nv, zk->shard_rank, zk->shard_count, o.msm_c, true, st, "7 (B2)", w_table);
    load_section<G1>(zk->C, need(secs, 8, "zkey"), 64, nv - zk->n_public - 1, zk->shard_rank, zk->shard_count, o.msm_c,
                     true, st, "8 (C)", w_table);
    // H scalars are dense (every window of every scalar is live): the one section where a window table pays.
    // opts.h_table / G16_H_TABLE: 1 = always, -1 = never, 0 = when it fits (see load_section)
    int h_table = o.h_table;
    if (const char* e = std::getenv("G16_H_TABLE")) h_table = std::atoi(e) > 0 ? 1 : -1;
    load_section<G1>(zk->H, need(secs, 9, "zkey"), 64, n, zk->shard_rank, zk->shard_count, o.msm_c, false, st, "9 (H)",
                     h_table > 0 ? 2 : h_table == 0 ? G16_H_TABLE_DEFAULT : 0);

    // section 4: coefficients -> CSR by (matrix, constraint)
    const Section& s4 = need(secs, 4, "zkey");
    if (s4.size < 4) throw FormatError("zkey: coefficient section too short");
    zk->n_coefs = rd32(s4.p);
    if (s4.size != 4 + zk->n_coefs * 44) throw FormatError("zkey: coefficient section has the wrong size");
    {
        DevBuf raw, counts, bad;
        raw.alloc(zk->n_coefs * 44);
        size_t nrows = 2 * size_t(n);
        counts.alloc((nrows + 1) * 4);
        bad.alloc(4);
        zk->row_ptr.alloc((nrows + 1) * 4);
        zk->col.alloc(zk->n_coefs * 4);
        zk->val.alloc(zk->n_coefs * 32);
        G16_CUDA(cudaMemsetAsync(counts.p, 0, (nrows + 1) * 4, st));
        G16_CUDA(cudaMemsetAsync(bad.p, 0, 4, st));
        if (zk->n_coefs) {
            G16_CUDA(cudaMemcpyAsync(raw.p, s4.p + 4, zk->n_coefs * 44, cudaMemcpyHostToDevice, st));
            unsigned grid = (unsigned)((zk->n_coefs + 255) / 256);
            r1cs_count_kernel<<<grid, 256, 0, st>>>(raw.as<uint8_t>(), zk->n_coefs, n, zk->n_vars, counts.as<uint32_t>(),
                                                    bad.as<uint32_t>());
        }
        size_t tb = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tb, counts.as<uint32_t>(), zk->row_ptr.as<uint32_t>(), (int)(nrows + 1));
        DevBuf tmp;
        tmp.alloc(tb + 16);
        G16_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, counts.as<uint32_t>(), zk->row_ptr.as<uint32_t>(),
                                               (int)(nrows + 1), st));
        G16_CUDA(cudaMemsetAsync(counts.p, 0, (nrows + 1) * 4, st));
        if (zk->n_coefs) {
            unsigned grid = (unsigned)((zk->n_coefs + 255) / 256);
            r1cs_fill_kernel<<<grid, 256, 0, st>>>(raw.as<uint8_t>(), zk->n_coefs, n, zk->row_ptr.as<uint32_t>(),
                                                   counts.as<uint32_t>(), zk->col.as<uint32_t>(), zk->val.as<fe_t>());
        }
        uint32_t hbad = 0;
        G16_CUDA(cudaMemcpyAsync(&hbad, bad.p, 4, cudaMemcpyDeviceToHost, st));
        G16_CUDA(cudaStreamSynchronize(st));
        if (hbad) throw FormatError("zkey: coefficient record out of range (matrix, constraint or signal index)");
    }
    zk->w.alloc(size_t(zk->n_vars) * 32);
    zk->pa.alloc(size_t(n) * 32);
    zk->pb.alloc(size_t(n) * 32);
    zk->pc.alloc(size_t(n) * 32);
    zk->dom.init(zk->logn, st);
    G16_CUDA(cudaStreamSynchronize(st));
    return zk.release();
}

}  // namespace g16

// ---------------------------------------------------------------------------------------------------------
// output encoding (row a16): the two byte layouts the reference's provers write
// ---------------------------------------------------------------------------------------------------------
namespace g16 {

struct J {  // minimal JSON tree: string leaf, list, or ordered object
    enum Kind { STR, LIST, OBJ } kind = STR;
    std::string s;
    std::vector<J> items;
    std::vector<std::string> keys;
    static J str(const std::string& v) {
        J j;
        j.s = v;
        return j;
    }
    static J list(std::vector<J> v) {
        J j;
        j.kind = LIST;
        j.items = std::move(v);
        return j;
    }
};

// rapidsnark: nlohmann compact dump; snarkjs: JSON.stringify(obj, null, 1); neither ends with a newline
static void dump(const J& j, int style, int depth, std::string& out) {
    if (j.kind == J::STR) {
        out += '"';
        out += j.s;
        out += '"';
        return;
    }
    char open = j.kind == J::LIST ? '[' : '{', close = j.kind == J::LIST ? ']' : '}';
    out += open;
    if (j.items.empty()) {
        out += close;
        return;
    }
    for (size_t i = 0; i < j.items.size(); i++) {
        if (i) out += ',';
        if (style == G16_STYLE_SNARKJS) {
            out += '\n';
            out.append(depth + 1, ' ');
        }
        if (j.kind == J::OBJ) {
            out += '"';
            out += j.keys[i];
            out += style == G16_STYLE_SNARKJS ? "\": " : "\":";
        }
        dump(j.items[i], style, depth + 1, out);
    }
    if (style == G16_STYLE_SNARKJS) {
        out += '\n';
        out.append(depth, ' ');
    }
    out += close;
}

static J g1_json(const HAffine<HFq>& a) {
    if (a.inf) return J::list({J::str("0"), J::str("1"), J::str("0")});
    return J::list({J::str(u256_to_dec(HFq::from_mont(a.x))), J::str(u256_to_dec(HFq::from_mont(a.y))), J::str("1")});
}
static J g2_json(const HAffine<HFq2>& a) {
    auto pair = [](const std::string& x, const std::string& y) { return J::list({J::str(x), J::str(y)}); };
    if (a.inf) return J::list({pair("0", "0"), pair("1", "0"), pair("0", "0")});
    return J::list({pair(u256_to_dec(HFq::from_mont(a.x.c0)), u256_to_dec(HFq::from_mont(a.x.c1))),
                    pair(u256_to_dec(HFq::from_mont(a.y.c0)), u256_to_dec(HFq::from_mont(a.y.c1))), pair("1", "0")});
}

static std::string proof_json(const HAffine<HFq>& a, const HAffine<HFq2>& b, const HAffine<HFq>& c, int style) {
    J o;
    o.kind = J::OBJ;
    o.keys = {"pi_a", "pi_b", "pi_c", "protocol"};
    o.items = {g1_json(a), g2_json(b), g1_json(c), J::str("groth16")};
    if (style == G16_STYLE_SNARKJS) {
        o.keys.push_back("curve");
        o.items.push_back(J::str("bn128"));
    }
    std::string out;
    dump(o, style, 0, out);
    return out;
}

static std::string public_json(const std::vector<uint8_t>& pub, int style) {
    std::vector<J> v;
    for (size_t i = 0; i + 32 <= pub.size(); i += 32) v.push_back(J::str(u256_to_dec(u256_from_bytes(&pub[i]))));
    std::string out;
    dump(J::list(std::move(v)), style, 0, out);
    return out;
}

// ---------------------------------------------------------------------------------------------------------
// witness (row a3)
// ---------------------------------------------------------------------------------------------------------
// header checks of a .wtns image; returns the payload (n_vars x 32 B, standard form)
static const uint8_t* witness_check(const g16_zkey* zk, const uint8_t* data, size_t len) {
    auto secs = parse_binfile(data, len, "wtns", 2);
    const Section& s1 = need(secs, 1, "wtns");
    if (s1.size < 4) throw FormatError("wtns: header section too short");
    uint32_t n8 = rd32(s1.p);
    if (s1.size != 4 + uint64_t(n8) + 4) throw FormatError("wtns: header section has the wrong size");
    if (n8 != 32 || !u256_eq(u256_from_bytes(s1.p + 4), HFr::mod()))
        throw StatusError(G16_ERR_CURVE, "Curve of the witness does not match the curve of the proving key");
    uint32_t nw = rd32(s1.p + 4 + n8);
    if (nw != zk->n_vars)
        throw StatusError(G16_ERR_WITNESS_LEN, "Invalid witness length. Circuit: " + std::to_string(zk->n_vars) +
                                                   ", witness: " + std::to_string(nw));
    const Section& s2 = need(secs, 2, "wtns");
    if (s2.size != uint64_t(nw) * 32) throw FormatError("wtns: data section has the wrong size");
    return s2.p;
}

static void witness_load(g16_zkey* zk, const uint8_t* data, size_t len) {
    struct { const uint8_t* p; } s2{witness_check(zk, data, len)};
    const uint32_t nw = zk->n_vars;
    G16_CUDA(cudaSetDevice(zk->device));
    G16_CUDA(cudaEventRecord(zk->ev[13], zk->s_main));
    G16_CUDA(cudaMemcpyAsync(zk->w.p, s2.p, size_t(nw) * 32, cudaMemcpyHostToDevice, zk->s_main));
    G16_CUDA(cudaEventRecord(zk->ev[14], zk->s_main));
    zk->public_bytes.assign(s2.p + 32, s2.p + 32 + size_t(zk->n_public) * 32);
    G16_CUDA(cudaStreamSynchronize(zk->s_main));
    float ms = 0;
    cudaEventElapsedTime(&ms, zk->ev[13], zk->ev[14]);
    zk->t_ms[0] = ms;
    zk->witness_loaded = true;
    zk->partial_done = false;
    zk->witness_enqueued = false;
}

// rows a4-a8 on s_main: leaves the H scalars (standard form) in zk->pa
static void enqueue_h_pipeline(g16_zkey* zk) {
    cudaStream_t s = zk->s_main;
    uint32_t n = zk->domain_size;
    r1cs_eval_kernel<<<(n + 127) / 128, 128, 0, s>>>(zk->row_ptr.as<uint32_t>(), zk->col.as<uint32_t>(),
                                                     zk->val.as<fe_t>(), zk->w.as<fe_t>(), n, zk->pa.as<fe_t>(),
                                                     zk->pb.as<fe_t>(), zk->pc.as<fe_t>());
    G16_CUDA(cudaGetLastError());
    G16_CUDA(cudaEventRecord(zk->ev[1], s));
    zk->dom.to_odd_coset(zk->pa.as<fe_t>(), s);
    zk->dom.to_odd_coset(zk->pb.as<fe_t>(), s);
    zk->dom.to_odd_coset(zk->pc.as<fe_t>(), s);
    ntt_join_kernel<<<(n + 255) / 256, 256, 0, s>>>(zk->pa.as<fe_t>(), zk->pb.as<fe_t>(), zk->pc.as<fe_t>(),
                                                    zk->pa.as<fe_t>(), n);
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(2);
    G16_CUDA(cudaEventRecord(zk->ev[2], s));
}

// coset evaluations of this shard's H range, produced elsewhere (transforms spread over ranks): device pointers to
// H.n elements each; the join (row a8) then runs on the range only
struct JoinedSlices {
    const fe_t *a, *b, *c;
    // transforms split in halves (g16_h_transform_half): a, b, c = the even-coefficient halves' slices, these the odd ones
    const fe_t *oa = nullptr, *ob = nullptr, *oc = nullptr;
};

// the device part of a proof: five partial MSM sums of this shard, left in the Msm objects' host buffers.
// The whole proof is ONE enqueue: no host thread and no read-back of a counter sits between the kernels (the MSMs size
// their launches from device memory, msm.cuh); the host only waits for the five results at the end, in the order they
// finish, and runs each finished MSM's window Horner chain while the others still compute.
static void prove_device(g16_zkey* zk, const JoinedSlices* js = nullptr, bool only_witness = false) {
    if (!zk->witness_loaded) throw std::runtime_error("no witness loaded");
    G16_CUDA(cudaSetDevice(zk->device));
    auto t0 = std::chrono::steady_clock::now();
    const uint32_t* w = zk->w.as<uint32_t>();
    // G16_SERIAL=1 (measurement): every kernel of the proof on ONE stream, i.e. the sum of the stand-alone phase times
    static const bool serial = [] { const char* e = std::getenv("G16_SERIAL"); return e && std::atoi(e) != 0; }();
    cudaStream_t s_pre[4], s_msm[5];
    for (int i = 0; i < 4; i++) s_pre[i] = serial ? zk->s_main : zk->s_pre[i];
    for (int i = 0; i < 5; i++) s_msm[i] = serial ? zk->s_main : zk->s_msm[i];
    try {
        if (!zk->witness_enqueued) G16_CUDA(cudaEventRecord(zk->ev[0], zk->s_main));
        // witness-scalar MSMs start at once; the H MSM waits for the NTT pipeline.  Preparation (counting sort of the
        // digits, HBM-bound and short) on streams of the transforms' priority so that it interleaves with the
        // integer-bound transforms; accumulations + reductions at middle priority.  B2 first: its G2 reductions are the
        // longest latency-bound tail.
        if (!zk->witness_enqueued)
            for (int i = 0; i < 4; i++) {
                G16_CUDA(cudaStreamWaitEvent(s_pre[i], zk->ev[0], 0));
                G16_CUDA(cudaEventRecord(zk->ev[3 + 2 * i], s_pre[i]));
            }
        // G16_GATE (scheduling experiment): bit 0 = the G1 witness accumulations wait for the transforms, bit 1 = B2 too.
        // The transforms then run without long-lived accumulation CTAs in their way, at the price of idle pipe slots.
        static const int gate_mask = [] { const char* e = std::getenv("G16_GATE"); return e ? std::atoi(e) : 0; }();
        const bool deferred = (gate_mask & 4) != 0;   // bit 2: enqueue the witness MSMs after the H chain
        cudaEvent_t ntt_done = zk->ev[2];
        const bool nogate = deferred || only_witness;   // nothing to wait for when the transforms are not part of this enqueue
        // G16_CHAIN=1 (scheduling experiment): the pipe-bound kernels (transforms, then the five accumulations) back to
        // back on the transform stream; preparation and reductions (latency-bound) on the side streams
        static const bool chain = [] { const char* e = std::getenv("G16_CHAIN"); return e && std::atoi(e) != 0; }();
        auto witness_msms = [&]() {
        if (chain) {
            zk->B2.msm.enqueue(zk->B2.bases.as<g2_aff>(), w + zk->B2.lo * 8, zk->B2.mask.as<uint32_t>(), zk->B2.n, s_pre[2], zk->s_main, nullptr, s_msm[2]);
            G16_CUDA(cudaEventRecord(zk->ev[8], s_msm[2]));
            zk->A.msm.enqueue(zk->A.bases.as<g1_aff>(), w + zk->A.lo * 8, zk->A.mask.as<uint32_t>(), zk->A.n, s_pre[0], zk->s_main, nullptr, s_msm[0]);
            G16_CUDA(cudaEventRecord(zk->ev[4], s_msm[0]));
            zk->B1.msm.enqueue(zk->B1.bases.as<g1_aff>(), w + zk->B1.lo * 8, zk->B1.mask.as<uint32_t>(), zk->B1.n, s_pre[1], zk->s_main, nullptr, s_msm[1]);
            G16_CUDA(cudaEventRecord(zk->ev[6], s_msm[1]));
            zk->C.msm.enqueue(zk->C.bases.as<g1_aff>(), w + (size_t(zk->n_public) + 1 + zk->C.lo) * 8, zk->C.mask.as<uint32_t>(), zk->C.n,
                              s_pre[3], zk->s_main, nullptr, s_msm[3]);
            G16_CUDA(cudaEventRecord(zk->ev[10], s_msm[3]));
            return;
        }
        if ((gate_mask & 8) && !nogate)   // bit 3: the witness MSMs' preparation waits for the transforms as well
            for (int i = 0; i < 4; i++) G16_CUDA(cudaStreamWaitEvent(s_pre[i], ntt_done, 0));
        zk->B2.msm.enqueue(zk->B2.bases.as<g2_aff>(), w + zk->B2.lo * 8, zk->B2.mask.as<uint32_t>(), zk->B2.n, s_pre[2], s_msm[2],
                           (gate_mask & 2) && !nogate ? ntt_done : nullptr);
        G16_CUDA(cudaEventRecord(zk->ev[8], s_msm[2]));
        zk->A.msm.enqueue(zk->A.bases.as<g1_aff>(), w + zk->A.lo * 8, zk->A.mask.as<uint32_t>(), zk->A.n, s_pre[0], s_msm[0],
                          (gate_mask & 1) && !nogate ? ntt_done : nullptr);
        G16_CUDA(cudaEventRecord(zk->ev[4], s_msm[0]));
        zk->B1.msm.enqueue(zk->B1.bases.as<g1_aff>(), w + zk->B1.lo * 8, zk->B1.mask.as<uint32_t>(), zk->B1.n, s_pre[1], s_msm[1],
                           (gate_mask & 1) && !nogate ? ntt_done : nullptr);
        G16_CUDA(cudaEventRecord(zk->ev[6], s_msm[1]));
        zk->C.msm.enqueue(zk->C.bases.as<g1_aff>(), w + (size_t(zk->n_public) + 1 + zk->C.lo) * 8, zk->C.mask.as<uint32_t>(), zk->C.n,
                          s_pre[3], s_msm[3], (gate_mask & 1) && !nogate ? ntt_done : nullptr);
        G16_CUDA(cudaEventRecord(zk->ev[10], s_msm[3]));
        };
        if (only_witness) {
            // fall through to the witness MSMs below
        } else if (js) {
            G16_CUDA(cudaEventRecord(zk->ev[1], zk->s_main));
            if (zk->H.n && js->oa) {
                ntt_join_halves_kernel<<<(unsigned)((zk->H.n + 255) / 256), 256, 0, zk->s_main>>>(
                    js->a, js->oa, js->b, js->ob, js->c, js->oc, zk->pa.as<fe_t>() + zk->H.lo, (uint32_t)zk->H.n,
                    (uint32_t)zk->H.lo, zk->domain_size / 2, zk->dom.tw.as<fe_t>());
                G16_CUDA(cudaGetLastError());
                G16_LAUNCHED(1);
            } else if (zk->H.n) {
                ntt_join_kernel<<<(unsigned)((zk->H.n + 255) / 256), 256, 0, zk->s_main>>>(
                    js->a, js->b, js->c, zk->pa.as<fe_t>() + zk->H.lo, (uint32_t)zk->H.n);
                G16_CUDA(cudaGetLastError());
                G16_LAUNCHED(1);
            }
            G16_CUDA(cudaEventRecord(zk->ev[2], zk->s_main));
        } else {
            enqueue_h_pipeline(zk);
        }
        if (only_witness) {   // g16_prove_begin: the witness MSMs only, nothing waited for
            witness_msms();
            zk->witness_enqueued = true;
            return;
        }
        if (!deferred && !zk->witness_enqueued) witness_msms();   // after the transforms are enqueued: ev[2] is this proof's record
        // H: preparation stays on the transform stream (highest priority, already ordered after the join); only the
        // accumulation and reductions move to the lowest-priority stream, filling the machine under everything else
        G16_CUDA(cudaEventRecord(zk->ev[11], zk->s_main));
        if (chain) {
            // H preparation on a side stream once the join is done; its accumulation last on the main stream
            G16_CUDA(cudaStreamWaitEvent(zk->s_hpre, zk->ev[2], 0));   // the join, not the accumulations queued behind it
            zk->H.msm.enqueue(zk->H.bases.as<g1_aff>(), zk->pa.as<uint32_t>() + zk->H.lo * 8, zk->H.mask.as<uint32_t>(), zk->H.n,
                              zk->s_hpre, zk->s_main, nullptr, s_msm[4]);
        } else
        zk->H.msm.enqueue(zk->H.bases.as<g1_aff>(), zk->pa.as<uint32_t>() + zk->H.lo * 8, zk->H.mask.as<uint32_t>(), zk->H.n,
                          zk->s_main, s_msm[4]);
        G16_CUDA(cudaEventRecord(zk->ev[12], s_msm[4]));
        if (deferred && !zk->witness_enqueued) witness_msms();
        zk->witness_enqueued = false;
        // results, in the order they are expected to finish; each finished MSM's Horner chain (255 doublings on the
        // host) overlaps the MSMs still running
        zk->A.msm.wait();  zk->fin_a = zk->A.msm.finish<HG1>();
        zk->B1.msm.wait(); zk->fin_b1 = zk->B1.msm.finish<HG1>();
        zk->C.msm.wait();  zk->fin_c = zk->C.msm.finish<HG1>();
        zk->B2.msm.wait(); zk->fin_b2 = zk->B2.msm.finish<HG2>();
        zk->H.msm.wait();
        for (int i = 0; i < 5; i++) G16_CUDA(cudaStreamSynchronize(s_msm[i]));
        G16_CUDA(cudaStreamSynchronize(zk->s_main));
    } catch (...) {
        // a failed enqueue leaves kernels of this proof in flight on other streams: drain them before the handle is
        // reused (the service keeps a key resident after a failed proof), and forget the half-made results
        cudaDeviceSynchronize();
        zk->partial_done = false;
        zk->witness_enqueued = false;
        throw;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, zk->ev[0], zk->ev[1]); zk->t_ms[1] = ms;
    cudaEventElapsedTime(&ms, zk->ev[1], zk->ev[2]); zk->t_ms[2] = ms;
    for (int i = 0; i < 5; i++) {
        cudaEventElapsedTime(&ms, zk->ev[3 + 2 * i], zk->ev[4 + 2 * i]);
        zk->t_ms[3 + i] = ms;
    }
    zk->t_ms[8] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    zk->partial_done = true;
}

static void export_partial(g16_zkey* zk, uint8_t* out) {
    auto t0 = std::chrono::steady_clock::now();
    HPoint<HFq> h = zk->H.msm.finish<HG1>();   // the other four chains were run in prove_device
    std::memcpy(out, &zk->fin_a, 128);
    std::memcpy(out + 128, &zk->fin_b1, 128);
    std::memcpy(out + 256, &zk->fin_c, 128);
    std::memcpy(out + 384, &h, 128);
    std::memcpy(out + 512, &zk->fin_b2, 256);
    zk->t_horner_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

static HPoint<HFq> g1_from_xyzz(const uint8_t* p) {
    return HPoint<HFq>{u256_from_bytes(p), u256_from_bytes(p + 32), u256_from_bytes(p + 64), u256_from_bytes(p + 96)};
}
static HPoint<HFq2> g2_from_xyzz(const uint8_t* p) {
    HPoint<HFq2> r;
    r.x = HFq2El{u256_from_bytes(p), u256_from_bytes(p + 32)};
    r.y = HFq2El{u256_from_bytes(p + 64), u256_from_bytes(p + 96)};
    r.zz = HFq2El{u256_from_bytes(p + 128), u256_from_bytes(p + 160)};
    r.zzz = HFq2El{u256_from_bytes(p + 192), u256_from_bytes(p + 224)};
    return r;
}

static U256 reduce_scalar(U256 a) {
    while (u256_cmp(a, HFr::mod()) >= 0) a = HFr::raw_sub(a, HFr::mod());
    return a;
}
// row a14: uniform in Fr by rejection, from the kernel's CSPRNG (getrandom(2); rapidsnark uses libsodium's
// randombytes_buf, which reads the same source).  The zero-knowledge property of every proof rests on these bytes,
// so a failing source is an error, never a fallback to a weaker generator.
static U256 random_scalar() {
    for (;;) {
        U256 a;
        uint8_t* p = reinterpret_cast<uint8_t*>(a.v);
        size_t got = 0;
        while (got < 32) {
            ssize_t k = getrandom(p + got, 32 - got, 0);
            if (k < 0) {
                if (errno == EINTR) continue;
                throw std::runtime_error("getrandom failed: no entropy source for the blinding factors");
            }
            got += size_t(k);
        }
        a.v[3] &= 0x3fffffffffffffffull;
        if (u256_cmp(a, HFr::mod()) < 0) return a;
    }
}

// row a14 + the part of row a15 that does not depend on the MSM results: r, s and the multiples of delta.
// Computed on host threads while the GPU works.
struct Blind {
    U256 r, s;
    HPoint<HFq> rd1, sd1, rsd1;
    HPoint<HFq2> sd2;
};
static Blind make_blind(const g16_zkey* zk, const g16_opts& o) {
    Blind b;
    if (o.has_rs) {
        b.r = reduce_scalar(u256_from_bytes(o.r));
        b.s = reduce_scalar(u256_from_bytes(o.s));
    } else {
        b.r = random_scalar();
        b.s = random_scalar();
    }
    U256 rs = HFr::from_mont(HFr::mul(HFr::to_mont(b.r), HFr::to_mont(b.s)));
    HPoint<HFq> d1 = HG1::from_affine(zk->delta1);
    HPoint<HFq2> d2 = HG2::from_affine(zk->delta2);
    auto f1 = std::async(std::launch::async, [&] { return HG1::mul(d1, b.r); });
    auto f2 = std::async(std::launch::async, [&] { return HG1::mul(d1, b.s); });
    auto f3 = std::async(std::launch::async, [&] { return HG1::mul(d1, rs); });
    b.sd2 = HG2::mul(d2, b.s);
    b.rd1 = f1.get();
    b.sd1 = f2.get();
    b.rsd1 = f3.get();
    return b;
}

// rest of row a15 + row a16
static void combine(g16_zkey* zk, const uint8_t* partials, int count, const g16_opts& o, const Blind& bl,
                    std::string& proof, std::string& pub) {
    auto t0 = std::chrono::steady_clock::now();
    HPoint<HFq> a = HG1::identity(), b1 = a, c = a, h = a;
    HPoint<HFq2> b2 = HG2::identity();
    for (int i = 0; i < count; i++) {
        const uint8_t* p = partials + size_t(i) * G16_PARTIAL_BYTES;
        a = HG1::add(a, g1_from_xyzz(p));
        b1 = HG1::add(b1, g1_from_xyzz(p + 128));
        c = HG1::add(c, g1_from_xyzz(p + 256));
        h = HG1::add(h, g1_from_xyzz(p + 384));
        b2 = HG2::add(b2, g2_from_xyzz(p + 512));
    }
    HPoint<HFq> pi_a = HG1::add(HG1::add(a, HG1::from_affine(zk->alpha1)), bl.rd1);
    HPoint<HFq> pb1 = HG1::add(HG1::add(b1, HG1::from_affine(zk->beta1)), bl.sd1);
    auto fsa = std::async(std::launch::async, [&] { return HG1::mul(pi_a, bl.s); });
    auto fb = std::async(std::launch::async, [&] {
        return HG2::to_affine(HG2::add(HG2::add(b2, HG2::from_affine(zk->beta2)), bl.sd2));
    });
    HPoint<HFq> rb1 = HG1::mul(pb1, bl.r);
    HPoint<HFq> pi_c = HG1::add(c, h);
    pi_c = HG1::add(pi_c, fsa.get());
    pi_c = HG1::add(pi_c, rb1);
    pi_c = HG1::add(pi_c, HG1::neg(bl.rsd1));
    HAffine<HFq> A = HG1::to_affine(pi_a), Cc = HG1::to_affine(pi_c);
    proof = proof_json(A, fb.get(), Cc, o.style);
    pub = public_json(zk->public_bytes, o.style);
    zk->t_ms[9] = zk->t_horner_ms + std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

static int emit(const std::string& proof, const std::string& pub, char* proof_json, size_t* proof_len,
                char* public_json, size_t* public_len) {
    if (!proof_len || !public_len) return fail(G16_ERR_GENERIC, "null length pointer");
    bool shortbuf = *proof_len < proof.size() || *public_len < pub.size() || !proof_json || !public_json;
    if (!shortbuf) {
        std::memcpy(proof_json, proof.data(), proof.size());
        std::memcpy(public_json, pub.data(), pub.size());
        if (*proof_len > proof.size()) proof_json[proof.size()] = 0;
        if (*public_len > pub.size()) public_json[pub.size()] = 0;
    }
    *proof_len = proof.size();
    *public_len = pub.size();
    return shortbuf ? fail(G16_ERR_SHORT_BUFFER, "output buffer too short") : G16_OK;
}

template <class F>
static int guarded(F&& f) {
    try {
        return f();
    } catch (const StatusError& e) {
        return fail(e.code, e.what());
    } catch (const FormatError& e) {
        return fail(G16_ERR_FORMAT, e.what());
    } catch (const std::exception& e) {
        std::string m = e.what();
        return fail(m.rfind("CUDA error", 0) == 0 ? G16_ERR_CUDA : G16_ERR_GENERIC, m);
    }
}

struct MappedFile {
    const uint8_t* p = nullptr;
    size_t len = 0;
    int fd = -1;
    explicit MappedFile(const char* path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) throw std::runtime_error(std::string("cannot open ") + path);
        struct stat st;
        if (fstat(fd, &st) != 0) throw std::runtime_error(std::string("cannot stat ") + path);
        len = (size_t)st.st_size;
        if (len) {
            void* m = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
            if (m == MAP_FAILED) throw std::runtime_error(std::string("cannot map ") + path);
            p = static_cast<const uint8_t*>(m);
        }
    }
    ~MappedFile() {
        if (p) munmap(const_cast<uint8_t*>(p), len);
        if (fd >= 0) ::close(fd);
    }
};

// written to a temporary beside the target and renamed, so that a concurrent reader (the workflow verifies right after
// the prover exits, scripts/g16_verify.sh) or a crash never sees a truncated proof.json
static void write_file(const char* path, const std::string& text) {
    std::string tmp = std::string(path) + ".tmp." + std::to_string((long)getpid());
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) throw std::runtime_error(std::string("cannot write ") + path);
    size_t w = std::fwrite(text.data(), 1, text.size(), f);
    if (std::fclose(f) != 0 || w != text.size()) {
        std::remove(tmp.c_str());
        throw std::runtime_error(std::string("short write to ") + path);
    }
    if (std::rename(tmp.c_str(), path) != 0) {
        std::remove(tmp.c_str());
        throw std::runtime_error(std::string("cannot rename onto ") + path);
    }
}

}  // namespace g16

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
extern "C" {

void g16_opts_default(g16_opts* o) {
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->device = -1;
    o->style = G16_STYLE_RAPIDSNARK;
    o->shard_count = 1;
}

const char* g16_last_error(void) { return g_last_error.c_str(); }
const char* g16_version(void) { return "g16b200 0.1 (sm_100a)"; }

int g16_zkey_open_mem(const uint8_t* data, size_t len, const g16_opts* opts, g16_zkey** out) {
    return guarded([&]() {
        if (!data || !out) return fail(G16_ERR_GENERIC, "null argument");
        *out = zkey_build(data, len, opts);
        return (int)G16_OK;
    });
}

int g16_zkey_open(const char* path, const g16_opts* opts, g16_zkey** out) {
    return guarded([&]() {
        if (!path || !out) return fail(G16_ERR_GENERIC, "null argument");
        MappedFile f(path);
        *out = zkey_build(f.p, f.len, opts);
        return (int)G16_OK;
    });
}

void g16_zkey_close(g16_zkey* zk) { delete zk; }

int g16_zkey_info(const g16_zkey* zk, uint32_t* n_vars, uint32_t* n_public, uint32_t* domain_size, uint64_t* n_coefs) {
    if (!zk) return fail(G16_ERR_GENERIC, "null handle");
    if (n_vars) *n_vars = zk->n_vars;
    if (n_public) *n_public = zk->n_public;
    if (domain_size) *domain_size = zk->domain_size;
    if (n_coefs) *n_coefs = zk->n_coefs;
    return G16_OK;
}

int g16_witness_load(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len) {
    return guarded([&]() {
        if (!zk || !wtns) return fail(G16_ERR_GENERIC, "null argument");
        witness_load(zk, wtns, wtns_len);
        return (int)G16_OK;
    });
}

// Double-buffered witness: the host->device copy of the next proof's witness runs on its own stream while the current
// proof computes; commit waits for it and swaps the buffers.
int g16_witness_prefetch(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len) {
    return guarded([&]() {
        if (!zk || !wtns) return fail(G16_ERR_GENERIC, "null argument");
        const uint8_t* payload = witness_check(zk, wtns, wtns_len);
        G16_CUDA(cudaSetDevice(zk->device));
        if (!zk->s_copy) G16_CUDA(cudaStreamCreateWithFlags(&zk->s_copy, cudaStreamNonBlocking));
        if (!zk->ev_copy) G16_CUDA(cudaEventCreateWithFlags(&zk->ev_copy, cudaEventDisableTiming));
        if (!zk->w_next.p) zk->w_next.alloc(size_t(zk->n_vars) * 32);
        G16_CUDA(cudaMemcpyAsync(zk->w_next.p, payload, size_t(zk->n_vars) * 32, cudaMemcpyHostToDevice, zk->s_copy));
        G16_CUDA(cudaEventRecord(zk->ev_copy, zk->s_copy));
        zk->next_public.assign(payload + 32, payload + 32 + size_t(zk->n_public) * 32);
        zk->next_ready = true;
        return (int)G16_OK;
    });
}

int g16_witness_commit(g16_zkey* zk) {
    return guarded([&]() {
        if (!zk) return fail(G16_ERR_GENERIC, "null handle");
        if (!zk->next_ready) return fail(G16_ERR_GENERIC, "no prefetched witness: call g16_witness_prefetch first");
        G16_CUDA(cudaSetDevice(zk->device));
        G16_CUDA(cudaEventSynchronize(zk->ev_copy));
        std::swap(zk->w.p, zk->w_next.p);
        std::swap(zk->w.bytes, zk->w_next.bytes);
        zk->public_bytes.swap(zk->next_public);
        zk->next_ready = false;
        zk->t_ms[0] = 0;
        zk->witness_loaded = true;
        zk->partial_done = false;
        return (int)G16_OK;
    });
}

int g16_prove_partial(g16_zkey* zk, uint8_t* partial_out) {
    return guarded([&]() {
        if (!zk || !partial_out) return fail(G16_ERR_GENERIC, "null argument");
        prove_device(zk);
        export_partial(zk, partial_out);
        return (int)G16_OK;
    });
}

int g16_prove_begin(g16_zkey* zk) {
    return guarded([&]() {
        if (!zk) return fail(G16_ERR_GENERIC, "null handle");
        if (zk->witness_enqueued) return (int)G16_OK;
        prove_device(zk, nullptr, true);
        return (int)G16_OK;
    });
}

int g16_h_transform(g16_zkey* zk, int which) {
    return guarded([&]() {
        if (!zk || (which & ~7)) return fail(G16_ERR_GENERIC, "bad argument");
        if (!zk->witness_loaded) return fail(G16_ERR_GENERIC, "no witness loaded");
        G16_CUDA(cudaSetDevice(zk->device));
        cudaStream_t s = zk->s_main;
        uint32_t n = zk->domain_size;
        if (!zk->witness_enqueued) G16_CUDA(cudaEventRecord(zk->ev[0], s));   // else g16_prove_begin marked the proof's start
        cudaEvent_t t0 = zk->ev[15];
        G16_CUDA(cudaEventRecord(t0, s));
        if (which) {   // a rank that owns no polynomial has nothing to evaluate
            r1cs_eval_kernel<<<(n + 127) / 128, 128, 0, s>>>(zk->row_ptr.as<uint32_t>(), zk->col.as<uint32_t>(),
                                                             zk->val.as<fe_t>(), zk->w.as<fe_t>(), n, zk->pa.as<fe_t>(),
                                                             zk->pb.as<fe_t>(), zk->pc.as<fe_t>());
            G16_CUDA(cudaGetLastError());
            G16_LAUNCHED(1);
        }
        G16_CUDA(cudaEventRecord(zk->ev[1], s));
        if (which & 1) zk->dom.to_odd_coset(zk->pa.as<fe_t>(), s);
        if (which & 2) zk->dom.to_odd_coset(zk->pb.as<fe_t>(), s);
        if (which & 4) zk->dom.to_odd_coset(zk->pc.as<fe_t>(), s);
        G16_CUDA(cudaEventRecord(zk->ev[2], s));
        G16_CUDA(cudaStreamSynchronize(s));
        float ms = 0;
        cudaEventElapsedTime(&ms, t0, zk->ev[1]); zk->t_transform_ms[0] = ms;
        cudaEventElapsedTime(&ms, zk->ev[1], zk->ev[2]); zk->t_transform_ms[1] = ms;
        return (int)G16_OK;
    });
}

int g16_h_transform_half(g16_zkey* zk, int poly, int parity) {
    return guarded([&]() {
        if (!zk || poly < 0 || poly > 2 || (parity & ~1)) return fail(G16_ERR_GENERIC, "bad argument");
        if (!zk->witness_loaded) return fail(G16_ERR_GENERIC, "no witness loaded");
        if (zk->logn < 2) return fail(G16_ERR_GENERIC, "a split transform needs a domain of at least 4 points");
        G16_CUDA(cudaSetDevice(zk->device));
        cudaStream_t s = zk->s_main;
        uint32_t n = zk->domain_size, half = n / 2;
        if (zk->dom_half.logn != zk->logn - 1) zk->dom_half.init(zk->logn - 1, s, false);   // first call only (allocates)
        if (!zk->witness_enqueued) G16_CUDA(cudaEventRecord(zk->ev[0], s));
        cudaEvent_t t0 = zk->ev[15];
        G16_CUDA(cudaEventRecord(t0, s));
        r1cs_eval_kernel<<<(n + 127) / 128, 128, 0, s>>>(zk->row_ptr.as<uint32_t>(), zk->col.as<uint32_t>(),
                                                         zk->val.as<fe_t>(), zk->w.as<fe_t>(), n, zk->pa.as<fe_t>(),
                                                         zk->pb.as<fe_t>(), zk->pc.as<fe_t>());
        G16_CUDA(cudaGetLastError());
        G16_CUDA(cudaEventRecord(zk->ev[1], s));
        fe_t* a = poly == 0 ? zk->pa.as<fe_t>() : poly == 1 ? zk->pb.as<fe_t>() : zk->pc.as<fe_t>();
        ntt_split_kernel<<<(half + 255) / 256, 256, 0, s>>>(a, half, parity, zk->dom.tw.as<fe_t>());
        G16_CUDA(cudaGetLastError());
        zk->dom_half.inverse_dif_with(a, zk->dom.scale.as<fe_t>() + size_t(parity) * half, s);
        zk->dom_half.forward_dit(a, s);
        G16_CUDA(cudaGetLastError());
        // the half twice, so that any range [lo, lo + count) of the buffer reads Y[k mod n/2]: slices cut like a whole transform's
        G16_CUDA(cudaMemcpyAsync(a + half, a, size_t(half) * sizeof(fe_t), cudaMemcpyDeviceToDevice, s));
        G16_LAUNCHED(2 + 2 * ntt_plan(zk->logn - 1).size());
        G16_CUDA(cudaEventRecord(zk->ev[2], s));
        G16_CUDA(cudaStreamSynchronize(s));
        float ms = 0;
        cudaEventElapsedTime(&ms, t0, zk->ev[1]); zk->t_transform_ms[0] = ms;
        cudaEventElapsedTime(&ms, zk->ev[1], zk->ev[2]); zk->t_transform_ms[1] = ms;
        return (int)G16_OK;
    });
}

int g16_prove_partial_joined_halves(g16_zkey* zk, const void* const d_halves[6], uint8_t* partial_out) {
    return guarded([&]() {
        if (!zk || !partial_out || !d_halves) return fail(G16_ERR_GENERIC, "null argument");
        for (int i = 0; i < 6; i++)
            if (zk->H.n && !d_halves[i]) return fail(G16_ERR_GENERIC, "null argument");
        JoinedSlices js{(const fe_t*)d_halves[0], (const fe_t*)d_halves[2], (const fe_t*)d_halves[4],
                        (const fe_t*)d_halves[1], (const fe_t*)d_halves[3], (const fe_t*)d_halves[5]};
        if (!zk->H.n) js.oa = nullptr;
        prove_device(zk, &js);
        zk->t_ms[1] = zk->t_transform_ms[0];
        zk->t_ms[2] += zk->t_transform_ms[1];
        export_partial(zk, partial_out);
        return (int)G16_OK;
    });
}

int g16_h_buffer(g16_zkey* zk, int poly, void** d_ptr, uint64_t* h_lo, uint64_t* h_n) {
    if (!zk || poly < 0 || poly > 2) return fail(G16_ERR_GENERIC, "bad argument");
    if (d_ptr) *d_ptr = poly == 0 ? zk->pa.p : poly == 1 ? zk->pb.p : zk->pc.p;
    if (h_lo) *h_lo = zk->H.lo;
    if (h_n) *h_n = zk->H.n;
    return G16_OK;
}

int g16_prove_partial_joined(g16_zkey* zk, const void* d_a, const void* d_b, const void* d_c, uint8_t* partial_out) {
    return guarded([&]() {
        if (!zk || !partial_out || (zk->H.n && (!d_a || !d_b || !d_c))) return fail(G16_ERR_GENERIC, "null argument");
        JoinedSlices js{(const fe_t*)d_a, (const fe_t*)d_b, (const fe_t*)d_c};
        prove_device(zk, &js);
        zk->t_ms[1] = zk->t_transform_ms[0];                // report the transform stage run before the exchange
        zk->t_ms[2] += zk->t_transform_ms[1];
        export_partial(zk, partial_out);
        return (int)G16_OK;
    });
}

int g16_prove_combine(g16_zkey* zk, const uint8_t* partials, int count, char* proof_json, size_t* proof_len,
                      char* public_json, size_t* public_len, const g16_opts* opts) {
    return guarded([&]() {
        if (!zk || !partials || count < 1) return fail(G16_ERR_GENERIC, "bad argument");
        g16_opts o;
        if (opts) o = *opts;
        else g16_opts_default(&o);
        std::string proof, pub;
        combine(zk, partials, count, o, make_blind(zk, o), proof, pub);
        return emit(proof, pub, proof_json, proof_len, public_json, public_len);
    });
}

int g16_prove_loaded(g16_zkey* zk, char* proof_json, size_t* proof_len, char* public_json, size_t* public_len,
                     const g16_opts* opts) {
    return guarded([&]() {
        if (!zk) return fail(G16_ERR_GENERIC, "null handle");
        if (zk->shard_count != 1)
            return fail(G16_ERR_GENERIC, "handle holds a shard: use g16_prove_partial + g16_prove_combine");
        g16_opts o;
        if (opts) o = *opts;
        else g16_opts_default(&o);
        uint8_t partial[G16_PARTIAL_BYTES];
        auto blind = std::async(std::launch::async, [&] { return make_blind(zk, o); });  // overlaps the GPU work
        prove_device(zk);
        export_partial(zk, partial);
        std::string proof, pub;
        combine(zk, partial, 1, o, blind.get(), proof, pub);
        return emit(proof, pub, proof_json, proof_len, public_json, public_len);
    });
}

int g16_prove(g16_zkey* zk, const uint8_t* wtns, size_t wtns_len, char* proof_json, size_t* proof_len,
              char* public_json, size_t* public_len, const g16_opts* opts) {
    int rc = g16_witness_load(zk, wtns, wtns_len);
    if (rc != G16_OK) return rc;
    return g16_prove_loaded(zk, proof_json, proof_len, public_json, public_len, opts);
}

int g16_prove_files(const char* zkey_path, const char* wtns_path, const char* proof_path, const char* public_path,
                    const g16_opts* opts) {
    return guarded([&]() {
        if (!zkey_path || !wtns_path || !proof_path || !public_path) return fail(G16_ERR_GENERIC, "null argument");
        g16_opts o;
        if (opts) o = *opts;
        else g16_opts_default(&o);
        o.shard_count = 1;
        if (o.h_table == 0) o.h_table = -1;   // one proof per key: building the window table costs more than it saves
        // G16_VERBOSE: where the wall time of one command-line call goes (the reference's figure is GNU time around the
        // whole process, scripts/lib/cmd_executor.sh:6,17)
        const bool verbose = std::getenv("G16_VERBOSE") != nullptr;
        auto t0 = std::chrono::steady_clock::now();
        auto lap = [&](const char* what) {
            if (!verbose) return;
            auto t1 = std::chrono::steady_clock::now();
            std::fprintf(stderr, "prover: %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
            t0 = t1;
        };
        MappedFile fz(zkey_path), fw(wtns_path);
        lap("map files");
        set_device(o.device);
        G16_CUDA(cudaFree(nullptr));
        lap("CUDA context");
        std::unique_ptr<g16_zkey> zk(zkey_build(fz.p, fz.len, &o));
        lap("key: parse + upload + tables");
        witness_load(zk.get(), fw.p, fw.len);
        lap("witness: parse + upload");
        uint8_t partial[G16_PARTIAL_BYTES];
        auto blind = std::async(std::launch::async, [&] { return make_blind(zk.get(), o); });
        prove_device(zk.get());
        export_partial(zk.get(), partial);
        std::string proof, pub;
        combine(zk.get(), partial, 1, o, blind.get(), proof, pub);
        lap("prove (first call, cold)");
        write_file(proof_path, proof);
        write_file(public_path, pub);
        zk.reset();
        lap("write outputs + release");
        return (int)G16_OK;
    });
}

int g16_h_scalars(g16_zkey* zk, uint8_t* out) {
    return guarded([&]() {
        if (!zk || !out) return fail(G16_ERR_GENERIC, "null argument");
        if (!zk->witness_loaded) return fail(G16_ERR_GENERIC, "no witness loaded");
        G16_CUDA(cudaSetDevice(zk->device));
        enqueue_h_pipeline(zk);
        G16_CUDA(cudaMemcpyAsync(out, zk->pa.p, size_t(zk->domain_size) * 32, cudaMemcpyDeviceToHost, zk->s_main));
        G16_CUDA(cudaStreamSynchronize(zk->s_main));
        return (int)G16_OK;
    });
}

int g16_h_pipeline_dev(g16_zkey* zk) {
    return guarded([&]() {
        if (!zk) return fail(G16_ERR_GENERIC, "null argument");
        if (!zk->witness_loaded) return fail(G16_ERR_GENERIC, "no witness loaded");
        G16_CUDA(cudaSetDevice(zk->device));
        G16_CUDA(cudaEventRecord(zk->ev[0], zk->s_main));
        enqueue_h_pipeline(zk);
        G16_CUDA(cudaStreamSynchronize(zk->s_main));
        float ms = 0;
        cudaEventElapsedTime(&ms, zk->ev[0], zk->ev[1]); zk->t_ms[1] = ms;
        cudaEventElapsedTime(&ms, zk->ev[1], zk->ev[2]); zk->t_ms[2] = ms;
        return (int)G16_OK;
    });
}

int g16_last_timings(const g16_zkey* zk, double out_ms[10]) {
    if (!zk || !out_ms) return fail(G16_ERR_GENERIC, "null argument");
    for (int i = 0; i < 10; i++) out_ms[i] = zk->t_ms[i];
    return G16_OK;
}

int g16_last_accumulate_ms(const g16_zkey* zk, double out[5]) {
    if (!zk || !out) return fail(G16_ERR_GENERIC, "null argument");
    out[0] = zk->A.msm.accumulate_ms();
    out[1] = zk->B1.msm.accumulate_ms();
    out[2] = zk->B2.msm.accumulate_ms();
    out[3] = zk->C.msm.accumulate_ms();
    out[4] = zk->H.msm.accumulate_ms();
    return G16_OK;
}

int g16_last_trace(const g16_zkey* zk, double out[30]) {
    if (!zk || !out) return fail(G16_ERR_GENERIC, "null argument");
    float t[6];
    auto put = [&](int i, const auto& msm) {
        msm.trace(zk->ev[0], t);
        for (int k = 0; k < 6; k++) out[6 * i + k] = t[k];
    };
    put(0, zk->A.msm); put(1, zk->B1.msm); put(2, zk->B2.msm); put(3, zk->C.msm); put(4, zk->H.msm);
    return G16_OK;
}

unsigned long long g16_launch_count(void) { return launch_counter().load(); }

int g16_last_msm_stats(const g16_zkey* zk, uint64_t out[25]) {
    if (!zk || !out) return fail(G16_ERR_GENERIC, "null argument");
    const MsmStats* st[5] = {&zk->A.msm.stats, &zk->B1.msm.stats, &zk->B2.msm.stats, &zk->C.msm.stats, &zk->H.msm.stats};
    for (int i = 0; i < 5; i++) {
        out[5 * i + 0] = st[i]->entries;
        out[5 * i + 1] = st[i]->tasks;
        out[5 * i + 2] = st[i]->cut_buckets;
        out[5 * i + 3] = (uint64_t)st[i]->c;
        out[5 * i + 4] = (uint64_t)st[i]->W;
    }
    return G16_OK;
}

}  // extern "C"
