// `snarkjs zkey new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>` on the GPU -- SURVEY.md 8(f) rank 2, first step.
//
// Replaces what scripts/g16_setup.sh:239-254 of the reference runs through snarkjs 0.7.2 (`groth16 setup`, src/zkey_new.js;
// the step before the contributions of csrc/setup.cpp).  From a circuit and a ceremony file prepared for phase 2 it
// builds the whole proving key: header, IC, coefficients, A, B1, B2, C, H and the contributions section with the
// circuit hash.  Everything heavy is a SPARSE LINEAR COMBINATION OF POINTS: with L_c the ceremony's Lagrange-form points
// of the circuit's domain (ptau sections 12-15),
//     A_s  = sum_{A[c][s]} coef L_c(tauG1)  (+ L_{nConstraints+s} for s <= nPublic)        B1_s, B2_s likewise over B
//     IC_s / C_s = sum_A coef L_c(beta tauG1) + sum_B coef L_c(alpha tauG1) + sum_C coef L_c(tauG1)  (+ the public rows)
//     H_i  = the odd Lagrange points of the domain of twice the size (copied);  gamma = delta = 1
// lincomb_chunk_kernel gives every chunk of at most 128 terms of one output point to a thread (signed double-and-add per
// term, XYZZ), lincomb_rows_kernel adds a point's chunks and converts to affine.  Integer-pipe bound like the prover.
// PARITY: oracle/zkey_new.py is the specification (byte for byte, tests/test_gpu_zkey_new.py); both are checked against the
// known-trapdoor factory, where gamma = delta = 1 makes `zkey new` a closed form.  The circuit hash is UNPINNED (no zkey
// in the reference to pin it on); see oracle/zkey_new.py.  C ABI in include/g16_b200.h.
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <thread>
#include <vector>

#include "../../include/g16_b200.h"
#include "internal.cuh"
#include "binfile_ro.hpp"
#include "setup_hash.hpp"

using namespace g16;
using namespace g16h;
using namespace g16files;

namespace {

constexpr uint32_t LC_CHUNK = 128;

struct RawScalar {
    uint32_t v[8];   // little-endian standard form, below r
};

// r and (r - 1) / 2, 32-bit limbs
__constant__ uint32_t ZN_R[8];
__constant__ uint32_t ZN_HALF[8];

__device__ __forceinline__ bool gt256(const uint32_t* a, const uint32_t* b) {
    for (int i = 7; i >= 0; i--)
        if (a[i] != b[i]) return a[i] > b[i];
    return false;
}

// partial[chunk] = sum over the chunk's terms of coef * table[src]
template <class C>
__global__ void __launch_bounds__(128)
lincomb_chunk_kernel(const uint64_t* __restrict__ order, const uint64_t* __restrict__ chunk_begin,
                     const uint32_t* __restrict__ chunk_len, size_t n_chunks,
                     const uint32_t* __restrict__ src, const RawScalar* __restrict__ coef,
                     const typename C::aff* __restrict__ table, typename C::pt* __restrict__ partial) {
    typedef typename C::field F;
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    size_t th = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (th >= n_chunks) return;
    const uint64_t t = order[th];   // cheap chunks first, full-width terms together at the end
    pt acc = C::identity();
    uint64_t at = chunk_begin[t];
    for (uint32_t e = 0; e < chunk_len[t]; e++, at++) {
        aff p = table[src[at]];
        if (C::is_inf(p)) continue;
        uint32_t k[8];
#pragma unroll
        for (int i = 0; i < 8; i++) k[i] = coef[at].v[i];
        if (gt256(k, ZN_HALF)) {   // r - k is the short one: add k P as -((r - k) P)
            uint32_t borrow = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                uint64_t d = uint64_t(ZN_R[i]) - k[i] - borrow;
                k[i] = uint32_t(d);
                borrow = uint32_t(d >> 63);
            }
            p.y = F::neg(p.y);
        }
        int top = -1;
        for (int i = 7; i >= 0 && top < 0; i--)
            if (k[i]) top = 32 * i + 31 - __clz(k[i]);
        if (top < 0) continue;     // coefficient zero
        if (top == 0) {            // +-1: the common case in a circuit
            C::madd(acc, p);
            continue;
        }
        pt q = C::from_affine(p);
        for (int bit = top - 1; bit >= 0; bit--) {
            q = C::dbl(q);
            if ((k[bit >> 5] >> (bit & 31)) & 1) C::madd(q, p);
        }
        C::add(acc, q);
    }
    partial[t] = acc;
}

// out[row] = affine(sum of the row's chunk partials); rows without terms are the point at infinity (all-zero record)
template <class C>
__global__ void __launch_bounds__(128)
lincomb_rows_kernel(const uint64_t* __restrict__ row_chunk, size_t n_rows, const typename C::pt* __restrict__ partial,
                    typename C::aff* __restrict__ out) {
    typedef typename C::field F;
    typedef typename C::fe fe;
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    size_t r = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (r >= n_rows) return;
    pt acc = C::identity();
    for (uint64_t c = row_chunk[r]; c < row_chunk[r + 1]; c++) C::add(acc, partial[c]);
    aff a;
    if (C::is_identity(acc)) {
        a.x = F::zero();
        a.y = F::zero();
    } else {
        fe u = F::inv(acc.zzz);
        fe zi = F::mul(u, acc.zz);
        a.x = F::mul(acc.x, F::sqr(zi));
        a.y = F::mul(acc.y, u);
    }
    out[r] = a;
}

// out[i] = tau^(i+n) G1 - tau^i G1, i < count: the points the circuit hash covers in place of section 9
__global__ void __launch_bounds__(128)
tau_diff_kernel(const g1_aff* __restrict__ mono, uint32_t n, uint32_t count, g1_aff* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    g1_aff hi = mono[i + n], lo = mono[i];
    g1_pt acc = G1::identity();
    if (!G1::is_inf(hi)) acc = G1::from_affine(hi);
    if (!G1::is_inf(lo)) {
        lo.y = Fq::neg(lo.y);
        G1::madd(acc, lo);
    }
    g1_aff a;
    if (G1::is_identity(acc)) {
        a.x = Fq::zero();
        a.y = Fq::zero();
    } else {
        fe_t u = Fq::inv(acc.zzz);
        fe_t zi = Fq::mul(u, acc.zz);
        a.x = Fq::mul(acc.x, Fq::sqr(zi));
        a.y = Fq::mul(acc.y, u);
    }
    out[i] = a;
}

struct RawBuf {   // bytes without the zero fill of a std::vector (these buffers are gigabytes and written in full)
    std::unique_ptr<uint8_t[]> p;
    size_t n = 0;
    explicit RawBuf(size_t bytes) : p(new uint8_t[bytes ? bytes : 1]), n(bytes) {}
    uint8_t* data() { return p.get(); }
    size_t size() const { return n; }
    void free() { p.reset(); n = 0; }
};

void read_points_into(const FileRO& f, const Sec& s, uint64_t first, uint64_t count, size_t rec, const char* what, uint8_t* dst) {
    if ((first + count) * rec > s.size)
        throw std::runtime_error(std::string("ptau: section ") + what + " is too short for this circuit (powers of tau too small)");
    f.read(s.off + first * rec, dst, count * rec);
}

// ---- host side: the term lists ---------------------------------------------------------------------------------------
struct Terms {   // one output section's terms, in file order
    std::vector<uint32_t> row, src;
    std::vector<RawScalar> coef;
    void add(uint32_t r, uint32_t s, const RawScalar& c) { row.push_back(r); src.push_back(s); coef.push_back(c); }
};

struct Csr {     // the same, grouped by row and cut into chunks of at most LC_CHUNK terms
    std::vector<uint32_t> src;
    std::vector<RawScalar> coef;
    std::vector<uint64_t> chunk_begin, row_chunk;
    std::vector<uint32_t> chunk_len;
    std::vector<uint64_t> order;   // chunks in the order threads take them: the cheap ones, then the full-width terms
};

// a term whose coefficient (or its negative) needs more than 40 doublings: it gets a chunk, i.e. a thread, of its own, and
// those chunks run together at the end of the launch -- a warp is then all +-1 / small constants or all long scalars,
// instead of 31 lanes waiting for the one that drew a 254-bit coefficient
bool heavy_term(const RawScalar& k) {
    static const U256 r = HFr::mod();
    U256 v;
    std::memcpy(v.v, k.v, 32);
    U256 neg = HFr::raw_sub(r, v);
    const U256& m = u256_cmp(v, neg) <= 0 ? v : neg;
    return m.v[3] || m.v[2] || m.v[1] || (m.v[0] >> 40);
}

Csr group_rows(const Terms& t, size_t n_rows) {
    Csr c;
    const size_t nt = t.row.size();
    std::vector<uint8_t> heavy(nt);
    for (size_t i = 0; i < nt; i++) heavy[i] = heavy_term(t.coef[i]) ? 1 : 0;
    // per row: light terms first, then heavy ones
    std::vector<uint64_t> start(n_rows + 1, 0), n_heavy(n_rows, 0);
    for (size_t i = 0; i < nt; i++) {
        start[t.row[i] + 1]++;
        n_heavy[t.row[i]] += heavy[i];
    }
    for (size_t r = 0; r < n_rows; r++) start[r + 1] += start[r];
    c.src.resize(nt);
    c.coef.resize(nt);
    std::vector<uint64_t> cur_l(n_rows), cur_h(n_rows);
    for (size_t r = 0; r < n_rows; r++) {
        cur_l[r] = start[r];
        cur_h[r] = start[r + 1] - n_heavy[r];
    }
    for (size_t i = 0; i < nt; i++) {
        uint64_t at = heavy[i] ? cur_h[t.row[i]]++ : cur_l[t.row[i]]++;
        c.src[at] = t.src[i];
        c.coef[at] = t.coef[i];
    }
    c.row_chunk.resize(n_rows + 1);
    std::vector<uint64_t> late;
    for (size_t r = 0; r < n_rows; r++) {
        c.row_chunk[r] = c.chunk_begin.size();
        const uint64_t light_end = start[r + 1] - n_heavy[r];
        for (uint64_t at = start[r]; at < light_end; at += LC_CHUNK) {
            c.order.push_back(c.chunk_begin.size());
            c.chunk_begin.push_back(at);
            c.chunk_len.push_back(uint32_t(std::min<uint64_t>(LC_CHUNK, light_end - at)));
        }
        for (uint64_t at = light_end; at < start[r + 1]; at++) {
            late.push_back(c.chunk_begin.size());
            c.chunk_begin.push_back(at);
            c.chunk_len.push_back(1);
        }
    }
    c.row_chunk[n_rows] = c.chunk_begin.size();
    c.order.insert(c.order.end(), late.begin(), late.end());
    return c;
}

template <class T>
void upload(DevBuf& d, const std::vector<T>& v) {
    d.alloc(v.size() * sizeof(T));
    if (!v.empty()) G16_CUDA(cudaMemcpy(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
}

// out = n_rows affine records (the zkey's own layout)
template <class C>
std::vector<uint8_t> run_lincomb(const Csr& c, const void* d_table, size_t n_rows) {
    typedef typename C::aff aff;
    typedef typename C::pt pt;
    std::vector<uint8_t> out(n_rows * sizeof(aff));
    if (!n_rows) return out;
    DevBuf d_src, d_coef, d_cb, d_cl, d_rc, d_part, d_out, d_order;
    upload(d_order, c.order);
    upload(d_src, c.src);
    upload(d_coef, c.coef);
    upload(d_cb, c.chunk_begin);
    upload(d_cl, c.chunk_len);
    upload(d_rc, c.row_chunk);
    size_t nch = c.chunk_begin.size();
    d_part.alloc((nch ? nch : 1) * sizeof(pt));
    d_out.alloc(n_rows * sizeof(aff));
    if (nch) {
        lincomb_chunk_kernel<C><<<(unsigned)((nch + 127) / 128), 128>>>(d_order.as<uint64_t>(), d_cb.as<uint64_t>(), d_cl.as<uint32_t>(), nch, d_src.as<uint32_t>(),
                                                                        d_coef.as<RawScalar>(), static_cast<const aff*>(d_table), d_part.as<pt>());
        G16_CUDA(cudaGetLastError());
    }
    lincomb_rows_kernel<C><<<(unsigned)((n_rows + 127) / 128), 128>>>(d_rc.as<uint64_t>(), n_rows, d_part.as<pt>(), d_out.as<aff>());
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(nch ? 2 : 1);
    G16_CUDA(cudaMemcpy(out.data(), d_out.p, out.size(), cudaMemcpyDeviceToHost));
    return out;
}

// ---- host side: hashing and writing ------------------------------------------------------------------------------------
void be32(uint8_t* out, const uint8_t* lem) {   // Montgomery little-endian -> standard big-endian
    U256 v = HFq::from_mont(u256_from_bytes(lem));
    for (int i = 0; i < 32; i++) out[i] = uint8_t(v.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
}

bool all_zero(const uint8_t* p, size_t n) {
    for (size_t i = 0; i < n; i++)
        if (p[i]) return false;
    return true;
}

// ffjavascript toRprUncompressed: G1 = x | y, G2 = x.c1 | x.c0 | y.c1 | y.c0, big-endian standard form; infinity = 0x40, zeros.
// The conversion (a Montgomery reduction per coordinate) runs on the host's cores; the hash itself is sequential.
void hash_points(g16setup::Blake2b& h, const uint8_t* recs, size_t count, bool g2, bool with_count = true) {
    if (with_count) {
        uint8_t be[4] = {uint8_t(count >> 24), uint8_t(count >> 16), uint8_t(count >> 8), uint8_t(count)};
        h.update(be, 4);
    }
    const size_t rec = g2 ? 128 : 64;
    std::vector<uint8_t> buf(count * rec);
    auto convert = [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; i++) {
            const uint8_t* p = recs + i * rec;
            uint8_t* u = &buf[i * rec];
            if (all_zero(p, rec)) {
                std::memset(u, 0, rec);
                u[0] = 0x40;
            } else if (g2) {
                be32(u, p + 32); be32(u + 32, p); be32(u + 64, p + 96); be32(u + 96, p + 64);
            } else {
                be32(u, p); be32(u + 32, p + 32);
            }
        }
    };
    unsigned nth = std::thread::hardware_concurrency();
    if (nth > 16) nth = 16;
    if (nth < 2 || count < 4096) convert(0, count);
    else {
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nth; k++) th.emplace_back(convert, count * k / nth, count * (k + 1) / nth);
        for (auto& x : th) x.join();
    }
    h.update(buf.data(), buf.size());
}

// The circuit hash is sequential over the sections in a fixed order, and each section is ready long before the next one:
// a worker thread hashes the jobs in the order they are pushed while the GPU computes the following section.
class HashWorker {
public:
    HashWorker() : h_(64), th_([this] { run(); }) {}
    ~HashWorker() { stop(); }
    void push(const uint8_t* recs, size_t count, bool g2, bool with_count = true) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            q_.push_back(Job{recs, count, g2, with_count});
        }
        cv_.notify_one();
    }
    std::vector<uint8_t> finish() {   // waits for every pushed job
        stop();
        if (err_) std::rethrow_exception(err_);
        return h_.digest();
    }

private:
    struct Job { const uint8_t* recs; size_t count; bool g2, with_count; };
    void stop() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            done_ = true;
        }
        cv_.notify_one();
        if (th_.joinable()) th_.join();
    }
    void run() {
        try {
            for (;;) {
                Job j;
                {
                    std::unique_lock<std::mutex> lk(mu_);
                    cv_.wait(lk, [this] { return done_ || !q_.empty(); });
                    if (q_.empty()) return;
                    j = q_.front();
                    q_.pop_front();
                }
                hash_points(h_, j.recs, j.count, j.g2, j.with_count);
            }
        } catch (...) {
            err_ = std::current_exception();
        }
    }
    g16setup::Blake2b h_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> q_;
    bool done_ = false;
    std::exception_ptr err_;
    std::thread th_;   // last member: starts after everything above exists
};

void use_device(const g16_opts* o) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw std::runtime_error("CUDA error: no CUDA device: libg16b200 has no CPU fallback");
    if (o && o->device >= 0) {
        if (o->device >= count) throw std::runtime_error("CUDA error: device ordinal out of range");
        G16_CUDA(cudaSetDevice(o->device));
    }
}

// generator records (Montgomery little-endian): G1 = (1, 2), G2 = the standard BN254 generator
void g1_generator(uint8_t out[64]) {
    u256_to_bytes(HFq::from_u64(1), out);
    u256_to_bytes(HFq::from_u64(2), out + 32);
}

void g2_generator(uint8_t out[128]) {
    static const U256 g[4] = {
        U256{{0x46debd5cd992f6edull, 0x674322d4f75edaddull, 0x426a00665e5c4479ull, 0x1800deef121f1e76ull}},   // x.c0
        U256{{0x97e485b7aef312c2ull, 0xf1aa493335a9e712ull, 0x7260bfb731fb5d25ull, 0x198e9393920d483aull}},   // x.c1
        U256{{0x4ce6cc0166fa7daaull, 0xe3d1e7690c43d37bull, 0x4aab71808dcb408full, 0x12c85ea5db8c6debull}},   // y.c0
        U256{{0x55acdadcd122975bull, 0xbc4b313370b38ef3ull, 0xec9e99ad690c3395ull, 0x090689d0585ff075ull}},   // y.c1
    };
    for (int i = 0; i < 4; i++) u256_to_bytes(HFq::to_mont(g[i]), out + 32 * i);
}

}  // namespace

extern "C" int g16_zkey_new(const char* r1cs_path, const char* ptau_path, const char* zkey_out, const g16_opts* opts,
                            uint8_t cs_hash_out[64]) {
    try {
        if (!r1cs_path || !ptau_path || !zkey_out) return fail(G16_ERR_GENERIC, "null argument");
        use_device(opts);
        const bool verbose = std::getenv("G16_VERBOSE") != nullptr;
        auto t_last = std::chrono::steady_clock::now();
        auto phase = [&](const char* what) {   // G16_VERBOSE=1: where the time goes, on stderr
            auto now = std::chrono::steady_clock::now();
            if (verbose) std::fprintf(stderr, "[zkey new] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
            t_last = now;
        };
        // ---- the circuit
        FileRO fr(r1cs_path);
        auto rs = scan_sections(fr, "r1cs");
        const Sec& r1 = need_sec(rs, 1, "r1cs");
        if (r1.size < 4 + 32 + 16 + 8 + 4) throw std::runtime_error("r1cs: header section too short");
        std::vector<uint8_t> rh(r1.size);
        fr.read(r1.off, rh.data(), rh.size());
        uint32_t n8;
        std::memcpy(&n8, rh.data(), 4);
        if (n8 != 32 || !u256_eq(u256_from_bytes(rh.data() + 4), HFr::mod()))
            return fail(G16_ERR_CURVE, "r1cs: the circuit's field is not the scalar field of BN254");
        uint32_t n_vars, n_out, n_pub_in, n_prv, n_cons;
        std::memcpy(&n_vars, &rh[36], 4);
        std::memcpy(&n_out, &rh[40], 4);
        std::memcpy(&n_pub_in, &rh[44], 4);
        std::memcpy(&n_prv, &rh[48], 4);
        std::memcpy(&n_cons, &rh[60], 4);
        const uint32_t n_pub = n_out + n_pub_in;
        if (n_vars == 0 || uint64_t(n_pub) + 1 > n_vars) throw std::runtime_error("r1cs: more public signals than wires");
        const uint64_t need = uint64_t(n_cons) + n_pub + 1;
        int k = 1;
        while ((uint64_t(1) << k) < need) k++;
        if (k > 28) throw std::runtime_error("r1cs: circuit too big for the curve's 2^28 domain");
        const uint64_t n = uint64_t(1) << k;
        // ---- the ceremony
        FileRO fp(ptau_path);
        auto ps = scan_sections(fp, "ptau");
        const Sec& p1 = need_sec(ps, 1, "ptau");
        if (p1.size < 4 + 32 + 8) throw std::runtime_error("ptau: header section too short");
        std::vector<uint8_t> ph(p1.size);
        fp.read(p1.off, ph.data(), ph.size());
        std::memcpy(&n8, ph.data(), 4);
        if (n8 != 32 || !u256_eq(u256_from_bytes(ph.data() + 4), HFq::mod()))
            return fail(G16_ERR_CURVE, "ptau: not a BN254 powers-of-tau file");
        uint32_t power;
        std::memcpy(&power, &ph[36], 4);
        if (uint32_t(k) > power)
            return fail(G16_ERR_GENERIC, "circuit too big for this power of tau ceremony. " + std::to_string(need) + " > 2**" + std::to_string(power));
        for (uint32_t t : {12u, 13u, 14u, 15u})
            if (!ps.count(t)) return fail(G16_ERR_FORMAT, "ptau: not prepared for phase 2 (section " + std::to_string(t) + " missing)");
        // Lagrange points of the circuit's domain; one G1 table [tauG1 | alphaTauG1 | betaTauG1] for the mixed section
        RawBuf tab1(3 * n * 64), tab2(n * 128);
        read_points_into(fp, ps[12], n - 1, n, 64, "12", tab1.data());
        read_points_into(fp, ps[14], n - 1, n, 64, "14", tab1.data() + n * 64);
        read_points_into(fp, ps[15], n - 1, n, 64, "15", tab1.data() + 2 * n * 64);
        read_points_into(fp, ps[13], n - 1, n, 128, "13", tab2.data());
        phase("ceremony points read");
        // ---- the terms
        const Sec& r2 = need_sec(rs, 2, "r1cs");
        std::vector<uint8_t> body(r2.size);
        fr.read(r2.off, body.data(), body.size());
        Terms tA, tB, tM;
        std::vector<uint8_t> sec4(4);
        uint32_t n_coefs = 0;
        const U256 rmod = HFr::mod();
        size_t pos = 0;
        const U256 r3 = HFr::to_mont(HFr::to_mont(HFr::to_mont(U256{{1, 0, 0, 0}})));   // R^3 mod r
        {   // an upper bound of the term count, so that the lists grow without reallocation
            size_t cap = body.size() / 36 + size_t(n_pub) + 1;
            for (Terms* t : {&tA, &tB, &tM}) { t->row.reserve(cap); t->src.reserve(cap); t->coef.reserve(cap); }
            sec4.reserve(4 + cap * 44);
        }
        auto coef_record = [&](uint32_t m, uint32_t c, uint32_t s, const U256& v) {
            size_t at = sec4.size();
            sec4.resize(at + 44);
            std::memcpy(&sec4[at], &m, 4);
            std::memcpy(&sec4[at + 4], &c, 4);
            std::memcpy(&sec4[at + 8], &s, 4);
            u256_to_bytes(HFr::mul(v, r3), &sec4[at + 12]);   // x R^2 (one Montgomery product by R^3): snarkjs's convention for section 4
            n_coefs++;
        };
        for (uint32_t c = 0; c < n_cons; c++) {
            for (int m = 0; m < 3; m++) {
                if (pos + 4 > body.size()) throw std::runtime_error("r1cs: constraints section is truncated");
                uint32_t cnt;
                std::memcpy(&cnt, &body[pos], 4);
                pos += 4;
                if (uint64_t(cnt) * 36 > body.size() - pos) throw std::runtime_error("r1cs: constraints section is truncated");
                for (uint32_t e = 0; e < cnt; e++, pos += 36) {
                    uint32_t s;
                    std::memcpy(&s, &body[pos], 4);
                    if (s >= n_vars) throw std::runtime_error("r1cs: wire index out of range in constraint " + std::to_string(c));
                    U256 v = u256_from_bytes(&body[pos + 4]);
                    if (u256_cmp(v, rmod) >= 0) throw std::runtime_error("r1cs: coefficient not below the field modulus");
                    RawScalar rsx;
                    std::memcpy(rsx.v, &body[pos + 4], 32);
                    if (m == 0) {
                        tA.add(s, c, rsx);
                        tM.add(s, uint32_t(2 * n + c), rsx);      // beta tauG1
                        coef_record(0, c, s, v);
                    } else if (m == 1) {
                        tB.add(s, c, rsx);
                        tM.add(s, uint32_t(n + c), rsx);          // alpha tauG1
                        coef_record(1, c, s, v);
                    } else {
                        tM.add(s, c, rsx);
                    }
                }
            }
        }
        if (pos != body.size()) throw std::runtime_error("r1cs: constraints section has the wrong size");
        body.clear();
        body.shrink_to_fit();
        RawScalar one{{1, 0, 0, 0, 0, 0, 0, 0}};
        for (uint32_t s = 0; s <= n_pub; s++) {   // the public-input rows
            tA.add(s, n_cons + s, one);
            tM.add(s, uint32_t(2 * n + n_cons + s), one);
            coef_record(0, n_cons + s, s, U256{{1, 0, 0, 0}});
        }
        std::memcpy(sec4.data(), &n_coefs, 4);
        phase("circuit parsed");
        // ---- the sections, on the GPU
        {
            U256 half = rmod;
            half.v[0] -= 1;
            for (int i = 0; i < 4; i++) half.v[i] = (half.v[i] >> 1) | (i < 3 ? half.v[i + 1] << 63 : 0);
            G16_CUDA(cudaMemcpyToSymbol(ZN_R, rmod.v, 32));
            G16_CUDA(cudaMemcpyToSymbol(ZN_HALF, half.v, 32));
        }
        DevBuf d_tab1, d_tab2;
        d_tab1.alloc(tab1.size());
        d_tab2.alloc(tab2.size());
        G16_CUDA(cudaMemcpy(d_tab1.p, tab1.data(), tab1.size(), cudaMemcpyHostToDevice));
        G16_CUDA(cudaMemcpy(d_tab2.p, tab2.data(), tab2.size(), cudaMemcpyHostToDevice));
        tab1.free();
        tab2.free();
        // Order of work = order of the circuit hash (header points | IC | tau differences | C | A | B1 | B2): the hash worker
        // takes each section as soon as it is back from the GPU.  The buffers live until the end of this function; the
        // worker is declared after them, so it is joined before they go.
        std::vector<uint8_t> secA, secB1, secB2, secM;
        uint8_t alpha1[64], beta1[64], beta2[128], g1[64], g2[128];
        fp.read(need_sec(ps, 4, "ptau").off, alpha1, 64);
        fp.read(need_sec(ps, 5, "ptau").off, beta1, 64);
        fp.read(need_sec(ps, 6, "ptau").off, beta2, 128);
        g1_generator(g1);
        g2_generator(g2);
        const size_t n_ic = size_t(n_pub) + 1;
        std::vector<uint8_t> hdiff((n - 1) * 64);
        RawBuf secH(n * 64);
        HashWorker hasher;
        hasher.push(alpha1, 1, false, false);
        hasher.push(beta1, 1, false, false);
        hasher.push(beta2, 1, true, false);
        hasher.push(g2, 1, true, false);
        hasher.push(g1, 1, false, false);
        hasher.push(g2, 1, true, false);
        {
            Csr c = group_rows(tM, n_vars);
            tM = Terms();
            phase("IC/C grouped");
            secM = run_lincomb<G1>(c, d_tab1.p, n_vars);
            phase("IC/C on the GPU");
        }
        hasher.push(secM.data(), n_ic, false);
        // the points the circuit hash covers for H: tau^(i+n) - tau^i
        if (n > 1) {
            RawBuf mono((2 * n - 1) * 64);
            read_points_into(fp, need_sec(ps, 2, "ptau"), 0, 2 * n - 1, 64, "2", mono.data());
            DevBuf d_mono, d_out;
            d_mono.alloc(mono.size());
            G16_CUDA(cudaMemcpy(d_mono.p, mono.data(), mono.size(), cudaMemcpyHostToDevice));
            d_out.alloc(hdiff.size());
            tau_diff_kernel<<<(unsigned)((n - 1 + 127) / 128), 128>>>(d_mono.as<g1_aff>(), (uint32_t)n, (uint32_t)(n - 1), d_out.as<g1_aff>());
            G16_CUDA(cudaGetLastError());
            G16_LAUNCHED(1);
            G16_CUDA(cudaMemcpy(hdiff.data(), d_out.p, hdiff.size(), cudaMemcpyDeviceToHost));
        }
        hasher.push(hdiff.data(), n - 1, false);
        hasher.push(secM.data() + n_ic * 64, n_vars - n_ic, false);
        phase("tau differences");
        {
            Csr c = group_rows(tA, n_vars);
            tA = Terms();
            phase("A grouped");
            secA = run_lincomb<G1>(c, d_tab1.p, n_vars);
            phase("A on the GPU");
        }
        hasher.push(secA.data(), n_vars, false);
        {
            Csr c = group_rows(tB, n_vars);
            tB = Terms();
            phase("B grouped");
            secB1 = run_lincomb<G1>(c, d_tab1.p, n_vars);
            phase("B1 on the GPU");
            hasher.push(secB1.data(), n_vars, false);
            secB2 = run_lincomb<G2>(c, d_tab2.p, n_vars);
            phase("B2 on the GPU");
        }
        hasher.push(secB2.data(), n_vars, true);
        d_tab1.free();
        d_tab2.free();
        // H: the odd Lagrange points of the domain of twice the size
        {
            RawBuf top(2 * n * 64);
            read_points_into(fp, ps[12], 2 * n - 1, 2 * n, 64, "12", top.data());
            for (uint64_t i = 0; i < n; i++) std::memcpy(secH.data() + i * 64, top.data() + (2 * i + 1) * 64, 64);
        }
        phase("H copied");
        // ---- circuit hash (what is left of it), header, file
        std::vector<uint8_t> cs = hasher.finish();
        phase("circuit hash");
        if (cs_hash_out) std::memcpy(cs_hash_out, cs.data(), 64);
        std::vector<uint8_t> hd;
        auto push = [&](const void* p, size_t len) { hd.insert(hd.end(), (const uint8_t*)p, (const uint8_t*)p + len); };
        uint32_t v32 = 32, dom = (uint32_t)n, proto = 1;
        uint8_t qb[32], rb[32];
        u256_to_bytes(HFq::mod(), qb);
        u256_to_bytes(rmod, rb);
        push(&v32, 4); push(qb, 32); push(&v32, 4); push(rb, 32);
        push(&n_vars, 4); push(&n_pub, 4); push(&dom, 4);
        push(alpha1, 64); push(beta1, 64); push(beta2, 128); push(g2, 128); push(g1, 64); push(g2, 128);
        cs.resize(68, 0);   // u32 0 contributions
        std::string tmp = std::string(zkey_out) + ".tmp." + std::to_string((long)getpid());
        {
            int fd = ::open(tmp.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
            if (fd < 0) throw std::runtime_error("cannot write " + tmp);
            auto put = [&](const void* data, uint64_t len) {
                const uint8_t* p = static_cast<const uint8_t*>(data);
                while (len) {
                    ssize_t w = ::write(fd, p, len > (uint64_t(1) << 30) ? size_t(1) << 30 : size_t(len));
                    if (w <= 0) { ::close(fd); ::unlink(tmp.c_str()); throw std::runtime_error("short write to " + tmp); }
                    p += w;
                    len -= (uint64_t)w;
                }
            };
            auto section = [&](uint32_t typ, const void* data, uint64_t size) {
                uint8_t h12[12];
                std::memcpy(h12, &typ, 4);
                std::memcpy(h12 + 4, &size, 8);
                put(h12, 12);
                put(data, size);
            };
            const uint8_t magic[12] = {'z', 'k', 'e', 'y', 1, 0, 0, 0, 10, 0, 0, 0};
            put(magic, 12);
            section(1, &proto, 4);
            section(2, hd.data(), hd.size());
            section(3, secM.data(), n_ic * 64);
            section(4, sec4.data(), sec4.size());
            section(5, secA.data(), secA.size());
            section(6, secB1.data(), secB1.size());
            section(7, secB2.data(), secB2.size());
            section(8, secM.data() + n_ic * 64, (n_vars - n_ic) * 64);
            section(9, secH.data(), secH.size());
            section(10, cs.data(), 68);
            if (::close(fd) != 0) { ::unlink(tmp.c_str()); throw std::runtime_error("short write to " + tmp); }
        }
        if (std::rename(tmp.c_str(), zkey_out) != 0) { ::unlink(tmp.c_str()); throw std::runtime_error(std::string("cannot rename onto ") + zkey_out); }
        phase("file written");
        return (int)G16_OK;
    } catch (const std::exception& e) {
        std::string m = e.what();
        return fail(m.rfind("CUDA error", 0) == 0 ? G16_ERR_CUDA : G16_ERR_FORMAT, "zkey new: " + m);
    }
}
