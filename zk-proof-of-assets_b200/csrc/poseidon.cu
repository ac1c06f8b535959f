// Poseidon (circom parameters, t = 3) and the anonymity-set Merkle tree on the GPU -- SURVEY.md 8(f) rank 4.
//
// Replaces the reference's scripts/merkle_tree.rs (light_poseidon Poseidon::<Fr>::new_circom(2) + rs_merkle; 2.5 h for a
// 10 M-leaf anonymity set on one core, scripts/README.md:913-915):
//   hasher   merkle_tree.rs:150-181   state [0, left, right], 8 full + 57 partial rounds of x^5, output state[0]
//   leaves   merkle_tree.rs:206-269   leaf = hash(address, balance); zero leaves up to the next power of two
//   tree     merkle_tree.rs:290-396   parent = hash(left, right); proof = siblings bottom-up, indices = index bits
// One thread per hash: 828 Montgomery products over Fr (195 round constants and the 3 x 3 MDS matrix in constant
// memory), the same 8 x 32-bit field arithmetic as the prover (field.cuh), integer-pipe bound.  The constants are not
// in the reference tree: they are regenerated on the host by the Poseidon paper's Grain-LFSR procedure and pinned on
// the reference's own roots and paths by the tests (tests/golden/merkle_vectors.json).  C ABI in include/g16_b200.h.
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "../../include/g16_b200.h"
#include "internal.cuh"

using namespace g16;
using namespace g16h;

// ---- g16_merkle_root_static: the tree's node levels in a fixed device array, for callers that cannot allocate ----
// Every level of a tree of up to G16_MERKLE_STATIC_MAX_ROWS leaves (2 * max * 32 bytes = 4 MiB).  Held by every GPU that
// loads the library until exit.  In host code the name below is a host-side stand-in, not a device address: the kernels
// are given the pointer cudaGetSymbolAddress returns, looked up once per device with that device current.
static __device__ __align__(256) unsigned char g_merkle_static_nodes[2 * size_t(G16_MERKLE_STATIC_MAX_ROWS) * 32];

namespace {

constexpr int POS_T = 3, POS_RF = 8, POS_RP = 57;
constexpr int POS_NC = (POS_RF + POS_RP) * POS_T;

__constant__ fe_t POS_C[POS_NC];
__constant__ fe_t POS_M[POS_T * POS_T];

__device__ __forceinline__ fe_t pow5(const fe_t& x) {
    fe_t x2 = Fr::sqr(x);
    return Fr::mul(Fr::sqr(x2), x);
}

// left, right: standard form; result: standard form
__device__ __forceinline__ fe_t poseidon2(const fe_t& left, const fe_t& right) {
    fe_t s0 = Fr::zero(), s1 = Fr::to_mont(left), s2 = Fr::to_mont(right);
#pragma unroll 1
    for (int r = 0; r < POS_RF + POS_RP; r++) {
        s0 = Fr::add(s0, POS_C[3 * r]);
        s1 = Fr::add(s1, POS_C[3 * r + 1]);
        s2 = Fr::add(s2, POS_C[3 * r + 2]);
        s0 = pow5(s0);
        if (r < POS_RF / 2 || r >= POS_RF / 2 + POS_RP) {
            s1 = pow5(s1);
            s2 = pow5(s2);
        }
        fe_t t0 = Fr::add(Fr::add(Fr::mul(POS_M[0], s0), Fr::mul(POS_M[1], s1)), Fr::mul(POS_M[2], s2));
        fe_t t1 = Fr::add(Fr::add(Fr::mul(POS_M[3], s0), Fr::mul(POS_M[4], s1)), Fr::mul(POS_M[5], s2));
        fe_t t2 = Fr::add(Fr::add(Fr::mul(POS_M[6], s0), Fr::mul(POS_M[7], s1)), Fr::mul(POS_M[8], s2));
        s0 = t0; s1 = t1; s2 = t2;
    }
    return Fr::from_mont(s0);
}

// out[i] = hash(in[2i], in[2i+1]), i < n; pad_from: outputs at or beyond it are the zero leaf (padding of the leaf level)
__global__ void __launch_bounds__(128) poseidon_pairs_kernel(const fe_t* __restrict__ in, fe_t* __restrict__ out, size_t n,
                                                            size_t n_out) {
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n_out) return;
    if (i >= n) {
        st_fe(out + i, Fr::zero());
        return;
    }
    st_fe(out + i, poseidon2(ld_fe(in + 2 * i), ld_fe(in + 2 * i + 1)));
}

// first index >= from with leaves[index] == target (the reference's forward scan, merkle_tree.rs:315-335)
__global__ void merkle_find_kernel(const fe_t* __restrict__ leaves, size_t n, size_t from, fe_t target,
                                   unsigned long long* __restrict__ found) {
    size_t i = from + blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    if (Fr::eq(ld_fe(leaves + i), target)) atomicMin(found, (unsigned long long)i);
}

// ---- the constants: Grain LFSR of the Poseidon reference generator (generate_parameters_grain.sage 1 0 254 3 8 57) ----
struct Grain {
    uint8_t st[80];
    int head = 0;   // st[(head + k) % 80] = bit k of the register
    Grain() {
        int pos = 0;
        auto put = [&](uint32_t v, int width) {
            for (int b = width - 1; b >= 0; b--) st[pos++] = (v >> b) & 1;
        };
        put(1, 2);              // prime field
        put(0, 4);              // S-box x^alpha
        put(254, 12);           // field size in bits
        put(POS_T, 12);
        put(POS_RF, 10);
        put(POS_RP, 10);
        for (int i = 0; i < 30; i++) st[pos++] = 1;
        for (int i = 0; i < 160; i++) clock();
    }
    int at(int k) const { return st[(head + k) % 80]; }
    int clock() {
        int b = at(62) ^ at(51) ^ at(38) ^ at(23) ^ at(13) ^ at(0);
        st[head] = (uint8_t)b;   // drop bit 0, append the new bit at the end
        head = (head + 1) % 80;
        return b;
    }
    int bit() {                  // self-shrinking output: a pair (keep, value)
        for (;;) {
            int keep = clock(), b = clock();
            if (keep) return b;
        }
    }
    U256 draw() {                // 254 bits, most significant first
        U256 v{{0, 0, 0, 0}};
        for (int i = 0; i < 254; i++) {
            for (int k = 3; k > 0; k--) v.v[k] = (v.v[k] << 1) | (v.v[k - 1] >> 63);
            v.v[0] = (v.v[0] << 1) | (uint64_t)bit();
        }
        return v;
    }
};

struct PoseidonParams {
    U256 c[POS_NC], m[POS_T * POS_T];   // Montgomery form
};

const PoseidonParams& host_params() {
    static PoseidonParams P = [] {
        PoseidonParams p;
        Grain g;
        int have = 0;
        while (have < POS_NC) {
            U256 v = g.draw();
            if (u256_cmp(v, HFr::mod()) < 0) p.c[have++] = HFr::to_mont(v);   // rejection sampling
        }
        for (;;) {   // Cauchy matrix 1 / (x_i + y_j) from 2t distinct field elements
            U256 v[2 * POS_T];
            bool ok = true;
            for (int i = 0; i < 2 * POS_T; i++) {
                v[i] = g.draw();
                while (u256_cmp(v[i], HFr::mod()) >= 0) v[i] = HFr::raw_sub(v[i], HFr::mod());
                for (int j = 0; j < i; j++) ok = ok && !u256_eq(v[i], v[j]);
                v[i] = HFr::to_mont(v[i]);
            }
            for (int i = 0; i < POS_T && ok; i++)
                for (int j = 0; j < POS_T; j++) {
                    U256 s = HFr::add(v[i], v[POS_T + j]);
                    if (u256_is_zero(s)) ok = false;
                    else p.m[i * POS_T + j] = HFr::inv(s);
                }
            if (ok) break;
        }
        return p;
    }();
    return P;
}

void upload_params() {   // once per device
    static std::mutex mu;
    static bool done[64] = {false};
    int dev = 0;
    G16_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    if (dev < 64 && done[dev]) return;
    const PoseidonParams& P = host_params();
    G16_CUDA(cudaMemcpyToSymbol(POS_C, P.c, sizeof(P.c)));
    G16_CUDA(cudaMemcpyToSymbol(POS_M, P.m, sizeof(P.m)));
    if (dev < 64) done[dev] = true;
}

void use_device(const g16_opts* o) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw std::runtime_error("CUDA error: no CUDA device: libg16b200 has no CPU fallback");
    if (o && o->device >= 0) {
        if (o->device >= count) throw std::runtime_error("CUDA error: device ordinal out of range");
        G16_CUDA(cudaSetDevice(o->device));
    }
}

template <class F>
int guarded(F&& f) {
    try {
        return f();
    } catch (const std::exception& e) {
        std::string m = e.what();
        return fail(m.rfind("CUDA error", 0) == 0 ? G16_ERR_CUDA : G16_ERR_GENERIC, m);
    }
}

void check_below_r(const uint8_t* p, size_t count) {
    for (size_t i = 0; i < count; i++)
        if (u256_cmp(u256_from_bytes(p + 32 * i), HFr::mod()) >= 0)
            throw std::runtime_error("poseidon: input " + std::to_string(i) + " is not below the field modulus");
}

void hash_level(const fe_t* in, fe_t* out, size_t n, size_t n_out, cudaStream_t s) {
    if (!n_out) return;
    poseidon_pairs_kernel<<<(unsigned)((n_out + 127) / 128), 128, 0, s>>>(in, out, n, n_out);
    G16_CUDA(cudaGetLastError());
    G16_LAUNCHED(1);
}

// ---- g16_merkle_root_static: set-up of the fixed device array declared at the top of this file ----
fe_t* static_nodes() {   // once per device; no allocation
    static std::mutex mu;
    static fe_t* addr[64] = {nullptr};
    int dev = 0;
    G16_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw std::runtime_error("CUDA error: device ordinal beyond the 64 this entry point tracks");
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_merkle_static_nodes);
        if (e != cudaSuccess || !p)
            throw std::runtime_error(std::string("CUDA error: cudaGetSymbolAddress(g_merkle_static_nodes): ") + cudaGetErrorString(e));
        addr[dev] = static_cast<fe_t*>(p);
    }
    return addr[dev];
}

}  // namespace

struct g16_merkle {
    int device = 0;
    uint64_t n_rows = 0, n_leaves = 0;   // rows hashed, leaves after padding (a power of two)
    uint32_t levels = 0;                 // node levels including the leaves and the root (what the reference prints as height)
    DevBuf nodes;                        // level 0 (leaves) first, then each level above; standard form
    std::vector<uint64_t> offset;        // first node of each level
    float build_ms = 0;
};

extern "C" {

int g16_poseidon_hash2(const uint8_t* pairs, size_t n, uint8_t* out, const g16_opts* opts) {
    return guarded([&]() {
        if ((!pairs || !out) && n) return fail(G16_ERR_GENERIC, "null argument");
        check_below_r(pairs, 2 * n);
        use_device(opts);
        upload_params();
        if (!n) return (int)G16_OK;
        DevBuf din, dout;
        din.alloc(n * 64);
        dout.alloc(n * 32);
        G16_CUDA(cudaMemcpy(din.p, pairs, n * 64, cudaMemcpyHostToDevice));
        hash_level(din.as<fe_t>(), dout.as<fe_t>(), n, n, 0);
        G16_CUDA(cudaMemcpy(out, dout.p, n * 32, cudaMemcpyDeviceToHost));
        return (int)G16_OK;
    });
}

int g16_merkle_build(const uint8_t* rows, size_t n_rows, const g16_opts* opts, g16_merkle** out) {
    return guarded([&]() {
        if (!out || (!rows && n_rows)) return fail(G16_ERR_GENERIC, "null argument");
        if (!n_rows) return fail(G16_ERR_GENERIC, "merkle: empty anonymity set");
        check_below_r(rows, 2 * n_rows);
        use_device(opts);
        upload_params();
        std::unique_ptr<g16_merkle> t(new g16_merkle());
        G16_CUDA(cudaGetDevice(&t->device));
        t->n_rows = n_rows;
        uint64_t size = 1;
        while (size < n_rows) size *= 2;
        t->n_leaves = size;
        for (uint64_t w = size, at = 0;; w /= 2) {
            t->offset.push_back(at);
            at += w;
            t->levels++;
            if (w == 1) break;
        }
        t->nodes.alloc((2 * size - 1) * 32);
        cudaEvent_t e0, e1;
        G16_CUDA(cudaEventCreate(&e0));
        G16_CUDA(cudaEventCreate(&e1));
        {
            // leaves in slabs, so that a 10 M-row set needs no second full-size staging buffer
            const size_t SLAB = size_t(1) << 22;
            DevBuf din;
            din.alloc((n_rows < SLAB ? n_rows : SLAB) * 64);
            fe_t* leaves = t->nodes.as<fe_t>();
            G16_CUDA(cudaEventRecord(e0));
            for (size_t off = 0; off < n_rows; off += SLAB) {
                size_t m = n_rows - off < SLAB ? n_rows - off : SLAB;
                G16_CUDA(cudaMemcpyAsync(din.p, rows + off * 64, m * 64, cudaMemcpyHostToDevice, 0));
                hash_level(din.as<fe_t>(), leaves + off, m, m, 0);
            }
            if (size > n_rows) G16_CUDA(cudaMemsetAsync(leaves + n_rows, 0, (size - n_rows) * 32, 0));
            for (uint32_t l = 1; l < t->levels; l++) {
                size_t w = size >> l;
                hash_level(t->nodes.as<fe_t>() + t->offset[l - 1], t->nodes.as<fe_t>() + t->offset[l], w, w, 0);
            }
            G16_CUDA(cudaEventRecord(e1));
            G16_CUDA(cudaEventSynchronize(e1));
        }
        cudaEventElapsedTime(&t->build_ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *out = t.release();
        return (int)G16_OK;
    });
}

void g16_merkle_close(g16_merkle* t) {
    if (t) cudaSetDevice(t->device);
    delete t;
}

int g16_merkle_info(const g16_merkle* t, uint64_t* n_rows, uint64_t* n_leaves, uint32_t* levels, double* build_ms) {
    if (!t) return fail(G16_ERR_GENERIC, "null handle");
    if (n_rows) *n_rows = t->n_rows;
    if (n_leaves) *n_leaves = t->n_leaves;
    if (levels) *levels = t->levels;
    if (build_ms) *build_ms = t->build_ms;
    return G16_OK;
}

int g16_merkle_node(const g16_merkle* t, uint32_t level, uint64_t index, uint8_t out[32]) {
    return guarded([&]() {
        if (!t || !out) return fail(G16_ERR_GENERIC, "null argument");
        if (level >= t->levels || index >= (t->n_leaves >> level)) return fail(G16_ERR_GENERIC, "merkle: node out of range");
        G16_CUDA(cudaSetDevice(t->device));
        G16_CUDA(cudaMemcpy(out, t->nodes.as<fe_t>() + t->offset[level] + index, 32, cudaMemcpyDeviceToHost));
        return (int)G16_OK;
    });
}

int g16_merkle_root(const g16_merkle* t, uint8_t out[32]) {
    if (!t) return fail(G16_ERR_GENERIC, "null handle");
    return g16_merkle_node(t, t->levels - 1, 0, out);
}

int g16_merkle_proof(const g16_merkle* t, uint64_t index, uint8_t* path_elements, uint8_t* path_indices) {
    return guarded([&]() {
        if (!t || !path_elements || !path_indices) return fail(G16_ERR_GENERIC, "null argument");
        if (index >= t->n_leaves) return fail(G16_ERR_GENERIC, "merkle: leaf out of range");
        G16_CUDA(cudaSetDevice(t->device));
        for (uint32_t l = 0; l + 1 < t->levels; l++) {
            G16_CUDA(cudaMemcpy(path_elements + 32 * l, t->nodes.as<fe_t>() + t->offset[l] + (index ^ 1), 32, cudaMemcpyDeviceToHost));
            path_indices[l] = uint8_t(index & 1);
            index >>= 1;
        }
        return (int)G16_OK;
    });
}

int g16_merkle_find(const g16_merkle* t, const uint8_t hash[32], uint64_t from, uint64_t* index) {
    return guarded([&]() {
        if (!t || !hash || !index) return fail(G16_ERR_GENERIC, "null argument");
        G16_CUDA(cudaSetDevice(t->device));
        *index = UINT64_MAX;
        if (from >= t->n_leaves) return (int)G16_OK;
        DevBuf found;
        found.alloc(8);
        G16_CUDA(cudaMemset(found.p, 0xff, 8));
        fe_t target;
        std::memcpy(&target, hash, 32);
        size_t span = t->n_leaves - from;
        merkle_find_kernel<<<(unsigned)((span + 255) / 256), 256>>>(t->nodes.as<fe_t>(), t->n_leaves, from, target,
                                                                     found.as<unsigned long long>());
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(1);
        unsigned long long h = 0;
        G16_CUDA(cudaMemcpy(&h, found.p, 8, cudaMemcpyDeviceToHost));
        *index = h;
        return (int)G16_OK;
    });
}

// The root of the anonymity-set tree with every buffer supplied: rows and root by the caller (device memory), the node
// levels in g_merkle_static_nodes.  Same kernels and level order as g16_merkle_build; calls no cudaMalloc / cudaFree,
// creates no stream or event, does not synchronise.  All calls on one GPU share the array: calls that overlap on the
// GPU race.  Nothing needs clearing between calls: the leaf launch writes all `size` leaves (hashes, then zero leaves),
// and each level above is written in full before it is read.
int g16_merkle_root_static(const void* d_rows, size_t n_rows, void* d_root, void* stream) {
    return guarded([&]() {
        if (!d_rows || !d_root) return fail(G16_ERR_GENERIC, "null argument");
        if (!n_rows) return fail(G16_ERR_GENERIC, "merkle: empty anonymity set");
        if (n_rows > size_t(G16_MERKLE_STATIC_MAX_ROWS))
            return fail(G16_ERR_GENERIC, "merkle: " + std::to_string(n_rows) + " rows exceed the " +
                                             std::to_string(G16_MERKLE_STATIC_MAX_ROWS) +
                                             " that g16_merkle_root_static's fixed device array holds; use g16_merkle_build");
        use_device(nullptr);   // the current device; fails loudly without one
        upload_params();       // cudaMemcpyToSymbol on a device's first call only
        fe_t* nodes = static_nodes();
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        size_t size = 1;
        while (size < n_rows) size *= 2;
        fe_t* root = static_cast<fe_t*>(d_root);
        // leaves: hashes of the rows, zero leaves up to `size` (written by the same launch, so stale contents never show)
        fe_t* level = size == 1 ? root : nodes;
        hash_level(static_cast<const fe_t*>(d_rows), level, n_rows, size, s);
        for (size_t w = size / 2; w >= 1; w /= 2) {
            fe_t* up = w == 1 ? root : level + 2 * w;
            hash_level(level, up, w, w, s);
            level = up;
        }
        return (int)G16_OK;
    });
}

}  // extern "C"
