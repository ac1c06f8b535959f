// SYNTH-1 fixture generator (bench/test support, C ABI in include/g16_synth.h).
// Mirrors oracle/factory.py bit for bit: same circuit, same trapdoor derivation, same zkey conventions
// (snarkjs `zkey new`).  Host: Fr scalars.  GPU: fixed-base scalar multiplications.
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/g16_synth.h"
#include "common.cuh"
#include "curve.cuh"
#include "hostfield.hpp"

using namespace g16;
using namespace g16h;

namespace {

thread_local std::string g_err;
int sfail(const std::string& m) {
    g_err = m;
    return 1;
}

// ---- hashing (oracle/factory.py mix64 / h64 / hfield / hsmall) ------------------------------------------------
inline uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
inline uint64_t h64(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) {
    return mix64(mix64(mix64(seed ^ (tag * 0xD6E8FEB86659FD93ull)) + i) + k);
}
inline U256 hfield(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) {
    U256 v;
    for (int j = 0; j < 4; j++) v.v[j] = h64(seed, tag, i, 4 * k + j);
    v.v[3] &= (1ull << 61) - 1;  // 253 bits
    return v;
}
inline uint64_t hsmall(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) { return (h64(seed, tag, i, k) & 0x7FFFFFFFull) | 1; }

enum { TAG_BIT = 1, TAG_SMALL = 2, TAG_F1 = 3, TAG_F2 = 4, TAG_FILL = 5, TAG_TRAP = 100 };
enum { CL_ONE, CL_BIT, CL_SMALL, CL_F1, CL_F2 };

struct Term {
    uint32_t wire;
    U256 coef;  // standard form
};
struct Cons {
    Term A[2], B[2], C[2];
    int nA = 0, nB = 0, nC = 0;
};

inline U256 u64v(uint64_t x) { return U256{{x, 0, 0, 0}}; }

struct Circuit {
    uint32_t n_vars, n_public, n_constraints;
    uint64_t seed;
    uint32_t nblocks;
    int log_n;
    uint32_t domain;

    uint32_t pick(uint32_t lo, uint32_t hi, uint64_t h) const {
        uint32_t q = uint32_t((h & 0xFFFFFFFFull) % nblocks);
        uint32_t cls = lo + uint32_t((h >> 32) % (hi - lo));
        return n_public + 1 + 20 * q + cls;
    }
    int wire_class(uint32_t i) const {
        if (i == 0) return CL_ONE;
        if (i <= n_public) return CL_F2;
        uint32_t u = (i - n_public - 1) % 20;
        return u < 11 ? CL_BIT : u < 18 ? CL_SMALL : u == 18 ? CL_F1 : CL_F2;
    }
    void constraint(uint32_t c, Cons& k) const {
        k.nA = k.nB = k.nC = 0;
        const uint64_t sd = seed;
        if (c < n_vars - 1) {
            uint32_t i = c + 1;
            int cl = wire_class(i);
            if (cl == CL_BIT) {
                U256 m1 = HFr::mod();
                m1.v[0] -= 1;
                k.A[k.nA++] = Term{i, u64v(1)};
                k.B[k.nB++] = Term{i, u64v(1)};
                k.B[k.nB++] = Term{0, m1};
                return;
            }
            if (cl == CL_SMALL) {
                uint32_t b0 = pick(0, 11, h64(sd, TAG_SMALL, i, 0));
                uint32_t b1 = pick(0, 11, h64(sd, TAG_SMALL, i, 1));
                k.A[k.nA++] = Term{b0, u64v(hsmall(sd, TAG_SMALL, i, 2))};
                k.A[k.nA++] = Term{b1, u64v(hsmall(sd, TAG_SMALL, i, 3))};
                k.B[k.nB++] = Term{0, u64v(hsmall(sd, TAG_SMALL, i, 4))};
                k.C[k.nC++] = Term{i, u64v(1)};
                return;
            }
            if (cl == CL_F1) {
                uint32_t s0 = pick(0, 18, h64(sd, TAG_F1, i, 0));
                uint32_t s1 = pick(11, 18, h64(sd, TAG_F1, i, 1));
                uint32_t t0 = pick(11, 18, h64(sd, TAG_F1, i, 2));
                k.A[k.nA++] = Term{s0, hfield(sd, TAG_F1, i, 1)};
                k.A[k.nA++] = Term{s1, hfield(sd, TAG_F1, i, 2)};
                k.B[k.nB++] = Term{t0, hfield(sd, TAG_F1, i, 3)};
                k.C[k.nC++] = Term{i, u64v(1)};
                return;
            }
            uint32_t s0 = pick(18, 19, h64(sd, TAG_F2, i, 0));
            uint32_t s1 = pick(0, 19, h64(sd, TAG_F2, i, 1));
            uint32_t t0 = pick(18, 19, h64(sd, TAG_F2, i, 2));
            uint32_t t1 = pick(0, 11, h64(sd, TAG_F2, i, 3));
            k.A[k.nA++] = Term{s0, hfield(sd, TAG_F2, i, 1)};
            k.A[k.nA++] = Term{s1, u64v(hsmall(sd, TAG_F2, i, 12))};
            k.B[k.nB++] = Term{t0, hfield(sd, TAG_F2, i, 2)};
            k.B[k.nB++] = Term{t1, u64v(hsmall(sd, TAG_F2, i, 13))};
            k.C[k.nC++] = Term{i, u64v(1)};
            return;
        }
        uint32_t s0 = pick(0, 20, h64(sd, TAG_FILL, c, 0));
        uint32_t s1 = pick(0, 20, h64(sd, TAG_FILL, c, 1));
        U256 a0 = hfield(sd, TAG_FILL, c, 1), a1 = u64v(hsmall(sd, TAG_FILL, c, 8));
        k.A[k.nA++] = Term{s0, a0};
        k.A[k.nA++] = Term{s1, a1};
        k.B[k.nB++] = Term{0, u64v(1)};
        k.C[k.nC++] = Term{s0, a0};
        k.C[k.nC++] = Term{s1, a1};
    }
};

// all host values below are Montgomery-form Fr unless named *_std
typedef U256 fr;
inline fr M(const U256& std_) { return HFr::to_mont(std_); }

fr root_of_unity(int log) {
    U256 e = HFr::mod();
    e.v[0] -= 1;
    for (int i = 0; i < 4; i++) e.v[i] = (e.v[i] >> 28) | (i < 3 ? e.v[i + 1] << 36 : 0);
    fr w = HFr::pow(HFr::from_u64(5), e);
    for (int i = 28; i > log; i--) w = HFr::sqr(w);
    return w;
}

// [L_c(tau)] on the 2^log_n domain: w^c (tau^n - 1) / (n (tau - w^c))   (oracle/factory.py lagrange_at)
std::vector<fr> lagrange_at(const fr& tau, int log_n) {
    size_t n = size_t(1) << log_n;
    fr w = root_of_unity(log_n);
    fr tn = tau;
    for (int i = 0; i < log_n; i++) tn = HFr::sqr(tn);
    fr zt = HFr::sub(tn, HFr::one());
    fr nn = HFr::from_u64(uint64_t(n));
    std::vector<fr> pw(n), den(n), pref(n + 1);
    pw[0] = HFr::one();
    for (size_t c = 1; c < n; c++) pw[c] = HFr::mul(pw[c - 1], w);
    pref[0] = HFr::one();
    for (size_t c = 0; c < n; c++) {
        den[c] = HFr::mul(nn, HFr::sub(tau, pw[c]));
        pref[c + 1] = HFr::mul(pref[c], den[c]);
    }
    fr inv = HFr::inv(pref[n]);
    std::vector<fr> out(n);
    for (size_t c = n; c-- > 0;) {
        out[c] = HFr::mul(HFr::mul(pw[c], zt), HFr::mul(inv, pref[c]));
        inv = HFr::mul(inv, den[c]);
    }
    return out;
}

}  // namespace

struct g16_synth {
    Circuit circ;
    uint64_t td_seed;
    fr tau, alpha, beta, gamma, delta;  // Montgomery
    std::vector<fr> w;                  // witness, Montgomery
    std::vector<fr> a, b, c;            // a_i(tau), b_i(tau), c_i(tau)
    std::vector<fr> hs;                 // H_j exponents
    fr At, Bt, Ct, priv;
    bool points_done = false;
    std::vector<uint8_t> vk_points;     // Montgomery records: alpha1 beta1 beta2 gamma2 delta1 delta2 IC...
};

namespace {

void compute_scalars(g16_synth* s) {
    Circuit& ci = s->circ;
    uint64_t ts = s->td_seed;
    auto trap = [&](uint64_t i) {
        U256 v = hfield(ts, TAG_TRAP, i, 0);
        v.v[0] |= 2;
        return M(v);
    };
    s->tau = trap(0);
    s->alpha = trap(1);
    s->beta = trap(2);
    s->gamma = trap(3);
    s->delta = trap(4);
    uint32_t nv = ci.n_vars;
    // witness, by dependency level of the wire classes
    s->w.assign(nv, HFr::zero());
    s->w[0] = HFr::one();
    Cons k;
    for (int level = CL_BIT; level <= CL_F2; level++) {
        for (uint32_t i = 1; i < nv; i++) {
            if (ci.wire_class(i) != level) continue;
            if (level == CL_BIT) {
                s->w[i] = (h64(ci.seed, TAG_BIT, i, 0) & 1) ? HFr::one() : HFr::zero();
                continue;
            }
            ci.constraint(i - 1, k);
            fr va = HFr::zero(), vb = HFr::zero();
            for (int t = 0; t < k.nA; t++) va = HFr::add(va, HFr::mul(M(k.A[t].coef), s->w[k.A[t].wire]));
            for (int t = 0; t < k.nB; t++) vb = HFr::add(vb, HFr::mul(M(k.B[t].coef), s->w[k.B[t].wire]));
            s->w[i] = HFr::mul(va, vb);
        }
    }
    // QAP polynomials at tau
    std::vector<fr> L = lagrange_at(s->tau, ci.log_n);
    s->a.assign(nv, HFr::zero());
    s->b.assign(nv, HFr::zero());
    s->c.assign(nv, HFr::zero());
    s->Ct = HFr::zero();
    for (uint32_t cc = 0; cc < ci.n_constraints; cc++) {
        ci.constraint(cc, k);
        fr va = HFr::zero(), vb = HFr::zero();
        for (int t = 0; t < k.nA; t++) {
            fr co = M(k.A[t].coef);
            s->a[k.A[t].wire] = HFr::add(s->a[k.A[t].wire], HFr::mul(co, L[cc]));
            va = HFr::add(va, HFr::mul(co, s->w[k.A[t].wire]));
        }
        for (int t = 0; t < k.nB; t++) {
            fr co = M(k.B[t].coef);
            s->b[k.B[t].wire] = HFr::add(s->b[k.B[t].wire], HFr::mul(co, L[cc]));
            vb = HFr::add(vb, HFr::mul(co, s->w[k.B[t].wire]));
        }
        for (int t = 0; t < k.nC; t++)
            s->c[k.C[t].wire] = HFr::add(s->c[k.C[t].wire], HFr::mul(M(k.C[t].coef), L[cc]));
        s->Ct = HFr::add(s->Ct, HFr::mul(HFr::mul(va, vb), L[cc]));
    }
    for (uint32_t i = 0; i <= ci.n_public; i++) s->a[i] = HFr::add(s->a[i], L[ci.n_constraints + i]);
    s->At = s->Bt = s->priv = HFr::zero();
    for (uint32_t i = 0; i < nv; i++) {
        s->At = HFr::add(s->At, HFr::mul(s->w[i], s->a[i]));
        s->Bt = HFr::add(s->Bt, HFr::mul(s->w[i], s->b[i]));
        if (i > ci.n_public) {
            fr mixed = HFr::add(HFr::add(HFr::mul(s->beta, s->a[i]), HFr::mul(s->alpha, s->b[i])), s->c[i]);
            s->priv = HFr::add(s->priv, HFr::mul(s->w[i], mixed));
        }
    }
    // H_j = L^(2n)_{2j+1}(tau) / delta
    std::vector<fr> L2 = lagrange_at(s->tau, ci.log_n + 1);
    fr dinv = HFr::inv(s->delta);
    s->hs.resize(ci.domain);
    for (uint32_t j = 0; j < ci.domain; j++) s->hs[j] = HFr::mul(L2[2 * size_t(j) + 1], dinv);
}

// ---- GPU fixed-base multiplication ----------------------------------------------------------------------------
constexpr int FB_WINDOWS = 32;  // 8-bit windows over 256 bits

template <class C>
__global__ void fb_table_kernel(typename C::aff gen, typename C::pt* table) {
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= FB_WINDOWS) return;
    typename C::pt base = C::from_affine(gen);
    for (int i = 0; i < 8 * w; i++) base = C::dbl(base);
    typename C::pt acc = base;
    table[w * 255] = acc;
    for (int d = 1; d < 255; d++) {
        C::add(acc, base);
        table[w * 255 + d] = acc;
    }
}

template <class C>
__device__ __forceinline__ typename C::aff xyzz_to_affine(const typename C::pt& p) {
    typedef typename C::field F;
    typename C::aff a;
    if (C::is_identity(p)) {
        a.x = F::zero();
        a.y = F::zero();
        return a;
    }
    // ZZ^3 = ZZZ^2  =>  1/ZZ = (ZZ/ZZZ)^2
    typename F::el t = F::inv(p.zzz);
    typename F::el u = F::mul(p.zz, t);
    a.x = F::mul(p.x, F::sqr(u));
    a.y = F::mul(p.y, t);
    return a;
}

template <class C>
__global__ void fb_affine_kernel(const typename C::pt* in, typename C::aff* out, size_t n) {
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i < n) out[i] = xyzz_to_affine<C>(in[i]);
}

template <class C>
__global__ void __launch_bounds__(128)
fb_mul_kernel(const typename C::aff* __restrict__ table, const uint32_t* __restrict__ scalars, size_t n,
              typename C::aff* __restrict__ out) {
    size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    typename C::pt acc = C::identity();
    for (int w = 0; w < FB_WINDOWS; w++) {
        uint32_t d = (scalars[i * 8 + (w >> 2)] >> (8 * (w & 3))) & 0xff;
        if (d) C::madd(acc, table[w * 255 + d - 1]);
    }
    out[i] = xyzz_to_affine<C>(acc);
}

template <class C>
struct FixedBase {
    DevBuf table;
    void init(const typename C::aff& gen) {
        DevBuf xyzz;
        xyzz.alloc(size_t(FB_WINDOWS) * 255 * sizeof(typename C::pt));
        table.alloc(size_t(FB_WINDOWS) * 255 * sizeof(typename C::aff));
        fb_table_kernel<C><<<1, 32>>>(gen, xyzz.as<typename C::pt>());
        fb_affine_kernel<C><<<(FB_WINDOWS * 255 + 63) / 64, 64>>>(xyzz.as<typename C::pt>(), table.as<typename C::aff>(),
                                                                  size_t(FB_WINDOWS) * 255);
        G16_CUDA(cudaGetLastError());
        G16_CUDA(cudaDeviceSynchronize());
    }
    // scalars: host, n x 32 B standard form; out: host, n records
    void mul(const uint8_t* scalars, size_t n, uint8_t* out) {
        const size_t CH = size_t(1) << 22;
        DevBuf ds, dout;
        ds.alloc((n < CH ? n : CH) * 32);
        dout.alloc((n < CH ? n : CH) * sizeof(typename C::aff));
        for (size_t off = 0; off < n; off += CH) {
            size_t m = n - off < CH ? n - off : CH;
            G16_CUDA(cudaMemcpy(ds.p, scalars + off * 32, m * 32, cudaMemcpyHostToDevice));
            fb_mul_kernel<C><<<(unsigned)((m + 127) / 128), 128>>>(table.as<typename C::aff>(), ds.as<uint32_t>(), m,
                                                                   dout.as<typename C::aff>());
            G16_CUDA(cudaGetLastError());
            G16_CUDA(cudaMemcpy(out + off * sizeof(typename C::aff), dout.p, m * sizeof(typename C::aff),
                                cudaMemcpyDeviceToHost));
        }
    }
};

g1_aff g1_generator() {
    g1_aff g;
    U256 one = HFq::one(), two = HFq::add(one, one);
    std::memcpy(&g.x, one.v, 32);
    std::memcpy(&g.y, two.v, 32);
    return g;
}
g2_aff g2_generator() {
    const U256 c[4] = {{{0x46debd5cd992f6edull, 0x674322d4f75edaddull, 0x426a00665e5c4479ull, 0x1800deef121f1e76ull}},
                       {{0x97e485b7aef312c2ull, 0xf1aa493335a9e712ull, 0x7260bfb731fb5d25ull, 0x198e9393920d483aull}},
                       {{0x4ce6cc0166fa7daaull, 0xe3d1e7690c43d37bull, 0x4aab71808dcb408full, 0x12c85ea5db8c6debull}},
                       {{0x55acdadcd122975bull, 0xbc4b313370b38ef3ull, 0xec9e99ad690c3395ull, 0x090689d0585ff075ull}}};
    g2_aff g;
    U256 m[4];
    for (int i = 0; i < 4; i++) m[i] = HFq::to_mont(c[i]);
    std::memcpy(&g.x.c0, m[0].v, 32);
    std::memcpy(&g.x.c1, m[1].v, 32);
    std::memcpy(&g.y.c0, m[2].v, 32);
    std::memcpy(&g.y.c1, m[3].v, 32);
    return g;
}

inline void wr32(uint8_t*& p, uint32_t v) {
    std::memcpy(p, &v, 4);
    p += 4;
}
inline void wr64(uint8_t*& p, uint64_t v) {
    std::memcpy(p, &v, 8);
    p += 8;
}
inline void wr256(uint8_t*& p, const U256& v) {
    std::memcpy(p, v.v, 32);
    p += 32;
}

uint64_t count_coefs(const Circuit& ci) {
    uint64_t n = 0;
    Cons k;
    for (uint32_t c = 0; c < ci.n_constraints; c++) {
        ci.constraint(c, k);
        n += k.nA + k.nB;
    }
    return n + ci.n_public + 1;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
namespace {

struct Tables {
    FixedBase<G1> g1;
    FixedBase<G2> g2;
    bool ready = false;
    void ensure() {
        if (ready) return;
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
            throw std::runtime_error("no CUDA device: the fixture generator needs the GPU for its point multiplications");
        g1.init(g1_generator());
        g2.init(g2_generator());
        ready = true;
    }
};
thread_local std::unique_ptr<Tables> g_tables;
Tables& tables() {
    if (!g_tables) g_tables.reset(new Tables());
    g_tables->ensure();
    return *g_tables;
}

template <class F>
int sguard(F&& f) {
    try {
        return f();
    } catch (const std::exception& e) {
        return sfail(e.what());
    }
}

constexpr uint64_t SEC2_BYTES = 4 + 32 + 4 + 32 + 12 + 64 + 64 + 128 + 128 + 64 + 128;

struct Layout {
    uint64_t ncoef, off[11], size[11], total;
};
Layout layout_of(const Circuit& ci) {
    Layout l;
    l.ncoef = count_coefs(ci);
    uint64_t nv = ci.n_vars, np = ci.n_public, n = ci.domain;
    l.size[1] = 4;
    l.size[2] = SEC2_BYTES;
    l.size[3] = (np + 1) * 64;
    l.size[4] = 4 + l.ncoef * 44;
    l.size[5] = nv * 64;
    l.size[6] = nv * 64;
    l.size[7] = nv * 128;
    l.size[8] = (nv - np - 1) * 64;
    l.size[9] = n * 64;
    l.size[10] = 68;
    uint64_t pos = 12;
    for (int s = 1; s <= 10; s++) {
        l.off[s] = pos + 12;  // payload offset
        pos += 12 + l.size[s];
    }
    l.total = pos;
    return l;
}

// scalars (Montgomery Fr) -> records at dst via the GPU
template <class C>
void emit_points(FixedBase<C>& fb, const std::vector<fr>& mont_scalars, uint8_t* dst) {
    std::vector<uint8_t> sd(mont_scalars.size() * 32);
    for (size_t i = 0; i < mont_scalars.size(); i++) {
        U256 v = HFr::from_mont(mont_scalars[i]);
        std::memcpy(&sd[i * 32], v.v, 32);
    }
    fb.mul(sd.data(), mont_scalars.size(), dst);
}

void write_zkey(g16_synth* s, uint8_t* out) {
    Circuit& ci = s->circ;
    Layout l = layout_of(ci);
    Tables& T = tables();
    uint8_t* p = out;
    std::memcpy(p, "zkey", 4);
    p += 4;
    wr32(p, 1);
    wr32(p, 10);
    for (int sec = 1; sec <= 10; sec++) {
        uint8_t* h = out + l.off[sec] - 12;
        wr32(h, (uint32_t)sec);
        wr64(h, l.size[sec]);
    }
    uint32_t nv = ci.n_vars, np = ci.n_public;
    p = out + l.off[1];
    wr32(p, 1);
    // header points: alpha1 beta1 delta1 (G1), beta2 gamma2 delta2 (G2)
    uint8_t g1h[3 * 64], g2h[3 * 128];
    emit_points<G1>(T.g1, {s->alpha, s->beta, s->delta}, g1h);
    emit_points<G2>(T.g2, {s->beta, s->gamma, s->delta}, g2h);
    p = out + l.off[2];
    wr32(p, 32);
    wr256(p, HFq::mod());
    wr32(p, 32);
    wr256(p, HFr::mod());
    wr32(p, nv);
    wr32(p, np);
    wr32(p, ci.domain);
    std::memcpy(p, g1h, 64); p += 64;            // alpha1
    std::memcpy(p, g1h + 64, 64); p += 64;       // beta1
    std::memcpy(p, g2h, 128); p += 128;          // beta2
    std::memcpy(p, g2h + 128, 128); p += 128;    // gamma2
    std::memcpy(p, g1h + 128, 64); p += 64;      // delta1
    std::memcpy(p, g2h + 256, 128); p += 128;    // delta2
    // mixed_i = beta a_i + alpha b_i + c_i
    fr dinv = HFr::inv(s->delta), ginv = HFr::inv(s->gamma);
    std::vector<fr> sc(np + 1);
    auto mixed = [&](uint32_t i) {
        return HFr::add(HFr::add(HFr::mul(s->beta, s->a[i]), HFr::mul(s->alpha, s->b[i])), s->c[i]);
    };
    for (uint32_t i = 0; i <= np; i++) sc[i] = HFr::mul(mixed(i), ginv);
    emit_points<G1>(T.g1, sc, out + l.off[3]);
    // section 4
    p = out + l.off[4];
    wr32(p, (uint32_t)l.ncoef);
    Cons k;
    auto rec = [&](uint32_t m, uint32_t c, uint32_t sig, const U256& coef_std) {
        wr32(p, m);
        wr32(p, c);
        wr32(p, sig);
        wr256(p, HFr::to_mont(HFr::to_mont(coef_std)));  // coef * R^2
    };
    for (uint32_t c = 0; c < ci.n_constraints; c++) {
        ci.constraint(c, k);
        for (int t = 0; t < k.nA; t++) rec(0, c, k.A[t].wire, k.A[t].coef);
        for (int t = 0; t < k.nB; t++) rec(1, c, k.B[t].wire, k.B[t].coef);
    }
    for (uint32_t i = 0; i <= np; i++) rec(0, ci.n_constraints + i, i, u64v(1));
    emit_points<G1>(T.g1, s->a, out + l.off[5]);
    emit_points<G1>(T.g1, s->b, out + l.off[6]);
    emit_points<G2>(T.g2, s->b, out + l.off[7]);
    sc.resize(nv - np - 1);
    for (uint32_t i = np + 1; i < nv; i++) sc[i - np - 1] = HFr::mul(mixed(i), dinv);
    emit_points<G1>(T.g1, sc, out + l.off[8]);
    emit_points<G1>(T.g1, s->hs, out + l.off[9]);
    std::memset(out + l.off[10], 0, 68);
    // keep the verification-key points (Montgomery records)
    s->vk_points.resize(64 + 3 * 128 + size_t(np + 1) * 64);
    uint8_t* v = s->vk_points.data();
    std::memcpy(v, g1h, 64);
    std::memcpy(v + 64, g2h, 3 * 128);
    std::memcpy(v + 64 + 384, out + l.off[3], size_t(np + 1) * 64);
    s->points_done = true;
}

void write_wtns(const g16_synth* s, uint8_t* out) {
    uint8_t* p = out;
    std::memcpy(p, "wtns", 4);
    p += 4;
    wr32(p, 2);
    wr32(p, 2);
    wr32(p, 1);
    wr64(p, 40);
    wr32(p, 32);
    wr256(p, HFr::mod());
    wr32(p, s->circ.n_vars);
    wr32(p, 2);
    wr64(p, uint64_t(s->circ.n_vars) * 32);
    for (uint32_t i = 0; i < s->circ.n_vars; i++) wr256(p, HFr::from_mont(s->w[i]));
}

void mont_record_to_std(const uint8_t* in, uint8_t* out, int words) {
    for (int i = 0; i < words; i++) {
        U256 v = HFq::from_mont(u256_from_bytes(in + 32 * i));
        std::memcpy(out + 32 * i, v.v, 32);
    }
}

}  // namespace

extern "C" {

const char* g16_synth_last_error(void) { return g_err.c_str(); }

int g16_synth_create(uint32_t n_vars, uint32_t n_public, uint32_t n_constraints, uint64_t circuit_seed,
                     uint64_t trapdoor_seed, g16_synth** out) {
    return sguard([&]() {
        if (!out) return sfail("null argument");
        if (n_vars < n_public + 21 || n_constraints + 1 < n_vars) return sfail("SYNTH-1 needs nVars >= nPublic+21 and nConstraints >= nVars-1");
        std::unique_ptr<g16_synth> s(new g16_synth());
        Circuit& ci = s->circ;
        ci.n_vars = n_vars;
        ci.n_public = n_public;
        ci.n_constraints = n_constraints;
        ci.seed = circuit_seed;
        ci.nblocks = (n_vars - n_public - 1) / 20;
        uint64_t need = uint64_t(n_constraints) + n_public + 1;
        ci.log_n = 1;
        while ((uint64_t(1) << ci.log_n) < need) ci.log_n++;
        if (ci.log_n > 27) return sfail("domain too large");
        ci.domain = 1u << ci.log_n;
        s->td_seed = trapdoor_seed;
        compute_scalars(s.get());
        *out = s.release();
        return 0;
    });
}

void g16_synth_destroy(g16_synth* s) { delete s; }

int g16_synth_sizes(const g16_synth* s, uint64_t* zkey_bytes, uint64_t* wtns_bytes, uint32_t* domain_size) {
    if (!s) return sfail("null handle");
    if (zkey_bytes) *zkey_bytes = layout_of(s->circ).total;
    if (wtns_bytes) *wtns_bytes = 12 + 12 + 40 + 12 + uint64_t(s->circ.n_vars) * 32;
    if (domain_size) *domain_size = s->circ.domain;
    return 0;
}

int g16_synth_write(g16_synth* s, uint8_t* zkey_out, uint8_t* wtns_out) {
    return sguard([&]() {
        if (!s) return sfail("null handle");
        if (wtns_out) write_wtns(s, wtns_out);
        if (zkey_out) write_zkey(s, zkey_out);
        return 0;
    });
}

int g16_synth_expected(g16_synth* s, const uint8_t r[32], const uint8_t sblind[32], uint8_t out[256]) {
    return sguard([&]() {
        if (!s || !r || !sblind || !out) return sfail("null argument");
        fr rr = M(u256_from_bytes(r)), ss = M(u256_from_bytes(sblind));
        fr ea = HFr::add(HFr::add(s->alpha, s->At), HFr::mul(rr, s->delta));
        fr eb = HFr::add(HFr::add(s->beta, s->Bt), HFr::mul(ss, s->delta));
        fr dinv = HFr::inv(s->delta);
        fr ec = HFr::mul(HFr::sub(HFr::add(s->priv, HFr::mul(s->At, s->Bt)), s->Ct), dinv);
        ec = HFr::add(ec, HFr::mul(ss, ea));
        ec = HFr::add(ec, HFr::mul(rr, eb));
        ec = HFr::sub(ec, HFr::mul(HFr::mul(rr, ss), s->delta));
        Tables& T = tables();
        uint8_t g1r[128], g2r[128];
        emit_points<G1>(T.g1, {ea, ec}, g1r);
        emit_points<G2>(T.g2, {eb}, g2r);
        mont_record_to_std(g1r, out, 2);
        mont_record_to_std(g2r, out + 64, 4);
        mont_record_to_std(g1r + 64, out + 192, 2);
        return 0;
    });
}

int g16_synth_vkey(g16_synth* s, uint8_t* out) {
    return sguard([&]() {
        if (!s || !out) return sfail("null argument");
        if (!s->points_done) return sfail("call g16_synth_write with a zkey buffer first");
        mont_record_to_std(s->vk_points.data(), out, (int)(s->vk_points.size() / 32));
        return 0;
    });
}

// The circuit as an .r1cs file and a powers-of-tau file for it under the generator's tau, alpha, beta: the inputs of
// `zkey new` (g16_zkey_new).  The ceremony file has power = log2(domain) + 1 and the iden3 layout, but ONLY the ranges
// `zkey new` reads for this circuit are populated (tauG1 monomial 0 .. 2n-2, first points of sections 4 and 5, betaG2,
// Lagrange levels log2 n of sections 12-15 and log2 n + 1 of section 12); everything else is a hole in a sparse file.
int g16_synth_write_setup_inputs(g16_synth* s, const char* r1cs_path, const char* ptau_path) {
    return sguard([&]() {
        if (!s || !r1cs_path || !ptau_path) return sfail("null argument");
        Circuit& ci = s->circ;
        Tables& T = tables();
        // ---- .r1cs: header, constraints, wire-to-label map
        {
            std::vector<uint8_t> body;
            body.reserve(size_t(ci.n_constraints) * 5 * 36);
            Cons k;
            auto put_lc = [&](const Term* t, int n) {
                uint32_t cnt = (uint32_t)n;
                body.insert(body.end(), (uint8_t*)&cnt, (uint8_t*)&cnt + 4);
                for (int i = 0; i < n; i++) {
                    body.insert(body.end(), (const uint8_t*)&t[i].wire, (const uint8_t*)&t[i].wire + 4);
                    body.insert(body.end(), (const uint8_t*)t[i].coef.v, (const uint8_t*)t[i].coef.v + 32);
                }
            };
            for (uint32_t c = 0; c < ci.n_constraints; c++) {
                ci.constraint(c, k);
                put_lc(k.A, k.nA);
                put_lc(k.B, k.nB);
                put_lc(k.C, k.nC);
            }
            std::vector<uint8_t> head(4 + 32 + 16 + 8 + 4);
            uint8_t* p = head.data();
            wr32(p, 32);
            std::memcpy(p, HFr::mod().v, 32);
            p += 32;
            wr32(p, ci.n_vars);
            wr32(p, 0);
            wr32(p, ci.n_public);
            wr32(p, ci.n_vars - 1 - ci.n_public);
            uint64_t labels = ci.n_vars;
            std::memcpy(p, &labels, 8);
            p += 8;
            wr32(p, ci.n_constraints);
            std::vector<uint64_t> map(ci.n_vars);
            for (uint32_t i = 0; i < ci.n_vars; i++) map[i] = i;
            FILE* f = std::fopen(r1cs_path, "wb");
            if (!f) return sfail(std::string("cannot write ") + r1cs_path);
            auto sec = [&](uint32_t typ, const void* data, uint64_t size) {
                std::fwrite(&typ, 4, 1, f);
                std::fwrite(&size, 8, 1, f);
                if (size) std::fwrite(data, 1, size, f);
            };
            uint32_t ver = 1, nsec = 3;
            std::fwrite("r1cs", 1, 4, f);
            std::fwrite(&ver, 4, 1, f);
            std::fwrite(&nsec, 4, 1, f);
            sec(1, head.data(), head.size());
            sec(2, body.data(), body.size());
            sec(3, map.data(), map.size() * 8);
            if (std::fclose(f) != 0) return sfail(std::string("short write to ") + r1cs_path);
        }
        // ---- .ptau (sparse)
        const int k = ci.log_n, power = k + 1;
        const uint64_t n = ci.domain, P = uint64_t(1) << power;
        struct S { uint32_t typ; uint64_t size, off; };
        std::vector<S> secs = {{1, 4 + 32 + 8, 0}, {2, (2 * P - 1) * 64, 0}, {3, P * 128, 0}, {4, P * 64, 0}, {5, P * 64, 0},
                               {6, 128, 0}, {7, 4, 0}, {12, (2 * P - 1 + 2 * P) * 64, 0}, {13, (2 * P - 1) * 128, 0},
                               {14, (2 * P - 1) * 64, 0}, {15, (2 * P - 1) * 64, 0}};
        uint64_t pos = 12;
        for (S& x : secs) { x.off = pos + 12; pos += 12 + x.size; }
        FILE* f = std::fopen(ptau_path, "wb");
        if (!f) return sfail(std::string("cannot write ") + ptau_path);
        auto at = [&](uint64_t off, const void* data, size_t len) {
            if (fseeko(f, (off_t)off, SEEK_SET) != 0 || std::fwrite(data, 1, len, f) != len) throw std::runtime_error(std::string("cannot write ") + ptau_path);
        };
        auto off_of = [&](uint32_t typ) { for (S& x : secs) if (x.typ == typ) return x.off; return uint64_t(0); };
        {
            uint32_t ver = 1, nsec = (uint32_t)secs.size();
            at(0, "ptau", 4);
            at(4, &ver, 4);
            at(8, &nsec, 4);
            for (S& x : secs) { at(x.off - 12, &x.typ, 4); at(x.off - 8, &x.size, 8); }
            uint8_t h[4 + 32 + 8];
            uint8_t* p = h;
            wr32(p, 32);
            std::memcpy(p, HFq::mod().v, 32);
            p += 32;
            wr32(p, (uint32_t)power);
            wr32(p, (uint32_t)power);
            at(off_of(1), h, sizeof h);
            uint32_t zero = 0;
            at(off_of(7), &zero, 4);
        }
        std::vector<uint8_t> rec;
        auto g1_points = [&](const std::vector<fr>& sc, uint64_t off) {
            rec.resize(sc.size() * 64);
            emit_points<G1>(T.g1, sc, rec.data());
            at(off, rec.data(), rec.size());
        };
        auto scaled = [&](const std::vector<fr>& v, const fr& m) {
            std::vector<fr> o(v.size());
            for (size_t i = 0; i < v.size(); i++) o[i] = HFr::mul(v[i], m);
            return o;
        };
        {   // monomial tauG1: tau^0 .. tau^(2n-2)
            std::vector<fr> tp(2 * n - 1);
            tp[0] = HFr::one();
            for (size_t i = 1; i < tp.size(); i++) tp[i] = HFr::mul(tp[i - 1], s->tau);
            g1_points(tp, off_of(2));
        }
        g1_points({s->alpha}, off_of(4));
        g1_points({s->beta}, off_of(5));
        {
            uint8_t b2[128];
            emit_points<G2>(T.g2, {s->beta}, b2);
            at(off_of(6), b2, 128);
        }
        {   // Lagrange level k of sections 12-15, level k+1 of section 12
            std::vector<fr> L = lagrange_at(s->tau, k);
            g1_points(L, off_of(12) + (n - 1) * 64);
            g1_points(scaled(L, s->alpha), off_of(14) + (n - 1) * 64);
            g1_points(scaled(L, s->beta), off_of(15) + (n - 1) * 64);
            rec.resize(L.size() * 128);
            emit_points<G2>(T.g2, L, rec.data());
            at(off_of(13) + (n - 1) * 128, rec.data(), rec.size());
            g1_points(lagrange_at(s->tau, k + 1), off_of(12) + (2 * n - 1) * 64);
        }
        // the file's full length (the tail may be a hole)
        if (fseeko(f, (off_t)(pos - 1), SEEK_SET) != 0 || std::fwrite("", 1, 1, f) != 1) { std::fclose(f); return sfail("cannot size the ptau file"); }
        if (std::fclose(f) != 0) return sfail(std::string("short write to ") + ptau_path);
        return 0;
    });
}

int g16_synth_fixed_base(int is_g2, const uint8_t* scalars, size_t n, uint8_t* out) {
    return sguard([&]() {
        if (!scalars || !out) return sfail("null argument");
        Tables& T = tables();
        if (is_g2) T.g2.mul(scalars, n, out);
        else T.g1.mul(scalars, n, out);
        return 0;
    });
}

}  // extern "C"
