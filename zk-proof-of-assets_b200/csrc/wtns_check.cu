// `snarkjs wtns check <circuit.r1cs> <witness.wtns>` on the GPU -- the witness check of scripts/g16_verify.sh:205-209.
//
// For every constraint c of the circuit: <A_c, w> * <B_c, w> == <C_c, w> over Fr (snarkjs 0.7.2 src/wtns_check.js; the
// reference runs it behind its -w flag before verifying a proof).  One thread per constraint: the three sparse dot
// products with the witness in Montgomery form (so that a coefficient in standard form times a witness value is one
// Montgomery product), one more product, one comparison; the lowest failing constraint index is kept with atomicMin.
// The host parses the .r1cs once into flat (wire, coefficient) arrays with three offsets per constraint.
// PARITY: oracle/wtns_check.py (Python big-int) gives the same verdict and the same first failing constraint
// (tests/test_gpu_wtns_check.py).  The wording of snarkjs' log lines is restated from the published source (UNPINNED: the
// reference commits no log of this step).  C ABI in include/g16_b200.h.
#include <cstring>
#include <memory>
#include <vector>

#include "../../include/g16_b200.h"
#include "binfile_ro.hpp"
#include "internal.cuh"

using namespace g16;
using namespace g16h;
using namespace g16files;

namespace {

__global__ void wtns_to_mont_kernel(const fe_t* __restrict__ w, uint32_t n, fe_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st_fe(out + i, Fr::to_mont(ld_fe(w + i)));
}

// off[3c], off[3c+1], off[3c+2], off[3c+3]: the term ranges of A_c, B_c, C_c
__global__ void __launch_bounds__(128)
wtns_check_kernel(const uint64_t* __restrict__ off, const uint32_t* __restrict__ wire, const fe_t* __restrict__ coef,
                  const fe_t* __restrict__ w_mont, uint32_t n_cons, unsigned long long* __restrict__ first_bad) {
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cons) return;
    fe_t v[3];
#pragma unroll 1
    for (int m = 0; m < 3; m++) {
        fe_t acc = Fr::zero();
        for (uint64_t t = off[3 * size_t(c) + m]; t < off[3 * size_t(c) + m + 1]; t++)
            acc = Fr::add(acc, Fr::mul(ld_fe(coef + t), ld_fe(w_mont + wire[t])));   // standard x Montgomery -> standard
        v[m] = acc;
    }
    fe_t ab = Fr::mul(Fr::to_mont(v[0]), v[1]);
    if (!Fr::eq(ab, v[2])) atomicMin(first_bad, (unsigned long long)c);
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

template <class T>
void upload(DevBuf& d, const std::vector<T>& v) {
    d.alloc(v.size() * sizeof(T));
    if (!v.empty()) G16_CUDA(cudaMemcpy(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
}

}  // namespace

extern "C" int g16_wtns_check(const char* r1cs_path, const char* wtns_path, const g16_opts* opts, int64_t* first_bad_constraint) {
    try {
        if (!r1cs_path || !wtns_path || !first_bad_constraint) return fail(G16_ERR_GENERIC, "null argument");
        *first_bad_constraint = -1;
        use_device(opts);
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
        uint32_t n_vars, n_cons;
        std::memcpy(&n_vars, &rh[36], 4);
        std::memcpy(&n_cons, &rh[60], 4);
        // ---- the witness (.wtns v2: sec1 { u32 n8 | prime | u32 nWitness }, sec2 values)
        FileRO fw(wtns_path);
        auto ws = scan_sections(fw, "wtns");
        const Sec& w1 = need_sec(ws, 1, "wtns");
        if (w1.size != 4 + 32 + 4) throw std::runtime_error("wtns: header section has the wrong size");
        uint8_t wh[40];
        fw.read(w1.off, wh, 40);
        std::memcpy(&n8, wh, 4);
        if (n8 != 32 || !u256_eq(u256_from_bytes(wh + 4), HFr::mod()))
            return fail(G16_ERR_CURVE, "Curve of the witness does not match the curve of the circuit");
        uint32_t nw;
        std::memcpy(&nw, wh + 36, 4);
        if (nw != n_vars)
            return fail(G16_ERR_WITNESS_LEN, "Invalid witness length. Circuit: " + std::to_string(n_vars) + ", witness: " + std::to_string(nw));
        const Sec& w2 = need_sec(ws, 2, "wtns");
        if (w2.size != uint64_t(nw) * 32) throw std::runtime_error("wtns: data section has the wrong size");
        std::vector<uint8_t> wit(w2.size);
        fw.read(w2.off, wit.data(), wit.size());
        const U256 rmod = HFr::mod();
        for (uint32_t i = 0; i < nw; i++)
            if (u256_cmp(u256_from_bytes(&wit[size_t(i) * 32]), rmod) >= 0)
                throw std::runtime_error("wtns: value " + std::to_string(i) + " is not below the field modulus");
        if (nw == 0 || !u256_eq(u256_from_bytes(wit.data()), U256{{1, 0, 0, 0}}))
            throw std::runtime_error("wtns: wire 0 is not the constant 1");
        // ---- the constraints, flattened
        const Sec& r2 = need_sec(rs, 2, "r1cs");
        std::vector<uint8_t> body(r2.size);
        fr.read(r2.off, body.data(), body.size());
        std::vector<uint64_t> off;
        std::vector<uint32_t> wire;
        std::vector<uint8_t> coef;
        off.reserve(3 * size_t(n_cons) + 1);
        wire.reserve(body.size() / 36);
        coef.reserve(body.size() / 36 * 32);
        size_t pos = 0;
        for (uint32_t c = 0; c < n_cons; c++) {
            for (int m = 0; m < 3; m++) {
                off.push_back(wire.size());
                if (pos + 4 > body.size()) throw std::runtime_error("r1cs: constraints section is truncated");
                uint32_t cnt;
                std::memcpy(&cnt, &body[pos], 4);
                pos += 4;
                if (uint64_t(cnt) * 36 > body.size() - pos) throw std::runtime_error("r1cs: constraints section is truncated");
                for (uint32_t e = 0; e < cnt; e++, pos += 36) {
                    uint32_t s;
                    std::memcpy(&s, &body[pos], 4);
                    if (s >= n_vars) throw std::runtime_error("r1cs: wire index out of range in constraint " + std::to_string(c));
                    if (u256_cmp(u256_from_bytes(&body[pos + 4]), rmod) >= 0) throw std::runtime_error("r1cs: coefficient not below the field modulus");
                    wire.push_back(s);
                    coef.insert(coef.end(), &body[pos + 4], &body[pos + 36]);
                }
            }
        }
        off.push_back(wire.size());
        if (pos != body.size()) throw std::runtime_error("r1cs: constraints section has the wrong size");
        body.clear();
        body.shrink_to_fit();
        if (!n_cons) return (int)G16_OK;
        // ---- the check
        DevBuf d_off, d_wire, d_coef, d_w, d_wm, d_bad;
        upload(d_off, off);
        upload(d_wire, wire);
        upload(d_coef, coef);
        upload(d_w, wit);
        d_wm.alloc(size_t(nw) * 32);
        d_bad.alloc(8);
        G16_CUDA(cudaMemset(d_bad.p, 0xff, 8));
        wtns_to_mont_kernel<<<(nw + 255) / 256, 256>>>(d_w.as<fe_t>(), nw, d_wm.as<fe_t>());
        wtns_check_kernel<<<(n_cons + 127) / 128, 128>>>(d_off.as<uint64_t>(), d_wire.as<uint32_t>(), d_coef.as<fe_t>(), d_wm.as<fe_t>(),
                                                         n_cons, d_bad.as<unsigned long long>());
        G16_CUDA(cudaGetLastError());
        G16_LAUNCHED(2);
        unsigned long long bad = 0;
        G16_CUDA(cudaMemcpy(&bad, d_bad.p, 8, cudaMemcpyDeviceToHost));
        if (bad != ~0ull) *first_bad_constraint = (int64_t)bad;
        return (int)G16_OK;
    } catch (const std::exception& e) {
        std::string m = e.what();
        return fail(m.rfind("CUDA error", 0) == 0 ? G16_ERR_CUDA : G16_ERR_FORMAT, "wtns check: " + m);
    }
}
