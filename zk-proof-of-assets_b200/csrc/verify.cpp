// g16_verify: the host-only entry point of the native verifier (csrc/verify*.hpp).  Compiled by g++, no CUDA.
#include <cstring>
#include <string>

#include "g16_b200.h"
#include "verify_groth16.hpp"

namespace g16 {
int fail(int code, const std::string& msg);  // prover.cu: sets g16_last_error() for this thread
}

extern "C" int g16_verify(const char* vkey_json, size_t vkey_len, const char* public_json, size_t public_len,
                          const char* proof_json, size_t proof_len, int* ok) {
    if (!vkey_json || !public_json || !proof_json || !ok) return g16::fail(G16_ERR_GENERIC, "null argument");
    *ok = 0;
    try {
        *ok = g16v::verify_groth16(vkey_json, vkey_len, public_json, public_len, proof_json, proof_len) ? 1 : 0;
        return G16_OK;
    } catch (const std::exception& e) {
        return g16::fail(G16_ERR_FORMAT, std::string("verify: ") + e.what());
    }
}

extern "C" int g16_sanitize(const char* vkey_json, size_t vkey_len, const char* public_json, size_t public_len,
                            const char* proof_json, size_t proof_len, char* out, size_t* out_len) {
    if (!vkey_json || !public_json || !proof_json || !out_len) return g16::fail(G16_ERR_GENERIC, "null argument");
    try {
        std::string s = g16v::sanitize_groth16(vkey_json, vkey_len, public_json, public_len, proof_json, proof_len);
        bool fits = out && *out_len >= s.size();
        if (fits) {
            std::memcpy(out, s.data(), s.size());
            if (*out_len > s.size()) out[s.size()] = 0;
        }
        *out_len = s.size();
        return fits ? (int)G16_OK : g16::fail(G16_ERR_SHORT_BUFFER, "output buffer too short");
    } catch (const std::exception& e) {
        return g16::fail(G16_ERR_FORMAT, std::string("sanitize: ") + e.what());
    }
}
