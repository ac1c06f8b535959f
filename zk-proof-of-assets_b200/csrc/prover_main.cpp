// `prover` -- command-line twin of rapidsnark's binary, the thing scripts/g16_prove.sh:246-252 execs:
//     prover <circuit.zkey> <witness.wtns> <proof.json> <public.json>
// (g16_prove.sh:195-199 requires the basename to be exactly "prover").  Exit status 0 = both files written;
// non-zero = failure with a message on stderr, which is all the caller's ERR trap looks at
// (scripts/lib/error_handling.sh:16-41).  Optional environment knobs never change the argv contract:
//   G16_DEVICE=<ordinal>   G16_STYLE=snarkjs|rapidsnark   G16_MSM_C=<bits>   G16_VERBOSE=1
//   G16_R=<decimal> G16_S=<decimal>   pin the blinding factors (test hook)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "g16_b200.h"

static bool dec_to_le32(const char* s, uint8_t out[32]) {
    std::memset(out, 0, 32);
    if (!*s) return false;
    for (; *s; s++) {
        if (*s < '0' || *s > '9') return false;
        unsigned carry = unsigned(*s - '0');
        for (int i = 0; i < 32; i++) {
            unsigned v = out[i] * 10u + carry;
            out[i] = uint8_t(v);
            carry = v >> 8;
        }
        if (carry) return false;
    }
    return true;
}

int main(int argc, char** argv) {
    if (argc != 5) {
        std::fprintf(stderr, "Invalid number of parameters\nUsage: prover <circuit.zkey> <witness.wtns> <proof.json> <public.json>\n");
        return 1;
    }
    g16_opts o;
    g16_opts_default(&o);
    if (const char* e = std::getenv("G16_DEVICE")) o.device = std::atoi(e);
    if (const char* e = std::getenv("G16_STYLE")) o.style = std::strcmp(e, "snarkjs") == 0 ? G16_STYLE_SNARKJS : G16_STYLE_RAPIDSNARK;
    if (const char* e = std::getenv("G16_MSM_C")) o.msm_c = std::atoi(e);
    const char* er = std::getenv("G16_R");
    const char* es = std::getenv("G16_S");
    if (er && es) {
        if (!dec_to_le32(er, o.r) || !dec_to_le32(es, o.s)) {
            std::fprintf(stderr, "prover: G16_R / G16_S must be decimal integers below 2^256\n");
            return 1;
        }
        o.has_rs = 1;
    }
    int rc = g16_prove_files(argv[1], argv[2], argv[3], argv[4], &o);
    if (rc != G16_OK) {
        std::fprintf(stderr, "prover: %s\n", g16_last_error());
        return rc;
    }
    if (std::getenv("G16_VERBOSE")) std::fprintf(stderr, "prover: %s wrote %s and %s\n", g16_version(), argv[3], argv[4]);
    return 0;
}
