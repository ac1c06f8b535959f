// `prover` -- command-line twin of rapidsnark's binary, the thing scripts/g16_prove.sh:246-252 execs:
//     prover <circuit.zkey> <witness.wtns> <proof.json> <public.json>
// (g16_prove.sh:195-199 requires the basename to be exactly "prover").  Exit status 0 = both files written;
// non-zero = failure with a message on stderr, which is all the caller's ERR trap looks at
// (scripts/lib/error_handling.sh:16-41).  Optional environment knobs never change the argv contract:
//   G16_DEVICE=<ordinal>   G16_STYLE=snarkjs|rapidsnark   G16_MSM_C=<bits>   G16_VERBOSE=1
//   G16_R=<decimal> G16_S=<decimal>   pin the blinding factors (test hook)
//   G16_SERVICE=<socket path>          send the request to a resident-key service (below) instead of proving in process
// Service mode (not part of the reference's contract, see csrc/service.hpp):
//     prover --serve [--socket PATH] [--devices 0,1,...] [--max-keys N]
//     prover --ping | --stats | --shutdown          (socket from --socket or G16_SERVICE)
// Verification (the argv of `snarkjs groth16 verify`, scripts/g16_verify.sh:212-216; no GPU needed):
//     prover --verify <verification_key.json> <public.json> <proof.json>     exit 0 + "OK!" | exit 1 + "Invalid proof"
//     prover --sanitize <verification_key.json> <public.json> <proof.json> <sanitized_proof.json>
//                                   the limb file scripts/sanitize_groth16_proof.py:112-123 writes, byte for byte
// Anonymity-set Merkle tree (the argv of the reference's `merkle-tree`, scripts/merkle_tree.rs:29-44):
//     prover --merkle-tree --anon-set <csv> --poa-input-data <json> --output-dir <dir>
// Witness check (the argv of `snarkjs wtns check`, scripts/g16_verify.sh:205-209):
//     prover --wtns-check <circuit.r1cs> <witness.wtns>       exit 0 + "WITNESS IS CORRECT" | exit 1 + "WITNESS CHECKING FAILED"
// Setup steps (snarkjs' argument order, scripts/g16_setup.sh:239-295):
//     prover --zkey-new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>
//     prover --zkey-contribute <in.zkey> <out.zkey> [--name=NAME] [-e=ENTROPY]
//     prover --zkey-beacon <in.zkey> <out.zkey> <beaconHash hex> <numIterationsExp> [-n=NAME]
//     prover --zkey-export-vkey <zkey> <verification_key.json>
//     prover --zkey-verify <circuit.r1cs> <pot_final.ptau> <circuit_final.zkey>      exit 0 + "ZKey Ok!" | exit 1 + the failed check
#include <unistd.h>

#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "g16_b200.h"
#include "service_handlers.hpp"

static const char* DEFAULT_SOCKET = "/tmp/g16b200.sock";

static std::string absolute(const char* p, bool must_exist) {
    char real[PATH_MAX];
    if (must_exist && ::realpath(p, real)) return real;
    if (p[0] == '/') return p;
    if (!::getcwd(real, sizeof real)) return p;
    return std::string(real) + "/" + p;
}

static int control(int argc, char** argv) {
    std::string sock = std::getenv("G16_SERVICE") ? std::getenv("G16_SERVICE") : DEFAULT_SOCKET;
    std::string devs;
    int max_keys = 4;
    std::string cmd = argv[1];
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--socket" && i + 1 < argc) sock = argv[++i];
        else if (a == "--devices" && i + 1 < argc) devs = argv[++i];
        else if (a == "--max-keys" && i + 1 < argc) max_keys = std::atoi(argv[++i]);
        else {
            std::fprintf(stderr, "prover: unknown option %s\n", a.c_str());
            return 1;
        }
    }
    if (cmd == "--serve") {
        std::vector<int> devices;
        if (devs.empty()) {
            if (const char* e = std::getenv("G16_DEVICE")) devs = e;
            else devs = "0";
        }
        size_t pos = 0;
        while (pos <= devs.size()) {
            size_t q = devs.find(',', pos);
            if (q == std::string::npos) q = devs.size();
            if (q > pos) devices.push_back(std::atoi(devs.substr(pos, q - pos).c_str()));
            pos = q + 1;
        }
        if (devices.empty()) devices.push_back(0);
        return g16svc::Server(sock, devices, max_keys).run();
    }
    std::string reply;
    std::string req = cmd == "--ping" ? "PING\n\n" : cmd == "--stats" ? "STATS\n\n" : "SHUTDOWN\n\n";
    if (!g16svc::exchange(sock, req, reply)) {
        std::fprintf(stderr, "prover: no service on %s\n", sock.c_str());
        return 1;
    }
    std::fputs(reply.c_str(), stdout);
    return reply.compare(0, 2, "OK") == 0 ? 0 : 1;
}

static int verify_cli(int argc, char** argv) {
    const bool sanitize = !std::strcmp(argv[1], "--sanitize");
    if (argc != (sanitize ? 6 : 5)) {
        std::fprintf(stderr, "Usage: prover --verify <verification_key.json> <public.json> <proof.json>\n"
                             "       prover --sanitize <verification_key.json> <public.json> <proof.json> <sanitized_proof.json>\n");
        return 1;
    }
    std::vector<uint8_t> f[3];
    std::string err;
    for (int i = 0; i < 3; i++)
        if (!g16svc::read_file(argv[2 + i], f[i], err)) {
            std::fprintf(stderr, "prover: %s\n", err.c_str());
            return 1;
        }
    if (sanitize) {
        std::vector<char> out(1 << 20);
        size_t n = out.size();
        int rc = g16_sanitize((const char*)f[0].data(), f[0].size(), (const char*)f[1].data(), f[1].size(),
                              (const char*)f[2].data(), f[2].size(), out.data(), &n);
        if (rc != G16_OK || !g16svc::write_file(argv[5], out.data(), n, err)) {
            std::fprintf(stderr, "prover: %s\n", rc != G16_OK ? g16_last_error() : err.c_str());
            return rc != G16_OK ? rc : 1;
        }
        return 0;
    }
    int ok = 0;
    int rc = g16_verify((const char*)f[0].data(), f[0].size(), (const char*)f[1].data(), f[1].size(),
                        (const char*)f[2].data(), f[2].size(), &ok);
    if (rc != G16_OK) {
        std::fprintf(stderr, "prover: %s\n", g16_last_error());
        return rc;
    }
    if (ok) std::printf("OK!\n");
    else std::fprintf(stderr, "Invalid proof\n");
    return ok ? 0 : 1;
}

// `merkle-tree` of the reference (scripts/merkle_tree.rs:29-44): -a/--anon-set, -p/--poa-input-data, -o/--output-dir
static int merkle_cli(int argc, char** argv) {
    const char *anon = nullptr, *poa = nullptr, *out = nullptr;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if ((a == "-a" || a == "--anon-set") && i + 1 < argc) anon = argv[++i];
        else if ((a == "-p" || a == "--poa-input-data") && i + 1 < argc) poa = argv[++i];
        else if ((a == "-o" || a == "--output-dir") && i + 1 < argc) out = argv[++i];
        else {
            std::fprintf(stderr, "Usage: prover --merkle-tree --anon-set <FILE_PATH> --poa-input-data <FILE_PATH> --output-dir <DIR_PATH>\n");
            return 2;
        }
    }
    if (!anon || !out) {
        std::fprintf(stderr, "Usage: prover --merkle-tree --anon-set <FILE_PATH> --poa-input-data <FILE_PATH> --output-dir <DIR_PATH>\n");
        return 2;
    }
    int rc = g16_merkle_files(anon, poa, out, nullptr);
    if (rc != G16_OK) {
        std::fprintf(stderr, "%s\n", g16_last_error());
        return 1;
    }
    return 0;
}

// the setup steps of scripts/g16_setup.sh:239-295 with snarkjs' argument order:
//   prover --zkey-new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>
//   prover --zkey-contribute <in.zkey> <out.zkey> [--name=NAME] [-e=ENTROPY]
//   prover --zkey-beacon <in.zkey> <out.zkey> <beaconHash hex> <numIterationsExp> [-n=NAME]
//   prover --zkey-export-vkey <zkey> <verification_key.json>
static int setup_cli(int argc, char** argv) {
    std::string cmd = argv[1];
    std::vector<std::string> pos;
    std::string name, entropy;
    bool has_entropy = false;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a.rfind("--name=", 0) == 0) name = a.substr(7);
        else if (a.rfind("-n=", 0) == 0) name = a.substr(3);
        else if (a.rfind("-e=", 0) == 0) { entropy = a.substr(3); has_entropy = true; }
        else pos.push_back(a);
    }
    auto hex = [](const uint8_t* h) {
        std::string s;
        char b[3];
        for (int i = 0; i < 64; i++) { std::snprintf(b, sizeof b, "%02x", h[i]); s += b; if (i % 4 == 3 && i != 63) s += i % 16 == 15 ? "\n\t\t" : " "; }
        return s;
    };
    uint8_t h[64];
    int rc;
    if (cmd == "--zkey-new" && pos.size() == 3) {
        rc = g16_zkey_new(pos[0].c_str(), pos[1].c_str(), pos[2].c_str(), nullptr, h);
        if (rc == G16_OK) std::printf("Circuit Hash: \n\t\t%s\n", hex(h).c_str());
    } else if (cmd == "--zkey-verify" && pos.size() == 3) {
        rc = g16_zkey_verify(pos[0].c_str(), pos[1].c_str(), pos[2].c_str(), nullptr);
        if (rc == G16_OK) std::printf("ZKey Ok!\n");
    } else if (cmd == "--zkey-contribute" && pos.size() == 2) {
        rc = g16_zkey_contribute(pos[0].c_str(), pos[1].c_str(), name.c_str(), has_entropy ? (const uint8_t*)entropy.data() : nullptr,
                                 entropy.size(), nullptr, h);
        if (rc == G16_OK) std::printf("Contribution Hash: \n\t\t%s\n", hex(h).c_str());
    } else if (cmd == "--zkey-beacon" && pos.size() == 4) {
        std::vector<uint8_t> b;
        for (size_t i = 0; i + 1 < pos[2].size(); i += 2) b.push_back((uint8_t)std::strtoul(pos[2].substr(i, 2).c_str(), nullptr, 16));
        rc = g16_zkey_beacon(pos[0].c_str(), pos[1].c_str(), name.c_str(), b.data(), b.size(), (uint32_t)std::atoi(pos[3].c_str()), nullptr, h);
        if (rc == G16_OK) std::printf("Contribution Hash: \n\t\t%s\n", hex(h).c_str());
    } else if (cmd == "--zkey-export-vkey" && pos.size() == 2) {
        std::vector<char> buf(1 << 16);
        size_t n = buf.size();
        rc = g16_zkey_export_vkey(pos[0].c_str(), buf.data(), &n);
        if (rc == G16_ERR_SHORT_BUFFER) {
            buf.resize(n + 1);
            n = buf.size();
            rc = g16_zkey_export_vkey(pos[0].c_str(), buf.data(), &n);
        }
        if (rc == G16_OK) {
            FILE* f = std::fopen(pos[1].c_str(), "wb");
            if (!f || std::fwrite(buf.data(), 1, n, f) != n || std::fclose(f) != 0) {
                std::fprintf(stderr, "cannot write %s\n", pos[1].c_str());
                return 1;
            }
        }
    } else {
        std::fprintf(stderr, "Usage: prover --zkey-new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>\n"
                             "       prover --zkey-contribute <in.zkey> <out.zkey> [--name=NAME] [-e=ENTROPY]\n"
                             "       prover --zkey-beacon <in.zkey> <out.zkey> <beaconHash hex> <numIterationsExp> [-n=NAME]\n"
                             "       prover --zkey-export-vkey <zkey> <verification_key.json>\n"
                             "       prover --zkey-verify <circuit.r1cs> <pot_final.ptau> <circuit_final.zkey>\n");
        return 2;
    }
    if (rc != G16_OK) {
        std::fprintf(stderr, "%s\n", g16_last_error());
        return 1;
    }
    return 0;
}

static int wtns_check_cli(int argc, char** argv) {
    if (argc != 4) {
        std::fprintf(stderr, "Usage: prover --wtns-check <circuit.r1cs> <witness.wtns>\n");
        return 2;
    }
    int64_t bad = -1;
    int rc = g16_wtns_check(argv[2], argv[3], nullptr, &bad);
    if (rc != G16_OK) {
        std::fprintf(stderr, "%s\n", g16_last_error());
        return rc;
    }
    if (bad >= 0) {
        std::printf("WITNESS CHECKING FAILED: constraint %lld is not satisfied\n", (long long)bad);
        return 1;
    }
    std::printf("WITNESS IS CORRECT\n");
    return 0;
}

int main(int argc, char** argv) {
    // every mode but --serve does ONE job in this process: the library's 32 hardware queues would save 0.7 ms of a proof
    // and cost 0.6 s of CUDA context creation.  Keep CUDA's default unless the user chose a value (the variable is read
    // at the first CUDA call, which comes later)
    if (std::getenv("G16_CONNECTIONS_DEFAULTED") && !(argc >= 2 && !std::strcmp(argv[1], "--serve")))
        ::setenv("CUDA_DEVICE_MAX_CONNECTIONS", "8", 1);
    if (argc >= 2 && !std::strcmp(argv[1], "--merkle-tree")) return merkle_cli(argc, argv);
    if (argc >= 2 && !std::strncmp(argv[1], "--zkey-", 7)) return setup_cli(argc, argv);
    if (argc >= 2 && !std::strcmp(argv[1], "--wtns-check")) return wtns_check_cli(argc, argv);
    if (argc >= 2 && (!std::strcmp(argv[1], "--verify") || !std::strcmp(argv[1], "--sanitize"))) return verify_cli(argc, argv);
    if (argc >= 2 && (!std::strcmp(argv[1], "--serve") || !std::strcmp(argv[1], "--ping") ||
                      !std::strcmp(argv[1], "--stats") || !std::strcmp(argv[1], "--shutdown")))
        return control(argc, argv);
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
        if (!g16svc::dec_to_le32(er, o.r) || !g16svc::dec_to_le32(es, o.s)) {
            std::fprintf(stderr, "prover: G16_R / G16_S must be decimal integers below 2^256\n");
            return 1;
        }
        o.has_rs = 1;
    }
    if (const char* sock = std::getenv("G16_SERVICE")) {
        std::string req = "PROVE\n" + absolute(argv[1], true) + "\n" + absolute(argv[2], true) + "\n" +
                          absolute(argv[3], false) + "\n" + absolute(argv[4], false) + "\n" +
                          (o.style == G16_STYLE_SNARKJS ? "1" : "0") + "\n" + (o.has_rs ? er : "-") + "\n" +
                          (o.has_rs ? es : "-") + "\n\n";
        std::string reply;
        if (g16svc::exchange(sock, req, reply)) {
            if (reply.compare(0, 2, "OK") == 0) {
                if (std::getenv("G16_VERBOSE")) std::fprintf(stderr, "prover: service %s", reply.c_str());
                return 0;
            }
            // "ERR <code> <message>"
            int code = 1;
            size_t sp = reply.find(' ', 4);
            if (reply.size() > 4) code = std::atoi(reply.c_str() + 4);
            std::string msg = sp == std::string::npos ? reply : reply.substr(sp + 1);
            while (!msg.empty() && msg.back() == '\n') msg.pop_back();
            std::fprintf(stderr, "prover: %s\n", msg.c_str());
            return code ? code : 1;
        }
        std::fprintf(stderr, "prover: no service on %s, proving in process\n", sock);
    }
    int rc = g16_prove_files(argv[1], argv[2], argv[3], argv[4], &o);
    if (rc != G16_OK) {
        std::fprintf(stderr, "prover: %s\n", g16_last_error());
        return rc;
    }
    if (std::getenv("G16_VERBOSE")) std::fprintf(stderr, "prover: %s wrote %s and %s\n", g16_version(), argv[3], argv[4]);
    // both files are complete and renamed into place: leave without tearing the CUDA context down (gigabytes of device
    // memory to free one allocation at a time, counted by the `time` the reference wraps around this command)
    std::fflush(nullptr);
    ::_exit(0);
}
