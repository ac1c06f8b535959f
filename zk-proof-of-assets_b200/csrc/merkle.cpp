// g16_merkle_files: the reference's `merkle-tree --anon-set F --poa-input-data F --output-dir D` as one call
// (scripts/merkle_tree.rs:398-421, run by scripts/full_workflow.sh:371-381).  Host side only: csv / JSON in,
// merkle_root.json + merkle_proofs.json out (serde_json pretty layout, scripts/README.md:633-673); every hash runs
// on the GPU through g16_merkle_* (csrc/poseidon.cu).  Compiled by g++, no CUDA.
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "g16_b200.h"
#include "verify_groth16.hpp"   // JParser, dec_u256, U256 helpers

namespace g16 {
int fail(int code, const std::string& msg);
}

namespace {
using namespace g16h;
using g16v::JParser;
using g16v::JV;

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

void write_file_atomic(const std::string& path, const std::string& text) {
    std::string tmp = path + ".tmp";
    {
        std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
        if (!f) throw std::runtime_error("cannot write " + path);
        f.write(text.data(), (std::streamsize)text.size());
        if (!f) throw std::runtime_error("short write to " + path);
    }
    if (std::rename(tmp.c_str(), path.c_str()) != 0) throw std::runtime_error("cannot rename onto " + path);
}

U256 parse_hex(const char* p, const char* e) {
    U256 r{{0, 0, 0, 0}};
    int digits = 0;
    for (; p < e; p++) {
        int d;
        if (*p >= '0' && *p <= '9') d = *p - '0';
        else if (*p >= 'a' && *p <= 'f') d = *p - 'a' + 10;
        else if (*p >= 'A' && *p <= 'F') d = *p - 'A' + 10;
        else throw std::runtime_error("anonymity set: bad hex digit in address");
        if (r.v[3] >> 60) throw std::runtime_error("anonymity set: address exceeds 256 bits");
        for (int k = 3; k > 0; k--) r.v[k] = (r.v[k] << 4) | (r.v[k - 1] >> 60);
        r.v[0] = (r.v[0] << 4) | (uint64_t)d;
        digits++;
    }
    if (!digits) throw std::runtime_error("anonymity set: empty address");
    return r;
}

U256 parse_dec(const char* p, const char* e) {
    JV v;
    v.kind = JV::NUM;
    v.s.assign(p, e);
    return g16v::dec_u256(v);
}

// csv with a heading line "address,eth_balance"; address = 0x-prefixed hex, balance = decimal (merkle_tree.rs:232-247)
std::vector<uint8_t> parse_anon_set(const std::string& text) {
    std::vector<uint8_t> rows;
    const char* p = text.data();
    const char* end = p + text.size();
    bool heading = true;
    while (p < end) {
        const char* nl = (const char*)std::memchr(p, '\n', size_t(end - p));
        const char* le = nl ? nl : end;
        const char* e = le;
        while (e > p && (e[-1] == '\r' || e[-1] == ' ')) e--;
        if (e > p) {
            if (heading) heading = false;
            else {
                const char* comma = (const char*)std::memchr(p, ',', size_t(e - p));
                if (!comma) throw std::runtime_error("anonymity set: expected address,eth_balance");
                const char* c2 = (const char*)std::memchr(comma + 1, ',', size_t(e - comma - 1));
                if (comma - p < 3 || p[0] != '0' || (p[1] != 'x' && p[1] != 'X')) throw std::runtime_error("anonymity set: address must start with 0x");
                U256 a = parse_hex(p + 2, comma), b = parse_dec(comma + 1, c2 ? c2 : e);
                size_t at = rows.size();
                rows.resize(at + 64);
                u256_to_bytes(a, &rows[at]);
                u256_to_bytes(b, &rows[at + 32]);
            }
        }
        p = nl ? nl + 1 : end;
    }
    return rows;
}

std::string big(const U256& v, int depth) {   // {"__bigint__": "..."} in serde_json pretty form at this depth
    std::string pad(size_t(2 * depth), ' ');
    return "{\n" + pad + "  \"__bigint__\": \"" + u256_to_dec(v) + "\"\n" + pad + "}";
}

}  // namespace

extern "C" int g16_merkle_files(const char* anon_set_csv, const char* poa_input_json, const char* output_dir,
                                const g16_opts* opts) {
    if (!anon_set_csv || !output_dir) return g16::fail(G16_ERR_GENERIC, "null argument");
    g16_merkle* tree = nullptr;
    try {
        std::printf("Initiating Merkle Tree build..\n");
        std::vector<uint8_t> rows = parse_anon_set(read_file(anon_set_csv));
        size_t n = rows.size() / 64;
        std::printf("Done creating %zu leaves\n", n);
        int rc = g16_merkle_build(rows.data(), n, opts, &tree);
        if (rc != G16_OK) return rc;
        uint64_t n_leaves = 0;
        uint32_t levels = 0;
        double ms = 0;
        g16_merkle_info(tree, nullptr, &n_leaves, &levels, &ms);
        std::printf("Number of leaves (after adding padding nodes): %llu\n", (unsigned long long)n_leaves);
        std::printf("Done creating Merkle tree of height %u (%.1f ms on the GPU)\n", levels, ms);
        uint8_t root[32];
        rc = g16_merkle_root(tree, root);
        if (rc != G16_OK) { g16_merkle_close(tree); return rc; }
        std::string dir = output_dir;
        std::string root_path = dir + "/merkle_root.json";
        write_file_atomic(root_path, big(u256_from_bytes(root), 0));
        std::printf("Root hash %s written to file \"%s\"\n", u256_to_dec(u256_from_bytes(root)).c_str(), root_path.c_str());

        // proofs for the owned accounts, in anonymity-set order (the reference scans forward only, merkle_tree.rs:315-335)
        std::string leaves_js, elems_js, idx_js;
        size_t owned = 0;
        if (poa_input_json) {
            std::string text = read_file(poa_input_json);
            JV doc = JParser(text.data(), text.size()).parse();
            const JV& atts = doc.at("accountAttestations");
            if (atts.kind != JV::LIST) throw std::runtime_error("PoA input: accountAttestations is not a list");
            uint64_t at = 0;
            std::vector<uint8_t> el(32 * (levels ? levels - 1 : 0)), ix(levels ? levels - 1 : 0);
            for (const JV& acc : atts.list) {
                const JV& data = acc.at("accountData");
                U256 a = g16v::dec_u256(data.at("address").at("__bigint__")), b = g16v::dec_u256(data.at("balance").at("__bigint__"));
                uint8_t pair[64], h[32];
                u256_to_bytes(a, pair);
                u256_to_bytes(b, pair + 32);
                rc = g16_poseidon_hash2(pair, 1, h, opts);
                if (rc != G16_OK) { g16_merkle_close(tree); return rc; }
                uint64_t found = 0;
                rc = g16_merkle_find(tree, h, at, &found);
                if (rc != G16_OK) { g16_merkle_close(tree); return rc; }
                if (found == UINT64_MAX)
                    throw std::runtime_error("Owned leaf " + u256_to_dec(a) + " at index " + std::to_string(owned) +
                                             " does not exist in the anonymity set");
                at = found;
                rc = g16_merkle_proof(tree, at, el.data(), ix.data());
                if (rc != G16_OK) { g16_merkle_close(tree); return rc; }
                if (owned) { leaves_js += ",\n"; elems_js += ",\n"; idx_js += ",\n"; }
                leaves_js += "    {\n      \"address\": " + big(a, 3) + ",\n      \"balance\": " + big(b, 3) +
                             ",\n      \"hash\": " + big(u256_from_bytes(h), 3) + "\n    }";
                std::string e = "    [", x = "    [";
                for (size_t l = 0; l < ix.size(); l++) {
                    e += std::string(l ? ",\n" : "\n") + "      " + big(u256_from_bytes(&el[32 * l]), 3);
                    x += std::string(l ? ",\n" : "\n") + "      " + std::to_string((int)ix[l]);
                }
                e += ix.empty() ? "]" : "\n    ]";
                x += ix.empty() ? "]" : "\n    ]";
                elems_js += e;
                idx_js += x;
                owned++;
            }
        }
        auto list = [&](const std::string& body) { return owned ? "[\n" + body + "\n  ]" : std::string("[]"); };
        std::string proofs = "{\n  \"leaves\": " + list(leaves_js) + ",\n  \"path_elements\": " + list(elems_js) +
                             ",\n  \"path_indices\": " + list(idx_js) + "\n}";
        write_file_atomic(dir + "/merkle_proofs.json", proofs);
        g16_merkle_close(tree);
        return G16_OK;
    } catch (const std::exception& e) {
        if (tree) g16_merkle_close(tree);
        return g16::fail(G16_ERR_FORMAT, std::string("merkle-tree: ") + e.what());
    }
}
