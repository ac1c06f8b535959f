// Phase-2 ceremony steps and verification-key export on the prover's own library -- SURVEY.md 8(f) rank 2, file level.
//
// Replaces what scripts/g16_setup.sh:254-295 of the reference runs through snarkjs 0.7.2:
//     snarkjs zkey contribute <in.zkey> <out.zkey> --name=.. -e=..                      -> g16_zkey_contribute
//     snarkjs zkey beacon <in.zkey> <out.zkey> <beaconHash> <numIterationsExp> -n=..    -> g16_zkey_beacon
//     snarkjs zkey export verificationkey <zkey> <vkey.json>                            -> g16_zkey_export_vkey
// The heavy part of a contribution is G.batchApplyKey over sections 8 (C) and 9 (H) (34 h / 290 GB for layer three in
// node, tests/4_sigs_2_batches_12_height/benchmarks.txt:59): that runs on the GPU through g16_batch_apply_key
// (csrc/ops.cu apply_key_kernel).  Everything here is host bookkeeping around it: the header's delta points, the section
// 10 record (layout of snarkjs zkey_utils.js writeMPCParams) and its Blake2b transcript.
// PARITY: oracle/setup.py is the specification this file matches byte for byte (tests/test_gpu_setup.py).  The value of
// delta', the point g1_s and hashToG2 are this repository's own Blake2b derivations where snarkjs draws from its ChaCha
// generator -- unpinned in the reference (no zkey is committed); see oracle/setup.py.  The vkey export's layout is
// pinned on the reference's committed vkey files.
#include <sys/random.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "g16_b200.h"
#include "setup_hash.hpp"
#include "verify_groth16.hpp"

namespace g16 {
int fail(int code, const std::string& msg);
}

namespace {
using namespace g16h;
using namespace g16setup;
typedef std::vector<uint8_t> bytes;

const char* TAG = "g16b200/phase2/";

bytes read_all(const char* path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    f.seekg(0, std::ios::end);
    std::streamoff n = f.tellg();
    f.seekg(0);
    bytes b((size_t)n);
    f.read(reinterpret_cast<char*>(b.data()), n);
    if (!f) throw std::runtime_error(std::string("cannot read ") + path);
    return b;
}

struct Sec {
    const uint8_t* p = nullptr;
    uint64_t size = 0;
};
void parse_zkey(const bytes& z, Sec secs[11]) {
    if (z.size() < 12 || std::memcmp(z.data(), "zkey", 4) != 0) throw std::runtime_error("not a zkey file");
    uint32_t nsec;
    std::memcpy(&nsec, z.data() + 8, 4);
    size_t pos = 12;
    for (uint32_t i = 0; i < nsec; i++) {
        if (pos + 12 > z.size()) throw std::runtime_error("zkey: truncated section header");
        uint32_t typ;
        uint64_t size;
        std::memcpy(&typ, z.data() + pos, 4);
        std::memcpy(&size, z.data() + pos + 4, 8);
        pos += 12;
        if (size > z.size() - pos) throw std::runtime_error("zkey: section runs past the end of the file");
        if (typ <= 10 && !secs[typ].p) secs[typ] = Sec{z.data() + pos, size};
        pos += size;
    }
    for (int t = 1; t <= 10; t++)
        if (!secs[t].p) throw std::runtime_error("zkey: missing section " + std::to_string(t));
    if (secs[2].size < 84 + 64 * 3 + 128 * 3) throw std::runtime_error("zkey: header section too short");
    if (secs[10].size < 68) throw std::runtime_error("zkey: contributions section too short");
}

// zkey point records (little-endian Montgomery, all-zero = infinity) <-> host points
HAffine<HFq> g1_rec(const uint8_t* p) {
    HAffine<HFq> a;
    a.x = u256_from_bytes(p);
    a.y = u256_from_bytes(p + 32);
    a.inf = u256_is_zero(a.x) && u256_is_zero(a.y);
    return a;
}
HAffine<HFq2> g2_rec(const uint8_t* p) {
    HAffine<HFq2> a;
    a.x = HFq2El{u256_from_bytes(p), u256_from_bytes(p + 32)};
    a.y = HFq2El{u256_from_bytes(p + 64), u256_from_bytes(p + 96)};
    a.inf = HFq2::is_zero(a.x) && HFq2::is_zero(a.y);
    return a;
}
void put_g1(uint8_t* p, const HAffine<HFq>& a) {
    std::memset(p, 0, 64);
    if (a.inf) return;
    u256_to_bytes(a.x, p);
    u256_to_bytes(a.y, p + 32);
}
void put_g2(uint8_t* p, const HAffine<HFq2>& a) {
    std::memset(p, 0, 128);
    if (a.inf) return;
    u256_to_bytes(a.x.c0, p);
    u256_to_bytes(a.x.c1, p + 32);
    u256_to_bytes(a.y.c0, p + 64);
    u256_to_bytes(a.y.c1, p + 96);
}
void be32(Blake2b& h, const U256& mont) {   // big-endian standard form
    U256 v = HFq::from_mont(mont);
    uint8_t b[32];
    for (int i = 0; i < 32; i++) b[i] = uint8_t(v.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
    h.update(b, 32);
}
void hash_g1(Blake2b& h, const HAffine<HFq>& a) {
    be32(h, a.x);
    be32(h, a.y);
}
void hash_g2(Blake2b& h, const HAffine<HFq2>& a) {   // Fq2 as c1 | c0
    be32(h, a.x.c1);
    be32(h, a.x.c0);
    be32(h, a.y.c1);
    be32(h, a.y.c0);
}
// one stored contribution: deltaAfter | g1_s | g1_sx (G1) | g2_spx (G2) | transcript[64]
void hash_pubkey(Blake2b& h, const uint8_t* rec) {
    hash_g1(h, g1_rec(rec));
    hash_g1(h, g1_rec(rec + 64));
    hash_g1(h, g1_rec(rec + 128));
    hash_g2(h, g2_rec(rec + 192));
    h.update(rec + 320, 64);
}

// little-endian bytes (any length) mod m, by Horner over the bytes from the top
U256 le_mod(const bytes& b, const U256& m, bool is_q) {
    U256 acc{{0, 0, 0, 0}};
    // shift-and-subtract, one bit at a time: acc < m < 2^254, so 2 acc + bit never overflows 256 bits
    for (size_t i = b.size(); i-- > 0;) {
        for (int bit = 7; bit >= 0; bit--) {
            // acc = 2 acc + bit
            uint64_t carry = (b[i] >> bit) & 1;
            for (int k = 0; k < 4; k++) {
                uint64_t nc = acc.v[k] >> 63;
                acc.v[k] = (acc.v[k] << 1) | carry;
                carry = nc;
            }
            if (u256_cmp(acc, m) >= 0) acc = is_q ? HFq::raw_sub(acc, m) : HFr::raw_sub(acc, m);
        }
    }
    return acc;
}

bytes blake(std::initializer_list<bytes> parts) {
    Blake2b h(64);
    for (const bytes& p : parts) h.update(p);
    return h.digest();
}
bytes tagged(const char* what) {
    std::string s = std::string(TAG) + what;
    return bytes(s.begin(), s.end());
}

// Fq2 square root by the complex method (q = 3 mod 4); false = not a square.  Same steps as oracle/setup.py f2_sqrt.
bool fq_is_square(const U256& a) {   // Montgomery in
    U256 e = HFq::mod();
    e.v[0] -= 1;
    for (int i = 0; i < 4; i++) e.v[i] = (e.v[i] >> 1) | (i < 3 ? e.v[i + 1] << 63 : 0);
    return HFq::eq(HFq::pow(a, e), HFq::one());
}
U256 fq_root(const U256& a) {   // a^((q+1)/4)
    U256 e = HFq::mod();
    e.v[0] += 1;   // q = 3 mod 4: no carry out of the low limb's low bits matters (q + 1 < 2^254)
    for (int r = 0; r < 2; r++)
        for (int i = 0; i < 4; i++) e.v[i] = (e.v[i] >> 1) | (i < 3 ? e.v[i + 1] << 63 : 0);
    return HFq::pow(a, e);
}
bool fq2_sqrt(const HFq2El& a, HFq2El& out) {
    if (u256_is_zero(a.c1)) {
        if (u256_is_zero(a.c0)) { out = HFq2::zero(); return true; }
        if (fq_is_square(a.c0)) out = HFq2El{fq_root(a.c0), HFq::zero()};
        else out = HFq2El{HFq::zero(), fq_root(HFq::neg(a.c0))};
        return true;
    }
    U256 norm = HFq::add(HFq::sqr(a.c0), HFq::sqr(a.c1));
    if (!fq_is_square(norm)) return false;
    U256 s = fq_root(norm);
    U256 half = HFq::inv(HFq::from_u64(2));
    U256 d = HFq::mul(HFq::add(a.c0, s), half);
    if (!fq_is_square(d)) {
        d = HFq::mul(HFq::sub(a.c0, s), half);
        if (!fq_is_square(d)) return false;
    }
    U256 x0 = fq_root(d);
    U256 x1 = HFq::mul(a.c1, HFq::inv(HFq::add(x0, x0)));
    out = HFq2El{x0, x1};
    return true;
}
bool fq_gt_half(const U256& mont) {   // standard-form value > (q-1)/2
    U256 v = HFq::from_mont(mont);
    U256 h = HFq::mod();
    h.v[0] -= 1;
    for (int i = 0; i < 4; i++) h.v[i] = (h.v[i] >> 1) | (i < 3 ? h.v[i + 1] << 63 : 0);
    return u256_cmp(v, h) > 0;
}

HAffine<HFq2> hash_to_g2(const bytes& transcript) {
    static const HFq2El b2 = HFq2::mul(HFq2El{HFq::from_u64(3), HFq::zero()}, HFq2::inv(HFq2El{HFq::from_u64(9), HFq::one()}));
    // cofactor of G2 in E'(Fq2): 2q - r
    U256 cof = HFq::mod();
    {
        unsigned __int128 c = 0;
        U256 two_q;
        for (int i = 0; i < 4; i++) {
            c += (unsigned __int128)cof.v[i] * 2;
            two_q.v[i] = (uint64_t)c;
            c >>= 64;
        }
        cof = HFr::raw_sub(two_q, HFr::mod());
    }
    for (uint32_t ctr = 0;; ctr++) {
        bytes c(4);
        std::memcpy(c.data(), &ctr, 4);
        U256 x0 = le_mod(blake({tagged("g2/c0"), transcript, c}), HFq::mod(), true);
        U256 x1 = le_mod(blake({tagged("g2/c1"), transcript, c}), HFq::mod(), true);
        HFq2El x{HFq::to_mont(x0), HFq::to_mont(x1)};
        HFq2El rhs = HFq2::add(HFq2::mul(HFq2::sqr(x), x), b2), y;
        if (!fq2_sqrt(rhs, y)) continue;
        bool negative = !u256_is_zero(y.c1) ? fq_gt_half(y.c1) : fq_gt_half(y.c0);
        if (negative) y = HFq2::neg(y);
        HAffine<HFq2> p{x, y, false};
        HAffine<HFq2> out = HG2::to_affine(HG2::mul(HG2::from_affine(p), cof));
        if (!out.inf) return out;
    }
}

void write_atomic(const char* path, const bytes& data) {
    std::string tmp = std::string(path) + ".tmp." + std::to_string((long)getpid());
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) throw std::runtime_error(std::string("cannot write ") + path);
    size_t w = std::fwrite(data.data(), 1, data.size(), f);
    if (std::fclose(f) != 0 || w != data.size() || std::rename(tmp.c_str(), path) != 0) {
        std::remove(tmp.c_str());
        throw std::runtime_error(std::string("cannot write ") + path);
    }
}

// the contribution itself; seed = the bytes every secret is derived from
int contribute(const char* in_path, const char* out_path, const char* name, const bytes& seed, const uint8_t* beacon_hash,
               size_t beacon_len, uint32_t beacon_exp, const g16_opts* opts, uint8_t contribution_hash[64]) {
    bytes z = read_all(in_path);
    Sec secs[11];
    parse_zkey(z, secs);
    const size_t o = 4 + 32 + 4 + 32 + 12;   // first point of the header section
    HAffine<HFq> delta1 = g1_rec(secs[2].p + o + 384);
    HAffine<HFq2> delta2 = g2_rec(secs[2].p + o + 448);
    // previous contributions
    const uint8_t* mp = secs[10].p;
    uint32_t ncontrib;
    std::memcpy(&ncontrib, mp + 64, 4);
    Blake2b th(64);
    th.update(mp, 64);   // csHash
    size_t pos = 68;
    for (uint32_t i = 0; i < ncontrib; i++) {
        if (pos + 388 > secs[10].size) throw std::runtime_error("zkey: contributions section is truncated");
        hash_pubkey(th, mp + pos);
        uint32_t plen;
        std::memcpy(&plen, mp + pos + 384, 4);
        pos += 388 + plen;
    }
    if (pos != secs[10].size) throw std::runtime_error("zkey: contributions section has trailing bytes");
    // secrets
    U256 d = le_mod(blake({tagged("delta"), seed}), HFr::mod(), false);
    U256 s = le_mod(blake({tagged("s"), seed}), HFr::mod(), false);
    if (u256_is_zero(d)) d = U256{{1, 0, 0, 0}};
    if (u256_is_zero(s)) s = U256{{1, 0, 0, 0}};
    HAffine<HFq> gen1{HFq::one(), HFq::add(HFq::one(), HFq::one()), false};
    HAffine<HFq> g1_s = HG1::to_affine(HG1::mul(HG1::from_affine(gen1), s));
    HAffine<HFq> g1_sx = HG1::to_affine(HG1::mul(HG1::from_affine(g1_s), d));
    hash_g1(th, g1_s);
    hash_g1(th, g1_sx);
    bytes transcript = th.digest();
    HAffine<HFq2> g2_sp = hash_to_g2(transcript);
    HAffine<HFq2> g2_spx = HG2::to_affine(HG2::mul(HG2::from_affine(g2_sp), d));
    HAffine<HFq> nd1 = HG1::to_affine(HG1::mul(HG1::from_affine(delta1), d));
    HAffine<HFq2> nd2 = HG2::to_affine(HG2::mul(HG2::from_affine(delta2), d));
    // the new contribution record
    bytes rec(384);
    put_g1(rec.data(), nd1);
    put_g1(rec.data() + 64, g1_s);
    put_g1(rec.data() + 128, g1_sx);
    put_g2(rec.data() + 192, g2_spx);
    std::memcpy(rec.data() + 320, transcript.data(), 64);
    bytes params;
    if (name && *name) {
        size_t nl = std::strlen(name);
        if (nl > 64) nl = 64;
        params.push_back(1);
        params.push_back((uint8_t)nl);
        params.insert(params.end(), name, name + nl);
    }
    if (beacon_hash) {
        params.push_back(2);
        params.push_back((uint8_t)beacon_exp);
        params.push_back(3);
        params.push_back((uint8_t)beacon_len);
        params.insert(params.end(), beacon_hash, beacon_hash + beacon_len);
    }
    // output file: sections 1..10 in order
    U256 dinv_m = HFr::inv(HFr::to_mont(d));
    U256 dinv = HFr::from_mont(dinv_m);
    uint8_t first[32], one[32] = {1};
    u256_to_bytes(dinv, first);
    bytes out;
    uint64_t total = 12;
    for (int t = 1; t <= 10; t++) total += 12 + secs[t].size;
    total += 388 + params.size();
    out.reserve(total);
    auto put = [&](const void* p, size_t n) { out.insert(out.end(), (const uint8_t*)p, (const uint8_t*)p + n); };
    uint32_t v1 = 1, v10 = 10;
    put("zkey", 4);
    put(&v1, 4);
    put(&v10, 4);
    for (uint32_t t = 1; t <= 10; t++) {
        uint64_t size = secs[t].size + (t == 10 ? 388 + params.size() : 0);
        put(&t, 4);
        put(&size, 8);
        size_t at = out.size();
        put(secs[t].p, secs[t].size);
        if (t == 2) {
            put_g1(out.data() + at + o + 384, nd1);
            put_g2(out.data() + at + o + 448, nd2);
        } else if (t == 8 || t == 9) {
            // G.batchApplyKey(section, 1/delta', 1) on the GPU
            size_t n = secs[t].size / 64;
            int rc = g16_batch_apply_key(0, secs[t].p, n, first, one, out.data() + at, opts);
            if (rc != G16_OK) return rc;
        } else if (t == 10) {
            uint32_t n1 = ncontrib + 1;
            std::memcpy(out.data() + at + 64, &n1, 4);
            put(rec.data(), rec.size());
            uint32_t plen = (uint32_t)params.size();
            put(&plen, 4);
            put(params.data(), params.size());
        }
    }
    write_atomic(out_path, out);
    if (contribution_hash) {
        Blake2b ch(64);
        hash_pubkey(ch, rec.data());
        bytes dg = ch.digest();
        std::memcpy(contribution_hash, dg.data(), 64);
    }
    return G16_OK;
}

std::string dec(const U256& mont) { return "\"" + u256_to_dec(HFq::from_mont(mont)) + "\""; }

// ---- zkey verify -----------------------------------------------------------------------------------------------------
// e(a1, b2) == e(a2, b1): (a1, a2) in G1 and (b1, b2) in G2 differ by the same factor.  One two-pair Miller loop.
bool same_ratio(const HAffine<HFq>& a1, const HAffine<HFq>& a2, const HAffine<HFq2>& b1, const HAffine<HFq2>& b2) {
    HAffine<HFq> na2 = a2;
    if (!na2.inf) na2.y = HFq::neg(na2.y);
    g16v::f12 f = g16v::miller_loop_multi({g16v::PairingInput{b2, a1}, g16v::PairingInput{b1, na2}});
    return g16v::f12_eq(g16v::final_exponentiate(f), g16v::f12_one());
}

bool g1_eq(const HAffine<HFq>& a, const HAffine<HFq>& b) {
    if (a.inf || b.inf) return a.inf == b.inf;
    return u256_eq(a.x, b.x) && u256_eq(a.y, b.y);
}

// the checks of snarkjs zkey_verify_frominit.js on this library's transcript conventions; "" = valid, else what failed
std::string verify_from_init(const bytes& init, const bytes& z, const g16_opts* opts) {
    Sec si[11], sz[11];
    parse_zkey(init, si);
    parse_zkey(z, sz);
    const size_t o = 4 + 32 + 4 + 32 + 12;
    auto same = [](const Sec& a, const Sec& b) { return a.size == b.size && std::memcmp(a.p, b.p, a.size) == 0; };
    if (!same(sz[1], si[1])) return "not a groth16 key";
    if (sz[2].size != si[2].size || std::memcmp(sz[2].p, si[2].p, o + 384) != 0)
        return "header differs from the initial key (sizes, alpha, beta or gamma)";
    const char* names[8] = {"", "", "", "IC", "coefficients", "A", "B1", "B2"};
    for (int t = 3; t <= 7; t++)
        if (!same(sz[t], si[t])) return "section " + std::to_string(t) + " (" + names[t] + ") differs from the initial key";
    if (std::memcmp(sz[10].p, si[10].p, 64) != 0) return "circuit hash differs from the initial key";
    HAffine<HFq> cur = g1_rec(si[2].p + o + 384);
    const HAffine<HFq> gen1{HFq::one(), HFq::add(HFq::one(), HFq::one()), false};
    const uint8_t* mp = sz[10].p;
    uint32_t ncontrib;
    std::memcpy(&ncontrib, mp + 64, 4);
    Blake2b th(64);
    th.update(mp, 64);
    size_t pos = 68;
    for (uint32_t i = 0; i < ncontrib; i++) {
        const std::string who = "contribution " + std::to_string(i + 1) + ": ";
        if (pos + 388 > sz[10].size) return "contributions section is truncated";
        const uint8_t* rec = mp + pos;
        uint32_t plen;
        std::memcpy(&plen, rec + 384, 4);
        if (plen > sz[10].size - pos - 388) return "contributions section is truncated";
        HAffine<HFq> after = g1_rec(rec), g1_s = g1_rec(rec + 64), g1_sx = g1_rec(rec + 128);
        HAffine<HFq2> g2_spx = g2_rec(rec + 192);
        if (after.inf || g1_s.inf || g1_sx.inf || g2_spx.inf) return who + "point at infinity";
        Blake2b t2 = th;
        hash_g1(t2, g1_s);
        hash_g1(t2, g1_sx);
        bytes transcript = t2.digest();
        if (std::memcmp(transcript.data(), rec + 320, 64) != 0) return who + "transcript mismatch";
        HAffine<HFq2> g2_sp = hash_to_g2(transcript);
        if (!same_ratio(g1_s, g1_sx, g2_sp, g2_spx)) return who + "invalid proof of knowledge";
        if (!same_ratio(cur, after, g2_sp, g2_spx)) return who + "delta does not follow from the previous one";
        // parameters: 1 len name | 2 exp | 3 len beaconHash
        const uint8_t* pp = rec + 388;
        const uint8_t* bh = nullptr;
        size_t bh_len = 0;
        int exp = -1;
        for (size_t k = 0; k < plen;) {
            uint8_t t = pp[k];
            if (t == 1 && k + 1 < plen && k + 2 + pp[k + 1] <= plen) k += 2 + pp[k + 1];
            else if (t == 2 && k + 1 < plen) { exp = pp[k + 1]; k += 2; }
            else if (t == 3 && k + 1 < plen && k + 2 + pp[k + 1] <= plen) { bh = pp + k + 2; bh_len = pp[k + 1]; k += 2 + bh_len; }
            else return who + "malformed parameters";
        }
        if (bh && exp >= 0) {
            if (exp > 63) return who + "beacon exponent above 63";
            bytes seed(bh, bh + bh_len);
            for (uint64_t it = 0; it < (uint64_t(1) << exp); it++) seed = sha256(seed.data(), seed.size());
            U256 d = le_mod(blake({tagged("delta"), seed}), HFr::mod(), false);
            U256 sc = le_mod(blake({tagged("s"), seed}), HFr::mod(), false);
            if (u256_is_zero(d)) d = U256{{1, 0, 0, 0}};
            if (u256_is_zero(sc)) sc = U256{{1, 0, 0, 0}};
            HAffine<HFq> want_s = HG1::to_affine(HG1::mul(HG1::from_affine(gen1), sc));
            HAffine<HFq> want_sx = HG1::to_affine(HG1::mul(HG1::from_affine(want_s), d));
            if (!g1_eq(g1_s, want_s) || !g1_eq(g1_sx, want_sx)) return who + "not the beacon it claims to be";
        }
        cur = after;
        hash_pubkey(th, rec);
        pos += 388 + plen;
    }
    if (pos != sz[10].size) return "contributions section has trailing bytes";
    HAffine<HFq> delta1 = g1_rec(sz[2].p + o + 384);
    HAffine<HFq2> delta2 = g2_rec(sz[2].p + o + 448);
    if (!g1_eq(delta1, cur)) return "delta in the header is not the last contribution's";
    HAffine<HFq2> gen2 = g2_rec(si[2].p + o + 448);   // the initial key's delta2 is the generator
    if (!same_ratio(gen1, delta1, gen2, delta2)) return "delta1 and delta2 differ";
    // sections 8 (C) and 9 (H): a random linear combination of the initial section and the same combination of the key's
    // differ by delta (two MSMs on the GPU per section, 128-bit coefficients from the kernel's CSPRNG)
    for (int t = 8; t <= 9; t++) {
        const std::string what = "section " + std::to_string(t) + (t == 8 ? " (C)" : " (H)");
        if (sz[t].size != si[t].size) return what + " has the wrong size";
        size_t n = sz[t].size / 64;
        if (!n) continue;
        bytes r(n * 32, 0);
        for (size_t k = 0; k < n; k++) {
            size_t got = 0;
            while (got < 16) {
                ssize_t w = getrandom(&r[k * 32 + got], 16 - got, 0);
                if (w < 0 && errno != EINTR) throw std::runtime_error("getrandom failed");
                if (w > 0) got += size_t(w);
            }
            r[k * 32] |= 1;
        }
        uint8_t a1[64], a2[64];
        int rc = g16_msm_g1_host(si[t].p, r.data(), n, a1, opts);
        if (rc == G16_OK) rc = g16_msm_g1_host(sz[t].p, r.data(), n, a2, opts);
        if (rc != G16_OK) throw std::runtime_error(g16_last_error());
        HAffine<HFq> p1 = g1_rec(a1), p2 = g1_rec(a2);
        if (p1.inf != p2.inf || (!p1.inf && !same_ratio(p1, p2, delta2, gen2))) return what + " is not the initial section divided by delta";
    }
    return "";
}

}  // namespace

extern "C" {

int g16_zkey_contribute(const char* zkey_in, const char* zkey_out, const char* name, const uint8_t* entropy, size_t entropy_len,
                        const g16_opts* opts, uint8_t contribution_hash[64]) {
    if (!zkey_in || !zkey_out) return g16::fail(G16_ERR_GENERIC, "null argument");
    try {
        bytes seed;
        if (entropy && entropy_len) seed.assign(entropy, entropy + entropy_len);
        else {   // no entropy given: 64 bytes from the kernel's CSPRNG
            seed.resize(64);
            size_t got = 0;
            while (got < 64) {
                ssize_t k = getrandom(seed.data() + got, 64 - got, 0);
                if (k < 0 && errno != EINTR) throw std::runtime_error("getrandom failed");
                if (k > 0) got += size_t(k);
            }
        }
        return contribute(zkey_in, zkey_out, name, seed, nullptr, 0, 0, opts, contribution_hash);
    } catch (const std::exception& e) {
        return g16::fail(G16_ERR_FORMAT, std::string("zkey contribute: ") + e.what());
    }
}

int g16_zkey_beacon(const char* zkey_in, const char* zkey_out, const char* name, const uint8_t* beacon_hash, size_t hash_len,
                    uint32_t num_iterations_exp, const g16_opts* opts, uint8_t contribution_hash[64]) {
    if (!zkey_in || !zkey_out || !beacon_hash || !hash_len) return g16::fail(G16_ERR_GENERIC, "null argument");
    if (hash_len > 255 || num_iterations_exp > 63) return g16::fail(G16_ERR_GENERIC, "zkey beacon: hash longer than 255 bytes or exponent above 63");
    try {
        bytes cur(beacon_hash, beacon_hash + hash_len);
        for (uint64_t i = 0; i < (uint64_t(1) << num_iterations_exp); i++) cur = sha256(cur.data(), cur.size());
        return contribute(zkey_in, zkey_out, name, cur, beacon_hash, hash_len, num_iterations_exp, opts, contribution_hash);
    } catch (const std::exception& e) {
        return g16::fail(G16_ERR_FORMAT, std::string("zkey beacon: ") + e.what());
    }
}

// `snarkjs zkey verify <circuit.r1cs> <pot.ptau> <circuit_final.zkey>` (scripts/g16_verify.sh:190-202): the initial key is made
// again from the circuit and the ceremony file (g16_zkey_new, on the GPU, into a temporary file), then the key under test must
// be that key plus a valid chain of contributions.
int g16_zkey_verify(const char* r1cs_path, const char* ptau_path, const char* zkey_path, const g16_opts* opts) {
    if (!r1cs_path || !ptau_path || !zkey_path) return g16::fail(G16_ERR_GENERIC, "null argument");
    const char* td = std::getenv("TMPDIR");
    std::string tmp = std::string(td && *td ? td : "/tmp") + "/g16_zkey_verify." + std::to_string((long)getpid()) + ".zkey";
    try {
        int rc = g16_zkey_new(r1cs_path, ptau_path, tmp.c_str(), opts, nullptr);
        if (rc != G16_OK) return rc;
        bytes init = read_all(tmp.c_str());
        std::remove(tmp.c_str());
        bytes z = read_all(zkey_path);
        std::string why = verify_from_init(init, z, opts);
        if (!why.empty()) return g16::fail(G16_ERR_GENERIC, "zkey verify: " + why);
        return G16_OK;
    } catch (const std::exception& e) {
        std::remove(tmp.c_str());
        return g16::fail(G16_ERR_FORMAT, std::string("zkey verify: ") + e.what());
    }
}

// verification_key.json as snarkjs writes it: JSON.stringify(vKey, null, 1), keys protocol, curve, nPublic, vk_alpha_1,
// vk_beta_2, vk_gamma_2, vk_delta_2, vk_alphabeta_12, IC (tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json)
int g16_zkey_export_vkey(const char* zkey_path, char* out, size_t* out_len) {
    if (!zkey_path || !out_len) return g16::fail(G16_ERR_GENERIC, "null argument");
    try {
        bytes z = read_all(zkey_path);
        Sec secs[11];
        parse_zkey(z, secs);
        uint32_t npub;
        std::memcpy(&npub, secs[2].p + 76, 4);
        const size_t o = 84;
        if (secs[3].size != uint64_t(npub + 1) * 64) throw std::runtime_error("zkey: IC section has the wrong size");
        HAffine<HFq> alpha1 = g1_rec(secs[2].p + o);
        HAffine<HFq2> beta2 = g2_rec(secs[2].p + o + 128), gamma2 = g2_rec(secs[2].p + o + 256), delta2 = g2_rec(secs[2].p + o + 448);
        auto g1 = [&](const HAffine<HFq>& a, const std::string& pad) {
            if (a.inf) return "[\n" + pad + " \"0\",\n" + pad + " \"1\",\n" + pad + " \"0\"\n" + pad + "]";
            return "[\n" + pad + " " + dec(a.x) + ",\n" + pad + " " + dec(a.y) + ",\n" + pad + " \"1\"\n" + pad + "]";
        };
        auto pair = [&](const std::string& a, const std::string& b, const std::string& pad) {
            return "[\n" + pad + " " + a + ",\n" + pad + " " + b + "\n" + pad + "]";
        };
        auto g2 = [&](const HAffine<HFq2>& a) {
            std::string pad = "  ";
            return "[\n" + pad + pair(dec(a.x.c0), dec(a.x.c1), pad) + ",\n" + pad + pair(dec(a.y.c0), dec(a.y.c1), pad) + ",\n" + pad +
                   pair("\"1\"", "\"0\"", pad) + "\n ]";
        };
        // e(alpha1, beta2) raised to snarkjs' fixed multiple 2x(6x^2+3x+1) (its final exponentiation's hard part), in the
        // tower basis Fq12 = Fq6[w]/(w^2 - v), Fq6 = Fq2[v]/(v^3 - (9+u)): coefficient (j, i) = w-power 2i + j
        g16v::f12 ab = g16v::final_exponentiate(g16v::miller_loop(beta2, beta2.inf, alpha1, alpha1.inf));
        {
            const unsigned __int128 x = 4965661367192848881ull;
            unsigned __int128 m = 2 * x;                 // 2x (6x^2 + 3x + 1) as a multi-word integer
            std::vector<uint64_t> e = {(uint64_t)m, (uint64_t)(m >> 64)};
            auto mul_small = [&](std::vector<uint64_t> a, unsigned __int128 k) {
                std::vector<uint64_t> r(a.size() + 2, 0);
                uint64_t k0 = (uint64_t)k, k1 = (uint64_t)(k >> 64);
                for (size_t i = 0; i < a.size(); i++) {
                    unsigned __int128 c = (unsigned __int128)a[i] * k0 + r[i];
                    r[i] = (uint64_t)c;
                    unsigned __int128 carry = c >> 64;
                    c = (unsigned __int128)a[i] * k1 + r[i + 1] + carry;
                    r[i + 1] = (uint64_t)c;
                    r[i + 2] += (uint64_t)(c >> 64);
                }
                while (r.size() > 1 && r.back() == 0) r.pop_back();
                return r;
            };
            unsigned __int128 inner = 6 * x * x + 3 * x + 1;   // < 2^128: 6 * (2^62.1)^2 = 2^126.8
            e = mul_small(e, inner);
            ab = g16v::f12_pow(ab, e);
        }
        static const U256 nine = HFq::from_u64(9);
        std::string s = "{\n \"protocol\": \"groth16\",\n \"curve\": \"bn128\",\n \"nPublic\": " + std::to_string(npub) + ",\n";
        s += " \"vk_alpha_1\": " + g1(alpha1, " ") + ",\n";
        s += " \"vk_beta_2\": " + g2(beta2) + ",\n \"vk_gamma_2\": " + g2(gamma2) + ",\n \"vk_delta_2\": " + g2(delta2) + ",\n";
        s += " \"vk_alphabeta_12\": [\n";
        for (int j = 0; j < 2; j++) {
            s += "  [\n";
            for (int i = 0; i < 3; i++) {
                int k = 2 * i + j;
                U256 c1 = ab[k + 6], c0 = HFq::add(ab[k], HFq::mul(nine, ab[k + 6]));
                s += "   " + pair(dec(c0), dec(c1), "   ") + (i < 2 ? ",\n" : "\n");
            }
            s += j == 0 ? "  ],\n" : "  ]\n";
        }
        s += " ],\n \"IC\": [\n";
        for (uint32_t i = 0; i <= npub; i++) s += "  " + g1(g1_rec(secs[3].p + 64 * size_t(i)), "  ") + (i < npub ? ",\n" : "\n");
        s += " ]\n}";
        bool fits = out && *out_len >= s.size();
        if (fits) {
            std::memcpy(out, s.data(), s.size());
            if (*out_len > s.size()) out[s.size()] = 0;
        }
        *out_len = s.size();
        return fits ? (int)G16_OK : g16::fail(G16_ERR_SHORT_BUFFER, "output buffer too short");
    } catch (const std::exception& e) {
        return g16::fail(G16_ERR_FORMAT, std::string("zkey export verificationkey: ") + e.what());
    }
}

// test hook: 0 = Blake2b-512, 1 = SHA-256
int g16_test_hash(int which, const uint8_t* data, size_t n, uint8_t* out) {
    if (!out || (!data && n)) return g16::fail(G16_ERR_GENERIC, "null argument");
    if (which == 0) {
        Blake2b h(64);
        h.update(data, n);
        bytes d = h.digest();
        std::memcpy(out, d.data(), 64);
    } else if (which == 1) {
        bytes d = sha256(data, n);
        std::memcpy(out, d.data(), 32);
    } else {   // 2: hashToG2 of a transcript -> the 128-byte point record
        put_g2(out, hash_to_g2(bytes(data, data + n)));
    }
    return G16_OK;
}

}  // extern "C"
