// Groth16 verification from the JSON texts snarkjs / rapidsnark / this prover write (see verify.hpp)
#pragma once
#include <map>
#include <memory>
#include <stdexcept>

#include "verify_pairing.hpp"

namespace g16v {

// ---- a JSON reader just large enough for vkey.json / proof.json / public.json --------------------------------
struct JV {
    enum Kind { STR, NUM, LIST, OBJ, OTHER } kind = OTHER;
    std::string s;                       // STR: the text; NUM: the literal
    std::vector<JV> list;                // LIST
    std::map<std::string, JV> obj;       // OBJ
    const JV& at(const std::string& k) const {
        auto it = obj.find(k);
        if (kind != OBJ || it == obj.end()) throw std::runtime_error("JSON: missing key " + k);
        return it->second;
    }
};

class JParser {
public:
    JParser(const char* p, size_t n) : p_(p), e_(p + n) {}
    JV parse() {
        JV v = value();
        ws();
        if (p_ != e_) throw std::runtime_error("JSON: trailing characters");
        return v;
    }

private:
    void ws() {
        while (p_ < e_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) p_++;
    }
    char peek() {
        ws();
        if (p_ >= e_) throw std::runtime_error("JSON: unexpected end");
        return *p_;
    }
    std::string str() {
        std::string out;
        p_++;  // opening quote
        while (p_ < e_ && *p_ != '"') {
            if (*p_ == '\\' && p_ + 1 < e_) p_++;
            out.push_back(*p_++);
        }
        if (p_ >= e_) throw std::runtime_error("JSON: unterminated string");
        p_++;
        return out;
    }
    JV value() {
        JV v;
        char c = peek();
        if (c == '"') {
            v.kind = JV::STR;
            v.s = str();
        } else if (c == '[') {
            v.kind = JV::LIST;
            p_++;
            if (peek() == ']') { p_++; return v; }
            for (;;) {
                v.list.push_back(value());
                char d = peek();
                p_++;
                if (d == ']') break;
                if (d != ',') throw std::runtime_error("JSON: expected , or ]");
            }
        } else if (c == '{') {
            v.kind = JV::OBJ;
            p_++;
            if (peek() == '}') { p_++; return v; }
            for (;;) {
                if (peek() != '"') throw std::runtime_error("JSON: expected a key");
                std::string k = str();
                if (peek() != ':') throw std::runtime_error("JSON: expected :");
                p_++;
                v.obj[k] = value();
                char d = peek();
                p_++;
                if (d == '}') break;
                if (d != ',') throw std::runtime_error("JSON: expected , or }");
            }
        } else {
            v.kind = (c == '-' || (c >= '0' && c <= '9')) ? JV::NUM : JV::OTHER;
            while (p_ < e_ && *p_ != ',' && *p_ != ']' && *p_ != '}' && *p_ != ' ' && *p_ != '\n' && *p_ != '\r' && *p_ != '\t')
                v.s.push_back(*p_++);
            if (v.s.empty()) throw std::runtime_error("JSON: unexpected character");
        }
        return v;
    }
    const char *p_, *e_;
};

// decimal text (a JSON string or number) -> integer below 2^256
inline U256 dec_u256(const JV& v) {
    if (v.kind != JV::STR && v.kind != JV::NUM) throw std::runtime_error("expected a decimal number");
    U256 r{{0, 0, 0, 0}};
    if (v.s.empty()) throw std::runtime_error("empty number");
    for (char c : v.s) {
        if (c < '0' || c > '9') throw std::runtime_error("not a decimal number: " + v.s);
        unsigned __int128 carry = unsigned(c - '0');
        for (int i = 0; i < 4; i++) {
            unsigned __int128 t = (unsigned __int128)r.v[i] * 10 + carry;
            r.v[i] = (uint64_t)t;
            carry = t >> 64;
        }
        if (carry) throw std::runtime_error("number exceeds 256 bits");
    }
    return r;
}

inline fq fq_of(const JV& v) {
    U256 a = dec_u256(v);
    while (u256_cmp(a, HFq::mod()) >= 0) a = HFq::raw_sub(a, HFq::mod());
    return HFq::to_mont(a);
}

struct G1J {
    HAffine<HFq> p;
    bool inf = false;
};
struct G2J {
    HAffine<HFq2> p;
    bool inf = false;
};

// ["x", "y", "1"]; z = "0" is the point at infinity (snarkjs G1.toObject)
inline G1J g1_of(const JV& v) {
    if (v.kind != JV::LIST || v.list.size() != 3) throw std::runtime_error("G1 point: expected [x, y, z]");
    G1J r;
    U256 z = dec_u256(v.list[2]);
    r.p.x = fq_of(v.list[0]);
    r.p.y = fq_of(v.list[1]);
    if (u256_is_zero(z)) r.inf = true;
    else if (!(z.v[0] == 1 && !z.v[1] && !z.v[2] && !z.v[3])) throw std::runtime_error("G1 point: z must be 1 or 0");
    r.p.inf = r.inf;
    return r;
}

// [["x.c0", "x.c1"], ["y.c0", "y.c1"], ["1", "0"]]
inline G2J g2_of(const JV& v) {
    if (v.kind != JV::LIST || v.list.size() != 3) throw std::runtime_error("G2 point: expected [x, y, z]");
    for (const JV& c : v.list)
        if (c.kind != JV::LIST || c.list.size() != 2) throw std::runtime_error("G2 point: coordinates are pairs");
    G2J r;
    r.p.x = HFq2El{fq_of(v.list[0].list[0]), fq_of(v.list[0].list[1])};
    r.p.y = HFq2El{fq_of(v.list[1].list[0]), fq_of(v.list[1].list[1])};
    U256 z0 = dec_u256(v.list[2].list[0]), z1 = dec_u256(v.list[2].list[1]);
    if (u256_is_zero(z0) && u256_is_zero(z1)) r.inf = true;
    else if (!(z0.v[0] == 1 && !z0.v[1] && !z0.v[2] && !z0.v[3] && u256_is_zero(z1)))
        throw std::runtime_error("G2 point: z must be [1, 0] or [0, 0]");
    r.p.inf = r.inf;
    return r;
}

inline bool on_curve_g1(const G1J& a) {  // y^2 = x^3 + 3
    if (a.inf) return true;
    fq rhs = HFq::add(HFq::mul(HFq::sqr(a.p.x), a.p.x), fq_small(3));
    return HFq::eq(HFq::sqr(a.p.y), rhs);
}

inline bool on_curve_g2(const G2J& a) {  // y^2 = x^3 + 3 / (9 + u)
    if (a.inf) return true;
    static const HFq2El b2 = HFq2::mul(HFq2El{fq_small(3), HFq::zero()}, HFq2::inv(HFq2El{fq_small(9), HFq::one()}));
    HFq2El rhs = HFq2::add(HFq2::mul(HFq2::sqr(a.p.x), a.p.x), b2);
    return HFq2::eq(HFq2::sqr(a.p.y), rhs);
}

// true iff the proof verifies; throws std::runtime_error on malformed input
inline bool verify_groth16(const char* vkey, size_t vkey_len, const char* pub, size_t pub_len, const char* proof,
                           size_t proof_len) {
    JV vk = JParser(vkey, vkey_len).parse(), pu = JParser(pub, pub_len).parse(), pr = JParser(proof, proof_len).parse();
    if (pu.kind != JV::LIST) throw std::runtime_error("public.json: expected a list");
    G1J A = g1_of(pr.at("pi_a")), C = g1_of(pr.at("pi_c")), alpha = g1_of(vk.at("vk_alpha_1"));
    G2J B = g2_of(pr.at("pi_b")), beta = g2_of(vk.at("vk_beta_2")), gamma = g2_of(vk.at("vk_gamma_2")),
        delta = g2_of(vk.at("vk_delta_2"));
    const JV& icj = vk.at("IC");
    if (icj.kind != JV::LIST || icj.list.size() != pu.list.size() + 1) return false;
    std::vector<G1J> ic;
    for (const JV& p : icj.list) ic.push_back(g1_of(p));
    if (!on_curve_g1(A) || !on_curve_g1(C) || !on_curve_g1(alpha)) return false;
    for (const G1J& p : ic)
        if (!on_curve_g1(p)) return false;
    if (!on_curve_g2(B) || !on_curve_g2(beta) || !on_curve_g2(gamma) || !on_curve_g2(delta)) return false;
    // vk_x = IC0 + sum pub_i IC_{i+1}
    auto lift = [](const G1J& a) { return HCurve<HFq>::from_affine(a.p); };
    HPoint<HFq> acc = lift(ic[0]);
    for (size_t i = 0; i < pu.list.size(); i++) {
        U256 s = dec_u256(pu.list[i]);
        if (u256_cmp(s, HFr::mod()) >= 0) return false;
        acc = HCurve<HFq>::add(acc, HCurve<HFq>::mul(lift(ic[i + 1]), s));
    }
    G1J vkx;
    vkx.p = HCurve<HFq>::to_affine(acc);
    vkx.inf = vkx.p.inf;
    G1J nA = A;
    nA.p.y = HFq::neg(A.p.y);
    // e(-A, B) e(alpha, beta) e(vk_x, gamma) e(C, delta) == 1: one Miller loop over the four pairs, one final exponentiation
    auto pair = [](const G2J& q, const G1J& p) {
        PairingInput x{q.p, p.p};
        x.q.inf = q.inf;
        x.p.inf = p.inf;
        return x;
    };
    f12 f = miller_loop_multi({pair(B, nA), pair(beta, alpha), pair(gamma, vkx), pair(delta, C)});
    return f12_eq(final_exponentiate(f), f12_one());
}

}  // namespace g16v

// ---- sanitize: proof + vkey + public -> the 43-bit x 6 limb input of the recursive verifier circuit ---------------
// Replaces scripts/sanitize_groth16_proof.py:39-123 (py_ecc pairing, 3.2-3.4 s per proof in the reference's logs).
// Output text = what Python's json.dump(dict) writes there: keys gamma2, delta2, negalfa1xbeta2, IC, negpa, pb, pc,
// pubInput, separators ", " and ": ", limbs and public inputs as JSON integers.
namespace g16v {

inline std::string limbs_json(const U256& x, int n = 43, int k = 6) {   // scripts/lib/field_helper.py numberToArray
    std::string out = "[";
    U256 a = x;
    for (int i = 0; i < k; i++) {
        uint64_t limb = a.v[0] & ((uint64_t(1) << n) - 1);
        // a >>= n
        for (int w = 0; w < 4; w++) a.v[w] = (a.v[w] >> n) | (w < 3 ? a.v[w + 1] << (64 - n) : 0);
        out += std::to_string(limb);
        out += i + 1 < k ? ", " : "]";
    }
    return out;
}
inline std::string fq_json(const fq& a) { return limbs_json(HFq::from_mont(a)); }
inline std::string fq2_json(const HFq2El& a) { return "[" + fq_json(a.c0) + ", " + fq_json(a.c1) + "]"; }

inline std::string sanitize_groth16(const char* vkey, size_t vkey_len, const char* pub, size_t pub_len, const char* proof,
                                    size_t proof_len) {
    JV vk = JParser(vkey, vkey_len).parse(), pu = JParser(pub, pub_len).parse(), pr = JParser(proof, proof_len).parse();
    if (pu.kind != JV::LIST) throw std::runtime_error("public.json: expected a list");
    G1J A = g1_of(pr.at("pi_a")), C = g1_of(pr.at("pi_c")), alpha = g1_of(vk.at("vk_alpha_1"));
    G2J B = g2_of(pr.at("pi_b")), beta = g2_of(vk.at("vk_beta_2")), gamma = g2_of(vk.at("vk_gamma_2")),
        delta = g2_of(vk.at("vk_delta_2"));
    const JV& icj = vk.at("IC");
    if (icj.kind != JV::LIST) throw std::runtime_error("vkey: IC must be a list");
    if (A.inf || B.inf || C.inf || alpha.inf || beta.inf || gamma.inf || delta.inf)
        throw std::runtime_error("sanitize: a point at infinity has no affine limbs");
    // e(-alpha, beta) in the basis the circuit uses: six Fq2 coefficients of w^0..w^5 with u = w^6 - 9
    G1J na = alpha;
    na.p.y = HFq::neg(alpha.p.y);
    f12 nab = final_exponentiate(miller_loop(beta.p, false, na.p, false));
    static const fq nine = fq_small(9);
    std::string s = "{\"gamma2\": [" + fq2_json(gamma.p.x) + ", " + fq2_json(gamma.p.y) + "], \"delta2\": [" +
                    fq2_json(delta.p.x) + ", " + fq2_json(delta.p.y) + "], \"negalfa1xbeta2\": [";
    for (int i = 0; i < 6; i++) {
        HFq2El c{HFq::add(nab[i], HFq::mul(nine, nab[i + 6])), nab[i + 6]};
        s += fq2_json(c) + (i < 5 ? ", " : "]");
    }
    s += ", \"IC\": [";
    for (size_t i = 0; i < icj.list.size(); i++) {
        G1J p = g1_of(icj.list[i]);
        if (p.inf) throw std::runtime_error("sanitize: IC point at infinity");
        s += "[" + fq_json(p.p.x) + ", " + fq_json(p.p.y) + "]" + (i + 1 < icj.list.size() ? ", " : "");
    }
    s += "], \"negpa\": [" + fq_json(A.p.x) + ", " + fq_json(HFq::neg(A.p.y)) + "], \"pb\": [" + fq2_json(B.p.x) + ", " +
         fq2_json(B.p.y) + "], \"pc\": [" + fq_json(C.p.x) + ", " + fq_json(C.p.y) + "], \"pubInput\": [";
    for (size_t i = 0; i < pu.list.size(); i++) {
        s += u256_to_dec(dec_u256(pu.list[i]));   // int(text): canonical digits
        if (i + 1 < pu.list.size()) s += ", ";
    }
    s += "]}";
    return s;
}

}  // namespace g16v
