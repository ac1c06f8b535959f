// request handlers of the resident-key service (service_server.hpp)
#pragma once
#include "service_server.hpp"

namespace g16svc {

inline std::string err_reply(int code, const std::string& msg) {
    std::string m = msg;
    for (auto& ch : m)
        if (ch == '\n') ch = ' ';
    return "ERR " + std::to_string(code) + " " + m + "\n";
}

inline std::string Server::prove(const std::vector<std::string>& f) {
    // f: PROVE zkey wtns proof public style r s
    if (f.size() != 8) return err_reply(G16_ERR_GENERIC, "PROVE takes 7 fields");
    for (int i = 1; i <= 4; i++)
        if (f[i].empty() || f[i][0] != '/') return err_reply(G16_ERR_GENERIC, "paths must be absolute");
    auto t0 = std::chrono::steady_clock::now();
    std::string err;
    std::string key = key_of(f[1], err);
    if (key.empty()) return err_reply(G16_ERR_GENERIC, err);
    g16_opts o;
    g16_opts_default(&o);
    o.style = f[5] == "1" ? G16_STYLE_SNARKJS : G16_STYLE_RAPIDSNARK;
    if (f[6] != "-" || f[7] != "-") {
        if (!dec_to_le32(f[6].c_str(), o.r) || !dec_to_le32(f[7].c_str(), o.s))
            return err_reply(G16_ERR_GENERIC, "r / s must be decimal integers below 2^256");
        o.has_rs = 1;
    }
    Mapped wtns;
    if (!wtns.open(f[2], err)) return err_reply(G16_ERR_GENERIC, err);

    bool must_open = false;
    std::shared_ptr<Resident> r = acquire(key, must_open);
    double open_ms = 0;
    if (must_open) {
        g16_opts oo;
        g16_opts_default(&oo);
        oo.device = r->device;
        auto to = std::chrono::steady_clock::now();
        int rc = g16_zkey_open(f[1].c_str(), &oo, &r->zk);
        open_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - to).count();
        if (rc != G16_OK) {
            std::string m = g16_last_error();
            r->zk = nullptr;
            release(r, true);
            n_errors_++;
            return err_reply(rc, m);
        }
        n_opens_++;
    }
    uint32_t nv = 0, npub = 0, dom = 0;
    uint64_t nc = 0;
    g16_zkey_info(r->zk, &nv, &npub, &dom, &nc);
    // snarkjs style costs up to 82 bytes per public signal (newline, indent, quotes, 77 digits, comma)
    std::vector<char> proof(4096), pub(256 + 96 * size_t(npub));
    size_t pl = proof.size(), ul = pub.size();
    auto tp = std::chrono::steady_clock::now();
    int rc = g16_prove(r->zk, wtns.p, wtns.n, proof.data(), &pl, pub.data(), &ul, &o);
    if (rc == G16_ERR_SHORT_BUFFER) {   // the required sizes were written back: resize and finish (witness is loaded)
        proof.resize(pl + 1);
        pub.resize(ul + 1);
        pl = proof.size();
        ul = pub.size();
        rc = g16_prove_loaded(r->zk, proof.data(), &pl, pub.data(), &ul, &o);
    }
    double prove_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp).count();
    std::string msg = rc == G16_OK ? "" : g16_last_error();
    int device = r->device;
    r->proofs++;
    // a failed proof (bad witness) leaves the key resident and usable; after a CUDA-class error the handle's device
    // state is not trusted any more and the resident copy is dropped
    release(r, rc == G16_ERR_CUDA);
    if (rc != G16_OK) {
        n_errors_++;
        return err_reply(rc, msg);
    }
    if (!write_file(f[3], proof.data(), pl, err) || !write_file(f[4], pub.data(), ul, err)) {
        n_errors_++;
        return err_reply(G16_ERR_GENERIC, err);
    }
    n_proofs_++;
    double total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    char line[256];
    std::snprintf(line, sizeof line, "OK device=%d resident=%d open_ms=%.1f prove_ms=%.2f total_ms=%.2f\n", device,
                  must_open ? 0 : 1, open_ms, prove_ms, total_ms);
    return line;
}

inline std::string Server::stats() {
    std::lock_guard<std::mutex> g(mu_);
    std::ostringstream os;
    os << "OK proofs=" << n_proofs_.load() << " opens=" << n_opens_.load() << " errors=" << n_errors_.load()
       << " resident=" << table_.size();
    for (auto& r : table_)
        os << " [dev" << r->device << (r->busy ? " busy " : " idle ") << r->proofs << " " << r->key.substr(0, r->key.find('|'))
           << "]";
    os << "\n";
    return os.str();
}

inline void Server::handle(int fd) {
    std::vector<std::string> f = split_lines(recv_until(fd, true));
    std::string reply;
    if (f.empty()) reply = err_reply(G16_ERR_GENERIC, "empty request");
    else if (f[0] == "PING") reply = std::string("OK ") + g16_version() + "\n";
    else if (f[0] == "STATS") reply = stats();
    else if (f[0] == "SHUTDOWN") {
        reply = "OK bye\n";
        stop_.store(true);
        send_all(fd, reply);
        ::shutdown(lfd_, SHUT_RDWR);   // wakes accept()
        ::close(lfd_);
        return;
    } else if (f[0] == "PROVE") {
        try {
            reply = prove(f);
        } catch (const std::exception& e) {
            reply = err_reply(G16_ERR_GENERIC, e.what());
        }
    } else reply = err_reply(G16_ERR_GENERIC, "unknown request " + f[0]);
    send_all(fd, reply);
}

}  // namespace g16svc
