// Resident-key prover service, host side above the C ABI (uses only g16_b200.h).
//
// Why: the reference's workflow execs `prover` once per batch with the SAME proving key (scripts/full_workflow.sh:
// 303-316, :503, :534) and launches the batches concurrently under GNU parallel (:552).  A 1-25 GB key is read, parsed
// and uploaded on every call, and concurrent processes have no rule for sharing GPUs.  `prover --serve` keeps parsed
// keys resident on the GPUs it owns and assigns devices; the four-argument `prover` command line becomes a thin
// client when G16_SERVICE=<socket path> is set.  The argv contract and the output bytes do not change.
//
// Wire format (AF_UNIX stream, one request per connection, fields are '\n'-terminated lines, blank line ends):
//   PROVE <zkey> <wtns> <proof.json> <public.json> <style 0|1> <r decimal or -> <s decimal or ->
//   PING | STATS | SHUTDOWN
// Reply: "OK <text>\n" or "ERR <g16_status> <message>\n".  Paths must be absolute (the client resolves them).
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <list>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "g16_b200.h"

namespace g16svc {

inline bool dec_to_le32(const char* s, uint8_t out[32]) {
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

inline bool read_file(const std::string& path, std::vector<uint8_t>& out, std::string& err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path + ": " + std::strerror(errno); return false; }
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? size_t(n) : 0);
    size_t got = out.empty() ? 0 : std::fread(out.data(), 1, out.size(), f);
    std::fclose(f);
    if (got != out.size()) { err = "short read on " + path; return false; }
    return true;
}

// read-only mapping of a whole file (the witness: up to 1.7 GB, copied once, by the H2D transfer itself)
struct Mapped {
    const uint8_t* p = nullptr;
    size_t n = 0;
    ~Mapped() {
        if (p && n) ::munmap(const_cast<uint8_t*>(p), n);
    }
    bool open(const std::string& path, std::string& err) {
        int fd = ::open(path.c_str(), O_RDONLY);
        struct stat st;
        if (fd < 0 || ::fstat(fd, &st) != 0) {
            err = "cannot open " + path + ": " + std::strerror(errno);
            if (fd >= 0) ::close(fd);
            return false;
        }
        n = size_t(st.st_size);
        if (n) {
            void* m = ::mmap(nullptr, n, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
            if (m == MAP_FAILED) {
                err = "cannot map " + path + ": " + std::strerror(errno);
                ::close(fd);
                n = 0;
                return false;
            }
            p = static_cast<const uint8_t*>(m);
        }
        ::close(fd);
        return true;
    }
};

// write to a temporary beside the target, then rename: a reader never sees a half-written proof
inline bool write_file(const std::string& path, const char* data, size_t n, std::string& err) {
    static std::atomic<unsigned long> seq{0};   // handler threads of one service process share the pid
    std::string tmp = path + ".tmp." + std::to_string(::getpid()) + "." + std::to_string(seq++);
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) { err = "cannot create " + tmp + ": " + std::strerror(errno); return false; }
    bool ok = std::fwrite(data, 1, n, f) == n;
    ok = (std::fclose(f) == 0) && ok;
    if (!ok || std::rename(tmp.c_str(), path.c_str()) != 0) {
        err = "cannot write " + path + ": " + std::strerror(errno);
        std::remove(tmp.c_str());
        return false;
    }
    return true;
}

inline bool send_all(int fd, const std::string& s) {
    size_t off = 0;
    while (off < s.size()) {
        ssize_t k = ::send(fd, s.data() + off, s.size() - off, MSG_NOSIGNAL);
        if (k <= 0) return false;
        off += size_t(k);
    }
    return true;
}

// reads until a blank line (request) or EOF (reply); bounded
inline std::string recv_until(int fd, bool blank_line_ends) {
    std::string s;
    char buf[4096];
    while (s.size() < (1u << 20)) {
        ssize_t k = ::recv(fd, buf, sizeof buf, 0);
        if (k <= 0) break;
        s.append(buf, size_t(k));
        if (blank_line_ends && s.size() >= 2 && s.compare(s.size() - 2, 2, "\n\n") == 0) break;
    }
    return s;
}

inline std::vector<std::string> split_lines(const std::string& s) {
    std::vector<std::string> v;
    std::istringstream is(s);
    std::string l;
    while (std::getline(is, l)) v.push_back(l);
    while (!v.empty() && v.back().empty()) v.pop_back();
    return v;
}

inline int connect_to(const std::string& path) {
    sockaddr_un a;
    std::memset(&a, 0, sizeof a);
    a.sun_family = AF_UNIX;
    if (path.size() >= sizeof a.sun_path) return -1;
    std::strcpy(a.sun_path, path.c_str());
    int fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    if (::connect(fd, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0) {
        ::close(fd);
        return -1;
    }
    return fd;
}

// one request/reply exchange; returns false if the service cannot be reached
inline bool exchange(const std::string& sock, const std::string& request, std::string& reply) {
    int fd = connect_to(sock);
    if (fd < 0) return false;
    bool ok = send_all(fd, request);
    if (ok) reply = recv_until(fd, false);
    ::close(fd);
    return ok && !reply.empty();
}

}  // namespace g16svc
