// `prover --serve`: the resident-key service itself (see service.hpp for the why and the wire format).
//
// Placement rule: a PROVE request takes an idle resident copy of its key if one exists; otherwise it opens the key
// on the device that holds the fewest keys and does not have this one yet (so the concurrent batches of
// scripts/full_workflow.sh:552 spread over the GPUs, each GPU paying the upload once); if every device already
// holds the key, it waits for a copy to become idle.  One prove in flight per resident copy (the C ABI's rule).
// Keys are identified by (real path, size, mtime); per device at most `max_keys` stay resident, least recently
// used idle copy evicted first.
#pragma once
#include "service.hpp"

namespace g16svc {

struct Resident {
    std::string key;
    int device = 0;
    g16_zkey* zk = nullptr;
    bool busy = false;       // opening or proving
    uint64_t last_used = 0;
    uint64_t proofs = 0;
};

class Server {
public:
    Server(std::string sock, std::vector<int> devices, int max_keys)
        : sock_(std::move(sock)), devices_(std::move(devices)), max_keys_(max_keys < 1 ? 1 : max_keys) {}

    int run() {
        sockaddr_un a;
        std::memset(&a, 0, sizeof a);
        a.sun_family = AF_UNIX;
        if (sock_.size() >= sizeof a.sun_path) { std::fprintf(stderr, "prover: socket path too long\n"); return 1; }
        std::strcpy(a.sun_path, sock_.c_str());
        if (connect_to(sock_) >= 0) { std::fprintf(stderr, "prover: a service already listens on %s\n", sock_.c_str()); return 1; }
        ::unlink(sock_.c_str());
        lfd_ = ::socket(AF_UNIX, SOCK_STREAM, 0);
        if (lfd_ < 0 || ::bind(lfd_, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0 || ::listen(lfd_, 64) != 0) {
            std::fprintf(stderr, "prover: cannot listen on %s: %s\n", sock_.c_str(), std::strerror(errno));
            return 1;
        }
        ::chmod(sock_.c_str(), 0600);   // requests name arbitrary output paths and can shut the service down: owner only
        std::fprintf(stderr, "prover: %s serving on %s, %zu device(s), %d key(s) per device\n", g16_version(),
                     sock_.c_str(), devices_.size(), max_keys_);
        while (!stop_.load()) {
            int fd = ::accept(lfd_, nullptr, nullptr);
            if (fd < 0) {
                if (stop_.load()) break;
                continue;
            }
            active_++;
            std::thread([this, fd] {
                handle(fd);
                ::close(fd);
                active_--;
            }).detach();
        }
        ::unlink(sock_.c_str());
        // wait for the handlers that are still proving or opening a key (a large proof can take a while); a handle
        // that is still busy after that is left to process exit rather than freed under a running proof
        for (int i = 0; i < 60000 && active_.load() > 0; i++) std::this_thread::sleep_for(std::chrono::milliseconds(10));
        std::lock_guard<std::mutex> g(mu_);
        for (auto& r : table_)
            if (r->zk && !r->busy) g16_zkey_close(r->zk);
        table_.clear();
        return 0;
    }

private:
    static std::string key_of(const std::string& path, std::string& err) {
        char real[4096];
        struct stat st;
        if (!::realpath(path.c_str(), real) || ::stat(real, &st) != 0) {
            err = "cannot open " + path + ": " + std::strerror(errno);
            return "";
        }
        return std::string(real) + "|" + std::to_string((long long)st.st_size) + "|" +
               std::to_string((long long)st.st_mtim.tv_sec) + "." + std::to_string((long long)st.st_mtim.tv_nsec);
    }

    // -> a resident copy marked busy (zk == nullptr: the caller must open it), or nullptr with err set
    std::shared_ptr<Resident> acquire(const std::string& key, bool& must_open) {
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            std::shared_ptr<Resident> idle;
            std::vector<int> holds(devices_.size(), 0), count(devices_.size(), 0);
            for (auto& r : table_) {
                size_t di = dev_index(r->device);
                count[di]++;
                if (r->key == key) {
                    holds[di] = 1;
                    if (!r->busy && r->zk && !idle) idle = r;
                }
            }
            if (idle) {
                idle->busy = true;
                idle->last_used = ++clock_;
                must_open = false;
                return idle;
            }
            int best = -1;
            for (size_t i = 0; i < devices_.size(); i++)
                if (!holds[i] && (best < 0 || count[i] < count[best])) best = int(i);
            if (best >= 0) {
                if (!evict_for(devices_[best])) {   // the device is full of busy keys: wait for one to be released
                    cv_.wait(lk);
                    continue;
                }
                auto r = std::make_shared<Resident>();
                r->key = key;
                r->device = devices_[best];
                r->busy = true;
                r->last_used = ++clock_;
                table_.push_back(r);
                must_open = true;
                return r;
            }
            cv_.wait(lk);  // every device holds this key and all copies are busy
        }
    }

    void release(const std::shared_ptr<Resident>& r, bool drop) {
        {
            std::lock_guard<std::mutex> g(mu_);
            r->busy = false;
            if (drop) {
                if (r->zk) g16_zkey_close(r->zk);
                r->zk = nullptr;
                table_.remove(r);
            }
        }
        cv_.notify_all();
    }

    // called with mu_ held: keep at most max_keys_ - 1 other copies on `device`, idle least-recently-used first
    // makes room for one more key on `device`; false = every resident key there is busy and the device is at --max-keys
    bool evict_for(int device) {
        for (;;) {
            int n = 0;
            std::shared_ptr<Resident> lru;
            for (auto& r : table_)
                if (r->device == device) {
                    n++;
                    if (!r->busy && (!lru || r->last_used < lru->last_used)) lru = r;
                }
            if (n < max_keys_) return true;
            if (!lru) return false;
            if (lru->zk) g16_zkey_close(lru->zk);
            table_.remove(lru);
        }
    }

    size_t dev_index(int device) const {
        for (size_t i = 0; i < devices_.size(); i++)
            if (devices_[i] == device) return i;
        return 0;
    }

    void handle(int fd);
    std::string prove(const std::vector<std::string>& f);
    std::string stats();

    std::string sock_;
    std::vector<int> devices_;
    int max_keys_;
    int lfd_ = -1;
    std::atomic<bool> stop_{false};
    std::atomic<int> active_{0};
    std::mutex mu_;
    std::condition_variable cv_;
    std::list<std::shared_ptr<Resident>> table_;
    uint64_t clock_ = 0;
    std::atomic<uint64_t> n_proofs_{0}, n_opens_{0}, n_errors_{0};
};

}  // namespace g16svc
