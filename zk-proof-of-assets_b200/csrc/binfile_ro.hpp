// Read-only access to iden3 binfiles (.r1cs, .ptau, .wtns) for the setup-side commands: the section table is read from the
// headers alone and payload ranges are fetched with pread, so a 100 GB ceremony file costs what is read from it.
// Host code, no CUDA.  Used by csrc/zkey_new.cu and csrc/wtns_check.cu.
#pragma once
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>

namespace g16files {

struct FileRO {
    int fd = -1;
    uint64_t size = 0;
    std::string path;
    explicit FileRO(const char* p) : path(p) {
        fd = ::open(p, O_RDONLY);
        if (fd < 0) throw std::runtime_error(std::string("cannot open ") + p);
        struct stat st;
        if (fstat(fd, &st) != 0) throw std::runtime_error(std::string("cannot stat ") + p);
        size = (uint64_t)st.st_size;
    }
    ~FileRO() { if (fd >= 0) ::close(fd); }
    void read(uint64_t off, void* dst, size_t len) const {
        if (off > size || len > size - off) throw std::runtime_error(path + ": read past the end of the file");
        uint8_t* d = static_cast<uint8_t*>(dst);
        while (len) {
            ssize_t k = ::pread(fd, d, len, (off_t)off);
            if (k <= 0) throw std::runtime_error("cannot read " + path);
            d += k; off += (uint64_t)k; len -= (size_t)k;
        }
    }
};

struct Sec {
    uint64_t off = 0, size = 0;
    bool present = false;
};

// iden3 binfile: magic[4] | u32 version | u32 nSections | { u32 type | u64 size | payload }*; only the headers are read
inline std::map<uint32_t, Sec> scan_sections(const FileRO& f, const char* magic) {
    uint8_t head[12];
    if (f.size < 12) throw std::runtime_error(f.path + ": too short for a binfile");
    f.read(0, head, 12);
    if (std::memcmp(head, magic, 4) != 0) throw std::runtime_error(f.path + ": not a ." + magic + " file");
    uint32_t nsec;
    std::memcpy(&nsec, head + 8, 4);
    std::map<uint32_t, Sec> out;
    uint64_t pos = 12;
    for (uint32_t i = 0; i < nsec; i++) {
        uint8_t h[12];
        f.read(pos, h, 12);
        uint32_t typ;
        uint64_t size;
        std::memcpy(&typ, h, 4);
        std::memcpy(&size, h + 4, 8);
        pos += 12;
        if (size > f.size - pos) throw std::runtime_error(f.path + ": section runs past the end of the file");
        if (!out[typ].present) out[typ] = Sec{pos, size, true};
        pos += size;
    }
    return out;
}

inline const Sec& need_sec(const std::map<uint32_t, Sec>& m, uint32_t typ, const std::string& what) {
    auto it = m.find(typ);
    if (it == m.end() || !it->second.present) throw std::runtime_error(what + ": missing section " + std::to_string(typ));
    return it->second;
}

}  // namespace g16files
