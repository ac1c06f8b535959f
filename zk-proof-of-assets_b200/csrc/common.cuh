// Shared host-side helpers: CUDA error check and an owning device buffer.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <stdexcept>
#include <string>

namespace g16 {

#define G16_CUDA(x)                                                                                     \
    do {                                                                                                \
        cudaError_t e_ = (x);                                                                           \
        if (e_ != cudaSuccess)                                                                          \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " +    \
                                     __FILE__ + ":" + std::to_string(__LINE__));                        \
    } while (0)

// number of this library's own kernels launched so far (bench.py's gpu_launches; cub's kernels are not counted)
inline std::atomic<unsigned long long>& launch_counter() {
    static std::atomic<unsigned long long> c{0};
    return c;
}
#define G16_LAUNCHED(n) (::g16::launch_counter() += (n))

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    void alloc(size_t n) {
        free();
        if (n == 0) n = 16;
        G16_CUDA(cudaMalloc(&p, n));
        bytes = n;
    }
    void free() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
    ~DevBuf() { free(); }
    DevBuf() {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

}  // namespace g16
