#include "../zk-proof-of-assets_b200/csrc/curve.cuh"
using namespace g16;
#ifndef VAR
#define VAR 0
#endif
template <class C, int THREADS>
__device__ __forceinline__ void bsum(typename C::pt& acc) {
    __shared__ typename C::pt sh[THREADS];
#if VAR == 1
    volatile uint4* d = reinterpret_cast<volatile uint4*>(&sh[threadIdx.x]);
    const uint4* s4 = reinterpret_cast<const uint4*>(&acc);
    for (int i = 0; i < (int)(sizeof(typename C::pt) / 16); i++) { uint4 v = s4[i]; d[i].x = v.x; d[i].y = v.y; d[i].z = v.z; d[i].w = v.w; }
#elif VAR == 2
    sh[threadIdx.x] = acc;
    asm volatile("" ::: "memory");
#else
    sh[threadIdx.x] = acc;
#endif
    __syncthreads();
    for (int s = THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) {
            typename C::pt a = sh[threadIdx.x];
            C::add(a, sh[threadIdx.x + s]);
            sh[threadIdx.x] = a;
        }
        __syncthreads();
    }
    acc = sh[0];
}
template <class C>
__global__ void tree(const typename C::aff* pts, size_t n, typename C::pt* out) {
#if VAR == 4
    typename C::pt acc = C::from_affine(pts[n]);
#else
    typename C::pt acc = C::identity();
#endif
    for (size_t i = threadIdx.x; i < n; i += 128) {
        typename C::pt p = C::from_affine(pts[i]);
        C::add(acc, p);
    }
#if VAR == 5
    if (n == 12345) acc.x.v[0] = 7;
#endif
#if VAR == 3
    asm volatile("" ::: "memory");
#endif
    bsum<C, 128>(acc);
    if (threadIdx.x == 0) *out = acc;
}
template __global__ void tree<G1>(const g1_aff*, size_t, g1_pt*);
template __global__ void tree<G2>(const g2_aff*, size_t, g2_pt*);
