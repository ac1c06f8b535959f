#include <cstdio>
#include <cstring>
#include "../zk-proof-of-assets_b200/csrc/msm.cuh"
using namespace g16;
template <class C>
__global__ void tree_dyn(const typename C::aff* pts, size_t n, typename C::pt* out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    typename C::pt* sh = reinterpret_cast<typename C::pt*>(smem_raw);
    typename C::pt acc = C::identity();
    for (size_t i = threadIdx.x; i < n; i += 128) {
        typename C::pt p = C::from_affine(pts[i]);
        C::add(acc, p);
    }
    block_sum_points<C, 128>(acc);
    if (threadIdx.x == 0) *out = acc;
}
template <class C>
__global__ void tree_nosum(const typename C::aff* pts, size_t n, typename C::pt* out) {
    typename C::pt acc = C::identity();
    for (size_t i = threadIdx.x; i < n; i += 128) {
        typename C::pt p = C::from_affine(pts[i]);
        C::add(acc, p);
    }
    if (threadIdx.x == 0) *out = acc;
}
template <class C>
__global__ void tree_static(const typename C::aff* pts, size_t n, typename C::pt* out) {
    __shared__ typename C::pt sh[128];
    typename C::pt acc = C::identity();
    for (size_t i = threadIdx.x; i < n; i += 128) {
        typename C::pt p = C::from_affine(pts[i]);
        C::add(acc, p);
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x == 0) *out = sh[0];
}
int main() {
    g1_aff h; memset(&h, 0, sizeof h);
    for (int i = 0; i < 8; i++) { h.x.v[i] = FqParams::r1(i); h.y.v[i] = FqParams::r2(i); }
    g1_aff* dq; g1_pt* dout; cudaMalloc(&dq, sizeof h); cudaMalloc(&dout, 8 * sizeof(g1_pt));
    cudaMemcpy(dq, &h, sizeof h, cudaMemcpyHostToDevice);
    cudaMemset(dout, 0, 8 * sizeof(g1_pt));
    tree_dyn<G1><<<1, 128, 128 * sizeof(g1_pt)>>>(dq, 1, dout);
    printf("launch dyn: %s\n", cudaGetErrorString(cudaGetLastError()));
    tree_nosum<G1><<<1, 128>>>(dq, 1, dout + 1);
    tree_static<G1><<<1, 128>>>(dq, 1, dout + 2);
    printf("launch: %s\n", cudaGetErrorString(cudaGetLastError()));
    g1_pt o[8]; cudaError_t e = cudaMemcpy(o, dout, sizeof o, cudaMemcpyDeviceToHost);
    printf("err=%s\n", cudaGetErrorString(e));
    for (int i = 0; i < 3; i++)
        printf("out[%d] x %s y %s  x0=%08x y0=%08x zz0=%08x zzz0=%08x\n", i, memcmp(&o[i].x, &h.x, 32) == 0 ? "ok" : "BAD", memcmp(&o[i].y, &h.y, 32) == 0 ? "ok" : "BAD", o[i].x.v[0], o[i].y.v[0], o[i].zz.v[0], o[i].zzz.v[0]);
}
