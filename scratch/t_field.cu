#include "../zk-proof-of-assets_b200/csrc/curve.cuh"
using namespace g16;
__global__ void k_mul(const fe_t* a, const fe_t* b, fe_t* o, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st_fe(o + i, Fq::mul(ld_fe(a + i), ld_fe(b + i)));
}
__global__ void k_madd(g1_pt* acc, const g1_aff* q, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { g1_pt p = acc[i]; G1::madd(p, q[i]); acc[i] = p; }
}
