#include "../zk-proof-of-assets_b200/csrc/msm.cuh"
#include "../zk-proof-of-assets_b200/csrc/ntt.cuh"
using namespace g16;
template class g16::Msm<G1>;
template class g16::Msm<G2>;
void f(fe_t* a){ NttDomain d; d.init(10,0); d.to_odd_coset(a,0);} 
