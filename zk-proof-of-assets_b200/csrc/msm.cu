// The one translation unit that instantiates the Pippenger pipeline for G1 and G2 (see msm.cuh).
#define G16_MSM_INSTANTIATE
#include "msm.cuh"

namespace g16 {
template class Msm<G1>;
template class Msm<G2>;
}  // namespace g16
