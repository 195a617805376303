// Kernel variants of the contraction engine for A_KIND 0, B_MAJOR 1 (see gemm_launch.cuh).
#define RTG_GEMM_INSTANTIATE
#include "gemm_launch.cuh"

namespace rtg {
template rtg_status launch_variants<0, 1>(Context*, int, int, const CUtensorMap&, const CUtensorMap&,
                                            const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                            const GemmArgs&, const ReduceArgs*);
}  // namespace rtg
