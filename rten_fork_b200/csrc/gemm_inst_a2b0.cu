// Kernel variants of the contraction engine for A_KIND 2, B_MAJOR 0 (see gemm_launch.cuh).
#define RTG_GEMM_INSTANTIATE
#include "gemm_launch.cuh"

namespace rtg {
template rtg_status launch_variants<2, 0>(Context*, int, int, const CUtensorMap&, const CUtensorMap&,
                                            const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                            const GemmArgs&, const ReduceArgs*);
}  // namespace rtg
