// k2_mc_small for f32 scores (see k2_mc_small.cuh)
#include "k2_mc_small.cuh"

namespace tmb {
int launch_mc_small_f32(const TbArgs& a, int64_t rows, int C, int micro, int small_bins, int low_k, int smem,
                         cudaStream_t st) {
  return launch_small_p<TMB_F32>(a, rows, C, micro, small_bins, low_k, smem, st);
}
}  // namespace tmb
