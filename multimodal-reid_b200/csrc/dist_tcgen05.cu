// placeholder until the tcgen05 engine lands
#include "common.cuh"
#include "kernels.h"
namespace reid {
int launch_dist_tc(const float* Ahi, const float* Alo, const float* ra, int m, const float* Bhi,
                   const float* Blo, const float* rb, int n, int dp, int metric, float* out,
                   long long ld_out, int* rowmax_key, cudaStream_t s) {
  return launch_dist_simt(Ahi, Alo, ra, m, Bhi, Blo, rb, n, dp, metric, out, ld_out, rowmax_key, s);
}
}  // namespace reid
