// K13 prologue (evaluate.py:112-114, reranking.py:36-39): row L2-normalise, zero-pad the feature
// width to the MMA K multiple, emit the TF32 hi / residual lo split the tensor-core contraction
// consumes, and the per-row squared norm / norm used by the distance epilogues.
#include "common.cuh"
#include "kernels.h"

namespace reid {

// one warp per row
__global__ void __launch_bounds__(256) prep_rows_kernel(const float* __restrict__ X, int n, int d,
                                                        long long ld_x, int normalize,
                                                        float* __restrict__ hi, float* __restrict__ lo,
                                                        int dp, float* __restrict__ sq,
                                                        float* __restrict__ nrm) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const int lane = lane_id();
  const float* x = X + (long long)row * ld_x;
  float ss = 0.0f;
  for (int k = lane; k < d; k += 32) {
    float v = x[k];
    ss = __fmaf_rn(v, v, ss);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss = __fadd_rn(ss, __shfl_xor_sync(FULL, ss, o));
  float denom = 1.0f;
  if (normalize) denom = fmaxf(sqrtf(ss), 1e-12f);     // F.normalize: x / max(||x||, eps)
  float ss2 = 0.0f;
  float* h = hi + (long long)row * dp;
  float* l = lo + (long long)row * dp;
  for (int k = lane; k < dp; k += 32) {
    float y = 0.0f;
    if (k < d) y = normalize ? __fdiv_rn(x[k], denom) : x[k];
    float yh = __uint_as_float(__float_as_uint(y) & 0xffffe000u);   // TF32: 1+8+10 bits
    h[k] = yh;
    l[k] = __fsub_rn(y, yh);                                         // exact residual
    ss2 = __fmaf_rn(y, y, ss2);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss2 = __fadd_rn(ss2, __shfl_xor_sync(FULL, ss2, o));
  if (lane == 0) {
    sq[row] = ss2;
    nrm[row] = sqrtf(ss2);
  }
}

int launch_prep(const float* X, int n, int d, long long ld_x, int normalize, float* hi, float* lo,
                int dp, float* sq, float* nrm, cudaStream_t s) {
  if (n == 0) return REID_OK;
  const int wpb = 8;
  prep_rows_kernel<<<cdiv(n, wpb), wpb * 32, 0, s>>>(X, n, d, ld_x, normalize, hi, lo, dp, sq, nrm);
  return check_launch("prep_rows_kernel");
}

}  // namespace reid
