// K13 prologue (evaluate.py:112-114, reranking.py:36-39): row L2-normalise, zero-pad the feature
// width to the MMA K-block multiple, emit the per-row squared norm / norm used by the distance
// epilogues, and split the rows into the float16 hi/lo operands of the tensor-core contraction:
//   x * 2^e = hi + lo,   hi = fp16(x * 2^e),   lo = fp16(x * 2^e - hi)
// with one power-of-two exponent e per matrix chosen so that max|x| * 2^e lies in [2^14, 2^15)
// (float16 keeps 11 significant bits down to 2^-14, so lo is a normal number for every element
// within 2^17 of the largest one).
#include "common.cuh"
#include "kernels.h"

namespace reid {

__device__ __forceinline__ int split_scale_exponent(uint32_t amax_bits) {   // == tc::scale_exponent
  if ((amax_bits & 0x7fffffffu) == 0) return 0;
  int E = (int)((amax_bits >> 23) & 0xff) - 127;
  int e = 14 - E;
  return max(-100, min(100, e));
}

// pass 1: one warp per row -> x32 (normalised, padded), sq, nrm, global max|x| (float bits)
// With a PRESET scale (amax == NULL, preset != NULL: the caller already knows a bound on max|x|, e.g. 1 for normalised
// rows) the float16 hi / lo operands are written here as well and the second pass is skipped: one read of X instead of a
// write + re-read of x32 (the prologue is replicated on every rank of a multi-GPU job, so its cost does not shrink with P).
// X2 != NULL: the row is the flip-TTA average (X + X2) / 2 of two embedding passes (prompt_learning.py:775-776,
// zero_shot_learning.py:121-125: `(emb + emb_aug) / 2` in float32 -- one rounded add, an exact halving), formed on
// the fly: the averaged matrix is never written or re-read.
__global__ void __launch_bounds__(256) prep_rows_kernel(const float* __restrict__ X, const float* __restrict__ X2,
                                                        long long ld_x2, int n, int d,
                                                        long long ld_x, int normalize,
                                                        float* __restrict__ x32, int dp,
                                                        float* __restrict__ sq, float* __restrict__ nrm,
                                                        uint32_t* __restrict__ amax, const uint32_t* __restrict__ preset,
                                                        __half* __restrict__ hi, __half* __restrict__ lo) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const int lane = lane_id();
  const float* x = X + (long long)row * ld_x;
  const float* x2 = X2 ? X2 + (long long)row * ld_x2 : nullptr;
  auto at = [&](int k) { return x2 ? __fmul_rn(__fadd_rn(x[k], x2[k]), 0.5f) : x[k]; };
  float ss = 0.0f;
  for (int k = lane; k < d; k += 32) {
    float v = at(k);
    ss = __fmaf_rn(v, v, ss);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss = __fadd_rn(ss, __shfl_xor_sync(FULL, ss, o));
  float denom = 1.0f;
  if (normalize) denom = fmaxf(sqrtf(ss), 1e-12f);     // F.normalize: x / max(||x||, eps)
  float ss2 = 0.0f, mx = 0.0f;
  float* o32 = x32 + (long long)row * dp;
  const float sc = preset ? exp2f((float)split_scale_exponent(*preset)) : 0.0f;      // exact power of two
  for (int k = lane; k < dp; k += 32) {
    float y = 0.0f;
    if (k < d) y = normalize ? __fdiv_rn(at(k), denom) : at(k);
    o32[k] = y;
    if (preset) {                                       // same arithmetic as split_f16_kernel
      const float a = y * sc;
      const __half h = __float2half_rn(a);
      hi[(long long)row * dp + k] = h;
      lo[(long long)row * dp + k] = __float2half_rn(__fsub_rn(a, __half2float(h)));
    }
    ss2 = __fmaf_rn(y, y, ss2);
    mx = fmaxf(mx, fabsf(y));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ss2 = __fadd_rn(ss2, __shfl_xor_sync(FULL, ss2, o));
    mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, o));
  }
  if (lane == 0) {
    sq[row] = ss2;
    nrm[row] = sqrtf(ss2);
    if (amax && mx == mx && mx < INFINITY) atomicMax(amax, __float_as_uint(mx));   // non-negative floats order as uints
  }
}

// pass 2: hi/lo float16 split with the matrix-wide exponent; 4 elements per thread
__global__ void __launch_bounds__(256) split_f16_kernel(const float* __restrict__ x32, long long total4,
                                                        const uint32_t* __restrict__ amax,
                                                        __half* __restrict__ hi, __half* __restrict__ lo) {
  const int e = split_scale_exponent(*amax);
  const float sc = exp2f((float)e);                    // exact power of two
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(x32)[i];
    const float a[4] = {v.x * sc, v.y * sc, v.z * sc, v.w * sc};
    __half h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = __float2half_rn(a[j]);
      l[j] = __float2half_rn(__fsub_rn(a[j], __half2float(h[j])));
    }
    reinterpret_cast<uint2*>(hi)[i] = make_uint2(
        (uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16),
        (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16));
    reinterpret_cast<uint2*>(lo)[i] = make_uint2(
        (uint32_t)__half_as_ushort(l[0]) | ((uint32_t)__half_as_ushort(l[1]) << 16),
        (uint32_t)__half_as_ushort(l[2]) | ((uint32_t)__half_as_ushort(l[3]) << 16));
  }
}

int launch_prep(const float* X, int n, int d, long long ld_x, int normalize, float* x32, __half* hi,
                __half* lo, int dp, float* sq, float* nrm, uint32_t* amax, cudaStream_t s) {
  return launch_prep_rows(X, nullptr, 0, n, d, ld_x, normalize, x32, hi, lo, dp, sq, nrm, amax, 0, s);
}

// preset_amax: *amax already holds the bits of a bound on max|x| of the WHOLE matrix these rows belong to (the streamed
// evaluator prepares the rows chunk by chunk, so the matrix-wide maximum is not known yet); it is used as is.
int launch_prep_rows(const float* X, const float* X2, long long ld_x2, int n, int d, long long ld_x, int normalize,
                     float* x32, __half* hi, __half* lo, int dp, float* sq, float* nrm, uint32_t* amax, int preset_amax,
                     cudaStream_t s) {
  if (!preset_amax) REID_CUDA(cudaMemsetAsync(amax, 0, sizeof(uint32_t), s));
  if (n == 0) return REID_OK;
  const int wpb = 8;
  prep_rows_kernel<<<cdiv(n, wpb), wpb * 32, 0, s>>>(X, X2, ld_x2, n, d, ld_x, normalize, x32, dp, sq, nrm,
                                                     preset_amax ? nullptr : amax, preset_amax ? amax : nullptr, hi, lo);
  int rc = check_launch("prep_rows_kernel");
  if (rc) return rc;
  if (preset_amax) return REID_OK;                      // operands written by the first pass
  const long long total4 = (long long)n * dp / 4;
  const int blocks = (int)((total4 + 255) / 256 < 148 * 8 ? (total4 + 255) / 256 : 148 * 8);
  split_f16_kernel<<<blocks, 256, 0, s>>>(x32, total4, amax, hi, lo);
  return check_launch("split_f16_kernel");
}

}  // namespace reid
