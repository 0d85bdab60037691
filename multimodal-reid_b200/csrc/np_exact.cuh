// Device restatements of the NumPy float32 / float16 primitives the reference's re-ranking
// relies on (reranking.py:70-71, :76, :88-95), so the sparse kernels reproduce its rounding
// chain bit-for-bit.  Mirrors oracle/reid_oracle.py::{exp_f32_restated, pairwise_sum_f32_restated}
// which are themselves checked against numpy in tests/test_oracle.py.
#pragma once
#include "common.cuh"

namespace reid {

// numpy's SIMD float32 exp: Cody-Waite reduction by ln2, (5,2) rational, scale by 2^q.
// Every operation is an explicitly-rounded intrinsic so nvcc can neither contract nor reorder.
__device__ __forceinline__ float np_expf(float x) {
  const float LOG2E = 1.44269504088896340736f;
  const float MAGIC = 12582912.0f;  // 0x1.8p23
  float q = __fmul_rn(x, LOG2E);
  q = __fsub_rn(__fadd_rn(q, MAGIC), MAGIC);
  float r = __fmaf_rn(q, -6.93145752e-1f, x);
  r = __fmaf_rn(q, -1.42860677e-6f, r);
  float num = 5.082762527590693718096e-04f;
  num = __fmaf_rn(num, r, 6.757896990527504603057e-03f);
  num = __fmaf_rn(num, r, 5.114512081637298353406e-02f);
  num = __fmaf_rn(num, r, 2.473615434895520810817e-01f);
  num = __fmaf_rn(num, r, 7.257664613233124478488e-01f);
  num = __fmaf_rn(num, r, 9.999999999980870924916e-01f);
  float den = __fmaf_rn(2.159509375685829852307e-02f, r, -2.742335390411667452936e-01f);
  den = __fmaf_rn(den, r, 1.0f);
  float y = __fdiv_rn(num, den);
  return ldexpf(y, (int)q);
}

// numpy pairwise float32 add-reduce, leaf (n <= 128).
__device__ __forceinline__ float np_sum_leaf(const float* a, int n) {
  if (n < 8) {
    if (n == 0) return 0.0f;
    float s = a[0];
    for (int i = 1; i < n; ++i) s = __fadd_rn(s, a[i]);
    return s;
  }
  float r0 = a[0], r1 = a[1], r2 = a[2], r3 = a[3], r4 = a[4], r5 = a[5], r6 = a[6], r7 = a[7];
  int i = 8;
  const int lim = n - (n & 7);
  for (; i < lim; i += 8) {
    r0 = __fadd_rn(r0, a[i + 0]); r1 = __fadd_rn(r1, a[i + 1]);
    r2 = __fadd_rn(r2, a[i + 2]); r3 = __fadd_rn(r3, a[i + 3]);
    r4 = __fadd_rn(r4, a[i + 4]); r5 = __fadd_rn(r5, a[i + 5]);
    r6 = __fadd_rn(r6, a[i + 6]); r7 = __fadd_rn(r7, a[i + 7]);
  }
  float res = __fadd_rn(__fadd_rn(__fadd_rn(r0, r1), __fadd_rn(r2, r3)),
                        __fadd_rn(__fadd_rn(r4, r5), __fadd_rn(r6, r7)));
  for (; i < n; ++i) res = __fadd_rn(res, a[i]);
  return res;
}

// Full pairwise reduction: above 128 elements split at n/2 rounded down to a multiple of 8.
// Explicit post-order walk (no device recursion).
__device__ inline float np_sum_f32(const float* a, int n) {
  if (n <= 128) return np_sum_leaf(a, n);
  int off[20], len[20];
  unsigned char stage[20];
  float left[20];
  int sp = 0;
  off[0] = 0; len[0] = n; stage[0] = 0;
  float res = 0.0f;
  while (sp >= 0) {
    if (len[sp] <= 128) { res = np_sum_leaf(a + off[sp], len[sp]); --sp; continue; }
    int n2 = len[sp] >> 1; n2 -= (n2 & 7);
    if (stage[sp] == 0) {
      stage[sp] = 1;
      off[sp + 1] = off[sp]; len[sp + 1] = n2; stage[sp + 1] = 0; ++sp;
    } else if (stage[sp] == 1) {
      left[sp] = res; stage[sp] = 2;
      off[sp + 1] = off[sp] + n2; len[sp + 1] = len[sp] - n2; stage[sp + 1] = 0; ++sp;
    } else {
      res = __fadd_rn(left[sp], res); --sp;
    }
  }
  return res;
}

// numpy float16 arithmetic = float32 IEEE op + round-to-nearest-even to float16.
__device__ __forceinline__ __half h_add(__half a, __half b) {
  return __float2half_rn(__fadd_rn(__half2float(a), __half2float(b)));
}
__device__ __forceinline__ __half h_min(__half a, __half b) {
  return __half2float(a) <= __half2float(b) ? a : b;     // values are finite and non-negative
}
// 1 - t / (2 - t) with each step rounded to float16 (reranking.py:93)
__device__ __forceinline__ __half jaccard_from_overlap(__half t) {
  float tf = __half2float(t);
  __half den = __float2half_rn(__fsub_rn(2.0f, tf));
  __half quo = __float2half_rn(__fdiv_rn(tf, __half2float(den)));
  return __float2half_rn(__fsub_rn(1.0f, __half2float(quo)));
}

}  // namespace reid
