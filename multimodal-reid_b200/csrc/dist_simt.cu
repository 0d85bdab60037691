// Distance contraction, CUDA-core engine (fp32 FMA on the float32 rows).  Bring-up and
// cross-check engine for the tcgen05 path in dist_tc2.cu; same fused epilogues:
//   metric 0: D = (|a|^2 + |b|^2) - 2 a.b          evaluate.py:7-13, reranking.py:36-41
//   metric 1: arccos(clip(a.b * (1/(|a||b|))))     evaluate.py:16-26
// plus an optional per-row running maximum (reranking.py:46 needs max_k D[k,i]).
#include "common.cuh"
#include "kernels.h"

namespace reid {

constexpr int TM = 128, TN = 128, TK = 16;

__device__ __forceinline__ float dist_epilogue(int metric, float acc, float ra, float rb) {
  if (metric == 0) return __fmaf_rn(-2.0f, acc, __fadd_rn(ra, rb));
  // reciprocal of the norm product, then multiply (not a division), clip, arccos
  float recip = __fdiv_rn(1.0f, __fmul_rn(ra, rb));
  float c = __fmul_rn(acc, recip);
  const float hi = (float)(1.0 - 0.00001), lo = (float)(-1.0 + 0.00001);   // np.clip bounds: float32(double)
  c = fminf(fmaxf(c, lo), hi);
  return acosf(c);
}

__global__ void __launch_bounds__(256)
dist_simt_kernel(const float* __restrict__ A, long long lda, const float* __restrict__ ra, int m,
                 const float* __restrict__ B, long long ldb, const float* __restrict__ rb, int n, int dp,
                 int metric, float* __restrict__ out, long long ld_out, int* __restrict__ rowmax_key) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;

  // each thread stages 2 float4 of A and 2 of B per K-slab: rows (tid>>2) and (tid>>2)+64
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  float4 pa[2], pb[2];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int r = m0 + lr + h * 64;
      if (r < m) {
        pa[h] = *reinterpret_cast<const float4*>(A + (long long)r * lda + k0 + lk);
      } else {
        pa[h] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      int c = n0 + lr + h * 64;
      if (c < n) {
        pb[h] = *reinterpret_cast<const float4*>(B + (long long)c * ldb + k0 + lk);
      } else {
        pb[h] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < dp; k0 += TK) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int r = lr + h * 64;
      As[lk + 0][r] = pa[h].x; As[lk + 1][r] = pa[h].y; As[lk + 2][r] = pa[h].z; As[lk + 3][r] = pa[h].w;
      Bs[lk + 0][r] = pb[h].x; Bs[lk + 1][r] = pb[h].y; Bs[lk + 2][r] = pb[h].z; Bs[lk + 3][r] = pb[h].w;
    }
    __syncthreads();
    if (k0 + TK < dp) fetch(k0 + TK);
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float a[8], b[8];
      *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      *reinterpret_cast<float4*>(b) = *reinterpret_cast<const float4*>(&Bs[k][tx * 8]);
      *reinterpret_cast<float4*>(b + 4) = *reinterpret_cast<const float4*>(&Bs[k][tx * 8 + 4]);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = __fmaf_rn(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  float rbv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int c = n0 + tx * 8 + j;
    rbv[j] = c < n ? rb[c] : 0.0f;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = m0 + ty * 8 + i;
    float rav = r < m ? ra[r] : 0.0f;
    float v[8];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[j] = dist_epilogue(metric, acc[i][j], rav, rbv[j]);
      if (n0 + tx * 8 + j < n) mx = fmaxf(mx, v[j]);
    }
    if (r < m) {
      float* o = out + (long long)r * ld_out + n0 + tx * 8;
      if (n0 + tx * 8 + 7 < n) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (n0 + tx * 8 + j < n) o[j] = v[j];
      }
    }
    if (rowmax_key) {
      // the 16 tx-threads of a row are 16 consecutive lanes
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, o));
      if (tx == 0 && r < m && mx > -INFINITY) atomicMax(rowmax_key + r, f2ordered_int(mx));
    }
  }
}

int launch_dist_simt(const float* A, long long lda, const float* ra, int m, const float* B, long long ldb,
                     const float* rb, int n, int dp, int metric, float* out, long long ld_out,
                     int* rowmax_key, cudaStream_t s) {
  if (m == 0 || n == 0) return REID_OK;
  REID_REQUIRE(dp % TK == 0, "dist_simt: padded width %d must be a multiple of %d", dp, TK);
  REID_REQUIRE(ld_out % 4 == 0 && lda % 4 == 0 && ldb % 4 == 0, "dist_simt: pitches must be multiples of 4");
  dim3 grid(cdiv(n, TN), cdiv(m, TM));
  dist_simt_kernel<<<grid, 256, 0, s>>>(A, lda, ra, m, B, ldb, rb, n, dp, metric, out, ld_out,
                                        rowmax_key);
  return check_launch("dist_simt_kernel");
}

}  // namespace reid
