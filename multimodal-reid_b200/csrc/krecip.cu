// K6 (reranking.py:51-71): k-reciprocal neighbour sets, the half-k expansion with the 2/3 rule,
// and the Gaussian-weighted sparse rows of V.
//
// The reference recomputes the half-depth reciprocal set of every candidate inside the loop over
// probes; that set depends on the candidate only, so it is built once per row here
// (reciprocal_sets_kernel, both depths in one sweep) and the per-probe kernel
// (expand_weights_kernel) only intersects and merges.  One warp per row; set algebra with
// ballots/shuffles; the neighbour table (N x K int32) is L2-resident.
#include "common.cuh"
#include "np_exact.cuh"
#include "kernels.h"

namespace reid {

// R_depth(i) = { nbr[i][r] : r < depth, i in nbr[nbr[i][r]][0:depth] }, kept in list order.
__global__ void __launch_bounds__(256)
reciprocal_sets_kernel(const int* __restrict__ nbr, int N, int K, int k1p, int h,
                       int* __restrict__ rfull, int* __restrict__ nfull,
                       int* __restrict__ rhalf, int* __restrict__ nhalf) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= N) return;
  const int lane = lane_id();
  const int* mine = nbr + (long long)row * K;
  int cf = 0, ch = 0;
  for (int r = 0; r < k1p; ++r) {
    const int c = mine[r];
    const int* theirs = nbr + (long long)c * K;
    bool hf = false, hh = false;
    for (int j = lane; j < k1p; j += 32) {
      bool eq = theirs[j] == row;
      hf |= eq;
      hh |= eq && (j < h);
    }
    const bool anyf = __any_sync(FULL, hf);
    const bool anyh = __any_sync(FULL, hh) && (r < h);
    if (lane == 0) {
      if (anyf) rfull[(long long)row * k1p + cf] = c;
      if (anyh) rhalf[(long long)row * h + ch] = c;
    }
    cf += anyf;
    ch += anyh;
  }
  if (lane == 0) { nfull[row] = cf; nhalf[row] = ch; }
}

// warp-level bitonic sort of P ints in shared memory (P power of two >= 32)
__device__ __forceinline__ void warp_bitonic_i32(int* a, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        int x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncwarp();
    }
  }
}

// per-warp shared memory: R list [k1p] | E list [capP] | weights [cap1]
__global__ void __launch_bounds__(128)
expand_weights_kernel(const int* __restrict__ rfull, const int* __restrict__ nfull,
                      const int* __restrict__ rhalf, const int* __restrict__ nhalf, int row0,
                      int row_stride, int rows, int k1p, int h, const float* __restrict__ D, long long ld,
                      const int* __restrict__ rowmax_key, int cap1, int capP,
                      int* __restrict__ Vc, __half* __restrict__ Vv, int* __restrict__ Vn) {
  extern __shared__ int smem_i[];
  const int wpb = blockDim.x >> 5;
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int lrow = blockIdx.x * wpb + w;          // local row: D / rowmax / V are indexed by it
  if (lrow >= rows) return;
  const int row = row0 + lrow * row_stride;       // global probe index
  const int per_warp = k1p + capP + cap1;
  int* R = smem_i + (long long)w * per_warp;
  int* E = R + k1p;
  float* W = reinterpret_cast<float*>(E + capP);

  const int nR = nfull[row];
  for (int j = lane; j < nR; j += 32) {
    int v = rfull[(long long)row * k1p + j];
    R[j] = v;
    E[j] = v;
  }
  __syncwarp();
  int nE = nR;
  for (int a = 0; a < nR; ++a) {
    const int c = R[a];
    const int nc = nhalf[c];
    int inter = 0;
    for (int j0 = 0; j0 < nc; j0 += 32) {
      const int j = j0 + lane;
      bool member = false;
      if (j < nc) {
        const int v = rhalf[(long long)c * h + j];
        for (int b = 0; b < nR; ++b) member |= (R[b] == v);
      }
      inter += __popc(__ballot_sync(FULL, member));
    }
    if (3 * inter > 2 * nc) {                      // len(intersect) > 2/3 * len(candidate set)
      for (int j = lane; j < nc; j += 32) E[nE + j] = rhalf[(long long)c * h + j];
      nE += nc;
    }
  }
  __syncwarp();
  // sort + unique (np.unique, reranking.py:69)
  int P = 32;
  while (P < nE) P <<= 1;
  for (int j = nE + lane; j < P; j += 32) E[j] = 0x7fffffff;
  __syncwarp();
  warp_bitonic_i32(E, P, lane);
  int nU = 0;
  for (int j0 = 0; j0 < nE; j0 += 32) {
    const int j = j0 + lane;
    int v = 0;
    bool keep = false;
    if (j < nE) {
      v = E[j];
      keep = (j == 0) || (E[j - 1] != v);
    }
    __syncwarp();
    const unsigned bal = __ballot_sync(FULL, keep);
    if (keep) E[nU + __popc(bal & ((1u << lane) - 1))] = v;
    nU += __popc(bal);
    __syncwarp();
  }
  // Gaussian weights on the normalised distance, numpy-exact exp and pairwise sum
  const float m = ordered_int2f(rowmax_key[lrow]);
  const float* drow = D + (long long)lrow * ld;
  for (int j = lane; j < nU; j += 32) {
    float o = __fdiv_rn(drow[E[j]], m);
    W[j] = np_expf(-o);
  }
  __syncwarp();
  const float s = np_sum_f32(W, nU);
  for (int j = lane; j < nU; j += 32) {
    Vc[(long long)lrow * cap1 + j] = E[j];
    Vv[(long long)lrow * cap1 + j] = __float2half_rn(__fdiv_rn(W[j], s));
  }
  if (lane == 0) Vn[lrow] = nU;
}

int launch_reciprocal_sets(const int* nbr, int N, int K, int k1, int* rfull, int* nfull, int* rhalf,
                           int* nhalf, cudaStream_t s) {
  if (N == 0) return REID_OK;
  const int k1p = k1 + 1, h = half_k(k1);
  REID_REQUIRE(K >= k1p, "reciprocal_sets: neighbour table width %d < k1+1=%d", K, k1p);
  const int wpb = 8;
  reciprocal_sets_kernel<<<cdiv(N, wpb), wpb * 32, 0, s>>>(nbr, N, K, k1p, h, rfull, nfull, rhalf,
                                                         nhalf);
  return check_launch("reciprocal_sets_kernel");
}

int launch_expand_weights(const int* rfull, const int* nfull, const int* rhalf, const int* nhalf,
                          int row0, int row_stride, int rows, int k1, const float* D, long long ld,
                          const int* rowmax_key, int* Vc, __half* Vv, int* Vn, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  const int k1p = k1 + 1, h = half_k(k1);
  const int cap1 = v_row_capacity(k1);
  int capP = 32;
  while (capP < cap1) capP <<= 1;
  const int per_warp = (k1p + capP + cap1) * 4;
  int wpb = 4;
  while (wpb > 1 && per_warp * wpb > 200 * 1024) wpb >>= 1;
  REID_REQUIRE(per_warp * wpb <= 227 * 1024, "expand_weights: k1=%d needs %d B of shared memory", k1,
               per_warp);
  const int smem = per_warp * wpb;
  REID_CUDA(cudaFuncSetAttribute(expand_weights_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  expand_weights_kernel<<<cdiv(rows, wpb), wpb * 32, smem, s>>>(rfull, nfull, rhalf, nhalf, row0,
                                                               row_stride, rows, k1p, h, D, ld,
                                                               rowmax_key, cap1, capP, Vc, Vv, Vn);
  return check_launch("expand_weights_kernel");
}

}  // namespace reid
