// K12 (evaluate.py:43-88): Market-1501 CMC / mAP from ranked gallery indices.
// One warp per query walks its ranked row 32 entries at a time; the junk filter (same pid AND same
// camera, evaluate.py:55), the kept-position counter and the match counter are warp ballots +
// popcount prefix scans; precision-at-match is accumulated in float64 like the reference.
#include "common.cuh"
#include "kernels.h"

namespace reid {

__global__ void __launch_bounds__(256)
score_rows_kernel(const int* __restrict__ ranked, int q, int g, const long long* __restrict__ q_pid,
                  const long long* __restrict__ g_pid, const long long* __restrict__ q_cam,
                  const long long* __restrict__ g_cam, int max_rank, double* __restrict__ ap,
                  int* __restrict__ first_hit, int* __restrict__ cmc_hist, int* __restrict__ n_valid) {
  const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= q) return;
  const int lane = lane_id();
  const unsigned le_mask = (lane == 31) ? FULL : ((2u << lane) - 1);
  const long long qp = q_pid[qi], qc = q_cam[qi];
  const int* r = ranked + (long long)qi * g;
  int kept = 0, hits = 0, first = -1;
  double acc = 0.0;
  for (int j0 = 0; j0 < g; j0 += 32) {
    const int j = j0 + lane;
    bool valid = false, match = false;
    if (j < g) {
      const int gi = r[j];
      const bool same = g_pid[gi] == qp;
      const bool junk = same && (g_cam[gi] == qc);
      valid = !junk;
      match = same && !junk;
    }
    const unsigned bv = __ballot_sync(FULL, valid);
    const unsigned bm = __ballot_sync(FULL, match);
    if (match) {
      const int pos = kept + __popc(bv & le_mask);       // 1-based rank among kept entries
      const int nth = hits + __popc(bm & le_mask);       // 1-based count of matches so far
      acc += (double)nth / (double)pos;
    }
    if (first < 0 && bm) {
      const int l = __ffs(bm) - 1;
      first = kept + __popc(bv & ((1u << l) - 1));       // 0-based rank of the first match
    }
    kept += __popc(bv);
    hits += __popc(bm);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  if (lane == 0) {
    if (hits > 0) {
      ap[qi] = acc / (double)hits;
      first_hit[qi] = first;
      if (first < max_rank) atomicAdd(cmc_hist + first, 1);
      atomicAdd(n_valid, 1);
    } else {
      ap[qi] = -1.0;                                     // query identity absent: skipped (:61-63)
      first_hit[qi] = -1;
    }
  }
}

// cmc[r] = float32(#queries whose first match is at rank <= r) / float32(n_valid);
// mAP = mean of the valid APs (float64, fixed-order tree).
__global__ void __launch_bounds__(1024)
score_finalize_kernel(const double* __restrict__ ap, int q, const int* __restrict__ cmc_hist,
                      int max_rank, const int* __restrict__ n_valid, float* __restrict__ cmc,
                      double* __restrict__ mAP) {
  __shared__ double red[1024];
  const int tid = threadIdx.x;
  double s = 0.0;
  for (int i = tid; i < q; i += 1024) {
    const double a = ap[i];
    if (a >= 0.0) s += a;
  }
  red[tid] = s;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  const int nv = *n_valid;
  if (tid == 0) {
    *mAP = nv > 0 ? red[0] / (double)nv : 0.0;
    int run = 0;
    const float nvf = (float)nv;
    for (int r = 0; r < max_rank; ++r) {
      run += cmc_hist[r];
      cmc[r] = nv > 0 ? __fdiv_rn((float)run, nvf) : 0.0f;
    }
  }
}

int launch_score(const int* ranked, int q, int g, const long long* q_pid, const long long* g_pid,
                 const long long* q_cam, const long long* g_cam, int max_rank, double* ap,
                 int* first_hit, int* cmc_hist, int* n_valid, cudaStream_t s) {
  if (q == 0) return REID_OK;
  score_rows_kernel<<<cdiv(q, 8), 256, 0, s>>>(ranked, q, g, q_pid, g_pid, q_cam, g_cam, max_rank, ap,
                                              first_hit, cmc_hist, n_valid);
  return check_launch("score_rows_kernel");
}

int launch_score_finalize(const double* ap, int q, const int* cmc_hist, int max_rank, const int* n_valid,
                          float* cmc, double* mAP, cudaStream_t s) {
  score_finalize_kernel<<<1, 1024, 0, s>>>(ap, q, cmc_hist, max_rank, n_valid, cmc, mAP);
  return check_launch("score_finalize_kernel");
}

}  // namespace reid
