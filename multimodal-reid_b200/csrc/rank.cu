// K11 (evaluate.py:40): full ranking of every row of the q x g distance matrix.
// Segmented LSD radix sort, one CTA per row, 8-bit digits, four passes, keys + column payload
// ping-ponged through global memory.  LSD passes are stable and the initial payload is the column
// index in ascending order, so equal distances keep ascending-index order (the tie-break the
// oracle uses; numpy's own argsort is unstable and unspecified there).
#include "common.cuh"
#include "kernels.h"

namespace reid {

constexpr int RK_T = 512;
constexpr int RK_W = RK_T / 32;

template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(RK_T)
rank_pass_kernel(const float* __restrict__ dist, long long ld_dist, const uint32_t* __restrict__ key_in,
                 const int* __restrict__ idx_in, uint32_t* __restrict__ key_out,
                 int* __restrict__ idx_out, int g, int shift) {
  __shared__ int hist[256];
  __shared__ int base_cur[256];
  __shared__ __align__(16) unsigned short whist[RK_W][256];
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const float* drow = dist + (long long)row * ld_dist;
  const uint32_t* kin = key_in + (long long)row * g;
  const int* iin = idx_in + (long long)row * g;
  uint32_t* kout = key_out + (long long)row * g;
  int* iout = idx_out + (long long)row * g;

  if (tid < 256) hist[tid] = 0;
  __syncthreads();
  // phase A: digit histogram of the row (warp-aggregated shared atomics)
  for (int j0 = 0; j0 < g; j0 += RK_T) {
    const int j = j0 + tid;
    int d = 0x1000 + lane;                       // inactive lanes: unique, matches nobody
    if (j < g) {
      uint32_t k = FIRST ? f2key(__fadd_rn(drow[j], 0.0f)) : kin[j];
      d = (k >> shift) & 255;
    }
    const unsigned peers = __match_any_sync(FULL, d);
    if (j < g && (peers & ((1u << lane) - 1)) == 0) atomicAdd(&hist[d], __popc(peers));
  }
  __syncthreads();
  // exclusive scan of the 256 counters (warps 0..7 hold 32 each) -> hist[d] = first output slot
  {
    __shared__ int wsum[8];
    int v = 0, x = 0;
    if (tid < 256) {
      v = hist[tid];
      x = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(FULL, x, o);
        if (lane >= o) x += y;
      }
      if (lane == 31) wsum[w] = x;
    }
    __syncthreads();
    if (tid < 256) {
      int add = 0;
      for (int ww = 0; ww < w; ++ww) add += wsum[ww];
      hist[tid] = x - v + add;
    }
    __syncthreads();
  }

  // phase B: stable scatter, tile by tile in row order
  for (int j0 = 0; j0 < g; j0 += RK_T) {
    reinterpret_cast<uint4*>(&whist[0][0])[tid] = make_uint4(0, 0, 0, 0);   // 8 KB / 512 threads
    __syncthreads();
    const int j = j0 + tid;
    const bool act = j < g;
    uint32_t k = 0;
    int payload = 0;
    int d = 0x1000 + lane;
    if (act) {
      if (FIRST) { k = f2key(__fadd_rn(drow[j], 0.0f)); payload = j; }
      else { k = kin[j]; payload = iin[j]; }
      d = (k >> shift) & 255;
    }
    const unsigned peers = __match_any_sync(FULL, d);
    const int rank_w = __popc(peers & ((1u << lane) - 1));
    if (act && rank_w == 0) whist[w][d] = (unsigned short)__popc(peers);
    __syncthreads();
    if (tid < 256) {
      int run = 0;
#pragma unroll
      for (int ww = 0; ww < RK_W; ++ww) {
        int c = whist[ww][tid];
        whist[ww][tid] = (unsigned short)run;
        run += c;
      }
      const int b = hist[tid];
      base_cur[tid] = b;
      hist[tid] = b + run;
    }
    __syncthreads();
    if (act) {
      const int pos = base_cur[d] + whist[w][d] + rank_w;
      if (!LAST) kout[pos] = k;
      iout[pos] = payload;
    }
    __syncthreads();
  }
}

int launch_rank_rows(const float* dist, int rows, int g, long long ld_dist, uint32_t* keyA, uint32_t* keyB,
                     int* idxA, int* idx_out, cudaStream_t s) {
  if (rows == 0 || g == 0) return REID_OK;
  static_assert(RK_W * 256 * 2 == RK_T * 16, "whist zeroing assumes one uint4 per thread");
  rank_pass_kernel<true, false><<<rows, RK_T, 0, s>>>(dist, ld_dist, nullptr, nullptr, keyA, idxA, g, 0);
  rank_pass_kernel<false, false><<<rows, RK_T, 0, s>>>(dist, ld_dist, keyA, idxA, keyB, idx_out, g, 8);
  rank_pass_kernel<false, false><<<rows, RK_T, 0, s>>>(dist, ld_dist, keyB, idx_out, keyA, idxA, g, 16);
  rank_pass_kernel<false, true><<<rows, RK_T, 0, s>>>(dist, ld_dist, keyA, idxA, keyB, idx_out, g, 24);
  count_launches(3);
  return check_launch("rank_pass_kernel");
}

}  // namespace reid
