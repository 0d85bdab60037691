// Shared helpers for the reid_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/reid_b200.h"

namespace reid {

void set_error(const char* fmt, ...);
int check_launch(const char* what);            // counts one launch; cudaGetLastError -> REID_OK / REID_ERR_CUDA
void count_launches(int n);                    // extra launches of a multi-kernel launcher

#define REID_CUDA(call)                                                              \
  do {                                                                               \
    cudaError_t _e = (call);                                                         \
    if (_e != cudaSuccess) {                                                         \
      reid::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return REID_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

#define REID_REQUIRE(cond, ...)                                                      \
  do {                                                                               \
    if (!(cond)) {                                                                   \
      reid::set_error(__VA_ARGS__);                                                  \
      return REID_ERR_ARG;                                                           \
    }                                                                                \
  } while (0)

constexpr unsigned FULL = 0xffffffffu;

// float <-> order-preserving unsigned key (ascending float == ascending key).  -0.0 must be
// canonicalised by the caller (x + 0.0f) so that it ties with +0.0 like numpy's sort does.
__device__ __forceinline__ uint32_t f2key(float f) {
  uint32_t b = __float_as_uint(f);
  return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  uint32_t b = k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu);
  return __uint_as_float(b);
}
// float -> signed int with the same ordering (for atomicMax on ints)
__device__ __forceinline__ int f2ordered_int(float f) {
  int b = __float_as_int(f);
  return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_int2f(int k) {
  return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff);
}
#define REID_ROWMAX_INIT ((int)0x80000000)

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long w = __shfl_xor_sync(FULL, v, o);
    v = w < v ? w : v;
  }
  return v;
}

// ---------------------------------------------------------------------------------------------
// Row ownership of a multi-rank job.  Two schemes, selected by the sign of `row_stride` wherever a kernel
// takes (row0, row_stride):
//   row_stride > 0   cyclic rows: local row l of rank row0 is global row row0 + l * row_stride (NCCL path, round 1)
//   row_stride < 0   snake blocks: P = -row_stride ranks own 256-row tile rows of the symmetric N x N matrix in
//                    boustrophedon order (tile rows 0..P-1 -> ranks 0..P-1, tile rows P..2P-1 -> ranks P-1..0, ...), so
//                    the upper-triangle tiles (tile row t holds T - t of them) are balanced to O(P / T^2); local row
//                    l is row (l & 255) of the rank's (l >> 8)-th tile row.  P = 1 is the identity in both schemes.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline int snake_owner(int t, int P) {
  const int pos = t % (2 * P);
  return pos < P ? pos : 2 * P - 1 - pos;
}
__host__ __device__ inline int snake_local_tile(int t, int P) {       // index of tile row t among its owner's tile rows
  const int pos = t % (2 * P);
  return 2 * (t / (2 * P)) + (pos >= P ? 1 : 0);
}
__host__ __device__ inline int snake_global_tile(int lt, int r, int P) {
  return 2 * P * (lt >> 1) + ((lt & 1) ? 2 * P - 1 - r : r);
}
// number of tile rows t < t_end owned by rank r
__host__ __device__ inline int snake_tiles_below(int t_end, int r, int P) {
  const int pos = t_end % (2 * P);
  return 2 * (t_end / (2 * P)) + (pos > r ? 1 : 0) + (pos > 2 * P - 1 - r ? 1 : 0);
}
__host__ __device__ inline long long own_row(int row0, int row_stride, int l) {
  if (row_stride >= 0) return row0 + (long long)l * row_stride;
  return (long long)snake_global_tile(l >> 8, row0, -row_stride) * 256 + (l & 255);
}
// local row of global row i on its owner under the snake scheme
__host__ __device__ inline int snake_local_row(int i, int P) { return snake_local_tile(i >> 8, P) * 256 + (i & 255); }

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline long long round_up(long long a, long long b) { return (a + b - 1) / b * b; }

}  // namespace reid
