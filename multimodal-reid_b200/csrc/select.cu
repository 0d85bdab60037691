// K4 + K5 (reranking.py:46, :48): the reference divides the N x N matrix by its column maxima,
// transposes, and fully argsorts every row, but only ever reads the first max(k1+1, k2) columns
// of that argsort.  Here: per-row maximum (fused into the GEMM epilogue on the main path, or the
// stand-alone kernel below for a caller-supplied matrix), then an exact per-row top-K selection
// over O[i, j] = D[i, j] / rowmax[i] with stable (value, index) ordering.  HBM-bound: one read
// of the row from DRAM, a second one out of L2.
#include "common.cuh"
#include "kernels.h"

namespace reid {

// ---------------------------------------------------------------------------------------------
// row maximum of a pitched matrix -> ordered-int key (same encoding the GEMM epilogue uses)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rowmax_kernel(const float* __restrict__ D, int rows, int cols, long long ld, int* __restrict__ key) {
  const int row = blockIdx.x;
  const float* p = D + (long long)row * ld;
  float mx = -INFINITY;
  for (int c = threadIdx.x * 4; c < cols; c += blockDim.x * 4) {
    if (c + 3 < cols) {
      float4 v = *reinterpret_cast<const float4*>(p + c);
      mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    } else {
      for (int j = c; j < cols; ++j) mx = fmaxf(mx, p[j]);
    }
  }
  mx = warp_max(mx);
  __shared__ float sm[8];
  if (lane_id() == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : -INFINITY;
    v = warp_max(v);
    if (threadIdx.x == 0) key[row] = f2ordered_int(v);
  }
}

int launch_rowmax(const float* D, int rows, int cols, long long ld, int* key, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(ld % 4 == 0, "rowmax: pitch %lld must be a multiple of 4", ld);
  rowmax_kernel<<<rows, 256, 0, s>>>(D, rows, cols, ld, key);
  return check_launch("rowmax_kernel");
}

// ---------------------------------------------------------------------------------------------
// exact top-K (smallest) per row with (value, column) ordering
// ---------------------------------------------------------------------------------------------
constexpr int SEL_CAP = 4096;   // candidate slots (8 B each)

__device__ __forceinline__ unsigned long long make_ck(float d, float m, int col) {
  float o = __fadd_rn(__fdiv_rn(d, m), 0.0f);          // O = D / max ; -0.0 -> +0.0
  return ((unsigned long long)f2key(o) << 32) | (unsigned)col;
}

// block-wide bitonic sort of P (power of two) 64-bit keys in shared memory, ascending
template <int T>
__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* a, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += T) {
        // pair index t -> element i with bit j clear
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        unsigned long long x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncthreads();
    }
  }
}

template <int T>
__global__ void __launch_bounds__(T)
select_topk_kernel(const float* __restrict__ D, int rows, int cols, long long ld,
                   const int* __restrict__ rowmax_key, int K, int* __restrict__ nbr,
                   float* __restrict__ nbr_val, int cap_override) {
  __shared__ unsigned long long buf[SEL_CAP];
  __shared__ unsigned long long tau_s;
  __shared__ int cnt_s;
  const int row = blockIdx.x;
  const int tid = threadIdx.x;
  const float* p = D + (long long)row * ld;
  const float m = ordered_int2f(rowmax_key[row]);
  const int cap = cap_override > 0 ? min(cap_override, SEL_CAP) : SEL_CAP;

  // pass 1: every thread keeps the minimum of its strided share of the row.  The K-th smallest
  // of those T minima bounds the K-th smallest element from above, and at most K threads own
  // anything at or below it.
  unsigned long long best = ~0ull;
  for (int c = tid * 4; c < cols; c += T * 4) {
    if (c + 3 < cols) {
      float4 v = *reinterpret_cast<const float4*>(p + c);
      unsigned long long k0 = make_ck(v.x, m, c), k1 = make_ck(v.y, m, c + 1),
                         k2 = make_ck(v.z, m, c + 2), k3 = make_ck(v.w, m, c + 3);
      k0 = k0 < k1 ? k0 : k1; k2 = k2 < k3 ? k2 : k3; k0 = k0 < k2 ? k0 : k2;
      best = best < k0 ? best : k0;
    } else {
      for (int j = c; j < cols; ++j) {
        unsigned long long k0 = make_ck(p[j], m, j);
        best = best < k0 ? best : k0;
      }
    }
  }
  buf[tid] = best;
  if (tid == 0) cnt_s = 0;
  __syncthreads();
  bitonic_sort_u64<T>(buf, T);
  if (tid == 0) tau_s = buf[K - 1];
  __syncthreads();
  const unsigned long long tau = tau_s;
  __syncthreads();

  // pass 2 (row now L2-resident): collect everything at or below the bound
  for (int c = tid * 4; c < cols; c += T * 4) {
    if (c + 3 < cols) {
      float4 v = *reinterpret_cast<const float4*>(p + c);
      float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        unsigned long long k0 = make_ck(vv[j], m, c + j);
        if (k0 <= tau) {
          int pos = atomicAdd(&cnt_s, 1);
          if (pos < cap) buf[pos] = k0;
        }
      }
    } else {
      for (int j = c; j < cols; ++j) {
        unsigned long long k0 = make_ck(p[j], m, j);
        if (k0 <= tau) {
          int pos = atomicAdd(&cnt_s, 1);
          if (pos < cap) buf[pos] = k0;
        }
      }
    }
  }
  __syncthreads();
  const int cnt = cnt_s;
  if (cnt <= cap) {
    int P = 32;
    while (P < cnt) P <<= 1;
    for (int t = cnt + tid; t < P; t += T) buf[t] = ~0ull;
    __syncthreads();
    bitonic_sort_u64<T>(buf, P);
    for (int t = tid; t < K; t += T) {
      unsigned long long k0 = buf[t];
      nbr[(long long)row * K + t] = (int)(unsigned)(k0 & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + t] = key2f((uint32_t)(k0 >> 32));
    }
    return;
  }

  // overflow (adversarial layouts / massive ties): K rounds of block-wide arg-min, each strictly
  // above the previous pick.  Slow, exact, essentially never taken.
  __shared__ unsigned long long red[T / 32];
  unsigned long long last = 0;
  for (int r = 0; r < K; ++r) {
    unsigned long long mine = ~0ull;
    for (int c = tid; c < cols; c += T) {
      unsigned long long k0 = make_ck(p[c], m, c);
      if ((r == 0 || k0 > last) && k0 < mine) mine = k0;
    }
    mine = warp_min_u64(mine);
    if ((tid & 31) == 0) red[tid >> 5] = mine;
    __syncthreads();
    if (tid < 32) {
      unsigned long long v = tid < (T / 32) ? red[tid] : ~0ull;
      v = warp_min_u64(v);
      if (tid == 0) red[0] = v;
    }
    __syncthreads();
    last = red[0];
    __syncthreads();
    if (tid == 0) {
      nbr[(long long)row * K + r] = (int)(unsigned)(last & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + r] = key2f((uint32_t)(last >> 32));
    }
  }
}

int launch_select_topk(const float* D, int rows, int cols, long long ld, const int* rowmax_key, int K,
                       int* nbr, float* nbr_val, int cap_override, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= 256, "select_topk: K=%d outside [1,256]", K);
  REID_REQUIRE(cols >= K, "select_topk: row length %d is shorter than K=%d", cols, K);
  REID_REQUIRE(ld % 4 == 0, "select_topk: pitch %lld must be a multiple of 4", ld);
  if (cols <= 32768) {
    select_topk_kernel<256><<<rows, 256, 0, s>>>(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val,
                                                 cap_override);
  } else {
    select_topk_kernel<1024><<<rows, 1024, 0, s>>>(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val,
                                                   cap_override);
  }
  return check_launch("select_topk_kernel");
}

}  // namespace reid
