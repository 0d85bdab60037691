// Neighbour lists of the N x N matrix from HALF of it (SURVEY 8 a13, step A of stream_rerank.cu): the matrix of
// reranking.py:36-46 is symmetric, so the panel sweep contracts only the rectangles on and right of the diagonal blocks
// and every element D[r, c] of a rectangle (r in the panel's rows, c right of them) is offered to BOTH rows it belongs
// to -- row r takes it through the hinted per-panel selection (select.cu), row c through the kernels here:
//   sym_group_thr_kernel   per 32-column chunk of the rectangle: the largest running threshold tau[c] of its columns
//   sym_push_kernel        one warp per rectangle row and 8 192 columns: a chunk whose recorded minimum (the contraction's epilogue) can
//                          matter to one of its columns is read, and every element <= tau[column] is appended to that
//                          column's candidate records ((D, row) keys; arrival order is irrelevant)
//   sym_fold_kernel        one warp per column with candidates: running list (the Kw nearest seen so far, ascending in
//                          (D, column) order) + candidates, sorted, first Kw kept; tau[c] = the new Kw-th distance
//   sym_merge_kernel       multi-rank: every rank's lists of a row (its share of the row's entries) -> the Kw nearest
// tau[c] is always the Kw-th smallest of entries row c really has, so no entry of the final list is ever above it: the
// lists are exactly those of selection over the full row (tests/test_gpu_retrieval.py compares with the materialised
// symmetric path bit for bit).  Values are the upper-triangle ones, as in the materialised path's upper-only form.
#include "common.cuh"
#include "kernels.h"

namespace reid {

namespace {
__device__ __forceinline__ int sy_thr_key(float t) { return t == INFINITY ? 0x7fffffff : f2ordered_int(t); }
__device__ __forceinline__ unsigned long long sy_key(float d, int col) {
  return ((unsigned long long)f2key(__fadd_rn(d, 0.0f)) << 32) | (unsigned)col;
}
__device__ __forceinline__ void sy_warp_bitonic(unsigned long long* a, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int p = i | j;
        const unsigned long long x = a[i], y = a[p];
        if ((x > y) == ((i & k) == 0)) { a[i] = y; a[p] = x; }
      }
      __syncwarp();
    }
  }
}
}  // namespace

__global__ void __launch_bounds__(256)
sym_group_thr_kernel(const float* __restrict__ tau, int c0, int n, int* __restrict__ cthr_key) {
  const int g = blockIdx.x * 8 + (threadIdx.x >> 5), lane = lane_id();
  if (g * 32 >= n) return;
  const int col = g * 32 + lane;
  int k = col < n ? sy_thr_key(tau[c0 + col]) : (int)0x80000000;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) k = max(k, __shfl_xor_sync(FULL, k, off));
  if (lane == 0) cthr_key[g] = k;
}

constexpr int SY_WPB = 8;
constexpr int SY_SEG = 256;        // chunks (of 32 columns) per warp: one rectangle row is shared by ceil(nch / 256) warps
__global__ void __launch_bounds__(SY_WPB * 32)
sym_push_kernel(const float* __restrict__ D, long long ld, int m, int row0, int n, int c0,
                const float* __restrict__ chunk_min, int cm_pitch, const float* __restrict__ tau,
                const int* __restrict__ cthr_key, int* __restrict__ count, unsigned long long* __restrict__ pbuf,
                int cap) {
  __shared__ unsigned short list_s[SY_WPB][SY_SEG];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int lrow = blockIdx.x * SY_WPB + w;
  if (lrow >= m) return;
  unsigned short* list = list_s[w];
  const float* p = D + (long long)lrow * ld;
  const float* cm = chunk_min + (long long)lrow * cm_pitch;
  const int nch = (n + 31) >> 5;
  const unsigned grow = (unsigned)(row0 + lrow);
  const int cb = blockIdx.y * SY_SEG, cend = min(nch, cb + SY_SEG);
  int nlist = 0;
  for (int cc = cb; cc < cend; cc += 32) {                 // warp-uniform trip count
    const int c = cc + lane;
    const bool take = c < cend && f2ordered_int(__ldg(cm + c)) <= __ldg(cthr_key + c);
    const unsigned b = __ballot_sync(FULL, take);
    if (take) list[nlist + __popc(b & ((1u << lane) - 1))] = (unsigned short)(c - cb);
    nlist += __popc(b);
  }
  __syncwarp();
  for (int l0 = 0; l0 < nlist; l0 += 8) {
    float v[8], tj[8];
    int col[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      col[u] = l0 + u < nlist ? (cb + list[l0 + u]) * 32 + lane : n;
      const bool in = col[u] < n;
      v[u] = in ? __ldg(p + col[u]) : INFINITY;
      tj[u] = in ? __ldg(tau + c0 + col[u]) : -INFINITY;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (col[u] < n && v[u] <= tj[u]) {
        const int pos = atomicAdd(count + col[u], 1);
        if (pos < cap) pbuf[(long long)col[u] * cap + pos] = sy_key(v[u], (int)grow);
      }
    }
  }
}

constexpr int SF_WPB = 4;
constexpr int SF_MAX = 1024;       // keys per warp: Kw + cap rounded up to a power of two
__global__ void __launch_bounds__(SF_WPB * 32)
sym_fold_kernel(int n, int c0, int Kw, float* __restrict__ Mv, int* __restrict__ Mi, float* __restrict__ tau,
                int* __restrict__ count, const unsigned long long* __restrict__ pbuf, int cap,
                int* __restrict__ overflow) {
  __shared__ unsigned long long ks[SF_WPB][SF_MAX];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int j = blockIdx.x * SF_WPB + w;
  if (j >= n) return;
  int cnt_all = lane == 0 ? count[j] : 0;                  // one reader: lane 0 resets the counter below
  cnt_all = __shfl_sync(FULL, cnt_all, 0);
  if (cnt_all == 0) return;
  const int cnt = min(cnt_all, cap);
  if (lane == 0) {
    if (cnt_all > cap) atomicMax(overflow, cnt_all);
    count[j] = 0;
  }
  unsigned long long* a = ks[w];
  const long long row = (long long)c0 + j;
  for (int i = lane; i < Kw; i += 32) a[i] = sy_key(Mv[row * Kw + i], Mi[row * Kw + i]);
  for (int i = lane; i < cnt; i += 32) a[Kw + i] = pbuf[(long long)j * cap + i];
  int P2 = 32;
  while (P2 < Kw + cnt) P2 <<= 1;
  for (int i = Kw + cnt + lane; i < P2; i += 32) a[i] = ~0ull;
  __syncwarp();
  sy_warp_bitonic(a, P2, lane);
  for (int i = lane; i < Kw; i += 32) {
    Mv[row * Kw + i] = key2f((uint32_t)(a[i] >> 32));
    Mi[row * Kw + i] = (int)(unsigned)(a[i] & 0xffffffffu);
  }
  if (lane == 0) tau[row] = fminf(tau[row], key2f((uint32_t)(a[Kw - 1] >> 32)));
}

// S lists per row ([S, rows, Kw], each ascending; unused slots hold +inf) -> the Kw smallest keys
__global__ void __launch_bounds__(SF_WPB * 32)
sym_merge_kernel(int rows, int S, int Kw, const float* __restrict__ Fv, const int* __restrict__ Fi,
                 float* __restrict__ out_v, int* __restrict__ out_i) {
  __shared__ unsigned long long ks[SF_WPB][SF_MAX];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int row = blockIdx.x * SF_WPB + w;
  if (row >= rows) return;
  unsigned long long* a = ks[w];
  const int tot = S * Kw;
  for (int i = lane; i < tot; i += 32) {
    const int s = i / Kw, e = i - s * Kw;
    const long long at = ((long long)s * rows + row) * Kw + e;
    a[i] = sy_key(Fv[at], Fi[at]);
  }
  int P2 = 32;
  while (P2 < tot) P2 <<= 1;
  for (int i = tot + lane; i < P2; i += 32) a[i] = ~0ull;
  __syncwarp();
  sy_warp_bitonic(a, P2, lane);
  for (int i = lane; i < Kw; i += 32) {
    out_v[(long long)row * Kw + i] = key2f((uint32_t)(a[i] >> 32));
    out_i[(long long)row * Kw + i] = (int)(unsigned)(a[i] & 0xffffffffu);
  }
}

// ---------------------------------------------------------------------------------------------
int launch_sym_push(const float* D, long long ld, int m, int row0, int n, int c0, const float* chunk_min, int cm_pitch,
                    const float* tau, int* cthr_key, int* count, unsigned long long* pbuf, int cap, cudaStream_t s) {
  if (m == 0 || n == 0) return REID_OK;
  REID_REQUIRE(cap >= 1 && cm_pitch >= (n + 31) / 32, "sym_push: cap=%d cm_pitch=%d for %d columns", cap, cm_pitch, n);
  sym_group_thr_kernel<<<cdiv((n + 31) / 32, 8), 256, 0, s>>>(tau, c0, n, cthr_key);
  int rc = check_launch("sym_group_thr_kernel");
  if (rc) return rc;
  const dim3 grid(cdiv(m, SY_WPB), cdiv((n + 31) / 32, SY_SEG));
  sym_push_kernel<<<grid, SY_WPB * 32, 0, s>>>(D, ld, m, row0, n, c0, chunk_min, cm_pitch, tau, cthr_key, count, pbuf, cap);
  return check_launch("sym_push_kernel");
}

int launch_sym_fold(int n, int c0, int Kw, float* Mv, int* Mi, float* tau, int* count, const unsigned long long* pbuf,
                    int cap, int* overflow, cudaStream_t s) {
  if (n == 0) return REID_OK;
  REID_REQUIRE(Kw >= 1 && Kw + cap <= SF_MAX, "sym_fold: Kw=%d + cap=%d exceeds %d keys", Kw, cap, SF_MAX);
  sym_fold_kernel<<<cdiv(n, SF_WPB), SF_WPB * 32, 0, s>>>(n, c0, Kw, Mv, Mi, tau, count, pbuf, cap, overflow);
  return check_launch("sym_fold_kernel");
}

int launch_sym_merge(int rows, int S, int Kw, const float* Fv, const int* Fi, float* out_v, int* out_i, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(S >= 1 && Kw >= 1 && S * Kw <= SF_MAX, "sym_merge: %d lists of %d exceed %d keys", S, Kw, SF_MAX);
  sym_merge_kernel<<<cdiv(rows, SF_WPB), SF_WPB * 32, 0, s>>>(rows, S, Kw, Fv, Fi, out_v, out_i);
  return check_launch("sym_merge_kernel");
}

}  // namespace reid
