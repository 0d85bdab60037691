// CMC / mAP for a q x g problem whose distance matrix is never materialised (evaluate.py:40-88 semantics).
// The matrix is produced panel by panel ([row chunk] x [column block]) by the contraction:
//   before   (host layer)          the distances of every query's same-identity gallery columns (its matches and its junk
//                                  = same identity AND same camera, evaluate.py:53-55) contracted as gathered pairs;
//            group_gather_kernel   the alternative: picked out of the panels of a first sweep (the re-ranked path does)
//            group_sort_kernel     per query: match keys (distance, column) sorted ascending, junk keys behind them
//   sweep    count_panel_kernel    streams each panel row once: an entry closer than the query's farthest match is binned
//                                  by the number of match keys below it (binary search over a handful of keys in shared
//                                  memory); bins are summed over the column blocks in global memory
//   finally  score_counts_kernel   prefix sums give every match's position among ALL entries; the junk in front of it is
//                                  subtracted -> kept rank -> AP (float64), first-hit rank -> CMC histogram
// Stable (distance, column) order like the rest of the path; counts are integers, so the result does not depend on the
// panel shape.  HBM-bound on the panel stream (4 bytes per pair per sweep); the contraction dominates.
#include "common.cuh"
#include "jaccard_common.cuh"
#include "kernels.h"

namespace reid {

// Same-identity gallery columns of one query are held in shared memory: the arrays are sized by the caller's longest
// identity group, rounded up to a power of two between 1 024 (the usual case: same footprint as a static array) and 8 192.
constexpr int SG_CAP_MIN = 1024, SG_CAP_MAX = 8192;
int stream_group_cap(int group_max) {
  int cap = SG_CAP_MIN;
  while (cap < group_max && cap < SG_CAP_MAX) cap <<= 1;
  return cap;
}

__global__ void __launch_bounds__(256)
group_lookup_kernel(const long long* __restrict__ q_pid, int nq, const unsigned long long* __restrict__ keys,
                    const long long* __restrict__ start, int T, int* __restrict__ glo, int* __restrict__ glen) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  const int slot = pid_find(keys, T, q_pid[i]);
  int lo = 0, len = 0;
  if (slot >= 0) { lo = (int)start[slot]; len = (int)(start[slot + 1] - start[slot]); }
  glo[i] = lo;
  glen[i] = len;
}

// one warp per row of the panel (rows = this chunk's queries, columns [c0, c1) of the gallery)
__global__ void __launch_bounds__(256)
group_gather_kernel(const float* __restrict__ panel, long long ld, int rows, int c0, int c1,
                    const int* __restrict__ glo, const int* __restrict__ glen, const long long* __restrict__ goff,
                    const int* __restrict__ list, float* __restrict__ gd) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lo = glo[row], len = glen[row];
  const long long off = goff[row];
  for (int e = lane_id(); e < len; e += 32) {
    const int col = list[lo + e];
    if (col >= c0 && col < c1) gd[off + e] = panel[(long long)row * ld + (col - c0)];
  }
}

__device__ __forceinline__ unsigned long long dist_key(float d, int col) {
  return ((unsigned long long)f2key(d + 0.0f) << 32) | (unsigned)col;
}

// one CTA per query: mkeys[goff .. goff+P) = sorted match keys, mkeys[goff+P .. goff+len) = junk keys; nmatch = P
__global__ void __launch_bounds__(256)
group_sort_kernel(int nq, const int* __restrict__ glo, const int* __restrict__ glen,
                  const long long* __restrict__ goff, const int* __restrict__ list, const float* __restrict__ gd,
                  const long long* __restrict__ q_cam, const long long* __restrict__ g_cam,
                  unsigned long long* __restrict__ mkeys, int* __restrict__ nmatch, int* __restrict__ overflow, int cap) {
  extern __shared__ __align__(16) unsigned char sg_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(sg_smem);            // [cap]
  __shared__ int s_p, s_j;
  const int qi = blockIdx.x, tid = threadIdx.x;
  const int lo = glo[qi], len = glen[qi];
  const long long off = goff[qi];
  if (len > cap) {
    if (tid == 0) { atomicMax(overflow, len); nmatch[qi] = 0; }
    return;
  }
  if (tid == 0) { s_p = 0; s_j = 0; }
  __syncthreads();
  const long long qc = q_cam[qi];
  // matches to the front of sk (any order), junk straight to the tail of the global segment
  for (int e = tid; e < len; e += 256) {
    const int col = list[lo + e];
    const unsigned long long k = dist_key(gd[off + e], col);
    if (g_cam[col] == qc) mkeys[off + len - 1 - atomicAdd(&s_j, 1)] = k;
    else sk[atomicAdd(&s_p, 1)] = k;
  }
  __syncthreads();
  const int P = s_p;
  int P2 = 1;
  while (P2 < P) P2 <<= 1;
  for (int i = P + tid; i < P2; i += 256) sk[i] = ~0ull;
  __syncthreads();
  for (int k = 2; k <= P2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int x = tid; x < (P2 >> 1); x += 256) {
        const int i = ((x & ~(j - 1)) << 1) | (x & (j - 1));
        const int p2 = i | j;
        const unsigned long long a = sk[i], b = sk[p2];
        if ((a > b) == ((i & k) == 0)) { sk[i] = b; sk[p2] = a; }
      }
      __syncthreads();
    }
  for (int i = tid; i < P; i += 256) mkeys[off + i] = sk[i];
  if (tid == 0) nmatch[qi] = P;
}

// one CTA per panel row: bins[goff + b] += #{entries of this block with exactly b match keys below them}, b < P.
// Entries at or below the farthest match's distance are counted in SG_NB cells over [0, dmax] (cell(d) is monotone in d,
// so an entry in a cell that holds no match lies on the same side of every match as the whole cell: one multiply, one
// convert, one shared-memory add); only entries that share a cell with a match take the exact (distance, column) search.
constexpr int SG_NB = 2048;
__device__ __forceinline__ int sg_cell(float d, float scale) {
  return (int)min(__float2uint_rz(__fmul_rn(d, scale)), (unsigned)(SG_NB - 1));       // negative -> 0
}

__global__ void __launch_bounds__(256)
count_panel_kernel(const float* __restrict__ panel, long long ld, int rows, int c0, int c1,
                   const long long* __restrict__ goff, const unsigned long long* __restrict__ mkeys,
                   const int* __restrict__ nmatch, int* __restrict__ bins, int cap) {
  extern __shared__ __align__(16) unsigned char sg_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(sg_smem);            // [cap]
  int* hist = reinterpret_cast<int*>(sg_smem + (size_t)cap * 8);                      // [cap]
  __shared__ int cells[SG_NB];
  const int row = blockIdx.x, tid = threadIdx.x;
  const int P = nmatch[row];
  if (P == 0) return;
  const long long off = goff[row];
  for (int i = tid; i < P; i += 256) { sk[i] = mkeys[off + i]; hist[i] = 0; }
  for (int i = tid; i < SG_NB; i += 256) cells[i] = 0;
  __syncthreads();
  const unsigned long long kmax = sk[P - 1];
  const float dmax = key2f((uint32_t)(kmax >> 32));
  const bool fast = dmax > 0.0f && dmax < INFINITY;      // otherwise (zero, inf, NaN) every entry is searched exactly
  float scale = 0.0f;
  if (fast) {
    scale = __fdiv_rn((float)SG_NB, dmax);
    if (!(scale < INFINITY)) scale = 0.0f;               // one cell: everything exact
    for (int i = tid; i < P; i += 256) cells[sg_cell(key2f((uint32_t)(sk[i] >> 32)), scale)] = (int)0x80000000;
  }
  __syncthreads();
  const float* r = panel + (long long)row * ld;
  const int n = c1 - c0;
  auto exact = [&](float d, int j) {
    const unsigned long long k = dist_key(d, c0 + j);
    if (k <= kmax) {
      int lo = 0, hi = P;                                // #{match keys < k}
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (sk[mid] < k) lo = mid + 1; else hi = mid;
      }
      atomicAdd(&hist[lo], 1);
    }
  };
  auto offer = [&](float d, int j) {
    if (fast) {
      if (d <= dmax && atomicAdd(&cells[sg_cell(d, scale)], 1) < 0) exact(d, j);
    } else if (!(d > dmax)) {
      exact(d, j);
    }
  };
  const int n4 = n >> 2;
  const float4* r4 = reinterpret_cast<const float4*>(r);
  for (int i = tid; i < n4; i += 256) {
    const float4 v = __ldg(r4 + i);
    if (!fast || fminf(fminf(v.x, v.y), fminf(v.z, v.w)) <= dmax) {
      offer(v.x, 4 * i); offer(v.y, 4 * i + 1); offer(v.z, 4 * i + 2); offer(v.w, 4 * i + 3);
    }
  }
  for (int j = 4 * n4 + tid; j < n; j += 256) offer(__ldg(r + j), j);
  __syncthreads();
  if (fast) {
    // a cell without a match: all its entries have exactly #{matches in lower cells} match keys below them
    for (int c = tid; c < SG_NB; c += 256) {
      const int cnt = cells[c];
      if (cnt > 0) {
        int lo = 0, hi = P;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (sg_cell(key2f((uint32_t)(sk[mid] >> 32)), scale) < c) lo = mid + 1; else hi = mid;
        }
        if (lo < P) atomicAdd(&hist[lo], cnt);
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < P; i += 256)
    if (hist[i]) bins[off + i] += hist[i];               // this CTA owns the row's bins: plain adds
}

// one thread per query
__global__ void __launch_bounds__(128)
score_counts_kernel(int nq, const int* __restrict__ glen, const long long* __restrict__ goff,
                    const unsigned long long* __restrict__ mkeys, const int* __restrict__ nmatch,
                    const int* __restrict__ bins, int max_rank, double* __restrict__ ap, int* __restrict__ first_hit,
                    int* __restrict__ cmc_hist, int* __restrict__ n_valid) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  const int P = nmatch[qi], len = glen[qi];
  if (P == 0) {                                          // identity absent from the gallery: skipped (evaluate.py:61-63)
    ap[qi] = -1.0;
    first_hit[qi] = -1;
    return;
  }
  const long long off = goff[qi];
  long long prefix = 0;
  double acc = 0.0;
  int first = 0;
  for (int i = 0; i < P; ++i) {
    prefix += bins[off + i];                             // entries at or before match i in (distance, column) order
    const unsigned long long k = mkeys[off + i];
    int junk_before = 0;
    for (int e = P; e < len; ++e) junk_before += mkeys[off + e] < k;
    const long long rank = prefix - junk_before;         // 1-based position among the kept (non-junk) entries
    if (i == 0) first = (int)(rank - 1);
    acc += (double)(i + 1) / (double)rank;
  }
  ap[qi] = acc / (double)P;
  first_hit[qi] = first;
  if (first < max_rank) atomicAdd(cmc_hist + first, 1);
  atomicAdd(n_valid, 1);
}

int launch_stream_group_lookup(const long long* q_pid, int nq, const unsigned long long* keys, const long long* start,
                               int T, int* glo, int* glen, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  group_lookup_kernel<<<cdiv(nq, 256), 256, 0, s>>>(q_pid, nq, keys, start, T, glo, glen);
  return check_launch("group_lookup_kernel");
}

int launch_stream_group_gather(const float* panel, long long ld, int rows, int c0, int c1, const int* glo,
                               const int* glen, const long long* goff, const int* list, float* gd, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  group_gather_kernel<<<cdiv(rows, 8), 256, 0, s>>>(panel, ld, rows, c0, c1, glo, glen, goff, list, gd);
  return check_launch("group_gather_kernel");
}

int launch_stream_group_sort(int nq, const int* glo, const int* glen, const long long* goff, const int* list,
                             const float* gd, const long long* q_cam, const long long* g_cam,
                             unsigned long long* mkeys, int* nmatch, int* overflow, int group_max, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  REID_REQUIRE(group_max <= SG_CAP_MAX, "stream_group_sort: an identity has %d gallery samples; at most %d per query",
               group_max, SG_CAP_MAX);
  const int cap = stream_group_cap(group_max);
  const size_t smem = (size_t)cap * 8;
  REID_CUDA(cudaFuncSetAttribute(group_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  group_sort_kernel<<<nq, 256, smem, s>>>(nq, glo, glen, goff, list, gd, q_cam, g_cam, mkeys, nmatch, overflow, cap);
  return check_launch("group_sort_kernel");
}

int launch_stream_count_panel(const float* panel, long long ld, int rows, int c0, int c1, const long long* goff,
                              const unsigned long long* mkeys, const int* nmatch, int* bins, int group_max,
                              cudaStream_t s) {
  if (rows == 0 || c1 <= c0) return REID_OK;
  REID_REQUIRE(group_max <= SG_CAP_MAX, "stream_count_panel: group_max=%d exceeds %d", group_max, SG_CAP_MAX);
  const int cap = stream_group_cap(group_max);
  const size_t smem = (size_t)cap * 12;
  REID_CUDA(cudaFuncSetAttribute(count_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  count_panel_kernel<<<rows, 256, smem, s>>>(panel, ld, rows, c0, c1, goff, mkeys, nmatch, bins, cap);
  return check_launch("count_panel_kernel");
}

int launch_stream_score(int nq, const int* glen, const long long* goff, const unsigned long long* mkeys,
                        const int* nmatch, const int* bins, int max_rank, double* ap, int* first_hit, int* cmc_hist,
                        int* n_valid, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  score_counts_kernel<<<cdiv(nq, 128), 128, 0, s>>>(nq, glen, goff, mkeys, nmatch, bins, max_rank, ap, first_hit,
                                                    cmc_hist, n_valid);
  return check_launch("score_counts_kernel");
}

}  // namespace reid
