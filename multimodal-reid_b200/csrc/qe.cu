// K7 (reranking.py:73-78): local query expansion.  The reference averages k2 dense N-wide float16
// rows per probe (42 % of its run time); only the union of the k2 sparse supports is non-zero, so
// each warp gathers the k2 sparse rows into shared memory, sorts the (column, neighbour-rank)
// pairs and reduces each column in neighbour order:  fp16( fp32-sum / fp32(k2) ).
#include "common.cuh"
#include "np_exact.cuh"
#include "kernels.h"

namespace reid {

// upper bound on |support(V_qe[i])| = sum of the k2 source row lengths
__global__ void qe_bound_kernel(const int* __restrict__ nbr, int K, int k2, const int* __restrict__ Vn,
                                int row0, int row_stride, int rows, int* __restrict__ ub) {
  int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= rows) return;
  const long long i = own_row(row0, row_stride, l);
  int t = 0;
  for (int r = 0; r < k2; ++r) t += Vn[nbr[i * K + r]];
  ub[l] = t;
}

// Exclusive scan int32 -> int64 offsets (+ total at out[n]); n is 2e4 .. 2e6 (row counts, identity-index slots).
// Every CTA scans tiles of 4 K entries (4 per thread).  With several CTAs each publishes its own total first (epoch
// flag per launch) and then adds up the totals of all CTAs before it, 32 at a time -- one hop instead of a chain.
// At most 128 CTAs, all resident at once, and no CTA waits for a later one, so the look-back cannot deadlock.
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = 1024 * SCAN_ITEMS;
constexpr int SCAN_MAX_CTAS = 128;
__device__ unsigned long long g_scan_sum[SCAN_MAX_CTAS];      // total of CTA b alone
__device__ unsigned int g_scan_epoch[SCAN_MAX_CTAS];

__global__ void __launch_bounds__(1024)
exclusive_scan_kernel(const int* __restrict__ in, int n, long long* __restrict__ out, unsigned int epoch,
                      int tiles_per_cta) {
  __shared__ long long warp_tot[32];
  __shared__ long long chunk_tot[2];
  __shared__ long long s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int first = blockIdx.x * tiles_per_cta * SCAN_TILE;
  const int last = min(n, first + tiles_per_cta * SCAN_TILE);
  // pass 1 (only when chained): this CTA's total, published before the prefix of the earlier CTAs is known
  long long prefix = 0;
  if (gridDim.x > 1) {
    long long mine = 0;
    for (int i = first + tid; i < last; i += 1024) mine += in[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(FULL, mine, o);
    if (lane == 0) warp_tot[w] = mine;
    __syncthreads();
    if (w == 0) {
      long long t = warp_tot[lane];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
      if (lane == 0) {
        *(volatile unsigned long long*)(g_scan_sum + blockIdx.x) = (unsigned long long)t;
        __threadfence();
        *(volatile unsigned int*)(g_scan_epoch + blockIdx.x) = epoch;
      }
      long long before = 0;
      for (int j0 = 0; j0 < (int)blockIdx.x; j0 += 32) {          // totals of the CTAs before this one
        const int j = j0 + lane;
        if (j < (int)blockIdx.x) {
          volatile unsigned int* ep = g_scan_epoch + j;
          while (*ep != epoch) __nanosleep(20);
          __threadfence();
          before += (long long)*(volatile unsigned long long*)(g_scan_sum + j);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(FULL, before, o);
      if (lane == 0) s_prefix = before;
    }
    __syncthreads();
    prefix = s_prefix;
  }
  long long carry = prefix;
  int pass = 0;
  for (int base = first; base < last; base += SCAN_TILE, ++pass) {
    const int i0 = base + tid * SCAN_ITEMS;
    int v[SCAN_ITEMS];
    if (i0 + SCAN_ITEMS <= last && ((reinterpret_cast<uintptr_t>(in + i0) & 15) == 0)) {
#pragma unroll
      for (int k = 0; k < SCAN_ITEMS; k += 4) {
        const int4 x = __ldg(reinterpret_cast<const int4*>(in + i0 + k));
        v[k] = x.x; v[k + 1] = x.y; v[k + 2] = x.z; v[k + 3] = x.w;
      }
    } else {
#pragma unroll
      for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = i0 + k < last ? in[i0 + k] : 0;
    }
    long long mine = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) mine += v[k];
    long long x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      long long y = __shfl_up_sync(FULL, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
      const long long t = warp_tot[lane];
      long long s2 = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        long long y = __shfl_up_sync(FULL, s2, o);
        if (lane >= o) s2 += y;
      }
      warp_tot[lane] = s2 - t;     // exclusive warp offsets
      if (lane == 31) chunk_tot[pass & 1] = s2;
    }
    __syncthreads();
    long long run = carry + warp_tot[w] + x - mine;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
      if (i0 + k < last) out[i0 + k] = run;
      run += v[k];
    }
    carry += chunk_tot[pass & 1];
    __syncthreads();
  }
  if (tid == 0 && blockIdx.x == gridDim.x - 1) out[n] = carry;
}

int launch_exclusive_scan(const int* in, int n, long long* out, cudaStream_t s) {
  static unsigned int epoch = 0;
  const int tiles = n > 0 ? (n + SCAN_TILE - 1) / SCAN_TILE : 1;
  const int per_cta = (tiles + SCAN_MAX_CTAS - 1) / SCAN_MAX_CTAS;
  const int ctas = (tiles + per_cta - 1) / per_cta;
  if (++epoch == 0) ++epoch;                       // 0 is the state of a fresh module
  exclusive_scan_kernel<<<ctas, 1024, 0, s>>>(in, n, out, epoch, per_cta);
  return check_launch("exclusive_scan_kernel");
}

__device__ __forceinline__ void warp_bitonic_u64(unsigned long long* a, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        unsigned long long x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncwarp();
    }
  }
}

// entry = column << 32 | neighbour-rank << 16 | fp16 bits
__global__ void __launch_bounds__(256)
qe_expand_kernel(const int* __restrict__ nbr, int K, int k2, const int* __restrict__ Vc,
                 const __half* __restrict__ Vv, const int* __restrict__ Vn, int cap1, int row0,
                 int row_stride, int rows, const long long* __restrict__ qoff, int cap_sort, int* __restrict__ Qc,
                 __half* __restrict__ Qv, int* __restrict__ Qn) {
  extern __shared__ unsigned long long smem_q[];
  const int wpb = blockDim.x >> 5;
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int lrow = blockIdx.x * wpb + w;          // local row: qoff / Qn are indexed by it
  if (lrow >= rows) return;
  const long long row = own_row(row0, row_stride, lrow);
  const long long orow = row_stride < 0 ? row : lrow;   // snake ownership (sharded path): qoff / Qn are GLOBAL tables
  unsigned long long* buf = smem_q + (long long)w * cap_sort * 2;
  const int* src = nbr + row * K;
  const long long out0 = qoff[orow];
  const float kf = (float)k2;

  int total = 0;
  for (int r = 0; r < k2; ++r) total += Vn[src[r]];

  if (total <= cap_sort) {
    // The k2 rows are sorted by column already: instead of sorting the concatenation, every entry finds its place in
    // the merged order by binary searches in the other k2 - 1 lists (entries of equal column stay in neighbour order,
    // which is the order the reference's mean adds them in).  ~4x fewer instructions than a bitonic network here.
    unsigned long long* stage = buf + cap_sort;           // second half of this warp's buffer: the lists as loaded
    int my_off = 0, my_len = 0;                           // lane r < k2: start and length of list r in `stage`
    {
      int fill = 0;
      for (int r = 0; r < k2; ++r) {
        const int s = src[r];
        const int ns = Vn[s];
        for (int j = lane; j < ns; j += 32) {
          unsigned c = (unsigned)Vc[(long long)s * cap1 + j];
          unsigned short hv = __half_as_ushort(Vv[(long long)s * cap1 + j]);
          stage[fill + j] = ((unsigned long long)c << 32) | ((unsigned long long)r << 16) | hv;
        }
        if (lane == r) { my_off = fill; my_len = ns; }
        fill += ns;
      }
    }
    __syncwarp();
    for (int j0 = 0; j0 < total; j0 += 32) {
      const int j = j0 + lane;
      const bool on = j < total;
      const unsigned long long e = on ? stage[j] : 0ull;
      const unsigned col = (unsigned)(e >> 32);
      const int r = (int)((e >> 16) & 0xffffu);
      int rank = 0;
      for (int r2 = 0; r2 < k2; ++r2) {
        const int off2 = __shfl_sync(FULL, my_off, r2), len2 = __shfl_sync(FULL, my_len, r2);
        if (!on) continue;
        if (r2 == r) { rank += j - off2; continue; }
        int lo = 0, hi = len2;                            // #{entries of list r2 with column < col}
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if ((unsigned)(stage[off2 + mid] >> 32) < col) lo = mid + 1; else hi = mid;
        }
        if (r2 < r && lo < len2 && (unsigned)(stage[off2 + lo] >> 32) == col) ++lo;   // an equal column of an earlier list
        rank += lo;
      }
      if (on) buf[rank] = e;
    }
    __syncwarp();
    int nout = 0;
    for (int j0 = 0; j0 < total; j0 += 32) {
      const int j = j0 + lane;
      bool head = false;
      unsigned col = 0;
      float acc = 0.0f;
      if (j < total) {
        const unsigned long long e = buf[j];
        col = (unsigned)(e >> 32);
        head = (j == 0) || ((unsigned)(buf[j - 1] >> 32) != col);
        if (head) {
          acc = __half2float(__ushort_as_half((unsigned short)(e & 0xffffu)));
          for (int t = j + 1; t < total; ++t) {
            const unsigned long long e2 = buf[t];
            if ((unsigned)(e2 >> 32) != col) break;
            acc = __fadd_rn(acc, __half2float(__ushort_as_half((unsigned short)(e2 & 0xffffu))));
          }
        }
      }
      const unsigned bal = __ballot_sync(FULL, head);
      if (head) {
        const long long o = out0 + nout + __popc(bal & ((1u << lane) - 1));
        Qc[o] = (int)col;
        Qv[o] = __float2half_rn(__fdiv_rn(acc, kf));
      }
      nout += __popc(bal);
    }
    if (lane == 0) Qn[orow] = nout;
    return;
  }

  // Oversized union (does not fit the per-warp sort buffer): k2-way merge straight from global
  // memory, one list head per lane.  Latency-bound, exact, rare.
  int pos = 0, len = 0, s = 0;
  if (lane < k2) { s = src[lane]; len = Vn[s]; }
  int nout = 0;
  while (true) {
    unsigned head = (lane < k2 && pos < len) ? (unsigned)Vc[(long long)s * cap1 + pos] : 0xffffffffu;
    unsigned mn = head;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = min(mn, __shfl_xor_sync(FULL, mn, o));
    if (mn == 0xffffffffu) break;
    const bool hit = head == mn;
    float v = hit ? __half2float(Vv[(long long)s * cap1 + pos]) : 0.0f;
    unsigned bal = __ballot_sync(FULL, hit);
    float acc = 0.0f;
    bool first = true;
    while (bal) {
      const int l = __ffs(bal) - 1;
      const float t = __shfl_sync(FULL, v, l);
      acc = first ? t : __fadd_rn(acc, t);
      first = false;
      bal &= bal - 1;
    }
    if (lane == 0) {
      Qc[out0 + nout] = (int)mn;
      Qv[out0 + nout] = __float2half_rn(__fdiv_rn(acc, kf));
    }
    ++nout;
    if (hit) ++pos;
  }
  if (lane == 0) Qn[orow] = nout;
}

int launch_qe_bound(const int* nbr, int K, int k2, const int* Vn, int row0, int row_stride, int rows,
                    int* ub, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(k2 >= 1 && k2 <= K && k2 <= 32, "qe: k2=%d outside [1, min(K=%d, 32)]", k2, K);
  qe_bound_kernel<<<cdiv(rows, 256), 256, 0, s>>>(nbr, K, k2, Vn, row0, row_stride, rows, ub);
  return check_launch("qe_bound_kernel");
}

int launch_qe_expand(const int* nbr, int K, int k2, const int* Vc, const __half* Vv, const int* Vn,
                     int cap1, int row0, int row_stride, int rows, const long long* qoff, int cap_sort,
                     int* Qc, __half* Qv, int* Qn, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(k2 >= 1 && k2 <= K && k2 <= 32, "qe: k2=%d outside [1, min(K=%d, 32)]", k2, K);
  REID_REQUIRE(cap_sort >= 32 && (cap_sort & (cap_sort - 1)) == 0 && cap_sort <= 8192,
               "qe: sort capacity %d must be a power of two in [32, 8192]", cap_sort);
  int wpb = 8;                                        // keep a CTA's merge buffers <= 32 KB so several CTAs share an SM
  while (wpb > 1 && (long long)wpb * cap_sort * 16 > 32 * 1024) wpb >>= 1;
  const int smem = wpb * cap_sort * 16;              // per warp: merged order | the lists as loaded
  REID_CUDA(cudaFuncSetAttribute(qe_expand_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  qe_expand_kernel<<<cdiv(rows, wpb), wpb * 32, smem, s>>>(nbr, K, k2, Vc, Vv, Vn, cap1, row0,
                                                          row_stride, rows, qoff, cap_sort, Qc, Qv, Qn);
  return check_launch("qe_expand_kernel");
}

}  // namespace reid
