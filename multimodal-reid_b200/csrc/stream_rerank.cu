// k-reciprocal re-ranking when neither the N x N nor the q x g matrix can exist (100k queries x 1M gallery;
// reranking.py:46-100 and evaluate.py:40-88 with D, V and the final matrix never dense).  The contraction produces
// panels; everything here works on lists:
//   lists_rekey_kernel      per row: its WIDE raw-distance neighbour list (top-Kw by (D, column), from the N x N panel
//                           sweep) re-keyed with the row's global maximum -> initial_rank[:, :K] in (O = D / max, column)
//                           order (reranking.py:46,48); the margin Kw - K covers distinct D whose quotients collide
//   (reciprocal sets, V weights, query expansion, inverted index: the kernels of the materialised path; the V weights read
//    their distances from the wide lists -- krecip.cu, wide-list mode)
//   overlap_hash_kernel     per query: the ordered float16 Jaccard chain (reranking.py:86-92) into a shared-memory hash
//                           table keyed by target instead of an N-bit bitmap (137 KB at N = 1.1 M); output = the query's
//                           gallery targets, sorted, with their overlap values
//   (distances of the overlap targets and of the same-identity columns: gathered from the q x g panels, stream_score.cu)
//   rr_prepare_kernel       per query: final values (reranking.py:93-95) of its same-identity columns, match keys sorted,
//                           raw-distance thresholds of every match for entries the Jaccard term does not touch
//   count_panel_rr_kernel   second q x g sweep: every entry binned by the matches it precedes AS IF untouched
//   score_counts_rr_kernel  per query: bins corrected by the special entries (same identity, overlap targets: counted
//                           as-if-untouched above, re-entered with their true final keys, junk dropped) -> kept ranks ->
//                           AP, first hit (evaluate.py:53-80)
//   rr_topk_kernel          per query: exact top-K of the re-ranked row = merge(top-Kd by raw distance, overlap targets)
#include "common.cuh"
#include "jaccard_common.cuh"
#include "kernels.h"

namespace reid {

__device__ __forceinline__ unsigned long long sr_ck(float d, float m, int col) {      // (O key, column), as select.cu
  const float o = __fadd_rn(__fdiv_rn(d, m), 0.0f);
  return ((unsigned long long)f2key(o) << 32) | (unsigned)col;
}
__device__ __forceinline__ void sr_warp_bitonic(unsigned long long* a, int P, int lane) {
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
template <int T>
__device__ __forceinline__ void sr_block_bitonic(unsigned long long* a, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += T) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int p = i | j;
        const unsigned long long x = a[i], y = a[p];
        if ((x > y) == ((i & k) == 0)) { a[i] = y; a[p] = x; }
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
constexpr int LR_WPB = 4;
__global__ void __launch_bounds__(LR_WPB * 32)
lists_rekey_kernel(const int* __restrict__ widx, const float* __restrict__ wval, int rows, int Kw,
                   const int* __restrict__ rowmax_key, int K, int* __restrict__ nbr, int* __restrict__ ambiguous) {
  __shared__ unsigned long long ks[LR_WPB][256];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int row = blockIdx.x * LR_WPB + w;
  if (row >= rows) return;
  unsigned long long* a = ks[w];
  const float m = ordered_int2f(rowmax_key[row]);
  int P2 = 32;
  while (P2 < Kw) P2 <<= 1;
  for (int i = lane; i < P2; i += 32)
    a[i] = i < Kw ? sr_ck(wval[(long long)row * Kw + i], m, widx[(long long)row * Kw + i]) : ~0ull;
  __syncwarp();
  sr_warp_bitonic(a, P2, lane);
  for (int i = lane; i < K; i += 32) nbr[(long long)row * K + i] = (int)(unsigned)(a[i] & 0xffffffffu);
  // every candidate beyond the raw-distance list has D >= the list's last D, hence O >= its O: the first K are exact
  // unless the K-th O equals the largest O of the list (then an unseen entry could tie with a smaller column)
  if (lane == 0 && Kw > K && (a[K - 1] >> 32) == (a[Kw - 1] >> 32)) atomicAdd(ambiguous, 1);
}

int launch_lists_rekey(const int* widx, const float* wval, int rows, int Kw, const int* rowmax_key, int K, int* nbr,
                       int* ambiguous, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= Kw && Kw <= 256, "lists_rekey: K=%d Kw=%d (K <= Kw <= 256)", K, Kw);
  lists_rekey_kernel<<<cdiv(rows, LR_WPB), LR_WPB * 32, 0, s>>>(widx, wval, rows, Kw, rowmax_key, K, nbr, ambiguous);
  return check_launch("lists_rekey_kernel");
}

// ---------------------------------------------------------------------------------------------
// Jaccard overlap of one query through the inverted index, hash-table form (any N)
// ---------------------------------------------------------------------------------------------
constexpr int OH_T = 256;
constexpr int OH_SLOTS = 16384;          // hash slots (power of two)
constexpr int OH_CAP = 8192;             // distinct gallery targets per query (half load)

__global__ void __launch_bounds__(OH_T)
overlap_hash_kernel(const int* __restrict__ Qc, const __half* __restrict__ Qv, const long long* __restrict__ qoff,
                    const int* __restrict__ Qn, const long long* __restrict__ colptr,
                    const unsigned long long* __restrict__ csc_ent, int q0, int q_stride, int q, int N, int cap_o,
                    int nrange, int* __restrict__ ocol, __half* __restrict__ ot, int* __restrict__ ocnt,
                    int* __restrict__ overflow) {
  extern __shared__ __align__(16) unsigned char oh_smem[];
  int* keys = reinterpret_cast<int*>(oh_smem);                                   // [OH_SLOTS] target or -1
  __half* tv = reinterpret_cast<__half*>(oh_smem + (size_t)OH_SLOTS * 4);         // [OH_SLOTS] running overlap
  unsigned long long* comp = reinterpret_cast<unsigned long long*>(oh_smem + (size_t)OH_SLOTS * 6);   // [OH_CAP]
  __shared__ int s_new, s_n, s_bad;
  const int qi = blockIdx.x, tid = threadIdx.x;
  const long long row = own_row(q0, q_stride, qi);
  const long long off = qoff[row];
  const int nq_cols = Qn[row];
  const int g = N - q;
  const int span = (g + nrange - 1) / nrange;               // gallery columns per pass
  int written = 0;                                           // entries already in the output (uniform)
  // A query at the 1M-gallery scale overlaps tens of thousands of gallery rows -- more than a shared-memory table
  // holds -- so the gallery is cut into `nrange` column ranges and the chain is run once per range: every pass walks
  // all of the query's column lists but keeps only the targets of its range (each target's additions still come in
  // ascending column order), then emits them sorted; ranges ascend, so the concatenation is sorted too.
  for (int rg = 0; rg < nrange; ++rg) {
    const int glo = q + rg * span, ghi = min(N, glo + span);
    for (int i = tid; i < OH_SLOTS; i += OH_T) { keys[i] = -1; tv[i] = __ushort_as_half(0); }
    if (tid == 0) { s_new = 0; s_n = 0; s_bad = 0; }
    __syncthreads();
    for (int ci = 0; ci < nq_cols; ++ci) {                   // ascending columns: the reference's order of additions
      const int c = Qc[off + ci];
      const float vq = __half2float(Qv[off + ci]);
      const long long p0 = colptr[c];
      const int len = (int)(colptr[c + 1] - p0);
      for (int p = tid; p < len; p += OH_T) {               // entries of one column hit distinct targets
        const unsigned long long ent = csc_ent[p0 + p];
        const int gt = (int)(unsigned)(ent & 0xffffffffu);
        if (gt < glo || gt >= ghi) continue;                 // only gallery columns (of this range) reach final[:q, q:]
        const float v = __half2float(__ushort_as_half((unsigned short)(ent >> 32)));
        unsigned slot = ((unsigned)gt * 2654435761u) >> (32 - 14);
        bool placed = false;
        for (int probe = 0; probe < OH_SLOTS; ++probe) {
          const int cur = atomicCAS(&keys[slot], -1, gt);
          if (cur == -1) { atomicAdd(&s_new, 1); placed = true; break; }
          if (cur == gt) { placed = true; break; }
          slot = (slot + 1) & (OH_SLOTS - 1);
        }
        if (!placed) { s_bad = 1; continue; }                // table full: reported below, the query is refused
        tv[slot] = __float2half_rn(__fadd_rn(__half2float(tv[slot]), fminf(vq, v)));   // fp16 add: fp32 add, RN to fp16
      }
      __syncthreads();                                       // the next column's additions come after this one's
    }
    const int nnew = s_new, bad = s_bad;                     // (the last column's barrier ordered these)
    if (bad || nnew > OH_CAP || written + nnew > cap_o) {
      if (tid == 0) { ocnt[qi] = -1; atomicAdd(overflow, 1); }
      return;
    }
    for (int i = tid; i < OH_SLOTS; i += OH_T) {
      const int gt = keys[i];
      if (gt >= 0) comp[atomicAdd(&s_n, 1)] = ((unsigned long long)(unsigned)(gt - q) << 16) | __half_as_ushort(tv[i]);
    }
    __syncthreads();
    const int n = s_n;
    int P2 = 32;
    while (P2 < n) P2 <<= 1;
    for (int i = n + tid; i < P2; i += OH_T) comp[i] = ~0ull;
    __syncthreads();
    sr_block_bitonic<OH_T>(comp, P2);
    for (int i = tid; i < n; i += OH_T) {
      const unsigned long long e = comp[i];
      ocol[(long long)qi * cap_o + written + i] = (int)(e >> 16);
      ot[(long long)qi * cap_o + written + i] = __ushort_as_half((unsigned short)(e & 0xffffu));
    }
    written += n;
    __syncthreads();                                         // comp / the table are reused by the next range
  }
  static_assert(OH_SLOTS == 1 << 14, "hash shift");
  if (tid == 0) ocnt[qi] = written;
}

int launch_overlap_hash(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn, const long long* colptr,
                        const unsigned long long* csc_ent, int q0, int q_stride, int nq, int q, int N, int cap_o,
                        int nrange, int* ocol, __half* ot, int* ocnt, int* overflow, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  REID_REQUIRE(cap_o >= 1 && q >= 0 && N > q && nrange >= 0, "overlap_hash: cap_o=%d q=%d N=%d", cap_o, q, N);
  // default: ranges sized so that a query filling its whole budget still puts only ~half of OH_CAP into one table; a
  // caller that knows how many targets its queries really touch passes fewer (every range costs one walk of the lists)
  if (nrange == 0) nrange = (cap_o + OH_CAP / 2 - 1) / (OH_CAP / 2);
  if (nrange > N - q) nrange = N - q;
  const size_t smem = (size_t)OH_SLOTS * 6 + (size_t)OH_CAP * 8;
  REID_CUDA(cudaFuncSetAttribute(overlap_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  overlap_hash_kernel<<<nq, OH_T, smem, s>>>(Qc, Qv, qoff, Qn, colptr, csc_ent, q0, q_stride, q, N, cap_o, nrange, ocol,
                                              ot, ocnt, overflow);
  return check_launch("overlap_hash_kernel");
}

// distances of a query's listed columns from one panel; the lists are ascending, so only the slice that falls into the
// panel's column block [c0, c1) is visited.  One warp per panel row.
__global__ void __launch_bounds__(256)
list_gather_sorted_kernel(const float* __restrict__ panel, long long ld, int rows, int c0, int c1, int cap_o,
                          const int* __restrict__ ocol, const int* __restrict__ ocnt, float* __restrict__ od) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int n = ocnt[row];
  const int* oc = ocol + (long long)row * cap_o;
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (oc[mid] < c0) lo = mid + 1; else hi = mid; }
  const int a = lo;
  hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (oc[mid] < c1) lo = mid + 1; else hi = mid; }
  for (int e = a + lane_id(); e < lo; e += 32)
    od[(long long)row * cap_o + e] = panel[(long long)row * ld + (oc[e] - c0)];
}

int launch_list_gather_sorted(const float* panel, long long ld, int rows, int c0, int c1, int cap_o, const int* ocol,
                              const int* ocnt, float* od, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  list_gather_sorted_kernel<<<cdiv(rows, 8), 256, 0, s>>>(panel, ld, rows, c0, c1, cap_o, ocol, ocnt, od);
  return check_launch("list_gather_sorted_kernel");
}

// ---------------------------------------------------------------------------------------------
// per-query preparation of the counting sweep
// ---------------------------------------------------------------------------------------------
// same-identity gallery columns per query: shared-memory arrays of stream_group_cap(longest group) entries, 1 024 ... 8 192
// (as stream_score.cu)
constexpr int RR_CAP_MAX = 8192;

// overlap value of gallery column `col` from the query's sorted target list (0 when the query does not touch it)
__device__ __forceinline__ __half rr_overlap(const int* __restrict__ oc, const __half* __restrict__ ot, int n, int col) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (oc[mid] < col) lo = mid + 1; else hi = mid;
  }
  return (lo < n && oc[lo] == col) ? ot[lo] : __ushort_as_half(0);
}

// bin of an entry that the Jaccard term does not touch: the number of matches it comes AFTER (jaccard_sparse.cu ss_zero_bin)
__device__ __forceinline__ int rr_zero_bin(float d, int col, const float* thrA, const float* thrB,
                                           const unsigned long long* mk, int P) {
  int lo = 0, hi = P;                                     // #{i : d > thrB[i]}
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (d > thrB[mid]) lo = mid + 1; else hi = mid;
  }
  int bin = lo;
  for (int i = lo; i < P && thrA[i] < d; ++i) {           // exact ties in final: the smaller column first
    if ((int)(unsigned)(mk[i] & 0xffffffffu) < col) bin = i + 1; else break;
  }
  return bin;
}

// one CTA per query.  Group arrays (glo/glen/goff/list/gd) as in stream_score.cu; overlap lists at qi * cap_o.
// Out, at goff[qi]: gkey[e] = true final key of group entry e (junk: bit 63 set... no -- see gjunk), gjunk[e];
// mk[0..P) sorted match keys, thrA/thrB[0..P) raw-distance thresholds (strict / non-strict) of every match, nmatch.
__global__ void __launch_bounds__(256)
rr_prepare_kernel(int nq, int q0, int q_stride, const int* __restrict__ glo, const int* __restrict__ glen,
                  const long long* __restrict__ goff, const int* __restrict__ list, const float* __restrict__ gd,
                  const long long* __restrict__ q_cam, const long long* __restrict__ g_cam,
                  const int* __restrict__ rowmax_key, float c1, float c2, int cap_o, const int* __restrict__ ocol,
                  const __half* __restrict__ ot, const int* __restrict__ ocnt, unsigned long long* __restrict__ gkey,
                  unsigned char* __restrict__ gjunk, unsigned long long* __restrict__ mk, float* __restrict__ thrA,
                  float* __restrict__ thrB, int* __restrict__ nmatch, int* __restrict__ overflow, int cap) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(rr_smem);            // [cap]
  __shared__ int s_p;
  const int qi = blockIdx.x, tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int lo = glo[qi], len = glen[qi];
  const long long off = goff[qi];
  const int no = ocnt[qi];
  if (len > cap || no < 0) {
    if (tid == 0) { if (len > cap) atomicMax(overflow, len); nmatch[qi] = 0; }
    return;
  }
  if (tid == 0) s_p = 0;
  __syncthreads();
  const float m = ordered_int2f(rowmax_key[own_row(q0, q_stride, qi)]);
  const float c1h = __half2float(__float2half_rn(c1));
  const long long qc = q_cam[qi];
  const int* oc = ocol + (long long)qi * cap_o;
  const __half* otv = ot + (long long)qi * cap_o;
  for (int e = tid; e < len; e += 256) {
    const int col = list[lo + e];
    const float f = final_value(rr_overlap(oc, otv, no, col), gd[off + e], m, c1h, c2);
    const unsigned long long k = final_key(f, col);
    const bool junk = g_cam[col] == qc;
    gkey[off + e] = k;
    gjunk[off + e] = junk ? 1 : 0;
    if (!junk) sk[atomicAdd(&s_p, 1)] = k;
  }
  __syncthreads();
  const int P = s_p;
  int P2 = 1;
  while (P2 < P) P2 <<= 1;
  for (int i = P + tid; i < P2; i += 256) sk[i] = ~0ull;
  __syncthreads();
  if (P2 > 1) sr_block_bitonic<256>(sk, P2);
  for (int i = tid; i < P; i += 256) mk[off + i] = sk[i];
  const bool mono = m > 0.0f && m < INFINITY && c2 > 0.0f && c2 < INFINITY;
  for (int x = w; x < 2 * P; x += 8) {                    // one warp per threshold
    const int i = x >> 1;
    const float f = key2f((uint32_t)(sk[i] >> 32));
    // degenerate maximum / lambda or a NaN final: no monotone shortcut -- thresholds that admit nothing (the entries
    // of such a row are then only its special ones; refused upstream by the non-finite check of the panel sweep)
    const float thr = (mono && f == f) ? final0_threshold_warp(f, (x & 1) != 0, m, c1h, c2) : -INFINITY;
    if (lane == 0) { if (x & 1) thrA[off + i] = thr; else thrB[off + i] = thr; }
  }
  if (tid == 0) nmatch[qi] = P;
}

// ---------------------------------------------------------------------------------------------
// second sweep: bins[goff + b] += #{entries of this block that come after exactly b matches, as if untouched}
// ---------------------------------------------------------------------------------------------
constexpr int RC_NB = 2048;
__device__ __forceinline__ int rc_cell(float d, float scale) {
  return (int)min(__float2uint_rz(__fmul_rn(d, scale)), (unsigned)(RC_NB - 1));       // negative, NaN -> 0
}

__global__ void __launch_bounds__(256)
count_panel_rr_kernel(const float* __restrict__ panel, long long ld, int rows, int c0, int c1,
                      const long long* __restrict__ goff, const unsigned long long* __restrict__ mk,
                      const float* __restrict__ thrA, const float* __restrict__ thrB, const int* __restrict__ nmatch,
                      int* __restrict__ bins, int cap) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(rr_smem);            // [cap] each
  float* sA = reinterpret_cast<float*>(rr_smem + (size_t)cap * 8);
  float* sB = sA + cap;
  int* hist = reinterpret_cast<int*>(sB + cap);
  __shared__ int cells[RC_NB];
  const int row = blockIdx.x, tid = threadIdx.x;
  const int P = nmatch[row];
  if (P == 0) return;
  const long long off = goff[row];
  for (int i = tid; i < P; i += 256) { sk[i] = mk[off + i]; sA[i] = thrA[off + i]; sB[i] = thrB[off + i]; hist[i] = 0; }
  for (int i = tid; i < RC_NB; i += 256) cells[i] = 0;
  __syncthreads();
  const float dthr = sB[P - 1];                          // an untouched entry above it comes after every match
  const bool fast = dthr > 0.0f && dthr < INFINITY;
  float scale = 0.0f;
  if (fast) {
    scale = __fdiv_rn((float)RC_NB, dthr);
    if (!(scale < INFINITY)) scale = 0.0f;
    for (int i = tid; i < P; i += 256) {                 // cells holding a threshold (or a tie range): ranked exactly
      const int hi = rc_cell(sB[i], scale);
      for (int x = rc_cell(sA[i], scale); x <= hi; ++x) cells[x] = (int)0x80000000;
    }
  }
  __syncthreads();
  const float* r = panel + (long long)row * ld;
  const int n = c1 - c0;
  auto offer = [&](float d, int j) {
    if (!(d <= dthr)) return;
    if (fast && atomicAdd(&cells[rc_cell(d, scale)], 1) >= 0) return;       // a clean cell: folded below
    const int bin = rr_zero_bin(d, c0 + j, sA, sB, sk, P);
    if (bin < P) atomicAdd(&hist[bin], 1);
  };
  const int n4 = n >> 2;
  const float4* r4 = reinterpret_cast<const float4*>(r);
  for (int i = tid; i < n4; i += 256) {
    const float4 v = __ldg(r4 + i);
    if (fminf(fminf(v.x, v.y), fminf(v.z, v.w)) <= dthr) {
      offer(v.x, 4 * i); offer(v.y, 4 * i + 1); offer(v.z, 4 * i + 2); offer(v.w, 4 * i + 3);
    }
  }
  for (int j = 4 * n4 + tid; j < n; j += 256) offer(__ldg(r + j), j);
  __syncthreads();
  if (fast) {
    for (int c = tid; c < RC_NB; c += 256) {              // a clean cell c: its entries come after #{i : cell(thrB[i]) < c} matches
      const int cnt = cells[c];
      if (cnt > 0) {
        int lo = 0, hi = P;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (rc_cell(sB[mid], scale) < c) lo = mid + 1; else hi = mid;
        }
        if (lo < P) atomicAdd(&hist[lo], cnt);
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < P; i += 256)
    if (hist[i]) bins[off + i] += hist[i];                // this CTA owns the row's bins: plain adds
}

// ---------------------------------------------------------------------------------------------
// kept ranks from the bins + the special entries' corrections
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int rr_find(const int* __restrict__ oc, int n, int col) {       // index in the sorted list or -1
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (oc[mid] < col) lo = mid + 1; else hi = mid;
  }
  return (lo < n && oc[lo] == col) ? lo : -1;
}

__global__ void __launch_bounds__(128)
score_counts_rr_kernel(int nq, int q0, int q_stride, const int* __restrict__ glo, const int* __restrict__ glen,
                       const long long* __restrict__ goff, const int* __restrict__ list, const float* __restrict__ gd,
                       const unsigned long long* __restrict__ gkey, const unsigned char* __restrict__ gjunk,
                       const unsigned long long* __restrict__ mk, const float* __restrict__ thrA,
                       const float* __restrict__ thrB, const int* __restrict__ nmatch, const int* __restrict__ bins,
                       int cap_o, const int* __restrict__ ocol, const __half* __restrict__ ot,
                       const float* __restrict__ od, const int* __restrict__ ocnt, const long long* __restrict__ q_pid,
                       const long long* __restrict__ g_pid, const int* __restrict__ rowmax_key, float c1, float c2,
                       int max_rank, double* __restrict__ ap, int* __restrict__ first_hit, int* __restrict__ cmc_hist,
                       int* __restrict__ n_valid, int cap) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(rr_smem);            // [cap] each
  float* sA = reinterpret_cast<float*>(rr_smem + (size_t)cap * 8);
  float* sB = sA + cap;
  int* corr = reinterpret_cast<int*>(sB + cap);
  const int qi = blockIdx.x, tid = threadIdx.x;
  const int P = nmatch[qi];
  if (P == 0) {                                          // identity absent from the gallery: skipped (evaluate.py:61-63)
    if (tid == 0) { ap[qi] = -1.0; first_hit[qi] = -1; }
    return;
  }
  const long long off = goff[qi];
  const int lo = glo[qi], len = glen[qi];
  for (int i = tid; i < P; i += 128) { sk[i] = mk[off + i]; sA[i] = thrA[off + i]; sB[i] = thrB[off + i]; corr[i] = 0; }
  __syncthreads();
  const float dthr = sB[P - 1];
  const float m = ordered_int2f(rowmax_key[own_row(q0, q_stride, qi)]);
  const float c1h = __half2float(__float2half_rn(c1));
  // every special entry was binned by the sweep as if the Jaccard term did not touch it: take that out, and put the
  // entry back where its true final key belongs (junk -- same identity and camera -- stays out: evaluate.py:55-58)
  auto fix = [&](float d, int col, unsigned long long key, bool junk) {
    if (d <= dthr) {
      const int b = rr_zero_bin(d, col, sA, sB, sk, P);
      if (b < P) atomicAdd(&corr[b], -1);
    }
    if (!junk) {
      const int lb = lower_bound_u64(sk, P, key);
      if (lb < P) atomicAdd(&corr[lb], 1);
    }
  };
  for (int e = tid; e < len; e += 128) fix(gd[off + e], list[lo + e], gkey[off + e], gjunk[off + e] != 0);
  const int no = ocnt[qi];
  const long long qp = q_pid[qi];
  const long long ob = (long long)qi * cap_o;
  for (int x = tid; x < no; x += 128) {
    const int col = ocol[ob + x];
    if (g_pid[col] == qp) continue;                       // in the identity group: handled above
    const float d = od[ob + x];
    fix(d, col, final_key(final_value(ot[ob + x], d, m, c1h, c2), col), false);
  }
  __syncthreads();
  if (tid == 0) {
    long long run = 0;
    double acc = 0.0;
    int first = 0;
    for (int i = 0; i < P; ++i) {
      run += (long long)bins[off + i] + corr[i];           // 1-based kept rank of match i
      if (i == 0) first = (int)(run - 1);
      acc += (double)(i + 1) / (double)run;
    }
    ap[qi] = acc / (double)P;
    first_hit[qi] = first;
    if (first < max_rank) atomicAdd(cmc_hist + first, 1);
    atomicAdd(n_valid, 1);
  }
}

// ---------------------------------------------------------------------------------------------
// exact top-K of the re-ranked row
// ---------------------------------------------------------------------------------------------
constexpr int RT_CAND = 8192;      // candidate keys per query in shared memory (64 KB)
__global__ void __launch_bounds__(256)
rr_topk_kernel(int nq, int q0, int q_stride, const int* __restrict__ tidx, const float* __restrict__ tval, int Kd, int g,
               int cap_o, const int* __restrict__ ocol, const __half* __restrict__ ot, const float* __restrict__ od,
               const int* __restrict__ ocnt, const int* __restrict__ rowmax_key, float c1, float c2, int K,
               int* __restrict__ out_idx, float* __restrict__ out_val, int* __restrict__ ambiguous) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(rt_smem);
  __shared__ int s_n;
  const int qi = blockIdx.x, tid = threadIdx.x;
  if (tid == 0) s_n = 0;
  __syncthreads();
  const int no = max(ocnt[qi], 0);
  const long long ob = (long long)qi * cap_o;
  const float m = ordered_int2f(rowmax_key[own_row(q0, q_stride, qi)]);
  const float c1h = __half2float(__float2half_rn(c1));
  // J <= 1 only lowers a final value, so the K nearest by raw distance all end at or below final0(D_K): nothing above
  // that bound can be among the K best of the re-ranked row -- the tens of thousands of overlap targets are filtered by it
  const float cut = final_value(__ushort_as_half(0), tval[(long long)qi * Kd + K - 1], m, c1h, c2);
  for (int x = tid; x < Kd; x += 256) {                   // nearest by raw distance: untouched ones rank by final0(D)
    const int col = tidx[(long long)qi * Kd + x];
    if (rr_find(ocol + ob, no, col) >= 0) continue;       // touched: enters below with its true value
    keys[atomicAdd(&s_n, 1)] = final_key(final_value(__ushort_as_half(0), tval[(long long)qi * Kd + x], m, c1h, c2), col);
  }
  for (int x = tid; x < no; x += 256) {
    const float f = final_value(ot[ob + x], od[ob + x], m, c1h, c2);
    if (!(f > cut)) {
      const int at = atomicAdd(&s_n, 1);
      if (at < RT_CAND) keys[at] = final_key(f, ocol[ob + x]);
    }
  }
  __syncthreads();
  const int n = min(s_n, RT_CAND);
  int P2 = 32;
  while (P2 < n) P2 <<= 1;
  for (int i = n + tid; i < P2; i += 256) keys[i] = ~0ull;
  __syncthreads();
  sr_block_bitonic<256>(keys, P2);
  for (int i = tid; i < K; i += 256) {
    out_idx[(long long)qi * K + i] = (int)(unsigned)(keys[i] & 0xffffffffu);
    out_val[(long long)qi * K + i] = key2f((uint32_t)(keys[i] >> 32));
  }
  // an untouched entry beyond the raw-distance list has final0 >= final0(last listed D): the first K are exact unless
  // the K-th value reaches that bound (or the candidate buffer overflowed)
  if (tid == 0) {
    bool amb = s_n > RT_CAND;
    if (Kd < g) {
      const float bound = final_value(__ushort_as_half(0), tval[(long long)qi * Kd + Kd - 1], m, c1h, c2);
      amb |= !(key2f((uint32_t)(keys[K - 1] >> 32)) < bound);
    }
    if (amb) atomicAdd(ambiguous, 1);
  }
}

// ---------------------------------------------------------------------------------------------
int launch_rr_prepare(int nq, int q0, int q_stride, const int* glo, const int* glen, const long long* goff,
                      const int* list, const float* gd, const long long* q_cam, const long long* g_cam,
                      const int* rowmax_key, float c1, float c2, int cap_o, const int* ocol, const __half* ot,
                      const int* ocnt, unsigned long long* gkey, unsigned char* gjunk, unsigned long long* mk,
                      float* thrA, float* thrB, int* nmatch, int* overflow, int group_max, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  REID_REQUIRE(group_max <= RR_CAP_MAX, "rr_prepare: an identity has %d gallery samples; at most %d per query", group_max,
               RR_CAP_MAX);
  const int cap = stream_group_cap(group_max);
  const size_t smem = (size_t)cap * 8;
  REID_CUDA(cudaFuncSetAttribute(rr_prepare_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  rr_prepare_kernel<<<nq, 256, smem, s>>>(nq, q0, q_stride, glo, glen, goff, list, gd, q_cam, g_cam, rowmax_key, c1, c2,
                                         cap_o, ocol, ot, ocnt, gkey, gjunk, mk, thrA, thrB, nmatch, overflow, cap);
  return check_launch("rr_prepare_kernel");
}

int launch_count_panel_rr(const float* panel, long long ld, int rows, int c0, int c1, const long long* goff,
                          const unsigned long long* mk, const float* thrA, const float* thrB, const int* nmatch,
                          int* bins, int group_max, cudaStream_t s) {
  if (rows == 0 || c1 <= c0) return REID_OK;
  REID_REQUIRE(ld % 4 == 0 && ((uintptr_t)panel & 15) == 0, "count_panel_rr: panel rows must be 16-byte aligned");
  REID_REQUIRE(group_max <= RR_CAP_MAX, "count_panel_rr: group_max=%d exceeds %d", group_max, RR_CAP_MAX);
  const int cap = stream_group_cap(group_max);
  const size_t smem = (size_t)cap * 20;
  REID_CUDA(cudaFuncSetAttribute(count_panel_rr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  count_panel_rr_kernel<<<rows, 256, smem, s>>>(panel, ld, rows, c0, c1, goff, mk, thrA, thrB, nmatch, bins, cap);
  return check_launch("count_panel_rr_kernel");
}

int launch_score_counts_rr(int nq, int q0, int q_stride, const int* glo, const int* glen, const long long* goff,
                           const int* list, const float* gd, const unsigned long long* gkey, const unsigned char* gjunk,
                           const unsigned long long* mk, const float* thrA, const float* thrB, const int* nmatch,
                           const int* bins, int cap_o, const int* ocol, const __half* ot, const float* od,
                           const int* ocnt, const long long* q_pid, const long long* g_pid, const int* rowmax_key,
                           float c1, float c2, int max_rank, double* ap, int* first_hit, int* cmc_hist, int* n_valid,
                           int group_max, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  REID_REQUIRE(group_max <= RR_CAP_MAX, "score_counts_rr: group_max=%d exceeds %d", group_max, RR_CAP_MAX);
  const int cap = stream_group_cap(group_max);
  const size_t smem = (size_t)cap * 20;
  REID_CUDA(cudaFuncSetAttribute(score_counts_rr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  score_counts_rr_kernel<<<nq, 128, smem, s>>>(nq, q0, q_stride, glo, glen, goff, list, gd, gkey, gjunk, mk, thrA, thrB,
                                              nmatch, bins, cap_o, ocol, ot, od, ocnt, q_pid, g_pid, rowmax_key, c1, c2,
                                              max_rank, ap, first_hit, cmc_hist, n_valid, cap);
  return check_launch("score_counts_rr_kernel");
}

int launch_rr_topk(int nq, int q0, int q_stride, const int* tidx, const float* tval, int Kd, int g, int cap_o,
                   const int* ocol, const __half* ot, const float* od, const int* ocnt, const int* rowmax_key, float c1,
                   float c2, int K, int* out_idx, float* out_val, int* ambiguous, cudaStream_t s) {
  if (nq == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= Kd && Kd <= g, "rr_topk: K=%d Kd=%d g=%d", K, Kd, g);
  REID_REQUIRE(Kd <= RT_CAND / 2, "rr_topk: Kd=%d too large", Kd);
  const size_t smem = (size_t)RT_CAND * 8;
  REID_CUDA(cudaFuncSetAttribute(rr_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  rr_topk_kernel<<<nq, 256, smem, s>>>(nq, q0, q_stride, tidx, tval, Kd, g, cap_o, ocol, ot, od, ocnt, rowmax_key, c1, c2,
                                      K, out_idx, out_val, ambiguous);
  return check_launch("rr_topk_kernel");
}

}  // namespace reid
