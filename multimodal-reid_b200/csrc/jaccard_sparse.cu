// K9 + K10 (+ K11 + K12), sparse formulation (reranking.py:84-100, evaluate.py:40-88).
//
// A query overlaps only a few thousand of the N probes (k1 = 20: ~2-5 k of 20-94 k), so the overlap row t[0:N]
// of jaccard_score.cu -- 2 N bytes of shared memory per query, one CTA per SM at MSMT17 scale -- is kept here as
//     bitmap[N bits]  +  prefix popcounts per 32-bit word  +  the non-zero values, packed
// with slot(g) = prefix[g >> 5] + popc(bitmap[g >> 5] below bit g).  That is small enough for many queries per
// SM, which is what the ordered float16 accumulation needs: it is one long dependent chain per query
// (~90 instructions per column, ~60-130 columns) that only other queries can hide.
//
//   overlap_sparse_kernel   four warps per query.  Pass A marks the targets of all the query's column lists in the
//                           bitmap (order-free), prefix-sums it, pass B walks the columns in ascending order
//                           (cp.async ring, one CTA barrier per column) adding min(V_qe[i,c], V_qe[g,c]) into the packed
//                           values with the reference's fp16 rounding after every add.  The record goes to global
//                           memory.  Queries touching more than `cap` targets are put on an overflow list and
//                           handled by the dense kernel (jaccard_score.cu).
//   score_sparse_kernel     one CTA per query: streams the query's row of D once; the bitmap byte of 8 consecutive
//                           entries says which of them overlap.  Same counting scheme as the dense kernel
//                           (sorted match keys, raw-D thresholds for non-overlapping entries, queues for the rare
//                           cases); optionally writes the mixed row.
#include "jaccard_common.cuh"
#include "kernels.h"

namespace reid {

constexpr int OS_RING = 4;      // column lists in flight per warp
constexpr int OS_SLOT = 128;    // ring slot capacity in entries; longer lists read their tail directly
constexpr int OS_CHUNK = 64;    // query columns staged per chunk

struct SparseRec {              // byte offsets inside one query's record
  size_t prefix, vals, bytes;
  int nw;
};
static inline __host__ __device__ SparseRec sparse_rec(int N, int cap) {
  SparseRec r;
  r.nw = (N + 31) >> 5;
  r.prefix = ((size_t)r.nw * 4 + 15) & ~(size_t)15;
  r.vals = r.prefix + (((size_t)r.nw * 2 + 15) & ~(size_t)15);
  r.bytes = r.vals + (((size_t)cap * 2 + 15) & ~(size_t)15);
  return r;
}

constexpr int OS_WARPS = 4;     // warps per query: entries of one column list hit distinct targets, so they run in parallel
constexpr int OS_T = 32 * OS_WARPS;

__global__ void __launch_bounds__(OS_T)
overlap_sparse_kernel(const int* __restrict__ Qc, const __half* __restrict__ Qv,
                      const long long* __restrict__ qoff, const int* __restrict__ Qn,
                      const long long* __restrict__ colptr, const unsigned long long* __restrict__ csc_ent,
                      int q0, int q_stride, int N, int cap, unsigned char* __restrict__ recs,
                      int* __restrict__ rec_count, int* __restrict__ over_list, int* __restrict__ over_n) {
  extern __shared__ __align__(16) unsigned char smem_o[];
  const SparseRec R = sparse_rec(N, cap);
  unsigned* bm = reinterpret_cast<unsigned*>(smem_o);
  unsigned short* pre = reinterpret_cast<unsigned short*>(smem_o + R.prefix);
  __half* vals = reinterpret_cast<__half*>(smem_o + R.vals);
  unsigned long long* ring = reinterpret_cast<unsigned long long*>(smem_o + R.bytes);
  long long* cstart = reinterpret_cast<long long*>(ring + OS_RING * OS_SLOT);
  int* clen = reinterpret_cast<int*>(cstart + OS_CHUNK);
  float* cval = reinterpret_cast<float*>(clen + OS_CHUNK);
  __shared__ int s_wtot[OS_WARPS];

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int qi = blockIdx.x;
  const long long row = own_row(q0, q_stride, qi);
  const long long o = qoff[row];
  const int nq = Qn[row];
  const int nw = R.nw;

  for (int i = tid; i < nw; i += OS_T) bm[i] = 0;

  auto stage = [&](int base, int nc) {
    __syncthreads();
    for (int i = tid; i < nc; i += OS_T) {
      const int c = Qc[o + base + i];
      const long long a = colptr[c];
      cstart[i] = a;
      clen[i] = (int)(colptr[c + 1] - a);
      cval[i] = __half2float(Qv[o + base + i]);
    }
    __syncthreads();
  };

  // ---------------------------------------------------------------- pass A: which targets does the query touch
  // (order-free: warp w takes the column lists w, w + OS_WARPS, ... four at a time per lane)
  for (int base = 0; base < nq; base += OS_CHUNK) {
    const int nc = min(OS_CHUNK, nq - base);
    stage(base, nc);
    for (int e0 = 4 * w; e0 < nc; e0 += 4 * OS_WARPS) {
      int g[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + u;
        const int len = e < nc ? clen[e] : 0;
        const unsigned long long* src = csc_ent + (e < nc ? cstart[e] : 0);
#pragma unroll
        for (int k = 0; k < 3; ++k) g[u][k] = lane + 32 * k < len ? (int)(unsigned)(src[lane + 32 * k] & 0xffffffffu) : -1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (g[u][k] >= 0) atomicOr(bm + (g[u][k] >> 5), 1u << (g[u][k] & 31));
        const int e = e0 + u;
        if (e < nc && clen[e] > 96) {
          const unsigned long long* src = csc_ent + cstart[e];
          for (int p = 96 + lane; p < clen[e]; p += 32) {
            const int gg = (int)(unsigned)(src[p] & 0xffffffffu);
            atomicOr(bm + (gg >> 5), 1u << (gg & 31));
          }
        }
      }
    }
  }
  __syncthreads();
  // exclusive prefix popcounts per word: every warp scans a contiguous quarter, then the quarters are offset
  const int per_w = (nw + OS_WARPS - 1) / OS_WARPS;
  const int wbeg = min(nw, w * per_w), wend = min(nw, wbeg + per_w);
  int run = 0;
  for (int w0 = wbeg; w0 < wend; w0 += 32) {
    const int i = w0 + lane;
    const int c = i < wend ? __popc(bm[i]) : 0;
    int incl = c;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int y = __shfl_up_sync(FULL, incl, off);
      if (lane >= off) incl += y;
    }
    if (i < wend) pre[i] = (unsigned short)min(run + incl - c, 65535);          // offset added below
    run += __shfl_sync(FULL, incl, 31);
  }
  if (lane == 0) s_wtot[w] = run;
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int k = 0; k < OS_WARPS; ++k) {
    if (k < w) before += s_wtot[k];
    total += s_wtot[k];
  }
  if (total > cap) {                                      // too dense for the packed form: the dense kernel takes it
    if (tid == 0) {
      rec_count[qi] = -1;
      over_list[atomicAdd(over_n, 1)] = qi;
    }
    return;
  }
  if (before)
    for (int i = wbeg + lane; i < wend; i += 32) pre[i] = (unsigned short)(pre[i] + before);
  for (int i = tid; i < total; i += OS_T) vals[i] = __ushort_as_half(0);

  // ---------------------------------------------------------------- pass B: ordered fp16 accumulation
  // columns one after the other (the reference's rounding sequence per target), the entries of a column across the CTA
  auto issue = [&](int c) {
    const int len = min(clen[c], OS_SLOT);
    const unsigned long long* src = csc_ent + cstart[c];
    unsigned long long* dst = ring + (c % OS_RING) * OS_SLOT;
    for (int p = tid; p < len; p += OS_T) cp_async8(dst + p, src + p);
  };
  auto add = [&](unsigned long long ent, float vi) {
    const unsigned g = (unsigned)(ent & 0xffffffffu);
    const unsigned wd = bm[g >> 5];
    const int slot = (int)pre[g >> 5] + __popc(wd & ((1u << (g & 31)) - 1));
    const float vg = __half2float(__ushort_as_half((unsigned short)(ent >> 32)));
    vals[slot] = __float2half_rn(__fadd_rn(__half2float(vals[slot]), fminf(vi, vg)));
  };
  static_assert(OS_SLOT == OS_T, "one ring entry per thread");
  for (int base = 0; base < nq; base += OS_CHUNK) {
    const int nc = min(OS_CHUNK, nq - base);
    stage(base, nc);                                               // also orders the prefix / zeroing above
#pragma unroll
    for (int s2 = 0; s2 < OS_RING - 1; ++s2) {
      if (s2 < nc) issue(s2);
      cp_async_commit();
    }
    for (int e = 0; e < nc; ++e) {
      cp_async_wait<OS_RING - 2>();                                // this thread's part of column e has landed
      __syncthreads();                                             // column e staged by all threads; column e-1 applied
      if (e + OS_RING - 1 < nc) issue(e + OS_RING - 1);            // into the slot column e-1 has just left
      cp_async_commit();
      const float vi = cval[e];
      const int len = clen[e];
      const unsigned long long* slotp = ring + (e % OS_RING) * OS_SLOT;
      if (tid < len) add(slotp[tid], vi);                          // len <= OS_SLOT entries: one per thread
      if (len > OS_SLOT) {
        const unsigned long long* src = csc_ent + cstart[e];
        for (int p = OS_SLOT + tid; p < len; p += OS_T) add(src[p], vi);
      }
    }
    cp_async_wait<0>();
  }
  __syncthreads();

  // ---------------------------------------------------------------- the record
  unsigned char* rec = recs + (size_t)qi * R.bytes;
  {
    uint4* dst = reinterpret_cast<uint4*>(rec);
    const uint4* src = reinterpret_cast<const uint4*>(smem_o);
    const size_t n16 = (R.vals + (((size_t)total * 2 + 15) & ~(size_t)15)) >> 4;
    for (size_t i = tid; i < n16; i += OS_T) dst[i] = src[i];
  }
  if (tid == 0) rec_count[qi] = total;
}

// ---------------------------------------------------------------------------------------------
// streaming mix + score from the sparse record
// ---------------------------------------------------------------------------------------------
constexpr int SS_T = 256;
constexpr int SS_CAP = 512;     // identity-group entries staged / matches ranked per pass over the row
constexpr int SS_QSLOTS = 64;   // per-warp queue capacity: untouched entries before the second-worst match (8 B)
constexpr int SS_QA = 128;      // per-warp queue capacity: overlapping entries (4 B; shares SS_QSLOTS * 8 bytes)
constexpr int SS_NB = 2048;     // cells of the raw-D histogram of untouched entries below the worst match's threshold

struct SsLayout {               // after bitmap + prefix
  static constexpr size_t mk = 0;                                      // u64[SS_CAP]
  static constexpr size_t gcol = mk + SS_CAP * 8;                      // int[SS_CAP]   column | junk << 31
  static constexpr size_t gd = gcol + SS_CAP * 4;                      // float[SS_CAP]
  static constexpr size_t hv = gd + SS_CAP * 4;                        // int[SS_CAP + 1]
  static constexpr size_t hp = hv + (SS_CAP + 1) * 4;                  // int[SS_CAP + 1]
  static constexpr size_t thrA = hp + (SS_CAP + 1) * 4;                // float[SS_CAP]
  static constexpr size_t thrB = thrA + SS_CAP * 4;                    // float[SS_CAP]
  static constexpr size_t queue = (thrB + SS_CAP * 4 + 15) & ~(size_t)15;
  static constexpr size_t hist = queue + (size_t)(SS_T / 32) * 2 * SS_QSLOTS * 8;   // int[SS_NB]
  static constexpr size_t bytes = hist + (size_t)SS_NB * 4;
};

// overlap value of probe g from the packed record (0 when the query does not touch g)
__device__ __forceinline__ __half sp_t(const unsigned* bm, const unsigned short* pre, const __half* __restrict__ vals,
                                       int g) {
  const unsigned wd = bm[g >> 5];
  const unsigned bit = 1u << (g & 31);
  if (!(wd & bit)) return __ushort_as_half(0);
  return vals[(int)pre[g >> 5] + __popc(wd & (bit - 1))];
}

// Overlap queue entries are probe indices; the raw distance is re-read (L1/L2: the line was streamed a moment ago)
// alongside the packed overlap value, so the two dependent-load latencies overlap.
static __device__ __noinline__ void sp_drain_overlap(const int* qa, int cnt, const unsigned* bm,
                                                     const unsigned short* pre, const __half* vals,
                                                     const float* __restrict__ drow, float m, float c1h, float c2, int q,
                                                     const unsigned long long* mk, int P, unsigned long long maxkey,
                                                     int* hv) {
  const int lane = threadIdx.x & 31;
  if (lane < cnt) {
    const int g = qa[lane];
    const float d = __ldg(drow + g);
    const float f = final_value(sp_t(bm, pre, vals, g), d, m, c1h, c2);
    const unsigned long long key = final_key(f, g - q);
    if (key <= maxkey) bin_add(hv, lower_bound_u64(mk, P, key));
  }
  __syncwarp();
}
// Untouched entries (overlap 0) rank by raw D alone.  Those at or below the worst match's threshold are counted in a
// histogram of SS_NB cells over [0, threshold]: cell(d) is monotone in d, so an entry in a cell that holds no match
// threshold lies on the same side of every threshold as the whole cell.  Cells that do hold one are marked (sign bit);
// their entries are ranked exactly by ss_zero_bin.
__device__ __forceinline__ int ss_cell(float d, float scale) {
  return (int)min(__float2uint_rz(__fmul_rn(d, scale)), (unsigned)(SS_NB - 1));     // negative, NaN -> 0
}
__device__ __forceinline__ int ss_zero_bin(float d, int col, const float* thrA, const float* thrB,
                                           const unsigned long long* mk, int P) {
  int lo = 0, hi = P;                                     // #{i : d > thrB[i]}; < P because d <= thrB[P-1]
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
__device__ __forceinline__ float ss_pick8(const float4& a, const float4& b, int e) {
  const float lo = (e & 2) ? ((e & 1) ? a.w : a.z) : ((e & 1) ? a.y : a.x);
  const float hi = (e & 2) ? ((e & 1) ? b.w : b.z) : ((e & 1) ? b.y : b.x);
  return (e & 4) ? hi : lo;
}

__global__ void __launch_bounds__(SS_T, 4)
score_sparse_kernel(const unsigned char* __restrict__ recs, const int* __restrict__ rec_count, int cap,
                    const float* __restrict__ D, long long ld, const int* __restrict__ rowmax_key, int q, int N,
                    float c1, float c2, float* __restrict__ out, long long ld_out, ScoreArgs sc) {
  constexpr int T = SS_T;
  const int qi = blockIdx.x;
  if (rec_count[qi] < 0) return;                          // on the overflow list: the dense kernel scores it
  extern __shared__ __align__(16) unsigned char smem_s[];
  const SparseRec R = sparse_rec(N, cap);
  unsigned* bm = reinterpret_cast<unsigned*>(smem_s);
  unsigned short* pre = reinterpret_cast<unsigned short*>(smem_s + R.prefix);
  unsigned char* aux = smem_s + R.vals;
  unsigned long long* mk = reinterpret_cast<unsigned long long*>(aux + SsLayout::mk);
  int* gcol = reinterpret_cast<int*>(aux + SsLayout::gcol);
  float* gd = reinterpret_cast<float*>(aux + SsLayout::gd);
  int* hv = reinterpret_cast<int*>(aux + SsLayout::hv);
  int* hp = reinterpret_cast<int*>(aux + SsLayout::hp);
  float* thrA = reinterpret_cast<float*>(aux + SsLayout::thrA);
  float* thrB = reinterpret_cast<float*>(aux + SsLayout::thrB);
  int* hist = reinterpret_cast<int*>(aux + SsLayout::hist);
  __shared__ int s_lo, s_len, s_total, s_cnt, s_first, s_zfast;

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  unsigned long long* wq = reinterpret_cast<unsigned long long*>(aux + SsLayout::queue) + w * 2 * SS_QSLOTS;
  const unsigned char* rec = recs + (size_t)qi * R.bytes;
  const __half* vals = reinterpret_cast<const __half*>(rec + R.vals);          // stays in global memory (L2)
  {
    const uint4* src = reinterpret_cast<const uint4*>(rec);
    uint4* dst = reinterpret_cast<uint4*>(smem_s);
    for (int i = tid; i < (int)(R.vals >> 4); i += T) dst[i] = __ldg(src + i);
  }
  const float m = ordered_int2f(rowmax_key[qi]);
  const float* drow = D + (long long)qi * ld;
  const float c1h = __half2float(__float2half_rn(c1));
  const bool scoring = sc.ap != nullptr;
  long long qc = 0;
  if (scoring) {
    qc = sc.q_cam[qi];
    if (tid == 0) {
      const int slot = pid_find(sc.keys, sc.T, sc.q_pid[qi]);
      int lo = 0, len = 0;
      if (slot >= 0) { lo = (int)sc.start[slot]; len = (int)(sc.start[slot + 1] - sc.start[slot]); }
      s_lo = lo; s_len = len; s_total = 0; s_first = 0x7fffffff;
    }
  }
  __syncthreads();

  auto tval = [&](int g) { return sp_t(bm, pre, vals, g); };
  // plain mix of the gallery columns (re_ranking()'s return value), coalesced scalar accesses
  auto mix_only = [&]() {
    float* orow = out + (long long)qi * ld_out;
    int g = q + tid;
    for (; g + 3 * T < N; g += 4 * T) {
      float dv[4];
      __half tv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { dv[u] = __ldg(drow + g + u * T); tv[u] = tval(g + u * T); }
#pragma unroll
      for (int u = 0; u < 4; ++u) orow[g + u * T - q] = final_value(tv[u], dv[u], m, c1h, c2);
    }
    for (; g < N; g += T) orow[g - q] = final_value(tval(g), __ldg(drow + g), m, c1h, c2);
  };
  if (!scoring) {
    if (out) mix_only();
    return;
  }
  const int* grp = sc.list + s_lo;      // gallery columns with the query's pid: each junk or a match
  const int L = s_len;
  {
    int matches = 0;
    for (int e = tid; e < L; e += T) {
      const int j = grp[e];
      const bool junk = sc.g_cam[j] == qc;
      matches += !junk;
      if (e < SS_CAP) { gcol[e] = j | (junk ? (int)0x80000000 : 0); gd[e] = __ldg(drow + q + j); }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) matches += __shfl_xor_sync(FULL, matches, off);
    if (lane == 0 && matches) atomicAdd(&s_total, matches);
  }
  __syncthreads();
  const int total = s_total;
  if (total == 0) {                     // query identity absent from the gallery: skipped (evaluate.py:61-63)
    if (tid == 0) { sc.ap[qi] = -1.0; sc.first_hit[qi] = -1; }
    if (out) mix_only();
    return;
  }

  double acc = 0.0;                     // sum over matches of (#matches up to p) / (kept rank of p); warp 0
  bool wrote = false;
  for (int e0 = 0; e0 < L; e0 += SS_CAP) {
    const int ns = min(SS_CAP, L - e0);
    // ---------------------------------------------------------------- this slice's matches, sorted
    if (e0 > 0) {
      for (int e = tid; e < ns; e += T) {
        const int j = grp[e0 + e];
        gcol[e] = j | (sc.g_cam[j] == qc ? (int)0x80000000 : 0);
        gd[e] = __ldg(drow + q + j);
      }
    }
    if (tid == 0) { s_cnt = 0; s_zfast = (m > 0.0f && m < INFINITY && c2 > 0.0f && c2 < INFINITY) ? 1 : 0; }
    for (int i = tid; i <= SS_CAP; i += T) { hv[i] = 0; hp[i] = 0; }
    __syncthreads();
    if (ns <= 32) {
      if (w == 0) {                     // the usual case: one warp sorts the keys in registers
        unsigned long long key = ~0ull;
        if (lane < ns) {
          const int j = gcol[lane];
          if (j >= 0) key = final_key(final_value(tval(q + j), gd[lane], m, c1h, c2), j);
        }
#pragma unroll
        for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
          for (int jj = k >> 1; jj > 0; jj >>= 1) {
            const unsigned long long other = __shfl_xor_sync(FULL, key, jj);
            const bool take_min = ((lane & k) == 0) == ((lane & jj) == 0);
            key = take_min ? (other < key ? other : key) : (other > key ? other : key);
          }
        }
        mk[lane] = key;
        const int cnt = __popc(__ballot_sync(FULL, key != ~0ull));
        if (lane == 0) s_cnt = cnt;
      }
      __syncthreads();
    } else {
      for (int e = tid; e < ns; e += T) {
        const int j = gcol[e];
        if (j >= 0) mk[atomicAdd(&s_cnt, 1)] = final_key(final_value(tval(q + j), gd[e], m, c1h, c2), j);
      }
      __syncthreads();
      const int Pn = s_cnt;
      int P2 = 32;
      while (P2 < Pn) P2 <<= 1;
      for (int i = Pn + tid; i < P2; i += T) mk[i] = ~0ull;
      __syncthreads();
      for (int k = 2; k <= P2; k <<= 1) {
        for (int jj = k >> 1; jj > 0; jj >>= 1) {
          for (int x = tid; x < (P2 >> 1); x += T) {
            const int i = ((x & ~(jj - 1)) << 1) | (x & (jj - 1));
            const int p2 = i | jj;
            const unsigned long long a = mk[i], b = mk[p2];
            const bool up = (i & k) == 0;
            if ((a > b) == up) { mk[i] = b; mk[p2] = a; }
          }
          __syncthreads();
        }
      }
    }
    const int P = s_cnt;
    if (P == 0) { __syncthreads(); continue; }
    // raw-D thresholds of the matches for non-overlapping entries (see jaccard_score.cu), one warp per threshold
    if (s_zfast) {
      for (int x = w; x < 2 * P; x += T / 32) {
        const int i = x >> 1;
        const float f = key2f((uint32_t)(mk[i] >> 32));
        if (f != f) { s_zfast = 0; continue; }
        const float thr = final0_threshold_warp(f, (x & 1) != 0, m, c1h, c2);
        if (lane == 0) { if (x & 1) thrA[i] = thr; else thrB[i] = thr; }
      }
    }
    __syncthreads();
    const bool zfast = s_zfast != 0;
    const unsigned long long maxkey = mk[P - 1];
    const bool use_hist = zfast && !(out && !wrote);
    float hscale = 0.0f;
    if (use_hist) {
      const float top = thrB[P - 1];
      if (top > 0.0f && top < INFINITY) hscale = __fdiv_rn((float)SS_NB, top);
      if (!(hscale < INFINITY)) hscale = 0.0f;                      // degenerate: one cell, every entry ranked exactly
      for (int x = tid; x < SS_NB; x += T) hist[x] = 0;
      __syncthreads();
      for (int i = tid; i < P; i += T) {                           // cells holding a threshold (or a tie range)
        const int hi = ss_cell(thrB[i], hscale);
        for (int x = ss_cell(thrA[i], hscale); x <= hi; ++x) hist[x] = (int)0x80000000;
      }
      __syncthreads();
    }

    // ---------------------------------------------------------------- one pass over the row
    if (out && !wrote) {
      float* orow = out + (long long)qi * ld_out;
      auto slow = [&](int g, float d) {
        const float f = final_value(tval(g), d, m, c1h, c2);
        const unsigned long long key = final_key(f, g - q);
        if (key <= maxkey) bin_add(hv, lower_bound_u64(mk, P, key));
        orow[g - q] = f;
      };
      int g = q + tid;
      for (; g + 3 * T < N; g += 4 * T) {
        float dv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) dv[u] = __ldg(drow + g + u * T);
#pragma unroll
        for (int u = 0; u < 4; ++u) slow(g + u * T, dv[u]);
      }
      for (; g < N; g += T) slow(g, __ldg(drow + g));
      wrote = true;
    } else {
      const float dthr = zfast ? thrB[P - 1] : -INFINITY;          // untouched and D > dthr: after every match
      int* qa = reinterpret_cast<int*>(wq);                        // SS_QA slots of 4 bytes
      int na = 0;
      const float4* d4 = reinterpret_cast<const float4*>(drow);
      auto ldg8 = [&](int g, float4& x0, float4& x1) {
        if (g < N) { x0 = __ldg(d4 + (g >> 2)); x1 = __ldg(d4 + (g >> 2) + 1); }
        else { x0 = make_float4(0.f, 0.f, 0.f, 0.f); x1 = x0; }
      };
      constexpr int STRIDE = T * 8;
      int gw = (q & ~7) + (tid & ~31) * 8;                        // warp-uniform base of this warp's group
      int gb = gw + lane * 8;                                     // this lane's 8 aligned entries
      float4 c0, c1v, n0, n1, f0, f1;                             // current, next, next-but-one
      ldg8(gb, c0, c1v);
      ldg8(gb + STRIDE, n0, n1);
      for (; gw < N; gw += STRIDE, gb += STRIDE) {
        ldg8(gb + 2 * STRIDE, f0, f1);
        // which of this lane's 8 entries are gallery columns, and which of those overlap
        unsigned vm = 0;
        if (gb < N) {
          vm = 0xffu;
          if (gb < q) vm &= 0xffu << (q - gb);                    // q - gb in [1, 7]: gb >= (q & ~7)
          if (gb + 8 > N) vm &= 0xffu >> (gb + 8 - N);
        }
        const unsigned tb = gb < N ? (bm[gb >> 5] >> (gb & 31)) & vm : 0u;
        const float dmin = fminf(fminf(fminf(c0.x, c0.y), fminf(c0.z, c0.w)), fminf(fminf(c1v.x, c1v.y), fminf(c1v.z, c1v.w)));
        unsigned ma = zfast ? tb : vm;
        unsigned mz = 0;
        if (zfast && vm != 0 && dmin <= dthr) {                   // something untouched ranks before the worst match
          const float dv[8] = {c0.x, c0.y, c0.z, c0.w, c1v.x, c1v.y, c1v.z, c1v.w};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            if ((((vm & ~tb) >> e) & 1u) && dv[e] <= dthr) {
              if (atomicAdd(hist + ss_cell(dv[e], hscale), 1) < 0) mz |= 1u << e;     // a cell that holds a threshold
            }
          }
        }
        if (__any_sync(FULL, ma != 0)) {
          const int mine = __popc(ma);
          int incl = mine;
#pragma unroll
          for (int off = 1; off < 32; off <<= 1) {
            const int y = __shfl_up_sync(FULL, incl, off);
            if (lane >= off) incl += y;
          }
          const int totA = __shfl_sync(FULL, incl, 31);
          if (na + totA > SS_QA && na > 0) {                      // make room: partial pass
            sp_drain_overlap(qa, na, bm, pre, vals, drow, m, c1h, c2, q, mk, P, maxkey, hv);
            na = 0;
          }
          if (totA <= SS_QA) {
            int pos = na + incl - mine;
            unsigned bits = ma;
            while (bits) {
              const int e = __ffs(bits) - 1;
              bits &= bits - 1;
              qa[pos++] = gb + e;
            }
            na += totA;
            __syncwarp();
            while (na >= 32) { na -= 32; sp_drain_overlap(qa + na, 32, bm, pre, vals, drow, m, c1h, c2, q, mk, P, maxkey, hv); }
          } else {                                                // dense stretch: one entry per lane at a time
#pragma unroll 1
            for (int e = 0; e < 8; ++e) {
              const bool on = (ma >> e) & 1u;
              const unsigned act = __ballot_sync(FULL, on);
              if (act) {
                if (on) qa[__popc(act & ((1u << lane) - 1))] = gb + e;
                __syncwarp();
                sp_drain_overlap(qa, __popc(act), bm, pre, vals, drow, m, c1h, c2, q, mk, P, maxkey, hv);
              }
            }
          }
        }
        while (mz) {                                              // divergent, rare: ranked against the thresholds
          const int e = __ffs(mz) - 1;
          mz &= mz - 1;
          const int bin = ss_zero_bin(ss_pick8(c0, c1v, e), gb + e - q, thrA, thrB, mk, P);
          if (bin < P) bin_add(hv, bin);
        }
        __syncwarp();
        c0 = n0; c1v = n1; n0 = f0; n1 = f1;
      }
      if (na > 0) sp_drain_overlap(qa, na, bm, pre, vals, drow, m, c1h, c2, q, mk, P, maxkey, hv);
    }
    __syncthreads();
    if (use_hist) {
      // fold the histogram into the match bins: a clean cell x belongs to bin #{i : cell(thrB[i]) < x}; every thread
      // walks SS_NB / T consecutive cells (bins are non-decreasing along them), then equal bins combine per warp
      constexpr int PER = SS_NB / T;
      int x = tid * PER;
      int lo = 0, hi = P;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (ss_cell(thrB[mid], hscale) < x) lo = mid + 1; else hi = mid;
      }
      int bin = lo, sum = 0;
      for (int k = 0; k < PER; ++k, ++x) {
        const int h = hist[x];
        if (h <= 0) continue;                                       // empty, or marked: its entries were ranked exactly
        while (bin < P && ss_cell(thrB[bin], hscale) < x) {
          if (sum) { atomicAdd(hv + bin, sum); sum = 0; }
          ++bin;
        }
        sum += h;
      }
      const unsigned peers = __match_any_sync(FULL, bin);
      const int tot = __reduce_add_sync(peers, sum);
      if (tot && lane == __ffs(peers) - 1 && bin < P) atomicAdd(hv + bin, tot);
    }
    // the whole identity group: junk leaves the kept bins, every match enters the match bins
    for (int e = tid; e < L; e += T) {
      int j;
      float d;
      if (e >= e0 && e < e0 + ns) { j = gcol[e - e0]; d = gd[e - e0]; }
      else { j = grp[e]; d = __ldg(drow + q + j); if (sc.g_cam[j] == qc) j |= (int)0x80000000; }
      const int col = j & 0x7fffffff;
      const unsigned long long key = final_key(final_value(tval(q + col), d, m, c1h, c2), col);
      if (key <= maxkey) {
        const int bin = lower_bound_u64(mk, P, key);
        if (j < 0) atomicSub(hv + bin, 1); else atomicAdd(hp + bin, 1);
      }
    }
    __syncthreads();
    // ranks of this slice's matches
    if (w == 0) {
      int carry_v = 0, carry_p = 0;
      for (int i0 = 0; i0 < P; i0 += 32) {
        const int i = i0 + lane;
        int xv = i < P ? hv[i] : 0, xp = i < P ? hp[i] : 0;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const int yv = __shfl_up_sync(FULL, xv, off), yp = __shfl_up_sync(FULL, xp, off);
          if (lane >= off) { xv += yv; xp += yp; }
        }
        xv += carry_v; xp += carry_p;
        if (i < P) {
          acc += (double)xp / (double)xv;       // xv = 1-based kept rank of match i, xp = matches up to it
          if (xp == 1) s_first = xv - 1;
        }
        carry_v = __shfl_sync(FULL, xv, 31);
        carry_p = __shfl_sync(FULL, xp, 31);
      }
    }
    __syncthreads();
  }
  if (out && !wrote) mix_only();
  if (w == 0) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(FULL, acc, off);
    if (lane == 0) {
      sc.ap[qi] = acc / (double)total;
      const int first = s_first;
      sc.first_hit[qi] = first;
      if (first < sc.max_rank) atomicAdd(sc.cmc_hist + first, 1);
      atomicAdd(sc.n_valid, 1);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
long long jaccard_sparse_ws_bytes(int nq, int N, int cap) {
  const SparseRec R = sparse_rec(N, cap);
  // records | rec_count[nq] | over_list[nq] | over_n[4]
  return (long long)R.bytes * nq + (long long)(2 * (long long)nq + 4) * 4 + 64;
}

bool jaccard_sparse_fits(int N, int cap) {
  const SparseRec R = sparse_rec(N, cap);
  const size_t k1 = R.bytes + (size_t)OS_RING * OS_SLOT * 8 + OS_CHUNK * 16;
  const size_t k2 = R.vals + SsLayout::bytes;
  return cap >= 64 && cap <= 65535 && k1 <= 200 * 1024 && k2 <= 200 * 1024;
}

int launch_jaccard_sparse(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                          const long long* colptr, const unsigned long long* csc_ent, const float* D, long long ld,
                          const int* rowmax_key, int q0, int q_stride, int nq_local, int q, int N, float c1, float c2,
                          float* out, long long ld_out, const ScoreArgs& sc, void* ws, int cap, int** over_list,
                          int** over_n, cudaStream_t s) {
  const SparseRec R = sparse_rec(N, cap);
  unsigned char* recs = static_cast<unsigned char*>(ws);
  int* rec_count = reinterpret_cast<int*>(recs + (((size_t)R.bytes * nq_local + 15) & ~(size_t)15));
  *over_list = rec_count + nq_local;
  *over_n = *over_list + nq_local;
  REID_CUDA(cudaMemsetAsync(*over_n, 0, 4, s));
  const size_t smem1 = R.bytes + (size_t)OS_RING * OS_SLOT * 8 + OS_CHUNK * 16;
  REID_CUDA(cudaFuncSetAttribute(overlap_sparse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
  overlap_sparse_kernel<<<nq_local, OS_T, smem1, s>>>(Qc, Qv, qoff, Qn, colptr, csc_ent, q0, q_stride, N, cap, recs,
                                                    rec_count, *over_list, *over_n);
  int rc = check_launch("overlap_sparse_kernel");
  if (rc) return rc;
  const size_t smem2 = R.vals + SsLayout::bytes;
  REID_CUDA(cudaFuncSetAttribute(score_sparse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  score_sparse_kernel<<<nq_local, SS_T, smem2, s>>>(recs, rec_count, cap, D, ld, rowmax_key, q, N, c1, c2, out, ld_out,
                                                   sc);
  return check_launch("score_sparse_kernel");
}

}  // namespace reid
