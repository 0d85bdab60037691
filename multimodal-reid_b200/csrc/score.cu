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


// ---------------------------------------------------------------------------------------------
// Fused K11 + K12 for the evaluator's hot path: CMC / mAP straight from the distance row, without
// materialising the ranking.  AP and the first-match rank only need, for every true match p of the
// query, (a) its 0-based position among the NON-junk gallery entries and (b) its position among the
// matches, both in the stable (distance, column) order evaluate.py:40 + a stable sort would give:
//     pos(p) = #{ valid j : (d_j, j) < (d_p, p) }
// The matches (a few dozen) are collected and sorted in shared memory; one streaming pass over the
// row bins every valid entry by binary search among them; prefix sums of the bins give the ranks.
// Reads: the distance row once (HBM) + the id vectors (L2).  Results are identical to
// rank_rows + score_rows.
// ---------------------------------------------------------------------------------------------
constexpr int SC_T = 256;
constexpr int SC_CAP = 1024;          // matches handled per pass over the row

__device__ __forceinline__ unsigned long long score_key(float d, int col) {
  return ((unsigned long long)f2key(__fadd_rn(d, 0.0f)) << 32) | (unsigned)col;
}

__global__ void __launch_bounds__(SC_T)
score_dist_kernel(const float* __restrict__ dist, int q, int g, long long ld,
                  const long long* __restrict__ q_pid, const long long* __restrict__ g_pid,
                  const long long* __restrict__ q_cam, const long long* __restrict__ g_cam, int max_rank,
                  double* __restrict__ ap, int* __restrict__ first_hit, int* __restrict__ cmc_hist,
                  int* __restrict__ n_valid) {
  extern __shared__ __align__(16) unsigned char smem_sc[];
  unsigned long long* mk = reinterpret_cast<unsigned long long*>(smem_sc);        // [SC_CAP] sorted match keys
  int* hv = reinterpret_cast<int*>(mk + SC_CAP);                                  // [SC_CAP + 1] kept-entry bins
  int* hp = hv + SC_CAP + 1;                                                      // [SC_CAP + 1] match bins
  unsigned* junk = reinterpret_cast<unsigned*>(hp + SC_CAP + 1);                  // [ceil(g/32)] junk bitmap
  unsigned* mbit = junk + ((g + 31) >> 5);                                        // [ceil(g/32)] match bitmap
  __shared__ int s_cnt, s_total, s_first;
  __shared__ int s_scan[SC_T / 32];

  const int qi = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const long long qp = q_pid[qi], qc = q_cam[qi];
  const float* drow = dist + (long long)qi * ld;
  const int gpad = (g + 31) & ~31;

  // phase 1: junk / match bitmaps and the number of matches
  if (tid == 0) { s_total = 0; s_first = 0x7fffffff; }
  __syncthreads();
  int my_matches = 0;
  for (int j = tid; j < gpad; j += SC_T) {
    bool isjunk = false, ismatch = false;
    if (j < g) {
      const bool same = g_pid[j] == qp;
      isjunk = same && (g_cam[j] == qc);
      ismatch = same && !isjunk;
    }
    const unsigned bj = __ballot_sync(FULL, isjunk), bm = __ballot_sync(FULL, ismatch);
    if (lane == 0) { junk[j >> 5] = bj; mbit[j >> 5] = bm; my_matches += __popc(bm); }
  }
  if (lane == 0 && my_matches) atomicAdd(&s_total, my_matches);
  __syncthreads();
  const int total = s_total;
  if (total == 0) {                                     // query identity absent: skipped (evaluate.py:61-63)
    if (tid == 0) { ap[qi] = -1.0; first_hit[qi] = -1; }
    return;
  }

  double acc = 0.0;          // sum over matches of (#matches up to p) / (kept rank of p); lives in warp 0
  int col_lo = 0;            // the matches are taken in column intervals of <= SC_CAP matches, one row pass each
  for (int done = 0; done < total;) {
    if (tid == 0) s_cnt = 0;
    for (int i = tid; i <= SC_CAP; i += SC_T) { hv[i] = 0; hp[i] = 0; }
    __syncthreads();
    int col_hi = g;
    for (int j0 = col_lo; j0 < g; j0 += SC_T) {
      const int j = j0 + tid;
      const bool ismatch = j < g && ((mbit[j >> 5] >> (j & 31)) & 1u);
      const unsigned bm = __ballot_sync(FULL, ismatch);
      if (lane == 0) s_scan[w] = __popc(bm);
      __syncthreads();
      const int cnt0 = s_cnt;
      int base = cnt0, tot = 0;
#pragma unroll
      for (int ww = 0; ww < SC_T / 32; ++ww) { if (ww < w) base += s_scan[ww]; tot += s_scan[ww]; }
      const bool fits = cnt0 + tot <= SC_CAP;
      if (fits && ismatch) mk[base + __popc(bm & ((1u << lane) - 1))] = score_key(drow[j], j);
      __syncthreads();
      if (!fits) { col_hi = j0; break; }
      if (tid == 0) s_cnt = cnt0 + tot;
    }
    __syncthreads();
    const int P = s_cnt;
    int P2 = 32;
    while (P2 < P) P2 <<= 1;
    for (int i = P + tid; i < P2; i += SC_T) mk[i] = ~0ull;
    __syncthreads();
    for (int k = 2; k <= P2; k <<= 1) {
      for (int jj = k >> 1; jj > 0; jj >>= 1) {
        for (int t = tid; t < (P2 >> 1); t += SC_T) {
          const int i = ((t & ~(jj - 1)) << 1) | (t & (jj - 1));
          const int p2 = i | jj;
          const unsigned long long x = mk[i], y = mk[p2];
          const bool up = (i & k) == 0;
          if ((x > y) == up) { mk[i] = y; mk[p2] = x; }
        }
        __syncthreads();
      }
    }
    // stream the row: bin every kept entry (and every match) by the number of chunk matches below it
    for (int j0 = 0; j0 < g; j0 += SC_T * 4) {
      const int j = j0 + tid * 4;
      if (j >= g) continue;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (j + 3 < g) {
        const float4 f = *reinterpret_cast<const float4*>(drow + j);
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
      } else {
        for (int e = 0; e < 4; ++e) if (j + e < g) v[e] = drow[j + e];
      }
      const unsigned jb = junk[j >> 5] >> (j & 31), mb = mbit[j >> 5] >> (j & 31);   // j % 4 == 0: same word
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (j + e >= g || ((jb >> e) & 1u)) continue;
        const unsigned long long key = score_key(v[e], j + e);
        int lo = 0, hi = P;                              // lo = #{chunk matches < key}
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (mk[mid] < key) lo = mid + 1; else hi = mid;
        }
        atomicAdd(&hv[lo], 1);
        if ((mb >> e) & 1u) atomicAdd(&hp[lo], 1);
      }
    }
    __syncthreads();
    // inclusive prefix sums over the P bins (one warp; P is small), then this chunk's AP terms
    if (w == 0) {
      int carry_v = 0, carry_p = 0;
      for (int i0 = 0; i0 < P; i0 += 32) {
        const int i = i0 + lane;
        int xv = i < P ? hv[i] : 0, xp = i < P ? hp[i] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int yv = __shfl_up_sync(FULL, xv, o), yp = __shfl_up_sync(FULL, xp, o);
          if (lane >= o) { xv += yv; xp += yp; }
        }
        xv += carry_v; xp += carry_p;
        if (i < P) {
          // kept entries up to and including match i = xv (its 1-based kept rank); matches likewise = xp
          acc += (double)xp / (double)xv;
          if (xp == 1) s_first = xv - 1;
        }
        carry_v = __shfl_sync(FULL, xv, 31);
        carry_p = __shfl_sync(FULL, xp, 31);
      }
    }
    __syncthreads();
    done += P;
    col_lo = col_hi;
  }
  if (w == 0) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
    if (lane == 0) {
      ap[qi] = acc / (double)total;
      const int first = s_first;
      first_hit[qi] = first;
      if (first < max_rank) atomicAdd(cmc_hist + first, 1);
      atomicAdd(n_valid, 1);
    }
  }
}

int launch_score_dist(const float* dist, int q, int g, long long ld, const long long* q_pid,
                      const long long* g_pid, const long long* q_cam, const long long* g_cam, int max_rank,
                      double* ap, int* first_hit, int* cmc_hist, int* n_valid, cudaStream_t s) {
  if (q == 0) return REID_OK;
  REID_REQUIRE(ld % 4 == 0, "score_dist: pitch %lld must be a multiple of 4", ld);
  const size_t smem = (size_t)SC_CAP * 8 + 2 * (size_t)(SC_CAP + 1) * 4 + 2 * (size_t)((g + 31) / 32) * 4 + 16;
  REID_REQUIRE(smem <= 200 * 1024, "score_dist: gallery of %d needs %zu B of shared memory", g, smem);
  REID_CUDA(cudaFuncSetAttribute(score_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  score_dist_kernel<<<q, SC_T, smem, s>>>(dist, q, g, ld, q_pid, g_pid, q_cam, g_cam, max_rank, ap, first_hit,
                                          cmc_hist, n_valid);
  return check_launch("score_dist_kernel");
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
