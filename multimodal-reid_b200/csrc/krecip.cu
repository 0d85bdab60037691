// K6 (reranking.py:51-71): k-reciprocal neighbour sets, the half-k expansion with the 2/3 rule,
// and the Gaussian-weighted sparse rows of V.
//
// The reference recomputes the half-depth reciprocal set of every candidate inside the loop over
// probes; that set depends on the candidate only, so it is built once per row here
// (reciprocal_sets_kernel, both depths in one sweep) and the per-probe kernel
// (expand_weights_kernel) only intersects and merges.  One warp per row; set algebra with
// ballots/shuffles; the neighbour table (N x K int32) is L2-resident.
#include "common.cuh"
#include "np_exact.cuh"
#include "kernels.h"

namespace reid {

// R_depth(i) = { nbr[i][r] : r < depth, i in nbr[nbr[i][r]][0:depth] }, kept in list order.
// One warp per row; the (k1+1)^2 (neighbour r, position j) probes of the neighbours' lists are independent loads
// spread over the lanes (a serial walk over r paid one L2 round trip per neighbour), hits are collected as bit masks
// and compacted in list order.
constexpr int RS_WPB = 8;
constexpr int RS_MAXK = 128;           // k1 + 1 <= 128
__global__ void __launch_bounds__(RS_WPB * 32)
reciprocal_sets_kernel(const int* __restrict__ nbr, int N, int K, int k1p, int h,
                       int* __restrict__ rfull, int* __restrict__ nfull,
                       int* __restrict__ rhalf, int* __restrict__ nhalf) {
  __shared__ unsigned bits[RS_WPB][2][RS_MAXK / 32];
  const int w = threadIdx.x >> 5;
  const int row = blockIdx.x * RS_WPB + w;
  if (row >= N) return;
  const int lane = lane_id();
  const int* mine = nbr + (long long)row * K;
  if (lane < 2 * (RS_MAXK / 32)) (&bits[w][0][0])[lane] = 0;
  __syncwarp();
  const int npairs = k1p * k1p;
  for (int p = lane; p < npairs; p += 32) {
    const int r = p / k1p, j = p - r * k1p;
    const int c = mine[r];
    if (nbr[(long long)c * K + j] == row) {
      atomicOr(&bits[w][0][r >> 5], 1u << (r & 31));
      if (j < h && r < h) atomicOr(&bits[w][1][r >> 5], 1u << (r & 31));
    }
  }
  __syncwarp();
  int cf = 0, ch = 0;
  for (int r0 = 0; r0 < k1p; r0 += 32) {
    const int r = r0 + lane;
    const bool f = r < k1p && ((bits[w][0][r0 >> 5] >> lane) & 1u);
    const bool hb = r < k1p && ((bits[w][1][r0 >> 5] >> lane) & 1u);
    const unsigned bf = __ballot_sync(FULL, f), bh = __ballot_sync(FULL, hb);
    const unsigned below = (1u << lane) - 1u;
    if (f) rfull[(long long)row * k1p + cf + __popc(bf & below)] = mine[r];
    if (hb) rhalf[(long long)row * h + ch + __popc(bh & below)] = mine[r];
    cf += __popc(bf);
    ch += __popc(bh);
  }
  if (lane == 0) { nfull[row] = cf; nhalf[row] = ch; }
}

// warp-level bitonic sort of P ints in shared memory (P power of two >= 32)
__device__ __forceinline__ void warp_bitonic_i32(int* a, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        int x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncwarp();
    }
  }
}

// per-warp shared memory: R list [k1p] | E list [capP] | weights [cap1] | 3 x [k1p] candidate bookkeeping
__global__ void __launch_bounds__(128)
expand_weights_kernel(const int* __restrict__ rfull, const int* __restrict__ nfull,
                      const int* __restrict__ rhalf, const int* __restrict__ nhalf, int row0,
                      int row_stride, int rows, int k1p, int h, const float* __restrict__ D, long long ld,
                      const int* __restrict__ rowmax_key, int cap1, int capP,
                      int* __restrict__ Vc, __half* __restrict__ Vv, int* __restrict__ Vn, int upper_only,
                      const float* const* __restrict__ panels, int out_global, const int* __restrict__ wl_idx,
                      const float* __restrict__ wl_val, int Kw, const float* __restrict__ x32, long long ldx,
                      const float* __restrict__ sq, int* __restrict__ wl_miss,
                      const unsigned long long* __restrict__ mt_keys, const float* __restrict__ mt_vals, int mt_n,
                      unsigned long long* __restrict__ miss_out, int miss_cap) {
  extern __shared__ int smem_i[];
  const int wpb = blockDim.x >> 5;
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int lrow = blockIdx.x * wpb + w;          // local row: D / rowmax / V are indexed by it
  if (lrow >= rows) return;
  const int row = (int)own_row(row0, row_stride, lrow);       // global probe index (common.cuh: cyclic or snake blocks)
  const long long vrow = out_global ? row : lrow;              // where the V row is written
  const int per_warp = k1p + capP + cap1 + 3 * k1p;
  int* R = smem_i + (long long)w * per_warp;
  int* E = R + k1p;
  float* W = reinterpret_cast<float*>(E + capP);

  const int nR = nfull[row];
  for (int j = lane; j < nR; j += 32) {
    int v = rfull[(long long)row * k1p + j];
    R[j] = v;
    E[j] = v;
  }
  __syncwarp();
  // 2/3 rule for every candidate c = R[a] (reranking.py:60-67), all (candidate, half-list entry) pairs in parallel:
  // lanes take pairs p = a * h + j; an entry counts if it is in R; candidate a is accepted if 3 * |Rc & R| > 2 * |Rc|
  int* cntS = reinterpret_cast<int*>(W + cap1);       // [k1p] |Rc & R|
  int* ncS = cntS + k1p;                              // [k1p] |Rc|
  int* offS = ncS + k1p;                              // [k1p] where an accepted Rc goes in E, -1 if rejected
  for (int a = lane; a < nR; a += 32) { cntS[a] = 0; ncS[a] = nhalf[R[a]]; }
  __syncwarp();
  const int npairs = nR * h;
  for (int p = lane; p < npairs; p += 32) {
    const int a = p / h, j = p - a * h;
    if (j < ncS[a]) {
      const int v = rhalf[(long long)R[a] * h + j];
      bool member = false;
      for (int b = 0; b < nR; ++b) member |= (R[b] == v);
      if (member) atomicAdd(&cntS[a], 1);
    }
  }
  __syncwarp();
  int nE = nR;
  for (int a0 = 0; a0 < nR; a0 += 32) {
    const int a = a0 + lane;
    const bool acc = a < nR && 3 * cntS[a] > 2 * ncS[a];       // len(intersect) > 2/3 * len(candidate set)
    const int len = acc ? ncS[a] : 0;
    int incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += y;
    }
    if (a < nR) offS[a] = acc ? nE + incl - len : -1;
    nE += __shfl_sync(FULL, incl, 31);
  }
  __syncwarp();
  for (int p = lane; p < npairs; p += 32) {
    const int a = p / h, j = p - a * h;
    if (offS[a] >= 0 && j < ncS[a]) E[offS[a] + j] = rhalf[(long long)R[a] * h + j];
  }
  __syncwarp();
  // sort + unique (np.unique, reranking.py:69)
  int P = 32;
  while (P < nE) P <<= 1;
  for (int j = nE + lane; j < P; j += 32) E[j] = 0x7fffffff;
  __syncwarp();
  warp_bitonic_i32(E, P, lane);
  int nU = 0;
  for (int j0 = 0; j0 < nE; j0 += 32) {
    const int j = j0 + lane;
    int v = 0;
    bool keep = false;
    if (j < nE) {
      v = E[j];
      keep = (j == 0) || (E[j - 1] != v);
    }
    __syncwarp();
    const unsigned bal = __ballot_sync(FULL, keep);
    if (keep) E[nU + __popc(bal & ((1u << lane) - 1))] = v;
    nU += __popc(bal);
    __syncwarp();
  }
  // Gaussian weights on the normalised distance, numpy-exact exp and pairwise sum
  const float m = ordered_int2f(rowmax_key[lrow]);
  const float* drow = D ? D + (long long)lrow * ld : nullptr;
  for (int j = lane; j < nU; j += 32) {
    const int e = E[j];                       // upper_only: left of the row's diagonal block only the mirror image exists
    float dv;
    if (wl_idx) {
      // streamed form (no matrix anywhere): D[row, e] is looked up in the row's wide raw-distance neighbour list -- the
      // exact value the contraction produced.  E_i consists of reciprocal neighbours of reciprocal neighbours, so a miss
      // is rare.  Misses are looked up in a second table (mt_keys = row << 32 | column, sorted): the caller runs this
      // kernel once to collect them (miss_out), contracts exactly those pairs on the tensor cores, and runs it again.
      // A pair in neither place is recomputed from the float32 features (~1e-6 off the tensor-core value) and counted.
      const int* li = wl_idx + (long long)row * Kw;
      int hit = -1;
      for (int t = 0; t < Kw; ++t)
        if (li[t] == e) { hit = t; break; }
      int mhit = -1;
      const unsigned long long pk = ((unsigned long long)(unsigned)row << 32) | (unsigned)e;
      if (hit < 0 && mt_keys) {
        int lo = 0, hi = mt_n;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (mt_keys[mid] < pk) lo = mid + 1; else hi = mid;
        }
        if (lo < mt_n && mt_keys[lo] == pk) mhit = lo;
      }
      if (hit >= 0) {
        dv = wl_val[(long long)row * Kw + hit];
      } else if (mhit >= 0) {
        dv = mt_vals[mhit];
      } else {
        const float* a = x32 + (long long)row * ldx;
        const float* b = x32 + (long long)e * ldx;
        float acc = 0.0f;
        for (int k = 0; k < (int)ldx; ++k) acc = __fmaf_rn(a[k], b[k], acc);
        dv = __fmaf_rn(-2.0f, acc, __fadd_rn(sq[row], sq[e]));
        const int pos = atomicAdd(wl_miss, 1);
        if (miss_out && pos < miss_cap) miss_out[pos] = pk;
      }
    } else if (upper_only && e < (row & ~255)) {
      // D[e, row]: on one rank a row of the same matrix; in a sharded job it sits in the panel of the rank that owns e's
      // tile row -- read over NVLink (a few dozen 4-byte loads per row; bit-identical: the element was computed once)
      const float* src = panels ? panels[snake_owner(e >> 8, -row_stride)] + (long long)snake_local_row(e, -row_stride) * ld
                                : D + (long long)e * ld;
      dv = src[row];
    } else {
      dv = drow[e];
    }
    float o = __fdiv_rn(dv, m);
    W[j] = np_expf(-o);
  }
  __syncwarp();
  const float s = np_sum_f32(W, nU);
  for (int j = lane; j < nU; j += 32) {
    Vc[vrow * cap1 + j] = E[j];
    Vv[vrow * cap1 + j] = __float2half_rn(__fdiv_rn(W[j], s));
  }
  if (lane == 0) Vn[vrow] = nU;
}

int launch_reciprocal_sets(const int* nbr, int N, int K, int k1, int* rfull, int* nfull, int* rhalf,
                           int* nhalf, cudaStream_t s) {
  if (N == 0) return REID_OK;
  const int k1p = k1 + 1, h = half_k(k1);
  REID_REQUIRE(K >= k1p, "reciprocal_sets: neighbour table width %d < k1+1=%d", K, k1p);
  REID_REQUIRE(k1p <= RS_MAXK, "reciprocal_sets: k1+1=%d exceeds %d", k1p, RS_MAXK);
  const int wpb = RS_WPB;
  reciprocal_sets_kernel<<<cdiv(N, wpb), wpb * 32, 0, s>>>(nbr, N, K, k1p, h, rfull, nfull, rhalf,
                                                         nhalf);
  return check_launch("reciprocal_sets_kernel");
}

int launch_expand_weights(const int* rfull, const int* nfull, const int* rhalf, const int* nhalf,
                          int row0, int row_stride, int rows, int k1, const float* D, long long ld,
                          const int* rowmax_key, int* Vc, __half* Vv, int* Vn, int upper_only,
                          const float* const* panels, int out_global, cudaStream_t s) {
  return launch_expand_weights_wl(rfull, nfull, rhalf, nhalf, row0, row_stride, rows, k1, D, ld, rowmax_key, Vc, Vv, Vn,
                                  upper_only, panels, out_global, nullptr, nullptr, 0, nullptr, 0, nullptr, nullptr,
                                  nullptr, nullptr, 0, nullptr, 0, s);
}

int launch_expand_weights_wl(const int* rfull, const int* nfull, const int* rhalf, const int* nhalf,
                             int row0, int row_stride, int rows, int k1, const float* D, long long ld,
                             const int* rowmax_key, int* Vc, __half* Vv, int* Vn, int upper_only,
                             const float* const* panels, int out_global, const int* wl_idx, const float* wl_val, int Kw,
                             const float* x32, long long ldx, const float* sq, int* wl_miss,
                             const unsigned long long* mt_keys, const float* mt_vals, int mt_n,
                             unsigned long long* miss_out, int miss_cap, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(D != nullptr || wl_idx != nullptr, "expand_weights: neither a distance matrix nor wide lists");
  REID_REQUIRE(!upper_only || (row0 == 0 && row_stride == 1) || (row_stride < 0 && panels),
               "expand_weights: upper_only needs all rows on one rank, or the sharded form with the peers' panels");
  const int k1p = k1 + 1, h = half_k(k1);
  const int cap1 = v_row_capacity(k1);
  int capP = 32;
  while (capP < cap1) capP <<= 1;
  const int per_warp = (k1p + capP + cap1 + 3 * k1p) * 4;
  int wpb = 4;
  while (wpb > 1 && per_warp * wpb > 200 * 1024) wpb >>= 1;
  REID_REQUIRE(per_warp * wpb <= 227 * 1024, "expand_weights: k1=%d needs %d B of shared memory", k1,
               per_warp);
  const int smem = per_warp * wpb;
  REID_CUDA(cudaFuncSetAttribute(expand_weights_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  expand_weights_kernel<<<cdiv(rows, wpb), wpb * 32, smem, s>>>(rfull, nfull, rhalf, nhalf, row0,
                                                               row_stride, rows, k1p, h, D, ld,
                                                               rowmax_key, cap1, capP, Vc, Vv, Vn, upper_only, panels,
                                                               out_global, wl_idx, wl_val, Kw, x32, ldx, sq, wl_miss,
                                                               mt_keys, mt_vals, mt_n, miss_out, miss_cap);
  return check_launch("expand_weights_kernel");
}

}  // namespace reid
