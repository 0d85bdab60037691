// Training-side distance ops, batch scale: m, n = 64 .. a few thousand rows, d <= 3072.
//   pair_dist_kernel    losses.py:14-28   dist = sqrt(clamp(|x|^2 + |y|^2 - 2 x.y, 1e-12)), norms fused (no prep pass)
//   hard_mine_kernel    losses.py:31-83   per anchor: farthest same-label column, nearest other-label column
//   mine_grad_kernel    autograd of the two selections: the gradient w.r.t. the distance matrix (two entries per row)
//   pair_grad_kernel    autograd of pair_dist: gx_i = sum_j W_ij (x_i - y_j), W = G / dist where the clamp is inactive
// These matrices are KBs to a few MB: every kernel here is latency-bound, so the design goal is few launches, enough
// CTAs to cover the SMs at B = 64 already, and fixed summation orders (no float atomics: results are run-to-run
// deterministic, which the reference's CUDA path is not).
#include "common.cuh"
#include "kernels.h"

namespace reid {

constexpr int PT = 32;     // output tile edge
constexpr int PK = 128;    // K slab: 16 + 16 independent loads per thread in flight, d / 128 dependent round trips

// 256 threads, 32 x 32 outputs, 2 x 2 per thread; rows/columns past the edge are staged as zeros.
__global__ void __launch_bounds__(256)
pair_dist_kernel(const float* __restrict__ X, long long ldx, int m, const float* __restrict__ Y, long long ldy, int n,
                 int d, float* __restrict__ out, long long ldo) {
  __shared__ float Xs[PK][PT + 1];
  __shared__ float Ys[PK][PT + 1];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * PT, n0 = blockIdx.x * PT;
  const int lr = tid >> 3, lk = (tid & 7) * 4;       // staging: row lr, 4 consecutive k in each 32-wide pass
  const int r_ld = m0 + lr, c_ld = n0 + lr;
  const float* xrow = X + (long long)min(r_ld, m - 1) * ldx;
  const float* yrow = Y + (long long)min(c_ld, n - 1) * ldy;
  float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
  float xx[2] = {0.f, 0.f}, yy[2] = {0.f, 0.f};
  for (int k0 = 0; k0 < d; k0 += PK) {
    float xv[PK / 8], yv[PK / 8];
#pragma unroll
    for (int p = 0; p < PK / 32; ++p)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k = k0 + p * 32 + lk + e;
        xv[p * 4 + e] = (r_ld < m && k < d) ? __ldg(xrow + k) : 0.f;
        yv[p * 4 + e] = (c_ld < n && k < d) ? __ldg(yrow + k) : 0.f;
      }
#pragma unroll
    for (int p = 0; p < PK / 32; ++p)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        Xs[p * 32 + lk + e][lr] = xv[p * 4 + e];
        Ys[p * 32 + lk + e][lr] = yv[p * 4 + e];
      }
    __syncthreads();
#pragma unroll 32
    for (int k = 0; k < PK; ++k) {
      const float a0 = Xs[k][ty], a1 = Xs[k][ty + 16], b0 = Ys[k][tx], b1 = Ys[k][tx + 16];
      acc[0][0] = __fmaf_rn(a0, b0, acc[0][0]);
      acc[0][1] = __fmaf_rn(a0, b1, acc[0][1]);
      acc[1][0] = __fmaf_rn(a1, b0, acc[1][0]);
      acc[1][1] = __fmaf_rn(a1, b1, acc[1][1]);
      xx[0] = __fmaf_rn(a0, a0, xx[0]);
      xx[1] = __fmaf_rn(a1, a1, xx[1]);
      yy[0] = __fmaf_rn(b0, b0, yy[0]);
      yy[1] = __fmaf_rn(b1, b1, yy[1]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int r = m0 + ty + 16 * i, c = n0 + tx + 16 * j;
      if (r < m && c < n) {
        // (xx + yy) - 2 * matmul: the doubling is exact, so one fma rounds like the reference's subtract
        const float v = __fmaf_rn(-2.0f, acc[i][j], __fadd_rn(xx[i], yy[j]));
        out[(long long)r * ldo + c] = __fsqrt_rn(fmaxf(v, 1e-12f));
      }
    }
}

// One warp per anchor row.  Positives: label equal (the anchor's own column included, as in the reference, unless
// skip_diag); negatives: label different.  Ties go to the lowest column (torch.max / torch.min on CPU keep the first).
__global__ void __launch_bounds__(256)
hard_mine_kernel(const float* __restrict__ D, long long ld, int m, int n, const int64_t* __restrict__ lab_r,
                 const int64_t* __restrict__ lab_c, int skip_diag, float* __restrict__ dist_ap,
                 float* __restrict__ dist_an, int64_t* __restrict__ p_ind, int64_t* __restrict__ n_ind) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= m) return;
  const int lane = lane_id();
  const int64_t me = lab_r[row];
  // keys: (ordered distance << 32) | tie-break, so that one 64-bit max / min decides value and column together
  unsigned long long best_p = 0ull, best_n = ~0ull;
  for (int c = lane; c < n; c += 32) {
    const float v = D[(long long)row * ld + c];
    const unsigned long long k = (unsigned long long)f2key(v + 0.0f) << 32;
    if (lab_c[c] == me) {
      if (!(skip_diag && c == row)) {
        const unsigned long long kp = k | (0xffffffffu - (unsigned)c);
        best_p = kp > best_p ? kp : best_p;
      }
    } else {
      const unsigned long long kn = k | (unsigned)c;
      best_n = kn < best_n ? kn : best_n;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long p = __shfl_xor_sync(FULL, best_p, o), q = __shfl_xor_sync(FULL, best_n, o);
    best_p = p > best_p ? p : best_p;
    best_n = q < best_n ? q : best_n;
  }
  if (lane == 0) {
    const bool hp = best_p != 0ull, hn = best_n != ~0ull;
    dist_ap[row] = hp ? key2f((uint32_t)(best_p >> 32)) : -INFINITY;
    dist_an[row] = hn ? key2f((uint32_t)(best_n >> 32)) : INFINITY;
    if (p_ind) p_ind[row] = hp ? (int64_t)(0xffffffffu - (uint32_t)best_p) : -1;
    if (n_ind) n_ind[row] = hn ? (int64_t)(uint32_t)best_n : -1;
  }
}

// G[i, j] = g_ap[i] [j == p_i] + g_an[i] [j == n_i]; every element written once (no memset, no atomics)
__global__ void __launch_bounds__(256)
mine_grad_kernel(const float* __restrict__ g_ap, const float* __restrict__ g_an, const int64_t* __restrict__ p_ind,
                 const int64_t* __restrict__ n_ind, int m, int n, float* __restrict__ G, long long ldg) {
  const int row = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  float v = 0.f;
  if (p_ind[row] == c) v += g_ap[row];
  if (n_ind[row] == c) v += g_an[row];
  G[(long long)row * ldg + c] = v;
}

// out[r, c] (+)= sum_t W(r,t) (P[r,c] - Q[t,c]),  W(r,t) = G(r,t) / dist(r,t) where dist > 1e-6 (clamp inactive), else 0.
// (r,t) addresses G and dist through two strides, so the same kernel serves gx (r = anchor) and gy (r = column).
// CTA = 8 rows x 128 feature columns, 128 threads; t walks in order: fixed summation order.
constexpr int GR = 8, GT = 64;
__global__ void __launch_bounds__(128)
pair_grad_kernel(const float* __restrict__ P, long long ldp, int rows, const float* __restrict__ Q, long long ldq,
                 int others, int d, const float* __restrict__ dist, const float* __restrict__ G, long long s_r,
                 long long s_t, long long g_r, long long g_t, float* __restrict__ out, long long ld_out,
                 int accumulate) {
  __shared__ float Ws[GR][GT];
  const int r0 = blockIdx.x * GR;
  const int c = blockIdx.y * 128 + threadIdx.x;
  float acc[GR], sw[GR];
#pragma unroll
  for (int i = 0; i < GR; ++i) acc[i] = 0.f, sw[i] = 0.f;
  for (int t0 = 0; t0 < others; t0 += GT) {
    for (int e = threadIdx.x; e < GR * GT; e += 128) {
      const int i = e / GT, t = e % GT;
      float w = 0.f;
      if (r0 + i < rows && t0 + t < others) {
        const float dv = dist[(long long)(r0 + i) * s_r + (long long)(t0 + t) * s_t];
        const float gv = G[(long long)(r0 + i) * g_r + (long long)(t0 + t) * g_t];
        w = dv > 1e-6f ? __fdiv_rn(gv, dv) : 0.f;
      }
      Ws[i][t] = w;
    }
    __syncthreads();
    const int tn = min(GT, others - t0);
    for (int t = 0; t < tn; ++t) {
      const float qv = c < d ? Q[(long long)(t0 + t) * ldq + c] : 0.f;
#pragma unroll
      for (int i = 0; i < GR; ++i) {
        const float w = Ws[i][t];
        acc[i] = __fmaf_rn(w, qv, acc[i]);
        sw[i] += w;
      }
    }
    __syncthreads();
  }
  if (c < d) {
#pragma unroll
    for (int i = 0; i < GR; ++i) {
      const int r = r0 + i;
      if (r < rows) {
        const float v = __fmaf_rn(sw[i], P[(long long)r * ldp + c], -acc[i]);
        float* o = out + (long long)r * ld_out + c;
        *o = accumulate ? *o + v : v;
      }
    }
  }
}

// Sparse backward of distance + mining taken together (the triplet losses never look at the rest of the matrix):
//   gx_i = wp_i (x_i - y_{p_i}) + wn_i (x_i - y_{n_i}),   wp = g_ap / dist_ap, wn = g_an / dist_an (0 where clamped / absent)
//   gy_j = sum over anchors i that selected column j of w_i (y_j - x_i), anchors in ascending order (deterministic)
// O((m + n) d) instead of the dense O(m n d).
__device__ __forceinline__ float pair_weight(float g, float dv, long long idx) {
  return (idx >= 0 && dv > 1e-6f) ? __fdiv_rn(g, dv) : 0.f;
}

__global__ void __launch_bounds__(256)
triplet_gx_kernel(const float* __restrict__ X, long long ldx, int m, const float* __restrict__ Y, long long ldy, int d,
                  const float* __restrict__ dist_ap, const float* __restrict__ dist_an,
                  const int64_t* __restrict__ p_ind, const int64_t* __restrict__ n_ind,
                  const float* __restrict__ g_ap, const float* __restrict__ g_an, float* __restrict__ gx,
                  long long ldgx) {
  const int i = blockIdx.x;
  const int c = blockIdx.y * 256 + threadIdx.x;
  if (c >= d) return;
  const long long p = p_ind[i], q = n_ind[i];
  const float wp = pair_weight(g_ap[i], dist_ap[i], p), wn = pair_weight(g_an[i], dist_an[i], q);
  const float xv = X[(long long)i * ldx + c];
  float v = 0.f;
  if (wp != 0.f) v = __fmaf_rn(wp, xv - Y[p * ldy + c], v);
  if (wn != 0.f) v = __fmaf_rn(wn, xv - Y[q * ldy + c], v);
  gx[(long long)i * ldgx + c] = v;
}

constexpr int TH = 512;    // hit-list capacity of one column's CTA (flushed when full)
__global__ void __launch_bounds__(256)
triplet_gy_kernel(const float* __restrict__ X, long long ldx, int m, const float* __restrict__ Y, long long ldy, int d,
                  const float* __restrict__ dist_ap, const float* __restrict__ dist_an,
                  const int64_t* __restrict__ p_ind, const int64_t* __restrict__ n_ind,
                  const float* __restrict__ g_ap, const float* __restrict__ g_an, float* __restrict__ gy,
                  long long ldgy, int accumulate) {
  __shared__ int hit_i[TH];
  __shared__ float hit_w[TH];
  __shared__ int warp_cnt[8];
  __shared__ int nhit;
  const int j = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) nhit = 0;
  __syncthreads();
  bool wrote = accumulate != 0;
  auto flush = [&]() {
    const int nh = nhit;
    for (int c = tid; c < d; c += 256) {
      const float yv = Y[(long long)j * ldy + c];
      float acc = 0.f;
      for (int h = 0; h < nh; ++h) acc = __fmaf_rn(hit_w[h], yv - X[(long long)hit_i[h] * ldx + c], acc);
      float* o = gy + (long long)j * ldgy + c;
      *o = wrote ? *o + acc : acc;
    }
    wrote = true;
    __syncthreads();
    if (tid == 0) nhit = 0;
    __syncthreads();
  };
  for (int i0 = 0; i0 < m; i0 += 256) {
    const int i = i0 + tid;
    float w = 0.f;
    if (i < m) {
      if (p_ind[i] == j) w += pair_weight(g_ap[i], dist_ap[i], j);
      if (n_ind[i] == j) w += pair_weight(g_an[i], dist_an[i], j);
    }
    const unsigned b = __ballot_sync(FULL, w != 0.f);
    if (lane == 0) warp_cnt[warp] = __popc(b);
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < warp) base += warp_cnt[k];
      total += warp_cnt[k];
    }
    if (nhit + total > TH) {      // uniform: nhit and total are the same for every thread
      flush();
    }
    if (w != 0.f) {
      const int slot = nhit + base + __popc(b & ((1u << lane) - 1));
      hit_i[slot] = i;
      hit_w[slot] = w;
    }
    __syncthreads();
    if (tid == 0) nhit += total;
    __syncthreads();
  }
  if (nhit > 0 || !wrote) flush();
}

// ---------------------------------------------------------------------------------------------
// Large batches (B >= ~1 024, cross-batch memories): the forward is a real GEMM (6.4 GFLOP at 1 024 x 1 024 x 3 072), so
// it goes through the evaluation path's tensor-core contraction -- rows split into fp16 hi/lo operands (prep.cu; one
// power-of-two scale per operand from its measured maximum), dist_tc2_kernel (x.y to ~2^-22 |x||y|, float32
// accumulation; upper triangle + mirrored stores when x and y are the same rows), then the clamped square root in place.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sqrt_clamp_kernel(float* __restrict__ D, long long ld, int m, int n4, int n, float floor_v) {
  const long long total = (long long)m * n4;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int r = (int)(i / n4), c = (int)(i % n4) * 4;
    float4* p = reinterpret_cast<float4*>(D + (long long)r * ld + c);        // pitch % 4 == 0: the pad columns exist
    float4 v = *p;
    v.x = __fsqrt_rn(fmaxf(v.x, floor_v));
    v.y = __fsqrt_rn(fmaxf(v.y, floor_v));
    v.z = __fsqrt_rn(fmaxf(v.z, floor_v));
    v.w = __fsqrt_rn(fmaxf(v.w, floor_v));
    if (c + 3 < n) {
      *p = v;
    } else {                                                                 // ragged last group: pad columns untouched
      float* q = reinterpret_cast<float*>(p);
      q[0] = v.x;
      if (c + 1 < n) q[1] = v.y;
      if (c + 2 < n) q[2] = v.z;
    }
  }
}

namespace {
struct PairTcPlan {
  int dp;
  bool same;
  float *x32a, *sqa, *nrma, *x32b, *sqb, *nrmb;
  __half *hia, *loa, *hib, *lob;
  uint32_t *amaxa, *amaxb;
};
size_t pair_tc_carve(PairTcPlan& p, unsigned char* base, int m, int n, int d, bool same) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    unsigned char* at = base ? base + off : nullptr;
    off = (off + bytes + 255) & ~(size_t)255;
    return at;
  };
  p.dp = padded_width(d);
  p.same = same;
  p.x32a = reinterpret_cast<float*>(take((size_t)m * p.dp * 4));
  p.hia = reinterpret_cast<__half*>(take((size_t)m * p.dp * 2));
  p.loa = reinterpret_cast<__half*>(take((size_t)m * p.dp * 2));
  p.sqa = reinterpret_cast<float*>(take((size_t)m * 4));
  p.nrma = reinterpret_cast<float*>(take((size_t)m * 4));
  p.amaxa = reinterpret_cast<uint32_t*>(take(16));
  if (same) {
    p.x32b = p.x32a; p.hib = p.hia; p.lob = p.loa; p.sqb = p.sqa; p.nrmb = p.nrma; p.amaxb = p.amaxa;
  } else {
    p.x32b = reinterpret_cast<float*>(take((size_t)n * p.dp * 4));
    p.hib = reinterpret_cast<__half*>(take((size_t)n * p.dp * 2));
    p.lob = reinterpret_cast<__half*>(take((size_t)n * p.dp * 2));
    p.sqb = reinterpret_cast<float*>(take((size_t)n * 4));
    p.nrmb = reinterpret_cast<float*>(take((size_t)n * 4));
    p.amaxb = reinterpret_cast<uint32_t*>(take(16));
  }
  return off;
}
}  // namespace

long long pair_dist_tc_workspace_bytes(int m, int n, int d, int same) {
  if (m < 0 || n < 0 || d < 1) return -1;
  PairTcPlan p{};
  return (long long)pair_tc_carve(p, nullptr, m, n, d, same != 0);
}

int launch_pair_dist_tc(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                        void* workspace, long long ws_bytes, float* out, long long ldo, cudaStream_t s) {
  if (m == 0 || n == 0) return REID_OK;
  const bool same = X == Y && ldx == ldy && m == n;
  PairTcPlan p{};
  const size_t need = pair_tc_carve(p, static_cast<unsigned char*>(workspace), m, n, d, same);
  REID_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && (size_t)ws_bytes >= need,
               "pair_dist_tc: workspace of %lld B (256-byte aligned) needed, got %lld", (long long)need, ws_bytes);
  REID_REQUIRE(ldo % 4 == 0 && ((uintptr_t)out & 15) == 0, "pair_dist_tc: output rows must be 16-byte aligned (pitch %lld)",
               ldo);
  int rc;
  if ((rc = launch_prep(X, m, d, ldx, 0, p.x32a, p.hia, p.loa, p.dp, p.sqa, p.nrma, p.amaxa, s))) return rc;
  if (!same && (rc = launch_prep(Y, n, d, ldy, 0, p.x32b, p.hib, p.lob, p.dp, p.sqb, p.nrmb, p.amaxb, s))) return rc;
  if ((rc = launch_dist_tc(p.hia, p.loa, p.dp, p.sqa, p.amaxa, m, p.hib, p.lob, p.dp, p.sqb, p.amaxb, n, p.dp, 0, out, ldo,
                           nullptr, nullptr, same ? 1 : 0, s))) return rc;
  const int n4 = (n + 3) / 4;
  const long long total = (long long)m * n4;
  const long long want = (total + 255) / 256;
  const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
  sqrt_clamp_kernel<<<blocks, 256, 0, s>>>(out, ldo, m, n4, n, 1e-12f);
  return check_launch("sqrt_clamp_kernel");
}

int launch_pair_dist(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d, float* out,
                     long long ldo, cudaStream_t s) {
  if (m == 0 || n == 0) return REID_OK;
  dim3 grid(cdiv(n, PT), cdiv(m, PT));
  pair_dist_kernel<<<grid, 256, 0, s>>>(X, ldx, m, Y, ldy, n, d, out, ldo);
  return check_launch("pair_dist_kernel");
}

int launch_hard_mine(const float* D, long long ld, int m, int n, const int64_t* lab_r, const int64_t* lab_c,
                     int skip_diag, float* dist_ap, float* dist_an, int64_t* p_ind, int64_t* n_ind, cudaStream_t s) {
  if (m == 0) return REID_OK;
  hard_mine_kernel<<<cdiv(m, 8), 256, 0, s>>>(D, ld, m, n, lab_r, lab_c, skip_diag, dist_ap, dist_an, p_ind, n_ind);
  return check_launch("hard_mine_kernel");
}

int launch_mine_grad(const float* g_ap, const float* g_an, const int64_t* p_ind, const int64_t* n_ind, int m, int n,
                     float* G, long long ldg, cudaStream_t s) {
  if (m == 0 || n == 0) return REID_OK;
  dim3 grid(cdiv(n, 256), m);
  mine_grad_kernel<<<grid, 256, 0, s>>>(g_ap, g_an, p_ind, n_ind, m, n, G, ldg);
  return check_launch("mine_grad_kernel");
}

int launch_pair_grad(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                     const float* dist, long long ldd, const float* G, long long ldg, float* gx, long long ldgx,
                     float* gy, long long ldgy, int accumulate_gy, cudaStream_t s) {
  if (d == 0) return REID_OK;
  if (gx && m > 0) {
    dim3 grid(cdiv(m, GR), cdiv(d, 128));
    pair_grad_kernel<<<grid, 128, 0, s>>>(X, ldx, m, Y, ldy, n, d, dist, G, ldd, 1, ldg, 1, gx, ldgx, 0);
    int rc = check_launch("pair_grad_kernel");
    if (rc != REID_OK) return rc;
  }
  if (gy && n > 0) {
    dim3 grid(cdiv(n, GR), cdiv(d, 128));
    pair_grad_kernel<<<grid, 128, 0, s>>>(Y, ldy, n, X, ldx, m, d, dist, G, 1, ldd, 1, ldg, gy, ldgy, accumulate_gy);
    return check_launch("pair_grad_kernel");
  }
  return REID_OK;
}

int launch_triplet_grad(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                        const float* dist_ap, const float* dist_an, const int64_t* p_ind, const int64_t* n_ind,
                        const float* g_ap, const float* g_an, float* gx, long long ldgx, float* gy, long long ldgy,
                        int accumulate_gy, cudaStream_t s) {
  if (m == 0 || d == 0) return REID_OK;
  if (gx) {
    dim3 grid(m, cdiv(d, 256));
    triplet_gx_kernel<<<grid, 256, 0, s>>>(X, ldx, m, Y, ldy, d, dist_ap, dist_an, p_ind, n_ind, g_ap, g_an, gx, ldgx);
    int rc = check_launch("triplet_gx_kernel");
    if (rc != REID_OK) return rc;
  }
  if (gy && n > 0) {
    triplet_gy_kernel<<<n, 256, 0, s>>>(X, ldx, m, Y, ldy, d, dist_ap, dist_an, p_ind, n_ind, g_ap, g_an, gy, ldgy,
                                        accumulate_gy);
    return check_launch("triplet_gy_kernel");
  }
  return REID_OK;
}

}  // namespace reid
