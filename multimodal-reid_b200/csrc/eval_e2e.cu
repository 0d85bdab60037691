// One call for the whole re-ranked evaluation on one GPU (reid_eval_e2e / reid_workspace_bytes):
// evaluate.py:110-134 with reranking=True -- features in, CMC / mAP out -- sequenced in C++ so that a caller in any
// language needs no copy of the host layer's orchestration: L2-normalise + split operands, symmetric upper-only
// contraction, scatter selection, reciprocal sets, V weights, k2 expansion, inverted index, identity index, fused
// Jaccard + lambda-mix + CMC/mAP, finalisation.  Every buffer comes from ONE caller-provided workspace (sized by
// eval_workspace_bytes from the structural bounds, so there is no host read and no allocation anywhere in the call: the
// whole evaluation is a fixed sequence of launches on the stream and can be captured in a CUDA graph as is).
#include "common.cuh"
#include "kernels.h"

namespace reid {

namespace {
struct Carver {                        // bump allocation over the caller's workspace, 256-byte aligned pieces
  unsigned char* base;
  size_t off = 0;
  explicit Carver(void* p) : base(static_cast<unsigned char*>(p)) {}
  template <typename T>
  T* take(size_t count) {
    const size_t at = off;
    off = (off + count * sizeof(T) + 255) & ~(size_t)255;
    return base ? reinterpret_cast<T*>(base + at) : nullptr;
  }
};

struct EvalPlan {
  int N, q, g, d, dp, K, k1, k2, h, cap1, max_rank, T, cap_sort, sparse_cap;
  long long ld, cpitch, pool;
  size_t sel_ws, jac_ws;
  // pieces
  float *x32, *sq, *nrm, *D, *cmin;
  __half *hi, *lo, *Vv, *Qv;
  uint32_t* amax;
  int *key, *nbr, *rfull, *nfull, *rhalf, *nhalf, *Vc, *Vn, *ub, *Qc, *Qn, *colcnt, *cursor, *pcnt, *pcursor, *plist,
      *first, *hist, *nvalid;
  long long *qoff, *colptr, *pstart;
  unsigned long long *csc, *pkeys;
  unsigned char *selws, *jacws;
  double* ap;
};

size_t carve(EvalPlan& p, void* ws) {
  Carver c(ws);
  const size_t N = (size_t)p.N;
  p.x32 = c.take<float>(N * p.dp);
  p.hi = c.take<__half>(N * p.dp);
  p.lo = c.take<__half>(N * p.dp);
  p.sq = c.take<float>(N);
  p.nrm = c.take<float>(N);
  p.amax = c.take<uint32_t>(4);
  p.D = c.take<float>(N * (size_t)p.ld);
  p.key = c.take<int>(N);
  p.cmin = c.take<float>(N * (size_t)p.cpitch);
  p.nbr = c.take<int>(N * p.K);
  p.selws = c.take<unsigned char>(p.sel_ws);
  p.rfull = c.take<int>(N * (p.k1 + 1));
  p.nfull = c.take<int>(N);
  p.rhalf = c.take<int>(N * p.h);
  p.nhalf = c.take<int>(N);
  p.Vc = c.take<int>(N * p.cap1);
  p.Vv = c.take<__half>(N * p.cap1);
  p.Vn = c.take<int>(N);
  p.ub = c.take<int>(N);
  p.qoff = c.take<long long>(N + 1);
  if (p.k2 != 1) {
    p.Qc = c.take<int>((size_t)p.pool);
    p.Qv = c.take<__half>((size_t)p.pool);
    p.Qn = c.take<int>(N);
  } else {
    p.Qc = p.Vc; p.Qv = p.Vv; p.Qn = p.Vn;
  }
  p.colcnt = c.take<int>(N);
  p.cursor = c.take<int>(N);
  p.colptr = c.take<long long>(N + 1);
  p.csc = c.take<unsigned long long>((size_t)p.pool);
  p.pkeys = c.take<unsigned long long>((size_t)p.T);
  p.pcnt = c.take<int>((size_t)p.T + 1);
  p.pcursor = c.take<int>((size_t)p.T + 1);
  p.pstart = c.take<long long>((size_t)p.T + 2);
  p.plist = c.take<int>((size_t)(p.g > 0 ? p.g : 1));
  p.jacws = c.take<unsigned char>(p.jac_ws);
  p.ap = c.take<double>((size_t)(p.q > 0 ? p.q : 1));
  p.first = c.take<int>((size_t)(p.q > 0 ? p.q : 1));
  p.hist = c.take<int>((size_t)p.max_rank + 1);
  p.nvalid = p.hist ? p.hist + p.max_rank : nullptr;
  return c.off;
}

int make_plan(EvalPlan& p, int N, int q, int d, int k1, int k2, int max_rank) {
  REID_REQUIRE(N >= 1 && q >= 0 && q < N && d >= 1 && k1 >= 1 && k2 >= 1 && max_rank >= 1, "eval: bad shape");
  p.N = N; p.q = q; p.g = N - q; p.d = d; p.dp = padded_width(d);
  p.k1 = k1; p.k2 = k2; p.K = k1 + 1 > k2 ? k1 + 1 : k2;
  REID_REQUIRE(N >= p.K, "eval: re-ranking needs at least max(k1+1, k2) = %d samples, got %d", p.K, N);
  REID_REQUIRE(p.K <= 256, "eval: max(k1+1, k2) = %d exceeds 256", p.K);
  p.h = half_k(k1); p.cap1 = v_row_capacity(k1);
  p.max_rank = max_rank < p.g ? max_rank : p.g;
  p.ld = (N + 31) / 32 * 32;
  p.cpitch = chunk_pitch(N);
  p.pool = k2 != 1 ? (long long)N * k2 * p.cap1 : (long long)N * p.cap1;    // structural: a union of k2 rows of <= cap1 entries
  p.sel_ws = select_scatter_workspace_bytes(N, 0);
  p.sparse_cap = 8192;
  const long long jw = launch_jaccard_ws_bytes(q, N, p.sparse_cap);
  p.jac_ws = jw > 0 ? (size_t)jw : 0;
  long long T = 64;
  while (T < 2 * (long long)p.g) T <<= 1;
  p.T = (int)T;
  int cs = 32;                                            // per-warp merge buffers of the expansion: a typical union
  while (cs < (k2 * (k1 + 1) * 3) / 2 && cs < 4096) cs <<= 1;
  p.cap_sort = cs;
  return REID_OK;
}

__global__ void eval_fill_kernel(int* p, long long n, int v) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] = v;
}
}  // namespace

long long eval_workspace_bytes(int N, int q, int d, int k1, int k2, int max_rank) {
  EvalPlan p{};
  if (make_plan(p, N, q, d, k1, k2, max_rank) != REID_OK) return -1;
  return (long long)carve(p, nullptr);
}

int eval_reranked(const float* feats, long long ld_x, int N, int q, int d, const long long* pids, const long long* cams,
                  int k1, int k2, double lambda_value, int normalize, int max_rank, void* workspace, long long ws_bytes,
                  float* cmc, double* mAP, int* n_valid, double* ap_out, int* first_out, cudaStream_t s) {
  EvalPlan p{};
  if (int rc = make_plan(p, N, q, d, k1, k2, max_rank)) return rc;
  REID_REQUIRE(max_rank <= p.g, "eval: max_rank=%d exceeds the gallery (%d); the caller clips it (evaluate.py:37-39)",
               max_rank, p.g);
  const size_t need = carve(p, workspace);
  REID_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && (size_t)ws_bytes >= need,
               "eval: workspace of %lld B (256-byte aligned) needed, got %lld", (long long)need, ws_bytes);
  int rc;
  // K13: prologue (rows bounded by 1 after normalisation: the operand scale is fixed up front, as the host layer does)
  if (normalize) {
    eval_fill_kernel<<<1, 32, 0, s>>>(reinterpret_cast<int*>(p.amax), 1, 0x3f800000);     // bits of 1.0f (graph-capturable)
    if ((rc = check_launch("eval_fill_kernel"))) return rc;
    if ((rc = launch_prep_rows(feats, nullptr, 0, N, d, ld_x, 1, p.x32, p.hi, p.lo, p.dp, p.sq, p.nrm, p.amax, 1, s))) return rc;
  } else {
    if ((rc = launch_prep(feats, N, d, ld_x, 0, p.x32, p.hi, p.lo, p.dp, p.sq, p.nrm, p.amax, s))) return rc;
  }
  // K3 + K4: symmetric contraction, upper-only form; row maxima and chunk minima from the epilogue
  eval_fill_kernel<<<cdiv(N, 256), 256, 0, s>>>(p.key, N, REID_ROWMAX_INIT);
  if ((rc = check_launch("eval_fill_kernel"))) return rc;
  if ((rc = launch_dist_tc(p.hi, p.lo, p.dp, p.sq, p.amax, N, p.hi, p.lo, p.dp, p.sq, p.amax, N, p.dp, 0, p.D, p.ld, p.key,
                           p.cmin, 2, s))) return rc;
  // K5: scatter selection
  if ((rc = launch_select_topk_scatter(p.D, N, p.ld, p.key, p.cmin, p.K, p.nbr, nullptr, 0, p.selws, p.sel_ws, s))) return rc;
  // K6
  if ((rc = launch_reciprocal_sets(p.nbr, N, p.K, k1, p.rfull, p.nfull, p.rhalf, p.nhalf, s))) return rc;
  if ((rc = launch_expand_weights(p.rfull, p.nfull, p.rhalf, p.nhalf, 0, 1, N, k1, p.D, p.ld, p.key, p.Vc, p.Vv, p.Vn, 1,
                                  nullptr, 0, s))) return rc;
  // K7
  if (k2 != 1) {
    if ((rc = launch_qe_bound(p.nbr, p.K, k2, p.Vn, 0, 1, N, p.ub, s))) return rc;
    if ((rc = launch_exclusive_scan(p.ub, N, p.qoff, s))) return rc;
    if ((rc = launch_qe_expand(p.nbr, p.K, k2, p.Vc, p.Vv, p.Vn, p.cap1, 0, 1, N, p.qoff, p.cap_sort, p.Qc, p.Qv, p.Qn, s)))
      return rc;
  } else {
    if ((rc = launch_ell_offsets(N, p.cap1, p.qoff, s))) return rc;
  }
  // K8
  REID_CUDA(cudaMemsetAsync(p.colcnt, 0, (size_t)N * 4, s));
  REID_CUDA(cudaMemsetAsync(p.cursor, 0, (size_t)N * 4, s));
  if ((rc = launch_csc_count(p.Qc, p.qoff, p.Qn, N, p.colcnt, s))) return rc;
  if ((rc = launch_exclusive_scan(p.colcnt, N, p.colptr, s))) return rc;
  if ((rc = launch_csc_fill(p.Qc, p.Qv, p.qoff, p.Qn, N, p.colptr, p.cursor, p.csc, s))) return rc;
  // identity index of the gallery + K9-K12 fused
  if ((rc = launch_pid_index(pids + q, p.g, p.T, p.pkeys, p.pcnt, p.pstart, p.pcursor, p.plist, s))) return rc;
  REID_CUDA(cudaMemsetAsync(p.hist, 0, (size_t)(p.max_rank + 1) * 4, s));
  // fp16 array * python float: the scalar is cast to fp16 first (reranking.py:95); fp32 array * python float: to fp32
  const float c1 = __half2float(__double2half(1.0 - lambda_value));
  const float c2 = (float)lambda_value;
  double* ap = ap_out ? ap_out : p.ap;
  int* first = first_out ? first_out : p.first;
  if ((rc = launch_jaccard_score(p.Qc, p.Qv, p.qoff, p.Qn, p.colptr, p.csc, p.D, p.ld, p.key, 0, 1, q, q, N, c1, c2, nullptr,
                                 0, nullptr, 0, pids, cams, cams + q, p.pkeys, p.pstart, p.plist, p.T, max_rank, ap, first,
                                 p.hist, p.nvalid, p.jac_ws ? p.jacws : nullptr, (long long)p.jac_ws, p.sparse_cap, s)))
    return rc;
  if ((rc = launch_score_finalize(ap, q, p.hist, max_rank, p.nvalid, cmc, mAP, s))) return rc;
  REID_CUDA(cudaMemcpyAsync(n_valid, p.nvalid, 4, cudaMemcpyDeviceToDevice, s));
  return REID_OK;
}

}  // namespace reid
