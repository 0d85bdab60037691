// extern "C" surface declared in include/reid_b200.h: argument validation + launch.
#include <stdarg.h>
#include <string.h>

#include "../../include/reid_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace reid {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;   // kernels enqueued by this library (host-side count)
void count_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }

int check_launch(const char* what) {
  count_launches(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s launch failed: %s", what, cudaGetErrorString(e));
    return REID_ERR_CUDA;
  }
  return REID_OK;
}

__global__ void fill_i32_kernel(int* p, int n, int v) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void rowmax_decode_kernel(const int* key, int n, float* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = ordered_int2f(key[i]);
}

// dst[i, j] (+)= src[j, i], 32x32 tiles through shared memory
__global__ void __launch_bounds__(256)
transpose_add_kernel(const float* __restrict__ src, int n, long long ld_src, float* __restrict__ dst,
                     long long ld_dst, int accumulate) {
  __shared__ float tile[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    int sr = bx + r, sc = by + tx;          // src row in [bx, bx+32), col in [by, by+32)
    tile[r][tx] = (sr < n && sc < n) ? src[(long long)sr * ld_src + sc] : 0.0f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    int dr = by + r, dc = bx + tx;          // dst[dr, dc] = src[dc, dr]
    if (dr < n && dc < n) {
      float v = tile[tx][r];
      float* p = dst + (long long)dr * ld_dst + dc;
      *p = accumulate ? __fadd_rn(*p, v) : v;
    }
  }
}

int launch_transpose_add(const float* src, int n, long long ld_src, float* dst, long long ld_dst,
                         int accumulate, cudaStream_t s) {
  if (n == 0) return REID_OK;
  dim3 grid(cdiv(n, 32), cdiv(n, 32));
  transpose_add_kernel<<<grid, 256, 0, s>>>(src, n, ld_src, dst, ld_dst, accumulate);
  return check_launch("transpose_add_kernel");
}

static int g_device_ok[64];   // 0 unknown, 1 ok, -1 unsupported

static int require_sm100() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("no CUDA device: %s", cudaGetErrorString(e));
    return REID_ERR_UNSUPPORTED;
  }
  if (dev >= 0 && dev < 64 && g_device_ok[dev] == 1) return REID_OK;
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess || major != 10) {
    set_error("reid_b200 needs an sm_100 (B200) device; device %d is compute capability %d.x and "
              "there is no CPU or other-architecture fallback", dev, major);
    return REID_ERR_UNSUPPORTED;
  }
  if (dev >= 0 && dev < 64) g_device_ok[dev] = 1;
  return REID_OK;
}

}  // namespace reid

using namespace reid;

#define ENTER()                       \
  do {                                \
    int _rc = require_sm100();        \
    if (_rc != REID_OK) return _rc;   \
  } while (0)
#define S(stream) reinterpret_cast<cudaStream_t>(stream)
#define H(p) reinterpret_cast<__half*>(p)
#define HC(p) reinterpret_cast<const __half*>(p)

extern "C" {

int reid_abi_version(void) { return REID_ABI_VERSION; }
#ifndef REID_SRC_HASH
#define REID_SRC_HASH "unknown"
#endif
const char* reid_source_hash(void) { return REID_SRC_HASH; }
const char* reid_last_error(void) { return g_err; }

int reid_device_check(int device) {
  int major = 0;
  cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (e != cudaSuccess) {
    set_error("cudaDeviceGetAttribute(%d): %s", device, cudaGetErrorString(e));
    return REID_ERR_UNSUPPORTED;
  }
  if (major != 10) {
    set_error("device %d is compute capability %d.x, need 10.x (sm_100a)", device, major);
    return REID_ERR_UNSUPPORTED;
  }
  return REID_OK;
}

int64_t reid_launch_count(void) { return (int64_t)__atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int reid_half_k(int k1) { return half_k(k1); }
int reid_v_row_capacity(int k1) { return v_row_capacity(k1); }
int reid_padded_width(int d) { return padded_width(d); }
int reid_chunk_pitch(int n) { return chunk_pitch(n); }

int reid_prepare_features(const float* X, int n, int d, int64_t ld_x, int normalize, float* x32, reid_half* hi,
                          reid_half* lo, int dpad, float* sq, float* nrm, uint32_t* amax, void* stream) {
  ENTER();
  REID_REQUIRE(n >= 0 && d >= 1 && ld_x >= d, "prepare_features: bad shape n=%d d=%d ld=%lld", n, d,
               (long long)ld_x);
  REID_REQUIRE(dpad >= d && dpad % 64 == 0, "prepare_features: dpad=%d must be a multiple of 64 >= d=%d",
               dpad, d);
  return launch_prep(X, n, d, ld_x, normalize, x32, H(hi), H(lo), dpad, sq, nrm, amax, S(stream));
}

int reid_prepare_rows(const float* X, int n, int d, int64_t ld_x, int normalize, float* x32, reid_half* hi,
                      reid_half* lo, int dpad, float* sq, float* nrm, uint32_t* amax, int preset_amax, void* stream) {
  ENTER();
  REID_REQUIRE(n >= 0 && d >= 1 && ld_x >= d, "prepare_rows: bad shape n=%d d=%d ld=%lld", n, d, (long long)ld_x);
  REID_REQUIRE(dpad >= d && dpad % 64 == 0, "prepare_rows: dpad=%d must be a multiple of 64 >= d=%d", dpad, d);
  return launch_prep_rows(X, nullptr, 0, n, d, ld_x, normalize, x32, H(hi), H(lo), dpad, sq, nrm, amax,
                          preset_amax ? 1 : 0, S(stream));
}

int reid_prepare_rows_tta(const float* X, const float* X_aug, int n, int d, int64_t ld_x, int64_t ld_aug, int normalize,
                          float* x32, reid_half* hi, reid_half* lo, int dpad, float* sq, float* nrm, uint32_t* amax,
                          int preset_amax, void* stream) {
  ENTER();
  REID_REQUIRE(n >= 0 && d >= 1 && ld_x >= d && ld_aug >= d && X_aug, "prepare_rows_tta: bad shape n=%d d=%d", n, d);
  REID_REQUIRE(dpad >= d && dpad % 64 == 0, "prepare_rows_tta: dpad=%d must be a multiple of 64 >= d=%d", dpad, d);
  return launch_prep_rows(X, X_aug, ld_aug, n, d, ld_x, normalize, x32, H(hi), H(lo), dpad, sq, nrm, amax,
                          preset_amax ? 1 : 0, S(stream));
}

int reid_note_replayed_launches(int n) {
  if (n > 0) count_launches(n);
  return REID_OK;
}

int reid_dist_clock_probe(int64_t* host_out4) {
  ENTER();
  return dist_clock_probe(reinterpret_cast<long long*>(host_out4));
}


int reid_dist(const reid_half* Ahi, const reid_half* Alo, int64_t lda, const float* ra, const uint32_t* a_amax,
              int m, const reid_half* Bhi, const reid_half* Blo, int64_t ldb, const float* rb,
              const uint32_t* b_amax, int n, int dpad, int metric, float* out, int64_t ld_out,
              int32_t* rowmax_key, float* chunk_min, int sym_mode, void* stream) {
  ENTER();
  REID_REQUIRE(metric == 0 || metric == 1, "dist: unknown metric %d", metric);
  REID_REQUIRE(m >= 0 && n >= 0 && dpad > 0 && ld_out >= n && lda >= dpad && ldb >= dpad, "dist: bad shape");
  return launch_dist_tc(HC(Ahi), HC(Alo), lda, ra, a_amax, m, HC(Bhi), HC(Blo), ldb, rb, b_amax, n, dpad, metric,
                        out, ld_out, rowmax_key, chunk_min, sym_mode, S(stream));
}

int reid_dist_block(const reid_half* Ahi, const reid_half* Alo, int64_t lda, const float* ra, const uint32_t* a_amax,
                    int m, const reid_half* Bhi, const reid_half* Blo, int64_t ldb, const float* rb,
                    const uint32_t* b_amax, int n, int dpad, int metric, float* out, int64_t ld_out,
                    int32_t* rowmax_key, float* chunk_min, int cm_pitch, int mode, int32_t* rowmax_key_cols,
                    float* chunk_min_cols, void* stream) {
  ENTER();
  REID_REQUIRE(metric == 0 || metric == 1, "dist_block: unknown metric %d", metric);
  REID_REQUIRE(m >= 0 && n >= 0 && dpad > 0 && ld_out >= n && lda >= dpad && ldb >= dpad, "dist_block: bad shape");
  REID_REQUIRE(mode == 0 || mode == 2 || mode == 3, "dist_block: mode %d (0 rectangle, 2 upper-only square, 3 above the diagonal)", mode);
  REID_REQUIRE(cm_pitch >= 0, "dist_block: negative chunk-minimum pitch");
  return launch_dist_tc_block(HC(Ahi), HC(Alo), lda, ra, a_amax, m, HC(Bhi), HC(Blo), ldb, rb, b_amax, n, dpad, metric,
                              out, ld_out, rowmax_key, chunk_min, cm_pitch, mode,
                              mode == 3 ? rowmax_key_cols : rowmax_key, mode == 3 ? chunk_min_cols : chunk_min, S(stream));
}

int reid_dist_fp32(const float* A, int64_t lda, const float* ra, int m, const float* B, int64_t ldb,
                   const float* rb, int n, int dpad, int metric, float* out, int64_t ld_out, int32_t* rowmax_key,
                   void* stream) {
  ENTER();
  REID_REQUIRE(metric == 0 || metric == 1, "dist: unknown metric %d", metric);
  REID_REQUIRE(m >= 0 && n >= 0 && dpad > 0 && ld_out >= n && lda >= dpad && ldb >= dpad, "dist: bad shape");
  return launch_dist_simt(A, lda, ra, m, B, ldb, rb, n, dpad, metric, out, ld_out, rowmax_key, S(stream));
}

int reid_rowmax_init(int32_t* rowmax_key, int rows, void* stream) {
  ENTER();
  if (rows == 0) return REID_OK;
  fill_i32_kernel<<<cdiv(rows, 256), 256, 0, S(stream)>>>(rowmax_key, rows, REID_ROWMAX_INIT);
  return check_launch("fill_i32_kernel");
}

int reid_rowmax(const float* D, int rows, int cols, int64_t ld, int32_t* rowmax_key, void* stream) {
  ENTER();
  return launch_rowmax(D, rows, cols, ld, rowmax_key, S(stream));
}

int reid_rowmax_decode(const int32_t* rowmax_key, int rows, float* out, void* stream) {
  ENTER();
  if (rows == 0) return REID_OK;
  rowmax_decode_kernel<<<cdiv(rows, 256), 256, 0, S(stream)>>>(rowmax_key, rows, out);
  return check_launch("rowmax_decode_kernel");
}

int reid_transpose_add(const float* src, int n, int64_t ld_src, float* dst, int64_t ld_dst, int accumulate,
                       void* stream) {
  ENTER();
  return launch_transpose_add(src, n, ld_src, dst, ld_dst, accumulate, S(stream));
}

int reid_topk_rows(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key, int K,
                   int32_t* nbr, float* nbr_val, void* stream) {
  ENTER();
  return launch_select_topk(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val, 0, S(stream));
}

int reid_topk_rows_hinted(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key,
                          const float* chunk_min, int K, int32_t* nbr, float* nbr_val, int cap, void* stream) {
  ENTER();
  REID_REQUIRE(chunk_min != nullptr, "topk_rows_hinted: chunk_min is NULL");
  return launch_select_topk_hinted(D, rows, cols, ld, rowmax_key, chunk_min, K, nbr, nbr_val, cap, 0, S(stream));
}

int reid_topk_rows_hinted_tri(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key,
                              const float* chunk_min, int K, int32_t* nbr, float* nbr_val, int cap, int upper_only,
                              void* stream) {
  ENTER();
  REID_REQUIRE(chunk_min != nullptr, "topk_rows_hinted: chunk_min is NULL");
  return launch_select_topk_hinted(D, rows, cols, ld, rowmax_key, chunk_min, K, nbr, nbr_val, cap, upper_only ? 1 : 0,
                                   S(stream));
}

int64_t reid_topk_scatter_workspace_bytes(int n, int cap) { return (int64_t)select_scatter_workspace_bytes(n, cap); }

int reid_topk_rows_scatter(const float* D, int n, int64_t ld, const int32_t* rowmax_key, const float* chunk_min, int K,
                           int32_t* nbr, float* nbr_val, int cap, void* workspace, int64_t workspace_bytes,
                           void* stream) {
  ENTER();
  REID_REQUIRE(chunk_min != nullptr, "topk_rows_scatter: chunk_min is NULL");
  REID_REQUIRE(workspace_bytes >= 0, "topk_rows_scatter: negative workspace size");
  return launch_select_topk_scatter(D, n, ld, rowmax_key, chunk_min, K, nbr, nbr_val, cap, workspace,
                                    (size_t)workspace_bytes, S(stream));
}

int reid_topk_rows_capped(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key, int K,
                          int32_t* nbr, float* nbr_val, int cap, void* stream) {
  ENTER();
  return launch_select_topk(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val, cap, S(stream));
}

int reid_krecip_sets(const int32_t* nbr, int N, int K, int k1, int32_t* rfull, int32_t* nfull,
                     int32_t* rhalf, int32_t* nhalf, void* stream) {
  ENTER();
  REID_REQUIRE(k1 >= 1, "krecip_sets: k1=%d", k1);
  return launch_reciprocal_sets(nbr, N, K, k1, rfull, nfull, rhalf, nhalf, S(stream));
}

int reid_krecip_V(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf,
                  int row0, int row_stride, int rows, int k1, const float* D, int64_t ld,
                  const int32_t* rowmax_key, int32_t* Vc, reid_half* Vv, int32_t* Vn, void* stream) {
  ENTER();
  REID_REQUIRE(k1 >= 1 && row_stride >= 1 && row0 >= 0, "krecip_V: bad window");
  return launch_expand_weights(rfull, nfull, rhalf, nhalf, row0, row_stride, rows, k1, D, ld, rowmax_key,
                               Vc, H(Vv), Vn, 0, nullptr, 0, S(stream));
}

int reid_krecip_V_tri(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf,
                      int rows, int k1, const float* D, int64_t ld, const int32_t* rowmax_key, int32_t* Vc,
                      reid_half* Vv, int32_t* Vn, int upper_only, void* stream) {
  ENTER();
  REID_REQUIRE(k1 >= 1, "krecip_V: bad k1");
  return launch_expand_weights(rfull, nfull, rhalf, nhalf, 0, 1, rows, k1, D, ld, rowmax_key, Vc, H(Vv), Vn,
                               upper_only ? 1 : 0, nullptr, 0, S(stream));
}

/* ---- sharded symmetric path (one process per GPU, peer memory) ---------------------------------------------- */
#define PEERS(p) reinterpret_cast<unsigned char* const*>(p)

int reid_dist_sharded(const reid_half* hi, const reid_half* lo, int64_t ld, const float* sq, const uint32_t* amax, int n,
                      int dpad, const int32_t* tiles, int ntiles, int world, float* out, int64_t ld_out, int nloc,
                      int32_t* rowmax_key, float* chunk_min, int cm_pitch, float* chunk_min_m, int cmm_pitch,
                      void* stream) {
  ENTER();
  REID_REQUIRE(n >= 0 && dpad > 0 && ld >= dpad && ld_out >= n && world >= 1 && ntiles >= 0, "dist_sharded: bad shape");
  return launch_dist_tc_sharded(HC(hi), HC(lo), ld, sq, amax, n, dpad, tiles, ntiles, world, out, ld_out, nloc,
                                rowmax_key, chunk_min, cm_pitch, chunk_min_m, cmm_pitch, S(stream));
}

int reid_shard_tau(int n, int rank, int world, const float* cm_d, int pitch_d, const float* cm_m, int pitch_m, int K,
                   int32_t* tau_key, void* stream) {
  ENTER();
  return launch_shard_tau(n, rank, world, cm_d, pitch_d, cm_m, pitch_m, K, reinterpret_cast<unsigned*>(tau_key),
                          S(stream));
}

int reid_shard_dstar(int n, int rank, int world, const int32_t* stats, int64_t slot_ints, int32_t* key_g, int32_t* key_l,
                     float* dstar, int32_t* count, int32_t* cthr_key, void* stream) {
  ENTER();
  return launch_shard_dstar(n, rank, world, stats, slot_ints, key_g, key_l, dstar, count, cthr_key, S(stream));
}

int reid_shard_push(const float* D, int nloc, int rank, int world, int n, int64_t ld, const float* cm_d, int pitch_d,
                    const float* dstar, const int32_t* cthr_key, int32_t* count, void* pbuf, int cap, void* stream) {
  ENTER();
  return launch_shard_push(D, nloc, rank, world, n, ld, cm_d, pitch_d, dstar, cthr_key, count,
                           static_cast<unsigned long long*>(pbuf), cap, S(stream));
}

int reid_shard_merge_local(const float* D, int64_t ld, int n, int rank, int world, const float* cm_d, int pitch_d,
                           const float* cm_m, int pitch_m, const int32_t* key_g, const float* dstar,
                           const int32_t* count, const void* pbuf, int cap, int K, void* const* peers,
                           int64_t inbox_off, int nloc_max, void* stream) {
  ENTER();
  return launch_shard_merge_local(D, ld, n, rank, world, cm_d, pitch_d, cm_m, pitch_m, key_g, dstar, count,
                                  static_cast<const unsigned long long*>(pbuf), cap, K, PEERS(peers), inbox_off,
                                  nloc_max, S(stream));
}

int reid_shard_merge_final(int nloc, int rank, int world, int K, const void* inbox, int nloc_max, void* const* peers,
                           int64_t nbr_off, float* nbr_val, void* stream) {
  ENTER();
  return launch_shard_merge_final(nloc, rank, world, K, static_cast<const unsigned long long*>(inbox), nloc_max,
                                  PEERS(peers), nbr_off, nbr_val, S(stream));
}

int reid_krecip_V_sharded(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf,
                          int rank, int world, int rows, int k1, const float* D, int64_t ld, const int32_t* rowmax_key,
                          const void* const* panels, int32_t* Vc, reid_half* Vv, int32_t* Vn, void* stream) {
  ENTER();
  REID_REQUIRE(k1 >= 1 && world >= 1 && rank >= 0 && rank < world && panels, "krecip_V_sharded: bad arguments");
  return launch_expand_weights(rfull, nfull, rhalf, nhalf, rank, -world, rows, k1, D, ld, rowmax_key, Vc, H(Vv), Vn, 1,
                               reinterpret_cast<const float* const*>(panels), 1, S(stream));
}

/* ---- re-ranking without any dense matrix --------------------------------------------------------- */
#define LL(p) reinterpret_cast<const long long*>(p)
#define ULL(p) reinterpret_cast<const unsigned long long*>(p)

int reid_lists_rekey(const int32_t* widx, const float* wval, int rows, int Kw, const int32_t* rowmax_key, int K,
                     int32_t* nbr, int32_t* ambiguous, void* stream) {
  ENTER();
  return launch_lists_rekey(widx, wval, rows, Kw, rowmax_key, K, nbr, ambiguous, S(stream));
}

int reid_krecip_V_lists(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf, int rows,
                        int k1, const int32_t* rowmax_key, const int32_t* wl_idx, const float* wl_val, int Kw,
                        const float* x32, int64_t ldx, const float* sq, int32_t* Vc, reid_half* Vv, int32_t* Vn,
                        int32_t* misses, const uint64_t* mt_keys, const float* mt_vals, int mt_n, uint64_t* miss_out,
                        int miss_cap, void* stream) {
  ENTER();
  REID_REQUIRE(k1 >= 1 && Kw >= 1 && wl_idx && wl_val && x32 && sq && misses, "krecip_V_lists: bad arguments");
  return launch_expand_weights_wl(rfull, nfull, rhalf, nhalf, 0, 1, rows, k1, nullptr, 0, rowmax_key, Vc, H(Vv), Vn, 0,
                                  nullptr, 0, wl_idx, wl_val, Kw, x32, ldx, sq, misses, ULL(mt_keys), mt_vals, mt_n,
                                  reinterpret_cast<unsigned long long*>(miss_out), miss_cap, S(stream));
}

int reid_overlap_lists(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                       const int64_t* colptr, const uint64_t* csc_ent, int q0, int q_stride, int nq, int q, int N,
                       int cap_o, int nrange, int32_t* ocol, reid_half* ot, int32_t* ocnt, int32_t* overflow,
                       void* stream) {
  ENTER();
  REID_REQUIRE(q_stride != 0 && nq >= 0 && q >= 0, "overlap_lists: bad window");
  return launch_overlap_hash(Qc, HC(Qv), LL(qoff), Qn, LL(colptr), ULL(csc_ent), q0, q_stride, nq, q, N, cap_o, nrange,
                             ocol, H(ot), ocnt, overflow, S(stream));
}

int reid_stream_list_gather(const float* panel, int64_t ld, int rows, int c0, int c1, int cap_o, const int32_t* ocol,
                            const int32_t* ocnt, float* od, void* stream) {
  ENTER();
  return launch_list_gather_sorted(panel, ld, rows, c0, c1, cap_o, ocol, ocnt, od, S(stream));
}

int reid_sym_push(const float* panel, int64_t ld, int rows, int row0, int cols, int c0, const float* chunk_min,
                  int cm_pitch, const float* tau, int32_t* cthr_key, int32_t* count, uint64_t* pbuf, int cap, void* stream) {
  ENTER();
  REID_REQUIRE(rows >= 0 && cols >= 0 && ld >= cols && row0 >= 0 && c0 >= 0, "sym_push: bad panel");
  return launch_sym_push(panel, ld, rows, row0, cols, c0, chunk_min, cm_pitch, tau, cthr_key, count,
                         reinterpret_cast<unsigned long long*>(pbuf), cap, S(stream));
}

int reid_sym_fold(int cols, int c0, int Kw, float* list_val, int32_t* list_idx, float* tau, int32_t* count,
                  const uint64_t* pbuf, int cap, int32_t* overflow, void* stream) {
  ENTER();
  REID_REQUIRE(cols >= 0 && c0 >= 0, "sym_fold: bad block");
  return launch_sym_fold(cols, c0, Kw, list_val, list_idx, tau, count, ULL(pbuf), cap, overflow, S(stream));
}

int reid_sym_merge(int rows, int nlists, int Kw, const float* lists_val, const int32_t* lists_idx, float* out_val,
                   int32_t* out_idx, void* stream) {
  ENTER();
  REID_REQUIRE(rows >= 0, "sym_merge: bad shape");
  return launch_sym_merge(rows, nlists, Kw, lists_val, lists_idx, out_val, out_idx, S(stream));
}

int reid_rr_prepare(int nq, int q0, int q_stride, const int32_t* glo, const int32_t* glen, const int64_t* goff,
                    const int32_t* list, const float* gd, const int64_t* q_cam, const int64_t* g_cam,
                    const int32_t* rowmax_key, float c1, float c2, int cap_o, const int32_t* ocol, const reid_half* ot,
                    const int32_t* ocnt, uint64_t* gkey, uint8_t* gjunk, uint64_t* mk, float* thrA, float* thrB,
                    int32_t* nmatch, int32_t* overflow, int group_max, void* stream) {
  ENTER();
  return launch_rr_prepare(nq, q0, q_stride, glo, glen, LL(goff), list, gd, LL(q_cam), LL(g_cam), rowmax_key, c1, c2,
                           cap_o, ocol, HC(ot), ocnt, reinterpret_cast<unsigned long long*>(gkey), gjunk,
                           reinterpret_cast<unsigned long long*>(mk), thrA, thrB, nmatch, overflow, group_max,
                           S(stream));
}

int reid_stream_count_panel_rr(const float* panel, int64_t ld, int rows, int c0, int c1, const int64_t* goff,
                               const uint64_t* mk, const float* thrA, const float* thrB, const int32_t* nmatch,
                               int32_t* bins, int group_max, void* stream) {
  ENTER();
  return launch_count_panel_rr(panel, ld, rows, c0, c1, LL(goff), ULL(mk), thrA, thrB, nmatch, bins, group_max, S(stream));
}

int reid_stream_score_rr(int nq, int q0, int q_stride, const int32_t* glo, const int32_t* glen, const int64_t* goff,
                         const int32_t* list, const float* gd, const uint64_t* gkey, const uint8_t* gjunk,
                         const uint64_t* mk, const float* thrA, const float* thrB, const int32_t* nmatch,
                         const int32_t* bins, int cap_o, const int32_t* ocol, const reid_half* ot, const float* od,
                         const int32_t* ocnt, const int64_t* q_pid, const int64_t* g_pid, const int32_t* rowmax_key,
                         float c1, float c2, int max_rank, double* ap, int32_t* first_hit, int32_t* cmc_hist,
                         int32_t* n_valid, int group_max, void* stream) {
  ENTER();
  return launch_score_counts_rr(nq, q0, q_stride, glo, glen, LL(goff), list, gd, ULL(gkey), gjunk, ULL(mk), thrA, thrB,
                                nmatch, bins, cap_o, ocol, HC(ot), od, ocnt, LL(q_pid), LL(g_pid), rowmax_key, c1, c2,
                                max_rank, ap, first_hit, cmc_hist, n_valid, group_max, S(stream));
}

int reid_rr_topk(int nq, int q0, int q_stride, const int32_t* tidx, const float* tval, int Kd, int g, int cap_o,
                 const int32_t* ocol, const reid_half* ot, const float* od, const int32_t* ocnt,
                 const int32_t* rowmax_key, float c1, float c2, int K, int32_t* out_idx, float* out_val,
                 int32_t* ambiguous, void* stream) {
  ENTER();
  return launch_rr_topk(nq, q0, q_stride, tidx, tval, Kd, g, cap_o, ocol, HC(ot), od, ocnt, rowmax_key, c1, c2, K,
                        out_idx, out_val, ambiguous, S(stream));
}

int64_t reid_eval_workspace_bytes(int N, int q, int d, int k1, int k2, int max_rank) {
  return eval_workspace_bytes(N, q, d, k1, k2, max_rank);
}

int reid_eval_reranked(const float* feats, int64_t ld, int N, int q, int d, const int64_t* pids, const int64_t* cams,
                       int k1, int k2, double lambda_value, int normalize, int max_rank, void* workspace,
                       int64_t workspace_bytes, float* cmc, double* mAP, int32_t* n_valid, double* ap, int32_t* first_hit,
                       void* stream) {
  ENTER();
  REID_REQUIRE(feats && pids && cams && cmc && mAP && n_valid && ld >= d, "eval_reranked: NULL argument or bad pitch");
  return eval_reranked(feats, ld, N, q, d, LL(pids), LL(cams), k1, k2, lambda_value, normalize, max_rank, workspace,
                       workspace_bytes, cmc, mAP, n_valid, ap, first_hit, S(stream));
}

int reid_fabric_alloc(int64_t bytes, void** out) { ENTER(); return fabric_alloc(bytes, out); }
int reid_fabric_free(void* p) { return fabric_free(p); }
int reid_fabric_export(void* p, void* handle64) { ENTER(); return fabric_export(p, handle64); }
int reid_fabric_open(const void* handle64, void** out) { ENTER(); return fabric_open(handle64, out); }
int reid_fabric_close(void* p) { return fabric_close(p); }
int reid_fabric_barrier(void* const* peers, int64_t flags_off, int rank, int world, int epoch, void* stream) {
  ENTER();
  return launch_fabric_barrier(PEERS(peers), flags_off, rank, world, (unsigned)epoch, S(stream));
}
int reid_fabric_bcast(void* const* peers, int rank, int world, int64_t offset, int64_t bytes, void* stream) {
  ENTER();
  return launch_fabric_bcast(PEERS(peers), rank, world, offset, bytes, S(stream));
}
int reid_fabric_bcast_rows(void* const* peers, int rank, int world, int64_t offset, int row_bytes, int n, void* stream) {
  ENTER();
  return launch_fabric_bcast_rows(PEERS(peers), rank, world, offset, row_bytes, n, S(stream));
}
int reid_fabric_bcast_ell(void* const* peers, int rank, int world, int64_t off_cnt, int64_t off_cols, int64_t off_vals,
                          int cap1, int n, void* stream) {
  ENTER();
  return launch_fabric_bcast_ell(PEERS(peers), rank, world, off_cnt, off_cols, off_vals, cap1, n, S(stream));
}
int reid_fabric_bcast_pool(void* const* peers, int rank, int world, int64_t off_cnt, int64_t off_cols, int64_t off_vals,
                           const int64_t* off, int n, void* stream) {
  ENTER();
  return launch_fabric_bcast_pool(PEERS(peers), rank, world, off_cnt, off_cols, off_vals,
                                  reinterpret_cast<const long long*>(off), n, S(stream));
}

int reid_qe_bound(const int32_t* nbr, int K, int k2, const int32_t* Vn, int row0, int row_stride, int rows,
                  int32_t* ub, void* stream) {
  ENTER();
  return launch_qe_bound(nbr, K, k2, Vn, row0, row_stride, rows, ub, S(stream));
}

int reid_exclusive_scan(const int32_t* in, int n, int64_t* out, void* stream) {
  ENTER();
  return launch_exclusive_scan(in, n, reinterpret_cast<long long*>(out), S(stream));
}

int reid_qe_k2_capped(const int32_t* nbr, int K, int k2, const int32_t* Vc, const reid_half* Vv,
                      const int32_t* Vn, int cap1, int row0, int row_stride, int rows, const int64_t* qoff,
                      int cap_sort, int32_t* Qc, reid_half* Qv, int32_t* Qn, void* stream) {
  ENTER();
  return launch_qe_expand(nbr, K, k2, Vc, HC(Vv), Vn, cap1, row0, row_stride, rows,
                          reinterpret_cast<const long long*>(qoff), cap_sort, Qc, H(Qv), Qn, S(stream));
}

int reid_qe_k2(const int32_t* nbr, int K, int k2, const int32_t* Vc, const reid_half* Vv, const int32_t* Vn,
               int cap1, int row0, int row_stride, int rows, const int64_t* qoff, int32_t* Qc,
               reid_half* Qv, int32_t* Qn, void* stream) {
  return reid_qe_k2_capped(nbr, K, k2, Vc, Vv, Vn, cap1, row0, row_stride, rows, qoff, 4096, Qc, Qv, Qn,
                           stream);
}

int reid_ell_offsets(int N, int pitch, int64_t* off, void* stream) {
  ENTER();
  return launch_ell_offsets(N, pitch, reinterpret_cast<long long*>(off), S(stream));
}

int reid_csc_count(const int32_t* Qc, const int64_t* qoff, const int32_t* Qn, int N, int32_t* colcnt,
                   void* stream) {
  ENTER();
  return launch_csc_count(Qc, reinterpret_cast<const long long*>(qoff), Qn, N, colcnt, S(stream));
}

int reid_csc_fill(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn, int N,
                  const int64_t* colptr, int32_t* cursor, uint64_t* csc_ent, void* stream) {
  ENTER();
  return launch_csc_fill(Qc, HC(Qv), reinterpret_cast<const long long*>(qoff), Qn, N,
                         reinterpret_cast<const long long*>(colptr), cursor,
                         reinterpret_cast<unsigned long long*>(csc_ent), S(stream));
}

int64_t reid_jaccard_ws_bytes(int nq, int N, int cap) {
  return (int64_t)launch_jaccard_ws_bytes(nq, N, cap);
}

int reid_jaccard_mix(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                     const int64_t* colptr, const uint64_t* csc_ent, const float* D,
                     int64_t ld, const int32_t* rowmax_key, int q0, int q_stride, int nq, int q, int N,
                     float c1, float c2, float* out, int64_t ld_out, reid_half* jaccard_out, int64_t ld_j,
                     void* ws, int64_t ws_bytes, int cap, void* stream) {
  ENTER();
  REID_REQUIRE(q >= 0 && q <= N && q_stride != 0, "jaccard_mix: bad shape q=%d N=%d", q, N);
  return launch_jaccard_mix(Qc, HC(Qv), reinterpret_cast<const long long*>(qoff), Qn,
                            reinterpret_cast<const long long*>(colptr),
                            reinterpret_cast<const unsigned long long*>(csc_ent), D, ld, rowmax_key, q0, q_stride, nq,
                            q, N, c1, c2, out, ld_out, H(jaccard_out), ld_j, ws, ws_bytes, cap, S(stream));
}

int reid_pid_index_slots(int g) {
  long long T = 64;
  while (T < 2 * (long long)g) T <<= 1;
  return (int)T;
}

int reid_pid_index(const int64_t* g_pid, int g, int T, uint64_t* keys, int32_t* cnt, int64_t* start,
                   int32_t* cursor, int32_t* list, void* stream) {
  ENTER();
  REID_REQUIRE(g >= 0, "pid_index: g=%d", g);
  return launch_pid_index(reinterpret_cast<const long long*>(g_pid), g, T,
                          reinterpret_cast<unsigned long long*>(keys), cnt, reinterpret_cast<long long*>(start),
                          cursor, list, S(stream));
}

int reid_jaccard_score(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                       const int64_t* colptr, const uint64_t* csc_ent, const float* D,
                       int64_t ld, const int32_t* rowmax_key, int q0, int q_stride, int nq, int q, int N, float c1,
                       float c2, float* out, int64_t ld_out, const int64_t* q_pid, const int64_t* q_cam,
                       const int64_t* g_cam, const uint64_t* pid_keys, const int64_t* pid_start,
                       const int32_t* pid_list, int pid_T, int max_rank, double* ap, int32_t* first_hit,
                       int32_t* cmc_hist, int32_t* n_valid, void* ws, int64_t ws_bytes, int cap, void* stream) {
  ENTER();
  REID_REQUIRE(q >= 0 && q <= N && q_stride != 0, "jaccard_score: bad shape q=%d N=%d", q, N);
  REID_REQUIRE(ap != nullptr, "jaccard_score: ap is NULL (use reid_jaccard_mix for the matrix alone)");
  return launch_jaccard_score(Qc, HC(Qv), reinterpret_cast<const long long*>(qoff), Qn,
                              reinterpret_cast<const long long*>(colptr),
                              reinterpret_cast<const unsigned long long*>(csc_ent), D, ld, rowmax_key,
                              q0, q_stride, nq, q, N, c1, c2, out, ld_out, nullptr, 0,
                              reinterpret_cast<const long long*>(q_pid), reinterpret_cast<const long long*>(q_cam),
                              reinterpret_cast<const long long*>(g_cam),
                              reinterpret_cast<const unsigned long long*>(pid_keys),
                              reinterpret_cast<const long long*>(pid_start), pid_list, pid_T, max_rank, ap,
                              first_hit, cmc_hist, n_valid, ws, ws_bytes, cap, S(stream));
}

int reid_rank_rows(const float* dist, int rows, int g, int64_t ld, uint32_t* keyA, uint32_t* keyB,
                   int32_t* idxA, int32_t* ranked, void* stream) {
  ENTER();
  REID_REQUIRE(ld >= g, "rank_rows: pitch %lld < row length %d", (long long)ld, g);
  return launch_rank_rows(dist, rows, g, ld, keyA, keyB, idxA, ranked, S(stream));
}

int reid_score_rows(const int32_t* ranked, int q, int g, const int64_t* q_pid, const int64_t* g_pid,
                    const int64_t* q_cam, const int64_t* g_cam, int max_rank, double* ap, int32_t* first_hit,
                    int32_t* cmc_hist, int32_t* n_valid, void* stream) {
  ENTER();
  REID_REQUIRE(max_rank >= 1 && max_rank <= g, "score_rows: max_rank=%d outside [1, g=%d]", max_rank, g);
  return launch_score(ranked, q, g, reinterpret_cast<const long long*>(q_pid),
                      reinterpret_cast<const long long*>(g_pid), reinterpret_cast<const long long*>(q_cam),
                      reinterpret_cast<const long long*>(g_cam), max_rank, ap, first_hit, cmc_hist, n_valid,
                      S(stream));
}

int reid_score_dist(const float* dist, int q, int g, int64_t ld, const int64_t* q_pid, const int64_t* g_pid,
                    const int64_t* q_cam, const int64_t* g_cam, int max_rank, double* ap, int32_t* first_hit,
                    int32_t* cmc_hist, int32_t* n_valid, void* stream) {
  ENTER();
  REID_REQUIRE(max_rank >= 1 && max_rank <= g, "score_dist: max_rank=%d outside [1, g=%d]", max_rank, g);
  REID_REQUIRE(ld >= g, "score_dist: pitch %lld < row length %d", (long long)ld, g);
  return launch_score_dist(dist, q, g, ld, reinterpret_cast<const long long*>(q_pid),
                           reinterpret_cast<const long long*>(g_pid), reinterpret_cast<const long long*>(q_cam),
                           reinterpret_cast<const long long*>(g_cam), max_rank, ap, first_hit, cmc_hist, n_valid,
                           S(stream));
}

int reid_score_finalize(const double* ap, int q, const int32_t* cmc_hist, int max_rank, const int32_t* n_valid,
                        float* cmc, double* mAP, void* stream) {
  ENTER();
  return launch_score_finalize(ap, q, cmc_hist, max_rank, n_valid, cmc, mAP, S(stream));
}

// ---- training-side distance ops
int reid_pair_dist(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d, float* out,
                   int64_t ld_out, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 0 && d >= 1 && ldx >= d && ldy >= d && ld_out >= n, "pair_dist: bad shape m=%d n=%d d=%d",
               m, n, d);
  return launch_pair_dist(x, ldx, m, y, ldy, n, d, out, ld_out, S(stream));
}

int64_t reid_pair_dist_tc_workspace_bytes(int m, int n, int d, int same) {
  return pair_dist_tc_workspace_bytes(m, n, d, same);
}

int reid_pair_dist_tc(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d, void* workspace,
                      int64_t workspace_bytes, float* out, int64_t ld_out, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 0 && d >= 1 && ldx >= d && ldy >= d && ld_out >= n, "pair_dist_tc: bad shape m=%d n=%d d=%d",
               m, n, d);
  return launch_pair_dist_tc(x, ldx, m, y, ldy, n, d, workspace, workspace_bytes, out, ld_out, S(stream));
}

int reid_hard_mine(const float* dist, int64_t ld, int m, int n, const int64_t* lab_r, const int64_t* lab_c,
                   int skip_diag, float* dist_ap, float* dist_an, int64_t* p_ind, int64_t* n_ind, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 1 && ld >= n, "hard_mine: bad shape m=%d n=%d ld=%lld", m, n, (long long)ld);
  return launch_hard_mine(dist, ld, m, n, lab_r, lab_c, skip_diag ? 1 : 0, dist_ap, dist_an, p_ind, n_ind, S(stream));
}

int reid_hard_mine_grad(const float* g_ap, const float* g_an, const int64_t* p_ind, const int64_t* n_ind, int m, int n,
                        float* G, int64_t ldg, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 0 && ldg >= n && m <= 65535, "hard_mine_grad: bad shape m=%d n=%d", m, n);
  return launch_mine_grad(g_ap, g_an, p_ind, n_ind, m, n, G, ldg, S(stream));
}

int reid_pair_dist_grad(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d,
                        const float* dist, int64_t ldd, const float* G, int64_t ldg, float* gx, int64_t ldgx,
                        float* gy, int64_t ldgy, int accumulate_gy, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 0 && d >= 1 && ldx >= d && ldy >= d && ldd >= n && ldg >= n,
               "pair_dist_grad: bad shape m=%d n=%d d=%d", m, n, d);
  REID_REQUIRE((!gx || ldgx >= d) && (!gy || ldgy >= d), "pair_dist_grad: gradient pitch below d=%d", d);
  return launch_pair_grad(x, ldx, m, y, ldy, n, d, dist, ldd, G, ldg, gx, ldgx, gy, ldgy, accumulate_gy ? 1 : 0,
                          S(stream));
}

int reid_triplet_grad(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d,
                      const float* dist_ap, const float* dist_an, const int64_t* p_ind, const int64_t* n_ind,
                      const float* g_ap, const float* g_an, float* gx, int64_t ldgx, float* gy, int64_t ldgy,
                      int accumulate_gy, void* stream) {
  ENTER();
  REID_REQUIRE(m >= 0 && n >= 0 && d >= 1 && ldx >= d && ldy >= d && m <= 0x7fffffff / 2,
               "triplet_grad: bad shape m=%d n=%d d=%d", m, n, d);
  REID_REQUIRE((!gx || ldgx >= d) && (!gy || ldgy >= d), "triplet_grad: gradient pitch below d=%d", d);
  return launch_triplet_grad(x, ldx, m, y, ldy, n, d, dist_ap, dist_an, p_ind, n_ind, g_ap, g_an, gx, ldgx, gy, ldgy,
                             accumulate_gy ? 1 : 0, S(stream));
}

// ---- CMC / mAP over a streamed q x g problem
#define LL(p) reinterpret_cast<const long long*>(p)
int reid_stream_group_lookup(const int64_t* q_pid, int nq, const uint64_t* pid_keys, const int64_t* pid_start, int pid_T,
                             int32_t* glo, int32_t* glen, void* stream) {
  ENTER();
  REID_REQUIRE(nq >= 0 && pid_T >= 1, "stream_group_lookup: bad shape");
  return launch_stream_group_lookup(LL(q_pid), nq, reinterpret_cast<const unsigned long long*>(pid_keys), LL(pid_start),
                                    pid_T, glo, glen, S(stream));
}

int reid_stream_group_gather(const float* panel, int64_t ld, int rows, int c0, int c1, const int32_t* glo,
                             const int32_t* glen, const int64_t* goff, const int32_t* pid_list, float* gd, void* stream) {
  ENTER();
  REID_REQUIRE(rows >= 0 && c0 >= 0 && c1 >= c0 && ld >= c1 - c0, "stream_group_gather: bad panel");
  return launch_stream_group_gather(panel, ld, rows, c0, c1, glo, glen, LL(goff), pid_list, gd, S(stream));
}

int reid_stream_group_sort(int nq, const int32_t* glo, const int32_t* glen, const int64_t* goff, const int32_t* pid_list,
                           const float* gd, const int64_t* q_cam, const int64_t* g_cam, uint64_t* mkeys, int32_t* nmatch,
                           int32_t* overflow, int group_max, void* stream) {
  ENTER();
  REID_REQUIRE(nq >= 0, "stream_group_sort: bad shape");
  return launch_stream_group_sort(nq, glo, glen, LL(goff), pid_list, gd, LL(q_cam), LL(g_cam),
                                  reinterpret_cast<unsigned long long*>(mkeys), nmatch, overflow, group_max,
                                  S(stream));
}

int reid_stream_count_panel(const float* panel, int64_t ld, int rows, int c0, int c1, const int64_t* goff,
                            const uint64_t* mkeys, const int32_t* nmatch, int32_t* bins, int group_max, void* stream) {
  ENTER();
  REID_REQUIRE(rows >= 0 && c0 >= 0 && c1 >= c0 && ld >= c1 - c0 && ld % 4 == 0, "stream_count_panel: bad panel");
  REID_REQUIRE(((uintptr_t)panel & 15) == 0, "stream_count_panel: panel must be 16-byte aligned");
  return launch_stream_count_panel(panel, ld, rows, c0, c1, LL(goff), reinterpret_cast<const unsigned long long*>(mkeys),
                                   nmatch, bins, group_max, S(stream));
}

int reid_stream_score(int nq, const int32_t* glen, const int64_t* goff, const uint64_t* mkeys, const int32_t* nmatch,
                      const int32_t* bins, int max_rank, double* ap, int32_t* first_hit, int32_t* cmc_hist,
                      int32_t* n_valid, void* stream) {
  ENTER();
  REID_REQUIRE(nq >= 0 && max_rank >= 1, "stream_score: bad shape");
  return launch_stream_score(nq, glen, LL(goff), reinterpret_cast<const unsigned long long*>(mkeys), nmatch, bins,
                             max_rank, ap, first_hit, cmc_hist, n_valid, S(stream));
}

}  // extern "C"