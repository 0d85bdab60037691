// Internal launcher declarations (one per kernel family).  The public surface is include/reid_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace reid {

// int(np.around(k1 / 2)) + 1 with round-half-to-even (reranking.py:60)
static inline int half_k(int k1) {
  int f = k1 / 2;
  if (k1 & 1) f += (f & 1);      // x.5 -> nearest even
  return f + 1;
}
// structural bound on |E_i|: |R| + |R| * h  with |R| <= k1+1
static inline int v_row_capacity(int k1) { return (k1 + 1) * (half_k(k1) + 1); }
static inline int padded_width(int d) { return (d + 63) / 64 * 64; }   // tensor-core K block
// row pitch of the 32-column chunk minima: 8 slots per 256-column tile, so a tile's slots are 16-byte aligned groups
static inline int chunk_pitch(int n) { return 8 * ((n + 255) / 256); }

int launch_prep_rows(const float* X, const float* X2, long long ld_x2, int n, int d, long long ld_x, int normalize,
                     float* x32, __half* hi, __half* lo, int dp, float* sq, float* nrm, uint32_t* amax, int preset_amax,
                     cudaStream_t s);
int launch_prep(const float* X, int n, int d, long long ld_x, int normalize, float* x32, __half* hi, __half* lo,
                int dp, float* sq, float* nrm, uint32_t* amax, cudaStream_t s);
int launch_dist_simt(const float* A, long long lda, const float* ra, int m, const float* B, long long ldb,
                     const float* rb, int n, int dp, int metric, float* out, long long ld_out,
                     int* rowmax_key, cudaStream_t s);
int launch_dist_tc_block(const __half* Ahi, const __half* Alo, long long lda, const float* ra, const uint32_t* a_amax,
                         int m, const __half* Bhi, const __half* Blo, long long ldb, const float* rb,
                         const uint32_t* b_amax, int n, int dp, int metric, float* out, long long ld_out,
                         int* rowmax_key, float* chunk_min, int cm_pitch, int mode, int* rowmax_key_m,
                         float* chunk_min_m, cudaStream_t s);
int launch_dist_tc(const __half* Ahi, const __half* Alo, long long lda, const float* ra, const uint32_t* a_amax,
                   int m, const __half* Bhi, const __half* Blo, long long ldb, const float* rb,
                   const uint32_t* b_amax, int n, int dp, int metric, float* out, long long ld_out,
                   int* rowmax_key, float* chunk_min, int sym_mode, cudaStream_t s);
int dist_clock_probe(long long* host_out4);
int launch_select_topk_hinted(const float* D, int rows, int cols, long long ld, const int* rowmax_key,
                              const float* chunk_min, int K, int* nbr, float* nbr_val, int cap_override,
                              int upper_only, cudaStream_t s);
int launch_rowmax(const float* D, int rows, int cols, long long ld, int* key, cudaStream_t s);
size_t select_scatter_workspace_bytes(int n, int cap_override);
int launch_select_topk_scatter(const float* D, int n, long long ld, const int* rowmax_key, const float* chunk_min,
                               int K, int* nbr, float* nbr_val, int cap_override, void* workspace,
                               size_t workspace_bytes, cudaStream_t s);
int launch_select_topk(const float* D, int rows, int cols, long long ld, const int* rowmax_key, int K,
                       int* nbr, float* nbr_val, int cap_override, cudaStream_t s);
int launch_reciprocal_sets(const int* nbr, int N, int K, int k1, int* rfull, int* nfull, int* rhalf,
                           int* nhalf, cudaStream_t s);
int launch_expand_weights(const int* rfull, const int* nfull, const int* rhalf, const int* nhalf, int row0,
                          int row_stride, int rows, int k1, const float* D, long long ld, const int* rowmax_key, int* Vc,
                          __half* Vv, int* Vn, int upper_only, const float* const* panels, int out_global,
                          cudaStream_t s);
int launch_expand_weights_wl(const int* rfull, const int* nfull, const int* rhalf, const int* nhalf, int row0,
                             int row_stride, int rows, int k1, const float* D, long long ld, const int* rowmax_key,
                             int* Vc, __half* Vv, int* Vn, int upper_only, const float* const* panels, int out_global,
                             const int* wl_idx, const float* wl_val, int Kw, const float* x32, long long ldx,
                             const float* sq, int* wl_miss, const unsigned long long* mt_keys, const float* mt_vals,
                             int mt_n, unsigned long long* miss_out, int miss_cap, cudaStream_t s);
// stream_rerank.cu (re-ranking without any dense matrix)
int launch_lists_rekey(const int* widx, const float* wval, int rows, int Kw, const int* rowmax_key, int K, int* nbr,
                       int* ambiguous, cudaStream_t s);
int launch_overlap_hash(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn, const long long* colptr,
                        const unsigned long long* csc_ent, int q0, int q_stride, int nq, int q, int N, int cap_o,
                        int nrange, int* ocol, __half* ot, int* ocnt, int* overflow, cudaStream_t s);
int launch_list_gather_sorted(const float* panel, long long ld, int rows, int c0, int c1, int cap_o, const int* ocol,
                              const int* ocnt, float* od, cudaStream_t s);
int launch_rr_prepare(int nq, int q0, int q_stride, const int* glo, const int* glen, const long long* goff,
                      const int* list, const float* gd, const long long* q_cam, const long long* g_cam,
                      const int* rowmax_key, float c1, float c2, int cap_o, const int* ocol, const __half* ot,
                      const int* ocnt, unsigned long long* gkey, unsigned char* gjunk, unsigned long long* mk,
                      float* thrA, float* thrB, int* nmatch, int* overflow, int group_max, cudaStream_t s);
int launch_count_panel_rr(const float* panel, long long ld, int rows, int c0, int c1, const long long* goff,
                          const unsigned long long* mk, const float* thrA, const float* thrB, const int* nmatch,
                          int* bins, int group_max, cudaStream_t s);
int launch_score_counts_rr(int nq, int q0, int q_stride, const int* glo, const int* glen, const long long* goff,
                           const int* list, const float* gd, const unsigned long long* gkey, const unsigned char* gjunk,
                           const unsigned long long* mk, const float* thrA, const float* thrB, const int* nmatch,
                           const int* bins, int cap_o, const int* ocol, const __half* ot, const float* od,
                           const int* ocnt, const long long* q_pid, const long long* g_pid, const int* rowmax_key,
                           float c1, float c2, int max_rank, double* ap, int* first_hit, int* cmc_hist, int* n_valid,
                           int group_max, cudaStream_t s);
int launch_rr_topk(int nq, int q0, int q_stride, const int* tidx, const float* tval, int Kd, int g, int cap_o,
                   const int* ocol, const __half* ot, const float* od, const int* ocnt, const int* rowmax_key, float c1,
                   float c2, int K, int* out_idx, float* out_val, int* ambiguous, cudaStream_t s);
int launch_dist_tc_sharded(const __half* hi, const __half* lo, long long ld, const float* sq, const uint32_t* amax,
                           int n, int dp, const int* tiles, int ntiles, int world, float* out, long long ld_out,
                           int nloc, int* rowmax_key, float* chunk_min, int cm_pitch, float* chunk_min_m,
                           int cmm_pitch, cudaStream_t s);
int launch_shard_tau(int n, int rank, int world, const float* cm_d, int pitch_d, const float* cm_m, int pitch_m, int K,
                     unsigned* tau_key, cudaStream_t s);
int launch_shard_dstar(int n, int rank, int world, const int* stats, long long slot, int* key_g, int* key_l,
                       float* dstar, int* count, int* cthr_key, cudaStream_t s);
int launch_shard_push(const float* D, int nloc, int rank, int world, int n, long long ld, const float* cm_d,
                      int pitch_d, const float* dstar, const int* cthr_key, int* count, unsigned long long* pbuf,
                      int cap, cudaStream_t s);
int launch_shard_merge_local(const float* D, long long ld, int n, int rank, int world, const float* cm_d, int pitch_d,
                             const float* cm_m, int pitch_m, const int* key_g, const float* dstar, const int* count,
                             const unsigned long long* pbuf, int cap, int K, unsigned char* const* peers,
                             long long inbox_off, int nloc_max, cudaStream_t s);
int launch_shard_merge_final(int nloc, int rank, int world, int K, const unsigned long long* inbox, int nloc_max,
                             unsigned char* const* peers, long long nbr_off, float* nbr_val, cudaStream_t s);
// eval_e2e.cu: the whole one-GPU evaluation as one call over a caller-provided workspace
long long eval_workspace_bytes(int N, int q, int d, int k1, int k2, int max_rank);
int eval_reranked(const float* feats, long long ld_x, int N, int q, int d, const long long* pids, const long long* cams,
                  int k1, int k2, double lambda_value, int normalize, int max_rank, void* workspace, long long ws_bytes,
                  float* cmc, double* mAP, int* n_valid, double* ap_out, int* first_out, cudaStream_t s);
// fabric.cu: peer-memory arena, flag barrier, copy kernels
int fabric_alloc(long long bytes, void** out);
int fabric_free(void* p);
int fabric_export(void* p, void* handle64);
int fabric_open(const void* handle64, void** out);
int fabric_close(void* p);
int launch_fabric_barrier(unsigned char* const* peers, long long flags_off, int rank, int world, unsigned epoch,
                          cudaStream_t s);
int launch_fabric_bcast(unsigned char* const* peers, int rank, int world, long long offset, long long bytes,
                        cudaStream_t s);
int launch_fabric_bcast_rows(unsigned char* const* peers, int rank, int world, long long offset, int row_bytes, int n,
                             cudaStream_t s);
int launch_fabric_bcast_ell(unsigned char* const* peers, int rank, int world, long long off_cnt, long long off_cols,
                            long long off_vals, int cap1, int n, cudaStream_t s);
int launch_fabric_bcast_pool(unsigned char* const* peers, int rank, int world, long long off_cnt, long long off_cols,
                             long long off_vals, const long long* off, int n, cudaStream_t s);
int launch_qe_bound(const int* nbr, int K, int k2, const int* Vn, int row0, int row_stride, int rows,
                    int* ub, cudaStream_t s);
int launch_exclusive_scan(const int* in, int n, long long* out, cudaStream_t s);
int launch_qe_expand(const int* nbr, int K, int k2, const int* Vc, const __half* Vv, const int* Vn, int cap1,
                     int row0, int row_stride, int rows, const long long* qoff, int cap_sort, int* Qc, __half* Qv, int* Qn,
                     cudaStream_t s);
int launch_ell_offsets(int N, int cap1, long long* off, cudaStream_t s);
int launch_csc_count(const int* Qc, const long long* qoff, const int* Qn, int N, int* colcnt,
                     cudaStream_t s);
int launch_csc_fill(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn, int N,
                    const long long* colptr, int* cursor, unsigned long long* csc_ent, cudaStream_t s);
int launch_jaccard_mix(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                       const long long* colptr, const unsigned long long* csc_ent, const float* D,
                       long long ld, const int* rowmax_key, int q0, int q_stride, int nq_local, int q, int N, float c1,
                       float c2, float* out, long long ld_out, __half* jaccard_out, long long ld_j, void* ws,
                       long long ws_bytes, int cap, cudaStream_t s);
int launch_pid_index(const long long* g_pid, int g, int T, unsigned long long* keys, int* cnt, long long* start,
                     int* cursor, int* list, cudaStream_t s);
int launch_jaccard_score(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                         const long long* colptr, const unsigned long long* csc_ent, const float* D,
                         long long ld, const int* rowmax_key, int q0, int q_stride, int nq_local, int q, int N,
                         float c1, float c2, float* out, long long ld_out, __half* jaccard_out, long long ld_j,
                         const long long* q_pid, const long long* q_cam, const long long* g_cam,
                         const unsigned long long* pid_keys, const long long* pid_start, const int* pid_list,
                         int pid_T, int max_rank, double* ap, int* first_hit, int* cmc_hist, int* n_valid,
                         void* ws, long long ws_bytes, int cap, cudaStream_t s);
long long launch_jaccard_ws_bytes(int nq, int N, int cap);   // 0 if the sparse path cannot be used for this N / cap
int launch_rank_rows(const float* dist, int rows, int g, long long ld_dist, uint32_t* keyA, uint32_t* keyB,
                     int* idxA, int* idx_out, cudaStream_t s);
int launch_score(const int* ranked, int q, int g, const long long* q_pid, const long long* g_pid,
                 const long long* q_cam, const long long* g_cam, int max_rank, double* ap, int* first_hit,
                 int* cmc_hist, int* n_valid, cudaStream_t s);
int launch_score_dist(const float* dist, int q, int g, long long ld, const long long* q_pid,
                      const long long* g_pid, const long long* q_cam, const long long* g_cam, int max_rank,
                      double* ap, int* first_hit, int* cmc_hist, int* n_valid, cudaStream_t s);
int launch_score_finalize(const double* ap, int q, const int* cmc_hist, int max_rank, const int* n_valid,
                          float* cmc, double* mAP, cudaStream_t s);
int launch_transpose_add(const float* src, int n, long long ld_src, float* dst, long long ld_dst, int accumulate,
                         cudaStream_t s);

// train.cu
long long pair_dist_tc_workspace_bytes(int m, int n, int d, int same);
int launch_pair_dist_tc(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                        void* workspace, long long ws_bytes, float* out, long long ldo, cudaStream_t s);
int launch_pair_dist(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d, float* out,
                     long long ldo, cudaStream_t s);
int launch_hard_mine(const float* D, long long ld, int m, int n, const int64_t* lab_r, const int64_t* lab_c,
                     int skip_diag, float* dist_ap, float* dist_an, int64_t* p_ind, int64_t* n_ind, cudaStream_t s);
int launch_mine_grad(const float* g_ap, const float* g_an, const int64_t* p_ind, const int64_t* n_ind, int m, int n,
                     float* G, long long ldg, cudaStream_t s);
int launch_pair_grad(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                     const float* dist, long long ldd, const float* G, long long ldg, float* gx, long long ldgx,
                     float* gy, long long ldgy, int accumulate_gy, cudaStream_t s);

int launch_triplet_grad(const float* X, long long ldx, int m, const float* Y, long long ldy, int n, int d,
                        const float* dist_ap, const float* dist_an, const int64_t* p_ind, const int64_t* n_ind,
                        const float* g_ap, const float* g_an, float* gx, long long ldgx, float* gy, long long ldgy,
                        int accumulate_gy, cudaStream_t s);

// stream_sym.cu (neighbour lists of the symmetric N x N matrix from its upper half)
int launch_sym_push(const float* D, long long ld, int m, int row0, int n, int c0, const float* chunk_min, int cm_pitch,
                    const float* tau, int* cthr_key, int* count, unsigned long long* pbuf, int cap, cudaStream_t s);
int launch_sym_fold(int n, int c0, int Kw, float* Mv, int* Mi, float* tau, int* count, const unsigned long long* pbuf,
                    int cap, int* overflow, cudaStream_t s);
int launch_sym_merge(int rows, int S, int Kw, const float* Fv, const int* Fi, float* out_v, int* out_i, cudaStream_t s);

// stream_score.cu (CMC / mAP without the q x g matrix)
int launch_stream_group_lookup(const long long* q_pid, int nq, const unsigned long long* keys, const long long* start,
                               int T, int* glo, int* glen, cudaStream_t s);
int launch_stream_group_gather(const float* panel, long long ld, int rows, int c0, int c1, const int* glo,
                               const int* glen, const long long* goff, const int* list, float* gd, cudaStream_t s);
int launch_stream_group_sort(int nq, const int* glo, const int* glen, const long long* goff, const int* list,
                             const float* gd, const long long* q_cam, const long long* g_cam,
                             unsigned long long* mkeys, int* nmatch, int* overflow, int group_max, cudaStream_t s);
int launch_stream_count_panel(const float* panel, long long ld, int rows, int c0, int c1, const long long* goff,
                              const unsigned long long* mkeys, const int* nmatch, int* bins, int group_max,
                              cudaStream_t s);
int stream_group_cap(int group_max);   // shared-memory entries per query for a longest identity group: 2^k in [1024, 8192]
int launch_stream_score(int nq, const int* glen, const long long* goff, const unsigned long long* mkeys,
                        const int* nmatch, const int* bins, int max_rank, double* ap, int* first_hit, int* cmc_hist,
                        int* n_valid, cudaStream_t s);

}  // namespace reid
