/* reid_b200 — C-ABI of the B200-native retrieval-evaluation hot path.
 *
 * The reference (SuperbTUM/Multimodal-ReID @ 8dc4505d) is pure Python and has no FFI; the entry
 * points below are the stage boundaries of its evaluate.py / reranking.py, each citing the
 * reference lines it replaces.  The Python host layer (multimodal-reid_b200/{evaluate,reranking,
 * engine}.py) binds them with ctypes; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller; nothing is allocated behind the
 *     caller's back and nothing is freed;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *     enqueued asynchronously on it;
 *   - return value 0 = success, negative = error (REID_ERR_*); reid_last_error() returns a
 *     thread-local description of the most recent failure;
 *   - matrices are row-major with an explicit pitch `ld` counted in elements;
 *   - "row window" arguments (row0, row_stride, rows) select the probe rows
 *     row0 + l*row_stride, l in [0, rows): this is how ranks own a cyclic share of the N probes
 *     (rank r of P passes row0=r, row_stride=P); outputs of such stages are indexed by l.
 *   - there is no CPU fallback: on a device that is not compute capability 10.x every entry point
 *     fails with REID_ERR_UNSUPPORTED.
 */
#ifndef REID_B200_H
#define REID_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define REID_API __attribute__((visibility("default")))
#else
#define REID_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define REID_ABI_VERSION 3

#define REID_OK 0
#define REID_ERR_CUDA (-1)
#define REID_ERR_ARG (-2)
#define REID_ERR_UNSUPPORTED (-3)
#define REID_ERR_OVERFLOW (-4)

#define REID_METRIC_SQEUCLIDEAN 0 /* evaluate.py:7-13  */
#define REID_METRIC_ARCCOS 1      /* evaluate.py:16-26 */

typedef uint16_t reid_half; /* IEEE binary16 bit pattern */

REID_API int reid_abi_version(void);
/* hash of the sources this binary was built from (multimodal-reid_b200/build.py source_hash); the loader refuses a
 * library whose hash differs from the sources beside it */
REID_API const char* reid_source_hash(void);
REID_API const char* reid_last_error(void);
/* 0 if `device` is an sm_100 part this library can run on, REID_ERR_UNSUPPORTED otherwise. */
REID_API int reid_device_check(int device);

/* number of kernels this library has enqueued since it was loaded (host-side counter) */
REID_API int64_t reid_launch_count(void);

/* int(np.around(k1/2)) + 1 with banker's rounding             (reranking.py:60) */
REID_API int reid_half_k(int k1);
/* structural bound on the support of one row of V: (k1+1) * (half_k + 1) */
REID_API int reid_v_row_capacity(int k1);
/* feature width padded to the tensor-core K block (64) */
REID_API int reid_padded_width(int d);
/* row pitch (floats) of reid_dist's chunk_min output for n columns: 8 slots per 256-column tile, >= ceil(n/32) */
REID_API int reid_chunk_pitch(int n);

/* K13  evaluate.py:112-114, reranking.py:36-39.  Row L2-normalise (x / max(|x|, 1e-12)) when
 * `normalize`, zero-pad to `dpad` (reid_padded_width) -> x32[n, dpad]; sq[i] = sum x^2 and
 * nrm[i] = sqrt(sq[i]) of the rows as written; amax[0] = float bits of max|x|; and the float16
 * operands of the tensor-core contraction  x * 2^e = hi + lo  (hi = fp16(x * 2^e), lo = fp16 of the
 * exact remainder; e = 14 - floor(log2 max|x|), one exponent per call). */
REID_API int reid_prepare_features(const float* X, int n, int d, int64_t ld_x, int normalize, float* x32,
                          reid_half* hi, reid_half* lo, int dpad, float* sq, float* nrm, uint32_t* amax,
                          void* stream);
/* the same for a range of rows of a larger matrix.  preset_amax != 0: *amax already holds (the float bits of) a bound on
 * max|x| over the WHOLE matrix and is used as is -- rows prepared in separate calls then share one scaling exponent */
REID_API int reid_prepare_rows(const float* X, int n, int d, int64_t ld_x, int normalize, float* x32, reid_half* hi,
                      reid_half* lo, int dpad, float* sq, float* nrm, uint32_t* amax, int preset_amax, void* stream);
/* the same with the flip-TTA average of the callers fused in front (prompt_learning.py:775-776,
 * zero_shot_learning.py:121-125: `(embeddings + embeddings_augmented) / 2` before get_cmc_map): every row is
 * (X[i] + X_aug[i]) / 2 in float32 -- one rounded add, an exact halving -- formed on the fly, then treated as above */
REID_API int reid_prepare_rows_tta(const float* X, const float* X_aug, int n, int d, int64_t ld_x, int64_t ld_aug,
                          int normalize, float* x32, reid_half* hi, reid_half* lo, int dpad, float* sq, float* nrm,
                          uint32_t* amax, int preset_amax, void* stream);

/* K1/K2/K3  evaluate.py:7-26, reranking.py:36-41.  out[i, j] for A rows i < m, B rows j < n.
 * metric 0: (ra[i] + rb[j]) - 2 a.b with ra/rb = squared norms; metric 1: arccos(clip(a.b *
 * (1 / (ra[i] * rb[j])))) with ra/rb = norms.  If rowmax_key != NULL the per-row maximum of
 * `out` is folded in with atomicMax on an order-preserving int encoding (initialise the array
 * with reid_rowmax_init).  If chunk_min != NULL (reid_dist only) it receives the minimum of every
 * 32-column chunk, [m, ceil(n/32)] with row pitch reid_chunk_pitch(n), for reid_topk_rows_hinted.  lda/ldb are row pitches in elements, so a rank's cyclic share of the
 * probes is the same buffers with base + rank*dpad and pitch world*dpad.
 *   reid_dist       tcgen05 tensor cores: float16 hi/lo operands, float32 accumulation in TMEM
 *                   (hi.hi + hi.lo + lo.hi), TMA in and out.  The product path.
 *   reid_dist_fp32  CUDA-core float32 FMA on x32: bring-up / cross-check engine. */
/* When A and B of reid_dist are the same rows (same pointers, pitches, norms; m == n) the contraction computes only
 * the tiles touching the upper triangle and stores every off-diagonal tile a second time transposed (with the
 * mirrored rows' maxima and chunk minima): outside its 256-wide diagonal blocks the N x N matrix of re-ranking is an
 * exact mirror, at half the tensor-core work.  `sym_mode` (an argument of every call -- no process-wide state, so two
 * engines / streams / threads cannot disturb each other): 1 as described; 0 = every tile computed; 2 = upper-only:
 * mirrored tiles are NOT stored (their row maxima and chunk minima still are) -- half the output stream; entries left of
 * a row's 256-wide diagonal block are then undefined and consumers must read D[col, row] instead (the _tri entry points
 * do).  Modes 1 and 2 apply when both operands are the same rows; a rectangle of different row sets always runs in full. */
/* reid_launch_count() counts the kernels this library enqueued (host side).  A caller that captured a sequence of
 * entry points in a CUDA graph reports the launches of every replay here, so that the count stays what ran on the GPU. */
REID_API int reid_note_replayed_launches(int n);
/* Diagnostics: with the environment variable REID_TC_DBG & 64 set, block 0 of reid_dist's kernel records
 * {clock64, globaltimer ns} at entry and exit; this copies the four values to HOST memory (synchronises). */
REID_API int reid_dist_clock_probe(int64_t* host_out4);
REID_API int reid_dist(const reid_half* Ahi, const reid_half* Alo, int64_t lda, const float* ra,
              const uint32_t* a_amax, int m, const reid_half* Bhi, const reid_half* Blo, int64_t ldb,
              const float* rb, const uint32_t* b_amax, int n, int dpad, int metric, float* out,
              int64_t ld_out, int32_t* rowmax_key, float* chunk_min, int sym_mode, void* stream);
/* One block of a larger symmetric distance matrix, for callers that compute it chunk by chunk (the streamed evaluator:
 * the features of evaluate.py:110 are still arriving over PCIe while the first blocks are contracted).  Block origins
 * must be multiples of 256 rows / columns of the enclosing matrix; out, rowmax_key and chunk_min point at the block's
 * origin inside the enclosing arrays; cm_pitch is the enclosing chunk-minimum row pitch (reid_chunk_pitch(N)).
 *   mode 0  plain rectangle
 *   mode 2  square diagonal block (same rows on both sides), upper-only form as sym_mode 2
 *   mode 3  rectangle strictly above the diagonal: stored as is; its mirror image's row maxima / chunk minima go to
 *           rowmax_key_cols[block column] and chunk_min_cols[block column * cm_pitch + block row / 32]
 * The blocks of one matrix give bit for bit the matrix a single reid_dist call gives. */
REID_API int reid_dist_block(const reid_half* Ahi, const reid_half* Alo, int64_t lda, const float* ra, const uint32_t* a_amax,
                    int m, const reid_half* Bhi, const reid_half* Blo, int64_t ldb, const float* rb,
                    const uint32_t* b_amax, int n, int dpad, int metric, float* out, int64_t ld_out,
                    int32_t* rowmax_key, float* chunk_min, int cm_pitch, int mode, int32_t* rowmax_key_cols,
                    float* chunk_min_cols, void* stream);
REID_API int reid_dist_fp32(const float* A, int64_t lda, const float* ra, int m, const float* B, int64_t ldb,
                   const float* rb, int n, int dpad, int metric, float* out, int64_t ld_out,
                   int32_t* rowmax_key, void* stream);
REID_API int reid_rowmax_init(int32_t* rowmax_key, int rows, void* stream);
/* K4 (first half)  reranking.py:46.  Stand-alone row maximum for a caller-supplied matrix. */
REID_API int reid_rowmax(const float* D, int rows, int cols, int64_t ld, int32_t* rowmax_key, void* stream);
/* decode the int encoding into floats (diagnostics / tests) */
REID_API int reid_rowmax_decode(const int32_t* rowmax_key, int rows, float* out, void* stream);
/* reranking.py:43-44 / :33-34.  dst[i, j] (+)= src[j, i]: the reference reads column i of the
 * N x N matrix for probe i, the kernels read row i. */
REID_API int reid_transpose_add(const float* src, int n, int64_t ld_src, float* dst, int64_t ld_dst,
                       int accumulate, void* stream);

/* K5  reranking.py:46,48.  For every row: the K smallest O[i, j] = D[i, j] / rowmax[i] in
 * ascending (value, column) order -> nbr[rows, K] (and their values if nbr_val != NULL). */
REID_API int reid_topk_rows(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key, int K,
                   int32_t* nbr, float* nbr_val, void* stream);
/* same result, reading only the chunks whose recorded minimum can hold a neighbour (~K of cols/32) */
REID_API int reid_topk_rows_hinted(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key,
                          const float* chunk_min, int K, int32_t* nbr, float* nbr_val, int cap, void* stream);
/* same for the square symmetric matrix of reid_dist's upper-only mode (sym_mode 2): entries left of
 * a row's 256-wide diagonal block are read from their mirror image D[col, row] */
REID_API int reid_topk_rows_hinted_tri(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key,
                              const float* chunk_min, int K, int32_t* nbr, float* nbr_val, int cap, int upper_only,
                              void* stream);
/* same result for the upper-only matrix in scatter form: every stored element is offered to both rows it belongs to
 * (candidate records through `workspace`, reid_topk_scatter_workspace_bytes(n, cap) bytes, 16-byte aligned), so the
 * unstored half is never gathered; three launches (thresholds, push, merge) */
REID_API int64_t reid_topk_scatter_workspace_bytes(int n, int cap);
REID_API int reid_topk_rows_scatter(const float* D, int n, int64_t ld, const int32_t* rowmax_key, const float* chunk_min,
                           int K, int32_t* nbr, float* nbr_val, int cap, void* workspace, int64_t workspace_bytes,
                           void* stream);
/* reid_topk_rows_scatter without any device allocation by caller or library.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs, outputs and bits as reid_topk_rows_scatter(cap = 0); the
 * workspace is a device array built into the library (8.03 MiB per GPU that loads it, held until exit), which limits
 * n to reid_topk_static_max_rows() = 4096 rows -- larger n is refused with REID_ERR_ARG and a message naming the limit.
 * The call performs no cudaMalloc / cudaFree and uses no allocating library routine.  Every call shares the one array
 * (its contents are those of the last call; what must start clean is re-initialised by the call itself), so calls on one
 * stream may follow each other freely, but calls that OVERLAP on the GPU (other streams, concurrent graph replays) race:
 * serialise them. */
REID_API int reid_topk_static_max_rows(void);
REID_API int reid_topk_rows_scatter_static(const float* D, int n, int64_t ld, const int32_t* rowmax_key,
                                  const float* chunk_min, int K, int32_t* nbr, float* nbr_val, void* stream);
/* test hook: same, with the candidate buffer capped so the exact fallback path is exercised */
REID_API int reid_topk_rows_capped(const float* D, int rows, int cols, int64_t ld, const int32_t* rowmax_key,
                          int K, int32_t* nbr, float* nbr_val, int cap, void* stream);

/* K6a  reranking.py:53-56, :60-64.  Reciprocal neighbour sets at depth k1+1 (rfull[N, k1+1],
 * nfull[N]) and at depth half_k (rhalf[N, half_k], nhalf[N]) for all N rows of nbr[N, K]. */
REID_API int reid_krecip_sets(const int32_t* nbr, int N, int K, int k1, int32_t* rfull, int32_t* nfull,
                     int32_t* rhalf, int32_t* nhalf, void* stream);
/* K6b  reranking.py:57-71.  2/3-rule expansion, unique, Gaussian weights -> V rows in ELL form
 * (pitch reid_v_row_capacity(k1)): Vc columns ascending, Vv float16 weights, Vn counts.  D and
 * rowmax_key are indexed by the local row l of the window. */
REID_API int reid_krecip_V(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf,
                  const int32_t* nhalf, int row0, int row_stride, int rows, int k1, const float* D,
                  int64_t ld, const int32_t* rowmax_key, int32_t* Vc, reid_half* Vv, int32_t* Vn,
                  void* stream);
/* all N rows on one rank, D optionally in upper-only form (see reid_topk_rows_hinted_tri) */
REID_API int reid_krecip_V_tri(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf,
                      const int32_t* nhalf, int rows, int k1, const float* D, int64_t ld,
                      const int32_t* rowmax_key, int32_t* Vc, reid_half* Vv, int32_t* Vn, int upper_only,
                      void* stream);

/* ---- The whole evaluation as ONE call (reid_eval_e2e / reid_workspace_bytes) ------------------------------
 * evaluate.py:110-134 with reranking=True on one GPU: feats [N, d] float32 (row pitch ld, queries first), pids / cams
 * [N] int64 -> cmc[max_rank] float32, mAP float64, n_valid (all device memory; evaluate.py:82 asserts n_valid > 0, the
 * caller checks it after reading the result back).  normalize != 0 applies F.normalize first (evaluate.py:112-114).
 * k1 / k2 / lambda_value as re_ranking (reranking.py:29; R1_mAP_eval.compute passes 50, 15, 0.3).  max_rank <= g: the
 * caller clips it and prints the note of evaluate.py:37-39.  Optional per-query outputs ap[q] (float64, -1 for queries
 * whose identity is absent from the gallery) and first_hit[q].  Every intermediate buffer lives in `workspace`
 * (256-byte aligned, reid_eval_workspace_bytes(...) bytes: sized by structural bounds, dominated by the 4 N^2-byte
 * distance matrix): the call allocates nothing, never synchronises and never reads the device back -- a fixed
 * sequence of ~30 launches on `stream` that can be captured in a CUDA graph.  Same kernels, same bits as the host
 * layer's one-rank path (tests/test_gpu_api.py). */
REID_API int64_t reid_eval_workspace_bytes(int N, int q, int d, int k1, int k2, int max_rank);
REID_API int reid_eval_reranked(const float* feats, int64_t ld, int N, int q, int d, const int64_t* pids,
                       const int64_t* cams, int k1, int k2, double lambda_value, int normalize, int max_rank,
                       void* workspace, int64_t workspace_bytes, float* cmc, double* mAP, int32_t* n_valid, double* ap,
                       int32_t* first_hit, void* stream);

/* ---- Re-ranking without any dense matrix (100k queries x 1M gallery; reranking.py:46-100 and
 * evaluate.py:40-88 with D, V and the final matrix never materialised).  The contraction produces panels
 * (reid_dist on row / column blocks); these entry points work on lists.  Query windows use (q0, q_stride) as elsewhere.
 *
 * reid_lists_rekey: reranking.py:46,48 from per-row WIDE raw-distance lists (widx/wval [rows, Kw]: the Kw nearest columns
 * by (D, column), e.g. from reid_topk_rows_hinted over the panels of an N x N sweep) and the rows' global maxima: the
 * first K columns in (D / max, column) order -> nbr[rows, K].  *ambiguous counts rows whose K-th quotient equals the
 * list's last one (an unseen column could tie): enlarge Kw.
 * reid_krecip_V_lists: reid_krecip_V for all rows with every distance D[i, e] looked up in row i's wide list, then in
 * the caller's table of extra pairs (mt_keys = row << 32 | column ascending, mt_vals; may be NULL).  A pair in neither is
 * counted in *misses, written to miss_out (up to miss_cap) and, for this call only, recomputed from the float32 features
 * x32 / squared norms sq: call once to collect the misses, contract those pairs (reid_dist on the gathered rows -- the
 * same arithmetic as the panels), call again with the table.
 * reid_overlap_lists: reranking.py:86-92 for a window of queries with a shared-memory hash table instead of an N-bit row:
 * per query the gallery columns (target - q) it overlaps, ascending, and the float16 overlap sums -> ocol / ot at
 * [l * cap_o ...], ocnt[l] (-1 and *overflow += 1 when a query touches more than cap_o gallery columns, or more than
 * 8192 inside one of the column ranges the gallery is processed in: nrange of them, 0 = ceil(cap_o / 4096); every range
 * costs one walk of the query's inverted lists, so a caller that knows the real list sizes passes fewer).
 * reid_rr_prepare: per query, from the distances of its same-identity gallery columns (group arrays as
 * reid_stream_group_*) and its overlap list: true final keys of the group (gkey, gjunk), the sorted match keys mk, and
 * for every match the raw-distance thresholds thrA / thrB (largest D whose untouched final value is < / <= the match's).
 * group_max (here and in the two calls below): the longest identity group, as for reid_stream_group_sort.
 * reid_stream_count_panel_rr: second sweep over the q x g panels: bins[goff + b] += entries that come after exactly b
 * matches when ranked as if the Jaccard term did not touch them.
 * reid_stream_score_rr: corrects the bins by the special entries (same identity, overlap targets; od = their exact
 * distances gathered from the panels) and turns them into AP / first hit / CMC histogram (evaluate.py:53-80).
 * reid_rr_topk: exact top-K of every re-ranked row = merge(tidx/tval: its Kd nearest by raw distance, overlap list);
 * *ambiguous counts rows whose K-th value reaches the untouched bound of the Kd-th distance (enlarge Kd). */
REID_API int reid_lists_rekey(const int32_t* widx, const float* wval, int rows, int Kw, const int32_t* rowmax_key, int K,
                     int32_t* nbr, int32_t* ambiguous, void* stream);
REID_API int reid_krecip_V_lists(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf,
                        int rows, int k1, const int32_t* rowmax_key, const int32_t* wl_idx, const float* wl_val, int Kw,
                        const float* x32, int64_t ldx, const float* sq, int32_t* Vc, reid_half* Vv, int32_t* Vn,
                        int32_t* misses, const uint64_t* mt_keys, const float* mt_vals, int mt_n, uint64_t* miss_out,
                        int miss_cap, void* stream);
REID_API int reid_overlap_lists(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                       const int64_t* colptr, const uint64_t* csc_ent, int q0, int q_stride, int nq, int q, int N,
                       int cap_o, int nrange, int32_t* ocol, reid_half* ot, int32_t* ocnt, int32_t* overflow,
                       void* stream);
/* od[l * cap_o + e] = panel[l, ocol[l * cap_o + e] - c0] for the listed columns inside the panel's block [c0, c1) */
REID_API int reid_stream_list_gather(const float* panel, int64_t ld, int rows, int c0, int c1, int cap_o,
                            const int32_t* ocol, const int32_t* ocnt, float* od, void* stream);
/* ---- neighbour lists of the symmetric N x N matrix from its upper half (step A of the streamed re-ranking) ----
 * The caller contracts only the diagonal blocks (reid_dist_block mode 2 + reid_topk_rows_scatter: the initial lists and
 * thresholds tau[row] = Kw-th distance) and the rectangles right of them (mode 3: the mirrored row maxima come with it).
 * Rows of a rectangle [rows x cols] at (row0, c0) select from it as from any panel (reid_topk_rows_hinted); its COLUMNS --
 * rows c0 .. c0+cols of the matrix -- receive every element <= tau[column]:
 * reid_sym_push: chunk_min = the rectangle's 32-column chunk minima (pitch cm_pitch); candidates (D, global row) go to
 *   pbuf[col * cap ...], count[col] (block-local column; zero before the first call); cthr_key: scratch, ceil(cols/32).
 * reid_sym_fold: per column with candidates: running list (list_val / list_idx [N, Kw], ascending by (D, column); unused
 *   slots +inf) + candidates -> the Kw nearest; tau[c0 + col] lowered; count reset; *overflow = largest count > cap seen
 *   (the caller then falls back to the full sweep).
 * reid_sym_merge: nlists partial lists per row ([nlists, rows, Kw]; one per rank) -> the Kw nearest. */
REID_API int reid_sym_push(const float* panel, int64_t ld, int rows, int row0, int cols, int c0, const float* chunk_min,
                  int cm_pitch, const float* tau, int32_t* cthr_key, int32_t* count, uint64_t* pbuf, int cap, void* stream);
REID_API int reid_sym_fold(int cols, int c0, int Kw, float* list_val, int32_t* list_idx, float* tau, int32_t* count,
                  const uint64_t* pbuf, int cap, int32_t* overflow, void* stream);
REID_API int reid_sym_merge(int rows, int nlists, int Kw, const float* lists_val, const int32_t* lists_idx, float* out_val,
                   int32_t* out_idx, void* stream);
REID_API int reid_rr_prepare(int nq, int q0, int q_stride, const int32_t* glo, const int32_t* glen, const int64_t* goff,
                    const int32_t* list, const float* gd, const int64_t* q_cam, const int64_t* g_cam,
                    const int32_t* rowmax_key, float c1, float c2, int cap_o, const int32_t* ocol, const reid_half* ot,
                    const int32_t* ocnt, uint64_t* gkey, uint8_t* gjunk, uint64_t* mk, float* thrA, float* thrB,
                    int32_t* nmatch, int32_t* overflow, int group_max, void* stream);
REID_API int reid_stream_count_panel_rr(const float* panel, int64_t ld, int rows, int c0, int c1, const int64_t* goff,
                               const uint64_t* mk, const float* thrA, const float* thrB, const int32_t* nmatch,
                               int32_t* bins, int group_max, void* stream);
REID_API int reid_stream_score_rr(int nq, int q0, int q_stride, const int32_t* glo, const int32_t* glen,
                         const int64_t* goff, const int32_t* list, const float* gd, const uint64_t* gkey,
                         const uint8_t* gjunk, const uint64_t* mk, const float* thrA, const float* thrB,
                         const int32_t* nmatch, const int32_t* bins, int cap_o, const int32_t* ocol, const reid_half* ot,
                         const float* od, const int32_t* ocnt, const int64_t* q_pid, const int64_t* g_pid,
                         const int32_t* rowmax_key, float c1, float c2, int max_rank, double* ap, int32_t* first_hit,
                         int32_t* cmc_hist, int32_t* n_valid, int group_max, void* stream);
REID_API int reid_rr_topk(int nq, int q0, int q_stride, const int32_t* tidx, const float* tval, int Kd, int g, int cap_o,
                 const int32_t* ocol, const reid_half* ot, const float* od, const int32_t* ocnt,
                 const int32_t* rowmax_key, float c1, float c2, int K, int32_t* out_idx, float* out_val,
                 int32_t* ambiguous, void* stream);

/* ---- Sharded symmetric path (no reference function: the reference is single-process) ----------------
 * One process per GPU.  Wherever an entry point takes (row0, row_stride) / (q0, q_stride), a NEGATIVE stride selects
 * "snake block" ownership: world = -stride ranks own the 256-row tile rows of the symmetric N x N matrix in
 * boustrophedon order (tile rows 0..P-1 -> ranks 0..P-1, P..2P-1 -> ranks P-1..0, ...), rank = row0; local row l is row
 * (l & 255) of the rank's (l >> 8)-th tile row.  A positive stride is the cyclic-row ownership of the NCCL path.
 *
 * reid_dist_sharded: reranking.py:36-41 for a rank's share -- the upper-triangle tiles (tile row <= tile column) of the
 * tile rows it owns; `tiles` = (tile row, tile column) pairs in global tile coordinates (device, int32 x 2 x ntiles).
 * out[nloc, ld_out]: the rank's rows, upper-only form.  rowmax_key[n] (global rows, reid_rowmax_init first): maximum
 * over the part of every row this rank computed, direct or mirrored.  chunk_min[nloc, cm_pitch]: 32-column chunk minima
 * of the stored rows (valid from the diagonal block on).  chunk_min_m[n, cmm_pitch]: chunk minima of the mirrored tiles,
 * entry [i, 8 * t' + k] = minimum over rows 32k..32k+31 of the rank's t'-th tile row of column i (valid for the tile
 * rows owned above row i's). */
REID_API int reid_dist_sharded(const reid_half* hi, const reid_half* lo, int64_t ld, const float* sq, const uint32_t* amax,
                      int n, int dpad, const int32_t* tiles, int ntiles, int world, float* out, int64_t ld_out,
                      int nloc, int32_t* rowmax_key, float* chunk_min, int cm_pitch, float* chunk_min_m, int cmm_pitch,
                      void* stream);
/* K5 (reranking.py:46,48) over the sharded matrix, in scatter form.  tau: per row, the K-th smallest chunk minimum this
 * rank holds (ordered key; 0xffffffff = fewer than K chunks).  dstar: stats = world slots of slot_ints ints, slot p =
 * rank p's {rowmax_key[n], tau_key[n]}; writes the global row maxima (key_g[n], and key_l[local row] for owned rows), the raw-distance
 * thresholds, zeroed counters and the 32-row group thresholds (cthr_key[reid_chunk_pitch(n)]).  push: candidates
 * (raw D bits << 32 | column) of ALL rows from the rank's stored elements into pbuf[n, cap].  merge_local: per row, the
 * K best (O key << 32 | column) records -> inbox of the row's owner: peers[owner] + inbox_off, laid out
 * [world][nloc_max][K] uint64.  merge_final: per owned row, the world x K records sorted; the first K columns are the
 * row's neighbour list, written to the table at peers[p] + nbr_off ([n, K] int32, global rows) of EVERY rank. */
REID_API int reid_shard_tau(int n, int rank, int world, const float* cm_d, int pitch_d, const float* cm_m, int pitch_m,
                   int K, int32_t* tau_key, void* stream);
REID_API int reid_shard_dstar(int n, int rank, int world, const int32_t* stats, int64_t slot_ints, int32_t* key_g,
                     int32_t* key_l, float* dstar, int32_t* count, int32_t* cthr_key, void* stream);
REID_API int reid_shard_push(const float* D, int nloc, int rank, int world, int n, int64_t ld, const float* cm_d,
                    int pitch_d, const float* dstar, const int32_t* cthr_key, int32_t* count, void* pbuf, int cap,
                    void* stream);
REID_API int reid_shard_merge_local(const float* D, int64_t ld, int n, int rank, int world, const float* cm_d, int pitch_d,
                           const float* cm_m, int pitch_m, const int32_t* key_g, const float* dstar,
                           const int32_t* count, const void* pbuf, int cap, int K, void* const* peers,
                           int64_t inbox_off, int nloc_max, void* stream);
REID_API int reid_shard_merge_final(int nloc, int rank, int world, int K, const void* inbox, int nloc_max,
                           void* const* peers, int64_t nbr_off, float* nbr_val, void* stream);
/* K6b for a rank's own rows of the sharded matrix: entries left of a row's diagonal block are read from the panel of the
 * rank owning the mirror image (panels[world] device pointers, same pitch ld everywhere).  V rows are written at their
 * GLOBAL row index (Vc / Vv [n, cap1], Vn [n]); D and rowmax_key are indexed by local row. */
REID_API int reid_krecip_V_sharded(const int32_t* rfull, const int32_t* nfull, const int32_t* rhalf, const int32_t* nhalf,
                          int rank, int world, int rows, int k1, const float* D, int64_t ld,
                          const int32_t* rowmax_key, const void* const* panels, int32_t* Vc, reid_half* Vv, int32_t* Vn,
                          void* stream);
/* Peer-memory fabric: an arena per rank (cudaMalloc; export -> 64-byte CUDA IPC handle; open maps a peer's arena into
 * this process with peer access enabled), identical layout on every rank.  `peers` is a DEVICE array of world arena base
 * pointers (own arena at index rank).  barrier: a kernel on the stream -- stores `epoch` into flags[rank] of every peer
 * (release.sys) and waits until all of its own flags reached it; flags = 64 uint32 at flags_off of every arena,
 * zero-initialised; epochs grow by one per barrier.  epoch == 0: the kernel keeps the count itself (word 32 of the rank's
 * own flag block) -- no per-call argument, so a step with barriers can be captured in a CUDA graph.  bcast*: copy this rank's part of a replicated table into every
 * peer's arena at the same offset: a byte range; the owned snake blocks of an [n, row_bytes] table; owned V rows in ELL
 * form (count + used prefix); owned rows of a CSR pool (cnt[n], int32 columns / float16 values at element offsets off[]). */
REID_API int reid_fabric_alloc(int64_t bytes, void** out);
REID_API int reid_fabric_free(void* p);
REID_API int reid_fabric_export(void* p, void* handle64);
REID_API int reid_fabric_open(const void* handle64, void** out);
REID_API int reid_fabric_close(void* p);
REID_API int reid_fabric_barrier(void* const* peers, int64_t flags_off, int rank, int world, int epoch, void* stream);
REID_API int reid_fabric_bcast(void* const* peers, int rank, int world, int64_t offset, int64_t bytes, void* stream);
REID_API int reid_fabric_bcast_rows(void* const* peers, int rank, int world, int64_t offset, int row_bytes, int n,
                           void* stream);
REID_API int reid_fabric_bcast_ell(void* const* peers, int rank, int world, int64_t off_cnt, int64_t off_cols,
                          int64_t off_vals, int cap1, int n, void* stream);
REID_API int reid_fabric_bcast_pool(void* const* peers, int rank, int world, int64_t off_cnt, int64_t off_cols,
                           int64_t off_vals, const int64_t* off, int n, void* stream);

/* K7  reranking.py:73-78.  ub[l] = upper bound of |supp V_qe[row_l]|; exclusive scan to offsets;
 * then the expansion itself into a CSR pool (Qc/Qv at qoff[l], Qn[l] entries, columns ascending).
 * V (Vc/Vv/Vn, ELL pitch cap1) and nbr are global (all N rows). */
REID_API int reid_qe_bound(const int32_t* nbr, int K, int k2, const int32_t* Vn, int row0, int row_stride,
                  int rows, int32_t* ub, void* stream);
REID_API int reid_exclusive_scan(const int32_t* in, int n, int64_t* out /* n+1 */, void* stream);
REID_API int reid_qe_k2(const int32_t* nbr, int K, int k2, const int32_t* Vc, const reid_half* Vv,
               const int32_t* Vn, int cap1, int row0, int row_stride, int rows, const int64_t* qoff,
               int32_t* Qc, reid_half* Qv, int32_t* Qn, void* stream);
REID_API int reid_qe_k2_capped(const int32_t* nbr, int K, int k2, const int32_t* Vc, const reid_half* Vv,
                      const int32_t* Vn, int cap1, int row0, int row_stride, int rows,
                      const int64_t* qoff, int cap_sort, int32_t* Qc, reid_half* Qv, int32_t* Qn,
                      void* stream);
/* offsets of an ELL matrix viewed as CSR: off[i] = i * pitch, i in [0, N] (k2 == 1: V_qe = V) */
REID_API int reid_ell_offsets(int N, int pitch, int64_t* off, void* stream);

/* K8  reranking.py:80-82.  Inverted index (CSC) of V_qe over all N rows: count, (scan), fill.
 * csc_ent[p] = row | (uint64)float16_bits << 32 for p in [colptr[c], colptr[c+1]): one 8-byte entry per
 * non-zero so that a column is one contiguous, 8-byte aligned stream for cp.async. */
REID_API int reid_csc_count(const int32_t* Qc, const int64_t* qoff, const int32_t* Qn, int N,
                   int32_t* colcnt /* zeroed, N */, void* stream);
REID_API int reid_csc_fill(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                  int N, const int64_t* colptr, int32_t* cursor /* zeroed, N */, uint64_t* csc_ent,
                  void* stream);

/* Workspace of the sparse Jaccard path (reid_jaccard_mix / reid_jaccard_score with ws != NULL): one packed
 * overlap record per query -- bitmap of the touched probes, prefix popcounts, up to `cap` float16 overlaps --
 * plus the list of queries that touch more than `cap` probes (those are finished by the dense kernel, one
 * 2N-byte shared-memory row per query).  Returns 0 when the sparse path cannot be used for this N / cap;
 * pass ws = NULL then (dense path for every query). */
REID_API int64_t reid_jaccard_ws_bytes(int nq, int N, int cap);

/* K9 + K10  reranking.py:84-100.  For the query window (global probe q0 + l*q_stride, l < nq):
 * float16 Jaccard chain through the inverted index and
 *   out[l, g - q] = fp32(fp16(J * fp16(c1))) + (D[l, g] / rowmax[l]) * c2   for g in [q, N)
 * with c1 = 1 - lambda (rounded to float16 by the caller) and c2 = float32(lambda).
 * jaccard_out (nullable, pitch ld_j, N columns) receives the float16 Jaccard rows for tests. */
REID_API int reid_jaccard_mix(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                     const int64_t* colptr, const uint64_t* csc_ent,
                     const float* D, int64_t ld, const int32_t* rowmax_key, int q0, int q_stride,
                     int nq, int q, int N, float c1, float c2, float* out, int64_t ld_out,
                     reid_half* jaccard_out, int64_t ld_j, void* ws, int64_t ws_bytes, int cap, void* stream);

/* Identity index for the fused scorer  evaluate.py:43, :53-55.  Groups the gallery columns by pid (open-addressing
 * hash on the 64-bit id, then a counting sort by slot) so that a query finds the handful of gallery entries
 * sharing its identity -- the only ones that can be junk or true matches -- without scanning g_pid.
 * T = reid_pid_index_slots(g) (a power of two >= 2g).  Workspace: keys[T] 64-bit, cnt[T+1], cursor[T+1],
 * start[T+2] (int64), list[g]. */
REID_API int reid_pid_index_slots(int g);
REID_API int reid_pid_index(const int64_t* g_pid, int g, int T, uint64_t* keys, int32_t* cnt, int64_t* start,
                   int32_t* cursor, int32_t* list, void* stream);

/* K9 + K10 + K11 + K12 fused (the evaluator's path)  reranking.py:84-100 then evaluate.py:40-88.
 * reid_jaccard_mix plus, per query, AP / first-match rank / CMC histogram computed from the mixed row while it
 * is produced: the kept rank of every true match is the number of non-junk entries at or before it in stable
 * (distance, column) order.  `out` may be NULL: then the q x g re-ranked matrix is never written.
 * q_pid / q_cam: ids of this window's queries [nq]; g_cam [g]; keys/start/list/T from reid_pid_index.
 * Score outputs as in reid_score_rows. */
REID_API int reid_jaccard_score(const int32_t* Qc, const reid_half* Qv, const int64_t* qoff, const int32_t* Qn,
                       const int64_t* colptr, const uint64_t* csc_ent,
                       const float* D, int64_t ld, const int32_t* rowmax_key, int q0, int q_stride,
                       int nq, int q, int N, float c1, float c2, float* out, int64_t ld_out,
                       const int64_t* q_pid, const int64_t* q_cam, const int64_t* g_cam,
                       const uint64_t* pid_keys, const int64_t* pid_start, const int32_t* pid_list, int pid_T,
                       int max_rank, double* ap, int32_t* first_hit, int32_t* cmc_hist, int32_t* n_valid,
                       void* ws, int64_t ws_bytes, int cap, void* stream);

/* K11  evaluate.py:40.  Stable full ranking of every row (ascending distance, ties by column).
 * Workspace: keyA, keyB, idxA of rows*g 32-bit words each. */
REID_API int reid_rank_rows(const float* dist, int rows, int g, int64_t ld, uint32_t* keyA, uint32_t* keyB,
                   int32_t* idxA, int32_t* ranked, void* stream);

/* K12  evaluate.py:43-88.  Per-query AP (float64; -1 = query skipped), first-match rank, and
 * the histogram / counter the CMC curve is built from (cmc_hist[max_rank] and n_valid zeroed by
 * the caller; they are plain sums, so ranks all-reduce them). */
REID_API int reid_score_rows(const int32_t* ranked, int q, int g, const int64_t* q_pid, const int64_t* g_pid,
                    const int64_t* q_cam, const int64_t* g_cam, int max_rank, double* ap,
                    int32_t* first_hit, int32_t* cmc_hist, int32_t* n_valid, void* stream);
/* K11 + K12 fused (the evaluator's path)  evaluate.py:40-88.  Same outputs as reid_rank_rows +
 * reid_score_rows, computed from the distance rows by counting: the kept rank of every true match is
 * the number of non-junk entries before it in stable (distance, column) order; no ranking is
 * materialised.  dist: [q, >= g] float32, pitch ld. */
REID_API int reid_score_dist(const float* dist, int q, int g, int64_t ld, const int64_t* q_pid, const int64_t* g_pid,
                    const int64_t* q_cam, const int64_t* g_cam, int max_rank, double* ap, int32_t* first_hit,
                    int32_t* cmc_hist, int32_t* n_valid, void* stream);
REID_API int reid_score_finalize(const double* ap, int q, const int32_t* cmc_hist, int max_rank,
                        const int32_t* n_valid, float* cmc, double* mAP, void* stream);

/* ---- training-side distance ops: batch scale, float32 CUDA cores, fixed summation order ---- */
/* losses.py:14-28 `euclidean_dist`.  out[i, j] = sqrt(max(|x_i|^2 + |y_j|^2 - 2 x_i.y_j, 1e-12)); the squared norms are
 * accumulated in the same pass (no prepared operands, any d, any pitch >= d). */
REID_API int reid_pair_dist(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d, float* out,
                   int64_t ld_out, void* stream);
/* The same op for large batches (B >= ~1024, cross-batch memories) on the tensor cores: rows split into the fp16 hi/lo
 * operands of the evaluation path, contracted by the tcgen05 kernel of reid_dist (x.y to ~2^-22 |x||y|; upper triangle +
 * mirrored stores when x and y are the same rows: pass the same pointer, pitch and count), clamped square root in place.
 * workspace: reid_pair_dist_tc_workspace_bytes(m, n, d, same) bytes, 256-byte aligned; out 16-byte aligned with a pitch
 * that is a multiple of 4.  No host synchronisation: capturable. */
REID_API int64_t reid_pair_dist_tc_workspace_bytes(int m, int n, int d, int same);
REID_API int reid_pair_dist_tc(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d, void* workspace,
                      int64_t workspace_bytes, float* out, int64_t ld_out, void* stream);
/* losses.py:31-83 `hard_example_mining` (and the masks of losses.py:126-139).  Per anchor row i of dist[m, n]:
 * dist_ap / p_ind = the largest entry among columns with lab_c == lab_r[i] (the anchor's own column i is left out when
 * skip_diag), dist_an / n_ind = the smallest among the others; ties keep the lowest column.  A row without a positive
 * (negative) gets -inf (+inf) and index -1 where the reference's view(N, -1) raises.  p_ind / n_ind may be NULL. */
REID_API int reid_hard_mine(const float* dist, int64_t ld, int m, int n, const int64_t* lab_r, const int64_t* lab_c,
                   int skip_diag, float* dist_ap, float* dist_an, int64_t* p_ind, int64_t* n_ind, void* stream);
/* autograd of reid_hard_mine: G[i, j] = g_ap[i] [j == p_ind[i]] + g_an[i] [j == n_ind[i]], every element written. */
REID_API int reid_hard_mine_grad(const float* g_ap, const float* g_an, const int64_t* p_ind, const int64_t* n_ind, int m,
                        int n, float* G, int64_t ldg, void* stream);
/* autograd of reid_pair_dist (sqrt, clamp, matmul and both norm terms of losses.py:22-27):
 * gx_i = sum_j W_ij (x_i - y_j), gy_j (+)= sum_i W_ij (y_j - x_i), W = G / dist where dist > 1e-6 (clamp inactive), else 0.
 * gx or gy may be NULL; accumulate_gy adds into gy (x and y the same tensor: pass gy = gx). */
REID_API int reid_pair_dist_grad(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d,
                        const float* dist, int64_t ldd, const float* G, int64_t ldg, float* gx, int64_t ldgx,
                        float* gy, int64_t ldgy, int accumulate_gy, void* stream);
/* reid_hard_mine_grad + reid_pair_dist_grad taken together for losses that only use the mined pairs (losses.py:96-108,
 * :141-151): gx_i = wp_i (x_i - y_p) + wn_i (x_i - y_n) with wp = g_ap / dist_ap, wn = g_an / dist_an (0 where the clamp
 * was active or the index is -1); gy_j (+)= the same terms seen from the selected columns, anchors in ascending order.
 * O((m + n) d) instead of O(m n d); no float atomics. */
REID_API int reid_triplet_grad(const float* x, int64_t ldx, int m, const float* y, int64_t ldy, int n, int d,
                      const float* dist_ap, const float* dist_an, const int64_t* p_ind, const int64_t* n_ind,
                      const float* g_ap, const float* g_an, float* gx, int64_t ldgx, float* gy, int64_t ldgy,
                      int accumulate_gy, void* stream);

/* ---- CMC / mAP when the q x g matrix is never materialised (evaluate.py:40-88 semantics) ----
 * The caller produces the matrix panel by panel (reid_dist on row chunks x column blocks).
 * Setup: reid_pid_index over the gallery ids; reid_stream_group_lookup -> for every query the slice [glo, glo+glen) of
 * pid_list holding the gallery columns of its identity; goff = exclusive prefix sum of glen (caller).
 * Match distances gd[goff[i] + e] = D[i, pid_list[glo[i] + e]]: either reid_stream_group_gather on every panel of a
 *          first sweep, or (what the host layer does) reid_dist on the gathered operand rows of those pairs, diagonal
 *          kept -- the same bits, and the counting below can then share the ONE sweep with the top-k selection;
 *          reid_stream_group_sort: per query the matches' (distance, column) keys ascending in mkeys[goff .. goff+nmatch),
 *          the junk keys (same camera, evaluate.py:55) behind them.  group_max = the longest identity group (max of
 *          glen; 0 = "at most 1024") sizes the per-query shared-memory arrays: 2^k entries, 1024 <= 2^k <= 8192; a larger
 *          group_max is refused, and a query whose group exceeds what was announced is reported through *overflow
 *          (largest group seen; 0 = none) and skipped.  The same group_max goes to every call of one evaluation.
 * Sweep: reid_stream_count_panel on every panel adds to bins[goff[i] + b] the number of entries of the panel row with
 *          exactly b match keys below them (entries beyond the farthest match are skipped with one compare);
 *          bins zeroed by the caller.  rows/glo/glen/goff/nmatch pointers are those of the panel's first row.
 * Finally reid_stream_score: kept rank of every match = prefix sum of its bins minus the junk in front of it -> AP
 *          (float64, -1 = skipped), first-hit rank, CMC histogram and n_valid exactly like reid_score_dist. */
REID_API int reid_stream_group_lookup(const int64_t* q_pid, int nq, const uint64_t* pid_keys, const int64_t* pid_start,
                             int pid_T, int32_t* glo, int32_t* glen, void* stream);
REID_API int reid_stream_group_gather(const float* panel, int64_t ld, int rows, int c0, int c1, const int32_t* glo,
                             const int32_t* glen, const int64_t* goff, const int32_t* pid_list, float* gd, void* stream);
REID_API int reid_stream_group_sort(int nq, const int32_t* glo, const int32_t* glen, const int64_t* goff,
                           const int32_t* pid_list, const float* gd, const int64_t* q_cam, const int64_t* g_cam,
                           uint64_t* mkeys, int32_t* nmatch, int32_t* overflow, int group_max, void* stream);
REID_API int reid_stream_count_panel(const float* panel, int64_t ld, int rows, int c0, int c1, const int64_t* goff,
                            const uint64_t* mkeys, const int32_t* nmatch, int32_t* bins, int group_max, void* stream);
REID_API int reid_stream_score(int nq, const int32_t* glen, const int64_t* goff, const uint64_t* mkeys,
                      const int32_t* nmatch, const int32_t* bins, int max_rank, double* ap, int32_t* first_hit,
                      int32_t* cmc_hist, int32_t* n_valid, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* REID_B200_H */
