// K4 + K5 (reranking.py:46, :48): the reference divides the N x N matrix by its column maxima,
// transposes, and fully argsorts every row, but only ever reads the first max(k1+1, k2) columns
// of that argsort.  Here: per-row maximum (fused into the GEMM epilogue on the main path, or the
// stand-alone kernel below for a caller-supplied matrix), then an exact per-row top-K selection
// over O[i, j] = D[i, j] / rowmax[i] with stable (value, index) ordering.  HBM-bound: one read
// of the row from DRAM, a second one out of L2.
#include "common.cuh"
#include "kernels.h"

namespace reid {

// ---------------------------------------------------------------------------------------------
// row maximum of a pitched matrix -> ordered-int key (same encoding the GEMM epilogue uses)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rowmax_kernel(const float* __restrict__ D, int rows, int cols, long long ld, int* __restrict__ key) {
  const int row = blockIdx.x;
  const float* p = D + (long long)row * ld;
  float mx = -INFINITY;
  for (int c = threadIdx.x * 4; c < cols; c += blockDim.x * 4) {
    if (c + 3 < cols) {
      float4 v = *reinterpret_cast<const float4*>(p + c);
      mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    } else {
      for (int j = c; j < cols; ++j) mx = fmaxf(mx, p[j]);
    }
  }
  mx = warp_max(mx);
  __shared__ float sm[8];
  if (lane_id() == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : -INFINITY;
    v = warp_max(v);
    if (threadIdx.x == 0) key[row] = f2ordered_int(v);
  }
}

int launch_rowmax(const float* D, int rows, int cols, long long ld, int* key, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(ld % 4 == 0, "rowmax: pitch %lld must be a multiple of 4", ld);
  rowmax_kernel<<<rows, 256, 0, s>>>(D, rows, cols, ld, key);
  return check_launch("rowmax_kernel");
}

// ---------------------------------------------------------------------------------------------
// exact top-K (smallest) per row with (value, column) ordering
// ---------------------------------------------------------------------------------------------
constexpr int SEL_CAP = 4096;   // candidate slots (8 B each)

__device__ __forceinline__ unsigned long long make_ck(float d, float m, int col) {
  float o = __fadd_rn(__fdiv_rn(d, m), 0.0f);          // O = D / max ; -0.0 -> +0.0
  return ((unsigned long long)f2key(o) << 32) | (unsigned)col;
}

// block-wide bitonic sort of P (power of two) 64-bit keys in shared memory, ascending
template <int T>
__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* a, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += T) {
        // pair index t -> element i with bit j clear
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        unsigned long long x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncthreads();
    }
  }
}

// block-wide bitonic sort of P (power of two) 32-bit keys in shared memory, ascending
template <int T>
__device__ __forceinline__ void bitonic_sort_u32(uint32_t* a, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += T) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        uint32_t x = a[i], y = a[p];
        bool up = (i & k) == 0;
        if ((x > y) == up) { a[i] = y; a[p] = x; }
      }
      __syncthreads();
    }
  }
}

// The ranking key is O = D / rowmax (reranking.py:46), an IEEE division per element.  Division by a
// positive constant is monotone, so the candidate filter runs on the raw D values with plain float
// compares and only the survivors (a few dozen per row) are divided:
//   pass 1  every thread keeps the minimum D of its strided share; the K-th smallest of those T
//           minima, tau_D, bounds the K-th smallest element from above (at most K threads own
//           anything at or below it);
//   bound   D* = the largest float whose quotient does not exceed tau_D / m  (>= tau_D; quotients
//           of neighbouring floats can coincide, and entries tied in O are ordered by column);
//   pass 2  (row now L2-resident) collect every D <= D*, form (O, column) keys, sort, emit K.
template <int T>
__global__ void __launch_bounds__(T)
select_topk_kernel(const float* __restrict__ D, int rows, int cols, long long ld,
                   const int* __restrict__ rowmax_key, int K, int* __restrict__ nbr,
                   float* __restrict__ nbr_val, int cap_override) {
  __shared__ unsigned long long buf[SEL_CAP];
  __shared__ float dstar_s;
  __shared__ int cnt_s;
  const int row = blockIdx.x;
  const int tid = threadIdx.x;
  const float* p = D + (long long)row * ld;
  const float m = ordered_int2f(rowmax_key[row]);
  const int cap = cap_override > 0 ? min(cap_override, SEL_CAP) : SEL_CAP;
  const bool monotone = m > 0.0f && m < INFINITY;      // otherwise (degenerate rows) filter on O itself

  uint32_t* kbuf = reinterpret_cast<uint32_t*>(buf);
  if (monotone) {
    float best = INFINITY;
    for (int c = tid * 4; c < cols; c += T * 4) {
      if (c + 3 < cols) {
        const float4 v = *reinterpret_cast<const float4*>(p + c);
        best = fminf(best, fminf(fminf(v.x, v.y), fminf(v.z, v.w)));
      } else {
        for (int j = c; j < cols; ++j) best = fminf(best, p[j]);
      }
    }
    kbuf[tid] = f2key(best);
  } else {
    unsigned long long best = ~0ull;
    for (int c = tid; c < cols; c += T) {
      const unsigned long long k0 = make_ck(p[c], m, c);
      best = best < k0 ? best : k0;
    }
    kbuf[tid] = (uint32_t)(best >> 32);                 // O key
  }
  if (tid == 0) cnt_s = 0;
  __syncthreads();
  bitonic_sort_u32<T>(kbuf, T);
  if (tid == 0) {
    float x = key2f(kbuf[K - 1]);                       // tau_D (monotone) or tau_O
    if (monotone) {
      const float tau_o = __fdiv_rn(x, m);
      for (int it = 0; it < 64; ++it) {                 // a handful of steps: quotients of adjacent floats rarely coincide
        const float nx = x >= 0.0f ? __uint_as_float(__float_as_uint(x) + 1)
                                   : (__float_as_uint(x) == 0x80000000u ? __uint_as_float(1u)
                                                                        : __uint_as_float(__float_as_uint(x) - 1));
        if (!(__fdiv_rn(nx, m) <= tau_o)) break;
        x = nx;
      }
    }
    dstar_s = x;
  }
  __syncthreads();
  const float dstar = dstar_s;
  __syncthreads();

  // pass 2: collect everything at or below the bound
  for (int c = tid * 4; c < cols; c += T * 4) {
    float vv[4];
    int nv = 4;
    if (c + 3 < cols) {
      const float4 v = *reinterpret_cast<const float4*>(p + c);
      vv[0] = v.x; vv[1] = v.y; vv[2] = v.z; vv[3] = v.w;
    } else {
      nv = cols - c;
      for (int j = 0; j < nv; ++j) vv[j] = p[c + j];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j >= nv) break;
      bool take;
      unsigned long long k0 = 0;
      if (monotone) {
        take = vv[j] <= dstar;
        if (take) k0 = make_ck(vv[j], m, c + j);
      } else {
        k0 = make_ck(vv[j], m, c + j);
        take = (uint32_t)(k0 >> 32) <= f2key(dstar);
      }
      if (take) {
        const int pos = atomicAdd(&cnt_s, 1);
        if (pos < cap) buf[pos] = k0;
      }
    }
  }
  __syncthreads();
  const int cnt = cnt_s;
  if (cnt <= cap) {
    int P = 32;
    while (P < cnt) P <<= 1;
    for (int t = cnt + tid; t < P; t += T) buf[t] = ~0ull;
    __syncthreads();
    bitonic_sort_u64<T>(buf, P);
    for (int t = tid; t < K; t += T) {
      unsigned long long k0 = buf[t];
      nbr[(long long)row * K + t] = (int)(unsigned)(k0 & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + t] = key2f((uint32_t)(k0 >> 32));
    }
    return;
  }

  // overflow (adversarial layouts / massive ties): K rounds of block-wide arg-min, each strictly
  // above the previous pick.  Slow, exact, essentially never taken.
  __shared__ unsigned long long red[T / 32];
  unsigned long long last = 0;
  for (int r = 0; r < K; ++r) {
    unsigned long long mine = ~0ull;
    for (int c = tid; c < cols; c += T) {
      unsigned long long k0 = make_ck(p[c], m, c);
      if ((r == 0 || k0 > last) && k0 < mine) mine = k0;
    }
    mine = warp_min_u64(mine);
    if ((tid & 31) == 0) red[tid >> 5] = mine;
    __syncthreads();
    if (tid < 32) {
      unsigned long long v = tid < (T / 32) ? red[tid] : ~0ull;
      v = warp_min_u64(v);
      if (tid == 0) red[0] = v;
    }
    __syncthreads();
    last = red[0];
    __syncthreads();
    if (tid == 0) {
      nbr[(long long)row * K + r] = (int)(unsigned)(last & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + r] = key2f((uint32_t)(last >> 32));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Top-K with chunk minima.  The tensor-core epilogue (dist_tc2.cu) records the minimum of every
// 32-column chunk of D.  K chunks whose minimum is <= tau hold K elements <= tau, so the K-th smallest of
// ANY K-or-more chunk minima bounds the K-th smallest element.  One warp per row:
//   pass 1  every lane streams its share of the row's chunk minima (coalesced, unrolled) and keeps its S
//           smallest in registers (32*S >= K); tau = the K-th smallest of those 32*S values, found by a
//           bitwise search with ballots on registers.  (The exact K-th smallest chunk minimum would be a
//           slightly tighter bound -- ~K qualifying chunks instead of ~1.6 K -- at the price of a shared-memory
//           copy of the row's minima and 32 passes over it.)
//   bound   D* = the largest float whose quotient by the row maximum does not exceed tau / m.
//   pass 2  chunk minima again (L1/L2 hits): chunks with minimum <= D* are listed, their 128-byte lines fetched
//           eight at a time, entries <= D* become (O, column) keys, sorted, first K emitted.
// The kernel reads N/32 minima + ~1.6 K lines per row instead of the row.
// ---------------------------------------------------------------------------------------------
constexpr int SELH_CAND = 256;      // candidate slots per warp
constexpr int SELH_LIST = 512;      // qualifying-chunk slots per warp
constexpr int SELH_WPB = 8;

__device__ __forceinline__ void warp_bitonic_u64_sel(unsigned long long* a, int P, int lane) {
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

// Pass 1 of the hinted selection, split into pieces so that the sharded selection (a rank sees only part of a row's
// chunk minima, in two arrays) can reuse them:
//   keep_chunks   folds chunk minima cm[c_begin .. c_end) into the lane's S smallest, ascending in keep[]
//   keep_kth      K-th smallest of the warp's 32*S kept values as an ordered key (0xffffffff when fewer than K chunks
//                 were offered: no bound)
//   dstar_of_tau  the raw-D threshold D* >= tau such that the row's K nearest in O = D / m order all have D <= D*
//                 (quotients of neighbouring floats can coincide); INFINITY when no monotone filter applies
template <int S>
__device__ __forceinline__ void keep_chunks(const float* __restrict__ cm, int c_begin, int c_end, float (&keep)[S]) {
  const int lane = lane_id();
  // lane l takes chunks l, l + 32, ...: neighbours cluster in adjacent chunks when the gallery is ordered by identity,
  // and a lane keeps only S values, so adjacent chunks must go to different lanes for the bound to stay tight
  for (int c = c_begin + lane; c < c_end; c += 8 * 32) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = c + u * 32 < c_end ? __ldg(cm + c + u * 32) : INFINITY;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      float x = v[u];
#pragma unroll
      for (int s = 0; s < S; ++s) { const float lo = fminf(keep[s], x); x = fmaxf(keep[s], x); keep[s] = lo; }
    }
  }
}

template <int S>
__device__ __forceinline__ uint32_t keep_kth(const float (&keep)[S], int nch, int K) {
  // K-th smallest of the 32*S kept values: largest prefix t with #{key < t} < K, built bit by bit
  uint32_t t = 0xffffffffu;                               // fewer chunks than K: every chunk is a candidate
  if (nch >= K) {
    uint32_t kk[S];
#pragma unroll
    for (int s = 0; s < S; ++s) kk[s] = f2key(keep[s]);
    t = 0;
    for (int bit = 31; bit >= 0; --bit) {
      const uint32_t trial = t | (1u << bit);
      int cnt = 0;
#pragma unroll
      for (int s = 0; s < S; ++s) cnt += __popc(__ballot_sync(FULL, kk[s] < trial));
      if (cnt < K) t = trial;                             // fewer than K keys below trial: the K-th is >= trial
    }
  }
  return t;
}

__device__ __forceinline__ float dstar_of_tau(uint32_t tau_key, float m) {
  float dstar = key2f(tau_key);
  if (tau_key != 0xffffffffu && m > 0.0f && m < INFINITY && dstar == dstar) {
    const float tau_o = __fdiv_rn(dstar, m);
    for (int it = 0; it < 64; ++it) {
      const float x = dstar;
      const float nx = x >= 0.0f ? __uint_as_float(__float_as_uint(x) + 1)
                                 : (__float_as_uint(x) == 0x80000000u ? __uint_as_float(1u)
                                                                      : __uint_as_float(__float_as_uint(x) - 1));
      if (!(__fdiv_rn(nx, m) <= tau_o)) break;
      dstar = nx;
    }
  } else {
    dstar = INFINITY;                                   // degenerate maximum: no monotone filter, take everything
  }
  return dstar;
}

// Pass 1 + bound for one row whose chunk minima lie in one array (one warp)
template <int S>
__device__ __forceinline__ float row_dstar(const float* __restrict__ cm, int nch, int K, float m) {
  float keep[S];
#pragma unroll
  for (int s = 0; s < S; ++s) keep[s] = INFINITY;
  keep_chunks<S>(cm, 0, nch, keep);
  return dstar_of_tau(keep_kth<S>(keep, nch, K), m);
}

template <int S>
__global__ void __launch_bounds__(SELH_WPB * 32)
select_topk_hinted_kernel(const float* __restrict__ D, int rows, int cols, long long ld,
                          const int* __restrict__ rowmax_key, const float* __restrict__ chunk_min, int nch,
                          int cm_pitch, int K, int* __restrict__ nbr, float* __restrict__ nbr_val, int cap, int list_cap,
                          int upper_only) {
  __shared__ unsigned long long cand_s[SELH_WPB][SELH_CAND];
  __shared__ int list_s[SELH_WPB][SELH_LIST];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int row = blockIdx.x * SELH_WPB + w;
  if (row >= rows) return;
  unsigned long long* cand = cand_s[w];
  int* list = list_s[w];
  const float* p = D + (long long)row * ld;
  // upper_only: D is the symmetric N x N matrix with only the tiles on and above the 256-wide diagonal blocks stored;
  // an entry left of this row's diagonal block is read from its mirror image D[col, row] (bit-identical by construction)
  const int tri0 = upper_only ? (row & ~255) : 0;
  auto ldD = [&](int col) { return col < tri0 ? __ldg(D + (long long)col * ld + row) : __ldg(p + col); };
  const float* cm = chunk_min + (long long)row * cm_pitch;
  const float m = ordered_int2f(rowmax_key[row]);

  const float dstar = row_dstar<S>(cm, nch, K, m);

  // pass 2a: list the qualifying chunks
  int nlist = 0;
  bool overflow = false;
  for (int c0 = 0; c0 < nch; c0 += 32) {
    const int c = c0 + lane;
    const bool take = c < nch && (__ldg(cm + c) <= dstar || dstar == INFINITY);
    const unsigned b = __ballot_sync(FULL, take);
    const int n_take = __popc(b);
    if (nlist + n_take > list_cap) { overflow = true; break; }
    if (take) list[nlist + __popc(b & ((1u << lane) - 1))] = c;
    nlist += n_take;
  }
  __syncwarp();
  // pass 2b: fetch their lines eight at a time, keep entries <= D*
  int cnt = 0;
  for (int l0 = 0; l0 < nlist && !overflow; l0 += 8) {
    float v[8];
    int col[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      col[u] = l0 + u < nlist ? list[l0 + u] * 32 + lane : cols;
      v[u] = col[u] < cols ? ldD(col[u]) : INFINITY;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const bool take = col[u] < cols && (v[u] <= dstar || dstar == INFINITY);
      const unsigned bt = __ballot_sync(FULL, take);
      const int n_take = __popc(bt);
      if (n_take == 0) continue;
      if (cnt + n_take > cap) { overflow = true; break; }
      if (take) cand[cnt + __popc(bt & ((1u << lane) - 1))] = make_ck(v[u], m, col[u]);
      cnt += n_take;
    }
  }
  __syncwarp();
  if (!overflow) {
    int P = 32;
    while (P < cnt) P <<= 1;
    for (int i = cnt + lane; i < P; i += 32) cand[i] = ~0ull;
    __syncwarp();
    warp_bitonic_u64_sel(cand, P, lane);
    for (int i = lane; i < K; i += 32) {
      const unsigned long long k0 = cand[i];
      nbr[(long long)row * K + i] = (int)(unsigned)(k0 & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + i] = key2f((uint32_t)(k0 >> 32));
    }
    return;
  }
  // overflow (massive ties): K rounds of warp-wide arg-min over the qualifying chunks, each strictly
  // above the previous pick.  Slow, exact, essentially never taken.
  unsigned long long last = 0;
  for (int r = 0; r < K; ++r) {
    unsigned long long mine = ~0ull;
    for (int cc = 0; cc < nch; ++cc) {
      if (!(__ldg(cm + cc) <= dstar || dstar == INFINITY)) continue;
      const int col = cc * 32 + lane;
      if (col < cols) {
        const unsigned long long k0 = make_ck(ldD(col), m, col);
        if ((r == 0 || k0 > last) && k0 < mine) mine = k0;
      }
    }
    mine = warp_min_u64(mine);
    last = mine;
    if (lane == 0) {
      nbr[(long long)row * K + r] = (int)(unsigned)(last & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + r] = key2f((uint32_t)(last >> 32));
    }
  }
}

int launch_select_topk_hinted(const float* D, int rows, int cols, long long ld, const int* rowmax_key,
                              const float* chunk_min, int K, int* nbr, float* nbr_val, int cap_override,
                              int upper_only, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= 256, "select_topk: K=%d outside [1,256]", K);
  REID_REQUIRE(cols >= K, "select_topk: row length %d is shorter than K=%d", cols, K);
  REID_REQUIRE(!upper_only || rows == cols, "select_topk: upper_only needs the square N x N matrix (%d x %d)", rows, cols);
  REID_REQUIRE(((uintptr_t)chunk_min & 15) == 0, "select_topk: chunk_min must be 16-byte aligned");
  const int nch = cdiv(cols, 32);
  const int cap = cap_override > 0 ? (cap_override < SELH_CAND ? cap_override : SELH_CAND) : SELH_CAND;
  const int list_cap = cap_override > 0 ? (cap_override < SELH_LIST ? cap_override : SELH_LIST) : SELH_LIST;
  REID_REQUIRE(cap >= K || cap_override > 0, "select_topk: K=%d exceeds the candidate buffer", K);
  const int grid = cdiv(rows, SELH_WPB), threads = SELH_WPB * 32;
#define REID_SELH(SS)                                                                                          \
  select_topk_hinted_kernel<SS><<<grid, threads, 0, s>>>(D, rows, cols, ld, rowmax_key, chunk_min, nch, chunk_pitch(cols), K, nbr, \
                                                         nbr_val, cap, list_cap, upper_only)
  // the warp keeps 32 * S chunk minima; their K-th smallest is the bound, so 32 * S must exceed K comfortably: at
  // K == 32 * S the "K-th smallest kept" is the LARGEST kept value -- a useless bound (measured: K = 64 with S = 2
  // sent most rows to the exact arg-min fallback, 41 s instead of 0.6 s over a 1.1 M x 1.1 M panel sweep)
  if (K <= 16) REID_SELH(1);
  else if (K <= 40) REID_SELH(2);
  else if (K <= 80) REID_SELH(4);
  else if (K <= 160) REID_SELH(8);
  else REID_SELH(16);
#undef REID_SELH
  return check_launch("select_topk_hinted_kernel");
}

// ---------------------------------------------------------------------------------------------
// Scatter form of the hinted selection for the upper-only symmetric matrix (reid_dist sym_mode 2): only the
// tiles on and above the 256-wide diagonal blocks of D are stored.  Reading the missing left part of a row through
// its mirror image is a sector gather (32 B fetched per 4 B used); instead every stored element is offered to BOTH
// rows it belongs to:
//   select_thr_kernel    one warp per row: D*[row] from the chunk minima (which the contraction records for the whole
//                        row, mirrored part included); the largest D* of every 32-row group -> cthr[group]
//   select_push_kernel   one warp per row c over its stored chunks (right of the diagonal block's left edge): a chunk
//                        is read if its minimum can matter to row c (<= D*[c]) or to one of its 32 columns
//                        (<= cthr[chunk]); element (c, j) becomes a candidate of row c if <= D*[c], and of row j
//                        (as column c) if j lies right of c's diagonal block and the value is <= D*[j]
//   select_merge_kernel  one warp per row: (O, column) keys of the row's candidates, sorted, first K emitted
// Candidate records are (raw D bits, column); their arrival order does not matter (the sort is total).  A row with
// more candidates than slots takes the exact arg-min rounds with mirror reads, as in the pull form.
// ---------------------------------------------------------------------------------------------
constexpr int SELP_LIST = 1024;     // listed chunks per warp (16-bit entries)
constexpr int SELP_STEP = 4 * 128;  // chunk minima examined per listing step: four float4 per lane

// threshold in the ordered-int domain; an unbounded row (D* = +inf) admits everything, NaN minima included
__device__ __forceinline__ int thr_key(float t) { return t == INFINITY ? 0x7fffffff : f2ordered_int(t); }


template <int S>
__global__ void __launch_bounds__(SELH_WPB * 32)
select_thr_kernel(int rows, const int* __restrict__ rowmax_key, const float* __restrict__ chunk_min, int nch,
                  int cm_pitch, int K, float* __restrict__ dstar_out, int* __restrict__ cthr_key,
                  int* __restrict__ count) {
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int row = blockIdx.x * SELH_WPB + w;
  if (row >= rows) return;
  const float m = ordered_int2f(rowmax_key[row]);
  const float ds = row_dstar<S>(chunk_min + (long long)row * cm_pitch, nch, K, m);
  if (lane == 0) {
    dstar_out[row] = ds;
    count[row] = 0;
    atomicMax(cthr_key + (row >> 5), thr_key(ds));
  }
}

__global__ void __launch_bounds__(SELH_WPB * 32, 5)
select_push_kernel(const float* __restrict__ D, int lrows, int row0, int row_stride, int n, long long ld,
                   const float* __restrict__ chunk_min, int nch, int cm_pitch, const float* __restrict__ dstar,
                   const int* __restrict__ cthr_key, int* __restrict__ count, unsigned long long* __restrict__ pbuf,
                   int cap, int pf) {
  __shared__ unsigned short list_s[SELH_WPB][SELP_LIST];
  __shared__ unsigned long long own_s[SELH_WPB][SELH_CAND];
  __shared__ int nlist_s[SELH_WPB];
  const int w = threadIdx.x >> 5, lane = lane_id();
  // D / chunk_min hold the `lrows` rows this rank stores (all n on one rank); thresholds, counters and candidate
  // records are indexed by GLOBAL row, so a rank of a sharded job pushes into the lists of rows it does not own
  const int lrow = blockIdx.x * SELH_WPB + w;
  if (lrow >= lrows) return;
  const int row = (int)own_row(row0, row_stride, lrow);
  unsigned short* list = list_s[w];
  unsigned long long* own = own_s[w];         // this row's own candidates: staged, appended with one atomic at the end
  const float* p = D + (long long)lrow * ld;
  const float* cm = chunk_min + (long long)lrow * cm_pitch;
  const float ds = dstar[row];
  const int dsk = thr_key(ds);
  const int tri0 = row & ~255;                // first stored column of this row
  const int push0 = tri0 + 256;               // columns from here on have this row in their unstored part
  const int pc0 = push0 >> 5;                 // first chunk whose columns take pushes (a multiple of 8)
  int nown = 0;
  if (lane == 0) nlist_s[w] = 0;
  __syncwarp();

  int c0 = tri0 >> 5;                         // a multiple of 8: float4 loads of chunk minima / thresholds are aligned
  while (c0 < nch) {
    // list the chunks that can matter to this row or to one of their columns: four consecutive chunks per lane and
    // load, four loads in flight; order within the list is irrelevant (the merge sorts)
    int nlist = 0;
    for (; c0 < nch && nlist <= SELP_LIST - SELP_STEP; c0 += SELP_STEP) {
      float4 cmv[4];
      int4 th[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c4 = c0 + (u * 32 + lane) * 4;
        cmv[u] = c4 < nch ? __ldg(reinterpret_cast<const float4*>(cm + c4)) : make_float4(0.f, 0.f, 0.f, 0.f);
        th[u] = (c4 < nch && c4 >= pc0) ? __ldg(reinterpret_cast<const int4*>(cthr_key + c4))
                                         : make_int4(dsk, dsk, dsk, dsk);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c4 = c0 + (u * 32 + lane) * 4;
        unsigned h = (f2ordered_int(cmv[u].x) <= max(dsk, th[u].x) ? 1u : 0u) |
                     (f2ordered_int(cmv[u].y) <= max(dsk, th[u].y) ? 2u : 0u) |
                     (f2ordered_int(cmv[u].z) <= max(dsk, th[u].z) ? 4u : 0u) |
                     (f2ordered_int(cmv[u].w) <= max(dsk, th[u].w) ? 8u : 0u);
        const int left = nch - c4;            // chunks of this float4 inside the row
        h = left >= 4 ? h : (left > 0 ? h & ((1u << left) - 1) : 0u);
        if (h) {
          int at = atomicAdd(&nlist_s[w], __popc(h));
          while (h) {
            const int c = c4 + __ffs(h) - 1;
            h &= h - 1;
            list[at++] = (unsigned short)c;
            // the chunk's line is needed a few microseconds from now: start the fetch
            if (pf == 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + (long long)c * 32));
            else if (pf == 2) asm volatile("prefetch.global.L1 [%0];" ::"l"(p + (long long)c * 32));
          }
        }
      }
      __syncwarp();
      nlist = nlist_s[w];
    }
    __syncwarp();
    if (lane == 0) nlist_s[w] = 0;
    __syncwarp();
    // fetch the listed lines eight at a time
    for (int l0 = 0; l0 < nlist; l0 += 8) {
      float v[8], tj[8];
      int col[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        col[u] = l0 + u < nlist ? list[l0 + u] * 32 + lane : n;
        const bool in = col[u] < n;
        v[u] = in ? __ldg(p + col[u]) : INFINITY;
        tj[u] = (in && col[u] >= push0) ? __ldg(dstar + col[u]) : -INFINITY;
      }
      // pushes first: independent atomics, all in flight before the first returned slot is used
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const bool push = col[u] < n && col[u] >= push0 && (v[u] <= tj[u] || tj[u] == INFINITY);
        tj[u] = __int_as_float(push ? atomicAdd(count + col[u], 1) : cap);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const bool take = col[u] < n && (v[u] <= ds || ds == INFINITY);
        const unsigned bt = __ballot_sync(FULL, take);
        const int at = nown + __popc(bt & ((1u << lane) - 1));
        if (take && at < cap) own[at] = ((unsigned long long)__float_as_uint(v[u]) << 32) | (unsigned)col[u];
        nown += __popc(bt);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int pos = __float_as_int(tj[u]);
        if (pos < cap)
          pbuf[(long long)col[u] * cap + pos] = ((unsigned long long)__float_as_uint(v[u]) << 32) | (unsigned)row;
      }
    }
    __syncwarp();
  }
  // append the staged candidates to this row's records (other rows' pushes land in the same list concurrently)
  if (nown > 0) {
    int base = 0;
    if (lane == 0) base = atomicAdd(count + row, nown);
    base = __shfl_sync(FULL, base, 0);
    if (nown <= cap) {
      unsigned long long* dst = pbuf + (long long)row * cap;
      for (int i = lane; i < nown; i += 32)
        if (base + i < cap) dst[base + i] = own[i];
    }
  }
}

__global__ void __launch_bounds__(SELH_WPB * 32)
select_merge_kernel(const float* __restrict__ D, int n, long long ld, const int* __restrict__ rowmax_key,
                    const float* __restrict__ chunk_min, int nch, int cm_pitch, const float* __restrict__ dstar_in,
                    const int* __restrict__ count, const unsigned long long* __restrict__ pbuf, int cap, int K,
                    int* __restrict__ nbr, float* __restrict__ nbr_val) {
  __shared__ unsigned long long cand_s[SELH_WPB][SELH_CAND];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int row = blockIdx.x * SELH_WPB + w;
  if (row >= n) return;
  unsigned long long* cand = cand_s[w];
  const float m = ordered_int2f(rowmax_key[row]);
  const int cnt = count[row];
  if (cnt <= cap && cnt >= K) {
    const unsigned long long* src = pbuf + (long long)row * cap;
    for (int i = lane; i < cnt; i += 32) {
      const unsigned long long e = src[i];
      cand[i] = make_ck(__uint_as_float((unsigned)(e >> 32)), m, (int)(unsigned)(e & 0xffffffffu));
    }
    int P = 32;
    while (P < cnt) P <<= 1;
    for (int i = cnt + lane; i < P; i += 32) cand[i] = ~0ull;
    __syncwarp();
    warp_bitonic_u64_sel(cand, P, lane);
    for (int i = lane; i < K; i += 32) {
      const unsigned long long k0 = cand[i];
      nbr[(long long)row * K + i] = (int)(unsigned)(k0 & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + i] = key2f((uint32_t)(k0 >> 32));
    }
    return;
  }
  // more candidates than slots (massive ties): K rounds of warp-wide arg-min, each strictly above the previous pick
  const float* p = D + (long long)row * ld;
  const float* cm = chunk_min + (long long)row * cm_pitch;
  const float dstar = dstar_in[row];
  const int tri0 = row & ~255;
  auto ldD = [&](int col) { return col < tri0 ? __ldg(D + (long long)col * ld + row) : __ldg(p + col); };
  unsigned long long last = 0;
  for (int r = 0; r < K; ++r) {
    unsigned long long mine = ~0ull;
    for (int cc = 0; cc < nch; ++cc) {
      if (!(__ldg(cm + cc) <= dstar || dstar == INFINITY)) continue;
      const int col = cc * 32 + lane;
      if (col < n) {
        const unsigned long long k0 = make_ck(ldD(col), m, col);
        if ((r == 0 || k0 > last) && k0 < mine) mine = k0;
      }
    }
    mine = warp_min_u64(mine);
    last = mine;
    if (lane == 0) {
      nbr[(long long)row * K + r] = (int)(unsigned)(last & 0xffffffffu);
      if (nbr_val) nbr_val[(long long)row * K + r] = key2f((uint32_t)(last >> 32));
    }
  }
}

// workspace: D*[n] f32 | count[n] i32 | (16-byte aligned) cthr[chunk_pitch(n)] i32 | (16-byte aligned) records[n * cap] u64
static inline size_t scatter_align16(size_t x) { return (x + 15) & ~(size_t)15; }
static inline size_t scatter_cthr_offset(int n) { return scatter_align16((size_t)8 * n); }
static inline size_t scatter_records_offset(int n) {
  return scatter_align16(scatter_cthr_offset(n) + (size_t)4 * chunk_pitch(n));
}
size_t select_scatter_workspace_bytes(int n, int cap_override) {
  const int cap = cap_override > 0 ? (cap_override < SELH_CAND ? cap_override : SELH_CAND) : SELH_CAND;
  return scatter_records_offset(n) + (size_t)8 * n * cap;
}

int launch_select_topk_scatter(const float* D, int n, long long ld, const int* rowmax_key, const float* chunk_min,
                               int K, int* nbr, float* nbr_val, int cap_override, void* workspace,
                               size_t workspace_bytes, cudaStream_t s) {
  if (n == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= 256, "select_topk: K=%d outside [1,256]", K);
  REID_REQUIRE(n >= K, "select_topk: row length %d is shorter than K=%d", n, K);
  const int cap = cap_override > 0 ? (cap_override < SELH_CAND ? cap_override : SELH_CAND) : SELH_CAND;
  REID_REQUIRE(cap >= K || cap_override > 0, "select_topk: K=%d exceeds the candidate buffer", K);
  REID_REQUIRE(workspace != nullptr && workspace_bytes >= select_scatter_workspace_bytes(n, cap_override),
               "select_topk_scatter: workspace of %zu bytes, %zu needed", workspace_bytes,
               select_scatter_workspace_bytes(n, cap_override));
  REID_REQUIRE(((uintptr_t)workspace & 15) == 0, "select_topk_scatter: workspace must be 16-byte aligned");
  REID_REQUIRE(((uintptr_t)chunk_min & 15) == 0, "select_topk_scatter: chunk_min must be 16-byte aligned");
  const int nch = cdiv(n, 32), pitch = chunk_pitch(n);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  float* dstar = reinterpret_cast<float*>(ws);
  int* count = reinterpret_cast<int*>(ws + (size_t)4 * n);
  int* cthr = reinterpret_cast<int*>(ws + scatter_cthr_offset(n));
  unsigned long long* pbuf = reinterpret_cast<unsigned long long*>(ws + scatter_records_offset(n));
  REID_REQUIRE(nch <= 65535, "select_topk_scatter: %d rows exceed the 16-bit chunk list", n);
  REID_CUDA(cudaMemsetAsync(cthr, 0x80, (size_t)4 * pitch, s));      // below every ordered float key
  const int grid = cdiv(n, SELH_WPB), threads = SELH_WPB * 32;
#define REID_SELT(SS) \
  select_thr_kernel<SS><<<grid, threads, 0, s>>>(n, rowmax_key, chunk_min, nch, pitch, K, dstar, cthr, count)
  // the warp keeps 32 * S chunk minima; their K-th smallest is the bound, so 32 * S must exceed K comfortably: at
  // K == 32 * S the "K-th smallest kept" is the LARGEST kept value -- a useless bound (measured: K = 64 with S = 2
  // sent most rows to the exact arg-min fallback, 41 s instead of 0.6 s over a 1.1 M x 1.1 M panel sweep)
  if (K <= 16) REID_SELT(1);
  else if (K <= 40) REID_SELT(2);
  else if (K <= 80) REID_SELT(4);
  else if (K <= 160) REID_SELT(8);
  else REID_SELT(16);
#undef REID_SELT
  if (int rc = check_launch("select_thr_kernel")) return rc;
  const char* pfe = getenv("REID_SEL_PF");
  select_push_kernel<<<grid, threads, 0, s>>>(D, n, 0, 1, n, ld, chunk_min, nch, pitch, dstar, cthr, count, pbuf, cap,
                                              pfe ? atoi(pfe) : 1);
  if (int rc = check_launch("select_push_kernel")) return rc;
  select_merge_kernel<<<grid, threads, 0, s>>>(D, n, ld, rowmax_key, chunk_min, nch, pitch, dstar, count, pbuf, cap, K,
                                               nbr, nbr_val);
  return check_launch("select_merge_kernel");
}


// ---------------------------------------------------------------------------------------------
// Sharded scatter selection (multi-rank symmetric contraction, launch_dist_tc_sharded).  Rank r holds of row i
//   * the direct part   D_r[local(i), j], j >= first column of i's diagonal block      -- only if r owns i's tile row
//   * the mirrored part D_r[local(j), i] for the rows j of every tile row r owns ABOVE i's
// and the matching chunk minima (cm_d suffix / cm_m compact prefix).  Every element of the symmetric matrix is computed by
// exactly one rank and offered to both rows it belongs to, so the union over ranks of the per-rank parts is the whole row.
//   shard_tau     per row: K-th smallest of the chunk minima this rank holds (an upper bound of the row's K-th element)
//   -- exchange: row maxima (max over ranks) and tau (min over ranks), 8 bytes per row and rank --
//   shard_dstar   D*[i] from the global tau and maximum (same rule as one rank), group thresholds, counters
//   select_push   as on one rank, over the rank's stored rows; candidates of ALL rows accumulate locally
//   merge_local   per row: sort this rank's candidates, its K best (O key, column) records go to the row owner's inbox
//   -- barrier --
//   merge_final   per owned row: the P x K records, sorted; first K = the row's neighbour list, written to every rank
// The (O key, column) order is total, so the lists do not depend on P or on arrival order: bit-identical to one rank.
// ---------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(SELH_WPB * 32)
shard_tau_kernel(int n, int rank, int P, const float* __restrict__ cm_d, int pitch_d, const float* __restrict__ cm_m,
                 int pitch_m, int K, unsigned* __restrict__ tau_key) {
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int i = blockIdx.x * SELH_WPB + w;
  if (i >= n) return;
  const int ti = i >> 8;
  const int npre = 8 * snake_tiles_below(ti, rank, P);
  float keep[S];
#pragma unroll
  for (int s = 0; s < S; ++s) keep[s] = INFINITY;
  keep_chunks<S>(cm_m + (long long)i * pitch_m, 0, npre, keep);
  int nch = npre;
  if (snake_owner(ti, P) == rank) {
    const int nchunks = (n + 31) >> 5;
    keep_chunks<S>(cm_d + (long long)snake_local_row(i, P) * pitch_d, ti * 8, nchunks, keep);
    nch += nchunks - ti * 8;
  }
  const uint32_t t = keep_kth<S>(keep, nch, K);
  if (lane == 0) tau_key[i] = t;
}

// stats: P slots of `slot` ints -- slot p = {row-maximum keys (ordered int) [n], tau keys (unsigned) [n]} of rank p
__global__ void __launch_bounds__(256)
shard_dstar_kernel(int n, int rank, int P, const int* __restrict__ stats, long long slot, int* __restrict__ key_g,
                   int* __restrict__ key_l, float* __restrict__ dstar, int* __restrict__ count,
                   int* __restrict__ cthr_key) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int mk = REID_ROWMAX_INIT;
  unsigned tk = 0xffffffffu;
  for (int p = 0; p < P; ++p) {
    mk = max(mk, stats[p * slot + i]);
    tk = min(tk, (unsigned)stats[p * slot + n + i]);
  }
  const float ds = dstar_of_tau(tk, ordered_int2f(mk));
  key_g[i] = mk;
  if (snake_owner(i >> 8, P) == rank) key_l[snake_local_row(i, P)] = mk;
  dstar[i] = ds;
  count[i] = 0;
  atomicMax(cthr_key + (i >> 5), thr_key(ds));
}

__global__ void __launch_bounds__(SELH_WPB * 32)
shard_merge_local_kernel(const float* __restrict__ D, long long ld, int n, int rank, int P,
                         const float* __restrict__ cm_d, int pitch_d, const float* __restrict__ cm_m, int pitch_m,
                         const int* __restrict__ key_g, const float* __restrict__ dstar_in,
                         const int* __restrict__ count, const unsigned long long* __restrict__ pbuf, int cap, int K,
                         unsigned char* const* __restrict__ peers, long long inbox_off, int nloc_max) {
  __shared__ unsigned long long cand_s[SELH_WPB][SELH_CAND];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int i = blockIdx.x * SELH_WPB + w;
  if (i >= n) return;
  unsigned long long* cand = cand_s[w];
  const int ti = i >> 8;
  const int owner = snake_owner(ti, P);
  unsigned long long* dst = reinterpret_cast<unsigned long long*>(peers[owner] + inbox_off) +
                            ((long long)rank * nloc_max + snake_local_row(i, P)) * K;
  const float m = ordered_int2f(key_g[i]);
  const int cnt = count[i];
  if (cnt <= cap) {
    const unsigned long long* src = pbuf + (long long)i * cap;
    for (int x = lane; x < cnt; x += 32) {
      const unsigned long long e = src[x];
      cand[x] = make_ck(__uint_as_float((unsigned)(e >> 32)), m, (int)(unsigned)(e & 0xffffffffu));
    }
    int P2 = 32;
    while (P2 < cnt) P2 <<= 1;
    for (int x = cnt + lane; x < P2; x += 32) cand[x] = ~0ull;
    __syncwarp();
    warp_bitonic_u64_sel(cand, P2, lane);
    for (int x = lane; x < K; x += 32) dst[x] = x < P2 ? cand[x] : ~0ull;
    return;
  }
  // more candidates than slots (massive ties): K rounds of warp-wide arg-min over everything this rank holds of the row
  const float ds = dstar_in[i];
  const int nchunks = (n + 31) >> 5;
  const bool mine = owner == rank;
  const int li = snake_local_row(i, P);
  const int npre = 8 * snake_tiles_below(ti, rank, P);
  const float* cmd = cm_d + (long long)li * pitch_d;
  const float* cmm = cm_m + (long long)i * pitch_m;
  unsigned long long last = 0;
  for (int r = 0; r < K; ++r) {
    unsigned long long best = ~0ull;
    if (mine) {
      for (int cc = ti * 8; cc < nchunks; ++cc) {
        if (!(__ldg(cmd + cc) <= ds || ds == INFINITY)) continue;
        const int col = cc * 32 + lane;
        if (col < n) {
          const unsigned long long k0 = make_ck(__ldg(D + (long long)li * ld + col), m, col);
          if ((r == 0 || k0 > last) && k0 < best) best = k0;
        }
      }
    }
    for (int c = 0; c < npre; ++c) {
      if (!(__ldg(cmm + c) <= ds || ds == INFINITY)) continue;
      const int lj = (c >> 3) * 256 + (c & 7) * 32 + lane;                       // local row of the mirrored element
      const int j = snake_global_tile(c >> 3, rank, P) * 256 + (c & 7) * 32 + lane;
      const unsigned long long k0 = make_ck(__ldg(D + (long long)lj * ld + i), m, j);
      if ((r == 0 || k0 > last) && k0 < best) best = k0;
    }
    best = warp_min_u64(best);
    if (best != ~0ull) last = best;
    if (lane == 0) dst[r] = best;
  }
}

constexpr int SHF_WPB = 4;
__global__ void __launch_bounds__(SHF_WPB * 32)
shard_merge_final_kernel(int nloc, int rank, int P, int K, const unsigned long long* __restrict__ inbox, int nloc_max,
                         int P2, unsigned char* const* __restrict__ peers, long long nbr_off,
                         float* __restrict__ nbr_val) {
  extern __shared__ unsigned long long fin_s[];
  const int w = threadIdx.x >> 5, lane = lane_id();
  const int l = blockIdx.x * SHF_WPB + w;
  if (l >= nloc) return;
  unsigned long long* cand = fin_s + (long long)w * P2;
  const int total = P * K;
  for (int x = lane; x < P2; x += 32)
    cand[x] = x < total ? inbox[((long long)(x / K) * nloc_max + l) * K + (x % K)] : ~0ull;
  __syncwarp();
  warp_bitonic_u64_sel(cand, P2, lane);
  const long long row = own_row(rank, -P, l);
  for (int x = lane; x < K; x += 32) {
    const unsigned long long k0 = cand[x];
    const int col = (int)(unsigned)(k0 & 0xffffffffu);
    for (int p = 0; p < P; ++p) reinterpret_cast<int*>(peers[p] + nbr_off)[row * K + x] = col;
    if (nbr_val) nbr_val[(long long)l * K + x] = key2f((uint32_t)(k0 >> 32));
  }
}

int launch_shard_tau(int n, int rank, int world, const float* cm_d, int pitch_d, const float* cm_m, int pitch_m, int K,
                     unsigned* tau_key, cudaStream_t s) {
  if (n == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= 256, "shard_tau: K=%d outside [1,256]", K);
  REID_REQUIRE(rank >= 0 && rank < world, "shard_tau: rank %d of %d", rank, world);
  const int grid = cdiv(n, SELH_WPB), threads = SELH_WPB * 32;
#define REID_SHT(SS) shard_tau_kernel<SS><<<grid, threads, 0, s>>>(n, rank, world, cm_d, pitch_d, cm_m, pitch_m, K, tau_key)
  // the warp keeps 32 * S chunk minima; their K-th smallest is the bound, so 32 * S must exceed K comfortably: at
  // K == 32 * S the "K-th smallest kept" is the LARGEST kept value -- a useless bound (measured: K = 64 with S = 2
  // sent most rows to the exact arg-min fallback, 41 s instead of 0.6 s over a 1.1 M x 1.1 M panel sweep)
  if (K <= 16) REID_SHT(1);
  else if (K <= 40) REID_SHT(2);
  else if (K <= 80) REID_SHT(4);
  else if (K <= 160) REID_SHT(8);
  else REID_SHT(16);
#undef REID_SHT
  return check_launch("shard_tau_kernel");
}

int launch_shard_dstar(int n, int rank, int world, const int* stats, long long slot, int* key_g, int* key_l,
                       float* dstar, int* count, int* cthr_key, cudaStream_t s) {
  if (n == 0) return REID_OK;
  REID_REQUIRE(slot >= 2ll * n, "shard_dstar: slot of %lld ints < 2 n", slot);
  REID_CUDA(cudaMemsetAsync(cthr_key, 0x80, (size_t)4 * chunk_pitch(n), s));      // below every ordered float key
  shard_dstar_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, rank, world, stats, slot, key_g, key_l, dstar, count, cthr_key);
  return check_launch("shard_dstar_kernel");
}

int launch_shard_push(const float* D, int nloc, int rank, int world, int n, long long ld, const float* cm_d,
                      int pitch_d, const float* dstar, const int* cthr_key, int* count, unsigned long long* pbuf,
                      int cap, cudaStream_t s) {
  if (nloc == 0) return REID_OK;
  REID_REQUIRE(cap >= 1 && cap <= SELH_CAND, "shard_push: cap=%d outside [1,%d]", cap, SELH_CAND);
  REID_REQUIRE(cdiv(n, 32) <= 65535, "shard_push: %d rows exceed the 16-bit chunk list", n);
  REID_REQUIRE(((uintptr_t)cm_d & 15) == 0 && pitch_d % 4 == 0 && ((uintptr_t)cthr_key & 15) == 0,
               "shard_push: chunk minima / thresholds must be 16-byte aligned");
  const char* pfe = getenv("REID_SEL_PF");
  select_push_kernel<<<cdiv(nloc, SELH_WPB), SELH_WPB * 32, 0, s>>>(D, nloc, rank, -world, n, ld, cm_d, cdiv(n, 32),
                                                                     pitch_d, dstar, cthr_key, count, pbuf, cap,
                                                                     pfe ? atoi(pfe) : 1);
  return check_launch("select_push_kernel");
}

int launch_shard_merge_local(const float* D, long long ld, int n, int rank, int world, const float* cm_d, int pitch_d,
                             const float* cm_m, int pitch_m, const int* key_g, const float* dstar, const int* count,
                             const unsigned long long* pbuf, int cap, int K, unsigned char* const* peers,
                             long long inbox_off, int nloc_max, cudaStream_t s) {
  if (n == 0) return REID_OK;
  REID_REQUIRE(cap >= 1 && cap <= SELH_CAND && K >= 1 && K <= 256, "shard_merge_local: cap=%d K=%d", cap, K);
  REID_REQUIRE(inbox_off % 8 == 0, "shard_merge_local: inbox offset must be 8-byte aligned");
  shard_merge_local_kernel<<<cdiv(n, SELH_WPB), SELH_WPB * 32, 0, s>>>(D, ld, n, rank, world, cm_d, pitch_d, cm_m,
                                                                        pitch_m, key_g, dstar, count, pbuf, cap, K,
                                                                        peers, inbox_off, nloc_max);
  return check_launch("shard_merge_local_kernel");
}

int launch_shard_merge_final(int nloc, int rank, int world, int K, const unsigned long long* inbox, int nloc_max,
                             unsigned char* const* peers, long long nbr_off, float* nbr_val, cudaStream_t s) {
  if (nloc == 0) return REID_OK;
  int P2 = 32;
  while (P2 < world * K) P2 <<= 1;
  const size_t smem = (size_t)SHF_WPB * P2 * 8;
  REID_REQUIRE(smem <= 200 * 1024, "shard_merge_final: %d ranks x K=%d do not fit shared memory", world, K);
  REID_REQUIRE(nbr_off % 4 == 0, "shard_merge_final: table offset must be 4-byte aligned");
  REID_CUDA(cudaFuncSetAttribute(shard_merge_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  shard_merge_final_kernel<<<cdiv(nloc, SHF_WPB), SHF_WPB * 32, smem, s>>>(nloc, rank, world, K, inbox, nloc_max, P2,
                                                                            peers, nbr_off, nbr_val);
  return check_launch("shard_merge_final_kernel");
}

int launch_select_topk(const float* D, int rows, int cols, long long ld, const int* rowmax_key, int K,
                       int* nbr, float* nbr_val, int cap_override, cudaStream_t s) {
  if (rows == 0) return REID_OK;
  REID_REQUIRE(K >= 1 && K <= 256, "select_topk: K=%d outside [1,256]", K);
  REID_REQUIRE(cols >= K, "select_topk: row length %d is shorter than K=%d", cols, K);
  REID_REQUIRE(ld % 4 == 0, "select_topk: pitch %lld must be a multiple of 4", ld);
  if (cols <= 32768) {
    select_topk_kernel<256><<<rows, 256, 0, s>>>(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val,
                                                 cap_override);
  } else {
    select_topk_kernel<1024><<<rows, 1024, 0, s>>>(D, rows, cols, ld, rowmax_key, K, nbr, nbr_val,
                                                   cap_override);
  }
  return check_launch("select_topk_kernel");
}

}  // namespace reid
