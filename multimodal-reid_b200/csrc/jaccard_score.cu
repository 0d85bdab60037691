// K9 + K10 (+ K11 + K12): the float16 Jaccard chain, the lambda mix and -- on the evaluator's path --
// CMC / mAP, one CTA per query, so that neither the q x N Jaccard matrix nor (when only the metrics
// are wanted) the q x g re-ranked matrix ever exists in memory (reranking.py:84-100,
// evaluate.py:40-88).
//
//   phase 1  t[0:N] (float16, shared memory) accumulates min(V_qe[i,c], V_qe[g,c]) over the query's
//            columns c in ascending order, a CTA barrier per column: every t[g] sees the reference's
//            rounding sequence (fp16 RN after every add).  Column lists are prefetched PF ahead.
//   phase 2  the gallery entries sharing the query's identity come from a hash index over g_pid
//            (pid_index_*, built once per evaluation): a few dozen columns, each junk (same camera,
//            evaluate.py:55) or a true match.  Their final distances are computed first and the
//            matches' (distance, column) keys sorted in shared memory.
//   phase 3  one streaming pass over the query's row of D: final = fp16(J * fp16(1-lambda)) +
//            O * fp32(lambda).  Scoring only needs  pos(p) = #{kept j : (final_j, j) <= (final_p, p)}
//            for every match p, so entries beyond the worst match are dropped with one compare
//            (non-overlapping entries, t = 0, have final monotone in D: a pre-computed threshold on
//            raw D decides without any arithmetic), the rest are binned among the sorted matches.
//            Junk entries are taken out of the bins afterwards.  With `out` the row is also written.
//   phase 4  prefix sums over the bins: kept rank and match count of every match -> AP (float64),
//            first-hit rank -> CMC histogram; identical to ranking the row explicitly.
#include "jaccard_common.cuh"
#include "kernels.h"

namespace reid {

// ---------------------------------------------------------------------------------------------
// identity index: open-addressing hash table pid -> slot, counting sort of the gallery by slot
// ---------------------------------------------------------------------------------------------
__global__ void pid_init_kernel(unsigned long long* keys, int* cnt, int* cursor, int T) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= T) { cnt[i] = 0; cursor[i] = 0; if (i < T) keys[i] = PID_EMPTY; }
}
__global__ void pid_insert_kernel(const long long* __restrict__ g_pid, int g, unsigned long long* keys, int* cnt,
                                  int T) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g) return;
  const unsigned long long k = (unsigned long long)g_pid[j];
  int slot = T;
  if (k != PID_EMPTY) {
    unsigned h = pid_hash(k) & (unsigned)(T - 1);
    for (;;) {
      const unsigned long long prev = atomicCAS(keys + h, PID_EMPTY, k);
      if (prev == PID_EMPTY || prev == k) { slot = (int)h; break; }
      h = (h + 1) & (unsigned)(T - 1);
    }
  }
  atomicAdd(cnt + slot, 1);
}
__global__ void pid_fill_kernel(const long long* __restrict__ g_pid, int g, const unsigned long long* __restrict__ keys,
                                const long long* __restrict__ start, int* cursor, int T, int* __restrict__ list) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g) return;
  const int slot = pid_find(keys, T, g_pid[j]);
  list[start[slot] + atomicAdd(cursor + slot, 1)] = j;
}

int launch_pid_index(const long long* g_pid, int g, int T, unsigned long long* keys, int* cnt, long long* start,
                     int* cursor, int* list, cudaStream_t s) {
  REID_REQUIRE(T >= 2 && (T & (T - 1)) == 0 && T >= 2 * (long long)g,
               "pid_index: table size %d must be a power of two >= 2 * g (g=%d)", T, g);
  pid_init_kernel<<<cdiv(T + 1, 256), 256, 0, s>>>(keys, cnt, cursor, T);
  int rc = check_launch("pid_init_kernel");
  if (rc) return rc;
  if (g > 0) {
    pid_insert_kernel<<<cdiv(g, 256), 256, 0, s>>>(g_pid, g, keys, cnt, T);
    if ((rc = check_launch("pid_insert_kernel"))) return rc;
  }
  if ((rc = launch_exclusive_scan(cnt, T + 1, start, s))) return rc;
  if (g > 0) {
    pid_fill_kernel<<<cdiv(g, 256), 256, 0, s>>>(g_pid, g, keys, start, cursor, T, list);
    if ((rc = check_launch("pid_fill_kernel"))) return rc;
  }
  return REID_OK;
}

// ---------------------------------------------------------------------------------------------
// fused Jaccard + mix (+ score)
// ---------------------------------------------------------------------------------------------
constexpr int JS_CHUNK = 256;   // query columns staged per chunk
constexpr int JS_RING = 8;      // column lists in flight (cp.async groups)
constexpr int JS_SLOT = 256;    // ring slot capacity in entries; longer lists read their tail directly
constexpr int JS_CAP = 512;     // identity-group entries staged / matches ranked per pass over the row
constexpr int JS_QSLOTS = 96;   // per-warp queue capacity (entries awaiting a dense 32-lane pass)

// shared memory after the overlap row t[Npad]
struct JsLayout {
  static constexpr size_t stage = 0;                                   // phase 1: cstart/clen/cval; later: mk
  static constexpr size_t stage_bytes = JS_CAP * 8 > JS_CHUNK * 14 ? JS_CAP * 8 : JS_CHUNK * 14;
  static constexpr size_t gcol = stage + stage_bytes;                  // int[JS_CAP]   column | junk << 31
  static constexpr size_t gd = gcol + JS_CAP * 4;                      // float[JS_CAP] D[i, q + column]
  static constexpr size_t work = gd + JS_CAP * 4;                      // phase 1: ring; phases 2-4: the rest
  static constexpr size_t ring_bytes = (size_t)JS_RING * JS_SLOT * 8;
  static constexpr size_t hv = work;                                   // int[JS_CAP + 1] kept-entry bins
  static constexpr size_t hp = hv + (JS_CAP + 1) * 4;                  // int[JS_CAP + 1] match bins
  static constexpr size_t thrA = hp + (JS_CAP + 1) * 4;                // float[JS_CAP]
  static constexpr size_t thrB = thrA + JS_CAP * 4;                    // float[JS_CAP]
  static constexpr size_t queue = (thrB + JS_CAP * 4 + 15) & ~(size_t)15;   // u64[warps][2][JS_QSLOTS]
  static constexpr size_t bytes(int warps) {
    const size_t a = queue + (size_t)warps * 2 * JS_QSLOTS * 8, b = work + ring_bytes;
    return a > b ? a : b;
  }
};

#ifdef REID_JS_PROFILE
__device__ unsigned long long g_js_prof[8];
#define JS_TICK(i) do { __syncthreads(); if (threadIdx.x == 0) { long long _n = clock64(); atomicAdd(&g_js_prof[i], (unsigned long long)(_n - _t0)); _t0 = _n; } } while (0)
#else
#define JS_TICK(i) do { } while (0)
#endif

// Dense 32-lane passes over a warp's queue (kept out of line: the streaming loop stays small enough for the
// instruction cache).  Entries: (float bits of D << 32) | g  resp. | column.
__device__ __noinline__ void js_drain_overlap(const unsigned long long* qa, int cnt, const __half* t, float m, float c1h,
                                              float c2, int q, const unsigned long long* mk, int P,
                                              unsigned long long maxkey, int* hv) {
  const int lane = threadIdx.x & 31;
  if (lane < cnt) {
    const unsigned long long e = qa[lane];
    const int g = (int)(unsigned)(e & 0xffffffffu);
    const float f = final_value(t[g], __uint_as_float((unsigned)(e >> 32)), m, c1h, c2);
    const unsigned long long key = final_key(f, g - q);
    if (key <= maxkey) bin_add(hv, lower_bound_u64(mk, P, key));
  }
  __syncwarp();
}
__device__ __noinline__ void js_drain_zero(const unsigned long long* qz, int cnt, const float* thrA, const float* thrB,
                                           const unsigned long long* mk, int P, int* hv) {
  const int lane = threadIdx.x & 31;
  if (lane < cnt) {
    const unsigned long long e = qz[lane];
    const int col = (int)(unsigned)(e & 0xffffffffu);
    const float d = __uint_as_float((unsigned)(e >> 32));
    int lo = 0, hi = P;                                   // #{i : d > thrB[i]}; < P because d <= thrB[P-1]
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (d > thrB[mid]) lo = mid + 1; else hi = mid;
    }
    int bin = lo;
    for (int i = lo; i < P && thrA[i] < d; ++i) {         // exact ties in final: the smaller column first
      if ((int)(unsigned)(mk[i] & 0xffffffffu) < col) bin = i + 1; else break;
    }
    if (bin < P) bin_add(hv, bin);
  }
  __syncwarp();
}

template <int T>
__device__ __forceinline__ void
js_dense_query(const int qi, const int* __restrict__ Qc, const __half* __restrict__ Qv,
               const long long* __restrict__ qoff, const int* __restrict__ Qn,
               const long long* __restrict__ colptr, const unsigned long long* __restrict__ csc_ent,
               const float* __restrict__ D, long long ld,
               const int* __restrict__ rowmax_key, int q0, int q_stride, int q, int N, int Npad, float c1,
               float c2, float* __restrict__ out, long long ld_out,
               __half* __restrict__ jaccard_out, long long ld_j, const ScoreArgs& sc) {
  extern __shared__ __align__(16) unsigned char smem_j[];
  __half* t = reinterpret_cast<__half*>(smem_j);
  unsigned char* aux = smem_j + (((size_t)Npad * 2 + 15) & ~(size_t)15);
  long long* cstart = reinterpret_cast<long long*>(aux + JsLayout::stage);      // phase 1 view of the stage area
  int* clen = reinterpret_cast<int*>(cstart + JS_CHUNK);
  __half* cval = reinterpret_cast<__half*>(clen + JS_CHUNK);
  unsigned long long* mk = reinterpret_cast<unsigned long long*>(aux + JsLayout::stage);   // phase 2-4 view
  int* gcol = reinterpret_cast<int*>(aux + JsLayout::gcol);
  float* gd = reinterpret_cast<float*>(aux + JsLayout::gd);
  int* hv = reinterpret_cast<int*>(aux + JsLayout::hv);
  int* hp = reinterpret_cast<int*>(aux + JsLayout::hp);
  float* thrA = reinterpret_cast<float*>(aux + JsLayout::thrA);
  float* thrB = reinterpret_cast<float*>(aux + JsLayout::thrB);
  __shared__ int s_lo, s_len, s_total, s_cnt, s_first, s_zfast;

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  unsigned long long* wq = reinterpret_cast<unsigned long long*>(aux + JsLayout::queue) + w * 2 * JS_QSLOTS;
  // qi: local query index (D / rowmax / out rows)
  const long long row = own_row(q0, q_stride, qi);        // global probe index (V_qe row)
#ifdef REID_JS_PROFILE
  long long _t0 = clock64();
#endif
  {
    uint4 z = make_uint4(0, 0, 0, 0);
    uint4* t4 = reinterpret_cast<uint4*>(t);
    for (int j = tid; j < (Npad >> 3); j += T) t4[j] = z;
  }
  const float m = ordered_int2f(rowmax_key[qi]);
  const float* drow = D + (long long)qi * ld;
  const float c1h = __half2float(__float2half_rn(c1));
  const bool scoring = sc.ap != nullptr;
  long long qc = 0;
  __syncthreads();                      // t is zero before anyone adds to it

  if (w == 0) {
    // ---------------------------------------------------------------- phase 1: ordered fp16 overlap
    // One warp walks the query's columns in ascending order.  Column lists (8-byte entries, contiguous in the
    // CSC) arrive through a ring of cp.async groups JS_RING columns ahead; every lane consumes the entries it
    // copied itself.  Program order plus a __syncwarp per column keeps every t[g] on the reference's rounding
    // sequence (fp16 RN after every add) with no CTA barrier on the path.
    unsigned long long* ring = reinterpret_cast<unsigned long long*>(aux + JsLayout::work);
    const long long o = qoff[row];
    const int nq = Qn[row];
    auto issue = [&](int c) {                              // this lane's entries of staged column c
      const int len = min(clen[c], JS_SLOT);
      const unsigned long long* src = csc_ent + cstart[c];
      unsigned long long* dst = ring + (c % JS_RING) * JS_SLOT;
      for (int p = lane; p < len; p += 32) cp_async8(dst + p, src + p);
    };
    for (int base = 0; base < nq; base += JS_CHUNK) {
      const int nc = min(JS_CHUNK, nq - base);
      __syncwarp();
      for (int i = lane; i < nc; i += 32) {
        const int c = Qc[o + base + i];
        const long long a = colptr[c];
        cstart[i] = a;
        clen[i] = (int)(colptr[c + 1] - a);
        cval[i] = Qv[o + base + i];
      }
      __syncwarp();
#pragma unroll
      for (int s = 0; s < JS_RING - 1; ++s) {
        if (s < nc) issue(s);
        cp_async_commit();
      }
      for (int e = 0; e < nc; ++e) {
        if (e + JS_RING - 1 < nc) issue(e + JS_RING - 1);
        cp_async_commit();
        cp_async_wait<JS_RING - 1>();                      // this lane's copies of column e have landed
        const float vi = __half2float(cval[e]);
        const int len = clen[e];
        const unsigned long long* slot = ring + (e % JS_RING) * JS_SLOT;
        const int lim = min(len, JS_SLOT);
        for (int p0 = 0; p0 < lim; p0 += 128) {            // four independent read-modify-write chains per lane
          unsigned long long ent[4];
          float tv[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * 32 + lane;
            ent[u] = p < lim ? slot[p] : ~0ull;
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
            tv[u] = ent[u] != ~0ull ? __half2float(t[(unsigned)(ent[u] & 0xffffffffu)]) : 0.0f;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (ent[u] != ~0ull) {
              const float vg = __half2float(__ushort_as_half((unsigned short)(ent[u] >> 32)));
              t[(unsigned)(ent[u] & 0xffffffffu)] = __float2half_rn(__fadd_rn(tv[u], fminf(vi, vg)));
            }
          }
        }
        if (len > JS_SLOT) {
          const unsigned long long* src = csc_ent + cstart[e];
          for (int p = JS_SLOT + lane; p < len; p += 32) {
            const unsigned long long ent = src[p];
            const unsigned g = (unsigned)(ent & 0xffffffffu);
            const float vg = __half2float(__ushort_as_half((unsigned short)(ent >> 32)));
            t[g] = __float2half_rn(__fadd_rn(__half2float(t[g]), fminf(vi, vg)));
          }
        }
        __syncwarp();
      }
      cp_async_wait<0>();
    }
  } else if (scoring && w == T / 32 - 1) {
    // Meanwhile the last warp looks the query's identity group up and stages its first slice (column, junk
    // flag, raw distance): these loads do not depend on t.
    qc = sc.q_cam[qi];
    int lo = 0, len = 0;
    if (lane == 0) {
      const int slot = pid_find(sc.keys, sc.T, sc.q_pid[qi]);
      if (slot >= 0) { lo = (int)sc.start[slot]; len = (int)(sc.start[slot + 1] - sc.start[slot]); }
    }
    lo = __shfl_sync(FULL, lo, 0);
    len = __shfl_sync(FULL, len, 0);
    const int* grp = sc.list + lo;
    int matches = 0;
    for (int e = lane; e < len; e += 32) {
      const int j = grp[e];
      const bool junk = sc.g_cam[j] == qc;
      matches += !junk;
      if (e < JS_CAP) { gcol[e] = j | (junk ? (int)0x80000000 : 0); gd[e] = __ldg(drow + q + j); }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) matches += __shfl_xor_sync(FULL, matches, off);
    if (lane == 0) { s_lo = lo; s_len = len; s_total = matches; s_first = 0x7fffffff; }
  }
  __syncthreads();
  JS_TICK(1);

  if (jaccard_out) {
    for (int g = tid; g < N; g += T) jaccard_out[(long long)qi * ld_j + g] = jaccard_from_overlap(t[g]);
  }

  // plain mix of the gallery columns (re_ranking()'s return value), coalesced scalar accesses
  auto mix_only = [&]() {
    float* orow = out + (long long)qi * ld_out;
    int g = q + tid;
    for (; g + 3 * T < N; g += 4 * T) {
      float dv[4];
      __half tv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { dv[u] = __ldg(drow + g + u * T); tv[u] = t[g + u * T]; }
#pragma unroll
      for (int u = 0; u < 4; ++u) orow[g + u * T - q] = final_value(tv[u], dv[u], m, c1h, c2);
    }
    for (; g < N; g += T) orow[g - q] = final_value(t[g], __ldg(drow + g), m, c1h, c2);
  };

  if (!scoring) {
    if (out) mix_only();
    return;
  }
  qc = sc.q_cam[qi];
  const int* grp = sc.list + s_lo;      // gallery columns with the query's pid: each junk or a match
  const int L = s_len;
  const int total = s_total;
  if (total == 0) {                     // query identity absent from the gallery: skipped (evaluate.py:61-63)
    if (tid == 0) { sc.ap[qi] = -1.0; sc.first_hit[qi] = -1; }
    if (out) mix_only();
    return;
  }

  double acc = 0.0;                     // sum over matches of (#matches up to p) / (kept rank of p); warp 0
  bool wrote = false;
  for (int e0 = 0; e0 < L; e0 += JS_CAP) {
    const int ns = min(JS_CAP, L - e0);
    // ---------------------------------------------------------------- phase 2: this slice's matches, sorted
    if (e0 > 0) {                       // slice 0 was staged by the last warp during phase 1
      for (int e = tid; e < ns; e += T) {
        const int j = grp[e0 + e];
        gcol[e] = j | (sc.g_cam[j] == qc ? (int)0x80000000 : 0);
        gd[e] = __ldg(drow + q + j);
      }
    }
    if (tid == 0) { s_cnt = 0; s_zfast = (m > 0.0f && m < INFINITY && c2 > 0.0f && c2 < INFINITY) ? 1 : 0; }
    for (int i = tid; i <= JS_CAP; i += T) { hv[i] = 0; hp[i] = 0; }
    __syncthreads();
    if (ns <= 32) {
      // the usual case: one warp sorts the slice's keys in registers (bitonic network over shuffles)
      if (w == 0) {
        unsigned long long key = ~0ull;
        if (lane < ns) {
          const int j = gcol[lane];
          if (j >= 0) key = final_key(final_value(t[q + j], gd[lane], m, c1h, c2), j);
        }
#pragma unroll
        for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
          for (int jj = k >> 1; jj > 0; jj >>= 1) {
            const unsigned long long other = __shfl_xor_sync(FULL, key, jj);
            const bool up = (lane & k) == 0, lower = (lane & jj) == 0;
            const bool take_min = up == lower;
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
        if (j >= 0) mk[atomicAdd(&s_cnt, 1)] = final_key(final_value(t[q + j], gd[e], m, c1h, c2), j);
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
    if (P == 0) { __syncthreads(); continue; }   // uniform: s_cnt is shared
    // Non-overlapping entries (t = 0) have final = fp16(1-lambda) + (D/max)*lambda, monotone in D, so their rank
    // among the matches follows from raw-D thresholds: match i is passed strictly when D > thrB[i]; it is tied in
    // final when thrA[i] < D <= thrB[i] (then the column decides).  One warp per threshold.
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
    JS_TICK(2);

    // ---------------------------------------------------------------- phase 3: one pass over the row
    if (out && !wrote) {
      auto slow = [&](__half th, float d, int g) -> float {
        const float f = final_value(th, d, m, c1h, c2);
        const unsigned long long key = final_key(f, g - q);
        if (key <= maxkey) bin_add(hv, lower_bound_u64(mk, P, key));
        return f;
      };
      float* orow = out + (long long)qi * ld_out;
      int g = q + tid;
      for (; g + 3 * T < N; g += 4 * T) {
        float dv[4];
        __half tv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { dv[u] = __ldg(drow + g + u * T); tv[u] = t[g + u * T]; }
#pragma unroll
        for (int u = 0; u < 4; ++u) orow[g + u * T - q] = slow(tv[u], dv[u], g + u * T);
      }
      for (; g < N; g += T) orow[g - q] = slow(t[g], __ldg(drow + g), g);
      wrote = true;
    } else {
      // Metrics only: 16 entries per thread per trip (two aligned groups of 8), next trip's D prefetched.
      // Per entry the common case is two compares: t == 0 and D beyond the worst match's threshold.  The hot
      // bin between the two worst matches is counted in a register.  Everything else is rare and goes through
      // two per-warp queues that are drained in dense 32-lane passes: overlapping entries (t != 0: full
      // arithmetic, 64-bit key search) and non-overlapping entries ranking before the second-worst match
      // (raw-D threshold search).
      const float dthr = zfast ? thrB[P - 1] : -INFINITY;          // t = 0 and D > dthr: after every match
      const float bsec = P >= 2 ? thrB[P - 2] : -INFINITY;
      const float atop = zfast ? thrA[P - 1] : -INFINITY;
      int cnt_top = 0;
      unsigned long long* qa = wq;
      unsigned long long* qz = wq + JS_QSLOTS;
      int na = 0, nzq = 0;
      const uint4* t4 = reinterpret_cast<const uint4*>(t);
      const float4* d4 = reinterpret_cast<const float4*>(drow);
      auto ldg8 = [&](int g, float4& x0, float4& x1) {
        if (g < N) { x0 = __ldg(d4 + (g >> 2)); x1 = __ldg(d4 + (g >> 2) + 1); }
        else { x0 = make_float4(0.f, 0.f, 0.f, 0.f); x1 = x0; }
      };
      constexpr int STRIDE = T * 8;
      int gw = (q & ~7) + (tid & ~31) * 8;                        // warp-uniform base of this warp's group
      int gb = gw + lane * 8;                                     // this lane's 8 aligned entries
      float4 c0, c1v, n0, n1, f0, f1;                             // current, next, next-but-one (prefetch depth 2)
      ldg8(gb, c0, c1v);
      ldg8(gb + STRIDE, n0, n1);
      for (; gw < N; gw += STRIDE, gb += STRIDE) {
        ldg8(gb + 2 * STRIDE, f0, f1);
        const uint4 tv = gb < N ? t4[gb >> 3] : make_uint4(0, 0, 0, 0);
        const float dmin = fminf(fminf(fminf(c0.x, c0.y), fminf(c0.z, c0.w)), fminf(fminf(c1v.x, c1v.y), fminf(c1v.z, c1v.w)));
        // group-level test first: nothing overlapping and nothing at or before the worst match -> next group
        const bool busy = gb < N && (((tv.x | tv.y | tv.z | tv.w) != 0) || !zfast || dmin <= dthr);
        if (__any_sync(FULL, busy)) {
          const unsigned tw[4] = {tv.x, tv.y, tv.z, tv.w};
          const float dv[8] = {c0.x, c0.y, c0.z, c0.w, c1v.x, c1v.y, c1v.z, c1v.w};
          const bool interior = gb >= q && gb + 8 <= N;
          // 8-bit masks of this lane's entries: overlapping / non-overlapping but ranked before the second-worst
          unsigned ma = 0, mz = 0;
          if (busy) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int g = gb + e;
              const bool valid = interior || (g >= q && g < N);
              const unsigned hb = (tw[e >> 1] >> ((e & 1) * 16)) & 0xffffu;
              const float d = dv[e];
              if (valid) {
                if (hb != 0 || !zfast) ma |= 1u << e;
                else if (d <= dthr) {
                  if (d > bsec && d <= atop) ++cnt_top; else mz |= 1u << e;
                }
              }
            }
          }
          if (__any_sync(FULL, (ma | mz) != 0)) {
            // warp prefix sums of both counts at once
            const unsigned mine = (unsigned)__popc(ma) | ((unsigned)__popc(mz) << 16);
            unsigned incl = mine;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
              const unsigned y = __shfl_up_sync(FULL, incl, off);
              if (lane >= off) incl += y;
            }
            const unsigned tot = __shfl_sync(FULL, incl, 31);
            const unsigned excl = incl - mine;
            const int totA = (int)(tot & 0xffffu), totZ = (int)(tot >> 16);
            if (totA) {
              if (na + totA > JS_QSLOTS && na > 0) { js_drain_overlap(qa, na, t, m, c1h, c2, q, mk, P, maxkey, hv); na = 0; }       // make room: partial pass
              if (na + totA <= JS_QSLOTS) {
                int pos = na + (int)(excl & 0xffffu);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if ((ma >> e) & 1u) qa[pos++] = ((unsigned long long)__float_as_uint(dv[e]) << 32) | (unsigned)(gb + e);
                na += totA;
                __syncwarp();
                while (na >= 32) { na -= 32; js_drain_overlap(qa + na, 32, t, m, c1h, c2, q, mk, P, maxkey, hv); }
              } else {                                            // dense row: one entry per lane at a time
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const bool on = (ma >> e) & 1u;
                  const unsigned act = __ballot_sync(FULL, on);
                  if (act) {
                    if (on) qa[na + __popc(act & ((1u << lane) - 1))] =
                        ((unsigned long long)__float_as_uint(dv[e]) << 32) | (unsigned)(gb + e);
                    __syncwarp();
                    js_drain_overlap(qa + na, __popc(act), t, m, c1h, c2, q, mk, P, maxkey, hv);
                  }
                }
              }
            }
            if (totZ) {
              if (nzq + totZ > JS_QSLOTS && nzq > 0) { js_drain_zero(qz, nzq, thrA, thrB, mk, P, hv); nzq = 0; }
              if (nzq + totZ <= JS_QSLOTS) {
                int pos = nzq + (int)(excl >> 16);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if ((mz >> e) & 1u) qz[pos++] = ((unsigned long long)__float_as_uint(dv[e]) << 32) | (unsigned)(gb + e - q);
                nzq += totZ;
                __syncwarp();
                while (nzq >= 32) { nzq -= 32; js_drain_zero(qz + nzq, 32, thrA, thrB, mk, P, hv); }
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const bool on = (mz >> e) & 1u;
                  const unsigned act = __ballot_sync(FULL, on);
                  if (act) {
                    if (on) qz[nzq + __popc(act & ((1u << lane) - 1))] =
                        ((unsigned long long)__float_as_uint(dv[e]) << 32) | (unsigned)(gb + e - q);
                    __syncwarp();
                    js_drain_zero(qz + nzq, __popc(act), thrA, thrB, mk, P, hv);
                  }
                }
              }
            }
          }
        }
        c0 = n0; c1v = n1; n0 = f0; n1 = f1;
      }
      if (na > 0) js_drain_overlap(qa, na, t, m, c1h, c2, q, mk, P, maxkey, hv);
      if (nzq > 0) js_drain_zero(qz, nzq, thrA, thrB, mk, P, hv);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) cnt_top += __shfl_xor_sync(FULL, cnt_top, off);
      if (lane == 0 && cnt_top) atomicAdd(hv + (P - 1), cnt_top);
    }
    __syncthreads();
    JS_TICK(3);
    // the whole identity group: junk leaves the kept bins, every match enters the match bins
    for (int e = tid; e < L; e += T) {
      int j;
      float d;
      if (e >= e0 && e < e0 + ns) { j = gcol[e - e0]; d = gd[e - e0]; }
      else { j = grp[e]; d = __ldg(drow + q + j); if (sc.g_cam[j] == qc) j |= (int)0x80000000; }
      const int col = j & 0x7fffffff;
      const unsigned long long key = final_key(final_value(t[q + col], d, m, c1h, c2), col);
      if (key <= maxkey) {
        const int bin = lower_bound_u64(mk, P, key);
        if (j < 0) atomicSub(hv + bin, 1); else atomicAdd(hp + bin, 1);
      }
    }
    __syncthreads();
    // ---------------------------------------------------------------- phase 4: ranks of this slice's matches
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
    JS_TICK(4);
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

// One CTA per query (over_list == NULL), or a persistent grid over the queries the sparse path handed back.
template <int T>
__global__ void __launch_bounds__(T)
jaccard_score_kernel(const int* __restrict__ Qc, const __half* __restrict__ Qv,
                     const long long* __restrict__ qoff, const int* __restrict__ Qn,
                     const long long* __restrict__ colptr, const unsigned long long* __restrict__ csc_ent,
                     const float* __restrict__ D, long long ld,
                     const int* __restrict__ rowmax_key, int q0, int q_stride, int q, int N, int Npad, float c1,
                     float c2, float* __restrict__ out, long long ld_out,
                     __half* __restrict__ jaccard_out, long long ld_j, ScoreArgs sc,
                     const int* __restrict__ over_list, const int* __restrict__ over_n) {
  if (over_list == nullptr) {
    js_dense_query<T>(blockIdx.x, Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld, rowmax_key, q0, q_stride, q, N, Npad, c1, c2,
                      out, ld_out, jaccard_out, ld_j, sc);
    return;
  }
  const int n = *over_n;
  for (int i = blockIdx.x; i < n; i += gridDim.x) {
    js_dense_query<T>(over_list[i], Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld, rowmax_key, q0, q_stride, q, N, Npad, c1,
                      c2, out, ld_out, jaccard_out, ld_j, sc);
    __syncthreads();
  }
}

// sparse path (jaccard_sparse.cu)
long long jaccard_sparse_ws_bytes(int nq, int N, int cap);
bool jaccard_sparse_fits(int N, int cap);
int launch_jaccard_sparse(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                          const long long* colptr, const unsigned long long* csc_ent, const float* D, long long ld,
                          const int* rowmax_key, int q0, int q_stride, int nq_local, int q, int N, float c1, float c2,
                          float* out, long long ld_out, const ScoreArgs& sc, void* ws, int cap, int** over_list,
                          int** over_n, cudaStream_t s);

// ws != NULL (ws_bytes >= jaccard_sparse_ws_bytes) selects the sparse path: per-query packed overlap records, with
// the dense kernel below run afterwards as a persistent grid over the queries that did not fit `cap`.
int launch_jaccard_score(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                         const long long* colptr, const unsigned long long* csc_ent, const float* D,
                         long long ld, const int* rowmax_key, int q0, int q_stride, int nq_local, int q, int N,
                         float c1, float c2, float* out, long long ld_out, __half* jaccard_out, long long ld_j,
                         const long long* q_pid, const long long* q_cam, const long long* g_cam,
                         const unsigned long long* pid_keys, const long long* pid_start, const int* pid_list,
                         int pid_T, int max_rank, double* ap, int* first_hit, int* cmc_hist, int* n_valid,
                         void* ws, long long ws_bytes, int cap, cudaStream_t s) {
  if (nq_local == 0) return REID_OK;
  const int Npad = (int)round_up(N, 8);
  const int threads = N <= 32768 ? 256 : 512;
  const size_t smem = (((size_t)Npad * 2 + 15) & ~(size_t)15) + JsLayout::bytes(threads / 32);
  REID_REQUIRE(smem <= 227 * 1024,
               "jaccard: N=%d needs %zu B of shared memory for the per-query overlap row (limit 227 KB => N <= ~110k); "
               "larger problems go through the list-based path (reid_overlap_lists / Engine.evaluate_streamed_reranked / "
               "retrieval.streamed_rerank_eval), which never holds a row of N entries",
               N, smem);
  REID_REQUIRE(ld % 4 == 0 && ld >= Npad, "jaccard: D pitch %lld must be a multiple of 4 and >= N rounded up to 8",
               ld);
  ScoreArgs sc{q_pid, q_cam, g_cam, pid_keys, pid_start, pid_list, pid_T, max_rank, ap, first_hit, cmc_hist, n_valid};
  if (ap) {
    REID_REQUIRE(q_pid && q_cam && g_cam && pid_keys && pid_start && pid_list && first_hit && cmc_hist && n_valid,
                 "jaccard_score: scoring needs the id vectors, the identity index and all outputs");
    REID_REQUIRE(max_rank >= 1 && max_rank <= N - q, "jaccard_score: max_rank=%d outside [1, g=%d]", max_rank, N - q);
  }
  int* over_list = nullptr;
  int* over_n = nullptr;
  int grid = nq_local;
  if (ws != nullptr && jaccard_out == nullptr) {
    REID_REQUIRE(jaccard_sparse_fits(N, cap), "jaccard: sparse capacity %d does not fit (N=%d)", cap, N);
    REID_REQUIRE(ws_bytes >= jaccard_sparse_ws_bytes(nq_local, N, cap),
                 "jaccard: workspace of %lld B is smaller than reid_jaccard_ws_bytes() = %lld B", ws_bytes,
                 jaccard_sparse_ws_bytes(nq_local, N, cap));
    const int rc = launch_jaccard_sparse(Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld, rowmax_key, q0, q_stride, nq_local, q,
                                         N, c1, c2, out, ld_out, sc, ws, cap, &over_list, &over_n, s);
    if (rc) return rc;
    int dev = 0, sms = 0;
    REID_CUDA(cudaGetDevice(&dev));
    REID_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    grid = nq_local < 2 * sms ? nq_local : 2 * sms;
  }
  if (N <= 32768) {
    REID_CUDA(cudaFuncSetAttribute(jaccard_score_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
    jaccard_score_kernel<256><<<grid, 256, smem, s>>>(Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld,
                                                      rowmax_key, q0, q_stride, q, N, Npad, c1, c2, out, ld_out,
                                                      jaccard_out, ld_j, sc, over_list, over_n);
  } else {
    REID_CUDA(cudaFuncSetAttribute(jaccard_score_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
    jaccard_score_kernel<512><<<grid, 512, smem, s>>>(Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld,
                                                      rowmax_key, q0, q_stride, q, N, Npad, c1, c2, out, ld_out,
                                                      jaccard_out, ld_j, sc, over_list, over_n);
  }
#ifdef REID_JS_PROFILE
  {
    cudaStreamSynchronize(s);
    unsigned long long h[8];
    cudaMemcpyFromSymbol(h, g_js_prof, sizeof(h));
    fprintf(stderr, "js_prof cycles/CTA: zero %.0f phase1 %.0f setup %.0f stream %.0f post %.0f (n=%d)\n",
            (double)h[0] / nq_local, (double)h[1] / nq_local, (double)h[2] / nq_local, (double)h[3] / nq_local,
            (double)h[4] / nq_local, nq_local);
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(g_js_prof, z, sizeof(z));
  }
#endif
  return check_launch("jaccard_score_kernel");
}

int launch_jaccard_mix(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                       const long long* colptr, const unsigned long long* csc_ent,
                       const float* D, long long ld, const int* rowmax_key, int q0, int q_stride,
                       int nq_local, int q, int N, float c1, float c2, float* out, long long ld_out, __half* jaccard_out,
                       long long ld_j, void* ws, long long ws_bytes, int cap, cudaStream_t s) {
  return launch_jaccard_score(Qc, Qv, qoff, Qn, colptr, csc_ent, D, ld, rowmax_key, q0, q_stride, nq_local,
                              q, N, c1, c2, out, ld_out, jaccard_out, ld_j, nullptr, nullptr, nullptr, nullptr,
                              nullptr, nullptr, 0, 0, nullptr, nullptr, nullptr, nullptr, ws, ws_bytes, cap, s);
}

long long launch_jaccard_ws_bytes(int nq, int N, int cap) {
  return jaccard_sparse_fits(N, cap) ? jaccard_sparse_ws_bytes(nq, N, cap) : 0;
}

}  // namespace reid
