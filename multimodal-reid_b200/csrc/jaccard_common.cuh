// Device helpers shared by the dense (jaccard_score.cu) and sparse (jaccard_sparse.cu) Jaccard / scoring kernels.
#pragma once
#include "common.cuh"
#include "np_exact.cuh"

namespace reid {

// ---------------------------------------------------------------------------------------------
// identity index lookup (built by launch_pid_index)
// ---------------------------------------------------------------------------------------------
constexpr unsigned long long PID_EMPTY = 0x8000000000000000ull;    // INT64_MIN: handled by slot T

__device__ __forceinline__ unsigned pid_hash(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return (unsigned)x;
}
__device__ __forceinline__ int pid_find(const unsigned long long* __restrict__ keys, int T, long long pid) {
  const unsigned long long k = (unsigned long long)pid;
  if (k == PID_EMPTY) return T;
  unsigned h = pid_hash(k) & (unsigned)(T - 1);
  for (int probe = 0; probe < T; ++probe) {
    const unsigned long long cur = keys[h];
    if (cur == k) return (int)h;
    if (cur == PID_EMPTY) return -1;
    h = (h + 1) & (unsigned)(T - 1);
  }
  return -1;
}


// ---------------------------------------------------------------------------------------------
// scoring arguments and the final-distance arithmetic
// ---------------------------------------------------------------------------------------------
struct ScoreArgs {
  const long long* q_pid;            // [nq] this launch's queries
  const long long* q_cam;            // [nq]
  const long long* g_cam;            // [g]
  const unsigned long long* keys;    // identity index (launch_pid_index)
  const long long* start;
  const int* list;
  int T;
  int max_rank;
  double* ap;                        // [nq]; NULL disables scoring
  int* first_hit;                    // [nq]
  int* cmc_hist;                     // [max_rank]
  int* n_valid;                      // [1]
};

// final[i, g] = fp32(fp16(J * fp16(1 - lambda))) + fp32(O * fp32(lambda))          (reranking.py:93-95)
__device__ __forceinline__ float final_value(__half tg, float d, float m, float c1h, float c2) {
  const float ol = __fmul_rn(__fdiv_rn(d, m), c2);
  if (__half_as_ushort(tg) == 0) return __fadd_rn(c1h, ol);       // J = 1 exactly, fp16(1 * c1h) = c1h
  const __half J = jaccard_from_overlap(tg);
  const __half Jl = __float2half_rn(__fmul_rn(__half2float(J), c1h));
  return __fadd_rn(__half2float(Jl), ol);
}
__device__ __forceinline__ unsigned long long final_key(float f, int col) {
  return ((unsigned long long)f2key(__fadd_rn(f, 0.0f)) << 32) | (unsigned)col;
}
__device__ __forceinline__ int lower_bound_u64(const unsigned long long* a, int n, unsigned long long key) {
  int lo = 0, hi = n;                                              // #{a[i] < key}
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// hv[bin] += 1 with the lanes that hit the same bin combined first
__device__ __forceinline__ void bin_add(int* hv, int bin) {
  const unsigned act = __activemask();
  const unsigned peers = __match_any_sync(act, bin);
  if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(hv + bin, __popc(peers));
}
// final0(x) = final_value(t = 0, x) is monotone non-decreasing in x for m > 0, c2 > 0 (every step is a
// monotone rounding).  thr(f, strict) = the largest float x, in the total order of f2ordered_int, with
// final0(x) <= f (or < f when strict); -inf if there is none.
__device__ __forceinline__ bool final0_pred(int x, float f, bool strict, float m, float c1h, float c2) {
  const float v = final_value(__ushort_as_half(0), ordered_int2f(x), m, c1h, c2);
  return strict ? v < f : v <= f;
}
// One warp, one threshold: the algebraic inverse (f - c1h) / c2 * m lands within a few ulps, so the 32 lanes
// test 32 consecutive floats around it at once; only if the flip is not inside that window does lane 0 bisect.
static __device__ __noinline__ float final0_threshold_warp(float f, bool strict, float m, float c1h, float c2) {
  const int lane = threadIdx.x & 31;
  const int lo0 = f2ordered_int(-INFINITY), hi0 = f2ordered_int(INFINITY);
  const float est = (f - c1h) / c2 * m;
  int x0 = (est == est) ? f2ordered_int(fminf(fmaxf(est, -1e30f), 1e30f)) : 0;
  const unsigned okm = __ballot_sync(FULL, final0_pred(x0 - 16 + lane, f, strict, m, c1h, c2));
  if ((okm & 1u) && !(okm >> 31) && ((okm + 1) & okm) == 0)       // 1...10...0: the flip is inside the window
    return ordered_int2f(x0 - 16 + __popc(okm) - 1);
  float r = -INFINITY;
  if (lane == 0 && final0_pred(lo0, f, strict, m, c1h, c2)) {
    int lo = lo0, hi = hi0;
    while (lo < hi) {
      const int mid = (int)(((long long)lo + (long long)hi + 1) >> 1);
      if (final0_pred(mid, f, strict, m, c1h, c2)) lo = mid; else hi = mid - 1;
    }
    r = ordered_int2f(lo);
  }
  return __shfl_sync(FULL, r, 0);
}

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
               "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }


}  // namespace reid
