// K8 + K9 + K10 (reranking.py:80-100): inverted index of V_qe, the float16 Jaccard chain and the
// lambda mix, fused so that the q x N Jaccard matrix never exists in memory.
//
//   * csc_count / csc_fill: CSR -> CSC of V_qe over all N rows (count, scan, scatter).  The order
//     of entries inside one column is irrelevant to the arithmetic (each target row receives at
//     most one contribution per column), so the scatter uses an atomic cursor.
//   * jaccard_mix_kernel: one CTA per query.  The running overlap t[0:N] lives in shared memory
//     as float16.  Query columns are visited in ascending order with a CTA barrier between them,
//     so every t[g] sees its additions in exactly the reference's order and rounding
//     (fp16 RN after every add).  Column lists are prefetched PF columns ahead so the barrier
//     cadence, not DRAM/L2 latency, sets the pace.  The epilogue streams the query's row of D
//     once and writes final = fp16(J * fp16(1-lambda)) + O * fp32(lambda) for gallery columns.
#include "common.cuh"
#include "np_exact.cuh"
#include "kernels.h"

namespace reid {

__global__ void ell_offsets_kernel(int N, int cap1, long long* __restrict__ off) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= N) off[i] = (long long)i * cap1;
}

int launch_ell_offsets(int N, int cap1, long long* off, cudaStream_t s) {
  ell_offsets_kernel<<<cdiv(N + 1, 256), 256, 0, s>>>(N, cap1, off);
  return check_launch("ell_offsets_kernel");
}

__global__ void __launch_bounds__(256)
csc_count_kernel(const int* __restrict__ Qc, const long long* __restrict__ qoff,
                 const int* __restrict__ Qn, int N, int* __restrict__ colcnt) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= N) return;
  const long long o = qoff[row];
  const int n = Qn[row];
  for (int j = lane_id(); j < n; j += 32) atomicAdd(colcnt + Qc[o + j], 1);
}

__global__ void __launch_bounds__(256)
csc_fill_kernel(const int* __restrict__ Qc, const __half* __restrict__ Qv,
                const long long* __restrict__ qoff, const int* __restrict__ Qn, int N,
                const long long* __restrict__ colptr, int* __restrict__ cursor,
                unsigned long long* __restrict__ csc_ent) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= N) return;
  const long long o = qoff[row];
  const int n = Qn[row];
  for (int j = lane_id(); j < n; j += 32) {
    const int c = Qc[o + j];
    const long long p = colptr[c] + atomicAdd(cursor + c, 1);
    csc_ent[p] = (unsigned long long)(unsigned)row | ((unsigned long long)__half_as_ushort(Qv[o + j]) << 32);
  }
}

int launch_csc_count(const int* Qc, const long long* qoff, const int* Qn, int N, int* colcnt,
                     cudaStream_t s) {
  if (N == 0) return REID_OK;
  csc_count_kernel<<<cdiv(N, 8), 256, 0, s>>>(Qc, qoff, Qn, N, colcnt);
  return check_launch("csc_count_kernel");
}

int launch_csc_fill(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn, int N,
                    const long long* colptr, int* cursor, unsigned long long* csc_ent, cudaStream_t s) {
  if (N == 0) return REID_OK;
  csc_fill_kernel<<<cdiv(N, 8), 256, 0, s>>>(Qc, Qv, qoff, Qn, N, colptr, cursor, csc_ent);
  return check_launch("csc_fill_kernel");
}

}  // namespace reid
