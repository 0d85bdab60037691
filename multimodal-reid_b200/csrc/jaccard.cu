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
                int* __restrict__ csc_row, __half* __restrict__ csc_val) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= N) return;
  const long long o = qoff[row];
  const int n = Qn[row];
  for (int j = lane_id(); j < n; j += 32) {
    const int c = Qc[o + j];
    const long long p = colptr[c] + atomicAdd(cursor + c, 1);
    csc_row[p] = row;
    csc_val[p] = Qv[o + j];
  }
}

int launch_csc_count(const int* Qc, const long long* qoff, const int* Qn, int N, int* colcnt,
                     cudaStream_t s) {
  if (N == 0) return REID_OK;
  csc_count_kernel<<<cdiv(N, 8), 256, 0, s>>>(Qc, qoff, Qn, N, colcnt);
  return check_launch("csc_count_kernel");
}

int launch_csc_fill(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn, int N,
                    const long long* colptr, int* cursor, int* csc_row, __half* csc_val,
                    cudaStream_t s) {
  if (N == 0) return REID_OK;
  csc_fill_kernel<<<cdiv(N, 8), 256, 0, s>>>(Qc, Qv, qoff, Qn, N, colptr, cursor, csc_row, csc_val);
  return check_launch("csc_fill_kernel");
}

constexpr int JC_CHUNK = 256;   // query columns staged per chunk
constexpr int JC_PF = 4;        // column lists prefetched ahead

template <int T>
__global__ void __launch_bounds__(T)
jaccard_mix_kernel(const int* __restrict__ Qc, const __half* __restrict__ Qv,
                   const long long* __restrict__ qoff, const int* __restrict__ Qn,
                   const long long* __restrict__ colptr, const int* __restrict__ csc_row,
                   const __half* __restrict__ csc_val, const float* __restrict__ D, long long ld,
                   const int* __restrict__ rowmax_key, int q0, int q_stride, int q, int N, int Npad, float c1,
                   float c2, float* __restrict__ out, long long ld_out,
                   __half* __restrict__ jaccard_out, long long ld_j) {
  extern __shared__ __align__(16) unsigned char smem_j[];
  __half* t = reinterpret_cast<__half*>(smem_j);
  long long* cstart = reinterpret_cast<long long*>(smem_j + (size_t)Npad * 2);
  int* clen = reinterpret_cast<int*>(cstart + JC_CHUNK);
  __half* cval = reinterpret_cast<__half*>(clen + JC_CHUNK);

  const int tid = threadIdx.x;
  const int qi = blockIdx.x;            // local query index (D / rowmax / out rows)
  const long long row = q0 + (long long)qi * q_stride;   // global probe index (V_qe row)
  {
    uint4 z = make_uint4(0, 0, 0, 0);
    uint4* t4 = reinterpret_cast<uint4*>(t);
    for (int j = tid; j < (Npad >> 3); j += T) t4[j] = z;
  }
  const long long o = qoff[row];
  const int nq = Qn[row];

  for (int base = 0; base < nq; base += JC_CHUNK) {
    const int nc = min(JC_CHUNK, nq - base);
    __syncthreads();
    if (tid < nc) {
      const int c = Qc[o + base + tid];
      const long long a = colptr[c];
      cstart[tid] = a;
      clen[tid] = (int)(colptr[c + 1] - a);
      cval[tid] = Qv[o + base + tid];
    }
    __syncthreads();

    int g_r[JC_PF];
    __half v_r[JC_PF];
#pragma unroll
    for (int s = 0; s < JC_PF; ++s) {
      g_r[s] = -1;
      v_r[s] = __ushort_as_half(0);
      if (s < nc && tid < clen[s]) {
        g_r[s] = csc_row[cstart[s] + tid];
        v_r[s] = csc_val[cstart[s] + tid];
      }
    }
    for (int e0 = 0; e0 < nc; e0 += JC_PF) {
#pragma unroll
      for (int s = 0; s < JC_PF; ++s) {
        const int e = e0 + s;
        if (e < nc) {
          const int g = g_r[s];
          const __half vg = v_r[s];
          const __half vi = cval[e];
          const int len = clen[e];
          const long long st = cstart[e];
          // refill this slot with column e + PF
          g_r[s] = -1;
          if (e + JC_PF < nc && tid < clen[e + JC_PF]) {
            g_r[s] = csc_row[cstart[e + JC_PF] + tid];
            v_r[s] = csc_val[cstart[e + JC_PF] + tid];
          }
          if (g >= 0) t[g] = h_add(t[g], h_min(vi, vg));
          for (int p = tid + T; p < len; p += T) {
            const int g2 = csc_row[st + p];
            t[g2] = h_add(t[g2], h_min(vi, csc_val[st + p]));
          }
          __syncthreads();
        }
      }
    }
  }
  __syncthreads();

  // lambda mix over the gallery columns [q, N)
  const float m = ordered_int2f(rowmax_key[qi]);
  const float* drow = D + (long long)qi * ld;
  const __half hc1 = __float2half_rn(c1);
  for (int g = q + tid; g < N; g += T) {
    const __half J = jaccard_from_overlap(t[g]);
    const __half Jl = __float2half_rn(__fmul_rn(__half2float(J), __half2float(hc1)));
    const float ol = __fmul_rn(__fdiv_rn(drow[g], m), c2);
    out[(long long)qi * ld_out + (g - q)] = __fadd_rn(__half2float(Jl), ol);
  }
  if (jaccard_out) {
    for (int g = tid; g < N; g += T) jaccard_out[(long long)qi * ld_j + g] = jaccard_from_overlap(t[g]);
  }
}

int launch_jaccard_mix(const int* Qc, const __half* Qv, const long long* qoff, const int* Qn,
                       const long long* colptr, const int* csc_row, const __half* csc_val,
                       const float* D, long long ld, const int* rowmax_key, int q0, int q_stride,
                       int nq_local, int q, int N, float c1, float c2, float* out, long long ld_out, __half* jaccard_out,
                       long long ld_j, cudaStream_t s) {
  if (nq_local == 0) return REID_OK;
  const int Npad = (int)round_up(N, 8);
  const size_t smem = (size_t)Npad * 2 + JC_CHUNK * (8 + 4 + 2) + 16;
  REID_REQUIRE(smem <= 227 * 1024,
               "jaccard_mix: N=%d needs %zu B of shared memory for the per-query overlap row "
               "(limit 227 KB => N <= ~114k); larger galleries need the streaming variant",
               N, smem);
  if (N <= 32768) {
    REID_CUDA(cudaFuncSetAttribute(jaccard_mix_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
    jaccard_mix_kernel<256><<<nq_local, 256, smem, s>>>(Qc, Qv, qoff, Qn, colptr, csc_row, csc_val, D,
                                                        ld, rowmax_key, q0, q_stride, q, N, Npad, c1, c2, out,
                                                        ld_out, jaccard_out, ld_j);
  } else {
    REID_CUDA(cudaFuncSetAttribute(jaccard_mix_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
    jaccard_mix_kernel<512><<<nq_local, 512, smem, s>>>(Qc, Qv, qoff, Qn, colptr, csc_row, csc_val, D,
                                                        ld, rowmax_key, q0, q_stride, q, N, Npad, c1, c2, out,
                                                        ld_out, jaccard_out, ld_j);
  }
  return check_launch("jaccard_mix_kernel");
}

}  // namespace reid
