// Distance contraction on the 5th-generation tensor cores (K1/K2/K3: evaluate.py:7-26,
// reranking.py:36-41).
//
// Operands are the float16 split of the (optionally L2-normalised) features produced by prep.cu:
// x * 2^e = hi + lo with hi = fp16(x * 2^e), lo = fp16(x * 2^e - hi), e chosen per matrix so that
// max|x| * 2^e lies in [2^14, 2^15).  The dot product is accumulated in float32 in TMEM as
//     hi_a.hi_b + hi_a.lo_b + lo_a.hi_b            (three kind::f16 MMAs per K step)
// which carries ~22 mantissa bits per factor (the dropped lo.lo term is 2^-22 relative), i.e. the
// reference's float32 sgemm accuracy at twice the TF32 MMA rate.  Powers of two scale exactly, so
// the epilogue's  (ra + rb) - 2 * acc * 2^-(ea+eb)  rounds exactly like the reference's addmm_.
//
// Structure (one CTA per SM, persistent over 128 x 256 output tiles):
//   warp 0      TMA producer: 4 boxes per K block (A_hi, A_lo: 128 x BK; B_hi, B_lo: 256 x BK fp16,
//               rows = one swizzle span) into a 192 KB shared-memory ring (2 stages of BK = 64; 4 of BK = 32
//               measured 7 % slower: the kernel is bound by shared-memory bandwidth, not latency),
//               completion on mbarriers
//   warp 1      allocates TMEM (2 accumulators x 256 columns) and issues tcgen05.mma
//               (cta_group::1, M=128, N=256, K=16) from one lane; tcgen05.commit frees ring slots
//               and publishes the finished accumulator
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns, fused distance formula, running row
//               maximum (atomicMax on an ordered-int key: reranking.py:46 needs max_k D[k, i]) and the
//               minimum of every 32-column chunk (lets the top-k selection skip ~97 % of the matrix),
//               128B-swizzled staging in shared memory, TMA store of 32 x 32 boxes (clipped at the
//               matrix edge by the tensor map), overlapped with the next tile's MMAs through the
//               second accumulator.
#include <cuda.h>

#include "common.cuh"
#include "kernels.h"

namespace reid {
namespace tc {

#ifndef REID_TC_BK
#define REID_TC_BK 64
#endif
constexpr int BM = 128, BN = 256, BK = REID_TC_BK, UK = 16;     // BK * 2 bytes = the TMA / UMMA swizzle width
constexpr int STAGES = BK == 64 ? 2 : 4;
static_assert(BK == 64 || BK == 32, "K block must be one 128-byte or one 64-byte swizzle row");
constexpr uint32_t A_BYTES = BM * BK * 2;                 // 16 KB
constexpr uint32_t B_BYTES = BN * BK * 2;                 // 32 KB
constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;   // 96 KB
constexpr uint32_t EPI_BOX_BYTES = 32 * 32 * 4;           // one warp's 32 x 32 float box
constexpr uint32_t EPI_BYTES = 4 * 2 * EPI_BOX_BYTES;     // 4 warps x 2 buffers
constexpr uint32_t RB_BYTES = 2 * BN * 4;
constexpr uint32_t SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + RB_BYTES + 128;
constexpr int THREADS = 192;
constexpr int RASTER_GROUP = 8;                           // m-blocks per raster group (L2 locality)

static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t bar, uint32_t dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile whose rows are exactly one swizzle span (BK * 2 bytes: 128 B or 64 B);
// 8-row groups are 8 * span bytes apart.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);        // start address
  d |= (uint64_t)((8 * BK * 2) >> 4) << 32;      // stride byte offset (between 8-row groups)
  d |= (uint64_t)1 << 46;                        // descriptor version (sm_100)
  d |= (uint64_t)(BK == 64 ? 2 : 4) << 61;       // SWIZZLE_128B / SWIZZLE_64B
  return d;
}
// kind::f16: A = B = F16, D = F32, both K-major, M = 128, N = 256
constexpr uint32_t IDESC = (1u << 4) | (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);

// power-of-two scale exponent shared with prep.cu: max|x| * 2^e in [2^14, 2^15)
__device__ __forceinline__ int scale_exponent(uint32_t amax_bits) {
  if ((amax_bits & 0x7fffffffu) == 0) return 0;
  int E = (int)((amax_bits >> 23) & 0xff) - 127;
  int e = 14 - E;
  return max(-100, min(100, e));
}

__device__ __forceinline__ void tile_coords(int t, int mblocks, int nblocks, int& mb, int& nb) {
  const int per_group = RASTER_GROUP * nblocks;
  const int grp = t / per_group;
  const int first = grp * RASTER_GROUP;
  const int gsz = min(RASTER_GROUP, mblocks - first);
  const int r = t - grp * per_group;
  mb = first + r % gsz;
  nb = r / gsz;
}

// Symmetric mode (A and B are the same rows: the N x N matrix of re-ranking).  D[i, j] = D[j, i], so only tiles that
// touch the upper triangle are computed -- (mb, nb) with mb <= 2 nb + 1 in units of BM = 128 rows and BN = 256
// columns -- and every tile strictly above the 256-wide diagonal band (mb < 2 nb) is also stored transposed, with
// the row maxima and 32-column chunk minima of the mirrored rows.  Each element is written exactly once.  Tiles are
// enumerated group by group (RASTER_GROUP m-blocks), column-block-major inside a group, like the full raster.
__host__ __device__ inline int sym_col_count(int g, int gsz, int nb) {       // tiles of group g in column block nb
  const int c = 2 * nb + 2 - RASTER_GROUP * g;
  return c < 0 ? 0 : (c > gsz ? gsz : c);
}
__host__ __device__ inline int sym_first_nb(int g) { return g == 0 ? 0 : (RASTER_GROUP / 2) * g; }
__host__ __device__ inline long long sym_group_tiles(int g, int mblocks, int nblocks) {
  const int gsz = min(RASTER_GROUP, mblocks - RASTER_GROUP * g);
  long long tot = 0;
  int nb = sym_first_nb(g);
  for (int s = 0; s < RASTER_GROUP / 2 && nb < nblocks; ++s, ++nb) tot += sym_col_count(g, gsz, nb);   // staircase
  if (nb < nblocks) tot += (long long)(nblocks - nb) * gsz;
  return tot;
}
struct SymCursor {           // t only grows between calls
  int g = 0;
  long long base = 0;
  __device__ void locate(long long t, int mblocks, int nblocks, int& mb, int& nb) {
    for (;;) {
      const long long c = sym_group_tiles(g, mblocks, nblocks);
      if (t < base + c) break;
      base += c;
      ++g;
    }
    const int gsz = min(RASTER_GROUP, mblocks - RASTER_GROUP * g);
    int r = (int)(t - base);
    nb = sym_first_nb(g);
    for (int s = 0; s < RASTER_GROUP / 2; ++s) {
      const int c = sym_col_count(g, gsz, nb);
      if (r < c) { mb = RASTER_GROUP * g + r; return; }
      r -= c;
      ++nb;
    }
    nb += r / gsz;
    mb = RASTER_GROUP * g + r % gsz;
  }
};

__global__ void __launch_bounds__(THREADS, 1)
dist_tc_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
               const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo,
               const __grid_constant__ CUtensorMap tmOut, const float* __restrict__ ra,
               const float* __restrict__ rb, const uint32_t* __restrict__ a_amax,
               const uint32_t* __restrict__ b_amax, int m, int n, int kblocks, int metric,
               int* __restrict__ rowmax_key, float* __restrict__ chunk_min, int nchunks, int symmetric,
               int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* epi_base = smem + STAGES * STAGE_BYTES;
  float* rb_s = reinterpret_cast<float*>(epi_base + EPI_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EPI_BYTES + RB_BYTES);
  // bars: STAGES full, STAGES empty, 2 tmem_full, 2 tmem_empty; then the TMEM base address
  static_assert((2 * STAGES + 4 + 1) * 8 <= 128, "barrier block overflows its 128 bytes");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  const uint32_t bar0 = smem_u32(bars);
  auto FULL_B = [&](int s) { return bar0 + 8u * s; };
  auto EMPTY_B = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto TFULL_B = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto TEMPTY_B = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mblocks = (m + BM - 1) / BM, nblocks = (n + BN - 1) / BN;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();          // swizzle-128B needs a 1024-byte aligned base
    tma_prefetch_desc(&tmAhi); tma_prefetch_desc(&tmAlo); tma_prefetch_desc(&tmBhi);
    tma_prefetch_desc(&tmBlo); tma_prefetch_desc(&tmOut);
    for (int s = 0; s < STAGES; ++s) { mbar_init(FULL_B(s), 1); mbar_init(EMPTY_B(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(TFULL_B(s), 1); mbar_init(TEMPTY_B(s), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0, phase = 0;
      SymCursor cur;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int mb, nb;
        if (symmetric) cur.locate(t, mblocks, nblocks, mb, nb); else tile_coords(t, mblocks, nblocks, mb, nb);
        const int m0 = mb * BM, n0 = nb * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(EMPTY_B(stage), phase ^ 1);
          const uint32_t base = smem_u32(smem) + stage * STAGE_BYTES;
          mbar_expect_tx(FULL_B(stage), STAGE_BYTES);
          tma_load_2d(&tmAhi, FULL_B(stage), base, kb * BK, m0);
          tma_load_2d(&tmAlo, FULL_B(stage), base + A_BYTES, kb * BK, m0);
          tma_load_2d(&tmBhi, FULL_B(stage), base + 2 * A_BYTES, kb * BK, n0);
          tma_load_2d(&tmBlo, FULL_B(stage), base + 2 * A_BYTES + B_BYTES, kb * BK, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      int stage = 0, phase = 0, it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
        const int as = it & 1, aph = (it >> 1) & 1;
        mbar_wait(TEMPTY_B(as), aph ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(FULL_B(stage), phase);
          tc_fence_after();
          const uint32_t base = smem_u32(smem) + stage * STAGE_BYTES;
          const uint64_t a_hi = smem_desc_sw128(base), a_lo = smem_desc_sw128(base + A_BYTES);
          const uint64_t b_hi = smem_desc_sw128(base + 2 * A_BYTES),
                         b_lo = smem_desc_sw128(base + 2 * A_BYTES + B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t adv = (uint64_t)((k * UK * 2) >> 4);      // +32 B per K step inside the swizzle atom
            tc_mma_f16(tmem_d, a_hi + adv, b_hi + adv, IDESC, (kb | k) != 0);
            tc_mma_f16(tmem_d, a_hi + adv, b_lo + adv, IDESC, 1);
            tc_mma_f16(tmem_d, a_lo + adv, b_hi + adv, IDESC, 1);
          }
          tc_commit(EMPTY_B(stage));                     // ring slot reusable once these MMAs retire
          if (kb == kblocks - 1) tc_commit(TFULL_B(as)); // accumulator complete
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may read
    const int ew = warp - 2;
    const int et = threadIdx.x - 64;           // 0..127
    uint8_t* my_epi = epi_base + ew * 2 * EPI_BOX_BYTES;
    const float inv_scale = exp2f((float)(-(scale_exponent(*a_amax) + scale_exponent(*b_amax))));
    const float neg2 = -2.0f * inv_scale;
    int it = 0, nstore = 0;
    SymCursor cur;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
      int mb, nb;
      if (symmetric) cur.locate(t, mblocks, nblocks, mb, nb); else tile_coords(t, mblocks, nblocks, mb, nb);
      const int m0 = mb * BM, n0 = nb * BN;
      const bool mirror = symmetric && mb < 2 * nb;
      const int as = it & 1, aph = (it >> 1) & 1;
      float* rbt = rb_s + (it & 1) * BN;
      for (int j = et; j < BN; j += 128) rbt[j] = (n0 + j < n) ? rb[n0 + j] : 0.0f;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const int row = m0 + quad * 32 + lane;
      const float rav = row < m ? ra[row] : 0.0f;
      float mx = -INFINITY;
      mbar_wait(TFULL_B(as), aph);
      tc_fence_after();
      const uint32_t tbase = tmem_base + as * BN + ((uint32_t)(quad * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tbase + c * 32, v);
        tmem_ld_wait();
        if (c == BN / 32 - 1) {                 // accumulator drained: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(TEMPTY_B(as));
        }
        const int col0 = n0 + c * 32;
        float o[32];
        if (metric == 0) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            o[j] = __fmaf_rn(neg2, __uint_as_float(v[j]), __fadd_rn(rav, rbt[c * 32 + j]));
        } else {
          const float hi = (float)(1.0 - 0.00001), lo = (float)(-1.0 + 0.00001);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float recip = __fdiv_rn(1.0f, __fmul_rn(rav, rbt[c * 32 + j]));
            float cs = __fmul_rn(__fmul_rn(__uint_as_float(v[j]), inv_scale), recip);
            cs = fminf(fmaxf(cs, lo), hi);
            o[j] = acosf(cs);
          }
        }
        float cmn = INFINITY;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < n) { mx = fmaxf(mx, o[j]); cmn = fminf(cmn, o[j]); }
        if (chunk_min && row < m && col0 < n) chunk_min[(long long)row * nchunks + (col0 >> 5)] = cmn;
        if (col0 < n && m0 + quad * 32 < m) {      // box at least partly inside the matrix
          uint8_t* buf = my_epi + (nstore & 1) * EPI_BOX_BYTES;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t* rowp = buf + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int chunk = j ^ (lane & 7);
            *reinterpret_cast<float4*>(rowp + chunk * 16) = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmOut, smem_u32(buf), col0, m0 + quad * 32);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++nstore;
          if (mirror) {
            // the same box transposed: out[col0 + j, m0 + quad*32 + lane] = o[j]; then lane L reads back row L of
            // the transposed box (= column col0 + L of this tile) for the mirrored row's maximum and chunk minimum
            uint8_t* tb = my_epi + (nstore & 1) * EPI_BOX_BYTES;
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              *reinterpret_cast<float*>(tb + j * 128 + (((lane >> 2) ^ (j & 7)) << 4) + ((lane & 3) << 2)) = o[j];
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmOut, smem_u32(tb), m0 + quad * 32, col0);
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            ++nstore;
            const int rows_ok = m - (m0 + quad * 32);               // valid rows of the direct box (>= 1 here)
            float tmx = -INFINITY, tmn = INFINITY;
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              const float4 v = *reinterpret_cast<const float4*>(tb + lane * 128 + ((c4 ^ (lane & 7)) << 4));
              const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (c4 * 4 + k < rows_ok) { tmx = fmaxf(tmx, e[k]); tmn = fminf(tmn, e[k]); }
            }
            const int trow = col0 + lane;
            if (trow < n) {
              if (chunk_min) chunk_min[(long long)trow * nchunks + ((m0 + quad * 32) >> 5)] = tmn;
              if (rowmax_key) atomicMax(rowmax_key + trow, f2ordered_int(tmx));
            }
          }
        }
      }
      if (rowmax_key && row < m && mx > -INFINITY) atomicMax(rowmax_key + row, f2ordered_int(mx));
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static int make_map(CUtensorMap* map, CUtensorMapDataType dt, int esize, const void* base, long long cols,
                    long long rows, long long ld_elems, int box_cols, int box_rows, const char* what,
                    CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return REID_ERR_UNSUPPORTED;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld_elems * esize};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(%s) failed with CUresult %d (base %p, %lld x %lld, ld %lld)", what, (int)r,
              base, rows, cols, ld_elems);
    return REID_ERR_CUDA;
  }
  return REID_OK;
}

}  // namespace tc

// Single-CTA variant (cta_group::1), kept for A/B measurements against the CTA-pair kernel (dist_tc2.cu):
// REID_TC_PAIR=0 in the environment routes reid_dist here.
int launch_dist_tc1(const __half* Ahi, const __half* Alo, long long lda, const float* ra, const uint32_t* a_amax,
                   int m, const __half* Bhi, const __half* Blo, long long ldb, const float* rb,
                   const uint32_t* b_amax, int n, int dp, int metric, float* out, long long ld_out,
                   int* rowmax_key, float* chunk_min, cudaStream_t s) {
  using namespace tc;
  if (m == 0 || n == 0) return REID_OK;
  REID_REQUIRE(dp % BK == 0, "dist_tc: padded width %d must be a multiple of %d", dp, BK);
  REID_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ld_out % 4 == 0,
               "dist_tc: row pitches must keep rows 16-byte aligned (lda=%lld ldb=%lld ld_out=%lld)", lda, ldb,
               ld_out);
  CUtensorMap mAhi, mAlo, mBhi, mBlo, mOut;
  const CUtensorMapSwizzle OPSWZ = BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  int rc;
  if ((rc = make_map(&mAhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Ahi, dp, m, lda, BK, BM, "A_hi", OPSWZ))) return rc;
  if ((rc = make_map(&mAlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Alo, dp, m, lda, BK, BM, "A_lo", OPSWZ))) return rc;
  if ((rc = make_map(&mBhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Bhi, dp, n, ldb, BK, BN, "B_hi", OPSWZ))) return rc;
  if ((rc = make_map(&mBlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Blo, dp, n, ldb, BK, BN, "B_lo", OPSWZ))) return rc;
  if ((rc = make_map(&mOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, n, m, ld_out, 32, 32, "out"))) return rc;
  int dev = 0, sms = 0;
  REID_CUDA(cudaGetDevice(&dev));
  REID_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  REID_CUDA(cudaFuncSetAttribute(dist_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  // same rows on both sides and a square output: compute the upper triangle only and mirror it
  const int symmetric = (Ahi == Bhi && Alo == Blo && lda == ldb && ra == rb && a_amax == b_amax && m == n &&
                         symmetric_contraction_enabled()) ? 1 : 0;
  const int mblocks = cdiv(m, BM), nblocks = cdiv(n, BN);
  long long ntiles = (long long)mblocks * nblocks;
  if (symmetric) {
    ntiles = 0;
    for (int g = 0; g * RASTER_GROUP < mblocks; ++g) ntiles += sym_group_tiles(g, mblocks, nblocks);
  }
  REID_REQUIRE(ntiles < (1ll << 31), "dist_tc: too many tiles");
  const int grid = ntiles < sms ? (int)ntiles : sms;
  dist_tc_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(mAhi, mAlo, mBhi, mBlo, mOut, ra, rb, a_amax, b_amax, m, n,
                                                  dp / BK, metric, rowmax_key, chunk_min, chunk_pitch(n), symmetric,
                                                  (int)ntiles);
  return check_launch("dist_tc_kernel");
}

}  // namespace reid
