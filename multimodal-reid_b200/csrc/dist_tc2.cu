// Distance contraction on CTA pairs (tcgen05.mma.cta_group::2) -- K1/K2/K3: evaluate.py:7-26, reranking.py:36-41.
//
// float16 hi/lo split operands, hi.hi + hi.lo + lo.hi accumulated in float32 in TMEM, fused distance epilogue, one
// 256 x 256 output tile per PAIR of SMs.  Why pairs: the single-CTA predecessor (round 1, retired) streamed
// 96 KB of operands per 64-wide K block into every SM (768 KB per 128 x 256 tile) and ncu / the store ablation showed
// it pinned at the L2 -> SM throughput cap (~12 TB/s) -- the tensor pipe waits for operands, and the 1.6 GB output
// stream competes for the same L2 ports.  A CTA pair shares the B operand: each CTA loads its own 128 rows of A and
// HALF of the 256 B rows (64 KB per K block), the pair's tensor cores read both halves.  Operand traffic per output
// element drops by a third and a ring stage by a third.
//
//   cluster (2,1,1): rank 0 = leader, rank 1 = peer; both run the same tile list (tile index = cluster id + k * clusters)
//   warp 0      TMA producer in BOTH CTAs: A_hi, A_lo rows m0 + 128 rank .., B_hi, B_lo rows n0 + 128 rank .. (4 boxes of
//               128 x 64 fp16 = 64 KB per stage); completion bytes go to the LEADER's full barrier (.cta_group::2)
//   warp 1      TMEM allocation (both CTAs); the leader's lane 0 issues tcgen05.mma.cta_group::2 (M = 256, N = 256,
//               K = 16), tcgen05.commit.multicast frees the ring slot in both CTAs and publishes the accumulator to
//               both epilogues
//   warps 2..9  epilogue of this CTA's 128 rows x 256 columns (two warps per TMEM lane quadrant), as in the single-CTA
//               kernel: fused formula, row maxima, 32-column chunk minima, TMA store of 32 x 32 boxes, mirrored tiles
//               (symmetric mode: tiles mb < nb of the upper triangle) stored transposed straight from registers (one
//               128-byte line per column); the mirrored rows' maxima / chunk minima come from a transposed pass
//               through the warp's staging box.  The peer's warps release the accumulator
//               with remote arrives on the leader's barrier.
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace reid {
namespace tc2 {

#ifndef REID_TC2_STAGES
#define REID_TC2_STAGES 3
#endif
constexpr int CM = 128;                  // rows of A (and of the accumulator) per CTA
constexpr int BM = 2 * CM;               // rows of one pair tile
constexpr int BN = 256;                  // columns of one pair tile
constexpr int HB = BN / 2;               // B rows each CTA stages
constexpr int BK = 64, UK = 16;          // BK * 2 bytes = one 128-byte swizzle row
constexpr int STAGES = REID_TC2_STAGES;
constexpr uint32_t OP_BYTES = CM * BK * 2;                  // 16 KB: one 128 x 64 fp16 box
constexpr uint32_t STAGE_BYTES = 4 * OP_BYTES;              // A_hi, A_lo, B_hi, B_lo
constexpr int EPI_WARPS = 8;
constexpr int EPI_BUFS = STAGES <= 2 ? 2 : 1;               // staging boxes per epilogue warp
constexpr uint32_t EPI_BOX_BYTES = 32 * 32 * 4;
constexpr uint32_t EPI_BYTES = EPI_WARPS * EPI_BUFS * EPI_BOX_BYTES;
constexpr uint32_t RB_BYTES = 2 * BN * 4;
constexpr uint32_t SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + RB_BYTES + 128;
constexpr int THREADS = 64 + 32 * EPI_WARPS;
constexpr int RASTER_GROUP = 8;                             // tile rows per raster group (L2 locality)
static_assert(HB == CM, "A and B boxes share one shape");
static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t n_clusters_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// arrive on a barrier of any CTA of the cluster (shared::cluster address from mapa).  Relaxed: the accumulator
// hand-over is ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync; a release at cluster scope would
// first drain this thread's outstanding global stores (ncu r01c: 24 % of the kernel's stall samples).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
// load into this CTA's shared memory, completion bytes on `cluster_bar` (the leader's barrier)
// (L2 evict_last: the operand panels are re-read by every tile of their row / column block while 1.6 GB of output
// streams through the same cache; measured -3 % cycles at Duke, -2 % at MSMT17 scale; hints on the stores did not help)
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t cluster_bar, uint32_t dst, int c0, int c1,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(cluster_bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
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
// arrives on the barrier at this shared-memory offset in BOTH CTAs once the MMAs issued so far have retired
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void tc_mma_pair_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
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

// K-major operand tile, rows = one 128-byte swizzle span, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);        // start address
  d |= (uint64_t)((8 * BK * 2) >> 4) << 32;      // stride byte offset (between 8-row groups)
  d |= (uint64_t)1 << 46;                        // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}
// kind::f16: A = B = F16, D = F32, both K-major, M = 256 (pair), N = 256
constexpr uint32_t IDESC = (1u << 4) | (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);

__device__ __forceinline__ int scale_exponent(uint32_t amax_bits) {      // shared with prep.cu
  if ((amax_bits & 0x7fffffffu) == 0) return 0;
  int E = (int)((amax_bits >> 23) & 0xff) - 127;
  int e = 14 - E;
  return max(-100, min(100, e));
}

// ---------------------------------------------------------------------------------------------
// tile order.  Full: groups of RASTER_GROUP tile rows, column-major inside a group.  Symmetric (A and B are the same
// rows): only tiles with mb <= nb, same grouping; tiles with mb < nb are also stored transposed.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline long long sym_group_tiles(int g, int T) {
  const int first = RASTER_GROUP * g;
  const int gsz = T - first < RASTER_GROUP ? T - first : RASTER_GROUP;
  return (long long)gsz * (gsz + 1) / 2 + (long long)(T - first - gsz) * gsz;
}
struct TileCursor {          // t only grows between calls
  int g = 0;
  long long base = 0;
  __device__ void locate(long long t, int symmetric, int mblocks, int nblocks, int& mb, int& nb) {
    if (!symmetric) {
      const int per_group = RASTER_GROUP * nblocks;
      const int grp = (int)(t / per_group);
      const int first = grp * RASTER_GROUP;
      const int gsz = min(RASTER_GROUP, mblocks - first);
      const int r = (int)(t - (long long)grp * per_group);
      mb = first + r % gsz;
      nb = r / gsz;
      return;
    }
    const int T = nblocks;
    for (;;) {
      const long long c = sym_group_tiles(g, T);
      if (t < base + c) break;
      base += c;
      ++g;
    }
    const int first = RASTER_GROUP * g;
    const int gsz = min(RASTER_GROUP, T - first);
    int r = (int)(t - base);
    const int stair = gsz * (gsz + 1) / 2;
    if (r < stair) {
      int j = 0;
      while (r >= j + 1) { r -= j + 1; ++j; }      // column first + j holds j + 1 tiles
      nb = first + j;
      mb = first + r;
    } else {
      r -= stair;
      nb = first + gsz + r / gsz;
      mb = first + r % gsz;
    }
  }
};

__device__ unsigned long long g_probe[4];      // clock probe (REID_TC_DBG & 64), see reid_dist_clock_probe
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
dist_tc2_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo,
                const __grid_constant__ CUtensorMap tmOut, const float* __restrict__ ra,
                const float* __restrict__ rb, const uint32_t* __restrict__ a_amax,
                const uint32_t* __restrict__ b_amax, int m, int n, int kblocks, int metric,
                int* __restrict__ rowmax_key, float* __restrict__ chunk_min, int nchunks, int symmetric,
                int ntiles, float* __restrict__ out, long long ld_out, int dbg,
                int* __restrict__ rowmax_key_m, float* __restrict__ chunk_min_m, int cmm_pitch,
                const int2* __restrict__ tile_list, int sh_world) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* epi_base = smem + STAGES * STAGE_BYTES;
  float* rb_s = reinterpret_cast<float*>(epi_base + EPI_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EPI_BYTES + RB_BYTES);
  // bars: STAGES full (used in the leader), STAGES empty, 2 tmem_full, 2 tmem_empty (leader); then the TMEM base
  static_assert((2 * STAGES + 4 + 1) * 8 <= 128, "barrier block overflows its 128 bytes");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  const uint32_t bar0 = smem_u32(bars);
  auto FULL_B = [&](int s) { return bar0 + 8u * s; };
  auto EMPTY_B = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto TFULL_B = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto TEMPTY_B = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int mblocks = (m + BM - 1) / BM, nblocks = (n + BN - 1) / BN;
  const int sym_cursor = symmetric == 3 ? 0 : symmetric;      // mode 3 walks the plain rectangle
  const int tile0 = (int)cluster_id_x(), tile_step = (int)n_clusters_x();

  if ((dbg & 64) && blockIdx.x == 0 && threadIdx.x == 0) { g_probe[0] = clock64(); g_probe[1] = globaltimer_ns(); }
  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();          // swizzle-128B needs a 1024-byte aligned base
    tma_prefetch_desc(&tmAhi); tma_prefetch_desc(&tmAlo); tma_prefetch_desc(&tmBhi);
    tma_prefetch_desc(&tmBlo); tma_prefetch_desc(&tmOut);
    for (int s = 0; s < STAGES; ++s) { mbar_init(FULL_B(s), 1); mbar_init(EMPTY_B(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(TFULL_B(s), 1); mbar_init(TEMPTY_B(s), 2 * EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();                                  // both CTAs' barriers are initialised before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0, phase = 0;
      TileCursor cur;
      const uint64_t pol_ld = l2_policy_evict_last();
      for (int t = tile0; t < ntiles; t += tile_step) {
        int mb, nb;
        if (tile_list) { const int2 tl = __ldg(tile_list + t); mb = tl.x; nb = tl.y; }
        else cur.locate(t, sym_cursor, mblocks, nblocks, mb, nb);
        const int m0 = mb * BM + (int)rank * CM, n0 = nb * BN + (int)rank * HB;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(EMPTY_B(stage), phase ^ 1);
          const uint32_t base = smem_u32(smem) + stage * STAGE_BYTES;
          const uint32_t full = mapa(FULL_B(stage), 0);
          if (rank == 0) mbar_expect_tx(FULL_B(stage), 2 * STAGE_BYTES);     // both CTAs' boxes
          tma_load_2d(&tmAhi, full, base, kb * BK, m0, pol_ld);
          tma_load_2d(&tmAlo, full, base + OP_BYTES, kb * BK, m0, pol_ld);
          tma_load_2d(&tmBhi, full, base + 2 * OP_BYTES, kb * BK, n0, pol_ld);
          tma_load_2d(&tmBlo, full, base + 3 * OP_BYTES, kb * BK, n0, pol_ld);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader only)
    if (lane == 0 && rank == 0) {
      int stage = 0, phase = 0, it = 0;
      for (int t = tile0; t < ntiles; t += tile_step, ++it) {
        const int as = it & 1, aph = (it >> 1) & 1;
        mbar_wait(TEMPTY_B(as), aph ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(FULL_B(stage), phase);
          tc_fence_after();
          const uint32_t base = smem_u32(smem) + stage * STAGE_BYTES;
          const uint64_t a_hi = smem_desc_sw128(base), a_lo = smem_desc_sw128(base + OP_BYTES);
          const uint64_t b_hi = smem_desc_sw128(base + 2 * OP_BYTES), b_lo = smem_desc_sw128(base + 3 * OP_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t adv = (uint64_t)((k * UK * 2) >> 4);      // +32 B per K step inside the swizzle atom
            tc_mma_pair_f16(tmem_d, a_hi + adv, b_hi + adv, IDESC, (kb | k) != 0);
            tc_mma_pair_f16(tmem_d, a_hi + adv, b_lo + adv, IDESC, 1);
            tc_mma_pair_f16(tmem_d, a_lo + adv, b_hi + adv, IDESC, 1);
          }
          tc_commit_pair(EMPTY_B(stage));                     // ring slot reusable in both CTAs
          if (kb == kblocks - 1) tc_commit_pair(TFULL_B(as)); // accumulator complete in both CTAs
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may read
    const int ew = warp - 2;                   // 0..7
    const int c_first = (ew >> 2) * (BN / 64); // this warp's four 32-column chunks of the tile
    const int et = threadIdx.x - 64;           // 0..255 = column of the tile this thread stages rb for
    static_assert(BN == 32 * EPI_WARPS, "one rb value per epilogue thread");
    uint8_t* my_epi = epi_base + ew * EPI_BUFS * EPI_BOX_BYTES;
    const float inv_scale = exp2f((float)(-(scale_exponent(*a_amax) + scale_exponent(*b_amax))));
    const float neg2 = -2.0f * inv_scale;
    const uint32_t tempty_leader0 = mapa(TEMPTY_B(0), 0), tempty_leader1 = mapa(TEMPTY_B(1), 0);
    const int row_in_tile = (int)rank * CM + quad * 32;
    int it = 0, nstore = 0;
    TileCursor cur;
    int mb = 0, nb = 0;
    // sharded symmetric mode (tile_list != NULL): this rank computes the upper-triangle tiles of ITS tile rows (snake
    // ownership, common.cuh); tile coordinates are global, A and B are all N rows, `out` / chunk_min hold only the
    // rank's own rows (local row = snake_local_tile(mb) * 256 + row in tile), rowmax_key is indexed by global row and
    // the mirrored rows' chunk minima go to chunk_min_m[global row, local tile row * 8 + 32-row group] (a compact prefix
    // per row: the tile rows this rank owns above it)
    auto locate = [&](int t, int& mb_, int& nb_) {
      if (tile_list) { const int2 tl = __ldg(tile_list + t); mb_ = tl.x; nb_ = tl.y; }
      else cur.locate(t, sym_cursor, mblocks, nblocks, mb_, nb_);
    };
    if (tile0 < ntiles) locate(tile0, mb, nb);
    // the row / column norms of a tile are fetched one tile ahead
    float rb_next = (nb * BN + et < n) ? rb[nb * BN + et] : 0.0f;
    float ra_next = (mb * BM + row_in_tile + lane < m) ? ra[mb * BM + row_in_tile + lane] : 0.0f;
    for (int t = tile0; t < ntiles; t += tile_step, ++it) {
      const int r0 = mb * BM + row_in_tile;    // first (global) row of this warp's 32
      const int lr0 = tile_list ? snake_local_tile(mb, sh_world) * BM + row_in_tile : r0;   // ... and where it is stored
      const int n0 = nb * BN;
      const bool mirror = symmetric == 3 || (symmetric && mb < nb);
      const bool store_mirror = symmetric == 1;   // mode 2: consumers read the lower triangle transposed
      const int as = it & 1, aph = (it >> 1) & 1;
      float* rbt = rb_s + (it & 1) * BN;
      rbt[et] = rb_next;
      const float rav = ra_next;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const int row = r0 + lane;
      const bool row_ok = row < m;
      {
        const int tn = t + tile_step;
        if (tn < ntiles) {
          locate(tn, mb, nb);
          rb_next = (nb * BN + et < n) ? rb[nb * BN + et] : 0.0f;
          ra_next = (mb * BM + row_in_tile + lane < m) ? ra[mb * BM + row_in_tile + lane] : 0.0f;
        }
      }
      float mx = -INFINITY;
      float cm4[BN / 64];                       // minima of this warp's four chunks: one 16-byte store per row and tile
#pragma unroll
      for (int k = 0; k < BN / 64; ++k) cm4[k] = INFINITY;
      mbar_wait(TFULL_B(as), aph);
      tc_fence_after();
      const uint32_t tbase = tmem_base + as * BN + ((uint32_t)(quad * 32) << 16);
#pragma unroll
      for (int cc = 0; cc < BN / 64; ++cc) {
        const int c = c_first + cc;
        uint32_t v[32];
        tmem_ld32(tbase + c * 32, v);
        tmem_ld_wait();
        if (cc == BN / 64 - 1) {                // this warp's part of the accumulator is drained
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(as ? tempty_leader1 : tempty_leader0);
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
        if (col0 >= n || r0 >= m) continue;                    // box entirely outside the matrix
        const int ncols = min(32, n - col0);                   // warp-uniform
        float cmn = INFINITY;
        if (ncols == 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j) { mx = fmaxf(mx, o[j]); cmn = fminf(cmn, o[j]); }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < ncols) { mx = fmaxf(mx, o[j]); cmn = fminf(cmn, o[j]); }
        }
        cm4[cc] = cmn;
        uint8_t* buf = my_epi + (EPI_BUFS == 2 ? (nstore & 1) : 0) * EPI_BOX_BYTES;
        if (lane == 0) {
          if (EPI_BUFS == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncwarp();
        float tmx = -INFINITY, tmn = INFINITY;
        if (mirror) {
          // mirrored rows' maximum / chunk minimum: the box goes through the staging buffer transposed (word (j, lane)
          // in row j, 16-byte chunks swizzled by row), lane L reads row L back = column col0 + L over the warp's rows
#pragma unroll
          for (int j = 0; j < 32; ++j)
            *reinterpret_cast<float*>(buf + j * 128 + (((lane >> 2) ^ (j & 7)) << 4) + ((lane & 3) << 2)) = o[j];
          __syncwarp();
          const int rows_ok = m - r0;                            // valid rows of the direct box (>= 1 here)
          if (rows_ok >= 32) {
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              const float4 x = *reinterpret_cast<const float4*>(buf + lane * 128 + ((c4 ^ (lane & 7)) << 4));
              tmx = fmaxf(fmaxf(tmx, fmaxf(x.x, x.y)), fmaxf(x.z, x.w));
              tmn = fminf(fminf(tmn, fminf(x.x, x.y)), fminf(x.z, x.w));
            }
          } else {
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              const float4 x = *reinterpret_cast<const float4*>(buf + lane * 128 + ((c4 ^ (lane & 7)) << 4));
              const float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (c4 * 4 + k < rows_ok) { tmx = fmaxf(tmx, e[k]); tmn = fminf(tmn, e[k]); }
            }
          }
          __syncwarp();
        }
        {
          uint8_t* rowp = buf + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int chunk = j ^ (lane & 7);
            *reinterpret_cast<float4*>(rowp + chunk * 16) = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmOut, smem_u32(buf), col0, lr0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++nstore;
        }
        if (mirror) {
          // the same box transposed, straight from registers: out[col0 + j, r0 + lane] = o[j], one 128-byte line per j
          float* orow = out + (long long)col0 * ld_out + row;
          if (!store_mirror) {
          } else if (ncols == 32 && row_ok) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              *orow = o[j];
              orow += ld_out;
            }
          } else if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) orow[(long long)j * ld_out] = o[j];
          }
          const int trow = col0 + lane;
          if (trow < n) {
            if (chunk_min_m) chunk_min_m[(long long)trow * cmm_pitch + (lr0 >> 5)] = tmn;
            if (rowmax_key_m) atomicMax(rowmax_key_m + trow, f2ordered_int(tmx));
          }
        }
      }
      if (chunk_min && row_ok)           // nchunks = 8 slots per tile column block: (n0 >> 5) + c_first is a multiple of 4
        *reinterpret_cast<float4*>(chunk_min + (long long)(lr0 + lane) * nchunks + (n0 >> 5) + c_first) =
            make_float4(cm4[0], cm4[1], cm4[2], cm4[3]);
      if (rowmax_key && row_ok && mx > -INFINITY) atomicMax(rowmax_key + row, f2ordered_int(mx));
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync();            // the peer's shared memory and the leader's barriers stay alive until both CTAs are done
  if ((dbg & 64) && blockIdx.x == 0 && threadIdx.x == 0) { g_probe[2] = clock64(); g_probe[3] = globaltimer_ns(); }
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
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
                    long long rows, long long ld_elems, int box_cols, int box_rows, const char* what) {
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
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(%s) failed with CUresult %d (base %p, %lld x %lld, ld %lld)", what, (int)r,
              base, rows, cols, ld_elems);
    return REID_ERR_CUDA;
  }
  return REID_OK;
}

}  // namespace tc2

int dist_clock_probe(long long* host_out4) {
  unsigned long long v[4];
  REID_CUDA(cudaMemcpyFromSymbol(v, tc2::g_probe, sizeof(v)));
  for (int i = 0; i < 4; ++i) host_out4[i] = (long long)v[i];
  return REID_OK;
}

int launch_dist_tc(const __half* Ahi, const __half* Alo, long long lda, const float* ra, const uint32_t* a_amax,
                   int m, const __half* Bhi, const __half* Blo, long long ldb, const float* rb,
                   const uint32_t* b_amax, int n, int dp, int metric, float* out, long long ld_out,
                   int* rowmax_key, float* chunk_min, int sym_mode, cudaStream_t s) {
  // sym_mode (per call, no process state): 0 every tile; 1 upper triangle computed + mirrored stores; 2 upper triangle
  // only.  1 and 2 apply when both operands are the same rows; a rectangle of two different row sets runs in full.
  REID_REQUIRE(sym_mode >= 0 && sym_mode <= 2, "dist_tc: sym_mode %d outside 0..2", sym_mode);
  return launch_dist_tc_block(Ahi, Alo, lda, ra, a_amax, m, Bhi, Blo, ldb, rb, b_amax, n, dp, metric, out, ld_out,
                              rowmax_key, chunk_min, 0, -1 - sym_mode, rowmax_key, chunk_min, s);
}

// One block of a larger distance matrix (the streamed evaluator computes D chunk by chunk while the features are still
// arriving).  mode -1 - sym (sym = 0, 1, 2): launch_dist_tc's whole-matrix call (symmetric schedule `sym` when both
// operands are the same rows); 0: plain rectangle; 2: square diagonal block, upper-only form; 3: rectangle that lies
// strictly above the diagonal of the enclosing symmetric matrix -- stored as is, and every tile also contributes the
// row maxima / chunk minima of its mirror image through rowmax_key_m / chunk_min_m (indexed by the block's columns).
// cm_pitch: row pitch of the chunk-minimum arrays (0: chunk_pitch(n)); all block origins must be multiples of 256.
int launch_dist_tc_block(const __half* Ahi, const __half* Alo, long long lda, const float* ra, const uint32_t* a_amax,
                         int m, const __half* Bhi, const __half* Blo, long long ldb, const float* rb,
                         const uint32_t* b_amax, int n, int dp, int metric, float* out, long long ld_out,
                         int* rowmax_key, float* chunk_min, int cm_pitch, int mode, int* rowmax_key_m,
                         float* chunk_min_m, cudaStream_t s) {
  using namespace tc2;
  if (m == 0 || n == 0) return REID_OK;
  REID_REQUIRE(dp % BK == 0, "dist_tc: padded width %d must be a multiple of %d", dp, BK);
  REID_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ld_out % 4 == 0,
               "dist_tc: row pitches must keep rows 16-byte aligned (lda=%lld ldb=%lld ld_out=%lld)", lda, ldb,
               ld_out);
  CUtensorMap mAhi, mAlo, mBhi, mBlo, mOut;
  int rc;
  if ((rc = make_map(&mAhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Ahi, dp, m, lda, BK, CM, "A_hi"))) return rc;
  if ((rc = make_map(&mAlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Alo, dp, m, lda, BK, CM, "A_lo"))) return rc;
  if ((rc = make_map(&mBhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Bhi, dp, n, ldb, BK, HB, "B_hi"))) return rc;
  if ((rc = make_map(&mBlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Blo, dp, n, ldb, BK, HB, "B_lo"))) return rc;
  if ((rc = make_map(&mOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, n, m, ld_out, 32, 32, "out"))) return rc;
  int dev = 0, sms = 0;
  REID_CUDA(cudaGetDevice(&dev));
  REID_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  REID_CUDA(cudaFuncSetAttribute(dist_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  // same rows on both sides and a square output: compute the upper triangle only and mirror it
  const bool same = Ahi == Bhi && Alo == Blo && lda == ldb && ra == rb && a_amax == b_amax && m == n;
  REID_REQUIRE(mode >= -3 && mode <= 3 && mode != 1, "dist_tc: unknown block mode %d", mode);
  REID_REQUIRE(mode != 2 || same, "dist_tc: the upper-only form needs the same rows on both sides");
  const int symmetric = mode >= 0 ? mode : (same ? -1 - mode : 0);
  const int pitch = cm_pitch > 0 ? cm_pitch : chunk_pitch(n);
  REID_REQUIRE(pitch % 4 == 0 && ((uintptr_t)chunk_min & 15) == 0, "dist_tc: chunk minima must be 16-byte aligned");
  const int mblocks = cdiv(m, BM), nblocks = cdiv(n, BN);
  long long ntiles = (long long)mblocks * nblocks;
  if (symmetric == 1 || symmetric == 2) {
    ntiles = 0;
    for (int g = 0; g * RASTER_GROUP < nblocks; ++g) ntiles += sym_group_tiles(g, nblocks);
  }
  REID_REQUIRE(ntiles < (1ll << 30), "dist_tc: too many tiles");
  const int pairs = sms / 2;
  const int clusters = ntiles < pairs ? (int)ntiles : pairs;
  const char* dbg_env = getenv("REID_TC_DBG");          // diagnostics only (scratch/dist_knobs.py)
  const int dbg = dbg_env ? atoi(dbg_env) : 0;
  dist_tc2_kernel<<<2 * clusters, THREADS, SMEM_BYTES, s>>>(mAhi, mAlo, mBhi, mBlo, mOut, ra, rb, a_amax, b_amax, m, n,
                                                            dp / BK, metric, rowmax_key, chunk_min, pitch,
                                                            symmetric, (int)ntiles, out, ld_out, dbg, rowmax_key_m,
                                                            chunk_min_m, pitch, nullptr, 1);
  return check_launch("dist_tc2_kernel");
}

// Sharded symmetric contraction of a multi-rank job: this rank computes the upper-triangle tiles (mb <= nb) of the
// tile rows it owns under the snake scheme (common.cuh) -- `tiles` is that list in global tile coordinates, built by
// the host (engine.py, cached per (N, world, rank)).  A = B = all n rows.  Stored: out[nloc, ld_out] = the rank's own
// rows in upper-only form (columns left of a row's 256-wide diagonal block are never written); chunk_min[nloc, cm_pitch]
// their chunk minima (defined from the diagonal block on); rowmax_key[n] (global rows) receives atomicMax from direct
// AND mirrored tiles -- the maximum over the part of each row this rank has seen; chunk_min_m[n, cmm_pitch] the chunk
// minima of the mirrored tiles, row i holding 8 * (number of owned tile rows above i's) valid leading entries.
int launch_dist_tc_sharded(const __half* hi, const __half* lo, long long ld, const float* sq, const uint32_t* amax,
                           int n, int dp, const int* tiles, int ntiles, int world, float* out, long long ld_out,
                           int nloc, int* rowmax_key, float* chunk_min, int cm_pitch, float* chunk_min_m,
                           int cmm_pitch, cudaStream_t s) {
  using namespace tc2;
  if (n == 0 || ntiles == 0) return REID_OK;
  REID_REQUIRE(dp % BK == 0 && ld % 8 == 0 && ld_out % 4 == 0, "dist_tc_sharded: bad pitches");
  REID_REQUIRE(cm_pitch % 4 == 0 && cm_pitch >= chunk_pitch(n) && ((uintptr_t)chunk_min & 15) == 0 && cmm_pitch >= 8,
               "dist_tc_sharded: chunk-minimum arrays must be 16-byte aligned with pitch >= %d", chunk_pitch(n));
  CUtensorMap mAhi, mAlo, mBhi, mBlo, mOut;
  int rc;
  if ((rc = make_map(&mAhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, hi, dp, n, ld, BK, CM, "A_hi"))) return rc;
  if ((rc = make_map(&mAlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, lo, dp, n, ld, BK, CM, "A_lo"))) return rc;
  if ((rc = make_map(&mBhi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, hi, dp, n, ld, BK, HB, "B_hi"))) return rc;
  if ((rc = make_map(&mBlo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, lo, dp, n, ld, BK, HB, "B_lo"))) return rc;
  if ((rc = make_map(&mOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, n, nloc, ld_out, 32, 32, "out"))) return rc;
  int dev = 0, sms = 0;
  REID_CUDA(cudaGetDevice(&dev));
  REID_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  REID_CUDA(cudaFuncSetAttribute(dist_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  const int pairs = sms / 2;
  const int clusters = ntiles < pairs ? ntiles : pairs;
  dist_tc2_kernel<<<2 * clusters, THREADS, SMEM_BYTES, s>>>(mAhi, mAlo, mBhi, mBlo, mOut, sq, sq, amax, amax, n, n,
                                                            dp / BK, 0, rowmax_key, chunk_min, cm_pitch, 2, ntiles, out,
                                                            ld_out, 0, rowmax_key, chunk_min_m, cmm_pitch,
                                                            reinterpret_cast<const int2*>(tiles), world);
  return check_launch("dist_tc2_kernel");
}

}  // namespace reid
