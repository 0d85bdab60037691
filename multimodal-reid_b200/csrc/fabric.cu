// Peer-memory fabric of a multi-rank job: one process per GPU on one NVSwitch box.  Every rank owns one
// "arena" (a plain cudaMalloc allocation, exported with CUDA IPC and mapped by all peers), laid out identically on all
// ranks.  Exchanges are written by kernels straight into the peers' arenas over NVLink -- the producing kernel's own
// stores (neighbour lists, candidate records) or the copy kernels below -- and ordered by a flag barrier that is itself a
// kernel on the stream: no host round trip, no library collective between the stages of one evaluation.
//
//   fabric_barrier_kernel   rank r stores the epoch into flags[r] of every peer (release, system scope) and spins until
//                           its own flags[0..P) have all reached the epoch (acquire).  One block of P threads.
//   fabric_bcast_kernel     arena[offset, offset + bytes) of this rank -> the same range of every peer (16-byte words)
//   fabric_bcast_rows       the rows this rank owns (snake blocks of 256, common.cuh) of a [n, row_bytes] table
//   fabric_bcast_ell        owned rows of the V table in ELL form: count + the used prefix of columns / values
//   fabric_bcast_pool       owned rows' segments [off[i], off[i] + cnt[i]) of a CSR pool (columns int32, values fp16)
#include "common.cuh"
#include "kernels.h"

namespace reid {

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// epoch == 0: the epoch is this rank's own count of barriers, kept in device memory beside its flags (word 32) and
// advanced by the kernel itself -- every rank runs the same sequence of barriers, so the counts agree everywhere, and the
// launch has no per-call argument: a step containing barriers can be captured in a CUDA graph and replayed.
__global__ void fabric_barrier_kernel(unsigned char* const* __restrict__ peers, long long flags_off, int rank, int P,
                                      unsigned epoch) {
  const int s = threadIdx.x;
  __shared__ unsigned e_s;
  if (epoch == 0) {
    if (s == 0) {
      unsigned* cnt = reinterpret_cast<unsigned*>(peers[rank] + flags_off) + 32;
      e_s = *cnt + 1;
      *cnt = e_s;
    }
    __syncthreads();
    epoch = e_s;
  }
  if (s < P) {
    __threadfence_system();          // everything this GPU wrote before the barrier is visible before the flag is
    st_release_sys(reinterpret_cast<unsigned*>(peers[s] + flags_off) + rank, epoch);
    const unsigned* mine = reinterpret_cast<const unsigned*>(peers[rank] + flags_off) + s;
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) __nanosleep(20);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
fabric_bcast_kernel(unsigned char* const* __restrict__ peers, int rank, int P, long long offset, long long words) {
  const uint4* src = reinterpret_cast<const uint4*>(peers[rank] + offset);
  const int per_peer = gridDim.x / (P - 1);                 // blocks working for one destination
  const int which = blockIdx.x / per_peer, b = blockIdx.x % per_peer;
  if (which >= P - 1) return;
  const int dstr = which < rank ? which : which + 1;
  uint4* dst = reinterpret_cast<uint4*>(peers[dstr] + offset);
  for (long long i = (long long)b * blockDim.x + threadIdx.x; i < words; i += (long long)per_peer * blockDim.x)
    dst[i] = src[i];
}

// one block per (owned tile row, destination): 256 rows x row_bytes contiguous bytes
__global__ void __launch_bounds__(256)
fabric_bcast_rows_kernel(unsigned char* const* __restrict__ peers, int rank, int P, long long offset, int row_bytes,
                         int n, int ntl) {
  const int lt = blockIdx.x % ntl, which = blockIdx.x / ntl;
  const int dstr = which < rank ? which : which + 1;
  const long long r0 = (long long)snake_global_tile(lt, rank, P) * 256;
  const long long r1 = r0 + 256 < n ? r0 + 256 : n;
  if (r0 >= n) return;
  const long long a = offset + r0 * row_bytes, bytes = (r1 - r0) * row_bytes;
  const unsigned char* src = peers[rank] + a;
  unsigned char* dst = peers[dstr] + a;
  if (((a | bytes) & 15) == 0) {
    for (long long i = threadIdx.x; i < bytes / 16; i += blockDim.x)
      reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
  } else {                                                   // row_bytes is a multiple of 4 (checked by the launcher)
    for (long long i = threadIdx.x; i < bytes / 4; i += blockDim.x)
      reinterpret_cast<unsigned*>(dst)[i] = reinterpret_cast<const unsigned*>(src)[i];
  }
}

// V rows (ELL: cnt[n], cols[n, cap1] int32, vals[n, cap1] fp16): one warp per (owned row, destination)
__global__ void __launch_bounds__(256)
fabric_bcast_ell_kernel(unsigned char* const* __restrict__ peers, int rank, int P, long long off_cnt, long long off_cols,
                        long long off_vals, int cap1, int n, int nloc) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int l = gw % nloc, which = gw / nloc;
  if (which >= P - 1) return;
  const int dstr = which < rank ? which : which + 1;
  const long long row = own_row(rank, -P, l);
  if (row >= n) return;
  const int cnt = reinterpret_cast<const int*>(peers[rank] + off_cnt)[row];
  const int* sc = reinterpret_cast<const int*>(peers[rank] + off_cols) + row * cap1;
  const __half* sv = reinterpret_cast<const __half*>(peers[rank] + off_vals) + row * cap1;
  int* dc = reinterpret_cast<int*>(peers[dstr] + off_cols) + row * cap1;
  __half* dv = reinterpret_cast<__half*>(peers[dstr] + off_vals) + row * cap1;
  for (int j = lane; j < cnt; j += 32) { dc[j] = sc[j]; dv[j] = sv[j]; }
  if (lane == 0) reinterpret_cast<int*>(peers[dstr] + off_cnt)[row] = cnt;
}

// CSR pool rows (off[n+1] int64 element offsets, identical on every rank; cnt[n]; cols int32; vals fp16).  The pool is
// laid out in global row order, so the rows of one owned tile row are ONE contiguous region [off[r0], off[r1]): a block per
// (owned tile row, destination) copies it with 16-byte stores (the slack between a row's count and its slot travels too --
// cheaper than a warp per row issuing 4- and 2-byte stores: 0.37 ms -> see DESIGN 5 at 8 ranks).
__device__ __forceinline__ void fabric_copy_bytes(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src,
                                                  long long bytes) {
  // dst and src have the same alignment (same offset in identically laid out arenas)
  long long head = (16 - ((uintptr_t)src & 15)) & 15;
  if (head > bytes) head = bytes;
  for (long long i = threadIdx.x; i < head; i += blockDim.x) dst[i] = src[i];
  const long long body = (bytes - head) >> 4;
  const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
  uint4* d4 = reinterpret_cast<uint4*>(dst + head);
  for (long long i = threadIdx.x; i < body; i += blockDim.x) d4[i] = s4[i];
  for (long long i = head + (body << 4) + threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
}

__global__ void __launch_bounds__(256)
fabric_bcast_pool_kernel(unsigned char* const* __restrict__ peers, int rank, int P, long long off_cnt,
                         long long off_cols, long long off_vals, const long long* __restrict__ off, int n, int ntl) {
  const int lt = blockIdx.x % ntl, which = blockIdx.x / ntl;
  const int dstr = which < rank ? which : which + 1;
  const long long r0 = (long long)snake_global_tile(lt, rank, P) * 256;
  if (r0 >= n) return;
  const long long r1 = r0 + 256 < n ? r0 + 256 : n;
  const long long a = off[r0], b = off[r1];
  fabric_copy_bytes(peers[dstr] + off_cols + a * 4, peers[rank] + off_cols + a * 4, (b - a) * 4);
  fabric_copy_bytes(peers[dstr] + off_vals + a * 2, peers[rank] + off_vals + a * 2, (b - a) * 2);
  fabric_copy_bytes(peers[dstr] + off_cnt + r0 * 4, peers[rank] + off_cnt + r0 * 4, (r1 - r0) * 4);
}

int fabric_alloc(long long bytes, void** out) {
  REID_REQUIRE(bytes > 0, "fabric_alloc: %lld bytes", bytes);
  REID_CUDA(cudaMalloc(out, (size_t)bytes));
  return REID_OK;
}
int fabric_free(void* p) {
  REID_CUDA(cudaFree(p));
  return REID_OK;
}
int fabric_export(void* p, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  REID_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  return REID_OK;
}
int fabric_open(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  REID_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return REID_OK;
}
int fabric_close(void* p) {
  REID_CUDA(cudaIpcCloseMemHandle(p));
  return REID_OK;
}

int launch_fabric_barrier(unsigned char* const* peers, long long flags_off, int rank, int world, unsigned epoch,
                          cudaStream_t s) {
  REID_REQUIRE(world >= 1 && world <= 32 && rank >= 0 && rank < world, "fabric_barrier: rank %d of %d", rank, world);
  fabric_barrier_kernel<<<1, 32, 0, s>>>(peers, flags_off, rank, world, epoch);
  return check_launch("fabric_barrier_kernel");
}

int launch_fabric_bcast(unsigned char* const* peers, int rank, int world, long long offset, long long bytes,
                        cudaStream_t s) {
  if (world <= 1 || bytes == 0) return REID_OK;
  REID_REQUIRE(offset % 16 == 0 && bytes % 16 == 0, "fabric_bcast: offset %lld / size %lld must be multiples of 16",
               offset, bytes);
  const long long words = bytes / 16;
  int per_peer = (int)((words + 1023) / 1024);
  const int cap = 592 / (world - 1) > 0 ? 592 / (world - 1) : 1;      // ~4 blocks per SM in total
  if (per_peer > cap) per_peer = cap;
  if (per_peer < 1) per_peer = 1;
  fabric_bcast_kernel<<<per_peer * (world - 1), 256, 0, s>>>(peers, rank, world, offset, words);
  return check_launch("fabric_bcast_kernel");
}

int launch_fabric_bcast_rows(unsigned char* const* peers, int rank, int world, long long offset, int row_bytes, int n,
                             cudaStream_t s) {
  if (world <= 1 || n == 0) return REID_OK;
  REID_REQUIRE(offset % 16 == 0 && row_bytes % 4 == 0, "fabric_bcast_rows: offset %lld / row size %d", offset, row_bytes);
  const int ntl = snake_tiles_below(cdiv(n, 256), rank, world);
  if (ntl == 0) return REID_OK;
  fabric_bcast_rows_kernel<<<ntl * (world - 1), 256, 0, s>>>(peers, rank, world, offset, row_bytes, n, ntl);
  return check_launch("fabric_bcast_rows_kernel");
}

static int owned_rows(int n, int rank, int world) {      // local row slots (the last tile row may be partial)
  const int T = cdiv(n, 256);
  int nloc = snake_tiles_below(T, rank, world) * 256;
  if (T > 0 && snake_owner(T - 1, world) == rank) nloc -= T * 256 - n;
  return nloc;
}

int launch_fabric_bcast_ell(unsigned char* const* peers, int rank, int world, long long off_cnt, long long off_cols,
                            long long off_vals, int cap1, int n, cudaStream_t s) {
  if (world <= 1 || n == 0) return REID_OK;
  const int nloc = owned_rows(n, rank, world);
  if (nloc == 0) return REID_OK;
  const long long warps = (long long)nloc * (world - 1);
  fabric_bcast_ell_kernel<<<cdiv(warps * 32, 256), 256, 0, s>>>(peers, rank, world, off_cnt, off_cols, off_vals, cap1, n,
                                                               nloc);
  return check_launch("fabric_bcast_ell_kernel");
}

int launch_fabric_bcast_pool(unsigned char* const* peers, int rank, int world, long long off_cnt, long long off_cols,
                             long long off_vals, const long long* off, int n, cudaStream_t s) {
  if (world <= 1 || n == 0) return REID_OK;
  const int ntl = snake_tiles_below(cdiv(n, 256), rank, world);
  if (ntl == 0) return REID_OK;
  fabric_bcast_pool_kernel<<<ntl * (world - 1), 256, 0, s>>>(peers, rank, world, off_cnt, off_cols, off_vals, off, n, ntl);
  return check_launch("fabric_bcast_pool_kernel");
}

}  // namespace reid
