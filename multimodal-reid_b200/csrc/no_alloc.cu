// reid_topk_rows_scatter_static: the scatter selection of select.cu (reid_topk_rows_scatter) for callers that cannot
// allocate device memory -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// The candidate records / thresholds / counters of the three launches live in a file-scope device array instead of a
// caller-provided workspace.  The array holds the workspace of the largest accepted problem, n <= 4 096 rows (8.03 MiB);
// every GPU that loads the library carries it until the process exits.  In HOST code the array's name is only a stand-in:
// its device address is fetched once per device, with that device current, by cudaGetSymbolAddress and that pointer is
// what the kernels get.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
// All calls share the array: the launcher re-initialises what must start clean (group thresholds by cudaMemsetAsync,
// per-row counters by select_thr_kernel), so back-to-back calls on one stream are independent, but two calls that
// overlap on the GPU (different streams, or concurrently replayed graphs) RACE on it -- serialise them.
#include "../../include/reid_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace reid {
namespace {

constexpr int NA_MAX_ROWS = 4096;
// select_scatter_workspace_bytes(4096, 0): D* and counters 8 n | group thresholds 4 * chunk_pitch(n) | records 8 n * 256
constexpr size_t NA_SCRATCH_BYTES = (size_t)8 * NA_MAX_ROWS + (size_t)4 * 8 * (NA_MAX_ROWS / 256) +
                                    (size_t)8 * NA_MAX_ROWS * 256;

static __device__ __align__(256) unsigned char g_scatter_scratch[NA_SCRATCH_BYTES];

constexpr int NA_MAX_DEVICES = 64;
void* g_scratch_ptr[NA_MAX_DEVICES];          // device address of the array on each device (0: not fetched yet)

// device address of the scratch array on the CURRENT device; fetched once per device
int scratch_address(void** out) {
  int dev = -1;
  REID_CUDA(cudaGetDevice(&dev));
  REID_REQUIRE(dev >= 0 && dev < NA_MAX_DEVICES, "no_alloc: device ordinal %d outside [0, %d)", dev, NA_MAX_DEVICES);
  void* p = __atomic_load_n(&g_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    REID_CUDA(cudaGetSymbolAddress(&p, g_scatter_scratch));
    REID_REQUIRE(p != nullptr && ((uintptr_t)p & 255) == 0, "no_alloc: scratch symbol resolved to %p", p);
    __atomic_store_n(&g_scratch_ptr[dev], p, __ATOMIC_RELEASE);   // same value whichever thread stores it
  }
  *out = p;
  return REID_OK;
}

}  // namespace
}  // namespace reid

using namespace reid;

extern "C" {

int reid_topk_static_max_rows(void) { return NA_MAX_ROWS; }

int reid_topk_rows_scatter_static(const float* D, int n, int64_t ld, const int32_t* rowmax_key, const float* chunk_min,
                                  int K, int32_t* nbr, float* nbr_val, void* stream) {
  int dev = -1;
  REID_CUDA(cudaGetDevice(&dev));
  if (int rc = reid_device_check(dev)) return rc;
  REID_REQUIRE(D != nullptr && chunk_min != nullptr && rowmax_key != nullptr && nbr != nullptr,
               "topk_rows_scatter_static: NULL argument");
  REID_REQUIRE(n >= 0 && ld >= n, "topk_rows_scatter_static: bad shape n=%d ld=%lld", n, (long long)ld);
  REID_REQUIRE(n <= NA_MAX_ROWS,
               "topk_rows_scatter_static: n=%d exceeds the %d rows the built-in scratch holds; use reid_topk_rows_scatter "
               "with a caller-provided workspace", n, NA_MAX_ROWS);
  const size_t need = select_scatter_workspace_bytes(n, 0);
  REID_REQUIRE(need <= NA_SCRATCH_BYTES, "topk_rows_scatter_static: %zu bytes of scratch needed, %zu built in", need,
               NA_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = scratch_address(&scratch)) return rc;
  return launch_select_topk_scatter(D, n, (long long)ld, rowmax_key, chunk_min, K, nbr, nbr_val, 0, scratch,
                                    NA_SCRATCH_BYTES, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
