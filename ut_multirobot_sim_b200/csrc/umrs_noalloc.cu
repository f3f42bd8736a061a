// umrs_scan_batch_reference_render_noalloc: the opt-in reference render (umrs_render.cu) for
// callers that must not allocate -- code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The render needs 104,000 bytes of global scratch per fan (candidate lines, rendered scene,
// cast constants + angular intervals + candidate lists).  The ordinary entry point sizes that
// scratch with cudaMalloc on demand; this one keeps it in a file-scope device array for at most
// UMRS_NOALLOC_MAX_FANS fans (6.35 MiB, held by every GPU that loads the library until exit).
//
//   * nothing here calls cudaMalloc / cudaFree or an allocating library routine;
//   * the array is SHARED by every call on a device and keeps the previous call's contents.
//     Nothing has to start at zero: render_fan writes every scratch word it later reads (the
//     ordinary entry point re-uses its scratch between launches in the same way), so no clear
//     is issued.  Calls that overlap on the GPU (different streams, no ordering) RACE on it;
//   * in host code the array's name is a host-side stand-in, not a device address: the address
//     the kernels get comes from cudaGetSymbolAddress, asked once per device with that device
//     current (below).  Do that first call outside stream capture (n_fans = 0 is enough).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/umrs.h"
#include "umrs_internal.h"

using namespace umrs;
struct umrs_map;
namespace umrs {
// umrs_capi.cu
int capi_cuda_fail(cudaError_t e, const char* what);
void capi_set_error(const char* msg);
void capi_map_segments(const umrs_map* m, const float4** segs_map_order, int32_t* n);
}  // namespace umrs

namespace {

constexpr int kLinesPerFan = 2000;                                   // the reference's candidate cap (vector_map.cpp:221)
constexpr size_t kLinesBytes = sizeof(float4) * kLinesPerFan;        // 32,000
constexpr size_t kCBytes = sizeof(float) * kRenderScratchCWords;     // 40,000
constexpr size_t kOffScene = kLinesBytes * UMRS_NOALLOC_MAX_FANS;
constexpr size_t kOffC = 2 * kLinesBytes * UMRS_NOALLOC_MAX_FANS;
constexpr size_t kScratchBytes = kOffC + kCBytes * UMRS_NOALLOC_MAX_FANS;
static_assert(kOffScene % 256 == 0 && kOffC % 256 == 0, "sub-arrays keep the 256-byte alignment");

// [lines: MAX_FANS x 2000 float4 | scene: the same | c: MAX_FANS x kRenderScratchCWords float]
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
unsigned char* g_scratch_addr[kMaxDevices];   // device address of g_noalloc_scratch, per device

// Device address of the scratch array on the CURRENT device (asked once per device).
int scratch_on_current_device(int* dev_out, unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return capi_cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) {
    capi_set_error("umrs_scan_batch_reference_render_noalloc: device ordinal out of range");
    return UMRS_ERR_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return capi_cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (ptr == nullptr) {
      capi_set_error("umrs_scan_batch_reference_render_noalloc: no device address for the scratch array");
      return UMRS_ERR_CUDA;
    }
    g_scratch_addr[dev] = static_cast<unsigned char*>(ptr);
  }
  *dev_out = dev;
  *out = g_scratch_addr[dev];
  return UMRS_OK;
}

}  // namespace

extern "C" int umrs_scan_batch_reference_render_noalloc(umrs_map* m, const float* obj_segs_dev, int32_t n_obj,
                                                        const float* fans_dev, int32_t n_fans,
                                                        float range_max, int32_t num_rays,
                                                        float* ranges_dev, void* stream) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || (n_fans > 0 && ranges_dev == NULL) ||
      (n_fans > 0 && fans_dev == NULL) || (n_obj > 0 && obj_segs_dev == NULL)) {
    capi_set_error("umrs_scan_batch_reference_render_noalloc: bad argument");
    return UMRS_ERR_INVALID;
  }
  if (n_fans > UMRS_NOALLOC_MAX_FANS) {
    char msg[160];
    snprintf(msg, sizeof(msg),
             "umrs_scan_batch_reference_render_noalloc: n_fans %d exceeds UMRS_NOALLOC_MAX_FANS (%d); "
             "split the batch or use umrs_scan_batch_reference_render_device",
             (int)n_fans, (int)UMRS_NOALLOC_MAX_FANS);
    capi_set_error(msg);
    return UMRS_ERR_INVALID;
  }
  int dev = 0;
  unsigned char* scratch = nullptr;
  const int st = scratch_on_current_device(&dev, &scratch);
  if (st != UMRS_OK) return st;
  if (n_fans == 0) return UMRS_OK;   // set-up only

  RenderScanParams p;
  memset(&p, 0, sizeof(p));
  capi_map_segments(m, &p.segs, &p.n_map);
  // the map's buffers must live on the device whose scratch array the launch is given
  cudaPointerAttributes attr;
  cudaError_t e = cudaPointerGetAttributes(&attr, p.segs);
  if (e != cudaSuccess) return capi_cuda_fail(e, "cudaPointerGetAttributes(map)");
  if (attr.type != cudaMemoryTypeDevice || attr.device != dev) {
    capi_set_error("umrs_scan_batch_reference_render_noalloc: the map's device is not the current device");
    return UMRS_ERR_INVALID;
  }
  p.obj_segs = reinterpret_cast<const float4*>(obj_segs_dev);
  p.n_obj = n_obj;
  p.fans = reinterpret_cast<const float4*>(fans_dev);
  p.n_fans = n_fans;
  p.range_max = range_max;
  p.num_rays = num_rays;
  p.ranges = ranges_dev;
  p.scratch_lines = reinterpret_cast<float4*>(scratch);
  p.scratch_scene = reinterpret_cast<float4*>(scratch + kOffScene);
  p.scratch_c = reinterpret_cast<float*>(scratch + kOffC);
  e = launch_reference_render(p, 0, n_fans, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return capi_cuda_fail(e, "reference render launch (no-alloc)");
  return UMRS_OK;
}
