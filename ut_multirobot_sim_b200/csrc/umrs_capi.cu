// C-ABI of libumrs over the sm_100a kernels (include/umrs.h).  Handles own their
// device memory; there is no CPU fallback anywhere in this file: without a
// usable device every compute entry point returns an error status.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <new>
#include <string>
#include <vector>

#include "../../include/umrs.h"
#include "umrs_host.h"
#include "umrs_internal.h"

namespace umrs {

static thread_local std::string g_last_cuda_error;
static std::atomic<uint64_t> g_launches{0};

void note_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

static int cuda_fail(cudaError_t e, const char* what) {
  g_last_cuda_error = std::string(what) + ": " + cudaGetErrorString(e);
  cudaGetLastError();  // clear the sticky flag where possible
  return e == cudaErrorMemoryAllocation ? UMRS_ERR_NOMEM : UMRS_ERR_CUDA;
}

#define UMRS_CUDA(call)                                     \
  do {                                                      \
    cudaError_t e__ = (call);                               \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call);   \
  } while (0)

// The kernels are built for sm_100a only.
static int ensure_device(int* device_out, int* smem_optin_out) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    g_last_cuda_error = e != cudaSuccess ? cudaGetErrorString(e) : "no CUDA device";
    cudaGetLastError();
    return UMRS_ERR_NO_DEVICE;
  }
  int dev = 0;
  UMRS_CUDA(cudaGetDevice(&dev));
  int major = 0, smem = 0;
  UMRS_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  UMRS_CUDA(cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if (major != 10) {
    g_last_cuda_error = "device is not sm_100 (kernels are sm_100a only)";
    return UMRS_ERR_NO_DEVICE;
  }
  if (device_out) *device_out = dev;
  if (smem_optin_out) *smem_optin_out = smem;
  return UMRS_OK;
}

template <typename T>
static int grow(T** ptr, size_t* cap, size_t need) {
  if (need <= *cap) return UMRS_OK;
  if (*ptr) cudaFree(*ptr);
  *ptr = nullptr;
  *cap = 0;
  UMRS_CUDA(cudaMalloc(reinterpret_cast<void**>(ptr), need * sizeof(T)));
  *cap = need;
  return UMRS_OK;
}

}  // namespace umrs

using namespace umrs;
struct umrs_map;
namespace umrs {
// shared with umrs_world.cu
int capi_cuda_fail(cudaError_t e, const char* what) { return cuda_fail(e, what); }
void capi_set_error(const char* msg) { g_last_cuda_error = msg; }
struct MapView { const float4* d_segs; int32_t n; int smem_optin; };
MapView capi_map_view(const umrs_map* m);
}  // namespace umrs

// ================================== map ==========================================
struct umrs_map {
  int device = 0;
  int smem_optin = 0;
  std::vector<float> host_segs;  // cleaned, xyxy
  int32_t n = 0;
  float4* d_segs = nullptr;
  cudaStream_t stream = nullptr;
  int32_t force_kernel = -1;
  // scratch for the host-buffer scan entry points
  float4* d_fans = nullptr; size_t fans_cap = 0;
  float4* d_obj = nullptr;  size_t obj_cap = 0;
  float* d_ranges = nullptr; size_t ranges_cap = 0;
  int32_t* d_idx = nullptr;  size_t idx_cap = 0;
};

static int map_from_clean(const float* segs, int32_t n, umrs_map** out) {
  int dev = 0, smem = 0;
  int st = ensure_device(&dev, &smem);
  if (st != UMRS_OK) return st;
  umrs_map* m = new (std::nothrow) umrs_map();
  if (!m) return UMRS_ERR_NOMEM;
  m->device = dev;
  m->smem_optin = smem;
  m->n = n;
  m->host_segs.assign(segs, segs + (size_t)n * 4);
  cudaError_t e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&m->d_segs), sizeof(float4) * (size_t)(n > 0 ? n : 1));
  if (e == cudaSuccess && n > 0)
    e = cudaMemcpy(m->d_segs, segs, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    int rc = cuda_fail(e, "umrs_map_create");
    umrs_map_destroy(m);
    return rc;
  }
  *out = m;
  return UMRS_OK;
}

namespace umrs {
MapView capi_map_view(const umrs_map* m) { return MapView{m->d_segs, m->n, m->smem_optin}; }
}  // namespace umrs

extern "C" {

const char* umrs_last_cuda_error(void) { return g_last_cuda_error.c_str(); }
uint64_t umrs_kernel_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int umrs_set_device(int32_t device) {
  UMRS_CUDA(cudaSetDevice(device));
  return UMRS_OK;
}
int umrs_get_device(int32_t* device) {
  if (device == NULL) return UMRS_ERR_INVALID;
  int d = 0;
  UMRS_CUDA(cudaGetDevice(&d));
  *device = d;
  return UMRS_OK;
}
int umrs_device_count(int32_t* count) {
  if (count == NULL) return UMRS_ERR_INVALID;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  *count = n;
  return UMRS_OK;
}

int umrs_map_create(const float* segs_xyxy, int32_t n, umrs_map** out) {
  if (out == NULL || n < 0 || (n > 0 && segs_xyxy == NULL)) return UMRS_ERR_INVALID;
  *out = NULL;
  return map_from_clean(segs_xyxy, n, out);
}

int umrs_map_load_file(const char* path, umrs_map** out) {
  if (out == NULL || path == NULL) return UMRS_ERR_INVALID;
  *out = NULL;
  std::vector<float> raw, clean;
  if (!LoadMapFile(path, &raw)) return UMRS_ERR_IO;  // the reference exit(1)s here
  CleanupSegments(raw.data(), (int)(raw.size() / 4), &clean);
  return map_from_clean(clean.data(), (int32_t)(clean.size() / 4), out);
}

int umrs_map_destroy(umrs_map* m) {
  if (m == NULL) return UMRS_OK;
  if (m->d_segs) cudaFree(m->d_segs);
  if (m->d_fans) cudaFree(m->d_fans);
  if (m->d_obj) cudaFree(m->d_obj);
  if (m->d_ranges) cudaFree(m->d_ranges);
  if (m->d_idx) cudaFree(m->d_idx);
  if (m->stream) cudaStreamDestroy(m->stream);
  delete m;
  return UMRS_OK;
}

int umrs_map_num_segments(const umrs_map* m, int32_t* n) {
  if (m == NULL || n == NULL) return UMRS_ERR_INVALID;
  *n = m->n;
  return UMRS_OK;
}

int umrs_map_get_segments(const umrs_map* m, float* segs, int32_t cap) {
  if (m == NULL || (segs == NULL && m->n > 0) || cap < m->n) return UMRS_ERR_INVALID;
  if (m->n > 0) memcpy(segs, m->host_segs.data(), sizeof(float) * 4 * (size_t)m->n);
  return UMRS_OK;
}

int umrs_map_set_scan_kernel(umrs_map* m, int32_t force) {
  if (m == NULL || force < -1 || force > 1) return UMRS_ERR_INVALID;
  m->force_kernel = force;
  return UMRS_OK;
}

// ================================== scans =========================================
int umrs_scan_batch_device(umrs_map* m, const float* obj_segs_dev, int32_t n_obj,
                           const float* fans_dev, int32_t n_fans, float range_max,
                           int32_t num_rays, float* ranges_dev, int32_t* hit_idx_dev,
                           void* stream) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || ranges_dev == NULL ||
      (n_fans > 0 && fans_dev == NULL) || (n_obj > 0 && obj_segs_dev == NULL))
    return UMRS_ERR_INVALID;
  const size_t smem = k1_smem_bytes(m->n + n_obj, 0, 0);
  if (smem > (size_t)m->smem_optin || m->force_kernel == 1) {
    g_last_cuda_error = "candidate set does not fit shared memory and the grid kernel is not built yet";
    return UMRS_ERR_UNSUPPORTED;
  }
  ScanBatchParams p;
  memset(&p, 0, sizeof(p));
  p.c.map_segs = m->d_segs;
  p.c.n_map = m->n;
  p.c.ranges = ranges_dev;
  p.c.hit_idx = hit_idx_dev;
  p.c.counters = nullptr;
  p.obj_segs = reinterpret_cast<const float4*>(obj_segs_dev);
  p.n_obj = n_obj;
  p.fans = reinterpret_cast<const float4*>(fans_dev);
  p.n_fans = n_fans;
  p.range_max = range_max;
  p.num_rays = num_rays;
  UMRS_CUDA(launch_scan_batch(p, stream ? (cudaStream_t)stream : m->stream));
  return UMRS_OK;
}

int umrs_scan_batch(umrs_map* m, const float* obj_segs_xyxy, int32_t n_obj,
                    const float* fans, int32_t n_fans, float range_max, int32_t num_rays,
                    float* ranges_out, int32_t* hit_idx_out) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || ranges_out == NULL ||
      (n_fans > 0 && fans == NULL) || (n_obj > 0 && obj_segs_xyxy == NULL))
    return UMRS_ERR_INVALID;
  if (n_fans == 0) return UMRS_OK;
  const size_t n_out = (size_t)n_fans * (size_t)num_rays;
  int st;
  if ((st = grow(&m->d_fans, &m->fans_cap, (size_t)n_fans)) != UMRS_OK) return st;
  if ((st = grow(&m->d_obj, &m->obj_cap, (size_t)(n_obj > 0 ? n_obj : 1))) != UMRS_OK) return st;
  if ((st = grow(&m->d_ranges, &m->ranges_cap, n_out)) != UMRS_OK) return st;
  if (hit_idx_out && (st = grow(&m->d_idx, &m->idx_cap, n_out)) != UMRS_OK) return st;
  UMRS_CUDA(cudaMemcpyAsync(m->d_fans, fans, sizeof(float4) * (size_t)n_fans,
                            cudaMemcpyHostToDevice, m->stream));
  if (n_obj > 0)
    UMRS_CUDA(cudaMemcpyAsync(m->d_obj, obj_segs_xyxy, sizeof(float4) * (size_t)n_obj,
                              cudaMemcpyHostToDevice, m->stream));
  st = umrs_scan_batch_device(m, reinterpret_cast<const float*>(m->d_obj), n_obj,
                              reinterpret_cast<const float*>(m->d_fans), n_fans, range_max,
                              num_rays, m->d_ranges, hit_idx_out ? m->d_idx : nullptr,
                              m->stream);
  if (st != UMRS_OK) return st;
  UMRS_CUDA(cudaMemcpyAsync(ranges_out, m->d_ranges, sizeof(float) * n_out,
                            cudaMemcpyDeviceToHost, m->stream));
  if (hit_idx_out)
    UMRS_CUDA(cudaMemcpyAsync(hit_idx_out, m->d_idx, sizeof(int32_t) * n_out,
                              cudaMemcpyDeviceToHost, m->stream));
  UMRS_CUDA(cudaStreamSynchronize(m->stream));
  return UMRS_OK;
}

int umrs_predicted_scan(umrs_map* m, const float* obj_segs_xyxy, int32_t n_obj,
                        float loc_x, float loc_y, float range_min, float range_max,
                        float angle_min, float angle_max, int32_t num_rays,
                        float* ranges_out, int32_t* hit_idx_out) {
  (void)range_min;  // unused by the reference too (vector_map.cpp:421)
  const float fan[4] = {loc_x, loc_y, angle_min, angle_max};
  return umrs_scan_batch(m, obj_segs_xyxy, n_obj, fan, 1, range_max, num_rays, ranges_out,
                         hit_idx_out);
}

}  // extern "C"
