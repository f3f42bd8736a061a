// C-ABI of libumrs over the sm_100a kernels (include/umrs.h).  Handles own their
// device memory; there is no CPU fallback anywhere in this file: without a
// usable device every compute entry point returns an error status.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

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
Tuning g_tuning;
// Bulk-copy staging when the "tma_stage" hook is on auto: off -- it lost the same-box A/B on every
// plan (profiles/r2_tma_ab.txt), and the shipped build does not even contain it (-DUMRS_TMA_STAGE).
constexpr int kTmaDefault = 0;

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
struct ScanPlan { MapSource src; int32_t cap; int32_t max_rows; int smem_optin; int32_t kernel; int32_t multipass; int32_t tma_stage; };
int capi_scan_plan(const umrs_map* m, float range_max, int32_t n_obj, int32_t max_rays,
                   int32_t n_obj_smem, int32_t n_ent, ScanPlan* out);
void capi_map_segments(const umrs_map* m, const float4** segs_map_order, int32_t* n);
void capi_map_retain(umrs_map* m, int delta);
int capi_read_flags(uint32_t* d_flags);
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
  // uniform grid (K2): segments re-ordered by the cell of their midpoint
  float4* d_sorted = nullptr;
  int32_t* d_orig = nullptr;
  int32_t* d_cell_start = nullptr;
  int32_t gx = 0, gy = 0, n_binned = 0;
  float gx0 = 0.f, gy0 = 0.f, cell = 1.f, inv_cell = 1.f, hmax_x = 0.f, hmax_y = 0.f;
  // scratch for the host-buffer scan entry points
  float4* d_fans = nullptr; size_t fans_cap = 0;
  float4* d_obj = nullptr;  size_t obj_cap = 0;
  float* d_ranges = nullptr; size_t ranges_cap = 0;
  int32_t* d_idx = nullptr;  size_t idx_cap = 0;
  // work accounting of the scan entry points (umrs_map_enable_counters)
  unsigned long long* d_counters = nullptr;
  bool counters_on = false;
  uint32_t* d_flags = nullptr;     // sticky device-side conditions (kFlagRowsTruncated)
  std::atomic<int> n_worlds{0};    // worlds created from this map (they use its device buffers)
  // pinned staging for small host-buffer scans (a pageable cudaMemcpyAsync stages and
  // synchronises inside the driver; a single LaserScan is 4 KB)
  float* h_stage = nullptr;        // [kStageFloats] outputs, then kStageIn floats of inputs
  // scratch of the opt-in reference render (allocated on first use)
  float4* d_render_lines = nullptr;
  float4* d_render_scene = nullptr;
  float* d_render_c = nullptr;
  int32_t render_cap = 0;          // fans the scratch holds
};

// Host-side, one-time (like VectorMap::Load/Cleanup): bin every segment by the
// grid cell of its midpoint; segments longer than two cells go to an
// always-visited tail.  Cell size: ~8 segments per cell on average.
static cudaError_t build_grid(umrs_map* m) {
  const int32_t n = m->n;
  const float* s = m->host_segs.data();
  float x0 = 0.f, y0 = 0.f, x1 = 1.f, y1 = 1.f;
  for (int32_t i = 0; i < n; ++i) {
    const float* q = s + 4 * i;
    const float lx = std::min(q[0], q[2]), hx = std::max(q[0], q[2]);
    const float ly = std::min(q[1], q[3]), hy = std::max(q[1], q[3]);
    if (i == 0) { x0 = lx; x1 = hx; y0 = ly; y1 = hy; }
    x0 = std::min(x0, lx); x1 = std::max(x1, hx);
    y0 = std::min(y0, ly); y1 = std::max(y1, hy);
  }
  const float w = std::max(x1 - x0, 1e-3f), h = std::max(y1 - y0, 1e-3f);
  float cell = sqrtf(8.0f * w * h / (float)std::max(n, 1));
  cell = std::max(cell, std::max(w, h) / 2048.0f);
  cell = std::max(cell, 1e-3f);
  x0 -= 1e-3f * cell + 1e-5f * fabsf(x0);
  y0 -= 1e-3f * cell + 1e-5f * fabsf(y0);
  const float inv_cell = 1.0f / cell;
  const int gx = std::max(1, (int)((x1 - x0) * inv_cell) + 1);
  const int gy = std::max(1, (int)((y1 - y0) * inv_cell) + 1);
  std::vector<int32_t> cell_of(n), count((size_t)gx * gy + 1, 0);
  float hmx = 0.f, hmy = 0.f;
  int32_t n_over = 0;
  for (int32_t i = 0; i < n; ++i) {
    const float* q = s + 4 * i;
    const float mx = 0.5f * (q[0] + q[2]), my = 0.5f * (q[1] + q[3]);
    const float hx = std::max(fabsf(q[0] - mx), fabsf(q[2] - mx));
    const float hy = std::max(fabsf(q[1] - my), fabsf(q[3] - my));
    if (!(hx <= cell && hy <= cell)) {  // also catches non-finite coordinates
      cell_of[i] = -1;
      ++n_over;
      continue;
    }
    hmx = std::max(hmx, hx);
    hmy = std::max(hmy, hy);
    int cx = (int)((mx - x0) * inv_cell), cy = (int)((my - y0) * inv_cell);
    cx = std::min(std::max(cx, 0), gx - 1);
    cy = std::min(std::max(cy, 0), gy - 1);
    cell_of[i] = cy * gx + cx;
    ++count[(size_t)cell_of[i] + 1];
  }
  for (size_t c = 0; c < (size_t)gx * gy; ++c) count[c + 1] += count[c];
  const int32_t n_binned = n - n_over;
  std::vector<int32_t> cursor(count.begin(), count.end() - 1), orig(std::max(n, 1));
  std::vector<float> sorted((size_t)std::max(n, 1) * 4);
  int32_t tail = n_binned;
  for (int32_t i = 0; i < n; ++i) {  // index order inside a cell
    const int32_t k = cell_of[i] >= 0 ? cursor[cell_of[i]]++ : tail++;
    orig[k] = i;
    memcpy(&sorted[(size_t)k * 4], s + 4 * i, sizeof(float) * 4);
  }
  m->gx = gx; m->gy = gy; m->n_binned = n_binned;
  m->gx0 = x0; m->gy0 = y0; m->cell = cell; m->inv_cell = inv_cell;
  m->hmax_x = hmx * 1.0001f + 1e-6f; m->hmax_y = hmy * 1.0001f + 1e-6f;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&m->d_sorted), sizeof(float4) * (size_t)std::max(n, 1));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&m->d_orig), sizeof(int32_t) * (size_t)std::max(n, 1));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&m->d_cell_start), sizeof(int32_t) * count.size());
  if (e == cudaSuccess && n > 0) e = cudaMemcpy(m->d_sorted, sorted.data(), sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && n > 0) e = cudaMemcpy(m->d_orig, orig.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(m->d_cell_start, count.data(), sizeof(int32_t) * count.size(), cudaMemcpyHostToDevice);
  return e;
}

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
  if (e == cudaSuccess) e = build_grid(m);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&m->d_flags), sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(m->d_flags, 0, sizeof(uint32_t));
  if (e != cudaSuccess) {
    int rc = cuda_fail(e, "umrs_map_create");
    umrs_map_destroy(m);
    return rc;
  }
  *out = m;
  return UMRS_OK;
}

namespace umrs {
// The cleaned map lines on the device in MAP ORDER (GetSceneLines order; the reference render).
void capi_map_segments(const umrs_map* m, const float4** segs_map_order, int32_t* n) {
  *segs_map_order = m->d_segs;
  *n = m->n;
}

void capi_map_retain(umrs_map* m, int delta) { m->n_worlds.fetch_add(delta); }

// Picks the kernel variant and the shared-memory plan for a launch against map m.
//   K1 (kernel 0): every fan streams the whole map;  K2 (kernel 1): grid rows.
// auto: K2 from 4096 map segments up.  cap = candidate slots per pass.

int capi_scan_plan(const umrs_map* m, float range_max, int32_t n_obj, int32_t max_rays,
                   int32_t n_obj_smem, int32_t n_ent, ScanPlan* out) {
  ScanPlan pl;
  memset(&pl, 0, sizeof(pl));
  const bool grid = m->force_kernel == 1 || (m->force_kernel < 0 && m->n >= (g_tuning.grid_from > 0 ? g_tuning.grid_from : 4096));
  pl.kernel = grid ? 1 : 0;
  pl.smem_optin = m->smem_optin;
  MapSource& s = pl.src;
  s.n = m->n;
  s.n_map_total = m->n;
  if (grid) {
    s.segs = m->d_sorted; s.orig = m->d_orig; s.n_binned = m->n_binned;
    s.cell_start = m->d_cell_start; s.gx = m->gx; s.gy = m->gy;
    s.x0 = m->gx0; s.y0 = m->gy0; s.cell = m->cell; s.inv_cell = m->inv_cell;
    s.hmax_x = m->hmax_x; s.hmax_y = m->hmax_y;
    // rows a fan's query box [ly +- (range_max + margin)] can overlap: the kernel's own margin
    // (grid_margin, umrs_internal.h) at the largest |ly| for which the box still meets the grid
    const float y_far = std::max(fabsf(m->gy0), fabsf(m->gy0 + (float)m->gy * m->cell)) + 2.0f * fabsf(range_max);
    const double reach = 2.0 * ((double)range_max + (double)grid_margin(m->hmax_y, m->cell, y_far, range_max));
    double rows = floor(reach * (double)m->inv_cell * 1.00001) + 3.0;
    if (!(rows < (double)m->gy + 1.0)) rows = (double)m->gy + 1.0;  // also NaN / inf range_max
    pl.max_rows = (int32_t)rows + 3;  // + long-segment tail + object row + spare
  } else {
    s.segs = m->d_segs; s.orig = nullptr; s.n_binned = m->n;
    s.cell_start = nullptr; s.gx = 0; s.gy = 0;
    pl.max_rows = 3;
  }
  // bucket entries are 16 * slot in 16 bits -> at most 4064 slots per pass
  const int cap_max = std::max(32, std::min(g_tuning.cap > 0 ? g_tuning.cap : 1536, 4064));
  int64_t need = (int64_t)m->n + n_obj;
  int32_t cap = (int32_t)std::min<int64_t>(std::max<int64_t>(need, 32), cap_max);
  cap = (cap + 31) & ~31;
  // the whole source set of any fan fits one pass <=> no per-ray quotient array
  int mp = need > cap ? 1 : 0;
  // shrink the per-pass slots until the CTA fits the opt-in shared memory
  while (cap > 32 && scan_smem_bytes(cap, pl.max_rows, max_rays, n_obj_smem, n_ent, mp, 0) > (size_t)m->smem_optin) {
    cap = ((cap / 2) + 31) & ~31;
    mp = 1;
  }
  pl.multipass = mp;
  if (scan_smem_bytes(cap, pl.max_rows, max_rays, n_obj_smem, n_ent, mp, 0) > (size_t)m->smem_optin) {
    capi_set_error("fan does not fit shared memory (too many rays / object lines per world)");
    return UMRS_ERR_UNSUPPORTED;
  }
  // Bulk-copy (TMA) staging of the source rows (umrs_kernels.cu: stage_*).  K1: the whole map, one copy,
  // only where every fan is single-pass; K2: one copy per grid row.  The staging area re-uses arrays that
  // are dead during the gather, so it costs at most a little padding.  Default: see kTmaDefault.
  const int want = g_tuning.tma_stage >= 0 ? g_tuning.tma_stage : kTmaDefault;
  int tma = 0;
  if (want && tma_stage_built() && m->n > 0) tma = grid ? 2 : (mp == 0 ? 1 : 0);
  if (tma && scan_smem_bytes(cap, pl.max_rows, max_rays, n_obj_smem, n_ent, mp, tma) > (size_t)m->smem_optin) tma = 0;
  pl.tma_stage = tma;
  pl.cap = cap;
  *out = pl;
  return UMRS_OK;
}
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
  if (!CleanupSegments(raw.data(), (int)(raw.size() / 4), &clean)) return UMRS_ERR_UNSUPPORTED;
  return map_from_clean(clean.data(), (int32_t)(clean.size() / 4), out);
}

int umrs_map_destroy(umrs_map* m) {
  if (m == NULL) return UMRS_OK;
  if (m->n_worlds.load() > 0) {
    g_last_cuda_error = "umrs_map_destroy: worlds created from this map still exist";
    return UMRS_ERR_INVALID;
  }
  if (m->stream) cudaStreamSynchronize(m->stream);
  if (m->d_flags) cudaFree(m->d_flags);
  if (m->d_segs) cudaFree(m->d_segs);
  if (m->d_sorted) cudaFree(m->d_sorted);
  if (m->d_orig) cudaFree(m->d_orig);
  if (m->d_cell_start) cudaFree(m->d_cell_start);
  if (m->d_fans) cudaFree(m->d_fans);
  if (m->d_obj) cudaFree(m->d_obj);
  if (m->d_ranges) cudaFree(m->d_ranges);
  if (m->d_idx) cudaFree(m->d_idx);
  if (m->d_counters) cudaFree(m->d_counters);
  if (m->h_stage) cudaFreeHost(m->h_stage);
  if (m->d_render_lines) cudaFree(m->d_render_lines);
  if (m->d_render_scene) cudaFree(m->d_render_scene);
  if (m->d_render_c) cudaFree(m->d_render_c);
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
static constexpr size_t kStageFloats = 64 * 1024;   // outputs staged through pinned memory up to 256 KB
static constexpr size_t kStageIn = 16 * 1024;       // inputs (fans + object lines) up to 64 KB

int umrs_map_enable_counters(umrs_map* m, int32_t on) {
  if (m == NULL) return UMRS_ERR_INVALID;
  if (on && m->d_counters == nullptr) {
    UMRS_CUDA(cudaMalloc(&m->d_counters, sizeof(unsigned long long) * kNumCounters));
    UMRS_CUDA(cudaMemset(m->d_counters, 0, sizeof(unsigned long long) * kNumCounters));
  }
  m->counters_on = on != 0;
  return UMRS_OK;
}

int umrs_map_read_counters(umrs_map* m, uint64_t out8[8]) {
  if (m == NULL || out8 == NULL) return UMRS_ERR_INVALID;
  memset(out8, 0, sizeof(uint64_t) * 8);
  if (m->d_counters == nullptr) return UMRS_OK;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(out8, m->d_counters, sizeof(uint64_t) * 8, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

static int read_flags(uint32_t* d_flags) {
  if (d_flags == nullptr) return UMRS_OK;
  uint32_t f = 0;
  UMRS_CUDA(cudaMemcpy(&f, d_flags, sizeof(f), cudaMemcpyDeviceToHost));
  if (f == 0) return UMRS_OK;
  UMRS_CUDA(cudaMemset(d_flags, 0, sizeof(f)));
  if (f & kFlagRowsTruncated) {
    g_last_cuda_error = "a fan touched more grid rows than the plan sized shared memory for (far rows not scanned)";
    return UMRS_ERR_UNSUPPORTED;
  }
  return UMRS_OK;
}

int umrs_map_check(umrs_map* m) {
  if (m == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());
  return read_flags(m->d_flags);
}

int umrs_tuning_set(const char* key, int32_t value) {
  if (key == NULL) return UMRS_ERR_INVALID;
  const std::string k(key);
  if (k == "cta_shape") { if (value < -1 || value > 6) return UMRS_ERR_INVALID; g_tuning.cta_shape = value; }
  else if (k == "cap") { if (value != 0 && (value < 32 || value > 4064)) return UMRS_ERR_INVALID; g_tuning.cap = value; }
  else if (k == "grid_from") { if (value < 0) return UMRS_ERR_INVALID; g_tuning.grid_from = value; }
  else if (k == "step_slices") { if (value < 0 || value > 15) return UMRS_ERR_INVALID; g_tuning.step_slices = value; }
  else if (k == "split_step") { if (value < -1 || value > 1) return UMRS_ERR_INVALID; g_tuning.split_step = value; }
  else if (k == "tma_stage") { if (value < -1 || value > 1) return UMRS_ERR_INVALID; g_tuning.tma_stage = value; }
  else if (k == "render_cta_fans") { if (value < 0) return UMRS_ERR_INVALID; g_tuning.render_cta_fans = value; }
  else return UMRS_ERR_INVALID;
  return UMRS_OK;
}

int umrs_tuning_get(const char* key, int32_t* value) {
  if (key == NULL || value == NULL) return UMRS_ERR_INVALID;
  const std::string k(key);
  if (k == "cta_shape") *value = g_tuning.cta_shape;
  else if (k == "cap") *value = g_tuning.cap;
  else if (k == "grid_from") *value = g_tuning.grid_from;
  else if (k == "step_slices") *value = g_tuning.step_slices;
  else if (k == "split_step") *value = g_tuning.split_step;
  else if (k == "tma_stage") *value = g_tuning.tma_stage;
  else if (k == "render_cta_fans") *value = g_tuning.render_cta_fans;
  else return UMRS_ERR_INVALID;
  return UMRS_OK;
}

int umrs_scan_batch_device(umrs_map* m, const float* obj_segs_dev, int32_t n_obj,
                           const float* fans_dev, int32_t n_fans, float range_max,
                           int32_t num_rays, float* ranges_dev, int32_t* hit_idx_dev,
                           void* stream) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || ranges_dev == NULL ||
      (n_fans > 0 && fans_dev == NULL) || (n_obj > 0 && obj_segs_dev == NULL))
    return UMRS_ERR_INVALID;
  if (num_rays > kMaxRays) {
    g_last_cuda_error = "num_rays above the per-fan limit";
    return UMRS_ERR_UNSUPPORTED;
  }
  ScanPlan pl;
  const int st = capi_scan_plan(m, range_max, n_obj, num_rays, 0, 0, &pl);
  if (st != UMRS_OK) return st;
  ScanBatchParams p;
  memset(&p, 0, sizeof(p));
  p.c.src = pl.src;
  p.c.cap = pl.cap;
  p.c.max_rows = pl.max_rows;
  p.c.max_rays = num_rays;
  p.c.multipass = pl.multipass;
  p.c.tma_stage = pl.tma_stage;
  p.c.ranges = ranges_dev;
  p.c.hit_idx = hit_idx_dev;
  p.c.counters = nullptr;
  p.c.flags = m->d_flags;
  if (m->counters_on && m->d_counters != nullptr) {
    p.c.counters = m->d_counters;
    UMRS_CUDA(cudaMemsetAsync(m->d_counters, 0, sizeof(unsigned long long) * kNumCounters,
                              stream ? (cudaStream_t)stream : m->stream));
  }
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
  // Small calls (one LaserScan is 4 KB) go through pinned staging: every copy is then truly
  // asynchronous and the call has ONE synchronisation point.
  const size_t n_in = 4 * ((size_t)n_fans + (size_t)n_obj);
  const size_t n_stage_out = n_out * (hit_idx_out ? 2 : 1);
  const bool staged = n_in <= kStageIn && n_stage_out <= kStageFloats;
  if (staged && m->h_stage == nullptr) {
    if (cudaHostAlloc(reinterpret_cast<void**>(&m->h_stage), sizeof(float) * (kStageFloats + kStageIn),
                      cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      m->h_stage = nullptr;
    }
  }
  const bool use_stage = staged && m->h_stage != nullptr;
  const float* src_fans = fans;
  const float* src_obj = obj_segs_xyxy;
  if (use_stage) {
    float* in = m->h_stage + kStageFloats;
    memcpy(in, fans, sizeof(float) * 4 * (size_t)n_fans);
    if (n_obj > 0) memcpy(in + 4 * (size_t)n_fans, obj_segs_xyxy, sizeof(float) * 4 * (size_t)n_obj);
    src_fans = in;
    src_obj = in + 4 * (size_t)n_fans;
  }
  UMRS_CUDA(cudaMemcpyAsync(m->d_fans, src_fans, sizeof(float4) * (size_t)n_fans,
                            cudaMemcpyHostToDevice, m->stream));
  if (n_obj > 0)
    UMRS_CUDA(cudaMemcpyAsync(m->d_obj, src_obj, sizeof(float4) * (size_t)n_obj,
                              cudaMemcpyHostToDevice, m->stream));
  st = umrs_scan_batch_device(m, reinterpret_cast<const float*>(m->d_obj), n_obj,
                              reinterpret_cast<const float*>(m->d_fans), n_fans, range_max,
                              num_rays, m->d_ranges, hit_idx_out ? m->d_idx : nullptr,
                              m->stream);
  if (st != UMRS_OK) return st;
  float* dst_r = use_stage ? m->h_stage : ranges_out;
  int32_t* dst_i = use_stage ? reinterpret_cast<int32_t*>(m->h_stage + n_out) : hit_idx_out;
  UMRS_CUDA(cudaMemcpyAsync(dst_r, m->d_ranges, sizeof(float) * n_out,
                            cudaMemcpyDeviceToHost, m->stream));
  if (hit_idx_out)
    UMRS_CUDA(cudaMemcpyAsync(dst_i, m->d_idx, sizeof(int32_t) * n_out,
                              cudaMemcpyDeviceToHost, m->stream));
  UMRS_CUDA(cudaStreamSynchronize(m->stream));
  if (use_stage) {
    memcpy(ranges_out, dst_r, sizeof(float) * n_out);
    if (hit_idx_out) memcpy(hit_idx_out, dst_i, sizeof(int32_t) * n_out);
  }
  // only the grid walk can raise a device-side condition (row table overflow)
  if (m->force_kernel == 1 || (m->force_kernel < 0 && m->n >= (g_tuning.grid_from > 0 ? g_tuning.grid_from : 4096)))
    return read_flags(m->d_flags);
  return UMRS_OK;
}

// ---- opt-in reference render (tier A) ------------------------------------------------
int umrs_scan_batch_reference_render_device(umrs_map* m, const float* obj_segs_dev, int32_t n_obj,
                                            const float* fans_dev, int32_t n_fans, float range_max,
                                            int32_t num_rays, float* ranges_dev, void* stream) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || ranges_dev == NULL ||
      (n_fans > 0 && fans_dev == NULL) || (n_obj > 0 && obj_segs_dev == NULL))
    return UMRS_ERR_INVALID;
  if (n_fans == 0) return UMRS_OK;
  // scratch for one launch: 104 KB per fan, sized to what this map has been asked for so far
  const int32_t want = n_fans < kRenderFansPerLaunch ? n_fans : kRenderFansPerLaunch;
  if (m->render_cap < want) {
    cudaStream_t s0 = stream ? (cudaStream_t)stream : m->stream;
    UMRS_CUDA(cudaStreamSynchronize(s0));   // an earlier launch may still use the old scratch
    if (m->d_render_lines) cudaFree(m->d_render_lines);
    if (m->d_render_scene) cudaFree(m->d_render_scene);
    if (m->d_render_c) cudaFree(m->d_render_c);
    m->d_render_lines = m->d_render_scene = nullptr;
    m->d_render_c = nullptr;
    m->render_cap = 0;
    const int32_t cap = want < 1024 ? 1024 : want;
    const size_t n = (size_t)cap * 2000;
    float4 *a = nullptr, *b = nullptr;
    float* c = nullptr;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&a), sizeof(float4) * n);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&b), sizeof(float4) * n);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&c), sizeof(float) * ((size_t)cap * kRenderScratchCWords));
    if (e != cudaSuccess) {
      if (a) cudaFree(a);
      if (b) cudaFree(b);
      return cuda_fail(e, "reference render scratch");
    }
    m->d_render_lines = a; m->d_render_scene = b; m->d_render_c = c;
    m->render_cap = cap;
  }
  RenderScanParams p;
  memset(&p, 0, sizeof(p));
  p.segs = m->d_segs;
  p.n_map = m->n;
  p.obj_segs = reinterpret_cast<const float4*>(obj_segs_dev);
  p.n_obj = n_obj;
  p.fans = reinterpret_cast<const float4*>(fans_dev);
  p.n_fans = n_fans;
  p.range_max = range_max;
  p.num_rays = num_rays;
  p.ranges = ranges_dev;
  p.scratch_lines = m->d_render_lines;
  p.scratch_scene = m->d_render_scene;
  p.scratch_c = m->d_render_c;
  cudaStream_t st = stream ? (cudaStream_t)stream : m->stream;
  for (int b = 0; b < n_fans; b += m->render_cap)   // stream order: one scratch, reused
    UMRS_CUDA(launch_reference_render(p, b, b + m->render_cap < n_fans ? b + m->render_cap : n_fans, st));
  return UMRS_OK;
}

int umrs_scan_batch_reference_render(umrs_map* m, const float* obj_segs_xyxy, int32_t n_obj,
                                     const float* fans, int32_t n_fans, float range_max,
                                     int32_t num_rays, float* ranges_out) {
  if (m == NULL || n_fans < 0 || num_rays <= 0 || n_obj < 0 || ranges_out == NULL ||
      (n_fans > 0 && fans == NULL) || (n_obj > 0 && obj_segs_xyxy == NULL))
    return UMRS_ERR_INVALID;
  if (n_fans == 0) return UMRS_OK;
  const size_t n_out = (size_t)n_fans * (size_t)num_rays;
  int st;
  if ((st = grow(&m->d_fans, &m->fans_cap, (size_t)n_fans)) != UMRS_OK) return st;
  if ((st = grow(&m->d_obj, &m->obj_cap, (size_t)(n_obj > 0 ? n_obj : 1))) != UMRS_OK) return st;
  if ((st = grow(&m->d_ranges, &m->ranges_cap, n_out)) != UMRS_OK) return st;
  UMRS_CUDA(cudaMemcpyAsync(m->d_fans, fans, sizeof(float4) * (size_t)n_fans,
                            cudaMemcpyHostToDevice, m->stream));
  if (n_obj > 0)
    UMRS_CUDA(cudaMemcpyAsync(m->d_obj, obj_segs_xyxy, sizeof(float4) * (size_t)n_obj,
                              cudaMemcpyHostToDevice, m->stream));
  st = umrs_scan_batch_reference_render_device(m, reinterpret_cast<const float*>(m->d_obj), n_obj,
                                               reinterpret_cast<const float*>(m->d_fans), n_fans,
                                               range_max, num_rays, m->d_ranges, m->stream);
  if (st != UMRS_OK) return st;
  UMRS_CUDA(cudaMemcpyAsync(ranges_out, m->d_ranges, sizeof(float) * n_out,
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

namespace umrs {
int capi_read_flags(uint32_t* d_flags) { return read_flags(d_flags); }
}  // namespace umrs
