// C-ABI, batched worlds (include/umrs.h: umrs_world_*).  Replaces
// Simulator::update + publishLaser (simulator.cpp:474-509, 393-426) for W
// independent worlds in lockstep.  Device state is ping-ponged between two
// buffers so that the fan CTAs of one world can all read the old state.
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include "../../include/umrs.h"
#include "umrs_host.h"
#include "umrs_internal.h"

using namespace umrs;

// from umrs_capi.cu
struct umrs_map;
namespace umrs {
int capi_cuda_fail(cudaError_t e, const char* what);
void capi_set_error(const char* msg);
struct ScanPlan { MapSource src; int32_t cap; int32_t max_rows; int smem_optin; int32_t kernel; int32_t multipass; int32_t tma_stage; };
int capi_scan_plan(const umrs_map* m, float range_max, int32_t n_obj, int32_t max_rays,
                   int32_t n_obj_smem, int32_t n_ent, ScanPlan* out);
void capi_map_segments(const umrs_map* m, const float4** segs_map_order, int32_t* n);
void capi_map_retain(umrs_map* m, int delta);
int capi_read_flags(uint32_t* d_flags);
}  // namespace umrs

#define UMRS_CUDA(call)                                          \
  do {                                                           \
    cudaError_t e__ = (call);                                    \
    if (e__ != cudaSuccess) return capi_cuda_fail(e__, #call);   \
  } while (0)

constexpr int kMaxSlices = 16;

struct umrs_world {
  umrs_world_desc desc;
  umrs_map* map = nullptr;   // retained: the plan points into its device buffers
  std::vector<umrs_robot_params> robots;
  std::vector<umrs_human_params> humans;
  ScanPlan plan;  // kernel variant + shared-memory plan, fixed at creation
  const float4* map_segs_in_map_order = nullptr;   // for the reference render
  int32_t n_map = 0;
  int32_t max_rays = 0;
  int32_t W = 0, R = 0, H = 0, O = 0;
  int32_t L_obj = 0, rays_per_world = 0;
  std::vector<int32_t> ent_off, ray_off, num_rays;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;   // device->host copies of finished slices
  cudaEvent_t slice_done[kMaxSlices] = {};
  // pipelined host-buffer ticks (umrs_world_step_async): copy of slice s of the previous tick,
  // and whole-tick completion, both recorded on the copy stream
  cudaEvent_t tick_done[4] = {};
  cudaEvent_t kernels_done = nullptr;   // the kernel(s) of a pipelined tick
  // Pipelined ticks alternate between TWO device output buffers, so that tick k+1's kernel never
  // waits for tick k's device->host copy: one launch + one copy per tick (no slices, a handful of
  // API calls), the copy engine busy back to back.  buf_free[b]: the last copy out of buffer b.
  cudaEvent_t buf_free[2] = {};
  bool buf_busy[2] = {false, false};    // a copy out of buffer b may still be in flight
  uint64_t async_ticks = 0;
  bool async_pending = false;     // device->host copies of a pipelined tick may still be running
  int out_buf = 0;                // the buffer the last step wrote (umrs_world_device_buffers, stats)
  uint64_t step = 0;
  int cur = 0;  // index of the "in" state buffer
  // device
  umrs_robot_params* d_robots = nullptr;
  umrs_human_params* d_humans = nullptr;
  float4* d_tmpl = nullptr;
  int16_t* d_line_ent = nullptr;
  int32_t* d_ent_off = nullptr;
  int32_t* d_ray_off = nullptr;
  int32_t* d_num_rays = nullptr;
  RobotState* d_rs[2] = {nullptr, nullptr};
  HumanState* d_hs[2] = {nullptr, nullptr};
  float* d_static = nullptr;
  float* d_cmds = nullptr;
  uint8_t* d_mask = nullptr;
  float* d_ranges = nullptr;      // output buffer 0 (every synchronous / device-resident step)
  int32_t* d_idx = nullptr;
  float* d_ranges2 = nullptr;     // output buffer 1 (allocated by the first pipelined tick)
  int32_t* d_idx2 = nullptr;
  float4* d_fan_params = nullptr;
  float4* d_obj_lines = nullptr;
  float* d_pose = nullptr;
  unsigned long long* d_counters = nullptr;
  // scratch of the reference render (desc.reference_render): kRenderFansPerLaunch fans
  float4* d_render_lines = nullptr;
  float4* d_render_scene = nullptr;
  float* d_render_c = nullptr;
  int32_t render_cap = 0;   // fans the render scratch holds (one launch)
  bool counters_on = false;
  double* d_stats = nullptr;
  uint32_t* d_flags = nullptr;   // sticky device-side conditions (kFlagRowsTruncated)
};

template <typename T>
static cudaError_t dalloc(T** p, size_t n) {
  return cudaMalloc(reinterpret_cast<void**>(p), sizeof(T) * (n > 0 ? n : 1));
}

extern "C" {

int umrs_world_destroy(umrs_world* w) {
  if (w == NULL) return UMRS_OK;
  // the caller's (pinned) output buffers may still be copy targets
  if (w->stream) cudaStreamSynchronize(w->stream);
  if (w->copy_stream) cudaStreamSynchronize(w->copy_stream);
  if (w->map) capi_map_retain(w->map, -1);
  if (w->d_flags) cudaFree(w->d_flags);
  if (w->kernels_done) cudaEventDestroy(w->kernels_done);
  void* ptrs[] = {w->d_robots, w->d_humans, w->d_tmpl, w->d_line_ent, w->d_ent_off,
                  w->d_ray_off, w->d_num_rays, w->d_rs[0], w->d_rs[1], w->d_hs[0],
                  w->d_hs[1], w->d_static, w->d_cmds, w->d_mask, w->d_ranges, w->d_idx, w->d_ranges2, w->d_idx2,
                  w->d_fan_params, w->d_obj_lines, w->d_pose, w->d_counters, w->d_stats,
                  w->d_render_lines, w->d_render_scene, w->d_render_c};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (w->stream) cudaStreamDestroy(w->stream);
  if (w->copy_stream) cudaStreamDestroy(w->copy_stream);
  for (cudaEvent_t e : w->slice_done)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : w->buf_free)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : w->tick_done)
    if (e) cudaEventDestroy(e);
  delete w;
  return UMRS_OK;
}

int umrs_world_create(umrs_map* map, const umrs_world_desc* d, umrs_world** out) {
  if (out == NULL) return UMRS_ERR_INVALID;
  *out = NULL;
  if (map == NULL || d == NULL || d->num_worlds <= 0 || d->robots_per_world <= 0 ||
      d->humans_per_world < 0 || d->static_objects_per_world < 0 || d->robots == NULL ||
      (d->humans_per_world > 0 && d->humans == NULL) || !(d->dt > 0.f))
    return UMRS_ERR_INVALID;
  umrs_world* w = new (std::nothrow) umrs_world();
  if (!w) return UMRS_ERR_NOMEM;
  w->desc = *d;
  w->W = d->num_worlds; w->R = d->robots_per_world;
  w->H = d->humans_per_world; w->O = d->static_objects_per_world;
  w->robots.assign(d->robots, d->robots + w->R);
  if (w->H > 0) w->humans.assign(d->humans, d->humans + w->H);
  w->desc.robots = nullptr;
  w->desc.humans = nullptr;
  capi_map_segments(map, &w->map_segs_in_map_order, &w->n_map);
  w->map = map;
  capi_map_retain(map, +1);

  // ---- object-line roster: [statics][humans][robot bodies] ---------------------
  std::vector<float> tmpl;      // xyxy per world-line
  std::vector<int16_t> line_ent;
  w->ent_off.assign(1, 0);
  auto add_entity = [&](const float* t, int n_lines) {
    const int e = (int)w->ent_off.size() - 1;
    for (int i = 0; i < n_lines; ++i) {
      tmpl.insert(tmpl.end(), t + 4 * i, t + 4 * i + 4);
      line_ent.push_back((int16_t)e);
    }
    w->ent_off.push_back(w->ent_off.back() + n_lines);
  };
  float sq[16];
  SquareTemplate(sq);
  for (int o = 0; o < w->O; ++o) add_entity(sq, 4);
  for (int h = 0; h < w->H; ++h) {
    if (w->humans[h].num_segments < 0 || w->humans[h].num_segments > 4096) {
      umrs_world_destroy(w);
      return UMRS_ERR_INVALID;
    }
    std::vector<float> ht;
    HumanTemplate(w->humans[h], &ht);
    add_entity(ht.data(), (int)(ht.size() / 4));
  }
  for (int r = 0; r < w->R; ++r) {
    float rt[16];
    RobotTemplate(w->robots[r], rt);
    add_entity(rt, w->robots[r].footprint_enabled ? 4 : 0);
  }
  w->L_obj = w->ent_off.back();
  if (w->O + w->H + w->R > 32767) {
    umrs_world_destroy(w);
    return UMRS_ERR_UNSUPPORTED;
  }

  // ---- rays -----------------------------------------------------------------------
  w->ray_off.assign(1, 0);
  for (int r = 0; r < w->R; ++r) {
    const int n = umrs_laser_num_rays(&w->robots[r].laser);
    if (n <= 0 || !(w->robots[r].laser.angle_max > w->robots[r].laser.angle_min)) {
      umrs_world_destroy(w);
      return UMRS_ERR_INVALID;
    }
    w->num_rays.push_back(n);
    w->ray_off.push_back(w->ray_off.back() + n);
  }
  w->rays_per_world = w->ray_off.back();

  float range_max = 0.f;
  for (int r = 0; r < w->R; ++r) {
    w->max_rays = w->num_rays[r] > w->max_rays ? w->num_rays[r] : w->max_rays;
    range_max = w->robots[r].laser.range_max > range_max ? w->robots[r].laser.range_max : range_max;
  }
  if (w->max_rays > kMaxRays || w->L_obj > 16384) {
    capi_set_error("too many rays per fan or object lines per world");
    umrs_world_destroy(w);
    return UMRS_ERR_UNSUPPORTED;
  }
  {
    const int st = capi_scan_plan(map, range_max, w->L_obj, w->max_rays, w->L_obj,
                                  w->O + w->H + w->R, &w->plan);
    if (st != UMRS_OK) {
      umrs_world_destroy(w);
      return st;
    }
  }

  // ---- device allocations -----------------------------------------------------------
  const size_t WR = (size_t)w->W * w->R, WH = (size_t)w->W * w->H, WO = (size_t)w->W * w->O;
  const size_t n_rays = (size_t)w->W * w->rays_per_world;
  cudaError_t e = cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking);
  auto ok = [&]() { return e == cudaSuccess; };
  if (ok()) e = cudaStreamCreateWithFlags(&w->copy_stream, cudaStreamNonBlocking);
  for (int k = 0; k < kMaxSlices; ++k)
    if (ok()) e = cudaEventCreateWithFlags(&w->slice_done[k], cudaEventDisableTiming);
  for (int k = 0; k < 2; ++k)
    if (ok()) e = cudaEventCreateWithFlags(&w->buf_free[k], cudaEventDisableTiming);
  for (int k = 0; k < 4; ++k)
    if (ok()) e = cudaEventCreateWithFlags(&w->tick_done[k], cudaEventDisableTiming);
  if (ok()) e = dalloc(&w->d_robots, (size_t)w->R);
  if (ok()) e = dalloc(&w->d_humans, (size_t)w->H);
  if (ok()) e = dalloc(&w->d_tmpl, (size_t)w->L_obj);
  if (ok()) e = dalloc(&w->d_line_ent, (size_t)w->L_obj);
  if (ok()) e = dalloc(&w->d_ent_off, w->ent_off.size());
  if (ok()) e = dalloc(&w->d_ray_off, w->ray_off.size());
  if (ok()) e = dalloc(&w->d_num_rays, (size_t)w->R);
  for (int b = 0; b < 2; ++b) {
    if (ok()) e = dalloc(&w->d_rs[b], WR);
    if (ok()) e = dalloc(&w->d_hs[b], WH);
  }
  if (ok()) e = dalloc(&w->d_static, WO * 3);
  if (ok()) e = dalloc(&w->d_cmds, WR * 3);
  if (ok()) e = dalloc(&w->d_mask, WR);
  if (ok()) e = dalloc(&w->d_ranges, n_rays);
  if (ok() && d->write_hit_idx && !d->reference_render) e = dalloc(&w->d_idx, n_rays);
  w->render_cap = (int32_t)(WR < (size_t)kRenderFansPerLaunch ? WR : (size_t)kRenderFansPerLaunch);
  if (ok() && d->reference_render) e = dalloc(&w->d_render_lines, (size_t)w->render_cap * 2000);
  if (ok() && d->reference_render) e = dalloc(&w->d_render_scene, (size_t)w->render_cap * 2000);
  if (ok() && d->reference_render) e = dalloc(&w->d_render_c, (size_t)w->render_cap * kRenderScratchCWords);
  if (ok()) e = dalloc(&w->d_fan_params, WR);
  if (ok()) e = dalloc(&w->d_obj_lines, (size_t)w->W * w->L_obj);
  if (ok()) e = dalloc(&w->d_pose, WR * 3);
  if (ok()) e = dalloc(&w->d_counters, (size_t)kNumCounters);
  if (ok()) e = dalloc(&w->d_stats, (size_t)4);
  if (ok()) e = dalloc(&w->d_flags, (size_t)1);
  if (ok()) e = cudaMemset(w->d_flags, 0, sizeof(uint32_t));
  if (ok()) e = cudaEventCreateWithFlags(&w->kernels_done, cudaEventDisableTiming);
  if (ok()) e = cudaMemcpy(w->d_robots, w->robots.data(), sizeof(umrs_robot_params) * w->R, cudaMemcpyHostToDevice);
  if (ok() && w->H) e = cudaMemcpy(w->d_humans, w->humans.data(), sizeof(umrs_human_params) * w->H, cudaMemcpyHostToDevice);
  if (ok() && w->L_obj) e = cudaMemcpy(w->d_tmpl, tmpl.data(), sizeof(float4) * w->L_obj, cudaMemcpyHostToDevice);
  if (ok() && w->L_obj) e = cudaMemcpy(w->d_line_ent, line_ent.data(), sizeof(int16_t) * w->L_obj, cudaMemcpyHostToDevice);
  if (ok()) e = cudaMemcpy(w->d_ent_off, w->ent_off.data(), sizeof(int32_t) * w->ent_off.size(), cudaMemcpyHostToDevice);
  if (ok()) e = cudaMemcpy(w->d_ray_off, w->ray_off.data(), sizeof(int32_t) * w->ray_off.size(), cudaMemcpyHostToDevice);
  if (ok()) e = cudaMemcpy(w->d_num_rays, w->num_rays.data(), sizeof(int32_t) * w->R, cudaMemcpyHostToDevice);
  if (ok()) e = cudaMemset(w->d_static, 0, sizeof(float) * (WO * 3 > 0 ? WO * 3 : 1));
  if (ok()) e = cudaMemset(w->d_counters, 0, sizeof(unsigned long long) * kNumCounters);
  if (ok()) e = cudaMemset(w->d_ranges, 0, sizeof(float) * n_rays);
  if (!ok()) {
    const int rc = capi_cuda_fail(e, "umrs_world_create");
    umrs_world_destroy(w);
    return rc;
  }
  *out = w;
  // default state: robots at the origin, humans at the desc's start/goal
  std::vector<float> zero(WR * 3, 0.f);
  int st = umrs_world_set_robot_state(w, zero.data(), NULL);
  if (st == UMRS_OK && w->H > 0) st = umrs_world_set_humans(w, NULL, NULL);
  if (st != UMRS_OK) {
    umrs_world_destroy(w);
    *out = NULL;
  }
  return st;
}

int umrs_world_rays_per_world(const umrs_world* w, int32_t* n) {
  if (w == NULL || n == NULL) return UMRS_ERR_INVALID;
  *n = w->rays_per_world;
  return UMRS_OK;
}

int umrs_world_object_lines_per_world(const umrs_world* w, int32_t* n) {
  if (w == NULL || n == NULL) return UMRS_ERR_INVALID;
  *n = w->L_obj;
  return UMRS_OK;
}

int umrs_world_scan_header(const umrs_world* w, int32_t robot, umrs_laser_scan_header* h) {
  if (w == NULL || h == NULL || robot < 0 || robot >= w->R) return UMRS_ERR_INVALID;
  const umrs_laser_params& l = w->robots[robot].laser;
  // simulator.cpp:150-159
  h->angle_min = l.angle_min; h->angle_max = l.angle_max;
  h->angle_increment = l.angle_increment;
  h->time_increment = 0.0f; h->scan_time = 0.05f;
  h->range_min = l.range_min; h->range_max = l.range_max;
  h->num_rays = w->num_rays[robot];
  h->ranges_offset = w->ray_off[robot];
  return UMRS_OK;
}

int umrs_world_set_robot_state(umrs_world* w, const float* pose, const float* vel) {
  if (w == NULL || pose == NULL) return UMRS_ERR_INVALID;
  const size_t WR = (size_t)w->W * w->R;
  std::vector<RobotState> s(WR);
  memset(s.data(), 0, sizeof(RobotState) * WR);
  for (size_t i = 0; i < WR; ++i) {
    for (int k = 0; k < 3; ++k) {
      s[i].pose[k] = pose[3 * i + k];
      s[i].vel[k] = vel ? vel[3 * i + k] : 0.f;
    }
    // DiffDriveModel keeps its own double copies (diff_drive_model.cpp:133-139)
    s[i].linear_vel = (double)s[i].vel[0];
    s[i].angular_vel = (double)s[i].vel[2];
    s[i].cmd_age = kAgeNever;  // t_last_cmd_(0)
  }
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(w->d_rs[w->cur], s.data(), sizeof(RobotState) * WR, cudaMemcpyHostToDevice));
  UMRS_CUDA(cudaMemcpy(w->d_pose, pose, sizeof(float) * WR * 3, cudaMemcpyHostToDevice));
  return UMRS_OK;
}

int umrs_world_get_robot_state(umrs_world* w, float* pose, float* vel) {
  if (w == NULL) return UMRS_ERR_INVALID;
  const size_t WR = (size_t)w->W * w->R;
  std::vector<RobotState> s(WR);
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(s.data(), w->d_rs[w->cur], sizeof(RobotState) * WR, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < WR; ++i)
    for (int k = 0; k < 3; ++k) {
      if (pose) pose[3 * i + k] = s[i].pose[k];
      if (vel) vel[3 * i + k] = s[i].vel[k];
    }
  return UMRS_OK;
}

int umrs_world_set_humans(umrs_world* w, const float* start, const float* goal) {
  if (w == NULL) return UMRS_ERR_INVALID;
  if (w->H == 0) return UMRS_OK;
  const size_t WH = (size_t)w->W * w->H;
  std::vector<HumanState> s(WH);
  memset(s.data(), 0, sizeof(HumanState) * WH);
  for (size_t i = 0; i < WH; ++i) {
    const umrs_human_params& p = w->humans[i % w->H];
    const float ds[3] = {p.start_x, p.start_y, p.start_theta};
    const float dg[3] = {p.goal_x, p.goal_y, p.goal_theta};
    for (int k = 0; k < 3; ++k) {
      s[i].start[k] = start ? start[3 * i + k] : ds[k];
      s[i].goal[k] = goal ? goal[3 * i + k] : dg[k];
      s[i].pose[k] = s[i].start[k];  // human_object.cpp:83-84
    }
  }
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(w->d_hs[w->cur], s.data(), sizeof(HumanState) * WH, cudaMemcpyHostToDevice));
  return UMRS_OK;
}

int umrs_world_get_human_poses(umrs_world* w, float* pose) {
  if (w == NULL || (pose == NULL && w->H > 0)) return UMRS_ERR_INVALID;
  const size_t WH = (size_t)w->W * w->H;
  if (WH == 0) return UMRS_OK;
  std::vector<HumanState> s(WH);
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(s.data(), w->d_hs[w->cur], sizeof(HumanState) * WH, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < WH; ++i)
    for (int k = 0; k < 3; ++k) pose[3 * i + k] = s[i].pose[k];
  return UMRS_OK;
}

int umrs_world_set_static_objects(umrs_world* w, const float* pose) {
  if (w == NULL || (pose == NULL && w->O > 0)) return UMRS_ERR_INVALID;
  if (w->O == 0) return UMRS_OK;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(w->d_static, pose, sizeof(float) * 3 * (size_t)w->W * w->O, cudaMemcpyHostToDevice));
  return UMRS_OK;
}

int umrs_world_get_object_lines(umrs_world* w, float* segs) {
  if (w == NULL || (segs == NULL && w->L_obj > 0)) return UMRS_ERR_INVALID;
  if (w->L_obj == 0) return UMRS_OK;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(segs, w->d_obj_lines, sizeof(float4) * (size_t)w->W * w->L_obj, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

}  // extern "C"

// ================================== stepping ========================================
// Launches worlds [first, first + count) of the current tick; `last` closes the tick
// (state buffers flip, step counter advances).
static int world_launch(umrs_world* w, const float* cmds_dev, const uint8_t* mask_dev,
                        cudaStream_t stream, int32_t first, int32_t count, bool last, int buf = 0) {
  WorldStepParams p;
  memset(&p, 0, sizeof(p));
  p.c.src = w->plan.src;
  p.c.cap = w->plan.cap;
  p.c.max_rows = w->plan.max_rows;
  p.c.max_rays = w->max_rays;
  p.c.multipass = w->plan.multipass;
  p.c.tma_stage = w->plan.tma_stage;
  p.c.ranges = buf ? w->d_ranges2 : w->d_ranges;
  p.c.hit_idx = buf ? w->d_idx2 : w->d_idx;
  p.c.counters = w->counters_on ? w->d_counters : nullptr;
  p.c.flags = w->d_flags;
  p.W = w->W; p.R = w->R; p.H = w->H; p.O = w->O;
  p.world_first = first; p.world_count = count;
  p.robots = w->d_robots;
  p.humans = w->d_humans;
  p.tmpl = w->d_tmpl;
  p.line_ent = w->d_line_ent;
  p.ent_off = w->d_ent_off;
  p.L_obj = w->L_obj;
  p.rs_in = w->d_rs[w->cur];
  p.rs_out = w->d_rs[w->cur ^ 1];
  p.hs_in = w->d_hs[w->cur];
  p.hs_out = w->d_hs[w->cur ^ 1];
  p.static_pose = w->d_static;
  p.cmds = cmds_dev;
  p.cmd_mask = mask_dev;
  p.ray_off = w->d_ray_off;
  p.num_rays = w->d_num_rays;
  p.rays_per_world = w->rays_per_world;
  p.fan_params = w->d_fan_params;
  p.obj_lines_out = w->d_obj_lines;
  p.robot_pose_out = w->d_pose;
  p.dt = w->desc.dt;
  p.key0 = (uint32_t)(w->desc.seed & 0xffffffffu);
  p.key1 = (uint32_t)(w->desc.seed >> 32);
  p.step_lo = (uint32_t)(w->step & 0xffffffffu);
  p.step_hi = (uint32_t)(w->step >> 32);
  p.world_base = w->desc.world_index_base;
  p.laser_noise = w->desc.laser_noise_enabled;
  p.motion_noise = w->desc.motion_noise_enabled;
  if (w->counters_on && first == 0)
    UMRS_CUDA(cudaMemsetAsync(w->d_counters, 0, sizeof(unsigned long long) * kNumCounters, stream));
  p.skip_scan = w->desc.reference_render ? 1 : 0;
  p.cars_only = w->H == 0 ? 1 : 0;
  for (const umrs_robot_params& rp : w->robots)
    if (rp.model_type != UMRS_ACKERMANN_DRIVE) p.cars_only = 0;
  UMRS_CUDA(launch_world_step(p, stream));
  if (w->desc.reference_render) {
    // the scans: the reference's own render over this launch's fans, on the fan parameters and
    // object lines the step kernel has just written (stream order)
    RenderScanParams q;
    memset(&q, 0, sizeof(q));
    q.segs = w->map_segs_in_map_order;
    q.n_map = w->n_map;
    q.fans = w->d_fan_params;
    q.ranges = buf ? w->d_ranges2 : w->d_ranges;
    q.scratch_lines = w->d_render_lines;
    q.scratch_scene = w->d_render_scene;
    q.scratch_c = w->d_render_c;
    q.world_R = w->R;
    q.world_obj_lines = w->d_obj_lines;
    q.L_obj = w->L_obj;
    q.ent_off = w->d_ent_off;
    q.first_robot_ent = w->O + w->H;
    q.robots = w->d_robots;
    q.ray_off = w->d_ray_off;
    q.num_rays_r = w->d_num_rays;
    q.rays_per_world = w->rays_per_world;
    q.laser_noise = w->desc.laser_noise_enabled;
    q.key0 = p.key0; q.key1 = p.key1; q.step_lo = p.step_lo; q.step_hi = p.step_hi;
    q.world_base = w->desc.world_index_base;
    const int f0 = first * w->R, f1 = (first + count) * w->R;
    for (int b = f0; b < f1; b += w->render_cap)
      UMRS_CUDA(launch_reference_render(q, b, b + w->render_cap < f1 ? b + w->render_cap : f1, stream));
  }
  if (last) {
    w->cur ^= 1;
    w->step += 1;
    w->out_buf = buf;
  }
  return UMRS_OK;
}

extern "C" {

int umrs_world_step_device(umrs_world* w, const float* cmds_dev, const uint8_t* mask_dev,
                           void* stream) {
  if (w == NULL) return UMRS_ERR_INVALID;
  cudaStream_t st = stream ? (cudaStream_t)stream : w->stream;
  // this launch overwrites output buffer 0: a device->host copy a pipelined host-buffer tick still
  // has in flight out of it must be done first (the event stays recorded; a wait clears the flag)
  if (w->buf_busy[0]) UMRS_CUDA(cudaStreamWaitEvent(st, w->buf_free[0], 0));
  return world_launch(w, cmds_dev, mask_dev, st, 0, w->W, true, 0);
}

}  // extern "C"

// The two device output buffers of the pipelined path exist from the first pipelined tick on.
static int ensure_second_buffers(umrs_world* w) {
  const size_t n_rays = (size_t)w->W * w->rays_per_world;
  if (w->d_ranges2 == nullptr) {
    UMRS_CUDA(dalloc(&w->d_ranges2, n_rays));
    UMRS_CUDA(cudaMemsetAsync(w->d_ranges2, 0, sizeof(float) * n_rays, w->stream));
  }
  if (w->d_idx != nullptr && w->d_idx2 == nullptr) UMRS_CUDA(dalloc(&w->d_idx2, n_rays));
  return UMRS_OK;
}

// One tick with host buffers, enqueued on the world's two streams (no synchronisation here).
//   blocking tick (umrs_world_step): cut into world slices on output buffer 0 -- the device->host
//     copy of slice k runs on the copy stream while slice k+1 computes, so the call costs about
//     max(kernel, PCIe) + one slice instead of their sum;
//   pipelined tick (umrs_world_step_async): ONE launch into the output buffer that is not being
//     copied, ONE copy behind it; tick k's copy runs under tick k+1's kernel, and the kernel only
//     ever waits for the copy of tick k-1 (long done in steady state).  Seven API calls per tick.
static int step_enqueue(umrs_world* w, const float* cmds, const uint8_t* cmd_mask,
                        float* ranges_out, int32_t* hit_idx_out, bool pipelined) {
  if (w == NULL) return UMRS_ERR_INVALID;
  if (hit_idx_out != NULL && w->d_idx == NULL) return UMRS_ERR_INVALID;   // also: reference-render worlds
  const size_t WR = (size_t)w->W * w->R;
  const size_t n_rays = (size_t)w->W * w->rays_per_world;
  if (!pipelined && w->async_pending) {
    // a blocking tick after pipelined ones: let their copies finish first
    UMRS_CUDA(cudaStreamSynchronize(w->copy_stream));
    w->async_pending = false;
    w->buf_busy[0] = w->buf_busy[1] = false;
  }
  int buf = 0;
  if (pipelined) {
    const int st2 = ensure_second_buffers(w);
    if (st2 != UMRS_OK) return st2;
    buf = (int)(w->async_ticks & 1);
    if (w->buf_busy[buf]) UMRS_CUDA(cudaStreamWaitEvent(w->stream, w->buf_free[buf], 0));
  }
  if (cmds)
    UMRS_CUDA(cudaMemcpyAsync(w->d_cmds, cmds, sizeof(float) * WR * 3, cudaMemcpyHostToDevice, w->stream));
  if (cmds && cmd_mask)
    UMRS_CUDA(cudaMemcpyAsync(w->d_mask, cmd_mask, WR, cudaMemcpyHostToDevice, w->stream));
  const float* dc = cmds ? w->d_cmds : nullptr;
  const uint8_t* dm = (cmds && cmd_mask) ? w->d_mask : nullptr;
  if (pipelined) {
    const int st = world_launch(w, dc, dm, w->stream, 0, w->W, true, buf);
    if (st != UMRS_OK) return st;
    // the tick is complete when its kernel AND its copies are
    UMRS_CUDA(cudaEventRecord(w->kernels_done, w->stream));
    UMRS_CUDA(cudaStreamWaitEvent(w->copy_stream, w->kernels_done, 0));
    if (ranges_out)
      UMRS_CUDA(cudaMemcpyAsync(ranges_out, buf ? w->d_ranges2 : w->d_ranges, sizeof(float) * n_rays,
                                cudaMemcpyDeviceToHost, w->copy_stream));
    if (hit_idx_out)
      UMRS_CUDA(cudaMemcpyAsync(hit_idx_out, buf ? w->d_idx2 : w->d_idx, sizeof(int32_t) * n_rays,
                                cudaMemcpyDeviceToHost, w->copy_stream));
    if (ranges_out != NULL || hit_idx_out != NULL) {
      UMRS_CUDA(cudaEventRecord(w->buf_free[buf], w->copy_stream));
      w->buf_busy[buf] = true;
      w->async_pending = true;
    }
    UMRS_CUDA(cudaEventRecord(w->tick_done[w->async_ticks & 3], w->copy_stream));
    w->async_ticks += 1;
    return UMRS_OK;
  }
  // blocking tick: 8 slices (first-slice latency matters) -- measured on B200
  int n_slices = 1;
  if (ranges_out != NULL && w->W >= 1024) n_slices = 8;
  if (g_tuning.step_slices > 0) n_slices = g_tuning.step_slices;   // umrs_tuning_set("step_slices")
  if (n_slices > w->W) n_slices = w->W;
  if (n_slices > kMaxSlices) n_slices = kMaxSlices;
  const size_t rpw = (size_t)w->rays_per_world;
  for (int s = 0; s < n_slices; ++s) {
    const int32_t first = (int32_t)((int64_t)w->W * s / n_slices);
    const int32_t end = (int32_t)((int64_t)w->W * (s + 1) / n_slices);
    int st = world_launch(w, dc, dm, w->stream, first, end - first, s == n_slices - 1, 0);
    if (st != UMRS_OK) return st;
    if (ranges_out == NULL && hit_idx_out == NULL) continue;
    UMRS_CUDA(cudaEventRecord(w->slice_done[s], w->stream));
    UMRS_CUDA(cudaStreamWaitEvent(w->copy_stream, w->slice_done[s], 0));
    const size_t off = (size_t)first * rpw, cnt = (size_t)(end - first) * rpw;
    if (ranges_out)
      UMRS_CUDA(cudaMemcpyAsync(ranges_out + off, w->d_ranges + off, sizeof(float) * cnt,
                                cudaMemcpyDeviceToHost, w->copy_stream));
    if (hit_idx_out)
      UMRS_CUDA(cudaMemcpyAsync(hit_idx_out + off, w->d_idx + off, sizeof(int32_t) * cnt,
                                cudaMemcpyDeviceToHost, w->copy_stream));
  }
  return UMRS_OK;
}

extern "C" {

int umrs_world_step(umrs_world* w, const float* cmds, const uint8_t* cmd_mask,
                    float* ranges_out, int32_t* hit_idx_out) {
  const int st = step_enqueue(w, cmds, cmd_mask, ranges_out, hit_idx_out, false);
  if (st != UMRS_OK) return st;
  UMRS_CUDA(cudaStreamSynchronize(w->stream));
  UMRS_CUDA(cudaStreamSynchronize(w->copy_stream));
  w->async_pending = false;
  w->buf_busy[0] = w->buf_busy[1] = false;
  if (w->plan.kernel == 1) return capi_read_flags(w->d_flags);   // only the grid walk can raise one
  return UMRS_OK;
}

int umrs_world_check(umrs_world* w) {
  if (w == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());
  return capi_read_flags(w->d_flags);
}

// ---- checkpoint / resume -----------------------------------------------------------------
namespace {
struct StateHeader {
  uint32_t magic, version;
  int32_t W, R, H, O;
  uint64_t seed, step;
  uint64_t bytes;
};
constexpr uint32_t kStateMagic = 0x53524D55u;   // "UMRS"
uint64_t state_bytes(const umrs_world* w) {
  return sizeof(StateHeader) + sizeof(RobotState) * (uint64_t)w->W * w->R +
         sizeof(HumanState) * (uint64_t)w->W * w->H + sizeof(float) * 3ull * (uint64_t)w->W * w->O;
}
}  // namespace

int umrs_world_state_size(const umrs_world* w, uint64_t* bytes) {
  if (w == NULL || bytes == NULL) return UMRS_ERR_INVALID;
  *bytes = state_bytes(w);
  return UMRS_OK;
}

int umrs_world_get_state(umrs_world* w, void* blob, uint64_t capacity) {
  if (w == NULL || blob == NULL || capacity < state_bytes(w)) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());   // the last step may have run on a caller stream
  StateHeader h;
  memset(&h, 0, sizeof(h));
  h.magic = kStateMagic; h.version = 1;
  h.W = w->W; h.R = w->R; h.H = w->H; h.O = w->O;
  h.seed = w->desc.seed; h.step = w->step; h.bytes = state_bytes(w);
  unsigned char* p = static_cast<unsigned char*>(blob);
  memcpy(p, &h, sizeof(h));
  p += sizeof(h);
  const size_t nr = sizeof(RobotState) * (size_t)w->W * w->R, nh = sizeof(HumanState) * (size_t)w->W * w->H;
  const size_t no = sizeof(float) * 3 * (size_t)w->W * w->O;
  UMRS_CUDA(cudaMemcpy(p, w->d_rs[w->cur], nr, cudaMemcpyDeviceToHost));
  p += nr;
  if (nh) UMRS_CUDA(cudaMemcpy(p, w->d_hs[w->cur], nh, cudaMemcpyDeviceToHost));
  p += nh;
  if (no) UMRS_CUDA(cudaMemcpy(p, w->d_static, no, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

int umrs_world_set_state(umrs_world* w, const void* blob, uint64_t bytes) {
  if (w == NULL || blob == NULL || bytes < sizeof(StateHeader)) return UMRS_ERR_INVALID;
  StateHeader h;
  memcpy(&h, blob, sizeof(h));
  if (h.magic != kStateMagic || h.version != 1 || h.W != w->W || h.R != w->R || h.H != w->H || h.O != w->O ||
      h.seed != w->desc.seed || h.bytes != state_bytes(w) || bytes < h.bytes) {
    capi_set_error("umrs_world_set_state: snapshot of another world shape or seed");
    return UMRS_ERR_INVALID;
  }
  UMRS_CUDA(cudaDeviceSynchronize());
  w->async_pending = false;
  w->buf_busy[0] = w->buf_busy[1] = false;
  const unsigned char* p = static_cast<const unsigned char*>(blob) + sizeof(h);
  const size_t WR = (size_t)w->W * w->R;
  const size_t nr = sizeof(RobotState) * WR, nh = sizeof(HumanState) * (size_t)w->W * w->H;
  const size_t no = sizeof(float) * 3 * (size_t)w->W * w->O;
  UMRS_CUDA(cudaMemcpy(w->d_rs[w->cur], p, nr, cudaMemcpyHostToDevice));
  // the convenience pose copy follows the state
  std::vector<float> pose(WR * 3);
  const RobotState* rs = reinterpret_cast<const RobotState*>(p);
  for (size_t i = 0; i < WR; ++i) {
    RobotState one;
    memcpy(&one, rs + i, sizeof(one));
    for (int k = 0; k < 3; ++k) pose[3 * i + k] = one.pose[k];
  }
  UMRS_CUDA(cudaMemcpy(w->d_pose, pose.data(), sizeof(float) * WR * 3, cudaMemcpyHostToDevice));
  p += nr;
  if (nh) UMRS_CUDA(cudaMemcpy(w->d_hs[w->cur], p, nh, cudaMemcpyHostToDevice));
  p += nh;
  if (no) UMRS_CUDA(cudaMemcpy(w->d_static, p, no, cudaMemcpyHostToDevice));
  w->step = h.step;
  return UMRS_OK;
}

int umrs_world_step_async(umrs_world* w, const float* cmds, const uint8_t* cmd_mask,
                          float* ranges_out, int32_t* hit_idx_out) {
  return step_enqueue(w, cmds, cmd_mask, ranges_out, hit_idx_out, true);
}

int umrs_world_wait_outputs(umrs_world* w, int32_t ticks_back) {
  if (w == NULL || ticks_back < 0 || ticks_back > 2) return UMRS_ERR_INVALID;
  if (w->async_ticks <= (uint64_t)ticks_back) return UMRS_OK;   // nothing that old was issued
  UMRS_CUDA(cudaEventSynchronize(w->tick_done[(w->async_ticks - 1 - (uint64_t)ticks_back) & 3]));
  return UMRS_OK;
}

int umrs_world_wait(umrs_world* w) {
  if (w == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaStreamSynchronize(w->stream));
  UMRS_CUDA(cudaStreamSynchronize(w->copy_stream));
  w->async_pending = false;
  w->buf_busy[0] = w->buf_busy[1] = false;
  return UMRS_OK;
}

int umrs_world_device_buffers(umrs_world* w, float** ranges_dev, int32_t** hit_idx_dev,
                              float** robot_pose_dev) {
  if (w == NULL) return UMRS_ERR_INVALID;
  if (ranges_dev) *ranges_dev = w->out_buf ? w->d_ranges2 : w->d_ranges;
  if (hit_idx_dev) *hit_idx_dev = w->out_buf ? w->d_idx2 : w->d_idx;
  if (robot_pose_dev) *robot_pose_dev = w->d_pose;
  return UMRS_OK;
}

int umrs_world_sync(umrs_world* w) {
  if (w == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaStreamSynchronize(w->stream));
  return UMRS_OK;
}

int umrs_world_step_count(const umrs_world* w, uint64_t* steps) {
  if (w == NULL || steps == NULL) return UMRS_ERR_INVALID;
  *steps = w->step;
  return UMRS_OK;
}

int umrs_world_enable_counters(umrs_world* w, int32_t on) {
  if (w == NULL) return UMRS_ERR_INVALID;
  w->counters_on = on != 0;
  return UMRS_OK;
}

int umrs_world_read_counters(umrs_world* w, uint64_t* tests, uint64_t* segs, uint64_t* pairs) {
  if (w == NULL) return UMRS_ERR_INVALID;
  unsigned long long h[4] = {0, 0, 0, 0};
  UMRS_CUDA(cudaDeviceSynchronize());  // the step may have run on a caller stream
  UMRS_CUDA(cudaMemcpy(h, w->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
  if (tests) *tests = h[0];
  if (segs) *segs = h[1];
  if (pairs) *pairs = h[2];
  return UMRS_OK;
}

int umrs_world_read_counters_ex(umrs_world* w, uint64_t out8[8]) {
  if (w == NULL || out8 == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(out8, w->d_counters, sizeof(uint64_t) * 8, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

int umrs_world_read_phase_cycles(umrs_world* w, uint64_t out8[8]) {
  if (w == NULL || out8 == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(out8, w->d_counters + 8, sizeof(uint64_t) * 8, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

// Fan parameters the last step actually used: [W][R][4] {loc_x, loc_y, angle_min,
// angle_max} (world frame).  Lets a checker feed the oracle the identical fan.
int umrs_world_get_fan_params(umrs_world* w, float* out) {
  if (w == NULL || out == NULL) return UMRS_ERR_INVALID;
  UMRS_CUDA(cudaDeviceSynchronize());
  UMRS_CUDA(cudaMemcpy(out, w->d_fan_params, sizeof(float4) * (size_t)w->W * w->R, cudaMemcpyDeviceToHost));
  return UMRS_OK;
}

int umrs_world_stats(umrs_world* w, double* out4) {
  if (w == NULL || out4 == NULL) return UMRS_ERR_INVALID;
  const unsigned long long inf_bits = 0x7f800000ull;
  double init[4] = {0.0, 0.0, 0.0, 0.0};
  memcpy(&init[1], &inf_bits, sizeof(double));
  UMRS_CUDA(cudaDeviceSynchronize());  // the last step may have run on a caller stream
  UMRS_CUDA(cudaMemcpyAsync(w->d_stats, init, sizeof(init), cudaMemcpyHostToDevice, w->stream));
  UMRS_CUDA(launch_stats(w->out_buf ? w->d_ranges2 : w->d_ranges, (size_t)w->W * w->rays_per_world, w->d_stats, w->stream));
  double raw[4];
  UMRS_CUDA(cudaMemcpyAsync(raw, w->d_stats, sizeof(raw), cudaMemcpyDeviceToHost, w->stream));
  UMRS_CUDA(cudaStreamSynchronize(w->stream));
  unsigned long long mnb, mxb;
  memcpy(&mnb, &raw[1], 8);
  memcpy(&mxb, &raw[2], 8);
  const uint32_t mn32 = (uint32_t)mnb, mx32 = (uint32_t)mxb;
  float mn, mx;
  memcpy(&mn, &mn32, 4);
  memcpy(&mx, &mx32, 4);
  out4[0] = raw[0];
  out4[1] = raw[0] > 0 ? (double)mn : 0.0;
  out4[2] = (double)mx;
  out4[3] = raw[3];
  return UMRS_OK;
}

}  // extern "C"
