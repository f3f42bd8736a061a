// Internal host<->kernel contract of libumrs (not part of the C-ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/umrs.h"

namespace umrs {

// ---- device-resident world state (one record per entity, 64 B) ---------------
// Mirrors the members the reference keeps per model instance:
// EntityBase::pose_ (entity_base.h:36), RobotModel::vel_ (robot_model.h:31),
// last_cmd_/target_*_vel_ + t_last_cmd_ (ackermann_model.h:15-16,
// diff_drive_model.h:28-29), DiffDriveModel::linear_vel_/angular_vel_ (double,
// diff_drive_model.h:30-31).
struct RobotState {
  float pose[3];
  float vel[3];
  float cmd[3];
  int32_t cmd_age;  // ticks since the last accepted command (sim-time watchdog)
  double linear_vel;
  double angular_vel;
  float pad[2];
};
static_assert(sizeof(RobotState) == 64, "RobotState must be 64 B");

// HumanObject members, human_object.h:40-52.
struct HumanState {
  float pose[3];
  float start[3];
  float goal[3];
  float trans_vel[2];
  float rot_vel;
  float pad[4];
};
static_assert(sizeof(HumanState) == 64, "HumanState must be 64 B");

constexpr int kAgeNever = 1 << 28;
#ifndef UMRS_SCAN_THREADS
#define UMRS_SCAN_THREADS 256
#endif
constexpr int kScanThreads = UMRS_SCAN_THREADS;  // threads of the widest fan CTA shape (umrs_kernels.cu: launchers)
constexpr int kMaxRays = 16384;    // per fan (bounded by the per-ray best arrays in smem)

// ---- where a fan's candidate segments come from ----------------------------------
// K1 ("shared-memory scan"): gx == 0, the whole cleaned map [0, n) in index order.
// K2 ("uniform-grid walk"):  the map re-ordered by the grid cell of each segment's
// midpoint (row-major cells), so that the cells [cx0, cx1] of one grid row are ONE
// contiguous range of `segs`; segments longer than two cells are kept in an
// always-visited tail [n_binned, n).  orig[k] = index of sorted segment k in the
// caller's map (the hit index the reference's GetSceneLines order defines,
// vector_map.cpp:200-214).
struct MapSource {
  const float4* segs;
  const int32_t* orig;        // nullptr: identity
  int32_t n;
  int32_t n_binned;           // == n when gx == 0
  const int32_t* cell_start;  // [gx * gy + 1], into segs
  int32_t gx, gy;
  float x0, y0, cell, inv_cell;
  float hmax_x, hmax_y;       // max |endpoint - midpoint| over binned segments
  int32_t n_map_total;        // object line o gets hit index n_map_total + o
};

constexpr int kNumCounters = 16;
constexpr uint32_t kFlagRowsTruncated = 1u;   // build_rows ran out of row slots: far grid rows not scanned

// Margin by which the grid walk widens its query box (K2): the largest midpoint-to-endpoint
// offset of a binned segment, plus a term that dominates every float rounding in the cell
// computation and in the host binning.  ONE definition for the kernel (build_rows) and for the
// host's bound on the rows a fan can touch (capi_scan_plan).
__host__ __device__ inline float grid_margin(float hmax, float cell, float abs_l, float range_max) {
  return hmax + 1e-3f * cell + 1e-5f * (abs_l + range_max);
}

// Process-wide tuning hooks (umrs_tuning_set); 0 / -1 = default.
struct Tuning {
  int cta_shape = -1;
  int cap = 0;
  int grid_from = 0;
  int step_slices = 0;
  int tma_stage = -1;   // -1 auto (the plan decides), 0 off, 1 on wherever the plan allows it
  int split_step = -1;       // world step as entity kernel + scan kernel: -1 auto (large batches), 0 never, 1 always
  int render_cta_fans = 0;   // reference render: batches up to this many fans run one 4-warp CTA per fan (0 = default)
};
extern Tuning g_tuning;

struct ScanCommon {
  MapSource src;
  float* ranges;                 // [fans][num_rays] or [W][rays_per_world]
  int32_t* hit_idx;              // nullable, same shape
  unsigned long long* counters;  // nullable: [kNumCounters] {tests, segs_prepared, chunk_seg_pairs, passes,
                                 // tie sweeps, spilled, sweep-all fans, -, 8 phase-cycle sums (dev builds)}
  uint32_t* flags;               // nullable: sticky device-side conditions (kFlagRowsTruncated)
  int32_t cap;                   // candidate slots per pass (smem sizing)
  int32_t max_rows;              // source rows a fan can touch (smem sizing)
  int32_t max_rays;              // largest num_rays of any fan of the launch (smem sizing)
  int32_t multipass;             // a fan's source set can exceed cap (keeps best_q in smem)
  int32_t tma_stage;             // 0: the gather reads the source rows with per-thread loads;
                                 // 1: K1, the whole map is staged into shared memory by ONE bulk copy
                                 //    (cp.async.bulk + mbarrier) issued before the entity steps;
                                 // 2: K2, one bulk copy per grid row, issued once the rows are known
  uint32_t smem_off[17];         // byte offsets of the shared-memory arrays (smem_layout) + total, filled by the
                                 // launch functions so that no thread recomputes the carve-up
};

struct ScanBatchParams {
  ScanCommon c;
  const float4* obj_segs;  // shared object lines (nullable)
  int32_t n_obj;
  const float4* fans;      // {loc_x, loc_y, angle_min, angle_max}
  int32_t n_fans;
  float range_max;
  int32_t num_rays;
};

struct WorldStepParams {
  ScanCommon c;
  int32_t W, R, H, O;
  int32_t world_first;              // first world of this launch (slices of one step)
  int32_t world_count;
  const umrs_robot_params* robots;  // [R]
  const umrs_human_params* humans;  // [H]
  const float4* tmpl;               // [L_obj] template line of world-line l
  const int16_t* line_ent;          // [L_obj] entity of world-line l
  const int32_t* ent_off;           // [O+H+R+1] first world-line of entity e
  int32_t L_obj;
  const RobotState* rs_in;
  RobotState* rs_out;
  const HumanState* hs_in;
  HumanState* hs_out;
  const float* static_pose;  // [W][O][3]
  const float* cmds;         // [W][R][3]
  const uint8_t* cmd_mask;   // nullable [W][R]
  const int32_t* ray_off;    // [R+1]
  const int32_t* num_rays;   // [R]
  int32_t rays_per_world;
  float4* fan_params;        // [W][R] {loc_x, loc_y, angle_min, angle_max} used
  float4* obj_lines_out;     // [W][L_obj]
  float* robot_pose_out;     // [W][R][3] (convenience copy for consumers)
  float dt;
  uint32_t key0, key1;
  uint32_t step_lo, step_hi;
  int32_t world_base;
  int32_t laser_noise, motion_noise;
  int32_t skip_scan;         // stop after the entity steps / object lines / fan (reference render follows)
  int32_t pre_stepped;       // set by the launcher: k_world_entities has run, the fan kernel only scans
  int32_t cars_only;         // every robot is an Ackermann car and there are no humans: the kernel instance
                             // without the other models' code is launched
};

// Opt-in reference render (umrs_render.cu): one warp per fan, scratch for the fans of ONE launch.
struct RenderScanParams {
  const float4* segs;      // cleaned map lines in MAP ORDER (GetSceneLines order)
  int32_t n_map;
  const float4* obj_segs;  // shared object lines (nullable)
  int32_t n_obj;
  const float4* fans;      // {loc_x, loc_y, angle_min, angle_max}
  int32_t n_fans;
  float range_max;
  int32_t num_rays;
  float* ranges;           // [n_fans][num_rays]
  float4* scratch_lines;   // [fans per launch][2000] candidate list, then the cast constants
  float4* scratch_scene;   // [fans per launch][2000]
  float* scratch_c;        // [fans per launch][kRenderScratchCWords]: cast constants, intervals, candidate list
  int32_t fan_begin;       // first fan of this launch (scratch slot = fan - fan_begin)
  // world mode (world_R > 0): fan g = (world g / R, robot g % R) of a world step; fans = the
  // step's fan_params, object lines per world minus the robot's own body, per-robot laser
  int32_t world_R;
  const float4* world_obj_lines;      // [W][L_obj]
  int32_t L_obj;
  const int32_t* ent_off;             // [entities + 1] first object line of entity e
  int32_t first_robot_ent;
  const umrs_robot_params* robots;    // [R]
  const int32_t* ray_off;             // [R]
  const int32_t* num_rays_r;          // [R]
  int32_t rays_per_world;
  int32_t laser_noise;
  uint32_t key0, key1, step_lo, step_hi;
  int32_t world_base;
};
constexpr int kRenderScratchCWords = 2000 + 2 * 2000 + 4 * 1000;   // per fan: float c[2000], uint2 iv[2000], 4 x u16 list[2000]
constexpr int kRenderFansPerLaunch = 16384;  // 1.7 GB of scratch (104 KB per fan); 4144 fans are resident at a time, the rest refill the SMs as fans finish
cudaError_t launch_reference_render(const RenderScanParams& p, int fan_begin, int fan_end,
                                    cudaStream_t stream);

// Dynamic shared memory of one fan CTA.
size_t scan_smem_bytes(int cap, int max_rows, int max_rays, int n_obj_lines, int n_ent,
                       int multipass, int tma_stage);
bool tma_stage_built();   // the bulk-copy staging variant is compiled in (-DUMRS_TMA_STAGE)

cudaError_t launch_scan_batch(const ScanBatchParams& p, cudaStream_t stream);
cudaError_t launch_world_step(const WorldStepParams& p, cudaStream_t stream);
cudaError_t launch_stats(const float* ranges, size_t n, double* out4_dev,
                         cudaStream_t stream);

// Launch accounting (umrs_kernel_launch_count).
void note_launch(int n = 1);

}  // namespace umrs
