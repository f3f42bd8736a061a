/* umrs.h -- C-ABI of the B200-native batched scan core ("umrs" = ut multi-robot
 * sim).  This is the drop-in boundary for ONE hot path of ut_multirobot_sim:
 * the simulated 2-D laser scan plus the pose integration that feeds it.
 *
 * Citations are file:line in the reference tree (ut-amrl/ut_multirobot_sim).
 * Each entry point names the reference interface it replaces:
 *
 *   umrs_map_create / umrs_map_load_file   VectorMap::lines after VectorMap::Load
 *                                          (src/simulator/vector_map.h:75,78;
 *                                           vector_map.cpp:397-411)
 *   umrs_predicted_scan                    VectorMap::GetPredictedScan
 *                                          (vector_map.h:66-72, vector_map.cpp:420-478)
 *                                          called from Simulator::publishLaser
 *                                          (simulator.cpp:405-411)
 *   umrs_scan_batch                        the same routine over many fans
 *   umrs_world_*                           Simulator::update + publishLaser
 *                                          (simulator.cpp:474-509, 393-426) with
 *                                          the EntityBase/RobotModel interface
 *                                          (entity_base.h:45-53, robot_model.h:36-37)
 *                                          batched over W independent worlds
 *
 * Conventions: extern "C"; plain pointers and sizes; every function returns an
 * int status (UMRS_OK == 0); no exceptions and no exit() cross the boundary (the
 * reference exit(1)s on a missing map, vector_map.cpp:399-402).  Handles are
 * not thread-safe (the reference is single-threaded); one handle belongs to
 * the CUDA device that was current when it was created.  There is no CPU
 * fallback: without a usable sm_100 device every compute entry point returns
 * UMRS_ERR_NO_DEVICE / UMRS_ERR_CUDA.
 */
#ifndef UMRS_H_
#define UMRS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UMRS_VERSION 100

/* ---- status codes --------------------------------------------------------- */
#define UMRS_OK 0
#define UMRS_ERR_INVALID 1     /* bad argument (null pointer, negative size ...) */
#define UMRS_ERR_CUDA 2        /* a CUDA call failed; see umrs_last_cuda_error() */
#define UMRS_ERR_NOMEM 3       /* host or device allocation failed */
#define UMRS_ERR_UNSUPPORTED 4 /* e.g. candidate set does not fit shared memory */
#define UMRS_ERR_NO_DEVICE 5   /* no CUDA device / not sm_100 */
#define UMRS_ERR_IO 6          /* map file could not be read */

int umrs_version(void);
const char* umrs_status_string(int status);
/* Text of the last CUDA error seen by this thread's calls ("" if none). */
const char* umrs_last_cuda_error(void);
/* Number of kernels THIS LIBRARY has launched since load (all handles). */
uint64_t umrs_kernel_launch_count(void);

/* Device selection for the calling thread (cudaSetDevice inside the library's
 * own CUDA runtime instance); handles created afterwards live on that device.
 * One process per GPU is the intended deployment. */
int umrs_set_device(int32_t device);
int umrs_get_device(int32_t* device);
int umrs_device_count(int32_t* count);

/* ---- parameter PODs (host shell fills them from its parsed Lua config) ----- */

/* Robot model kinds; strings as in simulator.cpp:128-142 / sim_config.lua:51-55. */
#define UMRS_ACKERMANN_DRIVE 0
#define UMRS_OMNIDIRECTIONAL_DRIVE 1
#define UMRS_DIFF_DRIVE 2
/* "ACKERMANN_DRIVE" | "OMNIDIRECTIONAL_DRIVE" | "DIFF_DRIVE" -> kind, else -1. */
int umrs_model_type_from_string(const char* robot_type);

/* Laser keys: simulator.cpp:70-72,87-93 (laser_loc.x/.y, laser_angle_min/max/
 * increment, laser_min_range, laser_max_range, laser_noise_stddev). */
typedef struct umrs_laser_params {
  float loc_x, loc_y;
  float angle_min, angle_max, angle_increment;
  float range_min, range_max;
  float noise_stddev;
} umrs_laser_params;

/* num_rays exactly as simulator.cpp:402-404:
 * int(1.0 + (angle_max - angle_min) / angle_increment), float subtraction and
 * division, double add. */
int umrs_laser_num_rays(const umrs_laser_params* laser);

/* One robot's parameters.  Keys cited per group; unused groups are ignored. */
typedef struct umrs_robot_params {
  int32_t model_type; /* UMRS_ACKERMANN_DRIVE ... */
  /* ackermann_model.cpp:16-23 (ak_*), defaults ut_automata_config.lua:4-10 */
  float ak_min_turn_radius, ak_max_speed, ak_max_accel;
  float ak_angular_error_bias, ak_angular_error_rate;
  /* diff_drive_model.cpp:42-51, defaults ut_jackal_config.lua:14-21 */
  int32_t invert_linear_vel_cmds, invert_angular_vel_cmds;
  float linear_pos_accel_limit, linear_neg_accel_limit;
  float angular_pos_accel_limit, angular_neg_accel_limit;
  float max_angular, max_linear_vel;
  /* omnidirectional_model.cpp:21-24, defaults cobot_config.lua:13-18 */
  float co_max_accel, co_max_angle_accel, co_max_speed, co_max_angle_vel;
  /* body (simulator.cpp:65-68; the reference only draws it, simulator.cpp:318-327,
   * 380-383).  footprint_enabled != 0 makes the body a 4-segment rectangle that
   * OTHER robots' scans see (new behaviour). */
  float car_length, car_width, rear_axle_offset;
  int32_t footprint_enabled;
  umrs_laser_params laser;
} umrs_robot_params;

/* human_object.cpp:30-43 (hu_*), defaults human_config.lua.  mode: 0 Singleshot,
 * 1 Repeat (human_object.h:33-36). */
typedef struct umrs_human_params {
  float radius;
  int32_t num_segments;
  float start_x, start_y, start_theta;
  float goal_x, goal_y, goal_theta;
  float max_speed, avg_speed, max_omega, avg_omega;
  float reach_goal_threshold;
  int32_t mode;
} umrs_human_params;

/* Defaults restating the reference's config files (values cited above). */
void umrs_default_laser(umrs_laser_params* out);            /* sim_config.lua:31,39-44 */
void umrs_default_jackal_laser(umrs_laser_params* out);     /* ut_jackal_config.lua:39,44-48 */
void umrs_default_ut_automata(umrs_robot_params* out);      /* ut_automata_config.lua + sim_config.lua:25-31 */
void umrs_default_ut_jackal(umrs_robot_params* out);        /* ut_jackal_config.lua */
void umrs_default_bwibot(umrs_robot_params* out);           /* bwibot_config.lua */
void umrs_default_cobot(umrs_robot_params* out);            /* cobot_config.lua */
void umrs_default_human(umrs_human_params* out);            /* human_config.lua */

/* LaserScan-shaped header, fields as Simulator::init sets them
 * (simulator.cpp:150-159): time_increment 0, scan_time 0.05, intensities empty. */
typedef struct umrs_laser_scan_header {
  float angle_min, angle_max, angle_increment;
  float time_increment, scan_time;
  float range_min, range_max;
  int32_t num_rays;       /* length of ranges[] */
  int32_t ranges_offset;  /* offset of this robot's ranges inside one world's block */
} umrs_laser_scan_header;

/* ---- map -------------------------------------------------------------------- */
typedef struct umrs_map umrs_map;

/* segs_xyxy: n segments {x0,y0,x1,y1}, ALREADY cleaned by the host shell
 * (VectorMap::Cleanup, vector_map.cpp:367-395).  Copied to the device. */
int umrs_map_create(const float* segs_xyxy, int32_t n, umrs_map** out);
/* Reads the reference's text format "%f,%f,%f,%f" per line (vector_map.cpp:404-407)
 * and runs the host-side restatement of Cleanup (umrs_map_cleanup). */
int umrs_map_load_file(const char* path, umrs_map** out);
int umrs_map_destroy(umrs_map* map);
int umrs_map_num_segments(const umrs_map* map, int32_t* n);
/* Copies the (cleaned) host segments out; cap in segments. */
int umrs_map_get_segments(const umrs_map* map, float* segs_xyxy, int32_t cap);

/* Host-only helper, no device needed: VectorMap::Cleanup + ShrinkLine
 * (vector_map.cpp:359-395) on raw segments.  out must hold cap segments;
 * *n_out receives the cleaned count (UMRS_ERR_INVALID if cap is too small). */
int umrs_map_cleanup(const float* segs_xyxy, int32_t n, float* out_xyxy,
                     int32_t cap, int32_t* n_out);
/* The same, choosing how the "first already-accepted line that crosses" search
 * (vector_map.cpp:377-385) is done: 0 auto, 1 the reference's quadratic loop,
 * 2 a uniform grid over the accepted lines' bounding boxes.  Every choice returns
 * the same segments in the same order, bit for bit (tests/test_oracle.py); the
 * grid turns the O(L^2) map load into ~O(L) for large maps. */
int umrs_map_cleanup_ex(const float* segs_xyxy, int32_t n, float* out_xyxy, int32_t cap,
                        int32_t* n_out, int32_t algorithm);

/* The same eight counters for the scan entry points below (for the
 * grid walk report rows visited and bytes of segment data fetched per ray). */
int umrs_map_enable_counters(umrs_map* map, int32_t on);
int umrs_map_read_counters(umrs_map* map, uint64_t out8[8]);

/* ---- drop-in scan -------------------------------------------------------------
 * Same seven arguments as VectorMap::GetPredictedScan (vector_map.h:66-72) plus
 * the object lines the reference keeps in VectorMap::object_lines
 * (vector_map.h:81).  Host buffers, synchronous.  ranges_out: num_rays floats.
 * hit_idx_out (nullable): index of the hit segment in map-then-object order
 * (GetSceneLines order, vector_map.cpp:200-214), -1 where nothing covers the
 * ray (range == range_max).  range_min is ignored, as in the reference. */
int umrs_predicted_scan(umrs_map* map, const float* obj_segs_xyxy, int32_t n_obj,
                        float loc_x, float loc_y, float range_min, float range_max,
                        float angle_min, float angle_max, int32_t num_rays,
                        float* ranges_out, int32_t* hit_idx_out);

/* Many fans against one map (+ one shared object-line set).  fans: n_fans x
 * {loc_x, loc_y, angle_min, angle_max}.  Output [n_fans][num_rays].  Host
 * buffers, synchronous.  Uses the shared-memory kernel when the candidate set
 * fits, else the uniform-grid walk. */
int umrs_scan_batch(umrs_map* map, const float* obj_segs_xyxy, int32_t n_obj,
                    const float* fans, int32_t n_fans, float range_max,
                    int32_t num_rays, float* ranges_out, int32_t* hit_idx_out);
/* Same with device-resident fans/outputs on a caller stream (cudaStream_t as
 * void*, NULL = default stream); asynchronous. */
int umrs_scan_batch_device(umrs_map* map, const float* obj_segs_dev, int32_t n_obj,
                           const float* fans_dev, int32_t n_fans, float range_max,
                           int32_t num_rays, float* ranges_dev, int32_t* hit_idx_dev,
                           void* stream);
/* Which kernel umrs_scan_batch* would pick for this map: 0 = shared-memory
 * (K1), 1 = uniform grid (K2).  force: -1 auto, 0, 1. */
int umrs_map_set_scan_kernel(umrs_map* map, int32_t force);

/* Opt-in REFERENCE RENDER: GetPredictedScan exactly as the reference computes it --
 * GetSceneLines, the SceneRender/TrimOcclusion visibility pass (vector_map.cpp:55-258)
 * and "first rendered piece that covers the ray" (:458-476) -- so the ranges are
 * bit-identical to the reference's own output, occlusion-trimming defects included
 * (human_object.cpp:106).  The entry points above return the exact nearest hit instead
 * and are ~100x faster; use this one where the reference's very numbers are needed.
 * One warp per fan, O(L^2) per fan like the reference; no hit index (the reference has
 * none).  The render stops at 2000 candidates like the reference (:221). */
int umrs_scan_batch_reference_render(umrs_map* map, const float* obj_segs_xyxy, int32_t n_obj,
                                     const float* fans, int32_t n_fans, float range_max,
                                     int32_t num_rays, float* ranges_out);
int umrs_scan_batch_reference_render_device(umrs_map* map, const float* obj_segs_dev, int32_t n_obj,
                                            const float* fans_dev, int32_t n_fans, float range_max,
                                            int32_t num_rays, float* ranges_dev, void* stream);

/* The reference render for callers that must not allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and the same
 * bits as umrs_scan_batch_reference_render_device, but the per-fan scratch (104,000 bytes a
 * fan) lives in a device array compiled into the library instead of cudaMalloc'd memory:
 *   - at most UMRS_NOALLOC_MAX_FANS fans per call; more returns UMRS_ERR_INVALID (text in
 *     umrs_last_cuda_error()) and launches nothing.  The array is 6,656,000 bytes, held by
 *     every GPU that loads the library until the process exits;
 *   - the call makes no cudaMalloc / cudaFree and calls no allocating routine; map, object
 *     lines, fans and ranges are the caller's device buffers on the CURRENT device;
 *   - every call on a device shares the one array: calls that may overlap on the GPU
 *     (different streams without an ordering between them) RACE.  Order them, e.g. on one
 *     stream.  Nothing in the array has to start at zero, so calls issue no clear;
 *   - the first call on a device looks the array's address up (cudaGetSymbolAddress); make
 *     that call outside stream capture -- n_fans = 0 does the set-up and launches nothing;
 *   - stream: cudaStream_t as void*, NULL = the legacy default stream; asynchronous. */
#define UMRS_NOALLOC_MAX_FANS 64
int umrs_scan_batch_reference_render_noalloc(umrs_map* map, const float* obj_segs_dev, int32_t n_obj,
                                             const float* fans_dev, int32_t n_fans, float range_max,
                                             int32_t num_rays, float* ranges_dev, void* stream);

/* ---- batched worlds ------------------------------------------------------------
 * W independent worlds stepping in lockstep; every world has the same robot
 * roster (robots[r] describes robot r of every world), H humans and O static
 * square objects (ShortTermObject, short_term_object.cpp:53-58), all against
 * one shared static map. */
typedef struct umrs_world umrs_world;

typedef struct umrs_world_desc {
  int32_t num_worlds;
  int32_t robots_per_world;
  int32_t humans_per_world;
  int32_t static_objects_per_world;
  const umrs_robot_params* robots; /* [robots_per_world] */
  const umrs_human_params* humans; /* [humans_per_world], per-world start/goal via set_humans */
  float dt;                        /* delta_t, sim_config.lua:17 */
  uint64_t seed;                   /* Philox key */
  int32_t laser_noise_enabled;     /* 0: ranges are GetPredictedScan's, untouched */
  int32_t motion_noise_enabled;    /* 0: Ackermann angular error draws are 0 */
  int32_t world_index_base;        /* global index of world 0 (multi-GPU shards) */
  int32_t write_hit_idx;           /* keep the hit-segment index buffer */
  int32_t reference_render;        /* 1: every scan goes through the reference's own render
                                    * (see umrs_scan_batch_reference_render): the reference's bits,
                                    * O(L^2) per scan, no hit indices */
} umrs_world_desc;

/* The world keeps using the map's device buffers: the map must outlive it
 * (umrs_map_destroy returns UMRS_ERR_INVALID while a world created from it exists). */
int umrs_world_create(umrs_map* map, const umrs_world_desc* desc, umrs_world** out);
int umrs_world_destroy(umrs_world* world);

/* Rays of one world = sum over robots of num_rays; ranges buffer is
 * [num_worlds][rays_per_world]. */
int umrs_world_rays_per_world(const umrs_world* world, int32_t* n);
int umrs_world_scan_header(const umrs_world* world, int32_t robot,
                           umrs_laser_scan_header* out);
/* Object lines of one world: [static objects][humans][robot footprints]. */
int umrs_world_object_lines_per_world(const umrs_world* world, int32_t* n);

/* EntityBase::SetPose / RobotModel::SetVel for every robot: [W][R][3] {x,y,theta}.
 * vel may be NULL (zeros).  Resets command state.  Host buffers. */
int umrs_world_set_robot_state(umrs_world* world, const float* pose_xyt,
                               const float* vel_xyt);
int umrs_world_get_robot_state(umrs_world* world, float* pose_xyt, float* vel_xyt);
/* start/goal: [W][H][3]; pose := start (HumanObject ctor, human_object.cpp:83-85).
 * NULL keeps the desc's values for every world. */
int umrs_world_set_humans(umrs_world* world, const float* start_xyt,
                          const float* goal_xyt);
int umrs_world_get_human_poses(umrs_world* world, float* pose_xyt);
/* [W][O][3] */
int umrs_world_set_static_objects(umrs_world* world, const float* pose_xyt);
/* Current object lines [W][L][4] (after the last step). */
int umrs_world_get_object_lines(umrs_world* world, float* segs_xyxy);

/* One tick = Simulator::update + publishLaser for every world.
 * cmds: [W][R][3] floats; Ackermann {velocity, curvature, -}
 * (AckermannCurvatureDriveMsg.msg:5,11), diff-drive {linear.x, angular.z, -},
 * omni {velocity_x, velocity_y, velocity_r}.  cmd_mask (nullable): [W][R] bytes,
 * 0 = no new command this tick (the watchdog then ages the old one in SIM time:
 * a command older than 0.1 s is zeroed, ackermann_model.cpp:55-58).
 * Host version copies cmds in and ranges (and idx, if non-NULL) out, synchronous. */
int umrs_world_step(umrs_world* world, const float* cmds, const uint8_t* cmd_mask,
                    float* ranges_out, int32_t* hit_idx_out);
/* The same tick, enqueued and not waited for: successive calls pipeline -- the device->host
 * copy of tick k runs while tick k+1 computes (the library alternates between two device
 * output buffers, one launch + one copy per tick) -- so a host shell can publish scan k while
 * the GPU works on k+1.  The caller's buffers belong to the library until waited for: use
 * (at least) two output buffers in turn; pinned memory keeps the copies asynchronous.
 * umrs_world_wait_outputs(w, back): blocks until the outputs of the tick issued `back`
 * async calls ago (0 = the latest, at most 2) are in host memory.  umrs_world_wait: all of it. */
int umrs_world_step_async(umrs_world* world, const float* cmds, const uint8_t* cmd_mask,
                          float* ranges_out, int32_t* hit_idx_out);
int umrs_world_wait_outputs(umrs_world* world, int32_t ticks_back);
int umrs_world_wait(umrs_world* world);
/* Device version: cmds/mask already on the device, results stay there.  Async on
 * stream (cudaStream_t as void*; NULL = the handle's own stream). */
int umrs_world_step_device(umrs_world* world, const float* cmds_dev,
                           const uint8_t* cmd_mask_dev, void* stream);
/* Device-resident outputs of the last step (owned by the handle). */
int umrs_world_device_buffers(umrs_world* world, float** ranges_dev,
                              int32_t** hit_idx_dev, float** robot_pose_dev);
/* Fan parameters the last step used: [W][R][4] {loc_x, loc_y, angle_min, angle_max}
 * in the world frame (laser origin simulator.cpp:398-400, angle_min + theta
 * :408-409), so a checker can hand the oracle the identical fan.  Host buffer. */
int umrs_world_get_fan_params(umrs_world* world, float* out);
int umrs_world_sync(umrs_world* world);
int umrs_world_step_count(const umrs_world* world, uint64_t* steps);

/* ---- checkpoint / resume -------------------------------------------------------
 * The reference can only inject a pose (/initialpose -> SetPose, simulator.cpp:191-196).
 * A snapshot here holds everything a BIT-REPRODUCIBLE resume needs: every robot's pose,
 * velocity, last command, command age and the diff-drive model's double-precision
 * velocities (diff_drive_model.h:28-31), every human's pose, (possibly swapped) start/goal
 * and velocities (human_object.h:40-52), the static objects' poses, and the step counter
 * that keys the Philox noise streams.  The blob is only valid for a world of the same
 * shape (worlds, robots, humans, objects) and seed; otherwise UMRS_ERR_INVALID.
 * Host buffers; both calls wait for the world's pending work first. */
int umrs_world_state_size(const umrs_world* world, uint64_t* bytes);
int umrs_world_get_state(umrs_world* world, void* blob, uint64_t capacity);
int umrs_world_set_state(umrs_world* world, const void* blob, uint64_t bytes);

/* Device-side conditions no status code could report at launch time, sticky until read:
 * UMRS_ERR_UNSUPPORTED when a fan of the uniform-grid walk touched more grid rows than the
 * plan sized shared memory for (the farthest rows were then NOT scanned).  The synchronous
 * host-buffer entry points check this themselves; call these after device/async steps. */
int umrs_map_check(umrs_map* map);
int umrs_world_check(umrs_world* world);

/* Work accounting for the roofline: candidate ray-segment tests executed and
 * candidate segments pre-processed by the scan kernel in the last step (counted
 * on the device when enabled; costs a few atomics per CTA). */
int umrs_world_enable_counters(umrs_world* world, int32_t on);
int umrs_world_read_counters(umrs_world* world, uint64_t* ray_seg_tests,
                             uint64_t* segs_prepared, uint64_t* chunk_seg_pairs);

/* All eight counters of the last step: {ray_seg_tests, segs_prepared,
 * chunk_seg_pairs, passes (summed over fans), tie sweeps, spilled-candidate
 * list entries, fans that swept every slot, source segments streamed by the
 * gather (low 40 bits) | source rows visited << 40}. */
int umrs_world_read_counters_ex(umrs_world* world, uint64_t out8[8]);
/* Development builds only (-DUMRS_PHASE_TIMING): clock cycles of every fan CTA's critical
 * path per phase, summed over the fans of the last step: {world phase A, setup, gather,
 * records, buckets, ray loop, finalise, spare}.  All zero in the shipped build. */
int umrs_world_read_phase_cycles(umrs_world* world, uint64_t out8[8]);

/* K3: per-shard statistics of the last step's ranges (count of hits, min, max,
 * sum), reduced on the device; 4 doubles out.  Host buffer. */
int umrs_world_stats(umrs_world* world, double* out4);

/* ---- config ingestion without Lua -------------------------------
 * Loads the flat subset of Lua used by the reference's config files (the .lua files under config/:
 * assignments, Vector2/Vector3/DegToRad helpers, tables, require) in order into one
 * global table -- config_reader::ConfigReader({env_config, robot_config}),
 * simulator.cpp:106 -- and answers dotted lookups like lua_script.h:82-119
 * ("laser_loc.x").  `root_dir` (nullable) is where require("config.sim_config") and
 * relative file names resolve (the reference runs from its repository root).
 * Constructs outside the subset fail the load with UMRS_ERR_UNSUPPORTED; an
 * unreadable file is UMRS_ERR_IO; umrs_config_last_error() has file:line text. */
typedef struct umrs_config umrs_config;
int umrs_config_load(const char* const* files, int32_t n_files, const char* root_dir,
                     umrs_config** out);
int umrs_config_destroy(umrs_config* cfg);
const char* umrs_config_last_error(void);
int umrs_config_has(const umrs_config* cfg, const char* key); /* 1 / 0 */
/* Typed getters (CONFIG_FLOAT/DOUBLE/INT/BOOL/STRING): UMRS_ERR_INVALID when the
 * key is missing or of another type; *out is then left untouched. */
int umrs_config_get_double(const umrs_config* cfg, const char* key, double* out);
int umrs_config_get_float(const umrs_config* cfg, const char* key, float* out);
int umrs_config_get_int(const umrs_config* cfg, const char* key, int32_t* out);
int umrs_config_get_bool(const umrs_config* cfg, const char* key, int32_t* out);
int umrs_config_get_string(const umrs_config* cfg, const char* key, char* buf, int32_t cap);
/* CONFIG_STRINGLIST / CONFIG_VECTOR3FLIST (robot_types, start_poses, human_config_list:
 * simulator.cpp:82,95-97). */
int umrs_config_list_length(const umrs_config* cfg, const char* key, int32_t* n);
int umrs_config_get_list_string(const umrs_config* cfg, const char* key, int32_t index,
                                char* buf, int32_t cap);
int umrs_config_get_list_vector3(const umrs_config* cfg, const char* key, int32_t index,
                                 float out3[3]);
/* The parameter PODs filled from the keys the reference reads (cited in
 * umrs_config.cpp).  A missing key becomes 0 like the reference's CONFIG_* default
 * and is counted in *n_missing (nullable).  footprint_enabled is left 0. */
int umrs_config_laser_params(const umrs_config* cfg, umrs_laser_params* out, int32_t* n_missing);
int umrs_config_robot_params(const umrs_config* cfg, const char* robot_type,
                             umrs_robot_params* out, int32_t* n_missing);
int umrs_config_human_params(const umrs_config* cfg, umrs_human_params* out, int32_t* n_missing);

/* ---- tuning hooks (experiments and tests; process-wide; the library never reads the
 * environment) -------------------------------------------------------------------------
 *   "cta_shape"    -1 auto | 0..6: force one fan-CTA shape (256x4, 192x5, 192x6, 128x8, 96x10, 512x2, 320x3);
 *                  a shape the plan's shared memory cannot seat is ignored
 *   "cap"          candidate slots per pass, 32..4064 (0 = default 1536)
 *   "grid_from"    map size from which the uniform-grid walk is picked (0 = default 4096)
 *   "step_slices"  world slices of a host-buffer tick, 1..15 (0 = auto)
 *   "tma_stage"    -1 auto | 0 | 1: stage the fan's source rows into shared memory with bulk copies
 *                  (cp.async.bulk + mbarrier) instead of per-thread loads; read when a plan is made
 *                  (world creation / scan call).  Only variant builds (-DUMRS_TMA_STAGE) contain the
 *                  code: measured slower on every plan (profiles/r2_tma_ab.txt), no effect otherwise
 *   "split_step"   -1 auto | 0 | 1: run a world tick as two launches (entity kernel, then the fan
 *                  kernel on the stepped poses) instead of the fused one; auto splits large batches
 *   "render_cta_fans"  reference render: batches of up to this many fans run one 4-warp CTA per fan
 *                  (the speculative form), larger ones a warp per fan (0 = default)
 * UMRS_ERR_INVALID for an unknown key or a value out of range. */
int umrs_tuning_set(const char* key, int32_t value);
int umrs_tuning_get(const char* key, int32_t* value);

/* Philox4x32-10 as used for the noise (host restatement for tests):
 * counter {c0,c1,c2,c3}, key {k0,k1} -> 4 words. */
void umrs_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif /* UMRS_H_ */
