// K1: shared-memory scan kernel (+ fused world step), sm_100a.
//
// One CTA = one ray fan (one robot's LaserScan of one world).  Replaces
//   VectorMap::GetPredictedScan   vector_map.cpp:420-478   (per-ray fill)
//   VectorMap::SceneRender        vector_map.cpp:217-258   (NOT reproduced: the
//       O(L^2) occlusion trimming is replaced by a per-ray minimum over all
//       candidates -- the same visible surface; SURVEY.md 8c "tier B")
//   VectorMap::GetSceneLines      vector_map.cpp:192-215   (AABB cull, order)
// and, in the world-step form, Simulator::update + publishLaser
//   (simulator.cpp:474-509, 393-426) with the model Step()s.
//
// Phases of a CTA:
//   A  (world step only) integrate every entity of the world, build the
//      world's object lines in shared memory
//   B  candidate pre-pass, one thread per candidate segment: AABB cull,
//      translate, two double atan2, unit normal -> {a0,a1,nx,ny | c | ray-index
//      span} records in shared memory
//   C  ray loop: a warp owns 32 consecutive rays (lane = ray); lanes first act
//      as 32 parallel cull testers over the span table (ballot), then every
//      surviving candidate is broadcast from shared memory and tested by all
//      32 rays with the reference's inclusive angular predicate
#include "umrs_device.cuh"
#include "umrs_internal.h"

namespace umrs {

// ---- shared-memory carve-up --------------------------------------------------
struct SmemView {
  float4* rec4;   // [n_cand] {a0, a1, nx, ny}
  float* recc;    // [n_cand] c
  int2* span;     // [n_cand] conservative ray-index interval {lo, hi}; lo > hi = dead
  float4* obj;    // [n_obj_lines] world object lines (world step only)
  float* ent;     // [n_ent][5] {x, y, theta, sin, cos} (world step only)
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

size_t k1_smem_bytes(int n_cand, int n_obj_lines, int n_ent) {
  size_t b = 0;
  b += align16(sizeof(float4) * (size_t)n_cand);
  b += align16(sizeof(float) * (size_t)n_cand);
  b += align16(sizeof(int2) * (size_t)n_cand);
  b += align16(sizeof(float4) * (size_t)n_obj_lines);
  b += align16(sizeof(float) * 5 * (size_t)n_ent);
  return b;
}

__device__ inline SmemView carve(unsigned char* base, int n_cand, int n_obj_lines) {
  SmemView v;
  size_t o = 0;
  v.rec4 = reinterpret_cast<float4*>(base + o); o += align16(sizeof(float4) * (size_t)n_cand);
  v.recc = reinterpret_cast<float*>(base + o);  o += align16(sizeof(float) * (size_t)n_cand);
  v.span = reinterpret_cast<int2*>(base + o);   o += align16(sizeof(int2) * (size_t)n_cand);
  v.obj = reinterpret_cast<float4*>(base + o);  o += align16(sizeof(float4) * (size_t)n_obj_lines);
  v.ent = reinterpret_cast<float*>(base + o);
  return v;
}

struct Fan {
  float lx, ly;        // laser origin
  float amin, amax;    // world-frame fan limits (angle_min + theta, ...)
  int num_rays;
  float range_max;
};

// ---- phase B: candidate pre-pass ----------------------------------------------
// Candidate j < n_map is map segment j; j >= n_map is object line j - n_map
// (GetSceneLines order).  Object lines [skip_lo, skip_hi) are the fan's own body.
__device__ void prepass(const SmemView& sm, const Fan& fan, const float4* __restrict__ map_segs,
                        int n_map, const float4* obj_segs, int n_obj, int skip_lo,
                        int skip_hi, unsigned long long* counters) {
  const float da = (fan.amax - fan.amin) / (float)fan.num_rays;  // vector_map.cpp:465
  // Conservative ray-index estimate; the exact decision is the float predicate
  // in the ray loop.  If the fan is degenerate or wider than one turn every
  // candidate spans all rays (plain brute force, still exact).
  const bool cull_ok = (da > 1e-7f) && ((fan.amax - fan.amin) <= 6.2832f);
  const float inv_da = cull_ok ? 1.0f / da : 0.f;
  int n_valid = 0;
  for (int j = threadIdx.x; j < n_map + n_obj; j += blockDim.x) {
    float4 s;
    bool keep = true;
    if (j < n_map) {
      s = __ldg(map_segs + j);
    } else {
      const int o = j - n_map;
      s = obj_segs[o];
      keep = !(o >= skip_lo && o < skip_hi);
    }
    keep = keep && aabb_keep(s, fan.lx, fan.ly, fan.range_max);
    SegCast k;
    k.valid = false;
    if (keep) k = make_cast(s, fan.lx, fan.ly);
    int2 sp = make_int2(1, 0);
    if (k.valid) {
      ++n_valid;
      if (cull_ok) {
        float u0 = k.a0 - fan.amin;
        u0 -= k2PiF * floorf(u0 / k2PiF);              // [0, 2pi)
        sp.x = (int)floorf(u0 * inv_da) - 2;
        sp.y = (int)ceilf((u0 + k.arc) * inv_da) + 2;
      } else {
        sp = make_int2(0, 0x3fffffff);
      }
    }
    sm.rec4[j] = make_float4(k.a0, k.a1, k.nx, k.ny);
    sm.recc[j] = k.c;
    sm.span[j] = sp;
  }
  if (counters != nullptr) {
    // one atomic per warp
    for (int o = 16; o > 0; o >>= 1) n_valid += __shfl_down_sync(0xffffffffu, n_valid, o);
    if ((threadIdx.x & 31) == 0 && n_valid) atomicAdd(counters + 1, (unsigned long long)n_valid);
  }
}

// ---- phase C: ray loop ----------------------------------------------------------
// Returns nothing; writes ranges/idx for the whole fan.  When `post` is set the
// publishLaser post-process (simulator.cpp:412-418) is applied with Philox noise.
struct PostProcess {
  bool enabled;
  float stddev;
  uint32_t key0, key1, step_lo, step_hi, fan_global;
};

__device__ void ray_loop(const SmemView& sm, const Fan& fan, int n_cand, float* __restrict__ ranges,
                         int32_t* __restrict__ hit_idx, const PostProcess& post,
                         unsigned long long* counters) {
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int n_warps = blockDim.x >> 5;
  const float span_f = fan.amax - fan.amin;
  const float da = span_f / (float)fan.num_rays;
  const bool cull_ok = (da > 1e-7f) && (span_f <= 6.2832f);
  // rays per full turn, bracketed for the wrapped image of a candidate's span
  const float period = cull_ok ? k2PiF / da : 0.f;
  const int p_hi = (int)ceilf(period) + 1, p_lo = (int)floorf(period) - 1;
  const int n_chunks = (fan.num_rays + 31) >> 5;
  unsigned long long n_tests = 0, n_pairs = 0;

  for (int chunk = warp; chunk < n_chunks; chunk += n_warps) {
    const int i = (chunk << 5) + lane;
    const bool live = i < fan.num_rays;
    const int cs = chunk << 5, ce = cs + 31;
    float a = __int_as_float(0x7fc00000);  // NaN: covers nothing
    float rx = 0.f, ry = 0.f;
    if (live) {
      a = ray_angle(fan.amin, da, i);                    // vector_map.cpp:467
      double sd, cd;
      sincos((double)a, &sd, &cd);                       // :472 double cos/sin
      rx = (float)cd;
      ry = (float)sd;
    }
    float best = __int_as_float(0x7f800000);  // +inf
    int best_j = -1;
    for (int base = 0; base < n_cand; base += 32) {
      const int j = base + lane;
      bool ov = false;
      if (j < n_cand) {
        const int2 sp = sm.span[j];
        ov = (sp.x <= ce && sp.y >= cs) || (sp.x - p_hi <= ce && sp.y - p_lo >= cs);
        ov = ov && (sp.x <= sp.y);
      }
      unsigned m = __ballot_sync(0xffffffffu, ov);
      n_pairs += __popc(m);
      while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        const int jj = base + b;
        const float4 r4 = sm.rec4[jj];                   // broadcast LDS.128
        if (cast_covers(r4.x, r4.y, a)) {                // :469-470
          const float den = r4.z * rx + r4.w * ry;       // n.dot(r)
          const float r = sm.recc[jj] / den;             // :473
          if (r < best) {                                // first strictly smaller wins
            best = r;
            best_j = jj;
          }
        }
      }
    }
    if (counters != nullptr) n_tests += (unsigned long long)0;  // filled below
    if (live) {
      float out = best_j >= 0 ? best : fan.range_max;    // :433 fill value
      if (post.enabled) {
        // simulator.cpp:412-418; range_max - 0.1 is a double expression
        if ((double)out > (double)fan.range_max - 0.1) {
          out = 0.f;
        } else {
          const uint4 w = philox4x32(
              make_uint4((uint32_t)(i >> 2), post.fan_global, post.step_lo,
                         post.step_hi ^ kStreamLaser),
              make_uint2(post.key0, post.key1));
          const float2 g = (i & 2) ? box_muller(w.z, w.w) : box_muller(w.x, w.y);
          const float gi = (i & 1) ? g.y : g.x;
          out = fmaxf(0.f, out + post.stddev * gi);
        }
      }
      ranges[i] = out;
      if (hit_idx != nullptr) hit_idx[i] = best_j;
    }
  }
  if (counters != nullptr) {
    // chunk-candidate pairs are warp-uniform; tests = pairs * 32 lanes
    if (lane == 0 && n_pairs) {
      atomicAdd(counters + 2, n_pairs);
      atomicAdd(counters + 0, n_pairs * 32ull);
    }
    (void)n_tests;
  }
}

// ---- scan-only kernel (umrs_predicted_scan / umrs_scan_batch) -------------------
__global__ void __launch_bounds__(kScanThreads)
k1_scan_batch(const __grid_constant__ ScanBatchParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_cand = p.c.n_map + p.n_obj;
  const SmemView sm = carve(smem_raw, n_cand, 0);
  for (int f = blockIdx.x; f < p.n_fans; f += gridDim.x) {
    const float4 fp = __ldg(p.fans + f);
    Fan fan;
    fan.lx = fp.x; fan.ly = fp.y; fan.amin = fp.z; fan.amax = fp.w;
    fan.num_rays = p.num_rays;
    fan.range_max = p.range_max;
    __syncthreads();  // previous fan's ray loop is done with the records
    prepass(sm, fan, p.c.map_segs, p.c.n_map, p.obj_segs, p.n_obj, 0, 0, p.c.counters);
    __syncthreads();
    PostProcess post;
    post.enabled = false;
    post.stddev = 0.f; post.key0 = post.key1 = post.step_lo = post.step_hi = post.fan_global = 0;
    const size_t off = (size_t)f * (size_t)p.num_rays;
    ray_loop(sm, fan, n_cand, p.c.ranges + off,
             p.c.hit_idx ? p.c.hit_idx + off : nullptr, post, p.c.counters);
  }
}

// ---- phase A: kinematic steps (per entity, one thread each) ------------------
__device__ __forceinline__ void boundf(float lo, float hi, float& x) {  // math_util.h:188-192
  if (x > hi) x = hi;
  if (x < lo) x = lo;
}
__device__ __forceinline__ int signf_(float v) { return (0.0f < v) - (v < 0.0f); }

// DriveCallback of each model: ackermann_model.cpp:41-50,
// diff_drive_model.cpp:170-197, omnidirectional_model.cpp:49-58.
__device__ void robot_command(RobotState& s, const umrs_robot_params& p, const float* cmd) {
  if (p.model_type == UMRS_ACKERMANN_DRIVE) {
    if (!isfinite(cmd[0]) || !isfinite(cmd[1])) return;
    s.cmd[0] = cmd[0]; s.cmd[1] = cmd[1]; s.cmd[2] = 0.f;
    s.cmd_age = 0;
  } else if (p.model_type == UMRS_DIFF_DRIVE) {
    double x = (double)cmd[0], z = (double)cmd[1];
    if (p.invert_linear_vel_cmds) x *= -1;
    if (p.invert_angular_vel_cmds) z *= -1;
    if ((double)p.max_linear_vel != 0.0) {
      if (fabs(x) > (double)p.max_linear_vel)
        x = (x > 0) ? (double)p.max_linear_vel : (double)-p.max_linear_vel;
    }
    if ((double)p.max_angular != 0.0) {
      if (fabs(z) > (double)p.max_angular)
        z = (z > 0) ? (double)p.max_angular : (double)-p.max_angular;
    }
    s.cmd[0] = (float)x; s.cmd[1] = (float)z; s.cmd[2] = 0.f;
    s.cmd_age = 0;
  } else {
    s.cmd[0] = cmd[0]; s.cmd[1] = cmd[1]; s.cmd[2] = cmd[2];
    s.cmd_age = 0;
  }
}

__device__ void accel_limit(double& v, float target, float pos_lim, float neg_lim) {
  // diff_drive_model.cpp:91-130 (limits per tick)
  if (v < (double)target) {
    if ((double)pos_lim == 0.0 || (double)target - v < (double)pos_lim) v = (double)target;
    else v += (double)pos_lim;
  } else if (v > (double)target) {
    if ((double)neg_lim == 0.0 || v - (double)target < (double)neg_lim) v = (double)target;
    else v -= (double)neg_lim;
  }
}

__device__ void robot_step(RobotState& s, const umrs_robot_params& p, double dt, float n1, float n2) {
  const bool stale = (double)s.cmd_age * dt > 0.1;  // kMaxCommandAge in sim time
  if (p.model_type == UMRS_ACKERMANN_DRIVE) {
    // ackermann_model.cpp:52-96
    if (stale) s.cmd[0] = 0.f;
    const float vel = s.vel[0];
    const float kEpsilonCurvature = (float)(1.0 / 1E3);
    float desired_vel = s.cmd[0];
    boundf(-p.ak_max_speed, p.ak_max_speed, desired_vel);
    const float max_curvature = (float)(1.0 / (double)p.ak_min_turn_radius);
    float desired_curvature = s.cmd[1];
    boundf(-max_curvature, max_curvature, desired_curvature);
    const bool linear_motion = fabsf(desired_curvature) < kEpsilonCurvature;
    const float dv_max = (float)(dt * (double)p.ak_max_accel);
    float bounded_dv = desired_vel - vel;
    boundf(-dv_max, dv_max, bounded_dv);
    s.vel[0] = vel + bounded_dv;
    const float dist = (float)((double)s.vel[0] * dt);
    float dx = 0.f, dy = 0.f, dtheta = 0.f;
    if (linear_motion) {
      dx = dist;
      dtheta = (float)(dt * (double)p.ak_angular_error_bias);
    } else {
      const float r = (float)(1.0 / (double)desired_curvature);
      const float arc = dist * desired_curvature;
      dtheta = (float)((double)arc + ((double)n1 * dt) * (double)p.ak_angular_error_bias +
                       (double)((n2 * p.ak_angular_error_rate) * fabsf(arc)));
      dx = r * sinf(dtheta);
      dy = (float)((double)r * (1.0 - (double)cosf(dtheta)));
    }
    const float sn = sinf(s.pose[2]), cs = cosf(s.pose[2]);
    s.pose[0] = s.pose[0] + (cs * dx - sn * dy);
    s.pose[1] = s.pose[1] + (sn * dx + cs * dy);
    s.pose[2] = angle_mod_f(s.pose[2] + dtheta);
  } else if (p.model_type == UMRS_DIFF_DRIVE) {
    // diff_drive_model.cpp:83-148
    if (stale) { s.cmd[0] = 0.f; s.cmd[1] = 0.f; }
    accel_limit(s.linear_vel, s.cmd[0], p.linear_pos_accel_limit, p.linear_neg_accel_limit);
    accel_limit(s.angular_vel, s.cmd[1], p.angular_pos_accel_limit, p.angular_neg_accel_limit);
    const float acc_times_dt = (float)(s.linear_vel - (double)s.vel[0]);
    const float dx = (float)((double)s.vel[0] * dt + (0.5 * (double)acc_times_dt) * dt);
    s.vel[0] = (float)s.linear_vel;
    s.vel[2] = (float)s.angular_vel;
    s.pose[0] = s.pose[0] + cosf(s.pose[2]) * dx;
    s.pose[1] = s.pose[1] + sinf(s.pose[2]) * dx;
    s.pose[2] = (float)angle_mod_d((double)s.pose[2] + (double)s.vel[2] * dt);
  } else {
    // omnidirectional_model.cpp:82-123
    if (stale) { s.cmd[0] = 0.f; s.cmd[1] = 0.f; s.cmd[2] = 0.f; }
    float dvx = s.cmd[0], dvy = s.cmd[1];
    float z = dvx * dvx + dvy * dvy;
    if (sqrtf(z) > p.co_max_speed) {
      if (z > 0.f) { const float n = sqrtf(z); dvx = dvx / n; dvy = dvy / n; }
      dvx = dvx * p.co_max_speed;
      dvy = dvy * p.co_max_speed;
    }
    float ax = dvx - s.vel[0], ay = dvy - s.vel[1];
    const float max_accel = (float)((double)p.co_max_accel * dt);
    z = ax * ax + ay * ay;
    if (sqrtf(z) > max_accel) {
      if (z > 0.f) { const float n = sqrtf(z); ax = ax / n; ay = ay / n; }
      ax = ax * max_accel;
      ay = ay * max_accel;
    }
    s.vel[0] = s.vel[0] + ax;
    s.vel[1] = s.vel[1] + ay;
    float desired_ang = s.cmd[2];
    if (fabsf(desired_ang) > p.co_max_angle_vel)
      desired_ang = (float)signf_(desired_ang) * p.co_max_angle_vel;
    const float max_angle_accel = (float)((double)p.co_max_angle_accel * dt);
    float delta_ang = desired_ang - s.vel[2];
    if (fabsf(desired_ang) > max_angle_accel)  // :115 quirk kept
      delta_ang = (float)signf_(delta_ang) * max_angle_accel;
    s.vel[2] = s.vel[2] + delta_ang;
    const float sn = sinf(s.pose[2]), cs = cosf(s.pose[2]);
    const float rx = cs * s.vel[0] - sn * s.vel[1];
    const float ry = sn * s.vel[0] + cs * s.vel[1];
    const float dtf = (float)dt;
    s.pose[0] = s.pose[0] + rx * dtf;
    s.pose[1] = s.pose[1] + ry * dtf;
    s.pose[2] = (float)angle_mod_d((double)s.pose[2] + (double)s.vel[2] * dt);
  }
  if (s.cmd_age < kAgeNever) s.cmd_age += 1;
}

// HumanObject::Step + CheckReachGoal, human_object.cpp:147-161,178-192.
__device__ void human_step(HumanState& s, const umrs_human_params& p, double dt) {
  const double avg_speed = (double)p.avg_speed, max_speed = (double)p.max_speed;
  float gx = s.goal[0] - s.pose[0], gy = s.goal[1] - s.pose[1];
  float z = gx * gx + gy * gy;
  if (z > 0.f) { const float n = sqrtf(z); gx = gx / n; gy = gy / n; }
  float vx = gx * (float)avg_speed, vy = gy * (float)avg_speed;
  z = vx * vx + vy * vy;
  if ((double)sqrtf(z) > max_speed) {
    if (z > 0.f) { const float n = sqrtf(z); vx = vx / n; vy = vy / n; }
    vx = vx * (float)max_speed;
    vy = vy * (float)max_speed;
  }
  s.trans_vel[0] = vx;
  s.trans_vel[1] = vy;
  s.pose[2] = (float)((double)s.pose[2] + (double)s.rot_vel * dt);
  s.pose[0] = s.pose[0] + vx * (float)dt;
  s.pose[1] = s.pose[1] + vy * (float)dt;
  const float ex = s.pose[0] - s.goal[0], ey = s.pose[1] - s.goal[1];
  if ((double)sqrtf(ex * ex + ey * ey) < (double)p.reach_goal_threshold) {
    if (p.mode == 0) {
      s.trans_vel[0] = 0.f; s.trans_vel[1] = 0.f; s.rot_vel = 0.f;
    } else if (p.mode == 1) {
      for (int k = 0; k < 3; ++k) { const float t = s.goal[k]; s.goal[k] = s.start[k]; s.start[k] = t; }
    }
  }
}

// ---- fused world-step kernel -------------------------------------------------
// grid = W * R fans; CTA (w, r).  Every CTA of world w integrates ALL entities of
// that world (R + H tiny steps, redundantly across the R CTAs) from the
// read-only "in" state so that no inter-CTA exchange is needed; the r == 0 CTA
// alone writes the "out" state (ping-pong buffers).  Order as in
// Simulator::Run (simulator.cpp:531-541): update() first, scan from the
// post-step pose against the post-step object lines.
__global__ void __launch_bounds__(kScanThreads)
k1_world_step(const __grid_constant__ WorldStepParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_cand = p.c.n_map + p.L_obj;
  const int n_ent = p.O + p.H + p.R;
  const SmemView sm = carve(smem_raw, n_cand, p.L_obj);
  const int w = blockIdx.x / p.R;
  const int r = blockIdx.x - w * p.R;
  const bool owner = (r == 0);
  const double dt = (double)p.dt;  // `const double& dt` bound to float CONFIG_DT
  const int t = threadIdx.x;

  // ---- phase A: entity steps -> sm.ent[e] = {x, y, theta, sin, cos} ----------
  for (int e = t; e < n_ent; e += blockDim.x) {
    float px, py, pth;
    if (e < p.O) {
      const float* sp = p.static_pose + ((size_t)w * p.O + e) * 3;
      px = sp[0]; py = sp[1]; pth = sp[2];
    } else if (e < p.O + p.H) {
      const int h = e - p.O;
      HumanState s = p.hs_in[(size_t)w * p.H + h];
      human_step(s, p.humans[h], dt);
      if (owner) p.hs_out[(size_t)w * p.H + h] = s;
      px = s.pose[0]; py = s.pose[1]; pth = s.pose[2];
    } else {
      const int rr = e - p.O - p.H;
      const size_t gi = (size_t)w * p.R + rr;
      RobotState s = p.rs_in[gi];
      const umrs_robot_params& rp = p.robots[rr];
      if (p.cmds != nullptr && (p.cmd_mask == nullptr || p.cmd_mask[gi] != 0)) {
        const float cmd[3] = {p.cmds[gi * 3 + 0], p.cmds[gi * 3 + 1], p.cmds[gi * 3 + 2]};
        robot_command(s, rp, cmd);
      }
      float n1 = 0.f, n2 = 0.f;
      if (p.motion_noise && rp.model_type == UMRS_ACKERMANN_DRIVE) {
        const uint4 wd = philox4x32(
            make_uint4(0u, (uint32_t)(p.world_base + w) * (uint32_t)p.R + (uint32_t)rr,
                       p.step_lo, p.step_hi ^ kStreamMotion),
            make_uint2(p.key0, p.key1));
        const float2 g = box_muller(wd.x, wd.y);
        n1 = g.x; n2 = g.y;
      }
      robot_step(s, rp, dt, n1, n2);
      if (owner) {
        p.rs_out[gi] = s;
        if (p.robot_pose_out != nullptr) {
          p.robot_pose_out[gi * 3 + 0] = s.pose[0];
          p.robot_pose_out[gi * 3 + 1] = s.pose[1];
          p.robot_pose_out[gi * 3 + 2] = s.pose[2];
        }
      }
      px = s.pose[0]; py = s.pose[1]; pth = s.pose[2];
    }
    // HumanObject::Transform rotation, human_object.cpp:132: R(AngleMod(angle))
    const float am = angle_mod_f(pth);
    float* en = sm.ent + 5 * e;
    en[0] = px; en[1] = py; en[2] = pth;
    en[3] = sinf(am); en[4] = cosf(am);
  }
  __syncthreads();

  // ---- object lines: R * template + T (human_object.cpp:135-138) --------------
  for (int l = t; l < p.L_obj; l += blockDim.x) {
    const int e = p.line_ent[l];
    const float4 tm = __ldg(p.tmpl + l);
    const float* en = sm.ent + 5 * e;
    const float sn = en[3], cs = en[4];
    float4 o;
    o.x = (cs * tm.x - sn * tm.y) + en[0];
    o.y = (sn * tm.x + cs * tm.y) + en[1];
    o.z = (cs * tm.z - sn * tm.w) + en[0];
    o.w = (sn * tm.z + cs * tm.w) + en[1];
    sm.obj[l] = o;
    if (owner && p.obj_lines_out != nullptr) p.obj_lines_out[(size_t)w * p.L_obj + l] = o;
  }

  // ---- fan of robot r (simulator.cpp:398-411) ----------------------------------
  const umrs_robot_params& me = p.robots[r];
  const float* en = sm.ent + 5 * (p.O + p.H + r);
  Fan fan;
  {
    // laserLoc = t + Rotation2Df(angle) * (laser_x, laser_y): sin/cos of the raw
    // pose angle (float overloads inside Eigen)
    const float sn = sinf(en[2]), cs = cosf(en[2]);
    fan.lx = en[0] + (cs * me.laser.loc_x - sn * me.laser.loc_y);
    fan.ly = en[1] + (sn * me.laser.loc_x + cs * me.laser.loc_y);
    fan.amin = me.laser.angle_min + en[2];
    fan.amax = me.laser.angle_max + en[2];
    fan.num_rays = p.num_rays[r];
    fan.range_max = me.laser.range_max;
  }
  if (t == 0 && p.fan_params != nullptr)
    p.fan_params[blockIdx.x] = make_float4(fan.lx, fan.ly, fan.amin, fan.amax);
  __syncthreads();

  const int skip_lo = p.ent_off[p.O + p.H + r], skip_hi = p.ent_off[p.O + p.H + r + 1];
  prepass(sm, fan, p.c.map_segs, p.c.n_map, sm.obj, p.L_obj, skip_lo, skip_hi, p.c.counters);
  __syncthreads();

  PostProcess post;
  post.enabled = p.laser_noise != 0;
  post.stddev = me.laser.noise_stddev;
  post.key0 = p.key0; post.key1 = p.key1;
  post.step_lo = p.step_lo; post.step_hi = p.step_hi;
  post.fan_global = (uint32_t)(p.world_base + w) * (uint32_t)p.R + (uint32_t)r;
  const size_t off = (size_t)w * (size_t)p.rays_per_world + (size_t)p.ray_off[r];
  ray_loop(sm, fan, n_cand, p.c.ranges + off,
           p.c.hit_idx ? p.c.hit_idx + off : nullptr, post, p.c.counters);
}

// ---- K3: statistics of a ranges buffer -------------------------------------------
// out4 = {count of finite positive ranges, min, max, sum}; one atomic set per CTA.
__global__ void k3_stats(const float* __restrict__ ranges, size_t n, double* out4) {
  double cnt = 0.0, sum = 0.0;
  float mn = __int_as_float(0x7f800000), mx = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    const float v = ranges[i];
    if (v > 0.f && v < __int_as_float(0x7f800000)) {
      cnt += 1.0; sum += (double)v;
      mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    sum += __shfl_down_sync(0xffffffffu, sum, o);
    mn = fminf(mn, __shfl_down_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(out4 + 0, cnt);
    atomicAdd(out4 + 3, sum);
    // positive floats order like their bit patterns
    atomicMin(reinterpret_cast<unsigned long long*>(out4 + 1),
              (unsigned long long)__float_as_uint(mn));
    atomicMax(reinterpret_cast<unsigned long long*>(out4 + 2),
              (unsigned long long)__float_as_uint(mx));
  }
}

// ---- launchers ----------------------------------------------------------------
static size_t g_k1_smem_configured = 0;

cudaError_t k1_configure(size_t smem_bytes) {
  if (smem_bytes <= g_k1_smem_configured) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(k1_scan_batch, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k1_world_step, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)smem_bytes);
  if (e != cudaSuccess) return e;
  g_k1_smem_configured = smem_bytes;
  return cudaSuccess;
}

cudaError_t launch_scan_batch(const ScanBatchParams& p, cudaStream_t stream) {
  if (p.n_fans <= 0) return cudaSuccess;
  const size_t smem = k1_smem_bytes(p.c.n_map + p.n_obj, 0, 0);
  cudaError_t e = k1_configure(smem);
  if (e != cudaSuccess) return e;
  k1_scan_batch<<<p.n_fans, kScanThreads, smem, stream>>>(p);
  note_launch();
  return cudaGetLastError();
}

cudaError_t launch_world_step(const WorldStepParams& p, cudaStream_t stream) {
  const int fans = p.W * p.R;
  if (fans <= 0) return cudaSuccess;
  const size_t smem = k1_smem_bytes(p.c.n_map + p.L_obj, p.L_obj, p.O + p.H + p.R);
  cudaError_t e = k1_configure(smem);
  if (e != cudaSuccess) return e;
  k1_world_step<<<fans, kScanThreads, smem, stream>>>(p);
  note_launch();
  return cudaGetLastError();
}

cudaError_t launch_stats(const float* ranges, size_t n, double* out4_dev, cudaStream_t stream) {
  // out4_dev must be pre-initialised by the caller: {0, bits(+inf), 0, 0}
  k3_stats<<<148 * 4, 256, 0, stream>>>(ranges, n, out4_dev);
  note_launch();
  return cudaGetLastError();
}

}  // namespace umrs
