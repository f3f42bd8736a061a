/* TEST INFRASTRUCTURE ONLY -- see umrs_oracle.h.  Plain-C restatement of the
 * reference algorithm; every function cites the reference file:line it
 * follows.  Compile with -ffp-contract=off (oracle/Makefile): float expressions
 * below are written so that each C operation is one IEEE operation in the type
 * the reference's C++ expression has (float unless a double operand promotes
 * it). */
#include "umrs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#define ORC_M_2PI (2.0 * M_PI) /* math_util.h:28-30 */

/* math_util::AngleMod<float>, math_util.h:55-59.  `rint` there is the C
 * ::rint(double) (only <cmath> is in scope), so the product T(M_2PI)*rint(..)
 * and the subtraction are evaluated in double and narrowed by the `-=`. */
static float angle_mod_f(float angle) {
  const float q = angle / (float)ORC_M_2PI;
  const double k = rint((double)q);
  return (float)((double)angle - (double)((float)ORC_M_2PI) * k);
}
/* math_util::AngleMod<double>. */
static double angle_mod_d(double angle) {
  return angle - ORC_M_2PI * rint(angle / ORC_M_2PI);
}

float orc_ray_angle(float angle_min, float angle_max, int32_t num_rays, int32_t i) {
  /* vector_map.cpp:465,467 */
  const float da = (angle_max - angle_min) / (float)num_rays;
  const float t = angle_min + (float)i * da;
  return angle_mod_f(t);
}

/* Per-candidate constants of the ray fill, vector_map.cpp:444-459,471. */
typedef struct {
  float a0, a1;
  int wraps;
  float nx, ny, c;
  int valid;
} orc_cast;

static orc_cast make_cast(const float s[4], float lx, float ly) {
  orc_cast k;
  memset(&k, 0, sizeof(k));
  /* :447-448  translate by -loc */
  float p0x = s[0] - lx, p0y = s[1] - ly;
  float p1x = s[2] - lx, p1y = s[3] - ly;
  /* :449-450  atan2 is the double C function (float arguments promote), result
   * narrowed into the float members a0/a1. */
  float a0 = (float)atan2((double)p0y, (double)p0x);
  float a1 = (float)atan2((double)p1y, (double)p1x);
  /* :451 */
  if (fabs((double)(a0 - a1)) < 0.0001) return k;
  /* :452 */
  const int wraps = fabs((double)(a1 - a0)) > M_PI;
  /* :453-457 */
  if ((wraps && a0 < a1) || (!wraps && a0 > a1)) {
    float t;
    t = a0; a0 = a1; a1 = t;
    t = p0x; p0x = p1x; p1x = t;
    t = p0y; p0y = p1y; p1y = t;
  }
  /* :471 UnitNormal = Perp(Dir()), line2d.h:62-69; normalized() divides each
   * component by sqrt(squaredNorm) when squaredNorm > 0. */
  float dx = p1x - p0x, dy = p1y - p0y;
  const float z = dx * dx + dy * dy;
  if (z > 0.0f) {
    const float len = sqrtf(z);
    dx = dx / len;
    dy = dy / len;
  }
  k.a0 = a0;
  k.a1 = a1;
  k.wraps = wraps;
  k.nx = -dy;
  k.ny = dx;
  /* :473 n.dot(l.line.p0) */
  k.c = k.nx * p0x + k.ny * p0y;
  k.valid = 1;
  return k;
}

/* :469-470 inclusive, wrap-aware */
static int cast_covers(const orc_cast* k, float a) {
  return (!k->wraps && k->a0 <= a && k->a1 >= a) ||
         (k->wraps && (k->a0 <= a || k->a1 >= a));
}

/* :472-473  r = (cos(a), sin(a)) through the double functions, narrowed into
 * Vector2f; range = n.dot(p0) / n.dot(r). */
static float cast_range(const orc_cast* k, float a) {
  const float rx = (float)cos((double)a);
  const float ry = (float)sin((double)a);
  return k->c / (k->nx * rx + k->ny * ry);
}

/* GetSceneLines AABB cull, vector_map.cpp:195-204. */
static int aabb_keep(const float s[4], float lx, float ly, float max_range) {
  const float x_min = lx - max_range, y_min = ly - max_range;
  const float x_max = lx + max_range, y_max = ly + max_range;
  if (s[0] < x_min && s[2] < x_min) return 0;
  if (s[1] < y_min && s[3] < y_min) return 0;
  if (s[0] > x_max && s[2] > x_max) return 0;
  if (s[1] > y_max && s[3] > y_max) return 0;
  return 1;
}

void orc_scan_tierb(const float* map_xyxy, int32_t n_map, const float* obj_xyxy,
                    int32_t n_obj, int32_t skip_lo, int32_t skip_hi, float loc_x,
                    float loc_y, float range_max, float angle_min, float angle_max,
                    int32_t num_rays, float* ranges, int32_t* idx) {
  const int32_t n_all = n_map + n_obj;
  orc_cast* casts = (orc_cast*)malloc(sizeof(orc_cast) * (size_t)(n_all > 0 ? n_all : 1));
  for (int32_t j = 0; j < n_all; ++j) {
    const float* s = j < n_map ? map_xyxy + 4 * j : obj_xyxy + 4 * (j - n_map);
    casts[j].valid = 0;
    if (j >= n_map && (j - n_map) >= skip_lo && (j - n_map) < skip_hi) continue;
    if (!aabb_keep(s, loc_x, loc_y, range_max)) continue;
    casts[j] = make_cast(s, loc_x, loc_y);
  }
  for (int32_t i = 0; i < num_rays; ++i) {
    const float a = orc_ray_angle(angle_min, angle_max, num_rays, i);
    /* Tier B: minimum over all candidates (first strictly smaller wins, so the
     * lowest index wins exact ties).  No clamp to range_max (class D7): a hit
     * beyond range_max inside the AABB is returned as is; only rays nothing
     * covers keep the fill value range_max (vector_map.cpp:433). */
    float best = INFINITY;
    int32_t best_j = -1;
    for (int32_t j = 0; j < n_all; ++j) {
      if (!casts[j].valid || !cast_covers(&casts[j], a)) continue;
      const float r = cast_range(&casts[j], a);
      if (r < best) {
        best = r;
        best_j = j;
      }
    }
    ranges[i] = best_j >= 0 ? best : range_max;
    if (idx) idx[i] = best_j;
  }
  free(casts);
}

float orc_candidate_range(const float seg_xyxy[4], float loc_x, float loc_y,
                          float angle_min, float angle_max, int32_t num_rays,
                          int32_t i) {
  const orc_cast k = make_cast(seg_xyxy, loc_x, loc_y);
  const float a = orc_ray_angle(angle_min, angle_max, num_rays, i);
  if (!k.valid || !cast_covers(&k, a)) return NAN;
  return cast_range(&k, a);
}

void orc_laser_postprocess(float* ranges, int32_t n, float range_max, float stddev,
                           const float* normals) {
  /* simulator.cpp:412-418.  `range_max - 0.1` is float - double -> double. */
  for (int32_t i = 0; i < n; ++i) {
    float r = ranges[i];
    if ((double)r > (double)range_max - 0.1) {
      ranges[i] = 0.0f;
      continue;
    }
    const float g = normals ? normals[i] : 0.0f;
    const float v = r + stddev * g;
    ranges[i] = v > 0.0f ? v : 0.0f; /* max<float>(0.0, v) */
  }
}

void orc_laser_origin(const float pose_xyt[3], float laser_x, float laser_y,
                      float out_xy[2]) {
  /* simulator.cpp:398-400: t + Rotation2Df(angle) * laserRobotLoc; Eigen's
   * Rotation2D takes sin/cos of a float through the float overloads. */
  const float s = sinf(pose_xyt[2]), c = cosf(pose_xyt[2]);
  out_xy[0] = pose_xyt[0] + (c * laser_x - s * laser_y);
  out_xy[1] = pose_xyt[1] + (s * laser_x + c * laser_y);
}

/* ---- map cleanup ---------------------------------------------------------- */
static float crossf(float ax, float ay, float bx, float by) {
  return ax * by - bx * ay; /* geometry.h:56-59 */
}
static float minf2(float a, float b) { return b < a ? b : a; } /* std::min */
static float maxf2(float a, float b) { return a < b ? b : a; } /* std::max */

/* Line<float>::Intersection(p2, p3, &out), line2d.h:139-160; l = {p0,p1}. */
static int seg_intersection(const float l[4], const float m[4], float out[2]) {
  const float p0x = l[0], p0y = l[1], p1x = l[2], p1y = l[3];
  const float p2x = m[0], p2y = m[1], p3x = m[2], p3y = m[3];
  if (minf2(p0x, p1x) > maxf2(p2x, p3x)) return 0;
  if (maxf2(p0x, p1x) < minf2(p2x, p3x)) return 0;
  if (minf2(p0y, p1y) > maxf2(p2y, p3y)) return 0;
  if (maxf2(p0y, p1y) < minf2(p2y, p3y)) return 0;
  const float d1x = p1x - p0x, d1y = p1y - p0y;
  const float d2x = p3x - p2x, d2y = p3y - p2y;
  if ((double)(crossf(d1x, d1y, p3x - p0x, p3y - p0y) *
               crossf(d1x, d1y, p2x - p0x, p2y - p0y)) > 0.0) return 0;
  if ((double)(crossf(d2x, d2y, p1x - p2x, p1y - p2y) *
               crossf(d2x, d2y, p0x - p2x, p0y - p2y)) > 0.0) return 0;
  const float d = crossf(d2x, d2y, -d1x, -d1y);
  if (d == 0.0f) return 0;
  const float tb = crossf(p0x - p2x, p0y - p2y, p0x - p1x, p0y - p1y) / d;
  out[0] = p2x + tb * d2x;
  out[1] = p2y + tb * d2y;
  return 1;
}

int32_t orc_line_intersection(const float a[4], const float b[4], float out[2]) {
  out[0] = 0.0f;
  out[1] = 0.0f;
  return seg_intersection(a, b, out);
}

static void seg_dir(const float l[4], float d[2]) { /* line2d.h:62-64 */
  float dx = l[2] - l[0], dy = l[3] - l[1];
  const float z = dx * dx + dy * dy;
  if (z > 0.0f) {
    const float n = sqrtf(z);
    dx = dx / n;
    dy = dy / n;
  }
  d[0] = dx;
  d[1] = dy;
}
static float seg_length(const float l[4]) { /* line2d.h:54-56: (p0 - p1).norm() */
  const float dx = l[0] - l[2], dy = l[1] - l[3];
  return sqrtf(dx * dx + dy * dy);
}

int32_t orc_map_cleanup(const float* in_xyxy, int32_t n, float* out_xyxy, int32_t cap) {
  /* vector_map.cpp:367-395 */
  const float kShrinkDistance = 1e-4f;
  const float kMinLineLength = 0.05f;
  size_t n_lines = (size_t)n, cap_lines = (size_t)(n > 0 ? 2 * n : 16);
  float* lines = (float*)malloc(sizeof(float) * 4 * cap_lines);
  memcpy(lines, in_xyxy, sizeof(float) * 4 * (size_t)n);
  size_t n_new = 0, cap_new = (size_t)(n > 0 ? n : 16);
  float* new_lines = (float*)malloc(sizeof(float) * 4 * cap_new);
  for (size_t i = 0; i < n_lines; ++i) {
    float l1[4];
    memcpy(l1, lines + 4 * i, sizeof(l1));
    if (seg_length(l1) < kMinLineLength) continue;
    int intersection = 0;
    float p[2];
    for (size_t j = 0; j < n_new; ++j) {
      if (seg_intersection(new_lines + 4 * j, l1, p)) {
        float dir[2];
        seg_dir(l1, dir);
        const float sx = kShrinkDistance * dir[0], sy = kShrinkDistance * dir[1];
        if (n_lines + 2 > cap_lines) {
          cap_lines *= 2;
          lines = (float*)realloc(lines, sizeof(float) * 4 * cap_lines);
        }
        float* a = lines + 4 * n_lines;
        a[0] = l1[0]; a[1] = l1[1]; a[2] = p[0] - sx; a[3] = p[1] - sy;
        a[4] = p[0] + sx; a[5] = p[1] + sy; a[6] = l1[2]; a[7] = l1[3];
        n_lines += 2;
        intersection = 1;
        break;
      }
    }
    if (!intersection) {
      if (n_new + 1 > cap_new) {
        cap_new *= 2;
        new_lines = (float*)realloc(new_lines, sizeof(float) * 4 * cap_new);
      }
      memcpy(new_lines + 4 * n_new, l1, sizeof(l1));
      ++n_new;
    }
  }
  /* ShrinkLine, vector_map.cpp:359-365 */
  for (size_t j = 0; j < n_new; ++j) {
    float* l = new_lines + 4 * j;
    const float len = seg_length(l);
    float dir[2];
    seg_dir(l, dir);
    if ((double)len < 2.0 * (double)kShrinkDistance) continue;
    l[0] = l[0] + kShrinkDistance * dir[0];
    l[1] = l[1] + kShrinkDistance * dir[1];
    l[2] = l[2] - kShrinkDistance * dir[0];
    l[3] = l[3] - kShrinkDistance * dir[1];
  }
  const int32_t out_n = (int32_t)n_new;
  if (out_n <= cap) memcpy(out_xyxy, new_lines, sizeof(float) * 4 * n_new);
  free(lines);
  free(new_lines);
  return out_n;
}

/* ---- kinematic steps -------------------------------------------------------------
 * Type discipline: the model sources include ros/ros.h, whose ros/duration.h
 * includes <math.h>; with libstdc++ (GCC >= 6) that puts the float overloads of
 * sin/cos/fabs/rint in the global namespace, so unqualified calls on float
 * arguments are the float functions there (unlike vector_map.cpp, which only
 * sees <cmath>).  `dt` is `const double&` bound to the float CONFIG_DT
 * (simulator.cpp:479), i.e. (double)dt_f.
 * The wall-clock command watchdog (GetMonotonicTime() > t_last_cmd_ + 0.1) is
 * restated in SIM time: age = ticks since the last accepted command. */
#define ORC_MAX_COMMAND_AGE 0.1
#define ORC_AGE_NEVER (1 << 28)

static void boundf(float lo, float hi, float* x) { /* math_util.h:188-192 */
  if (*x > hi) *x = hi;
  if (*x < lo) *x = lo;
}
static int signf_(float v) { return (0.0f < v) - (v < 0.0f); } /* math_util.h:203-206 */

void orc_robot_init(orc_robot_state* s, const float pose_xyt[3]) {
  memset(s, 0, sizeof(*s));
  memcpy(s->pose, pose_xyt, sizeof(float) * 3);
  s->cmd_age = ORC_AGE_NEVER; /* t_last_cmd_(0): older than any threshold */
}

void orc_robot_command(orc_robot_state* s, const umrs_robot_params* p, const float cmd[3]) {
  if (p->model_type == UMRS_ACKERMANN_DRIVE) {
    /* ackermann_model.cpp:41-50: non-finite commands are ignored. */
    if (!isfinite(cmd[0]) || !isfinite(cmd[1])) return;
    s->cmd[0] = cmd[0];
    s->cmd[1] = cmd[1];
    s->cmd[2] = 0.0f;
    s->cmd_age = 0;
  } else if (p->model_type == UMRS_DIFF_DRIVE) {
    /* diff_drive_model.cpp:170-197 */
    double x = (double)cmd[0], z = (double)cmd[1];
    if (p->invert_linear_vel_cmds) x *= -1;
    if (p->invert_angular_vel_cmds) z *= -1;
    if ((double)p->max_linear_vel != 0.0) {
      if (fabs(x) > (double)p->max_linear_vel)
        x = (x > 0) ? (double)p->max_linear_vel : (double)-p->max_linear_vel;
    }
    if ((double)p->max_angular != 0.0) {
      if (fabs(z) > (double)p->max_angular)
        z = (z > 0) ? (double)p->max_angular : (double)-p->max_angular;
    }
    s->cmd[0] = (float)x; /* target_linear_vel_ (float member) */
    s->cmd[1] = (float)z; /* target_angular_vel_ */
    s->cmd[2] = 0.0f;
    s->cmd_age = 0;
  } else {
    /* omnidirectional_model.cpp:49-58: the non-finite check only prints; the
     * command is stored regardless. */
    s->cmd[0] = cmd[0];
    s->cmd[1] = cmd[1];
    s->cmd[2] = cmd[2];
    s->cmd_age = 0;
  }
}

static void ackermann_step(orc_robot_state* s, const umrs_robot_params* p, double dt,
                           float n1, float n2) {
  /* ackermann_model.cpp:52-96 */
  if ((double)s->cmd_age * dt > ORC_MAX_COMMAND_AGE) s->cmd[0] = 0.0f; /* :55-58 */
  const float vel = s->vel[0];
  const float kEpsilonCurvature = (float)(1.0 / 1E3);
  float desired_vel = s->cmd[0];
  boundf(-p->ak_max_speed, p->ak_max_speed, &desired_vel);
  const float max_curvature = (float)(1.0 / (double)p->ak_min_turn_radius);
  float desired_curvature = s->cmd[1];
  boundf(-max_curvature, max_curvature, &desired_curvature);
  const int linear_motion = fabsf(desired_curvature) < kEpsilonCurvature;
  const float dv_max = (float)(dt * (double)p->ak_max_accel);
  float bounded_dv = desired_vel - vel;
  boundf(-dv_max, dv_max, &bounded_dv);
  s->vel[0] = vel + bounded_dv;
  const float dist = (float)((double)s->vel[0] * dt);
  float dx = 0.0f, dy = 0.0f, dtheta = 0.0f;
  if (linear_motion) {
    dx = dist;
    dtheta = (float)(dt * (double)p->ak_angular_error_bias);
  } else {
    const float r = (float)(1.0 / (double)desired_curvature);
    const float arc = dist * desired_curvature;
    dtheta = (float)((double)arc + ((double)n1 * dt) * (double)p->ak_angular_error_bias +
                     (double)((n2 * p->ak_angular_error_rate) * fabsf(arc)));
    dx = r * sinf(dtheta);
    dy = (float)((double)r * (1.0 - (double)cosf(dtheta)));
  }
  /* :94 Rotation2Df(angle) * d_vector */
  const float sn = sinf(s->pose[2]), cs = cosf(s->pose[2]);
  s->pose[0] = s->pose[0] + (cs * dx - sn * dy);
  s->pose[1] = s->pose[1] + (sn * dx + cs * dy);
  s->pose[2] = angle_mod_f(s->pose[2] + dtheta);
}

static void accel_limit(double* v, float target, float pos_lim, float neg_lim) {
  /* diff_drive_model.cpp:91-130: limits are per TICK (not multiplied by dt). */
  if (*v < (double)target) {
    if ((double)pos_lim == 0.0 || (double)target - *v < (double)pos_lim) *v = (double)target;
    else *v += (double)pos_lim;
  } else if (*v > (double)target) {
    if ((double)neg_lim == 0.0 || *v - (double)target < (double)neg_lim) *v = (double)target;
    else *v -= (double)neg_lim;
  }
}

static void diff_drive_step(orc_robot_state* s, const umrs_robot_params* p, double dt) {
  /* diff_drive_model.cpp:83-148 */
  if ((double)s->cmd_age * dt > ORC_MAX_COMMAND_AGE) { /* :85-88 */
    s->cmd[0] = 0.0f;
    s->cmd[1] = 0.0f;
  }
  accel_limit(&s->linear_vel, s->cmd[0], p->linear_pos_accel_limit, p->linear_neg_accel_limit);
  accel_limit(&s->angular_vel, s->cmd[1], p->angular_pos_accel_limit, p->angular_neg_accel_limit);
  const float acc_times_dt = (float)(s->linear_vel - (double)s->vel[0]); /* :135 */
  const float dx = (float)((double)s->vel[0] * dt + (0.5 * (double)acc_times_dt) * dt);
  s->vel[0] = (float)s->linear_vel;
  s->vel[2] = (float)s->angular_vel;
  /* :140 Heading(angle) * dx */
  s->pose[0] = s->pose[0] + cosf(s->pose[2]) * dx;
  s->pose[1] = s->pose[1] + sinf(s->pose[2]) * dx;
  /* :141 float + float*double -> AngleMod<double>, narrowed on assignment */
  s->pose[2] = (float)angle_mod_d((double)s->pose[2] + (double)s->vel[2] * dt);
}

static void omni_step(orc_robot_state* s, const umrs_robot_params* p, double dt) {
  /* omnidirectional_model.cpp:82-123 */
  if ((double)s->cmd_age * dt > ORC_MAX_COMMAND_AGE) { /* :85-90 */
    s->cmd[0] = 0.0f;
    s->cmd[1] = 0.0f;
    s->cmd[2] = 0.0f;
  }
  float dvx = s->cmd[0], dvy = s->cmd[1];
  float z = dvx * dvx + dvy * dvy;
  if (sqrtf(z) > p->co_max_speed) { /* :93-96 */
    if (z > 0.0f) {
      const float n = sqrtf(z);
      dvx = dvx / n;
      dvy = dvy / n;
    }
    dvx = dvx * p->co_max_speed;
    dvy = dvy * p->co_max_speed;
  }
  float ax = dvx - s->vel[0], ay = dvy - s->vel[1]; /* :99 */
  const float max_accel = (float)((double)p->co_max_accel * dt);
  z = ax * ax + ay * ay;
  if (sqrtf(z) > max_accel) {
    if (z > 0.0f) {
      const float n = sqrtf(z);
      ax = ax / n;
      ay = ay / n;
    }
    ax = ax * max_accel;
    ay = ay * max_accel;
  }
  s->vel[0] = s->vel[0] + ax;
  s->vel[1] = s->vel[1] + ay;
  float desired_ang = s->cmd[2]; /* :109-118 */
  if (fabsf(desired_ang) > p->co_max_angle_vel)
    desired_ang = (float)signf_(desired_ang) * p->co_max_angle_vel;
  const float max_angle_accel = (float)((double)p->co_max_angle_accel * dt);
  float delta_ang = desired_ang - s->vel[2];
  /* :115 tests |desired_ang_vel| (not |delta|) against the accel cap -- kept. */
  if (fabsf(desired_ang) > max_angle_accel)
    delta_ang = (float)signf_(delta_ang) * max_angle_accel;
  s->vel[2] = s->vel[2] + delta_ang;
  /* :120 (Rotation2Df(angle) * vel.t) * dt, dt narrowed to float by Eigen */
  const float sn = sinf(s->pose[2]), cs = cosf(s->pose[2]);
  const float rx = cs * s->vel[0] - sn * s->vel[1];
  const float ry = sn * s->vel[0] + cs * s->vel[1];
  const float dtf = (float)dt;
  s->pose[0] = s->pose[0] + rx * dtf;
  s->pose[1] = s->pose[1] + ry * dtf;
  s->pose[2] = (float)angle_mod_d((double)s->pose[2] + (double)s->vel[2] * dt);
}

void orc_robot_step(orc_robot_state* s, const umrs_robot_params* p, float dt_f,
                    float n1, float n2) {
  const double dt = (double)dt_f;
  if (p->model_type == UMRS_ACKERMANN_DRIVE) ackermann_step(s, p, dt, n1, n2);
  else if (p->model_type == UMRS_DIFF_DRIVE) diff_drive_step(s, p, dt);
  else omni_step(s, p, dt);
  if (s->cmd_age < ORC_AGE_NEVER) s->cmd_age += 1;
}

void orc_robots_tick(orc_robot_state* s, int32_t n, const umrs_robot_params* p,
                     const float* cmds, float dt, float* fans) {
  for (int32_t i = 0; i < n; ++i) {
    orc_robot_command(&s[i], p, cmds + 3 * i);
    orc_robot_step(&s[i], p, dt, 0.0f, 0.0f);
    orc_laser_origin(s[i].pose, p->laser.loc_x, p->laser.loc_y, fans + 4 * i);
    fans[4 * i + 2] = p->laser.angle_min + s[i].pose[2]; /* simulator.cpp:408 */
    fans[4 * i + 3] = p->laser.angle_max + s[i].pose[2]; /* :409 */
  }
}

/* ---- humans and footprints --------------------------------------------------- */
void orc_human_init(orc_human_state* s, const umrs_human_params* p) {
  /* human_object.cpp:83-85 */
  memset(s, 0, sizeof(*s));
  s->start[0] = p->start_x; s->start[1] = p->start_y; s->start[2] = p->start_theta;
  s->goal[0] = p->goal_x; s->goal[1] = p->goal_y; s->goal[2] = p->goal_theta;
  memcpy(s->pose, s->start, sizeof(s->pose));
}

void orc_human_step(orc_human_state* s, const umrs_human_params* p, float dt_f) {
  /* human_object.cpp:147-161.  avg_speed_/max_speed_ are double members; Eigen
   * narrows a double scalar factor to float before multiplying. */
  const double dt = (double)dt_f;
  const double avg_speed = (double)p->avg_speed, max_speed = (double)p->max_speed;
  float gx = s->goal[0] - s->pose[0], gy = s->goal[1] - s->pose[1];
  float z = gx * gx + gy * gy;
  if (z > 0.0f) {
    const float n = sqrtf(z);
    gx = gx / n;
    gy = gy / n;
  }
  float vx = gx * (float)avg_speed, vy = gy * (float)avg_speed;
  z = vx * vx + vy * vy;
  if ((double)sqrtf(z) > max_speed) { /* :153-155 */
    if (z > 0.0f) {
      const float n = sqrtf(z);
      vx = vx / n;
      vy = vy / n;
    }
    vx = vx * (float)max_speed;
    vy = vy * (float)max_speed;
  }
  s->trans_vel[0] = vx;
  s->trans_vel[1] = vy;
  /* :157 pose_.Set(angle + rot_vel_*dt, translation + trans_vel_*dt) */
  s->pose[2] = (float)((double)s->pose[2] + s->rot_vel * dt);
  s->pose[0] = s->pose[0] + vx * (float)dt;
  s->pose[1] = s->pose[1] + vy * (float)dt;
  /* CheckReachGoal, :178-192 (threshold is a double member) */
  const float ex = s->pose[0] - s->goal[0], ey = s->pose[1] - s->goal[1];
  if ((double)sqrtf(ex * ex + ey * ey) < (double)p->reach_goal_threshold) {
    if (p->mode == 0) {
      s->trans_vel[0] = 0.0f;
      s->trans_vel[1] = 0.0f;
      s->rot_vel = 0.0;
    } else if (p->mode == 1) {
      float t[3];
      memcpy(t, s->goal, sizeof(t));
      memcpy(s->goal, s->start, sizeof(t));
      memcpy(s->start, t, sizeof(t));
    }
  }
}

void orc_human_template(const umrs_human_params* p, float* tmpl) {
  /* human_object.cpp:92-112 */
  const float r = p->radius;
  const int num_segments = (int)p->num_segments; /* CONFIG_FLOAT narrowed to int */
  const float angle_increment = (float)(2 * M_PI / num_segments);
  float v0x = r, v0y = 0.0f, v1x = 0.0f, v1y = 0.0f;
  const float eps = 0.001f;
  int k = 0;
  for (int i = 1; i < num_segments; i++) {
    const float a = angle_increment * (float)i;
    const float sn = sinf(a), cs = cosf(a);
    v1x = cs * r - sn * 0.0f;
    v1y = sn * r + cs * 0.0f;
    float ex = v1x - v0x, ey = v1y - v0y;
    const float z = ex * ex + ey * ey;
    if (z > 0.0f) {
      const float n = sqrtf(z);
      ex = ex / n;
      ey = ey / n;
    }
    ex = ex * eps;
    ey = ey * eps;
    tmpl[4 * k + 0] = v0x + ex; tmpl[4 * k + 1] = v0y + ey;
    tmpl[4 * k + 2] = v1x - ex; tmpl[4 * k + 3] = v1y - ey;
    ++k;
    v0x = v1x;
    v0y = v1y;
  }
  tmpl[4 * k + 0] = v1x; tmpl[4 * k + 1] = v1y;
  tmpl[4 * k + 2] = r;   tmpl[4 * k + 3] = 0.0f;
}

void orc_transform_lines(const float* tmpl, int32_t n, const float pose[3], float* out) {
  /* HumanObject::Transform, human_object.cpp:131-139 (same arithmetic in
   * short_term_object.cpp:60-66): R(AngleMod(angle)) * p + T */
  const float a = angle_mod_f(pose[2]);
  const float sn = sinf(a), cs = cosf(a);
  for (int32_t i = 0; i < 2 * n; ++i) {
    const float x = tmpl[2 * i], y = tmpl[2 * i + 1];
    out[2 * i] = (cs * x - sn * y) + pose[0];
    out[2 * i + 1] = (sn * x + cs * y) + pose[1];
  }
}

void orc_square_template(float* t) {
  /* short_term_object.cpp:53-58: double r, eps; Vector2f(double, double) narrows
   * each coordinate after the double subtraction. */
  const double r = 0.5, eps = 0.001;
  const float seg[16] = {
      (float)(r - eps), (float)r, (float)(-r + eps), (float)r,
      (float)-r, (float)(r - eps), (float)-r, (float)(-r + eps),
      (float)(-r + eps), (float)-r, (float)(r - eps), (float)-r,
      (float)r, (float)(-r + eps), (float)r, (float)(r - eps)};
  memcpy(t, seg, sizeof(seg));
}

void orc_robot_template(const umrs_robot_params* p, float* t) {
  /* NEW (no reference behaviour): body rectangle in the robot frame.  The
   * reference draws its cube centred at pose - heading*rear_axle_offset with
   * scale (car_length, car_width) (simulator.cpp:318-327,380-383); same
   * corner-gap convention as the square above, traversal order identical. */
  const float hl = 0.5f * p->car_length, hw = 0.5f * p->car_width;
  const float cx = -p->rear_axle_offset;
  const float eps = 0.001f;
  const float seg[16] = {
      cx + (hl - eps), hw, cx + (-hl + eps), hw,
      cx - hl, hw - eps, cx - hl, -hw + eps,
      cx + (-hl + eps), -hw, cx + (hl - eps), -hw,
      cx + hl, -hw + eps, cx + hl, hw - eps};
  memcpy(t, seg, sizeof(seg));
}

/* ---- Philox4x32-10 (Salmon et al., SC'11) ---------------------------------- */
void orc_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void orc_box_muller(uint32_t u0, uint32_t u1, float out[2]) {
  /* u in (0,1): (top 24 bits + 0.5) * 2^-24; angle 2*pi*y - pi in (-pi, pi) */
  const float x = ((float)(u0 >> 8) + 0.5f) * (1.0f / 16777216.0f);
  const float y = ((float)(u1 >> 8) + 0.5f) * (1.0f / 16777216.0f);
  const float r = sqrtf(-2.0f * logf(x));
  const float th = 6.28318530717958647692f * y - 3.14159265358979323846f;
  out[0] = r * cosf(th);
  out[1] = r * sinf(th);
}
