// Host-only parts of libumrs: parameter defaults restating the reference's Lua
// config files, the one-time map preparation (VectorMap::Load/Cleanup), and
// small utilities.  No CUDA here.  Citations: file:line in the reference tree.
#include <math.h>
#include <cmath>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../include/umrs.h"
#include "umrs_host.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace umrs {

// Lua numbers are doubles; CONFIG_FLOAT narrows to float (config_reader
// types/config_numeric.h).  DegToRad in the Lua files is math.pi * d / 180.
static float LuaDegToRad(double d) { return (float)(M_PI * d / 180.0); }

// ---- VectorMap::Cleanup, vector_map.cpp:367-395 (host, one-time) --------------
namespace {
struct Seg { float x0, y0, x1, y1; };

inline float Cross(float ax, float ay, float bx, float by) { return ax * by - bx * ay; }  // geometry.h:56-59
inline float Min(float a, float b) { return b < a ? b : a; }
inline float Max(float a, float b) { return a < b ? b : a; }

// Line<float>::Intersection, line2d.h:139-160 (this = l, argument = m).
bool Intersection(const Seg& l, const Seg& m, float* px, float* py) {
  if (Min(l.x0, l.x1) > Max(m.x0, m.x1)) return false;
  if (Max(l.x0, l.x1) < Min(m.x0, m.x1)) return false;
  if (Min(l.y0, l.y1) > Max(m.y0, m.y1)) return false;
  if (Max(l.y0, l.y1) < Min(m.y0, m.y1)) return false;
  const float d1x = l.x1 - l.x0, d1y = l.y1 - l.y0;
  const float d2x = m.x1 - m.x0, d2y = m.y1 - m.y0;
  if ((double)(Cross(d1x, d1y, m.x1 - l.x0, m.y1 - l.y0) *
               Cross(d1x, d1y, m.x0 - l.x0, m.y0 - l.y0)) > 0.0) return false;
  if ((double)(Cross(d2x, d2y, l.x1 - m.x0, l.y1 - m.y0) *
               Cross(d2x, d2y, l.x0 - m.x0, l.y0 - m.y0)) > 0.0) return false;
  const float d = Cross(d2x, d2y, -d1x, -d1y);
  if (d == 0.0f) return false;
  const float tb = Cross(l.x0 - m.x0, l.y0 - m.y0, l.x0 - l.x1, l.y0 - l.y1) / d;
  *px = m.x0 + tb * d2x;
  *py = m.y0 + tb * d2y;
  return true;
}
void Dir(const Seg& l, float* dx, float* dy) {  // line2d.h:62-64
  float x = l.x1 - l.x0, y = l.y1 - l.y0;
  const float z = x * x + y * y;
  if (z > 0.0f) {
    const float n = sqrtf(z);
    x = x / n;
    y = y / n;
  }
  *dx = x;
  *dy = y;
}
float Length(const Seg& l) {  // line2d.h:54-56
  const float x = l.x0 - l.x1, y = l.y0 - l.y1;
  return sqrtf(x * x + y * y);
}
}  // namespace

// Same result as the quadratic loop below, found through a uniform grid over the
// kept segments' bounding boxes (Cleanup is O(L^2) and dominates map
// loading from ~20k segments up).  What must be reproduced is "the FIRST kept line, in
// insertion order, that Intersection() accepts" (vector_map.cpp:377-385).
// Intersection()'s broad phase rejects pairs whose bounding boxes do not overlap with
// exact float comparisons (line2d.h:143-146), so only kept lines whose box overlaps
// l1's can be accepted; two overlapping boxes always share a grid cell because the cell
// index is a monotone function of the coordinate.  Candidates are therefore the kept
// lines registered in the cells l1's box touches (plus the few that span too many
// cells), and the lowest kept index among the accepted ones is the reference's choice.
namespace {
// The reference's loop does not terminate on some inputs: where the 1e-4 m cut
// around a crossing is below the float spacing of the coordinates (beyond ~1 km) the
// pieces pushed back equal the line that was split, forever.  The queue is capped (far
// above anything a terminating run reaches) and the call fails instead of hanging.
// A piece equal to the line it was cut from meets the same kept lines again: a certain cycle.
inline bool SameSeg(const Seg& a, const Seg& b) {
  return a.x0 == b.x0 && a.y0 == b.y0 && a.x1 == b.x1 && a.y1 == b.y1;
}
inline size_t CleanupQueueLimit(int n) { return (size_t)256 * (size_t)(n > 0 ? n : 0) + 65536; }

struct CleanupGrid {
  float x0, y0, inv;
  int gx, gy;
  std::vector<std::vector<int>> cells;
  std::vector<int> wide;          // kept lines spanning more than kMaxSpan cells
  std::vector<unsigned> stamp;    // per kept line: last query that tested it
  unsigned query = 0;
  static constexpr int kMaxSpan = 256;

  int Cx(float x) const {
    const float f = (x - x0) * inv;
    if (!(f > 0.f)) return 0;
    return f >= (float)gx ? gx - 1 : (int)f;
  }
  int Cy(float y) const {
    const float f = (y - y0) * inv;
    if (!(f > 0.f)) return 0;
    return f >= (float)gy ? gy - 1 : (int)f;
  }
  void Box(const Seg& s, int* cx0, int* cx1, int* cy0, int* cy1) const {
    *cx0 = Cx(Min(s.x0, s.x1)); *cx1 = Cx(Max(s.x0, s.x1));
    *cy0 = Cy(Min(s.y0, s.y1)); *cy1 = Cy(Max(s.y0, s.y1));
  }
  void Insert(const Seg& s, int kept_index) {
    int cx0, cx1, cy0, cy1;
    Box(s, &cx0, &cx1, &cy0, &cy1);
    stamp.push_back(0u);
    if ((long)(cx1 - cx0 + 1) * (cy1 - cy0 + 1) > kMaxSpan) {
      wide.push_back(kept_index);
      return;
    }
    for (int cy = cy0; cy <= cy1; ++cy)
      for (int cx = cx0; cx <= cx1; ++cx) cells[(size_t)cy * gx + cx].push_back(kept_index);
  }
};

bool AllFinite(const float* v, size_t n) {
  for (size_t i = 0; i < n; ++i)
    if (!std::isfinite(v[i])) return false;
  return true;
}
}  // namespace

bool CleanupSegmentsGrid(const float* in_xyxy, int n, std::vector<float>* out_xyxy) {
  const float kShrinkDistance = 1e-4f;
  const float kMinLineLength = 0.05f;
  std::vector<Seg> lines(n);
  if (n > 0) memcpy(lines.data(), in_xyxy, sizeof(Seg) * (size_t)n);
  CleanupGrid g;
  float xmin = 0.f, xmax = 0.f, ymin = 0.f, ymax = 0.f;
  for (int i = 0; i < n; ++i) {
    const Seg& s = lines[i];
    if (i == 0) { xmin = xmax = s.x0; ymin = ymax = s.y0; }
    xmin = Min(xmin, Min(s.x0, s.x1)); xmax = Max(xmax, Max(s.x0, s.x1));
    ymin = Min(ymin, Min(s.y0, s.y1)); ymax = Max(ymax, Max(s.y0, s.y1));
  }
  const float ext = Max(Max(xmax - xmin, ymax - ymin), 1e-3f);
  int side = (int)(sqrt((double)(n > 0 ? n : 1)) * 0.75) + 1;
  if (side > 4096) side = 4096;
  const float cell = ext / (float)side;
  g.x0 = xmin; g.y0 = ymin; g.inv = 1.0f / cell;
  g.gx = (int)((xmax - xmin) * g.inv) + 1;
  g.gy = (int)((ymax - ymin) * g.inv) + 1;
  g.cells.resize((size_t)g.gx * g.gy);
  std::vector<Seg> kept;
  kept.reserve(n);
  const size_t limit = CleanupQueueLimit(n);
  for (size_t i = 0; i < lines.size(); ++i) {  // the list grows while iterating
    if (lines.size() > limit) return false;
    const Seg l1 = lines[i];
    if (Length(l1) < kMinLineLength) continue;
    int best = -1;
    float bpx = 0.f, bpy = 0.f;
    ++g.query;
    auto consider = [&](int k) {
      if (g.stamp[(size_t)k] == g.query) return;
      g.stamp[(size_t)k] = g.query;
      if (best >= 0 && k > best) return;
      float px, py;
      if (Intersection(kept[(size_t)k], l1, &px, &py)) { best = k; bpx = px; bpy = py; }
    };
    int cx0, cx1, cy0, cy1;
    g.Box(l1, &cx0, &cx1, &cy0, &cy1);
    for (int cy = cy0; cy <= cy1; ++cy)
      for (int cx = cx0; cx <= cx1; ++cx)
        for (int k : g.cells[(size_t)cy * g.gx + cx]) consider(k);
    for (int k : g.wide) consider(k);
    if (best >= 0) {
      float dx, dy;
      Dir(l1, &dx, &dy);
      const float sx = kShrinkDistance * dx, sy = kShrinkDistance * dy;
      lines.push_back(Seg{l1.x0, l1.y0, bpx - sx, bpy - sy});
      lines.push_back(Seg{bpx + sx, bpy + sy, l1.x1, l1.y1});
      if (SameSeg(lines[lines.size() - 2], l1) || SameSeg(lines.back(), l1)) return false;
    } else {
      g.Insert(l1, (int)kept.size());
      kept.push_back(l1);
    }
  }
  for (Seg& l : kept) {  // ShrinkLine, vector_map.cpp:359-365
    const float len = Length(l);
    float dx, dy;
    Dir(l, &dx, &dy);
    if ((double)len < 2.0 * (double)kShrinkDistance) continue;
    l.x0 = l.x0 + kShrinkDistance * dx;
    l.y0 = l.y0 + kShrinkDistance * dy;
    l.x1 = l.x1 - kShrinkDistance * dx;
    l.y1 = l.y1 - kShrinkDistance * dy;
  }
  out_xyxy->resize(kept.size() * 4);
  if (!kept.empty()) memcpy(out_xyxy->data(), kept.data(), sizeof(Seg) * kept.size());
  return true;
}

// algorithm: 0 auto (grid from 512 segments up, if every coordinate is finite),
// 1 the reference's quadratic loop, 2 grid.
bool CleanupSegmentsWith(const float* in_xyxy, int n, int algorithm, std::vector<float>* out_xyxy) {
  const bool finite = AllFinite(in_xyxy, (size_t)(n > 0 ? n : 0) * 4);
  if (algorithm == 2 ? finite : (algorithm == 0 && n >= 512 && finite)) {
    return CleanupSegmentsGrid(in_xyxy, n, out_xyxy);
  }
  return CleanupSegmentsQuadratic(in_xyxy, n, out_xyxy);
}

bool CleanupSegments(const float* in_xyxy, int n, std::vector<float>* out_xyxy) {
  return CleanupSegmentsWith(in_xyxy, n, 0, out_xyxy);
}

bool CleanupSegmentsQuadratic(const float* in_xyxy, int n, std::vector<float>* out_xyxy) {
  const float kShrinkDistance = 1e-4f;
  const float kMinLineLength = 0.05f;
  std::vector<Seg> lines(n);
  if (n > 0) memcpy(lines.data(), in_xyxy, sizeof(Seg) * (size_t)n);
  std::vector<Seg> kept;
  kept.reserve(n);
  const size_t limit = CleanupQueueLimit(n);
  for (size_t i = 0; i < lines.size(); ++i) {  // the list grows while iterating
    if (lines.size() > limit) return false;
    const Seg l1 = lines[i];
    if (Length(l1) < kMinLineLength) continue;
    bool split = false;
    for (const Seg& l2 : kept) {
      float px, py;
      if (Intersection(l2, l1, &px, &py)) {
        float dx, dy;
        Dir(l1, &dx, &dy);
        const float sx = kShrinkDistance * dx, sy = kShrinkDistance * dy;
        lines.push_back(Seg{l1.x0, l1.y0, px - sx, py - sy});
        lines.push_back(Seg{px + sx, py + sy, l1.x1, l1.y1});
        if (SameSeg(lines[lines.size() - 2], l1) || SameSeg(lines.back(), l1)) return false;
        split = true;
        break;
      }
    }
    if (!split) kept.push_back(l1);
  }
  for (Seg& l : kept) {  // ShrinkLine, vector_map.cpp:359-365
    const float len = Length(l);
    float dx, dy;
    Dir(l, &dx, &dy);
    if ((double)len < 2.0 * (double)kShrinkDistance) continue;
    l.x0 = l.x0 + kShrinkDistance * dx;
    l.y0 = l.y0 + kShrinkDistance * dy;
    l.x1 = l.x1 - kShrinkDistance * dx;
    l.y1 = l.y1 - kShrinkDistance * dy;
  }
  out_xyxy->resize(kept.size() * 4);
  if (!kept.empty()) memcpy(out_xyxy->data(), kept.data(), sizeof(Seg) * kept.size());
  return true;
}

// VectorMap::Load, vector_map.cpp:397-411: "%f,%f,%f,%f" until the first mismatch.
bool LoadMapFile(const char* path, std::vector<float>* raw_xyxy) {
  FILE* fid = fopen(path, "r");
  if (fid == NULL) return false;
  raw_xyxy->clear();
  float x1 = 0, y1 = 0, x2 = 0, y2 = 0;
  while (fscanf(fid, "%f,%f,%f,%f", &x1, &y1, &x2, &y2) == 4) {
    raw_xyxy->push_back(x1);
    raw_xyxy->push_back(y1);
    raw_xyxy->push_back(x2);
    raw_xyxy->push_back(y2);
  }
  fclose(fid);
  return true;
}

// ---- footprint templates ----------------------------------------------------------
// HumanObject ctor, human_object.cpp:92-112.
void HumanTemplate(const umrs_human_params& p, std::vector<float>* tmpl) {
  const float r = p.radius;
  const int num_segments = (int)p.num_segments;
  tmpl->assign((size_t)(num_segments > 0 ? num_segments : 0) * 4, 0.f);
  if (num_segments <= 0) return;
  const float angle_increment = (float)(2 * M_PI / num_segments);
  float v0x = r, v0y = 0.f, v1x = 0.f, v1y = 0.f;
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
    float* t = tmpl->data() + 4 * k++;
    t[0] = v0x + ex; t[1] = v0y + ey; t[2] = v1x - ex; t[3] = v1y - ey;
    v0x = v1x;
    v0y = v1y;
  }
  float* t = tmpl->data() + 4 * k;
  t[0] = v1x; t[1] = v1y; t[2] = r; t[3] = 0.f;
}

// ShortTermObject, short_term_object.cpp:53-58 (double r, eps; narrowed per coordinate).
void SquareTemplate(float* t) {
  const double r = 0.5, eps = 0.001;
  const float seg[16] = {
      (float)(r - eps), (float)r, (float)(-r + eps), (float)r,
      (float)-r, (float)(r - eps), (float)-r, (float)(-r + eps),
      (float)(-r + eps), (float)-r, (float)(r - eps), (float)-r,
      (float)r, (float)(-r + eps), (float)r, (float)(r - eps)};
  memcpy(t, seg, sizeof(seg));
}

// Robot body rectangle (new behaviour; the reference only draws this box,
// simulator.cpp:318-327,380-383).  Same traversal and 1 mm corner gaps as the square.
void RobotTemplate(const umrs_robot_params& p, float* t) {
  const float hl = 0.5f * p.car_length, hw = 0.5f * p.car_width;
  const float cx = -p.rear_axle_offset;
  const float eps = 0.001f;
  const float seg[16] = {
      cx + (hl - eps), hw, cx + (-hl + eps), hw,
      cx - hl, hw - eps, cx - hl, -hw + eps,
      cx + (-hl + eps), -hw, cx + (hl - eps), -hw,
      cx + hl, -hw + eps, cx + hl, hw - eps};
  memcpy(t, seg, sizeof(seg));
}

}  // namespace umrs

// ================================ C-ABI (host-only part) ============================
extern "C" {

int umrs_version(void) { return UMRS_VERSION; }

const char* umrs_status_string(int status) {
  switch (status) {
    case UMRS_OK: return "ok";
    case UMRS_ERR_INVALID: return "invalid argument";
    case UMRS_ERR_CUDA: return "CUDA error";
    case UMRS_ERR_NOMEM: return "out of memory";
    case UMRS_ERR_UNSUPPORTED: return "unsupported configuration";
    case UMRS_ERR_NO_DEVICE: return "no usable sm_100 CUDA device";
    case UMRS_ERR_IO: return "I/O error";
    default: return "unknown status";
  }
}

int umrs_model_type_from_string(const char* s) {
  if (s == NULL) return -1;
  if (strcmp(s, "ACKERMANN_DRIVE") == 0) return UMRS_ACKERMANN_DRIVE;
  if (strcmp(s, "OMNIDIRECTIONAL_DRIVE") == 0) return UMRS_OMNIDIRECTIONAL_DRIVE;
  if (strcmp(s, "DIFF_DRIVE") == 0) return UMRS_DIFF_DRIVE;
  return -1;
}

int umrs_laser_num_rays(const umrs_laser_params* l) {
  if (l == NULL) return -1;
  // simulator.cpp:402-404: float subtraction and division, double add, int cast
  return (int)(1.0 + (l->angle_max - l->angle_min) / l->angle_increment);
}

void umrs_default_laser(umrs_laser_params* o) {
  o->loc_x = 0.2f; o->loc_y = 0.f;                       // sim_config.lua:31
  o->angle_min = umrs::LuaDegToRad(-135.0);              // :40
  o->angle_max = umrs::LuaDegToRad(135.0);               // :41
  o->angle_increment = umrs::LuaDegToRad(0.25);          // :42
  o->range_min = 0.4f; o->range_max = 100.0f;            // :43-44
  o->noise_stddev = 0.01f;                               // :39
}

void umrs_default_jackal_laser(umrs_laser_params* o) {
  o->loc_x = 0.07f; o->loc_y = 0.f;                      // ut_jackal_config.lua:39
  o->angle_min = umrs::LuaDegToRad(-180.0);              // :44
  o->angle_max = umrs::LuaDegToRad(180.0);               // :45
  o->angle_increment = umrs::LuaDegToRad(0.1);           // :46
  o->range_min = 0.4f; o->range_max = 150.0f;            // :47-48
  o->noise_stddev = 0.01f;                               // sim_config.lua:39 (not overridden)
}

static void zero_robot(umrs_robot_params* o) {
  memset(o, 0, sizeof(*o));
  // keys a config file leaves unset keep every model's own defaults below
  o->ak_min_turn_radius = 0.98f; o->ak_max_speed = 1.2f; o->ak_max_accel = 3.0f;
  o->ak_angular_error_bias = 0.f; o->ak_angular_error_rate = 0.1f;
  o->linear_pos_accel_limit = 3.0f; o->linear_neg_accel_limit = 3.0f;
  o->angular_pos_accel_limit = 3.0f; o->angular_neg_accel_limit = 3.0f;
  o->max_angular = 3.0f; o->max_linear_vel = 2.0f;
  o->co_max_accel = 3.0f; o->co_max_angle_accel = (float)M_PI;
  o->co_max_speed = 1.2f; o->co_max_angle_vel = (float)M_PI;
  o->footprint_enabled = 1;
}

void umrs_default_ut_automata(umrs_robot_params* o) {
  zero_robot(o);
  o->model_type = UMRS_ACKERMANN_DRIVE;
  // ut_automata_config.lua:4-10 (+ require sim_config: car dims, laser)
  o->car_width = 0.281f; o->car_length = 0.535f;         // sim_config.lua:25-26
  o->rear_axle_offset = -0.162f;                         // :30
  umrs_default_laser(&o->laser);
}

void umrs_default_ut_jackal(umrs_robot_params* o) {
  zero_robot(o);
  o->model_type = UMRS_DIFF_DRIVE;
  o->max_linear_vel = 2.0f;                              // ut_jackal_config.lua:21
  o->car_width = 0.43f; o->car_length = 0.50f;           // :40-41
  o->rear_axle_offset = 0.f;                             // :35
  umrs_default_jackal_laser(&o->laser);
}

void umrs_default_bwibot(umrs_robot_params* o) {
  zero_robot(o);
  o->model_type = UMRS_DIFF_DRIVE;
  o->max_linear_vel = 3.0f;                              // bwibot_config.lua:9
  o->car_width = 0.34f; o->car_length = 0.34f;           // :27-28
  o->rear_axle_offset = 0.f;                             // :23
  umrs_default_laser(&o->laser);                         // sim_config.lua:39-44
  o->laser.loc_x = 0.15f; o->laser.loc_y = 0.f;          // bwibot_config.lua:26
}

void umrs_default_cobot(umrs_robot_params* o) {
  zero_robot(o);
  o->model_type = UMRS_OMNIDIRECTIONAL_DRIVE;
  // cobot_config.lua:13-18; body = square of side 2 * co_base_radius (:7)
  o->car_width = 0.4f; o->car_length = 0.4f; o->rear_axle_offset = 0.f;
  umrs_default_laser(&o->laser);
}

void umrs_default_human(umrs_human_params* o) {
  // human_config.lua:2-28
  o->radius = 0.1f; o->num_segments = 20;
  o->start_x = 39.f; o->start_y = 18.7f; o->start_theta = 0.f;
  o->goal_x = 42.f; o->goal_y = 22.7f; o->goal_theta = 0.f;
  o->max_speed = 1.5f; o->avg_speed = 1.f; o->max_omega = 0.2f; o->avg_omega = 0.f;
  o->reach_goal_threshold = 0.3f;
  o->mode = 1;
}

int umrs_map_cleanup(const float* segs_xyxy, int32_t n, float* out_xyxy, int32_t cap,
                     int32_t* n_out) {
  return umrs_map_cleanup_ex(segs_xyxy, n, out_xyxy, cap, n_out, 0);
}

int umrs_map_cleanup_ex(const float* segs_xyxy, int32_t n, float* out_xyxy, int32_t cap,
                        int32_t* n_out, int32_t algorithm) {
  if (n < 0 || (n > 0 && segs_xyxy == NULL) || n_out == NULL || algorithm < 0 || algorithm > 2)
    return UMRS_ERR_INVALID;
  std::vector<float> out;
  *n_out = 0;
  // the split queue ran away: an input on which the reference's own loop never ends
  if (!umrs::CleanupSegmentsWith(segs_xyxy, n, algorithm, &out)) return UMRS_ERR_UNSUPPORTED;
  *n_out = (int32_t)(out.size() / 4);
  if (*n_out > cap || (out_xyxy == NULL && *n_out > 0)) return UMRS_ERR_INVALID;
  if (!out.empty()) memcpy(out_xyxy, out.data(), sizeof(float) * out.size());
  return UMRS_OK;
}

void umrs_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

}  // extern "C"
