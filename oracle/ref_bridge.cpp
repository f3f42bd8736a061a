// TEST INFRASTRUCTURE ONLY -- never linked into or called by the product.
//
// extern "C" bridge over the UNMODIFIED reference scan path, compiled in place
// from /root/reference (see oracle/Makefile, target _ref/libumrs_ref.so):
//   src/simulator/vector_map.cpp            VectorMap::{Load,Cleanup,GetSceneLines,
//                                           SceneRender,GetPredictedScan}
//   submodules/shared/util/timer.cc         CumulativeFunctionTimer (static in the scan)
//   submodules/shared/math/{line2d,geometry,math_util}.h   primitives
// against the Eigen/glog stand-ins in oracle/shim.  This is "tier A" of the
// parity contract and the CPU baseline of kind "reference".
//
// Nothing here restates reference logic; every function only marshals plain
// arrays into the reference's own types and calls the reference's own code.

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "eigen3/Eigen/Dense"
#include "shared/math/geometry.h"
#include "shared/math/line2d.h"
#include "shared/math/math_util.h"
#include "shared/math/poses_2d.h"
#include "vector_map.h"
#include "simulator/short_term_object.h"

using Eigen::Vector2f;
using geometry::Line2f;
using vector_map::VectorMap;

namespace {
void FillLines(const float* xyxy, int n, std::vector<Line2f>* out) {
  out->clear();
  out->reserve(n);
  for (int i = 0; i < n; ++i) {
    out->push_back(Line2f(Vector2f(xyxy[4 * i], xyxy[4 * i + 1]),
                          Vector2f(xyxy[4 * i + 2], xyxy[4 * i + 3])));
  }
}
int DumpLines(const std::vector<Line2f>& lines, float* out, int cap) {
  const int n = static_cast<int>(lines.size());
  for (int i = 0; i < n && i < cap; ++i) {
    out[4 * i + 0] = lines[i].p0.x();
    out[4 * i + 1] = lines[i].p0.y();
    out[4 * i + 2] = lines[i].p1.x();
    out[4 * i + 3] = lines[i].p1.y();
  }
  return n;
}
}  // namespace

extern "C" {

// ---- map ------------------------------------------------------------------
// VectorMap::Load (vector_map.cpp:397-411): text "%f,%f,%f,%f" + Cleanup().
void* umrs_ref_map_load(const char* path) {
  FILE* f = fopen(path, "r");
  if (f == NULL) return NULL;  // the reference would exit(1) (vector_map.cpp:399-402)
  fclose(f);
  VectorMap* m = new VectorMap();
  m->Load(path);
  return m;
}

// Raw segments; cleanup != 0 runs VectorMap::Cleanup (vector_map.cpp:367-395).
void* umrs_ref_map_from_segments(const float* xyxy, int n, int cleanup) {
  VectorMap* m = new VectorMap();
  FillLines(xyxy, n, &m->lines);
  if (cleanup) m->Cleanup();
  return m;
}

void umrs_ref_map_free(void* map) { delete static_cast<VectorMap*>(map); }

int umrs_ref_map_num_lines(void* map) {
  return static_cast<int>(static_cast<VectorMap*>(map)->lines.size());
}

int umrs_ref_map_get_lines(void* map, float* out, int cap) {
  return DumpLines(static_cast<VectorMap*>(map)->lines, out, cap);
}

// map_.object_lines as Simulator::update fills it (simulator.cpp:501-506).
void umrs_ref_map_set_object_lines(void* map, const float* xyxy, int n) {
  FillLines(xyxy, n, &static_cast<VectorMap*>(map)->object_lines);
}

// VectorMap::GetSceneLines (vector_map.cpp:192-215): the candidate order that
// defines hit indices.
int umrs_ref_scene_lines(void* map, float x, float y, float max_range,
                         float* out, int cap) {
  std::vector<Line2f> l;
  static_cast<VectorMap*>(map)->GetSceneLines(Vector2f(x, y), max_range, &l);
  return DumpLines(l, out, cap);
}

// VectorMap::SceneRender (vector_map.cpp:217-258): visible sub-segments.
int umrs_ref_scene_render(void* map, float x, float y, float max_range,
                          float angle_min, float angle_max, float* out, int cap) {
  std::vector<Line2f> l;
  static_cast<VectorMap*>(map)->SceneRender(Vector2f(x, y), max_range, angle_min,
                                            angle_max, &l);
  return DumpLines(l, out, cap);
}

// ---- the scan: tier A --------------------------------------------------------
// VectorMap::GetPredictedScan (vector_map.cpp:420-478), verbatim call.
void umrs_ref_predicted_scan(void* map, float x, float y, float range_min,
                             float range_max, float angle_min, float angle_max,
                             int num_rays, float* ranges_out) {
  std::vector<float> scan;
  static_cast<VectorMap*>(map)->GetPredictedScan(
      Vector2f(x, y), range_min, range_max, angle_min, angle_max, num_rays, &scan);
  memcpy(ranges_out, scan.data(), sizeof(float) * scan.size());
}

// Harness-parallel batch for the CPU baseline: fans[f] = {x, y, angle_min,
// angle_max}; every thread owns a private VectorMap copy (the routine holds a
// static timer and mutates nothing else, vector_map.cpp:427; the unsynchronised
// statistics counters are the only shared state).  object lines (optional) are
// per fan: obj_xyxy + 4*obj_offset[f] .. 4*obj_offset[f+1].
void umrs_ref_scan_batch(void* map, const float* fans, int n_fans,
                         const float* obj_xyxy, const int32_t* obj_offset,
                         float range_min, float range_max, int num_rays,
                         int n_threads, float* ranges_out) {
  const VectorMap* src = static_cast<VectorMap*>(map);
  if (n_threads < 1) n_threads = 1;
  auto work = [&](int tid) {
    VectorMap local(src->lines);
    std::vector<float> scan;
    for (int f = tid; f < n_fans; f += n_threads) {
      if (obj_xyxy != NULL && obj_offset != NULL) {
        FillLines(obj_xyxy + 4 * obj_offset[f], obj_offset[f + 1] - obj_offset[f],
                  &local.object_lines);
      }
      local.GetPredictedScan(Vector2f(fans[4 * f], fans[4 * f + 1]), range_min,
                             range_max, fans[4 * f + 2], fans[4 * f + 3], num_rays,
                             &scan);
      memcpy(ranges_out + static_cast<size_t>(f) * num_rays, scan.data(),
             sizeof(float) * num_rays);
    }
  };
  if (n_threads == 1) {
    work(0);
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < n_threads; ++t) pool.emplace_back(work, t);
  for (auto& t : pool) t.join();
}

// ---- primitives (for the gtest known-answer tables, line2d_tests.cc:32-87) ---
int umrs_ref_line_intersects(const float* a, const float* b) {
  return Line2f(a[0], a[1], a[2], a[3]).Intersects(Vector2f(b[0], b[1]),
                                                   Vector2f(b[2], b[3]));
}
int umrs_ref_line_intersection(const float* a, const float* b, float* out) {
  Vector2f p(0, 0);
  const bool r = Line2f(a[0], a[1], a[2], a[3])
                     .Intersection(Vector2f(b[0], b[1]), Vector2f(b[2], b[3]), &p);
  out[0] = p.x();
  out[1] = p.y();
  return r;
}
int umrs_ref_ray_intersects(const float* a, float px, float py, float dx, float dy) {
  return Line2f(a[0], a[1], a[2], a[3]).RayIntersects(Vector2f(px, py), Vector2f(dx, dy));
}
void umrs_ref_ray_intersection(const float* a, float px, float py, float dx,
                               float dy, float* out) {
  const Vector2f p = Line2f(a[0], a[1], a[2], a[3])
                         .RayIntersection(Vector2f(px, py), Vector2f(dx, dy));
  out[0] = p.x();
  out[1] = p.y();
}
void umrs_ref_unit_normal(const float* a, float* out) {
  const Vector2f n = Line2f(a[0], a[1], a[2], a[3]).UnitNormal();
  out[0] = n.x();
  out[1] = n.y();
}
float umrs_ref_cross(float ax, float ay, float bx, float by) {
  return geometry::Cross(Vector2f(ax, ay), Vector2f(bx, by));
}
float umrs_ref_angle_mod_f(float a) { return math_util::AngleMod<float>(a); }
double umrs_ref_angle_mod_d(double a) { return math_util::AngleMod<double>(a); }
float umrs_ref_angle_diff_f(float a, float b) { return math_util::AngleDiff<float>(a, b); }
void umrs_ref_heading_f(float a, float* out) {
  const Vector2f h = geometry::Heading<float>(a);
  out[0] = h.x();
  out[1] = h.y();
}
// Pose2Df::ApplyPose (poses_2d.h:96-100).
void umrs_ref_apply_pose(float* pose_xyt, const float* other_xyt) {
  pose_2d::Pose2Df p(pose_xyt[2], Vector2f(pose_xyt[0], pose_xyt[1]));
  p.ApplyPose(pose_2d::Pose2Df(other_xyt[2], Vector2f(other_xyt[0], other_xyt[1])));
  pose_xyt[0] = p.translation.x();
  pose_xyt[1] = p.translation.y();
  pose_xyt[2] = p.angle;
}

// ShortTermObject(config) (short_term_object.cpp:39-67): template square and the
// pose lines at whichever of its two hard-coded spots srand(time(0)) picked.
// out_template / out_pose_lines: 4 segments each; out_pose: x, y, theta.
int umrs_ref_short_term_object(float* out_template, float* out_pose_lines,
                               float* out_pose) {
  ShortTermObject obj("short_term_config.lua");
  DumpLines(obj.GetTemplateLines(), out_template, 4);
  const int n = DumpLines(obj.GetLines(), out_pose_lines, 4);
  const pose_2d::Pose2Df p = obj.GetPose();
  out_pose[0] = p.translation.x();
  out_pose[1] = p.translation.y();
  out_pose[2] = p.angle;
  return n;
}

}  // extern "C"
