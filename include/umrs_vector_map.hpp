// umrs_vector_map.hpp -- header-only C++ host adapter over the C-ABI (umrs.h) with the
// reference's own signatures for the scan path, so that the call sites in
// ut_multirobot_sim compile against it unchanged:
//
//   vector_map::VectorMap::Load             vector_map.h:75     (simulator.cpp:173)
//   vector_map::VectorMap::lines            vector_map.h:78
//   vector_map::VectorMap::object_lines     vector_map.h:81     (simulator.cpp:501-506)
//   vector_map::VectorMap::GetPredictedScan vector_map.h:66-72  (simulator.cpp:405-411)
//
// The vector / line types are template parameters so that this header needs
// neither Eigen nor the reference's headers: on a ROS machine instantiate
// umrs::VectorMapT<Eigen::Vector2f, geometry::Line2f>; the tests use PODs.
// Vec2 needs x(), y(); Line needs members p0, p1 of type Vec2 and a
// constructor Line(Vec2, Vec2).
//
// Error behaviour: the reference exit(1)s on a missing map (vector_map.cpp:399-402)
// and has no other failure path; here a failed call throws std::runtime_error
// carrying umrs_status_string() + umrs_last_cuda_error() (define
// UMRS_ADAPTER_EXIT_ON_ERROR to get the reference's exit(1) instead).
#ifndef UMRS_VECTOR_MAP_HPP_
#define UMRS_VECTOR_MAP_HPP_

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "umrs.h"

namespace umrs {

template <class Vec2, class Line>
class VectorMapT {
 public:
  VectorMapT() {}
  explicit VectorMapT(const std::string& name) { Load(name); }
  ~VectorMapT() { umrs_map_destroy(map_); }
  VectorMapT(const VectorMapT&) = delete;
  VectorMapT& operator=(const VectorMapT&) = delete;

  // VectorMap::Load: text "%f,%f,%f,%f" per line + Cleanup() (vector_map.cpp:397-411).
  void Load(const std::string& file) {
    umrs_map_destroy(map_);
    map_ = nullptr;
    Check(umrs_map_load_file(file.c_str(), &map_), "umrs_map_load_file");
    file_name = file;
    int32_t n = 0;
    Check(umrs_map_num_segments(map_, &n), "umrs_map_num_segments");
    std::vector<float> flat(static_cast<size_t>(n) * 4);
    Check(umrs_map_get_segments(map_, flat.data(), n), "umrs_map_get_segments");
    lines.clear();
    for (int32_t i = 0; i < n; ++i)
      lines.push_back(Line(Vec2(flat[4 * i], flat[4 * i + 1]), Vec2(flat[4 * i + 2], flat[4 * i + 3])));
  }

  // VectorMap::GetPredictedScan (vector_map.h:66-72): same seven arguments, same
  // resize-and-fill of *scan; reads `object_lines` like the reference.
  void GetPredictedScan(const Vec2& loc, float range_min, float range_max, float angle_min,
                        float angle_max, int num_rays, std::vector<float>* scan) {
    flat_obj_.resize(object_lines.size() * 4);
    for (size_t i = 0; i < object_lines.size(); ++i) {
      flat_obj_[4 * i + 0] = object_lines[i].p0.x();
      flat_obj_[4 * i + 1] = object_lines[i].p0.y();
      flat_obj_[4 * i + 2] = object_lines[i].p1.x();
      flat_obj_[4 * i + 3] = object_lines[i].p1.y();
    }
    scan->resize(static_cast<size_t>(num_rays));
    if (reference_render) {
      // the reference's own SceneRender/TrimOcclusion pass on the GPU: its very bits
      const float fan[4] = {loc.x(), loc.y(), angle_min, angle_max};
      Check(umrs_scan_batch_reference_render(map_, flat_obj_.empty() ? nullptr : flat_obj_.data(),
                                             static_cast<int32_t>(object_lines.size()), fan, 1, range_max,
                                             num_rays, scan->data()),
            "umrs_scan_batch_reference_render");
      return;
    }
    Check(umrs_predicted_scan(map_, flat_obj_.empty() ? nullptr : flat_obj_.data(),
                              static_cast<int32_t>(object_lines.size()), loc.x(), loc.y(), range_min,
                              range_max, angle_min, angle_max, num_rays, scan->data(), nullptr),
          "umrs_predicted_scan");
  }

  umrs_map* handle() const { return map_; }

  std::vector<Line> lines;         // vector_map.h:78 (read-only mirror of the device map)
  std::vector<Line> object_lines;  // vector_map.h:81
  std::string file_name;           // vector_map.h:79
  // false (default): exact nearest hit per ray.  true: bit-identical to the reference's
  // GetPredictedScan, occlusion-trimming defects included (slower: O(L^2) per scan).
  bool reference_render = false;

 private:
  static void Check(int status, const char* where) {
    if (status == UMRS_OK) return;
    const std::string msg = std::string(where) + ": " + umrs_status_string(status) + " [" +
                            umrs_last_cuda_error() + "]";
#ifdef UMRS_ADAPTER_EXIT_ON_ERROR
    std::fprintf(stderr, "%s\n", msg.c_str());
    std::exit(1);
#else
    throw std::runtime_error(msg);
#endif
  }
  umrs_map* map_ = nullptr;
  std::vector<float> flat_obj_;
};

// Fills a sensor_msgs::LaserScan-shaped message the way Simulator::init sets its template
// (simulator.cpp:150-159) and publishLaser its ranges (simulator.cpp:419-424), from one robot's
// header (umrs_world_scan_header) and one world's block of the ranges buffer.  LaserScanMsg needs
// the members angle_min, angle_max, angle_increment, time_increment, scan_time, range_min,
// range_max, ranges (vector<float>) and intensities (any container with clear()).
template <class LaserScanMsg>
void FillLaserScan(const umrs_laser_scan_header& h, const float* world_ranges, LaserScanMsg* msg) {
  msg->angle_min = h.angle_min;
  msg->angle_max = h.angle_max;
  msg->angle_increment = h.angle_increment;
  msg->time_increment = h.time_increment;
  msg->scan_time = h.scan_time;
  msg->range_min = h.range_min;
  msg->range_max = h.range_max;
  msg->ranges.assign(world_ranges + h.ranges_offset, world_ranges + h.ranges_offset + h.num_rays);
  msg->intensities.clear();
}

}  // namespace umrs
#endif  // UMRS_VECTOR_MAP_HPP_
