// Drives include/umrs_vector_map.hpp the way Simulator::init / publishLaser drive
// the reference's VectorMap (simulator.cpp:173, 405-411).  Usage:
//   adapter_demo <map.txt> <x> <y> <range_max> <angle_min> <angle_max> <num_rays> [obj x0 y0 x1 y1]...
// prints one range per line (%.9g).  Exit code 3 = the library reported an error.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cstring>

#include "umrs_robot_model.hpp"
#include "umrs_vector_map.hpp"

struct Vec2 {
  float x_, y_;
  Vec2() : x_(0), y_(0) {}
  Vec2(float x, float y) : x_(x), y_(y) {}
  float x() const { return x_; }
  float y() const { return y_; }
};
struct Line {
  Vec2 p0, p1;
  Line(const Vec2& a, const Vec2& b) : p0(a), p1(b) {}
};

// stand-in for pose_2d::Pose2Df (poses_2d.h:32-104)
struct Pose {
  Vec2 translation;
  float angle;
  Pose(float a, const Vec2& t) : translation(t), angle(a) {}
};

// adapter_demo model <map.txt> <ut_automata|ut_jackal|cobot> <x> <y> <theta> <steps> <c0> <c1> <c2>
// drives include/umrs_robot_model.hpp the way Simulator::update drives a motion model
// (simulator.cpp:479-486); prints "pose x y theta vx vy w" per step, then the last scan.
static int RunModel(int argc, char** argv) {
  if (argc < 11) return 2;
  umrs::VectorMapT<Vec2, Line> map_;
  map_.Load(argv[2]);
  umrs_robot_params rp;
  if (!std::strcmp(argv[3], "ut_jackal")) umrs_default_ut_jackal(&rp);
  else if (!std::strcmp(argv[3], "cobot")) umrs_default_cobot(&rp);
  else umrs_default_ut_automata(&rp);
  umrs::RobotModelT<Pose, Vec2, Line> model(map_.handle(), rp, 0.025f);
  model.SetPose(Pose(atof(argv[6]), Vec2(atof(argv[4]), atof(argv[5]))));
  const int steps = atoi(argv[7]);
  for (int s = 0; s < steps; ++s) {
    if (s % 3 != 2) model.DriveCallback(atof(argv[8]), atof(argv[9]), atof(argv[10]));   // a gap every third tick
    model.Step(0.025f);
    const Pose p = model.GetPose(), v = model.GetVel();
    std::printf("pose %.9g %.9g %.9g %.9g %.9g %.9g\n", p.translation.x(), p.translation.y(), p.angle,
                v.translation.x(), v.translation.y(), v.angle);
  }
  std::printf("body %zu rays %d\n", model.GetLines().size(), model.header().num_rays);
  for (float r : model.ranges()) std::printf("%.9g\n", r);
  return 0;
}

// stand-in for sensor_msgs::LaserScan (the fields simulator.cpp:150-159 sets)
struct LaserScanPod {
  float angle_min, angle_max, angle_increment, time_increment, scan_time, range_min, range_max;
  std::vector<float> ranges, intensities;
};

static int CheckFillLaserScan() {
  umrs_laser_scan_header h = {};
  h.angle_min = -1.f; h.angle_max = 1.f; h.angle_increment = 0.5f; h.scan_time = 0.05f;
  h.range_min = 0.4f; h.range_max = 30.f; h.num_rays = 4; h.ranges_offset = 2;
  const float block[8] = {9, 9, 1, 2, 3, 4, 9, 9};
  LaserScanPod m;
  m.intensities.push_back(1.f);
  umrs::FillLaserScan(h, block, &m);
  return (m.ranges.size() == 4 && m.ranges[0] == 1.f && m.ranges[3] == 4.f && m.intensities.empty() &&
          m.angle_increment == 0.5f && m.scan_time == 0.05f && m.range_max == 30.f) ? 0 : 1;
}

int main(int argc, char** argv) {
  if (CheckFillLaserScan() != 0) return 4;
  if (argc > 1 && !std::strcmp(argv[1], "model")) {
    try {
      return RunModel(argc, argv);
    } catch (const std::exception& e) {
      std::fprintf(stderr, "%s\n", e.what());
      return 3;
    }
  }
  if (argc < 8) return 2;
  try {
    umrs::VectorMapT<Vec2, Line> map_;
    map_.Load(argv[1]);
    for (int k = 8; k + 3 < argc; k += 4)
      map_.object_lines.push_back(Line(Vec2(atof(argv[k]), atof(argv[k + 1])),
                                       Vec2(atof(argv[k + 2]), atof(argv[k + 3]))));
    std::vector<float> ranges;
    map_.GetPredictedScan(Vec2(atof(argv[2]), atof(argv[3])), 0.4f, atof(argv[4]), atof(argv[5]),
                          atof(argv[6]), atoi(argv[7]), &ranges);
    std::printf("lines %zu\n", map_.lines.size());
    for (float r : ranges) std::printf("%.9g\n", r);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 3;
  }
  return 0;
}
