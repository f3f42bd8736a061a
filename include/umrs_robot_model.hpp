// C++ host adapter: one robot of the reference's simulator behind the reference's own entity
// interface, over the C-ABI.  Mirrors
//   EntityBase::Step / SetPose / GetPose / GetLines        entity_base.h:45-53
//   robot_model::RobotModel::SetVel / GetVel               robot_model.h:36-37
// as Simulator::update drives them (simulator.cpp:479-486), for the three model families
// (MakeMotionModel, simulator.cpp:128-142: the type lives in umrs_robot_params::model_type).
//
// The library steps a robot and scans from its new pose in ONE fused kernel, so Step() also
// produces the tick's LaserScan: ranges() holds what publishLaser would publish
// (simulator.cpp:393-426) and header() the message template (simulator.cpp:150-159).
// Commands arrive like the reference's ROS callbacks (DriveCallback stores them; Step consumes):
//   Ackermann  {velocity, curvature}          AckermannCurvatureDriveMsg.msg:5,11
//   diff drive {linear.x, angular.z}          geometry_msgs/Twist
//   omni       {velocity_x, velocity_y, velocity_r}
//
// Pose needs members `translation` (with x(), y()) and `angle` and a constructor
// Pose(angle, Vec2(x, y)) -- pose_2d::Pose2Df (poses_2d.h:32-104) on a ROS machine; Line a
// constructor Line(Vec2, Vec2).  Errors throw std::runtime_error (the reference exit(1)s).
#ifndef UMRS_ROBOT_MODEL_HPP_
#define UMRS_ROBOT_MODEL_HPP_

#include <stdexcept>
#include <string>
#include <vector>

#include "umrs.h"

namespace umrs {

template <class Pose, class Vec2, class Line>
class RobotModelT {
 public:
  RobotModelT(umrs_map* map, const umrs_robot_params& params, float dt, uint64_t seed = 0,
              bool laser_noise = false, bool motion_noise = false)
      : params_(params), dt_(dt) {
    umrs_world_desc d = umrs_world_desc();
    d.num_worlds = 1;
    d.robots_per_world = 1;
    d.robots = &params_;
    d.dt = dt;
    d.seed = seed;
    d.laser_noise_enabled = laser_noise ? 1 : 0;
    d.motion_noise_enabled = motion_noise ? 1 : 0;
    Check(umrs_world_create(map, &d, &world_), "umrs_world_create");
    Check(umrs_world_scan_header(world_, 0, &header_), "umrs_world_scan_header");
    int32_t n = 0;
    Check(umrs_world_object_lines_per_world(world_, &n), "umrs_world_object_lines_per_world");
    body_lines_ = n;
    ranges_.assign(static_cast<size_t>(header_.num_rays), header_.range_max);
  }
  ~RobotModelT() { umrs_world_destroy(world_); }
  RobotModelT(const RobotModelT&) = delete;
  RobotModelT& operator=(const RobotModelT&) = delete;

  // the reference's drive callbacks: the newest command wins, Step() consumes it
  void DriveCallback(float a, float b, float c = 0.f) {
    cmd_[0] = a; cmd_[1] = b; cmd_[2] = c;
    have_cmd_ = true;
  }

  // EntityBase::Step (entity_base.h:47): one delta_t of the model -- and the scan from the new pose
  void Step(const double& dt) {
    if (!(static_cast<float>(dt) == dt_)) throw std::runtime_error("RobotModelT::Step: dt differs from the one it was built with");
    const uint8_t mask = have_cmd_ ? 1 : 0;
    Check(umrs_world_step(world_, cmd_, &mask, ranges_.data(), nullptr), "umrs_world_step");
    have_cmd_ = false;
  }

  void SetPose(const Pose& pose) {   // entity_base.h:49
    const float p[3] = {pose.translation.x(), pose.translation.y(), pose.angle};
    float q[3], v[3];
    Check(umrs_world_get_robot_state(world_, q, v), "umrs_world_get_robot_state");
    Check(umrs_world_set_robot_state(world_, p, v), "umrs_world_set_robot_state");
  }
  Pose GetPose() {                   // entity_base.h:51
    float p[3], v[3];
    Check(umrs_world_get_robot_state(world_, p, v), "umrs_world_get_robot_state");
    return Pose(p[2], Vec2(p[0], p[1]));
  }
  void SetVel(const Pose& vel) {     // robot_model.h:36
    float p[3], v[3];
    Check(umrs_world_get_robot_state(world_, p, v), "umrs_world_get_robot_state");
    const float nv[3] = {vel.translation.x(), vel.translation.y(), vel.angle};
    Check(umrs_world_set_robot_state(world_, p, nv), "umrs_world_set_robot_state");
  }
  Pose GetVel() {                    // robot_model.h:37
    float p[3], v[3];
    Check(umrs_world_get_robot_state(world_, p, v), "umrs_world_get_robot_state");
    return Pose(v[2], Vec2(v[0], v[1]));
  }
  // EntityBase::GetLines (entity_base.h:53): the body rectangle at the current pose (the
  // reference's robots have no lines of their own; empty when footprint_enabled is 0)
  std::vector<Line> GetLines() {
    std::vector<Line> out;
    if (body_lines_ <= 0) return out;
    std::vector<float> flat(static_cast<size_t>(body_lines_) * 4);
    Check(umrs_world_get_object_lines(world_, flat.data()), "umrs_world_get_object_lines");
    for (int i = 0; i < body_lines_; ++i)
      out.push_back(Line(Vec2(flat[4 * i], flat[4 * i + 1]), Vec2(flat[4 * i + 2], flat[4 * i + 3])));
    return out;
  }

  const std::vector<float>& ranges() const { return ranges_; }     // LaserScan::ranges of the last Step
  const umrs_laser_scan_header& header() const { return header_; }
  umrs_world* handle() const { return world_; }

 private:
  static void Check(int status, const char* where) {
    if (status == UMRS_OK) return;
    throw std::runtime_error(std::string(where) + ": " + umrs_status_string(status) + " [" +
                             umrs_last_cuda_error() + "]");
  }
  umrs_robot_params params_;
  float dt_;
  umrs_world* world_ = nullptr;
  umrs_laser_scan_header header_;
  int32_t body_lines_ = 0;
  float cmd_[3] = {0.f, 0.f, 0.f};
  bool have_cmd_ = false;
  std::vector<float> ranges_;
};

}  // namespace umrs
#endif  // UMRS_ROBOT_MODEL_HPP_
