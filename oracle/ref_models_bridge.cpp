// TEST INFRASTRUCTURE ONLY -- never linked into or called by the product.
//
// extern "C" bridge over the UNMODIFIED reference kinematic models, compiled in
// place from /root/reference into _ref/libumrs_ref.so (oracle/Makefile):
//   src/simulator/ackermann_model.cpp         ackermann::AckermannModel
//   src/simulator/diff_drive_model.cpp        diffdrive::DiffDriveModel
//   src/simulator/omnidirectional_model.cpp   omnidrive::OmnidirectionalModel
//   src/simulator/human_object.cpp            human::HumanObject
//   src/simulator/robot_model.cpp             robot_model::RobotModel
// against the stand-ins in oracle/shim for the headers this container lacks
// (roscpp, tf, the generated message structs, config_reader, Eigen).  The
// stand-ins only carry data: ConfigReader slots are filled from here instead of
// a Lua file, and NodeHandle::subscribe keeps each model's private DriveCallback
// so that a drive message is delivered to it the way a ROS spin would.
//
// Nothing here restates reference logic.  It pins the oracle's restated Step()s
// (umrs_oracle.c: ackermann_step, diff_drive_step, omni_step, orc_human_step).

#include <cstdint>
#include <cstring>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "config_reader/config_reader.h"
#include "simulator/ackermann_model.h"
#include "simulator/diff_drive_model.h"
#include "simulator/human_object.h"
#include "simulator/omnidirectional_model.h"

using Eigen::Vector2f;
using pose_2d::Pose2Df;

namespace {

struct ModelBox {
  int type;   // umrs_robot_params::model_type = the reference's RobotType order
              // (simulator.h:58-62): 0 Ackermann, 1 omnidirectional, 2 diff drive
  ros::NodeHandle nh;
  std::unique_ptr<robot_model::RobotModel> model;
};

const std::vector<std::string> kNoFiles;

}  // namespace

extern "C" {

// ---- config slots (what ConfigReader would have read from config/*.lua) ----------
void umrs_ref_config_set_float(const char* key, float v) { config_reader::Slot<float>(key) = v; }
void umrs_ref_config_set_int(const char* key, int v) { config_reader::Slot<int>(key) = v; }
void umrs_ref_config_set_bool(const char* key, int v) { config_reader::Slot<bool>(key) = (v != 0); }
void umrs_ref_config_set_string(const char* key, const char* v) {
  config_reader::Slot<std::string>(key) = v;
}

// ---- robot models: MakeMotionModel, simulator.cpp:128-142 ----------------------
void* umrs_ref_model_create(int type) {
  ModelBox* b = new ModelBox();
  b->type = type;
  if (type == 0) b->model.reset(new ackermann::AckermannModel(kNoFiles, &b->nh));
  else if (type == 2) b->model.reset(new diffdrive::DiffDriveModel(kNoFiles, &b->nh, ""));
  else b->model.reset(new omnidrive::OmnidirectionalModel(kNoFiles, &b->nh));
  return b;
}
void umrs_ref_model_free(void* h) { delete static_cast<ModelBox*>(h); }

void umrs_ref_model_set_pose(void* h, const float* xyt) {
  static_cast<ModelBox*>(h)->model->SetPose(Pose2Df(xyt[2], Vector2f(xyt[0], xyt[1])));
}
void umrs_ref_model_get_pose(void* h, float* xyt) {
  const Pose2Df p = static_cast<ModelBox*>(h)->model->GetPose();
  xyt[0] = p.translation.x(); xyt[1] = p.translation.y(); xyt[2] = p.angle;
}
void umrs_ref_model_set_vel(void* h, const float* xyt) {
  static_cast<ModelBox*>(h)->model->SetVel(Pose2Df(xyt[2], Vector2f(xyt[0], xyt[1])));
}
void umrs_ref_model_get_vel(void* h, float* xyt) {
  const Pose2Df p = static_cast<ModelBox*>(h)->model->GetVel();
  xyt[0] = p.translation.x(); xyt[1] = p.translation.y(); xyt[2] = p.angle;
}

// One drive message through the model's own (private) DriveCallback.
//   Ackermann {velocity, curvature}; diff drive {linear.x, angular.z};
//   omnidirectional {velocity_x, velocity_y, velocity_r}
void umrs_ref_model_command(void* h, const float* cmd) {
  ModelBox* b = static_cast<ModelBox*>(h);
  if (b->type == 0) {
    ut_multirobot_sim::AckermannCurvatureDriveMsg m;
    m.velocity = cmd[0];
    m.curvature = cmd[1];
    b->nh.deliver(&m);
  } else if (b->type == 2) {
    geometry_msgs::Twist m;
    m.linear.x = cmd[0];
    m.angular.z = cmd[1];
    b->nh.deliver(&m);
  } else {
    ut_multirobot_sim::CobotDriveMsg m;
    m.velocity_x = cmd[0];
    m.velocity_y = cmd[1];
    m.velocity_r = cmd[2];
    b->nh.deliver(&m);
  }
}

void umrs_ref_model_step(void* h, double dt) { static_cast<ModelBox*>(h)->model->Step(dt); }

// The stream AckermannModel's own members would produce (ackermann_model.h:17-18:
// a default-seeded std::default_random_engine behind normal_distribution<float>(0, 1)),
// so that a test can feed the oracle's explicit noise arguments the same draws.
void umrs_ref_normal_stream(int n, float* out) {
  std::default_random_engine rng;
  std::normal_distribution<float> d(0, 1);
  for (int i = 0; i < n; ++i) out[i] = d(rng);
}

// ---- human: human_object.cpp ----------------------------------------------------
void* umrs_ref_human_create() { return new human::HumanObject(kNoFiles); }
void umrs_ref_human_free(void* h) { delete static_cast<human::HumanObject*>(h); }
void umrs_ref_human_step(void* h, double dt) { static_cast<human::HumanObject*>(h)->Step(dt); }
void umrs_ref_human_set_vel(void* h, float vx, float vy, double rot) {
  static_cast<human::HumanObject*>(h)->SetVel(Vector2f(vx, vy), rot);
}
void umrs_ref_human_set_pose(void* h, const float* xyt) {
  static_cast<human::HumanObject*>(h)->SetPose(Pose2Df(xyt[2], Vector2f(xyt[0], xyt[1])));
}
void umrs_ref_human_get_pose(void* h, float* xyt) {
  const Pose2Df p = static_cast<human::HumanObject*>(h)->GetPose();
  xyt[0] = p.translation.x(); xyt[1] = p.translation.y(); xyt[2] = p.angle;
}
static int DumpHumanLines(const std::vector<geometry::Line2f>& lines, float* out, int cap) {
  const int n = static_cast<int>(lines.size());
  for (int i = 0; i < n && i < cap; ++i) {
    out[4 * i + 0] = lines[i].p0.x(); out[4 * i + 1] = lines[i].p0.y();
    out[4 * i + 2] = lines[i].p1.x(); out[4 * i + 3] = lines[i].p1.y();
  }
  return n;
}
int umrs_ref_human_lines(void* h, float* out, int cap) {
  return DumpHumanLines(static_cast<human::HumanObject*>(h)->GetLines(), out, cap);
}
int umrs_ref_human_template(void* h, float* out, int cap) {
  return DumpHumanLines(static_cast<human::HumanObject*>(h)->GetTemplateLines(), out, cap);
}

}  // extern "C"
