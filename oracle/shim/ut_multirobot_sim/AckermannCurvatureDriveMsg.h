// TEST INFRASTRUCTURE ONLY (oracle build): msg/AckermannCurvatureDriveMsg.msg as the
// struct genmsg would emit (Header header; float32 velocity; float32 curvature).
#ifndef UMRS_ORACLE_ACKERMANN_MSG_SHIM
#define UMRS_ORACLE_ACKERMANN_MSG_SHIM
#include "std_msgs/Header.h"
namespace ut_multirobot_sim {
struct AckermannCurvatureDriveMsg {
  std_msgs::Header header;
  float velocity;
  float curvature;
  AckermannCurvatureDriveMsg() : velocity(0.f), curvature(0.f) {}
};
}  // namespace ut_multirobot_sim
#endif
