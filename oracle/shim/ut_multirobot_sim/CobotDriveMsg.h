// TEST INFRASTRUCTURE ONLY (oracle build): msg/CobotDriveMsg.msg (float32 fields).
#ifndef UMRS_ORACLE_COBOT_DRIVE_MSG_SHIM
#define UMRS_ORACLE_COBOT_DRIVE_MSG_SHIM
#include "std_msgs/Header.h"
namespace ut_multirobot_sim {
struct CobotDriveMsg {
  std_msgs::Header header;
  float velocity_x, velocity_y, velocity_r;
  float transMaxAcceleration, transMaxDeceleration, transMaxVelocity;
  float rotMaxAcceleration, rotMaxDeceleration, rotMaxVelocity;
  CobotDriveMsg()
      : velocity_x(0.f), velocity_y(0.f), velocity_r(0.f), transMaxAcceleration(0.f),
        transMaxDeceleration(0.f), transMaxVelocity(0.f), rotMaxAcceleration(0.f),
        rotMaxDeceleration(0.f), rotMaxVelocity(0.f) {}
};
}  // namespace ut_multirobot_sim
#endif
