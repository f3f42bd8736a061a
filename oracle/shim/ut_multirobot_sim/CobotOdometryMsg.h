// TEST INFRASTRUCTURE ONLY (oracle build): msg/CobotOdometryMsg.msg (float32 fields, uint8 status).
#ifndef UMRS_ORACLE_COBOT_ODOM_MSG_SHIM
#define UMRS_ORACLE_COBOT_ODOM_MSG_SHIM
#include <stdint.h>
#include "std_msgs/Header.h"
namespace ut_multirobot_sim {
struct CobotOdometryMsg {
  std_msgs::Header header;
  float dr, dx, dy, v0, v1, v2, v3, vr, vx, vy, VBatt;
  uint8_t status;
};
}  // namespace ut_multirobot_sim
#endif
