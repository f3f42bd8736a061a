// TEST INFRASTRUCTURE ONLY (oracle build): the one tf helper diff_drive_model.cpp calls.
#ifndef UMRS_ORACLE_TF_SHIM
#define UMRS_ORACLE_TF_SHIM
#include <cmath>
#include "geometry_msgs/Quaternion.h"
namespace tf {
inline geometry_msgs::Quaternion createQuaternionMsgFromYaw(double yaw) {
  geometry_msgs::Quaternion q;
  q.z = std::sin(yaw * 0.5);
  q.w = std::cos(yaw * 0.5);
  return q;
}
}  // namespace tf
#endif
