// TEST INFRASTRUCTURE ONLY (oracle build): nav_msgs/Odometry as plain structs.
#ifndef UMRS_ORACLE_NAV_MSGS_ODOMETRY_SHIM
#define UMRS_ORACLE_NAV_MSGS_ODOMETRY_SHIM
#include <string>
#include "geometry_msgs/Twist.h"
#include "std_msgs/Header.h"
namespace geometry_msgs {
struct Pose { Point position; Quaternion orientation; };
struct PoseWithCovariance { Pose pose; double covariance[36]; };
struct TwistWithCovariance { Twist twist; double covariance[36]; };
}  // namespace geometry_msgs
namespace nav_msgs {
struct Odometry {
  std_msgs::Header header;
  std::string child_frame_id;
  geometry_msgs::PoseWithCovariance pose;
  geometry_msgs::TwistWithCovariance twist;
};
}  // namespace nav_msgs
#endif
