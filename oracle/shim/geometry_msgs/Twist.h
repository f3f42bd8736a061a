// TEST INFRASTRUCTURE ONLY (oracle build): geometry_msgs/Twist (float64 fields).
#ifndef UMRS_ORACLE_GEOMETRY_MSGS_TWIST_SHIM
#define UMRS_ORACLE_GEOMETRY_MSGS_TWIST_SHIM
#include "geometry_msgs/Quaternion.h"
namespace geometry_msgs {
struct Twist {
  Vector3 linear, angular;
};
}  // namespace geometry_msgs
#endif
