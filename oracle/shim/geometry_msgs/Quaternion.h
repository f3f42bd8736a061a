// TEST INFRASTRUCTURE ONLY (oracle build): geometry_msgs/Quaternion (float64 fields).
#ifndef UMRS_ORACLE_GEOMETRY_MSGS_QUATERNION_SHIM
#define UMRS_ORACLE_GEOMETRY_MSGS_QUATERNION_SHIM
namespace geometry_msgs {
struct Quaternion {
  double x, y, z, w;
  Quaternion() : x(0), y(0), z(0), w(0) {}
};
struct Vector3 {
  double x, y, z;
  Vector3() : x(0), y(0), z(0) {}
};
typedef Vector3 Point;
}  // namespace geometry_msgs
#endif
