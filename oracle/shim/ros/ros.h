// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for the slice of roscpp the
// reference's kinematic models touch: NodeHandle::subscribe keeps the (private)
// member callback so that the bridge can deliver a drive message the way a ROS
// spin would; advertise/publish keep nothing.
//
// <math.h> is included on purpose: roscpp's own headers pull it in
// (ros/duration.h), which is what makes the models' unqualified sin/cos/fabs
// calls on floats resolve to the float overloads under libstdc++.
#ifndef UMRS_ORACLE_ROS_SHIM
#define UMRS_ORACLE_ROS_SHIM
#include <math.h>
#include <stdint.h>

#include <functional>
#include <string>

namespace ros {

struct Time {
  double sec;
  Time() : sec(0.0) {}
  static Time now() { return Time(); }
};

class Subscriber {};

class Publisher {
 public:
  template <class M> void publish(const M&) const {}
};

class NodeHandle {
 public:
  // the newest subscription's callback, type-erased: deliver(&msg)
  std::function<void(const void*)> deliver;
  std::string topic;

  template <class M, class T>
  Subscriber subscribe(const std::string& name, uint32_t, void (T::*fp)(const M&), T* obj) {
    topic = name;
    deliver = [fp, obj](const void* m) { (obj->*fp)(*static_cast<const M*>(m)); };
    return Subscriber();
  }
  template <class M>
  Publisher advertise(const std::string&, uint32_t, bool = false) { return Publisher(); }
};

}  // namespace ros
#endif
