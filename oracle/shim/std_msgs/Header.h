// TEST INFRASTRUCTURE ONLY (oracle build): std_msgs/Header as a plain struct.
#ifndef UMRS_ORACLE_STD_MSGS_HEADER_SHIM
#define UMRS_ORACLE_STD_MSGS_HEADER_SHIM
#include <stdint.h>
#include <string>
#include "ros/ros.h"
namespace std_msgs {
struct Header {
  uint32_t seq;
  ros::Time stamp;
  std::string frame_id;
  Header() : seq(0) {}
};
}  // namespace std_msgs
#endif
