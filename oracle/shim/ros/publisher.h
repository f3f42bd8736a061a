// TEST INFRASTRUCTURE ONLY (oracle build).  See ros.h in this directory.
#include "ros/ros.h"
