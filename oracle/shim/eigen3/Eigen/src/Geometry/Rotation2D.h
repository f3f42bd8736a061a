// TEST INFRASTRUCTURE ONLY (oracle build).  The models include this internal Eigen path.
#include "../../Geometry"
