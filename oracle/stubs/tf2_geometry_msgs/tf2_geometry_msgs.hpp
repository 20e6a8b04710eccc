// TEST INFRASTRUCTURE ONLY: forwards to the one stub header (see ros_stub_core.hpp).
#include "ros_stub_core.hpp"
