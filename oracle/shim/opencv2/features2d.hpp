// TEST INFRASTRUCTURE - forwards to the minimal cv:: shim (see opencv2/core.hpp)
#pragma once
#include <opencv2/core.hpp>
