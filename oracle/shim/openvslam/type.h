// TEST INFRASTRUCTURE - stands in for openvslam/type.h (which pulls in Eigen): feature/orb_extractor.cc includes it but
// uses nothing from it except M_PI via <cmath>.
#pragma once
#include <cmath>
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
