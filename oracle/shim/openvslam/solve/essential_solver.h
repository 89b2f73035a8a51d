// TEST INFRASTRUCTURE - stand-in for solve/essential_solver.h: match/robust.cc includes it for match_frame_and_keyframe
// (RANSAC over the brute-force matches, :180-213), which the oracle never calls. The stub reports "no valid solution".
#pragma once
#include "openvslam/type.h"
#include <utility>
#include <vector>
namespace openvslam {
namespace solve {
class essential_solver {
public:
    essential_solver(const eigen_alloc_vector<Vec3_t>&, const eigen_alloc_vector<Vec3_t>&, const std::vector<std::pair<int, int>>&) {}
    void find_via_ransac(unsigned int, bool = true) {}
    bool solution_is_valid() const { return false; }
    std::vector<bool> get_inlier_matches() const { return {}; }
};
}  // namespace solve
}  // namespace openvslam
