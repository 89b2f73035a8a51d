// Drop-in body of openvslam::match::robust::brute_force_match (reference: match/robust.h:19-30, match/robust.cc:215-321).
// The reference's robust.h, data/frame.h, data/keyframe.h and data/landmark.h are used UNCHANGED. Build: compile this file and
// compile the reference's match/robust.cc with -DLPSLAM_CUDA_BRUTE_FORCE, with `#ifndef LPSLAM_CUDA_BRUTE_FORCE` / `#endif`
// put around its definition of brute_force_match: match_for_triangulation, match_frame_and_keyframe
// and check_epipolar_constraint stay the reference's.
#include "openvslam/data/frame.h"
#include "openvslam/data/keyframe.h"
#include "openvslam/data/landmark.h"
#include "openvslam/match/robust.h"

#include "lpslam_cuda.h"

#include <stdexcept>
#include <string>

namespace openvslam {
namespace match {

unsigned int robust::brute_force_match(data::frame& frm, data::keyframe* keyfrm, std::vector<std::pair<int, int>>& matches) {
    const unsigned int n1 = frm.num_keypts_, n2 = keyfrm->num_keypts_;
    const auto lms_2 = keyfrm->get_landmarks();
    std::vector<uint8_t> valid_2(n2);  // robust.cc:240-246: rows without a live landmark are skipped
    for (unsigned int i = 0; i < n2; ++i) valid_2[i] = lms_2.at(i) && !lms_2.at(i)->will_be_erased();
    std::vector<float> ang_1, ang_2;  // only read when check_orientation_ (robust.cc:293-297 uses keypts_, not undist_keypts_)
    if (check_orientation_) {
        ang_1.resize(n1);
        ang_2.resize(n2);
        for (unsigned int i = 0; i < n1; ++i) ang_1[i] = frm.keypts_.at(i).angle;
        for (unsigned int i = 0; i < n2; ++i) ang_2[i] = keyfrm->keypts_.at(i).angle;
    }
    // descriptor matrices are n x 32 CV_8U; a matrix with padded rows is copied into a dense block first
    const cv::Mat d1 = frm.descriptors_.step == 32 ? frm.descriptors_ : frm.descriptors_.clone();
    const cv::Mat d2 = keyfrm->descriptors_.step == 32 ? keyfrm->descriptors_ : keyfrm->descriptors_.clone();
    std::vector<int32_t> match_2_in_1(n1, -1);
    uint32_t n = 0;
    if (lpslam_match_bruteforce(d1.data, n1, d2.data, n2, valid_2.data(), check_orientation_ ? ang_1.data() : nullptr,
                                check_orientation_ ? ang_2.data() : nullptr, lowe_ratio_, check_orientation_, match_2_in_1.data(), &n) < 0)
        throw std::runtime_error(std::string("liblpslam_cuda: lpslam_match_bruteforce: ") + lpslam_cuda_last_error());
    matches.clear();  // robust.cc:310-318: pairs in order of the frame index
    matches.reserve(n);
    for (unsigned int i = 0; i < n1; ++i)
        if (match_2_in_1[i] >= 0) matches.emplace_back(std::make_pair(static_cast<int>(i), match_2_in_1[i]));
    return n;
}

} // namespace match
} // namespace openvslam
