// TEST INFRASTRUCTURE - stand-in for data/map_database.h: data/landmark.cc only calls erase_landmark (prepare_for_erasing / replace).
#pragma once
#include <set>
namespace openvslam {
namespace data {
class landmark;
class map_database {
public:
    void erase_landmark(landmark* lm) { erased_.insert(lm); }
    std::set<landmark*> erased_;
};
}  // namespace data
}  // namespace openvslam
