// TEST INFRASTRUCTURE - the two container types of FBoW (3rd/FBoW/include/fbow/fbow.h) that the matchers iterate:
// BoWFeatVector = node id -> keypoint indices, BoWVector = word id -> weight, both ordered maps.
#pragma once
#include <cstdint>
#include <map>
#include <vector>
namespace fbow {
struct BoWVector : std::map<uint32_t, float> {};
struct BoWFeatVector : std::map<uint32_t, std::vector<uint32_t>> {};
}  // namespace fbow
