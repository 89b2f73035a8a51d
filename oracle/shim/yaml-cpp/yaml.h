// TEST INFRASTRUCTURE - stub of yaml-cpp for compiling the reference's feature/orb_params.{h,cc} unmodified.
// The oracle never builds parameters from YAML; a Node therefore always yields the caller's default.
#pragma once
#include <string>
#include <vector>
namespace YAML {
class Node {
public:
    Node operator[](const std::string&) const { return Node(); }
    Node operator[](const char*) const { return Node(); }
    template<typename T>
    T as(const T& fallback) const { return fallback; }
};
}  // namespace YAML
