// TEST INFRASTRUCTURE - stub of yaml-cpp for compiling the reference's feature/orb_params.{h,cc}, camera/base.{h,cc} and
// camera/perspective.{h,cc} unmodified. The oracle never builds parameters from YAML: a Node is always undefined, yields
// the caller's default (or a value-initialised T) and converts to false.
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
    template<typename T>
    T as() const { return T(); }
    explicit operator bool() const { return false; }
    bool operator!() const { return true; }
};
}  // namespace YAML
