// TEST INFRASTRUCTURE - logging calls of the reference compile to nothing in the oracle build.
#pragma once
namespace spdlog {
template<typename... A> inline void debug(const A&...) {}
template<typename... A> inline void info(const A&...) {}
template<typename... A> inline void warn(const A&...) {}
template<typename... A> inline void error(const A&...) {}
template<typename... A> inline void critical(const A&...) {}
template<typename... A> inline void trace(const A&...) {}
}  // namespace spdlog
