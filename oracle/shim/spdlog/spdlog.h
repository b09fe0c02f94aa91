// TEST INFRASTRUCTURE.  Stand-in for the spdlog dependency of the reference's
// gpu/raytracer.cu (FetchContent in gpu/CMakeLists.txt:19-32, not fetchable offline): the
// three calls it makes (gpu/raytracer.cu:24-25,203) become stderr prints / no-ops.
#pragma once
#include <cstdio>
namespace spdlog {
template <class... A> inline void error(const char *fmt, A &&...) { std::fprintf(stderr, "[reference error] %s\n", fmt); }
template <class... A> inline void debug(const char *, A &&...) {}
template <class... A> inline void info(const char *, A &&...) {}
}  // namespace spdlog
