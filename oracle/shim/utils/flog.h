// TEST INFRASTRUCTURE ONLY (oracle).  Silent stand-in for the reference's
// logger (core/src/utils/flog.h) so the header-only DSP blocks build without
// flog.cpp; the hot path never depends on log output.
#pragma once
#include <string>
namespace flog {
    template <typename... Args> inline void debug(const char*, Args...) {}
    template <typename... Args> inline void info(const char*, Args...) {}
    template <typename... Args> inline void warn(const char*, Args...) {}
    template <typename... Args> inline void error(const char*, Args...) {}
    template <typename... Args> inline void exception(const char*, Args...) {}
}
