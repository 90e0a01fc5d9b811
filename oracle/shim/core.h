// TEST INFRASTRUCTURE ONLY (oracle).  Shadows the reference's core/src/core.h,
// which drags in the config/module/CLI managers; dsp/block.h:9 only needs
// SetThreadName (core.h:43-49).
#pragma once
#include <string>
#ifdef __linux__
#include <sys/prctl.h>
#endif
inline void SetThreadName(const std::string& threadName) {
#ifdef __linux__
    prctl(PR_SET_NAME, threadName.c_str(), 0, 0, 0);
#endif
}
