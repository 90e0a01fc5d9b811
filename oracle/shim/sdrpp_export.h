// TEST INFRASTRUCTURE ONLY (oracle).  Shadows the CMake-generated export header.
#pragma once
#define SDRPP_EXPORT extern
