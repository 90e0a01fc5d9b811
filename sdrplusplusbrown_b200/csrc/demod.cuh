// Demodulators attached to VFO outputs (FM / AM / SSB + AGC), stage 2 of the path.
#pragma once
#include "common.cuh"

namespace b200 {
struct VfoBank;
struct DemodState;
void demodDestroy(DemodState* d);
int processDemods(VfoBank& bank, cudaStream_t s);
} // namespace b200
