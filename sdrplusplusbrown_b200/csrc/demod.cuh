// Demodulators attached to VFO outputs (FM / AM / SSB + AGC), stage 2 of the path.
#pragma once
#include "common.cuh"

namespace b200 {
struct VfoBank;
struct DemodState;
void demodDestroy(DemodState* d);
int processDemods(VfoBank& bank, cudaStream_t s);
// carries an SSB demodulator's internal rotator phase across a FAST <-> EXACT switch
int demodSwitchRotatorMode(DemodState* d, int newMode);
} // namespace b200
