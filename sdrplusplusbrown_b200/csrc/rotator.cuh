// Rotator phase source shared by every kernel that applies FrequencyXlator (channel/frequency_xlator.h:43-56).
//
// The reference's phasor is VOLK's generic rotator: phase *= phaseDelta per sample in fp32,
// phase /= |phase| after every 512 samples of a call (and after a ragged tail).  All kernels here
// evaluate   phase[m] = chk[m >> 9] * wtab[m & 511]   where
//   wtab[j] = phaseDelta^j (complex power of the fp32-rounded step, evaluated in float64), j < 512
//   chk[c]  = the phase at sample 512*c of the current block ("checkpoint").
// Two ways to fill chk[]:
//   EXACT  one GPU thread per VFO replays the reference recursion bit for bit (IEEE fp32 mul/add/div, the
//          same hypotf) and records the phase at every renormalisation point.  Sequential: ~8 cycles/sample.
//   FAST   closed form chk[c] = exp(i*(angle0 + theta*512*c)) in float64, theta = the step calibrated once per
//          retune against the exact recursion (host_design.h calibrateRotatorStep).
// Everything downstream of chk[] is identical in both modes.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kRotChunk = 512;       // ROTATOR_RELOAD of VOLK's generic kernel
constexpr int kRotChunkLog = 9;

struct RotJob {
    float2* chk;          // [count/512 + 2]
    float2* state;        // exact mode: the reference's fp32 `phase` (read at start, written at end)
    double angle0, theta; // fast mode
    float dRe, dIm;       // phaseDelta (fp32, as the reference stores it)
    int count;
    int exact;
};

int launchCheckpoints(const RotJob* dJobs, int nJobs, int nExact, int maxCount, cudaStream_t s);
void makeRotatorTable(float dRe, float dIm, float2* out512);   // host

// phase at sample m >= 0 of the block
__device__ __forceinline__ float2 rotPhaseAt(const float2* __restrict__ chk, const float2* __restrict__ wtab, int m) {
    const float2 c = __ldg(chk + (m >> kRotChunkLog));
    const float2 w = __ldg(wtab + (m & (kRotChunk - 1)));
    return make_float2(c.x * w.x - c.y * w.y, c.x * w.y + c.y * w.x);
}

} // namespace b200
