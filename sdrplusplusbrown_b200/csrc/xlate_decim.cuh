// Fused frequency translation + first decimating FIR stage (K1+K2 of SURVEY §2.2), many VFOs per launch.
#pragma once
#include <vector>

#include "common.cuh"

namespace b200 {

struct Stage;
struct Vfo;

// per-call descriptor of one VFO's fused first stage
struct FusedJob {
    const cf32* in;        // shared input block (count samples)
    cf32* out;             // nOut results
    const cf32* hist;      // H rotated samples carried from the previous call
    cf32* histNext;        // H rotated samples to carry
    double angle0, theta;  // rotator angle at sample 0, calibrated step
    int planIndex;         // which pow-2 plan's first stage (selects taps/D/T at compile time)
    int offset;            // DecimatingFIR::offset
    int count, nOut;
};

bool fusedStageSupported(const Stage& st);
void fillFusedJob(FusedJob& f, const Stage& st, const Vfo& v, const cf32* dIn, int count);
int launchFused(const FusedJob* dJobs, const std::vector<FusedJob>& hostJobs, cudaStream_t s);

} // namespace b200
