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
    const cf32* hist;      // T-1 rotated samples carried from the previous call
    cf32* histNext;        // T-1 rotated samples to carry
    const float2* chk;     // rotator checkpoints of this block (rotator.cuh)
    const float2* wtab;    // phaseDelta^j, j < 512
    const float2* hWtab;   // HOST copy of that table (read by the launcher only; may be NULL: tables then come from `wtab`)
    double thetaEff, logMag;   // arg / log|.| of phaseDelta, for the backward extension into the history region
    int ratio;             // pow-2 plan (selects D, T and the taps at compile time)
    int offset;            // DecimatingFIR::offset
    int count, nOut;
};

extern bool fusedEnabled;   // test hook: false forces the unfused xlate -> generic FIR path
extern int fusedTileSamples; // tuning hook: 0 = by VFO count, else input samples per CTA tile
extern int fusedVfosPerPass; // tuning hook: 0 = by code size, else VFOs accumulated per register pass of the grouped kernel
extern bool fusedLeaveRoom;  // the next first-stage launches keep part of every SM free for the deferred tail of the previous block
extern bool fusedParamTables; // rotation tables in the kernel parameter block, read through uniform loads (B200_OPT_PARAM_TABLES)
extern bool fusedSplitRuns;  // tuning hook: split a plan whose loop body exceeds the instruction cache into two launches
extern int fusedVfosPerCta;  // tuning hook: VFOs a CTA of the grouped kernel serves from one staged tile
bool fusedStageSupported(const Vfo& v);
void fillFusedJob(FusedJob& f, const Stage& st, const Vfo& v, const cf32* dIn, int count);
int launchFused(const FusedJob* dJobs, const std::vector<FusedJob>& hostJobs, cudaStream_t s);

} // namespace b200
