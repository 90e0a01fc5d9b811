// Generic "FIR job" machinery shared by the VFO bank, the front-end decimator and the demodulators.
//
// A job is one streaming filter stage of the reference (DecimatingFIR, PolyphaseResampler or FIR)
// with its carried history.  Its input is the concatenation [history (H = taps-1 samples) | data
// (nIn samples)], exactly the reference's work buffer (filter/fir.h:25-27,66), but held as two
// segments so no memcpy/memmove of the block is needed; the history is double-buffered and
// refreshed by one kernel per process call.
#pragma once
#include "common.cuh"

namespace b200 {

enum JobKind : int { JOB_DECIM_FIR = 0, JOB_POLYPHASE = 1 };

// Device-visible description of one job for one process call (uploaded per call).
struct FirJob {
    const void* hist;      // H samples carried from previous calls
    const void* data;      // nIn new samples
    void* out;             // nOut results
    void* histNext;        // H samples to carry into the next call
    const float* taps;     // DECIM_FIR: T taps; POLYPHASE: bank [L][rowStride]
    int H;                 // history length = taps-1 (tapsPerPhase-1 for polyphase)
    int T;                 // dot-product length (taps, or tapsPerPhase)
    int D;                 // decimation (DECIM_FIR); 1 for a plain FIR
    int offset;            // position of output 0 in [hist|data] coordinates (DecimatingFIR::offset / PolyphaseResampler::offset)
    int nIn, nOut;
    int kind;
    int L, M, phase0, rowStride;   // POLYPHASE only
    int elemFloats;        // 2 = complex samples, 1 = real samples
    int plan;              // how a CTA works on this job (firJobPlan): 0 = one output per thread;
                           // DECIM_FIR > 0: 8 outputs per thread, taps split over `plan` adjacent lanes;
                           // POLYPHASE: 8 = eight lanes per output, 2 = one output per thread by residue class (short
                           // banks), 3 = two or three outputs of a class per thread sharing their samples
};

// chooses FirJob::plan from the job's shape (host side, called by Stage::fillJob)
void firJobPlan(FirJob& j);

// Call on the jobs of one launch before they are uploaded: plan-table jobs fall back to one output per thread when the
// whole launch would have too few CTAs to fill the GPU (a single VFO: 512 fat CTAs measured slower than 2048 thin ones).
void firJobsBalance(FirJob* jobs, int nJobs);

// CTAs job j occupies along grid.x; sorting a launch's jobs by it (descending) lets launchFirJobs leave the empty CTAs out
int firJobTiles(const FirJob& j);
void firJobsSortByTiles(FirJob* jobs, int nJobs);

// Launchers (all jobs of one "level" in a single launch; grid.y = job index)
// every job of a launch must have the same elemFloats; h_jobs is the host copy of d_jobs (grid and shared-memory sizing)
int launchFirJobs(const FirJob* d_jobs, const FirJob* h_jobs, int nJobs, int elemFloats, cudaStream_t s);
int launchHistoryUpdate(const FirJob* d_jobs, int nJobs, int maxH, cudaStream_t s);

} // namespace b200
