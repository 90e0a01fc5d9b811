// RxVFO bank: host-side state mirror + launch orchestration (device kernels in xlate.cu / fir_jobs.cu / demod.cu).
#pragma once
#include <memory>
#include <vector>

#include "common.cuh"
#include "fir_jobs.cuh"
#include "host_design.h"

namespace b200 {

// One streaming filter stage with its carried state (device history ping-pong + integer state on host).
struct Stage {
    int kind = JOB_DECIM_FIR;
    int T = 0, H = 0, D = 1;
    int L = 1, M = 1, rowStride = 0;
    int offset = 0, phase = 0;        // DecimatingFIR::offset / PolyphaseResampler::{offset,phase}
    int elemFloats = 2;
    std::vector<float> hostTaps;      // prototype taps (for introspection)
    float* dTaps = nullptr;           // device taps (FIR) or bank (polyphase)
    void* dHist[2] = { nullptr, nullptr };
    int cur = 0;
    size_t histCapBytes = 0;
    void* dOut = nullptr;
    int outCap = 0;

    int create(int kind, const std::vector<float>& taps, int D, int L, int M, int elemFloats, int outCap);
    void destroy();
    int resetState();                                   // zero history, offset = phase = 0
    int setTapsKeepHistory(const std::vector<float>& taps);   // FIR::setTaps (fir.h:33-54)
    int ensureOutCap(int cap);
    int outCountFor(int nIn) const;
    void advance(int nIn);                              // integer state after a block of nIn
    void fillJob(FirJob& j, const void* data, int nIn) const;
};

// per-call rotation job (unfused path)
struct XlateJob {
    const cf32* in;
    cf32* out;
    double angle0;    // rotator angle at sample 0 of this block
    double theta;     // calibrated step per sample
    float stepRe, stepIm;
    int n;
    int pad;
};

struct DemodState;   // demod.cu

struct Vfo {
    bool live = false;
    double outSR = 0, bw = 0, offset = 0;
    // FrequencyXlator state (frequency_xlator.h:15-29,70-71)
    float dRe = 1.f, dIm = 0.f;      // phaseDelta as the reference stores it (fp32)
    double theta = 0.0;              // calibrated effective step
    double angle = 0.0;              // current phase angle (|phase| == 1)
    bool calDirty = true;
    // RationalResampler configuration
    int mode = 3, ratio = 1, interp = 1, decim = 1, resampTapCount = 0;
    std::vector<float> resampTaps;
    std::vector<Stage> decimStages;
    Stage poly;  bool hasPoly = false;
    Stage chan;  bool chanCreated = false; bool filterNeeded = false;
    cf32* dZ = nullptr; int zCap = 0;      // rotated block (unfused path)
    // last process call
    int lastOut = 0;
    const cf32* lastOutPtr = nullptr;
    DemodState* demod = nullptr;
};

struct VfoBank {
    int device = 0;
    double inSR = 0;
    int maxBlock = 0;
    std::vector<Vfo> vfos;
    int calSteps = 1 << 20;
    cudaStream_t lastStream = nullptr;
    bool everRan = false;
    // per-call descriptors
    PinnedRing ring;
    void* dDesc = nullptr; size_t dDescBytes = 0;
    // host staging for process_host
    cf32* hPinnedIn = nullptr; size_t hPinnedInCap = 0;
    cf32* dStageIn = nullptr; size_t dStageInCap = 0;
    void* hPinnedOut = nullptr; size_t hPinnedOutCap = 0;

    int init(double inSR, int maxBlock, int device);
    void destroy();
    int addVfo(double outSR, double bw, double offset);
    int removeVfo(int v);
    int liveCount() const;
    int check(int v) const;
    int syncLast();
    int configureXlator(Vfo& v);
    int reconfigureResampler(Vfo& v);            // RationalResampler::reconfigure
    int regenChannelTaps(Vfo& v, bool create);   // RxVFO::generateTaps + filter.setTaps
    int resetVfo(Vfo& v);
    int calibrateDirty();
    int process(const cf32* dIn, int count, cudaStream_t s);
    int maxOut(int v, int count) const;
};

int launchXlate(const XlateJob* dJobs, int nJobs, int maxN, cudaStream_t s);

} // namespace b200

// the opaque handle of include/b200dsp.h
struct b200_vfo_bank {
    b200::VfoBank impl;
    bool ownedByFrontend = false;
};
