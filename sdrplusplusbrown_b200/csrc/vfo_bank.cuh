// RxVFO bank: host-side state mirror + launch orchestration (device kernels in xlate.cu / fir_jobs.cu / demod.cu).
#pragma once
#include <functional>
#include <memory>
#include <vector>

#include "common.cuh"
#include "fir_jobs.cuh"
#include "rotator.cuh"
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
    void* dOutAlt = nullptr;          // second output buffer of a first stage whose consumers run one block behind (deferred tail)
    int outCap = 0;

    bool planTaps = false;            // taps are one of the pow-2 plan tables (decim_tables.inc): the kernel may use its constant-bank copy
    int create(int kind, const std::vector<float>& taps, int D, int L, int M, int elemFloats, int outCap, bool planTaps = false);
    void destroy();
    int resetState();                                   // zero history, offset = phase = 0
    int setTapsKeepHistory(const std::vector<float>& taps);   // FIR::setTaps (fir.h:33-54)
    int ensureOutCap(int cap);
    int flipOut();                                      // swap dOut and dOutAlt (allocated on first use)
    int outCountFor(int nIn) const;
    void advance(int nIn);                              // integer state after a block of nIn
    void fillJob(FirJob& j, const void* data, int nIn) const;
};

// per-call rotation job (unfused path); phase = chk[m >> 9] * wtab[m & 511] (rotator.cuh)
struct XlateJob {
    const cf32* in;
    cf32* out;
    const float2* chk;
    const float2* wtab;
    int n;
    int pad;
};

struct DemodState;   // demod.cu

struct Vfo {
    bool live = false;
    double outSR = 0, bw = 0, offset = 0;
    // FrequencyXlator state (frequency_xlator.h:15-29,70-71)
    float dRe = 1.f, dIm = 0.f;      // phaseDelta as the reference stores it (fp32)
    double theta = 0.0;              // FAST mode: calibrated effective step
    double angle = 0.0;              // FAST mode: current phase angle (|phase| == 1)
    bool calDirty = true;
    bool tableDirty = true;          // wtab must be re-uploaded (phaseDelta changed)
    float2* dChk = nullptr; int chkCap = 0;   // per-block phase checkpoints
    float2* dWtab = nullptr;                  // phaseDelta^j, j < 512
    std::vector<float2> hWtab;                // the same table on the host (the bank kernel takes it in its parameter block)
    float2* dRotState = nullptr;              // EXACT mode: the reference's fp32 phase
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
    int rotatorMode = 0;             // B200_ROTATOR_FAST / B200_ROTATOR_EXACT
    void* dDemodDesc = nullptr; size_t dDemodDescBytes = 0;
    cudaStream_t lastStream = nullptr;
    bool everRan = false;
    cudaEvent_t evStage1 = nullptr;  // recorded behind the full-rate stage of every process() call (the front end starts its FFT branch there)
    bool descViaParams = false;      // descriptor uploads avoid the copy engine (uploadDesc); set by the pipelined host path
    // per-call descriptors
    PinnedRing ring;
    void* dDesc = nullptr; size_t dDescBytes = 0;
    // B200_OPT_DEFER_TAIL: the stages behind the full-rate one (and the demodulators) of block k run on tailStream beside
    // the full-rate stage of block k+1; the caller's stream joins them when it asks for an output pointer.
    cudaStream_t tailStream = nullptr;
    cudaEvent_t evTail[2] = { nullptr, nullptr };
    bool tailPending[2] = { false, false };
    int tailPar = 0, lastTailPar = 0;
    bool tailUnjoined = false;
    void* dTailDesc = nullptr; size_t dTailDescBytes = 0;
    int joinTail();                  // lastStream waits for the newest deferred tail
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
    int allocRotator(Vfo& v);
    void freeRotator(Vfo& v);
    int getPhase(Vfo& v, double* re, double* im);
    int setPhase(Vfo& v, double re, double im);
    int setRotatorMode(int mode);
    int reconfigureResampler(Vfo& v);            // RationalResampler::reconfigure
    int regenChannelTaps(Vfo& v, bool create);   // RxVFO::generateTaps + filter.setTaps
    int resetVfo(Vfo& v);
    int calibrateDirty();
    // afterStage1: called once, right behind the launch of the full-rate first-stage kernels (the front end enqueues its
    // FFT branch there when the branch order asks for it)
    int process(const cf32* dIn, int count, cudaStream_t s, const std::function<int()>* afterStage1 = nullptr);
    int maxOut(int v, int count) const;
};

extern bool deferTail;
extern bool fusedLeaveRoomEnabled;
int launchXlate(const XlateJob* dJobs, int nJobs, int maxN, cudaStream_t s);

} // namespace b200

// the opaque handle of include/b200dsp.h
struct b200_vfo_bank {
    b200::VfoBank impl;
    bool ownedByFrontend = false;
};
