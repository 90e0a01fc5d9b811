// Windowed forward FFT -> log power, batched over frames (K8+K9+K10 of SURVEY §2.2).
#pragma once
#include <vector>

#include "common.cuh"

namespace b200 {

extern int fftCtasPerSm;
extern int fftDirectMaxLog;
extern int fftDirectPrefetch;
extern int branchOrder;    // frontend.cu
extern int hostChunkMiB;   // frontend.cu

// One frame of the Reshaper (keep nz samples every I): the samples may straddle the carry buffer
// and the current block, so a frame is [src0[0..len0) | src1[0..nz-len0)].
struct FrameDesc {
    const cf32* src0;
    const cf32* src1;
    int len0;
    int pad;
};

struct FftPlan {
    int device = 0;
    int N = 0, logN = 0;
    int N1 = 1, N2 = 0;          // N = N1 * N2; N1 == 1 -> single pass
    bool direct = false;         // single pass by fftDirectKernel (three register-resident stages)
    float2* dTwN1 = nullptr;     // W_N1^j, j < N1/2
    float2* dTwN2 = nullptr;     // W_N2^j, j < N2/2
    float2* dTwLo = nullptr;     // W_N^j,      j < S      (inter-pass twiddle, two-level)
    float2* dTwHi = nullptr;     // W_N^(j*S),  j < N/S
    int S = 0, logS = 0;
    cf32* dScratch = nullptr;    // pass-1 output, frames x N (L2-resident ring)
    int scratchFrames = 0;

    int create(int N, int device);
    void destroy();
    int ensureScratch(int frames);
    // frames: device array of FrameDesc; window: device float[nz] (may be NULL = no window);
    // outDb: device float[nFrames][N] (may be NULL), outSpec: device cf32[nFrames][N] (may be NULL)
    int run(const FrameDesc* dFrames, int nFrames, const float* dWindow, int nz, float* outDb, cf32* outSpec, cudaStream_t s);
    int runPasses(const FrameDesc* dFrames, int nFrames, const float* dWindow, int nz, float* outDb, cf32* outSpec, cudaStream_t s);
};

} // namespace b200
