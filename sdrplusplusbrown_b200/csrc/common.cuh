// Shared declarations for libb200dsp (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/b200dsp.h"

namespace b200 {

struct cf32 { float re, im; };

void setError(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void countLaunch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

#define B200_CUDA(call)                                                                         \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            ::b200::setError("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return B200_ECUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define B200_CHECK_LAUNCH()                                                                     \
    do {                                                                                        \
        cudaError_t e__ = cudaGetLastError();                                                   \
        if (e__ != cudaSuccess) {                                                               \
            ::b200::setError("kernel launch failed at %s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return B200_ECUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define B200_TRY(expr)                         \
    do {                                       \
        int rc__ = (expr);                     \
        if (rc__ != B200_OK) { return rc__; }  \
    } while (0)

#ifdef __CUDACC__
// Blackwell's packed FP32 pair arithmetic (fma.rn.f32x2 -> SASS FFMA2): two IEEE FMAs in one issue slot.  A scalar
// multiplier is given as the pair (h, h); ptxas folds it into the instruction's broadcast operand form.
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b); rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
#endif

// RAII device selection
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) { ok = false; }
    }
    ~DeviceGuard() { if (prev >= 0) { cudaSetDevice(prev); } }
};

// Pinned host staging ring for per-call descriptor uploads (stream-ordered reuse).
struct PinnedRing {
    static constexpr int kSlots = 8;
    void* host[kSlots] = {};
    cudaEvent_t done[kSlots] = {};
    bool used[kSlots] = {};
    size_t bytes = 0;
    int next = 0;
    int init(size_t nbytes);
    void destroy();
    // returns a host slot that no in-flight copy still reads; *slot receives its index
    int acquire(size_t nbytes, void** hostPtr, int* slot);
    int release(int slot, cudaStream_t s);   // record the slot's event after the H2D copy was enqueued
};

// Measurement hooks (bench.py's roofline entries): when enabled, the launches of a kernel family are bracketed by CUDA
// events on their launch stream.  Families: PROF_STAGE1 = the fused xlate + first-stage decimation kernel (the
// full-rate kernel of the channelizer), PROF_FFT = the window + FFT + log-power passes of one FFT-branch call,
// PROF_DC = the DC blocker pass.  `bytes` / `flops` are the ALGORITHMIC traffic and arithmetic of the launch.
enum ProfKind { PROF_STAGE1 = 0, PROF_FFT = 1, PROF_DC = 2, PROF_KINDS = 3 };
bool profEnabled();
int profBegin(int kind, cudaStream_t s, double bytes, double flops, size_t* slot);
int profEnd(int kind, size_t slot, cudaStream_t s);

// Small host -> device transfer of per-call descriptors.  viaParams = false: cudaMemcpyAsync from the pinned slot.
// viaParams = true: the bytes ride in the parameter block of a tiny kernel that stores them (several launches for
// more than 3.9 KB), so the transfer goes through the stream's command buffer instead of the copy engine -- the
// pipelined host path uses this while the copy engine is busy with the block copies queued ahead.
int uploadDesc(void* dDst, const void* hSrc, size_t bytes, cudaStream_t s, bool viaParams);

} // namespace b200
