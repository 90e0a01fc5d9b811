// dsp::compression::ExperimentalFFTCompressor on the device (the SDR++ server's sample stream).
//
// Reference: core/src/dsp/compression/experimental_fft_compressor.h -- setSampleRate :66-88 (frame length
// 2^floor(log2(sampleRate / 20)), Blackman-like window), process :198-427 (per frame: spectrum and windowed spectrum, halves
// swapped, power in dB; ten frames of delay; mean of the last <= 21 dB rows), filterSignal :116-177 (noise floor from
// moving variance / moving averages / hole interpolation / sampled percentiles, bins that do not rise above it zeroed),
// estimateNoise :181-189, amplitude -> fourth root :257-266; wired in core/src/server.cpp:101-110 between the resampler
// and the PCM sample-packet compressor (b200_stream_compress here).
//
// The two transforms of a frame are the power-of-two FFT passes of fft.cu.  The noise-floor tracker is a chain of running
// sums (centeredSma, core/src/utils/arrays.cpp:474-495) and hole interpolations (:433-470) whose fp32 rounding depends on
// the order of summation, with thresholds taken on the results: it runs in ONE thread block in the reference's order --
// element-wise steps on all threads, every running sum on one thread over differences staged in shared memory -- so that
// for the same dB rows every bin is kept or zeroed exactly as in the reference.  A frame is 27-55 ms of signal.  Transmit mode (setTxMode) is not served.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

#include "common.cuh"
#include "fft.cuh"

namespace b200 {

namespace {

constexpr int kMinRecents = 10;          // frames of delay (experimental_fft_compressor.h:48)
constexpr int kMagRows = 2 * kMinRecents + 1;
constexpr int kNoisePoints = 16;         // noiseNPoints
constexpr int kSignalWidth = 300;
constexpr int kTrackThreads = 1024;

__device__ __forceinline__ float log2Clamped(float v) {
    const float r = log2f(v);
    return isinf(r) ? copysignf(127.0f, r) : r;
}

// swapfft (utils/arrays.cpp:1128-1134) + volk_32fc_s32f_power_spectrum_32f (norm = n); freqOut may be NULL
__global__ void fcSpectrumKernel(const cf32* __restrict__ spec, cf32* __restrict__ freqOut, float* __restrict__ magOut, int n) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) { return; }
    const int half = n / 2;
    const cf32 v = spec[(j < half) ? j + half : j - half];
    if (freqOut) { freqOut[j] = v; }
    const float inv = 1.0f / (float)n;
    const float r = __fmul_rn(v.re, inv), q = __fmul_rn(v.im, inv);
    magOut[j] = __fmul_rn(3.01029995663981209120f, log2Clamped(__fadd_rn(__fmul_rn(r, r), __fmul_rn(q, q))));
}

// mean of the kept dB rows, oldest first (addeach from zeros, then times (float)(1.0 / rows))
__global__ void fcMeanKernel(const float* __restrict__ ring, int head, int count, int cap, float inv, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    float acc = 0.0f;
    int r = head;
    for (int t = 0; t < count; t++) {
        acc = __fadd_rn(acc, ring[(size_t)r * n + i]);
        r = (r + 1 == cap) ? 0 : r + 1;
    }
    out[i] = __fmul_rn(acc, inv);
}

constexpr int kSmaChunk = 8192;           // running-sum elements staged in shared memory at a time

// centeredSma (utils/arrays.cpp:474-495) by the whole block.  The running total is a sequential fp32 recurrence
// (total += in[i] - in[i - w], in that order), so ONE thread walks it -- but over shared memory: all threads first form the
// differences of a chunk (exact, element-wise), thread 0 adds them up leaving the total BEFORE each step in place, all
// threads divide and store.  in and out must not overlap.
__device__ void blockCenteredSma(const float* __restrict__ in, float* __restrict__ out, int limit, int winsize, float* sm, float* sTotal) {
    const int tid = threadIdx.x, win2 = winsize / 2;
    const float fw = (float)winsize;
    if (tid == 0) {
        float total = 0.0f;
        for (int i = 0; i < winsize; i++) { total = __fadd_rn(total, in[i]); }
        *sTotal = total;
    }
    __syncthreads();
    for (int c0 = winsize; c0 < limit; c0 += kSmaChunk) {
        const int m = min(kSmaChunk, limit - c0);
        for (int k = tid; k < m; k += kTrackThreads) { sm[k] = __fsub_rn(in[c0 + k], in[c0 + k - winsize]); }
        __syncthreads();
        if (tid == 0) {
            float total = *sTotal;
            for (int k = 0; k < m; k++) { const float before = total; total = __fadd_rn(total, sm[k]); sm[k] = before; }
            *sTotal = total;
        }
        __syncthreads();
        for (int k = tid; k < m; k += kTrackThreads) { out[c0 + k - win2] = __fdiv_rn(sm[k], fw); }
        __syncthreads();
    }
    const float head = out[winsize - win2], tail = out[limit - win2 - 1];
    __syncthreads();
    for (int i = tid; i < winsize - win2; i += kTrackThreads) { out[i] = head; }
    for (int i = limit - win2 + tid; i < limit; i += kTrackThreads) { out[i] = tail; }
    __syncthreads();
}

// linearInterpolateHoles (utils/arrays.cpp:433-470) by the whole block: src = the array with its zeros, dst receives the
// filled array.  Every kept element fills the gap to its left neighbour with the reference's running sum (sequential per
// gap, so one thread per gap); the first / last kept ones also fill the ends.  Nothing kept: dst = src.
__device__ void blockFillHoles(const float* __restrict__ src, float* __restrict__ dst, int n) {
    const int tid = threadIdx.x;
    for (int q = tid; q < n; q += kTrackThreads) { dst[q] = src[q]; }
    __syncthreads();                                     // (the gap fillers below write other threads' elements)
    for (int q = tid; q < n; q += kTrackThreads) {
        const float val = src[q];
        if (val == 0.0f) { continue; }
        int last = q - 1;
        while (last >= 0 && src[last] == 0.0f) { last--; }
        if (last < 0) { for (int w = q - 1; w >= 0; w--) { dst[w] = val; } }
        else if (q - last > 1) {
            const float d = __fdiv_rn(__fsub_rn(val, src[last]), (float)(q - last));
            float running = src[last];
            for (int w = last + 1; w < q; w++) { running = __fadd_rn(running, d); dst[w] = running; }
        }
        int next = q + 1;
        while (next < n && src[next] == 0.0f) { next++; }
        if (next >= n) { for (int w = q + 1; w < n; w++) { dst[w] = val; } }
    }
    __syncthreads();
}

// percentile::percentile_sampling (utils/strings.h:153-175) over kthSmallest / partition (:93-140), whose pivot is held in
// an int (`int x = arr[r]`); one thread, 100 sampled values
__device__ float seqPercentileSampling(const float* arr, int n, double p) {
    if (n == 0) { return 0.0f; }
    float s[200];
    int m = n;
    if (n > 200) {
        const float step = __fdiv_rn((float)(n - 1), 100.0f);
        for (int z = 0; z < 100; z++) { s[z] = arr[(int)floorf(__fmul_rn(step, (float)z))]; }
        m = 100;
    }
    else { for (int z = 0; z < n; z++) { s[z] = arr[z]; } }
    int k = (int)((double)(m - 1) * p);
    int l = 0, r = m - 1;
    for (;;) {
        if (!(k > 0 && k <= r - l + 1)) { return s[0]; }
        const int x = (int)s[r];
        int i = l;
        for (int j = l; j <= r - 1; j++) {
            if (s[j] <= (float)x) { const float t = s[i]; s[i] = s[j]; s[j] = t; i++; }
        }
        { const float t = s[i]; s[i] = s[r]; s[r] = t; }
        if (i - l == k - 1) { return s[i]; }
        if (i - l > k - 1) { r = i - 1; continue; }
        k = k - i + l - 1;
        l = i + 1;
    }
}

struct TrackArgs {
    const float* win;        // mean windowed dB row
    const float* clr;        // mean plain dB row
    cf32* frame;             // the frame going out (ten frames old), filtered and companded in place
    float* a; float* b; float* c; float* d; float* e;      // scratch, n floats each
    float* prevAllowance;    // carried between frames
    float* noiseFigure;      // 30 values
    const int* masked;       // pairs (from, to) in Hz
    int nMasked;
    int n, largeTick;
    float hzTick;
    double lossRate;
    int filter;              // lossRate > 0
};

// filterSignal :116-177 + estimateNoise :181-189 + the amplitude companding :257-266, one block
__global__ void __launch_bounds__(kTrackThreads) fcTrackKernel(TrackArgs t) {
    const int tid = threadIdx.x, n = t.n;
    __shared__ float sm[kSmaChunk];
    __shared__ float sTotal;
    __shared__ double sAllowance;
    __shared__ float sCmaxAllow;
    if (t.filter) {
        blockCenteredSma(t.win, t.a, n, kNoisePoints, sm, &sTotal);                 // a = mean over 16 bins of the windowed row
        for (int i = tid; i < n; i += kTrackThreads) { const float e = __fsub_rn(t.win[i], t.a[i]); t.b[i] = __fmul_rn(e, e); }
        __syncthreads();
        blockCenteredSma(t.b, t.a, n, kNoisePoints, sm, &sTotal);                   // a = moving variance
        blockCenteredSma(t.win, t.c, n, t.largeTick, sm, &sTotal);                  // c = long moving average of the row
        if (tid == 0) {
            const double newAllowance = t.lossRate * (double)seqPercentileSampling(t.a, n, .15);
            const double allowance = __dadd_rn(__dmul_rn(newAllowance, 0.1), __dmul_rn((double)*t.prevAllowance, 0.9));     // (no contraction)
            *t.prevAllowance = (float)allowance;
            sAllowance = allowance;
        }
        __syncthreads();
        const double allowance = sAllowance;
        for (int i = tid; i < n; i += kTrackThreads) { if ((double)t.a[i] > allowance) { t.c[i] = 0.0f; } }
        __syncthreads();
        blockFillHoles(t.c, t.e, n);
        blockCenteredSma(t.e, t.b, n, t.largeTick, sm, &sTotal);                    // b = cleared noise floor
        blockCenteredSma(t.b, t.d, n, 5 * t.largeTick, sm, &sTotal);                // d = its longer average
        for (int i = tid; i < n; i += kTrackThreads) { t.d[i] = fabsf(__fsub_rn(t.b[i], t.d[i])); }
        __syncthreads();
        if (tid == 0) { sCmaxAllow = seqPercentileSampling(t.d, n, .15); }
        __syncthreads();
        const float cmaxAllow = sCmaxAllow;
        for (int i = tid; i < n; i += kTrackThreads) { if (t.d[i] > cmaxAllow) { t.b[i] = 0.0f; } }
        __syncthreads();
        blockFillHoles(t.b, t.e, n);
        blockCenteredSma(t.e, t.c, n, t.largeTick, sm, &sTotal);                    // c = the noise floor
        // bins above the floor by more than the allowance, and the caller's masked ranges, widened by 37 bins
        for (int i = tid; i < n; i += kTrackThreads) { t.a[i] = ((double)t.clr[i] > (double)t.c[i] + allowance) ? 1.0f : 0.0f; }
        __syncthreads();
        for (int m = 0; m + 1 < t.nMasked; m += 2) {
            const int from = (int)((float)(n / 2) + __fdiv_rn((float)t.masked[m], t.hzTick));
            const int to = (int)((float)(n / 2) + __fdiv_rn((float)t.masked[m + 1], t.hzTick));
            for (int j = from + tid; j < to; j += kTrackThreads) { if (j >= 0 && j < n) { t.a[j] = 1.0f; } }
        }
        __syncthreads();
        blockCenteredSma(t.a, t.b, n, kSignalWidth / 8, sm, &sTotal);
        for (int i = tid; i < n; i += kTrackThreads) { if (t.b[i] == 0.0f) { t.frame[i] = cf32{ 0.0f, 0.0f }; } }
        const int slice = n / 30;
        if (tid < 30) { t.noiseFigure[tid] = __fadd_rn(7.0f, t.c[tid * slice + slice / 2]); }
        __syncthreads();
    }
    // amplitude to its fourth root, for the 8-bit scaling that follows
    for (int i = tid; i < n; i += kTrackThreads) {
        cf32 v = t.frame[i];
        if (v.re == 0.0f && v.im == 0.0f) { continue; }
        const float amp = sqrtf(__fadd_rn(__fmul_rn(v.re, v.re), __fmul_rn(v.im, v.im)));
        const float g = __fdiv_rn(sqrtf(sqrtf(amp)), amp);
        t.frame[i] = cf32{ __fmul_rn(v.re, g), __fmul_rn(v.im, g) };
    }
}

inline int blocksFor(int n, int t) { return (n + t - 1) / t; }

} // namespace

struct FftCompressor {
    int device = 0;
    int sampleRate = 0, N = 0, maxBlock = 0;
    double lossRate = 1.0;
    float hzTick = 0; int largeTick = 0;
    int freqCount = 0, freqHead = 0, magCount = 0, magHead = 0;
    bool haveNoise = false;
    cudaStream_t lastStream = nullptr; bool everRan = false;
    cf32* dFifo[2] = { nullptr, nullptr }; int fifoCap = 0, fifoLen = 0, cur = 0;
    float* dWin = nullptr;
    cf32* dSpec = nullptr; cf32* dFreqRing = nullptr; cf32* dOut = nullptr;
    float* dClrRing = nullptr; float* dWinRing = nullptr; float* dClrAvg = nullptr; float* dWinAvg = nullptr;
    float* dScratch = nullptr;       // 5 x N
    float* dState = nullptr;         // [0] prevAllowance, [1..30] noise figure
    int* dMasked = nullptr; int nMasked = 0;
    FrameDesc* dFrames = nullptr;
    cf32* dStage = nullptr;
    FftPlan plan;

    int init(int sampleRate, double lossRate, int maxBlock, int device);
    void destroy();
    int syncLast() { if (everRan) { B200_CUDA(cudaStreamSynchronize(lastStream)); } return B200_OK; }
    int process(const cf32* dIn, int count, int* outCount, cudaStream_t s);
};

int FftCompressor::init(int sampleRate_, double lossRate_, int maxBlock_, int device_) {
    device = device_; sampleRate = sampleRate_; lossRate = lossRate_; maxBlock = maxBlock_;
    DeviceGuard g(device);
    if (!g.ok) { setError("cannot select CUDA device %d", device); return B200_ECUDA; }
    int n = sampleRate * 50 / 1000;                                      // :70-71
    if (n < 1) { setError("fft compressor: sample rate %d too low", sampleRate); return B200_EINVAL; }
    n = (int)pow(2, floor(log2((double)n)));
    hzTick = ((float)sampleRate) / n;
    const int smallTick = (int)(kSignalWidth / hzTick);
    largeTick = smallTick * 10;
    if (n < 64 || 5 * largeTick > n || largeTick < 1) { setError("fft compressor: sample rate %d gives frames of %d bins: too short for its %d-bin averages", sampleRate, n, 5 * largeTick); return B200_EINVAL; }
    N = n;
    B200_TRY(plan.create(N, device));
    fifoCap = 4 * std::max(maxBlock, N);
    for (int k = 0; k < 2; k++) { B200_CUDA(cudaMalloc((void**)&dFifo[k], (size_t)fifoCap * sizeof(cf32))); }
    B200_CUDA(cudaMalloc((void**)&dWin, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dSpec, (size_t)N * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dOut, (size_t)N * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dFreqRing, (size_t)kMinRecents * N * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dClrRing, (size_t)kMagRows * N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dWinRing, (size_t)kMagRows * N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dClrAvg, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dWinAvg, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dScratch, (size_t)5 * N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dState, 32 * sizeof(float)));
    B200_CUDA(cudaMemset(dState, 0, 32 * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dMasked, 256 * sizeof(int)));
    B200_CUDA(cudaMalloc((void**)&dFrames, 2 * sizeof(FrameDesc)));
    std::vector<float> win((size_t)N);
    for (int i = 0; i < N; i++) {                                        // blackmanWindowElement(i + 5, fftSize + 10) :15-27
        const double factor = 2.0 * M_PI * (i + 5) / (N + 10 - 1);
        win[i] = (float)(0.42 - 0.5 * std::cos(factor) + 0.08 * std::cos(2 * factor));
    }
    B200_CUDA(cudaMemcpy(dWin, win.data(), win.size() * sizeof(float), cudaMemcpyHostToDevice));
    return B200_OK;
}

void FftCompressor::destroy() {
    DeviceGuard g(device);
    if (everRan) { cudaStreamSynchronize(lastStream); }
    void* ptrs[] = { dFifo[0], dFifo[1], dWin, dSpec, dOut, dFreqRing, dClrRing, dWinRing, dClrAvg, dWinAvg, dScratch, dState, dMasked, dFrames, dStage };
    for (void* p : ptrs) { if (p) { cudaFree(p); } }
    plan.destroy();
}

// run() :431-452 with process() :198-427: the block joins the input buffer, then at most ONE frame is taken from it
int FftCompressor::process(const cf32* dIn, int count, int* outCount, cudaStream_t s) {
    *outCount = 0;
    if (count < 0 || count > maxBlock) { setError("fft compressor: block of %d samples exceeds max_block %d", count, maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(device);
    if (everRan && s != lastStream) { B200_TRY(syncLast()); }
    lastStream = s; everRan = true;
    if (fifoLen + count > fifoCap) { setError("fft compressor: input buffer overflow (one frame of %d samples leaves per call)", N); return B200_ETOOBIG; }
    if (count > 0) { B200_CUDA(cudaMemcpyAsync(dFifo[cur] + fifoLen, dIn, (size_t)count * sizeof(cf32), cudaMemcpyDeviceToDevice, s)); }
    fifoLen += count;
    if (fifoLen < N) { return B200_OK; }
    // ---- the frame's spectrum (-> delay line, dB row) and windowed spectrum (-> dB row)
    FrameDesc fd; fd.src0 = dFifo[cur]; fd.src1 = dFifo[cur]; fd.len0 = N; fd.pad = 0;
    B200_TRY(uploadDesc(dFrames, &fd, sizeof(fd), s, true));
    const int freqSlot = (freqHead + freqCount) % kMinRecents;           // the delay line never overflows: one leaves per frame once full
    int magSlot;
    if (magCount < kMagRows) { magSlot = (magHead + magCount) % kMagRows; magCount++; }
    else { magSlot = magHead; magHead = (magHead + 1) % kMagRows; }
    freqCount++;
    B200_TRY(plan.run(dFrames, 1, nullptr, N, nullptr, dSpec, s));
    fcSpectrumKernel<<<blocksFor(N, 256), 256, 0, s>>>(dSpec, dFreqRing + (size_t)freqSlot * N, dClrRing + (size_t)magSlot * N, N);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(plan.run(dFrames, 1, dWin, N, nullptr, dSpec, s));
    fcSpectrumKernel<<<blocksFor(N, 256), 256, 0, s>>>(dSpec, nullptr, dWinRing + (size_t)magSlot * N, N);
    B200_CHECK_LAUNCH(); countLaunch();
    // the frame leaves the input buffer
    const int rest = fifoLen - N;
    B200_CUDA(cudaMemcpyAsync(dFifo[cur ^ 1], dFifo[cur] + N, (size_t)rest * sizeof(cf32), cudaMemcpyDeviceToDevice, s));
    cur ^= 1; fifoLen = rest;
    if (freqCount < kMinRecents) { return B200_OK; }
    // ---- the oldest frame goes out, filtered by what the last rows say about the noise floor
    B200_CUDA(cudaMemcpyAsync(dOut, dFreqRing + (size_t)freqHead * N, (size_t)N * sizeof(cf32), cudaMemcpyDeviceToDevice, s));
    const float inv = (float)(1.0 / (float)magCount);
    fcMeanKernel<<<blocksFor(N, 256), 256, 0, s>>>(dWinRing, magHead, magCount, kMagRows, inv, dWinAvg, N);
    B200_CHECK_LAUNCH(); countLaunch();
    fcMeanKernel<<<blocksFor(N, 256), 256, 0, s>>>(dClrRing, magHead, magCount, kMagRows, inv, dClrAvg, N);
    B200_CHECK_LAUNCH(); countLaunch();
    TrackArgs t;
    t.win = dWinAvg; t.clr = dClrAvg; t.frame = dOut;
    t.a = dScratch; t.b = dScratch + N; t.c = dScratch + 2 * (size_t)N; t.d = dScratch + 3 * (size_t)N; t.e = dScratch + 4 * (size_t)N;
    t.prevAllowance = dState; t.noiseFigure = dState + 1;
    t.masked = dMasked; t.nMasked = nMasked;
    t.n = N; t.largeTick = largeTick; t.hzTick = hzTick; t.lossRate = lossRate; t.filter = lossRate > 0 ? 1 : 0;
    fcTrackKernel<<<1, kTrackThreads, 0, s>>>(t);
    B200_CHECK_LAUNCH(); countLaunch();
    if (t.filter) { haveNoise = true; }
    freqHead = (freqHead + 1) % kMinRecents; freqCount--;               // cleanFreqDomain.erase(begin) :412
    if (magCount > 2 * kMinRecents) { magHead = (magHead + 1) % kMagRows; magCount--; }      // :413-416
    *outCount = N;
    return B200_OK;
}

} // namespace b200

using namespace b200;

struct b200_fftcomp { FftCompressor impl; };

#define FC_OR_FAIL(c) \
    if (!(c)) { setError("NULL fft compressor handle"); return B200_EINVAL; }

extern "C" {

int b200_fftcomp_create(b200_fftcomp** out, int sample_rate, double loss_rate, int max_block, int device) {
    if (!out || sample_rate <= 0 || max_block <= 0) { setError("bad arguments to b200_fftcomp_create"); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { setError("no CUDA device available: libb200dsp has no CPU fallback"); return B200_ECUDA; }
    if (device < 0 || device >= ndev) { setError("device %d out of range (%d devices)", device, ndev); return B200_EINVAL; }
    auto* c = new (std::nothrow) b200_fftcomp;
    if (!c) { return B200_ENOMEM; }
    const int rc = c->impl.init(sample_rate, loss_rate, max_block, device);
    if (rc != B200_OK) { c->impl.destroy(); delete c; return rc; }
    *out = c;
    return B200_OK;
}

void b200_fftcomp_destroy(b200_fftcomp* c) {
    if (!c) { return; }
    c->impl.destroy();
    delete c;
}

int b200_fftcomp_fft_size(b200_fftcomp* c) { return c ? c->impl.N : 0; }
int b200_fftcomp_set_loss_rate(b200_fftcomp* c, double loss_rate) { FC_OR_FAIL(c); c->impl.lossRate = loss_rate; return B200_OK; }

int b200_fftcomp_set_masked(b200_fftcomp* c, const int32_t* from_to_hz, int n) {
    FC_OR_FAIL(c);
    if (n < 0 || n > 256 || (n > 0 && !from_to_hz)) { setError("fft compressor: up to 128 masked ranges"); return B200_EINVAL; }
    DeviceGuard g(c->impl.device);
    B200_TRY(c->impl.syncLast());
    if (n > 0) { B200_CUDA(cudaMemcpy(c->impl.dMasked, from_to_hz, (size_t)n * sizeof(int), cudaMemcpyHostToDevice)); }
    c->impl.nMasked = n;
    return B200_OK;
}

int b200_fftcomp_process(b200_fftcomp* c, const float* d_in, int count, const float** d_out, int* out_count, void* stream) {
    FC_OR_FAIL(c);
    if (!out_count || (count > 0 && !d_in)) { setError("bad arguments to b200_fftcomp_process"); return B200_EINVAL; }
    const int rc = c->impl.process((const cf32*)d_in, count, out_count, (cudaStream_t)stream);
    if (d_out) { *d_out = (const float*)c->impl.dOut; }
    return rc;
}

int b200_fftcomp_process_host(b200_fftcomp* c, const float* h_in, int count, float* h_out, int* out_count, void* stream) {
    FC_OR_FAIL(c);
    if (!out_count || (count > 0 && !h_in)) { setError("bad arguments to b200_fftcomp_process_host"); return B200_EINVAL; }
    FftCompressor& f = c->impl;
    if (count < 0 || count > f.maxBlock) { setError("fft compressor: block of %d samples exceeds max_block %d", count, f.maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(f.device);
    cudaStream_t s = (cudaStream_t)stream;
    B200_TRY(f.syncLast());
    if (!f.dStage) { B200_CUDA(cudaMalloc((void**)&f.dStage, (size_t)f.maxBlock * sizeof(cf32))); }
    if (count > 0) { B200_CUDA(cudaMemcpyAsync(f.dStage, h_in, (size_t)count * sizeof(cf32), cudaMemcpyHostToDevice, s)); }
    B200_TRY(f.process(f.dStage, count, out_count, s));
    if (*out_count > 0 && h_out) { B200_CUDA(cudaMemcpyAsync(h_out, f.dOut, (size_t)*out_count * sizeof(cf32), cudaMemcpyDeviceToHost, s)); }
    B200_CUDA(cudaStreamSynchronize(s));
    return B200_OK;
}

int b200_fftcomp_sync(b200_fftcomp* c) { FC_OR_FAIL(c); DeviceGuard g(c->impl.device); return c->impl.syncLast(); }

// noiseFigure (30 values, server.cpp:470-474 sends them to the client): returns the count, 0 before the first filtered frame
int b200_fftcomp_noise_figure(b200_fftcomp* c, float* h_out, int cap) {
    FC_OR_FAIL(c);
    DeviceGuard g(c->impl.device);
    B200_TRY(c->impl.syncLast());
    if (!c->impl.haveNoise) { return 0; }
    if (h_out && cap > 0) { B200_CUDA(cudaMemcpy(h_out, c->impl.dState + 1, (size_t)std::min(cap, 30) * sizeof(float), cudaMemcpyDeviceToHost)); }
    return 30;
}

} // extern "C"
