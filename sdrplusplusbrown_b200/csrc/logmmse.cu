// The logmmse module's front-end pre-processor on the device: wide-band log-MMSE spectral noise
// reduction of the IQ stream.
//
// Reference: misc_modules/noise_reduction_logmmse/src/if_nr.h:63-113 (dsp::IFNRLogMMSE::process: sample FIFO, first call
// after 28 hops learns the noise from 12 frames, every call de-noises all complete frames and multiplies by 4), added to
// the front end's chain by the module (main.cpp:165); logmmse.h:244-296 (logmmse_sample), :298-405 (logmmse_all: Hann
// frames of Slen = even(floor(0.02 * rate)) samples, hop Slen / 2, transform length 2 * Slen, a-priori SNR by decision
// direction, gain A * exp(E1(v) / 2), overlap-add), :117-143 (noise / deviation histories of 200 frames and their running
// sums), :150-242 (wide-band noise tracker: bins whose deviation is below the mode of the deviation histogram carry
// the noise estimate, the others are interpolated), bgnoise.h:27-59, core/src/utils/arrays.cpp:60-431 (E1 table),
// :433-470 (hole interpolation), :1068-1092 (moving average).
//
// The frames of a call depend on each other (decision-directed SNR, overlap-add, running sums), so the stream is walked
// frame by frame; inside a frame every stage is one launch over the 2 * Slen bins.  The transform length is not a power
// of two (96 000 at 2.4 MS/s): ChirpZ (chirpz.cu).  Every per-bin recurrence is evaluated with the reference's own
// sequence of IEEE single-precision operations (no contraction), so all that separates the results from the
// reference's is the rounding of the transforms (and of log10f in the noise tracker's histogram).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

#include "chirpz.cuh"

namespace b200 {

namespace {

constexpr int kE1Entries = 3001;
constexpr float kErased = 1e9f;           // BackgroundNoiseCaltulator::ERASED_SAMPLE
constexpr int kBuckets = 1000;

// exponential integral E1 in float64 (power series below 1, continued fraction above): the reference tabulates
// scipy.special.expn(1, x) at x = (i / 1000)^4 as floats (arrays.cpp:62-420); the table is rebuilt from the function
double expInt1(double x) {
    if (x <= 1.0) {
        double sum = 0.0, term = 1.0;
        for (int k = 1; k < 200; k++) {
            term *= -x / (double)k;
            const double add = -term / (double)k;
            sum += add;
            if (fabs(add) < 1e-18 * fabs(sum)) { break; }
        }
        return -0.57721566490153286061 - log(x) + sum;
    }
    double b = x + 1.0, c = 1e300, d = 1.0 / b, h = d;
    for (int i = 1; i < 2000; i++) {
        const double an = -(double)i * (double)i;
        b += 2.0;
        d = 1.0 / (an * d + b);
        c = b + an / c;
        const double del = c * d;
        h *= del;
        if (fabs(del - 1.0) < 1e-16) { break; }
    }
    return h * exp(-x);
}

__device__ __forceinline__ float e1Lookup(const float* __restrict__ t, float x) {       // EXPN::expn__ (arrays.cpp:421-427)
    const float index = __fmul_rn(sqrtf(sqrtf(x)), 1000.0f);
    const float iindex = floorf(index);
    const float f = __fsub_rn(index, iindex);
    int i = (int)iindex;
    i = min(max(i, 0), kE1Entries - 2);       // (the reference reads past its table for arguments beyond 81)
    return __fadd_rn(t[i], __fmul_rn(__fsub_rn(t[i + 1], t[i]), f));
}

__device__ __forceinline__ float fastExp(float x) {                                     // volk_32f_expfast_32f, generic
    const float fact = 8388608.0f / 0.6931471805f, bias = 1065353216.0f - 60801.0f;
    return __int_as_float(__float2int_rz(__fadd_rn(__fmul_rn(fact, x), bias)));
}

// seg[i] = complex_t{win[i], 0} * x[i] (muleach, arrays.cpp:817-827), i < Slen
__global__ void lmWindowKernel(const cf32* __restrict__ x, const float* __restrict__ win, cf32* __restrict__ seg, int slen) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= slen) { return; }
    const cf32 v = x[i];
    const float w = win[i];
    seg[i] = cf32{ __fsub_rn(__fmul_rn(w, v.re), __fmul_rn(0.0f, v.im)), __fadd_rn(__fmul_rn(0.0f, v.re), __fmul_rn(w, v.im)) };
}

__global__ void lmMagnitudeKernel(const cf32* __restrict__ spec, float* __restrict__ sig, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const cf32 v = spec[i];
    sig[i] = sqrtf(__fadd_rn(__fmul_rn(v.re, v.re), __fmul_rn(v.im, v.im)));
}

// logmmse.h:318-322: a zero magnitude takes its left neighbour's (already repaired) value.  In place: whichever thread
// repairs a bin first writes the value any other thread would find further to the left.
__global__ void lmZeroFixKernel(float* sig, int n) {
    const int z = blockIdx.x * blockDim.x + threadIdx.x;
    if (z < 1 || z >= n || sig[z] != 0.0f) { return; }
    int j = z - 1;
    while (j > 0 && sig[j] == 0.0f) { j--; }
    sig[z] = sig[j];
}

// add_noise_history (logmmse.h:117-143): running sums of the last `cap` magnitude frames and of their squared deviation
// from the running mean; `full` = the ring already holds cap frames, slot = where this frame goes (the oldest one's place)
__global__ void lmHistoryKernel(const float* __restrict__ noise, float* __restrict__ sums, float* __restrict__ devs,
                                float* __restrict__ noiseRing, float* __restrict__ devRing, int slot, int full, float invCount, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const size_t at = (size_t)slot * n + i;
    const float v = noise[i];
    float s = __fadd_rn(sums[i], v);
    if (full) { s = __fsub_rn(s, noiseRing[at]); }
    noiseRing[at] = v;
    sums[i] = s;
    const float d = __fsub_rn(v, __fmul_rn(s, invCount));
    const float dd = __fmul_rn(d, d);
    float e = __fadd_rn(devs[i], dd);
    if (full) { e = __fsub_rn(e, devRing[at]); }
    devRing[at] = dd;
    devs[i] = e;
}

__global__ void lmAccumulateKernel(float* __restrict__ acc, const float* __restrict__ v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { acc[i] = __fadd_rn(acc[i], v[i]); }
}

// the tail of logmmse_sample (logmmse.h:280-285): mean of the 12 frames, npmavg(., 120) -- a running sum whose rounding
// depends on the order, so one thread walks it (once per noise sampling) --, squared
__global__ void lmInitialNoiseKernel(const float* __restrict__ mean, float invFrames, float* __restrict__ tmp, float* __restrict__ mu2, int n) {
    if (blockIdx.x != 0 || threadIdx.x != 0) { return; }
    for (int i = 0; i < n; i++) { tmp[i] = __fmul_rn(mean[i], invFrames); }
    const int windowSize = 120, ws2 = 60;
    float sum = 0.0f, count = 0.0f;
    int o = 0;
    for (int ix = 0; ix < n + ws2; ix++) {
        if (ix < n) { sum = __fadd_rn(sum, tmp[ix]); count = __fadd_rn(count, 1.0f); }
        if (ix > windowSize) { count = __fsub_rn(count, 1.0f); sum = __fsub_rn(sum, tmp[(int)__fsub_rn((float)ix, count)]); }
        if (ix >= ws2) { const float a = __fdiv_rn(sum, count); mu2[o++] = __fmul_rn(a, a); }
    }
}

// the per-bin gain (logmmse.h:326-352) and the shaped spectrum; anyZero[0] tells whether the previous frame's Xk held a
// zero (npall, :329), anyZero[1] collects the same for this frame's
__global__ void lmGainKernel(const float* __restrict__ sig, const cf32* __restrict__ spec, const float* __restrict__ mu2,
                             float* __restrict__ xkPrev, cf32* __restrict__ shaped, const float* __restrict__ e1,
                             const int* __restrict__ anyZeroIn, int* __restrict__ anyZeroOut, float aa, float ksiMin, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const float s = sig[i], m2 = mu2[i];
    float g = __fdiv_rn(__fmul_rn(s, s), m2);
    if (g > 40.0f) { g = 40.0f; }
    float m = __fadd_rn(g, -1.0f);
    if (m < 0.0f) { m = 0.0f; }
    const float oneMinus = __fsub_rn(1.0f, aa);
    float ksi;
    if (*anyZeroIn) { ksi = __fadd_rn(__fmul_rn(m, oneMinus), aa); }
    else {
        ksi = __fadd_rn(__fdiv_rn(__fmul_rn(xkPrev[i], aa), m2), __fmul_rn(m, oneMinus));
        if (ksi < ksiMin) { ksi = ksiMin; }
    }
    const float A = __fdiv_rn(ksi, __fadd_rn(ksi, 1.0f));
    const float vk = __fmul_rn(A, g);
    const float hw = __fmul_rn(A, fastExp(__fmul_rn(e1Lookup(e1, vk), 0.5f)));
    const float so = __fmul_rn(s, hw);
    const float xk = __fmul_rn(so, so);
    xkPrev[i] = xk;
    if (xk == 0.0f) { *anyZeroOut = 1; }
    const cf32 v = spec[i];
    shaped[i] = cf32{ __fsub_rn(__fmul_rn(hw, v.re), __fmul_rn(0.0f, v.im)), __fadd_rn(__fmul_rn(0.0f, v.re), __fmul_rn(hw, v.im)) };
}

// overlap-add (logmmse.h:357-362) and the module's fixed gain of 4 (if_nr.h:103-107)
__global__ void lmOverlapKernel(const cf32* __restrict__ back, cf32* __restrict__ xOld, cf32* __restrict__ out, int len1) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len1) { return; }
    const cf32 o = xOld[i], b = back[i];
    out[i] = cf32{ __fmul_rn(__fadd_rn(o.re, b.re), 4.0f), __fmul_rn(__fadd_rn(o.im, b.im), 4.0f) };
    xOld[i] = back[len1 + i];
}

// ---- noise tracker (update_noise_mu2, wide-band branch, logmmse.h:200-240)
// avg = sums / nframes, dv = (devs / nframes)^2 with the middle 30 % of the table (the band edges) erased
__global__ void lmNoiseStatsKernel(const float* __restrict__ sums, const float* __restrict__ devs, float inv, float* __restrict__ avg,
                                   float* __restrict__ dv, int n) {
    const int z = blockIdx.x * blockDim.x + threadIdx.x;
    if (z >= n) { return; }
    avg[z] = __fmul_rn(sums[z], inv);
    const float hi = __fmul_rn(devs[z], inv);
    dv[z] = (abs(z - n / 2) < n * 15 / 100) ? kErased : __fmul_rn(hi, hi);
}

__device__ __forceinline__ void atomicMinFloat(float* addr, float v) {
    int* a = reinterpret_cast<int*>(addr);
    int old = *a;
    while (__int_as_float(old) > v) { const int prev = atomicCAS(a, old, __float_as_int(v)); if (prev == old) { break; } old = prev; }
}
__device__ __forceinline__ void atomicMaxFloat(float* addr, float v) {
    int* a = reinterpret_cast<int*>(addr);
    int old = *a;
    while (__int_as_float(old) < v) { const int prev = atomicCAS(a, old, __float_as_int(v)); if (prev == old) { break; } old = prev; }
}

// BackgroundNoiseCaltulator::addFrame (bgnoise.h:36-49): log10 of the kept bins, their range ...
__global__ void lmLogRangeKernel(const float* __restrict__ dv, float* __restrict__ lg, float* __restrict__ range, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo = kErased, hi = -kErased;
    if (i < n) {
        const float q = dv[i];
        float l = kErased;
        if (q != kErased) { l = log10f(q); lo = fminf(lo, l); hi = fmaxf(hi, l); }
        lg[i] = l;
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (lo != kErased) { atomicMinFloat(range, lo); }
        if (hi != -kErased) { atomicMaxFloat(range + 1, hi); }
    }
}

// ... and their histogram over 1000 buckets (bgnoise.h:50-54; the largest value falls one past the end there and is not counted)
__global__ void lmHistogramKernel(const float* __restrict__ dv, const float* __restrict__ lg, const float* __restrict__ range,
                                  int* __restrict__ buckets, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || dv[i] == kErased) { return; }
    const float minn = range[0], width = __fsub_rn(range[1], range[0]);
    const int b = __float2int_rz(__fmul_rn(1000.0f, __fdiv_rn(__fsub_rn(lg[i], minn), width)));
    if (b >= 0 && b < kBuckets) { atomicAdd(buckets + b, 1); }
}

// bins quieter than the limit carry avg^2, the others 0 (logmmse.h:222-229)
__global__ void lmNoiseApplyKernel(const float* __restrict__ avg, const float* __restrict__ dv, float limit, float* __restrict__ holes,
                                   int* __restrict__ anyKept, int n) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) { return; }
    float v = 0.0f;
    if (dv[q] < limit) { v = __fmul_rn(avg[q], avg[q]); }
    holes[q] = v;
    if (v != 0.0f) { *anyKept = 1; }
}

// linearInterpolateHoles (arrays.cpp:433-470): every kept bin fills the gap to its left neighbour with the reference's
// running sum (its rounding depends on the order, so one thread walks one gap); the first / last kept bins also fill the
// table's ends.  With no kept bin the previous estimate stays (logmmse.h:231-233).
__global__ void lmFillHolesKernel(const float* __restrict__ holes, const int* __restrict__ anyKept, float* __restrict__ mu2, int n) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n || !*anyKept) { return; }
    const float val = holes[q];
    if (val == 0.0f) { return; }
    mu2[q] = val;
    int last = q - 1;
    while (last >= 0 && holes[last] == 0.0f) { last--; }
    if (last < 0) { for (int w = q - 1; w >= 0; w--) { mu2[w] = val; } }
    else if (q - last > 1) {
        const float d = __fdiv_rn(__fsub_rn(val, holes[last]), (float)(q - last));
        float running = holes[last];
        for (int w = last + 1; w < q; w++) { running = __fadd_rn(running, d); mu2[w] = running; }
    }
    int next = q + 1;
    while (next < n && holes[next] == 0.0f) { next++; }
    if (next >= n) { for (int w = q + 1; w < n; w++) { mu2[w] = val; } }
}

inline int blocksFor(int n, int t) { return (n + t - 1) / t; }

} // namespace

struct LogMmse {
    int device = 0;
    int rate = 0, maxBlock = 0;
    bool hold = false, needReset = true, sampled = false;
    // geometry
    int Slen = 0, len1 = 0, len2 = 0, nFFT = 0, histCap = 0;
    // state
    int histCount = 0, histHead = 0;      // ring: the oldest frame sits at histHead
    int bgFrameCount = 0; double bgLastNoise = kErased;
    int parity = 0;
    cudaStream_t lastStream = nullptr; bool everRan = false;
    // device memory
    cf32* dFifo[2] = { nullptr, nullptr }; int fifoCap = 0, fifoLen = 0, cur = 0;
    float* dWin = nullptr; float* dE1 = nullptr;
    float* dSums = nullptr; float* dDevs = nullptr; float* dNoiseRing = nullptr; float* dDevRing = nullptr;
    float* dMu2 = nullptr; float* dXk = nullptr; cf32* dXOld = nullptr;
    cf32* dSeg = nullptr; cf32* dSpec = nullptr; cf32* dShaped = nullptr; cf32* dBack = nullptr;
    float* dSig = nullptr; float* dAvg = nullptr; float* dDv = nullptr; float* dLg = nullptr; float* dHoles = nullptr; float* dMean = nullptr;
    float* dRange = nullptr; int* dBuckets = nullptr; int* dFlags = nullptr;      // flags: [0..1] anyZero ping-pong, [2] anyKept
    cf32* dOut = nullptr; int outCap = 0;
    cf32* dStage = nullptr; int stageCap = 0;
    ChirpZ cz;

    int init(int rate, int maxBlock, int device);
    void destroy();
    int syncLast() { if (everRan) { B200_CUDA(cudaStreamSynchronize(lastStream)); } return B200_OK; }
    void resetState();
    int magnitudeOf(const cf32* x, cudaStream_t s);           // window, forward transform, |.|  -> dSpec, dSig
    int pushHistory(cudaStream_t s);
    int sample(cudaStream_t s);
    int updateNoise(cudaStream_t s);
    int process(const cf32* dIn, int count, int* outCount, cudaStream_t s);
};

int LogMmse::init(int rate_, int maxBlock_, int device_) {
    device = device_; rate = rate_; maxBlock = maxBlock_;
    DeviceGuard g(device);
    if (!g.ok) { setError("cannot select CUDA device %d", device); return B200_ECUDA; }
    Slen = (int)floor(0.02 * rate);                 // logmmse.h:245-251
    if (Slen % 2 == 1) { Slen++; }
    if (Slen < 4) { setError("logmmse: sample rate %d too low", rate); return B200_EINVAL; }
    len1 = (int)floor(Slen * 50 / 100);
    len2 = Slen - len1;
    nFFT = 2 * Slen;
    histCap = nFFT < 1000 ? 2000 : 200;             // noise_history_len (logmmse.h:57-63)
    B200_TRY(cz.init(nFFT, device));
    const int fram = rate / 100;
    fifoCap = fram * 28 + Slen + 2 * maxBlock + 64;
    outCap = fifoCap;
    for (int k = 0; k < 2; k++) { B200_CUDA(cudaMalloc((void**)&dFifo[k], (size_t)fifoCap * sizeof(cf32))); }
    B200_CUDA(cudaMalloc((void**)&dOut, (size_t)outCap * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dWin, (size_t)Slen * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dE1, kE1Entries * sizeof(float)));
    float** perBin[] = { &dSums, &dDevs, &dMu2, &dXk, &dSig, &dAvg, &dDv, &dLg, &dHoles, &dMean };
    for (float** p : perBin) { B200_CUDA(cudaMalloc((void**)p, (size_t)nFFT * sizeof(float))); }
    B200_CUDA(cudaMalloc((void**)&dNoiseRing, (size_t)histCap * nFFT * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dDevRing, (size_t)histCap * nFFT * sizeof(float)));
    cf32** perBinC[] = { &dSpec, &dShaped, &dBack };
    for (cf32** p : perBinC) { B200_CUDA(cudaMalloc((void**)p, (size_t)nFFT * sizeof(cf32))); }
    B200_CUDA(cudaMalloc((void**)&dSeg, (size_t)Slen * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dXOld, (size_t)len1 * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dRange, 2 * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dBuckets, kBuckets * sizeof(int)));
    B200_CUDA(cudaMalloc((void**)&dFlags, 4 * sizeof(int)));
    // nphanning scaled to the hop (logmmse.h:262-264): fp32 in the reference's order
    std::vector<float> win((size_t)Slen), scaled((size_t)Slen);
    for (int i = 0; i < Slen; i++) { win[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * i / (Slen - 1))); }
    float total = 0;
    for (int i = 0; i < Slen; i++) { scaled[i] = win[i] * (float)len2; }
    for (int i = 0; i < Slen; i++) { total += win[i]; }
    const float invTotal = (float)(1.0 / total);
    for (int i = 0; i < Slen; i++) { win[i] = scaled[i] * invTotal; }
    B200_CUDA(cudaMemcpy(dWin, win.data(), win.size() * sizeof(float), cudaMemcpyHostToDevice));
    std::vector<float> e1(kE1Entries);
    e1[0] = (float)expInt1(1e-16);
    for (int i = 1; i < kE1Entries; i++) { const double r = (double)i / 1000.0; e1[i] = (float)expInt1(r * r * r * r); }
    B200_CUDA(cudaMemcpy(dE1, e1.data(), e1.size() * sizeof(float), cudaMemcpyHostToDevice));
    resetState();
    return B200_OK;
}

void LogMmse::destroy() {
    DeviceGuard g(device);
    if (everRan) { cudaStreamSynchronize(lastStream); }
    void* ptrs[] = { dFifo[0], dFifo[1], dOut, dWin, dE1, dSums, dDevs, dMu2, dXk, dSig, dAvg, dDv, dLg, dHoles, dMean, dNoiseRing, dDevRing,
                     dSpec, dShaped, dBack, dSeg, dXOld, dRange, dBuckets, dFlags, dStage };
    for (void* p : ptrs) { if (p) { cudaFree(p); } }
    cz.destroy();
}

// SavedParamsC::reset (logmmse.h:103-115) + the emptied FIFO (if_nr.h:64-77)
void LogMmse::resetState() {
    fifoLen = 0; sampled = false;
    histCount = 0; histHead = 0;
    bgFrameCount = 0; bgLastNoise = kErased;
}

int LogMmse::magnitudeOf(const cf32* x, cudaStream_t s) {
    lmWindowKernel<<<blocksFor(Slen, 256), 256, 0, s>>>(x, dWin, dSeg, Slen);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(cz.run(dSeg, Slen, dSpec, false, 1.0f, s));         // the plan zero-pads to nFFT (arrays.cpp:533-535)
    lmMagnitudeKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dSpec, dSig, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

int LogMmse::pushHistory(cudaStream_t s) {
    if (hold) { return B200_OK; }
    const int full = histCount >= histCap ? 1 : 0;
    const int slot = full ? histHead : (histHead + histCount) % histCap;
    if (full) { histHead = (histHead + 1) % histCap; } else { histCount++; }
    const float inv = (float)(1.0 / (float)histCount);           // div(sums, size): times (float)(1.0 / size)
    lmHistoryKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dSig, dSums, dDevs, dNoiseRing, dDevRing, slot, full, inv, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

// logmmse_sample (logmmse.h:244-296)
int LogMmse::sample(cudaStream_t s) {
    const int noiseFrames = 12;
    histCount = 0; histHead = 0;
    B200_CUDA(cudaMemsetAsync(dSums, 0, (size_t)nFFT * sizeof(float), s));
    B200_CUDA(cudaMemsetAsync(dDevs, 0, (size_t)nFFT * sizeof(float), s));
    B200_CUDA(cudaMemsetAsync(dMean, 0, (size_t)nFFT * sizeof(float), s));
    for (int j = 0; j < Slen * noiseFrames; j += Slen) {
        B200_TRY(magnitudeOf(dFifo[cur] + j, s));
        B200_TRY(pushHistory(s));
        lmAccumulateKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dMean, dSig, nFFT);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    lmInitialNoiseKernel<<<1, 32, 0, s>>>(dMean, (float)(1.0 / (float)noiseFrames), dAvg, dMu2, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_CUDA(cudaMemsetAsync(dXOld, 0, (size_t)len1 * sizeof(cf32), s));
    // Xk_prev = zeros: the first frame takes the branch without it (npall is false)
    const int flags[4] = { 1, 1, 0, 0 };
    B200_CUDA(cudaMemcpyAsync(dFlags, flags, sizeof(flags), cudaMemcpyHostToDevice, s));
    parity = 0;
    sampled = true;
    return B200_OK;
}

int LogMmse::updateNoise(cudaStream_t s) {
    if (!(histCount > 100 && !hold)) { return B200_OK; }
    const float inv = 1 / (float)histCount;
    lmNoiseStatsKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dSums, dDevs, inv, dAvg, dDv, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    // BackgroundNoiseCaltulator::addFrame: a new estimate with every 10th call, the last one in between (bgnoise.h:28-31)
    if (!(bgFrameCount > 0 && bgFrameCount % 10 != 0)) {
        const float init[2] = { kErased, -kErased };
        B200_CUDA(cudaMemcpyAsync(dRange, init, sizeof(init), cudaMemcpyHostToDevice, s));
        B200_CUDA(cudaMemsetAsync(dBuckets, 0, kBuckets * sizeof(int), s));
        lmLogRangeKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dDv, dLg, dRange, nFFT);
        B200_CHECK_LAUNCH(); countLaunch();
        lmHistogramKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dDv, dLg, dRange, dBuckets, nFFT);
        B200_CHECK_LAUNCH(); countLaunch();
        int buckets[kBuckets]; float range[2];
        B200_CUDA(cudaMemcpyAsync(buckets, dBuckets, sizeof(buckets), cudaMemcpyDeviceToHost, s));
        B200_CUDA(cudaMemcpyAsync(range, dRange, sizeof(range), cudaMemcpyDeviceToHost, s));
        B200_CUDA(cudaStreamSynchronize(s));                 // 4 KB, once per ten calls: the mode decides the next launches' limit
        const int ix = (int)(std::max_element(buckets, buckets + kBuckets) - buckets);
        const float minn = range[0], width = range[1] - range[0];
        const double maxf = pow(10, ((((double)ix) / kBuckets) * width + minn));
        if (bgLastNoise == kErased) { bgLastNoise = maxf; }
        else { bgLastNoise = 0.9 * bgLastNoise + 0.1 * maxf; }
    }
    bgFrameCount++;
    const float limit = (float)bgLastNoise;
    B200_CUDA(cudaMemsetAsync(dFlags + 2, 0, sizeof(int), s));
    lmNoiseApplyKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dAvg, dDv, limit, dHoles, dFlags + 2, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    lmFillHolesKernel<<<blocksFor(nFFT, 128), 128, 0, s>>>(dHoles, dFlags + 2, dMu2, nFFT);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

// IFNRLogMMSE::process (if_nr.h:63-113) with logmmse_all (logmmse.h:298-405) inside
int LogMmse::process(const cf32* dIn, int count, int* outCount, cudaStream_t s) {
    *outCount = 0;
    if (count < 0 || count > maxBlock) { setError("logmmse: block of %d samples exceeds max_block %d", count, maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(device);
    if (everRan && s != lastStream) { B200_TRY(syncLast()); }
    lastStream = s; everRan = true;
    if (needReset) { needReset = false; resetState(); }
    if (fifoLen + count > fifoCap) { setError("logmmse: sample buffer overflow"); return B200_ETOOBIG; }
    if (count > 0) { B200_CUDA(cudaMemcpyAsync(dFifo[cur] + fifoLen, dIn, (size_t)count * sizeof(cf32), cudaMemcpyDeviceToDevice, s)); }
    fifoLen += count;
    const int noiseFrames = 12, fram = rate / 100;
    int demand = fram * 2;
    if (!sampled) { demand = fram * (noiseFrames + 2) * 2; }
    if (fifoLen < demand) { return B200_OK; }
    if (!sampled) { B200_TRY(sample(s)); }
    const int nFrames = fifoLen / len2 - Slen / len2;
    B200_TRY(updateNoise(s));
    const float aa = 0.98f;
    const float ksiMin = (float)::pow(10, -25.0 / 10.0);
    for (int f = 0; f < nFrames; f++) {
        const int k = f * len2;
        B200_TRY(magnitudeOf(dFifo[cur] + k, s));
        lmZeroFixKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dSig, nFFT);
        B200_CHECK_LAUNCH(); countLaunch();
        B200_TRY(pushHistory(s));
        B200_CUDA(cudaMemsetAsync(dFlags + (parity ^ 1), 0, sizeof(int), s));
        lmGainKernel<<<blocksFor(nFFT, 256), 256, 0, s>>>(dSig, dSpec, dMu2, dXk, dShaped, dE1, dFlags + parity, dFlags + (parity ^ 1), aa, ksiMin, nFFT);
        B200_CHECK_LAUNCH(); countLaunch();
        parity ^= 1;
        B200_TRY(cz.run(dShaped, nFFT, dBack, true, 1.0f / (float)nFFT, s));      // reverse plan: backward transform / N (arrays.cpp:539-541)
        lmOverlapKernel<<<blocksFor(len1, 256), 256, 0, s>>>(dBack, dXOld, dOut + k, len1);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    const int produced = std::max(0, nFrames * len2);
    const int rest = fifoLen - produced;
    if (produced > 0) {
        B200_CUDA(cudaMemcpyAsync(dFifo[cur ^ 1], dFifo[cur] + produced, (size_t)rest * sizeof(cf32), cudaMemcpyDeviceToDevice, s));
        cur ^= 1;
        fifoLen = rest;
    }
    *outCount = produced;
    return B200_OK;
}

} // namespace b200

using namespace b200;

struct b200_logmmse { LogMmse impl; };

#define LM_OR_FAIL(d) \
    if (!(d)) { setError("NULL logmmse handle"); return B200_EINVAL; }

extern "C" {

int b200_logmmse_create(b200_logmmse** out, int sample_rate, int max_block, int device) {
    if (!out || sample_rate <= 0 || max_block <= 0) { setError("bad arguments to b200_logmmse_create"); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { setError("no CUDA device available: libb200dsp has no CPU fallback"); return B200_ECUDA; }
    if (device < 0 || device >= ndev) { setError("device %d out of range (%d devices)", device, ndev); return B200_EINVAL; }
    auto* d = new (std::nothrow) b200_logmmse;
    if (!d) { return B200_ENOMEM; }
    const int rc = d->impl.init(sample_rate, max_block, device);
    if (rc != B200_OK) { d->impl.destroy(); delete d; return rc; }
    *out = d;
    return B200_OK;
}

void b200_logmmse_destroy(b200_logmmse* d) {
    if (!d) { return; }
    d->impl.destroy();
    delete d;
}

int b200_logmmse_reset(b200_logmmse* d) { LM_OR_FAIL(d); d->impl.needReset = true; return B200_OK; }
int b200_logmmse_set_hold(b200_logmmse* d, int hold) { LM_OR_FAIL(d); d->impl.hold = hold != 0; return B200_OK; }
int b200_logmmse_fft_size(b200_logmmse* d) { return d ? d->impl.nFFT : 0; }
int b200_logmmse_max_out(b200_logmmse* d) { return d ? d->impl.outCap : 0; }

int b200_logmmse_process(b200_logmmse* d, const float* d_in, int count, const float** d_out, int* out_count, void* stream) {
    LM_OR_FAIL(d);
    if (!out_count || (count > 0 && !d_in)) { setError("bad arguments to b200_logmmse_process"); return B200_EINVAL; }
    const int rc = d->impl.process((const cf32*)d_in, count, out_count, (cudaStream_t)stream);
    if (d_out) { *d_out = (const float*)d->impl.dOut; }
    return rc;
}

int b200_logmmse_process_host(b200_logmmse* d, const float* h_in, int count, float* h_out, int out_cap, int* out_count, void* stream) {
    LM_OR_FAIL(d);
    if (!out_count || (count > 0 && !h_in)) { setError("bad arguments to b200_logmmse_process_host"); return B200_EINVAL; }
    LogMmse& lm = d->impl;
    if (count < 0 || count > lm.maxBlock) { setError("logmmse: block of %d samples exceeds max_block %d", count, lm.maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(lm.device);
    cudaStream_t s = (cudaStream_t)stream;
    B200_TRY(lm.syncLast());
    if (!lm.dStage) { B200_CUDA(cudaMalloc((void**)&lm.dStage, (size_t)lm.maxBlock * sizeof(cf32))); lm.stageCap = lm.maxBlock; }
    if (count > 0) { B200_CUDA(cudaMemcpyAsync(lm.dStage, h_in, (size_t)count * sizeof(cf32), cudaMemcpyHostToDevice, s)); }
    B200_TRY(lm.process(lm.dStage, count, out_count, s));
    if (*out_count > out_cap) { setError("logmmse: %d output samples do not fit the buffer of %d", *out_count, out_cap); return B200_ETOOBIG; }
    if (*out_count > 0 && h_out) { B200_CUDA(cudaMemcpyAsync(h_out, lm.dOut, (size_t)*out_count * sizeof(cf32), cudaMemcpyDeviceToHost, s)); }
    B200_CUDA(cudaStreamSynchronize(s));
    return B200_OK;
}

int b200_logmmse_sync(b200_logmmse* d) { LM_OR_FAIL(d); DeviceGuard g(d->impl.device); return d->impl.syncLast(); }

// SavedParamsC::noise_mu2 to the host: returns its length, 0 before the noise was sampled
int b200_logmmse_noise(b200_logmmse* d, float* h_out, int cap) {
    LM_OR_FAIL(d);
    DeviceGuard g(d->impl.device);
    B200_TRY(d->impl.syncLast());
    if (!d->impl.sampled) { return 0; }
    if (h_out && cap > 0) { B200_CUDA(cudaMemcpy(h_out, d->impl.dMu2, (size_t)std::min(cap, d->impl.nFFT) * sizeof(float), cudaMemcpyDeviceToHost)); }
    return d->impl.nFFT;
}

} // extern "C"
