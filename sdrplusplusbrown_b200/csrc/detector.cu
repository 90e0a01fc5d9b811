// dsp::detector::SignalDetector on the device.
//
// Reference: core/src/dsp/detector/signal_detector.cpp -- run() :359-414 collects sampleRate / 10 samples, applies a
// Blackman window, transforms, swaps the halves, takes 20 log10(|X| + 1e-10) and hands the row to addSingleFFTRow
// :454-481, which turns it into "line candidates" (getLineCandidates :151-267: local-mean removal :16-50, dominant line
// spacing per bin :77-148, then per 100-bin section the spacing's best phase and the four lines before its strongest
// one), keeps 100 such rows and every tenth row averages them over time and smooths over three bins
// (aggregateAndDetect :416-452) into sigs_smoothed, which the main window plots (gui/main_window.cpp:525-526).
//
// The transform length is not a power of two (240 000 points at 2.4 MS/s), so the row is evaluated as a chirp-z
// (Bluestein) circular convolution of length M = 2^k >= 2N - 1 on the power-of-two FFT passes of fft.cu:
//     X[k] = w[k] * sum_n (x[n] win[n] w[n]) conj(w[k - n]),      w[n] = exp(-i pi n^2 / N)
// -- two forward transforms of length M per row (the inverse one runs as a forward transform of the conjugate), the
// chirp filter's spectrum is computed once per sample rate.  Everything after the dB row is evaluated in the reference's
// own operation order with IEEE single-precision adds / multiplies / divides (no contraction), so that for the same dB
// row the candidate rows, their history and sigs_smoothed equal the reference's bit for bit; the dB row itself agrees to
// FFT rounding (the reference's FFTW is single precision too).
//
// One deviation: when the sample rate changes the row length, the reference keeps its older, differently sized candidate
// rows in the history (and averages over whatever rows cover a bin); here the history restarts.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

#include "detector.cuh"

namespace b200 {

namespace {

constexpr int kNormHalf = 150;           // WINDOW / 2 of normalizeMagnitudes (:17)
constexpr int kLagMin = 8, kLagMax = 35; // the spacings getLineCandidates asks for (:167)
constexpr int kSection = 100, kSectionStep = 50;

__global__ void sdChirpInKernel(const cf32* __restrict__ x, const float* __restrict__ win, const float2* __restrict__ chirp,
                                cf32* __restrict__ a, int n, int m) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) { return; }
    cf32 o{ 0.f, 0.f };
    if (i < n) {
        const cf32 v = x[i];
        const float w = win[i];
        const float xr = __fmul_rn(v.re, w), xi = __fmul_rn(v.im, w);       // the reference's windowed fp32 sample (:376-379)
        const float2 c = chirp[i];
        o.re = xr * c.x - xi * c.y;
        o.im = xr * c.y + xi * c.x;
    }
    a[i] = o;
}

// out = conj(A * B): the inverse transform of A B is the conjugate of the forward transform of this, over M
__global__ void sdSpecMulKernel(const cf32* __restrict__ A, const cf32* __restrict__ B, cf32* __restrict__ out, int m) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) { return; }
    const cf32 a = A[i], b = B[i];
    out[i] = cf32{ a.re * b.re - a.im * b.im, -(a.re * b.im + a.im * b.re) };
}

// swapfft (utils/arrays.cpp:1128-1134) + npabsolute + 20 log10(. + 1e-10) (:388-394)
__global__ void sdRowKernel(const cf32* __restrict__ c, const float2* __restrict__ chirp, float* __restrict__ row, int n, float invM) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) { return; }
    const int half = n / 2;
    const int k = (j < half) ? j + half : ((j < 2 * half) ? j - half : j);
    const cf32 v = c[k];
    const float yr = v.re * invM, yi = -v.im * invM;
    const float2 w = chirp[k];
    const float re = yr * w.x - yi * w.y, im = yr * w.y + yi * w.x;
    const float mag = sqrtf(__fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im)));
    row[j] = __fmul_rn(20.0f, log10f(__fadd_rn(mag, 1e-10f)));
}

// normalizeMagnitudes (:16-50): row minus the mean of the (up to) 301 bins around each bin, summed in index order
__global__ void __launch_bounds__(256) sdNormalizeKernel(const float* __restrict__ row, float* __restrict__ norm, int n) {
    __shared__ float t[256 + 2 * kNormHalf];
    const int base = blockIdx.x * 256 - kNormHalf;
    for (int k = threadIdx.x; k < 256 + 2 * kNormHalf; k += 256) {
        const int j = base + k;
        t[k] = (j >= 0 && j < n) ? row[j] : 0.f;
    }
    __syncthreads();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) { return; }
    const int lo = max(0, i - kNormHalf), hi = min(n - 1, i + kNormHalf);
    float sum = 0.f;
    for (int j = lo; j <= hi; j++) { sum = __fadd_rn(sum, t[j - base]); }
    const float mean = __fdiv_rn(sum, (float)(hi - lo + 1));
    norm[i] = __fsub_rn(t[i - base], mean);
}

__device__ __forceinline__ float positivePart(float v) { return (0.0f < v) ? v : 0.0f; }

// findDominantHarmonicIntervals (:77-148) for spacings 8..35: the first spacing with the largest product of positive
// parts; also presets the candidate row to -20 (:158)
__global__ void sdLagKernel(const float* __restrict__ s, int* __restrict__ lag, float* __restrict__ cand, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const float a = positivePart(s[i]);
    float best = -INFINITY;
    int bestD = kLagMin;
#pragma unroll
    for (int d = kLagMin; d <= kLagMax; d++) {
        const float v = (i < n - d) ? __fmul_rn(a, positivePart(s[i + d])) : 0.0f;
        if (v > best) { best = v; bestD = d; }
    }
    lag[i] = bestD;
    cand[i] = -20.0f;
}

// the section loop of getLineCandidates (:176-264), one thread per 100-bin section (sections overlap by 50 bins; two
// sections that mark the same bin write the same value)
__global__ void sdSectionKernel(const float* __restrict__ norm, const int* __restrict__ lag, float* __restrict__ cand, int nSections) {
    const int sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= nSections) { return; }
    const int p = sidx * kSectionStep;
    const float* sub = norm + p;
    // median of the section's spacings: element 50 of the sorted 100 (:62-74) from a histogram over the 28 possible values
    int hist[kLagMax - kLagMin + 1];
#pragma unroll
    for (int v = 0; v <= kLagMax - kLagMin; v++) { hist[v] = 0; }
    for (int k = 0; k < kSection; k++) { hist[lag[p + k] - kLagMin]++; }
    int period = kLagMax, acc = 0;
    for (int v = 0; v <= kLagMax - kLagMin; v++) {
        acc += hist[v];
        if (acc > kSection / 2) { period = kLagMin + v; break; }
    }
    // level summed per phase of that spacing (:213-221); bin x belongs to phase (x + period - 1) mod period
    float score[kLagMax];
    for (int k = 0; k < period; k++) { score[k] = 0.0f; }
    int off = period - 1;
    for (int x = 0; x < kSection; x++) {
        score[off] = __fadd_rn(score[off], sub[x]);
        off = (off + 1 == period) ? 0 : off + 1;
    }
    int phase = 0;
    for (int k = 1; k < period; k++) { if (score[k] > score[phase]) { phase = k; } }
    // strongest line at that phase (:231-247), then the four lines before it (:249-258)
    int nIn = 0, maxi = 0;
    float top = sub[phase];
    for (int ix = phase; ix < kSection; ix += period, nIn++) {
        if (sub[ix] > top) { top = sub[ix]; maxi = nIn; }
    }
    for (int q = max(0, maxi - 4); q < maxi; q++) {
        const int ix = phase + q * period;
        cand[p + ix] = -sub[ix];
    }
}

// aggregateAndDetect (:416-452): mean over the kept rows, oldest first
__global__ void sdMeanKernel(const float* __restrict__ hist, int head, int count, int n, float* __restrict__ mean) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n) { return; }
    float sum = 0.0f;
    int r = head;
    for (int t = 0; t < count; t++) {
        sum = __fadd_rn(sum, hist[(size_t)r * n + f]);
        r = (r + 1 == SignalDetector::kHistoryRows) ? 0 : r + 1;
    }
    mean[f] = __fdiv_rn(sum, (float)count);
}

__global__ void sdSmoothKernel(const float* __restrict__ m, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    if (n == 1) { out[i] = m[i]; return; }        // (the reference reads past the end here)
    if (i == 0) { out[i] = __fdiv_rn(__fadd_rn(m[0], m[1]), 2.0f); }
    else if (i == n - 1) { out[i] = __fdiv_rn(__fadd_rn(m[n - 2], m[n - 1]), 2.0f); }
    else { out[i] = __fdiv_rn(__fadd_rn(__fadd_rn(m[i - 1], m[i]), m[i + 1]), 3.0f); }
}

inline int blocksFor(int n, int t) { return (n + t - 1) / t; }

} // namespace

int SignalDetector::init(int device_) {
    device = device_;
    return B200_OK;
}

void SignalDetector::freeBuffers() {
    void* ptrs[] = { dBuf, dWin, dChirp, dA, dB, dC, dRow, dNorm, dLag, dHist, dMean, dSmooth, dFrames };
    for (void* p : ptrs) { if (p) { cudaFree(p); } }
    dBuf = nullptr; dWin = nullptr; dChirp = nullptr; dA = nullptr; dB = nullptr; dC = nullptr; dRow = nullptr; dNorm = nullptr;
    dLag = nullptr; dHist = nullptr; dMean = nullptr; dSmooth = nullptr; dFrames = nullptr;
    plan.destroy();
}

void SignalDetector::destroy() {
    DeviceGuard g(device);
    if (everRan) { cudaStreamSynchronize(lastStream); }
    freeBuffers();
    if (dStage) { cudaFree(dStage); dStage = nullptr; stageCap = 0; }
    N = 0; M = 0;
}

// setSampleRate :297-307 + updateFFTSize :328-347 + generateWindow :349-357
int SignalDetector::setSampleRate(double sr) {
    if (sampleRate == sr) { return B200_OK; }
    if (sr > 0) {
        const double nd = sr * (1 / 10.0);
        if (nd > (double)(1 << 24)) { setError("signal detector: %g S/s needs a %g-point transform; at most %d points are served", sr, nd, 1 << 24); return B200_ETOOBIG; }
    }
    DeviceGuard g(device);
    B200_TRY(syncLast());
    sampleRate = sr;
    clear();
    if (sr <= 0) { return B200_OK; }
    const int n = (int)(sr * (1 / 10.0));
    if (n == N) { return B200_OK; }
    freeBuffers();
    N = n; M = 0;
    histCount = 0; histHead = 0; haveResult = false;
    if (N <= 0) { return B200_OK; }
    const int rc = allocate();
    if (rc != B200_OK) { freeBuffers(); N = 0; M = 0; }      // a detector without buffers passes samples through and makes no rows
    return rc;
}

int SignalDetector::allocate() {
    int m = 32;
    while (m < 2 * N - 1) { m <<= 1; }
    M = m;
    B200_TRY(plan.create(M, device));
    B200_TRY(plan.ensureScratch(1));
    B200_CUDA(cudaMalloc((void**)&dBuf, (size_t)N * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dWin, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dChirp, (size_t)N * sizeof(float2)));
    B200_CUDA(cudaMalloc((void**)&dA, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dB, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dC, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dRow, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dNorm, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dLag, (size_t)N * sizeof(int)));
    B200_CUDA(cudaMalloc((void**)&dHist, (size_t)kHistoryRows * N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dMean, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dSmooth, (size_t)N * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&dFrames, sizeof(FrameDesc)));
    // window::blackman(i, N) (dsp/window/blackman.h:5-8 over cosine.h:7-16), evaluated in double and stored as float
    std::vector<float> win((size_t)N);
    const double coefs[3] = { 0.42, 0.5, 0.08 };
    for (int i = 0; i < N; i++) {
        double w = 0.0, sign = 1.0;
        for (int k = 0; k < 3; k++) { w += sign * coefs[k] * cos((double)k * 2.0 * 3.14159265358979323846 * (double)i / (double)N); sign = -sign; }
        win[i] = (float)w;
    }
    // chirp w[n] = exp(-i pi n^2 / N): the angle from n^2 mod 2N in integers, so that it stays exact for long rows
    std::vector<float2> chirp((size_t)N);
    std::vector<cf32> filt((size_t)M, cf32{ 0.f, 0.f });
    for (int i = 0; i < N; i++) {
        const unsigned long long r = ((unsigned long long)i * (unsigned long long)i) % (2ULL * (unsigned long long)N);
        const double a = 3.14159265358979323846 * (double)r / (double)N;
        chirp[i] = make_float2((float)cos(a), (float)-sin(a));
        const cf32 b{ (float)cos(a), (float)sin(a) };           // conj(w[n]) at lags +n and -n
        filt[i] = b;
        if (i > 0) { filt[M - i] = b; }
    }
    B200_CUDA(cudaMemcpy(dWin, win.data(), win.size() * sizeof(float), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(dChirp, chirp.data(), chirp.size() * sizeof(float2), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(dA, filt.data(), filt.size() * sizeof(cf32), cudaMemcpyHostToDevice));
    FrameDesc fd; fd.src0 = dA; fd.src1 = dA; fd.len0 = M; fd.pad = 0;
    B200_CUDA(cudaMemcpy(dFrames, &fd, sizeof(fd), cudaMemcpyHostToDevice));
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dB, nullptr));
    B200_CUDA(cudaStreamSynchronize(nullptr));
    return B200_OK;
}

// setCenterFrequency :309-321
void SignalDetector::setCenterFrequency(double f) {
    if (centerFrequency == f) { return; }
    centerFrequency = f;
    clear();
}

int SignalDetector::rowFrom(const cf32* dSamples, cudaStream_t s) {
    sdChirpInKernel<<<blocksFor(M, 256), 256, 0, s>>>(dSamples, dWin, dChirp, dA, N, M);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dC, s));
    sdSpecMulKernel<<<blocksFor(M, 256), 256, 0, s>>>(dC, dB, dA, M);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dC, s));
    sdRowKernel<<<blocksFor(N, 256), 256, 0, s>>>(dC, dChirp, dRow, N, 1.0f / (float)M);
    B200_CHECK_LAUNCH(); countLaunch();
    return addRow(dRow, s);
}

// addSingleFFTRow :454-481
int SignalDetector::addRow(const float* dRowDb, cudaStream_t s) {
    int slot;
    if (histCount < kHistoryRows) { slot = (histHead + histCount) % kHistoryRows; histCount++; }
    else { slot = histHead; histHead = (histHead + 1) % kHistoryRows; }
    float* cand = dHist + (size_t)slot * N;
    sdNormalizeKernel<<<blocksFor(N, 256), 256, 0, s>>>(dRowDb, dNorm, N);
    B200_CHECK_LAUNCH(); countLaunch();
    sdLagKernel<<<blocksFor(N, 256), 256, 0, s>>>(dNorm, dLag, cand, N);
    B200_CHECK_LAUNCH(); countLaunch();
    if (N >= kSection) {      // shorter rows (and rows no longer than the largest spacing) stay at -20 (:170-173, :88-91)
        const int nSections = (N - kSection) / kSectionStep + 1;
        sdSectionKernel<<<blocksFor(nSections, 64), 64, 0, s>>>(dNorm, dLag, cand, nSections);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    sinceDetect++;
    rows++;
    if (histCount > kMinDetectRows && sinceDetect >= kDetectInterval) {
        sdMeanKernel<<<blocksFor(N, 256), 256, 0, s>>>(dHist, histHead, histCount, N, dMean);
        B200_CHECK_LAUNCH(); countLaunch();
        sdSmoothKernel<<<blocksFor(N, 256), 256, 0, s>>>(dMean, dSmooth, N);
        B200_CHECK_LAUNCH(); countLaunch();
        haveResult = true;
        sinceDetect = 0;
    }
    return B200_OK;
}

// run() :359-414: the block itself passes through unchanged, so there is nothing to write back
int SignalDetector::process(const cf32* dIn, int count, cudaStream_t s) {
    if (!enabled || N <= 0 || count <= 0) { return B200_OK; }
    DeviceGuard g(device);
    if (everRan && s != lastStream) { B200_TRY(syncLast()); }
    lastStream = s; everRan = true;
    while (count > 0) {
        if (pos == 0 && count >= N) {            // a whole row inside the block: transform it where it lies
            B200_TRY(rowFrom(dIn, s));
            dIn += N; count -= N;
            continue;
        }
        const int take = std::min(N - pos, count);
        B200_CUDA(cudaMemcpyAsync(dBuf + pos, dIn, (size_t)take * sizeof(cf32), cudaMemcpyDeviceToDevice, s));
        pos += take; dIn += take; count -= take;
        if (pos >= N) {
            B200_TRY(rowFrom(dBuf, s));
            pos = 0;
        }
    }
    return B200_OK;
}

} // namespace b200

using namespace b200;

#define SD_OR_FAIL(d) \
    if (!(d)) { setError("NULL signal detector handle"); return B200_EINVAL; }

extern "C" {

int b200_sigdet_create(b200_sigdet** out, double sample_rate, int device) {
    if (!out) { setError("bad arguments to b200_sigdet_create"); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { setError("no CUDA device available: libb200dsp has no CPU fallback"); return B200_ECUDA; }
    if (device < 0 || device >= ndev) { setError("device %d out of range (%d devices)", device, ndev); return B200_EINVAL; }
    auto* d = new (std::nothrow) b200_sigdet;
    if (!d) { return B200_ENOMEM; }
    int rc = d->impl.init(device);
    if (rc == B200_OK) { rc = d->impl.setSampleRate(sample_rate); }
    if (rc != B200_OK) { d->impl.destroy(); delete d; return rc; }
    *out = d;
    return B200_OK;
}

void b200_sigdet_destroy(b200_sigdet* d) {
    if (!d || d->ownedByFrontend) { return; }
    d->impl.destroy();
    delete d;
}

int b200_sigdet_set_sample_rate(b200_sigdet* d, double sample_rate) { SD_OR_FAIL(d); return d->impl.setSampleRate(sample_rate); }
int b200_sigdet_set_center_frequency(b200_sigdet* d, double f) { SD_OR_FAIL(d); d->impl.setCenterFrequency(f); return B200_OK; }
int b200_sigdet_set_enabled(b200_sigdet* d, int enabled) { SD_OR_FAIL(d); d->impl.enabled = enabled != 0; return B200_OK; }
int b200_sigdet_is_enabled(b200_sigdet* d) { return (d && d->impl.enabled) ? 1 : 0; }
int b200_sigdet_fft_size(b200_sigdet* d) { return d ? d->impl.N : 0; }
int64_t b200_sigdet_rows(b200_sigdet* d) { return d ? d->impl.rows : 0; }

int b200_sigdet_process(b200_sigdet* d, const float* d_in, int count, void* stream) {
    SD_OR_FAIL(d);
    if (count < 0 || (count > 0 && !d_in)) { setError("bad arguments to b200_sigdet_process"); return B200_EINVAL; }
    return d->impl.process((const cf32*)d_in, count, (cudaStream_t)stream);
}

// the same for a block in HOST memory (the stand-alone block on dsp::stream buffers): staged on the device first
int b200_sigdet_process_host(b200_sigdet* d, const float* h_in, int count, void* stream) {
    SD_OR_FAIL(d);
    if (count < 0 || (count > 0 && !h_in)) { setError("bad arguments to b200_sigdet_process_host"); return B200_EINVAL; }
    SignalDetector& sd = d->impl;
    if (!sd.enabled || sd.N <= 0 || count == 0) { return B200_OK; }
    DeviceGuard g(sd.device);
    cudaStream_t s = (cudaStream_t)stream;
    // the staging buffer is reused: the previous block's kernels must have read it
    B200_TRY(sd.syncLast());
    if (count > sd.stageCap) {
        if (sd.dStage) { cudaFree(sd.dStage); sd.dStage = nullptr; sd.stageCap = 0; }
        B200_CUDA(cudaMalloc((void**)&sd.dStage, (size_t)count * sizeof(cf32)));
        sd.stageCap = count;
    }
    B200_CUDA(cudaMemcpyAsync(sd.dStage, h_in, (size_t)count * sizeof(cf32), cudaMemcpyHostToDevice, s));
    return sd.process(sd.dStage, count, s);
}

int b200_sigdet_add_row(b200_sigdet* d, const float* d_row_db, void* stream) {
    SD_OR_FAIL(d);
    if (!d_row_db || d->impl.N <= 0) { setError("bad arguments to b200_sigdet_add_row"); return B200_EINVAL; }
    DeviceGuard g(d->impl.device);
    cudaStream_t s = (cudaStream_t)stream;
    if (d->impl.everRan && s != d->impl.lastStream) { B200_TRY(d->impl.syncLast()); }
    d->impl.lastStream = s; d->impl.everRan = true;
    return d->impl.addRow(d_row_db, s);
}

int b200_sigdet_sync(b200_sigdet* d) { SD_OR_FAIL(d); DeviceGuard g(d->impl.device); return d->impl.syncLast(); }

int b200_sigdet_result(b200_sigdet* d, float* h_out, int cap) {
    SD_OR_FAIL(d);
    DeviceGuard g(d->impl.device);
    B200_TRY(d->impl.syncLast());
    if (!d->impl.haveResult) { return 0; }
    if (h_out && cap > 0) { B200_CUDA(cudaMemcpy(h_out, d->impl.dSmooth, (size_t)std::min(cap, d->impl.N) * sizeof(float), cudaMemcpyDeviceToHost)); }
    return d->impl.N;
}

const float* b200_sigdet_result_ptr(b200_sigdet* d) { return (d && d->impl.haveResult) ? d->impl.dSmooth : nullptr; }
const float* b200_sigdet_last_row_ptr(b200_sigdet* d) { return d ? d->impl.dRow : nullptr; }

int b200_sigdet_history(b200_sigdet* d, int row, float* h_out, int cap) {
    SD_OR_FAIL(d);
    DeviceGuard g(d->impl.device);
    B200_TRY(d->impl.syncLast());
    if (h_out && cap > 0 && row >= 0 && row < d->impl.histCount) {
        B200_CUDA(cudaMemcpy(h_out, d->impl.histRow(row), (size_t)std::min(cap, d->impl.N) * sizeof(float), cudaMemcpyDeviceToHost));
    }
    return d->impl.histCount;
}

} // extern "C"
