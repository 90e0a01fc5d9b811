// Windowed forward FFT -> log power.  Replaces IQFrontEnd::handler (signal_path/iq_frontend.cpp:267-287):
//   :271  fftIn[i] = data[i] * window[i]   (window carries the (-1)^i fft-shift, :303-311; tail zero, :317)
//   :274  forward c2c FFT of size N (FFTW in the reference, utils/arrays.cpp:533-546)
//   :282  out[k] = 10*log10((re^2+im^2)/N^2) (VOLK power spectrum, log2 clamped at +-127)
// N <= 8192: one CTA per frame, whole transform in shared memory.
// N >  8192: N = N1 x N2 in two kernels; the N1 x N2 intermediate lives in a small scratch ring that
// stays resident in the 126 MB L2, so HBM sees the samples once and the dB row once.
// Radix-2 DIT on bit-reversed shared-memory rows, twiddles from float64-designed tables.
#include "fft.cuh"

#include <cmath>

namespace b200 {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

__device__ __forceinline__ float2 frameSample(const FrameDesc& f, const float* __restrict__ win, int i, int nz) {
    if (i >= nz) { return make_float2(0.f, 0.f); }
    const float2 x = (i < f.len0) ? __ldg(reinterpret_cast<const float2*>(f.src0) + i)
                                  : __ldg(reinterpret_cast<const float2*>(f.src1) + (i - f.len0));
    const float w = win ? __ldg(win + i) : 1.0f;
    return make_float2(x.x * w, x.y * w);
}

// nb independent length-L transforms stored as rows of pitch LP in shared memory, inputs already in
// bit-reversed order.  tw[j] = exp(-2*pi*i*j/L), j < L/2.
__device__ void fftRows(float2* s, int L, int logL, int LP, int nb, const float2* __restrict__ tw) {
    const int halfL = L >> 1;
    const int total = nb * halfL;
    for (int lg = 1; lg <= logL; lg++) {
        const int half = 1 << (lg - 1);
        const int twStride = L >> lg;
        for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
            const int b = idx / halfL;
            const int r = idx - b * halfL;
            const int k = r & (half - 1);
            const int i0 = b * LP + ((r >> (lg - 1)) << lg) + k;
            const int i1 = i0 + half;
            const float2 w = __ldg(tw + k * twStride);
            const float2 a = s[i0];
            const float2 t = cmul(s[i1], w);
            s[i0] = make_float2(a.x + t.x, a.y + t.y);
            s[i1] = make_float2(a.x - t.x, a.y - t.y);
        }
        __syncthreads();
    }
}

__device__ __forceinline__ float powerDb(float2 X, float invN) {
    const float r = X.x * invN, q = X.y * invN;
    float l = log2f(r * r + q * q);
    if (isinf(l)) { l = copysignf(127.0f, l); }
    return 3.01029995663981209120f * l;
}

// ---- N <= 8192: one CTA per frame
__global__ void __launch_bounds__(kThreads) fftSingleKernel(const FrameDesc* __restrict__ frames, const float* __restrict__ win, int nz,
                                                            int N, int logN, const float2* __restrict__ tw, float* __restrict__ outDb,
                                                            cf32* __restrict__ outSpec) {
    extern __shared__ __align__(16) float2 sm[];
    const FrameDesc f = frames[blockIdx.x];
    for (int i = threadIdx.x; i < N; i += kThreads) {
        sm[__brev((unsigned)i) >> (32 - logN)] = frameSample(f, win, i, nz);
    }
    __syncthreads();
    fftRows(sm, N, logN, N, 1, tw);
    const float invN = 1.0f / (float)N;
    for (int k = threadIdx.x; k < N; k += kThreads) {
        const float2 X = sm[k];
        if (outDb) { outDb[(size_t)blockIdx.x * N + k] = powerDb(X, invN); }
        if (outSpec) { reinterpret_cast<float2*>(outSpec)[(size_t)blockIdx.x * N + k] = X; }
    }
}

// ---- N = N1 * N2, pass 1: C adjacent columns n2 per CTA, length-N1 transforms over n1 (input stride N2),
// then the inter-pass twiddle W_N^(n2*k1); result T[k1][n2].
__global__ void __launch_bounds__(kThreads) fftPass1Kernel(const FrameDesc* __restrict__ frames, const float* __restrict__ win, int nz,
                                                           int N1, int logN1, int N2, int C, const float2* __restrict__ twN1,
                                                           const float2* __restrict__ twLo, const float2* __restrict__ twHi, int logS,
                                                           cf32* __restrict__ scratch, int scratchFrames) {
    extern __shared__ __align__(16) float2 sm[];
    const FrameDesc f = frames[blockIdx.y];
    const int c0 = blockIdx.x * C;
    const int LP = N1 + 1;
    const int total = N1 * C;
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int c = idx % C, n1 = idx / C;
        sm[c * LP + (__brev((unsigned)n1) >> (32 - logN1))] = frameSample(f, win, n1 * N2 + c0 + c, nz);
    }
    __syncthreads();
    fftRows(sm, N1, logN1, LP, C, twN1);
    float2* T = reinterpret_cast<float2*>(scratch) + (size_t)(blockIdx.y % scratchFrames) * ((size_t)N1 * N2);
    const int maskS = (1 << logS) - 1;
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int c = idx % C, k1 = idx / C;
        const int m = (c0 + c) * k1;                     // < N
        const float2 w = cmul(__ldg(twLo + (m & maskS)), __ldg(twHi + (m >> logS)));
        T[(size_t)k1 * N2 + c0 + c] = cmul(sm[c * LP + k1], w);
    }
}

// pass 2: R adjacent rows k1 per CTA, length-N2 transforms over n2; X[k1 + N1*k2].
__global__ void __launch_bounds__(kThreads) fftPass2Kernel(int N1, int N2, int logN2, int R, const float2* __restrict__ twN2,
                                                           const cf32* __restrict__ scratch, int scratchFrames, int frameBase,
                                                           float* __restrict__ outDb, cf32* __restrict__ outSpec) {
    extern __shared__ __align__(16) float2 sm[];
    const int frame = frameBase + blockIdx.y;
    const int r0 = blockIdx.x * R;
    const int LP = N2 + 1;
    const int total = N2 * R;
    const float2* T = reinterpret_cast<const float2*>(scratch) + (size_t)(frame % scratchFrames) * ((size_t)N1 * N2);
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int n2 = idx % N2, r = idx / N2;
        sm[r * LP + (__brev((unsigned)n2) >> (32 - logN2))] = T[(size_t)(r0 + r) * N2 + n2];
    }
    __syncthreads();
    fftRows(sm, N2, logN2, LP, R, twN2);
    const size_t N = (size_t)N1 * N2;
    const float invN = 1.0f / (float)N;
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int r = idx % R, k2 = idx / R;
        const float2 X = sm[r * LP + k2];
        const size_t k = (size_t)k2 * N1 + r0 + r;
        if (outDb) { outDb[(size_t)frame * N + k] = powerDb(X, invN); }
        if (outSpec) { reinterpret_cast<float2*>(outSpec)[(size_t)frame * N + k] = X; }
    }
}

int uploadTwiddles(float2** d, int count, int N, int stride) {
    std::vector<float2> h((size_t)std::max(count, 1));
    for (int j = 0; j < count; j++) {
        const double a = -2.0 * 3.14159265358979323846 * (double)j * (double)stride / (double)N;
        h[j] = make_float2((float)cos(a), (float)sin(a));
    }
    B200_CUDA(cudaMalloc((void**)d, h.size() * sizeof(float2)));
    B200_CUDA(cudaMemcpy(*d, h.data(), h.size() * sizeof(float2), cudaMemcpyHostToDevice));
    return B200_OK;
}

int pickBatch(int L, int want, size_t smemCap) {
    int b = want;
    while (b > 1 && (size_t)b * (L + 1) * sizeof(float2) > smemCap) { b >>= 1; }
    return b;
}

constexpr size_t kSmemCap = 160 * 1024;

} // namespace

int FftPlan::create(int N_, int device_) {
    destroy();
    device = device_;
    N = N_;
    logN = 0;
    while ((1 << logN) < N) { logN++; }
    if ((1 << logN) != N || logN < 5 || logN > 22) { setError("fft size %d is not a power of two in [32, 4194304]", N); return B200_EINVAL; }
    if (logN <= 13) {
        N1 = 1; N2 = N;
        B200_TRY(uploadTwiddles(&dTwN2, N / 2, N, 1));
    }
    else {
        const int l2 = (logN + 1) / 2;
        N2 = 1 << l2; N1 = N >> l2;
        B200_TRY(uploadTwiddles(&dTwN1, N1 / 2, N1, 1));
        B200_TRY(uploadTwiddles(&dTwN2, N2 / 2, N2, 1));
        logS = l2; S = 1 << logS;
        B200_TRY(uploadTwiddles(&dTwLo, S, N, 1));
        B200_TRY(uploadTwiddles(&dTwHi, N >> logS, N, S));
    }
    return B200_OK;
}

void FftPlan::destroy() {
    float2** t[] = { &dTwN1, &dTwN2, &dTwLo, &dTwHi };
    for (auto p : t) { if (*p) { cudaFree(*p); *p = nullptr; } }
    if (dScratch) { cudaFree(dScratch); dScratch = nullptr; }
    scratchFrames = 0;
}

int FftPlan::ensureScratch(int frames) {
    if (N1 == 1) { return B200_OK; }
    // keep the ring small enough to stay in L2 (~126 MB): at most 32 MB of intermediates
    int want = (int)std::max<size_t>(1, std::min<size_t>((size_t)frames, (32u << 20) / ((size_t)N * sizeof(cf32))));
    if (want <= scratchFrames) { return B200_OK; }
    if (dScratch) { cudaFree(dScratch); dScratch = nullptr; }
    B200_CUDA(cudaMalloc((void**)&dScratch, (size_t)want * N * sizeof(cf32)));
    scratchFrames = want;
    return B200_OK;
}

int FftPlan::run(const FrameDesc* dFrames, int nFrames, const float* dWindow, int nz, float* outDb, cf32* outSpec, cudaStream_t s) {
    if (nFrames <= 0) { return B200_OK; }
    if (N1 == 1) {
        const size_t smem = (size_t)N * sizeof(float2);
        B200_CUDA(cudaFuncSetAttribute(fftSingleKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fftSingleKernel<<<nFrames, kThreads, smem, s>>>(dFrames, dWindow, nz, N, logN, dTwN2, outDb, outSpec);
        B200_CHECK_LAUNCH();
        countLaunch();
        return B200_OK;
    }
    B200_TRY(ensureScratch(nFrames));
    int logN1 = 0, logN2 = 0;
    while ((1 << logN1) < N1) { logN1++; }
    while ((1 << logN2) < N2) { logN2++; }
    const int C = pickBatch(N1, 16, kSmemCap);
    const int R = pickBatch(N2, 32, kSmemCap);
    const size_t smem1 = (size_t)C * (N1 + 1) * sizeof(float2), smem2 = (size_t)R * (N2 + 1) * sizeof(float2);
    B200_CUDA(cudaFuncSetAttribute(fftPass1Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    B200_CUDA(cudaFuncSetAttribute(fftPass2Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    // frames are processed in groups that fit the scratch ring, pass 2 right behind pass 1 (L2 reuse)
    for (int f0 = 0; f0 < nFrames; f0 += scratchFrames) {
        const int nf = std::min(scratchFrames, nFrames - f0);
        // the ring index is (frame % scratchFrames); groups start at multiples of scratchFrames
        fftPass1Kernel<<<dim3(N2 / C, nf), kThreads, smem1, s>>>(dFrames + f0, dWindow, nz, N1, logN1, N2, C, dTwN1, dTwLo, dTwHi, logS,
                                                                 dScratch, scratchFrames);
        B200_CHECK_LAUNCH();
        fftPass2Kernel<<<dim3(N1 / R, nf), kThreads, smem2, s>>>(N1, N2, logN2, R, dTwN2, dScratch, scratchFrames, f0, outDb, outSpec);
        B200_CHECK_LAUNCH();
        countLaunch(2);
    }
    return B200_OK;
}

} // namespace b200
