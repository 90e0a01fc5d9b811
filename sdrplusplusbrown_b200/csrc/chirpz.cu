// Forward / backward DFT of any length on the power-of-two FFT passes (see chirpz.cuh).
#include <cmath>
#include <vector>

#include "chirpz.cuh"

namespace b200 {

namespace {

// a[i] = x[i] w[i] (conjugated input for the backward transform), zero beyond nIn
__global__ void czInKernel(const cf32* __restrict__ x, int nIn, const float2* __restrict__ chirp, cf32* __restrict__ a, int m, int conjIn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) { return; }
    cf32 o{ 0.f, 0.f };
    if (i < nIn) {
        const cf32 v = x[i];
        const float xr = v.re, xi = conjIn ? -v.im : v.im;
        const float2 c = chirp[i];
        o.re = xr * c.x - xi * c.y;
        o.im = xr * c.y + xi * c.x;
    }
    a[i] = o;
}

// out = conj(A * B): the inverse transform of A B is the conjugate of the forward transform of this, over M
__global__ void czMulKernel(const cf32* __restrict__ A, const cf32* __restrict__ B, cf32* __restrict__ out, int m) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) { return; }
    const cf32 a = A[i], b = B[i];
    out[i] = cf32{ a.re * b.re - a.im * b.im, -(a.re * b.im + a.im * b.re) };
}

__global__ void czOutKernel(const cf32* __restrict__ c, const float2* __restrict__ chirp, cf32* __restrict__ out, int n, float scale, int conjOut) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) { return; }
    const cf32 v = c[k];
    const float yr = v.re * scale, yi = -v.im * scale;
    const float2 w = chirp[k];
    const float re = yr * w.x - yi * w.y, im = yr * w.y + yi * w.x;
    out[k] = cf32{ re, conjOut ? -im : im };
}

inline int blocksFor(int n, int t) { return (n + t - 1) / t; }

} // namespace

int ChirpZ::init(int n, int device_) {
    destroy();
    device = device_;
    if (n <= 0) { setError("transform length %d", n); return B200_EINVAL; }
    if (n > (1 << 24)) { setError("transform of %d points: at most %d are served", n, 1 << 24); return B200_ETOOBIG; }
    N = n;
    int m = 32;
    while (m < 2 * N - 1) { m <<= 1; }
    M = m;
    B200_TRY(plan.create(M, device));
    B200_CUDA(cudaMalloc((void**)&dChirp, (size_t)N * sizeof(float2)));
    B200_CUDA(cudaMalloc((void**)&dA, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dB, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dC, (size_t)M * sizeof(cf32)));
    B200_CUDA(cudaMalloc((void**)&dFrames, sizeof(FrameDesc)));
    // the angle from n^2 mod 2N in integers, so that it stays exact for long transforms
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
    B200_CUDA(cudaMemcpy(dChirp, chirp.data(), chirp.size() * sizeof(float2), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(dA, filt.data(), filt.size() * sizeof(cf32), cudaMemcpyHostToDevice));
    FrameDesc fd; fd.src0 = dA; fd.src1 = dA; fd.len0 = M; fd.pad = 0;
    B200_CUDA(cudaMemcpy(dFrames, &fd, sizeof(fd), cudaMemcpyHostToDevice));
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dB, nullptr));
    B200_CUDA(cudaStreamSynchronize(nullptr));
    return B200_OK;
}

void ChirpZ::destroy() {
    void* ptrs[] = { dChirp, dA, dB, dC, dFrames };
    for (void* p : ptrs) { if (p) { cudaFree(p); } }
    dChirp = nullptr; dA = nullptr; dB = nullptr; dC = nullptr; dFrames = nullptr;
    plan.destroy();
    N = 0; M = 0;
}

int ChirpZ::run(const cf32* dIn, int nIn, cf32* dOut, bool backward, float scale, cudaStream_t s) {
    czInKernel<<<blocksFor(M, 256), 256, 0, s>>>(dIn, nIn < N ? nIn : N, dChirp, dA, M, backward ? 1 : 0);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dC, s));
    czMulKernel<<<blocksFor(M, 256), 256, 0, s>>>(dC, dB, dA, M);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(plan.run(dFrames, 1, nullptr, M, nullptr, dC, s));
    czOutKernel<<<blocksFor(N, 256), 256, 0, s>>>(dC, dChirp, dOut, N, scale / (float)M, backward ? 1 : 0);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

} // namespace b200
