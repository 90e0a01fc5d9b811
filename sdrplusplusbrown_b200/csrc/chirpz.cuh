// Forward DFT of any length N on the power-of-two FFT passes of fft.cu (Bluestein's chirp-z form):
//     X[k] = w[k] * sum_n (x[n] w[n]) conj(w[k - n]),      w[n] = exp(-i pi n^2 / N)
// i.e. a circular convolution of length M = 2^k >= 2N - 1: two forward transforms of length M per call (the inverse one
// runs as a forward transform of the conjugate); the chirp filter's spectrum is computed once per length.
// The backward DFT is conj(forward(conj(x))).  Used where the reference hands FFTW a length that is not a power of two
// (logmmse: 2 * floor(0.02 * sampleRate), misc_modules/noise_reduction_logmmse/src/logmmse.h:274-276).
#pragma once
#include "common.cuh"
#include "fft.cuh"

namespace b200 {

struct ChirpZ {
    int device = 0;
    int N = 0, M = 0;
    float2* dChirp = nullptr;    // w[n], n < N
    cf32* dA = nullptr;          // M
    cf32* dB = nullptr;          // M: spectrum of the chirp filter
    cf32* dC = nullptr;          // M
    FrameDesc* dFrames = nullptr;
    FftPlan plan;

    int init(int n, int device);
    void destroy();
    // out[k] = scale * DFT(in)[k]; in is read as in[i] for i < nIn (zero beyond), i.e. the zero-padding of a shorter
    // input is free.  backward = true gives the unnormalised backward DFT times scale.  in / out may not alias dA / dC.
    int run(const cf32* dIn, int nIn, cf32* dOut, bool backward, float scale, cudaStream_t s);
};

} // namespace b200
