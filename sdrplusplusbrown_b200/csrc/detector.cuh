// dsp::detector::SignalDetector on the device: the front end's detector pre-processor
// (reference: core/src/dsp/detector/signal_detector.cpp, wired in signal_path/iq_frontend.cpp:37-44).
#pragma once
#include <cstdint>

#include "common.cuh"
#include "fft.cuh"

namespace b200 {

struct SignalDetector {
    static constexpr int kHistoryRows = 100;     // N_FFT_ROWS: 10 s of rows (signal_detector.h:103)
    static constexpr int kMinDetectRows = 20;    // MIN_DETECT_FFT_ROWS (strictly more are needed)
    static constexpr int kDetectInterval = 10;   // DETECT_INTERVAL_FRAMES

    int device = 0;
    double sampleRate = 0.0, centerFrequency = 0.0;
    bool enabled = true;
    int N = 0;                   // (int)(sampleRate / 10): any length
    int M = 0;                   // power of two >= 2N - 1: the circular convolution length of the chirp-z evaluation
    int pos = 0;                 // samples waiting in dBuf
    int sinceDetect = 0;
    int64_t rows = 0;
    int histCount = 0, histHead = 0;
    bool haveResult = false;
    cudaStream_t lastStream = nullptr; bool everRan = false;

    cf32* dBuf = nullptr;        // N samples of the row being collected
    float* dWin = nullptr;       // Blackman window, N
    float2* dChirp = nullptr;    // exp(-i pi n^2 / N), n < N
    cf32* dA = nullptr;          // M: chirped, windowed, zero-padded samples; later conj(A B)
    cf32* dB = nullptr;          // M: spectrum of the chirp filter
    cf32* dC = nullptr;          // M: transform output
    float* dRow = nullptr;       // N: dB row
    float* dNorm = nullptr;      // N: row minus its local mean
    int* dLag = nullptr;         // N: dominant line spacing per bin
    float* dHist = nullptr;      // kHistoryRows x N candidate rows (ring)
    float* dMean = nullptr;      // N
    float* dSmooth = nullptr;    // N: sigs_smoothed
    FrameDesc* dFrames = nullptr;
    cf32* dStage = nullptr; int stageCap = 0;    // b200_sigdet_process_host
    FftPlan plan;

    int init(int device);
    void destroy();
    void freeBuffers();
    int syncLast() { if (everRan) { B200_CUDA(cudaStreamSynchronize(lastStream)); } return B200_OK; }
    int setSampleRate(double sr);
    int allocate();              // buffers, window, chirp and chirp-filter spectrum for the current N
    void setCenterFrequency(double f);
    void clear() { pos = 0; sinceDetect = 0; }
    int process(const cf32* dIn, int count, cudaStream_t s);
    int rowFrom(const cf32* dSamples, cudaStream_t s);       // window + transform + dB row, then addRow
    int addRow(const float* dRowDb, cudaStream_t s);         // addSingleFFTRow
    float* histRow(int chronological) { return dHist + (size_t)((histHead + chronological) % kHistoryRows) * N; }
};

} // namespace b200

struct b200_sigdet { b200::SignalDetector impl; bool ownedByFrontend = false; };
