// TEST INFRASTRUCTURE ONLY (oracle, reference side).  Part of oracle/_ref/libref_oracle.so.
//
// The reference's own SignalDetector, compiled from the sources where they lie:
//   $REF_ROOT/core/src/dsp/detector/signal_detector.cpp   (included below, so that its file-local helpers
//                                                          getLineCandidates / normalizeMagnitudes are reachable)
//   $REF_ROOT/core/src/utils/arrays.cpp                   (FFT plan + array helpers; linked by oracle/Makefile)
// against oracle/shim (generic-C VOLK stand-in; <fftw3.h> served by the float64 DFT of oracle/fft64.h).
// The detector's run() is driven through its real streams, single-threaded: one swap() into the input, one run(),
// one read()/flush() of the output per block.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <iomanip>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>
#include <thread>
#include <mutex>
#include <condition_variable>

#include <dsp/processor.h>
#include <utils/arrays.h>

#define private public
#include <dsp/detector/signal_detector.cpp>
#undef private

#include "oracle_api.h"

namespace {
struct RefSigDet {
    dsp::stream<dsp::complex_t> in;
    dsp::detector::SignalDetector det;
    int rows = 0;
};
}

extern "C" {

void* orc_sigdet_create(double sampleRate) {
    auto* h = new RefSigDet;
    h->det.init(&h->in);
    h->det.setSampleRate(sampleRate);
    return h;
}
void orc_sigdet_destroy(void* h) { delete (RefSigDet*)h; }
int orc_sigdet_fft_size(void* h) { return ((RefSigDet*)h)->det.fftSize; }
void orc_sigdet_set_center_frequency(void* h, double f) { ((RefSigDet*)h)->det.setCenterFrequency(f); }

int orc_sigdet_process(void* hv, int count, const float* in) {
    auto* h = (RefSigDet*)hv;
    const auto* x = (const dsp::complex_t*)in;
    int done = 0;
    while (done < count) {
        const int n = std::min(count - done, (int)STREAM_BUFFER_SIZE);
        memcpy(h->in.writeBuf, x + done, (size_t)n * sizeof(dsp::complex_t));
        h->in.swap(n);
        // a row completes each time the sample buffer fills (bufferPos returns to 0)
        const int posBefore = h->det.bufferPos;
        h->det.run();
        if (h->det.fftSize > 0 && h->det.isEnabled()) { h->rows += (posBefore + n) / h->det.fftSize; }
        h->det.out.read();
        h->det.out.flush();
        done += n;
    }
    return h->rows;
}

int orc_sigdet_result(void* hv, float* out, int cap) {
    auto& r = ((RefSigDet*)hv)->det.sigs_smoothed;
    const int n = (int)r.size();
    if (out) { memcpy(out, r.data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}

int orc_sigdet_history(void* hv, int row, float* out, int cap) {
    auto& hst = ((RefSigDet*)hv)->det.suppressedCarrierCandidates;
    if (out && row >= 0 && row < (int)hst.size()) {
        memcpy(out, hst[row]->data(), std::min((size_t)cap, hst[row]->size()) * sizeof(float));
    }
    return (int)hst.size();
}

void orc_sigdet_row(const float* iq, int N, float* rowDb) {
    // the row part of run() with the detector's own members: window, plan, swapfft, npabsolute, 20 log10
    dsp::detector::SignalDetector d;
    d.sampleRate = (double)N * 10.0;
    d.fftSize = N;
    d.fftWindowBuf = new float[N];
    d.fftInArray = std::make_shared<std::vector<dsp::complex_t>>(N);
    d.fftPlan = dsp::arrays::allocateFFTWPlan(false, N);
    d.generateWindow();
    const auto* x = (const dsp::complex_t*)iq;
    auto& inVec = *d.fftInArray;
    for (int j = 0; j < N; j++) {
        inVec[j].re = x[j].re * d.fftWindowBuf[j];
        inVec[j].im = x[j].im * d.fftWindowBuf[j];
    }
    dsp::arrays::npfftfft(d.fftInArray, d.fftPlan);
    dsp::arrays::swapfft(d.fftPlan->getOutput());
    auto mag = dsp::arrays::npabsolute(d.fftPlan->getOutput());
    for (int j = 0; j < N; j++) { rowDb[j] = 20.0f * log10f(mag->at(j) + 1e-10f); }
    delete[] d.fftWindowBuf;
    d.fftWindowBuf = nullptr;
}

void orc_sigdet_candidates(const float* rowDb, int N, float* out) {
    auto c = dsp::detector::getLineCandidates(dsp::detector::ArrayView<float>(rowDb, (size_t)N));
    memcpy(out, c->data(), c->size() * sizeof(float));
}

void orc_sigdet_add_row(void* hv, const float* rowDb, int N) {
    auto* h = (RefSigDet*)hv;
    h->det.addSingleFFTRow(dsp::detector::ArrayView<float>(rowDb, (size_t)N));
    h->rows++;
}

}
