// TEST INFRASTRUCTURE ONLY (oracle, reference side).  Part of oracle/_ref/libref_oracle.so.
//
// The reference's own dsp::compression::ExperimentalFFTCompressor (core/src/dsp/compression/experimental_fft_compressor.h),
// the block the SDR++ server puts between its resampler and the PCM sample-packet compressor (core/src/server.cpp:101-110),
// compiled from the header where it lies (with utils/arrays.cpp for the array helpers; FFTW served by oracle/fft64.h).
// Driven through its own buffers: samples are appended to `inputBuffer`, process() is called once per block as run() does
// (:431-452) and what it leaves in out.writeBuf is returned.
#include <algorithm>
#include <cstring>
#include <vector>

#include <dsp/compression/experimental_fft_compressor.h>

#include "oracle_api.h"

extern "C" {

void* orc_fftcomp_create(int sampleRate, double lossRate) {
    auto* c = new dsp::compression::ExperimentalFFTCompressor;
    c->setSampleRate(sampleRate);
    c->lossRate = lossRate;
    c->setEnabled(true);
    return c;
}
void orc_fftcomp_destroy(void* h) { delete (dsp::compression::ExperimentalFFTCompressor*)h; }
int orc_fftcomp_fft_size(void* h) { return ((dsp::compression::ExperimentalFFTCompressor*)h)->fftSize; }
void orc_fftcomp_set_masked(void* h, const int* fromTo, int n) {
    ((dsp::compression::ExperimentalFFTCompressor*)h)->setMaskedFrequencies(std::vector<int32_t>(fromTo, fromTo + n));
}
int orc_fftcomp_process(void* hv, int count, const float* in, float* out) {
    auto* c = (dsp::compression::ExperimentalFFTCompressor*)hv;
    const auto* x = (const dsp::complex_t*)in;
    c->inputBuffer.insert(c->inputBuffer.end(), x, x + count);
    const int n = c->process();
    if (n > 0) { memcpy(out, c->out.writeBuf, (size_t)n * sizeof(dsp::complex_t)); }
    return n;
}
int orc_fftcomp_noise_figure(void* hv, float* out, int cap) {
    auto* c = (dsp::compression::ExperimentalFFTCompressor*)hv;
    const int n = (int)c->noiseFigure.size();
    if (out) { memcpy(out, c->noiseFigure.data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}

}
