// dsp::compression::ExperimentalFFTCompressor on the GPU (reference: core/src/dsp/compression/experimental_fft_compressor.h),
// the block the SDR++ server runs between its resampler and the PCM sample-packet compressor (core/src/server.cpp:101-110).
// Same members the server touches -- init, setSampleRate / getSampleRate, setEnabled / isEnabled, setMaskedFrequencies,
// setTxMode, the public `lossRate`, `fftSize`, `noiseFigure` + `sharedDataLock` (server.cpp:470-474), run() -- over
// b200_fftcomp_* of libb200dsp.  Disabled it copies its input through (:435-441); enabled it emits one frame of
// `fftSize` frequency-domain samples per `fftSize` input samples, ten frames late.  setTxMode is recorded only: the device
// implementation keeps filtering as in receive mode (the reference drops the signal mask while transmitting, :152-159).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "../processor.h"

namespace dsp::compression {

    class ExperimentalFFTCompressor : public Processor<complex_t, complex_t> {
        using base_type = Processor<complex_t, complex_t>;
    public:
        int sampleRate = 0;
        int fftSize = 1024;
        double lossRate = 1.0;
        std::vector<float> noiseFigure;
        std::mutex sharedDataLock;
        bool enabled = false;
        bool txMode = false;

        ExperimentalFFTCompressor() {}
        ExperimentalFFTCompressor(stream<complex_t>* in) { init(in); }
        ~ExperimentalFFTCompressor() override {
            if (base_type::_block_init) { base_type::stop(); }
            if (comp) { b200_fftcomp_destroy(comp); }
        }

        int getSampleRate() { return sampleRate; }
        void setSampleRate(int sr) {
            std::lock_guard<std::mutex> lk(mtx);
            if (sr == sampleRate) { return; }
            if (comp) { b200_fftcomp_destroy(comp); comp = nullptr; }
            sampleRate = sr;
            if (b200_fftcomp_create(&comp, sr, lossRate, STREAM_BUFFER_SIZE, 0) != B200_OK) { fprintf(stderr, "ExperimentalFFTCompressor: %s\n", b200_last_error()); comp = nullptr; return; }
            fftSize = b200_fftcomp_fft_size(comp);
            if (!masked.empty()) { b200_fftcomp_set_masked(comp, masked.data(), (int)masked.size()); }
            std::lock_guard<std::mutex> dl(sharedDataLock);
            noiseFigure.clear();
        }

        void setEnabled(bool e) { enabled = e; }
        bool isEnabled() { return enabled; }
        void setTxMode(int tx) { txMode = tx; }
        void setMaskedFrequencies(const std::vector<int32_t> maskedFrequencies) {
            std::lock_guard<std::mutex> lk(mtx);
            masked = maskedFrequencies;
            if (comp) { b200_fftcomp_set_masked(comp, masked.data(), (int)masked.size()); }
        }

        int run() override {
            int count = base_type::_in->read();
            if (count < 0) { return -1; }
            if (!enabled || !comp) {
                memcpy(this->out.writeBuf, _in->readBuf, count * sizeof(complex_t));
                base_type::_in->flush();
                if (count > 0) { if (!base_type::out.swap(count)) { return -1; } }
                return count;
            }
            int outCount = 0;
            {
                std::lock_guard<std::mutex> lk(mtx);
                b200_fftcomp_set_loss_rate(comp, lossRate);
                if (b200_fftcomp_process_host(comp, reinterpret_cast<const float*>(_in->readBuf), count, reinterpret_cast<float*>(this->out.writeBuf), &outCount, nullptr) != B200_OK) {
                    fprintf(stderr, "ExperimentalFFTCompressor: %s\n", b200_last_error());
                    outCount = 0;
                }
                if (outCount) {
                    float nf[30];
                    const int k = b200_fftcomp_noise_figure(comp, nf, 30);
                    if (k > 0) { std::lock_guard<std::mutex> dl(sharedDataLock); noiseFigure.assign(nf, nf + k); }
                }
            }
            base_type::_in->flush();
            if (outCount) { if (!base_type::out.swap(outCount)) { return -1; } }
            return outCount;
        }

    private:
        std::mutex mtx;
        b200_fftcomp* comp = nullptr;
        std::vector<int32_t> masked;
    };
}
