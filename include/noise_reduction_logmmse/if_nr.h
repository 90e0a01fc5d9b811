// dsp::IFNRLogMMSE on the GPU (reference: misc_modules/noise_reduction_logmmse/src/if_nr.h:15-184), the processor the
// logmmse module adds to the front end's pre-processing chain (main.cpp:165: sigpath::iqFrontEnd.addPreprocessor(&ifnrProcessor, false)).
// Same members a caller touches -- init / setInput / setHold / reset / process / run, `bypass`, `stopReason`,
// `params.forceSampleRate` / `params.hold` -- over b200_logmmse_* of libb200dsp.  The block works on host dsp::stream
// buffers like every plug-in pre-processor (IQFrontEnd::addPreprocessor in include/dsp/iq_frontend.h): a block goes to the
// device, the de-noised frames come back (0 samples until the noise has been sampled, then whole hops).
//
// The reference asks the application for the rate (sigpath::iqFrontEnd.getSampleRate(), if_nr.h:72) unless
// params.forceSampleRate is set; here `sampleRateSource` stands for that global: point it at the front end,
//     nr.sampleRateSource = [&] { return (int)iqFrontEnd.getSampleRate(); };
#pragma once
#include <cstdio>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>

#include "../dsp/processor.h"

namespace dsp {

    class IFNRLogMMSE : public Processor<complex_t, complex_t> {
        using base_type = Processor<complex_t, complex_t>;
    public:
        IFNRLogMMSE() {}
        ~IFNRLogMMSE() override {
            if (base_type::_block_init) { base_type::stop(); }
            if (nr) { b200_logmmse_destroy(nr); }
        }

        void init(stream<complex_t>* in) override { base_type::init(in); }

        void setInput(stream<complex_t>* in) override {
            if (!_block_init) { init(in); }
            else { Processor::setInput(in); }
        }

        struct Params { int forceSampleRate = 0; bool hold = false; } params;
        std::function<int()> sampleRateSource;
        int freq = 192000;
        std::mutex freqMutex;
        bool bypass = true;
        std::string stopReason = "";

        void setHold(bool hold) { params.hold = hold; }
        void reset() { shouldReset = true; }

        // if_nr.h:63-113
        void process(complex_t* in, int count, complex_t* out, int& outCount) {
            std::lock_guard<std::mutex> lk(freqMutex);
            outCount = 0;
            if (shouldReset) {
                shouldReset = false;
                int f = params.forceSampleRate;
                if (f == 0 && sampleRateSource) { f = sampleRateSource(); }
                if (f <= 0) { fprintf(stderr, "IFNRLogMMSE: no sample rate (set params.forceSampleRate or sampleRateSource)\n"); return; }
                if (!nr || f != freq) {
                    if (nr) { b200_logmmse_destroy(nr); nr = nullptr; }
                    freq = f;
                    if (b200_logmmse_create(&nr, freq, STREAM_BUFFER_SIZE, 0) != B200_OK) { fprintf(stderr, "IFNRLogMMSE: %s\n", b200_last_error()); nr = nullptr; return; }
                }
                else { b200_logmmse_reset(nr); }
            }
            if (!nr) { return; }
            b200_logmmse_set_hold(nr, params.hold ? 1 : 0);
            if (b200_logmmse_process_host(nr, reinterpret_cast<const float*>(in), count, reinterpret_cast<float*>(out), STREAM_BUFFER_SIZE, &outCount, nullptr) != B200_OK) {
                fprintf(stderr, "IFNRLogMMSE: %s\n", b200_last_error());
                outCount = 0;
            }
        }

        int run() override {
            int count = _in->read();
            if (count < 0) { return -1; }
            int outCount = 0;
            process(_in->readBuf, count, out.writeBuf, outCount);
            _in->flush();
            if (!out.swap(outCount)) { return -1; }
            return count;
        }

    private:
        b200_logmmse* nr = nullptr;
        bool shouldReset = true;
    };
}
