// dsp::detector::SignalDetector on the GPU (reference: core/src/dsp/detector/signal_detector.{h,cpp}).
// Same public surface -- init / setSampleRate / setCenterFrequency / setEnabled / isEnabled / run and the public
// `sigs_smoothed` the main window plots (gui/main_window.cpp:525-526) -- over b200_sigdet_* of libb200dsp.
//
// Two ways to run it:
//   * as the reference does inside IQFrontEnd: `detectorPreprocessor` is the last block of the pre-processing chain
//     (signal_path/iq_frontend.cpp:37-44).  The GPU front end owns that detector on the device (b200_frontend_detector)
//     and feeds it the pre-processed block where it lies; this object is attach()ed to it, forwards the setters and
//     receives sigs_smoothed from the front end's worker after every block that completed a detection;
//   * stand-alone on dsp::stream buffers (init(&stream), start()): run() passes the block through unchanged and hands a
//     copy to its own device detector (b200_sigdet_process_host).
#pragma once
#include <cstring>
#include <mutex>
#include <vector>

#include "../processor.h"

namespace dsp::detector {

    class SignalDetector : public Processor<complex_t, complex_t> {
        using base_type = Processor<complex_t, complex_t>;
    public:
        std::vector<float> sigs_smoothed;       // Smoothed signal vector for detection (signal_detector.h:82)

        SignalDetector() {}
        ~SignalDetector() override {
            if (base_type::_block_init) { base_type::stop(); }
            if (own) { b200_sigdet_destroy(own); }
        }

        // the reference's front end calls init(NULL): the block then only lives inside the chain
        void init(stream<complex_t>* in) override { base_type::init(in); }

        void setSampleRate(double sr) {
            std::lock_guard<std::mutex> lk(mtx);
            sampleRate = sr;
            if (shared) { return; }              // the front end keeps its detector at the effective sample rate itself
            if (!own) {
                if (sr > 0 && b200_sigdet_create(&own, sr, 0) != B200_OK) { fprintf(stderr, "SignalDetector: %s\n", b200_last_error()); own = nullptr; }
                if (own) { b200_sigdet_set_enabled(own, enabled ? 1 : 0); }
            }
            else if (b200_sigdet_set_sample_rate(own, sr) != B200_OK) { fprintf(stderr, "SignalDetector: %s\n", b200_last_error()); }
        }

        // (the GUI thread calls this while the worker runs, display.cpp:151: a shared detector is touched under the
        // front end's lock, which its worker holds for the GPU call of a block)
        void setCenterFrequency(double f) {
            std::unique_lock<std::mutex> fl;
            if (feMtx) { fl = std::unique_lock<std::mutex>(*feMtx); }
            std::lock_guard<std::mutex> lk(mtx);
            if (det()) { b200_sigdet_set_center_frequency(det(), f); }
        }

        void setEnabled(bool e) {
            std::unique_lock<std::mutex> fl;
            if (feMtx) { fl = std::unique_lock<std::mutex>(*feMtx); }
            std::lock_guard<std::mutex> lk(mtx);
            enabled = e;
            if (det()) { b200_sigdet_set_enabled(det(), e ? 1 : 0); }
        }
        bool isEnabled() const { return enabled; }

        // stand-alone block: signal_detector.cpp:359-414
        int run() override {
            int count = base_type::_in->read();
            if (count < 0) { return -1; }
            memcpy(base_type::out.writeBuf, base_type::_in->readBuf, count * sizeof(complex_t));
            {
                std::lock_guard<std::mutex> lk(mtx);
                if (own && enabled) {
                    if (b200_sigdet_process_host(own, reinterpret_cast<const float*>(base_type::_in->readBuf), count, nullptr) != B200_OK) {
                        fprintf(stderr, "SignalDetector: %s\n", b200_last_error());
                    }
                    pull(own);
                }
            }
            base_type::_in->flush();
            if (!base_type::out.swap(count)) { return -1; }
            return count;
        }

        // ---- glue to IQFrontEnd (include/dsp/iq_frontend.h)
        void attach(b200_sigdet* frontEndDetector, std::mutex* frontEndMtx) {
            std::lock_guard<std::mutex> lk(mtx);
            shared = frontEndDetector; feMtx = frontEndMtx; seenRows = 0;
            enabled = b200_sigdet_is_enabled(shared) != 0;
        }
        // called by the front end's worker after a block (under the front end's lock): fetch sigs_smoothed when rows completed
        void refresh() { std::lock_guard<std::mutex> lk(mtx); if (shared) { pull(shared); } }
        int fftSize() { std::lock_guard<std::mutex> lk(mtx); return det() ? b200_sigdet_fft_size(det()) : 0; }
        long long rows() { std::lock_guard<std::mutex> lk(mtx); return det() ? (long long)b200_sigdet_rows(det()) : 0; }

    private:
        b200_sigdet* det() { return shared ? shared : own; }
        void pull(b200_sigdet* d) {
            const long long r = (long long)b200_sigdet_rows(d);
            // detections run with every 10th row (DETECT_INTERVAL_FRAMES): nothing new before that many rows have passed
            if (r == seenRows) { return; }
            seenRows = r;
            const int n = b200_sigdet_result(d, nullptr, 0);
            if (n <= 0) { return; }
            sigs_smoothed.resize((size_t)n);
            b200_sigdet_result(d, sigs_smoothed.data(), n);
        }

        std::mutex mtx;
        b200_sigdet* own = nullptr;
        b200_sigdet* shared = nullptr;
        std::mutex* feMtx = nullptr;
        double sampleRate = 0.0;
        bool enabled = true;
        long long seenRows = 0;
    };
}
