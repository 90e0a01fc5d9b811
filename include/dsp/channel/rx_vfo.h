// dsp::channel::RxVFO on the GPU (reference: core/src/dsp/channel/rx_vfo.h:6-135).
// Same constructor, setters, process() and run(); the arithmetic runs in libb200dsp on a one-VFO bank.
// The rotator runs in B200_ROTATOR_EXACT mode: the phase recursion of the reference's FrequencyXlator is
// replayed bit for bit, so a stream processed here tracks the CPU block for any offset and length.
// (Many VFOs on one input should share a bank -- see IQFrontEnd::addVFO in ../iq_frontend.h -- so that the
// input block crosses PCIe once; a stand-alone RxVFO pays one H2D copy per VFO.)
#pragma once
#include <cstdio>
#include <vector>

#include "../processor.h"

namespace dsp::channel {
    class RxVFO : public Processor<complex_t, complex_t> {
        using base_type = Processor<complex_t, complex_t>;
    public:
        RxVFO() {}
        RxVFO(stream<complex_t>* in, double inSamplerate, double outSamplerate, double bandwidth, double offset) {
            init(in, inSamplerate, outSamplerate, bandwidth, offset);
        }
        ~RxVFO() override {
            if (!base_type::_block_init) { return; }
            base_type::stop();
            if (ownsBank && bank) { b200_vfo_bank_destroy(bank); }
        }

        void init(stream<complex_t>* in, double inSamplerate, double outSamplerate, double bandwidth, double offset) {
            if (b200_vfo_bank_create(&bank, inSamplerate, STREAM_BUFFER_SIZE, 0) != B200_OK) { fail("b200_vfo_bank_create"); }
            ownsBank = true;
            b200_vfo_bank_set_rotator_mode(bank, B200_ROTATOR_EXACT);
            vfo = b200_vfo_bank_add(bank, outSamplerate, bandwidth, offset);
            if (vfo < 0) { fail("b200_vfo_bank_add"); }
            base_type::init(in);
        }

        // attach to a VFO that lives in a front end's shared bank (used by IQFrontEnd::addVFO): the front end's
        // worker fills `out`, so this object runs no thread of its own and shares the front end's lock
        void attach(b200_vfo_bank* sharedBank, int index, std::mutex* frontEndMtx) {
            bank = sharedBank; vfo = index; ownsBank = false; attached = true; sharedMtx = frontEndMtx;
            base_type::init(nullptr);
        }
        void start() override { if (!attached) { base_type::start(); } }
        void stop() override { if (!attached) { base_type::stop(); } }

        void setInSamplerate(double inSamplerate) {
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            base_type::tempStop();
            { std::lock_guard<std::mutex> lk2(lock()); if (b200_vfo_bank_set_in_samplerate(bank, inSamplerate) != B200_OK) { fail("setInSamplerate"); } }
            base_type::tempStart();
        }
        void setOutSamplerate(double outSamplerate, double bandwidth) {
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            base_type::tempStop();
            { std::lock_guard<std::mutex> lk2(lock()); if (b200_vfo_set_out_samplerate(bank, vfo, outSamplerate, bandwidth) != B200_OK) { fail("setOutSamplerate"); } }
            base_type::tempStart();
        }
        void setBandwidth(double bandwidth) {
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            std::lock_guard<std::mutex> lk2(lock());
            if (b200_vfo_set_bandwidth(bank, vfo, bandwidth) != B200_OK) { fail("setBandwidth"); }
        }
        void setOffset(double offset) {
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            std::lock_guard<std::mutex> lk2(lock());
            if (b200_vfo_set_offset(bank, vfo, offset) != B200_OK) { fail("setOffset"); }
        }
        void reset() {
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            base_type::tempStop();
            { std::lock_guard<std::mutex> lk2(lock()); b200_vfo_reset(bank, vfo); }
            base_type::tempStart();
        }

        // synchronous: count input samples -> returns the number of output samples written to `out`
        int process(int count, const complex_t* in, complex_t* out) {
            std::lock_guard<std::mutex> lk(lock());
            hostOuts.assign((size_t)vfo + 1, nullptr);     // tables indexed by the VFO's slot in its bank, whatever that is
            hostCounts.assign((size_t)vfo + 1, 0);
            hostOuts[vfo] = reinterpret_cast<float*>(out);
            if (b200_vfo_bank_process_host(bank, reinterpret_cast<const float*>(in), count, hostOuts.data(), hostCounts.data(), nullptr) != B200_OK) { fail("process"); }
            return hostCounts[vfo];
        }

        int run() override {
            int count = base_type::_in->read();
            if (count < 0) { return -1; }
            int outCount = process(count, base_type::_in->readBuf, base_type::out.writeBuf);
            base_type::_in->flush();
            if (outCount) {
                if (!base_type::out.swap(outCount)) { return -1; }
            }
            return outCount;
        }

        b200_vfo_bank* bankHandle() const { return bank; }
        int index() const { return vfo; }
        int pending = 0;   // output samples of the front end's last block, waiting to be swapped out

    protected:
        [[noreturn]] static void fail(const char* what) {
            fprintf(stderr, "RxVFO: %s failed: %s\n", what, b200_last_error());
            abort();
        }
        b200_vfo_bank* bank = nullptr;
        int vfo = -1;
        bool ownsBank = false, attached = false;
        std::mutex procMtx;
        std::vector<float*> hostOuts;
        std::vector<int> hostCounts;
        std::mutex* sharedMtx = nullptr;
        std::mutex& lock() { return sharedMtx ? *sharedMtx : procMtx; }
    };
}
