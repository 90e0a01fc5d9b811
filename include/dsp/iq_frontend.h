// IQFrontEnd on the GPU (reference: core/src/signal_path/iq_frontend.{h,cpp}).
// Same public surface: init(...) with the FFT acquire/release callbacks, the sample-rate / decimation /
// DC-block / invert / FFT setters, addVFO/removeVFO returning dsp::channel::RxVFO*, bind/unbindIQStream,
// start/stop.  One worker thread replaces the reference's SampleFrameBuffer + chain + Splitter + Reshaper +
// Handler + per-VFO threads (iq_frontend.cpp:223-261): it takes a block from the input stream, runs
// b200_frontend_process_host (pre-processing, every FFT frame of the block, every VFO), hands each dB row to
// acquireFFTBuffer/releaseFFTBuffer and swaps each VFO's `out` stream.
//
// Locks.  `mtx` guards the handle, the VFO map, the bound-stream list and the FFT geometry; the worker holds it for
// the GPU call of a block only.  `rowMtx` is held by the worker from that call until the block's dB rows have been
// handed to acquire/release; the FFT-geometry setters pass through it, so once one of them returns no row of the old
// size is delivered any more (the reference: tempStop of the reshaper and the FFT sink in updateFFTPath,
// iq_frontend.cpp:289-325).  `deliverMtx` is held by the worker for the whole block, until the last output stream
// has been swapped; removeVFO and unbindIQStream first stop the writers of every output stream -- what
// Splitter::unbindStream's tempStop does (routing/splitter.h:14-28) -- so a worker waiting on ANY consumer comes
// back, then pass through deliverMtx: when they return the worker holds no pointer to the removed VFO / stream.
// Lock order: deliverMtx, rowMtx, mtx.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../utils/event.h"
#include "buffer/frame_buffer.h"
#include "chain.h"
#include "channel/rx_vfo.h"
#include "detector/signal_detector.h"

class IQFrontEnd {
public:
    ~IQFrontEnd() {
        if (!_init) { return; }
        stop();
        for (auto& kv : vfos) { delete kv.second; }
        b200_frontend_destroy(fe);
    }

    enum FFTWindow { RECTANGULAR, BLACKMAN, NUTTALL };

    void init(dsp::stream<dsp::complex_t>* in, double sampleRate, bool buffering, int decimRatio, bool dcBlocking, int fftSize,
              double fftRate, FFTWindow fftWindow, float* (*acquireFFTBuffer)(void* ctx), void (*releaseFFTBuffer)(void* ctx), void* fftCtx) {
        _in = in; _buffering = buffering;
        // the reference always routes the source through its SampleFrameBuffer (a plain copy when buffering is off);
        // here the FIFO is only in the path while buffering is on
        inBuf.init(in);
        inBuf.bypass = false;
        preproc.init(head()); workIn = head(); _sampleRate = sampleRate; _decimRatio = decimRatio; _fftSize = fftSize; _fftRate = fftRate; _fftWindow = fftWindow;
        _acquire = acquireFFTBuffer; _release = releaseFFTBuffer; _ctx = fftCtx;
        if (b200_frontend_create(&fe, sampleRate, decimRatio, dcBlocking ? 1 : 0, 0, fftSize, fftRate, (int)fftWindow,
                                 STREAM_BUFFER_SIZE, 0) != B200_OK) { fail("b200_frontend_create"); }
        bank = b200_frontend_vfo_bank(fe);
        b200_vfo_bank_set_rotator_mode(bank, B200_ROTATOR_EXACT);
        // the detector is the last block of the pre-processing chain, added disabled (iq_frontend.cpp:37-44): it lives on
        // the device behind the GPU's decimator / DC blocker / conjugate, this object is its host face
        detectorPreprocessor.init(NULL);
        if (b200_sigdet* d = b200_frontend_detector(fe)) { detectorPreprocessor.attach(d, &mtx); }
        else { fail("b200_frontend_detector"); }
        resizeRows();
        _init = true;
    }

    void setInput(dsp::stream<dsp::complex_t>* in) {
        const bool was = running;
        if (was) { stop(); }
        _in = in;
        inBuf.setInput(in);
        preproc.setInput(head(), [this](dsp::stream<dsp::complex_t>* out) { workIn = out; });
        if (!userBlocks) { workIn = head(); }
        if (was) { start(); }
    }

    // Plug-in pre-processors (reference: iq_frontend.cpp:327-342, e.g. the logmmse noise reduction module).  They are
    // host blocks, so they run on their own threads in a dsp::chain in FRONT of the GPU stage; the built-in
    // decimator / DC blocker / conjugate run on the GPU after them (the reference orders plug-ins after its
    // built-in blocks -- the difference only matters for a plug-in that does not commute with those three).
    void addPreprocessor(dsp::Processor<dsp::complex_t, dsp::complex_t>* processor, bool enabled) {
        pauseFor([&] { preproc.addBlock(processor, false); userBlocks++; if (enabled) { preproc.enableBlock(processor, [this](dsp::stream<dsp::complex_t>* o) { workIn = o; }); } });
    }
    void removePreprocessor(dsp::Processor<dsp::complex_t, dsp::complex_t>* processor) {
        pauseFor([&] { preproc.removeBlock(processor, [this](dsp::stream<dsp::complex_t>* o) { workIn = o; }); userBlocks--; });
    }
    void togglePreprocessor(dsp::Processor<dsp::complex_t, dsp::complex_t>* processor, bool enabled) {
        if (processor == &detectorPreprocessor) {        // display.cpp:144: no host block to re-plumb, the device detector is switched
            { std::lock_guard<std::mutex> lk(mtx); if (b200_frontend_set_detector(fe, enabled ? 1 : 0) != B200_OK) { fail("togglePreprocessor(detector)"); } }
            detectorPreprocessor.setEnabled(enabled);
            return;
        }
        pauseFor([&] { preproc.setBlockEnabled(processor, enabled, [this](dsp::stream<dsp::complex_t>* o) { workIn = o; }); });
    }
    void setSampleRate(double sampleRate) {
        {
            std::lock_guard<std::mutex> rl(rowMtx);
            std::lock_guard<std::mutex> lk(mtx);
            _sampleRate = sampleRate;
            if (b200_frontend_set_sample_rate(fe, sampleRate) != B200_OK) { fail("setSampleRate"); }
            resizeRows();
        }
        onEffectiveSampleRateChange.emit(_sampleRate / _decimRatio);     // outside the lock: handlers may call back in
    }
    double getSampleRate() { return _sampleRate / _decimRatio; }
    // IQFrontEnd::setBuffering (iq_frontend.cpp:118-120): switch the input FIFO in or out
    void setBuffering(bool enabled) {
        if (enabled == _buffering) { return; }
        pauseFor([&] {
            _buffering = enabled;
            preproc.setInput(head(), [this](dsp::stream<dsp::complex_t>* out) { workIn = out; });
            if (!userBlocks) { workIn = head(); }
        });
    }
    void setDecimation(int ratio) {
        {
            std::lock_guard<std::mutex> rl(rowMtx);
            std::lock_guard<std::mutex> lk(mtx);
            _decimRatio = ratio;
            if (b200_frontend_set_decimation(fe, ratio) != B200_OK) { fail("setDecimation"); }
            resizeRows();
        }
        onEffectiveSampleRateChange.emit(_sampleRate / _decimRatio);
    }
    void setInvertIQ(bool enabled) { std::lock_guard<std::mutex> lk(mtx); b200_frontend_set_invert_iq(fe, enabled); }
    void setDCBlocking(bool enabled) { std::lock_guard<std::mutex> lk(mtx); b200_frontend_set_dc_blocking(fe, enabled); }

    // extra consumers of the pre-processed IQ (recorder, exporters): reference iq_frontend.cpp:151-157
    void bindIQStream(dsp::stream<dsp::complex_t>* s) { std::lock_guard<std::mutex> lk(mtx); bound.push_back(s); }
    void unbindIQStream(dsp::stream<dsp::complex_t>* s) {
        detach([&] { bound.erase(std::remove(bound.begin(), bound.end(), s), bound.end()); });
    }

    dsp::channel::RxVFO* addVFO(std::string name, double sampleRate, double bandwidth, double offset) {
        std::lock_guard<std::mutex> lk(mtx);
        if (vfos.find(name) != vfos.end()) { fprintf(stderr, "[IQFrontEnd] Tried to add VFO with existing name.\n"); return nullptr; }
        if ((int)vfos.size() >= kMaxVfos) { fprintf(stderr, "[IQFrontEnd] addVFO: the bank is full (%d VFOs).\n", kMaxVfos); return nullptr; }
        const int idx = b200_vfo_bank_add(bank, sampleRate, bandwidth, offset);
        if (idx < 0) { fprintf(stderr, "[IQFrontEnd] addVFO failed: %s\n", b200_last_error()); return nullptr; }
        auto* v = new dsp::channel::RxVFO();
        v->attach(bank, idx, &mtx);
        vfos[name] = v;
        return v;
    }
    // IQFrontEnd::removeVFO (iq_frontend.cpp:181-202).  The reference stops the VFO's own thread, unbinds its input
    // from the Splitter and deletes it while everything else keeps running; the same holds here whether the VFO's
    // reader is still consuming or has already been stopped by the module (the usual order in a module's destructor).
    void removeVFO(std::string name) {
        dsp::channel::RxVFO* v = nullptr;
        detach([&] {
            auto it = vfos.find(name);
            if (it == vfos.end()) { fprintf(stderr, "[IQFrontEnd] Tried to remove a VFO that doesn't exist.\n"); return; }
            v = it->second;
            vfos.erase(it);
            b200_vfo_bank_remove(bank, v->index());
        });
        delete v;
    }

    void setFFTSize(int size) { std::lock_guard<std::mutex> rl(rowMtx); std::lock_guard<std::mutex> lk(mtx); _fftSize = size; if (b200_frontend_set_fft_size(fe, size) != B200_OK) { fail("setFFTSize"); } resizeRows(); }
    void setFFTRate(double rate) { std::lock_guard<std::mutex> rl(rowMtx); std::lock_guard<std::mutex> lk(mtx); _fftRate = rate; if (b200_frontend_set_fft_rate(fe, rate) != B200_OK) { fail("setFFTRate"); } resizeRows(); }
    double getFFTRate() { return _fftRate; }
    void setFFTWindow(FFTWindow w) { std::lock_guard<std::mutex> lk(mtx); _fftWindow = w; if (b200_frontend_set_fft_window(fe, (int)w) != B200_OK) { fail("setFFTWindow"); } }
    void flushInputBuffer() { inBuf.flush(); }

    void start() {
        if (running || !_in) { return; }
        running = true;
        workIn->clearReadStop();
        if (_buffering) { inBuf.start(); }
        preproc.start();
        worker = std::thread([this] { while (step() >= 0) {} });
    }
    void stop() {
        if (!running) { return; }
        workIn->stopReader();
        for (auto& kv : vfos) { kv.second->out.stopWriter(); }
        for (auto* s : bound) { s->stopWriter(); }
        if (worker.joinable()) { worker.join(); }
        preproc.stop();
        if (_buffering) { inBuf.stop(); }
        workIn->clearReadStop();
        for (auto& kv : vfos) { kv.second->out.clearWriteStop(); }
        for (auto* s : bound) { s->clearWriteStop(); }
        running = false;
    }

    double getEffectiveSamplerate() { return b200_frontend_effective_samplerate(fe); }
    Event<double> onEffectiveSampleRateChange;     // emitted by setSampleRate / setDecimation (iq_frontend.cpp:99,129)
    dsp::detector::SignalDetector detectorPreprocessor;   // signal_path/iq_frontend.h:77
    b200_frontend* handle() { return fe; }

protected:
    // one block: the work of SampleFrameBuffer::run .. Splitter::run .. RxVFO::run / IQFrontEnd::handler
    int step() {
        dsp::stream<dsp::complex_t>* in = workIn;
        const int count = in->read();
        if (count < 0) { return -1; }
        std::lock_guard<std::mutex> dl(deliverMtx);        // held until this block's streams are handed out
        std::unique_lock<std::mutex> rl(rowMtx);           // held until this block's rows are handed out
        std::vector<dsp::channel::RxVFO*> live;
        std::vector<dsp::stream<dsp::complex_t>*> boundNow;
        int nFrames = 0, preCount = 0, rowLen = 0;
        {
            std::lock_guard<std::mutex> lk(mtx);
            int top = -1;
            for (auto& kv : vfos) { top = std::max(top, kv.second->index()); }
            outs.assign((size_t)top + 1, nullptr);
            counts.assign((size_t)top + 1, 0);
            for (auto& kv : vfos) { outs[kv.second->index()] = reinterpret_cast<float*>(kv.second->out.writeBuf); }
            if (rows.size() != (size_t)maxRows * _fftSize) { rows.assign((size_t)maxRows * _fftSize, 0.0f); }
            if (b200_frontend_process_host(fe, reinterpret_cast<const float*>(in->readBuf), count, rows.data(), maxRows, outs.data(), nullptr,
                                           counts.data(), nullptr) != B200_OK) { fail("process"); }
            nFrames = std::min(b200_frontend_frame_count(fe), maxRows);
            detectorPreprocessor.refresh();                // sigs_smoothed, when the block completed detector rows
            rowLen = _fftSize;                             // the geometry these rows were made with (its setters wait on rowMtx)
            for (auto& kv : vfos) { kv.second->pending = counts[kv.second->index()]; live.push_back(kv.second); }
            boundNow = bound;
            if (!boundNow.empty()) {
                const float* dPre = nullptr;
                b200_frontend_preproc_out(fe, &dPre, &preCount);
                for (auto* s : boundNow) { b200_device_to_host(s->writeBuf, dPre, (size_t)preCount * sizeof(dsp::complex_t), nullptr); }
            }
        }
        in->flush();
        for (int f = 0; f < nFrames; f++) {
            float* dst = _acquire ? _acquire(_ctx) : nullptr;
            if (dst) { memcpy(dst, rows.data() + (size_t)f * rowLen, (size_t)rowLen * sizeof(float)); }
            if (_release) { _release(_ctx); }    // always called, even when acquire returned NULL (iq_frontend.cpp:281-286)
        }
        rl.unlock();
        // A swap that comes back false belongs to a stream whose writer was stopped: a VFO being removed, a stream being
        // unbound, or stop() -- in every case this consumer is skipped and the others are still served.  stop() also
        // stops the input's reader, so the next read() ends the worker; nothing else does.
        for (auto* v : live) { if (v->pending > 0) { (void)v->out.swap(v->pending); } }
        for (auto* s : boundNow) { if (preCount > 0) { (void)s->swap(preCount); } }
        return count;
    }

    // Take a consumer out while the worker runs: stop the writers of every output stream (a worker waiting for any
    // consumer returns from swap() and skips it; that block is lost for the skipped consumers, as it is behind the
    // reference's Splitter), wait until the worker is between blocks, change the lists, release the writers again.
    template <typename F>
    void detach(F change) {
        std::vector<dsp::stream<dsp::complex_t>*> stopped;
        {
            std::lock_guard<std::mutex> lk(mtx);
            for (auto& kv : vfos) { stopped.push_back(&kv.second->out); }
            for (auto* s : bound) { stopped.push_back(s); }
        }
        for (auto* s : stopped) { s->stopWriter(); }
        {
            std::lock_guard<std::mutex> dl(deliverMtx);
            std::lock_guard<std::mutex> lk(mtx);
            change();
        }
        // (a removed VFO is deleted by the caller right after this; clearing the flag of its stream first is harmless)
        if (running) { for (auto* s : stopped) { s->clearWriteStop(); } }
    }

    template <typename F>
    void pauseFor(F change) {
        const bool was = running;
        if (was) { stop(); }
        change();
        if (was) { start(); }
    }

    void resizeRows() {
        const double eff = _sampleRate / _decimRatio;
        const int interval = std::max(1, (int)(eff / _fftRate + 0.5));
        maxRows = STREAM_BUFFER_SIZE / interval + 2;     // `rows` itself is resized by the worker before its next block
    }

    [[noreturn]] static void fail(const char* what) {
        fprintf(stderr, "IQFrontEnd: %s failed: %s\n", what, b200_last_error());
        abort();
    }

    static constexpr int kMaxVfos = 4096;
    b200_frontend* fe = nullptr;
    b200_vfo_bank* bank = nullptr;
    dsp::stream<dsp::complex_t>* _in = nullptr;
    dsp::stream<dsp::complex_t>* workIn = nullptr;      // what the worker reads: _in, or the tail of the plug-in chain
    dsp::chain<dsp::complex_t> preproc;
    dsp::buffer::SampleFrameBuffer<dsp::complex_t> inBuf;
    bool _buffering = false;
    dsp::stream<dsp::complex_t>* head() { return _buffering ? &inBuf.out : _in; }   // what the plug-in chain reads
    int userBlocks = 0;
    std::map<std::string, dsp::channel::RxVFO*> vfos;
    std::vector<dsp::stream<dsp::complex_t>*> bound;
    std::vector<float> rows;                            // worker only
    std::vector<float*> outs;                           // worker only: per-VFO-index host pointers / counts of a block
    std::vector<int> counts;
    int maxRows = 0;
    double _sampleRate = 0, _fftRate = 0;
    int _decimRatio = 1, _fftSize = 0;
    FFTWindow _fftWindow = NUTTALL;
    float* (*_acquire)(void*) = nullptr;
    void (*_release)(void*) = nullptr;
    void* _ctx = nullptr;
    std::mutex mtx;
    std::mutex rowMtx;
    std::mutex deliverMtx;
    std::thread worker;
    bool running = false, _init = false;
};
