// dsp::block — thread-per-block lifecycle (reference: core/src/dsp/block.h:20-150):
// start()/stop() own a worker thread that loops `while (run() >= 0)`; tempStop()/tempStart() are
// depth-counted pauses used by setters; doStop() raises the stop flags of every registered stream, joins the
// worker and clears the flags again.
#pragma once
#include <algorithm>
#include <cassert>
#include <mutex>
#include <thread>
#include <vector>

#include "stream.h"
#include "types.h"

namespace dsp {
    class generic_block {
    public:
        virtual ~generic_block() {}
        virtual void start() {}
        virtual void stop() {}
        virtual int run() { return -1; }
    };

    class block : public generic_block {
    public:
        ~block() override {
            if (!_block_init) { return; }
            stop();
            _block_init = false;
        }

        void start() override {
            assert(_block_init);
            std::lock_guard<std::recursive_mutex> lk(ctrlMtx);
            if (running) { return; }
            running = true;
            doStart();
        }

        void stop() override {
            assert(_block_init);
            std::lock_guard<std::recursive_mutex> lk(ctrlMtx);
            if (!running) { return; }
            doStop();
            running = false;
        }

        void tempStop() {
            assert(_block_init);
            if (tempStopDepth++) { return; }
            if (running && !tempStopped) { doStop(); tempStopped = true; }
        }

        void tempStart() {
            assert(_block_init);
            if (!tempStopDepth || --tempStopDepth) { return; }
            if (tempStopped) { doStart(); tempStopped = false; }
        }

        int run() override = 0;

    protected:
        virtual void doStart() { worker = std::thread([this] { while (run() >= 0) {} }); }

        virtual void doStop() {
            for (auto* in : inputs) { if (in) { in->stopReader(); } }
            for (auto* out : outputs) { out->stopWriter(); }
            if (worker.joinable()) { worker.join(); }
            for (auto* in : inputs) { if (in) { in->clearReadStop(); } }
            for (auto* out : outputs) { out->clearWriteStop(); }
        }

        void registerInput(untyped_stream* s) {
            if (inputs.size() == 1 && s) { abort(); }    // one input per block, like the reference (block.h:119-125)
            inputs.push_back(s);
        }
        void unregisterInput(untyped_stream* s) { inputs.erase(std::remove(inputs.begin(), inputs.end(), s), inputs.end()); }
        void registerOutput(untyped_stream* s) { outputs.push_back(s); }
        void unregisterOutput(untyped_stream* s) { outputs.erase(std::remove(outputs.begin(), outputs.end(), s), outputs.end()); }

        bool _block_init = false;
        std::recursive_mutex ctrlMtx;
        std::vector<untyped_stream*> inputs, outputs;
        bool running = false, tempStopped = false;
        int tempStopDepth = 0;
        std::thread worker;
    };
}
