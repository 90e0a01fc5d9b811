// dsp::buffer::SampleFrameBuffer<T> — decouples a bursty source from the DSP graph (reference:
// core/src/dsp/buffer/frame_buffer.h:8-131; IQFrontEnd::setBuffering, signal_path/iq_frontend.cpp:98-105).
// bypass = true: every block is copied straight from the input stream to `out`.  bypass = false: the input side
// queues up to kFrames blocks and a second thread feeds `out`, so the source never waits on the graph for longer
// than one copy.  Same members as the reference (init / setInput / flush / run / out / bypass); the queue is a
// deque of vectors instead of 32 preallocated million-sample buffers, and when it is full the OLDEST block is
// dropped, which is what the reference's wrap-around of its write cursor amounts to.
#pragma once
#include <condition_variable>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include "../block.h"

namespace dsp::buffer {
    template <class T>
    class SampleFrameBuffer : public block {
    public:
        static constexpr int kFrames = 32;

        SampleFrameBuffer() {}
        explicit SampleFrameBuffer(stream<T>* in) { init(in); }
        ~SampleFrameBuffer() override {
            if (!_block_init) { return; }
            block::stop();
            _block_init = false;
        }

        void init(stream<T>* in) {
            _in = in;
            registerInput(_in);
            registerOutput(&out);
            _block_init = true;
        }

        void setInput(stream<T>* in) {
            assert(_block_init);
            std::lock_guard<std::recursive_mutex> lk(ctrlMtx);
            tempStop();
            unregisterInput(_in);
            _in = in;
            registerInput(_in);
            tempStart();
        }

        // forget everything queued (IQFrontEnd::flushInputBuffer)
        void flush() {
            std::lock_guard<std::mutex> lk(qMtx);
            queue.clear();
        }

        int run() override {
            const int count = _in->read();
            if (count < 0) { return -1; }
            if (bypass) {
                memcpy(out.writeBuf, _in->readBuf, (size_t)count * sizeof(T));
                _in->flush();
                return out.swap(count) ? count : -1;
            }
            {
                std::lock_guard<std::mutex> lk(qMtx);
                if ((int)queue.size() >= kFrames - 1) { queue.pop_front(); }
                queue.emplace_back(_in->readBuf, _in->readBuf + count);
            }
            qCv.notify_all();
            _in->flush();
            return count;
        }

        stream<T> out{ "frame_buffer.out" };
        bool bypass = false;

    protected:
        void doStart() override {
            block::doStart();
            feeder = std::thread([this] { feed(); });
        }

        void doStop() override {
            { std::lock_guard<std::mutex> lk(qMtx); stopFeeder = true; }
            qCv.notify_all();
            block::doStop();                       // raises the stream stop flags, joins the input-side worker
            if (feeder.joinable()) { feeder.join(); }
            stopFeeder = false;
        }

    private:
        void feed() {
            while (true) {
                std::vector<T> blk;
                {
                    std::unique_lock<std::mutex> lk(qMtx);
                    qCv.wait(lk, [this] { return !queue.empty() || stopFeeder; });
                    if (stopFeeder) { return; }
                    blk.swap(queue.front());
                    queue.pop_front();
                }
                memcpy(out.writeBuf, blk.data(), blk.size() * sizeof(T));
                if (!out.swap((int)blk.size())) { return; }
            }
        }

        stream<T>* _in = nullptr;
        std::deque<std::vector<T>> queue;
        std::mutex qMtx;
        std::condition_variable qCv;
        std::thread feeder;
        bool stopFeeder = false;
    };
}
