// dsp::stream<T> — the double-buffer rendezvous between a writer thread and a reader thread
// (reference: core/src/dsp/stream.h:35-218).  Same public contract:
//   writer: fill writeBuf, swap(n) -> false when stopped        (stream.h:88-113)
//   reader: n = read() (-1 when stopped), use readBuf, flush()  (stream.h:115-153)
//   stopWriter/clearWriteStop/stopReader/clearReadStop          (stream.h:155-186)
// Difference: the two buffers are page-locked host memory from libb200dsp (b200_host_alloc), so blocks handed to
// the GPU front end are copied at full PCIe rate without an intermediate staging copy.
#pragma once
#include <condition_variable>
#include <cstdlib>
#include <mutex>

#include "../b200dsp.h"

#define STREAM_BUFFER_SIZE 1000000   // samples per buffer (reference: stream.h:14)

namespace dsp {
    class untyped_stream {
    public:
        virtual ~untyped_stream() {}
        virtual bool swap(int) { return false; }
        virtual int read() { return -1; }
        virtual void flush() {}
        virtual void stopWriter() {}
        virtual void clearWriteStop() {}
        virtual void stopReader() {}
        virtual void clearReadStop() {}
    };

    template <class T>
    class stream : public untyped_stream {
    public:
        stream() { allocate(STREAM_BUFFER_SIZE); }
        explicit stream(const char* name) : origin(name) { allocate(STREAM_BUFFER_SIZE); }
        stream(const stream&) = delete;
        stream& operator=(const stream&) = delete;
        ~stream() override { release(); }

        void setBufferSize(int samples) {
            release();
            allocate(samples);
        }

        // writer side: publish `size` samples of writeBuf; blocks until the reader has flushed the previous block
        bool swap(int size) override {
            {
                std::unique_lock<std::mutex> lk(swapMtx);
                swapCv.wait(lk, [this] { return canSwap || writerStop; });
                if (writerStop) { return false; }
                dataSize = size;
                T* t = writeBuf; writeBuf = readBuf; readBuf = t;
                canSwap = false;
            }
            {
                std::lock_guard<std::mutex> lk(readyMtx);
                dataReady = true;
            }
            readyCv.notify_all();
            return true;
        }

        // reader side: wait for a block; -1 once stopReader() was called
        int read() override {
            if (!readBuf) { return -1; }
            std::unique_lock<std::mutex> lk(readyMtx);
            readyCv.wait(lk, [this] { return dataReady || readerStop; });
            return readerStop ? -1 : dataSize;
        }

        void flush() override {
            {
                std::lock_guard<std::mutex> lk(readyMtx);
                dataReady = false;
            }
            {
                std::lock_guard<std::mutex> lk(swapMtx);
                canSwap = true;
            }
            swapCv.notify_all();
        }

        void stopWriter() override {
            { std::lock_guard<std::mutex> lk(swapMtx); writerStop = true; }
            swapCv.notify_all();
        }
        void clearWriteStop() override { std::lock_guard<std::mutex> lk(swapMtx); writerStop = false; }
        void stopReader() override {
            { std::lock_guard<std::mutex> lk(readyMtx); readerStop = true; }
            readyCv.notify_all();
        }
        void clearReadStop() override { std::lock_guard<std::mutex> lk(readyMtx); readerStop = false; }

        void free() { release(); }

        T* writeBuf = nullptr;
        T* readBuf = nullptr;
        const char* origin = "stream";

    private:
        void allocate(int samples) {
            capacity = samples;
            bufA = static_cast<T*>(b200_host_alloc((size_t)samples * sizeof(T)));
            bufB = static_cast<T*>(b200_host_alloc((size_t)samples * sizeof(T)));
            if (!bufA || !bufB) { abort(); }     // the reference aborts on allocation failure too (stream.h:58-63)
            writeBuf = bufA; readBuf = bufB;
        }
        void release() {
            if (bufA) { b200_host_free(bufA); bufA = nullptr; }
            if (bufB) { b200_host_free(bufB); bufB = nullptr; }
            writeBuf = readBuf = nullptr;
        }

        T* bufA = nullptr;
        T* bufB = nullptr;
        int capacity = 0;
        std::mutex swapMtx, readyMtx;
        std::condition_variable swapCv, readyCv;
        bool canSwap = true, dataReady = false, readerStop = false, writerStop = false;
        int dataSize = 0;
    };
}
