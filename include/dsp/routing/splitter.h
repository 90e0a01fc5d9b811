// dsp::routing::Splitter<T> — copies every input block to each bound stream (reference:
// core/src/dsp/routing/splitter.h:6-97).  Host-side fan-out for consumers that still want a dsp::stream; the GPU
// path itself has no fan-out copy (every device consumer reads the same buffer).
#pragma once
#include <algorithm>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <vector>

#include "../sink.h"

namespace dsp::routing {
    template <class T>
    class Splitter : public Sink<T> {
        using base_type = Sink<T>;

    public:
        Splitter() {}
        explicit Splitter(stream<T>* in) { base_type::init(in); }

        void bindStream(stream<T>* s) {
            assert(base_type::_block_init);
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            if (isBound(s)) { throw std::runtime_error("[Splitter] Tried to bind stream to that is already bound"); }
            base_type::tempStop();
            base_type::registerOutput(s);
            streams.push_back(s);
            base_type::tempStart();
        }

        void unbindStream(stream<T>* s) {
            assert(base_type::_block_init);
            std::lock_guard<std::recursive_mutex> lk(base_type::ctrlMtx);
            if (!isBound(s)) { throw std::runtime_error("[Splitter] Tried to unbind stream to that isn't bound"); }
            base_type::tempStop();
            streams.erase(std::find(streams.begin(), streams.end(), s));
            base_type::unregisterOutput(s);
            base_type::tempStart();
        }

        // called with every block before it is handed on (the reference's tap for recorders / meters)
        void setHook(const std::function<void(T*, int)>& h) { hook = h; }

        int run() override {
            const int count = base_type::_in->read();
            if (count < 0) { return -1; }
            if (hook) { hook(base_type::_in->readBuf, count); }
            for (auto* s : streams) {
                memcpy(s->writeBuf, base_type::_in->readBuf, (size_t)count * sizeof(T));
                if (!s->swap(count)) { return -1; }
            }
            base_type::_in->flush();
            return count;
        }

    private:
        bool isBound(stream<T>* s) const { return std::find(streams.begin(), streams.end(), s) != streams.end(); }

        std::vector<stream<T>*> streams;
        std::function<void(T*, int)> hook;
    };
}
