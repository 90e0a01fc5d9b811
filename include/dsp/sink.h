// dsp::Sink<T> — a block with one input stream and no output (reference: core/src/dsp/sink.h:5-41).
#pragma once
#include <cstdlib>

#include "block.h"

namespace dsp {
    template <class T>
    class Sink : public block {
    public:
        Sink() {}
        explicit Sink(stream<T>* in) { init(in); }
        ~Sink() override {}

        virtual void init(stream<T>* in) {
            if (_block_init) { abort(); }          // double init is a programming error in the reference too
            _in = in;
            registerInput(_in);
            _block_init = true;
        }

        virtual void setInput(stream<T>* in) {
            assert(_block_init);
            std::lock_guard<std::recursive_mutex> lk(ctrlMtx);
            tempStop();
            unregisterInput(_in);
            _in = in;
            if (_in) { registerInput(_in); }
            tempStart();
        }

        int run() override = 0;

    protected:
        stream<T>* _in = nullptr;
    };
}
