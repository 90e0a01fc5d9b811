// dsp::Processor<I, O> — one input stream, one owned output stream (reference: core/src/dsp/processor.h:40-109).
#pragma once
#include "block.h"

namespace dsp {
    template <class I, class O>
    class Processor : public block {
    public:
        Processor() {}
        explicit Processor(stream<I>* in) { init(in); }
        ~Processor() override {}

        virtual void init(stream<I>* in) {
            _in = in;
            registerInput(_in);
            registerOutput(&out);
            _block_init = true;
        }

        virtual void setInput(stream<I>* in) {
            assert(_block_init);
            std::lock_guard<std::recursive_mutex> lk(ctrlMtx);
            tempStop();
            unregisterInput(_in);
            _in = in;
            if (_in) { registerInput(_in); }
            tempStart();
        }

        int run() override = 0;

        stream<O> out{ "processor.out" };

    protected:
        stream<I>* _in = nullptr;
    };
}
