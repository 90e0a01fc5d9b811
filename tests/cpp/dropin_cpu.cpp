// CPU-only checks of the drop-in runtime headers (include/dsp/{stream,block,processor}.h): the stream
// rendezvous, stop semantics and the block lifecycle behave as the reference's (stream.h:88-186, block.h:28-109).
#include <atomic>
#include <cstdio>
#include <numeric>

#include <dsp/chain.h>

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

// a Processor that doubles its input
class Doubler : public dsp::Processor<float, float> {
public:
    int run() override {
        int n = _in->read();
        if (n < 0) { return -1; }
        for (int i = 0; i < n; i++) { out.writeBuf[i] = 2.0f * _in->readBuf[i]; }
        _in->flush();
        if (!out.swap(n)) { return -1; }
        return n;
    }
};

int main() {
    // 1. writer/reader rendezvous carries sizes and data in order
    {
        dsp::stream<float> s;
        std::atomic<long> sum{ 0 };
        std::thread reader([&] {
            for (;;) {
                int n = s.read();
                if (n < 0) { break; }
                for (int i = 0; i < n; i++) { sum += (long)s.readBuf[i]; }
                s.flush();
            }
        });
        long expect = 0;
        for (int b = 1; b <= 50; b++) {
            for (int i = 0; i < b * 10; i++) { s.writeBuf[i] = (float)(i % 7); expect += i % 7; }
            CHECK(s.swap(b * 10));
        }
        // wait until the last block was consumed, then stop the reader
        while (sum.load() != expect) { std::this_thread::yield(); }
        s.stopReader();
        reader.join();
        CHECK(sum.load() == expect);
        s.clearReadStop();
    }
    // 2. a stopped writer returns false from swap, a stopped reader -1 from read
    {
        dsp::stream<float> s;
        CHECK(s.swap(1));              // first swap never blocks
        s.stopWriter();
        CHECK(!s.swap(1));
        s.clearWriteStop();
        s.stopReader();
        CHECK(s.read() == -1);
        s.clearReadStop();
        CHECK(s.read() == 1);          // the block published above is still there
    }
    // 3. block lifecycle: start/stop, tempStop/tempStart nesting, data flows through a Processor
    {
        dsp::stream<float> in;
        Doubler d;
        d.init(&in);
        d.start();
        d.start();                     // idempotent
        for (int k = 0; k < 3; k++) {
            for (int i = 0; i < 100; i++) { in.writeBuf[i] = (float)i; }
            CHECK(in.swap(100));
            CHECK(d.out.read() == 100);
            CHECK(d.out.readBuf[99] == 198.0f);
            d.out.flush();
            if (k == 1) { d.tempStop(); d.tempStop(); d.tempStart(); d.tempStart(); }   // nested pause resumes the worker
        }
        d.stop();
        d.stop();
    }
    // 4. chain: blocks can be enabled / disabled while running; `out` always names the stream to read
    {
        dsp::stream<float> in;
        Doubler a, b;
        a.init(nullptr); b.init(nullptr);
        dsp::chain<float> ch(&in);
        dsp::stream<float>* tail = ch.out;
        auto onOut = [&](dsp::stream<float>* o) { tail = o; };
        CHECK(ch.out == &in);
        ch.addBlock(&a, false);
        ch.addBlock(&b, false);
        ch.start();
        auto pump = [&](float expectGain) {
            in.writeBuf[0] = 3.0f;
            if (!in.swap(1)) { return false; }
            if (tail->read() != 1) { return false; }
            const bool ok = tail->readBuf[0] == 3.0f * expectGain;
            tail->flush();
            return ok;
        };
        CHECK(pump(1.0f));                                  // nothing enabled: output is the input stream itself
        ch.enableBlock(&b, onOut);   CHECK(tail == &b.out); CHECK(pump(2.0f));
        ch.enableBlock(&a, onOut);   CHECK(tail == &b.out); CHECK(pump(4.0f));   // a is linked in front of b
        ch.disableBlock(&b, onOut);  CHECK(tail == &a.out); CHECK(pump(2.0f));
        ch.setBlockEnabled(&a, false, onOut); CHECK(tail == &in); CHECK(pump(1.0f));
        bool threw = false;
        try { ch.addBlock(&a, false); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);                                       // misuse raises std::runtime_error like the reference (chain.h:35)
        ch.stop();
    }
    printf("dropin_cpu ok\n");
    return 0;
}
