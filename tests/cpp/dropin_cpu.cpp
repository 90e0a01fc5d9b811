// CPU-only checks of the drop-in runtime headers (include/dsp/{stream,block,processor}.h): the stream
// rendezvous, stop semantics and the block lifecycle behave as the reference's (stream.h:88-186, block.h:28-109).
#include <atomic>
#include <cstdio>
#include <numeric>

#include <dsp/processor.h>

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
    printf("dropin_cpu ok\n");
    return 0;
}
