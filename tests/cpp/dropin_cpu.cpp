// CPU-only checks of the drop-in runtime headers (include/dsp/{stream,block,processor}.h): the stream
// rendezvous, stop semantics and the block lifecycle behave as the reference's (stream.h:88-186, block.h:28-109).
#include <atomic>
#include <cstdio>
#include <numeric>

#include <dsp/buffer/frame_buffer.h>
#include <dsp/chain.h>
#include <dsp/routing/splitter.h>
#include <utils/event.h>
#include <utils/wav.h>
#include <stdexcept>
#include <vector>

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
    {   // Splitter: every bound stream receives every block; the hook sees it first; double bind throws
        dsp::stream<float> in, o1, o2;
        dsp::routing::Splitter<float> sp(&in);
        int hooked = 0;
        sp.setHook([&](float* p, int n) { hooked += n; });
        sp.bindStream(&o1);
        sp.bindStream(&o2);
        bool threw = false;
        try { sp.bindStream(&o1); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        sp.start();
        for (int b = 0; b < 3; b++) {
            for (int i = 0; i < 50; i++) { in.writeBuf[i] = (float)(b * 100 + i); }
            CHECK(in.swap(50));
            CHECK(o1.read() == 50); CHECK(o1.readBuf[49] == (float)(b * 100 + 49)); o1.flush();
            CHECK(o2.read() == 50); CHECK(o2.readBuf[0] == (float)(b * 100)); o2.flush();
        }
        sp.stop();
        CHECK(hooked == 150);
        sp.unbindStream(&o2);
        threw = false;
        try { sp.unbindStream(&o2); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
    }
    {   // SampleFrameBuffer: bypass copies block for block; buffered mode lets the writer run ahead of the reader
        dsp::stream<float> in;
        dsp::buffer::SampleFrameBuffer<float> fb(&in);
        fb.bypass = true;
        fb.start();
        in.writeBuf[0] = 7.0f; CHECK(in.swap(1));
        CHECK(fb.out.read() == 1); CHECK(fb.out.readBuf[0] == 7.0f); fb.out.flush();
        fb.stop();
        fb.bypass = false;
        fb.start();
        for (int b = 0; b < 8; b++) {                          // nobody reads `out` yet: the source is never blocked for long
            in.writeBuf[0] = (float)b;
            CHECK(in.swap(1));
        }
        for (int b = 0; b < 8; b++) {
            CHECK(fb.out.read() == 1);
            CHECK(fb.out.readBuf[0] == (float)b);               // in order, nothing lost below the FIFO depth
            fb.out.flush();
        }
        fb.flush();
        fb.stop();
    }
    {   // Event: handlers run in bind order with their context; a handler may unbind itself while the event is emitted
        Event<double> ev;
        static double seen[2];
        static Event<double>* evp;
        static EventHandler<double> h1, h2;
        seen[0] = seen[1] = 0.0;
        evp = &ev;
        h1 = EventHandler<double>([](double v, void* c) { ((double*)c)[0] += v; evp->unbindHandler(&h1); }, seen);
        h2 = EventHandler<double>([](double v, void* c) { ((double*)c)[1] += v; }, seen);
        ev.bindHandler(&h1);
        ev.bindHandler(&h2);
        ev.emit(2.5);
        ev.emit(1.0);
        CHECK(seen[0] == 2.5);
        CHECK(seen[1] == 3.5);
    }
    {   // wav::Reader: header fields, looping readSamples, plain readSamples2, rejection of a non-RIFF file
        const char* path = "/tmp/b200_dropin_test.wav";
        {
            FILE* f = fopen(path, "wb");
            CHECK(f != nullptr);
            const uint32_t sr = 2400000, nbytes = 6 * 4;
            const uint16_t pcm = 1, ch = 2, bps = 4, depth = 16;
            const uint32_t fileSize = 36 + nbytes, fmtLen = 16, bytesPerSec = sr * 4;
            fwrite("RIFF", 1, 4, f); fwrite(&fileSize, 4, 1, f); fwrite("WAVE", 1, 4, f); fwrite("fmt ", 1, 4, f); fwrite(&fmtLen, 4, 1, f);
            fwrite(&pcm, 2, 1, f); fwrite(&ch, 2, 1, f); fwrite(&sr, 4, 1, f); fwrite(&bytesPerSec, 4, 1, f); fwrite(&bps, 2, 1, f); fwrite(&depth, 2, 1, f);
            fwrite("data", 1, 4, f); fwrite(&nbytes, 4, 1, f);
            const int16_t iq[12] = { 1, -1, 2, -2, 3, -3, 4, -4, 5, -5, 6, -6 };
            fwrite(iq, 2, 12, f);
            fclose(f);
        }
        wav::Reader r(path);
        CHECK(r.isValid());
        CHECK(r.getSampleRate() == 2400000 && r.getChannelCount() == 2 && r.getBitDepth() == 16);
        CHECK(r.getSampleType() == wav::SAMP_TYPE_INT16);
        int16_t buf[16];
        r.readSamples(buf, 16 * sizeof(int16_t));             // 12 values in the file: wraps to the first sample
        CHECK(buf[0] == 1 && buf[11] == -6 && buf[12] == 1 && buf[15] == -2);
        r.rewind();
        CHECK(r.readSamples2(buf, sizeof(buf)) == 12 * sizeof(int16_t));
        r.close();
        FILE* g = fopen(path, "wb"); fwrite("not a wav file, only some text that is long enough", 1, 50, g); fclose(g);
        wav::Reader bad(path);
        CHECK(!bad.isValid() && bad.error == "signature mismatch");
        wav::Reader missing("/tmp/b200_no_such_file.wav");
        CHECK(!missing.isValid() && missing.error == "cannot open file");
        remove(path);
    }
    {   // wav::Writer: container layout (checked byte for byte against the reference's writer by the Python side), the
        // setters' refusal while a file is open, and a round trip through wav::Reader
        const char* pf = "/tmp/b200_dropin_writer_f32.wav";
        const char* pi = "/tmp/b200_dropin_writer_i16.wav";
        std::vector<float> x(2 * 1000);
        for (int i = 0; i < 2000; i++) { x[i] = (float)i / 2000.0f - 0.5f; }
        wav::Writer w(2, 48000, wav::FORMAT_WAV, wav::SAMP_TYPE_FLOAT32);
        CHECK(w.open(pf));
        bool threw = false;
        try { w.setChannels(1); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        w.write(x.data(), 600);
        w.write(x.data() + 1200, 400);
        CHECK(w.getSamplesWritten() == 1000);
        w.close();
        CHECK(!w.isOpen());
        w.setSampleType(wav::SAMP_TYPE_INT16);
        w.setChannels(1);
        w.setSamplerate(15000);
        CHECK(w.open(pi));
        threw = false;
        try { w.write(x.data(), 10); } catch (const std::runtime_error&) { threw = true; }     // PCM types take device-converted bytes
        CHECK(threw);
        const int16_t pcm[5] = { 100, -200, 300, -400, 500 };
        w.writePcm(pcm, 5);
        w.close();
        wav::Reader r(pi);
        CHECK(r.isValid() && r.getSampleRate() == 15000 && r.getChannelCount() == 1 && r.getBitDepth() == 16 && r.getSampleType() == wav::SAMP_TYPE_INT16);
        int16_t back[5];
        CHECK(r.readSamples2(back, sizeof(back)) == sizeof(back) && back[0] == 100 && back[4] == 500);
        r.close();
        remove(pi);
    }
    printf("dropin_cpu ok\n");
    return 0;
}
