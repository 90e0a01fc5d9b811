// GPU check of IQFrontEnd::removeVFO / unbindIQStream / setFFTSize while the front end is running
// (include/dsp/iq_frontend.h; reference behaviour: iq_frontend.cpp:181-202 -- the removed VFO's thread is stopped,
// its input unbound from the Splitter, the object deleted, and everything else keeps running).
//   * VFO "B" has no reader at all: the worker ends up blocked in B's out.swap().  removeVFO("B") from another thread
//     must return, and the other VFOs and the FFT rows must keep flowing.
//   * VFO "A" is consumed until right before it is removed (its reader is stopped the way a module stops its chain).
//   * a bound IQ stream is unbound while its reader has stalled.
//   * setFFTSize while rows are being delivered: no row of the old size may arrive once the setter has returned.
//   usage: dropin_remove_gpu <input.cf32> <nSamples> <blockSize>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <dsp/iq_frontend.h>

static std::vector<float> g_row(8192, 0.0f);
static std::atomic<int> g_rows{ 0 };
static float* acquire(void*) { return g_row.data(); }
static void release(void*) { g_rows++; }

static bool waitFor(std::atomic<int>& v, int target, int ms) {
    for (int i = 0; i < ms && v.load() < target; i++) { std::this_thread::sleep_for(std::chrono::milliseconds(1)); }
    return v.load() >= target;
}

int main(int argc, char** argv) {
    if (argc < 4) { return 2; }
    const int n = atoi(argv[2]), blk = atoi(argv[3]);
    std::vector<dsp::complex_t> x(n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(x.data(), sizeof(dsp::complex_t), n, f) != (size_t)n) { fprintf(stderr, "cannot read input\n"); return 2; }
    fclose(f);

    dsp::stream<dsp::complex_t> src;
    IQFrontEnd fe;
    fe.init(&src, 2.4e6, false, 1, false, 8192, 200.0, IQFrontEnd::NUTTALL, acquire, release, nullptr);   // one row per block
    dsp::channel::RxVFO* a = fe.addVFO("A", 50000.0, 12500.0, 100e3);
    dsp::channel::RxVFO* b = fe.addVFO("B", 50000.0, 12500.0, -50e3);
    dsp::channel::RxVFO* c = fe.addVFO("C", 24000.0, 2800.0, 300e3);
    if (!a || !b || !c) { return 3; }
    dsp::stream<dsp::complex_t> tap;                       // recorder-style consumer of the pre-processed IQ
    fe.bindIQStream(&tap);
    fe.start();

    std::atomic<int> gotA{ 0 }, gotC{ 0 }, gotTap{ 0 };
    auto drain = [](dsp::stream<dsp::complex_t>* s, std::atomic<int>* cnt, int limit) {
        for (int k = 0; limit < 0 || k < limit; k++) {
            int m = s->read();
            if (m < 0) { break; }
            *cnt += m;
            s->flush();
        }
    };
    std::thread sinkA(drain, &a->out, &gotA, -1);
    std::thread sinkC(drain, &c->out, &gotC, -1);
    std::thread sinkTap(drain, &tap, &gotTap, 2);          // reads two blocks, then stalls for good

    std::atomic<bool> feeding{ true };
    std::atomic<int> fed{ 0 };
    std::thread feeder([&] {
        for (size_t i = 0; feeding.load(); i = (i + blk) % (size_t)(n - blk)) {
            memcpy(src.writeBuf, x.data() + i, blk * sizeof(dsp::complex_t));
            if (!src.swap(blk)) { break; }
            fed++;
        }
    });

    // 1. the worker blocks on B (no reader) after B's first block, or on the stalled tap; everything stalls
    std::this_thread::sleep_for(std::chrono::milliseconds(300));
    const int c0 = gotC.load(), r0 = g_rows.load();
    fe.removeVFO("B");                                     // must come back although the worker sits in B's swap()
    fe.unbindIQStream(&tap);                               // ... or in the tap's
    tap.stopReader();
    sinkTap.join();
    tap.clearReadStop();
    if (!waitFor(gotC, c0 + 5 * (blk / 100), 10000)) { fprintf(stderr, "C starved after removeVFO(B): %d -> %d\n", c0, gotC.load()); return 4; }
    if (!waitFor(g_rows, r0 + 5, 10000)) { fprintf(stderr, "FFT rows stopped after removeVFO(B)\n"); return 5; }

    // 2. A is consumed until right before its removal (a module stops its own chain, then deletes the VFO)
    a->out.stopReader();
    sinkA.join();
    const int c1 = gotC.load();
    fe.removeVFO("A");
    fe.removeVFO("A");                                     // a second removal only complains
    if (!waitFor(gotC, c1 + 5 * (blk / 100), 10000)) { fprintf(stderr, "C starved after removeVFO(A)\n"); return 6; }

    // 3. FFT size change under load: after the setter returns, every delivered row has the new size.  The callback's
    // buffer keeps its 8192 floats, but the upper half is now a sentinel that only a stale, old-size row would touch.
    fe.setFFTSize(4096);
    for (int i = 4096; i < 8192; i++) { g_row[i] = -12345.0f; }
    const int r1 = g_rows.load();
    if (!waitFor(g_rows, r1 + 5, 10000)) { fprintf(stderr, "FFT rows stopped after setFFTSize\n"); return 7; }
    for (int i = 4096; i < 8192; i++) { if (g_row[i] != -12345.0f) { fprintf(stderr, "a row of the old size was delivered after setFFTSize\n"); return 8; } }

    // 4. a VFO added while running is served
    dsp::channel::RxVFO* d = fe.addVFO("D", 50000.0, 12500.0, 0.0);
    if (!d) { return 9; }
    std::atomic<int> gotD{ 0 };
    std::thread sinkD(drain, &d->out, &gotD, -1);
    if (!waitFor(gotD, 3 * (blk / 48), 10000)) { fprintf(stderr, "the VFO added while running got no data\n"); return 10; }

    feeding = false;
    c->out.stopReader(); d->out.stopReader();
    sinkC.join(); sinkD.join();
    src.stopWriter();
    feeder.join();
    src.clearWriteStop();
    fe.stop();
    printf("dropin_remove_gpu ok: fed %d blocks, C %d samples, rows %d\n", fed.load(), gotC.load(), g_rows.load());
    return 0;
}
