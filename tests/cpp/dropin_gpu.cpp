// GPU check of the drop-in IQFrontEnd + RxVFO classes (include/dsp/iq_frontend.h, include/dsp/channel/rx_vfo.h):
// a source thread feeds blocks through dsp::stream, the front end delivers FFT rows through the acquire/release
// callbacks and VFO output through the RxVFO's `out` stream, exactly as plugins consume them in the reference
// (radio_module.h:62,76-107).  Results are written to a file; tests/test_dropin_cpp.py compares them with the oracle.
//   usage: dropin_gpu <input.cf32> <nSamples> <blockSize> <output.bin>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <dsp/iq_frontend.h>

static std::vector<float> g_row(8192);
static std::vector<float> g_rows;
static int g_acquires = 0, g_releases = 0;
static float* acquire(void*) { g_acquires++; return g_row.data(); }
static void release(void*) { g_releases++; g_rows.insert(g_rows.end(), g_row.begin(), g_row.end()); }

int main(int argc, char** argv) {
    if (argc < 5) { return 2; }
    const int n = atoi(argv[2]), blk = atoi(argv[3]);
    std::vector<dsp::complex_t> x(n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(x.data(), sizeof(dsp::complex_t), n, f) != (size_t)n) { fprintf(stderr, "cannot read input\n"); return 2; }
    fclose(f);

    dsp::stream<dsp::complex_t> src;
    IQFrontEnd fe;
    fe.init(&src, 2.4e6, false, 1, false, 8192, 20.0, IQFrontEnd::NUTTALL, acquire, release, nullptr);
    dsp::channel::RxVFO* vfo = fe.addVFO("Radio", 50000.0, 12500.0, 100e3);
    if (!vfo) { return 3; }
    if (fe.addVFO("Radio", 50000.0, 12500.0, 0.0) != nullptr) { return 4; }   // duplicate name is refused
    fe.start();

    std::vector<dsp::complex_t> got;
    std::atomic<int> gotCount{ 0 };
    std::thread sink([&] {
        for (;;) {
            int c = vfo->out.read();
            if (c < 0) { break; }
            got.insert(got.end(), vfo->out.readBuf, vfo->out.readBuf + c);
            gotCount += c;
            vfo->out.flush();
        }
    });
    const int outPerBlock = blk / 48;                     // 2.4 MS/s -> 50 kS/s
    for (int i = 0; i < n; i += blk) {
        const int c = std::min(blk, n - i);
        memcpy(src.writeBuf, x.data() + i, c * sizeof(dsp::complex_t));
        if (i == 5 * blk) {
            // retune from the "GUI thread" while the worker runs; wait for the first five blocks' output so the
            // block at which the new offset takes effect is deterministic
            while (gotCount.load() < 5 * outPerBlock) { std::this_thread::sleep_for(std::chrono::milliseconds(1)); }
            vfo->setOffset(-50e3);
        }
        if (!src.swap(c)) { return 5; }
    }
    const int expectOut = (n / blk) * outPerBlock;
    for (int spin = 0; spin < 20000 && gotCount.load() < expectOut; spin++) { std::this_thread::sleep_for(std::chrono::milliseconds(1)); }
    // the input FIFO (IQFrontEnd::setBuffering): the source may now run ahead of the worker; a short burst below the
    // FIFO depth must still come out complete and in order
    const int before = gotCount.load();
    const size_t rowsBefore = g_rows.size();
    const int acqBefore = g_acquires, relBefore = g_releases;
    fe.setBuffering(true);
    const int burst = 8;
    for (int b = 0; b < burst; b++) {
        memcpy(src.writeBuf, x.data() + (size_t)b * blk, blk * sizeof(dsp::complex_t));
        if (!src.swap(blk)) { return 6; }
    }
    for (int spin = 0; spin < 20000 && gotCount.load() < before + burst * outPerBlock; spin++) { std::this_thread::sleep_for(std::chrono::milliseconds(1)); }
    if (gotCount.load() != before + burst * outPerBlock) { fprintf(stderr, "buffered burst: %d of %d samples\n", gotCount.load() - before, burst * outPerBlock); return 7; }
    // effective-sample-rate event
    static double announced = 0.0;
    EventHandler<double> onSr([](double v, void*) { announced = v; }, nullptr);
    fe.onEffectiveSampleRateChange.bindHandler(&onSr);
    vfo->out.stopReader();
    sink.join();
    fe.stop();
    fe.setSampleRate(2.4e6);
    if (announced != 2.4e6) { return 8; }
    fe.onEffectiveSampleRateChange.unbindHandler(&onSr);
    got.resize((size_t)before);                           // the file compared by the Python test holds the first phase only
    g_rows.resize(rowsBefore); g_acquires = acqBefore; g_releases = relBefore;

    FILE* o = fopen(argv[4], "wb");
    int hdr[4] = { (int)got.size(), (int)(g_rows.size() / 8192), g_acquires, g_releases };
    fwrite(hdr, sizeof(int), 4, o);
    fwrite(got.data(), sizeof(dsp::complex_t), got.size(), o);
    fwrite(g_rows.data(), sizeof(float), g_rows.size(), o);
    fclose(o);
    printf("dropin_gpu ok: %d vfo samples, %d fft rows\n", hdr[0], hdr[1]);
    return 0;
}
