// GPU check of the drop-in dsp::detector::SignalDetector (include/dsp/detector/signal_detector.h), both ways it runs:
//   1. as IQFrontEnd::detectorPreprocessor (reference: signal_path/iq_frontend.cpp:37-44, toggled like
//      gui/menus/display.cpp:144): off at first, toggled on, fed through the running front end;
//   2. stand-alone on dsp::stream buffers (reference: signal_detector.cpp:359-414): passes the block through unchanged.
// Both write sigs_smoothed (gui/main_window.cpp:525-526 reads it) for the Python side to compare with the oracle.
//   usage: dropin_detector_gpu <input.cf32> <nSamples> <blockSize> <sampleRate> <out.bin>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <dsp/iq_frontend.h>

static std::vector<float> g_row(4096, 0.0f);
static float* acquire(void*) { return g_row.data(); }
static void release(void*) {}

int main(int argc, char** argv) {
    if (argc < 6) { return 2; }
    const int n = atoi(argv[2]), blk = atoi(argv[3]);
    const double sr = atof(argv[4]);
    std::vector<dsp::complex_t> x(n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(x.data(), sizeof(dsp::complex_t), n, f) != (size_t)n) { fprintf(stderr, "cannot read input\n"); return 2; }
    fclose(f);

    // ---- 0. toggled while the front end runs: nothing while off (the block fed before is known to be through once the
    // next swap has returned), rows once on
    {
        dsp::stream<dsp::complex_t> src;
        IQFrontEnd fe;
        fe.init(&src, sr, false, 1, false, 4096, 20.0, IQFrontEnd::NUTTALL, acquire, release, nullptr);
        if (fe.detectorPreprocessor.isEnabled()) { fprintf(stderr, "the detector must start disabled\n"); return 3; }
        if (fe.detectorPreprocessor.fftSize() != (int)(sr * (1 / 10.0))) { fprintf(stderr, "wrong row length %d\n", fe.detectorPreprocessor.fftSize()); return 3; }
        fe.start();
        const int rowLen = fe.detectorPreprocessor.fftSize();
        for (int k = 0; k < 2; k++) {
            memcpy(src.writeBuf, x.data(), 2 * rowLen * sizeof(dsp::complex_t));
            if (!src.swap(k == 0 ? 2 * rowLen : 16)) { return 4; }
        }
        if (fe.detectorPreprocessor.rows() != 0) { fprintf(stderr, "a disabled detector made rows\n"); return 3; }
        fe.togglePreprocessor(&fe.detectorPreprocessor, true);
        if (!fe.detectorPreprocessor.isEnabled()) { return 5; }
        for (int k = 0; k < 3; k++) {
            memcpy(src.writeBuf, x.data(), 2 * rowLen * sizeof(dsp::complex_t));
            if (!src.swap(k < 2 ? 2 * rowLen : 16)) { return 4; }
        }
        src.stopWriter();
        fe.stop();
        src.clearWriteStop();
        if (fe.detectorPreprocessor.rows() < 3) { fprintf(stderr, "the detector toggled on made %lld rows\n", fe.detectorPreprocessor.rows()); return 3; }
        fe.togglePreprocessor(&fe.detectorPreprocessor, false);
        if (fe.detectorPreprocessor.isEnabled()) { return 5; }
    }

    // ---- 1. inside the front end, on from the first sample
    std::vector<float> viaFrontEnd;
    long long feRows = 0;
    {
        dsp::stream<dsp::complex_t> src;
        IQFrontEnd fe;
        fe.init(&src, sr, false, 1, false, 4096, 20.0, IQFrontEnd::NUTTALL, acquire, release, nullptr);
        fe.togglePreprocessor(&fe.detectorPreprocessor, true);
        fe.detectorPreprocessor.setCenterFrequency(7.1e6);
        fe.start();
        for (int i = 0; i + blk <= n; i += blk) {
            memcpy(src.writeBuf, x.data() + i, blk * sizeof(dsp::complex_t));
            if (!src.swap(blk)) { return 4; }
        }
        // one more small block: once its swap has returned, the worker has finished the block before it
        memcpy(src.writeBuf, x.data(), 16 * sizeof(dsp::complex_t));
        if (!src.swap(16)) { return 4; }
        src.stopWriter();
        fe.stop();
        src.clearWriteStop();
        viaFrontEnd = fe.detectorPreprocessor.sigs_smoothed;
        feRows = fe.detectorPreprocessor.rows();
    }

    // ---- 2. stand-alone block
    std::vector<float> standAlone;
    long long saRows = 0;
    bool passThrough = true;
    {
        dsp::stream<dsp::complex_t> src;
        dsp::detector::SignalDetector det;
        det.init(&src);
        det.setSampleRate(sr);
        det.start();
        for (int i = 0; i + blk <= n; i += blk) {
            memcpy(src.writeBuf, x.data() + i, blk * sizeof(dsp::complex_t));
            if (!src.swap(blk)) { return 6; }
            if (det.out.read() != blk) { return 7; }
            if (memcmp(det.out.readBuf, x.data() + i, blk * sizeof(dsp::complex_t)) != 0) { passThrough = false; }
            det.out.flush();
        }
        det.stop();
        standAlone = det.sigs_smoothed;
        saRows = det.rows();
    }
    if (!passThrough) { fprintf(stderr, "the detector changed the samples it passes through\n"); return 8; }

    FILE* o = fopen(argv[5], "wb");
    const int hdr[4] = { (int)viaFrontEnd.size(), (int)standAlone.size(), (int)feRows, (int)saRows };
    fwrite(hdr, sizeof(int), 4, o);
    fwrite(viaFrontEnd.data(), sizeof(float), viaFrontEnd.size(), o);
    fwrite(standAlone.data(), sizeof(float), standAlone.size(), o);
    fclose(o);
    printf("dropin_detector_gpu ok: %lld rows through the front end, %lld stand-alone\n", feRows, saRows);
    return 0;
}
