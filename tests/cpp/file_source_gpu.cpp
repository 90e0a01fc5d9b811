// A baseband WAV file played through the front end, as in the reference's in-app test (tests/baseband_file_test.cpp:
// File source -> IQFrontEnd -> radio VFO): WavFileSource (include/dsp/source/file_source.h) once as the reference's
// worker thread feeding IQFrontEnd through its dsp::stream, once pumping the file's int16 blocks straight into a
// b200_frontend.  The VFO outputs and dB rows of both go to a file; tests/test_dropin_cpp.py compares them with the oracle.
//   usage: file_source_gpu <in.wav> <blocks> <out.bin>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <dsp/iq_frontend.h>
#include <dsp/source/file_source.h>

static std::vector<float> g_row(8192);
static std::vector<float> g_rows;
static float* acquire(void*) { return g_row.data(); }
static void release(void*) { g_rows.insert(g_rows.end(), g_row.begin(), g_row.end()); }

int main(int argc, char** argv) {
    if (argc < 4) { return 2; }
    const int blocks = atoi(argv[2]);
    dsp::source::WavFileSource src;
    if (!src.open(argv[1])) { fprintf(stderr, "open: %s\n", src.error.c_str()); return 3; }
    const double sr = src.getSampleRate();
    const int n = src.blockSize();

    // ---- 1. the reference's arrangement: source worker -> stream -> IQFrontEnd -> RxVFO::out
    std::vector<dsp::complex_t> gotA;
    {
        IQFrontEnd fe;
        fe.init(&src.stream, sr, false, 1, false, 8192, 20.0, IQFrontEnd::NUTTALL, acquire, release, nullptr);
        dsp::channel::RxVFO* vfo = fe.addVFO("Radio", 24000.0, 2800.0, 10e3);
        if (!vfo) { return 4; }
        fe.start();
        std::atomic<int> blocksSeen{ 0 };
        std::thread sink([&] {
            for (;;) {
                int c = vfo->out.read();
                if (c < 0) { break; }
                gotA.insert(gotA.end(), vfo->out.readBuf, vfo->out.readBuf + c);
                vfo->out.flush();
                if (++blocksSeen >= blocks) { break; }
            }
        });
        src.start();
        sink.join();
        src.stop();
        vfo->out.stopReader();
        fe.stop();
    }
    const size_t rowsA = g_rows.size() / 8192;

    // ---- 2. the same file pumped into a C-ABI front end, int16 over PCIe
    src.rewind();
    std::vector<dsp::complex_t> gotB;
    std::vector<float> rowsB;
    {
        b200_frontend* fe = nullptr;
        if (b200_frontend_create(&fe, sr, 1, 0, 0, 8192, 20.0, B200_WINDOW_NUTTALL, n, 0) != B200_OK) { fprintf(stderr, "%s\n", b200_last_error()); return 5; }
        b200_vfo_bank* bank = b200_frontend_vfo_bank(fe);
        b200_vfo_bank_set_rotator_mode(bank, B200_ROTATOR_EXACT);
        const int v = b200_vfo_bank_add(bank, 24000.0, 2800.0, 10e3);
        std::vector<float> out((size_t)b200_vfo_max_out(bank, v, n) * 2), frames((size_t)4 * 8192);
        float* outs[1] = { out.data() };
        int counts[1] = { 0 };
        const int rc = src.pump(fe, blocks, frames.data(), 4, outs, nullptr, counts, [&](int) {
            const dsp::complex_t* p = reinterpret_cast<const dsp::complex_t*>(out.data());
            gotB.insert(gotB.end(), p, p + counts[0]);
            const int nf = b200_frontend_frame_count(fe);
            rowsB.insert(rowsB.end(), frames.begin(), frames.begin() + (size_t)nf * 8192);
        });
        if (rc != B200_OK) { fprintf(stderr, "pump: %s\n", b200_last_error()); return 6; }
        b200_frontend_destroy(fe);
    }

    FILE* o = fopen(argv[3], "wb");
    int hdr[4] = { (int)gotA.size(), (int)gotB.size(), (int)rowsA, (int)(rowsB.size() / 8192) };
    fwrite(hdr, sizeof(int), 4, o);
    fwrite(gotA.data(), sizeof(dsp::complex_t), gotA.size(), o);
    fwrite(gotB.data(), sizeof(dsp::complex_t), gotB.size(), o);
    fwrite(g_rows.data(), sizeof(float), rowsA * 8192, o);
    fwrite(rowsB.data(), sizeof(float), rowsB.size(), o);
    fclose(o);
    printf("file_source_gpu ok: %d / %d vfo samples, %d / %d rows, block %d\n", hdr[0], hdr[1], hdr[2], hdr[3], n);
    return 0;
}
