// GPU check of the drop-in dsp::IFNRLogMMSE (include/noise_reduction_logmmse/if_nr.h) as a stand-alone block on
// dsp::stream buffers (reference: misc_modules/noise_reduction_logmmse/src/if_nr.h:63-160): blocks in, de-noised
// frames out -- nothing at first, then a burst, then whole hops -- written to a file for the Python side to compare
// with the oracle.   usage: dropin_logmmse_gpu <input.cf32> <nSamples> <blockSize> <sampleRate> <out.bin>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <noise_reduction_logmmse/if_nr.h>

int main(int argc, char** argv) {
    if (argc < 6) { return 2; }
    const int n = atoi(argv[2]), blk = atoi(argv[3]), sr = atoi(argv[4]);
    std::vector<dsp::complex_t> x(n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(x.data(), sizeof(dsp::complex_t), n, f) != (size_t)n) { fprintf(stderr, "cannot read input\n"); return 2; }
    fclose(f);

    dsp::stream<dsp::complex_t> src;
    dsp::IFNRLogMMSE nr;
    nr.sampleRateSource = [sr] { return sr; };          // the application's sigpath::iqFrontEnd.getSampleRate()
    nr.init(&src);
    nr.start();
    std::vector<dsp::complex_t> y;
    std::vector<int> counts;
    for (int i = 0; i + blk <= n; i += blk) {
        memcpy(src.writeBuf, x.data() + i, blk * sizeof(dsp::complex_t));
        if (!src.swap(blk)) { return 3; }
        const int m = nr.out.read();
        if (m < 0) { return 4; }
        counts.push_back(m);
        y.insert(y.end(), nr.out.readBuf, nr.out.readBuf + m);
        nr.out.flush();
    }
    nr.stop();
    FILE* o = fopen(argv[5], "wb");
    const int hdr[2] = { (int)counts.size(), (int)y.size() };
    fwrite(hdr, sizeof(int), 2, o);
    fwrite(counts.data(), sizeof(int), counts.size(), o);
    fwrite(y.data(), sizeof(dsp::complex_t), y.size(), o);
    fclose(o);
    printf("dropin_logmmse_gpu ok: %d blocks in, %d samples out\n", (int)counts.size(), (int)y.size());
    return 0;
}
