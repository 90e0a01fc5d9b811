// Compile check of the drop-in dsp::compression::ExperimentalFFTCompressor (include/dsp/compression/experimental_fft_compressor.h)
// and its disabled pass-through path, which needs no GPU (reference: experimental_fft_compressor.h:435-441).
#include <cstdio>
#include <dsp/compression/experimental_fft_compressor.h>

int main() {
    dsp::stream<dsp::complex_t> src;
    dsp::compression::ExperimentalFFTCompressor c(&src);
    if (c.isEnabled()) { return 1; }
    c.start();
    for (int i = 0; i < 100; i++) { src.writeBuf[i] = dsp::complex_t{ (float)i, -(float)i }; }
    if (!src.swap(100)) { return 2; }
    if (c.out.read() != 100) { return 3; }
    for (int i = 0; i < 100; i++) { if (c.out.readBuf[i].re != (float)i || c.out.readBuf[i].im != -(float)i) { return 4; } }
    c.out.flush();
    c.stop();
    printf("dropin_compressor_cpu ok\n");
    return 0;
}
