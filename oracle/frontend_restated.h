// TEST INFRASTRUCTURE ONLY (oracle).
//
// CPU restatement of the parts of the IQ-front-end path that cannot be compiled
// from the reference in this image: signal_path/iq_frontend.cpp pulls GL through
// gui/gui.h (:5), utils/arrays.cpp needs FFTW3f (absent, un-pinned), and
// gui/widgets/waterfall.cpp needs ImGui + OpenGL.  These are plain CPU
// functions in the reference (no shaders); each one below cites the lines it
// follows.  Shared by ref_harness.cpp and oracle_port.cpp.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orc_restated {

// IQFrontEnd::genReshapeParams — signal_path/iq_frontend.h:88-92
inline void reshapeParams(double sampleRate, int size, double rate, int* skip, int* nz) {
    int fftInterval = (int)std::round(sampleRate / rate);
    *nz = std::min<int>(fftInterval, size);
    *skip = fftInterval - *nz;
}

// Forward DFT standing in for fftwf_plan_dft_1d(N, in, out, FFTW_FORWARD, FFTW_ESTIMATE) +
// fftwf_execute (utils/arrays.cpp:522,537): unnormalised, sign -1.  FFTW3f is a
// third-party dependency absent from /root/reference; it is restated as an
// iterative radix-2 transform carried out in float64 and narrowed to fp32 at
// the end, i.e. the mathematically exact DFT to fp32 precision (fftwf itself is
// only accurate to ~1e-7*sqrt(log2 N) relative).  N must be a power of two.
inline void fftForwardF64(std::vector<double>& re, std::vector<double>& im) {
    const size_t n = re.size();
    // bit reversal
    for (size_t i = 1, j = 0; i < n; i++) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) { j ^= bit; }
        j ^= bit;
        if (i < j) { std::swap(re[i], re[j]); std::swap(im[i], im[j]); }
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        const size_t half = len >> 1;
        std::vector<double> wr(half), wi(half);
        for (size_t k = 0; k < half; k++) {
            const double a = -2.0 * M_PI * (double)k / (double)len;
            wr[k] = std::cos(a); wi[k] = std::sin(a);
        }
        for (size_t i = 0; i < n; i += len) {
            for (size_t k = 0; k < half; k++) {
                const size_t a = i + k, b = i + k + half;
                const double tr = re[b] * wr[k] - im[b] * wi[k];
                const double ti = re[b] * wi[k] + im[b] * wr[k];
                re[b] = re[a] - tr; im[b] = im[a] - ti;
                re[a] += tr; im[a] += ti;
            }
        }
    }
}

// IQFrontEnd::handler — signal_path/iq_frontend.cpp:267-287.
//   :271 fftIn[i] = data[i] * window[i]  (fp32 multiply, i < nz; tail stays zero, :317)
//   :274 forward FFT of size N
//   :282 out[k] = 10*log10((re^2+im^2)/N^2) via the VOLK power-spectrum kernel
// The caller passes the power-spectrum routine so that ref_harness uses the shim's
// and the port uses its own restatement.
template <typename PowerSpectrumFn>
inline void fftFrame(const float* iq, const float* window, int nz, int N, float* outDb, float* spectrum, PowerSpectrumFn ps) {
    std::vector<double> re(N, 0.0), im(N, 0.0);
    for (int i = 0; i < nz; i++) {
        const float r = iq[2 * i] * window[i];
        const float q = iq[2 * i + 1] * window[i];
        re[i] = r; im[i] = q;
    }
    fftForwardF64(re, im);
    std::vector<float> spec(2 * (size_t)N);
    for (int k = 0; k < N; k++) { spec[2 * k] = (float)re[k]; spec[2 * k + 1] = (float)im[k]; }
    if (spectrum) { memcpy(spectrum, spec.data(), sizeof(float) * 2 * (size_t)N); }
    if (outDb) { ps(outDb, spec.data(), (float)N, N); }
}

// doZoom index walk — gui/widgets/waterfall.cpp:98-152.  fp32 running start index.
inline void zoomIndices(int offset, int width, int inSize, int outSize, int* sIdOut, int* uFactorOut) {
    if (width > 524288) { width = 524288; }   // :100-102
    if (offset < 0) { offset = 0; }           // :103-106
    float factor = (float)width / (float)outSize;
    int sFactor = (int)ceilf(factor);
    float id = (float)offset;
    for (int i = 0; i < outSize; i++) {
        int sId = (int)id;
        int uFactor = (sId + sFactor > inSize) ? inSize - sId : sFactor;
        sIdOut[i] = sId;
        uFactorOut[i] = uFactor;
        id += factor;
    }
}

// doZoom — gui/widgets/waterfall.cpp:98-154: max over uFactor bins from sId (max is
// order-independent, so the 8-way unrolled reduction of :127-149 needs no restating).
inline void doZoom(int offset, int width, int inSize, int outSize, const float* in, float* out) {
    std::vector<int> sId(outSize), uF(outSize);
    zoomIndices(offset, width, inSize, outSize, sId.data(), uF.data());
    for (int i = 0; i < outSize; i++) {
        float m = -INFINITY;
        for (int j = 0; j < uF[i]; j++) { m = std::max<float>(m, in[sId[i] + j]); }
        out[i] = m;
    }
}

// pushFFT palette index — gui/widgets/waterfall.cpp:1229-1235, WATERFALL_RESOLUTION = 1000 (waterfall.h)
inline void paletteIndices(const float* row, int n, float wfMin, float wfMax, int* idx) {
    const float dataRange = wfMax - wfMin;
    for (int j = 0; j < n; j++) {
        float pixel = (std::clamp<float>(row[j], wfMin, wfMax) - wfMin) / dataRange;
        idx[j] = (int)(pixel * (1000 - 1));
    }
}

// WaterFall::updatePallette — gui/widgets/waterfall.cpp:1282-1295
inline void makePalette(const float* colors, int colorCount, uint32_t* lut) {
    const int RES = 1000;
    for (int i = 0; i < RES; i++) {
        // position of LUT entry i along the colour list, in fp32 like the reference
        const float pos = ((float)i / (float)RES) * colorCount;
        const int lo = std::clamp<int>((int)floorf(pos), 0, colorCount - 1);
        const int hi = std::clamp<int>((int)ceilf(pos), 0, colorCount - 1);
        const float ratio = pos - lo;
        uint32_t ch[3];
        for (int c = 0; c < 3; c++) {
            // the blend is evaluated in double (1.0 - ratio) and narrowed to float, then truncated
            float v = (colors[lo * 3 + c] * (1.0 - ratio)) + (colors[hi * 3 + c] * (ratio));
            ch[c] = (uint32_t)v;
        }
        lut[i] = ((uint32_t)255 << 24) | (ch[2] << 16) | (ch[1] << 8) | ch[0];
    }
}

// WaterFall::rawFFTIndex — gui/widgets/waterfall.cpp:761-763
inline int rawFFTIndex(double frequency, double wholeBandwidth, int rawFFTSize) {
    return std::clamp<int>((int)(((frequency / (wholeBandwidth / 2.0)) * (double)(rawFFTSize / 2)) + (rawFFTSize / 2)), 0, rawFFTSize);
}

// WaterFall::calculateVFOSignalInfo — gui/widgets/waterfall.cpp:688-759.  strength / snr of one VFO from one raw dB
// row: the peak inside the VFO band against the mean of the two side bands (each one bandwidth wide), corrected by how
// far that mean sits above the quietest quarter of the side-band bins.  Returns false where the reference does.
inline bool vfoSignalInfo(const float* fftLine, int rawFFTSize, double wholeBandwidth, double centerOffset, double bandwidth,
                          float* strength, float* snr) {
    const int minSide = rawFFTIndex(centerOffset - bandwidth, wholeBandwidth, rawFFTSize);
    const int minOff = rawFFTIndex(centerOffset - bandwidth / 2.0, wholeBandwidth, rawFFTSize);
    const int maxOff = rawFFTIndex(centerOffset + bandwidth / 2.0, wholeBandwidth, rawFFTSize);
    const int maxSide = rawFFTIndex(centerOffset + bandwidth, wholeBandwidth, rawFFTSize);
    double avg = 0, qavg = 0;
    std::vector<float> side;
    for (int i = minSide; i < minOff; i++) { side.push_back(fftLine[i]); avg += fftLine[i]; }
    for (int i = maxOff + 1; i < maxSide; i++) { side.push_back(fftLine[i]); avg += fftLine[i]; }
    if (side.empty()) { return false; }
    const size_t avgCount = side.size();
    std::sort(side.begin(), side.end());
    const size_t lower = side.size() / 4;
    if (lower == 0) { return false; }
    const float kth = side[lower];
    for (size_t i = 0; i < side.size(); i++) { if (side[i] <= kth) { qavg += side[i]; } }   // ties above the quartile are included, as in the reference
    qavg /= (double)lower;
    avg /= (double)avgCount;
    const double avgdiff = avg - qavg;
    float mx = -INFINITY;
    for (int i = minOff; i <= maxOff && i < rawFFTSize; i++) { if (fftLine[i] > mx) { mx = fftLine[i]; } }
    *strength = (float)(mx - avgdiff);
    *snr = (float)(mx - avg - avgdiff);
    return true;
}

// pushFFT smoothing and peak hold — gui/widgets/waterfall.cpp:1247-1252,1269-1272 (buffers start at -1000, :1073,1090).
// One zoomed row in, smoothed row out (in place), state carried in smoothingBuf / hold.
inline void smoothAndHold(float* latest, float* smoothingBuf, float* hold, int width, bool smoothing, float alpha, bool doHold, float holdSpeed) {
    if (smoothing) {
        const float beta = 1.0f - alpha;
        for (int i = 0; i < width; i++) {
            const float a = latest[i] * alpha;
            const float b = smoothingBuf[i] * beta;
            smoothingBuf[i] = a + b;
            latest[i] = smoothingBuf[i];
        }
    }
    if (doHold) {
        for (int i = 1; i < width; i++) { hold[i] = std::max<float>(latest[i], hold[i] - holdSpeed); }
    }
}

} // namespace orc_restated
