// TEST INFRASTRUCTURE ONLY (oracle port).  Part of oracle/libport_oracle.so; never linked into, or loaded by, the product.
//
// Self-contained restatement (no reference includes) of dsp::compression::ExperimentalFFTCompressor
// (core/src/dsp/compression/experimental_fft_compressor.h): setSampleRate :66-88, process :198-427, filterSignal :116-177,
// estimateNoise :181-189; helpers centeredSma / movingVariance (core/src/utils/arrays.cpp:474-507), linearInterpolateHoles
// (:433-470), percentile_sampling / kthSmallest / partition (core/src/utils/strings.h:93-175), swapfft (arrays.cpp:1128-1134),
// VOLK's generic power spectrum (oracle/shim/volk/volk.h).  The block turns every fftSize input samples into one frame of
// fftSize frequency-domain samples, ten frames late: bins that do not rise above the tracked noise floor are zeroed, the
// others are amplitude-companded (|X| -> |X|^(1/4)), ready for the int8 sample packets of the server.
// Pinned bit for bit against the reference's header compiled into oracle/_ref (ref_compressor.cpp) by
// tests/test_oracle_compressor.py.  The DFT is oracle/fft64.h on both sides; all else is sequential fp32 / fp64 in
// reference order (-ffp-contract=off).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <vector>

#include "fft64.h"
#include "oracle_api.h"

namespace {

struct cf { float re, im; };

float log2Clamped(float x) {                                  // as oracle/shim/volk/volk.h: log2 with 0 / inf clamped to +-127
    const float l = log2f(x);
    return std::isinf(l) ? copysignf(127.0f, l) : l;
}

std::vector<float> centeredSma(const std::vector<float>& in, int winsize) {     // arrays.cpp:474-495
    const int limit = (int)in.size();
    std::vector<float> out((size_t)limit, 0.0f);
    float total = 0;
    const int win2 = winsize / 2;
    for (int i = 0; i < winsize; i++) { total += in[i]; }
    for (int i = winsize; i < limit; i++) {
        out[i - win2] = total / winsize;
        total += in[i] - in[i - winsize];
    }
    for (int i = 0; i < winsize - win2; i++) { out[i] = out[winsize - win2]; }
    for (int i = limit - win2; i < limit; i++) { out[i] = out[limit - win2 - 1]; }
    return out;
}

std::vector<float> movingVariance(const std::vector<float>& in, int winsize) {  // arrays.cpp:497-507
    const std::vector<float> mean = centeredSma(in, winsize);
    std::vector<float> sq(mean.size());
    for (size_t i = 0; i < mean.size(); i++) { sq[i] = (in[i] - mean[i]) * (in[i] - mean[i]); }
    return centeredSma(sq, winsize);
}

bool fillHoles(float* arr, int n) {                           // arrays.cpp:433-470
    int first = -1, last = -1;
    for (int q = 0; q < n; q++) {
        const float val = arr[q];
        if (first < 0 && val != 0) { first = q; }
        if (val != 0) {
            if (last != -1 && q - last > 1) {
                const float d = (val - arr[last]) / (q - last);
                float running = arr[last];
                for (int w = last + 1; w < q; w++) { running += d; arr[w] = running; }
            }
            last = q;
        }
    }
    if (first < 0 || last < 0) { return false; }
    for (int q = first - 1; q >= 0; q--) { arr[q] = arr[first]; }
    for (int q = last + 1; q < n; q++) { arr[q] = arr[last]; }
    return true;
}

// strings.h:93-175.  The pivot is held in an int there (`int x = arr[r]`): kept, it decides which element comes out.
float kthSmallest(float* arr, int l, int r, int k) {
    for (;;) {
        if (!(k > 0 && k <= r - l + 1)) { return arr[0]; }
        const int x = (int)arr[r];
        int i = l;
        for (int j = l; j <= r - 1; j++) {
            if (arr[j] <= x) { std::swap(arr[i], arr[j]); i++; }
        }
        std::swap(arr[i], arr[r]);
        const int index = i;
        if (index - l == k - 1) { return arr[index]; }
        if (index - l > k - 1) { r = index - 1; continue; }
        k = k - index + l - 1;
        l = index + 1;
    }
}

float percentileSampling(std::vector<float> arr, double p) {
    int n = (int)arr.size();
    if (n == 0) { return 0; }
    const int targetPoints = 100;
    float* data = arr.data();
    std::vector<float> sampled;
    if (n > 2 * targetPoints) {
        const float step = (n - 1) / (float)targetPoints;
        sampled.resize(targetPoints);
        for (int z = 0; z < targetPoints; z++) { sampled[z] = data[(int)std::floor(step * z)]; }
        data = sampled.data();
        n = (int)sampled.size();
    }
    const double k = (n - 1) * p;
    return kthSmallest(data, 0, n - 1, (int)k);
}

struct FftCompressor {
    int sampleRate = 0, fftSize = 1024;
    double lossRate = 1.0;
    const int sliceMsec = 50, minRecents = 10, noiseNPoints = 16, signalWidth = 300;
    float hzTick = 0; int smallTick = 0, largeTick = 0;
    float prevAllowance = 0;
    std::vector<float> window, noiseFigure;
    std::vector<int32_t> masked;
    std::vector<cf> input;
    std::deque<std::vector<cf>> cleanFreq;
    std::deque<std::vector<float>> cleanMag, windowedMag;
    orc_fft64::Plan plan;

    void setSampleRate(int sr) {                              // :66-88
        if (sr == sampleRate) { return; }
        input.clear();
        sampleRate = sr;
        fftSize = sr * sliceMsec / 1000;
        fftSize = (int)pow(2, floor(log2(fftSize)));
        plan.init(fftSize, -1);
        window.resize((size_t)fftSize);
        for (int i = 0; i < fftSize; i++) {                   // blackmanWindowElement(i + 5, fftSize + 10) :15-27
            const double factor = 2.0 * M_PI * (i + 5) / (fftSize + 10 - 1);
            window[i] = (float)(0.42 - 0.5 * std::cos(factor) + 0.08 * std::cos(2 * factor));
        }
        hzTick = ((float)sr) / fftSize;
        smallTick = (int)(signalWidth / hzTick);
        largeTick = smallTick * 10;
        noiseFigure.clear();
    }

    void transform(const std::vector<cf>& in, std::vector<cf>& out) const {
        std::vector<orc_fft64::cd> a((size_t)fftSize), b((size_t)fftSize);
        for (int i = 0; i < fftSize; i++) { a[i] = orc_fft64::cd(in[i].re, in[i].im); }
        plan.run(a.data(), b.data());
        out.resize((size_t)fftSize);
        for (int i = 0; i < fftSize; i++) { out[i] = cf{ (float)b[i].real(), (float)b[i].imag() }; }
        const int half = fftSize / 2;                         // swapfft
        for (int i = 0; i < half; i++) { std::swap(out[i], out[i + half]); }
    }

    std::vector<float> powerSpectrum(const std::vector<cf>& s) const {
        std::vector<float> p(s.size());
        const float inv = 1.0f / (float)fftSize;
        for (size_t i = 0; i < s.size(); i++) {
            const float r = s[i].re * inv, q = s[i].im * inv;
            p[i] = 3.01029995663981209120f * log2Clamped(r * r + q * q);
        }
        return p;
    }

    static std::vector<float> meanOf(const std::deque<std::vector<float>>& rows, int n) {
        std::vector<float> acc((size_t)n, 0.0f);
        for (auto& r : rows) { for (int i = 0; i < n; i++) { acc[i] = acc[i] + r[i]; } }
        const float inv = (float)(1.0 / (float)rows.size());
        for (int i = 0; i < n; i++) { acc[i] = acc[i] * inv; }
        return acc;
    }

    std::vector<float> filterSignal(const std::vector<float>& windowedMags, const std::vector<float>& clearMags, cf* frame) {   // :116-177
        const std::vector<float> mvar = movingVariance(windowedMags, noiseNPoints);
        const double newAllowance = lossRate * percentileSampling(mvar, .15);
        const double allowance = newAllowance * 0.1 + prevAllowance * 0.9;
        prevAllowance = (float)allowance;
        std::vector<float> cma = centeredSma(windowedMags, largeTick);
        for (int i = 0; i < fftSize; i++) { if (mvar[i] > allowance) { cma[i] = 0; } }
        fillHoles(cma.data(), (int)cma.size());
        cma = centeredSma(cma, largeTick);
        const std::vector<float> cmax = centeredSma(cma, 5 * largeTick);
        std::vector<float> diff((size_t)fftSize);
        for (int i = 0; i < fftSize; i++) { diff[i] = (float)fabs(cma[i] - cmax[i]); }
        const float cmaxAllow = percentileSampling(diff, .15);
        for (int i = 0; i < fftSize; i++) { if (diff[i] > cmaxAllow) { cma[i] = 0; } }
        fillHoles(cma.data(), (int)cma.size());
        cma = centeredSma(cma, largeTick);
        std::vector<float> mask((size_t)fftSize, 0.0f);
        for (int i = 0; i < fftSize; i++) { if (clearMags[i] > cma[i] + allowance) { mask[i] = 1; } }
        for (size_t i = 0; i + 1 < masked.size(); i += 2) {
            const int tickFrom = (int)(fftSize / 2 + masked[i] / hzTick), tickTo = (int)(fftSize / 2 + masked[i + 1] / hzTick);
            for (int j = tickFrom; j < tickTo; j++) { if (j >= 0 && j < fftSize) { mask[j] = 1; } }
        }
        mask = centeredSma(mask, signalWidth / 8);
        for (int i = 0; i < fftSize; i++) { if (mask[i] == 0) { frame[i] = cf{ 0, 0 }; } }
        return cma;
    }

    int process(cf* out) {                                    // :198-427 (receive mode)
        if ((int)input.size() < fftSize) { return 0; }
        std::vector<cf> frame(input.begin(), input.begin() + fftSize), spec;
        input.erase(input.begin(), input.begin() + fftSize);
        transform(frame, spec);
        cleanFreq.push_back(spec);
        cleanMag.push_back(powerSpectrum(spec));
        for (int i = 0; i < fftSize; i++) { frame[i] = cf{ frame[i].re * window[i], frame[i].im * window[i] }; }
        transform(frame, spec);
        windowedMag.push_back(powerSpectrum(spec));
        if ((int)cleanFreq.size() < minRecents) { return 0; }
        memcpy(out, cleanFreq[0].data(), (size_t)fftSize * sizeof(cf));
        const std::vector<float> windowedSpectrum = meanOf(windowedMag, fftSize), clearSpectrum = meanOf(cleanMag, fftSize);
        if (lossRate > 0) {
            const std::vector<float> nf = filterSignal(windowedSpectrum, clearSpectrum, out);
            const int nslices = 30, slice = fftSize / nslices;  // estimateNoise :181-189
            noiseFigure.assign(nslices, 0.0f);
            for (int i = 0; i < nslices; i++) { noiseFigure[i] = 7 + nf[(size_t)i * slice + slice / 2]; }
        }
        for (int i = 0; i < fftSize; i++) {                   // amplitude to its fourth root (:257-266)
            if (out[i].re == 0 && out[i].im == 0) { continue; }
            const float amp = sqrtf((out[i].re * out[i].re) + (out[i].im * out[i].im));
            const float namp = sqrtf(sqrtf(amp));
            const float g = namp / amp;
            out[i].re *= g; out[i].im *= g;
        }
        cleanFreq.pop_front();
        if ((int)cleanMag.size() > minRecents * 2) { cleanMag.pop_front(); windowedMag.pop_front(); }
        return fftSize;
    }
};

} // namespace

extern "C" {

void* orc_fftcomp_create(int sampleRate, double lossRate) {
    auto* c = new FftCompressor;
    c->setSampleRate(sampleRate);
    c->lossRate = lossRate;
    return c;
}
void orc_fftcomp_destroy(void* h) { delete (FftCompressor*)h; }
int orc_fftcomp_fft_size(void* h) { return ((FftCompressor*)h)->fftSize; }
void orc_fftcomp_set_masked(void* h, const int* fromTo, int n) { ((FftCompressor*)h)->masked.assign(fromTo, fromTo + n); }
int orc_fftcomp_process(void* hv, int count, const float* in, float* out) {
    auto* c = (FftCompressor*)hv;
    const cf* x = (const cf*)in;
    c->input.insert(c->input.end(), x, x + count);
    return c->process((cf*)out);
}
int orc_fftcomp_noise_figure(void* hv, float* out, int cap) {
    auto& v = ((FftCompressor*)hv)->noiseFigure;
    const int n = (int)v.size();
    if (out) { memcpy(out, v.data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}

}
