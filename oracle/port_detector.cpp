// TEST INFRASTRUCTURE ONLY (oracle port).  Part of oracle/libport_oracle.so; never linked into, or loaded by, the product.
//
// Self-contained restatement (no reference includes) of dsp::detector::SignalDetector
// (core/src/dsp/detector/signal_detector.cpp), the front end's detector pre-processor
// (signal_path/iq_frontend.cpp:37-44).  Pinned bit for bit against the reference's own sources compiled into
// oracle/_ref (ref_detector.cpp) by tests/test_oracle_detector.py.  The DFT is oracle/fft64.h on both sides (FFTW is
// un-vendored); all other arithmetic is sequential fp32 in reference order (-ffp-contract=off).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <limits>
#include <vector>

#include "fft64.h"
#include "oracle_api.h"

namespace {

struct cf { float re, im; };

// window::blackman (dsp/window/blackman.h:5-8 over cosine.h:7-16): 0.42 - 0.5 cos(2 pi n / N) + 0.08 cos(4 pi n / N)
double blackmanAt(double n, double N) {
    const double coefs[3] = { 0.42, 0.5, 0.08 };
    double win = 0.0, sign = 1.0;
    for (int i = 0; i < 3; i++) {
        win += sign * coefs[i] * cos((double)i * 2.0 * 3.14159265358979323846 * n / N);
        sign = -sign;
    }
    return win;
}

// run(), the part executed once the sample buffer is full (:373-397)
void rowOf(const cf* x, int N, float* rowDb) {
    orc_fft64::Plan plan;
    plan.init(N, -1);
    std::vector<orc_fft64::cd> in(N), spec(N);
    for (int j = 0; j < N; j++) {
        const float w = (float)blackmanAt(j, N);                         // generateWindow :349-357 (stored as float)
        in[j] = orc_fft64::cd((double)(x[j].re * w), (double)(x[j].im * w));
    }
    plan.run(in.data(), spec.data());
    std::vector<cf> y(N);
    for (int j = 0; j < N; j++) { y[j] = cf{ (float)spec[j].real(), (float)spec[j].imag() }; }
    const int half = N / 2;                                              // swapfft (utils/arrays.cpp:1128-1134): an odd last bin stays
    for (int j = 0; j < half; j++) { std::swap(y[j], y[j + half]); }
    for (int j = 0; j < N; j++) {
        const float mag = sqrtf(y[j].re * y[j].re + y[j].im * y[j].im);  // npabsolute -> generic volk_32fc_magnitude_32f
        rowDb[j] = 20.0f * log10f(mag + 1e-10f);
    }
}

// normalizeMagnitudes (:16-50): subtract the mean of the 301 bins around each bin (fewer at the ends)
void subtractLocalMean(const float* row, int n, float* out) {
    for (int i = 0; i < n; i++) {
        const int lo = std::max(0, i - 150), hi = std::min(n - 1, i + 150);
        float sum = 0.0f;
        int cnt = 0;
        for (int j = lo; j <= hi; j++) { sum += row[j]; cnt++; }
        const float mean = (cnt > 0) ? sum / cnt : 0.0f;
        out[i] = row[i] - mean;
    }
}

inline float positivePart(float v) { return (0.0f < v) ? v : 0.0f; }      // std::max<float>(0.0f, v)

// findDominantHarmonicIntervals (:77-148) as called with lags 8..35: per bin, the lag whose product of positive parts
// is the largest (first of equals; bins too close to the end count as 0).  false = the reference's "invalid range" exit.
bool strongestLag(const float* s, int n, int dMin, int dMax, int* lag) {
    if (n == 0 || dMin <= 0 || dMax < dMin || dMax >= n) { return false; }
    for (int i = 0; i < n; i++) {
        float best = -std::numeric_limits<float>::infinity();
        int bestK = 0;
        for (int d = dMin; d <= dMax; d++) {
            const float v = (i < n - d) ? positivePart(s[i]) * positivePart(s[i + d]) : 0.0f;
            if (v > best) { best = v; bestK = d - dMin; }
        }
        lag[i] = dMin + bestK;
    }
    return true;
}

// getLineCandidates (:151-267)
void lineCandidates(const float* rowDb, int n, float* out) {
    for (int i = 0; i < n; i++) { out[i] = -20.0f; }
    if (n == 0) { return; }
    std::vector<float> norm(n);
    subtractLocalMean(rowDb, n, norm.data());
    std::vector<int> lag(n);
    if (!strongestLag(norm.data(), n, 8, 35, lag.data()) || n < 100) { return; }
    for (int p = 0; p + 100 <= n; p += 50) {
        const float* sub = norm.data() + p;
        int sorted[100];
        for (int k = 0; k < 100; k++) { sorted[k] = lag[p + k]; }
        std::sort(sorted, sorted + 100);
        const int period = sorted[50];                                    // nth_element at n / 2 (:62-74)
        if (period <= 0) { continue; }
        std::vector<float> score(period, 0.0f);
        for (int x = 0; x < 100; x++) { score[(x + period - 1) % period] += sub[x]; }
        int phase = 0;
        for (int k = 1; k < period; k++) { if (score[k] > score[phase]) { phase = k; } }       // max_element: first maximum
        int nIn = 0, maxi = 0;
        for (int ix = phase; ix < 100; ix += period, nIn++) {
            if (sub[ix] > sub[phase + maxi * period]) { maxi = nIn; }
        }
        for (int q = maxi - 4; q < maxi; q++) {
            if (q < 0 || q >= nIn) { continue; }
            const int ix = p + phase + q * period;
            out[ix] = -norm[ix];
        }
    }
}

struct SigDet {
    double sampleRate = 0.0, centerFrequency = 0.0;
    int N = 0, pos = 0, sinceDetect = 0, rows = 0;
    std::vector<cf> buf;
    std::deque<std::vector<float>> history;       // suppressedCarrierCandidates, at most 100 rows
    std::vector<float> smoothed;                  // sigs_smoothed

    void setSampleRate(double sr) {               // :297-307 + updateFFTSize :328-347
        if (sampleRate == sr) { return; }
        sampleRate = sr;
        if (sr > 0) {
            N = (int)(sr * (1 / 10.0));
            buf.assign((size_t)N, cf{ 0.0f, 0.0f });
        }
        pos = 0; sinceDetect = 0;
    }

    void aggregate() {                            // aggregateAndDetect :416-452
        if (history.empty()) { return; }
        const size_t bins = history[0].size();
        if (bins == 0) { return; }
        std::vector<float> mean(bins, 0.0f);
        for (size_t f = 0; f < bins; f++) {
            float sum = 0.0f;
            int cnt = 0;
            for (auto& r : history) { if (f < r.size()) { sum += r[f]; cnt++; } }
            mean[f] = (cnt > 0) ? sum / cnt : 0.0f;
        }
        smoothed.resize(bins);
        for (size_t i = 0; i < bins; i++) {
            if (i == 0) { smoothed[i] = (mean[i] + mean[i + 1]) / 2.0f; }
            else if (i == bins - 1) { smoothed[i] = (mean[i - 1] + mean[i]) / 2.0f; }
            else { smoothed[i] = (mean[i - 1] + mean[i] + mean[i + 1]) / 3.0f; }
        }
    }

    void addRow(const float* rowDb, int n) {      // addSingleFFTRow :454-481
        std::vector<float> c(n);
        lineCandidates(rowDb, n, c.data());
        history.push_back(std::move(c));
        while (history.size() > 100) { history.pop_front(); }
        sinceDetect++;
        rows++;
        if (history.size() > 20 && sinceDetect >= 10) { aggregate(); sinceDetect = 0; }
    }

    void process(int count, const cf* in) {       // run() :359-414 (the block itself passes through unchanged)
        if (N <= 0) { return; }
        std::vector<float> row((size_t)N);
        for (int i = 0; i < count; i++) {
            if (pos < N) { buf[pos++] = in[i]; }
            if (pos >= N) {
                rowOf(buf.data(), N, row.data());
                addRow(row.data(), N);
                pos = 0;
            }
        }
    }
};

} // namespace

extern "C" {

void* orc_sigdet_create(double sampleRate) { auto* d = new SigDet; d->setSampleRate(sampleRate); return d; }
void orc_sigdet_destroy(void* h) { delete (SigDet*)h; }
int orc_sigdet_fft_size(void* h) { return ((SigDet*)h)->N; }
void orc_sigdet_set_center_frequency(void* hv, double f) {
    auto* h = (SigDet*)hv;
    if (h->centerFrequency == f) { return; }
    h->centerFrequency = f;
    h->pos = 0; h->sinceDetect = 0;               // clear() :323-326
}
int orc_sigdet_process(void* h, int count, const float* in) { ((SigDet*)h)->process(count, (const cf*)in); return ((SigDet*)h)->rows; }
int orc_sigdet_result(void* hv, float* out, int cap) {
    auto& r = ((SigDet*)hv)->smoothed;
    const int n = (int)r.size();
    if (out) { memcpy(out, r.data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}
int orc_sigdet_history(void* hv, int row, float* out, int cap) {
    auto& hst = ((SigDet*)hv)->history;
    if (out && row >= 0 && row < (int)hst.size()) { memcpy(out, hst[row].data(), std::min((size_t)cap, hst[row].size()) * sizeof(float)); }
    return (int)hst.size();
}
void orc_sigdet_row(const float* iq, int N, float* rowDb) { rowOf((const cf*)iq, N, rowDb); }
void orc_sigdet_candidates(const float* rowDb, int N, float* out) { lineCandidates(rowDb, N, out); }
void orc_sigdet_add_row(void* h, const float* rowDb, int N) { ((SigDet*)h)->addRow(rowDb, N); }

}
