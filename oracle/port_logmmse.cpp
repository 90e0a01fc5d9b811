// TEST INFRASTRUCTURE ONLY (oracle port).  Part of oracle/libport_oracle.so; never linked into, or loaded by, the product.
//
// Self-contained restatement (no reference includes) of the logmmse module's front-end pre-processor:
//   dsp::IFNRLogMMSE::process            misc_modules/noise_reduction_logmmse/src/if_nr.h:63-113
//   LogMMSE::logmmse_sample / _all       .../logmmse.h:244-296, 298-405  (wide-band branch: IFNRLogMMSE forces it, if_nr.h:19-21)
//   SavedParamsC::add_noise_history      .../logmmse.h:117-143
//   SavedParamsC::update_noise_mu2       .../logmmse.h:150-242 (wide-band branch :200-240)
//   BackgroundNoiseCaltulator::addFrame  .../bgnoise.h:27-59
//   array helpers it calls               core/src/utils/arrays.cpp (nphanning :729, npsum :738, npmavg :1068, expn :60-431,
//                                        linearInterpolateHoles :433-470, npall :842, npminimum_ :911, npmaximum_ :975)
// Pinned against the reference's own sources compiled into oracle/_ref (ref_logmmse.cpp) by tests/test_oracle_logmmse.py.
// The DFT is oracle/fft64.h on both sides (FFTW is un-vendored), `npexp` is VOLK's generic fast exp as stated in
// oracle/shim/volk/volk.h; all other arithmetic is sequential fp32 in reference order (-ffp-contract=off).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <vector>

#include "fft64.h"
#include "oracle_api.h"

namespace {

struct cf { float re, im; };

// ------------------------------------------------------------------ exponential integral E1 (scipy.special.expn(1, x))
// The reference tabulates E1 at x = (i / 1000)^4, i = 0 .. 3000 (entry 0 at 1e-16), as floats, and interpolates linearly
// in x^(1/4) (arrays.cpp:60-431).  The table is rebuilt here from the function: power series below 1, continued fraction
// (modified Lentz) above, both in float64 to ~1e-15, then narrowed -- tests check the result against the reference's table.
double expInt1(double x) {
    if (x <= 1.0) {
        double sum = 0.0, term = 1.0;
        for (int k = 1; k < 200; k++) {
            term *= -x / (double)k;
            const double add = -term / (double)k;
            sum += add;
            if (fabs(add) < 1e-18 * fabs(sum)) { break; }
        }
        return -0.57721566490153286061 - log(x) + sum;
    }
    const double tiny = 1e-300;
    double b = x + 1.0, c = 1.0 / tiny, d = 1.0 / b, h = d;
    for (int i = 1; i < 2000; i++) {
        const double an = -(double)i * (double)i;
        b += 2.0;
        d = 1.0 / (an * d + b);
        c = b + an / c;
        const double del = c * d;
        h *= del;
        if (fabs(del - 1.0) < 1e-16) { break; }
    }
    return h * exp(-x);
}

const std::vector<float>& e1Table() {
    static std::vector<float> t;
    if (t.empty()) {
        t.resize(3001);
        t[0] = (float)expInt1(1e-16);
        for (int i = 1; i <= 3000; i++) {
            const double r = (double)i / 1000.0;
            t[i] = (float)expInt1(r * r * r * r);
        }
    }
    return t;
}

float expnLookup(float x) {                                   // EXPN::expn__ (arrays.cpp:61-427)
    const std::vector<float>& t = e1Table();
    const float index = sqrtf(sqrtf(x)) * 1000;
    const float iindex = floorf(index);
    const float f = index - iindex;
    const size_t i = (size_t)iindex;
    return t[i] + (t[i + 1] - t[i]) * f;
}

void fastExp(float* out, const float* in, int n) {            // volk_32f_expfast_32f, generic (oracle/shim/volk/volk.h)
    const float fact = 8388608.0f / 0.6931471805f, bias = 1065353216.0f - 60801.0f;
    for (int i = 0; i < n; i++) {
        const int bits = (int)(fact * in[i] + bias);
        memcpy(&out[i], &bits, sizeof(float));
    }
}

void movingAverage(std::vector<float>& v, int windowSize) {   // npmavg (arrays.cpp:1068-1092), "only even window sizes"
    std::vector<float> out;
    out.reserve(v.size());
    float sum = 0, count = 0;
    const int ws2 = windowSize / 2;
    for (int ix = 0; ix < (int)v.size() + ws2; ix++) {
        if (ix < (int)v.size()) { sum += v[ix]; count++; }
        if (ix > windowSize) { count--; sum -= v[(size_t)(ix - count)]; }
        if (ix >= ws2) { out.push_back(sum / count); }
    }
    v.swap(out);
}

bool fillHoles(float* arr, int n) {                           // linearInterpolateHoles (arrays.cpp:433-470)
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

// BackgroundNoiseCaltulator (bgnoise.h): the mode of the log10 histogram of the kept bins, every 10th frame, smoothed
struct BackgroundNoise {
    static constexpr float kErased = 1e9f;
    double lastNoise = kErased;
    int frameCount = 0;
    void reset() { lastNoise = kErased; frameCount = 0; }
    float addFrame(const std::vector<float>& frame) {
        if (frameCount > 0 && frameCount % 10 != 0) { frameCount++; return (float)lastNoise; }
        frameCount++;
        float minn = kErased, maxx = -kErased;
        std::vector<float> lg;
        lg.reserve(frame.size());
        for (float q : frame) {
            if (q != kErased) {
                q = log10f(q);
                minn = std::min<float>(minn, q);
                maxx = std::max<float>(maxx, q);
                lg.push_back(q);
            }
        }
        const float width = maxx - minn;
        std::vector<int> buckets(1000, 0);
        for (float f : lg) {
            const int b = (int)(1000 * ((f - minn) / width));
            if (b >= 0 && b < 1000) { buckets[b]++; }          // (the largest value lands one past the end in the reference)
        }
        const int ix = (int)(std::max_element(buckets.begin(), buckets.end()) - buckets.begin());
        const double maxf = pow(10, ((((double)ix) / 1000) * width + minn));
        if (lastNoise == kErased) { lastNoise = maxf; }
        else { lastNoise = 0.9 * lastNoise + 0.1 * maxf; }
        return (float)lastNoise;
    }
};

struct LogMmse {
    int rate = 0;
    bool hold = false, needReset = true, haveFifo = false, sampled = false;
    std::vector<cf> fifo;                                       // worker1c
    // SavedParamsC
    int Slen = 0, len1 = 0, len2 = 0, nFFT = 0;
    std::vector<float> win, sums, devs, noiseMu2, XkPrev;
    std::deque<std::vector<float>> noiseHist, devHist;
    std::vector<cf> xOld;
    float aa = 0.98f, ksiMin = 0;
    BackgroundNoise bg;
    orc_fft64::Plan fwd, bwd;

    size_t historyLen() const { return nFFT < 1000 ? 2000 : 200; }

    void transform(const orc_fft64::Plan& p, const std::vector<cf>& in, std::vector<cf>& out) const {   // fftwPlanImplFFTW::npfftfft (arrays.cpp:533-546)
        std::vector<orc_fft64::cd> a((size_t)nFFT, orc_fft64::cd(0, 0)), b((size_t)nFFT);
        for (size_t i = 0; i < in.size() && i < (size_t)nFFT; i++) { a[i] = orc_fft64::cd(in[i].re, in[i].im); }
        p.run(a.data(), b.data());
        out.resize((size_t)nFFT);
        for (int i = 0; i < nFFT; i++) { out[i] = cf{ (float)b[i].real(), (float)b[i].imag() }; }
        if (p.sign > 0) {                                       // div_ by the length: complex multiply by (1 / N, 0)
            const float s = 1.0f / (float)nFFT;
            for (auto& v : out) { v = cf{ v.re * s - v.im * 0.0f, v.re * 0.0f + v.im * s }; }
        }
    }

    void windowed(const cf* x, std::vector<cf>& out) const {    // muleach(win, x[k : k + Slen]): complex_t{w, 0} * x
        out.resize((size_t)Slen);
        for (int i = 0; i < Slen; i++) { out[i] = cf{ (win[i] * x[i].re) - (0.0f * x[i].im), (0.0f * x[i].re) + (win[i] * x[i].im) }; }
    }

    static void magnitude(const std::vector<cf>& s, std::vector<float>& m) {
        m.resize(s.size());
        for (size_t i = 0; i < s.size(); i++) { m[i] = sqrtf(s[i].re * s[i].re + s[i].im * s[i].im); }
    }

    void addNoiseHistory(const std::vector<float>& noise) {     // logmmse.h:117-143
        if (hold) { return; }
        noiseHist.push_back(noise);
        for (int i = 0; i < nFFT; i++) { sums[i] = sums[i] + noise[i]; }
        while (noiseHist.size() > historyLen()) {
            for (int i = 0; i < nFFT; i++) { sums[i] = sums[i] - noiseHist.front()[i]; }
            noiseHist.pop_front();
        }
        const float inv = (float)(1.0 / (float)noiseHist.size());
        std::vector<float> diff((size_t)nFFT);
        for (int i = 0; i < nFFT; i++) { const float d = noise[i] - sums[i] * inv; diff[i] = d * d; }
        for (int i = 0; i < nFFT; i++) { devs[i] = devs[i] + diff[i]; }
        devHist.push_back(std::move(diff));
        while (devHist.size() > historyLen()) {
            for (int i = 0; i < nFFT; i++) { devs[i] = devs[i] - devHist.front()[i]; }
            devHist.pop_front();
        }
    }

    void updateNoise() {                                        // update_noise_mu2, wide-band branch (logmmse.h:200-240)
        const size_t nframes = noiseHist.size();
        if (!(nframes > 100 && !hold)) { return; }
        const std::vector<float> keep = noiseMu2;
        const float inv = 1 / (float)nframes;
        std::vector<float> avg((size_t)nFFT), dv((size_t)nFFT);
        for (int i = 0; i < nFFT; i++) { avg[i] = sums[i] * inv; }
        for (int i = 0; i < nFFT; i++) { const float hi = devs[i] * inv; dv[i] = hi * hi; }
        for (int z = 0; z < nFFT; z++) { if (abs(z - nFFT / 2) < nFFT * 15 / 100) { dv[z] = BackgroundNoise::kErased; } }
        std::fill(noiseMu2.begin(), noiseMu2.end(), 0.0f);
        const float limit = bg.addFrame(dv);
        for (int q = 0; q < nFFT; q++) { if (dv[q] < limit) { noiseMu2[q] = avg[q] * avg[q]; } }
        if (!fillHoles(noiseMu2.data(), nFFT)) { noiseMu2 = keep; }
    }

    void sample(int noiseFrames) {                              // logmmse_sample (logmmse.h:244-296)
        Slen = (int)floor(0.02 * rate);
        if (Slen % 2 == 1) { Slen++; }
        len1 = (int)floor(Slen * 50 / 100);
        noiseHist.clear(); devHist.clear();
        len2 = Slen - len1;
        win.resize((size_t)Slen);
        for (int i = 0; i < Slen; i++) { win[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * i / (Slen - 1))); }   // nphanning
        float total = 0;
        std::vector<float> scaled((size_t)Slen);
        for (int i = 0; i < Slen; i++) { scaled[i] = win[i] * (float)len2; }
        for (int i = 0; i < Slen; i++) { total += win[i]; }
        const float invTotal = (float)(1.0 / total);
        for (int i = 0; i < Slen; i++) { win[i] = scaled[i] * invTotal; }
        nFFT = 2 * Slen;
        fwd.init(nFFT, -1); bwd.init(nFFT, +1);
        sums.assign((size_t)nFFT, 0.0f); devs.assign((size_t)nFFT, 0.0f);
        std::vector<float> mean((size_t)nFFT, 0.0f), noise;
        std::vector<cf> seg, spec;
        for (int j = 0; j < Slen * noiseFrames; j += Slen) {
            windowed(fifo.data() + j, seg);
            transform(fwd, seg, spec);
            magnitude(spec, noise);
            addNoiseHistory(noise);
            for (int i = 0; i < nFFT; i++) { mean[i] = mean[i] + noise[i]; }
        }
        const float invFrames = (float)(1.0 / (float)noiseFrames);
        noiseMu2.resize((size_t)nFFT);
        for (int i = 0; i < nFFT; i++) { noiseMu2[i] = mean[i] * invFrames; }
        movingAverage(noiseMu2, 120);
        for (int i = 0; i < nFFT; i++) { noiseMu2[i] = noiseMu2[i] * noiseMu2[i]; }
        XkPrev.assign((size_t)len1, 0.0f);
        xOld.assign((size_t)len1, cf{ 0.0f, 0.0f });
        ksiMin = (float)::pow(10, -25.0 / 10.0);
        sampled = true;
    }

    void all(std::vector<cf>& result) {                         // logmmse_all (logmmse.h:298-405)
        const int nFrames = (int)(floor((double)(fifo.size() / (size_t)len2)) - floor((double)(Slen / len2)));
        updateNoise();
        result.assign((size_t)std::max(0, nFrames * len2), cf{ 0.0f, 0.0f });
        std::vector<cf> seg, spec, shaped, back;
        std::vector<float> sig, tmp((size_t)nFFT), gain((size_t)nFFT), ksi((size_t)nFFT), gammak((size_t)nFFT), A((size_t)nFFT);
        for (int k = 0; k < nFrames * len2; k += len2) {
            windowed(fifo.data() + k, seg);
            transform(fwd, seg, spec);
            magnitude(spec, sig);
            for (int z = 1; z < nFFT; z++) { if (sig[z] == 0) { sig[z] = sig[z - 1]; } }
            addNoiseHistory(sig);
            for (int i = 0; i < nFFT; i++) {
                float g = (sig[i] * sig[i]) / noiseMu2[i];
                if (g > 40) { g = 40; }
                gammak[i] = g;
            }
            bool allNonZero = true;                              // npall
            for (float v : XkPrev) { if (v == 0) { allNonZero = false; break; } }
            const float oneMinus = 1 - aa;
            for (int i = 0; i < nFFT; i++) {
                float m = gammak[i] + -1.0f;
                if (m < 0) { m = 0; }
                if (!allNonZero) { ksi[i] = m * oneMinus + aa; }
                else {
                    float v = (XkPrev[i] * aa) / noiseMu2[i] + m * oneMinus;
                    if (v < ksiMin) { v = ksiMin; }
                    ksi[i] = v;
                }
            }
            for (int i = 0; i < nFFT; i++) { A[i] = ksi[i] / (ksi[i] + 1.0f); }
            for (int i = 0; i < nFFT; i++) { tmp[i] = expnLookup(A[i] * gammak[i]) * 0.5f; }
            fastExp(gain.data(), tmp.data(), nFFT);
            for (int i = 0; i < nFFT; i++) { gain[i] = A[i] * gain[i]; }
            XkPrev.resize((size_t)nFFT);
            shaped.resize((size_t)nFFT);
            for (int i = 0; i < nFFT; i++) {
                const float s = sig[i] * gain[i];
                XkPrev[i] = s * s;
                shaped[i] = cf{ (gain[i] * spec[i].re) - (0.0f * spec[i].im), (0.0f * spec[i].re) + (gain[i] * spec[i].im) };
            }
            transform(bwd, shaped, back);
            for (int i = 0; i < len1; i++) { result[(size_t)k + i] = cf{ xOld[i].re + back[i].re, xOld[i].im + back[i].im }; }
            for (int i = 0; i < len1; i++) { xOld[i] = back[(size_t)len1 + i]; }
        }
    }

    int process(const cf* in, int count, cf* out) {             // IFNRLogMMSE::process (if_nr.h:63-113)
        if (needReset) { needReset = false; haveFifo = false; }
        if (!haveFifo) {
            fifo.clear(); haveFifo = true;
            noiseHist.clear(); devHist.clear(); XkPrev.clear(); noiseMu2.clear(); xOld.clear(); sampled = false; bg.reset();   // params.reset()
        }
        fifo.insert(fifo.end(), in, in + count);
        const int noiseFrames = 12, fram = rate / 100;
        int demand = fram * 2;
        if (!sampled) { demand = fram * (noiseFrames + 2) * 2; }
        if ((int)fifo.size() < demand) { return 0; }
        if (!sampled) { sample(noiseFrames); }
        std::vector<cf> rv;
        all(rv);
        for (size_t i = 0; i < rv.size(); i++) { out[i] = cf{ rv[i].re * (float)4.0, rv[i].im * (float)4.0 }; }
        fifo.erase(fifo.begin(), fifo.begin() + (long)rv.size());
        return (int)rv.size();
    }
};

} // namespace

extern "C" {

void* orc_logmmse_create(int sampleRate) { auto* p = new LogMmse; p->rate = sampleRate; return p; }
void orc_logmmse_destroy(void* h) { delete (LogMmse*)h; }
void orc_logmmse_reset(void* h) { ((LogMmse*)h)->needReset = true; }
void orc_logmmse_set_hold(void* h, int hold) { ((LogMmse*)h)->hold = hold != 0; }
int orc_logmmse_process(void* h, int count, const float* in, float* out) { return ((LogMmse*)h)->process((const cf*)in, count, (cf*)out); }
int orc_logmmse_noise(void* hv, float* out, int cap) {
    auto& v = ((LogMmse*)hv)->noiseMu2;
    const int n = (int)v.size();
    if (out) { memcpy(out, v.data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}
float orc_expn(float x) { return expnLookup(x); }

}
