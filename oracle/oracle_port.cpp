// TEST INFRASTRUCTURE ONLY (oracle port).  Never linked into, or loaded by, the product:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may use it, and only as the checker / the reported CPU baseline.
//
// Self-contained CPU restatement (no reference includes) of the SDR++Brown
// IQ-front-end hot path, function by function, citing the reference lines each
// one follows.  It exists because /root/reference (and so oracle/_ref) is not
// present on the GPU box.  Parity pin: tests/test_oracle_cpu.py checks this file
// bit-for-bit against oracle/_ref/libref_oracle.so (the reference's own headers
// compiled in place) whenever that library is available, and against the
// committed vectors in tests/golden/ everywhere.  Caveat stated once: the
// reference's arithmetic lives in VOLK / FFTW3f, which are un-vendored and
// un-pinned (core/CMakeLists.txt:160-171) and the reference has no golden
// vectors for this path, so "reference" here means the reference's blocks on
// VOLK's generic kernels (oracle/shim/volk/volk.h).
//
// All arithmetic is sequential IEEE fp32 in reference order; build with
// -ffp-contract=off (oracle/Makefile).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <vector>

#include "frontend_restated.h"
#include "oracle_api.h"

namespace {

struct cf { float re, im; };
const double PI_D = 3.14159265358979323846;   // DB_M_PI, dsp/math/constants.h:3
const float PI_F = 3.1415926535f;            // FL_M_PI, dsp/math/constants.h:4

// ------------------------------------------------------------------ decimation plans
// pass 1: the coefficient tables
#define DECIM_TABLE_U32(name, count) static const uint32_t name[count] = {
#define DECIM_ROW(...) __VA_ARGS__,
#define DECIM_TABLE_END };
#define DECIM_PLAN_BEGIN(ratio, nstages)
#define DECIM_STAGE(index, decim, name, ntaps)
#define DECIM_PLAN_END
#include "decim_tables.inc"
#undef DECIM_TABLE_U32
#undef DECIM_ROW
#undef DECIM_TABLE_END
#undef DECIM_PLAN_BEGIN
#undef DECIM_STAGE
#undef DECIM_PLAN_END

struct StageDef { int decim; const uint32_t* taps; int ntaps; };
struct PlanDef { int ratio; int nstages; StageDef stages[4]; };
// pass 2: the plans
static const PlanDef kPlans[] = {
#define DECIM_TABLE_U32(name, count)
#define DECIM_ROW(...)
#define DECIM_TABLE_END
#define DECIM_PLAN_BEGIN(ratio, nstages) { ratio, nstages, {
#define DECIM_STAGE(index, decim, name, ntaps) { decim, name, ntaps },
#define DECIM_PLAN_END } },
#include "decim_tables.inc"
#undef DECIM_TABLE_U32
#undef DECIM_ROW
#undef DECIM_TABLE_END
#undef DECIM_PLAN_BEGIN
#undef DECIM_STAGE
#undef DECIM_PLAN_END
};
const int kNumPlans = sizeof(kPlans) / sizeof(kPlans[0]);   // 13 -> max ratio 8192 (power_decimator.h:29-31)

const PlanDef* findPlan(int ratio) {
    for (int i = 0; i < kNumPlans; i++) { if (kPlans[i].ratio == ratio) { return &kPlans[i]; } }
    return nullptr;
}

// ------------------------------------------------------------------ windows & taps
// window::cosine / nuttall / blackman — dsp/window/cosine.h:7-16, nuttall.h:5-8, blackman.h:5-8
double cosineWindow(double n, double N, const double* coefs, int coefCount) {
    double win = 0.0, sign = 1.0;
    for (int i = 0; i < coefCount; i++) {
        win += sign * coefs[i] * cos((double)i * 2.0 * PI_D * n / N);
        sign = -sign;
    }
    return win;
}
double nuttall(double n, double N) {
    const double c[] = { 0.355768, 0.487396, 0.144232, 0.012604 };
    return cosineWindow(n, N, c, 4);
}
double blackman(double n, double N) {
    const double c[] = { 0.42, 0.5, 0.08 };
    return cosineWindow(n, N, c, 3);
}
double sincD(double x) { return (x == 0.0) ? 1.0 : (sin(x) / x); }                       // math/sinc.h:5-7
double hzToRads(double f, double sr) { return 2.0 * PI_D * (f / sr); }                   // math/hz_to_rads.h:6-8
int estimateTapCount(double transWidth, double sr) { return (int)(3.8 * sr / transWidth); } // taps/estimate_tap_count.h:4-6

// taps::windowedSinc<float> — taps/windowed_sinc.h:9-33 (window evaluated at t - half, count)
template <typename W>
std::vector<float> windowedSinc(int count, double omega, W window) {
    std::vector<float> taps(count);
    double half = (double)count / 2.0;
    double corr = 1.0 * omega / PI_D;
    for (int i = 0; i < count; i++) {
        double t = (double)i - half + 0.5;
        taps[i] = (float)(sincD(t * omega) * window(t - half, (double)count) * corr);
    }
    return taps;
}
// taps::lowPass — taps/low_pass.h:8-15
std::vector<float> lowPass(double cutoff, double transWidth, double sr) {
    int count = estimateTapCount(transWidth, sr);
    return windowedSinc(count, hzToRads(cutoff, sr), [](double n, double N) { return nuttall(n, N); });
}
// taps::highPass — taps/high_pass.h:8-15
std::vector<float> highPass(double cutoff, double transWidth, double sr) {
    int count = estimateTapCount(transWidth, sr);
    return windowedSinc(count, hzToRads((sr / 2.0) - cutoff, sr), [](double n, double N) {
        return nuttall(n, N) * (((int)round(n) % 2) ? -1.0f : 1.0f);
    });
}
// taps::bandPass<float> — taps/band_pass.h:10-24
std::vector<float> bandPass(double bandStart, double bandStop, double transWidth, double sr) {
    float offsetOmega = (float)hzToRads((bandStart + bandStop) / 2.0, sr);
    int count = estimateTapCount(transWidth, sr);
    return windowedSinc(count, hzToRads((bandStop - bandStart) / 2.0, sr), [=](double n, double N) {
        return (double)(2.0f * cosf(offsetOmega * (float)n)) * nuttall(n, N);
    });
}

// ------------------------------------------------------------------ VOLK generic kernels (see oracle/shim/volk/volk.h)
inline cf dotCF(const cf* x, const float* taps, int n) {          // volk_32fc_32f_dot_prod_32fc
    float sr = 0.0f, si = 0.0f;
    for (int i = 0; i < n; i++) { sr += x[i].re * taps[i]; si += x[i].im * taps[i]; }
    return cf{ sr, si };
}
inline float dotFF(const float* x, const float* taps, int n) {    // volk_32f_x2_dot_prod_32f
    float s = 0.0f;
    for (int i = 0; i < n; i++) { s += x[i] * taps[i]; }
    return s;
}
void powerSpectrum(float* out, const float* X, float norm, int n) {   // volk_32fc_s32f_power_spectrum_32f
    const float inv = 1.0f / norm;
    for (int i = 0; i < n; i++) {
        const float r = X[2 * i] * inv, q = X[2 * i + 1] * inv;
        float l = log2f(r * r + q * q);
        if (std::isinf(l)) { l = copysignf(127.0f, l); }
        out[i] = 3.01029995663981209120f * l;
    }
}

// ------------------------------------------------------------------ filter::FIR / DecimatingFIR
// filter/fir.h:19-92 and filter/decimating_fir.h:45-68.  History = last taps-1 inputs,
// zero at start; correlation form (no tap reversal).
template <typename D>
struct Fir {
    std::vector<float> taps;
    std::vector<D> buf;      // [history (taps-1) | block]
    int decim = 1;
    int offset = 0;          // DecimatingFIR::offset
    void init(const std::vector<float>& t, int d = 1) {
        taps = t; decim = d; offset = 0;
        buf.assign(taps.size() + 4096, D{});   // grows on demand (the reference allocates STREAM_BUFFER_SIZE + 64000 up front, fir.h:25)
    }
    void ensure(int count) { size_t need = taps.size() + (size_t)count; if (buf.size() < need) { buf.resize(need, D{}); } }
    void reset() { std::fill(buf.begin(), buf.begin() + (taps.size() - 1), D{}); offset = 0; }
    // FIR::setTaps — fir.h:33-54: keep the most recent history aligned to the new length
    void setTaps(const std::vector<float>& t) {
        int oldTC = (int)taps.size();
        taps = t;
        int newTC = (int)taps.size();
        ensure(oldTC);
        if (newTC < oldTC) {
            memmove(buf.data(), &buf[oldTC - newTC], (newTC - 1) * sizeof(D));
        }
        else if (newTC > oldTC) {
            memmove(&buf[newTC - oldTC], buf.data(), (oldTC - 1) * sizeof(D));
            std::fill(buf.begin(), buf.begin() + (newTC - oldTC), D{});
        }
    }
    static D dot(const cf* x, const float* t, int n) { return dotCF(x, t, n); }
    static float dot(const float* x, const float* t, int n) { return dotFF(x, t, n); }
    // FIR::process — fir.h:64-92
    int processFir(int count, const D* in, D* out) {
        const int nt = (int)taps.size();
        ensure(count);
        memcpy(&buf[nt - 1], in, count * sizeof(D));
        for (int i = 0; i < count; i++) { out[i] = dot(&buf[i], taps.data(), nt); }
        memmove(buf.data(), &buf[count], (nt - 1) * sizeof(D));
        return count;
    }
    // DecimatingFIR::process — decimating_fir.h:45-68
    int processDecim(int count, const D* in, D* out) {
        const int nt = (int)taps.size();
        ensure(count);
        memcpy(&buf[nt - 1], in, count * sizeof(D));
        int outCount = 0;
        for (; offset < count; offset += decim) { out[outCount++] = dot(&buf[offset], taps.data(), nt); }
        offset -= count;
        memmove(buf.data(), &buf[count], (nt - 1) * sizeof(D));
        return outCount;
    }
};

// ------------------------------------------------------------------ multirate::PowerDecimator — power_decimator.h:51-67,92-110
struct PowerDecim {
    int ratio = 1;
    std::vector<Fir<cf>> firs;
    void setRatio(int r) {
        ratio = r;
        firs.clear();
        if (ratio > 1) {
            const PlanDef* p = findPlan(ratio);
            firs.resize(p->nstages);
            for (int i = 0; i < p->nstages; i++) {
                std::vector<float> t(p->stages[i].ntaps);
                memcpy(t.data(), p->stages[i].taps, sizeof(float) * t.size());
                firs[i].init(t, p->stages[i].decim);
            }
        }
    }
    void reset() { for (auto& f : firs) { f.reset(); } }
    int process(int count, const cf* in, cf* out) {
        if (ratio == 1) { memcpy(out, in, count * sizeof(cf)); return count; }
        const cf* data = in;
        for (auto& f : firs) { count = f.processDecim(count, data, out); data = out; }
        return count;
    }
};

// ------------------------------------------------------------------ multirate::PolyphaseResampler — polyphase_resampler.h:69-99, polyphase_bank.h:14-47
struct Polyphase {
    int interp = 1, decim = 1, tapsPerPhase = 1;
    std::vector<float> bank;   // [phase][tapsPerPhase]
    std::vector<cf> buf;
    int phase = 0, offset = 0;
    void setRatio(int L, int M, const std::vector<float>& taps) {
        interp = L; decim = M;
        tapsPerPhase = ((int)taps.size() + L - 1) / L;
        bank.assign((size_t)L * tapsPerPhase, 0.0f);
        int tot = L * tapsPerPhase;
        for (int i = 0; i < tot; i++) {
            bank[(size_t)((L - 1) - (i % L)) * tapsPerPhase + (i / L)] = (i < (int)taps.size()) ? taps[i] : 0.0f;
        }
        if (buf.size() < (size_t)tapsPerPhase + 4096) { buf.resize((size_t)tapsPerPhase + 4096, cf{}); }
        reset();
    }
    void reset() { std::fill(buf.begin(), buf.begin() + (tapsPerPhase - 1), cf{}); phase = 0; offset = 0; }
    int process(int count, const cf* in, cf* out) {
        int outCount = 0;
        if (buf.size() < (size_t)tapsPerPhase + count) { buf.resize((size_t)tapsPerPhase + count, cf{}); }
        memcpy(&buf[tapsPerPhase - 1], in, count * sizeof(cf));
        while (offset < count) {
            out[outCount++] = dotCF(&buf[offset], &bank[(size_t)phase * tapsPerPhase], tapsPerPhase);
            phase += decim;
            offset += phase / interp;
            phase = phase % interp;
        }
        offset -= count;
        memmove(buf.data(), &buf[count], (tapsPerPhase - 1) * sizeof(cf));
        return outCount;
    }
};

// ------------------------------------------------------------------ multirate::RationalResampler — rational_resampler.h:90-104,128-173
struct Rational {
    enum Mode { BOTH, DECIM_ONLY, RESAMP_ONLY, NONE };
    Mode mode = NONE;
    PowerDecim decim;
    Polyphase resamp;
    std::vector<float> rtaps;
    double inSR = 1, outSR = 1;
    void init(double in, double out) {
        inSR = in; outSR = out;
        decim.setRatio(2);                                     // dummy init, :32
        rtaps = lowPass(0.25, 0.1, 1.0);                       // dummy init, :31-33
        resamp.setRatio(1, 1, rtaps);
        reconfigure();
    }
    void reconfigure() {
        const int maxRatio = 1 << kNumPlans;
        int predecPower = std::min<int>((int)floor(log2(inSR / outSR)), maxRatio);
        int predecRatio = (predecPower > 0) ? std::min<int>(1 << std::min(predecPower, 30), maxRatio) : 1;   // only used when useDecim
        double intSR = inSR;
        bool useDecim = (inSR > outSR && predecPower > 0);
        if (useDecim) { intSR = inSR / (double)predecRatio; decim.setRatio(predecRatio); }
        int IntSR = (int)round(intSR);
        int OutSR = (int)round(outSR);
        int g = std::gcd(IntSR, OutSR);
        int interp = OutSR / g;
        int dec = IntSR / g;
        if (interp == dec) { mode = useDecim ? DECIM_ONLY : NONE; return; }
        double tapSamplerate = intSR * (double)interp;
        double tapBandwidth = std::min<double>(inSR, outSR) / 2.0;
        double tapTransWidth = tapBandwidth * 0.1;
        rtaps = lowPass(tapBandwidth, tapTransWidth, tapSamplerate);
        for (auto& t : rtaps) { t *= (float)interp; }
        resamp.setRatio(interp, dec, rtaps);
        mode = useDecim ? BOTH : RESAMP_ONLY;
    }
    void reset() { decim.reset(); resamp.reset(); }
    int process(int count, const cf* in, cf* out) {
        switch (mode) {
        case BOTH: count = decim.process(count, in, out); return resamp.process(count, out, out);
        case DECIM_ONLY: return decim.process(count, in, out);
        case RESAMP_ONLY: return resamp.process(count, in, out);
        case NONE: memcpy(out, in, count * sizeof(cf)); return count;
        }
        return count;
    }
};

// ------------------------------------------------------------------ channel::FrequencyXlator — frequency_xlator.h:15-56 on the generic VOLK rotator
struct Xlator {
    float pr = 1.0f, pi = 0.0f;   // phase
    float dr = 1.0f, di = 0.0f;   // phaseDelta
    void setOffsetRads(double w) { dr = (float)cos(w); di = (float)sin(w); }
    void setOffset(double offset, double sr) { setOffsetRads(hzToRads(offset, sr)); }
    void reset() { pr = 1.0f; pi = 0.0f; }
    void process(int n, const cf* in, cf* out) {
        const int RELOAD = 512;
        int done = 0;
        for (int b = 0; b < n / RELOAD; b++) {
            for (int j = 0; j < RELOAD; j++, done++) { step(in[done], out[done]); }
            renorm();
        }
        int tail = n % RELOAD;
        for (int j = 0; j < tail; j++, done++) { step(in[done], out[done]); }
        if (tail) { renorm(); }
    }
    inline void step(const cf& x, cf& y) {
        const float xr = x.re, xi = x.im;
        y.re = xr * pr - xi * pi;
        y.im = xr * pi + xi * pr;
        const float nr = pr * dr - pi * di;
        const float ni = pr * di + pi * dr;
        pr = nr; pi = ni;
    }
    inline void renorm() { const float h = hypotf(pr, pi); pr /= h; pi /= h; }
};

// ------------------------------------------------------------------ channel::RxVFO — rx_vfo.h:19-121
struct Vfo {
    Xlator xlator;
    Rational resamp;
    Fir<cf> filter;
    std::vector<float> ftaps;
    bool filterNeeded = false;
    double inSR, outSR, bw, offset;
    void genTaps() { double w = bw / 2.0; ftaps = lowPass(w, w * 0.1, outSR); }
    void init(double in, double out, double bandwidth, double off) {
        inSR = in; outSR = out; bw = bandwidth; offset = off;
        filterNeeded = (bw != outSR);
        xlator.setOffset(-offset, inSR);
        resamp.init(inSR, outSR);
        genTaps();
        filter.init(ftaps);
    }
    int process(int count, const cf* in, cf* out) {
        xlator.process(count, in, out);
        if (!filterNeeded) { return resamp.process(count, out, out); }
        count = resamp.process(count, out, out);
        filter.processFir(count, out, out);
        return count;
    }
};

// ------------------------------------------------------------------ correction::DCBlocker — dc_blocker.h:54-60
struct DcBlockC {
    float rate; cf off{ 0, 0 };
    void process(int n, const cf* in, cf* out) {
        for (int i = 0; i < n; i++) {
            out[i].re = in[i].re - off.re; out[i].im = in[i].im - off.im;
            off.re += out[i].re * rate; off.im += out[i].im * rate;
        }
    }
};
struct DcBlockF {
    float rate; float off = 0.0f;
    void process(int n, const float* in, float* out) {
        for (int i = 0; i < n; i++) { out[i] = in[i] - off; off += out[i] * rate; }
    }
};

// ------------------------------------------------------------------ loop::AGC — agc.h:13-27,85-139
template <typename T> struct Agc {
    float setPoint, attack, invAttack, decay, invDecay, maxGain, maxOutputAmp, initGain;
    int startEnvelope = 0;
    const int totalEnvelopeLength = 4800;
    float amp = 1.0f;
    void init(double sp, double att, double dec, double mg, double moa, double ig) {
        setPoint = (float)sp; attack = (float)att; invAttack = 1.0f - attack;
        decay = (float)dec; invDecay = 1.0f - decay; maxGain = (float)mg; maxOutputAmp = (float)moa; initGain = (float)ig;
        startEnvelope = 0;
        amp = setPoint / initGain;
    }
    void reset() { amp = setPoint / initGain; startEnvelope = 0; }
    static float ampOf(const cf& v) { return sqrtf((v.re * v.re) + (v.im * v.im)); }   // complex_t::amplitude, types.h:77-79 (float sqrt of float)
    static float ampOf(float v) { return fabsf(v); }
    static cf scale(const cf& v, float g, float e) { cf t{ v.re * g, v.im * g }; return cf{ t.re * e, t.im * e }; }
    static float scale(float v, float g, float e) { return v * g * e; }
    int process(int count, const T* in, T* out) {
        float envelope = 1.0f;
        if (attack <= 0) { std::copy(in, in + count, out); return count; }
        for (int i = 0; i < count; i++) {
            float inAmp = ampOf(in[i]), gain = 1.0f;
            if (inAmp != 0.0f) {
                float namp = (inAmp > amp) ? ((amp * invAttack) + (inAmp * attack)) : ((amp * invDecay) + (inAmp * decay));
                if (!std::isnan(namp)) { amp = namp; gain = std::min<float>(setPoint / amp, maxGain); }
            }
            else { gain = 1.0f; }
            if (startEnvelope < totalEnvelopeLength) { envelope = startEnvelope / (float)totalEnvelopeLength; }
            startEnvelope++;
            out[i] = scale(in[i], gain, envelope);
        }
        return count;
    }
};

// ------------------------------------------------------------------ demod::Quadrature + FM — quadrature.h:39-46, fm.h:86-145
inline float normalizePhase(float d) {            // math/normalize_phase.h
    if (d > PI_F) { d -= 2.0f * PI_F; }
    else if (d <= -PI_F) { d += 2.0f * PI_F; }
    return d;
}
struct FmDemod {
    float invDeviation; float phase = 0.0f;
    bool filtering = false;
    Fir<float> fir;
    std::vector<float> tmp;
    void init(double sr, double bw, bool lp, bool hp) {
        invDeviation = (float)(1.0 / hzToRads(bw / 2.0, sr));
        std::vector<float> t;
        if (lp && hp) { t = bandPass(300.0, bw / 2.0, 100.0, sr); }
        else if (hp) { t = highPass(300.0, 100.0, sr); }
        else if (lp) { t = lowPass(bw / 2.0, (bw / 2.0) * 0.1, sr); }
        else { t = { 1.0f }; }
        filtering = lp || hp;
        fir.init(t);
        tmp.resize(65536);
    }
    int process(int count, const cf* in, float* outStereo) {
        if ((int)tmp.size() < count) { tmp.resize(count); }
        for (int i = 0; i < count; i++) {
            float cphase = atan2f(in[i].im, in[i].re);   // complex_t::phase, types.h:55-57
            tmp[i] = normalizePhase(cphase - phase) * invDeviation;
            phase = cphase;
        }
        if (filtering) { fir.processFir(count, tmp.data(), tmp.data()); }
        for (int i = 0; i < count; i++) { outStereo[2 * i] = tmp[i]; outStereo[2 * i + 1] = tmp[i]; }
        return count;
    }
};

// ------------------------------------------------------------------ demod::BroadcastFM — broadcast_fm.h:34-215
// taps::bandPass<complex_t> (band_pass.h:10-24 over windowed_sinc.h:26-29): the sinc prototype is shifted to the band
// centre by a complex exponential; every product is formed in float, in the order the complex_t operators give it
std::vector<cf> bandPassComplex(double bandStart, double bandStop, double transWidth, double sr, bool oddTapCount) {
    float offsetOmega = (float)hzToRads((bandStart + bandStop) / 2.0, sr);
    int count = estimateTapCount(transWidth, sr);
    if (oddTapCount && !(count % 2)) { count++; }
    const double omega = hzToRads((bandStop - bandStart) / 2.0, sr);
    std::vector<cf> taps(count);
    const double half = (double)count / 2.0;
    const double corr = 1.0 * omega / PI_D;
    for (int i = 0; i < count; i++) {
        const double t = (double)i - half + 0.5;
        const double n = t - half;
        const float x = -offsetOmega * (float)n;
        const float wnd = (float)nuttall(n, (double)count);
        const cf w{ cosf(x) * wnd, sinf(x) * wnd };                          // phasor * (double)nuttall
        const cf c{ (float)sincD(t * omega), 0.0f };
        const cf p{ (c.re * w.re) - (c.im * w.im), (c.im * w.re) + (c.re * w.im) };   // complex_t::operator*(complex_t)
        taps[i] = cf{ p.re * (float)corr, p.im * (float)corr };
    }
    return taps;
}
struct WfmDemod {
    float invDeviation = 0.f, qphase = 0.0f;
    bool stereo = false, lowPassOn = true;
    std::vector<cf> pilotTaps; std::vector<cf> pilotBuf;       // FIR<complex_t, complex_t>: [history | block]
    float alpha = 0.f, beta = 0.f, pllPhase = 0.f, pllFreq = 0.f, minFreq = 0.f, maxFreq = 0.f;
    int delay = 0; std::vector<float> lprBuf; std::vector<cf> lmrBuf;   // math::Delay buffers: [delay samples | block]
    Fir<float> alFir, arFir;
    std::vector<float> mpx, l, r, lmr; std::vector<cf> rtoc, pil, vco, lmrD;
    bool rdsOn = false; Xlator rdsXlator; Rational rdsResamp; int rdsCount = 0;   // broadcast_fm.h:52-53
    // RDS branch (broadcast_fm.h:169-175 stereo, :197-206 mono): the MPX as complex, down by 57 kHz, resampled to 5 kS/s
    void rds(int count, cf* rdsOut) {
        rdsXlator.process(count, rtoc.data(), rtoc.data());
        rdsCount = rdsResamp.process(count, rtoc.data(), rdsOut);
    }
    void init(double deviation, double sr, bool st, bool lp, bool rdsOut = false) {
        stereo = st; lowPassOn = lp; rdsOn = rdsOut;
        rdsXlator.setOffset(-57000.0, sr);
        rdsResamp.init(sr, 5000.0);
        invDeviation = (float)(1.0 / hzToRads(deviation, sr));
        pilotTaps = bandPassComplex(18750.0, 19250.0, 3000.0, sr, true);
        pilotBuf.assign(pilotTaps.size() + 4096, cf{ 0.f, 0.f });
        // PLL::init -> PhaseControlLoop<float>::criticallyDamped (phase_control_loop.h:30-35), evaluated in float
        {
            const float bandwidth = (float)(25000.0 / sr);
            const float damp = sqrt(2.0) / 2.0;
            const float den = (1.0 + 2.0 * damp * bandwidth + bandwidth * bandwidth);
            alpha = (4 * damp * bandwidth) / den;
            beta = (4 * bandwidth * bandwidth) / den;
        }
        pllPhase = 0.0f; pllFreq = (float)hzToRads(19000.0, sr);
        minFreq = (float)hzToRads(18750.0, sr); maxFreq = (float)hzToRads(19250.0, sr);
        delay = (((int)pilotTaps.size() - 1) / 2) + 1;
        lprBuf.assign(delay + 4096, 0.f); lmrBuf.assign(delay + 4096, cf{ 0.f, 0.f });
        std::vector<float> at = lowPass(15000.0, 4000.0, sr);
        alFir.init(at); arFir.init(at);
    }
    template <typename V> static void fit(V& v, size_t n) { if (v.size() < n) { v.resize(n); } }
    int process(int count, const cf* in, float* outStereo, cf* rdsOut = nullptr) {
        rdsCount = 0;
        fit(mpx, count); fit(l, count); fit(r, count); fit(lmr, count); fit(rtoc, count); fit(pil, count); fit(vco, count); fit(lmrD, count);
        for (int i = 0; i < count; i++) {
            const float cphase = atan2f(in[i].im, in[i].re);
            mpx[i] = normalizePhase(cphase - qphase) * invDeviation;
            qphase = cphase;
        }
        if (!stereo) {
            if (rdsOn) {
                for (int i = 0; i < count; i++) { rtoc[i] = cf{ mpx[i], 0.0f }; }
                rds(count, rdsOut);
            }
            if (lowPassOn) { alFir.processFir(count, mpx.data(), mpx.data()); }
            for (int i = 0; i < count; i++) { outStereo[2 * i] = mpx[i]; outStereo[2 * i + 1] = mpx[i]; }
            return count;
        }
        for (int i = 0; i < count; i++) { rtoc[i] = cf{ mpx[i], 0.0f }; }
        // pilot band-pass: complex taps x complex data (volk_32fc_x2_dot_prod_32fc, generic)
        const int nt = (int)pilotTaps.size();
        fit(pilotBuf, (size_t)nt + count);
        memcpy(&pilotBuf[nt - 1], rtoc.data(), count * sizeof(cf));
        for (int i = 0; i < count; i++) {
            float sr = 0.0f, si = 0.0f;
            for (int k = 0; k < nt; k++) {
                const cf a = pilotBuf[i + k], b = pilotTaps[k];
                sr += a.re * b.re - a.im * b.im;
                si += a.re * b.im + a.im * b.re;
            }
            pil[i] = cf{ sr, si };
        }
        memmove(pilotBuf.data(), &pilotBuf[count], (nt - 1) * sizeof(cf));
        // PLL::process (pll.h:64-70) over PhaseControlLoop::advance (phase_control_loop.h:58-66), phase kept in [-pi, pi]
        for (int i = 0; i < count; i++) {
            vco[i] = cf{ cosf(pllPhase), sinf(pllPhase) };
            const float err = normalizePhase(atan2f(pil[i].im, pil[i].re) - pllPhase);
            pllFreq += beta * err;
            if (pllFreq > maxFreq) { pllFreq = maxFreq; } else if (pllFreq < minFreq) { pllFreq = minFreq; }
            pllPhase += pllFreq + (alpha * err);
            while (pllPhase > PI_F) { pllPhase -= (PI_F - (-PI_F)); }
            while (pllPhase < -PI_F) { pllPhase += (PI_F - (-PI_F)); }
        }
        // Delay (math/delay.h:44-52) of the MPX and of its complex copy
        fit(lprBuf, (size_t)delay + count); fit(lmrBuf, (size_t)delay + count);
        memcpy(&lprBuf[delay], mpx.data(), count * sizeof(float));
        memcpy(mpx.data(), lprBuf.data(), count * sizeof(float));
        memmove(lprBuf.data(), &lprBuf[count], delay * sizeof(float));
        memcpy(&lmrBuf[delay], rtoc.data(), count * sizeof(cf));
        memcpy(lmrD.data(), lmrBuf.data(), count * sizeof(cf));
        memmove(lmrBuf.data(), &lmrBuf[count], delay * sizeof(cf));
        for (int i = 0; i < count; i++) {
            const cf c{ vco[i].re, -vco[i].im };
            cf a = lmrD[i];
            a = cf{ a.re * c.re - a.im * c.im, a.re * c.im + a.im * c.re };
            a = cf{ a.re * c.re - a.im * c.im, a.re * c.im + a.im * c.re };
            lmr[i] = a.re * 2.0f;
            l[i] = mpx[i] + lmr[i];
            r[i] = mpx[i] - lmr[i];
        }
        if (rdsOn) { rds(count, rdsOut); }   // rtoc still holds the undelayed MPX
        if (lowPassOn) { alFir.processFir(count, l.data(), l.data()); arFir.processFir(count, r.data(), r.data()); }
        for (int i = 0; i < count; i++) { outStereo[2 * i] = l[i]; outStereo[2 * i + 1] = r[i]; }
        return count;
    }
};

// ------------------------------------------------------------------ demod::SSB — ssb.h:23-36,82-123
struct SsbDemod {
    Xlator xl; Agc<float> agc;
    std::vector<cf> xbuf; std::vector<float> abuf;
    void init(int mode, double bw, double sr, double att, double dec) {
        double tr = (mode == 0) ? bw / 2.0 : (mode == 1) ? -bw / 2.0 : 0.0;
        xl.setOffset(tr, sr);
        agc.init(1.0, att, dec, 10e6, 10.0, INFINITY);
        xbuf.resize(65536); abuf.resize(65536);
    }
    int process(int count, const cf* in, float* outStereo) {
        if ((int)xbuf.size() < count) { xbuf.resize(count); abuf.resize(count); }
        xl.process(count, in, xbuf.data());
        for (int i = 0; i < count; i++) { abuf[i] = xbuf[i].re; }
        agc.process(count, abuf.data(), abuf.data());
        for (int i = 0; i < count; i++) { outStereo[2 * i] = abuf[i]; outStereo[2 * i + 1] = abuf[i]; }
        return count;
    }
};

// ------------------------------------------------------------------ demod::AM — am.h:29-47,101-133
struct AmDemod {
    int agcMode;
    Agc<cf> carrierAgc; Agc<float> audioAgc; DcBlockF dc; Fir<float> lpf;
    std::vector<cf> cbuf; std::vector<float> abuf;
    void init(int mode, double bw, double att, double dec, double dcRate, double sr) {
        agcMode = mode;
        carrierAgc.init(1.0, att, dec, 10e6, 10.0, INFINITY);
        audioAgc.init(1.0, att, dec, 10e6, 10.0, INFINITY);
        dc.rate = (float)dcRate;
        lpf.init(lowPass(bw / 2.0, (bw / 2.0) * 0.1, sr));
        cbuf.resize(65536); abuf.resize(65536);
    }
    int process(int count, const cf* in, float* outStereo) {
        if ((int)cbuf.size() < count) { cbuf.resize(count); abuf.resize(count); }
        if (agcMode == 0) { carrierAgc.process(count, in, cbuf.data()); in = cbuf.data(); }
        for (int i = 0; i < count; i++) { abuf[i] = sqrtf(in[i].re * in[i].re + in[i].im * in[i].im); }
        dc.process(count, abuf.data(), abuf.data());
        if (agcMode == 1) { audioAgc.process(count, abuf.data(), abuf.data()); }
        lpf.processFir(count, abuf.data(), abuf.data());
        for (int i = 0; i < count; i++) { outStereo[2 * i] = abuf[i]; outStereo[2 * i + 1] = abuf[i]; }
        return count;
    }
};

// ------------------------------------------------------------------ radio IF chain: NoiseBlanker then Squelch
// (noise_reduction/noise_blanker.h:38-56: running mean amplitude, samples whose amplitude exceeds level x mean are
// scaled down to the mean; noise_reduction/squelch.h:55-69: the block passes when 10*log10(mean |x|) >= level)
struct IfChain {
    bool nbOn = false, sqOn = false;
    float rate = 0.f, invRate = 1.f, nbLevel = 0.f, amp = 1.0f, sqLevel = 0.f;
    void init(int nb, double nbRate, double nbLvl, int sq, double sqLvl) {
        nbOn = nb != 0; sqOn = sq != 0;
        rate = nbRate; invRate = 1.0f - rate; nbLevel = nbLvl; sqLevel = sqLvl;
    }
    int process(int count, const cf* in, cf* out) {
        for (int i = 0; i < count; i++) { out[i] = in[i]; }
        if (nbOn) {
            for (int i = 0; i < count; i++) {
                const float inAmp = sqrt((out[i].re * out[i].re) + (out[i].im * out[i].im));
                float gain = 1.0f;
                if (inAmp != 0.0f) {
                    amp = (amp * invRate) + (inAmp * rate);
                    const float excess = inAmp / amp;
                    if (excess > nbLevel) { gain = 1.0f / excess; }
                }
                out[i].re = out[i].re * gain; out[i].im = out[i].im * gain;
            }
        }
        if (sqOn && count > 0) {
            float sum = 0.0f;
            for (int i = 0; i < count; i++) { sum += sqrtf(out[i].re * out[i].re + out[i].im * out[i].im); }   // generic VOLK magnitude + accumulator
            sum /= (float)count;
            if (!(10.0f * log10f(sum) >= sqLevel)) { for (int i = 0; i < count; i++) { out[i].re = 0.f; out[i].im = 0.f; } }
        }
        return count;
    }
};

// ------------------------------------------------------------------ noise_reduction::FMIF — fm_if.h:45-80,106-124
// Per sample: window the newest `bins` samples (Nuttall over bins-1, volk_32fc_32f_multiply_32fc), forward DFT, keep the
// strongest bin only (generic volk_32fc_magnitude_32f + volk_32f_index_max_32u: the first of equal maxima), backward DFT,
// take element bins/2.  The DFTs follow oracle/shim/fftw3.h: float64 sums in index order, narrowed to fp32 once.
struct FmIf {
    int bins = 0;
    std::vector<cf> buf; std::vector<float> win; std::vector<double> cs, sn;
    void init(int b) {
        bins = b;
        buf.assign((size_t)bins - 1 + 4096, cf{ 0.f, 0.f });
        win.resize(bins);
        for (int i = 0; i < bins; i++) { win[i] = nuttall(i, bins - 1); }
        cs.resize(bins); sn.resize(bins);
        for (int m = 0; m < bins; m++) {
            const double a = 2.0 * 3.14159265358979323846 * (double)m / (double)bins;
            cs[m] = cos(a); sn[m] = sin(a);
        }
    }
    int process(int count, const cf* in, cf* out) {
        if (buf.size() < (size_t)bins - 1 + count) { buf.resize((size_t)bins - 1 + count); }
        memcpy(&buf[bins - 1], in, (size_t)count * sizeof(cf));
        std::vector<cf> w(bins), X(bins);
        for (int i = 0; i < count; i++) {
            for (int n = 0; n < bins; n++) { w[n] = cf{ buf[i + n].re * win[n], buf[i + n].im * win[n] }; }
            for (int k = 0; k < bins; k++) {
                double sr = 0.0, si = 0.0;
                for (int n = 0; n < bins; n++) {
                    const int m = (int)(((long long)k * n) % bins);
                    const double c = cs[m], s = -sn[m];
                    const double xr = w[n].re, xi = w[n].im;
                    sr += xr * c - xi * s;
                    si += xr * s + xi * c;
                }
                X[k] = cf{ (float)sr, (float)si };
            }
            int idx = 0; float best = sqrtf(X[0].re * X[0].re + X[0].im * X[0].im);
            for (int k = 1; k < bins; k++) {
                const float a = sqrtf(X[k].re * X[k].re + X[k].im * X[k].im);
                if (a > best) { best = a; idx = k; }
            }
            const int m = (int)(((long long)(bins / 2) * idx) % bins);
            const double xr = X[idx].re, xi = X[idx].im;
            out[i] = cf{ (float)(xr * cs[m] - xi * sn[m]), (float)(xr * sn[m] + xi * cs[m]) };
        }
        memmove(buf.data(), &buf[count], (size_t)(bins - 1) * sizeof(cf));
        return count;
    }
};

// ------------------------------------------------------------------ radio AF chain: RationalResampler<stereo_t> + Deemphasis<stereo_t>
// (radio_module.h:101-107; filter/deephasis.h:58-76,89-92).  stereo_t goes through the same complex-by-real dot product.
struct AfPost {
    Rational resamp; bool deOn = false; float alpha = 0.f; cf last{ 0.f, 0.f };
    void init(double inSR, double outSR, double tau) {
        resamp.init(inSR, outSR);
        deOn = tau > 0;
        float dt = 1.0f / outSR;                 // updateAlpha: float dt, then float alpha = dt / (tau + dt) evaluated in double
        alpha = dt / ((tau > 0 ? tau : 50e-6) + dt);
    }
    int process(int count, const cf* in, cf* out) {
        int n = resamp.process(count, in, out);
        if (deOn && n > 0) {
            out[0].re = (alpha * out[0].re) + ((1 - alpha) * last.re);
            out[0].im = (alpha * out[0].im) + ((1 - alpha) * last.im);
            for (int i = 1; i < n; i++) {
                out[i].re = (alpha * out[i].re) + ((1 - alpha) * out[i - 1].re);
                out[i].im = (alpha * out[i].im) + ((1 - alpha) * out[i - 1].im);
            }
            last = out[n - 1];
        }
        return n;
    }
};

// ------------------------------------------------------------------ IQFrontEnd pre-processing chain — iq_frontend.cpp:29-44
struct Preproc {
    PowerDecim decim; DcBlockC dc; bool useDecim, useDc, useConj;
    std::vector<cf> tmp;
};

} // namespace

extern "C" {

const char* orc_kind(void) { return "port"; }
// (the threaded graph is measured with the reference's own blocks only)
double orc_graph_run(double, int, const double*, const double*, const double*, int, int, const float*) { return -1.0; }

void* orc_vfo_create(double inSR, double outSR, double bandwidth, double offset) {
    auto* v = new Vfo; v->init(inSR, outSR, bandwidth, offset); return v;
}
void orc_vfo_destroy(void* h) { delete (Vfo*)h; }
int orc_vfo_process(void* h, int count, const float* in, float* out) { return ((Vfo*)h)->process(count, (const cf*)in, (cf*)out); }
void orc_vfo_set_offset(void* h, double offset) { auto* v = (Vfo*)h; v->offset = offset; v->xlator.setOffset(-offset, v->inSR); }   // rx_vfo.h:72-77
void orc_vfo_set_bandwidth(void* h, double bw) {                                                                                 // rx_vfo.h:60-70
    auto* v = (Vfo*)h; v->bw = bw; v->filterNeeded = (v->bw != v->outSR);
    if (v->filterNeeded) { v->genTaps(); v->filter.setTaps(v->ftaps); }
}
void orc_vfo_set_in_samplerate(void* h, double sr) {                                                                             // rx_vfo.h:35-43
    auto* v = (Vfo*)h; v->inSR = sr; v->xlator.setOffset(-v->offset, sr); v->resamp.inSR = sr; v->resamp.reconfigure();
}
void orc_vfo_set_out_samplerate(void* h, double sr, double bw) {                                                                 // rx_vfo.h:45-58
    auto* v = (Vfo*)h; v->outSR = sr; v->bw = bw; v->filterNeeded = (bw != sr);
    v->resamp.outSR = sr; v->resamp.reconfigure();
    if (v->filterNeeded) { v->genTaps(); v->filter.setTaps(v->ftaps); }
}
void orc_vfo_reset(void* h) { auto* v = (Vfo*)h; v->xlator.reset(); v->resamp.reset(); v->filter.reset(); }                      // rx_vfo.h:79-87
void orc_vfo_get_phase(void* h, float* reim) { auto* v = (Vfo*)h; reim[0] = v->xlator.pr; reim[1] = v->xlator.pi; }
void orc_vfo_set_phase(void* h, float re, float im) { auto* v = (Vfo*)h; v->xlator.pr = re; v->xlator.pi = im; }
void orc_vfo_get_phase_delta(void* h, float* reim) { auto* v = (Vfo*)h; reim[0] = v->xlator.dr; reim[1] = v->xlator.di; }
void orc_vfo_describe(void* h, int* info) {
    auto* v = (Vfo*)h; auto& r = v->resamp;
    const bool hasDecim = (r.mode == Rational::BOTH || r.mode == Rational::DECIM_ONLY);
    const bool hasResamp = (r.mode == Rational::BOTH || r.mode == Rational::RESAMP_ONLY);
    info[0] = (int)r.mode;
    info[1] = hasDecim ? r.decim.ratio : 1;
    info[2] = hasResamp ? r.resamp.interp : 1;
    info[3] = hasResamp ? r.resamp.decim : 1;
    info[4] = hasResamp ? (int)r.rtaps.size() : 0;
    info[5] = hasResamp ? r.resamp.tapsPerPhase : 0;
    info[6] = v->filterNeeded ? (int)v->ftaps.size() : 0;
}
int orc_vfo_get_resamp_taps(void* h, float* out, int cap) {
    auto* v = (Vfo*)h; int n = (int)v->resamp.rtaps.size();
    if (out) { memcpy(out, v->resamp.rtaps.data(), sizeof(float) * std::min(n, cap)); }
    return n;
}
int orc_vfo_get_chan_taps(void* h, float* out, int cap) {
    auto* v = (Vfo*)h; int n = (int)v->ftaps.size();
    if (out) { memcpy(out, v->ftaps.data(), sizeof(float) * std::min(n, cap)); }
    return n;
}

void* orc_decim_create(int ratio) { auto* d = new PowerDecim; d->setRatio(ratio); return d; }
void orc_decim_destroy(void* h) { delete (PowerDecim*)h; }
int orc_decim_process(void* h, int count, const float* in, float* out) { return ((PowerDecim*)h)->process(count, (const cf*)in, (cf*)out); }
int orc_decim_plan(int ratio, int* decims, int* tapCounts, int cap) {
    const PlanDef* p = findPlan(ratio);
    if (!p) { return 0; }
    for (int i = 0; i < p->nstages && i < cap; i++) { decims[i] = p->stages[i].decim; tapCounts[i] = p->stages[i].ntaps; }
    return p->nstages;
}
int orc_decim_stage_taps(int ratio, int stage, float* out, int cap) {
    const PlanDef* p = findPlan(ratio);
    if (!p || stage >= p->nstages) { return 0; }
    if (out) { memcpy(out, p->stages[stage].taps, sizeof(float) * std::min(p->stages[stage].ntaps, cap)); }
    return p->stages[stage].ntaps;
}

void* orc_dcblock_create(double rate) { auto* d = new DcBlockC; d->rate = (float)rate; return d; }
void orc_dcblock_destroy(void* h) { delete (DcBlockC*)h; }
int orc_dcblock_process(void* h, int count, const float* in, float* out) { ((DcBlockC*)h)->process(count, (const cf*)in, (cf*)out); return count; }

void* orc_preproc_create(double sampleRate, int decimRatio, int dcBlocking, int invertIQ) {
    auto* p = new Preproc;
    double effectiveSr = sampleRate / decimRatio;
    p->decim.setRatio(decimRatio);
    p->dc.rate = (float)(50.0 / effectiveSr);      // genDCBlockRate, iq_frontend.h:84-86
    p->useDecim = decimRatio > 1; p->useDc = dcBlocking != 0; p->useConj = invertIQ != 0;
    p->tmp.resize(65536);
    return p;
}
void orc_preproc_destroy(void* h) { delete (Preproc*)h; }
int orc_preproc_process(void* h, int count, const float* in, float* out) {
    auto* p = (Preproc*)h; cf* o = (cf*)out;
    if ((int)p->tmp.size() < count) { p->tmp.resize(count); }
    memcpy(o, in, sizeof(cf) * count);
    if (p->useDecim) { count = p->decim.process(count, o, p->tmp.data()); memcpy(o, p->tmp.data(), sizeof(cf) * count); }
    if (p->useDc) { p->dc.process(count, o, o); }
    if (p->useConj) { for (int i = 0; i < count; i++) { o[i].im = -o[i].im; } }   // math/conjugate.h:12-15
    return count;
}

int orc_lowpass_taps(double cutoff, double transWidth, double sampleRate, float* out, int cap) {
    auto t = lowPass(cutoff, transWidth, sampleRate);
    if (out) { memcpy(out, t.data(), sizeof(float) * std::min<int>((int)t.size(), cap)); }
    return (int)t.size();
}
void orc_fft_window(int type, int nz, float* out) {   // iq_frontend.cpp:303-311: (float)(double window * float sign)
    for (int i = 0; i < nz; i++) {
        float sgn = (i % 2) ? -1.0f : 1.0f;
        if (type == 0) { out[i] = 1.0f * sgn; }
        else if (type == 1) { out[i] = (float)(blackman(i, nz) * sgn); }
        else { out[i] = (float)(nuttall(i, nz) * sgn); }
    }
}
void orc_reshape_params(double sampleRate, int fftSize, double fftRate, int* skip, int* nz) { orc_restated::reshapeParams(sampleRate, fftSize, fftRate, skip, nz); }
void orc_fft_frame(const float* iq, const float* window, int nz, int N, float* outDb, float* spectrum) {
    orc_restated::fftFrame(iq, window, nz, N, outDb, spectrum, powerSpectrum);
}
void orc_do_zoom(int offset, int width, int inSize, int outSize, const float* in, float* out) { orc_restated::doZoom(offset, width, inSize, outSize, in, out); }
void orc_zoom_indices(int offset, int width, int inSize, int outSize, int* sId, int* uFactor) { orc_restated::zoomIndices(offset, width, inSize, outSize, sId, uFactor); }
void orc_palette_indices(const float* row, int n, float wfMin, float wfMax, int* idx) { orc_restated::paletteIndices(row, n, wfMin, wfMax, idx); }
int orc_vfo_signal_info(const float* row, int rawFFTSize, double wholeBandwidth, double centerOffset, double bandwidth, float* strength, float* snr) {
    return orc_restated::vfoSignalInfo(row, rawFFTSize, wholeBandwidth, centerOffset, bandwidth, strength, snr) ? 1 : 0;
}
void orc_smooth_hold(float* latest, float* smoothingBuf, float* hold, int width, int smoothing, float alpha, int doHold, float holdSpeed) {
    orc_restated::smoothAndHold(latest, smoothingBuf, hold, width, smoothing != 0, alpha, doHold != 0, holdSpeed);
}
void orc_make_palette(const float* colors, int colorCount, unsigned int* lut) { orc_restated::makePalette(colors, colorCount, lut); }

void* orc_fm_create(double sr, double bw, int lp, int hp) { auto* d = new FmDemod; d->init(sr, bw, lp != 0, hp != 0); return d; }
void orc_fm_destroy(void* h) { delete (FmDemod*)h; }
int orc_fm_process(void* h, int count, const float* in, float* out) { return ((FmDemod*)h)->process(count, (const cf*)in, out); }
void* orc_ssb_create(int mode, double bw, double sr, double att, double dec) { auto* d = new SsbDemod; d->init(mode, bw, sr, att, dec); return d; }
void orc_ssb_destroy(void* h) { delete (SsbDemod*)h; }
int orc_ssb_process(void* h, int count, const float* in, float* out) { return ((SsbDemod*)h)->process(count, (const cf*)in, out); }
void* orc_am_create(int agcMode, double bw, double att, double dec, double dcRate, double sr) { auto* d = new AmDemod; d->init(agcMode, bw, att, dec, dcRate, sr); return d; }
void orc_am_destroy(void* h) { delete (AmDemod*)h; }
int orc_am_process(void* h, int count, const float* in, float* out) { return ((AmDemod*)h)->process(count, (const cf*)in, out); }
void* orc_wfm_create(double deviation, double samplerate, int stereo, int lowPass) { auto* d = new WfmDemod; d->init(deviation, samplerate, stereo != 0, lowPass != 0); return d; }
void orc_wfm_destroy(void* h) { delete (WfmDemod*)h; }
int orc_wfm_process(void* h, int count, const float* in, float* out) { return ((WfmDemod*)h)->process(count, (const cf*)in, out); }
void* orc_wfm_create_rds(double deviation, double samplerate, int stereo, int lowPass) { auto* d = new WfmDemod; d->init(deviation, samplerate, stereo != 0, lowPass != 0, true); return d; }
int orc_wfm_process_rds(void* h, int count, const float* in, float* out, float* rdsOut) { auto* d = (WfmDemod*)h; d->process(count, (const cf*)in, out, (cf*)rdsOut); return d->rdsCount; }
void* orc_fmif_create(int bins) { auto* f = new FmIf; f->init(bins); return f; }
void orc_fmif_destroy(void* h) { delete (FmIf*)h; }
int orc_fmif_process(void* h, int count, const float* in, float* out) { return ((FmIf*)h)->process(count, (const cf*)in, (cf*)out); }
int orc_volume_process(double volume, int muted, int count, const float* in, float* out) {   // audio/volume.h:13-16,47-50
    const float v = powf(volume, 2);
    const float g = muted ? 0.0f : v;
    for (int i = 0; i < 2 * count; i++) { out[i] = in[i] * g; }
    return count;
}
// wav::Writer over riff::Writer (utils/wav.cpp:36-98,148-182; utils/riff.cpp:17-141): "RIFF" <size> "WAVE", a 16-byte "fmt "
// chunk (codec 1 PCM / 3 float, channels, rate, bytes per second, bytes per frame, bit depth), then one "data" chunk; chunk
// sizes are patched in when the chunks end.  Conversions: uint8 (x*127)+128 truncated, int16 / int32 by VOLK's generic
// convert (scale by 32767 / 2147483647, clamp, round to nearest even), float32 as is.
// SampleStreamCompressor / SampleStreamDecompressor (compression/sample_stream_compressor.h:30-62, _decompressor.h:14-35)
int orc_stream_compress(int count, int pcmType, const float* in, unsigned char* out) {
    uint16_t z = 0, t = (uint16_t)pcmType;
    memcpy(out, &z, 2); memcpy(out + 2, &t, 2);
    if (pcmType == 2) {
        const float zero = 0.0f;
        memcpy(out + 4, &zero, 4);
        memcpy(out + 8, in, (size_t)count * 8);
        return 8 + count * 8;
    }
    float maxVal = 1.0f;
    if (count > 0) {                                   // generic volk_32f_index_max_32u: strict >, first of equals
        maxVal = in[0];
        for (int i = 1; i < 2 * count; i++) { if (in[i] > maxVal) { maxVal = in[i]; } }
    }
    memcpy(out + 4, &maxVal, 4);
    if (pcmType == 0) {
        const float sc = 128.0f / maxVal;
        int8_t* d = (int8_t*)(out + 8);
        for (int i = 0; i < 2 * count; i++) {
            float r = in[i] * sc;
            if (r > 127.0f) { r = 127.0f; } else if (r < -128.0f) { r = -128.0f; }
            d[i] = (int8_t)rintf(r);
        }
        return 8 + count * 2;
    }
    const float sc = 32768.0f / maxVal;
    int16_t* d = (int16_t*)(out + 8);
    for (int i = 0; i < 2 * count; i++) {
        float r = in[i] * sc;
        if (r > 32767.0f) { r = 32767.0f; } else if (r < -32768.0f) { r = -32768.0f; }
        d[i] = (int16_t)rintf(r);
    }
    return 8 + count * 4;
}
int orc_stream_decompress(int nbytes, const unsigned char* in, float* out) {
    uint16_t t; float scaler;
    memcpy(&t, in + 2, 2); memcpy(&scaler, in + 4, 4);
    if (t == 2) { memcpy(out, in + 8, (size_t)nbytes - 8); return (nbytes - 8) / 8; }
    if (t == 1) {
        const int n = (nbytes - 8) / 4; const float sc = 32768.0f / scaler; const int16_t* d = (const int16_t*)(in + 8);
        for (int i = 0; i < 2 * n; i++) { out[i] = (float)d[i] / sc; }
        return n;
    }
    if (t == 0) {
        const int n = (nbytes - 8) / 2; const float sc = 128.0f / scaler; const int8_t* d = (const int8_t*)(in + 8);
        for (int i = 0; i < 2 * n; i++) { out[i] = (float)d[i] / sc; }
        return n;
    }
    return 0;
}
int orc_wav_write(const char* path, int channels, double samplerate, int sampleType, const float* samples, int count, int blockLen) {
    FILE* f = fopen(path, "wb");
    if (!f) { return -1; }
    const int bits = (sampleType == 0) ? 8 : (sampleType == 1) ? 16 : 32;
    const uint32_t bytesPerSamp = (uint32_t)(bits / 8) * (uint32_t)channels;
    uint32_t dataBytes = 0;
    auto u32 = [&](uint32_t v) { fwrite(&v, 4, 1, f); };
    auto u16 = [&](uint16_t v) { fwrite(&v, 2, 1, f); };
    fwrite("RIFF", 1, 4, f); u32(0); fwrite("WAVE", 1, 4, f);
    fwrite("fmt ", 1, 4, f); u32(16);
    u16(sampleType == 3 ? 3 : 1); u16((uint16_t)channels); u32((uint32_t)(uint64_t)samplerate);
    u32((uint32_t)(bytesPerSamp * (uint64_t)samplerate)); u16((uint16_t)bytesPerSamp); u16((uint16_t)bits);
    fwrite("data", 1, 4, f); u32(0);
    std::vector<uint8_t> b8; std::vector<int16_t> b16; std::vector<int32_t> b32;
    for (int i = 0; i < count; i += blockLen) {
        const int n = std::min(blockLen, count - i), t = n * channels;
        const float* x = samples + (size_t)i * channels;
        if (sampleType == 0) {
            b8.resize(t);
            for (int k = 0; k < t; k++) { b8[k] = (x[k] * 127.0f) + 128.0f; }
            fwrite(b8.data(), 1, t, f);
        }
        else if (sampleType == 1) {
            b16.resize(t);
            for (int k = 0; k < t; k++) {
                float r = x[k] * 32767.0f;
                if (r > 32767.0f) { r = 32767.0f; } else if (r < -32768.0f) { r = -32768.0f; }
                b16[k] = (int16_t)rintf(r);
            }
            fwrite(b16.data(), 2, t, f);
        }
        else if (sampleType == 2) {
            b32.resize(t);
            for (int k = 0; k < t; k++) {
                const float r = x[k] * 2147483647.0f;
                b32[k] = (r >= 2147483648.0f) ? 2147483647 : (r < -2147483648.0f) ? (int32_t)(-2147483647 - 1) : (int32_t)rintf(r);
            }
            fwrite(b32.data(), 4, t, f);
        }
        else { fwrite(x, 4, t, f); }
        dataBytes += (uint32_t)n * bytesPerSamp;
    }
    // endChunk(data), endRIFF: sizes patched; the RIFF size counts the form, the fmt chunk and the data chunk with their headers
    fseek(f, 40, SEEK_SET); u32(dataBytes);
    fseek(f, 4, SEEK_SET); u32(4 + (8 + 16) + (8 + dataBytes));
    fclose(f);
    return 0;
}
void* orc_ifchain_create(int nbOn, double nbRate, double nbLevel, int sqOn, double sqLevel) { auto* c = new IfChain; c->init(nbOn, nbRate, nbLevel, sqOn, sqLevel); return c; }
void orc_ifchain_destroy(void* h) { delete (IfChain*)h; }
int orc_ifchain_process(void* h, int count, const float* in, float* out) { return ((IfChain*)h)->process(count, (const cf*)in, (cf*)out); }
void* orc_afpost_create(double inSR, double outSR, double tau) { auto* a = new AfPost; a->init(inSR, outSR, tau); return a; }
void orc_afpost_destroy(void* h) { delete (AfPost*)h; }
int orc_afpost_process(void* h, int count, const float* in, float* out) { return ((AfPost*)h)->process(count, (const cf*)in, (cf*)out); }
void* orc_agc_create(double sp, double att, double dec, double mg, double moa, double ig) { auto* a = new Agc<float>; a->init(sp, att, dec, mg, moa, ig); return a; }
void orc_agc_destroy(void* h) { delete (Agc<float>*)h; }
int orc_agc_process(void* h, int count, const float* in, float* out) { return ((Agc<float>*)h)->process(count, in, out); }

} // extern "C"
