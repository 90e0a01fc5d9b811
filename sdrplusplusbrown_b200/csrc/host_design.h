// Host-side filter/window design and integer state machines for the B200 IQ front end.
// Double-precision math narrowed to fp32 exactly where the reference narrows, so tap tables
// come out bit-identical to the reference's (checked against the oracle in tests/test_vfo_gpu.py: test_vfo_matches_oracle_per_block).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

namespace b200 {
namespace design {

constexpr double kPi = 3.14159265358979323846;   // DB_M_PI (dsp/math/constants.h:3)

// ---- power-of-two decimation plans (multirate/decim/plans.h:36-140), coefficients as bit patterns
struct StageDef { int decim; const uint32_t* taps; int ntaps; };
struct PlanDef { int ratio; int nstages; StageDef stages[4]; };

namespace tables {
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
constexpr int kNumPlans = sizeof(kPlans) / sizeof(kPlans[0]);
} // namespace tables

inline int maxDecimRatio() { return 1 << tables::kNumPlans; }   // PowerDecimator::getMaxRatio, power_decimator.h:29-31
inline const PlanDef* findPlan(int ratio) {
    for (int i = 0; i < tables::kNumPlans; i++) { if (tables::kPlans[i].ratio == ratio) { return &tables::kPlans[i]; } }
    return nullptr;
}
inline bool validDecimRatio(int ratio) { return ratio == 1 || findPlan(ratio) != nullptr; }   // power_decimator.h:112-115

// ---- windows (dsp/window/cosine.h:7-16; divisor is N, not N-1)
inline double cosineWindow(double n, double N, const double* c, int nc) {
    double w = 0.0, sgn = 1.0;
    for (int i = 0; i < nc; i++) { w += sgn * c[i] * cos((double)i * 2.0 * kPi * n / N); sgn = -sgn; }
    return w;
}
inline double nuttall(double n, double N) { const double c[] = { 0.355768, 0.487396, 0.144232, 0.012604 }; return cosineWindow(n, N, c, 4); }
inline double blackman(double n, double N) { const double c[] = { 0.42, 0.5, 0.08 }; return cosineWindow(n, N, c, 3); }

inline double hzToRads(double f, double sr) { return 2.0 * kPi * (f / sr); }   // math/hz_to_rads.h:6-8

// ---- windowed-sinc designers (taps/windowed_sinc.h:9-33, low_pass.h, high_pass.h, band_pass.h)
template <typename W>
inline std::vector<float> windowedSinc(int count, double omega, W window) {
    std::vector<float> t((size_t)std::max(count, 0));
    const double half = (double)count / 2.0;
    const double corr = omega / kPi;
    for (int i = 0; i < count; i++) {
        const double x = (double)i - half + 0.5;
        const double s = (x * omega == 0.0) ? 1.0 : sin(x * omega) / (x * omega);
        t[i] = (float)(s * window(x - half, (double)count) * corr);
    }
    return t;
}
inline int estimateTapCount(double transWidth, double sr) { return (int)(3.8 * sr / transWidth); }
inline std::vector<float> lowPass(double cutoff, double transWidth, double sr) {
    return windowedSinc(estimateTapCount(transWidth, sr), hzToRads(cutoff, sr), [](double n, double N) { return nuttall(n, N); });
}
inline std::vector<float> highPass(double cutoff, double transWidth, double sr) {
    return windowedSinc(estimateTapCount(transWidth, sr), hzToRads((sr / 2.0) - cutoff, sr),
                        [](double n, double N) { return nuttall(n, N) * (((int)round(n) % 2) ? -1.0f : 1.0f); });
}
inline std::vector<float> bandPass(double lo, double hi, double transWidth, double sr) {
    const float w0 = (float)hzToRads((lo + hi) / 2.0, sr);
    return windowedSinc(estimateTapCount(transWidth, sr), hzToRads((hi - lo) / 2.0, sr),
                        [=](double n, double N) { return (double)(2.0f * cosf(w0 * (float)n)) * nuttall(n, N); });
}

// taps::bandPass<complex_t> (taps/band_pass.h:10-24 over windowed_sinc.h:26-29): the low-pass prototype shifted to the
// band centre by a complex exponential; the products are formed in float in the order the reference's complex_t
// operators give them.  Real and imaginary parts are returned separately (the pilot filter runs as two real FIRs).
inline void bandPassComplex(double lo, double hi, double transWidth, double sr, bool oddTapCount, std::vector<float>& re, std::vector<float>& im) {
    const float w0 = (float)hzToRads((lo + hi) / 2.0, sr);
    int count = estimateTapCount(transWidth, sr);
    if (oddTapCount && !(count % 2)) { count++; }
    re.assign((size_t)std::max(count, 0), 0.0f); im.assign((size_t)std::max(count, 0), 0.0f);
    const double omega = hzToRads((hi - lo) / 2.0, sr);
    const double half = (double)count / 2.0, corr = 1.0 * omega / kPi;
    for (int i = 0; i < count; i++) {
        const double t = (double)i - half + 0.5, n = t - half;
        const float x = -w0 * (float)n;
        const float wnd = (float)nuttall(n, (double)count);
        const float wr = cosf(x) * wnd, wi = sinf(x) * wnd;
        const double a = t * omega;
        const float cr = (float)((a == 0.0) ? 1.0 : sin(a) / a), ci = 0.0f;          // math::sinc (math/sinc.h:5-7)
        const float pr = (cr * wr) - (ci * wi), pi = (ci * wr) + (cr * wi);
        re[i] = pr * (float)corr; im[i] = pi * (float)corr;
    }
}

// ---- FFT window with the fft-shift alternation (iq_frontend.cpp:303-311)
inline std::vector<float> fftWindow(int type, int nz) {
    std::vector<float> w((size_t)nz);
    for (int i = 0; i < nz; i++) {
        const float sgn = (i % 2) ? -1.0f : 1.0f;
        if (type == 0) { w[i] = sgn; }
        else if (type == 1) { w[i] = (float)(blackman(i, nz) * sgn); }
        else { w[i] = (float)(nuttall(i, nz) * sgn); }
    }
    return w;
}
// IQFrontEnd::genReshapeParams (iq_frontend.h:88-92)
inline void reshapeParams(double sr, int size, double rate, int& skip, int& nz) {
    const int interval = (int)std::round(sr / rate);
    nz = std::min<int>(interval, size);
    skip = interval - nz;
}

// ---- RationalResampler::reconfigure (rational_resampler.h:128-173), as a pure function
struct ResampPlan {
    int mode;        // 0 BOTH, 1 DECIM_ONLY, 2 RESAMP_ONLY, 3 NONE
    bool useDecim;   // whether decim.setRatio() was (re)issued
    int ratio;       // power-of-two pre-decimation (valid when useDecim)
    int interp, decim;
    std::vector<float> taps;   // prototype low-pass * interp (empty when the resampler is skipped)
};
inline ResampPlan planResampler(double inSR, double outSR) {
    ResampPlan p;
    const int maxRatio = maxDecimRatio();
    const int predecPower = std::min<int>((int)floor(log2(inSR / outSR)), maxRatio);
    p.useDecim = (inSR > outSR && predecPower > 0);
    p.ratio = p.useDecim ? std::min<int>(1 << std::min(predecPower, 30), maxRatio) : 1;
    const double intSR = p.useDecim ? inSR / (double)p.ratio : inSR;
    const int IntSR = (int)round(intSR), OutSR = (int)round(outSR);
    const int g = std::gcd(IntSR, OutSR);
    p.interp = OutSR / g;
    p.decim = IntSR / g;
    if (p.interp == p.decim) { p.mode = p.useDecim ? 1 : 3; return p; }
    const double tapSR = intSR * (double)p.interp;
    const double tapBW = std::min<double>(inSR, outSR) / 2.0;
    p.taps = lowPass(tapBW, tapBW * 0.1, tapSR);
    for (auto& t : p.taps) { t *= (float)p.interp; }
    p.mode = p.useDecim ? 0 : 2;
    return p;
}
// buildPolyphaseBank (polyphase_bank.h:14-47); rows padded with zeros to rowStride floats
inline std::vector<float> polyphaseBank(const std::vector<float>& taps, int L, int& tapsPerPhase, int& rowStride) {
    tapsPerPhase = ((int)taps.size() + L - 1) / L;
    rowStride = (tapsPerPhase + 3) & ~3;
    std::vector<float> bank((size_t)L * rowStride, 0.0f);
    const int tot = L * tapsPerPhase;
    for (int i = 0; i < tot; i++) {
        bank[(size_t)((L - 1) - (i % L)) * rowStride + (i / L)] = (i < (int)taps.size()) ? taps[i] : 0.0f;
    }
    return bank;
}

// ---- integer state machines (bit-exact counts / decimation phase)
// DecimatingFIR::process loop bounds (decimating_fir.h:51-62)
inline int decimOutCount(int offset, int decim, int nIn) { return (offset < nIn) ? (nIn - offset + decim - 1) / decim : 0; }
inline int decimNextOffset(int offset, int decim, int nIn) { return offset + decimOutCount(offset, decim, nIn) * decim - nIn; }
// PolyphaseResampler::process loop (polyphase_resampler.h:75-93): before output k the walk is at
// total = phase0 + k*M  ->  buffer offset offset0 + total / L, bank row total % L.
inline int polyOutCount(int phase0, int offset0, int L, int M, int nIn) {
    if (offset0 >= nIn) { return 0; }
    const int64_t R = (int64_t)nIn - 1 - offset0;
    const int64_t num = (R + 1) * (int64_t)L - phase0;   // need phase0 + k*M < (R+1)*L
    return (int)((num + M - 1) / M);
}
inline void polyNextState(int& phase, int& offset, int L, int M, int nIn) {
    const int n = polyOutCount(phase, offset, L, M, nIn);
    const int64_t total = (int64_t)phase + (int64_t)n * M;
    offset = (int)(offset + total / L - nIn);
    phase = (int)(total % L);
}

// ---- rotator calibration.  The reference's phasor is the fp32 recursion of VOLK's generic
// rotator (phase *= phaseDelta per sample, phase /= |phase| every 512 samples;
// channel/frequency_xlator.h:47-51).  Its mean step differs from atan2(di, dr) by a
// deterministic rounding bias ~1e-10 rad/sample; running the exact recursion once for `steps`
// samples recovers the effective step to ~1e-13 rad/sample.  Plain IEEE fp32, no FMA.
inline double calibrateRotatorStep(float dr, float di, int steps) {
    float pr = 1.0f, pi = 0.0f;
    for (int done = 0; done < steps;) {
        const int chunk = std::min(512, steps - done);
        for (int j = 0; j < chunk; j++) {
            const float nr = pr * dr - pi * di;
            const float ni = pr * di + pi * dr;
            pr = nr; pi = ni;
        }
        const float h = hypotf(pr, pi);
        pr /= h; pi /= h;
        done += chunk;
    }
    const double thetaEff = atan2((double)di, (double)dr);
    const double end = atan2((double)pi, (double)pr);
    const double k = std::round((thetaEff * steps - end) / (2.0 * kPi));
    return (end + 2.0 * kPi * k) / (double)steps;
}

// ---- doZoom index walk (gui/widgets/waterfall.cpp:98-152): fp32 running start index
inline void zoomTable(int offset, int width, int inSize, int outSize, std::vector<int>& sid, std::vector<int>& uf) {
    if (width > 524288) { width = 524288; }
    if (offset < 0) { offset = 0; }
    const float factor = (float)width / (float)outSize;
    const int sFactor = (int)ceilf(factor);
    float id = (float)offset;
    sid.resize(outSize); uf.resize(outSize);
    for (int i = 0; i < outSize; i++) {
        const int s = (int)id;
        sid[i] = s;
        uf[i] = (s + sFactor > inSize) ? inSize - s : sFactor;
        id += factor;
    }
}

} // namespace design
} // namespace b200
