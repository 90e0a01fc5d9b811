// TEST INFRASTRUCTURE ONLY (oracle). Not linked into the product library.
//
// Stand-in for <fftw3.h> (single precision API).  FFTW3f is an un-vendored, un-pinned third-party
// dependency of the reference (core/CMakeLists.txt:160-164) and is not installed in this image.  This header
// serves the one call site on the radio IF chain that uses FFTW directly:
//   dsp/noise_reduction/fm_if.h:106-124   fftwf_malloc / fftwf_plan_dft_1d / fftwf_execute / fftwf_destroy_plan / fftwf_free
// The transform is the published definition  Y[k] = sum_n X[n] exp(-+ 2 pi i k n / N)  (unnormalised, sign -1 for
// FFTW_FORWARD, +1 for FFTW_BACKWARD), evaluated directly in float64 and narrowed to fp32 once per output: more
// accurate than fftwf itself (whose error is ~1e-7 relative), so it stands in the middle of what real FFTW builds
// return.  O(N^2) up to 64 points (that call site uses N <= 32).
// The other call site, utils/arrays.cpp:509-551 (fftwPlanImplFFTW: the SignalDetector's and logmmse's plans, which holds
// the plan as `fftwf_plan_s*`), runs thousands to millions of points: those go through the float64 mixed-radix
// recursion of oracle/fft64.h -- same definition, same single narrowing to fp32.
#pragma once
#include <cmath>
#include <cstdlib>
#include <vector>

#include "../fft64.h"

typedef float fftwf_complex[2];
#define FFTW_FORWARD (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_ESTIMATE (1U << 6)

#define oracle_fftwf_plan_s fftwf_plan_s
struct fftwf_plan_s {
    int n;
    int sign;
    fftwf_complex* in;
    fftwf_complex* out;
    std::vector<double> cs, sn;   // cos / sin of 2 pi m / n, m < n
    orc_fft64::Plan fast;         // n > 64
};
typedef oracle_fftwf_plan_s* fftwf_plan;

static inline void* fftwf_malloc(size_t n) {
    void* p = nullptr;
    if (posix_memalign(&p, 64, n ? n : 64) != 0) { return nullptr; }
    return p;
}
static inline void fftwf_free(void* p) { free(p); }

static inline fftwf_plan fftwf_plan_dft_1d(int n, fftwf_complex* in, fftwf_complex* out, int sign, unsigned) {
    auto* p = new oracle_fftwf_plan_s;
    p->n = n; p->sign = sign; p->in = in; p->out = out;
    if (n > 64) { p->fast.init(n, sign); return p; }
    p->cs.resize(n); p->sn.resize(n);
    for (int m = 0; m < n; m++) {
        const double a = 2.0 * 3.14159265358979323846 * (double)m / (double)n;
        p->cs[m] = cos(a); p->sn[m] = sin(a);
    }
    return p;
}
static inline void fftwf_destroy_plan(fftwf_plan p) { delete p; }

static inline void fftwf_execute(const fftwf_plan p) {
    const int n = p->n;
    if (n > 64) {
        std::vector<orc_fft64::cd> x(n), y(n);
        for (int j = 0; j < n; j++) { x[j] = orc_fft64::cd(p->in[j][0], p->in[j][1]); }
        p->fast.run(x.data(), y.data());
        for (int k = 0; k < n; k++) { p->out[k][0] = (float)y[k].real(); p->out[k][1] = (float)y[k].imag(); }
        return;
    }
    std::vector<double> yr(n), yi(n);
    for (int k = 0; k < n; k++) {
        double sr = 0.0, si = 0.0;
        for (int j = 0; j < n; j++) {
            const int m = (int)(((long long)k * j) % n);
            const double c = p->cs[m], s = (p->sign < 0) ? -p->sn[m] : p->sn[m];
            const double xr = p->in[j][0], xi = p->in[j][1];
            sr += xr * c - xi * s;
            si += xr * s + xi * c;
        }
        yr[k] = sr; yi[k] = si;
    }
    for (int k = 0; k < n; k++) { p->out[k][0] = (float)yr[k]; p->out[k][1] = (float)yi[k]; }
}
