// TEST INFRASTRUCTURE ONLY (oracle).  Not linked into the product library.
//
// Float64 DFT of any length, the FFTW stand-in of both oracle libraries for transforms too long for the O(N^2) sum of
// oracle/shim/fftw3.h: the SignalDetector's sampleRate / 10 points (dsp/detector/signal_detector.cpp:330-346) and logmmse's
// 2 * floor(0.02 * sampleRate) (misc_modules/noise_reduction_logmmse/src/logmmse.h:256,274).  FFTW3f is an un-vendored,
// un-pinned dependency of the reference (core/CMakeLists.txt:160-164); what is restated here is the published definition
//     Y[k] = sum_n X[n] exp(sign * 2 pi i k n / N)      (unnormalised; sign -1 forward, +1 backward)
// evaluated as a decimation-in-time mixed-radix recursion over the prime factors of N in float64, every twiddle taken
// from one table of cos / sin (2 pi m / N) computed in float64.  The caller narrows to fp32 once per output, so the result
// is closer to the exact DFT than any fftwf build and sits in the middle of what real FFTW installations return.
#pragma once
#include <cmath>
#include <complex>
#include <vector>

namespace orc_fft64 {

typedef std::complex<double> cd;

struct Plan {
    int n = 0;
    int sign = -1;
    std::vector<cd> tw;            // exp(sign * 2 pi i m / n), m < n
    std::vector<int> factors;      // prime factors of n, smallest first

    void init(int n_, int sign_) {
        n = n_; sign = sign_;
        tw.resize(n > 0 ? n : 0);
        for (int m = 0; m < n; m++) {
            const double a = 2.0 * 3.14159265358979323846 * (double)m / (double)n;
            tw[m] = cd(cos(a), sign < 0 ? -sin(a) : sin(a));
        }
        factors.clear();
        int r = n;
        for (int p = 2; (long long)p * p <= r; p++) { while (r % p == 0) { factors.push_back(p); r /= p; } }
        if (r > 1) { factors.push_back(r); }
    }

    // out[0..len) = DFT of in[0], in[stride], ...; level = index into `factors`; twStep = n / len
    void rec(const cd* in, cd* out, int len, int stride, int level, std::vector<cd>& tmp) const {
        if (len == 1) { out[0] = in[0]; return; }
        const int p = factors[level], m = len / p;
        const int twStep = n / len;
        for (int r = 0; r < p; r++) { rec(in + (size_t)r * stride, out + (size_t)r * m, m, stride * p, level + 1, tmp); }
        const int pStep = n / p;
        cd* t = tmp.data() + (size_t)level * 2 * (size_t)maxFactor();
        cd* y = t + maxFactor();
        for (int k = 0; k < m; k++) {
            for (int r = 0; r < p; r++) { t[r] = out[(size_t)r * m + k] * tw[((long long)r * k % len) * twStep]; }
            for (int q = 0; q < p; q++) {
                cd acc = t[0];
                for (int r = 1; r < p; r++) { acc += t[r] * tw[((long long)r * q % p) * pStep]; }
                y[q] = acc;
            }
            for (int q = 0; q < p; q++) { out[(size_t)q * m + k] = y[q]; }
        }
    }

    int maxFactor() const { int f = 1; for (int p : factors) { if (p > f) { f = p; } } return f; }

    void run(const cd* in, cd* out) const {
        if (n <= 0) { return; }
        std::vector<cd> tmp((factors.size() + 1) * 2 * (size_t)maxFactor());
        rec(in, out, n, 1, 0, tmp);
    }
};

} // namespace orc_fft64
