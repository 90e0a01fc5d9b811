// TEST INFRASTRUCTURE ONLY (oracle). Not linked into the product library.
//
// Stand-in for <volk/volk.h>.  VOLK is an un-vendored, un-pinned third-party
// dependency of the reference (core/CMakeLists.txt:169, pkg_check_modules(VOLK
// REQUIRED volk)); it is not installed in this image.  This header restates the
// *generic* (non-SIMD) kernels of VOLK's public API for exactly the call sites
// on the IQ-front-end path, so that the reference's own header-only DSP blocks
// compile unchanged against it.  Every routine is plain sequential IEEE fp32
// (compile with -ffp-contract=off so no FMA contraction sneaks in).
//
// Call sites served (reference file:line):
//   dsp/buffer/buffer.h:20-34                  volk_malloc / volk_free / volk_get_alignment
//   dsp/channel/frequency_xlator.h:47-51       rotator / rotator2
//   dsp/filter/fir.h:71-78, decimating_fir.h:53-59, multirate/polyphase_resampler.h:78-81   dot products
//   dsp/math/conjugate.h:13                    conjugate
//   signal_path/iq_frontend.cpp:271,282        window multiply, power spectrum
//   dsp/demod/am.h:109,120                     magnitude
//   dsp/convert/complex_to_real.h:15, mono_to_stereo.h:13   deinterleave / interleave
//   dsp/noise_reduction/fm_if.h:51,58-59       window multiply, magnitude, index of the maximum
//   utils/wav.cpp:161,165                      float -> int16 / int32 PCM
//   dsp/compression/sample_stream_compressor.h:47-59, sample_stream_decompressor.h:24-32   server stream PCM
#pragma once
#include <complex>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cstdint>

typedef std::complex<float> lv_32fc_t;
#define lv_cmake(r, i) lv_32fc_t((r), (i))
#define lv_creal(x) ((x).real())
#define lv_cimag(x) ((x).imag())

// Pretend to be an old VOLK so frequency_xlator.h picks the by-value rotator
// (the by-pointer rotator2 of VOLK >= 3.1 has identical arithmetic).
#ifndef VOLK_VERSION
#define VOLK_VERSION 020500
#endif

static inline size_t volk_get_alignment(void) { return 64; }
static inline void* volk_malloc(size_t size, size_t alignment) {
    void* p = nullptr;
    if (size == 0) size = alignment;
    if (posix_memalign(&p, alignment, size) != 0) return nullptr;
    return p;
}
static inline void volk_free(void* p) { free(p); }

// ---- rotator: out = in * phase; phase *= inc; renormalise every 512 samples
// and once more after a ragged tail (counters restart on every call).
static inline void oracle_rotator_core(lv_32fc_t* out, const lv_32fc_t* in, lv_32fc_t inc, lv_32fc_t* phase, unsigned int n) {
    float pr = phase->real(), pi = phase->imag();
    const float dr = inc.real(), di = inc.imag();
    unsigned int done = 0;
    const unsigned int RELOAD = 512;
    const unsigned int full = n / RELOAD;
    for (unsigned int b = 0; b < full; b++) {
        for (unsigned int j = 0; j < RELOAD; j++, done++) {
            const float xr = in[done].real(), xi = in[done].imag();
            out[done] = lv_32fc_t(xr * pr - xi * pi, xr * pi + xi * pr);
            const float nr = pr * dr - pi * di;
            const float ni = pr * di + pi * dr;
            pr = nr; pi = ni;
        }
        const float h = hypotf(pr, pi);
        pr /= h; pi /= h;
    }
    unsigned int tail = n % RELOAD;
    for (unsigned int j = 0; j < tail; j++, done++) {
        const float xr = in[done].real(), xi = in[done].imag();
        out[done] = lv_32fc_t(xr * pr - xi * pi, xr * pi + xi * pr);
        const float nr = pr * dr - pi * di;
        const float ni = pr * di + pi * dr;
        pr = nr; pi = ni;
    }
    if (tail) {
        const float h = hypotf(pr, pi);
        pr /= h; pi /= h;
    }
    *phase = lv_32fc_t(pr, pi);
}
static inline void volk_32fc_s32fc_x2_rotator_32fc(lv_32fc_t* out, const lv_32fc_t* in, lv_32fc_t inc, lv_32fc_t* phase, unsigned int n) {
    oracle_rotator_core(out, in, inc, phase, n);
}
static inline void volk_32fc_s32fc_x2_rotator2_32fc(lv_32fc_t* out, const lv_32fc_t* in, const lv_32fc_t* inc, lv_32fc_t* phase, unsigned int n) {
    oracle_rotator_core(out, in, *inc, phase, n);
}

// ---- dot products: one running fp32 sum per component, in index order
static inline void volk_32fc_32f_dot_prod_32fc(lv_32fc_t* result, const lv_32fc_t* x, const float* taps, unsigned int n) {
    float sr = 0.0f, si = 0.0f;
    const float* p = reinterpret_cast<const float*>(x);
    for (unsigned int i = 0; i < n; i++) {
        sr += p[2 * i] * taps[i];
        si += p[2 * i + 1] * taps[i];
    }
    *result = lv_32fc_t(sr, si);
}
static inline void volk_32f_x2_dot_prod_32f(float* result, const float* a, const float* b, unsigned int n) {
    float s = 0.0f;
    for (unsigned int i = 0; i < n; i++) { s += a[i] * b[i]; }
    *result = s;
}
static inline void volk_32fc_x2_dot_prod_32fc(lv_32fc_t* result, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    float sr = 0.0f, si = 0.0f;
    for (unsigned int i = 0; i < n; i++) {
        const float ar = a[i].real(), ai = a[i].imag(), br = b[i].real(), bi = b[i].imag();
        sr += ar * br - ai * bi;
        si += ar * bi + ai * br;
    }
    *result = lv_32fc_t(sr, si);
}

// ---- element-wise helpers
static inline void volk_32fc_conjugate_32fc(lv_32fc_t* out, const lv_32fc_t* in, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = lv_32fc_t(in[i].real(), -in[i].imag()); }
}
static inline void volk_32fc_32f_multiply_32fc(lv_32fc_t* out, const lv_32fc_t* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = lv_32fc_t(a[i].real() * b[i], a[i].imag() * b[i]); }
}
static inline void volk_32fc_x2_multiply_32fc(lv_32fc_t* out, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) {
        const float ar = a[i].real(), ai = a[i].imag(), br = b[i].real(), bi = b[i].imag();
        out[i] = lv_32fc_t(ar * br - ai * bi, ar * bi + ai * br);
    }
}
static inline void volk_32fc_magnitude_32f(float* out, const lv_32fc_t* in, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) {
        const float r = in[i].real(), q = in[i].imag();
        out[i] = sqrtf(r * r + q * q);
    }
}
// index of the maximum: the first of equal maxima wins (strict >), as in VOLK's generic kernel
static inline void volk_32f_index_max_32u(uint32_t* target, const float* src, uint32_t n) {
    if (n == 0) { return; }
    float best = src[0];
    uint32_t idx = 0;
    for (uint32_t i = 1; i < n; i++) {
        if (src[i] > best) { best = src[i]; idx = i; }
    }
    *target = idx;
}
static inline void volk_32fc_deinterleave_real_32f(float* out, const lv_32fc_t* in, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = in[i].real(); }
}
static inline void volk_32f_x2_interleave_32fc(lv_32fc_t* out, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = lv_32fc_t(a[i], b[i]); }
}
static inline void volk_32f_s32f_multiply_32f(float* out, const float* a, float s, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = a[i] * s; }
}
static inline void volk_32f_x2_add_32f(float* out, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = a[i] + b[i]; }
}
static inline void volk_32f_x2_subtract_32f(float* out, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = a[i] - b[i]; }
}
static inline void volk_32f_x2_multiply_32f(float* out, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = a[i] * b[i]; }
}
// ---- used by utils/arrays.cpp only (the SignalDetector's and logmmse's array helpers; arrays.cpp:801,837,933,1036)
static inline void volk_32fc_x2_add_32fc(lv_32fc_t* out, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = lv_32fc_t(a[i].real() + b[i].real(), a[i].imag() + b[i].imag()); }
}
static inline void volk_32f_x2_divide_32f(float* out, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = a[i] / b[i]; }
}
static inline void volk_32fc_s32fc_multiply_32fc(lv_32fc_t* out, const lv_32fc_t* a, const lv_32fc_t s, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) {
        const float ar = a[i].real(), ai = a[i].imag(), br = s.real(), bi = s.imag();
        out[i] = lv_32fc_t(ar * br - ai * bi, ar * bi + ai * br);
    }
}
// VOLK's generic "fast exp" is Schraudolph's construction: the float whose bit pattern is the integer
// (2^23 / ln 2) * x + (127 * 2^23 - 60801); a few per cent of relative error, monotonic.
static inline void volk_32f_expfast_32f(float* out, const float* in, unsigned int n) {
    const float fact = 8388608.0f / 0.6931471805f, bias = 1065353216.0f - 60801.0f;
    for (unsigned int i = 0; i < n; i++) {
        const int bits = (int)(fact * in[i] + bias);
        float f;
        memcpy(&f, &bits, sizeof(f));
        out[i] = f;
    }
}
static inline void volk_32f_accumulator_s32f(float* result, const float* in, unsigned int n) {
    float s = 0.0f;
    for (unsigned int i = 0; i < n; i++) { s += in[i]; }
    *result = s;
}
// float -> integer PCM as VOLK's generic kernels do it: scale, clamp to the integer type's range, round to nearest even
// (call site: utils/wav.cpp:161,165)
static inline void volk_32f_s32f_convert_16i(int16_t* out, const float* in, float scalar, unsigned int n) {
    const float lo = -32768.0f, hi = 32767.0f;
    for (unsigned int i = 0; i < n; i++) {
        float r = in[i] * scalar;
        if (r > hi) { r = hi; } else if (r < lo) { r = lo; }
        out[i] = (int16_t)rintf(r);
    }
}
static inline void volk_32f_s32f_convert_8i(int8_t* out, const float* in, float scalar, unsigned int n) {
    const float lo = -128.0f, hi = 127.0f;
    for (unsigned int i = 0; i < n; i++) {
        float r = in[i] * scalar;
        if (r > hi) { r = hi; } else if (r < lo) { r = lo; }
        out[i] = (int8_t)rintf(r);
    }
}
static inline void volk_8i_s32f_convert_32f(float* out, const int8_t* in, float scalar, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = (float)in[i] / scalar; }
}
static inline void volk_32f_s32f_convert_32i(int32_t* out, const float* in, float scalar, unsigned int n) {
    const float lo = -2147483648.0f, hi = 2147483647.0f;   // (float)INT_MAX is 2^31
    for (unsigned int i = 0; i < n; i++) {
        float r = in[i] * scalar;
        if (r >= hi) { out[i] = 2147483647; } else if (r < lo) { out[i] = (int32_t)(-2147483647 - 1); } else { out[i] = (int32_t)rintf(r); }
    }
}
static inline void volk_16i_s32f_convert_32f(float* out, const int16_t* in, float scalar, unsigned int n) {
    for (unsigned int i = 0; i < n; i++) { out[i] = (float)in[i] / scalar; }
}

// ---- power spectrum: 10*log10(|x/norm|^2) done as (10/log2(10)) * log2f(.)
// with log2 of zero/inf clamped to -/+127 (so an all-zero bin reads ~ -382.3 dB).
static inline float oracle_log2f_clamped(float v) {
    const float r = log2f(v);
    return std::isinf(r) ? copysignf(127.0f, r) : r;
}
static inline void volk_32fc_s32f_power_spectrum_32f(float* out, const lv_32fc_t* in, float norm, unsigned int n) {
    const float inv = 1.0f / norm;
    for (unsigned int i = 0; i < n; i++) {
        const float r = in[i].real() * inv;
        const float q = in[i].imag() * inv;
        out[i] = 3.01029995663981209120f * oracle_log2f_clamped(r * r + q * q);
    }
}
