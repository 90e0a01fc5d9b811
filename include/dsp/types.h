// Sample types of the drop-in DSP headers (layout-compatible with the reference's dsp::complex_t /
// dsp::stereo_t, core/src/dsp/types.h:6-94: two packed floats).
#pragma once
#include <cmath>

namespace dsp {
    struct complex_t {
        float re, im;
        complex_t operator+(const complex_t& b) const { return { re + b.re, im + b.im }; }
        complex_t operator-(const complex_t& b) const { return { re - b.re, im - b.im }; }
        complex_t operator*(float b) const { return { re * b, im * b }; }
        complex_t operator*(const complex_t& b) const { return { re * b.re - im * b.im, im * b.re + re * b.im }; }
        complex_t conj() const { return { re, -im }; }
        float amplitude() const { return sqrtf(re * re + im * im); }
        float phase() const { return atan2f(im, re); }
    };
    struct stereo_t {
        float l, r;
    };
}
