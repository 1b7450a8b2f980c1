// RGB spectrum arithmetic used while reading scene parameters.  The renderer
// works in RGB (the reference's default build: `typedef RGBSpectrum Spectrum`,
// src/core/pbrt.h:128-131); spectral parameter data is converted once, at
// parse time, the way RGBSpectrum::FromSampled does (src/core/spectrum.h:481-502).
#ifndef PG_HOST_SPECTRUM_H
#define PG_HOST_SPECTRUM_H

#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

namespace pg {

struct RGB {
    float c[3] = {0, 0, 0};
    RGB() {}
    explicit RGB(float v) { c[0] = c[1] = c[2] = v; }
    RGB(float r, float g, float b) { c[0] = r; c[1] = g; c[2] = b; }
    float &operator[](int i) { return c[i]; }
    float operator[](int i) const { return c[i]; }
    bool isBlack() const { return c[0] == 0 && c[1] == 0 && c[2] == 0; }
    RGB operator*(const RGB &o) const { return RGB(c[0] * o.c[0], c[1] * o.c[1], c[2] * o.c[2]); }
    RGB operator*(float s) const { return RGB(c[0] * s, c[1] * s, c[2] * s); }
    // CoefficientSpectrum::Clamp (spectrum.h:250-257)
    RGB clamp(float lo = 0, float hi = INFINITY) const {
        RGB r;
        for (int i = 0; i < 3; ++i) r.c[i] = c[i] < lo ? lo : (c[i] > hi ? hi : c[i]);
        return r;
    }
    // RGBSpectrum::y (spectrum.h:476-480)
    float y() const { return 0.212671f * c[0] + 0.715160f * c[1] + 0.072169f * c[2]; }
};

// spectrum.h:62-76
inline void xyzToRGB(const float xyz[3], float rgb[3]) {
    rgb[0] = 3.240479f * xyz[0] - 1.537150f * xyz[1] - 0.498535f * xyz[2];
    rgb[1] = -0.969256f * xyz[0] + 1.875991f * xyz[1] + 0.041556f * xyz[2];
    rgb[2] = 0.055648f * xyz[0] - 0.204043f * xyz[1] + 1.057311f * xyz[2];
}
inline void rgbToXYZ(const float rgb[3], float xyz[3]) {
    xyz[0] = 0.412453f * rgb[0] + 0.357580f * rgb[1] + 0.180423f * rgb[2];
    xyz[1] = 0.212671f * rgb[0] + 0.715160f * rgb[1] + 0.072169f * rgb[2];
    xyz[2] = 0.019334f * rgb[0] + 0.119193f * rgb[1] + 0.950227f * rgb[2];
}

RGB rgbFromXYZ(const float xyz[3]);
// (lambda, value) samples -> RGB through the CIE matching functions.
RGB rgbFromSampled(const float *lambda, const float *v, int n);
// blackbody (T kelvin, scale) -> RGB, ParamSet::AddBlackbodySpectrum (paramset.cpp:134-150)
RGB rgbFromBlackbody(float T, float scale);

// pbrt.h FindInterval: largest index with pred true, clamped to [0, size-2].
template <typename Pred>
int findInterval(int size, const Pred &pred) {
    int first = 0, len = size;
    while (len > 0) {
        int half = len >> 1, middle = first + half;
        if (pred(middle)) {
            first = middle + 1;
            len -= half + 1;
        } else
            len = half;
    }
    return std::min(std::max(first - 1, 0), size - 2);
}

}  // namespace pg
#endif
