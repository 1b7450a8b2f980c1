#include "spectrum.h"

namespace pg {

#include "cie_data.inc"

static const int kNCIE = 471;
static const float kCIEYIntegral = 106.856895f;  // spectrum.h:83

static inline float lerpf(float t, float a, float b) { return (1 - t) * a + t * b; }

RGB rgbFromXYZ(const float xyz[3]) {
    RGB r;
    xyzToRGB(xyz, r.c);
    return r;
}

// InterpolateSpectrumSamples, spectrum.cpp:179-188
static float interpolateSamples(const float *lambda, const float *vals, int n, float l) {
    if (l <= lambda[0]) return vals[0];
    if (l >= lambda[n - 1]) return vals[n - 1];
    int offset = findInterval(n, [&](int i) { return lambda[i] <= l; });
    float t = (l - lambda[offset]) / (lambda[offset + 1] - lambda[offset]);
    return lerpf(t, vals[offset], vals[offset + 1]);
}

RGB rgbFromSampled(const float *lambda, const float *v, int n) {
    // unsorted input is sorted by (lambda, value) pairs first (spectrum.cpp:41-57)
    bool sorted = true;
    for (int i = 0; i < n - 1; ++i)
        if (lambda[i] > lambda[i + 1]) sorted = false;
    if (!sorted) {
        std::vector<std::pair<float, float>> sv;
        for (int i = 0; i < n; ++i) sv.push_back({lambda[i], v[i]});
        std::sort(sv.begin(), sv.end());
        std::vector<float> sl(n), vv(n);
        for (int i = 0; i < n; ++i) sl[i] = sv[i].first, vv[i] = sv[i].second;
        return rgbFromSampled(sl.data(), vv.data(), n);
    }
    float xyz[3] = {0, 0, 0};
    for (int i = 0; i < kNCIE; ++i) {
        float val = interpolateSamples(lambda, v, n, float(360 + i));
        xyz[0] += val * kCIEX[i];
        xyz[1] += val * kCIEY[i];
        xyz[2] += val * kCIEZ[i];
    }
    float scale = float(float(360 + kNCIE - 1) - float(360)) / float(kCIEYIntegral * kNCIE);
    xyz[0] *= scale;
    xyz[1] *= scale;
    xyz[2] *= scale;
    return rgbFromXYZ(xyz);
}

// Blackbody / BlackbodyNormalized, spectrum.cpp:939-964.  The constants are
// doubles narrowed to float and the products mix float and double exactly as
// the reference's expressions do.
static void blackbody(const float *lambda, int n, float T, float *Le) {
    if (T <= 0) {
        for (int i = 0; i < n; ++i) Le[i] = 0.f;
        return;
    }
    const float c = 299792458;
    const float h = 6.62606957e-34;
    const float kb = 1.3806488e-23;
    for (int i = 0; i < n; ++i) {
        float l = lambda[i] * 1e-9;
        float lambda5 = (l * l) * (l * l) * l;
        Le[i] = (2 * h * c * c) / (lambda5 * (std::exp((h * c) / (l * kb * T)) - 1));
    }
}

RGB rgbFromBlackbody(float T, float scale) {
    std::vector<float> lam(kNCIE), v(kNCIE);
    for (int i = 0; i < kNCIE; ++i) lam[i] = float(360 + i);
    blackbody(lam.data(), kNCIE, T, v.data());
    float lambdaMax = 2.8977721e-3 / T * 1e9;
    float maxL;
    blackbody(&lambdaMax, 1, T, &maxL);
    for (int i = 0; i < kNCIE; ++i) v[i] /= maxL;
    return rgbFromSampled(lam.data(), v.data(), kNCIE) * scale;
}

}  // namespace pg
