// Camera-ray generation and RGB spectrum helpers.
//   Sampler::GetCameraSample                    core/sampler.cpp:46-53
//   PerspectiveCamera::GenerateRayDifferential  cameras/perspective.cpp:100-153
//   Camera::GenerateWvls                        core/camera.cpp:62-76
//   Transform::operator()(Ray)                  core/transform.h:383-400
//   ConcentricSampleDisk                        core/sampling.cpp:113-130
// Ray differentials only feed texture filtering; every texture on this path is
// constant, so they are not generated (not observable in the image).
#ifndef PG_DEVICE_CAMERA_H
#define PG_DEVICE_CAMERA_H

#include "sampler.h"
#include "shapes.h"

namespace pg {

// ---- RGBSpectrum (core/spectrum.h:130-230, 478-481).  A type of its own:
// Spectrum / Float divides each channel, while Vector3f / Float multiplies by
// the rounded reciprocal (geometry.h:244-248), and the two round differently.
struct Spec {
    float r, g, b;
};
PG_HD Spec spec(float r, float g, float b) { Spec s; s.r = r; s.g = g; s.b = b; return s; }
PG_HD Spec spec(float v) { return spec(v, v, v); }
PG_HD Spec loadSpec(const float *p) { return spec(PG_LD(p), PG_LD(p + 1), PG_LD(p + 2)); }
PG_HD Spec operator+(Spec a, Spec b) { return spec(a.r + b.r, a.g + b.g, a.b + b.b); }
PG_HD Spec operator-(Spec a, Spec b) { return spec(a.r - b.r, a.g - b.g, a.b - b.b); }
PG_HD Spec operator-(Spec a) { return spec(-a.r, -a.g, -a.b); }
PG_HD Spec clampSpec(Spec a) {  // CoefficientSpectrum::Clamp(0, Infinity), spectrum.h:244-251
    return spec(a.r < 0.f ? 0.f : a.r, a.g < 0.f ? 0.f : a.g, a.b < 0.f ? 0.f : a.b);
}
PG_HD Spec operator*(Spec a, Spec b) { return spec(a.r * b.r, a.g * b.g, a.b * b.b); }
PG_HD Spec operator*(Spec a, float s) { return spec(a.r * s, a.g * s, a.b * s); }
PG_HD Spec operator*(float s, Spec a) { return spec(a.r * s, a.g * s, a.b * s); }
PG_HD Spec operator/(Spec a, float s) { return spec(a.r / s, a.g / s, a.b / s); }
PG_HD Spec operator/(Spec a, Spec b) { return spec(a.r / b.r, a.g / b.g, a.b / b.b); }
PG_HD Spec &operator+=(Spec &a, Spec b) { a = a + b; return a; }
PG_HD Spec &operator*=(Spec &a, Spec b) { a = a * b; return a; }
PG_HD Spec &operator*=(Spec &a, float s) { a = a * s; return a; }
PG_HD Spec &operator/=(Spec &a, float s) { a = a / s; return a; }
PG_HD bool isBlack(Spec s) { return s.r == 0.f && s.g == 0.f && s.b == 0.f; }
PG_HD float luminance(Spec s) { return 0.212671f * s.r + 0.715160f * s.g + 0.072169f * s.b; }
PG_HD float maxComponentValue(Spec s) { return maxf(s.r, maxf(s.g, s.b)); }
PG_HD bool hasNaNs(Spec s) { return s.r != s.r || s.g != s.g || s.b != s.b; }
PG_HD void rgbToXYZ(const float rgb[3], float xyz[3]) {  // spectrum.h:64-68
    xyz[0] = 0.412453f * rgb[0] + 0.357580f * rgb[1] + 0.180423f * rgb[2];
    xyz[1] = 0.212671f * rgb[0] + 0.715160f * rgb[1] + 0.072169f * rgb[2];
    xyz[2] = 0.019334f * rgb[0] + 0.119193f * rgb[1] + 0.950227f * rgb[2];
}

PG_HD void concentricSampleDisk(float u0, float u1, float *dx, float *dy) {
    float ox = 2.f * u0 - 1, oy = 2.f * u1 - 1;
    if (ox == 0 && oy == 0) {
        *dx = 0, *dy = 0;
        return;
    }
    float theta, r;
    if (fabsf(ox) > fabsf(oy)) {
        r = ox;
        theta = kPiOver4 * (oy / ox);
    } else {
        r = oy;
        theta = kPiOver2 - kPiOver4 * (ox / oy);
    }
    *dx = r * cosf(theta);
    *dy = r * sinf(theta);
}

struct CameraParams {
    M44 cameraToWorld, rasterToCamera;
    float lensRadius, focalDistance;
};

struct CameraSample {
    float pFilmX, pFilmY, time, pLensX, pLensY, wvl;
};

// GetCameraSample: 6 dimensions, in this order.
// needTimeAndLens = false: the camera is a pinhole and nothing moves, so the time and lens samples (dimensions
// 2, 3, 4) have no reader; the sampler's dimension still advances past them, their values are not computed.
// needWavelength = false: no material of the scene is dispersive, so the hero wavelength (dimension 5) has no reader either.
PG_HD CameraSample getCameraSample(const HaltonParams &h, PathSampler &s, int px, int py, bool needTimeAndLens = true, bool needWavelength = true) {
    CameraSample cs;
    float u0, u1;
    get2D(h, s, &u0, &u1);
    cs.pFilmX = (float)px + u0;
    cs.pFilmY = (float)py + u1;
    if (needTimeAndLens) {
        cs.time = get1D(h, s);
        get2D(h, s, &cs.pLensX, &cs.pLensY);
    } else {
        cs.time = cs.pLensX = cs.pLensY = 0;
        s.dim += 3;  // dimensions 2..4 lie below GlobalSampler's array dimensions: no jump to account for
    }
    if (needWavelength) cs.wvl = get1D(h, s);
    else {
        if (s.dim >= kArrayStartDim && s.dim < h.arrayEndDim) s.dim = h.arrayEndDim;  // GlobalSampler::Get1D's jump over the sample arrays
        ++s.dim;
        cs.wvl = 0;
    }
    return cs;
}

// Returns the world-space ray; *wvl0 = the hero wavelength wvls[0].
PG_HD Ray generateCameraRay(const CameraParams &c, const CameraSample &cs, float *wvl0) {
    V3 pCamera = xformPoint(c.rasterToCamera, v3(cs.pFilmX, cs.pFilmY, 0));
    Ray ray;
    ray.o = v3(0, 0, 0);
    ray.d = normalize(pCamera);
    if (c.lensRadius > 0) {
        float lx, ly;
        concentricSampleDisk(cs.pLensX, cs.pLensY, &lx, &ly);
        lx *= c.lensRadius, ly *= c.lensRadius;
        float ft = c.focalDistance / ray.d.z;
        V3 pFocus = ray.o + ray.d * ft;
        ray.o = v3(lx, ly, 0);
        ray.d = normalize(pFocus - ray.o);
    }
    *wvl0 = 400.f + 300.f * cs.wvl;  // SampleUniformSpectrum, spectrum.h:528-530
    // CameraToWorld(ray): Transform::operator()(const Ray &) offsets the origin
    // along d by the transform's rounding error bound (transform.h:383-400).
    V3 oError;
    V3 o = xformPointErr(c.cameraToWorld, ray.o, &oError);
    V3 d = xformVector(c.cameraToWorld, ray.d);
    float lengthSquared_ = lengthSquared(d);
    float tMax = kInfinity;
    if (lengthSquared_ > 0) {
        float dt = dot(vabs(d), oError) / lengthSquared_;
        o += d * dt;
        tMax -= dt;
    }
    ray.o = o;
    ray.d = d;
    ray.tMax = tMax;
    return ray;
}

}  // namespace pg
#endif
