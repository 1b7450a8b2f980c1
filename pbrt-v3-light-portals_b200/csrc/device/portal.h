// Portal-restricted sampling of a PortalArealight.
//   PortalArealight::EstimateDirect (selection)   lights/portal_arealight.cpp:29-113
//   AAPortal::InFront / SamplePortal / SampleProj portals/aaportal.cpp:73-83, 97-99, 114-159
//   AAPlaneShape::Sample / Normal / InFront       shapes/plane.cpp:57-83, 109-115
// Behaviours that look like bugs are kept, because parity is against what the
// reference computes: InFrustum is always true; the
// projection sampler uses u.x for both in-plane axes, mixes world and object
// coordinates, returns an un-normalised wi and can return a negative pdf
// (rejected by the caller's `pdf > 0`); SampleUniformPortal is not divided by
// the portal-selection probability.  The one deliberate difference: the
// selected portal is a per-path local instead of the shared mutable member
// `selectedPortal`, which removes the data race (Q7) without changing what a
// single thread computes.
#ifndef PG_DEVICE_PORTAL_H
#define PG_DEVICE_PORTAL_H

#include "lights.h"

namespace pg {

struct PortalRec {  // AAPortal::portal, an AAPlaneShape built with the light's transforms
    PlaneParams pl;
};
PG_HD PortalRec loadPortal(const float *rec) {
    PortalRec p;
    p.pl = loadPlane(rec);
    return p;
}

// AAPortal::InFront
PG_HD bool portalInFront(const PortalRec &po, const M44 &lightW2O, V3 p) {
    V3 q = xformPoint(lightW2O, p);
    return po.pl.facingFw ? (q[po.pl.ax] > po.pl.lo[po.pl.ax]) : (q[po.pl.ax] < po.pl.lo[po.pl.ax]);
}

// Portal choice.  Returns: -2 = behind every portal (fall back to light
// sampling), -1 = contributes nothing, else the selected portal and its pdf.
PG_HD_NOINLINE int portalSelect(const float *portals, int nPortals, const M44 &lightW2O, V3 p, float u, float *portalPdf) {
    bool behindAll = true;
    float sum = 0;
    for (int i = 0; i < nPortals; ++i)
        if (portalInFront(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i), lightW2O, p)) {
            behindAll = false;
            sum += 1;
        }
    if (behindAll) return -2;
    if (sum == 0) return -1;
    // Distribution1D over dist[i] = (inFront ? 1 : 0) / sum, then SampleDiscrete(u)
    float cdf = 0, funcInt = 0;
    for (int i = 0; i < nPortals; ++i) {
        float d = portalInFront(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i), lightW2O, p) ? 1.f / sum : 0.f;
        funcInt = funcInt + d / nPortals;
    }
    int offset = 0;
    float selFunc = 0;
    // FindInterval: the last i with cdf[i] <= u, clamped to [0, n-1]
    for (int i = 0; i < nPortals; ++i) {
        float d = portalInFront(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i), lightW2O, p) ? 1.f / sum : 0.f;
        float cdfI = funcInt == 0 ? (float)i / (float)nPortals : cdf / funcInt;
        if (cdfI <= u) {
            offset = i;
            selFunc = d;
        }
        cdf = cdf + d / nPortals;
    }
    *portalPdf = (funcInt > 0) ? selFunc / (funcInt * nPortals) : 0;
    return offset;
}

// AAPortal::SamplePortal
PG_HD void portalSampleUniform(const PortalRec &po, const M44 &lightO2W, V3 refP, float u0, float u1, V3 *wi, float *pdf) {
    V3 loW = xformPoint(lightO2W, po.pl.lo), hiW = xformPoint(lightO2W, po.pl.hi);
    V3 sp = v3(0, 0, 0);
    sp.set(po.pl.ax, loW[po.pl.ax]);
    sp.set(po.pl.ax0, loW[po.pl.ax0] + (hiW[po.pl.ax0] - loW[po.pl.ax0]) * u0);
    sp.set(po.pl.ax1, loW[po.pl.ax1] + (hiW[po.pl.ax1] - loW[po.pl.ax1]) * u1);
    *wi = normalize(sp - refP);
    *pdf = distanceSquared(refP, sp) / (absDot(aaplaneNormal(po.pl), -*wi) * po.pl.area);
}

// AAPortal::SampleProj
PG_HD_NOINLINE void portalSampleProj(const PortalRec &po, const PlaneParams &light, const M44 &lightW2O, V3 refP, float u0, V3 *wi,
                            float *pdf) {
    V3 dLo = normalize(refP - light.lo);
    V3 dHi = normalize(refP - light.hi);
    if (dLo.z == 0 || dHi.z == 0) {
        *pdf = 0;
        return;
    }
    float tLo = (po.pl.lo[po.pl.ax] - light.lo[light.ax]) / dLo[light.ax];
    float tHi = (po.pl.lo[po.pl.ax] - light.hi[light.ax]) / dHi[light.ax];
    V3 projLo = light.lo + dLo * tLo;
    V3 projHi = light.hi + dHi * tHi;
    V3 isectHi = vmax(po.pl.lo, projLo);
    V3 isectLo = vmin(po.pl.hi, projHi);
    float isectLen0 = isectHi[po.pl.ax0] - isectLo[po.pl.ax0];
    float isectLen1 = isectHi[po.pl.ax1] - isectLo[po.pl.ax1];
    V3 sampled = v3(0, 0, 0);
    sampled.set(po.pl.ax, po.pl.lo[po.pl.ax]);
    sampled.set(po.pl.ax0, isectLo[po.pl.ax0] + u0 * isectLen0);
    sampled.set(po.pl.ax1, isectLo[po.pl.ax1] + u0 * isectLen1);
    V3 sampledWorld = xformPoint(lightW2O, sampled);
    *wi = sampledWorld - refP;
    *pdf = distanceSquared(refP, sampled) / (absDot(aaplaneNormal(po.pl), -*wi) * (isectLen0 * isectLen1));
}

// ---- Corrected portal sampling (NOT in the reference; strategies "solidangle" and "solidangle_mis").
// what the fork's TODOs point at (aaportal.cpp:162 "TODO implement for MIS", the
// commented-out MIS estimators of integrator.cpp:263-513).  The portal rectangle is sampled uniformly in
// SOLID ANGLE as seen from the shading point (Urena, Fajardo, King: "An Area-Preserving Parametrization for
// Spherical Rectangles", EGSR 2013), so the estimator has no distance^2 / cosine factor, and the portal is
// picked with probability proportional to its solid angle, which the estimate IS divided by.
struct SphRect {
    V3 o, x, y, z;
    float z0, x0, y0, x1, y1;
    float b0, b1, k, S;  // S = solid angle
};
PG_HD float safeAcos(float v) { return acosf(clampf(v, -1.f, 1.f)); }
// the world-space rectangle of a portal: corner and the two edge vectors
PG_HD void portalWorldRect(const PortalRec &po, const M44 &lightO2W, V3 *corner, V3 *ex, V3 *ey) {
    V3 lo = po.pl.lo, a = po.pl.lo, b = po.pl.lo;
    a.set(po.pl.ax0, po.pl.hi[po.pl.ax0]);
    b.set(po.pl.ax1, po.pl.hi[po.pl.ax1]);
    *corner = xformPoint(lightO2W, lo);
    *ex = xformPoint(lightO2W, a) - *corner;
    *ey = xformPoint(lightO2W, b) - *corner;
}
PG_HD SphRect sphRectInit(V3 s, V3 ex, V3 ey, V3 o) {
    SphRect q;
    q.o = o;
    float exl = length(ex), eyl = length(ey);
    q.x = ex / exl;
    q.y = ey / eyl;
    q.z = cross(q.x, q.y);
    V3 d = s - o;
    q.z0 = dot(d, q.z);
    if (q.z0 > 0) {
        q.z = -q.z;
        q.z0 = -q.z0;
    }
    q.x0 = dot(d, q.x);
    q.y0 = dot(d, q.y);
    q.x1 = q.x0 + exl;
    q.y1 = q.y0 + eyl;
    V3 v00 = v3(q.x0, q.y0, q.z0), v01 = v3(q.x0, q.y1, q.z0), v10 = v3(q.x1, q.y0, q.z0), v11 = v3(q.x1, q.y1, q.z0);
    V3 n0 = normalize(cross(v00, v10)), n1 = normalize(cross(v10, v11)), n2 = normalize(cross(v11, v01)), n3 = normalize(cross(v01, v00));
    float g0 = safeAcos(-dot(n0, n1)), g1 = safeAcos(-dot(n1, n2)), g2 = safeAcos(-dot(n2, n3)), g3 = safeAcos(-dot(n3, n0));
    q.b0 = n0.z;
    q.b1 = n2.z;
    q.k = 2 * kPi - g2 - g3;
    q.S = g0 + g1 - q.k;
    if (!(q.z0 < 0) || !(q.S > 0)) q.S = 0;  // the point lies in the rectangle's plane (or numerics gave up): nothing to sample
    return q;
}
PG_HD V3 sphRectSample(const SphRect &q, float u, float v) {
    float au = u * q.S + q.k;
    float fu = (cosf(au) * q.b0 - q.b1) / sinf(au);
    float cu = 1 / sqrtf(fu * fu + q.b0 * q.b0) * (fu > 0 ? 1.f : -1.f);
    cu = clampf(cu, -1.f, 1.f);
    float xu = -(cu * q.z0) / sqrtf(maxf(1e-12f, 1 - cu * cu));
    xu = clampf(xu, q.x0, q.x1);
    float d = sqrtf(xu * xu + q.z0 * q.z0);
    float h0 = q.y0 / sqrtf(d * d + q.y0 * q.y0), h1 = q.y1 / sqrtf(d * d + q.y1 * q.y1);
    float hv = h0 + v * (h1 - h0), hv2 = hv * hv;
    float yv = (hv2 < 1 - 1e-6f) ? (hv * d) / sqrtf(1 - hv2) : q.y1;
    yv = clampf(yv, q.y0, q.y1);
    return q.o + xu * q.x + yv * q.y + q.z0 * q.z;
}
// does the direction wi from o pass through the rectangle?
PG_HD bool rectContainsDirection(V3 s, V3 ex, V3 ey, V3 o, V3 wi) {
    V3 n = cross(ex, ey);
    float denom = dot(n, wi);
    if (denom == 0) return false;
    float t = dot(n, s - o) / denom;
    if (!(t > 0)) return false;
    V3 h = o + t * wi - s;
    float a = dot(h, ex) / dot(ex, ex), b = dot(h, ey) / dot(ey, ey);
    return a >= 0 && a <= 1 && b >= 0 && b <= 1;
}

// AAPortal's viewing frustum (aaportal.cpp:14-57, 101-112): four planes, each through one edge of the portal and the
// OPPOSITE edge of the light (spanned by the two diagonals corner-of-portal minus opposite-corner-of-light that end on
// that edge) -- outside it no point of the light is seen through the portal.  The reference builds the planes and then
// disables the test (`return true;`, Q1).  It is used here only by the strategy "solidangle_mis3", as a cull of the
// portal TECHNIQUE inside a multiple-importance combination whose other two techniques cover every direction: a portal
// that fails the test is neither sampled nor counted in the portal pdf, so the test cannot bias the estimate, it only
// stops samples from being spent on portals that cannot show the light.  Everything is evaluated in the light's object
// space (the reference mixes the portal's object-space corners with a world-space point; its scenes' transforms are
// identities).
PG_HD V3 aaplaneCorner(const PlaneParams &pl, int k) {  // AAPlaneShape::V0..V3, plane.cpp:85-107
    V3 v = pl.lo;
    if (k == 1 || k == 2) v.set(pl.ax1, pl.hi[pl.ax1]);
    if (k == 2 || k == 3) v.set(pl.ax0, pl.hi[pl.ax0]);
    if (k == 2) v = pl.hi;
    return v;
}
PG_HD_NOINLINE bool portalInFrustum(const PortalRec &po, const PlaneParams &light, const M44 &lightW2O, V3 p) {
    V3 q = xformPoint(lightW2O, p);
    V3 fd[4], pc[4];
    for (int k = 0; k < 4; ++k) {
        pc[k] = aaplaneCorner(po.pl, k);
        fd[k] = normalize(pc[k] - aaplaneCorner(light, (k + 2) & 3));
    }
    bool res = true;
    for (int k = 0; k < 4; ++k) {
        int k1 = (k + 1) & 3;
        V3 fn = po.pl.facingFw ? cross(fd[k], fd[k1]) : cross(fd[k1], fd[k]);
        V3 fp = (pc[k] + pc[k1]) / 2;
        res &= dot(fp - q, fn) >= 0;
    }
    return res;
}

// Solid angles of the portals in front of p; returns their sum.  With `pick` in [0, 1): *sel = the portal drawn
// with probability omega_i / sum and *rect = its spherical rectangle.
// `frustumOf` != nullptr (the light's own plane): portals whose viewing frustum does not hold p are left out as well.
PG_HD_NOINLINE float portalSolidAngles(const float *portals, int nPortals, const M44 &lightO2W, const M44 &lightW2O, V3 p, float pick, int *sel,
                                       SphRect *rect, const PlaneParams *frustumOf = nullptr) {
    float sum = 0;
    for (int i = 0; i < nPortals; ++i) {
        PortalRec po = loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i);
        if (!portalInFront(po, lightW2O, p)) continue;
        if (frustumOf && !portalInFrustum(po, *frustumOf, lightW2O, p)) continue;
        V3 s, ex, ey;
        portalWorldRect(po, lightO2W, &s, &ex, &ey);
        sum += sphRectInit(s, ex, ey, p).S;
    }
    *sel = -1;
    if (!(sum > 0)) return 0;
    float target = pick * sum, acc = 0;
    for (int i = 0; i < nPortals; ++i) {
        PortalRec po = loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i);
        if (!portalInFront(po, lightW2O, p)) continue;
        if (frustumOf && !portalInFrustum(po, *frustumOf, lightW2O, p)) continue;
        V3 s, ex, ey;
        portalWorldRect(po, lightO2W, &s, &ex, &ey);
        SphRect q = sphRectInit(s, ex, ey, p);
        if (!(q.S > 0)) continue;
        *sel = i;
        *rect = q;
        acc += q.S;
        if (target < acc) break;
    }
    return sum;
}
// is wi (from p) inside one of the portals in front of p?
PG_HD bool portalsContainDirection(const float *portals, int nPortals, const M44 &lightO2W, const M44 &lightW2O, V3 p, V3 wi,
                                   const PlaneParams *frustumOf = nullptr) {
    for (int i = 0; i < nPortals; ++i) {
        PortalRec po = loadPortal(portals + PBRTGPU_PORTAL_STRIDE * i);
        if (!portalInFront(po, lightW2O, p)) continue;
        if (frustumOf && !portalInFrustum(po, *frustumOf, lightW2O, p)) continue;
        V3 s, ex, ey;
        portalWorldRect(po, lightO2W, &s, &ex, &ey);
        if (rectContainsDirection(s, ex, ey, p, wi)) return true;
    }
    return false;
}

}  // namespace pg
#endif
