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

}  // namespace pg
#endif
