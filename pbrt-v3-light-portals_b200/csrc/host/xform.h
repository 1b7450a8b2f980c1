// Host-only transform construction (Translate/Scale/Rotate/LookAt/Perspective,
// 4x4 product and Gauss-Jordan inverse) with the reference's operation order,
// so that world-space vertices and camera matrices come out bit-identical
// (src/core/transform.cpp:40-260, transform.h:86-93).
#ifndef PG_HOST_XFORM_H
#define PG_HOST_XFORM_H

#include <cmath>
#include <cstring>
#include <utility>

#include "../common/pg_math.h"
#include "params.h"
#include "scene.h"

namespace pg {

inline M44 m44Identity() {
    M44 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) r.m[i][j] = (i == j) ? 1.f : 0.f;
    return r;
}
inline M44 m44(float a00, float a01, float a02, float a03, float a10, float a11, float a12, float a13, float a20,
               float a21, float a22, float a23, float a30, float a31, float a32, float a33) {
    M44 r = {{{a00, a01, a02, a03}, {a10, a11, a12, a13}, {a20, a21, a22, a23}, {a30, a31, a32, a33}}};
    return r;
}
inline M44 m44Mul(const M44 &a, const M44 &b) {  // transform.h:86-93
    M44 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j] + a.m[i][3] * b.m[3][j];
    return r;
}
inline M44 m44Transpose(const M44 &a) {
    M44 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i];
    return r;
}
// transform.cpp:82-134 (full-pivot Gauss-Jordan; note the double 1./pivot)
inline M44 m44Inverse(const M44 &m) {
    int indxc[4], indxr[4];
    int ipiv[4] = {0, 0, 0, 0};
    float minv[4][4];
    memcpy(minv, m.m, 16 * sizeof(float));
    for (int i = 0; i < 4; i++) {
        int irow = 0, icol = 0;
        float big = 0.f;
        for (int j = 0; j < 4; j++) {
            if (ipiv[j] != 1) {
                for (int k = 0; k < 4; k++) {
                    if (ipiv[k] == 0) {
                        if (std::abs(minv[j][k]) >= big) {
                            big = float(std::abs(minv[j][k]));
                            irow = j;
                            icol = k;
                        }
                    } else if (ipiv[k] > 1)
                        hostError("Singular matrix in MatrixInvert");
                }
            }
        }
        ++ipiv[icol];
        if (irow != icol)
            for (int k = 0; k < 4; ++k) std::swap(minv[irow][k], minv[icol][k]);
        indxr[i] = irow;
        indxc[i] = icol;
        if (minv[icol][icol] == 0.f) hostError("Singular matrix in MatrixInvert");
        float pivinv = 1. / minv[icol][icol];
        minv[icol][icol] = 1.;
        for (int j = 0; j < 4; j++) minv[icol][j] *= pivinv;
        for (int j = 0; j < 4; j++) {
            if (j != icol) {
                float save = minv[j][icol];
                minv[j][icol] = 0;
                for (int k = 0; k < 4; k++) minv[j][k] -= minv[icol][k] * save;
            }
        }
    }
    for (int j = 3; j >= 0; j--) {
        if (indxr[j] != indxc[j])
            for (int k = 0; k < 4; k++) std::swap(minv[k][indxr[j]], minv[k][indxc[j]]);
    }
    M44 r;
    memcpy(r.m, minv, 16 * sizeof(float));
    return r;
}

inline Xform xfIdentity() { return Xform{m44Identity(), m44Identity()}; }
inline Xform xfFromMatrix(const M44 &m) { return Xform{m, m44Inverse(m)}; }
inline Xform xfInverse(const Xform &t) { return Xform{t.mInv, t.m}; }
inline Xform xfMul(const Xform &a, const Xform &b) {  // transform.cpp:254-256
    return Xform{m44Mul(a.m, b.m), m44Mul(b.mInv, a.mInv)};
}
inline Xform xfTranslate(V3 d) {
    return Xform{m44(1, 0, 0, d.x, 0, 1, 0, d.y, 0, 0, 1, d.z, 0, 0, 0, 1),
                 m44(1, 0, 0, -d.x, 0, 1, 0, -d.y, 0, 0, 1, -d.z, 0, 0, 0, 1)};
}
inline Xform xfScale(float x, float y, float z) {
    return Xform{m44(x, 0, 0, 0, 0, y, 0, 0, 0, 0, z, 0, 0, 0, 0, 1),
                 m44(1 / x, 0, 0, 0, 0, 1 / y, 0, 0, 0, 0, 1 / z, 0, 0, 0, 0, 1)};
}
inline Xform xfRotate(float theta, V3 axis) {  // transform.cpp:186-206
    V3 a = normalize(axis);
    float sinTheta = std::sin(radians(theta));
    float cosTheta = std::cos(radians(theta));
    M44 m = m44Identity();
    m.m[0][0] = a.x * a.x + (1 - a.x * a.x) * cosTheta;
    m.m[0][1] = a.x * a.y * (1 - cosTheta) - a.z * sinTheta;
    m.m[0][2] = a.x * a.z * (1 - cosTheta) + a.y * sinTheta;
    m.m[0][3] = 0;
    m.m[1][0] = a.x * a.y * (1 - cosTheta) + a.z * sinTheta;
    m.m[1][1] = a.y * a.y + (1 - a.y * a.y) * cosTheta;
    m.m[1][2] = a.y * a.z * (1 - cosTheta) - a.x * sinTheta;
    m.m[1][3] = 0;
    m.m[2][0] = a.x * a.z * (1 - cosTheta) - a.y * sinTheta;
    m.m[2][1] = a.y * a.z * (1 - cosTheta) + a.x * sinTheta;
    m.m[2][2] = a.z * a.z + (1 - a.z * a.z) * cosTheta;
    m.m[2][3] = 0;
    return Xform{m, m44Transpose(m)};
}
inline Xform xfLookAt(V3 pos, V3 look, V3 up) {  // transform.cpp:208-241
    M44 c2w = m44Identity();
    c2w.m[0][3] = pos.x;
    c2w.m[1][3] = pos.y;
    c2w.m[2][3] = pos.z;
    c2w.m[3][3] = 1;
    V3 dir = normalize(look - pos);
    if (length(cross(normalize(up), dir)) == 0) {
        hostError("\"up\" vector (%f, %f, %f) and viewing direction (%f, %f, %f) passed to LookAt are pointing in "
                  "the same direction.  Using the identity transformation.",
                  up.x, up.y, up.z, dir.x, dir.y, dir.z);
        return xfIdentity();
    }
    V3 right = normalize(cross(normalize(up), dir));
    V3 newUp = cross(dir, right);
    c2w.m[0][0] = right.x; c2w.m[1][0] = right.y; c2w.m[2][0] = right.z; c2w.m[3][0] = 0.;
    c2w.m[0][1] = newUp.x; c2w.m[1][1] = newUp.y; c2w.m[2][1] = newUp.z; c2w.m[3][1] = 0.;
    c2w.m[0][2] = dir.x; c2w.m[1][2] = dir.y; c2w.m[2][2] = dir.z; c2w.m[3][2] = 0.;
    return Xform{m44Inverse(c2w), c2w};
}
inline Xform xfPerspective(float fov, float n, float f) {  // transform.cpp:1097-1106
    M44 persp = m44(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, f / (f - n), -f * n / (f - n), 0, 0, 1, 0);
    float invTanAng = 1 / std::tan(radians(fov) / 2);
    return xfMul(xfScale(invTanAng, invTanAng, 1), xfFromMatrix(persp));
}
inline bool xfSwapsHandedness(const Xform &t) {  // transform.cpp:258-263
    const M44 &m = t.m;
    float det = m.m[0][0] * (m.m[1][1] * m.m[2][2] - m.m[1][2] * m.m[2][1]) -
                m.m[0][1] * (m.m[1][0] * m.m[2][2] - m.m[1][2] * m.m[2][0]) +
                m.m[0][2] * (m.m[1][0] * m.m[2][1] - m.m[1][1] * m.m[2][0]);
    return det < 0;
}
inline Bounds3 xfBounds(const Xform &t, const Bounds3 &b) {  // transform.cpp:243-252
    const M44 &M = t.m;
    Bounds3 ret;
    V3 p0 = xformPoint(M, v3(b.pMin.x, b.pMin.y, b.pMin.z));
    ret.pMin = ret.pMax = p0;
    ret = unionB(ret, xformPoint(M, v3(b.pMax.x, b.pMin.y, b.pMin.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMin.x, b.pMax.y, b.pMin.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMin.x, b.pMin.y, b.pMax.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMin.x, b.pMax.y, b.pMax.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMax.x, b.pMax.y, b.pMin.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMax.x, b.pMin.y, b.pMax.z)));
    ret = unionB(ret, xformPoint(M, v3(b.pMax.x, b.pMax.y, b.pMax.z)));
    return ret;
}

}  // namespace pg
#endif
