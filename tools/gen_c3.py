"""BASELINE.json's configuration C3 as SURVEY.md 8(d) states it: the synthetic "skylight room".

    python tools/gen_c3.py OUT.pbrt [--strategy light|portal|projection|...] [--spp 16] [--res 512] [--window ceiling|wall_x|wall_y]

A matte room [-2,2] x [-2,2] x [0,3] (Kd 0.5, 12 wall/ceiling quads), floor clutter of 20 000 triangles from a
seeded height field (seed 7, a 100 x 100-quad grid over the floor, smooth hills up to 0.45 high, a third of the
patches plastic), a 1.5 x 1.5 hole, an `aaplane` emitter outside under `AreaLightSource "portal"` (L = 10,
twosided) with one AA portal in the hole, perspective camera inside, box filter, Halton, Integrator "path"
maxdepth 5.  Deterministic: the text depends only on the arguments.

window = ceiling   hole in the ceiling, emitter [-3,3]^2 at z = 6, portal axis 2 (the configuration SURVEY names first)
window = wall_x    hole in the wall x = 2, emitter at x = 5, portal axis 0
window = wall_y    hole in the wall y = 2, emitter at y = 5, portal axis 1
The wall variants exercise AAPlaneShape's quirk: Intersect hard-codes dpdu / dpdv, so the hit
normal is +-z for every axis and a wall window emits towards the room only because the light is two-sided.
The portal sits 1 mm outside the wall it pierces: coplanar, AAPortal::InFront of a point ON that wall would be
decided by the last bit of the interpolated hit point (in the reference as well).
"""
import argparse

import numpy as np


def _quad(p):
    return 'Shape "trianglemesh" "integer indices" [0 1 2 2 3 0] "point P" [%s]\n' % " ".join("%g" % v for v in p)


def _clutter(seed=7, n=100):
    """n x n quads over the floor: 2 n^2 triangles, split into a matte and a plastic mesh by 10 x 10-quad patch."""
    rng = np.random.default_rng(seed)
    coarse = rng.random((12, 12))
    u = np.linspace(0, 10, n + 1)
    i0 = np.floor(u).astype(int)
    f = u - i0
    s = f * f * (3 - 2 * f)
    rows = coarse[i0][:, None, :] * (1 - s)[:, None, None] + coarse[i0 + 1][:, None, :] * s[:, None, None]  # (n+1, 1, 12) along y
    h = rows[:, 0, :][:, i0] * (1 - s)[None, :] + rows[:, 0, :][:, i0 + 1] * s[None, :]                       # (n+1, n+1)
    h = 0.45 * h + 0.03 * rng.random((n + 1, n + 1))
    xs = np.linspace(-1.98, 1.98, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    verts = np.stack([X, Y, h], -1).reshape(-1, 3).astype(np.float32)
    j, i = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    v00 = (j * (n + 1) + i).ravel()
    tris = np.stack([np.stack([v00, v00 + 1, v00 + n + 2], -1), np.stack([v00, v00 + n + 2, v00 + n + 1], -1)], 1)
    plastic = rng.random((n // 10, n // 10))[(j // 10).ravel(), (i // 10).ravel()] < 1 / 3
    out = []
    for sel in (~plastic, plastic):
        fidx = tris[sel].reshape(-1, 3)
        used, inv = np.unique(fidx, return_inverse=True)
        out.append((verts[used], inv.reshape(-1, 3)))
    return out


def scene_text(strategy="portal", spp=16, res=512, window="ceiling", maxdepth=5):
    t = []
    t.append('LookAt 1.6 -1.7 1.4  -0.2 0.3 0.9  0 0 1\nCamera "perspective" "float fov" [70]\n')
    t.append('Film "image" "string filename" ["c3.pfm"] "integer xresolution" [%d] "integer yresolution" [%d]\n' % (res, res))
    t.append('Sampler "halton" "integer pixelsamples" [%d]\nIntegrator "path" "integer maxdepth" [%d]\nWorldBegin\n' % (spp, maxdepth))
    e = 0.001
    if window == "ceiling":
        portal, emitter, axis = (-0.75, -0.75, 3 + e, 0.75, 0.75, 3 + e), ((-3, -3, 6), (3, 3, 6)), 2
    elif window == "wall_x":
        portal, emitter, axis = (2 + e, -0.75, 1, 2 + e, 0.75, 2.5), ((5, -3, -1), (5, 3, 5)), 0
    elif window == "wall_y":
        portal, emitter, axis = (-0.75, 2 + e, 1, 0.75, 2 + e, 2.5), ((-3, 5, -1), (3, 5, 5)), 1
    else:
        raise ValueError(window)
    t.append('AttributeBegin\n  AreaLightSource "portal" "rgb L" [10 10 10] "bool twosided" "true"\n'
             '    "string portalData" "((AA %g %g %g %g %g %g %d -))" "string strategy" "%s"\n' % (portal + (axis, strategy)))
    t.append('  Shape "aaplane" "point lo" [%g %g %g] "point hi" [%g %g %g] "integer axis" [%d]\nAttributeEnd\n' % (emitter[0] + emitter[1] + (axis,)))
    t.append('Material "matte" "rgb Kd" [0.5 0.5 0.5]\n')
    t.append(_quad([-2, -2, 0, 2, -2, 0, 2, 2, 0, -2, 2, 0]))       # floor
    t.append(_quad([-2, -2, 0, -2, 2, 0, -2, 2, 3, -2, -2, 3]))     # wall x = -2
    t.append(_quad([-2, -2, 0, -2, -2, 3, 2, -2, 3, 2, -2, 0]))     # wall y = -2
    def holed(fix, a0, a1, b0, b1, h):  # a quad in the plane `fix` with the hole h = (a_lo, a_hi, b_lo, b_hi): four strips
        def q(alo, ahi, blo, bhi):
            return _quad(fix(alo, blo) + fix(alo, bhi) + fix(ahi, bhi) + fix(ahi, blo))
        return q(a0, h[0], b0, b1) + q(h[1], a1, b0, b1) + q(h[0], h[1], b0, h[2]) + q(h[0], h[1], h[3], b1)
    ceiling = lambda a, b: [a, b, 3]
    wall_x = lambda a, b: [2, a, b]
    wall_y = lambda a, b: [a, 2, b]
    t.append(holed(ceiling, -2, 2, -2, 2, (-0.75, 0.75, -0.75, 0.75)) if window == "ceiling" else _quad([-2, -2, 3, -2, 2, 3, 2, 2, 3, 2, -2, 3]))
    t.append(holed(wall_x, -2, 2, 0, 3, (-0.75, 0.75, 1, 2.5)) if window == "wall_x" else _quad([2, -2, 0, 2, -2, 3, 2, 2, 3, 2, 2, 0]))
    t.append(holed(wall_y, -2, 2, 0, 3, (-0.75, 0.75, 1, 2.5)) if window == "wall_y" else _quad([-2, 2, 0, 2, 2, 0, 2, 2, 3, -2, 2, 3]))
    mats = ['Material "matte" "rgb Kd" [0.55 0.5 0.4]\n', 'Material "plastic" "rgb Kd" [0.6 0.3 0.2] "rgb Ks" [0.2 0.2 0.2] "float roughness" [0.2]\n']
    for mat, (v, f) in zip(mats, _clutter()):
        t.append(mat)
        t.append('Shape "trianglemesh" "integer indices" [%s]\n  "point P" [%s]\n' %
                 (" ".join(map(str, f.ravel())), " ".join("%.9g" % x for x in v.ravel())))
    t.append("WorldEnd\n")
    return "".join(t)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--strategy", default="portal")
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--res", type=int, default=512)
    ap.add_argument("--window", default="ceiling")
    a = ap.parse_args()
    open(a.out, "w").write(scene_text(a.strategy, a.spp, a.res, a.window))
    print("wrote", a.out)
