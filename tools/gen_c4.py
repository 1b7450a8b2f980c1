"""Writes BASELINE.json's configuration C4: a synthetic many-triangle scene with
several portal skylights.  Deterministic (hash noise, no RNG state).

    python tools/gen_c4.py OUTDIR [--quads 2236] [--res 3840 2160] [--spp 16]

OUTDIR/c4.pbrt + two binary PLY meshes (a displaced quads x quads grid, 2 * quads^2 triangles,
split into a matte and a plastic mesh by a hash of the 32 x 32-quad patch) inside a large matte
box whose ceiling has four square holes; above each hole an aaplane emitter under
`AreaLightSource "portal"` with one AA portal in the hole, strategy "portal".
quads = 2236 gives 10.0 M triangles.
"""
import argparse
import os

import numpy as np


def hash01(ix, iy, seed):
    """Integer hash -> [0, 1), vectorised (Wang-style mix of the two lattice coordinates)."""
    h = (ix.astype(np.uint64) * np.uint64(73856093)) ^ (iy.astype(np.uint64) * np.uint64(19349663)) ^ np.uint64(seed * 83492791)
    h = (h ^ (h >> np.uint64(13))) * np.uint64(0x5BD1E995) & np.uint64(0xFFFFFFFF)
    h = h ^ (h >> np.uint64(15))
    return (h & np.uint64(0xFFFFFF)).astype(np.float64) / float(1 << 24)


def value_noise(x, y, seed):
    x0, y0 = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64)
    fx, fy = x - x0, y - y0
    sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
    a, b = hash01(x0, y0, seed), hash01(x0 + 1, y0, seed)
    c, d = hash01(x0, y0 + 1, seed), hash01(x0 + 1, y0 + 1, seed)
    return (a * (1 - sx) + b * sx) * (1 - sy) + (c * (1 - sx) + d * sx) * sy


def write_ply(path, verts, faces):
    with open(path, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
                 "element face %d\nproperty list uchar int vertex_indices\nend_header\n" % (len(verts), len(faces))).encode())
        f.write(np.ascontiguousarray(verts, dtype="<f4").tobytes())
        rec = np.empty(len(faces), dtype=[("n", "u1"), ("i", "<i4", (3,))])
        rec["n"] = 3
        rec["i"] = faces
        f.write(rec.tobytes())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("outdir")
    ap.add_argument("--quads", type=int, default=2236)
    ap.add_argument("--res", type=int, nargs=2, default=[3840, 2160])
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--seed", type=int, default=1234)
    args = ap.parse_args()
    os.makedirs(args.outdir, exist_ok=True)
    n = args.quads
    half = 20.0  # the terrain covers [-20, 20]^2
    u = np.linspace(-half, half, n + 1)
    X, Y = np.meshgrid(u, u, indexing="xy")
    Z = np.zeros_like(X)
    for octave, (freq, amp) in enumerate(((0.15, 2.0), (0.6, 0.7), (2.5, 0.2), (9.0, 0.05))):
        Z += amp * value_noise((X + half) * freq, (Y + half) * freq, args.seed + octave)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3)
    j, i = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    v00 = (j * (n + 1) + i).ravel()
    tris = np.stack([np.stack([v00, v00 + 1, v00 + n + 2], -1), np.stack([v00, v00 + n + 2, v00 + n + 1], -1)], 1)  # (quads, 2, 3)
    plastic = hash01((i // 32).ravel(), (j // 32).ravel(), args.seed + 99) < 0.5
    for name, sel in (("c4_matte.ply", ~plastic), ("c4_plastic.ply", plastic)):
        f = tris[sel].reshape(-1, 3)
        used, inv = np.unique(f, return_inverse=True)
        write_ply(os.path.join(args.outdir, name), verts[used], inv.reshape(-1, 3).astype(np.int32))
    H = 12.0  # ceiling height
    holes = [(-10, -10), (10, -10), (-10, 10), (10, 10)]
    hs = 3.0  # half size of a hole
    with open(os.path.join(args.outdir, "c4.pbrt"), "w") as f:
        f.write('LookAt -17 -17 7  4 4 1  0 0 1\nCamera "perspective" "float fov" [60]\n')
        f.write('Film "image" "string filename" ["c4.pfm"] "integer xresolution" [%d] "integer yresolution" [%d]\n' % tuple(args.res))
        f.write('Sampler "halton" "integer pixelsamples" [%d]\nIntegrator "path" "integer maxdepth" [5]\n' % args.spp)
        f.write('Accelerator "bvh" "string splitmethod" "sah"\nWorldBegin\n')
        for (cx, cy) in holes:
            f.write('AttributeBegin\n  AreaLightSource "portal" "rgb L" [18 17 15] "bool twosided" "true"\n'
                    '    "string portalData" "((AA %g %g %g %g %g %g 2 -))" "string strategy" "portal"\n'
                    '  Shape "aaplane" "point lo" [%g %g %g] "point hi" [%g %g %g] "integer axis" [2]\nAttributeEnd\n'
                    % (cx - hs, cy - hs, H + 0.001, cx + hs, cy + hs, H + 0.001, cx - 3 * hs, cy - 3 * hs, H + 6, cx + 3 * hs, cy + 3 * hs, H + 6))
        f.write('Material "matte" "rgb Kd" [0.55 0.55 0.5]\nShape "plymesh" "string filename" "c4_matte.ply"\n')
        f.write('Material "plastic" "rgb Kd" [0.25 0.4 0.3] "rgb Ks" [0.3 0.3 0.3] "float roughness" [0.15]\nShape "plymesh" "string filename" "c4_plastic.ply"\n')
        f.write('Material "matte" "rgb Kd" [0.6 0.6 0.6]\n')
        B = half

        def quad(p):
            f.write('Shape "trianglemesh" "integer indices" [0 1 2 2 3 0] "point P" [%s]\n' % " ".join("%g" % v for v in p))
        quad([-B, -B, -1, B, -B, -1, B, B, -1, -B, B, -1])            # floor under the terrain
        quad([-B, -B, -1, -B, B, -1, -B, B, H, -B, -B, H])            # walls
        quad([B, -B, -1, B, -B, H, B, B, H, B, B, -1])
        quad([-B, B, -1, B, B, -1, B, B, H, -B, B, H])
        quad([-B, -B, -1, -B, -B, H, B, -B, H, B, -B, -1])
        # ceiling: a 5 x 5 arrangement of strips around the four holes
        xs = [-B, -10 - hs, -10 + hs, 10 - hs, 10 + hs, B]
        for a in range(5):
            for b in range(5):
                if a in (1, 3) and b in (1, 3):
                    continue  # a hole
                quad([xs[a], xs[b], H, xs[a], xs[b + 1], H, xs[a + 1], xs[b + 1], H, xs[a + 1], xs[b], H])
        f.write('WorldEnd\n')
    print("wrote %s: %d triangles" % (os.path.join(args.outdir, "c4.pbrt"), 2 * n * n))


if __name__ == "__main__":
    main()
