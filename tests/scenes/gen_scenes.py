"""Writes the small synthetic .pbrt scenes the parity tests feed to BOTH the
reference (through oracle/_ref) and this implementation.  Deterministic
(seeded); the generated files are committed so the tests do not need to run
this script.  Usage: python tests/scenes/gen_scenes.py
"""
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def fmt(a):
    return " ".join("%.9g" % float(x) for x in np.asarray(a).ravel())


def tri_soup(path, n_tris=3000, seed=11):
    """Random small triangles in a box, a sphere, an aaplane and two big walls:
    every primitive kind the BVH leaf loop knows, with shared vertices and
    axis-aligned faces so that edge/tie cases occur."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-10, 10, (n_tris, 1, 3))
    tri = c + rng.normal(0, 0.6, (n_tris, 3, 3))
    P = tri.reshape(-1, 3).astype(np.float32)
    idx = np.arange(3 * n_tris)
    # a welded grid (shared edges and vertices -> exact edge hits)
    g = 24
    xs = np.linspace(-10, 10, g + 1)
    gv = np.array([[x, -10.0, z] for z in xs for x in xs], np.float32)
    gi = []
    for j in range(g):
        for i in range(g):
            a = j * (g + 1) + i
            gi += [a, a + 1, a + g + 2, a, a + g + 2, a + g + 1]
    with open(path, "w") as f:
        f.write('LookAt 0 5 -40  0 0 0  0 1 0\nCamera "perspective" "float fov" [40]\n')
        f.write('Film "image" "integer xresolution" [64] "integer yresolution" [48] "string filename" "tri_soup.pfm"\n')
        f.write('Sampler "halton" "integer pixelsamples" [4]\nIntegrator "path" "integer maxdepth" [3]\nWorldBegin\n')
        f.write('AttributeBegin\n  AreaLightSource "diffuse" "rgb L" [8 8 8]\n  Translate 0 14 0\n  Shape "sphere" "float radius" [2.5]\nAttributeEnd\n')
        f.write('AttributeBegin\n  AreaLightSource "diffuse" "rgb L" [5 4 3] "bool twosided" "true"\n'
                '  Shape "aaplane" "point lo" [-4 12 -4] "point hi" [4 12 4] "integer axis" 1\nAttributeEnd\n')
        f.write('Material "matte" "rgb Kd" [.6 .6 .6]\n')
        f.write('Shape "trianglemesh" "integer indices" [%s]\n  "point P" [%s]\n' % (fmt(idx).replace(".0", ""), fmt(P)))
        f.write('Material "plastic" "rgb Kd" [.3 .5 .3] "rgb Ks" [.4 .4 .4] "float roughness" [.1]\n')
        f.write('Shape "trianglemesh" "integer indices" [%s]\n  "point P" [%s]\n' % (" ".join(map(str, gi)), fmt(gv)))
        f.write('AttributeBegin\n  Rotate 30 0 1 0\n  Scale 1.5 1 1\n  Material "matte" "rgb Kd" [.7 .3 .3]\n'
                '  Shape "trianglemesh" "integer indices" [0 1 2 2 3 0] "point P" [-12 -10 12  12 -10 12  12 12 12  -12 12 12]\nAttributeEnd\n')
        f.write('WorldEnd\n')


if __name__ == "__main__":
    tri_soup(os.path.join(HERE, "tri_soup.pbrt"))
    print("wrote", os.path.join(HERE, "tri_soup.pbrt"))
