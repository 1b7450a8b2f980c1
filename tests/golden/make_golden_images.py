"""Generates tests/golden/images.npz: the reference binary's own renders
(oracle/_ref/pbrt, built from /root/reference by oracle/Makefile; one thread,
float PFM output) of the small scenes in tests/scenes.  Run in the development
container, where /root/reference exists:
    make -C oracle ref && python tests/golden/make_golden_images.py
The tier-2 parity tests compare the CUDA path with these images.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from dev_compare import read_pfm  # noqa: E402

# name -> (scene file, text substitutions)
CASES = {
    "box_glass": ("box_glass.pbrt", {}),
    "sphere_lights": ("sphere_lights.pbrt", {}),
    "portal_room_light": ("portal_room.pbrt", {"STRATEGY": "light"}),
    "portal_room_portal": ("portal_room.pbrt", {"STRATEGY": "portal"}),
    "portal_room_projection": ("portal_room.pbrt", {"STRATEGY": "projection"}),
    "tri_soup": ("tri_soup.pbrt", {}),
    # Integrator "directlighting": UniformSampleAllLights with sample arrays, and UniformSampleOneLight
    "direct_all": ("direct_mixed.pbrt", {"DIRECT_STRATEGY": "all"}),
    "direct_one": ("direct_mixed.pbrt", {"DIRECT_STRATEGY": "one"}),
    # Material "uber" with all five BxDFs (specular transmission for opacity, Lambert, microfacet, specular reflection,
    # specular transmission), under the path integrator
    "uber_path": ("direct_mixed.pbrt", {'Integrator "directlighting" "integer maxdepth" [3] "string strategy" "DIRECT_STRATEGY"': 'Integrator "path" "integer maxdepth" [6]',
                                        'Material "plastic" "rgb Kd" [.3 .5 .3] "rgb Ks" [.4 .4 .4] "float roughness" [.1]':
                                        'Material "uber" "rgb Kd" [.3 .5 .3] "rgb Ks" [.3 .3 .3] "rgb Kr" [.4 .4 .4] "rgb Kt" [.5 .5 .6] "rgb opacity" [.8 .8 .8] "float roughness" [.15] "float index" [1.4]'}),
    # the portal room under directlighting "all" (how the fork's own scenes use portals)
    "portal_room_direct": ("portal_room.pbrt", {"STRATEGY": "portal", 'Integrator "path" "integer maxdepth" [4]': 'Integrator "directlighting"'}),
    # directlighting over glass and uber with Kt: Li recurses into reflection AND transmission (directlighting.cpp:91-95)
    "direct_glass_all": ("direct_glass.pbrt", {}),
    # material-less boundaries: more wavefront iterations than maxdepth + 1 (the batch runs rounds until its queue is empty)
    "null_shells": ("null_shells.pbrt", {}),
    "direct_glass_one": ("direct_glass.pbrt", {'"integer maxdepth" [5] "string strategy" "all"': '"integer maxdepth" [7] "string strategy" "one"'}),
}


def scene_text(case):
    fn, subs = CASES[case]
    text = open(os.path.join(ROOT, "tests", "scenes", fn)).read()
    for k, v in subs.items():
        text = text.replace(k, v)
    return text


def main():
    """No arguments: every case.  With case names: only those, merged into the existing file."""
    pbrt = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    out = {}
    only = sys.argv[1:]
    if only:
        out = dict(np.load(os.path.join(ROOT, "tests", "golden", "images.npz")))
    for case in (only or CASES):
        with tempfile.TemporaryDirectory() as d:
            sp = os.path.join(d, "s.pbrt")
            open(sp, "w").write(scene_text(case))
            img = os.path.join(d, "o.pfm")
            subprocess.check_call([pbrt, "--nthreads", "1", "--quiet", "--minloglevel", "2", "--outfile", img, sp],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            out[case] = read_pfm(img)
            print(case, out[case].shape, float(out[case].mean()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "images.npz"), **out)


if __name__ == "__main__":
    main()
