"""The multi-GPU plan on the CPU: world size 2, gloo.  Tiles t with
t % WORLD_SIZE == RANK per rank, one reduce(sum) of the float4 film to rank 0."""
import os
import socket
import subprocess
import sys

import numpy as np

import util
from conftest import GOLDEN, ROOT

sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden_images import scene_text  # noqa: E402


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_ranks_reduce_to_the_reference_image(pkg, hostsim_lib, tmp_path):
    out = str(tmp_path / "img.npz")
    port = free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), "box_glass", out], env=env))
    for p in procs:
        assert p.wait(timeout=300) == 0
    got = np.load(out)
    ref = np.load(os.path.join(GOLDEN, "images.npz"))["box_glass"]
    s = util.image_stats(got["rgb"], ref)
    assert s["rel_rmse"] <= 1e-5, s
    hs = pkg.HostScene(text=scene_text("box_glass"), base_dir=os.path.join(ROOT, "tests", "scenes"))
    sb = hs.params.sample_bounds
    assert int(got["camera_rays"]) == (sb[2] - sb[0]) * (sb[3] - sb[1]) * hs.params.spp
