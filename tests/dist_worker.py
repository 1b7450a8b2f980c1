"""Worker of tests/test_distributed.py: one rank of the multi-GPU plan, on the CPU.
Each rank renders its tile shard (shard_index = RANK, shard_count = WORLD_SIZE) with the
CPU-stepped device code, the films are summed onto rank 0 with torch.distributed (gloo here,
NCCL on the GPUs: bench.py), and rank 0 writes the image."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden"), os.path.join(ROOT, "tools")]
import util  # noqa: E402
from dev_compare import hostsim_lib  # noqa: E402
from make_golden_images import scene_text  # noqa: E402


def main():
    case, out = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    hs = pk.HostScene(text=scene_text(case), base_dir=os.path.join(ROOT, "tests", "scenes"))
    p = hs.params.copy()
    p.shard_index, p.shard_count = rank, world
    import ctypes as C
    lib = C.CDLL(os.path.join(ROOT, "tests", "_build", "libhostsim.so"))
    lib.hostsim_create.restype = C.c_void_p
    lib.hostsim_create.argtypes = [C.c_void_p]
    lib.hostsim_destroy.argtypes = [C.c_void_p]
    _, film, st = util.hostsim_render(lib, hs, p, threads=2)
    t = torch.from_numpy(film.copy())
    rays = torch.tensor([st[0]], dtype=torch.int64)
    dist.reduce(t, 0)
    dist.reduce(rays, 0)
    if rank == 0:
        np.savez(out, rgb=hs.film_to_rgb(t.numpy()), camera_rays=int(rays.item()))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
