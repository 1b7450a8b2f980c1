"""bench.py -- samples/s of the path integrator on BASELINE.json's configuration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--spp-per-step S]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (configs[1]): scenes/killeroo-simple.pbrt with Integrator "path" (maxdepth 5), Halton,
1920x1080 -- the 1024-spp job of BASELINE.json cut into steps.  One step = one pass of the hot
path (SamplerIntegrator::Render's loop + PathIntegrator::Li) over one batch of camera samples:
the full 1920x1080 frame at S x N consecutive Halton sample numbers (S = --spp-per-step, N =
GPUs), the frame's 16x16 sample tiles sharded round-robin over the N ranks (weak scaling: each
rank always renders S spp worth of a full frame), then one NCCL reduce of the float4 film.
Step k uses sample numbers [k*S*N, (k+1)*S*N) of the same 1024-spp sequence.

  value   device-timed (CUDA events, max over ranks), scene resident in HBM
  e2e     the same step through the C ABI with host buffers: pbrtgpu_upload_scene (H2D of the
          flattened scene) + pbrtgpu_render (kernels + D2H of the film), wall-clocked
  roofline   the dominant kernel (k_trace_closest), timed with CUDA events in a separate
          profile-mode pass; algorithmic bytes from SURVEY.md 8(d) with V and T counted on the
          device in that pass
  cpu_baseline / --impl reference   the reference's own binary (oracle/_ref/pbrt, built from
          /root/reference by oracle/Makefile) on all host cores, on a bounded sample of the workload
"""
import argparse
import ctypes as C
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "path samples per second (PathIntegrator::Li evaluations), killeroo-simple 1920x1080"
UNIT = "Msamples/s"
RES = (1920, 1080)
TOTAL_SPP = 1024


WORKLOAD = "killeroo"


def prepare_c4(quads):
    """Config C4 (tools/gen_c4.py): generated on the box, once (local rank 0), the other ranks wait for it."""
    d = os.path.join(tempfile.gettempdir(), "pbrtgpu_c4_%d" % quads)
    done = os.path.join(d, "done")
    if int(os.environ.get("LOCAL_RANK", "0")) == 0 and not os.path.exists(done):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_c4.py"), d, "--quads", str(quads)], stdout=subprocess.DEVNULL)
        open(done, "w").write("ok")
    while not os.path.exists(done):
        time.sleep(0.5)
    return os.path.join(d, "c4.pbrt")


def find_scene():
    if WORKLOAD.startswith("c4"):
        quads = int(WORKLOAD.split(":")[1]) if ":" in WORKLOAD else 2236
        return prepare_c4(quads), "synthetic C4 scene (tools/gen_c4.py, %d triangles, 4 portal skylights, strategy portal)" % (2 * quads * quads)
    p = os.path.join(ROOT, "oracle", "_ref", "scenes", "killeroo-path.pbrt")
    if os.path.exists(p):
        return p, "killeroo-simple.pbrt (reference scene file, Integrator line set to \"path\")"
    p = os.path.join(ROOT, "tests", "scenes", "killeroo_like.pbrt")
    if os.path.exists(p):
        return p, "killeroo_like.pbrt (synthetic stand-in of the same shape: the reference scene is not staged)"
    raise SystemExit("no benchmark scene: run __graft_entry__.build() where /root/reference exists, or tests/scenes/gen_scenes.py")


def scene_variant(path, spp, res):
    text = open(path).read()
    text = re.sub(r'"integer pixelsamples"\s*\[?\s*\d+\s*\]?', '"integer pixelsamples" [%d]' % spp, text)
    text = re.sub(r'"integer xresolution"\s*\[?\s*\d+\s*\]?', '"integer xresolution" [%d]' % res[0], text)
    text = re.sub(r'"integer yresolution"\s*\[?\s*\d+\s*\]?', '"integer yresolution" [%d]' % res[1], text)
    return text


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for i, nm in enumerate(names):
                if len(r) > 3 + i and r[3 + i].lower().startswith("active"):
                    reasons.add(nm)
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None, "reasons": sorted(reasons),
                "samples": len(self.rows)}


def run_reference_cpu(scene_path, spp, res, threads):
    """Times oracle/_ref/pbrt on `threads` host threads.  Render time = whole run minus a run of the
    same scene cropped to one pixel (parse + loop subdivision + BVH build)."""
    pbrt = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    if not os.path.exists(pbrt):
        return None
    text = scene_variant(scene_path, spp, res)
    d = os.path.dirname(scene_path)
    out = {}
    for tag, extra in (("full", ""), ("setup", ' "float cropwindow" [0 %g 0 %g]' % (1.0 / res[0], 1.0 / res[1]))):
        t = text.replace('Film "image"', 'Film "image"' + extra, 1)
        tmp = os.path.join(d, "_bench_%d_%s.pbrt" % (os.getpid(), tag))
        open(tmp, "w").write(t)
        try:
            t0 = time.perf_counter()
            subprocess.check_call([pbrt, "--nthreads", str(threads), "--quiet", "--minloglevel", "2", "--outfile",
                                   os.path.join(tempfile.gettempdir(), "_bench_%d.pfm" % os.getpid()), tmp],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            out[tag] = time.perf_counter() - t0
        finally:
            os.unlink(tmp)
    render_s = max(out["full"] - out["setup"], 1e-6)
    return {"samples": res[0] * res[1] * spp, "render_s": render_s, "whole_s": out["full"], "setup_s": out["setup"]}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene_path, scene_desc = find_scene()
    cores = os.cpu_count() or 1
    res = (RES[0] // 2, RES[1] // 2)  # bounded sample: quarter frame, 8 spp
    spp = 8
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "pbrt")):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/pbrt is not built (run __graft_entry__.build() where /root/reference exists)"}))
        return
    for _ in range(args.warmup):
        run_reference_cpu(scene_path, 1, (res[0] // 4, res[1] // 4), cores)
    tot_s, tot_n = 0.0, 0
    for _ in range(args.steps):
        r = run_reference_cpu(scene_path, spp, res, cores)
        tot_s += r["render_s"]
        tot_n += r["samples"]
    v = tot_n / tot_s / 1e6
    sample = "%dx%d at %d spp per step (render time = whole run minus a 1-pixel run), %d threads" % (res[0], res[1], spp, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "reference scene file", "config": {"workload": scene_desc + ", %dx%d path integrator maxdepth 5, halton" % RES, "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--spp-per-step", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="killeroo", help="killeroo (BASELINE.json configs[1], the default) | c4[:quads] (configs[3])")
    args = ap.parse_args()
    global WORKLOAD, RES, TOTAL_SPP, METRIC
    WORKLOAD = args.workload
    if WORKLOAD.startswith("c4"):
        RES, TOTAL_SPP = (3840, 2160), 4096
        METRIC = "path samples per second (PathIntegrator::Li evaluations), synthetic 10M-triangle portal scene 3840x2160"
    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the path integrator has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    scene_path, scene_desc = find_scene()
    S = args.spp_per_step
    hs = pk.HostScene(text=scene_variant(scene_path, TOTAL_SPP, RES), base_dir=os.path.dirname(scene_path))
    ctx = pk.GpuContext(local)
    ctx.upload(hs)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    params = hs.params.copy()
    params.shard_index, params.shard_count = rank, world
    cb = params.cropped_bounds
    n_pix = (cb[2] - cb[0]) * (cb[3] - cb[1])
    film_t = torch.zeros(n_pix * 4, dtype=torch.float32, device="cuda")
    sb = params.sample_bounds
    samples_per_step = (sb[2] - sb[0]) * (sb[3] - sb[1]) * S * world  # whole job, all ranks

    def step(k):
        params.spp_begin = (k * S * world) % TOTAL_SPP
        params.spp_end = params.spp_begin + S * world
        _, st = ctx.render_device(params)
        if world > 1:
            ctx.copy_film_device(film_t.data_ptr(), n_pix * 16)
            dist.reduce(film_t, 0)
        return st

    for k in range(args.warmup):
        step(k)
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if rank == 0 else None
    if clocks:
        clocks.start()
        time.sleep(0.3)
    # the barrier comes AFTER rank 0's sampler start-up: a rank that enters the timed region early spends the
    # difference waiting in its first film reduce (this cost 115 ms per run, 14 ms per step at 8 steps, until r01h)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    launches, rays = 0, 0
    for k in range(args.steps):
        st = step(args.warmup + k)
        launches += st.kernel_launches
        rays += st.closest_rays + st.shadow_rays
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    tot = torch.tensor([float(launches), float(rays)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    ms = float(ms.item())
    clk = clocks.finish() if clocks else None
    value = samples_per_step * args.steps / (ms / 1e3) / 1e6

    # ---- e2e: host buffers through the C ABI (scene H2D + render + film D2H), wall clock
    view = hs.view
    scene_bytes = (view.n_nodes * (24 + 4 + 2 + 1) + view.n_prims * (1 + 1 + 4 * 4 + 36 + (36 if view.tri_n else 0) + (24 if view.tri_uv else 0)))
    e2e_steps = max(2, min(args.steps, 4))
    film_host = torch.empty((cb[3] - cb[1], cb[2] - cb[0], 4), dtype=torch.float32, pin_memory=True).numpy()  # pinned result buffer
    ctx.upload(hs)
    ctx.render(params, out=film_host)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e2e_step_ms = []
    for k in range(e2e_steps):
        ts = time.perf_counter()
        params.spp_begin = (k * S * world) % TOTAL_SPP
        params.spp_end = params.spp_begin + S * world
        ctx.upload(hs)
        film, _ = ctx.render(params, out=film_host)
        e2e_step_ms.append(round((time.perf_counter() - ts) * 1e3, 2))
    t_e2e = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * e2e_steps / float(t_e2e.item()) / 1e6

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel: one profile-mode step (event pair around every launch)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    ctx.set_profile(True)
    params.spp_begin, params.spp_end = 0, S * world
    _, pst = ctx.render_device(params)
    prof = ctx.get_profile()
    ctx.set_profile(False)
    tc = prof["trace_closest"]
    total_ms = sum(p["ms"] for p in prof.values())
    # The dominant kernel is the traversal launch (k_ptrace_closest for the camera rays, then one
    # k_ptrace_fused per depth: path rays + next-event rays + shadow rays).  Algorithmic bytes
    # (SURVEY.md 8(d)): 32 B ray in + 16 B hit out per closest-hit ray, 32 B in + 1 B out per any-hit
    # ray, 32 B per node test (sizeof LinearBVHNode), 48 B per primitive test (indices + vertices).
    path_rays = pst.closest_rays - pst.nee_closest_rays
    node_tests = sum(prof[k]["node_tests"] for k in ("trace_closest", "trace_shadow", "trace_nee"))
    prim_tests = sum(prof[k]["prim_tests"] for k in ("trace_closest", "trace_shadow", "trace_nee"))
    alg_bytes = 48.0 * pst.closest_rays + 33.0 * pst.shadow_rays + 32.0 * node_tests + 48.0 * prim_tests
    all_rays = pst.closest_rays + pst.shadow_rays
    achieved = alg_bytes / (tc["ms"] / 1e3) / 1e9 if tc["ms"] > 0 else 0.0
    traffic, traffic_src = None, None
    import glob
    for tp in sorted(glob.glob(os.path.join(ROOT, "profiles", "*traffic.json"))) if WORKLOAD == "killeroo" else []:
        tj = json.load(open(tp))
        traffic, traffic_src = tj["dram_bytes_read"] + tj["dram_bytes_write"], os.path.basename(tp) + ": " + tj["capture"]
    roofline = {"kernel": "k_ptrace_fused (+ k_ptrace_closest for the camera rays)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "alg_bytes_per_launch": alg_bytes / max(1, tc["launches"]), "peak_source": peak_src, "launches": tc["launches"], "avg_launch_ms": tc["ms"] / max(1, tc["launches"]),
                "node_tests_per_ray": node_tests / max(1, all_rays), "prim_tests_per_ray": prim_tests / max(1, all_rays),
                "rays_in_profile_step": {"path_closest": int(path_rays), "nee_closest": int(pst.nee_closest_rays), "shadow_any": int(pst.shadow_rays),
                                         "camera": int(pst.camera_rays)},
                "kernel_ms_in_profile_step": {k: p["ms"] for k, p in prof.items()},
                "share_of_step": {k: (p["ms"] / total_ms if total_ms else 0) for k, p in prof.items()}}

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
           "data": "reference scene file" if "reference scene" in scene_desc else "synthetic",
           "config": {"workload": scene_desc + ", %dx%d" % RES + ", Integrator path maxdepth 5, halton; one step = %d spp of the full frame per GPU "
                      "(sample numbers of the %d-spp job), 16x16 tiles sharded over ranks, film reduce per step" % (S, TOTAL_SPP),
                      "spp_per_step_per_gpu": S, "l2": "wavefront state of a step (%.0f MB) and the film exceed L2; no flush needed" %
                      (min(samples_per_step // world, 64 << 20) * 290 / 1e6), "rays_per_sample": float(tot[1].item()) / (samples_per_step * args.steps)},
           "gpu_launches": int(tot[0].item()), "clocks": clk,
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(scene_bytes), "d2h_bytes_per_step": int(n_pix * 16), "steps": e2e_steps, "step_ms": e2e_step_ms},
           "roofline": roofline}
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        r = run_reference_cpu(scene_path, 8, (RES[0] // 2, RES[1] // 2), cores)
        if r:
            out["cpu_baseline"] = {"value": r["samples"] / r["render_s"] / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
                                   "sample": "960x540 at 8 spp, oracle/_ref/pbrt --nthreads %d, render time %.2f s (whole run %.2f s minus setup %.2f s)"
                                   % (cores, r["render_s"], r["whole_s"], r["setup_s"])}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
