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

def ncu_kernel_figures(kernel_substr, workload_tag=None):
    """Lanes per instruction, issue rate and L2 hit rate of a kernel from the newest committed ncu summary under
    profiles/ (tools/ncu_summary.py over an `ncu --set full` capture of this same bench command): the average over
    the captured launches of that kernel, weighted by their duration.  workload_tag: only summaries whose file name
    carries it (the C4 captures: *_c4_*; Cornell: *_c1_*), else only those that carry no workload tag (killeroo)."""
    import glob
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu*summary*.txt"))):
        if workload_tag is None and any(t in os.path.basename(path) for t in ("_c4_", "_c1_")):
            continue
        if workload_tag is not None and ("_%s_" % workload_tag) not in os.path.basename(path):
            continue
        blocks, cur, rec = [], None, {}
        for line in list(open(path)) + ["----"]:
            if line.startswith("----"):
                if cur and kernel_substr in cur and {"lanes", "issue", "ms"} <= set(rec):
                    blocks.append(rec)
                cur, rec = line, {}
            elif "gpu__time_duration.sum" in line:
                rec["ms"] = float(line.split()[-1])
            elif "smsp__thread_inst_executed_per_inst_executed.ratio" in line:
                rec["lanes"] = float(line.split()[-1])
            elif "sm__inst_executed.sum.pct_of_peak_sustained_elapsed" in line:
                rec["issue"] = float(line.split()[-1]) / 100.0
            elif "lts__t_sector_hit_rate.pct" in line:
                rec["l2_hit"] = float(line.split()[-1]) / 100.0
        if blocks:
            w = sum(r["ms"] for r in blocks)
            best = {k: sum(r.get(k, 0.0) * r["ms"] for r in blocks) / w for k in ("lanes", "issue", "l2_hit")}
            best.update(source=os.path.basename(path), launches=len(blocks))
    return best


def reference_image(text, base_dir, threads):
    """One render of the reference binary (float PFM), for the rel_rmse of other_configs."""
    pbrt = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    if not os.path.exists(pbrt):
        return None
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from dev_compare import read_pfm
    tmp = os.path.join(base_dir, "_bench_oc_%d.pbrt" % os.getpid())
    out = os.path.join(tempfile.gettempdir(), "_bench_oc_%d.pfm" % os.getpid())
    open(tmp, "w").write(text)
    try:
        subprocess.check_call([pbrt, "--nthreads", str(threads), "--quiet", "--minloglevel", "2", "--outfile", out, tmp],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return read_pfm(out)
    finally:
        os.unlink(tmp)
        if os.path.exists(out):
            os.unlink(out)


def other_configs(pk, device):
    """BASELINE.json's other configurations, after the timed region (N = 1): short device-timed runs at the named
    size, each with a check against the reference -- images: rel RMSE against oracle/_ref/pbrt at 8 spp (the full-size
    comparisons are tests/test_baseline_scenes.py); ray batches: IDs / t bits / occlusion flags of the first 2 M rays of
    the timed batch against the reference's BVHAccel (tools/bench_rays.py checks whole 100 M batches)."""
    import numpy as np
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import gen_c3
    import util
    cores = os.cpu_count() or 1
    out = {}
    ctx = pk.GpuContext(device)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)

    def image_config(name, text_of_spp, base_dir, spp):
        hs = pk.HostScene(text=text_of_spp(spp), base_dir=base_dir)
        ctx.upload(hs)
        best = None
        for _ in range(3):
            _, st = ctx.render_device(hs.params)
            best = st.device_ms if best is None else min(best, st.device_ms)
        ctx.set_profile(True)
        ctx.render_device(hs.params)
        prof = ctx.get_profile()
        ctx.set_profile(False)
        tot = sum(v["ms"] for v in prof.values()) or 1.0
        sbb = hs.params.sample_bounds
        samples = (sbb[2] - sbb[0]) * (sbb[3] - sbb[1]) * hs.params.spp
        rec = {"value": samples / best / 1e3, "unit": UNIT, "samples": samples, "device_ms": best,
               "rays_per_sample": (st.closest_rays + st.shadow_rays) / samples, "share_of_step": {k: v["ms"] / tot for k, v in prof.items()}}
        small = pk.HostScene(text=text_of_spp(8), base_dir=base_dir)
        ref = reference_image(text_of_spp(8), base_dir, cores)
        if ref is not None:
            ctx.upload(small)
            film, _ = ctx.render(small.params)
            s = util.image_stats(small.film_to_rgb(film), ref)
            rec["rel_rmse"], rec["mean_rel"], rec["checked"] = s["rel_rmse"], s["mean_rel"], "same scene at 8 spp against oracle/_ref/pbrt (%d threads)" % cores
        out[name] = rec

    ref_scenes = os.path.join(ROOT, "oracle", "_ref", "scenes")
    c1 = os.path.join(ref_scenes, "cornell-path.pbrt")
    if os.path.exists(c1):
        image_config("C1 cornell_dielectric path maxdepth 5, 512x512, 64 spp", lambda spp: scene_variant(c1, spp, (512, 512)), ref_scenes, 64)
    for strategy in ("light", "portal", "projection"):
        image_config("C3 skylight room (tools/gen_c3.py), 512x512, 64 spp, strategy " + strategy,
                     lambda spp, s=strategy: gen_c3.scene_text(s, spp, 512, "ceiling"), tempfile.gettempdir(), 64)
    kil = os.path.join(ref_scenes, "killeroo-path.pbrt")
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if os.path.exists(kil):
        hs = pk.HostScene(kil)
        ctx.upload(hs)
        n, k = 10_000_000, 2_000_000
        box = ([-100, -100, -150], [100, 150, 50])
        sets = {"incoherent closest": (lambda: util.rays_incoherent_closest(hs.view, n, seed=42), False),
                "toward-object closest": (lambda: util.rays_toward_box(n, box[0], box[1], seed=44), False),
                "incoherent any-hit": (lambda: util.rays_shadow(hs.view, n, seed=43), True),
                "object-bound any-hit": (lambda: util.rays_shadow(hs.view, n, seed=46, lo=np.array(box[0], float) * 1.5, hi=np.array(box[1], float) * 1.5), True)}
        shm = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        for name, (make, any_hit) in sets.items():
            rays = make()
            d_rays = torch.from_numpy(rays).cuda()
            d_out = torch.empty(n * (1 if any_hit else 16), dtype=torch.uint8, device="cuda")
            ms = min(ctx.intersect_device(d_rays.data_ptr(), n, d_out.data_ptr(), any_hit) for _ in range(4))
            rec = {"value": n / ms / 1e3, "unit": "Mrays/s", "rays": n, "device_ms": ms}
            if os.path.exists(harness):
                rp, op = os.path.join(shm, "_oc_rays_%d.bin" % os.getpid()), os.path.join(shm, "_oc_out_%d.bin" % os.getpid())
                rays[:k].tofile(rp)
                try:
                    subprocess.run([harness, "intersect", kil, rp, op, "any" if any_hit else "closest", str(cores)], stdout=subprocess.DEVNULL,
                                   stderr=subprocess.DEVNULL, check=True)
                    got = d_out.cpu().numpy()
                    if any_hit:
                        same = bool(np.array_equal(got[:k], np.fromfile(op, np.uint8)))
                    else:
                        want = np.fromfile(op, np.uint32).reshape(-1, 2)
                        g = got.view(np.uint32).reshape(-1, 4)[:k]
                        hit = want[:, 0] != 0xFFFFFFFF
                        same = bool(np.array_equal(g[:, 0], want[:, 0]) and np.array_equal(g[hit, 1], want[hit, 1]))
                    rec["bit_exact"], rec["rays_checked"] = same, k
                finally:
                    for f in (rp, op):
                        if os.path.exists(f):
                            os.unlink(f)
            out["C5 killeroo BVH, 10 M rays, " + name] = rec
            del d_rays, d_out
    ctx.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--spp-per-step", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the short runs of C1 / C3 / C5 after the timed region (N = 1 only)")
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
    # The film of step k is accumulated straight in film_t[k % 2] (pbrtgpu_render_device_into: no staging copy, no host
    # synchronisation) and handed to NCCL as it is; the reduce is issued asynchronously, so it runs on NCCL's stream
    # while the kernels of step k + 1 run on ours.  Two buffers: step k + 2 waits for the reduce of step k.
    film_t = [torch.zeros(n_pix * 4, dtype=torch.float32, device="cuda") for _ in range(2)]
    pending = [None, None]
    sb = params.sample_bounds
    samples_per_step = (sb[2] - sb[0]) * (sb[3] - sb[1]) * S * world  # whole job, all ranks

    def set_range(k):
        params.spp_begin = (k * S * world) % TOTAL_SPP
        params.spp_end = params.spp_begin + S * world

    def step(k):
        set_range(k)
        b = k % 2
        if pending[b] is not None:
            pending[b].wait()  # orders our stream behind the reduce that last read this buffer (issued two steps ago)
            pending[b] = None
        ctx.render_device_into(params, film_t[b].data_ptr())
        if world > 1:
            pending[b] = dist.reduce(film_t[b], 0, async_op=True)

    def drain():
        for b in range(2):
            if pending[b] is not None:
                pending[b].wait()
                pending[b] = None

    for k in range(args.warmup):
        step(k)
    drain()
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
    launches0 = ctx.kernel_launches()
    e0.record(stream)
    for k in range(args.steps):
        step(args.warmup + k)
    drain()  # the last reduces are inside the timed region
    e1.record(stream)
    torch.cuda.synchronize()
    launches = ctx.kernel_launches() - launches0
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    tot = torch.tensor([float(launches)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    ms = float(ms.item())
    clk = clocks.finish() if clocks else None
    value = samples_per_step * args.steps / (ms / 1e3) / 1e6

    # ---- N > 1: is the film NCCL reduced the image?  One more step; rank 0 then renders the SAME sample numbers alone
    # (shard_count 1) and compares (the N-GPU image equals the 1-GPU image up to float summation order).
    film_check = None
    if world > 1:
        set_range(0)
        ctx.render_device_into(params, film_t[0].data_ptr())
        dist.reduce(film_t[0], 0)
        torch.cuda.synchronize()
        if rank == 0:
            solo = params.copy()
            solo.shard_index, solo.shard_count = 0, 1
            ctx.render_device_into(solo, film_t[1].data_ptr())
            torch.cuda.synchronize()
            a, b = film_t[0].double().view(-1, 4), film_t[1].double().view(-1, 4)
            rel_rmse = float(torch.sqrt(((a[:, :3] - b[:, :3]) ** 2).mean()) / b[:, :3].abs().mean())
            w_equal = bool(torch.equal(film_t[0].view(-1, 4)[:, 3], film_t[1].view(-1, 4)[:, 3]))
            film_check = {"rel_rmse_vs_1gpu_render_of_same_samples": rel_rmse, "filter_weight_sums_equal": w_equal,
                          "spp_compared": S * world, "tolerance": 1e-6, "ok": bool(rel_rmse <= 1e-6 and w_equal)}
        dist.barrier()

    # ---- e2e: the same step through the C ABI with HOST buffers, wall clock, per rank: H2D of the flattened scene
    # (pbrtgpu_upload_scene), the kernels, at N > 1 the NCCL reduce of the film, D2H of the film into pinned memory
    # (rank 0 at N > 1: the other ranks' films end in the reduce).  upload and render are also timed separately.
    view = hs.view
    scene_bytes = (view.n_nodes * (24 + 4 + 2 + 1) + view.n_prims * (1 + 1 + 4 * 4 + 36 + (36 if view.tri_n else 0) + (24 if view.tri_uv else 0)))
    e2e_steps = max(2, min(args.steps, 4))
    film_host_t = torch.empty((cb[3] - cb[1], cb[2] - cb[0], 4), dtype=torch.float32, pin_memory=True)
    film_host = film_host_t.numpy()

    def e2e_step(k):
        set_range(k)
        t0 = time.perf_counter()
        ctx.upload(hs)
        t1 = time.perf_counter()
        if world == 1:
            ctx.render(params, out=film_host)
        else:
            ctx.render_device_into(params, film_t[0].data_ptr())
            dist.reduce(film_t[0], 0)
            if rank == 0:
                film_host_t.view(-1).copy_(film_t[0], non_blocking=True)
            torch.cuda.synchronize()
        t2 = time.perf_counter()
        return (t1 - t0) * 1e3, (t2 - t1) * 1e3

    e2e_step(0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e2e_upload_ms, e2e_render_ms = [], []
    for k in range(e2e_steps):
        u, r = e2e_step(k)
        e2e_upload_ms.append(round(u, 2))
        e2e_render_ms.append(round(r, 2))
    t_e2e = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * e2e_steps / float(t_e2e.item()) / 1e6
    e2e_sync_value, e2e_pipeline = None, None

    # ---- e2e as a renderer would run it: TWO steps in flight.  Two contexts (each with its own copy of the scene, its own
    # wave buffers and stream) take the steps in turn through the same calls -- pbrtgpu_upload_scene from host memory,
    # pbrtgpu_render_device_into, at N > 1 the NCCL reduce of the film, the film download into pinned memory (rank 0 at N > 1)
    # -- so the scene upload of step k + 1 and the reduce / download of step k - 1 run under the kernels of step k, and the
    # tail of one step's launches is filled by the other's.  Every step's H2D and D2H stay inside the timed region.  The
    # one-step-at-a-time figure above is kept as `sync_value`.  PBRTGPU_BENCH_E2E_PIPELINE=0 skips this.
    if os.environ.get("PBRTGPU_BENCH_E2E_PIPELINE", "1") != "0":
        ctx_b = pk.GpuContext(local)
        ctx_b.upload(hs)
        lanes = [(ctx, torch.cuda.Stream()), (ctx_b, torch.cuda.Stream())]
        host_films = [film_host_t, torch.empty_like(film_host_t).pin_memory()]
        done = [torch.cuda.Event(), torch.cuda.Event()]
        for (c, st) in lanes:
            c.set_stream(st.cuda_stream)

        def pipe_step(k):
            i = k % 2
            c, st = lanes[i]
            done[i].synchronize()  # this lane's previous film has reached the host; its buffers are free
            set_range(k)
            c.upload(hs)
            c.render_device_into(params, film_t[i].data_ptr())
            with torch.cuda.stream(st):
                if world > 1:
                    dist.reduce(film_t[i], 0, async_op=True).wait()  # NCCL's stream behind the render, this stream behind NCCL
                if rank == 0:
                    host_films[i].view(-1).copy_(film_t[i], non_blocking=True)
                done[i].record(st)

        for i in range(2):
            done[i].record(lanes[i][1])
        for k in range(2):  # warm-up: the second context's buffers, both lanes primed
            pipe_step(k)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        n_pipe = 2 * e2e_steps
        t0 = time.perf_counter()
        for k in range(n_pipe):
            pipe_step(k)
        torch.cuda.synchronize()
        t_pipe = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t_pipe, op=dist.ReduceOp.MAX)
        t_pipe = float(t_pipe.item())
        e2e_sync_value = e2e_value
        e2e_value = samples_per_step * n_pipe / t_pipe / 1e6
        ctx_b.close()
        ctx.set_stream(stream.cuda_stream)
        ctx.upload(hs)
        # what the pipeline delivered for its last step against a one-step-at-a-time render (and reduce) of the same sample numbers
        last = host_films[(n_pipe - 1) % 2].clone()
        set_range(n_pipe - 1)
        if world == 1:
            ctx.render(params, out=film_host)
        else:
            ctx.render_device_into(params, film_t[0].data_ptr())
            dist.reduce(film_t[0], 0)
            if rank == 0:
                film_host_t.view(-1).copy_(film_t[0])
            torch.cuda.synchronize()
        if rank == 0:
            a, b = last.double().view(-1, 4), film_host_t.double().view(-1, 4)
            pipe_rel = float(torch.sqrt(((a[:, :3] - b[:, :3]) ** 2).mean()) / b[:, :3].abs().mean())
            e2e_pipeline = {"how": "two steps in flight: two contexts (own scene copy, wave buffers and stream), taking the steps in turn",
                            "steps": n_pipe, "wall_ms": round(t_pipe * 1e3, 1),
                            "film_rel_rmse_vs_synchronous_render": pipe_rel, "ok": bool(pipe_rel <= 1e-6 and torch.equal(a[:, 3], b[:, 3]))}

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
    tag = "c4" if WORKLOAD.startswith("c4") else "killeroo"
    for tp in sorted(glob.glob(os.path.join(ROOT, "profiles", "*traffic*.json"))):
        tj = json.load(open(tp))
        if tj.get("workload", "killeroo") == tag:
            traffic, traffic_src = tj["dram_bytes_read"] + tj["dram_bytes_write"], os.path.basename(tp) + ": " + tj["capture"]
            # launches differ in size (first depth against last, one batch against another): carry the capture's DRAM bytes per
            # millisecond over to this run's average launch
            if tj.get("duration_ms") and tc["launches"] and tc["ms"] > 0:
                traffic = traffic / tj["duration_ms"] * (tc["ms"] / tc["launches"])
    alg_per_launch = alg_bytes / max(1, tc["launches"])
    ncu = ncu_kernel_figures("k_ptrace_fused", "c4" if tag == "c4" else None)
    # What bounds the kernel: on killeroo the BVH (9 MB) is served by L1 / L2 and DRAM sees 5 % of the algorithmic bytes --
    # the limiter is the SM's issue rate times the share of lanes that do work.  `achieved / peak / frac` stay the
    # HBM-normalised figures of SURVEY.md 8(d), the common yardstick; they are not a claim that the kernel is at 63 %
    # of a memory roofline.
    roofline = {"kernel": "k_ptrace_fused (+ k_ptrace_closest for the camera rays)",
                "bound": "issue" if tag == "killeroo" else "hbm latency (L2 hit 59 %, 1.06 TB/s of 32-byte DRAM sectors under the launch: profiles/r02o_ncu_c4_trace_summary.txt)", "yardstick": "hbm (SURVEY.md 8(d) algorithmic bytes / measured copy bandwidth)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "issue_frac": ncu["issue"] if ncu else None, "lanes_per_inst": ncu["lanes"] if ncu else None,
                "thread_inst_frac": (ncu["issue"] * ncu["lanes"] / 32.0) if ncu else None, "l2_hit": ncu.get("l2_hit") if ncu else None,
                "ncu_source": ncu["source"] if ncu else None,
                "traffic": traffic, "traffic_ratio": (traffic / alg_per_launch) if traffic else None, "traffic_source": traffic_src,
                "alg_bytes_per_launch": alg_per_launch, "peak_source": peak_src, "launches": tc["launches"], "avg_launch_ms": tc["ms"] / max(1, tc["launches"]),
                "node_tests_per_ray": node_tests / max(1, all_rays), "prim_tests_per_ray": prim_tests / max(1, all_rays),
                "rays_in_profile_step": {"path_closest": int(path_rays), "nee_closest": int(pst.nee_closest_rays), "shadow_any": int(pst.shadow_rays),
                                         "camera": int(pst.camera_rays)},
                "kernel_ms_in_profile_step": {k: p["ms"] for k, p in prof.items()},
                "share_of_step": {k: (p["ms"] / total_ms if total_ms else 0) for k, p in prof.items()}}

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
           "data": "reference scene file" if "reference scene" in scene_desc else "synthetic",
           "config": {"workload": scene_desc + ", %dx%d" % RES + ", Integrator path maxdepth 5, halton; one step = %d spp of the full frame per GPU "
                      "(sample numbers of the %d-spp job), 16x16 tiles sharded over ranks, film reduce per step (asynchronous: it overlaps the next step)" % (S, TOTAL_SPP),
                      "spp_per_step_per_gpu": S, "l2": "wavefront state of a step (%.0f MB) and the film exceed L2; no flush needed" %
                      (min(samples_per_step // world, 64 << 20) * 290 / 1e6),
                      "rays_per_sample": float(all_rays) / (samples_per_step / world) if world == 1 else float(all_rays) * world / samples_per_step},
           "gpu_launches": int(tot[0].item()), "clocks": clk,
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(scene_bytes), "d2h_bytes_per_step": int(n_pix * 16), "steps": e2e_steps,
                   "upload_ms": e2e_upload_ms, "render_ms": e2e_render_ms, "sync_value": e2e_sync_value, "pipeline": e2e_pipeline,
                   "includes": "scene H2D + kernels + film D2H" + (" + NCCL reduce of the film; times are rank 0's, the value is the max over ranks" if world > 1 else "")},
           "roofline": roofline}
    if film_check is not None:
        out["film_check"] = film_check
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        r = run_reference_cpu(scene_path, 8, (RES[0] // 2, RES[1] // 2), cores)
        if r:
            out["cpu_baseline"] = {"value": r["samples"] / r["render_s"] / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
                                   "sample": "960x540 at 8 spp, oracle/_ref/pbrt --nthreads %d, render time %.2f s (whole run %.2f s minus setup %.2f s)"
                                   % (cores, r["render_s"], r["whole_s"], r["setup_s"])}
    if world == 1 and not args.no_other_configs and WORKLOAD == "killeroo":
        ctx.close()
        try:
            out["other_configs"] = other_configs(pk, local)
        except Exception as e:  # the headline line must not be lost to a side run
            out["other_configs"] = {"error": repr(e)}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
