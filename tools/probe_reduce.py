"""Where the per-step film reduce of bench.py spends its time at N > 1 (run under torchrun on the GPU box)."""
import importlib, os, sys, time
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pk = importlib.import_module("pbrt-v3-light-portals_b200")
scene_path, _ = bench.find_scene()
hs = pk.HostScene(text=bench.scene_variant(scene_path, 1024, bench.RES), base_dir=os.path.dirname(scene_path))
ctx = pk.GpuContext(local)
ctx.upload(hs)
stream = torch.cuda.current_stream()
ctx.set_stream(stream.cuda_stream)
p = hs.params.copy()
p.shard_index, p.shard_count = rank, world
cb = p.cropped_bounds
n_pix = (cb[2] - cb[0]) * (cb[3] - cb[1])
film = torch.zeros(n_pix * 4, dtype=torch.float32, device="cuda")
mode = sys.argv[1] if len(sys.argv) > 1 else "reduce"
for k in range(8):
    p.spp_begin, p.spp_end = 16 * world * k, 16 * world * (k + 1)
    t0 = time.perf_counter()
    _, st = ctx.render_device(p)
    t1 = time.perf_counter()
    ctx.copy_film_device(film.data_ptr(), n_pix * 16)
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if mode == "barrier_first":
        dist.barrier()
    t2 = time.perf_counter()
    ea.record()
    dist.reduce(film, 0)
    eb.record()
    t3 = time.perf_counter()
    torch.cuda.synchronize()
    t4 = time.perf_counter()
    print("rank %d step %d: render host %.1f ms (device %.1f), reduce call %.2f ms host, %.2f ms on the stream, sync %.1f ms" %
          (rank, k, (t1 - t0) * 1e3, st.device_ms, (t3 - t2) * 1e3, ea.elapsed_time(eb), (t4 - t3) * 1e3), flush=True)
dist.destroy_process_group()
