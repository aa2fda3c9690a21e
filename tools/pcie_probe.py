"""H2D / D2H bandwidth of pinned buffers on the box (what bounds bench.py's e2e).

One process: `python tools/pcie_probe.py`.  All links at once: launch under torchrun
(`python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py`): every rank binds to
its GPU's NUMA node like bench.py does, the ranks start each copy loop together (gloo barrier) and rank 0 prints one JSON
line with the per-rank and aggregate rates, so that the 8-GPU e2e number can be read against what the host can feed."""
import json, os, sys, time
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("gloo")
    try:
        import bench
        bench.bind_near_gpu(local)
    except Exception as e:                                    # the probe still says something without the binding
        print(f"rank {rank}: no NUMA binding ({e})", file=sys.stderr)

def barrier():
    if world > 1: dist.barrier()

n = 315 * 1000 * 1000                                         # one tum step's images per GPU (1024 x 640 x 480)
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 4, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n // 4, dtype=torch.uint8, device="cuda")
out = {}
for name, src, dst in (("h2d", h, d), ("d2h", d, h)):
    for _ in range(2): dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize(); barrier()
    t = time.perf_counter()
    for _ in range(5): dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    out[name + "_GBps"] = n / ((time.perf_counter() - t) / 5) / 1e9
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); barrier(); t = time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
out["h2d_315MB_plus_d2h_79MB_ms"] = (time.perf_counter() - t) / 5 * 1e3
if world > 1:
    allout = [None] * world
    dist.all_gather_object(allout, out)
    if rank == 0:
        print(json.dumps({"ranks": world, "concurrent": True,
                          "h2d_GBps_per_rank": [round(o["h2d_GBps"], 1) for o in allout],
                          "d2h_GBps_per_rank": [round(o["d2h_GBps"], 1) for o in allout],
                          "h2d_GBps_total": round(sum(o["h2d_GBps"] for o in allout), 1),
                          "d2h_GBps_total": round(sum(o["d2h_GBps"] for o in allout), 1),
                          "both_ways_ms_per_rank": [round(o["h2d_315MB_plus_d2h_79MB_ms"], 2) for o in allout]}))
    dist.destroy_process_group()
else:
    print(json.dumps({"ranks": 1, **{k: round(v, 2) for k, v in out.items()}}))
