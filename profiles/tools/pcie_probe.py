"""Per-rank host<->device copy bandwidth with all ranks copying at once (names the link that bounds the host-buffer `e2e` figure).
launch: python -m torch.distributed.run --nproc-per-node N profiles/tools/pcie_probe.py"""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
res = {}
for name, mb in (("obs_like_13MB", 13.2), ("act_like_4MB", 4.2), ("bulk_256MB", 256.0)):
    n = int(mb * 1e6) // 4
    h = torch.empty(n, dtype=torch.float32).pin_memory()
    d = torch.empty(n, dtype=torch.float32, device="cuda")
    for direction in ("d2h", "h2d"):
        def once():
            (h.copy_(d, non_blocking=True) if direction == "d2h" else d.copy_(h, non_blocking=True))
        for _ in range(3):
            once()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        reps = 20
        for _ in range(reps):
            once()
        b.record()
        torch.cuda.synchronize()
        res[f"{name}_{direction}_GBps"] = round(n * 4 * reps / (a.elapsed_time(b) * 1e-3) / 1e9, 2)
t = torch.tensor(list(res.values()), device="cuda", dtype=torch.float64)
mn = t.clone()
if world > 1:
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
if rank == 0:
    print(json.dumps({"world": world, "per_rank_min": dict(zip(res, mn.tolist())), "aggregate": dict(zip(res, [round(x, 1) for x in t.tolist()]))}))
if world > 1:
    dist.destroy_process_group()
