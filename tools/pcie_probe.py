"""Copy-only micro-benchmark of the box's host<->device path: every rank copies a
pinned 1 GiB buffer host->device and another device->host CONCURRENTLY on two streams, all ranks at once;
prints per-rank and aggregate GB/s. Launch: python -m torch.distributed.run --nproc-per-node N tools/pcie_probe.py"""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 27  # doubles: 1 GiB
h_in = torch.empty(n, dtype=torch.float64).pin_memory(); h_out = torch.empty(n, dtype=torch.float64).pin_memory()
h_in.fill_(1.0)
d_in = torch.empty(n, dtype=torch.float64, device="cuda"); d_out = torch.ones(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
res = {}
for mode in ("h2d", "d2h", "both"):
    for rep in range(3):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(4):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    gb = 4 * 8 * n / 1e9 * (2 if mode == "both" else 1)
    res[mode] = gb / dt
t = torch.tensor([res["h2d"], res["d2h"], res["both"]], device="cuda", dtype=torch.float64)
if world > 1:
    tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
else:
    tmin = t
if rank == 0:
    print("ranks %d: aggregate GB/s h2d %.1f d2h %.1f both directions %.1f | slowest rank h2d %.1f d2h %.1f both %.1f | cpu affinity %d cores" % (
        world, t[0], t[1], t[2], tmin[0], tmin[1], tmin[2], len(os.sched_getaffinity(0))))
if world > 1:
    dist.destroy_process_group()
