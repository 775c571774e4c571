#!/usr/bin/env python
"""Plain pinned host<->device copy bandwidth of the box, per process and in aggregate: the ceiling of
bench.py's `e2e` leg, which moves 48 MB host->device and 54 MB device->host per step.

    python scripts/h2d_microbench.py                                   # one process, GPU 0
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/h2d_microbench.py

One `cudaMemcpyAsync` per buffer (torch `copy_(non_blocking=True)` on a pinned tensor), nothing batched, no kernels:
  h2d      48 MB pinned -> device, back to back on one stream
  d2h      54 MB device -> pinned
  duplex   both at once on two streams (what a pipelined e2e step does)
Every rank binds to its GPU's CPU set first (NVML), like bench.py does for its pinned buffers.  Rank 0 prints one JSON line
with per-rank and aggregate GB/s (max-over-ranks time), plus the NUMA node of each GPU if sysfs shows it.
"""
from __future__ import annotations

import json
import os
import sys
import time

import torch
import torch.distributed as dist

H2D_BYTES, D2H_BYTES = 48_000_000, 54_067_200


def main():
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    numa = None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        p = f"/sys/bus/pci/devices/{bus.lower()[-12:]}/numa_node"
        numa = int(open(p).read()) if os.path.exists(p) else None
    except Exception:
        pass
    hin = torch.empty(H2D_BYTES, dtype=torch.uint8, pin_memory=True).fill_(1)
    hout = torch.empty(D2H_BYTES, dtype=torch.uint8, pin_memory=True)
    din, dout = torch.empty(H2D_BYTES, dtype=torch.uint8, device=dev), torch.ones(D2H_BYTES, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, reps=40):
        for _ in range(5):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            worst = float(t.item())
        else:
            worst = dt
        barrier()
        return dt / reps, worst / reps

    def f_h2d():
        with torch.cuda.stream(s1):
            din.copy_(hin, non_blocking=True)

    def f_d2h():
        with torch.cuda.stream(s2):
            hout.copy_(dout, non_blocking=True)

    def f_duplex():
        f_h2d()
        f_d2h()

    rec = {}
    for name, fn, nbytes in (("h2d", f_h2d, H2D_BYTES), ("d2h", f_d2h, D2H_BYTES), ("duplex", f_duplex, H2D_BYTES + D2H_BYTES)):
        mine, worst = timed(fn)
        rec[name] = {"my_GBs": nbytes / mine / 1e9, "aggregate_GBs": world * nbytes / worst / 1e9, "ms_per_step_worst_rank": worst * 1e3}
    rows = [None] * world
    row = {"rank": rank, "numa_node": numa, **{k: round(v["my_GBs"], 2) for k, v in rec.items()}}
    if world > 1:
        dist.all_gather_object(rows, row)
    else:
        rows = [row]
    if rank == 0:
        steps_per_s = 1e3 / rec["duplex"]["ms_per_step_worst_rank"]
        print(json.dumps({"n_procs": world, "h2d_bytes": H2D_BYTES, "d2h_bytes": D2H_BYTES,
                          "aggregate_GBs": {k: round(v["aggregate_GBs"], 2) for k, v in rec.items()},
                          "e2e_ceiling_clips_per_s": round(world * 100 * steps_per_s),
                          "per_rank": rows, "host_cpus": os.cpu_count(),
                          "note": "e2e_ceiling = 100 clips per duplex step (48 MB in + 54 MB out) at the worst rank's rate x ranks"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
