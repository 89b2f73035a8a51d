#!/usr/bin/env python
"""Pinned host -> device copy ceiling of the box with 1, 2, 4, 8 GPUs copying CONCURRENTLY (the ceiling of bench.py's end-to-end leg:
59 MB in and ~10 MB out per 32-pair batch and GPU). One process per GPU, all started on a barrier, each copying its own 59 MB pinned
buffer to its own GPU back to back for about a second, optionally with the 10.6 MB device -> host read-back on a second stream.
usage: pcie_rate.py [max_gpus=all]   -> one JSON line (also what profiles/r2_pcie_rate.json holds)"""
import json
import multiprocessing as mp
import os
import sys
import time


def worker(gpu, n, mode, start, out):
    import torch
    torch.cuda.set_device(gpu)
    x = torch.empty(59_000_000, dtype=torch.uint8).pin_memory()
    y = torch.empty_like(x, device="cuda")
    z = torch.empty(10_600_000, dtype=torch.uint8, device="cuda")
    zh = torch.empty(10_600_000, dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(5):
        with torch.cuda.stream(s1):
            y.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    start.wait()
    reps = 60
    t = time.perf_counter()
    for _ in range(reps):
        with torch.cuda.stream(s1):
            y.copy_(x, non_blocking=True)
        if mode == "both":
            with torch.cuda.stream(s2):
                zh.copy_(z, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / reps
    out.put((gpu, 59e-3 / dt))


def main():
    import torch
    have = torch.cuda.device_count()
    top = min(have, int(sys.argv[1]) if len(sys.argv) > 1 else have)
    ctx = mp.get_context("spawn")
    res = {"cpus": os.cpu_count(), "gpus_visible": have, "numa_nodes": len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
           if os.path.isdir("/sys/devices/system/node") else None, "runs": []}
    for n in [k for k in (1, 2, 4, 8) if k <= top]:
        for mode in ("h2d", "both"):
            start, out = ctx.Barrier(n), ctx.Queue()
            ps = [ctx.Process(target=worker, args=(g, n, mode, start, out)) for g in range(n)]
            for p in ps:
                p.start()
            rates = sorted(out.get() for _ in ps)
            for p in ps:
                p.join()
            per = [round(r, 1) for _, r in rates]
            res["runs"].append({"gpus": n, "mode": mode, "h2d_gbs_per_gpu": per, "h2d_gbs_total": round(sum(per), 1), "h2d_gbs_min": min(per)})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
