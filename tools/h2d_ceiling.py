#!/usr/bin/env python
"""Host-link ceiling probe: what N concurrent pinned host->device streams reach on this box with NO kernels running.

    python tools/h2d_ceiling.py [--ns 1,2,4,8] [--mb 15.36] [--d2h-mb 2.05] [--seconds 1.0] [--placement spread|linear]

For every N it starts N processes (one per GPU, like the bench's ranks), each with `--depth` pinned buffers of `--mb`
megabytes (the bench's e2e arm: 8 x 120 k points x 16 B = 15.36 MB per step) and one CUDA graph per buffer holding just
the H2D copy (and optionally the D2H of `--d2h-mb`), replays them round-robin on `--depth` streams for `--seconds`
after a warm-up, and prints one JSON line per N: per-rank GB/s, aggregate GB/s, the placement used.  bench.py reads the
committed result (profiles/r2_h2d_ceiling.json) to report `e2e.roofline` = {bound: "host-link", ...} per N.

No product kernels run here on purpose: this is the denominator, not the measurement.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time


def spread_device(local, world, n_visible):
    """Same placement rule as 3d_object_detection_b200.sharding.place_devices: alternate the two host bridges."""
    if world >= n_visible or n_visible % 2:
        return local
    half = n_visible // 2
    return (local % 2) * half + local // 2


def worker(rank, world, args, barrier, out_q):
    import torch
    n_vis = torch.cuda.device_count()
    dev_idx = spread_device(rank, world, n_vis) if args.placement == "spread" else rank
    torch.cuda.set_device(dev_idx)
    dev = torch.device("cuda", dev_idx)
    n_h2d = int(args.mb * 1e6) // 4
    n_d2h = int(args.d2h_mb * 1e6) // 4
    pipes = []
    for _ in range(args.depth):
        p = dict(stream=torch.cuda.Stream(dev), h_in=torch.empty(n_h2d, dtype=torch.float32).pin_memory(),
                 d_in=torch.empty(n_h2d, dtype=torch.float32, device=dev))
        p["h_in"].normal_()
        if n_d2h:
            p["d_out"] = torch.zeros(n_d2h, dtype=torch.float32, device=dev)
            p["h_out"] = torch.empty(n_d2h, dtype=torch.float32).pin_memory()
        pipes.append(p)

    def body(p):
        p["d_in"].copy_(p["h_in"], non_blocking=True)
        if n_d2h:
            p["h_out"].copy_(p["d_out"], non_blocking=True)

    for p in pipes:
        with torch.cuda.stream(p["stream"]):
            body(p)
        torch.cuda.synchronize(dev)
        p["graph"] = torch.cuda.CUDAGraph()
        with torch.cuda.graph(p["graph"], stream=p["stream"]):
            body(p)
    torch.cuda.synchronize(dev)

    def run_for(seconds):
        n = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            for _ in range(8 * args.depth):
                p = pipes[n % args.depth]
                with torch.cuda.stream(p["stream"]):
                    p["graph"].replay()
                n += 1
            torch.cuda.synchronize(dev)
        return n

    run_for(max(0.75, args.seconds / 2))          # the pinned-copy path needs ~0.5 s to reach its steady rate
    barrier.wait()
    cur = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(cur)
    for p in pipes:
        p["stream"].wait_event(e0)
    n = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < args.seconds:
        for _ in range(8 * args.depth):
            p = pipes[n % args.depth]
            with torch.cuda.stream(p["stream"]):
                p["graph"].replay()
            n += 1
        for p in pipes:                            # bounded queue depth, like a serving loop
            p["stream"].synchronize()
    for p in pipes:
        cur.wait_stream(p["stream"])
    e1.record(cur)
    torch.cuda.synchronize(dev)
    sec = e0.elapsed_time(e1) * 1e-3
    out_q.put(dict(rank=rank, device=dev_idx, copies=n, seconds=sec, h2d_gbs=n * n_h2d * 4 / sec / 1e9,
                   d2h_gbs=n * n_d2h * 4 / sec / 1e9))
    barrier.wait()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ns", default="1,2,4,8")
    ap.add_argument("--mb", type=float, default=15.36)
    ap.add_argument("--d2h-mb", type=float, default=0.0)
    ap.add_argument("--depth", type=int, default=3)
    ap.add_argument("--seconds", type=float, default=1.0)
    ap.add_argument("--placement", default="spread", choices=["spread", "linear"])
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    n_vis = torch.cuda.device_count()
    if n_vis == 0:
        raise SystemExit("h2d_ceiling: no CUDA device")
    ctx = mp.get_context("spawn")
    lines = []
    for n in [int(x) for x in args.ns.split(",")]:
        if n > n_vis:
            continue
        barrier, q = ctx.Barrier(n), ctx.Queue()
        procs = [ctx.Process(target=worker, args=(r, n, args, barrier, q)) for r in range(n)]
        for p in procs:
            p.start()
        res = sorted((q.get(timeout=300) for _ in range(n)), key=lambda r: r["rank"])
        for p in procs:
            p.join(timeout=60)
        line = dict(n_gpus=n, visible_gpus=n_vis, placement=args.placement, mb_per_copy=args.mb, d2h_mb=args.d2h_mb,
                    depth=args.depth, devices=[r["device"] for r in res],
                    per_rank_h2d_gbs=[round(r["h2d_gbs"], 2) for r in res],
                    aggregate_h2d_gbs=round(sum(r["h2d_gbs"] for r in res), 2),
                    aggregate_d2h_gbs=round(sum(r["d2h_gbs"] for r in res), 2),
                    min_rank_h2d_gbs=round(min(r["h2d_gbs"] for r in res), 2),
                    note="copy-only CUDA graphs on pinned buffers, no kernels; aggregate = sum over ranks of bytes / own time")
        print(json.dumps(line), flush=True)
        lines.append(line)
    if args.out:
        with open(args.out, "w") as fh:
            json.dump({"host_cores": os.cpu_count(), "lines": lines}, fh, indent=1)


if __name__ == "__main__":
    sys.exit(main())
