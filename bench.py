#!/usr/bin/env python
"""bench.py -- points/s and frames/s of the point-cloud front end (voxelise -> decorate -> BEV scatter).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--grid 20cm] [--batch 8]

A "step" is one pass of the hot path over one batch of B synthetic frames per GPU (BASELINE.json configs[1]:
seven_20cm/eight_20cm, batch 8 x 120k-point frames, 20 cm pillars).  Frames are independent, so N GPUs run
N replicas on disjoint frames with no collective on the data path (weak scaling).  Rank 0 prints ONE JSON
line.  `value` is device-resident throughput (inputs in HBM, whole step replayed as one CUDA graph); `e2e`
goes through the public Python API with host (pinned) points, H2D and D2H inside the timed region;
`roofline` is the dominant kernel (canvas fill) against the measured HBM copy bandwidth; `cpu_baseline` /
`--impl reference` time the CPU oracle port of the reference path on the host cores.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "voxelize+decorate+scatter throughput"
UNIT = "points/s"
FRAME_KINDS = ["S1", "S2", "S3"]          # clustered / ring (overflows max_voxels) / uniform (overflows)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--grid", default="20cm", help="20cm | 10cm | 16cm | nuscene or a config stem")
    ap.add_argument("--batch", type=int, default=0, help="frames per GPU per step (default: 8; 10cm: 32)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU-baseline sample budget")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-single", action="store_true", help="skip the single-frame drop-in latency measurement")
    ap.add_argument("--no-pfn", action="store_true", help="skip the fused-vs-torch PFN side measurement")
    ap.add_argument("--no-anchor-mask", action="store_true", help="skip the anchor-mask (row N2) side measurement")
    ap.add_argument("--no-postprocess", action="store_true", help="skip the detection post-processing (row N4) side measurement")
    ap.add_argument("--e2e-eager", action="store_true", help="e2e arm: launch every copy / kernel eagerly instead of one graph per pipeline")
    ap.add_argument("--no-affinity", action="store_true",
                    help="do not pin the rank to the host cores NVML reports as local to its GPU (NUMA-local pinned buffers)")
    ap.add_argument("--device-pipelines", type=int, default=0,
                    help="device-resident arm: front ends (workspace + outputs + graph each) in flight, batches dealt round-robin; "
                         "2 lets the latency-bound voxelise chain of batch k+1 run under the HBM-bound canvas fill of batch k (+7 %% "
                         "at 20 cm x 8, +11 %% at nuScenes x 8, neutral at 10 cm x 32), 3 adds another 8 %% on canvases below 1 GB "
                         "(nuScenes x 8); 0 (default) = 3 below 1 GB of canvas per batch, else 2; the strictly serial figure is "
                         "always reported under 'serial'")
    ap.add_argument("--chain-priority", action="store_true",
                    help="capture the voxelise chain of every device pipeline on a high-priority stream (see PillarFrontEnd.capture)")
    ap.add_argument("--pipelines", type=int, default=3, help="e2e: independent stream/workspace sets in flight")
    ap.add_argument("--min-seconds", type=float, default=0.0,
                    help="raise --steps until the timed region lasts at least this long (BASELINE configs[4]: >= 2 s)")
    ap.add_argument("--pinned-frames", action="store_true",
                    help="--sharded: hand the frames over as pinned torch tensors (no host staging copy) instead of pageable numpy")
    ap.add_argument("--stage-threads", type=int, default=0, help="--sharded: host threads per GPU for the pageable -> pinned copy (0 = the package's rule)")
    ap.add_argument("--sharded", action="store_true",
                    help="single-process mode: ONE process drives --gpus N GPUs through the package's ShardedFrontEnd "
                         "(one worker thread per GPU, numpy frames in, frame b -> GPU b mod N) instead of one rank per GPU")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def spread_device(local, world, n_visible):
    """GPU for local rank `local`: the package's bridge-aware placement (sharding.place_devices; PF_DEVICE_MAP=linear
    keeps LOCAL_RANK -> GPU)."""
    sh = importlib.import_module("3d_object_detection_b200.sharding")
    if world > n_visible:
        return local
    return sh.place_devices(world, n_visible, os.environ.get("PF_DEVICE_MAP", "spread"))[local]


def workload_config(grid, cfg, batch):
    """The `config` object of the JSON line: ONE definition for both arms, so that the driver's same_config check compares
    like with like.  Arm-specific details (pipelining, parallelism, measured voxel counts) live under `detail`."""
    return {"workload": f"{grid}: batch {batch} x {cfg['frame_points']}-point synthetic frames per GPU (S1/S2/S3 mix), "
                        f"voxelize+decorate+scatter, canvas ({batch},64,{cfg['_nx']},{cfg['_ny']}) fp32",
            "grid": grid, "batch_per_gpu": batch, "frame_points": cfg["frame_points"], "max_voxels": cfg["max_voxels"],
            "max_num_points": cfg["max_num_points"],
            "l2": f"each step writes a {4 * 64 * cfg['_nx'] * cfg['_ny'] * batch / 1e6:.0f} MB canvas (> 126 MB L2), which "
                  f"evicts the previous step's lines; no separate flush",
            "e2e_io": "host points in (pinned), voxel counts + coordinates out; the canvas stays on the GPU for the backbone"}


def load_workload(pfm, grid, batch_arg):
    cfg = pfm.builtin_config(grid)
    cfg["_name"] = grid
    geom = pfm.derive_grid(cfg)
    cfg["_nx"], cfg["_ny"] = int(geom.grid_size[0]), int(geom.grid_size[1])
    batch = batch_arg or (32 if "10cm" in grid else 8)
    return cfg, geom, batch


def make_batch(pfm, cfg, geom, batch, rank):
    """B seeded frames for this rank (kinds cycle S1,S2,S3), concatenated, + int32 offsets."""
    synth = importlib.import_module("3d_object_detection_b200.synth")
    cfgmod = importlib.import_module("3d_object_detection_b200.config")
    gkey = cfgmod.CONFIG_GRID.get(cfg["_name"], cfg["_name"])
    n = cfg["frame_points"]
    frames = []
    kinds = os.environ.get("PF_KINDS", "").split(",") if os.environ.get("PF_KINDS") else FRAME_KINDS   # probe aid
    for b in range(batch):
        kind = kinds[b % len(kinds)]
        seed = synth.frame_seed(50 + rank, b)
        f = synth.make_frame(kind, geom, n, seed, synth.S1_CELLS.get(gkey, 12000))
        frames.append(np.ascontiguousarray(f[:n], np.float32))
    sizes = [len(f) for f in frames]
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    return frames, np.concatenate(frames), offs


def bind_to_gpu_cores(index):
    """Best effort: restrict this process to the host cores NVML lists as near GPU `index`, BEFORE the pinned staging
    buffers are allocated, so that they land on the GPU's NUMA node.  Returns the cores used, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        near = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cores = sorted(near & os.sched_getaffinity(0))
        if cores:
            os.sched_setaffinity(0, cores)
            return cores
    except Exception:
        pass
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                break
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path (C, one frame per OpenMP thread) on the host cores
def cpu_arm(cfg, geom, frames, steps, warmup, budget_s, max_threads=None):
    """K steps, each step = the batch of B frames the GPU arm runs.  The port is frame-parallel (one frame per OpenMP
    thread), so m = cores // B steps run side by side (m * B frames per call) to occupy the host; a step is still B frames
    and `steps_run` counts steps.  Stops early (>= 3 calls) when `budget_s` is spent."""
    from oracle import oracle          # bench.py's cpu_baseline / --impl reference legs may run the oracle
    b = len(frames)
    cores = os.cpu_count() or 1
    if max_threads:
        cores = min(cores, max_threads)
    m = max(1, cores // b)
    threads = min(cores, m * b)
    consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)

    def tile(k):
        fr = frames * k
        sizes = [len(f) for f in fr]
        return (oracle.FrontendBatch(geom, cfg["max_voxels"], cfg["max_num_points"], frames[0].shape[1], k * b, 64),
                np.concatenate(fr), np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64),
                np.random.default_rng(1).standard_normal((k * b * cfg["max_voxels"], 64)).astype(np.float32))

    fb, pts, offs64, pillar = tile(m)
    for _ in range(max(1, (warmup + m - 1) // m)):
        fb.run(pts, offs64, consts, pillar, threads)
    done, total = 0, 0.0
    t_start = time.perf_counter()
    calls = 0
    while done < steps:
        k = min(m, steps - done)
        if k != m:
            fb, pts, offs64, pillar = tile(k)
            fb.run(pts, offs64, consts, pillar, min(threads, k * b))          # first touch of the new buffers, untimed
        t0 = time.perf_counter()
        fb.run(pts, offs64, consts, pillar, min(threads, k * b))
        total += time.perf_counter() - t0
        done += k
        calls += 1
        if budget_s and time.perf_counter() - t_start > budget_s and calls >= 3:
            break
    n_pts = sum(len(f) for f in frames) * done
    return {"value": n_pts / total, "unit": UNIT, "cores": threads, "kind": "port",
            "frames_per_s": b * done / total, "ms_per_step": 1e3 * total / done,
            "steps_run": done, "host_cores": os.cpu_count() or 1,
            "sample": f"{done} steps of {b} frames x {len(frames[0])} points ({cfg['_name']} grid), {m} steps side by side "
                      f"per call, C oracle port of numba points_to_voxels + decoration + scatter, one frame per thread on "
                      f"{threads} of {os.cpu_count() or 1} host cores"}


def cpu_post_arm(cfg, geom, reps=3):
    """cpu_baseline leg of the post-processing side measurement: the numpy restatement of Inference.infer_gpu's per-class
    loop (oracle.postprocess_frame) on ONE frame of the sparse population, one host core."""
    from oracle import oracle          # bench.py's cpu_baseline leg may run the oracle
    synth = importlib.import_module("3d_object_detection_b200.synth")
    fmap = (int(geom.grid_size[0]) // 2, int(geom.grid_size[1]) // 2)
    anchors, ranges = synth.anchors_3d(geom, fmap)
    cls, box, dirp, mask = synth.post_preds(1, len(anchors), 7, mean_logit=-9.6)
    lim = [float(x) for x in cfg.get("center_limit", [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0])]
    oracle.postprocess_frame(cls[0], box[0], dirp[0], mask[0], anchors, ranges, lim)
    t0 = time.perf_counter()
    for _ in range(reps):
        n = len(oracle.postprocess_frame(cls[0], box[0], dirp[0], mask[0], anchors, ranges, lim)[0])
    return {"ms_per_frame": 1e3 * (time.perf_counter() - t0) / reps, "cores": 1, "kind": "port", "detections": n,
            "sample": f"{reps} runs of one frame, sparse population, numpy restatement of Inference.infer_gpu's per-class loop"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the Python/numba reference
    cannot travel to the GPU box) on the host cores.  Rank 0 only.  Same `config`, `steps` and `warmup` as the b200 arm."""
    if rank != 0:
        return
    pfm = importlib.import_module("3d_object_detection_b200")
    cfg, geom, batch = load_workload(pfm, args.grid, args.batch)
    frames, _, _ = make_batch(pfm, cfg, geom, batch, 0)
    W = max(args.warmup, 3)
    r = cpu_arm(cfg, geom, frames, args.steps, W, budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": r["steps_run"], "warmup": W, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "frames_per_s": r["frames_per_s"],
            "config": workload_config(args.grid, cfg, batch),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "host_cores")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, local):
    import torch
    import torch.distributed as dist
    pfm = importlib.import_module("3d_object_detection_b200")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the front end has no CPU fallback)")
    local = spread_device(local, world, torch.cuda.device_count())
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bound = None if args.no_affinity else bind_to_gpu_cores(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cfg, geom, B = load_workload(pfm, args.grid, args.batch)
    n_pts_frame = cfg["frame_points"]
    frames, pts_cat, offs = make_batch(pfm, cfg, geom, B, rank)
    n_points = int(pts_cat.shape[0])
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---------------- device-resident arm: whole step as one CUDA graph ----------------
    # (a) strictly serial: one front end, one graph, step k+1 starts when step k has written its canvas;
    # (b) pipelined (the reported `value`): D front ends (own workspace, outputs and graph each) on D streams, batches
    #     dealt round-robin, so the HBM-bound canvas fill of batch k overlaps the latency-bound voxelisation of
    #     batch k+1 -- what a serving loop does.  Every batch still runs every kernel; K steps = K batches.
    fe = pfm.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n_pts_frame, device=local)
    pts_d = torch.from_numpy(pts_cat).to(dev)
    offs_d = torch.from_numpy(offs).to(dev)
    pillar = torch.randn((B * cfg["max_voxels"], 64), device=dev)      # PFN output stand-in (not on the path)
    fe.capture(pts_d, offs_d, pillar)
    for _ in range(W):
        fe.replay()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K):
        fe.replay()
    ev1.record()
    barrier()
    t_serial = ev0.elapsed_time(ev1) * 1e-3 / K        # seconds per step, strictly serial
    if args.min_seconds > 0 and t_serial * K < args.min_seconds:      # same K on every rank: the estimate is reduced with MAX
        est = importlib.import_module("3d_object_detection_b200.sharding").reduce_max_time(t_serial, dev)
        K = int(np.ceil(args.min_seconds / est * 1.05))
    counts = fe.voxel_num.cpu().numpy()
    bytes_alg = fe.bytes_algorithmic(n_points, counts)

    d_dev = args.device_pipelines if args.device_pipelines > 0 else (3 if 4 * 64 * cfg["_nx"] * cfg["_ny"] * B < 1e9 else 2)
    if d_dev > 1 and args.chain_priority:
        fe.capture(pts_d, offs_d, pillar, chain_priority=True)
    dpipes = [dict(fe=fe, stream=torch.cuda.Stream(dev))]
    for d in range(1, d_dev):
        fe_d = pfm.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n_pts_frame, device=local)
        fe_d.capture(pts_d, offs_d, pillar, chain_priority=args.chain_priority)
        dpipes.append(dict(fe=fe_d, stream=torch.cuda.Stream(dev)))
    cur = torch.cuda.current_stream(dev)

    def dev_step(i):
        p = dpipes[i % d_dev]
        with torch.cuda.stream(p["stream"]):
            p["fe"].replay()

    for i in range(max(W, d_dev)):
        dev_step(i)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0.record(cur)
    for p in dpipes:
        p["stream"].wait_event(ev0)
    for i in range(K):
        dev_step(i)
    for p in dpipes:
        cur.wait_stream(p["stream"])
    ev1.record(cur)
    barrier()
    t_dev = ev0.elapsed_time(ev1) * 1e-3
    for p in dpipes[1:]:
        assert p["fe"].voxel_num.cpu().numpy().tolist() == counts.tolist()

    # ---------------- per-stage timing (eager, CUDA events on the launch stream) ----------------
    KS = min(K, 300)
    st = [torch.cuda.Event(enable_timing=True) for _ in range(3 * KS)]
    for _ in range(3):
        fe.step(pts_d, offs_d, pillar_feats=pillar)
    torch.cuda.synchronize(dev)
    for i in range(KS):
        st[3 * i].record()
        fe.voxelize_decorate(pts_d, offs_d)
        st[3 * i + 1].record()
        fe.scatter(pillar)
        st[3 * i + 2].record()
    torch.cuda.synchronize(dev)
    t_vox = float(np.mean([st[3 * i].elapsed_time(st[3 * i + 1]) for i in range(KS)])) * 1e-3
    t_scat = float(np.mean([st[3 * i + 1].elapsed_time(st[3 * i + 2]) for i in range(KS)])) * 1e-3

    # the same fill launched back to back on its own (stand-alone pf_scatter on the batch's coordinates: k_map_index + the
    # fill), for the reading of roofline.frac: inside the step a launch starts with the L2 full of the gather's dirty
    # lines and has to write those back before its own bytes find room
    fill_alone = None
    if rank == 0:
        cnts = fe.voxel_num.cpu().numpy()
        rows = torch.cat([torch.arange(b * cfg["max_voxels"], b * cfg["max_voxels"] + int(cnts[b]), device=dev) for b in range(B)])
        co_c, fe_c = fe.coors[rows].contiguous(), pillar[rows].contiguous()
        sc_mod = pfm.PointPillarsScatter(B, geom.grid_size, 64)
        with torch.no_grad():
            for _ in range(3):
                sc_mod(fe_c, co_c)
            torch.cuda.synchronize(dev)
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n_fa = 20
            f0.record()
            for _ in range(n_fa):
                sc_mod(fe_c, co_c)
            f1.record()
            torch.cuda.synchronize(dev)
        fill_alone = f0.elapsed_time(f1) / n_fa * 1e-3
        del sc_mod, co_c, fe_c, rows

    # ---------------- next-row N1: the learned PFN in the loop, fused vs torch (reported beside, not the headline) ----
    with_pfn = None
    if not args.no_pfn:
        net = pfm.PointNet(4, cfg["voxel_size"], geom.offset).to(dev).eval()

        def unfused():
            fe.voxelize_decorate(pts_d, offs_d)
            fe.scatter(net.features_to_pillars(fe.features))

        def t_of(fn, n):
            with torch.no_grad():
                for _ in range(3):
                    fn()
                torch.cuda.synchronize(dev)
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                for _ in range(n):
                    fn()
                a1.record()
                torch.cuda.synchronize(dev)
            return a0.elapsed_time(a1) / n

        n_pfn = max(5, min(K, 40))
        fe_p = pfm.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n_pts_frame, device=local, want_voxels=False)
        fe_p.capture(pts_d, offs_d, pfn=net)          # the deployed form: one CUDA graph, no voxel / feature tensors
        with_pfn = {"fused_graph_ms_per_step": t_of(fe_p.replay, max(n_pfn, 20)),
                    "fused_ms_per_step": t_of(lambda: fe.step(pts_d, offs_d, pfn=net), n_pfn),
                    "torch_pfn_ms_per_step": t_of(unfused, max(3, n_pfn // 4)),
                    "note": "random-init PointNet.eval(); fused_graph = graph replay of voxelise + decorate + PFN + fill without "
                            "materialising voxels / decorated rows; the other two eager; fused = pf_voxelize_pfn (decorate + Conv1d 9->64 + folded BN "
                            "+ ReLU + max in the gather kernel) + fill; torch = CUDA decoration -> pfn_layers in torch -> fill"}
        del net, fe_p

    # ---------------- next-row N2: anchor occupancy mask of the batch from the device coordinates ------------------
    anchor = None
    if rank == 0 and not args.no_anchor_mask:
        synth = importlib.import_module("3d_object_detection_b200.synth")
        amod = importlib.import_module("3d_object_detection_b200.anchor_mask")
        fmap = (int(geom.grid_size[0]) // 2, int(geom.grid_size[1]) // 2)         # feature_map_size of the configs
        am = amod.AnchorMask(synth.anchors_bv(geom, fmap), voxel_size=np.asarray(cfg["voxel_size"], np.float32),
                             offset=geom.offset, grid_size=geom.grid_size, device=local)
        fe.voxelize_decorate(pts_d, offs_d, keep_cell_map=False)      # no scatter follows: do not keep the cell map
        for _ in range(3):
            mask_d = fe.anchor_mask(am)
        torch.cuda.synchronize(dev)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_am = max(5, min(K, 40))
        a0.record()
        for _ in range(n_am):
            mask_d = fe.anchor_mask(am)
        a1.record()
        torch.cuda.synchronize(dev)
        anchor = {"ms_per_batch": a0.elapsed_time(a1) / n_am, "anchors": am.n_anchors, "frames": B,
                  "set_fraction": float(mask_d.float().mean().item()),
                  "note": "pf_anchor_mask on the device coordinates of the batch: memset + count + two parallel scans + "
                          "four-corner lookup, mask (frames, anchors) bool left on the device"}
        del am, mask_d

    # ---------------- next-row N4: detection post-processing of the batch (score mask, top-k, decode, NMS, range mask) --
    post = None
    if rank == 0 and not args.no_postprocess:
        try:
            synth = importlib.import_module("3d_object_detection_b200.synth")
            imod = importlib.import_module("3d_object_detection_b200.inference")
            fmap = (int(geom.grid_size[0]) // 2, int(geom.grid_size[1]) // 2)
            anchors3, ranges3 = synth.anchors_3d(geom, fmap)
            lim = [float(x) for x in cfg.get("center_limit", [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0])]
            Bp = max(1, min(B, 8, 12_000_000 // len(anchors3)))       # bounds the host time spent generating head outputs
            pp = imod.PostProcessor(anchors3, ranges3, lim, max_frames=Bp, device=local)
            # synthetic head outputs: ~300 anchors per class and frame above the 0.05 score threshold ("sparse", a trained
            # head) and ~6 % of all anchors ("dense"); random-init heads have no meaningful population of their own
            post = {"anchors": len(anchors3), "frames": Bp, "launches": 6}
            for label, mean_logit in (("sparse", -9.6), ("dense", -6.0)):
                ins = [torch.from_numpy(x).to(dev) for x in synth.post_preds(Bp, len(anchors3), 7, mean_logit=mean_logit)]
                for _ in range(3):
                    pp(*ins)
                torch.cuda.synchronize(dev)
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n_pp = max(5, min(K, 40))
                a0.record()
                for _ in range(n_pp):
                    pp(*ins)
                a1.record()
                torch.cuda.synchronize(dev)
                post[label] = {"ms_per_batch": a0.elapsed_time(a1) / n_pp,
                               "detections_per_frame": float(pp.det_num[:Bp].float().mean().item())}
                del ins
            post["note"] = ("pf_postprocess: anchor mask + sigmoid + score mask -> candidate lists, radix select + bitonic "
                            "sort of the top 1000 per class and frame, box decode + standup boxes, NMS, direction flip + "
                            "range mask; detections left on the device")
            del pp
        except Exception as exc:      # a side measurement must not take the headline down with it
            post = {"error": f"{type(exc).__name__}: {exc}"}

    # ---------------- drop-in single-frame latency: the reference's own call sequence (InferData.get + forward) -------
    single = None
    if rank == 0 and not args.no_single:
        vg = pfm.VoxelGenerator(dict(cfg), device=local, max_points_per_frame=n_pts_frame)
        net1 = pfm.PointNet(4, cfg["voxel_size"], geom.offset).to(dev).eval()
        scat1 = pfm.PointPillarsScatter(1, geom.grid_size, 64)
        frame0 = frames[0]
        to_torch = importlib.import_module("3d_object_detection_b200.collate").example_convert_to_torch

        def one_frame(device_outputs):
            v, c, n = vg.generate(frame0, device_outputs=device_outputs)       # numpy in (pageable), like dataset.py:215
            if not device_outputs:                      # the package's drop-in for example_convert_to_torch, utils.py:7-20
                ex = to_torch({"voxels": v, "coordinates": c, "num_points_per_voxel": n}, device=dev)
                v, c, n = ex["voxels"], ex["coordinates"], ex["num_points_per_voxel"]
            with torch.no_grad():
                canvas1 = scat1(net1(v, n, c).reshape(-1, 64), c)
            torch.cuda.synchronize(dev)
            return canvas1

        lat = {}
        for key, dev_out in (("numpy_outputs", False), ("device_outputs", True)):
            for _ in range(5):
                one_frame(dev_out)
            t0 = time.perf_counter()
            reps = 30
            for _ in range(reps):
                one_frame(dev_out)
            lat[key + "_ms_per_frame"] = 1e3 * (time.perf_counter() - t0) / reps
        def fused_frame():
            canvas1, _, _ = vg.generate_and_scatter(frame0, net1)
            torch.cuda.synchronize(dev)
            return canvas1

        def gen_only(dev_out):
            return vg.generate(frame0, device_outputs=dev_out)

        for key, fn in (("fused_generate_and_scatter_ms_per_frame", fused_frame),
                        ("generate_only_numpy_ms_per_frame", lambda: gen_only(False)),
                        ("generate_only_device_ms_per_frame", lambda: gen_only(True))):
            with torch.no_grad():
                for _ in range(5):
                    fn()
                t0 = time.perf_counter()
                for _ in range(30):
                    fn()
            lat[key] = 1e3 * (time.perf_counter() - t0) / 30
        single = dict(lat, frame_points=int(len(frame0)),
                      api="numpy_outputs / device_outputs: the reference's call sequence, VoxelGenerator.generate(numpy points) "
                          "[-> example_convert_to_torch] -> PointNet.eval() (fused PFN) -> PointPillarsScatter, batch 1, host wall "
                          "clock incl. staging, H2D, the voxel-count read-back and a final synchronize; fused_generate_and_scatter: "
                          "the same work as ONE CUDA graph (VoxelGenerator.generate_and_scatter); generate_only_*: generate() alone")
        vg.close()

    # ---------------- end-to-end arm: public API, pinned host points in, coords/counts out ----------------
    depth = max(1, args.pipelines)
    pipes = []
    for d in range(depth):
        p_fe = dpipes[d]["fe"] if d < d_dev else pfm.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n_pts_frame,
                                                                     device=local)
        pipes.append(dict(fe=p_fe, stream=torch.cuda.Stream(dev),
                          h_pts=torch.from_numpy(pts_cat).pin_memory(), d_pts=torch.empty_like(pts_d),
                          h_cnt=torch.empty((B,), dtype=torch.int32).pin_memory(),
                          h_coors=torch.empty((B * cfg["max_voxels"], 4), dtype=torch.int32).pin_memory()))
    h2d = pts_cat.nbytes
    d2h = pipes[0]["h_cnt"].numel() * 4 + pipes[0]["h_coors"].numel() * 4

    def e2e_body(p):
        p["d_pts"].copy_(p["h_pts"], non_blocking=True)
        p["fe"].step(p["d_pts"], offs_d, pillar_feats=pillar)
        p["h_cnt"].copy_(p["fe"].voxel_num, non_blocking=True)
        p["h_coors"].copy_(p["fe"].coors, non_blocking=True)

    # each pipeline's whole step (H2D of the pinned points, the eight kernels, D2H of counts and coordinates) is one
    # CUDA graph on the pipeline's stream, like the device arm; --e2e-eager launches it call by call instead
    if not args.e2e_eager:
        for p in pipes:
            with torch.cuda.stream(p["stream"]):
                for _ in range(2):
                    e2e_body(p)
            torch.cuda.synchronize(dev)
            p["graph"] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(p["graph"], stream=p["stream"]):
                e2e_body(p)
        torch.cuda.synchronize(dev)

    def e2e_step(i):
        p = pipes[i % depth]
        with torch.cuda.stream(p["stream"]):
            if args.e2e_eager:
                e2e_body(p)
            else:
                p["graph"].replay()

    # the PCIe link and the pinned-copy path need far longer to reach steady state than the kernels do
    # (tools/e2e_probe.py: the first ~50 copies run at 20-27 GB/s, later ones at ~55 GB/s): untimed warm-up
    # ... and every pipeline has its own pinned buffer: warm up by step count AND by time (>= 100 steps per pipeline,
    # >= 0.75 s), untimed
    w_e2e = max(W, 100 * depth)
    t_w0 = time.perf_counter()
    n_w = 0
    while n_w < w_e2e or time.perf_counter() - t_w0 < 0.75:
        for _ in range(depth * 8):
            e2e_step(n_w)
            n_w += 1
        torch.cuda.synchronize(dev)
    w_e2e = n_w
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(cur)
    for p in pipes:
        p["stream"].wait_event(e0)
    for i in range(K):
        e2e_step(i)
    for p in pipes:
        cur.wait_stream(p["stream"])
    e1.record(cur)
    barrier()
    t_e2e_local = e0.elapsed_time(e1) * 1e-3
    clocks = sampler.stop()
    assert pipes[0]["h_cnt"].numpy().tolist() == counts.tolist()

    # ---------------- reduce over ranks: MAX time, SUM work ----------------
    sh = importlib.import_module("3d_object_detection_b200.sharding")
    t_dev = sh.reduce_max_time(t_dev, dev)
    t_e2e = sh.reduce_max_time(t_e2e_local, dev)
    tot_pts, tot_frames = sh.reduce_sum([n_points * K, B * K], dev)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    fill_variant = int(importlib.import_module("3d_object_detection_b200._lib").load().pf_debug_set_fill_variant(
        int(os.environ["PF_FILL_VARIANT"]) if os.environ.get("PF_FILL_VARIANT", "").isdigit() else -1))
    return dict(fill_alone=fill_alone, fill_variant=fill_variant, cfg=cfg, geom=geom, frames=frames, B=B, K=K, W=W, t_dev=t_dev, t_e2e=t_e2e, t_vox=t_vox,
                t_scat=t_scat, with_pfn=with_pfn, anchor=anchor, post=post, single=single, bound=bound, w_e2e=w_e2e, t_serial=t_serial, d_dev=d_dev, counts=counts, bytes_alg=bytes_alg, n_points=n_points, tot_pts=tot_pts,
                tot_frames=tot_frames, clocks=clocks, h2d=h2d, d2h=d2h, fe=fe, depth=depth, world=world)


def run_sharded(args):
    """--sharded: the package-level multi-GPU runner.  One process, one worker thread per GPU (ShardedFrontEnd), host
    numpy frames in (pageable, as a dataset hands them over), frame b -> GPU b mod N, pinned staging + one CUDA graph per
    worker step, voxel counts back.  Wall clock over K steps is the end-to-end figure; the device figure is the MAX over
    workers of their summed CUDA-event step times."""
    import torch
    pfm = importlib.import_module("3d_object_detection_b200")
    G = args.gpus
    cfg, geom, B = load_workload(pfm, args.grid, args.batch)
    per_dev = [make_batch(pfm, cfg, geom, B, r)[0] for r in range(G)]
    frames = [per_dev[b % G][b // G] for b in range(G * B)]          # frame b lands on worker b mod G
    n_points = sum(len(f) for f in frames)
    nbytes = sum(f.nbytes for f in frames)
    if args.pinned_frames:
        frames = [torch.from_numpy(f).pin_memory() for f in frames]
    K, W = args.steps, max(args.warmup, 3)
    with pfm.ShardedFrontEnd(cfg, devices=G, frames_per_device=B, max_points_per_frame=cfg["frame_points"], depth=3,
                             stage_threads=args.stage_threads or None) as sfe:
        for _ in range(W):
            sfe.run(frames)
        sampler = ClockSampler(sfe.devices[0])
        sampler.start()
        t0 = time.perf_counter()
        pending = [sfe.submit(frames) for _ in range(K)]
        res = [[f.result() for f in futs] for futs in pending]
        wall = time.perf_counter() - t0
        clocks = sampler.stop()
        dev_s = max(sum(step[i].device_ms for step in res) for i in range(G)) * 1e-3    # graphs of one worker may overlap
        counts = [int(c) for r in res[-1] for c in r.counts]
        devices = list(sfe.devices)
    line = {"metric": METRIC, "value": n_points * K / wall, "unit": UNIT, "n_gpus": G, "steps": K, "warmup": W,
            "ms_per_step": 1e3 * wall / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "frames_per_s": G * B * K / wall,
            "mode": "single process, ShardedFrontEnd: one worker thread per GPU, no torch.distributed",
            "config": workload_config(args.grid, cfg, B),
            "detail": {"devices": devices, "voxels_per_frame_last_step": counts,
                       "value_is": "points / host wall clock over all K steps (this mode starts from pageable host frames, so "
                                   "there is no device-resident figure); a worker's graphs overlap (one stream per slot), its "
                                   "busiest-GPU sum of CUDA-event graph times per step is below",
                       "max_worker_graph_ms_per_step": 1e3 * dev_s / K},
            "e2e": {"value": n_points * K / wall, "unit": UNIT, "frames_per_s": G * B * K / wall, "ms_per_step": 1e3 * wall / K,
                    "h2d_bytes_per_step": int(nbytes // G), "d2h_bytes_per_step": 4 * B,
                    "frames": "pinned torch tensors (no staging copy)" if args.pinned_frames else "pageable numpy arrays (staged into pinned memory by pf_stage_copy)",
                    "api": "ShardedFrontEnd.submit(list of host frames): [host staging copy into pinned memory,] one H2D per frame, "
                           "voxelise + decorate + canvas fill, D2H of the voxel counts; host wall clock over all K steps"},
            "gpu_launches": K * G * (pfm.PillarFrontEnd.n_kernels_voxelize(cfg["frame_points"]) + 1),
            "clocks": clocks}
    print(json.dumps(line), flush=True)


def peak_hbm_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def host_link_ceiling(n_gpus):
    """Aggregate copy-only pinned H2D GB/s for N concurrent GPUs from the committed probe (tools/h2d_ceiling.py), else None."""
    for name in ("r2_h2d_ceiling.json", f"r2_h2d_ceiling_n{n_gpus}.json"):
        try:
            d = json.loads((ROOT / "profiles" / name).read_text())
            for ln in d["lines"]:
                if ln["n_gpus"] == n_gpus:
                    return dict(ln, source=f"profiles/{name}")
        except Exception:
            continue
    return None


def fill_kernel_name(variant, cells):
    """Which canvas-fill kernel pf_scatter_from_map launched (the rule of run_fill, csrc/pf_scatter.cu)."""
    if variant < 0:
        variant = 6 if cells >= 1 << 25 else (0 if cells <= 1310720 else 1)
    names = {0: "k_canvas_fill<1,1,4> (one warp per 128 cells, 16-byte streaming stores)",
             1: "k_canvas_fill<1,1,8> (one warp per 256 cells, 32-byte streaming stores)",
             6: "k_canvas_fill<1,1,16,4,2304> (one warp per 512 cells, 32-byte streaming stores)"}
    return names.get(variant, f"canvas fill variant {variant} (pf_debug_set_fill_variant)") + ", pf_scatter_from_map"


def ncu_traffic(grid, batch):
    """DRAM bytes of one canvas-fill launch from the committed ncu capture of this workload (shipped kernel set), else None."""
    try:
        t = json.loads((ROOT / "profiles" / "r2_traffic.json").read_text()).get(f"{grid}:{batch}")
        return None if t is None else float(t["dram_read_bytes"]) + float(t["dram_write_bytes"])
    except Exception:
        return None


def main():
    args = parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.sharded:
        if rank == 0:
            run_sharded(args)
        return
    r = run_b200(args, rank, world, local)
    if rank != 0:
        return
    cfg, B, K = r["cfg"], r["B"], r["K"]
    value = r["tot_pts"] / r["t_dev"]
    e2e_value = r["tot_pts"] / r["t_e2e"]
    peak, peak_src = peak_hbm_gbs()
    scat_bytes = r["bytes_alg"]["scatter"]
    achieved = scat_bytes / r["t_scat"] / 1e9
    fe = r["fe"]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": r["world"], "steps": K, "warmup": r["W"],
        "ms_per_step": 1e3 * r["t_dev"] / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "frames_per_s": r["tot_frames"] / r["t_dev"],
        "device_pipelines": r["d_dev"],
        "serial": {"ms_per_step": 1e3 * r["t_serial"], "value": r["n_points"] / r["t_serial"], "unit": UNIT,
                   "note": "rank 0, one front end, step k+1 starts after step k's canvas is complete"},
        "config": workload_config(args.grid, cfg, B),
        "detail": {"voxels_per_frame": r["counts"].tolist(), "parallelism": f"frame-sharded x{r['world']}, no collective",
                   "pipelining": f"{r['d_dev']} batches in flight per GPU (own workspace, outputs and CUDA graph each); "
                                 f"serial figure under 'serial'", "fill_variant": r["fill_variant"]},
        "e2e": {"value": e2e_value, "unit": UNIT, "frames_per_s": r["tot_frames"] / r["t_e2e"],
                "ms_per_step": 1e3 * r["t_e2e"] / K, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                "api": "PillarFrontEnd.step on pinned host points; D2H of voxel_num + coordinates; canvas stays "
                       "on the GPU for the backbone, as in the reference", "pipelines": r["depth"], "warmup_steps": r["w_e2e"],
                "launch": "eager" if args.e2e_eager else "one CUDA graph per pipeline (H2D + kernels + D2H)"},
        "gpu_launches": K * (fe.n_kernels_voxelize(cfg["frame_points"]) + fe.N_KERNELS_SCATTER),
        "clocks": r["clocks"],
        "stages_ms": {"voxelize_decorate": 1e3 * r["t_vox"], "scatter": 1e3 * r["t_scat"]},
        "bytes_per_step": r["bytes_alg"],
        "device_of_rank0": r["fe"].device.index,
        "host_binding": {"cores": len(r["bound"]), "first": r["bound"][0], "last": r["bound"][-1]} if r["bound"] else None,
        "with_pfn": r["with_pfn"],
        "anchor_mask": r["anchor"],
        "postprocess": r["post"],
        "single_frame_api": r["single"],
        "e2e_hbm_frac": r["bytes_alg"]["e2e"] / (r["t_dev"] / K) / 1e9 / peak,
        "e2e_hbm_frac_serial": r["bytes_alg"]["e2e"] / r["t_serial"] / 1e9 / peak,
        "roofline": {"kernel": fill_kernel_name(r["fill_variant"], B * int(r["geom"].grid_size[0]) * int(r["geom"].grid_size[1])),
                     "bound": "hbm", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(args.grid, B),
                     "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / 8000.0,
                     "bytes_per_launch": scat_bytes, "ms_per_launch": 1e3 * r["t_scat"],
                     "back_to_back": None if not r["fill_alone"] else {
                         "ms_per_launch": 1e3 * r["fill_alone"], "frac": scat_bytes / r["fill_alone"] / 1e9 / peak,
                         "note": "the same canvas through the stand-alone pf_scatter (k_map_index + the same fill kernel, fresh canvas "
                                 "allocation included), 20 launches back to back: no gather in front, so the L2 does not start full of "
                                 "another kernel's dirty lines"}},
    }
    link = host_link_ceiling(r["world"])
    if link is not None:
        # the timing rule is MAX over ranks, so the rank on the slowest host bridge sets the pace: the ceiling of an N-rank
        # e2e step is N x the slowest rank's copy-only rate (at N = 8 four GPUs share each bridge unevenly: 20.5 / 36 GB/s)
        peak = r["world"] * link["min_rank_h2d_gbs"]
        got = r["world"] * r["h2d"] / (r["t_e2e"] / K) / 1e9
        line["e2e"]["roofline"] = {"bound": "host-link", "peak_gbs": peak, "achieved_gbs": got, "frac": got / peak,
                                   "aggregate_copy_only_gbs": link["aggregate_h2d_gbs"], "slowest_rank_copy_only_gbs": link["min_rank_h2d_gbs"],
                                   "peak_source": link["source"],
                                   "note": "pinned host->device GB/s of the e2e arm (all ranks) against N x the slowest rank of N "
                                           "copy-only streams with no kernels running (tools/h2d_ceiling.py); the e2e arm also moves "
                                           "its D2H bytes over the same link, the probe does not"}
    if not args.no_cpu_baseline and r["world"] >= 1:
        c = cpu_arm(cfg, r["geom"], r["frames"], steps=50, warmup=1, budget_s=args.cpu_seconds)
        line["cpu_baseline"] = {k: c[k] for k in ("value", "unit", "cores", "kind", "sample", "frames_per_s", "host_cores")}
        c1 = cpu_arm(cfg, r["geom"], r["frames"][:1], steps=6, warmup=1, budget_s=min(6.0, args.cpu_seconds), max_threads=1)
        line["cpu_baseline_1thread"] = {k: c1[k] for k in ("value", "unit", "cores", "kind", "sample", "frames_per_s")}
        if isinstance(line.get("postprocess"), dict) and "sparse" in line["postprocess"]:
            try:
                line["postprocess"]["cpu_baseline"] = cpu_post_arm(cfg, r["geom"])
            except Exception as exc:
                line["postprocess"]["cpu_baseline"] = {"error": f"{type(exc).__name__}: {exc}"}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
