"""Frame-wise sharding across GPUs: frames are independent, so rank r of G owns a subset of
the frames of every batch and runs its own PillarFrontEnd replica; there is no collective on the data path.
Only timings / counters are reduced, over whatever process group the launcher set up (NCCL on GPUs, gloo in
the CPU tests)."""
from __future__ import annotations


def shard_frames(n_frames: int, rank: int, world_size: int, mode: str = "block"):
    """Indices of the frames rank `rank` owns.  "block": contiguous ceil-split; "cyclic": b -> b mod G."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if mode == "cyclic":
        return list(range(rank, n_frames, world_size))
    if mode != "block":
        raise ValueError(f"unknown sharding mode {mode!r}")
    base, extra = divmod(n_frames, world_size)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def reduce_max_time(seconds: float, device=None) -> float:
    """MAX over ranks of a per-rank duration (the multi-GPU timing rule); identity without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(values, device=None):
    """SUM over ranks of a list of counters (frames, points, bytes)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [float(v) for v in values]
    t = torch.tensor(list(values), dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


# ---------------------------------------------------------------------------------------------------------
# Package-level multi-GPU runner: one worker thread per GPU, frames b -> GPU b mod G,
# no NCCL.  The reference pins everything to cuda:0 (train.py:27, 190, 276); this is the frame-sharded replacement.

def place_devices(n_workers: int, n_visible: int, mode: str = "spread"):
    """GPU index for each of `n_workers` workers.  "spread": when fewer workers than GPUs run, alternate between the
    two halves of the box (0, n/2, 1, n/2+1, ...) so that 2 or 4 workers do not all hang off the same host bridge
    (an 8-GPU HGX box has two, ~105 GB/s of host traffic each: profiles/r1_topology_8gpu.txt); "linear": 0, 1, 2, ..."""
    if n_workers < 1 or n_workers > n_visible:
        raise ValueError(f"{n_workers} workers for {n_visible} visible GPUs")
    if mode == "linear" or n_workers >= n_visible or n_visible % 2:
        return list(range(n_workers))
    if mode != "spread":
        raise ValueError(f"unknown placement {mode!r}")
    half = n_visible // 2
    return [(r % 2) * half + r // 2 for r in range(n_workers)]


class ShardResult:
    """What one GPU produced for its share of a submitted batch.  `canvas` (n, C, nx, ny), `coors` (n * max_voxels, 4)
    and `voxel_num` (n,) are CUDA tensors on `device`, views of the worker's buffers: valid until that worker has been
    handed `depth - 1` further batches (the worker stages batch k+1 while batch k runs, so depth >= 2; the default is 3).
    `counts` is `voxel_num` on the host."""

    __slots__ = ("device", "frame_ids", "canvas", "coors", "voxel_num", "counts", "device_ms")

    def __init__(self, device, frame_ids, canvas, coors, voxel_num, counts, device_ms):
        self.device, self.frame_ids, self.canvas, self.coors = device, frame_ids, canvas, coors
        self.voxel_num, self.counts, self.device_ms = voxel_num, counts, device_ms


class _GpuWorker:
    """One GPU: `depth` PillarFrontEnd replicas (workspace, outputs, pinned staging, one stream and one CUDA graph
    each), used in turn; the graph holds H2D of the staged frames + offsets, voxelise + decorate (+ fused PFN), canvas
    fill, D2H of the voxel counts; the points themselves are copied in front of it, one copy per frame."""

    def __init__(self, config, device, frames_per_device, max_points_per_frame, n_channels, pfn, depth, stage_threads):
        import copy

        import torch

        from . import _lib
        from .frontend import PillarFrontEnd
        self.torch, self._lib = torch, _lib
        self.device = int(device)
        self.B, self.cap, self.C = int(frames_per_device), int(max_points_per_frame), int(n_channels)
        self.threads = stage_threads
        dev = torch.device("cuda", self.device)
        self.dev = dev
        self.F = int(config.get("num_point_features", 4))
        self.slots = []
        with torch.cuda.device(dev):
            self.pfn = copy.deepcopy(pfn).to(dev).eval() if pfn is not None else None
            for _ in range(depth):
                fe = PillarFrontEnd(config, max_batch=self.B, max_points_per_frame=self.cap, device=self.device,
                                    n_channels=self.C, want_voxels=False)
                s = dict(fe=fe, h_pts=torch.empty((self.B * self.cap, self.F), dtype=torch.float32).pin_memory(),
                         h_offs=torch.zeros((self.B + 1,), dtype=torch.int32).pin_memory(),
                         h_cnt=torch.zeros((self.B,), dtype=torch.int32).pin_memory(),
                         d_pts=torch.empty((self.B * self.cap, self.F), dtype=torch.float32, device=dev),
                         d_offs=torch.zeros((self.B + 1,), dtype=torch.int32, device=dev),
                         pillar=None if self.pfn is not None else torch.zeros((self.B * fe.max_voxels, self.C), device=dev),
                         e0=torch.cuda.Event(enable_timing=True), e1=torch.cuda.Event(enable_timing=True),
                         stream=torch.cuda.Stream(dev))       # one stream per slot: the H2D of batch k+1 runs under the kernels of batch k
                s["np_pts"], s["np_offs"], s["np_cnt"] = s["h_pts"].numpy(), s["h_offs"].numpy(), s["h_cnt"].numpy()
                with torch.cuda.stream(s["stream"]), torch.no_grad():
                    self._body(s)
                    s["stream"].synchronize()
                    s["graph"] = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(s["graph"], stream=s["stream"]):
                        self._body(s)
                self.slots.append(s)
        self.turn = 0

    def _body(self, s):
        # the points are copied in front of the graph, frame by frame, from wherever they live (launch())
        s["d_offs"].copy_(s["h_offs"], non_blocking=True)
        if self.pfn is not None:
            s["fe"].step(s["d_pts"], s["d_offs"], pfn=self.pfn)
        else:
            s["fe"].step(s["d_pts"], s["d_offs"], pillar_feats=s["pillar"])
        s["h_cnt"].copy_(s["fe"].voxel_num, non_blocking=True)

    def launch(self, frame_ids, frames):
        """Move `frames` (at most B arrays of shape (N, F), float32) onto the GPU and enqueue the next slot's graph;
        returns a token for :meth:`finish`.  Does not wait for the GPU.  A frame may be
          * a numpy array (pageable host memory, what a dataset hands over): copied into the slot's pinned staging
            buffer, then one asynchronous H2D;
          * a pinned torch CPU tensor: one asynchronous H2D straight from it, no staging copy (the caller must not
            rewrite it before the batch has finished);
          * a CUDA tensor on this GPU: a device-to-device copy.
        All of it is ONE library call (pf_upload_frames), so the worker thread gives the interpreter lock up for the whole
        copy phase."""
        import ctypes as C
        torch = self.torch
        s = self.slots[self.turn % len(self.slots)]
        self.turn += 1
        n = len(frames)
        if n > self.B:
            raise ValueError(f"{n} frames for a worker built for {self.B}")
        ptrs, rows, kinds, keep = (C.c_void_p * self.B)(), (C.c_int64 * self.B)(), (C.c_int32 * self.B)(), []
        total = 0
        for b, f in enumerate(frames):
            is_t = torch.is_tensor(f)
            if (f.dtype != torch.float32) if is_t else (str(getattr(f, "dtype", None)) != "float32"):
                raise TypeError(f"frames must be float32, got {getattr(f, 'dtype', type(f).__name__)}")
            m = int(f.shape[0])
            if f.ndim != 2 or int(f.shape[1]) != self.F:
                raise ValueError(f"frame of shape {tuple(f.shape)} for {self.F} floats per point")
            total += m
            if is_t:
                f = f.contiguous()
                if f.is_cuda and f.device.index != self.device:
                    f = f.to(self.dev)
                ptrs[b], kinds[b] = f.data_ptr(), 2 if f.is_cuda else (1 if f.is_pinned() else 0)
            else:
                if not f.flags.c_contiguous:
                    f = f.copy()
                ptrs[b], kinds[b] = f.ctypes.data, 0
            keep.append(f)                                   # the (possibly re-laid-out) sources must outlive the copies
            rows[b] = m
        if total > self.B * self.cap:
            raise ValueError(f"{total} points for a worker built for {self.B} x {self.cap}")
        if any(int(r) > self.cap for r in rows):
            raise ValueError(f"a frame exceeds the capacity of {self.cap} points")
        torch.cuda.set_device(self.dev)                      # thread-local: this worker thread only
        torch.cuda.set_stream(s["stream"])
        st = s["stream"]
        s["e0"].record(st)
        self._lib.check(self._lib.load().pf_upload_frames(
            s["d_pts"].data_ptr(), s["h_pts"].data_ptr(), ptrs, rows, kinds, n, self.B, 4 * self.F, self.threads,
            s["h_offs"].data_ptr(), st.cuda_stream), "pf_upload_frames")
        s["graph"].replay()
        s["e1"].record(st)
        s["keep"] = keep
        return s, list(frame_ids), n

    def finish(self, token):
        s, frame_ids, n = token
        s["e1"].synchronize()
        s["keep"] = None
        fe = s["fe"]
        return ShardResult(self.device, frame_ids, fe.canvas[:n], fe.coors[:n * fe.max_voxels], fe.voxel_num[:n],
                           s["np_cnt"][:n].copy(), s["e0"].elapsed_time(s["e1"]))

    def run(self, frame_ids, frames):
        return self.finish(self.launch(frame_ids, frames))

    def close(self):
        for s in self.slots:
            s["graph"] = None
            s["fe"].close()
        self.slots = []


class ShardedFrontEnd:
    """Frame-sharded front end over the GPUs of one box: ``submit(frames)`` deals frame ``b`` to GPU ``b mod G`` (one
    worker thread per GPU -- the ctypes calls and CUDA waits release the GIL), every GPU runs its own replica of the
    whole path on its share (pinned staging, one CUDA graph per step), and nothing crosses between GPUs: no NCCL, the
    canvases stay where that GPU's backbone replica would read them.

        sfe = ShardedFrontEnd(config, devices=8, frames_per_device=8, pfn=point_net)
        futures = sfe.submit(list_of_64_numpy_frames)        # one Future per GPU -> ShardResult
        results = sfe.run(list_of_64_numpy_frames)           # the same, waited for, in GPU order

    `pfn`: a PointNet (eval mode) replicated onto every GPU and fused into the voxeliser; without it the canvas is
    filled from each worker's `pillar` stand-in tensor (what bench.py's headline does: the PFN is not on the path)."""

    def __init__(self, config, devices=None, frames_per_device=8, max_points_per_frame=None, n_channels=64, pfn=None,
                 depth=3, placement="spread", stage_threads=None, _worker_factory=None):
        import queue
        import threading
        if _worker_factory is None:
            import torch
            n_vis = torch.cuda.device_count()
            if n_vis == 0:
                from . import _lib
                raise _lib.PillarFrontError("no CUDA device: the pillar front end has no CPU fallback")
            if devices is None:
                devices = n_vis
            if isinstance(devices, int):
                devices = place_devices(devices, n_vis, placement)
            if stage_threads is None:      # host threads per worker for the pageable -> pinned copy: the workers already run
                import os                 # side by side, so from 4 GPUs on each copies with its own thread only
                cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
                stage_threads = 1 if len(devices) >= 4 else max(1, min(4, cores // (2 * len(devices))))
            cap = int(max_points_per_frame or config.get("frame_points", 0) or 0)
            if cap < 1:
                raise ValueError("max_points_per_frame is required")
            # graphs are captured here, one device after the other (capture is not safe against concurrent CUDA calls)
            self.workers = [_GpuWorker(dict(config), d, frames_per_device, cap, n_channels, pfn, depth, stage_threads)
                            for d in devices]
        else:
            self.workers = [_worker_factory(d) for d in devices]
        self.devices = list(devices)
        self.B = int(frames_per_device)
        self._queues = [queue.Queue() for _ in self.workers]
        self._threads = [threading.Thread(target=self._loop, args=(i,), daemon=True) for i in range(len(self.workers))]
        for t in self._threads:
            t.start()

    def _loop(self, i):
        """Worker thread of GPU i.  With a two-phase worker (launch / finish) the staging copy of the next batch overlaps
        the GPU work of the current one: a batch is finished -- its Future resolved -- only after the next one, if any is
        queued, has been staged and enqueued."""
        import queue
        w, q = self.workers[i], self._queues[i]
        two_phase = hasattr(w, "launch") and hasattr(w, "finish")
        pending = None                       # (future, token) of the batch running on the GPU
        while True:
            try:
                job = q.get(block=pending is None)
            except queue.Empty:
                job = False                  # nothing queued right now
            nxt = None
            if job not in (None, False):
                fut, ids, frames = job
                if fut.set_running_or_notify_cancel():
                    try:
                        if two_phase:
                            nxt = (fut, w.launch(ids, frames))
                        else:
                            fut.set_result(w.run(ids, frames))
                    except BaseException as exc:      # hand every failure to the waiter instead of killing the worker
                        fut.set_exception(exc)
            if pending is not None:
                pfut, token = pending
                try:
                    pfut.set_result(w.finish(token))
                except BaseException as exc:
                    pfut.set_exception(exc)
            pending = nxt
            if job is None:
                return

    def assignment(self, n_frames: int):
        """frame ids per worker for a batch of n_frames: b -> worker b mod G."""
        g = len(self.workers)
        if n_frames > g * self.B:
            raise ValueError(f"{n_frames} frames exceed {g} GPUs x {self.B} frames per step")
        return [list(range(i, n_frames, g)) for i in range(g)]

    def submit(self, frames):
        """Deal the frames out; returns one Future per worker (None for a worker that got no frame)."""
        from concurrent.futures import Future
        futs = []
        for i, ids in enumerate(self.assignment(len(frames))):
            if not ids:
                futs.append(None)
                continue
            fut = Future()
            self._queues[i].put((fut, ids, [frames[b] for b in ids]))
            futs.append(fut)
        return futs

    def run(self, frames):
        return [f.result() if f is not None else None for f in self.submit(frames)]

    def close(self):
        for q in self._queues:
            q.put(None)
        for t in self._threads:
            t.join(timeout=10)
        for w in self.workers:
            if hasattr(w, "close"):
                w.close()
        self.workers = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
