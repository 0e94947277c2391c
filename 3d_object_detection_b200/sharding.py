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
    `counts` is `voxel_num` on the host; `device_ms` the CUDA-event time of the batch (upload + graph; -1 if the slot had
    already been reused when the batch was waited for)."""

    __slots__ = ("device", "frame_ids", "counts", "_fe", "_n", "_ev")

    def __init__(self, device, frame_ids, fe, n, counts, ev):
        self.device, self.frame_ids, self.counts = device, frame_ids, counts
        self._fe, self._n, self._ev = fe, n, ev

    @property
    def canvas(self):
        return self._fe.canvas[:self._n]

    @property
    def coors(self):
        return self._fe.coors[:self._n * self._fe.max_voxels]

    @property
    def voxel_num(self):
        return self._fe.voxel_num[:self._n]

    @property
    def device_ms(self):
        return self._ev


class _GpuWorker:
    """One GPU: `depth` PillarFrontEnd replicas (workspace, outputs, pinned staging, one stream and one CUDA graph
    each), used in turn by the GPU's native runner (pf_runner_*); the graph holds H2D of the staged frames + offsets, voxelise + decorate (+ fused PFN), canvas
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
        import ctypes as C
        # the native runner: one host thread of the library per GPU uploads, launches and is waited on -- no interpreter
        # lock on the data path.  Slot k of the runner = self.slots[k]: its instantiated graph, stream and buffers.
        L = _lib.load()
        self._runner = C.c_void_p()
        _lib.check(L.pf_runner_create(self.device, depth, self.B, 4 * self.F, self.threads, C.byref(self._runner)), "pf_runner_create")
        for k, s in enumerate(self.slots):
            _lib.check(L.pf_runner_set_slot(self._runner, k, s["graph"].raw_cuda_graph_exec(), s["stream"].cuda_stream,
                                            s["d_pts"].data_ptr(), s["h_pts"].data_ptr(), s["h_offs"].data_ptr()), "pf_runner_set_slot")

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
          (it must be complete: the copy runs on the runner's stream, not on the one that produced the tensor).
        Only pointers are handed to the GPU's native runner (pf_runner_submit); its thread does the copies and the launch."""
        import ctypes as C
        torch = self.torch
        n = len(frames)
        if n > self.B:
            raise ValueError(f"{n} frames for a worker built for {self.B}")
        ptrs, rows, kinds, keep = (C.c_void_p * self.B)(), (C.c_int64 * self.B)(), (C.c_int32 * self.B)(), []
        total = 0
        f32 = torch.float32
        for b, f in enumerate(frames):
            if isinstance(f, torch.Tensor):
                if f.dtype != f32:
                    raise TypeError(f"frames must be float32, got {f.dtype}")
                if not f.is_contiguous():
                    f = f.contiguous()
                if f.is_cuda and f.device.index != self.device:
                    f = f.to(self.dev)
                ptrs[b], kinds[b] = f.data_ptr(), 2 if f.is_cuda else -1      # host tensor: pinned or not, the library asks the driver
            else:
                if str(getattr(f, "dtype", None)) != "float32":
                    raise TypeError(f"frames must be float32, got {getattr(f, 'dtype', type(f).__name__)}")
                if not f.flags.c_contiguous:
                    f = f.copy()
                ptrs[b], kinds[b] = f.ctypes.data, 0
            shp = f.shape
            if len(shp) != 2 or int(shp[1]) != self.F:
                raise ValueError(f"frame of shape {tuple(shp)} for {self.F} floats per point")
            keep.append(f)                                   # the (possibly re-laid-out) sources must outlive the copies
            rows[b] = int(shp[0])
            total += int(shp[0])
        if total > self.B * self.cap:
            raise ValueError(f"{total} points for a worker built for {self.B} x {self.cap}")
        if any(int(r) > self.cap for r in rows):
            raise ValueError(f"a frame exceeds the capacity of {self.cap} points")
        L = self._lib.load()
        ticket = L.pf_runner_submit(self._runner, ptrs, rows, kinds, n)
        if ticket < 0:
            raise self._lib.PillarFrontError(f"pf_runner_submit failed (status {ticket}): {L.pf_runner_last_error().decode()}")
        return ticket, keep, list(frame_ids), n

    def finish(self, token):
        import ctypes as C
        ticket, _keep, frame_ids, n = token
        L = self._lib.load()
        ms = C.c_float(0.0)
        rc = L.pf_runner_wait(self._runner, ticket, C.byref(ms))      # blocks in the library: the interpreter lock is free
        if rc != 0:
            raise self._lib.PillarFrontError(f"batch {ticket} on GPU {self.device} failed (status {rc}): {L.pf_runner_last_error().decode()}")
        s = self.slots[ticket % len(self.slots)]
        return ShardResult(self.device, frame_ids, s["fe"], n, s["np_cnt"][:n].copy(), float(ms.value))

    def run(self, frame_ids, frames):
        return self.finish(self.launch(frame_ids, frames))

    def close(self):
        if getattr(self, "_runner", None):
            self._lib.load().pf_runner_destroy(self._runner)      # drains the queue, joins the thread
            self._runner = None
        for s in self.slots:
            s["graph"] = None
            s["fe"].close()
        self.slots = []


class _NativeFuture:
    """Future-like handle of one batch on one GPU's native runner: result() waits inside the library."""

    __slots__ = ("_worker", "_token", "_result", "_exc")

    def __init__(self, worker, token):
        self._worker, self._token, self._result, self._exc = worker, token, None, None

    def result(self, timeout=None):
        if self._token is not None:
            try:
                self._result = self._worker.finish(self._token)
            except BaseException as exc:
                self._exc = exc
            self._token = None               # releases the frames kept alive for the upload
        if self._exc is not None:
            raise self._exc
        return self._result

    def done(self):
        return self._token is None


class ShardedFrontEnd:
    """Frame-sharded front end over the GPUs of one box: ``submit(frames)`` deals frame ``b`` to GPU ``b mod G``, every GPU
    runs its own replica of the whole path on its share (pinned staging, one CUDA graph per step), and nothing crosses
    between GPUs: no NCCL, the canvases stay where that GPU's backbone replica would read them.  Each GPU is
    driven by ONE NATIVE THREAD of the library (``pf_runner_*``, csrc/pf_runner.cu): ``submit`` hands it pointers and
    returns, the thread uploads the frames and launches the graph, ``result()`` waits inside the library -- the interpreter
    lock is not on the data path.

        sfe = ShardedFrontEnd(config, devices=8, frames_per_device=8, pfn=point_net)
        futures = sfe.submit(list_of_64_frames)              # one future per GPU -> ShardResult
        results = sfe.run(list_of_64_frames)                 # the same, waited for, in GPU order

    `pfn`: a PointNet (eval mode) replicated onto every GPU and fused into the voxeliser; without it the canvas is
    filled from each worker's `pillar` stand-in tensor (what bench.py's headline does: the PFN is not on the path).
    A result's tensors are views of one of the worker's `depth` buffer sets: valid until `depth` further batches have been
    submitted to that GPU."""

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
            if stage_threads is None:      # host threads per GPU for the pageable -> pinned copy: the runners already work
                import os                 # side by side; with 8 GPUs the copies are bound by host memory (~215 GB/s of traffic
                cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
                # on the pool's box): 1 / 2 threads per GPU = 3.97 / 4.48 G points/s (profiles/r2_sharded_n8_nt1_t2_v3.json)
                stage_threads = max(1, min(4, cores // (2 * len(devices))))
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
        # workers without a native runner (the scheduling tests' stand-ins) get one Python thread each
        self._native = all(hasattr(w, "launch") and hasattr(w, "finish") for w in self.workers)
        self._queues, self._threads = [], []
        if not self._native:
            self._queues = [queue.Queue() for _ in self.workers]
            self._threads = [threading.Thread(target=self._loop, args=(i,), daemon=True) for i in range(len(self.workers))]
            for t in self._threads:
                t.start()

    def _loop(self, i):
        w, q = self.workers[i], self._queues[i]
        while True:
            job = q.get()
            if job is None:
                return
            fut, ids, frames = job
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                fut.set_result(w.run(ids, frames))
            except BaseException as exc:      # hand every failure to the waiter instead of killing the worker
                fut.set_exception(exc)

    def assignment(self, n_frames: int):
        """frame ids per worker for a batch of n_frames: b -> worker b mod G."""
        g = len(self.workers)
        if n_frames > g * self.B:
            raise ValueError(f"{n_frames} frames exceed {g} GPUs x {self.B} frames per step")
        return [list(range(i, n_frames, g)) for i in range(g)]

    def submit(self, frames):
        """Deal the frames out; returns one future per worker (None for a worker that got no frame).  A frame that
        cannot be accepted (dtype, shape, capacity) surfaces when that worker's future is asked for its result."""
        from concurrent.futures import Future
        futs = []
        for i, ids in enumerate(self.assignment(len(frames))):
            if not ids:
                futs.append(None)
                continue
            mine = [frames[b] for b in ids]
            if self._native:
                fut = _NativeFuture(self.workers[i], None)
                try:
                    fut._token = self.workers[i].launch(ids, mine)
                except BaseException as exc:
                    fut._exc = exc
            else:
                fut = Future()
                self._queues[i].put((fut, ids, mine))
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
