"""CPU tests of the host logic: config/geometry, drop-in side effects, the C-ABI library's exports, sharding."""
import ctypes
import importlib
import re
from pathlib import Path

import numpy as np
import pytest

from tests import util
from tests.util import cfgmod

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol():
    """The built .so loads and exports exactly what include/pillarfront.h declares (no compute without a GPU)."""
    L = importlib.import_module("3d_object_detection_b200._lib")
    header = (ROOT / "include" / "pillarfront.h").read_text()
    declared = sorted(set(re.findall(r"PF_API[^;(]*?\b(pf_\w+)\s*\(", header)))
    assert declared == sorted(L.EXPORTS)
    lib = L.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.pf_version() == int(re.search(r"#define PF_VERSION (\d+)", header).group(1))
    # argument validation works without a device: null descriptor -> PF_ERR_INVALID_ARGUMENT + message
    out = ctypes.c_void_p()
    assert lib.pf_create(None, 1, 1, 0, ctypes.byref(out)) == -1
    assert b"null" in lib.pf_last_error_string()
    assert ctypes.sizeof(L.GridDesc) == 48


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
def test_geometry_and_config_side_effects(pf, gkey):
    g = util.load_small(gkey, "s1")
    cfg = cfgmod.builtin_config(gkey)
    raw_vs = list(cfg["voxel_size"])
    vg = pf.VoxelGenerator(cfg)                        # constructing needs no GPU
    for got, key in [(vg.grid_size, "grid_size"), (vg.offset, "offset"), (vg.voxel_size, "voxel_size"),
                     (vg.detection_range, "detection_range")]:
        assert got.dtype == g[key].dtype and np.array_equal(got, g[key]), key
    # F8: the four keys the reference writes back
    assert cfg["grid_size"] is vg.grid_size and cfg["detection_offset"] is vg.offset
    assert np.array_equal(cfg["detection_range"], vg.detection_range)
    assert np.array_equal(cfg["detection_range_diff"], vg.grid_size.astype(np.float32) * vg.voxel_size)
    assert cfg["voxel_size"] == raw_vs                 # left alone, PointNet reads the Python floats (F13)
    assert vg.max_voxels == cfg["max_voxels"] and vg.max_num_points == cfg["max_num_points"]


def test_nuscene_offset_is_fp32_exact(pf):
    geom = cfgmod.derive_grid(cfgmod.builtin_config("nuscene"))
    assert geom.grid_size.tolist() == [512, 480, 1]
    assert geom.offset[1] == np.float32(-38.399998) and geom.offset.dtype == np.float32


def test_load_config_tolerates_trailing_comma(tmp_path):
    p = tmp_path / "nuscene.json"
    p.write_text('{"max_voxels": 12000, "voxel_size": [0.16, 0.16, 6,],\n "eval_step": 100,\n}\n')
    cfg = cfgmod.load_config(p)
    assert cfg["max_voxels"] == 12000 and cfg["voxel_size"] == [0.16, 0.16, 6]
    for stem, key in cfgmod.CONFIG_GRID.items():
        assert key in cfgmod.GRIDS, stem


def test_no_cpu_fallback(pf):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    vg = pf.VoxelGenerator(cfgmod.builtin_config("20cm"))
    with pytest.raises(pf.PillarFrontError):
        vg.generate(np.zeros((3, 4), np.float32))
    with pytest.raises(pf.PillarFrontError):
        pf.PointPillarsScatter(1, np.array([8, 8, 1]), 4)(torch.zeros(2, 4), torch.zeros(2, 3, dtype=torch.int32))
    with pytest.raises(pf.PillarFrontError):
        pf.decorate(torch.zeros(2, 3, 4), torch.ones(2, dtype=torch.int32), torch.zeros(2, 3, dtype=torch.int32), .2, .2, 0., 0.)


def test_no_cpu_fallback_next_rows(pf):
    """Anchor mask and NMS (rows N2 / N4) also refuse to run without a device."""
    import importlib
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    am = importlib.import_module("3d_object_detection_b200.anchor_mask")
    nm = importlib.import_module("3d_object_detection_b200.nms")
    with pytest.raises(pf.PillarFrontError):
        am.AnchorMask(np.zeros((4, 4), np.float32), voxel_size=np.float32([.2, .2, 11]), offset=np.float32([0, 0, 0]),
                      grid_size=np.array([8, 8, 1]))
    with pytest.raises(pf.PillarFrontError):
        nm.nms_gpu(np.zeros((3, 5), np.float32), 0.1)
    with pytest.raises(pf.PillarFrontError):
        nm.nms(np.zeros((3, 4), np.float32), np.zeros(3, np.float32), 1000, 300, 0.1)


def test_c_abi_argument_validation_needs_no_device():
    """Bad arguments are rejected with a negative pf_status and a message before any CUDA call is made."""
    L = importlib.import_module("3d_object_detection_b200._lib").load()
    one = ctypes.c_void_p(256)      # non-null, aligned, never dereferenced: validation fails first
    cases = [
        (L.pf_anchor_mask(one, 1, None, 5, 0, 1, 8, 8, one, 3, one, one, None), "coors_stride"),
        (L.pf_anchor_mask(one, 3, None, 5, 0, 2, 8, 8, one, 3, one, one, None), "column 3"),
        (L.pf_anchor_mask(one, 4, None, 5, 7, 2, 8, 8, one, 3, one, one, None), "voxel_num"),
        (L.pf_anchor_mask(one, 3, None, 5, 0, 1, 0, 8, one, 3, one, one, None), "grid"),
        (L.pf_anchor_mask(one, 3, None, 5, 0, 1, 8, 8, ctypes.c_void_p(260), 3, one, one, None), "aligned"),
        (L.pf_nms(one, one, None, 1, 5, 5000, 0.1, 10, one, one, one, None), "max_segment_boxes"),
        (L.pf_nms(one, one, None, 2, 5, 100, 0.1, 10, one, one, one, None), "seg_offsets"),
        (L.pf_nms(one, one, None, 1, 500, 100, 0.1, 10, one, one, one, None), "max_segment_boxes"),
        (L.pf_nms(one, one, None, 1, 5, 100, 0.1, 0, one, one, one, None), "post_max"),
        (L.pf_nms(one, one, None, 1, 5, 100, 0.1, 10, None, one, one, None), "scratch"),
        (L.pf_voxelize(None, one, one, 1, 5, 4, 0, 3, one, one, one, one, None), "null handle"),
        (L.pf_pfn(one, one, one, 3, 5, 15, 4, .2, .2, 0., 0., one, one, one, 32, one, None), "n_ch"),
        (L.pf_decorate(one, one, one, 1, 5, 15, 4, .2, .2, 0., 0., one, None), "coors_stride"),
        (L.pf_scatter_backward(one, one, 3, 5, 0, 64, 8, 8, one, 64, 1, None), "bad sizes"),
    ]
    for rc, needle in cases:
        assert rc < 0, needle
    # the message belongs to the last failing call of this thread
    assert L.pf_scatter_backward(one, one, 3, 5, 0, 64, 8, 8, one, 64, 1, None) < 0
    assert b"bad sizes" in L.pf_last_error_string()
    assert L.pf_nms(one, one, None, 1, 5, 5000, 0.1, 10, one, one, one, None) < 0
    assert b"max_segment_boxes" in L.pf_last_error_string()
    assert L.pf_nms_scratch_bytes(0, 100) == 0 and L.pf_nms_scratch_bytes(1, 5000) == 0
    assert L.pf_nms_scratch_bytes(24, 1000) >= 24 * 1000 * (16 + 4 + 16 * 8)
    # empty work is accepted without a device
    assert L.pf_nms(None, None, None, 0, 0, 100, 0.1, 10, None, None, None, None) == 0
    assert L.pf_anchor_mask(None, 3, None, 0, 0, 0, 8, 8, None, 0, None, None, None) == 0


def test_postprocess_abi_validation_and_refusal(pf):
    """pf_postprocess / pf_post_scratch_*: bad descriptors are rejected without a device; the scratch layout is ordered
    and 256-byte aligned; the Python mirror refuses to run without a device or on config['device'] = 'cpu'."""
    import torch
    Lm = importlib.import_module("3d_object_detection_b200._lib")
    L = Lm.load()
    lim = [-80, -80, -10, 80, 80, 10]
    good = Lm.make_post_desc([[0, 100], [100, 150]], 1, 1000, 300, 0.05, 0.1, lim)
    one = ctypes.c_void_p(256)
    bad = [
        (Lm.make_post_desc([[0, 100], [50, 150]], 1, 1000, 300, 0.05, 0.1, lim), 150, "overlap"),
        (Lm.make_post_desc([[0, 100], [100, 151]], 1, 1000, 300, 0.05, 0.1, lim), 150, "class_range"),
        (Lm.make_post_desc([[0, 100]], 1, 4097, 300, 0.05, 0.1, lim), 150, "pre_max"),
        (Lm.make_post_desc([[0, 100]], 1, 1000, 1025, 0.05, 0.1, lim), 150, "post_max"),
        (Lm.make_post_desc([[0, 100]], 0, 1000, 300, 0.05, 0.1, lim), 150, "n_cls_channels"),
    ]
    for d, A, needle in bad:
        assert L.pf_post_scratch_bytes(ctypes.byref(d), 1, A) == 0, needle
        assert needle.encode() in L.pf_last_error_string(), (needle, L.pf_last_error_string())
        assert L.pf_postprocess(ctypes.byref(d), one, one, one, None, one, 1, A, one, one, one, one, None, one, None) < 0
    assert L.pf_postprocess(ctypes.byref(good), one, one, one, None, one, 1, 150, None, one, one, one, None, one, None) < 0
    assert b"scratch" in L.pf_last_error_string()
    assert L.pf_postprocess(ctypes.byref(good), one, one, one, None, one, 1, 150, one, None, one, one, None, one, None) < 0
    assert L.pf_postprocess(ctypes.byref(good), one, one, one, None, one, 0, 150, None, None, None, None, None, None, None) == 0
    offs = (ctypes.c_size_t * len(Lm.POST_REGION_NAMES))()
    total = L.pf_post_scratch_layout(ctypes.byref(good), 8, 150, offs)
    o = list(offs)
    assert o[0] == 0 and o == sorted(o) and all(x % 256 == 0 for x in o) and total > o[-1]
    assert total == L.pf_post_scratch_bytes(ctypes.byref(good), 8, 150)
    assert o[1] - o[0] >= 8 * 150 * 8 and total - o[-1] >= L.pf_nms_scratch_bytes(16, 1000)
    with pytest.raises(ValueError):
        Lm.make_post_desc([[0, 1]] * 9, 1, 1000, 300, 0.05, 0.1, lim)
    inf = importlib.import_module("3d_object_detection_b200.inference")

    class A_:
        anchors = np.zeros((150, 7), np.float32)
        class_masks = {"vehicle": [0, 100], "pedestrian": [100, 150]}
    with pytest.raises(pf.PillarFrontError):
        inf.Inference(dict(device="cpu", center_limit=lim, detect_class=["vehicle", "pedestrian"]), A_())
    if not torch.cuda.is_available():
        with pytest.raises(pf.PillarFrontError):
            inf.PostProcessor(A_.anchors, [[0, 100], [100, 150]], lim)
        i = inf.Inference(dict(device=0, center_limit=lim, detect_class=["vehicle", "pedestrian"]), A_())
        assert (i._nms_pre_max_size, i._nms_post_max_size, i._nms_iou_threshold, i._nms_score_threshold) == (1000, 300, 0.1, 0.05)
        with pytest.raises(pf.PillarFrontError):
            i.infer_gpu({}, dict(cls_preds=torch.zeros(1, 150, 1), box_preds=torch.zeros(1, 150, 7), dir_preds=torch.zeros(1, 150, 2)))
    for dev in (0, "0", "cuda:1", "cuda", torch.device("cuda:0")):     # what configs/*.json and ros_node.py pass
        i = inf.Inference(dict(device=dev, center_limit=lim, detect_class=["vehicle", "pedestrian"]), A_())
        assert i.device.type == "cuda" and i.device.index == (1 if dev == "cuda:1" else 0)
    assert inf.Inference.infer is inf.Inference.infer_gpu and inf.Inference.infer_torch is inf.Inference.infer_gpu
    anno = inf.get_start_result_anno()
    assert set(anno) == {"name", "truncated", "occluded", "alpha", "bbox", "dimensions", "location", "rotation_y", "score"}
    assert anno["bbox"].shape == (0, 4) and anno["location"].shape == (0, 3)


def test_synthetic_anchors_3d_layout():
    """synth.anchors_3d lays the anchors out like AnchorAssigner (all cells of one size / yaw contiguously, x slow) and
    its class ranges tile the anchor array; the full-size golden case pins it to the reference's own array by hash."""
    geom = cfgmod.derive_grid(cfgmod.builtin_config("20cm"))
    synth = importlib.import_module("3d_object_detection_b200.synth")
    a, ranges = synth.anchors_3d(geom, (6, 5))
    assert a.shape == (6 * 5 * 9, 7) and a.dtype == np.float32
    assert ranges == [[0, 180], [180, 210], [210, 270]]
    assert np.all(a[:30, 3:6] == np.float32([4.6, 2.1, 1.8])) and np.all(a[:30, 6] == 0) and np.all(a[30:60, 6] == np.float32(np.pi / 2))
    assert a[0, 0] == a[4, 0] and a[0, 1] != a[1, 1] and a[5, 0] != a[0, 0]          # y fast, x slow
    assert np.allclose(a[:, 2], a[:, 5] / 2)


def test_product_never_imports_oracle():
    for path in (ROOT / "3d_object_detection_b200").rglob("*.py"):
        text = path.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), path
        # nor the host emulation of the post-processing kernels (tests/emul/): that is a checker, never a fallback
        assert "emul" not in text and "PF_HOST_EMUL" not in text, path
    build_py = (ROOT / "3d_object_detection_b200" / "build.py").read_text()
    assert "PF_HOST_EMUL" not in build_py
    for path in (ROOT / "3d_object_detection_b200" / "csrc").glob("*.cu"):
        assert "PF_HOST_EMUL" not in path.read_text(), path      # only the shared kernel header knows the switch


def test_shard_frames(pf):
    for n in (0, 1, 7, 8, 64):
        for g in (1, 2, 4, 8):
            for mode in ("block", "cyclic"):
                parts = [pf.shard_frames(n, r, g, mode) for r in range(g)]
                assert sorted(sum(parts, [])) == list(range(n))
                assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        pf.shard_frames(4, 2, 2)


def test_place_devices(pf):
    """Bridge-aware placement (moved out of bench.py): fewer workers than GPUs alternate between the two halves."""
    assert pf.place_devices(4, 8) == [0, 4, 1, 5] and pf.place_devices(2, 8) == [0, 4]
    assert pf.place_devices(8, 8) == list(range(8)) and pf.place_devices(2, 4) == [0, 2]
    assert pf.place_devices(1, 8) == [0] and pf.place_devices(2, 3) == [0, 1]
    assert sorted(pf.place_devices(6, 8)) == [0, 1, 2, 4, 5, 6]
    assert pf.place_devices(4, 8, "linear") == [0, 1, 2, 3]
    with pytest.raises(ValueError):
        pf.place_devices(9, 8)


def test_sharded_front_end_scheduling(pf):
    """ShardedFrontEnd without a GPU: frame b goes to worker b mod G, one worker thread per device runs its share,
    results come back per device with the frame ids they cover, worker failures surface in the Future."""
    import threading
    import time

    class FakeWorker:
        def __init__(self, device):
            self.device, self.calls, self.threads = device, [], set()

        def run(self, ids, frames):
            self.threads.add(threading.get_ident())
            time.sleep(0.05)
            if any(f is None for f in frames):
                raise RuntimeError("bad frame")
            self.calls.append(list(ids))
            return (self.device, list(ids), [int(f.sum()) for f in frames])

    frames = [np.full((3, 4), b, np.float32) for b in range(11)]
    sfe = pf.ShardedFrontEnd({}, devices=[0, 4, 1], frames_per_device=4, _worker_factory=FakeWorker)
    assert sfe.assignment(11) == [[0, 3, 6, 9], [1, 4, 7, 10], [2, 5, 8]]
    assert sfe.assignment(2) == [[0], [1], []]
    with pytest.raises(ValueError):
        sfe.assignment(13)
    t0 = time.perf_counter()
    res = sfe.run(frames)
    assert time.perf_counter() - t0 < 0.14                      # three 50 ms jobs ran side by side, not in sequence
    assert [r[0] for r in res] == [0, 4, 1]
    for r, ids in zip(res, sfe.assignment(11)):
        assert r[1] == ids and r[2] == [12 * b for b in ids]
    futs = sfe.submit(frames[:2])
    assert futs[2] is None and futs[0].result()[1] == [0] and futs[1].result()[1] == [1]
    main = threading.get_ident()
    assert all(len(w.threads) == 1 and main not in w.threads for w in sfe.workers)
    bad = sfe.submit([frames[0], None, frames[2]])
    assert bad[0].result()[1] == [0]
    with pytest.raises(RuntimeError, match="bad frame"):
        bad[1].result()
    assert sfe.run(frames[:3])[1][1] == [1]                     # the worker survived the failure
    sfe.close()


_GLOO_WORKER = r"""
import importlib, json, os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
sh = importlib.import_module("3d_object_detection_b200.sharding")
dist.init_process_group("gloo")
r, g = dist.get_rank(), dist.get_world_size()
mine = sh.shard_frames(13, r, g, "block")
t = sh.reduce_max_time(0.25 * (r + 1))
tot = sh.reduce_sum([len(mine), 1000 * len(mine)])
dist.barrier()
if r == 0:
    print(json.dumps({"t": t, "tot": tot, "world": g}))
dist.destroy_process_group()
"""


def test_sharded_reductions_gloo_world2(tmp_path):
    """N > 1 path of bench.py on CPU: two gloo ranks shard 13 frames, MAX the time, SUM the work."""
    import json
    import subprocess
    import sys
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", str(script), str(ROOT)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0, res.stderr[-2000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == 2 and out["t"] == pytest.approx(0.5) and out["tot"] == [13.0, 13000.0]


def test_bench_reference_arm_runs_on_cpu():
    """bench.py --impl reference needs no GPU: one JSON line with the contract's keys."""
    import json
    import subprocess
    import sys
    res = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "points/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0


def test_collate_matches_reference_layout():
    """Row N3: merge_second_batch on numpy examples is the reference's layout (frame id padded as the LAST coordinate
    column only when the batch has more than one example, framework/utils.py:34-46); torch CPU tensors take the same
    path the CUDA tensors take."""
    import importlib
    torch = pytest.importorskip("torch")
    col = importlib.import_module("3d_object_detection_b200.collate")
    rng = np.random.default_rng(3)
    ex = []
    for v in (5, 0, 7):
        ex.append({"voxels": rng.random((v, 4, 4), dtype=np.float32),
                   "num_points_per_voxel": rng.integers(1, 5, v).astype(np.int32),
                   "coordinates": rng.integers(0, 50, (v, 3)).astype(np.int32),
                   "anchors_mask": rng.random(11) > 0.5, "match_indices_num": np.array([v], np.int32),
                   "image_idx": np.int64(v)})
    m = col.merge_second_batch(ex)
    assert m["voxels"].shape == (12, 4, 4) and m["num_points_per_voxel"].shape == (12,)
    assert m["coordinates"].shape == (12, 4) and m["coordinates"].dtype == np.int32
    assert np.array_equal(m["coordinates"][:, 3], [0] * 5 + [2] * 7)
    assert np.array_equal(m["coordinates"][:5, :3], ex[0]["coordinates"])
    assert m["anchors_mask"].shape == (3, 11) and m["match_indices_num"].tolist() == [5, 0, 7]
    assert m["image_idx"].tolist() == [5, 0, 7]
    one = col.merge_second_batch(ex[:1])
    assert one["coordinates"].shape == (5, 3)                       # no pad for a batch of one
    # tensors: same values, tensor outputs
    ext = [{k: (torch.from_numpy(v) if isinstance(v, np.ndarray) and k != "image_idx" else v) for k, v in e.items()}
           for e in ex]
    mt = col.merge_second_batch(ext)
    for k in ("voxels", "num_points_per_voxel", "coordinates", "anchors_mask", "match_indices_num"):
        assert torch.is_tensor(mt[k]) and np.array_equal(mt[k].numpy(), m[k]), k
    conv = col.example_convert_to_torch(m, device="cpu")
    assert conv["voxels"].dtype == torch.float32 and conv["coordinates"].dtype == torch.int32
    assert conv["anchors_mask"].dtype == torch.bool and conv["image_idx"] is m["image_idx"]


def test_collate_matches_reference_golden():
    """The same function against the output of the reference's own merge_second_batch (oracle/make_golden_collate.py)."""
    import importlib
    col = importlib.import_module("3d_object_detection_b200.collate")
    z = np.load(Path(__file__).resolve().parent / "golden" / "collate.npz")
    keys = ("voxels", "num_points_per_voxel", "coordinates", "anchors_mask", "match_indices_num")
    ex = [{k: z[f"in{i}_{k}"] for k in keys} for i in range(3)]
    for tag, batch in (("merged", ex), ("single", ex[:1])):
        got = col.merge_second_batch(batch)
        for k in keys:
            want = z[f"{tag}_{k}"]
            assert got[k].dtype == want.dtype and got[k].shape == want.shape and np.array_equal(got[k], want), (tag, k)


def test_bench_rank_placement():
    """bench.spread_device: fewer ranks than GPUs alternate between the two halves of the box; a full box, an odd
    box, or PF_DEVICE_MAP=linear keep LOCAL_RANK -> GPU."""
    import os
    bench = importlib.import_module("bench")
    old = os.environ.pop("PF_DEVICE_MAP", None)
    try:
        assert [bench.spread_device(r, 4, 8) for r in range(4)] == [0, 4, 1, 5]
        assert [bench.spread_device(r, 2, 8) for r in range(2)] == [0, 4]
        assert [bench.spread_device(r, 8, 8) for r in range(8)] == list(range(8))
        assert [bench.spread_device(r, 2, 4) for r in range(2)] == [0, 2]
        assert bench.spread_device(0, 1, 8) == 0 and bench.spread_device(1, 2, 3) == 1
        assert sorted(bench.spread_device(r, 6, 8) for r in range(6)) == [0, 1, 2, 4, 5, 6]     # a permutation, no clash
        os.environ["PF_DEVICE_MAP"] = "linear"
        assert [bench.spread_device(r, 4, 8) for r in range(4)] == [0, 1, 2, 3]
    finally:
        os.environ.pop("PF_DEVICE_MAP", None)
        if old is not None:
            os.environ["PF_DEVICE_MAP"] = old


def test_stage_copy_streaming_stores_copy_exactly():
    """pf_stage_copy (non-temporal stores into the pinned staging buffer): every size / misalignment / thread count copies
    exactly the requested bytes and nothing around them.  Pure host code, no GPU."""
    L = importlib.import_module("3d_object_detection_b200._lib").load()
    rng = np.random.default_rng(0)
    for n in (0, 1, 15, 64, 4095, 4096, 4097, 100003, 600000, 1920001):
        for off_s, off_d in ((0, 0), (3, 0), (0, 5), (7, 9)):
            src = rng.integers(0, 255, n + 16, dtype=np.uint8)
            dst = np.zeros(n + 32, np.uint8)
            for thr in (1, 4, 16):
                dst[:] = 0
                assert L.pf_stage_copy(dst.ctypes.data + off_d, src.ctypes.data + off_s, n, thr) == 0
                assert np.array_equal(dst[off_d:off_d + n], src[off_s:off_s + n]), (n, off_s, off_d, thr)
                assert not dst[:off_d].any() and not dst[off_d + n:].any()
