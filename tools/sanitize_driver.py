"""Small end-to-end driver for compute-sanitizer (memcheck / racecheck / initcheck): every kernel family of the library at
sizes that finish in seconds under the tool, each result checked against the oracle so a silent corruption also fails.

    compute-sanitizer --tool memcheck python tools/sanitize_driver.py
"""
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pf = importlib.import_module("3d_object_detection_b200")
synth = importlib.import_module("3d_object_detection_b200.synth")
_lib = importlib.import_module("3d_object_detection_b200._lib")
from oracle import oracle  # noqa: E402  (checker only)

L = _lib.load()
cfg = pf.builtin_config("20cm")
cfg["max_voxels"] = 900
geom = pf.derive_grid(cfg)
P = cfg["max_num_points"]

# 1. voxelise: several k_assign / k_tile_rank tiles, an overflowing and a short frame, eager single-frame path twice
vg = pf.VoxelGenerator(dict(cfg))
for seed, kind, n in ((1, "S1", 9000), (2, "S3", 7000), (3, "S2", 4096), (4, "S1", 3)):
    pts = synth.make_frame(kind, geom, max(n, 8), seed, 700)[:n]
    pts = np.ascontiguousarray(pts, np.float32)
    v, c, k = vg.generate(pts)
    wv, wc, wk = oracle.generate(geom, pts, cfg["max_voxels"], P)
    assert np.array_equal(k, wk) and np.array_equal(c, wc) and np.array_equal(v.view(np.uint32), wv.view(np.uint32)), kind
    dv, dc, dk = vg.generate(torch.from_numpy(pts).cuda())
    assert np.array_equal(dk.cpu().numpy(), wk)
vg.close()
print("voxelise ok")

# 2. batched pipeline: voxelise + decorate + fill (every fill variant), compact and slab, then the fused PFN
frames = [np.ascontiguousarray(synth.make_frame(k, geom, n, 10 + i, 500), np.float32) for i, (k, n) in enumerate([("S1", 6000), ("S3", 5000), ("S4", 2500)])]
sizes = [len(f) for f in frames]
pts = torch.from_numpy(np.concatenate(frames)).cuda()
offs = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)).cuda()
consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)
for compact in (False, True):
    fe = pf.PillarFrontEnd(cfg, max_batch=3, max_points_per_frame=max(sizes), device=0, compact=compact)
    pillar = torch.randn((3 * cfg["max_voxels"], 64), device="cuda")
    for variant in (0, 1, 2, 4, 6):
        assert L.pf_debug_set_fill_variant(variant) == variant
        canvas = fe.step(pts, offs, pillar_feats=pillar)
        counts = fe.voxel_num.cpu().numpy()
        starts = np.concatenate([[0], np.cumsum(counts)]) if compact else np.arange(4) * cfg["max_voxels"]
        for f, fr in enumerate(frames):
            v, c, k = oracle.generate(geom, fr, cfg["max_voxels"], P)
            sl = slice(int(starts[f]), int(starts[f]) + len(k))
            assert np.array_equal(fe.coors[sl].cpu().numpy()[:, :3], c)
            feats = fe.features[sl].cpu().numpy()
            want = oracle.decorate(v, k, c, *consts)
            ok = np.isnan(want) | (np.abs(feats - want) <= oracle.decorate_tolerance(v))
            assert ok.all()
            assert np.array_equal(canvas[f].cpu().numpy(), oracle.scatter(pillar[sl].cpu().numpy(), c, fe.nx, fe.ny)[0]), (variant, f)
    L.pf_debug_set_fill_variant(-1)
    net = pf.PointNet(4, cfg["voxel_size"], geom.offset).cuda().eval()
    with torch.no_grad():
        fe.step(pts, offs, pfn=net)
    fe.close()
print("pipeline ok")

# 3. stand-alone scatter + backward, decorate
g = np.load(ROOT / "tests" / "golden" / "small_nuscene_s1.npz")
sc = pf.PointPillarsScatter(1, g["grid_size"], 64)
ft = torch.from_numpy(g["pillar_feats"]).cuda().requires_grad_(True)
out = sc(ft, torch.from_numpy(g["coors"]).cuda())
out.sum().backward()
assert torch.equal(ft.grad, torch.ones_like(ft))
print("scatter ok")

# 4. anchor mask, NMS, post-processing (small)
amod = importlib.import_module("3d_object_detection_b200.anchor_mask")
bv = synth.anchors_bv(geom, (50, 50))
am = amod.AnchorMask(bv, voxel_size=np.asarray(cfg["voxel_size"], np.float32), offset=geom.offset, grid_size=geom.grid_size)
fe = pf.PillarFrontEnd(cfg, max_batch=3, max_points_per_frame=max(sizes), device=0)
fe.voxelize_decorate(pts, offs, keep_cell_map=False)
mask = fe.anchor_mask(am).cpu().numpy()
counts = fe.voxel_num.cpu().numpy()
for f in range(3):
    c = fe.frame_outputs(f, counts)[1].cpu().numpy()[:, :3]
    assert np.array_equal(mask[f], oracle.anchors_mask(c, geom.grid_size[:2], am.anchors_coors))
fe.close()
nmod = importlib.import_module("3d_object_detection_b200.nms")
dets = synth.nms_dets(200, 5)
assert nmod.nms_gpu(dets, 0.1) == oracle.nms_gpu(dets, np.float32(0.1))
imod = importlib.import_module("3d_object_detection_b200.inference")
anchors, ranges = synth.anchors_3d(geom, (20, 20))
cls, box, dirp, amask = synth.post_preds(2, len(anchors), 21, mean_logit=-4.0)
post = imod.PostProcessor(anchors, ranges, [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0], max_frames=2, pre_max_size=100, post_max_size=30)
res = post(*(torch.from_numpy(x).cuda() for x in (cls, box, dirp, amask)))
torch.cuda.synchronize()
assert int(res[4].sum().item()) >= 0
print("anchor mask / nms / post ok")
print("SANITIZE DRIVER DONE")
