"""Shared helpers for the test-suite: golden fixtures, hashing, synthetic frames."""
import hashlib
import importlib
import json
from pathlib import Path

import numpy as np

GOLD = Path(__file__).resolve().parent / "golden"
cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")

GRID_KEYS = sorted(cfgmod.GRIDS)
ANCHOR_NAMES = ["pillar_small", "pillar_sparse", "grid3d"]
SMALL_NAMES = ["s1", "s1_overflow", "s3_overflow", "s2_ring", "s4_edges", "dense_cap", "n0", "n5"]


def sha(a):
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.view(np.uint8).reshape(-1).tobytes() if a.size else b"").hexdigest()


def load_small(gkey, name):
    z = np.load(GOLD / f"small_{gkey}_{name}.npz")
    return {k: z[k] for k in z.files}


def small_config(gkey, g):
    cfg = cfgmod.builtin_config(gkey)
    cfg["max_voxels"] = int(g["max_voxels"])
    cfg["max_num_points"] = int(g["max_points"])
    return cfg


def full_hashes():
    return json.loads((GOLD / "full_hashes.json").read_text())


def full_frame(gkey, kind):
    """Regenerate the seeded full-size frame a full_hashes.json entry was made from."""
    h = full_hashes()[f"{gkey}/{kind}"]
    cfg = cfgmod.builtin_config(gkey)
    geom = cfgmod.derive_grid(cfg)
    pts = synth.make_frame(kind, geom, cfg["frame_points"], h["seed"], synth.S1_CELLS[gkey])
    return cfg, geom, pts, h


def assert_vox_equal(got, want, what=""):
    gv, gc, gn = got
    wv, wc, wn = want
    assert gn.shape == wn.shape, f"{what}: voxel count {gn.shape} != {wn.shape}"
    assert np.array_equal(gn, wn), f"{what}: num_points_per_voxel differs"
    assert np.array_equal(gc, wc), f"{what}: coors differ"
    # bit-exact incl. NaN payloads: compare raw words
    assert np.array_equal(np.ascontiguousarray(gv).view(np.uint32), np.ascontiguousarray(wv).view(np.uint32)), \
        f"{what}: voxels differ"


def assert_features_close(got, want, tol, what=""):
    """|got - want| <= tol elementwise; NaN payloads (e.g. a NaN intensity) must be NaN on both sides."""
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    gn, wn = np.isnan(got), np.isnan(want)
    assert np.array_equal(gn, wn), f"{what}: NaN pattern differs"
    with np.errstate(invalid="ignore"):
        ok = (np.abs(got.astype(np.float64) - want) <= tol) | gn
    assert np.all(ok), f"{what}: {np.count_nonzero(~ok)} elements out of tolerance"


EXTRA_NAMES = ["3d_overflow", "3d_fits", "3d_f6", "2d_f3"]


def load_extra(name):
    """Reference outputs for geometries no shipped config uses (nz > 1, F != 4): oracle/make_golden_extra.py."""
    z = np.load(GOLD / f"extra_{name}.npz")
    g = {k: z[k] for k in z.files}
    cfg = dict(detection_range=[float(x) for x in g["detection_range"]], voxel_size=[float(x) for x in g["voxel_size_cfg"]],
               max_voxels=int(g["max_voxels"]), max_num_points=int(g["max_points"]),
               num_point_features=int(g["points"].shape[1]), batch_size=1, device=0)
    return g, cfg


def load_anchor(name):
    z = np.load(GOLD / f"anchor_{name}.npz")
    return {k: z[k] for k in z.files}


def post_standup_tol(boxes):
    """Tolerance of a standup-box coordinate: the corners are centre + rotated half sizes, each a few ulp of exp / sin /
    cos away from numpy's -> 2e-6 relative to the largest operand."""
    b = np.abs(np.asarray(boxes, np.float64))
    return 2e-6 * np.maximum(1.0, b[:, [0, 1, 3, 4]].max(axis=1))


def assert_post_boxes_close(got, want, what="", final=False):
    """Decoded boxes (x, y, z, l, w, h, r): x, y and r are exact fp32 restatements of the numpy arithmetic (bit-exact);
    l, w, h go through exp (2e-6 relative) and z subtracts h / 2 (2e-6 of the larger operand)."""
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    for col in (0, 1, 6):
        assert np.array_equal(got[:, col].view(np.uint32), np.ascontiguousarray(want[:, col]).view(np.uint32)), \
            f"{what}: column {col} is not bit-exact"
    w64 = want.astype(np.float64)
    assert np.all(np.abs(got[:, 3:6] - w64[:, 3:6]) <= 2e-6 * np.abs(w64[:, 3:6])), f"{what}: dimensions out of tolerance"
    ztol = 2e-6 * np.maximum(1.0, np.maximum(np.abs(w64[:, 2]), np.abs(w64[:, 5])))
    assert np.all(np.abs(got[:, 2] - w64[:, 2]) <= ztol), f"{what}: z out of tolerance"


POST_CASES = ["small", "topk", "range", "empty_class", "full20"]


def load_post(name):
    """Reference-made vectors of the detection post-processing (oracle/make_golden_post.py).  The full-size case
    commits hashes instead of 100 MB of inputs: they are regenerated from the seed and must hash the same."""
    import pytest
    z = np.load(GOLD / f"post_{name}.npz")
    g = {k: z[k] for k in z.files}
    if "anchors" not in g:
        cfg = cfgmod.builtin_config("20cm")
        anchors, ranges = synth.anchors_3d(cfgmod.derive_grid(cfg), tuple(int(x) for x in g["feature_map"]))
        cls, box, dirp, mask = synth.post_preds(1, len(anchors), int(g["seed"]), mean_logit=float(g["mean_logit"]))
        assert np.array_equal(np.asarray(ranges), g["class_ranges"])
        for key, arr in (("sha_anchors", anchors), ("sha_cls", cls), ("sha_box", box), ("sha_dir", dirp), ("sha_mask", mask)):
            if sha(arr) != str(g[key]):
                pytest.xfail(f"synthetic {key[4:]} not bit-identical on this host (numpy stream / libm differs): the "
                             "reference-made full-size vector cannot be applied here -- reported, never silently skipped")
        g.update(anchors=anchors, cls_preds=cls, box_preds=box, dir_preds=dirp, anchors_mask=mask)
    return g


def load_backward(case):
    """Reference-made gradient of the BEV scatter (row A7): returns (npz, grad_canvas regenerated from the seed, coors)."""
    import pytest
    z = np.load(GOLD / "scatter_backward.npz")
    grid = z[case + "_grid"]
    gc = np.random.default_rng(int(z[case + "_seed"])).standard_normal((1, 64, int(grid[0]), int(grid[1])), dtype=np.float32)
    if sha(gc) != str(z[case + "_grad_canvas_sha"]):
        pytest.xfail("seeded grad_canvas not bit-identical on this host (numpy stream differs): the reference-made "
                     "gradient cannot be applied here -- reported, never silently skipped")
    return z, gc, z[case + "_coors"]
