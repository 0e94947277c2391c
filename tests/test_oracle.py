"""CPU tests: the oracle (C restatement + data-parallel spec) against the REFERENCE's golden vectors."""
import numpy as np
import pytest

from oracle import oracle, spec_parallel
from tests import util
from tests.util import cfgmod, synth


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("name", util.SMALL_NAMES)
def test_c_oracle_voxelize_matches_reference(gkey, name):
    g = util.load_small(gkey, name)
    cfg = util.small_config(gkey, g)
    geom = cfgmod.derive_grid(cfg)
    assert np.array_equal(geom.grid_size, g["grid_size"])
    assert np.array_equal(geom.offset.view(np.uint32), g["offset"].view(np.uint32))
    assert np.array_equal(geom.detection_range.view(np.uint32), g["detection_range"].view(np.uint32))
    got = oracle.generate(geom, g["points"], cfg["max_voxels"], cfg["max_num_points"])
    util.assert_vox_equal(got, (g["voxels"], g["coors"], g["num"]), f"{gkey}/{name}")


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("name", util.SMALL_NAMES)
def test_parallel_spec_matches_reference(gkey, name):
    g = util.load_small(gkey, name)
    cfg = util.small_config(gkey, g)
    geom = cfgmod.derive_grid(cfg)
    want = (g["voxels"], g["coors"], g["num"])
    util.assert_vox_equal(spec_parallel.voxelize_sorted(geom, g["points"], cfg["max_voxels"],
                                                        cfg["max_num_points"]), want, "sorted spec")
    for tile in (64, 2048):
        util.assert_vox_equal(spec_parallel.voxelize_tiled(geom, g["points"], cfg["max_voxels"],
                                                           cfg["max_num_points"], tile=tile), want,
                              f"tiled spec tile={tile}")


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("kind", ["S1", "S2", "S3", "S4"])
def test_c_oracle_full_size_hashes(gkey, kind):
    cfg, geom, pts, h = util.full_frame(gkey, kind)
    if util.sha(pts) != h["points_sha"]:
        pytest.skip("synthetic frame not bit-identical on this host (libm/SIMD differs); hash pin not applicable")
    v, c, n = oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"])
    assert len(n) == h["V"] and int(n.sum()) == h["kept"]
    assert util.sha(v) == h["voxels_sha"] and util.sha(c) == h["coors_sha"] and util.sha(n) == h["num_sha"]


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("name", [n for n in util.SMALL_NAMES if n != "n0"])
def test_oracle_decorate_matches_reference(gkey, name):
    g = util.load_small(gkey, name)
    cfg = util.small_config(gkey, g)
    geom = cfgmod.derive_grid(cfg)
    consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)
    ref = g["features"]
    tol = oracle.decorate_tolerance(g["voxels"])            # 1e-6 * max(1, |xyz|max) per pillar (H4)
    f64 = oracle.decorate_f64(g["voxels"], g["num"], g["coors"], *consts)
    # the fp64 yardstick also sees the rounding of float(c)*v (+ offset) at grid-extent magnitude (F13)
    tol64 = np.maximum(tol, 1e-6 * float(np.abs(geom.detection_range).max()))
    for impl in (oracle.decorate, oracle.decorate_numpy):
        got = impl(g["voxels"], g["num"], g["coors"], *consts)
        assert got.shape == ref.shape and got.dtype == np.float32
        both_nan = np.isnan(got) & np.isnan(ref)
        assert np.all((np.abs(got.astype(np.float64) - ref) <= tol) | both_nan), impl.__name__
        assert np.all((np.abs(got - f64) <= tol64) | both_nan)
    # the pass-through columns and the padding rows are exact
    got = oracle.decorate(g["voxels"], g["num"], g["coors"], *consts)
    assert np.array_equal(got[..., :4].view(np.uint32), g["voxels"].view(np.uint32))


@pytest.mark.parametrize("gkey,name", [("20cm", "s1"), ("20cm", "s4_edges"), ("nuscene", "s1"),
                                       ("nuscene", "s4_edges")])
def test_oracle_scatter_matches_reference(gkey, name):
    g = util.load_small(gkey, name)
    nx, ny = int(g["grid_size"][0]), int(g["grid_size"][1])
    canvas = oracle.scatter(g["pillar_feats"], g["coors"], nx, ny)
    assert tuple(canvas.shape) == tuple(g["canvas_shape"])
    assert util.sha(canvas) == str(g["canvas_sha"])
    cm = oracle.scatter_cmajor(np.ascontiguousarray(g["pillar_feats"].T), g["coors"], nx, ny)
    assert util.sha(cm) == str(g["canvas_sha"])
    # batched contract: batch id in the last coords column, per-frame result stacked
    co4 = np.concatenate([np.pad(g["coors"], ((0, 0), (0, 1)), constant_values=b) for b in range(2)])
    two = oracle.scatter(np.concatenate([g["pillar_feats"], 2 * g["pillar_feats"]]), co4, nx, ny, batch=2)
    assert util.sha(two[0:1]) == str(g["canvas_sha"]) and np.array_equal(two[1], 2 * two[0])
    # backward = gather
    gb = oracle.scatter_backward(canvas, g["coors"])
    assert np.array_equal(gb, g["pillar_feats"])


BACKWARD_CASES = ["nuscene_s1", "20cm_s4_edges", "20cm_dup"]


@pytest.mark.parametrize("case", BACKWARD_CASES)
def test_oracle_scatter_backward_matches_reference(case):
    """Row A7: the oracle's gather against the gradient the REFERENCE's PointPillarsScatter returned through autograd
    (oracle/make_golden_backward.py)."""
    z, gc, coors = util.load_backward(case)
    got = oracle.scatter_backward(gc, coors)
    assert np.array_equal(got.view(np.uint32), z[case + "_grad_feats"].view(np.uint32))


def test_micro_cases():
    """Hand-checkable cases of SURVEY.md §8c on the 20 cm grid."""
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)

    def cells(rows):
        return oracle.generate(geom, np.asarray(rows, np.float32), 16000, 15)[1].tolist()

    assert cells([[-80, -80, -2.5, 0]]) == [[0, 0, 0]]
    assert cells([[80, 80, 8.5, 0]]) == []
    assert cells([[79.99999, 79.99999, 8.4999, 0]]) == []          # fp32 79.99999+80 rounds to 160.0
    assert cells([[-0.0, 0.0, 0, 0]]) == [[400, 400, 0]]
    assert cells([[np.nan, 0, 0, 0], [0, np.inf, 0, 0], [0, 0, -np.inf, 0], [3e9, 0, 0, 0], [-3e9, 0, 0, 0]]) == []
    # F5: after max_voxels cells exist the first NEW cell stops everything, even points of existing voxels
    v, c, n = oracle.generate(geom, np.asarray([[0, 0, 0, 1], [1, 0, 0, 2], [2, 0, 0, 3], [0, 0, 0, 4]],
                                               np.float32), 2, 15)
    assert c.tolist() == [[400, 400, 0], [405, 400, 0]] and n.tolist() == [1, 1]
    # F6: a cell with more than P points keeps the first P in input order
    pts = np.zeros((40, 4), np.float32)
    pts[:, 3] = np.arange(40)
    v, c, n = oracle.generate(geom, pts, 16000, 15)
    assert n.tolist() == [15] and v[0, :, 3].tolist() == list(range(15))


@pytest.mark.parametrize("name", util.EXTRA_NAMES)
def test_oracle_extra_geometries_match_reference(name):
    """nz > 1 grids and F != 4 point rows, pinned by the reference's own outputs."""
    g, cfg = util.load_extra(name)
    geom = cfgmod.derive_grid(cfg)
    assert np.array_equal(geom.grid_size, g["grid_size"])
    assert np.array_equal(geom.offset.view(np.uint32), g["offset"].view(np.uint32))
    want = (g["voxels"], g["coors"], g["num"])
    util.assert_vox_equal(oracle.generate(geom, g["points"], cfg["max_voxels"], cfg["max_num_points"]), want, name)
    util.assert_vox_equal(spec_parallel.voxelize_sorted(geom, g["points"], cfg["max_voxels"], cfg["max_num_points"]),
                          want, f"{name} sorted spec")
    util.assert_vox_equal(spec_parallel.voxelize_tiled(geom, g["points"], cfg["max_voxels"], cfg["max_num_points"],
                                                       tile=256), want, f"{name} tiled spec")


def test_oracle_vs_parallel_spec_random_geometries():
    """Property test (hypothesis): on random small grids / caps / clouds the sequential C oracle and the order-free
    parallel statement agree bit for bit, and the documented invariants hold."""
    hyp = pytest.importorskip("hypothesis")
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(seed=st.integers(0, 2 ** 31 - 1), nx=st.integers(1, 12), ny=st.integers(1, 12), nz=st.integers(1, 4),
           max_voxels=st.integers(1, 60), max_points=st.integers(1, 6), n=st.integers(0, 300), f=st.integers(3, 6))
    def run(seed, nx, ny, nz, max_voxels, max_points, n, f):
        rng = np.random.default_rng(seed)
        vs = np.array([0.5, 0.25, 1.0])
        lo = np.array([-1.0, 2.0, -3.0])
        cfg = dict(detection_range=list(lo) + list(lo + vs * [nx, ny, nz]), voxel_size=list(vs),
                   max_voxels=max_voxels, max_num_points=max_points, num_point_features=f)
        geom = cfgmod.derive_grid(cfg)
        span = vs * [nx, ny, nz]
        pts = (lo - 0.2 * span + rng.random((n, 3)) * 1.4 * span).astype(np.float32)
        pts = np.ascontiguousarray(np.concatenate([pts, rng.random((n, f - 3), dtype=np.float32)], axis=1))
        if n:   # duplicates and exact cell boundaries
            pts[rng.integers(0, n, n // 4)] = pts[rng.integers(0, n, n // 4)]
            k = rng.integers(0, n, n // 8)
            pts[k, 0] = (lo[0] + vs[0] * rng.integers(0, nx + 1, len(k))).astype(np.float32)
        v, c, m = oracle.generate(geom, pts, max_voxels, max_points)
        util.assert_vox_equal(spec_parallel.voxelize_sorted(geom, pts, max_voxels, max_points), (v, c, m), "sorted")
        util.assert_vox_equal(spec_parallel.voxelize_tiled(geom, pts, max_voxels, max_points, tile=32), (v, c, m), "tiled")
        assert len(m) <= max_voxels and (len(m) == 0 or (m.min() >= 1 and m.max() <= max_points))
        assert len(np.unique(c, axis=0)) == len(c)
        assert np.all((c >= 0) & (c < geom.grid_size))

    run()


@pytest.mark.parametrize("name", util.ANCHOR_NAMES)
def test_anchor_mask_oracle_matches_reference_golden(name):
    """Row N2: the restated get_anchor_coor / create_mask arithmetic against vectors made by the reference's own
    functions (oracle/make_golden_anchor.py)."""
    g = util.load_anchor(name)
    ac = oracle.anchor_coor(g["anchors_bv"], g["voxel_size"], g["offset"], g["grid_size"])
    assert ac.dtype == np.int32 and np.array_equal(ac, g["anchor_coors"])
    mask = oracle.anchors_mask(g["coors"], g["grid_size"][:2], g["anchor_coors"])
    assert mask.dtype == np.bool_ and np.array_equal(mask, g["mask"])
    assert 0 < mask.sum() < mask.size or name == "grid3d"


@pytest.mark.parametrize("name", util.ANCHOR_NAMES)
def test_get_anchor_coor_host_mirror_matches_reference_golden(name):
    """The product's init-time host function (numpy only, no device work) reproduces the reference's rectangles."""
    import importlib
    am = importlib.import_module("3d_object_detection_b200.anchor_mask")
    g = util.load_anchor(name)
    got = am.get_anchor_coor(g["anchors_bv"], g["voxel_size"], g["offset"], g["grid_size"])
    assert got.dtype == np.int32 and np.array_equal(got, g["anchor_coors"])


def test_synthetic_anchor_population():
    """synth.anchors_bv yields the reference's anchor count (9 per feature-map cell) inside the detection range."""
    geom = cfgmod.derive_grid(cfgmod.builtin_config("20cm"))
    bv = util.synth.anchors_bv(geom, (40, 40))
    assert bv.shape == (9 * 1600, 4) and bv.dtype == np.float32
    assert np.all(bv[:, 2] > bv[:, 0]) and np.all(bv[:, 3] > bv[:, 1])
    ac = oracle.anchor_coor(bv, geom.voxel_size[:2].astype(np.float32), geom.offset[:2], geom.grid_size)
    assert ac.min() >= 0 and np.all(ac[:, 2] < geom.grid_size[0]) and np.all(ac[:, 3] < geom.grid_size[1])


NMS_CASES = ["n1", "n7", "n64", "n65", "n200", "n300_t5"]


@pytest.mark.parametrize("name", NMS_CASES)
def test_nms_oracle_matches_reference_golden(name):
    """Row N4: the restated standup-box NMS against the reference's own nms_gpu run under numba's CUDA simulator
    (oracle/make_golden_nms.py)."""
    z = np.load(util.GOLD / "nms.npz")
    dets, thresh, want = z[f"{name}_dets"], z[f"{name}_thresh"], z[f"{name}_keep"]
    got = oracle.nms_gpu(dets, thresh)
    assert got == want.tolist()
    k = oracle.nms(dets[:, :4], dets[:, 4], 1000, 5, thresh)
    assert k.dtype == np.int64 and k.tolist() == want.tolist()[:5]
    assert oracle.nms(dets[:0, :4], dets[:0, 4], 1000, 300, thresh) is None


@pytest.mark.parametrize("name", ["t01", "t05", "t03"])
def test_nms_oracle_float64_rule_matches_compiled_numba(name):
    """The PRECISION of the overlap test: decisive pairs within a few float32 ulps of the threshold, expected keep list
    from the reference's own nms_gpu with its iou_device compiled by real numba (oracle/make_golden_nms_f64.py); an
    all-float32 evaluation gives `keep_if_float32`, which differs in every decisive pair."""
    z = np.load(util.GOLD / "nms_f64.npz")
    dets, thresh = z[f"{name}_dets"], z[f"{name}_thresh"]
    want, f32 = z[f"{name}_keep"].tolist(), z[f"{name}_keep_if_float32"].tolist()
    assert want != f32
    assert oracle.nms_gpu(dets, thresh) == want


def test_nms_ptx_evidence_of_the_reference_kernel():
    """What numba's CUDA target makes of the reference's iou_device / nms_kernel source (committed by the generator):
    float32 edge differences, everything from the `+ 1` on in float64, float64 comparison with the threshold."""
    import json
    ev = json.loads((util.GOLD / "nms_iou_ptx_ops.json").read_text())
    assert ev["numba_return_type"] == "float64"
    assert {"sub.f32", "cvt.f64.f32", "add.f64", "mul.f64", "div.rn.f64"} <= set(ev["float_ptx_ops"])
    assert not any(op.endswith(".f32") and op.split(".")[0] in ("div", "mul", "add") for op in ev["float_ptx_ops"])
    assert "setp.gt.f64" in ev["kernel_float_ptx_ops"]
    assert not any(op.startswith("setp") and op.endswith("f32") for op in ev["kernel_float_ptx_ops"])


# ---- detection post-processing (row N4 around the NMS) ---------------------------------------------------------------

POST_CASES = util.POST_CASES
_post_golden = util.load_post


@pytest.mark.parametrize("name", POST_CASES)
def test_postprocess_oracle_matches_reference(name):
    """oracle.postprocess_frame against the annotations of the reference's own Inference.infer_gpu
    (oracle/make_golden_post.py): same detections in the same order, boxes bit-exact (the restatement runs the same numpy
    operations in the same container), scores to 1 ulp (the reference's sigmoid is torch's, the oracle's numpy's)."""
    g = _post_golden(name)
    b, s, c, a = oracle.postprocess_frame(g["cls_preds"][0], g["box_preds"][0], g["dir_preds"][0], g["anchors_mask"][0],
                                          g["anchors"], g["class_ranges"].tolist(), g["center_limit"].tolist(), 0.05,
                                          int(g["pre_max"]), int(g["post_max"]), 0.1)
    names = [str(g["detect_class"][k]) for k in c]
    assert names == [str(x) for x in g["name"]]
    want = np.concatenate([g["location"], g["dimensions"], g["rotation_y"].reshape(-1, 1)], 1).astype(np.float32)
    assert np.array_equal(b.view(np.uint32), want.view(np.uint32)), "boxes are not bit-identical to the reference's"
    assert np.allclose(s, g["score"], rtol=2e-7, atol=0)
    assert len(b) > 0


def test_postprocess_cases_exercise_the_branches():
    """The committed cases cover what they are named for."""
    g = _post_golden("topk")
    detail = []
    oracle.postprocess_frame(g["cls_preds"][0], g["box_preds"][0], g["dir_preds"][0], g["anchors_mask"][0], g["anchors"],
                             g["class_ranges"].tolist(), g["center_limit"].tolist(), 0.05, int(g["pre_max"]),
                             int(g["post_max"]), 0.1, detail=detail)
    assert all(len(d["sel"]) == int(g["pre_max"]) for d in detail)          # more candidates than pre_max in every class
    assert any(len(d["keep"]) == int(g["post_max"]) for d in detail)        # and the post_max cut is reached
    g = _post_golden("range")
    detail = []
    b = oracle.postprocess_frame(g["cls_preds"][0], g["box_preds"][0], g["dir_preds"][0], g["anchors_mask"][0], g["anchors"],
                                 g["class_ranges"].tolist(), g["center_limit"].tolist(), 0.05, 1000, 300, 0.1, detail=detail)[0]
    assert 0 < len(b) < sum(len(d["keep"]) for d in detail)                 # the range mask drops some kept boxes
    g = _post_golden("empty_class")
    assert "pedestrian" not in [str(x) for x in g["name"]]
