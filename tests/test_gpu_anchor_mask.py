"""GPU parity tests for the anchor occupancy mask (row N2, pf_anchor_mask): bit-exact against the reference-made
golden vectors and the oracle."""
import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import cfgmod, synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _am(pf):
    import importlib
    return importlib.import_module("3d_object_detection_b200.anchor_mask")


@pytest.mark.parametrize("name", util.ANCHOR_NAMES)
def test_mask_matches_reference_golden(pf, name):
    g = util.load_anchor(name)
    am = _am(pf)
    a = am.AnchorMask(g["anchors_bv"], voxel_size=g["voxel_size"], offset=g["offset"], grid_size=g["grid_size"])
    assert np.array_equal(a.anchors_coors, g["anchor_coors"])
    for rep in range(2):
        got = a.create_mask(g["coors"], g["grid_size"], g["voxel_size"], g["offset"])          # reference call shape
        assert got.dtype == np.bool_ and got.shape == g["mask"].shape
        assert np.array_equal(got, g["mask"]), f"{name} rep {rep}: {np.flatnonzero(got != g['mask'])[:8]}"
    # the free function with the reference's argument list, torch coordinates
    got = am.fused_get_anchors_mask_gpu(torch.from_numpy(g["coors"]).cuda(), tuple(g["grid_size"][:-1]),
                                        a.anchors_coors_cuda, a.anchors_mask_cuda)
    assert np.array_equal(got, g["mask"])
    with pytest.raises(pf.PillarFrontError):
        a.create_mask(g["coors"], g["grid_size"], g["voxel_size"], g["offset"], gpu=False)


def test_integral_table_is_the_double_running_sum(pf):
    g = util.load_anchor("grid3d")                      # several voxels per (x, y) column: counts > 1
    am = _am(pf)
    ac = torch.from_numpy(g["anchor_coors"]).cuda()
    nx, ny = int(g["grid_size"][0]), int(g["grid_size"][1])
    mask, integ = am.anchor_mask_device(torch.from_numpy(g["coors"]).cuda(), (nx, ny), ac)
    dense = np.zeros((nx, ny), np.int64)
    np.add.at(dense, (g["coors"][:, 0], g["coors"][:, 1]), 1)
    assert dense.max() > 1
    table = am.integral_table(integ, len(g["coors"]))
    assert table.dtype == np.uint16 and np.array_equal(table[0], dense.cumsum(0).cumsum(1))
    assert np.array_equal(mask[0].cpu().numpy(), g["mask"])
    # more than 65535 rows in one frame -> the int32 table (duplicated pillars just raise the counts)
    many = torch.from_numpy(np.tile(g["coors"], (30, 1))).cuda()
    assert many.shape[0] > 65535
    mask2, integ2 = am.anchor_mask_device(many, (nx, ny), ac)
    table2 = am.integral_table(integ2, many.shape[0])
    assert table2.dtype == np.int32 and np.array_equal(table2[0], 30 * dense.cumsum(0).cumsum(1))
    assert np.array_equal(mask2[0].cpu().numpy(), g["mask"])


def test_empty_frame_and_no_anchors(pf):
    g = util.load_anchor("pillar_small")
    am = _am(pf)
    a = am.AnchorMask(g["anchors_bv"], voxel_size=g["voxel_size"], offset=g["offset"], grid_size=g["grid_size"])
    got = a.create_mask(np.zeros((0, 3), np.int32), g["grid_size"], g["voxel_size"], g["offset"])
    assert got.shape == g["mask"].shape and not got.any()
    one = a.create_mask(g["coors"][:1], g["grid_size"], g["voxel_size"], g["offset"])
    assert np.array_equal(one, oracle.anchors_mask(g["coors"][:1], g["grid_size"][:2], g["anchor_coors"]))
    none = am.AnchorMask(np.zeros((0, 4), np.float32), voxel_size=g["voxel_size"], offset=g["offset"],
                         grid_size=g["grid_size"])
    assert none.create_mask(g["coors"], g["grid_size"], g["voxel_size"], g["offset"]).shape == (0,)


def test_corner_index_rules(pf):
    """Rectangles that touch the grid border, degenerate ones, and a negative maximum (numpy wrap-around)."""
    am = _am(pf)
    nx, ny = 37, 53                                      # not multiples of 32: partial strips in both scans
    rng = np.random.default_rng(5)
    coors = np.stack([rng.integers(0, nx, 400), rng.integers(0, ny, 400), np.zeros(400, np.int64)], 1).astype(np.int32)
    coors = np.unique(coors, axis=0)
    lo = np.stack([rng.integers(0, nx, 3000), rng.integers(0, ny, 3000)], 1)
    hi = np.minimum(lo + np.stack([rng.integers(0, 9, 3000), rng.integers(0, 9, 3000)], 1), [nx - 1, ny - 1])
    ac = np.concatenate([lo, hi], 1).astype(np.int32)
    ac = np.concatenate([ac, [[0, 0, nx - 1, ny - 1], [0, 0, 0, 0], [nx - 1, ny - 1, nx - 1, ny - 1],
                              [0, 0, -1, -1], [0, 3, -2, 20], [5, 0, 20, -3]]]).astype(np.int32)
    want = oracle.anchors_mask(coors, (nx, ny), ac)       # numpy wraps the negative indices like numba does
    mask, _ = am.anchor_mask_device(torch.from_numpy(coors).cuda(), (nx, ny), torch.from_numpy(ac).cuda())
    assert np.array_equal(mask[0].cpu().numpy(), want)


@pytest.mark.parametrize("gkey,fmap", [("20cm", (400, 400)), ("nuscene", (256, 240))])
def test_batch_from_front_end_matches_oracle_per_frame(pf, gkey, fmap):
    """Full-size grids: the mask of every frame of a voxelised batch (slab layout, device coordinates, no PCIe)
    equals the oracle run on that frame's coordinates; 1.44 M anchors at 20 cm like the reference's set."""
    cfg = cfgmod.builtin_config(gkey)
    geom = cfgmod.derive_grid(cfg)
    B, n = 3, 60000
    frames = [synth.make_frame(k, geom, n, 77 + i) for i, k in enumerate(["S1", "S2", "S3"])]
    frames[1] = frames[1][:2000]                          # a sparse frame
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = torch.tensor(np.concatenate([[0], np.cumsum([len(f) for f in frames])]), dtype=torch.int32).cuda()
    fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n, want_voxels=False)
    fe.voxelize_decorate(pts, offs, keep_cell_map=False)
    bv = synth.anchors_bv(geom, fmap)
    a = _am(pf).AnchorMask(bv, voxel_size=np.asarray(cfg["voxel_size"], np.float32), offset=geom.offset,
                           grid_size=geom.grid_size)
    assert a.n_anchors == 9 * fmap[0] * fmap[1]
    ac = oracle.anchor_coor(bv, np.asarray(cfg["voxel_size"], np.float32), geom.offset, geom.grid_size)
    assert np.array_equal(a.anchors_coors, ac)
    counts = fe.voxel_num.cpu().numpy()
    for rep in range(2):
        mask = fe.anchor_mask(a).cpu().numpy()
        for f in range(B):
            c = fe.frame_outputs(f, counts)[1].cpu().numpy()
            want = oracle.anchors_mask(c[:, :3], geom.grid_size[:2], ac)
            assert np.array_equal(mask[f], want), f"{gkey} frame {f} rep {rep}"
            assert 0 < want.sum() < want.size


def test_merged_batch_layout(pf):
    """merge_second_batch layout (frame id in the last column, framework/utils.py:38-41) without voxel counts, and
    the voxeliser's compact layout with them."""
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 3000
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, 9000, 31 + i) for i, k in enumerate(["S1", "S3"])]
    vg = pf.VoxelGenerator(cfg)
    coors = vg.generate_batch(frames, device_outputs=False)["coordinates"]
    assert coors.shape[1] == 4 and set(np.unique(coors[:, 3])) == {0, 1}
    bv = synth.anchors_bv(geom, (100, 100))
    a = _am(pf).AnchorMask(bv, config=cfg)
    got = a.create_mask_batch(torch.from_numpy(coors).cuda(), None, 2, 0).cpu().numpy()
    for f in range(2):
        want = oracle.anchors_mask(coors[coors[:, 3] == f][:, :3], geom.grid_size[:2], a.anchors_coors)
        assert np.array_equal(got[f], want)
    # compact device layout: rows beyond sum(voxel_num) are garbage and must be ignored
    padded = np.concatenate([coors, np.full((500, 4), 7, np.int32)])
    vnum = torch.tensor([(coors[:, 3] == 0).sum(), (coors[:, 3] == 1).sum()], dtype=torch.int32).cuda()
    got2 = a.create_mask_batch(torch.from_numpy(padded).cuda(), vnum, 2, 0).cpu().numpy()
    assert np.array_equal(got2, got)
