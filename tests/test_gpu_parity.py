"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the oracle and the reference's
golden vectors.  Integer / index / copy work is bit-exact; the fp32 decoration is held to
1e-6 * max(1, |xyz|max of the pillar)."""
import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import cfgmod, synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def pfm(pf):
    assert torch.cuda.is_available()
    return pf


def _np(t):
    return t.detach().cpu().numpy()


# ------------------------------------------------------------------------------------------------ voxelise
@pytest.mark.parametrize("gkey", util.GRID_KEYS)
def test_voxelize_small_golden(pfm, gkey):
    """Every small fixture of one grid through ONE generator instance (also exercises the grid self-reset)."""
    by_caps = {}
    for name in util.SMALL_NAMES:
        g = util.load_small(gkey, name)
        by_caps.setdefault((int(g["max_voxels"]), int(g["max_points"])), []).append((name, g))
    for (mv, mp), cases in by_caps.items():
        cfg = cfgmod.builtin_config(gkey)
        cfg["max_voxels"], cfg["max_num_points"] = mv, mp
        vg = pfm.VoxelGenerator(cfg)
        for rep in range(2):
            for name, g in cases:
                got = vg.generate(g["points"])
                util.assert_vox_equal(got, (g["voxels"], g["coors"], g["num"]), f"{gkey}/{name} rep{rep}")
        vg.close()


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("kind", ["S1", "S2", "S3", "S4"])
def test_voxelize_full_size(pfm, gkey, kind):
    cfg, geom, pts, h = util.full_frame(gkey, kind)
    vg = pfm.VoxelGenerator_trt(cfg)
    v, c, n, vn = vg.generate(pts)
    want = oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"])
    assert vn == len(want[2])
    util.assert_vox_equal((v, c, n), want, f"{gkey}/{kind}")
    # device in -> device out, same bits
    dv, dc, dn, _ = vg.generate(torch.from_numpy(pts).cuda())
    assert dv.is_cuda and np.array_equal(_np(dv).view(np.uint32), v.view(np.uint32))
    assert np.array_equal(_np(dc), c) and np.array_equal(_np(dn), n)
    vg.close()
    # the pin to the REFERENCE's own outputs at full size: loud when it cannot be applied (never a silent pass)
    if util.sha(pts) != h["points_sha"]:
        pytest.xfail(f"{gkey}/{kind}: the synthetic frame is not bit-identical on this host (numpy / libm differ from the "
                     "build container), so the committed reference hashes do not apply; oracle parity above still held")
    assert util.sha(v) == h["voxels_sha"] and util.sha(c) == h["coors_sha"] and util.sha(n) == h["num_sha"]


def test_voxelize_feature_counts_and_strides(pfm):
    """F = 3, 5, 7 use the generic (non-float4) load path."""
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 400
    geom = cfgmod.derive_grid(cfg)
    base = synth.make_frame("S1", geom, 5000, 77, 600)
    rng = np.random.default_rng(5)
    for f in (3, 5, 7):
        pts = np.ascontiguousarray(np.concatenate([base[:, :3], rng.random((5000, f - 3), dtype=np.float32)], axis=1))
        vg = pfm.VoxelGenerator(cfg)
        util.assert_vox_equal(vg.generate(pts), oracle.generate(geom, pts, 400, 15), f"F={f}")
        vg.close()


@pytest.mark.parametrize("name", util.EXTRA_NAMES)
def test_voxelize_extra_geometries(pfm, name):
    """True 3-D grids (nz > 1) and F = 3 / 6 point rows against the reference's own outputs."""
    g, cfg = util.load_extra(name)
    vg = pfm.VoxelGenerator(cfg)
    for rep in range(2):
        util.assert_vox_equal(vg.generate(g["points"]), (g["voxels"], g["coors"], g["num"]), f"{name} rep{rep}")
    out = vg.generate_batch([g["points"], g["points"][: len(g["points"]) // 3], g["points"]], device_outputs=False)
    geom = cfgmod.derive_grid(dict(cfg))
    refs = [oracle.generate(geom, p, cfg["max_voxels"], cfg["max_num_points"])
            for p in (g["points"], g["points"][: len(g["points"]) // 3], g["points"])]
    assert out["voxel_num"].tolist() == [len(r[2]) for r in refs]
    util.assert_vox_equal((out["voxels"], out["coordinates"][:, :3], out["num_points_per_voxel"]),
                          tuple(np.concatenate([r[k] for r in refs]) for k in range(3)), f"{name} batch")
    vg.close()


def test_voxelize_random_geometries(pfm):
    """30 seeded random grids / caps / feature counts / cloud sizes (incl. empty, duplicate and boundary points)."""
    rng = np.random.default_rng(2024)
    for case in range(30):
        nx, ny, nz = (int(rng.integers(1, 40)), int(rng.integers(1, 40)), int(rng.integers(1, 5)))
        vs = np.array([0.5, 0.25, 1.0])
        lo = np.array([-3.0, 1.0, -2.0])
        f = int(rng.integers(3, 7))
        cfg = dict(detection_range=list(lo) + list(lo + vs * [nx, ny, nz]), voxel_size=list(vs),
                   max_voxels=int(rng.integers(1, 400)), max_num_points=int(rng.integers(1, 40)),
                   num_point_features=f, batch_size=1, device=0)
        geom = cfgmod.derive_grid(dict(cfg))
        n = int(rng.choice([0, 1, 7, 300, 5000, 40000]))
        span = vs * [nx, ny, nz]
        pts = (lo - 0.2 * span + rng.random((n, 3)) * 1.4 * span).astype(np.float32)
        pts = np.ascontiguousarray(np.concatenate([pts, rng.random((n, f - 3), dtype=np.float32)], axis=1))
        if n > 8:
            pts[rng.integers(0, n, n // 4)] = pts[rng.integers(0, n, n // 4)]
            k = rng.integers(0, n, n // 8)
            pts[k, 1] = (lo[1] + vs[1] * rng.integers(0, ny + 1, len(k))).astype(np.float32)
        vg = pfm.VoxelGenerator(cfg)
        want = oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"])
        util.assert_vox_equal(vg.generate(pts), want, f"case {case}: grid {(nx, ny, nz)} F={f} n={n} cfg={cfg}")
        vg.close()


def test_voxelize_batch_collated(pfm):
    """generate_batch == merge_second_batch of per-frame results (frame id in the LAST coors column)."""
    cfg = cfgmod.builtin_config("16cm")
    cfg["max_voxels"] = 900
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, n, 300 + i, 700) for i, (k, n) in
              enumerate([("S1", 9000), ("S3", 2500), ("S2", 4096), ("S4", 3000), ("S1", 1)])]
    frames.insert(2, np.zeros((0, 4), np.float32))
    vg = pfm.VoxelGenerator(cfg)
    out = vg.generate_batch(frames, device_outputs=False)
    refs = [oracle.generate(geom, f, 900, 15) for f in frames]
    assert out["voxel_num"].tolist() == [len(r[2]) for r in refs]
    want_v = np.concatenate([r[0] for r in refs])
    want_c = np.concatenate([np.pad(r[1], ((0, 0), (0, 1)), constant_values=i) for i, r in enumerate(refs)])
    want_n = np.concatenate([r[2] for r in refs])
    util.assert_vox_equal((out["voxels"], out["coordinates"], out["num_points_per_voxel"]), (want_v, want_c, want_n))
    # the staged host-frame path again: a bigger batch (staging grows), a smaller one right after (staging reused while the
    # previous DMA may still be in flight), a Fortran-ordered frame, and a mixed list (CUDA tensor inside -> device path)
    import torch
    big = [synth.make_frame("S2", geom, 40000, 400 + i, 700) for i in range(3)]
    for batch in (big, frames[:2], [np.asfortranarray(frames[0]), frames[1]], [frames[0], torch.from_numpy(frames[1]).cuda()]):
        got = vg.generate_batch(batch, device_outputs=False)
        rb = [oracle.generate(geom, np.ascontiguousarray(f.cpu().numpy() if hasattr(f, "cpu") else f), 900, 15) for f in batch]
        assert got["voxel_num"].tolist() == [len(r[2]) for r in rb]
        util.assert_vox_equal((got["voxels"], got["coordinates"], got["num_points_per_voxel"]),
                              (np.concatenate([r[0] for r in rb]),
                               np.concatenate([np.pad(r[1], ((0, 0), (0, 1)), constant_values=i) for i, r in enumerate(rb)]),
                               np.concatenate([r[2] for r in rb])))
    vg.close()


def test_voxelize_properties_full_size(pfm):
    """Size-independent properties at BASELINE.json's full frame size (20 cm, 120k points, overflow case)."""
    cfg, geom, pts, _ = util.full_frame("20cm", "S2")
    vg = pfm.VoxelGenerator(cfg)
    v, c, n = vg.generate(pts)
    # idempotence
    v2, c2, n2 = vg.generate(pts)
    util.assert_vox_equal((v2, c2, n2), (v, c, n), "repeat")
    # coordinates unique and in range; counts in [1, P]; padding rows zero
    lin = c[:, 0].astype(np.int64) * geom.grid_size[1] + c[:, 1]
    assert len(np.unique(lin)) == len(lin) and np.all((c >= 0) & (c < geom.grid_size))
    assert n.min() >= 1 and n.max() <= cfg["max_num_points"]
    pad = np.arange(cfg["max_num_points"])[None, :] >= n[:, None]
    assert not v[pad].any()
    # every stored point lies in its voxel's cell
    q = np.floor((v[..., :3] - geom.offset) / geom.voxel_size).astype(np.int64)
    assert np.array_equal(q[~pad], np.repeat(c, n, axis=0))
    # permuting the points after the cut-off index i* does not change the output (the loop never reaches them)
    from oracle import spec_parallel
    cell, _ = spec_parallel.quantize(geom, pts)
    ins = np.nonzero(cell >= 0)[0]
    _, first_pos = np.unique(cell[ins], return_index=True)
    firsts = np.sort(ins[first_pos])
    assert len(firsts) > cfg["max_voxels"], "S2 must overflow max_voxels"
    istar = int(firsts[cfg["max_voxels"]])
    tail = pts[istar + 1:].copy()
    np.random.default_rng(0).shuffle(tail)
    pts2 = np.ascontiguousarray(np.concatenate([pts[:istar + 1], tail]))
    util.assert_vox_equal(vg.generate(pts2), (v, c, n), "tail shuffle")
    # appending a copy of a stored point right after it puts the copy in the next slot (or drops it when full)
    k = int(np.argmax(n < cfg["max_num_points"]))
    src = int(np.nonzero((pts == v[k, n[k] - 1]).all(1))[0][0])
    pts3 = np.ascontiguousarray(np.insert(pts, src + 1, pts[src], axis=0))
    v3, c3, n3 = vg.generate(pts3)
    util.assert_vox_equal((v3, c3, n3), oracle.generate(geom, pts3, cfg["max_voxels"], cfg["max_num_points"]), "dup")
    vg.close()


def test_voxelize_rejects_bad_input(pfm):
    vg = pfm.VoxelGenerator(cfgmod.builtin_config("20cm"))
    with pytest.raises(TypeError):
        vg.generate(np.zeros((4, 4), np.float64))
    with pytest.raises(ValueError):
        vg.generate(np.zeros((4,), np.float32))
    with pytest.raises(ValueError):
        vg.generate(np.zeros((4, 2), np.float32))
    v, c, n = vg.generate(np.zeros((0, 4), np.float32))
    assert v.shape == (0, 15, 4) and c.shape == (0, 3) and n.shape == (0,)
    vg.close()


def test_oversized_shapes_are_refused_not_crashed(pfm):
    """(max_points, features) whose gather / decoration kernel cannot fit a CTA's shared memory are refused at
    validation time with PF_ERR_UNSUPPORTED, not with a raw CUDA launch error; shapes inside the limit still work."""
    big_v = torch.zeros((3, 4096, 4), device="cuda")
    n = torch.ones((3,), dtype=torch.int32, device="cuda")
    c = torch.zeros((3, 3), dtype=torch.int32, device="cuda")
    with pytest.raises(pfm.PillarFrontError, match="shared memory"):
        pfm.decorate(big_v, n, c, 0.2, 0.2, 0.1, 0.1)
    cfg = dict(detection_range=[0.0, 0.0, 0.0, 8.0, 8.0, 1.0], voxel_size=[1.0, 1.0, 1.0], max_voxels=10,
               max_num_points=255, num_point_features=16, batch_size=1, device=0)
    vg = pfm.VoxelGenerator(dict(cfg))
    with pytest.raises(pfm.PillarFrontError, match="shared memory"):
        vg.generate(np.zeros((5, 16), np.float32))
    vg.close()
    cfg["max_num_points"] = 120                      # 8 * 120 * 37 * 4 bytes = 142 KB: fits
    geom = cfgmod.derive_grid(dict(cfg))
    pts = np.random.default_rng(1).random((500, 16), dtype=np.float32) * 8
    vg = pfm.VoxelGenerator(dict(cfg))
    util.assert_vox_equal(vg.generate(pts), oracle.generate(geom, pts, 10, 120), "P=120 F=16")
    vg.close()
    ok_v = torch.rand((5, 600, 4), device="cuda")   # F = 4, P > 512 -> generic kernel, 8 * 600 * 13 * 4 = 250 KB: refused
    with pytest.raises(pfm.PillarFrontError, match="shared memory"):
        pfm.decorate(ok_v, torch.full((5,), 600, dtype=torch.int32, device="cuda"), torch.zeros((5, 3), dtype=torch.int32, device="cuda"), 0.2, 0.2, 0.1, 0.1)


# ------------------------------------------------------------------------------------------------ decorate
@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("name", ["s1", "s2_ring", "s4_edges", "dense_cap", "n5"])
def test_decorate_matches_reference(pfm, gkey, name):
    g = util.load_small(gkey, name)
    cfg = util.small_config(gkey, g)
    geom = cfgmod.derive_grid(cfg)
    consts = pfm.decorate_constants(cfg["voxel_size"], geom.offset)
    assert consts == oracle.decorate_consts(cfg["voxel_size"], geom.offset)
    got = _np(pfm.decorate(torch.from_numpy(g["voxels"]).cuda(), torch.from_numpy(g["num"]).cuda(),
                           torch.from_numpy(g["coors"]).cuda(), *consts))
    tol = oracle.decorate_tolerance(g["voxels"])
    for want in (g["features"], oracle.decorate(g["voxels"], g["num"], g["coors"], *consts)):
        both_nan = np.isnan(got) & np.isnan(want)
        assert got.shape == want.shape
        assert np.all((np.abs(got.astype(np.float64) - want) <= tol) | both_nan)
    assert np.array_equal(got[..., :4].view(np.uint32), g["voxels"].view(np.uint32))     # pass-through is exact
    # the centre-offset columns see identical rounding steps -> bit-exact against the C oracle
    want = oracle.decorate(g["voxels"], g["num"], g["coors"], *consts)
    assert np.array_equal(got[..., 7:9], want[..., 7:9])


def test_pointnet_module(pfm):
    """PointNet keeps the reference's state-dict keys and forward result shape."""
    cfg = cfgmod.builtin_config("20cm")
    vgcfg = dict(cfg)
    vg = pfm.VoxelGenerator(vgcfg)
    net = pfm.PointNet(cfg["num_point_features"], cfg["voxel_size"], vgcfg["detection_offset"]).cuda().eval()
    assert sorted(net.state_dict()) == sorted(
        ["pfn_layers.0.weight", "pfn_layers.1.weight", "pfn_layers.1.bias", "pfn_layers.1.running_mean",
         "pfn_layers.1.running_var", "pfn_layers.1.num_batches_tracked"])
    pts = synth.make_frame("S1", vg.geometry, 20000, 9, 3000)
    v, c, n = vg.generate(pts, device_outputs=True)
    with torch.no_grad():
        out = net(v, n, c)
        feats = torch.from_numpy(oracle.decorate(_np(v), _np(n), _np(c), net.vx, net.vy, net.x_offset,
                                                 net.y_offset)).cuda()
        want = torch.max(net.pfn_layers(feats.permute(0, 2, 1).contiguous()).permute(0, 2, 1), dim=1)[0]
    assert out.shape == (len(n), 64)
    assert torch.allclose(out, want, rtol=1e-4, atol=1e-4)
    vg.close()


# ------------------------------------------------------------------------------------------------ single-frame fast path
def test_generate_outputs_are_owned_by_the_caller(pfm):
    """The numpy arrays generate() returns are views of pinned blocks from a recycling pool: results a caller keeps must
    stay intact over any number of later calls (more of them than the pool holds), a block is reused only after every
    array from it has been released, and device outputs are fresh tensors."""
    import gc
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 1200
    geom = cfgmod.derive_grid(cfg)
    vg = pfm.VoxelGenerator(dict(cfg))
    frames = [np.ascontiguousarray(synth.make_frame(["S1", "S3", "S2"][i % 3], geom, 3000 + 517 * i, 800 + i, 400), np.float32)
              for i in range(12)]
    want = [oracle.generate(geom, f, 1200, 15) for f in frames]
    kept = [vg.generate(f) for f in frames]                    # 12 live results > the 8 pooled blocks -> pageable fall-back
    for i, (g, w) in enumerate(zip(kept, want)):
        util.assert_vox_equal(g, w, f"kept result {i}")
    assert len(vg._pool) == 8 and not any(b.free() for b in vg._pool)
    first_block = vg._pool[0]
    del kept, g
    gc.collect()
    assert all(b.free() for b in vg._pool)
    again = vg.generate(frames[3])
    assert again[0].base is first_block.nv                     # a released block is handed out again
    util.assert_vox_equal(again, want[3], "recycled block")
    dev = [vg.generate(f, device_outputs=True) for f in frames[:3]]
    for i, d in enumerate(dev):                                # device outputs: fresh tensors, untouched by later calls
        util.assert_vox_equal(tuple(_np(t) for t in d), want[i], f"device result {i}")
    assert len({d[0].data_ptr() for d in dev}) == 3
    # a longer frame (another capacity bucket) and another feature count on the same instance
    big = np.ascontiguousarray(synth.make_frame("S3", geom, 40000, 5), np.float32)
    util.assert_vox_equal(vg.generate(big), oracle.generate(geom, big, 1200, 15), "second capacity bucket")
    five = np.ascontiguousarray(np.concatenate([frames[0], frames[0][:, :1]], axis=1))
    util.assert_vox_equal(vg.generate(five), oracle.generate(geom, five, 1200, 15), "F = 5 on the same instance")
    util.assert_vox_equal(vg.generate(frames[1]), want[1], "back to F = 4")
    vg.close()


def test_generate_and_scatter_matches_the_module_sequence(pfm):
    """generate_and_scatter (one graph: H2D, voxelise, decorate + fused PFN, fill) against generate -> PointNet.eval() ->
    PointPillarsScatter on the same frames, frame after frame on one instance (sizes change between calls)."""
    cfg = cfgmod.builtin_config("nuscene")
    cfg["max_voxels"] = 2000
    geom = cfgmod.derive_grid(cfg)
    torch.manual_seed(3)
    net = pfm.PointNet(4, cfg["voxel_size"], geom.offset).cuda().eval()
    sc = pfm.PointPillarsScatter(1, geom.grid_size, 64)
    vg = pfm.VoxelGenerator(dict(cfg))
    for i, (kind, n) in enumerate([("S1", 9000), ("S3", 6000), ("S1", 0), ("S2", 12800), ("S1", 9000)]):
        pts = np.ascontiguousarray(synth.make_frame(kind, geom, max(n, 8), 900 + i, 700)[:n], np.float32)
        with torch.no_grad():
            canvas, coors, v = vg.generate_and_scatter(pts, net)
            got = canvas.clone()
            wv, wc, wn = oracle.generate(geom, pts, 2000, 100)
            assert v == len(wn) and np.array_equal(_np(coors)[:, :3], wc)
            dv, dc, dn = vg.generate(pts, device_outputs=True)
            want = sc(net(dv, dn, dc).reshape(-1, 64), dc) if v else torch.zeros_like(got)
        assert torch.equal(got != 0, want != 0) or torch.equal(got == 0, want == 0), f"frame {i}: occupancy pattern"
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-5), f"frame {i} ({kind}, {n} points): {(got - want).abs().max()}"
    vg.close()
