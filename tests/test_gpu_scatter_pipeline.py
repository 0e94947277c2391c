"""GPU parity tests for the BEV scatter (bit-exact copy work) and the fused batched pipeline."""
import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import cfgmod, synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _np(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("gkey,name", [("20cm", "s1"), ("20cm", "s4_edges"), ("nuscene", "s1"), ("nuscene", "s4_edges")])
def test_scatter_matches_reference_golden(pf, gkey, name):
    g = util.load_small(gkey, name)
    nx, ny = int(g["grid_size"][0]), int(g["grid_size"][1])
    feats = torch.from_numpy(g["pillar_feats"]).cuda()
    coors = torch.from_numpy(g["coors"]).cuda()
    sc = pf.PointPillarsScatter(batch_size=1, output_shape=g["grid_size"], num_input_features=64)
    for rep in range(2):                                     # second call checks the map put itself back
        canvas = _np(sc(feats, coors))
        assert canvas.shape == tuple(g["canvas_shape"])
        assert util.sha(canvas) == str(g["canvas_sha"]), f"rep {rep}"
    # channel-major variants
    cm = pf.Scatter(g["grid_size"], 64).run(feats.t().contiguous(), coors)
    assert util.sha(_np(cm)) == str(g["canvas_sha"])
    cm3 = pf.Scatter(g["grid_size"], 64).run(feats.t().contiguous()[None], coors)
    assert util.sha(_np(cm3)) == str(g["canvas_sha"])
    pc = pf.Scatter_cuda(g["grid_size"], 64)
    out = pc.run(feats.t().contiguous(), coors)
    assert out.data_ptr() == pc.canvas_d.data_ptr() and util.sha(_np(out)) == str(g["canvas_sha"])
    out = pc.run((2 * feats).t().contiguous()[:, :7], coors[:7])          # persistent canvas is fully rewritten
    assert np.array_equal(_np(out), oracle.scatter_cmajor(2 * g["pillar_feats"].T[:, :7], g["coors"][:7], nx, ny))


def test_scatter_batched_and_backward(pf):
    """B > 1: per-frame batch-1 rule, frame id in the last coords column; gradient = gather."""
    cfg = cfgmod.builtin_config("nuscene")
    geom = cfgmod.derive_grid(cfg)
    nx, ny = int(geom.grid_size[0]), int(geom.grid_size[1])
    rng = np.random.default_rng(3)
    B, C = 3, 64
    coors, feats = [], []
    for b, v in enumerate([700, 0, 1300]):
        cells = rng.choice(nx * ny, size=v, replace=False)
        coors.append(np.stack([cells // ny, cells % ny, np.zeros(v, np.int64), np.full(v, b)], 1).astype(np.int32))
        feats.append(rng.standard_normal((v, C)).astype(np.float32))
    coors, feats = np.concatenate(coors), np.concatenate(feats)
    want = oracle.scatter(feats, coors, nx, ny, batch=B)
    sc = pf.PointPillarsScatter(batch_size=B, output_shape=geom.grid_size, num_input_features=C)
    ft = torch.from_numpy(feats).cuda().requires_grad_(True)
    canvas = sc(ft, torch.from_numpy(coors).cuda())
    assert np.array_equal(_np(canvas), want)
    gc = torch.from_numpy(rng.standard_normal(want.shape).astype(np.float32)).cuda()
    canvas.backward(gc)
    assert np.array_equal(_np(ft.grad), oracle.scatter_backward(_np(gc), coors, batch=B))
    # odd channel count and a plane that is not a multiple of 128 cells -> generic load path / partial warps
    nx2, ny2 = 37, 52
    v = 300
    cells = rng.choice(nx2 * ny2, size=v, replace=False)
    co = np.stack([cells // ny2, cells % ny2, np.zeros(v, np.int64)], 1).astype(np.int32)
    fe = rng.standard_normal((v, 6)).astype(np.float32)
    sc2 = pf.PointPillarsScatter(1, np.array([nx2, ny2, 1]), 6)
    assert np.array_equal(_np(sc2(torch.from_numpy(fe).cuda(), torch.from_numpy(co).cuda())),
                          oracle.scatter(fe, co, nx2, ny2))
    nx3, ny3 = 5, 7          # plane % 4 != 0 -> scalar fall-back kernel
    co3 = np.array([[0, 0, 0], [4, 6, 0], [2, 3, 0], [2, 3, 0]], np.int32)      # duplicate cell: last row wins
    fe3 = rng.standard_normal((4, 3)).astype(np.float32)
    sc3 = pf.PointPillarsScatter(1, np.array([nx3, ny3, 1]), 3)
    for _ in range(2):
        assert np.array_equal(_np(sc3(torch.from_numpy(fe3).cuda(), torch.from_numpy(co3).cuda())),
                              oracle.scatter(fe3, co3, nx3, ny3))


@pytest.fixture
def fill_variant(request):
    """Select a canvas-fill kernel (pf_debug_set_fill_variant) for one test, then restore the library default."""
    import importlib
    L = importlib.import_module("3d_object_detection_b200._lib").load()
    assert L.pf_debug_set_fill_variant(request.param) == request.param
    yield request.param
    L.pf_debug_set_fill_variant(-1)


@pytest.mark.parametrize("fill_variant", [0, 1, 2, 3, 4, 5, 6], indirect=True)
def test_fill_variants_write_the_same_canvas(pf, fill_variant):
    """16-byte stores, 32-byte stores and the two TMA bulk-store kernels must produce identical canvases: the
    reference-made hashes, a dense block (every cell of many 128 / 256-cell chunks occupied -> multi-pass staging),
    a plane whose last chunk is partial (1000 x 1000), few channels, a batch, and the fused pipeline's cell map."""
    for gkey, name in (("20cm", "s1"), ("nuscene", "s4_edges")):
        g = util.load_small(gkey, name)
        sc = pf.PointPillarsScatter(batch_size=1, output_shape=g["grid_size"], num_input_features=64)
        feats, coors = torch.from_numpy(g["pillar_feats"]).cuda(), torch.from_numpy(g["coors"]).cuda()
        for rep in range(2):
            assert util.sha(_np(sc(feats, coors))) == str(g["canvas_sha"]), f"{gkey}/{name} rep {rep}"
    rng = np.random.default_rng(17 + fill_variant)
    nx = ny = 1000                                           # 16 cm grid: 1e6 cells = 7812.5 chunks of 128
    for C, B in ((64, 2), (6, 3), (3, 1)):
        coors, feats = [], []
        for b in range(B):
            bx, by = np.meshgrid(np.arange(300, 340), np.arange(500 + b, 548 + b), indexing="ij")     # dense 40 x 48 block
            dense = bx.ravel() * ny + by.ravel()
            tail = np.arange(nx * ny - 70, nx * ny)          # the partial last chunk, fully occupied
            rest = rng.choice(nx * ny - 100, size=3000, replace=False)
            cells = np.unique(np.concatenate([dense, tail, rest]))
            rng.shuffle(cells)
            coors.append(np.stack([cells // ny, cells % ny, np.zeros_like(cells), np.full_like(cells, b)], 1).astype(np.int32))
            feats.append(rng.standard_normal((len(cells), C)).astype(np.float32))
        coors, feats = np.concatenate(coors), np.concatenate(feats)
        sc = pf.PointPillarsScatter(batch_size=B, output_shape=np.array([nx, ny, 1]), num_input_features=C)
        for rep in range(2):
            got = sc(torch.from_numpy(feats).cuda(), torch.from_numpy(coors if B > 1 else coors[:, :3]).cuda())
            assert np.array_equal(_np(got), oracle.scatter(feats, coors, nx, ny, batch=B)), f"C={C} B={B} rep {rep}"
    # the fused pipeline (MODE 1: the voxeliser's grid is the map), compact layout
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 3000
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, n, 70 + i, 9) for i, (k, n) in enumerate([("S1", 9000), ("S3", 5000), ("S1", 4000)])]
    fe = pf.PillarFrontEnd(cfg, max_batch=3, max_points_per_frame=9000, device=0, compact=True)
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = torch.tensor(np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.int32)).cuda()
    pillar = torch.randn((3 * 3000, 64), device="cuda")
    for rep in range(2):
        canvas = fe.step(pts, offs, pillar_feats=pillar)
        counts = _np(fe.voxel_num)
        starts = np.concatenate([[0], np.cumsum(counts)])
        for f in range(3):
            sl = slice(int(starts[f]), int(starts[f + 1]))
            want = oracle.scatter(_np(pillar[sl]), _np(fe.coors[sl])[:, :3], fe.nx, fe.ny)[0]
            assert np.array_equal(_np(canvas[f]), want), f"pipeline frame {f} rep {rep}"
    fe.close()


@pytest.mark.parametrize("case", ["nuscene_s1", "20cm_s4_edges", "20cm_dup"])
def test_scatter_backward_matches_reference_golden(pf, case):
    """Row A7 pinned: pf_scatter_backward through autograd against the gradient the REFERENCE's own
    PointPillarsScatter returned for the same coordinates and grad_canvas (oracle/make_golden_backward.py)."""
    z, gc, coors = util.load_backward(case)
    grid = z[case + "_grid"]
    want = z[case + "_grad_feats"]
    feats = torch.zeros((len(coors), 64), device="cuda").requires_grad_(True)
    sc = pf.PointPillarsScatter(batch_size=1, output_shape=grid, num_input_features=64)
    sc(feats, torch.from_numpy(coors).cuda()).backward(torch.from_numpy(gc).cuda())
    assert np.array_equal(_np(feats.grad).view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("gkey,compact", [("20cm", False), ("20cm", True), ("nuscene", False), ("10cm", True)])
def test_pipeline_batched(pf, gkey, compact):
    """voxelise -> decorate -> scatter for a ragged batch in one launch sequence, slab and collated layouts,
    eagerly and as a replayed CUDA graph."""
    cfg = cfgmod.builtin_config(gkey)
    cfg["max_voxels"] = 2500
    geom = cfgmod.derive_grid(cfg)
    P, mv = cfg["max_num_points"], cfg["max_voxels"]
    spec = [("S1", 30000), ("S3", 9000), ("S2", 20032), ("S4", 6000), ("S1", 0), ("S1", 17)]
    frames = [synth.make_frame(k, geom, max(n, 8), 40 + i, 2000)[:n if k != "S4" else None] for i, (k, n) in enumerate(spec)]
    frames = [np.ascontiguousarray(f, np.float32) for f in frames]
    B = len(frames)
    sizes = [len(f) for f in frames]
    fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=max(sizes), device=0, compact=compact)
    dev = fe.device
    pts = torch.from_numpy(np.concatenate(frames)).to(dev)
    offs = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32), device=dev)
    pillar = torch.randn((B * mv, 64), device=dev)
    consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)
    refs = [oracle.generate(geom, f, mv, P) for f in frames]
    counts_want = [len(r[2]) for r in refs]
    starts = np.concatenate([[0], np.cumsum(counts_want)]) if compact else np.arange(B + 1) * mv

    def check(canvas):
        counts = _np(fe.voxel_num)
        assert counts.tolist() == counts_want
        pil = _np(pillar)
        for f, (v, c, n) in enumerate(refs):
            sl = slice(int(starts[f]), int(starts[f]) + len(n))
            util.assert_vox_equal((_np(fe.voxels[sl]), _np(fe.coors[sl])[:, :3], _np(fe.num_points[sl])), (v, c, n), f"frame {f}")
            assert np.all(_np(fe.coors[sl])[:, 3] == f)
            want_f = oracle.decorate(v, n, c, *consts)
            util.assert_features_close(_np(fe.features[sl]), want_f, oracle.decorate_tolerance(v), f"features {f}")
            assert np.array_equal(_np(canvas[f]), oracle.scatter(pil[sl], c, fe.nx, fe.ny)[0]), f"canvas {f}"

    for _ in range(2):
        check(fe.step(pts, offs, pillar_feats=pillar))
    fe.capture(pts, offs, pillar)
    for _ in range(2):
        fe.canvas.fill_(7.0)
        fe.features.fill_(7.0)
        check(fe.replay())
    fe.close()


def test_pipeline_full_size_checksums(pf):
    """BASELINE.json configs[1] shape (8 frames x 120k points, 20 cm): size-independent checks at full size."""
    cfg = cfgmod.builtin_config("20cm")
    geom = cfgmod.derive_grid(cfg)
    B, n = 8, cfg["frame_points"]
    kinds = ["S1", "S2", "S3", "S1", "S2", "S3", "S1", "S4"]
    frames = [synth.make_frame(k, geom, n, 900 + i, synth.S1_CELLS["20cm"])[:n] for i, k in enumerate(kinds)]
    fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=n, device=0)
    dev = fe.device
    pts = torch.from_numpy(np.concatenate(frames)).to(dev)
    offs = torch.arange(0, (B + 1) * n, n, dtype=torch.int32, device=dev)
    pillar = torch.randn((B * cfg["max_voxels"], 64), device=dev)
    canvas = fe.step(pts, offs, pillar_feats=pillar)
    counts = _np(fe.voxel_num)
    for f in (0, 1, 7):                                      # three frames against the oracle, bit-exact
        v, c, nn = oracle.generate(geom, frames[f], cfg["max_voxels"], cfg["max_num_points"])
        gv, gc, gn, _ = fe.frame_outputs(f, counts)
        util.assert_vox_equal((_np(gv), _np(gc)[:, :3], _np(gn)), (v, c, nn), f"frame {f}")
    # canvas: gathering it back at the coordinates returns the pillar features; everything else is zero
    for f in range(B):
        gv, gc, gn, _ = fe.frame_outputs(f, counts)
        rows = pillar[f * cfg["max_voxels"]: f * cfg["max_voxels"] + int(counts[f])]
        got = canvas[f][:, gc[:, 0].long(), gc[:, 1].long()].t()
        assert torch.equal(got, rows)
        assert torch.count_nonzero(canvas[f]).item() == torch.count_nonzero(rows).item()
    # second run on the same handle gives the same bits (grid put itself back)
    c1 = canvas.clone()
    assert torch.equal(fe.step(pts, offs, pillar_feats=pillar), c1)
    fe.close()


# ------------------------------------------------------------------------------------------------ fused PFN (N1)
def _random_pointnet(pf, cfg, geom, seed):
    torch.manual_seed(seed)
    net = pf.PointNet(cfg["num_point_features"], cfg["voxel_size"], geom.offset).cuda()
    bn = net.pfn_layers[1]
    with torch.no_grad():   # non-trivial running statistics and affine terms, some shifts positive (padding can win the max)
        bn.running_mean.copy_(torch.randn(64) * 0.5)
        bn.running_var.copy_(torch.rand(64) * 2 + 0.1)
        bn.weight.copy_(torch.randn(64))
        bn.bias.copy_(torch.randn(64) * 0.5)
        net.pfn_layers[0].weight.mul_(3.0)
    return net.eval()


@pytest.mark.parametrize("gkey", util.GRID_KEYS)
@pytest.mark.parametrize("name", ["s1", "s2_ring", "dense_cap", "n5"])
def test_fused_pfn_matches_reference_math(pf, gkey, name):
    """pf_pfn (decorate + conv + folded BN + ReLU + max in one kernel) against an fp64 restatement evaluated on the
    REFERENCE's own decorated features, and against the unfused torch modules."""
    g = util.load_small(gkey, name)
    cfg = util.small_config(gkey, g)
    geom = cfgmod.derive_grid(cfg)
    net = _random_pointnet(pf, cfg, geom, 3)
    conv, bn = net.pfn_layers[0], net.pfn_layers[1]
    v, n, c = (torch.from_numpy(g[k]).cuda() for k in ("voxels", "num", "coors"))
    with torch.no_grad():
        got = net(v, n, c).reshape(len(g["num"]), 64).cpu().numpy()
        unfused = net.features_to_pillars(net.decorate(v, n, c)).cpu().numpy()
    finite = np.isfinite(g["features"]).all(axis=(1, 2))           # s4-style NaN payloads are not in these fixtures
    want = oracle.pfn(g["features"], conv.weight.detach().cpu().numpy(), bn.weight.detach().cpu().numpy(),
                      bn.bias.detach().cpu().numpy(), bn.running_mean.cpu().numpy(), bn.running_var.cpu().numpy(), bn.eps)
    scale = np.abs(g["features"]).reshape(len(finite), -1).max(axis=1, keepdims=True) + 1.0
    # fp32 conv of 9 terms of magnitude <= scale, times a BN gain of up to ~10: 1e-4 relative to the operand magnitude
    assert np.all(np.abs(got[finite] - want[finite]) <= 1e-4 * scale[finite] + 1e-4 * np.abs(want[finite])), \
        np.abs(got - want).max()
    assert np.allclose(got, unfused, rtol=1e-4, atol=1e-4 * float(scale.max()))
    assert (got >= 0).all()


def test_pointnet_eval_mode_is_differentiable(pf):
    """under model.eval() with autograd on (frozen-BN fine-tuning, saliency) the module must stay
    differentiable like the reference's; the fused kernel (no backward) may only run when no gradient is recorded."""
    g = util.load_small("20cm", "s1")
    cfg = util.small_config("20cm", g)
    geom = cfgmod.derive_grid(cfg)
    net = _random_pointnet(pf, cfg, geom, 9)                 # .eval()
    v, n, c = (torch.from_numpy(g[k]).cuda() for k in ("voxels", "num", "coors"))
    with torch.no_grad():
        fused = net(v, n, c)
        assert not net.wants_grad(v) and net.fusable(v)
    assert fused.grad_fn is None and net.wants_grad(v)
    out = net(v, n, c)                                       # grad mode on, parameters require grad -> torch tail
    assert out.grad_fn is not None
    assert torch.allclose(out, fused, rtol=1e-4, atol=1e-3)
    w = net.pfn_layers[0].weight
    (ga,) = torch.autograd.grad(out.square().sum(), w)
    ref = net.features_to_pillars(net.decorate(v, n, c))
    (gb,) = torch.autograd.grad(ref.square().sum(), w)
    assert torch.equal(ga, gb)
    for p in net.parameters():                               # frozen parameters: nothing to record -> fused again
        p.requires_grad_(False)
    assert net(v, n, c).grad_fn is None


@pytest.mark.parametrize("gkey", ["20cm", "nuscene"])
def test_pipeline_with_fused_pfn(pf, gkey):
    """PillarFrontEnd.step(pfn=PointNet.eval()) = pf_voxelize_pfn + fill, against voxelise -> torch PFN -> scatter."""
    cfg = cfgmod.builtin_config(gkey)
    cfg["max_voxels"] = 700
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, n, 40 + i, 500) for i, (k, n) in enumerate([("S1", 9000), ("S3", 3000), ("S2", 8192)])]
    net = _random_pointnet(pf, cfg, geom, 5)
    fe = pf.PillarFrontEnd(cfg, max_batch=3, max_points_per_frame=9000, device=0)
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = torch.tensor(np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.int32)).cuda()
    with torch.no_grad():
        canvas = fe.step(pts, offs, pfn=net).clone()
        fused_rows = fe.pillar_feats.clone()
        fe.voxelize_decorate(pts, offs, keep_cell_map=False)
        want_rows = net.features_to_pillars(fe.features)
    counts = fe.voxel_num.cpu().numpy()
    for f in range(3):
        lo = f * cfg["max_voxels"]
        a, b = fused_rows[lo:lo + counts[f]], want_rows[lo:lo + counts[f]]
        assert torch.allclose(a, b, rtol=1e-4, atol=2e-3), (a - b).abs().max()
        coors = fe.coors[lo:lo + counts[f]].cpu().numpy()
        want_canvas = oracle.scatter(a.cpu().numpy(), coors[:, :3], fe.nx, fe.ny)[0]
        assert np.array_equal(canvas[f].cpu().numpy(), want_canvas)      # the fill itself stays bit-exact
    # the same step captured into a CUDA graph, without materialising voxels / decorated rows, on new contents
    fe2 = pf.PillarFrontEnd(cfg, max_batch=3, max_points_per_frame=9000, device=0, want_voxels=False)
    pts2 = torch.zeros_like(pts)
    fe2.capture(pts2, offs, pfn=net)
    pts2.copy_(pts)
    for rep in range(2):
        got = fe2.replay()
        assert torch.equal(got, canvas), f"graph replay {rep}"
        assert torch.equal(fe2.voxel_num, fe.voxel_num)
        for f in range(3):
            sl = slice(f * cfg["max_voxels"], f * cfg["max_voxels"] + int(counts[f]))
            assert torch.equal(fe2.pillar_feats[sl], fused_rows[sl])
    fe2.close()
    fe.close()


def test_collate_of_device_outputs_equals_generate_batch(pf):
    """Row N3: merge_second_batch over per-frame device outputs == the collated layout generate_batch writes directly."""
    import importlib
    col = importlib.import_module("3d_object_detection_b200.collate")
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 2500
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, 7000, 91 + i) for i, k in enumerate(["S1", "S3", "S2"])]
    vg = pf.VoxelGenerator(cfg)
    ex = []
    for f in frames:
        v, c, n = vg.generate(f, device_outputs=True)
        ex.append({"voxels": v.clone(), "coordinates": c.clone(), "num_points_per_voxel": n.clone()})
    merged = col.merge_second_batch(ex)
    batch = vg.generate_batch(frames)
    assert merged["coordinates"].is_cuda and merged["coordinates"].shape[1] == 4
    for k in ("voxels", "coordinates", "num_points_per_voxel"):
        assert torch.equal(merged[k], batch[k]), k
    conv = col.example_convert_to_torch(merged, device=merged["voxels"].device)
    assert conv["voxels"].data_ptr() == merged["voxels"].data_ptr()          # no copy for resident tensors


def test_kept_cell_map_must_be_consumed(pf):
    """A voxelise call that keeps its cell map must be followed by exactly one scatter (or a discard): the C ABI
    refuses the next voxelise otherwise instead of silently producing empty frames."""
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 500
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame("S1", geom, 4000, 61, 300), synth.make_frame("S3", geom, 3000, 62)]
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = torch.tensor([0, 4000, 7000], dtype=torch.int32).cuda()
    fe = pf.PillarFrontEnd(cfg, max_batch=2, max_points_per_frame=4000, device=0)
    pillar = torch.randn((2 * 500, 64), device="cuda")
    want = fe.step(pts, offs, pillar_feats=pillar).clone()
    want_counts = fe.voxel_num.clone()
    fe.voxelize_decorate(pts, offs)                                   # keeps the map ...
    with pytest.raises(pf.PillarFrontError, match="not consumed"):
        fe.voxelize_decorate(pts, offs)                               # ... so this is refused
    with pytest.raises(pf.PillarFrontError, match="covers 2 frames"):
        fe.scatter(pillar, n_frames=1)
    fe.discard_cell_map()
    with pytest.raises(pf.PillarFrontError, match="no cell map"):
        fe.scatter(pillar)
    fe.voxelize_decorate(pts, offs, keep_cell_map=False)              # nothing pending after this one
    fe.voxelize_decorate(pts, offs, keep_cell_map=False)
    assert torch.equal(fe.voxel_num, want_counts)
    got = fe.step(pts, offs, pillar_feats=pillar)                     # and the handle still works, bit for bit
    assert torch.equal(got, want) and torch.equal(fe.voxel_num, want_counts)
    fe.close()


def test_sharded_front_end_matches_oracle(pf):
    """ShardedFrontEnd on the visible GPUs (one here, eight on a full box): frame b -> GPU b mod G, every canvas and
    every coordinate list equal to the oracle's for that frame, ragged batch, two submissions in flight per worker."""
    n_gpu = torch.cuda.device_count()
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 1500
    geom = cfgmod.derive_grid(cfg)
    B = 3
    n_frames = n_gpu * B - 1
    frames = [np.ascontiguousarray(synth.make_frame(["S1", "S3", "S2"][i % 3], geom, 3000 + 700 * (i % 4), 500 + i, 400))
              for i in range(n_frames)]
    with pf.ShardedFrontEnd(cfg, devices=n_gpu, frames_per_device=B, max_points_per_frame=6000, depth=2) as sfe:
        assert len(sfe.workers) == n_gpu and sorted(sfe.devices) == list(range(n_gpu))
        stand_in = []
        for w in sfe.workers:                                 # seeded pillar stand-ins, as bench.py feeds the fill
            for s in w.slots:
                s["pillar"].copy_(torch.randn(s["pillar"].shape, generator=torch.Generator().manual_seed(w.device)).to(s["pillar"].device))
            stand_in.append(w.slots[0]["pillar"].cpu().numpy())
        first = sfe.submit(frames)
        second = sfe.submit([torch.from_numpy(f).pin_memory() for f in frames[::-1]])      # pinned tensors: no staging copy
        for order, futs in ((frames, first), (frames[::-1], second)):
            for wi, fut in enumerate(futs):
                r = fut.result()
                assert r.device == sfe.devices[wi] and r.canvas.device.index == r.device
                for k, b in enumerate(r.frame_ids):
                    v, c, n = oracle.generate(geom, order[b], cfg["max_voxels"], cfg["max_num_points"])
                    assert int(r.counts[k]) == len(n)
                    rows = slice(k * cfg["max_voxels"], k * cfg["max_voxels"] + len(n))
                    assert np.array_equal(_np(r.coors[rows])[:, :3], c)
                    want = oracle.scatter(stand_in[wi][rows], c, int(geom.grid_size[0]), int(geom.grid_size[1]))[0]
                    assert np.array_equal(_np(r.canvas[k]), want), f"frame {b} on GPU {r.device}"


def test_overlong_frame_is_truncated_on_the_device(pf):
    """the kernels read the frame length from the device-side offsets but the grids are sized from the host's
    max_frame_points; a frame longer than that bound is processed up to the bound, deterministically, and the handle stays
    usable (no stale counts, no stale cell ids)."""
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 5000
    geom = cfgmod.derive_grid(cfg)
    long_f = np.ascontiguousarray(synth.make_frame("S3", geom, 9000, 31), np.float32)
    short_f = np.ascontiguousarray(synth.make_frame("S1", geom, 3000, 32, 300), np.float32)
    fe = pf.PillarFrontEnd(cfg, max_batch=2, max_points_per_frame=9000, device=0)
    pts = torch.from_numpy(np.concatenate([long_f, short_f])).cuda()
    offs = torch.tensor([0, 9000, 12000], dtype=torch.int32).cuda()
    pillar = torch.randn((2 * 5000, 64), device="cuda")
    for bound, frames in ((4096, (long_f[:4096], short_f)), (9000, (long_f, short_f)), (2048, (long_f[:2048], short_f[:2048]))):
        canvas = fe.step(pts, offs, pillar_feats=pillar, max_frame_points=bound)
        counts = _np(fe.voxel_num)
        for f, fr in enumerate(frames):
            v, c, n = oracle.generate(geom, fr, 5000, 15)
            assert counts[f] == len(n), f"bound {bound} frame {f}"
            gv, gc, gn, _ = fe.frame_outputs(f, counts)
            util.assert_vox_equal((_np(gv), _np(gc)[:, :3], _np(gn)), (v, c, n), f"bound {bound} frame {f}")
            assert np.array_equal(_np(canvas[f]), oracle.scatter(_np(pillar[f * 5000: f * 5000 + len(n)]), c, fe.nx, fe.ny)[0])
    fe.close()


def test_pdl_switch_is_bit_identical(pf):
    """Programmatic dependent launch of the chain (pf_debug_set_pdl, off by default): same outputs eager and in a captured
    graph, frames of different lengths so that whole CTAs of the dependent grids have nothing to do."""
    import importlib
    L = importlib.import_module("3d_object_detection_b200._lib").load()
    cfg = cfgmod.builtin_config("20cm")
    cfg["max_voxels"] = 3000
    geom = cfgmod.derive_grid(cfg)
    frames = [synth.make_frame(k, geom, n, 170 + i, 9)[:n] for i, (k, n) in enumerate([("S1", 40000), ("S3", 700), ("S2", 21000)])]
    frames.append(np.zeros((0, 4), np.float32))                      # an empty frame: every CTA of it exits at once
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = torch.tensor(np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.int32)).cuda()
    pillar = torch.randn((4 * 3000, 64), device="cuda")
    got = {}
    try:
        for on in (0, 1):
            assert L.pf_debug_set_pdl(on) == on
            fe = pf.PillarFrontEnd(cfg, max_batch=4, max_points_per_frame=40000, device=0)
            canvas = fe.step(pts, offs, pillar_feats=pillar)
            eager = [t.clone() for t in (fe.voxels, fe.coors, fe.num_points, fe.voxel_num, fe.features, canvas)]
            fe.capture(pts, offs, pillar)
            for t in (fe.voxels, fe.features, fe.canvas):
                t.fill_(7.0)
            canvas = fe.replay()
            graph = [t.clone() for t in (fe.voxels, fe.coors, fe.num_points, fe.voxel_num, fe.features, canvas)]
            counts = _np(fe.voxel_num)
            mv = cfg["max_voxels"]
            for a, b in zip(eager, graph):
                if a.shape[0] == 4 * mv:         # slab layout: only the first voxel_num[f] rows of a frame are defined
                    for f in range(4):
                        sl = slice(f * mv, f * mv + int(counts[f]))
                        assert torch.equal(a[sl].view(torch.int32), b[sl].view(torch.int32))
                else:
                    assert torch.equal(a.view(torch.int32), b.view(torch.int32))
            got[on] = (graph, counts)
            fe.close()
    finally:
        assert L.pf_debug_set_pdl(-1) == 2          # "auto" is a state of its own ...
        assert L.pf_debug_set_pdl(0) == 0           # ... and off is the default
    assert got[0][1].tolist() == got[1][1].tolist() and got[0][1][3] == 0
    mv = cfg["max_voxels"]
    for a, b in zip(got[0][0], got[1][0]):
        if a.shape[0] == 4 * mv:
            for f in range(4):
                sl = slice(f * mv, f * mv + int(got[0][1][f]))
                assert torch.equal(a[sl].view(torch.int32), b[sl].view(torch.int32))
        else:
            assert torch.equal(a.view(torch.int32), b.view(torch.int32))
