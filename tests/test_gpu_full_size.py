"""GPU parity at the sizes the bench lines are published for (BASELINE.json configs[1]-[4]):

    10 cm  x 32 frames x 300 k points   (4 700+ tiles: more CTAs than are resident, so the tile chain really waits)
    nuScenes x 8 frames x 350 k points  (P = 100, byte counters saturating at 255)
    20 cm  x 64 frames x 120 k points   (the streaming 64-frame batch)

through ``PillarFrontEnd.step`` and its CUDA-graph replay with the FULL ``max_voxels`` of the config, EVERY frame
compared with the oracle: voxels / coordinates / counts / canvas bit-exact, the decoration at the H4 tolerance
(1e-6 * max(1, |xyz|max of the pillar)); and the same for the fused ``pf_voxelize_pfn`` path.  The oracle is pinned
to the reference at these frame sizes by tests/golden/full_hashes.json (tests/test_oracle.py)."""
import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import cfgmod, synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

KINDS = ["S1", "S2", "S3", "S4"]
CASES = [("10cm", 32), ("nuscene", 8), ("20cm", 64), ("20cm", 8), ("16cm", 8)]


def _np(t):
    return t.detach().cpu().numpy()


def _frames(gkey, batch, seed0):
    cfg = cfgmod.builtin_config(gkey)
    geom = cfgmod.derive_grid(cfg)
    n = cfg["frame_points"]
    frames = []
    for b in range(batch):
        kind = KINDS[b % len(KINDS)]
        f = synth.make_frame(kind, geom, n, seed0 + b, synth.S1_CELLS[gkey])
        frames.append(np.ascontiguousarray(f[:n], np.float32))       # S4 splices 42 edge rows in: keep the first n
    return cfg, geom, frames


def _offsets(frames, dev):
    return torch.tensor(np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.int32), device=dev)


def _refs(cfg, geom, frames):
    return [oracle.generate(geom, f, cfg["max_voxels"], cfg["max_num_points"]) for f in frames]


def _check_step(fe, cfg, geom, refs, pillar, canvas, what, features=True, voxels=True):
    mv = cfg["max_voxels"]
    consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)
    counts = _np(fe.voxel_num)
    assert counts.tolist() == [len(r[2]) for r in refs], f"{what}: voxel counts"
    for f, (v, c, n) in enumerate(refs):
        sl = slice(f * mv, f * mv + len(n))
        tag = f"{what} frame {f}"
        co = _np(fe.coors[sl])
        assert np.array_equal(co[:, :3], c), f"{tag}: coors differ"
        assert np.all(co[:, 3] == f), f"{tag}: frame id column"
        assert np.array_equal(_np(fe.num_points[sl]), n), f"{tag}: num_points differ"
        if voxels:
            assert np.array_equal(_np(fe.voxels[sl]).view(np.uint32), v.view(np.uint32)), f"{tag}: voxels differ"
        if features:
            util.assert_features_close(_np(fe.features[sl]), oracle.decorate(v, n, c, *consts), oracle.decorate_tolerance(v), tag)
        want = torch.from_numpy(oracle.scatter(_np(pillar[sl]), c, fe.nx, fe.ny)[0]).to(canvas.device)
        assert torch.equal(canvas[f].view(torch.int32), want.view(torch.int32)), f"{tag}: canvas differs"
        del want


@pytest.mark.parametrize("gkey,batch", CASES)
def test_step_full_size_every_frame(pf, gkey, batch):
    cfg, geom, frames = _frames(gkey, batch, 7000)
    n = cfg["frame_points"]
    refs = _refs(cfg, geom, frames)
    assert any(len(r[2]) == cfg["max_voxels"] for r in refs), "the mix must contain overflowing frames"
    fe = pf.PillarFrontEnd(cfg, max_batch=batch, max_points_per_frame=n, device=0)
    dev = fe.device
    pts = torch.from_numpy(np.concatenate(frames)).to(dev)
    offs = _offsets(frames, dev)
    pillar = torch.randn((batch * cfg["max_voxels"], 64), device=dev)
    canvas = fe.step(pts, offs, pillar_feats=pillar)
    _check_step(fe, cfg, geom, refs, pillar, canvas, f"{gkey}x{batch} eager")
    # graph replay on NEW contents of the static input tensor: frames rotated by one, outputs poisoned first
    fe.capture(pts, offs, pillar)
    rot = frames[1:] + frames[:1]
    pts.copy_(torch.from_numpy(np.concatenate(rot)))
    offs.copy_(_offsets(rot, dev))
    fe.canvas.fill_(7.0)
    fe.features.fill_(7.0)
    fe.voxels.fill_(7.0)
    fe.coors.fill_(-7)
    canvas = fe.replay()
    _check_step(fe, cfg, geom, refs[1:] + refs[:1], pillar, canvas, f"{gkey}x{batch} graph")
    fe.close()


@pytest.mark.parametrize("gkey,batch", [("10cm", 32), ("nuscene", 8), ("20cm", 64)])
def test_fused_pfn_full_size_every_frame(pf, gkey, batch):
    """pf_voxelize_pfn at the published sizes: voxels / coors / counts bit-exact, pillar features against the fp64
    restatement of the PFN evaluated on the oracle's decoration, canvas = bit-exact scatter of the device's rows."""
    cfg, geom, frames = _frames(gkey, batch, 8000)
    n, mv = cfg["frame_points"], cfg["max_voxels"]
    refs = _refs(cfg, geom, frames)
    torch.manual_seed(11)
    net = pf.PointNet(cfg["num_point_features"], cfg["voxel_size"], geom.offset).cuda()
    bn = net.pfn_layers[1]
    with torch.no_grad():
        bn.running_mean.copy_(torch.randn(64) * 0.5)
        bn.running_var.copy_(torch.rand(64) * 2 + 0.1)
        bn.weight.copy_(torch.randn(64))
        bn.bias.copy_(torch.randn(64) * 0.5)
    net.eval()
    fe = pf.PillarFrontEnd(cfg, max_batch=batch, max_points_per_frame=n, device=0)
    dev = fe.device
    pts = torch.from_numpy(np.concatenate(frames)).to(dev)
    offs = _offsets(frames, dev)
    with torch.no_grad():
        canvas = fe.step(pts, offs, pfn=net)
    rows = fe.pillar_feats
    _check_step(fe, cfg, geom, refs, rows, canvas, f"{gkey}x{batch} fused pfn", features=False)
    consts = oracle.decorate_consts(cfg["voxel_size"], geom.offset)
    conv = net.pfn_layers[0]
    wts = [t.detach().cpu().numpy() for t in (conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var)]
    for f in range(0, batch, max(1, batch // 8)):            # the fp64 restatement is slow: 8 frames spread over the batch
        v, c, nn = refs[f]
        feats = oracle.decorate(v, nn, c, *consts)
        finite = np.isfinite(feats).all(axis=(1, 2))
        want = oracle.pfn(feats, *wts, bn.eps)
        got = _np(rows[f * mv: f * mv + len(nn)])
        scale = np.abs(np.where(np.isfinite(feats), feats, 0)).reshape(len(nn), -1).max(axis=1, keepdims=True) + 1.0
        ok = np.abs(got[finite] - want[finite]) <= 1e-4 * scale[finite] + 1e-4 * np.abs(want[finite])
        assert np.all(ok), f"frame {f}: {np.abs(got[finite] - want[finite]).max()}"
    fe.close()
