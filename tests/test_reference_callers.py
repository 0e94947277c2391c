"""The reference's OWN callers driven with the swapped classes, in the build container.

``InferData.get`` (framework/dataset.py:213-229) and ``PointPillars.forward`` (networks/pointpillars8_shared.py:361-382)
are imported from /root/reference and handed this package's ``VoxelGenerator`` / ``PointNet`` / ``PointPillarsScatter``
in place of the reference's.  There is no GPU here, so the three classes are sub-classed with their CUDA calls stubbed by
the CPU oracle (test infrastructure); everything else -- constructors, attributes, config write-back, argument order,
dict keys, dtypes, shapes, ``state_dict`` keys, autograd -- is the product code the reference's callers really touch.
The GPU tests own the numbers; this test owns the call sequence.  Skipped where /root/reference does not exist (the GPU
box)."""
import copy
import importlib
import os
import sys

import numpy as np
import pytest

from oracle import oracle
from tests.util import cfgmod, synth

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference checkout is only in the build container")
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ref():
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import types
    for name in ("matplotlib", "matplotlib.pyplot"):          # framework/dataset.py imports pyplot for a debug plot only
        try:
            importlib.import_module(name)
        except ModuleNotFoundError:
            sys.modules[name] = types.ModuleType(name)
    mods = {name: importlib.import_module(name) for name in
            ("framework.voxel_generator", "framework.dataset", "framework.utils", "networks.pointpillars8_shared")}
    return mods


@pytest.fixture(scope="module")
def swapped():
    """This package's classes with the device calls replaced by the oracle."""
    pkg = importlib.import_module("3d_object_detection_b200")
    vgmod = importlib.import_module("3d_object_detection_b200.voxel_generator")
    ops = importlib.import_module("3d_object_detection_b200.pillar_ops")

    class VG(vgmod.VoxelGenerator):
        def _require_cuda(self):
            pass

        def _generate_frame(self, points, n, n_feat, device_outputs):
            v, c, k = oracle.generate(self.geometry, points, self.max_voxels, self.max_num_points)
            return (v, c, k), len(k)

    class PN(ops.PointNet):
        def fusable(self, voxels=None):
            return False

        def decorate(self, voxels, num, coors):
            return torch.from_numpy(oracle.decorate(voxels.numpy(), num.numpy(), coors.numpy(), self.vx, self.vy,
                                                    self.x_offset, self.y_offset))

    class SC(ops.PointPillarsScatter):
        def forward(self, voxel_features, coords):
            class Fn(torch.autograd.Function):
                @staticmethod
                def forward(ctx, feats):
                    ctx.coords = coords
                    return torch.from_numpy(oracle.scatter(feats.detach().numpy(), coords.numpy(), self.nx, self.ny,
                                                           batch=self.batch_size))

                @staticmethod
                def backward(ctx, g):
                    return torch.from_numpy(oracle.scatter_backward(g.numpy(), ctx.coords.numpy(), batch=self.batch_size))
            return Fn.apply(voxel_features)

    return pkg, VG, PN, SC


def _config():
    return dict(detection_range=[-8.0, -8.0, -2.5, 8.0, 8.0, 8.5], voxel_size=[0.2, 0.2, 11], max_voxels=900,
                max_num_points=15, num_point_features=4, batch_size=1, device="cpu", create_mask_gpu=0)


class _Anchors:
    """Stands in for AnchorAssigner.create_mask (row N2 has its own reference-made vectors): records its arguments."""

    def create_mask(self, coors, grid_size, voxel_size, offset, gpu=False):
        self.seen = (coors, grid_size, voxel_size, offset, gpu)
        return np.ones(50, dtype=bool)


def test_infer_data_get_with_swapped_voxel_generator(ref, swapped):
    pkg, VG, _, _ = swapped
    cfg_ref, cfg_new = _config(), _config()
    vg_ref = ref["framework.voxel_generator"].VoxelGenerator(cfg_ref)
    vg_new = VG(cfg_new, device=0)          # config['device'] is 'cpu' for the reference's tensors; the product refuses it
    for key in ("detection_range", "detection_offset", "detection_range_diff", "grid_size"):       # F8 write-back
        assert np.array_equal(cfg_ref[key], cfg_new[key]) and cfg_ref[key].dtype == cfg_new[key].dtype, key
    for attr in ("voxel_size", "detection_range", "offset", "grid_size", "max_num_points", "max_voxels"):
        a, b = getattr(vg_ref, attr), getattr(vg_new, attr)
        assert np.array_equal(a, b) and np.asarray(a).dtype == np.asarray(b).dtype, attr
    geom = cfgmod.derive_grid(_config())
    pts = synth.make_frame("S1", geom, 6000, 5, 700)
    InferData = ref["framework.dataset"].InferData
    an_ref, an_new = _Anchors(), _Anchors()
    ex_ref = InferData(cfg_ref, vg_ref, an_ref).get(pts)
    d_new = InferData(cfg_new, vg_new, an_new)
    ex_new = d_new.get(pts)
    assert list(ex_ref) == list(ex_new) == ["voxels", "coordinates", "num_points_per_voxel", "anchors_mask"]
    for k in ex_ref:
        assert ex_ref[k].dtype == ex_new[k].dtype and ex_ref[k].shape == ex_new[k].shape, k
        assert torch.equal(ex_ref[k], ex_new[k]), k
    for a, b in zip(an_ref.seen[:4], an_new.seen[:4]):           # what create_mask is handed
        assert np.array_equal(a, b) and a.dtype == b.dtype
    assert d_new.voxel_time > 0 and vg_new.voxel_time > 0
    raw = InferData(cfg_new, vg_new, an_new).get(pts, toTorch=False)
    assert isinstance(raw["voxels"], np.ndarray) and raw["coordinates"].dtype == np.int32


def test_pointpillars_forward_with_swapped_modules(ref, swapped, monkeypatch):
    pkg, VG, PN, SC = swapped
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)     # PointPillars.forward times every stage
    cfg = _config()
    vg = VG(cfg, device=0)
    shared = ref["networks.pointpillars8_shared"]
    torch.manual_seed(0)
    net_ref = shared.PointPillars(cfg).eval()
    net_new = copy.deepcopy(net_ref)
    # the two one-line swaps of INTEGRATION.md §2
    net_new.pillar_point_net = PN(cfg["num_point_features"], cfg["voxel_size"], cfg["detection_offset"])
    net_new.middle_feature_extractor = SC(batch_size=cfg["batch_size"], output_shape=cfg["grid_size"],
                                          num_input_features=net_new.pillar_point_net.out_channels)
    assert list(net_new.state_dict()) == list(net_ref.state_dict())          # checkpoints keep loading, strict
    net_new.load_state_dict(net_ref.state_dict(), strict=True)
    net_new.eval()
    geom = cfgmod.derive_grid(_config())
    pts = synth.make_frame("S1", geom, 5000, 9, 600)
    example = ref["framework.dataset"].InferData(cfg, vg, _Anchors()).get(pts)
    with torch.no_grad():
        out_ref, out_new = net_ref(example), net_new(example)
    assert list(out_ref) == list(out_new) == ["cls_preds", "box_preds", "dir_preds"]
    for k in out_ref:
        assert out_ref[k].shape == out_new[k].shape
        assert torch.allclose(out_ref[k], out_new[k], rtol=1e-4, atol=1e-5), (k, (out_ref[k] - out_new[k]).abs().max())
    assert net_new.pfn_time > 0 and net_new.scatter_time > 0
    # training direction: the scatter is differentiable w.r.t. the pillar features, gradients reach the PFN weights
    net_ref.train(), net_new.train()
    loss_ref = sum(v.square().mean() for v in net_ref(example).values())
    loss_new = sum(v.square().mean() for v in net_new(example).values())
    loss_ref.backward(), loss_new.backward()
    g_ref = net_ref.pillar_point_net.pfn_layers[0].weight.grad
    g_new = net_new.pillar_point_net.pfn_layers[0].weight.grad
    assert g_new is not None and torch.allclose(g_ref, g_new, rtol=1e-3, atol=1e-6)
