"""Pillar decoration and BEV scatter with the reference's module interfaces, backed by libpillarfront.so.

Mirrors (reference paths relative to /root/reference):

* :func:`decorate`            <- networks/pointpillars8_shared.py:32-54 (and the 8 inlined copies listed in
                                 SURVEY.md §2a) -- ``[x,y,z,i, x-mean, y-mean, z-mean, x-cx, y-cy]``, padded
                                 rows zero.
* :class:`PointNet`           <- networks/pointpillars8_shared.py:10-60 (same constructor, same
                                 ``pfn_layers`` state-dict keys, same forward signature/result).
* :class:`PointPillarsScatter` <- networks/pointpillars8_shared.py:63-111 (differentiable w.r.t.
                                 ``voxel_features``).
* :class:`Scatter`            <- networks/pointpillars8_export.py:34-51 / pointpillars_trt.py:26-44
                                 (channel-major ``(C, V)`` or ``(1, C, V)`` features, ``run``).
* :class:`Scatter_cuda`       <- networks/pointpillars8_trt.py:139-205 (persistent canvas, ``run`` on device
                                 tensors).

None of the modules registers parameters or buffers beyond the reference's ``pfn_layers`` so checkpoints keep
loading with ``strict=True``.  Batched scatter follows the batch-1 index rule ``x*ny + y`` per
frame with the frame id read from the LAST coords column (the reference's own B>1 branch
indexes ``y*ny + z`` and is never exercised by a shipped config).
"""
from __future__ import annotations

import numpy as np
import torch
from torch import nn

from . import _lib
from .config import GridGeometry


def decorate_constants(voxel_size, offset):
    """(vx, vy, x_offset, y_offset) formed by the same expressions as PointNet.__init__
    (pointpillars8_shared.py:16-19); numpy/Python scalar typing as in SURVEY.md F13, then rounded to fp32."""
    vx = voxel_size[0]
    vy = voxel_size[1]
    x_offset = vx / 2 + offset[0]
    y_offset = vy / 2 + offset[1]
    return (float(np.float32(vx)), float(np.float32(vy)), float(np.float32(x_offset)), float(np.float32(y_offset)))


def _require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise _lib.PillarFrontError(f"{name} must be a CUDA tensor: the pillar front end has no CPU fallback")


def decorate(voxels, num_points, coors, vx, vy, x_offset, y_offset):
    """(V,P,F) voxels, (V,) counts, (V,>=2) coords -> (V,P,F+5) decorated features (fp32, CUDA)."""
    _require_cuda(voxels, "voxels")
    if voxels.dtype != torch.float32:
        raise TypeError(f"voxels must be float32, got {voxels.dtype}")
    if voxels.dim() != 3:
        raise ValueError(f"voxels must be (V, P, F), got {tuple(voxels.shape)}")
    v, p, f = voxels.shape
    voxels = voxels.contiguous()
    num_points = num_points.to(device=voxels.device, dtype=torch.int32).contiguous()
    coors = coors.to(device=voxels.device, dtype=torch.int32).contiguous()
    if num_points.shape[0] != v or coors.shape[0] != v or coors.dim() != 2 or coors.shape[1] < 2:
        raise ValueError("num_points / coors do not match voxels")
    out = torch.empty((v, p, f + 5), dtype=torch.float32, device=voxels.device)
    lib = _lib.load()
    with _lib.on_device(voxels.device):
        stream = _lib.raw_stream(voxels.device.index)
        _lib.check(lib.pf_decorate(voxels.data_ptr(), num_points.data_ptr(), coors.data_ptr(), int(coors.shape[1]),
                                   v, p, f, vx, vy, x_offset, y_offset, out.data_ptr(), stream), "pf_decorate")
    return out


def fold_pfn(pfn_layers):
    """Fold ``pfn_layers`` = Conv1d(9->64, k=1, bias=False) + BatchNorm1d (eval) + ReLU of the reference
    (pointpillars8_shared.py:24-28) into (weight (64, 9), scale (64,), shift (64,)) fp32 contiguous tensors:
    ``relu(bn(conv(x))) == relu(scale * (weight @ x) + shift)`` with scale = gamma / sqrt(var + eps),
    shift = beta - mean * scale."""
    conv, bn = pfn_layers[0], pfn_layers[1]
    if conv.bias is not None or conv.kernel_size != (1,) or conv.out_channels != 64:
        raise ValueError("fold_pfn expects Conv1d(C_in, 64, kernel_size=1, bias=False)")
    with torch.no_grad():
        w = conv.weight.detach().reshape(conv.out_channels, conv.in_channels).to(torch.float32).contiguous()
        gamma = bn.weight.detach().float() if bn.affine else torch.ones_like(bn.running_var)
        beta = bn.bias.detach().float() if bn.affine else torch.zeros_like(bn.running_var)
        scale = (gamma / torch.sqrt(bn.running_var.detach().float() + bn.eps)).contiguous()
        shift = (beta - bn.running_mean.detach().float() * scale).contiguous()
    return w, scale, shift


def pfn_fused(voxels, num_points, coors, vx, vy, x_offset, y_offset, weight, scale, shift):
    """decorate + Conv1d/BN(eval)/ReLU + max over the points in ONE kernel (pf_pfn): (V,P,4) -> (V,64)."""
    _require_cuda(voxels, "voxels")
    if voxels.dtype != torch.float32 or voxels.dim() != 3 or voxels.shape[2] != 4:
        raise ValueError(f"pfn_fused needs float32 (V, P, 4) voxels, got {voxels.dtype} {tuple(voxels.shape)}")
    v, p, f = voxels.shape
    voxels = voxels.contiguous()
    num_points = num_points.to(device=voxels.device, dtype=torch.int32).contiguous()
    coors = coors.to(device=voxels.device, dtype=torch.int32).contiguous()
    out = torch.empty((v, 64), dtype=torch.float32, device=voxels.device)
    lib = _lib.load()
    with _lib.on_device(voxels.device):
        stream = _lib.raw_stream(voxels.device.index)
        _lib.check(lib.pf_pfn(voxels.data_ptr(), num_points.data_ptr(), coors.data_ptr(), int(coors.shape[1]), v, p, f,
                              vx, vy, x_offset, y_offset, weight.data_ptr(), scale.data_ptr(), shift.data_ptr(), 64,
                              out.data_ptr(), stream), "pf_pfn")
    return out


class PointNet(nn.Module):
    """Pillar feature net: CUDA decoration + the reference's learned ``pfn_layers``.  In eval mode with 4-float
    points and no gradient wanted (``torch.no_grad()``, as every inference loop of the reference runs it, or frozen
    parameters) the whole forward (decorate, conv, folded BN, ReLU, max) is one fused CUDA kernel (pf_pfn); in
    training mode (batch statistics) or whenever autograd records the call, the PFN runs in torch on the
    CUDA-decorated features, so eval-mode gradients match the reference's."""

    def __init__(self, num_input_features, voxel_size, offset):
        super().__init__()
        self.name = "PointNet"
        self.vx, self.vy, self.x_offset, self.y_offset = decorate_constants(voxel_size, offset)
        self.out_channels = 64
        self.pfn_layers = nn.Sequential(
            nn.Conv1d(num_input_features + 5, self.out_channels, kernel_size=1, padding=0, bias=False),
            nn.BatchNorm1d(self.out_channels),
            nn.ReLU(True))

    def decorate(self, voxels, num_point_per_voxel, coors):
        return decorate(voxels, num_point_per_voxel, coors, self.vx, self.vy, self.x_offset, self.y_offset)

    def features_to_pillars(self, features):
        """The learned tail of forward() on already decorated (V, P, 9) features, in torch (any mode)."""
        x = self.pfn_layers(features.permute(0, 2, 1).contiguous()).permute(0, 2, 1).contiguous()
        return torch.max(x, dim=1, keepdim=True)[0].squeeze(1)

    def folded(self):
        """(weight, scale, shift) of fold_pfn, cached until a parameter or running statistic changes."""
        conv, bn = self.pfn_layers[0], self.pfn_layers[1]
        ts = [conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var]
        key = tuple((t.data_ptr(), t._version) for t in ts if t is not None)
        if getattr(self, "_fold_key", None) != key:
            self._fold = fold_pfn(self.pfn_layers)
            self._fold_key = key
        return self._fold

    def wants_grad(self, voxels=None) -> bool:
        """True when autograd would record this forward: grad mode on and the voxels or a PFN parameter require grad.
        The fused kernel has no backward, so it may only replace the torch tail when this is False (the reference
        module is differentiable in eval mode too: frozen-BN fine-tuning, saliency maps)."""
        if not torch.is_grad_enabled():
            return False
        if voxels is not None and voxels.requires_grad:
            return True
        return any(p.requires_grad for p in self.pfn_layers.parameters())

    def fusable(self, voxels=None) -> bool:
        return not self.training and not self.wants_grad(voxels)

    def forward(self, voxels, num_point_per_voxel, coors):
        if (self.fusable(voxels) and voxels.is_cuda and voxels.dtype == torch.float32 and voxels.dim() == 3
                and voxels.shape[2] == 4 and self.pfn_layers[0].weight.is_cuda):
            w, sc, sh = self.folded()
            return pfn_fused(voxels, num_point_per_voxel, coors, self.vx, self.vy, self.x_offset, self.y_offset,
                             w, sc, sh).squeeze()
        features = self.decorate(voxels, num_point_per_voxel, coors)
        x = self.pfn_layers(features.permute(0, 2, 1).contiguous()).permute(0, 2, 1).contiguous()
        return torch.max(x, dim=1, keepdim=True)[0].squeeze()


# -------------------------------------------------------------------------------------------------------
_SCATTER_HANDLES = {}


def _scatter_handle(nx: int, ny: int, batch: int, device: torch.device) -> "_lib.Handle":
    """Workspace holding the cell->row map for a stand-alone scatter: one per (device, STREAM, plane) -- a handle is
    not re-entrant (include/pillarfront.h), and two modules running on different streams must not share one map."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, _lib.raw_stream(idx), nx, ny)
    h = _SCATTER_HANDLES.get(key)
    if h is None or h.max_frames < batch:
        if h is not None:
            torch.cuda.synchronize(idx)
            h.close()
        geom = GridGeometry(voxel_size=np.ones(3, np.float32), detection_range=np.zeros(6, np.float32),
                            offset=np.zeros(3, np.float32), range_diff=np.zeros(3, np.float32),
                            grid_size=np.array([nx, ny, 1], np.int32))
        h = _lib.Handle(geom, 1, 1, 4, batch, 1, idx)
        _SCATTER_HANDLES[key] = h
    return h


def _scatter_raw(feats, sv, sc, coors, n_vox, batch, n_ch, nx, ny, out=None):
    dev = feats.device
    h = _scatter_handle(nx, ny, batch, dev)
    canvas = out if out is not None else torch.empty((batch, n_ch, nx, ny), dtype=torch.float32, device=dev)
    with _lib.on_device(dev):
        stream = _lib.raw_stream(dev.index)
        _lib.check(h.lib.pf_scatter(h.ptr, feats.data_ptr(), sv, sc, coors.data_ptr(), int(coors.shape[1]),
                                    n_vox, batch, n_ch, canvas.data_ptr(), stream), "pf_scatter")
    return canvas


class _ScatterFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, voxel_features, coords, batch, nx, ny):
        feats = voxel_features.contiguous()
        coords = coords.to(dtype=torch.int32).contiguous()
        ctx.save_for_backward(coords)
        ctx.dims = (batch, feats.shape[1], nx, ny)
        return _scatter_raw(feats, feats.shape[1], 1, coords, feats.shape[0], batch, feats.shape[1], nx, ny)

    @staticmethod
    def backward(ctx, grad_canvas):
        (coords,) = ctx.saved_tensors
        batch, n_ch, nx, ny = ctx.dims
        g = grad_canvas.contiguous().to(torch.float32)
        n_vox = coords.shape[0]
        out = torch.empty((n_vox, n_ch), dtype=torch.float32, device=g.device)
        lib = _lib.load()
        with _lib.on_device(g.device):
            stream = _lib.raw_stream(g.device.index)
            _lib.check(lib.pf_scatter_backward(g.data_ptr(), coords.data_ptr(), int(coords.shape[1]), n_vox, batch,
                                               n_ch, nx, ny, out.data_ptr(), n_ch, 1, stream), "pf_scatter_backward")
        return out, None, None, None, None


class PointPillarsScatter(nn.Module):
    def __init__(self, batch_size, output_shape, num_input_features=64):
        super().__init__()
        self.name = "PointPillarsScatter"
        self.output_shape = output_shape
        self.nx = int(output_shape[0])
        self.ny = int(output_shape[1])
        self.num_channels = num_input_features
        self.batch_size = batch_size

    def forward(self, voxel_features, coords):
        _require_cuda(voxel_features, "voxel_features")
        if voxel_features.dtype != torch.float32:
            raise TypeError(f"voxel_features must be float32, got {voxel_features.dtype}")
        if voxel_features.dim() != 2 or voxel_features.shape[1] != self.num_channels:
            raise ValueError(f"voxel_features must be (V, {self.num_channels}), got {tuple(voxel_features.shape)}")
        if self.batch_size > 1 and coords.shape[1] < 4:
            raise ValueError("batch_size > 1 needs coords (V, 4) with the frame id in the last column")
        coords = coords.to(device=voxel_features.device)
        if not (torch.is_grad_enabled() and voxel_features.requires_grad):      # inference: no autograd node to build
            feats = voxel_features.contiguous()
            coords = coords.to(dtype=torch.int32).contiguous()
            return _scatter_raw(feats, feats.shape[1], 1, coords, feats.shape[0], int(self.batch_size), feats.shape[1],
                                self.nx, self.ny)
        return _ScatterFn.apply(voxel_features, coords, int(self.batch_size), self.nx, self.ny)


class Scatter:
    """Channel-major variant: ``run((C,V) | (1,C,V) features, (V,3) coords) -> (1, C, nx, ny)``."""

    def __init__(self, output_shape, num_input_features=64):
        self.output_shape = output_shape
        self.nx = int(output_shape[0])
        self.ny = int(output_shape[1])
        self.num_channels = num_input_features

    def run(self, voxel_features, coords):
        _require_cuda(voxel_features, "voxel_features")
        f = voxel_features
        if f.dim() == 3 and f.shape[0] == 1:
            f = f[0]
        if f.dim() != 2 or f.shape[0] != self.num_channels or f.dtype != torch.float32:
            raise ValueError(f"voxel_features must be float32 ({self.num_channels}, V), got {tuple(voxel_features.shape)}")
        f = f.contiguous()
        coords = coords.to(device=f.device, dtype=torch.int32).contiguous()
        n_vox = f.shape[1]
        return _scatter_raw(f, 1, n_vox, coords, n_vox, 1, self.num_channels, self.nx, self.ny)


class Scatter_cuda:
    """Persistent-canvas variant (pointpillars8_trt.py:139-205): ``run`` rewrites and returns the same canvas."""

    def __init__(self, output_shape, num_input_features=64, max_num_pillars=16000, device="cuda:0"):
        self.nx = int(output_shape[0])
        self.ny = int(output_shape[1])
        self.num_channels = num_input_features
        self.max_num_pillars = max_num_pillars
        self.canvas_d = torch.zeros((1, num_input_features, self.nx, self.ny), dtype=torch.float32, device=device)
        self.canvas_size = self.canvas_d.numel()

    def run(self, voxel_d, coors_d):
        f = voxel_d.contiguous()
        coords = coors_d.to(dtype=torch.int32).contiguous()
        n_vox = f.shape[1]
        return _scatter_raw(f, 1, n_vox, coords, n_vox, 1, self.num_channels, self.nx, self.ny, out=self.canvas_d)

    def toHost(self, canvas_d):
        return canvas_d.detach().cpu().numpy().reshape(-1)
