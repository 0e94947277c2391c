"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Reference-made vector for row A7 (gradient of the BEV scatter).

Run in the build container only (needs /root/reference and torch-CPU):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_backward

Calls ``.backward()`` through the REFERENCE's own ``PointPillarsScatter.forward``
(networks/pointpillars8_shared.py:76-111; the ``canvas[:, indices] = voxels`` index_put at :100) on the coordinates
of committed small fixtures, with a seeded ``grad_canvas``, and stores the gradient it returns for
``voxel_features`` in tests/golden/scatter_backward.npz.  The grad_canvas itself (up to 63 MB) is not committed: the
test regenerates it from the seed and checks its sha256 first.
"""
from __future__ import annotations

import numpy as np

from oracle.make_golden import GOLD, ref_modules, sha

CASES = [("nuscene", "s1"), ("20cm", "s4_edges"), ("20cm", "dup")]


def grad_canvas(seed, shape):
    return np.random.default_rng(seed).standard_normal(shape, dtype=np.float32)


def main():
    import torch
    _, _, PointPillarsScatter = ref_modules()
    out = {}
    for ci, (gkey, name) in enumerate(CASES):
        if name == "dup":            # the same cell twice: index_put's gradient is a plain gather for BOTH rows
            g = np.load(GOLD / f"small_{gkey}_s1.npz")
            coors = np.ascontiguousarray(np.concatenate([g["coors"][:50], g["coors"][10:30]]))
        else:
            g = np.load(GOLD / f"small_{gkey}_{name}.npz")
            coors = g["coors"]
        grid = g["grid_size"]
        seed = 4200 + ci
        feats = np.random.default_rng(seed + 100).standard_normal((len(coors), 64), dtype=np.float32)
        ft = torch.from_numpy(feats.copy()).requires_grad_()
        sc = PointPillarsScatter(batch_size=1, output_shape=grid, num_input_features=64)
        canvas = sc(ft, torch.from_numpy(coors.copy()))
        gc = grad_canvas(seed, tuple(canvas.shape))
        canvas.backward(torch.from_numpy(gc))
        key = f"{gkey}_{name}"
        out[key + "_coors"] = coors
        out[key + "_grid"] = grid
        out[key + "_seed"] = np.int64(seed)
        out[key + "_grad_canvas_sha"] = np.array(sha(gc))
        out[key + "_grad_feats"] = ft.grad.numpy().copy()
        print(key, "V =", len(coors), "grad_feats", ft.grad.shape, "canvas", tuple(canvas.shape))
    np.savez_compressed(GOLD / "scatter_backward.npz", **out)


if __name__ == "__main__":
    main()
