"""ORACLE -- TEST INFRASTRUCTURE ONLY.  How representative is the C port as the CPU baseline?  Times the REFERENCE's own
numba VoxelGenerator.generate and torch-CPU PointPillarsScatter against the C port (oracle/pillar_oracle.c) on the
same frames, one thread, in the build container (needs /root/reference).

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.compare_ref_speed
"""
from __future__ import annotations

import importlib
import time

import numpy as np

from oracle import oracle
from oracle.make_golden import ref_modules

cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")


def best(fn, n=7):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3


def main():
    import torch
    torch.set_num_threads(1)
    VoxelGenerator, _, Scatter = ref_modules()
    for gkey in ("20cm", "10cm", "nuscene"):
        cfg = cfgmod.builtin_config(gkey)
        geom = cfgmod.derive_grid(cfg)
        n = cfg["frame_points"]
        vg = VoxelGenerator(dict(cfg))
        for kind in ("S1", "S2"):
            pts = synth.make_frame(kind, geom, n, 42, synth.S1_CELLS.get(gkey, 12000))
            vg.generate(pts)                                                    # JIT warm-up
            t_ref = best(lambda: vg.generate(pts))
            t_port = best(lambda: oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"]))
            v, c, m = oracle.generate(geom, pts, cfg["max_voxels"], cfg["max_num_points"])
            feats = np.random.default_rng(0).standard_normal((len(m), 64)).astype(np.float32)
            sc = Scatter(1, geom.grid_size, 64)
            tf, tc = torch.from_numpy(feats), torch.from_numpy(c)
            t_ref_s = best(lambda: sc(tf, tc), 3)
            t_port_s = best(lambda: oracle.scatter(feats, c, int(geom.grid_size[0]), int(geom.grid_size[1])), 3)
            print(f"{gkey:8s} {kind}  N={len(pts):6d} V={len(m):5d}  generate: reference numba {t_ref:7.2f} ms, C port {t_port:7.2f} ms"
                  f"   scatter: reference torch-CPU(1 thread) {t_ref_s:7.2f} ms, C port {t_port_s:7.2f} ms")


if __name__ == "__main__":
    main()
