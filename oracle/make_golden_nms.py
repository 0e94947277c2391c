"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Golden vectors for the standup-box NMS (row N4) from the REFERENCE's own
framework/nms.py, executed under numba's CUDA simulator (NUMBA_ENABLE_CUDASIM=1; the build container has no GPU).

The simulator runs nms_kernel / iou_device (framework/nms.py:102-147) as Python on numpy scalars, where
`float32 - float32 + 1` stays float32; compiled numba promotes the `+ 1` to float64 (checked with the same
expression under @njit: signature (float32, float32) -> float64).  The vectors therefore pin the logic (score
order, the `+ 1` box convention, strict `>` threshold, the 64-wide bit mask, nms_postprocess) while the precision of
the IoU is pinned by numba's typing rule; the generator asserts that no pair of a committed case lies within 1e-5 of
the threshold, so both precisions give the same answer on them.

    NUMBA_ENABLE_CUDASIM=1 PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_nms     (needs /root/reference)
"""
import os
import sys

os.environ["NUMBA_ENABLE_CUDASIM"] = "1"

import numpy as np  # noqa: E402

import importlib  # noqa: E402

from oracle.make_golden import GOLD, REF  # noqa: E402

make_dets = importlib.import_module("3d_object_detection_b200.synth").nms_dets

CASES = [("n1", 1, 0.1, 11), ("n7", 7, 0.1, 12), ("n64", 64, 0.1, 13), ("n65", 65, 0.3, 14), ("n200", 200, 0.1, 15),
         ("n300_t5", 300, 0.5, 16)]


def min_gap(dets, thresh):
    b = dets[:, :4].astype(np.float64)
    w = np.maximum(np.minimum(b[:, None, 2], b[None, :, 2]) - np.maximum(b[:, None, 0], b[None, :, 0]) + 1, 0)
    h = np.maximum(np.minimum(b[:, None, 3], b[None, :, 3]) - np.maximum(b[:, None, 1], b[None, :, 1]) + 1, 0)
    s = (b[:, 2] - b[:, 0] + 1) * (b[:, 3] - b[:, 1] + 1)
    iou = w * h / (s[:, None] + s[None, :] - w * h)
    return float(np.abs(iou - thresh).min())


def main():
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    from framework.nms import nms_gpu
    out = {}
    for name, n, thresh, seed in CASES:
        dets = make_dets(n, seed)
        assert len(np.unique(dets[:, 4])) == n
        gap = min_gap(dets, thresh)
        assert gap > 1e-5, (name, gap)
        keep = np.asarray(nms_gpu(dets, np.float32(thresh)), dtype=np.int64)
        out[f"{name}_dets"] = dets
        out[f"{name}_thresh"] = np.float32(thresh)
        out[f"{name}_keep"] = keep
        print(f"nms {name:8s} n={n:4d} thresh={thresh} kept={len(keep)} closest |iou - thresh| = {gap:.2e}")
    np.savez_compressed(GOLD / "nms.npz", **out)


if __name__ == "__main__":
    main()
