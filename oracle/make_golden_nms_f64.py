"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Pins the PRECISION of the NMS overlap test (row N4) to the reference's own
`iou_device` (framework/nms.py:102-113) as REAL numba compiles it -- the half oracle/make_golden_nms.py leaves open, because
numba's CUDA simulator evaluates that function on numpy float32 scalars.

Run in the build container only (needs /root/reference and numba; no GPU needed):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_nms_f64

Two things are produced, both from the reference's source text of `iou_device`, never from a restatement:

1. `tests/golden/nms_iou_ptx_ops.json`: the function compiled by numba's CUDA target (`numba.cuda.compile_ptx`, device
   function, signature `(float32[:], float32[:])` as the reference declares it; needs NVVM, not a GPU).  Stored: numba's
   inferred return type and the set of floating-point PTX instructions.  It shows float32 edge differences (`sub.f32`),
   the `+ 1` and everything after it in float64 (`cvt.f64.f32`, `add.f64`, `mul.f64`, `div.rn.f64`), one contraction
   (`Sa + Sb` = `fma.rn.f64(aw, ah, Sb)`), and in the kernel a float64 comparison with the threshold (`setp.gt.f64`).
2. `tests/golden/nms_f64.npz`: detections whose decisive pairs have an IoU within a few float32 ulps of the threshold,
   chosen so that a float32 evaluation of the same formula and the float64 one DISAGREE about `iou > threshold` for every
   pair, plus pairs whose IoU equals the threshold exactly (strict `>`).  The expected keep list comes from the
   reference's own `nms_gpu` (argsort, `nms_kernel`, `nms_postprocess`, framework/nms.py:6-44, 85-147) with the kernel run
   by the simulator and `iou_device` replaced by the SAME source compiled by real numba for the host (`numba.njit`; the
   type inference that yields float64 is shared by the host and the CUDA targets, item 1 shows the CUDA side).  The
   generator asserts that an all-float32 evaluation gives a different keep list, i.e. that the vector discriminates.
"""
import json
import os
import re
import subprocess
import sys

import numpy as np

from oracle.make_golden import GOLD, REF

NMS_SRC = os.path.join(REF, "framework", "nms.py")


def ref_source(name):
    """Source text of a @cuda.jit function of the reference's framework/nms.py, decorator stripped."""
    src = open(NMS_SRC).read()
    m = re.search(r"^@cuda\.jit\([^\n]*\)\ndef %s\(.*?(?=^\S|\Z)" % name, src, re.S | re.M)
    body = m.group(0)
    return body[body.index("def "):]


def iou_source():
    return ref_source("iou_device")


_FLOAT_OP = r"\b((?:div|add|sub|mul|fma|max|min|cvt|setp)[\w.]*\.f(?:32|64))\b"


def ptx_evidence():
    """Runs WITHOUT the simulator (own process): numba.cuda.compile_ptx of the reference's iou_device, and of its whole
    nms_kernel (the nested device-function compile asks for the current device's compute capability, so that one lookup
    is answered with (9, 0) -- there is no device here)."""
    import types

    import numba
    import numba.cuda.dispatcher as disp
    from numba import cuda, float32, int64, uint64
    ns = {}
    exec(iou_source(), ns)
    ptx, res = cuda.compile_ptx(ns["iou_device"], (float32[:], float32[:]), device=True, cc=(9, 0))
    ops = sorted(set(re.findall(_FLOAT_OP, ptx)))
    disp.get_current_device = lambda: types.SimpleNamespace(compute_capability=(9, 0))
    ns = {"cuda": cuda, "numba": numba}
    exec(iou_source(), ns)
    ns["iou_device"] = cuda.jit(device=True, inline=True)(ns["iou_device"])
    exec(ref_source("nms_kernel"), ns)
    kptx, _ = cuda.compile_ptx(ns["nms_kernel"], (int64, float32, float32[:], uint64[:]), cc=(9, 0))
    kops = re.findall(_FLOAT_OP, kptx)
    return {"function": "framework/nms.py:102-113 iou_device", "signature": "(float32[:], float32[:])",
            "numba_return_type": str(res), "float_ptx_ops": ops,
            "float_ptx_sequence": re.findall(_FLOAT_OP, ptx),
            "kernel": "framework/nms.py:116-147 nms_kernel, signature (int64, float32, float32[:], uint64[:])",
            "kernel_float_ptx_ops": {op: kops.count(op) for op in sorted(set(kops))},
            "numba_version": numba.__version__,
            "how": "numba.cuda.compile_ptx(cc=(9, 0)) on the reference's source text, no GPU"}


def iou32(a, b):
    """The same formula with every intermediate rounded to float32 (what the simulator does)."""
    f = np.float32
    left, right = np.maximum(a[0], b[0]), np.minimum(a[2], b[2])
    top, bottom = np.maximum(a[1], b[1]), np.minimum(a[3], b[3])
    w = np.maximum(f(f(right - left) + f(1)), f(0))
    h = np.maximum(f(f(bottom - top) + f(1)), f(0))
    inter = f(w * h)
    sa = f(f(f(a[2] - a[0]) + f(1)) * f(f(a[3] - a[1]) + f(1)))
    sb = f(f(f(b[2] - b[0]) + f(1)) * f(f(b[3] - b[1]) + f(1)))
    return f(inter / f(f(sa + sb) - inter))


def find_pair(iou64, thresh, origin, rng):
    """One pair (a, b) of float32 boxes near `origin` with (iou64 > t) != (iou32 > t): bisect a shift of b to the crossing,
    then walk the float32 neighbours of the shift."""
    t = np.float64(np.float32(thresh))
    f = np.float32
    while True:
        wh = np.abs(rng.normal([4.5, 2.0], [0.8, 0.4])) + 0.5
        x0, y0 = f(origin[0] + rng.uniform(-3, 3)), f(origin[1] + rng.uniform(-3, 3))
        a = np.array([x0, y0, f(x0 + f(wh[0])), f(y0 + f(wh[1]))], np.float32)
        wb = np.abs(rng.normal([4.5, 2.0], [0.8, 0.4])) + 0.5
        dy = f(rng.uniform(-0.5, 0.5))

        def box_b(dx):
            bx, by = f(a[0] + dx), f(a[1] + dy)
            return np.array([bx, by, f(bx + f(wb[0])), f(by + f(wb[1]))], np.float32)

        lo, hi = f(0.0), f(wh[0] + 2.0)                 # the IoU falls as b slides out
        if not (iou64(a, box_b(lo)) > t > iou64(a, box_b(hi))):
            continue
        for _ in range(40):
            mid = f((lo + hi) / 2)
            if iou64(a, box_b(mid)) > t:
                lo = mid
            else:
                hi = mid
        dx = lo
        for _ in range(64):
            dx = np.nextafter(dx, f(-100))
        for _ in range(128):
            b = box_b(dx)
            d64, d32 = bool(iou64(a, b) > t), bool(iou32(a, b) > f(thresh))
            if d64 != d32:
                return a, b, d64
            dx = np.nextafter(dx, f(100))


def build_case(iou64, thresh, n_pairs, seed):
    """dets: decisive pairs on a coarse lattice (no interaction across pairs); scores distinct, either box may lead."""
    rng = np.random.default_rng(seed)
    scores = rng.permutation(2 * n_pairs + 4).astype(np.float64) / (2 * n_pairs + 4) * 0.9 + 0.05
    rows, n_sup64 = [], 0
    for k in range(n_pairs):
        a, b, sup64 = find_pair(iou64, thresh, (64.0 * (k % 8), 64.0 * (k // 8)), rng)
        rows.append([*a, scores[2 * k]])
        rows.append([*b, scores[2 * k + 1]])
        n_sup64 += sup64
    dets = np.asarray(rows, np.float64).astype(np.float32)
    assert np.array_equal(dets[:, :4], np.asarray(rows)[:, :4])        # boxes are float32 values already
    if thresh == 0.5:            # an exact tie: IoU == threshold in float64 and in float32 -> strict `>` keeps both
        far = np.float32(1000.0)
        ties = np.array([[far, far, far + 9, far + 9, scores[-1]], [far, far, far + 9, far + 4, scores[-2]]], np.float32)
        assert iou64(ties[0, :4], ties[1, :4]) == 0.5
        dets = np.concatenate([dets, ties])
    return dets, n_pairs, n_sup64


def keep32(dets, thresh):
    """Greedy NMS with the all-float32 overlap test (the answer a float32 implementation would give)."""
    order = np.argsort(dets[:, 4], kind="stable")[::-1]
    b = dets[order, :4]
    n = len(b)
    removed, keep = np.zeros(n, bool), []
    for i in range(n):
        if removed[i]:
            continue
        keep.append(int(order[i]))
        for j in range(i + 1, n):
            if iou32(b[i], b[j]) > np.float32(thresh):
                removed[j] = True
    return keep


def main():
    if "--ptx" in sys.argv:
        print(json.dumps(ptx_evidence()))
        return
    env = {k: v for k, v in os.environ.items() if k != "NUMBA_ENABLE_CUDASIM"}
    ev = json.loads(subprocess.check_output([sys.executable, "-m", "oracle.make_golden_nms_f64", "--ptx"], env=env,
                                            text=True).strip().splitlines()[-1])
    assert ev["numba_return_type"] == "float64", ev
    assert "div.rn.f64" in ev["float_ptx_ops"] and "add.f64" in ev["float_ptx_ops"] and "sub.f32" in ev["float_ptx_ops"], ev
    assert not any(op.startswith("div") and op.endswith("f32") for op in ev["float_ptx_ops"]), ev
    kops = ev["kernel_float_ptx_ops"]
    assert "setp.gt.f64" in kops and not any(op.startswith("setp") and op.endswith("f32") for op in kops), kops
    with open(GOLD / "nms_iou_ptx_ops.json", "w") as fh:
        json.dump(ev, fh, indent=1)
    print("ptx:", ev["numba_return_type"], ev["float_ptx_ops"], "kernel:", kops)

    os.environ["NUMBA_ENABLE_CUDASIM"] = "1"      # the kernel runs in the simulator, iou_device as compiled host code
    import numba
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    ns = {}
    exec(iou_source(), ns)
    compiled = numba.njit("(float32[:], float32[:])")(ns["iou_device"])
    assert str(compiled.nopython_signatures[0].return_type) == "float64"

    def iou64(a, b):
        return compiled(np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32))

    import framework.nms as ref_nms

    def as_np(x):                                  # the simulator hands slices of its fake device arrays
        return np.ascontiguousarray(getattr(x, "_ary", x), np.float32)

    # np.float64, not a Python float: numpy would round a weak Python scalar to the float32 threshold's type before the
    # comparison, compiled numba compares in float64 (setp.gt.f64 in the kernel's PTX)
    ref_nms.iou_device = lambda a, b: np.float64(compiled(as_np(a), as_np(b)))
    out = {}
    for name, thresh, n_pairs, seed in [("t01", 0.1, 48, 501), ("t05", 0.5, 40, 502), ("t03", 0.3, 70, 503)]:
        dets, n_disc, n_sup64 = build_case(iou64, thresh, n_pairs, seed)
        assert len(np.unique(dets[:, 4])) == len(dets)
        keep = [int(k) for k in ref_nms.nms_gpu(dets, np.float32(thresh))]
        k32 = keep32(dets, thresh)
        assert keep != k32, "the vector must tell the float64 rule from a float32 one"
        out[f"{name}_dets"], out[f"{name}_thresh"] = dets, np.float32(thresh)
        out[f"{name}_keep"], out[f"{name}_keep_if_float32"] = np.asarray(keep, np.int64), np.asarray(k32, np.int64)
        print(f"nms_f64 {name}: {len(dets)} boxes, {n_disc} decisive pairs ({n_sup64} suppress under float64), kept {len(keep)} "
              f"(a float32 rule would keep {len(k32)}; lists differ in {len(set(keep) ^ set(k32))} boxes)")
    np.savez_compressed(GOLD / "nms_f64.npz", **out)


if __name__ == "__main__":
    main()
