"""GPU parity tests for the standup-box NMS (row N4, pf_nms): bit-exact against vectors from the reference's own
nms_gpu (numba CUDA simulator) and against the oracle."""
import importlib

import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import synth

make_dets = synth.nms_dets

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

CASES = ["n1", "n7", "n64", "n65", "n200", "n300_t5"]


def _nms(pf):
    return importlib.import_module("3d_object_detection_b200.nms")


@pytest.mark.parametrize("name", CASES)
def test_matches_reference_golden(pf, name):
    z = np.load(util.GOLD / "nms.npz")
    dets, thresh, want = z[f"{name}_dets"], float(z[f"{name}_thresh"]), z[f"{name}_keep"].tolist()
    m = _nms(pf)
    assert m.nms_gpu(dets, thresh) == want
    assert m.nms_gpu(torch.from_numpy(dets).cuda(), thresh) == want
    got = m.nms(dets[:, :4], dets[:, 4], 1000, 5, thresh)
    assert got.dtype == np.int64 and got.tolist() == want[:5]
    got_t = m.nms(torch.from_numpy(dets[:, :4]).cuda(), torch.from_numpy(dets[:, 4]).cuda(), 1000, 300, thresh)
    assert got_t.is_cuda and got_t.dtype == torch.int64 and got_t.cpu().tolist() == want[:300]


@pytest.mark.parametrize("name", ["t01", "t05", "t03"])
def test_float64_rule_matches_compiled_numba(pf, name):
    """Decisive pairs within a few float32 ulps of the threshold (and one exact tie): the keep list of the reference's
    nms_gpu with iou_device compiled by real numba (oracle/make_golden_nms_f64.py; float32 edge differences, `+ 1` and
    the quotient in float64, float64 `>`).  An all-float32 overlap test would return `keep_if_float32` instead."""
    z = np.load(util.GOLD / "nms_f64.npz")
    dets, thresh = z[f"{name}_dets"], float(z[f"{name}_thresh"])
    want, f32 = z[f"{name}_keep"].tolist(), z[f"{name}_keep_if_float32"].tolist()
    m = _nms(pf)
    got = m.nms_gpu(dets, thresh)
    assert got != f32, "the device evaluates the overlap in float32"
    assert got == want
    assert m.nms_gpu(torch.from_numpy(dets).cuda(), thresh) == want


@pytest.mark.parametrize("n,thresh,seed", [(1000, 0.1, 1), (1000, 0.5, 2), (1023, 0.3, 3), (1025, 0.1, 4), (2500, 0.1, 5),
                                           (4096, 0.7, 6), (129, 0.0, 7), (333, 0.99, 8)])
def test_matches_oracle_at_size(pf, n, thresh, seed):
    """pre_max_size = 1000 is the reference's working point; larger segments use the global-memory walk."""
    dets = make_dets(n, seed)
    want = oracle.nms_gpu(dets, np.float32(thresh))
    got = _nms(pf).nms_gpu(dets, thresh)
    assert got == want, f"first difference at {next(i for i, (a, b) in enumerate(zip(got, want)) if a != b) if len(got) == len(want) else (len(got), len(want))}"
    assert 0 < len(want) <= n


def test_edge_cases(pf):
    m = _nms(pf)
    assert m.nms_gpu(np.zeros((0, 5), np.float32), 0.1) == []
    assert m.nms(np.zeros((0, 4), np.float32), np.zeros((0,), np.float32), 1000, 300, 0.1) is None
    # identical boxes: only the best survives; disjoint boxes: all survive, in score order
    same = np.tile(np.array([[0, 0, 4, 2, 0]], np.float32), (70, 1))
    same[:, 4] = np.linspace(0.1, 0.9, 70)
    assert m.nms_gpu(same, 0.5) == [69]
    far = same.copy()
    far[:, 0] += 100 * np.arange(70)
    far[:, 2] += 100 * np.arange(70)
    assert m.nms_gpu(far, 0.5) == list(range(69, -1, -1))
    # equal scores: higher index first (stable ascending argsort read backwards), as the oracle defines it
    tie = far.copy()
    tie[:, 4] = 0.5
    assert m.nms_gpu(tie, 0.5) == oracle.nms_gpu(tie, np.float32(0.5)) == list(range(69, -1, -1))
    # negative and zero scores order correctly
    neg = far[:5].copy()
    neg[:, 4] = [-1.0, 0.0, 2.0, -0.5, 1.0]
    assert m.nms_gpu(neg, 0.5) == [2, 4, 1, 3, 0]
    with pytest.raises(ValueError):
        m.nms_gpu(np.zeros((5000, 5), np.float32), 0.1)


def test_batched_segments(pf):
    """All (class, frame) segments of a batch in one call: equals the per-segment oracle, -1 padded, no sync."""
    m = _nms(pf)
    sizes = [1000, 0, 37, 640, 1, 999, 64, 128]
    thresh, post = 0.1, 300
    dets = [make_dets(n, 40 + i) if n else np.zeros((0, 5), np.float32) for i, n in enumerate(sizes)]
    cat = np.concatenate(dets)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    boxes = torch.from_numpy(cat[:, :4].copy()).cuda()
    scores = torch.from_numpy(cat[:, 4].copy()).cuda()
    for rep in range(2):
        keep, num = m.nms_device(boxes, scores, torch.from_numpy(offs).cuda(), 1000, thresh, post)
        keep, num = keep.cpu().numpy(), num.cpu().numpy()
        for s, d in enumerate(dets):
            want = oracle.nms_gpu(d, np.float32(thresh))[:post] if len(d) else []
            assert num[s] == len(want), (s, rep)
            assert keep[s, :len(want)].tolist() == want and np.all(keep[s, len(want):] == -1), (s, rep)
    # a segment above the bound is flagged, the others are unaffected
    keep, num = m.nms_device(boxes, scores, torch.from_numpy(offs).cuda(), 700, thresh, post)
    num = num.cpu().numpy()
    assert num[0] == -1 and num[5] == -1 and num[3] == len(oracle.nms_gpu(dets[3], np.float32(thresh))[:post])


def test_negative_threshold_and_touching_boxes(pf):
    """The quick reject of disjoint pairs is only valid for thresholds >= 0; boxes exactly one unit apart have a zero
    side under the reference's `+ 1` rule, boxes less than one unit apart still intersect."""
    m = _nms(pf)
    dets = synth.nms_dets(200, 77)
    for thresh in (-0.5, 0.0, 1e-6):
        assert m.nms_gpu(dets, thresh) == oracle.nms_gpu(dets, np.float32(thresh)), thresh
    a = np.array([[0, 0, 4, 2, 0.9], [5, 0, 9, 2, 0.8], [4.5, 0, 8.5, 2, 0.7], [0, 3, 4, 5, 0.6], [0, 2.25, 4, 4.25, 0.5]],
                 np.float32)
    for thresh in (0.0, 0.01, 0.05, 0.2):
        assert m.nms_gpu(a, thresh) == oracle.nms_gpu(a, np.float32(thresh)), thresh
