"""CPU tests of the post-processing KERNEL LOGIC (csrc/pf_post_kernels.cuh): the same kernel bodies, compiled with g++
against tests/emul/cuda_emul.h and run on host threads, against the oracle.  Candidate lists, radix select, bitonic sort,
gather + decode and the ordered compaction are integer / index work and must agree exactly; the decoded numbers go
through glibc's expf / sinf / cosf here (CUDA's on the device), so they are held to the tolerance of the GPU tests."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.emul import build as emul_build
from tests.util import synth

LIMIT = [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0]


@pytest.fixture(scope="module")
def emul():
    L = C.CDLL(str(emul_build.build()))
    L.emul_post_front.restype = C.c_int
    L.emul_post_finish.restype = C.c_int
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def run_emul(L, cls, box, dirp, mask, anchors, ranges, limit, thr, pre_max, post_max, iou_thr, force_scalar=0):
    """score + select + decode on the emulated kernels, the oracle's NMS in place of pf_nms (whose kernels use warp
    shuffles and are covered on the GPU), then the emulated finish kernel."""
    B, A = cls.shape[0], cls.shape[1]
    nc = len(ranges)
    S = B * nc
    rng = np.ascontiguousarray(np.asarray(ranges, np.int32).reshape(-1))
    rows = S * pre_max
    cand = np.zeros(max(B * A, 1), np.uint64)
    cand_count = np.full(S, -7, np.int32)
    sel_key = np.zeros(rows, np.uint64)
    sel_num = np.full(S, -7, np.int32)
    seg_off = np.full(S + 1, -7, np.int32)
    sel_anchor = np.full(rows, -7, np.int32)
    sel_score = np.full(rows, np.nan, np.float32)
    sel_box = np.full((rows, 7), np.nan, np.float32)
    sel_su = np.full((rows, 4), np.nan, np.float32)
    sel_dir = np.full(rows, -7, np.int32)
    sel_rank = np.full(rows, np.nan, np.float32)
    nms_boxes = np.full((rows, 4), np.nan, np.float32)
    nms_order = np.full(rows, -7, np.int32)
    m8 = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    L.emul_post_front(C.c_int(nc), _p(rng), C.c_int(cls.shape[2]), C.c_int(pre_max), C.c_float(thr), _p(cls), _p(box), _p(dirp),
                      _p(m8) if m8 is not None else None, _p(anchors), C.c_int(B), C.c_longlong(A), _p(cand), _p(cand_count),
                      _p(sel_key), _p(sel_num), _p(seg_off), _p(sel_anchor), _p(sel_score), _p(sel_box), _p(sel_su),
                      _p(sel_dir), _p(sel_rank), _p(nms_boxes), _p(nms_order), C.c_int(force_scalar))
    for s in range(S):          # what pf_nms's sort would have produced: the boxes in order, the identity permutation
        lo, n = seg_off[s], sel_num[s]
        assert np.array_equal(nms_order[s * pre_max: s * pre_max + n], np.arange(n))
        assert np.array_equal(nms_boxes[s * pre_max: s * pre_max + n].view(np.uint32), sel_su[lo:lo + n].view(np.uint32))
    keep = np.full((S, post_max), -1, np.int32)
    keep_num = np.zeros(S, np.int32)
    for s in range(S):
        lo, hi = seg_off[s], seg_off[s + 1]
        if hi > lo:
            k = oracle.nms(sel_su[lo:hi], sel_rank[lo:hi], pre_max, post_max, iou_thr)
            if k is not None:
                keep[s, :len(k)] = k
                keep_num[s] = len(k)
    det_box = np.full((B, nc * post_max, 7), np.nan, np.float32)
    det_score = np.full((B, nc * post_max), np.nan, np.float32)
    det_class = np.full((B, nc * post_max), -7, np.int32)
    det_anchor = np.full((B, nc * post_max), -7, np.int32)
    det_num = np.full(B, -7, np.int32)
    lim = np.asarray(limit, np.float64)
    L.emul_post_finish(C.c_int(nc), _p(rng), C.c_int(pre_max), C.c_int(post_max), _p(lim), C.c_int(B), _p(seg_off), _p(keep),
                       _p(keep_num), _p(sel_box), _p(sel_score), _p(sel_dir), _p(sel_anchor), _p(det_box), _p(det_score),
                       _p(det_class), _p(det_anchor), _p(det_num))
    return dict(cand_count=cand_count, sel_num=sel_num, seg_off=seg_off, sel_anchor=sel_anchor, sel_score=sel_score,
                sel_box=sel_box, sel_su=sel_su, sel_dir=sel_dir, keep=keep, keep_num=keep_num, det_box=det_box,
                det_score=det_score, det_class=det_class, det_anchor=det_anchor, det_num=det_num)


def check_against_oracle(got, cls, box, dirp, mask, anchors, ranges, limit, thr, pre_max, post_max, iou_thr):
    B, nc = cls.shape[0], len(ranges)
    for f in range(B):
        detail = []
        wb, ws, wc, wa = oracle.postprocess_frame(cls[f], box[f], dirp[f], None if mask is None else mask[f], anchors, ranges,
                                                  limit, thr, pre_max, post_max, iou_thr, detail=detail)
        for c in range(nc):
            s = f * nc + c
            d = detail[c]
            lo = got["seg_off"][s]
            n = len(d["sel"])
            assert got["sel_num"][s] == n and got["seg_off"][s + 1] - lo == n
            assert np.array_equal(got["sel_anchor"][lo:lo + n], ranges[c][0] + d["sel"]), f"selection differs (frame {f}, class {c})"
            if n == 0:
                continue
            util.assert_post_boxes_close(got["sel_box"][lo:lo + n], d["boxes"], f"decoded boxes ({f}, {c})")
            assert np.all(np.abs(got["sel_su"][lo:lo + n] - d["standup"]) <= util.post_standup_tol(d["boxes"])[:, None])
            assert np.array_equal(got["sel_dir"][lo:lo + n], d["dir"])
            assert np.allclose(got["sel_score"][lo:lo + n], d["scores"], rtol=2e-6, atol=0)
            assert got["keep"][s, :got["keep_num"][s]].tolist() == d["keep"].tolist()
        n = got["det_num"][f]
        assert n == len(wb), (n, len(wb))
        assert np.array_equal(got["det_class"][f, :n], wc) and np.array_equal(got["det_anchor"][f, :n], wa)
        util.assert_post_boxes_close(got["det_box"][f, :n], wb, f"detections of frame {f}", final=True)
        assert np.allclose(got["det_score"][f, :n], ws, rtol=2e-6, atol=0)


@pytest.mark.parametrize("fmap,B,mean_logit,pre_max,post_max,with_mask,seed", [
    ((12, 10), 2, -5.0, 1000, 300, True, 1),     # fewer candidates than pre_max: no radix select
    ((16, 16), 1, -1.0, 100, 30, True, 2),       # radix select: ~1000 candidates per class, keep 100
    ((16, 16), 2, 0.0, 64, 64, False, 3),        # no anchor mask
    ((20, 20), 1, -1.0, 1000, 300, True, 4),     # the vehicle class holds ~1400 candidates: select 1000 of them
    ((6, 5), 1, -9.0, 1000, 300, True, 5),       # hardly any candidate
    ((24, 24), 1, 1.0, 2000, 700, True, 6),      # > 1024 kept boxes in the frame: the finish kernel walks two chunks
])
def test_emulated_kernels_match_oracle(emul, fmap, B, mean_logit, pre_max, post_max, with_mask, seed):
    cfg = util.cfgmod.builtin_config("20cm")
    geom = util.cfgmod.derive_grid(cfg)
    anchors, ranges = synth.anchors_3d(geom, fmap)
    cls, box, dirp, mask = synth.post_preds(B, len(anchors), seed, mean_logit=mean_logit)
    limit = LIMIT if seed != 2 else [-10.0, -12.0, 0.5, 5.0, 2.2, 2.0]
    for force_scalar in (0, 1):     # the vectorised and the generic score kernel
        got = run_emul(emul, cls, box, dirp, mask if with_mask else None, anchors, ranges, limit, 0.05, pre_max, post_max, 0.1,
                       force_scalar)
        check_against_oracle(got, cls, box, dirp, mask if with_mask else None, anchors, ranges, limit, 0.05, pre_max, post_max, 0.1)
        if seed == 6:
            assert got["det_num"][0] > 1024


def test_emulated_select_with_ties_and_specials(emul):
    """Equal logits (ties broken by the anchor index), a NaN logit, +-inf, -0.0 / +0.0, two class channels."""
    cfg = util.cfgmod.builtin_config("20cm")
    geom = util.cfgmod.derive_grid(cfg)
    anchors, ranges = synth.anchors_3d(geom, (8, 8))
    A = len(anchors)
    rng = np.random.default_rng(9)
    cls = rng.choice(np.array([-4.0, -2.0, -1.0, 0.5, 3.0], np.float32), size=(1, A, 2)).astype(np.float32)
    cls[0, 5, 0] = np.nan
    cls[0, 6, 1] = np.nan
    cls[0, 7] = np.inf
    cls[0, 8] = -np.inf
    cls[0, 9] = (-0.0, -5.0)
    cls[0, 10] = (0.0, -5.0)
    _, box, dirp, mask = synth.post_preds(1, A, 10)
    mask[0, :16] = True
    dirp[0, :50, 1] = dirp[0, :50, 0]             # direction ties -> label 0
    got = run_emul(emul, cls, box, dirp, mask, anchors, ranges, LIMIT, 0.05, 40, 10, 0.1)
    check_against_oracle(got, cls, box, dirp, mask, anchors, ranges, LIMIT, 0.05, 40, 10, 0.1)
    first = got["sel_anchor"][:got["sel_num"][0]].tolist()
    assert first[0] == 7 and 5 not in first and 6 not in first and 8 not in first      # +inf first; NaN and -inf never


def test_emulated_score_mask_around_the_threshold(emul):
    """Logits within 0.01 of the threshold's logit: inside +-2e-3 the kernel evaluates the sigmoid, outside it decides by
    comparison; both must give the oracle's score mask (logits within 2e-7 of a tie in the score are excluded by the
    generator)."""
    cfg = util.cfgmod.builtin_config("20cm")
    geom = util.cfgmod.derive_grid(cfg)
    anchors, ranges = synth.anchors_3d(geom, (12, 12))
    cls, box, dirp, mask = synth.post_preds(1, len(anchors), 11)
    synth.post_threshold_band(cls, 0.05, 0.01, seed=12)
    for force_scalar in (0, 1):
        got = run_emul(emul, cls, box, dirp, mask, anchors, ranges, LIMIT, 0.05, 1000, 300, 0.1, force_scalar)
        check_against_oracle(got, cls, box, dirp, mask, anchors, ranges, LIMIT, 0.05, 1000, 300, 0.1)
    assert 0.3 < got["cand_count"].sum() / mask.sum() < 0.7


def _select(L, lists, pre_max):
    A = max(max((len(x) for x in lists), default=1), 1)
    S = len(lists)
    cand = np.zeros(S * A, np.uint64)
    cnt = np.zeros(S, np.int32)
    for s_, x in enumerate(lists):
        cand[s_ * A: s_ * A + len(x)] = x
        cnt[s_] = len(x)
    sel_key = np.zeros(S * pre_max, np.uint64)
    sel_num = np.full(S, -7, np.int32)
    L.emul_post_select(C.c_int(pre_max), C.c_int(S), C.c_longlong(A), _p(cand), _p(cnt), _p(sel_key), _p(sel_num))
    return [sel_key[s_ * pre_max: s_ * pre_max + sel_num[s_]] for s_ in range(S)]


def test_emulated_radix_select_on_adversarial_keys(emul):
    """k_post_select alone against numpy's sort: keys that differ only in their lowest bits, only in their highest, a
    boundary that falls inside a run of equal upper words (the search has to go through all eight digits), M = pre_max,
    M = pre_max + 1, a single candidate, none, and random 64-bit keys.  Keys are unique, as the kernel's are."""
    emul.emul_post_select.restype = C.c_int
    rng = np.random.default_rng(3)
    hi = np.uint64(0xBF8ABCDE) << np.uint64(32)
    lists = [
        hi | rng.permutation(5000).astype(np.uint64),                                         # one upper word, 5000 lower words
        (rng.permutation(3000).astype(np.uint64) << np.uint64(40)) | np.uint64(7),            # spread over the top digits
        np.concatenate([hi | np.arange(1, 700, dtype=np.uint64), (hi + (np.uint64(1) << np.uint64(32))) | np.arange(1, 50, dtype=np.uint64)]),
        rng.permutation(300).astype(np.uint64) + np.uint64(1),                                 # M = pre_max
        rng.permutation(301).astype(np.uint64) + np.uint64(1),                                 # M = pre_max + 1
        np.array([12345], np.uint64),
        np.zeros(0, np.uint64),
        np.unique(rng.integers(1, 2 ** 63, 20000, dtype=np.uint64)),
        (np.uint64(0xFFFFFFFF) << np.uint64(32)) | rng.permutation(2048).astype(np.uint64),   # all-ones upper word
    ]
    lists[7] = rng.permutation(lists[7])
    for pre_max in (300, 1, 1000):
        got = _select(emul, lists, pre_max)
        for x, g in zip(lists, got):
            want = np.sort(x)[::-1][:pre_max]
            assert np.array_equal(g, want), (pre_max, len(x))


def test_emulated_kernels_fuzz(emul):
    """Seeded random configurations (feature map, batch, candidate density, pre / post max, score threshold incl. the two
    ends where the logit pre-filter is off, 1-3 class channels, reordered class ranges with unassigned anchors between
    them, quantised logits with many exact ties): both score kernels, every stage against the oracle."""
    cfg = util.cfgmod.builtin_config("20cm")
    geom = util.cfgmod.derive_grid(cfg)
    rng = np.random.default_rng(20261105)
    for _ in range(8):
        fmap = (int(rng.integers(2, 12)), int(rng.integers(2, 12)))
        B = int(rng.integers(1, 3))
        mean = float(rng.choice([-8.0, -5.0, -3.0, -1.0, 1.0, 4.0]))
        pre_max = int(rng.choice([1, 2, 7, 33, 64, 100, 257, 1000, 1500]))
        post_max = int(rng.choice([1, 3, 10, 64, 300, 1000]))
        thr = float(rng.choice([0.05, 0.3, 0.5, 0.9, 1e-7, 0.9995]))
        nch = int(rng.choice([1, 1, 2, 3]))
        limit = LIMIT if rng.random() < 0.5 else [-10.0, -12.0, 0.5, 5.0, 2.2, 2.0]
        anchors, ranges = synth.anchors_3d(geom, fmap)
        if rng.random() < 0.3:
            ranges = [ranges[2], [ranges[0][0], ranges[0][1] - 3], ranges[1]]
        seed = int(rng.integers(1 << 30))
        cls, box, dirp, mask = synth.post_preds(B, len(anchors), seed, mean_logit=mean)
        if nch > 1:
            cls = np.concatenate([cls] + [synth.post_preds(B, len(anchors), seed + k, mean_logit=mean)[0] for k in range(1, nch)], axis=2)
        if rng.random() < 0.3:
            cls = np.round(cls * 2) / 2
        with np.errstate(over="ignore"):      # keep the scores off the threshold: glibc's and numpy's exp differ by an ulp
            sc = 1.0 / (1.0 + np.exp(-cls.max(axis=2).astype(np.float64)))
        cls[np.abs(sc - np.float64(np.float32(thr))) < 1e-6 * max(thr, 1e-7) + 1e-9] += 0.25
        m = mask if rng.random() < 0.7 else None
        cls = np.ascontiguousarray(cls)
        for force_scalar in (0, 1):
            got = run_emul(emul, cls, box, dirp, m, anchors, ranges, limit, thr, pre_max, post_max, 0.1, force_scalar)
            check_against_oracle(got, cls, box, dirp, m, anchors, ranges, limit, thr, pre_max, post_max, 0.1)


@pytest.mark.parametrize("name", ["small", "topk", "range", "empty_class"])
def test_emulated_kernels_match_reference_golden(emul, name):
    """The kernel bodies (host emulation) against the annotations of the reference's own Inference.infer_gpu."""
    g = util.load_post(name)
    ranges = g["class_ranges"].tolist()
    pre_max, post_max = int(g["pre_max"]), int(g["post_max"])
    got = run_emul(emul, g["cls_preds"], g["box_preds"], g["dir_preds"], g["anchors_mask"], g["anchors"], ranges,
                   g["center_limit"].tolist(), 0.05, pre_max, post_max, 0.1)
    n = int(got["det_num"][0])
    assert [str(g["detect_class"][k]) for k in got["det_class"][0, :n]] == [str(x) for x in g["name"]]
    want = np.concatenate([g["location"], g["dimensions"], g["rotation_y"].reshape(-1, 1)], 1).astype(np.float32)
    util.assert_post_boxes_close(got["det_box"][0, :n], want, name)
    assert np.allclose(got["det_score"][0, :n], g["score"], rtol=2e-6, atol=0)
