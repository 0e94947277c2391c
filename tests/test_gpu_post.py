"""GPU parity tests for the detection post-processing (row N4 around the NMS, pf_postprocess): against vectors made by
the reference's own Inference.infer_gpu (oracle/make_golden_post.py) and, stage by stage, against the oracle.

Bars: the selection (which anchors, in which order), the direction labels, the NMS keep lists, the class / anchor of every
detection and the x, y, rotation_y columns are exact; l, w, h go through exp and z subtracts h / 2 (2e-6 relative to the
larger operand), the standup boxes through sin / cos (2e-6 of the largest operand); scores 2e-6 relative."""
import importlib

import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.util import synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

GOLDEN = ["small", "topk", "range", "empty_class"]
LIMIT = [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0]


def _inf(pf):
    return importlib.import_module("3d_object_detection_b200.inference")


class _Assigner:
    def __init__(self, anchors, ranges, names):
        self.anchors = anchors
        self.class_masks = {str(n): [int(lo), int(hi)] for n, (lo, hi) in zip(names, ranges)}


_golden_inputs = util.load_post


@pytest.mark.parametrize("name", GOLDEN + ["full20"])
def test_matches_reference_golden(pf, name):
    """Inference.infer_gpu of this package against the annotations the reference's Inference.infer_gpu produced."""
    g = _golden_inputs(name)
    cfg = dict(device=0, center_limit=[float(x) for x in g["center_limit"]], detect_class=[str(x) for x in g["detect_class"]],
               batch_size=1)
    inf = _inf(pf).Inference(cfg, _Assigner(g["anchors"], g["class_ranges"], g["detect_class"]))
    inf._nms_pre_max_size = int(g["pre_max"])
    inf._nms_post_max_size = int(g["post_max"])
    example = {"anchors_mask": torch.from_numpy(g["anchors_mask"]).cuda()}
    preds = {"cls_preds": torch.from_numpy(g["cls_preds"]).cuda(), "box_preds": torch.from_numpy(g["box_preds"]).cuda(),
             "dir_preds": torch.from_numpy(g["dir_preds"]).cuda()}
    annos = inf.infer_gpu(example, preds)
    assert len(annos) == 1
    a = annos[0]
    assert a["name"].tolist() == [str(x) for x in g["name"]], "detections (class sequence) differ"
    got = np.concatenate([a["location"].reshape(-1, 3), a["dimensions"].reshape(-1, 3), a["rotation_y"].reshape(-1, 1)], 1)
    want = np.concatenate([g["location"], g["dimensions"], g["rotation_y"].reshape(-1, 1)], 1)
    util.assert_post_boxes_close(got.astype(np.float32), want.astype(np.float32), name)
    assert np.allclose(a["score"], g["score"], rtol=2e-6, atol=0)
    assert set(a) == {"name", "truncated", "occluded", "alpha", "bbox", "dimensions", "location", "rotation_y", "score"}


def stagewise_check(post, B, cls, box, dirp, mask, anchors, ranges, limit, thr, pre_max, post_max, iou_thr, outs):
    """Every stage of the device pipeline against the oracle; the stages behind a floating-point one are fed the device's
    own numbers, so that each comparison is exact wherever the arithmetic is."""
    nc = len(ranges)
    S = B * nc
    reg = {k: post.region(k, B).cpu().numpy() for k in ("seg_offsets", "sel_num", "cand_count", "sel_anchor", "sel_score",
                                                         "sel_box", "sel_standup", "sel_dir", "sel_rank", "keep", "keep_num")}
    keep = reg["keep"][:S * post_max].reshape(S, post_max)
    det_boxes, det_scores, det_class, det_anchor, det_num = (t.cpu().numpy() for t in outs)
    n_sel = n_keep = n_det = 0
    for f in range(B):
        rows = []
        for c, (lo, hi) in enumerate(ranges):
            s = f * nc + c
            sel, scores = oracle.post_select(cls[f, lo:hi], None if mask is None else mask[f, lo:hi], thr, pre_max)
            o, n = int(reg["seg_offsets"][s]), len(sel)
            assert reg["sel_num"][s] == n and reg["seg_offsets"][s + 1] - o == n, f"segment ({f}, {c}): {reg['sel_num'][s]} != {n}"
            assert np.array_equal(reg["sel_anchor"][o:o + n], lo + sel), f"selection differs in segment ({f}, {c})"
            n_sel += n
            if n == 0:
                assert reg["keep_num"][s] == 0
                continue
            want_boxes = oracle.box_decode(box[f, lo:hi][sel], anchors[lo:hi][sel])
            g_boxes, g_su = reg["sel_box"][o:o + n], reg["sel_standup"][o:o + n]
            util.assert_post_boxes_close(g_boxes, want_boxes, f"decoded boxes ({f}, {c})")
            want_su = oracle.standup_boxes(want_boxes[:, [0, 1, 3, 4, 6]])
            assert np.all(np.abs(g_su - want_su.astype(np.float64)) <= util.post_standup_tol(want_boxes)[:, None]), \
                f"standup boxes ({f}, {c})"
            d = dirp[f, lo:hi][sel]
            assert np.array_equal(reg["sel_dir"][o:o + n], (d[:, 1] > d[:, 0]).astype(np.int32))
            assert np.allclose(reg["sel_score"][o:o + n], scores, rtol=2e-6, atol=0)
            assert np.array_equal(reg["sel_rank"][o:o + n], np.arange(pre_max, pre_max - n, -1, dtype=np.float32))
            # NMS on the device's own standup boxes: exact
            want_keep = oracle.nms(g_su, reg["sel_rank"][o:o + n], pre_max, post_max, iou_thr)
            want_keep = np.zeros(0, np.int64) if want_keep is None else want_keep
            k = int(reg["keep_num"][s])
            assert keep[s, :k].tolist() == want_keep.tolist(), f"NMS keep list differs in segment ({f}, {c})"
            assert np.all(keep[s, k:] == -1)
            n_keep += k
            if k:
                fb, fs, rm = oracle.post_finish(g_boxes, reg["sel_score"][o:o + n], reg["sel_dir"][o:o + n], want_keep, limit)
                rows.append((fb, fs, np.full(len(fb), c, np.int32), (lo + sel[want_keep][rm]).astype(np.int32)))
        n = int(det_num[f])
        want_n = sum(len(r[0]) for r in rows)
        assert n == want_n, f"frame {f}: {n} detections, oracle {want_n}"
        if n:
            wb = np.concatenate([r[0] for r in rows])
            assert np.array_equal(det_boxes[f, :n].view(np.uint32), wb.view(np.uint32)), f"frame {f}: detections differ"
            assert np.array_equal(det_scores[f, :n].view(np.uint32), np.concatenate([r[1] for r in rows]).view(np.uint32))
            assert np.array_equal(det_class[f, :n], np.concatenate([r[2] for r in rows]))
            assert np.array_equal(det_anchor[f, :n], np.concatenate([r[3] for r in rows]))
        n_det += n
    return n_sel, n_keep, n_det


def run_case(pf, fmap, B, seed, mean_logit, pre_max, post_max, with_mask=True, limit=LIMIT, max_frames=None, edit=None):
    cfg = util.cfgmod.builtin_config("20cm")
    anchors, ranges = synth.anchors_3d(util.cfgmod.derive_grid(cfg), fmap)
    cls, box, dirp, mask = synth.post_preds(B, len(anchors), seed, mean_logit=mean_logit)
    if edit is not None:
        edit(cls, box, dirp, mask)
    if not with_mask:
        mask = None
    post = _inf(pf).PostProcessor(anchors, ranges, limit, max_frames=max_frames or B, n_cls_channels=cls.shape[2],
                                  pre_max_size=pre_max, post_max_size=post_max, iou_threshold=0.1, score_threshold=0.05)
    outs = post(torch.from_numpy(cls).cuda(), torch.from_numpy(box).cuda(), torch.from_numpy(dirp).cuda(),
                None if mask is None else torch.from_numpy(mask).cuda())
    torch.cuda.synchronize()
    counts = stagewise_check(post, B, cls, box, dirp, mask, anchors, ranges, limit, 0.05, pre_max, post_max, 0.1, outs)
    return post, counts, (cls, box, dirp, mask, anchors, ranges)


@pytest.mark.parametrize("fmap,B,seed,mean_logit,pre_max,post_max,with_mask", [
    ((20, 16), 1, 31, -5.0, 1000, 300, True),       # fewer candidates than pre_max everywhere
    ((40, 40), 3, 32, -3.0, 1000, 300, True),       # the vehicle class exceeds pre_max: radix select
    ((40, 40), 2, 33, 2.0, 256, 100, False),        # nearly every anchor is a candidate, no anchor mask
    ((64, 64), 4, 34, -6.0, 1000, 300, True),
    ((8, 8), 2, 35, -12.0, 1000, 300, True),        # almost no candidates; empty segments
    ((100, 100), 1, 36, 0.0, 4096, 1024, True),     # the largest sizes the ABI admits
])
def test_stagewise_against_oracle(pf, fmap, B, seed, mean_logit, pre_max, post_max, with_mask):
    _, (n_sel, n_keep, n_det), _ = run_case(pf, fmap, B, seed, mean_logit, pre_max, post_max, with_mask)
    if mean_logit > -10:
        assert n_sel > 0 and n_keep > 0 and n_det > 0


def test_range_mask_and_reuse(pf):
    """A tight center_limit drops detections (the reference's any(xyz > low) & any(lwh < high) rule); a second call on
    the same PostProcessor with a smaller batch reuses scratch and outputs."""
    limit = [-10.0, -12.0, 0.5, 5.0, 2.2, 2.0]
    post, (n_sel, n_keep, n_det), (cls, box, dirp, mask, anchors, ranges) = run_case(pf, (32, 32), 3, 41, -4.0, 500, 200, True, limit)
    assert 0 < n_det < n_keep
    outs = post(torch.from_numpy(cls[1:2]).cuda(), torch.from_numpy(box[1:2]).cuda(), torch.from_numpy(dirp[1:2]).cuda(),
                torch.from_numpy(mask[1:2]).cuda())
    torch.cuda.synchronize()
    stagewise_check(post, 1, cls[1:2], box[1:2], dirp[1:2], mask[1:2], anchors, ranges, limit, 0.05, 500, 200, 0.1, outs)


def test_ties_and_special_values(pf):
    """Equal logits are ordered by the anchor index; NaN logits and -inf never pass; +inf comes first; two class
    channels (max over the channels); direction ties give label 0."""
    def edit(cls, box, dirp, mask):
        rng = np.random.default_rng(5)
        cls[...] = rng.choice(np.array([-4.0, -2.0, -1.0, 0.5, 3.0], np.float32), size=cls.shape)
        cls[0, 5] = np.nan
        cls[0, 7] = np.inf
        cls[0, 8] = -np.inf
        cls[0, 9] = -0.0
        cls[0, 10] = 0.0
        mask[0, :16] = True
        dirp[0, :200, 1] = dirp[0, :200, 0]
    post, _, _ = run_case(pf, (16, 16), 2, 51, -3.0, 64, 32, True, edit=edit)
    first = post.region("sel_anchor", 2).cpu().numpy()[:3].tolist()
    assert first[0] == 7


def test_score_mask_around_the_threshold(pf):
    """Logits within 0.01 of the threshold's logit (the kernel evaluates the sigmoid inside +-2e-3 of it and decides by
    comparison outside): same score mask as the oracle, for the vectorised and the generic score kernel."""
    def edit(cls, box, dirp, mask):
        synth.post_threshold_band(cls, 0.05, 0.01, seed=12)
    post, (n_sel, _, _), (cls, _, _, mask, _, _) = run_case(pf, (48, 48), 2, 52, -3.0, 4096, 300, True, edit=edit)
    assert 0.3 < n_sel / mask.sum() < 0.7
    run_case(pf, (47, 47), 1, 53, -3.0, 4096, 300, True, edit=edit)       # 47 * 47 * 9 anchors: A % 4 != 0, generic kernel


def test_two_class_channels(pf):
    cfg = util.cfgmod.builtin_config("20cm")
    anchors, ranges = synth.anchors_3d(util.cfgmod.derive_grid(cfg), (24, 24))
    A = len(anchors)
    cls1, box, dirp, mask = synth.post_preds(2, A, 61, mean_logit=-4.0)
    cls = np.concatenate([cls1, synth.post_preds(2, A, 62, mean_logit=-4.0)[0]], axis=2)
    cls[0, 3, 1] = np.nan
    post = _inf(pf).PostProcessor(anchors, ranges, LIMIT, max_frames=2, n_cls_channels=2, pre_max_size=300, post_max_size=100)
    outs = post(torch.from_numpy(cls).cuda(), torch.from_numpy(box).cuda(), torch.from_numpy(dirp).cuda(),
                torch.from_numpy(mask).cuda())
    torch.cuda.synchronize()
    stagewise_check(post, 2, cls, box, dirp, mask, anchors, ranges, LIMIT, 0.05, 300, 100, 0.1, outs)


def test_full_size_batch(pf):
    """BASELINE.json's 20 cm configuration: 8 frames x 1.44 M anchors (400 x 400 x 9), the anchor mask produced by
    pf_anchor_mask's layout (bool (B, A)), ~40 k candidates in the vehicle class of every frame."""
    _, (n_sel, n_keep, n_det), _ = run_case(pf, (400, 400), 8, 71, -6.0, 1000, 300, True)
    assert n_sel == 8 * 3 * 1000 and n_det == n_keep > 8 * 3 * 50


def test_capturable_and_idempotent(pf):
    """The call allocates nothing and synchronises nowhere: it can be captured into a CUDA graph, and replaying it gives
    the same detections."""
    cfg = util.cfgmod.builtin_config("20cm")
    anchors, ranges = synth.anchors_3d(util.cfgmod.derive_grid(cfg), (50, 50))
    cls, box, dirp, mask = (torch.from_numpy(a).cuda() for a in synth.post_preds(2, len(anchors), 81, mean_logit=-4.0))
    post = _inf(pf).PostProcessor(anchors, ranges, LIMIT, max_frames=2)
    ref = [t.clone() for t in post(cls, box, dirp, mask)]
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        post(cls, box, dirp, mask)
        s.synchronize()
        with torch.cuda.graph(g, stream=s):
            post(cls, box, dirp, mask)
    for t in (post.det_boxes, post.det_scores, post.det_num):
        t.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    n = ref[4].cpu().numpy()
    assert np.array_equal(post.det_num.cpu().numpy(), n) and n.min() > 0
    for f in range(2):
        assert torch.equal(post.det_boxes[f, :n[f]], ref[0][f, :n[f]]) and torch.equal(post.det_scores[f, :n[f]], ref[1][f, :n[f]])


def test_argument_validation(pf):
    m = _inf(pf)
    cfg = util.cfgmod.builtin_config("20cm")
    anchors, ranges = synth.anchors_3d(util.cfgmod.derive_grid(cfg), (8, 8))
    with pytest.raises(Exception):
        m.PostProcessor(anchors, ranges, LIMIT, pre_max_size=5000)
    with pytest.raises(Exception):
        m.PostProcessor(anchors, [[0, 100], [50, 200]], LIMIT)
    post = m.PostProcessor(anchors, ranges, LIMIT)
    cls, box, dirp, mask = (torch.from_numpy(a).cuda() for a in synth.post_preds(1, len(anchors), 1))
    with pytest.raises(ValueError):
        post(cls[:, :-1], box, dirp, mask)
    with pytest.raises(TypeError):
        post(cls.double(), box, dirp, mask)
    with pytest.raises(ValueError):
        post(torch.cat([cls, cls]), torch.cat([box, box]), torch.cat([dirp, dirp]), None)     # batch 2 > max_frames 1
