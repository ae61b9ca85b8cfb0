"""GPU parity of the box-head post-processor (softmax, class-wise decode, clip, threshold,
per-class NMS, detections_per_img cut) against the oracle: boxes, scores, labels and proposal
rows bit-exact, in the reference's order (class-major, descending score inside a class)."""
import numpy as np
import pytest
import torch

import oracle
import roipath_b200 as rp
from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _boxlists(props, counts, im_wh):
    out, o = [], 0
    for b, n in enumerate(counts):
        out.append(rp.BoxList(T(props[o:o + n]), (int(im_wh[b, 0]), int(im_wh[b, 1]))))
        o += n
    return out


def _check(res, lg, reg, pr, counts, wh, **kw):
    o = 0
    for b, n in enumerate(counts):
        ref = oracle.box_post_image(lg[o:o + n], reg[o:o + n], pr[o:o + n], wh[b], **kw)
        o += n
        np.testing.assert_array_equal(res[b].get_field("labels").cpu().numpy(), ref[2])
        np.testing.assert_array_equal(res[b].get_field("rows").cpu().numpy(), ref[3])
        np.testing.assert_array_equal(res[b].get_field("scores").cpu().numpy(), ref[1])
        np.testing.assert_array_equal(res[b].bbox.cpu().numpy(), ref[0])


@pytest.mark.parametrize("n_images,per_image,C", [(1, 1, 81), (2, 37, 81), (3, 512, 81), (8, 1000, 81),
                                                  (2, 300, 2), (2, 700, 151)])
@pytest.mark.parametrize("ties", [False, True])
def test_post_processor_matches_oracle(n_images, per_image, C, ties):
    lg, reg, pr, counts, wh = synth.box_head_batch(n_images, per_image, num_classes=C,
                                                   seed=per_image + C, ties=ties)
    post = rp.PostProcessor(0.05, 0.5, 100)
    res = post((T(lg), T(reg)), _boxlists(pr, counts, wh))
    assert len(res) == n_images
    _check(res, lg, reg, pr, counts, wh)


@pytest.mark.parametrize("dets,thresh,nms", [(0, 0.05, 0.5), (10, 0.3, 0.3), (100, 0.0, 0.7), (5, 0.9999, 0.5),
                                             (100, 0.05, 0.0), (0, 0.2, -1.0)])  # nms <= 0: NMS skipped
def test_parameters(dets, thresh, nms):
    lg, reg, pr, counts, wh = synth.box_head_batch(2, 400, num_classes=21, seed=4, ties=True)
    post = rp.PostProcessor(thresh, nms, dets)
    res = post((T(lg), T(reg)), _boxlists(pr, counts, wh))
    _check(res, lg, reg, pr, counts, wh, score_thresh=thresh, nms_thresh=nms, dets=dets)


def test_ties_beyond_the_slack_slots():
    # every foreground row identical in score: the kthvalue cut keeps far more than dets
    lg, reg, pr, counts, wh = synth.box_head_batch(1, 600, num_classes=5, seed=2, ties=True)
    lg[:] = lg[0]
    lg[:, 1] = 3.0
    pr = synth.proposals(600, 77, smin=20.0, smax=60.0, adversarial=False)  # mostly disjoint
    reg[:] = 0
    post = rp.PostProcessor(0.05, 0.5, 10)
    res = post((T(lg), T(reg)), _boxlists(pr, counts, wh))
    assert len(res[0]) > 10 + 64
    _check(res, lg, reg, pr, counts, wh, dets=10)


def test_candidate_slots_overflow_is_repeated():
    # threshold 0: every (row, class) pair is a candidate, far more than the first-try slots
    lg, reg, pr, counts, wh = synth.box_head_batch(2, 700, num_classes=21, seed=5)
    post = rp.PostProcessor(0.0, 0.5, 100)
    post.cand_slots_per_row = 1
    res = post((T(lg), T(reg)), _boxlists(pr, counts, wh))
    _check(res, lg, reg, pr, counts, wh, score_thresh=0.0)


def test_cls_agnostic_regression():
    lg, reg, pr, counts, wh = synth.box_head_batch(2, 300, num_classes=21, seed=6)
    reg = np.ascontiguousarray(reg[:, :8])  # [R, 2*4]: background + one shared box
    post = rp.PostProcessor(0.05, 0.5, 100, cls_agnostic_bbox_reg=True)
    res = post((T(lg), T(reg)), _boxlists(pr, counts, wh))
    _check(res, lg, reg, pr, counts, wh, cls_agnostic=True)


def test_probabilities_and_filter_results_twin():
    lg, reg, pr, counts, wh = synth.box_head_batch(1, 500, num_classes=21, seed=8, ties=True)
    n = counts[0]
    offs = T(np.array([0, n], np.int32))
    r = rp.box_post(T(lg), T(reg), T(pr), offs, T(wh), n, 0.05, 0.5, 100)
    np.testing.assert_array_equal(r["prob"].cpu().numpy(), oracle.softmax_rows(lg))
    # filter_results on a decoded BoxList (the bbox-aug entry) gives the same detections
    post = rp.PostProcessor(0.05, 0.5, 100)
    dec = rp.box_decode(T(reg), T(pr), (10.0, 10.0, 5.0, 5.0))
    bl = post.prepare_boxlist(dec, r["prob"], (int(wh[0, 0]), int(wh[0, 1]))).clip_to_image(remove_empty=False)
    twin = post.filter_results(bl, 21)
    ref = oracle.box_post_image(lg, reg, pr, wh[0])
    np.testing.assert_array_equal(twin.bbox.cpu().numpy(), ref[0])
    np.testing.assert_array_equal(twin.get_field("scores").cpu().numpy(), ref[1])
    np.testing.assert_array_equal(twin.get_field("labels").cpu().numpy(), ref[2])


def test_errors_are_loud():
    lg, reg, pr, counts, wh = synth.box_head_batch(1, 20, num_classes=5, seed=1)
    offs = T(np.array([0, 20], np.int32))
    with pytest.raises(ValueError):
        rp.box_post(T(lg), T(reg[:, :8]), T(pr), offs, T(wh), 20, 0.05, 0.5, 100)
    with pytest.raises(ValueError):
        rp.box_post(T(lg), T(reg), T(pr), offs, T(wh), 20, 0.05, float("nan"), 100)
    with pytest.raises(Exception):
        rp.box_post(torch.from_numpy(lg), T(reg), T(pr), offs, T(wh), 20, 0.05, 0.5, 100)


def test_filter_results_v2_matches_numpy_restatement():
    # inference.py:198-243: one ml_nms over all (row, class) candidates + topk; box column = label - 1
    lg, reg, pr, counts, wh = synth.box_head_batch(1, 400, num_classes=21, seed=12, ties=True)
    C = 21
    post = rp.PostProcessor(0.05, 0.5, 30)
    prob = oracle.softmax_rows(lg)
    dec = rp.box_decode(T(reg), T(pr), (10.0, 10.0, 5.0, 5.0))
    bl = post.prepare_boxlist(dec, T(prob), (int(wh[0, 0]), int(wh[0, 1]))).clip_to_image(remove_empty=False)
    got = post.filter_results_v2(bl, C)
    boxes = bl.bbox.cpu().numpy().reshape(-1, C, 4)
    sc = prob[:, 1:]
    r, c = np.nonzero(sc > np.float32(0.05))
    cb, cs, cl = boxes[r, c], sc[r, c], (c + 1).astype(np.float32)
    keep = oracle.ml_nms(cb, cs, cl, 0.5)[:30]
    np.testing.assert_array_equal(got.bbox.cpu().numpy(), cb[keep])
    np.testing.assert_array_equal(got.get_field("scores").cpu().numpy(), cs[keep])
    np.testing.assert_array_equal(got.get_field("labels").cpu().numpy(), (c + 1)[keep])
