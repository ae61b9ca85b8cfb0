"""GPU parity of the other RoI-family operators (ROIPool fwd/bwd, boxlist_iou, EmbedMask
mask_prob) against the oracle: bit-exact except the atomically accumulated ROIPool gradient."""
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


@pytest.mark.parametrize("size,scale,shape", [(7, 0.0625, (2, 16, 50, 84)), ((3, 5), 0.125, (2, 3, 100, 168)),
                                               (14, 0.25, (1, 4, 200, 336)), (2, 0.03125, (3, 8, 25, 42))])
def test_roi_pool_matches_oracle(size, scale, shape):
    rng = np.random.default_rng(shape[2])
    feat = rng.standard_normal(shape).astype(np.float32)
    rois = synth.rois_for_images(shape[0], 100, seed=3)
    rois[5, 1:] = (-50.0, -40.0, -20.0, -10.0)   # entirely outside: empty bins -> 0 / -1
    rois[6, 1:] = (300.0, 200.0, 100.0, 50.0)    # malformed: forced to 1x1
    f = T(feat).requires_grad_(True)
    out, arg = rp.roi_pool(f, T(rois), size, scale, return_argmax=True)
    o_ref, a_ref = oracle.roi_pool_forward(feat, rois, size, scale)
    np.testing.assert_array_equal(arg.cpu().numpy(), a_ref)
    np.testing.assert_array_equal(out.detach().cpu().numpy(), o_ref)
    g = rng.standard_normal(o_ref.shape).astype(np.float32)
    out.backward(T(g))
    g_ref = oracle.roi_pool_backward(g, rois, a_ref, shape)
    np.testing.assert_allclose(f.grad.cpu().numpy(), g_ref, rtol=1e-5, atol=1e-5)
    mod = rp.ROIPool(size, scale)
    np.testing.assert_array_equal(mod(T(feat), T(rois)).cpu().numpy(), o_ref)
    assert rp.roi_pool(T(feat), T(rois[:0]), size, scale).shape[0] == 0


@pytest.mark.parametrize("n,m", [(1, 1), (37, 5), (2000, 33), (513, 1000)])
def test_boxlist_iou_matches_oracle(n, m):
    a, b = synth.proposals(n, n, adversarial=n >= 8), synth.proposals(m, m + 1, adversarial=False)
    A, B = rp.BoxList(T(a), (synth.IMG_W, synth.IMG_H)), rp.BoxList(T(b), (synth.IMG_W, synth.IMG_H))
    np.testing.assert_array_equal(rp.boxlist_iou(A, B).cpu().numpy(), oracle.boxlist_iou(a, b))
    with pytest.raises(RuntimeError):
        rp.boxlist_iou(A, rp.BoxList(T(b), (10, 10)))
    assert rp.boxlist_iou(A, rp.BoxList(T(b[:0]), (synth.IMG_W, synth.IMG_H))).shape == (n, 0)


@pytest.mark.parametrize("n_center,h,w,dim", [(1, 8, 8, 4), (6, 40, 56, 8), (40, 100, 136, 32)])
def test_mask_prob_matches_oracle(n_center, h, w, dim):
    ep, ec, sg, bx, ar, mw = synth.embed_mask_inputs(n_center, h, w, dim, seed=dim)
    got = rp.mask_prob_cuda(T(ep), T(ec), T(sg), T(bx), T(ar), int(ar.sum()), mw)
    np.testing.assert_array_equal(got.cpu().numpy(), oracle.mask_prob(ep, ec, sg, bx, ar, mw))
    assert rp.mask_prob_cuda(T(ep), T(ec[:0]), T(sg[:0]), T(bx[:0]), T(ar[:0]), 0, mw).numel() == 0


@pytest.mark.parametrize("h,w,padding", [(100, 136, 0.0), (37, 53, 0.1), (8, 8, 0.5)])
def test_box_masks_match_oracle(h, w, padding):
    rng = np.random.default_rng(h)
    n = 9
    x1 = rng.uniform(-5, w - 2, n); y1 = rng.uniform(-5, h - 2, n)
    boxes = np.stack([x1, y1, x1 + rng.uniform(0, w, n), y1 + rng.uniform(0, h, n)], 1).astype(np.float32)
    boxes[0] = (3.0, 2.0, 3.0, 2.0)
    got = rp.boxes_to_masks(T(boxes), h, w, padding)
    assert got.dtype == torch.bool
    np.testing.assert_array_equal(got.cpu().numpy().astype(np.uint8), oracle.box_masks(boxes, h, w, padding)[1])
    masks = rng.uniform(0, 1, (n, h, w)).astype(np.float32)
    out, crop = rp.crop_by_box(T(masks), T(boxes), padding)
    o_ref, c_ref = oracle.box_masks(boxes, h, w, padding, crop=True, masks=masks)
    np.testing.assert_array_equal(crop.cpu().numpy().astype(np.uint8), c_ref)
    np.testing.assert_array_equal(out.cpu().numpy(), o_ref)
    # the numpy form of boxlist_ops.py:237-261
    b = boxes
    bw, bh = b[:, 2] - b[:, 0], b[:, 3] - b[:, 1]
    X1 = np.clip(b[:, 0] - 1 - bw * np.float32(padding), 0, None); X2 = np.clip(b[:, 2] + 1 + bw * np.float32(padding), None, w)
    Y1 = np.clip(b[:, 1] - 1 - bh * np.float32(padding), 0, None); Y2 = np.clip(b[:, 3] + 1 + bh * np.float32(padding), None, h)
    cols = np.arange(w, dtype=np.float32)[None, None, :]; rows = np.arange(h, dtype=np.float32)[None, :, None]
    ref = (cols >= X1[:, None, None]) & (cols < X2[:, None, None]) & (rows >= Y1[:, None, None]) & (rows < Y2[:, None, None])
    np.testing.assert_array_equal(got.cpu().numpy(), ref)


def test_affine_paste_bit_exact():
    # rp_affine_paste against the oracle (pinned to cv2.warpAffine in tests/test_oracle_golden.py) and,
    # when cv2 is importable on the box, against cv2 itself
    from tests.test_oracle_golden import _affine_cases
    rng = np.random.default_rng(5)
    n, width, height = 16, 333, 250
    e = np.exp(rng.standard_normal((n, 64, 64, 2)).astype(np.float32))
    pred = (e / e.sum(-1, keepdims=True)).astype(np.float32)
    M = _affine_cases(rng, n)
    got = rp.affine_paste(T(pred), M, width, height).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.affine_paste(pred, M, width, height))
    one = rp.warpAffine(T(pred[3]), np.vstack([M[3], [0, 0, 1]]), width, height).cpu().numpy()
    np.testing.assert_array_equal(one, got[3])
    try:
        import cv2
    except ImportError:
        return
    for i in range(n):
        ref = cv2.warpAffine(pred[i], M[i], (width, height), borderMode=cv2.BORDER_CONSTANT,
                             flags=cv2.WARP_INVERSE_MAP + cv2.INTER_LINEAR)[:, :, 1]
        np.testing.assert_array_equal(got[i], (ref > 0.5).astype(np.uint8))


def test_embed_mask_stage_matches_oracle():
    # rp_embed_masks / rp_resize_threshold against the composed oracle (inference.py:153-202)
    rng = np.random.default_rng(9)
    dim, m_h, m_w, n, stride = 32, 100, 136, 24, 4.0
    pe = (rng.standard_normal((dim, m_h, m_w)) * 0.3).astype(np.float32)
    ce = (rng.standard_normal((n, dim)) * 0.3).astype(np.float32)
    mg = rng.uniform(0.05, 2.0, n).astype(np.float32)
    bx = synth.proposals(n, 3, img_w=int(m_w * stride), img_h=int(m_h * stride), smin=20, smax=300, adversarial=False)
    bx[0] = (0, 0, m_w * stride - 1, m_h * stride - 1)
    bx[1] = (50.5, 60.25, 50.9, 60.75)       # sub-pixel box
    bx[2] = (-30, -20, 40, 50)               # partly outside
    exp = oracle.embed_masks(pe, ce, mg, bx, stride)
    got = rp.embed_masks(T(pe), T(ce), T(mg), T(bx), stride)
    np.testing.assert_array_equal(got.cpu().numpy(), exp)
    # the unfused mirror of compute_mask_prob + boxes_to_masks gives the same product
    stage = rp.EmbedMaskMaskStage(8, 2, 0.5, True)
    prob = stage.compute_mask_prob(T(pe), T(ce), T(mg), T(bx) / stride)
    twin = prob * rp.boxes_to_masks(T(bx) / stride, m_h, m_w).float()
    np.testing.assert_array_equal(twin.cpu().numpy(), exp)
    in_h, in_w = 390, 540
    up = rp.resize_threshold(got, stride, 0.5, in_h, in_w).cpu().numpy()
    np.testing.assert_array_equal(up, oracle.resize_threshold(exp, int(m_h * stride), int(m_w * stride), 0.5, in_h, in_w))
    bl = rp.BoxList(T(bx), (in_w, in_h))
    bl.add_field("proposal_embed", T(ce))
    bl.add_field("proposal_margin", T(mg))
    bl.add_field("labels", torch.ones(n, device=DEV))
    bl.add_field("scores", torch.ones(n, device=DEV))
    res = stage.forward_for_mask([bl], T(pe)[None])
    np.testing.assert_array_equal(res[0].get_field("mask").cpu().numpy().astype(np.uint8), up)
