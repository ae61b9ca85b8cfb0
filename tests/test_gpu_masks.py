"""GPU parity of the Mask R-CNN mask post-processing (sigmoid + label select, Masker /
paste-back) against the oracle: binary masks bit-exact."""
import numpy as np
import pytest
import torch

import oracle
import roipath_b200 as rp
from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def T(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return t if dtype is None else t.to(dtype)


def test_mask_select_sigmoid_bit_exact():
    rng = np.random.default_rng(0)
    logits = (rng.standard_normal((23, 81, 28, 28)) * 4).astype(np.float32)
    logits[0, 0, 0, :6] = [-200, 200, 0, -87.5, 88.8, np.float32(1e-30)]
    labels = rng.integers(0, 81, 23).astype(np.int32)
    labels[0] = 0
    got = rp.mask_select_sigmoid(T(logits), T(labels)).cpu().numpy()
    np.testing.assert_array_equal(got[:, 0], oracle.mask_select_sigmoid(logits, labels))


@pytest.mark.parametrize("hw", [(800, 1333), (427, 641), (64, 48), (33, 16), (600, 1000)])
@pytest.mark.parametrize("thresh", [0.5, -1.0])
def test_paste_bit_exact(hw, thresh):
    h, w = hw
    masks, boxes = synth.paste_inputs(40, 200, img_w=w, img_h=h)
    if thresh < 0:
        masks = masks * (masks > 0.6)  # exact zeros so that (mask*255) != 0 is not all-true
    got = rp.paste_masks(T(masks), T(boxes), h, w, thresh, 1).cpu().numpy()
    exp = oracle.paste_masks(masks, boxes, h, w, thresh, 1)
    assert got.dtype == np.uint8 and got.max() <= 1
    np.testing.assert_array_equal(got, exp)
    assert 0 < exp.mean() < 0.6


def test_masker_and_postprocessor_interfaces():
    h, w = 800, 1333
    rng = np.random.default_rng(2)
    per = [7, 0, 12]
    lists, probs, logits_all = [], [], []
    for i, n in enumerate(per):
        m, b = synth.paste_inputs(max(n, 8), 300 + i, img_w=w, img_h=h)
        bl = rp.BoxList(T(b[:n]).reshape(-1, 4), (w, h))
        bl.add_field("labels", T(rng.integers(1, 81, n).astype(np.int64)))
        lists.append(bl)
        probs.append(m[:n])
        logits_all.append((rng.standard_normal((n, 81, 28, 28)) * 3).astype(np.float32))
    masker = rp.Masker(threshold=0.5, padding=1)
    res = masker([T(p).reshape(-1, 1, 28, 28) for p in probs], lists)
    for i, n in enumerate(per):
        assert tuple(res[i].shape)[:2] == (n, 1)
        if n:
            assert tuple(res[i].shape) == (n, 1, h, w) and res[i].dtype == torch.bool
            exp = oracle.paste_masks(probs[i], lists[i].bbox.cpu().numpy(), h, w, 0.5, 1)
            np.testing.assert_array_equal(res[i][:, 0].cpu().numpy().astype(np.uint8), exp)
    # single-instance entry point
    one = rp.paste_mask_in_image(T(probs[0][0]), lists[0].bbox[0], h, w)
    np.testing.assert_array_equal(one.cpu().numpy(), res[0][0, 0].cpu().numpy())
    # MaskPostProcessor: sigmoid + label select + Masker
    pp = rp.MaskPostProcessor(masker)
    out = pp(T(np.concatenate(logits_all)), lists)
    for i, n in enumerate(per):
        if not n:
            continue
        lab = lists[i].get_field("labels").cpu().numpy().astype(np.int32)
        prob = oracle.mask_select_sigmoid(logits_all[i], lab)
        exp = oracle.paste_masks(prob, lists[i].bbox.cpu().numpy(), h, w, 0.5, 1)
        np.testing.assert_array_equal(out[i].get_field("mask")[:, 0].cpu().numpy().astype(np.uint8), exp)


def test_full_size_properties():
    # BASELINE cfg 5 size for one image: 100 instances at 1333x800.  Outside the expanded box the
    # output is zero, and pasting is independent per instance (permutation equivariance).
    h, w = 800, 1333
    masks, boxes = synth.paste_inputs(100, 200)
    out = rp.paste_masks(T(masks), T(boxes), h, w)
    assert tuple(out.shape) == (100, h, w)
    perm = np.random.default_rng(0).permutation(100)
    out2 = rp.paste_masks(T(masks[perm]), T(boxes[perm]), h, w)
    assert torch.equal(out[torch.from_numpy(perm).to(DEV)], out2)
    exp = oracle.paste_masks(masks[:10], boxes[:10], h, w)
    np.testing.assert_array_equal(out[:10].cpu().numpy(), exp)
    ys = torch.arange(h, device=DEV).view(1, h, 1)
    xs = torch.arange(w, device=DEV).view(1, 1, w)
    b = T(boxes)
    outside = (xs < (b[:, 0] - 40).view(-1, 1, 1)) | (xs > (b[:, 2] + 40).view(-1, 1, 1)) | \
              (ys < (b[:, 1] - 40).view(-1, 1, 1)) | (ys > (b[:, 3] + 40).view(-1, 1, 1))
    assert int((out.bool() & outside).sum()) == 0


@pytest.mark.parametrize("shape", [(3, 550, 550), (5, 9, 13), (2, 800, 1333), (1, 1, 7), (4, 16, 64)])
def test_pack_mask_bits_matches_numpy(shape):
    # rp_pack_mask_bits (the final gather's payload) against numpy.packbits, LSB first
    rng = np.random.default_rng(shape[1])
    m = (rng.random(shape) > 0.6).astype(np.uint8)
    got = rp.parallel.pack_masks(T(m)).cpu().numpy()
    exp = np.packbits(m.reshape(shape[0], -1), axis=1, bitorder="little")
    np.testing.assert_array_equal(got, exp)
    back = rp.parallel.unpack_masks(T(exp), shape[1], shape[2]).cpu().numpy()
    np.testing.assert_array_equal(back, m)


@pytest.mark.parametrize("n,H,W,crop,scaled", [(9, 23, 31, True, True), (100, 138, 138, True, True),
                                               (37, 138, 138, False, False), (1, 5, 7, True, False),
                                               (0, 20, 20, True, False)])
def test_lincomb_mask_loss_matches_oracle(n, H, W, crop, scaled):
    # rp_lincomb_mask_loss (multibox_loss.py:616-660 forward + backward) against the oracle, 1e-5
    rng = np.random.default_rng(n + H)
    D = 32
    proto = np.maximum(rng.standard_normal((H, W, D)), 0).astype(np.float32)
    coef = np.tanh(rng.standard_normal((n, D))).astype(np.float32)
    mask_t = (rng.random((H, W, n)) > 0.6).astype(np.float32)
    c = rng.uniform(0.2, 0.8, (n, 2))
    wh = rng.uniform(0.1, 0.5, (n, 2))
    boxes = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    scale = None
    if scaled:
        scale = (H * W / ((boxes[:, 2] - boxes[:, 0]) * W) / ((boxes[:, 3] - boxes[:, 1]) * H)).astype(np.float32)
    bx = boxes if crop else None
    loss, dp, dc = rp.lincomb_mask_loss_raw(T(proto), T(coef), T(mask_t), T(bx) if bx is not None else None,
                                            T(scale) if scale is not None else None)
    if n == 0:
        assert loss.numel() == 0 and float(dp.abs().max()) == 0.0
        return
    el, edp, edc = oracle.lincomb_mask_loss(proto, coef, mask_t, bx, scale)
    np.testing.assert_allclose(loss.cpu().numpy(), el, rtol=1e-5)
    np.testing.assert_allclose(dp.cpu().numpy(), edp, rtol=1e-5, atol=1e-5 * np.abs(edp).max())
    np.testing.assert_allclose(dc.cpu().numpy(), edc, rtol=1e-5, atol=1e-5 * np.abs(edc).max())


def test_lincomb_mask_loss_autograd():
    # the autograd wrapper: gradients of 0.5 * total reach proto and coef scaled by the upstream factor
    rng = np.random.default_rng(0)
    proto = torch.from_numpy(np.maximum(rng.standard_normal((20, 24, 32)), 0).astype(np.float32)).to(DEV).requires_grad_(True)
    coef = torch.from_numpy(np.tanh(rng.standard_normal((6, 32))).astype(np.float32)).to(DEV).requires_grad_(True)
    mask_t = T((rng.random((20, 24, 6)) > 0.5).astype(np.float32))
    boxes = T(np.tile(np.array([[0.1, 0.2, 0.8, 0.9]], np.float32), (6, 1)))
    total, per = rp.lincomb_mask_loss(proto, coef, mask_t, boxes, normalize_emulate_roi_pooling=True)
    (0.5 * total).backward()
    _, dp, dc = rp.lincomb_mask_loss_raw(proto.detach(), coef.detach(), mask_t, boxes,
                                         torch.full((6,), 20 * 24 / (0.7 * 24) / (0.7 * 20), device=DEV))
    torch.testing.assert_close(proto.grad, 0.5 * dp, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(coef.grad, 0.5 * dc, rtol=1e-5, atol=1e-5)
    assert abs(float(total) - float(per.sum())) <= 1e-5 * abs(float(total))
