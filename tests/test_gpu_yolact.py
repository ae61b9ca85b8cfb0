"""GPU parity of the Yolact post-processing path (decode, FastNMS, cross-class FastNMS,
traditional NMS, mask assembly / postprocess) against the oracle: indices and binary masks
bit-exact."""
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


@pytest.fixture(scope="module")
def priors():
    p = synth.yolact_priors()
    assert p.shape == (19248, 4)
    return p


def _candidates(seed, priors, conf_thresh=0.05):
    loc, conf, mask, proto = synth.yolact_image(seed, priors)
    boxes = oracle.yolact_decode(loc, priors)
    fg = np.ascontiguousarray(conf[:, 1:].T)           # [80, P]
    keep = oracle.detect_filter(fg, conf_thresh)
    return boxes[keep], mask[keep], np.ascontiguousarray(fg[:, keep]), proto, (loc, conf, mask)


def test_decode_bit_exact(priors):
    loc = synth.yolact_image(1, priors)[0]
    got = rp.decode(T(loc), T(priors)).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.yolact_decode(loc, priors))


@pytest.mark.parametrize("seed", [100, 101])
def test_fast_nms_matches_oracle(seed, priors):
    boxes, coeff, scores, _, _ = _candidates(seed, priors)
    assert 100 < boxes.shape[0] < 5000
    det = rp.Detect(81, 0, 200, 0.05, 0.5)
    for second in (False, True):
        ob, oc, ocl, osc, oi = oracle.fast_nms(boxes, coeff, scores, 0.5, 200, 100, second, 0.05)
        b, m, c, s = det.fast_nms(T(boxes), T(coeff), T(scores), 0.5, 200, second)
        np.testing.assert_array_equal(c.cpu().numpy(), ocl)
        np.testing.assert_array_equal(s.cpu().numpy(), osc)
        np.testing.assert_array_equal(b.cpu().numpy(), ob)
        np.testing.assert_array_equal(m.cpu().numpy(), oc)
    ob, oc, ocl, osc, oi = oracle.cc_fast_nms(boxes, coeff, scores, 0.5, 200)
    b, m, c, s = det.cc_fast_nms(T(boxes), T(coeff), T(scores), 0.5, 200)
    np.testing.assert_array_equal(c.cpu().numpy(), ocl)
    np.testing.assert_array_equal(s.cpu().numpy(), osc)
    np.testing.assert_array_equal(b.cpu().numpy(), ob)
    np.testing.assert_array_equal(m.cpu().numpy(), oc)


def test_fast_nms_small_ties_and_degenerate():
    rng = np.random.default_rng(3)
    n, C, D = 37, 5, 8
    boxes = rng.uniform(0, 1, (n, 4)).astype(np.float32)
    boxes[:, 2:] = boxes[:, :2] + rng.uniform(0.05, 0.4, (n, 2)).astype(np.float32)
    boxes[5] = boxes[6] = [0.3, 0.3, 0.3, 0.3]        # two zero-area boxes: 0/0 -> NaN -> dropped
    boxes[9] = boxes[8]                               # exact duplicate
    coeff = rng.standard_normal((n, D)).astype(np.float32)
    scores = (np.floor(rng.uniform(0, 1, (C, n)) * 8) / 8).astype(np.float32)  # heavy ties
    det = rp.Detect(C + 1, 0, 200, 0.05, 0.5, max_num_detections=20)
    for top_k in (200, 16):
        o = oracle.fast_nms(boxes, coeff, scores, 0.5, top_k, 20, False, 0.05)
        b, m, c, s = det.fast_nms(T(boxes), T(coeff), T(scores), 0.5, top_k)
        np.testing.assert_array_equal(c.cpu().numpy(), o[2])
        np.testing.assert_array_equal(s.cpu().numpy(), o[3])
        np.testing.assert_array_equal(b.cpu().numpy(), o[0])
        np.testing.assert_array_equal(m.cpu().numpy(), o[1])


def test_fast_nms_batched_matches_per_image(priors):
    cands = [_candidates(200 + i, priors) for i in range(3)]
    offs = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
    boxes = np.concatenate([c[0] for c in cands])
    coeff = np.concatenate([c[1] for c in cands])
    scores = np.concatenate([c[2] for c in cands], 1)
    r = rp.fast_nms_batched(T(boxes), T(coeff), T(scores), T(offs), 0.5, 200, 100)
    cnt = r["count"].cpu().numpy()
    for i, c in enumerate(cands):
        o = oracle.fast_nms(c[0], c[1], c[2], 0.5, 200, 100)
        assert cnt[i] == len(o[3])
        np.testing.assert_array_equal(r["score"][i, :cnt[i]].cpu().numpy(), o[3])
        np.testing.assert_array_equal(r["class"][i, :cnt[i]].cpu().numpy(), o[2])
        np.testing.assert_array_equal(r["index"][i, :cnt[i]].cpu().numpy(), o[4])
        np.testing.assert_array_equal(r["box"][i, :cnt[i]].cpu().numpy(), o[0])


@pytest.mark.parametrize("out_hw", [(550, 550), (137, 301), (138, 138)])
def test_mask_assembly_bit_exact(out_hw, priors):
    boxes, coeff, scores, proto, _ = _candidates(300, priors)
    ob, oc, ocl, osc, oi = oracle.fast_nms(boxes, coeff, scores, 0.5, 200, 100)
    # add hostile boxes: reversed corners, outside the unit square, zero size
    ob = ob.copy()
    ob[0] = [0.9, 0.8, 0.1, 0.2]
    ob[1] = [-0.2, -0.1, 1.3, 1.2]
    ob[2] = [0.5, 0.5, 0.5, 0.5]
    h, w = out_hw
    for crop in (True, False):
        m_o, b_o, low_o = oracle.mask_assemble(proto, oc, ob, h, w, crop, return_low=True)
        m, b, low = rp.mask_assemble(T(proto), T(oc), T(ob), h, w, crop, return_low=True)
        np.testing.assert_array_equal(low.cpu().numpy(), low_o)
        np.testing.assert_array_equal(b.cpu().numpy(), b_o)
        got = m.cpu().numpy()
        assert got.dtype == np.uint8 and got.max() <= 1
        np.testing.assert_array_equal(got, m_o)
        if crop:
            assert 0 < got.mean() < 0.5


def test_detect_and_postprocess_end_to_end(priors):
    B = 2
    imgs = [synth.yolact_image(400 + i, priors) for i in range(B)]
    pred = {"loc": T(np.stack([i[0] for i in imgs])), "conf": T(np.stack([i[1] for i in imgs])),
            "mask": T(np.stack([i[2] for i in imgs])), "priors": T(priors),
            "proto": T(np.stack([i[3] for i in imgs]))}
    det = rp.Detect(81, 0, 200, 0.05, 0.5)
    det.use_fast_nms = True
    out = det(pred, None)
    for i in range(B):
        loc, conf, mask, proto = imgs[i]
        boxes = oracle.yolact_decode(loc, priors)
        fg = np.ascontiguousarray(conf[:, 1:].T)
        keep = oracle.detect_filter(fg, 0.05)
        ob, oc, ocl, osc, oi = oracle.fast_nms(boxes[keep], mask[keep], np.ascontiguousarray(fg[:, keep]),
                                               0.5, 200, 100)
        m_o, b_o = oracle.mask_assemble(proto, oc, ob, 550, 550, True)
        classes, scores, iboxes, masks = rp.postprocess(out, 550, 550, batch_idx=i)
        np.testing.assert_array_equal(classes.cpu().numpy(), ocl)
        np.testing.assert_array_equal(scores.cpu().numpy(), osc)
        np.testing.assert_array_equal(iboxes.cpu().numpy(), b_o)
        assert masks.dtype == torch.bool
        np.testing.assert_array_equal(masks.cpu().numpy().astype(np.uint8), m_o)
    # no candidate above the threshold -> None -> empty outputs (detection.py:98-99)
    det2 = rp.Detect(81, 0, 200, 2.0, 0.5)
    det2.use_fast_nms = True
    o2 = det2(pred, None)
    assert o2[0]["detection"] is None
    assert all(t.numel() == 0 for t in rp.postprocess(o2, 550, 550))


def test_traditional_nms_matches_cython_semantics(priors):
    boxes, coeff, scores, _, _ = _candidates(500, priors)
    det = rp.Detect(81, 0, 200, 0.05, 0.5)
    b, m, c, s = det.traditional_nms(T(boxes), T(coeff), T(scores), 0.5, 0.05)
    # oracle restatement of detection.py:191-237
    scaled = (boxes * np.float32(550)).astype(np.float32)
    idx_l, cls_l, scr_l = [], [], []
    for k in range(scores.shape[0]):
        sel = np.nonzero(scores[k] > np.float32(0.05))[0]
        if len(sel) == 0:
            continue
        dets = np.concatenate([scaled[sel], scores[k][sel][:, None]], 1).astype(np.float32)
        keep = oracle.cython_nms(dets, 0.5)
        idx_l.append(sel[keep]); cls_l.append(np.full(len(keep), k)); scr_l.append(scores[k][sel][keep])
    idx, cls, scr = np.concatenate(idx_l), np.concatenate(cls_l), np.concatenate(scr_l)
    order = oracle.argsort_desc(scr)[:100]
    np.testing.assert_array_equal(s.cpu().numpy(), scr[order])
    np.testing.assert_array_equal(c.cpu().numpy(), cls[order])
    np.testing.assert_array_equal(m.cpu().numpy(), coeff[idx[order]])


def test_full_size_batch64_properties(priors):
    # BASELINE cfg 4 size: 64 images through FastNMS + assembly in batched launches.
    B = 64
    base = [_candidates(100 + (i % 4), priors) for i in range(4)]
    cands = [base[i % 4] for i in range(B)]
    offs = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
    boxes = T(np.concatenate([c[0] for c in cands]))
    coeff = T(np.concatenate([c[1] for c in cands]))
    scores = T(np.concatenate([c[2] for c in cands], 1))
    proto = T(np.stack([c[3] for c in cands]))
    r = rp.fast_nms_batched(boxes, coeff, scores, T(offs), 0.5, 200, 100)
    cnt = r["count"].cpu().numpy()
    assert np.all(cnt == cnt[0:4].tolist() * 16)       # identical images -> identical results
    sc = r["score"].cpu().numpy()
    for i in range(B):
        assert np.all(sc[i, :cnt[i] - 1] >= sc[i, 1:cnt[i]])
    det_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    sel = torch.cat([torch.arange(cnt[i], device=DEV) + i * 100 for i in range(B)])
    cf = r["mask"].reshape(B * 100, -1)[sel].contiguous()
    bx = r["box"].reshape(B * 100, 4)[sel].contiguous()
    masks, ib = rp.mask_assemble(proto, cf, bx, 550, 550, True, det_offsets=T(det_off))
    m0 = oracle.mask_assemble(cands[0][3], cf[:cnt[0]].cpu().numpy(), bx[:cnt[0]].cpu().numpy(), 550, 550, True)[0]
    np.testing.assert_array_equal(masks[:cnt[0]].cpu().numpy(), m0)
    # image 4 is a copy of image 0: checksum of checksums
    a = masks[:cnt[0]].sum(dtype=torch.int64)
    b4 = masks[det_off[4]: det_off[5]].sum(dtype=torch.int64)
    assert int(a) == int(b4) and int(a) > 0


@pytest.mark.parametrize("n,levels", [(700, 16), (3000, 1024), (7000, 64), (260, 4)])
def test_fast_nms_select_ties_large(n, levels):
    """The fused top-k selection (radix select + ordered tie pick + bitonic sort) must equal the
    first top_k entries of a stable descending sort: heavy ties, n above top_k, n above the
    shared-memory key cache (7000), two images of different sizes in one launch."""
    rng = np.random.default_rng(n)
    C, D = 6, 8
    sizes = [n, max(201, n // 3)]
    cands = []
    for m in sizes:
        ctr = rng.uniform(0.1, 0.9, (m, 2)).astype(np.float32)
        wh = rng.uniform(0.02, 0.3, (m, 2)).astype(np.float32)
        boxes = np.concatenate([ctr - wh / 2, ctr + wh / 2], 1).astype(np.float32)
        coeff = rng.standard_normal((m, D)).astype(np.float32)
        scores = (np.floor(rng.uniform(0, 1, (C, m)) * levels) / levels).astype(np.float32)
        scores[0, : m // 2] = 0.0  # a class where more than half of the candidates tie at zero
        cands.append((boxes, coeff, scores))
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    r = rp.fast_nms_batched(T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands])),
                            T(np.concatenate([c[2] for c in cands], 1)), T(offs), 0.5, 200, 100)
    cnt = r["count"].cpu().numpy()
    for i, c in enumerate(cands):
        o = oracle.fast_nms(c[0], c[1], c[2], 0.5, 200, 100)
        assert cnt[i] == len(o[3])
        np.testing.assert_array_equal(r["score"][i, :cnt[i]].cpu().numpy(), o[3])
        np.testing.assert_array_equal(r["class"][i, :cnt[i]].cpu().numpy(), o[2])
        np.testing.assert_array_equal(r["index"][i, :cnt[i]].cpu().numpy(), o[4])


def _batch(priors, seeds):
    imgs = [synth.yolact_image(s, priors) for s in seeds]
    return imgs, {"loc": T(np.stack([i[0] for i in imgs])), "conf": T(np.stack([i[1] for i in imgs])),
                  "mask": T(np.stack([i[2] for i in imgs])), "priors": T(priors),
                  "proto": T(np.stack([i[3] for i in imgs]))}


def test_detect_prefilter_matches_oracle(priors):
    imgs, pred = _batch(priors, [500, 501, 502])
    conf = pred["conf"].clone()
    conf[1, :, 1:] = 0.0  # image 1 has no candidate at all
    for cap in (None, 3000):  # 3000 < the number of candidates: truncated run, n_cand reports it
        r = rp.detect_prefilter(pred["loc"], conf, pred["mask"], pred["priors"], 0.05, cap)
        off = r["offsets"].cpu().numpy()
        n_cand = int(r["n_cand"].item())
        total = 0
        for b, (loc, cf, mask, _) in enumerate(imgs):
            fg = np.ascontiguousarray(cf[:, 1:].T) if b != 1 else np.zeros((80, priors.shape[0]), np.float32)
            keep = oracle.detect_filter(fg, 0.05)
            total += len(keep)
            lo, hi = off[b], off[b + 1]
            if cap is None:
                assert hi - lo == len(keep)
            keep = keep[: hi - lo]
            np.testing.assert_array_equal(r["prior"][lo:hi].cpu().numpy(), keep)
            np.testing.assert_array_equal(r["box"][lo:hi].cpu().numpy(), oracle.yolact_decode(loc, priors)[keep])
            np.testing.assert_array_equal(r["mask"][lo:hi].cpu().numpy(), mask[keep])
            np.testing.assert_array_equal(r["score"][:, lo:hi].cpu().numpy(), fg[:, keep])
        assert n_cand == total and off[0] == 0
        if cap is not None:
            assert n_cand > cap and off[-1] == cap


@pytest.mark.parametrize("cross_class", [False, True])
def test_detect_batched_matches_oracle(priors, cross_class):
    imgs, pred = _batch(priors, [510, 511, 512, 513, 514])
    pred["conf"][3, :, 1:] = 0.0
    det = rp.Detect(81, 0, 200, 0.05, 0.5)
    det.use_fast_nms = True
    det.use_cross_class_nms = cross_class
    out = det(pred, None)
    assert out[3]["detection"] is None
    for b, (loc, cf, mask, _) in enumerate(imgs):
        if b == 3:
            continue
        fg = np.ascontiguousarray(cf[:, 1:].T)
        keep = oracle.detect_filter(fg, 0.05)
        bx = oracle.yolact_decode(loc, priors)[keep]
        sc = np.ascontiguousarray(fg[:, keep])
        if cross_class:
            ob, oc, ocl, osc, _ = oracle.cc_fast_nms(bx, mask[keep], sc, 0.5, 200)
        else:
            ob, oc, ocl, osc, _ = oracle.fast_nms(bx, mask[keep], sc, 0.5, 200, 100)
        d = out[b]["detection"]
        np.testing.assert_array_equal(d["class"].cpu().numpy(), ocl)
        np.testing.assert_array_equal(d["score"].cpu().numpy(), osc)
        np.testing.assert_array_equal(d["box"].cpu().numpy(), ob)
        np.testing.assert_array_equal(d["mask"].cpu().numpy(), oc)


def test_detect_batched_repeats_when_the_candidate_slots_overflow(priors, monkeypatch):
    imgs, pred = _batch(priors, [520, 521])
    calls = []
    real = rp.package.yolact.detect_prefilter

    def small_first(loc, conf, mask, pri, thr, cap=None):
        calls.append(cap)
        return real(loc, conf, mask, pri, thr, cap if cap else 300)

    monkeypatch.setattr(rp.package.yolact, "detect_prefilter", small_first)
    det = rp.Detect(81, 0, 200, 0.05, 0.5)
    det.use_fast_nms = True
    out = det(pred, None)
    assert len(calls) == 2 and calls[0] is None and calls[1] > 300
    loc, cf, mask, _ = imgs[1]
    fg = np.ascontiguousarray(cf[:, 1:].T)
    keep = oracle.detect_filter(fg, 0.05)
    ob, oc, ocl, osc, _ = oracle.fast_nms(oracle.yolact_decode(loc, priors)[keep], mask[keep],
                                          np.ascontiguousarray(fg[:, keep]), 0.5, 200, 100)
    np.testing.assert_array_equal(out[1]["detection"]["score"].cpu().numpy(), osc)
    np.testing.assert_array_equal(out[1]["detection"]["box"].cpu().numpy(), ob)
