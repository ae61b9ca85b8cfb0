"""GPU parity of the greedy-NMS family against the oracle and the reference's golden vectors:
keep indices bit-exact, including order and stable-sort tie-breaking."""
import os

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


def test_reference_golden_vectors(golden_dir):
    k = np.load(os.path.join(golden_dir, "nms_kat.npz"))
    for thr, gt in zip(k["thresholds5"], k["keep5"]):
        keep = rp.nms(T(k["boxes5"]), T(k["scores5"]), float(thr)).cpu().numpy()
        np.testing.assert_array_equal(np.sort(keep), gt[gt >= 0])  # detectron tests/test_nms.py:53
    keep = rp.nms(T(k["boxes53"]), T(k["scores53"]), 0.5).cpu().numpy()
    np.testing.assert_array_equal(np.sort(keep), k["keep53"])     # tests/test_nms.py:184-213
    np.testing.assert_array_equal(keep, oracle.nms(k["boxes53"], k["scores53"], 0.5))


@pytest.mark.parametrize("n", [1, 2, 63, 64, 65, 257, 2000, 5000])
@pytest.mark.parametrize("tie_heavy", [False, True])
def test_matches_oracle(n, tie_heavy):
    b, s, _ = synth.nms_boxes(n, 7 + n, tie_heavy=tie_heavy)
    for thr in (0.7, 0.5, 0.3):
        keep = rp.nms(T(b), T(s), thr).cpu().numpy()
        np.testing.assert_array_equal(keep, oracle.nms(b, s, thr))


def test_threshold_compared_as_double():
    # two boxes whose IoU is exactly float32(0.3): (double)iou > 0.3 holds (nms.py:34)
    f = np.float32(0.3)
    # search a pair with iou == f is brittle; instead check the comparator against the oracle on
    # a dense family of near-threshold pairs
    rng = np.random.default_rng(0)
    n = 4000
    base = np.array([100, 100, 199, 199], np.float32)
    shift = rng.uniform(0, 60, (n, 1)).astype(np.float32)
    b = np.tile(base, (n, 1)) + np.concatenate([shift, 0 * shift, shift, 0 * shift], 1)
    s = np.linspace(1, 0.1, n).astype(np.float32)
    for thr in (0.3, 0.5, 0.7, 1 / 3):
        np.testing.assert_array_equal(rp.nms(T(b), T(s), thr).cpu().numpy(), oracle.nms(b, s, thr))


def test_ml_nms_and_max_proposals():
    b, s, l = synth.nms_boxes(3000, 11)
    l = (l % 5).astype(np.float32)
    keep = rp.ml_nms(T(b), T(s), T(l), 0.5).cpu().numpy()
    np.testing.assert_array_equal(keep, oracle.ml_nms(b, s, l, 0.5))
    bl = rp.BoxList(T(b), (synth.IMG_W, synth.IMG_H))
    bl.add_field("scores", T(s))
    bl.add_field("labels", T(l).long())
    for mp in (-1, 1, 100, 1000):
        o = oracle.nms(b, s, 0.7, max_keep=mp)
        r = rp.boxlist_nms(bl, 0.7, max_proposals=mp)
        np.testing.assert_array_equal(r.bbox.cpu().numpy(), b[o])
        np.testing.assert_array_equal(r.get_field("scores").cpu().numpy(), s[o])
    r = rp.boxlist_ml_nms(bl, 0.5, max_proposals=50)
    np.testing.assert_array_equal(r.bbox.cpu().numpy(), b[oracle.ml_nms(b, s, l, 0.5, max_keep=50)])
    assert rp.boxlist_nms(bl, 0.0) is bl  # boxlist_ops.py:24-25


def test_batched_ragged_and_empty():
    sizes = [0, 1, 700, 64, 0, 1300, 5]
    bs, ss = [], []
    for i, n in enumerate(sizes):
        b, s, _ = synth.nms_boxes(max(n, 1), 100 + i)
        bs.append(b[:n]); ss.append(s[:n])
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    keep, nk = rp.nms_batched(T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs), 0.7, n_max=max(sizes))
    keep, nk = keep.cpu().numpy(), nk.cpu().numpy()
    for i, n in enumerate(sizes):
        o = oracle.nms(bs[i], ss[i], 0.7) if n else np.zeros((0,), np.int32)
        assert nk[i] == len(o)
        np.testing.assert_array_equal(keep[offs[i]: offs[i] + nk[i]], o)
    lists = []
    for i, n in enumerate(sizes):
        bl = rp.BoxList(T(bs[i]).reshape(-1, 4), (synth.IMG_W, synth.IMG_H))
        bl.add_field("scores", T(ss[i]))
        lists.append(bl)
    res = rp.boxlist_nms_batched(lists, 0.7, max_proposals=100)
    for i, n in enumerate(sizes):
        o = oracle.nms(bs[i], ss[i], 0.7, max_keep=100) if n else np.zeros((0,), np.int64)
        np.testing.assert_array_equal(res[i].bbox.cpu().numpy().reshape(-1, 4), bs[i][o].reshape(-1, 4))


def test_cython_variant(golden_dir):
    c = np.load(os.path.join(golden_dir, "cython_nms_ref.npz"))
    for key, thr in (("keep_3", 0.3), ("keep_5", 0.5), ("keep_7", 0.7)):
        got = rp.cython_nms(T(c["dets"]), float(np.float32(thr))).cpu().numpy()
        np.testing.assert_array_equal(got, c[key])  # the reference's own cython_nms.pyx output
    b, s, _ = synth.nms_boxes(1500, 3, tie_heavy=True)
    dets = np.concatenate([b, s[:, None]], 1)
    np.testing.assert_array_equal(rp.cython_nms(T(dets), 0.5).cpu().numpy(), oracle.cython_nms(dets, 0.5))


def test_full_size_20k_batch8():
    # BASELINE cfg 2 at full size: 20k boxes x batch 8, against the oracle and through properties
    B, n = 8, 20000
    bs, ss = zip(*[synth.nms_boxes(n, 7 + i)[:2] for i in range(B)])
    offs = (np.arange(B + 1) * n).astype(np.int32)
    for mk in (-1, 1000):
        keep, nk = rp.nms_batched(T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs), 0.7,
                                  max_keep=mk, n_max=n)
        keep, nk = keep.cpu().numpy(), nk.cpu().numpy()
        for i in (0, B - 1):
            o = oracle.nms(bs[i], ss[i], 0.7, max_keep=mk)
            np.testing.assert_array_equal(keep[offs[i]: offs[i] + nk[i]], o)
        for i in range(B):
            k = keep[offs[i]: offs[i] + nk[i]]
            sc = ss[i][k]
            assert np.all(sc[:-1] >= sc[1:])          # sortedness
            assert len(np.unique(k)) == len(k)
            if mk < 0:                                # idempotence: NMS of the survivors keeps all
                k2 = rp.nms(T(bs[i][k]), T(ss[i][k]), 0.7).cpu().numpy()
                np.testing.assert_array_equal(k2, np.arange(len(k)))


def test_n_total_and_n_max_may_be_upper_bounds():
    # a producer with device-side counts (rp_rpn_pre_nms) hands over padded arrays: rows past
    # offsets[B] must be ignored, whatever they hold
    B, n = 3, 700
    bs, ss = zip(*[synth.nms_boxes(n - 50 * i, 40 + i, tie_heavy=True)[:2] for i in range(B)])
    lens = [b.shape[0] for b in bs]
    boxes, scores = np.concatenate(bs), np.concatenate(ss)
    pad = 900
    boxes_p = np.concatenate([boxes, np.full((pad, 4), 1e9, np.float32)])
    scores_p = np.concatenate([scores, np.full((pad,), 7.0, np.float32)])  # higher than any real score
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    keep, n_keep = rp.nms_batched(T(boxes_p), T(scores_p), T(offs), 0.5, n_max=1024)
    keep, n_keep = keep.cpu().numpy(), n_keep.cpu().numpy()
    for i in range(B):
        ref = oracle.nms(bs[i], ss[i], 0.5)
        assert n_keep[i] == len(ref)
        np.testing.assert_array_equal(keep[offs[i]: offs[i] + n_keep[i]], ref)


def test_iou_comparator_on_exact_threshold_pairs():
    """Exhaustive family of INTEGER-coordinate box pairs (legacy +1 areas): box A = 10 x 10 at the origin
    against every box B inside [0, 12]^2 -- 8281 two-box problems whose IoU values are small exact
    rationals, many of them EXACTLY 0.3, 0.5, 0.25, 1/3 or within one ulp of a threshold.  The
    reference compares (double)iou_f32 > thr (det/layers/nms.py:23,34): float32(30/100) > 0.3 holds,
    0.5 > 0.5 does not.  The kernel's comparator (fminf/fmaxf, division-free pre-filter, t_cmp =
    float_le(thr)) must agree with the oracle's compare-and-select on every one of them."""
    bs = []
    for x1 in range(13):
        for x2 in range(x1, 13):
            for y1 in range(13):
                for y2 in range(y1, 13):
                    bs.append((x1, y1, x2, y2))
    B = len(bs)
    boxes = np.zeros((B, 2, 4), np.float32)
    boxes[:, 0] = (0, 0, 9, 9)
    boxes[:, 1] = np.asarray(bs, np.float32)
    scores = np.tile(np.array([0.9, 0.5], np.float32), (B, 1))
    offs = (np.arange(B + 1) * 2).astype(np.int32)
    flat_b, flat_s = boxes.reshape(-1, 4), scores.reshape(-1)
    # the family does contain exact-threshold pairs
    inter = (np.minimum(9, boxes[:, 1, 2]) - boxes[:, 1, 0] + 1).clip(0) * (np.minimum(9, boxes[:, 1, 3]) - boxes[:, 1, 1] + 1).clip(0)
    area_b = (boxes[:, 1, 2] - boxes[:, 1, 0] + 1) * (boxes[:, 1, 3] - boxes[:, 1, 1] + 1)
    iou = inter / (100 + area_b - inter)
    for thr in (0.3, 0.5, 0.25):
        assert (np.abs(iou - thr) < 1e-7).sum() >= 3
    for thr in (0.3, 0.5, 0.7, 0.25, 1 / 3, 0.1, 0.9):
        keep, nk = rp.nms_batched(T(flat_b), T(flat_s), T(offs), thr, n_max=2)
        nk = nk.cpu().numpy()
        exp = np.array([len(oracle.nms(boxes[i], scores[i], thr)) for i in range(B)])
        bad = np.nonzero(nk != exp)[0]
        assert bad.size == 0, "thr %r: %d pairs disagree, first B=%s iou=%r" % (thr, bad.size, bs[bad[0]], iou[bad[0]])
        # the cython variant's comparator: iou >= float(thr)
        keep, nk = rp.nms_batched(T(flat_b), T(flat_s), T(offs), thr, n_max=2, flags=1)  # RP_NMS_GE
        exp_ge = np.array([len(oracle.cython_nms(np.concatenate([boxes[i], scores[i][:, None]], 1), thr)) for i in range(B)])
        assert np.array_equal(nk.cpu().numpy(), exp_ge)


def _chain_boxes(n, step):
    # box i = box 0 shifted by i * step pixels: every box overlaps its neighbours only, so the greedy
    # result alternates along one long dependency chain (the worst case of the band fixpoint)
    x = (np.arange(n) * step).astype(np.float32)
    b = np.stack([x, np.zeros(n, np.float32), x + 99.0, np.full(n, 49.0, np.float32)], 1)
    s = np.linspace(1.0, 0.1, n).astype(np.float32)
    return b, s


_NMS_MODES = {
    "default": {},                                        # dataflow kernel for few words, else grid barriers
    "dataflow": {"RP_NMS_FLOW_WORDS": "1000000"},         # dataflow kernel for every batch of <= 16 problems
    "barriers": {"RP_NMS_FLOW": "0"},                     # one cooperative launch, grid barriers
    "launches": {"RP_NMS_FUSED": "0"},                    # one resolve + one suppress launch per band
}


@pytest.mark.parametrize("mode", sorted(_NMS_MODES))
def test_one_launch_and_per_band_launch_paths(monkeypatch, mode):
    # The band loop has three drivers; all must give the oracle's keeps, bit for bit.
    for k, v in _NMS_MODES[mode].items():
        monkeypatch.setenv(k, v)
    rp.reload_env()
    try:
        b, s, _ = synth.nms_boxes(5000, 11, tie_heavy=True)
        for thr, mk in ((0.7, -1), (0.5, 300), (0.3, -1)):
            keep, nk = rp.nms_batched(T(b), T(s), T(np.array([0, 5000], np.int32)), thr, max_keep=mk)
            np.testing.assert_array_equal(keep.cpu().numpy()[: int(nk[0])], oracle.nms(b, s, thr, max_keep=mk))
        # ragged batch with labels, problems of 1 .. 3 bands
        lens = [2500, 0, 1100, 64, 3000]
        bs, ss, ls = zip(*[synth.nms_boxes(max(n, 1), 60 + i) for i, n in enumerate(lens)])
        bs = [x[:n] for x, n in zip(bs, lens)]
        ss = [x[:n] for x, n in zip(ss, lens)]
        ls = [x[:n] for x, n in zip(ls, lens)]
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        keep, nk = rp.nms_batched(T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs), 0.5,
                                  labels=T(np.concatenate(ls).astype(np.float32)), max_keep=200)
        keep, nk = keep.cpu().numpy(), nk.cpu().numpy()
        for i, n in enumerate(lens):
            ref = oracle.nms(bs[i], ss[i], 0.5, labels=ls[i], max_keep=200) if n else np.zeros(0, np.int32)
            np.testing.assert_array_equal(keep[offs[i]: offs[i] + nk[i]], ref)
        # one dependency chain through 3 bands: hundreds of fixpoint rounds per band
        cb, cs = _chain_boxes(2500, 20.0)
        for thr in (0.5, 0.3):
            keep = rp.nms(T(cb), T(cs), thr).cpu().numpy()
            np.testing.assert_array_equal(keep, oracle.nms(cb, cs, thr))
        dets = np.concatenate([cb, cs[:, None]], 1)
        np.testing.assert_array_equal(rp.cython_nms(T(dets), 0.5).cpu().numpy(), oracle.cython_nms(dets, 0.5))
        # 20 problems (more than the dataflow kernel takes), two bands each
        bs, ss = zip(*[synth.nms_boxes(1500, 90 + i)[:2] for i in range(20)])
        offs = (np.arange(21) * 1500).astype(np.int32)
        keep, nk = rp.nms_batched(T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs), 0.6, n_max=1500)
        keep, nk = keep.cpu().numpy(), nk.cpu().numpy()
        for i in (0, 7, 19):
            np.testing.assert_array_equal(keep[offs[i]: offs[i] + nk[i]], oracle.nms(bs[i], ss[i], 0.6))
    finally:
        for k in _NMS_MODES[mode]:
            monkeypatch.delenv(k)
        rp.reload_env()


@pytest.mark.parametrize("mode", ["default", "dataflow", "barriers"])
def test_random_batches_stress(monkeypatch, mode):
    # the one-launch drivers synchronise CTAs through counters in global memory: many random shapes
    # (ragged, empty, 1-6 bands, with and without labels / caps), each launched several times in a row
    for k, v in _NMS_MODES[mode].items():
        monkeypatch.setenv(k, v)
    rp.reload_env()
    try:
        rng = np.random.default_rng(1234)
        for case in range(14):
            B = int(rng.integers(1, 14))
            lens = [int(rng.choice([0, 1, 63, 64, 65, 300, 1024, 1025, 1500, 2500, 4000, 6000])) for _ in range(B)]
            if case % 4 == 0:
                lens = [int(x) for x in rng.integers(0, 1200, B)]
            parts = [synth.nms_boxes(max(n, 1), 1000 + 31 * case + i, tie_heavy=bool(i & 1)) for i, n in enumerate(lens)]
            bs = [p[0][:n] for p, n in zip(parts, lens)]
            ss = [p[1][:n] for p, n in zip(parts, lens)]
            ls = [(p[2][:n] % 3).astype(np.float32) for p, n in zip(parts, lens)]
            offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
            if int(offs[-1]) == 0:
                continue
            thr = float(rng.choice([0.3, 0.5, 0.7]))
            mk = int(rng.choice([-1, -1, 50, 700]))
            use_labels = bool(case % 3 == 1)
            bt, st, ot = T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs)
            lt = T(np.concatenate(ls)) if use_labels else None
            refs = [oracle.nms(bs[i], ss[i], thr, labels=ls[i] if use_labels else None, max_keep=mk)
                    if lens[i] else np.zeros(0, np.int32) for i in range(B)]
            for rep in range(3):
                keep, nk = rp.nms_batched(bt, st, ot, thr, labels=lt, max_keep=mk)
                keep, nk = keep.cpu().numpy(), nk.cpu().numpy()
                for i in range(B):
                    np.testing.assert_array_equal(keep[offs[i]: offs[i] + nk[i]], refs[i],
                                                  err_msg="case %d rep %d problem %d lens %s" % (case, rep, i, lens))
    finally:
        for k in _NMS_MODES[mode]:
            monkeypatch.delenv(k)
        rp.reload_env()
