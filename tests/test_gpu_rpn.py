"""GPU parity of the RPN proposal stage (BoxCoder.decode, sigmoid + top-k + gather + decode +
clip + small-box removal, chained into the batched NMS) against the oracle: boxes, scores and
indices bit-exact, including the tie order of equal scores."""
import os

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


def test_box_decode_reference_kat(golden_dir):
    # detectron.jittor/tests/test_box_coder.py:11-108
    k = np.load(os.path.join(golden_dir, "box_coder_kat.npz"))
    coder = rp.BoxCoder(weights=(1.0, 1.0, 1.0, 1.0))
    out = coder.decode(T(k["deltas"]), T(k["bbox"])).cpu().numpy()
    np.testing.assert_allclose(out, k["gt_bbox"], atol=float(k["atol"]))
    np.testing.assert_array_equal(out, oracle.box_decode(k["deltas"], k["bbox"]))


@pytest.mark.parametrize("n,K", [(1, 1), (255, 1), (1000, 81), (4097, 2)])
def test_box_decode_matches_oracle(n, K):
    rng = np.random.default_rng(n + K)
    x1 = rng.uniform(0, 900, n); y1 = rng.uniform(0, 600, n)
    boxes = np.stack([x1, y1, x1 + rng.uniform(0, 400, n), y1 + rng.uniform(0, 300, n)], 1).astype(np.float32)
    codes = (rng.standard_normal((n, 4 * K)) * 2.0).astype(np.float32)
    codes[::7, 2] = 50.0
    w = (10.0, 10.0, 5.0, 5.0)
    out = rp.BoxCoder(w).decode(T(codes), T(boxes)).cpu().numpy()
    np.testing.assert_array_equal(out, oracle.box_decode(codes, boxes, w))
    assert rp.box_decode(T(codes[:0]), T(boxes[:0])).shape == (0, 4 * K)


CASES = [  # N, A, H, W, stride, top_n, min_size, ties
    (1, 3, 5, 7, 64, 1000, 0.0, False),       # fewer anchors than top_n
    (2, 3, 25, 42, 32, 1000, 0.0, False),
    (2, 3, 50, 84, 16, 1000, 8.0, True),
    (3, 3, 100, 168, 8, 2000, 4.0, True),
    (2, 1, 13, 9, 16, 50, 16.0, True),
    (2, 5, 31, 17, 16, 257, 2.0, False),      # A not 3, k not a multiple of the CTA width
    (1, 3, 200, 336, 4, 6000, 0.0, False),    # the P2 level of a 800 x 1344 image, training top_n 2000+
]


@pytest.mark.parametrize("case", CASES)
def test_pre_nms_matches_oracle(case):
    N, A, H, W, stride, top_n, min_size, ties = case
    anchors, obj, reg, im_wh = synth.rpn_level(N, A, H, W, stride, seed=H * W + A, ties=ties)
    boxes, scores, index, offsets, k = rp.rpn_pre_nms(T(obj), T(reg), T(anchors), T(im_wh), top_n, min_size)
    assert k == min(top_n, A * H * W)
    off = offsets.cpu().numpy()
    ref = oracle.rpn_pre_nms(obj, reg, anchors, im_wh, top_n, min_size)
    assert off[0] == 0
    bx, sc, ix = boxes.cpu().numpy(), scores.cpu().numpy(), index.cpu().numpy()
    for b in range(N):
        rb, rs, ri = ref[b]
        assert off[b + 1] - off[b] == len(ri)
        sl = slice(off[b], off[b + 1])
        np.testing.assert_array_equal(ix[sl], ri)
        np.testing.assert_array_equal(sc[sl], rs)
        np.testing.assert_array_equal(bx[sl], rb)
    if min_size > 0:
        assert off[N] < N * k  # the filter removed something: the packed layout is exercised


def test_pre_nms_levels_matches_oracle():
    N, A = 3, 3
    shapes = [(50, 84, 16), (25, 42, 32), (13, 21, 64), (2, 3, 128)]
    lv = [synth.rpn_level(N, A, H, W, s, seed=40 + s, ties=(s == 32)) for H, W, s in shapes]
    im_wh = lv[0][3]
    boxes, scores, index, offsets, ks = rp.rpn_pre_nms_levels(
        [T(l[1]) for l in lv], [T(l[2]) for l in lv], [T(l[0]) for l in lv], T(im_wh), 600, 6.0)
    assert ks == [600, 600, 600, 18]
    off = offsets.cpu().numpy()
    bx, sc, ix = boxes.cpu().numpy(), scores.cpu().numpy(), index.cpu().numpy()
    assert bx.shape[0] == N * sum(ks) and off[0] == 0
    for l, lvl in enumerate(lv):
        ref = oracle.rpn_pre_nms(lvl[1], lvl[2], lvl[0], im_wh, 600, 6.0)
        for b in range(N):
            sl = slice(off[l * N + b], off[l * N + b + 1])
            np.testing.assert_array_equal(ix[sl], ref[b][2])
            np.testing.assert_array_equal(sc[sl], ref[b][1])
            np.testing.assert_array_equal(bx[sl], ref[b][0])


def test_select_levels_matches_host_branch():
    # rp_rpn_select_levels against the BoxList form of inference.py:175-183, tie-heavy scores
    N, A = 3, 3
    shapes = [(25, 42, 32), (13, 21, 64), (7, 11, 128)]
    lv = [synth.rpn_level(N, A, H, W, s, seed=60 + s, ties=True) for H, W, s in shapes]
    im_wh = lv[0][3]
    anchors = [[rp.BoxList(T(l[0][b]), (int(im_wh[b, 0]), int(im_wh[b, 1]))) for l in lv] for b in range(N)]
    objs, regs = [T(l[1]) for l in lv], [T(l[2]) for l in lv]
    for top in (50, 400, 5000):
        post = rp.RPNPostProcessor(1000, 300, 0.7, 0, fpn_post_nms_top_n=top)
        post.eval()
        fused = post(anchors, objs, regs)
        st = post._launch(list(zip(*anchors)), objs, regs)
        host = post.select_over_all_levels([rp.cat_boxlist(bl) for bl in zip(*post._collect(st))])
        for f, h in zip(fused, host):
            np.testing.assert_array_equal(f.bbox.cpu().numpy(), h.bbox.cpu().numpy())
            np.testing.assert_array_equal(f.get_field("objectness").cpu().numpy(),
                                          h.get_field("objectness").cpu().numpy())


def _anchor_lists(anchors, im_wh):
    return [rp.BoxList(T(anchors[b]), (int(im_wh[b, 0]), int(im_wh[b, 1])), mode="xyxy")
            for b in range(anchors.shape[0])]


def _oracle_level(obj, reg, anchors, im_wh, top_n, min_size, thr, post):
    out = []
    for rb, rs, _ in oracle.rpn_pre_nms(obj, reg, anchors, im_wh, top_n, min_size):
        keep = oracle.nms(rb, rs, thr, max_keep=post) if len(rb) else np.zeros((0,), np.int32)
        out.append((rb[keep], rs[keep]))
    return out


@pytest.mark.parametrize("ties", [False, True])
def test_single_feature_map_matches_oracle(ties):
    N, A, H, W, stride = 3, 3, 50, 84, 16
    anchors, obj, reg, im_wh = synth.rpn_level(N, A, H, W, stride, seed=11, ties=ties)
    post = rp.RPNPostProcessor(pre_nms_top_n=2000, post_nms_top_n=300, nms_thresh=0.7, min_size=4)
    res = post.forward_for_single_feature_map(_anchor_lists(anchors, im_wh), T(obj), T(reg))
    ref = _oracle_level(obj, reg, anchors, im_wh, 2000, 4.0, 0.7, 300)
    assert len(res) == N
    for b in range(N):
        assert res[b].size == (int(im_wh[b, 0]), int(im_wh[b, 1])) and res[b].mode == "xyxy"
        np.testing.assert_array_equal(res[b].bbox.cpu().numpy(), ref[b][0])
        np.testing.assert_array_equal(res[b].get_field("objectness").cpu().numpy(), ref[b][1])
        assert len(res[b]) <= 300


def test_all_levels_matches_oracle():
    # execute(): 5 FPN levels, image-major anchor lists, test-time per-image top-k over levels
    N, A = 2, 3
    shapes = [(200, 336, 4), (100, 168, 8), (50, 84, 16), (25, 42, 32), (13, 21, 64)]
    levels = [synth.rpn_level(N, A, H, W, s, seed=100 + s) for H, W, s in shapes]
    im_wh = levels[0][3]
    anchors = [[rp.BoxList(T(lv[0][b]), (int(im_wh[b, 0]), int(im_wh[b, 1]))) for lv in levels]
               for b in range(N)]
    post = rp.RPNPostProcessor(pre_nms_top_n=1000, post_nms_top_n=1000, nms_thresh=0.7, min_size=0,
                               fpn_post_nms_top_n=1000)
    post.eval()
    res = post(anchors, [T(lv[1]) for lv in levels], [T(lv[2]) for lv in levels])
    per_level = [_oracle_level(lv[1], lv[2], lv[0], lv[3], 1000, 0.0, 0.7, 1000) for lv in levels]
    for b in range(N):
        rb = np.concatenate([pl[b][0] for pl in per_level], 0)
        rs = np.concatenate([pl[b][1] for pl in per_level], 0)
        order = np.argsort(-rs, kind="stable")[:1000]
        np.testing.assert_array_equal(res[b].bbox.cpu().numpy(), rb[order])
        np.testing.assert_array_equal(res[b].get_field("objectness").cpu().numpy(), rs[order])


def test_training_mode_adds_gt_and_selects_per_batch():
    N, A, H, W, stride = 2, 3, 25, 42, 32
    lv = [synth.rpn_level(N, A, H, W, stride, seed=5), synth.rpn_level(N, A, 13, 21, 64, seed=6)]
    im_wh = lv[0][3]
    sizes = [(int(im_wh[b, 0]), int(im_wh[b, 1])) for b in range(N)]
    anchors = [[rp.BoxList(T(l[0][b]), sizes[b]) for l in lv] for b in range(N)]
    targets = [rp.BoxList(T(synth.proposals(3, 50 + b, adversarial=False)), sizes[b]) for b in range(N)]
    post = rp.RPNPostProcessor(2000, 2000, 0.7, 0, fpn_post_nms_top_n=500, fpn_post_nms_per_batch=True)
    post.train()
    res = post(anchors, [T(l[1]) for l in lv], [T(l[2]) for l in lv], targets)
    assert sum(len(r) for r in res) == 500 + 3 * N
    for b in range(N):
        obj = res[b].get_field("objectness").cpu().numpy()
        assert np.all(obj[-3:] == 1.0)
        np.testing.assert_array_equal(res[b].bbox[-3:].cpu().numpy(), targets[b].bbox.cpu().numpy())


def test_errors_are_loud():
    anchors, obj, reg, im_wh = synth.rpn_level(1, 3, 5, 7, 64, seed=1)
    with pytest.raises(ValueError):
        rp.rpn_pre_nms(T(obj), T(reg[:, :8]), T(anchors), T(im_wh), 100, 0)
    with pytest.raises(Exception):
        rp.rpn_pre_nms(obj, reg, anchors, im_wh, 100, 0)  # numpy / CPU input: no fallback


def test_sort_path_and_select_path_agree(monkeypatch):
    # RP_RPN_SELECT=0 keeps the whole-batch radix sort; the default per-problem radix select must
    # give the same candidates in the same order (massive ties: logits quantised to 1/4)
    N, A, H, W, stride = 2, 3, 100, 168, 8
    anchors, obj, reg, im_wh = synth.rpn_level(N, A, H, W, stride, seed=5, ties=True)
    obj[0, :, :40] = 3.0  # one huge plateau of equal scores: the cut falls inside it
    outs = []
    for sel in ("1", "0"):
        monkeypatch.setenv("RP_RPN_SELECT", sel)
        rp.reload_env()
        b, s, i, o, k = rp.rpn_pre_nms(T(obj), T(reg), T(anchors), T(im_wh), 1000, 0.0)
        outs.append((b.cpu().numpy(), s.cpu().numpy(), i.cpu().numpy(), o.cpu().numpy()))
    monkeypatch.delenv("RP_RPN_SELECT")
    rp.reload_env()
    for x, y in zip(*outs):
        np.testing.assert_array_equal(x, y)
    ref = oracle.rpn_pre_nms(obj, reg, anchors, im_wh, 1000, 0.0)
    np.testing.assert_array_equal(outs[0][2][: len(ref[0][2])], ref[0][2])


def test_anchors_are_not_cached_by_address():
    # a landscape batch followed by a portrait batch with the same anchor count must be
    # decoded against its own anchors (the reference builds new anchor tensors on every forward)
    post = rp.RPNPostProcessor(pre_nms_top_n=200, post_nms_top_n=50, nms_thresh=0.7, min_size=0)
    res = []
    for (H, W, seed) in ((20, 30, 1), (30, 20, 2)):
        anchors, obj, reg, im_wh = synth.rpn_level(1, 3, H, W, 16, seed=seed)
        a_t = T(anchors)  # same size both times: the allocator hands back the same address
        out = post.forward_for_single_feature_map(_anchor_lists(anchors, im_wh), T(obj), T(reg))
        ref = _oracle_level(obj, reg, anchors, im_wh, 200, 0.0, 0.7, 50)
        np.testing.assert_array_equal(out[0].bbox.cpu().numpy(), ref[0][0])
        res.append(out)
        del a_t


def test_select_levels_one_launch_kernel_rounds_and_radix_path(monkeypatch):
    # rp_rpn_select_levels: per image the survivors of all levels, stable-sorted by descending score,
    # first top_n.  20 000 survivors per image (more than the 8192 the one-launch kernel sorts at a time:
    # it works in rounds), tie-heavy scores; RP_RPN_SEL_FUSED=0 is the whole-batch radix sort.
    rng = np.random.default_rng(5)
    L, N, per = 4, 3, 6000
    S = L * N
    lens = rng.integers(per - 2000, per + 1, S)
    lens[5] = 0
    offs = np.concatenate([[0], np.cumsum(np.full(S, per))]).astype(np.int32)  # padded segments
    n_tot = int(offs[-1])
    boxes = rng.uniform(0, 500, (n_tot, 4)).astype(np.float32)
    scores = (np.floor(rng.uniform(0, 1, n_tot) * 512) / 512).astype(np.float32)
    keep = np.zeros(n_tot, np.int32)
    n_keep = np.zeros(S, np.int32)
    for s in range(S):
        k = rng.permutation(lens[s])[: max(0, lens[s] - 100)].astype(np.int32)
        keep[offs[s]: offs[s] + len(k)] = k
        n_keep[s] = len(k)
    for top in (1000, 2000):
        ref_b, ref_s = [], []
        for b in range(N):
            src = np.concatenate([offs[l * N + b] + keep[offs[l * N + b]: offs[l * N + b] + n_keep[l * N + b]]
                                  for l in range(L)])
            order = np.argsort(-scores[src], kind="stable")[:top]
            ref_b.append(boxes[src[order]])
            ref_s.append(scores[src[order]])
        for fused in ("1", "0"):
            monkeypatch.setenv("RP_RPN_SEL_FUSED", fused)
            rp.reload_env()
            try:
                ob, osc, cnt = rp.rpn_select_levels(T(boxes), T(scores), T(offs), T(keep), T(n_keep), L, N, top)
            finally:
                monkeypatch.delenv("RP_RPN_SEL_FUSED")
                rp.reload_env()
            ob, osc, cnt = ob.cpu().numpy(), osc.cpu().numpy(), cnt.cpu().numpy()
            for b in range(N):
                assert cnt[b] == len(ref_s[b])
                np.testing.assert_array_equal(osc[b, : cnt[b]], ref_s[b])
                np.testing.assert_array_equal(ob[b, : cnt[b]], ref_b[b])
