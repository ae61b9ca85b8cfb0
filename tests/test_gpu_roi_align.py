"""GPU parity of ROIAlign fwd/bwd and the fused Pooler against the oracle, through the C ABI.

Tolerances (north_star: "ROIAlign within 1e-5 relative in fp32"):
  mode 1 (reference operand order): bit-exact against the oracle / the reference-text fixture;
  mode 0 (tiled, separable order):  |a-b| <= 1e-5*|b| + 1e-5*S, S = the magnitude scale of the
                                    data (1 for unit-variance features, max|grad| for bwd).
"""
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


def _close(a, b, scale=1.0):
    np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-5 * scale)


def _rois(rng, K, N):
    ctr = rng.uniform(0, 1, (K, 2)) * [1344, 800]
    wh = np.exp(rng.uniform(np.log(2), np.log(1200), (K, 2)))
    r = np.concatenate([rng.integers(0, N, (K, 1)), ctr - wh / 2, ctr + wh / 2], 1).astype(np.float32)
    r[:6, 1:] = [[10, 20, 300, 400], [0, 0, 1343, 799], [50, 50, 50, 50], [-40, -30, 100, 90],
                 [1200, 700, 1500, 900], [-500, -500, -300, -300]]
    return r


def test_reference_text_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "roi_align_ref.npz"))
    feat, rois = g["feat"], g["rois"]
    for tag in ("p7_s2", "p14_s2", "p14_s0", "p3x5_s3"):
        ph, pw, scale, sr = g[tag + "_cfg"]
        P = (int(ph), int(pw))
        exact = rp.roi_align_forward(T(feat), T(rois), P, float(scale), float(sr), mode=1).cpu().numpy()
        np.testing.assert_array_equal(exact, g[tag + "_out"])
        fast = rp.roi_align_forward(T(feat), T(rois), P, float(scale), float(sr), mode=0).cpu().numpy()
        _close(fast, g[tag + "_out"])
        for mode in (0, 1):
            gin = rp.roi_align_backward(T(g[tag + "_dout"]), T(rois), feat.shape, float(scale),
                                        float(sr), mode=mode).cpu().numpy()
            _close(gin, g[tag + "_gin"], scale=np.abs(g[tag + "_gin"]).max())


@pytest.mark.parametrize("P,sr", [((7, 7), 2), ((14, 14), 2), ((14, 14), 0), ((7, 7), 0),
                                  ((3, 5), 3), ((14, 14), 1), ((28, 28), 2)])
@pytest.mark.parametrize("C", [37, 64])
def test_single_level(P, sr, C):
    rng = np.random.default_rng(0)
    N = 2
    feat = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    rois = _rois(rng, 200, N)
    a = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
    exact = rp.roi_align_forward(T(feat), T(rois), P, 1 / 16, sr, mode=1).cpu().numpy()
    np.testing.assert_array_equal(exact, a)
    fast = rp.roi_align_forward(T(feat), T(rois), P, 1 / 16, sr, mode=0).cpu().numpy()
    _close(fast, a)
    d = rng.standard_normal(a.shape).astype(np.float32)
    ga = oracle.roi_align_backward(d, rois, feat.shape, 1 / 16, sr)
    for mode in (0, 1):
        gb = rp.roi_align_backward(T(d), T(rois), feat.shape, 1 / 16, sr, mode=mode).cpu().numpy()
        _close(gb, ga, scale=np.abs(ga).max())


def test_oversize_rois_take_the_direct_path():
    # P2-scale map with image-sized RoIs: tiles of up to 200x336 do not fit shared memory
    rng = np.random.default_rng(4)
    feat = rng.standard_normal((1, 8, 200, 336)).astype(np.float32)
    rois = np.array([[0, 0, 0, 1343, 799], [0, 100, 50, 1200, 700], [0, 3, 3, 40, 40]], np.float32)
    for sr in (2, 0):
        a = oracle.roi_align_forward(feat, rois, (14, 14), 0.25, sr)
        b = rp.roi_align_forward(T(feat), T(rois), (14, 14), 0.25, sr).cpu().numpy()
        _close(b, a)
        d = rng.standard_normal(a.shape).astype(np.float32)
        ga = oracle.roi_align_backward(d, rois, feat.shape, 0.25, sr)
        gb = rp.roi_align_backward(T(d), T(rois), feat.shape, 0.25, sr).cpu().numpy()
        _close(gb, ga, scale=np.abs(ga).max())


def test_empty_and_invalid():
    feat = torch.randn(1, 8, 20, 30, device=DEV)
    out = rp.roi_align(feat, torch.zeros((0, 5), device=DEV), (7, 7), 0.25, 2)
    assert tuple(out.shape) == (0, 8, 7, 7)
    # batch index out of range -> zeros, no fault
    rois = torch.tensor([[3, 0, 0, 10, 10.0], [0, 0, 0, 10, 10]], device=DEV)
    out = rp.roi_align(feat, rois, (7, 7), 0.25, 2)
    assert float(out[0].abs().max()) == 0.0 and float(out[1].abs().max()) > 0.0


def test_pooler_matches_oracle_and_keeps_roi_order():
    N, C = 2, 32
    feats = synth.fpn_features(N, C)
    rois = synth.rois_for_images(N, 300)
    boxes = [rp.BoxList(T(rois[rois[:, 0] == i][:, 1:]), (synth.IMG_W, synth.IMG_H)) for i in range(N)]
    lv_o = oracle.level_map(rois[:, 1:], 2, 5)
    for res in (7, 14):
        pooler = rp.Pooler((res, res), synth.FPN_SCALES, 2)
        lv = pooler.map_levels(boxes).cpu().numpy()
        np.testing.assert_array_equal(lv, lv_o)
        assert len(np.unique(lv)) == 4
        a = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, lv_o, (res, res), 2)
        xs = [T(f).requires_grad_(True) for f in feats]
        out = pooler(xs + [torch.zeros(1, device=DEV)], boxes)  # a 5th map (P6) is ignored
        _close(out.detach().cpu().numpy(), a)
        d = np.random.default_rng(5).standard_normal(a.shape).astype(np.float32)
        out.backward(T(d))
        ga = oracle.pooler_backward(d, rois, lv_o, [f.shape for f in feats], synth.FPN_SCALES, 2)
        for x, g in zip(xs, ga):
            _close(x.grad.cpu().numpy(), g, scale=max(1.0, np.abs(g).max()))


def test_level_mapper_with_area():
    rois = synth.rois_for_images(2, 200)
    boxes = [rp.BoxList(T(rois[rois[:, 0] == i][:, 1:]), (synth.IMG_W, synth.IMG_H)) for i in range(2)]
    lv = rp.LevelMapperwithArea(2, 5)(boxes).cpu().numpy()
    area = np.full((rois.shape[0],), float(synth.IMG_W * synth.IMG_H), np.float32)
    np.testing.assert_array_equal(lv, oracle.level_map_area(rois[:, 1:], area, 2, 5))


def test_single_level_pooler_c4():
    rng = np.random.default_rng(6)
    feat = rng.standard_normal((2, 48, 50, 84)).astype(np.float32)
    rois = synth.rois_for_images(2, 100)
    boxes = [rp.BoxList(T(rois[rois[:, 0] == i][:, 1:]), (synth.IMG_W, synth.IMG_H)) for i in range(2)]
    pooler = rp.Pooler((14, 14), (1 / 16,), 0)
    out = pooler([T(feat)], boxes).cpu().numpy()
    _close(out, oracle.roi_align_forward(feat, rois, (14, 14), 1 / 16, 0))


def test_full_size_properties():
    # BASELINE cfg 3 geometry (1000 RoIs/image, 256 channels, 14x14) on 2 images: constants are
    # preserved where every sample is inside the map, and the operator is linear.
    N, C = 2, 256
    rois = T(synth.rois_for_images(N, 1000))
    levels = rp.LevelMapper(2, 5).map_rois(rois)
    shapes = synth.fpn_shapes(N, C)
    g = torch.Generator(device=DEV).manual_seed(1)
    f1 = [torch.randn(s, device=DEV, generator=g) for s in shapes]
    f2 = [torch.randn(s, device=DEV, generator=g) for s in shapes]
    o1 = rp.pooler_forward(f1, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    o2 = rp.pooler_forward(f2, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    o12 = rp.pooler_forward([a + 2 * b for a, b in zip(f1, f2)], synth.FPN_SCALES, rois, levels, (14, 14), 2)
    assert torch.allclose(o12, o1 + 2 * o2, rtol=1e-4, atol=1e-4)
    ones = [torch.full(s, 3.0, device=DEV) for s in shapes]
    oc = rp.pooler_forward(ones, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    assert torch.allclose(oc, torch.full_like(oc, 3.0), rtol=1e-6)
    # tiled and reference-order kernels agree at full size
    oe = rp.pooler_forward(f1, synth.FPN_SCALES, rois, levels, (14, 14), 2, mode=1)
    assert torch.allclose(o1, oe, rtol=1e-5, atol=1e-5)
    # adjoint identity <P f, d> == <f, P^T d>
    d = torch.randn(o1.shape, device=DEV, generator=g)
    grads = rp.pooler_backward(d, rois, levels, shapes, synth.FPN_SCALES, 2)
    lhs = float((o1.double() * d.double()).sum())
    rhs = float(sum((a.double() * b.double()).sum() for a, b in zip(f1, grads)))
    assert abs(lhs - rhs) <= 1e-5 * max(abs(lhs), abs(rhs), 1.0) * 10


@pytest.mark.parametrize("impl", ["3", "4", "5", "6", "7"])
@pytest.mark.parametrize("P,sr", [((14, 14), 2), ((7, 7), 2), ((14, 14), 0), ((3, 5), 3)])
def test_forward_kernel_variants(impl, P, sr, monkeypatch):
    """The forward kernels (RP_ROI_IMPL=3: column walk; =4: banded, v3 body as its fallback; =5:
    merged taps, reference-order fallback; =6: tensor-core kernel for whole 8-channel groups and a fixed
    sampling grid, the SIMT kernel otherwise; =7: two chained warp-level matrix products, column walk for
    tiles above 48 rows / columns; adaptive sampling always takes the column walk except in 7)
    against the oracle, on aligned (W % 4 == 0), 8-byte (W = 42) and odd-width maps."""
    monkeypatch.setenv("RP_ROI_IMPL", impl)
    rp.reload_env()
    rng = np.random.default_rng(11)
    for (H, W), C, cpc in [((50, 84), 37, "16"), ((25, 42), 8, "64"), ((13, 19), 5, "64"), ((3, 9), 4, "64")]:
        monkeypatch.setenv("RP_ROI_CPC", cpc)
        rp.reload_env()
        feat = rng.standard_normal((2, C, H, W)).astype(np.float32)
        rois = _rois(rng, 90, 2)
        rois[:, 1:] *= np.float32(W * 16 / 1344)
        exp = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
        out = rp.roi_align(T(feat), T(rois), P, 1 / 16, sr).cpu().numpy()
        _close(out, exp)


@pytest.mark.parametrize("impl", ["4", "5", "6", "7"])
def test_banded_pooler_multi_level(impl, monkeypatch):
    monkeypatch.setenv("RP_ROI_IMPL", impl)
    rp.reload_env()
    N, C = 2, 24
    feats = synth.fpn_features(N, C)
    rois = synth.rois_for_images(N, 200)
    lv = oracle.level_map(rois[:, 1:], 2, 5)
    for P in [(14, 14), (7, 7)]:
        exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, lv, P, 2)
        out = rp.pooler_forward([T(f) for f in feats], synth.FPN_SCALES, T(rois), T(lv), P, 2)
        _close(out.cpu().numpy(), exp)


def test_backward_kernels_agree_with_oracle_on_wide_bins(monkeypatch):
    """Gather backward (default) and the sweep kernel on RoIs whose bins are wider than the
    short-tap form allows (reference-order fallback inside the gather kernel), plus adaptive
    sampling (sweep kernel in a second launch)."""
    rng = np.random.default_rng(3)
    feat_shape = (2, 6, 60, 90)
    rois = _rois(rng, 40, 2)
    rois[:8, 1:] = np.array([0.0, 0.0, 1400.0, 940.0], np.float32)  # 87 x 59 px at 1/16: 29 px bins at 3x2
    for P, sr in [((2, 3), 2), ((7, 7), 2), ((7, 7), 0)]:
        d = rng.standard_normal((rois.shape[0], feat_shape[1]) + P).astype(np.float32)
        exp = oracle.roi_align_backward(d, rois, feat_shape, 1 / 16, sr)
        for impl in ("2", "1"):
            monkeypatch.setenv("RP_ROI_BWD_IMPL", impl)
            rp.reload_env()
            got = rp.roi_align_backward(T(d), T(rois), feat_shape, 1 / 16, sr).cpu().numpy()
            np.testing.assert_allclose(got, exp, rtol=1e-5, atol=1e-5 * max(1.0, np.abs(exp).max()))


def test_pooler_execute_host_pipeline():
    """Host-buffer entry (three-stream pipeline over images) == the resident-tensor call."""
    N, C = 5, 8
    feats = synth.fpn_features(N, C)
    rois = synth.rois_for_images(N, 40)
    pooler = rp.Pooler((7, 7), synth.FPN_SCALES, 2)
    h_feats = [torch.from_numpy(f).pin_memory() for f in feats]
    h_rois = torch.from_numpy(rois)
    boxes_host = [h_rois[h_rois[:, 0] == i][:, 1:].contiguous() for i in range(N)]
    boxes_host[2] = boxes_host[2][:0]  # an image without proposals
    out = pooler.execute_host(h_feats, boxes_host)
    bl = [rp.BoxList(b.to(DEV), (synth.IMG_W, synth.IMG_H)) for b in boxes_host]
    with torch.no_grad():
        ref = pooler([f.to(DEV) for f in h_feats], bl)
    assert out.shape == ref.shape
    np.testing.assert_array_equal(out.numpy(), ref.cpu().numpy())
    # second call reuses the staging buffers
    out2 = pooler.execute_host(h_feats, boxes_host)
    np.testing.assert_array_equal(out2.numpy(), ref.cpu().numpy())


@pytest.mark.parametrize("P,sr", [((14, 14), 2), ((7, 7), 2), ((14, 14), 0), ((3, 5), 3)])
@pytest.mark.parametrize("C", [37, 64, 8, 100])
def test_channels_last_entry_matches_oracle(P, sr, C):
    # torch.channels_last feature maps take the NHWC kernels; same oracle, same bar.  C % 4 == 0 with
    # sampling_ratio 2 is the separable kernel (the RoIs span tiles that fit 32, 16, 8 and 4 channels
    # per chunk and tiles that do not fit at all: reference-order path; C = 100: a last chunk of 4
    # channels; pooled width 7: 8 column lanes), everything else the gather kernel (16-byte taps)
    rng = np.random.default_rng(2)
    N = 2
    feat = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    rois = _rois(rng, 150, N)
    exp = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
    x = T(feat).contiguous(memory_format=torch.channels_last)
    assert not x.is_contiguous()
    got = rp.roi_align_forward(x, T(rois), P, 1 / 16, sr).cpu().numpy()
    _close(got, exp)


def test_channels_last_pooler_matches_oracle():
    N, C = 2, 40
    feats = synth.fpn_features(N, C)
    rois = synth.rois_for_images(N, 200)
    lv = oracle.level_map(rois[:, 1:], 2, 5)
    exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, lv, (14, 14), 2)
    xs = [T(f).contiguous(memory_format=torch.channels_last) for f in feats]
    got = rp.pooler_forward(xs, synth.FPN_SCALES, T(rois), T(lv), (14, 14), 2).cpu().numpy()
    _close(got, exp)


def test_channels_last_tma_variant_matches_oracle(monkeypatch):
    # RP_NHWC_TMA=1: the tile of every chunk arrives by cp.async.bulk.tensor boxes into two buffers
    monkeypatch.setenv("RP_NHWC_TMA", "1")
    rp.reload_env()
    try:
        N, C = 2, 72
        feats = synth.fpn_features(N, C)
        rois = synth.rois_for_images(N, 300)
        lv = oracle.level_map(rois[:, 1:], 2, 5)
        xs = [T(f).contiguous(memory_format=torch.channels_last) for f in feats]
        for P in ((14, 14), (7, 7)):
            exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, lv, P, 2)
            got = rp.pooler_forward(xs, synth.FPN_SCALES, T(rois), T(lv), P, 2).cpu().numpy()
            _close(got, exp)
        rng = np.random.default_rng(4)
        feat = rng.standard_normal((N, 64, 50, 84)).astype(np.float32)
        r2 = _rois(rng, 150, N)  # includes RoIs wider than every box width and outside the map
        exp = oracle.roi_align_forward(feat, r2, (14, 14), 1 / 16, 2)
        x = T(feat).contiguous(memory_format=torch.channels_last)
        _close(rp.roi_align_forward(x, T(r2), (14, 14), 1 / 16, 2).cpu().numpy(), exp)
    finally:
        monkeypatch.delenv("RP_NHWC_TMA")
        rp.reload_env()


@pytest.mark.parametrize("mode", ["0", "1"])
def test_chained_mma_forward_modes(mode, monkeypatch):
    """RP_ROI_IMPL=7 (roi_fwd_mma.cuh) in both precision modes (0: tf32 + one bf16 correction product,
    1: 3xTF32), with channel slices that leave warps with uneven channel counts, on every tile class
    (1..6 blocks per axis, wider tiles take the column walk) and with adaptive sampling."""
    monkeypatch.setenv("RP_ROI_IMPL", "7")
    monkeypatch.setenv("RP_ROI_MMA_MODE", mode)
    rng = np.random.default_rng(5)
    for cpc, C, (H, W) in [("8", 13, (100, 168)), ("24", 37, (50, 84)), ("256", 20, (25, 42))]:
        monkeypatch.setenv("RP_ROI_MMA_CPC", cpc)
        rp.reload_env()
        feat = rng.standard_normal((2, C, H, W)).astype(np.float32)
        K = 120
        rois = np.zeros((K, 5), np.float32)
        rois[:, 0] = rng.integers(0, 2, K)
        side = np.exp(rng.uniform(np.log(4), np.log(60 * 16), (K, 2)))  # 0.25 .. 60 feature pixels
        ctr = rng.uniform(0, 1, (K, 2)) * [W * 16, H * 16]
        rois[:, 1:3] = ctr - side / 2
        rois[:, 3:5] = ctr + side / 2
        for P, sr in [((14, 14), 2), ((7, 7), 2), ((16, 16), 1), ((14, 14), 0), ((5, 9), 3)]:
            exp = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
            out = rp.roi_align(T(feat), T(rois), P, 1 / 16, sr).cpu().numpy()
            _close(out, exp)


@pytest.mark.parametrize("knobs", [{"RP_ROI_PLANES": "0"}, {"RP_ROI_TAP3": "0"}, {"RP_ROI_YMERGE": "0"},
                                   {"RP_ROI_PLANES": "0", "RP_ROI_TAP3": "0", "RP_ROI_YMERGE": "0"},
                                   {"RP_ROI_DBUF": "1"}, {"RP_ROI_DBUF": "1", "RP_ROI_PLANES": "0"}])
def test_default_forward_with_each_step_switched_off(knobs, monkeypatch):
    """The default forward call = column walk with three-tap rows and merged y-samples + the whole-plane
    kernel on the coarse levels.  Every step has a switch; with any of them off (and with the opt-in
    double-buffered chunk loop on) the call must still match the oracle: on the FPN pyramid (plane kernel
    takes P4 / P5 there), on one image (its RoIs are dealt to several CTAs) and on RoIs with bins wider
    than 2 feature pixels (no three-tap rows), 14x14 and 7x7."""
    for k, v in knobs.items():
        monkeypatch.setenv(k, v)
    rp.reload_env()
    for N, per in [(3, 150), (1, 300)]:
        C = 20
        feats = synth.fpn_features(N, C)
        rois = synth.rois_for_images(N, per)
        rois[::7, 3] += 300.0  # wide bins on every level
        rois[::11, 4] += 250.0
        lv = oracle.level_map(rois[:, 1:], 2, 5)
        for P in [(14, 14), (7, 7)]:
            exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, lv, P, 2)
            out = rp.pooler_forward([T(f) for f in feats], synth.FPN_SCALES, T(rois), T(lv), P, 2)
            _close(out.cpu().numpy(), exp)
