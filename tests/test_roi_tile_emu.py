"""The tiled ROIAlign algorithm (plans + sweeps of csrc/roi_tile.h, the code the CUDA kernels
run) emulated on the host against the oracle.  Tolerance: the north_star's 1e-5 relative,
applied as |a-b| <= 1e-5 * (|b| + 1) for unit-variance features (the separable evaluation
order differs from the reference's per-sample order; observed error ~6e-7)."""
import numpy as np
import pytest

import oracle
from tests import emu, synth


def _close(a, b, scale=1.0):
    np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-5 * scale)


def _rois(rng, K, N):
    ctr = rng.uniform(0, 1, (K, 2)) * [1344, 800]
    wh = np.exp(rng.uniform(np.log(2), np.log(1200), (K, 2)))
    r = np.concatenate([rng.integers(0, N, (K, 1)), ctr - wh / 2, ctr + wh / 2], 1).astype(np.float32)
    r[:6, 1:] = [[10, 20, 300, 400], [0, 0, 1343, 799], [50, 50, 50, 50], [-40, -30, 100, 90],
                 [1200, 700, 1500, 900], [-500, -500, -300, -300]]
    return r


@pytest.mark.parametrize("P,sr", [((7, 7), 2), ((14, 14), 2), ((14, 14), 0), ((7, 7), 0),
                                  ((3, 5), 3), ((14, 14), 1)])
def test_single_level_forward_backward(P, sr):
    rng = np.random.default_rng(0)
    N, C = 2, 37  # C not a multiple of 4: exercises the ragged channel tail
    feat = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    rois = _rois(rng, 120, N)
    a = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
    d = rng.standard_normal(a.shape).astype(np.float32)
    ga = oracle.roi_align_backward(d, rois, feat.shape, 1 / 16, sr)
    for vs, cpc in [(8192, None), (2048, 16), (600, 12)]:
        b, _ = emu.pooler_forward([feat], [1 / 16], rois, None, P, sr, vs, cpc)
        _close(b, a)
        gb, _ = emu.pooler_backward(d, [feat.shape], [1 / 16], rois, None, sr, vs, cpc)
        _close(gb[0], ga, scale=np.abs(ga).max())


def test_direct_path_is_bit_exact():
    rng = np.random.default_rng(1)
    feat = rng.standard_normal((1, 8, 50, 84)).astype(np.float32)
    rois = _rois(rng, 40, 1)
    a = oracle.roi_align_forward(feat, rois, (7, 7), 1 / 16, 2)
    b, nd = emu.pooler_forward([feat], [1 / 16], rois, None, (7, 7), 2, vs_floats=2, c_per_cta=8)
    assert nd == 40 - 1  # everything but the RoI that lies entirely outside the map
    np.testing.assert_array_equal(a, b)


def test_multi_level_pooler():
    rng = np.random.default_rng(2)
    N, C = 2, 16
    shapes = [(N, C, 200, 336), (N, C, 100, 168), (N, C, 50, 84), (N, C, 25, 42)]
    feats = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    rois = synth.rois_for_images(N, 150)
    levels = oracle.level_map(rois[:, 1:], 2, 5)
    assert set(np.unique(levels)) == {0, 1, 2, 3}
    a = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    b, nd = emu.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    assert nd == 0
    _close(b, a)
    d = rng.standard_normal(a.shape).astype(np.float32)
    ga = oracle.pooler_backward(d, rois, levels, shapes, synth.FPN_SCALES, 2)
    gb, _ = emu.pooler_backward(d, shapes, synth.FPN_SCALES, rois, levels, 2)
    for x, y in zip(gb, ga):
        _close(x, y, scale=max(1.0, np.abs(y).max()))


def test_linearity_and_constant_preservation():
    # size-independent properties: pooling a constant map gives the constant wherever all
    # samples are inside the map; the operator is linear in the features
    rng = np.random.default_rng(3)
    N, C = 1, 8
    f1 = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    f2 = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    rois = synth.rois_for_images(1, 64)
    o1, _ = emu.pooler_forward([f1], [1 / 16], rois, None, (7, 7), 2)
    o2, _ = emu.pooler_forward([f2], [1 / 16], rois, None, (7, 7), 2)
    o12, _ = emu.pooler_forward([f1 + 2 * f2], [1 / 16], rois, None, (7, 7), 2)
    np.testing.assert_allclose(o12, o1 + 2 * o2, rtol=1e-4, atol=1e-5)
    ones = np.full((N, C, 50, 84), 3.0, np.float32)
    oc, _ = emu.pooler_forward([ones], [1 / 16], rois, None, (7, 7), 2)
    np.testing.assert_allclose(oc, 3.0, rtol=1e-6)


# ---- banded (v4) forward: same plan / walk code as rb_pool_fwd_kernel -----------------------------
@pytest.mark.parametrize("P,sr", [((7, 7), 2), ((14, 14), 2), ((14, 14), 0), ((3, 5), 3), ((14, 14), 1),
                                  ((28, 28), 2)])
def test_band_forward_single_level(P, sr):
    rng = np.random.default_rng(5)
    N, C = 2, 21
    feat = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
    rois = _rois(rng, 120, N)
    a = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
    for smem, cpc, nt in [(10240, None, 128), (2600, 8, 128), (1200, 5, 64)]:
        b, nf = emu.band_forward([feat], [1 / 16], rois, None, P, sr, smem, cpc, nt)
        _close(b, a)
        assert nf < 120 * (1 if cpc is None else -(-C // cpc))  # the band path is exercised


def test_band_forward_multi_level_and_odd_maps():
    rng = np.random.default_rng(6)
    N, C = 2, 16
    shapes = [(N, C, 200, 336), (N, C, 100, 168), (N, C, 50, 84), (N, C, 25, 42)]
    feats = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    rois = synth.rois_for_images(N, 150)
    levels = oracle.level_map(rois[:, 1:], 2, 5)
    a = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    b, nf = emu.band_forward(feats, synth.FPN_SCALES, rois, levels, (14, 14), 2)
    assert nf == 0  # every FPN-mapped RoI has a band form at 14 bins, sampling 2
    _close(b, a)
    # maps whose rows are not 16-byte aligned (8- and 4-byte staging) and tiny maps (fallback)
    for H, W in [(25, 42), (13, 19), (4, 4), (3, 9)]:
        f = rng.standard_normal((1, 6, H, W)).astype(np.float32)
        r = _rois(rng, 60, 1)
        r[:, 1:] *= np.float32(W * 16 / 1344)
        a = oracle.roi_align_forward(f, r, (7, 7), 1 / 16, 2)
        b, _ = emu.band_forward([f], [1 / 16], r, None, (7, 7), 2)
        _close(b, a)


def test_band_plan_properties():
    # merged weights of a bin sum to the number of in-range samples; band starts are monotone
    import ctypes
    rng = np.random.default_rng(7)
    f = np.ones((1, 4, 50, 84), np.float32)
    rois = _rois(rng, 200, 1)
    out, _ = emu.band_forward([f], [1 / 16], rois, None, (14, 14), 2)
    ref = oracle.roi_align_forward(f, rois, (14, 14), 1 / 16, 2)
    np.testing.assert_allclose(out, ref, rtol=2e-6, atol=2e-6)
