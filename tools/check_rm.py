"""Chained-MMA Pooler forward (RP_ROI_IMPL=7) against the oracle and the tile-walk kernel, then timing (GPU box).
    python tools/check_rm.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def run(impl, fn):
    os.environ["RP_ROI_IMPL"] = str(impl)
    rp.reload_env()
    out = fn()
    torch.cuda.synchronize()
    return out


def report(tag, got, exp):
    err = np.abs(got - exp)
    tol = 1e-5 * np.abs(exp) + 1e-5
    bad = int((err > tol).sum())
    rel = (err / (np.abs(exp) + 1.0)).max()
    print("%-40s max|err| %.3e  max err/(|ref|+1) %.3e  violations %d / %d  %s"
          % (tag, err.max(), rel, bad, err.size, "OK" if bad == 0 else "FAIL"))
    if bad:
        for i in np.argwhere(err > tol)[:5]:
            print("   at", tuple(i), "got", got[tuple(i)], "exp", exp[tuple(i)])
    return bad


rng = np.random.default_rng(11)
fails = 0
for P, sr in [((14, 14), 2), ((7, 7), 2), ((14, 14), 0), ((3, 5), 3), ((16, 16), 1)]:
    for (H, W), C in [((50, 84), 37), ((25, 42), 8), ((13, 19), 5), ((3, 9), 4), ((100, 168), 16)]:
        feat = rng.standard_normal((2, C, H, W)).astype(np.float32)
        rois = np.zeros((90, 5), np.float32)
        rois[:, 0] = rng.integers(0, 2, 90)
        x1 = rng.uniform(-20, W * 16 - 8, 90)
        y1 = rng.uniform(-20, H * 16 - 8, 90)
        rois[:, 1] = x1
        rois[:, 2] = y1
        rois[:, 3] = x1 + rng.uniform(1, W * 8, 90)
        rois[:, 4] = y1 + rng.uniform(1, H * 8, 90)
        exp = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
        got = run(7, lambda: rp.roi_align(T(feat), T(rois), P, 1 / 16, sr)).cpu().numpy()
        fails += report("P %s sr %d map %dx%d C %d" % (P, sr, H, W, C), got, exp)

# the headline workload: 7 against 3 (and a sample against the oracle), both precision modes
images = 16
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(images, 256)]
rois = torch.from_numpy(synth.rois_for_images(images, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
fe = [f.cpu().numpy() for f in feats]
sub = slice(0, 64)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for res in (14, 7):
    out = torch.empty((rois.shape[0], 256, res, res), device=dev)
    fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (res, res), 2, out=out)
    exp = oracle.pooler_forward(fe, synth.FPN_SCALES, rois[sub].cpu().numpy(), levels[sub].cpu().numpy(), (res, res), 2)
    a = run(3, fn).clone()
    print("headline %dx%d  impl 3: %.3f ms" % (res, res, timeit(fn)))
    for mode in ("0", "1") + (("2",) if os.environ.get("RM_WITH_MODE2") else ()):  # mode 2 needs a -DRM_WITH_MODE2 build
        for cpc in ("128", "256"):
            os.environ["RP_ROI_MMA_MODE"] = mode
            os.environ["RP_ROI_MMA_CPC"] = cpc
            b = run(7, fn).clone()
            d = (a - b).abs()
            ms = timeit(fn)
            print("  impl 7 mode %s cpc %s: %.3f ms   vs impl 3: max|diff| %.3e  max diff/(|ref|+1) %.3e"
                  % (mode, cpc, ms, float(d.max()), float((d / (a.abs() + 1)).max())))
        bad = report("  mode %s first 64 RoIs vs oracle" % mode, b[sub].cpu().numpy(), exp)
        fails += bad if mode != "2" else 0  # mode 2 (single tf32 product) is a measurement, not a product path
sys.exit(1 if fails else 0)
