"""Tensor-core Pooler forward (RP_ROI_IMPL=6) against the oracle and the SIMT kernel, then timing (GPU box).
    python tools/check_rx.py [reps]"""
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
os.environ["RP_RX_DEBUG"] = "1"
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def dump_timeouts():
    import ctypes
    buf = (ctypes.c_int * 128)()
    if rp.lib().rp_debug_rx_timeouts(buf, 128):
        names = {1: "plan_full", 2: "plan_empty", 3: "tile_empty", 4: "d_empty", 5: "v_full", 6: "d_full",
                 7: "v_empty/v_freed", 8: "tile_full", 9: "issued"}
        for w in range(16):
            e = buf[8 * w:8 * w + 5]
            if e[0]:
                print("  warp %2d block %3d waits on %-10s parity %d counter %d" % (w, e[1], names.get(e[2], e[2]), e[3], e[4]))


def dump_profile():
    """RX_PROFILE build: cycles per warp of block 0 spent in each wait (lane 0), whole loop, work."""
    import ctypes
    buf = (ctypes.c_int * 1024)()
    if not rp.lib().rp_debug_rx_timeouts(buf, 1024):
        return
    a = np.frombuffer(buf, dtype=np.int32).copy()[256:256 + 32 * 12 * 2].view(np.int64).reshape(32, 12)
    if not a.any():
        return
    roles = {0: "TMA", 1: "TMA", 2: "PLAN", 3: "PLAN", 4: "EPI", 5: "EPI", 6: "EPI", 7: "EPI", 8: "FREE"}
    names = ["total", "plan_full", "plan_empty", "tile_empty", "d_empty", "v_full", "d_full", "v_freed",
             "tile_full", "issued", "work", "units"]
    print("  warp role  " + " ".join("%10s" % n for n in names))
    for w in range(32):
        if a[w, 0]:
            print("  %4d %-5s " % (w, roles.get(w, "Y")) + " ".join("%10d" % v for v in a[w]))


def run(impl, fn):
    os.environ["RP_ROI_IMPL"] = str(impl)
    rp.reload_env()
    try:
        out = fn()
        torch.cuda.synchronize()
    except Exception as ex:
        print("CUDA failure:", str(ex).splitlines()[0])
        dump_timeouts()
        sys.exit(2)
    return out


def report(tag, got, exp):
    err = np.abs(got - exp)
    tol = 1e-5 * np.abs(exp) + 1e-5
    bad = int((err > tol).sum())
    print("%-34s max|err| %.3e  violations %d / %d  %s" % (tag, err.max(), bad, err.size, "OK" if bad == 0 else "FAIL"))
    if bad:
        idx = np.argwhere(err > tol)[:5]
        for i in idx:
            print("   at", tuple(i), "got", got[tuple(i)], "exp", exp[tuple(i)])
    return bad == 0


ok = True
rng = np.random.default_rng(0)
# ---- 0: tiny cases first (one RoI, 8 channels), growing
feat0 = rng.standard_normal((1, 64, 50, 84)).astype(np.float32)
for tag, r0, c0 in (("one RoI 8 ch", [[0, 10, 20, 300, 400]], 8), ("one RoI 64 ch", [[0, 10, 20, 300, 400]], 64),
                    ("wide RoI (two K halves)", [[0, 100, 100, 800, 400]], 16),
                    ("tall RoI", [[0, 100, 10, 400, 790]], 16),
                    ("4 RoIs", [[0, 10, 20, 300, 400], [0, 0, 0, 50, 60], [0, 500, 300, 900, 600], [0, 33.3, 47.7, 101.9, 222.2]], 16)):
    r0 = np.asarray(r0, np.float32)
    f0 = np.ascontiguousarray(feat0[:, :c0])
    exp = oracle.roi_align_forward(f0, r0, (14, 14), 1 / 16, 2)
    got = run(6, lambda: rp.roi_align_forward(T(f0), T(r0), (14, 14), 1 / 16, 2, mode=0)).cpu().numpy()
    ok &= report("0 " + tag, got, exp)
    sys.stdout.flush()
# ---- A: single level, random RoIs of every size (most large ones take the in-kernel fallback)
N, C = 2, 64
feat = rng.standard_normal((N, C, 50, 84)).astype(np.float32)
K = 200
ctr = rng.uniform(0, 1, (K, 2)) * [1344, 800]
wh = np.exp(rng.uniform(np.log(2), np.log(1200), (K, 2)))
rois = np.concatenate([rng.integers(0, N, (K, 1)), ctr - wh / 2, ctr + wh / 2], 1).astype(np.float32)
rois[:6, 1:] = [[10, 20, 300, 400], [0, 0, 1343, 799], [50, 50, 50, 50], [-40, -30, 100, 90],
                [1200, 700, 1500, 900], [-500, -500, -300, -300]]
for P, sr in (((14, 14), 2), ((7, 7), 2), ((14, 14), 1), ((3, 5), 3), ((16, 16), 2)):
    exp = oracle.roi_align_forward(feat, rois, P, 1 / 16, sr)
    got = run(6, lambda: rp.roi_align_forward(T(feat), T(rois), P, 1 / 16, sr, mode=0)).cpu().numpy()
    ok &= report("A single level P=%s sr=%d" % (P, sr), got, exp)

# ---- B: multi-level Pooler, 2 images x 300 RoIs, 16 and 136 channels
for C in (16, 136):
    feats = synth.fpn_features(2, C)
    rois = synth.rois_for_images(2, 300)
    levels = rp.LevelMapper(2, 5).map_rois(T(rois))
    exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, levels.cpu().numpy(), (14, 14), 2)
    xs = [T(f) for f in feats]
    got = run(6, lambda: rp.pooler_forward(xs, synth.FPN_SCALES, T(rois), levels, (14, 14), 2)).cpu().numpy()
    ok &= report("B pooler C=%d 14x14" % C, got, exp)
    exp = oracle.pooler_forward(feats, synth.FPN_SCALES, rois, levels.cpu().numpy(), (7, 7), 2)
    got = run(6, lambda: rp.pooler_forward(xs, synth.FPN_SCALES, T(rois), levels, (7, 7), 2)).cpu().numpy()
    ok &= report("B pooler C=%d 7x7" % C, got, exp)

# ---- C: the bench workload: impl 6 against impl 3, and timing
images = 16
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(images, 256)]
rois = torch.from_numpy(synth.rois_for_images(images, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
for res in (14, 7):
    o3 = torch.empty((rois.shape[0], 256, res, res), device=dev)
    o6 = torch.empty_like(o3)
    run(3, lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (res, res), 2, out=o3))
    run(6, lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (res, res), 2, out=o6))
    d = (o3 - o6).abs()
    tol = 1e-5 * o3.abs() + 1e-5
    bad = int((d > tol).sum().item())
    print("C bench %dx%d impl 6 vs impl 3: max|diff| %.3e, violations %d %s" % (res, res, d.max().item(), bad, "OK" if bad == 0 else "FAIL"))
    ok &= bad == 0
    for impl, o in ((3, o3), (6, o6)):
        os.environ["RP_ROI_IMPL"] = str(impl)
        rp.reload_env()
        fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (res, res), 2, out=o)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print("  impl %d %dx%d: %.3f ms" % (impl, res, res, e0.elapsed_time(e1) / reps))
        if impl == 6:
            dump_profile()
print("ALL OK" if ok else "SOME FAILED")
sys.exit(0 if ok else 1)
