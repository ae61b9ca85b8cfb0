"""Yolact post-processing (FastNMS + mask assembly at 550x550), batch 64 (BASELINE configs[3]).
python tools/run_yolact_post.py   (under `ncu --metrics gpu__time_duration.sum` for the per-kernel split)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
B = 64
reps = int(os.environ.get("REPS", "10"))
priors = synth.yolact_priors()
base = []
for i in range(4):
    loc, conf, mask, proto = synth.yolact_image(100 + i, priors)
    boxes = rp.decode(T(loc), T(priors)).cpu().numpy()
    fg = np.ascontiguousarray(conf[:, 1:].T)
    kp = np.nonzero(fg.max(0) > np.float32(0.05))[0]
    base.append((boxes[kp], mask[kp], np.ascontiguousarray(fg[:, kp]), proto))
cands = [base[i % 4] for i in range(B)]
offs_np = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
yb, yc = T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands]))
ys = T(np.concatenate([c[2] for c in cands], 1))
yp = T(np.stack([c[3] for c in cands]))
yo = T(offs_np)
masks_buf = torch.empty((B * 100, 550, 550), dtype=torch.uint8, device=dev)
slot_off = T((np.arange(B + 1) * 100).astype(np.int32))
n_max = int(np.diff(offs_np).max())
print("candidates per image", np.diff(offs_np)[:4], "n_max", n_max)


def fast():
    return rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100, n_max=n_max)


def post():
    r = fast()
    rp.mask_assemble(yp, r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4), 550, 550, True,
                     det_offsets=slot_off, out=masks_buf)
    return r


for name, fn in (("fast_nms", fast), ("post (fast_nms + assembly)", post)):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print("%s: median %.4f ms per batch of %d = %.0f images/s" % (name, float(np.median(ts)), B, B / np.median(ts) * 1e3))
