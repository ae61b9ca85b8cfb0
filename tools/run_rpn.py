"""RPN proposal stage alone (5 levels x 8 images, top-n 1000), for launch lists / ncu captures.
    python tools/run_rpn.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
N, A, top_n = 8, 3, 1000
lv = [synth.rpn_level(N, A, H, W, s, seed=100 + s)
      for H, W, s in ((200, 336, 4), (100, 168, 8), (50, 84, 16), (25, 42, 32), (13, 21, 64))]
objs, regs, ancs, wh = [T(l[1]) for l in lv], [T(l[2]) for l in lv], [T(l[0]) for l in lv], T(lv[0][3])


def step():
    bx, sc, _, off, ks = rp.rpn_pre_nms_levels(objs, regs, ancs, wh, top_n, 0.0)
    keep, nk = rp.nms_batched(bx, sc, off, 0.7, max_keep=top_n, n_max=max(ks),
                              flags=rp.package._lib.RP_NMS_PRESORTED)  # as RPNPostProcessor calls it
    return rp.rpn_select_levels(bx, sc, off, keep, nk, len(lv), N, top_n)


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    r = step()
e1.record()
torch.cuda.synchronize()
print("rpn device part %.3f ms, counts %s" % (e0.elapsed_time(e1) / reps, r[2].cpu().tolist()))
