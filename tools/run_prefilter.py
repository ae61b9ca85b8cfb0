"""Yolact Detect pre-stage alone (batch 64), for launch lists / ncu captures.
    python tools/run_prefilter.py [reps]"""
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
priors = synth.yolact_priors()
B = 64
imgs = [synth.yolact_image(100 + i, priors) for i in range(4)]
loc = T(np.stack([imgs[i % 4][0] for i in range(B)]))
conf = T(np.stack([imgs[i % 4][1] for i in range(B)]))
mask = T(np.stack([imgs[i % 4][2] for i in range(B)]))
pri = T(priors)
for _ in range(reps):
    r = rp.detect_prefilter(loc, conf, mask, pri, 0.05)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    r = rp.detect_prefilter(loc, conf, mask, pri, 0.05)
e1.record()
torch.cuda.synchronize()
print("prefilter %.3f ms, kept %d" % (e0.elapsed_time(e1) / reps, int(r["n_cand"].item())))
