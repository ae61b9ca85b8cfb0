"""Run the batched NMS once per configuration (GPU box; used under ncu for the launch list)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for B in (8, 1):
    bs, ss = zip(*[synth.nms_boxes(n, 7 + i)[:2] for i in range(B)])
    boxes, scores = T(np.concatenate(bs)), T(np.concatenate(ss))
    offs = T((np.arange(B + 1) * n).astype(np.int32))
    for mk in (-1, 1000):
        for _ in range(reps):
            keep, nk = rp.nms_batched(boxes, scores, offs, 0.7, max_keep=mk, n_max=n)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            keep, nk = rp.nms_batched(boxes, scores, offs, 0.7, max_keep=mk, n_max=n)
        e1.record()
        torch.cuda.synchronize()
        print("n=%d B=%d max_keep=%d: %.3f ms/call, kept %s" % (n, B, mk, e0.elapsed_time(e1) / reps, nk.cpu().tolist()), flush=True)
