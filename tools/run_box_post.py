"""Box-head post-processor alone (8 images x 1000 proposals x 81 classes), for launch lists.
    python tools/run_box_post.py [reps] [cand_slots_per_row]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
slots = int(sys.argv[2]) if len(sys.argv) > 2 else 80
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
lg, reg, pr, counts, wh = synth.box_head_batch(8, 1000, seed=3)
lgt, regt, prt, wht = T(lg), T(reg), T(pr), T(wh)
offs = T(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))
R = lg.shape[0]
fn = lambda: rp.box_post(lgt, regt, prt, offs, wht, 1000, 0.05, 0.5, 100, cand_cap=min(R * 80, R * slots))
for _ in range(3):
    r = fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    r = fn()
e1.record()
torch.cuda.synchronize()
print("box_post %.3f ms (cand_cap %d, candidates %d, kept %s)"
      % (e0.elapsed_time(e1) / reps, r["cand_cap"], int(r["n_cand"].item()), r["count"].cpu().tolist()))
