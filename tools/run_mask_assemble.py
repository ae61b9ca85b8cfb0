"""Mask assembly alone (batch 64, 550x550), for timing / ncu captures.
    python tools/run_mask_assemble.py [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
priors = synth.yolact_priors()
B = 64
base = []
for i in range(4):
    loc, conf, mask, proto = synth.yolact_image(100 + i, priors)
    bx = rp.decode(T(loc), T(priors)).cpu().numpy()
    fg = np.ascontiguousarray(conf[:, 1:].T)
    kp = np.nonzero(fg.max(0) > np.float32(0.05))[0]
    base.append((bx[kp], mask[kp], np.ascontiguousarray(fg[:, kp]), proto))
cands = [base[i % 4] for i in range(B)]
offs = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
r = rp.fast_nms_batched(T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands])),
                        T(np.concatenate([c[2] for c in cands], 1)), T(offs), 0.5, 200, 100)
yp = T(np.stack([c[3] for c in cands]))
out = torch.empty((B * 100, 550, 550), dtype=torch.uint8, device=dev)
slot = T((np.arange(B + 1) * 100).astype(np.int32))
cf, bx = r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4)
fn = lambda: rp.mask_assemble(yp, cf, bx, 550, 550, True, det_offsets=slot, out=out)
for _ in range(3):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    fn()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print("mask_assemble %.3f ms  (%d set pixels, checksum %d)" % (ms, int(out.sum()), int(out[::97].long().sum())))
