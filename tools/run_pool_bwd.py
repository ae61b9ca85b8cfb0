"""Pooler backward alone at the default knobs (16 images x 512 RoIs 7x7; 2048 RoIs 14x14).
    python tools/run_pool_bwd.py [reps]"""
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
N = 16
shapes = synth.fpn_shapes(N, 256)
rois = torch.from_numpy(synth.rois_for_images(N, 512)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
grad_bytes = sum(int(np.prod(s)) for s in shapes) * 4
for res, k in ((7, int(rois.shape[0])), (14, 2048)):
    d = torch.randn((k, 256, res, res), device=dev)
    fn = lambda: rp.pooler_backward(d, rois[:k], levels[:k], shapes, synth.FPN_SCALES, 2)
    for _ in range(10):  # the first launches after start-up run at ramping clocks
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g = fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("pooler bwd %dx%d x %d: %.3f ms (%.0f GB/s incl. zero fill), |g| %.4e"
          % (res, res, k, ms, (d.numel() * 4 + grad_bytes) / ms / 1e6, float(sum(x.abs().sum() for x in g))))
