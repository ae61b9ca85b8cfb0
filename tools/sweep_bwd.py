"""Time the Pooler backward launch under different knobs (GPU box only)."""
import itertools
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda", 0)
N = 16
shapes = synth.fpn_shapes(N, 256)
rois = torch.from_numpy(synth.rois_for_images(N, 512)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
ref = {}
for res, k in ((7, rois.shape[0]), (14, 2048)):
    d = torch.randn((k, 256, res, res), device=dev)
    for thr, cc, vs, cpc in itertools.product((256, 128), (32, 16, 8), (8192, 4096), (128, 64)):
        os.environ.update(RP_ROI_BWD_THREADS=str(thr), RP_ROI_BWD_CC=str(cc), RP_ROI_VS_FLOATS=str(vs),
                          RP_ROI_CPC=str(cpc))
        fn = lambda: rp.pooler_backward(d, rois[:k], levels[:k], shapes, synth.FPN_SCALES, 2)
        g = fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g = fn()
        e1.record()
        torch.cuda.synchronize()
        if res not in ref:
            ref[res] = [x.clone() for x in g]
        err = max(float((a - b).abs().max()) for a, b in zip(g, ref[res]))
        print("res %d thr %d cc %d vs %d cpc %d : %.3f ms  maxdiff %.1e" % (res, thr, cc, vs, cpc, e0.elapsed_time(e1) / 5, err), flush=True)
