"""Time the fused Pooler forward launch under different launch-plan knobs (GPU box only).

    python tools/sweep_roi.py [--images 16] [--res 14]

The knobs are the RP_ROI_* environment variables read by rt_plan_launch (csrc/roi_align.cu)."""
import argparse
import itertools
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=16)
    ap.add_argument("--res", type=int, default=14)
    ap.add_argument("--cpc", default="32,64,128,256")
    ap.add_argument("--u", default="1,2,4")
    ap.add_argument("--tile", default="12288")
    ap.add_argument("--bwd", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    shapes = synth.fpn_shapes(a.images, 256)
    gen = torch.Generator(device=dev).manual_seed(1234)
    feats = [torch.randn(s, device=dev, generator=gen) for s in shapes]
    rois = torch.from_numpy(synth.rois_for_images(a.images, 1000)).to(dev)
    levels = rp.LevelMapper(2, 5).map_rois(rois)
    K = rois.shape[0]
    out = torch.empty((K, 256, a.res, a.res), device=dev)
    ref = None
    for tile, cpc, u in itertools.product(a.tile.split(","), a.cpc.split(","), a.u.split(",")):
        os.environ["RP_ROI_TILE_FLOATS"] = tile
        os.environ["RP_ROI_CPC"] = cpc
        os.environ["RP_ROI_U"] = u
        fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (a.res, a.res), 2, out=out)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        if ref is None:
            ref = out.clone()
            err = 0.0
        else:
            err = float((out - ref).abs().max())
        gbs = out.numel() * 4 / ms / 1e6
        print("tile=%s cpc=%s U=%s : %.3f ms  %.2f M RoIs/s  out-only %.0f GB/s  maxdiff %.2e"
              % (tile, cpc, u, ms, K / ms / 1e3, gbs, err), flush=True)


if __name__ == "__main__":
    main()
