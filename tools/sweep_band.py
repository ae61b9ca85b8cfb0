"""Time the banded (v4) Pooler forward launch against v3 under different knobs (GPU box only).

    python tools/sweep_band.py [--images 16] [--res 14] [--kb 32,40,56] [--cpc 64,128,256]

Knobs are the RP_ROI_* environment variables read in csrc/roi_align.cu (rt_forward)."""
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
    ap.add_argument("--kb", default="32,40,56")
    ap.add_argument("--cpc", default="64,128,256")
    ap.add_argument("--cfg", default="0")
    ap.add_argument("--nbuf", default="2")
    ap.add_argument("--tma", default="0")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    shapes = synth.fpn_shapes(a.images, 256)
    gen = torch.Generator(device=dev).manual_seed(1234)
    feats = [torch.randn(s, device=dev, generator=gen) for s in shapes]
    rois = torch.from_numpy(synth.rois_for_images(a.images, 1000)).to(dev)
    levels = rp.LevelMapper(2, 5).map_rois(rois)
    K = rois.shape[0]
    out = torch.empty((K, 256, a.res, a.res), device=dev)

    def run(tag):
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
        return ms

    os.environ["RP_ROI_IMPL"] = "3"
    ms = run("v3")
    ref = out.clone()
    print("v3 default: %.3f ms  %.2f M RoIs/s" % (ms, K / ms / 1e3), flush=True)
    os.environ["RP_ROI_IMPL"] = "4"
    for tma, nbuf, cfg, kb, cpc in itertools.product(a.tma.split(","), a.nbuf.split(","), a.cfg.split(","),
                                                     a.kb.split(","), a.cpc.split(",")):
        os.environ["RP_ROI_TMA"] = tma
        os.environ["RP_ROI_BAND_NBUF"] = nbuf
        os.environ["RP_ROI_BAND_CFG"] = cfg
        os.environ["RP_ROI_BAND_SMEM_KB"] = kb
        os.environ["RP_ROI_CPC"] = cpc
        out.fill_(7e30)
        ms = run("v4")
        err = float(((out - ref).abs() / (ref.abs() + 1)).max())
        print("v4 tma=%s nbuf=%s cfg=%s kb=%s cpc=%s : %.3f ms  %.2f M RoIs/s  out-only %.0f GB/s  max rel diff vs v3 %.2e"
              % (tma, nbuf, cfg, kb, cpc, ms, K / ms / 1e3, out.numel() * 4 / ms / 1e6, err), flush=True)


if __name__ == "__main__":
    main()
