"""The other RoI-family operators alone, realistic sizes, for timing / ncu:
ROIPool fwd/bwd, boxlist_iou, EmbedMask mask stage, Pose2Seg affine paste, lincomb mask loss, mask
bit-packing.    python tools/run_misc.py [reps]"""
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
rng = np.random.default_rng(0)


def timed(tag, fn, nbytes=None):
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
    extra = "  %.0f GB/s" % (nbytes / (ms * 1e-3) / 1e9) if nbytes else ""
    print("%-44s %.3f ms%s" % (tag, ms, extra))


# ROIPool: one FPN level (P4, 16 images), 8000 RoIs, 7x7x256
feat = torch.randn((16, 256, 50, 84), device=dev)
rois = T(synth.rois_for_images(16, 500))
out, arg = rp.roi_pool(feat, rois, (7, 7), 1 / 16, return_argmax=True)
timed("roi_pool fwd 8000 x 7x7x256", lambda: rp.roi_pool(feat, rois, (7, 7), 1 / 16, return_argmax=True),
      out.numel() * 8)
# boxlist_iou 2000 x 1000
a = rp.BoxList(T(synth.proposals(2000, 1)), (1333, 800))
b = rp.BoxList(T(synth.proposals(1000, 2)), (1333, 800))
timed("boxlist_iou 2000 x 1000", lambda: rp.boxlist_iou(a, b), 2000 * 1000 * 4)
# EmbedMask mask stage: 100 proposals, 32-dim embedding at 200 x 336, upsample x4, crop 800 x 1333
dim, m_h, m_w, n = 32, 200, 336, 100
pe = torch.randn((dim, m_h, m_w), device=dev) * 0.3
ce = torch.randn((n, dim), device=dev) * 0.3
mg = torch.rand((n,), device=dev) + 0.05
bx = T(synth.proposals(n, 3, smin=24, smax=500, adversarial=False))
masks = rp.embed_masks(pe, ce, mg, bx, 4.0)
timed("embed_masks 100 x 200x336 (dim 32)", lambda: rp.embed_masks(pe, ce, mg, bx, 4.0), n * m_h * m_w * 4)
timed("resize_threshold x4 -> 100 x 800x1333", lambda: rp.resize_threshold(masks, 4.0, 0.5, 800, 1333),
      n * 800 * 1333)
# Pose2Seg paste: 8 instances, 64x64x2 maps into 512 x 512 (the model's input size), and 100 into 800 x 1333
pred = torch.softmax(torch.randn((100, 64, 64, 2), device=dev), -1)
M = np.tile(np.array([[0.1, 0.0, -5.0], [0.0, 0.1, -3.0]]), (100, 1, 1))
timed("affine_paste 8 x 512x512", lambda: rp.affine_paste(pred[:8], M[:8], 512, 512), 8 * 512 * 512)
timed("affine_paste 100 x 800x1333", lambda: rp.affine_paste(pred, M, 1333, 800), 100 * 800 * 1333)
# lincomb mask loss: 100 positives, 138 x 138 x 32 prototypes
proto = torch.relu(torch.randn((138, 138, 32), device=dev))
coef = torch.tanh(torch.randn((100, 32), device=dev))
mt = (torch.rand((138, 138, 100), device=dev) > 0.6).float()
gb = T(np.clip(synth.proposals(100, 5, img_w=1, img_h=1, smin=0.1, smax=0.6, adversarial=False), 0, 1))
timed("lincomb_mask_loss fwd+bwd n=100 138x138x32", lambda: rp.lincomb_mask_loss_raw(proto, coef, mt, gb, None),
      138 * 138 * (100 + 64) * 4)
# bit-packing of 6400 masks of 550 x 550
mk = (torch.rand((1600, 550, 550), device=dev) > 0.5).to(torch.uint8)
timed("pack_masks 1600 x 550x550", lambda: rp.parallel.pack_masks(mk), mk.numel() * 1.125)
