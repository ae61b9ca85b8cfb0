"""Paste-back timing (BASELINE configs[4]: 100 instances per image at 800x1333) beside the write speed of
the same bytes (torch zero_ = a fill kernel): python tools/run_paste.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731


def med(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


m_, b_ = [T(x) for x in synth.paste_inputs(100, 200)]
one = torch.empty((100, 800, 1333), dtype=torch.uint8, device=dev)
nbytes = one.numel()
t = med(lambda: rp.paste_masks(m_, b_, 800, 1333, out=one))
print("paste 100 instances into one 107 MB buffer (L2 resident): %.4f ms = %.0f GB/s" % (t, nbytes / t / 1e6))
t = med(lambda: one.zero_())
print("zero_ of the same buffer: %.4f ms = %.0f GB/s" % (t, nbytes / t / 1e6))
big = torch.empty((32, 100, 800, 1333), dtype=torch.uint8, device=dev)  # 3.4 GB: 32 images, distinct outputs


def paste32():
    for i in range(32):
        rp.paste_masks(m_, b_, 800, 1333, out=big[i])


t = med(paste32, n=5, warm=2)
print("paste 32 images into distinct buffers (3.4 GB): %.4f ms = %.0f GB/s, %.4f ms/image" % (t, big.numel() / t / 1e6, t / 32))
t = med(lambda: big.zero_(), n=5, warm=2)
print("zero_ of the 3.4 GB: %.4f ms = %.0f GB/s" % (t, big.numel() / t / 1e6))
