"""Single-image NMS latency: one problem of n boxes, thr 0.7, no keep cap.
python tools/run_nms_single.py            -> median ms per n
(run it under `ncu --metrics gpu__time_duration.sum` for the per-kernel split)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

def stamps():
    """cycle stamps of the one-launch kernel (library built with RP_EXTRA_NVCC=-DNMS_PROFILE)"""
    import ctypes
    h = rp.lib() if callable(getattr(rp, "lib", None)) else rp.lib
    if not hasattr(h, "rp_debug_nms_prof"):
        return
    buf = (ctypes.c_longlong * 32)()
    torch.cuda.synchronize()
    h.rp_debug_nms_prof(buf)
    v = list(buf)
    print("   band 3, CTA 0: resolve %d (prologue %d, wait+load %d, fixpoint %d, emit %d)  then: barrier %d  suppress %d  barrier %d | signal %d cycles"
          % (v[1] - v[0], v[10] - v[0], v[11] - v[10], v[12] - v[11], v[1] - v[12], v[2] - v[1], v[3] - v[2], v[4] - v[3], v[5] - v[1]))
    print("      fixpoint: all-box rounds %d, compaction %d, compact rounds %d cycles; %d boxes compacted"
          % (v[13] - v[11], v[14] - v[13], v[12] - v[14], v[20]))


dev = torch.device("cuda:0")
reps = int(os.environ.get("REPS", "20"))
for n in (2000, 5000, 20000):
    b, s = synth.nms_boxes(n, 7)[:2]
    b, s = torch.from_numpy(b).to(dev), torch.from_numpy(s).to(dev)
    o = torch.tensor([0, n], dtype=torch.int32, device=dev)
    for mk in (-1, 1000):
        for _ in range(3):
            keep, nk = rp.nms_batched(b, s, o, 0.7, max_keep=mk, n_max=n)
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rp.nms_batched(b, s, o, 0.7, max_keep=mk, n_max=n)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print("n=%d max_keep=%d kept=%d  median %.4f ms  min %.4f ms" % (n, mk, int(nk[0]), float(np.median(ts)), min(ts)))
        if n >= 5000:
            stamps()

# batch of 8 images (BASELINE configs[1]): ms per image
for n in (5000, 20000):
    bs, ss = zip(*[synth.nms_boxes(n, 7 + i)[:2] for i in range(8)])
    b8 = torch.from_numpy(np.concatenate(bs)).to(dev)
    s8 = torch.from_numpy(np.concatenate(ss)).to(dev)
    o8 = torch.from_numpy((np.arange(9) * n).astype(np.int32)).to(dev)
    for mk in (-1, 1000):
        for _ in range(3):
            rp.nms_batched(b8, s8, o8, 0.7, max_keep=mk, n_max=n)
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rp.nms_batched(b8, s8, o8, 0.7, max_keep=mk, n_max=n)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print("batch 8 x n=%d max_keep=%d  median %.4f ms = %.4f ms/image" % (n, mk, float(np.median(ts)), float(np.median(ts)) / 8))
