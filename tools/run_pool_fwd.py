"""Run the fused Pooler forward launch a few times (GPU box; used under ncu).
    python tools/run_pool_fwd.py [images] [res] [reps]   (knobs: RP_ROI_* environment variables)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

images = int(sys.argv[1]) if len(sys.argv) > 1 else 16
res = int(sys.argv[2]) if len(sys.argv) > 2 else 14
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(images, 256)]
if os.environ.get("LAYOUT", "nchw") == "nhwc":  # torch.channels_last: the NHWC entry
    feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
rois = torch.from_numpy(synth.rois_for_images(images, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
out = torch.empty((rois.shape[0], 256, res, res), device=dev)
fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (res, res), 2, out=out)
for _ in range(3):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    fn()
e1.record()
torch.cuda.synchronize()
print("pooler fwd %s %dx%d, %d RoIs: %.3f ms" % (os.environ.get("LAYOUT", "nchw"), res, res, rois.shape[0], e0.elapsed_time(e1) / reps))
