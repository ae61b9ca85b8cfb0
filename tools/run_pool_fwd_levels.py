"""Per-level timing of the fused Pooler forward on the bench workload (GPU box).
    python tools/run_pool_fwd_levels.py [images] [res] [reps]
Prints, per FPN level, the RoI count, the launch time with only that level's RoIs, and
us per RoI; then the full launch.  """
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

images = int(sys.argv[1]) if len(sys.argv) > 1 else 16
res = int(sys.argv[2]) if len(sys.argv) > 2 else 14
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(images, 256)]
rois = torch.from_numpy(synth.rois_for_images(images, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)


def timed(r, lv):
    out = torch.empty((r.shape[0], 256, res, res), device=dev)
    fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, r, lv, (res, res), 2, out=out)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for l in range(4):
    m = levels == l
    r, lv = rois[m].contiguous(), levels[m].contiguous()
    ms = timed(r, lv)
    print("level %d: %5d RoIs  %.3f ms  %.3f us/RoI" % (l, r.shape[0], ms, 1e3 * ms / max(1, r.shape[0])))
ms = timed(rois, levels)
print("all    : %5d RoIs  %.3f ms  %.3f us/RoI" % (rois.shape[0], ms, 1e3 * ms / rois.shape[0]))
