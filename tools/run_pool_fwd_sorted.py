import os, sys, torch
ROOT = "/root/repo" if os.path.exists("/root/repo/tests") else os.getcwd()
sys.path.insert(0, ROOT)
import roipath_b200 as rp
from tests import synth
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(16, 256)]
rois = torch.from_numpy(synth.rois_for_images(16, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
def timed(r, lv, reps=10):
    out = torch.empty((r.shape[0], 256, 14, 14), device=dev)
    fn = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, r, lv, (14, 14), 2, out=out)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
print("original order      %.3f ms" % timed(rois, levels))
key = levels.long() * 10**9 + rois[:, 0].long() * 10**6 + (rois[:, 2].clamp(0, 999)).long() * 1000 + (rois[:, 1].clamp(0, 1343) / 2).long()
o = torch.argsort(key)
print("sorted (level,img,y,x) %.3f ms" % timed(rois[o].contiguous(), levels[o].contiguous()))
key = levels.long() * 10**9 + rois[:, 0].long() * 10**6
o = torch.argsort(key, stable=True)
print("sorted (level,img)  %.3f ms" % timed(rois[o].contiguous(), levels[o].contiguous()))
o = torch.argsort(rois[:, 0].long() * 10 + levels.long(), stable=True)
print("sorted (img,level)  %.3f ms" % timed(rois[o].contiguous(), levels[o].contiguous()))
area = (rois[:, 3] - rois[:, 1]) * (rois[:, 4] - rois[:, 2])
o = torch.argsort(-area)
print("sorted by area desc %.3f ms" % timed(rois[o].contiguous(), levels[o].contiguous()))
