import os, sys, torch
sys.path.insert(0, os.getcwd())
import roipath_b200 as rp
from tests import synth
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(1234)
feats = [torch.randn(s, device=dev, generator=gen) for s in synth.fpn_shapes(16, 256)]
rois = torch.from_numpy(synth.rois_for_images(16, 1000)).to(dev)
levels = rp.LevelMapper(2, 5).map_rois(rois)
m = levels == int(sys.argv[1]) if len(sys.argv) > 1 else 0
r, lv = rois[m].contiguous(), levels[m].contiguous()
out = torch.empty((r.shape[0], 256, 14, 14), device=dev)
for _ in range(4):
    rp.pooler_forward(feats, synth.FPN_SCALES, r, lv, (14, 14), 2, out=out)
torch.cuda.synchronize()
print(r.shape[0])
