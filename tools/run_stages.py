"""Run every other stage of the path once or twice at its BASELINE configuration (GPU box; used
under ncu for per-kernel captures, and plain for a timing table).

    python tools/run_stages.py [reps]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import roipath_b200 as rp  # noqa: E402
from tests import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def timed(name, fn, nbytes=None):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("%-44s %8.3f ms%s" % (name, ms, ("  %7.0f GB/s" % (nbytes / ms / 1e6)) if nbytes else ""), flush=True)


# Pooler backward, train configuration: 16 images x 512 RoIs at 7x7 (box head), 256 channels
N = 16
shapes = synth.fpn_shapes(N, 256)
rois = T(synth.rois_for_images(N, 512))
levels = rp.LevelMapper(2, 5).map_rois(rois)
for res in (7, 14):
    k = rois.shape[0] if res == 7 else 2048
    d = torch.randn((k, 256, res, res), device=dev)
    grad_bytes = sum(int(np.prod(s)) for s in shapes) * 4
    timed("pooler bwd %dx%d, %d RoIs" % (res, res, k),
          lambda: rp.pooler_backward(d, rois[:k], levels[:k], shapes, synth.FPN_SCALES, 2),
          d.numel() * 4 + grad_bytes)
    del d

# greedy NMS, 8 images x 20k boxes
B, n = 8, 20000
bs, ss = zip(*[synth.nms_boxes(n, 7 + i)[:2] for i in range(B)])
boxes, scores = T(np.concatenate(bs)), T(np.concatenate(ss))
offs = T((np.arange(B + 1) * n).astype(np.int32))
timed("nms 8 x 20000, thr 0.7", lambda: rp.nms_batched(boxes, scores, offs, 0.7, n_max=n))

# Yolact post, batch 64
priors = synth.yolact_priors()
B = 64
base = []
for i in range(4):
    loc, conf, mask, proto = synth.yolact_image(100 + i, priors)
    bx = rp.decode(T(loc), T(priors)).cpu().numpy()
    fg = np.ascontiguousarray(conf[:, 1:].T)
    kp = np.nonzero(fg.max(0) > np.float32(0.05))[0]
    base.append((bx[kp], mask[kp], np.ascontiguousarray(fg[:, kp]), proto))
cands = [base[i % 4] for i in range(B)]
offs_np = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
yb, yc = T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands]))
ys = T(np.concatenate([c[2] for c in cands], 1))
yp = T(np.stack([c[3] for c in cands]))
yo = T(offs_np)
masks_buf = torch.empty((B * 100, 550, 550), dtype=torch.uint8, device=dev)
slot_off = T((np.arange(B + 1) * 100).astype(np.int32))
res = {}


def fnms():
    res["r"] = rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100, n_max=int(np.diff(offs_np).max()))


timed("fast_nms batch 64", fnms)
r = res["r"]
timed("mask_assemble batch 64, 550x550",
      lambda: rp.mask_assemble(yp, r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4), 550, 550,
                               True, det_offsets=slot_off, out=masks_buf),
      B * (138 * 138 * 32 * 4) + B * 100 * 550 * 550)
del masks_buf, yp

# paste-back, 100 instances at 1333x800
pm, pb = synth.paste_inputs(100, 200)
pmt, pbt = T(pm), T(pb)
pout = torch.empty((100, 800, 1333), dtype=torch.uint8, device=dev)
timed("paste_masks 100 x 800x1333", lambda: rp.paste_masks(pmt, pbt, 800, 1333, out=pout),
      100 * 28 * 28 * 4 + 100 * 800 * 1333)

# RPN proposal stage, 5 FPN levels of an 800 x 1344 batch, test configuration
# (PRE_NMS_TOP_N_TEST 1000, POST_NMS_TOP_N_TEST 1000, FPN_POST_NMS_TOP_N_TEST 1000, thr 0.7)
N, A = 8, 3
shapes5 = [(200, 336, 4), (100, 168, 8), (50, 84, 16), (25, 42, 32), (13, 21, 64)]
lv = [synth.rpn_level(N, A, H, W, s, seed=100 + s) for H, W, s in shapes5]
im_wh = lv[0][3]
anchors = [[rp.BoxList(T(l[0][b]), (int(im_wh[b, 0]), int(im_wh[b, 1]))) for l in lv] for b in range(N)]
objs, regs = [T(l[1]) for l in lv], [T(l[2]) for l in lv]
anc_t = [T(l[0]) for l in lv]
wh_t = T(im_wh)
head_bytes = sum(l[1].nbytes + l[2].nbytes + l[0].nbytes for l in lv)
for top_n in (1000, 2000):
    post = rp.RPNPostProcessor(top_n, top_n, 0.7, 0, fpn_post_nms_top_n=top_n)
    post.eval()

    def device_part():
        bx, sc, ix, off, ks = rp.rpn_pre_nms_levels(objs, regs, anc_t, wh_t, top_n, 0.0)
        keep, nk = rp.nms_batched(bx, sc, off, 0.7, max_keep=top_n, n_max=max(ks))
        rp.rpn_select_levels(bx, sc, off, keep, nk, len(lv), N, top_n)

    def per_level():
        for a, o, r in zip(anc_t, objs, regs):
            bx, sc, ix, off, k = rp.rpn_pre_nms(o, r, a, wh_t, top_n, 0.0)
            rp.nms_batched(bx, sc, off, 0.7, max_keep=top_n, n_max=k)

    timed("rpn pre-NMS 5 levels x %d images, top %d" % (N, top_n),
          lambda: rp.rpn_pre_nms_levels(objs, regs, anc_t, wh_t, top_n, 0.0), head_bytes)
    timed("rpn pre-NMS + NMS, one call per level", per_level)
    timed("rpn pre-NMS + NMS + select, all levels fused", device_part)
    timed("RPNPostProcessor.execute (BoxLists out)", lambda: post(anchors, objs, regs))

# box-head post-processor, 8 images x 1000 proposals, 81 classes (thr 0.05, NMS 0.5, 100 per image)
lg, reg, pr, counts, wh = synth.box_head_batch(8, 1000, seed=3)
lgt, regt, prt, wht = T(lg), T(reg), T(pr), T(wh)
offs_bh = T(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))
bls, o = [], 0
for b, n in enumerate(counts):
    bls.append(rp.BoxList(prt[o:o + n], (int(wh[b, 0]), int(wh[b, 1]))))
    o += n
bh = rp.PostProcessor(0.05, 0.5, 100)
timed("box post device part, 8 x 1000 x 81", lambda: rp.box_post(lgt, regt, prt, offs_bh, wht, 1000, 0.05, 0.5, 100),
      lg.nbytes + reg.nbytes + pr.nbytes)
timed("PostProcessor.execute (BoxLists out)", lambda: bh((lgt, regt), bls))

# Yolact Detect for the batch from the raw head outputs: prefilter (decode + class-max + conf
# filter + gathers over the 399 MB conf tensor) -> FastNMS, batch 64
B = 64
imgs = [synth.yolact_image(100 + i, priors) for i in range(4)]
loc64 = T(np.stack([imgs[i % 4][0] for i in range(B)]))
conf64 = T(np.stack([imgs[i % 4][1] for i in range(B)]))
mask64 = T(np.stack([imgs[i % 4][2] for i in range(B)]))
pri_t = T(priors)
timed("yolact prefilter batch 64 (19248 priors x 81)",
      lambda: rp.detect_prefilter(loc64, conf64, mask64, pri_t, 0.05), conf64.numel() * 4 + loc64.numel() * 4)
det = rp.Detect(81, 0, 200, 0.05, 0.5)
det.use_fast_nms = True
pred = {"loc": loc64, "conf": conf64, "mask": mask64, "priors": pri_t}
timed("Detect.__call__ batch 64 (prefilter + FastNMS, dicts out)", lambda: det(pred, None))
