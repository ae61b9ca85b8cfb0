"""The other figures of the compound metric (BASELINE.json: NMS ms @20k boxes, Yolact post imgs/s,
paste-back, plus the widened stages), measured after bench.py's timed region.

Every configuration is sharded BY IMAGE across the ranks (parallel.shard_range; no data-path
collective): configs[1] NMS batch 8, configs[3] Yolact post batch 64, configs[4] paste-back batch
128.  Each rank times its own shard with CUDA events (median of n), the per-op times are reduced
with MAX over the ranks, and rank 0 derives the whole-job figures from the global batch sizes.
`gather` runs parallel.gather_detections (the path's ONE exchange: kept detections + bit-packed
550x550 masks to rank 0 over NCCL) on the Yolact shard outputs and reports its time and bytes
separately -- it is outside every timed kernel figure.
`cpu` holds the oracle (CPU restatement) timed on the host cores on a bounded sample of the same
inputs, with the number of threads used.
"""
import os
import time

import numpy as np

from tests import synth

NMS_BATCH = 8
YOLACT_BATCH = 64
PASTE_BATCH = 128
CHANNELS = 256
SAMPLING = 2


def _med(torch, fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize()
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def local_times(rp, torch, dev, rank, world, images_per_gpu):
    """Times (ms) of this rank's shard of every configuration, plus the Yolact shard's outputs for
    the gather.  Keys are identical on every rank."""
    ms = {}
    keep = {}
    med = lambda fn, n=10, warm=3: _med(torch, fn, n, warm)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    shard = lambda n: rp.parallel.shard_range(n, rank, world)

    # --- Pooler backward, train shape of BASELINE configs[2] (weak: every rank its own 16 images)
    shapes = synth.fpn_shapes(images_per_gpu, CHANNELS)
    rois_b = T(synth.rois_for_images(images_per_gpu, 512))
    lv_b = rp.LevelMapper(2, 5).map_rois(rois_b)
    for res, k in ((7, int(rois_b.shape[0])), (14, 2048)):
        d = torch.randn((k, CHANNELS, res, res), device=dev)
        ms["pooler_bwd_%dx%d" % (res, res)] = med(
            lambda: rp.pooler_backward(d, rois_b[:k], lv_b[:k], shapes, synth.FPN_SCALES, SAMPLING), n=5)
        del d
    # --- Pooler forward on torch.channels_last feature maps (the NHWC entry), same workload as the headline
    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    feats_cl = [torch.randn(s_, device=dev, generator=gen).contiguous(memory_format=torch.channels_last)
                for s_ in shapes]
    rois_f = T(synth.rois_for_images(images_per_gpu, 1000))
    lv_f = rp.LevelMapper(2, 5).map_rois(rois_f)
    out_f = torch.empty((rois_f.shape[0], CHANNELS, 14, 14), device=dev)
    ms["pooler_fwd_nhwc_14x14"] = med(
        lambda: rp.pooler_forward(feats_cl, synth.FPN_SCALES, rois_f, lv_f, (14, 14), SAMPLING, out=out_f), n=5)
    del feats_cl, out_f
    # --- NMS sweep, batch 8 sharded by image (BASELINE configs[1]); thr 0.7
    b0, b1 = shard(NMS_BATCH)
    nb = max(1, b1 - b0)
    for n in (2000, 5000, 10000, 20000):
        bs, ss = zip(*[synth.nms_boxes(n, 7 + b0 + i)[:2] for i in range(nb)])
        boxes, scores = T(np.concatenate(bs)), T(np.concatenate(ss))
        offs = T((np.arange(nb + 1) * n).astype(np.int32))
        for mk in (-1, 1000):
            ms["nms_n%d_max%d" % (n, mk)] = med(
                lambda: rp.nms_batched(boxes, scores, offs, 0.7, max_keep=mk, n_max=n)) if b1 > b0 else 0.0
        if n == 20000:
            b1_, s1_ = T(bs[0]), T(ss[0])
            o1 = T(np.array([0, n], np.int32))
            ms["nms_single_n20000"] = med(lambda: rp.nms_batched(b1_, s1_, o1, 0.7, n_max=n))
        del boxes, scores
    # --- Yolact post (FastNMS + assembly at 550x550), batch 64 sharded by image (configs[3])
    priors = synth.yolact_priors()
    y0, y1 = shard(YOLACT_BATCH)
    B = max(1, y1 - y0)
    base = []
    for i in range(4):  # decode + conf filter prepare the inputs; they are outside the timed op
        loc, conf, mask, proto = synth.yolact_image(100 + i, priors)
        boxes = rp.decode(T(loc), T(priors)).cpu().numpy()
        fg = np.ascontiguousarray(conf[:, 1:].T)
        kp = np.nonzero(fg.max(0) > np.float32(0.05))[0]
        base.append((boxes[kp], mask[kp], np.ascontiguousarray(fg[:, kp]), proto))
    cands = [base[(y0 + i) % 4] for i in range(B)]
    offs_np = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
    yb, yc = T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands]))
    ys = T(np.concatenate([c[2] for c in cands], 1))
    yp = T(np.stack([c[3] for c in cands]))
    yo = T(offs_np)
    masks_buf = torch.empty((B * 100, 550, 550), dtype=torch.uint8, device=dev)
    slot_off = T((np.arange(B + 1) * 100).astype(np.int32))
    n_max = int(np.diff(offs_np).max())

    def yolact_post():
        r = rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100, n_max=n_max)
        # padded layout: every image owns 100 detection slots (unused slots hold zero boxes)
        rp.mask_assemble(yp, r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4), 550, 550,
                         True, det_offsets=slot_off, out=masks_buf)
        return r

    ms["yolact_post"] = med(yolact_post, n=5, warm=2)
    ms["yolact_fast_nms"] = med(lambda: rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100))
    r = yolact_post()
    cf, bx = r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4)
    ms["mask_assemble"] = med(
        lambda: rp.mask_assemble(yp, cf, bx, 550, 550, True, det_offsets=slot_off, out=masks_buf), n=5, warm=2)
    keep["yolact"] = (r, masks_buf, y0, B)
    keep["yolact_cpu_inputs"] = base[0]
    del yp
    # --- Yolact Detect pre-stage from the raw head outputs (decode + class-max + filter + gathers)
    raw = [synth.yolact_image(100 + i, priors) for i in range(4)]
    loc64 = T(np.stack([raw[(y0 + i) % 4][0] for i in range(B)]))
    conf64 = T(np.stack([raw[(y0 + i) % 4][1] for i in range(B)]))
    mask64 = T(np.stack([raw[(y0 + i) % 4][2] for i in range(B)]))
    pri_t = T(priors)
    ms["yolact_prefilter"] = med(lambda: rp.detect_prefilter(loc64, conf64, mask64, pri_t, 0.05), n=5, warm=2)
    keep["prefilter_bytes_per_image"] = (conf64[0].numel() + loc64[0].numel()) * 4
    del loc64, conf64, mask64
    # --- paste-back, 100 instances per image at 1333x800, batch 128 sharded by image (configs[4]);
    #     one launch per image (the output of one image is 107 MB)
    p0, p1 = shard(PASTE_BATCH)
    pin = [synth.paste_inputs(100, 200 + (p0 + i) % 4) for i in range(min(4, max(1, p1 - p0)))]
    pts = [(T(m_), T(b_)) for m_, b_ in pin]
    pout = torch.empty((100, 800, 1333), dtype=torch.uint8, device=dev)

    def paste_shard():
        for i in range(p1 - p0):
            m_, b_ = pts[i % len(pts)]
            rp.paste_masks(m_, b_, 800, 1333, out=pout)

    ms["paste_shard"] = med(paste_shard, n=3, warm=1)
    ms["paste_one_image"] = med(lambda: rp.paste_masks(pts[0][0], pts[0][1], 800, 1333, out=pout))
    keep["paste_cpu_inputs"] = pin[0]
    # --- RPN proposal stage (8 images sharded) and box-head post-processor (8 images sharded)
    r0, r1 = shard(8)
    N, A = max(1, r1 - r0), 3
    lv = [synth.rpn_level(N, A, H, W, st, seed=100 + st + r0)
          for H, W, st in ((200, 336, 4), (100, 168, 8), (50, 84, 16), (25, 42, 32), (13, 21, 64))]
    im_wh = lv[0][3]
    anchors = [[rp.BoxList(T(l[0][b]), (int(im_wh[b, 0]), int(im_wh[b, 1]))) for l in lv] for b in range(N)]
    objs, regs, ancs, wh = [T(l[1]) for l in lv], [T(l[2]) for l in lv], [T(l[0]) for l in lv], T(im_wh)
    post = rp.RPNPostProcessor(1000, 1000, 0.7, 0, fpn_post_nms_top_n=1000)
    post.eval()

    def rpn_device():
        bx_, sc_, _, off_, ks = rp.rpn_pre_nms_levels(objs, regs, ancs, wh, 1000, 0.0)
        kp_, nk_ = rp.nms_batched(bx_, sc_, off_, 0.7, max_keep=1000, n_max=max(ks),
                                  flags=rp.package._lib.RP_NMS_PRESORTED)  # as RPNPostProcessor._launch does
        rp.rpn_select_levels(bx_, sc_, off_, kp_, nk_, len(lv), N, 1000)

    ms["rpn_post_device"] = med(rpn_device)
    ms["rpn_post_execute"] = med(lambda: post(anchors, objs, regs))
    lg, reg, pr, counts, bwh = synth.box_head_batch(N, 1000, seed=3 + r0)
    lgt, regt, prt, bwht = T(lg), T(reg), T(pr), T(bwh)
    boffs = T(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))
    ms["box_post_device"] = med(lambda: rp.box_post(lgt, regt, prt, boffs, bwht, 1000, 0.05, 0.5, 100))
    return ms, keep


def gather_run(rp, torch, dist, dev, rank, world, keep):
    """parallel.gather_detections on the Yolact shard outputs: kept detections + bit-packed 550x550
    masks to rank 0 (NCCL all_gather of counts, then of the payload).  Outside every timed figure;
    reported separately.  At world == 1 only the packing kernel runs."""
    r, masks_buf, y0, B = keep["yolact"]
    cnt = r["count"].cpu().numpy() if "count" in r else np.full((B,), 100)
    sel = torch.cat([torch.arange(b * 100, b * 100 + int(cnt[b]), device=dev) for b in range(B)]) \
        if B else torch.zeros((0,), dtype=torch.int64, device=dev)
    ids = torch.cat([torch.full((int(cnt[b]),), y0 + b, dtype=torch.int64, device=dev) for b in range(B)])
    cls = r["class"].reshape(B * 100)[sel]
    sc = r["score"].reshape(B * 100)[sel]
    bx = r["box"].reshape(B * 100, 4)[sel]
    mk = masks_buf[sel]
    m = int(sel.numel())
    pack_ms = _med(torch, lambda: rp.parallel.pack_masks(mk), n=5, warm=1)
    out = {"detections_this_rank": m, "mask_bytes_raw": int(mk.numel()),
           "payload_bytes_this_rank": int(m * (8 + 8 + 4 + 16) + m * ((550 * 550 + 7) // 8)),
           "pack_masks_ms": pack_ms, "pack_masks_gbs": mk.numel() * 1.125 / (pack_ms * 1e-3) / 1e9 if m else None}
    if world > 1:
        rp.parallel.gather_detections(ids, cls, sc, bx, mk)  # warm-up: NCCL sets its channels up lazily
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g = rp.parallel.gather_detections(ids, cls, sc, bx, mk)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["gather_ms"] = float(t.item()) * 1e3
        if rank == 0:
            out["gathered_detections"] = int(g["score"].shape[0])
            out["gathered_images"] = int(torch.unique(g["image_id"]).numel())
            ok = bool((g["mask"][:m] == mk).all().item()) if m else True
            out["gather_roundtrip_exact"] = ok
        out["backend"] = "nccl"
    else:
        out["gather_ms"] = None
        out["backend"] = "none (1 rank: nothing to exchange)"
    return out


def cpu_figures(keep, cores):
    """The oracle (CPU restatement) beside each figure, on a bounded sample, on rank 0's host cores."""
    import oracle
    out = {"cores_available": cores}
    # NMS: one image per call (the restatement is a scalar loop: 1 thread), 20k and 2k boxes
    for n in (2000, 20000):
        b, s, _ = synth.nms_boxes(n, 7)
        t0 = time.perf_counter()
        oracle.nms(b, s, 0.7)
        out["nms_ms_per_image_n%d" % n] = (time.perf_counter() - t0) * 1e3
    out["nms_threads"] = 1
    b, s, _ = synth.nms_boxes(5000, 7)
    dets = np.concatenate([b, s[:, None]], 1).astype(np.float32)
    t0 = time.perf_counter()
    oracle.cython_nms(dets, 0.7)
    out["cython_nms_ms_per_image_n5000"] = (time.perf_counter() - t0) * 1e3
    # Yolact post: FastNMS + assembly at 550x550 of one image (assembly is OpenMP-parallel over
    # detections inside the restatement)
    bxs, cf, sc, proto = keep["yolact_cpu_inputs"]
    t0 = time.perf_counter()
    ob, oc, _, _, _ = oracle.fast_nms(bxs, cf, sc, 0.5, 200, 100)
    t1 = time.perf_counter()
    oracle.mask_assemble(proto, oc, ob, 550, 550, crop=True)
    t2 = time.perf_counter()
    out["yolact_fast_nms_ms_per_image"] = (t1 - t0) * 1e3
    out["yolact_assemble_ms_per_image"] = (t2 - t1) * 1e3
    out["yolact_post_images_per_s"] = 1.0 / (t2 - t0)
    out["yolact_threads"] = int(os.environ.get("OMP_NUM_THREADS", cores))
    # paste-back: 100 instances at 1333x800, one image
    pm, pb = keep["paste_cpu_inputs"]
    t0 = time.perf_counter()
    oracle.paste_masks(pm, pb, 800, 1333)
    out["paste_ms_per_image_100inst"] = (time.perf_counter() - t0) * 1e3
    out["paste_threads"] = int(os.environ.get("OMP_NUM_THREADS", cores))
    out["kind"] = "port (oracle/roipath_oracle.c); cython_nms: the reference's own .pyx restated"
    return out


def derive(ms, world, images_per_gpu, keep):
    """Whole-job figures from the MAX-over-ranks times (ms) of the per-rank shards."""
    out = {"sharding": "by image over %d rank(s); times are the max over ranks" % world}
    shapes = synth.fpn_shapes(images_per_gpu, CHANNELS)
    grad_bytes = sum(int(np.prod(sh)) for sh in shapes) * 4
    for res, k in ((7, images_per_gpu * 512), (14, 2048)):
        t = ms["pooler_bwd_%dx%d" % (res, res)]
        out["pooler_bwd_%dx%d_ms" % (res, res)] = t
        out["pooler_bwd_%dx%d_gbs_per_gpu" % (res, res)] = (k * CHANNELS * res * res * 4 + grad_bytes) / (t * 1e-3) / 1e9
    out["pooler_fwd_nhwc_14x14_ms"] = ms["pooler_fwd_nhwc_14x14"]
    out["pooler_fwd_nhwc_layout"] = "features torch.channels_last [N,H,W,C] fp32, output [K,C,14,14]"
    per_rank = -(-NMS_BATCH // world)
    for n in (2000, 5000, 10000, 20000):
        for mk in (-1, 1000):
            t = ms["nms_n%d_max%d" % (n, mk)]
            # per image of the batch of 8: the shard of the slowest rank holds ceil(8 / world) images
            out["nms_ms_per_image_n%d_max%d" % (n, mk)] = t / per_rank
            out["nms_batch8_ms_n%d_max%d" % (n, mk)] = t
    out["nms_ms_single_image_n20000"] = ms["nms_single_n20000"]
    t = ms["yolact_post"]
    out["yolact_post_images_per_s"] = YOLACT_BATCH / (t * 1e-3)
    out["yolact_post_ms_per_batch64"] = t
    out["yolact_fast_nms_ms_per_batch64"] = ms["yolact_fast_nms"]
    ta = ms["mask_assemble"]
    out["mask_assemble_ms_per_batch64"] = ta
    out["mask_assemble_gbs"] = (YOLACT_BATCH * (138 * 138 * 32 * 4 + 100 * 32 * 4 + 1600)
                                + YOLACT_BATCH * 100 * 550 * 550) / (ta * 1e-3) / 1e9
    tf = ms["yolact_prefilter"]
    out["yolact_prefilter_ms_per_batch64"] = tf
    out["yolact_prefilter_gbs"] = YOLACT_BATCH * keep["prefilter_bytes_per_image"] / (tf * 1e-3) / 1e9
    tp = ms["paste_shard"]
    out["paste_batch128_ms"] = tp
    out["paste_images_per_s_batch128"] = PASTE_BATCH / (tp * 1e-3)
    out["paste_gbs"] = PASTE_BATCH * (100 * 28 * 28 * 4 + 100 * 800 * 1333) / (tp * 1e-3) / 1e9
    out["paste_ms_per_image_100inst"] = ms["paste_one_image"]
    out["rpn_post_device_ms_per_batch8"] = ms["rpn_post_device"]
    out["rpn_post_execute_ms_per_batch8"] = ms["rpn_post_execute"]
    out["box_post_device_ms_per_batch8"] = ms["box_post_device"]
    return out


def run(rp, torch, dist, dev, rank, world, images_per_gpu, cores):
    """All ranks call this; rank 0 gets the dict for the bench line, the others None."""
    ms, keep = local_times(rp, torch, dev, rank, world, images_per_gpu)
    keys = sorted(ms)
    t = torch.tensor([ms[k] for k in keys], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = dict(zip(keys, t.tolist()))
    gather = gather_run(rp, torch, dist, dev, rank, world, keep)
    if rank != 0:
        return None
    out = derive(ms_max, world, images_per_gpu, keep)
    out["gather"] = gather
    out["cpu"] = cpu_figures(keep, cores)
    return out
