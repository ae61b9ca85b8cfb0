"""bench.py -- the hot path's headline metric on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[2], the configuration the ROIAlign metric is quoted on): the
Mask R-CNN R-50-FPN mask-head RoI path -- Pooler forward, 14x14x256, sampling_ratio 2, over a
batch of 16 images at 800x1333 (padded 800x1344; P2..P5 feature maps, 256 channels, fp32) with
1000 proposals per image.  One "step" = one pass of Pooler.execute over the batch: the level
mapping launch plus the fused pooling launch (16 000 RoIs).  Images are sharded across GPUs
with no data-path collective (weak scaling: every rank owns 16 images); value = RoIs/s of the
whole job.  Inputs are resident in HBM for `value`; `e2e` times the same call from pinned HOST
buffers with the H2D of the feature maps / RoIs and the D2H of the pooled tensor inside the
timed region.  The NMS / Yolact / paste-back figures of the compound metric are measured
after the timed region and reported under "extra".
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from tests import synth  # noqa: E402  (seeded synthetic inputs shared with the tests)

IMAGES_PER_GPU = 16
ROIS_PER_IMAGE = 1000
CHANNELS = 256
RES = 14
SAMPLING = 2
METRIC = "ROIAlign RoIs/s (14x14x256)"


def workload_config(n_gpus):
    return {
        "workload": "Mask R-CNN R-50-FPN mask-head RoI path: Pooler fwd 14x14x256, sampling 2, "
                    "batch 16 images/GPU at 800x1333, 1000 RoIs/image (BASELINE configs[2])",
        "images_per_gpu": IMAGES_PER_GPU, "rois_per_image": ROIS_PER_IMAGE,
        "channels": CHANNELS, "resolution": RES, "sampling_ratio": SAMPLING,
        "layout": "NCHW fp32", "sharding": "images across %d GPU(s), no collective" % n_gpus,
        "l2": "inputs (1.46 GB features) + output (3.2 GB) per step exceed the 126 MB L2",
    }


# ------------------------------------------------------------------------------------------
# algorithmic bytes: 4*K*C*PH*PW (output) + 20*K (rois) + 4*U (distinct
# feature elements touched).  U is counted here with numpy from the sampling lattice.
# ------------------------------------------------------------------------------------------
def _axis_taps(start, binsz, pooled, grid, extent):
    p = np.repeat(np.arange(pooled, dtype=np.float32), grid)
    i = np.tile(np.arange(grid, dtype=np.float32), pooled)
    v = np.float32(start) + p * np.float32(binsz) + (i + np.float32(.5)) * np.float32(binsz) / np.float32(grid)
    ok = ~((v < -1.0) | (v > extent))
    v = np.maximum(v[ok], 0)
    lo = np.minimum(v.astype(np.int64), extent - 1)
    hi = np.minimum(lo + 1, extent - 1)
    return np.unique(np.concatenate([lo, hi]))


def distinct_elements_touched(rois, levels, shapes, scales, res, sampling):
    marks = [np.zeros((s[0], s[2], s[3]), bool) for s in shapes]
    for r, l in zip(rois, levels):
        b = int(r[0])
        sc = np.float32(scales[l])
        H, W = shapes[l][2], shapes[l][3]
        x1, y1, x2, y2 = [np.float32(v) * sc for v in r[1:]]
        rw = max(float(x2 - x1), 1.0)
        rh = max(float(y2 - y1), 1.0)
        ys = _axis_taps(y1, np.float32(rh) / np.float32(res), res, sampling, H)
        xs = _axis_taps(x1, np.float32(rw) / np.float32(res), res, sampling, W)
        if len(ys) and len(xs):
            marks[l][b][np.ix_(ys, xs)] = True
    return int(sum(int(m.sum()) for m in marks)) * shapes[0][1]


def numpy_levels(boxes, k_min=2, k_max=5, s0=224.0, lvl0=4, eps=1e-6):
    area = (boxes[:, 2] - boxes[:, 0] + 1) * (boxes[:, 3] - boxes[:, 1] + 1)
    t = np.sqrt(area.astype(np.float32)) / np.float32(s0) + np.float32(eps)
    lv = np.floor(lvl0 + np.log2(t.astype(np.float64)))
    return (np.clip(lv, k_min, k_max) - k_min).astype(np.int32)


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own ROIAlign kernel text compiled as C++ (oracle/_ref), else the port
# ------------------------------------------------------------------------------------------
def cpu_pooler(feats, rois, levels, threads):
    """Pooler forward on the host cores: RoIs split across `threads` python threads, each
    calling the (GIL-releasing) C library on its slice."""
    import oracle
    ref = oracle.ref_lib()
    kind = "reference" if ref is not None else "port"
    fwd = oracle.ref_roi_align_forward if ref is not None else oracle.roi_align_forward
    K = rois.shape[0]
    out = np.zeros((K, feats[0].shape[1], RES, RES), np.float32)
    if ref is None:  # the port is OpenMP-parallel inside: one call per level
        t0 = time.perf_counter()
        for l in range(len(feats)):
            idx = np.nonzero(levels == l)[0]
            if len(idx):
                out[idx] = fwd(feats[l], rois[idx], (RES, RES), synth.FPN_SCALES[l], SAMPLING)
        return time.perf_counter() - t0, kind, out
    jobs = []
    for l in range(len(feats)):
        idx = np.nonzero(levels == l)[0]
        for part in np.array_split(idx, max(1, threads)):
            if len(part):
                jobs.append((l, part))

    def work(job):
        l, part = job
        out[part] = fwd(feats[l], rois[part], (RES, RES), synth.FPN_SCALES[l], SAMPLING)

    t0 = time.perf_counter()
    ts = [threading.Thread(target=lambda js=jobs[i::threads]: [work(j) for j in js]) for i in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    return time.perf_counter() - t0, kind, out


def cpu_sample(n_images=1):
    feats = synth.fpn_features(n_images, CHANNELS)
    rois = synth.rois_for_images(n_images, ROIS_PER_IMAGE)
    levels = numpy_levels(rois[:, 1:])
    return feats, rois, levels


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = len(os.sched_getaffinity(0))
    feats, rois, levels = cpu_sample(1)
    sample = "1 image x %d RoIs of the same workload per step" % ROIS_PER_IMAGE
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_pooler(feats, rois, levels, cores)
    steps = max(1, min(args.steps, 5))
    times = []
    kind = "port"
    for _ in range(steps):
        dt, kind, _ = cpu_pooler(feats, rois, levels, cores)
        times.append(dt)
    total = float(sum(times))
    value = rois.shape[0] * steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "RoIs/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": value, "unit": "RoIs/s", "cores": cores, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "RoIs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def time_steps(fn, steps, stream_sync, barrier):
    import torch
    barrier()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1)  # ms


def extras(rp, torch, dev):
    """The other figures of the compound metric, each timed with CUDA events (median of 10)."""
    out = {}

    def med(fn, n=10, warm=3):
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

    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    # --- Pooler backward, train shape of BASELINE configs[2]: 16 images x 512 RoIs (7x7, box head)
    #     and 2048 positives (14x14, mask head); bytes = dout read + every level gradient written
    shapes = synth.fpn_shapes(IMAGES_PER_GPU, CHANNELS)
    rois_b = T(synth.rois_for_images(IMAGES_PER_GPU, 512))
    lv_b = rp.LevelMapper(2, 5).map_rois(rois_b)
    grad_bytes = sum(int(np.prod(sh)) for sh in shapes) * 4
    for res, k in ((7, int(rois_b.shape[0])), (14, 2048)):
        d = torch.randn((k, CHANNELS, res, res), device=dev)
        ms = med(lambda: rp.pooler_backward(d, rois_b[:k], lv_b[:k], shapes, synth.FPN_SCALES, SAMPLING), n=5)
        out["pooler_bwd_%dx%d_ms" % (res, res)] = ms
        out["pooler_bwd_%dx%d_gbs" % (res, res)] = (d.numel() * 4 + grad_bytes) / (ms * 1e-3) / 1e9
        del d
    # --- NMS sweep, batch 8 (BASELINE configs[1]); thr 0.7
    for n in (2000, 5000, 10000, 20000):
        B = 8
        bs, ss = zip(*[synth.nms_boxes(n, 7 + i)[:2] for i in range(B)])
        boxes, scores = T(np.concatenate(bs)), T(np.concatenate(ss))
        offs = T((np.arange(B + 1) * n).astype(np.int32))
        for mk in (-1, 1000):
            ms = med(lambda: rp.nms_batched(boxes, scores, offs, 0.7, max_keep=mk, n_max=n))
            out["nms_ms_per_image_n%d_max%d" % (n, mk)] = ms / B
        if n == 20000:
            b1, s1 = T(bs[0]), T(ss[0])
            o1 = T(np.array([0, n], np.int32))
            out["nms_ms_single_image_n20000"] = med(lambda: rp.nms_batched(b1, s1, o1, 0.7, n_max=n))
        del boxes, scores
    # --- Yolact post (FastNMS + assembly at 550x550), batch 64 (BASELINE configs[3])
    priors = synth.yolact_priors()
    B = 64
    base = []
    for i in range(4):  # decode + conf filter prepare the inputs; they are outside the timed op
        loc, conf, mask, proto = synth.yolact_image(100 + i, priors)
        boxes = rp.decode(T(loc), T(priors)).cpu().numpy()
        fg = np.ascontiguousarray(conf[:, 1:].T)
        kp = np.nonzero(fg.max(0) > np.float32(0.05))[0]
        base.append((boxes[kp], mask[kp], np.ascontiguousarray(fg[:, kp]), proto))
    cands = [base[i % 4] for i in range(B)]
    offs_np = np.concatenate([[0], np.cumsum([c[0].shape[0] for c in cands])]).astype(np.int32)
    yb, yc = T(np.concatenate([c[0] for c in cands])), T(np.concatenate([c[1] for c in cands]))
    ys = T(np.concatenate([c[2] for c in cands], 1))
    yp = T(np.stack([c[3] for c in cands]))
    yo = T(offs_np)
    masks_buf = torch.empty((B * 100, 550, 550), dtype=torch.uint8, device=dev)
    slot_off = T((np.arange(B + 1) * 100).astype(np.int32))

    def yolact_post():
        r = rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100, n_max=int(np.diff(offs_np).max()))
        # padded layout: every image owns 100 detection slots (unused slots hold zero boxes)
        rp.mask_assemble(yp, r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4), 550, 550,
                         True, det_offsets=slot_off, out=masks_buf)

    ms = med(yolact_post, n=5, warm=2)
    out["yolact_post_images_per_s"] = B / (ms * 1e-3)
    out["yolact_post_ms_per_batch64"] = ms
    r = rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100)
    out["yolact_fast_nms_ms_per_batch64"] = med(lambda: rp.fast_nms_batched(yb, yc, ys, yo, 0.5, 200, 100))
    cf, bx = r["mask"].reshape(B * 100, -1), r["box"].reshape(B * 100, 4)
    ms_a = med(lambda: rp.mask_assemble(yp, cf, bx, 550, 550, True, det_offsets=slot_off, out=masks_buf), n=5, warm=2)
    out["mask_assemble_ms_per_batch64"] = ms_a
    out["mask_assemble_gbs"] = (B * (138 * 138 * 32 * 4 + 100 * 32 * 4 + 1600) + B * 100 * 550 * 550) / (ms_a * 1e-3) / 1e9
    del masks_buf, yp
    # --- Yolact Detect pre-stage from the raw head outputs, batch 64:
    #     decode + class-max + conf filter + gathers over the 399 MB conf tensor
    raw = [synth.yolact_image(100 + i, priors) for i in range(4)]
    loc64 = T(np.stack([raw[i % 4][0] for i in range(B)]))
    conf64 = T(np.stack([raw[i % 4][1] for i in range(B)]))
    mask64 = T(np.stack([raw[i % 4][2] for i in range(B)]))
    pri_t = T(priors)
    ms_f = med(lambda: rp.detect_prefilter(loc64, conf64, mask64, pri_t, 0.05), n=5, warm=2)
    out["yolact_prefilter_ms_per_batch64"] = ms_f
    out["yolact_prefilter_gbs"] = (conf64.numel() + loc64.numel()) * 4 / (ms_f * 1e-3) / 1e9
    del loc64, conf64, mask64
    # --- paste-back, 100 instances at 1333x800 (BASELINE configs[4], one image per launch)
    pm, pb = synth.paste_inputs(100, 200)
    pmt, pbt = T(pm), T(pb)
    pout = torch.empty((100, 800, 1333), dtype=torch.uint8, device=dev)
    ms_p = med(lambda: rp.paste_masks(pmt, pbt, 800, 1333, out=pout))
    out["paste_ms_per_image_100inst"] = ms_p
    out["paste_gbs"] = (100 * 28 * 28 * 4 + 100 * 800 * 1333) / (ms_p * 1e-3) / 1e9
    # --- RPN proposal stage: 5 FPN levels of 8 images at 800 x 1344, test
    #     configuration (pre/post/fpn top-n 1000, thr 0.7): pre-NMS + NMS + per-image selection,
    #     device part (no host read) and the whole RPNPostProcessor.execute (BoxLists out)
    N, A = 8, 3
    lv = [synth.rpn_level(N, A, H, W, st, seed=100 + st)
          for H, W, st in ((200, 336, 4), (100, 168, 8), (50, 84, 16), (25, 42, 32), (13, 21, 64))]
    im_wh = lv[0][3]
    anchors = [[rp.BoxList(T(l[0][b]), (int(im_wh[b, 0]), int(im_wh[b, 1]))) for l in lv] for b in range(N)]
    objs, regs, ancs, wh = [T(l[1]) for l in lv], [T(l[2]) for l in lv], [T(l[0]) for l in lv], T(im_wh)
    post = rp.RPNPostProcessor(1000, 1000, 0.7, 0, fpn_post_nms_top_n=1000)
    post.eval()

    def rpn_device():
        bx, sc, _, off, ks = rp.rpn_pre_nms_levels(objs, regs, ancs, wh, 1000, 0.0)
        keep, nk = rp.nms_batched(bx, sc, off, 0.7, max_keep=1000, n_max=max(ks))
        rp.rpn_select_levels(bx, sc, off, keep, nk, len(lv), N, 1000)

    out["rpn_post_device_ms_per_batch8"] = med(rpn_device)
    out["rpn_post_execute_ms_per_batch8"] = med(lambda: post(anchors, objs, regs))
    # --- box-head post-processor: 8 images x 1000 proposals x 81 classes
    lg, reg, pr, counts, bwh = synth.box_head_batch(8, 1000, seed=3)
    lgt, regt, prt, bwht = T(lg), T(reg), T(pr), T(bwh)
    boffs = T(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))
    out["box_post_device_ms_per_batch8"] = med(
        lambda: rp.box_post(lgt, regt, prt, boffs, bwht, 1000, 0.05, 0.5, 100))
    return out


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    import roipath_b200 as rp

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        barrier = lambda: dist.barrier()
    else:
        barrier = lambda: None
    rp.lib()

    # ---- inputs for this rank's shard of images (seeds depend on the global image index)
    shapes = synth.fpn_shapes(IMAGES_PER_GPU, CHANNELS)
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    feats = [torch.randn(s, device=dev, dtype=torch.float32, generator=gen) for s in shapes]
    rois_np = synth.rois_for_images(IMAGES_PER_GPU, ROIS_PER_IMAGE, seed=4321 + rank * IMAGES_PER_GPU)
    rois = torch.from_numpy(rois_np).to(dev)
    K = rois.shape[0]
    mapper = rp.LevelMapper(2, 5)
    out = torch.empty((K, CHANNELS, RES, RES), dtype=torch.float32, device=dev)

    def step():
        levels = mapper.map_rois(rois)
        rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (RES, RES), SAMPLING, out=out)

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = time_steps(step, args.steps, None, barrier)
    clocks = sampler.stop()
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * K * args.steps / (ms_max * 1e-3)

    # ---- dominant kernel alone (the fused pooling launch), for the roofline
    levels = mapper.map_rois(rois)
    kern = lambda: rp.pooler_forward(feats, synth.FPN_SCALES, rois, levels, (RES, RES), SAMPLING, out=out)
    ms_kernel = time_steps(kern, args.steps, None, barrier) / args.steps

    # ---- e2e: host pinned buffers in, pooled tensor out to the host, copies inside the region
    e2e = None
    if rank == 0 or world > 1:
        h_feats = [torch.empty(s, dtype=torch.float32).pin_memory() for s in shapes]
        for hf, f in zip(h_feats, feats):
            hf.copy_(f)
        h_rois = torch.from_numpy(rois_np).pin_memory()
        h_out = torch.empty((K, CHANNELS, RES, RES), dtype=torch.float32).pin_memory()
        boxes_host = [h_rois[h_rois[:, 0] == i][:, 1:] for i in range(IMAGES_PER_GPU)]
        pooler = rp.Pooler((RES, RES), synth.FPN_SCALES, SAMPLING)

        boxes_host = [b.contiguous().pin_memory() for b in boxes_host]

        def e2e_step():
            # Pooler.execute_host: host buffers in, host buffer out; per image the H2D copy, the
            # pooling launch and the D2H copy run on three streams (all inside the timed region)
            with torch.no_grad():
                pooler.execute_host(h_feats, boxes_host, h_out)

        e2e_step()
        torch.cuda.synchronize()
        n_e2e = max(1, min(args.steps, 3))
        ms_e2e = time_steps(e2e_step, n_e2e, None, barrier)
        te = torch.tensor([ms_e2e], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        h2d = int(sum(hf.numel() * 4 for hf in h_feats) + h_rois[:, 1:].numel() * 4)
        d2h = int(h_out.numel() * 4)
        e2e = {"value": world * K * n_e2e / (float(te.item()) * 1e-3), "unit": "RoIs/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": n_e2e}
        del h_feats, h_out

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak = float(json.load(open(peaks_path))["hbm_gbs"])
        peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    lv_np = levels.cpu().numpy()
    U = distinct_elements_touched(rois_np, lv_np, shapes, synth.FPN_SCALES, RES, SAMPLING)
    out_bytes = 4 * K * CHANNELS * RES * RES
    alg_bytes = out_bytes + 20 * K + 4 * U
    achieved = alg_bytes / (ms_kernel * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get("rt_pool_fwd_kernel_dram_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "rt_pool_fwd_kernel", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "output_bytes_per_launch": out_bytes,
                "distinct_feature_bytes": 4 * U, "kernel_ms": ms_kernel,
                "output_only_gbs": out_bytes / (ms_kernel * 1e-3) / 1e9,
                "level_histogram": np.bincount(lv_np, minlength=4).tolist()}

    # ---- CPU baseline on this box's host cores (bounded sample)
    cores = len(os.sched_getaffinity(0))
    cf, cr, cl = cpu_sample(1)
    dt, kind, cpu_out = cpu_pooler(cf, cr, cl, cores)
    dt2, _, _ = cpu_pooler(cf, cr, cl, cores)
    dt = min(dt, dt2)
    cpu_baseline = {"value": cr.shape[0] / dt, "unit": "RoIs/s", "cores": cores, "kind": kind,
                    "sample": "1 image x %d RoIs of the same workload, best of 2" % ROIS_PER_IMAGE}

    line = {
        "metric": METRIC, "value": value, "unit": "RoIs/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world), "roofline": roofline, "cpu_baseline": cpu_baseline,
        "e2e": e2e, "gpu_launches": 2 * args.steps, "clocks": clocks,
    }
    if not args.no_extras and world == 1:
        try:
            line["extra"] = extras(rp, torch, dev)
        except Exception as ex:  # the headline line must still be printed
            line["extra"] = {"error": repr(ex)}
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
