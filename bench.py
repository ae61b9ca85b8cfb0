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

    for _ in range(args.warmup):
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
        per_rank = [te.clone() for _ in range(world)]
        if world > 1:
            dist.all_gather(per_rank, te)
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        h2d = int(sum(hf.numel() * 4 for hf in h_feats) + h_rois[:, 1:].numel() * 4)
        d2h = int(h_out.numel() * 4)

        # host ceiling: the same pinned H2D + D2H bytes on two streams with NO kernel, all ranks at once
        d_out_tmp = out
        s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

        def copies_only():
            with torch.cuda.stream(s_in):
                for hf, f in zip(h_feats, feats):
                    f.copy_(hf, non_blocking=True)
            with torch.cuda.stream(s_out):
                h_out.copy_(d_out_tmp, non_blocking=True)
            torch.cuda.current_stream(dev).wait_stream(s_in)
            torch.cuda.current_stream(dev).wait_stream(s_out)

        copies_only()
        torch.cuda.synchronize()
        ms_copy = time_steps(copies_only, n_e2e, None, barrier)
        tc = torch.tensor([ms_copy], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        e2e = {"value": world * K * n_e2e / (float(te.item()) * 1e-3), "unit": "RoIs/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": n_e2e,
               "per_rank_rois_per_s": [K * n_e2e / (float(t.item()) * 1e-3) for t in per_rank],
               "host_ceiling": {
                   "what": "the same pinned H2D + D2H bytes per rank on two streams, no kernel, all ranks "
                           "at once (max over ranks)",
                   "rois_per_s": world * K * n_e2e / (float(tc.item()) * 1e-3),
                   "aggregate_gbs": world * (h2d + d2h) * n_e2e / (float(tc.item()) * 1e-3) / 1e9}}
        del h_feats, h_out

    # ---- the other figures of the compound metric, sharded by image over the ranks (every N)
    extra = None
    cores = len(os.sched_getaffinity(0))
    if not args.no_extras:
        import bench_extras
        try:
            extra = bench_extras.run(rp, torch, dist, dev, rank, world, IMAGES_PER_GPU, cores)
        except Exception as ex:  # the headline line must still be printed
            extra = {"error": repr(ex)}
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
    # the kernel the forward call launches: the SIMT column walk by default, the tensor-core kernel when
    # RP_ROI_IMPL=6 is set; `traffic` is that kernel's ncu DRAM bytes per launch (same workload), taken
    # from the round's capture listed in profiles/roofline_traffic.json -- null without a capture
    impl = os.environ.get("RP_ROI_IMPL", "3")
    kernel_name = {"3": "rt_pool_fwd_kernel", "4": "rb_pool_fwd_kernel", "5": "rg_pool_fwd_kernel",
                   "6": "rx_pool_fwd_kernel", "7": "rm_pool_fwd_kernel"}.get(impl, "rt_pool_fwd_kernel")
    planes = impl == "3" and os.environ.get("RP_ROI_PLANES", "1") != "0"
    if planes:  # the call is three launches: sample tables, whole-plane kernel (coarse levels), column walk
        kernel_name = "rt_pool_fwd_kernel+rp_plane_fwd_kernel"
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = tj.get(kernel_name + "_dram_bytes_per_launch")
        traffic_src = tj.get(kernel_name + "_capture")
    roofline = {"bound": "hbm", "kernel": kernel_name, "traffic_source": traffic_src, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "output_bytes_per_launch": out_bytes,
                "distinct_feature_bytes": 4 * U, "kernel_ms": ms_kernel,
                "output_only_gbs": out_bytes / (ms_kernel * 1e-3) / 1e9,
                "level_histogram": np.bincount(lv_np, minlength=4).tolist()}

    # ---- CPU baseline on this box's host cores (bounded sample)
    cf, cr, cl = cpu_sample(1)
    dt, kind, cpu_out = cpu_pooler(cf, cr, cl, cores)
    dt2, _, _ = cpu_pooler(cf, cr, cl, cores)
    dt = min(dt, dt2)
    cpu_baseline = {"value": cr.shape[0] / dt, "unit": "RoIs/s", "cores": cores, "kind": kind,
                    "sample": "1 image x %d RoIs of the same workload, best of 2" % ROIS_PER_IMAGE}
    # ---- parity at the headline configuration: the GPU path on the CPU arm's own inputs (1 image x
    #      1000 RoIs x 256 channels x 14x14) against the CPU arm's output (reference kernel text when
    #      kind == "reference"); north_star bar: 1e-5 relative in fp32
    g_feats = [torch.from_numpy(f).to(dev) for f in cf]
    g_rois = torch.from_numpy(cr).to(dev)
    g_out = rp.pooler_forward(g_feats, synth.FPN_SCALES, g_rois, mapper.map_rois(g_rois), (RES, RES), SAMPLING)
    g_np = g_out.cpu().numpy()
    err = np.abs(g_np - cpu_out)
    tol = 1e-5 * np.abs(cpu_out) + 1e-5
    parity = {"against": kind, "n": int(err.size), "max_abs": float(err.max()),
              "max_rel_where_abs_ref_ge_0.1": float((err / np.abs(cpu_out))[np.abs(cpu_out) >= 0.1].max()),
              "violations": int((err > tol).sum()), "tol": "1e-5*|ref| + 1e-5",
              "levels_equal": bool(np.array_equal(mapper.map_rois(g_rois).cpu().numpy(), cl))}
    del g_feats, g_out

    line = {
        "metric": METRIC, "value": value, "unit": "RoIs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world), "roofline": roofline, "cpu_baseline": cpu_baseline,
        "parity": parity, "e2e": e2e, "gpu_launches": (4 if planes else 2) * args.steps, "clocks": clocks,
    }
    if extra is not None:
        line["extra"] = extra
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
