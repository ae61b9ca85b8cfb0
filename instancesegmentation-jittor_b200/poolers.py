"""The multi-level FPN Pooler with the reference's interface, as ONE fused launch.

Mirrors detectron.jittor/detectron/modeling/poolers.py: LevelMapper :11-43,
LevelMapperwithArea :46-73, Pooler :76-156, make_pooler :159-168.

The reference loops over levels with nonzero -> gather -> ROIAlign -> scatter (4 host
synchronisations and 8 tensor passes per call).  Here the level of every RoI is computed on
the device (rp_level_map*), and one kernel pools every RoI from the map of its level and
writes it to its own output row; the backward pass scatters into all level gradients in one
launch as well.
"""
import ctypes
import math

import torch

from . import _lib

_I32 = torch.int32
_F32 = torch.float32


def convert_to_roi_format(boxes):
    """poolers.py:110-121: [sum K, 5] = (image index, x1, y1, x2, y2)."""
    dev = boxes[0].bbox.device
    concat = torch.cat([b.bbox.float() for b in boxes], dim=0)
    ids = torch.cat([torch.full((len(b), 1), float(i), dtype=_F32, device=dev)
                     for i, b in enumerate(boxes)], dim=0)
    return torch.cat([ids, concat], dim=1).contiguous()


def _level_map(rois, k_min, k_max, s0, lvl0, eps):
    K = rois.shape[0]
    levels = torch.empty((max(K, 1),), dtype=_I32, device=rois.device)
    if K:
        rc = _lib.lib().rp_level_map(_lib.dptr(rois, _F32, "rois"), K, int(k_min), int(k_max),
                                     float(s0), int(lvl0), float(eps), _lib.dptr(levels, _I32),
                                     _lib.stream_ptr(rois.device))
        _lib.check(rc, "rp_level_map")
    return levels[:K]


def _level_map_area(rois, img_area, k_min, k_max, eps):
    K = rois.shape[0]
    levels = torch.empty((max(K, 1),), dtype=_I32, device=rois.device)
    if K:
        rc = _lib.lib().rp_level_map_area(_lib.dptr(rois, _F32, "rois"),
                                          _lib.dptr(img_area, _F32, "img_area"), K, int(k_min),
                                          int(k_max), float(eps), _lib.dptr(levels, _I32),
                                          _lib.stream_ptr(rois.device))
        _lib.check(rc, "rp_level_map_area")
    return levels[:K]


class LevelMapper(object):
    """poolers.py:11-43 (FPN paper Eqn.(1); box area with the legacy +1)."""

    def __init__(self, k_min, k_max, canonical_scale=224, canonical_level=4, eps=1e-6):
        self.k_min = int(round(k_min, 4))
        self.k_max = int(round(k_max, 4))
        self.s0 = canonical_scale
        self.lvl0 = canonical_level
        self.eps = eps

    def __call__(self, boxlists):
        return self.map_rois(convert_to_roi_format(boxlists))

    def map_rois(self, rois):
        return _level_map(rois, self.k_min, self.k_max, self.s0, self.lvl0, self.eps)


class LevelMapperwithArea(object):
    """poolers.py:46-73 (CenterMask): ceil(k_max - log2(image_area / box_area + eps))."""

    def __init__(self, k_min, k_max, eps=1e-6):
        self.k_min = int(round(k_min, 4))
        self.k_max = int(round(k_max, 4))
        self.eps = eps

    def __call__(self, boxlists):
        rois = convert_to_roi_format(boxlists)
        img_area = torch.cat([b.image_area() for b in boxlists]).contiguous()
        return _level_map_area(rois, img_area, self.k_min, self.k_max, self.eps)


def _is_channels_last(t):
    """A [N,C,H,W] tensor whose memory is [N,H,W,C] (torch.channels_last) and not also NCHW-contiguous."""
    return (t.ndim == 4 and t.is_cuda and t.dtype == _F32 and not t.is_contiguous()
            and t.is_contiguous(memory_format=torch.channels_last))


def _host_arrays(feats, scales):
    L = len(feats)
    ptrs = (ctypes.c_void_p * L)(*[f.data_ptr() for f in feats])
    Hs = (ctypes.c_int * L)(*[int(f.shape[2]) for f in feats])
    Ws = (ctypes.c_int * L)(*[int(f.shape[3]) for f in feats])
    sc = (ctypes.c_float * L)(*[float(s) for s in scales])
    return ptrs, Hs, Ws, sc


def pooler_forward(feats, scales, rois, levels, output_size, sampling_ratio, mode=0, out=None):
    """feats: list of [N,C,H_l,W_l] f32 CUDA tensors; rois [K,5]; levels [K] int32 (or None
    when there is one level) -> [K,C,PH,PW]."""
    L = len(feats)
    if L < 1 or L > _lib.RP_MAX_LEVELS:
        raise ValueError("between 1 and %d levels" % _lib.RP_MAX_LEVELS)
    nhwc = all(_is_channels_last(f) for f in feats)
    if not nhwc:
        for f in feats:
            _lib.dptr(f, _F32, "feature map")
    N, C = int(feats[0].shape[0]), int(feats[0].shape[1])
    PH, PW = (output_size if isinstance(output_size, (tuple, list)) else (output_size, output_size))
    K = int(rois.shape[0])
    if out is None:
        out = torch.empty((K, C, int(PH), int(PW)), dtype=_F32, device=rois.device)
    if K == 0:
        return out
    ptrs, Hs, Ws, sc = _host_arrays(feats, scales)
    if nhwc:  # torch.channels_last feature maps: [N,H,W,C] in memory, 16-byte taps of 4 channels
        if mode != 0:
            raise ValueError("mode 1 (reference operand order) exists for NCHW feature maps only")
        rc = _lib.lib().rp_pooler_forward_nhwc(
            ptrs, Hs, Ws, sc, L, N, C, _lib.dptr(rois, _F32, "rois"), _lib.dptr(levels, _I32, "levels"),
            K, int(PH), int(PW), float(sampling_ratio), _lib.dptr(out, _F32, "out"),
            _lib.stream_ptr(rois.device))
        _lib.check(rc, "rp_pooler_forward_nhwc")
        return out
    rc = _lib.lib().rp_pooler_forward(
        ptrs, Hs, Ws, sc, L, N, C, _lib.dptr(rois, _F32, "rois"), _lib.dptr(levels, _I32, "levels"),
        K, int(PH), int(PW), float(sampling_ratio), _lib.dptr(out, _F32, "out"), int(mode),
        _lib.stream_ptr(rois.device))
    _lib.check(rc, "rp_pooler_forward")
    return out


def pooler_backward(dout, rois, levels, feat_shapes, scales, sampling_ratio, mode=0):
    """Adjoint of pooler_forward w.r.t. the feature maps: list of [N,C,H_l,W_l] gradients."""
    L = len(feat_shapes)
    dev = dout.device
    grads = [torch.empty(tuple(int(v) for v in s), dtype=_F32, device=dev) for s in feat_shapes]
    N, C = int(feat_shapes[0][0]), int(feat_shapes[0][1])
    K, _, PH, PW = dout.shape
    ptrs, Hs, Ws, sc = _host_arrays(grads, scales)
    rc = _lib.lib().rp_pooler_backward(
        _lib.dptr(dout, _F32, "dout"), _lib.dptr(rois, _F32, "rois"),
        _lib.dptr(levels, _I32, "levels"), int(K), Hs, Ws, sc, L, N, C, int(PH), int(PW),
        float(sampling_ratio), ptrs, int(mode), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_pooler_backward")
    return grads


class _PoolerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rois, levels, output_size, scales, sampling_ratio, *feats):
        feats = [_lib.as_f32(f, "feature map") for f in feats]
        ctx.save_for_backward(rois, levels)
        ctx.cfg = ([tuple(f.shape) for f in feats], tuple(scales), float(sampling_ratio))
        return pooler_forward(feats, scales, rois, levels, output_size, sampling_ratio)

    @staticmethod
    def backward(ctx, dout):
        rois, levels = ctx.saved_tensors
        shapes, scales, sr = ctx.cfg
        grads = pooler_backward(dout.contiguous(), rois, levels, shapes, scales, sr)
        return (None, None, None, None, None) + tuple(grads)


class Pooler(torch.nn.Module):
    """Drop-in for detectron.modeling.poolers.Pooler (poolers.py:76-156)."""

    def __init__(self, output_size, scales, sampling_ratio, lvl_map_func="MASKRCNNLevelMapFunc"):
        super().__init__()
        self.output_size = output_size
        self.scales = tuple(float(s) for s in scales)
        self.sampling_ratio = sampling_ratio
        # poolers.py:105-106 (the 4-decimal rounding is LevelMapper's :25-26 in the reference)
        lvl_min = -math.log(scales[0], 2.0)
        lvl_max = -math.log(scales[-1], 2.0)
        cls = LevelMapper if lvl_map_func == "MASKRCNNLevelMapFunc" else LevelMapperwithArea
        self.map_levels = cls(lvl_min, lvl_max)

    def convert_to_roi_format(self, boxes):
        return convert_to_roi_format(boxes)

    def forward(self, x, boxes):
        """x: list of per-level feature maps; boxes: list[BoxList] -> [sum K, C, R, R] in the
        original RoI order (poolers.py:123-156)."""
        num_levels = len(self.scales)
        rois = self.convert_to_roi_format(boxes)
        feats = list(x)[:num_levels]  # zip(x, poolers) silently drops extra maps (:148)
        if num_levels == 1:
            levels = torch.zeros((rois.shape[0],), dtype=_I32, device=rois.device)
        else:
            levels = self.map_levels(boxes)
        return _PoolerFn.apply(rois, levels, self.output_size, self.scales, self.sampling_ratio,
                               *feats)

    execute = forward

    def execute_host(self, x_host, boxes_host, out_host=None, device=None):
        """The same pooling for HOST-resident inputs, pipelined image by image.

        x_host: list of per-level CPU tensors [N,C,H_l,W_l] (pinned for asynchronous copies);
        boxes_host: list of N CPU tensors [k_i,4] (xyxy); out_host: optional pinned CPU tensor
        [sum k_i, C, R, R].  Three streams overlap the host->device copy of image i+1, the
        pooling launch of image i and the device->host copy of image i-1 (images are independent
        on this path), with two device staging buffers per direction.  Forward
        only; returns out_host after the last copy has finished."""
        if not isinstance(self.map_levels, LevelMapper):
            raise ValueError("execute_host supports the MASKRCNN level mapping only")
        dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        L = len(self.scales)
        feats = list(x_host)[:L]
        N, C = int(feats[0].shape[0]), int(feats[0].shape[1])
        PH, PW = (self.output_size if isinstance(self.output_size, (tuple, list))
                  else (self.output_size, self.output_size))
        counts = [int(b.shape[0]) for b in boxes_host]
        K, kmax = sum(counts), max(counts + [1])
        if out_host is None:
            out_host = torch.empty((K, C, PH, PW), dtype=_F32).pin_memory()
        key = (str(dev), tuple(tuple(f.shape[1:]) for f in feats), kmax, PH, PW)
        st = getattr(self, "_host_pipe", None)
        if st is None or st["key"] != key:
            st = {"key": key,
                  "feat": [[torch.empty((1,) + tuple(f.shape[1:]), dtype=_F32, device=dev) for f in feats]
                           for _ in range(2)],
                  "rois": [torch.zeros((kmax, 5), dtype=_F32, device=dev) for _ in range(2)],
                  "out": [torch.empty((kmax, C, PH, PW), dtype=_F32, device=dev) for _ in range(2)],
                  "s_in": torch.cuda.Stream(dev), "s_run": torch.cuda.Stream(dev),
                  "s_out": torch.cuda.Stream(dev)}
            self._host_pipe = st
        s_in, s_run, s_out = st["s_in"], st["s_run"], st["s_out"]
        cur = torch.cuda.current_stream(dev)
        for s in (s_in, s_run, s_out):
            s.wait_stream(cur)
        ev_in = [None, None]
        ev_run = [None, None]
        ev_out = [None, None]
        row = 0
        for i in range(N):
            b = i & 1
            k = counts[i]
            with torch.cuda.stream(s_in):
                if ev_run[b] is not None:
                    s_in.wait_event(ev_run[b])  # the staging maps of slot b have been consumed
                for l in range(L):
                    st["feat"][b][l].copy_(feats[l][i:i + 1], non_blocking=True)
                if k:
                    st["rois"][b][:k, 1:].copy_(boxes_host[i], non_blocking=True)
                ev_in[b] = s_in.record_event()
            with torch.cuda.stream(s_run):
                s_run.wait_event(ev_in[b])
                if ev_out[b] is not None:
                    s_run.wait_event(ev_out[b])  # the staging output of slot b has been copied out
                if k:
                    rois = st["rois"][b][:k]
                    levels = (self.map_levels.map_rois(rois) if L > 1
                              else torch.zeros((k,), dtype=_I32, device=dev))
                    pooler_forward(st["feat"][b], self.scales, rois, levels, (PH, PW),
                                   self.sampling_ratio, out=st["out"][b][:k])
                ev_run[b] = s_run.record_event()
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_run[b])
                if k:
                    out_host[row:row + k].copy_(st["out"][b][:k], non_blocking=True)
                ev_out[b] = s_out.record_event()
            row += k
        for s in (s_in, s_run, s_out):
            cur.wait_stream(s)
        s_out.synchronize()
        return out_host


def make_pooler(resolution, scales, sampling_ratio):
    """poolers.py:159-168 without the yacs cfg object (not installed here): pass
    cfg.MODEL[head].POOLER_RESOLUTION / POOLER_SCALES / POOLER_SAMPLING_RATIO directly."""
    return Pooler(output_size=(resolution, resolution), scales=scales, sampling_ratio=sampling_ratio)
