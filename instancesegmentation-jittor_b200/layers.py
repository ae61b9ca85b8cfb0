"""ROIAlign and NMS with the reference's call signatures, on torch CUDA tensors.

Mirrors detectron.jittor/detectron/layers/roi_align.py:298-328 (roi_align, ROIAlign) and
detectron.jittor/detectron/layers/nms.py:12-40 (nms, ml_nms).  Every call goes through the
C ABI of libroipath_b200.so; there is no torch / CPU fallback.
"""
import ctypes

import torch

from . import _lib

_I32 = torch.int32
_F32 = torch.float32


def _pair(v):
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


# --------------------------------------------------------------------------- ROIAlign
def roi_align_forward(input, rois, output_size, spatial_scale, sampling_ratio, mode=0, out=None):
    """feat [N,C,H,W] f32, rois [K,5] f32 -> [K,C,PH,PW] f32 (reference: jt.code forward)."""
    PH, PW = _pair(output_size)
    N, C, H, W = input.shape
    K = rois.shape[0]
    if out is None:
        out = torch.empty((K, C, PH, PW), dtype=_F32, device=input.device)
    if K == 0:  # roi_align.py:304-305
        return out
    if (input.ndim == 4 and not input.is_contiguous()
            and input.is_contiguous(memory_format=torch.channels_last) and mode == 0):
        # torch.channels_last input: the NHWC entry (16-byte taps of 4 channels), output stays [K,C,PH,PW]
        import ctypes
        rc = _lib.lib().rp_roi_align_forward_nhwc(
            ctypes.c_void_p(input.data_ptr()), N, C, H, W, _lib.dptr(rois, _F32, "rois"), K, PH, PW,
            float(spatial_scale), float(sampling_ratio), _lib.dptr(out, _F32, "out"),
            _lib.stream_ptr(input.device))
        _lib.check(rc, "rp_roi_align_forward_nhwc")
        return out
    rc = _lib.lib().rp_roi_align_forward(
        _lib.dptr(input, _F32, "input"), N, C, H, W, _lib.dptr(rois, _F32, "rois"), K, PH, PW,
        float(spatial_scale), float(sampling_ratio), _lib.dptr(out, _F32, "out"), int(mode),
        _lib.stream_ptr(input.device))
    _lib.check(rc, "rp_roi_align_forward")
    return out


def roi_align_backward(dout, rois, input_shape, spatial_scale, sampling_ratio, mode=0):
    """dout [K,C,PH,PW] -> grad_input [N,C,H,W] (reference: cuda_grad_src[0])."""
    N, C, H, W = [int(v) for v in input_shape]
    K, _, PH, PW = dout.shape
    grad = torch.empty((N, C, H, W), dtype=_F32, device=dout.device)
    rc = _lib.lib().rp_roi_align_backward(
        _lib.dptr(dout, _F32, "dout"), _lib.dptr(rois, _F32, "rois"), K, N, C, H, W, PH, PW,
        float(spatial_scale), float(sampling_ratio), _lib.dptr(grad, _F32, "grad_input"),
        int(mode), _lib.stream_ptr(dout.device))
    _lib.check(rc, "rp_roi_align_backward")
    return grad


class _ROIAlignFn(torch.autograd.Function):
    # gradient only w.r.t. input: cuda_grad_src = [K2, '', ''] (roi_align.py:189,296)
    @staticmethod
    def forward(ctx, input, rois, output_size, spatial_scale, sampling_ratio):
        input = _lib.as_f32(input, "input")
        rois = _lib.as_f32(rois, "rois")
        ctx.save_for_backward(rois)
        ctx.cfg = (tuple(input.shape), float(spatial_scale), float(sampling_ratio))
        return roi_align_forward(input, rois, output_size, spatial_scale, sampling_ratio)

    @staticmethod
    def backward(ctx, dout):
        (rois,) = ctx.saved_tensors
        shape, scale, sr = ctx.cfg
        grad = roi_align_backward(dout.contiguous(), rois, shape, scale, sr)
        return grad, None, None, None, None


def roi_align(input, rois, output_size, spatial_scale, sampling_ratio):
    """Drop-in for detectron.layers.roi_align (roi_align.py:298-307)."""
    return _ROIAlignFn.apply(input, rois, output_size, spatial_scale, sampling_ratio)


class ROIAlign(torch.nn.Module):
    """Drop-in for detectron.layers.ROIAlign (roi_align.py:310-328)."""

    def __init__(self, output_size, spatial_scale, sampling_ratio):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale
        self.sampling_ratio = sampling_ratio

    def forward(self, input, rois):
        return roi_align(input, rois, self.output_size, self.spatial_scale, self.sampling_ratio)

    execute = forward  # the reference's jittor.Module entry point name

    def __repr__(self):
        return "%s(output_size=%s, spatial_scale=%s, sampling_ratio=%s)" % (
            self.__class__.__name__, self.output_size, self.spatial_scale, self.sampling_ratio)


# --------------------------------------------------------------------------- NMS
def nms_batched(boxes, scores, offsets, threshold, labels=None, max_keep=-1, flags=0,
                n_max=None, offsets_host=None):
    """B independent greedy-NMS problems in one set of launches, NO host synchronisation.

    boxes [n,4] f32, scores [n] f32, labels [n] f32 or None, offsets [B+1] int32 (device).
    Returns (keep [n] int32, n_keep [B] int32): problem b's kept indices, relative to
    offsets[b], are keep[offsets[b] : offsets[b] + n_keep[b]].
    n_max (max problem size) may be passed to avoid reading offsets on the host.
    """
    n_total = int(boxes.shape[0])
    B = int(offsets.shape[0]) - 1
    dev = boxes.device
    if n_max is None:
        oh = offsets_host if offsets_host is not None else offsets.cpu()
        n_max = int((oh[1:] - oh[:-1]).max()) if B > 0 else 0
    keep = torch.empty((max(n_total, 1),), dtype=_I32, device=dev)
    n_keep = torch.empty((B,), dtype=_I32, device=dev)
    L = _lib.lib()
    nbytes = L.rp_nms_workspace_bytes(n_total, int(n_max), B)
    ws = _lib.workspace(nbytes, dev, "nms")
    rc = L.rp_nms_batched(
        _lib.dptr(boxes, _F32, "boxes"), _lib.dptr(scores, _F32, "scores"),
        _lib.dptr(labels, _F32, "labels"), _lib.dptr(offsets, _I32, "offsets"), B, n_total,
        int(n_max), float(threshold), int(max_keep), int(flags), _lib.dptr(keep, _I32),
        _lib.dptr(n_keep, _I32), ctypes.c_void_p(ws.data_ptr()), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_nms_batched")
    return keep[:n_total], n_keep


def _nms_single(boxes, scores, labels, threshold, max_keep=-1, flags=0):
    boxes = _lib.as_f32(boxes, "boxes")
    scores = _lib.as_f32(scores, "scores").reshape(-1)
    n = int(boxes.shape[0])
    if n == 0:
        return torch.empty((0,), dtype=torch.int64, device=boxes.device)
    if labels is not None:
        labels = _lib.as_f32(labels, "labels").reshape(-1)
    offsets = torch.tensor([0, n], dtype=_I32, device=boxes.device)
    keep, n_keep = nms_batched(boxes, scores, offsets, threshold, labels, max_keep, flags, n_max=n)
    m = int(n_keep.item())  # the keep list has a data-dependent length: one sync, as in jt.nms
    return keep[:m].long()


def nms(dets, scores, threshold):
    """Drop-in for detectron.layers.nms (nms.py:12-14): keep indices, descending score."""
    return _nms_single(dets, scores, None, threshold)


def ml_nms(boxes, scores, labels, threshold):
    """Drop-in for detectron.layers.ml_nms (nms.py:38-40)."""
    return _nms_single(boxes, scores, labels, threshold)


def cython_nms(dets, thresh):
    """Drop-in for Yolact.jittor/utils/cython_nms.pyx:24 nms(dets[n,5], thresh): ascending
    indices of the survivors, '>=' comparator, +1 areas."""
    dets = _lib.as_f32(dets, "dets")
    return _nms_single(dets[:, :4].contiguous(), dets[:, 4].contiguous(), None, thresh,
                       flags=_lib.RP_NMS_CYTHON)


# ------------------------------------------------------------------------------- ROIPool
class _ROIPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, rois, output_size, spatial_scale):
        input = _lib.as_f32(input, "input")
        rois = _lib.as_f32(rois, "rois")
        PH, PW = _pair(output_size)
        N, C, H, W = (int(v) for v in input.shape)
        K = int(rois.shape[0])
        out = torch.zeros((K, C, PH, PW), dtype=_F32, device=input.device)
        arg = torch.zeros((K, C, PH, PW), dtype=_I32, device=input.device)
        rc = _lib.lib().rp_roi_pool_forward(
            _lib.dptr(input, _F32, "input"), N, C, H, W, _lib.dptr(rois, _F32, "rois"), K, PH, PW,
            float(spatial_scale), _lib.dptr(out), _lib.dptr(arg), _lib.stream_ptr(input.device))
        _lib.check(rc, "rp_roi_pool_forward")
        ctx.save_for_backward(rois, arg)
        ctx.shape = (N, C, H, W)
        ctx.mark_non_differentiable(arg)
        return out, arg

    @staticmethod
    def backward(ctx, dout, _darg):
        rois, arg = ctx.saved_tensors
        N, C, H, W = ctx.shape
        K, _, PH, PW = (int(v) for v in arg.shape)
        dout = _lib.as_f32(dout, "dout")
        grad = torch.empty((N, C, H, W), dtype=_F32, device=dout.device)
        rc = _lib.lib().rp_roi_pool_backward(
            _lib.dptr(dout, _F32), _lib.dptr(rois, _F32), _lib.dptr(arg, _I32), K, N, C, H, W, PH, PW,
            _lib.dptr(grad), _lib.stream_ptr(dout.device))
        _lib.check(rc, "rp_roi_pool_backward")
        return grad, None, None, None


def roi_pool(input, rois, output_size, spatial_scale, return_argmax=False):
    """Drop-in for detectron.layers.roi_pool (roi_pool.py:154-162)."""
    out, arg = _ROIPoolFn.apply(input, rois, output_size, spatial_scale)
    return (out, arg) if return_argmax else out


class ROIPool(torch.nn.Module):
    """roi_pool.py:166-183."""

    def __init__(self, output_size, spatial_scale):
        super().__init__()
        self.output_size = output_size
        self.spatial_scale = spatial_scale

    def forward(self, input, rois):
        return roi_pool(input, rois, self.output_size, self.spatial_scale)

    execute = forward

    def __repr__(self):
        return "%s(output_size=%s, spatial_scale=%s)" % (self.__class__.__name__, self.output_size,
                                                         self.spatial_scale)


# ------------------------------------------------------------------------------- EmbedMask
def mask_prob_cuda(embed_pixel, embed_center, sigma_center, boxes, box_areas, area_sum, mask_width):
    """Drop-in for detectron.layers.mask_prob.mask_prob_cuda (mask_prob.py:84-103): probs
    [num_center, num_pixel] (the reference declares the shape transposed but indexes it this way,
    :56).  area_sum is accepted for signature parity; the launch does not need it."""
    assert embed_pixel.ndim == 2, "embed_pixel should be MxDim"
    assert embed_center.ndim == 2, "embed_center should be NxDim"
    assert sigma_center.ndim == 1, "sigma_center should be N"
    assert embed_pixel.shape[1] == embed_center.shape[1], "Dim should the same"
    assert embed_center.shape[0] == sigma_center.shape[0], "center number should be the same"
    assert embed_center.shape[0] == boxes.shape[0], "center number and box number should be the same"
    M, N = int(embed_pixel.shape[0]), int(embed_center.shape[0])
    embed_pixel = _lib.as_f32(embed_pixel, "embed_pixel")
    if M * N == 0:
        return torch.empty((0,), dtype=_F32, device=embed_pixel.device)
    out = torch.empty((N, M), dtype=_F32, device=embed_pixel.device)
    rc = _lib.lib().rp_mask_prob(
        _lib.dptr(embed_pixel, _F32), _lib.dptr(_lib.as_f32(embed_center, "embed_center"), _F32),
        _lib.dptr(_lib.as_f32(sigma_center, "sigma_center"), _F32),
        _lib.dptr(boxes.to(_I32).contiguous(), _I32), _lib.dptr(box_areas.to(_I32).contiguous(), _I32),
        M, N, int(embed_pixel.shape[1]), int(mask_width), _lib.dptr(out), _lib.stream_ptr(out.device))
    _lib.check(rc, "rp_mask_prob")
    return out
