"""EmbedMask mask stage: the host logic around mask_prob.

Reference: detectron.jittor/detectron/modeling/rpn/embed_mask/inference.py
  :153-169  compute_mask_prob  (int cast of the mask boxes, +-2 growth, clamps, areas, mask_prob_cuda)
  :171-216  forward_for_mask   (boxes / stride, boxes_to_masks, prob * box mask, optional upsample +
                                threshold + crop, BoxList fields mask / labels / scores / stride / mask_th)
The reference transposes pixel_embed to [HW, dim] and makes three passes over n x HW floats per
image; rp_embed_masks does the whole product in one launch on the [dim, m_h, m_w] tensor.
"""
import torch

from . import _lib
from .structures import BoxList

_F32 = torch.float32


def embed_masks(pixel_embed, proposal_embed, proposal_margin, boxes, stride):
    """masks [n, m_h, m_w] = mask_prob * boxes_to_masks(boxes / stride) (inference.py:193-197)."""
    pixel_embed = _lib.as_f32(pixel_embed, "pixel_embed")
    dim, m_h, m_w = (int(v) for v in pixel_embed.shape)
    n = int(proposal_embed.shape[0])
    out = torch.empty((n, m_h, m_w), dtype=_F32, device=pixel_embed.device)
    if n == 0:
        return out
    rc = _lib.lib().rp_embed_masks(
        _lib.dptr(pixel_embed, _F32), _lib.dptr(_lib.as_f32(proposal_embed, "proposal_embed"), _F32),
        _lib.dptr(_lib.as_f32(proposal_margin, "proposal_margin"), _F32),
        _lib.dptr(_lib.as_f32(boxes, "boxes"), _F32), float(stride), n, dim, m_h, m_w,
        _lib.dptr(out, _F32), _lib.stream_ptr(out.device))
    _lib.check(rc, "rp_embed_masks")
    return out


def resize_threshold(masks, scale_factor, thresh, crop_h, crop_w):
    """nn.interpolate(masks, scale_factor, 'bilinear', align_corners=False) > thresh, [:crop_h, :crop_w]
    (inference.py:199-202) -> uint8 [n, crop_h, crop_w]."""
    masks = _lib.as_f32(masks, "masks")
    n, h, w = (int(v) for v in masks.shape)
    H, W = int(h * scale_factor), int(w * scale_factor)
    crop_h, crop_w = min(int(crop_h), H), min(int(crop_w), W)
    out = torch.empty((n, crop_h, crop_w), dtype=torch.uint8, device=masks.device)
    if n == 0:
        return out
    import ctypes
    rc = _lib.lib().rp_resize_threshold(_lib.dptr(masks, _F32), n, h, w, H, W, float(thresh), crop_h, crop_w,
                                        ctypes.c_void_p(out.data_ptr()), _lib.stream_ptr(out.device))
    _lib.check(rc, "rp_resize_threshold")
    return out


class EmbedMaskMaskStage(object):
    """forward_for_mask of EmbedMaskPostProcessor (inference.py:171-216) with its constructor values
    fpn_strides[0], mask_scale_factor, mask_th, post_process_masks."""

    def __init__(self, fpn_stride0, mask_scale_factor, mask_th, post_process_masks):
        self.stride = float(fpn_stride0) / float(mask_scale_factor)
        self.mask_th = float(mask_th)
        self.post_process_masks = bool(post_process_masks)

    def compute_mask_prob(self, pixel_embed, proposal_embed, proposal_margin, mask_boxes):
        """inference.py:153-169 (boxes already divided by the stride); the box-mask product is NOT applied."""
        from .layers import mask_prob_cuda
        dim, m_h, m_w = (int(v) for v in pixel_embed.shape)
        pe = pixel_embed.reshape(dim, m_h * m_w).t().contiguous()
        b = mask_boxes.to(torch.int32)
        b[:, 0] = torch.clamp(b[:, 0] - 2, min=0)
        b[:, 1] = torch.clamp(b[:, 1] - 2, min=0)
        b[:, 2] = torch.clamp(b[:, 2] + 2, max=m_w)
        b[:, 3] = torch.clamp(b[:, 3] + 2, max=m_h)
        areas = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
        prob = mask_prob_cuda(pe, proposal_embed, proposal_margin, b, areas, None, m_w)
        return prob.view(-1, m_h, m_w)

    def forward_for_mask(self, boxlists, pixel_embed):
        N, dim, m_h, m_w = (int(v) for v in pixel_embed.shape)
        out = []
        dev = pixel_embed.device
        for im in range(N):
            bl = boxlists[im]
            input_w, input_h = bl.size
            pe = bl.get_field("proposal_embed")
            new = BoxList(bl.bbox, bl.size, mode="xyxy")
            new.add_field("labels", bl.get_field("labels"))
            new.add_field("scores", bl.get_field("scores"))
            post = self.post_process_masks
            if int(pe.shape[0]) == 0:
                new.add_field("mask", torch.empty((0,), device=dev))
            else:
                masks = embed_masks(pixel_embed[im], pe, bl.get_field("proposal_margin"), bl.bbox, self.stride)
                if post:
                    masks = resize_threshold(masks, self.stride, self.mask_th, input_h, input_w).bool()
                new.add_field("mask", masks)
            new.add_field("stride", torch.tensor([1.0 if post else self.stride], device=dev))
            new.add_field("mask_th", torch.tensor([0.0 if post else self.mask_th], device=dev))
            out.append(new)
        return out
