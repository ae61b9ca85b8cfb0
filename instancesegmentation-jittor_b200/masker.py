"""Mask R-CNN mask post-processing with the reference's interface.

Mirrors detectron.jittor/detectron/modeling/roi_heads/mask_head/inference.py:
MaskPostProcessor :12-60, expand_boxes :90-104, expand_masks :107-115,
paste_mask_in_image :118-162, Masker :165-205.

The reference pastes one instance at a time in a python loop with a .numpy() sync per
instance; here all instances of an image (or of a batch of same-sized images) are pasted by
one launch that writes every output byte exactly once.
"""
import torch

from . import _lib
from .structures import BoxList

_I32 = torch.int32
_F32 = torch.float32


def mask_select_sigmoid(logits, labels):
    """inference.py:38-45: sigmoid(logits)[i, labels[i]] -> [K,1,M,M]."""
    logits = _lib.as_f32(logits, "logits")
    K, NC, M, M2 = [int(v) for v in logits.shape]
    assert M == M2
    labels = labels.to(_I32).contiguous()
    out = torch.empty((K, 1, M, M), dtype=_F32, device=logits.device)
    if K:
        rc = _lib.lib().rp_mask_select_sigmoid(_lib.dptr(logits, _F32), _lib.dptr(labels, _I32), K,
                                               NC, M, _lib.dptr(out, _F32),
                                               _lib.stream_ptr(logits.device))
        _lib.check(rc, "rp_mask_select_sigmoid")
    return out


def paste_masks(masks, boxes, im_h, im_w, thresh=0.5, padding=1, out=None):
    """masks [n,M,M] (or [n,1,M,M]) probabilities, boxes [n,4] xyxy -> uint8 [n,im_h,im_w]."""
    masks = _lib.as_f32(masks, "masks")
    boxes = _lib.as_f32(boxes, "boxes")
    if masks.ndim == 4:
        masks = masks[:, 0].contiguous()
    n, M = int(masks.shape[0]), int(masks.shape[-1])
    if out is None:
        out = torch.empty((n, int(im_h), int(im_w)), dtype=torch.uint8, device=masks.device)
    if n:
        rc = _lib.lib().rp_paste_masks(_lib.dptr(masks, _F32), _lib.dptr(boxes, _F32), n, M,
                                       int(im_h), int(im_w), float(thresh), int(padding),
                                       _lib.dptr(out, torch.uint8), _lib.stream_ptr(masks.device))
        _lib.check(rc, "rp_paste_masks")
    return out


def paste_mask_in_image(mask, box, im_h, im_w, thresh=0.5, padding=1):
    """Drop-in for inference.py:118-162: one [M,M] mask, one box -> bool [im_h,im_w]."""
    return paste_masks(mask.reshape(1, mask.shape[-2], mask.shape[-1]), box.reshape(1, 4), im_h,
                       im_w, thresh, padding)[0].bool()


class Masker(object):
    """Drop-in for inference.py:165-205."""

    def __init__(self, threshold=0.5, padding=1):
        self.threshold = threshold
        self.padding = padding

    def forward_single_image(self, masks, boxes):
        boxes = boxes.convert("xyxy")
        im_w, im_h = boxes.size
        if boxes.bbox.shape[0] == 0:
            return masks.new_empty((0, 1, masks.shape[-2], masks.shape[-1]))
        res = paste_masks(masks, boxes.bbox, im_h, im_w, self.threshold, self.padding)
        return res.bool().unsqueeze(1)

    def __call__(self, masks, boxes):
        if isinstance(boxes, BoxList):
            boxes = [boxes]
        assert len(boxes) == len(masks), "Masks and boxes should have the same length."
        results = []
        for mask, box in zip(masks, boxes):
            assert mask.shape[0] == len(box), "Number of objects should be the same."
            results.append(self.forward_single_image(mask, box))
        return results


class MaskPostProcessor(torch.nn.Module):
    """Drop-in for inference.py:12-60."""

    def __init__(self, masker=None):
        super().__init__()
        self.masker = masker

    def forward(self, x, boxes):
        labels = torch.cat([b.get_field("labels") for b in boxes], dim=0)
        mask_prob = mask_select_sigmoid(x, labels)
        boxes_per_image = [len(b) for b in boxes]
        mask_prob = mask_prob.split(boxes_per_image, dim=0)
        if self.masker:
            mask_prob = self.masker(mask_prob, boxes)
        results = []
        for prob, box in zip(mask_prob, boxes):
            bbox = BoxList(box.bbox, box.size, mode="xyxy")
            for field in box.fields():
                bbox.add_field(field, box.get_field(field))
            bbox.add_field("mask", prob)
            results.append(bbox)
        return results

    execute = forward
