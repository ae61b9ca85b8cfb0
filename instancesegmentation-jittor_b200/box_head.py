"""Box-head post-processor: softmax, class-wise decode, clip, score threshold, per-class NMS and
the detections_per_img cut for a whole batch in one chain of launches.

Mirrors detectron.jittor/detectron/modeling/roi_heads/box_head/inference.py (PostProcessor:
__init__ :18-43, execute :45-90, prepare_boxlist :92-111, filter_results :112-195,
make_roi_box_post_processor :244-266) with the same names, arguments and results.  The
reference loops over images and, per image, over the 80 classes (nonzero, gather, boxlist_nms,
cat) with a device synchronisation per image; here the batch is ONE rp_box_post call (its NMS
step covers all N*(C-1) problems at once) and the host reads one vector of counts back.
"""
import ctypes

import torch

from . import _lib, layers
from .rpn import BoxCoder, box_decode
from .structures import BoxList

_F32, _I32 = torch.float32, torch.int32


def box_post(class_logits, box_regression, proposals, row_offsets, im_wh, r_max, score_thresh,
             nms_thresh, detections_per_img, weights=(10.0, 10.0, 5.0, 5.0),
             bbox_xform_clip=BoxCoder((1, 1, 1, 1)).bbox_xform_clip, cls_agnostic=False,
             cand_cap=None, out_cap=None):
    """class_logits [R,C], box_regression [R,4C] (or [R,4k] with cls_agnostic), proposals [R,4],
    row_offsets [N+1] int32 (device), im_wh [N,2].  Returns a dict of device tensors:
    boxes [N,out_cap,4], scores, labels, rows [N,out_cap], count [N], n_cand [1], prob [R,C].
    No host synchronisation."""
    class_logits = _lib.as_f32(class_logits, "class_logits")
    box_regression = _lib.as_f32(box_regression, "box_regression")
    proposals = _lib.as_f32(proposals, "proposals")
    im_wh = _lib.as_f32(im_wh, "im_wh")
    if class_logits.ndim != 2 or box_regression.ndim != 2 or proposals.ndim != 2:
        raise ValueError("class_logits [R,C], box_regression [R,4C], proposals [R,4]")
    R, C = (int(v) for v in class_logits.shape)
    N = int(row_offsets.shape[0]) - 1
    reg_cols = int(box_regression.shape[1])
    if int(box_regression.shape[0]) != R or tuple(proposals.shape) != (R, 4):
        raise ValueError("box_regression / proposals rows do not match class_logits")
    if not cls_agnostic and reg_cols != 4 * C:
        raise ValueError("box_regression should be [%d,%d], got %s" % (R, 4 * C, tuple(box_regression.shape)))
    dev = class_logits.device
    dets = int(detections_per_img)
    if cand_cap is None:
        cand_cap = R * (C - 1)
    if out_cap is None:
        out_cap = max(2 * dets, dets + 64) if dets > 0 else int(r_max) * (C - 1)
    out = dict(
        boxes=torch.empty((N, out_cap, 4), dtype=_F32, device=dev),
        scores=torch.empty((N, out_cap), dtype=_F32, device=dev),
        labels=torch.empty((N, out_cap), dtype=_I32, device=dev),
        rows=torch.empty((N, out_cap), dtype=_I32, device=dev),
        count=torch.empty((N,), dtype=_I32, device=dev),
        n_cand=torch.empty((1,), dtype=_I32, device=dev),
        prob=torch.empty((R, C), dtype=_F32, device=dev), out_cap=out_cap, cand_cap=cand_cap)
    if R == 0:
        out["count"].zero_()
        out["n_cand"].zero_()
        return out
    lib = _lib.lib()
    nbytes = lib.rp_box_post_workspace_bytes(R, C, N, int(r_max), int(cand_cap))
    ws = _lib.workspace(nbytes, dev, "box_post")
    wx, wy, ww, wh = (float(w) for w in weights)
    rc = lib.rp_box_post(
        _lib.dptr(class_logits, _F32), _lib.dptr(box_regression, _F32), _lib.dptr(proposals, _F32),
        _lib.dptr(row_offsets, _I32, "row_offsets"), _lib.dptr(im_wh, _F32), R, C, N, int(r_max),
        reg_cols, int(bool(cls_agnostic)), float(score_thresh), float(nms_thresh), dets, wx, wy, ww,
        wh, float(bbox_xform_clip), int(cand_cap), _lib.dptr(out["prob"], _F32),
        _lib.dptr(out["boxes"], _F32), _lib.dptr(out["scores"], _F32), _lib.dptr(out["labels"], _I32),
        _lib.dptr(out["rows"], _I32), _lib.dptr(out["count"], _I32), int(out_cap),
        _lib.dptr(out["n_cand"], _I32), ctypes.c_void_p(ws.data_ptr()), ws.numel(),
        _lib.stream_ptr(dev))
    _lib.check(rc, "rp_box_post")
    return out


class PostProcessor(torch.nn.Module):
    """box_head/inference.py:12-195."""

    def __init__(self, score_thresh=0.05, nms=0.5, detections_per_img=100, box_coder=None,
                 cls_agnostic_bbox_reg=False, bbox_aug_enabled=False):
        super().__init__()
        self.score_thresh = score_thresh
        self.nms = nms
        self.detections_per_img = detections_per_img
        if box_coder is None:
            box_coder = BoxCoder(weights=(10.0, 10.0, 5.0, 5.0))
        self.box_coder = box_coder
        self.cls_agnostic_bbox_reg = cls_agnostic_bbox_reg
        self.bbox_aug_enabled = bbox_aug_enabled
        self.cand_slots_per_row = 8  # first-try candidate slots per proposal (see forward)

    def forward(self, x, boxes):
        """x = (class_logits [R,C], box_regression [R,4C]); boxes: list[BoxList], one per image.
        Returns list[BoxList] with fields scores, labels (and the proposal row in "rows")."""
        class_logits, box_regression = x
        image_shapes = [b.size for b in boxes]
        boxes_per_image = [len(b) for b in boxes]
        if self.bbox_aug_enabled:
            return self._decode_only(class_logits, box_regression, boxes)
        dev = class_logits.device
        concat = torch.cat([b.bbox for b in boxes], dim=0)
        offs = [0]
        for n in boxes_per_image:
            offs.append(offs[-1] + n)
        row_offsets = torch.tensor(offs, dtype=_I32).to(dev, non_blocking=True)
        im_wh = torch.tensor([[float(s[0]), float(s[1])] for s in image_shapes], dtype=_F32).to(
            dev, non_blocking=True)
        R = offs[-1]
        box_regression = box_regression.reshape(R, -1)
        kw = dict(weights=self.box_coder.weights, bbox_xform_clip=self.box_coder.bbox_xform_clip,
                  cls_agnostic=self.cls_agnostic_bbox_reg)
        C = int(class_logits.shape[1])
        # candidate slots: a few per proposal is plenty for a trained head (2-3 % of the R*(C-1)
        # pairs pass 0.05); the call reports the real number and is repeated when it was short
        cand_cap = min(R * (C - 1), max(self.cand_slots_per_row * R, 4096))
        r_max = max(boxes_per_image + [1])
        r = box_post(class_logits, box_regression, concat, row_offsets, im_wh, r_max,
                     self.score_thresh, self.nms, self.detections_per_img, cand_cap=cand_cap, **kw)
        host = torch.cat([r["count"], r["n_cand"]]).cpu().tolist()  # the one host read
        counts, n_cand = host[:-1], host[-1]
        if n_cand > cand_cap or max(counts + [0]) > r["out_cap"]:
            # short of candidate slots, or more score ties than slack output slots: repeat once
            # with exact room
            r = box_post(class_logits, box_regression, concat, row_offsets, im_wh, r_max,
                         self.score_thresh, self.nms, self.detections_per_img,
                         cand_cap=max(n_cand, 1), out_cap=max(counts + [r["out_cap"]]), **kw)
            host = torch.cat([r["count"], r["n_cand"]]).cpu().tolist()
            counts = host[:-1]
            if max(counts + [0]) > r["out_cap"]:  # the first run was truncated: now the counts are exact
                r = box_post(class_logits, box_regression, concat, row_offsets, im_wh, r_max,
                             self.score_thresh, self.nms, self.detections_per_img,
                             cand_cap=max(n_cand, 1), out_cap=max(counts), **kw)
        results = []
        for b, c in enumerate(counts):
            bl = BoxList(r["boxes"][b, :c], image_shapes[b], mode="xyxy")
            bl.add_field("scores", r["scores"][b, :c])
            bl.add_field("labels", r["labels"][b, :c])
            bl.add_field("rows", r["rows"][b, :c])
            results.append(bl)
        return results

    execute = forward

    def _decode_only(self, class_logits, box_regression, boxes):
        """bbox_aug_enabled: decode + clip every class, no filtering (inference.py:82-86)."""
        prob = torch.softmax(_lib.as_f32(class_logits, "class_logits"), -1)
        concat = torch.cat([b.bbox for b in boxes], dim=0)
        reg = box_regression.reshape(concat.shape[0], -1)
        if self.cls_agnostic_bbox_reg:
            reg = reg[:, -4:].contiguous()
        props = self.box_coder.decode(reg, concat)
        if self.cls_agnostic_bbox_reg:
            props = props.repeat(1, prob.shape[1])
        sizes = [len(b) for b in boxes]
        out = []
        for p, s, b in zip(props.split(sizes, 0), prob.split(sizes, 0), boxes):
            out.append(self.prepare_boxlist(p, s, b.size).clip_to_image(remove_empty=False))
        return out

    def prepare_boxlist(self, boxes, scores, image_shape):
        boxlist = BoxList(boxes.reshape(-1, 4), image_shape, mode="xyxy")
        boxlist.add_field("scores", scores.reshape(-1))
        return boxlist

    def filter_results(self, boxlist, num_classes):
        """inference.py:112-195 on an already decoded BoxList ([n*C,4] boxes, [n*C] scores): the
        80 per-class NMS calls become ONE labelled NMS; its keep list (descending score) is put
        back into class-major order with a stable sort."""
        boxes = boxlist.bbox.reshape(-1, num_classes, 4)
        scores = boxlist.get_field("scores").reshape(-1, num_classes)
        inds = (scores[:, 1:] > self.score_thresh).nonzero()
        labels = (inds[:, 1] + 1)
        sc = scores[inds[:, 0], labels].contiguous()
        bx = boxes[inds[:, 0], labels].contiguous()
        if bx.shape[0] > 0:
            keep = layers.ml_nms(bx, sc, labels.float(), self.nms)
            keep = keep[torch.sort(labels[keep], stable=True)[1]]
            bx, sc, labels = bx[keep], sc[keep], labels[keep]
        result = BoxList(bx, boxlist.size, mode="xyxy")
        result.add_field("scores", sc)
        result.add_field("labels", labels.int())
        n = len(result)
        if n > self.detections_per_img > 0:
            thresh = torch.kthvalue(sc, n - self.detections_per_img + 1)[0]
            result = result[(sc >= thresh).nonzero().squeeze(1)]
        return result


    def filter_results_v2(self, boxlist, num_classes):
        """inference.py:198-243: ONE multi-label NMS over every (row, class) above the threshold, then
        the detections_per_img highest scores (topk, sorted).  Kept as the reference has it, including
        its indexing: the box of candidate (row, label) is boxes[row, label - 1] (:218, the column index
        of scores[:, 1:] is used unshifted).  boxlist_ml_nms already returns descending scores (stable),
        so topk(k) is its first k rows."""
        boxes = boxlist.bbox.reshape(-1, num_classes, 4)
        scores = boxlist.get_field("scores").reshape(-1, num_classes)[:, 1:]
        inds = (scores > self.score_thresh).nonzero()
        labels = inds[:, 1] + 1
        sc = scores[inds[:, 0], inds[:, 1]].contiguous()
        bx = boxes[inds[:, 0], inds[:, 1], :].reshape(-1, 4).contiguous()
        if bx.shape[0] > 0 and self.nms > 0:
            keep = layers.ml_nms(bx, sc, labels.float(), self.nms)
            bx, sc, labels = bx[keep], sc[keep], labels[keep]
        if bx.shape[0] > self.detections_per_img > 0:
            if not self.nms > 0:  # no NMS ran: order by score first
                order = torch.sort(sc, descending=True, stable=True)[1]
                bx, sc, labels = bx[order], sc[order], labels[order]
            k = self.detections_per_img
            bx, sc, labels = bx[:k], sc[:k], labels[:k]
        result = BoxList(bx, boxlist.size, mode="xyxy")
        result.add_field("scores", sc)
        result.add_field("labels", labels)
        return result


def make_roi_box_post_processor(cfg):
    """inference.py:244-266; cfg is any object with the MODEL.* / TEST.* attributes."""
    rh = cfg.MODEL.ROI_HEADS
    return PostProcessor(rh.SCORE_THRESH, rh.NMS, rh.DETECTIONS_PER_IMG,
                         BoxCoder(weights=rh.BBOX_REG_WEIGHTS), cfg.MODEL.CLS_AGNOSTIC_BBOX_REG,
                         cfg.TEST.BBOX_AUG.ENABLED)
