"""Yolact post-processing with the reference's interface: decode, Detect (FastNMS /
cross-class FastNMS / traditional NMS), crop, sanitize_coordinates, postprocess.

Mirrors Yolact.jittor/layers/box_utils.py (decode :272-316, sanitize_coordinates :333-351,
crop :355-378, jaccard :52-78), layers/functions/detection.py (Detect :11-237) and
layers/output_utils.py (postprocess :17-124, lincomb branch).

The reference takes its constants from the global `cfg` (data/config.py: max_num_detections
:423, nms_top_k :449, max_size :666); here they are constructor / keyword arguments with the
yolact_resnet50 values as defaults.  All arithmetic runs in libroipath_b200.so.
"""
import ctypes

import torch

from . import _lib, layers

_I32 = torch.int32
_F32 = torch.float32


def decode(loc, priors):
    """box_utils.py:272-316: SSD regressors -> relative point-form boxes [n,4]."""
    loc = _lib.as_f32(loc, "loc")
    priors = _lib.as_f32(priors, "priors")
    out = torch.empty_like(loc)
    n = int(loc.shape[0])
    if n:
        rc = _lib.lib().rp_yolact_decode(_lib.dptr(loc, _F32), _lib.dptr(priors, _F32), n,
                                         _lib.dptr(out, _F32), _lib.stream_ptr(loc.device))
        _lib.check(rc, "rp_yolact_decode")
    return out


def sanitize_coordinates(_x1, _x2, img_size, padding=0, cast=True):
    """box_utils.py:333-351 (elementwise on [n] tensors; kept for API parity -- the fused
    kernels apply the same arithmetic internally)."""
    _x1 = _x1 * img_size
    _x2 = _x2 * img_size
    if cast:
        _x1 = _x1.int()
        _x2 = _x2.int()
    x1 = torch.minimum(_x1, _x2)
    x2 = torch.maximum(_x1, _x2)
    x1 = torch.clamp(x1 - padding, min=0)
    x2 = torch.clamp(x2 + padding, max=img_size)
    return x1, x2


def fast_nms_batched(boxes, coeff, scores, offsets, iou_threshold=0.5, top_k=200, max_det=100,
                     second_threshold=False, conf_thresh=0.05, cross_class=False, n_max=None):
    """FastNMS for a batch of images without host synchronisation.

    boxes [n,4], coeff [n,D], scores [C,n] (class-major), offsets [B+1] int32 (device).
    Returns dict of padded outputs: box [B,max_det,4], mask [B,max_det,D], class [B,max_det]
    int32, score [B,max_det], index [B,max_det] int32, count [B] int32.
    """
    dev = boxes.device
    n_total = int(boxes.shape[0])
    C = int(scores.shape[0])
    D = int(coeff.shape[1])
    B = int(offsets.shape[0]) - 1
    if cross_class:
        max_det = top_k
    ob = torch.zeros((B, max_det, 4), dtype=_F32, device=dev)
    oc = torch.zeros((B, max_det, D), dtype=_F32, device=dev)
    ocl = torch.zeros((B, max_det), dtype=_I32, device=dev)
    osc = torch.zeros((B, max_det), dtype=_F32, device=dev)
    oi = torch.zeros((B, max_det), dtype=_I32, device=dev)
    cnt = torch.zeros((B,), dtype=_I32, device=dev)
    L = _lib.lib()
    nbytes = L.rp_fast_nms_workspace_bytes(n_total, int(n_max or n_total), B, C, int(top_k))
    ws = _lib.workspace(nbytes, dev, "fast_nms")
    rc = L.rp_fast_nms_batched(
        _lib.dptr(boxes, _F32, "boxes"), _lib.dptr(coeff, _F32, "coeff"),
        _lib.dptr(scores, _F32, "scores"), _lib.dptr(offsets, _I32, "offsets"), B, n_total,
        int(n_max or n_total), C, D, float(iou_threshold), int(top_k), int(max_det),
        int(bool(second_threshold)), float(conf_thresh), int(bool(cross_class)),
        _lib.dptr(ob), _lib.dptr(oc), _lib.dptr(ocl), _lib.dptr(osc), _lib.dptr(oi),
        _lib.dptr(cnt), ctypes.c_void_p(ws.data_ptr()), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_fast_nms_batched")
    return {"box": ob, "mask": oc, "class": ocl, "score": osc, "index": oi, "count": cnt}


def detect_prefilter(loc, conf, mask, priors, conf_thresh, cand_cap=None):
    """Detect pre-stage for a whole batch (detection.py:64-98) in three launches: decode,
    class-max over the foreground classes, `> conf_thresh`, ordered gathers.  No host sync.

    loc [B,P,4], conf [B,P,NC], mask [B,P,D], priors [P,4].  Returns a dict of device tensors:
    box [cap,4], mask [cap,D], score [NC-1,cap], prior [cap] int32, offsets [B+1] int32,
    n_cand [1] int32 (the real number of kept priors: > cap means the run was truncated).
    """
    loc = _lib.as_f32(loc, "loc")
    conf = _lib.as_f32(conf, "conf")
    mask = _lib.as_f32(mask, "mask")
    priors = _lib.as_f32(priors, "priors")
    B, P = int(loc.shape[0]), int(loc.shape[1])
    NC, D = int(conf.shape[-1]), int(mask.shape[-1])
    if conf.numel() != B * P * NC or mask.numel() != B * P * D or priors.numel() != P * 4:
        raise ValueError("loc [B,P,4], conf [B,P,NC], mask [B,P,D], priors [P,4]")
    dev = loc.device
    cap = int(cand_cap) if cand_cap else B * min(P, 4096)
    out = dict(box=torch.empty((cap, 4), dtype=_F32, device=dev),
               mask=torch.empty((cap, D), dtype=_F32, device=dev),
               score=torch.empty((NC - 1, cap), dtype=_F32, device=dev),
               prior=torch.empty((cap,), dtype=_I32, device=dev),
               offsets=torch.empty((B + 1,), dtype=_I32, device=dev),
               n_cand=torch.empty((1,), dtype=_I32, device=dev), cap=cap)
    L = _lib.lib()
    ws = _lib.workspace(L.rp_yolact_prefilter_workspace_bytes(B, P, cap), dev, "yolact_prefilter")
    rc = L.rp_yolact_prefilter(
        _lib.dptr(loc, _F32), _lib.dptr(conf, _F32), _lib.dptr(mask, _F32), _lib.dptr(priors, _F32),
        B, P, NC, D, float(conf_thresh), cap, _lib.dptr(out["box"]), _lib.dptr(out["mask"]),
        _lib.dptr(out["score"]), _lib.dptr(out["prior"]), _lib.dptr(out["offsets"]),
        _lib.dptr(out["n_cand"]), ctypes.c_void_p(ws.data_ptr()), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_yolact_prefilter")
    return out


class Detect(object):
    """Drop-in for layers.functions.detection.Detect (detection.py:11-237)."""

    def __init__(self, num_classes, bkg_label, top_k, conf_thresh, nms_thresh,
                 max_num_detections=100, max_size=550):
        self.num_classes = num_classes
        self.background_label = bkg_label
        self.top_k = top_k
        self.nms_thresh = nms_thresh
        if nms_thresh <= 0:
            raise ValueError("nms_threshold must be non negative.")
        self.conf_thresh = conf_thresh
        self.max_num_detections = max_num_detections  # cfg.max_num_detections
        self.max_size = max_size                      # cfg.max_size
        self.use_cross_class_nms = False
        self.use_fast_nms = False

    def __call__(self, predictions, net=None):
        """detection.py:32-78.  predictions: dict with 'loc' [B,P,4], 'conf' [B,P,num_classes],
        'mask' [B,P,D], 'priors' [P,4], optional 'proto' [B,h,w,D]."""
        loc_data = predictions["loc"]
        conf_data = predictions["conf"]
        mask_data = predictions["mask"]
        prior_data = predictions["priors"]
        proto_data = predictions.get("proto")
        out = []
        batch_size = loc_data.shape[0]
        num_priors = prior_data.shape[0]
        if self.use_fast_nms:
            return self._call_batched(loc_data, conf_data, mask_data, prior_data, proto_data, net)
        conf_preds = conf_data.view(batch_size, num_priors, self.num_classes).transpose(2, 1)
        for batch_idx in range(batch_size):
            decoded_boxes = decode(loc_data[batch_idx], prior_data)
            result = self.detect(batch_idx, conf_preds, decoded_boxes, mask_data)
            if result is not None and proto_data is not None:
                result["proto"] = proto_data[batch_idx]
            out.append({"detection": result, "net": net})
        return out

    def _call_batched(self, loc_data, conf_data, mask_data, prior_data, proto_data, net):
        """The FastNMS branches for the whole batch: rp_yolact_prefilter -> rp_fast_nms_batched
        chained through device-side offsets; ONE host read (counts) for all images."""
        B, P = int(loc_data.shape[0]), int(prior_data.shape[0])
        conf = conf_data.reshape(B, P, self.num_classes)
        cap = None
        while True:
            pre = detect_prefilter(loc_data, conf, mask_data, prior_data, self.conf_thresh, cap)
            r = fast_nms_batched(pre["box"], pre["mask"], pre["score"], pre["offsets"],
                                 self.nms_thresh, self.top_k, self.max_num_detections, False,
                                 self.conf_thresh, self.use_cross_class_nms, n_max=min(P, pre["cap"]))
            host = torch.cat([r["count"], pre["n_cand"], pre["offsets"]]).cpu().tolist()
            counts, n_cand, offs = host[:B], host[B], host[B + 1:]
            if n_cand <= pre["cap"]:
                break
            cap = n_cand  # more candidates than slots: repeat once with exact room
        out = []
        for b in range(B):
            m = counts[b]
            if offs[b + 1] - offs[b] == 0:  # detection.py:92-93: no candidate -> None
                result = None
            else:
                result = {"box": r["box"][b, :m], "mask": r["mask"][b, :m],
                          "class": r["class"][b, :m].long(), "score": r["score"][b, :m]}
                if proto_data is not None:
                    result["proto"] = proto_data[b]
            out.append({"detection": result, "net": net})
        return out

    def detect(self, batch_idx, conf_preds, decoded_boxes, mask_data, inst_data=None):
        """detection.py:81-110 (class-max over the foreground rows, conf filter, NMS)."""
        cur_scores = conf_preds[batch_idx, 1:]
        conf_scores = cur_scores.max(dim=0).values
        keep = conf_scores > self.conf_thresh
        scores = cur_scores[:, keep].contiguous()
        boxes = decoded_boxes[keep].contiguous()
        masks = mask_data[batch_idx][keep].contiguous()
        if scores.shape[1] == 0:
            return None
        if self.use_fast_nms:
            if self.use_cross_class_nms:
                boxes, masks, classes, scores = self.cc_fast_nms(boxes, masks, scores,
                                                                 self.nms_thresh, self.top_k)
            else:
                boxes, masks, classes, scores = self.fast_nms(boxes, masks, scores,
                                                              self.nms_thresh, self.top_k)
        else:
            boxes, masks, classes, scores = self.traditional_nms(boxes, masks, scores,
                                                                 self.nms_thresh, self.conf_thresh)
        return {"box": boxes, "mask": masks, "class": classes, "score": scores}

    def _single(self, boxes, masks, scores, iou_threshold, top_k, second_threshold, cross_class):
        boxes = _lib.as_f32(boxes, "boxes")
        masks = _lib.as_f32(masks, "masks")
        scores = _lib.as_f32(scores, "scores")
        n = int(boxes.shape[0])
        offsets = torch.tensor([0, n], dtype=_I32, device=boxes.device)
        r = fast_nms_batched(boxes, masks, scores, offsets, iou_threshold, top_k,
                             self.max_num_detections, second_threshold, self.conf_thresh,
                             cross_class, n_max=n)
        m = int(r["count"].item())
        return r["box"][0, :m], r["mask"][0, :m], r["class"][0, :m].long(), r["score"][0, :m]

    def fast_nms(self, boxes, masks, scores, iou_threshold=0.5, top_k=200, second_threshold=False):
        """detection.py:139-189.  boxes [n,4], masks [n,D], scores [C,n]."""
        return self._single(boxes, masks, scores, iou_threshold, top_k, second_threshold, False)

    def cc_fast_nms(self, boxes, masks, scores, iou_threshold=0.5, top_k=200):
        """detection.py:113-137."""
        return self._single(boxes, masks, scores, iou_threshold, top_k, False, True)

    def traditional_nms(self, boxes, masks, scores, iou_threshold=0.5, conf_thresh=0.05):
        """detection.py:191-237: per-class cython NMS on boxes * max_size (all classes are ONE
        batched launch here), then the global descending sort and top max_num_detections."""
        boxes = _lib.as_f32(boxes, "boxes")
        scores = _lib.as_f32(scores, "scores")
        num_classes, n = scores.shape
        dev = boxes.device
        scaled = boxes * self.max_size
        conf_mask = scores > conf_thresh                     # [C,n]
        counts = conf_mask.sum(dim=1)                        # [C]
        cls_of, idx_of = conf_mask.nonzero(as_tuple=True)    # class-major, ascending index
        if idx_of.numel() == 0:
            e = torch.empty((0,), device=dev)
            return boxes[:0], masks[:0], e.long(), e
        offsets = torch.zeros((num_classes + 1,), dtype=_I32, device=dev)
        offsets[1:] = counts.cumsum(0).to(_I32)
        pb = scaled[idx_of].contiguous()
        ps = scores[cls_of, idx_of].contiguous()
        keep, n_keep = layers.nms_batched(pb, ps, offsets, float(iou_threshold),
                                          None, -1, _lib.RP_NMS_CYTHON, n_max=int(counts.max()))
        offs = offsets.cpu().tolist()
        nk = n_keep.cpu().tolist()
        idx_lst, cls_lst, scr_lst = [], [], []
        for c in range(num_classes):
            if nk[c] == 0:
                continue
            k = keep[offs[c]: offs[c] + nk[c]].long() + offs[c]
            idx_lst.append(idx_of[k])
            cls_lst.append(cls_of[k])
            scr_lst.append(ps[k])
        idx = torch.cat(idx_lst)
        classes = torch.cat(cls_lst)
        sc = torch.cat(scr_lst)
        sc, idx2 = sc.sort(dim=0, descending=True, stable=True)
        idx2 = idx2[: self.max_num_detections]
        sc = sc[: self.max_num_detections]
        idx = idx[idx2]
        classes = classes[idx2]
        return scaled[idx] / self.max_size, masks[idx], classes, sc


def mask_assemble(proto, coeff, boxes, out_h, out_w, crop_masks=True, det_offsets=None,
                  return_low=False, out=None):
    """Fused lincomb -> sigmoid -> crop -> bilinear upsample -> >0.5 (output_utils.py:71-96)
    plus the absolute int32 boxes (:99-101).

    proto [ph,pw,D] or [B,ph,pw,D]; coeff [m,D]; boxes [m,4] relative.  det_offsets [B+1]
    int32 (device) maps detections to images when proto is batched.
    Returns (masks uint8 [m,out_h,out_w], boxes int32 [m,4][, low f32 [m,ph,pw]])."""
    proto = _lib.as_f32(proto, "proto")
    coeff = _lib.as_f32(coeff, "coeff")
    boxes = _lib.as_f32(boxes, "boxes")
    if proto.ndim == 3:
        proto = proto.unsqueeze(0)
    B, ph, pw, D = [int(v) for v in proto.shape]
    m = int(coeff.shape[0])
    dev = proto.device
    masks = out if out is not None else torch.empty((m, int(out_h), int(out_w)), dtype=torch.uint8, device=dev)
    ob = torch.empty((m, 4), dtype=_I32, device=dev)
    low = torch.empty((m, ph, pw), dtype=_F32, device=dev) if return_low else None
    if m:
        rc = _lib.lib().rp_mask_assemble(
            _lib.dptr(proto, _F32), _lib.dptr(coeff, _F32), _lib.dptr(boxes, _F32),
            _lib.dptr(det_offsets, _I32, "det_offsets"), B, m, ph, pw, D, int(out_h), int(out_w),
            int(bool(crop_masks)), _lib.dptr(masks, torch.uint8), _lib.dptr(ob, _I32),
            _lib.dptr(low, _F32), _lib.stream_ptr(dev))
        _lib.check(rc, "rp_mask_assemble")
    return (masks, ob, low) if return_low else (masks, ob)


def crop(masks, boxes, padding=1):
    """box_utils.py:355-378 on [h,w,n] masks (API parity; postprocess uses the fused kernel)."""
    h, w, n = masks.shape
    x1, x2 = sanitize_coordinates(boxes[:, 0], boxes[:, 2], w, padding, cast=False)
    y1, y2 = sanitize_coordinates(boxes[:, 1], boxes[:, 3], h, padding, cast=False)
    rows = torch.arange(w, dtype=x1.dtype, device=masks.device).view(1, -1, 1).expand(h, w, n)
    cols = torch.arange(h, dtype=x1.dtype, device=masks.device).view(-1, 1, 1).expand(h, w, n)
    crop_mask = (rows >= x1.view(1, 1, -1)) & (rows < x2.view(1, 1, -1)) & \
                (cols >= y1.view(1, 1, -1)) & (cols < y2.view(1, 1, -1))
    return masks * crop_mask.float()


def postprocess(det_output, w, h, batch_idx=0, interpolation_mode="bilinear",
                visualize_lincomb=False, crop_masks=True, score_threshold=0):
    """Drop-in for layers.output_utils.postprocess (output_utils.py:17-124), lincomb masks.
    Returns classes [m], scores [m], boxes int32 [m,4], masks bool [m,h,w]."""
    if interpolation_mode != "bilinear":
        raise ValueError("only the reference's default 'bilinear' interpolation is implemented")
    dets = det_output[batch_idx]["detection"]
    if dets is None:
        e = torch.empty((0,))
        return [e] * 4
    if score_threshold > 0:
        keep = dets["score"] > score_threshold
        for k in dets:
            if k != "proto":
                dets[k] = dets[k][keep]
        if dets["score"].shape[0] == 0:
            e = torch.empty((0,))
            return [e] * 4
    classes, boxes, scores, coeff = dets["class"], dets["box"], dets["score"], dets["mask"]
    masks, iboxes = mask_assemble(dets["proto"], coeff.contiguous(), boxes.contiguous(), h, w,
                                  crop_masks)
    return classes, scores, iboxes, masks.bool()


# ------------------------------------------------------------------------------------------------
# lincomb mask loss (training twin of the assembly; multibox_loss.py:616-660)
# ------------------------------------------------------------------------------------------------
def lincomb_mask_loss_raw(proto, coef, mask_t, boxes=None, scale=None):
    """One launch: per-instance losses [n] and the gradients of their sum (d_proto [H,W,D],
    d_coef [n,D]).  proto [H,W,D], coef [n,D], mask_t [H,W,n], boxes [n,4] relative xyxy or None
    (no crop), scale [n] or None."""
    proto = _lib.as_f32(proto, "proto")
    H, W, D = (int(v) for v in proto.shape)
    n = int(coef.shape[0])
    dev = proto.device
    loss = torch.empty((n,), dtype=_F32, device=dev)
    d_proto = torch.empty((H, W, D), dtype=_F32, device=dev)
    d_coef = torch.empty((n, D), dtype=_F32, device=dev)
    if n:
        coef = _lib.as_f32(coef, "coef")
        mask_t = _lib.as_f32(mask_t, "mask_t")
        if tuple(mask_t.shape) != (H, W, n) or tuple(coef.shape) != (n, D):
            raise ValueError("mask_t must be [H,W,n] and coef [n,D]")
        boxes = _lib.as_f32(boxes, "boxes") if boxes is not None else None
        scale = _lib.as_f32(scale, "scale") if scale is not None else None
    rc = _lib.lib().rp_lincomb_mask_loss(
        _lib.dptr(proto, _F32), _lib.dptr(coef, _F32) if n else None, _lib.dptr(mask_t, _F32) if n else None,
        _lib.dptr(boxes, _F32) if n else None, _lib.dptr(scale, _F32) if n else None, H, W, D, n,
        _lib.dptr(loss, _F32), _lib.dptr(d_proto, _F32), _lib.dptr(d_coef, _F32), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_lincomb_mask_loss")
    return loss, d_proto, d_coef


class _LincombMaskLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, proto, coef, mask_t, boxes, scale):
        loss, d_proto, d_coef = lincomb_mask_loss_raw(proto.detach(), coef.detach(), mask_t, boxes, scale)
        ctx.save_for_backward(d_proto, d_coef)
        ctx.mark_non_differentiable(loss)
        return loss.sum(), loss

    @staticmethod
    def backward(ctx, g_total, _g_per_instance):
        d_proto, d_coef = ctx.saved_tensors
        return g_total * d_proto, g_total * d_coef, None, None, None


def lincomb_mask_loss(proto, coef, mask_t, boxes=None, normalize_emulate_roi_pooling=False, weight=1.0):
    """The per-image body of MultiBoxLoss.lincomb_mask_loss (multibox_loss.py:616-660) for the
    sigmoid activation: returns (sum of the per-instance losses -- differentiable w.r.t. proto and
    coef --, per-instance losses).  boxes = pos_gt_box_t (relative xyxy) when mask_proto_crop;
    normalize_emulate_roi_pooling divides instance i by gt_w*gt_h (in mask pixels) and multiplies by
    H*W when cropping (:645-649); `weight` folds old_num_pos/num_pos (:652-653)."""
    H, W = int(proto.shape[0]), int(proto.shape[1])
    n = int(coef.shape[0])
    scale = None
    if normalize_emulate_roi_pooling or weight != 1.0:
        scale = torch.full((n,), float(weight), dtype=_F32, device=proto.device)
        if normalize_emulate_roi_pooling:
            if boxes is None:
                raise ValueError("normalize_emulate_roi_pooling needs the gt boxes")
            w_roi = float(H * W)  # mask_proto_crop is on whenever boxes are given
            gw = (boxes[:, 2] - boxes[:, 0]).to(_F32) * W
            gh = (boxes[:, 3] - boxes[:, 1]).to(_F32) * H
            scale = scale / gw / gh * w_roi
    return _LincombMaskLoss.apply(proto, coef, mask_t, boxes, scale)
