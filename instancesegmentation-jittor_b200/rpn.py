"""RPN proposal selection in front of the NMS: BoxCoder.decode and RPNPostProcessor.

Mirrors detectron.jittor/detectron/modeling/box_coder.py (BoxCoder.decode :53-94) and
detectron.jittor/detectron/modeling/rpn/inference.py (RPNPostProcessor: __init__ :21-53,
add_gt_proposals :55-74, forward_for_single_feature_map :76-125, execute :127-155,
select_over_all_levels :157-188, make_rpn_postprocessor :191-213) with the same names,
arguments and results.  Where the reference runs permute / sigmoid / topk / two gathers /
decode per level and then a per-image python loop of clip, remove_small_boxes and
boxlist_nms, this module makes ONE rp_rpn_pre_nms_levels call (all levels, all images, one
sort), ONE rp_nms_batched call over the L*N problems and, at test time, ONE
rp_rpn_select_levels call, chained through device-side offsets; the host reads one small
vector of counts back at the end.
"""
import ctypes
import math

import torch

from . import _lib, layers
from .structures import BoxList, cat_boxlist

_F32, _I32 = torch.float32, torch.int32


def box_decode(rel_codes, boxes, weights=(1.0, 1.0, 1.0, 1.0), bbox_xform_clip=math.log(1000.0 / 16)):
    """rel_codes [n,4K], boxes [n,4] -> decoded boxes [n,4K] (box_coder.py:53-94)."""
    rel_codes = _lib.as_f32(rel_codes, "rel_codes")
    boxes = _lib.as_f32(boxes, "boxes")
    if rel_codes.ndim != 2 or rel_codes.shape[1] % 4 != 0 or rel_codes.shape[1] == 0:
        raise ValueError("rel_codes should be [n,4K], got %s" % (tuple(rel_codes.shape),))
    n, K = int(rel_codes.shape[0]), int(rel_codes.shape[1]) // 4
    if boxes.ndim != 2 or boxes.shape[0] != n or boxes.shape[1] != 4:
        raise ValueError("boxes should be [%d,4], got %s" % (n, tuple(boxes.shape)))
    out = torch.empty_like(rel_codes)
    wx, wy, ww, wh = (float(w) for w in weights)
    rc = _lib.lib().rp_box_decode(_lib.dptr(rel_codes, _F32), _lib.dptr(boxes, _F32), n, K, wx, wy,
                                  ww, wh, float(bbox_xform_clip), _lib.dptr(out, _F32),
                                  _lib.stream_ptr(rel_codes.device))
    _lib.check(rc, "rp_box_decode")
    return out


class BoxCoder(object):
    """box_coder.py:7-20.  encode (:22-51) is a training-target helper outside the path; it is
    kept as plain tensor arithmetic so that the class is a drop-in."""

    def __init__(self, weights, bbox_xform_clip=math.log(1000.0 / 16)):
        self.weights = weights
        self.bbox_xform_clip = bbox_xform_clip

    def encode(self, reference_boxes, proposals):
        TO_REMOVE = 1
        ex_w = proposals[:, 2] - proposals[:, 0] + TO_REMOVE
        ex_h = proposals[:, 3] - proposals[:, 1] + TO_REMOVE
        ex_cx = proposals[:, 0] + 0.5 * ex_w
        ex_cy = proposals[:, 1] + 0.5 * ex_h
        gt_w = reference_boxes[:, 2] - reference_boxes[:, 0] + TO_REMOVE
        gt_h = reference_boxes[:, 3] - reference_boxes[:, 1] + TO_REMOVE
        gt_cx = reference_boxes[:, 0] + 0.5 * gt_w
        gt_cy = reference_boxes[:, 1] + 0.5 * gt_h
        wx, wy, ww, wh = self.weights
        return torch.stack((wx * (gt_cx - ex_cx) / ex_w, wy * (gt_cy - ex_cy) / ex_h,
                            ww * torch.log(gt_w / ex_w), wh * torch.log(gt_h / ex_h)), dim=1)

    def decode(self, rel_codes, boxes):
        return box_decode(rel_codes, boxes, self.weights, self.bbox_xform_clip)


def rpn_pre_nms(objectness, box_regression, anchors, im_wh, pre_nms_top_n, min_size,
                weights=(1.0, 1.0, 1.0, 1.0), bbox_xform_clip=math.log(1000.0 / 16)):
    """One level, N images: sigmoid + top-k + gather + decode + clip + small-box removal.

    objectness [N,A,H,W], box_regression [N,4A,H,W], anchors [N,H*W*A,4] (the reference's
    flattened (h,w,a) order), im_wh [N,2] float (width, height).
    Returns (boxes [N*k,4], scores [N*k], index [N*k] int32, offsets [N+1] int32, k): image b
    owns rows offsets[b]:offsets[b+1], in rank order.  No host synchronisation.
    """
    objectness = _lib.as_f32(objectness, "objectness")
    box_regression = _lib.as_f32(box_regression, "box_regression")
    anchors = _lib.as_f32(anchors, "anchors")
    im_wh = _lib.as_f32(im_wh, "im_wh")
    if objectness.ndim != 4:
        raise ValueError("objectness should be [N,A,H,W]")
    N, A, H, W = (int(v) for v in objectness.shape)
    if tuple(box_regression.shape) != (N, 4 * A, H, W):
        raise ValueError("box_regression should be [%d,%d,%d,%d], got %s"
                         % (N, 4 * A, H, W, tuple(box_regression.shape)))
    if anchors.numel() != N * A * H * W * 4:
        raise ValueError("anchors should be [%d,%d,4], got %s" % (N, A * H * W, tuple(anchors.shape)))
    if im_wh.numel() != 2 * N:
        raise ValueError("im_wh should be [%d,2]" % N)
    dev = objectness.device
    k = min(int(pre_nms_top_n), A * H * W)
    boxes = torch.empty((N * k, 4), dtype=_F32, device=dev)
    scores = torch.empty((N * k,), dtype=_F32, device=dev)
    index = torch.empty((N * k,), dtype=_I32, device=dev)
    offsets = torch.empty((N + 1,), dtype=_I32, device=dev)
    L = _lib.lib()
    nbytes = L.rp_rpn_pre_nms_workspace_bytes(N, A, H, W, k)
    ws = _lib.workspace(nbytes, dev, "rpn")
    wx, wy, ww, wh = (float(w) for w in weights)
    rc = L.rp_rpn_pre_nms(
        _lib.dptr(objectness, _F32), _lib.dptr(box_regression, _F32), _lib.dptr(anchors, _F32),
        _lib.dptr(im_wh, _F32), N, A, H, W, int(pre_nms_top_n), float(min_size), wx, wy, ww, wh,
        float(bbox_xform_clip), _lib.dptr(boxes, _F32), _lib.dptr(scores, _F32),
        _lib.dptr(index, _I32), _lib.dptr(offsets, _I32), ctypes.c_void_p(ws.data_ptr()),
        ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_rpn_pre_nms")
    return boxes, scores, index, offsets, k


def rpn_pre_nms_levels(objectness, box_regression, anchors, im_wh, pre_nms_top_n, min_size,
                       weights=(1.0, 1.0, 1.0, 1.0), bbox_xform_clip=math.log(1000.0 / 16)):
    """All FPN levels of a batch in one set of launches (one sort for everything).

    objectness / box_regression / anchors: lists over levels of [N,A_l,H_l,W_l] /
    [N,4A_l,H_l,W_l] / [N,H_l*W_l*A_l,4].  Returns (boxes, scores, index, offsets [L*N+1], ks):
    problem s = l*N + b owns rows offsets[s]:offsets[s+1]; ks[l] = min(top_n, A_l*H_l*W_l).
    """
    L = len(objectness)
    if L < 1 or L > _lib.RP_MAX_LEVELS or len(box_regression) != L or len(anchors) != L:
        raise ValueError("between 1 and %d levels, one tensor of each kind per level" % _lib.RP_MAX_LEVELS)
    objectness = [_lib.as_f32(o, "objectness") for o in objectness]
    box_regression = [_lib.as_f32(r, "box_regression") for r in box_regression]
    anchors = [_lib.as_f32(a, "anchors") for a in anchors]
    im_wh = _lib.as_f32(im_wh, "im_wh")
    N = int(objectness[0].shape[0])
    As, Hs, Ws, ks = [], [], [], []
    for o, r, a in zip(objectness, box_regression, anchors):
        if o.ndim != 4 or int(o.shape[0]) != N:
            raise ValueError("objectness should be [N,A,H,W] with the same N on every level")
        _, A, H, W = (int(v) for v in o.shape)
        if tuple(r.shape) != (N, 4 * A, H, W):
            raise ValueError("box_regression should be [%d,%d,%d,%d], got %s"
                             % (N, 4 * A, H, W, tuple(r.shape)))
        if a.numel() != N * A * H * W * 4:
            raise ValueError("anchors should be [%d,%d,4], got %s" % (N, A * H * W, tuple(a.shape)))
        As.append(A), Hs.append(H), Ws.append(W), ks.append(min(int(pre_nms_top_n), A * H * W))
    if im_wh.numel() != 2 * N:
        raise ValueError("im_wh should be [%d,2]" % N)
    dev = objectness[0].device
    cap = N * sum(ks)
    boxes = torch.empty((cap, 4), dtype=_F32, device=dev)
    scores = torch.empty((cap,), dtype=_F32, device=dev)
    index = torch.empty((cap,), dtype=_I32, device=dev)
    offsets = torch.empty((L * N + 1,), dtype=_I32, device=dev)
    arr_i = lambda v: (ctypes.c_int * L)(*v)
    arr_p = lambda ts: (ctypes.c_void_p * L)(*[_lib.dptr(t, _F32).value for t in ts])
    cA, cH, cW = arr_i(As), arr_i(Hs), arr_i(Ws)
    lib = _lib.lib()
    nbytes = lib.rp_rpn_pre_nms_levels_workspace_bytes(L, N, cA, cH, cW, int(pre_nms_top_n))
    ws = _lib.workspace(nbytes, dev, "rpn")
    wx, wy, ww, wh = (float(w) for w in weights)
    rc = lib.rp_rpn_pre_nms_levels(
        arr_p(objectness), arr_p(box_regression), arr_p(anchors), cA, cH, cW, L,
        _lib.dptr(im_wh, _F32), N, int(pre_nms_top_n), float(min_size), wx, wy, ww, wh,
        float(bbox_xform_clip), _lib.dptr(boxes, _F32), _lib.dptr(scores, _F32),
        _lib.dptr(index, _I32), _lib.dptr(offsets, _I32), ctypes.c_void_p(ws.data_ptr()),
        ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_rpn_pre_nms_levels")
    return boxes, scores, index, offsets, ks


def rpn_select_levels(boxes, scores, offsets, keep, n_keep, L, N, top_n):
    """Per image, the top_n best of the boxes its L problems kept (inference.py:175-183).
    Returns (out_boxes [N,top_n,4], out_scores [N,top_n], out_count [N] int32); no host sync."""
    dev = boxes.device
    cap = int(boxes.shape[0])
    out_boxes = torch.empty((N, int(top_n), 4), dtype=_F32, device=dev)
    out_scores = torch.empty((N, int(top_n)), dtype=_F32, device=dev)
    out_count = torch.empty((N,), dtype=_I32, device=dev)
    lib = _lib.lib()
    nbytes = lib.rp_rpn_select_workspace_bytes(cap, L, N)
    ws = _lib.workspace(nbytes, dev, "rpn_select")
    rc = lib.rp_rpn_select_levels(
        _lib.dptr(boxes, _F32), _lib.dptr(scores, _F32), _lib.dptr(offsets, _I32),
        _lib.dptr(keep, _I32), _lib.dptr(n_keep, _I32), L, N, cap, int(top_n),
        _lib.dptr(out_boxes, _F32), _lib.dptr(out_scores, _F32), _lib.dptr(out_count, _I32),
        ctypes.c_void_p(ws.data_ptr()), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "rp_rpn_select_levels")
    return out_boxes, out_scores, out_count


class RPNPostProcessor(torch.nn.Module):
    """rpn/inference.py:15-188.  Device work for the whole batch and all levels: one
    rp_rpn_pre_nms_levels, one rp_nms_batched over the L*N problems and (test time) one
    rp_rpn_select_levels; the host reads N (or L*N) counts back ONCE to slice BoxLists."""

    def __init__(self, pre_nms_top_n, post_nms_top_n, nms_thresh, min_size, box_coder=None,
                 fpn_post_nms_top_n=None, fpn_post_nms_per_batch=True):
        super().__init__()
        self.pre_nms_top_n = pre_nms_top_n
        self.post_nms_top_n = post_nms_top_n
        self.nms_thresh = nms_thresh
        self.min_size = min_size
        if box_coder is None:
            box_coder = BoxCoder(weights=(1.0, 1.0, 1.0, 1.0))
        self.box_coder = box_coder
        if fpn_post_nms_top_n is None:
            fpn_post_nms_top_n = post_nms_top_n
        self.fpn_post_nms_top_n = fpn_post_nms_top_n
        self.fpn_post_nms_per_batch = fpn_post_nms_per_batch

    def is_training(self):
        return self.training

    def add_gt_proposals(self, proposals, targets):
        gt_boxes = [t.copy_with_fields([]) for t in targets]
        for g in gt_boxes:
            g.add_field("objectness", torch.ones(len(g), dtype=_F32, device=g.bbox.device))
        return [cat_boxlist((p, g)) for p, g in zip(proposals, gt_boxes)]

    def _stack_anchors(self, anchors):
        """list[BoxList] of one level -> [N, H*W*A, 4].  The reference builds new anchor tensors on
        every forward (anchor_generator.py grid_anchors), so nothing is cached by address here: one
        stack per level is a single small launch."""
        return torch.stack([a.bbox for a in anchors], dim=0).to(_F32).contiguous()

    # ---- device part: nothing here waits for the GPU
    def _launch(self, anchors_by_level, objectness, box_regression):
        L, N = len(objectness), int(objectness[0].shape[0])
        dev = objectness[0].device
        shapes = [a.size for a in anchors_by_level[0]]
        if len(shapes) != N:
            raise ValueError("%d anchor lists for %d images" % (len(shapes), N))
        stacked = [self._stack_anchors(a) for a in anchors_by_level]
        im_wh = torch.tensor([[float(s[0]), float(s[1])] for s in shapes], dtype=_F32).to(
            dev, non_blocking=True)
        bc = self.box_coder
        boxes, scores, index, offsets, ks = rpn_pre_nms_levels(
            objectness, box_regression, stacked, im_wh, self.pre_nms_top_n, self.min_size,
            bc.weights, bc.bbox_xform_clip)
        keep = n_keep = None
        if self.nms_thresh > 0:  # boxlist_nms returns its input otherwise (boxlist_ops.py:24-25)
            # the candidates of every (level, image) come out of the top-k in descending score order
            # (ties by ascending anchor index), i.e. as the reference's argsort would leave them
            keep, n_keep = layers.nms_batched(
                boxes, scores, offsets, self.nms_thresh, None,
                max_keep=self.post_nms_top_n if self.post_nms_top_n > 0 else -1, n_max=max(ks),
                flags=_lib.RP_NMS_PRESORTED)
        return dict(boxes=boxes, scores=scores, offsets=offsets, keep=keep, n_keep=n_keep,
                    shapes=shapes, N=N, L=L)

    @staticmethod
    def _collect(st):
        """One device->host copy (offsets + keep counts), then BoxLists per (level, image)."""
        parts = [st["offsets"]] + ([st["n_keep"]] if st["n_keep"] is not None else [])
        host = torch.cat(parts).cpu().tolist()
        S = st["L"] * st["N"]
        off, nk = host[:S + 1], host[S + 1:]
        out = []
        for l in range(st["L"]):
            row = []
            for b in range(st["N"]):
                s = l * st["N"] + b
                if st["keep"] is None:
                    bx, sc = st["boxes"][off[s]: off[s + 1]], st["scores"][off[s]: off[s + 1]]
                else:
                    idx = st["keep"][off[s]: off[s] + nk[s]].long() + off[s]
                    bx, sc = st["boxes"][idx], st["scores"][idx]
                bl = BoxList(bx, st["shapes"][b], mode="xyxy")
                bl.add_field("objectness", sc)
                row.append(bl)
            out.append(row)
        return out

    def forward_for_single_feature_map(self, anchors, objectness, box_regression):
        """anchors: list[BoxList]; objectness [N,A,H,W]; box_regression [N,4A,H,W]."""
        st = self._launch([list(anchors)], [objectness], [box_regression])
        return self._collect(st)[0]

    def forward(self, anchors, objectness, box_regression, targets=None):
        """anchors: list[list[BoxList]] (image-major); objectness, box_regression: list per level."""
        num_levels = len(objectness)
        st = self._launch(list(zip(*anchors)), list(objectness), list(box_regression))
        per_batch = self.is_training() and self.fpn_post_nms_per_batch
        if num_levels > 1 and not per_batch and st["keep"] is not None:
            ob, osc, cnt = rpn_select_levels(st["boxes"], st["scores"], st["offsets"], st["keep"],
                                             st["n_keep"], st["L"], st["N"], self.fpn_post_nms_top_n)
            boxlists = []
            for b, c in enumerate(cnt.cpu().tolist()):  # the one host read of the stage
                bl = BoxList(ob[b, :c], st["shapes"][b], mode="xyxy")
                bl.add_field("objectness", osc[b, :c])
                boxlists.append(bl)
        else:
            sampled = self._collect(st)
            boxlists = [cat_boxlist(bl) for bl in zip(*sampled)]
            if num_levels > 1:
                boxlists = self.select_over_all_levels(boxlists)
        if self.is_training() and targets is not None:
            boxlists = self.add_gt_proposals(boxlists, targets)
        return boxlists

    execute = forward

    def select_over_all_levels(self, boxlists):
        """inference.py:157-188 on BoxLists (the training per-batch branch, and the fallback of the
        test-time branch).  Stable: equal scores keep the lower index."""
        num_images = len(boxlists)
        if self.is_training() and self.fpn_post_nms_per_batch:
            objectness = torch.cat([b.get_field("objectness") for b in boxlists], dim=0)
            box_sizes = [len(b) for b in boxlists]
            post = min(self.fpn_post_nms_top_n, int(objectness.shape[0]))
            inds_sorted = torch.sort(objectness, descending=True, stable=True)[1][:post]
            inds_mask = torch.zeros(objectness.shape, dtype=torch.bool, device=objectness.device)
            inds_mask[inds_sorted] = True
            inds_mask = inds_mask.split(box_sizes, dim=0)
            for i in range(num_images):
                boxlists[i] = boxlists[i][inds_mask[i]]
        else:
            for i in range(num_images):
                objectness = boxlists[i].get_field("objectness")
                post = min(self.fpn_post_nms_top_n, int(objectness.shape[0]))
                inds_sorted = torch.sort(objectness, descending=True, stable=True)[1][:post]
                boxlists[i] = boxlists[i][inds_sorted]
        return boxlists


def make_rpn_postprocessor(config, rpn_box_coder, is_train):
    """inference.py:191-213; config is any object with the MODEL.RPN.* attributes."""
    rpn = config.MODEL.RPN
    return RPNPostProcessor(
        pre_nms_top_n=rpn.PRE_NMS_TOP_N_TRAIN if is_train else rpn.PRE_NMS_TOP_N_TEST,
        post_nms_top_n=rpn.POST_NMS_TOP_N_TRAIN if is_train else rpn.POST_NMS_TOP_N_TEST,
        nms_thresh=rpn.NMS_THRESH, min_size=rpn.MIN_SIZE, box_coder=rpn_box_coder,
        fpn_post_nms_top_n=rpn.FPN_POST_NMS_TOP_N_TRAIN if is_train else rpn.FPN_POST_NMS_TOP_N_TEST,
        fpn_post_nms_per_batch=rpn.FPN_POST_NMS_PER_BATCH)
