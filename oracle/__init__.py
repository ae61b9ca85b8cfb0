"""numpy front-end of the CPU oracle (oracle/roipath_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package never imports it.

All arrays are C-contiguous float32 / int32 numpy arrays; shapes follow the reference:
  feat [N,C,H,W], rois [K,5] (batch_idx,x1,y1,x2,y2), boxes [n,4] xyxy.
"""
import ctypes
import os

import numpy as np

from . import build as _build

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int)
_u8p = ctypes.POINTER(ctypes.c_uint8)


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_f32p)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_i32p)


_lib = None
_ref = None
_cy = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(_build.build_oracle())
        _lib.orc_roi_align_touched.restype = ctypes.c_longlong
        for name in ("orc_expf", "orc_sigmoidf", "orc_float_le", "orc_float_ge", "orc_iou_plus1"):
            getattr(_lib, name).restype = ctypes.c_float
        _lib.orc_expf.argtypes = [ctypes.c_float]
        _lib.orc_sigmoidf.argtypes = [ctypes.c_float]
        _lib.orc_float_le.argtypes = [ctypes.c_double]
        _lib.orc_float_ge.argtypes = [ctypes.c_double]
    return _lib


def ref_lib():
    """The reference's own RoIAlign kernel text compiled as C++ (None if unavailable)."""
    global _ref
    if _ref is None:
        paths = _build.build_ref()
        p = paths.get("libroialign_ref")
        _ref = ctypes.CDLL(p) if p and os.path.exists(p) else False
    return _ref or None


def ref_cython_nms():
    """The reference's cython_nms.nms compiled here (None if unavailable)."""
    global _cy
    if _cy is None:
        paths = _build.build_ref()
        p = paths.get("cython_nms_ref")
        if p and os.path.exists(p):
            import importlib.util
            spec = importlib.util.spec_from_file_location("cython_nms_ref", p)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            _cy = mod.nms
        else:
            _cy = False
    return _cy or None


# --------------------------------------------------------------------------- ROIAlign
def _roi_align_fwd(fn, feat, rois, output_size, spatial_scale, sampling_ratio):
    feat, fp = _f(feat)
    rois, rp = _f(rois)
    N, C, H, W = feat.shape
    K = rois.shape[0]
    PH, PW = (output_size, output_size) if np.isscalar(output_size) else output_size
    out = np.zeros((K, C, PH, PW), np.float32)
    if K:
        fn(fp, N, C, H, W, rp, K, PH, PW, ctypes.c_float(spatial_scale),
           ctypes.c_float(sampling_ratio), out.ctypes.data_as(_f32p))
    return out


def _roi_align_bwd(fn, dout, rois, feat_shape, spatial_scale, sampling_ratio):
    dout, dp = _f(dout)
    rois, rp = _f(rois)
    N, C, H, W = feat_shape
    K, _, PH, PW = dout.shape
    g = np.zeros((N, C, H, W), np.float32)
    fn(dp, rp, K, N, C, H, W, PH, PW, ctypes.c_float(spatial_scale),
       ctypes.c_float(sampling_ratio), g.ctypes.data_as(_f32p))
    return g


def roi_align_forward(feat, rois, output_size, spatial_scale, sampling_ratio):
    return _roi_align_fwd(lib().orc_roi_align_forward, feat, rois, output_size, spatial_scale,
                          sampling_ratio)


def roi_align_backward(dout, rois, feat_shape, spatial_scale, sampling_ratio):
    return _roi_align_bwd(lib().orc_roi_align_backward, dout, rois, feat_shape, spatial_scale,
                          sampling_ratio)


def ref_roi_align_forward(feat, rois, output_size, spatial_scale, sampling_ratio):
    return _roi_align_fwd(ref_lib().ref_roi_align_forward, feat, rois, output_size,
                          spatial_scale, sampling_ratio)


def ref_roi_align_backward(dout, rois, feat_shape, spatial_scale, sampling_ratio):
    return _roi_align_bwd(ref_lib().ref_roi_align_backward, dout, rois, feat_shape,
                          spatial_scale, sampling_ratio)


def roi_align_touched(feat_shape, rois, output_size, spatial_scale, sampling_ratio):
    """Distinct feature elements read by a forward call (SURVEY.md 8(d) 'U')."""
    rois, rp = _f(rois)
    N, C, H, W = feat_shape
    PH, PW = (output_size, output_size) if np.isscalar(output_size) else output_size
    return int(lib().orc_roi_align_touched(N, C, H, W, rp, rois.shape[0], PH, PW,
                                           ctypes.c_float(spatial_scale),
                                           ctypes.c_float(sampling_ratio)))


# --------------------------------------------------------------------------- Pooler
def level_map(boxes, k_min, k_max, canonical_scale=224, canonical_level=4, eps=1e-6):
    boxes, bp = _f(boxes)
    lv = np.zeros((boxes.shape[0],), np.int32)
    lib().orc_level_map(bp, boxes.shape[0], int(k_min), int(k_max),
                        ctypes.c_float(canonical_scale), int(canonical_level),
                        ctypes.c_float(eps), lv.ctypes.data_as(_i32p))
    return lv


def level_map_area(boxes, img_area, k_min, k_max, eps=1e-6):
    boxes, bp = _f(boxes)
    img_area, ap = _f(img_area)
    lv = np.zeros((boxes.shape[0],), np.int32)
    lib().orc_level_map_area(bp, ap, boxes.shape[0], int(k_min), int(k_max), ctypes.c_float(eps),
                             lv.ctypes.data_as(_i32p))
    return lv


def pooler_forward(feats, scales, rois, levels, output_size, sampling_ratio):
    feats = [np.ascontiguousarray(f, np.float32) for f in feats]
    L = len(feats)
    N, C = feats[0].shape[:2]
    rois, rp = _f(rois)
    levels, lp = _i(levels)
    K = rois.shape[0]
    PH, PW = (output_size, output_size) if np.isscalar(output_size) else output_size
    fptr = (_f32p * L)(*[f.ctypes.data_as(_f32p) for f in feats])
    Hs = (ctypes.c_int * L)(*[f.shape[2] for f in feats])
    Ws = (ctypes.c_int * L)(*[f.shape[3] for f in feats])
    sc = (ctypes.c_float * L)(*[float(s) for s in scales])
    out = np.zeros((K, C, PH, PW), np.float32)
    lib().orc_pooler_forward(fptr, Hs, Ws, sc, L, N, C, rp, lp, K, PH, PW,
                             ctypes.c_float(sampling_ratio), out.ctypes.data_as(_f32p))
    return out


def pooler_backward(dout, rois, levels, feat_shapes, scales, sampling_ratio):
    dout, dp = _f(dout)
    rois, rp = _f(rois)
    levels, lp = _i(levels)
    L = len(feat_shapes)
    N, C = feat_shapes[0][:2]
    K, _, PH, PW = dout.shape
    grads = [np.zeros(s, np.float32) for s in feat_shapes]
    gptr = (_f32p * L)(*[g.ctypes.data_as(_f32p) for g in grads])
    Hs = (ctypes.c_int * L)(*[s[2] for s in feat_shapes])
    Ws = (ctypes.c_int * L)(*[s[3] for s in feat_shapes])
    sc = (ctypes.c_float * L)(*[float(s) for s in scales])
    lib().orc_pooler_backward(dp, rp, lp, K, Hs, Ws, sc, L, N, C, PH, PW,
                              ctypes.c_float(sampling_ratio), gptr)
    return grads


# --------------------------------------------------------------------------- NMS
def nms(boxes, scores, threshold, labels=None, max_keep=-1):
    """Greedy NMS (+1, strict >, stable descending sort). Returns keep (int32, desc score)."""
    boxes, bp = _f(boxes)
    scores, sp = _f(scores)
    n = boxes.shape[0]
    keep = np.zeros((max(n, 1),), np.int32)
    if labels is not None:
        labels, lp = _f(labels)
    else:
        lp = None
    m = lib().orc_nms(bp, sp, n, ctypes.c_double(threshold), lp, int(max_keep),
                      keep.ctypes.data_as(_i32p))
    return keep[:m].copy()


def ml_nms(boxes, scores, labels, threshold, max_keep=-1):
    return nms(boxes, scores, threshold, labels=labels, max_keep=max_keep)


def cython_nms(dets, thresh):
    dets, dp = _f(dets)
    n = dets.shape[0]
    keep = np.zeros((max(n, 1),), np.int32)
    m = lib().orc_cython_nms(dp, n, ctypes.c_float(thresh), keep.ctypes.data_as(_i32p))
    return keep[:m].copy()


def argsort_desc(key):
    key, kp = _f(key)
    order = np.zeros((key.shape[0],), np.int32)
    lib().orc_argsort_desc(kp, key.shape[0], order.ctypes.data_as(_i32p))
    return order


# --------------------------------------------------------------------------- RPN
def box_decode(rel_codes, boxes, weights=(1.0, 1.0, 1.0, 1.0), clip=float(np.log(1000.0 / 16))):
    rel_codes, cp = _f(rel_codes)
    boxes, bp = _f(boxes)
    n, K = rel_codes.shape[0], rel_codes.shape[1] // 4
    out = np.zeros_like(rel_codes)
    cf = ctypes.c_float
    lib().orc_box_decode(cp, bp, n, K, cf(weights[0]), cf(weights[1]), cf(weights[2]),
                         cf(weights[3]), cf(clip), out.ctypes.data_as(_f32p))
    return out


def rpn_pre_nms(objectness, box_regression, anchors, im_wh, pre_nms_top_n, min_size,
                weights=(1.0, 1.0, 1.0, 1.0), clip=float(np.log(1000.0 / 16))):
    """Per image lists of (boxes [m,4], scores [m], index [m]) in rank order."""
    objectness, op = _f(objectness)
    box_regression, rp = _f(box_regression)
    anchors, ap = _f(anchors)
    im_wh, wp = _f(im_wh)
    N, A, H, W = objectness.shape
    k = min(int(pre_nms_top_n), A * H * W)
    boxes = np.zeros((N, k, 4), np.float32)
    scores = np.zeros((N, k), np.float32)
    index = np.zeros((N, k), np.int32)
    counts = np.zeros((N,), np.int32)
    cf = ctypes.c_float
    lib().orc_rpn_pre_nms(op, rp, ap, wp, N, A, H, W, int(pre_nms_top_n), cf(min_size),
                          cf(weights[0]), cf(weights[1]), cf(weights[2]), cf(weights[3]), cf(clip),
                          boxes.ctypes.data_as(_f32p), scores.ctypes.data_as(_f32p),
                          index.ctypes.data_as(_i32p), counts.ctypes.data_as(_i32p))
    return [(boxes[b, :counts[b]].copy(), scores[b, :counts[b]].copy(), index[b, :counts[b]].copy())
            for b in range(N)]


# --------------------------------------------------------------------------- box head
def softmax_rows(x):
    x, xp = _f(x)
    out = np.zeros_like(x)
    lib().orc_softmax_rows(xp, x.shape[0], x.shape[1], out.ctypes.data_as(_f32p))
    return out


def box_post_image(logits, reg, props, im_wh, score_thresh=0.05, nms_thresh=0.5, dets=100,
                   weights=(10.0, 10.0, 5.0, 5.0), clip=float(np.log(1000.0 / 16)),
                   cls_agnostic=False):
    """One image of PostProcessor.execute + filter_results -> (boxes, scores, labels, rows)."""
    logits, lp = _f(logits)
    reg, rp_ = _f(reg)
    props, pp = _f(props)
    n, C = logits.shape
    cap = max(n * (C - 1), 1)
    ob = np.zeros((cap, 4), np.float32)
    osc = np.zeros((cap,), np.float32)
    ol = np.zeros((cap,), np.int32)
    orow = np.zeros((cap,), np.int32)
    cf = ctypes.c_float
    fn = lib().orc_box_post_image
    fn.restype = ctypes.c_int
    m = fn(lp, rp_, pp, n, C, reg.shape[1], int(bool(cls_agnostic)), cf(im_wh[0]), cf(im_wh[1]),
           cf(score_thresh), ctypes.c_double(nms_thresh), int(dets), cf(weights[0]), cf(weights[1]),
           cf(weights[2]), cf(weights[3]), cf(clip), ob.ctypes.data_as(_f32p),
           osc.ctypes.data_as(_f32p), ol.ctypes.data_as(_i32p), orow.ctypes.data_as(_i32p))
    return ob[:m].copy(), osc[:m].copy(), ol[:m].copy(), orow[:m].copy()


# --------------------------------------------------------------------------- other RoI ops
def boxlist_iou(a, b):
    a, ap = _f(a)
    b, bp = _f(b)
    out = np.zeros((a.shape[0], b.shape[0]), np.float32)
    lib().orc_boxlist_iou(ap, a.shape[0], bp, b.shape[0], out.ctypes.data_as(_f32p))
    return out


def roi_pool_forward(feat, rois, output_size, spatial_scale):
    feat, fp = _f(feat)
    rois, rp_ = _f(rois)
    N, C, H, W = feat.shape
    PH, PW = output_size if isinstance(output_size, (tuple, list)) else (output_size, output_size)
    K = rois.shape[0]
    out = np.zeros((K, C, PH, PW), np.float32)
    arg = np.zeros((K, C, PH, PW), np.int32)
    lib().orc_roi_pool_forward(fp, N, C, H, W, rp_, K, PH, PW, ctypes.c_float(spatial_scale),
                               out.ctypes.data_as(_f32p), arg.ctypes.data_as(_i32p))
    return out, arg


def roi_pool_backward(dout, rois, argmax, feat_shape):
    dout, dp = _f(dout)
    rois, rp_ = _f(rois)
    argmax, ap = _i(argmax)
    N, C, H, W = feat_shape
    K, _, PH, PW = dout.shape
    grad = np.zeros((N, C, H, W), np.float32)
    lib().orc_roi_pool_backward(dp, rp_, ap, K, N, C, H, W, PH, PW, grad.ctypes.data_as(_f32p))
    return grad


def mask_prob(embed_pixel, embed_center, sigma_center, boxes, box_areas, mask_width):
    ep, epp = _f(embed_pixel)
    ec, ecp = _f(embed_center)
    sg, sgp = _f(sigma_center)
    bx, bxp = _i(boxes)
    ar, arp = _i(box_areas)
    out = np.zeros((ec.shape[0], ep.shape[0]), np.float32)
    lib().orc_mask_prob(epp, ecp, sgp, bxp, arp, ep.shape[0], ec.shape[0], ep.shape[1],
                        int(mask_width), out.ctypes.data_as(_f32p))
    return out


def box_masks(boxes, h, w, padding=0.0, crop=False, masks=None):
    boxes, bp = _f(boxes)
    n = boxes.shape[0]
    ind = np.zeros((n, h, w), np.uint8)
    mi = mo = None
    mip = mop = None
    if masks is not None:
        mi, mip = _f(masks)
        mo = np.zeros_like(mi)
        mop = mo.ctypes.data_as(_f32p)
    lib().orc_box_masks(bp, n, int(h), int(w), ctypes.c_float(padding), int(bool(crop)), mip, mop,
                        ind.ctypes.data_as(_u8p))
    return mo, ind


# --------------------------------------------------------------------------- Yolact
def yolact_decode(loc, priors):
    loc, lp = _f(loc)
    priors, pp = _f(priors)
    out = np.zeros_like(loc)
    lib().orc_yolact_decode(lp, pp, loc.shape[0], out.ctypes.data_as(_f32p))
    return out


def jaccard(a, b):
    a, ap = _f(a)
    b, bp = _f(b)
    out = np.zeros((a.shape[0], b.shape[0]), np.float32)
    lib().orc_jaccard(ap, a.shape[0], bp, b.shape[0], out.ctypes.data_as(_f32p))
    return out


def fast_nms(boxes, coeff, scores, iou_threshold=0.5, top_k=200, max_det=100,
             second_threshold=False, conf_thresh=0.05):
    """Returns (boxes[m,4], coeff[m,D], classes[m], scores[m], index[m])."""
    boxes, bp = _f(boxes)
    coeff, cp = _f(coeff)
    scores, sp = _f(scores)
    C, n = scores.shape
    D = coeff.shape[1]
    ob = np.zeros((max_det, 4), np.float32)
    oc = np.zeros((max_det, D), np.float32)
    ocl = np.zeros((max_det,), np.int32)
    osc = np.zeros((max_det,), np.float32)
    oi = np.zeros((max_det,), np.int32)
    m = lib().orc_fast_nms(bp, cp, sp, n, C, D, ctypes.c_float(iou_threshold), int(top_k),
                           int(max_det), int(bool(second_threshold)), ctypes.c_float(conf_thresh),
                           ob.ctypes.data_as(_f32p), oc.ctypes.data_as(_f32p),
                           ocl.ctypes.data_as(_i32p), osc.ctypes.data_as(_f32p),
                           oi.ctypes.data_as(_i32p))
    return ob[:m].copy(), oc[:m].copy(), ocl[:m].copy(), osc[:m].copy(), oi[:m].copy()


def cc_fast_nms(boxes, coeff, scores, iou_threshold=0.5, top_k=200):
    boxes, bp = _f(boxes)
    coeff, cp = _f(coeff)
    scores, sp = _f(scores)
    C, n = scores.shape
    D = coeff.shape[1]
    k = max(min(n, top_k), 1)
    ob = np.zeros((k, 4), np.float32)
    oc = np.zeros((k, D), np.float32)
    ocl = np.zeros((k,), np.int32)
    osc = np.zeros((k,), np.float32)
    oi = np.zeros((k,), np.int32)
    m = lib().orc_cc_fast_nms(bp, cp, sp, n, C, D, ctypes.c_float(iou_threshold), int(top_k),
                              ob.ctypes.data_as(_f32p), oc.ctypes.data_as(_f32p),
                              ocl.ctypes.data_as(_i32p), osc.ctypes.data_as(_f32p),
                              oi.ctypes.data_as(_i32p))
    return ob[:m].copy(), oc[:m].copy(), ocl[:m].copy(), osc[:m].copy(), oi[:m].copy()


def detect_filter(conf_fg, conf_thresh):
    conf_fg, cp = _f(conf_fg)
    C, P = conf_fg.shape
    keep = np.zeros((max(P, 1),), np.int32)
    m = lib().orc_detect_filter(cp, C, P, ctypes.c_float(conf_thresh), keep.ctypes.data_as(_i32p))
    return keep[:m].copy()


# --------------------------------------------------------------------------- masks
def bilinear_resize(img, out_h, out_w):
    img, ip = _f(img)
    out = np.zeros((out_h, out_w), np.float32)
    lib().orc_bilinear_resize(ip, img.shape[0], img.shape[1], out.ctypes.data_as(_f32p), out_h,
                              out_w)
    return out


def mask_assemble(proto, coeff, boxes, out_h, out_w, crop=True, return_low=False):
    """Returns (masks u8 [m,out_h,out_w], boxes int32 [m,4][, low f32 [m,ph,pw]])."""
    proto, pp = _f(proto)
    coeff, cp = _f(coeff)
    boxes, bp = _f(boxes)
    ph, pw, D = proto.shape
    m = coeff.shape[0]
    masks = np.zeros((m, out_h, out_w), np.uint8)
    ob = np.zeros((m, 4), np.int32)
    low = np.zeros((m, ph, pw), np.float32) if return_low else None
    lib().orc_mask_assemble(pp, cp, bp, m, ph, pw, D, out_h, out_w, int(bool(crop)),
                            masks.ctypes.data_as(_u8p), ob.ctypes.data_as(_i32p),
                            low.ctypes.data_as(_f32p) if return_low else None)
    return (masks, ob, low) if return_low else (masks, ob)


def mask_select_sigmoid(logits, labels):
    logits, lp = _f(logits)
    labels, lb = _i(labels)
    K, NC, M, _ = logits.shape
    out = np.zeros((K, M, M), np.float32)
    lib().orc_mask_select_sigmoid(lp, lb, K, NC, M, out.ctypes.data_as(_f32p))
    return out


def paste_masks(masks, boxes, im_h, im_w, thresh=0.5, padding=1):
    """masks [n,M,M] probabilities, boxes [n,4] -> u8 [n,im_h,im_w]."""
    masks, mp = _f(masks)
    boxes, bp = _f(boxes)
    n, M, _ = masks.shape
    out = np.zeros((n, im_h, im_w), np.uint8)
    lib().orc_paste_masks(mp, bp, n, M, im_h, im_w, ctypes.c_float(thresh), int(padding),
                          out.ctypes.data_as(_u8p))
    return out


def expf(x):
    return float(lib().orc_expf(ctypes.c_float(x)))


def sigmoidf(x):
    return float(lib().orc_sigmoidf(ctypes.c_float(x)))


def lincomb_mask_loss(proto, coef, mask_t, boxes=None, scale=None):
    """Returns (loss [n], d_proto [H,W,D], d_coef [n,D]) of sum_i loss[i]."""
    proto, pp = _f(proto)
    coef, cp = _f(coef)
    mask_t, mp = _f(mask_t)
    H, W, D = proto.shape
    n = coef.shape[0]
    bp = sp = None
    if boxes is not None:
        boxes, bp = _f(boxes)
    if scale is not None:
        scale, sp = _f(scale)
    loss = np.zeros((n,), np.float32)
    dpr = np.zeros((H, W, D), np.float32)
    dco = np.zeros((n, D), np.float32)
    lib().orc_lincomb_mask_loss(pp, cp, mp, bp, sp, H, W, D, n, loss.ctypes.data_as(_f32p),
                                dpr.ctypes.data_as(_f32p), dco.ctypes.data_as(_f32p))
    return loss, dpr, dco


def affine_paste(pred, M, width, height, channel=1, return_values=False):
    """pred [n,h,w,C] float32, M [n,2,3] float64 (dst -> src) -> u8 [n,height,width] (value > 0.5)."""
    pred, pp = _f(pred)
    M = np.ascontiguousarray(M, np.float64).reshape(-1, 6)
    n, h, w, C = pred.shape
    out = np.zeros((n, height, width), np.uint8)
    vals = np.zeros((n, height, width), np.float32) if return_values else None
    lib().orc_affine_paste(pp, M.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), n, h, w, C, int(channel),
                           int(width), int(height), out.ctypes.data_as(_u8p),
                           vals.ctypes.data_as(_f32p) if return_values else None)
    return (out, vals) if return_values else out


def embed_masks(pixel_embed, proposal_embed, proposal_margin, boxes, stride):
    """det/modeling/rpn/embed_mask/inference.py:153-169 + :193-197 composed from the restated pieces:
    pixel_embed [dim,m_h,m_w], boxes [n,4] image pixels -> masks [n,m_h,m_w] float."""
    pe = np.ascontiguousarray(pixel_embed, np.float32)
    dim, m_h, m_w = pe.shape
    n = proposal_embed.shape[0]
    mask_boxes = (np.asarray(boxes, np.float32) / np.float32(stride)).astype(np.float32)
    ib = mask_boxes.astype(np.int32)  # .int(): truncation toward zero
    ib[:, 0] = np.maximum(ib[:, 0] - 2, 0)
    ib[:, 1] = np.maximum(ib[:, 1] - 2, 0)
    ib[:, 2] = np.minimum(ib[:, 2] + 2, m_w)
    ib[:, 3] = np.minimum(ib[:, 3] + 2, m_h)
    areas = ((ib[:, 2] - ib[:, 0]) * (ib[:, 3] - ib[:, 1])).astype(np.int32)
    ep = np.ascontiguousarray(pe.reshape(dim, m_h * m_w).T)
    prob = mask_prob(ep, proposal_embed, proposal_margin, ib, areas, m_w).reshape(n, m_h, m_w)
    _, ind = box_masks(mask_boxes, m_h, m_w, 0.0, crop=False)
    return (prob * ind.astype(np.float32)).astype(np.float32)


def resize_threshold(masks, H, W, thresh, crop_h, crop_w):
    """inference.py:199-202: bilinear resize (align_corners=False) > thresh, cropped."""
    out = np.zeros((masks.shape[0], crop_h, crop_w), np.uint8)
    for i in range(masks.shape[0]):
        out[i] = (bilinear_resize(masks[i], H, W) > np.float32(thresh))[:crop_h, :crop_w]
    return out
