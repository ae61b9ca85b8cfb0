"""Golden fixtures produced by EXECUTING the reference's own Python text under the torch-backed
`jittor` stand-in (oracle/jt_shim) -- run in the BUILD container (needs /root/reference):

    python tests/golden/make_golden_jt.py

Writes tests/golden/jt_*.npz; tests/test_oracle_golden.py requires the C oracle to reproduce them
(keeps / classes / level ids / binary masks bit-exact, float masks to 1e-6).  What is executed:

  jt_yolact.npz   Yolact.jittor/layers/box_utils.py            jaccard, crop, sanitize_coordinates
                  Yolact.jittor/layers/functions/detection.py  Detect.fast_nms, Detect.cc_fast_nms
                  Yolact.jittor/layers/output_utils.py         postprocess (lincomb branch)
  jt_detectron.npz
                  detectron/modeling/poolers.py                LevelMapper, LevelMapperwithArea
                  detectron/structures/bounding_box.py         BoxList.area / image_area / clip
                  detectron/layers/nms.py                      nms, ml_nms (_ml_nms + jt.candidate)
                  detectron/structures/boxlist_ops.py          boxlist_nms, boxlist_ml_nms
                  detectron/modeling/roi_heads/mask_head/inference.py
                                                               expand_boxes, expand_masks,
                                                               paste_mask_in_image, Masker
The semantics the stand-in assumes for Jittor's own ops are listed in oracle/jt_shim/jittor/__init__.py.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "jt_shim"))
REF = "/root/reference"
YOL = os.path.join(REF, "Yolact.jittor")
DET = os.path.join(REF, "detectron.jittor", "detectron")

import jittor as jt  # noqa: E402  (the stand-in)
from tests import synth  # noqa: E402


def _pkg(name, path=None):
    m = types.ModuleType(name)
    m.__path__ = [path] if path else []
    sys.modules[name] = m
    return m


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


class _Cfg(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def load_yolact():
    """The three Yolact files, with `data` (config) and `utils` (timer) replaced by the handful
    of values the executed functions read (Yolact.jittor/data/config.py:306-312,423,463)."""
    class _Env:
        def __init__(self, *_):
            pass

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    utils = _pkg("utils")
    utils.timer = types.SimpleNamespace(env=_Env)
    aug = types.ModuleType("utils.augmentations")
    aug.Resize = object
    sys.modules["utils.augmentations"] = aug
    data = _pkg("data")
    mask_type = _Cfg(direct=0, lincomb=1)
    data.mask_type = mask_type
    data.activation_func = _Cfg(sigmoid=jt.sigmoid)
    data.MEANS, data.STD = (0, 0, 0), (1, 1, 1)
    data.cfg = _Cfg(max_num_detections=100, max_size=550, mask_type=mask_type.lincomb,
                    eval_mask_branch=True, mask_proto_debug=False, use_maskiou=False,
                    mask_proto_mask_activation=jt.sigmoid, rescore_mask=False, rescore_bbox=False,
                    mask_size=16)
    _pkg("yol_layers", os.path.join(YOL, "layers"))
    _pkg("yol_layers.functions", os.path.join(YOL, "layers", "functions"))
    bu = _load("yol_layers.box_utils", os.path.join(YOL, "layers", "box_utils.py"))
    dt = _load("yol_layers.functions.detection", os.path.join(YOL, "layers", "functions", "detection.py"))
    ou = _load("yol_layers.output_utils", os.path.join(YOL, "layers", "output_utils.py"))
    return bu, dt, ou, data.cfg


def load_detectron():
    _pkg("detectron")
    layers = _pkg("detectron.layers")
    nms_mod = _load("detectron.layers.nms", os.path.join(DET, "layers", "nms.py"))
    layers.nms, layers.ml_nms, layers.ROIAlign = nms_mod.nms, nms_mod.ml_nms, None
    misc = types.ModuleType("detectron.layers.misc")
    # detectron/layers/misc.py:68-74: non-empty input goes straight to nn.interpolate
    misc.interpolate = lambda input, size=None, scale_factor=None, mode="nearest", align_corners=None: \
        jt.nn.interpolate(input, size, scale_factor, mode, align_corners)
    sys.modules["detectron.layers.misc"] = misc
    _pkg("detectron.structures", os.path.join(DET, "structures"))
    bb = _load("detectron.structures.bounding_box", os.path.join(DET, "structures", "bounding_box.py"))
    ops = _load("detectron.structures.boxlist_ops", os.path.join(DET, "structures", "boxlist_ops.py"))
    _pkg("detectron.modeling", os.path.join(DET, "modeling"))
    _load("detectron.modeling.utils", os.path.join(DET, "modeling", "utils.py"))
    pl = _load("detectron.modeling.poolers", os.path.join(DET, "modeling", "poolers.py"))
    _pkg("detectron.modeling.roi_heads", os.path.join(DET, "modeling", "roi_heads"))
    _pkg("detectron.modeling.roi_heads.mask_head", os.path.join(DET, "modeling", "roi_heads", "mask_head"))
    mi = _load("detectron.modeling.roi_heads.mask_head.inference",
               os.path.join(DET, "modeling", "roi_heads", "mask_head", "inference.py"))
    return nms_mod, bb, ops, pl, mi


def N(v):
    return v.numpy() if hasattr(v, "numpy") else np.asarray(v)


# --------------------------------------------------------------------------------------------
def make_yolact():
    bu, dt, ou, cfg = load_yolact()
    out = {}
    rng = np.random.default_rng(20261105)

    # ---- jaccard on relative boxes (degenerate / touching boxes included)
    a = rng.uniform(0, 1, (40, 4)).astype(np.float32)
    a = np.stack([np.minimum(a[:, 0], a[:, 2]), np.minimum(a[:, 1], a[:, 3]),
                  np.maximum(a[:, 0], a[:, 2]), np.maximum(a[:, 1], a[:, 3])], 1)
    a[3] = a[2]                       # identical boxes
    a[5, 2] = a[5, 0]                 # zero width
    a[7] = (0.25, 0.25, 0.5, 0.5)
    a[8] = (0.5, 0.25, 0.75, 0.5)     # touching
    out["jac_boxes"] = a
    with np.errstate(all="ignore"):
        out["jac_iou"] = N(bu.jaccard(jt.array(a), jt.array(a)))

    # ---- sanitize_coordinates (cast and no cast, swapped inputs, padding)
    x1 = rng.uniform(-0.2, 1.2, 64).astype(np.float32)
    x2 = rng.uniform(-0.2, 1.2, 64).astype(np.float32)
    out["san_x1"], out["san_x2"] = x1, x2
    for name, (size, pad, cast) in {"a": (138, 1, False), "b": (550, 0, False), "c": (550, 0, True)}.items():
        lo, hi = bu.sanitize_coordinates(jt.array(x1), jt.array(x2), size, pad, cast=cast)
        out["san_%s_lo" % name], out["san_%s_hi" % name] = N(lo).astype(np.float32), N(hi).astype(np.float32)
        out["san_%s_args" % name] = np.array([size, pad, int(cast)])

    # ---- FastNMS / cross-class FastNMS on two candidate sets (one tie-heavy)
    priors = synth.yolact_priors()
    import oracle
    for tag, seed, ties in (("a", 11, False), ("b", 12, True)):
        loc, conf, mask, proto = synth.yolact_image(seed, priors)
        boxes_all = oracle.yolact_decode(loc, priors)          # decode is pinned by its own test
        fg = np.ascontiguousarray(conf[:, 1:].T)               # [80, P]
        keep = np.nonzero(fg.max(0) > 0.05)[0][:600]
        b, m, s = boxes_all[keep], mask[keep], np.ascontiguousarray(fg[:, keep])
        if ties:
            s = (np.floor(s * 64) / 64).astype(np.float32)
        det = dt.Detect(81, 0, 200, 0.05, 0.5)
        rb, rm, rc, rs = det.fast_nms(jt.array(b), jt.array(m), jt.array(s), 0.5, 200)
        out["fn_%s_boxes" % tag], out["fn_%s_coeff" % tag], out["fn_%s_scores" % tag] = b, m, s
        out["fn_%s_out_boxes" % tag], out["fn_%s_out_coeff" % tag] = N(rb), N(rm)
        out["fn_%s_out_classes" % tag], out["fn_%s_out_scores" % tag] = N(rc).astype(np.int32), N(rs)
        cb, cm, cc, cs = det.cc_fast_nms(jt.array(b), jt.array(m), jt.array(s), 0.5, 200)
        out["cc_%s_out_boxes" % tag], out["cc_%s_out_coeff" % tag] = N(cb), N(cm)
        out["cc_%s_out_classes" % tag], out["cc_%s_out_scores" % tag] = N(cc).astype(np.int32), N(cs)
        if tag == "a":
            # ---- postprocess (lincomb): proto @ coeff^T -> sigmoid -> crop -> resize -> > 0.5
            n = min(12, rb.shape[0])
            for ptag, (w, h) in {"p": (101, 77), "q": (138, 138), "r": (301, 137)}.items():
                dets = {"class": rc[:n], "box": jt.array(N(rb)[:n].copy()), "score": rs[:n],
                        "mask": rm[:n], "proto": jt.array(proto)}
                cls_, sc_, bx_, mk_ = ou.postprocess([{"net": None, "detection": dets}], w, h)
                out["pp_%s_wh" % ptag] = np.array([w, h])
                out["pp_%s_masks" % ptag] = np.packbits(N(mk_).astype(np.uint8), axis=None)
                out["pp_%s_boxes" % ptag] = N(bx_).astype(np.int32)
            # the prototypes are synth.yolact_image(11, priors)[3]: regenerated by the test, only a
            # checksum is stored
            out["pp_proto_seed"], out["pp_proto_sum"] = np.array([seed]), np.array([proto.astype(np.float64).sum()])
            out["pp_coeff"], out["pp_in_boxes"] = N(rm)[:n], N(rb)[:n]
            # the low-resolution stage alone (matmul + sigmoid + crop), float
            low = bu.crop(jt.sigmoid(jt.matmul(jt.array(proto), jt.array(N(rm)[:n]).transpose(1, 0))),
                          jt.array(N(rb)[:n]))
            out["pp_low"] = N(low.permute(2, 0, 1))[:4]
    np.savez_compressed(os.path.join(HERE, "jt_yolact.npz"), **out)
    print("jt_yolact.npz:", {k: v.shape for k, v in out.items() if k.startswith(("fn_a_out", "cc_a_out", "pp_p"))})


# --------------------------------------------------------------------------------------------
def make_detectron():
    nms_mod, bb, ops, pl, mi = load_detectron()
    out = {}
    rng = np.random.default_rng(20261106)

    # ---- LevelMapper / LevelMapperwithArea on BoxLists (reference area(): +1 convention)
    boxes = synth.proposals(400, 5, adversarial=True)
    # exact level boundaries: sqrt(area) = 112, 224, 448 (+1 convention => side - 1)
    boxes[:3] = [[10, 10, 10 + 111, 10 + 111], [10, 10, 10 + 223, 10 + 223], [10, 10, 10 + 447, 10 + 447]]
    bl = bb.BoxList(jt.array(boxes), (synth.IMG_W, synth.IMG_H), mode="xyxy")
    out["lm_boxes"] = boxes
    out["lm_levels"] = N(pl.LevelMapper(2, 5)([bl])).astype(np.int32)
    out["lm_levels_36"] = N(pl.LevelMapper(3, 6, canonical_scale=160, canonical_level=3)([bl])).astype(np.int32)
    out["lm_img_area"] = N(bl.image_area()).astype(np.float32) if hasattr(bl, "image_area") else \
        np.full((boxes.shape[0],), synth.IMG_W * synth.IMG_H, np.float32)
    if hasattr(bl, "image_area"):
        out["lm_levels_area"] = N(pl.LevelMapperwithArea(2, 5)([bl])).astype(np.int32)
    out["bl_area"] = N(bl.area()).astype(np.float32)

    # ---- nms / ml_nms / boxlist_nms / boxlist_ml_nms
    for tag, n, seed, ties in (("a", 300, 3, False), ("b", 300, 4, True)):
        b, s, l = synth.nms_boxes(n, seed, clusters=24, tie_heavy=ties)
        l = (l % 5).astype(np.float32)
        out["nms_%s_boxes" % tag], out["nms_%s_scores" % tag], out["nms_%s_labels" % tag] = b, s, l
        for thr in (0.3, 0.5, 0.7):
            k = nms_mod.nms(jt.array(b), jt.array(s), thr)
            out["nms_%s_keep_%d" % (tag, int(thr * 10))] = N(k).astype(np.int32)
            k = nms_mod.ml_nms(jt.array(b), jt.array(s), jt.array(l), thr)
            out["mlnms_%s_keep_%d" % (tag, int(thr * 10))] = N(k).astype(np.int32)
        bl = bb.BoxList(jt.array(b), (synth.IMG_W, synth.IMG_H), mode="xyxy")
        bl.add_field("scores", jt.array(s))
        bl.add_field("labels", jt.array(l))
        r = ops.boxlist_nms(bl, 0.5, max_proposals=40)
        out["blnms_%s_boxes" % tag], out["blnms_%s_scores" % tag] = N(r.bbox), N(r.get_field("scores"))
        r = ops.boxlist_ml_nms(bl, 0.5, max_proposals=40)
        out["blmlnms_%s_boxes" % tag], out["blmlnms_%s_scores" % tag] = N(r.bbox), N(r.get_field("scores"))
        out["blmlnms_%s_labels" % tag] = N(r.get_field("labels"))

    # ---- paste_mask_in_image through Masker.forward_single_image
    for tag, (im_w, im_h, n, seed) in {"a": (200, 150, 10, 1), "b": (333, 127, 10, 2)}.items():
        masks, boxes = synth.paste_inputs(n, seed, img_w=im_w, img_h=im_h)
        bl = bb.BoxList(jt.array(boxes), (im_w, im_h), mode="xyxy")
        for thr in (0.5, 0.3):
            res = mi.Masker(threshold=thr, padding=1).forward_single_image(jt.array(masks[:, None]), bl)
            out["paste_%s_out_%d" % (tag, int(thr * 10))] = np.packbits(N(res)[:, 0].astype(np.uint8), axis=None)
        out["paste_%s_masks" % tag], out["paste_%s_boxes" % tag] = masks, boxes
        out["paste_%s_hw" % tag] = np.array([im_h, im_w])
    eb = mi.expand_boxes(jt.array(boxes), 30.0 / 28.0)
    out["expand_boxes_in"], out["expand_boxes_out"] = boxes, N(eb)
    np.savez_compressed(os.path.join(HERE, "jt_detectron.npz"), **out)
    print("jt_detectron.npz:", sorted(out.keys())[:8], "...")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    make_yolact()
    make_detectron()
