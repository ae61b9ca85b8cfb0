"""Generate the committed golden fixtures under tests/golden/.

Run in the BUILD container (needs /root/reference); the GPU box only reads the .npz files.

  nms_kat.npz        -- the reference's own known-answer vectors, parsed out of
                        detectron.jittor/tests/test_nms.py:11-58 (5 boxes, 5 thresholds) and
                        :60-217 (53 boxes, thr 0.5 -> 26 sorted keep indices).
  roi_align_ref.npz  -- inputs + outputs of the reference's RoIAlign forward/backward kernel
                        text (detectron.jittor/detectron/layers/roi_align.py:116-169,190-274)
                        compiled as C++ by oracle/build.py (oracle/_ref/libroialign_ref.so).
  cython_nms_ref.npz -- inputs + keeps of the reference's Yolact.jittor/utils/cython_nms.pyx
                        compiled here (oracle/_ref/cython_nms_ref*.so).
  box_coder_kat.npz  -- the reference's known-answer vectors for BoxCoder.decode, parsed out of
                        detectron.jittor/tests/test_box_coder.py:11-108 (5 boxes, 5 deltas,
                        5 expected boxes, assert_allclose atol 1e-4).
"""
import ast
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def _np_array_literals(path):
    """All `np.array([...])` list literals in a python file, in source order."""
    tree = ast.parse(open(path).read())
    out = []
    for node in ast.walk(tree):
        if (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute)
                and node.func.attr == "array" and node.args
                and isinstance(node.args[0], ast.List)):
            try:
                out.append((node.lineno, ast.literal_eval(node.args[0])))
            except ValueError:
                pass
    out.sort(key=lambda t: t[0])
    return [v for _, v in out]


def make_nms_kat():
    lits = _np_array_literals(os.path.join(REF, "detectron.jittor/tests/test_nms.py"))
    # source order: inputs(25 numbers), [np.array(gt_index) is not a literal], boxes 53x4,
    # scores 53, gt_indices 26
    flat = [np.asarray(l) for l in lits]
    inputs = next(a for a in flat if a.size == 25).astype(np.float32).reshape(-1, 5)
    boxes53 = next(a for a in flat if a.ndim == 2 and a.shape == (53, 4)).astype(np.float32)
    scores53 = next(a for a in flat if a.ndim == 1 and a.size == 53).astype(np.float32)
    gt53 = next(a for a in flat if a.ndim == 1 and a.size == 26).astype(np.int64)
    np.savez(os.path.join(HERE, "nms_kat.npz"),
             boxes5=inputs[:, :4], scores5=inputs[:, 4],
             thresholds5=np.array([0.1, 0.3, 0.5, 0.8, 0.9]),
             keep5=np.array([[1, 3, -1, -1, -1], [1, 3, -1, -1, -1], [1, 3, -1, -1, -1],
                             [1, 2, 3, 4, -1], [0, 1, 2, 3, 4]], np.int64),
             boxes53=boxes53, scores53=scores53, keep53=gt53)


def make_box_coder_kat():
    lits = _np_array_literals(os.path.join(REF, "detectron.jittor/tests/test_box_coder.py"))
    assert len(lits) == 3 and all(len(l) == 20 for l in lits)
    bbox, deltas, gt = (np.asarray(l, np.float32).reshape(-1, 4) for l in lits)  # source order
    np.savez(os.path.join(HERE, "box_coder_kat.npz"), bbox=bbox, deltas=deltas, gt_bbox=gt,
             atol=np.float32(1e-4))


def make_roi_align_ref():
    import oracle
    rng = np.random.default_rng(20261104)
    feat = rng.standard_normal((2, 6, 25, 42)).astype(np.float32)
    rois = np.array([
        [0, 10, 20, 300, 400],
        [1, 0, 0, 1343, 799],      # full image
        [0, 33.3, 47.7, 101.9, 222.2],
        [1, 50, 50, 50, 50],       # zero size
        [0, -40, -30, 100, 90],    # negative origin
        [1, 1200, 700, 1500, 900],  # past the edge
        [0, 5, 5, 6, 6],           # 1 px
    ], np.float32)
    cases = {}
    for tag, P, scale, sr in [("p7_s2", (7, 7), 1 / 32, 2), ("p14_s2", (14, 14), 1 / 32, 2),
                              ("p14_s0", (14, 14), 1 / 32, 0), ("p3x5_s3", (3, 5), 1 / 32, 3)]:
        out = oracle.ref_roi_align_forward(feat, rois, P, scale, sr)
        dout = rng.standard_normal(out.shape).astype(np.float32)
        gin = oracle.ref_roi_align_backward(dout, rois, feat.shape, scale, sr)
        cases[tag + "_cfg"] = np.array([P[0], P[1], scale, sr], np.float64)
        cases[tag + "_out"] = out
        cases[tag + "_dout"] = dout
        cases[tag + "_gin"] = gin
    np.savez_compressed(os.path.join(HERE, "roi_align_ref.npz"), feat=feat, rois=rois, **cases)


def make_cython_nms_ref():
    import oracle
    cnms = oracle.ref_cython_nms()
    rng = np.random.default_rng(7)
    n = 300
    ctr = rng.uniform(0, 500, (n, 2))
    wh = rng.uniform(10, 120, (n, 2))
    dets = np.concatenate([ctr - wh / 2, ctr + wh / 2, rng.permutation(n)[:, None] / n + 0.001],
                          1).astype(np.float32)
    out = {"dets": dets}
    for thr in (0.3, 0.5, 0.7):
        out["keep_%d" % int(thr * 10)] = np.asarray(cnms(dets, np.float32(thr)), np.int64)
    kat = np.load(os.path.join(HERE, "nms_kat.npz"))
    d5 = np.concatenate([kat["boxes5"], kat["scores5"][:, None]], 1).astype(np.float32)
    out["dets5"] = d5
    out["keep5_5"] = np.asarray(cnms(d5, np.float32(0.5)), np.int64)
    np.savez_compressed(os.path.join(HERE, "cython_nms_ref.npz"), **out)


if __name__ == "__main__":
    make_nms_kat()
    make_box_coder_kat()
    make_roi_align_ref()
    make_cython_nms_ref()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
