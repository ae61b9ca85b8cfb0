"""A minimal torch-backed stand-in for the `jittor` package -- TEST INFRASTRUCTURE ONLY.

Used only by tests/golden/make_golden_jt.py (in the build container) to EXECUTE the reference's own
Python text (Yolact.jittor/layers/box_utils.py, layers/functions/detection.py,
layers/output_utils.py; detectron.jittor/detectron/modeling/poolers.py, layers/nms.py,
structures/bounding_box.py, structures/boxlist_ops.py,
modeling/roi_heads/mask_head/inference.py, ...) and record its outputs as golden fixtures for the
C oracle.  Nothing under instancesegmentation-jittor_b200/ imports it.

Jittor itself (version unpinned by the reference, detectron.jittor/setup.py:7) is not installable
in this container.  The reference files run here are line-for-line ports of dbolya/yolact and
facebookresearch/maskrcnn-benchmark, whose semantics are torch's; detectron.jittor/tests/test_diff.py
compares against torch outputs.  The op semantics this stand-in ASSUMES (and that the fixtures
therefore pin the oracle to):
  * argsort(descending=True) is stable (equal keys keep ascending index order) and returns
    (indices, sorted values); argmax returns (indices, values); Var.max(dim) returns values only;
  * int32() truncates toward zero; clamp(min_v, max_v); minimum/maximum propagate like torch;
  * nn.interpolate(mode='bilinear', align_corners=False) == torch.nn.functional.interpolate
    (half-pixel centres, source index clamped at 0, neighbour index clamped at size-1);
  * jt.sigmoid == torch.sigmoid, matmul == torch.matmul (CPU fp32);
  * jt.candidate(x, fail_cond): the documented greedy loop -- row i is selected iff no previously
    selected row j makes fail_cond(i, j) true; the condition string is evaluated in float32 with
    C promotion rules for the threshold literal (double);
  * jt.nms(dets, thr): sort by score descending (stable), then candidate() with the legacy "+1"
    IoU > thr test (the same condition detectron/layers/nms.py:27-33 spells out for _ml_nms).
"""
import re
import types

import numpy as np
import torch

float32 = "float32"
int32 = "int32"
flags = types.SimpleNamespace(use_cuda=0)


class Var(torch.Tensor):
    # ---- dtype casts (jittor spells them as methods) ----
    def int32(self):
        return _v(torch.Tensor.to(self, torch.int32))

    def int64(self):
        return _v(torch.Tensor.to(self, torch.int64))

    def float32(self):
        return _v(torch.Tensor.to(self, torch.float32))

    def float64(self):
        return _v(torch.Tensor.to(self, torch.float64))

    def uint8(self):
        return _v(torch.Tensor.to(self, torch.uint8))

    def numpy(self):
        return self.as_subclass(torch.Tensor).detach().cpu().numpy()

    # ---- reductions with jittor's return conventions ----
    def max(self, dim=None, keepdims=False):
        t = self.as_subclass(torch.Tensor)
        if dim is None:
            return _v(t.max())
        return _v(t.max(dim=dim, keepdim=keepdims).values)

    def min(self, dim=None, keepdims=False):
        t = self.as_subclass(torch.Tensor)
        if dim is None:
            return _v(t.min())
        return _v(t.min(dim=dim, keepdim=keepdims).values)

    def argmax(self, dim, keepdims=False):
        t = self.as_subclass(torch.Tensor)
        # first maximum, like a sequential scan
        vals = t.max(dim=dim, keepdim=True).values
        first = (t == vals).to(torch.int64).argmax(dim=dim, keepdim=keepdims)
        v = vals if keepdims else vals.squeeze(dim)
        return _v(first), _v(v)

    def argsort(self, dim=-1, descending=False):
        t = self.as_subclass(torch.Tensor)
        vals, idx = torch.sort(t, dim=dim, descending=descending, stable=True)
        return _v(idx), _v(vals)

    def sort(self, dim=-1, descending=False):
        t = self.as_subclass(torch.Tensor)
        vals, idx = torch.sort(t, dim=dim, descending=descending, stable=True)
        return _v(vals), _v(idx)

    def sum(self, dim=None, keepdims=False):
        t = self.as_subclass(torch.Tensor)
        return _v(t.sum() if dim is None else t.sum(dim=dim, keepdim=keepdims))

    def prod(self, dim=None):
        t = self.as_subclass(torch.Tensor)
        return _v(t.prod() if dim is None else t.prod(dim=dim))

    def expand(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        return _v(self.as_subclass(torch.Tensor).expand(*shape))

    def view(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        return _v(self.as_subclass(torch.Tensor).reshape(*shape))

    def sync(self):
        return self

    @property
    def ndim(self):
        return self.dim()


def _v(t):
    if isinstance(t, Var):
        return t
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(t)
    return t.as_subclass(Var)


def array(x, dtype=None):
    if isinstance(x, torch.Tensor):
        return _v(x.clone())
    a = np.asarray(x)
    if a.dtype == np.float64 and dtype is None:
        a = a.astype(np.float32)  # jittor keeps python floats as float32
    t = torch.from_numpy(np.ascontiguousarray(a))
    return _v(t)


def _dt(dtype):
    if dtype is None:
        return torch.float32
    if isinstance(dtype, torch.dtype):
        return dtype
    return {"float32": torch.float32, "float": torch.float32, "float64": torch.float64,
            "int32": torch.int32, "int": torch.int32, "int64": torch.int64, "bool": torch.bool,
            "uint8": torch.uint8}[str(dtype).replace("torch.", "")]


def _shape(args):
    if len(args) == 1 and isinstance(args[0], (tuple, list, torch.Size)):
        return tuple(int(v) for v in args[0])
    return tuple(int(v) for v in args)


def zeros(*shape, dtype=None):
    return _v(torch.zeros(_shape(shape), dtype=_dt(dtype)))


def ones(*shape, dtype=None):
    return _v(torch.ones(_shape(shape), dtype=_dt(dtype)))


def zeros_like(x):
    return _v(torch.zeros_like(x.as_subclass(torch.Tensor)))


def arange(*args, dtype=None, device=None):
    t = torch.arange(*args)
    if dtype is not None:
        t = t.to(_dt(dtype))
    elif t.dtype == torch.int64:
        t = t.to(torch.int32)  # jittor's default integer type
    return _v(t)


def _un(fn):
    return lambda x: _v(fn(x.as_subclass(torch.Tensor)))


sqrt, log2, log, exp, floor, ceil, sigmoid, abs = (_un(f) for f in (
    torch.sqrt, torch.log2, torch.log, torch.exp, torch.floor, torch.ceil, torch.sigmoid, torch.abs))


def _t(x):
    return x.as_subclass(torch.Tensor) if isinstance(x, torch.Tensor) else x


def minimum(a, b):
    return _v(torch.minimum(_t(a), _t(b)))


def maximum(a, b):
    return _v(torch.maximum(_t(a), _t(b)))


def clamp(x, min_v=None, max_v=None):
    return _v(torch.clamp(_t(x), min=min_v, max=max_v))


def matmul(a, b):
    return _v(torch.matmul(_t(a), _t(b)))


def max(x, dim=None, keepdims=False):  # noqa: A001  (jittor's name)
    return _v(x).max(dim, keepdims)


def sum(x, dim=None, keepdims=False):  # noqa: A001
    return _v(x).sum(dim, keepdims)


def argsort(x, dim=-1, descending=False):
    return _v(x).argsort(dim, descending)


def stack(xs, dim=0):
    return _v(torch.stack([_t(x) for x in xs], dim=dim))


def concat(xs, dim=0):
    return _v(torch.cat([_t(x) for x in xs], dim=dim))


def where(c, a, b):
    return _v(torch.where(_t(c), _t(a), _t(b)))


def no_grad():
    return torch.no_grad()


contrib = types.SimpleNamespace(concat=concat)


def candidate(x, fail_cond):
    """row i is selected iff no previously selected row j satisfies fail_cond(i, j)."""
    a = _t(x).detach().cpu().numpy()
    expr = fail_cond.replace("(Tx)0", "ZERO").replace("&&", " and ").replace("||", " or ")
    expr = re.sub(r"@x\((\w+),(\d+)\)", r"X[\1,\2]", expr)
    code = compile(expr, "<fail_cond>", "eval")
    env = {"X": a, "ZERO": a.dtype.type(0), "max": lambda p, q: p if p > q else q,
           "min": lambda p, q: p if p < q else q}
    sel = []
    with np.errstate(all="ignore"):
        for i in range(a.shape[0]):
            ok = True
            for j in sel:
                env["i"], env["j"] = i, j
                if eval(code, env):
                    ok = False
                    break
            if ok:
                sel.append(i)
    return _v(torch.tensor(sel, dtype=torch.int64))


def nms(dets, thresh):
    """jittor.nms: dets [n,5] = x1,y1,x2,y2,score -> kept indices in descending-score order."""
    d = _v(dets)
    order, _ = d[:, 4].argsort(0, descending=True)
    d = d[order]
    s1 = "(@x(j,2)-@x(j,0)+1)*(@x(j,3)-@x(j,1)+1)"
    s2 = "(@x(i,2)-@x(i,0)+1)*(@x(i,3)-@x(i,1)+1)"
    iw = "max((Tx)0,min(@x(j,2),@x(i,2))-max(@x(j,0),@x(i,0))+1)"
    ih = "max((Tx)0,min(@x(j,3),@x(i,3))-max(@x(j,1),@x(i,1))+1)"
    inter = ih + "*" + iw
    iou = inter + "/(" + s1 + "+" + s2 + "-" + inter + ")"
    sel = candidate(d, iou + ">" + str(thresh))
    return order[sel]


class Module(object):
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return self.execute(*a, **k)


from . import nn, init  # noqa: E402,F401
