"""jittor.nn stand-in (see jittor/__init__.py): only what the reference's post-processing text uses."""
import torch
import torch.nn.functional as F

from . import Module, _t, _v  # noqa: F401


def interpolate(X, size=None, scale_factor=None, mode="bilinear", align_corners=False):
    x = _t(X)
    if size is not None:
        size = tuple(int(s) for s in size)
    if mode == "nearest":
        return _v(F.interpolate(x, size=size, scale_factor=scale_factor, mode=mode))
    return _v(F.interpolate(x, size=size, scale_factor=scale_factor, mode=mode,
                            align_corners=align_corners))


def softmax(x, dim=None):
    return _v(torch.softmax(_t(x), dim=dim))


def relu(x):
    return _v(torch.relu(_t(x)))
