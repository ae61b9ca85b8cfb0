"""Pose2Seg paste-back on the GPU.

Reference: Pose2Seg.jittor/modeling/build_model.py:31-43 `warpAffine(predmap, H_e2e, width, height)`
-- cv2.warpAffine with WARP_INVERSE_MAP + INTER_LINEAR on the host, channel 1, > 0.5 -> uint8 --
called once per instance after a device->host copy of every instance's 64x64x2 softmax map.
Here all instances of an image are pasted by one launch and the maps never leave the device.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def affine_paste(pred, H_e2e, width, height, channel=1, out=None):
    """pred [n,h,w,C] float32 CUDA (HWC softmax maps); H_e2e [n,2,3] or [n,3,3] (output pixel ->
    source map, the matrix the reference passes with WARP_INVERSE_MAP; numpy or tensor, float64);
    returns uint8 [n,height,width] masks (value of `channel` > 0.5), bit-identical to cv2."""
    pred = _lib.as_f32(pred, "pred")
    if pred.ndim != 4:
        raise ValueError("pred must be [n,h,w,C]")
    n, h, w, C = (int(v) for v in pred.shape)
    M = torch.as_tensor(np.asarray(H_e2e.cpu() if isinstance(H_e2e, torch.Tensor) else H_e2e, np.float64))
    M = M.reshape(n, -1, 3)[:, :2, :].contiguous().to(pred.device)
    if out is None:
        out = torch.empty((n, int(height), int(width)), dtype=torch.uint8, device=pred.device)
    rc = _lib.lib().rp_affine_paste(_lib.dptr(pred, torch.float32), ctypes.c_void_p(M.data_ptr()), n, h, w, C,
                                    int(channel), int(width), int(height),
                                    ctypes.c_void_p(out.data_ptr()), _lib.stream_ptr(pred.device))
    _lib.check(rc, "rp_affine_paste")
    return out


def warpAffine(predmap, H_e2e, width, height):
    """Drop-in for build_model.py:31-43 (one instance): predmap [h,w,2] CUDA tensor -> uint8 [height,width]."""
    return affine_paste(predmap[None], np.asarray(H_e2e, np.float64)[None, 0:2], width, height, channel=1)[0]
