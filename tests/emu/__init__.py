"""Host emulation of the tiled ROIAlign kernels (test infrastructure): builds
tests/emu/roi_tile_emu.cpp with g++ and exposes numpy wrappers."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libroi_tile_emu.so")
SRC = os.path.join(HERE, "roi_tile_emu.cpp")
HDR = os.path.join(HERE, "..", "..", "instancesegmentation-jittor_b200", "csrc", "roi_tile.h")

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int)
_lib = None


def lib():
    global _lib
    if _lib is None:
        os.makedirs(os.path.dirname(SO), exist_ok=True)
        if (not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(HDR))):
            subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-shared", "-fPIC",
                            "-fvisibility=hidden", "-o", SO, SRC], check=True)
        _lib = ctypes.CDLL(SO)
    return _lib


def _levels_args(arrs, scales):
    L = len(arrs)
    ptr = (_f32p * L)(*[a.ctypes.data_as(_f32p) for a in arrs])
    Hs = (ctypes.c_int * L)(*[a.shape[2] for a in arrs])
    Ws = (ctypes.c_int * L)(*[a.shape[3] for a in arrs])
    sc = (ctypes.c_float * L)(*[float(s) for s in scales])
    return L, ptr, Hs, Ws, sc


def pooler_forward(feats, scales, rois, levels, output_size, sampling_ratio, vs_floats=12288,
                   c_per_cta=None, U=2):
    feats = [np.ascontiguousarray(f, np.float32) for f in feats]
    rois = np.ascontiguousarray(rois, np.float32)
    L, ptr, Hs, Ws, sc = _levels_args(feats, scales)
    N, C = feats[0].shape[:2]
    K = rois.shape[0]
    PH, PW = output_size
    out = np.full((K, C, PH, PW), 7e30, np.float32)
    lv = None
    if levels is not None:
        levels = np.ascontiguousarray(levels, np.int32)
        lv = levels.ctypes.data_as(_i32p)
    nd = lib().emu_pooler_forward(ptr, Hs, Ws, sc, L, N, C, rois.ctypes.data_as(_f32p), lv, K, PH,
                                  PW, ctypes.c_float(sampling_ratio), out.ctypes.data_as(_f32p),
                                  int(vs_floats), int(c_per_cta or C), int(U))
    return out, nd


def band_forward(feats, scales, rois, levels, output_size, sampling_ratio, smem_floats=10240,
                 c_per_cta=None, nthreads=128):
    """The banded (v4) forward kernel emulated on the host; returns (out, number of
    (RoI, channel slice) pairs that had no band form)."""
    feats = [np.ascontiguousarray(f, np.float32) for f in feats]
    rois = np.ascontiguousarray(rois, np.float32)
    L, ptr, Hs, Ws, sc = _levels_args(feats, scales)
    N, C = feats[0].shape[:2]
    K = rois.shape[0]
    PH, PW = output_size
    out = np.full((K, C, PH, PW), 7e30, np.float32)
    lv = None
    if levels is not None:
        levels = np.ascontiguousarray(levels, np.int32)
        lv = levels.ctypes.data_as(_i32p)
    nf = lib().emu_band_forward(ptr, Hs, Ws, sc, L, N, C, rois.ctypes.data_as(_f32p), lv, K, PH, PW,
                                ctypes.c_float(sampling_ratio), out.ctypes.data_as(_f32p),
                                int(smem_floats), int(c_per_cta or C), int(nthreads))
    return out, nf


def pooler_backward(dout, feat_shapes, scales, rois, levels, sampling_ratio, vs_floats=8192,
                    c_per_cta=None):
    dout = np.ascontiguousarray(dout, np.float32)
    rois = np.ascontiguousarray(rois, np.float32)
    grads = [np.zeros(s, np.float32) for s in feat_shapes]
    L, ptr, Hs, Ws, sc = _levels_args(grads, scales)
    N, C = feat_shapes[0][:2]
    K, _, PH, PW = dout.shape
    lv = None
    if levels is not None:
        levels = np.ascontiguousarray(levels, np.int32)
        lv = levels.ctypes.data_as(_i32p)
    nd = lib().emu_pooler_backward(dout.ctypes.data_as(_f32p), rois.ctypes.data_as(_f32p), lv, K,
                                   Hs, Ws, sc, L, N, C, PH, PW, ctypes.c_float(sampling_ratio),
                                   ptr, int(vs_floats), int(c_per_cta or C))
    return grads, nd
