"""ctypes binding of libroipath_b200.so (the C ABI declared in include/roipath.h).

The product path has NO fallback: if the CUDA library is missing and cannot be built, or a
tensor is not a CUDA float32/int32 tensor, the call raises.  Nothing here imports oracle/.
"""
import ctypes
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libroipath_b200.so")

RP_MAX_LEVELS = 8
RP_NMS_GE = 1
RP_NMS_ASCENDING = 2
RP_NMS_TIES_HIGH_FIRST = 4
RP_NMS_CYTHON = RP_NMS_GE | RP_NMS_ASCENDING | RP_NMS_TIES_HIGH_FIRST
RP_NMS_PRESORTED = 8

_c = ctypes
_vp, _i, _f, _d, _sz = _c.c_void_p, _c.c_int, _c.c_float, _c.c_double, _c.c_size_t

# name -> (restype, argtypes); mirrors include/roipath.h one to one
SIGNATURES = {
    "rp_version": (_i, []),
    "rp_built_for_sm": (_i, []),
    "rp_debug_rx_timeouts": (_i, [_vp, _i]),
    "rp_reload_env": (None, []),
    "rp_roi_align_forward": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _f, _f, _vp, _i, _vp]),
    "rp_roi_align_backward": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _f, _vp, _i, _vp]),
    "rp_roi_align_forward_nhwc": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _f, _f, _vp, _vp]),
    "rp_pooler_forward_nhwc": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _f, _vp, _vp]),
    "rp_level_map": (_i, [_vp, _i, _i, _i, _f, _i, _f, _vp, _vp]),
    "rp_level_map_area": (_i, [_vp, _vp, _i, _i, _i, _f, _vp, _vp]),
    "rp_pooler_forward": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _f, _vp, _i, _vp]),
    "rp_pooler_backward": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp, _i, _vp]),
    "rp_nms_workspace_bytes": (_sz, [_i, _i, _i]),
    "rp_nms_batched": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _d, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "rp_nms_static_capacity_bytes": (_sz, []),
    "rp_nms_batched_static": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _d, _i, _i, _vp, _vp, _vp]),
    "rp_yolact_decode": (_i, [_vp, _vp, _i, _vp, _vp]),
    "rp_fast_nms_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "rp_fast_nms_batched": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _i, _i, _i, _f, _i,
                                 _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "rp_mask_assemble": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "rp_mask_select_sigmoid": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "rp_paste_masks": (_i, [_vp, _vp, _i, _i, _i, _i, _f, _i, _vp, _vp]),
    "rp_lincomb_mask_loss": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "rp_affine_paste": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "rp_embed_masks": (_i, [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, _vp, _vp]),
    "rp_resize_threshold": (_i, [_vp, _i, _i, _i, _i, _i, _f, _i, _i, _vp, _vp]),
    "rp_pack_mask_bits": (_i, [_vp, _c.c_longlong, _c.c_longlong, _vp, _vp]),
    "rp_box_decode": (_i, [_vp, _vp, _i, _i, _f, _f, _f, _f, _f, _vp, _vp]),
    "rp_rpn_pre_nms_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "rp_rpn_pre_nms": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _f, _f, _f, _f,
                            _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "rp_rpn_pre_nms_levels_workspace_bytes": (_sz, [_i, _i, _vp, _vp, _vp, _i]),
    "rp_rpn_pre_nms_levels": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _f, _f, _f, _f, _f,
                                   _f, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "rp_yolact_prefilter_workspace_bytes": (_sz, [_i, _i, _i]),
    "rp_yolact_prefilter": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _vp, _vp, _vp, _vp, _vp,
                                 _vp, _vp, _sz, _vp]),
    "rp_box_post_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "rp_box_post": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _d, _i, _f, _f, _f, _f,
                         _f, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "rp_box_masks": (_i, [_vp, _i, _i, _i, _f, _i, _vp, _vp, _vp, _vp]),
    "rp_boxlist_iou": (_i, [_vp, _i, _vp, _i, _vp, _vp]),
    "rp_roi_pool_forward": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _f, _vp, _vp, _vp]),
    "rp_roi_pool_backward": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "rp_mask_prob": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "rp_rpn_select_workspace_bytes": (_sz, [_i, _i, _i]),
    "rp_rpn_select_levels": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz,
                                  _vp]),
}

_lib = None


class RoiPathError(RuntimeError):
    pass


def lib():
    """Load (building in-tree with nvcc if absent) the CUDA library.  Raises if impossible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("_roipath_build", os.path.join(HERE, "build.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build()
        except Exception as e:  # no fallback: the CUDA library IS the product
            raise RoiPathError(
                "libroipath_b200.so is missing and could not be built (%s). "
                "Run `python instancesegmentation-jittor_b200/build.py`." % (e,))
    L = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if L.rp_version() != 1:
        raise RoiPathError("libroipath_b200.so ABI version mismatch")
    if L.rp_built_for_sm() != 100:
        raise RoiPathError("libroipath_b200.so was not built for sm_100a (built for sm_%d)" % L.rp_built_for_sm())
    _lib = L
    return _lib


def reload_env():
    """Make the library re-read its RP_* knobs (they are read once per process otherwise)."""
    lib().rp_reload_env()


_ERR = {-1: "invalid argument", -2: "workspace too small", -3: "size exceeds a documented limit"}


def check(rc, what):
    if rc == 0:
        return
    if rc < 0:
        raise ValueError("%s: %s (rc=%d)" % (what, _ERR.get(rc, "error"), rc))
    raise RoiPathError("%s: CUDA error %d" % (what, rc))


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def dptr(t, dtype=None, name="tensor"):
    """Device pointer of a contiguous CUDA tensor (None -> NULL).  No silent copies."""
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise RoiPathError("%s must live on a CUDA device: this path has no CPU fallback" % name)
    if dtype is not None and t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    return ctypes.c_void_p(t.data_ptr())


def as_f32(t, name="tensor"):
    """float32 contiguous view/copy on the same CUDA device (device-side cast only)."""
    if not t.is_cuda:
        raise RoiPathError("%s must live on a CUDA device: this path has no CPU fallback" % name)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


_ws_cache = {}


def workspace(nbytes, device, tag="default"):
    """A scratch buffer per (device index, tag, current stream) that only grows.  Keyed on the
    stream so that two streams never share keys / bit words; a buffer that is replaced by a
    larger one is handed back to the allocator only after the work queued on its stream."""
    dev = torch.device(device)
    index = dev.index if dev.index is not None else torch.cuda.current_device()
    stream = torch.cuda.current_stream(index)
    key = (index, tag, stream.cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        if buf is not None:
            buf.record_stream(stream)
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=torch.device("cuda", index))
        _ws_cache[key] = buf
    return buf
