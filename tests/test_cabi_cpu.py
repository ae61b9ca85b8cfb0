"""CPU checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads without a
GPU, exports exactly the symbols include/roipath.h declares, and the host layer refuses to
run without CUDA tensors (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

import roipath_b200 as rp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "roipath.h")


def _declared():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"^RP_API\s+\w+\s+(rp_\w+)\s*\(", text, flags=re.M)))


def test_header_declares_the_whole_path():
    names = _declared()
    for must in ("rp_roi_align_forward", "rp_roi_align_backward", "rp_pooler_forward",
                 "rp_pooler_backward", "rp_level_map", "rp_nms_batched", "rp_fast_nms_batched",
                 "rp_mask_assemble", "rp_paste_masks", "rp_mask_select_sigmoid", "rp_yolact_decode",
                 "rp_box_decode", "rp_rpn_pre_nms", "rp_rpn_pre_nms_workspace_bytes",
                 "rp_rpn_pre_nms_levels", "rp_rpn_select_levels", "rp_box_post",
                 "rp_yolact_prefilter", "rp_boxlist_iou", "rp_roi_pool_forward", "rp_roi_pool_backward",
                 "rp_mask_prob", "rp_box_masks"):
        assert must in names


def test_library_exports_every_declared_symbol():
    L = rp.lib()
    path = rp.package._lib.LIB_PATH
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (rp_\w+)", out))
    assert exported == set(_declared())
    assert set(rp.package._lib.SIGNATURES) == set(_declared())
    assert L.rp_version() == 1
    assert L.rp_built_for_sm() == 100


def test_library_holds_sm100a_code_and_async_copies():
    path = rp.package._lib.LIB_PATH
    out = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    assert "LDGSTS" in sass  # cp.async global->shared tile staging in the ROIAlign forward kernel


def test_argument_errors_need_no_gpu():
    L = rp.lib()
    # null pointers / bad sizes are rejected before any CUDA call
    assert L.rp_roi_align_forward(None, 1, 1, 4, 4, None, 3, 7, 7, 1.0, 2.0, None, 0, None) == -1
    assert L.rp_nms_batched(None, None, None, None, 0, 0, 0, 0.5, -1, 0, None, None, None, 0, None) == -1
    assert L.rp_mask_assemble(None, None, None, None, 1, 5, 138, 138, 32, 550, 550, 1, None, None,
                              None, None) == -1
    assert L.rp_paste_masks(None, None, 3, 28, 800, 1333, 0.5, 1, None, None) == -1
    # banded NMS: the workspace is O(n), no n x n/64 bit matrix (400 MB at 8 x 20k boxes)
    ws = L.rp_nms_workspace_bytes(160000, 20000, 8)
    assert 160000 * 40 < ws < 64 * 1024 * 1024
    assert L.rp_fast_nms_workspace_bytes(500, 500, 1, 80, 200) > 0
    assert L.rp_box_decode(None, None, 3, 1, 1.0, 1.0, 1.0, 1.0, 4.0, None, None) == -1
    assert L.rp_box_decode(None, None, 0, 1, 1.0, 1.0, 1.0, 1.0, 4.0, None, None) == 0
    assert L.rp_rpn_pre_nms(None, None, None, None, 1, 3, 4, 4, 10, 0.0, 1.0, 1.0, 1.0, 1.0, 4.0,
                            None, None, None, None, None, 0, None) == -1
    # RPN proposal stage: O(N*A*H*W) sort buffers + O(N*k) slots
    ws = L.rp_rpn_pre_nms_workspace_bytes(2, 3, 200, 336, 2000)
    assert 2 * 3 * 200 * 336 * 24 < ws < 32 * 1024 * 1024
    assert L.rp_rpn_pre_nms_workspace_bytes(0, 3, 4, 4, 10) == 0
    A3 = (ctypes.c_int * 2)(3, 3)
    H2 = (ctypes.c_int * 2)(50, 25)
    W2 = (ctypes.c_int * 2)(84, 42)
    assert L.rp_rpn_pre_nms_levels_workspace_bytes(2, 4, A3, H2, W2, 1000) > 4 * 3 * (50 * 84 + 25 * 42) * 24
    assert L.rp_rpn_pre_nms_levels_workspace_bytes(9, 4, A3, H2, W2, 1000) == 0  # > RP_MAX_LEVELS
    assert L.rp_rpn_select_levels(None, None, None, None, None, 5, 8, 1000, 100, None, None, None,
                                  None, 0, None) == -1
    assert L.rp_yolact_prefilter_workspace_bytes(64, 19248, 65536) >= 64 * 19248 + 4 * 65536
    assert L.rp_yolact_prefilter(None, None, None, None, 64, 19248, 81, 32, 0.05, 65536, None, None,
                                 None, None, None, None, None, 0, None) == -1
    assert L.rp_box_masks(None, 3, 10, 10, 0.0, 0, None, None, None, None) == -1
    assert L.rp_boxlist_iou(None, 5, None, 5, None, None) == -1
    assert L.rp_boxlist_iou(None, 0, None, 5, None, None) == 0
    assert L.rp_roi_pool_forward(None, 1, 1, 4, 4, None, 3, 7, 7, 1.0, None, None, None) == -1
    assert L.rp_mask_prob(None, None, None, None, None, 100, 3, 200, 10, None, None) == -3  # dim > 128
    assert L.rp_box_post_workspace_bytes(8000, 81, 8, 1000, 8000 * 80) > 8000 * 80 * 28
    assert L.rp_box_post(None, None, None, None, None, 10, 81, 1, 10, 324, 0, 0.05, 0.5, 100, 10.0,
                         10.0, 5.0, 5.0, 4.0, 800, None, None, None, None, None, None, 200, None,
                         None, 0, None) == -1


def test_no_cpu_fallback():
    x = torch.zeros(1, 4, 8, 8)
    rois = torch.tensor([[0, 0, 0, 4, 4.0]])
    with pytest.raises(rp.RoiPathError):
        rp.roi_align(x, rois, (7, 7), 1.0, 2)
    with pytest.raises(rp.RoiPathError):
        rp.nms(torch.zeros(3, 4), torch.zeros(3), 0.5)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "instancesegmentation-jittor_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "libroipath_oracle" not in text and "roipath_oracle.c" not in text, f
