"""rp_nms_batched_static: the NMS entry point that allocates nothing (its workspace is a fixed array in
device memory, include/roipath.h).  Called through the C ABI (ctypes), checked against the existing entry
point rp_nms_batched on the same inputs -- keep indices and counts bit-exact, the bar of every NMS test --
on two calls in a row (the array keeps the first call's contents), and against the oracle.

This file sorts last on purpose (pytest collects files by name)."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
import roipath_b200 as rp
from tests import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
_lib = rp.package._lib


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _problem(lens, seed, tie_heavy):
    parts = [synth.nms_boxes(max(n, 1), seed + i, tie_heavy=tie_heavy) for i, n in enumerate(lens)]
    bs = [p[0][:n] for p, n in zip(parts, lens)]
    ss = [p[1][:n] for p, n in zip(parts, lens)]
    ls = [(p[2][:n] % 3).astype(np.float32) for p, n in zip(parts, lens)]
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    return bs, ss, ls, offs


def _call_existing(L, boxes, scores, labels, offsets, B, n_total, n_max, thr, max_keep, flags):
    keep = torch.full((max(n_total, 1),), -7, dtype=torch.int32, device=DEV)
    n_keep = torch.full((B,), -7, dtype=torch.int32, device=DEV)
    nbytes = L.rp_nms_workspace_bytes(n_total, n_max, B)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=DEV)
    rc = L.rp_nms_batched(_lib.dptr(boxes), _lib.dptr(scores), _lib.dptr(labels), _lib.dptr(offsets), B,
                          n_total, n_max, float(thr), max_keep, flags, _lib.dptr(keep), _lib.dptr(n_keep),
                          ctypes.c_void_p(ws.data_ptr()), nbytes, _lib.stream_ptr(torch.device(DEV)))
    assert rc == 0
    torch.cuda.synchronize()
    return keep.cpu().numpy(), n_keep.cpu().numpy()


def _call_static(L, boxes, scores, labels, offsets, B, n_total, n_max, thr, max_keep, flags, keep, n_keep):
    # outputs are allocated by the caller BEFORE the call; the call itself allocates nothing
    keep.fill_(-7)
    n_keep.fill_(-7)
    rc = L.rp_nms_batched_static(_lib.dptr(boxes), _lib.dptr(scores), _lib.dptr(labels), _lib.dptr(offsets),
                                 B, n_total, n_max, float(thr), max_keep, flags, _lib.dptr(keep),
                                 _lib.dptr(n_keep), _lib.stream_ptr(torch.device(DEV)))
    assert rc == 0
    torch.cuda.synchronize()
    return keep.cpu().numpy(), n_keep.cpu().numpy()


def test_static_workspace_entry_matches_existing_entry_twice_in_a_row():
    L = rp.lib()
    cap = L.rp_nms_static_capacity_bytes()
    assert cap == 4 << 20
    # call 1: one image of 20 000 boxes (20 bands, the dataflow kernel), no cap
    # call 2: a ragged batch with labels and a keep cap (different layout of the same array)
    # call 3: call 1 again (the array now holds call 2's state)
    cases = [([20000], 7, False, 0.7, -1, False, 0),
             ([1500, 0, 700, 64, 2500, 1], 40, True, 0.5, 200, True, 0),
             ([20000], 7, False, 0.7, -1, False, 0),
             ([3000], 3, True, 0.5, -1, False, _lib.RP_NMS_CYTHON)]
    for lens, seed, ties, thr, mk, use_labels, flags in cases:
        bs, ss, ls, offs = _problem(lens, seed, ties)
        B, n_total, n_max = len(lens), int(offs[-1]), int(max(lens))
        assert L.rp_nms_workspace_bytes(n_total, n_max, B) <= cap
        boxes, scores, offsets = T(np.concatenate(bs)), T(np.concatenate(ss)), T(offs)
        labels = T(np.concatenate(ls)) if use_labels else None
        keep = torch.empty((n_total,), dtype=torch.int32, device=DEV)
        n_keep = torch.empty((B,), dtype=torch.int32, device=DEV)
        ref_keep, ref_n = _call_existing(L, boxes, scores, labels, offsets, B, n_total, n_max, thr, mk, flags)
        for rep in range(2):  # two calls in a row on the shared array
            got_keep, got_n = _call_static(L, boxes, scores, labels, offsets, B, n_total, n_max, thr, mk,
                                           flags, keep, n_keep)
            np.testing.assert_array_equal(got_n, ref_n)
            for i in range(B):
                np.testing.assert_array_equal(got_keep[offs[i]: offs[i] + got_n[i]],
                                              ref_keep[offs[i]: offs[i] + ref_n[i]])
        if flags == 0:  # and the oracle itself, for the first and the last problem
            for i in (0, B - 1):
                if lens[i]:
                    o = oracle.nms(bs[i], ss[i], thr, labels=ls[i] if use_labels else None, max_keep=mk)
                    np.testing.assert_array_equal(got_keep[offs[i]: offs[i] + got_n[i]], o)


def test_static_workspace_entry_rejects_sizes_beyond_its_array():
    L = rp.lib()
    cap = L.rp_nms_static_capacity_bytes()
    n, B = 20000, 8  # 9.9 MiB of workspace
    assert L.rp_nms_workspace_bytes(n * B, n, B) > cap
    boxes = torch.zeros((n * B, 4), device=DEV)
    scores = torch.zeros((n * B,), device=DEV)
    offsets = T((np.arange(B + 1) * n).astype(np.int32))
    keep = torch.empty((n * B,), dtype=torch.int32, device=DEV)
    n_keep = torch.empty((B,), dtype=torch.int32, device=DEV)
    rc = L.rp_nms_batched_static(_lib.dptr(boxes), _lib.dptr(scores), None, _lib.dptr(offsets), B, n * B, n,
                                 0.7, -1, 0, _lib.dptr(keep), _lib.dptr(n_keep),
                                 _lib.stream_ptr(torch.device(DEV)))
    assert rc == -3  # RP_E_LIMIT: size exceeds a documented limit
    with pytest.raises(ValueError, match="limit"):
        _lib.check(rc, "rp_nms_batched_static")
    # bad arguments are still RP_E_ARG
    assert L.rp_nms_batched_static(None, None, None, None, 0, 0, 0, 0.5, -1, 0, None, None, None) == -1
