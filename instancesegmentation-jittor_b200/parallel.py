"""Image sharding across the GPUs of one box and the final gather of kept detections.

The path shards by image with no data-path collective: every op is per image
(rois[:,0] picks the image's map, det/layers/roi_align.py:130,144; NMS per image,
det/modeling/rpn/inference.py:112-124; FastNMS / assembly per batch_idx,
yol/layers/functions/detection.py:68-76; paste per image, mask_head/inference.py:201-204).
The reference itself has no collective.  The only exchange here is one variable-length gather
of the kept detections and bit-packed masks to rank 0, once per batch, outside the timed kernels.
"""
import torch
import torch.distributed as dist


def shard_range(n_images, rank, world):
    """Contiguous block of images owned by `rank`: [begin, end)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n_images, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def rebase_rois(rois, begin):
    """RoIs [K,5] with GLOBAL image ids -> ids local to a shard that starts at image `begin`."""
    out = rois.clone()
    out[:, 0] -= begin
    return out


def pack_masks(masks_u8):
    """uint8 0/1 masks [m, H, W] -> bit-packed uint8 [m, ceil(H*W/8)] (LSB first).  CUDA tensors go
    through rp_pack_mask_bits (one pass, npix bytes in, npix/8 out); the torch form below only
    serves CPU tensors (the gloo tests of the gather logic)."""
    m = masks_u8.shape[0]
    if masks_u8.is_cuda:
        import ctypes
        from . import _lib
        src = masks_u8.to(torch.uint8).contiguous()
        npix = int(src[0].numel()) if m else 0
        out = torch.empty((m, (npix + 7) // 8), dtype=torch.uint8, device=src.device)
        if m and npix:
            rc = _lib.lib().rp_pack_mask_bits(ctypes.c_void_p(src.data_ptr()), m, npix,
                                              ctypes.c_void_p(out.data_ptr()), _lib.stream_ptr(src.device))
            _lib.check(rc, "rp_pack_mask_bits")
        return out
    flat = masks_u8.reshape(m, -1).to(torch.uint8)
    pad = (-flat.shape[1]) % 8
    if pad:
        flat = torch.nn.functional.pad(flat, (0, pad))
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.int32, device=flat.device)
    return (flat.reshape(m, -1, 8).to(torch.int32) * w).sum(-1).to(torch.uint8)


def unpack_masks(packed, h, w):
    """Inverse of pack_masks."""
    m = packed.shape[0]
    sh = torch.arange(8, dtype=torch.int32, device=packed.device)
    bits = (packed.to(torch.int32).unsqueeze(-1) >> sh) & 1
    return bits.reshape(m, -1)[:, : h * w].reshape(m, h, w).to(torch.uint8)


def _all_gather_var(t, counts, group):
    """all_gather of per-rank tensors whose first dimension differs (counts[r] rows on rank r)."""
    world = len(counts)
    cap = max(max(counts), 1)
    buf = t.new_zeros((cap,) + tuple(t.shape[1:]))
    buf[: t.shape[0]] = t
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf, group=group)
    return [o[:c] for o, c in zip(outs, counts)]


def gather_detections(image_ids, classes, scores, boxes, masks=None, group=None, dst=0):
    """Gather this rank's kept detections to rank `dst`.

    image_ids [m] int64 (GLOBAL image index), classes [m] int64, scores [m] fp32, boxes [m,4],
    masks optional uint8 [m,H,W] (sent bit-packed).  Counts are exchanged first, then the
    payload (variable length).  Returns a dict on rank `dst` (rows ordered by rank, i.e. by image
    block), None elsewhere.  With no process group initialised it returns the inputs unchanged.
    """
    if not (dist.is_available() and dist.is_initialized()):
        return {"image_id": image_ids, "class": classes, "score": scores, "box": boxes, "mask": masks}
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = scores.device
    cnt = torch.tensor([scores.shape[0]], dtype=torch.int64, device=dev)
    all_cnt = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(all_cnt, cnt, group=group)
    counts = [int(c.item()) for c in all_cnt]
    parts = {
        "image_id": _all_gather_var(image_ids.to(torch.int64), counts, group),
        "class": _all_gather_var(classes.to(torch.int64), counts, group),
        "score": _all_gather_var(scores, counts, group),
        "box": _all_gather_var(boxes, counts, group),
    }
    hw = None
    if masks is not None:
        hw = (int(masks.shape[1]), int(masks.shape[2]))
        parts["mask"] = _all_gather_var(pack_masks(masks), counts, group)
    if rank != dst:
        return None
    out = {k: torch.cat(v, 0) for k, v in parts.items()}
    if hw is not None:
        out["mask"] = unpack_masks(out["mask"], hw[0], hw[1])
    else:
        out["mask"] = None
    return out
