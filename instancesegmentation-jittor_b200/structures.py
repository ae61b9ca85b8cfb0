"""Minimal BoxList and the boxlist ops on the hot path, with the reference's semantics.

Mirrors detectron.jittor/detectron/structures/bounding_box.py (BoxList: __init__ :19-41,
fields :50-64, convert :65-94, __getitem__ :253-257, clip_to_image :263-290, area :292-304,
copy_with_fields :306-316) and structures/boxlist_ops.py (boxlist_nms :12-34,
boxlist_ml_nms :36-60, remove_small_boxes :64-76, cat_boxlist :127-152).
Only what Pooler / Masker / the NMS call sites need is carried over; bbox tensors are torch
CUDA tensors.
"""
import torch

from . import _lib, layers

TO_REMOVE = 1


class BoxList(object):
    def __init__(self, bbox, image_size, mode="xyxy"):
        if mode not in ("xyxy", "xywh"):
            raise ValueError("mode should be 'xyxy' or 'xywh'")
        if not isinstance(bbox, torch.Tensor):
            bbox = torch.as_tensor(bbox, dtype=torch.float32)
        if bbox.numel() == 0:
            bbox = bbox.reshape(0, 4)
        if bbox.ndim != 2 or bbox.shape[-1] != 4:
            raise ValueError("bbox should be [n,4], got %s" % (tuple(bbox.shape),))
        self.bbox = bbox
        self.size = image_size  # (image_width, image_height)
        self.mode = mode
        self.extra_fields = {}

    # ---- fields
    def add_field(self, field, field_data):
        self.extra_fields[field] = field_data

    def get_field(self, field):
        return self.extra_fields[field]

    def has_field(self, field):
        return field in self.extra_fields

    def fields(self):
        return list(self.extra_fields.keys())

    def _copy_extra_fields(self, bbox):
        for k, v in bbox.extra_fields.items():
            self.extra_fields[k] = v

    def to(self, device):
        out = BoxList(self.bbox.to(device), self.size, self.mode)
        for k, v in self.extra_fields.items():
            out.add_field(k, v.to(device) if hasattr(v, "to") else v)
        return out

    # ---- geometry
    def _split_into_xyxy(self):
        b = self.bbox
        if self.mode == "xyxy":
            return b[:, 0:1], b[:, 1:2], b[:, 2:3], b[:, 3:4]
        xmin, ymin, w, h = b[:, 0:1], b[:, 1:2], b[:, 2:3], b[:, 3:4]
        return xmin, ymin, xmin + (w - TO_REMOVE).clamp(min=0), ymin + (h - TO_REMOVE).clamp(min=0)

    def convert(self, mode):
        if mode not in ("xyxy", "xywh"):
            raise ValueError("mode should be 'xyxy' or 'xywh'")
        if mode == self.mode:
            return self
        xmin, ymin, xmax, ymax = self._split_into_xyxy()
        if mode == "xyxy":
            bbox = torch.cat((xmin, ymin, xmax, ymax), dim=-1)
        else:
            bbox = torch.cat((xmin, ymin, xmax - xmin + TO_REMOVE, ymax - ymin + TO_REMOVE), dim=-1)
        out = BoxList(bbox, self.size, mode=mode)
        out._copy_extra_fields(self)
        return out

    def __getitem__(self, item):
        out = BoxList(self.bbox[item], self.size, self.mode)
        for k, v in self.extra_fields.items():
            out.add_field(k, v[item])
        return out

    def __len__(self):
        return self.bbox.shape[0]

    def clip_to_image(self, remove_empty=True):
        if self.bbox.numel() == 0:
            return self
        b = self.bbox
        b[:, 0].clamp_(min=0, max=self.size[0] - TO_REMOVE)
        b[:, 1].clamp_(min=0, max=self.size[1] - TO_REMOVE)
        b[:, 2].clamp_(min=0, max=self.size[0] - TO_REMOVE)
        b[:, 3].clamp_(min=0, max=self.size[1] - TO_REMOVE)
        if remove_empty:
            keep = (b[:, 3] > b[:, 1]) & (b[:, 2] > b[:, 0])
            return self[keep]
        return self

    def area(self):
        box = self.bbox
        if self.mode == "xyxy":
            return (box[:, 2] - box[:, 0] + TO_REMOVE) * (box[:, 3] - box[:, 1] + TO_REMOVE)
        return box[:, 2] * box[:, 3]

    def image_area(self):
        return torch.full((len(self),), float(self.size[0] * self.size[1]), dtype=torch.float32,
                          device=self.bbox.device)

    def copy_with_fields(self, fields, skip_missing=False):
        out = BoxList(self.bbox, self.size, self.mode)
        if not isinstance(fields, (list, tuple)):
            fields = [fields]
        for field in fields:
            if self.has_field(field):
                out.add_field(field, self.get_field(field))
            elif not skip_missing:
                raise KeyError("Field '{}' not found in {}".format(field, self))
        return out

    def __repr__(self):
        return "BoxList(num_boxes={}, image_width={}, image_height={}, mode={})".format(
            len(self), self.size[0], self.size[1], self.mode)


def boxlist_nms(boxlist, nms_thresh, max_proposals=-1, score_field="scores"):
    """boxlist_ops.py:12-34.  The keep[:max_proposals] truncation is pushed into the scan
    kernel (it stops after max_proposals keeps), which yields the same list."""
    if nms_thresh <= 0 or len(boxlist) == 0:
        return boxlist
    mode = boxlist.mode
    boxlist = boxlist.convert("xyxy")
    keep = layers._nms_single(boxlist.bbox, boxlist.get_field(score_field), None, nms_thresh,
                              max_keep=max_proposals if max_proposals > 0 else -1)
    return boxlist[keep].convert(mode)


def boxlist_ml_nms(boxlist, nms_thresh, max_proposals=-1, score_field="scores",
                   label_field="labels"):
    """boxlist_ops.py:36-60 (labels compared as floats, :56)."""
    if nms_thresh <= 0:
        return boxlist
    mode = boxlist.mode
    boxlist = boxlist.convert("xyxy")
    keep = layers._nms_single(boxlist.bbox, boxlist.get_field(score_field),
                              boxlist.get_field(label_field).float(), nms_thresh,
                              max_keep=max_proposals if max_proposals > 0 else -1)
    return boxlist[keep].convert(mode)


def boxlist_nms_batched(boxlists, nms_thresh, max_proposals=-1, score_field="scores",
                        label_field=None):
    """All images' (or all (level, image) pairs') NMS problems in ONE set of launches with a
    single host read of the keep counts -- the B200 form of the per-image python loop at
    det/modeling/rpn/inference.py:112-124.  Returns a list of BoxLists."""
    if nms_thresh <= 0:
        return list(boxlists)
    lens = [len(b) for b in boxlists]
    if sum(lens) == 0:
        return list(boxlists)
    modes = [b.mode for b in boxlists]
    xyxy = [b.convert("xyxy") for b in boxlists]
    dev = xyxy[0].bbox.device
    boxes = torch.cat([b.bbox.float() for b in xyxy], 0).contiguous()
    scores = torch.cat([b.get_field(score_field).float().reshape(-1) for b in xyxy], 0).contiguous()
    labels = None
    if label_field is not None:
        labels = torch.cat([b.get_field(label_field).float().reshape(-1) for b in xyxy], 0).contiguous()
    offs = [0]
    for n in lens:
        offs.append(offs[-1] + n)
    offsets = torch.tensor(offs, dtype=torch.int32, device=dev)
    keep, n_keep = layers.nms_batched(boxes, scores, offsets, nms_thresh, labels,
                                      max_keep=max_proposals if max_proposals > 0 else -1,
                                      n_max=max(lens))
    counts = n_keep.cpu().tolist()  # one sync for the whole batch
    out = []
    for i, b in enumerate(xyxy):
        k = keep[offs[i]: offs[i] + counts[i]].long()
        out.append(b[k].convert(modes[i]))
    return out


def remove_small_boxes(boxlist, min_size):
    """boxlist_ops.py:64-76."""
    xywh = boxlist.convert("xywh").bbox
    keep = (xywh[:, 2] >= min_size) & (xywh[:, 3] >= min_size)
    return boxlist[keep]


def boxlist_iou(boxlist1, boxlist2):
    """boxlist_ops.py:79-116: [N,M] IoU with the legacy +1 (one launch, no [N,M,2] temporaries)."""
    if boxlist1.size != boxlist2.size:
        raise RuntimeError("boxlists should have same image size, got {}, {}".format(boxlist1, boxlist2))
    b1 = _lib.as_f32(boxlist1.convert("xyxy").bbox, "boxlist1")
    b2 = _lib.as_f32(boxlist2.convert("xyxy").bbox, "boxlist2")
    N, M = int(b1.shape[0]), int(b2.shape[0])
    out = torch.empty((N, M), dtype=torch.float32, device=b1.device)
    rc = _lib.lib().rp_boxlist_iou(_lib.dptr(b1), N, _lib.dptr(b2), M, _lib.dptr(out),
                                   _lib.stream_ptr(b1.device))
    _lib.check(rc, "rp_boxlist_iou")
    return out


def _box_masks(boxes, h, w, padding, crop, masks=None):
    boxes = _lib.as_f32(boxes, "boxes")
    n = int(boxes.shape[0])
    dev = boxes.device
    ind = torch.empty((n, int(h), int(w)), dtype=torch.uint8, device=dev)
    out = None
    if masks is not None:
        masks = _lib.as_f32(masks, "masks")
        out = torch.empty_like(masks)
    rc = _lib.lib().rp_box_masks(_lib.dptr(boxes), n, int(h), int(w), float(padding), int(crop),
                                 _lib.dptr(masks), _lib.dptr(out), _lib.dptr(ind),
                                 _lib.stream_ptr(dev))
    _lib.check(rc, "rp_box_masks")
    return out, ind.bool()


def boxes_to_masks(boxes, h, w, padding=0.0):
    """boxlist_ops.py:237-261: [n,h,w] bool indicator of every box grown by 1 px + padding * size."""
    return _box_masks(boxes, h, w, padding, 0)[1]


def crop_by_box(masks, box, padding=0.0):
    """boxlist_ops.py:262-281: (masks * indicator, indicator)."""
    n, h, w = masks.shape
    return _box_masks(box, h, w, padding, 1, masks)


def cat_boxlist(bboxes):
    """boxlist_ops.py:127-152."""
    assert isinstance(bboxes, (list, tuple)) and len(bboxes) > 0
    size, mode, fields = bboxes[0].size, bboxes[0].mode, set(bboxes[0].fields())
    assert all(b.size == size and b.mode == mode and set(b.fields()) == fields for b in bboxes)
    out = BoxList(torch.cat([b.bbox for b in bboxes], dim=0), size, mode)
    for f in fields:
        out.add_field(f, torch.cat([b.get_field(f) for b in bboxes], dim=0))
    return out
