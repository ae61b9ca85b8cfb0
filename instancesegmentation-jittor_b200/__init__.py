"""b200-roi-path: the RoI / NMS / mask hot path of Jittor/InstanceSegmentation-jittor,
hand-written for B200 (sm_100a) behind the reference's Python call signatures.

Every operator of the path calls libroipath_b200.so through ctypes (include/roipath.h); no kernel has
a torch-op or CPU fallback (a missing library raises RoiPathError) and nothing under oracle/ is ever
imported.  torch supplies device memory, streams and, on four non-default configuration branches that
no benchmark touches, the index glue AROUND the CUDA NMS (`Detect.traditional_nms`,
`PostProcessor._decode_only` / `filter_results` / `filter_results_v2` on already decoded BoxLists, the
training branch of `RPNPostProcessor.select_over_all_levels`: nonzero / stable sort / softmax on the device).
"""
from . import _lib
from ._lib import RoiPathError, lib, reload_env
from .layers import (ROIAlign, ROIPool, cython_nms, mask_prob_cuda, ml_nms, nms, nms_batched, roi_align,
                     roi_pool,
                     roi_align_backward, roi_align_forward)
from . import parallel
from .pose2seg import affine_paste, warpAffine
from .embed_mask import EmbedMaskMaskStage, embed_masks, resize_threshold
from .masker import (Masker, MaskPostProcessor, mask_select_sigmoid, paste_mask_in_image,
                     paste_masks)
from .poolers import (LevelMapper, LevelMapperwithArea, Pooler, convert_to_roi_format,
                      make_pooler, pooler_backward, pooler_forward)
from .box_head import PostProcessor, box_post, make_roi_box_post_processor
from .rpn import (BoxCoder, RPNPostProcessor, box_decode, make_rpn_postprocessor, rpn_pre_nms,
                  rpn_pre_nms_levels, rpn_select_levels)
from .structures import (BoxList, boxes_to_masks, crop_by_box, boxlist_iou, boxlist_ml_nms, boxlist_nms, boxlist_nms_batched, cat_boxlist,
                         remove_small_boxes)
from .yolact import (Detect, crop, decode, detect_prefilter, fast_nms_batched, lincomb_mask_loss,
                     lincomb_mask_loss_raw, mask_assemble, postprocess,
                     sanitize_coordinates)

__all__ = [n for n in dir() if not n.startswith("_")]
