/*
 * roipath.h -- C ABI of libroipath_b200.so: the B200 (sm_100a) implementation of the
 * RoI / NMS / mask hot path of Jittor/InstanceSegmentation-jittor.
 *
 * The reference has no C ABI for this path: its operators are Python callables that hand
 * inline CUDA text to jt.code / jt.candidate.  Each entry point below names the reference
 * interface it replaces (paths relative to the reference checkout; det/ =
 * detectron.jittor/detectron/, yol/ = Yolact.jittor/).  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add at each of those sites.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the parameter is documented "host";
 *   - the caller owns all memory (inputs, outputs, workspaces); nothing is retained;
 *   - `stream` is a cudaStream_t passed as void*; every call is asynchronous on it and
 *     performs no host synchronisation;
 *   - return value: 0 = ok, < 0 = invalid argument (RP_E_*), > 0 = a cudaError_t;
 *   - tensors are dense, row-major, float32 unless stated; feature maps are NCHW.
 */
#ifndef ROIPATH_H_
#define ROIPATH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RP_OK 0
#define RP_E_ARG (-1)       /* null pointer / non-positive size / unsupported value */
#define RP_E_WORKSPACE (-2) /* workspace too small */
#define RP_E_LIMIT (-3)     /* size exceeds a documented limit */

#define RP_MAX_LEVELS 8

/* exported with default visibility (the library is built with -fvisibility=hidden) */
#if defined(__GNUC__)
#define RP_API __attribute__((visibility("default")))
#else
#define RP_API
#endif

/* ABI version of this header; rp_version() returns the one the library was built with. */
#define RP_ABI_VERSION 1
RP_API int rp_version(void);
/* Compute capability the kernels were compiled for (100 for sm_100a). */
RP_API int rp_built_for_sm(void);
/* The RP_* tuning knobs (environment variables) are read once per process, at their first use;
 * rp_reload_env() forgets the values read so far (tests and tuning tools). */
RP_API void rp_reload_env(void);
/* Diagnostic.  The warp-specialised Pooler forward never waits unboundedly: a barrier that does not
 * complete makes the kernel trap (a CUDA error, not a hang).  With RP_RX_DEBUG=1 in the environment
 * every warp records {1, block, wait code, parity, counter} in a host-mapped buffer before the trap;
 * this copies the first n ints (8 per warp, 16 warps) to out and returns 1, or 0 when nothing was set up. */
RP_API int rp_debug_rx_timeouts(int* out, int n);

/* ------------------------------------------------------------------------------------
 * ROIAlign (legacy / non-aligned), forward and backward.
 * Replaces: roi_align(input, rois, output_size, spatial_scale, sampling_ratio)
 *           det/layers/roi_align.py:298-307 (jt.code over RoIAlignForward :116-169 and
 *           RoIAlignBackwardFeature :190-274).
 *   feat [N,C,H,W]; rois [K,5] = (batch_idx, x1, y1, x2, y2); out [K,C,PH,PW].
 *   sampling_ratio > 0: fixed grid; <= 0: adaptive ceil(roi/pooled) (double in forward,
 *   float in backward, as in the reference).
 *   mode 0 = tiled fast path (separable evaluation order, result within 1e-5 relative of
 *   the reference); mode 1 = one thread per output in the reference's exact operand order.
 * rp_roi_align_backward zero-fills grad_input itself (reference: cudaMemsetAsync :285).
 * ------------------------------------------------------------------------------------ */
RP_API int rp_roi_align_forward(const float* feat, int N, int C, int H, int W, const float* rois, int K,
                         int PH, int PW, float spatial_scale, float sampling_ratio, float* out,
                         int mode, void* stream);

RP_API int rp_roi_align_backward(const float* dout, const float* rois, int K, int N, int C, int H, int W,
                          int PH, int PW, float spatial_scale, float sampling_ratio,
                          float* grad_input, int mode, void* stream);

/* ------------------------------------------------------------------------------------
 * FPN level mapping and the fused multi-level Pooler.
 * Replaces: LevelMapper.__call__ det/modeling/poolers.py:31-43 (area with +1,
 *           det/structures/bounding_box.py:292-304), LevelMapperwithArea.__call__ :62-73,
 *           Pooler.execute :123-156 (per-level nonzero/gather/ROIAlign/scatter loop).
 *   rois [K,5]; levels [K] int32 in [0,L).  One launch pools every RoI from the map of its
 *   level and writes it to its ORIGINAL row of out [K,C,PH,PW].
 *   feats/H/W/scales and grads are HOST arrays of length L (L <= RP_MAX_LEVELS).
 * ------------------------------------------------------------------------------------ */
RP_API int rp_level_map(const float* rois, int K, int k_min, int k_max, float canonical_scale,
                 int canonical_level, float eps, int* levels, void* stream);

RP_API int rp_level_map_area(const float* rois, const float* img_area, int K, int k_min, int k_max,
                      float eps, int* levels, void* stream);

RP_API int rp_pooler_forward(const float* const* feats, const int* H, const int* W, const float* scales,
                      int L, int N, int C, const float* rois, const int* levels, int K, int PH,
                      int PW, float sampling_ratio, float* out, int mode, void* stream);

RP_API int rp_pooler_backward(const float* dout, const float* rois, const int* levels, int K,
                       const int* H, const int* W, const float* scales, int L, int N, int C,
                       int PH, int PW, float sampling_ratio, float* const* grads, int mode,
                       void* stream);

/* Channels-last entries (north_star: "NHWC-aware bilinear gathers"; the reference's layout is NCHW,
 * det/layers/roi_align.py:116-169, so these are an ADDITIONAL path for channels-last backbones):
 * feat / feats[l] are [N, H_l, W_l, C] fp32, every tap is one 16-byte load of 4 channels; the output
 * stays [K, C, PH, PW].  Same semantics and tolerance as mode 0 of the NCHW entries. */
RP_API int rp_roi_align_forward_nhwc(const float* feat, int N, int C, int H, int W, const float* rois,
                                     int K, int PH, int PW, float spatial_scale, float sampling_ratio,
                                     float* out, void* stream);
RP_API int rp_pooler_forward_nhwc(const float* const* feats, const int* H, const int* W,
                                  const float* scales, int L, int N, int C, const float* rois,
                                  const int* levels, int K, int PH, int PW, float sampling_ratio,
                                  float* out, void* stream);

/* ------------------------------------------------------------------------------------
 * Greedy NMS family.
 * Replaces: nms(dets, scores, thr) det/layers/nms.py:12-14 (-> jittor.nms),
 *           ml_nms / _ml_nms :18-40 (jt.argsort + jt.candidate with the +1 IoU fail
 *           condition :28-34), the keep[:max_proposals] truncation of boxlist_nms
 *           det/structures/boxlist_ops.py:31-32, and the comparator variant of
 *           yol/utils/cython_nms.pyx:24-74 (flags & RP_NMS_GE).
 * A batch is B independent problems ("segments"): problem b owns rows
 * [offsets[b], offsets[b+1]) of boxes/scores/labels (offsets: DEVICE int32 [B+1]).
 *   boxes [n,4] xyxy; scores [n]; labels [n] float or NULL; thr compared as the reference
 *   does ((double)iou > thr, or iou >= (float)thr with RP_NMS_GE).
 *   keep [n] int32: for problem b, keep[offsets[b] + i], i < n_keep[b], are the kept
 *   indices RELATIVE to offsets[b], in descending-score order (stable: equal scores keep
 *   ascending index); n_keep [B] int32.  max_keep > 0 stops after that many.
 *   n_total >= offsets[B] and n_max >= max segment length are passed by the host; both may
 *   be upper bounds (rows in [offsets[B], n_total) must be readable and are ignored), so a
 *   producer with device-side counts (rp_rpn_pre_nms) chains without a host read.
 * flags: RP_NMS_GE               suppress when iou >= (float)thr (cython_nms.pyx:69)
 *        RP_NMS_ASCENDING        return each keep list sorted by ascending index (:74)
 *        RP_NMS_TIES_HIGH_FIRST  equal scores: higher index first (argsort()[::-1], :35)
 *        RP_NMS_CYTHON           all three: the yol/utils/cython_nms.pyx contract
 *        RP_NMS_PRESORTED        the caller guarantees that the scores of every problem are
 *                                non-increasing (the stable argsort is then the identity and is
 *                                skipped: rp_rpn_pre_nms emits its candidates in that order).
 *                                Not with RP_NMS_TIES_HIGH_FIRST.  An input that breaks the promise
 *                                is processed in the given order (greedy NMS in input order).
 * ------------------------------------------------------------------------------------ */
#define RP_NMS_GE 1
#define RP_NMS_ASCENDING 2
#define RP_NMS_TIES_HIGH_FIRST 4
#define RP_NMS_CYTHON (RP_NMS_GE | RP_NMS_ASCENDING | RP_NMS_TIES_HIGH_FIRST)
#define RP_NMS_PRESORTED 8

RP_API size_t rp_nms_workspace_bytes(int n_total, int n_max, int B);

RP_API int rp_nms_batched(const float* boxes, const float* scores, const float* labels,
                   const int* offsets, int B, int n_total, int n_max, double thr, int max_keep,
                   int flags, int* keep, int* n_keep, void* workspace, size_t workspace_bytes,
                   void* stream);

/* rp_nms_batched_static: rp_nms_batched without a workspace argument, for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO
 * counterpart in the reference.  The workspace is
 * a fixed array of rp_nms_static_capacity_bytes() = 4 MiB of device memory that every GPU which loads
 * the library holds until the process exits; the call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  Shapes whose rp_nms_workspace_bytes(n_total, n_max, B) exceeds the
 * capacity return RP_E_LIMIT (one image of 20 000 boxes needs 1.3 MiB, 8 x 1000 boxes 2.8 MiB).
 * All calls share the array: calls that may overlap on the GPU (different streams or host threads)
 * race; serialise them.  The array keeps the previous call's contents, the call clears what must
 * start at zero.  Same arguments, results and error codes as rp_nms_batched otherwise. */
RP_API size_t rp_nms_static_capacity_bytes(void);

RP_API int rp_nms_batched_static(const float* boxes, const float* scores, const float* labels,
                          const int* offsets, int B, int n_total, int n_max, double thr,
                          int max_keep, int flags, int* keep, int* n_keep, void* stream);

/* ------------------------------------------------------------------------------------
 * Yolact: box decode, detection pre-filter, FastNMS.
 * Replaces: decode(loc, priors) yol/layers/box_utils.py:272-316;
 *           Detect.detect 'Slices' yol/layers/functions/detection.py:81-99;
 *           Detect.fast_nms :139-189 and Detect.cc_fast_nms :113-137
 *           (jaccard / intersect yol/layers/box_utils.py:31-78).
 * ------------------------------------------------------------------------------------ */
RP_API int rp_yolact_decode(const float* loc, const float* priors, int n, float* out, void* stream);

/* Batched FastNMS.  Image b owns candidates [offsets[b], offsets[b+1]) (DEVICE int32 [B+1])
 * of boxes [n_total,4] (relative xyxy) and coeff [n_total,D]; scores is [C, n_total]
 * (class-major, row stride n_total).  Per image at most max_det detections, sorted by
 * descending score, are written to row b of
 *   out_boxes [B,max_det,4], out_coeff [B,max_det,D], out_classes [B,max_det] int32,
 *   out_scores [B,max_det], out_index [B,max_det] int32 (candidate index relative to
 *   offsets[b]; may be NULL), n_out [B] int32.
 * cross_class != 0 selects cc_fast_nms (max_det is then ignored up to top_k).
 * Limits: top_k <= 256. */
/* Detect pre-stage for a whole batch: Detect.__call__ / detect 'Slices'
 * yol/layers/functions/detection.py:64-98 -- decode (box_utils.py:272-316), class-max over the
 * foreground classes, conf_scores > conf_thresh, and the gathers of boxes / mask coefficients /
 * scores, without the boolean-mask synchronisation.
 *   loc [B,P,4]; conf [B,P,NC] (class 0 = background); mask [B,P,D]; priors [P,4].
 *   Kept priors of image b, in prior order, are rows [out_offsets[b], out_offsets[b+1]) of
 *   out_boxes [cand_cap,4], out_coeff [cand_cap,D], out_prior [cand_cap] (prior index) and
 *   COLUMNS of out_scores [NC-1, cand_cap] (row c = class c+1) -- exactly the
 *   (boxes, coeff, scores, offsets, n_total = cand_cap, C = NC-1) inputs of rp_fast_nms_batched.
 *   out_offsets [B+1] DEVICE int32; n_cand [1] DEVICE: the real number of kept priors; when it
 *   exceeds cand_cap the run was truncated and the caller repeats it with more room. */
RP_API size_t rp_yolact_prefilter_workspace_bytes(int B, int P, int cand_cap);

RP_API int rp_yolact_prefilter(const float* loc, const float* conf, const float* mask,
                        const float* priors, int B, int P, int NC, int D, float conf_thresh,
                        int cand_cap, float* out_boxes, float* out_coeff, float* out_scores,
                        int* out_prior, int* out_offsets, int* n_cand, void* workspace,
                        size_t workspace_bytes, void* stream);

RP_API size_t rp_fast_nms_workspace_bytes(int n_total, int n_max, int B, int C, int top_k);

RP_API int rp_fast_nms_batched(const float* boxes, const float* coeff, const float* scores,
                        const int* offsets, int B, int n_total, int n_max, int C, int D,
                        float iou_thr, int top_k, int max_det, int second_threshold,
                        float conf_thresh, int cross_class, float* out_boxes, float* out_coeff,
                        int* out_classes, float* out_scores, int* out_index, int* n_out,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Prototype x coefficient mask assembly, fused with sigmoid, crop, bilinear upsample and
 * threshold.
 * Replaces: postprocess(...) lincomb branch yol/layers/output_utils.py:60-101 with
 *           crop / sanitize_coordinates yol/layers/box_utils.py:333-378.
 * Image b: proto [B,ph,pw,D] (NHWC), and detections [det_offsets[b], det_offsets[b+1])
 * (DEVICE int32 [B+1]; or NULL with B == 1 and m_total detections) of coeff [m_total,D]
 * and boxes [m_total,4] (relative xyxy).
 *   out_masks [m_total,out_h,out_w] uint8 (0/1); out_boxes [m_total,4] int32 (may be NULL);
 *   low_out [m_total,ph,pw] float (may be NULL): cropped low-resolution probabilities.
 * Limits: D <= 64.
 * ------------------------------------------------------------------------------------ */
RP_API int rp_mask_assemble(const float* proto, const float* coeff, const float* boxes,
                     const int* det_offsets, int B, int m_total, int ph, int pw, int D, int out_h,
                     int out_w, int crop, uint8_t* out_masks, int32_t* out_boxes, float* low_out,
                     void* stream);

/* ------------------------------------------------------------------------------------
 * Mask R-CNN mask post-processing.
 * Replaces: MaskPostProcessor.execute det/modeling/roi_heads/mask_head/inference.py:38-45
 *           (sigmoid + label channel select) and Masker / paste_mask_in_image :90-162,
 *           :175-205 (pad, expand box, int32 truncation, bilinear resize, threshold, paste).
 *   logits [K,NC,M,M], labels [K] int32 -> prob [K,M,M].
 *   masks [n,M,M] probabilities, boxes [n,4] xyxy -> out [n,im_h,im_w] uint8 (0/1), fully
 *   written (zero outside the box).  thresh < 0 selects the (mask*255) != 0 branch.
 * ------------------------------------------------------------------------------------ */
RP_API int rp_mask_select_sigmoid(const float* logits, const int* labels, int K, int NC, int M,
                           float* out, void* stream);

RP_API int rp_paste_masks(const float* masks, const float* boxes, int n, int M, int im_h, int im_w,
                   float thresh, int padding, uint8_t* out, void* stream);

/* ------------------------------------------------------------------------------------
 * Yolact lincomb mask loss, forward + backward in one launch (the training twin of
 * rp_mask_assemble).
 * Replaces, per image: yol/layers/modules/multibox_loss.py:616-660 -- pred = sigmoid(proto @ coef^T),
 *           crop(pred, boxes) (box_utils.py:355-378, padding 1; boxes NULL = mask_proto_crop off),
 *           binary_cross_entropy(clamp(pred,0,1), mask_t) (:8-9), per-instance sum times scale[i]
 *           (scale NULL = 1; with mask_proto_normalize_emulate_roi_pooling the caller passes
 *           H*W / (gt_w*W) / (gt_h*H), :645-649) -- and the autograd backward of sum_i loss[i].
 *   proto [H,W,D], coef [n,D], mask_t [H,W,n] (0/1 targets), boxes [n,4] relative xyxy, scale [n]
 *   -> loss [n], d_proto [H,W,D] (fully written), d_coef [n,D].
 * Limits: D in {8,16,32,64}; n*(D+134) floats of shared memory (n <= ~280 at D = 32).
 * ------------------------------------------------------------------------------------ */
RP_API int rp_lincomb_mask_loss(const float* proto, const float* coef, const float* mask_t,
                                const float* boxes, const float* scale, int H, int W, int D, int n,
                                float* loss, float* d_proto, float* d_coef, void* stream);

/* ------------------------------------------------------------------------------------
 * Pose2Seg paste-back (BASELINE configs[4] names Pose2Seg).
 * Replaces: warpAffine(predmap, H_e2e, width, height) Pose2Seg.jittor/modeling/build_model.py:31-43 =
 *           cv2.warpAffine(predmap, H_e2e[0:2], (width, height), WARP_INVERSE_MAP + INTER_LINEAR,
 *           BORDER_CONSTANT)[:, :, channel] > 0.5, bit for bit (OpenCV's 10-bit fixed-point source
 *           coordinates and 32x32 bilinear weight table).
 *   pred [n,h,w,C] float (HWC, as the seg head's softmax output), M [n,6] DOUBLE row-major 2x3
 *   (output pixel -> source), out [n,height,width] uint8 (0/1).
 * ------------------------------------------------------------------------------------ */
RP_API int rp_affine_paste(const float* pred, const double* M, int n, int h, int w, int C, int channel,
                           int width, int height, uint8_t* out, void* stream);

/* ------------------------------------------------------------------------------------
 * EmbedMask mask stage.
 * Replaces, per image: compute_mask_prob + the box-mask product of forward_for_mask,
 *           det/modeling/rpn/embed_mask/inference.py:153-169,193-197 (boxes / stride, int cast, +-2
 *           growth, clamp; layers/mask_prob.py:36-57; boxlist_ops.py:237-261 boxes_to_masks), and the
 *           post_process_masks branch :199-202 (bilinear upsample by the stride, > mask_th, crop).
 *   rp_embed_masks: pixel_embed [dim,m_h,m_w], proposal_embed [n,dim], proposal_margin [n],
 *                   boxes [n,4] xyxy in image pixels -> masks [n,m_h,m_w] float (fully written).
 *   rp_resize_threshold: masks [n,h,w] float -> out [n,crop_h,crop_w] uint8 = (bilinear resize to
 *                   (H,W), align_corners=False) > thresh, rows < crop_h, columns < crop_w.
 * Limits: dim <= 128, n <= 65535.
 * ------------------------------------------------------------------------------------ */
RP_API int rp_embed_masks(const float* pixel_embed, const float* proposal_embed, const float* proposal_margin,
                          const float* boxes, float stride, int n, int dim, int m_h, int m_w, float* masks,
                          void* stream);
RP_API int rp_resize_threshold(const float* masks, int n, int h, int w, int H, int W, float thresh,
                               int crop_h, int crop_w, uint8_t* out, void* stream);

/* Bit-packing of binary masks for the final gather of kept detections (the one exchange of the
 * image-sharded path; the reference is single-process and has no counterpart).
 *   masks [m, npix] uint8 (0 / non-zero) -> out [m, ceil(npix/8)] uint8, pixel i in bit (i & 7) of byte i >> 3. */
RP_API int rp_pack_mask_bits(const uint8_t* masks, long long m, long long npix, uint8_t* out, void* stream);

/* ------------------------------------------------------------------------------------
 * RPN proposal stage in front of the NMS (the caller of rp_nms_batched).
 * Replaces: BoxCoder.decode det/modeling/box_coder.py:53-94 (rp_box_decode), and in
 *           RPNPostProcessor.forward_for_single_feature_map det/modeling/rpn/inference.py:83-117
 *           everything before boxlist_nms: permute_and_flatten (rpn/utils.py:10-15), sigmoid,
 *           topk(pre_nms_top_n, sorted), the regression / anchor gathers, decode,
 *           clip_to_image(remove_empty=False) bounding_box.py:263-277 and remove_small_boxes
 *           boxlist_ops.py:64-76.
 * rp_box_decode: rel_codes [n,4K] (K codes per reference box, the 0::4 .. 3::4 strides of
 *   :73-76), boxes [n,4] xyxy (all 16-byte aligned) -> out [n,4K].
 * rp_rpn_pre_nms, one FPN level, N images:
 *   objectness [N,A,H,W]; box_regression [N,4A,H,W]; anchors [N,H*W*A,4] in the reference's
 *   flattened (h,w,a) order; im_wh [N,2] float (width, height of each image).
 *   k = min(pre_nms_top_n, A*H*W) candidates per image, descending sigmoid score, equal
 *   scores by ascending flattened index.
 *   out_offsets [N+1] DEVICE int32: image b owns rows [out_offsets[b], out_offsets[b+1]) of
 *   out_boxes [N*k,4], out_scores [N*k], out_index [N*k] (flattened anchor index; may be
 *   NULL), in rank order with the small boxes removed -- the layout rp_nms_batched reads.
 * rp_rpn_pre_nms_levels: the same for L <= RP_MAX_LEVELS levels in ONE set of launches (one
 *   sort for the whole batch).  objectness / box_regression / anchors / A / H / W are HOST
 *   arrays of length L.  Problem s = l*N + b (level-major) owns rows
 *   [out_offsets[s], out_offsets[s+1]); out_offsets [L*N+1]; outputs hold
 *   sum_l N*k_l rows, k_l = min(pre_nms_top_n, A_l*H_l*W_l).
 * rp_rpn_select_levels: RPNPostProcessor.select_over_all_levels, test-time branch
 *   inference.py:175-183: per image, the top_n best of the boxes its L problems kept
 *   (offsets / keep / n_keep as returned by rp_nms_batched for the L*N problems above),
 *   descending score, equal scores in (level, keep rank) order = the cat_boxlist order :143-144.
 *   n_cap = rows of boxes / scores / keep.  out_boxes [N*top_n,4], out_scores [N*top_n]: image
 *   b owns rows [b*top_n, b*top_n + out_count[b]).
 *   One launch (one CTA per image: radix select + sort of the selected keys) for top_n <= 4096,
 *   else a stable radix sort of the whole batch in the workspace; identical results.
 * ------------------------------------------------------------------------------------ */
RP_API int rp_box_decode(const float* rel_codes, const float* boxes, int n, int K, float wx,
                  float wy, float ww, float wh, float bbox_xform_clip, float* out, void* stream);

RP_API size_t rp_rpn_pre_nms_workspace_bytes(int N, int A, int H, int W, int pre_nms_top_n);

RP_API int rp_rpn_pre_nms(const float* objectness, const float* box_regression, const float* anchors,
                   const float* im_wh, int N, int A, int H, int W, int pre_nms_top_n,
                   float min_size, float wx, float wy, float ww, float wh, float bbox_xform_clip,
                   float* out_boxes, float* out_scores, int* out_index, int* out_offsets,
                   void* workspace, size_t workspace_bytes, void* stream);

RP_API size_t rp_rpn_pre_nms_levels_workspace_bytes(int L, int N, const int* A, const int* H,
                                             const int* W, int pre_nms_top_n);

RP_API int rp_rpn_pre_nms_levels(const float* const* objectness, const float* const* box_regression,
                          const float* const* anchors, const int* A, const int* H, const int* W,
                          int L, const float* im_wh, int N, int pre_nms_top_n, float min_size,
                          float wx, float wy, float ww, float wh, float bbox_xform_clip,
                          float* out_boxes, float* out_scores, int* out_index, int* out_offsets,
                          void* workspace, size_t workspace_bytes, void* stream);

RP_API size_t rp_rpn_select_workspace_bytes(int n_cap, int L, int N);

RP_API int rp_rpn_select_levels(const float* boxes, const float* scores, const int* offsets,
                         const int* keep, const int* n_keep, int L, int N, int n_cap, int top_n,
                         float* out_boxes, float* out_scores, int* out_count, void* workspace,
                         size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Box-head post-processor (the caller of the per-class NMS).
 * Replaces: PostProcessor.execute det/modeling/roi_heads/box_head/inference.py:45-90 (softmax,
 *           class-wise BoxCoder.decode, clip_to_image) and filter_results :112-195 (per class
 *           j >= 1: score > score_thresh, boxlist_nms; cat over classes; when more than
 *           detections_per_img survive, keep those with score >= the detections_per_img-th
 *           largest score -- jt.kthvalue :180-187, ties all kept).
 *   class_logits [R,C]; box_regression [R,reg_cols] (reg_cols = 4C, or any multiple of 4 with
 *   cls_agnostic != 0: the LAST four columns are used for every class, :66-67,71-72);
 *   proposals [R,4] xyxy; image b owns rows [row_offsets[b], row_offsets[b+1]) (DEVICE int32
 *   [N+1]); r_max >= the largest image's row count; im_wh [N,2] float (width, height).
 *   cand_cap bounds the number of (row, class) pairs above the threshold that are processed
 *   (R*(C-1) is always enough); n_cand [1] DEVICE receives the real number -- a caller that
 *   passed a smaller cap compares and repeats the call when n_cand > cand_cap.
 *   prob_out [R,C] softmax probabilities (may be NULL).
 *   Image b's detections are rows [b*out_cap, b*out_cap + min(out_count[b], out_cap)) of
 *   out_boxes [N*out_cap,4], out_scores, out_labels (int32 class), out_rows (int32 proposal
 *   row relative to row_offsets[b]; may be NULL), class-major, descending score inside a
 *   class; out_count[b] is the full count (it can exceed detections_per_img on ties).
 * nms_thresh <= 0 skips NMS (boxlist_nms returns its input, boxlist_ops.py:24-25): every candidate
 * is kept in nonzero() order; the score threshold and the detections_per_img cut still apply.
 * Limits: N*(C-1) <= 65535.
 * ------------------------------------------------------------------------------------ */
RP_API size_t rp_box_post_workspace_bytes(int R, int C, int N, int r_max, int cand_cap);

RP_API int rp_box_post(const float* class_logits, const float* box_regression, const float* proposals,
                const int* row_offsets, const float* im_wh, int R, int C, int N, int r_max,
                int reg_cols, int cls_agnostic, float score_thresh, double nms_thresh,
                int detections_per_img, float wx, float wy, float ww, float wh,
                float bbox_xform_clip, int cand_cap, float* prob_out, float* out_boxes,
                float* out_scores, int* out_labels, int* out_rows, int* out_count, int out_cap,
                int* n_cand, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Other RoI-family operators.
 * rp_boxlist_iou: boxlist_iou det/structures/boxlist_ops.py:79-116 -- boxes1 [N,4], boxes2 [M,4]
 *   xyxy (16-byte aligned) -> out [N,M] = inter / (area1 + area2 - inter), TO_REMOVE = 1.
 * rp_roi_pool_forward / _backward: roi_pool det/layers/roi_pool.py:16-88,101-148 -- feat
 *   [N,C,H,W], rois [K,5] -> out [K,C,PH,PW] (max, 0 for an empty bin) and argmax [K,C,PH,PW]
 *   int32 (offset inside the plane, -1 when nothing was pooled); the backward zero-fills
 *   grad_input [N,C,H,W] (:143) and adds dout at the argmax positions.
 * rp_mask_prob: mask_prob_cuda det/layers/mask_prob.py:9-103 (EmbedMask) -- embed_pixel
 *   [num_pixel,dim], embed_center [num_center,dim], sigma_center [num_center], boxes
 *   [num_center,4] int32 (x1,y1,x2,y2, end exclusive), box_areas [num_center] int32 ->
 *   probs [num_center,num_pixel] = exp(-|e_p - e_c|^2 * sigma) inside each centre's box, 0
 *   elsewhere (the launch zero-fills, :60).  Limits: dim <= 128, num_center <= 65535.
 * ------------------------------------------------------------------------------------ */
/* rp_box_masks: boxes_to_masks det/structures/boxlist_ops.py:237-261 (crop = 0: indicator of the
 *   box grown by 1 px + padding * size, upper clamp w / h) and crop_by_box :262-281 (crop = 1:
 *   upper clamp w-1 / h-1, masks_out = masks_in * indicator).  boxes [n,4]; indicator [n,h,w]
 *   uint8 (may be NULL when masks_out is given); masks_in / masks_out [n,h,w] float or NULL. */
RP_API int rp_box_masks(const float* boxes, int n, int h, int w, float padding, int crop,
                 const float* masks_in, float* masks_out, uint8_t* indicator, void* stream);

RP_API int rp_boxlist_iou(const float* boxes1, int N, const float* boxes2, int M, float* out,
                   void* stream);

RP_API int rp_roi_pool_forward(const float* feat, int N, int C, int H, int W, const float* rois, int K,
                        int PH, int PW, float spatial_scale, float* out, int* argmax,
                        void* stream);

RP_API int rp_roi_pool_backward(const float* dout, const float* rois, const int* argmax, int K, int N,
                         int C, int H, int W, int PH, int PW, float* grad_input, void* stream);

RP_API int rp_mask_prob(const float* embed_pixel, const float* embed_center, const float* sigma_center,
                 const int* boxes, const int* box_areas, int num_pixel, int num_center, int dim,
                 int mask_width, float* probs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ROIPATH_H_ */
