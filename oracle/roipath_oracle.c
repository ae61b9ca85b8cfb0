/*
 * roipath_oracle.c -- CPU restatement of the reference's RoI / NMS / mask hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline.
 * The product path (instancesegmentation-jittor_b200/) never imports it and fails
 * loudly when its CUDA library is missing.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the reference checkout of Jittor/InstanceSegmentation-jittor):
 *   det/ = detectron.jittor/detectron/     yol/ = Yolact.jittor/
 *
 * Pinning status (see DESIGN.md "Oracle"):
 *   - orc_roi_align_forward/backward: pinned bit-for-bit against the reference's own
 *     kernel text compiled as C++ (oracle/_ref, built by oracle/build_ref.py) and
 *     against torchvision.ops.roi_align(aligned=False).
 *   - orc_nms: pinned against the golden vectors of
 *     detectron.jittor/tests/test_nms.py:11-58,60-217 (keep sets).
 *   - orc_cython_nms: pinned against the reference's cython_nms.pyx compiled here.
 *   - orc_box_decode: pinned against the known-answer vectors of
 *     detectron.jittor/tests/test_box_coder.py (tests/golden/box_coder_kat.npz, atol 1e-4).
 *   - orc_rpn_pre_nms: the decode inside is pinned as above; the sort's tie order, the clip and
 *     the small-box test follow the source text (PARITY UNPINNED for those).
 *   - orc_box_post_image / orc_softmax_rows: PARITY UNPINNED (no known-answer test in the
 *     reference); the decode and the NMS inside are the pinned functions above.
 *   - everything else (ml_nms label term, FastNMS, jaccard, crop, assembly, bilinear
 *     resize, paste-back, level mapping): PARITY UNPINNED -- the reference holds no
 *     known-answer test and Jittor is not installable; the restatement follows the
 *     reference source text plus the choices listed in DESIGN.md.
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/build.py).
 * -ffp-contract=off is REQUIRED: results are defined as uncontracted IEEE fp32 in the
 * written operand order, fmaf() only where written.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/roipath_math.h"

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* ROIAlign (legacy, non-aligned)                                             */
/* ------------------------------------------------------------------------- */

typedef struct {
  int valid; /* 0: coordinate outside [-1, extent] -> sample contributes 0 */
  int lo, hi;
  float l, h; /* weight of hi tap, weight of lo tap */
} orc_tap;

/* One axis of the sampling lattice of one RoI.
 * det/layers/roi_align.py:156-159 (coordinate), :20 (range test), :25-45 (clamp),
 * :47-49 (weights).  Same for the backward text :243-246, :71-103. */
static void orc_axis_taps(float start, float bin, int pooled, int grid, int extent,
                          orc_tap* t) {
  for (int p = 0; p < pooled; ++p) {
    for (int i = 0; i < grid; ++i) {
      float v = start + (float)p * bin + ((float)i + .5f) * bin / (float)grid;
      orc_tap* o = &t[p * grid + i];
      if (v < -1.0f || v > (float)extent) {
        o->valid = 0;
        o->lo = o->hi = -1;
        o->l = o->h = 0.f;
        continue;
      }
      if (v <= 0) v = 0;
      int lo = (int)v, hi;
      if (lo >= extent - 1) {
        hi = lo = extent - 1;
        v = (float)lo;
      } else {
        hi = lo + 1;
      }
      float l = v - (float)lo;
      o->valid = 1;
      o->lo = lo;
      o->hi = hi;
      o->l = l;
      o->h = 1.0f - l;
    }
  }
}

/* RoI geometry of the FORWARD kernel: width/height are promoted to double by
 * max(float, 1.) (det/layers/roi_align.py:139-140) and the adaptive grid is
 * ceil(double / int) (:147-148). */
static void orc_roi_geom_fwd(const float* roi, float scale, float sampling_ratio, int PH,
                             int PW, float* sw, float* sh, float* bw, float* bh, int* gw,
                             int* gh) {
  float roi_start_w = roi[1] * scale;
  float roi_start_h = roi[2] * scale;
  float roi_end_w = roi[3] * scale;
  float roi_end_h = roi[4] * scale;
  double roi_width = (double)(roi_end_w - roi_start_w);
  roi_width = roi_width > 1. ? roi_width : 1.; /* max(float, 1.) -> double */
  double roi_height = (double)(roi_end_h - roi_start_h);
  roi_height = roi_height > 1. ? roi_height : 1.;
  *bh = (float)roi_height / (float)PH;
  *bw = (float)roi_width / (float)PW;
  *gh = (sampling_ratio > 0) ? (int)(double)sampling_ratio : (int)ceil(roi_height / PH);
  *gw = (sampling_ratio > 0) ? (int)(double)sampling_ratio : (int)ceil(roi_width / PW);
  *sw = roi_start_w;
  *sh = roi_start_h;
}

/* RoI geometry of the BACKWARD kernel: all float (det/layers/roi_align.py:226-229,
 * :238-239). */
static void orc_roi_geom_bwd(const float* roi, float scale, float sampling_ratio, int PH,
                             int PW, float* sw, float* sh, float* bw, float* bh, int* gw,
                             int* gh) {
  float roi_start_w = roi[1] * scale;
  float roi_start_h = roi[2] * scale;
  float roi_end_w = roi[3] * scale;
  float roi_end_h = roi[4] * scale;
  float roi_width = roi_end_w - roi_start_w;
  roi_width = roi_width > 1.f ? roi_width : 1.f; /* max(float, (float)1.) */
  float roi_height = roi_end_h - roi_start_h;
  roi_height = roi_height > 1.f ? roi_height : 1.f;
  *bh = roi_height / (float)PH;
  *bw = roi_width / (float)PW;
  *gh = (sampling_ratio > 0) ? (int)sampling_ratio : (int)ceilf(roi_height / (float)PH);
  *gw = (sampling_ratio > 0) ? (int)sampling_ratio : (int)ceilf(roi_width / (float)PW);
  *sw = roi_start_w;
  *sh = roi_start_h;
}

/* det/layers/roi_align.py:116-169 (RoIAlignForward) + :14-60 (bilinear_interpolate).
 * feat [N,C,H,W], rois [K,5] = (batch_idx,x1,y1,x2,y2), out [K,C,PH,PW]. */
ORC_API void orc_roi_align_forward(const float* feat, int N, int C, int H, int W,
                                   const float* rois, int K, int PH, int PW, float scale,
                                   float sampling_ratio, float* out) {
  (void)N;
#pragma omp parallel for schedule(dynamic, 1)
  for (int k = 0; k < K; ++k) {
    const float* roi = rois + (size_t)k * 5;
    int b = (int)roi[0];
    float sw, sh, bw, bh;
    int gw, gh;
    orc_roi_geom_fwd(roi, scale, sampling_ratio, PH, PW, &sw, &sh, &bw, &bh, &gw, &gh);
    if (gw < 0) gw = 0;
    if (gh < 0) gh = 0;
    orc_tap* ty = (orc_tap*)malloc(sizeof(orc_tap) * (size_t)(PH * gh + 1));
    orc_tap* tx = (orc_tap*)malloc(sizeof(orc_tap) * (size_t)(PW * gw + 1));
    orc_axis_taps(sh, bh, PH, gh, H, ty);
    orc_axis_taps(sw, bw, PW, gw, W, tx);
    const float count = (float)(gh * gw);
    for (int c = 0; c < C; ++c) {
      const float* plane = feat + ((size_t)b * C + c) * H * W;
      float* o = out + ((size_t)k * C + c) * PH * PW;
      for (int ph = 0; ph < PH; ++ph) {
        for (int pw = 0; pw < PW; ++pw) {
          float acc = 0.f;
          for (int iy = 0; iy < gh; ++iy) {
            const orc_tap* y = &ty[ph * gh + iy];
            for (int ix = 0; ix < gw; ++ix) {
              const orc_tap* x = &tx[pw * gw + ix];
              float val = 0.f;
              if (y->valid && x->valid) {
                float v1 = plane[y->lo * W + x->lo];
                float v2 = plane[y->lo * W + x->hi];
                float v3 = plane[y->hi * W + x->lo];
                float v4 = plane[y->hi * W + x->hi];
                float w1 = y->h * x->h, w2 = y->h * x->l, w3 = y->l * x->h,
                      w4 = y->l * x->l;
                val = w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4;
              }
              acc += val;
            }
          }
          o[ph * PW + pw] = acc / count;
        }
      }
    }
    free(ty);
    free(tx);
  }
}

/* det/layers/roi_align.py:190-274 (RoIAlignBackwardFeature) + :62-111.
 * grad_input [N,C,H,W] is zero-filled here (reference: cudaMemsetAsync :285).
 * Accumulation order = ascending flat output index, the order a serial run of the
 * reference loop produces (the reference's atomicAdd order is unspecified, so the
 * CUDA path is compared with a tolerance, not bitwise). */
ORC_API void orc_roi_align_backward(const float* dout, const float* rois, int K, int N,
                                    int C, int H, int W, int PH, int PW, float scale,
                                    float sampling_ratio, float* grad_input) {
  memset(grad_input, 0, sizeof(float) * (size_t)N * C * H * W);
  /* parallel over channels: every (n,c) plane is owned by one thread => deterministic */
#pragma omp parallel for schedule(static)
  for (int c = 0; c < C; ++c) {
    orc_tap* ty = NULL;
    orc_tap* tx = NULL;
    size_t cap_y = 0, cap_x = 0;
    for (int k = 0; k < K; ++k) {
      const float* roi = rois + (size_t)k * 5;
      int b = (int)roi[0];
      float sw, sh, bw, bh;
      int gw, gh;
      orc_roi_geom_bwd(roi, scale, sampling_ratio, PH, PW, &sw, &sh, &bw, &bh, &gw, &gh);
      if (gw < 0) gw = 0;
      if (gh < 0) gh = 0;
      if ((size_t)(PH * gh + 1) > cap_y) {
        cap_y = (size_t)(PH * gh + 1);
        ty = (orc_tap*)realloc(ty, sizeof(orc_tap) * cap_y);
      }
      if ((size_t)(PW * gw + 1) > cap_x) {
        cap_x = (size_t)(PW * gw + 1);
        tx = (orc_tap*)realloc(tx, sizeof(orc_tap) * cap_x);
      }
      orc_axis_taps(sh, bh, PH, gh, H, ty);
      orc_axis_taps(sw, bw, PW, gw, W, tx);
      const float count = (float)(gh * gw);
      float* plane = grad_input + ((size_t)b * C + c) * H * W;
      const float* g = dout + ((size_t)k * C + c) * PH * PW;
      for (int ph = 0; ph < PH; ++ph) {
        for (int pw = 0; pw < PW; ++pw) {
          float top = g[ph * PW + pw];
          for (int iy = 0; iy < gh; ++iy) {
            const orc_tap* y = &ty[ph * gh + iy];
            for (int ix = 0; ix < gw; ++ix) {
              const orc_tap* x = &tx[pw * gw + ix];
              if (!(y->valid && x->valid)) continue;
              float w1 = y->h * x->h, w2 = y->h * x->l, w3 = y->l * x->h, w4 = y->l * x->l;
              plane[y->lo * W + x->lo] += top * w1 / count;
              plane[y->lo * W + x->hi] += top * w2 / count;
              plane[y->hi * W + x->lo] += top * w3 / count;
              plane[y->hi * W + x->hi] += top * w4 / count;
            }
          }
        }
      }
    }
    free(ty);
    free(tx);
  }
}

/* Number of DISTINCT feature elements (n,c,y,x) read by a forward call: the "U" of
 * SURVEY.md 8(d)'s algorithmic-bytes formula.  All channels of an image touch the
 * same (y,x) set, so U = C * sum over images of |union of taps|. */
ORC_API long long orc_roi_align_touched(int N, int C, int H, int W, const float* rois, int K,
                                        int PH, int PW, float scale, float sampling_ratio) {
  unsigned char* mark = (unsigned char*)calloc((size_t)N * H * W, 1);
  for (int k = 0; k < K; ++k) {
    const float* roi = rois + (size_t)k * 5;
    int b = (int)roi[0];
    if (b < 0 || b >= N) continue;
    float sw, sh, bw, bh;
    int gw, gh;
    orc_roi_geom_fwd(roi, scale, sampling_ratio, PH, PW, &sw, &sh, &bw, &bh, &gw, &gh);
    if (gw < 0) gw = 0;
    if (gh < 0) gh = 0;
    orc_tap* ty = (orc_tap*)malloc(sizeof(orc_tap) * (size_t)(PH * gh + 1));
    orc_tap* tx = (orc_tap*)malloc(sizeof(orc_tap) * (size_t)(PW * gw + 1));
    orc_axis_taps(sh, bh, PH, gh, H, ty);
    orc_axis_taps(sw, bw, PW, gw, W, tx);
    unsigned char* m = mark + (size_t)b * H * W;
    for (int iy = 0; iy < PH * gh; ++iy) {
      if (!ty[iy].valid) continue;
      for (int ix = 0; ix < PW * gw; ++ix) {
        if (!tx[ix].valid) continue;
        m[ty[iy].lo * W + tx[ix].lo] = 1;
        m[ty[iy].lo * W + tx[ix].hi] = 1;
        m[ty[iy].hi * W + tx[ix].lo] = 1;
        m[ty[iy].hi * W + tx[ix].hi] = 1;
      }
    }
    free(ty);
    free(tx);
  }
  long long u = 0;
  for (size_t i = 0; i < (size_t)N * H * W; ++i) u += mark[i];
  free(mark);
  return u * (long long)C;
}

/* ------------------------------------------------------------------------- */
/* FPN level mapping + Pooler                                                 */
/* ------------------------------------------------------------------------- */

/* det/modeling/poolers.py:31-43 (LevelMapper.__call__), area with +1
 * (det/structures/bounding_box.py:292-304).  boxes [K,4] xyxy. */
ORC_API void orc_level_map(const float* boxes, int K, int k_min, int k_max, float s0, int lvl0,
                           float eps, int* levels) {
  for (int k = 0; k < K; ++k) {
    const float* b = boxes + (size_t)k * 4;
    levels[k] = rp_fpn_level(b[0], b[1], b[2], b[3], k_min, k_max, s0, lvl0, eps);
  }
}

/* det/modeling/poolers.py:62-73 (LevelMapperwithArea.__call__). img_area [K]. */
ORC_API void orc_level_map_area(const float* boxes, const float* img_area, int K, int k_min,
                                int k_max, float eps, int* levels) {
  for (int k = 0; k < K; ++k) {
    const float* b = boxes + (size_t)k * 4;
    levels[k] = rp_fpn_level_area(b[0], b[1], b[2], b[3], img_area[k], k_min, k_max, eps);
  }
}

/* det/modeling/poolers.py:123-156 (Pooler.execute): every RoI is pooled from the map of
 * its level and written to its ORIGINAL row.  rois [K,5]; levels [K] in [0,L). */
ORC_API void orc_pooler_forward(const float* const* feats, const int* H, const int* W,
                                const float* scales, int L, int N, int C, const float* rois,
                                const int* levels, int K, int PH, int PW, float sampling_ratio,
                                float* out) {
  memset(out, 0, sizeof(float) * (size_t)K * C * PH * PW);
  for (int k = 0; k < K; ++k) {
    int l = levels[k];
    if (l < 0 || l >= L) continue;
    orc_roi_align_forward(feats[l], N, C, H[l], W[l], rois + (size_t)k * 5, 1, PH, PW,
                          scales[l], sampling_ratio, out + (size_t)k * C * PH * PW);
  }
}

/* Adjoint of Pooler.execute w.r.t. the feature maps (Jittor autograd of
 * det/modeling/poolers.py:148-155 through roi_align.py:189-296). grads[l] is
 * zero-filled and accumulated over the RoIs of level l in ascending RoI order. */
ORC_API void orc_pooler_backward(const float* dout, const float* rois, const int* levels, int K,
                                 const int* H, const int* W, const float* scales, int L, int N,
                                 int C, int PH, int PW, float sampling_ratio,
                                 float* const* grads) {
  for (int l = 0; l < L; ++l) {
    int n_l = 0;
    for (int k = 0; k < K; ++k) n_l += (levels[k] == l);
    float* r = (float*)malloc(sizeof(float) * 5 * (size_t)(n_l + 1));
    float* d = (float*)malloc(sizeof(float) * (size_t)(n_l + 1) * C * PH * PW);
    int j = 0;
    for (int k = 0; k < K; ++k) {
      if (levels[k] != l) continue;
      memcpy(r + (size_t)j * 5, rois + (size_t)k * 5, sizeof(float) * 5);
      memcpy(d + (size_t)j * C * PH * PW, dout + (size_t)k * C * PH * PW,
             sizeof(float) * (size_t)C * PH * PW);
      ++j;
    }
    orc_roi_align_backward(d, r, n_l, N, C, H[l], W[l], PH, PW, scales[l], sampling_ratio,
                           grads[l]);
    free(r);
    free(d);
  }
}

/* ------------------------------------------------------------------------- */
/* Greedy NMS family                                                          */
/* ------------------------------------------------------------------------- */

typedef struct {
  float key;
  int idx;
} orc_kv;

static int orc_cmp_desc_stable(const void* a, const void* b) {
  const orc_kv* x = (const orc_kv*)a;
  const orc_kv* y = (const orc_kv*)b;
  if (x->key > y->key) return -1;
  if (x->key < y->key) return 1;
  return (x->idx > y->idx) - (x->idx < y->idx);
}

static int orc_cmp_asc_stable(const void* a, const void* b) {
  const orc_kv* x = (const orc_kv*)a;
  const orc_kv* y = (const orc_kv*)b;
  if (x->key < y->key) return -1;
  if (x->key > y->key) return 1;
  return (x->idx > y->idx) - (x->idx < y->idx);
}

/* Stable descending argsort (jt.argsort(descending=True), det/layers/nms.py:25;
 * equal scores keep ascending original index -- DESIGN.md choice (1)). */
ORC_API void orc_argsort_desc(const float* key, int n, int* order) {
  orc_kv* kv = (orc_kv*)malloc(sizeof(orc_kv) * (size_t)(n + 1));
  for (int i = 0; i < n; ++i) {
    kv[i].key = key[i];
    kv[i].idx = i;
  }
  qsort(kv, (size_t)n, sizeof(orc_kv), orc_cmp_desc_stable);
  for (int i = 0; i < n; ++i) order[i] = kv[i].idx;
  free(kv);
}

/* det/layers/nms.py:12-14 (nms -> jittor.nms) and :18-40 (_ml_nms / ml_nms):
 *   order = argsort(score, descending); candidate i (in that order) is kept unless an
 *   already-kept j has  iou(+1) > thr  [&& label_i == label_j].
 * The threshold is pasted into the generated C as a double literal (nms.py:23,34), so the
 * comparison is (double)iou > thr  <=>  iou > rp_float_le(thr).
 * boxes [n,4] xyxy, scores [n], labels [n] or NULL.  keep receives original indices in
 * descending-score order; returns their number.  (The caller's keep[:max_proposals]
 * truncation of det/structures/boxlist_ops.py:31-32 is applied via max_keep > 0.) */
ORC_API int orc_nms(const float* boxes, const float* scores, int n, double thr,
                    const float* labels, int max_keep, int* keep) {
  if (n <= 0) return 0;
  int* order = (int*)malloc(sizeof(int) * (size_t)n);
  float* area = (float*)malloc(sizeof(float) * (size_t)n);
  orc_argsort_desc(scores, n, order);
  for (int i = 0; i < n; ++i) {
    const float* b = boxes + (size_t)i * 4;
    area[i] = rp_area_plus1(b[0], b[1], b[2], b[3]);
  }
  const float tf = rp_float_le(thr);
  int m = 0;
  for (int r = 0; r < n; ++r) {
    int i = order[r];
    const float* bi = boxes + (size_t)i * 4;
    int ok = 1;
    for (int q = 0; q < m; ++q) {
      int j = keep[q];
      if (labels && !(labels[i] == labels[j])) continue;
      const float* bj = boxes + (size_t)j * 4;
      float iou =
          rp_iou_plus1(bj[0], bj[1], bj[2], bj[3], area[j], bi[0], bi[1], bi[2], bi[3], area[i]);
      if (iou > tf) {
        ok = 0;
        break;
      }
    }
    if (ok) {
      keep[m++] = i;
      if (max_keep > 0 && m >= max_keep) break;
    }
  }
  free(order);
  free(area);
  return m;
}

/* yol/utils/cython_nms.pyx:24-74: CPU greedy NMS, +1 areas, suppress when ovr >= thresh
 * (float compare), order = scores.argsort()[::-1] (restated as: stable ascending argsort,
 * reversed -- numpy's default sort is unstable, so tie order is this repo's choice),
 * returns ASCENDING original indices of the survivors.  dets [n,5]. */
ORC_API int orc_cython_nms(const float* dets, int n, float thresh, int* keep) {
  if (n <= 0) return 0;
  orc_kv* kv = (orc_kv*)malloc(sizeof(orc_kv) * (size_t)n);
  float* area = (float*)malloc(sizeof(float) * (size_t)n);
  unsigned char* sup = (unsigned char*)calloc((size_t)n, 1);
  for (int i = 0; i < n; ++i) {
    const float* d = dets + (size_t)i * 5;
    kv[i].key = d[4];
    kv[i].idx = i;
    area[i] = rp_area_plus1(d[0], d[1], d[2], d[3]);
  }
  qsort(kv, (size_t)n, sizeof(orc_kv), orc_cmp_asc_stable);
  for (int a = n - 1; a >= 0; --a) {
    int i = kv[a].idx;
    if (sup[i]) continue;
    const float* di = dets + (size_t)i * 5;
    for (int b2 = a - 1; b2 >= 0; --b2) {
      int j = kv[b2].idx;
      if (sup[j]) continue;
      const float* dj = dets + (size_t)j * 5;
      float ovr =
          rp_iou_plus1(di[0], di[1], di[2], di[3], area[i], dj[0], dj[1], dj[2], dj[3], area[j]);
      if (ovr >= thresh) sup[j] = 1;
    }
  }
  int m = 0;
  for (int i = 0; i < n; ++i)
    if (!sup[i]) keep[m++] = i;
  free(kv);
  free(area);
  free(sup);
  return m;
}

/* ------------------------------------------------------------------------- */
/* Yolact: decode, jaccard, FastNMS                                           */
/* ------------------------------------------------------------------------- */

/* yol/layers/box_utils.py:272-316 (decode, SSD regressors, variances .1/.2):
 *   c = p.xy + loc.xy*0.1*p.wh ; wh = p.wh*exp(loc.wh*0.2) ; x1y1 = c - wh/2 ; x2y2 = wh + x1y1
 * exp is rp_expf (DESIGN.md choice (4)). loc, priors, out: [n,4]. */
ORC_API void orc_yolact_decode(const float* loc, const float* priors, int n, float* out) {
  for (int i = 0; i < n; ++i) {
    const float* l = loc + (size_t)i * 4;
    const float* p = priors + (size_t)i * 4;
    float cx = p[0] + l[0] * 0.1f * p[2];
    float cy = p[1] + l[1] * 0.1f * p[3];
    float w = p[2] * rp_expf(l[2] * 0.2f);
    float h = p[3] * rp_expf(l[3] * 0.2f);
    float x1 = cx - w / 2.0f;
    float y1 = cy - h / 2.0f;
    out[(size_t)i * 4 + 0] = x1;
    out[(size_t)i * 4 + 1] = y1;
    out[(size_t)i * 4 + 2] = w + x1;
    out[(size_t)i * 4 + 3] = h + y1;
  }
}

/* yol/layers/box_utils.py:52-78 (jaccard, iscrowd=False). a [A,4], b [B,4] -> out [A,B]. */
ORC_API void orc_jaccard(const float* a, int A, const float* b, int B, float* out) {
  for (int i = 0; i < A; ++i) {
    const float* p = a + (size_t)i * 4;
    float pa = rp_area_rel(p[0], p[1], p[2], p[3]);
    for (int j = 0; j < B; ++j) {
      const float* q = b + (size_t)j * 4;
      float qa = rp_area_rel(q[0], q[1], q[2], q[3]);
      out[(size_t)i * B + j] = rp_jaccard(p[0], p[1], p[2], p[3], pa, q[0], q[1], q[2], q[3], qa);
    }
  }
}

/* column max of triu(iou, 1) with NaN propagation (DESIGN.md choice (6)):
 * iou_max[j] = max(0, max_{i<j} iou[i][j]); NaN if any term is NaN. */
static void orc_fast_nms_colmax(const float* bx, int k, float* iou_max) {
  float* area = (float*)malloc(sizeof(float) * (size_t)(k + 1));
  for (int i = 0; i < k; ++i)
    area[i] = rp_area_rel(bx[i * 4 + 0], bx[i * 4 + 1], bx[i * 4 + 2], bx[i * 4 + 3]);
  for (int j = 0; j < k; ++j) {
    float m = 0.f;
    int nan = 0;
    for (int i = 0; i < j; ++i) {
      float v = rp_jaccard(bx[i * 4 + 0], bx[i * 4 + 1], bx[i * 4 + 2], bx[i * 4 + 3], area[i],
                           bx[j * 4 + 0], bx[j * 4 + 1], bx[j * 4 + 2], bx[j * 4 + 3], area[j]);
      if (v != v) nan = 1;
      if (v > m) m = v;
    }
    iou_max[j] = nan ? NAN : m;
  }
  free(area);
}

/* yol/layers/functions/detection.py:139-189 (Detect.fast_nms):
 *   per class: stable desc sort, top_k; jaccard; triu(1); column max; keep iou_max <= thr
 *   [and score > conf_thresh if second_threshold]; flatten kept in (class, rank) order;
 *   stable desc sort by score; first max_det.
 * boxes [n,4], coeff [n,D], scores [C,n].  Outputs sized max_det. Returns m. */
ORC_API int orc_fast_nms(const float* boxes, const float* coeff, const float* scores, int n, int C,
                         int D, float iou_thr, int top_k, int max_det, int second_threshold,
                         float conf_thresh, float* out_boxes, float* out_coeff, int* out_classes,
                         float* out_scores, int* out_index) {
  int k = n < top_k ? n : top_k;
  if (k <= 0 || C <= 0) return 0;
  int* order = (int*)malloc(sizeof(int) * (size_t)n);
  float* bx = (float*)malloc(sizeof(float) * 4 * (size_t)k);
  float* iou_max = (float*)malloc(sizeof(float) * (size_t)k);
  orc_kv* cand = (orc_kv*)malloc(sizeof(orc_kv) * (size_t)C * k);
  int* cand_cls = (int*)malloc(sizeof(int) * (size_t)C * k);
  int* cand_src = (int*)malloc(sizeof(int) * (size_t)C * k);
  int nc = 0;
  for (int c = 0; c < C; ++c) {
    const float* sc = scores + (size_t)c * n;
    orc_argsort_desc(sc, n, order);
    for (int r = 0; r < k; ++r) memcpy(bx + r * 4, boxes + (size_t)order[r] * 4, sizeof(float) * 4);
    orc_fast_nms_colmax(bx, k, iou_max);
    for (int r = 0; r < k; ++r) {
      int keep = iou_max[r] <= iou_thr;
      if (second_threshold) keep = keep && (sc[order[r]] > conf_thresh);
      if (!keep) continue;
      cand[nc].key = sc[order[r]];
      cand[nc].idx = nc;
      cand_cls[nc] = c;
      cand_src[nc] = order[r];
      ++nc;
    }
  }
  qsort(cand, (size_t)nc, sizeof(orc_kv), orc_cmp_desc_stable);
  int m = nc < max_det ? nc : max_det;
  for (int r = 0; r < m; ++r) {
    int q = cand[r].idx;
    int src = cand_src[q];
    memcpy(out_boxes + (size_t)r * 4, boxes + (size_t)src * 4, sizeof(float) * 4);
    memcpy(out_coeff + (size_t)r * D, coeff + (size_t)src * D, sizeof(float) * (size_t)D);
    out_classes[r] = cand_cls[q];
    out_scores[r] = cand[r].key;
    if (out_index) out_index[r] = src;
  }
  free(order);
  free(bx);
  free(iou_max);
  free(cand);
  free(cand_cls);
  free(cand_src);
  return m;
}

/* yol/layers/functions/detection.py:113-137 (Detect.cc_fast_nms): classes collapsed by
 * argmax (first maximum on ties), one sort, top_k, column max, keep <= thr.  Outputs are
 * in descending-score order, at most top_k of them. */
ORC_API int orc_cc_fast_nms(const float* boxes, const float* coeff, const float* scores, int n,
                            int C, int D, float iou_thr, int top_k, float* out_boxes,
                            float* out_coeff, int* out_classes, float* out_scores,
                            int* out_index) {
  if (n <= 0 || C <= 0) return 0;
  float* best = (float*)malloc(sizeof(float) * (size_t)n);
  int* cls = (int*)malloc(sizeof(int) * (size_t)n);
  for (int i = 0; i < n; ++i) {
    float b = scores[i];
    int bc = 0;
    for (int c = 1; c < C; ++c) {
      float v = scores[(size_t)c * n + i];
      if (v > b) {
        b = v;
        bc = c;
      }
    }
    best[i] = b;
    cls[i] = bc;
  }
  int* order = (int*)malloc(sizeof(int) * (size_t)n);
  orc_argsort_desc(best, n, order);
  int k = n < top_k ? n : top_k;
  float* bx = (float*)malloc(sizeof(float) * 4 * (size_t)k);
  float* iou_max = (float*)malloc(sizeof(float) * (size_t)k);
  for (int r = 0; r < k; ++r) memcpy(bx + r * 4, boxes + (size_t)order[r] * 4, sizeof(float) * 4);
  orc_fast_nms_colmax(bx, k, iou_max);
  int m = 0;
  for (int r = 0; r < k; ++r) {
    if (!(iou_max[r] <= iou_thr)) continue;
    int src = order[r];
    memcpy(out_boxes + (size_t)m * 4, boxes + (size_t)src * 4, sizeof(float) * 4);
    memcpy(out_coeff + (size_t)m * D, coeff + (size_t)src * D, sizeof(float) * (size_t)D);
    out_classes[m] = cls[src];
    out_scores[m] = best[src];
    if (out_index) out_index[m] = src;
    ++m;
  }
  free(best);
  free(cls);
  free(order);
  free(bx);
  free(iou_max);
  return m;
}

/* yol/layers/functions/detection.py:81-99 (Detect.detect, 'Slices'): class-max over the
 * foreground rows of conf [C+1? no: rows 1..], keep priors with max > conf_thresh, in prior
 * order.  conf_fg [C,P] (row 0 = class 1).  Writes kept prior indices; returns their count. */
ORC_API int orc_detect_filter(const float* conf_fg, int C, int P, float conf_thresh,
                              int* keep_idx) {
  int m = 0;
  for (int p = 0; p < P; ++p) {
    float b = conf_fg[p];
    for (int c = 1; c < C; ++c) {
      float v = conf_fg[(size_t)c * P + p];
      if (v > b) b = v;
    }
    if (b > conf_thresh) keep_idx[m++] = p;
  }
  return m;
}

/* ------------------------------------------------------------------------- */
/* Bilinear resize, mask assembly, paste-back                                 */
/* ------------------------------------------------------------------------- */

/* One bilinear sample of a [h,w] image at output pixel (Y,X) of an [H,W] resize,
 * align_corners=False, half-pixel centres, indices clamped (DESIGN.md choice (2)):
 *   out = h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11). */
static inline float orc_bilinear_at(const float* img, int h, int w, float sy, float sx, int Y,
                                    int X) {
  float fy = rp_resize_src(Y, sy);
  float fx = rp_resize_src(X, sx);
  int y0 = (int)fy, x0 = (int)fx;
  if (y0 > h - 1) y0 = h - 1;
  if (x0 > w - 1) x0 = w - 1;
  int y1 = y0 + (y0 < h - 1 ? 1 : 0);
  int x1 = x0 + (x0 < w - 1 ? 1 : 0);
  float h1 = fy - (float)y0, h0 = 1.0f - h1;
  float w1 = fx - (float)x0, w0 = 1.0f - w1;
  float top = w0 * img[y0 * w + x0] + w1 * img[y0 * w + x1];
  float bot = w0 * img[y1 * w + x0] + w1 * img[y1 * w + x1];
  return h0 * top + h1 * bot;
}

/* jittor nn.interpolate(mode='bilinear', align_corners=False) as used at
 * yol/layers/output_utils.py:93 and det/.../mask_head/inference.py:140. in [h,w] -> out [H,W]. */
ORC_API void orc_bilinear_resize(const float* in, int h, int w, float* out, int H, int W) {
  float sy = (float)h / (float)H, sx = (float)w / (float)W;
  for (int Y = 0; Y < H; ++Y)
    for (int X = 0; X < W; ++X) out[(size_t)Y * W + X] = orc_bilinear_at(in, h, w, sy, sx, Y, X);
}

/* yol/layers/box_utils.py:333-351 (sanitize_coordinates, cast=False). */
static inline void orc_sanitize(float a, float b, int img_size, int padding, float* lo,
                                float* hi) {
  float x1 = a * (float)img_size;
  float x2 = b * (float)img_size;
  float mn = rp_fminf(x1, x2);
  float mx = rp_fmaxf(x1, x2);
  mn = mn - (float)padding;
  if (mn < 0.f) mn = 0.f;
  mx = mx + (float)padding;
  if (mx > (float)img_size) mx = (float)img_size;
  *lo = mn;
  *hi = mx;
}

/* yol/layers/output_utils.py:60-101 (postprocess, lincomb branch) with
 * yol/layers/box_utils.py:355-378 (crop):
 *   low[y][x] = sigmoid( fma-chain_k proto[y][x][k]*coeff[j][k] )  (k ascending, acc0 = 0)
 *   crop: low *= (x >= x1 && x < x2 && y >= y1 && y < y2), box padded by 1 px
 *   up = bilinear resize to (out_h,out_w) ; mask = up > 0.5
 * and :99-101: boxes -> absolute, sorted, clamped, int32 truncation.
 * proto [ph,pw,D]; coeff [m,D]; boxes [m,4] relative xyxy;
 * out_masks [m,out_h,out_w] u8 ; out_boxes [m,4] int32.
 * low_out (optional, [m,ph,pw]) receives the cropped low-resolution probabilities. */
ORC_API void orc_mask_assemble(const float* proto, const float* coeff, const float* boxes, int m,
                               int ph, int pw, int D, int out_h, int out_w, int crop,
                               uint8_t* out_masks, int32_t* out_boxes, float* low_out) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int j = 0; j < m; ++j) {
    float* low = (float*)malloc(sizeof(float) * (size_t)ph * pw);
    const float* cf = coeff + (size_t)j * D;
    const float* b = boxes + (size_t)j * 4;
    float x1, x2, y1, y2;
    orc_sanitize(b[0], b[2], pw, 1, &x1, &x2);
    orc_sanitize(b[1], b[3], ph, 1, &y1, &y2);
    for (int y = 0; y < ph; ++y) {
      for (int x = 0; x < pw; ++x) {
        const float* p = proto + ((size_t)y * pw + x) * D;
        float acc = 0.f;
        for (int k = 0; k < D; ++k) acc = fmaf(p[k], cf[k], acc);
        float v = rp_sigmoidf(acc);
        if (crop) {
          int in = ((float)x >= x1) && ((float)x < x2) && ((float)y >= y1) && ((float)y < y2);
          v = v * (in ? 1.0f : 0.0f);
        }
        low[y * pw + x] = v;
      }
    }
    if (low_out) memcpy(low_out + (size_t)j * ph * pw, low, sizeof(float) * (size_t)ph * pw);
    float sy = (float)ph / (float)out_h, sx = (float)pw / (float)out_w;
    uint8_t* o = out_masks + (size_t)j * out_h * out_w;
    for (int Y = 0; Y < out_h; ++Y)
      for (int X = 0; X < out_w; ++X)
        o[(size_t)Y * out_w + X] = orc_bilinear_at(low, ph, pw, sy, sx, Y, X) > 0.5f;
    if (out_boxes) {
      float ax1, ax2, ay1, ay2;
      orc_sanitize(b[0], b[2], out_w, 0, &ax1, &ax2);
      orc_sanitize(b[1], b[3], out_h, 0, &ay1, &ay2);
      out_boxes[j * 4 + 0] = (int32_t)ax1;
      out_boxes[j * 4 + 1] = (int32_t)ay1;
      out_boxes[j * 4 + 2] = (int32_t)ax2;
      out_boxes[j * 4 + 3] = (int32_t)ay2;
    }
    free(low);
  }
}

/* det/modeling/roi_heads/mask_head/inference.py:38-45 (MaskPostProcessor.execute):
 * prob[i] = sigmoid(logits[i, labels[i]]).  logits [K,NC,M,M] -> out [K,M,M]. */
ORC_API void orc_mask_select_sigmoid(const float* logits, const int* labels, int K, int NC, int M,
                                     float* out) {
  for (int i = 0; i < K; ++i) {
    const float* src = logits + ((size_t)i * NC + labels[i]) * M * M;
    for (int t = 0; t < M * M; ++t) out[(size_t)i * M * M + t] = rp_sigmoidf(src[t]);
  }
}

/* det/modeling/roi_heads/mask_head/inference.py:90-162
 * (expand_boxes, expand_masks, paste_mask_in_image) and :175-189 (Masker.forward_single_image):
 *   padded = zero-pad(mask, padding); scale = (M+2p)/M
 *   box' = centre +- half*scale ; int32 truncation ; w = x2-x1+1, h = y2-y1+1, >= 1
 *   r = bilinear resize of padded to (h,w) ; r > thresh   (thresh < 0: (r*255) != 0)
 *   paste into [im_h,im_w] at the box, clipped to the image.
 * A box that lies entirely outside the image yields an empty mask (the reference's python
 * slicing is ill-defined there).
 * masks [n,M,M] probabilities; boxes [n,4] xyxy; out [n,im_h,im_w] u8. */
ORC_API void orc_paste_masks(const float* masks, const float* boxes, int n, int M, int im_h,
                             int im_w, float thresh, int padding, uint8_t* out) {
  int P = M + 2 * padding;
  float scale = (float)((double)P / (double)M);
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < n; ++i) {
    uint8_t* o = out + (size_t)i * im_h * im_w;
    memset(o, 0, (size_t)im_h * im_w);
    float* pad = (float*)calloc((size_t)P * P, sizeof(float));
    for (int y = 0; y < M; ++y)
      memcpy(pad + (size_t)(y + padding) * P + padding, masks + ((size_t)i * M + y) * M,
             sizeof(float) * (size_t)M);
    const float* b = boxes + (size_t)i * 4;
    float w_half = (b[2] - b[0]) * .5f;
    float h_half = (b[3] - b[1]) * .5f;
    float x_c = (b[2] + b[0]) * .5f;
    float y_c = (b[3] + b[1]) * .5f;
    w_half *= scale;
    h_half *= scale;
    int bx0 = (int)(x_c - w_half), bx1 = (int)(x_c + w_half);
    int by0 = (int)(y_c - h_half), by1 = (int)(y_c + h_half);
    int w = bx1 - bx0 + 1, h = by1 - by0 + 1;
    if (w < 1) w = 1;
    if (h < 1) h = 1;
    int x_0 = bx0 > 0 ? bx0 : 0;
    int x_1 = bx1 + 1 < im_w ? bx1 + 1 : im_w;
    int y_0 = by0 > 0 ? by0 : 0;
    int y_1 = by1 + 1 < im_h ? by1 + 1 : im_h;
    float sy = (float)P / (float)h, sx = (float)P / (float)w;
    for (int Y = y_0; Y < y_1; ++Y) {
      int my = Y - by0;
      if (my < 0 || my >= h) continue;
      for (int X = x_0; X < x_1; ++X) {
        int mx = X - bx0;
        if (mx < 0 || mx >= w) continue;
        float v = orc_bilinear_at(pad, P, P, sy, sx, my, mx);
        o[(size_t)Y * im_w + X] = thresh >= 0 ? (v > thresh) : ((v * 255.0f) != 0.0f);
      }
    }
    free(pad);
  }
}

/* ------------------------------------------------------------------------- */
/* RPN proposal stage in front of the NMS                                     */
/* ------------------------------------------------------------------------- */

/* det/modeling/box_coder.py:53-94 BoxCoder.decode.  rel_codes [n,4K], boxes [n,4], out [n,4K].
 * TO_REMOVE = 1 (:65-67); d* = code / weight (:72-75); dw, dh clamped from above (:77-78);
 * centre and size (:80-83); x2 / y2 carry the "- 1" (:90-93).  exp is rp_expf, the function the
 * CUDA side shares (pinned against the reference's known-answer test at atol 1e-4 and against
 * libm in tests/test_oracle_golden.py). */
ORC_API void orc_box_decode(const float* rel_codes, const float* boxes, int n, int K, float wx,
                            float wy, float ww, float wh, float clip, float* out) {
  for (int i = 0; i < n; ++i) {
    const float* b = boxes + (size_t)i * 4;
    const float widths = b[2] - b[0] + 1.0f;
    const float heights = b[3] - b[1] + 1.0f;
    const float ctr_x = b[0] + 0.5f * widths;
    const float ctr_y = b[1] + 0.5f * heights;
    for (int c = 0; c < K; ++c) {
      const float* d = rel_codes + ((size_t)i * K + c) * 4;
      float* o = out + ((size_t)i * K + c) * 4;
      const float dx = d[0] / wx, dy = d[1] / wy;
      float dw = d[2] / ww, dh = d[3] / wh;
      if (dw > clip) dw = clip;
      if (dh > clip) dh = clip;
      const float pcx = dx * widths + ctr_x;
      const float pcy = dy * heights + ctr_y;
      const float pw = rp_expf(dw) * widths;
      const float ph = rp_expf(dh) * heights;
      o[0] = pcx - 0.5f * pw;
      o[1] = pcy - 0.5f * ph;
      o[2] = pcx + 0.5f * pw - 1.0f;
      o[3] = pcy + 0.5f * ph - 1.0f;
    }
  }
}

/* det/modeling/rpn/inference.py:83-111, one level, up to (not including) boxlist_nms:
 *   permute_and_flatten (rpn/utils.py:10-15): element e = (h*W + w)*A + a of image b is
 *   objectness[b][a][h][w] and box_regression[b][4a .. 4a+3][h][w];
 *   sigmoid (:89); topk(pre_nms_top_n, sorted=True) (:95-96) -- restated as a stable
 *   descending sort, equal scores by ascending e (tie order is this repo's choice, PARITY
 *   UNPINNED for ties); gathers (:100-104); decode (:106-108);
 *   clip_to_image(remove_empty=False) det/structures/bounding_box.py:263-277: clamp to
 *   [0, w-1] x [0, h-1]; remove_small_boxes det/structures/boxlist_ops.py:64-76: keep when
 *   x2-x1+1 >= min_size and y2-y1+1 >= min_size.
 * anchors [N, H*W*A, 4]; im_wh [N,2] (width, height).  Outputs are padded per image:
 * out_boxes [N,k,4], out_scores [N,k], out_index [N,k], counts [N]; k = min(top_n, A*H*W). */
ORC_API void orc_rpn_pre_nms(const float* objectness, const float* box_regression,
                             const float* anchors, const float* im_wh, int N, int A, int H, int W,
                             int pre_nms_top_n, float min_size, float wx, float wy, float ww,
                             float wh, float clip, float* out_boxes, float* out_scores,
                             int* out_index, int* counts) {
  const int HW = H * W;
  const int per = A * HW;
  const int k = pre_nms_top_n < per ? pre_nms_top_n : per;
  float* flat = (float*)malloc(sizeof(float) * (size_t)per);
  int* order = (int*)malloc(sizeof(int) * (size_t)per);
  for (int b = 0; b < N; ++b) {
    const float* obj = objectness + (size_t)b * per;
    const float* reg = box_regression + (size_t)b * per * 4;
    const float* anc = anchors + (size_t)b * per * 4;
    for (int hw = 0; hw < HW; ++hw)
      for (int a = 0; a < A; ++a) flat[hw * A + a] = rp_sigmoidf(obj[(size_t)a * HW + hw]);
    orc_argsort_desc(flat, per, order);
    int m = 0;
    for (int j = 0; j < k; ++j) {
      const int e = order[j];
      const int hw = e / A, a = e % A;
      float code[4], box[4];
      for (int c = 0; c < 4; ++c) code[c] = reg[((size_t)a * 4 + c) * HW + hw];
      orc_box_decode(code, anc + (size_t)e * 4, 1, 1, wx, wy, ww, wh, clip, box);
      const float xmax = im_wh[2 * b] - 1.0f, ymax = im_wh[2 * b + 1] - 1.0f;
      box[0] = fminf(fmaxf(box[0], 0.0f), xmax);
      box[1] = fminf(fmaxf(box[1], 0.0f), ymax);
      box[2] = fminf(fmaxf(box[2], 0.0f), xmax);
      box[3] = fminf(fmaxf(box[3], 0.0f), ymax);
      const float ws = box[2] - box[0] + 1.0f, hs = box[3] - box[1] + 1.0f;
      if (ws >= min_size && hs >= min_size) {
        const size_t o = (size_t)b * k + m;
        memcpy(out_boxes + o * 4, box, sizeof(box));
        out_scores[o] = flat[e];
        out_index[o] = e;
        ++m;
      }
    }
    counts[b] = m;
  }
  free(flat);
  free(order);
}

/* ------------------------------------------------------------------------- */
/* Box-head post-processor                                                    */
/* ------------------------------------------------------------------------- */

/* nn.softmax(class_logits, -1) det/modeling/roi_heads/box_head/inference.py:58, restated as
 * exp(x - max) / sum with the sum taken in class order (PARITY UNPINNED: Jittor's reduction
 * order is not visible in the reference tree). */
ORC_API void orc_softmax_rows(const float* x, int R, int C, float* out) {
  for (int i = 0; i < R; ++i) {
    const float* r = x + (size_t)i * C;
    float* o = out + (size_t)i * C;
    float m = r[0];
    for (int c = 1; c < C; ++c) m = fmaxf(m, r[c]);
    float sum = 0.0f;
    for (int c = 0; c < C; ++c) {
      o[c] = rp_expf(r[c] - m);
      sum += o[c];
    }
    for (int c = 0; c < C; ++c) o[c] = o[c] / sum;
  }
}

static int orc_cmp_float_asc(const void* a, const void* b) {
  const float x = *(const float*)a, y = *(const float*)b;
  return (x > y) - (x < y);
}

/* One image of PostProcessor.execute + filter_results, inference.py:45-90,112-195.
 *   logits [n,C], reg [n,reg_cols], props [n,4]; softmax (:58); decode of every class's codes
 *   with the proposal (:68-70; cls_agnostic: the last four columns for every class :66-67,
 *   71-72); clip_to_image(remove_empty=False) (:80); per class j = 1..C-1 (:126-168):
 *   inds = nonzero(scores[:,j] > score_thresh) in ascending row order, boxlist_nms(thr)
 *   -> kept in descending score order, labels = j; cat over classes (:170); when
 *   number_of_detections > detections_per_img > 0 (:178): image_thresh =
 *   kthvalue(scores, n - dets + 1) = the dets-th largest score, keep = scores >= image_thresh in
 *   their order (:180-187).
 * The score threshold is compared in float32 (PARITY UNPINNED).
 * Outputs sized n*(C-1): out_boxes [.,4], out_scores, out_labels, out_rows; returns the count. */
ORC_API int orc_box_post_image(const float* logits, const float* reg, const float* props, int n, int C,
                               int reg_cols, int cls_agnostic, float im_w, float im_h,
                               float score_thresh, double nms_thresh, int dets, float wx, float wy,
                               float ww, float wh, float clip, float* out_boxes, float* out_scores,
                               int* out_labels, int* out_rows) {
  if (n <= 0) return 0;
  float* prob = (float*)malloc(sizeof(float) * (size_t)n * C);
  float* cb = (float*)malloc(sizeof(float) * 4 * (size_t)n);
  float* cs = (float*)malloc(sizeof(float) * (size_t)n);
  int* crow = (int*)malloc(sizeof(int) * (size_t)n);
  int* keep = (int*)malloc(sizeof(int) * (size_t)n);
  orc_softmax_rows(logits, n, C, prob);
  int m = 0;
  for (int j = 1; j < C; ++j) {
    int nc = 0;
    for (int i = 0; i < n; ++i) {
      const float p = prob[(size_t)i * C + j];
      if (!(p > score_thresh)) continue;
      const float* code = reg + (size_t)i * reg_cols + (cls_agnostic ? reg_cols - 4 : 4 * j);
      float* b = cb + (size_t)nc * 4;
      orc_box_decode(code, props + (size_t)i * 4, 1, 1, wx, wy, ww, wh, clip, b);
      b[0] = fminf(fmaxf(b[0], 0.0f), im_w - 1.0f);
      b[1] = fminf(fmaxf(b[1], 0.0f), im_h - 1.0f);
      b[2] = fminf(fmaxf(b[2], 0.0f), im_w - 1.0f);
      b[3] = fminf(fmaxf(b[3], 0.0f), im_h - 1.0f);
      cs[nc] = p;
      crow[nc] = i;
      ++nc;
    }
    if (nc == 0) continue;
    int nk;
    if (nms_thresh > 0) {
      nk = orc_nms(cb, cs, nc, nms_thresh, NULL, -1, keep);
    } else { /* det/structures/boxlist_ops.py:24-25: the boxlist is returned unchanged */
      nk = nc;
      for (int q = 0; q < nc; ++q) keep[q] = q;
    }
    for (int q = 0; q < nk; ++q) {
      memcpy(out_boxes + (size_t)m * 4, cb + (size_t)keep[q] * 4, sizeof(float) * 4);
      out_scores[m] = cs[keep[q]];
      out_labels[m] = j;
      out_rows[m] = crow[keep[q]];
      ++m;
    }
  }
  if (dets > 0 && m > dets) {
    float* sorted = (float*)malloc(sizeof(float) * (size_t)m);
    memcpy(sorted, out_scores, sizeof(float) * (size_t)m);
    qsort(sorted, (size_t)m, sizeof(float), orc_cmp_float_asc);
    const float image_thresh = sorted[m - dets]; /* kthvalue(k = m - dets + 1), 1-based */
    free(sorted);
    int w = 0;
    for (int q = 0; q < m; ++q) {
      if (out_scores[q] >= image_thresh) {
        if (w != q) {
          memmove(out_boxes + (size_t)w * 4, out_boxes + (size_t)q * 4, sizeof(float) * 4);
          out_scores[w] = out_scores[q];
          out_labels[w] = out_labels[q];
          out_rows[w] = out_rows[q];
        }
        ++w;
      }
    }
    m = w;
  }
  free(prob);
  free(cb);
  free(cs);
  free(crow);
  free(keep);
  return m;
}

/* ------------------------------------------------------------------------- */
/* Other RoI-family operators                                                 */
/* ------------------------------------------------------------------------- */

/* det/structures/boxlist_ops.py:79-116 boxlist_iou: area with TO_REMOVE = 1
 * (bounding_box.py:292-304), wh = clamp(rb - lt + 1, 0), iou = inter / (a1 + a2 - inter). */
ORC_API void orc_boxlist_iou(const float* a, int N, const float* b, int M, float* out) {
  for (int i = 0; i < N; ++i) {
    const float* p = a + (size_t)i * 4;
    const float area1 = (p[2] - p[0] + 1.0f) * (p[3] - p[1] + 1.0f);
    for (int j = 0; j < M; ++j) {
      const float* q = b + (size_t)j * 4;
      const float area2 = (q[2] - q[0] + 1.0f) * (q[3] - q[1] + 1.0f);
      const float w = fmaxf(fminf(p[2], q[2]) - fmaxf(p[0], q[0]) + 1.0f, 0.0f);
      const float h = fmaxf(fminf(p[3], q[3]) - fmaxf(p[1], q[1]) + 1.0f, 0.0f);
      const float inter = w * h;
      out[(size_t)i * M + j] = inter / (area1 + area2 - inter);
    }
  }
}

/* det/layers/roi_pool.py:16-88 RoIPoolForwardKernel: round() of the scaled corners, malformed
 * RoIs forced to 1x1, bin size = (float)(1.0 * extent / pooled) (a double quotient, :55-56),
 * floor / ceil bin edges clipped to the map, empty bin -> 0 / -1, otherwise the first maximum
 * in (h, w) scan order above -99999999. */
ORC_API void orc_roi_pool_forward(const float* feat, int N, int C, int H, int W, const float* rois,
                                  int K, int PH, int PW, float scale, float* out, int* argmax) {
  (void)N;
  for (int n = 0; n < K; ++n) {
    const float* r = rois + (size_t)n * 5;
    const int bi = (int)r[0];
    const int sw = (int)roundf(r[1] * scale), sh = (int)roundf(r[2] * scale);
    const int ew = (int)roundf(r[3] * scale), eh = (int)roundf(r[4] * scale);
    const int rw = ew - sw + 1 > 1 ? ew - sw + 1 : 1, rh = eh - sh + 1 > 1 ? eh - sh + 1 : 1;
    const float bh = (float)(1.0 * rh / PH), bw = (float)(1.0 * rw / PW);
    for (int c = 0; c < C; ++c) {
      const float* plane = feat + ((size_t)bi * C + c) * H * W;
      for (int ph = 0; ph < PH; ++ph)
        for (int pw = 0; pw < PW; ++pw) {
          int hs = (int)floorf(ph * bh), ws = (int)floorf(pw * bw);
          int he = (int)ceilf((ph + 1) * bh), we = (int)ceilf((pw + 1) * bw);
          hs = hs + sh < 0 ? 0 : (hs + sh > H ? H : hs + sh);
          he = he + sh < 0 ? 0 : (he + sh > H ? H : he + sh);
          ws = ws + sw < 0 ? 0 : (ws + sw > W ? W : ws + sw);
          we = we + sw < 0 ? 0 : (we + sw > W ? W : we + sw);
          const int empty = (he <= hs) || (we <= ws);
          float maxval = empty ? 0.0f : -99999999.0f;
          int maxidx = -1;
          for (int h = hs; h < he; ++h)
            for (int w = ws; w < we; ++w)
              if (plane[h * W + w] > maxval) {
                maxval = plane[h * W + w];
                maxidx = h * W + w;
              }
          const size_t o = (((size_t)n * C + c) * PH + ph) * PW + pw;
          out[o] = maxval;
          argmax[o] = maxidx;
        }
    }
  }
}

/* det/layers/roi_pool.py:101-148 RoIPoolBackwardKernel (sequential accumulation order). */
ORC_API void orc_roi_pool_backward(const float* dout, const float* rois, const int* argmax, int K,
                                   int N, int C, int H, int W, int PH, int PW, float* grad) {
  memset(grad, 0, sizeof(float) * (size_t)N * C * H * W);
  for (int n = 0; n < K; ++n) {
    const int bi = (int)rois[(size_t)n * 5];
    for (int c = 0; c < C; ++c)
      for (int q = 0; q < PH * PW; ++q) {
        const size_t o = ((size_t)n * C + c) * PH * PW + q;
        if (argmax[o] != -1) grad[((size_t)bi * C + c) * H * W + argmax[o]] += dout[o];
      }
  }
}

/* det/layers/mask_prob.py:9-57 MaskProbForward: for centre n, its first box_areas[n] pixels in
 * row-major order of the box (x1,y1,x2,y2 end exclusive): p = exp(-sum_d (e_p - e_c)^2 * sigma);
 * the sum in dimension order, uncontracted; exp is rp_expf (PARITY UNPINNED: the reference's
 * expf / FMA contraction are nvcc's). */
ORC_API void orc_mask_prob(const float* ep, const float* ec, const float* sigma, const int* boxes,
                           const int* areas, int num_pixel, int num_center, int dim, int mask_width,
                           float* probs) {
  memset(probs, 0, sizeof(float) * (size_t)num_pixel * num_center);
  for (int n = 0; n < num_center; ++n) {
    const int x1 = boxes[4 * n], y1 = boxes[4 * n + 1], bw = boxes[4 * n + 2] - x1;
    if (bw <= 0) continue;
    for (int q = 0; q < areas[n]; ++q) {
      const int pid = (q / bw + y1) * mask_width + (q % bw + x1);
      if (pid < 0 || pid >= num_pixel) continue;
      float norm2 = 0.0f;
      for (int d = 0; d < dim; ++d) {
        const float t = ep[(size_t)pid * dim + d] - ec[(size_t)n * dim + d];
        norm2 = norm2 + t * t;
      }
      probs[(size_t)n * num_pixel + pid] = rp_expf(-norm2 * sigma[n]);
    }
  }
}

/* det/structures/boxlist_ops.py:237-261 boxes_to_masks (crop = 0) and :262-281 crop_by_box
 * (crop = 1), in the written operand order; indicator [n,h,w] uint8, masks_out = masks_in * ind. */
ORC_API void orc_box_masks(const float* boxes, int n, int h, int w, float padding, int crop,
                           const float* masks_in, float* masks_out, uint8_t* ind) {
  for (int i = 0; i < n; ++i) {
    const float* b = boxes + (size_t)i * 4;
    const float bw = b[2] - b[0], bh = b[3] - b[1];
    float x1, x2, y1, y2;
    if (crop) {
      x1 = fmaxf(b[0] - bw * padding - 1.0f, 0.0f);
      x2 = fminf(b[2] + bw * padding + 1.0f, (float)(w - 1));
      y1 = fmaxf(b[1] - bh * padding - 1.0f, 0.0f);
      y2 = fminf(b[3] + bh * padding + 1.0f, (float)(h - 1));
    } else {
      x1 = fmaxf(b[0] - 1.0f - bw * padding, 0.0f);
      x2 = fminf(b[2] + 1.0f + bw * padding, (float)w);
      y1 = fmaxf(b[1] - 1.0f - bh * padding, 0.0f);
      y2 = fminf(b[3] + 1.0f + bh * padding, (float)h);
    }
    for (int r = 0; r < h; ++r)
      for (int c = 0; c < w; ++c) {
        const size_t o = ((size_t)i * h + r) * w + c;
        const int in = ((float)c >= x1) && ((float)c < x2) && ((float)r >= y1) && ((float)r < y2);
        if (ind) ind[o] = (uint8_t)in;
        if (masks_out) masks_out[o] = masks_in[o] * (in ? 1.0f : 0.0f);
      }
  }
}

/* ------------------------------------------------------------------------- */
/* helpers exported for tests                                                 */
/* ------------------------------------------------------------------------- */
ORC_API float orc_expf(float x) { return rp_expf(x); }
ORC_API float orc_sigmoidf(float x) { return rp_sigmoidf(x); }
ORC_API float orc_float_le(double d) { return rp_float_le(d); }
ORC_API float orc_float_ge(double d) { return rp_float_ge(d); }
ORC_API float orc_iou_plus1(const float* a, const float* b) {
  return rp_iou_plus1(a[0], a[1], a[2], a[3], rp_area_plus1(a[0], a[1], a[2], a[3]), b[0], b[1],
                      b[2], b[3], rp_area_plus1(b[0], b[1], b[2], b[3]));
}

/* yol/layers/modules/multibox_loss.py:616-660 (lincomb_mask_loss, sigmoid activation), per image:
 *   pred = sigmoid(proto @ coef^T) (:616-618); crop(pred, boxes) when boxes != NULL (box_utils.py:355-378);
 *   bce = -(t*log(max(p,1e-20)) + (1-t)*log(max(1-p,1e-20))), p = clamp(pred,0,1) (:8-9,:632);
 *   loss[i] = scale[i] * sum_{y,x} bce[y,x,i]  (scale NULL = 1; :645-653 are folded into scale by the caller)
 * and the gradients of sum_i loss[i] w.r.t. proto [H,W,D] and coef [n,D] (what autograd produces:
 * the max(.,1e-20) clamps pass no gradient below the clamp, the crop zeroes it outside the box).
 * Sums are accumulated in double (the checker should be the more exact side). */
ORC_API void orc_lincomb_mask_loss(const float* proto, const float* coef, const float* mask_t,
                                   const float* boxes, const float* scale, int H, int W, int D, int n,
                                   float* loss, float* d_proto, float* d_coef) {
  double* lacc = (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
  double* dc = (double*)calloc((size_t)(n > 0 ? n : 1) * D, sizeof(double));
  float* cb = (float*)malloc(sizeof(float) * 4 * (size_t)(n > 0 ? n : 1));
  for (int i = 0; i < n; ++i) {
    float x1 = 0.f, x2 = (float)W, y1 = 0.f, y2 = (float)H;
    if (boxes) {
      orc_sanitize(boxes[4 * i], boxes[4 * i + 2], W, 1, &x1, &x2);
      orc_sanitize(boxes[4 * i + 1], boxes[4 * i + 3], H, 1, &y1, &y2);
    }
    cb[4 * i] = x1; cb[4 * i + 1] = x2; cb[4 * i + 2] = y1; cb[4 * i + 3] = y2;
  }
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {
      const size_t pix = (size_t)y * W + x;
      const float* pr = proto + pix * D;
      double* dp = (double*)calloc((size_t)D, sizeof(double));
      for (int i = 0; i < n; ++i) {
        const float* c = coef + (size_t)i * D;
        float z = 0.f;
        for (int k = 0; k < D; ++k) z = fmaf(pr[k], c[k], z);
        const float t = mask_t[pix * n + i];
        const int inside = (float)x >= cb[4 * i] && (float)x < cb[4 * i + 1] && (float)y >= cb[4 * i + 2] &&
                           (float)y < cb[4 * i + 3];
        const float p = inside ? 1.0f / (1.0f + expf(-z)) : 0.f;
        const float q = 1.0f - p;
        const float bce = -(t * logf(p > 1e-20f ? p : 1e-20f) + (1.0f - t) * logf(q > 1e-20f ? q : 1e-20f));
        const float s = scale ? scale[i] : 1.0f;
        lacc[i] += (double)s * (double)bce;
        if (inside) {
          const float gpos = p > 1e-20f ? -t * q : 0.f;
          const float gneg = q > 1e-20f ? (1.0f - t) * p : 0.f;
          const double g = (double)s * (double)(gpos + gneg);
          for (int k = 0; k < D; ++k) {
            dp[k] += g * c[k];
            dc[(size_t)i * D + k] += g * pr[k];
          }
        }
      }
      for (int k = 0; k < D; ++k) d_proto[pix * D + k] = (float)dp[k];
      free(dp);
    }
  for (int i = 0; i < n; ++i) loss[i] = (float)lacc[i];
  for (size_t i = 0; i < (size_t)n * D; ++i) d_coef[i] = (float)dc[i];
  free(lacc); free(dc); free(cb);
}

/* Pose2Seg.jittor/modeling/build_model.py:31-43 (warpAffine): cv2.warpAffine(predmap, H_e2e[0:2],
 * (width, height), WARP_INVERSE_MAP + INTER_LINEAR, BORDER_CONSTANT), channel `c` of the result,
 * > 0.5 -> 1 else 0.  pred [n,h,w,C] float; M [n,6] double (dst -> src); out [n,height,width] u8;
 * val_out (optional) receives the interpolated float values.  Arithmetic: include/roipath_math.h
 * (OpenCV's 10-bit fixed-point coordinates and 32x32 weight table); pinned against cv2 itself in
 * tests/test_oracle_golden.py. */
ORC_API void orc_affine_paste(const float* pred, const double* M, int n, int h, int w, int C, int c,
                              int width, int height, uint8_t* out, float* val_out) {
  for (int i = 0; i < n; ++i) {
    const float* src = pred + (size_t)i * h * w * C;
    const double* Mi = M + 6 * (size_t)i;
    for (int y = 0; y < height; ++y)
      for (int x = 0; x < width; ++x) {
        int sx, sy, fx, fy;
        rp_warp_affine_coord(Mi, x, y, &sx, &sy, &fx, &fy);
        const float v = rp_warp_affine_sample(src, h, w, C, c, sx, sy, fx, fy);
        const size_t o = ((size_t)i * height + y) * width + x;
        out[o] = v > 0.5f ? 1 : 0;
        if (val_out) val_out[o] = v;
      }
  }
}
