/*
 * roipath_math.h -- arithmetic shared, token for token, by the CUDA kernels
 * (instancesegmentation-jittor_b200/csrc) and the CPU oracle (oracle/).
 *
 * Every function here is written so that gcc (-ffp-contract=off) and nvcc
 * (--fmad=false) produce bit-identical fp32 results: only IEEE-754 correctly
 * rounded +,-,*,/ and explicit fmaf() are used, no libm transcendental.  This is
 * what lets the parity tests demand bit-exact keep lists and binary masks.
 *
 * Reference sites whose arithmetic is restated here (paths relative to the
 * reference checkout):
 *   - IoU with the legacy +1 convention:
 *       detectron.jittor/detectron/layers/nms.py:28-34
 *       Yolact.jittor/utils/cython_nms.pyx:31,61-68
 *   - IoU without +1 on relative coordinates:
 *       Yolact.jittor/layers/box_utils.py:31-78 (intersect / jaccard)
 *   - sigmoid:  Yolact.jittor/data/config.py:374,463 (mask_proto_mask_activation),
 *               detectron .../roi_heads/mask_head/inference.py:38
 *   - FPN level heuristic: detectron.jittor/detectron/modeling/poolers.py:31-43,62-73
 *
 * exp/sigmoid/log2 live in un-vendored Jittor (libm / CUDA libdevice); the
 * versions below are this repo's own definitions ("parity unpinned vs live
 * Jittor").
 */
#ifndef ROIPATH_MATH_H_
#define ROIPATH_MATH_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define RP_HD __host__ __device__ __forceinline__
#else
#define RP_HD static inline
#endif

RP_HD float rp_bits_to_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, sizeof f);
  return f;
#endif
}

RP_HD uint32_t rp_float_to_bits(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, sizeof u);
  return u;
#endif
}

RP_HD float rp_fminf(float a, float b) { return a < b ? a : b; }
RP_HD float rp_fmaxf(float a, float b) { return a > b ? a : b; }

/* exp(x) in fp32: Cody-Waite reduction by ln2 (hi/lo), degree-6 polynomial,
 * scaling by 2^n in two exact steps.  Max error ~1 ulp on [-87, 88]. */
RP_HD float rp_expf(float x) {
  if (x != x) return x;
  if (x > 88.72283f) return rp_bits_to_float(0x7f800000u);
  if (x < -87.0f) return 0.0f;
  float n = rintf(x * 1.44269504088896341f);
  float r = fmaf(n, -0.693145751953125f, x);
  r = fmaf(n, -1.42860682030941723e-6f, r);
  float p = 1.9875691500e-4f;
  p = fmaf(p, r, 1.3981999507e-3f);
  p = fmaf(p, r, 8.3334519073e-3f);
  p = fmaf(p, r, 4.1665795894e-2f);
  p = fmaf(p, r, 1.6666665459e-1f);
  p = fmaf(p, r, 5.0000001201e-1f);
  p = fmaf(p, r * r, r);
  p = p + 1.0f;
  int ni = (int)n;
  int a = ni >> 1;
  int b = ni - a;
  float sa = rp_bits_to_float((uint32_t)(a + 127) << 23);
  float sb = rp_bits_to_float((uint32_t)(b + 127) << 23);
  return (p * sa) * sb;
}

/* sigmoid(x) = 1 / (1 + exp(-x)), IEEE division. */
RP_HD float rp_sigmoidf(float x) { return 1.0f / (1.0f + rp_expf(-x)); }

/* floor(log2(t)) for a finite positive float, exactly (exponent extraction,
 * subnormals included).  Returns INT32_MIN for t <= 0 or NaN. */
RP_HD int rp_floor_log2f(float t) {
  if (!(t > 0.0f)) return INT32_MIN;
  uint32_t u = rp_float_to_bits(t);
  int e = (int)(u >> 23);
  if (e == 255) return 128; /* +inf */
  if (e != 0) return e - 127;
  /* subnormal: position of the top mantissa bit */
  uint32_t m = u & 0x7fffffu;
  int k = -127;
  while (!(m & 0x400000u)) {
    m <<= 1;
    --k;
  }
  return k;
}

/* Legacy "+1" IoU, operand order of nms.py:28-33:
 *   inter = inter_h * inter_w ; iou = inter / ((area_j + area_i) - inter)
 * a* is box j (already kept), b* is box i (candidate); the expression is
 * symmetric under fp32 commutativity so the roles do not matter. */
RP_HD float rp_area_plus1(float x1, float y1, float x2, float y2) {
  return ((x2 - x1) + 1.0f) * ((y2 - y1) + 1.0f);
}

RP_HD float rp_iou_plus1(float ax1, float ay1, float ax2, float ay2, float aarea,
                         float bx1, float by1, float bx2, float by2, float barea) {
  float iw = rp_fmaxf(0.0f, (rp_fminf(ax2, bx2) - rp_fmaxf(ax1, bx1)) + 1.0f);
  float ih = rp_fmaxf(0.0f, (rp_fminf(ay2, by2) - rp_fmaxf(ay1, by1)) + 1.0f);
  float inter = ih * iw;
  return inter / ((aarea + barea) - inter);
}

/* Yolact jaccard on relative point-form boxes (no +1), box_utils.py:31-78:
 *   inter = clamp(min(a.x2,b.x2)-max(a.x1,b.x1),0) * clamp(min(a.y2,b.y2)-max(a.y1,b.y1),0)
 *   union = (area_a + area_b) - inter ; 0/0 -> NaN is kept. */
RP_HD float rp_area_rel(float x1, float y1, float x2, float y2) {
  return (x2 - x1) * (y2 - y1);
}

RP_HD float rp_jaccard(float ax1, float ay1, float ax2, float ay2, float aarea,
                       float bx1, float by1, float bx2, float by2, float barea) {
  float dx = rp_fmaxf(rp_fminf(ax2, bx2) - rp_fmaxf(ax1, bx1), 0.0f);
  float dy = rp_fmaxf(rp_fminf(ay2, by2) - rp_fmaxf(ay1, by1), 0.0f);
  float inter = dx * dy;
  return inter / ((aarea + barea) - inter);
}

/* FPN level of a box (poolers.py:31-43): floor(lvl0 + log2(sqrt(area)/s0 + eps)),
 * clamped to [k_min,k_max], minus k_min.  floor(log2(.)) is evaluated exactly. */
RP_HD int rp_fpn_level(float x1, float y1, float x2, float y2, int k_min, int k_max,
                       float s0, int lvl0, float eps) {
  float s = sqrtf(rp_area_plus1(x1, y1, x2, y2));
  float t = s / s0 + eps;
  int fl = rp_floor_log2f(t);
  int lvl;
  if (fl == INT32_MIN) {
    lvl = k_min;
  } else {
    lvl = lvl0 + fl;
    if (lvl < k_min) lvl = k_min;
    if (lvl > k_max) lvl = k_max;
  }
  return lvl - k_min;
}

/* CenterMask level (poolers.py:62-73): ceil(k_max - log2(img_area/box_area + eps)). */
RP_HD int rp_fpn_level_area(float x1, float y1, float x2, float y2, float img_area,
                            int k_min, int k_max, float eps) {
  float a = rp_area_plus1(x1, y1, x2, y2);
  float t = img_area / a + eps;
  int fl = rp_floor_log2f(t);
  int lvl;
  if (fl == INT32_MIN) {
    lvl = k_max;
  } else {
    /* ceil(k - L) = k - floor(L) for integer k and every real L = log2(t) */
    lvl = k_max - fl;
    if (lvl < k_min) lvl = k_min;
    if (lvl > k_max) lvl = k_max;
  }
  return lvl - k_min;
}

/* Half-pixel bilinear source coordinate, align_corners=False (PyTorch / Caffe2
 * convention): src = max(0, scale*(dst+0.5) - 0.5), scale = in/out. */
RP_HD float rp_resize_src(int dst, float scale) {
  float s = scale * ((float)dst + 0.5f) - 0.5f;
  return s < 0.0f ? 0.0f : s;
}

/* Largest float <= d  (so that  (double)f > d  <=>  f > rp_float_below(d)). */
RP_HD float rp_float_le(double d) {
  float f = (float)d;
  if ((double)f > d) {
    uint32_t u = rp_float_to_bits(f);
    if (f > 0.0f) u -= 1u;
    else if (f < 0.0f) u += 1u;
    else u = 0x80000001u; /* just below zero */
    f = rp_bits_to_float(u);
  }
  return f;
}

/* Smallest float >= d  (so that  (double)f >= d  <=>  f >= rp_float_ge(d)). */
RP_HD float rp_float_ge(double d) {
  float f = (float)d;
  if ((double)f < d) {
    uint32_t u = rp_float_to_bits(f);
    if (f > 0.0f) u += 1u;
    else if (f < 0.0f) u -= 1u;
    else u = 0x00000001u;
    f = rp_bits_to_float(u);
  }
  return f;
}

/* ---- cv2.warpAffine(flags = WARP_INVERSE_MAP + INTER_LINEAR, borderMode = BORDER_CONSTANT, 0) for
 * float32 images, as Pose2Seg.jittor/modeling/build_model.py:31-43 calls it.  OpenCV (imgwarp.cpp,
 * warpAffine / remapBilinear) evaluates the source coordinate in 10-bit fixed point
 *   X = (cvRound((M[1]*y + M[2]) * 1024) + 16 + cvRound(M[0]*x * 1024)) >> 5       (same for Y with M[3..5])
 * so that X >> 5 is the integer source column (saturated to int16) and X & 31 indexes a table of 32x32
 * bilinear weights  w = {(1-fy)(1-fx), (1-fy)fx, fy(1-fx), fy*fx},  f = i * (1/32) in float; the value is
 * S00*w0 + S01*w1 + S10*w2 + S11*w3 in float, taps outside the image read 0. */
RP_HD int rp_cv_round(double v) {
#if defined(__CUDA_ARCH__)
  return __double2int_rn(v);
#else
  return (int)lrint(v);
#endif
}

RP_HD int rp_sat_short(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

RP_HD void rp_warp_affine_coord(const double* M, int x, int y, int* sx, int* sy, int* fx, int* fy) {
  const int X0 = rp_cv_round((M[1] * (double)y + M[2]) * 1024.0) + 16;
  const int Y0 = rp_cv_round((M[4] * (double)y + M[5]) * 1024.0) + 16;
  const int X = (X0 + rp_cv_round(M[0] * (double)x * 1024.0)) >> 5;
  const int Y = (Y0 + rp_cv_round(M[3] * (double)x * 1024.0)) >> 5;
  *sx = rp_sat_short(X >> 5);
  *sy = rp_sat_short(Y >> 5);
  *fx = X & 31;
  *fy = Y & 31;
}

/* channel c of an interleaved [h, w, C] float image */
RP_HD float rp_warp_affine_sample(const float* src, int h, int w, int C, int c, int sx, int sy, int fx,
                                  int fy) {
  if (sx >= w || sx + 1 < 0 || sy >= h || sy + 1 < 0) return 0.0f;
  const float scale = 1.0f / 32.0f;
  const float x = (float)fx * scale, y = (float)fy * scale;
  const float x0 = 1.0f - x, y0 = 1.0f - y;
  const float w0 = y0 * x0, w1 = y0 * x, w2 = y * x0, w3 = y * x;
  const int okx0 = sx >= 0 && sx < w, okx1 = sx + 1 >= 0 && sx + 1 < w;
  const int oky0 = sy >= 0 && sy < h, oky1 = sy + 1 >= 0 && sy + 1 < h;
  const float v0 = (okx0 && oky0) ? src[((size_t)sy * w + sx) * C + c] : 0.0f;
  const float v1 = (okx1 && oky0) ? src[((size_t)sy * w + sx + 1) * C + c] : 0.0f;
  const float v2 = (okx0 && oky1) ? src[((size_t)(sy + 1) * w + sx) * C + c] : 0.0f;
  const float v3 = (okx1 && oky1) ? src[((size_t)(sy + 1) * w + sx + 1) * C + c] : 0.0f;
  return v0 * w0 + v1 * w1 + v2 * w2 + v3 * w3;
}

#endif /* ROIPATH_MATH_H_ */
