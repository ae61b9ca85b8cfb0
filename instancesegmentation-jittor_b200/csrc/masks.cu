// masks.cu -- mask assembly (Yolact) and mask paste-back (Mask R-CNN) for sm_100a.
//
// Reference being replaced (paths relative to the reference checkout):
//   yol/layers/output_utils.py:60-101   postprocess, lincomb branch: proto @ coeff^T ->
//                                       sigmoid -> crop -> permute -> bilinear resize -> > 0.5
//                                       (cuBLAS SGEMM + ~12 elementwise passes over
//                                       [138,138,m] / [m,550,550] fp32 temporaries)
//   yol/layers/box_utils.py:333-378     sanitize_coordinates / crop
//   det/modeling/roi_heads/mask_head/inference.py:38-45    sigmoid + label select
//   det/modeling/roi_heads/mask_head/inference.py:90-162,175-205  expand / paste_mask_in_image
//                                       (python loop with a .numpy() sync per instance)
//
// Both kernels are bandwidth kernels: the only large tensor is the uint8 output, written
// once with 4-pixel packed stores; nothing of size [m,H,W] fp32 is ever materialised.
// All arithmetic that decides a mask bit (fma chain, rp_sigmoidf, crop compare, bilinear)
// is the oracle's, operation for operation, so the binary masks are bit-exact.
#include "common.cuh"

namespace {

// --------------------------------------------------------------------------------------
// shared pieces
// --------------------------------------------------------------------------------------
// yol/layers/box_utils.py:333-351 sanitize_coordinates(cast=False)
__device__ __forceinline__ void mk_sanitize(float a, float b, int img_size, int padding, float* lo,
                                            float* hi) {
  const float x1 = a * (float)img_size;
  const float x2 = b * (float)img_size;
  float mn = rp_fminf(x1, x2);
  float mx = rp_fmaxf(x1, x2);
  mn = mn - (float)padding;
  if (mn < 0.f) mn = 0.f;
  mx = mx + (float)padding;
  if (mx > (float)img_size) mx = (float)img_size;
  *lo = mn;
  *hi = mx;
}

// One axis of the half-pixel bilinear resize (align_corners=False), indices clamped.
struct MkTap {
  int i0, i1;
  float w0, w1;
};
__device__ __forceinline__ MkTap mk_tap(int dst, float scale, int in_size) {
  MkTap t;
  const float f = rp_resize_src(dst, scale);
  int i0 = (int)f;
  if (i0 > in_size - 1) i0 = in_size - 1;
  t.i0 = i0;
  t.i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  t.w1 = f - (float)i0;
  t.w0 = 1.0f - t.w1;
  return t;
}

// store 4 consecutive mask bytes at p (any alignment), n_valid of them inside the row
__device__ __forceinline__ void mk_store4(uint8_t* p, uint32_t packed, int n_valid) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  if (n_valid >= 4 && (a & 3) == 0) {
    __stcs(reinterpret_cast<uint32_t*>(p), packed);
  } else if (n_valid >= 4 && (a & 1) == 0) {
    __stcs(reinterpret_cast<unsigned short*>(p), (unsigned short)(packed & 0xffffu));
    __stcs(reinterpret_cast<unsigned short*>(p + 2), (unsigned short)(packed >> 16));
  } else {
    for (int i = 0; i < n_valid && i < 4; ++i) p[i] = (uint8_t)((packed >> (8 * i)) & 0xffu);
  }
}

// --------------------------------------------------------------------------------------
// Yolact mask assembly
// --------------------------------------------------------------------------------------
constexpr int MA_THREADS = 256;
constexpr int MA_LIST = 128;  // detections per tile kept in the shared list (more: the plain loop)
constexpr int MA_TW = 128;  // output tile width  (32 threads x 4 pixels)
constexpr int MA_TH = 32;   // output tile height (8 thread rows x 4 rows)
#ifndef MA_MINB
#define MA_MINB 4  // resident CTAs per SM asked of the register allocator
#endif
constexpr int MA_DMAX = 64;

// CTA = (output tile, image).  The tile's prototype footprint is staged once, k-major, in
// shared memory and reused by every detection of the image.
__global__ void __launch_bounds__(MA_THREADS, MA_MINB)
mask_assemble_kernel(const float* __restrict__ proto, const float* __restrict__ coeff,
                     const float* __restrict__ boxes, const int* __restrict__ det_offsets,
                     int m_total, int ph, int pw, int D, int out_h, int out_w, int crop,
                     int tiles_x, int fp_rows_max, int fp_cols_max,
                     uint8_t* __restrict__ out_masks, int prefilled) {
  extern __shared__ __align__(16) float ma_smem[];
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int tile_y = blockIdx.x / tiles_x, tile_x = blockIdx.x - tile_y * tiles_x;
  const int Y0 = tile_y * MA_TH, X0 = tile_x * MA_TW;
  const int Y1 = min(out_h, Y0 + MA_TH) - 1, X1 = min(out_w, X0 + MA_TW) - 1;
  const int j0 = det_offsets ? det_offsets[b] : 0;
  const int j1 = det_offsets ? det_offsets[b + 1] : m_total;
  if (j0 >= j1) return;
  const float sy = (float)ph / (float)out_h, sx = (float)pw / (float)out_w;
  // footprint of the tile in the prototype grid
  const MkTap ty0 = mk_tap(Y0, sy, ph), ty1 = mk_tap(Y1, sy, ph);
  const MkTap tx0 = mk_tap(X0, sx, pw), tx1 = mk_tap(X1, sx, pw);
  const int fy0 = ty0.i0, fy1 = ty1.i1, fx0 = tx0.i0, fx1 = tx1.i1;
  // detections whose crop box reaches this tile (round 2: 30 % of the kernel's instructions were every
  // thread re-deriving every detection's crop box): listed once per CTA with their sanitized boxes
  __shared__ int s_list[MA_LIST];
  __shared__ float4 s_box[MA_LIST];
  __shared__ int s_nlist;
  bool use_list = false;
  if (prefilled && crop) {
    // the output was zero-filled by the caller: a tile no detection's crop box reaches has
    // nothing to write and does not even stage the prototypes
    if (tid == 0) s_nlist = 0;
    __syncthreads();
    for (int j = j0 + tid; j < j1; j += MA_THREADS) {
      const float* bxj = boxes + (size_t)j * 4;
      float cx1, cx2, cy1, cy2;
      mk_sanitize(bxj[0], bxj[2], pw, 1, &cx1, &cx2);
      mk_sanitize(bxj[1], bxj[3], ph, 1, &cy1, &cy2);
      if (!(!((float)fx1 >= cx1) || !((float)fx0 < cx2) || !((float)fy1 >= cy1) || !((float)fy0 < cy2))) {
        const int pos = atomicAdd(&s_nlist, 1);
        if (pos < MA_LIST) {
          s_list[pos] = j;
          s_box[pos] = make_float4(cx1, cx2, cy1, cy2);
        }
      }
    }
    __syncthreads();
    if (s_nlist == 0) return;
    use_list = s_nlist <= MA_LIST;
  }
  const int n_iter = use_list ? s_nlist : j1 - j0;
  const int frows = fy1 - fy0 + 1, fcols = fx1 - fx0 + 1;
  const int npix = frows * fcols;
  const int pixp = (fp_rows_max * fp_cols_max) | 1;  // odd pitch: conflict-free k-major staging
  float* proto_s = ma_smem;                       // [D][pixp]
  float* low_s = proto_s + (size_t)D * pixp;      // [npix]
  float* coef_s = low_s + ((fp_rows_max * fp_cols_max + 3) & ~3);  // [D]
  const float* proto_b = proto + (size_t)b * ph * pw * D;
  // stage: consecutive threads read consecutive k of one pixel (coalesced 4*D bytes)
  for (int i = tid; i < npix * D; i += MA_THREADS) {
    const int pix = i / D, k = i - pix * D;
    const int py = pix / fcols, px = pix - py * fcols;
    proto_s[(size_t)k * pixp + pix] = __ldg(proto_b + ((size_t)(fy0 + py) * pw + (fx0 + px)) * D + k);
  }
  // this thread's output pixels: 4 columns x 4 rows
  const int tx = tid & 31, tyq = tid >> 5;
  const int Xb = X0 + tx * 4;
  MkTap xt[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    xt[i] = mk_tap(min(Xb + i, out_w - 1), sx, pw);
    xt[i].i0 -= fx0;
    xt[i].i1 -= fx0;
  }
  __syncthreads();

  for (int it = 0; it < n_iter; ++it) {
    int j;
    float cx1, cx2, cy1, cy2;
    bool empty;
    if (use_list) {
      j = s_list[it];
      const float4 bx4 = s_box[it];
      cx1 = bx4.x;
      cx2 = bx4.y;
      cy1 = bx4.z;
      cy2 = bx4.w;
      empty = false;
    } else {
      j = j0 + it;
      const float* bxj = boxes + (size_t)j * 4;
      mk_sanitize(bxj[0], bxj[2], pw, 1, &cx1, &cx2);
      mk_sanitize(bxj[1], bxj[3], ph, 1, &cy1, &cy2);
      // no prototype pixel of the footprint inside the crop box => the whole tile is zero
      empty = crop && (!((float)fx1 >= cx1) || !((float)fx0 < cx2) || !((float)fy1 >= cy1) ||
                       !((float)fy0 < cy2));
    }
    uint8_t* out_j = out_masks + (size_t)j * out_h * out_w;
    if (!empty) {
      if (tid < D) coef_s[tid] = coeff[(size_t)j * D + tid];
      __syncthreads();
      for (int pix = tid; pix < npix; pix += MA_THREADS) {
        const int py = pix / fcols, px = pix - py * fcols;
        const int gy = fy0 + py, gx = fx0 + px;
        const bool in = !crop || (((float)gx >= cx1) && ((float)gx < cx2) &&
                                  ((float)gy >= cy1) && ((float)gy < cy2));
        float v = 0.f;
        if (in) {
          float acc = 0.f;
          if ((D & 3) == 0) {  // four coefficients per (broadcast) shared load; same k order
            const float* pp = proto_s + pix;
#pragma unroll 2
            for (int k = 0; k < D; k += 4) {
              const float4 c4 = *reinterpret_cast<const float4*>(coef_s + k);
              acc = fmaf(pp[(size_t)k * pixp], c4.x, acc);
              acc = fmaf(pp[(size_t)(k + 1) * pixp], c4.y, acc);
              acc = fmaf(pp[(size_t)(k + 2) * pixp], c4.z, acc);
              acc = fmaf(pp[(size_t)(k + 3) * pixp], c4.w, acc);
            }
          } else {
            for (int k = 0; k < D; ++k) acc = fmaf(proto_s[(size_t)k * pixp + pix], coef_s[k], acc);
          }
          v = rp_sigmoidf(acc);
        }
        low_s[pix] = v;
      }
      __syncthreads();
    }
    if (empty && prefilled) continue;  // zeros are already there
#pragma unroll
    for (int ry = 0; ry < 4; ++ry) {
      const int Y = Y0 + tyq * 4 + ry;
      if (Y > Y1 || Xb > X1) continue;
      uint32_t packed = 0;
      if (!empty) {
        const MkTap yt = mk_tap(Y, sy, ph);
        const float* r0 = low_s + (yt.i0 - fy0) * fcols;
        const float* r1 = low_s + (yt.i1 - fy0) * fcols;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float top = xt[i].w0 * r0[xt[i].i0] + xt[i].w1 * r0[xt[i].i1];
          const float bot = xt[i].w0 * r1[xt[i].i0] + xt[i].w1 * r1[xt[i].i1];
          const float val = yt.w0 * top + yt.w1 * bot;
          if (val > 0.5f) packed |= (1u << (8 * i));
        }
      }
      mk_store4(out_j + (size_t)Y * out_w + Xb, packed, X1 - Xb + 1);
    }
    if (!empty) __syncthreads();  // low_s / coef_s are rewritten by the next detection
  }
}

// Fallback for resize ratios whose per-tile prototype footprint does not fit shared memory
// (down-sampling, which Yolact only meets for images smaller than the 138x138 prototypes):
// one thread per 4 output pixels, the four low-resolution taps of each pixel recomputed from
// global memory.  Same operations per mask bit as the tiled kernel => same bits.
__device__ __forceinline__ float ma_low_value(const float* __restrict__ proto_b,
                                              const float* __restrict__ cf, int y, int x, int pw,
                                              int D, int crop, float cx1, float cx2, float cy1,
                                              float cy2) {
  const bool in = !crop || (((float)x >= cx1) && ((float)x < cx2) && ((float)y >= cy1) &&
                            ((float)y < cy2));
  if (!in) return 0.f;
  const float* p = proto_b + ((size_t)y * pw + x) * D;
  float acc = 0.f;
  for (int k = 0; k < D; ++k) acc = fmaf(__ldg(p + k), __ldg(cf + k), acc);
  return rp_sigmoidf(acc);
}

__global__ void __launch_bounds__(256)
mask_assemble_direct_kernel(const float* __restrict__ proto, const float* __restrict__ coeff,
                            const float* __restrict__ boxes, const int* __restrict__ det_offsets,
                            int B, int m_total, int ph, int pw, int D, int out_h, int out_w,
                            int crop, uint8_t* __restrict__ out_masks) {
  const int groups = (out_w + 3) >> 2;
  const size_t per_det = (size_t)groups * out_h;
  const size_t total = per_det * m_total;
  const float sy = (float)ph / (float)out_h, sx = (float)pw / (float)out_w;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int j = (int)(i / per_det);
    const size_t rem = i - (size_t)j * per_det;
    const int Y = (int)(rem / groups);
    const int X = (int)(rem - (size_t)Y * groups) * 4;
    int b = 0;
    if (det_offsets) {
      int lo = 0, hi = B - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (det_offsets[mid] <= j)
          lo = mid;
        else
          hi = mid - 1;
      }
      b = lo;
    }
    const float* proto_b = proto + (size_t)b * ph * pw * D;
    const float* cf = coeff + (size_t)j * D;
    const float* bxj = boxes + (size_t)j * 4;
    float cx1, cx2, cy1, cy2;
    mk_sanitize(bxj[0], bxj[2], pw, 1, &cx1, &cx2);
    mk_sanitize(bxj[1], bxj[3], ph, 1, &cy1, &cy2);
    const MkTap yt = mk_tap(Y, sy, ph);
    uint32_t packed = 0;
    for (int q = 0; q < 4 && X + q < out_w; ++q) {
      const MkTap xt = mk_tap(X + q, sx, pw);
      const float v00 = ma_low_value(proto_b, cf, yt.i0, xt.i0, pw, D, crop, cx1, cx2, cy1, cy2);
      const float v01 = ma_low_value(proto_b, cf, yt.i0, xt.i1, pw, D, crop, cx1, cx2, cy1, cy2);
      const float v10 = ma_low_value(proto_b, cf, yt.i1, xt.i0, pw, D, crop, cx1, cx2, cy1, cy2);
      const float v11 = ma_low_value(proto_b, cf, yt.i1, xt.i1, pw, D, crop, cx1, cx2, cy1, cy2);
      const float top = xt.w0 * v00 + xt.w1 * v01;
      const float bot = xt.w0 * v10 + xt.w1 * v11;
      const float val = yt.w0 * top + yt.w1 * bot;
      if (val > 0.5f) packed |= (1u << (8 * q));
    }
    mk_store4(out_masks + ((size_t)j * out_h + Y) * out_w + X, packed, out_w - X);
  }
}

// boxes -> absolute int32 (yol/layers/output_utils.py:99-101) and optional low-res output
__global__ void mask_boxes_kernel(const float* __restrict__ boxes, int m, int out_h, int out_w,
                                  int32_t* __restrict__ out_boxes) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const float* b = boxes + (size_t)j * 4;
  float ax1, ax2, ay1, ay2;
  mk_sanitize(b[0], b[2], out_w, 0, &ax1, &ax2);
  mk_sanitize(b[1], b[3], out_h, 0, &ay1, &ay2);
  out_boxes[j * 4 + 0] = (int32_t)ax1;
  out_boxes[j * 4 + 1] = (int32_t)ay1;
  out_boxes[j * 4 + 2] = (int32_t)ax2;
  out_boxes[j * 4 + 3] = (int32_t)ay2;
}

__global__ void mask_low_kernel(const float* __restrict__ proto, const float* __restrict__ coeff,
                                const float* __restrict__ boxes, const int* __restrict__ det_offsets,
                                int B, int m_total, int ph, int pw, int D, int crop,
                                float* __restrict__ low_out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t total = (size_t)m_total * ph * pw;
  if (i >= total) return;
  const int x = (int)(i % pw);
  const int y = (int)((i / pw) % ph);
  const int j = (int)(i / ((size_t)pw * ph));
  int b = 0;
  if (det_offsets) {
    int lo = 0, hi = B - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (det_offsets[mid] <= j)
        lo = mid;
      else
        hi = mid - 1;
    }
    b = lo;
  }
  const float* p = proto + (((size_t)b * ph + y) * pw + x) * D;
  const float* cf = coeff + (size_t)j * D;
  float acc = 0.f;
  for (int k = 0; k < D; ++k) acc = fmaf(p[k], cf[k], acc);
  float v = rp_sigmoidf(acc);
  if (crop) {
    const float* bx = boxes + (size_t)j * 4;
    float x1, x2, y1, y2;
    mk_sanitize(bx[0], bx[2], pw, 1, &x1, &x2);
    mk_sanitize(bx[1], bx[3], ph, 1, &y1, &y2);
    const bool in = ((float)x >= x1) && ((float)x < x2) && ((float)y >= y1) && ((float)y < y2);
    v = v * (in ? 1.0f : 0.0f);
  }
  low_out[i] = v;
}

// --------------------------------------------------------------------------------------
// Mask R-CNN: sigmoid + label select, paste-back
// --------------------------------------------------------------------------------------
__global__ void mask_select_sigmoid_kernel(const float* __restrict__ logits,
                                           const int* __restrict__ labels, int K, int NC, int MM,
                                           float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)K * MM) return;
  const int k = (int)(i / MM);
  const int t = (int)(i - (size_t)k * MM);
  const int l = labels[k];
  const float v = (l >= 0 && l < NC) ? logits[((size_t)k * NC + l) * MM + t] : 0.f;
  out[i] = rp_sigmoidf(v);
}

constexpr int PM_THREADS = 256;
constexpr int PM_ROWS = 16;  // output rows per CTA
constexpr int PM_MAXP = 66;  // padded mask side limit (M + 2*padding)

// CTA = (row band, instance).  Every byte of the band is written exactly once.
__global__ void __launch_bounds__(PM_THREADS)
paste_masks_kernel(const float* __restrict__ masks, const float* __restrict__ boxes, int M,
                   int im_h, int im_w, float thresh, int padding, float scale,
                   uint8_t* __restrict__ out) {
  __shared__ float pad_s[PM_MAXP * PM_MAXP];
  const int tid = threadIdx.x;
  const int i = blockIdx.y;
  const int Yb = blockIdx.x * PM_ROWS;
  const int P = M + 2 * padding;
  // det/.../mask_head/inference.py:90-104 expand_boxes, :126 int32 truncation, :129-132
  const float* b = boxes + (size_t)i * 4;
  float w_half = (b[2] - b[0]) * .5f;
  float h_half = (b[3] - b[1]) * .5f;
  const float x_c = (b[2] + b[0]) * .5f;
  const float y_c = (b[3] + b[1]) * .5f;
  w_half *= scale;
  h_half *= scale;
  const int bx0 = (int)(x_c - w_half), bx1 = (int)(x_c + w_half);
  const int by0 = (int)(y_c - h_half), by1 = (int)(y_c + h_half);
  int w = bx1 - bx0 + 1, h = by1 - by0 + 1;
  if (w < 1) w = 1;
  if (h < 1) h = 1;
  const int x_0 = max(bx0, 0), x_1 = min(bx1 + 1, im_w);
  const int y_0 = max(by0, 0), y_1 = min(by1 + 1, im_h);
  uint8_t* out_i = out + (size_t)i * im_h * im_w;
  const int Ye = min(im_h, Yb + PM_ROWS);
  const bool touches = (Yb < y_1) && (Ye > y_0) && (x_0 < x_1);
  if (touches) {
    for (int t = tid; t < P * P; t += PM_THREADS) {
      const int py = t / P, px = t - py * P;
      const int my = py - padding, mx = px - padding;
      pad_s[t] = (my >= 0 && my < M && mx >= 0 && mx < M)
                     ? __ldg(masks + ((size_t)i * M + my) * M + mx)
                     : 0.f;
    }
    __syncthreads();
  }
  const float sy = (float)P / (float)h, sx = (float)P / (float)w;
  const int groups = (im_w + 3) >> 2;  // 4-pixel groups per row
  for (int g = tid; g < groups * (Ye - Yb); g += PM_THREADS) {
    const int ry = g / groups;
    const int Y = Yb + ry;
    const int X = (g - ry * groups) * 4;
    uint32_t packed = 0;
    if (touches && Y >= y_0 && Y < y_1 && X + 3 >= x_0 && X < x_1) {
      const int my = Y - by0;
      if (my >= 0 && my < h) {
        const MkTap yt = mk_tap(my, sy, P);
        const float* r0 = pad_s + yt.i0 * P;
        const float* r1 = pad_s + yt.i1 * P;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int Xq = X + q;
          const int mx = Xq - bx0;
          if (Xq >= x_0 && Xq < x_1 && mx >= 0 && mx < w) {
            const MkTap xt = mk_tap(mx, sx, P);
            const float top = xt.w0 * r0[xt.i0] + xt.w1 * r0[xt.i1];
            const float bot = xt.w0 * r1[xt.i0] + xt.w1 * r1[xt.i1];
            const float v = yt.w0 * top + yt.w1 * bot;
            const bool on = (thresh >= 0.f) ? (v > thresh) : ((v * 255.0f) != 0.0f);
            if (on) packed |= (1u << (8 * q));
          }
        }
      }
    }
    mk_store4(out_i + (size_t)Y * im_w + X, packed, im_w - X);
  }
}

// Flat variant: the instance plane [im_h * im_w] is written as 16-byte words regardless of the
// row length (1333-byte rows make 3 rows in 4 unaligned for word stores in the band kernel above).
// CTA = (PF_CHUNK bytes of the plane, instance).  Chunks whose rows miss the pasted box -- most of
// them: a box covers a few percent of the image -- are a pure streaming zero fill; the others
// evaluate the same per-pixel arithmetic as paste_masks_kernel.  Needs a 16-byte aligned output
// and im_h * im_w % 16 == 0.
constexpr int PF_CHUNK = 16384;

__global__ void __launch_bounds__(PM_THREADS)
paste_masks_flat_kernel(const float* __restrict__ masks, const float* __restrict__ boxes, int M,
                        int im_h, int im_w, float thresh, int padding, float scale,
                        uint8_t* __restrict__ out, int sub, int chunk) {
  __shared__ float pad_s[PM_MAXP * PM_MAXP];
  const int tid = threadIdx.x;
  const int i = blockIdx.y;
  const int P = M + 2 * padding;
  const float* b = boxes + (size_t)i * 4;
  float w_half = (b[2] - b[0]) * .5f;
  float h_half = (b[3] - b[1]) * .5f;
  const float x_c = (b[2] + b[0]) * .5f;
  const float y_c = (b[3] + b[1]) * .5f;
  w_half *= scale;
  h_half *= scale;
  const int bx0 = (int)(x_c - w_half), bx1 = (int)(x_c + w_half);
  const int by0 = (int)(y_c - h_half), by1 = (int)(y_c + h_half);
  int w = bx1 - bx0 + 1, h = by1 - by0 + 1;
  if (w < 1) w = 1;
  if (h < 1) h = 1;
  const int x_0 = max(bx0, 0), x_1 = min(min(bx1 + 1, im_w), bx0 + w);
  const int y_0 = max(by0, 0), y_1 = min(min(by1 + 1, im_h), by0 + h);
  const int plane = im_h * im_w;
  uint8_t* out_i = out + (size_t)i * plane;
  const float sy = (float)P / (float)h, sx = (float)P / (float)w;
  bool staged = false;
  // `sub` consecutive chunks per CTA: the box arithmetic above (a dependent global load) is paid once
  // per `sub * PF_CHUNK` bytes, and a launch has few enough CTAs to be resident at once
  for (int cb = 0; cb < sub; ++cb) {
    const int f0 = (blockIdx.x * sub + cb) * chunk;
    if (f0 >= plane) break;
    const int f1 = min(plane, f0 + chunk);
    const int Yfirst = f0 / im_w, Ylast = (f1 - 1) / im_w;
    const bool touches = (Yfirst < y_1) && (Ylast >= y_0) && (x_0 < x_1);
    if (!touches) {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      for (int f = f0 + tid * 16; f < f1; f += PM_THREADS * 16) __stcs(reinterpret_cast<uint4*>(out_i + f), z);
      continue;
    }
    if (!staged) {  // (CTA-uniform)
      for (int t = tid; t < P * P; t += PM_THREADS) {
        const int py = t / P, px = t - py * P;
        const int my = py - padding, mx = px - padding;
        pad_s[t] = (my >= 0 && my < M && mx >= 0 && mx < M) ? __ldg(masks + ((size_t)i * M + my) * M + mx) : 0.f;
      }
      __syncthreads();
      staged = true;
    }
    // zero the whole chunk at fill speed, then (after a barrier: same bytes, other threads) write the
    // box pixels of the chunk's rows, spread evenly over the threads in groups of 4 pixels -- only
    // 1-2 % of an image lies inside a box, and a thread that met a box while walking its own 16-byte
    // groups used to hold the whole CTA back
    {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      for (int f = f0 + tid * 16; f < f1; f += PM_THREADS * 16) __stcs(reinterpret_cast<uint4*>(out_i + f), z);
    }
    __syncthreads();
    const int Ya = max(y_0, Yfirst), Yb = min(y_1 - 1, Ylast);
    const int gpr = (x_1 - x_0 + 3) >> 2;  // groups per row
    for (int t = tid; t < (Yb - Ya + 1) * gpr; t += PM_THREADS) {
      const int ry = t / gpr, g = t - ry * gpr;
      const int Y = Ya + ry, X0 = x_0 + 4 * g;
      const MkTap yt = mk_tap(Y - by0, sy, P);
      const float* r0 = pad_s + yt.i0 * P;
      const float* r1 = pad_s + yt.i1 * P;
      const int fb = Y * im_w + X0;
      uint32_t packed = 0u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int X = X0 + q;
        if (X < x_1 && fb + q >= f0 && fb + q < f1) {
          const MkTap xt = mk_tap(X - bx0, sx, P);
          const float top = xt.w0 * r0[xt.i0] + xt.w1 * r0[xt.i1];
          const float bot = xt.w0 * r1[xt.i0] + xt.w1 * r1[xt.i1];
          const float v = yt.w0 * top + yt.w1 * bot;
          const bool on = (thresh >= 0.f) ? (v > thresh) : ((v * 255.0f) != 0.0f);
          if (on) packed |= 1u << (8 * q);
        }
      }
      if (packed) {  // (pixels that are off, or outside the chunk, are zero already / not ours)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if ((packed >> (8 * q)) & 1u) out_i[fb + q] = (uint8_t)1;
      }
    }
  }
}

// ---- bit-packing of binary masks for the final gather (parallel.gather_detections) ----------------
// masks [m, npix] uint8 (0 / non-zero) -> bits [m, ceil(npix / 8)], pixel i of a mask in bit (i & 7) of
// byte i >> 3.  One thread per 32 pixels: eight 4-byte loads (rows of 4-byte multiples) or 32 byte
// loads, one 4-byte / up to four 1-byte stores.  HBM bound: npix bytes in, npix / 8 out.
__global__ void __launch_bounds__(256)
pack_mask_bits_kernel(const uint8_t* __restrict__ masks, long long m, long long npix,
                      uint8_t* __restrict__ out) {
  const long long words = (npix + 31) >> 5;   // 32-pixel groups per mask
  const long long obytes = (npix + 7) >> 3;   // output bytes per mask
  const bool vec = (npix & 3) == 0 && (reinterpret_cast<uintptr_t>(masks) & 3) == 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m * words;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / words, w = i - r * words;
    const uint8_t* src = masks + r * npix + (w << 5);
    const long long left = npix - (w << 5);
    uint32_t bits = 0;
    if (vec && left >= 32) {
      const uint32_t* s4 = reinterpret_cast<const uint32_t*>(src);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t v = __ldcs(s4 + j);
        bits |= ((v & 0xffu) ? 1u : 0u) << (4 * j);
        bits |= ((v & 0xff00u) ? 1u : 0u) << (4 * j + 1);
        bits |= ((v & 0xff0000u) ? 1u : 0u) << (4 * j + 2);
        bits |= ((v & 0xff000000u) ? 1u : 0u) << (4 * j + 3);
      }
    } else {
      const int n = left < 32 ? (int)left : 32;
      for (int j = 0; j < n; ++j) bits |= (src[j] ? 1u : 0u) << j;
    }
    uint8_t* dst = out + r * obytes + (w << 2);
    const long long bleft = obytes - (w << 2);
    if (bleft >= 4 && ((reinterpret_cast<uintptr_t>(dst) & 3) == 0)) {
      *reinterpret_cast<uint32_t*>(dst) = bits;
    } else {
      const int nb = bleft < 4 ? (int)bleft : 4;
      for (int j = 0; j < nb; ++j) dst[j] = (uint8_t)(bits >> (8 * j));
    }
  }
}

}  // namespace


extern "C" {

int rp_mask_assemble(const float* proto, const float* coeff, const float* boxes,
                     const int* det_offsets, int B, int m_total, int ph, int pw, int D, int out_h,
                     int out_w, int crop, uint8_t* out_masks, int32_t* out_boxes, float* low_out,
                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (m_total == 0) return RP_OK;
  if (!proto || !coeff || !boxes || !out_masks || B < 1 || m_total < 0 || ph < 1 || pw < 1 ||
      D < 1 || out_h < 1 || out_w < 1)
    return RP_E_ARG;
  if (!det_offsets && B != 1) return RP_E_ARG;
  if (D > MA_DMAX) return RP_E_LIMIT;
  // prototype footprint of one output tile (+2: the hi tap and the clamp)
  const double sy = (double)ph / out_h, sx = (double)pw / out_w;
  int fr = (int)(sy * (MA_TH - 1)) + 3, fc = (int)(sx * (MA_TW - 1)) + 3;
  if (fr > ph) fr = ph;
  if (fc > pw) fc = pw;
  const size_t pix = (size_t)fr * fc;
  const size_t smem = sizeof(float) * ((size_t)D * (pix | 1) + ((pix + 3) & ~(size_t)3) + MA_DMAX);
  if (B > 65535) return RP_E_LIMIT;
  if (smem > 200 * 1024) {
    const size_t total = (size_t)((out_w + 3) >> 2) * out_h * m_total;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)RP_NUM_SMS * 64) blocks = (size_t)RP_NUM_SMS * 64;
    mask_assemble_direct_kernel<<<(unsigned)blocks, 256, 0, st>>>(
        proto, coeff, boxes, det_offsets, B, m_total, ph, pw, D, out_h, out_w, crop, out_masks);
    RP_LAUNCH_CHECK();
  } else {
    if (smem > 48 * 1024)
      RP_SET_MAX_SMEM(mask_assemble_kernel, smem);
    const int tiles_x = (out_w + MA_TW - 1) / MA_TW, tiles_y = (out_h + MA_TH - 1) / MA_TH;
    dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)B);
    // With cropping most (detection, tile) pairs lie outside the crop box: zero-fill the whole
    // output at memset speed first and let the kernel write only the tiles a box reaches (row
    // lengths such as 550 make half of the tile rows unaligned for word stores otherwise).
    const int prefill = crop && rp_env_int("RP_MASK_PREFILL", 1);
    if (prefill)
      RP_CUDA_TRY(cudaMemsetAsync(out_masks, 0, (size_t)m_total * out_h * out_w, st));
    mask_assemble_kernel<<<grid, MA_THREADS, smem, st>>>(proto, coeff, boxes, det_offsets, m_total,
                                                         ph, pw, D, out_h, out_w, crop, tiles_x,
                                                         fr, fc, out_masks, prefill);
    RP_LAUNCH_CHECK();
  }
  if (out_boxes) {
    mask_boxes_kernel<<<(m_total + 127) / 128, 128, 0, st>>>(boxes, m_total, out_h, out_w,
                                                             out_boxes);
    RP_LAUNCH_CHECK();
  }
  if (low_out) {
    const size_t total = (size_t)m_total * ph * pw;
    mask_low_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
        proto, coeff, boxes, det_offsets, B, m_total, ph, pw, D, crop, low_out);
    RP_LAUNCH_CHECK();
  }
  return RP_OK;
}

int rp_mask_select_sigmoid(const float* logits, const int* labels, int K, int NC, int M,
                           float* out, void* stream) {
  if (K == 0) return RP_OK;
  if (!logits || !labels || !out || K < 0 || NC < 1 || M < 1) return RP_E_ARG;
  const size_t total = (size_t)K * M * M;
  mask_select_sigmoid_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      logits, labels, K, NC, M * M, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_paste_masks(const float* masks, const float* boxes, int n, int M, int im_h, int im_w,
                   float thresh, int padding, uint8_t* out, void* stream) {
  if (n == 0) return RP_OK;
  if (!masks || !boxes || !out || n < 0 || M < 1 || im_h < 1 || im_w < 1 || padding < 0)
    return RP_E_ARG;
  const int P = M + 2 * padding;
  if (P > PM_MAXP) return RP_E_LIMIT;
  if (n > 65535) return RP_E_LIMIT;
  const float scale = (float)((double)P / (double)M);
  const long long plane = (long long)im_h * im_w;
  if ((plane & 15) == 0 && im_w >= 16 && plane < 0x7fff0000LL && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
      rp_env_int("RP_PASTE_FLAT", 1)) {
    int sub = rp_env_int("RP_PASTE_SUB", 1);
    if (sub < 1) sub = 1;
    int chunk = rp_env_int("RP_PASTE_CHUNK", PF_CHUNK);
    if (chunk < 4096 || (chunk & 4095)) chunk = PF_CHUNK;  // whole passes of 256 threads x 16 bytes
    const long long per_cta = (long long)chunk * sub;
    dim3 grid((unsigned)((plane + per_cta - 1) / per_cta), (unsigned)n);
    paste_masks_flat_kernel<<<grid, PM_THREADS, 0, (cudaStream_t)stream>>>(masks, boxes, M, im_h, im_w,
                                                                            thresh, padding, scale, out, sub, chunk);
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  dim3 grid((unsigned)((im_h + PM_ROWS - 1) / PM_ROWS), (unsigned)n);
  paste_masks_kernel<<<grid, PM_THREADS, 0, (cudaStream_t)stream>>>(masks, boxes, M, im_h, im_w,
                                                                     thresh, padding, scale, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_pack_mask_bits(const uint8_t* masks, long long m, long long npix, uint8_t* out, void* stream) {
  if (m == 0 || npix == 0) return RP_OK;
  if (!masks || !out || m < 0 || npix < 0) return RP_E_ARG;
  const long long groups = m * ((npix + 31) >> 5);
  long long blocks = (groups + 255) / 256;
  if (blocks > (long long)RP_NUM_SMS * 32) blocks = (long long)RP_NUM_SMS * 32;
  pack_mask_bits_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(masks, m, npix, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
