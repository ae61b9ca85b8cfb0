// roi_misc.cu -- the other RoI-family kernels of SURVEY.md 8(f) rank 4, sm_100a:
//   boxlist_iou   det/structures/boxlist_ops.py:79-116   N x M IoU (+1) for the training Matcher
//   ROIPool       det/layers/roi_pool.py:16-88 (forward, argmax), :101-148 (backward)
//   mask_prob     det/layers/mask_prob.py:9-57            EmbedMask pixel/centre margin probability
#include "common.cuh"

namespace {

// ---- boxlist_iou: CTA = 32 x 8 tile of the [N, M] matrix, rows' boxes and areas in registers -----
__global__ void __launch_bounds__(256)
iou_matrix_kernel(const float* __restrict__ a, int N, const float* __restrict__ b, int M,
                  float* __restrict__ out) {
  const int j = blockIdx.x * 32 + (threadIdx.x & 31);
  const int i0 = (blockIdx.y * 8 + (threadIdx.x >> 5)) * 4;
  if (j >= M) return;
  const float4 q = *reinterpret_cast<const float4*>(b + (size_t)j * 4);
  const float area2 = (q.z - q.x + 1.0f) * (q.w - q.y + 1.0f);  // BoxList.area(), TO_REMOVE = 1
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = i0 + r;
    if (i >= N) return;
    const float4 p = *reinterpret_cast<const float4*>(a + (size_t)i * 4);
    const float area1 = (p.z - p.x + 1.0f) * (p.w - p.y + 1.0f);
    const float w = fmaxf(fminf(p.z, q.z) - fmaxf(p.x, q.x) + 1.0f, 0.0f);  // :106-111
    const float h = fmaxf(fminf(p.w, q.w) - fmaxf(p.y, q.y) + 1.0f, 0.0f);
    const float inter = w * h;
    __stcs(out + (size_t)i * M + j, inter / (area1 + area2 - inter));  // :114
  }
}

// ---- ROIPool ------------------------------------------------------------------------------
// One thread per output element, index order (n, c, ph, pw) as the reference: neighbouring lanes
// read neighbouring bins of one plane.  round() is C round (half away from zero), the bin maths
// is float as in the kernel text (roi_pool.py:44-69).
__global__ void __launch_bounds__(256)
roi_pool_fwd_kernel(const float* __restrict__ feat, const float* __restrict__ rois, size_t total,
                    int C, int H, int W, int PH, int PW, float scale, float* __restrict__ out,
                    int* __restrict__ argmax) {
  const size_t index = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (index >= total) return;
  const int pw = (int)(index % PW);
  const int ph = (int)((index / PW) % PH);
  const int c = (int)((index / PW / PH) % C);
  const int n = (int)(index / PW / PH / C);
  const float* r = rois + (size_t)n * 5;
  const int bi = (int)r[0];
  const int sw = (int)roundf(r[1] * scale), sh = (int)roundf(r[2] * scale);
  const int ew = (int)roundf(r[3] * scale), eh = (int)roundf(r[4] * scale);
  const int rw = max(ew - sw + 1, 1), rh = max(eh - sh + 1, 1);
  const float bh = (float)(1.0 * rh / PH), bw = (float)(1.0 * rw / PW);  // double quotient, :55-56
  int hs = (int)floorf(ph * bh), ws = (int)floorf(pw * bw);
  int he = (int)ceilf((ph + 1) * bh), we = (int)ceilf((pw + 1) * bw);
  hs = min(max(hs + sh, 0), H);
  he = min(max(he + sh, 0), H);
  ws = min(max(ws + sw, 0), W);
  we = min(max(we + sw, 0), W);
  const bool empty = (he <= hs) || (we <= ws);
  float maxval = empty ? 0.0f : -99999999.0f;
  int maxidx = -1;
  const float* plane = feat + ((size_t)bi * C + c) * H * W;
  for (int h = hs; h < he; ++h)
    for (int w = ws; w < we; ++w) {
      const float v = plane[h * W + w];
      if (v > maxval) {
        maxval = v;
        maxidx = h * W + w;
      }
    }
  out[index] = maxval;
  argmax[index] = maxidx;
}

__global__ void __launch_bounds__(256)
roi_pool_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ rois,
                    const int* __restrict__ argmax, size_t total, int C, int H, int W, int PH, int PW,
                    float* __restrict__ grad) {
  const size_t index = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (index >= total) return;
  const int c = (int)((index / PW / PH) % C);
  const int n = (int)(index / PW / PH / C);
  const int a = argmax[index];
  if (a < 0) return;
  const int bi = (int)rois[(size_t)n * 5];
  atomicAdd(grad + ((size_t)bi * C + c) * H * W + a, dout[index]);
}

// ---- EmbedMask mask_prob --------------------------------------------------------------------
// CTA = (slice of the box's pixels, centre): no per-thread linear search for the centre
// (mask_prob.py:25-32); the centre's embedding sits in shared memory.
constexpr int MP_MAXDIM = 128;
__global__ void __launch_bounds__(256)
mask_prob_kernel(const float* __restrict__ embed_pixel, const float* __restrict__ embed_center,
                 const float* __restrict__ sigma_center, const int* __restrict__ boxes,
                 const int* __restrict__ box_areas, int num_pixel, int mask_width, int dim,
                 float* __restrict__ probs) {
  __shared__ float ec[MP_MAXDIM];
  const int n = blockIdx.y;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) ec[d] = embed_center[(size_t)n * dim + d];
  __syncthreads();
  const int area = box_areas[n];
  const int x1 = boxes[4 * n], y1 = boxes[4 * n + 1];
  const int bw = boxes[4 * n + 2] - x1;
  if (bw <= 0) return;
  const float sg = sigma_center[n];
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < area; q += gridDim.x * blockDim.x) {
    const int x = q % bw + x1, y = q / bw + y1;
    const int pid = y * mask_width + x;
    if (pid < 0 || pid >= num_pixel) continue;
    const float* ep = embed_pixel + (size_t)pid * dim;
    float norm2 = 0.0f;
    for (int d = 0; d < dim; ++d) {
      const float t = ep[d] - ec[d];
      norm2 = norm2 + t * t;
    }
    probs[(size_t)n * num_pixel + pid] = rp_expf(-norm2 * sg);
  }
}


// ---- EmbedMask box masks --------------------------------------------------------------------
// boxes_to_masks (boxlist_ops.py:237-261) and crop_by_box (:262-281): a padded-box indicator per
// (instance, pixel); crop_by_box also multiplies a float mask by it.  The two differ in the order
// of the padding arithmetic and in the upper clamp (w / h against w-1 / h-1); `crop` selects.
// CTA = (256-pixel run, instance): rows are written in plane order (coalesced bytes / floats).
__global__ void __launch_bounds__(256)
box_mask_kernel(const float* __restrict__ boxes, int h, int w, float padding, int crop,
                const float* __restrict__ masks_in, float* __restrict__ masks_out,
                uint8_t* __restrict__ ind_out) {
  const int n = blockIdx.y;
  const float* bx = boxes + (size_t)n * 4;
  const float bw = bx[2] - bx[0], bh = bx[3] - bx[1];
  float x1, x2, y1, y2;
  if (crop) {
    x1 = fmaxf(bx[0] - bw * padding - 1.0f, 0.0f);
    x2 = fminf(bx[2] + bw * padding + 1.0f, (float)(w - 1));
    y1 = fmaxf(bx[1] - bh * padding - 1.0f, 0.0f);
    y2 = fminf(bx[3] + bh * padding + 1.0f, (float)(h - 1));
  } else {
    x1 = fmaxf(bx[0] - 1.0f - bw * padding, 0.0f);
    x2 = fminf(bx[2] + 1.0f + bw * padding, (float)w);
    y1 = fmaxf(bx[1] - 1.0f - bh * padding, 0.0f);
    y2 = fminf(bx[3] + 1.0f + bh * padding, (float)h);
  }
  const size_t plane = (size_t)h * w;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane;
       i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / w), c = (int)(i - (size_t)r * w);
    const bool in = ((float)c >= x1) && ((float)c < x2) && ((float)r >= y1) && ((float)r < y2);
    const size_t o = (size_t)n * plane + i;
    if (ind_out) ind_out[o] = in ? 1 : 0;
    if (masks_out) masks_out[o] = masks_in[o] * (in ? 1.0f : 0.0f);
  }
}

// ---- Pose2Seg paste-back: cv2.warpAffine(WARP_INVERSE_MAP + INTER_LINEAR) + "> 0.5" -----------------
// Pose2Seg.jittor/modeling/build_model.py:31-43.  One thread per 4 output pixels of one instance; the
// source map (64 x 64 x 2 floats = 32 KB) stays in L1/L2; OpenCV's fixed-point coordinates and weight
// table (include/roipath_math.h, shared with the oracle, which is pinned against cv2 itself).
// HBM-bound on the output: width*height bytes per instance, written as 4-byte words.
__global__ void __launch_bounds__(256)
affine_paste_kernel(const float* __restrict__ pred, const double* __restrict__ M, int h, int w, int C, int c,
                    int width, int height, uint8_t* __restrict__ out) {
  const int i = blockIdx.y;
  const float* src = pred + (size_t)i * h * w * C;
  double m[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) m[k] = M[6 * (size_t)i + k];
  const size_t plane = (size_t)width * height;
  uint8_t* dst = out + (size_t)i * plane;
  const bool aligned = ((reinterpret_cast<uintptr_t>(dst) & 3) == 0);
  for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g * 4 < plane; g += (size_t)gridDim.x * blockDim.x) {
    uint32_t packed = 0;
    const size_t p0 = g * 4;
    const int nv = plane - p0 < 4 ? (int)(plane - p0) : 4;
    int y = (int)(p0 / width), x = (int)(p0 - (size_t)y * width);
    // rp_warp_affine_coord with its per-row terms (X0, Y0) hoisted: they change only when the row does
    int X0 = rp_cv_round((m[1] * (double)y + m[2]) * 1024.0) + 16;
    int Y0 = rp_cv_round((m[4] * (double)y + m[5]) * 1024.0) + 16;
    for (int j = 0; j < nv; ++j) {
      const int X = (X0 + rp_cv_round(m[0] * (double)x * 1024.0)) >> 5;
      const int Y = (Y0 + rp_cv_round(m[3] * (double)x * 1024.0)) >> 5;
      const float v = rp_warp_affine_sample(src, h, w, C, c, rp_sat_short(X >> 5), rp_sat_short(Y >> 5), X & 31, Y & 31);
      packed |= (v > 0.5f ? 1u : 0u) << (8 * j);
      if (++x == width) {
        x = 0;
        ++y;
        X0 = rp_cv_round((m[1] * (double)y + m[2]) * 1024.0) + 16;
        Y0 = rp_cv_round((m[4] * (double)y + m[5]) * 1024.0) + 16;
      }
    }
    if (aligned && nv == 4) {
      __stcs(reinterpret_cast<uint32_t*>(dst + p0), packed);
    } else {
      for (int j = 0; j < nv; ++j) dst[p0 + j] = (uint8_t)(packed >> (8 * j));
    }
  }
}

// ---- EmbedMask mask stage, fused -------------------------------------------------
// det/modeling/rpn/embed_mask/inference.py:153-169 (compute_mask_prob) and :193-197 (forward_for_mask):
//   mask_boxes = boxes / stride; b = int(mask_boxes) grown by 2 and clamped to the map (:157-161);
//   prob[n][pixel] = exp(-|e_pixel - e_center|^2 * margin) inside b, 0 elsewhere (layers/mask_prob.py:36-57);
//   masks = prob * boxes_to_masks(mask_boxes, m_h, m_w)   (boxlist_ops.py:237-261, padding 0).
// The reference transposes pixel_embed to [HW, dim] first and makes three passes over n x HW floats;
// here pixel_embed stays [dim, m_h, m_w] (lanes along x read each embedding plane coalesced), the
// output is zero-filled once and only the grown box of each instance is visited.
__global__ void __launch_bounds__(256)
embed_masks_kernel(const float* __restrict__ pixel_embed, const float* __restrict__ embed_center,
                   const float* __restrict__ margin, const float* __restrict__ boxes, float inv_stride_num,
                   float stride, int dim, int m_h, int m_w, float* __restrict__ masks) {
  __shared__ float ec[MP_MAXDIM];
  const int n = blockIdx.y;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) ec[d] = embed_center[(size_t)n * dim + d];
  __syncthreads();
  (void)inv_stride_num;
  const float fx1 = boxes[4 * n] / stride, fy1 = boxes[4 * n + 1] / stride;
  const float fx2 = boxes[4 * n + 2] / stride, fy2 = boxes[4 * n + 3] / stride;
  // int() truncation, +-2, clamps (:157-161)
  int x1 = (int)fx1 - 2, y1 = (int)fy1 - 2, x2 = (int)fx2 + 2, y2 = (int)fy2 + 2;
  x1 = x1 < 0 ? 0 : x1;
  y1 = y1 < 0 ? 0 : y1;
  x2 = x2 > m_w ? m_w : x2;
  y2 = y2 > m_h ? m_h : y2;
  const int bw = x2 - x1, bh = y2 - y1;
  if (bw <= 0 || bh <= 0) return;
  // boxes_to_masks indicator on the float mask boxes
  const float ix1 = fmaxf(fx1 - 1.0f, 0.0f), ix2 = fminf(fx2 + 1.0f, (float)m_w);
  const float iy1 = fmaxf(fy1 - 1.0f, 0.0f), iy2 = fminf(fy2 + 1.0f, (float)m_h);
  const float sg = margin[n];
  const size_t HW = (size_t)m_h * m_w;
  const int area = bw * bh;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < area; q += gridDim.x * blockDim.x) {
    const int yy = q / bw, x = q - yy * bw + x1, y = yy + y1;
    const size_t pid = (size_t)y * m_w + x;
    float norm2 = 0.0f;
    for (int d = 0; d < dim; ++d) {
      const float t = pixel_embed[(size_t)d * HW + pid] - ec[d];
      norm2 = norm2 + t * t;
    }
    const bool in = ((float)x >= ix1) && ((float)x < ix2) && ((float)y >= iy1) && ((float)y < iy2);
    masks[(size_t)n * HW + pid] = rp_expf(-norm2 * sg) * (in ? 1.0f : 0.0f);
  }
}

// masks [n,h,w] float -> bilinear x (H/h, W/w), align_corners=False -> > thresh -> [:crop_h, :crop_w] as
// uint8 (inference.py:199-202: nn.interpolate(scale_factor=stride) > mask_th, cropped to the input size).
// One thread per 4 output pixels; HBM-bound on the output bytes.
__global__ void __launch_bounds__(256)
resize_threshold_kernel(const float* __restrict__ masks, int h, int w, int H, int W, float thresh, int crop_h,
                        int crop_w, uint8_t* __restrict__ out) {
  const int n = blockIdx.y;
  const float* img = masks + (size_t)n * h * w;
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const size_t plane = (size_t)crop_h * crop_w;
  uint8_t* dst = out + (size_t)n * plane;
  const bool aligned = ((reinterpret_cast<uintptr_t>(dst) & 3) == 0);
  for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g * 4 < plane; g += (size_t)gridDim.x * blockDim.x) {
    const size_t p0 = g * 4;
    const int nv = plane - p0 < 4 ? (int)(plane - p0) : 4;
    int Y = (int)(p0 / crop_w), X = (int)(p0 - (size_t)Y * crop_w);
    uint32_t packed = 0;
    for (int j = 0; j < nv; ++j) {
      const float fy = rp_resize_src(Y, sy), fx = rp_resize_src(X, sx);
      int y0 = (int)fy, x0 = (int)fx;
      if (y0 > h - 1) y0 = h - 1;
      if (x0 > w - 1) x0 = w - 1;
      const int y1 = y0 + (y0 < h - 1 ? 1 : 0), x1 = x0 + (x0 < w - 1 ? 1 : 0);
      const float h1 = fy - (float)y0, h0 = 1.0f - h1, w1 = fx - (float)x0, w0 = 1.0f - w1;
      const float top = w0 * img[y0 * w + x0] + w1 * img[y0 * w + x1];
      const float bot = w0 * img[y1 * w + x0] + w1 * img[y1 * w + x1];
      packed |= ((h0 * top + h1 * bot) > thresh ? 1u : 0u) << (8 * j);
      if (++X == crop_w) {
        X = 0;
        ++Y;
      }
    }
    if (aligned && nv == 4) {
      __stcs(reinterpret_cast<uint32_t*>(dst + p0), packed);
    } else {
      for (int j = 0; j < nv; ++j) dst[p0 + j] = (uint8_t)(packed >> (8 * j));
    }
  }
}

}  // namespace

extern "C" {

int rp_boxlist_iou(const float* boxes1, int N, const float* boxes2, int M, float* out, void* stream) {
  if (N == 0 || M == 0) return RP_OK;
  if (!boxes1 || !boxes2 || !out || N < 0 || M < 0) return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(boxes1) | reinterpret_cast<uintptr_t>(boxes2)) & 15) != 0)
    return RP_E_ARG;
  const dim3 grid((unsigned)((M + 31) / 32), (unsigned)((N + 31) / 32));
  if (grid.y > 65535) return RP_E_LIMIT;
  iou_matrix_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(boxes1, N, boxes2, M, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_roi_pool_forward(const float* feat, int N, int C, int H, int W, const float* rois, int K, int PH,
                        int PW, float spatial_scale, float* out, int* argmax, void* stream) {
  if (K == 0) return RP_OK;
  if (!feat || !rois || !out || !argmax || N < 1 || C < 1 || H < 1 || W < 1 || K < 0 || PH < 1 ||
      PW < 1)
    return RP_E_ARG;
  if ((size_t)H * W > 0x7fffffffull) return RP_E_LIMIT;
  const size_t total = (size_t)K * C * PH * PW;
  roi_pool_fwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      feat, rois, total, C, H, W, PH, PW, spatial_scale, out, argmax);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_roi_pool_backward(const float* dout, const float* rois, const int* argmax, int K, int N, int C,
                         int H, int W, int PH, int PW, float* grad_input, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (!grad_input || N < 1 || C < 1 || H < 1 || W < 1 || K < 0 || PH < 1 || PW < 1) return RP_E_ARG;
  RP_CUDA_TRY(cudaMemsetAsync(grad_input, 0, sizeof(float) * (size_t)N * C * H * W, st));  // :143
  if (K == 0) return RP_OK;
  if (!dout || !rois || !argmax) return RP_E_ARG;
  const size_t total = (size_t)K * C * PH * PW;
  roi_pool_bwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(dout, rois, argmax, total, C, H,
                                                                       W, PH, PW, grad_input);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_mask_prob(const float* embed_pixel, const float* embed_center, const float* sigma_center,
                 const int* boxes, const int* box_areas, int num_pixel, int num_center, int dim,
                 int mask_width, float* probs, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (num_pixel < 0 || num_center < 0 || dim < 1 || mask_width < 1) return RP_E_ARG;
  if (dim > MP_MAXDIM || num_center > 65535) return RP_E_LIMIT;
  if ((size_t)num_pixel * num_center == 0) return RP_OK;
  if (!embed_pixel || !embed_center || !sigma_center || !boxes || !box_areas || !probs) return RP_E_ARG;
  RP_CUDA_TRY(cudaMemsetAsync(probs, 0, sizeof(float) * (size_t)num_pixel * num_center, st));  // :60
  mask_prob_kernel<<<dim3(32, (unsigned)num_center), 256, 0, st>>>(
      embed_pixel, embed_center, sigma_center, boxes, box_areas, num_pixel, mask_width, dim, probs);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_box_masks(const float* boxes, int n, int h, int w, float padding, int crop,
                 const float* masks_in, float* masks_out, uint8_t* indicator, void* stream) {
  if (n == 0) return RP_OK;
  if (!boxes || n < 0 || h < 1 || w < 1 || n > 65535) return RP_E_ARG;
  if (!indicator && !masks_out) return RP_E_ARG;
  if (masks_out && !masks_in) return RP_E_ARG;
  const size_t plane = (size_t)h * w;
  unsigned gx = (unsigned)((plane + 255) / 256);
  if (gx > 4096) gx = 4096;
  box_mask_kernel<<<dim3(gx, (unsigned)n), 256, 0, (cudaStream_t)stream>>>(boxes, h, w, padding, crop ? 1 : 0,
                                                                         masks_in, masks_out, indicator);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_affine_paste(const float* pred, const double* M, int n, int h, int w, int C, int channel, int width,
                    int height, uint8_t* out, void* stream) {
  if (n == 0) return RP_OK;
  if (!pred || !M || !out || n < 0 || h < 1 || w < 1 || C < 1 || channel < 0 || channel >= C || width < 1 ||
      height < 1)
    return RP_E_ARG;
  if (n > 65535 || h > 32767 || w > 32767) return RP_E_LIMIT;
  const size_t groups = ((size_t)width * height + 3) / 4;
  unsigned gx = (unsigned)((groups + 255) / 256);
  if (gx > 2048) gx = 2048;
  affine_paste_kernel<<<dim3(gx, (unsigned)n), 256, 0, (cudaStream_t)stream>>>(pred, M, h, w, C, channel, width,
                                                                              height, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_embed_masks(const float* pixel_embed, const float* proposal_embed, const float* proposal_margin,
                   const float* boxes, float stride, int n, int dim, int m_h, int m_w, float* masks,
                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) return RP_OK;
  if (!pixel_embed || !proposal_embed || !proposal_margin || !boxes || !masks || n < 0 || dim < 1 || m_h < 1 ||
      m_w < 1 || !(stride > 0))
    return RP_E_ARG;
  if (dim > MP_MAXDIM || n > 65535) return RP_E_LIMIT;
  RP_CUDA_TRY(cudaMemsetAsync(masks, 0, sizeof(float) * (size_t)n * m_h * m_w, st));
  embed_masks_kernel<<<dim3(16, (unsigned)n), 256, 0, st>>>(pixel_embed, proposal_embed, proposal_margin, boxes,
                                                            0.f, stride, dim, m_h, m_w, masks);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_resize_threshold(const float* masks, int n, int h, int w, int H, int W, float thresh, int crop_h,
                        int crop_w, uint8_t* out, void* stream) {
  if (n == 0) return RP_OK;
  if (!masks || !out || n < 0 || h < 1 || w < 1 || H < 1 || W < 1 || crop_h < 1 || crop_w < 1 || crop_h > H ||
      crop_w > W)
    return RP_E_ARG;
  if (n > 65535) return RP_E_LIMIT;
  const size_t groups = ((size_t)crop_h * crop_w + 3) / 4;
  unsigned gx = (unsigned)((groups + 255) / 256);
  if (gx > 2048) gx = 2048;
  resize_threshold_kernel<<<dim3(gx, (unsigned)n), 256, 0, (cudaStream_t)stream>>>(masks, h, w, H, W, thresh,
                                                                                  crop_h, crop_w, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
