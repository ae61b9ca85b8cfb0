// mask_loss.cu -- the training twin of Yolact's mask assembly: lincomb_mask_loss forward + backward.
//
// Reference being replaced (per image, Yolact.jittor/layers/modules/multibox_loss.py:616-660):
//   pred = sigmoid(proto @ coef^T)                      [H, W, n]      (:616-618)
//   pred = crop(pred, gt boxes)   (mask_proto_crop)     box_utils.py:355-378, padding 1
//   bce  = -(t*log(max(p,1e-20)) + (1-t)*log(max(1-p,1e-20))),  p = clamp(pred, 0, 1)   (:8-9, :632)
//   per instance:  loss_i = scale_i * sum_{y,x} bce[y,x,i]      (:645-649 gives scale_i =
//                  H*W / (gt_w*W) / (gt_h*H) with mask_proto_normalize_emulate_roi_pooling, else 1;
//                  the caller folds old_num_pos/num_pos and mask_alpha/H/W into scale_i too)
// and its backward: d proto [H,W,D], d coef [n,D] of  sum_i loss_i.
//
// One launch does both directions (the reference's autograd re-reads pred, mask_t and proto several
// times).  CTA = a tile of LM_PIX pixels.  The tile's prototype rows and ALL coefficients live in
// shared memory; thread = pixel: for every instance the K = D dot product, sigmoid, crop, BCE and the
// logit gradient g = scale_i * ((1-t)*p*[1-p > 1e-20] - t*(1-p)*[p > 1e-20]) (zero outside the crop),
// d proto accumulated in registers (written once, no atomics); g goes to a shared [pixel][instance]
// tile from which thread (i, k) reduces d coef over the tile's pixels and issues ONE atomicAdd per
// (instance, k, CTA).  HBM-bound: mask_t (H*W*n floats) is read once, proto once, d proto written once.
#include "common.cuh"

namespace {

constexpr int LM_PIX = 128;     // pixels per CTA
constexpr int LM_THREADS = 128;
constexpr int LM_MAXD = 64;

__device__ __forceinline__ void lm_sanitize(float a, float b, int img_size, int padding, float* lo,
                                            float* hi) {
  const float x1 = a * (float)img_size, x2 = b * (float)img_size;
  float mn = fminf(x1, x2), mx = fmaxf(x1, x2);
  mn = mn - (float)padding;
  if (mn < 0.f) mn = 0.f;
  mx = mx + (float)padding;
  if (mx > (float)img_size) mx = (float)img_size;
  *lo = mn;
  *hi = mx;
}

template <int D>
__global__ void __launch_bounds__(LM_THREADS)
lincomb_loss_kernel(const float* __restrict__ proto, const float* __restrict__ coef,
                    const float* __restrict__ mask_t, const float* __restrict__ boxes,
                    const float* __restrict__ scale, int H, int W, int n, float* __restrict__ loss,
                    float* __restrict__ d_proto, float* __restrict__ d_coef) {
  extern __shared__ __align__(16) float lm_smem[];
  float* coef_s = lm_smem;                 // [n][D]
  float* crop_s = coef_s + n * D;          // [n][4]: x1, x2, y1, y2 (or the whole map)
  float* scale_s = crop_s + n * 4;         // [n]
  float* lsum_s = scale_s + n;             // [n] loss partials of this CTA
  float* proto_s = lsum_s + n;             // [LM_PIX][D + 1]
  float* g_s = proto_s + LM_PIX * (D + 1); // [LM_PIX][n]  (n odd-padded pitch)
  const int gp = n | 1;
  const int tid = threadIdx.x;
  const int HW = H * W;
  const int p0 = blockIdx.x * LM_PIX;
  const int npix = min(LM_PIX, HW - p0);

  for (int i = tid; i < n * D; i += LM_THREADS) coef_s[i] = coef[i];
  for (int i = tid; i < n; i += LM_THREADS) {
    float x1 = 0.f, x2 = (float)W, y1 = 0.f, y2 = (float)H;
    if (boxes) {
      const float* b = boxes + 4 * i;
      lm_sanitize(b[0], b[2], W, 1, &x1, &x2);
      lm_sanitize(b[1], b[3], H, 1, &y1, &y2);
    }
    crop_s[4 * i + 0] = x1;
    crop_s[4 * i + 1] = x2;
    crop_s[4 * i + 2] = y1;
    crop_s[4 * i + 3] = y2;
    scale_s[i] = scale ? scale[i] : 1.0f;
    lsum_s[i] = 0.f;
  }
  for (int i = tid; i < npix * D; i += LM_THREADS) {
    const int p = i / D, k = i - p * D;
    proto_s[p * (D + 1) + k] = proto[(size_t)p0 * D + i];
  }
  __syncthreads();

  {
    // every thread runs the loop (the warp reduction below needs whole warps); threads past the last
    // pixel of the map work on a clamped pixel and contribute nothing
    const bool active = tid < npix;
    const int pix = active ? p0 + tid : p0;
    const int y = pix / W, x = pix - y * W;
    const float fx = (float)x, fy = (float)y;
    float pr[D], dp[D];
#pragma unroll
    for (int k = 0; k < D; ++k) {
      pr[k] = proto_s[(active ? tid : 0) * (D + 1) + k];
      dp[k] = 0.f;
    }
    const float* mt = mask_t + (size_t)pix * n;
    for (int i = 0; i < n; ++i) {
      const float* c = coef_s + i * D;
      float z = 0.f;
#pragma unroll
      for (int k = 0; k < D; ++k) z = fmaf(pr[k], c[k], z);
      const float t = mt[i];
      const bool inside = fx >= crop_s[4 * i] && fx < crop_s[4 * i + 1] && fy >= crop_s[4 * i + 2] &&
                          fy < crop_s[4 * i + 3];
      const float p = inside ? 1.0f / (1.0f + expf(-z)) : 0.f;
      const float q = 1.0f - p;
      const float bce = -(t * logf(fmaxf(p, 1e-20f)) + (1.0f - t) * logf(fmaxf(q, 1e-20f)));
      const float s = scale_s[i];
      {  // one shared atomic per warp, not per thread (all lanes hit the same word)
        float part = active ? s * bce : 0.f;
#pragma unroll
        for (int d_ = 16; d_; d_ >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d_);
        if ((tid & 31) == 0) atomicAdd(&lsum_s[i], part);
      }
      float g = 0.f;
      if (inside && active) {
        const float gpos = p > 1e-20f ? -t * q : 0.f;           // d/dz of -t log p
        const float gneg = q > 1e-20f ? (1.0f - t) * p : 0.f;   // d/dz of -(1-t) log(1-p)
        g = s * (gpos + gneg);
      }
      if (active) g_s[tid * gp + i] = g;
#pragma unroll
      for (int k = 0; k < D; ++k) dp[k] = fmaf(g, c[k], dp[k]);
    }
    float* dpo = d_proto + (size_t)pix * D;
#pragma unroll
    for (int k = 0; k < D; ++k)
      if (active) proto_s[tid * (D + 1) + k] = dp[k];  // staged for a coalesced write
    (void)dpo;
  }
  __syncthreads();
  // d proto tile: coalesced
  for (int i = tid; i < npix * D; i += LM_THREADS) {
    const int p = i / D, k = i - p * D;
    const float v = proto_s[p * (D + 1) + k];
    d_proto[(size_t)p0 * D + i] = v;
  }
  // loss partials
  for (int i = tid; i < n; i += LM_THREADS) atomicAdd(&loss[i], lsum_s[i]);
  __syncthreads();
  // d coef: thread (i, k) sums g[p][i] * proto[p][k] over the tile; proto must be re-staged (the
  // buffer holds d proto now)
  for (int i = tid; i < npix * D; i += LM_THREADS) {
    const int p = i / D, k = i - p * D;
    proto_s[p * (D + 1) + k] = proto[(size_t)p0 * D + i];
  }
  __syncthreads();
  for (int o = tid; o < n * D; o += LM_THREADS) {
    const int i = o / D, k = o - i * D;
    float acc = 0.f;
    for (int p = 0; p < npix; ++p) acc = fmaf(g_s[p * gp + i], proto_s[p * (D + 1) + k], acc);
    atomicAdd(&d_coef[o], acc);
  }
}

}  // namespace

extern "C" {

int rp_lincomb_mask_loss(const float* proto, const float* coef, const float* mask_t, const float* boxes,
                         const float* scale, int H, int W, int D, int n, float* loss, float* d_proto,
                         float* d_coef, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (H < 1 || W < 1 || D < 1 || n < 0 || !proto || !d_proto) return RP_E_ARG;
  if (n == 0) {
    RP_CUDA_TRY(cudaMemsetAsync(d_proto, 0, sizeof(float) * (size_t)H * W * D, st));
    return RP_OK;
  }
  if (!coef || !mask_t || !loss || !d_coef) return RP_E_ARG;
  if (D != 32 && D != 64 && D != 16 && D != 8) return RP_E_LIMIT;
  const size_t smem = sizeof(float) * ((size_t)n * D + 6 * (size_t)n + (size_t)LM_PIX * (D + 1) +
                                       (size_t)LM_PIX * (n | 1));
  if (smem > 200 * 1024) return RP_E_LIMIT;
  RP_CUDA_TRY(cudaMemsetAsync(loss, 0, sizeof(float) * (size_t)n, st));
  RP_CUDA_TRY(cudaMemsetAsync(d_coef, 0, sizeof(float) * (size_t)n * D, st));
  const int grid = (H * W + LM_PIX - 1) / LM_PIX;
#define LM_LAUNCH(D_)                                                                              \
  do {                                                                                             \
    RP_SET_MAX_SMEM(lincomb_loss_kernel<D_>, smem);                                                 \
    lincomb_loss_kernel<D_><<<grid, LM_THREADS, smem, st>>>(proto, coef, mask_t, boxes, scale, H, W, n, \
                                                            loss, d_proto, d_coef);                \
  } while (0)
  switch (D) {
    case 8: LM_LAUNCH(8); break;
    case 16: LM_LAUNCH(16); break;
    case 32: LM_LAUNCH(32); break;
    default: LM_LAUNCH(64); break;
  }
#undef LM_LAUNCH
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
