// rpn.cu -- the RPN proposal stage in front of the NMS for sm_100a.
//
// Reference being replaced (paths relative to the reference checkout):
//   det/modeling/rpn/inference.py:83-117   RPNPostProcessor.forward_for_single_feature_map, up to
//                                          (not including) boxlist_nms
//   det/modeling/rpn/utils.py:10-15        permute_and_flatten: (N,A*C,H,W) -> (N, H*W*A, C)
//   det/modeling/box_coder.py:53-94        BoxCoder.decode (legacy +1 widths, -1 on x2 / y2)
//   det/structures/bounding_box.py:263-277 clip_to_image(remove_empty=False)
//   det/structures/boxlist_ops.py:64-76    remove_small_boxes
//
// The reference runs ~12 tensor ops per level (permute, sigmoid, topk, two fancy-index gathers,
// decode, then per image clip / where / index).  Here, per level and for the whole batch:
//   1. keys    64-bit key = (image << 32) | order-reversed bits of sigmoid(logit); the value is
//              the flat anchor index in the reference's (h, w, a) order
//   2. sort    one stable radix sort (cub) = topk(sorted=True) with ties by ascending index
//   3. decode  one CTA per image walks its first k ranks in order: gather regression + anchor
//              straight from the NCHW head outputs, decode, clip, flag small boxes, ordered
//              compaction (block scan) into the image's slots
//   4. pack    offsets = exclusive scan of the per-image counts; boxes / scores / anchor indices
//              are packed contiguously so that rp_nms_batched consumes them as they are.
// No host synchronisation; everything downstream (NMS) reads device-side offsets.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int RPN_THREADS = 256;

__device__ __forceinline__ uint32_t rpn_desc_bits(float s) {
  if (s == 0.0f) s = 0.0f;
  uint32_t u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ~u;
}

// box_coder.py:53-94 for one box, operation for operation (no FMA contraction: --fmad=false)
__device__ __forceinline__ float4 rpn_decode_one(float4 a, float4 d, float wx, float wy, float ww,
                                                 float wh, float clip) {
  const float widths = a.z - a.x + 1.0f;
  const float heights = a.w - a.y + 1.0f;
  const float ctr_x = a.x + 0.5f * widths;
  const float ctr_y = a.y + 0.5f * heights;
  const float dx = d.x / wx, dy = d.y / wy;
  float dw = d.z / ww, dh = d.w / wh;
  dw = fminf(dw, clip);
  dh = fminf(dh, clip);
  const float pcx = dx * widths + ctr_x;
  const float pcy = dy * heights + ctr_y;
  const float pw = rp_expf(dw) * widths;
  const float ph = rp_expf(dh) * heights;
  float4 o;
  o.x = pcx - 0.5f * pw;
  o.y = pcy - 0.5f * ph;
  o.z = pcx + 0.5f * pw - 1.0f;
  o.w = pcy + 0.5f * ph - 1.0f;
  return o;
}

// rel_codes [n, 4K] (K boxes per reference box: the class-specific regression of the box head,
// box_coder.py:73-76 strides 0::4 .. 3::4), boxes [n, 4]
__global__ void rpn_box_decode_kernel(const float* __restrict__ codes, const float* __restrict__ boxes,
                                      size_t nk, int K, float wx, float wy, float ww, float wh,
                                      float clip, float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nk) return;
  const float4 d = *reinterpret_cast<const float4*>(codes + i * 4);
  const float4 a = *reinterpret_cast<const float4*>(boxes + (i / K) * 4);
  *reinterpret_cast<float4*>(out + i * 4) = rpn_decode_one(a, d, wx, wy, ww, wh, clip);
}

// element e of image b in the reference's flattened order is (h, w, a) = (e / (W*A), ., e % A);
// its logit lives at objectness[b][a][h][w].
__global__ void rpn_keys_kernel(const float* __restrict__ objectness, int N, int A, int HW,
                                uint64_t* __restrict__ keys, int* __restrict__ vals) {
  const size_t per = (size_t)A * HW;
  const size_t f = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= per * N) return;
  // threads walk the NCHW tensor (coalesced reads); the key goes to the element's flattened slot
  const int b = (int)(f / per);
  const size_t r = f - (size_t)b * per;
  const int a = (int)(r / HW);
  const int hw = (int)(r - (size_t)a * HW);
  const int e = hw * A + a;
  const float s = rp_sigmoidf(objectness[f]);
  const size_t o = (size_t)b * per + e;
  keys[o] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)rpn_desc_bits(s);
  vals[o] = e;
}

__device__ __forceinline__ int rpn_block_excl_scan(int v, int* warp_sums, int* total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += t;
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < RPN_THREADS / 32; ++w) {
    const int sw = warp_sums[w];
    if (w < wid) base += sw;
    tot += sw;
  }
  __syncthreads();
  *total = tot;
  return base + inc - v;
}

// One CTA per image: ranks 0..k-1 in order -> padded slots [b*k + j], count[b].
__global__ void __launch_bounds__(RPN_THREADS)
rpn_decode_kernel(const float* __restrict__ objectness, const float* __restrict__ regression,
                  const float* __restrict__ anchors, const float* __restrict__ im_wh,
                  const int* __restrict__ sorted_vals, int A, int HW, int k, float min_size,
                  float wx, float wy, float ww, float wh, float clip, float* __restrict__ pad_boxes,
                  float* __restrict__ pad_scores, int* __restrict__ pad_index,
                  int* __restrict__ counts) {
  __shared__ int wsum[RPN_THREADS / 32];
  const int b = blockIdx.x;
  const size_t per = (size_t)A * HW;
  const float im_w = im_wh[2 * b], im_h = im_wh[2 * b + 1];
  const float* obj_b = objectness + (size_t)b * per;
  const float* reg_b = regression + (size_t)b * per * 4;
  const float* anc_b = anchors + (size_t)b * per * 4;
  int base = 0;
  for (int j0 = 0; j0 < k; j0 += RPN_THREADS) {
    const int j = j0 + threadIdx.x;
    int ok = 0, e = 0;
    float4 bx = make_float4(0.f, 0.f, 0.f, 0.f);
    float sc = 0.f;
    if (j < k) {
      e = sorted_vals[(size_t)b * per + j];
      const int hw = e / A, a = e - hw * A;
      sc = rp_sigmoidf(obj_b[(size_t)a * HW + hw]);
      float4 d;  // box_regression[b][a*4 + c][h][w]
      d.x = reg_b[((size_t)a * 4 + 0) * HW + hw];
      d.y = reg_b[((size_t)a * 4 + 1) * HW + hw];
      d.z = reg_b[((size_t)a * 4 + 2) * HW + hw];
      d.w = reg_b[((size_t)a * 4 + 3) * HW + hw];
      const float4 an = *reinterpret_cast<const float4*>(anc_b + (size_t)e * 4);
      bx = rpn_decode_one(an, d, wx, wy, ww, wh, clip);
      // clip_to_image: clamp(min 0, max size - 1)
      bx.x = fminf(fmaxf(bx.x, 0.f), im_w - 1.0f);
      bx.y = fminf(fmaxf(bx.y, 0.f), im_h - 1.0f);
      bx.z = fminf(fmaxf(bx.z, 0.f), im_w - 1.0f);
      bx.w = fminf(fmaxf(bx.w, 0.f), im_h - 1.0f);
      // remove_small_boxes on the xywh view (legacy +1)
      const float ws = bx.z - bx.x + 1.0f, hs = bx.w - bx.y + 1.0f;
      ok = (ws >= min_size) && (hs >= min_size);
    }
    int tot;
    const int ex = rpn_block_excl_scan(ok, wsum, &tot);
    if (ok) {
      const size_t o = (size_t)b * k + base + ex;
      *reinterpret_cast<float4*>(pad_boxes + o * 4) = bx;
      pad_scores[o] = sc;
      pad_index[o] = e;
    }
    base += tot;
  }
  if (threadIdx.x == 0) counts[b] = base;
}

__global__ void rpn_offsets_kernel(const int* __restrict__ counts, int N, int* __restrict__ offsets) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int acc = 0;
    for (int b = 0; b < N; ++b) {
      offsets[b] = acc;
      acc += counts[b];
    }
    offsets[N] = acc;
  }
}

__global__ void rpn_pack_kernel(const float* __restrict__ pad_boxes, const float* __restrict__ pad_scores,
                                const int* __restrict__ pad_index, const int* __restrict__ offsets,
                                int k, float* __restrict__ out_boxes, float* __restrict__ out_scores,
                                int* __restrict__ out_index) {
  const int b = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int off = offsets[b], cnt = offsets[b + 1] - off;
  if (j >= cnt) return;
  const size_t s = (size_t)b * k + j, o = (size_t)off + j;
  *reinterpret_cast<float4*>(out_boxes + o * 4) = *reinterpret_cast<const float4*>(pad_boxes + s * 4);
  out_scores[o] = pad_scores[s];
  if (out_index) out_index[o] = pad_index[s];
}

struct RpnWorkspace {
  uint64_t *keys_in, *keys_out;
  int *vals_in, *vals_out;
  float *pad_boxes, *pad_scores;
  int *pad_index, *counts;
  void* cub_temp;
  size_t cub_bytes, total;
};

size_t rpn_cub_bytes(size_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int*)nullptr, (int*)nullptr, (int)n);
  return bytes;
}

RpnWorkspace rpn_carve(void* base, size_t n_total, int N, int k) {
  RpnWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (void*)((char*)base + off) : nullptr;
    off += rp_align_up(bytes, 256);
    return p;
  };
  w.keys_in = (uint64_t*)take(8 * n_total);
  w.keys_out = (uint64_t*)take(8 * n_total);
  w.vals_in = (int*)take(4 * n_total);
  w.vals_out = (int*)take(4 * n_total);
  w.pad_boxes = (float*)take(16 * (size_t)N * k);
  w.pad_scores = (float*)take(4 * (size_t)N * k);
  w.pad_index = (int*)take(4 * (size_t)N * k);
  w.counts = (int*)take(4 * (size_t)N);
  w.cub_bytes = rpn_cub_bytes(n_total);
  w.cub_temp = take(w.cub_bytes);
  w.total = off;
  return w;
}

int rpn_bits(size_t v) {
  int b = 1;
  while (b < 32 && ((size_t)1 << b) < v) ++b;
  return b;
}

}  // namespace

extern "C" {

int rp_box_decode(const float* rel_codes, const float* boxes, int n, int K, float wx, float wy,
                  float ww, float wh, float bbox_xform_clip, float* out, void* stream) {
  if (n == 0) return RP_OK;
  if (!rel_codes || !boxes || !out || n < 0 || K < 1) return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(rel_codes) | reinterpret_cast<uintptr_t>(boxes) |
        reinterpret_cast<uintptr_t>(out)) & 15) != 0)
    return RP_E_ARG;
  const size_t nk = (size_t)n * K;
  rpn_box_decode_kernel<<<(unsigned)((nk + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      rel_codes, boxes, nk, K, wx, wy, ww, wh, bbox_xform_clip, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

size_t rp_rpn_pre_nms_workspace_bytes(int N, int A, int H, int W, int k) {
  if (N < 1 || A < 1 || H < 1 || W < 1 || k < 1) return 0;
  return rpn_carve(nullptr, (size_t)N * A * H * W, N, k).total;
}

int rp_rpn_pre_nms(const float* objectness, const float* box_regression, const float* anchors,
                   const float* im_wh, int N, int A, int H, int W, int pre_nms_top_n, float min_size,
                   float wx, float wy, float ww, float wh, float bbox_xform_clip, float* out_boxes,
                   float* out_scores, int* out_index, int* out_offsets, void* workspace,
                   size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (N < 1 || A < 1 || H < 1 || W < 1 || pre_nms_top_n < 1) return RP_E_ARG;
  if (!objectness || !box_regression || !anchors || !im_wh || !out_boxes || !out_scores ||
      !out_offsets || !workspace)
    return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(anchors) | reinterpret_cast<uintptr_t>(out_boxes)) & 15) != 0)
    return RP_E_ARG;
  const size_t per = (size_t)A * H * W;
  const size_t n_total = per * N;
  if (n_total > 0x7fffffffull) return RP_E_LIMIT;
  const int k = (int)((size_t)pre_nms_top_n < per ? (size_t)pre_nms_top_n : per);
  RpnWorkspace w = rpn_carve(workspace, n_total, N, k);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  rpn_keys_kernel<<<(unsigned)((n_total + 255) / 256), 256, 0, st>>>(objectness, N, A, H * W, w.keys_in,
                                                                    w.vals_in);
  RP_LAUNCH_CHECK();
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out, w.vals_in,
                                              w.vals_out, (int)n_total, 0, 32 + rpn_bits((size_t)N), st));
  rpn_decode_kernel<<<N, RPN_THREADS, 0, st>>>(objectness, box_regression, anchors, im_wh, w.vals_out, A,
                                               H * W, k, min_size, wx, wy, ww, wh, bbox_xform_clip,
                                               w.pad_boxes, w.pad_scores, w.pad_index, w.counts);
  RP_LAUNCH_CHECK();
  rpn_offsets_kernel<<<1, 32, 0, st>>>(w.counts, N, out_offsets);
  RP_LAUNCH_CHECK();
  if (N > 65535) return RP_E_LIMIT;
  rpn_pack_kernel<<<dim3((unsigned)((k + 255) / 256), (unsigned)N), 256, 0, st>>>(
      w.pad_boxes, w.pad_scores, w.pad_index, out_offsets, k, out_boxes, out_scores, out_index);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
