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


// ---- all FPN levels of a batch in one set of launches --------------------------------------
// Problem ("segment") s = l * N + b is level l of image b.  Level l holds per_l = A_l*H_l*W_l
// anchors per image and keeps k_l = min(pre_nms_top_n, per_l) of them.  Elements of all segments
// live in one array (segment s at elem_base[l] + b*per_l), their candidate slots in another
// (segment s at slot_base[l] + b*k_l).
struct RpnLevels {
  const float* obj[RP_MAX_LEVELS];
  const float* reg[RP_MAX_LEVELS];
  const float* anc[RP_MAX_LEVELS];
  int A[RP_MAX_LEVELS], HW[RP_MAX_LEVELS], k[RP_MAX_LEVELS];
  int elem_base[RP_MAX_LEVELS + 1];  // prefix of N * per_l
  int slot_base[RP_MAX_LEVELS + 1];  // prefix of N * k_l
  int L, N;
};

// element e of a segment in the reference's flattened order is (h, w, a) = (e / (W*A), ., e % A);
// its logit lives at objectness[b][a][h][w].
__global__ void rpn_keys_kernel(const __grid_constant__ RpnLevels lv, uint64_t* __restrict__ keys,
                                int* __restrict__ vals) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= lv.elem_base[lv.L]) return;
  int l = 0;
  while (l + 1 < lv.L && f >= lv.elem_base[l + 1]) ++l;
  // threads walk each NCHW tensor (coalesced reads); the key goes to the element's flattened slot
  const int A = lv.A[l], HW = lv.HW[l];
  const int per = A * HW;
  const int r0 = f - lv.elem_base[l];
  const int b = r0 / per;
  const int r = r0 - b * per;
  const int a = r / HW;
  const int hw = r - a * HW;
  const int e = hw * A + a;
  const float s = rp_sigmoidf(lv.obj[l][r0]);
  const int o = lv.elem_base[l] + b * per + e;
  keys[o] = ((uint64_t)(uint32_t)(l * lv.N + b) << 32) | (uint64_t)rpn_desc_bits(s);
  vals[o] = e;
}

template <int T>
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
  for (int w = 0; w < T / 32; ++w) {
    const int sw = warp_sums[w];
    if (w < wid) base += sw;
    tot += sw;
  }
  __syncthreads();
  *total = tot;
  return base + inc - v;
}

// One CTA per segment: ranks 0..k-1 in order -> the segment's slots, counts[s].
__global__ void __launch_bounds__(RPN_THREADS)
rpn_decode_kernel(const __grid_constant__ RpnLevels lv, const float* __restrict__ im_wh,
                  const int* __restrict__ sorted_vals, float min_size, float wx, float wy, float ww,
                  float wh, float clip, float* __restrict__ pad_boxes,
                  float* __restrict__ pad_scores, int* __restrict__ pad_index,
                  int* __restrict__ counts) {
  __shared__ int wsum[RPN_THREADS / 32];
  const int s = blockIdx.x;
  const int l = s / lv.N, b = s - l * lv.N;
  const int A = lv.A[l], HW = lv.HW[l], k = lv.k[l];
  const size_t per = (size_t)A * HW;
  const float im_w = im_wh[2 * b], im_h = im_wh[2 * b + 1];
  const float* obj_b = lv.obj[l] + (size_t)b * per;
  const float* reg_b = lv.reg[l] + (size_t)b * per * 4;
  const float* anc_b = lv.anc[l] + (size_t)b * per * 4;
  const int* vals_s = sorted_vals + lv.elem_base[l] + (size_t)b * per;
  const size_t slot0 = (size_t)lv.slot_base[l] + (size_t)b * k;
  int base = 0;
  for (int j0 = 0; j0 < k; j0 += RPN_THREADS) {
    const int j = j0 + threadIdx.x;
    int ok = 0, e = 0;
    float4 bx = make_float4(0.f, 0.f, 0.f, 0.f);
    float sc = 0.f;
    if (j < k) {
      e = vals_s[j];
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
    const int ex = rpn_block_excl_scan<RPN_THREADS>(ok, wsum, &tot);
    if (ok) {
      const size_t o = slot0 + base + ex;
      *reinterpret_cast<float4*>(pad_boxes + o * 4) = bx;
      pad_scores[o] = sc;
      pad_index[o] = e;
    }
    base += tot;
  }
  if (threadIdx.x == 0) counts[s] = base;
}

// offsets[0..S] = exclusive scan of counts[0..S) (one CTA; S = L*N is small)
__global__ void __launch_bounds__(RPN_THREADS)
rpn_offsets_kernel(const int* __restrict__ counts, int S, int* __restrict__ offsets) {
  __shared__ int wsum[RPN_THREADS / 32];
  int base = 0;
  for (int s0 = 0; s0 < S; s0 += RPN_THREADS) {
    const int s = s0 + threadIdx.x;
    const int v = s < S ? counts[s] : 0;
    int tot;
    const int ex = rpn_block_excl_scan<RPN_THREADS>(v, wsum, &tot);
    if (s < S) offsets[s] = base + ex;
    base += tot;
  }
  if (threadIdx.x == 0) offsets[S] = base;
}

__global__ void rpn_pack_kernel(const __grid_constant__ RpnLevels lv,
                                const float* __restrict__ pad_boxes,
                                const float* __restrict__ pad_scores,
                                const int* __restrict__ pad_index, const int* __restrict__ offsets,
                                float* __restrict__ out_boxes, float* __restrict__ out_scores,
                                int* __restrict__ out_index) {
  const int s = blockIdx.y;
  const int l = s / lv.N, b = s - l * lv.N;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int off = offsets[s], cnt = offsets[s + 1] - off;
  if (j >= cnt) return;
  const size_t src = (size_t)lv.slot_base[l] + (size_t)b * lv.k[l] + j, o = (size_t)off + j;
  *reinterpret_cast<float4*>(out_boxes + o * 4) =
      *reinterpret_cast<const float4*>(pad_boxes + src * 4);
  out_scores[o] = pad_scores[src];
  if (out_index) out_index[o] = pad_index[src];
}

// ---- per-image top-k over the levels' NMS survivors (inference.py:175-183) ------------------
// seg_pos[s]: where segment s = (l, b) starts in the image-major concatenation
// (cat_boxlist over levels, inference.py:143-144); img_off[b]: where image b starts.
__global__ void __launch_bounds__(RPN_THREADS)
rpn_sel_scan_kernel(const int* __restrict__ n_keep, int L, int N, int* __restrict__ img_off,
                    int* __restrict__ seg_pos) {
  __shared__ int wsum[RPN_THREADS / 32];
  int base = 0;
  for (int b0 = 0; b0 < N; b0 += RPN_THREADS) {
    const int b = b0 + threadIdx.x;
    int v = 0;
    if (b < N)
      for (int l = 0; l < L; ++l) v += n_keep[l * N + b];
    int tot;
    const int ex = rpn_block_excl_scan<RPN_THREADS>(v, wsum, &tot);
    if (b < N) {
      int p = base + ex;
      img_off[b] = p;
      for (int l = 0; l < L; ++l) {
        seg_pos[l * N + b] = p;
        p += n_keep[l * N + b];
      }
    }
    base += tot;
  }
  if (threadIdx.x == 0) img_off[N] = base;
}

__global__ void rpn_sel_keys_kernel(const float* __restrict__ scores, const int* __restrict__ offsets,
                                    const int* __restrict__ keep, const int* __restrict__ img_off,
                                    const int* __restrict__ seg_pos, int L, int N, int n_cap,
                                    uint64_t* __restrict__ keys, int* __restrict__ vals) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_cap) return;
  if (p >= img_off[N]) {  // slack behind the last image
    keys[p] = (uint64_t)(uint32_t)N << 32;
    vals[p] = 0;
    return;
  }
  int lo = 0, hi = N - 1;  // largest b with img_off[b] <= p
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (img_off[mid] <= p)
      lo = mid;
    else
      hi = mid - 1;
  }
  const int b = lo;
  int l = 0;  // last level whose start is <= p (empty levels share a start: take the last)
  for (int q = 1; q < L; ++q)
    if (seg_pos[q * N + b] <= p) l = q;
  const int s = l * N + b;
  const int src = offsets[s] + keep[offsets[s] + (p - seg_pos[s])];
  keys[p] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)rpn_desc_bits(scores[src]);
  vals[p] = src;
}

__global__ void rpn_sel_emit_kernel(const float* __restrict__ boxes, const float* __restrict__ scores,
                                    const int* __restrict__ img_off, const int* __restrict__ sorted_vals,
                                    int top_n, float* __restrict__ out_boxes,
                                    float* __restrict__ out_scores, int* __restrict__ out_count) {
  const int b = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int off = img_off[b];
  int cnt = img_off[b + 1] - off;
  if (cnt > top_n) cnt = top_n;
  if (j == 0) out_count[b] = cnt;
  if (j >= cnt) return;
  const int src = sorted_vals[off + j];
  const size_t o = (size_t)b * top_n + j;
  *reinterpret_cast<float4*>(out_boxes + o * 4) = *reinterpret_cast<const float4*>(boxes + (size_t)src * 4);
  out_scores[o] = scores[src];
}

size_t rpn_cub_bytes(size_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int*)nullptr, (int*)nullptr, (int)n);
  return bytes;
}

struct RpnWorkspace {
  uint64_t *keys_in, *keys_out;
  int *vals_in, *vals_out;
  float *pad_boxes, *pad_scores;
  int *pad_index, *counts;
  void* cub_temp;
  size_t cub_bytes, total;
};

RpnWorkspace rpn_carve(void* base, size_t n_elems, size_t n_slots, int S) {
  RpnWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (void*)((char*)base + off) : nullptr;
    off += rp_align_up(bytes, 256);
    return p;
  };
  w.keys_in = (uint64_t*)take(8 * n_elems);
  w.keys_out = (uint64_t*)take(8 * n_elems);
  w.vals_in = (int*)take(4 * n_elems);
  w.vals_out = (int*)take(4 * n_elems);
  w.pad_boxes = (float*)take(16 * n_slots);
  w.pad_scores = (float*)take(4 * n_slots);
  w.pad_index = (int*)take(4 * n_slots);
  w.counts = (int*)take(4 * (size_t)S);
  w.cub_bytes = rpn_cub_bytes(n_elems);
  w.cub_temp = take(w.cub_bytes);
  w.total = off;
  return w;
}

struct RpnSelWorkspace {
  uint64_t *keys_in, *keys_out;
  int *vals_in, *vals_out, *img_off, *seg_pos;
  void* cub_temp;
  size_t cub_bytes, total;
};

RpnSelWorkspace rpn_sel_carve(void* base, size_t n_cap, int L, int N) {
  RpnSelWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (void*)((char*)base + off) : nullptr;
    off += rp_align_up(bytes, 256);
    return p;
  };
  w.keys_in = (uint64_t*)take(8 * n_cap);
  w.keys_out = (uint64_t*)take(8 * n_cap);
  w.vals_in = (int*)take(4 * n_cap);
  w.vals_out = (int*)take(4 * n_cap);
  w.img_off = (int*)take(4 * (size_t)(N + 1));
  w.seg_pos = (int*)take(4 * (size_t)L * N);
  w.cub_bytes = rpn_cub_bytes(n_cap);
  w.cub_temp = take(w.cub_bytes);
  w.total = off;
  return w;
}

int rpn_bits(size_t v) {  // bits needed for ids 0 .. v-1
  int b = 1;
  while (b < 31 && ((size_t)1 << b) < v) ++b;
  return b;
}

// sizes of a set of levels; returns 0 when an argument is out of range
int rpn_fill_levels(RpnLevels& lv, int L, int N, const int* A, const int* H, const int* W,
                    int pre_nms_top_n) {
  if (L < 1 || L > RP_MAX_LEVELS || N < 1 || !A || !H || !W || pre_nms_top_n < 1) return 0;
  lv.L = L;
  lv.N = N;
  size_t eb = 0, sb = 0;
  for (int l = 0; l < L; ++l) {
    if (A[l] < 1 || H[l] < 1 || W[l] < 1) return 0;
    const size_t per = (size_t)A[l] * H[l] * W[l];
    const size_t k = (size_t)pre_nms_top_n < per ? (size_t)pre_nms_top_n : per;
    lv.A[l] = A[l];
    lv.HW[l] = H[l] * W[l];
    lv.k[l] = (int)k;
    lv.elem_base[l] = (int)eb;
    lv.slot_base[l] = (int)sb;
    eb += per * N;
    sb += k * N;
    if (eb > 0x7fffffffull) return -1;
  }
  lv.elem_base[L] = (int)eb;
  lv.slot_base[L] = (int)sb;
  return 1;
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

size_t rp_rpn_pre_nms_levels_workspace_bytes(int L, int N, const int* A, const int* H, const int* W,
                                             int pre_nms_top_n) {
  RpnLevels lv;
  if (rpn_fill_levels(lv, L, N, A, H, W, pre_nms_top_n) != 1) return 0;
  return rpn_carve(nullptr, (size_t)lv.elem_base[L], (size_t)lv.slot_base[L], L * N).total;
}

int rp_rpn_pre_nms_levels(const float* const* objectness, const float* const* box_regression,
                          const float* const* anchors, const int* A, const int* H, const int* W,
                          int L, const float* im_wh, int N, int pre_nms_top_n, float min_size,
                          float wx, float wy, float ww, float wh, float bbox_xform_clip,
                          float* out_boxes, float* out_scores, int* out_index, int* out_offsets,
                          void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  RpnLevels lv;
  const int ok = rpn_fill_levels(lv, L, N, A, H, W, pre_nms_top_n);
  if (ok == 0) return RP_E_ARG;
  if (ok < 0) return RP_E_LIMIT;
  if (!objectness || !box_regression || !anchors || !im_wh || !out_boxes || !out_scores ||
      !out_offsets || !workspace)
    return RP_E_ARG;
  if ((reinterpret_cast<uintptr_t>(out_boxes) & 15) != 0) return RP_E_ARG;
  int kmax = 0;
  for (int l = 0; l < L; ++l) {
    if (!objectness[l] || !box_regression[l] || !anchors[l]) return RP_E_ARG;
    if ((reinterpret_cast<uintptr_t>(anchors[l]) & 15) != 0) return RP_E_ARG;
    lv.obj[l] = objectness[l];
    lv.reg[l] = box_regression[l];
    lv.anc[l] = anchors[l];
    if (lv.k[l] > kmax) kmax = lv.k[l];
  }
  const int S = L * N;
  if (S > 65535) return RP_E_LIMIT;
  const int n_elems = lv.elem_base[L];
  RpnWorkspace w = rpn_carve(workspace, (size_t)n_elems, (size_t)lv.slot_base[L], S);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  rpn_keys_kernel<<<(unsigned)((n_elems + 255) / 256), 256, 0, st>>>(lv, w.keys_in, w.vals_in);
  RP_LAUNCH_CHECK();
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out, w.vals_in,
                                              w.vals_out, n_elems, 0, 32 + rpn_bits((size_t)S), st));
  rpn_decode_kernel<<<S, RPN_THREADS, 0, st>>>(lv, im_wh, w.vals_out, min_size, wx, wy, ww, wh,
                                               bbox_xform_clip, w.pad_boxes, w.pad_scores,
                                               w.pad_index, w.counts);
  RP_LAUNCH_CHECK();
  rpn_offsets_kernel<<<1, RPN_THREADS, 0, st>>>(w.counts, S, out_offsets);
  RP_LAUNCH_CHECK();
  rpn_pack_kernel<<<dim3((unsigned)((kmax + 255) / 256), (unsigned)S), 256, 0, st>>>(
      lv, w.pad_boxes, w.pad_scores, w.pad_index, out_offsets, out_boxes, out_scores, out_index);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

size_t rp_rpn_pre_nms_workspace_bytes(int N, int A, int H, int W, int pre_nms_top_n) {
  return rp_rpn_pre_nms_levels_workspace_bytes(1, N, &A, &H, &W, pre_nms_top_n);
}

int rp_rpn_pre_nms(const float* objectness, const float* box_regression, const float* anchors,
                   const float* im_wh, int N, int A, int H, int W, int pre_nms_top_n, float min_size,
                   float wx, float wy, float ww, float wh, float bbox_xform_clip, float* out_boxes,
                   float* out_scores, int* out_index, int* out_offsets, void* workspace,
                   size_t workspace_bytes, void* stream) {
  return rp_rpn_pre_nms_levels(&objectness, &box_regression, &anchors, &A, &H, &W, 1, im_wh, N,
                               pre_nms_top_n, min_size, wx, wy, ww, wh, bbox_xform_clip, out_boxes,
                               out_scores, out_index, out_offsets, workspace, workspace_bytes, stream);
}

size_t rp_rpn_select_workspace_bytes(int n_cap, int L, int N) {
  if (n_cap < 1 || L < 1 || N < 1) return 0;
  return rpn_sel_carve(nullptr, (size_t)n_cap, L, N).total;
}

int rp_rpn_select_levels(const float* boxes, const float* scores, const int* offsets, const int* keep,
                         const int* n_keep, int L, int N, int n_cap, int top_n, float* out_boxes,
                         float* out_scores, int* out_count, void* workspace, size_t workspace_bytes,
                         void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (L < 1 || N < 1 || n_cap < 1 || top_n < 1 || N > 65535) return RP_E_ARG;
  if (!boxes || !scores || !offsets || !keep || !n_keep || !out_boxes || !out_scores || !out_count ||
      !workspace)
    return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(boxes) | reinterpret_cast<uintptr_t>(out_boxes)) & 15) != 0)
    return RP_E_ARG;
  RpnSelWorkspace w = rpn_sel_carve(workspace, (size_t)n_cap, L, N);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  rpn_sel_scan_kernel<<<1, RPN_THREADS, 0, st>>>(n_keep, L, N, w.img_off, w.seg_pos);
  RP_LAUNCH_CHECK();
  rpn_sel_keys_kernel<<<(unsigned)((n_cap + 255) / 256), 256, 0, st>>>(
      scores, offsets, keep, w.img_off, w.seg_pos, L, N, n_cap, w.keys_in, w.vals_in);
  RP_LAUNCH_CHECK();
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out, w.vals_in,
                                              w.vals_out, n_cap, 0, 32 + rpn_bits((size_t)N + 1), st));
  rpn_sel_emit_kernel<<<dim3((unsigned)((top_n + 255) / 256), (unsigned)N), 256, 0, st>>>(
      boxes, scores, w.img_off, w.vals_out, top_n, out_boxes, out_scores, out_count);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
