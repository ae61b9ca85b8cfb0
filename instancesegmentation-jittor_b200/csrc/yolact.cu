// yolact.cu -- Yolact post-processing front half for sm_100a: SSD box decode and FastNMS.
//
// Reference being replaced (paths relative to the reference checkout):
//   yol/layers/box_utils.py:272-316            decode(loc, priors)
//   yol/layers/functions/detection.py:139-189  Detect.fast_nms  (~15 tensor ops that each
//                                              materialise [80,200,200(,2)] temporaries)
//   yol/layers/functions/detection.py:113-137  Detect.cc_fast_nms
//   yol/layers/box_utils.py:31-78              intersect / jaccard (no +1, relative coords)
//
// FastNMS pipeline for a whole batch of images (no host synchronisation, no temporaries of
// size [C,k,k]):
//   1. keys   64-bit key = ((image*C + class) << 32) | order-reversed score bits
//   2. sort   one stable radix sort (cub) => per (image,class) descending stable order
//   3. class  one CTA per (image,class): top_k boxes into shared memory, thread j scans the
//             rows i<j of its column of the upper-triangular IoU matrix keeping a running max
//             (NaN-propagating, as triu_ + max do), keep = max <= thr
//   4. sort   kept candidates of an image, keyed (image << 32) | score bits, stable => the
//             reference's (class, rank) flattening order among ties
//   5. emit   first max_det per image: boxes, coefficients, classes, scores
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int FN_MAX_TOPK = 256;

__device__ __forceinline__ uint32_t fn_desc_bits(float s) {
  if (s == 0.0f) s = 0.0f;
  uint32_t u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ~u;
}

__device__ __forceinline__ int fn_find_segment(const int* __restrict__ offsets, int B, int i) {
  int lo = 0, hi = B - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (offsets[mid] <= i)
      lo = mid;
    else
      hi = mid - 1;
  }
  return lo;
}

__global__ void yolact_decode_kernel(const float* __restrict__ loc, const float* __restrict__ priors,
                                     int n, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 l = *reinterpret_cast<const float4*>(loc + (size_t)i * 4);
  const float4 p = *reinterpret_cast<const float4*>(priors + (size_t)i * 4);
  const float cx = p.x + l.x * 0.1f * p.z;
  const float cy = p.y + l.y * 0.1f * p.w;
  const float w = p.z * rp_expf(l.z * 0.2f);
  const float h = p.w * rp_expf(l.w * 0.2f);
  const float x1 = cx - w / 2.0f;
  const float y1 = cy - h / 2.0f;
  float4 o;
  o.x = x1;
  o.y = y1;
  o.z = w + x1;
  o.w = h + y1;
  *reinterpret_cast<float4*>(out + (size_t)i * 4) = o;
}

// per-class keys: element f = c * n_total + j
__global__ void fn_keys_kernel(const float* __restrict__ scores, const int* __restrict__ offsets,
                               int B, int C, int n_total, uint64_t* __restrict__ keys,
                               int* __restrict__ vals) {
  const size_t f = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= (size_t)C * n_total) return;
  const int c = (int)(f / n_total);
  const int j = (int)(f - (size_t)c * n_total);
  if (j >= offsets[B]) {  // n_total may be an upper bound: slack sorts behind every problem
    keys[f] = (uint64_t)(uint32_t)(B * C) << 32;
    vals[f] = 0;
    return;
  }
  const int b = fn_find_segment(offsets, B, j);
  keys[f] = ((uint64_t)(uint32_t)(b * C + c) << 32) | (uint64_t)fn_desc_bits(scores[f]);
  vals[f] = j - offsets[b];
}

// cross-class keys: per candidate max / first argmax over classes
__global__ void fn_cc_keys_kernel(const float* __restrict__ scores, const int* __restrict__ offsets,
                                  int B, int C, int n_total, uint64_t* __restrict__ keys,
                                  int* __restrict__ vals, float* __restrict__ best_out,
                                  int* __restrict__ cls_out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_total) return;
  if (j >= offsets[B]) {  // n_total may be an upper bound (rp_yolact_prefilter's cand_cap)
    keys[j] = (uint64_t)(uint32_t)B << 32;
    vals[j] = 0;
    return;
  }
  float best = scores[j];
  int bc = 0;
  for (int c = 1; c < C; ++c) {
    const float v = scores[(size_t)c * n_total + j];
    if (v > best) {
      best = v;
      bc = c;
    }
  }
  const int b = fn_find_segment(offsets, B, j);
  keys[j] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)fn_desc_bits(best);
  vals[j] = j - offsets[b];
  best_out[j] = best;
  cls_out[j] = bc;
}

// Column max of triu(jaccard, 1) for the first k sorted boxes; returns whether rank r is kept.
// bx: shared [k] boxes, ar: shared [k] areas.
__device__ __forceinline__ bool fn_keep_rank(const float4* bx, const float* ar, int r,
                                             float iou_thr) {
  const float4 q = bx[r];
  const float qa = ar[r];
  float m = 0.f;
  bool nan = false;
  for (int i = 0; i < r; ++i) {
    const float4 p = bx[i];
    const float v = rp_jaccard(p.x, p.y, p.z, p.w, ar[i], q.x, q.y, q.z, q.w, qa);
    if (v != v) nan = true;
    if (v > m) m = v;
  }
  return !nan && (m <= iou_thr);
}

// grid (C, B), FN_MAX_TOPK threads.
__global__ void __launch_bounds__(FN_MAX_TOPK)
fn_class_kernel(const float* __restrict__ boxes, const float* __restrict__ scores,
                const int* __restrict__ offsets, const int* __restrict__ sorted_vals, int C,
                int n_total, int top_k, float iou_thr, int second_threshold, float conf_thresh,
                uint64_t* __restrict__ keys2, int* __restrict__ vals2,
                float* __restrict__ cand_score, int* __restrict__ cand_src) {
  __shared__ float4 bx[FN_MAX_TOPK];
  __shared__ float ar[FN_MAX_TOPK];
  const int c = blockIdx.x, b = blockIdx.y;
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
  const int k = min(n, top_k);
  const int r = threadIdx.x;
  const size_t seg = (size_t)C * off + (size_t)c * n;  // start of (b,c) in the sorted arrays
  int src = 0;
  float sc = 0.f;
  if (r < k) {
    src = sorted_vals[seg + r];
    const float4 p = *reinterpret_cast<const float4*>(boxes + (size_t)(off + src) * 4);
    bx[r] = p;
    ar[r] = rp_area_rel(p.x, p.y, p.z, p.w);
    sc = scores[(size_t)c * n_total + off + src];
  }
  __syncthreads();
  if (r >= top_k) return;
  bool keep = false;
  if (r < k) {
    keep = fn_keep_rank(bx, ar, r, iou_thr);
    if (second_threshold) keep = keep && (sc > conf_thresh);
  }
  const size_t o = ((size_t)b * C + c) * top_k + r;
  keys2[o] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)fn_desc_bits(keep ? sc : -INFINITY);
  vals2[o] = c * top_k + r;
  cand_score[o] = keep ? sc : -INFINITY;
  cand_src[o] = src;
}

// ------------------------------------------------------------------------------------------
// Fused top-k selection + FastNMS of one (class, image): replaces keys + batch-wide radix sort +
// fn_class_kernel (the sort of all C * n scores was 60 % of FastNMS).  The CTA
//   1. loads the class's n scores as order-reversed keys (shared-memory cache for n <= FN_CACHE),
//   2. finds the top_k-th smallest key T by a 4-pass byte-wise radix select (256-bin shared
//      histogram, block scan over the bins),
//   3. collects every key < T and, in ascending candidate order, the first (top_k - #less) keys
//      == T  -- exactly the first top_k entries of a stable descending sort,
//   4. bitonic-sorts the <= 256 (key, candidate) pairs, and
//   5. continues as fn_class_kernel: boxes to shared memory, column max of triu(IoU), keep.
// ------------------------------------------------------------------------------------------
constexpr int FN_CACHE = 6144;

__device__ __forceinline__ int fn_block_excl_scan(int v, int* warp_sums, int* total) {
  // exclusive scan of one int per thread over FN_MAX_TOPK (256) threads
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
  for (int w = 0; w < FN_MAX_TOPK / 32; ++w) {
    const int sw = warp_sums[w];
    if (w < wid) base += sw;
    tot += sw;
  }
  __syncthreads();
  *total = tot;
  return base + inc - v;
}

__global__ void __launch_bounds__(FN_MAX_TOPK)
fn_class_select_kernel(const float* __restrict__ boxes, const float* __restrict__ scores,
                       const int* __restrict__ offsets, int C, int n_total, int top_k, float iou_thr,
                       int second_threshold, float conf_thresh, uint64_t* __restrict__ keys2,
                       int* __restrict__ vals2, float* __restrict__ cand_score,
                       int* __restrict__ cand_src) {
  __shared__ uint32_t kc[FN_CACHE];
  __shared__ int hist[256];
  __shared__ int wsum[FN_MAX_TOPK / 32];
  __shared__ unsigned long long sel[FN_MAX_TOPK];
  __shared__ float4 bx[FN_MAX_TOPK];
  __shared__ float ar[FN_MAX_TOPK];
  __shared__ int s_bin, s_kk, s_cnt;
  const int c = blockIdx.x, b = blockIdx.y;
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
  const int k = min(n, top_k);
  const int tid = threadIdx.x;
  const float* sc_c = scores + (size_t)c * n_total + off;
  const bool cached = n <= FN_CACHE;
  if (cached)
    for (int j = tid; j < n; j += FN_MAX_TOPK) kc[j] = fn_desc_bits(sc_c[j]);
  sel[tid] = ~0ull;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  auto key_at = [&](int j) -> uint32_t { return cached ? kc[j] : fn_desc_bits(sc_c[j]); };

  if (n <= top_k) {
    for (int j = tid; j < n; j += FN_MAX_TOPK) sel[j] = ((unsigned long long)key_at(j) << 32) | (uint32_t)j;
  } else {
    // ---- radix select: T = the k-th smallest key, kk = how many keys == T belong to the top k
    uint32_t prefix = 0, mask = 0;
    int kk = k;
    for (int pass = 3; pass >= 0; --pass) {
      hist[tid] = 0;
      __syncthreads();
      const int sh = 8 * pass;
      for (int j = tid; j < n; j += FN_MAX_TOPK) {
        const uint32_t key = key_at(j);
        if ((key & mask) == prefix) atomicAdd(&hist[(key >> sh) & 255u], 1);
      }
      __syncthreads();
      const int h = hist[tid];
      int tot;
      const int ex = fn_block_excl_scan(h, wsum, &tot);
      if (ex < kk && kk <= ex + h) {
        s_bin = tid;
        s_kk = kk - ex;
      }
      __syncthreads();
      prefix |= (uint32_t)s_bin << sh;
      mask |= 255u << sh;
      kk = s_kk;
      __syncthreads();
    }
    const uint32_t T = prefix;
    const int n_less = k - kk;
    // ---- collect: all keys < T (any order), then the first kk keys == T in candidate order
    int eq_base = 0;
    for (int j0 = 0; j0 < n; j0 += FN_MAX_TOPK) {
      const int j = j0 + tid;
      uint32_t key = 0xffffffffu;
      bool lt = false, eq = false;
      if (j < n) {
        key = key_at(j);
        lt = key < T;
        eq = key == T;
      }
      if (lt) sel[atomicAdd(&s_cnt, 1)] = ((unsigned long long)key << 32) | (uint32_t)j;
      int tot;
      const int ex = fn_block_excl_scan(eq ? 1 : 0, wsum, &tot);
      if (eq && eq_base + ex < kk) sel[n_less + eq_base + ex] = ((unsigned long long)key << 32) | (uint32_t)j;
      eq_base += tot;
    }
  }
  __syncthreads();
  // ---- bitonic sort of the 256 (key, candidate) pairs, ascending = descending score, stable
  for (int size = 2; size <= FN_MAX_TOPK; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int partner = tid ^ stride;
      if (partner > tid) {
        const unsigned long long a = sel[tid], d = sel[partner];
        const bool up = (tid & size) == 0;
        if ((a > d) == up) {
          sel[tid] = d;
          sel[partner] = a;
        }
      }
      __syncthreads();
    }
  }
  // ---- FastNMS on the sorted top k (as fn_class_kernel)
  const int r = tid;
  int src = 0;
  float sc = 0.f;
  if (r < k) {
    src = (int)(uint32_t)(sel[r] & 0xffffffffull);
    const float4 p = *reinterpret_cast<const float4*>(boxes + (size_t)(off + src) * 4);
    bx[r] = p;
    ar[r] = rp_area_rel(p.x, p.y, p.z, p.w);
    sc = sc_c[src];
  }
  __syncthreads();
  if (r >= top_k) return;
  bool keep = false;
  if (r < k) {
    keep = fn_keep_rank(bx, ar, r, iou_thr);
    if (second_threshold) keep = keep && (sc > conf_thresh);
  }
  const size_t o = ((size_t)b * C + c) * top_k + r;
  keys2[o] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)fn_desc_bits(keep ? sc : -INFINITY);
  vals2[o] = c * top_k + r;
  cand_score[o] = keep ? sc : -INFINITY;
  cand_src[o] = src;
}

// grid B; emits the first max_det kept candidates of the image.
__global__ void fn_emit_kernel(const float* __restrict__ boxes, const float* __restrict__ coeff,
                               const int* __restrict__ offsets, const int* __restrict__ sorted_vals2,
                               const float* __restrict__ cand_score, const int* __restrict__ cand_src,
                               int C, int D, int top_k, int max_det, float* __restrict__ out_boxes,
                               float* __restrict__ out_coeff, int* __restrict__ out_classes,
                               float* __restrict__ out_scores, int* __restrict__ out_index,
                               int* __restrict__ n_out) {
  const int b = blockIdx.x;
  const int off = offsets[b];
  const size_t per = (size_t)C * top_k;
  const int lim = (int)min((size_t)max_det, per);
  __shared__ int s_count;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  for (int r = threadIdx.x; r < lim; r += blockDim.x) {
    const int q = sorted_vals2[b * per + r];
    const float sc = cand_score[b * per + q];
    if (sc == -INFINITY) continue;  // sentinels sort last: everything before them is kept
    atomicAdd(&s_count, 1);
    const int src = cand_src[b * per + q];
    const size_t o = (size_t)b * max_det + r;
    const float4 p = *reinterpret_cast<const float4*>(boxes + (size_t)(off + src) * 4);
    *reinterpret_cast<float4*>(out_boxes + o * 4) = p;
    for (int d = 0; d < D; ++d) out_coeff[o * D + d] = coeff[(size_t)(off + src) * D + d];
    out_classes[o] = q / top_k;
    out_scores[o] = sc;
    if (out_index) out_index[o] = src;
  }
  __syncthreads();
  if (threadIdx.x == 0) n_out[b] = s_count;
}

// cross-class: grid B, FN_MAX_TOPK threads; kept ranks are compacted in rank order.
__global__ void __launch_bounds__(FN_MAX_TOPK)
fn_cc_kernel(const float* __restrict__ boxes, const float* __restrict__ coeff,
             const int* __restrict__ offsets, const int* __restrict__ sorted_vals,
             const float* __restrict__ best, const int* __restrict__ cls, int D, int top_k,
             int max_det, float iou_thr, float* __restrict__ out_boxes,
             float* __restrict__ out_coeff, int* __restrict__ out_classes,
             float* __restrict__ out_scores, int* __restrict__ out_index,
             int* __restrict__ n_out) {
  __shared__ float4 bx[FN_MAX_TOPK];
  __shared__ float ar[FN_MAX_TOPK];
  __shared__ int s_warp[FN_MAX_TOPK / 32];
  const int b = blockIdx.x;
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
  const int k = min(n, top_k);
  const int r = threadIdx.x;
  int src = 0;
  if (r < k) {
    src = sorted_vals[off + r];
    const float4 p = *reinterpret_cast<const float4*>(boxes + (size_t)(off + src) * 4);
    bx[r] = p;
    ar[r] = rp_area_rel(p.x, p.y, p.z, p.w);
  }
  __syncthreads();
  const bool keep = (r < k) && fn_keep_rank(bx, ar, r, iou_thr);
  const unsigned bal = __ballot_sync(0xffffffffu, keep);
  const int lane = r & 31, wid = r >> 5;
  if (lane == 0) s_warp[wid] = __popc(bal);
  __syncthreads();
  int pos = __popc(bal & ((1u << lane) - 1u));
  int total = 0;
  for (int w = 0; w < FN_MAX_TOPK / 32; ++w) {
    if (w < wid) pos += s_warp[w];
    total += s_warp[w];
  }
  if (keep && pos < max_det) {
    const size_t o = (size_t)b * max_det + pos;
    *reinterpret_cast<float4*>(out_boxes + o * 4) = bx[r];
    for (int d = 0; d < D; ++d) out_coeff[o * D + d] = coeff[(size_t)(off + src) * D + d];
    out_classes[o] = cls[off + src];
    out_scores[o] = best[off + src];
    if (out_index) out_index[o] = src;
  }
  if (r == 0) n_out[b] = min(total, max_det);
}

struct FnWorkspace {
  uint64_t *keys_in, *keys_out, *keys2_in, *keys2_out;
  int *vals_in, *vals_out, *vals2_in, *vals2_out;
  float* cand_score;
  int* cand_src;
  float* best;
  int* cls;
  void* cub_temp;
  size_t cub_bytes;
  size_t total;
};

size_t fn_cub_bytes(size_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint64_t*)nullptr,
                                  (uint64_t*)nullptr, (const int*)nullptr, (int*)nullptr, (int)n, 0,
                                  64, (cudaStream_t)0);
  return bytes;
}

FnWorkspace fn_carve(void* base, int n_total, int B, int C, int top_k) {
  FnWorkspace w;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t at = o;
    o = rp_align_up(o + bytes, 256);
    return base ? (void*)((char*)base + at) : (void*)nullptr;
  };
  // the batch-wide sort arrays hold C * n_total keys only on the legacy sort path; the default
  // (fused select) path sorts nothing of that size, the cross-class path sorts n_total keys
  const size_t nt = (size_t)(n_total > 0 ? n_total : 1);
  const size_t n1 = rp_env_int("RP_FASTNMS_SELECT", 1) ? nt : (size_t)C * nt;
  const size_t n2 = (size_t)B * C * top_k;
  w.keys_in = (uint64_t*)take(n1 * 8);
  w.keys_out = (uint64_t*)take(n1 * 8);
  w.vals_in = (int*)take(n1 * 4);
  w.vals_out = (int*)take(n1 * 4);
  w.keys2_in = (uint64_t*)take(n2 * 8);
  w.keys2_out = (uint64_t*)take(n2 * 8);
  w.vals2_in = (int*)take(n2 * 4);
  w.vals2_out = (int*)take(n2 * 4);
  w.cand_score = (float*)take(n2 * 4);
  w.cand_src = (int*)take(n2 * 4);
  w.best = (float*)take((size_t)(n_total > 0 ? n_total : 1) * 4);
  w.cls = (int*)take((size_t)(n_total > 0 ? n_total : 1) * 4);
  const size_t c1 = fn_cub_bytes(n1), c2 = fn_cub_bytes(n2);
  w.cub_bytes = c1 > c2 ? c1 : c2;
  w.cub_temp = take(w.cub_bytes);
  w.total = o;
  return w;
}

int fn_bits(size_t v) {
  int b = 1;
  while (((size_t)1 << b) < v) ++b;
  return b;
}


// ---- Detect pre-stage (detection.py:64-98) for a whole batch ---------------------------------
// decode + class-max over the foreground classes + "> conf_thresh" + ordered compaction of the
// kept priors, with the [C, n] score layout FastNMS reads.  A chunk is YP_CHUNK consecutive
// priors of one image.
constexpr int YP_CHUNK = 256;

// pass A: flags[b*P + p] = max_c conf[b][p][1 + c] > thr; chunk_count[b*nchunks + chunk].
// The chunk's conf rows are one contiguous run of YP_CHUNK*NC floats: it is streamed into shared
// memory with 16-byte loads (all of them in flight at once), then thread t takes the maximum of row
// t (row stride NC: conflict-free when NC is odd, 2-way at worst otherwise).
__global__ void __launch_bounds__(YP_CHUNK)
yp_flag_kernel(const float* __restrict__ conf, int P, int NC, float thr,
               uint8_t* __restrict__ flags, int* __restrict__ chunk_count) {
  extern __shared__ __align__(16) float yp_rows[];
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int p0 = chunk * YP_CHUNK;
  const int rows = min(YP_CHUNK, P - p0);
  const float* src = conf + ((size_t)b * P + p0) * NC;
  const int nfl = rows * NC;
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const int nv = nfl >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(yp_rows);
#pragma unroll 8
    for (int i = threadIdx.x; i < nv; i += YP_CHUNK) d4[i] = __ldcs(s4 + i);
    for (int i = (nv << 2) + threadIdx.x; i < nfl; i += YP_CHUNK) yp_rows[i] = src[i];
  } else {
    for (int i = threadIdx.x; i < nfl; i += YP_CHUNK) yp_rows[i] = src[i];
  }
  __syncthreads();
  int f = 0;
  if (threadIdx.x < rows) {
    const float* row = yp_rows + threadIdx.x * NC;
    float m = -INFINITY;
    for (int c = 1; c < NC; ++c) m = fmaxf(m, row[c]);
    f = m > thr;
    flags[(size_t)b * P + p0 + threadIdx.x] = (uint8_t)f;
  }
  const int cnt = __syncthreads_count(f);
  if (threadIdx.x == 0) chunk_count[b * gridDim.x + chunk] = cnt;
}

// pass B: exclusive scan over the B*nchunks counts (image-major); offsets[b] = image starts,
// everything clamped to cand_cap; n_cand = the unclamped total.  Thread t owns a contiguous run.
__global__ void __launch_bounds__(FN_MAX_TOPK)
yp_scan_kernel(const int* __restrict__ chunk_count, int B, int nchunks, int cand_cap,
               int* __restrict__ chunk_off, int* __restrict__ offsets, int* __restrict__ n_cand) {
  __shared__ int wsum[FN_MAX_TOPK / 32];
  const int S = B * nchunks;
  const int run = (S + FN_MAX_TOPK - 1) / FN_MAX_TOPK;
  const int s0 = threadIdx.x * run, s1 = min(S, s0 + run);
  int sum = 0;
  for (int s = s0; s < s1; ++s) sum += chunk_count[s];
  int tot;
  int acc = fn_block_excl_scan(sum, wsum, &tot);
  for (int s = s0; s < s1; ++s) {
    const int o = min(acc, cand_cap);
    chunk_off[s] = o;
    if (s % nchunks == 0) offsets[s / nchunks] = o;
    acc += chunk_count[s];
  }
  if (threadIdx.x == 0) {
    offsets[B] = min(tot, cand_cap);
    *n_cand = tot;
  }
}

// pass C: thread per prior -> its slot; decoded box, prior index and image of the slot
__global__ void __launch_bounds__(YP_CHUNK)
yp_index_kernel(const float* __restrict__ loc, const float* __restrict__ priors,
                const uint8_t* __restrict__ flags, const int* __restrict__ chunk_off, int P,
                int cand_cap, float* __restrict__ out_boxes, int* __restrict__ slot_prior,
                int* __restrict__ slot_img) {
  __shared__ int wsum[YP_CHUNK / 32];
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int p = chunk * YP_CHUNK + threadIdx.x;
  const int f = (p < P) ? (int)flags[(size_t)b * P + p] : 0;
  int tot;
  const int ex = fn_block_excl_scan(f, wsum, &tot);
  const int dst = chunk_off[b * gridDim.x + chunk] + ex;
  if (f && dst < cand_cap) {  // a clamped run drops the tail (the caller repeats with more room)
    const float4 l = *reinterpret_cast<const float4*>(loc + ((size_t)b * P + p) * 4);
    const float4 pr = *reinterpret_cast<const float4*>(priors + (size_t)p * 4);
    // decode(), box_utils.py:272-316 -- same operation order as yolact_decode_kernel
    const float cx = pr.x + l.x * 0.1f * pr.z;
    const float cy = pr.y + l.y * 0.1f * pr.w;
    const float w = pr.z * rp_expf(l.z * 0.2f);
    const float h = pr.w * rp_expf(l.w * 0.2f);
    const float x1 = cx - w / 2.0f;
    const float y1 = cy - h / 2.0f;
    *reinterpret_cast<float4*>(out_boxes + (size_t)dst * 4) = make_float4(x1, y1, w + x1, h + y1);
    slot_prior[dst] = p;
    slot_img[dst] = b;
  }
}

// pass D: 32 consecutive slots per CTA.  Each warp copies whole rows (coefficients straight
// through, conf rows into shared memory); the scores then leave transposed with the slots along
// the lanes, i.e. as full 128-byte lines of out_scores [NC-1, cand_cap].
constexpr int YP_GROUP = 32;
__global__ void __launch_bounds__(YP_CHUNK)
yp_gather_kernel(const float* __restrict__ conf, const float* __restrict__ mask,
                 const int* __restrict__ offsets, const int* __restrict__ slot_prior,
                 const int* __restrict__ slot_img, int B, int P, int NC, int D, int cand_cap,
                 float* __restrict__ out_coeff, float* __restrict__ out_scores) {
  extern __shared__ __align__(16) float yp_kept[];  // [YP_GROUP][NC]
  const int total = offsets[B];
  const int d0 = blockIdx.x * YP_GROUP;
  if (d0 >= total) return;
  const int cnt = min(YP_GROUP, total - d0);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int e = wid; e < cnt; e += YP_CHUNK / 32) {
    const size_t row = (size_t)slot_img[d0 + e] * P + slot_prior[d0 + e];
    const float* mrow = mask + row * D;
    for (int d = lane; d < D; d += 32) out_coeff[(size_t)(d0 + e) * D + d] = mrow[d];
    const float* crow = conf + row * NC;
    for (int c = lane; c < NC; c += 32) yp_kept[e * NC + c] = crow[c];
  }
  __syncthreads();
  if (lane < cnt)
    for (int c = 1 + wid; c < NC; c += YP_CHUNK / 32)
      out_scores[(size_t)(c - 1) * cand_cap + d0 + lane] = yp_kept[lane * NC + c];
}

struct YpWorkspace {
  uint8_t* flags;
  int *chunk_count, *chunk_off, *slot_img;
  size_t total;
};

YpWorkspace yp_carve(void* base, int B, int P, int cand_cap) {
  YpWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (void*)((char*)base + off) : nullptr;
    off += rp_align_up(bytes, 256);
    return p;
  };
  const int nchunks = (P + YP_CHUNK - 1) / YP_CHUNK;
  w.flags = (uint8_t*)take((size_t)B * P);
  w.chunk_count = (int*)take(4 * (size_t)B * nchunks);
  w.chunk_off = (int*)take(4 * (size_t)B * nchunks);
  w.slot_img = (int*)take(4 * (size_t)cand_cap);
  w.total = off;
  return w;
}

}  // namespace

extern "C" {

size_t rp_yolact_prefilter_workspace_bytes(int B, int P, int cand_cap) {
  if (B < 1 || P < 1 || cand_cap < 1) return 0;
  return yp_carve(nullptr, B, P, cand_cap).total;
}

int rp_yolact_prefilter(const float* loc, const float* conf, const float* mask, const float* priors,
                        int B, int P, int NC, int D, float conf_thresh, int cand_cap,
                        float* out_boxes, float* out_coeff, float* out_scores, int* out_prior,
                        int* out_offsets, int* n_cand, void* workspace, size_t workspace_bytes,
                        void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (B < 1 || P < 1 || NC < 2 || D < 1 || cand_cap < 1 || B > 65535) return RP_E_ARG;
  if (!loc || !conf || !mask || !priors || !out_boxes || !out_coeff || !out_scores || !out_prior ||
      !out_offsets || !n_cand || !workspace)
    return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(loc) | reinterpret_cast<uintptr_t>(priors) |
        reinterpret_cast<uintptr_t>(out_boxes)) & 15) != 0)
    return RP_E_ARG;
  YpWorkspace w = yp_carve(workspace, B, P, cand_cap);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  const int nchunks = (P + YP_CHUNK - 1) / YP_CHUNK;
  const dim3 grid((unsigned)nchunks, (unsigned)B);
  const size_t flag_smem = (size_t)YP_CHUNK * NC * sizeof(float);
  if (flag_smem > 200 * 1024) return RP_E_LIMIT;  // NC <= 200 classes
  RP_SET_MAX_SMEM(yp_flag_kernel, flag_smem);
  yp_flag_kernel<<<grid, YP_CHUNK, flag_smem, st>>>(conf, P, NC, conf_thresh, w.flags, w.chunk_count);
  RP_LAUNCH_CHECK();
  yp_scan_kernel<<<1, FN_MAX_TOPK, 0, st>>>(w.chunk_count, B, nchunks, cand_cap, w.chunk_off,
                                            out_offsets, n_cand);
  RP_LAUNCH_CHECK();
  yp_index_kernel<<<grid, YP_CHUNK, 0, st>>>(loc, priors, w.flags, w.chunk_off, P, cand_cap, out_boxes,
                                             out_prior, w.slot_img);
  RP_LAUNCH_CHECK();
  yp_gather_kernel<<<(cand_cap + YP_GROUP - 1) / YP_GROUP, YP_CHUNK,
                     (size_t)YP_GROUP * NC * sizeof(float), st>>>(
      conf, mask, out_offsets, out_prior, w.slot_img, B, P, NC, D, cand_cap, out_coeff, out_scores);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_yolact_decode(const float* loc, const float* priors, int n, float* out, void* stream) {
  if (n == 0) return RP_OK;
  if (!loc || !priors || !out || n < 0) return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(loc) | reinterpret_cast<uintptr_t>(priors) |
        reinterpret_cast<uintptr_t>(out)) & 15) != 0)
    return RP_E_ARG;
  yolact_decode_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(loc, priors, n, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

size_t rp_fast_nms_workspace_bytes(int n_total, int n_max, int B, int C, int top_k) {
  (void)n_max;
  if (n_total < 0 || B < 1 || C < 1 || top_k < 1) return 0;
  return fn_carve(nullptr, n_total, B, C, top_k).total;
}

int rp_fast_nms_batched(const float* boxes, const float* coeff, const float* scores,
                        const int* offsets, int B, int n_total, int n_max, int C, int D,
                        float iou_thr, int top_k, int max_det, int second_threshold,
                        float conf_thresh, int cross_class, float* out_boxes, float* out_coeff,
                        int* out_classes, float* out_scores, int* out_index, int* n_out,
                        void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  (void)n_max;
  if (B < 1 || C < 1 || D < 1 || top_k < 1 || max_det < 1 || n_total < 0 || !offsets || !n_out)
    return RP_E_ARG;
  if (top_k > FN_MAX_TOPK) return RP_E_LIMIT;
  if (n_total == 0) {
    RP_CUDA_TRY(cudaMemsetAsync(n_out, 0, sizeof(int) * (size_t)B, st));
    return RP_OK;
  }
  if (!boxes || !coeff || !scores || !out_boxes || !out_coeff || !out_classes || !out_scores ||
      !workspace)
    return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(boxes) | reinterpret_cast<uintptr_t>(out_boxes)) & 15) != 0)
    return RP_E_ARG;
  if ((size_t)C * n_total > 0x7fffffffull || (size_t)B * C * top_k > 0x7fffffffull)
    return RP_E_LIMIT;
  FnWorkspace w = fn_carve(workspace, n_total, B, C, top_k);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  const int T = 256;
  if (cross_class) {
    fn_cc_keys_kernel<<<(n_total + T - 1) / T, T, 0, st>>>(scores, offsets, B, C, n_total,
                                                          w.keys_in, w.vals_in, w.best, w.cls);
    RP_LAUNCH_CHECK();
    RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out,
                                                w.vals_in, w.vals_out, n_total, 0,
                                                32 + fn_bits((size_t)B + 1), st));
    fn_cc_kernel<<<B, FN_MAX_TOPK, 0, st>>>(boxes, coeff, offsets, w.vals_out, w.best, w.cls, D,
                                            top_k, max_det, iou_thr, out_boxes, out_coeff,
                                            out_classes, out_scores, out_index, n_out);
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  if (B > 65535) return RP_E_LIMIT;
  if (rp_env_int("RP_FASTNMS_SELECT", 1)) {
    // fused per-(class, image) top-k selection + FastNMS: no batch-wide sort of the C * n scores
    fn_class_select_kernel<<<dim3((unsigned)C, (unsigned)B), FN_MAX_TOPK, 0, st>>>(
        boxes, scores, offsets, C, n_total, top_k, iou_thr, second_threshold, conf_thresh, w.keys2_in,
        w.vals2_in, w.cand_score, w.cand_src);
    RP_LAUNCH_CHECK();
  } else {
    const size_t n1 = (size_t)C * n_total;
    fn_keys_kernel<<<(unsigned)((n1 + T - 1) / T), T, 0, st>>>(scores, offsets, B, C, n_total,
                                                              w.keys_in, w.vals_in);
    RP_LAUNCH_CHECK();
    RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out,
                                                w.vals_in, w.vals_out, (int)n1, 0,
                                                32 + fn_bits((size_t)B * C + 1), st));
    fn_class_kernel<<<dim3((unsigned)C, (unsigned)B), FN_MAX_TOPK, 0, st>>>(
        boxes, scores, offsets, w.vals_out, C, n_total, top_k, iou_thr, second_threshold,
        conf_thresh, w.keys2_in, w.vals2_in, w.cand_score, w.cand_src);
    RP_LAUNCH_CHECK();
  }
  const size_t n2 = (size_t)B * C * top_k;
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys2_in, w.keys2_out,
                                              w.vals2_in, w.vals2_out, (int)n2, 0,
                                              32 + fn_bits((size_t)B), st));
  fn_emit_kernel<<<B, 128, 0, st>>>(boxes, coeff, offsets, w.vals2_out, w.cand_score, w.cand_src,
                                    C, D, top_k, max_det, out_boxes, out_coeff, out_classes,
                                    out_scores, out_index, n_out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
