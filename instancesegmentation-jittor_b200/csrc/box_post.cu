// box_post.cu -- the box-head post-processor around the per-class NMS,
// sm_100a.
//
// Reference being replaced (paths relative to the reference checkout):
//   det/modeling/roi_heads/box_head/inference.py:45-90   PostProcessor.execute: softmax,
//        class-wise BoxCoder.decode, per-image split, clip_to_image
//   det/modeling/roi_heads/box_head/inference.py:112-195 filter_results: per class j >= 1
//        score > thresh -> nonzero -> boxlist_nms -> labels; cat; kthvalue cut at
//        detections_per_img (ties with the k-th score are all kept, :180-187)
// The reference launches ~80 nonzero + 80 NMS + 80 gathers per image and synchronises after the
// nonzeros.  Here, for the whole batch:
//   1. softmax      one warp per proposal row; also counts the candidates of every
//                   (image, class) problem
//   2. offsets      exclusive scan of the counts (image-major, class ascending)
//   3. candidates   one warp per problem: ordered compaction of its passing rows; only those are
//                   decoded and clipped
//   4. NMS          ONE rp_nms_batched call over the N*(C-1) problems
//   5. finish       one CTA per image: the detections_per_img-th largest kept score by radix
//                   select, then ordered emission (class-major, descending score in a class)
// No host synchronisation anywhere; the caller reads out_count / n_cand once.
#include "common.cuh"

namespace {

constexpr int BP_THREADS = 256;

__device__ __forceinline__ uint32_t bp_desc_bits(float s) {
  if (s == 0.0f) s = 0.0f;
  uint32_t u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ~u;
}

// box_coder.py:53-94 for one box (same operation order as rpn.cu / the oracle; --fmad=false)
__device__ __forceinline__ float4 bp_decode_one(float4 a, float4 d, float wx, float wy, float ww,
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

__device__ __forceinline__ int bp_find_image(const int* __restrict__ row_offsets, int N, int i) {
  int lo = 0, hi = N - 1;  // largest b with row_offsets[b] <= i
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (row_offsets[mid] <= i)
      lo = mid;
    else
      hi = mid - 1;
  }
  return lo;
}

template <int T>
__device__ __forceinline__ int bp_block_excl_scan(int v, int* warp_sums, int* total) {
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

// softmax(class_logits, -1): max, exp(x - max), sum in class order, divide (inference.py:58).
// One warp per row; the row's exponentials sit in shared memory so that lane 0 can add them in
// class order (the order the oracle uses).
__global__ void __launch_bounds__(BP_THREADS)
bp_softmax_kernel(const float* __restrict__ logits, const int* __restrict__ row_offsets, int R, int C,
                  int N, float thr, float* __restrict__ prob, int* __restrict__ counts) {
  extern __shared__ float bp_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int i = blockIdx.x * (BP_THREADS / 32) + wid;
  if (i >= R) return;
  float* e = bp_smem + (size_t)wid * C;
  const float* x = logits + (size_t)i * C;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, x[c]);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, d));
  for (int c = lane; c < C; c += 32) e[c] = rp_expf(x[c] - m);
  __syncwarp();
  float sum = 0.0f;
  if (lane == 0)
    for (int c = 0; c < C; ++c) sum += e[c];
  sum = __shfl_sync(0xffffffffu, sum, 0);
  const int b = bp_find_image(row_offsets, N, i);
  for (int c = lane; c < C; c += 32) {
    const float p = e[c] / sum;
    prob[(size_t)i * C + c] = p;
    if (c >= 1 && p > thr) atomicAdd(&counts[b * (C - 1) + (c - 1)], 1);
  }
}

// seg_off[0..S] = min(exclusive scan of counts, cand_cap); n_cand = the unclamped total
__global__ void __launch_bounds__(BP_THREADS)
bp_offsets_kernel(const int* __restrict__ counts, int S, int cand_cap, int* __restrict__ seg_off,
                  int* __restrict__ n_cand) {
  __shared__ int wsum[BP_THREADS / 32];
  int base = 0;
  for (int s0 = 0; s0 < S; s0 += BP_THREADS) {
    const int s = s0 + threadIdx.x;
    const int v = s < S ? counts[s] : 0;
    int tot;
    const int ex = bp_block_excl_scan<BP_THREADS>(v, wsum, &tot);
    if (s < S) seg_off[s] = min(base + ex, cand_cap);
    base += tot;
  }
  if (threadIdx.x == 0) {
    seg_off[S] = min(base, cand_cap);
    *n_cand = base;
  }
}

// One warp per problem s = b*(C-1) + (j-1): rows of image b with prob[., j] > thr, ascending row
// (the order of nonzero(), inference.py:128), decoded with the class's codes and clipped.
__global__ void __launch_bounds__(BP_THREADS)
bp_candidates_kernel(const float* __restrict__ prob, const float* __restrict__ reg,
                     const float* __restrict__ proposals, const int* __restrict__ row_offsets,
                     const float* __restrict__ im_wh, const int* __restrict__ seg_off, int C, int N,
                     int reg_cols, int agnostic, float thr, float wx, float wy, float ww, float wh,
                     float clip, float* __restrict__ cand_boxes, float* __restrict__ cand_scores,
                     int* __restrict__ cand_row) {
  const int lane = threadIdx.x & 31;
  const int s = blockIdx.x * (BP_THREADS / 32) + (threadIdx.x >> 5);
  if (s >= N * (C - 1)) return;
  const int b = s / (C - 1), j = s - b * (C - 1) + 1;
  const int r0 = row_offsets[b], r1 = row_offsets[b + 1];
  const int off = seg_off[s], limit = seg_off[s + 1] - off;
  if (limit <= 0) return;
  const float xmax = im_wh[2 * b] - 1.0f, ymax = im_wh[2 * b + 1] - 1.0f;
  const int code_col = agnostic ? reg_cols - 4 : 4 * j;
  int base = 0;
  // four 32-row steps per trip: the strided probability loads of a trip are all in flight before the
  // first ballot (a problem is one warp walking ~1000 rows: latency, not bandwidth)
  for (int i0 = r0; i0 < r1 && base < limit; i0 += 128) {
    float pv[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int i = i0 + 32 * t + lane;
      pv[t] = (i < r1) ? prob[(size_t)i * C + j] : -1.0f;
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int i = i0 + 32 * t + lane;
      const float p = pv[t];
      const bool ok = (i < r1) && (p > thr);
      const unsigned m = __ballot_sync(0xffffffffu, ok);
      const int pos = base + __popc(m & ((1u << lane) - 1u));
      if (ok && pos < limit) {
        const float4 d = *reinterpret_cast<const float4*>(reg + (size_t)i * reg_cols + code_col);
        const float4 a = *reinterpret_cast<const float4*>(proposals + (size_t)i * 4);
        float4 o = bp_decode_one(a, d, wx, wy, ww, wh, clip);
        o.x = fminf(fmaxf(o.x, 0.f), xmax);  // clip_to_image(remove_empty=False), :80
        o.y = fminf(fmaxf(o.y, 0.f), ymax);
        o.z = fminf(fmaxf(o.z, 0.f), xmax);
        o.w = fminf(fmaxf(o.w, 0.f), ymax);
        *reinterpret_cast<float4*>(cand_boxes + (size_t)(off + pos) * 4) = o;
        cand_scores[off + pos] = p;
        cand_row[off + pos] = i - r0;
      }
      base += __popc(m);
    }
  }
}

// One CTA per image: the kept boxes of its C-1 problems in class order; when there are more than
// `dets` of them, only those whose score >= the dets-th largest score (kthvalue, :180-187).
__global__ void __launch_bounds__(BP_THREADS)
bp_finish_kernel(const float* __restrict__ cand_boxes, const float* __restrict__ cand_scores,
                 const int* __restrict__ cand_row, const int* __restrict__ seg_off,
                 const int* __restrict__ keep, const int* __restrict__ n_keep, int C, int dets,
                 int out_cap, float* __restrict__ out_boxes, float* __restrict__ out_scores,
                 int* __restrict__ out_labels, int* __restrict__ out_rows,
                 int* __restrict__ out_count) {
  extern __shared__ int bp_pre[];  // [C] prefix of the kept counts over this image's classes
  __shared__ int wsum[BP_THREADS / 32];
  __shared__ int hist[256];
  __shared__ int s_bin, s_kk;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int S0 = b * (C - 1);
  int n = 0;
  for (int c0 = 0; c0 < C - 1; c0 += BP_THREADS) {
    const int c = c0 + tid;
    const int v = c < C - 1 ? n_keep[S0 + c] : 0;
    int tot;
    const int ex = bp_block_excl_scan<BP_THREADS>(v, wsum, &tot);
    if (c < C - 1) bp_pre[c] = n + ex;
    n += tot;
  }
  if (tid == 0) bp_pre[C - 1] = n;
  __syncthreads();
  // kept entry q of the image -> (class slot c, row of the candidate arrays)
  auto locate = [&](int q, int& c) -> int {
    int lo = 0, hi = C - 2;  // largest c with pre[c] <= q
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (bp_pre[mid] <= q)
        lo = mid;
      else
        hi = mid - 1;
    }
    c = lo;
    const int so = seg_off[S0 + lo];
    return so + keep[so + (q - bp_pre[lo])];
  };
  uint32_t T = 0xffffffffu;  // keep entries whose descending key <= T
  if (dets > 0 && n > dets) {
    uint32_t prefix = 0, mask = 0;
    int kk = dets;
    for (int pass = 3; pass >= 0; --pass) {
      hist[tid] = 0;
      __syncthreads();
      const int sh = 8 * pass;
      for (int q = tid; q < n; q += BP_THREADS) {
        int c;
        const uint32_t key = bp_desc_bits(cand_scores[locate(q, c)]);
        if ((key & mask) == prefix) atomicAdd(&hist[(key >> sh) & 255u], 1);
      }
      __syncthreads();
      const int h = hist[tid];
      int tot;
      const int ex = bp_block_excl_scan<BP_THREADS>(h, wsum, &tot);
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
    T = prefix;
  }
  int base = 0;
  for (int q0 = 0; q0 < n; q0 += BP_THREADS) {
    const int q = q0 + tid;
    int ok = 0, c = 0, src = 0;
    if (q < n) {
      src = locate(q, c);
      ok = bp_desc_bits(cand_scores[src]) <= T;
    }
    int tot;
    const int ex = bp_block_excl_scan<BP_THREADS>(ok, wsum, &tot);
    const int pos = base + ex;
    if (ok && pos < out_cap) {
      const size_t o = (size_t)b * out_cap + pos;
      *reinterpret_cast<float4*>(out_boxes + o * 4) =
          *reinterpret_cast<const float4*>(cand_boxes + (size_t)src * 4);
      out_scores[o] = cand_scores[src];
      out_labels[o] = c + 1;
      if (out_rows) out_rows[o] = cand_row[src];
    }
    base += tot;
  }
  if (tid == 0) out_count[b] = base;
}

struct BpWorkspace {
  int *counts, *seg_off, *cand_row, *keep, *n_keep;
  float *cand_boxes, *cand_scores, *prob;
  void* nms;
  size_t nms_bytes, total;
};

// nms_thresh <= 0: boxlist_nms returns its input unchanged (boxlist_ops.py:24-25), i.e. every
// candidate of the segment is kept in candidate (= nonzero()) order.
__global__ void bp_identity_keep_kernel(const int* __restrict__ seg_off, int S, int* __restrict__ keep,
                                        int* __restrict__ n_keep) {
  const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (s >= S) return;
  const int o = seg_off[s], n = seg_off[s + 1] - o;
  for (int i = lane; i < n; i += 32) keep[o + i] = i;
  if (lane == 0) n_keep[s] = n;
}

BpWorkspace bp_carve(void* base, int R, int C, int N, int r_max, int cand_cap, bool need_prob) {
  BpWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (void*)((char*)base + off) : nullptr;
    off += rp_align_up(bytes, 256);
    return p;
  };
  const int S = N * (C - 1);
  w.counts = (int*)take(4 * (size_t)S);
  w.seg_off = (int*)take(4 * (size_t)(S + 1));
  w.n_keep = (int*)take(4 * (size_t)S);
  w.cand_boxes = (float*)take(16 * (size_t)cand_cap);
  w.cand_scores = (float*)take(4 * (size_t)cand_cap);
  w.cand_row = (int*)take(4 * (size_t)cand_cap);
  w.keep = (int*)take(4 * (size_t)cand_cap);
  w.prob = need_prob ? (float*)take(4 * (size_t)R * C) : nullptr;
  w.nms_bytes = rp_nms_workspace_bytes(cand_cap, r_max < cand_cap ? r_max : cand_cap, S);
  w.nms = take(w.nms_bytes);
  w.total = off;
  return w;
}

}  // namespace

extern "C" {

size_t rp_box_post_workspace_bytes(int R, int C, int N, int r_max, int cand_cap) {
  if (R < 1 || C < 2 || N < 1 || r_max < 1 || cand_cap < 1) return 0;
  return bp_carve(nullptr, R, C, N, r_max, cand_cap, true).total;
}

int rp_box_post(const float* class_logits, const float* box_regression, const float* proposals,
                const int* row_offsets, const float* im_wh, int R, int C, int N, int r_max,
                int reg_cols, int cls_agnostic, float score_thresh, double nms_thresh,
                int detections_per_img, float wx, float wy, float ww, float wh,
                float bbox_xform_clip, int cand_cap, float* prob_out, float* out_boxes,
                float* out_scores, int* out_labels, int* out_rows, int* out_count, int out_cap,
                int* n_cand, void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (R < 1 || C < 2 || N < 1 || r_max < 1 || cand_cap < 1 || out_cap < 1 || nms_thresh != nms_thresh)
    return RP_E_ARG;
  if (!class_logits || !box_regression || !proposals || !row_offsets || !im_wh || !out_boxes ||
      !out_scores || !out_labels || !out_count || !n_cand || !workspace)
    return RP_E_ARG;
  if (reg_cols < 4 || (reg_cols & 3) || (!cls_agnostic && reg_cols != 4 * C)) return RP_E_ARG;
  if (((reinterpret_cast<uintptr_t>(box_regression) | reinterpret_cast<uintptr_t>(proposals) |
        reinterpret_cast<uintptr_t>(out_boxes)) & 15) != 0)
    return RP_E_ARG;
  if (C > 8192 || (size_t)N * (C - 1) > 65535) return RP_E_LIMIT;
  BpWorkspace w = bp_carve(workspace, R, C, N, r_max, cand_cap, prob_out == nullptr);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  float* prob = prob_out ? prob_out : w.prob;
  const int S = N * (C - 1);
  RP_CUDA_TRY(cudaMemsetAsync(w.counts, 0, sizeof(int) * (size_t)S, st));
  const int wpb = BP_THREADS / 32;
  bp_softmax_kernel<<<(R + wpb - 1) / wpb, BP_THREADS, (size_t)wpb * C * sizeof(float), st>>>(
      class_logits, row_offsets, R, C, N, score_thresh, prob, w.counts);
  RP_LAUNCH_CHECK();
  bp_offsets_kernel<<<1, BP_THREADS, 0, st>>>(w.counts, S, cand_cap, w.seg_off, n_cand);
  RP_LAUNCH_CHECK();
  bp_candidates_kernel<<<(S + wpb - 1) / wpb, BP_THREADS, 0, st>>>(
      prob, box_regression, proposals, row_offsets, im_wh, w.seg_off, C, N, reg_cols,
      cls_agnostic ? 1 : 0, score_thresh, wx, wy, ww, wh, bbox_xform_clip, w.cand_boxes,
      w.cand_scores, w.cand_row);
  RP_LAUNCH_CHECK();
  const int n_max = r_max < cand_cap ? r_max : cand_cap;
  if (nms_thresh > 0) {
    const int rc = rp_nms_batched(w.cand_boxes, w.cand_scores, nullptr, w.seg_off, S, cand_cap, n_max,
                                  nms_thresh, -1, 0, w.keep, w.n_keep, w.nms, w.nms_bytes, stream);
    if (rc) return rc;
  } else {
    bp_identity_keep_kernel<<<(S + wpb - 1) / wpb, BP_THREADS, 0, st>>>(w.seg_off, S, w.keep, w.n_keep);
    RP_LAUNCH_CHECK();
  }
  bp_finish_kernel<<<N, BP_THREADS, (size_t)C * sizeof(int), st>>>(
      w.cand_boxes, w.cand_scores, w.cand_row, w.seg_off, w.keep, w.n_keep, C, detections_per_img,
      out_cap, out_boxes, out_scores, out_labels, out_rows, out_count);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
