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


// ---- per-problem top-k without a sort of all anchors -------------------------------------------
// topk(pre_nms_top_n, sorted=True) of every (level, image) (rpn/inference.py:83-111) as a byte-wise
// radix SELECT spread over the whole GPU, then a small sort of the k survivors per problem:
//   rs_keys     coalesced walk of the NCHW objectness tensors: 32-bit key (descending sigmoid bits) written
//               at the element's flattened (h, w, a) position + histogram of the top byte (shared-memory
//               histogram per 1024-element chunk, flushed with a few global atomics)
//   rs_hist x3  every chunk first derives, from the previous pass's histogram, which bin holds the k-th key
//               of its problem (all chunks of a problem compute the same thing), then histograms the next byte
//               of the keys that match the prefix so far
//   rs_ties     the k-th key T is known: per chunk, how many keys equal T (ties are taken in ascending
//               flattened index = ascending position)
//   rs_collect  keys < T go to the problem's list through an atomic cursor; a chunk takes its ties only while
//               (ties in earlier chunks + rank inside the chunk) < needed
//   rs_sort     one CTA per problem: bitonic sort of (key << 32 | index) in shared memory, indices out
// Replaces the 64-bit cub::DeviceRadixSort over all L*N*A*H*W anchors of the batch (38 % of the stage's
// device time in round 1: 5 onesweep passes over 2.15 M pairs).
constexpr int RS_CAP = 4096;     // pre_nms_top_n handled here; larger values keep the sort path
constexpr int RS_CHUNK = 1024;   // elements per CTA (256 threads x 4)

struct RsPlan {
  int chunk_base[RP_MAX_LEVELS + 1];  // prefix of N * chunks_per_problem[l]
  int cpp[RP_MAX_LEVELS];             // chunks per problem of level l
  uint32_t* keys;                     // [n_elems] in flattened order
  int* hist;                          // [4][S][256]
  int* state;                         // [4][S][2]: prefix, remaining rank after the pass
  int* tie_cnt;                       // [n_chunks]
  int* cursor;                        // [S]
  unsigned long long* list;           // [S][RS_CAP]
};

struct RsWhere {
  int l, b, j, s, per, base;  // level, image, chunk inside the problem, problem id, elements, first element
};

__device__ __forceinline__ RsWhere rs_where(const RpnLevels& lv, const RsPlan& pl, int chunk) {
  RsWhere w;
  int l = 0;
  while (l + 1 < lv.L && chunk >= pl.chunk_base[l + 1]) ++l;
  const int r = chunk - pl.chunk_base[l];
  w.l = l;
  w.b = r / pl.cpp[l];
  w.j = r - w.b * pl.cpp[l];
  w.s = l * lv.N + w.b;
  w.per = lv.A[l] * lv.HW[l];
  w.base = lv.elem_base[l] + w.b * w.per;
  return w;
}

// The bin of `hist` (256 counts) that holds the rank-th smallest (1-based) value, and how many values
// lie in lower bins; evaluated by the whole CTA (256 threads), result in every thread.
__device__ __forceinline__ void rs_pick(const int* __restrict__ hist, int rank, int* s_scan, int* bin,
                                        int* before, int* count) {
  const int t = threadIdx.x;
  const int v = hist[t];
  // inclusive scan over 256 threads
  int inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc, d);
    if ((t & 31) >= d) inc += u;
  }
  if ((t & 31) == 31) s_scan[t >> 5] = inc;
  __syncthreads();
  int wbase = 0;
  for (int w = 0; w < (t >> 5); ++w) wbase += s_scan[w];
  inc += wbase;
  const int exc = inc - v;
  __syncthreads();
  if (exc < rank && inc >= rank) {
    s_scan[8] = t;
    s_scan[9] = exc;
    s_scan[10] = v;
  }
  __syncthreads();
  *bin = s_scan[8];
  *before = s_scan[9];
  *count = s_scan[10];
  __syncthreads();
}

// prefix / remaining rank of problem s after passes 3 .. pass+1 (recomputed by every chunk; chunk 0 stores it)
__device__ __forceinline__ void rs_state(const RpnLevels& lv, const RsPlan& pl, const RsWhere& w, int S, int pass,
                                         int* s_scan, uint32_t* prefix, int* rank, int* last_count) {
  uint32_t pre = 0;
  int kk = lv.k[w.l], cnt = 0;
  if (pass < 3) {
    if (pass < 2) {  // state after the pass before the previous one was stored by chunk 0 of that launch
      pre = (uint32_t)pl.state[((pass + 2) * S + w.s) * 2];
      kk = pl.state[((pass + 2) * S + w.s) * 2 + 1];
    }
    int bin, before;
    rs_pick(pl.hist + ((size_t)(pass + 1) * S + w.s) * 256, kk, s_scan, &bin, &before, &cnt);
    pre |= (uint32_t)bin << (8 * (pass + 1));
    kk -= before;
    if (w.j == 0 && threadIdx.x == 0) {
      pl.state[((pass + 1) * S + w.s) * 2] = (int)pre;
      pl.state[((pass + 1) * S + w.s) * 2 + 1] = kk;
    }
  }
  *prefix = pre;
  *rank = kk;
  *last_count = cnt;
}

__global__ void __launch_bounds__(256)
rs_keys_kernel(const __grid_constant__ RpnLevels lv, const __grid_constant__ RsPlan pl, int S) {
  __shared__ int hist[256];
  const RsWhere w = rs_where(lv, pl, blockIdx.x);
  hist[threadIdx.x] = 0;
  __syncthreads();
  const int A = lv.A[w.l], HW = lv.HW[w.l];
  const float* obj_b = lv.obj[w.l] + (size_t)w.b * w.per;
#pragma unroll
  for (int q = 0; q < RS_CHUNK / 256; ++q) {
    const int i = w.j * RS_CHUNK + q * 256 + threadIdx.x;  // memory order (a, hw)
    if (i < w.per) {
      const int a = i / HW, hw = i - a * HW;
      const uint32_t key = rpn_desc_bits(rp_sigmoidf(obj_b[i]));
      pl.keys[w.base + hw * A + a] = key;
      atomicAdd(&hist[key >> 24], 1);
    }
  }
  __syncthreads();
  const int v = hist[threadIdx.x];
  if (v) atomicAdd(&pl.hist[((size_t)3 * S + w.s) * 256 + threadIdx.x], v);
}

__global__ void __launch_bounds__(256)
rs_hist_kernel(const __grid_constant__ RpnLevels lv, const __grid_constant__ RsPlan pl, int S, int pass) {
  __shared__ int hist[256];
  __shared__ int s_scan[12];
  const RsWhere w = rs_where(lv, pl, blockIdx.x);
  uint32_t prefix;
  int rank, cnt;
  rs_state(lv, pl, w, S, pass, s_scan, &prefix, &rank, &cnt);
  hist[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t mask = 0xffffffffu << (8 * (pass + 1));
  const int sh = 8 * pass;
#pragma unroll
  for (int q = 0; q < RS_CHUNK / 256; ++q) {
    const int e = w.j * RS_CHUNK + q * 256 + threadIdx.x;
    if (e < w.per) {
      const uint32_t key = pl.keys[w.base + e];
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> sh) & 255u], 1);
    }
  }
  __syncthreads();
  const int v = hist[threadIdx.x];
  if (v) atomicAdd(&pl.hist[((size_t)pass * S + w.s) * 256 + threadIdx.x], v);
}

// pass == -1 in rs_state terms: all four bytes known after this derivation
__device__ __forceinline__ void rs_final(const RpnLevels& lv, const RsPlan& pl, const RsWhere& w, int S, int* s_scan,
                                         uint32_t* T, int* need, int* equal) {
  uint32_t pre = (uint32_t)pl.state[(1 * S + w.s) * 2];
  int kk = pl.state[(1 * S + w.s) * 2 + 1];
  int bin, before, cnt;
  rs_pick(pl.hist + ((size_t)0 * S + w.s) * 256, kk, s_scan, &bin, &before, &cnt);
  *T = pre | (uint32_t)bin;
  *need = kk - before;
  *equal = cnt;
}

__global__ void __launch_bounds__(256)
rs_ties_kernel(const __grid_constant__ RpnLevels lv, const __grid_constant__ RsPlan pl, int S) {
  __shared__ int s_scan[12];
  __shared__ int s_n;
  const RsWhere w = rs_where(lv, pl, blockIdx.x);
  uint32_t T;
  int need, equal;
  rs_final(lv, pl, w, S, s_scan, &T, &need, &equal);
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  int n = 0;
#pragma unroll
  for (int q = 0; q < RS_CHUNK / 256; ++q) {
    const int e = w.j * RS_CHUNK + q * 256 + threadIdx.x;
    if (e < w.per && pl.keys[w.base + e] == T) ++n;
  }
  if (n) atomicAdd(&s_n, n);
  __syncthreads();
  if (threadIdx.x == 0) pl.tie_cnt[blockIdx.x] = s_n;
}

__global__ void __launch_bounds__(256)
rs_collect_kernel(const __grid_constant__ RpnLevels lv, const __grid_constant__ RsPlan pl, int S) {
  __shared__ int s_scan[12];
  __shared__ int s_w[8];
  const RsWhere w = rs_where(lv, pl, blockIdx.x);
  uint32_t T;
  int need, equal;
  rs_final(lv, pl, w, S, s_scan, &T, &need, &equal);
  // ties in the earlier chunks of this problem
  int earlier = 0;
  {
    const int c0 = blockIdx.x - w.j;
    int part = 0;
    for (int c = threadIdx.x; c < w.j; c += 256) part += pl.tie_cnt[c0 + c];
#pragma unroll
    for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = part;
    __syncthreads();
    for (int i = 0; i < 8; ++i) earlier += s_w[i];
    __syncthreads();
  }
  unsigned long long* list = pl.list + (size_t)w.s * RS_CAP;
  // the chunk's elements in ascending position: thread t takes positions 4t .. 4t+3
  const int e0 = w.j * RS_CHUNK + threadIdx.x * 4;
  uint32_t key[4];
  int tie = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    key[q] = (e0 + q < w.per) ? pl.keys[w.base + e0 + q] : 0xffffffffu;
    if (e0 + q < w.per && key[q] == T) ++tie;
  }
  // exclusive scan of the tie counts over the CTA
  int inc = tie;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc, d);
    if ((threadIdx.x & 31) >= d) inc += u;
  }
  if ((threadIdx.x & 31) == 31) s_w[threadIdx.x >> 5] = inc;
  __syncthreads();
  int rank = earlier + inc - tie;
  for (int i = 0; i < (threadIdx.x >> 5); ++i) rank += s_w[i];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (e0 + q >= w.per) continue;
    bool take = key[q] < T;
    if (key[q] == T) {
      take = rank < need;
      ++rank;
    }
    if (take) {
      const int pos = atomicAdd(&pl.cursor[w.s], 1);
      if (pos < RS_CAP) list[pos] = ((unsigned long long)key[q] << 32) | (uint32_t)(e0 + q);
    }
  }
}

__global__ void __launch_bounds__(1024)
rs_sort_kernel(const __grid_constant__ RpnLevels lv, const __grid_constant__ RsPlan pl, int* __restrict__ vals_out) {
  __shared__ unsigned long long list[RS_CAP];
  const int s = blockIdx.x;
  const int l = s / lv.N, b = s - l * lv.N;
  const int k = lv.k[l], per = lv.A[l] * lv.HW[l];
  int np2 = 1;
  while (np2 < k) np2 <<= 1;
  const unsigned long long* src = pl.list + (size_t)s * RS_CAP;
  for (int i = threadIdx.x; i < np2; i += 1024) list[i] = i < k ? src[i] : ~0ull;
  __syncthreads();
  for (int size = 2; size <= np2; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < (np2 >> 1); t += 1024) {
        const int lo = ((t / stride) * stride * 2) + (t % stride), hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const unsigned long long x = list[lo], y = list[hi];
        if ((x > y) == up) {
          list[lo] = y;
          list[hi] = x;
        }
      }
      __syncthreads();
    }
  int* dst = vals_out + lv.elem_base[l] + (size_t)b * per;
  for (int j = threadIdx.x; j < k; j += 1024) dst[j] = (int)(uint32_t)(list[j] & 0xffffffffull);
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

// The same selection in ONE launch, one CTA per image.  Every survivor of the image (all levels, in
// cat_boxlist order) gets the key (descending-score bits << 32 | position): a total order whose first
// top_n elements are what the stable radix sort of the whole batch puts first.  The CTA finds the
// top_n-th smallest key by an 8-pass radix select (shared histograms), collects the keys up to it and
// sorts just those with a bitonic network (<= 4096 keys; the source row is looked up by position
// afterwards, so only keys move).  Keys and source rows of up to RSEL_CAP survivors are cached in
// shared memory; a larger image recomputes them from global memory in every pass (slow, correct).
constexpr int RSEL_CAP = 8192;
constexpr int RSEL_TOP = 4096;
constexpr int RSEL_THREADS = 1024;

struct RselImage {
  const float* scores;
  const int *offsets, *keep;
  const int* lvl_start;  // shared
  int L, N, b;
  __device__ __forceinline__ int src_at(int pos) const {
    int l = 0;  // last level whose start is <= pos (empty levels share a start: take the last)
    for (int q = 1; q < L; ++q)
      if (lvl_start[q] <= pos) l = q;
    const int s = l * N + b;
    return offsets[s] + keep[offsets[s] + (pos - lvl_start[l])];
  }
  __device__ __forceinline__ uint64_t key_at(int pos) const {
    return ((uint64_t)rpn_desc_bits(scores[src_at(pos)]) << 32) | (uint32_t)pos;
  }
};

__global__ void __launch_bounds__(RSEL_THREADS)
rpn_sel_image_kernel(const float* __restrict__ boxes, const float* __restrict__ scores,
                     const int* __restrict__ offsets, const int* __restrict__ keep,
                     const int* __restrict__ n_keep, int L, int N, int top_n,
                     float* __restrict__ out_boxes, float* __restrict__ out_scores,
                     int* __restrict__ out_count) {
  extern __shared__ __align__(16) unsigned char rsel_smem[];
  uint64_t* sk = reinterpret_cast<uint64_t*>(rsel_smem);   // [RSEL_CAP] keys by position
  uint64_t* sel = sk + RSEL_CAP;                           // [RSEL_TOP] selected keys
  int* sv = reinterpret_cast<int*>(sel + RSEL_TOP);        // [RSEL_CAP] source rows by position
  __shared__ int lvl_start[RP_MAX_LEVELS + 1];
  __shared__ int hist[256];
  __shared__ int s_digit, s_k, s_nsel;
  const int b = blockIdx.x, tid = threadIdx.x;
  if (tid == 0) {
    int p = 0;
    for (int l = 0; l < L; ++l) {
      lvl_start[l] = p;
      p += n_keep[l * N + b];
    }
    lvl_start[L] = p;
    s_nsel = 0;
  }
  __syncthreads();
  const int cnt = lvl_start[L];
  const RselImage im{scores, offsets, keep, lvl_start, L, N, b};
  const bool cached = cnt <= RSEL_CAP;
  if (cached) {
    for (int pos = tid; pos < cnt; pos += RSEL_THREADS) {
      const int src = im.src_at(pos);
      sv[pos] = src;
      sk[pos] = ((uint64_t)rpn_desc_bits(scores[src]) << 32) | (uint32_t)pos;
    }
    __syncthreads();
  }
  const int n_out = min(cnt, top_n);
  uint64_t kth = ~0ull;  // every key is <= kth
  if (cnt > top_n) {
    // radix select of the top_n-th smallest key, most significant byte first
    uint64_t prefix = 0ull;
    int k = top_n;  // rank (1-based) among the keys that match the prefix
    for (int shift = 56; shift >= 0; shift -= 8) {
      if (tid < 256) hist[tid] = 0;
      __syncthreads();
      const uint64_t mask = shift == 56 ? 0ull : (~0ull << (shift + 8));
      for (int pos = tid; pos < cnt; pos += RSEL_THREADS) {
        const uint64_t key = cached ? sk[pos] : im.key_at(pos);
        if ((key & mask) == prefix) atomicAdd(&hist[(int)((key >> shift) & 0xffull)], 1);
      }
      __syncthreads();
      if (tid < 32) {
        // 8 bins per lane: find the digit whose cumulative count reaches k
        int c[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          c[i] = hist[tid * 8 + i];
          sum += c[i];
        }
        int incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, incl, d);
          if (tid >= d) incl += t;
        }
        const int excl = incl - sum;
        if (excl < k && k <= incl) {
          int run = excl;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (run < k && k <= run + c[i]) {
              s_digit = tid * 8 + i;
              s_k = k - run;
            }
            run += c[i];
          }
        }
      }
      __syncthreads();
      prefix |= (uint64_t)s_digit << shift;
      k = s_k;
    }
    kth = prefix;
  }
  // collect the keys up to kth (exactly n_out of them: keys are unique), sort, emit
  for (int pos = tid; pos < cnt; pos += RSEL_THREADS) {
    const uint64_t key = cached ? sk[pos] : im.key_at(pos);
    if (key <= kth) sel[atomicAdd(&s_nsel, 1)] = key;
  }
  int P2 = 32;
  while (P2 < n_out) P2 <<= 1;
  __syncthreads();
  for (int i = n_out + tid; i < P2; i += RSEL_THREADS) sel[i] = ~0ull;
  __syncthreads();
  for (int k = 2; k <= P2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < (P2 >> 1); i += RSEL_THREADS) {
        const int lo = ((i / j) * (j << 1)) + (i & (j - 1)), hi = lo + j;
        const uint64_t a = sel[lo], c = sel[hi];
        if ((a > c) == ((lo & k) == 0)) {
          sel[lo] = c;
          sel[hi] = a;
        }
      }
      __syncthreads();
    }
  }
  if (tid == 0) out_count[b] = n_out;
  for (int j = tid; j < n_out; j += RSEL_THREADS) {
    const int pos = (int)(uint32_t)sel[j];
    const int src = cached ? sv[pos] : im.src_at(pos);
    const size_t o = (size_t)b * top_n + j;
    *reinterpret_cast<float4*>(out_boxes + o * 4) = *reinterpret_cast<const float4*>(boxes + (size_t)src * 4);
    out_scores[o] = scores[src];
  }
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
  RsPlan pl;
  int n_chunks = 0;
  for (int l = 0; l < RP_MAX_LEVELS; ++l) pl.cpp[l] = 0;
  for (int l = 0; l < L; ++l) {
    pl.chunk_base[l] = n_chunks;
    pl.cpp[l] = (lv.A[l] * lv.HW[l] + RS_CHUNK - 1) / RS_CHUNK;
    n_chunks += N * pl.cpp[l];
  }
  for (int l = L; l <= RP_MAX_LEVELS; ++l) pl.chunk_base[l] = n_chunks;
  // scratch carved from the sort path's key buffers (8 bytes per element each)
  const size_t small_bytes = rp_align_up(sizeof(int) * ((size_t)4 * S * 256 + (size_t)4 * S * 2 + S), 256);
  const size_t need_in = rp_align_up((size_t)n_elems * 4, 256) + small_bytes + rp_align_up(4 * (size_t)n_chunks, 256);
  const size_t need_out = sizeof(unsigned long long) * (size_t)S * RS_CAP;
  if (kmax <= RS_CAP && need_in <= 8 * (size_t)n_elems && need_out <= 8 * (size_t)n_elems &&
      rp_env_int("RP_RPN_SELECT", 1)) {
    // per-problem radix select + small sort (no pass over all anchors of the batch with 64-bit keys)
    char* base = reinterpret_cast<char*>(w.keys_in);
    pl.keys = reinterpret_cast<uint32_t*>(base);
    char* small = base + rp_align_up((size_t)n_elems * 4, 256);
    pl.hist = reinterpret_cast<int*>(small);
    pl.state = pl.hist + (size_t)4 * S * 256;
    pl.cursor = pl.state + (size_t)4 * S * 2;
    pl.tie_cnt = reinterpret_cast<int*>(small + small_bytes);
    pl.list = reinterpret_cast<unsigned long long*>(w.keys_out);
    RP_CUDA_TRY(cudaMemsetAsync(small, 0, small_bytes, st));
    rs_keys_kernel<<<n_chunks, 256, 0, st>>>(lv, pl, S);
    for (int pass = 2; pass >= 0; --pass) rs_hist_kernel<<<n_chunks, 256, 0, st>>>(lv, pl, S, pass);
    rs_ties_kernel<<<n_chunks, 256, 0, st>>>(lv, pl, S);
    rs_collect_kernel<<<n_chunks, 256, 0, st>>>(lv, pl, S);
    rs_sort_kernel<<<S, 1024, 0, st>>>(lv, pl, w.vals_out);
    RP_LAUNCH_CHECK();
  } else {
    rpn_keys_kernel<<<(unsigned)((n_elems + 255) / 256), 256, 0, st>>>(lv, w.keys_in, w.vals_in);
    RP_LAUNCH_CHECK();
    RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out, w.vals_in,
                                                w.vals_out, n_elems, 0, 32 + rpn_bits((size_t)S), st));
  }
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
  if (L <= RP_MAX_LEVELS && top_n <= RSEL_TOP && rp_env_int("RP_RPN_SEL_FUSED", 1)) {
    const size_t smem = (size_t)RSEL_CAP * (sizeof(uint64_t) + sizeof(int)) + (size_t)RSEL_TOP * sizeof(uint64_t);
    RP_SET_MAX_SMEM(rpn_sel_image_kernel, smem);
    rpn_sel_image_kernel<<<N, RSEL_THREADS, smem, st>>>(boxes, scores, offsets, keep, n_keep, L, N, top_n,
                                                        out_boxes, out_scores, out_count);
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
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
