// nms.cu -- batched greedy NMS / multi-label NMS / cython-variant NMS for sm_100a.
//
// Reference being replaced (paths relative to the reference checkout):
//   det/layers/nms.py:12-14      nms -> jittor.nms
//   det/layers/nms.py:18-40      _ml_nms: jt.argsort(desc) + jt.candidate(fail_cond) with the
//                                legacy +1 IoU (:28-33) and the label term (:34)
//   det/structures/boxlist_ops.py:12-60   boxlist_nms / boxlist_ml_nms (keep[:max_proposals])
//   yol/utils/cython_nms.pyx:24-74        CPU greedy NMS, ">=" comparator, ascending output
//
// Pipeline (all problems of a batch in the same launches, no host synchronisation):
//   1. keys    64-bit key = (problem id << 32) | order-reversed score bits, value = element
//   2. sort    ONE stable LSD radix sort over the whole batch (cub::DeviceRadixSort) -- the
//              stable descending score order of jt.argsort; ties keep ascending index
//   3. gather  boxes (+area, +label) into rank order
//   4. mask    upper-triangular suppression bit matrix in 64x64 tiles: lanes = 32 consecutive
//              rows, the 64 column boxes of the tile broadcast from shared memory.  The IoU
//              test is exact: a pair is rejected without a division when inter < t_lo * union
//              (t_lo just below the threshold), everything else takes the IEEE division.
//   5. scan    one CTA per problem walks the ranks in 64-box chunks: a single thread resolves
//              the chunk against its 64x64 diagonal block, all threads OR the kept rows into
//              the removed bitmap; stops at max_keep.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

constexpr int NMS_TILE = 64;
constexpr int NMS_COLS_PER_CTA = 4;  // column tiles per CTA in the mask kernel
constexpr int NMS_SCAN_THREADS = 512;

__device__ __forceinline__ uint32_t nms_desc_bits(float s) {
  if (s == 0.0f) s = 0.0f;  // -0 and +0 tie, as in a comparison sort
  uint32_t u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending-orderable
  return ~u;                                       // descending
}

// problem of global element i: largest b with offsets[b] <= i
__device__ __forceinline__ int nms_find_segment(const int* __restrict__ offsets, int B, int i) {
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

__global__ void nms_keys_kernel(const float* __restrict__ scores, const int* __restrict__ offsets,
                                int B, int n_total, int ties_high_first,
                                uint64_t* __restrict__ keys, int* __restrict__ vals) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_total) return;
  const int b = nms_find_segment(offsets, B, p);
  // With ties_high_first the segment is fed to the stable sort in reversed element order, so
  // equal scores come out with the HIGHER index first (argsort()[::-1] of cython_nms.pyx:35).
  const int e = ties_high_first ? (offsets[b] + offsets[b + 1] - 1 - p) : p;
  keys[p] = ((uint64_t)(uint32_t)b << 32) | (uint64_t)nms_desc_bits(scores[e]);
  vals[p] = e;
}

struct __align__(16) NmsBox {
  float x1, y1, x2, y2;
};

__global__ void nms_gather_kernel(const float* __restrict__ boxes, const float* __restrict__ labels,
                                  const int* __restrict__ sorted_vals, int n_total,
                                  NmsBox* __restrict__ sboxes, float* __restrict__ sarea,
                                  float* __restrict__ slabel) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_total) return;
  const int e = sorted_vals[p];
  const float4 b = *reinterpret_cast<const float4*>(boxes + (size_t)e * 4);
  NmsBox o{b.x, b.y, b.z, b.w};
  sboxes[p] = o;
  sarea[p] = rp_area_plus1(b.x, b.y, b.z, b.w);
  if (labels) slabel[p] = labels[e];
}

// One CTA: 64 rows x (NMS_COLS_PER_CTA * 64) columns of one problem's bit matrix.
// thread = (column tile j = tid / 64, row r = tid % 64).
template <bool GE, bool LABELS>
__global__ void __launch_bounds__(NMS_TILE * NMS_COLS_PER_CTA)
nms_mask_kernel(const NmsBox* __restrict__ sboxes, const float* __restrict__ sarea,
                const float* __restrict__ slabel, const int* __restrict__ offsets,
                const size_t* __restrict__ mask_off, float t_cmp, float t_lo,
                uint64_t* __restrict__ mask) {
  const int b = blockIdx.z;
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
  const int W = (n + 63) >> 6;
  const int rb = blockIdx.y;
  const int cb0 = blockIdx.x * NMS_COLS_PER_CTA;
  if (rb >= W || cb0 >= W) return;
  if (cb0 + NMS_COLS_PER_CTA - 1 < rb) return;  // entirely below the diagonal

  __shared__ NmsBox cbox[NMS_TILE * NMS_COLS_PER_CTA];
  __shared__ float carea[NMS_TILE * NMS_COLS_PER_CTA];
  __shared__ float clabel[NMS_TILE * NMS_COLS_PER_CTA];
  const int tid = threadIdx.x;
  {
    const int c = cb0 * 64 + tid;
    if (c < n) {
      cbox[tid] = sboxes[off + c];
      carea[tid] = sarea[off + c];
      if (LABELS) clabel[tid] = slabel[off + c];
    }
  }
  __syncthreads();
  const int j = tid >> 6;
  const int r = tid & 63;
  const int cb = cb0 + j;
  const int row = rb * 64 + r;
  if (cb >= W || cb < rb || row >= n) return;
  const NmsBox a = sboxes[off + row];
  const float aarea = sarea[off + row];
  float alabel = 0.f;
  if (LABELS) alabel = slabel[off + row];
  const int ncol = min(64, n - cb * 64);
  const int start = (cb == rb) ? r + 1 : 0;  // strict upper triangle
  uint64_t bits = 0;
  const NmsBox* cbx = cbox + j * 64;
  const float* car = carea + j * 64;
  const float* clb = clabel + j * 64;
  for (int c = start; c < ncol; ++c) {
    const NmsBox q = cbx[c];
    const float iw = rp_fmaxf(0.0f, (rp_fminf(a.x2, q.x2) - rp_fmaxf(a.x1, q.x1)) + 1.0f);
    const float ih = rp_fmaxf(0.0f, (rp_fminf(a.y2, q.y2) - rp_fmaxf(a.y1, q.y1)) + 1.0f);
    const float inter = ih * iw;
    const float uni = (aarea + car[c]) - inter;
    // exact filter: fl(inter/uni) > t (or >= t) is impossible when inter < t_lo*uni with
    // uni > 0 (t_lo*uni <= 0 <= inter when uni <= 0, so such pairs fall through).
    if (inter < t_lo * uni) continue;
    const float iou = inter / uni;
    bool sup = GE ? (iou >= t_cmp) : (iou > t_cmp);
    if (LABELS) sup = sup && (alabel == clb[c]);
    if (sup) bits |= (1ull << c);
  }
  mask[mask_off[b] + (size_t)row * W + cb] = bits;
}

// One CTA per problem.  keep receives relative element indices in rank order.
__global__ void __launch_bounds__(NMS_SCAN_THREADS)
nms_scan_kernel(const uint64_t* __restrict__ mask, const size_t* __restrict__ mask_off,
                const int* __restrict__ offsets, const int* __restrict__ sorted_vals,
                int max_keep, int ascending, int* __restrict__ keep, int* __restrict__ n_keep,
                uint8_t* __restrict__ flags) {
  extern __shared__ uint64_t removed[];  // W words
  __shared__ uint64_t diag[64];
  __shared__ uint64_t s_kept;
  __shared__ int s_count;
  const int b = blockIdx.x;
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
  const int W = (n + 63) >> 6;
  const int tid = threadIdx.x;
  const uint64_t* M = mask + mask_off[b];
  for (int w = tid; w < W; w += NMS_SCAN_THREADS) removed[w] = 0;
  if (tid == 0) s_count = 0;
  if (ascending)
    for (int i = tid; i < n; i += NMS_SCAN_THREADS) flags[off + i] = 0;
  __syncthreads();
  for (int c = 0; c < W; ++c) {
    const int nrow = min(64, n - c * 64);
    if (tid < nrow) diag[tid] = M[(size_t)(c * 64 + tid) * W + c];
    __syncthreads();
    if (tid == 0) {
      uint64_t rem = removed[c];
      const uint64_t valid = (nrow == 64) ? ~0ull : ((1ull << nrow) - 1ull);
      uint64_t alive = ~rem & valid;
      uint64_t kept = 0;
      int count = s_count;
      while (alive) {
        const int i = __ffsll((long long)alive) - 1;
        kept |= (1ull << i);
        ++count;
        if (max_keep > 0 && count >= max_keep) break;
        rem |= diag[i];
        alive &= ~rem;
        alive &= ~(1ull << i);
      }
      s_kept = kept;
    }
    __syncthreads();
    const uint64_t kept = s_kept;
    const int base = s_count;
    // emit the kept boxes of this chunk (rank order)
    if (tid < 64 && ((kept >> tid) & 1ull)) {
      const int pos = base + __popcll(kept & ((1ull << tid) - 1ull));
      const int e = sorted_vals[off + c * 64 + tid] - off;
      if (ascending)
        flags[off + e] = 1;
      else
        keep[off + pos] = e;
    }
    const int total = base + __popcll(kept);
    const bool done = (max_keep > 0 && total >= max_keep);
    if (!done) {
      // OR the kept rows into the removed bitmap, words right of the diagonal
      for (int w = c + 1 + tid; w < W; w += NMS_SCAN_THREADS) {
        uint64_t acc = removed[w];
        uint64_t k = kept;
        while (k) {
          const int i = __ffsll((long long)k) - 1;
          k &= k - 1;
          acc |= M[(size_t)(c * 64 + i) * W + w];
        }
        removed[w] = acc;
      }
    }
    __syncthreads();
    if (tid == 0) s_count = total;
    if (done) break;
    __syncthreads();
  }
  __syncthreads();
  if (!ascending) {
    if (tid == 0) n_keep[b] = s_count;
    return;
  }
  // ascending output (cython_nms.pyx:74 np.where(suppressed == 0)): ordered compaction of flags
  __shared__ int s_warp[NMS_SCAN_THREADS / 32];
  __shared__ int s_base;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += NMS_SCAN_THREADS) {
    const int i = i0 + tid;
    const int f = (i < n) ? flags[off + i] : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f);
    const int lane = tid & 31, wid = tid >> 5;
    if (lane == 0) s_warp[wid] = __popc(bal);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < wid; ++w) before += s_warp[w];
    if (f) keep[off + before + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < NMS_SCAN_THREADS / 32; ++w) t += s_warp[w];
      s_base += t;
    }
    __syncthreads();
  }
  if (tid == 0) n_keep[b] = s_base;
}

// prefix of per-problem mask sizes (B is small: serial)
__global__ void nms_mask_offsets_kernel(const int* __restrict__ offsets, int B,
                                        size_t* __restrict__ mask_off) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    size_t acc = 0;
    for (int b = 0; b < B; ++b) {
      mask_off[b] = acc;
      const size_t n = (size_t)(offsets[b + 1] - offsets[b]);
      acc += n * ((n + 63) >> 6);
    }
    mask_off[B] = acc;
  }
}

struct NmsWorkspace {
  uint64_t *keys_in, *keys_out;
  int *vals_in, *vals_out;
  NmsBox* sboxes;
  float *sarea, *slabel;
  size_t* mask_off;
  uint8_t* flags;
  void* cub_temp;
  size_t cub_bytes;
  uint64_t* mask;
  size_t mask_words;
  size_t total;
};

size_t nms_cub_bytes(int n_total) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint64_t*)nullptr,
                                  (uint64_t*)nullptr, (const int*)nullptr, (int*)nullptr, n_total,
                                  0, 64, (cudaStream_t)0);
  return bytes;
}

// Upper bound of sum_b n_b * ceil(n_b / 64) given only (n_total, n_max, B): the sum is convex
// in the segment lengths, so it is maximal when as many segments as possible have n_max rows.
size_t nms_mask_words_bound(int n_total, int n_max, int B) {
  if (n_max <= 0 || n_total <= 0) return 0;
  const size_t wmax = ((size_t)n_max + 63) >> 6;
  size_t full = (size_t)n_total / (size_t)n_max;
  if (full > (size_t)B) full = (size_t)B;
  const size_t rest = (size_t)n_total - full * (size_t)n_max;
  return full * (size_t)n_max * wmax + rest * ((rest + 63) >> 6) + 64;
}

NmsWorkspace nms_carve(void* base, int n_total, int n_max, int B) {
  NmsWorkspace w;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t at = o;
    o = rp_align_up(o + bytes, 256);
    return base ? (void*)((char*)base + at) : (void*)nullptr;
  };
  const size_t n = (size_t)(n_total > 0 ? n_total : 1);
  w.keys_in = (uint64_t*)take(n * 8);
  w.keys_out = (uint64_t*)take(n * 8);
  w.vals_in = (int*)take(n * 4);
  w.vals_out = (int*)take(n * 4);
  w.sboxes = (NmsBox*)take(n * sizeof(NmsBox));
  w.sarea = (float*)take(n * 4);
  w.slabel = (float*)take(n * 4);
  w.mask_off = (size_t*)take(((size_t)B + 1) * sizeof(size_t));
  w.flags = (uint8_t*)take(n);
  w.cub_bytes = nms_cub_bytes(n_total > 0 ? n_total : 1);
  w.cub_temp = take(w.cub_bytes);
  w.mask_words = nms_mask_words_bound(n_total, n_max, B);
  w.mask = (uint64_t*)take(w.mask_words * 8);
  w.total = o;
  return w;
}

}  // namespace

extern "C" {

size_t rp_nms_workspace_bytes(int n_total, int n_max, int B) {
  if (n_total < 0 || n_max < 0 || B < 1) return 0;
  return nms_carve(nullptr, n_total, n_max, B).total;
}

int rp_nms_batched(const float* boxes, const float* scores, const float* labels,
                   const int* offsets, int B, int n_total, int n_max, double thr, int max_keep,
                   int flags, int* keep, int* n_keep, void* workspace, size_t workspace_bytes,
                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (B < 1 || !offsets || !n_keep || n_total < 0 || n_max < 0 || n_max > n_total) return RP_E_ARG;
  if (n_total == 0) {
    RP_CUDA_TRY(cudaMemsetAsync(n_keep, 0, sizeof(int) * (size_t)B, st));
    return RP_OK;
  }
  if (!boxes || !scores || !keep || !workspace) return RP_E_ARG;
  if ((reinterpret_cast<uintptr_t>(boxes) & 15) != 0) return RP_E_ARG;  // float4 gathers
  if (n_max > 65535 * 64) return RP_E_LIMIT;
  NmsWorkspace w = nms_carve(workspace, n_total, n_max, B);
  if (w.total > workspace_bytes) return RP_E_WORKSPACE;
  const bool ge = (flags & RP_NMS_GE) != 0;
  const int ascending = (flags & RP_NMS_ASCENDING) ? 1 : 0;
  // Reference comparators: (double)iou > thr (thr pasted as a double literal, nms.py:23,34)
  // <=> iou > float_le(thr);  cython: float iou >= float thresh (cython_nms.pyx:24,69).
  const float t_cmp = ge ? (float)thr : rp_float_le(thr);
  float t_lo;
  if (t_cmp > 0.0f && t_cmp < 3.0e38f)
    t_lo = t_cmp * 0.999999f;  // 2^-20 margin >> the three roundings involved
  else
    t_lo = -INFINITY;  // never reject early: every pair takes the exact division
  const int T = 256;
  const int nb = (n_total + T - 1) / T;
  nms_keys_kernel<<<nb, T, 0, st>>>(scores, offsets, B, n_total,
                                    (flags & RP_NMS_TIES_HIGH_FIRST) ? 1 : 0, w.keys_in, w.vals_in);
  RP_LAUNCH_CHECK();
  int seg_bits = 1;
  while ((1 << seg_bits) < B) ++seg_bits;
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out,
                                              w.vals_in, w.vals_out, n_total, 0, 32 + seg_bits,
                                              st));
  nms_gather_kernel<<<nb, T, 0, st>>>(boxes, labels, w.vals_out, n_total, w.sboxes, w.sarea,
                                      w.slabel);
  RP_LAUNCH_CHECK();
  nms_mask_offsets_kernel<<<1, 32, 0, st>>>(offsets, B, w.mask_off);
  RP_LAUNCH_CHECK();
  const int Wmax = (n_max + 63) / 64;
  dim3 grid((unsigned)((Wmax + NMS_COLS_PER_CTA - 1) / NMS_COLS_PER_CTA), (unsigned)Wmax,
            (unsigned)B);
  if (grid.z > 65535) return RP_E_LIMIT;
  const int mt = NMS_TILE * NMS_COLS_PER_CTA;
  if (labels) {
    if (ge)
      nms_mask_kernel<true, true><<<grid, mt, 0, st>>>(w.sboxes, w.sarea, w.slabel, offsets,
                                                       w.mask_off, t_cmp, t_lo, w.mask);
    else
      nms_mask_kernel<false, true><<<grid, mt, 0, st>>>(w.sboxes, w.sarea, w.slabel, offsets,
                                                        w.mask_off, t_cmp, t_lo, w.mask);
  } else {
    if (ge)
      nms_mask_kernel<true, false><<<grid, mt, 0, st>>>(w.sboxes, w.sarea, w.slabel, offsets,
                                                        w.mask_off, t_cmp, t_lo, w.mask);
    else
      nms_mask_kernel<false, false><<<grid, mt, 0, st>>>(w.sboxes, w.sarea, w.slabel, offsets,
                                                         w.mask_off, t_cmp, t_lo, w.mask);
  }
  RP_LAUNCH_CHECK();
  const size_t smem = sizeof(uint64_t) * (size_t)Wmax;
  if (smem > 200 * 1024) return RP_E_LIMIT;
  if (smem > 48 * 1024)
    RP_CUDA_TRY(cudaFuncSetAttribute(nms_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)smem));
  nms_scan_kernel<<<B, NMS_SCAN_THREADS, smem, st>>>(w.mask, w.mask_off, offsets, w.vals_out,
                                                     max_keep, ascending, keep, n_keep, w.flags);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
