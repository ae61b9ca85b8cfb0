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
//   4. bands   ranks are processed in bands of <= 1024 boxes (see "Banded greedy NMS" below):
//              per band one `resolve` launch (greedy inside the band from its diagonal bit block,
//              warp fixpoint on REDUX OR reductions) and one GPU-wide `suppress` launch (every
//              later, still alive box against the boxes the band kept; early exit).  The IoU
//              test is exact: a pair is rejected without a division when inter < t_lo * union
//              (t_lo just below the threshold), everything else takes the IEEE division.
//              No n x n/64 bit matrix exists; workspace is O(n).
//   5. finish  ordered compaction of the kept flags: keep indices in descending-score order (or
//              ascending index for the cython variant), n_keep, truncation at max_keep.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

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
  if (p >= offsets[B]) {
    // n_total may be an upper bound of offsets[B] (a device-side count the host never reads,
    // e.g. the output of rp_rpn_pre_nms): the slack rows sort behind the last problem
    keys[p] = ((uint64_t)(uint32_t)B << 32);
    vals[p] = p;
    return;
  }
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

// ------------------------------------------------------------------------------------------
// Banded greedy NMS.  Ranks are processed in bands of BR <= 1024 boxes.  Per band b:
//   resolve   (one CTA per problem) greedy NMS inside the band from its BR x BR diagonal
//             bit block, starting from the removed bits earlier bands left on it
//   suppress  (whole GPU) every later, still alive box is tested against the <= BR boxes the
//             band KEPT (early exit on the first hit) and its removed bit is set; the same
//             launch also computes the diagonal block of band b+1, which does not depend on state
// Only pairs (kept box, later alive box) and the diagonal blocks are ever evaluated -- with the
// usual 70-95 % suppression that is a fraction of the n(n-1)/2 pairs of a full bit matrix, and
// no n x n/64 matrix is materialised.
// ------------------------------------------------------------------------------------------
constexpr int NMS_BAND_MAX = 1024;

// the reference's IoU test (det/layers/nms.py:28-34), exact: pairs with inter < t_lo * union
// are rejected without a division, the rest take the IEEE division the oracle takes.
template <bool GE>
__device__ __forceinline__ bool nms_pair(const NmsBox& a, float aarea, const NmsBox& q, float qarea,
                                         float t_cmp, float t_lo) {
  // fminf/fmaxf (one FMNMX each) instead of the oracle's compare-and-select: the two differ
  // only in the sign of a zero result (which the +1 absorbs) and for NaN coordinates.
  const float iw = fmaxf(0.0f, (fminf(a.x2, q.x2) - fmaxf(a.x1, q.x1)) + 1.0f);
  const float ih = fmaxf(0.0f, (fminf(a.y2, q.y2) - fmaxf(a.y1, q.y1)) + 1.0f);
  const float inter = ih * iw;
  const float uni = (aarea + qarea) - inter;
  // exact filter: fl(inter/uni) > t (or >= t) is impossible when inter < t_lo*uni with
  // uni > 0 (t_lo*uni <= 0 <= inter when uni <= 0, so such pairs fall through).
  if (inter < t_lo * uni) return false;
  const float iou = inter / uni;
  return GE ? (iou >= t_cmp) : (iou > t_cmp);
}

__device__ __forceinline__ uint64_t nms_warp_or64(uint64_t v) {
  const uint32_t lo = __reduce_or_sync(0xffffffffu, (uint32_t)v);
  const uint32_t hi = __reduce_or_sync(0xffffffffu, (uint32_t)(v >> 32));
  return ((uint64_t)hi << 32) | lo;
}

struct NmsBand {
  const NmsBox* sboxes;   // rank order, all problems
  const float* sarea;
  const float* slabel;
  const int* offsets;
  const int* sorted_vals;
  uint64_t* removed;      // [B][wmax]
  uint64_t* D;            // [B][BR][bw]  diagonal block of the current band (upper triangle)
  NmsBox* kbox;           // [B][BR]      boxes the current band kept
  float* karea;
  float* klabel;
  int* kcount;            // [B] kept in the current band
  int* count;             // [B] kept so far
  int* done;              // [B] max_keep reached
  int wmax, BR, bw;
  float t_cmp, t_lo;
};

// Diagonal block of `band`: CTA = 64 rows x 4 column tiles (as thread = (tile j, row r)).
template <bool GE, bool LABELS>
__device__ __forceinline__ void nms_diag_part(const NmsBand& p, int seg, int band, int blk,
                                              NmsBox* cbox, float* carea, float* clabel) {
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int base = band * p.BR;
  if (base >= n) return;
  const int groups = (p.bw + 3) >> 2;
  const int rb = blk / groups, cb0 = (blk - rb * groups) * 4;
  if (rb >= p.bw || cb0 + 3 < rb) return;
  const int tid = threadIdx.x;
  {
    const int c = base + cb0 * 64 + tid;
    if (cb0 * 64 + tid < p.BR && c < n) {
      cbox[tid] = p.sboxes[off + c];
      carea[tid] = p.sarea[off + c];
      if (LABELS) clabel[tid] = p.slabel[off + c];
    }
  }
  __syncthreads();
  const int j = tid >> 6, r = tid & 63;
  const int cb = cb0 + j;
  const int rrow = rb * 64 + r;  // row inside the band
  const int row = base + rrow;
  if (cb >= p.bw || cb < rb || row >= n) return;
  const NmsBox a = p.sboxes[off + row];
  const float aarea = p.sarea[off + row];
  float alabel = 0.f;
  if (LABELS) alabel = p.slabel[off + row];
  int ncol = n - (base + cb * 64);
  ncol = ncol > 64 ? 64 : ncol;
  uint64_t up = 0;
  const NmsBox* cbx = cbox + j * 64;
  const float* car = carea + j * 64;
  const float* clb = clabel + j * 64;
  for (int c = (cb == rb) ? r + 1 : 0; c < ncol; ++c) {  // strict upper triangle
    bool sup = nms_pair<GE>(a, aarea, cbx[c], car[c], p.t_cmp, p.t_lo);
    if (LABELS) sup = sup && (alabel == clb[c]);
    if (sup) up |= (1ull << c);
  }
  p.D[((size_t)seg * p.BR + rrow) * p.bw + cb] = up;
}

template <bool GE, bool LABELS>
__global__ void __launch_bounds__(256)
nms_diag_kernel(NmsBand p, int band) {
  __shared__ NmsBox cbox[256];
  __shared__ float carea[256];
  __shared__ float clabel[256];
  nms_diag_part<GE, LABELS>(p, blockIdx.y, band, blockIdx.x, cbox, carea, clabel);
}

// Blocks [0, nsup): 64 later boxes each against the boxes band `band` kept.
// Blocks [nsup, ...): diagonal block of band + 1.
template <bool GE, bool LABELS>
__global__ void __launch_bounds__(256)
nms_suppress_kernel(NmsBand p, int band, int nsup) {
  __shared__ NmsBox kb[NMS_BAND_MAX];
  __shared__ float ka[NMS_BAND_MAX];
  __shared__ float kl[LABELS ? NMS_BAND_MAX : 1];
  __shared__ unsigned long long s_new;
  const int seg = blockIdx.y;
  const int tid = threadIdx.x;
  if ((int)blockIdx.x >= nsup) {
    nms_diag_part<GE, LABELS>(p, seg, band + 1, blockIdx.x - nsup, kb, ka, kl);
    return;
  }
  if (p.done[seg]) return;
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int wj = (band + 1) * p.bw + blockIdx.x;
  if (wj * 64 >= n) return;
  const int kc = p.kcount[seg];
  if (kc == 0) return;
  for (int i = tid; i < kc; i += 256) {
    kb[i] = p.kbox[(size_t)seg * p.BR + i];
    ka[i] = p.karea[(size_t)seg * p.BR + i];
    if (LABELS) kl[i] = p.klabel[(size_t)seg * p.BR + i];
  }
  if (tid == 0) s_new = 0ull;
  __syncthreads();
  const int jj = tid & 63, part = tid >> 6;
  const int j = wj * 64 + jj;
  const uint64_t word = p.removed[(size_t)seg * p.wmax + wj];
  if (j < n && !((word >> jj) & 1ull)) {
    const NmsBox q = p.sboxes[off + j];
    const float qarea = p.sarea[off + j];
    float qlabel = 0.f;
    if (LABELS) qlabel = p.slabel[off + j];
    bool hit = false;
    for (int i = part; i < kc && !hit; i += 4) {
      bool sup = nms_pair<GE>(kb[i], ka[i], q, qarea, p.t_cmp, p.t_lo);
      if (LABELS) sup = sup && (kl[i] == qlabel);
      hit = sup;
    }
    if (hit) atomicOr(&s_new, 1ull << jj);
  }
  __syncthreads();
  if (tid == 0 && s_new) p.removed[(size_t)seg * p.wmax + wj] = word | s_new;
}

// Greedy NMS inside band `band`, one CTA per problem.
__global__ void __launch_bounds__(1024)
nms_resolve_kernel(NmsBand p, int band, int max_keep, int ascending, int* __restrict__ keep,
                   uint8_t* __restrict__ flags) {
  extern __shared__ __align__(16) uint64_t nr_smem[];
  uint64_t* Ds = nr_smem;               // [BR][bw]
  uint64_t* rem = Ds + p.BR * p.bw;     // [bw]
  uint64_t* keptm = rem + p.bw;         // [bw]
  int* cbase = reinterpret_cast<int*>(keptm + p.bw);  // [bw + 1]
  const int seg = blockIdx.x;
  const int tid = threadIdx.x;
  if (tid == 0) p.kcount[seg] = 0;
  if (p.done[seg]) return;
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int base = band * p.BR;
  if (base >= n) return;
  const int rows = min(p.BR, n - base);
  const int chunks = (rows + 63) >> 6;
  {
    const uint64_t* Dg = p.D + (size_t)seg * p.BR * p.bw;
    if ((p.bw & 1) == 0) {  // 16-byte copies (rows * bw is even, both bases 16-byte aligned)
      const uint4* g4 = reinterpret_cast<const uint4*>(Dg);
      uint4* s4 = reinterpret_cast<uint4*>(Ds);
      const int n4 = (rows * p.bw) >> 1;
#pragma unroll 8
      for (int i = tid; i < n4; i += blockDim.x) s4[i] = g4[i];
    } else {
      for (int i = tid; i < rows * p.bw; i += blockDim.x) Ds[i] = Dg[i];
    }
    if (tid < p.bw) {
      rem[tid] = p.removed[(size_t)seg * p.wmax + band * p.bw + tid];
      keptm[tid] = 0ull;
    }
  }
  __syncthreads();
  const int count0 = p.count[seg];
  if (tid < 32) {
    const int l = tid;
    int count = count0;
    for (int c = 0; c < chunks; ++c) {
      const int nrow = min(64, rows - c * 64);
      const uint64_t valid = (nrow == 64) ? ~0ull : ((1ull << nrow) - 1ull);
      // Greedy rule as a warp-parallel fixpoint.  Lane l holds the diagonal-block rows of boxes
      // l and l+32 (bits = the lower-scored boxes of the chunk each one overlaps).  A box is
      // suppressed once a KEPT higher-scored box overlaps it, and kept once no undecided
      // higher-scored box overlaps it; the lowest undecided box is always decided, and chains
      // of overlapping boxes inside 64 ranks are short.
      uint64_t undec = ~rem[c] & valid;
      uint64_t kept = 0ull;
      const uint64_t r0 = (l < nrow) ? Ds[(c * 64 + l) * p.bw + c] : 0ull;
      const uint64_t r1 = (l + 32 < nrow) ? Ds[(c * 64 + 32 + l) * p.bw + c] : 0ull;
      while (undec) {
        const uint64_t ks = (((kept >> l) & 1ull) ? r0 : 0ull) | (((kept >> (l + 32)) & 1ull) ? r1 : 0ull);
        const uint64_t us = (((undec >> l) & 1ull) ? r0 : 0ull) | (((undec >> (l + 32)) & 1ull) ? r1 : 0ull);
        const uint64_t sup = nms_warp_or64(ks);  // overlapped by a kept box
        const uint64_t pot = nms_warp_or64(us);  // overlapped by an undecided box
        const uint64_t ns = undec & sup;
        const uint64_t nk = undec & ~sup & ~pot;
        kept |= nk;
        undec &= ~(nk | ns);
      }
      bool stop = false;
      if (max_keep > 0) {
        const int allowed = max_keep - count;
        while (__popcll(kept) > allowed) kept &= ~(1ull << (63 - __clzll((long long)kept)));
        stop = (count + __popcll(kept) >= max_keep);
      }
      if (l == 0) {
        keptm[c] = kept;
        cbase[c] = count - count0;
      }
      count += __popcll(kept);
      if (stop) break;
      // removed bits of the later chunks of this band: OR of the kept rows, word by word
      {
        const bool k0 = (kept >> l) & 1ull, k1 = (kept >> (l + 32)) & 1ull;
        const uint64_t* d0 = Ds + (c * 64 + l) * p.bw;
        const uint64_t* d1 = Ds + (c * 64 + 32 + l) * p.bw;
        for (int w = c + 1; w < chunks; ++w) {
          const uint64_t v = (k0 ? d0[w] : 0ull) | (k1 ? d1[w] : 0ull);
          const uint64_t o = nms_warp_or64(v);
          if (l == 0) rem[w] |= o;
        }
      }
      __syncwarp();
    }
    if (l == 0) {
      cbase[p.bw] = count - count0;
      p.kcount[seg] = count - count0;
      p.count[seg] = count;
      if (max_keep > 0 && count >= max_keep) p.done[seg] = 1;
    }
  }
  __syncthreads();
  // emit: keep list in rank order, and the band's kept boxes for the suppress launch
  for (int t = tid; t < rows; t += blockDim.x) {
    const int c = t >> 6, bit = t & 63;
    const uint64_t kept = keptm[c];
    if ((kept >> bit) & 1ull) {
      const int pos = cbase[c] + __popcll(kept & ((1ull << bit) - 1ull));
      const int rank = off + base + t;
      const int e = p.sorted_vals[rank] - off;
      if (ascending)
        flags[off + e] = 1;
      else
        keep[off + count0 + pos] = e;
      p.kbox[(size_t)seg * p.BR + pos] = p.sboxes[rank];
      p.karea[(size_t)seg * p.BR + pos] = p.sarea[rank];
      if (p.slabel) p.klabel[(size_t)seg * p.BR + pos] = p.slabel[rank];
    }
  }
}

// n_keep, and the ascending-index output of the cython variant (cython_nms.pyx:74
// np.where(suppressed == 0)): ordered compaction of the kept flags.
__global__ void __launch_bounds__(NMS_SCAN_THREADS)
nms_finish_kernel(const int* __restrict__ offsets, const int* __restrict__ count, int ascending,
                  const uint8_t* __restrict__ flags, int* __restrict__ keep,
                  int* __restrict__ n_keep) {
  const int b = blockIdx.x;
  const int tid = threadIdx.x;
  if (!ascending) {
    if (tid == 0) n_keep[b] = count[b];
    return;
  }
  const int off = offsets[b];
  const int n = offsets[b + 1] - off;
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

struct NmsWorkspace {
  uint64_t *keys_in, *keys_out;
  int *vals_in, *vals_out;
  NmsBox* sboxes;
  float *sarea, *slabel;
  uint8_t* flags;
  void* cub_temp;
  size_t cub_bytes;
  // banded scan state; [zero_begin, zero_begin + zero_bytes) is cleared at the start of a call
  uint64_t* removed;
  int *count, *done, *kcount;
  char* zero_begin;
  size_t zero_bytes;
  uint64_t* D;
  NmsBox* kbox;
  float *karea, *klabel;
  int BR, bw, wmax;
  size_t total;
};

size_t nms_cub_bytes(int n_total) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs((void*)nullptr, bytes, (const uint64_t*)nullptr,
                                  (uint64_t*)nullptr, (const int*)nullptr, (int*)nullptr, n_total,
                                  0, 64, (cudaStream_t)0);
  return bytes;
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
  const size_t nb = (size_t)(B > 0 ? B : 1);
  w.wmax = (n_max + 63) / 64;
  if (w.wmax < 1) w.wmax = 1;
  w.bw = w.wmax < NMS_BAND_MAX / 64 ? w.wmax : NMS_BAND_MAX / 64;
  w.BR = w.bw * 64;
  w.keys_in = (uint64_t*)take(n * 8);
  w.keys_out = (uint64_t*)take(n * 8);
  w.vals_in = (int*)take(n * 4);
  w.vals_out = (int*)take(n * 4);
  w.sboxes = (NmsBox*)take(n * sizeof(NmsBox));
  w.sarea = (float*)take(n * 4);
  w.slabel = (float*)take(n * 4);
  w.flags = (uint8_t*)take(n);
  w.cub_bytes = nms_cub_bytes(n_total > 0 ? n_total : 1);
  w.cub_temp = take(w.cub_bytes);
  const size_t z0 = o;
  w.removed = (uint64_t*)take(nb * (size_t)w.wmax * 8);
  w.count = (int*)take(nb * 4);
  w.done = (int*)take(nb * 4);
  w.kcount = (int*)take(nb * 4);
  w.zero_begin = base ? (char*)base + z0 : nullptr;
  w.zero_bytes = o - z0;
  w.D = (uint64_t*)take(nb * (size_t)w.BR * w.bw * 8);
  w.kbox = (NmsBox*)take(nb * (size_t)w.BR * sizeof(NmsBox));
  w.karea = (float*)take(nb * (size_t)w.BR * 4);
  w.klabel = (float*)take(nb * (size_t)w.BR * 4);
  w.total = o;
  return w;
}

template <bool GE, bool LABELS>
int nms_run_bands(const NmsBand& p, int B, int n_max, int max_keep, int ascending, int* keep,
                  uint8_t* flags, cudaStream_t st) {
  const int bands = (n_max + p.BR - 1) / p.BR;
  const int groups = (p.bw + 3) / 4;
  const int ndiag = p.bw * groups;
  const size_t smem = sizeof(uint64_t) * ((size_t)p.BR * p.bw + 2 * (size_t)p.bw) +
                      sizeof(int) * ((size_t)p.bw + 2);
  RP_SET_MAX_SMEM(nms_resolve_kernel, smem);
  const int rthreads = p.BR >= 1024 ? 1024 : (p.BR < 128 ? 128 : p.BR);
  nms_diag_kernel<GE, LABELS><<<dim3((unsigned)ndiag, (unsigned)B), 256, 0, st>>>(p, 0);
  RP_LAUNCH_CHECK();
  for (int b = 0; b < bands; ++b) {
    nms_resolve_kernel<<<B, rthreads, smem, st>>>(p, b, max_keep, ascending, keep, flags);
    RP_LAUNCH_CHECK();
    if (b + 1 < bands) {
      const int nsup = p.wmax - (b + 1) * p.bw;
      nms_suppress_kernel<GE, LABELS><<<dim3((unsigned)(nsup + ndiag), (unsigned)B), 256, 0, st>>>(
          p, b, nsup);
      RP_LAUNCH_CHECK();
    }
  }
  return RP_OK;
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
  while ((1 << seg_bits) < B + 1) ++seg_bits;  // B problems + the slack id B
  RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out,
                                              w.vals_in, w.vals_out, n_total, 0, 32 + seg_bits,
                                              st));
  nms_gather_kernel<<<nb, T, 0, st>>>(boxes, labels, w.vals_out, n_total, w.sboxes, w.sarea,
                                      w.slabel);
  RP_LAUNCH_CHECK();
  RP_CUDA_TRY(cudaMemsetAsync(w.zero_begin, 0, w.zero_bytes, st));
  if (ascending) RP_CUDA_TRY(cudaMemsetAsync(w.flags, 0, (size_t)n_total, st));
  if (B > 65535) return RP_E_LIMIT;
  NmsBand p;
  p.sboxes = w.sboxes;
  p.sarea = w.sarea;
  p.slabel = labels ? w.slabel : nullptr;
  p.offsets = offsets;
  p.sorted_vals = w.vals_out;
  p.removed = w.removed;
  p.D = w.D;
  p.kbox = w.kbox;
  p.karea = w.karea;
  p.klabel = w.klabel;
  p.kcount = w.kcount;
  p.count = w.count;
  p.done = w.done;
  p.wmax = w.wmax;
  p.BR = w.BR;
  p.bw = w.bw;
  p.t_cmp = t_cmp;
  p.t_lo = t_lo;
  int rc;
  if (labels)
    rc = ge ? nms_run_bands<true, true>(p, B, n_max, max_keep, ascending, keep, w.flags, st)
            : nms_run_bands<false, true>(p, B, n_max, max_keep, ascending, keep, w.flags, st);
  else
    rc = ge ? nms_run_bands<true, false>(p, B, n_max, max_keep, ascending, keep, w.flags, st)
            : nms_run_bands<false, false>(p, B, n_max, max_keep, ascending, keep, w.flags, st);
  if (rc) return rc;
  nms_finish_kernel<<<B, NMS_SCAN_THREADS, 0, st>>>(offsets, w.count, ascending, w.flags, keep,
                                                    n_keep);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"
