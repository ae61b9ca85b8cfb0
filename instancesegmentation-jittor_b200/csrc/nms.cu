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

#include <mutex>

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

// zero8 / n_zero8: the scan state of this call (removed bits, counters), cleared here instead of by a
// separate memset launch when it is small enough for the grid.
__global__ void nms_keys_kernel(const float* __restrict__ scores, const int* __restrict__ offsets,
                                int B, int n_total, int ties_high_first,
                                uint64_t* __restrict__ keys, int* __restrict__ vals,
                                uint64_t* __restrict__ zero8, int n_zero8) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n_zero8) zero8[p] = 0ull;
  if (p >= n_total) return;
  if (p >= offsets[B]) {
    // n_total may be an upper bound of offsets[B] (a device-side count the host never reads,
    // e.g. the output of rp_rpn_pre_nms): the slack rows sort behind the last problem (and, with
    // the low word all ones, behind every score when only the low 32 bits are sorted: B == 1)
    keys[p] = ((uint64_t)(uint32_t)B << 32) | 0xffffffffull;
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

// RP_NMS_PRESORTED: rank order = input order; also clears the scan state
__global__ void nms_iota_kernel(int n_total, int* __restrict__ vals, uint64_t* __restrict__ zero8,
                                int n_zero8) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n_zero8) zero8[p] = 0ull;
  if (p < n_total) vals[p] = p;
}

struct __align__(16) NmsBox {
  float x1, y1, x2, y2;
};

__global__ void nms_gather_kernel(const float* __restrict__ boxes, const float* __restrict__ labels,
                                  const int* __restrict__ sorted_vals, int n_total,
                                  const int* __restrict__ offsets, int B,
                                  NmsBox* __restrict__ sboxes, float* __restrict__ sarea,
                                  float* __restrict__ slabel) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_total || p >= offsets[B]) return;  // the slack rows belong to no problem
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
// The diagonal block is stored "who can suppress me": row j, word w holds the HIGHER-ranked boxes
// w*64 .. w*64+63 of the band that overlap box j (lower triangle; IoU is symmetric), word-major
// (D[w][j]) so that both the tile writes and the per-box reads of the resolve step are coalesced.
// ------------------------------------------------------------------------------------------
constexpr int NMS_BAND_MAX = 1024;

#ifdef NMS_PROFILE
// cycle stamps of CTA 0, thread 0 in band 3 of the one-launch kernel (tools/run_nms_single.py)
__device__ long long nms_prof[32];
__device__ int nms_prof_on;
#define NMS_STAMP(i)                                                        \
  do {                                                                      \
    if (nms_prof_on && blockIdx.x == 0 && threadIdx.x == 0) nms_prof[i] = clock64(); \
  } while (0)
#else
#define NMS_STAMP(i) \
  do {               \
  } while (0)
#endif

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
  uint64_t* D;            // [B][bw][BR]  diagonal block of the current band (lower triangle)
  NmsBox* kbox;           // [B][BR]      boxes the current band kept
  float* karea;
  float* klabel;
  int* kcount;            // [B] kept in the current band
  int* count;             // [B] kept so far
  int* done;              // [B] max_keep reached
  int wmax, BR, bw;
  float t_cmp, t_lo;
};

// 64 x 64 tiles (row block rb, column block cb <= rb) of the diagonal block of `band`: every group of
// 256 threads of the CTA takes one tile (t_first + group * t_step, `count` tiles in all, count <=
// blockDim / 256); thread = (row r, column quarter q), 16 pair tests, the quarters of a row are merged
// through a shared word.  bw (bw + 1) / 2 tiles per problem.  The whole CTA must call this together.
// Shared arrays: 64 entries per group.
template <bool GE, bool LABELS>
__device__ __forceinline__ void nms_diag_tiles(const NmsBand& p, int seg, int band, int t_first,
                                               int t_step, int count, NmsBox* cbox, float* carea,
                                               float* clabel, unsigned long long* wordS) {
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int base = band * p.BR;
  const int g = threadIdx.x >> 8, tid = threadIdx.x & 255;
  int tile = t_first + g * t_step;
  int rb = 0;
  while (tile >= rb + 1) {
    tile -= rb + 1;
    ++rb;
  }
  const int cb = tile;
  const bool valid = g < count && base < n && rb < p.bw && base + rb * 64 < n;
  cbox += g * 64;
  carea += g * 64;
  clabel += g * 64;
  wordS += g * 64;
  if (valid && tid < 64) {
    const int c = base + cb * 64 + tid;  // c < n: the column block is not after the row block
    if (c < n) {
      cbox[tid] = p.sboxes[off + c];
      carea[tid] = p.sarea[off + c];
      if (LABELS) clabel[tid] = p.slabel[off + c];
    }
    wordS[tid] = 0ull;
  }
  __syncthreads();
  const int r = tid & 63, q = tid >> 6;
  const int row = base + rb * 64 + r;
  if (valid && row < n) {
    const NmsBox a = p.sboxes[off + row];
    const float aarea = p.sarea[off + row];
    float alabel = 0.f;
    if (LABELS) alabel = p.slabel[off + row];
    // columns of strictly higher rank only
    const int cend = min((q + 1) * 16, (cb == rb) ? r : 64);
    uint64_t bits = 0;
    for (int c = q * 16; c < cend; ++c) {
      bool sup = nms_pair<GE>(cbox[c], carea[c], a, aarea, p.t_cmp, p.t_lo);
      if (LABELS) sup = sup && (alabel == clabel[c]);
      if (sup) bits |= (1ull << c);
    }
    if (bits) atomicOr(&wordS[r], (unsigned long long)bits);
  }
  __syncthreads();
  if (valid && tid < 64 && base + rb * 64 + tid < n)
    p.D[((size_t)seg * p.bw + cb) * p.BR + rb * 64 + tid] = wordS[tid];
}

// tiles of the diagonal block of `band` that hold any box of problem seg (tiles are numbered row block
// by row block, so these are the first ones)
__device__ __forceinline__ int nms_tiles_needed(const NmsBand& p, int seg, int band) {
  const int n = p.offsets[seg + 1] - p.offsets[seg];
  const int rows = min(p.BR, n - band * p.BR);
  if (rows <= 0) return 0;
  const int nrb = (rows + 63) >> 6;
  return nrb * (nrb + 1) / 2;
}

template <bool GE, bool LABELS>
__global__ void __launch_bounds__(256)
nms_diag_kernel(NmsBand p, int band) {
  __shared__ NmsBox cbox[64];
  __shared__ float carea[64];
  __shared__ float clabel[64];
  __shared__ unsigned long long wordS[64];
  nms_diag_tiles<GE, LABELS>(p, blockIdx.y, band, blockIdx.x, 1, 1, cbox, carea, clabel, wordS);
}

// Blocks [0, nsup): 64 later boxes each against the boxes band `band` kept, 16 threads per box (the
// kept list of a full band is 1024 boxes: 64 tests per thread, early exit on the first hit).
// Blocks [nsup, ...): the tiles of the diagonal block of band + 1.
constexpr int NMS_SUP_PARTS = 16;
constexpr int NMS_SUP_THREADS = 64 * NMS_SUP_PARTS;

struct NmsSupSmem {
  NmsBox kb[NMS_BAND_MAX];
  float ka[NMS_BAND_MAX];
  float kl[NMS_BAND_MAX];
  unsigned long long s_new;
  unsigned long long wordS[256];
  unsigned bcast;
  // multi-word pass
  unsigned long long mw_word[16], mw_alive[16], mw_new[16];
  int mw_pre[17];
};
constexpr int NMS_MW = 16;  // words per multi-word pass

// Work item x of problem seg in the suppress step of `band` (the CTA returns together).
// The still-alive boxes of the 64-box word are compacted and each gets P = 16, 32 or 64 threads
// (as many as 1024 threads allow), which split the kept list between them.  `cached` names the
// problem whose kept list is already in shared memory (-1: none; a CTA of the one-launch kernel
// works through several words of the same problem).
template <bool GE, bool LABELS>
__device__ __forceinline__ void nms_suppress_body(const NmsBand& p, int seg, int band, int x, int nsup,
                                                  NmsSupSmem& sm, int& cached) {
  const int tid = threadIdx.x;
  if (x >= nsup) {
    nms_diag_tiles<GE, LABELS>(p, seg, band + 1, (x - nsup) * 4, 1, 4, sm.kb, sm.ka, sm.kl, sm.wordS);
    cached = -1;  // the tiles used the kept-list buffers
    return;
  }
  const int wj = (band + 1) * p.bw + x;
  const int done = p.done[seg];
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int kc = p.kcount[seg];
  const uint64_t word = p.removed[(size_t)seg * p.wmax + wj];
  if (done || wj * 64 >= n || kc == 0) return;
  const int nv = min(64, n - wj * 64);
  const uint64_t alive = ~word & (nv == 64 ? ~0ull : ((1ull << nv) - 1ull));
  const int na = __popcll(alive);
  if (na == 0) return;
  if (cached != seg) {
    for (int i = tid; i < kc; i += NMS_SUP_THREADS) {
      sm.kb[i] = p.kbox[(size_t)seg * p.BR + i];
      sm.ka[i] = p.karea[(size_t)seg * p.BR + i];
      if (LABELS) sm.kl[i] = p.klabel[(size_t)seg * p.BR + i];
    }
    cached = seg;
  }
  if (tid == 0) sm.s_new = 0ull;
  __syncthreads();
  const int lp = na <= 16 ? 6 : (na <= 32 ? 5 : 4);  // log2 of the threads per alive box
  const int P = 1 << lp;
  const int slot = tid >> lp, part = tid & (P - 1);
  if (slot < na) {
    const uint32_t lo = (uint32_t)alive, hi = (uint32_t)(alive >> 32);
    const int cl = __popc(lo);
    const int jj = slot < cl ? (int)__fns(lo, 0, slot + 1) : 32 + (int)__fns(hi, 0, slot - cl + 1);
    const int j = wj * 64 + jj;
    const NmsBox q = p.sboxes[off + j];
    const float qarea = p.sarea[off + j];
    float qlabel = 0.f;
    if (LABELS) qlabel = p.slabel[off + j];
    bool hit = false;
    for (int i = part; i < kc && !hit; i += P) {
      bool sup = nms_pair<GE>(sm.kb[i], sm.ka[i], q, qarea, p.t_cmp, p.t_lo);
      if (LABELS) sup = sup && (sm.kl[i] == qlabel);
      hit = sup;
    }
    if (hit) atomicOr(&sm.s_new, 1ull << jj);
  }
  __syncthreads();
  if (tid == 0 && sm.s_new) p.removed[(size_t)seg * p.wmax + wj] = word | sm.s_new;
}

// Up to NMS_MW words (w_first, w_first + w_step, ...) of ONE problem in one pass: the alive boxes of
// all words are compacted together and share the 1024 threads, P = 1024 / alive threads per box (at
// most 64), which split the kept list between them.  One set of round trips per pass instead of per
// word.  Everything that does not depend on the kept list of the current band -- the words, the
// compaction, the candidates' own boxes -- is fetched BEFORE `wait()` (dataflow kernel: the resolver
// may still be running); wait() returns true when the problem needs no further work (CTA-uniform).
// Does not look at done[seg] itself.  Returns wait()'s verdict.
struct NmsGoOn {
  __device__ __forceinline__ bool operator()() const { return false; }
};

template <bool GE, bool LABELS, class Wait>
__device__ __forceinline__ bool nms_suppress_multi(const NmsBand& p, int seg, int w_first, int w_step,
                                                   int nw, NmsSupSmem& sm, int& cached, const Wait& wait) {
  const int tid = threadIdx.x;
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  if (tid < 32) {
    uint64_t word = 0ull, alive = 0ull;
    const int wk = w_first + tid * w_step;
    if (tid < nw && wk * 64 < n) {
      word = p.removed[(size_t)seg * p.wmax + wk];
      const int nv = min(64, n - wk * 64);
      alive = ~word & (nv == 64 ? ~0ull : ((1ull << nv) - 1ull));
    }
    int c = __popcll(alive);
#pragma unroll
    for (int d = 1; d < NMS_MW; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, c, d);
      if (tid >= d) c += t;
    }
    if (tid < NMS_MW) {
      sm.mw_word[tid] = word;
      sm.mw_alive[tid] = alive;
      sm.mw_new[tid] = 0ull;
      sm.mw_pre[tid + 1] = c;
    }
    if (tid == 0) sm.mw_pre[0] = 0;
  }
  __syncthreads();
  const int na = sm.mw_pre[NMS_MW];  // <= 1024: 16 words x 64 boxes
  // Nothing alive: no work, but the caller's "done" signal must still come after the band's resolve
  // (the dataflow counters are running totals: an item of a later step that reports early would let
  // the resolver's wait for an earlier step pass before that step's words are final).
  if (na == 0) return wait();
  int P = NMS_SUP_THREADS / na;
  if (P > 64) P = 64;
  const int slot = tid / P, part = tid - slot * P;
  int k = 0, jj = 0;
  NmsBox q{0.f, 0.f, 0.f, 0.f};
  float qarea = 0.f, qlabel = 0.f;
  if (slot < na) {
#pragma unroll
    for (int d = NMS_MW / 2; d; d >>= 1)
      if (sm.mw_pre[k + d] <= slot) k += d;
    const uint64_t a = sm.mw_alive[k];
    const int ls = slot - sm.mw_pre[k];
    const uint32_t lo = (uint32_t)a, hi = (uint32_t)(a >> 32);
    const int cl = __popc(lo);
    jj = ls < cl ? (int)__fns(lo, 0, ls + 1) : 32 + (int)__fns(hi, 0, ls - cl + 1);
    const int j = (w_first + k * w_step) * 64 + jj;
    q = p.sboxes[off + j];
    qarea = p.sarea[off + j];
    if (LABELS) qlabel = p.slabel[off + j];
  }
  if (wait()) return true;
  const int kc = p.kcount[seg];
  if (cached != seg) {
    // the whole buffer, without waiting for kc (entries past kc are never read)
    for (int i = tid; i < p.BR; i += NMS_SUP_THREADS) {
      sm.kb[i] = p.kbox[(size_t)seg * p.BR + i];
      sm.ka[i] = p.karea[(size_t)seg * p.BR + i];
      if (LABELS) sm.kl[i] = p.klabel[(size_t)seg * p.BR + i];
    }
    cached = seg;
    __syncthreads();
  }
  if (slot < na) {
    bool hit = false;
    for (int i = part; i < kc && !hit; i += P) {
      bool sup = nms_pair<GE>(sm.kb[i], sm.ka[i], q, qarea, p.t_cmp, p.t_lo);
      if (LABELS) sup = sup && (sm.kl[i] == qlabel);
      hit = sup;
    }
    if (hit) atomicOr(&sm.mw_new[k], 1ull << jj);
  }
  __syncthreads();
  if (tid < nw && sm.mw_new[tid])
    p.removed[(size_t)seg * p.wmax + w_first + tid * w_step] = sm.mw_word[tid] | sm.mw_new[tid];
  return false;
}

template <bool GE, bool LABELS>
__global__ void __launch_bounds__(NMS_SUP_THREADS)
nms_suppress_kernel(NmsBand p, int band, int nsup) {
  __shared__ NmsSupSmem sm;
  int cached = -1;
  nms_suppress_body<GE, LABELS>(p, blockIdx.y, band, blockIdx.x, nsup, sm, cached);
}

// Greedy NMS inside band `band`, one CTA per problem, one THREAD per box, as a band-wide fixpoint.
// Thread j holds row j of the diagonal block in registers (the higher-ranked boxes of the band that
// overlap box j).  Two bit sets over the band live in shared memory: kept and undecided.  Per round
//   box j is suppressed  once a KEPT higher-ranked box overlaps it,
//   box j is kept        once no UNDECIDED higher-ranked box overlaps it (and no kept one does);
// the highest-ranked undecided box is always decided, so the loop ends, and the number of rounds is
// the longest alternating chain of overlaps in the band (a handful), not the number of kept boxes.
// The greedy result is unique, so this is the reference's sequential loop (det/layers/nms.py:13-45).
constexpr int NMS_RES_WORDS = NMS_BAND_MAX / 64;
constexpr int NMS_RES_COMPACT = 256;  // undecided boxes at which the rounds switch to a compacted set
struct NmsResSmem {
  __align__(8) uint32_t keptS[2][NMS_BAND_MAX / 32];
  __align__(8) uint32_t undecS[2][NMS_BAND_MAX / 32];
  int wsum[NMS_BAND_MAX / 32];
  uint64_t rows[NMS_RES_WORDS][NMS_RES_COMPACT];  // diagonal rows of the compacted boxes, word-major
  short ids[NMS_RES_COMPACT];
};

// blockDim >= BR, a multiple of 32 (threads past the band idle along)
struct NmsNoWait {
  __device__ __forceinline__ void operator()() const {}
};

// `wait` (called by the whole CTA) returns once the removed bits of this band are final; everything
// before it -- the diagonal rows, the boxes' own records -- is fetched while the suppress step of the
// previous band may still be running (one-launch dataflow kernel)
// Returns (CTA-uniform) whether the keep cap of the problem has been reached.
template <class Wait>
__device__ __forceinline__ bool nms_resolve_body(const NmsBand& p, int seg, int band, int max_keep,
                                                 int ascending, int* __restrict__ keep,
                                                 uint8_t* __restrict__ flags, NmsResSmem& sm,
                                                 const Wait& wait) {
  uint32_t (*keptS)[NMS_BAND_MAX / 32] = sm.keptS;
  uint32_t (*undecS)[NMS_BAND_MAX / 32] = sm.undecS;
  int* wsum = sm.wsum;
  const int tid = threadIdx.x;
  // independent loads first: one L2 round trip, not four
  const int done = p.done[seg];
  const int off = p.offsets[seg];
  const int n = p.offsets[seg + 1] - off;
  const int count0 = p.count[seg];
  const int base = band * p.BR;
  if (done || base >= n) {  // (the dataflow kernel never gets here: its resolver stops first)
    if (tid == 0) p.kcount[seg] = 0;
    return done != 0;
  }
  const int rows = min(p.BR, n - base);
  const int lane = tid & 31, warp = tid >> 5;
  const int cj = tid >> 6;  // the box's 64-rank chunk: only words 0..cj of its row can be set
  const bool live = tid < rows;
  uint64_t T[NMS_RES_WORDS];
  {
    const uint64_t* col = p.D + (size_t)seg * p.bw * p.BR + tid;
#pragma unroll
    for (int w = 0; w < NMS_RES_WORDS; ++w) T[w] = (live && w <= cj) ? col[(size_t)w * p.BR] : 0ull;
  }
  // the box's own record, for the outputs (fetched now, used after the fixpoint)
  int my_e = 0;
  NmsBox my_box{0.f, 0.f, 0.f, 0.f};
  float my_area = 0.f, my_label = 0.f;
  if (live) {
    const int rank = off + base + tid;
    my_e = p.sorted_vals[rank] - off;
    my_box = p.sboxes[rank];
    my_area = p.sarea[rank];
    if (p.slabel) my_label = p.slabel[rank];
  }
  NMS_STAMP(10);
  wait();
  bool my_undec = false;
  if (live) {
    const uint64_t rem = p.removed[(size_t)seg * p.wmax + band * p.bw + cj];
    my_undec = !((rem >> (tid & 63)) & 1ull);
  }
  {
    const unsigned ub = __ballot_sync(0xffffffffu, my_undec);
    if (lane == 0) {
      undecS[0][warp] = ub;
      keptS[0][warp] = 0u;
    }
  }
  __syncthreads();
  NMS_STAMP(11);
  // Rounds over all boxes (thread = box, double-buffered bit sets written by ballots) until at most
  // NMS_RES_COMPACT boxes are undecided; then those boxes (typically a quarter of the band after the
  // first round, scattered over all warps) are moved to the first threads, diagonal rows and all, and
  // the remaining rounds keep only a few warps busy (bit sets updated with shared atomics).
  int cur = 0;
  int left;
  for (int round = 0;; ++round) {
    const unsigned ub = undecS[cur][warp];
    bool nk = false, ns = false;
    if (ub) {
      const uint64_t* k64 = reinterpret_cast<const uint64_t*>(keptS[cur]);
      const uint64_t* u64 = reinterpret_cast<const uint64_t*>(undecS[cur]);
      uint64_t supd = 0ull, potd = 0ull;
      if (round == 0) {  // nothing is kept yet
#pragma unroll
        for (int w = 0; w < NMS_RES_WORDS; ++w)
          if (w <= cj) potd |= T[w] & u64[w];
      } else {
#pragma unroll
        for (int w = 0; w < NMS_RES_WORDS; ++w) {
          if (w <= cj) {  // warp-uniform
            supd |= T[w] & k64[w];
            potd |= T[w] & u64[w];
          }
        }
      }
      if (my_undec) {
        ns = supd != 0ull;
        nk = !ns && potd == 0ull;
      }
    }
    const unsigned bk = __ballot_sync(0xffffffffu, nk);
    const unsigned bs = __ballot_sync(0xffffffffu, ns);
    if (nk || ns) my_undec = false;
    const unsigned nub = ub & ~(bk | bs);
    if (lane == 0) {
      keptS[cur ^ 1][warp] = keptS[cur][warp] | bk;
      undecS[cur ^ 1][warp] = nub;
      wsum[warp] = __popc(nub);
    }
    cur ^= 1;
    left = __syncthreads_count(my_undec);
    if (left <= NMS_RES_COMPACT) break;
  }
  NMS_STAMP(13);
#ifdef NMS_PROFILE
  if (nms_prof_on && blockIdx.x == 0 && threadIdx.x == 0) nms_prof[20] = left;
#endif
  if (left > 0) {
    // compaction: slot = rank of the box among the undecided ones
    {
      const unsigned ub = undecS[cur][warp];
      const int nwarps_ = blockDim.x >> 5;
      const int wc = lane < nwarps_ ? wsum[lane] : 0;
      const int before = (int)__reduce_add_sync(0xffffffffu, (unsigned)(lane < warp ? wc : 0));
      if (my_undec) {
        const int sl = before + __popc(ub & ((1u << lane) - 1u));
        sm.ids[sl] = (short)tid;
#pragma unroll
        for (int w = 0; w < NMS_RES_WORDS; ++w) sm.rows[w][sl] = T[w];
      }
    }
    __syncthreads();
    const bool mine = tid < left;
    int myj = 0;
    unsigned nzw = 0u;  // words of the box's row that have any bit: a handful of the 16
    if (mine) {
      myj = sm.ids[tid];
#pragma unroll
      for (int w = 0; w < NMS_RES_WORDS; ++w) nzw |= (sm.rows[w][tid] != 0ull) ? (1u << w) : 0u;
    }
    my_undec = mine;
    NMS_STAMP(14);
    const uint64_t* k64 = reinterpret_cast<const uint64_t*>(keptS[cur]);
    const uint64_t* u64 = reinterpret_cast<const uint64_t*>(undecS[cur]);
    if (warp * 32 < left) {
      // only the warps that hold compacted boxes take part: named barrier 1 over those threads.
      // Few warps run here, so a round costs its instruction count: rows are read back from shared
      // memory word by word, non-empty words only.
      const unsigned nthr = (unsigned)((left + 31) & ~31);
      for (;;) {
        bool nk = false, ns = false;
        if (my_undec) {
          uint64_t supd = 0ull, potd = 0ull;
          for (unsigned m = nzw; m; m &= m - 1u) {
            const int w = __ffs(m) - 1;
            const uint64_t t = sm.rows[w][tid];
            supd |= t & k64[w];
            potd |= t & u64[w];
          }
          ns = supd != 0ull;
          nk = !ns && potd == 0ull;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory");  // every box has read this round's sets
        if (nk) atomicOr(&keptS[cur][myj >> 5], 1u << (myj & 31));
        if (nk || ns) {
          atomicAnd(&undecS[cur][myj >> 5], ~(1u << (myj & 31)));
          my_undec = false;
        }
        unsigned more;
        asm volatile(
            "{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %1, 0;\n\tbar.red.or.pred q, 1, %2, p;\n\t"
            "selp.u32 %0, 1, 0, q;\n\t}"
            : "=r"(more)
            : "r"((unsigned)my_undec), "r"(nthr)
            : "memory");
        if (!more) break;
      }
    }
    __syncthreads();
  }
  NMS_STAMP(12);
  // positions in rank order, the max_keep cap, outputs
  const unsigned kb = keptS[cur][warp];
  if (lane == 0) wsum[warp] = __popc(kb);
  __syncthreads();
  const int nwarps = blockDim.x >> 5;
  const int wc = lane < nwarps ? wsum[lane] : 0;
  const int before = (int)__reduce_add_sync(0xffffffffu, (unsigned)(lane < warp ? wc : 0));
  int total = (int)__reduce_add_sync(0xffffffffu, (unsigned)wc);
  const int allowed = max_keep > 0 ? max_keep - count0 : 0x7fffffff;
  if (total > allowed) total = allowed;
  if (tid == 0) {
    p.kcount[seg] = total;
    p.count[seg] = count0 + total;
    if (max_keep > 0 && count0 + total >= max_keep) p.done[seg] = 1;
  }
  if ((kb >> lane) & 1u) {
    const int pos = before + __popc(kb & ((1u << lane) - 1u));
    if (pos < allowed) {
      if (ascending)
        flags[off + my_e] = 1;
      else
        keep[off + count0 + pos] = my_e;
      p.kbox[(size_t)seg * p.BR + pos] = my_box;
      p.karea[(size_t)seg * p.BR + pos] = my_area;
      if (p.slabel) p.klabel[(size_t)seg * p.BR + pos] = my_label;
    }
  }
  return max_keep > 0 && count0 + total >= max_keep;
}

__global__ void __launch_bounds__(NMS_BAND_MAX)
nms_resolve_kernel(NmsBand p, int band, int max_keep, int ascending, int* __restrict__ keep,
                   uint8_t* __restrict__ flags) {
  __shared__ NmsResSmem sm;
  nms_resolve_body(p, blockIdx.x, band, max_keep, ascending, keep, flags, sm, NmsNoWait());
}

// ------------------------------------------------------------------------------------------
// All bands in ONE cooperative launch (one CTA per SM, 1024 threads): per band the problems are
// resolved by the first B CTAs, a grid barrier, the suppress / next-diagonal work items are spread
// over the grid, a grid barrier.  Replaces 2 launches per band (each ~3 us of launch + drain on an
// otherwise idle GPU) by two ~1 us barriers; the single-image n = 20000 case has 20 bands.
// The barrier is a monotonic arrival counter in the (zeroed) workspace; the launch is cooperative,
// so all CTAs are resident, and the wait is bounded (trap instead of a hang).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void nms_grid_barrier(unsigned* bar, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    unsigned spins = 0;
    while (*reinterpret_cast<volatile unsigned*>(bar) < target) {
      if (++spins > (1u << 23)) __trap();  // seconds: a resident grid never gets here
    }
    __threadfence();
  }
  __syncthreads();
}

union NmsFusedSmem {
  NmsResSmem res;
  NmsSupSmem sup;
};

// CTA c works for problem c % B (B <= grid) as member c / B of that problem's group, or for the
// problems c, c + grid, ... alone: a CTA keeps one kept list in shared memory per step.
template <bool GE, bool LABELS>
__global__ void __launch_bounds__(NMS_BAND_MAX)
nms_bands_kernel(NmsBand p, int B, int bands, int max_keep, int ascending, int* __restrict__ keep,
                 uint8_t* __restrict__ flags, int* __restrict__ n_keep, unsigned* __restrict__ bar) {
  __shared__ NmsFusedSmem sm;
  const int G = (int)gridDim.x, c = (int)blockIdx.x;
  unsigned phase = 0;
  const int ndiag = p.bw * (p.bw + 1) / 2;
  const bool grouped = B <= G;
  const int seg0 = grouped ? c % B : c;
  const int seg_step = grouped ? B : G;            // (grouped: the loop below runs once)
  const int rnk = grouped ? c / B : 0;             // member of the problem's group
  const int cnt = grouped ? (G - seg0 + B - 1) / B : 1;
  // diagonal block of band 0
  for (int seg = seg0; seg < B; seg += seg_step) {
    const int nt = nms_tiles_needed(p, seg, 0);
    for (int t = rnk; t < nt; t += 4 * cnt) {
      nms_diag_tiles<GE, LABELS>(p, seg, 0, t, cnt, (nt - t + cnt - 1) / cnt, sm.sup.kb, sm.sup.ka,
                                 sm.sup.kl, sm.sup.wordS);
      __syncthreads();
    }
    if (grouped) break;
  }
  nms_grid_barrier(bar, ++phase * G);
  for (int b = 0; b < bands; ++b) {
#ifdef NMS_PROFILE
    if (threadIdx.x == 0 && blockIdx.x == 0) nms_prof_on = (b == 3);
    __syncthreads();
#endif
    NMS_STAMP(0);
    for (int seg = c; seg < B; seg += G) {
      nms_resolve_body(p, seg, b, max_keep, ascending, keep, flags, sm.res, NmsNoWait());
      __syncthreads();
    }
    if (b + 1 == bands) break;
    NMS_STAMP(1);
    nms_grid_barrier(bar, ++phase * G);
    NMS_STAMP(2);
    {
      // all problems capped or out of boxes: nothing left to do (the flags are stable until the
      // next resolve step, so the whole grid takes this exit together)
      int open_problems = 0;
      for (int seg = (int)threadIdx.x; seg < B; seg += (int)blockDim.x) {
        const int n = p.offsets[seg + 1] - p.offsets[seg];
        if (!p.done[seg] && (b + 1) * p.BR < n) open_problems = 1;
      }
      if (!__syncthreads_or(open_problems)) break;
    }
    const int w0 = (b + 1) * p.bw;
    for (int seg = seg0; seg < B; seg += seg_step) {
      const int nt = p.done[seg] ? 0 : nms_tiles_needed(p, seg, b + 1);
      for (int t = rnk; t < nt; t += 4 * cnt) {
        nms_diag_tiles<GE, LABELS>(p, seg, b + 1, t, cnt, (nt - t + cnt - 1) / cnt, sm.sup.kb,
                                   sm.sup.ka, sm.sup.kl, sm.sup.wordS);
        __syncthreads();
      }
      if (!p.done[seg]) {
        int cached = -1;  // the tiles used the kept-list buffers
        const int n_seg = p.offsets[seg + 1] - p.offsets[seg];
        const int wlim = min(p.wmax, (n_seg + 63) >> 6);  // the problem's own words
        for (int w = w0 + rnk; w < wlim; w += cnt * NMS_MW) {
          int nw = (wlim - w + cnt - 1) / cnt;
          if (nw > NMS_MW) nw = NMS_MW;
          (void)nms_suppress_multi<GE, LABELS>(p, seg, w, cnt, nw, sm.sup, cached, NmsGoOn());
          __syncthreads();
        }
      }
      if (grouped) break;
    }
    NMS_STAMP(3);
    nms_grid_barrier(bar, ++phase * G);
    NMS_STAMP(4);
  }
  if (!ascending) {
    __syncthreads();
    for (int seg = c; seg < B; seg += G)
      if (threadIdx.x == 0) n_keep[seg] = p.count[seg];
  }
}

// ------------------------------------------------------------------------------------------
// Few problems, many bands (one image of 20000 proposals = 20 bands): the two grid barriers per band
// of nms_bands_kernel leave 147 CTAs idle while one resolves.  The dataflow kernel gives every
// problem its own resolver CTA and lets the other CTAs (workers) own the removed words and diagonal
// tiles statically (word (problem, w) -> worker (problem * wmax + w) mod NW, so the updates of a word
// are ordered by program order).  Monotonic counters connect them:
//   R[s]   bands of problem s resolved          (resolver -> workers: kept list of band b is ready)
//   Dc[par][s]  diagonal tiles finished, bands of parity par   (workers -> resolver)
//   Uc[s]       "urgent" word items finished: the words of band b+1 in the suppress step of band b
//   Fc[par][s]  word items finished, steps of parity par (the kept-list and diagonal buffers are
//               double buffered; a buffer is reused two bands later, once its step has drained)
// (Dc and Fc are kept per parity because a fast worker may already count items of step b+1 while a
// slow one is still inside step b: a single running total could reach its target early.)
// The resolver of band b+1 waits only for the 16 urgent words of step b, which their owners process
// first; the rest of step b overlaps with the resolve of band b+1.  The launch is cooperative (all
// CTAs resident), every wait is bounded.
// ------------------------------------------------------------------------------------------
struct NmsFlow {
  unsigned *R, *Uc, *Dc[2], *Fc[2];
  uint64_t* D[2];
  NmsBox* kbox[2];
  float *karea[2], *klabel[2];
  int* kcount[2];
};

__device__ __forceinline__ void nms_wait_ge(const unsigned* ctr, unsigned target) {
  if (threadIdx.x == 0) {
    unsigned spins = 0, v;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
      if (v >= target) break;
      if (++spins > (1u << 23)) __trap();  // seconds: a resident grid never gets here
    }
    __threadfence();
  }
  __syncthreads();
}
// the same, handing the value seen to the whole CTA (through `slot`, free again after the next barrier)
__device__ __forceinline__ unsigned nms_wait_ge_value(const unsigned* ctr, unsigned target, unsigned* slot) {
  if (threadIdx.x == 0) {
    unsigned spins = 0, v;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
      if (v >= target) break;
      if (++spins > (1u << 23)) __trap();
    }
    __threadfence();
    *slot = v;
  }
  __syncthreads();
  const unsigned v = *slot;
  __syncthreads();
  return v;
}
// after the CTA's writes: counter a (and b, if given) += the given amounts
__device__ __forceinline__ void nms_signal_add2(unsigned* a, unsigned na, unsigned* b, unsigned nb) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (na) atomicAdd(a, na);
    if (nb) atomicAdd(b, nb);
  }
}
// after the CTA's writes
__device__ __forceinline__ void nms_signal_add(unsigned* ctr) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
  }
}
__device__ __forceinline__ void nms_signal_max(unsigned* ctr, unsigned v) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicMax(ctr, v);
  }
}

// worker side: the kept list of the band is ready; true = the problem has reached its keep cap
struct NmsFlowWaitR {
  const unsigned* r;
  unsigned target, bands;
  unsigned* slot;
  __device__ __forceinline__ bool operator()() const {
    return nms_wait_ge_value(r, target, slot) > bands;
  }
};

struct NmsFlowWait {
  const unsigned *uc, *fc;
  unsigned ut, ft;
  int band;
  __device__ __forceinline__ void operator()() const {
    if (band >= 1) nms_wait_ge(uc, ut);
    if (band >= 2) nms_wait_ge(fc, ft);
  }
};

template <bool GE, bool LABELS>
__global__ void __launch_bounds__(NMS_BAND_MAX)
nms_flow_kernel(NmsBand p, NmsFlow f, int B, int bands, int max_keep, int ascending,
                int* __restrict__ keep, uint8_t* __restrict__ flags, int* __restrict__ n_keep) {
  __shared__ NmsFusedSmem sm;
  const int G = gridDim.x, NW = G - B;
  const int ndiag = p.bw * (p.bw + 1) / 2;
  {  // diagonal block of band 0: all CTAs
    NmsBand q = p;
    q.D = f.D[0];
    for (int it = blockIdx.x; it < B * ndiag; it += G) {
      const int seg = it / ndiag;
      nms_diag_tiles<GE, LABELS>(q, seg, 0, it - seg * ndiag, 1, 1, sm.sup.kb, sm.sup.ka, sm.sup.kl, sm.sup.wordS);
      nms_signal_add(&f.Dc[0][seg]);
    }
    __syncthreads();
  }
  if ((int)blockIdx.x < B) {
    // ---- resolver of problem seg
    const int seg = blockIdx.x;
    const int n = p.offsets[seg + 1] - p.offsets[seg];
    unsigned ut = 0, fsum[2] = {0u, 0u};  // urgent items of steps < b; word items of the steps of each parity
    for (int b = 0; b < bands; ++b) {
      const int par = b & 1;
      NmsBand q = p;
      q.D = f.D[par];
      q.kbox = f.kbox[par];
      q.karea = f.karea[par];
      q.klabel = f.klabel[par];
      q.kcount = f.kcount[par];
#ifdef NMS_PROFILE
      if (threadIdx.x == 0 && blockIdx.x == 0) nms_prof_on = (b == 3);
      __syncthreads();
#endif
      nms_wait_ge(&f.Dc[par][seg], (unsigned)((b / 2 + 1) * ndiag));
      NmsFlowWait wait{&f.Uc[seg], &f.Fc[par][seg], ut, fsum[par], b};
      NMS_STAMP(0);
      const bool capped = nms_resolve_body(q, seg, b, max_keep, ascending, keep, flags, sm.res, wait);
      NMS_STAMP(1);
      const bool last = (b + 1 == bands) || capped || (b + 1) * p.BR >= n;
      // R = bands: all bands resolved; bands + 1: the keep cap was reached, later steps are void
      nms_signal_max(&f.R[seg], (unsigned)(capped ? bands + 1 : (last ? bands : b + 1)));
      NMS_STAMP(5);
      if (last) break;
      const int items = p.wmax - (b + 1) * p.bw;  // word items of step b (> 0: the problem goes on)
      ut += (unsigned)(items < p.bw ? items : p.bw);
      fsum[par] += (unsigned)items;
    }
    __syncthreads();
    if (!ascending && threadIdx.x == 0) n_keep[seg] = p.count[seg];
    return;
  }
  // ---- worker v
  const int v = blockIdx.x - B;
  unsigned capped_mask = 0u;
  const unsigned all_mask = (B >= 32) ? 0xffffffffu : ((1u << B) - 1u);
  for (int b = 0; b + 1 < bands; ++b) {
    {  // tiles of band b+1, into the buffer band b-1 used
      NmsBand q = p;
      q.D = f.D[(b + 1) & 1];
      for (int gt = v; gt < B * ndiag; gt += NW) {
        const int seg = gt / ndiag;
        if ((capped_mask >> seg) & 1u) continue;
        if (b >= 1 && nms_wait_ge_value(&f.R[seg], (unsigned)b, &sm.sup.bcast) > (unsigned)bands) {
          capped_mask |= 1u << seg;
          continue;
        }
        nms_diag_tiles<GE, LABELS>(q, seg, b + 1, gt - seg * ndiag, 1, 1, sm.sup.kb, sm.sup.ka, sm.sup.kl,
                                   sm.sup.wordS);
        nms_signal_add(&f.Dc[(b + 1) & 1][seg]);
      }
      __syncthreads();
    }
    const int par = b & 1;
    NmsBand q = p;
    q.kbox = f.kbox[par];
    q.karea = f.karea[par];
    q.klabel = f.klabel[par];
    q.kcount = f.kcount[par];
    const int w0 = (b + 1) * p.bw;
    const int w1 = w0 + p.bw < p.wmax ? w0 + p.bw : p.wmax;
    int cached = -1;
    for (int sweep = 0; sweep < 2; ++sweep) {  // the words of band b+1 first
      const int lo = sweep ? w1 : w0, hi = sweep ? p.wmax : w1;
      for (int seg = 0; seg < B; ++seg) {
        if ((capped_mask >> seg) & 1u) continue;
        const int r = (int)(((long long)seg * p.wmax + lo) % NW);
        int w = lo + (v - r + NW) % NW;
        while (w < hi) {
          int nw = (hi - w + NW - 1) / NW;
          if (nw > NMS_MW) nw = NMS_MW;
          NmsFlowWaitR waitR{&f.R[seg], (unsigned)(b + 1), (unsigned)bands, &sm.sup.bcast};
          if (nms_suppress_multi<GE, LABELS>(q, seg, w, NW, nw, sm.sup, cached, waitR)) {
            capped_mask |= 1u << seg;  // cap reached: nobody will read these words again
            break;
          }
          nms_signal_add2(&f.Uc[seg], sweep == 0 ? (unsigned)nw : 0u, &f.Fc[par][seg], (unsigned)nw);
          w += nw * NW;
        }
      }
    }
    if (capped_mask == all_mask) return;  // (uniform: every thread sees the same R values)
    __syncthreads();
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

constexpr int NMS_FLOW_MAXB = 16;  // problems the dataflow kernel takes (one resolver CTA each)

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
  unsigned* bar;  // grid-barrier arrival counter of the one-launch path
  unsigned* flow_ctr;  // [6][B] counters of the dataflow kernel
  int* kcount2;
  uint64_t* D2;
  NmsBox* kbox2;
  float *karea2, *klabel2;
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
  w.bar = (unsigned*)take(256);
  w.flow_ctr = (unsigned*)take(nb * 24);
  w.kcount2 = (int*)take(nb * 4);
  w.zero_begin = base ? (char*)base + z0 : nullptr;
  w.zero_bytes = o - z0;
  w.D = (uint64_t*)take(nb * (size_t)w.BR * w.bw * 8);
  w.kbox = (NmsBox*)take(nb * (size_t)w.BR * sizeof(NmsBox));
  w.karea = (float*)take(nb * (size_t)w.BR * 4);
  w.klabel = (float*)take(nb * (size_t)w.BR * 4);
  // second buffer set of the dataflow kernel (few problems, many bands)
  const size_t nf = nb <= NMS_FLOW_MAXB ? nb : 0;
  w.D2 = (uint64_t*)take(nf * (size_t)w.BR * w.bw * 8);
  w.kbox2 = (NmsBox*)take(nf * (size_t)w.BR * sizeof(NmsBox));
  w.karea2 = (float*)take(nf * (size_t)w.BR * 4);
  w.klabel2 = (float*)take(nf * (size_t)w.BR * 4);
  w.total = o;
  return w;
}

// cooperative-launch capacity (CTAs of 1024 threads, at most one per SM) of a one-launch kernel
int nms_coop_capacity(const void* kernel, int slot) {
  static std::mutex mu;
  static int cap[8][64];
  static bool known[8][64];
  int dev = 0;
  (void)cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || slot < 0 || slot >= 8) return 0;
  std::lock_guard<std::mutex> lock(mu);
  if (!known[slot][dev]) {
    int coop = 0, sms = 0, occ = 0;
    (void)cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    (void)cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, NMS_BAND_MAX, 0) != cudaSuccess) {
      (void)cudaGetLastError();
      occ = 0;
    }
    cap[slot][dev] = (coop && occ >= 1) ? sms : 0;  // one CTA per SM: synchronisation is the cost
    known[slot][dev] = true;
  }
  return cap[slot][dev];
}

// *need_finish = 0 when n_keep has been written already
template <bool GE, bool LABELS>
int nms_run_bands(const NmsBand& p, const NmsWorkspace& w, int B, int n_max, int max_keep, int ascending,
                  int* keep, uint8_t* flags, int* n_keep, int* need_finish, cudaStream_t st) {
  const int bands = (n_max + p.BR - 1) / p.BR;
  const int ndiag = p.bw * (p.bw + 1) / 2;  // lower-triangle tiles of a diagonal block
  const int slot = (GE ? 1 : 0) + (LABELS ? 2 : 0);
  *need_finish = 1;
  // dataflow kernel: latency-bound shapes (few words per band step in all); else grid barriers
  const int flow_words = rp_env_int("RP_NMS_FLOW_WORDS", 400);
  if (bands >= 2 && B <= NMS_FLOW_MAXB && (long long)B * p.wmax <= flow_words &&
      rp_env_int("RP_NMS_FUSED", 1) && rp_env_int("RP_NMS_FLOW", 1)) {
    const int cap = nms_coop_capacity((const void*)nms_flow_kernel<GE, LABELS>, 4 + slot);
    if (cap >= B + 64) {
      NmsBand pc = p;
      NmsFlow f;
      f.R = w.flow_ctr;
      f.Uc = w.flow_ctr + B;
      f.Dc[0] = w.flow_ctr + 2 * B;
      f.Dc[1] = w.flow_ctr + 3 * B;
      f.Fc[0] = w.flow_ctr + 4 * B;
      f.Fc[1] = w.flow_ctr + 5 * B;
      f.D[0] = w.D;
      f.D[1] = w.D2;
      f.kbox[0] = w.kbox;
      f.kbox[1] = w.kbox2;
      f.karea[0] = w.karea;
      f.karea[1] = w.karea2;
      f.klabel[0] = w.klabel;
      f.klabel[1] = w.klabel2;
      f.kcount[0] = w.kcount;
      f.kcount[1] = w.kcount2;
      int B_ = B, bands_ = bands, mk = max_keep, asc = ascending;
      void* args[] = {&pc, &f, &B_, &bands_, &mk, &asc, &keep, &flags, &n_keep};
      RP_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)nms_flow_kernel<GE, LABELS>, dim3((unsigned)cap),
                                              dim3(NMS_BAND_MAX), args, 0, st));
      *need_finish = ascending ? 1 : 0;
      return RP_OK;
    }
  }
  const int cap = rp_env_int("RP_NMS_FUSED", 1) ? nms_coop_capacity((const void*)nms_bands_kernel<GE, LABELS>, slot) : 0;
  if (cap > 0) {
    long long items = (long long)B * ndiag;
    if (bands > 1) items = (long long)B * (p.wmax - p.bw + ndiag);
    if (items < B) items = B;
    const int G = (int)(items < cap ? items : cap);
    NmsBand pc = p;
    unsigned* bar = w.bar;
    int B_ = B, bands_ = bands, mk = max_keep, asc = ascending;
    void* args[] = {&pc, &B_, &bands_, &mk, &asc, &keep, &flags, &n_keep, &bar};
    RP_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)nms_bands_kernel<GE, LABELS>, dim3((unsigned)G),
                                            dim3(NMS_BAND_MAX), args, 0, st));
    *need_finish = ascending ? 1 : 0;
    return RP_OK;
  }
  nms_diag_kernel<GE, LABELS><<<dim3((unsigned)ndiag, (unsigned)B), 256, 0, st>>>(p, 0);
  RP_LAUNCH_CHECK();
  for (int b = 0; b < bands; ++b) {
    nms_resolve_kernel<<<B, p.BR, 0, st>>>(p, b, max_keep, ascending, keep, flags);
    RP_LAUNCH_CHECK();
    if (b + 1 < bands) {
      const int nsup = p.wmax - (b + 1) * p.bw;
      nms_suppress_kernel<GE, LABELS>
          <<<dim3((unsigned)(nsup + (ndiag + 3) / 4), (unsigned)B), NMS_SUP_THREADS, 0, st>>>(p, b, nsup);
      RP_LAUNCH_CHECK();
    }
  }
  return RP_OK;
}

}  // namespace

extern "C" {

#ifdef NMS_PROFILE
__attribute__((visibility("default"))) int rp_debug_nms_prof(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, nms_prof, sizeof(long long) * 32);
}
#endif

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
  const bool zero_in_keys = (w.zero_bytes % 8 == 0) && (w.zero_bytes / 8 <= (size_t)nb * T);
  if (flags & RP_NMS_PRESORTED) {
    if (flags & RP_NMS_TIES_HIGH_FIRST) return RP_E_ARG;
    nms_iota_kernel<<<nb, T, 0, st>>>(n_total, w.vals_out, reinterpret_cast<uint64_t*>(w.zero_begin),
                                      zero_in_keys ? (int)(w.zero_bytes / 8) : 0);
    RP_LAUNCH_CHECK();
  } else {
    // (cub::DeviceSegmentedRadixSort, one CTA per problem, was tried for small problems: 6 launches,
    // 150 us for the 1280 problems of the box head -- slower than the whole-batch sort)
    nms_keys_kernel<<<nb, T, 0, st>>>(scores, offsets, B, n_total,
                                      (flags & RP_NMS_TIES_HIGH_FIRST) ? 1 : 0, w.keys_in, w.vals_in,
                                      reinterpret_cast<uint64_t*>(w.zero_begin),
                                      zero_in_keys ? (int)(w.zero_bytes / 8) : 0);
    RP_LAUNCH_CHECK();
    int seg_bits = 0;  // one problem: its rows and the slack rows differ in the low word already
    if (B > 1) {
      seg_bits = 1;
      while ((1 << seg_bits) < B + 1) ++seg_bits;  // B problems + the slack id B
    }
    RP_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.cub_temp, w.cub_bytes, w.keys_in, w.keys_out,
                                                w.vals_in, w.vals_out, n_total, 0, 32 + seg_bits,
                                                st));
  }
  nms_gather_kernel<<<nb, T, 0, st>>>(boxes, labels, w.vals_out, n_total, offsets, B, w.sboxes, w.sarea,
                                      w.slabel);
  RP_LAUNCH_CHECK();
  if (!zero_in_keys) RP_CUDA_TRY(cudaMemsetAsync(w.zero_begin, 0, w.zero_bytes, st));
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
  int rc, need_finish = 1;
  if (labels)
    rc = ge ? nms_run_bands<true, true>(p, w, B, n_max, max_keep, ascending, keep, w.flags, n_keep, &need_finish, st)
            : nms_run_bands<false, true>(p, w, B, n_max, max_keep, ascending, keep, w.flags, n_keep, &need_finish, st);
  else
    rc = ge ? nms_run_bands<true, false>(p, w, B, n_max, max_keep, ascending, keep, w.flags, n_keep, &need_finish, st)
            : nms_run_bands<false, false>(p, w, B, n_max, max_keep, ascending, keep, w.flags, n_keep, &need_finish, st);
  if (rc) return rc;
  if (need_finish)
    nms_finish_kernel<<<B, NMS_SCAN_THREADS, 0, st>>>(offsets, w.count, ascending, w.flags, keep,
                                                      n_keep);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// rp_nms_batched_static: rp_nms_batched for callers that may not allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The workspace is a fixed array in device memory that every GPU
// which loads the library holds until the process exits.  In host code the array's NAME is only a
// host-side stand-in, not a device address: the address is fetched once per device, with that device
// current, by cudaGetSymbolAddress and handed to the kernels as an ordinary pointer.
// All calls share the array: calls that overlap on the GPU (two streams, two host threads) race.
// The array keeps the previous call's contents; rp_nms_batched clears the part that must start at
// zero (removed bits, counters, barrier words) at the start of every call, on the stream.
// ------------------------------------------------------------------------------------------
#define RP_NMS_STATIC_BYTES (4u << 20)
static __device__ __align__(256) unsigned char nms_static_scratch[RP_NMS_STATIC_BYTES];

namespace {
// device address of nms_static_scratch on the current device (nullptr + *rc on failure)
void* nms_static_ptr(int* rc) {
  static std::mutex mu;
  static void* ptr[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    (void)cudaGetLastError();
    *rc = e != cudaSuccess ? (int)e : RP_E_LIMIT;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, nms_static_scratch);
    if (e != cudaSuccess || !p) {
      (void)cudaGetLastError();
      *rc = e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
      return nullptr;
    }
    ptr[dev] = p;
  }
  *rc = RP_OK;
  return ptr[dev];
}
}  // namespace

extern "C" {

size_t rp_nms_static_capacity_bytes(void) { return (size_t)RP_NMS_STATIC_BYTES; }

int rp_nms_batched_static(const float* boxes, const float* scores, const float* labels,
                          const int* offsets, int B, int n_total, int n_max, double thr, int max_keep,
                          int flags, int* keep, int* n_keep, void* stream) {
  if (B < 1 || n_total < 0 || n_max < 0) return RP_E_ARG;
  // the same sizing rule as rp_nms_workspace_bytes; too large for the fixed array: RP_E_LIMIT
  const size_t need = nms_carve(nullptr, n_total, n_max, B).total;
  if (need > (size_t)RP_NMS_STATIC_BYTES) return RP_E_LIMIT;
  int rc = RP_OK;
  void* ws = nms_static_ptr(&rc);
  if (!ws) return rc;
  return rp_nms_batched(boxes, scores, labels, offsets, B, n_total, n_max, thr, max_keep, flags, keep,
                        n_keep, ws, (size_t)RP_NMS_STATIC_BYTES, stream);
}

}  // extern "C"
