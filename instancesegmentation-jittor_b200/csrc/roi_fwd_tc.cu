// roi_fwd_tc.cu -- ROIAlign / Pooler forward on the sm_100a tensor cores ("rx" kernel).
//
// Reference being replaced: det/layers/roi_align.py:116-169 (RoIAlignForward, one thread per output,
// 16 scalar gathers) behind det/modeling/poolers.py:123-156.
//
// The separable form  out[c,ph,pw] = 1/(gh*gw) * sum_y sum_x Ay[ph,y] * Ax[pw,x] * F[c,y,x]
// (roi_tile.h) is evaluated as
//   y-pass (SIMT)    V[(ph,c)][x] = sum_y Ay[ph,y] * F[c,y,x]      4-row bands, one 16-byte strip per thread
//   x-pass (tcgen05) D[(ph,c)][pw] = sum_x V[(ph,c)][x] * Ax'[pw,x]  M=128 (16 bins x 8 channels), N=16
// with the x-pass in 3xTF32 (V_hi*Ax_hi + V_lo*Ax_hi + V_hi*Ax_lo; the tensor core truncates fp32 to
// tf32, measured with tools/umma_probe.cu), fp32 accumulation in TMEM: ~3e-7 relative error.
//
// The RoI's tile is processed in column blocks of 16 ("K quarters", up to 4): a V tile is 128 rows x
// 64 bytes (K-major SWIZZLE_64B, hi and lo planes: 16 KB), the accumulator collects the quarters.
// V rows are ordered (ph, c): the 8 channels x 4 strips a warp stores for one bin are one whole
// 512-byte swizzle atom (conflict-free), and the epilogue's packed [c][ph][pw] stores are conflict-free too.
//
// One persistent CTA per SM, 32 warps with fixed roles, connected by mbarriers:
//   warps 0-1   TMA producers (sub-chunk q belongs to producer q % 2): per sub-chunk (<= 8 channels x R
//               rows x 16 columns) a few cp.async.bulk.tensor boxes; the tensor maps are (x, plane, y) so
//               that a box lands [row][channel][x]: the 16-byte strips of one row of all channels are one
//               contiguous, conflict-free 512-byte run for the y-pass.  Warp 1 also owns the TMEM allocation
//   warps 2-3   planners (even / odd work items): the y band plan (4 merged weights per bin, rotated to
//               the register ring, per row segment) and the dense x-weight matrices Ax' (hi / lo per
//               quarter, K-major SWIZZLE_64B, 1/(gh*gw) folded in)
//   warps 4-7   epilogue: tcgen05.ld (one accumulator row = 16 pw per lane), tcgen05.st of zeros (the next
//               user of the accumulator always accumulates, so K quarters may arrive in any order) ->
//               packed [c][ph][pw] rows in shared memory -> one cp.async.bulk store of 8 channels x PH x PW
//               floats (contiguous in out)
//   warp 8      V-slot release: turns MMA completions into a sequence number the y-pass warps can test
//   warps 9-31  y-pass: unit = (sub-chunk, row segment) = one warp pass of <= 32 strips; ring of 4 rows;
//               every finished bin is split into tf32 hi / lo and stored into the A operand.  The warp
//               whose unit completes a V tile (shared atomic count of the units written) issues that
//               tile's MMAs (one lane; 3 per 8 columns), tcgen05.commit frees the V slot and arrives on the
//               accumulator barrier -- there is no dedicated MMA warp
// The input tile never passes through the LSU on its way in (TMA), the x-pass and the V read-back are
// gone from the LSU / FMA pipes, and the output leaves through the bulk-copy engine.
// RoIs without a band form, wider than 64 columns or taller than the tile slot are evaluated in the
// reference's operand order by the epilogue warps (same CTA, same launch).
#include <cuda.h>

#include <mutex>

#include "common.cuh"
#include "roi_levels.h"
#include "roi_tile.h"

namespace {

constexpr int RX_THREADS = 1024;
constexpr int RX_W_TMA = 0, RX_W_MMA = 1, RX_W_PLAN = 2, RX_W_EPI0 = 4, RX_W_FREE = 8, RX_W_Y0 = 9;
constexpr int RX_NY = 23;              // y-pass warps
constexpr int RX_S = 6;                // tile slots
constexpr int RX_SLOT = 14336;         // bytes per tile slot: 8 channels x 28 rows x 16 columns
constexpr int RX_SV = 6;               // V slots
constexpr int RX_VBYTES = 16384;       // V tile: hi plane (128 rows x 64 B) + lo plane
constexpr int RX_SD = 4;               // accumulator tiles in TMEM (16 columns each)
constexpr int RX_NP = 2;               // plans in flight (one per planner warp)
constexpr int RX_CS = 128;             // channels per work item (RoI x channel slice)
constexpr int RX_RMAX = 64;            // tile rows
constexpr int RX_OSTAGE = 8 * 16 * 16 * 4;  // out stage: 8 channels x <= 16x16 floats (one buffer)
constexpr int RX_NMAP = 18;            // tensor maps per level: 2 widths x 3 row counts x 3 channel counts
constexpr int RX_NPROD = 2;             // TMA producer warps (warps 0, 1): sub-chunk q belongs to producer q % 2
constexpr int RX_PLAN_ARRIVALS = RX_NPROD + RX_NY + 4 + 1;  // TMA warps, y-pass warps, epilogue warps, V release
static_assert(RX_S % RX_NPROD == 0, "each producer owns whole tile slots");

struct RxMeta {
  int kind;  // 0 tensor-core path, 1 reference-order fallback, 2 zero fill
  int level, R, nq, nkq, cs, y0, x0a, plane0, nmt;  // nq: 16-byte strips per tile row, nkq: 16-column quarters
  int H, W, nch;
  float* out;          // &out[(k*C + c_begin) * P]
  const float* featb;  // &feat[level][(b*C + c_begin) * H*W]
  RtGeom g;
  // the bins of a RoI are cut into nseg row segments, each walked by its own warp
  int nseg, lseg, seg_b0[5], seg_r0[4], seg_rows[4];
  RbF4 wrot[16];
  uint32_t mw[4][RX_RMAX / 4 + 2];
};

constexpr int RX_MAXL = 5;  // levels with tensor maps (kernel parameter space: 11.5 KB)
struct RxMaps {
  CUtensorMap m[RX_MAXL][RX_NMAP];
};

struct RxBars {
  uint64_t plan_full[RX_NP], plan_empty[RX_NP];
  uint64_t tile_full[RX_S], tile_empty[RX_S];
  uint64_t v_full[RX_SV], v_empty[RX_SV];
  uint64_t d_full[RX_SD], d_empty[RX_SD];
  // software sequence numbers for the hand-offs that several warps wait on out of order (an mbarrier
  // parity cannot tell phase k from phase k + 2): sub-chunks issued by the TMA warp, V tiles whose
  // MMAs have completed
  uint32_t issued[2], v_freed, d_freed;
  uint32_t vcount[RX_SV];  // channel-segments written into each V slot (32 = complete)
};

// shared-memory layout (offsets from a 1024-byte aligned base)
constexpr int RX_OFF_V = 0;
constexpr int RX_OFF_B = RX_OFF_V + RX_SV * RX_VBYTES;
constexpr int RX_OFF_TILE = RX_OFF_B + RX_NP * 8192;
constexpr int RX_OFF_OST = RX_OFF_TILE + RX_S * RX_SLOT;
constexpr int RX_OFF_META = RX_OFF_OST + RX_OSTAGE;
constexpr int RX_OFF_AY = RX_OFF_META + RX_NP * 1024;
constexpr int RX_OFF_BARS = RX_OFF_AY + RX_NP * 1024;
constexpr int RX_SMEM = RX_OFF_BARS + 512 + 1024;  // + alignment slack
static_assert(sizeof(RxMeta) <= 1024, "RxMeta");
static_assert(sizeof(RbAxis) <= 1024, "RbAxis");
static_assert(sizeof(RxBars) <= 512, "RxBars");
static_assert(RX_SMEM <= 227 * 1024, "shared memory");

// Bounded wait.  On a timeout the warp records (block, code, parity, counter) in the host-mapped
// diagnostic buffer (when one was passed), gives the other warps time to record theirs, and traps:
// a kernel that cannot make progress fails loudly instead of hanging the GPU.
__device__ int* rx_dbg = nullptr;
__device__ __noinline__ void rx_timeout(uint32_t code, uint32_t parity, uint32_t counter) {
  int* d = rx_dbg;
  if (d) {
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
      d[8 * w + 1] = (int)blockIdx.x;
      d[8 * w + 2] = (int)code;
      d[8 * w + 3] = (int)parity;
      d[8 * w + 4] = (int)counter;
      d[8 * w + 0] = 1;
      __threadfence_system();
    }
    for (int i = 0; i < 2000; ++i) __nanosleep(100000);
  }
  __trap();
}
#ifdef RX_PROFILE
#define RX_PROF_T0 const long long rx_t0_ = clock64();
#define RX_PROF_ADD(code) rx_prof_acc[(code)] += clock64() - rx_t0_;
__device__ long long* rx_prof_ptr() {
  __shared__ long long acc[32][12];  // [warp][wait code]; index 0: whole role loop, 10: work, 11: units
  return acc[threadIdx.x >> 5];
}
#define rx_prof_acc (rx_prof_ptr())
#else
#define RX_PROF_T0
#define RX_PROF_ADD(code)
#endif
__device__ __forceinline__ void rx_wait(uint64_t* bar, uint32_t parity, uint32_t code = 0, uint32_t counter = 0) {
  RX_PROF_T0
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(rp_smem_addr(bar)), "r"(parity)
        : "memory");
    if (spin > (1u << 20)) rx_timeout(code, parity, counter);
  }
#ifdef RX_PROFILE
  if ((threadIdx.x & 31) == 0) { RX_PROF_ADD(code) }
#endif
}
__device__ __forceinline__ void rx_publish(uint32_t* ctr, uint32_t v) {
  asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(rp_smem_addr(ctr)), "r"(v) : "memory");
}
// wait until *ctr >= target (sequence numbers: compared modulo 2^32)
__device__ __forceinline__ void rx_wait_ge(uint32_t* ctr, uint32_t target, uint32_t code, uint32_t counter) {
  RX_PROF_T0
  for (uint32_t spin = 0;; ++spin) {
    uint32_t v;
    asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(rp_smem_addr(ctr)) : "memory");
    if ((int32_t)(v - target) >= 0) {
#ifdef RX_PROFILE
      if ((threadIdx.x & 31) == 0) { RX_PROF_ADD(code) }
#endif
      return;
    }
    __nanosleep(spin < 4 ? 64 : 256);  // polling warps must not take the issue slots / the LSU of the working warps
    if (spin > (1u << 22)) rx_timeout(code, target, counter);
  }
}
__device__ __forceinline__ void rx_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rp_smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void rx_arrive_n(uint64_t* bar, uint32_t n) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(rp_smem_addr(bar)), "r"(n) : "memory");
}
__device__ __forceinline__ void rx_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   rp_smem_addr(bar))
               : "memory");
}
// K-major SWIZZLE_64B operand of 16 floats per row (validated by tools/umma_probe.cu): layout type 4,
// stride between 8-row groups 512 bytes; a k-step of 8 floats advances the start address by 32 bytes
__device__ __forceinline__ uint64_t rx_desc_sw64(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)4 << 61);
}
// byte offset of (row, k) in that tile: the 16-byte chunk index is XORed with bits 1-2 of the row
__device__ __forceinline__ uint32_t rx_sw64(int row, int k) {
  return (uint32_t)(row * 64 + ((((k >> 2) ^ ((row >> 1) & 3)) << 4) | ((k & 3) << 2)));
}
// exact floor(j / d) for j * d < 2^32, with m = rx_magic(d), d >= 2
__device__ __forceinline__ uint32_t rx_magic(uint32_t d) { return 0xffffffffu / d + 1u; }
__device__ __forceinline__ float rx_tf32_hi(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }

// ---- planner ---------------------------------------------------------------------------------------
__device__ __forceinline__ void rx_build_plan(RxMeta* m, unsigned char* bmat, RbAxis* ay, int lane,
                                              const RtLevels& lv, int N, int C,
                                              const float* __restrict__ rois,
                                              const int* __restrict__ levels, int k, int c_begin, int PH,
                                              int PW, float sampling_ratio, float* __restrict__ out) {
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  const int nch = min(RX_CS, C - c_begin);
  int kind = 0;
  if (level < 0 || level >= lv.L || b < 0 || b >= N) kind = 2;
  const int lvl = kind == 2 ? 0 : level;
  const int H = lv.H[lvl], W = lv.W[lvl];
  RtGeom g = rt_geom_fwd(roi, lv.scale[lvl], sampling_ratio, PH, PW);
  if (kind == 0 && !(rb_axis_candidate(g.bh, g.gh, PH, H) && g.gw >= 1 && g.gw <= RB_MAXG && W >= 4)) kind = 1;

  // y axis: band plan (one bin per lane)
  int oky = 1;
  if (kind == 0 && lane < PH) oky = rb_build_bin(ay, lane, g.sh, g.bh, g.gh, H);
  if (!__all_sync(0xffffffffu, oky)) kind = kind == 0 ? 1 : kind;
  __syncwarp();
  int R = 4;
  if (kind == 0) {
    int fin = 1;
    if (lane == 0) fin = rb_finish_axis(ay, PH);
    fin = __shfl_sync(0xffffffffu, fin, 0);
    __syncwarp();
    R = ay->ext;
    if (!fin || R > RX_RMAX) kind = 1;
  }
  // x axis: extent of the valid taps
  int xmin = 0x7fffffff, xmax = -1;
  if (kind == 0 && lane < PW) {
    for (int i = 0; i < g.gw; ++i) {
      int lo, hi;
      float wlo, whi;
      if (rt_axis_tap(g.sw, g.bw, lane, i, g.gw, W, &lo, &hi, &wlo, &whi)) {
        xmin = min(xmin, lo);
        xmax = max(xmax, hi);
      }
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    xmin = min(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
    xmax = max(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
  }
  int x0a = 0, Kp = 8;
  if (kind == 0) {
    if (xmax < 0) {
      kind = 2;  // no x sample inside the map: the output is zero
    } else {
      x0a = xmin & ~3;
      Kp = (xmax - x0a + 1 + 7) & ~7;
      if (Kp > 64) kind = 1;
    }
  }
  int cs = 8;
  if (kind == 0) {
    // channels per sub-chunk (one 16-column quarter of cs channels must fit a tile slot)
    const int Re = (R + 1) & ~1;
    if (Re * 16 * 8 * 4 <= RX_SLOT)
      cs = 8;
    else if (Re * 16 * 4 * 4 <= RX_SLOT)
      cs = 4;
    else if (Re * 16 * 2 * 4 <= RX_SLOT)
      cs = 2;
    else
      kind = 1;
  }
  if (kind == 0) {
    const int nkq = (Kp + 15) >> 4;
    // y: segments, weights rotated to the ring of the bin's segment (slot = segment row & 3), bins
    // completed per segment row
    const int nseg = PH >= 8 ? 4 : (PH >= 4 ? 2 : 1);
    // segment sg covers bins [sg*PH/nseg, (sg+1)*PH/nseg) and the tile rows its bands touch
    auto seg_of = [&](int bin) {
      int sg = 0;
      while (sg + 1 < nseg && bin >= ((sg + 1) * PH) / nseg) ++sg;
      return sg;
    };
    if (lane < PH) {
      const int sg = seg_of(lane);
      const int r0 = ay->start[(sg * PH) / nseg];
      const int rot = (ay->start[lane] - r0) & 3;
      const RbF4 w = ay->w[lane];
      RbF4 r;
      r.x = rot == 0 ? w.x : (rot == 1 ? w.w : (rot == 2 ? w.z : w.y));
      r.y = rot == 0 ? w.y : (rot == 1 ? w.x : (rot == 2 ? w.w : w.z));
      r.z = rot == 0 ? w.z : (rot == 1 ? w.y : (rot == 2 ? w.x : w.w));
      r.w = rot == 0 ? w.w : (rot == 1 ? w.z : (rot == 2 ? w.y : w.x));
      m->wrot[lane] = r;
    }
    if (lane < RX_RMAX / 4 + 2) {
      for (int sg = 0; sg < nseg; ++sg) {
        const int bb0 = (sg * PH) / nseg, bb1 = ((sg + 1) * PH) / nseg, r0 = ay->start[bb0];
        uint32_t word = 0;
        for (int bn = bb0; bn < bb1; ++bn) {
          const int r = ay->start[bn] - r0 + 3;
          if ((r >> 2) == lane) word += 1u << (8 * (r & 3));
        }
        m->mw[sg][lane] = word;
      }
    }
    if (lane == 0) {
      m->nseg = nseg;
      m->lseg = nseg == 4 ? 2 : (nseg == 2 ? 1 : 0);
      for (int sg = 0; sg < nseg; ++sg) {
        const int bb0 = (sg * PH) / nseg, bb1 = ((sg + 1) * PH) / nseg;
        m->seg_b0[sg] = bb0;
        m->seg_r0[sg] = ay->start[bb0];
        m->seg_rows[sg] = ay->start[bb1 - 1] + 4 - ay->start[bb0];
      }
      m->seg_b0[nseg] = PH;
    }
    // x: dense weights per quarter [16 pw rows][16 columns], hi plane at +0 and lo plane at +1024 of
    // the quarter's 2 KB; accumulated in full precision in the hi planes, then split
    float4* bz = reinterpret_cast<float4*>(bmat);
    for (int i = lane; i < nkq * 128; i += 32) bz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    if (lane < PW) {
      for (int i = 0; i < g.gw; ++i) {
        int lo, hi;
        float wlo, whi;
        if (rt_axis_tap(g.sw, g.bw, lane, i, g.gw, W, &lo, &hi, &wlo, &whi)) {
          const int c0 = lo - x0a, c1 = hi - x0a;
          float* p0 = reinterpret_cast<float*>(bmat + (c0 >> 4) * 2048 + rx_sw64(lane, c0 & 15));
          *p0 += wlo;
          float* p1 = reinterpret_cast<float*>(bmat + (c1 >> 4) * 2048 + rx_sw64(lane, c1 & 15));
          *p1 += whi;
        }
      }
    }
    __syncwarp();
    const float inv = 1.0f / (float)(g.gh * g.gw);
    for (int i = lane; i < nkq * 256; i += 32) {  // 16 rows x 16 floats per quarter
      const int kq = i >> 8, e = i & 255;
      float* ph_ = reinterpret_cast<float*>(bmat + kq * 2048) + e;
      const float v = *ph_ * inv;
      const float hi = rx_tf32_hi(v);
      *ph_ = hi;
      ph_[256] = v - hi;  // lo plane: +1024 bytes
    }
    if (lane == 0) {
      m->nkq = nkq;
      m->nq = Kp >> 2;
      m->cs = cs;
      m->R = R;
      m->y0 = ay->lo0;
      m->x0a = x0a;
    }
  }
  if (lane == 0) {
    m->kind = kind;
    m->level = lvl;
    m->plane0 = b * C + c_begin;
    m->nmt = nch >> 3;
    m->nch = nch;
    m->H = H;
    m->W = W;
    m->out = out + ((size_t)k * C + c_begin) * P;
    m->featb = kind == 2 ? nullptr : lv.feat[lvl] + ((size_t)b * C + c_begin) * H * W;
    m->g = g;
  }
  rp_fence_proxy_async_smem();
  __syncwarp();
}

// ---- y-pass of one unit (sub-chunk x row segment) by one warp ---------------------------------------
// tile: [rows][cs channels][nqh strips] RbF4 (nqh <= 4, so cs * nqh <= 32: one item per lane);
// V row of (ph, channel c) is ph * 8 + c: 64 bytes, 16-byte chunk s at (s ^ ((row >> 1) & 3)).
__device__ __forceinline__ void rx_ypass(const RbF4* __restrict__ tile, int R, int cs, int nqh,
                                         const uint32_t* __restrict__ mw, const RbF4* __restrict__ wrot,
                                         int ph0, unsigned char* vhi, int cbase, int lane) {
  const int nitems = cs * nqh;
  if (lane >= nitems) return;
  const int c = nqh == 4 ? lane >> 2 : lane >> 1;
  const int s = lane - c * nqh;
  const int ch = cbase + c;
  const RbF4* rows = tile + lane;
  // byte offset of row (ph, ch), chunk s:  (ph*8 + ch)*64 + ((s ^ ((ph*8 + ch) >> 1 & 3)) << 4); ph*8 is a
  // multiple of 8, so the XOR term depends on ch only
  unsigned char* vdst = vhi + ch * 64 + ((s ^ ((ch >> 1) & 3)) << 4) + ph0 * 512;
  const RbF4* wr = wrot + ph0;
  RbF4 W0, W1, W2, W3;
  W0.x = W0.y = W0.z = W0.w = 0.f;
  W1 = W0;
  W2 = W0;
  W3 = W0;
#define RX_BINS(cnt)                                                                             \
  for (uint32_t m_ = (cnt); m_ != 0; --m_) {                                                    \
    const RbF4 w_ = *wr++;                                                                      \
    float4 a_;                                                                                  \
    a_.x = fmaf(w_.w, W3.x, fmaf(w_.z, W2.x, fmaf(w_.y, W1.x, w_.x * W0.x)));                   \
    a_.y = fmaf(w_.w, W3.y, fmaf(w_.z, W2.y, fmaf(w_.y, W1.y, w_.x * W0.y)));                   \
    a_.z = fmaf(w_.w, W3.z, fmaf(w_.z, W2.z, fmaf(w_.y, W1.z, w_.x * W0.z)));                   \
    a_.w = fmaf(w_.w, W3.w, fmaf(w_.z, W2.w, fmaf(w_.y, W1.w, w_.x * W0.w)));                   \
    float4 h_;                                                                                  \
    h_.x = rx_tf32_hi(a_.x);                                                                    \
    h_.y = rx_tf32_hi(a_.y);                                                                    \
    h_.z = rx_tf32_hi(a_.z);                                                                    \
    h_.w = rx_tf32_hi(a_.w);                                                                    \
    *reinterpret_cast<float4*>(vdst) = h_;                                                      \
    *reinterpret_cast<float4*>(vdst + 8192) =                                                   \
        make_float4(a_.x - h_.x, a_.y - h_.y, a_.z - h_.z, a_.w - h_.w);                        \
    vdst += 512;                                                                                \
  }
  for (int r = 0; r < R; r += 4) {
    const uint32_t mword = mw[r >> 2];
    W0 = rows[0];
    RX_BINS(mword & 0xffu)
    if (r + 1 < R) W1 = rows[nitems];
    RX_BINS((mword >> 8) & 0xffu)
    if (r + 2 < R) W2 = rows[2 * nitems];
    RX_BINS((mword >> 16) & 0xffu)
    if (r + 3 < R) W3 = rows[3 * nitems];
    RX_BINS(mword >> 24)
    rows += 4 * nitems;
  }
#undef RX_BINS
}

__global__ void __launch_bounds__(RX_THREADS, 1)
rx_pool_fwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   float* __restrict__ out, const __grid_constant__ RxMaps maps, int n_items, int ablate) {
  extern __shared__ unsigned char rx_smem_raw[];
  unsigned char* smem = rx_smem_raw + ((1024u - (rp_smem_addr(rx_smem_raw) & 1023u)) & 1023u);
  RxBars* bars = reinterpret_cast<RxBars*>(smem + RX_OFF_BARS);
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int P = PH * PW;

  if (tid == 0) {
    for (int i = 0; i < RX_NP; ++i) {
      rp_mbar_init(&bars->plan_full[i], 1);
      rp_mbar_init(&bars->plan_empty[i], RX_PLAN_ARRIVALS);
    }
    for (int i = 0; i < RX_S; ++i) {
      rp_mbar_init(&bars->tile_full[i], 1);
      rp_mbar_init(&bars->tile_empty[i], 4);
    }
    for (int i = 0; i < RX_SV; ++i) {
      rp_mbar_init(&bars->v_full[i], 1);  // unused since the last y-pass warp of a tile issues its MMAs
      rp_mbar_init(&bars->v_empty[i], 1);
      bars->vcount[i] = 0;
    }
    for (int i = 0; i < RX_SD; ++i) {
      rp_mbar_init(&bars->d_full[i], 4);  // one commit per K quarter (+ 4 - nkq plain arrivals)
      rp_mbar_init(&bars->d_empty[i], 4);
    }
    bars->issued[0] = bars->issued[1] = 0;
    bars->v_freed = 0;
    bars->d_freed = 0;
  }
  if (warp == RX_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(
                     rp_smem_addr(&tmem_base_s))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
#ifdef RX_PROFILE
  if (lane < 12) rx_prof_acc[lane] = 0;
  __syncwarp();
  const long long rx_role_t0 = clock64();
#endif

#define RX_ITEM_LOOP for (int it = blockIdx.x, p = 0; it < n_items; it += gridDim.x, ++p)
#define RX_META()                                                     \
  const int ps = p % RX_NP;                                           \
  rx_wait(&bars->plan_full[ps], (uint32_t)(p / RX_NP) & 1u, 1, p);    \
  const RxMeta* M = reinterpret_cast<const RxMeta*>(smem + RX_OFF_META + ps * 1024);

  if (warp == RX_W_PLAN || warp == RX_W_PLAN + 1) {
    // ===== planners: warp RX_W_PLAN builds the even work items (plan slot 0), the next warp the odd ones
    const int pw_ = warp - RX_W_PLAN;
    RbAxis* ay = reinterpret_cast<RbAxis*>(smem + RX_OFF_AY + pw_ * 1024);
    RX_ITEM_LOOP {
      const int ps = p % RX_NP;
      if (ps != pw_) continue;
      rx_wait(&bars->plan_empty[ps], ((uint32_t)(p / RX_NP) & 1u) ^ 1u, 2, p);
      const int sl = it / K, k = it - sl * K;
#ifdef RX_PROFILE
      const long long pl_t0 = clock64();
#endif
      rx_build_plan(reinterpret_cast<RxMeta*>(smem + RX_OFF_META + ps * 1024), smem + RX_OFF_B + ps * 8192,
                    ay, lane, lv, N, C, rois, levels, k, sl * RX_CS, PH, PW, sampling_ratio, out);
#ifdef RX_PROFILE
      if (lane == 0) {
        rx_prof_acc[10] += clock64() - pl_t0;
        rx_prof_acc[11] += 1;
      }
#endif
      if (lane == 0) rx_arrive(&bars->plan_full[ps]);
    }
  } else if (warp < RX_NPROD) {
    // ===== TMA producers: warp w issues the sub-chunks q == w (mod RX_NPROD) into the slots it owns =====
    uint32_t q = 0;
    RX_ITEM_LOOP {
      RX_META()
      if (M->kind == 0) {
        const int R = M->R, cs = M->cs, nkq = M->nkq, nq = M->nq, nsc = 8 / cs;
        const int Re = (R + 1) & ~1, n8 = Re >> 3, rem = Re & 7;
        const int nbox = n8 + ((rem & 4) ? 1 : 0) + ((rem & 2) ? 1 : 0);
        // this lane's box: row offset and row-count index (0: 8 rows, 1: 4, 2: 2)
        int brow = lane * 8, bri = 0;
        if (lane >= n8) {
          brow = n8 * 8;
          if ((rem & 4) && lane == n8) {
            bri = 1;
          } else {
            bri = 2;
            brow += rem & 4;
          }
        }
        const CUtensorMap* lmaps = maps.m[M->level];
        const int ci = cs == 8 ? 0 : (cs == 4 ? 1 : 2);
        for (int mt = 0; mt < M->nmt; ++mt)
          for (int kq = 0; kq < nkq; ++kq) {
            const int nqh = min(4, nq - 4 * kq);  // 2 or 4 strips: box width 8 or 16 floats
            const CUtensorMap* tm = lmaps + (((nqh >> 1) - 1) * 3 + bri) * 3 + ci;
            const uint32_t row_bytes = (uint32_t)(cs * nqh * 16);
            for (int sc = 0; sc < nsc; ++sc, ++q) {
              if ((int)(q % RX_NPROD) != warp) continue;
              const int ts = q % RX_S;
              rx_wait(&bars->tile_empty[ts], ((q / RX_S) & 1u) ^ 1u, 3, q);
              if (lane == 0) rp_mbar_arrive_expect_tx(&bars->tile_full[ts], (uint32_t)Re * row_bytes);
              __syncwarp();
              if (lane < nbox)
                rp_tma_load_3d(smem + RX_OFF_TILE + ts * RX_SLOT + (uint32_t)brow * row_bytes, tm,
                               M->x0a + 16 * kq, M->plane0 + mt * 8 + sc * cs, M->y0 + brow,
                               &bars->tile_full[ts]);
              __syncwarp();
              if (lane == 0) rx_publish(&bars->issued[q % RX_NPROD], q / RX_NPROD + 1);
            }
          }
      }
      __syncwarp();
      if (lane == 0) rx_arrive(&bars->plan_empty[ps]);
    }
  } else if (warp == RX_W_FREE) {
    // ===== V-slot release: the only waiter of v_empty, in order; publishes the count of finished V tiles
    uint32_t v = 0;
    RX_ITEM_LOOP {
      RX_META()
      if (M->kind == 0) {
        const int nv = M->nmt * M->nkq;
        for (int i = 0; i < nv; ++i, ++v) {
          rx_wait(&bars->v_empty[v % RX_SV], (v / RX_SV) & 1u, 7, v);
          if (lane == 0) rx_publish(&bars->v_freed, v + 1u);
        }
      }
      __syncwarp();
      if (lane == 0) rx_arrive(&bars->plan_empty[ps]);
    }
  } else if (warp >= RX_W_EPI0 && warp < RX_W_EPI0 + 4) {
    // ===== epilogue =====
    const int quad = warp & 3, etid = tid - RX_W_EPI0 * 32;
    uint64_t evict_first;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(evict_first));
    uint32_t d = 0;
    // accumulators start at zero and are zeroed again after every drain: the K quarters of a tile are
    // issued by whichever y-pass warp finishes each quarter last, in any order, always accumulating
    for (int i = 0; i < RX_SD; ++i) {
      const uint32_t za = tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)(i * 16);
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(za),
          "r"(0u)
          : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (etid == 0) rx_publish(&bars->d_freed, (uint32_t)RX_SD);  // tiles 0 .. SD-1 may be issued
    RX_ITEM_LOOP {
      RX_META()
      if (M->kind == 0) {
        float* outp = M->out;
        for (int mt = 0; mt < M->nmt; ++mt, ++d) {
          const int ds = d % RX_SD;
          rx_wait(&bars->d_full[ds], (d / RX_SD) & 1u, 6, d);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          uint32_t r[16];
          const uint32_t taddr = tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)(ds * 16);
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
              "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
                "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
                "=r"(r[14]), "=r"(r[15])
              : "r"(taddr)
              : "memory");
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          asm volatile(
              "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(
                  taddr),
              "r"(0u)
              : "memory");
          asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          // the stage buffer was last read by the bulk store of the previous tile
          if (etid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (etid == 0) rx_publish(&bars->d_freed, d + 1u + (uint32_t)RX_SD);  // slot ds is drained and zero
          unsigned char* stage = smem + RX_OFF_OST;
          const int row = quad * 32 + lane, ph = row >> 3, c = row & 7;  // V / accumulator rows are (ph, c)
          if (ph < PH) {
            float* dst = reinterpret_cast<float*>(stage) + c * P + ph * PW;
            if ((PW & 1) == 0) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (2 * j < PW)
                  *reinterpret_cast<float2*>(dst + 2 * j) =
                      make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < PW) dst[j] = __uint_as_float(r[j]);
            }
          }
          rp_fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (etid == 0) {
            // the output is written once and never read here: evict-first keeps the feature planes in L2
            asm volatile(
                "cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(
                    outp + (size_t)mt * 8 * P),
                "r"(rp_smem_addr(stage)), "r"((uint32_t)(8 * P * 4)), "l"(evict_first)
                : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else if (M->kind == 1) {
        const int total = M->nch * P, H = M->H, W = M->W;
        const size_t HW = (size_t)H * W;
        const RtGeom g = M->g;
        for (int i = etid; i < total; i += 128) {
          const int c = i / P, rem = i - c * P;
          const int ph = rem / PW, pw = rem - ph * PW;
          M->out[i] = rt_direct_output(M->featb + (size_t)c * HW, H, W, g, ph, pw);
        }
      } else {
        const int total = M->nch * P;
        for (int i = etid; i < total; i += 128) M->out[i] = 0.f;
      }
      __syncwarp();
      if (lane == 0) rx_arrive(&bars->plan_empty[ps]);
    }
    if (etid == 0) rp_bulk_wait_all();
  } else if (warp >= RX_W_Y0) {
    // ===== y-pass =====
    const int yw = warp - RX_W_Y0;
    uint32_t q0 = 0, v0 = 0, u0 = 0, d0 = 0;  // sub-chunks, V tiles, work units, accumulator tiles before this item
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
    RX_ITEM_LOOP {
      RX_META()
      if (M->kind == 0) {
        // unit j of the item = ((tile * nsc + sc) * nseg + sg), tile = mt * nkq + kq; nsc and nseg are
        // powers of two: this warp's units (u0 + j == yw mod NY) are decoded, not walked
        const int cs = M->cs, nkq = M->nkq, nq = M->nq, nseg = M->nseg;
        const int lseg = M->lseg, lsc = cs == 8 ? 0 : (cs == 4 ? 1 : 2);
        const int ntiles = M->nmt * nkq;
        const uint32_t b_s = rp_smem_addr(smem + RX_OFF_B + ps * 8192);
        const int nunits = ntiles << (lsc + lseg);
        for (int j = (int)((uint32_t)(yw + RX_NY) - u0 % RX_NY) % RX_NY; j < nunits; j += RX_NY) {
          const int sg = j & (nseg - 1);
          const int qi = j >> lseg, sc = qi & ((1 << lsc) - 1);
          const int vi = qi >> lsc, kq = vi % nkq;
          const uint32_t q = q0 + (uint32_t)qi, v = v0 + (uint32_t)vi;
          const int vs = v % RX_SV, ts = q % RX_S;
          const int nqh = min(4, nq - 4 * kq);
          rx_wait_ge(&bars->v_freed, v + 1u - (uint32_t)RX_SV, 7, v);  // the MMAs of tile v - SV are done
          rx_wait_ge(&bars->issued[q % RX_NPROD], q / RX_NPROD + 1u, 9, q);  // the slot's barrier is in the phase of sub-chunk q
          rx_wait(&bars->tile_full[ts], (q / RX_S) & 1u, 8, q);
#ifdef RX_PROFILE
          const long long yp_t0 = clock64();
#endif
          if (!(ablate & 2))  // timing ablation (RP_RX_ABLATE bit 1): no y-pass at all, wrong results
          rx_ypass(reinterpret_cast<const RbF4*>(smem + RX_OFF_TILE + ts * RX_SLOT) + M->seg_r0[sg] * cs * nqh,
                   M->seg_rows[sg], cs, nqh, M->mw[sg], M->wrot, M->seg_b0[sg], smem + RX_OFF_V + vs * RX_VBYTES,
                   sc * cs, lane);
          rp_fence_proxy_async_smem();
          __syncwarp();
#ifdef RX_PROFILE
          if (lane == 0) {
            rx_prof_acc[10] += clock64() - yp_t0;
            rx_prof_acc[11] += 1;
          }
#endif
          // hand the tile slot back; count this unit's channel-segments into the V slot (32 = the tile
          // is complete): the warp that completes it issues the tile's MMAs itself
          uint32_t done = 0;
          if (lane == 0) {
            rx_arrive_n(&bars->tile_empty[ts], (uint32_t)(4 >> lseg));
            const uint32_t add = (uint32_t)(cs * (4 >> lseg));
            uint32_t old;
            asm volatile("atom.add.acq_rel.cta.shared::cta.u32 %0, [%1], %2;"
                         : "=r"(old)
                         : "r"(rp_smem_addr(&bars->vcount[vs])), "r"(add)
                         : "memory");
            done = old + add;
          }
          done = __shfl_sync(0xffffffffu, done, 0);
          if (done == 32u) {
            const int mt = vi / nkq;
            const uint32_t d = d0 + (uint32_t)mt;
            const int ds = d % RX_SD;
            rx_wait_ge(&bars->d_freed, d + 1u, 4, d);  // accumulator slot drained and zeroed
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (lane == 0) {
              bars->vcount[vs] = 0;
              const int nks = nqh >> 1;
              const uint32_t a_s = rp_smem_addr(smem + RX_OFF_V + vs * RX_VBYTES);
              const uint64_t dah = rx_desc_sw64(a_s), dal = rx_desc_sw64(a_s + 8192);
              const uint64_t dbh = rx_desc_sw64(b_s + kq * 2048), dbl = rx_desc_sw64(b_s + kq * 2048 + 1024);
              const uint32_t dtm = tmem + (uint32_t)(ds * 16);
              for (int ks = 0; ks < nks; ++ks) {
                const uint64_t adv = (uint64_t)(ks * 2);
#define RX_MMA(DA, DB)                                                                        \
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"                                   \
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(dtm),   \
               "l"((DA) + adv), "l"((DB) + adv), "r"(idesc), "r"(1u)                         \
               : "memory");
                RX_MMA(dah, dbh)
                if (!(ablate & 1)) {  // timing ablation (RP_RX_ABLATE bit 0): hi*hi only, wrong results
                  RX_MMA(dal, dbh)
                  RX_MMA(dah, dbl)
                }
#undef RX_MMA
              }
              rx_commit(&bars->v_empty[vs]);
              rx_commit(&bars->d_full[ds]);  // 4 arrivals complete the accumulator: one commit per quarter ...
              if (kq == 0 && nkq < 4) rx_arrive_n(&bars->d_full[ds], (uint32_t)(4 - nkq));  // ... and the rest here
            }
            __syncwarp();
          }
        }
        d0 += (uint32_t)M->nmt;
        u0 += (uint32_t)nunits;
        q0 += (uint32_t)(nunits >> lseg);
        v0 += (uint32_t)ntiles;
      }
      __syncwarp();
      if (lane == 0) rx_arrive(&bars->plan_empty[ps]);
    }
  }
#undef RX_ITEM_LOOP
#undef RX_META
#ifdef RX_PROFILE
  if (lane == 0) rx_prof_acc[0] = clock64() - rx_role_t0;
  __syncwarp();
  if (blockIdx.x == 0 && rx_dbg && lane < 12)
    reinterpret_cast<long long*>(rx_dbg + 256)[warp * 12 + lane] = rx_prof_acc[lane];
#endif
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == RX_W_MMA)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

// Rows of `W` floats -> rows of `pitch` floats (pitch % 4 == 0, zero padded): a level whose rows are not
// 16-byte multiples (P5 of an 800 x 1344 batch is 25 x 42) cannot be a TMA tensor as it lies; it is
// re-pitched into a stream-ordered scratch buffer first (a few MB; the big levels never need it).
__global__ void rx_repitch_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t rows, int W,
                                  int pitch) {
  const size_t total = rows * (size_t)pitch;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / (size_t)pitch;
    const int x = (int)(i - r * (size_t)pitch);
    dst[i] = x < W ? src[r * (size_t)W + x] : 0.f;
  }
}

// ---- host: tensor maps (x, plane, y) per level, kept in device memory ------------------------------
typedef CUresult (*rx_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

rx_encode_fn rx_encoder() {
  static rx_encode_fn fn = []() -> rx_encode_fn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    return reinterpret_cast<rx_encode_fn>(p);
  }();
  return fn;
}

struct RxMapEntry {
  const float* ptr;
  long long planes;
  int H, W, pitch, ok;
  unsigned long long stamp;
  CUtensorMap m[RX_NMAP];
};

// Fills maps->m[l] for level l; a small host cache keeps repeated calls on the same tensors free of
// encoding (18 cuTensorMapEncodeTiled calls per level otherwise).  Returns 0 on failure.
int rx_level_maps(RxMaps* maps, int l, const float* ptr, long long planes, int H, int W, int pitch) {
  static std::mutex mu;
  static RxMapEntry cache[32];
  static unsigned long long clock_ = 0;
  rx_encode_fn enc = rx_encoder();
  if (!enc) return 0;
  std::lock_guard<std::mutex> lock(mu);
  int victim = 0;
  for (int i = 0; i < 32; ++i) {
    RxMapEntry& e = cache[i];
    if (e.ok && e.ptr == ptr && e.planes == planes && e.H == H && e.W == W && e.pitch == pitch) {
      e.stamp = ++clock_;
      memcpy(maps->m[l], e.m, sizeof(e.m));
      return 1;
    }
    if (e.stamp < cache[victim].stamp) victim = i;
  }
  RxMapEntry& e = cache[victim];
  e.ok = 0;
  static const int rows_of[3] = {8, 4, 2};
  const cuuint64_t gdim[3] = {(cuuint64_t)W, (cuuint64_t)planes, (cuuint64_t)H};
  const cuuint64_t gstride[2] = {(cuuint64_t)H * pitch * 4, (cuuint64_t)pitch * 4};
  const cuuint32_t estride[3] = {1, 1, 1};
  for (int wi = 0; wi < 2; ++wi)
    for (int ri = 0; ri < 3; ++ri)
      for (int ci = 0; ci < 3; ++ci) {
        const cuuint32_t box[3] = {(cuuint32_t)(8 * (wi + 1)), (cuuint32_t)(8 >> ci), (cuuint32_t)rows_of[ri]};
        const CUresult r = enc(&e.m[(wi * 3 + ri) * 3 + ci], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                               const_cast<float*>(ptr), gdim, gstride, box, estride,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return 0;
      }
  e.ptr = ptr;
  e.planes = planes;
  e.H = H;
  e.W = W;
  e.pitch = pitch;
  e.ok = 1;
  e.stamp = ++clock_;
  memcpy(maps->m[l], e.m, sizeof(e.m));
  return 1;
}

}  // namespace

static int* rx_dbg_host_ptr = nullptr;
// diagnostic (RP_RX_DEBUG=1): per warp {set, block, wait code, parity, counter} of the last barrier timeout
extern "C" RP_API int rp_debug_rx_timeouts(int* out, int n) {
  if (!rx_dbg_host_ptr || !out) return 0;
  for (int i = 0; i < n && i < 1024; ++i) out[i] = rx_dbg_host_ptr[i];
  return 1;
}

int rx_forward(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K, int PH,
               int PW, float sampling_ratio, float* out, cudaStream_t st) {
  // eligibility: fixed sampling grid, bins fit 16 accumulator rows / columns, whole 8-channel groups;
  // a level whose rows or base are not 16-byte aligned is re-pitched into a scratch copy for the TMA
  if (!(sampling_ratio > 0) || (int)sampling_ratio > RB_MAXG || PH > 16 || PW > 16 || (C & 7)) return 1;
  const long long planes = (long long)N * C;
  if (planes > 0x7fffffffLL) return 1;
  if (lv.L > RX_MAXL) return 1;
  for (int l = 0; l < lv.L; ++l)
    if (lv.H[l] < 1 || lv.W[l] < 4 || (reinterpret_cast<uintptr_t>(lv.feat[l]) & 3)) return 1;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(rx_pool_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RX_SMEM);
  });
  if (attr_err != cudaSuccess) return (int)attr_err;
  const int nsl = (C + RX_CS - 1) / RX_CS;
  const long long n_items = (long long)K * nsl;
  if (n_items > 0x7fffffffLL) return RP_E_LIMIT;
  static RxMaps maps;  // 11.5 KB: filled under the lock below, copied into the launch by value
  static std::mutex launch_mu;
  std::lock_guard<std::mutex> lock(launch_mu);
  float* scratch[RX_MAXL] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  auto release = [&]() {
    for (int l = 0; l < RX_MAXL; ++l)
      if (scratch[l]) (void)cudaFreeAsync(scratch[l], st);
  };
  for (int l = 0; l < lv.L; ++l) {
    const float* ptr = lv.feat[l];
    int pitch = lv.W[l];
    if ((lv.W[l] & 3) || (reinterpret_cast<uintptr_t>(ptr) & 15)) {
      pitch = (lv.W[l] + 3) & ~3;
      const size_t rows = (size_t)planes * lv.H[l];
      if (cudaMallocAsync(reinterpret_cast<void**>(&scratch[l]), rows * pitch * sizeof(float), st) != cudaSuccess) {
        (void)cudaGetLastError();
        scratch[l] = nullptr;
        release();
        return 1;
      }
      size_t blocks = (rows * pitch + 255) / 256;
      if (blocks > (size_t)RP_NUM_SMS * 16) blocks = (size_t)RP_NUM_SMS * 16;
      rx_repitch_kernel<<<(unsigned)blocks, 256, 0, st>>>(ptr, scratch[l], rows, lv.W[l], pitch);
      ptr = scratch[l];
    }
    if (!rx_level_maps(&maps, l, ptr, planes, lv.H[l], lv.W[l], pitch)) {
      release();
      return 1;
    }
  }
  static int* dbg_host = nullptr;
  if (!dbg_host && rp_env_int("RP_RX_DEBUG", 0)) {
    int* dev = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&dbg_host), 4096, cudaHostAllocMapped) == cudaSuccess &&
        cudaHostGetDevicePointer(reinterpret_cast<void**>(&dev), dbg_host, 0) == cudaSuccess) {
      memset(dbg_host, 0, 4096);
      cudaMemcpyToSymbol(rx_dbg, &dev, sizeof(dev));
      rx_dbg_host_ptr = dbg_host;
    }
  }
  const int grid = (int)(n_items < RP_NUM_SMS ? n_items : RP_NUM_SMS);
  rx_pool_fwd_kernel<<<grid, RX_THREADS, RX_SMEM, st>>>(lv, N, C, rois, levels, K, PH, PW, sampling_ratio,
                                                       out, maps, (int)n_items, rp_env_int("RP_RX_ABLATE", 0));
  const cudaError_t launch_err = cudaGetLastError();
  release();
  if (launch_err != cudaSuccess) return (int)launch_err;
  return RP_OK;
}
