// roi_align.cu -- ROIAlign forward/backward and the fused multi-level FPN Pooler for sm_100a.
//
// Reference being replaced (paths relative to the reference checkout):
//   det/layers/roi_align.py:116-169  RoIAlignForward          (1 thread / output, 16 scalar gathers)
//   det/layers/roi_align.py:190-274  RoIAlignBackwardFeature  (16 global atomics / output)
//   det/modeling/poolers.py:123-156  Pooler.execute           (per-level nonzero/gather/scatter loop)
//   det/modeling/poolers.py:31-43,62-73  LevelMapper / LevelMapperwithArea
//
// Design (see DESIGN.md "ROIAlign"): one CTA per RoI (x an optional channel split).  The CTA
// builds the two 1-D sweep plans of the RoI once (roi_tile.h), then for each chunk of
// channels
//   phase 1  lanes along x, U1 channels per thread: one monotone sweep down the RoI's
//            feature tile straight from global memory (coalesced row segments, every tile
//            element read exactly once, D1*U1 loads in flight per thread) -> V[c][ph][x] in smem
//   phase 2  thread = (c, ph) row of V: one monotone sweep along x from shared memory
//            -> out tile [c][ph][pw] in smem
//   store    the chunk's output is contiguous in global memory: one TMA bulk store
//            (cp.async.bulk shared->global) issued by a single thread.
// All RoIs of all FPN levels go through ONE launch: the level only selects the base pointer,
// extent and scale, and every RoI writes its own output row (no nonzero / gather / scatter).
// Backward is the exact transpose with the RoI's gradient tile accumulated privately and
// flushed with ONE coalesced atomicAdd per tile element (instead of 16 per output element).
#include <cuda.h>  // CUtensorMap types only; the encoder is fetched through the runtime

#include <mutex>

#include "common.cuh"
#include "roi_levels.h"
#include "roi_tile.h"

namespace {

constexpr int RT_THREADS = 256;
#ifndef RP_ROI_IMPL_DEFAULT
#define RP_ROI_IMPL_DEFAULT 3  // forward kernel: 3 = tile walk, 4 = banded, 5 = merged taps
#endif
#ifndef RP_ROI_TAP3_DEFAULT
#define RP_ROI_TAP3_DEFAULT 1  // three-tap rows in the sample walk (merged x weights): 2.24 -> 2.11 ms
#endif
#ifndef RP_ROI_WALK_TABS_DEFAULT
#define RP_ROI_WALK_TABS_DEFAULT 0  // the column walk reads the tables kernel's sample tables
#endif
#ifndef RP_ROI_YMERGE_DEFAULT
#define RP_ROI_YMERGE_DEFAULT 1  // merged y-samples in the walk table: 1.995 -> 1.93 ms
#endif
#ifndef RP_ROI_PLANES_DEFAULT
#define RP_ROI_PLANES_DEFAULT 1  // whole-plane kernel for the coarse levels (roi_fwd_plane.cuh): 2.11 -> 1.98 ms
#endif
constexpr int RT_U1 = 4;  // channels per thread, phase 1
constexpr int RT_U2 = 2;  // channels per thread, phase 2
constexpr int RT_CC_CAP = 32;  // max channels per chunk (bounds the out tile)

__device__ __forceinline__ void rt_zero_fill(float* dst, size_t n) {
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = 0.f;
}

// ------------------------------------------------------------------------------------------
// forward kernel (v3): TMA-staged raw tile, one thread per (output column, U channels, segment)
// ------------------------------------------------------------------------------------------
// Per RoI the CTA builds, in parallel, the y-sample table (weights + tile row, one 16-byte
// entry per sample) and the x-sample table in shared memory.  Then, per chunk of channels:
//   stage    the RoI's feature tile [cc][R][twa] goes global -> shared.  When the map's rows
//            are 16-byte aligned every tile row is ONE TMA bulk copy (cp.async.bulk, SASS
//            UBLKCP) whose completion is counted on an mbarrier; otherwise 4-byte cp.async.
//   compute  roi_tile.h rt_fwd_item: thread (pw, U channels, segment of output rows) walks the
//            tile rows once; finished bins are scaled and streamed to global memory
//            (lanes = consecutive pw: 4*PW-byte runs per channel).
// The CTA has PW * gpb compute threads (gpb a power of two), so the item count per chunk,
// gpb = ncg channel groups x nseg segments, always fills the CTA: a large tile (few channels
// fit) is cut into more segments instead of idling threads.
constexpr int RT3_MAX_THREADS = 256;

__device__ __forceinline__ void rt_cp_async16(void* sdst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(rp_smem_addr(sdst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void rt_cp_async4(void* sdst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(rp_smem_addr(sdst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void rt_cp_async8(void* sdst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(rp_smem_addr(sdst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void rt_cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
// exact floor(j / d) for j * d < 2^32, with m = rt_magic(d)
__device__ __forceinline__ uint32_t rt_magic(uint32_t d) { return 0xffffffffu / d + 1u; }
__device__ __forceinline__ uint32_t rt_div(uint32_t j, uint32_t d, uint32_t m) {
  return d == 1 ? j : __umulhi(j, m);
}

// Stage cc channels of a [R rows][vw columns of VEC floats] tile: warp w owns channels w, w+nw, ...
// and its lanes walk the channel's columns row-major (coalesced row segments; one magic
// division per copy, no per-copy channel arithmetic).  Shared rows are contiguous (pitch
// vw * VEC), channel pitch chp.
template <int VEC>
__device__ __forceinline__ void rt_stage_tile(float* tile, const float* __restrict__ src, int cc,
                                              int R, int vw, uint32_t magicV, int W, size_t HW,
                                              int chp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int per_ch = R * vw;
  for (int c = warp; c < cc; c += nw) {
    const float* g = src + (size_t)c * HW;
    float* s = tile + c * chp;
    for (int i = lane; i < per_ch; i += 32) {
      const int y = (int)rt_div((uint32_t)i, (uint32_t)vw, magicV);
      const int q = i - y * vw;
      const float* gp = g + (y * W + q * VEC);
      if (VEC == 4)
        rt_cp_async16(s + i * 4, gp);
      else if (VEC == 2)
        rt_cp_async8(s + i * 2, gp);
      else
        rt_cp_async4(s + i, gp);
    }
  }
}

// Per-thread staging constants of one RoI (computed once, used for every chunk): the lanes of a
// warp form rpw = 32 / vpr rows of vpr staging vectors (VEC floats each) and step down the tile
// rpw rows at a time; warp w stages channels w, w + nw, ...  One cp.async, one 64-bit and one
// 32-bit add per copy; no division inside the chunk loop.
struct RbStageLane {
  int active;     // this lane copies (lanes past rpw * vpr idle)
  int vrow;       // first tile row of the lane
  int rpw;        // rows per warp step
  int zero;       // the lane's columns lie past the end of the map row: store zeros instead
  size_t goff;    // vrow * W + col
  uint32_t soff;  // byte offset of (vrow, col) inside a channel's tile
};

// vpr staging vectors of `vec` floats per tile row; tile rows are rpitch floats apart
__device__ __forceinline__ RbStageLane rb_stage_lane(int vec, int vpr, int rpitch, int W, int wlim) {
  RbStageLane l;
  const int lane = threadIdx.x & 31;
  int rpw = 32 / vpr;
  if (rpw < 1) rpw = 1;  // rows wider than 32 vectors (vec < 4 only): not on the band path
  l.rpw = rpw;
  l.vrow = lane / vpr;
  const int col = (lane - l.vrow * vpr) * vec;
  l.active = l.vrow < rpw;
  l.zero = col >= wlim;
  l.goff = (size_t)l.vrow * W + col;
  l.soff = 4u * (uint32_t)(l.vrow * rpitch + col);
  return l;
}

template <int VEC>
__device__ __forceinline__ void rb_stage_tile(uint32_t tile_s, const float* __restrict__ src, int cc,
                                              int R, int rpitch, int W, size_t HW, int chp,
                                              const RbStageLane& l) {
  if (!l.active) return;
  const int warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const size_t gstep = (size_t)l.rpw * W;
  const uint32_t sstep = 4u * (uint32_t)(l.rpw * rpitch);
  const float* gc = src + (size_t)warp * HW + l.goff;
  uint32_t sc = tile_s + 4u * (uint32_t)(warp * chp) + l.soff;
  const size_t gcs = (size_t)nw * HW;
  const uint32_t scs = 4u * (uint32_t)(nw * chp);
  for (int c = warp; c < cc; c += nw, gc += gcs, sc += scs) {
    const float* gp = gc;
    uint32_t sp = sc;
    if (VEC == 4 || !l.zero) {
      for (int y = l.vrow; y < R; y += l.rpw) {
        if (VEC == 4)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sp), "l"(gp) : "memory");
        else if (VEC == 2)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sp), "l"(gp) : "memory");
        else
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sp), "l"(gp) : "memory");
        gp += gstep;
        sp += sstep;
      }
    } else {  // columns past the end of the map row: never weighted, but keep them finite
      for (int y = l.vrow; y < R; y += l.rpw) {
#pragma unroll
        for (int e = 0; e < VEC; ++e)
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(sp + 4u * e), "f"(0.f) : "memory");
        sp += sstep;
      }
    }
  }
}

// global sink of the finished bins of one item: out[c_u][bin][pw], streaming stores
template <int U>
struct FwdGlobalSink {
  float* optr;     // &out[k][c of channel 0][b0][pw], advanced by PW per bin
  int cstride;     // ncg * PH * PW
  int PW, nvalid;
  float inv;  // 1 / (gy * gx): within 1 ulp of the reference's division
  __device__ __forceinline__ void operator()(int, const float (&acc)[U]) {
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u < nvalid) {
        __stcs(optr + u * cstride, acc[u] * inv);
      }
    optr += PW;
  }
};

#ifndef RT_ST_POLICY
#define RT_ST_POLICY 0  // output stores of the sample walk: 0 = streaming (.cs), 1 = plain, 2 = .cg (measured: see DESIGN 4.1)
#endif
#if RT_ST_POLICY == 0
#define RT_ST_OUT(p, v) __stcs(p, v)
#elif RT_ST_POLICY == 1
#define RT_ST_OUT(p, v) (*(p) = (v))
#elif RT_ST_POLICY == 2
#define RT_ST_OUT(p, v) __stcg(p, v)
#else  // 3: timing experiment, results wrong: the value is computed but (practically) never stored
#define RT_ST_OUT(p, v) do { const float rt_v_ = (v); if (rt_v_ == 12345.678f) *(p) = rt_v_; } while (0)
#endif
// rt_fwd_item_fsm (roi_tile.h) with the address arithmetic done by hand: the same operations on the
// same values in the same order (bit-identical results), but
//   * tile and sample-table accesses are ld.shared with 32-bit addresses: ONE add per (row, channel,
//     tap pair) instead of the 64-bit index * 4 + base the generic-pointer form compiles to,
//   * the U output pointers are kept as running 64-bit pointers (one add per emitted row each).
// ncu on the generic form: IMAD 19 % and LEA / IADD3 5 % of the instruction stream were addresses.
__device__ __forceinline__ float rt_lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float rt_lds_f32_4(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+4];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ RtYSample rt_lds_sample(uint32_t a) {
  RtYSample e;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=f"(e.wlo), "=f"(e.whi), "=r"(e.row), "=r"(e.flags)
               : "r"(a));
  return e;
}

// TAP3: the two x-samples of every output column of this RoI lie in the same or in adjacent tile
// columns (rt_x_tap3: true for every bin of at most 2 feature pixels), so a row needs 3 loads and 3 FMAs
// per channel with merged weights instead of 4 + 4 (reassociation only, ~1e-7 relative); the third
// load repeats column lo+1 with weight 0 when both samples share a column pair.
template <int U, bool TAP3 = false>
__device__ __forceinline__ void rt_fwd_item_fsm_s(uint32_t tile_a, uint32_t ustride4, int nvalid,
                                                  uint32_t ytab_a, const RtW* xw, const short* xo_t,
                                                  int gy, uint32_t tw4, int pw, int b0, int b1,
                                                  float* optr, int cstride, int PW, float inv) {
  if (b0 >= b1) return;
  RtW w0 = xw[pw * 2], w1 = xw[pw * 2 + 1];
  float w3a = 0.f, w3b = 0.f, w3c = 0.f;
  if (TAP3) {
    const bool same = xo_t[pw * 2 + 1] == xo_t[pw * 2];
    w3a = same ? w0.lo + w1.lo : w0.lo;
    w3b = same ? w0.hi + w1.hi : w0.hi + w1.lo;
    w3c = same ? 0.f : w1.hi;
  }
  uint32_t st = ytab_a + (uint32_t)(b0 * gy) * 16u;
  int bins_left = b1 - b0;
  /* the first sample of a segment: its row pair is computed from scratch, flags ignored */
  RtYSample e = rt_lds_sample(st);
  st += 16u;
  uint32_t a0[U], a1[U];
  float* op[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint32_t b = tile_a + (u < nvalid ? (uint32_t)u * ustride4 : 0u) + (uint32_t)e.row * tw4;
    a0[u] = b + 4u * (uint32_t)xo_t[pw * 2];
    a1[u] = b + 4u * (uint32_t)xo_t[pw * 2 + 1];  // TAP3: its +4 is the third column (lo+2, or lo+1 again)
    op[u] = optr + (size_t)u * cstride;
  }
  float Ta[U], Tb[U], acc[U];
#define RS_XROW(T)                                                                              \
  _Pragma("unroll") for (int u = 0; u < U; ++u) {                                               \
    if (TAP3)                                                                                   \
      T[u] = fmaf(w3c, rt_lds_f32_4(a1[u]),                                                     \
                  fmaf(w3b, rt_lds_f32_4(a0[u]), fmaf(w3a, rt_lds_f32(a0[u]), 0.f)));           \
    else                                                                                        \
      T[u] = fmaf(w1.hi, rt_lds_f32_4(a1[u]),                                                   \
                  fmaf(w1.lo, rt_lds_f32(a1[u]),                                                \
                       fmaf(w0.hi, rt_lds_f32_4(a0[u]), fmaf(w0.lo, rt_lds_f32(a0[u]), 0.f)))); \
    a0[u] += tw4;                                                                               \
    a1[u] += tw4;                                                                               \
  }
#define RS_JUMP(LO, HI)                                      \
  {                                                          \
    const uint32_t skip = (uint32_t)(e.flags >> 8) * tw4;    \
    _Pragma("unroll") for (int u = 0; u < U; ++u) {          \
      a0[u] += skip;                                         \
      a1[u] += skip;                                         \
    }                                                        \
    RS_XROW(LO)                                              \
    RS_XROW(HI)                                              \
  }
#define RS_YFMA(LO, HI)                                      \
  _Pragma("unroll") for (int u = 0; u < U; ++u) acc[u] =     \
      fmaf(e.whi, HI[u], fmaf(e.wlo, LO[u], acc[u]));
#define RS_EMIT()                                            \
  if (e.flags & RT_F_EMIT) {                                  \
    _Pragma("unroll") for (int u = 0; u < U; ++u) {          \
      if (u < nvalid) RT_ST_OUT(op[u], acc[u] * inv);        \
      op[u] += PW;                                           \
      acc[u] = 0.f;                                          \
    }                                                        \
    if (--bins_left == 0) return;                            \
  }
  RS_XROW(Ta)
  RS_XROW(Tb)
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = 0.f;
  RS_YFMA(Ta, Tb)
  RS_EMIT()
  for (;;) {
    /* state A: Ta = lo row, Tb = hi row */
    for (;;) {
      e = rt_lds_sample(st);
      st += 16u;
      if (e.flags & RT_F_ADV) {
        RS_XROW(Ta) /* Ta becomes the new hi row */
        RS_YFMA(Tb, Ta)
        RS_EMIT()
        break;
      }
      if (e.flags & RT_F_JUMP) RS_JUMP(Ta, Tb)
      RS_YFMA(Ta, Tb)
      RS_EMIT()
    }
    /* state B: Tb = lo row, Ta = hi row */
    for (;;) {
      e = rt_lds_sample(st);
      st += 16u;
      if (e.flags & RT_F_ADV) {
        RS_XROW(Tb)
        RS_YFMA(Ta, Tb)
        RS_EMIT()
        break;
      }
      if (e.flags & RT_F_JUMP) RS_JUMP(Tb, Ta)
      RS_YFMA(Tb, Ta)
      RS_EMIT()
    }
  }
#undef RS_XROW
#undef RS_JUMP
#undef RS_YFMA
#undef RS_EMIT
}

// Sample tables of one RoI as rp_tables_kernel (roi_fwd_plane.cuh) leaves them in global memory: the values
// rt3_roi_forward builds itself, with ABSOLUTE map rows / columns, out-of-range samples already parked.
constexpr int RP_NS = 32;  // samples per axis (bins * 2 <= 32)
struct __align__(16) RpTab {
  RtYSample y[RP_NS];  // absolute row of the lo tap, weights, walk flags
  RtW xw[RP_NS];
  short xo[RP_NS];     // absolute column of the lo tap
  int meta[4];         // [0]: 1 tables valid, 0 every output is zero, 2 not built; [1]: three-tap rows allowed;
                       // [2]: first | last << 16 row that holds a lo tap; [3]: the same for columns
};
static_assert(sizeof(RpTab) % 16 == 0, "RpTab is copied in 16-byte pieces");

// The v3 evaluation of one (RoI, channel slice) by the whole CTA; any CTA size that is a multiple
// of gw.  Also the fallback of the banded kernel below for RoIs without a band form.
template <int U, bool LEAN = false>
__device__ __forceinline__ void rt3_roi_forward(unsigned char* rt_smem, const RtLevels& lv, int N,
                                                int C, const float* __restrict__ rois,
                                                const int* __restrict__ levels, int PH, int PW,
                                                float sampling_ratio, float* __restrict__ out,
                                                int c_per_cta, int tile_floats, int gpb, int gw,
                                                int nbuf = 1, const RpTab* __restrict__ tabs = nullptr) {
  // layout: ytab[RT_SMAX] | xw[RT_SMAX] | xo[RT_SMAX] | tile[tile_floats]
  // nbuf bit 4: three-tap rows where the RoI allows them.
  // nbuf bit 5: y-merge (below).
  // (nbuf & 3) == 2: the tile buffer is cut into two halves and the copies of chunk i+1 are issued before
  // chunk i is walked, so they fly under the walk (a timing ablation put the exposed copies at 0.6 of
  // the kernel's 2.2 ms: DESIGN.md 4.1); a RoI whose tile does not fit a half keeps the single buffer
  RtYSample* ytab = reinterpret_cast<RtYSample*>(rt_smem);
  RtW* xw_s = reinterpret_cast<RtW*>(ytab + RT_SMAX);
  short* xo_s = reinterpret_cast<short*>(xw_s + RT_SMAX);
  float* tile = reinterpret_cast<float*>(xo_s + RT_SMAX);
  __shared__ int s_ext[4];  // y: min lo, max lo ; x: min lo, max lo

  const int tid = threadIdx.x;
  const int nthreads = blockDim.x;
  const int k = blockIdx.x;
  const int c_begin = blockIdx.y * c_per_cta;
  const int c_end = min(C, c_begin + c_per_cta);
  if (c_begin >= c_end) return;
  const int nch = c_end - c_begin;
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  float* out_k = out + ((size_t)k * C + c_begin) * P;
  if (level < 0 || level >= lv.L || b < 0 || b >= N) {
    rt_zero_fill(out_k, (size_t)nch * P);
    return;
  }
  const int H = lv.H[level], W = lv.W[level];
  const size_t HW = (size_t)H * W;
  const float* feat_b = lv.feat[level] + ((size_t)b * C + c_begin) * HW;
  const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  const int SY = PH * g.gh, SX = PW * g.gw;
  const bool fits = (SY <= RT_SMAX) && (SX <= RT_SMAX) && SY > 0 && SX > 0 && H >= 2 && W >= 2;

  // ---- tables: read from rp_tables_kernel's copy when there is one (one warp built them once per RoI),
  //      else taps of every y- and x-sample in parallel, extents by shared atomics ------------------
  const RpTab* tb = tabs ? tabs + k : nullptr;
  const bool from_tabs = tb && fits && g.gh == 2 && g.gw == 2 && SY <= RP_NS && SX <= RP_NS && tb->meta[0] != 2;
  int y0, yl, x0, xl;
  if (from_tabs) {
    if (tid < 32) {
      if (tid < SY) ytab[tid] = tb->y[tid];
    } else if (tid < 64) {
      const int t = tid - 32;
      if (t < SX) {
        xw_s[t] = tb->xw[t];
        xo_s[t] = tb->xo[t];
      }
    }
    const int valid = tb->meta[0], m2 = tb->meta[2], m3 = tb->meta[3];
    y0 = m2 & 0xffff;
    yl = valid ? (m2 >> 16) : -1;
    x0 = m3 & 0xffff;
    xl = valid ? (m3 >> 16) : -1;
    __syncthreads();
  } else {
    if (tid < 4) s_ext[tid] = (tid & 1) ? -1 : 0x7fffffff;
    __syncthreads();
    if (fits) {
      for (int s = tid; s < SY + SX; s += nthreads) {
        const bool isy = s < SY;
        const int t = isy ? s : s - SY;
        const int grid = isy ? g.gh : g.gw;
        const int p = t / grid, i = t - p * grid;
        int lo = 0;
        float wlo = 0.f, whi = 0.f;
        const int ok = isy ? rt_axis_tap_pair(g.sh, g.bh, p, i, grid, H, &lo, &wlo, &whi)
                           : rt_axis_tap_pair(g.sw, g.bw, p, i, grid, W, &lo, &wlo, &whi);
        if (ok) {
          atomicMin(&s_ext[isy ? 0 : 2], lo);
          atomicMax(&s_ext[isy ? 1 : 3], lo);
        }
        if (isy) {
          ytab[t].wlo = wlo;
          ytab[t].whi = whi;
          ytab[t].row = lo;
          ytab[t].flags = (i == grid - 1) ? RT_F_EMIT : 0;
        } else {
          xw_s[t].lo = wlo;
          xw_s[t].hi = whi;
          xo_s[t] = (short)lo;
        }
      }
    }
    __syncthreads();
    y0 = s_ext[0], yl = s_ext[1], x0 = s_ext[2], xl = s_ext[3];
  }
  if (fits && (yl < 0 || xl < 0)) {  // no sample inside the map on some axis
    rt_zero_fill(out_k, (size_t)nch * P);
    return;
  }
  const int R = yl - y0 + 2;  // the hi tap of every sample is lo + 1
  // staging width: TMA rows need 16-byte aligned row starts (W % 4 == 0, aligned base pointer);
  // maps with even W are staged with 8-byte cp.async, anything else with 4-byte cp.async
  const uintptr_t base_addr = reinterpret_cast<uintptr_t>(lv.feat[level]);
  const int vecw = (((W & 3) == 0) && ((base_addr & 15) == 0)) ? 4
                   : (((W & 1) == 0) && ((base_addr & 7) == 0)) ? 2 : 1;
  const int xa0 = x0 & ~(vecw - 1);
  const int twa = ((xl + 1) / vecw - xa0 / vecw + 1) * vecw;
  RtTilePlan tp;
  tp.cc = 0;
  int nbufs = 1, buf_floats = tile_floats;
  if (fits && (nbuf & 3) == 2) {
    const int half = (tile_floats / 2) & ~31;
    tp = rt_plan_tile(R, twa, half, gpb, gw, U, nch, PH);
    if (tp.cc >= 1 && tp.cc < nch) {  // more than one chunk: worth the second buffer
      nbufs = 2;
      buf_floats = half;
    } else {
      tp.cc = 0;
    }
  }
  if (fits && nbufs == 1) tp = rt_plan_tile(R, twa, tile_floats, gpb, gw, U, nch, PH);
  if (tp.cc < 1) {
    // RoI too large for the tile buffer (or a degenerate map): reference-order evaluation
    const int total = nch * P;
    for (int i = tid; i < total; i += nthreads) {
      const int c = i / P;
      const int rem = i - c * P;
      const int ph = rem / PW, pw = rem - ph * PW;
      out_k[i] = rt_direct_output(feat_b + (size_t)c * HW, H, W, g, ph, pw);
    }
    return;
  }
  // tile-relative rows / columns; out-of-range samples (zero weights) park on the nearest one
  for (int s = tid; s < SY + SX; s += nthreads) {
    if (s < SY) {
      const int lo = ytab[s].row;
      ytab[s].row = from_tabs ? lo - y0 : (lo >= 0 ? lo - y0 : (lo == -1 ? 0 : yl - y0));
    } else {
      const int t = s - SY;
      const int lo = xo_s[t];
      xo_s[t] = (short)(from_tabs ? lo - xa0 : (lo >= 0 ? lo - xa0 : (lo == -1 ? x0 - xa0 : xl - xa0)));
    }
  }
  __syncthreads();
  // y-merge: when the two y-samples of a bin share a row pair their weights are added and the bin keeps
  // ONE table entry (every second bin on the fine level: fewer entries, loads and FMAs in the walk;
  // reassociation only).  Only when every item walks its whole column (one segment: the walk then starts
  // at entry 0 and stops on the bin count, so no per-bin start index is needed).
  if ((nbuf & 32) && LEAN && g.gh == 2 && g.gw == 2 && tp.nseg == 1 && SY <= 32 && tid < 32) {
    const int t = tid;
    RtYSample e = ytab[t < SY ? t : 0];
    const int rown = __shfl_down_sync(0xffffffffu, e.row, 1);
    const float wlon = __shfl_down_sync(0xffffffffu, e.wlo, 1), whin = __shfl_down_sync(0xffffffffu, e.whi, 1);
    const int m = (t < SY) && !(t & 1) && (rown == e.row);  // decided on the bin's first sample
    const int mprev = __shfl_up_sync(0xffffffffu, m, 1);
    const bool keep = (t < SY) && !((t & 1) && mprev);
    if (m) {
      e.wlo += wlon;
      e.whi += whin;
    }
    e.flags = ((t & 1) || m) ? RT_F_EMIT : 0;  // last entry of its bin
    const unsigned km = __ballot_sync(0xffffffffu, keep);
    const int pos = __popc(km & ((1u << t) - 1u));
    __syncwarp();
    if (keep) ytab[pos] = e;
  }
  __syncthreads();
  // how the live row pair moves before each y-sample (rows are final now)
  for (int s = tid; s < SY; s += nthreads) {
    const int cur = ytab[s].row;
    const int prev = s ? ytab[s - 1].row : cur;
    ytab[s].flags = rt_ysample_flags(cur, prev, ytab[s].flags & RT_F_EMIT);
  }
  // three-tap rows (rt_fwd_item_fsm_s<U, true>): every output column's two x-samples in the same or in
  // adjacent tile columns -- CTA-uniform, decided here
  int x3ok = 1;
  if (tid < PW && g.gw == 2) {
    const int d = xo_s[2 * tid + 1] - xo_s[2 * tid];
    x3ok = (d == 0) || (d == 1);
  }
  const bool tap3 = __syncthreads_and(x3ok) && g.gw == 2 && (nbuf & 16);

  // ---- this thread's item: (pw, channel group cg, segment seg) -------------------------------
  const int chp = tp.chp;
  const int grp = tid / gw;
  const int pw = tid - grp * gw;
  const bool active = (grp < gpb) && (pw < PW);
  const int ncg_shift = 31 - __clz(tp.ncg);
  const int cg = grp & (tp.ncg - 1);
  const int seg = grp >> ncg_shift;
  const int b0 = active ? seg * PH / tp.nseg : 0;
  const int b1 = active ? (seg + 1) * PH / tp.nseg : 0;
  const int count = g.gh * g.gw;
  const float inv = 1.0f / (float)count;
  const float* src0 = feat_b + (size_t)y0 * W + xa0;
  const int vw = twa / vecw;  // staging vectors (of vecw floats) per tile row
  const bool lane_stage = vw <= 32;  // lanes form whole rows: no division per copy
  const uint32_t magicV = rt_magic((uint32_t)vw);
  const RbStageLane sl = rb_stage_lane(vecw, vw, twa, W, 0x7fffffff);
  const uint32_t tile_s = rp_smem_addr(tile);

  // copies of the chunk that starts at channel c0 into buffer b
  auto stage_chunk = [&](int c0, int b) {
    const int cc = min(tp.cc, nch - c0);
    const float* src = src0 + (size_t)c0 * HW;
    float* tb = tile + b * buf_floats;
    const uint32_t tb_s = tile_s + 4u * (uint32_t)(b * buf_floats);
    if (lane_stage) {
      if (vecw == 4)
        rb_stage_tile<4>(tb_s, src, cc, R, twa, W, HW, chp, sl);
      else if (vecw == 2)
        rb_stage_tile<2>(tb_s, src, cc, R, twa, W, HW, chp, sl);
      else
        rb_stage_tile<1>(tb_s, src, cc, R, twa, W, HW, chp, sl);
    } else if (vecw == 4)
      rt_stage_tile<4>(tb, src, cc, R, vw, magicV, W, HW, chp);
    else if (vecw == 2)
      rt_stage_tile<2>(tb, src, cc, R, vw, magicV, W, HW, chp);
    else
      rt_stage_tile<1>(tb, src, cc, R, vw, magicV, W, HW, chp);
  };
  if (nbufs == 2) {
    stage_chunk(0, 0);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int ib = 0;
  for (int c0 = 0; c0 < nch; c0 += tp.cc) {
    const int cc = min(tp.cc, nch - c0);
    if (nbufs == 2) {
      const bool more = c0 + tp.cc < nch;
#ifdef RT3_ABLATE_NOSTAGE
      if (false)
#endif
      if (more) stage_chunk(c0 + tp.cc, ib ^ 1);
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (more)
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      else
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    } else {
#ifdef RT3_ABLATE_NOSTAGE  // timing experiment, results wrong: only the first chunk of a CTA is staged
      if (c0 == 0)
#endif
      stage_chunk(c0, 0);
      rt_cp_async_wait_all();
    }
    __syncthreads();
    const float* tile_b = tile + ib * buf_floats;
    const uint32_t tile_bs = tile_s + 4u * (uint32_t)(ib * buf_floats);
    if (b0 < b1) {
      int nvalid = (cc - cg + tp.ncg - 1) >> ncg_shift;
      if (nvalid > U) nvalid = U;
      if (nvalid > 0) {
        FwdGlobalSink<U> sink{out_k + (size_t)(c0 + cg) * P + b0 * PW + pw, tp.ncg * P, PW,
                              nvalid, inv};
        const float* tile_g = tile_b + cg * chp;
        const int ustride = tp.ncg * chp;
        if (g.gw == 2 && LEAN && tap3)
          rt_fwd_item_fsm_s<U, true>(tile_bs + 4u * (uint32_t)(cg * chp), 4u * (uint32_t)ustride, nvalid,
                                     rp_smem_addr(ytab), xw_s, xo_s, g.gh, 4u * (uint32_t)twa, pw, b0, b1,
                                     out_k + (size_t)(c0 + cg) * P + b0 * PW + pw, tp.ncg * P, PW, inv);
        else if (g.gw == 2 && LEAN)
          rt_fwd_item_fsm_s<U>(tile_bs + 4u * (uint32_t)(cg * chp), 4u * (uint32_t)ustride, nvalid,
                               rp_smem_addr(ytab), xw_s, xo_s, g.gh, 4u * (uint32_t)twa, pw, b0, b1,
                               out_k + (size_t)(c0 + cg) * P + b0 * PW + pw, tp.ncg * P, PW, inv);
        else if (g.gw == 2)
          rt_fwd_item_fsm<U>(tile_g, ustride, nvalid, ytab, xw_s, xo_s, g.gh, twa, pw, b0, b1, sink);
        else
          rt_fwd_item<0, U>(tile_g, ustride, nvalid, ytab, xw_s, xo_s, g.gh, g.gw, twa, pw, b0, b1, sink);
      }
    }
    __syncthreads();  // the buffer is overwritten by the copies issued in the next iteration
    if (nbufs == 2) ib ^= 1;
  }
}

constexpr int RT3_TABLE_BYTES =
    sizeof(RtYSample) * RT_SMAX + sizeof(RtW) * RT_SMAX + sizeof(short) * RT_SMAX;

#include "roi_fwd_mma.cuh"  // rm_pool_fwd_kernel: both passes on the warp-level tensor path
#include "roi_fwd_plane.cuh"  // rp_plane_fwd_kernel: whole planes of the coarse levels staged once

#ifndef RT3_MINB
#define RT3_MINB 4  // resident CTAs per SM the column-walk kernel is compiled for (64 registers)
#endif
template <int U, bool LEAN>
__global__ void __launch_bounds__(RT3_MAX_THREADS, RT3_MINB)
rt_pool_fwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   float* __restrict__ out, int c_per_cta, int tile_floats, int gpb, int gw, int nbuf,
                   const int* __restrict__ skip, const RpTab* __restrict__ tabs) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  if (skip && skip[blockIdx.x] >= 0) return;  // taken by the whole-plane kernel (roi_fwd_plane.cuh)
  rt3_roi_forward<U, LEAN>(rt_smem, lv, N, C, rois, levels, PH, PW, sampling_ratio, out, c_per_cta,
                     tile_floats, gpb, gw, nbuf, tabs);
}

// ------------------------------------------------------------------------------------------
// forward kernel (v4, banded): see roi_tile.h "Forward (v4)".  CTA = one RoI x one channel
// slice, RB_THREADS threads, 5 CTAs per SM.  Per chunk of cc channels:
//   stage   the RoI's tile [cc][R rows][nq 16-byte strips] global -> shared with cp.async
//           (issued right after the previous chunk's y-pass, so it overlaps the x-pass);
//   y-pass  one thread per (channel, strip, segment): rb_ywalk reads the strip once and writes
//           V[c][ph][4 columns] (16-byte shared stores);
//   x-pass  thread = (pw, row slot): 4 shared loads + 4 FMA per output, thread t writes float t
//           of the chunk's contiguous output block (128-byte lines per warp, streaming).
// RoIs without a band form (bins wider than 4 feature pixels at sampling 2, adaptive sampling
// of large RoIs, maps smaller than 4x4) are evaluated by the v3 body with the same CTA.
// ------------------------------------------------------------------------------------------
constexpr int RB_THREADS = 128;

// CTA-uniform launch geometry of one RoI, written by thread 0.
struct RbGeo {
  int ok, cc, nseg, nq, vp, chp4, xa0, vecw;
  int tma;        // tile staged by TMA boxes: layout [segment][row][channel][4nq]
  int segoff4[RB_MAXSEG];  // first 16-byte unit of each segment's rows (TMA layout)
  int rows_total;
};

// TMA descriptors of the feature maps: per level one 3-D tensor (W, H, N*C planes) and one
// descriptor per box width 4*nq floats (nq = 1..16), box = (4nq, 1 row, RB_BOXC planes).  A box
// lands in shared memory as [RB_BOXC][4nq]: consecutive channels' strips are contiguous, so the
// strip loads of a warp are one conflict-free 512-byte run.  Levels whose rows are not 16-byte
// multiples (W % 4 != 0) cannot be TMA tensors and keep the cp.async path.
constexpr int RB_TMA_LEVELS = 5;
constexpr int RB_TMA_NQ = 16;
constexpr int RB_BOXC = 8;
struct RbMaps {
  CUtensorMap m[RB_TMA_LEVELS][RB_TMA_NQ];
  int ok[RB_TMA_LEVELS];
};

template <int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
rb_pool_fwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   float* __restrict__ out, int c_per_cta, int smem_floats, int gpb, int gw, int nbuf,
                   const __grid_constant__ RbMaps maps) {
  extern __shared__ __align__(128) unsigned char rt_smem[];
  __shared__ RbPlan plan;
  __shared__ RbGeo geo;
  __shared__ int s_fin[2];
  __shared__ __align__(8) uint64_t s_bar[2];

  const int tid = threadIdx.x;
  const int k = blockIdx.x;
  const int c_begin = blockIdx.y * c_per_cta;
  const int c_end = min(C, c_begin + c_per_cta);
  if (c_begin >= c_end) return;
  const int nch = c_end - c_begin;
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  float* out_k = out + ((size_t)k * C + c_begin) * P;
  if (level < 0 || level >= lv.L || b < 0 || b >= N) {
    rt_zero_fill(out_k, (size_t)nch * P);
    return;
  }
  const int H = lv.H[level], W = lv.W[level];
  const size_t HW = (size_t)H * W;
  const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  if (tid == 0) {
    rp_mbar_init(&s_bar[0], 1);
    rp_mbar_init(&s_bar[1], 1);
  }

  // ---- band plan of the RoI (CTA-uniform) ----------------------------------------------------
  int ok = rb_axis_candidate(g.bh, g.gh, PH, H) && rb_axis_candidate(g.bw, g.gw, PW, W);
  if (ok) {
    if (tid < PH)
      ok = rb_build_bin(&plan.ay, tid, g.sh, g.bh, g.gh, H);
    else if (tid >= 32 && tid < 32 + PW)
      ok = rb_build_bin(&plan.ax, tid - 32, g.sw, g.bw, g.gw, W);
  }
  ok = __syncthreads_and(ok);
  if (ok) {
    if (tid == 0) s_fin[0] = rb_finish_axis(&plan.ay, PH);
    if (tid == 32) s_fin[1] = rb_finish_axis(&plan.ax, PW);
    __syncthreads();
    if (tid == 0) {
      RbGeo ge;
      ge.ok = s_fin[0] && s_fin[1];
      ge.cc = 0;
      ge.nseg = 1;
      const int x0 = plan.ax.lo0, xe = x0 + plan.ax.ext, R = plan.ay.ext;
      const uintptr_t base_addr = reinterpret_cast<uintptr_t>(lv.feat[level]);
      ge.vecw = (((W & 3) == 0) && ((base_addr & 15) == 0)) ? 4
                : (((W & 1) == 0) && ((base_addr & 7) == 0)) ? 2 : 1;
      ge.xa0 = x0 & ~(ge.vecw - 1);
      ge.nq = (xe - ge.xa0 + 3) >> 2;
      int vpu = ge.nq;  // V row pitch: vpu 16-byte units, vpu >= nq, vpu mod 8 in {3, 5}: the two
      while ((vpu & 7) != 3 && (vpu & 7) != 5) ++vpu;  // V rows of an x-pass warp sit 12 / 20 banks apart
      ge.vp = 4 * vpu;
      // channel pitch in 16-byte units, == nq (mod 8): the strips of consecutive channels
      // continue the banks
      ge.chp4 = R * ge.nq + ((ge.nq - R * ge.nq) & 7);
      ge.tma = 0;
      ge.segoff4[0] = ge.segoff4[1] = 0;
      ge.rows_total = R;
      if (ge.ok && level < RB_TMA_LEVELS && maps.ok[level]) {
        // TMA staging: the innermost box coordinate must be a multiple of 16 bytes (measured:
        // unaligned x raises an illegal-instruction fault), so the tile starts at x0 & ~3
        const int nq_t = (xe - (x0 & ~3) + 3) >> 2;
        if (nq_t <= RB_TMA_NQ) {
          int vpu_t = nq_t;
          while ((vpu_t & 7) != 3 && (vpu_t & 7) != 5) ++vpu_t;
          int cc_t = 0, nseg_t = 1;
          rb_plan_chunk(smem_floats, nbuf * (R + 4) * 4 * nq_t + PH * 4 * vpu_t, nq_t, nch, PH, NT,
                        &cc_t, &nseg_t, RB_BOXC);
          if (cc_t >= RB_BOXC) {
            ge.tma = 1;
            ge.xa0 = x0 & ~3;
            ge.nq = nq_t;
            ge.vp = 4 * vpu_t;
            ge.cc = cc_t;
            ge.nseg = nseg_t;
          }
        }
      }
      if (ge.tma) {
        rb_plan_segments(&plan, PH, ge.nseg);
        int acc = 0;
        for (int sg = 0; sg < ge.nseg; ++sg) {
          ge.segoff4[sg] = acc * ge.cc * ge.nq;
          acc += plan.seg_rows[sg];
        }
        ge.rows_total = acc;
      } else {
        if (ge.ok && (4 / ge.vecw) * ge.nq <= 32) {
          rb_plan_chunk(smem_floats, nbuf * 4 * ge.chp4 + PH * ge.vp, ge.nq, nch, PH, NT, &ge.cc,
                        &ge.nseg);
          ge.ok = ge.cc >= 1;
        } else {
          ge.ok = 0;
        }
        if (ge.ok) rb_plan_segments(&plan, PH, ge.nseg);
      }
      geo = ge;
    }
    __syncthreads();
    ok = geo.ok;
  }
  if (!ok) {
    rt3_roi_forward<2>(rt_smem, lv, N, C, rois, levels, PH, PW, sampling_ratio, out, c_per_cta,
                       (smem_floats * 4 - RT3_TABLE_BYTES) / 4, gpb, gw);
    return;
  }
  const int cc = geo.cc, nseg = geo.nseg, nq = geo.nq, vp = geo.vp, chp4 = geo.chp4;
  const int xa0 = geo.xa0, vecw = geo.vecw, R = plan.ay.ext, y0 = plan.ay.lo0;
  {
    const int count = g.gh * g.gw;
    if (tid < 32) rb_plan_bin(&plan, tid, PH, PW, nseg, plan.ax.lo0 - xa0, 1.0f / (float)count);
    const int nword = RB_MAXE / 4 + 1;
    if (tid >= 32 && tid < 32 + nword) {
      for (int s = 0; s < nseg; ++s)
        if ((tid - 32) * 4 < plan.seg_rows[s] + 4) rb_plan_mword(&plan, s, tid - 32);
    }
  }
  __syncthreads();
  // 3-tap x bands (every RoI narrower than 28 feature pixels): skip the 4th load
  const int tap3 = __syncthreads_and(tid >= PW || plan.xw[tid].w == 0.f);

  // TMA boxes need a 128-byte aligned destination; the dynamic window itself is only 16-byte
  // aligned, so the tile starts at the next 128-byte boundary (the launch adds the slack)
  float* tile = reinterpret_cast<float*>(rt_smem + ((128u - (rp_smem_addr(rt_smem) & 127u)) & 127u));
  const uint32_t tile_s = rp_smem_addr(tile);
  const int chp = 4 * chp4;
  const int tma = geo.tma;
  const int rp4 = tma ? cc * nq : nq;  // tile row pitch in 16-byte units
  const int tsz = tma ? geo.rows_total * cc * 4 * nq : cc * chp;  // floats per tile buffer
  float* V = tile + nbuf * tsz;
  const float* src0 = lv.feat[level] + ((size_t)b * C + c_begin) * HW + (size_t)y0 * W + xa0;
  const RbStageLane sl = rb_stage_lane(vecw, (4 / vecw) * nq, 4 * nq, W, W - xa0);
  // x-pass mapping: groups of xg lanes (16 when PW <= 16, so a warp holds two whole V rows)
  const int xg = PW <= 16 ? 16 : PW;
  const int nslots = NT / xg;
  const int pw = tid % xg, slot = tid / xg;
  const bool xactive = pw < PW && slot < nslots;
  const RbF4 xw = plan.xw[xactive ? pw : 0];
  const int xo = plan.xo[xactive ? pw : 0];
  // y-pass mapping: strip s = (segment, channel, 4-column strip) in units of the FULL chunk, so
  // the decode is done once; channels past the end of a short last chunk are skipped
  const int per_seg = cc * nq, nstrip = per_seg * nseg;
  int ys_c = 0, ys_rows = 0, ys_toff = 0, ys_voff = 0, ys_b0 = 0, ys_seg = 0;
  if (tid < nstrip) {
    ys_seg = tid >= per_seg ? 1 : 0;
    const int rem = tid - ys_seg * per_seg;
    ys_c = rem / nq;
    const int q = rem - ys_c * nq;
    ys_b0 = plan.seg_b0[ys_seg];
    ys_rows = plan.seg_rows[ys_seg];
    ys_toff = tma ? geo.segoff4[ys_seg] + ys_c * nq + q : ys_c * chp4 + plan.seg_r0[ys_seg] * nq + q;
    ys_voff = (ys_c * PH + ys_b0) * vp + 4 * q;
  }

  // TMA staging constants: box i = (tile row i / ng, channel group i % ng)
  const int ng = cc / RB_BOXC;
  const int nbox = geo.rows_total * ng;
  const int plane0 = b * C + c_begin;
  const int x0t = xa0;
  const int seg1_at = nseg > 1 ? plan.seg_rows[0] : 0x7fffffff;
  const int yseg0 = y0 + plan.seg_r0[0];
  const int yseg1 = nseg > 1 ? y0 + plan.seg_r0[1] - plan.seg_rows[0] : 0;
  const void* tmap = tma ? (const void*)&maps.m[level][nq - 1] : nullptr;
  auto stage = [&](int c0, int ccur, int buf) {
    if (tma) {
      if (tid < 32) {
        uint64_t* bar = &s_bar[buf];
        if (tid == 0) rp_mbar_arrive_expect_tx(bar, (uint32_t)nbox * (uint32_t)(RB_BOXC * 16 * nq));
        __syncwarp();
        float* dst = tile + buf * tsz;
        for (int i = tid; i < nbox; i += 32) {
          const int ri = i / ng, gi = i - ri * ng;
          const int y = ri < seg1_at ? yseg0 + ri : yseg1 + ri;
          rp_tma_load_3d(dst + (size_t)(ri * cc + gi * RB_BOXC) * 4 * nq, tmap, x0t, y,
                         plane0 + c0 + gi * RB_BOXC, bar);
        }
      }
      return;
    }
    const float* src = src0 + (size_t)c0 * HW;
    const uint32_t dst_s = tile_s + 4u * (uint32_t)(buf * tsz);
    if (vecw == 4)
      rb_stage_tile<4>(dst_s, src, ccur, R, 4 * nq, W, HW, chp, sl);
    else if (vecw == 2)
      rb_stage_tile<2>(dst_s, src, ccur, R, 4 * nq, W, HW, chp, sl);
    else
      rb_stage_tile<1>(dst_s, src, ccur, R, 4 * nq, W, HW, chp, sl);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  const int tsz4 = tsz / 4;
  stage(0, min(cc, nch), 0);
  for (int c0 = 0, j = 0; c0 < nch; c0 += cc, ++j) {
    const int ccur = min(cc, nch - c0);
    const bool more = c0 + cc < nch;
    const int buf = nbuf == 2 ? (j & 1) : 0;
    // double-buffered: the next chunk's copies are in flight during this chunk's y- and x-pass
    // (its buffer was last read by the y-pass before the previous chunk's second barrier)
    if (nbuf == 2 && more) stage(c0 + cc, min(cc, nch - c0 - cc), buf ^ 1);
    if (tma) {
      rp_mbar_wait_bounded(&s_bar[buf], (uint32_t)(nbuf == 2 ? (j >> 1) : j) & 1u);
    } else if (nbuf == 2 && more) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();  // tile(c0) has landed; every thread is done with V of the previous chunk
    // ---- y-pass ------------------------------------------------------------------------------
    const RbF4* tile4 = reinterpret_cast<const RbF4*>(tile) + buf * tsz4;
    if (tid < nstrip && ys_c < ccur)
      rb_ywalk(tile4 + ys_toff, rp4, ys_rows, plan.mw[ys_seg], plan.wrot + ys_b0, V + ys_voff, vp);
    for (int s = tid + NT; s < nstrip; s += NT) {  // more strips than threads
      const int seg = s >= per_seg ? 1 : 0;
      const int rem = s - seg * per_seg;
      const int c = rem / nq, q = rem - c * nq;
      const int b0 = plan.seg_b0[seg];
      if (c < ccur)
        rb_ywalk(tile4 + (tma ? geo.segoff4[seg] + c * nq + q : c * chp4 + plan.seg_r0[seg] * nq + q),
                 rp4, plan.seg_rows[seg], plan.mw[seg], plan.wrot + b0, V + (c * PH + b0) * vp + 4 * q,
                 vp);
    }
    __syncthreads();  // V complete, tile free
    if (nbuf == 1 && more) stage(c0 + cc, min(cc, nch - c0 - cc), 0);
    // ---- x-pass ------------------------------------------------------------------------------
    if (xactive) {
      const float* v = V + slot * vp + xo;
      float* o = out_k + (size_t)c0 * P + slot * PW + pw;
      const int nrow = ccur * PH;
      const int vstep = nslots * vp, ostep = nslots * PW;
      if (tap3) {
#pragma unroll 2
        for (int r = slot; r < nrow; r += nslots) {
          __stcs(o, fmaf(xw.z, v[2], fmaf(xw.y, v[1], xw.x * v[0])));
          v += vstep;
          o += ostep;
        }
      } else {
#pragma unroll 2
        for (int r = slot; r < nrow; r += nslots) {
          __stcs(o, fmaf(xw.w, v[3], fmaf(xw.z, v[2], fmaf(xw.y, v[1], xw.x * v[0]))));
          v += vstep;
          o += ostep;
        }
      }
    }
  }
}

// mode 1: one thread per output element, reference operand order (bit-exact vs the oracle).
__global__ void __launch_bounds__(256)
rt_pool_fwd_exact_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                         const int* __restrict__ levels, size_t total, int PH, int PW,
                         float sampling_ratio, float* __restrict__ out) {
  const int P = PH * PW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int pw = (int)(i % PW);
    const int ph = (int)((i / PW) % PH);
    const int c = (int)((i / P) % C);
    const int k = (int)(i / ((size_t)P * C));
    const float* roi = rois + (size_t)k * 5;
    const int level = levels ? levels[k] : 0;
    const int b = (int)roi[0];
    if (level < 0 || level >= lv.L || b < 0 || b >= N) {
      out[i] = 0.f;
      continue;
    }
    const int H = lv.H[level], W = lv.W[level];
    const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
    out[i] = rt_direct_output(lv.feat[level] + ((size_t)b * C + c) * H * W, H, W, g, ph, pw);
  }
}

// ------------------------------------------------------------------------------------------
// backward functors
// ------------------------------------------------------------------------------------------
template <int U>
struct BwdDoutSource {  // phase A: dout tile [c][ph][bin=pw] in smem
  const float* base;    // &d_s[(q*PH + ph)*PW]
  int cstride;
  int nvalid;
  __device__ __forceinline__ void operator()(int bin, float (&g)[U]) const {
#pragma unroll
    for (int u = 0; u < U; ++u) g[u] = (u < nvalid) ? base[u * cstride + bin] : 0.f;
  }
};

template <int U>
struct BwdHEmit {  // phase A: Hs[c][ph][x]
  float* base;     // &Hs[(q*PH + ph)*twp]
  int cstride;
  int nvalid;
  __device__ __forceinline__ void operator()(int r, const float (&v)[U]) const {
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u < nvalid) base[u * cstride + r] = v[u];
  }
};

template <int U>
struct BwdHSource {  // phase B: Hs[c][bin=ph][x]
  const float* base; // &Hs[(q*PH)*twp + x]
  int bstride;       // twp
  int cstride;
  int nvalid;
  __device__ __forceinline__ void operator()(int bin, float (&g)[U]) const {
    const float* p = base + bin * bstride;
#pragma unroll
    for (int u = 0; u < U; ++u) g[u] = (u < nvalid) ? p[u * cstride] : 0.f;
  }
};

template <int U>
struct BwdGradEmit {  // phase B: one atomicAdd per tile element, lanes along x (coalesced RED)
  float* base;        // &grad[b][c0+q][lo0y][lo0x + x]
  size_t cstride;
  int W;
  int nvalid;
  float inv;
  float count;
  __device__ __forceinline__ void operator()(int r, const float (&v)[U]) const {
    float* p = base + (size_t)r * W;
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u < nvalid) {
        const float val = (inv != 0.f) ? v[u] * inv : v[u] / count;
#if defined(RT_BWD_ABLATE_STORE)  // timing ablation only (wrong results): plain store instead of RED
        if (val != 0.f) p[(size_t)u * cstride] = val;
#elif defined(RT_BWD_ABLATE_NOFLUSH)  // timing ablation only: no global traffic at all
        if (val == 123.456f) p[(size_t)u * cstride] = val;
#else
        if (val != 0.f) atomicAdd(p + (size_t)u * cstride, val);
#endif
      }
  }
};

// Reference-order backward for one output element (det/layers/roi_align.py:62-111,:236-272).
__device__ __forceinline__ void rt_direct_backward(float* plane, int H, int W, const RtGeom& g,
                                                   int ph, int pw, float top) {
  const float count = (float)(g.gh * g.gw);
  for (int iy = 0; iy < g.gh; ++iy) {
    int ylo, yhi;
    float hy, ly;
    const int vy = rt_axis_tap(g.sh, g.bh, ph, iy, g.gh, H, &ylo, &yhi, &hy, &ly);
    for (int ix = 0; ix < g.gw; ++ix) {
      int xlo, xhi;
      float hx, lx;
      const int vx = rt_axis_tap(g.sw, g.bw, pw, ix, g.gw, W, &xlo, &xhi, &hx, &lx);
      if (!(vy && vx)) continue;
      const float w1 = hy * hx, w2 = hy * lx, w3 = ly * hx, w4 = ly * lx;
      atomicAdd(plane + ylo * W + xlo, top * w1 / count);
      atomicAdd(plane + ylo * W + xhi, top * w2 / count);
      atomicAdd(plane + yhi * W + xlo, top * w3 / count);
      atomicAdd(plane + yhi * W + xhi, top * w4 / count);
    }
  }
}

// ------------------------------------------------------------------------------------------
// backward kernel, gather form (default).  Per axis every pooled bin touches a short run of
// consecutive tile rows / columns (at most RG_TAPS) with merged weights; first and last touched
// rows are non-decreasing in the bin index, so the bins that touch one tile row are a contiguous
// range.  Per chunk of channels:
//   stage    dout[c][ph][pw] (contiguous) -> shared
//   phase A  thread = (ph, x), RG_U channels: H[c][ph][x] = sum_{pw covering x} wx * dout
//   phase B  thread = (y, x),  RG_U channels: g = sum_{ph covering y} wy * H[c][ph][x];
//            one RED per tile element, lanes along x
// Every tile element is produced exactly once and the weights / ranges are fetched once for
// RG_U channels.  RoIs without a band form are left to the sweep kernel below (second launch).
// ------------------------------------------------------------------------------------------
#ifndef RG_UCH
#define RG_UCH 8
#endif
constexpr int RG_U = RG_UCH;  // channels per thread
constexpr int RG_TAPS = 8;  // tile rows / columns one bin may touch (merged taps)

struct RgAxis {
  int lo0;             // feature row / column of tile row / column 0
  int ext;             // tile extent
  int start[RB_MAXP];  // first touched row of the bin: absolute until rg_finish_axis, then
                       // tile-relative; non-decreasing (bins without a sample: len 0)
  int len[RB_MAXP];    // touched rows: start .. start + len - 1, at most RG_TAPS
  float w[RB_MAXP][RG_TAPS];  // merged weights of those rows
};

struct RgAxes {
  RgAxis ay, ax;
  int covy[RB_MAXE], covx[RB_MAXE];  // first | last << 8 covering bin of a tile row / column
};

// One bin (independent of the others).  0 when its taps span more than RG_TAPS rows.
__device__ __forceinline__ int rg_build_bin(RgAxis* a, int b, float start, float bin, int grid,
                                            int extent) {
  int first = 0x7fffffff, last = -1;
  for (int i = 0; i < grid; ++i) {
    int lo, hi;
    float wlo, whi;
    if (rt_axis_tap(start, bin, b, i, grid, extent, &lo, &hi, &wlo, &whi)) {
      first = lo < first ? lo : first;
      last = hi > last ? hi : last;
    }
  }
#pragma unroll
  for (int j = 0; j < RG_TAPS; ++j) a->w[b][j] = 0.f;
  if (last < 0) {
    a->start[b] = -1;
    a->len[b] = 0;
    return 1;
  }
  a->start[b] = first;
  a->len[b] = last - first + 1;
  if (last - first + 1 > RG_TAPS) return 0;
  for (int i = 0; i < grid; ++i) {
    int lo, hi;
    float wlo, whi;
    if (rt_axis_tap(start, bin, b, i, grid, extent, &lo, &hi, &wlo, &whi)) {
      a->w[b][lo - first] += wlo;  // only this thread touches row b of w
      a->w[b][hi - first] += whi;
    }
  }
  return 1;
}

// Serial pass: empty bins take a neighbour's start (they cover nothing), starts become
// tile-relative.  0 when the tile is too large.
__device__ __forceinline__ int rg_finish_axis(RgAxis* a, int n) {
  int prev = -1, end = 0;
  for (int b = 0; b < n; ++b)
    if (a->start[b] >= 0) {
      prev = a->start[b];
      break;
    }
  if (prev < 0) prev = 0;  // nothing in range
  const int lo0 = prev;
  for (int b = 0; b < n; ++b) {
    if (a->start[b] < 0)
      a->start[b] = prev;
    else
      prev = a->start[b];
    const int e = a->start[b] + a->len[b];
    end = e > end ? e : end;
    a->start[b] -= lo0;
  }
  a->lo0 = lo0;
  a->ext = end > lo0 ? end - lo0 : 0;
  return a->ext <= RB_MAXE;
}

// CTA-uniform (blockDim.x >= 64): builds the merged-tap axes of the RoI; 1 when both exist.
__device__ __forceinline__ int rg_build_axes(RgAxes* pl, int* fin, const RtGeom& g, int PH, int PW,
                                             int H, int W) {
  if (g.gh < 1 || g.gh > RB_MAXG || g.gw < 1 || g.gw > RB_MAXG || PH > RB_MAXP || PW > RB_MAXP)
    return 0;
  const int tid = threadIdx.x;
  int ok = 1;
  if (tid < PH)
    ok = rg_build_bin(&pl->ay, tid, g.sh, g.bh, g.gh, H);
  else if (tid >= 32 && tid < 32 + PW)
    ok = rg_build_bin(&pl->ax, tid - 32, g.sw, g.bw, g.gw, W);
  if (!__syncthreads_and(ok)) return 0;
  if (tid == 0) fin[0] = rg_finish_axis(&pl->ay, PH);
  if (tid == 32) fin[1] = rg_finish_axis(&pl->ax, PW);
  __syncthreads();
  return fin[0] && fin[1];
}

#ifndef RG_MINB
#define RG_MINB 4  // resident 256-thread CTAs per SM asked of the register allocator (<= 64 registers)
#endif
__global__ void __launch_bounds__(RT_THREADS, RG_MINB)
rg_pool_bwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   const float* __restrict__ dout, int c_per_cta, int buf_floats, int cc_cap,
                   int direct_fallback) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  RgAxes* pl = reinterpret_cast<RgAxes*>(rt_smem);
  float* d_s = reinterpret_cast<float*>(rt_smem + ((sizeof(RgAxes) + 15) / 16) * 16);
  __shared__ int fin[2];

  const int tid = threadIdx.x;
  const int k = blockIdx.x;
  const int c_begin = blockIdx.y * c_per_cta;
  const int c_end = min(C, c_begin + c_per_cta);
  if (c_begin >= c_end) return;
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  if (level < 0 || level >= lv.L || b < 0 || b >= N) return;
  const int H = lv.H[level], W = lv.W[level];
  const size_t HW = (size_t)H * W;
  const RtGeom g = rt_geom_bwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  const int banded = rg_build_axes(pl, fin, g, PH, PW, H, W);
  if (!banded && !direct_fallback) return;  // the sweep kernel (second launch) owns this RoI
  int cc_max = 0, exty = 0, extx = 0;
  if (banded) {
    exty = pl->ay.ext;
    extx = pl->ax.ext;
    if (exty == 0 || extx == 0) return;  // no sample inside the map: the gradient is zero
    cc_max = (buf_floats - 4) / (P + PH * extx);  // whole groups of RG_U channels (H is group-padded)
    if (cc_max > cc_cap) cc_max = cc_cap;
    cc_max &= ~(RG_U - 1);
  }
  if (cc_max < RG_U) {
    // no short-tap form (one-launch configuration) or tile too large for the buffer: rare,
    // reference-order scatter by the whole CTA
    const float* dk = dout + ((size_t)k * C + c_begin) * P;
    float* gb = lv.grad[level] + ((size_t)b * C + c_begin) * HW;
    const int total = (c_end - c_begin) * P;
    for (int i = tid; i < total; i += blockDim.x) {
      const int c = i / P;
      const int rem = i - c * P;
      const int ph = rem / PW, pw = rem - ph * PW;
      rt_direct_backward(gb + (size_t)c * HW, H, W, g, ph, pw, dk[i]);
    }
    return;
  }
  // covering bin ranges of every tile row / column
  for (int i = tid; i < exty + extx; i += blockDim.x) {
    const bool isy = i < exty;
    const RgAxis* a = isy ? &pl->ay : &pl->ax;
    const int e = isy ? i : i - exty, n = isy ? PH : PW;
    int fb = n, lb = -1;
    for (int p = 0; p < n; ++p) {
      const int st = a->start[p];
      if (st <= e && e < st + a->len[p]) {
        if (fb == n) fb = p;
        lb = p;
      }
    }
    (isy ? pl->covy : pl->covx)[e] = fb | (lb << 8);
  }
  const int count = g.gh * g.gw;
  const float inv = rt_exact_inv(count);
  const float fcount = (float)count;
  const float* wyf = &pl->ay.w[0][0];
  const float* wxf = &pl->ax.w[0][0];
  float* tile0 = lv.grad[level] + ((size_t)b * C + c_begin) * HW + (size_t)pl->ay.lo0 * W + pl->ax.lo0;
  const float* dout_k = dout + ((size_t)k * C + c_begin) * P;
  const int nA = PH * extx, nB = exty * extx;
  const uint32_t mx = rt_magic((uint32_t)extx);

  for (int c0 = 0; c0 < c_end - c_begin; c0 += cc_max) {
    const int cc = min(cc_max, c_end - c_begin - c0);
    float* Hs = d_s + ((cc_max * P + 3) & ~3);  // [group][RG_U/4 planes][PH][extx][4 channels]
    {
      const float* src = dout_k + (size_t)c0 * P;
      const int nfl = cc * P;
      if (((reinterpret_cast<uintptr_t>(src) & 15) == 0) && (nfl & 3) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(d_s);
        for (int i = tid; i < (nfl >> 2); i += blockDim.x) d4[i] = __ldg(s4 + i);
      } else {
        for (int i = tid; i < nfl; i += blockDim.x) d_s[i] = __ldg(src + i);
      }
    }
    __syncthreads();  // also orders the cov arrays before their first use
    const int ngrp = (cc + RG_U - 1) / RG_U;
    const int hgroup = PH * extx * RG_U;  // group pitch of Hs
    // ---- phase A ---------------------------------------------------------------------------
    for (int grp = 0; grp < ngrp; ++grp) {
      const int cb = grp * RG_U;
      const int nvalid = min(RG_U, cc - cb);
      for (int rem = tid; rem < nA; rem += blockDim.x) {
        const int ph = (int)rt_div((uint32_t)rem, (uint32_t)extx, mx);
        const int x = rem - ph * extx;
        const int cv = pl->covx[x];
        const int fb = cv & 255, lb = cv >> 8;
        float acc[RG_U];
#pragma unroll
        for (int u = 0; u < RG_U; ++u) acc[u] = 0.f;
        const float* dp = d_s + cb * P + ph * PW;
        if (nvalid == RG_U) {
          for (int p = fb; p <= lb; ++p) {
            const float w = wxf[p * RG_TAPS + (x - pl->ax.start[p])];
            const float* dq = dp + p;
#pragma unroll
            for (int u = 0; u < RG_U; ++u) acc[u] = fmaf(w, dq[u * P], acc[u]);
          }
        } else {
          for (int p = fb; p <= lb; ++p) {
            const float w = wxf[p * RG_TAPS + (x - pl->ax.start[p])];
#pragma unroll
            for (int u = 0; u < RG_U; ++u)
              if (u < nvalid) acc[u] = fmaf(w, dp[u * P + p], acc[u]);
          }
        }
        // channels past nvalid hold zeros: the group is stored whole (two 16-byte stores)
        // H is kept as RG_U/4 planes of 4 channels: consecutive lanes (x) are 16 bytes apart in a
        // plane, which keeps the 16-byte shared accesses conflict-free (32-byte strides are 2-way)
        float4* hp = reinterpret_cast<float4*>(Hs + grp * hgroup) + ph * extx + x;
#pragma unroll
        for (int v = 0; v < RG_U / 4; ++v)
          hp[v * nA] = make_float4(acc[4 * v], acc[4 * v + 1], acc[4 * v + 2], acc[4 * v + 3]);
      }
    }
    __syncthreads();
    // ---- phase B ---------------------------------------------------------------------------
    for (int grp = 0; grp < ngrp; ++grp) {
      const int cb = grp * RG_U;
      const int nvalid = min(RG_U, cc - cb);
      float* gbase = tile0 + (size_t)(c0 + cb) * HW;
      for (int rem = tid; rem < nB; rem += blockDim.x) {
        const int y = (int)rt_div((uint32_t)rem, (uint32_t)extx, mx);
        const int x = rem - y * extx;
        const int cv = pl->covy[y];
        const int fb = cv & 255, lb = cv >> 8;
        float acc[RG_U];
#pragma unroll
        for (int u = 0; u < RG_U; ++u) acc[u] = 0.f;
        const float4* hp = reinterpret_cast<const float4*>(Hs + grp * hgroup) + x;
        for (int q = fb; q <= lb; ++q) {
          const float w = wyf[q * RG_TAPS + (y - pl->ay.start[q])];
          const float4* hq = hp + q * extx;
#pragma unroll
          for (int v = 0; v < RG_U / 4; ++v) {
            const float4 h = hq[v * nA];
            acc[4 * v] = fmaf(w, h.x, acc[4 * v]);
            acc[4 * v + 1] = fmaf(w, h.y, acc[4 * v + 1]);
            acc[4 * v + 2] = fmaf(w, h.z, acc[4 * v + 2]);
            acc[4 * v + 3] = fmaf(w, h.w, acc[4 * v + 3]);
          }
        }
        if (fb > lb) continue;  // a tile row no bin touches carries no gradient
        float* gp = gbase + (size_t)y * W + x;
        // one RED per tile element, unconditionally: a per-element "!= 0" test costs a divergent
        // branch each, and zeros only arise from zero dout here
        if (nvalid == RG_U && inv != 0.f) {
#pragma unroll
          for (int u = 0; u < RG_U; ++u) {
            atomicAdd(gp, acc[u] * inv);
            gp += HW;
          }
        } else {
#pragma unroll
          for (int u = 0; u < RG_U; ++u)
            if (u < nvalid) atomicAdd(gp + (size_t)u * HW, (inv != 0.f) ? acc[u] * inv : acc[u] / fcount);
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// forward kernel, merged-tap form (v5): the same per-axis plan as the gather backward (RgAxis:
// every pooled bin reads a short run of consecutive tile rows with merged weights).  CTA = (RoI,
// channel slice); per chunk of cc channels (whole groups of RG_U):
//   stage   the RoI's tile [cc][exty][extx] global -> shared, 4-byte cp.async; a thread keeps its
//           tile positions and only steps the channel
//   y-pass  thread = (ph, x) x RG_U channels: V[ph][x][c8] = sum_j wy[ph][j] * T[c][sy+j][x]
//   x-pass  thread = (ph, pw) x RG_U channels: out = sum_j wx[pw][j] * V[ph][sx+j][c8] / count;
//           V is channel-minor (two 16-byte loads per tap), consecutive threads write consecutive
//           output floats of a channel plane (streaming stores).
// Weights, starts and lengths are fetched once for RG_U channels.  RoIs without a short-tap form
// (rare with a fixed sampling grid) are evaluated in reference order by the same CTA.
// ------------------------------------------------------------------------------------------
#ifndef RGF_MINB
#define RGF_MINB 4
#endif
template <bool DIRECT>  // DIRECT: no staged tile, the y-pass reads the feature map through L1
__global__ void __launch_bounds__(RT_THREADS, RGF_MINB)
rg_pool_fwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   float* __restrict__ out, int c_per_cta, int buf_floats, int cc_cap) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  RgAxes* pl = reinterpret_cast<RgAxes*>(rt_smem);
  float* T = reinterpret_cast<float*>(rt_smem + ((sizeof(RgAxes) + 15) / 16) * 16);
  __shared__ int fin[2];

  const int tid = threadIdx.x;
  const int k = blockIdx.x;
  const int c_begin = blockIdx.y * c_per_cta;
  const int c_end = min(C, c_begin + c_per_cta);
  if (c_begin >= c_end) return;
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  float* out_k = out + ((size_t)k * C + c_begin) * P;
  if (level < 0 || level >= lv.L || b < 0 || b >= N) {
    rt_zero_fill(out_k, (size_t)(c_end - c_begin) * P);
    return;
  }
  const int H = lv.H[level], W = lv.W[level];
  const size_t HW = (size_t)H * W;
  const float* feat_b = lv.feat[level] + ((size_t)b * C + c_begin) * HW;
  const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  const int banded = rg_build_axes(pl, fin, g, PH, PW, H, W);
  int cc_max = 0, exty = 0, extx = 0;
  if (banded) {
    exty = pl->ay.ext;
    extx = pl->ax.ext;
    if (exty == 0 || extx == 0) {  // every sample lies outside the map
      rt_zero_fill(out_k, (size_t)(c_end - c_begin) * P);
      return;
    }
    cc_max = (buf_floats - 4) / ((DIRECT ? 0 : exty * extx) + PH * extx);
    if (cc_max > cc_cap) cc_max = cc_cap;
    cc_max &= ~(RG_U - 1);
  }
  if (cc_max < RG_U) {  // no short-tap form, or the tile does not fit: reference order (rare)
    const int total = (c_end - c_begin) * P;
    for (int i = tid; i < total; i += blockDim.x) {
      const int c = i / P;
      const int rem = i - c * P;
      const int ph = rem / PW, pw = rem - ph * PW;
      out_k[i] = rt_direct_output(feat_b + (size_t)c * HW, H, W, g, ph, pw);
    }
    return;
  }
  const int count = g.gh * g.gw;
  const float inv = rt_exact_inv(count);
  const float fcount = (float)count;
  const float* wyf = &pl->ay.w[0][0];
  const float* wxf = &pl->ax.w[0][0];
  const int nT = exty * extx, nV = PH * extx;
  const uint32_t mx = rt_magic((uint32_t)extx), mpw = rt_magic((uint32_t)PW);
  const float* tile_g = feat_b + (size_t)pl->ay.lo0 * W + pl->ax.lo0;
  float* V = T + (DIRECT ? 0 : ((cc_max * nT + 3) & ~3));  // [group][RG_U/4 planes][PH][extx][4 channels]
  const int vgroup = nV * RG_U;

  for (int c0 = 0; c0 < c_end - c_begin; c0 += cc_max) {
    const int cc = min(cc_max, c_end - c_begin - c0);
    const int ngrp = (cc + RG_U - 1) / RG_U;
    // ---- stage -------------------------------------------------------------------------------
    if (!DIRECT) {
      for (int pos = tid; pos < nT; pos += blockDim.x) {
        const int y = (int)rt_div((uint32_t)pos, (uint32_t)extx, mx);
        const int x = pos - y * extx;
        const float* src = tile_g + (size_t)c0 * HW + (size_t)y * W + x;
        uint32_t dst = rp_smem_addr(T + pos);
        for (int c = 0; c < cc; ++c) {
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
          src += HW;
          dst += (uint32_t)nT * 4u;
        }
      }
      if (cc < ngrp * RG_U)  // the last group's missing channels read as zeros
        for (int i = cc * nT + tid; i < ngrp * RG_U * nT; i += blockDim.x) T[i] = 0.f;
      rt_cp_async_wait_all();
      __syncthreads();
    }
    // ---- y-pass ------------------------------------------------------------------------------
    for (int grp = 0; grp < ngrp; ++grp) {
      for (int rem = tid; rem < nV; rem += blockDim.x) {
        const int ph = (int)rt_div((uint32_t)rem, (uint32_t)extx, mx);
        const int x = rem - ph * extx;
        const int ly = pl->ay.len[ph];
        const float* wp = wyf + ph * RG_TAPS;
        float acc[RG_U];
#pragma unroll
        for (int u = 0; u < RG_U; ++u) acc[u] = 0.f;
        if (DIRECT) {
          const int cb = grp * RG_U;
          const int nvalid = min(RG_U, cc - cb);
          const float* gp = tile_g + (size_t)(c0 + cb) * HW + (size_t)pl->ay.start[ph] * W + x;
          for (int j = 0; j < ly; ++j) {
            const float w = wp[j];
            const float* gq = gp + (size_t)j * W;
#pragma unroll
            for (int u = 0; u < RG_U; ++u)
              if (u < nvalid) acc[u] = fmaf(w, __ldg(gq + (size_t)u * HW), acc[u]);
          }
        } else {
          const float* tp = T + (grp * RG_U * exty + pl->ay.start[ph]) * extx + x;
          for (int j = 0; j < ly; ++j) {
            const float w = wp[j];
            const float* tq = tp + j * extx;
#pragma unroll
            for (int u = 0; u < RG_U; ++u) acc[u] = fmaf(w, tq[u * nT], acc[u]);
          }
        }
        float4* vp = reinterpret_cast<float4*>(V + grp * vgroup) + rem;  // planes of 4 channels
#pragma unroll
        for (int v = 0; v < RG_U / 4; ++v)
          vp[v * nV] = make_float4(acc[4 * v], acc[4 * v + 1], acc[4 * v + 2], acc[4 * v + 3]);
      }
    }
    __syncthreads();
    // ---- x-pass ------------------------------------------------------------------------------
    for (int grp = 0; grp < ngrp; ++grp) {
      const int cb = grp * RG_U;
      const int nvalid = min(RG_U, cc - cb);
      float* ob = out_k + (size_t)(c0 + cb) * P;
      for (int i = tid; i < P; i += blockDim.x) {
        const int ph = (int)rt_div((uint32_t)i, (uint32_t)PW, mpw);
        const int pw = i - ph * PW;
        const int lx = pl->ax.len[pw];
        const float4* vq = reinterpret_cast<const float4*>(V + grp * vgroup) + ph * extx + pl->ax.start[pw];
        const float* wp = wxf + pw * RG_TAPS;
        float acc[RG_U];
#pragma unroll
        for (int u = 0; u < RG_U; ++u) acc[u] = 0.f;
        for (int j = 0; j < lx; ++j) {
          const float w = wp[j];
#pragma unroll
          for (int v = 0; v < RG_U / 4; ++v) {
            const float4 h = vq[v * nV + j];
            acc[4 * v] = fmaf(w, h.x, acc[4 * v]);
            acc[4 * v + 1] = fmaf(w, h.y, acc[4 * v + 1]);
            acc[4 * v + 2] = fmaf(w, h.z, acc[4 * v + 2]);
            acc[4 * v + 3] = fmaf(w, h.w, acc[4 * v + 3]);
          }
        }
        float* op = ob + i;
        if (nvalid == RG_U && inv != 0.f) {
#pragma unroll
          for (int u = 0; u < RG_U; ++u) {
            __stcs(op, acc[u] * inv);
            op += P;
          }
        } else {
#pragma unroll
          for (int u = 0; u < RG_U; ++u)
            if (u < nvalid) __stcs(op + (size_t)u * P, (inv != 0.f) ? acc[u] * inv : acc[u] / fcount);
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// backward kernel (transpose of the forward sweeps)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RT_THREADS)
rt_pool_bwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   const float* __restrict__ dout, int c_per_cta, int vs_floats, int cc_cap,
                   int skip_banded) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  __shared__ RgAxes rg_axes_s;
  __shared__ int rg_fin[2];
  RtAxisPlan* py = reinterpret_cast<RtAxisPlan*>(rt_smem);
  RtAxisPlan* px = py + 1;
  float* Hs = reinterpret_cast<float*>(rt_smem + ((2 * sizeof(RtAxisPlan) + 15) / 16) * 16);
  float* d_s = Hs + vs_floats;
  __shared__ int s_ok[2];

  const int tid = threadIdx.x;
  const int k = blockIdx.x;
  const int c_begin = blockIdx.y * c_per_cta;
  const int c_end = min(C, c_begin + c_per_cta);
  if (c_begin >= c_end) return;
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  if (level < 0 || level >= lv.L || b < 0 || b >= N) return;
  const int H = lv.H[level], W = lv.W[level];
  const size_t HW = (size_t)H * W;
  float* grad_b = lv.grad[level] + ((size_t)b * C + c_begin) * HW;
  const float* dout_k = dout + ((size_t)k * C + c_begin) * P;
  const RtGeom g = rt_geom_bwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  // second launch of the default path: only the RoIs the gather kernel returned from
  if (skip_banded && rg_build_axes(&rg_axes_s, rg_fin, g, PH, PW, H, W)) return;

  if (tid == 0) s_ok[0] = rt_build_axis_plan(py, g.sh, g.bh, PH, g.gh, H);
  if (tid == 32) s_ok[1] = rt_build_axis_plan(px, g.sw, g.bw, PW, g.gw, W);
  __syncthreads();

  const int tiled = s_ok[0] && s_ok[1];
  const int R = py->n, tw = px->n;
  if (tiled && (R == 0 || tw == 0)) return;
  const int twp = tw | 1;
  int cc_max = tiled ? (vs_floats / (PH * twp)) : 0;
  if (cc_max > cc_cap) cc_max = cc_cap;
  cc_max &= ~3;
  if (cc_max < 4) {
    const int total = (c_end - c_begin) * P;
    for (int i = tid; i < total; i += blockDim.x) {
      const int c = i / P;
      const int rem = i - c * P;
      const int ph = rem / PW, pw = rem - ph * PW;
      rt_direct_backward(grad_b + (size_t)c * HW, H, W, g, ph, pw, dout_k[i]);
    }
    return;
  }

  const int count = g.gh * g.gw;
  const float inv = rt_exact_inv(count);
  float* tile0 = grad_b + (size_t)py->lo0 * W + px->lo0;

  for (int c0 = 0; c0 < c_end - c_begin; c0 += cc_max) {
    const int cc = min(cc_max, c_end - c_begin - c0);
    // ---- stage the chunk's dout (contiguous) --------------------------------------------
    {
      const float* src = dout_k + (size_t)c0 * P;
      const int nfl = cc * P;
      for (int i = tid; i < nfl; i += blockDim.x) d_s[i] = __ldg(src + i);
    }
    __syncthreads();
    // ---- phase A: transpose of the column sweep: dout[c][ph][:] -> Hs[c][ph][x] ------------
    {
      const int nq = (cc + RT_U2 - 1) / RT_U2;
      const int items = nq * PH;
      for (int it = tid; it < items; it += blockDim.x) {
        const int q = it / PH;
        const int ph = it - q * PH;
        int nvalid = (cc - q + nq - 1) / nq;
        if (nvalid > RT_U2) nvalid = RT_U2;
        BwdDoutSource<RT_U2> src{d_s + (size_t)(q * PH + ph) * PW, nq * PH * PW, nvalid};
        BwdHEmit<RT_U2> emit{Hs + (size_t)(q * PH + ph) * twp, nq * PH * twp, nvalid};
        rt_sweep_bwd<RT_U2>(*px, src, emit);
      }
    }
    __syncthreads();
    // ---- phase B: transpose of the row sweep: Hs[c][:][x] -> grad tile, one RED / element ---
    {
      const int nq = (cc + RT_U1 - 1) / RT_U1;
      const int items = nq * tw;
      for (int it = tid; it < items; it += blockDim.x) {
        const int q = it / tw;
        const int x = it - q * tw;
        int nvalid = (cc - q + nq - 1) / nq;
        if (nvalid > RT_U1) nvalid = RT_U1;
        BwdHSource<RT_U1> src{Hs + (size_t)q * PH * twp + x, twp, nq * PH * twp, nvalid};
        BwdGradEmit<RT_U1> emit{tile0 + (size_t)(c0 + q) * HW + x, (size_t)nq * HW, W, nvalid,
                                inv, (float)count};
        rt_sweep_bwd<RT_U1>(*py, src, emit);
      }
    }
    __syncthreads();
  }
}

// mode 1: the reference's own scheme (one thread per output element, 4 atomics per sample).
__global__ void __launch_bounds__(256)
rt_pool_bwd_exact_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                         const int* __restrict__ levels, size_t total, int PH, int PW,
                         float sampling_ratio, const float* __restrict__ dout) {
  const int P = PH * PW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int pw = (int)(i % PW);
    const int ph = (int)((i / PW) % PH);
    const int c = (int)((i / P) % C);
    const int k = (int)(i / ((size_t)P * C));
    const float* roi = rois + (size_t)k * 5;
    const int level = levels ? levels[k] : 0;
    const int b = (int)roi[0];
    if (level < 0 || level >= lv.L || b < 0 || b >= N) continue;
    const int H = lv.H[level], W = lv.W[level];
    const RtGeom g = rt_geom_bwd(roi, lv.scale[level], sampling_ratio, PH, PW);
    rt_direct_backward(lv.grad[level] + ((size_t)b * C + c) * H * W, H, W, g, ph, pw, dout[i]);
  }
}

// ------------------------------------------------------------------------------------------
// level mapping
// ------------------------------------------------------------------------------------------
__global__ void rt_level_map_kernel(const float* __restrict__ rois, int K, int k_min, int k_max,
                                    float s0, int lvl0, float eps, int* __restrict__ levels) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const float* b = rois + (size_t)k * 5 + 1;
  levels[k] = rp_fpn_level(b[0], b[1], b[2], b[3], k_min, k_max, s0, lvl0, eps);
}

__global__ void rt_level_map_area_kernel(const float* __restrict__ rois,
                                         const float* __restrict__ img_area, int K, int k_min,
                                         int k_max, float eps, int* __restrict__ levels) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const float* b = rois + (size_t)k * 5 + 1;
  levels[k] = rp_fpn_level_area(b[0], b[1], b[2], b[3], img_area[k], k_min, k_max, eps);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct RtLaunch {
  int csplit, c_per_cta, vs_floats, tile_floats, U, bwd_cc, bwd_threads;
  int gpb, gw, threads_fwd;  // forward: thread groups per CTA, lanes per group (>= PW), CTA size
  size_t smem, smem_fwd;
};

RtLaunch rt_plan_launch(int K, int C, int PH, int PW) {
  RtLaunch l;
  // backward: H buffer, enough for RT_CC_CAP channels of a 27-wide tile at 14 bins
  // measured (tools/sweep_bwd.py, 16 images x 512 RoIs): 7x7 is fastest with 128 threads and a 16 KB
  // H buffer (2.12 vs 2.68 ms), 14x14 with 256 threads and 16 channels per chunk (1.59 vs 1.69 ms)
  const bool small_bins = PH * PW <= 64;
  l.vs_floats = rp_env_int("RP_ROI_VS_FLOATS", small_bins ? 4096 : 8192) & ~3;
  const int min_vs = PH * 3 * 4;  // at least 4 channels of a 2-wide tile
  if (l.vs_floats < min_vs) l.vs_floats = min_vs;
  // forward: raw feature tile buffer (48 KB => 4 resident CTAs per SM)
  l.tile_floats = rp_env_int("RP_ROI_TILE_FLOATS", 12288) & ~3;
  if (l.tile_floats < 256) l.tile_floats = 256;
  l.U = rp_env_int("RP_ROI_U", 2);
  // forward CTA: gpb groups of gw lanes; gw = PW rounded up to a power of two when that is < 32
  // (whole groups per warp => the groups of a warp can be kept on disjoint banks), gpb the
  // largest power of two that fits 256 threads
  l.gw = PW;
  if (PW < 32) {
    l.gw = 1;
    while (l.gw < PW) l.gw *= 2;
  }
  l.gpb = 1;
  while (l.gw * l.gpb * 2 <= RT3_MAX_THREADS) l.gpb *= 2;
  l.threads_fwd = (l.gw * l.gpb + 31) / 32 * 32;
  // channel split: the slow grid dimension walks channel slices, so that the feature planes
  // of one slice (C/csplit of every image) are what competes for L2 at any one time
  int cpc = rp_env_int("RP_ROI_CPC", 128);
  if (cpc < 4) cpc = 4;
  if (cpc > C) cpc = C;
  cpc = (cpc + 3) & ~3;
  l.c_per_cta = cpc;
  l.csplit = (C + cpc - 1) / cpc;
  l.bwd_cc = rp_env_int("RP_ROI_BWD_CC", small_bins ? RT_CC_CAP : 16);
  if (l.bwd_cc < 4) l.bwd_cc = 4;
  if (l.bwd_cc > RT_CC_CAP) l.bwd_cc = RT_CC_CAP;
  l.bwd_cc &= ~3;
  l.bwd_threads = rp_env_int("RP_ROI_BWD_THREADS", small_bins ? 128 : RT_THREADS);
  if (l.bwd_threads < 64) l.bwd_threads = 64;
  if (l.bwd_threads > RT_THREADS) l.bwd_threads = RT_THREADS;
  l.bwd_threads &= ~31;
  l.smem = rp_align_up(2 * sizeof(RtAxisPlan), 16) + sizeof(float) * (size_t)l.vs_floats +
           sizeof(float) * (size_t)l.bwd_cc * PH * PW;
  l.smem_fwd = sizeof(RtYSample) * RT_SMAX + sizeof(RtW) * RT_SMAX + sizeof(short) * RT_SMAX +
               sizeof(float) * (size_t)l.tile_floats;
  return l;
}

// ---- TMA descriptors: encoded on the host once per (base pointer, planes, H, W), cached ---------
typedef CUresult (*rb_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

rb_encode_fn rb_encoder() {
  static rb_encode_fn fn = []() -> rb_encode_fn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    return reinterpret_cast<rb_encode_fn>(p);
  }();
  return fn;
}

struct RbMapEntry {
  const float* ptr;
  long long planes;
  int H, W, ok;
  CUtensorMap m[RB_TMA_NQ];
};

// Fills maps->m[l] / ok[l] for level l; a small cache keeps repeated calls free of encoding.
void rb_level_maps(RbMaps* maps, int l, const float* ptr, long long planes, int H, int W) {
  static std::mutex mu;
  static RbMapEntry cache[16];
  static int next = 0;
  maps->ok[l] = 0;
  if (!ptr || (W & 3) || (reinterpret_cast<uintptr_t>(ptr) & 15) || planes < RB_BOXC ||
      planes > 0x7fffffffLL)
    return;
  rb_encode_fn enc = rb_encoder();
  if (!enc) return;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < 16; ++i) {
    const RbMapEntry& e = cache[i];
    if (e.ptr == ptr && e.planes == planes && e.H == H && e.W == W) {
      if (e.ok) {
        memcpy(maps->m[l], e.m, sizeof(e.m));
        maps->ok[l] = 1;
      }
      return;
    }
  }
  RbMapEntry& e = cache[next];
  next = (next + 1) & 15;
  e.ptr = ptr;
  e.planes = planes;
  e.H = H;
  e.W = W;
  e.ok = 1;
  const cuuint64_t gdim[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
  const cuuint64_t gstride[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};
  const cuuint32_t estride[3] = {1, 1, 1};
  for (int nq = 1; nq <= RB_TMA_NQ && e.ok; ++nq) {
    const cuuint32_t box[3] = {(cuuint32_t)(4 * nq), 1, (cuuint32_t)RB_BOXC};
    const CUresult r = enc(&e.m[nq - 1], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr),
                           gdim, gstride, box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) e.ok = 0;
  }
  if (e.ok) {
    memcpy(maps->m[l], e.m, sizeof(e.m));
    maps->ok[l] = 1;
  }
}

// Descriptors of the chained-MMA kernel (roi_fwd_mma.cuh RmMaps): per level and (row pitch class, row
// blocks) one box pitch x 8*KY x 1 plane.  Cached per (base pointer, planes, H, W).
struct RmMapEntry {
  const float* ptr;
  long long planes;
  int H, W, ok;
  CUtensorMap m[RM_NPITCH][RM_KMAX];
};

void rm_level_maps(RmMaps* maps, int l, const float* ptr, long long planes, int H, int W) {
  static std::mutex mu;
  static RmMapEntry cache[16];
  static int next = 0;
  maps->ok[l] = 0;
  if (!ptr || (W & 3) || (reinterpret_cast<uintptr_t>(ptr) & 15) || planes < 1 || planes > 0x7fffffffLL)
    return;
  rb_encode_fn enc = rb_encoder();
  if (!enc) return;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < 16; ++i) {
    const RmMapEntry& e = cache[i];
    if (e.ptr == ptr && e.planes == planes && e.H == H && e.W == W) {
      if (e.ok) {
        memcpy(maps->m[l], e.m, sizeof(e.m));
        maps->ok[l] = 1;
      }
      return;
    }
  }
  RmMapEntry& e = cache[next];
  next = (next + 1) & 15;
  e.ptr = ptr;
  e.planes = planes;
  e.H = H;
  e.W = W;
  e.ok = 1;
  const cuuint64_t gdim[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
  const cuuint64_t gstride[2] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4};
  const cuuint32_t estride[3] = {1, 1, 1};
  static const int pitches[RM_NPITCH] = {8, 24, 40, 56};
  for (int pc = 0; pc < RM_NPITCH && e.ok; ++pc)
    for (int ky = 1; ky <= RM_KMAX && e.ok; ++ky) {
      const cuuint32_t box[3] = {(cuuint32_t)pitches[pc], (cuuint32_t)(8 * ky), 1};
      const CUresult r = enc(&e.m[pc][ky - 1], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr),
                             gdim, gstride, box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) e.ok = 0;
    }
  if (e.ok) {
    memcpy(maps->m[l], e.m, sizeof(e.m));
    maps->ok[l] = 1;
  }
}

// Stream-ordered scratch of the forward call from a pool of this library's own (one per device) that
// keeps its memory across synchronisations: the default pool trims to zero at every sync, which turned
// the per-call allocation into a 0.1 ms driver call in the bench's step loop.
cudaError_t rt_scratch_alloc(void** p, size_t bytes, cudaStream_t st) {
  static std::mutex mu;
  static cudaMemPool_t pools[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaMallocAsync(p, bytes, st);
  cudaMemPool_t pool;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!pools[dev]) {
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      e = cudaMemPoolCreate(&pools[dev], &props);
      if (e != cudaSuccess) {
        pools[dev] = nullptr;
        (void)cudaGetLastError();
        return cudaMallocAsync(p, bytes, st);
      }
      unsigned long long keep = ~0ull;
      (void)cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
    }
    pool = pools[dev];
  }
  return cudaMallocFromPoolAsync(p, bytes, pool, st);
}

template <int U>
int rt_launch_fwd(const RtLaunch& l, const RtLevels& lv, int N, int C, const float* rois,
                  const int* levels, int K, int PH, int PW, float sampling_ratio, float* out,
                  cudaStream_t st, const int* skip = nullptr, const RpTab* tabs = nullptr) {
  dim3 grid((unsigned)K, (unsigned)l.csplit);
  const int nbuf = (rp_env_int("RP_ROI_DBUF", 0) ? 2 : 1)     // two half-size tile buffers: copies under the walk
                   | (rp_env_int("RP_ROI_TAP3", RP_ROI_TAP3_DEFAULT) ? 16 : 0)   // three-tap rows
                   | (rp_env_int("RP_ROI_YMERGE", RP_ROI_YMERGE_DEFAULT) ? 32 : 0);  // one entry per bin where both y-samples share a row pair
  if (rp_env_int("RP_ROI_LEAN", 1)) {  // hand-written shared-memory addressing in the sample walk
    RP_SET_MAX_SMEM((rt_pool_fwd_kernel<U, true>), l.smem_fwd);
    rt_pool_fwd_kernel<U, true><<<grid, l.threads_fwd, l.smem_fwd, st>>>(
        lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, l.c_per_cta, l.tile_floats, l.gpb, l.gw, nbuf, skip, tabs);
  } else {
    RP_SET_MAX_SMEM((rt_pool_fwd_kernel<U, false>), l.smem_fwd);
    rt_pool_fwd_kernel<U, false><<<grid, l.threads_fwd, l.smem_fwd, st>>>(
        lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, l.c_per_cta, l.tile_floats, l.gpb, l.gw, nbuf, skip, tabs);
  }
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rt_fill_levels(RtLevels* lv, const float* const* feats, float* const* grads, const int* H,
                   const int* W, const float* scales, int L) {
  if (L < 1 || L > RP_MAX_LEVELS || !H || !W || !scales) return RP_E_ARG;
  lv->L = L;
  for (int l = 0; l < RP_MAX_LEVELS; ++l) {
    lv->feat[l] = nullptr;
    lv->grad[l] = nullptr;
    lv->H[l] = lv->W[l] = 0;
    lv->scale[l] = 0.f;
  }
  for (int l = 0; l < L; ++l) {
    if (H[l] < 1 || W[l] < 1) return RP_E_ARG;
    if (feats) {
      if (!feats[l]) return RP_E_ARG;
      lv->feat[l] = feats[l];
    }
    if (grads) {
      if (!grads[l]) return RP_E_ARG;
      lv->grad[l] = grads[l];
    }
    lv->H[l] = H[l];
    lv->W[l] = W[l];
    lv->scale[l] = scales[l];
  }
  return RP_OK;
}

int rt_forward(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K,
               int PH, int PW, float sampling_ratio, float* out, int mode, cudaStream_t st) {
  if (K == 0) return RP_OK;
  if (!rois || !out || N < 1 || C < 1 || K < 0 || PH < 1 || PW < 1) return RP_E_ARG;
  if ((size_t)PH * PW > 4096) return RP_E_LIMIT;
  if (mode == 1 || (mode == 0 && PW > RT3_MAX_THREADS)) {
    const size_t total = (size_t)K * C * PH * PW;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)RP_NUM_SMS * 32) blocks = (size_t)RP_NUM_SMS * 32;
    rt_pool_fwd_exact_kernel<<<(unsigned)blocks, 256, 0, st>>>(lv, N, C, rois, levels, total, PH,
                                                               PW, sampling_ratio, out);
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  if (mode != 0) return RP_E_ARG;
  if (rp_env_int("RP_ROI_IMPL", RP_ROI_IMPL_DEFAULT) == 6) {
    // tensor-core kernel (roi_fwd_tc.cu); 1 = configuration not eligible, use the SIMT kernels
    const int rc = rx_forward(lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, st);
    if (rc != 1) return rc;
  }
  const RtLaunch l = rt_plan_launch(K, C, PH, PW);
  if (rp_env_int("RP_ROI_IMPL", RP_ROI_IMPL_DEFAULT) == 5 && sampling_ratio > 0 &&
      (int)sampling_ratio <= RB_MAXG && PH <= RB_MAXP && PW <= RB_MAXP) {
    // merged-tap kernel (fixed sampling grid); RoIs without a short-tap form: reference order
    int threads = rp_env_int("RP_ROI_G_THREADS", 256) & ~31;
    if (threads < 64) threads = 64;
    if (threads > RT_THREADS) threads = RT_THREADS;
    int gfloats = rp_env_int("RP_ROI_G_FLOATS", 12288) & ~3;
    if (gfloats < 2048) gfloats = 2048;
    int gcc = rp_env_int("RP_ROI_G_CC", 32) & ~(RG_U - 1);
    if (gcc < RG_U) gcc = RG_U;
    const size_t gsmem = rp_align_up(sizeof(RgAxes), 16) + sizeof(float) * (size_t)gfloats;
    if (gsmem > 227 * 1024) return RP_E_LIMIT;
    dim3 grid((unsigned)K, (unsigned)l.csplit);
    if (rp_env_int("RP_ROI_G_DIRECT", 0)) {
      RP_SET_MAX_SMEM(rg_pool_fwd_kernel<true>, gsmem);
      rg_pool_fwd_kernel<true><<<grid, threads, gsmem, st>>>(lv, N, C, rois, levels, K, PH, PW,
                                                             sampling_ratio, out, l.c_per_cta, gfloats, gcc);
    } else {
      RP_SET_MAX_SMEM(rg_pool_fwd_kernel<false>, gsmem);
      rg_pool_fwd_kernel<false><<<grid, threads, gsmem, st>>>(lv, N, C, rois, levels, K, PH, PW,
                                                              sampling_ratio, out, l.c_per_cta, gfloats, gcc);
    }
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  if (rp_env_int("RP_ROI_IMPL", RP_ROI_IMPL_DEFAULT) == 4 && PW <= 128) {
    // banded kernel; RoIs without a band form run the v3 body inside the same CTA
    const int cfg = rp_env_int("RP_ROI_BAND_CFG", 0);  // 0: 128 thr x5, 1: 128 x8, 2: 256 x4, 3: 256 x3
    const int nt = cfg >= 2 ? 256 : 128;
    int gw = PW, gpb = 1;
    if (PW < 32) {
      gw = 1;
      while (gw < PW) gw *= 2;
    }
    while (gw * gpb * 2 <= nt) gpb *= 2;
    static const int dflt_kb[4] = {40, 24, 52, 70};
    int kb = rp_env_int("RP_ROI_BAND_SMEM_KB", dflt_kb[cfg & 3]);
    if (kb < 16) kb = 16;
    if (kb > 200) kb = 200;
    const int smem = kb * 1024;
    const int nbuf = rp_env_int("RP_ROI_BAND_NBUF", 1) == 2 ? 2 : 1;  // double buffering measured slower (smaller chunks)
    dim3 grid((unsigned)K, (unsigned)l.csplit);
    RbMaps maps;
    memset(&maps, 0, sizeof(maps));
    if (rp_env_int("RP_ROI_TMA", 0))  // opt-in: single-buffered TMA staging measured slower (2.44 vs 2.29 ms)
      for (int i = 0; i < lv.L && i < RB_TMA_LEVELS; ++i)
        rb_level_maps(&maps, i, lv.feat[i], (long long)N * C, lv.H[i], lv.W[i]);
#define RB_LAUNCH(NT_, MB_)                                                                      \
  do {                                                                                           \
    RP_SET_MAX_SMEM((rb_pool_fwd_kernel<NT_, MB_>), smem);                                       \
    rb_pool_fwd_kernel<NT_, MB_><<<grid, NT_, smem, st>>>(lv, N, C, rois, levels, K, PH, PW,     \
                                                          sampling_ratio, out, l.c_per_cta,      \
                                                          (smem - 128) / 4, gpb, gw, nbuf, maps); \
  } while (0)
    switch (cfg & 3) {
      case 1: RB_LAUNCH(128, 8); break;
      case 2: RB_LAUNCH(256, 4); break;
      case 3: RB_LAUNCH(256, 3); break;
      default: RB_LAUNCH(128, 5); break;
    }
#undef RB_LAUNCH
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  if (l.smem_fwd > 227 * 1024) return RP_E_LIMIT;
  if (rp_env_int("RP_ROI_IMPL", RP_ROI_IMPL_DEFAULT) == 7 && PH <= 16 && PW <= 16 &&
      l.threads_fwd == RM_THREADS) {
    // chained-MMA kernel (roi_fwd_mma.cuh); tiles it does not take run the tile-walk body in the same CTA
    const size_t smem = RM_SMEM_BYTES > l.smem_fwd ? RM_SMEM_BYTES : l.smem_fwd;
    int cpc = rp_env_int("RP_ROI_MMA_CPC", 256);  // measured: 256 is 8 % faster than 128 (set-up is ~3 us per CTA)
    if (cpc < 8) cpc = 8;
    if (cpc > C) cpc = C;
    dim3 grid((unsigned)K, (unsigned)((C + cpc - 1) / cpc));
    RmMaps maps;
    memset(&maps, 0, sizeof(maps));
    if (rp_env_int("RP_ROI_MMA_TMA", 1))  // levels with 16-byte rows are staged by TMA boxes, the others by cp.async
      for (int i = 0; i < lv.L && i < RM_TMA_LEVELS; ++i)
        rm_level_maps(&maps, i, lv.feat[i], (long long)N * C, lv.H[i], lv.W[i]);
    if (rp_env_int("RP_ROI_MMA_MODE", 0) == 0) {  // 0: bf16 correction product, 1: 3xTF32
      RP_SET_MAX_SMEM(rm_pool_fwd_kernel<0>, smem);
      rm_pool_fwd_kernel<0><<<grid, RM_THREADS, smem, st>>>(lv, N, C, rois, levels, K, PH, PW, sampling_ratio,
                                                             out, cpc, l.tile_floats, l.gpb, l.gw, rp_env_int("RP_ROI_MMA_ABLATE", 0), maps);
#ifdef RM_WITH_MODE2  // build with RP_EXTRA_NVCC=-DRM_WITH_MODE2: single tf32 product (~1e-3), measurement only
    } else if (rp_env_int("RP_ROI_MMA_MODE", 0) == 2) {
      RP_SET_MAX_SMEM(rm_pool_fwd_kernel<2>, smem);
      rm_pool_fwd_kernel<2><<<grid, RM_THREADS, smem, st>>>(lv, N, C, rois, levels, K, PH, PW, sampling_ratio,
                                                             out, cpc, l.tile_floats, l.gpb, l.gw, rp_env_int("RP_ROI_MMA_ABLATE", 0), maps);
#endif
    } else {
      RP_SET_MAX_SMEM(rm_pool_fwd_kernel<1>, smem);
      rm_pool_fwd_kernel<1><<<grid, RM_THREADS, smem, st>>>(lv, N, C, rois, levels, K, PH, PW, sampling_ratio,
                                                             out, cpc, l.tile_floats, l.gpb, l.gw, rp_env_int("RP_ROI_MMA_ABLATE", 0), maps);
    }
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  // Whole-plane kernel for the coarse levels (roi_fwd_plane.cuh): their RoIs leave the column walk.
  const int* skip = nullptr;
  const RpTab* wtabs = nullptr;  // sample tables the column walk reads instead of building its own
  void* scratch = nullptr;
  if (rp_env_int("RP_ROI_PLANES", RP_ROI_PLANES_DEFAULT) && sampling_ratio == 2.0f && PH <= 16 && PW <= 16 &&
      K <= 262144) {  // 848 bytes of tables per RoI in stream-ordered scratch
    int max_groups = 0;
    for (int i = 0; i < lv.L; ++i) {
      const long long hw = (long long)lv.H[i] * lv.W[i];
      if (lv.H[i] < 2 || lv.W[i] < 2 || hw > RP_PLANE_MAX) continue;
      int cg = (int)(RP_PLANE_FLOATS / hw);
      if (cg > C) cg = C;
      cg &= ~1;
      if (cg < 2) cg = C >= 2 ? 2 : 1;
      const int groups = (C + cg - 1) / cg;
      if (groups > max_groups) max_groups = groups;
    }
    if (max_groups > 0 && (long long)N * max_groups <= 0x7fffffffLL) {
      const size_t tb = rp_align_up(sizeof(RpTab) * (size_t)K, 256);
      const size_t kb = rp_align_up(sizeof(int) * (size_t)K, 256);
      const int nkeys = lv.L * N;
      const size_t nwin = ((size_t)K + RP_LIST - 1) / RP_LIST;
      const size_t wb = (nwin * nkeys <= (1u << 20)) ? sizeof(int) * nwin * nkeys : 0;  // window counts (optional)
      if (rt_scratch_alloc(&scratch, tb + kb + wb, st) != cudaSuccess) {
        (void)cudaGetLastError();
        scratch = nullptr;  // no scratch: the column walk takes every RoI
      } else {
        RpTab* tabs = reinterpret_cast<RpTab*>(scratch);
        int* key = reinterpret_cast<int*>(reinterpret_cast<char*>(scratch) + tb);
        int* wcnt = wb ? reinterpret_cast<int*>(reinterpret_cast<char*>(scratch) + tb + kb) : nullptr;
        if (wcnt && cudaMemsetAsync(wcnt, 0, wb, st) != cudaSuccess) {
          (void)cudaGetLastError();
          wcnt = nullptr;  // without the counts the plane kernel visits every window
        }
        rp_tables_kernel<<<(unsigned)((K + 7) / 8), 256, 0, st>>>(lv, N, rois, levels, K, PH, PW, sampling_ratio,
                                                                  tabs, key, rp_env_int("RP_ROI_YMERGE", RP_ROI_YMERGE_DEFAULT),
                                                                  wcnt, nkeys);
        RP_LAUNCH_CHECK();
        RP_SET_MAX_SMEM(rp_plane_fwd_kernel, RP_SMEM_BYTES);
        // enough CTAs for two waves of the 2 x 148 resident ones: with few images the RoIs of one
        // (level, image) are dealt to up to 8 CTAs
        int slices = 1;
        while (slices < 8 && (long long)N * max_groups * slices < 4 * RP_NUM_SMS) slices *= 2;
        rp_plane_fwd_kernel<<<dim3((unsigned)(N * max_groups), (unsigned)lv.L, (unsigned)slices), RP_THREADS, RP_SMEM_BYTES, st>>>(
            lv, N, C, key, tabs, K, PH, PW, out, max_groups, rp_env_int("RP_ROI_TAP3", RP_ROI_TAP3_DEFAULT), wcnt, nkeys);
        RP_LAUNCH_CHECK();
        skip = key;
        if (rp_env_int("RP_ROI_WALK_TABS", RP_ROI_WALK_TABS_DEFAULT)) wtabs = tabs;
      }
    }
  }
  int rc;
  switch (l.U) {
    case 1: rc = rt_launch_fwd<1>(l, lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, st, skip, wtabs); break;
    case 4: rc = rt_launch_fwd<4>(l, lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, st, skip, wtabs); break;
    default: rc = rt_launch_fwd<2>(l, lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, st, skip, wtabs); break;
  }
  if (scratch) (void)cudaFreeAsync(scratch, st);
  return rc;
}

int rt_backward(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K,
                int PH, int PW, float sampling_ratio, const float* dout, int mode,
                cudaStream_t st) {
  for (int l = 0; l < lv.L; ++l)
    RP_CUDA_TRY(cudaMemsetAsync(lv.grad[l], 0, sizeof(float) * (size_t)N * C * lv.H[l] * lv.W[l], st));
  if (K == 0) return RP_OK;
  if (!rois || !dout || N < 1 || C < 1 || K < 0 || PH < 1 || PW < 1) return RP_E_ARG;
  if ((size_t)PH * PW > 4096) return RP_E_LIMIT;
  if (mode == 1) {
    const size_t total = (size_t)K * C * PH * PW;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)RP_NUM_SMS * 32) blocks = (size_t)RP_NUM_SMS * 32;
    rt_pool_bwd_exact_kernel<<<(unsigned)blocks, 256, 0, st>>>(lv, N, C, rois, levels, total, PH,
                                                               PW, sampling_ratio, dout);
    RP_LAUNCH_CHECK();
    return RP_OK;
  }
  if (mode != 0) return RP_E_ARG;
  const RtLaunch l = rt_plan_launch(K, C, PH, PW);
  if (l.smem > 227 * 1024) return RP_E_LIMIT;
  dim3 grid((unsigned)K, (unsigned)l.csplit);
  const int gather = rp_env_int("RP_ROI_BWD_IMPL", 2) == 2 && PH <= RB_MAXP && PW <= RB_MAXP;
  if (gather) {
    int gthreads = rp_env_int("RP_ROI_BWD_G_THREADS", 256) & ~31;
    if (gthreads < 64) gthreads = 64;
    if (gthreads > RT_THREADS) gthreads = RT_THREADS;
    int gfloats = rp_env_int("RP_ROI_BWD_G_FLOATS", 8192) & ~3;
    if (gfloats < 1024) gfloats = 1024;
    int gcc = rp_env_int("RP_ROI_BWD_G_CC", 32) & ~3;
    if (gcc < 4) gcc = 4;
    const size_t gsmem = rp_align_up(sizeof(RgAxes), 16) + sizeof(float) * (size_t)gfloats;
    if (gsmem > 227 * 1024) return RP_E_LIMIT;
    RP_SET_MAX_SMEM(rg_pool_bwd_kernel, gsmem);
    // with a fixed sampling grid <= RB_MAXG only bins wider than ~6 feature pixels lack a short-tap
    // form: too rare for a second launch, the gather kernel scatters them in reference order itself
    const int one_launch = sampling_ratio > 0 && (int)sampling_ratio <= RB_MAXG &&
                           rp_env_int("RP_ROI_BWD_SWEEP_FALLBACK", 0) == 0;
    rg_pool_bwd_kernel<<<grid, gthreads, gsmem, st>>>(lv, N, C, rois, levels, K, PH, PW, sampling_ratio,
                                                      dout, l.c_per_cta, gfloats, gcc, one_launch);
    RP_LAUNCH_CHECK();
    if (one_launch) return RP_OK;
  }
  RP_SET_MAX_SMEM(rt_pool_bwd_kernel, l.smem);
  rt_pool_bwd_kernel<<<grid, l.bwd_threads, l.smem, st>>>(lv, N, C, rois, levels, K, PH, PW,
                                                          sampling_ratio, dout, l.c_per_cta,
                                                          l.vs_floats, l.bwd_cc, gather);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
extern "C" {

int rp_roi_align_forward(const float* feat, int N, int C, int H, int W, const float* rois, int K,
                         int PH, int PW, float spatial_scale, float sampling_ratio, float* out,
                         int mode, void* stream) {
  if (K == 0) return RP_OK;
  if (!feat) return RP_E_ARG;
  RtLevels lv;
  const float* feats[1] = {feat};
  int rc = rt_fill_levels(&lv, feats, nullptr, &H, &W, &spatial_scale, 1);
  if (rc) return rc;
  return rt_forward(lv, N, C, rois, nullptr, K, PH, PW, sampling_ratio, out, mode,
                    (cudaStream_t)stream);
}

int rp_roi_align_backward(const float* dout, const float* rois, int K, int N, int C, int H, int W,
                          int PH, int PW, float spatial_scale, float sampling_ratio,
                          float* grad_input, int mode, void* stream) {
  if (!grad_input || N < 1 || C < 1) return RP_E_ARG;
  RtLevels lv;
  float* grads[1] = {grad_input};
  int rc = rt_fill_levels(&lv, nullptr, grads, &H, &W, &spatial_scale, 1);
  if (rc) return rc;
  return rt_backward(lv, N, C, rois, nullptr, K, PH, PW, sampling_ratio, dout, mode,
                     (cudaStream_t)stream);
}

int rp_level_map(const float* rois, int K, int k_min, int k_max, float canonical_scale,
                 int canonical_level, float eps, int* levels, void* stream) {
  if (K == 0) return RP_OK;
  if (!rois || !levels || K < 0 || k_max < k_min) return RP_E_ARG;
  rt_level_map_kernel<<<(K + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      rois, K, k_min, k_max, canonical_scale, canonical_level, eps, levels);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_level_map_area(const float* rois, const float* img_area, int K, int k_min, int k_max,
                      float eps, int* levels, void* stream) {
  if (K == 0) return RP_OK;
  if (!rois || !img_area || !levels || K < 0 || k_max < k_min) return RP_E_ARG;
  rt_level_map_area_kernel<<<(K + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      rois, img_area, K, k_min, k_max, eps, levels);
  RP_LAUNCH_CHECK();
  return RP_OK;
}

int rp_pooler_forward(const float* const* feats, const int* H, const int* W, const float* scales,
                      int L, int N, int C, const float* rois, const int* levels, int K, int PH,
                      int PW, float sampling_ratio, float* out, int mode, void* stream) {
  if (K == 0) return RP_OK;
  if (!feats || (L > 1 && !levels)) return RP_E_ARG;
  RtLevels lv;
  int rc = rt_fill_levels(&lv, feats, nullptr, H, W, scales, L);
  if (rc) return rc;
  return rt_forward(lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, mode,
                    (cudaStream_t)stream);
}

int rp_pooler_backward(const float* dout, const float* rois, const int* levels, int K,
                       const int* H, const int* W, const float* scales, int L, int N, int C,
                       int PH, int PW, float sampling_ratio, float* const* grads, int mode,
                       void* stream) {
  if (!grads || N < 1 || C < 1 || (L > 1 && K > 0 && !levels)) return RP_E_ARG;
  RtLevels lv;
  int rc = rt_fill_levels(&lv, nullptr, grads, H, W, scales, L);
  if (rc) return rc;
  return rt_backward(lv, N, C, rois, levels, K, PH, PW, sampling_ratio, dout, mode,
                     (cudaStream_t)stream);
}

int rp_roi_align_forward_nhwc(const float* feat, int N, int C, int H, int W, const float* rois, int K,
                              int PH, int PW, float spatial_scale, float sampling_ratio, float* out,
                              void* stream) {
  if (K == 0) return RP_OK;
  if (!feat) return RP_E_ARG;
  RtLevels lv;
  const float* feats[1] = {feat};
  int rc = rt_fill_levels(&lv, feats, nullptr, &H, &W, &spatial_scale, 1);
  if (rc) return rc;
  return rn_forward_nhwc(lv, N, C, rois, nullptr, K, PH, PW, sampling_ratio, out, (cudaStream_t)stream);
}

int rp_pooler_forward_nhwc(const float* const* feats, const int* H, const int* W, const float* scales,
                           int L, int N, int C, const float* rois, const int* levels, int K, int PH,
                           int PW, float sampling_ratio, float* out, void* stream) {
  if (K == 0) return RP_OK;
  if (!feats || (L > 1 && !levels)) return RP_E_ARG;
  RtLevels lv;
  int rc = rt_fill_levels(&lv, feats, nullptr, H, W, scales, L);
  if (rc) return rc;
  return rn_forward_nhwc(lv, N, C, rois, levels, K, PH, PW, sampling_ratio, out, (cudaStream_t)stream);
}

}  // extern "C"
