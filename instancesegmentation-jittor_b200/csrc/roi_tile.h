/*
 * roi_tile.h -- the per-RoI arithmetic of the tiled ROIAlign / Pooler kernels, written as
 * __host__ __device__ functions so that the CUDA kernels (roi_align.cu) and the host
 * emulation used by the CPU tests (tests/emu/roi_tile_emu.cpp) run the SAME code.
 *
 * What is computed (reference: det/layers/roi_align.py:116-169 forward, :190-274 backward):
 * the average over a gh x gw lattice of bilinear samples per output bin.  Because the
 * sample coordinates, their validity test (:20), the clamps (:25-45) and the bilinear
 * weights (:47-53) all factor per axis, the forward operator is separable:
 *
 *     out[c,ph,pw] = 1/(gh*gw) * sum_y sum_x Ay[ph,y] * Ax[pw,x] * F[c,y,x]
 *
 * with Ay/Ax holding <= 2 taps per sample.  The kernels evaluate it in two 1-D passes
 * (rows first, straight from global memory with lanes along x; columns second, from
 * shared memory), each a single monotone sweep over the tile that touches every tile
 * element exactly once.  The summation order therefore differs from the reference's
 * (w1*v1+w2*v2+w3*v3+w4*v4 per sample); results agree to ~1e-6 relative (tests: 1e-5).
 * The backward pass is the exact transpose of the same two sweeps.
 */
#ifndef RP_ROI_TILE_H_
#define RP_ROI_TILE_H_

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD static inline
#endif

#define RT_SMAX 128 /* max samples per axis (pooled * grid) handled by the tiled path */
#define RT_RMAX 128 /* max tile rows / columns handled by the tiled path */

/* RoI geometry: start, bin size and sampling grid per axis. */
struct RtGeom {
  float sw, sh; /* roi_start_w / roi_start_h */
  float bw, bh; /* bin_size_w / bin_size_h */
  int gw, gh;   /* roi_bin_grid_w / roi_bin_grid_h */
};

/* Forward geometry: roi_width/height promoted to double by max(float, 1.)
 * (det/layers/roi_align.py:139-140), adaptive grid = ceil(double / int) (:147-148). */
RT_HD RtGeom rt_geom_fwd(const float* roi, float scale, float sampling_ratio, int PH, int PW) {
  RtGeom g;
  const float roi_start_w = roi[1] * scale;
  const float roi_start_h = roi[2] * scale;
  const float roi_end_w = roi[3] * scale;
  const float roi_end_h = roi[4] * scale;
  double roi_width = (double)(roi_end_w - roi_start_w);
  roi_width = roi_width > 1. ? roi_width : 1.;
  double roi_height = (double)(roi_end_h - roi_start_h);
  roi_height = roi_height > 1. ? roi_height : 1.;
  g.bh = (float)roi_height / (float)PH;
  g.bw = (float)roi_width / (float)PW;
  g.gh = (sampling_ratio > 0) ? (int)(double)sampling_ratio : (int)ceil(roi_height / PH);
  g.gw = (sampling_ratio > 0) ? (int)(double)sampling_ratio : (int)ceil(roi_width / PW);
  if (g.gh < 0) g.gh = 0;
  if (g.gw < 0) g.gw = 0;
  g.sw = roi_start_w;
  g.sh = roi_start_h;
  return g;
}

/* Backward geometry: everything in float (det/layers/roi_align.py:226-229, :238-239). */
RT_HD RtGeom rt_geom_bwd(const float* roi, float scale, float sampling_ratio, int PH, int PW) {
  RtGeom g;
  const float roi_start_w = roi[1] * scale;
  const float roi_start_h = roi[2] * scale;
  const float roi_end_w = roi[3] * scale;
  const float roi_end_h = roi[4] * scale;
  float roi_width = roi_end_w - roi_start_w;
  roi_width = roi_width > 1.f ? roi_width : 1.f;
  float roi_height = roi_end_h - roi_start_h;
  roi_height = roi_height > 1.f ? roi_height : 1.f;
  g.bh = roi_height / (float)PH;
  g.bw = roi_width / (float)PW;
  g.gh = (sampling_ratio > 0) ? (int)sampling_ratio : (int)ceilf(roi_height / (float)PH);
  g.gw = (sampling_ratio > 0) ? (int)sampling_ratio : (int)ceilf(roi_width / (float)PW);
  if (g.gh < 0) g.gh = 0;
  if (g.gw < 0) g.gw = 0;
  g.sw = roi_start_w;
  g.sh = roi_start_h;
  return g;
}

/* One sample of one axis: coordinate (roi_align.py:156-159), range test (:20), clamps
 * (:25-45), weights (:47-49).  Returns 0 for an out-of-range sample. */
RT_HD int rt_axis_tap(float start, float bin, int p, int i, int grid, int extent, int* lo,
                      int* hi, float* wlo, float* whi) {
  float v = start + (float)p * bin + ((float)i + .5f) * bin / (float)grid;
  if (v < -1.0f || v > (float)extent) return 0;
  if (v <= 0) v = 0;
  int l = (int)v, h;
  if (l >= extent - 1) {
    h = l = extent - 1;
    v = (float)l;
  } else {
    h = l + 1;
  }
  const float f = v - (float)l;
  *lo = l;
  *hi = h;
  *whi = f;        /* ly / lx : weight of the hi tap */
  *wlo = 1.0f - f; /* hy / hx : weight of the lo tap */
  return 1;
}

/* The sweep plan of one axis of one RoI.
 *   tile row r (0 <= r < n) is feature row lo0 + r;
 *   samples (rng[r] & 0xffff) .. (rng[r] >> 16)-1 have their lo tap on tile row r and their hi tap on r+1
 *   (or on r with weight 0 when clamped at the border);
 *   out-of-range samples carry zero weights and ride on the first / last row.
 * The sample -> row map is monotone because the sample coordinate is monotone in s. */
struct __attribute__((aligned(8))) RtW {
  float lo, hi; /* weight of the lo tap (reference hy / hx), of the hi tap (ly / lx) */
};

struct RtAxisPlan {
  int lo0;
  int n; /* 0 when no sample is in range: the RoI's output is all zero */
  int S; /* samples = pooled * grid */
  int g; /* samples per bin */
  RtW w[RT_SMAX];
  uint32_t rng[RT_RMAX + 1]; /* row r: first sample | (one past last sample) << 16 */
  short lo[RT_SMAX];
};

/* Serial plan construction (one thread per axis).  Returns 0 when the axis does not fit the
 * tiled path (too many samples or rows): the caller takes the direct path. */
RT_HD int rt_build_axis_plan(RtAxisPlan* p, float start, float bin, int pooled, int grid,
                             int extent) {
  const int S = pooled * grid;
  p->S = S;
  p->g = grid;
  p->n = 0;
  p->lo0 = 0;
  if (S > RT_SMAX) return 0;
  int first_valid = -1, last_valid = -1, lo0 = 0, last_lo = 0, last_hi = -1;
  int pb = 0, ib = 0;
  for (int s = 0; s < S; ++s) {
    int lo = 0, hi = 0;
    float wlo = 0.f, whi = 0.f;
    if (rt_axis_tap(start, bin, pb, ib, grid, extent, &lo, &hi, &wlo, &whi)) {
      if (first_valid < 0) {
        first_valid = s;
        lo0 = lo;
      }
      last_valid = s;
      last_lo = lo;
      last_hi = hi;
    }
    p->w[s].lo = wlo;
    p->w[s].hi = whi;
    p->lo[s] = (short)(lo - lo0);
    if (++ib == grid) {
      ib = 0;
      ++pb;
    }
  }
  if (first_valid < 0) return 1; /* n = 0 */
  const int n = last_hi - lo0 + 1;
  if (n > RT_RMAX) return 0;
  p->lo0 = lo0;
  p->n = n;
  int cur = 0, first = 0;
  for (int s = 0; s < S; ++s) {
    int idx = (s < first_valid) ? 0 : (s > last_valid ? last_lo - lo0 : (int)p->lo[s]);
    while (cur < idx) {
      p->rng[cur] = (uint32_t)first | ((uint32_t)s << 16);
      first = s;
      ++cur;
    }
  }
  while (cur < n) {
    p->rng[cur] = (uint32_t)first | ((uint32_t)S << 16);
    first = S;
    ++cur;
  }
  return 1;
}

/* Transposed sweep (backward): bins are read, tile rows are produced.
 *   src(b, g)     loads the gradient of bin b into g[0..U)
 *   emit(r, val)  receives the finished gradient of tile row r (each row exactly once)
 * Row r collects  sum_{s: lo(s)=r} wl[s]*G[bin(s)]  +  sum_{s: lo(s)=r-1} wh[s]*G[bin(s)]. */
template <int U, class Source, class Emit>
RT_HD void rt_sweep_bwd(const RtAxisPlan& p, Source& src, Emit& emit) {
  float g[U], cur[U], nxt[U];
#pragma unroll
  for (int u = 0; u < U; ++u) g[u] = cur[u] = nxt[u] = 0.f;
  const int n = p.n;
  const int grid = p.g;
  int cnt = grid, bin = -1; /* forces a load at the first sample */
  for (int r = 0; r < n; ++r) {
    const uint32_t rg = p.rng[r];
    const int sb = (int)(rg & 0xffffu), se = (int)(rg >> 16);
    for (int s = sb; s < se; ++s) {
      if (cnt == grid) {
        cnt = 0;
        ++bin;
        src(bin, g);
      }
      ++cnt;
      const RtW w = p.w[s];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        cur[u] = fmaf(w.lo, g[u], cur[u]);
        nxt[u] = fmaf(w.hi, g[u], nxt[u]);
      }
    }
    emit(r, cur);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      cur[u] = nxt[u];
      nxt[u] = 0.f;
    }
  }
}

/* ------------------------------------------------------------------------------------------
 * Forward (v3): column walk over a raw feature tile held in shared memory.
 *
 * The CTA of one RoI stages the RoI's feature tile [cc channels][R rows][twa columns] and
 * every thread owns one ITEM = (output column pw, U channels, a contiguous range of output
 * rows [b0,b1) = "segment").  The item walks the tile rows of its segment once, top to bottom:
 * per row it forms the horizontally interpolated value T(row) of its column (2 taps per
 * x-sample, x taps in registers), and the y-samples consume the rolling pair
 * (T(row), T(row+1)).  Segments exist so that RoIs with large tiles (few channels fit in
 * shared memory) still give every thread of the CTA an item.
 * ------------------------------------------------------------------------------------------ */

/* A tap pair that never reads outside [0, extent): the reference's border clamp
 * (hi = lo = extent-1, weight of hi = 0, roi_align.py:33-45) is re-expressed on the pair
 * (extent-2, extent-1) with weights (0, 1) -- the same value, and lo+1 is always valid.
 * Needs extent >= 2.  Returns 0 for an out-of-range sample (zero weights, lo = -1 / -2). */
RT_HD int rt_axis_tap_pair(float start, float bin, int p, int i, int grid, int extent, int* lo,
                           float* wlo, float* whi) {
  int l = 0, h = 0;
  float a = 0.f, b = 0.f;
  const int ok = rt_axis_tap(start, bin, p, i, grid, extent, &l, &h, &a, &b);
  if (ok && h == l) { /* clamped at the far border: a = 1, b = 0 on (l, l) */
    l = l - 1;
    b = a;
    a = 0.f;
  }
  if (!ok) { /* -1: before the map, -2: past it (the caller parks it on the nearest tile row) */
    const float v = start + (float)p * bin + ((float)i + .5f) * bin / (float)grid;
    l = (v < -1.0f) ? -1 : -2;
  }
  *lo = l;
  *wlo = a;
  *whi = b;
  return ok;
}

struct __attribute__((aligned(16))) RtYSample {
  float wlo, whi; /* weights of tile rows (row, row + 1) */
  int row;        /* tile-relative row of the lo tap; monotone in the sample index */
  int flags;      /* RT_F_EMIT on the last y-sample of an output row, RT_F_ADV / RT_F_JUMP */
};

/* How one RoI's work is cut into chunks and items (all CTA-uniform). */
struct RtTilePlan {
  int R, twa, chp; /* tile rows, row pitch (floats), channel pitch (floats) */
  int nseg, ncg;   /* segments per column, channel groups per chunk: nseg * ncg == gpb */
  int cc;          /* channels staged per chunk (<= ncg * U) ; 0 => direct path */
};

RT_HD int rt_pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

/* gpb = thread groups per CTA (a power of two; each group = gw >= PW lanes, gw a power of two
 * when it is < 32), U = channels per thread, c_cta = channels this CTA must produce,
 * tile_floats = shared-memory tile budget. */
RT_HD RtTilePlan rt_plan_tile(int R, int twa, int tile_floats, int gpb, int gw, int U, int c_cta,
                              int PH) {
  RtTilePlan t;
  t.R = R;
  t.twa = twa;
  /* channel pitch == gw (mod 32) floats: the 32/gw channel groups that share a warp start gw
   * banks apart, so their loads do not collide (exactly when the tile is <= gw wide); stays a
   * multiple of 4 (16-byte rows) when twa is one */
  int chp = R * twa;
  if (gw >= 4 && gw < 32 && (twa & 3) == 0) {
    chp += (gw - chp) & 31;
    if (chp > tile_floats) chp = R * twa;
  }
  t.chp = chp;
  const int cc_fit = chp > 0 ? tile_floats / chp : 0;
  if (cc_fit < 1) {
    t.nseg = t.ncg = 1;
    t.cc = 0;
    return t;
  }
  const int seg_cap = rt_pow2_ceil(PH) < gpb ? rt_pow2_ceil(PH) : gpb;
  const int c_need = (c_cta + U - 1) / U * U;
  const int c_lim = cc_fit < c_need ? cc_fit : c_need;
  int nseg = 1;
  while (nseg < seg_cap && (gpb / nseg) * U > c_lim) nseg <<= 1;
  t.nseg = nseg;
  t.ncg = gpb / nseg;
  t.cc = t.ncg * U < cc_fit ? t.ncg * U : cc_fit;
  return t;
}

/* One item.  tile_g = tile + cg * chp (channel u of the item lives at tile_g + u * ustride,
 * channels >= nvalid alias channel 0 and are not emitted).  xw / xo_t: weights and
 * tile-relative lo column of the x-samples; ytab: the y-samples.  sink(bin, acc) receives the
 * unscaled sum over the gy*gx samples of output row `bin`, for bin = b0 .. b1-1 in order. */
template <int G, int U, class Sink>
RT_HD void rt_fwd_item(const float* tile_g, int ustride, int nvalid, const RtYSample* ytab,
                       const RtW* xw, const short* xo_t, int gy, int gx, int twa, int pw, int b0,
                       int b1, Sink& sink) {
  float hx[G > 0 ? G : 1], lx[G > 0 ? G : 1];
  int xo[G > 0 ? G : 1];
  if (G > 0) {
#pragma unroll
    for (int i = 0; i < G; ++i) {
      const RtW w = xw[pw * G + i];
      hx[i] = w.lo;
      lx[i] = w.hi;
      xo[i] = xo_t[pw * G + i];
    }
  }
  const float* base[U];
#pragma unroll
  for (int u = 0; u < U; ++u) base[u] = tile_g + (u < nvalid ? u * ustride : 0);

  auto rowT = [&](int row, float(&T)[U]) {
    const int ro = row * twa;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const float* p = base[u] + ro;
      float t = 0.f;
      if (G > 0) {
#pragma unroll
        for (int i = 0; i < G; ++i) {
          t = fmaf(hx[i], p[xo[i]], t);
          t = fmaf(lx[i], p[xo[i] + 1], t);
        }
      } else {
        for (int i = 0; i < gx; ++i) {
          const RtW w = xw[pw * gx + i];
          const int o = xo_t[pw * gx + i];
          t = fmaf(w.lo, p[o], t);
          t = fmaf(w.hi, p[o + 1], t);
        }
      }
      T[u] = t;
    }
  };

  if (b0 >= b1) return;
  float Tlo[U], Thi[U], acc[U];
  int s = b0 * gy;
  int row = ytab[s].row;
  rowT(row, Tlo);
  rowT(row + 1, Thi);
  for (int bin = b0; bin < b1; ++bin) {
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = 0.f;
    for (int i = 0; i < gy; ++i, ++s) {
      const RtYSample e = ytab[s];
      const int d = e.row - row;
      if (d != 0) {
        row = e.row;
        if (d == 1) {
#pragma unroll
          for (int u = 0; u < U; ++u) Tlo[u] = Thi[u];
        } else {
          rowT(row, Tlo);
        }
        rowT(row + 1, Thi);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = fmaf(e.whi, Thi[u], fmaf(e.wlo, Tlo[u], acc[u]));
    }
    sink(bin, acc);
  }
}

/* The same item for gx == 2 (x taps in registers, any gy), written for the instruction count.
 * Each y-sample record carries flags that say how the live row pair moves before the sample
 * (CTA-uniform, so every branch below is uniform):
 *   RT_F_ADV   advance by exactly one row: ONE new x-interpolated row is computed.  The two
 *              live rows sit in two register sets whose roles alternate, so the walk is a
 *              two-state machine (which set is the lo row) and nothing is ever moved;
 *   RT_F_JUMP  advance by d >= 2 rows: both rows are recomputed, (d - 2) is in flags >> 8;
 *   RT_F_EMIT  the sample is the last of its output row.
 * Identical arithmetic, operation for operation, to rt_fwd_item<2, U>. */
#define RT_F_ADV 1
#define RT_F_EMIT 2
#define RT_F_JUMP 4

/* flags of sample s from its own row, the previous sample's row and the emit bit */
RT_HD int rt_ysample_flags(int row, int prev_row, int emit) {
  const int d = row - prev_row;
  int f = emit ? RT_F_EMIT : 0;
  if (d == 1) f |= RT_F_ADV;
  if (d >= 2) f |= RT_F_JUMP | ((d - 2) << 8);
  return f;
}

template <int U, class Sink>
RT_HD void rt_fwd_item_fsm(const float* tile_g, int ustride, int nvalid, const RtYSample* ytab,
                           const RtW* xw, const short* xo_t, int gy, int twa, int pw, int b0,
                           int b1, Sink& sink) {
  if (b0 >= b1) return;
  const RtW w0 = xw[pw * 2], w1 = xw[pw * 2 + 1];
  const int xo0 = xo_t[pw * 2], xo1 = xo_t[pw * 2 + 1];
  const RtYSample* st = ytab + b0 * gy;
  int bins_left = b1 - b0;
  /* the first sample of a segment: its row pair is computed from scratch, flags ignored */
  const int r0 = st->row;
  const float* p0[U];
  const float* p1[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const float* b = tile_g + (u < nvalid ? u * ustride : 0) + r0 * twa;
    p0[u] = b + xo0;
    p1[u] = b + xo1;
  }
  float Ta[U], Tb[U], acc[U];
#define RT_XROW(T)                                                                              \
  _Pragma("unroll") for (int u = 0; u < U; ++u) {                                               \
    T[u] = fmaf(w1.hi, p1[u][1],                                                                \
                fmaf(w1.lo, p1[u][0], fmaf(w0.hi, p0[u][1], fmaf(w0.lo, p0[u][0], 0.f))));      \
    p0[u] += twa;                                                                               \
    p1[u] += twa;                                                                               \
  }
#define RT_JUMP(LO, HI)                                                                         \
  {                                                                                             \
    const int skip = (e.flags >> 8) * twa;                                                       \
    _Pragma("unroll") for (int u = 0; u < U; ++u) {                                             \
      p0[u] += skip;                                                                            \
      p1[u] += skip;                                                                            \
    }                                                                                           \
    RT_XROW(LO)                                                                                 \
    RT_XROW(HI)                                                                                 \
  }
#define RT_YFMA(LO, HI)                                                                         \
  _Pragma("unroll") for (int u = 0; u < U; ++u) acc[u] =                                        \
      fmaf(e.whi, HI[u], fmaf(e.wlo, LO[u], acc[u]));
#define RT_EMIT()                                                                               \
  if (e.flags & RT_F_EMIT) {                                                                     \
    sink(0, acc);                                                                               \
    _Pragma("unroll") for (int u = 0; u < U; ++u) acc[u] = 0.f;                                 \
    if (--bins_left == 0) return;                                                               \
  }
  RT_XROW(Ta)
  RT_XROW(Tb)
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = 0.f;
  {
    const RtYSample e = *st++;
    RT_YFMA(Ta, Tb)
    RT_EMIT()
  }
  for (;;) {
    /* state A: Ta = lo row, Tb = hi row */
    for (;;) {
      const RtYSample e = *st++;
      if (e.flags & RT_F_ADV) {
        RT_XROW(Ta) /* Ta becomes the new hi row */
        RT_YFMA(Tb, Ta)
        RT_EMIT()
        break;
      }
      if (e.flags & RT_F_JUMP) RT_JUMP(Ta, Tb)
      RT_YFMA(Ta, Tb)
      RT_EMIT()
    }
    /* state B: Tb = lo row, Ta = hi row */
    for (;;) {
      const RtYSample e = *st++;
      if (e.flags & RT_F_ADV) {
        RT_XROW(Tb)
        RT_YFMA(Ta, Tb)
        RT_EMIT()
        break;
      }
      if (e.flags & RT_F_JUMP) RT_JUMP(Tb, Ta)
      RT_YFMA(Tb, Ta)
      RT_EMIT()
    }
  }
#undef RT_XROW
#undef RT_JUMP
#undef RT_YFMA
#undef RT_EMIT
}

/* ------------------------------------------------------------------------------------------
 * Forward (v4, "banded"): every output bin of an axis reads a BAND of at most 4 consecutive
 * tile rows (columns).  With `grid` samples per bin spaced bin/grid apart, the lo taps of a bin
 * span ceil((grid-1)*bin/grid) rows and every sample adds its hi tap, so the band form exists
 * whenever (grid-1)*bin/grid <= 2 -- e.g. sampling_ratio 2 with bins up to 4 feature pixels
 * (RoIs up to 56 pixels on their FPN level at 14 bins; the FPN level mapping keeps them below
 * 45).  The per-sample weights of a bin are MERGED per row: out = sum_j w[j] * F[start + j],
 * which differs from the reference's per-sample order only by reassociation (~1e-7 relative).
 *
 * The two passes are then
 *   y-pass  thread = one 4-column strip of one channel: it reads its strip top to bottom ONCE
 *           (16-byte shared loads, a 4-deep register ring: row r lives in slot r & 3) and, as
 *           soon as row start+3 of a bin has arrived, forms V[bin][4 columns] with the bin's
 *           weights rotated to the ring (static register names, no moves, uniform branches);
 *   x-pass  thread = output column pw: out[ph][pw] = sum_j xw[pw][j] * V[ph][xo[pw] + j],
 *           consecutive threads write consecutive output floats.
 * Everything below is CTA-uniform data built once per RoI.
 * ------------------------------------------------------------------------------------------ */
#define RB_MAXP 32 /* pooled bins per axis */
#define RB_MAXG 8  /* samples per bin per axis */
#define RB_MAXE 96 /* tile rows / columns */
#define RB_MAXSEG 2

struct __attribute__((aligned(16))) RbF4 {
  float x, y, z, w;
};

struct RbAxis {
  int lo0; /* feature row / column of tile row / column 0 */
  int ext; /* tile extent: last band start + 4 */
  int start[RB_MAXP]; /* band start per bin: absolute (-1 = no sample in range) until
                         rb_finish_axis, tile-relative afterwards; non-decreasing */
  RbF4 w[RB_MAXP];    /* merged weights of rows start .. start+3 */
};

/* Cheap necessary condition, evaluated identically by every thread. */
RT_HD int rb_axis_candidate(float bin, int grid, int pooled, int extent) {
  if (grid < 1 || grid > RB_MAXG || pooled < 1 || pooled > RB_MAXP || extent < 4) return 0;
  return (float)(grid - 1) * bin / (float)grid <= 2.0f;
}

/* One bin (independent of the other bins).  Returns 0 when the bin's taps do not fit a band. */
RT_HD int rb_build_bin(RbAxis* a, int b, float start, float bin, int grid, int extent) {
  int first = 0x7fffffff, last = -1;
  for (int i = 0; i < grid; ++i) {
    int lo, hi;
    float wlo, whi;
    if (rt_axis_tap(start, bin, b, i, grid, extent, &lo, &hi, &wlo, &whi)) {
      first = lo < first ? lo : first;
      last = hi > last ? hi : last;
    }
  }
  float w0 = 0.f, w1 = 0.f, w2 = 0.f, w3 = 0.f;
  int st = -1, ok = 1;
  if (last >= 0) {
    st = first <= extent - 4 ? first : extent - 4;
    if (last - st > 3) {
      ok = 0;
    } else {
      for (int i = 0; i < grid; ++i) {
        int lo, hi;
        float wlo, whi;
        if (rt_axis_tap(start, bin, b, i, grid, extent, &lo, &hi, &wlo, &whi)) {
          const int jl = lo - st, jh = hi - st;
          w0 += (jl == 0 ? wlo : 0.f) + (jh == 0 ? whi : 0.f);
          w1 += (jl == 1 ? wlo : 0.f) + (jh == 1 ? whi : 0.f);
          w2 += (jl == 2 ? wlo : 0.f) + (jh == 2 ? whi : 0.f);
          w3 += (jl == 3 ? wlo : 0.f) + (jh == 3 ? whi : 0.f);
        }
      }
    }
  }
  a->start[b] = st;
  a->w[b].x = w0;
  a->w[b].y = w1;
  a->w[b].z = w2;
  a->w[b].w = w3;
  return ok;
}

/* Serial pass over the bins: bins without a sample ride on a neighbour's band (zero weights),
 * starts become tile-relative.  Returns 0 when the tile is too large. */
RT_HD int rb_finish_axis(RbAxis* a, int n) {
  int prev = -1;
  for (int b = 0; b < n; ++b)
    if (a->start[b] >= 0) {
      prev = a->start[b];
      break;
    }
  if (prev < 0) prev = 0; /* nothing in range: all weights are zero, any band will do */
  for (int b = 0; b < n; ++b) {
    if (a->start[b] < 0)
      a->start[b] = prev;
    else
      prev = a->start[b];
  }
  a->lo0 = a->start[0];
  a->ext = a->start[n - 1] + 4 - a->lo0;
  for (int b = 0; b < n; ++b) a->start[b] -= a->lo0;
  return a->ext <= RB_MAXE;
}

/* CTA-uniform plan of one RoI. */
struct RbPlan {
  RbAxis ay, ax;
  RbF4 wrot[RB_MAXP];                     /* y weights rotated to the ring of the bin's segment */
  RbF4 xw[RB_MAXP];                       /* x weights times 1/(gy*gx) */
  int xo[RB_MAXP];                        /* first V column of the band of pw */
  uint32_t mw[RB_MAXSEG][RB_MAXE / 4 + 1]; /* byte k of word i: bins completed by row 4i+k */
  int seg_b0[RB_MAXSEG + 1];              /* bins of segment s: seg_b0[s] .. seg_b0[s+1]-1 */
  int seg_r0[RB_MAXSEG];                  /* first tile row of the segment */
  int seg_rows[RB_MAXSEG];                /* rows the segment walks */
};

RT_HD void rb_plan_segments(RbPlan* p, int PH, int nseg) {
  for (int s = 0; s <= nseg; ++s) p->seg_b0[s] = s * PH / nseg;
  for (int s = 0; s < nseg; ++s) {
    p->seg_r0[s] = p->ay.start[p->seg_b0[s]];
    p->seg_rows[s] = p->ay.start[p->seg_b0[s + 1] - 1] + 4 - p->seg_r0[s];
  }
}

/* Per-bin derived data; b < max(PH, PW).  xoff = tile column of feature column ax.lo0. */
RT_HD void rb_plan_bin(RbPlan* p, int b, int PH, int PW, int nseg, int xoff, float inv) {
  if (b < PH) {
    int s = 0;
    while (s + 1 < nseg && b >= p->seg_b0[s + 1]) ++s;
    const int rot = (p->ay.start[b] - p->seg_r0[s]) & 3; /* ring slot of the band's first row */
    const RbF4 w = p->ay.w[b];
    RbF4 r;
    r.x = rot == 0 ? w.x : (rot == 1 ? w.w : (rot == 2 ? w.z : w.y));
    r.y = rot == 0 ? w.y : (rot == 1 ? w.x : (rot == 2 ? w.w : w.z));
    r.z = rot == 0 ? w.z : (rot == 1 ? w.y : (rot == 2 ? w.x : w.w));
    r.w = rot == 0 ? w.w : (rot == 1 ? w.z : (rot == 2 ? w.y : w.x));
    p->wrot[b] = r;
  }
  if (b < PW) {
    const RbF4 w = p->ax.w[b];
    RbF4 r;
    r.x = w.x * inv;
    r.y = w.y * inv;
    r.z = w.z * inv;
    r.w = w.w * inv;
    p->xw[b] = r;
    p->xo[b] = p->ax.start[b] + xoff;
  }
}

/* Word i of segment s: how many bins complete at rows 4i .. 4i+3 of the segment. */
RT_HD void rb_plan_mword(RbPlan* p, int s, int i) {
  uint32_t m = 0;
  for (int b = p->seg_b0[s]; b < p->seg_b0[s + 1]; ++b) {
    const int r = p->ay.start[b] - p->seg_r0[s] + 3;
    if ((r >> 2) == i) m += 1u << (8 * (r & 3));
  }
  p->mw[s][i] = m;
}

/* y-pass of one strip: rows = the segment's rows of this strip (pitch rp4 RbF4), wr / v start at
 * the segment's first bin; V rows are vp floats apart. */
RT_HD void rb_ywalk(const RbF4* rows, int rp4, int nrows, const uint32_t* mw, const RbF4* wr,
                    float* v, int vp) {
  RbF4 W0, W1, W2, W3;
  W0.x = W0.y = W0.z = W0.w = 0.f;
  W1 = W0;
  W2 = W0;
  W3 = W0;
#define RB_BINS(cnt)                                                                             \
  for (uint32_t m_ = (cnt); m_ != 0; --m_) {                                                    \
    const RbF4 w_ = *wr++;                                                                      \
    RbF4 a_;                                                                                    \
    a_.x = fmaf(w_.w, W3.x, fmaf(w_.z, W2.x, fmaf(w_.y, W1.x, w_.x * W0.x)));                   \
    a_.y = fmaf(w_.w, W3.y, fmaf(w_.z, W2.y, fmaf(w_.y, W1.y, w_.x * W0.y)));                   \
    a_.z = fmaf(w_.w, W3.z, fmaf(w_.z, W2.z, fmaf(w_.y, W1.z, w_.x * W0.z)));                   \
    a_.w = fmaf(w_.w, W3.w, fmaf(w_.z, W2.w, fmaf(w_.y, W1.w, w_.x * W0.w)));                   \
    *reinterpret_cast<RbF4*>(v) = a_;                                                           \
    v += vp;                                                                                    \
  }
  for (int r = 0; r < nrows; r += 4) {
    const uint32_t m = mw[r >> 2];
    W0 = rows[0];
    RB_BINS(m & 0xffu)
    if (r + 1 < nrows) W1 = rows[rp4];
    RB_BINS((m >> 8) & 0xffu)
    if (r + 2 < nrows) W2 = rows[2 * rp4];
    RB_BINS((m >> 16) & 0xffu)
    if (r + 3 < nrows) W3 = rows[3 * rp4];
    RB_BINS(m >> 24)
    rows += 4 * rp4;
  }
#undef RB_BINS
}

/* Channels per chunk and y segments for one RoI (CTA-uniform).  per_ch = shared-memory floats
 * one channel needs (tile + V), nq = strips per channel, nt = threads. */
RT_HD void rb_plan_chunk(int budget_floats, int per_ch, int nq, int nch, int PH, int nt, int* cc_out,
                         int* nseg_out, int cmul = 1) {
  int cc_fit = per_ch > 0 ? budget_floats / per_ch : 0;
  if (cmul > 1) { /* chunks of whole channel groups (TMA boxes); the last one may run past nch */
    cc_fit = cc_fit / cmul * cmul;
    const int nch_up = (nch + cmul - 1) / cmul * cmul;
    if (cc_fit > nch_up) cc_fit = nch_up;
    int best_cc = 0, best_seg = 1;
    float best = -1.f;
    for (int nseg = 1; nseg <= RB_MAXSEG && nseg <= PH; ++nseg) {
      for (int cc = cmul; cc <= cc_fit; cc += cmul) {
        const int items = cc * nq * nseg;
        const int nchunk = (nch + cc - 1) / cc;
        /* thread utilisation of a full chunk x how well the chunks cover nch */
        const float util = (float)items / (float)((items + nt - 1) / nt * nt);
        const float cover = (float)nch / (float)(nchunk * cc);
        const float score = util * cover * (nseg == 1 ? 1.0f : 0.88f) * (1.0f - 0.25f / (float)cc);
        if (score > best) {
          best = score;
          best_cc = cc;
          best_seg = nseg;
        }
      }
    }
    *cc_out = best_cc;
    *nseg_out = best_seg;
    return;
  }
  if (cc_fit > nch) cc_fit = nch;
  int best_cc = cc_fit, best_seg = 1;
  float best = -1.f;
  for (int nseg = 1; nseg <= RB_MAXSEG && nseg <= PH; ++nseg) {
    if (cc_fit < 1) break;
    const int per = nq * nseg;
    int rounds = cc_fit * per / nt;
    if (rounds < 1) rounds = 1;
    int cc = rounds * nt / per;
    if (cc > cc_fit) cc = cc_fit;
    if (cc < 1) cc = 1;
    /* even chunks over the CTA's channels */
    const int nchunk = (nch + cc - 1) / cc;
    cc = (nch + nchunk - 1) / nchunk;
    const int items = cc * per;
    const float util = (float)items / (float)((items + nt - 1) / nt * nt);
    const float score = util * (nseg == 1 ? 1.0f : 0.88f); /* a second segment re-reads 3 rows */
    if (score > best) {
      best = score;
      best_cc = cc;
      best_seg = nseg;
    }
  }
  *cc_out = best_cc;
  *nseg_out = best_seg;
}

/* 1/count when that is exact (count a power of two), else 0 => divide. */
RT_HD float rt_exact_inv(int count) {
  return (count > 0 && (count & (count - 1)) == 0) ? 1.0f / (float)count : 0.0f;
}

/* One output in the reference's own operand order (det/layers/roi_align.py:14-60,150-165):
 * used by mode 1 and for RoIs that do not fit the tiled path. */
RT_HD float rt_direct_output(const float* plane, int H, int W, const RtGeom& g, int ph, int pw) {
  float out = 0.f;
  for (int iy = 0; iy < g.gh; ++iy) {
    int ylo, yhi;
    float hy, ly;
    const int vy = rt_axis_tap(g.sh, g.bh, ph, iy, g.gh, H, &ylo, &yhi, &hy, &ly);
    for (int ix = 0; ix < g.gw; ++ix) {
      int xlo, xhi;
      float hx, lx;
      const int vx = rt_axis_tap(g.sw, g.bw, pw, ix, g.gw, W, &xlo, &xhi, &hx, &lx);
      float val = 0.f;
      if (vy && vx) {
        const float v1 = plane[ylo * W + xlo];
        const float v2 = plane[ylo * W + xhi];
        const float v3 = plane[yhi * W + xlo];
        const float v4 = plane[yhi * W + xhi];
        const float w1 = hy * hx, w2 = hy * lx, w3 = ly * hx, w4 = ly * lx;
        val = w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4;
      }
      out += val;
    }
  }
  return out / (float)(g.gh * g.gw);
}

#endif /* RP_ROI_TILE_H_ */
