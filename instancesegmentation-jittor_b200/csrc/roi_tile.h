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

/* Forward sweep along one axis.
 *   ld(r, v)      loads tile row r (0 <= r < p.n) into v[0..U)
 *   sink(b, acc)  receives the finished bin b
 * Rows are fetched D ahead through a statically indexed register ring, so D*U loads are in
 * flight per thread; every branch depends only on the plan (uniform across the CTA). */
template <int U, int D, class Loader, class Sink>
RT_HD void rt_sweep_fwd(const RtAxisPlan& p, Loader& ld, Sink& sink) {
  float v[D][U];
  float acc[U];
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = 0.f;
  const int n = p.n;
  const int g = p.g;
#pragma unroll
  for (int j = 0; j < D; ++j) {
    if (j < n) {
      ld(j, v[j]);
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) v[j][u] = 0.f;
    }
  }
  int cnt = 0, bin = 0;
  for (int r = 0; r < n; r += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const int rr = r + j;
      if (rr < n) {
        const int jn = (j + 1) % D;
        const uint32_t rg = p.rng[rr];
        const int sb = (int)(rg & 0xffffu), se = (int)(rg >> 16);
        for (int s = sb; s < se; ++s) {
          const RtW w = p.w[s];
#pragma unroll
          for (int u = 0; u < U; ++u) acc[u] = fmaf(w.hi, v[jn][u], fmaf(w.lo, v[j][u], acc[u]));
          if (++cnt == g) {
            sink(bin, acc);
#pragma unroll
            for (int u = 0; u < U; ++u) acc[u] = 0.f;
            cnt = 0;
            ++bin;
          }
        }
        const int nr = rr + D;
        if (nr < n) {
          ld(nr, v[j]);
        } else {
#pragma unroll
          for (int u = 0; u < U; ++u) v[j][u] = 0.f;
        }
      }
    }
  }
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
 * Forward, v2: column walk over a raw tile held in shared memory.
 * y-plan: one entry per y-sample: weights of the (row, row+1) pair and the number of tile rows
 * to advance before the sample (monotone => the tile is walked once, top to bottom).
 * ------------------------------------------------------------------------------------------ */
struct __attribute__((aligned(16))) RtYSample {
  float wlo, whi;
  int adv;
  int pad;
};

/* Finish a y-plan serially: ylo[s] = lo row of sample s (or -1 when out of range), y0 = the
 * smallest lo.  (The kernel fills wlo/whi/ylo in parallel and calls this on one thread.) */
RT_HD void rt_finish_yplan(RtYSample* ytab, const short* ylo, int S, int y0) {
  int cur = 0;
  for (int s = 0; s < S; ++s) {
    const int lo = ylo[s];
    int adv = 0;
    if (lo >= 0) {
      adv = (lo - y0) - cur;
      cur = lo - y0;
    }
    ytab[s].adv = adv;
  }
}

/* One item = (output column pw, U channels q + u*nq) of one channel chunk.
 *   tile [cc][R rows][twa] with channel pitch chp; xw/xo: weights and tile-relative lo column
 *   of the x-samples (pw*gx + i); the hi tap is column lo+1 (weight 0 where it was clamped).
 *   sink(bin, acc) receives each finished output bin (unscaled sum over the gy*gx samples). */
template <int G, int U, class Sink>
RT_HD void rt_fwd_item(const float* tile, const RtYSample* ytab, const RtW* xw, const short* xo_t,
                       int S, int gy, int gx, int R, int twa, int chp, int q, int nq, int nvalid,
                       int pw, Sink& sink) {
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
  for (int u = 0; u < U; ++u) base[u] = tile + (u < nvalid ? (q + u * nq) : q) * chp;

  auto rowT = [&](int row, float(&T)[U]) {
    if (row < R) {
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
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) T[u] = 0.f;
    }
  };

  float Tlo[U], Thi[U], acc[U];
  rowT(0, Tlo);
  rowT(1, Thi);
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = 0.f;
  int row = 0, cnt = 0, bin = 0;
  for (int s = 0; s < S; ++s) {
    const RtYSample e = ytab[s];
    if (e.adv == 1) {
      ++row;
#pragma unroll
      for (int u = 0; u < U; ++u) Tlo[u] = Thi[u];
      rowT(row + 1, Thi);
    } else if (e.adv > 1) {
      row += e.adv;
      rowT(row, Tlo);
      rowT(row + 1, Thi);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = fmaf(e.whi, Thi[u], fmaf(e.wlo, Tlo[u], acc[u]));
    if (++cnt == gy) {
      sink(bin, acc);
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = 0.f;
      cnt = 0;
      ++bin;
    }
  }
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
