// roi_tile_emu.cpp -- HOST emulation of the tiled ROIAlign kernels (test infrastructure).
//
// Runs the same roi_tile.h code the CUDA kernels run (plans + sweeps), with the CTA's thread
// loops executed serially and each __syncthreads() replaced by finishing the loop.  It lets
// the CPU test-suite check the tiled algorithm (item mapping, channel chunking, plan
// construction, forward and transposed sweeps) against the oracle without a GPU.
// The chunk / item arithmetic below mirrors rt_pool_fwd_kernel / rt_pool_bwd_kernel in
// instancesegmentation-jittor_b200/csrc/roi_align.cu line for line.
#include <algorithm>
#include <cstdint>
#include <cmath>
#include <cstring>
#include <vector>

#include "../../instancesegmentation-jittor_b200/csrc/roi_tile.h"

namespace {
constexpr int U1 = 4, D1 = 8, U2 = 2, D2 = 4, CC_CAP = 32;

template <int U>
struct RowLoader {
  const float* base;
  size_t cstride;
  int W, nvalid;
  void operator()(int r, float (&v)[U]) const {
    for (int u = 0; u < U; ++u) v[u] = (u < nvalid) ? base[(size_t)r * W + (size_t)u * cstride] : 0.f;
  }
};
template <int U>
struct VSink {
  float* base;
  int bstride, cstride, nvalid;
  void operator()(int bin, const float (&acc)[U]) const {
    for (int u = 0; u < U; ++u)
      if (u < nvalid) base[bin * bstride + u * cstride] = acc[u];
  }
};
template <int U>
struct VLoader {
  const float* base;
  int cstride, nvalid;
  void operator()(int r, float (&v)[U]) const {
    for (int u = 0; u < U; ++u) v[u] = (u < nvalid) ? base[u * cstride + r] : 0.f;
  }
};
template <int U>
struct OutSink {
  float* base;
  int cstride, nvalid;
  float inv, count;
  void operator()(int bin, const float (&acc)[U]) const {
    for (int u = 0; u < U; ++u)
      if (u < nvalid) base[u * cstride + bin] = (inv != 0.f) ? acc[u] * inv : acc[u] / count;
  }
};
template <int U>
struct DoutSource {
  const float* base;
  int cstride, nvalid;
  void operator()(int bin, float (&g)[U]) const {
    for (int u = 0; u < U; ++u) g[u] = (u < nvalid) ? base[u * cstride + bin] : 0.f;
  }
};
template <int U>
struct HEmit {
  float* base;
  int cstride, nvalid;
  void operator()(int r, const float (&v)[U]) const {
    for (int u = 0; u < U; ++u)
      if (u < nvalid) base[u * cstride + r] = v[u];
  }
};
template <int U>
struct HSource {
  const float* base;
  int bstride, cstride, nvalid;
  void operator()(int bin, float (&g)[U]) const {
    for (int u = 0; u < U; ++u) g[u] = (u < nvalid) ? base[bin * bstride + u * cstride] : 0.f;
  }
};
template <int U>
struct GradEmit {
  float* base;
  size_t cstride;
  int W, nvalid;
  float inv, count;
  void operator()(int r, const float (&v)[U]) const {
    for (int u = 0; u < U; ++u)
      if (u < nvalid) {
        const float val = (inv != 0.f) ? v[u] * inv : v[u] / count;
        if (val != 0.f) base[(size_t)r * W + (size_t)u * cstride] += val;
      }
  }
};
}  // namespace

template <int U>
struct EmuOutSink {
  float* optr;
  size_t cstride;
  int PW, nvalid;
  float inv;
  void operator()(int, const float (&acc)[U]) {
    for (int u = 0; u < U; ++u)
      if (u < nvalid) optr[(size_t)u * cstride] = acc[u] * inv;
    optr += PW;
  }
};

// levels == NULL: single level.  feats/H/W/scales are arrays of length L.
// Mirrors rt_pool_fwd_kernel<U> (v3); the CTA's threads run serially.
template <int U>
static int emu_pooler_forward_t(const float* const* feats, const int* Hs, const int* Ws,
                                const float* scales, int L, int N, int C, const float* rois,
                                const int* levels, int K, int PH, int PW, float sampling_ratio,
                                float* out, int tile_floats, int c_per_cta) {
  const int P = PH * PW;
  int n_direct = 0;
  int gw = PW;
  if (PW < 32) {
    gw = 1;
    while (gw < PW) gw *= 2;
  }
  int gpb = 1;
  while (gw * gpb * 2 <= 256) gpb *= 2;
  std::vector<float> tile(tile_floats);
  std::vector<RtYSample> ytab(RT_SMAX);
  std::vector<RtW> xw(RT_SMAX);
  std::vector<short> xo(RT_SMAX);
  for (int k = 0; k < K; ++k) {
    for (int c_begin = 0; c_begin < C; c_begin += c_per_cta) {
      const int c_end = std::min(C, c_begin + c_per_cta);
      const int nch = c_end - c_begin;
      const float* roi = rois + (size_t)k * 5;
      const int level = levels ? levels[k] : 0;
      const int b = (int)roi[0];
      float* out_k = out + ((size_t)k * C + c_begin) * P;
      if (level < 0 || level >= L || b < 0 || b >= N) {
        std::fill(out_k, out_k + (size_t)nch * P, 0.f);
        continue;
      }
      const int H = Hs[level], W = Ws[level];
      const size_t HW = (size_t)H * W;
      const float* feat_b = feats[level] + ((size_t)b * C + c_begin) * HW;
      const RtGeom g = rt_geom_fwd(roi, scales[level], sampling_ratio, PH, PW);
      const int SY = PH * g.gh, SX = PW * g.gw;
      const bool fits = (SY <= RT_SMAX) && (SX <= RT_SMAX) && SY > 0 && SX > 0 && H >= 2 && W >= 2;
      int ext[4] = {0x7fffffff, -1, 0x7fffffff, -1};
      if (fits) {
        for (int s = 0; s < SY + SX; ++s) {
          const bool isy = s < SY;
          const int t = isy ? s : s - SY;
          const int grid = isy ? g.gh : g.gw;
          const int p = t / grid, i = t - p * grid;
          int lo = 0;
          float wlo = 0.f, whi = 0.f;
          const int ok = isy ? rt_axis_tap_pair(g.sh, g.bh, p, i, grid, H, &lo, &wlo, &whi)
                             : rt_axis_tap_pair(g.sw, g.bw, p, i, grid, W, &lo, &wlo, &whi);
          if (ok) {
            ext[isy ? 0 : 2] = std::min(ext[isy ? 0 : 2], lo);
            ext[isy ? 1 : 3] = std::max(ext[isy ? 1 : 3], lo);
          }
          if (isy) {
            ytab[t].wlo = wlo;
            ytab[t].whi = whi;
            ytab[t].row = lo;
            ytab[t].flags = (i == grid - 1) ? RT_F_EMIT : 0;
          } else {
            xw[t].lo = wlo;
            xw[t].hi = whi;
            xo[t] = (short)lo;
          }
        }
      }
      const int y0 = ext[0], yl = ext[1], x0 = ext[2], xl = ext[3];
      if (fits && (yl < 0 || xl < 0)) {
        std::fill(out_k, out_k + (size_t)nch * P, 0.f);
        continue;
      }
      const int R = yl - y0 + 2;
      const int vecw = (W & 3) == 0 ? 4 : ((W & 1) == 0 ? 2 : 1);
      const int xa0 = x0 & ~(vecw - 1);
      const int twa = ((xl + 1) / vecw - xa0 / vecw + 1) * vecw;
      RtTilePlan tp;
      tp.cc = 0;
      if (fits) tp = rt_plan_tile(R, twa, tile_floats, gpb, gw, U, nch, PH);
      if (tp.cc < 1) {
        ++n_direct;
        const int total = nch * P;
        for (int i = 0; i < total; ++i) {
          const int c = i / P, rem = i - c * P, ph = rem / PW, pw = rem - ph * PW;
          out_k[i] = rt_direct_output(feat_b + (size_t)c * HW, H, W, g, ph, pw);
        }
        continue;
      }
      for (int s = 0; s < SY + SX; ++s) {
        if (s < SY) {
          const int lo = ytab[s].row;
          ytab[s].row = lo >= 0 ? lo - y0 : (lo == -1 ? 0 : yl - y0);
        } else {
          const int t = s - SY;
          const int lo = xo[t];
          xo[t] = (short)(lo >= 0 ? lo - xa0 : (lo == -1 ? x0 - xa0 : xl - xa0));
        }
      }
      {
        std::vector<int> fl(SY);
        for (int s = 0; s < SY; ++s)
          fl[s] = rt_ysample_flags(ytab[s].row, s ? ytab[s - 1].row : ytab[s].row, ytab[s].flags & RT_F_EMIT);
        for (int s = 0; s < SY; ++s) ytab[s].flags = fl[s];
      }
      const int chp = tp.chp;
      const int count = g.gh * g.gw;
      const float inv = 1.0f / (float)count;
      const float* src0 = feat_b + (size_t)y0 * W + xa0;
      for (int c0 = 0; c0 < nch; c0 += tp.cc) {
        const int cc = std::min(tp.cc, nch - c0);
        std::fill(tile.begin(), tile.end(), NAN);  // poison: anything read must have been staged
        for (int c = 0; c < cc; ++c)
          for (int y = 0; y < R; ++y)
            std::memcpy(&tile[c * chp + y * twa], src0 + (size_t)(c0 + c) * HW + (size_t)y * W,
                        sizeof(float) * twa);
        for (int tid = 0; tid < gw * gpb; ++tid) {
          const int grp = tid / gw, pw = tid - grp * gw;
          if (pw >= PW) continue;
          const int cg = grp & (tp.ncg - 1), seg = grp / tp.ncg;
          const int b0 = seg * PH / tp.nseg, b1 = (seg + 1) * PH / tp.nseg;
          if (b0 >= b1) continue;
          int nvalid = (cc - cg + tp.ncg - 1) / tp.ncg;
          if (nvalid > U) nvalid = U;
          if (nvalid <= 0) continue;
          EmuOutSink<U> sink{out_k + (size_t)(c0 + cg) * P + b0 * PW + pw, (size_t)tp.ncg * P, PW,
                             nvalid, inv};
          const float* tile_g = tile.data() + cg * chp;
          const int ustride = tp.ncg * chp;
          if (g.gw == 2)
            rt_fwd_item_fsm<U>(tile_g, ustride, nvalid, ytab.data(), xw.data(), xo.data(), g.gh, twa, pw, b0, b1, sink);
          else
            rt_fwd_item<0, U>(tile_g, ustride, nvalid, ytab.data(), xw.data(), xo.data(), g.gh, g.gw, twa, pw, b0, b1, sink);
        }
      }
    }
  }
  return n_direct;
}

extern "C" __attribute__((visibility("default"))) int emu_pooler_forward(
    const float* const* feats, const int* Hs, const int* Ws, const float* scales, int L, int N,
    int C, const float* rois, const int* levels, int K, int PH, int PW, float sampling_ratio,
    float* out, int tile_floats, int c_per_cta, int U) {
  switch (U) {
    case 1: return emu_pooler_forward_t<1>(feats, Hs, Ws, scales, L, N, C, rois, levels, K, PH, PW, sampling_ratio, out, tile_floats, c_per_cta);
    case 4: return emu_pooler_forward_t<4>(feats, Hs, Ws, scales, L, N, C, rois, levels, K, PH, PW, sampling_ratio, out, tile_floats, c_per_cta);
    default: return emu_pooler_forward_t<2>(feats, Hs, Ws, scales, L, N, C, rois, levels, K, PH, PW, sampling_ratio, out, tile_floats, c_per_cta);
  }
}

// Mirrors rb_pool_fwd_kernel (v4, banded): the same plan functions, chunking, strip walk and
// x-pass; the CTA's threads run serially.  Returns the number of (RoI, slice) pairs without a
// band form (evaluated in the reference's order here; by the v3 body on the GPU).
extern "C" __attribute__((visibility("default"))) int emu_band_forward(
    const float* const* feats, const int* Hs, const int* Ws, const float* scales, int L, int N,
    int C, const float* rois, const int* levels, int K, int PH, int PW, float sampling_ratio,
    float* out, int smem_floats, int c_per_cta, int nthreads) {
  const int P = PH * PW;
  int n_fallback = 0;
  std::vector<float> smem((size_t)smem_floats + 4);
  float* base = smem.data();
  while (reinterpret_cast<uintptr_t>(base) & 15) ++base;
  for (int k = 0; k < K; ++k) {
    for (int c_begin = 0; c_begin < C; c_begin += c_per_cta) {
      const int c_end = std::min(C, c_begin + c_per_cta);
      const int nch = c_end - c_begin;
      const float* roi = rois + (size_t)k * 5;
      const int level = levels ? levels[k] : 0;
      const int b = (int)roi[0];
      float* out_k = out + ((size_t)k * C + c_begin) * P;
      if (level < 0 || level >= L || b < 0 || b >= N) {
        std::fill(out_k, out_k + (size_t)nch * P, 0.f);
        continue;
      }
      const int H = Hs[level], W = Ws[level];
      const size_t HW = (size_t)H * W;
      const RtGeom g = rt_geom_fwd(roi, scales[level], sampling_ratio, PH, PW);
      static RbPlan plan;
      int ok = rb_axis_candidate(g.bh, g.gh, PH, H) && rb_axis_candidate(g.bw, g.gw, PW, W);
      if (ok) {
        for (int t = 0; t < PH; ++t) ok &= rb_build_bin(&plan.ay, t, g.sh, g.bh, g.gh, H);
        for (int t = 0; t < PW; ++t) ok &= rb_build_bin(&plan.ax, t, g.sw, g.bw, g.gw, W);
      }
      if (ok) ok = rb_finish_axis(&plan.ay, PH) & rb_finish_axis(&plan.ax, PW);
      int cc = 0, nseg = 1, nq = 0, vp = 0, chp4 = 0, xa0 = 0, vecw = 1, R = 0, y0 = 0;
      if (ok) {
        y0 = plan.ay.lo0;
        R = plan.ay.ext;
        const int x0 = plan.ax.lo0, xe = x0 + plan.ax.ext;
        vecw = (W & 3) == 0 ? 4 : ((W & 1) == 0 ? 2 : 1);
        xa0 = x0 & ~(vecw - 1);
        nq = (xe - xa0 + 3) >> 2;
        {
          int vpu = nq;
          while ((vpu & 7) != 3 && (vpu & 7) != 5) ++vpu;
          vp = 4 * vpu;
        }
        chp4 = R * nq + ((nq - R * nq) & 7);
        if ((4 / vecw) * nq <= 32) {
          rb_plan_chunk(smem_floats, 4 * chp4 + PH * vp, nq, nch, PH, nthreads, &cc, &nseg);
          ok = cc >= 1;
        } else {
          ok = 0;
        }
      }
      if (!ok) {
        ++n_fallback;
        const float* feat_b = feats[level] + ((size_t)b * C + c_begin) * HW;
        for (int i = 0; i < nch * P; ++i) {
          const int c = i / P, rem = i - c * P, ph = rem / PW, pw = rem - ph * PW;
          out_k[i] = rt_direct_output(feat_b + (size_t)c * HW, H, W, g, ph, pw);
        }
        continue;
      }
      rb_plan_segments(&plan, PH, nseg);
      for (int t = 0; t < 32; ++t)
        rb_plan_bin(&plan, t, PH, PW, nseg, plan.ax.lo0 - xa0, 1.0f / (float)(g.gh * g.gw));
      for (int sgm = 0; sgm < nseg; ++sgm)
        for (int i = 0; i * 4 < plan.seg_rows[sgm] + 4; ++i) rb_plan_mword(&plan, sgm, i);
      float* tile = base;
      const int chp = 4 * chp4;
      float* V = tile + cc * chp;
      const float* src0 = feats[level] + ((size_t)b * C + c_begin) * HW + (size_t)y0 * W + xa0;
      const int wlim = W - xa0;
      const int xg = PW <= 16 ? 16 : PW;
      const int nslots = nthreads / xg;
      int tap3 = 1;
      for (int t = 0; t < PW; ++t) tap3 &= plan.xw[t].w == 0.f;
      for (int c0 = 0; c0 < nch; c0 += cc) {
        const int ccur = std::min(cc, nch - c0);
        std::fill(base, base + smem_floats, NAN);  // poison: anything read must have been written
        for (int c = 0; c < ccur; ++c)
          for (int y = 0; y < R; ++y)
            for (int col = 0; col < 4 * nq; ++col)
              tile[c * chp + y * 4 * nq + col] =
                  col < wlim ? src0[(size_t)(c0 + c) * HW + (size_t)y * W + col] : 0.f;
        const int per_seg = ccur * nq, total = per_seg * nseg;
        for (int sidx = 0; sidx < total; ++sidx) {
          const int seg = sidx >= per_seg ? 1 : 0;
          const int rem = sidx - seg * per_seg;
          const int c = rem / nq, q = rem - c * nq;
          const int b0 = plan.seg_b0[seg];
          rb_ywalk(reinterpret_cast<const RbF4*>(tile) + c * chp4 + plan.seg_r0[seg] * nq + q, nq,
                   plan.seg_rows[seg], plan.mw[seg], plan.wrot + b0, V + (c * PH + b0) * vp + 4 * q, vp);
        }
        for (int tid = 0; tid < nthreads; ++tid) {
          const int pw = tid % xg, slot = tid / xg;
          if (!(pw < PW && slot < nslots)) continue;
          const RbF4 w = plan.xw[pw];
          const float* v = V + slot * vp + plan.xo[pw];
          float* o = out_k + (size_t)c0 * P + slot * PW + pw;
          for (int r = slot; r < ccur * PH; r += nslots) {
            *o = tap3 ? fmaf(w.z, v[2], fmaf(w.y, v[1], w.x * v[0]))
                      : fmaf(w.w, v[3], fmaf(w.z, v[2], fmaf(w.y, v[1], w.x * v[0])));
            v += nslots * vp;
            o += nslots * PW;
          }
        }
      }
    }
  }
  return n_fallback;
}

extern "C" {

// grads[l] must be zero-filled by the caller.
__attribute__((visibility("default"))) int emu_pooler_backward(
    const float* dout, const float* rois, const int* levels, int K, const int* Hs, const int* Ws,
    const float* scales, int L, int N, int C, int PH, int PW, float sampling_ratio,
    float* const* grads, int vs_floats, int c_per_cta) {
  const int P = PH * PW;
  int n_direct = 0;
  std::vector<float> Hbuf(vs_floats), d_s((size_t)CC_CAP * P);
  for (int k = 0; k < K; ++k) {
    for (int c_begin = 0; c_begin < C; c_begin += c_per_cta) {
      const int c_end = std::min(C, c_begin + c_per_cta);
      const float* roi = rois + (size_t)k * 5;
      const int level = levels ? levels[k] : 0;
      const int b = (int)roi[0];
      if (level < 0 || level >= L || b < 0 || b >= N) continue;
      const int H = Hs[level], W = Ws[level];
      const size_t HW = (size_t)H * W;
      float* grad_b = grads[level] + ((size_t)b * C + c_begin) * HW;
      const float* dout_k = dout + ((size_t)k * C + c_begin) * P;
      const RtGeom g = rt_geom_bwd(roi, scales[level], sampling_ratio, PH, PW);
      RtAxisPlan py, px;
      const int ok0 = rt_build_axis_plan(&py, g.sh, g.bh, PH, g.gh, H);
      const int ok1 = rt_build_axis_plan(&px, g.sw, g.bw, PW, g.gw, W);
      const int tiled = ok0 && ok1;
      const int R = py.n, tw = px.n;
      if (tiled && (R == 0 || tw == 0)) continue;
      const int twp = tw | 1;
      int cc_max = tiled ? (vs_floats / (PH * twp)) : 0;
      if (cc_max > CC_CAP) cc_max = CC_CAP;
      cc_max &= ~3;
      const int count = g.gh * g.gw;
      if (cc_max < 4) {
        ++n_direct;
        const int total = (c_end - c_begin) * P;
        for (int i = 0; i < total; ++i) {
          const int c = i / P, rem = i - c * P, ph = rem / PW, pw = rem - ph * PW;
          float* plane = grad_b + (size_t)c * HW;
          const float top = dout_k[i];
          for (int iy = 0; iy < g.gh; ++iy) {
            int ylo, yhi;
            float hy, ly;
            const int vy = rt_axis_tap(g.sh, g.bh, ph, iy, g.gh, H, &ylo, &yhi, &hy, &ly);
            for (int ix = 0; ix < g.gw; ++ix) {
              int xlo, xhi;
              float hx, lx;
              const int vx = rt_axis_tap(g.sw, g.bw, pw, ix, g.gw, W, &xlo, &xhi, &hx, &lx);
              if (!(vy && vx)) continue;
              plane[ylo * W + xlo] += top * (hy * hx) / (float)count;
              plane[ylo * W + xhi] += top * (hy * lx) / (float)count;
              plane[yhi * W + xlo] += top * (ly * hx) / (float)count;
              plane[yhi * W + xhi] += top * (ly * lx) / (float)count;
            }
          }
        }
        continue;
      }
      const float inv = rt_exact_inv(count);
      float* tile0 = grad_b + (size_t)py.lo0 * W + px.lo0;
      for (int c0 = 0; c0 < c_end - c_begin; c0 += cc_max) {
        const int cc = std::min(cc_max, c_end - c_begin - c0);
        std::memcpy(d_s.data(), dout_k + (size_t)c0 * P, sizeof(float) * (size_t)cc * P);
        std::fill(Hbuf.begin(), Hbuf.end(), 1e30f);
        {
          const int nq = (cc + U2 - 1) / U2, items = nq * PH;
          for (int it = 0; it < items; ++it) {
            const int q = it / PH, ph = it - q * PH;
            int nvalid = (cc - q + nq - 1) / nq;
            if (nvalid > U2) nvalid = U2;
            DoutSource<U2> src{d_s.data() + (size_t)(q * PH + ph) * PW, nq * PH * PW, nvalid};
            HEmit<U2> emit{Hbuf.data() + (size_t)(q * PH + ph) * twp, nq * PH * twp, nvalid};
            rt_sweep_bwd<U2>(px, src, emit);
          }
        }
        {
          const int nq = (cc + U1 - 1) / U1, items = nq * tw;
          for (int it = 0; it < items; ++it) {
            const int q = it / tw, x = it - q * tw;
            int nvalid = (cc - q + nq - 1) / nq;
            if (nvalid > U1) nvalid = U1;
            HSource<U1> src{Hbuf.data() + (size_t)q * PH * twp + x, twp, nq * PH * twp, nvalid};
            GradEmit<U1> emit{tile0 + (size_t)(c0 + q) * HW + x, (size_t)nq * HW, W, nvalid, inv,
                              (float)count};
            rt_sweep_bwd<U1>(py, src, emit);
          }
        }
      }
    }
  }
  return n_direct;
}

}  // extern "C"
