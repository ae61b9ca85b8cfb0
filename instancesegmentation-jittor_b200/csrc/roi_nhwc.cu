// roi_nhwc.cu -- channels-last (NHWC) entry of the ROIAlign / Pooler forward (north_star: "vectorised
// NHWC-aware bilinear gathers"; SURVEY.md 7 "Layout": the reference layout is NCHW, this is an
// additional fast path for callers whose backbone runs channels-last).
//
// Reference semantics: det/layers/roi_align.py:116-169 (the same per-sample arithmetic; rt_axis_tap).
// Features [N, H_l, W_l, C] fp32; output [K, C, PH, PW] (the reference's layout, so that the heads
// that follow are unchanged).
//
// CTA = (RoI, group of 32 channels), 256 threads = 8 channel quads x 32 bin slots.  Every tap of a
// sample is ONE 16-byte load of 4 channels; the 8 quads of a bin read 128 contiguous bytes, so the
// global / L1 traffic is whole sectors.  The per-axis sample tables (tap rows / columns, weights,
// validity) are built once per CTA in shared memory; the 32 x PH*PW output block of the CTA is
// contiguous in `out` and leaves through a shared-memory stage as full 128-byte lines.
// LSU-bound (16 vector loads per output quad); no staging of the feature tile is needed because
// neighbouring bins hit the same lines in L1.
#include "common.cuh"
#include "roi_levels.h"
#include "roi_tile.h"

namespace {

constexpr int RN_CG = 32;        // channels per CTA
constexpr int RN_THREADS = 256;
constexpr int RN_MAXS = 128;     // samples per axis (pooled * grid)

struct __align__(16) RnTap {
  int lo, hi;       // tap row / column (valid samples), or -1
  float wlo, whi;
};

__global__ void __launch_bounds__(RN_THREADS)
rn_pool_fwd_nhwc_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                        const int* __restrict__ levels, int PH, int PW, float sampling_ratio,
                        float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char rn_smem[];
  RnTap* ytab = reinterpret_cast<RnTap*>(rn_smem);
  RnTap* xtab = ytab + RN_MAXS;
  float* stage = reinterpret_cast<float*>(xtab + RN_MAXS);  // [RN_CG][P]
  const int tid = threadIdx.x;
  const int k = blockIdx.x;
  const int c0 = blockIdx.y * RN_CG;
  const int nch = min(RN_CG, C - c0);
  const int P = PH * PW;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  float* out_k = out + ((size_t)k * C + c0) * P;
  if (level < 0 || level >= lv.L || b < 0 || b >= N) {
    for (int i = tid; i < nch * P; i += RN_THREADS) out_k[i] = 0.f;
    return;
  }
  const int H = lv.H[level], W = lv.W[level];
  const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  const int SY = PH * g.gh, SX = PW * g.gw;
  const float* feat_b = lv.feat[level] + (size_t)b * H * W * C + c0;
  if (SY > RN_MAXS || SX > RN_MAXS || SY <= 0 || SX <= 0) {
    // adaptive sampling of a very large RoI: reference order, one thread per output
    for (int i = tid; i < nch * P; i += RN_THREADS) {
      const int c = i / P, rem = i - c * P;
      const int ph = rem / PW, pw = rem - ph * PW;
      float acc = 0.f;
      for (int iy = 0; iy < g.gh; ++iy) {
        int ylo, yhi;
        float hy, ly;
        const int vy = rt_axis_tap(g.sh, g.bh, ph, iy, g.gh, H, &ylo, &yhi, &hy, &ly);
        for (int ix = 0; ix < g.gw; ++ix) {
          int xlo, xhi;
          float hx, lx;
          const int vx = rt_axis_tap(g.sw, g.bw, pw, ix, g.gw, W, &xlo, &xhi, &hx, &lx);
          if (vy && vx) {
            const float v1 = feat_b[((size_t)ylo * W + xlo) * C + c], v2 = feat_b[((size_t)ylo * W + xhi) * C + c];
            const float v3 = feat_b[((size_t)yhi * W + xlo) * C + c], v4 = feat_b[((size_t)yhi * W + xhi) * C + c];
            acc += hy * hx * v1 + hy * lx * v2 + ly * hx * v3 + ly * lx * v4;
          }
        }
      }
      out_k[i] = acc / (float)(g.gh * g.gw);
    }
    return;
  }
  for (int s = tid; s < SY + SX; s += RN_THREADS) {
    const bool isy = s < SY;
    const int t = isy ? s : s - SY;
    const int grid = isy ? g.gh : g.gw;
    const int p = t / grid, i = t - p * grid;
    RnTap e;
    const int ok = isy ? rt_axis_tap(g.sh, g.bh, p, i, grid, H, &e.lo, &e.hi, &e.wlo, &e.whi)
                       : rt_axis_tap(g.sw, g.bw, p, i, grid, W, &e.lo, &e.hi, &e.wlo, &e.whi);
    if (!ok) {
      e.lo = e.hi = -1;
      e.wlo = e.whi = 0.f;
    }
    (isy ? ytab : xtab)[t] = e;
  }
  __syncthreads();
  const int cq = tid & 7, slot = tid >> 3;  // channel quad, bin slot
  const bool vec = ((C & 3) == 0) && ((reinterpret_cast<uintptr_t>(lv.feat[level]) & 15) == 0) && (cq * 4 + 3 < nch);
  const float inv = 1.0f / (float)(g.gh * g.gw);
  for (int bin = slot; bin < P; bin += RN_THREADS / 8) {
    const int ph = bin / PW, pw = bin - ph * PW;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int iy = 0; iy < g.gh; ++iy) {
      const RnTap ty = ytab[ph * g.gh + iy];
      if (ty.lo < 0) continue;
      const float* r0 = feat_b + (size_t)ty.lo * W * C;
      const float* r1 = feat_b + (size_t)ty.hi * W * C;
      for (int ix = 0; ix < g.gw; ++ix) {
        const RnTap tx = xtab[pw * g.gw + ix];
        if (tx.lo < 0) continue;
        const float w1 = ty.wlo * tx.wlo, w2 = ty.wlo * tx.whi, w3 = ty.whi * tx.wlo, w4 = ty.whi * tx.whi;
        if (vec) {
          const float4 v1 = __ldg(reinterpret_cast<const float4*>(r0 + (size_t)tx.lo * C) + cq);
          const float4 v2 = __ldg(reinterpret_cast<const float4*>(r0 + (size_t)tx.hi * C) + cq);
          const float4 v3 = __ldg(reinterpret_cast<const float4*>(r1 + (size_t)tx.lo * C) + cq);
          const float4 v4 = __ldg(reinterpret_cast<const float4*>(r1 + (size_t)tx.hi * C) + cq);
          acc.x = fmaf(w1, v1.x, fmaf(w2, v2.x, fmaf(w3, v3.x, fmaf(w4, v4.x, acc.x))));
          acc.y = fmaf(w1, v1.y, fmaf(w2, v2.y, fmaf(w3, v3.y, fmaf(w4, v4.y, acc.y))));
          acc.z = fmaf(w1, v1.z, fmaf(w2, v2.z, fmaf(w3, v3.z, fmaf(w4, v4.z, acc.z))));
          acc.w = fmaf(w1, v1.w, fmaf(w2, v2.w, fmaf(w3, v3.w, fmaf(w4, v4.w, acc.w))));
        } else {
          float* a = &acc.x;
          for (int u = 0; u < 4; ++u) {
            const int c = cq * 4 + u;
            if (c < nch)
              a[u] += w1 * r0[(size_t)tx.lo * C + c] + w2 * r0[(size_t)tx.hi * C + c] +
                      w3 * r1[(size_t)tx.lo * C + c] + w4 * r1[(size_t)tx.hi * C + c];
          }
        }
      }
    }
    // stage[c][bin]: the 8 quads of a bin hit banks 4*cq*P ... apart; P = 196 -> conflict-free enough
    stage[(cq * 4 + 0) * P + bin] = acc.x * inv;
    stage[(cq * 4 + 1) * P + bin] = acc.y * inv;
    stage[(cq * 4 + 2) * P + bin] = acc.z * inv;
    stage[(cq * 4 + 3) * P + bin] = acc.w * inv;
  }
  __syncthreads();
  for (int i = tid; i < nch * P; i += RN_THREADS) __stcs(out_k + i, stage[i]);
}

}  // namespace

int rn_forward_nhwc(const RtLevels& lv, int N, int C, const float* rois, const int* levels, int K, int PH,
                    int PW, float sampling_ratio, float* out, cudaStream_t st) {
  if (K == 0) return RP_OK;
  if (!rois || !out || N < 1 || C < 1 || K < 0 || PH < 1 || PW < 1) return RP_E_ARG;
  const size_t smem = 2 * sizeof(RnTap) * RN_MAXS + sizeof(float) * (size_t)RN_CG * PH * PW;
  if (smem > 200 * 1024) return RP_E_LIMIT;
  RP_SET_MAX_SMEM(rn_pool_fwd_nhwc_kernel, smem);
  dim3 grid((unsigned)K, (unsigned)((C + RN_CG - 1) / RN_CG));
  rn_pool_fwd_nhwc_kernel<<<grid, RN_THREADS, smem, st>>>(lv, N, C, rois, levels, PH, PW, sampling_ratio, out);
  RP_LAUNCH_CHECK();
  return RP_OK;
}
