// roi_fwd_mma.cuh -- ROIAlign / Pooler forward as two chained warp-level matrix products
// ("rm", RP_ROI_IMPL=7).  Included by roi_align.cu (inside its anonymous namespace, after
// rt3_roi_forward, which serves the RoIs this form does not take).
//
// Reference semantics: det/layers/roi_align.py:116-169.  Per axis the sample taps and weights
// factor (roi_tile.h), so with DENSE per-RoI weight matrices
//
//     out[c] (PH x PW) = Wy (PH x R) . T[c] (R x W) . Wx (W x PW)
//
// where T[c] is the RoI's feature tile of channel c, Wy / Wx hold the merged bilinear weights of
// every bin (<= 2 * grid nonzeros per row; the 1 / count average is folded into Wy).  Both
// products run on the legacy tensor path (mma.sync, SASS HMMA; measured on B200 with
// tools/mma_probe.cu: one m16n8k8 tf32 per 8 cycles per SM sub-partition), chained in
// registers: the accumulator fragment of the y-product IS the A fragment of the x-product
// (the contraction index may be permuted freely), so the intermediate never touches shared
// memory and every tile element is read from shared memory exactly once.
//
// Precision: fp32 operands are split x = hi + lo, hi rounded to tf32 (11 significant bits),
// and every product is evaluated as  hi.hi (tf32 m16n8k8)  +  [lo | hi] . [hi ; lo] (ONE bf16
// m16n8k16: the two correction terms concatenated along the contraction index; they only need
// ~8 significant bits).  Dropped term lo.lo ~ 2^-22; observed error against the oracle ~1e-6
// relative (tests: 1e-5 * |ref| + 1e-5).  Two HMMA per 16 x 8 x 8 block instead of three.
//
// CTA = (RoI, channel slice), 8 warps, 2 CTAs per SM (126 registers: the Wy fragments of up to 6 row
// blocks stay in registers).  Set-up (whole CTA): the sample tables of the tile-walk kernel, the two
// dense weight matrices in shared memory, the x-product B fragments of every (column block, pw block)
// pair in fragment order in shared memory (one 16-byte load per use), the y-product A fragments in
// registers.  Steady state (per warp, no CTA barrier): a warp owns channels w, w + 8, ...; the
// channel's tile [KY*8 rows][pitch] arrives in a private ring of 2-4 buffers (pitch = 8 or 24 mod 32
// floats: the B-fragment loads of a warp hit 32 distinct banks) either as ONE TMA box per channel
// (cp.async.bulk.tensor.3d on the warp's mbarrier; levels whose rows are 16-byte multiples; rows and
// columns beyond the map arrive as zeros) or through cp.async copies of the lanes (other levels; the
// rows / columns that pad the tile to whole 8 x 8 blocks are zero-filled once); for every 8-column
// block: y-product over the KY row blocks, split, x-product into the pw blocks whose band covers the
// column block; finally 8-byte streaming stores.
//
// Not taken (evaluated by the tile-walk body in the same launch): tiles above 48 rows or 48 (aligned)
// columns or whose double buffer exceeds the ring, PH or PW above 16, sample lattices above RT_SMAX.
// One semantic difference for NON-FINITE feature values: a dense product spreads an Inf / NaN inside
// the staged box to every bin of its row / column block, the reference only to the bins that sample it.
//
// Measured (B200, headline workload 16 x 1000 RoIs, 14x14x256; DESIGN.md 4.1, profiles/ncu_rm_fwd_r02.txt):
// 2.08 ms (MODE 0) / 2.30 ms (MODE 1) against 2.20 ms for the tile-walk kernel, max error against the
// oracle 7.0e-6 (MODE 0) / 2.4e-6 (MODE 1) where the tile walk has 7e-7.  The tensor pipe is 28 % busy
// and removing the HMMAs altogether changes the time by 4 %: splitting and packing the operands
// (VIADD, LOP3, FADD, F2FP: 140 warp instructions per (RoI, channel) of 307) costs more issue slots than
// the FFMA + LDS of the tile walk it replaces (122).  Opt-in; not the default.

// row pitch (floats) of a staged tile of KX column blocks: 8 or 24 mod 32, >= 8 * KX
#define RM_PITCH(KX) ((KX) == 1 ? 8 : ((KX) <= 3 ? 24 : ((KX) <= 5 ? 40 : 56)))
constexpr int RM_WARPS = 8;
constexpr int RM_THREADS = RM_WARPS * 32;
constexpr int RM_KMAX = 6;                   // 8-row / 8-column blocks per tile axis
constexpr int RM_WARP_FLOATS = 3072;         // private staging ring of one warp (12 KB)
constexpr int RM_WY_PITCH = RM_KMAX * 8 + 4, RM_WX_PITCH = 20;  // fragment loads of a warp hit distinct banks
constexpr int RM_W_FLOATS = 16 * RM_WY_PITCH + RM_KMAX * 8 * RM_WX_PITCH;  // Wy[16][52] | Wx[48][20]
constexpr int RM_BFRAG_U4 = RM_KMAX * 2 * 32;
// TMA descriptors (host-encoded, roi_align.cu rm_level_maps): per level the 3-D tensor (W, H, N*C
// planes) with one box per (row pitch class, row blocks): pitch x 8*KY x 1 plane, dense in shared
// memory, so the box IS the [KY*8][pitch] buffer the fragments are read from; out-of-range rows and
// columns arrive as zeros (no pad zeroing), in-range ones beyond the tile meet zero weights.
constexpr int RM_TMA_LEVELS = 5;
constexpr int RM_NPITCH = 4;
struct RmMaps {
  CUtensorMap m[RM_TMA_LEVELS][RM_NPITCH][RM_KMAX];
  int ok[RM_TMA_LEVELS];
};
#define RM_PITCH_CLASS(KX) ((KX) == 1 ? 0 : ((KX) <= 3 ? 1 : ((KX) <= 5 ? 2 : 3)))  // x-product B fragments: [column block][pw block][lane] x 16 bytes

__device__ __forceinline__ void rm_mma_tf32(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                            uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void rm_mma_bf16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                            uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// MODE 0: x = hi + lo with hi = x ROUNDED to tf32 (11 significant bits, half away), corrections in
//         ONE bf16 m16n8k16 (two HMMA per block).  MODE 1: hi = x TRUNCATED to tf32 (what the tensor
//         core does to a raw fp32 operand), corrections as two more tf32 m16n8k8 (classic 3xTF32,
//         three HMMA per block, no packing).  lo is exact in both.  The tensor core ignores the low
//         13 bits of a tf32 operand, so hi is passed unmasked.
// MODE 2: ONE tf32 product, no split (~1e-3 relative: NOT within the 1e-5 bar; a measurement of what the
//         fp32 emulation costs, never selected unless RP_ROI_MMA_MODE=2 is set explicitly).
template <int MODE>
__device__ __forceinline__ void rm_split(float x, uint32_t& hi, float& lo) {
  hi = MODE == 0 ? __float_as_uint(x) + 0x1000u : __float_as_uint(x);
  lo = MODE == 2 ? 0.f : x - __uint_as_float(hi & 0xffffe000u);
}
// {lower 16 bits = bf16(a), upper 16 bits = bf16(b)}
__device__ __forceinline__ uint32_t rm_pack_bf16(float a, float b) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}
// correction operands of the B side for an element pair (x0, x1) with low parts (l0, l1)
template <int MODE>
__device__ __forceinline__ void rm_corr_b(float x0, float x1, float l0, float l1, uint32_t (&c)[2]) {
  if (MODE == 0) {
    c[0] = rm_pack_bf16(x0, x1);  // k-slots 2q, 2q+1: meet lo(A)
    c[1] = rm_pack_bf16(l0, l1);  // k-slots 2q+8, 2q+9: meet A
  } else if (MODE == 2) {
    c[0] = c[1] = 0u;
  } else {
    c[0] = __float_as_uint(l0);
    c[1] = __float_as_uint(l1);
  }
}
// One 16 x 8 x 8 block product d += A . B in fp32-equivalent precision: A = (ah, ac), B = (bh, bc).
template <int MODE>
__device__ __forceinline__ void rm_block(float (&d)[4], const uint32_t (&ah)[4], const uint32_t (&ac)[4],
                                         const uint32_t (&bh)[2], const uint32_t (&bc)[2]) {
  rm_mma_tf32(d, ah[0], ah[1], ah[2], ah[3], bh[0], bh[1]);
  if (MODE == 0) {
    rm_mma_bf16(d, ac[0], ac[1], ac[2], ac[3], bc[0], bc[1]);
  } else if (MODE == 1) {
    rm_mma_tf32(d, ac[0], ac[1], ac[2], ac[3], bh[0], bh[1]);  // lo(A) . hi(B)
    rm_mma_tf32(d, ah[0], ah[1], ah[2], ah[3], bc[0], bc[1]);  // hi(A) . lo(B)
  }
}
__device__ __forceinline__ void rm_tma_issue(uint32_t dst, const void* tmap, int x, int y, int z, uint32_t bar,
                                             uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void rm_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > (1u << 22)) __trap();  // a copy that never lands must not hang the GPU
  }
}
__device__ __forceinline__ float rm_lds(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}

// Per-lane staging constants of one RoI: the 32 lanes of a warp form rpw rows of vpr vectors and step
// down the tile rpw rows at a time (ncopy copies per lane and channel, no division in the loop).
struct RmStage {
  int ncopy;       // copies of this lane per channel (0: idle lane)
  size_t goff;     // vrow * W + col
  size_t gstep;    // rpw * W
  uint32_t soff;   // byte offset of (vrow, col) in a channel buffer
  uint32_t sstep;  // rpw * pitch, bytes
};

template <int VEC>
__device__ __forceinline__ void rm_stage_channel(uint32_t dst, const float* __restrict__ src,
                                                 const RmStage& s) {
  const float* gp = src + s.goff;
  uint32_t sp = dst + s.soff;
#pragma unroll 1
  for (int n = s.ncopy; n > 0; --n) {
    if (VEC == 4)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sp), "l"(gp) : "memory");
    else if (VEC == 2)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sp), "l"(gp) : "memory");
    else
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sp), "l"(gp) : "memory");
    gp += s.gstep;
    sp += s.sstep;
  }
}

__device__ __forceinline__ void rm_stage_any(int vecw, uint32_t dst, const float* __restrict__ src,
                                             const RmStage& s) {
  if (vecw == 4)
    rm_stage_channel<4>(dst, src, s);
  else if (vecw == 2)
    rm_stage_channel<2>(dst, src, s);
  else
    rm_stage_channel<1>(dst, src, s);
}

// What the channel loop needs besides its fragments (CTA-uniform unless noted).
struct RmRoi {
  const float* src;  // &feat[b][c_begin][y0][xa0]
  float* out;        // &out[k][c_begin][0][0]
  size_t HW;
  int nch, R, W, vecw, P, PW;
  uint32_t chp4;     // channel-buffer pitch in bytes
  int nst;           // ring depth
  int ablate;        // timing experiments (results wrong): 1 = stage the first channels only, 2 = no stores
  uint32_t ring;     // this warp's ring (shared address)           [per warp]
  uint32_t wy;       // shared address of Wy[16][RM_WY_PITCH]
  uint32_t bfrag;    // shared address of the x-product B fragments (this lane's slot of pair 0)
  uint32_t band;     // bit xb*2+pb: the (column block, pw block) pair has a nonzero weight
  const void* tmap;  // TMA descriptor of this tile class, or nullptr: cp.async staging
  int tx, ty, tz;    // box origin: column, row, plane of this warp's first channel
  uint32_t bars;     // this warp's mbarriers, one per ring buffer (shared address)           [per warp]
};

// Channels a warp works on at once.  Two (for tiles small enough for four ring buffers and a second set
// of accumulators: KY <= 4 && 4 * KY * 8 * RM_PITCH(KX) <= RM_WARP_FLOATS) was measured SLOWER on the
// headline workload (2.33 against 2.13 ms), so one; the loop is written for any U.
#ifndef RM_UNROLL2
#define RM_UNROLL2 0
#endif
__host__ __device__ constexpr int rm_unroll(int KY, int KX) {
  return (RM_UNROLL2 && KY <= 4 && 4 * KY * 8 * RM_PITCH(KX) <= RM_WARP_FLOATS) ? 2 : 1;
}

// The steady state of one warp for a tile of KY x KX blocks.
template <int MODE, int KY, int KX>
__device__ __forceinline__ void rm_channels(const RmRoi& r, const RmStage& st) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, q = lane & 3;

  // ---- constant fragments (always split with rounding) ----------------------------------------
  // y-product A = Wy (16 x 8 per row block): tf32 {a0..a3} = (g, q), (g+8, q), (g, q+4), (g+8, q+4).
  // Correction operand, MODE 0 (bf16 m16n8k16): k-slots 2q, 2q+1 = lo of columns q, q+4 (they meet
  // the tile values), slots 2q+8, 2q+9 = the weights themselves (they meet lo of the tile values);
  // MODE 1: the four lo parts in tf32 fragment order.
  uint32_t ah[KY][4], ab[KY][4];
#pragma unroll
  for (int yb = 0; yb < KY; ++yb) {
    const uint32_t base = r.wy + 4u * (uint32_t)(g * RM_WY_PITCH + yb * 8 + q);
    const float w00 = rm_lds(base), w10 = rm_lds(base + 4u * 8 * RM_WY_PITCH);
    const float w01 = rm_lds(base + 16u), w11 = rm_lds(base + 4u * 8 * RM_WY_PITCH + 16u);
    float l00, l10, l01, l11;
    rm_split<0>(w00, ah[yb][0], l00);
    rm_split<0>(w10, ah[yb][1], l10);
    rm_split<0>(w01, ah[yb][2], l01);
    rm_split<0>(w11, ah[yb][3], l11);
    if (MODE == 0) {
      ab[yb][0] = rm_pack_bf16(l00, l01);
      ab[yb][1] = rm_pack_bf16(l10, l11);
      ab[yb][2] = rm_pack_bf16(w00, w01);
      ab[yb][3] = rm_pack_bf16(w10, w11);
    } else {
      ab[yb][0] = __float_as_uint(l00);
      ab[yb][1] = __float_as_uint(l10);
      ab[yb][2] = __float_as_uint(l01);
      ab[yb][3] = __float_as_uint(l11);
    }
  }
  // x-product B = Wx: built once per CTA in shared memory (rm_build_bfrags), one 16-byte load per
  // (column block, pw block) pair inside the band.
  const uint32_t band = r.band;

  // ---- output addressing: (ph, pw) = (g | g+8, pb*8 + 2q | +1) ------------------------------
  const int PW = r.PW;
  const int PH = r.P / PW;
  const bool pair = (PW & 1) == 0;  // 8-byte stores need even rows
  const int ooff0 = g * PW + 2 * q, ooff1 = ooff0 + 8 * PW;  // pw block 1: + 8 floats
  uint32_t okm = 0;  // bit 2 * (pb * 2 + h): first element in range, bit 2 * (..) + 1: second
#pragma unroll
  for (int pb = 0; pb < 2; ++pb)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int ph = g + 8 * h, pw = pb * 8 + 2 * q;
      if (ph < PH && pw < PW) okm |= 1u << (2 * (pb * 2 + h));
      if (ph < PH && pw + 1 < PW) okm |= 2u << (2 * (pb * 2 + h));
    }

  // ---- channel ring: U channels per iteration (independent instruction streams for the scheduler) ----
  constexpr uint32_t pitch4 = 4u * RM_PITCH(KX);
  constexpr int U = rm_unroll(KY, KX);
  const int nst = r.nst;  // ring slots of U buffers each
  const int nmine = (r.nch - warp + RM_WARPS - 1) / RM_WARPS;  // channels warp, warp + 8, ...
  const int niter = (nmine + U - 1) / U;
  const float* gnext = r.src + (size_t)warp * r.HW;  // next channel to stage
  const size_t cstep = (size_t)RM_WARPS * r.HW;
  const uint32_t ring_end = r.ring + (uint32_t)(nst * U) * r.chp4;
  uint32_t snext = r.ring;  // buffer of the next channel to stage
  int staged = 0;           // channels staged so far
  const bool tma = r.tmap != nullptr;
  int znext = r.tz;           // plane of the next channel to stage (TMA)
  uint32_t bnext = r.bars;    // its mbarrier
  uint32_t bcur = r.bars;     // mbarrier of the next buffer to consume
  const uint32_t bars_end = r.bars + 8u * (uint32_t)(nst * U);
  uint32_t phase = 0;         // parity to wait for on bcur (flips every trip round the ring)
  for (int sl = 0; sl < nst - 1; ++sl) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (staged < nmine) {
        if (!tma)
          rm_stage_any(r.vecw, snext, gnext, st);
        else if (lane == 0)
          rm_tma_issue(snext, r.tmap, r.tx, r.ty, znext, bnext, r.chp4);
      }
      ++staged;
      gnext += cstep;
      snext += r.chp4;
      znext += RM_WARPS;
      bnext += 8u;
    }
    if (!tma) asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // B-fragment addresses inside a buffer: rows q and q+4 of the row block, column g of the column block
  const uint32_t frag = 4u * (uint32_t)g + (uint32_t)q * pitch4;
  uint32_t scur = r.ring + frag;
  const uint32_t scur_end = ring_end + frag;
  float* optr = r.out + (size_t)warp * r.P;
  const size_t ostep = (size_t)RM_WARPS * r.P;
  int left = nmine;  // channels not yet computed
#pragma unroll 1
  for (int i = 0; i < niter; ++i) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (staged < nmine && !(r.ablate & 1)) {
        if (!tma)
          rm_stage_any(r.vecw, snext, gnext, st);
        else if (lane == 0)
          rm_tma_issue(snext, r.tmap, r.tx, r.ty, znext, bnext, r.chp4);
      }
      ++staged;
      gnext += cstep;
      snext += r.chp4;
      znext += RM_WARPS;
      bnext += 8u;
    }
    if (snext == ring_end) {
      snext = r.ring;
      bnext = r.bars;
    }
    if (!tma) {
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (nst == 2)
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      else if (nst == 3)
        asm volatile("cp.async.wait_group 2;" ::: "memory");
      else
        asm volatile("cp.async.wait_group 3;" ::: "memory");
      __syncwarp();
    } else if (!(r.ablate & 1) || i < nst - 1) {
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (u < left) rm_mbar_wait(bcur + 8u * u, phase);
    }
    bcur += 8u * U;
    if (bcur == bars_end) {
      bcur = r.bars;
      phase ^= 1u;
    }
    const uint32_t t0 = scur;
    scur += (uint32_t)U * r.chp4;
    if (scur == scur_end) scur = r.ring + frag;

    float o[U][2][4];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int pb = 0; pb < 2; ++pb)
#pragma unroll
        for (int e = 0; e < 4; ++e) o[u][pb][e] = 0.f;
#pragma unroll
    for (int xb = 0; xb < KX; ++xb) {
      float v[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int e = 0; e < 4; ++e) v[u][e] = 0.f;
#pragma unroll
        for (int yb = 0; yb < KY; ++yb) {
          const uint32_t a = t0 + (uint32_t)u * r.chp4 + 32u * xb + (uint32_t)(yb * 8) * pitch4;
          const float e0 = rm_lds(a), e1 = rm_lds(a + 4u * pitch4);
          uint32_t th[2], tc[2];
          float l0, l1;
          rm_split<MODE>(e0, th[0], l0);
          rm_split<MODE>(e1, th[1], l1);
          rm_corr_b<MODE>(e0, e1, l0, l1, tc);
          rm_block<MODE>(v[u], ah[yb], ab[yb], th, tc);
        }
      }
      // the B fragments of this column block (shared by the U channels)
      uint32_t bh[2][2], bc[2][2];
#pragma unroll
      for (int pb = 0; pb < 2; ++pb)
        if (band & (1u << (xb * 2 + pb)))
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(bh[pb][0]), "=r"(bh[pb][1]), "=r"(bc[pb][0]), "=r"(bc[pb][1])
                       : "r"(r.bfrag + 512u * (uint32_t)(xb * 2 + pb)));
#pragma unroll
      for (int u = 0; u < U; ++u) {
        // v = V[ph = g | g+8][column 2q | 2q+1 of block xb] -> A operand of the x-product
        // (tf32 fragment order a0..a3 = v0, v2, v1, v3)
        uint32_t vh[4], vc[4];
        float vl[4];
        rm_split<MODE>(v[u][0], vh[0], vl[0]);
        rm_split<MODE>(v[u][2], vh[1], vl[1]);
        rm_split<MODE>(v[u][1], vh[2], vl[2]);
        rm_split<MODE>(v[u][3], vh[3], vl[3]);
        if (MODE == 0) {
          vc[0] = rm_pack_bf16(vl[0], vl[2]);
          vc[1] = rm_pack_bf16(vl[1], vl[3]);
          vc[2] = rm_pack_bf16(v[u][0], v[u][1]);
          vc[3] = rm_pack_bf16(v[u][2], v[u][3]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) vc[e] = __float_as_uint(vl[e]);
        }
#pragma unroll
        for (int pb = 0; pb < 2; ++pb)
          if (band & (1u << (xb * 2 + pb))) rm_block<MODE>(o[u][pb], vh, vc, bh[pb], bc[pb]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u < left) {
        float* p0 = optr + ooff0;
        float* p1 = optr + ooff1;
        if (r.ablate & 2) {
          if (o[u][0][0] + o[u][0][1] + o[u][0][2] + o[u][0][3] + o[u][1][0] + o[u][1][1] + o[u][1][2] + o[u][1][3] == 12345.678f) *p0 = 0.f;
        } else if (pair) {
          if (okm & 0x01u) asm volatile("st.global.cs.v2.f32 [%0], {%1, %2};" ::"l"(p0), "f"(o[u][0][0]), "f"(o[u][0][1]) : "memory");
          if (okm & 0x04u) asm volatile("st.global.cs.v2.f32 [%0], {%1, %2};" ::"l"(p1), "f"(o[u][0][2]), "f"(o[u][0][3]) : "memory");
          if (okm & 0x10u) asm volatile("st.global.cs.v2.f32 [%0+32], {%1, %2};" ::"l"(p0), "f"(o[u][1][0]), "f"(o[u][1][1]) : "memory");
          if (okm & 0x40u) asm volatile("st.global.cs.v2.f32 [%0+32], {%1, %2};" ::"l"(p1), "f"(o[u][1][2]), "f"(o[u][1][3]) : "memory");
        } else {
#pragma unroll
          for (int pb = 0; pb < 2; ++pb)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float* p = (h ? p1 : p0) + 8 * pb;
              const int sh = 2 * (pb * 2 + h);
              if (okm & (1u << sh)) asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(o[u][pb][2 * h]) : "memory");
              if (okm & (2u << sh)) asm volatile("st.global.cs.f32 [%0+4], %1;" ::"l"(p), "f"(o[u][pb][2 * h + 1]) : "memory");
            }
        }
      }
      optr += ostep;
    }
    left -= U;
    __syncwarp();  // every lane is done reading the buffers that the next iteration's copies overwrite
  }
  if (!tma) asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int MODE, int KY>
__device__ __forceinline__ void rm_dispatch_kx(int KX, const RmRoi& r, const RmStage& st) {
  switch (KX) {
    case 1: rm_channels<MODE, KY, 1>(r, st); break;
    case 2: rm_channels<MODE, KY, 2>(r, st); break;
    case 3: rm_channels<MODE, KY, 3>(r, st); break;
    case 4: rm_channels<MODE, KY, 4>(r, st); break;
    case 5: rm_channels<MODE, KY, 5>(r, st); break;
    default: rm_channels<MODE, KY, 6>(r, st); break;
  }
}

// The x-product B fragments of every (column block, pw block) pair, written once per CTA: pair p is
// built by warp p % 8; k-slot q <-> column 2q, slot q+4 <-> column 2q+1 of the block (the order the
// y-product's accumulator holds them); n = g.  Returns nothing; *band_s collects the nonzero pairs.
template <int MODE>
__device__ __forceinline__ void rm_build_bfrags(const float* wx, uint4* bfrag, int KX, int* band_s) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  for (int p = warp; p < 2 * KX; p += RM_WARPS) {
    const int xb = p >> 1, pb = p & 1;
    const float* w = wx + (xb * 8 + 2 * q) * RM_WX_PITCH + pb * 8 + g;
    const float w0 = w[0], w1 = w[RM_WX_PITCH];
    uint32_t bh[2], bc[2];
    float l0, l1;
    rm_split<0>(w0, bh[0], l0);
    rm_split<0>(w1, bh[1], l1);
    rm_corr_b<MODE>(w0, w1, l0, l1, bc);
    bfrag[p * 32 + lane] = make_uint4(bh[0], bh[1], bc[0], bc[1]);
    if (__any_sync(0xffffffffu, w0 != 0.f || w1 != 0.f) && lane == 0) atomicOr(band_s, 1 << p);
  }
}

template <int MODE>
__global__ void __launch_bounds__(RM_THREADS, 2)
rm_pool_fwd_kernel(RtLevels lv, int N, int C, const float* __restrict__ rois,
                   const int* __restrict__ levels, int K, int PH, int PW, float sampling_ratio,
                   float* __restrict__ out, int c_per_cta, int tile_floats, int gpb, int gw, int ablate,
                   const __grid_constant__ RmMaps maps) {
  extern __shared__ __align__(16) unsigned char rt_smem[];
  // layout: [tile-walk tables: ytab | xw | xo] [Wy | Wx] [8 warp rings]
  RtYSample* ytab = reinterpret_cast<RtYSample*>(rt_smem);
  RtW* xw_s = reinterpret_cast<RtW*>(ytab + RT_SMAX);
  short* xo_s = reinterpret_cast<short*>(xw_s + RT_SMAX);
  float* wmat = reinterpret_cast<float*>(rt_smem + RT3_TABLE_BYTES);
  uint4* bfrag = reinterpret_cast<uint4*>(wmat + RM_W_FLOATS);
  float* rings = reinterpret_cast<float*>(bfrag + RM_BFRAG_U4);
  rings += ((128u - (rp_smem_addr(rings) & 127u)) & 127u) >> 2;  // TMA boxes land on 128-byte boundaries
  __shared__ int s_ext[5];  // y: min lo, max lo; x: min lo, max lo; band bits
  __shared__ __align__(8) uint64_t s_bar[RM_WARPS][4];  // per warp: one mbarrier per ring buffer (TMA staging)

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
  const int SY = PH * g.gh, SX = PW * g.gw;
  const bool fits = (SY <= RT_SMAX) && (SX <= RT_SMAX) && SY > 0 && SX > 0 && H >= 2 && W >= 2 &&
                    PH <= 16 && PW <= 16;

  // zero the weight matrices while the taps are computed
  for (int i = tid; i < RM_W_FLOATS / 4; i += RM_THREADS)
    reinterpret_cast<float4*>(wmat)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid < 4) s_ext[tid] = (tid & 1) ? -1 : 0x7fffffff;
  if (tid == 4) s_ext[4] = 0;
  __syncthreads();
  if (fits) {
    for (int s = tid; s < SY + SX; s += RM_THREADS) {
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
        ytab[t].row = ok ? lo : -1;
      } else {
        xw_s[t].lo = wlo;
        xw_s[t].hi = whi;
        xo_s[t] = (short)(ok ? lo : -1);
      }
    }
  }
  __syncthreads();
  const int y0 = s_ext[0], yl = s_ext[1], x0 = s_ext[2], xl = s_ext[3];
  if (fits && (yl < 0 || xl < 0)) {  // no sample inside the map on some axis
    rt_zero_fill(out_k, (size_t)nch * P);
    return;
  }
  const int R = yl - y0 + 2;
  const uintptr_t base_addr = reinterpret_cast<uintptr_t>(lv.feat[level]);
  const int vecw = (((W & 3) == 0) && ((base_addr & 15) == 0)) ? 4
                   : (((W & 1) == 0) && ((base_addr & 7) == 0)) ? 2 : 1;
  const int xa0 = x0 & ~(vecw - 1);
  const int twa = ((xl + 1) / vecw - xa0 / vecw + 1) * vecw;
  const int KY = (R + 7) >> 3, KX = (twa + 7) >> 3;
  const int pitch = RM_PITCH(KX);
  const int chp = KY * 8 * pitch;
  if (!fits || KY > RM_KMAX || KX > RM_KMAX || 2 * chp > RM_WARP_FLOATS || twa / vecw > 32) {
    __syncthreads();  // the tile-walk body rebuilds its tables over the same shared memory
    rt3_roi_forward<2, true>(rt_smem, lv, N, C, rois, levels, PH, PW, sampling_ratio, out, c_per_cta,
                             tile_floats, gpb, gw);
    return;
  }

  // ---- dense weight matrices: one thread per output row / column, samples in order --------------
  const float inv = 1.0f / (float)(g.gh * g.gw);
  if (tid < PH) {
    float* row = wmat + tid * RM_WY_PITCH;
    for (int i = 0; i < g.gh; ++i) {
      const RtYSample e = ytab[tid * g.gh + i];
      if (e.row >= 0) {
        row[e.row - y0] += e.wlo * inv;
        row[e.row - y0 + 1] += e.whi * inv;
      }
    }
  } else if (tid >= 32 && tid < 32 + PW) {
    const int pw = tid - 32;
    float* col = wmat + 16 * RM_WY_PITCH + pw;
    for (int i = 0; i < g.gw; ++i) {
      const int lo = xo_s[pw * g.gw + i];
      const RtW w = xw_s[pw * g.gw + i];
      if (lo >= 0) {
        col[(lo - xa0) * RM_WX_PITCH] += w.lo;
        col[(lo - xa0 + 1) * RM_WX_PITCH] += w.hi;
      }
    }
  }
  __syncthreads();
  rm_build_bfrags<MODE>(wmat + 16 * RM_WY_PITCH, bfrag, KX, &s_ext[4]);
  __syncthreads();

  RmRoi r;
  r.src = lv.feat[level] + ((size_t)b * C + c_begin) * HW + (size_t)y0 * W + xa0;
  r.out = out_k;
  r.HW = HW;
  r.nch = nch;
  r.R = R;
  r.W = W;
  r.vecw = vecw;
  r.P = P;
  r.PW = PW;
  r.ablate = ablate;
  r.band = (uint32_t)s_ext[4];
  const bool tma = level < RM_TMA_LEVELS && maps.ok[level] && !(ablate & 4);
  r.tmap = tma ? (const void*)&maps.m[level][RM_PITCH_CLASS(KX)][KY - 1] : nullptr;
  r.tx = xa0;
  r.ty = y0;
  r.tz = b * C + c_begin + (tid >> 5);
  r.bars = rp_smem_addr(&s_bar[tid >> 5][0]);
  if (tma && (tid & 31) == 0) {
    for (int i = 0; i < 4; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(r.bars + 8u * i) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  r.chp4 = 4u * (uint32_t)chp;
  const int U = rm_unroll(KY, KX);
  int nst = RM_WARP_FLOATS / (chp * U);  // ring slots of U buffers
  r.nst = nst > 4 / U ? 4 / U : nst;  // at most 4 buffers (and mbarriers) per warp
  r.ring = rp_smem_addr(rings + (tid >> 5) * RM_WARP_FLOATS);
  r.wy = rp_smem_addr(wmat);
  r.bfrag = rp_smem_addr(bfrag + (tid & 31));

  RmStage st;
  {
    const int lane = tid & 31;
    const int vpr = twa / vecw;  // <= 32 (checked above)
    const int rpw = 32 / vpr;
    const int vrow = lane / vpr;
    const int col = (lane - vrow * vpr) * vecw;
    st.ncopy = (vrow < rpw && vrow < R) ? (R - vrow + rpw - 1) / rpw : 0;
    st.goff = (size_t)vrow * W + col;
    st.gstep = (size_t)rpw * W;
    st.soff = 4u * (uint32_t)(vrow * pitch + col);
    st.sstep = 4u * (uint32_t)(rpw * pitch);
    // the rows / columns that pad the tile to whole blocks live in the last row / column block of
    // every ring buffer: zero them once (the copies never write them)
    const int RP = KY * 8, CP = KX * 8;
    for (int bf = 0; bf < (tma ? 0 : r.nst * U); ++bf) {
      float* bp = rings + (tid >> 5) * RM_WARP_FLOATS + bf * chp;
      for (int c = lane; c < CP; c += 32)
        for (int row = R; row < RP; ++row) bp[row * pitch + c] = 0.f;
      if (twa < CP) {
        const int c = twa + (lane & 7);
        if (c < CP)
          for (int row = lane >> 3; row < R; row += 4) bp[row * pitch + c] = 0.f;
      }
    }
    __syncwarp();
  }
  switch (KY) {
    case 1: rm_dispatch_kx<MODE, 1>(KX, r, st); break;
    case 2: rm_dispatch_kx<MODE, 2>(KX, r, st); break;
    case 3: rm_dispatch_kx<MODE, 3>(KX, r, st); break;
    case 4: rm_dispatch_kx<MODE, 4>(KX, r, st); break;
    case 5: rm_dispatch_kx<MODE, 5>(KX, r, st); break;
    default: rm_dispatch_kx<MODE, 6>(KX, r, st); break;
  }
}

constexpr size_t RM_SMEM_BYTES = RT3_TABLE_BYTES + sizeof(float) * (size_t)RM_W_FLOATS +
                                 sizeof(uint4) * (size_t)RM_BFRAG_U4 +
                                 sizeof(float) * (size_t)(RM_WARPS * RM_WARP_FLOATS) + 128;
static_assert(2 * (RM_SMEM_BYTES + 512 + 1024) <= 233472, "two CTAs per SM");
