// roi_fwd_plane.cuh -- whole-plane Pooler forward for the coarse FPN levels ("rp", RP_ROI_PLANES).
// Included by roi_align.cu (inside its anonymous namespace, after rt_fwd_item_fsm_s).
//
// Why: a timing ablation of the column-walk kernel puts the per-RoI tile copies at 0.6 of its 2.2 ms,
// and copying them under the walk (two buffers) does not hide them: it is the THROUGHPUT of 10 GB of
// overlapping tiles through L2 -> L1 -> shared memory, not their latency.  On the coarse
// levels the RoIs of one image cover the plane many times over (P4: 200 RoIs x ~21 x 21 pixels on a
// 50 x 84 plane = 21x, P5 29x), so there the PLANES are staged instead, once:
//
//   rp_tables_kernel   one warp per RoI: the sample tables of the column walk (det/layers/roi_align.py:
//                      14-60,150-165 per axis, exactly the values rt3_roi_forward builds), with ABSOLUTE
//                      plane rows / columns, written to global memory once instead of once per channel
//                      slice; key[k] = level * N + image for the RoIs this path takes, -1 otherwise.
//   rp_plane_fwd_kernel  CTA = (level, image, group of CG channels), 16 warps, 2 CTAs per SM: stages the
//                      CG planes (contiguous in NCHW) once, collects the RoIs of its (level, image) from
//                      key[], and its warps then work on their own: a warp takes a RoI, loads its tables
//                      into its private slot and walks every channel pair with rt_fwd_item_fsm_s -- the
//                      same walk, the same operations in the same order as the column-walk kernel, so the
//                      same values -- lanes = 2 channel pairs x 14 output columns.  No CTA barrier after the
//                      planes have landed, no copy per RoI.
//   the column-walk kernel skips the RoIs with key[k] >= 0.
//
// Taken: levels whose plane fits the CG budget (H * W <= RP_PLANE_MAX), sampling_ratio == 2, PH and PW <= 16,
// valid image index.  Everything else stays with the column-walk kernel.

constexpr int RP_WARPS = 16;
constexpr int RP_THREADS = RP_WARPS * 32;
constexpr int RP_PLANE_FLOATS = 16896;  // plane buffer of a CTA (66 KB): CG = RP_PLANE_FLOATS / (H * W) channels
constexpr int RP_PLANE_MAX = RP_PLANE_FLOATS / 2;  // a plane must allow a channel pair
constexpr int RP_LIST = 1024;           // RoIs of one (level, image) handled per round
// RpTab (the per-RoI sample tables) is defined in roi_align.cu, next to rt3_roi_forward, which reads it too.

__device__ __forceinline__ bool rp_level_planar(const RtLevels& lv, int level) {
  return level >= 0 && level < lv.L && lv.H[level] >= 2 && lv.W[level] >= 2 &&
         lv.H[level] * lv.W[level] <= RP_PLANE_MAX;
}

// One warp per RoI.
__global__ void __launch_bounds__(256)
rp_tables_kernel(RtLevels lv, int N, const float* __restrict__ rois, const int* __restrict__ levels, int K,
                 int PH, int PW, float sampling_ratio, RpTab* __restrict__ tabs, int* __restrict__ key,
                 int ymerge, int* __restrict__ wcnt, int nkeys) {
  const int lane = threadIdx.x & 31;
  const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (k >= K) return;
  const float* roi = rois + (size_t)k * 5;
  const int level = levels ? levels[k] : 0;
  const int b = (int)roi[0];
  RpTab* t = tabs + k;
  if (level < 0 || level >= lv.L || b < 0 || b >= N || lv.H[level] < 2 || lv.W[level] < 2) {
    if (lane == 0) {
      key[k] = -1;
      t->meta[0] = 2;  // no tables: the column walk decides on its own
    }
    return;
  }
  const bool planar = rp_level_planar(lv, level);  // whole-plane kernel; else only the tables are used
  ymerge = ymerge && planar;  // the column walk merges itself, where its items walk whole columns
  const int H = lv.H[level], W = lv.W[level];
  const RtGeom g = rt_geom_fwd(roi, lv.scale[level], sampling_ratio, PH, PW);
  const int SY = PH * 2, SX = PW * 2;  // the launcher guarantees sampling_ratio == 2
  // lane s < SY: y-sample s; every lane also x-sample s (SX <= 32)
  int ylo = 0, xlo = 0;
  float ywl = 0.f, ywh = 0.f, xwl = 0.f, xwh = 0.f;
  const int yok = lane < SY ? rt_axis_tap_pair(g.sh, g.bh, lane >> 1, lane & 1, 2, H, &ylo, &ywl, &ywh) : 0;
  const int xok = lane < SX ? rt_axis_tap_pair(g.sw, g.bw, lane >> 1, lane & 1, 2, W, &xlo, &xwl, &xwh) : 0;
  const int y0 = __reduce_min_sync(0xffffffffu, yok ? ylo : 0x7fffffff);
  const int yl = __reduce_max_sync(0xffffffffu, yok ? ylo : -1);
  const int x0 = __reduce_min_sync(0xffffffffu, xok ? xlo : 0x7fffffff);
  const int xl = __reduce_max_sync(0xffffffffu, xok ? xlo : -1);
  // wcnt[window][key]: how many RoIs of each (level, image) the window of RP_LIST keys holds, so that the
  // plane kernel only visits the windows that concern it
  if (lane == 0 && wcnt && planar) atomicAdd(&wcnt[(size_t)(k / RP_LIST) * nkeys + level * N + b], 1);
  if (yl < 0 || xl < 0) {  // no sample inside the map on some axis: zeros (roi_align.py:20)
    if (lane == 0) {
      t->meta[0] = 0;
      t->meta[2] = 0;
      t->meta[3] = 0;
      key[k] = planar ? level * N + b : -1;
    }
    return;
  }
  // out-of-range samples carry zero weights and park on the nearest row / column that has a sample
  const int yrow = (lane < SY) ? (yok ? ylo : (ylo == -1 ? y0 : yl)) : yl;
  const int xcol = xok ? xlo : (xlo == -1 ? x0 : xl);
  {
    // y-merge (see rt3_roi_forward): the two samples of a bin that share a row pair become one entry
    const int rown = __shfl_down_sync(0xffffffffu, yrow, 1);
    const float wlon = __shfl_down_sync(0xffffffffu, ywl, 1), whin = __shfl_down_sync(0xffffffffu, ywh, 1);
    const int m = ymerge && (lane < SY) && !(lane & 1) && (rown == yrow);
    const int mprev = __shfl_up_sync(0xffffffffu, m, 1);
    const bool keep = (lane < SY) && !((lane & 1) && mprev);
    const unsigned km = __ballot_sync(0xffffffffu, keep);
    const int pos = __popc(km & ((1u << lane) - 1u));
    // row of the entry before this one in the compacted sequence
    const int src = pos > 0 ? (int)__fns(km, 0, pos) : lane;
    const int prev = __shfl_sync(0xffffffffu, yrow, src);
    if (keep) {
      RtYSample e;
      e.wlo = m ? ywl + wlon : ywl;
      e.whi = m ? ywh + whin : ywh;
      e.row = yrow;
      e.flags = rt_ysample_flags(yrow, pos ? prev : yrow, (lane & 1) || m);
      t->y[pos] = e;
    }
  }
  if (lane < SX) {
    RtW w;
    w.lo = xwl;
    w.hi = xwh;
    t->xw[lane] = w;
    t->xo[lane] = (short)xcol;
  }
  // three-tap rows: the two x-samples of every column in the same or in adjacent plane columns
  const int xnext = __shfl_down_sync(0xffffffffu, xcol, 1);
  const int d = xnext - xcol;
  const int x3 = __all_sync(0xffffffffu, (lane >= SX) || (lane & 1) || d == 0 || d == 1);
  if (lane == 0) {
    t->meta[0] = 1;
    t->meta[1] = x3;
    t->meta[2] = y0 | (yl << 16);
    t->meta[3] = x0 | (xl << 16);
    key[k] = planar ? level * N + b : -1;
  }
}

__global__ void __launch_bounds__(RP_THREADS, 2)
rp_plane_fwd_kernel(RtLevels lv, int N, int C, const int* __restrict__ key, const RpTab* __restrict__ tabs,
                    int K, int PH, int PW, float* __restrict__ out, int max_groups, int tap3,
                    const int* __restrict__ wcnt, int nkeys) {
  extern __shared__ __align__(16) unsigned char rp_smem[];
  // layout: planes[RP_PLANE_FLOATS] | tables[2 * RP_WARPS] | list[RP_LIST]
  float* planes = reinterpret_cast<float*>(rp_smem);
  RpTab* slots = reinterpret_cast<RpTab*>(planes + RP_PLANE_FLOATS);
  int* list = reinterpret_cast<int*>(slots + 2 * RP_WARPS);
  __shared__ int s_n;
  __shared__ __align__(8) uint64_t s_bar;  // completion of the planes' bulk copy

  const int level = blockIdx.y;
  if (!rp_level_planar(lv, level)) return;
  const int H = lv.H[level], W = lv.W[level], HW = H * W;
  int CG = RP_PLANE_FLOATS / HW;
  if (CG > C) CG = C;
  CG &= ~1;  // channel pairs
  if (CG < 2) CG = C >= 2 ? 2 : 1;
  const int groups = (C + CG - 1) / CG;
  const int n = blockIdx.x / max_groups, cgp = blockIdx.x - n * max_groups;
  if (cgp >= groups || n >= N) return;
  const int c0 = cgp * CG;
  const int nch = min(CG, C - c0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int P = PH * PW;
  const int mykey = level * N + n;

  // ---- the planes of this CTA's channels: one contiguous run of the NCHW map, fetched by ONE TMA bulk
  //      copy (cp.async.bulk on an mbarrier, SASS UBLKCP) when the run is 16-byte aligned and a multiple of
  //      16 bytes (P4: 4 x 50 x 84, P5: 16 x 25 x 42 floats = 67 200 bytes both), else by cp.async ----------
  const float* src = lv.feat[level] + ((size_t)n * C + c0) * HW;
  const int total = nch * HW;
  const bool v16 = ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
  const bool bulk = v16 && (total & 3) == 0 && total * 4 < (1 << 20);
  const uint32_t planes_s = rp_smem_addr(planes);
  if (bulk) {
    if (tid == 0) {
      rp_mbar_init(&s_bar, 1);
      rp_mbar_arrive_expect_tx(&s_bar, (uint32_t)total * 4u);
      rp_bulk_load(planes, src, (uint32_t)total * 4u, &s_bar);
    }
  } else {
    const int n16 = v16 ? (total >> 2) : 0;
    for (int i = tid; i < n16; i += RP_THREADS)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(planes_s + 16u * (uint32_t)i), "l"(src + 4 * i) : "memory");
    for (int i = 4 * n16 + tid; i < total; i += RP_THREADS)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(planes_s + 4u * (uint32_t)i), "l"(src + i) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  }

  // lanes: 16 per channel pair (PW <= 16 of them walk a column each); the two halves of a warp take
  // different channel pairs of the SAME RoI, so both run the same sample sequence (no divergence)
  const int half = lane >> 4, pw = lane & 15;
  const bool active = pw < PW;
  RpTab* slot = slots + 2 * warp;  // two slots per warp

  bool landed = false;
  for (int base = 0; base < K; base += RP_LIST) {
    // ---- the RoIs of this (level, image) among the next RP_LIST keys ----------------------------
    if (wcnt && wcnt[(size_t)(base / RP_LIST) * nkeys + mykey] == 0) continue;  // CTA-uniform
    if (tid == 0) s_n = 0;
    __syncthreads();
    const int end = min(K, base + RP_LIST);
    for (int kb = base; kb < end; kb += RP_THREADS) {
      const int k = kb + tid;
      // gridDim.z > 1 (few images): RoI k belongs to the CTA with blockIdx.z == k % gridDim.z (the list
      // order depends on warp timing, so the deal is by RoI index, not by list position)
      const bool hit = k < end && key[k] == mykey && (k % (int)gridDim.z) == (int)blockIdx.z;
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (m) {
        int pos = 0;
        if (lane == 0) pos = atomicAdd(&s_n, __popc(m));
        pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1u));
        if (hit) list[pos] = k;
      }
    }
    if (!landed && !bulk) asm volatile("cp.async.wait_group 0;" ::: "memory");  // the planes
    __syncthreads();  // the list is complete; s_bar is initialised for every thread
    if (!landed && bulk) rp_mbar_wait_bounded(&s_bar, 0);
    landed = true;
    const int cnt = s_n;
    // ---- warps on their own from here -------------------------------------------------------
    // tables of the warp's first RoI, then always one RoI ahead (two slots per warp)
    auto fetch = [&](int i, int sl) {
      if (i < cnt) {
        const char* g = reinterpret_cast<const char*>(tabs + list[i]);
        const uint32_t d = rp_smem_addr(slot + sl);
        for (int j = lane; j < (int)(sizeof(RpTab) / 16); j += 32)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + 16u * (uint32_t)j), "l"(g + 16 * j) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int sl = 0;
    fetch(warp, 0);
    for (int i = warp; i < cnt; i += RP_WARPS, sl ^= 1) {
      const int k = list[i];
      fetch(i + RP_WARPS, sl ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
      const RpTab* tb = slot + sl;
      float* out_k = out + ((size_t)k * C + c0) * P;
      if (tb->meta[0] == 0) {
        for (int j = lane; j < nch * P; j += 32) out_k[j] = 0.f;
      } else if (active) {
        const uint32_t ytab_s = rp_smem_addr(tb->y);
        if (tap3 && tb->meta[1]) {
          for (int c = 2 * half; c < nch; c += 4)
            rt_fwd_item_fsm_s<2, true>(planes_s + 4u * (uint32_t)(c * HW), 4u * (uint32_t)HW, min(2, nch - c),
                                       ytab_s, tb->xw, tb->xo, 2, 4u * (uint32_t)W, pw, 0, PH,
                                       out_k + (size_t)c * P + pw, P, PW, 0.25f);
        } else {
          for (int c = 2 * half; c < nch; c += 4)
            rt_fwd_item_fsm_s<2>(planes_s + 4u * (uint32_t)(c * HW), 4u * (uint32_t)HW, min(2, nch - c), ytab_s,
                                 tb->xw, tb->xo, 2, 4u * (uint32_t)W, pw, 0, PH,
                                 out_k + (size_t)c * P + pw, P, PW, 0.25f);
        }
      }
      __syncwarp();  // the slot is refilled two RoIs later
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // the list is rebuilt by the next round
  }
  // no RoI of this (level, image): drain the copies before the CTA's shared memory is given away
  if (!landed && !bulk) asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (!landed && bulk) {
    __syncthreads();
    rp_mbar_wait_bounded(&s_bar, 0);
  }
}

constexpr size_t RP_SMEM_BYTES =
    sizeof(float) * (size_t)RP_PLANE_FLOATS + sizeof(RpTab) * (size_t)(2 * RP_WARPS) + sizeof(int) * (size_t)RP_LIST;
