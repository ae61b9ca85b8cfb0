// common.cuh -- small helpers shared by the kernels of libroipath_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/roipath.h"
#include "../../include/roipath_math.h"

#define RP_NUM_SMS 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

#define RP_CUDA_TRY(expr)                  \
  do {                                     \
    cudaError_t rp_err__ = (expr);         \
    if (rp_err__ != cudaSuccess) {         \
      (void)cudaGetLastError();            \
      return (int)rp_err__;                \
    }                                      \
  } while (0)

#define RP_LAUNCH_CHECK()                  \
  do {                                     \
    cudaError_t rp_err__ = cudaGetLastError(); \
    if (rp_err__ != cudaSuccess) return (int)rp_err__; \
  } while (0)

// Tuning knobs (RP_* environment variables): read ONCE per process, at their first use, into a table in
// capi.cu, so that the library's behaviour cannot change under a running program; rp_reload_env()
// (C ABI; tests and tuning tools) drops the table.  A knob that is not set costs one table lookup.
int rp_env_int(const char* name, int dflt);

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) only when the kernel needs more than has already been
// set for it on the current device (once per kernel in the steady state, not once per launch).  Keyed by
// the kernel's address (capi.cu), not by call site: several call sites may launch the same kernel.
int rp_smem_ensure(const void* kernel, int bytes);  // 0 or a cudaError_t
#define RP_SET_MAX_SMEM(kernel, bytes)                                                 \
  do {                                                                                 \
    const int rp_rc__ = rp_smem_ensure(reinterpret_cast<const void*>(kernel), (int)(bytes)); \
    if (rp_rc__) return rp_rc__;                                                        \
  } while (0)

static inline size_t rp_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__device__ __forceinline__ uint32_t rp_smem_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// Make this thread's generic-proxy shared-memory writes visible to the async proxy (TMA).
__device__ __forceinline__ void rp_fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// TMA 1-D bulk store shared -> global (SASS: UBLKCP).  dst and src 16-byte aligned,
// bytes a multiple of 16.  Issued by ONE thread; completion tracked by the bulk group.
__device__ __forceinline__ void rp_bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(rp_smem_addr(ssrc)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// Wait until every committed bulk store of this thread has finished READING shared memory.
__device__ __forceinline__ void rp_bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

__device__ __forceinline__ void rp_bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- mbarrier + TMA 1-D bulk load global -> shared (SASS: UBLKCP) -------------------------
__device__ __forceinline__ void rp_mbar_init(uint64_t* bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rp_smem_addr(bar)), "r"(arrivals)
               : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// One arrival that also announces `bytes` of asynchronous copies to wait for in this phase.
__device__ __forceinline__ void rp_mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rp_smem_addr(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void rp_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "RP_MBAR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra RP_MBAR_DONE;\n"
      "bra RP_MBAR_WAIT;\n"
      "RP_MBAR_DONE:\n"
      "}\n" ::"r"(rp_smem_addr(bar)),
      "r"(parity)
      : "memory");
}
// dst / src 16-byte aligned, bytes a multiple of 16; completion is counted on `bar`.
__device__ __forceinline__ void rp_bulk_load(void* sdst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          rp_smem_addr(sdst)),
      "l"(gsrc), "r"(bytes), "r"(rp_smem_addr(bar))
      : "memory");
}

// Streaming (evict-first) stores for write-once outputs.
__device__ __forceinline__ void rp_st_cs_u32(uint32_t* p, uint32_t v) { __stcs(p, v); }

// ---- TMA tensor load (cp.async.bulk.tensor, SASS UTMALDG): one 3-D box global -> shared --------
// dst 128-byte aligned; completion (box bytes, out-of-bounds parts included) counted on `bar`.
__device__ __forceinline__ void rp_tma_load_3d(void* sdst, const void* tmap, int x, int y, int z,
                                               uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(rp_smem_addr(sdst)),
      "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(rp_smem_addr(bar))
      : "memory");
}
// Bounded wait: traps instead of hanging the GPU if the expected bytes never arrive.
__device__ __forceinline__ void rp_mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
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
    if (spin > (1u << 24)) __trap();
  }
}
