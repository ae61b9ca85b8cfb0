// tools/mma_probe.cu -- issue rate of the legacy warp-level tensor path (mma.sync m16n8k8 tf32) on sm_100a.
// Decides whether the chained-MMA Pooler forward (roi_fwd_mma.cu) can be bound by something other than
// the tensor pipe.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe tools/mma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NACC>
__global__ void probe(float* out, int iters, long long* cyc) {
  uint32_t a[4], b[2];
  for (int i = 0; i < 4; ++i) a[i] = __float_as_uint(1.0f + threadIdx.x * 1e-3f + i);
  for (int i = 0; i < 2; ++i) b[i] = __float_as_uint(0.5f + threadIdx.x * 1e-3f + i);
  float acc[NACC][4];
  for (int j = 0; j < NACC; ++j)
    for (int i = 0; i < 4; ++i) acc[j][i] = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < NACC; ++j) mma_tf32(acc[j], a, b);
  }
  __syncthreads();
  const long long t1 = clock64();
  float s = 0.f;
  for (int j = 0; j < NACC; ++j)
    for (int i = 0; i < 4; ++i) s += acc[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NACC>
void run(int warps, int iters) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(float) * 148 * 1024);
  cudaMallocManaged(&cyc, sizeof(long long));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  probe<NACC><<<148, warps * 32>>>(out, 10, cyc);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  probe<NACC><<<148, warps * 32>>>(out, iters, cyc);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double mmas_per_smsp = (double)iters * NACC * warps / 4.0;
  const double flop = 148.0 * warps * (double)iters * NACC * 2.0 * 16 * 8 * 8;
  printf("warps/SM %2d  acc %d  cycles %lld  cycles per MMA per SMSP %.2f  %.1f TFLOP/s  (%s)\n", warps, NACC,
         *cyc, (double)*cyc / mmas_per_smsp, flop / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  const int iters = 20000;
  for (int w : {4, 8, 16, 32}) {
    run<1>(w, iters);
    run<2>(w, iters);
    run<4>(w, iters);
    run<8>(w, iters);
  }
  return 0;
}
