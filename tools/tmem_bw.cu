// Micro-benchmark: TMEM -> register read throughput (tcgen05.ld 32x32b) per SM on B200.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bw tools/tmem_bw.cu && ./tmem_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../graphembedding_b200/csrc/sgnn_tc.cuh"

template <int X>
__device__ __forceinline__ float ld_cols(uint32_t taddr);
template <>
__device__ __forceinline__ float ld_cols<32>(uint32_t taddr) {
  float v[32];
  tc::tmem_ld32(taddr, v);
  tc::tmem_ld_wait();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) s += v[i];
  return s;
}
template <>
__device__ __forceinline__ float ld_cols<128>(uint32_t taddr) {
  float v[128];
#pragma unroll
  for (int c = 0; c < 4; ++c) tc::tmem_ld32(taddr + c * 32, v + c * 32);
  tc::tmem_ld_wait();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 128; ++i) s += v[i];
  return s;
}

template <int X>
__global__ void __launch_bounds__(512, 1) bw_kernel(int reps, float* out, long long* cycles) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tc::tmem_alloc(&slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) * X) % 512;
  float s = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) s += ld_cols<X>(base);
  __syncthreads();
  const long long t1 = clock64();
  if (s == 123.456f) out[0] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(slot, 512);
  (void)lane;
}

template <int X>
void run(int warps) {
  float* out; long long* cyc;
  cudaMalloc(&out, 4); cudaMalloc(&cyc, 8 * 148);
  const int reps = 2000;
  bw_kernel<X><<<148, warps * 32, 0>>>(reps, out, cyc);
  bw_kernel<X><<<148, warps * 32, 0>>>(reps, out, cyc);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  const double bytes = (double)reps * warps * 32 * X * 4;
  printf("x%-3d warps=%2d  cycles=%lld  bytes/cycle/SM=%.1f  (%s)\n", X, warps, h[0], bytes / h[0], cudaGetErrorString(e));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int w : {4, 8, 16}) run<32>(w);
  for (int w : {4, 8}) run<128>(w);
  return 0;
}
