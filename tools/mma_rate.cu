// tcgen05.mma issue rate of ONE thread for the operand layouts and accumulator patterns of this repo's kernels:
// K-major NO-SWIZZLE core-matrix tiles (what sgnn_hist_tc.cuh / sgnn_allpairs_tc.cuh stage with one bulk copy per tile)
// against K-major SWIZZLE_128B rows, M = 128, N in {64, 128, 256}, K = 16 per instruction; chains of CHAIN
// accumulating MMAs per TMEM buffer, rotating over NBUF buffers.  Prints cycles per MMA (clock64 around issue + commit
// + wait) and checks the SWIZZLE_128B result against the host.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mma_rate tools/mma_rate.cu && tools/mma_rate
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_bf16.h>
#include "../graphembedding_b200/csrc/sgnn_tc.cuh"

__device__ __forceinline__ uint64_t desc_sw128(uint32_t smem_addr) {   // K-major, 128-byte rows, 8-row atoms of 1024 bytes
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;                  // LBO: unused for one swizzle atom along K
  d |= (uint64_t)(1024 >> 4) << 32;        // SBO: 8-row group stride
  d |= (uint64_t)1 << 46;                  // descriptor version
  d |= (uint64_t)2 << 61;                  // SWIZZLE_128B
  return d;
}

struct Args { int mode, N, chain, nbuf, tiles, acc_first; };   // acc_first: the first MMA of a chain accumulates too (timing only)

// out[0] = cycles, out[1..] = D of buffer 0 after ONE chain (check): [128 lanes][N] floats at out + 16
__global__ void __launch_bounds__(128) rate_kernel(Args a, const uint8_t* __restrict__ Ag, const uint8_t* __restrict__ Bg,
                                                   long long* __restrict__ cyc, float* __restrict__ D) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                 // 16 KB either way
  uint8_t* sB = smem + 16384;         // N rows x 128 B <= 32 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 16384 / 16; i += 128) reinterpret_cast<uint4*>(sA)[i] = reinterpret_cast<const uint4*>(Ag)[i];
  for (int i = threadIdx.x; i < 32768 / 16; i += 128) reinterpret_cast<uint4*>(sB)[i] = reinterpret_cast<const uint4*>(Bg)[i];
  if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::mbar_init(bar + 1, 1); tc::fence_barrier_init(); }
  tc::fence_proxy_async_smem();
  if (warp == 0) tc::tmem_alloc(slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t idesc = tc::umma_idesc_bf16(128, a.N);
  if (threadIdx.x == 0) {
    // K slices: no-swizzle -- slice s at + s * (rows * 32 B) ([2 chunks][rows][16 B]); SW128 -- slice s at + 32 s bytes in the row
    uint64_t ad[4], bd[4];
    for (int s = 0; s < 4; ++s) {
      if (a.mode == 0) {
        ad[s] = tc::umma_desc(tc::smem_u32(sA) + s * 4096, 128 * 16, 128);
        bd[s] = tc::umma_desc(tc::smem_u32(sB) + s * a.N * 32, a.N * 16, 128);
      } else {
        ad[s] = desc_sw128(tc::smem_u32(sA) + s * 32);
        bd[s] = desc_sw128(tc::smem_u32(sB) + s * 32);
      }
    }
    // check pass: one chain of 4 slices into buffer 0
    for (int s = 0; s < 4; ++s) tc::umma_bf16(tmem, ad[s], bd[s], idesc, s > 0);
    tc::umma_commit(bar);
    tc::mbar_wait(bar, 0);
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  for (int c0 = 0; c0 < a.N; c0 += 8) {
    float v[8];
    tc::tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tc::tmem_ld_wait();
    for (int j = 0; j < 8; ++j) D[(size_t)(warp * 32 + lane) * a.N + c0 + j] = v[j];
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  if (threadIdx.x == 0) {
    uint64_t ad[4], bd[4];
    for (int s = 0; s < 4; ++s) {
      if (a.mode == 0) {
        ad[s] = tc::umma_desc(tc::smem_u32(sA) + s * 4096, 128 * 16, 128);
        bd[s] = tc::umma_desc(tc::smem_u32(sB) + s * a.N * 32, a.N * 16, 128);
      } else {
        ad[s] = desc_sw128(tc::smem_u32(sA) + s * 32);
        bd[s] = desc_sw128(tc::smem_u32(sB) + s * 32);
      }
    }
    const long long t0 = clock64();
    for (int t = 0; t < a.tiles; ++t) {
      const uint32_t d = tmem + (uint32_t)(t & (a.nbuf - 1)) * (uint32_t)a.N;   // (nbuf is a power of two: no division on the lone issuing thread)
      for (int p = 0; p < a.chain; ++p) tc::umma_bf16(d, ad[p & 3], bd[(p + 1) & 3], idesc, (p > 0) | a.acc_first);
    }
    const long long t1 = clock64();
    tc::umma_commit(bar + 1);
    tc::mbar_wait(bar + 1, 0);
    const long long t2 = clock64();
    cyc[0] = t1 - t0; cyc[1] = t2 - t0;
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

static float aval(int r, int k) { return (float)(((r * 7 + k * 3) % 11) - 5); }
static float bval(int n, int k) { return (float)(((n * 5 + k) % 7) - 3); }

int main() {
  const int Ns[3] = {64, 128, 256};
  long long* cyc; float* D; uint8_t *Ag, *Bg;
  cudaMalloc(&cyc, 16); cudaMalloc(&D, 128 * 256 * 4); cudaMalloc(&Ag, 16384); cudaMalloc(&Bg, 32768);
  const size_t smem = 16384 + 32768 + 64;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int mode = 0; mode < 2; ++mode)
    for (int ni = 0; ni < 3; ++ni) {
      const int N = Ns[ni];
      std::vector<__nv_bfloat16> A(8192, __float2bfloat16(0.f)), B(16384, __float2bfloat16(0.f));
      for (int r = 0; r < 128; ++r)
        for (int k = 0; k < 64; ++k) {
          size_t o;
          if (mode == 0) o = (size_t)(k / 16) * 2048 + (size_t)((k % 16) / 8) * 1024 + (size_t)r * 8 + (k % 8);
          else o = (size_t)r * 64 + (size_t)((((k / 8) ^ (r & 7))) * 8) + (k % 8);
          A[o] = __float2bfloat16(aval(r, k));
        }
      for (int n = 0; n < N; ++n)
        for (int k = 0; k < 64; ++k) {
          size_t o;
          if (mode == 0) o = (size_t)(k / 16) * (N * 16) + (size_t)((k % 16) / 8) * (N * 8) + (size_t)n * 8 + (k % 8);
          else o = (size_t)n * 64 + (size_t)((((k / 8) ^ (n & 7))) * 8) + (k % 8);
          B[o] = __float2bfloat16(bval(n, k));
        }
      cudaMemcpy(Ag, A.data(), 16384, cudaMemcpyHostToDevice);
      cudaMemcpy(Bg, B.data(), 32768, cudaMemcpyHostToDevice);
      // (chain, buffers, first MMA accumulates): the histogram kernel's, independent MMAs, one long chain, then the two
      // that separate the cost of a NEW accumulator from the cost of NOT accumulating
      const int pats[5][3] = {{6, 4, 0}, {1, 4, 0}, {24, 1, 0}, {1, 4, 1}, {1, 1, 0}};
      for (int pi = 0; pi < 5; ++pi) {
        if (mode == 1 && pi > 0) continue;
        Args a{mode, N, pats[pi][0], std::min(pats[pi][1], 512 / N), 96 / pats[pi][0] * 4, pats[pi][2]};
        cudaMemset(cyc, 0, 16);
        rate_kernel<<<1, 128, smem>>>(a, Ag, Bg, cyc, D);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d N %d: %s\n", mode, N, cudaGetErrorString(e)); return 1; }
        long long h[2]; cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
        std::vector<float> d((size_t)128 * N);
        cudaMemcpy(d.data(), D, d.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r = 0; r < 128; ++r)
          for (int n = 0; n < N; ++n) {
            float want = 0.f;
            for (int k = 0; k < 64; ++k) want += aval(r, k) * bval(n, k);
            if (d[(size_t)r * N + n] != want) ++bad;
          }
        const int mmas = a.tiles * a.chain;
        printf("%-10s N %3d chain %2d bufs %d acc_first %d: %4d MMAs, issue %6.1f cyc/MMA, complete %6.1f cyc/MMA (floor %d), check: %d wrong of %d\n",
               mode ? "SW128" : "no-swizzle", N, a.chain, a.nbuf, a.acc_first, mmas, (double)h[0] / mmas, (double)h[1] / mmas, N / 2, bad, 128 * N);
      }
    }
  return 0;
}
