// Ranking kernels: descending top-k / full ranking per query row with the reference's tie
// order.  src/results.py:221-224 ranks by `argsort(sim, kind='mergesort')[:, ::-1]`: a
// stable ascending sort reversed, so equal scores come out HIGHER INDEX FIRST.  Packing
// (orderable(score) << 32 | index) into one 64-bit key and sorting keys descending
// reproduces exactly that order with integer compares only.
#pragma once
#include "sgnn_common.cuh"

__device__ __forceinline__ unsigned long long rank_key(float f, int idx) {
  f += 0.f;  // -0.0 -> +0.0 (numpy compares them equal)
  unsigned int u = __float_as_uint(f);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned int)idx;
}

__device__ __forceinline__ unsigned long long block_max_u64(unsigned long long v,
                                                            unsigned long long* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t > v ? t : v;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  v = lane < nw ? s_red[lane] : 0ull;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t > v ? t : v;
  }
  return v;
}

// One CTA per row.  k selection rounds; round r returns the largest key strictly below
// the previous winner (keys are unique because the index is part of the key).
__global__ void __launch_bounds__(256)
topk_kernel(const float* __restrict__ sim, const int32_t* __restrict__ col_ids, int n, int64_t ld, int k,
            int row_in_smem, int32_t* __restrict__ ids, float* __restrict__ vals) {
  extern __shared__ __align__(16) float s_row[];
  __shared__ unsigned long long s_red[32];
  const float* r = sim + (size_t)blockIdx.x * ld;
  // merge mode: candidate i carries an explicit (global) id; ties still break on that id
  const int32_t* cid = col_ids ? col_ids + (size_t)blockIdx.x * ld : nullptr;
  if (row_in_smem) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_row[i] = r[i];
    __syncthreads();
    r = s_row;
  }
  unsigned long long prev = ~0ull;
  for (int round = 0; round < k; ++round) {
    unsigned long long best = 0ull;
    int best_i = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const unsigned long long key = rank_key(r[i], cid ? cid[i] : i);
      if (key < prev && key > best) { best = key; best_i = i; }
    }
    const unsigned long long mine = best;
    best = block_max_u64(best, s_red);
    if (mine == best && best != 0ull) {  // keys are unique: exactly one owner
      ids[(size_t)blockIdx.x * k + round] = (int)(best & 0xffffffffull);
      if (vals) vals[(size_t)blockIdx.x * k + round] = r[best_i];
    }
    prev = best;
  }
}

// Full descending ranking of one row per CTA: bitonic sort of 64-bit keys in shared memory.
__global__ void __launch_bounds__(1024)
sort_rows_kernel(const float* __restrict__ sim, int n, int n_pad, int64_t ld,
                 int32_t* __restrict__ ids) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  const float* r = sim + (size_t)blockIdx.x * ld;
  for (int i = threadIdx.x; i < n_pad; i += blockDim.x) s_keys[i] = i < n ? rank_key(r[i], i) : 0ull;
  __syncthreads();
  for (int size = 2; size <= n_pad; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < (n_pad >> 1); t += blockDim.x) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool desc = (lo & size) == 0;  // descending blocks first -> overall descending
        const unsigned long long a = s_keys[lo], b = s_keys[hi];
        if ((a < b) == desc) { s_keys[lo] = b; s_keys[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    ids[(size_t)blockIdx.x * n + i] = (int)(s_keys[i] & 0xffffffffull);
}
