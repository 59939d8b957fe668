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
  if (f != f) u = 0xFFFFFFFFu;   // NaN of either sign sorts above every number (np.argsort puts NaN last, the ranking reverses it)
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

__device__ __forceinline__ float rank_key_score(unsigned long long key) {
  unsigned int u = (unsigned int)(key >> 32);
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}

// Top-k for k <= 256, two streaming passes over the row instead of k.  Pass 1: every thread keeps
// the largest (score, index) key of its slice.  Each warp sorts its 32 slice maxima (register bitonic
// sort over shuffles) and publishes its j-th largest, j = ceil(k / 8); the minimum L of the eight
// published keys has at least 8 j >= k elements >= it, so L is a lower bound of the row's k-th
// largest key.  Pass 2 gathers the (few) keys >= L into shared memory, where each candidate finds its
// exact rank by counting.  Keys carry the index, so ties cost nothing extra.  Rows with more than
// TOPK_CAND survivors fall back to k selection rounds.  Rows are read as float4 when aligned.
constexpr int TOPK_CAND = 1024;

__device__ __forceinline__ unsigned int orderable_u32(float f) {
  f += 0.f;   // -0.0 -> +0.0 (numpy compares them equal)
  const unsigned int u = __float_as_uint(f);
  if (f != f) return 0xFFFFFFFFu;   // NaN of either sign sorts above every number, as in numpy's order reversed
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void __launch_bounds__(256)
topk_select_kernel(const float* __restrict__ sim, const int32_t* __restrict__ col_ids, int n, int64_t ld,
                   int k, int32_t* __restrict__ ids, float* __restrict__ vals) {
  __shared__ unsigned long long s_cand[TOPK_CAND];
  __shared__ unsigned long long s_red[32];
  __shared__ unsigned long long s_warp[8];
  __shared__ int s_cnt;
  const float* r = sim + (size_t)blockIdx.x * ld;
  const int32_t* cid = col_ids ? col_ids + (size_t)blockIdx.x * ld : nullptr;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool vec = (ld & 3) == 0 && (reinterpret_cast<uintptr_t>(sim) & 15) == 0;
  const int n4 = vec ? n >> 2 : 0;
  const float4* r4 = reinterpret_cast<const float4*>(r);
  // slice maximum: element positions grow inside a thread, so ">=" keeps the higher position on ties
  unsigned int bu = 0u;
  int bi = -1;
  auto see = [&](float f, int i) {
    const unsigned int u = orderable_u32(f);
    if (u >= bu) { bu = u; bi = i; }
  };
#pragma unroll 4
  for (int i = tid; i < n4; i += 256) {
    const float4 v = r4[i];
    see(v.x, 4 * i); see(v.y, 4 * i + 1); see(v.z, 4 * i + 2); see(v.w, 4 * i + 3);
  }
  for (int i = 4 * n4 + tid; i < n; i += 256) see(r[i], i);
  unsigned long long key = 0ull;
  if (bi >= 0) {
    key = ((unsigned long long)bu << 32) | (unsigned int)(cid ? cid[bi] : bi);
    if (cid) {   // merge mode: ties break on the carried (global) id, not on the position
      // the slice maximum by (score, id) may be another position with the same score: rescan the ties
      for (int i = tid; i < n; i += 256)
        if (orderable_u32(r[i]) == bu) {
          const unsigned long long kk = ((unsigned long long)bu << 32) | (unsigned int)cid[i];
          key = kk > key ? kk : key;
        }
    }
  }
  if (tid == 0) s_cnt = 0;
  // descending bitonic sort of the warp's 32 keys
#pragma unroll
  for (int k2 = 2; k2 <= 32; k2 <<= 1) {
#pragma unroll
    for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, j2);
      const bool take_max = ((lane & j2) == 0) == ((lane & k2) == 0);
      key = take_max ? (key > other ? key : other) : (key < other ? key : other);
    }
  }
  const int j = (k + 7) >> 3;   // 1..32
  const unsigned long long tw = __shfl_sync(0xffffffffu, key, j - 1);
  if (lane == 0) s_warp[warp] = tw;
  __syncthreads();
  unsigned long long thr = s_warp[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) thr = s_warp[w] < thr ? s_warp[w] : thr;
  auto push = [&](float f, int i) {
    const unsigned int u = orderable_u32(f);
    if (u >= (unsigned int)(thr >> 32)) {   // cheap 32-bit pre-filter, then the exact 64-bit test
      const unsigned long long kk = ((unsigned long long)u << 32) | (unsigned int)(cid ? cid[i] : i);
      if (kk >= thr) {
        const int pos = atomicAdd(&s_cnt, 1);
        if (pos < TOPK_CAND) s_cand[pos] = kk;
      }
    }
  };
#pragma unroll 4
  for (int i = tid; i < n4; i += 256) {
    const float4 v = r4[i];
    push(v.x, 4 * i); push(v.y, 4 * i + 1); push(v.z, 4 * i + 2); push(v.w, 4 * i + 3);
  }
  for (int i = 4 * n4 + tid; i < n; i += 256) push(r[i], i);
  __syncthreads();
  const int c = s_cnt;
  int32_t* out_i = ids + (size_t)blockIdx.x * k;
  float* out_v = vals ? vals + (size_t)blockIdx.x * k : nullptr;
  if (c <= TOPK_CAND) {
    for (int q = tid; q < c; q += 256) {
      const unsigned long long kq = s_cand[q];
      int rank = 0;
      for (int t = 0; t < c; ++t) rank += s_cand[t] > kq ? 1 : 0;
      if (rank < k) {
        out_i[rank] = (int)(kq & 0xffffffffull);
        if (out_v) out_v[rank] = rank_key_score(kq);
      }
    }
    return;
  }
  unsigned long long prev = ~0ull;   // pathological rows: k selection rounds
  for (int round = 0; round < k; ++round) {
    unsigned long long bk = 0ull;
    for (int i = tid; i < n; i += 256) {
      const unsigned long long kk = rank_key(r[i], cid ? cid[i] : i);
      if (kk < prev && kk > bk) bk = kk;
    }
    bk = block_max_u64(bk, s_red);
    if (tid == 0) {
      out_i[round] = (int)(bk & 0xffffffffull);
      if (out_v) out_v[round] = rank_key_score(bk);
    }
    prev = bk;
  }
}

// Top-k with slice-maximum hints.  The tensor-core scorer also emits, for every query row, the
// largest rank key (score, then graph id) of every 32-column slice (smax [m][n_slices]).  The k-th
// largest slice key is a lower bound of the row's k-th largest key (k different slices each hold
// an element >= it), and a slice whose largest key is below the bound holds no top-k element: the
// kernel reads n / 32 hints plus a handful of slices instead of the row (twice).  Because the
// keys carry the graph id, rows full of equal scores cost nothing extra.  Bound, candidate ranking
// and tie order are those of topk_select_kernel.
#ifdef SGNN_TILE_TRACE
__device__ int g_topk_dbg[4096 * 4];   // per row: candidates, hot slices, clock cycles, fell back
#endif
__global__ void __launch_bounds__(256)
topk_hint_kernel(const float* __restrict__ sim, const unsigned long long* __restrict__ smax, int n_slices, int n,
                 int64_t ld, int k, int32_t* __restrict__ ids, float* __restrict__ vals) {
  __shared__ unsigned long long s_cand[TOPK_CAND];
  __shared__ unsigned long long s_red[32];
  __shared__ unsigned long long s_top[8][32];
  __shared__ unsigned long long s_thr;
  __shared__ int s_cnt;
#ifdef SGNN_TILE_TRACE
  const long long dbg_t0 = clock64();
  int dbg_hot = 0;
#endif
  const float* r = sim + (size_t)blockIdx.x * ld;
  const unsigned long long* h = smax + (size_t)blockIdx.x * n_slices;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // thread maximum over its hints (threads own disjoint slices); the EXACT k-th largest of the 256
  // thread maxima is the bound: each warp sorts its 32 keys, and a key of global rank < k only has
  // keys of local rank < k above it, so counting over every warp's first min(k, 32) keys is exact
  unsigned long long key = 0ull;
  for (int i = tid; i < n_slices; i += 256) {
    const unsigned long long u = h[i];
    key = u > key ? u : key;
  }
  if (tid == 0) { s_cnt = 0; s_thr = 0ull; }
#pragma unroll
  for (int k2 = 2; k2 <= 32; k2 <<= 1) {
#pragma unroll
    for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, j2);
      const bool take_max = ((lane & j2) == 0) == ((lane & k2) == 0);
      key = take_max ? (key > other ? key : other) : (key < other ? key : other);
    }
  }
  const int kk32 = k < 32 ? k : 32;
  if (lane < kk32) s_top[warp][lane] = key;
  __syncthreads();
  if (lane < kk32 && key != 0ull) {
    int above = 0;
    for (int w = 0; w < 8; ++w)
      for (int l = 0; l < kk32; ++l) above += s_top[w][l] > key ? 1 : 0;
    if (above == k - 1) s_thr = key;   // keys are unique: one writer
  }
  __syncthreads();
  const unsigned long long thr = s_thr;
  const unsigned int thr_u = (unsigned int)(thr >> 32);
  // slices that can hold a top-k element; a warp reads up to four of them per step, lane = element
  for (int s0 = warp * 32; s0 < n_slices; s0 += 256) {
    const int sl = s0 + lane;
    const bool hot = sl < n_slices && thr != 0ull && h[sl] >= thr;
    unsigned int todo = __ballot_sync(0xffffffffu, hot);
#ifdef SGNN_TILE_TRACE
    dbg_hot += __popc(todo);
#endif
    while (todo) {
      int idx[4];
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        idx[u] = -1;
        if (todo) {
          idx[u] = (s0 + __ffs(todo) - 1) * 32 + lane;
          todo &= todo - 1;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = idx[u] >= 0 && idx[u] < n ? r[idx[u]] : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (idx[u] >= 0 && idx[u] < n) {
          const unsigned int ou = orderable_u32(v[u]);
          if (ou >= thr_u) {
            const unsigned long long kk = ((unsigned long long)ou << 32) | (unsigned int)idx[u];
            if (kk >= thr) {
              const int pos = atomicAdd(&s_cnt, 1);
              if (pos < TOPK_CAND) s_cand[pos] = kk;
            }
          }
        }
      }
    }
  }
  __syncthreads();
  const int c = s_cnt;
#ifdef SGNN_TILE_TRACE
  if (blockIdx.x < 4096 && lane == 0) {
    if (tid == 0) { g_topk_dbg[blockIdx.x * 4] = c; g_topk_dbg[blockIdx.x * 4 + 2] = (int)(clock64() - dbg_t0);
                    g_topk_dbg[blockIdx.x * 4 + 3] = !(c >= k && c <= TOPK_CAND); g_topk_dbg[blockIdx.x * 4 + 1] = 0; }
    __threadfence_block();
  }
  __syncthreads();
  if (blockIdx.x < 4096 && lane == 0) atomicAdd(&g_topk_dbg[blockIdx.x * 4 + 1], dbg_hot);
#endif
  int32_t* out_i = ids + (size_t)blockIdx.x * k;
  float* out_v = vals ? vals + (size_t)blockIdx.x * k : nullptr;
  if (c >= k && c <= TOPK_CAND) {
    for (int q = tid; q < c; q += 256) {
      const unsigned long long kq = s_cand[q];
      int rank = 0;
      for (int t = 0; t < c; ++t) rank += s_cand[t] > kq ? 1 : 0;
      if (rank < k) {
        out_i[rank] = (int)(kq & 0xffffffffull);
        if (out_v) out_v[rank] = rank_key_score(kq);
      }
    }
    return;
  }
  unsigned long long prev = ~0ull;   // no usable bound: k selection rounds over the row
  for (int round = 0; round < k; ++round) {
    unsigned long long bk = 0ull;
    for (int i = tid; i < n; i += 256) {
      const unsigned long long kk = rank_key(r[i], i);
      if (kk < prev && kk > bk) bk = kk;
    }
    bk = block_max_u64(bk, s_red);
    if (tid == 0) {
      out_i[round] = (int)(bk & 0xffffffffull);
      if (out_v) out_v[round] = rank_key_score(bk);
    }
    prev = bk;
  }
}

// Warp-per-row version of the hinted top-k for k <= 32 and rows of <= 1024 * HMAX columns: the
// hints of a row sit in registers (HMAX per lane), the bound is the k-th largest LANE maximum
// (lanes own disjoint slices), the hot slices are read four at a time, and the handful of
// candidates is ranked by counting -- no CTA barrier anywhere, so a row costs about one chain of
// three dependent memory accesses.  Rows whose candidates overflow fall back to k selection rounds
// by the same warp.
constexpr int TOPK_WARP_CAND = 128;
constexpr int TOPK_WARP_HOT = 96;

template <int HMAX>
__global__ void __launch_bounds__(256)
topk_hint_warp_kernel(const float* __restrict__ sim, const unsigned long long* __restrict__ smax, int n_slices, int m,
                      int n, int64_t ld, int k, int32_t* __restrict__ ids, float* __restrict__ vals) {
  __shared__ unsigned long long s_cand[8][TOPK_WARP_CAND];
  __shared__ int s_hot[8][TOPK_WARP_HOT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = blockIdx.x * 8 + warp;
  if (row >= m) return;
  const float* r = sim + (size_t)row * ld;
  const unsigned long long* h = smax + (size_t)row * n_slices;
  unsigned long long hk[HMAX];
#pragma unroll
  for (int j = 0; j < HMAX; ++j) hk[j] = j * 32 + lane < n_slices ? h[j * 32 + lane] : 0ull;
  unsigned long long key = hk[0];
#pragma unroll
  for (int j = 1; j < HMAX; ++j) key = hk[j] > key ? hk[j] : key;
#pragma unroll
  for (int k2 = 2; k2 <= 32; k2 <<= 1) {
#pragma unroll
    for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, j2);
      const bool take_max = ((lane & j2) == 0) == ((lane & k2) == 0);
      key = take_max ? (key > other ? key : other) : (key < other ? key : other);
    }
  }
  const unsigned long long thr = __shfl_sync(0xffffffffu, key, k - 1);
  const unsigned int thr_u = (unsigned int)(thr >> 32);
  const unsigned int lt = (1u << lane) - 1u;
  int n_hot = 0;
  if (thr != 0ull) {
#pragma unroll
    for (int j = 0; j < HMAX; ++j) {
      const bool hot = hk[j] >= thr;
      const unsigned int mk = __ballot_sync(0xffffffffu, hot);
      const int pos = n_hot + __popc(mk & lt);
      if (hot && pos < TOPK_WARP_HOT) s_hot[warp][pos] = j * 32 + lane;
      n_hot += __popc(mk);
    }
  }
  __syncwarp();
  int c = 0;
  bool ok = thr != 0ull && n_hot <= TOPK_WARP_HOT;
  if (ok) {
    for (int t0 = 0; t0 < n_hot; t0 += 4) {
      int idx[4];
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        idx[u] = t0 + u < n_hot ? s_hot[warp][t0 + u] * 32 + lane : n;
        v[u] = idx[u] < n ? r[idx[u]] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        bool take = false;
        unsigned long long kk = 0ull;
        if (idx[u] < n) {
          const unsigned int ou = orderable_u32(v[u]);
          kk = ((unsigned long long)ou << 32) | (unsigned int)idx[u];
          take = ou >= thr_u && kk >= thr;
        }
        const unsigned int mk = __ballot_sync(0xffffffffu, take);
        const int pos = c + __popc(mk & lt);
        if (take && pos < TOPK_WARP_CAND) s_cand[warp][pos] = kk;
        c += __popc(mk);
      }
    }
    ok = c >= k && c <= TOPK_WARP_CAND;
  }
  __syncwarp();
  int32_t* out_i = ids + (size_t)row * k;
  float* out_v = vals ? vals + (size_t)row * k : nullptr;
  if (ok) {
    for (int q = lane; q < c; q += 32) {
      const unsigned long long kq = s_cand[warp][q];
      int rank = 0;
      for (int t = 0; t < c; ++t) rank += s_cand[warp][t] > kq ? 1 : 0;
      if (rank < k) {
        out_i[rank] = (int)(kq & 0xffffffffull);
        if (out_v) out_v[rank] = rank_key_score(kq);
      }
    }
    return;
  }
  unsigned long long prev = ~0ull;   // k selection rounds by this warp
  for (int round = 0; round < k; ++round) {
    unsigned long long bk = 0ull;
    for (int i = lane; i < n; i += 32) {
      const unsigned long long kk = rank_key(r[i], i);
      if (kk < prev && kk > bk) bk = kk;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long t = __shfl_xor_sync(0xffffffffu, bk, o);
      bk = t > bk ? t : bk;
    }
    if (lane == 0) {
      out_i[round] = (int)(bk & 0xffffffffull);
      if (out_v) out_v[round] = rank_key_score(bk);
    }
    prev = bk;
  }
}

// ---------------------------------------------------------------------------------------
// Bitonic sort of n_pad (power of two, 16 <= n_pad <= 16 * blockDim.x) 64-bit keys in shared memory; DESC selects the
// direction.  Register-blocked: a thread takes 2^NB keys whose indices differ in NB consecutive index bits, runs the NB
// compare-exchange steps of those strides in registers and writes them back, so a merge phase of log2(size) steps costs
// ceil(log2(size) / 4) trips through shared memory (32 for 16 384 keys instead of 105), and the phases of size <= 16
// are one trip.  While sorting, logical key i lives in slot bitonic_slot(i): a rotation inside its 16-key block by the
// block number, so both a thread's 16 CONTIGUOUS keys (stride-1 trip) and 16 lanes' consecutive keys (all other trips)
// fall into 16 different 8-byte banks.  The caller FILLS s_keys[bitonic_slot(i)]; the result is in linear order.
__device__ __forceinline__ int bitonic_slot(int i) { return (i & ~15) | ((i + (i >> 4)) & 15); }

template <bool DESC>
__device__ __forceinline__ void bitonic_cex(unsigned long long& a, unsigned long long& b, bool first) {
  if (((a < b) == first) == DESC) { const unsigned long long t = a; a = b; b = t; }
}

// the NB steps of strides 2^(ls+NB-1) ... 2^ls of the merge phase `size`
template <bool DESC, int NB>
__device__ __forceinline__ void bitonic_trip(unsigned long long* s_keys, int n_pad, int size, int ls) {
  constexpr int E = 1 << NB;
  for (int g = threadIdx.x; g < (n_pad >> NB); g += blockDim.x) {
    const int base = ((g >> ls) << (ls + NB)) | (g & ((1 << ls) - 1));
    const bool first = (base & size) == 0;   // the bit `size` lies above every stride bit: one direction per thread
    unsigned long long v[E];
#pragma unroll
    for (int j = 0; j < E; ++j) v[j] = s_keys[bitonic_slot(base + (j << ls))];
#pragma unroll
    for (int k = NB - 1; k >= 0; --k)
#pragma unroll
      for (int j = 0; j < E; ++j)
        if (!(j & (1 << k))) bitonic_cex<DESC>(v[j], v[j | (1 << k)], first);
#pragma unroll
    for (int j = 0; j < E; ++j) s_keys[bitonic_slot(base + (j << ls))] = v[j];
  }
  __syncthreads();
}

template <bool DESC>
__device__ __forceinline__ void block_bitonic_u64(unsigned long long* s_keys, int n_pad) {
  // merge phases of size 2, 4, 8, 16: a thread sorts 16 contiguous keys in registers
  for (int g = threadIdx.x; g < (n_pad >> 4); g += blockDim.x) {
    const int base = g << 4;
    unsigned long long v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = s_keys[bitonic_slot(base + j)];
#pragma unroll
    for (int size = 2; size <= 16; size <<= 1)
#pragma unroll
      for (int stride = size >> 1; stride > 0; stride >>= 1)
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (!(j & stride)) bitonic_cex<DESC>(v[j], v[j | stride], ((base + j) & size) == 0);
#pragma unroll
    for (int j = 0; j < 16; ++j) s_keys[bitonic_slot(base + j)] = v[j];
  }
  __syncthreads();
  for (int L = 5; (1 << L) <= n_pad; ++L) {
    const int size = 1 << L, r = L & 3;
    if (r == 1) bitonic_trip<DESC, 1>(s_keys, n_pad, size, L - 1);
    else if (r == 2) bitonic_trip<DESC, 2>(s_keys, n_pad, size, L - 2);
    else if (r == 3) bitonic_trip<DESC, 3>(s_keys, n_pad, size, L - 3);
    for (int ls = L - r - 4; ls >= 0; ls -= 4) bitonic_trip<DESC, 4>(s_keys, n_pad, size, ls);
  }
  // back to linear order (a permutation inside every 16-key block)
  unsigned long long v[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    const int i = threadIdx.x + c * blockDim.x;
    if (i < n_pad) v[c] = s_keys[bitonic_slot(i)];
  }
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    const int i = threadIdx.x + c * blockDim.x;
    if (i < n_pad) s_keys[i] = v[c];
  }
  __syncthreads();
}

// Full descending ranking of one row per CTA: bitonic sort of 64-bit keys in shared memory (block_bitonic_u64).
__global__ void __launch_bounds__(1024)
sort_rows_kernel(const float* __restrict__ sim, int n, int n_pad, int64_t ld,
                 int32_t* __restrict__ ids) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  const float* r = sim + (size_t)blockIdx.x * ld;
  for (int i = threadIdx.x; i < n_pad; i += blockDim.x) s_keys[bitonic_slot(i)] = i < n ? rank_key(r[i], i) : 0ull;
  __syncthreads();
  block_bitonic_u64<true>(s_keys, n_pad);
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    ids[(size_t)blockIdx.x * n + i] = (int)(s_keys[i] & 0xffffffffull);
}

// ---------------------------------------------------------------------------------------
// Ranking metrics, one CTA (1024 threads) per query row, everything in shared memory.
// precision@k (src/metrics.py:14-35 with results.py:51-82 top-k semantics): the true top-k is
// tie-inclusive, the predicted top-k is the first k of the descending ranking.  Item j is in
// both sets iff k >= need_j = max(tie-inclusive true position, predicted position), so ONE pass
// gives |true_topk(k) & pred_topk(k)| for every k.  hit_sum[i] (u64) += hits of this row at ks[i].
// smem: keys[n_pad] u64 | pos[n] i32 | hist[n + 2] i32
__global__ void __launch_bounds__(1024)
prec_at_k_kernel(const float* __restrict__ true_sim, const float* __restrict__ pred_sim, int n, int n_pad,
                 const int32_t* __restrict__ ks, int n_ks, unsigned long long* __restrict__ hit_sum) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  int* s_pos = reinterpret_cast<int*>(s_keys + n_pad);
  int* s_hist = s_pos + n;
  __shared__ int s_warp[32];
  const float* t = true_sim + (size_t)blockIdx.x * n;
  const float* p = pred_sim + (size_t)blockIdx.x * n;
  const int tid = threadIdx.x;
  for (int i = tid; i < n_pad; i += blockDim.x) s_keys[bitonic_slot(i)] = i < n ? rank_key(t[i], i) : 0ull;
  for (int i = tid; i < n + 2; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  block_bitonic_u64<true>(s_keys, n_pad);
  // tie-inclusive position = 1 + start of the run of equal scores: lower bound by binary search
  for (int i = tid; i < n; i += blockDim.x) {
    const unsigned int sc = (unsigned int)(s_keys[i] >> 32);
    int lo = 0, hi = i;   // first index whose score == sc (scores descending)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if ((unsigned int)(s_keys[mid] >> 32) > sc) lo = mid + 1; else hi = mid;
    }
    s_pos[(int)(s_keys[i] & 0xffffffffull)] = lo + 1;
  }
  __syncthreads();
  for (int i = tid; i < n_pad; i += blockDim.x) s_keys[bitonic_slot(i)] = i < n ? rank_key(p[i], i) : 0ull;
  __syncthreads();
  block_bitonic_u64<true>(s_keys, n_pad);
  for (int i = tid; i < n; i += blockDim.x) {
    const int j = (int)(s_keys[i] & 0xffffffffull);
    atomicAdd(&s_hist[max(s_pos[j], i + 1)], 1);
  }
  __syncthreads();
  // inclusive prefix sum of hist[0..n] in place (chunk per thread + block scan of chunk sums)
  const int per = (n + 1 + blockDim.x - 1) / blockDim.x;
  const int b0 = tid * per, b1 = min(n + 1, b0 + per);
  int sum = 0;
  for (int i = b0; i < b1; ++i) sum += s_hist[i];
  int incl = sum;
  const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  int run = incl - sum + (warp > 0 ? s_warp[warp - 1] : 0);
  for (int i = b0; i < b1; ++i) { run += s_hist[i]; s_hist[i] = run; }
  __syncthreads();
  for (int i = tid; i < n_ks; i += blockDim.x) atomicAdd(&hit_sum[i], (unsigned long long)s_hist[ks[i]]);
}

// prec[i] = hit_sum[i] / (ks[i] * m), in place over the u64 counters
__global__ void prec_finish_kernel(const int32_t* __restrict__ ks, int n_ks, int m, double* prec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_ks) return;
  const unsigned long long h = reinterpret_cast<unsigned long long*>(prec)[i];
  prec[i] = (double)h / ((double)ks[i] * (double)m);
}

__device__ __forceinline__ long long block_sum_i64(long long v, long long* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  v = lane < nw ? s_red[lane] : 0ll;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// tie pairs contributed by the elements of mask word t (kendall_tau_kernel): every set bit i adds the number of
// consecutive set bits ending at i, counted across word boundaries
__device__ __forceinline__ long long tie_pairs_of_word(const unsigned int* mask, int t) {
  unsigned int w = mask[t];
  if (w == 0u) return 0ll;
  long long carry = 0;
  if (w & 1u)   // the run at bit 0 continues the set bits at the top of the words before
    for (int j = t - 1; j >= 0; --j) {
      const int lead = __clz(~mask[j]);   // set bits from bit 31 downward (32 for an all-ones word)
      carry += lead;
      if (lead < 32) break;
    }
  long long sum = 0;
  int consumed = 0;
  while (w != 0u) {
    const int tz = __ffs(w) - 1;
    w >>= tz; consumed += tz;
    const int r = w == 0xffffffffu ? 32 : __ffs(~w) - 1;   // run length (the shifted-in top bits are zero)
    sum += (long long)r * (consumed == 0 ? carry : 0ll) + (long long)r * (r + 1) / 2;
    w = r == 32 ? 0u : w >> r;
    consumed += r;
  }
  return sum;
}

// Kendall tau-b of (true row, predicted row), exact integer pair counts (Knight's algorithm):
// sort by (x, y); tie pairs from run lengths; discordant pairs = strict inversions of the y
// sequence, counted by a bottom-up merge sort whose merges are rank computations by binary
// search.  counts = {concordant, discordant, ties only in x, ties only in y, joint ties}.
// smem: keys[n_pad] u64, later re-used as two u32 ping-pong arrays of n_pad entries each.
__global__ void __launch_bounds__(1024)
kendall_tau_kernel(const float* __restrict__ true_sim, const float* __restrict__ pred_sim, int n, int n_pad,
                   long long* __restrict__ counts, double* __restrict__ tau) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ long long s_red[32];
  const float* x = true_sim + (size_t)blockIdx.x * n;
  const float* y = pred_sim + (size_t)blockIdx.x * n;
  const int tid = threadIdx.x;
  for (int i = tid; i < n_pad; i += blockDim.x)
    s_keys[bitonic_slot(i)] = i < n ? (((unsigned long long)orderable_u32(x[i]) << 32) | orderable_u32(y[i])) : ~0ull;
  __syncthreads();
  block_bitonic_u64<false>(s_keys, n_pad);   // ascending; padding (all ones) goes last
  // tie pairs of a sorted sequence = sum_i (#elements before i inside i's run of equal keys): bit i of a mask says
  // "element i equals element i - 1" (one ballot per 32 consecutive elements), and the sum over a run of r set bits
  // that continues c set bits of the words before it is r c + r (r + 1) / 2 -- no binary searches
  __shared__ unsigned int s_mx[512], s_mxy[512];   // n_pad <= 16384
  const int n_words = (n + 31) >> 5;
  for (int i = tid; i < n_words * 32; i += blockDim.x) {   // (whole warps: i / 32 is warp-uniform)
    bool ex = false, exy = false;
    if (i >= 1 && i < n) {
      const unsigned long long k = s_keys[i], kp = s_keys[i - 1];
      ex = (unsigned int)(k >> 32) == (unsigned int)(kp >> 32);
      exy = k == kp;
    }
    const unsigned int bx = __ballot_sync(0xffffffffu, ex), bxy = __ballot_sync(0xffffffffu, exy);
    if ((tid & 31) == 0) { s_mx[i >> 5] = bx; s_mxy[i >> 5] = bxy; }
  }
  __syncthreads();
  long long nx = tid < n_words ? tie_pairs_of_word(s_mx, tid) : 0ll;
  long long nxy = tid < n_words ? tie_pairs_of_word(s_mxy, tid) : 0ll;
  nx = block_sum_i64(nx, s_red);
  nxy = block_sum_i64(nxy, s_red);
  // compact the y keys (low words) of the sorted sequence into a u32 array over the same memory
  unsigned int* bufA = reinterpret_cast<unsigned int*>(s_keys);
  unsigned int* bufB = bufA + n_pad;
  {
    unsigned int tmp[16];   // n_pad <= 16384 = 16 * 1024
    int c = 0;
    for (int i = tid; i < n_pad; i += blockDim.x) tmp[c++] = (unsigned int)(s_keys[i] & 0xffffffffull);
    __syncthreads();
    c = 0;
    for (int i = tid; i < n_pad; i += blockDim.x) bufA[i] = tmp[c++];
    __syncthreads();
  }
  long long dis = 0;
  unsigned int* src = bufA;
  unsigned int* dst = bufB;
  for (int w = 1; w < n; w <<= 1) {
    for (int i = tid; i < n; i += blockDim.x) {
      const int blk = i / (2 * w) * (2 * w);
      const int mid = min(blk + w, n), end = min(blk + 2 * w, n);
      const unsigned int v = src[i];
      if (i < mid) {              // left element: goes after the right elements strictly below it
        int lo = mid, hi = end;
        while (lo < hi) { const int md = (lo + hi) >> 1; if (src[md] < v) lo = md + 1; else hi = md; }
        dst[i + (lo - mid)] = v;
      } else {                    // right element: goes after the left elements <= it (stable)
        int lo = blk, hi = mid;
        while (lo < hi) { const int md = (lo + hi) >> 1; if (src[md] <= v) lo = md + 1; else hi = md; }
        dst[(i - mid) + lo] = v;
        dis += mid - lo;          // left elements strictly greater: inversions
      }
    }
    __syncthreads();
    unsigned int* sw = src; src = dst; dst = sw;
  }
  dis = block_sum_i64(dis, s_red);
  for (int i = tid; i < n_words * 32; i += blockDim.x) {
    const bool ey = i >= 1 && i < n && src[i] == src[i - 1];
    const unsigned int by = __ballot_sync(0xffffffffu, ey);
    if ((tid & 31) == 0) s_mx[i >> 5] = by;
  }
  __syncthreads();
  long long ny = tid < n_words ? tie_pairs_of_word(s_mx, tid) : 0ll;
  ny = block_sum_i64(ny, s_red);
  if (tid == 0) {
    const long long tot = (long long)n * (n - 1) / 2;
    const long long con = tot - nx - ny + nxy - dis;
    long long* c = counts + (size_t)blockIdx.x * 5;
    c[0] = con; c[1] = dis; c[2] = nx - nxy; c[3] = ny - nxy; c[4] = nxy;
    if (tau) {
      const double den = sqrt((double)(con + dis + (nx - nxy))) * sqrt((double)(con + dis + (ny - nxy)));
      tau[blockIdx.x] = den > 0.0 ? (double)(con - dis) / den : nan("");
    }
  }
}


// precision@k for SMALL k from two top-k lists (src/metrics.py:14-35): the tie-inclusive true top-k set of a row is
// { j : true[j] >= k-th largest true value }, the predicted top-k the first k ids of the predicted order, so
//   prec@k = mean_rows |{ i < k : true[row][pred_ids[row][i]] >= true_vals[row][k - 1] }| / k.
// One CTA per k of the list; the hit counts are integers, so the result is exact and deterministic (no sort of the
// 10 000-wide rows: 2 top-k launches + this kernel instead of two full row sorts).
__global__ void __launch_bounds__(256)
prec_from_topk_kernel(const float* __restrict__ true_sim, int64_t ld, const float* __restrict__ true_vals,
                      const int32_t* __restrict__ pred_ids, int m, int kmax, const int32_t* __restrict__ ks, int n_ks,
                      double* __restrict__ prec) {
  __shared__ long long s_sum[8];
  const int ki = blockIdx.x;
  if (ki >= n_ks) return;
  const int k = ks[ki];
  long long hits = 0;
  if (k >= 1 && k <= kmax)
    for (int r = threadIdx.x; r < m; r += blockDim.x) {
      const float thr = true_vals[(size_t)r * kmax + (k - 1)];
      const int32_t* ids = pred_ids + (size_t)r * kmax;
      const float* row = true_sim + (size_t)r * ld;
      int c = 0;
      for (int i = 0; i < k; ++i) c += row[ids[i]] >= thr ? 1 : 0;
      hits += c;
    }
  for (int o = 16; o > 0; o >>= 1) hits += __shfl_down_sync(0xffffffffu, hits, o);
  if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = hits;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < 8; ++w) t += s_sum[w];
    prec[ki] = (k >= 1 && k <= kmax) ? (double)t / ((double)k * (double)m) : 0.0 / 0.0;
  }
}
