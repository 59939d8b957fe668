// Top-k candidate exchange between the GPUs of one node over NVLink / NVSwitch peer memory.
//
// All-pairs ranking shards the DATABASE over ranks; every rank ends a step with a local
// top-k [m][k] per query and needs the k best of the world * k candidates.  The exchange is a
// candidate all-gather of m * k * 8 bytes per rank -- 80 KB: pure latency.  Instead of a library
// all-gather plus glue kernels, every rank PUSHES its candidates straight into a slot of every
// peer's exchange buffer (remote stores over NVLink into memory opened through CUDA IPC), publishes
// a per-sender epoch flag with release semantics at system scope, and the merge kernel of every
// rank acquires the world flags and ranks the candidates by counting.  Two kernels, no host
// round trip, no staging copies.
//
// Exchange buffer of one rank (device memory of that rank, opened by every peer):
//   u64 flags[2][PEER_MAX_WORLD]            epoch of the newest complete push of sender r, per parity
//   u64 bflags[PEER_MAX_WORLD]              device-side barrier (sgnn_peer_barrier): arrival epoch of rank r
//   u64 cand[2][world][m_cap * k_cap]       (orderable score << 32 | global id) keys, per parity
// Parity = epoch & 1 double-buffers the slots: a sender can only be one epoch ahead of a receiver
// (it needs the receiver's flag of epoch e before it finishes epoch e), so the slot it overwrites at
// epoch e + 2 has been merged by then.
#pragma once
#include <cstdio>
#include "sgnn_rank.cuh"

constexpr int PEER_MAX_WORLD = 16;
constexpr size_t PEER_HEADER_BYTES = 3 * PEER_MAX_WORLD * sizeof(unsigned long long);

struct PeerPtrs {
  unsigned long long* base[PEER_MAX_WORLD];   // exchange buffer of every rank (own included)
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Wait for a peer's flag to reach `epoch`.  Bounded: a rank that died or diverged (unequal call counts -- every rank must
// make the same sequence of collective calls, include/sgnn.h) would otherwise hang its peers inside a kernel for ever.
// After 30 s of waiting the kernel traps: the stream fails with a launch error that the next
// synchronising call reports, instead of a silent hang.
__device__ unsigned long long g_peer_spin_limit_ns = 30ull * 1000000000ull;
__device__ __forceinline__ unsigned long long peer_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void peer_wait_flag(const unsigned long long* f, unsigned long long epoch, unsigned ns) {
  if (ld_acquire_sys(f) >= epoch) return;
  const unsigned long long t0 = peer_now_ns(), limit = g_peer_spin_limit_ns;
  while (ld_acquire_sys(f) < epoch) {
    __nanosleep(ns);
    if (peer_now_ns() - t0 > limit) {
      printf("sgnn peer exchange: waited %llu s for a peer's flag (epoch %llu) -- a rank died or the ranks made different "
             "sequences of collective calls\n", limit / 1000000000ull, epoch);
      __trap();
    }
  }
}

// Stream-ordered barrier between the ranks, no host involvement: arrive at every peer, wait for every peer.
// Epochs live in DEVICE memory (epoch_ctr[0] candidates, [1] gradients, [2] barrier) and are advanced by the
// kernels themselves, so the launches carry no per-step host argument and can sit inside a CUDA graph.
__global__ void peer_barrier_kernel(int rank, int world, PeerPtrs P, unsigned long long* epoch_ctr) {
  const unsigned long long epoch = *epoch_ctr + 1;
  __syncwarp();
  if (threadIdx.x == 0) *epoch_ctr = epoch;
  if ((int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(P.base[threadIdx.x] + 2 * PEER_MAX_WORLD + rank, epoch);
    const unsigned long long* f = P.base[rank] + 2 * PEER_MAX_WORLD + threadIdx.x;
    peer_wait_flag(f, epoch, 32);
  }
}

// Every rank: local candidates -> slot `rank` of every peer, then the epoch flag.  done_counter is a
// local device word (zero before and after each launch).
__global__ void __launch_bounds__(256)
peer_push_kernel(const float* __restrict__ vals, const int32_t* __restrict__ ids, int count, int id_offset,
                 int rank, int world, int64_t slot_elems, PeerPtrs P, unsigned long long* epoch_ctr,
                 unsigned int* __restrict__ done_counter) {
  const unsigned long long epoch = *epoch_ctr + 1;   // advanced by the last CTA, after every CTA has read it
  const int parity = (int)(epoch & 1ull);
  const size_t slot = PEER_HEADER_BYTES / 8 + ((size_t)parity * world + rank) * slot_elems;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const unsigned long long key = rank_key(vals[i], ids[i] + id_offset);
    for (int r = 0; r < world; ++r) P.base[r][slot + i] = key;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done_counter, 1u);
    if (prev == gridDim.x - 1) {   // last CTA: every candidate of this rank is visible system-wide
      *done_counter = 0u;
      __threadfence_system();
      for (int r = 0; r < world; ++r) st_release_sys(P.base[r] + parity * PEER_MAX_WORLD + rank, epoch);
      *epoch_ctr = epoch;
    }
  }
}

// Every rank: wait for the world flags of this epoch, then top-k of the world * k candidates of
// each row (keys are unique: the global id is part of the key).  One warp per row.
__global__ void __launch_bounds__(256)
peer_merge_kernel(const unsigned long long* __restrict__ mine, int m, int k, int world, int64_t slot_elems,
                  const unsigned long long* __restrict__ epoch_ctr, int32_t* __restrict__ ids,
                  float* __restrict__ vals) {
  extern __shared__ unsigned long long s_keys[];   // [8 warps][world * k]
  const unsigned long long epoch = *epoch_ctr;     // the push kernel ahead of us in the stream advanced it
  const int parity = (int)(epoch & 1ull);
  if (threadIdx.x < world) {
    const unsigned long long* f = mine + parity * PEER_MAX_WORLD + threadIdx.x;
    peer_wait_flag(f, epoch, 64);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = blockIdx.x * 8 + warp;
  if (row >= m) return;
  const int c = world * k;
  unsigned long long* sk = s_keys + (size_t)warp * c;
  const unsigned long long* cand = mine + PEER_HEADER_BYTES / 8 + (size_t)parity * world * slot_elems;
  for (int t = lane; t < c; t += 32) {
    const int r = t / k, j = t - r * k;
    sk[t] = __ldcg(cand + (size_t)r * slot_elems + (size_t)row * k + j);   // written by peers: bypass L1
  }
  __syncwarp();
  for (int t = lane; t < c; t += 32) {
    const unsigned long long kq = sk[t];
    int rk = 0;
    for (int u = 0; u < c; ++u) rk += sk[u] > kq ? 1 : 0;
    if (rk < k) {
      ids[(size_t)row * k + rk] = (int)(kq & 0xffffffffull);
      if (vals) vals[(size_t)row * k + rk] = rank_key_score(kq);
    }
  }
}

// ---- data-parallel training: gradient all-reduce fused into the optimiser ----------------------
// The flat gradient blob is 11-40 KB: again pure latency.  Every rank pushes its blob into slot
// `rank` of every peer's gradient region (second region of the same exchange buffer), and the Adam
// kernel of every rank sums the world slots IN RANK ORDER (so all ranks apply bit-identical updates)
// as it reads them -- there is no separate reduced-gradient pass and no library collective.
// Region (after the candidate region, `goff` u64 words from the base):
//   u64 gflags[2][PEER_MAX_WORLD]; float grad[2][world][n_pad]
__global__ void __launch_bounds__(256)
peer_grad_push_kernel(const float* __restrict__ grads, int n, int n_pad, int rank, int world, size_t goff,
                      PeerPtrs P, unsigned long long* epoch_ctr, unsigned int* __restrict__ done_counter) {
  const unsigned long long epoch = *epoch_ctr + 1;
  const int parity = (int)(epoch & 1ull);
  const size_t slot = (size_t)(parity * world + rank) * n_pad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float g = grads[i];
    for (int r = 0; r < world; ++r)
      reinterpret_cast<float*>(P.base[r] + goff + 2 * PEER_MAX_WORLD)[slot + i] = g;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done_counter, 1u);
    if (prev == gridDim.x - 1) {
      *done_counter = 0u;
      __threadfence_system();
      for (int r = 0; r < world; ++r) st_release_sys(P.base[r] + goff + parity * PEER_MAX_WORLD + rank, epoch);
      *epoch_ctr = epoch;
    }
  }
}

// weight decay gradient + TF-1 Adam (as adam_kernel) on the sum of the world gradient slots
__global__ void __launch_bounds__(256)
adam_peer_kernel(int n, int n_pad, int world, const unsigned long long* __restrict__ mine, size_t goff,
                 const unsigned long long* __restrict__ epoch_ctr, float* __restrict__ params,
                 float* __restrict__ grads,
                 float* __restrict__ m, float* __restrict__ v, float grad_scale, float wd, float lr_t, float b1,
                 float b2, float eps, float* __restrict__ loss) {
  const unsigned long long epoch = *epoch_ctr;
  const int parity = (int)(epoch & 1ull);
  if (threadIdx.x < world) {
    const unsigned long long* f = mine + goff + parity * PEER_MAX_WORLD + threadIdx.x;
    peer_wait_flag(f, epoch, 64);
  }
  __syncthreads();
  const float* slots = reinterpret_cast<const float*>(mine + goff + 2 * PEER_MAX_WORLD) + (size_t)parity * world * n_pad;
  float wdl = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float gs = 0.f;
    for (int r = 0; r < world; ++r) gs += __ldcg(slots + (size_t)r * n_pad + i);   // rank order: same sum everywhere
    const float p = params[i];
    const float g = gs * grad_scale + wd * p;
    wdl = fmaf(p, p, wdl);
    const float mi = b1 * m[i] + (1.f - b1) * g;
    const float vi = b2 * v[i] + (1.f - b2) * g * g;
    m[i] = mi; v[i] = vi;
    params[i] = p - lr_t * mi / (sqrtf(vi) + eps);
    grads[i] = 0.f;
  }
  if (loss) {
    wdl = warp_sum(wdl);
    if ((threadIdx.x & 31) == 0 && wdl != 0.f) atomicAdd(loss + 1, 0.5f * wd * wdl);
  }
}
