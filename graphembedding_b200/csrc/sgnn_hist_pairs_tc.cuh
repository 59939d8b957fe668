// Node-node similarity histogram of LISTED pairs of LARGE graphs on the tensor cores (Web-shaped workloads, BASELINE
// configs[4]: 100-500 nodes per graph).  Same layer as sgnn_hist_tc.cuh (SURVEY a24 -- BUILD-DEFINED; pin =
// oracle/siamese_oracle.py: pair_histogram):  s_ij = sigmoid(H1_i . H2_j),  counts[p][min(B - 1, floor(s_ij B))] += 1.
//
// Where the all-pairs kernel packs many small graphs into one 128 x 128 accumulator and has to sort its cells by
// (graph, query) afterwards, a large pair fills whole accumulators by itself: every graph is cut into tiles of 128
// node rows (bf16 three-way split, K-major, both the plain form and the form scaled by -log2 e, written once per
// call), a work unit is (pair p, row tile ti of g1, column tile tj of g2), and ALL cells of a unit belong to the
// same pair.  So the epilogue is just the cell loop of sgnn_hist_tc.cuh (ht_segment: ex2, one reciprocal per two
// cells, round-down fma = bin, one-hot nibble word from the shared-memory table) over 14 columns at a time, the
// nibble words folded into 16 byte counters in two registers, and per unit one warp reduction (REDUX on widened
// 16-bit fields) + at most 16 integer atomics per warp onto counts[p][:].  A last pass puts the sigma == 1 cells
// (which the nibble table drops) into the last bin from the cell count and writes the normalised histogram.
//
// Roles (576 threads, 1 CTA per SM, persistent over a contiguous range of units): warps 0-15 = four epilogue groups
// of four warps (one TMEM lane quadrant each; group e takes the units it % 4 == e, TMEM buffer e), warp 16 = bulk
// copy producer (A tile once per run of column tiles, B tile per unit), warp 17 = MMA issuer (six bf16 products).
#pragma once
#include "sgnn_hist_tc.cuh"

constexpr int HP_STAGES = 4;            // B-tile ring
constexpr int HP_TILE_BYTES = 3 * 4096; // three parts, one K step (node rows of <= 16 values)
constexpr int HP_THREADS = HT_THREADS;
constexpr int HP_SEG = 14;              // cells per nibble word (<= 15; even: the paired form has no tail cell)
constexpr int HP_SMEM_BYTES = (2 + HP_STAGES) * HP_TILE_BYTES + (4 + 2 * HP_STAGES + 2 * HT_BUFS) * 8 + 16 + 32 * 8;

struct HpGeom {
  int n_graphs, n_pairs, d, bins;
  uint32_t lut_bias;   // see ht_count
};

// exclusive prefix sums by one CTA.  MODE 0: tiles per graph, ceil(nodes / 128) (a = node_off);
// MODE 1: units per pair, tiles(a[i]) * tiles(b[i]) (a = g1, b = g2).  out[n] = total.
template <int MODE>
__global__ void __launch_bounds__(1024)
hp_scan_kernel(const int32_t* __restrict__ a, const int32_t* __restrict__ b, const int32_t* __restrict__ tile_off, int n,
               int32_t* __restrict__ out) {
  __shared__ int wsum[32];
  __shared__ int base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int i = i0 + tid;
    int v = 0;
    if (i < n) {
      if (MODE == 0) v = (a[i + 1] - a[i] + 127) >> 7;
      else v = (tile_off[a[i] + 1] - tile_off[a[i]]) * (tile_off[b[i] + 1] - tile_off[b[i]]);
    }
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      wsum[lane] = w;
    }
    __syncthreads();
    const int before = base + (warp > 0 ? wsum[warp - 1] : 0) + x - v;
    if (i < n) out[i] = before;
    __syncthreads();
    if (tid == 1023) base = before + v;
    __syncthreads();
  }
  if (tid == 0) out[n] = base;
}

// operand tiles of every graph: tile t = rows [128 k, 128 k + 128) of graph g; At plain, Bt scaled by -log2(e)
__global__ void __launch_bounds__(128)
hp_prep_kernel(const HpGeom G, const float* __restrict__ nodes, const int32_t* __restrict__ node_off,
               const int32_t* __restrict__ tile_off, __nv_bfloat16* __restrict__ At, __nv_bfloat16* __restrict__ Bt) {
  __shared__ int s_g;
  const int n_tiles = tile_off[G.n_graphs];
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    if (threadIdx.x == 0) {   // last graph whose first tile is <= t (graphs without nodes own no tile)
      int lo = 0, hi = G.n_graphs;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (tile_off[mid + 1] <= t) lo = mid + 1; else hi = mid;
      }
      s_g = lo;
    }
    __syncthreads();
    const int g = s_g;
    const int r = (t - tile_off[g]) * 128 + (int)threadIdx.x;
    const int n = node_off[g + 1] - node_off[g];
    const float* src = r < n ? nodes + (size_t)(node_off[g] + r) * G.d : nullptr;
    const size_t o = (size_t)t * (HP_TILE_BYTES / 2) + (size_t)threadIdx.x * 8;
    ht_write_row(src, G.d, 1.f, 3, 1, At + o);
    ht_write_row(src, G.d, -1.4426950408889634f, 3, 1, Bt + o);
    __syncthreads();
  }
}

// position of a role in the unit list: pair p, row tile ti of g1, column tile tj of g2
struct HpWalk {
  int p, ti, tj, T1, T2, ta, tb, n1, n2;
  const int32_t* tile_off; const int32_t* node_off; const int32_t* unit_off; const int32_t* g1; const int32_t* g2;
  int n_pairs;
  __device__ __forceinline__ void load_pair() {
    const int a = g1[p], b = g2[p];
    ta = tile_off[a]; T1 = tile_off[a + 1] - ta;
    tb = tile_off[b]; T2 = tile_off[b + 1] - tb;
    n1 = node_off[a + 1] - node_off[a];
    n2 = node_off[b + 1] - node_off[b];
  }
  __device__ __forceinline__ void seek(int u) {   // u < unit_off[n_pairs]
    int lo = 0, hi = n_pairs - 1;
    while (lo < hi) {   // last pair with unit_off[p] <= u
      const int mid = (lo + hi + 1) >> 1;
      if (unit_off[mid] <= u) lo = mid; else hi = mid - 1;
    }
    p = lo;
    while (unit_off[p + 1] == unit_off[p]) ++p;   // (pairs with an empty graph own no unit; u < total keeps p in range)
    load_pair();
    const int local = u - unit_off[p];
    ti = local / T2; tj = local - ti * T2;
  }
  __device__ __forceinline__ void next() {   // only called when another unit follows
    if (++tj < T2) return;
    tj = 0;
    if (++ti < T1) return;
    ti = 0;
    do { ++p; } while (unit_off[p + 1] == unit_off[p]);
    load_pair();
  }
};

__global__ void __launch_bounds__(HP_THREADS, 1)
hist_pairs_tc_kernel(const HpGeom G, const __nv_bfloat16* __restrict__ At, const __nv_bfloat16* __restrict__ Bt,
                     const int32_t* __restrict__ tile_off, const int32_t* __restrict__ unit_off,
                     const int32_t* __restrict__ node_off, const int32_t* __restrict__ g1,
                     const int32_t* __restrict__ g2, int32_t* __restrict__ counts) {
  extern __shared__ __align__(1024) uint8_t hp_smem[];
  uint8_t* sA = hp_smem;                              // [2] tiles
  uint8_t* sB = sA + 2 * HP_TILE_BYTES;               // [HP_STAGES] tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + HP_STAGES * HP_TILE_BYTES);
  uint64_t* a_full = bars;                            // [2]
  uint64_t* a_empty = bars + 2;                       // [2]
  uint64_t* b_full = bars + 4;                        // [HP_STAGES]
  uint64_t* b_empty = b_full + HP_STAGES;             // [HP_STAGES]
  uint64_t* t_full = b_empty + HP_STAGES;             // [HT_BUFS]
  uint64_t* t_empty = t_full + HT_BUFS;               // [HT_BUFS]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + HT_BUFS);
  unsigned long long* s_lut = reinterpret_cast<unsigned long long*>(tmem_slot + 2);   // [32] one-hot nibble words

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 32) s_lut[threadIdx.x] = threadIdx.x < 16 ? 1ull << (4 * threadIdx.x) : 0ull;
  const long long total = unit_off[G.n_pairs];
  const int u0 = (int)(total * blockIdx.x / gridDim.x), u1 = (int)(total * (blockIdx.x + 1) / gridDim.x);

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { tc::mbar_init(a_full + s, 1); tc::mbar_init(a_empty + s, 1); }
    for (int s = 0; s < HP_STAGES; ++s) { tc::mbar_init(b_full + s, 1); tc::mbar_init(b_empty + s, 1); }
    for (int b = 0; b < HT_BUFS; ++b) { tc::mbar_init(t_full + b, 1); tc::mbar_init(t_empty + b, 4); }
    tc::fence_barrier_init();
  }
  if (warp == HT_PROD_WARP) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (u0 < u1) {
    HpWalk W;
    W.tile_off = tile_off; W.node_off = node_off; W.unit_off = unit_off; W.g1 = g1; W.g2 = g2; W.n_pairs = G.n_pairs;
    W.seek(u0);
    if (warp == HT_PROD_WARP) {
      // ===== producer: the A tile when a run of column tiles begins, one B tile per unit =====
      if (lane == 0) {
        int run = 0, useq = 0;
        for (int u = u0; u < u1; ++u, ++useq) {
          if (u == u0 || W.tj == 0) {
            const int sl = run & 1;
            if (run >= 2) tc::mbar_wait(a_empty + sl, ((run >> 1) - 1) & 1);
            tc::mbar_arrive_expect_tx(a_full + sl, HP_TILE_BYTES);
            tc::bulk_g2s(sA + sl * HP_TILE_BYTES, reinterpret_cast<const uint8_t*>(At) + (size_t)(W.ta + W.ti) * HP_TILE_BYTES,
                         HP_TILE_BYTES, a_full + sl);
            ++run;
          }
          const int s = useq % HP_STAGES;
          if (useq >= HP_STAGES) tc::mbar_wait(b_empty + s, ((useq / HP_STAGES) - 1) & 1);
          tc::mbar_arrive_expect_tx(b_full + s, HP_TILE_BYTES);
          tc::bulk_g2s(sB + s * HP_TILE_BYTES, reinterpret_cast<const uint8_t*>(Bt) + (size_t)(W.tb + W.tj) * HP_TILE_BYTES,
                       HP_TILE_BYTES, b_full + s);
          if (u + 1 < u1) W.next();
        }
      }
    } else if (warp == HT_MMA_WARP) {
      // ===== MMA issuer (one thread) =====
      if (lane == 0) {
        const uint32_t idesc = tc::umma_idesc_bf16(128, 128);
        const uint64_t a_desc0 = tc::umma_desc(tc::smem_u32(sA), 128 * 16, 128);
        const uint64_t b_desc0 = tc::umma_desc(tc::smem_u32(sB), 128 * 16, 128);
        constexpr uint32_t tile16 = HP_TILE_BYTES >> 4, part16 = 256u;
        const uint32_t pa = 0x201100u, pb = 0x021010u;   // (a part, b part) of the six products, smallest terms last
        int run = 0, useq = 0;
        for (int u = u0; u < u1; ++u, ++useq) {
          if (u == u0 || W.tj == 0) {
            tc::mbar_wait(a_full + (run & 1), (run >> 1) & 1);
            ++run;
          }
          const int sl = (run - 1) & 1;
          const int s = useq % HP_STAGES;
          tc::mbar_wait(b_full + s, (useq / HP_STAGES) & 1);
          const int buf = useq & (HT_BUFS - 1);
          if (useq >= HT_BUFS) tc::mbar_wait(t_empty + buf, ((useq / HT_BUFS) - 1) & 1);
          tc::tc_fence_after();
          const uint64_t ad = a_desc0 + (uint64_t)((uint32_t)sl * tile16), bd = b_desc0 + (uint64_t)((uint32_t)s * tile16);
          const uint32_t d_addr = tmem_base + (uint32_t)buf * 128u;
#pragma unroll
          for (int p = 0; p < 6; ++p)
            tc::umma_bf16(d_addr, ad + ((pa >> (4 * p)) & 15u) * part16, bd + ((pb >> (4 * p)) & 15u) * part16, idesc,
                          p > 0 ? 1u : 0u);
          tc::umma_commit(t_full + buf);
          tc::umma_commit(b_empty + s);
          if (W.tj == W.T2 - 1 || u + 1 == u1) tc::umma_commit(a_empty + sl);   // the run's last unit: the A slot is free
          if (u + 1 < u1) W.next();
        }
      }
    } else {
      // ===== epilogue =====
      const int grp = warp >> 2, quad = warp & 3;
      const int row = quad * 32 + lane;
      const float fb = (float)G.bins, ci = 1.f / (fb * (1.f + 4.8e-7f));
      const uint32_t lutc = tc::smem_u32(s_lut) - G.lut_bias;
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)grp * 128u;
      constexpr unsigned long long M4 = 0x0F0F0F0F0F0F0F0Full;
      int useq = 0;
      for (int u = u0; u < u1; ++u, ++useq) {
        if ((useq & (HT_BUFS - 1)) == grp) {
          const int rows = min(128, W.n1 - W.ti * 128), nc = min(128, W.n2 - W.tj * 128);
          tc::mbar_wait(t_full + grp, (useq / HT_BUFS) & 1);
          tc::tc_fence_after();
          unsigned long long ev = 0ull, od = 0ull;
          const bool warp_has_rows = quad * 32 < rows, rv = row < rows;
          auto release = [&]() {
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(t_empty + grp);
          };
          auto fold = [&](unsigned long long w) {
            if (rv) { ev += w & M4; od += (w >> 4) & M4; }
          };
          if (!warp_has_rows) release();
          else {
            int c = 0;
#pragma unroll 1
            while (c < nc) {
              const int rem = nc - c;
              if (rem >= 2 * HP_SEG) {
                float v[2 * HP_SEG];
                tc::tmem_ld_n<2 * HP_SEG>(taddr + (uint32_t)c, v);
                tc::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 2 * HP_SEG; ++i) asm volatile("" : "+f"(v[i]));   // uses stay behind the wait
                c += 2 * HP_SEG;
                if (c >= nc) release();
                fold(ht_segment<HP_SEG, true>(v, fb, ci, lutc));
                fold(ht_segment<HP_SEG, true>(v + HP_SEG, fb, ci, lutc));
              } else if (rem >= HP_SEG) {
                float v[HP_SEG];
                tc::tmem_ld_n<HP_SEG>(taddr + (uint32_t)c, v);
                tc::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < HP_SEG; ++i) asm volatile("" : "+f"(v[i]));
                c += HP_SEG;
                if (c >= nc) release();
                fold(ht_segment<HP_SEG, true>(v, fb, ci, lutc));
              } else {
                float v[16];
                // fewer than HP_SEG columns left: load 16 (columns beyond the tile's width hold stale but finite-or-not data
                // that is never binned) and bin the first `rem` through the generic per-cell form
                tc::tmem_ld_n<16>(taddr + (uint32_t)min(c, 112), v);
                tc::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) asm volatile("" : "+f"(v[i]));
                const int sh = c - min(c, 112);
                release();
                unsigned long long w = 0ull;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  if (i >= sh && i < sh + rem) {
                    const float uu = ht_ex2(fminf(v[i], 126.f));
                    ht_count(w, ht_bin_bits(ht_rcp(1.f + uu) + 2.4e-7f, fb), lutc);
                  }
                }
                c = nc;
                fold(w);
              }
            }
            // 16 byte counters (<= 128 each) -> 16-bit fields -> sum over the warp -> one atomic per bin
            const uint32_t w32[4] = {(uint32_t)ev, (uint32_t)(ev >> 32), (uint32_t)od, (uint32_t)(od >> 32)};
            uint32_t ra[4], rb[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              ra[k] = __reduce_add_sync(0xffffffffu, w32[k] & 0x00FF00FFu);
              rb[k] = __reduce_add_sync(0xffffffffu, (w32[k] >> 8) & 0x00FF00FFu);
            }
            if (lane < G.bins) {   // bin l: parity -> ev / od, byte j = l >> 1 of that 64-bit word
              const int j = lane >> 1, k = (lane & 1) * 2 + (j >> 2), byte = j & 3;
              uint32_t x = 0u;
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                if (kk == k) x = (byte & 1) ? rb[kk] : ra[kk];
              }
              const uint32_t val = (x >> (16 * (byte >> 1))) & 0xFFFFu;
              if (val) atomicAdd(counts + (size_t)W.p * G.bins + lane, (int)val);
            }
          }
        }
        if (u + 1 < u1) W.next();
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == HT_PROD_WARP) tc::tmem_dealloc(tmem_base, 512);
}

// the cells the nibble table dropped (sigma == 1, or sigma B == bins when bins < 16) belong to the last bin; normalise
__global__ void __launch_bounds__(256)
hp_finalize_kernel(const HpGeom G, const int32_t* __restrict__ node_off, const int32_t* __restrict__ g1,
                   const int32_t* __restrict__ g2, int32_t* __restrict__ counts, float* __restrict__ hist) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= G.n_pairs) return;
  const int n1 = node_off[g1[p] + 1] - node_off[g1[p]], n2 = node_off[g2[p] + 1] - node_off[g2[p]];
  const int cells = n1 * n2;
  int32_t* c = counts + (size_t)p * G.bins;
  int seen = 0;
  for (int b = 0; b < G.bins; ++b) seen += c[b];
  c[G.bins - 1] += cells - seen;
  if (hist) {
    const float inv = cells > 0 ? 1.f / (float)cells : 0.f;
    for (int b = 0; b < G.bins; ++b) hist[(size_t)p * G.bins + b] = (float)c[b] * inv;
  }
}
