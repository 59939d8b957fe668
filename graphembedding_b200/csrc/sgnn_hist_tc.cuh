// Node-node similarity histogram of ALL (query, database) pairs on the 5th-gen tensor cores.
//
// SimGNN-style histogram layer (SURVEY a24 -- BUILD-DEFINED, the reference snapshot stops at the NTN,
// layers_factory.py:20-33; pin = oracle/siamese_oracle.py: pair_histogram):
//   s_ij = sigmoid(Hq_i . Hg_j),  bin = min(B - 1, floor(s_ij * B)),  counts[q][g][bin] += 1.
// Every (database node row, query node row) cell of the product  [sum N_db x d] . [d x sum N_q]  belongs to
// exactly one (database graph, query graph) pair, so the whole layer is ONE dense GEMM whose epilogue bins the
// accumulator instead of storing it:
//
//   * database side = TMEM lanes: tile t holds the whole graphs whose first node row lies in [S t, S (t + 1)),
//     S = 129 - max_nodes, i.e. <= 128 consecutive node rows (closed form, no scan; ~93 % of the lanes used);
//   * query side = accumulator columns: queries are bucketed by node count (counting sort), a column tile
//     holds floor(128 / n) queries of exactly n nodes, so the epilogue of a tile is compiled for n (<= 15):
//     every thread walks the tile segment by segment with its n cells in registers;
//   * operands are bf16 splits x = p1 + p2 (+ p3) in the UMMA K-major no-swizzle core-matrix order (one tile =
//     one bulk copy); with three parts the six products p1q1 p1q2 p2q1 p2q2 p1q3 p3q1 accumulate in fp32 in
//     TMEM to ~2^-24 relative, i.e. the bins agree with an fp32 dot product (bin indices are integers: the
//     usual bf16x3 error of 2^-16 would move ~2 % of the pairs' counts).  The query rows are pre-scaled by
//     -log2(e), so the accumulator is the argument of ex2 directly;
//   * per cell the epilogue issues  ex2, 1 + u, rcp  (the sigmoid), one  fma.rm(sigma, B, 2^23)  whose low
//     mantissa bits ARE floor(sigma * B), and three integer instructions that add 1 to a 4-bit field of a
//     64-bit register:  t = 1 << bin,  t2 = t * t,  nib += (u64) t2 * t2  (bin == B == 16 for sigma == 1
//     falls off the end of the register and is restored from the known cell count);
//   * a thread's 16 nibble counters of one (lane, query) go to shared memory as ONE 8-byte store; after the
//     tile the 128 threads of the epilogue group turn pair-parallel: thread p sums the <= max_nodes lane words
//     of pair (graph, query), widens them to bytes and writes the pair's 16 counts as one 16-byte store.
//
// Roles (576 threads, 1 CTA per SM, persistent over a contiguous range of (A block, B tile) units): warps
// 0-15 = four epilogue groups of four warps (group e owns TMEM buffer e: lanes 32 (w % 4)..; the counter words are
// double-buffered, one barrier per batch), warp 16 = bulk-copy producer, warp 17 = MMA issuer.
#pragma once
#include "sgnn_common.cuh"
#include "sgnn_tc.cuh"
#include "sgnn_hist_tc_geom.h"

__device__ __forceinline__ void ht_split3(float x, __nv_bfloat16& p1, __nv_bfloat16& p2, __nv_bfloat16& p3) {
  p1 = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(p1);
  p2 = __float2bfloat16_rn(r1);
  p3 = __float2bfloat16_rn(r1 - __bfloat162float(p2));
}

// one node row (d floats, scaled) -> its 16-byte K chunks of every part, at `dst` + part * part_stride + kc * chunk_stride
// (element units); rows of the K-major no-swizzle layout are 8 bf16 = 16 bytes
__device__ __forceinline__ void ht_write_row(const float* __restrict__ src, int d, float scale, int parts, int ksteps,
                                             __nv_bfloat16* __restrict__ dst) {
  const int nkc = 2 * ksteps;
  for (int kc = 0; kc < nkc; ++kc) {
    __align__(16) __nv_bfloat16 p[3][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = kc * 8 + j;
      const float x = (src != nullptr && c < d) ? src[c] * scale : 0.f;
      ht_split3(x, p[0][j], p[1][j], p[2][j]);
    }
    for (int pt = 0; pt < parts; ++pt)
      *reinterpret_cast<uint4*>(dst + ((size_t)pt * nkc + kc) * 1024) = *reinterpret_cast<const uint4*>(p[pt]);
  }
}

// ---- query side: counting sort by node count, tile table ---------------------------------------------------
__global__ void __launch_bounds__(1024)
ht_sort_q_kernel(const int32_t* __restrict__ node_off_q, int m, int max_btiles, int32_t* __restrict__ qmeta,
                 int32_t* __restrict__ status) {
  __shared__ int hist[16], start[16], tstart[16], tcount[16], running[16];
  __shared__ int wtot[32][16], wpre[32][16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 16) { hist[tid] = 0; running[tid] = 0; }
  __syncthreads();
  bool bad = false;
  for (int i = tid; i < m; i += 1024) {
    const int s = node_off_q[i + 1] - node_off_q[i];
    if (s < 0 || s > HT_MAXN) bad = true;
    else if (s > 0) atomicAdd(&hist[s], 1);
  }
  if (bad) atomicExch(status, 1);   // a query graph beyond HT_MAXN nodes: the caller checks max_nodes first, so this is a guard
  __syncthreads();
  if (tid == 0) {
    int o = 0, t = 0;
    for (int s = HT_MAXN; s >= 1; --s) {
      const int per = 128 / s;
      start[s] = o; o += hist[s];
      tstart[s] = t; tcount[s] = (hist[s] + per - 1) / per; t += tcount[s];
    }
    start[0] = o; tstart[0] = t; tcount[0] = 0;
    qmeta[0] = min(t, max_btiles);
  }
  __syncthreads();
  int32_t* bt_n = qmeta + 4;
  int32_t* bt_cnt = bt_n + max_btiles;
  int32_t* bt_first = bt_cnt + max_btiles;
  int32_t* q_sorted = bt_first + max_btiles;
  for (int s = 1; s <= HT_MAXN; ++s) {
    const int per = 128 / s;
    for (int j = tid; j < tcount[s]; j += 1024) {
      const int u = tstart[s] + j;
      if (u < max_btiles) {
        bt_n[u] = s; bt_first[u] = start[s] + j * per; bt_cnt[u] = min(per, hist[s] - j * per);
      }
    }
  }
  // stable ranks inside every size bucket, 1024 queries per round
  for (int base = 0; base < m; base += 1024) {
    const int i = base + tid;
    int s = 0;
    if (i < m) { s = node_off_q[i + 1] - node_off_q[i]; if (s < 0 || s > HT_MAXN) s = 0; }
    if (tid < 512) wtot[tid >> 4][tid & 15] = 0;
    __syncthreads();
    const unsigned same = __match_any_sync(0xffffffffu, s);
    const int prior = __popc(same & ((1u << lane) - 1u));
    if (lane == 31 - __clz(same)) wtot[warp][s] = __popc(same);
    __syncthreads();
    if (tid < 512) {
      const int w = tid >> 4, ss = tid & 15;
      int acc = 0;
      for (int w2 = 0; w2 < w; ++w2) acc += wtot[w2][ss];
      wpre[w][ss] = acc;
    }
    __syncthreads();
    if (s > 0) q_sorted[start[s] + running[s] + wpre[warp][s] + prior] = i;
    __syncthreads();
    if (tid < 16) running[tid] += wpre[31][tid] + wtot[31][tid];
    __syncthreads();
  }
}

// B operand: one CTA per column tile, thread = accumulator column (query node row), rows scaled by -log2(e)
__global__ void __launch_bounds__(128)
ht_prep_q_kernel(const HistTcGeom G, const float* __restrict__ nodes_q, const int32_t* __restrict__ node_off_q,
                 const int32_t* __restrict__ qmeta, __nv_bfloat16* __restrict__ Bt) {
  const int u = blockIdx.x;
  if (u >= qmeta[0]) return;
  const int32_t* bt_n = qmeta + 4;
  const int32_t* bt_cnt = bt_n + G.max_btiles;
  const int32_t* bt_first = bt_cnt + G.max_btiles;
  const int32_t* q_sorted = bt_first + G.max_btiles;
  const int c = threadIdx.x;
  const int n = bt_n[u], seg = c / n, r = c - seg * n;
  const float* src = nullptr;
  if (seg < bt_cnt[u]) {
    const int q = q_sorted[bt_first[u] + seg];
    src = nodes_q + (size_t)(node_off_q[q] + r) * G.d;
  }
  ht_write_row(src, G.d, -1.4426950408889634f, G.parts, G.ksteps,
               Bt + (size_t)u * (ht_tile_bytes(G) / 2) + (size_t)c * 8);
}

// A operand: one CTA per lane tile; also the tile table a_gfirst[t] = first graph whose node offset >= S t
__global__ void __launch_bounds__(128)
ht_prep_db_kernel(const HistTcGeom G, const float* __restrict__ nodes_db, const int32_t* __restrict__ node_off_db,
                  int32_t* __restrict__ qmeta, int32_t* __restrict__ ameta, __nv_bfloat16* __restrict__ At,
                  int32_t* __restrict__ status) {
  __shared__ int s_g[2];
  const int t = blockIdx.x;
  const int total = node_off_db[G.n];
  const int n_at = min(G.max_atiles, total / G.S + 1);
  if (t == 0 && threadIdx.x == 0) qmeta[1] = n_at;
  if (t >= n_at) return;
  if (threadIdx.x < 64) {   // lower_bound over the graph start offsets: warp w searches key (t + w) S, 32 probes per step
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;   // (3 dependent loads for 10 000 graphs instead of 14)
    const int key = (t + w) * G.S;
    int lo = 0, hi = G.n;   // answer in [lo, hi]
    while (lo < hi) {
      const int span = hi - lo;
      const int step = (span + 32) / 33;                     // probes at lo + step (lane + 1) - 1 ...
      const int pos = min(hi - 1, lo + step * (lane + 1) - 1);
      const bool less = node_off_db[pos] < key;              // monotone in pos
      const unsigned bal = __ballot_sync(0xffffffffu, less);
      const int nless = __popc(bal);                         // probes 0 .. nless - 1 are < key
      const int new_lo = nless == 0 ? lo : min(hi - 1, lo + step * nless - 1) + 1;
      const int new_hi = nless == 32 ? hi : min(hi - 1, lo + step * (nless + 1) - 1);
      lo = new_lo; hi = max(new_lo, new_hi);
    }
    if (lane == 0) s_g[w] = lo;
  }
  __syncthreads();
  const int g0 = s_g[0], g1 = s_g[1];
  if (threadIdx.x == 0) {
    ameta[t] = g0;
    if (t == n_at - 1) ameta[n_at] = G.n;   // (the last tile takes every remaining graph)
  }
  const int row0 = node_off_db[g0];
  const int row1 = (t == n_at - 1) ? total : node_off_db[g1];
  if (threadIdx.x == 0 && row1 - row0 > 128) atomicExch(status, 2);   // max_nodes understated by the caller
  const int row = row0 + (int)threadIdx.x;
  const float* src = row < row1 ? nodes_db + (size_t)row * G.d : nullptr;
  ht_write_row(src, G.d, 1.f, G.parts, G.ksteps, At + (size_t)t * (ht_tile_bytes(G) / 2) + (size_t)threadIdx.x * 8);
}

// ---- epilogue helpers ---------------------------------------------------------------------------------------
__device__ __forceinline__ float ht_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ht_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// low mantissa bits of the result = floor(sigma * fb) (round-down fma onto 2^23, where one ulp is 1)
__device__ __forceinline__ uint32_t ht_bin_bits(float sigma, float fb) {
  float y;
  asm("fma.rm.f32 %0, %1, %2, 0f4B000000;" : "=f"(y) : "f"(sigma), "f"(fb));
  return __float_as_uint(y);
}
// nib += 16^bin: bin in the low 5 bits of `bits` (= 0x4B000000 + bin).  The one-hot nibble word comes from a 32-entry
// table in shared memory (entries >= 16 are 0: sigma == 1 adds nothing): one address computation + one LDS.64
// instead of a shift and two multiplies, and it moves the work from the ALU / FMA pipes to the idle LSU.  lutc = table
// address - ((0x4B000000 << 3) mod 2^32), so that bits * 8 + lutc is the entry's address (32-bit wrap-around).
__device__ __forceinline__ void ht_count(unsigned long long& nib, uint32_t bits, uint32_t lutc) {
  unsigned long long one;
  asm("ld.shared.u64 %0, [%1];" : "=l"(one) : "r"(bits * 8u + lutc));
  nib += one;
}

// the n cells of one (lane, query) segment -> 16 nibble counters
// fb = number of bins, ci = 1 / (fb (1 + eps)), eps = 4 ulp: the paired form's operand scale
template <int N, bool PAIRED>
__device__ __forceinline__ unsigned long long ht_segment(const float* v, float fb, float ci, uint32_t lutc) {
  unsigned long long nib = 0ull;
  constexpr float BIAS = 2.4e-7f;
  if constexpr (!PAIRED) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const float u = ht_ex2(v[i]);                 // exp(-s)
      const float sg = ht_rcp(1.f + u);             // sigmoid(s); 1 / inf = 0
      ht_count(nib, ht_bin_bits(sg, fb), lutc);
    }
  } else {
    // one reciprocal per two cells.  With v = (1 + u) ci:  r = 1 / (v1 v2),  r v2 = 1 / v1 = fb (1 + eps) sigma1 and
    // r v1 = fb (1 + eps) sigma2 -- the bin is the round-down fma of that onto 2^23 (one ulp = 1), so a pair costs
    // fma.f32x2, mul, rcp, fma.rm.f32x2.  The arguments are clamped so that the product stays finite; eps (4 ulp,
    // above the error of the approximate reciprocal and the two roundings) keeps sigmoid(0) = 1/2 on the upper side
    // of its edge.  Which sigma belongs to which cell does not matter, both go into the same histogram.
#pragma unroll
    for (int i = 0; i + 1 < N; i += 2) {
      unsigned long long u12, v12, y12, c2, r2, m2;
      asm("mov.b64 %0, {%1, %2};" : "=l"(u12) : "f"(ht_ex2(fminf(v[i], 60.f))), "f"(ht_ex2(fminf(v[i + 1], 60.f))));
      asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(ci));
      asm("fma.rn.f32x2 %0, %1, %2, %2;" : "=l"(v12) : "l"(u12), "l"(c2));
      float v1, v2;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(v1), "=f"(v2) : "l"(v12));
      const float r = ht_rcp(v1 * v2);
      asm("mov.b64 %0, {%1, %1};" : "=l"(r2) : "f"(r));
      asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(8388608.f));
      asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(y12) : "l"(r2), "l"(v12), "l"(m2));
      uint32_t y1, y2;
      asm("mov.b64 {%0, %1}, %2;" : "=r"(y1), "=r"(y2) : "l"(y12));
      ht_count(nib, y1, lutc);
      ht_count(nib, y2, lutc);
    }
    if constexpr (N & 1) {
      const float u = ht_ex2(v[N - 1]);
      ht_count(nib, ht_bin_bits(ht_rcp(1.f + u) + BIAS, fb), lutc);
    }
  }
  return nib;
}

// two count bytes of `word` (PRMT selector sel) as a packed pair of exact fp16 values
__device__ __forceinline__ uint32_t ht_half2(uint32_t word, uint32_t sel) {
  uint32_t r = __byte_perm(word, 0x64646464u, sel);
  asm("sub.rn.f16x2 %0, %0, %1;" : "+r"(r) : "r"(0x64006400u));
  return r;
}

// ceil(65536 / n) for n = 1..16 by table: (p * ht_rcp16(n)) >> 16 == p / n for p < 4096 (an integer division by a run-time
// value costs ~25 instructions, and the pair phase needs one per round)
__constant__ uint32_t c_ht_rcp16[17] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4096};
__device__ __forceinline__ uint32_t ht_rcp16(int n) { return c_ht_rcp16[n]; }

struct HtTileCtx {
  uint32_t taddr;                 // this warp's lanes, this tile's TMEM buffer
  int n_seg, bt_first;            // B tile: segments, first sorted query
  int g0, ng, row0;               // A tile: first graph, graphs, first node row
  unsigned long long* s_cnt;      // [HT_SEGB][128] of this epilogue group
  uint64_t* t_empty;
  int bar_id;
  float fb, ci;                   // bins; 1 / (bins (1 + 4 ulp))
  uint32_t lutc;                  // one-hot table (ht_count)
  int bins;
  const int32_t* q_sorted;
  const int32_t* node_off_db;
  uint8_t* cnt_out;               // [m][n][16] bytes, or fp16 operand tiles (m_pad > 0)
  int n;
  int m_pad;                      // > 0: counts leave as the A operand of the head kernel's second MMA (ht_cnt_tiled_bytes)
};

// One tile for one epilogue group (HT_GROUP_WARPS warps: 4 TMEM lane quadrants x 1 or 2 halves; the warps of a quadrant
// split the tile's segments).  row = this thread's TMEM lane (0..127), tid = its index in the group, half = 0 / 1.
// batch: the group's running batch counter -- the counter words are double-buffered on it, so ONE barrier per batch
// is enough (the words of batch b are rewritten in batch b + 2, behind the barrier of batch b + 1).
template <int N, bool PAIRED>
__device__ __forceinline__ void ht_tile(const HtTileCtx& C, int row, int tid, int half, int lane, int& batch) {
  // the first pair this thread reduces: its lane range is fetched now, long before it is needed
  int pre_l0 = 0, pre_l1 = 0;
  const int pre_gi = (int)(((uint32_t)tid * ht_rcp16(max(1, min(HT_SEGB, C.n_seg)))) >> 16);   // graph of this thread's first pair (first batch)
  if (pre_gi < C.ng) {
    pre_l0 = C.node_off_db[C.g0 + pre_gi] - C.row0;
    pre_l1 = C.node_off_db[C.g0 + pre_gi + 1] - C.row0;
  }
  const uint32_t ng = (uint32_t)C.ng;
  if (C.n_seg == 0) {   // (never produced by the tile table; keeps the barrier protocol whole)
    tc::tc_fence_before();
    __syncwarp();
    if (lane == 0) tc::mbar_arrive(C.t_empty);
  }
  for (int s0 = 0; s0 < C.n_seg; s0 += HT_SEGB, ++batch) {
    const int ns = min(HT_SEGB, C.n_seg - s0);
    unsigned long long* const cnt = C.s_cnt + (batch & 1) * (HT_SEGB * HT_CSTRIDE);
    // two segments per iteration (one TMEM wait and one loop overhead for 2 N cells); iterations alternate
    // between the two halves of the quadrant
    const bool last_batch = s0 + ns == C.n_seg;
    bool released = false;
    constexpr int STEP = 2 * (HT_GROUP_WARPS / 4);
#pragma unroll 1
    for (int sg = 2 * half; sg < ns; sg += STEP) {
      const bool last_read = last_batch && sg + STEP >= ns;   // this warp's last TMEM read of the tile
      if (sg + 1 < ns) {
        float v[2 * N];
        tc::tmem_ld_n<2 * N>(C.taddr + (uint32_t)((s0 + sg) * N), v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 2 * N; ++i) asm volatile("" : "+f"(v[i]));   // uses stay behind the wait
        if (last_read) {   // hand the buffer back to the MMA issuer before binning the last cells: the ~1 500 cycles of
          tc::tc_fence_before();   // the next tile's six chained MMAs overlap them and the pair phase
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(C.t_empty);
          released = true;
        }
        const unsigned long long w0 = ht_segment<N, PAIRED>(v, C.fb, C.ci, C.lutc), w1 = ht_segment<N, PAIRED>(v + N, C.fb, C.ci, C.lutc);
        cnt[sg * HT_CSTRIDE + row] = w0;
        cnt[(sg + 1) * HT_CSTRIDE + row] = w1;
      } else {
        float v[N];
        tc::tmem_ld_n<N>(C.taddr + (uint32_t)((s0 + sg) * N), v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < N; ++i) asm volatile("" : "+f"(v[i]));
        if (last_read) {
          tc::tc_fence_before();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(C.t_empty);
          released = true;
        }
        cnt[sg * HT_CSTRIDE + row] = ht_segment<N, PAIRED>(v, C.fb, C.ci, C.lutc);
      }
    }
    if (last_batch && !released) {   // (a warp without segments in the last batch)
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(C.t_empty);
    }
    asm volatile("bar.sync %0, %1;" ::"r"(C.bar_id), "n"(32 * HT_GROUP_WARPS) : "memory");
    // ---- pair-parallel: thread p owns (graph gi, segment sg) of this batch ----
    // pair p = (graph p / ns, segment p % ns): the threads of a warp share two or three graphs, so their lane loops
    // have (nearly) the same length -- with the graph index fastest every warp ran for its longest graph (avg 13 of
    // 32 threads active in the loop).  The counter rows are HT_CSTRIDE = 129 words apart: 32 segments of one lane
    // sit in different banks.
    const uint32_t n_pairs = ng * (uint32_t)ns;
    const uint32_t rcp_ns = ht_rcp16(ns);
    for (uint32_t p = (uint32_t)tid; p < n_pairs; p += 32u * HT_GROUP_WARPS) {
      const uint32_t gi = (p * rcp_ns) >> 16, sg = p - gi * (uint32_t)ns;   // p / ns, p % ns (exact: p < 4096, ns <= 16)
      int l0, l1;
      if (s0 == 0 && p == (uint32_t)tid) { l0 = pre_l0; l1 = pre_l1; }   // the prefetched range
      else {
        l0 = C.node_off_db[C.g0 + (int)gi] - C.row0;
        l1 = C.node_off_db[C.g0 + (int)gi + 1] - C.row0;
      }
      unsigned long long ev = 0ull, od = 0ull;
      const unsigned long long* col = cnt + sg * HT_CSTRIDE;
      for (int l = l0; l < l1; ++l) {
        const unsigned long long w = col[l];
        ev += w & 0x0F0F0F0F0F0F0F0Full;
        od += (w >> 4) & 0x0F0F0F0F0F0F0F0Full;
      }
      const uint32_t el = (uint32_t)ev, eh = (uint32_t)(ev >> 32), ol = (uint32_t)od, oh = (uint32_t)(od >> 32);
      uint4 o;
      o.x = __byte_perm(el, ol, 0x5140); o.y = __byte_perm(el, ol, 0x7362);
      o.z = __byte_perm(eh, oh, 0x5140); o.w = __byte_perm(eh, oh, 0x7362);
      // sigma == 1 gives field `bins`: for 16 bins it fell off the register (restored from the cell count), for
      // fewer bins it sits in the next field; either way those cells belong to the last bin
      int seen = __dp4a(o.x, 0x01010101u, 0u);
      seen = __dp4a(o.y, 0x01010101u, (unsigned)seen);
      seen = __dp4a(o.z, 0x01010101u, (unsigned)seen);
      seen = __dp4a(o.w, 0x01010101u, (unsigned)seen);
      uint32_t extra = (uint32_t)((l1 - l0) * N - seen);
      auto word = [&](int wi) { return wi == 0 ? o.x : (wi == 1 ? o.y : (wi == 2 ? o.z : o.w)); };
      auto add_field = [&](int idx, uint32_t val) {
        const uint32_t a = val << ((idx & 3) * 8);
        const int wi = idx >> 2;
        o.x += wi == 0 ? a : 0u; o.y += wi == 1 ? a : 0u; o.z += wi == 2 ? a : 0u; o.w += wi == 3 ? a : 0u;
      };
      if (C.bins < HT_MAX_BINS) {
        const uint32_t over = (word(C.bins >> 2) >> ((C.bins & 3) * 8)) & 0xFFu;
        add_field(C.bins, 0u - over);
        extra += over;
      }
      if (extra) add_field(C.bins - 1, extra);   // (rare: only cells whose sigmoid rounds to 1)
      const int q = C.q_sorted[C.bt_first + s0 + (int)sg];
      if (C.m_pad > 0) {
        // fp16 K-major operand tiles [g / 128][q][k chunk][g % 128][8 bins]: half(1024 + c) has the bits 0x6400 | c, so
        // a count byte becomes an exact fp16 with one PRMT and half a packed subtraction
        const int g = C.g0 + (int)gi;
        uint4* dst = reinterpret_cast<uint4*>(C.cnt_out) + ((size_t)(g >> 7) * C.m_pad + q) * 256 + (g & 127);
        dst[0] = make_uint4(ht_half2(o.x, 0x4140), ht_half2(o.x, 0x4342), ht_half2(o.y, 0x4140), ht_half2(o.y, 0x4342));
        dst[128] = make_uint4(ht_half2(o.z, 0x4140), ht_half2(o.z, 0x4342), ht_half2(o.w, 0x4140), ht_half2(o.w, 0x4342));
      } else {
        *reinterpret_cast<uint4*>(C.cnt_out + ((size_t)q * C.n + (size_t)(C.g0 + (int)gi)) * HT_CNT_BYTES) = o;
      }
    }
  }
}

template <bool PAIRED>
__device__ __forceinline__ void ht_tile_dispatch(int N, const HtTileCtx& C, int row, int tid, int half, int lane, int& batch) {
  switch (N) {
    case 1: ht_tile<1, PAIRED>(C, row, tid, half, lane, batch); break;
    case 2: ht_tile<2, PAIRED>(C, row, tid, half, lane, batch); break;
    case 3: ht_tile<3, PAIRED>(C, row, tid, half, lane, batch); break;
    case 4: ht_tile<4, PAIRED>(C, row, tid, half, lane, batch); break;
    case 5: ht_tile<5, PAIRED>(C, row, tid, half, lane, batch); break;
    case 6: ht_tile<6, PAIRED>(C, row, tid, half, lane, batch); break;
    case 7: ht_tile<7, PAIRED>(C, row, tid, half, lane, batch); break;
    case 8: ht_tile<8, PAIRED>(C, row, tid, half, lane, batch); break;
    case 9: ht_tile<9, PAIRED>(C, row, tid, half, lane, batch); break;
    case 10: ht_tile<10, PAIRED>(C, row, tid, half, lane, batch); break;
    case 11: ht_tile<11, PAIRED>(C, row, tid, half, lane, batch); break;
    case 12: ht_tile<12, PAIRED>(C, row, tid, half, lane, batch); break;
    case 13: ht_tile<13, PAIRED>(C, row, tid, half, lane, batch); break;
    case 14: ht_tile<14, PAIRED>(C, row, tid, half, lane, batch); break;
    default: ht_tile<15, PAIRED>(C, row, tid, half, lane, batch); break;
  }
}

// ---- the persistent kernel ----------------------------------------------------------------------------------
template <bool PAIRED>
__global__ void __launch_bounds__(HT_THREADS, 1)
hist_tc_kernel(const HistTcGeom G, const __nv_bfloat16* __restrict__ At, const __nv_bfloat16* __restrict__ Bt,
               const int32_t* __restrict__ qmeta, const int32_t* __restrict__ ameta,
               const int32_t* __restrict__ node_off_db, uint8_t* __restrict__ cnt_out) {
  extern __shared__ __align__(1024) uint8_t ht_smem[];
  const int tile_bytes = ht_tile_bytes(G);
  uint8_t* sA = ht_smem;
  uint8_t* sB = sA + HT_GA * tile_bytes;
  unsigned long long* s_cnt = reinterpret_cast<unsigned long long*>(sB + HT_STAGES * tile_bytes);   // [HT_GROUPS][2][HT_SEGB][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_cnt + HT_GROUPS * 2 * HT_SEGB * HT_CSTRIDE);
  uint64_t* a_full = bars;                    // [1]
  uint64_t* a_free = bars + 1;                // [1]
  uint64_t* b_full = bars + 2;                // [HT_STAGES]
  uint64_t* b_empty = b_full + HT_STAGES;     // [HT_STAGES]
  uint64_t* t_full = b_empty + HT_STAGES;     // [HT_BUFS]
  uint64_t* t_empty = t_full + HT_BUFS;       // [HT_BUFS]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + HT_BUFS);
  unsigned long long* s_lut = reinterpret_cast<unsigned long long*>(tmem_slot + 2);   // [32] one-hot nibble words

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 32) s_lut[threadIdx.x] = threadIdx.x < 16 ? 1ull << (4 * threadIdx.x) : 0ull;
  const int n_b = qmeta[0], n_at = qmeta[1];
  const int n_ab = (n_at + HT_GA - 1) / HT_GA;
  const long long total_units = (long long)n_ab * n_b;
  const long long u0 = total_units * blockIdx.x / gridDim.x, u1 = total_units * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    tc::mbar_init(a_full, 1);
    tc::mbar_init(a_free, 1);
    for (int s = 0; s < HT_STAGES; ++s) { tc::mbar_init(b_full + s, 1); tc::mbar_init(b_empty + s, 1); }
    for (int b = 0; b < HT_BUFS; ++b) { tc::mbar_init(t_full + b, 1); tc::mbar_init(t_empty + b, HT_GROUP_WARPS); }
    tc::fence_barrier_init();
  }
  if (warp == HT_PROD_WARP) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (u0 < u1) {
    const int ab0 = (int)(u0 / n_b), b0 = (int)(u0 - (long long)ab0 * n_b);
    if (warp == HT_PROD_WARP) {
      // ===== producer: A block whenever it changes, one B tile per unit =====
      if (lane == 0) {
        int ab = ab0, b = b0, ab_seq = 0, useq = 0, cur_ab = -1;
        for (long long u = u0; u < u1; ++u) {
          if (ab != cur_ab) {
            if (ab_seq > 0) tc::mbar_wait(a_free, (ab_seq - 1) & 1);   // every MMA on the previous block has completed
            const int ga = min(HT_GA, n_at - ab * HT_GA);
            tc::mbar_arrive_expect_tx(a_full, ga * tile_bytes);
            for (int a = 0; a < ga; ++a)
              tc::bulk_g2s(sA + a * tile_bytes, reinterpret_cast<const uint8_t*>(At) + (size_t)(ab * HT_GA + a) * tile_bytes,
                           tile_bytes, a_full);
            cur_ab = ab; ++ab_seq;
          }
          const int s = useq % HT_STAGES;
          if (useq >= HT_STAGES) tc::mbar_wait(b_empty + s, ((useq / HT_STAGES) - 1) & 1);
          tc::mbar_arrive_expect_tx(b_full + s, tile_bytes);
          tc::bulk_g2s(sB + s * tile_bytes, reinterpret_cast<const uint8_t*>(Bt) + (size_t)b * tile_bytes, tile_bytes,
                       b_full + s);
          ++useq;
          if (++b == n_b) { b = 0; ++ab; }
        }
      }
    } else if (warp == HT_MMA_WARP) {
      // ===== MMA issuer (one thread) =====
      if (lane == 0) {
        const uint32_t idesc = tc::umma_idesc_bf16(128, 128);
        const uint64_t a_desc0 = tc::umma_desc(tc::smem_u32(sA), 128 * 16, 128);
        const uint64_t b_desc0 = tc::umma_desc(tc::smem_u32(sB), 128 * 16, 128);
        const uint32_t tile16 = (uint32_t)tile_bytes >> 4;            // descriptor units of 16 bytes
        const uint32_t part16 = (uint32_t)G.ksteps * 256u, kstep16 = 256u;
        const int nprod = G.parts == 1 ? 1 : (G.parts == 2 ? 3 : 6);
        // (a part, b part) of the products, smallest terms last
        const uint32_t pa = 0x201100u;   // nibbles: 0,0,1,1,0,2  -> a parts of product 0..5
        const uint32_t pb = 0x021010u;   // nibbles: 0,1,0,1,2,0  -> b parts
        int ab = ab0, b = b0, ab_seq = 0, useq = 0, cur_ab = -1, it = 0;
        for (long long u = u0; u < u1; ++u) {
          const int ga = min(HT_GA, n_at - ab * HT_GA);
          if (ab != cur_ab) {
            tc::mbar_wait(a_full, ab_seq & 1);
            cur_ab = ab; ++ab_seq;
          }
          const int s = useq % HT_STAGES;
          tc::mbar_wait(b_full + s, (useq / HT_STAGES) & 1);
          tc::tc_fence_after();
          const uint64_t bd = b_desc0 + (uint64_t)((uint32_t)s * tile16);
          for (int a = 0; a < ga; ++a, ++it) {
            const int buf = it & (HT_BUFS - 1);
            if (it >= HT_BUFS) {
              tc::mbar_wait(t_empty + buf, ((it / HT_BUFS) - 1) & 1);
              tc::tc_fence_after();
            }
            const uint64_t ad = a_desc0 + (uint64_t)((uint32_t)a * tile16);
            const uint32_t d_addr = tmem_base + (uint32_t)buf * 128u;
            uint32_t acc = 0;
            for (int p = 0; p < nprod; ++p) {
              const uint32_t ao = ((pa >> (4 * p)) & 15u) * part16, bo = ((pb >> (4 * p)) & 15u) * part16;
              for (int ks = 0; ks < G.ksteps; ++ks) {
                tc::umma_bf16(d_addr, ad + ao + (uint64_t)(ks * kstep16), bd + bo + (uint64_t)(ks * kstep16), idesc, acc);
                acc = 1;
              }
            }
            tc::umma_commit(t_full + buf);
          }
          tc::umma_commit(b_empty + s);   // this unit's B tile is free once its MMAs have completed
          ++useq;
          const bool last_of_block = (b + 1 == n_b) || (u + 1 == u1);
          if (last_of_block) tc::umma_commit(a_free);
          if (++b == n_b) { b = 0; ++ab; }
        }
      }
    } else {
      // ===== epilogue: HT_GROUPS groups of HT_GROUP_WARPS warps; group g takes the tiles it = g, g + HT_GROUPS, ...
      // (TMEM buffer it % 4).  Measured (r2j): two groups of eight warps, each alternating between two TMEM buffers so
      // that the MMA of its next tile is hidden, LOSE to four groups of four (0.76 vs 0.65 ms): the 256-thread barrier
      // and the 8-warp pair phase cost more than the ~9 % of the time a group waits for its accumulator. =====
      const int grp = warp / HT_GROUP_WARPS, quad = warp & 3, half = (warp % HT_GROUP_WARPS) >> 2;
      const int row = quad * 32 + lane, tid = (warp % HT_GROUP_WARPS) * 32 + lane;
      const int32_t* bt_n = qmeta + 4;
      const int32_t* bt_cnt = bt_n + G.max_btiles;
      const int32_t* bt_first = bt_cnt + G.max_btiles;
      HtTileCtx C;
      C.s_cnt = s_cnt + grp * (2 * HT_SEGB * HT_CSTRIDE);   // two batches deep
      C.bar_id = 1 + grp;
      C.fb = (float)G.bins; C.bins = G.bins;
      C.ci = 1.f / (C.fb * (1.f + 4.8e-7f));
      C.lutc = tc::smem_u32(s_lut) - G.lut_bias;   // (a kernel argument: ptxas must not split it into base + offset + immediate)
      C.q_sorted = bt_first + G.max_btiles;
      C.node_off_db = node_off_db;
      C.cnt_out = cnt_out; C.n = G.n; C.m_pad = G.cnt_m_pad;
      int ab = ab0, b = b0, it = 0, batch = 0;
      for (long long u = u0; u < u1; ++u) {
        const int ga = min(HT_GA, n_at - ab * HT_GA);
        // (ga <= HT_GROUPS: at most one tile of a unit belongs to this group -- no walk over the other groups' tiles)
        static_assert(HT_GA <= HT_GROUPS && (HT_GROUPS & (HT_GROUPS - 1)) == 0, "one tile per unit and group");
        const int a = (grp - it) & (HT_GROUPS - 1);
        const int it_unit = it;
        it += ga;
        if (a < ga) {
          const int it = it_unit + a;
          const int buf = it & (HT_BUFS - 1);
          const int t = ab * HT_GA + a;
          C.g0 = ameta[t]; C.ng = ameta[t + 1] - C.g0; C.row0 = node_off_db[C.g0];
          C.n_seg = bt_cnt[b]; C.bt_first = bt_first[b];
          C.taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)buf * 128u;
          C.t_empty = t_empty + buf;
          tc::mbar_wait(t_full + buf, (it / HT_BUFS) & 1);
          tc::tc_fence_after();
          ht_tile_dispatch<PAIRED>(bt_n[b], C, row, tid, half, lane, batch);
        }
        if (++b == n_b) { b = 0; ++ab; }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == HT_PROD_WARP) tc::tmem_dealloc(tmem_base, 512);
}
