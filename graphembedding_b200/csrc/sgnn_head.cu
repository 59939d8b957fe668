// All-pairs scoring for models with a SimGNN-style head (include/sgnn.h: NTN K-vector | node-node
// similarity histogram -> Dense layers).  The head and the histogram are BUILD-DEFINED (absent from the
// reference snapshot); the NTN slab is layers.py:282-310, the Dense rows
// layers.py:192-205, the m x n loop train.py:57-61.
//
// CUDA-core kernel: thread = one database graph.  The graph's pooled embedding lives in registers, its
// last node-level rows in a private shared-memory stripe (stride = 4 * odd words, so 16-byte reads by the
// 32 lanes of a warp hit 32 different bank groups).  The CTA walks a block of queries; per query the
// factorised NTN tables P'/C (sgnn_allpairs.cuh), the query's node rows and the head weights are read as
// warp-uniform broadcasts.  Per pair:
//   f_k   = U_k * inneract(C[q][k] + P'[q][k][:] . x_g)                       (2 D K flops)
//   s_ij  = sigmoid(Hq_i . Hg_j), bin = min(bins - 1, floor(s_ij * bins))     (2 N_q N_g d flops)
//   head  = Dense stack on [f | counts / (N_q N_g)]
// The histogram arithmetic (sequential fmaf dot, 1 / (1 + expf(-s)), floorf) is the same as in
// pair_histogram_kernel and the training kernel, so all three agree on every bin index; zero padding only
// ever adds fmaf(x, 0, s) = s terms.  Per-thread bin counters are a private shared-memory column.
#include <algorithm>
#include "sgnn_common.cuh"
#include "sgnn_internal.h"

struct HeadPlan {
  int T;             // database graphs per CTA (the CTA has 2 T threads: two query lanes per graph)
  int S;             // words per database-graph stripe (node rows, stride ND each)
  int nmax_q;        // node rows reserved per staged query
  int hw_floats;     // head weight region
  int q_floats;      // one staged query: P' [K][DEP] | C [K] | rows [nmax_q + 1][ND] | n (1, padded to 4)
  int w0, b0;        // offsets inside the head weight region: layer 0 weights [head_in][H0], bias [H0]
  int wl, bl;        // later layers: [n_head-1][H0][H0], [n_head-1][H0]
};

constexpr int HEAD_QC = 4;   // queries staged per round (two per query lane)

template <int H0>
__host__ __device__ inline void head_plan_weights(const ModelDev& M, HeadPlan& P) {
  int o = round_up4(M.K);          // U
  P.w0 = o; o += M.head_in * H0;
  P.b0 = o; o += H0;
  P.wl = o; o += (M.n_head - 1) * H0 * H0;
  P.bl = o; o += (M.n_head - 1) * H0;
  P.hw_floats = o;
}

// Bin counters of one pair, in registers.  A cell adds 1 to a 4-bit field of `nib` (16 bins per 64-bit word);
// every 12 cells the nibbles are folded into 8-bit fields (even / odd bins), and those spill into the thread's
// shared-memory column only when a pair has more than ~240 cells (graphs beyond AIDS size).
template <int NB>
struct BinCounters {
  static constexpr int W = NB / 16;
  unsigned long long nib[W], ev[W], od[W];
  int since_nib, since_byte;
  bool spilled;
  __device__ __forceinline__ void reset() {
#pragma unroll
    for (int w = 0; w < W; ++w) nib[w] = ev[w] = od[w] = 0ull;
    since_nib = since_byte = 0; spilled = false;
  }
  __device__ __forceinline__ void add(int bin, bool valid) {
    if (W == 1) {
      nib[0] += (unsigned long long)valid << (bin * 4);
    } else {
#pragma unroll
      for (int w = 0; w < W; ++w) nib[w] += (unsigned long long)(valid && (bin >> 4) == w) << ((bin & 15) * 4);
    }
  }
  __device__ __forceinline__ void fold() {
#pragma unroll
    for (int w = 0; w < W; ++w) {
      ev[w] += nib[w] & 0x0F0F0F0F0F0F0F0Full;
      od[w] += (nib[w] >> 4) & 0x0F0F0F0F0F0F0F0Full;
      nib[w] = 0ull;
    }
    since_byte += since_nib; since_nib = 0;
  }
  __device__ __forceinline__ int byte_of(int b) const {   // b: compile-time after unrolling
    const unsigned long long v = (b & 1) ? od[b >> 4] : ev[b >> 4];
    return (int)((v >> (((b & 15) >> 1) * 8)) & 0xFFull);
  }
  // f(b, count) for every bin b < bins, then the byte fields are cleared
  template <typename F>
  __device__ __forceinline__ void drain(int bins, F f) {
    for (int b = 0; b < bins; ++b) {
      int v = 0;
#pragma unroll
      for (int c = 0; c < NB; ++c) if (c == b) v = byte_of(c);
      f(b, v);
    }
#pragma unroll
    for (int w = 0; w < W; ++w) ev[w] = od[w] = 0ull;
    since_byte = 0;
  }
  __device__ __forceinline__ void spill(int* col, int T, int bins) {
    const bool first = !spilled;
    drain(bins, [&](int b, int v) { col[b * T] = (first ? 0 : col[b * T]) + v; });
    spilled = true;
  }
  // after every group of up to 4 cells
  __device__ __forceinline__ void tick(int* col, int T, int bins) {
    since_nib += 4;
    if (since_nib > 11) {
      fold();
      if (since_byte > 240) spill(col, T, bins);
    }
  }
  // same, draining into f when the byte fields fill up
  template <typename F>
  __device__ __forceinline__ void tick_drain(int bins, F f) {
    since_nib += 4;
    if (since_nib > 11) {
      fold();
      if (since_byte > 240) drain(bins, f);
    }
  }
};

// PRE: the bin counts of every pair were computed beforehand (tensor-core histogram, sgnn_hist_tc.cuh) and are read
// from cnt_in [m][n][16 bytes]; the node rows are then not staged and the cell loop is skipped.
template <int DEP, int H0, int ND, int NB, bool PRE>
__global__ void __launch_bounds__(256)
allpairs_head_kernel(const __grid_constant__ ModelDev M, const __grid_constant__ HeadPlan HP,
                     const float* __restrict__ params, const float* __restrict__ Pq, const float* __restrict__ Cq,
                     const float* __restrict__ emb_db, const float* __restrict__ nodes_q,
                     const int32_t* __restrict__ node_off_q, const float* __restrict__ nodes_db,
                     const int32_t* __restrict__ node_off_db, int m, int n, int q_per_block,
                     float* __restrict__ sim, int64_t ld, const uint8_t* __restrict__ cnt_in) {
  extern __shared__ __align__(16) float smem[];
  const int NT = blockDim.x, tid = threadIdx.x, T = HP.T;
  const int gl = tid % T, ql = tid / T;   // graph lane, query lane (uniform per warp: T is a multiple of 32)
  const int K = M.K, D = M.emb_dim, d = M.node_dim, bins = M.hist_bins;
  float* s_hw = smem;
  float* s_q = s_hw + HP.hw_floats;                       // [HEAD_QC][q_floats]
  float* s_rows = s_q + HEAD_QC * HP.q_floats;            // [T][S]
  // behind the stripes: ND words of padding, then [bins][2 T] spill columns (only when a pair can exceed 240 cells)
  int* s_cnt = reinterpret_cast<int*>(s_rows + (size_t)T * HP.S + ND);

  // ---- head weights, zero-padded to H0 columns (and H0 rows for the later layers) ----
  for (int i = tid; i < HP.hw_floats; i += NT) s_hw[i] = 0.f;
  __syncthreads();
  for (int k = tid; k < K; k += NT) s_hw[k] = params[M.u_off + k];
  {
    const LayerDev& L = M.Hd[0];
    for (int i = tid; i < L.din * L.dout; i += NT) {
      const int k = i / L.dout, j = i - k * L.dout;
      s_hw[HP.w0 + k * H0 + j] = params[L.w_off + i];
    }
    if (L.bias) for (int j = tid; j < L.dout; j += NT) s_hw[HP.b0 + j] = params[L.b_off + j];
  }
  for (int l = 1; l < M.n_head; ++l) {
    const LayerDev& L = M.Hd[l];
    for (int i = tid; i < L.din * L.dout; i += NT) {
      const int k = i / L.dout, j = i - k * L.dout;
      s_hw[HP.wl + ((l - 1) * H0 + k) * H0 + j] = params[L.w_off + i];
    }
    if (L.bias) for (int j = tid; j < L.dout; j += NT) s_hw[HP.bl + (l - 1) * H0 + j] = params[L.b_off + j];
  }

  // ---- this CTA's database graphs: embedding -> registers, node rows -> private stripes ----
  const int g0 = blockIdx.x * T;
  const int g = g0 + gl;
  float x[DEP];
#pragma unroll
  for (int b = 0; b < DEP; ++b) x[b] = (g < n && b < D) ? emb_db[(size_t)g * D + b] : 0.f;
  const int n_g = (g < n && bins > 0) ? node_off_db[g + 1] - node_off_db[g] : 0;
  if (bins > 0 && !PRE) {
    const int lane = tid & 31, warp = tid >> 5, nw = NT >> 5;
    for (int gi = warp; gi < T && g0 + gi < n; gi += nw) {
      const int base = node_off_db[g0 + gi];
      const int cnt = node_off_db[g0 + gi + 1] - base;
      float* dst = s_rows + (size_t)gi * HP.S;
      // (for an odd node count the 2 x 2 cell blocks read one row past the last: whatever lies there -- the
      // next stripe or the tail padding -- is computed on and discarded by the validity flags)
      for (int idx = lane; idx < cnt * ND; idx += 32) {
        const int r = idx / ND, c = idx - r * ND;
        dst[idx] = c < d ? nodes_db[(size_t)(base + r) * d + c] : 0.f;
      }
    }
  }
  const float* my_rows = s_rows + (size_t)gl * HP.S;
  int* my_cnt = s_cnt + tid;
  const int act = M.pair_act;
  const int q_begin = blockIdx.y * q_per_block;
  const int q_end = min(m, q_begin + q_per_block);

  for (int q0 = q_begin; q0 < q_end; q0 += HEAD_QC) {
    const int nq = min(HEAD_QC, q_end - q0);
    __syncthreads();
    for (int qi = 0; qi < nq; ++qi) {
      float* sq = s_q + qi * HP.q_floats;
      const float4* src = reinterpret_cast<const float4*>(Pq + (size_t)(q0 + qi) * K * DEP);
      float4* dst = reinterpret_cast<float4*>(sq);
      for (int i = tid; i < K * DEP / 4; i += NT) dst[i] = src[i];
      float* sc = sq + K * DEP;
      for (int k = tid; k < K; k += NT) sc[k] = Cq[(size_t)(q0 + qi) * K + k];
      float* sr = sc + round_up4(K);
      if (bins > 0) {
        const int base = node_off_q[q0 + qi];
        const int cnt = node_off_q[q0 + qi + 1] - base;
        if (!PRE)
          for (int idx = tid; idx < (cnt + 1) * ND; idx += NT) {
            const int r = idx / ND, c = idx - r * ND;
            sr[idx] = (r < cnt && c < d) ? nodes_q[(size_t)(base + r) * d + c] : 0.f;
          }
        if (tid == 0) reinterpret_cast<int*>(sr + (HP.nmax_q + 1) * ND)[0] = cnt;
      }
    }
    __syncthreads();
    if (g >= n) continue;
    for (int qi = ql; qi < nq; qi += 2) {
      const float* sq = s_q + qi * HP.q_floats;
      const float* sc = sq + K * DEP;
      const float* sr = sc + round_up4(K);
      const int n_q = bins > 0 ? reinterpret_cast<const int*>(sr + (HP.nmax_q + 1) * ND)[0] : 0;
      float acc[H0];
#pragma unroll
      for (int j = 0; j < H0; ++j) acc[j] = 0.f;
      // ---- NTN K-vector, folded straight into head layer 0 (row k of W0) ----
#pragma unroll 2
      for (int k = 0; k < K; ++k) {
        const float4* p4 = reinterpret_cast<const float4*>(sq + k * DEP);
        float mk = sc[k];
#pragma unroll
        for (int b = 0; b < DEP / 4; ++b) {
          const float4 p = p4[b];
          mk = fmaf(p.x, x[4 * b + 0], mk);
          mk = fmaf(p.y, x[4 * b + 1], mk);
          mk = fmaf(p.z, x[4 * b + 2], mk);
          mk = fmaf(p.w, x[4 * b + 3], mk);
        }
        const float f = s_hw[k] * sgnn_act(mk, act, M.yeta);
        const float4* w4 = reinterpret_cast<const float4*>(s_hw + HP.w0 + k * H0);
#pragma unroll
        for (int j = 0; j < H0 / 4; ++j) {
          const float4 w = w4[j];
          acc[4 * j + 0] = fmaf(f, w.x, acc[4 * j + 0]);
          acc[4 * j + 1] = fmaf(f, w.y, acc[4 * j + 1]);
          acc[4 * j + 2] = fmaf(f, w.z, acc[4 * j + 2]);
          acc[4 * j + 3] = fmaf(f, w.w, acc[4 * j + 3]);
        }
      }
      // ---- node-node similarity histogram ----
      if (PRE && bins > 0) {
        const uint4 cw = *reinterpret_cast<const uint4*>(cnt_in + ((size_t)(q0 + qi) * n + g) * 16);
        const int cells = n_q * n_g;
        const float inv_cells = cells > 0 ? 1.f / (float)cells : 0.f;
#pragma unroll
        for (int b = 0; b < 16; ++b) {
          if (b < bins) {
            const unsigned wv = (b >> 2) == 0 ? cw.x : ((b >> 2) == 1 ? cw.y : ((b >> 2) == 2 ? cw.z : cw.w));
            const float hb = (float)((wv >> ((b & 3) * 8)) & 0xFFu) * inv_cells;
            const float4* w4 = reinterpret_cast<const float4*>(s_hw + HP.w0 + (K + b) * H0);
#pragma unroll
            for (int j = 0; j < H0 / 4; ++j) {
              const float4 w = w4[j];
              acc[4 * j + 0] = fmaf(hb, w.x, acc[4 * j + 0]);
              acc[4 * j + 1] = fmaf(hb, w.y, acc[4 * j + 1]);
              acc[4 * j + 2] = fmaf(hb, w.z, acc[4 * j + 2]);
              acc[4 * j + 3] = fmaf(hb, w.w, acc[4 * j + 3]);
            }
          }
        }
      }
      // 2 x 2 blocks of cells = four independent chains
      if (!PRE && bins > 0) {
        BinCounters<NB> bc;
        bc.reset();
        const float fb = (float)bins;
        const int bmax = bins - 1;
        for (int j = 0; j < n_g; j += 2) {
          float ha[ND], hb[ND];
          const float4* ra = reinterpret_cast<const float4*>(my_rows + j * ND);
          const float4* rb = reinterpret_cast<const float4*>(my_rows + (j + 1) * ND);
#pragma unroll
          for (int c = 0; c < ND / 4; ++c) {
            const float4 u = ra[c], v = rb[c];
            ha[4 * c] = u.x; ha[4 * c + 1] = u.y; ha[4 * c + 2] = u.z; ha[4 * c + 3] = u.w;
            hb[4 * c] = v.x; hb[4 * c + 1] = v.y; hb[4 * c + 2] = v.z; hb[4 * c + 3] = v.w;
          }
          const bool vj = j + 1 < n_g;
          for (int i = 0; i < n_q; i += 2) {
            const float4* qa = reinterpret_cast<const float4*>(sr + i * ND);
            const float4* qb = reinterpret_cast<const float4*>(sr + (i + 1) * ND);
            float s00 = 0.f, s01 = 0.f, s10 = 0.f, s11 = 0.f;
#pragma unroll
            for (int c = 0; c < ND / 4; ++c) {
              const float4 u = qa[c], v = qb[c];
              s00 = fmaf(u.x, ha[4 * c], s00); s01 = fmaf(u.x, hb[4 * c], s01);
              s10 = fmaf(v.x, ha[4 * c], s10); s11 = fmaf(v.x, hb[4 * c], s11);
              s00 = fmaf(u.y, ha[4 * c + 1], s00); s01 = fmaf(u.y, hb[4 * c + 1], s01);
              s10 = fmaf(v.y, ha[4 * c + 1], s10); s11 = fmaf(v.y, hb[4 * c + 1], s11);
              s00 = fmaf(u.z, ha[4 * c + 2], s00); s01 = fmaf(u.z, hb[4 * c + 2], s01);
              s10 = fmaf(v.z, ha[4 * c + 2], s10); s11 = fmaf(v.z, hb[4 * c + 2], s11);
              s00 = fmaf(u.w, ha[4 * c + 3], s00); s01 = fmaf(u.w, hb[4 * c + 3], s01);
              s10 = fmaf(v.w, ha[4 * c + 3], s10); s11 = fmaf(v.w, hb[4 * c + 3], s11);
            }
            const bool vi = i + 1 < n_q;
            bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s00) * fb)), true);
            bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s01) * fb)), vj);
            bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s10) * fb)), vi);
            bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s11) * fb)), vi && vj);
            bc.tick(my_cnt, NT, bins);
          }
        }
        bc.fold();
        const int cells = n_q * n_g;
        const float inv_cells = cells > 0 ? 1.f / (float)cells : 0.f;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          if (b < bins) {
            const int cnt = bc.byte_of(b) + (bc.spilled ? my_cnt[b * NT] : 0);
            const float hb = (float)cnt * inv_cells;
            const float4* w4 = reinterpret_cast<const float4*>(s_hw + HP.w0 + (K + b) * H0);
#pragma unroll
            for (int j = 0; j < H0 / 4; ++j) {
              const float4 w = w4[j];
              acc[4 * j + 0] = fmaf(hb, w.x, acc[4 * j + 0]);
              acc[4 * j + 1] = fmaf(hb, w.y, acc[4 * j + 1]);
              acc[4 * j + 2] = fmaf(hb, w.z, acc[4 * j + 2]);
              acc[4 * j + 3] = fmaf(hb, w.w, acc[4 * j + 3]);
            }
          }
        }
      
      }
      // ---- rest of the head ----
      {
        const int a0 = M.Hd[0].act;
#pragma unroll
        for (int j = 0; j < H0; ++j) acc[j] = sgnn_act(acc[j] + s_hw[HP.b0 + j], a0, M.yeta);
      }
      for (int l = 1; l < M.n_head; ++l) {
        const float* w = s_hw + HP.wl + (l - 1) * H0 * H0;
        const float* bl = s_hw + HP.bl + (l - 1) * H0;
        const int din = M.Hd[l].din, dout = M.Hd[l].dout, al = M.Hd[l].act;
        float nxt[H0];
#pragma unroll
        for (int j = 0; j < H0; ++j) nxt[j] = 0.f;
#pragma unroll
        for (int k = 0; k < H0; ++k) {
          if (k < din) {
            const float4* w4 = reinterpret_cast<const float4*>(w + k * H0);
#pragma unroll
            for (int j = 0; j < H0 / 4; ++j) {
              if (4 * j < dout) {
                const float4 v = w4[j];
                nxt[4 * j + 0] = fmaf(acc[k], v.x, nxt[4 * j + 0]);
                nxt[4 * j + 1] = fmaf(acc[k], v.y, nxt[4 * j + 1]);
                nxt[4 * j + 2] = fmaf(acc[k], v.z, nxt[4 * j + 2]);
                nxt[4 * j + 3] = fmaf(acc[k], v.w, nxt[4 * j + 3]);
              }
            }
          }
        }
#pragma unroll
        for (int j = 0; j < H0; ++j) acc[j] = sgnn_act(nxt[j] + bl[j], al, M.yeta);
      }
      sim[(size_t)(q0 + qi) * ld + g] = sgnn_act(acc[0], M.final_act, M.yeta);
    }
  }
}

// Pairwise node-node similarity histogram of LISTED pairs for graphs of any size (Web-shaped: 100..500 nodes,
// ~10^5 cells per pair).  CTA per pair; the second graph's node rows are staged in shared memory in tiles of up to
// 512 (d <= 16) or 256 rows; a work unit = two rows of the first graph (registers) x a chunk of 32 staged rows, processed as
// 2 x 2 blocks of cells (four independent chains); consecutive lanes take consecutive row pairs of the SAME chunk,
// so the staged rows are warp-uniform broadcasts.  Bins are counted in registers (BinCounters) and drained into
// CTA-wide shared counters.  Same cell arithmetic as the other two binning kernels.
constexpr int HIST_TILE_FLOATS = 8192;   // staged rows: 512 at ND = 16, 256 at ND = 32 (32 KB static)
constexpr int HIST_THREADS = 256;

template <int ND, int NB>
__global__ void __launch_bounds__(HIST_THREADS)
pair_histogram_tiled_kernel(const float* __restrict__ node_emb, int d, const int32_t* __restrict__ node_off,
                            const int32_t* __restrict__ g1, const int32_t* __restrict__ g2, int n_pairs, int bins,
                            float* __restrict__ hist, int32_t* __restrict__ counts) {
  constexpr int HIST_TJ = HIST_TILE_FLOATS / ND;
  __shared__ __align__(16) float s_h2[(HIST_TJ + 1) * ND];
  __shared__ int s_cnt[32];
  const int tid = threadIdx.x;
  const float fb = (float)bins;
  const int bmax = bins - 1;
  for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
    const int a0 = node_off[g1[p]], n1 = node_off[g1[p] + 1] - a0;
    const int b0 = node_off[g2[p]], n2 = node_off[g2[p] + 1] - b0;
    if (tid < 32) s_cnt[tid] = 0;
    BinCounters<NB> bc;
    bc.reset();
    auto to_cta = [&](int b, int v) { if (v) atomicAdd(&s_cnt[b], v); };
    const int nip = (n1 + 1) >> 1;
    for (int j0 = 0; j0 < n2; j0 += HIST_TJ) {
      const int tj = min(HIST_TJ, n2 - j0);
      __syncthreads();
      for (int idx = tid; idx < (tj + 1) * ND; idx += HIST_THREADS) {   // one zero row behind the tile (odd counts)
        const int r = idx / ND, c = idx - r * ND;
        s_h2[idx] = (r < tj && c < d) ? node_emb[(size_t)(b0 + j0 + r) * d + c] : 0.f;
      }
      __syncthreads();
      const int nchunks = (tj + 31) >> 5;
      for (int u = tid; u < nip * nchunks; u += HIST_THREADS) {
        const int jc = u / nip, ip = u - jc * nip;
        const int i = 2 * ip;
        const bool vi = i + 1 < n1;
        float ha[ND], hb[ND];
        if ((d & 3) == 0) {   // rows are 16-byte aligned: vector loads (row i + 1 of the last odd row is zeroed below)
          const float4* ra = reinterpret_cast<const float4*>(node_emb + (size_t)(a0 + i) * d);
          const float4* rb = reinterpret_cast<const float4*>(node_emb + (size_t)(a0 + (vi ? i + 1 : i)) * d);
#pragma unroll
          for (int c = 0; c < ND / 4; ++c) {
            const float4 u = 4 * c < d ? __ldg(ra + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 v = (vi && 4 * c < d) ? __ldg(rb + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            ha[4 * c] = u.x; ha[4 * c + 1] = u.y; ha[4 * c + 2] = u.z; ha[4 * c + 3] = u.w;
            hb[4 * c] = v.x; hb[4 * c + 1] = v.y; hb[4 * c + 2] = v.z; hb[4 * c + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int c = 0; c < ND; ++c) {
            ha[c] = c < d ? node_emb[(size_t)(a0 + i) * d + c] : 0.f;
            hb[c] = (vi && c < d) ? node_emb[(size_t)(a0 + i + 1) * d + c] : 0.f;
          }
        }
        const int jend = min(tj, (jc + 1) << 5);
        for (int j = jc << 5; j < jend; j += 2) {
          const float4* qa = reinterpret_cast<const float4*>(s_h2 + j * ND);
          const float4* qb = reinterpret_cast<const float4*>(s_h2 + (j + 1) * ND);
          float s00 = 0.f, s01 = 0.f, s10 = 0.f, s11 = 0.f;   // s[i-row][j-row]
#pragma unroll
          for (int c = 0; c < ND / 4; ++c) {
            const float4 x = qa[c], y = qb[c];
            s00 = fmaf(ha[4 * c], x.x, s00); s01 = fmaf(ha[4 * c], y.x, s01);
            s10 = fmaf(hb[4 * c], x.x, s10); s11 = fmaf(hb[4 * c], y.x, s11);
            s00 = fmaf(ha[4 * c + 1], x.y, s00); s01 = fmaf(ha[4 * c + 1], y.y, s01);
            s10 = fmaf(hb[4 * c + 1], x.y, s10); s11 = fmaf(hb[4 * c + 1], y.y, s11);
            s00 = fmaf(ha[4 * c + 2], x.z, s00); s01 = fmaf(ha[4 * c + 2], y.z, s01);
            s10 = fmaf(hb[4 * c + 2], x.z, s10); s11 = fmaf(hb[4 * c + 2], y.z, s11);
            s00 = fmaf(ha[4 * c + 3], x.w, s00); s01 = fmaf(ha[4 * c + 3], y.w, s01);
            s10 = fmaf(hb[4 * c + 3], x.w, s10); s11 = fmaf(hb[4 * c + 3], y.w, s11);
          }
          const bool vj = j + 1 < jend;
          bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s00) * fb)), true);
          bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s01) * fb)), vj);
          bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s10) * fb)), vi);
          bc.add(min(bmax, (int)floorf(sgnn_hist_sigmoid(s11) * fb)), vi && vj);
          bc.tick_drain(bins, to_cta);
        }
      }
    }
    bc.fold();
    bc.drain(bins, to_cta);
    __syncthreads();
    if (tid < bins) {
      const int cells = n1 * n2;
      const int tot = s_cnt[tid];
      if (counts) counts[(size_t)p * bins + tid] = tot;
      if (hist) hist[(size_t)p * bins + tid] = cells > 0 ? (float)tot * (1.f / (float)cells) : 0.f;
    }
    __syncthreads();
  }
}

// returns cudaErrorInvalidValue when d > 32 (the caller keeps the generic kernel)
cudaError_t sgnn_launch_pair_histogram_tiled(const float* node_emb, int d, const int32_t* node_off, const int32_t* g1,
                                             const int32_t* g2, int n_pairs, int bins, float* hist, int32_t* counts,
                                             int sm_count, cudaStream_t st) {
  if (d > 32 || bins > 32) return cudaErrorInvalidValue;
  const int blocks = std::min(n_pairs, sm_count * 8);
  if (d <= 16 && bins <= 16)
    pair_histogram_tiled_kernel<16, 16><<<blocks, HIST_THREADS, 0, st>>>(node_emb, d, node_off, g1, g2, n_pairs, bins, hist, counts);
  else if (d <= 16)
    pair_histogram_tiled_kernel<16, 32><<<blocks, HIST_THREADS, 0, st>>>(node_emb, d, node_off, g1, g2, n_pairs, bins, hist, counts);
  else if (bins <= 16)
    pair_histogram_tiled_kernel<32, 16><<<blocks, HIST_THREADS, 0, st>>>(node_emb, d, node_off, g1, g2, n_pairs, bins, hist, counts);
  else
    pair_histogram_tiled_kernel<32, 32><<<blocks, HIST_THREADS, 0, st>>>(node_emb, d, node_off, g1, g2, n_pairs, bins, hist, counts);
  return cudaGetLastError();
}

template <int DEP, int H0, int ND, int NB, bool PRE = false>
static cudaError_t launch_head(const ModelDev& M, const float* params, const float* Pq, const float* Cq,
                               const float* emb_db, const float* nodes_q, const int32_t* node_off_q, int max_nodes_q,
                               const float* nodes_db, const int32_t* node_off_db, int max_nodes_db, int m, int n,
                               float* sim, int64_t ld, int sm_count, int smem_optin, cudaStream_t st,
                               const uint8_t* cnt_in = nullptr) {
  HeadPlan P;
  head_plan_weights<H0>(M, P);
  const bool hist = M.hist_bins > 0;
  P.nmax_q = hist ? std::max(1, max_nodes_q) : 0;
  if (PRE) P.nmax_q = 0;   // node rows are not staged: the record keeps only the node count behind one padding row
  P.q_floats = M.K * DEP + round_up4(M.K) + (hist ? (P.nmax_q + 1) * ND + 4 : 0);
  const int rows = (hist && !PRE) ? std::max(1, max_nodes_db) * ND : 0;
  P.S = (hist && !PRE) ? (rows + 7) / 8 * 8 + 4 : 0;   // 4 * odd words: conflict-free 16-byte reads across a warp
  // (the 2 x 2 cell blocks tick the counters for the padded cells of odd node counts too: 15 x 15 graphs tick 256 times)
  const bool can_spill = hist && !PRE && (int64_t)((max_nodes_q + 1) & ~1) * ((max_nodes_db + 1) & ~1) > 240;   // BinCounters byte fields hold 255
  auto bytes_for = [&](int t) {
    return ((size_t)P.hw_floats + (size_t)HEAD_QC * P.q_floats + (size_t)t * P.S + ND +
            (can_spill ? (size_t)M.hist_bins * 2 * t : 0)) * sizeof(float);
  };
  // stripes per CTA: the choice that keeps the most warps resident (shared memory: 1 KB reserved per CTA;
  // registers: 4 CTAs of 128 threads or 2 of 256)
  int T = 0, best = 0;
  for (int t = 128; t >= 32; t >>= 1) {
    if (bytes_for(t) > (size_t)smem_optin) continue;
    const int by_smem = (int)(((size_t)smem_optin + 1024) / (bytes_for(t) + 1024));
    const int warps = std::min(by_smem, 512 / (2 * t)) * (2 * t / 32);
    if (warps > best) { best = warps; T = t; }
  }
  if (!T) return cudaErrorInvalidValue;
  P.T = T;
  const size_t smem = bytes_for(T);
  auto kern = allpairs_head_kernel<DEP, H0, ND, NB, PRE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int gx = (n + T - 1) / T;
  // ~24 CTAs per SM over the whole grid (6+ waves of the resident set, so the last wave costs a few per cent);
  // a CTA keeps its database stripes for all its queries -- re-staging them per CTA is ~20 bytes per pair from L2
  int gy = std::max(1, std::min((m + HEAD_QC - 1) / HEAD_QC, (sm_count * 24 + gx - 1) / gx));
  const int qpb = ((m + gy - 1) / gy + HEAD_QC - 1) / HEAD_QC * HEAD_QC;
  gy = (m + qpb - 1) / qpb;
  kern<<<dim3(gx, gy), 2 * T, smem, st>>>(M, P, params, Pq, Cq, emb_db, nodes_q, node_off_q, nodes_db, node_off_db,
                                          m, n, qpb, sim, ld, cnt_in);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_allpairs_head_counts(const ModelDev& M, int DEP, const float* params, const float* Pq,
                                             const float* Cq, const float* emb_db, const int32_t* node_off_q,
                                             const int32_t* node_off_db, const uint8_t* cnt_in, int m, int n,
                                             float* sim, int64_t ld, int sm_count, int smem_optin, cudaStream_t st) {
  if (M.n_head < 1 || M.hist_bins < 1 || M.hist_bins > 16) return cudaErrorInvalidValue;
  int hmax = 0;
  for (int l = 0; l < M.n_head; ++l) {
    hmax = std::max(hmax, M.Hd[l].dout);
    if (l > 0) hmax = std::max(hmax, M.Hd[l].din);
  }
  if (hmax > 32 || (DEP != 16 && DEP != 32)) return cudaErrorInvalidValue;
#define SGNN_HEADC_ARGS M, params, Pq, Cq, emb_db, nullptr, node_off_q, 0, nullptr, node_off_db, 0, m, n, sim, ld, \
                        sm_count, smem_optin, st, cnt_in
  if (DEP == 16) return hmax <= 16 ? launch_head<16, 16, 16, 16, true>(SGNN_HEADC_ARGS) : launch_head<16, 32, 16, 16, true>(SGNN_HEADC_ARGS);
  return hmax <= 16 ? launch_head<32, 16, 16, 16, true>(SGNN_HEADC_ARGS) : launch_head<32, 32, 16, 16, true>(SGNN_HEADC_ARGS);
#undef SGNN_HEADC_ARGS
}

cudaError_t sgnn_launch_allpairs_head(const ModelDev& M, int DEP, const float* params, const float* Pq, const float* Cq,
                                      const float* emb_db, const float* nodes_q, const int32_t* node_off_q,
                                      int max_nodes_q, const float* nodes_db, const int32_t* node_off_db,
                                      int max_nodes_db, int m, int n, float* sim, int64_t ld, int sm_count,
                                      int smem_optin, cudaStream_t st) {
  if (M.n_head < 1) return cudaErrorInvalidValue;
  int hmax = 0;
  for (int l = 0; l < M.n_head; ++l) {
    hmax = std::max(hmax, M.Hd[l].dout);
    if (l > 0) hmax = std::max(hmax, M.Hd[l].din);
  }
  const int nd = M.node_dim;
  if (hmax > 32 || nd > 32 || (DEP != 16 && DEP != 32)) return cudaErrorInvalidValue;
#define SGNN_HEAD_ARGS M, params, Pq, Cq, emb_db, nodes_q, node_off_q, max_nodes_q, nodes_db, node_off_db, \
                       max_nodes_db, m, n, sim, ld, sm_count, smem_optin, st
  const int key = (DEP == 32 ? 4 : 0) | (hmax > 16 ? 2 : 0) | (nd > 16 ? 1 : 0);
  if (M.hist_bins <= 16) {
    switch (key) {
      case 0: return launch_head<16, 16, 16, 16>(SGNN_HEAD_ARGS);
      case 1: return launch_head<16, 16, 32, 16>(SGNN_HEAD_ARGS);
      case 2: return launch_head<16, 32, 16, 16>(SGNN_HEAD_ARGS);
      case 3: return launch_head<16, 32, 32, 16>(SGNN_HEAD_ARGS);
      case 4: return launch_head<32, 16, 16, 16>(SGNN_HEAD_ARGS);
      case 5: return launch_head<32, 16, 32, 16>(SGNN_HEAD_ARGS);
      case 6: return launch_head<32, 32, 16, 16>(SGNN_HEAD_ARGS);
      default: return launch_head<32, 32, 32, 16>(SGNN_HEAD_ARGS);
    }
  }
  switch (key) {
    case 0: return launch_head<16, 16, 16, 32>(SGNN_HEAD_ARGS);
    case 1: return launch_head<16, 16, 32, 32>(SGNN_HEAD_ARGS);
    case 2: return launch_head<16, 32, 16, 32>(SGNN_HEAD_ARGS);
    case 3: return launch_head<16, 32, 32, 32>(SGNN_HEAD_ARGS);
    case 4: return launch_head<32, 16, 16, 32>(SGNN_HEAD_ARGS);
    case 5: return launch_head<32, 16, 32, 32>(SGNN_HEAD_ARGS);
    case 6: return launch_head<32, 32, 16, 32>(SGNN_HEAD_ARGS);
    default: return launch_head<32, 32, 32, 32>(SGNN_HEAD_ARGS);
  }
#undef SGNN_HEAD_ARGS
}
