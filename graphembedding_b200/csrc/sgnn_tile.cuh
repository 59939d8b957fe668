// Node-per-thread embedding kernel for small graphs (N <= 32).
//
// A warp owns a TILE of up to 32 consecutive node rows = `gpw` whole graphs of the block-diagonal
// batch (gpw = floor(32 / max_nodes)); lane l is node l of the tile.  Every thread carries its
// node's output row in registers (JT accumulators at a time):
//   layer 0 (one-hot X, graphs.py:55-62):  out_i = act(sum_e A_hat[i,e] * W0[label(e), :] + b)   -- a row gather of W0
//   GCN l>0 (layers.py:91-118):            T_i = H_i . W  (own row x broadcast weight rows), then
//                                          out_i = act(sum_e A_hat[i,e] * T_e + b) through shared memory
//   Dense (layers.py:192-205):             out_i = act(H_i . W + b)
// so the feature x weight products run at full lane utilisation with no per-row predication, and
// the only cross-lane traffic is the neighbour gather (warp-synchronous shared memory).  Weights
// are re-staged once per CTA in a padded, 16-byte aligned layout (EmbPlan) so that weight rows are
// read as broadcast float4.  Pooling (Padding / Average / Attention, layers.py:136-160, 223-227)
// is done per graph by the same warp.
#pragma once
#include <cuda_bf16.h>
#include "sgnn_common.cuh"
#include "sgnn_tc_geom.h"

// Build with -DSGNN_TILE_TRACE to stamp the phases of warp 0's first tile of every CTA (tools/tile_probe.py)
#ifdef SGNN_TILE_TRACE
__device__ unsigned long long g_tile_trace[1024 * 8];
#define TILE_STAMP(i)                                                                   \
  do {                                                                                  \
    if (threadIdx.x == 0 && blockIdx.x < 1024) {                                        \
      unsigned long long t_;                                                            \
      if ((i) == 0 || (i) == 7) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));  \
      else t_ = clock64();                                                              \
      g_tile_trace[blockIdx.x * 8 + (i)] = t_;                                          \
    }                                                                                   \
  } while (0)
#else
#define TILE_STAMP(i)
#endif

constexpr int TILE_CSR_CAP = 160;   // staged nnz per tile (falls back to global reads beyond)
constexpr int TILE_MAX_WARPS = 12;  // warps (= tiles in flight) per CTA; the launcher picks <= this

struct EmbPlan {
  int w_s[SGNN_MAX_LAYERS];    // offset (floats) of layer l's padded weight matrix in shared memory
  int b_s[SGNN_MAX_LAYERS];    // offset of the padded bias (zeros when the layer has none)
  int ldw[SGNN_MAX_LAYERS];    // padded row stride of the weight matrix
  int att_s;                   // attention weights [d][d] (row stride d), -1 if unused
  int w_floats;                // total staged floats (multiple of 4)
  int ldh;                     // row stride of activation buffer A (holds every layer's output; ldh % 8 == 4)
  int ldb;                     // row stride of buffer B (outputs of layers l >= 1 only; ldb % 8 == 4)
  int gpw;                     // graphs per warp tile
  int warp_floats;             // per-warp shared memory (floats)
};

inline EmbPlan make_emb_plan(const ModelDev& M, int max_nodes) {
  EmbPlan P;
  int o = 0, dmax = 4, dmax1 = 4;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    int ldw = round_up(L.dout, 4);
    if (l == 0 && ldw % 8 == 0) ldw += 4;   // gathered rows: spread consecutive labels over the banks
    const int rows = l == 0 ? L.din : round_up(L.din, 4);
    P.ldw[l] = ldw; P.w_s[l] = o; o += rows * ldw;
    P.b_s[l] = o; o += round_up(L.dout, 4);
    dmax = dmax > L.dout ? dmax : L.dout;
    if (l > 0) dmax1 = dmax1 > L.dout ? dmax1 : L.dout;
  }
  P.att_s = -1;
  if (M.pool_kind == SGNN_LAYER_ATTENTION) {
    P.att_s = o; o += round_up(M.node_dim * M.node_dim, 4);
    dmax1 = dmax1 > M.node_dim ? dmax1 : M.node_dim;   // pooling scratch lives in the idle buffer
  }
  P.w_floats = round_up(o, 4);
  P.ldh = round_up(dmax, 4);
  if (P.ldh % 8 == 0) P.ldh += 4;
  P.ldb = round_up(dmax1, 4);
  if (P.ldb % 8 == 0) P.ldb += 4;
  int gpw = max_nodes > 0 ? 32 / max_nodes : 1;
  P.gpw = gpw < 1 ? 1 : (gpw > 16 ? 16 : gpw);
  P.warp_floats = 32 * P.ldh + 32 * P.ldb + 2 * TILE_CSR_CAP + 32 + 32;
  return P;
}

__device__ __forceinline__ void tile_cp_async4(float* dst_smem, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
               : "memory");
}
__device__ __forceinline__ void tile_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tile_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Weights of the node layers (+ attention) into the padded shared layout, asynchronously: the
// caller overlaps the first tile's CSR loads with the copy and waits (cp.async wait + CTA barrier)
// before the first layer.  Warp-per-row, lane-per-column: no index arithmetic beyond adds.
__device__ __forceinline__ void tile_stage_weights_async(const ModelDev& M, const EmbPlan& P,
                                                         const float* __restrict__ params, float* s_w) {
  float4* z = reinterpret_cast<float4*>(s_w);
  for (int i = threadIdx.x; i < P.w_floats / 4; i += blockDim.x) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();   // the zero fill is ordered before the asynchronous element copies
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    const float* src = params + L.w_off;
    float* dst = s_w + P.w_s[l];
    for (int k = warp; k < L.din; k += nw)
      for (int j = lane; j < L.dout; j += 32) tile_cp_async4(dst + k * P.ldw[l] + j, src + k * L.dout + j);
    if (L.bias && warp == (l % nw))
      for (int j = lane; j < L.dout; j += 32) tile_cp_async4(s_w + P.b_s[l] + j, params + L.b_off + j);
  }
  if (P.att_s >= 0)
    for (int i = threadIdx.x; i < M.node_dim * M.node_dim; i += blockDim.x)
      tile_cp_async4(s_w + P.att_s + i, params + M.pool_w_off + i);
  tile_cp_async_commit();
}

struct TileCtx {
  int n_rows;          // node rows in this tile (<= 32)
  bool on;             // this lane owns a node
  int g_off;           // lane of the first node of my graph
  int e0, e1;          // my row's entries, relative to the tile's first entry
  const int* col;      // staged (shared) or global, indexed relative to the tile's first entry
  const float* val;
  const int* lab;      // [32] labels of the tile's nodes (shared)
};

// sigmoid / tanh / sim_kernel are rare on node layers: keep their (large) code out of line, once
__device__ __noinline__ float tile_act_slow(float x, int act) { return sgnn_act(x, act, 0.f); }

template <int JT>
__device__ __forceinline__ void tile_store_row(float* dst, float (&acc)[JT], const float* bias, int j0, int dout,
                                               int act) {
  if (bias) {
#pragma unroll
    for (int jj = 0; jj < JT; jj += 4) {
      if (j0 + jj < dout) {   // warp-uniform
        const float4 b4 = *reinterpret_cast<const float4*>(bias + j0 + jj);   // padded with zeros
        acc[jj] += b4.x; acc[jj + 1] += b4.y; acc[jj + 2] += b4.z; acc[jj + 3] += b4.w;
      }
    }
  }
  if (act == SGNN_ACT_RELU) {
#pragma unroll
    for (int jj = 0; jj < JT; ++jj) acc[jj] = fmaxf(acc[jj], 0.f);
  } else if (act != SGNN_ACT_IDENTITY) {
#pragma unroll
    for (int jj = 0; jj < JT; ++jj)
      if (j0 + jj < dout) acc[jj] = tile_act_slow(acc[jj], act);
  }
#pragma unroll
  for (int jj = 0; jj < JT; jj += 4) {
    if (j0 + jj < dout) {
      float4 o;
      o.x = acc[jj];
      o.y = j0 + jj + 1 < dout ? acc[jj + 1] : 0.f;   // padded columns stay exactly zero
      o.z = j0 + jj + 2 < dout ? acc[jj + 2] : 0.f;
      o.w = j0 + jj + 3 < dout ? acc[jj + 3] : 0.f;
      *reinterpret_cast<float4*>(dst + j0 + jj) = o;
    }
  }
}

// acc[jj] += sum_e val[e] * Src[row(e)][j0 + jj]; rows are weight rows (by label) or neighbour rows
template <int JT, bool BY_LABEL>
__device__ __forceinline__ void tile_gather(const TileCtx& T, const float* Src, int ld, int j0, int dout,
                                            float (&acc)[JT]) {
  for (int e = T.e0; e < T.e1; ++e) {
    const int nb = T.g_off + T.col[e];
    const float v = T.val[e];
    const float* row = Src + (BY_LABEL ? T.lab[nb] : nb) * ld + j0;
#pragma unroll
    for (int jj = 0; jj < JT; jj += 4) {
      if (j0 + jj < dout) {
        const float4 w = *reinterpret_cast<const float4*>(row + jj);
        acc[jj + 0] = fmaf(v, w.x, acc[jj + 0]);
        acc[jj + 1] = fmaf(v, w.y, acc[jj + 1]);
        acc[jj + 2] = fmaf(v, w.z, acc[jj + 2]);
        acc[jj + 3] = fmaf(v, w.w, acc[jj + 3]);
      }
    }
  }
}

// acc[jj] = sum_k hrow[k] * W[k][j0 + jj]   (own row x broadcast weight rows)
template <int JT>
__device__ __forceinline__ void tile_row_gemm(const float* hrow, int din, const float* W, int ldw, int j0, int dout,
                                              float (&acc)[JT]) {
  for (int k = 0; k < din; k += 4) {
    const float4 h4 = *reinterpret_cast<const float4*>(hrow + k);
    const float hk[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float* wr = W + (k + kk) * ldw + j0;
#pragma unroll
      for (int jj = 0; jj < JT; jj += 4) {
        if (j0 + jj < dout) {
          const float4 w = *reinterpret_cast<const float4*>(wr + jj);
          acc[jj + 0] = fmaf(hk[kk], w.x, acc[jj + 0]);
          acc[jj + 1] = fmaf(hk[kk], w.y, acc[jj + 1]);
          acc[jj + 2] = fmaf(hk[kk], w.z, acc[jj + 2]);
          acc[jj + 3] = fmaf(hk[kk], w.w, acc[jj + 3]);
        }
      }
    }
  }
}

// One node layer of one tile with JT accumulators per pass.  first: gather of W rows by label into
// cur.  GCN: cur -> (gemm) oth -> (aggregate) cur.  Dense: cur -> oth (caller swaps).
template <int JT>
__device__ __forceinline__ void tile_layer(const LayerDev& L, bool first, const TileCtx& T, const float* W, int ldw,
                                        const float* b, float* cur, int ldc, float* oth, int ldo, int lane) {
  const bool gcn = L.kind == SGNN_LAYER_GCN;
  if (first) {
    if (T.on)
      for (int j0 = 0; j0 < L.dout; j0 += JT) {
        float acc[JT];
#pragma unroll
        for (int jj = 0; jj < JT; ++jj) acc[jj] = 0.f;
        tile_gather<JT, true>(T, W, ldw, j0, L.dout, acc);
        tile_store_row<JT>(cur + lane * ldc, acc, b, j0, L.dout, L.act);
      }
    __syncwarp();
    return;
  }
  if (T.on)
    for (int j0 = 0; j0 < L.dout; j0 += JT) {
      float acc[JT];
#pragma unroll
      for (int jj = 0; jj < JT; ++jj) acc[jj] = 0.f;
      tile_row_gemm<JT>(cur + lane * ldc, round_up4(L.din), W, ldw, j0, L.dout, acc);
      tile_store_row<JT>(oth + lane * ldo, acc, gcn ? nullptr : b, j0, L.dout, gcn ? (int)SGNN_ACT_IDENTITY : L.act);
    }
  __syncwarp();
  if (!gcn) return;
  if (T.on)
    for (int j0 = 0; j0 < L.dout; j0 += JT) {
      float acc[JT];
#pragma unroll
      for (int jj = 0; jj < JT; ++jj) acc[jj] = 0.f;
      tile_gather<JT, false>(T, oth, ldo, j0, L.dout, acc);
      tile_store_row<JT>(cur + lane * ldc, acc, b, j0, L.dout, L.act);
    }
  __syncwarp();
}

// all node layers of one tile; returns the buffer holding the last layer's output rows (stride *ld_out).
// Buffer A (stride ldh) can hold any layer's output; buffer B (stride ldb) only those of layers >= 1.
__device__ __forceinline__ float* tile_node_layers(const ModelDev& M, const EmbPlan& P, const TileCtx& T,
                                                   const float* s_w, float* bufA, float* bufB, int lane,
                                                   int* ld_out) {
  float* cur = bufA;
  float* oth = bufB;
  int ldc = P.ldh, ldo = P.ldb;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    const float* W = s_w + P.w_s[l];
    const float* b = L.bias ? s_w + P.b_s[l] : nullptr;
    if (L.dout > 4) tile_layer<16>(L, l == 0, T, W, P.ldw[l], b, cur, ldc, oth, ldo, lane);
    else tile_layer<4>(L, l == 0, T, W, P.ldw[l], b, cur, ldc, oth, ldo, lane);
    if (l > 0 && L.kind != SGNN_LAYER_GCN) {
      float* sw = cur; cur = oth; oth = sw;
      const int t = ldc; ldc = ldo; ldo = t;
    }
  }
  *ld_out = ldc;
  return cur;
}

// Optional fused output: the database-side operand of the tensor-core all-pairs scorer (what tc_prep_kernel's
// database branch builds from emb: bf16 hi / lo split of [x_g, 1, 0...] in the UMMA K-major core-matrix order,
// sgnn_allpairs_tc.cuh) written by the warp that just pooled the graphs -- one kernel and one launch gap less on
// the critical path of the ranking step.
struct TileOperandOut {
  __nv_bfloat16* A;   // NULL = off
  int KCH, parts;     // 16-byte K chunks per part; 1 = bf16, 2 = bf16x3 (hi + lo)
  // QUERY side (what tc_prep_kernel's query branch / tc_coef_kernel + tc_prep_q_kernel build): every CTA computes
  // the (D + 1) x C coefficient matrix of the operand's columns into shared memory while its weights are in flight,
  // and the warp that pooled a tile's graphs writes their operand rows -- the whole query branch of the ranking
  // step is then ONE kernel (each extra graph node on that branch costs ~3 us of launch latency).
  __nv_bfloat16* B;   // NULL = off
  int K, KC, QH;      // NTN feature maps, accumulator columns per query, queries per 128-row B tile
  int wm_off;         // float offset of the coefficient matrix inside dynamic shared memory
};

__global__ void __launch_bounds__(TILE_MAX_WARPS * 32, 2)
embed_tile_kernel(const __grid_constant__ ModelDev M, const __grid_constant__ EmbPlan P, const BatchDev B,
                  const float* __restrict__ params, float* __restrict__ emb, float* __restrict__ node_out,
                  const TileOperandOut ao) {
  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;
  TILE_STAMP(0);
  TILE_STAMP(1);
  tile_stage_weights_async(M, P, params, s_w);
  TILE_STAMP(2);
  if (ao.B) {   // coefficient matrix of the query-side operand, pass 1: the plain columns (one parameter each);
    const int D1 = M.emb_dim + 1, C = ao.KCH * 8 * ao.KC;   // the relu sum columns follow the first CTA barrier
    float* Wm = smem + ao.wm_off;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < D1 * C; idx += blockDim.x) {
      const int a = idx / C, c = idx - a * C;
      const int kk = c / ao.KC, k = c - kk * ao.KC;
      Wm[idx] = k < ao.K ? tc_prep_coef(M, params, M.emb_dim, ao.K, a, kk, k) : 0.f;
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float* my = smem + P.w_floats + warp * P.warp_floats;
  float* bufA = my;
  float* bufB = bufA + 32 * P.ldh;
  int* s_col = reinterpret_cast<int*>(bufB + 32 * P.ldb);
  float* s_val = reinterpret_cast<float*>(s_col + TILE_CSR_CAP);
  int* s_lab = reinterpret_cast<int*>(s_val + TILE_CSR_CAP);
  float* s_att = reinterpret_cast<float*>(s_lab + 32);
  const int gpw = P.gpw;
  const int n_tiles = (B.n_graphs + gpw - 1) / gpw;
  const int E = M.emb_dim, d = M.node_dim;
  bool w_ready = false;
  for (int t = blockIdx.x * nw + warp; ; t += gridDim.x * nw) {
    const bool live = t < n_tiles;     // warps without a tile still take part in the one CTA barrier below
    TileCtx T;
    int my_no = 0, base = 0, g0 = 0, g1 = 0, gi = 0;
    if (live) {
      g0 = t * gpw; g1 = min(B.n_graphs, g0 + gpw);
      // lanes 0..gpw hold the node offsets of the tile's graphs
      my_no = B.node_off[min(g0 + lane, g1)];
      base = __shfl_sync(0xffffffffu, my_no, 0);
      const int n_rows = __shfl_sync(0xffffffffu, my_no, g1 - g0) - base;
      T.n_rows = n_rows; T.on = lane < n_rows;
      T.g_off = 0;
      for (int k = 1; k < g1 - g0; ++k) {
        const int off = __shfl_sync(0xffffffffu, my_no, k) - base;
        if (lane >= off) { gi = k; T.g_off = off; }
      }
      const int rp_lo = B.rowptr[base + min(lane, n_rows)];
      const int rp_hi = B.rowptr[base + min(lane + 1, n_rows)];
      const int lab = T.on ? B.labels[base + lane] : 0;
      const int tile_e0 = __shfl_sync(0xffffffffu, rp_lo, 0);
      const int tile_nnz = __shfl_sync(0xffffffffu, rp_hi, 31) - tile_e0;   // lane 31 read rowptr[base + n_rows]
      T.e0 = rp_lo - tile_e0; T.e1 = T.on ? rp_hi - tile_e0 : T.e0;
      s_lab[lane] = lab;
      if (tile_nnz <= TILE_CSR_CAP) {
        for (int e = lane; e < tile_nnz; e += 32) { s_col[e] = B.col[tile_e0 + e]; s_val[e] = B.val[tile_e0 + e]; }
        T.col = s_col; T.val = s_val;
      } else {
        T.col = B.col + tile_e0; T.val = B.val + tile_e0;
      }
      T.lab = s_lab;
      __syncwarp();
    }
    if (!w_ready) {   // the first tile's CSR loads overlapped the weight copy
      tile_cp_async_wait_all();
      __syncthreads();
      w_ready = true;
      if (ao.B && ao.KC > ao.K) {   // pass 2: sum columns sum_k w_k Wm[a][(kk, k)], same order as tc_prep_kernel
        const int D1 = M.emb_dim + 1, C = ao.KCH * 8 * ao.KC, groups = ao.KCH * 8;
        float* Wm = smem + ao.wm_off;
        const bool ref = M.pair_kind == SGNN_LAYER_DOT || M.ntn_mode == SGNN_NTN_REFERENCE;
        for (int it = threadIdx.x; it < D1 * groups; it += blockDim.x) {
          const int a = it / groups, col = (it - a * groups) * ao.KC;
          float v = 0.f;
          for (int kq = 0; kq < ao.K; ++kq) v = fmaf(ref ? 1.f : params[M.u_off + kq], Wm[a * C + col + kq], v);
          Wm[a * C + col + ao.K] = v;
        }
        __syncthreads();
      }
    }
    if (!live) break;
    if (t == blockIdx.x * nw) TILE_STAMP(3);
    int ldh;
    float* H = tile_node_layers(M, P, T, s_w, bufA, bufB, lane, &ldh);
    if (t == blockIdx.x * nw) TILE_STAMP(4);
    float* scratch = H == bufA ? bufB : bufA;
    if (node_out) {   // last node-level rows of the tile (the histogram layer's input), [total_nodes][node_dim]
      __syncwarp();
      for (int idx = lane; idx < T.n_rows * d; idx += 32) {
        const int r = idx / d, c = idx - r * d;
        node_out[(size_t)(base + r) * d + c] = H[r * ldh + c];
      }
    }
    // ---- pooling, per graph of the tile ----
    if (M.pool_kind == SGNN_LAYER_PADDING) {
      for (int k = 0; k < g1 - g0; ++k) {
        const int off = __shfl_sync(0xffffffffu, my_no, k) - base;
        const int n = __shfl_sync(0xffffffffu, my_no, k + 1) - base - off;
        float* out = emb + (size_t)(g0 + k) * E;
        for (int idx = lane; idx < E; idx += 32) {
          const int p = idx / d, c = idx - p * d;
          out[idx] = p < n ? H[(off + p) * ldh + c] : M.pool_fparam;
        }
      }
    } else {
      // Average / Attention: lane idx owns output (graph k, column j); graph extents come by shuffle
      const int cnt = (g1 - g0) * d;
      float* s_mean = scratch;                 // [gpw][d]   (Attention only)
      float* s_c = scratch + cnt;              // [gpw][d]
      const bool att = M.pool_kind == SGNN_LAYER_ATTENTION;
      for (int i0 = 0; i0 < cnt; i0 += 32) {
        const int idx = i0 + lane;
        const bool ok = idx < cnt;
        const int k = ok ? idx / d : 0, j = idx - k * d;
        const int off = __shfl_sync(0xffffffffu, my_no, k) - base;
        const int n = __shfl_sync(0xffffffffu, my_no, k + 1) - base - off;
        if (ok) {
          float s = 0.f;
          for (int i = 0; i < n; ++i) s += H[(off + i) * ldh + j];
          s /= (float)n;
          if (att) s_mean[idx] = s; else emb[(size_t)(g0 + k) * E + j] = s;
        }
      }
      if (att) {   // layers.py:154-160
        const float* Wa = s_w + P.att_s;
        __syncwarp();
        for (int idx = lane; idx < cnt; idx += 32) {
          const int k = idx / d, j = idx - k * d;
          float s = 0.f;
          for (int kk = 0; kk < d; ++kk) s = fmaf(s_mean[k * d + kk], Wa[kk * d + j], s);
          s_c[idx] = tanhf(s);
        }
        __syncwarp();
        if (T.on) {
          float s = 0.f;
          for (int kk = 0; kk < d; ++kk) s = fmaf(H[lane * ldh + kk], s_c[gi * d + kk], s);
          s_att[lane] = 1.f / (1.f + expf(-s));
        }
        __syncwarp();
        for (int i0 = 0; i0 < cnt; i0 += 32) {
          const int idx = i0 + lane;
          const bool ok = idx < cnt;
          const int k = ok ? idx / d : 0, j = idx - k * d;
          const int off = __shfl_sync(0xffffffffu, my_no, k) - base;
          const int n = __shfl_sync(0xffffffffu, my_no, k + 1) - base - off;
          if (ok) {
            float s = 0.f;
            for (int i = 0; i < n; ++i) s = fmaf(s_att[off + i], H[(off + i) * ldh + j], s);
            emb[(size_t)(g0 + k) * E + j] = s;
          }
        }
      }
    }
    __syncwarp();
    if (ao.A) {   // the rows this warp just wrote (ordered by the __syncwarp above), re-read through L2
      for (int item = lane; item < (g1 - g0) * ao.KCH; item += 32) {
        const int gi = item / ao.KCH, kc = item - gi * ao.KCH;
        const int g = g0 + gi;
        __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int b = kc * 8 + j;
          const float v = b < E ? __ldcg(emb + (size_t)g * E + b) : (b == E ? 1.f : 0.f);
          hi[j] = __float2bfloat16_rn(v);
          lo[j] = __float2bfloat16_rn(v - __bfloat162float(hi[j]));
        }
        const size_t base = ((size_t)((g >> 7) * ao.parts) * ao.KCH + kc) * 1024 + (size_t)(g & 127) * 8;
        *reinterpret_cast<uint4*>(ao.A + base) = *reinterpret_cast<const uint4*>(hi);
        if (ao.parts == 2)
          *reinterpret_cast<uint4*>(ao.A + base + (size_t)ao.KCH * 1024) = *reinterpret_cast<const uint4*>(lo);
      }
    }
    if (ao.B) {   // operand rows of the tile's queries: lane = column, same summation order as tc_prep_q_kernel
      const int C = ao.KCH * 8 * ao.KC, ng = g1 - g0;
      const float* Wm = smem + ao.wm_off;
      float* xs = scratch;   // [ng][E]: the rows this warp just wrote, one L2 round trip
      for (int i = lane; i < ng * E; i += 32) xs[i] = __ldcg(emb + (size_t)g0 * E + i);
      __syncwarp();
      const size_t lo_off = (size_t)ao.KCH * TC_NPAD * 8;
      for (int c = lane; c < C; c += 32) {
        const int kk = c / ao.KC, k = c - kk * ao.KC;
        const float cst = Wm[E * C + c];
        for (int q0 = 0; q0 < ng; q0 += 4) {
          float acc[4] = {cst, cst, cst, cst};
          for (int a = 0; a < E; ++a) {
            const float w = Wm[a * C + c];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j] = fmaf(q0 + j < ng ? xs[(q0 + j) * E + a] : 0.f, w, acc[j]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (q0 + j < ng) {
              const int q = g0 + q0 + j;
              const int qb = q / ao.QH, row = (q - qb * ao.QH) * ao.KC + k;
              __nv_bfloat16* dst = ao.B + ((size_t)(qb * ao.parts) * ao.KCH + (kk >> 3)) * (TC_NPAD * 8) + (size_t)row * 8 + (kk & 7);
              const __nv_bfloat16 hi = __float2bfloat16_rn(acc[j]);
              *dst = hi;
              if (ao.parts == 2) dst[lo_off] = __float2bfloat16_rn(acc[j] - __bfloat162float(hi));
            }
          }
        }
      }
      __syncwarp();
    }
    if (t == blockIdx.x * nw) TILE_STAMP(5);
  }
  TILE_STAMP(6);
  TILE_STAMP(7);
}
