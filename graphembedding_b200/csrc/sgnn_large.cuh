// Large-graph embedding path (N > 32 nodes: AIDS10k-shaped beyond 32, Web-shaped 100..500+): one CTA
// per graph, node rows striped over the CTA's threads, the same node-per-thread row arithmetic as the
// tile kernel (sgnn_tile.cuh: own row x broadcast weight rows, neighbour gather through CSR).  Row
// buffers live in a global workspace (2 x total_nodes x ldh floats; L2-resident for these sizes) because
// a 500-node graph's activations do not fit next to the weights in shared memory for wide layers; the
// gather of neighbour rows is the HBM/L2-bound stage of this path (CSR SpMM, 16-byte loads).
// Reference math: layers.py:91-118 (GCN), :192-205 (Dense), :136-140 (Average), :154-160 (Attention).
#pragma once
#include "sgnn_tile.cuh"

constexpr int LARGE_THREADS = 256;

template <int JT>
__device__ __forceinline__ void large_layer_rows(const LayerDev& L, int phase, const TileCtx& T, const float* W, int ldw,
                                                 const float* b, const float* src, float* dst, int ldh, int row) {
  // phase 0: first layer (gather W rows by label) -> dst ; phase 1: gemm src row -> dst row
  // phase 2: aggregate neighbour rows of src -> dst (+ bias, act)
  const bool gcn = L.kind == SGNN_LAYER_GCN;
  for (int j0 = 0; j0 < L.dout; j0 += JT) {
    float acc[JT];
#pragma unroll
    for (int jj = 0; jj < JT; ++jj) acc[jj] = 0.f;
    if (phase == 0) {
      tile_gather<JT, true>(T, W, ldw, j0, L.dout, acc);
      tile_store_row<JT>(dst + (size_t)row * ldh, acc, b, j0, L.dout, L.act);
    } else if (phase == 1) {
      tile_row_gemm<JT>(src + (size_t)row * ldh, round_up4(L.din), W, ldw, j0, L.dout, acc);
      tile_store_row<JT>(dst + (size_t)row * ldh, acc, gcn ? nullptr : b, j0, L.dout,
                         gcn ? (int)SGNN_ACT_IDENTITY : L.act);
    } else {
      tile_gather<JT, false>(T, src, ldh, j0, L.dout, acc);
      tile_store_row<JT>(dst + (size_t)row * ldh, acc, b, j0, L.dout, L.act);
    }
  }
}

__device__ __noinline__ void large_layer_phase(const LayerDev& L, int phase, const BatchDev& B, int base, int n, int e0,
                                               const float* W, int ldw, const float* b, const float* src, float* dst,
                                               int ldh) {
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    TileCtx T;
    T.n_rows = n; T.on = true; T.g_off = 0;
    T.e0 = B.rowptr[base + r] - e0; T.e1 = B.rowptr[base + r + 1] - e0;
    T.col = B.col + e0; T.val = B.val + e0; T.lab = B.labels + base;
    if (L.dout > 16) large_layer_rows<32>(L, phase, T, W, ldw, b, src, dst, ldh, r);
    else if (L.dout > 4) large_layer_rows<16>(L, phase, T, W, ldw, b, src, dst, ldh, r);
    else large_layer_rows<4>(L, phase, T, W, ldw, b, src, dst, ldh, r);
  }
  __syncthreads();
}

// graphs listed in `which` (indices into the batch; nullptr = all graphs)
__global__ void __launch_bounds__(LARGE_THREADS)
embed_large_kernel(const __grid_constant__ ModelDev M, const __grid_constant__ EmbPlan P, const BatchDev B,
                   const int32_t* __restrict__ which, int n_which, const float* __restrict__ params,
                   float* __restrict__ bufA, float* __restrict__ bufB, float* __restrict__ emb,
                   float* __restrict__ node_out) {
  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;
  float* s_red = smem + P.w_floats;          // [2 * d + 8 warps * d] pooling scratch, also raw staging
  tile_stage_weights_async(M, P, params, s_w);
  tile_cp_async_wait_all();
  __syncthreads();
  const int ldh = P.ldh, d = M.node_dim, E = M.emb_dim;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int gi = blockIdx.x; gi < n_which; gi += gridDim.x) {
    const int g = which ? which[gi] : gi;
    const int base = B.node_off[g], n = B.node_off[g + 1] - base;
    const int e0 = B.rowptr[base];
    float* cur = bufA + (size_t)base * ldh;   // rows are addressed graph-locally
    float* oth = bufB + (size_t)base * ldh;
    for (int l = 0; l < M.n_node; ++l) {
      const LayerDev& L = M.L[l];
      const float* W = s_w + P.w_s[l];
      const float* b = L.bias ? s_w + P.b_s[l] : nullptr;
      if (l == 0) {
        large_layer_phase(L, 0, B, base, n, e0, W, P.ldw[l], b, nullptr, cur, ldh);
      } else if (L.kind == SGNN_LAYER_GCN) {
        large_layer_phase(L, 1, B, base, n, e0, W, P.ldw[l], b, cur, oth, ldh);
        large_layer_phase(L, 2, B, base, n, e0, W, P.ldw[l], b, oth, cur, ldh);
      } else {
        large_layer_phase(L, 1, B, base, n, e0, W, P.ldw[l], b, cur, oth, ldh);
        float* sw = cur; cur = oth; oth = sw;
      }
    }
    const float* H = cur;
    if (node_out) {   // node-level output of the last node layer, [total_nodes][node_dim] (histogram feature)
      for (int idx = threadIdx.x; idx < n * d; idx += blockDim.x) {
        const int i = idx / d, c = idx - i * d;
        node_out[(size_t)(base + i) * d + c] = H[(size_t)i * ldh + c];
      }
    }
    if (!emb) { __syncthreads(); continue; }
    float* out = emb + (size_t)g * E;
    if (M.pool_kind == SGNN_LAYER_PADDING) {
      for (int idx = threadIdx.x; idx < E; idx += blockDim.x) {
        const int p = idx / d, c = idx - p * d;
        out[idx] = p < n ? H[(size_t)p * ldh + c] : M.pool_fparam;
      }
    } else {
      // column sums by warp-striped rows, combined in a FIXED order (deterministic)
      float* s_mean = s_red; float* s_c = s_red + d; float* s_part = s_red + 2 * d;   // [nw][d]
      const bool att = M.pool_kind == SGNN_LAYER_ATTENTION;
      for (int pass = 0; pass < (att ? 2 : 1); ++pass) {
        for (int j = lane; j < d; j += 32) {
          float s = 0.f;
          for (int i = warp; i < n; i += nw) {
            float wgt = 1.f;
            if (pass == 1) {   // attention weight of node i: sigmoid(H_i . c)
              float t = 0.f;
              for (int k = 0; k < d; ++k) t = fmaf(H[(size_t)i * ldh + k], s_c[k], t);
              wgt = 1.f / (1.f + expf(-t));
            }
            s = fmaf(wgt, H[(size_t)i * ldh + j], s);
          }
          s_part[warp * d + j] = s;
        }
        __syncthreads();
        for (int j = threadIdx.x; j < d; j += blockDim.x) {
          float s = 0.f;
          for (int w = 0; w < nw; ++w) s += s_part[w * d + j];
          if (!att) out[j] = s / (float)n;
          else if (pass == 0) s_mean[j] = s / (float)n;
          else out[j] = s;
        }
        __syncthreads();
        if (att && pass == 0) {
          const float* Wa = s_w + P.att_s;
          for (int j = threadIdx.x; j < d; j += blockDim.x) {
            float s = 0.f;
            for (int k = 0; k < d; ++k) s = fmaf(s_mean[k], Wa[k * d + j], s);
            s_c[j] = tanhf(s);
          }
          __syncthreads();
        }
      }
    }
    __syncthreads();
  }
}

// Pairwise node-node similarity histogram (SimGNN's histogram feature; ABSENT from the reference snapshot,
// SURVEY F1 -- build-defined): for a pair (G1, G2) with node embeddings H1 [N1 x d], H2 [N2 x d],
//   s_ij = sigmoid(H1_i . H2_j),  bin_ij = min(bins - 1, floor(s_ij * bins))   (fixed width over [0, 1],
//   tf.histogram_fixed_width semantics),  hist[b] = #{bin_ij == b} / (N1 N2).
// One CTA per pair; a warp scores 32 (i, j) cells at a time and counts them per bin with ballot + popc
// (lane b accumulates bin b), so there are no shared-memory atomics.  bins <= 32.
__global__ void __launch_bounds__(128)
pair_histogram_kernel(const float* __restrict__ node_emb, int d, const int32_t* __restrict__ node_off,
                      const int32_t* __restrict__ g1, const int32_t* __restrict__ g2, int n_pairs, int bins,
                      float* __restrict__ hist, int32_t* __restrict__ counts) {
  __shared__ int s_cnt[4][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
    const int a0 = node_off[g1[p]], n1 = node_off[g1[p] + 1] - a0;
    const int b0 = node_off[g2[p]], n2 = node_off[g2[p] + 1] - b0;
    const int cells = n1 * n2;
    int mine = 0;   // lane b: cells of this warp that fell into bin b
    for (int c0 = warp * 32; c0 < cells; c0 += 128) {
      const int c = c0 + lane;
      int bin = -1;
      if (c < cells) {
        const int i = c / n2, j = c - i * n2;
        const float* x = node_emb + (size_t)(a0 + i) * d;
        const float* y = node_emb + (size_t)(b0 + j) * d;
        float s = 0.f;
        for (int k = 0; k < d; ++k) s = fmaf(x[k], y[k], s);
        bin = min(bins - 1, (int)floorf(sgnn_hist_sigmoid(s) * (float)bins));
      }
      for (int b = 0; b < bins; ++b) {
        const unsigned m = __ballot_sync(0xffffffffu, bin == b);
        if (lane == b) mine += __popc(m);
      }
    }
    s_cnt[warp][lane] = mine;
    __syncthreads();
    if (threadIdx.x < bins) {
      const int tot = s_cnt[0][lane] + s_cnt[1][lane] + s_cnt[2][lane] + s_cnt[3][lane];
      if (counts) counts[(size_t)p * bins + lane] = tot;
      if (hist) hist[(size_t)p * bins + lane] = cells > 0 ? (float)tot * (1.f / (float)cells) : 0.f;
    }
    __syncthreads();
  }
}
