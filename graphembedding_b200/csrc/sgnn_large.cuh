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
                   float* __restrict__ bufA, float* __restrict__ bufB, float* __restrict__ emb) {
  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;
  float* s_red = smem + P.w_floats;          // [2 * d + 8 warps * d] pooling scratch, also raw staging
  tile_stage_weights(M, P, params, s_w, s_red);
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
