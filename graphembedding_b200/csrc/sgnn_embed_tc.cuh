// Graph embedding with the per-layer feature x weight products on the tensor cores (north_star's
// first contraction): node rows of whole graphs are packed into tiles of <= 128 rows (the SAME closed-form tiling as
// the histogram's database operand: tile t = the graphs whose first node row lies in [S t, S (t + 1)), S = 129 -
// max_nodes), a GCN layer  H' = act(A_hat (H W) + b)  (layers.py:91-118) is
//   * layer 0 (one-hot input): a gather of W0 rows by neighbour label -- no product to make;
//   * layers l >= 1: ONE tcgen05 GEMM per tile, [128 rows x din] . [din x dout] with the activations written by their
//     threads as a 3-way bf16 split in the UMMA K-major layout and SIX products accumulated in fp32 in TMEM (the same
//     split as the histogram: ~2^-24, i.e. fp32 accuracy -- the histogram's integer bins sit downstream), then the
//     block-diagonal A_hat applied OUTSIDE the MMA: thread = node row, neighbour rows of H W gathered from shared
//     memory through the tile's CSR.
// Pooling (Average / Attention / Padding: layers.py:136-140, 154-160, 223-227) runs on the tile's final rows in shared
// memory; the node rows are written for the histogram layer.  Roles: 128 row threads + one MMA-issuing warp; two CTAs
// per SM.  Served: node stacks of GCN layers only (first one sparse), widths 16 / 32 / 48 / 64, graphs of <= 64
// nodes, no dropout, a tensor-core precision; everything else keeps embed_tile_kernel / embed_large_kernel.
#pragma once
#include "sgnn_common.cuh"
#include "sgnn_tc.cuh"

constexpr int ETC_ROW_WARPS = 4;           // one per TMEM lane quadrant: thread = node row of the tile
constexpr int ETC_THREADS = 32 * (ETC_ROW_WARPS + 1);
constexpr int ETC_MAXD = 64;
constexpr int ETC_ECAP = 1024;   // CSR entries of a tile staged in shared memory (a tile beyond that reads them from global memory)

struct EtcPlan {
  int S, n_tiles;
  int d0;                         // width of layer 0
  int dmax_in, dmax_out;          // over the dense layers l >= 1
  int w0_off, bias_off[SGNN_MAX_LAYERS], wt_off[SGNN_MAX_LAYERS];   // byte offsets into shared memory
  int a_off, xw_off, h_off, watt_off, csr_off, misc_off, total_bytes;
  int xw_ld, h_ld;
};

__host__ __device__ inline int etc_round16(int x) { return (x + 15) / 16 * 16; }

inline bool etc_supported(const ModelDev& M, int max_nodes) {
  if (M.n_node < 1 || max_nodes < 1 || max_nodes > 64 || M.keep < 1.f) return false;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    if (L.kind != SGNN_LAYER_GCN || L.drop) return false;
    if (L.dout % 16 != 0 || L.dout > ETC_MAXD || L.dout < 16) return false;
    if (l > 0 && L.din != M.L[l - 1].dout) return false;
  }
  if (M.pool_kind != SGNN_LAYER_AVERAGE && M.pool_kind != SGNN_LAYER_ATTENTION && M.pool_kind != SGNN_LAYER_PADDING) return false;
  return M.node_dim == M.L[M.n_node - 1].dout && M.input_dim <= 64;
}

inline EtcPlan etc_plan(const ModelDev& M, int max_nodes, int64_t total_rows) {
  EtcPlan P;
  P.S = 129 - max_nodes;
  P.n_tiles = (int)(total_rows / P.S + 1);
  P.d0 = M.L[0].dout;
  P.dmax_in = 16; P.dmax_out = 16;
  for (int l = 1; l < M.n_node; ++l) { P.dmax_in = std::max(P.dmax_in, M.L[l].din); P.dmax_out = std::max(P.dmax_out, M.L[l].dout); }
  int o = 0;
  P.w0_off = o; o += M.input_dim * (P.d0 + 1) * 4;
  for (int l = 0; l < M.n_node; ++l) { P.bias_off[l] = o; o += M.L[l].dout * 4; }
  o = (o + 15) / 16 * 16;
  for (int l = 0; l < M.n_node; ++l) {
    P.wt_off[l] = o;
    if (l > 0) o += 3 * (M.L[l].din / 8) * M.L[l].dout * 16;
  }
  o = (o + 127) / 128 * 128;
  P.a_off = o; o += 3 * (P.dmax_in / 8) * 128 * 16;
  P.xw_ld = std::max(P.dmax_out, 2 * M.node_dim) + 1;   // (the pooling scratch [graphs][d] x 2 lives here afterwards)
  P.xw_off = o; o += 128 * P.xw_ld * 4;
  P.h_ld = M.node_dim + 1;
  P.h_off = o; o += 128 * P.h_ld * 4;
  P.watt_off = o; o += M.pool_kind == SGNN_LAYER_ATTENTION ? M.node_dim * M.node_dim * 4 : 0;
  P.csr_off = o; o += (3 * ETC_ECAP + 128 + 129) * 4;   // neighbour row, value, layer-0 table offset per entry; labels; row pointers
  P.misc_off = o; o += (129 + 128 + 128 + 4) * 4 + 32;   // goff, gid, attention weights of the rows, tile bounds, mbarrier, TMEM slot
  P.total_bytes = o;
  return P;
}

// 16 fp32 values of row r (columns [jc, jc + 16) of a din-wide activation) -> 3-way bf16 split in the A operand tiles
__device__ __forceinline__ void etc_store_split(uint8_t* sA, int nkc, int r, int jc, const float* v) {
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    __align__(16) __nv_bfloat16 p[3][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x = v[half * 8 + j];
      p[0][j] = __float2bfloat16_rn(x);
      const float r1 = x - __bfloat162float(p[0][j]);
      p[1][j] = __float2bfloat16_rn(r1);
      p[2][j] = __float2bfloat16_rn(r1 - __bfloat162float(p[1][j]));
    }
    const int kc = jc / 8 + half;
#pragma unroll
    for (int pt = 0; pt < 3; ++pt)
      *reinterpret_cast<uint4*>(sA + ((size_t)(pt * nkc + kc) * 128 + r) * 16) = *reinterpret_cast<const uint4*>(p[pt]);
  }
}

__global__ void __launch_bounds__(ETC_THREADS, 2)
embed_tc_kernel(const __grid_constant__ ModelDev M, const __grid_constant__ EtcPlan P, const BatchDev B, int total_rows,
                const float* __restrict__ params, float* __restrict__ emb, float* __restrict__ node_out,
                int32_t* __restrict__ status) {
  extern __shared__ __align__(1024) uint8_t etc_smem[];
  float* sW0 = reinterpret_cast<float*>(etc_smem + P.w0_off);
  uint8_t* sA = etc_smem + P.a_off;
  float* sXW = reinterpret_cast<float*>(etc_smem + P.xw_off);
  float* sH = reinterpret_cast<float*>(etc_smem + P.h_off);
  float* sWatt = reinterpret_cast<float*>(etc_smem + P.watt_off);
  int* s_col = reinterpret_cast<int*>(etc_smem + P.csr_off);     // [ECAP] tile-local neighbour row of every entry
  float* s_val = reinterpret_cast<float*>(s_col + ETC_ECAP);     // [ECAP]
  int* s_lw = reinterpret_cast<int*>(s_val + ETC_ECAP);          // [ECAP] offset of the neighbour's label row in sW0
  int* s_lab = s_lw + ETC_ECAP;                                  // [128]
  int* s_rp = s_lab + 128;                                       // [129]
  int* s_goff = reinterpret_cast<int*>(etc_smem + P.misc_off);   // [129]
  int* s_gid = s_goff + 129;                                       // [128]
  float* s_att = reinterpret_cast<float*>(s_gid + 128);            // [128]
  int* s_tile = reinterpret_cast<int*>(s_att + 128);               // [2] first graph of this tile / of the next
  uint64_t* mbar = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(s_tile + 4) + 7) & ~uintptr_t(7));
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nl = M.n_node, d = M.node_dim;
  // ---- once per CTA: layer-0 table, biases, dense weights as 3-way bf16 split B tiles, attention weights ----
  {
    const LayerDev& L0 = M.L[0];
    for (int i = tid; i < M.input_dim * P.d0; i += ETC_THREADS) {
      const int a = i / P.d0, j = i - a * P.d0;
      sW0[a * (P.d0 + 1) + j] = params[L0.w_off + i];
    }
    for (int l = 0; l < nl; ++l) {
      float* sb = reinterpret_cast<float*>(etc_smem + P.bias_off[l]);
      for (int j = tid; j < M.L[l].dout; j += ETC_THREADS) sb[j] = M.L[l].bias ? params[M.L[l].b_off + j] : 0.f;
    }
    for (int l = 1; l < nl; ++l) {
      const LayerDev& L = M.L[l];
      uint8_t* sWt = etc_smem + P.wt_off[l];
      const int nkc = L.din / 8;
      for (int i = tid; i < L.din * L.dout; i += ETC_THREADS) {
        const int k = i / L.dout, n = i - k * L.dout;
        float w = params[L.w_off + i];
#pragma unroll
        for (int pt = 0; pt < 3; ++pt) {
          const __nv_bfloat16 hv = __float2bfloat16_rn(w);
          *reinterpret_cast<__nv_bfloat16*>(sWt + ((size_t)(pt * nkc + (k >> 3)) * L.dout + n) * 16 + (k & 7) * 2) = hv;
          w -= __bfloat162float(hv);
        }
      }
    }
    if (M.pool_kind == SGNN_LAYER_ATTENTION)
      for (int i = tid; i < d * d; i += ETC_THREADS) sWatt[i] = params[M.pool_w_off + i];
    if (tid == 0) { tc::mbar_init(mbar, 1); tc::fence_barrier_init(); }
    if (warp == ETC_ROW_WARPS) tc::tmem_alloc(tmem_slot, 64);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
  }
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t phase = 0;

  for (int t = blockIdx.x; t < P.n_tiles; t += gridDim.x) {
    // ---- tile bounds: graphs whose first row lies in [S t, S (t + 1)) (the last tile takes every remaining graph) ----
    if (tid < 2) {
      const int key = (t + tid) * P.S;
      int lo = 0, hi = B.n_graphs;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (B.node_off[mid] < key) lo = mid + 1; else hi = mid;
      }
      s_tile[tid] = (tid == 1 && t == P.n_tiles - 1) ? B.n_graphs : lo;
    }
    __syncthreads();
    const int g0 = s_tile[0], g1 = s_tile[1], ng = g1 - g0;
    if (ng <= 0) { __syncthreads(); continue; }
    const int row0 = B.node_off[g0];
    const int rows = B.node_off[g1] - row0;
    if (rows > 128) {   // max_nodes understated by the caller: flagged, the tile is skipped
      if (tid == 0 && status) atomicExch(status, 3);
      __syncthreads();
      continue;
    }
    for (int i = tid; i <= ng; i += ETC_THREADS) s_goff[i] = B.node_off[g0 + i] - row0;
    // the tile's CSR slice is contiguous: staged once (coalesced), every later access is a shared-memory read
    const int e_lo = B.rowptr[row0], nE = B.rowptr[row0 + rows] - e_lo;
    const bool staged = nE <= ETC_ECAP;
    if (staged) {
      for (int i = tid; i < nE; i += ETC_THREADS) { s_col[i] = B.col[e_lo + i]; s_val[i] = B.val[e_lo + i]; }
    }
    for (int i = tid; i < rows; i += ETC_THREADS) s_lab[i] = B.labels[row0 + i];
    for (int i = tid; i <= rows; i += ETC_THREADS) s_rp[i] = B.rowptr[row0 + i] - e_lo;
    __syncthreads();
    // thread = node row r of the tile (measured: four threads per row, one 16-column chunk each, 17 warps and one
    // CTA per SM is slower still -- the barriers between the phases of a tile cost more than the shorter chunks save)
    const bool rowthread = warp < ETC_ROW_WARPS;
    const int r = (warp & 3) * 32 + lane, chunk = 0;
    const bool valid = rowthread && r < rows;
    if (valid && chunk == 0) {
      int lo = 0, hi = ng - 1;
      while (lo < hi) {   // last graph with goff <= r
        const int mid = (lo + hi + 1) >> 1;
        if (s_goff[mid] <= r) lo = mid; else hi = mid - 1;
      }
      s_gid[r] = lo;
      const int gb = s_goff[lo];
      if (staged)   // CSR columns are graph-local: make them tile-local, and resolve the layer-0 table row
        for (int e = s_rp[r]; e < s_rp[r + 1]; ++e) {
          const int nb = gb + s_col[e];
          s_col[e] = nb;
          s_lw[e] = s_lab[nb] * (P.d0 + 1);
        }
    }
    __syncthreads();
    int gbase = 0, e0 = 0, e1 = 0;
    if (valid) { gbase = s_goff[s_gid[r]]; e0 = s_rp[r]; e1 = s_rp[r + 1]; }
    const int* gcol = B.col + e_lo; const float* gval = B.val + e_lo;
    // ---- layer 0: gather of W0 rows by neighbour label ----
    {
      const LayerDev& L = M.L[0];
      const float* sb = reinterpret_cast<const float*>(etc_smem + P.bias_off[0]);
      const bool last = nl == 1;
      const int nkc = last ? 0 : M.L[1].din / 8;
      if (rowthread) {
        for (int jc = 0; jc < L.dout; jc += 16) {
          float acc[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = 0.f;
          if (valid)
            for (int e = e0; e < e1; ++e) {
              const float w = staged ? s_val[e] : gval[e];
              const float* wr = sW0 + (staged ? s_lw[e] : s_lab[gbase + gcol[e]] * (P.d0 + 1)) + jc;
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] = fmaf(w, wr[j], acc[j]);
            }
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = valid ? sgnn_act(acc[j] + sb[jc + j], L.act, M.yeta) : 0.f;
          if (!last) etc_store_split(sA, nkc, r, jc, acc);
          else {
#pragma unroll
            for (int j = 0; j < 16; ++j) sH[r * P.h_ld + jc + j] = acc[j];
          }
        }
      }
    }
    // ---- dense GCN layers: MMA (H W), then A_hat through the tile's CSR ----
    for (int l = 1; l < nl; ++l) {
      const LayerDev& L = M.L[l];
      const bool last = l == nl - 1;
      tc::fence_proxy_async_smem();
      tc::tc_fence_before();
      __syncthreads();
      if (tid == 32 * ETC_ROW_WARPS) {   // the MMA issuer: six products of the 3-way split, K = din
        tc::tc_fence_after();
        const int nkc = L.din / 8, nks = L.din / 16;
        const uint32_t idesc = tc::umma_idesc_bf16(128, L.dout);
        const uint64_t ad = tc::umma_desc(tc::smem_u32(sA), 128 * 16, 128);
        const uint64_t bd = tc::umma_desc(tc::smem_u32(etc_smem + P.wt_off[l]), (uint32_t)L.dout * 16, 128);
        const uint32_t a_part = (uint32_t)(nkc * 128 * 16) >> 4, b_part = (uint32_t)(nkc * L.dout * 16) >> 4;
        const uint32_t a_ks = (2 * 128 * 16) >> 4, b_ks = (uint32_t)(2 * L.dout * 16) >> 4;
        const uint32_t pa = 0x201100u, pb = 0x021010u;
        uint32_t accum = 0;
        for (int p = 0; p < 6; ++p)
          for (int ks = 0; ks < nks; ++ks) {
            tc::umma_bf16(tmem_base, ad + ((pa >> (4 * p)) & 15u) * a_part + (uint64_t)(ks * a_ks),
                          bd + ((pb >> (4 * p)) & 15u) * b_part + (uint64_t)(ks * b_ks), idesc, accum);
            accum = 1;
          }
        tc::umma_commit(mbar);
      }
      if (rowthread) {
        tc::mbar_wait(mbar, phase);
        tc::tc_fence_after();
        for (int jc = 0; jc < L.dout; jc += 16) {
          float v[16];
          tc::tmem_ld16(taddr + (uint32_t)jc, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) sXW[r * P.xw_ld + jc + j] = v[j];
        }
        tc::tc_fence_before();
      }
      phase ^= 1u;
      __syncthreads();
      if (rowthread) {
        const float* sb = reinterpret_cast<const float*>(etc_smem + P.bias_off[l]);
        const int nkc = last ? 0 : M.L[l + 1].din / 8;
        for (int jc = 0; jc < L.dout; jc += 16) {
          float acc[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = 0.f;
          if (valid)
            for (int e = e0; e < e1; ++e) {
              const float w = staged ? s_val[e] : gval[e];
              const float* xr = sXW + (staged ? s_col[e] : gbase + gcol[e]) * P.xw_ld + jc;
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] = fmaf(w, xr[j], acc[j]);
            }
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = valid ? sgnn_act(acc[j] + sb[jc + j], L.act, M.yeta) : 0.f;
          if (!last) {
            // (the MMA that read sA has completed: its commit was waited for above)
            etc_store_split(sA, nkc, r, jc, acc);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) sH[r * P.h_ld + jc + j] = acc[j];
          }
        }
      }
    }
    __syncthreads();
    // ---- outputs: node rows (coalesced), pooling per graph ----
    if (node_out)
      for (int i = tid; i < rows * d; i += ETC_THREADS) {
        const int rr = i / d, c = i - rr * d;
        node_out[(size_t)(row0 + rr) * d + c] = sH[rr * P.h_ld + c];
      }
    if (emb) {
      const int E = M.emb_dim;
      if (M.pool_kind == SGNN_LAYER_PADDING) {
        for (int i = tid; i < ng * E; i += ETC_THREADS) {
          const int gi = i / E, idx = i - gi * E;
          const int p = idx / d, c = idx - p * d;
          const int n = s_goff[gi + 1] - s_goff[gi];
          emb[(size_t)(g0 + gi) * E + idx] = p < n ? sH[(s_goff[gi] + p) * P.h_ld + c] : M.pool_fparam;
        }
      } else {
        float* s_mean = sXW;                 // [ng][d]   (the XW rows are dead)
        float* s_c = sXW + 128 * d;          // [ng][d]
        const bool att = M.pool_kind == SGNN_LAYER_ATTENTION;
        for (int i = tid; i < ng * d; i += ETC_THREADS) {
          const int gi = i / d, j = i - gi * d;
          const int a = s_goff[gi], b = s_goff[gi + 1];
          float s = 0.f;
          for (int rr = a; rr < b; ++rr) s += sH[rr * P.h_ld + j];
          s = b > a ? s / (float)(b - a) : 0.f;
          if (att) s_mean[i] = s; else emb[(size_t)(g0 + gi) * E + j] = s;
        }
        if (att) {
          __syncthreads();
          for (int i = tid; i < ng * d; i += ETC_THREADS) {
            const int gi = i / d, j = i - gi * d;
            float s = 0.f;
            for (int k = 0; k < d; ++k) s = fmaf(s_mean[gi * d + k], sWatt[k * d + j], s);
            s_c[i] = tanhf(s);
          }
          __syncthreads();
          if (valid && chunk == 0) {
            const float* c = s_c + s_gid[r] * d;
            float tt = 0.f;
            for (int k = 0; k < d; ++k) tt = fmaf(sH[r * P.h_ld + k], c[k], tt);
            s_att[r] = 1.f / (1.f + expf(-tt));
          }
          __syncthreads();
          for (int i = tid; i < ng * d; i += ETC_THREADS) {
            const int gi = i / d, j = i - gi * d;
            float s = 0.f;
            for (int rr = s_goff[gi]; rr < s_goff[gi + 1]; ++rr) s = fmaf(s_att[rr], sH[rr * P.h_ld + j], s);
            emb[(size_t)(g0 + gi) * E + j] = s;
          }
        }
      }
    }
    __syncthreads();
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == ETC_ROW_WARPS) tc::tmem_dealloc(tmem_base, 64);
}
