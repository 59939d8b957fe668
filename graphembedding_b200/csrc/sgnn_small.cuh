// Small-graph kernels (N <= NMAX in {16, 32}): one warp owns one graph (embedding) or one
// pair (training).  Lanes map to OUTPUT FEATURE COLUMNS, node rows live in per-warp shared
// memory, so every inner product is "uniform broadcast row value x per-lane weight" with
// warp-uniform control flow: no shuffles in the GEMMs, neighbour aggregation is a CSR walk
// over shared memory.  Reference math: model/Siamese/layers.py:91-118 (GCN), :136-140
// (Average), :154-160 (Attention), :192-205 (Dense), :223-227 (Padding), :249-252 (Dot),
// :282-310 (NTN); loss model_mse.py:145-151.
#pragma once
#include "sgnn_common.cuh"

struct GraphS {
  int n, nnz;
  const int* rp;     // [n+1] row offsets local to the graph
  const int* lab;    // [n]
  const int* col;    // [nnz] local node index
  const float* val;  // [nnz]
};

template <int NMAX>
struct SmallGeom {
  static constexpr int EMAX = NMAX * 8;                       // staged nnz capacity
  static constexpr int CSR_INTS = (NMAX + 1) + NMAX + EMAX;   // rp, lab, col
  static constexpr int CSR_WORDS = ((CSR_INTS + EMAX) + 3) / 4 * 4;
};

template <int NMAX>
__device__ __forceinline__ GraphS load_graph(const BatchDev& B, int g, float* scratch, int lane) {
  using Geo = SmallGeom<NMAX>;
  int* s_rp = reinterpret_cast<int*>(scratch);
  int* s_lab = s_rp + (NMAX + 1);
  int* s_col = s_lab + NMAX;
  float* s_val = scratch + Geo::CSR_INTS;
  const int base = B.node_off[g];
  const int n = B.node_off[g + 1] - base;
  const int e0 = B.rowptr[base];
  const int nnz = B.rowptr[base + n] - e0;
  for (int i = lane; i <= n; i += 32) s_rp[i] = B.rowptr[base + i] - e0;
  for (int i = lane; i < n; i += 32) s_lab[i] = B.labels[base + i];
  GraphS G;
  G.n = n; G.nnz = nnz; G.rp = s_rp; G.lab = s_lab;
  if (nnz <= Geo::EMAX) {
    for (int e = lane; e < nnz; e += 32) { s_col[e] = B.col[e0 + e]; s_val[e] = B.val[e0 + e]; }
    G.col = s_col; G.val = s_val;
  } else {
    G.col = B.col + e0; G.val = B.val + e0;
  }
  __syncwarp();
  return G;
}

// Out(i, j) = sum_k H[i*ldh + k] * W[k*dout + j], delivered through epi(i, j, value).
// lane = j; NMAX row accumulators in registers; H rows read as broadcast float4.
template <int NMAX, typename Epi>
__device__ __forceinline__ void warp_gemm(const float* __restrict__ H, int ldh, int n, int din,
                                          const float* __restrict__ W, int dout, int lane, Epi epi) {
  if (dout <= 8) {  // thin output: lane = (row, col)
    for (int idx = lane; idx < n * dout; idx += 32) {
      const int i = idx / dout, j = idx - i * dout;
      float s = 0.f;
      for (int k = 0; k < din; ++k) s = fmaf(H[i * ldh + k], W[k * dout + j], s);
      epi(i, j, s);
    }
    return;
  }
  for (int j0 = 0; j0 < dout; j0 += 32) {
    const int j = j0 + lane;
    const bool on = j < dout;
    const int jc = on ? j : 0;
    float acc[NMAX];
#pragma unroll
    for (int i = 0; i < NMAX; ++i) acc[i] = 0.f;
    int k = 0;
    for (; k + 4 <= din; k += 4) {
      const float w0 = W[(k + 0) * dout + jc], w1 = W[(k + 1) * dout + jc];
      const float w2 = W[(k + 2) * dout + jc], w3 = W[(k + 3) * dout + jc];
#pragma unroll
      for (int i = 0; i < NMAX; ++i) {
        if (i < n) {
          const float4 h = *reinterpret_cast<const float4*>(H + i * ldh + k);
          acc[i] = fmaf(h.x, w0, acc[i]);
          acc[i] = fmaf(h.y, w1, acc[i]);
          acc[i] = fmaf(h.z, w2, acc[i]);
          acc[i] = fmaf(h.w, w3, acc[i]);
        }
      }
    }
    for (; k < din; ++k) {
      const float w0 = W[k * dout + jc];
#pragma unroll
      for (int i = 0; i < NMAX; ++i)
        if (i < n) acc[i] = fmaf(H[i * ldh + k], w0, acc[i]);
    }
    if (on) {
#pragma unroll
      for (int i = 0; i < NMAX; ++i)
        if (i < n) epi(i, j, acc[i]);
    }
  }
}

// Out[i][j] = act( sum_{e in row i} val[e] * T[col[e]][j] + b[j] )      (A_hat . XW + b)
__device__ __forceinline__ void warp_aggregate(const GraphS& G, const float* T, int ldt, int dout,
                                               const float* bias, int act, float* Out, int ldo,
                                               int lane) {
  for (int j = lane; j < dout; j += 32) {
    const float bj = bias ? bias[j] : 0.f;
    for (int i = 0; i < G.n; ++i) {
      float s = 0.f;
      for (int e = G.rp[i]; e < G.rp[i + 1]; ++e) s = fmaf(G.val[e], T[G.col[e] * ldt + j], s);
      Out[i * ldo + j] = sgnn_act(s + bj, act, 0.f);
    }
  }
}

// One node-level layer (GCN or Dense).  Hin: [n][ldin]; T: scratch [n][ldt]; Hout: [n][ldo].
// drop (training / single-pair evaluation with FLAGS.dropout > 0): dropout on the layer input.
// Layer 0 masks whole one-hot rows (sparse_dropout: one draw per non-zero, layers.py:332-338) through
// nmask[NMAX]; later layers mask Hin IN PLACE (so the saved activation is the masked one the weight
// gradient needs).
template <int NMAX>
__device__ __forceinline__ void node_layer_fwd(const LayerDev& L, int l, const GraphS& G,
                                               const float* W, float* Hin, int ldin, float* T,
                                               int ldt, float* Hout, int ldo, int lane,
                                               bool drop = false, uint32_t dkey = 0, float keep = 1.f,
                                               float inv_keep = 1.f, float* nmask = nullptr) {
  const float* w = W + L.w_off;
  const float* b = L.bias ? W + L.b_off : nullptr;
  if (drop && l > 0) {
    for (int idx = lane; idx < G.n * L.din; idx += 32) {
      const int i = idx / L.din, k = idx - i * L.din;
      Hin[i * ldin + k] *= drop_mask(dkey, i, k, keep, inv_keep);
    }
    __syncwarp();
  }
  if (L.kind == SGNN_LAYER_GCN) {
    if (l == 0) {
      if (nmask) {
        for (int i = lane; i < G.n; i += 32) nmask[i] = drop ? drop_mask(dkey, i, 0, keep, inv_keep) : 1.f;
        __syncwarp();
      }
      // sparse one-hot input (graphs.py:55-62): X.W is a row gather of W
      for (int j = lane; j < L.dout; j += 32) {
        const float bj = b ? b[j] : 0.f;
        for (int i = 0; i < G.n; ++i) {
          float s = 0.f;
          for (int e = G.rp[i]; e < G.rp[i + 1]; ++e) {
            const int c = G.col[e];
            const float v = nmask ? G.val[e] * nmask[c] : G.val[e];
            s = fmaf(v, w[G.lab[c] * L.dout + j], s);
          }
          Hout[i * ldo + j] = sgnn_act(s + bj, L.act, 0.f);
        }
      }
    } else {
      warp_gemm<NMAX>(Hin, ldin, G.n, L.din, w, L.dout, lane,
                      [&](int i, int j, float v) { T[i * ldt + j] = v; });
      __syncwarp();
      warp_aggregate(G, T, ldt, L.dout, b, L.act, Hout, ldo, lane);
    }
  } else {  // Dense
    const int act = L.act;
    warp_gemm<NMAX>(Hin, ldin, G.n, L.din, w, L.dout, lane, [&](int i, int j, float v) {
      Hout[i * ldo + j] = sgnn_act(v + (b ? b[j] : 0.f), act, 0.f);
    });
  }
  __syncwarp();
}

// Pooling / padding: node matrix H [n][node_dim] -> emb[emb_dim].  save: [2*d + NMAX]
// floats (mean, c, att) kept for the attention backward.
__device__ __forceinline__ void pool_fwd(const ModelDev& M, const GraphS& G, const float* W,
                                         const float* H, int ld, float* emb, float* save, int lane) {
  const int d = M.node_dim, n = G.n;
  if (M.pool_kind == SGNN_LAYER_PADDING) {
    for (int idx = lane; idx < M.emb_dim; idx += 32) {
      const int p = idx / d, c = idx - p * d;
      emb[idx] = p < n ? H[p * ld + c] : M.pool_fparam;
    }
  } else if (M.pool_kind == SGNN_LAYER_AVERAGE) {
    for (int j = lane; j < d; j += 32) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s += H[i * ld + j];
      emb[j] = s / (float)n;
    }
  } else {  // Attention (layers.py:154-160)
    float* mean = save; float* c = save + d; float* att = save + 2 * d;
    const float* w = W + M.pool_w_off;
    for (int j = lane; j < d; j += 32) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s += H[i * ld + j];
      mean[j] = s / (float)n;
    }
    __syncwarp();
    for (int j = lane; j < d; j += 32) {
      float s = 0.f;
      for (int k = 0; k < d; ++k) s = fmaf(mean[k], w[k * d + j], s);
      c[j] = tanhf(s);
    }
    __syncwarp();
    for (int i = lane; i < n; i += 32) {
      float s = 0.f;
      for (int k = 0; k < d; ++k) s = fmaf(H[i * ld + k], c[k], s);
      att[i] = 1.f / (1.f + expf(-s));
    }
    __syncwarp();
    for (int j = lane; j < d; j += 32) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s = fmaf(att[i], H[i * ld + j], s);
      emb[j] = s;
    }
  }
  __syncwarp();
}

// Node layers + pooling of one graph.  bufs: 3 x [NMAX][ld] ping/pong/scratch.
// If acts != nullptr every layer output is written to acts + L.out_off (row stride act_ld)
// and kept (training); otherwise two buffers alternate.
template <int NMAX>
__device__ __forceinline__ void graph_fwd(const ModelDev& M, const GraphS& G, const float* W,
                                          float* bufs, float* acts, float* emb, float* save,
                                          int lane, uint32_t dbase = 0, int slot = 0, float* nmask = nullptr) {
  const int ld = M.ld;
  float* T = bufs + 2 * NMAX * ld;
  float* Hin = nullptr;
  int ldin = 0;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    float* Hout; int ldo;
    if (acts) { Hout = acts + L.out_off; ldo = M.act_ld; }
    else { Hout = bufs + (l & 1) * NMAX * ld; ldo = ld; }
    const bool drop = nmask != nullptr && L.drop;   // nmask is only supplied by the dropout-enabled callers
    const uint32_t dkey = drop ? drop_layer_key(dbase, slot, L.ref) : 0u;
    node_layer_fwd<NMAX>(L, l, G, W, Hin, ldin, T, ld, Hout, ldo, lane, drop, dkey, M.keep, M.inv_keep, nmask);
    Hin = Hout; ldin = ldo;
  }
  pool_fwd(M, G, W, Hin, ldin, emb, save, lane);
}

// NTN / Dot on one pair.  Wp: [b][a*K + k] permuted copy of weights_W (conflict-free for
// u[a][k] = sum_b W[a][b][k] x2[b]).  ubuf [D*K], mbuf [K], tbuf [K] are kept for backward.
__device__ __forceinline__ float pair_fwd(const ModelDev& M, const float* W, const float* Wp,
                                          const float* x1, const float* x2, float* ubuf,
                                          float* mbuf, float* tbuf, int lane, float* st_out,
                                          float* su_out) {
  const int D = M.D, K = M.K;
  if (M.pair_kind == SGNN_LAYER_DOT) {
    float s = 0.f;
    for (int a = lane; a < D; a += 32) s = fmaf(x1[a], x2[a], s);
    return warp_sum(s);
  }
  const int DK = D * K;
  for (int ak = lane; ak < DK; ak += 32) {
    float s = 0.f;
    for (int b = 0; b < D; ++b) s = fmaf(Wp[b * DK + ak], x2[b], s);
    ubuf[ak] = s;
  }
  __syncwarp();
  const float* V = W + M.v_off;
  const float* U = W + M.u_off;
  float st = 0.f, su = 0.f, sut = 0.f;
  for (int k = lane; k < K; k += 32) {
    float s = M.pair_bias ? W[M.b_off + k] : 0.f;
    for (int a = 0; a < D; ++a) s = fmaf(x1[a], ubuf[a * K + k], s);
    for (int c = 0; c < D; ++c) {
      s = fmaf(V[k * 2 * D + c], x1[c], s);
      s = fmaf(V[k * 2 * D + D + c], x2[c], s);
    }
    const float t = sgnn_act(s, M.pair_act, M.yeta);
    mbuf[k] = s; tbuf[k] = t;
    st += t; su += U[k]; sut = fmaf(U[k], t, sut);
  }
  st = warp_sum(st); su = warp_sum(su); sut = warp_sum(sut);
  if (st_out) { *st_out = st; *su_out = su; }
  __syncwarp();
  return M.ntn_mode == SGNN_NTN_REFERENCE ? su * st : sut;  // F5: layers.py:305-308
}

__device__ __forceinline__ void stage_weights(const ModelDev& M, const float* params, float* s_w,
                                              float* s_wp) {
  // every load of a thread is issued before its first store: the kernels that stage ~4 400 parameters with 256-512
  // threads paid one memory latency per loop trip here (ncu r2t: 13 % of train_team_kernel's samples on the store that
  // waits for its load); 16-byte loads where the block is aligned
  const int nt = blockDim.x;
  if ((reinterpret_cast<uintptr_t>(params) & 15) == 0 && (reinterpret_cast<uintptr_t>(s_w) & 15) == 0) {
    const int n4 = M.n_params >> 2;
    const float4* p4 = reinterpret_cast<const float4*>(params);
    float4* s4 = reinterpret_cast<float4*>(s_w);
    for (int i0 = threadIdx.x; i0 < n4; i0 += 4 * nt) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i0 + u * nt < n4) v[u] = __ldg(p4 + i0 + u * nt);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i0 + u * nt < n4) s4[i0 + u * nt] = v[u];
    }
    for (int i = 4 * n4 + threadIdx.x; i < M.n_params; i += nt) s_w[i] = params[i];
  } else {
    for (int i = threadIdx.x; i < M.n_params; i += nt) s_w[i] = params[i];
  }
  if (s_wp && M.pair_kind == SGNN_LAYER_NTN) {
    const int D = M.D, K = M.K, DK = D * K, n = D * DK;
    const float* w = params + M.w_off;
    for (int i0 = threadIdx.x; i0 < n; i0 += 8 * nt) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) if (i0 + u * nt < n) v[u] = __ldg(w + i0 + u * nt);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * nt;
        if (i < n) {
          const int a = i / DK, r = i - a * DK, b = r / K, k = r - b * K;
          s_wp[b * DK + a * K + k] = v[u];
        }
      }
    }
  }
}

__host__ __device__ inline int embed_warp_floats(const ModelDev& M, int nmax, int csr_words) {
  return 3 * nmax * M.ld + csr_words + ((M.emb_dim + 3) / 4 * 4) + ((2 * M.node_dim + nmax + 3) / 4 * 4);
}

// ---------------------------------------------------------------------------------------
// embed: emb[g][:] for every graph in the batch
// ---------------------------------------------------------------------------------------
template <int NMAX>
__global__ void __launch_bounds__(256)
embed_kernel(const __grid_constant__ ModelDev M, const BatchDev B, const float* __restrict__ params,
             float* __restrict__ emb) {
  extern __shared__ __align__(16) float smem[];
  using Geo = SmallGeom<NMAX>;
  float* s_w = smem;
  const int wfl = (M.n_params + 3) / 4 * 4;
  stage_weights(M, params, s_w, nullptr);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float* my = smem + wfl + warp * embed_warp_floats(M, NMAX, Geo::CSR_WORDS);
  float* bufs = my;
  float* csr = bufs + 3 * NMAX * M.ld;
  float* s_emb = csr + Geo::CSR_WORDS;
  float* save = s_emb + (M.emb_dim + 3) / 4 * 4;
  for (int g = blockIdx.x * wpb + warp; g < B.n_graphs; g += gridDim.x * wpb) {
    GraphS G = load_graph<NMAX>(B, g, csr, lane);
    graph_fwd<NMAX>(M, G, s_w, bufs, nullptr, s_emb, save, lane);
    for (int i = lane; i < M.emb_dim; i += 32) emb[(size_t)g * M.emb_dim + i] = s_emb[i];
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------
// pair scoring from precomputed embeddings: warp per pair
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
pair_score_kernel(const __grid_constant__ ModelDev M, const float* __restrict__ params,
                  const float* __restrict__ emb, const int32_t* __restrict__ g1,
                  const int32_t* __restrict__ g2, int n_pairs, float* __restrict__ raw,
                  float* __restrict__ sim) {
  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;
  const int wfl = (M.n_params + 3) / 4 * 4;
  const int DK = M.D * M.K;
  float* s_wp = s_w + wfl;
  const int wpfl = (M.pair_kind == SGNN_LAYER_NTN) ? (M.D * DK + 3) / 4 * 4 : 0;
  stage_weights(M, params, s_w, s_wp);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int Dp = (M.D + 3) / 4 * 4, Kp = (M.K + 3) / 4 * 4;
  const int per = 2 * Dp + ((DK + 3) / 4 * 4) + 2 * Kp;
  float* my = s_wp + wpfl + warp * per;
  float* x1 = my; float* x2 = x1 + Dp; float* ub = x2 + Dp; float* mb = ub + (DK + 3) / 4 * 4;
  float* tb = mb + Kp;
  for (int p = blockIdx.x * wpb + warp; p < n_pairs; p += gridDim.x * wpb) {
    const float* e1 = emb + (size_t)g1[p] * M.emb_dim;
    const float* e2 = emb + (size_t)g2[p] * M.emb_dim;
    for (int i = lane; i < M.D; i += 32) { x1[i] = e1[i]; x2[i] = e2[i]; }
    __syncwarp();
    const float r = pair_fwd(M, s_w, s_wp, x1, x2, ub, mb, tb, lane, nullptr, nullptr);
    if (lane == 0) {
      if (raw) raw[p] = r;
      if (sim) sim[p] = sgnn_act(r, M.final_act, M.yeta);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------
// training: one warp per pair, forward + backward fused; parameter gradients accumulate in
// a per-CTA shared buffer (red.shared) and are flushed with one global atomic per
// parameter per CTA.
// ---------------------------------------------------------------------------------------
struct TrainLayout {
  int acts, csr, save, emb, ubuf, vbuf, mbuf, tbuf, dmbuf, dx, gbuf, nmask, total;
};

__host__ __device__ inline TrainLayout train_layout(const ModelDev& M, int nmax, int csr_words) {
  TrainLayout t;
  int o = 0;
  const int r4 = 3;
  t.acts = o; o += 2 * nmax * M.act_ld;
  t.csr = o; o += 2 * csr_words;
  t.save = o; o += 2 * ((2 * M.node_dim + nmax + r4) & ~r4);
  t.emb = o; o += 2 * ((M.emb_dim + r4) & ~r4);
  const int DK = (M.D * M.K + r4) & ~r4, Kp = (M.K + r4) & ~r4, Dp = (M.D + r4) & ~r4;
  t.ubuf = o; o += DK;
  t.vbuf = o; o += DK;
  t.mbuf = o; o += Kp;
  t.tbuf = o; o += Kp;
  t.dmbuf = o; o += Kp;
  t.dx = o; o += 2 * ((M.emb_dim + r4) & ~r4) > 2 * Dp ? 2 * ((M.emb_dim + r4) & ~r4) : 2 * Dp;
  t.gbuf = o; o += 3 * nmax * M.ld;
  t.nmask = o; o += 2 * nmax;   // layer-0 node masks of the two arms (dropout)
  t.total = o;
  return t;
}

// transposed copies of every dense weight ([dout][din]) for the input-gradient GEMMs
__device__ __forceinline__ void stage_transposed(const ModelDev& M, const float* s_w, float* s_aux) {
  for (int l = 1; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    for (int i = threadIdx.x; i < L.din * L.dout; i += blockDim.x) {
      const int k = i / L.dout, j = i - k * L.dout;
      s_aux[L.wt_off + j * L.din + k] = s_w[L.w_off + i];
    }
  }
  if (M.pool_kind == SGNN_LAYER_ATTENTION) {
    const int d = M.node_dim;
    for (int i = threadIdx.x; i < d * d; i += blockDim.x) {
      const int k = i / d, j = i - k * d;
      s_aux[M.pool_wt_off + j * d + k] = s_w[M.pool_w_off + i];
    }
  }
}

template <int NMAX>
__device__ __forceinline__ void pool_bwd(const ModelDev& M, const GraphS& G, const float* W,
                                         const float* Wt_att, const float* H, int ldh,
                                         const float* demb, const float* save, float* Gout, int ld,
                                         float* tmp, float* s_g, int lane) {
  const int d = M.node_dim, n = G.n;
  if (M.pool_kind == SGNN_LAYER_PADDING) {
    for (int idx = lane; idx < n * d; idx += 32) {
      const int p = idx / d, c = idx - p * d;
      Gout[p * ld + c] = demb[idx];
    }
  } else if (M.pool_kind == SGNN_LAYER_AVERAGE) {
    const float inv = 1.f / (float)n;
    for (int j = lane; j < d; j += 32) {
      const float v = demb[j] * inv;
      for (int i = 0; i < n; ++i) Gout[i * ld + j] = v;
    }
  } else {
    const float* mean = save; const float* c = save + d; const float* att = save + 2 * d;
    float* dz = tmp; float* dq = tmp + NMAX; float* dmean = dq + d;
    for (int i = lane; i < n; i += 32) {
      float s = 0.f;
      for (int j = 0; j < d; ++j) s = fmaf(demb[j], H[i * ldh + j], s);
      dz[i] = s * att[i] * (1.f - att[i]);
    }
    __syncwarp();
    for (int k = lane; k < d; k += 32) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s = fmaf(dz[i], H[i * ldh + k], s);
      dq[k] = s * (1.f - c[k] * c[k]);
    }
    __syncwarp();
    for (int j = lane; j < d; j += 32) {
      const float q = dq[j];
      for (int k = 0; k < d; ++k) atomicAdd(&s_g[M.pool_w_off + k * d + j], mean[k] * q);
    }
    for (int k = lane; k < d; k += 32) {
      float s = 0.f;
      for (int j = 0; j < d; ++j) s = fmaf(Wt_att[j * d + k], dq[j], s);
      dmean[k] = s / (float)n;
    }
    __syncwarp();
    for (int j = lane; j < d; j += 32) {
      const float de = demb[j], cj = c[j], dm = dmean[j];
      for (int i = 0; i < n; ++i) Gout[i * ld + j] = att[i] * de + dz[i] * cj + dm;
    }
  }
  __syncwarp();
}

// Backward through the node layers of one graph.  Gcur holds dL/d(output of last node
// layer) [n][ld]; Gnxt and T are scratch of the same shape.
template <int NMAX>
__device__ __forceinline__ void graph_bwd(const ModelDev& M, const GraphS& G, const float* W,
                                          const float* s_aux, const float* acts, float* Gcur,
                                          float* Gnxt, float* T, float* s_g, int lane,
                                          bool dropping = false, uint32_t dbase = 0, int slot = 0,
                                          const float* nmask = nullptr) {
  const int ld = M.ld, n = G.n;
  for (int l = M.n_node - 1; l >= 0; --l) {
    const LayerDev& L = M.L[l];
    const float* Hout = acts + L.out_off;
    const int dout = L.dout;
    // If the NEXT node layer dropped its input, the saved Hout is y * mask (mask in {0, 1/keep}) and the
    // incoming gradient is w.r.t. the masked value: dL/dy = g * mask, and y = saved * keep where kept.
    const bool masked = dropping && l + 1 < M.n_node && M.L[l + 1].drop;
    const uint32_t mkey = masked ? drop_layer_key(dbase, slot, M.L[l + 1].ref) : 0u;
    // dZ = dOut * act'(out); bias gradient
    for (int j = lane; j < dout; j += 32) {
      float sb = 0.f;
      for (int i = 0; i < n; ++i) {
        float y = Hout[i * M.act_ld + j];
        float g = Gcur[i * ld + j];
        if (masked) { g *= drop_mask(mkey, i, j, M.keep, M.inv_keep); y *= M.keep; }
        g *= sgnn_act_grad(y, 0.f, L.act, 0.f);
        Gcur[i * ld + j] = g;
        sb += g;
      }
      if (L.bias) atomicAdd(&s_g[L.b_off + j], sb);
    }
    __syncwarp();
    const float* dXW = Gcur;
    if (L.kind == SGNN_LAYER_GCN) {  // A_hat is symmetric: dXW = A_hat . dZ
      warp_aggregate(G, Gcur, ld, dout, nullptr, SGNN_ACT_IDENTITY, T, ld, lane);
      __syncwarp();
      dXW = T;
    }
    if (l == 0) {
      // one-hot input: dW[label_i][:] += mask_i * dXW[i][:]
      for (int j = lane; j < dout; j += 32)
        for (int i = 0; i < n; ++i)
          atomicAdd(&s_g[L.w_off + G.lab[i] * dout + j], nmask ? nmask[i] * dXW[i * ld + j] : dXW[i * ld + j]);
    } else {
      const float* Hin = acts + M.L[l - 1].out_off;
      const int din = L.din;
      // dW[k][j] += sum_i Hin[i][k] * dXW[i][j]     (lane = j, dXW column in registers)
      for (int j0 = 0; j0 < dout; j0 += 32) {
        const int j = j0 + lane;
        if (j < dout) {
          float t[NMAX];
#pragma unroll
          for (int i = 0; i < NMAX; ++i) t[i] = i < n ? dXW[i * ld + j] : 0.f;
          for (int k = 0; k < din; ++k) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < NMAX; ++i)
              if (i < n) s = fmaf(Hin[i * M.act_ld + k], t[i], s);
            atomicAdd(&s_g[L.w_off + k * dout + j], s);
          }
        }
      }
      // dHin[i][k] = sum_j dXW[i][j] * W[k][j]  == GEMM with the transposed copy
      warp_gemm<NMAX>(dXW, ld, n, dout, s_aux + L.wt_off, din, lane,
                      [&](int i, int k, float v) { Gnxt[i * ld + k] = v; });
      __syncwarp();
      float* sw = Gcur; Gcur = Gnxt; Gnxt = sw;
    }
    __syncwarp();
  }
}

template <int NMAX, bool BWD>
__global__ void __launch_bounds__(128)
train_pairs_kernel(const __grid_constant__ ModelDev M, const BatchDev B,
                   const float* __restrict__ params, const int32_t* __restrict__ g1,
                   const int32_t* __restrict__ g2, const float* __restrict__ target_sim, int n_pairs,
                   float inv_global_batch, float tsum, float tsqsum, float* __restrict__ grads,
                   float* __restrict__ loss, float* __restrict__ raw_out,
                   float* __restrict__ sim_out, int dropping, uint32_t dbase, int p0, int pstride) {
  extern __shared__ __align__(16) float smem[];
  using Geo = SmallGeom<NMAX>;
  const int wfl = (M.n_params + 3) / 4 * 4;
  float* s_w = smem;
  float* s_g = s_w + wfl;
  float* s_wp = s_g + wfl;
  const int DK = M.D * M.K;
  const int wpfl = (M.pair_kind == SGNN_LAYER_NTN) ? (M.D * DK + 3) / 4 * 4 : 0;
  float* s_aux = s_wp + wpfl;
  float* warp_base = s_aux + ((M.aux_floats + 3) / 4 * 4);
  stage_weights(M, params, s_w, s_wp);
  for (int i = threadIdx.x; i < wfl; i += blockDim.x) s_g[i] = 0.f;
  __syncthreads();
  stage_transposed(M, s_w, s_aux);
  __syncthreads();

  const TrainLayout TL = train_layout(M, NMAX, Geo::CSR_WORDS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float* my = warp_base + warp * TL.total;
  const int embp = (M.emb_dim + 3) & ~3;
  const int savep = (2 * M.node_dim + NMAX + 3) & ~3;
  float loss_acc = 0.f;

  for (int p = blockIdx.x * wpb + warp; p < n_pairs; p += gridDim.x * wpb) {
    GraphS Gs[2];
    float* acts[2]; float* embs[2]; float* saves[2]; float* nmask[2];
    const int gid[2] = {g1[p], g2[p]};
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      acts[s] = my + TL.acts + s * NMAX * M.act_ld;
      embs[s] = my + TL.emb + s * embp;
      saves[s] = my + TL.save + s * savep;
      nmask[s] = dropping ? my + TL.nmask + s * NMAX : nullptr;
      Gs[s] = load_graph<NMAX>(B, gid[s], my + TL.csr + s * Geo::CSR_WORDS, lane);
      graph_fwd<NMAX>(M, Gs[s], s_w, my + TL.gbuf, acts[s], embs[s], saves[s], lane, dbase, 2 * (p0 + p * pstride) + s, nmask[s]);
      if (dropping && M.pair_drop) {   // dropout on the pair layer's inputs (layers.py:287-288)
        const uint32_t key = drop_layer_key(dbase, 2 * (p0 + p * pstride) + s, M.pair_ref);
        for (int a = lane; a < M.emb_dim; a += 32) embs[s][a] *= drop_mask(key, 0, a, M.keep, M.inv_keep);
        __syncwarp();
      }
    }
    float st = 0.f, su = 0.f;
    float* ub = my + TL.ubuf; float* vb = my + TL.vbuf; float* mb = my + TL.mbuf;
    float* tb = my + TL.tbuf; float* dmb = my + TL.dmbuf;
    const float* x1 = embs[0]; const float* x2 = embs[1];
    const float raw = pair_fwd(M, s_w, s_wp, x1, x2, ub, mb, tb, lane, &st, &su);
    const float pr = sgnn_act(raw, M.final_act, M.yeta);
    const float tgt = target_sim ? target_sim[p] : 0.f;
    float dp, lc;
    if (M.loss_mode == SGNN_LOSS_REFERENCE) {
      // (1/2B) sum_i (s_i - p)^2 summed over the B targets (F6, model_mse.py:145-151)
      dp = pr - tsum * inv_global_batch;
      lc = 0.5f * (pr * pr - 2.f * pr * tsum * inv_global_batch + tsqsum * inv_global_batch);
    } else {
      dp = (pr - tgt) * inv_global_batch;
      lc = 0.5f * (pr - tgt) * (pr - tgt) * inv_global_batch;
    }
    loss_acc += lc;
    if (lane == 0) {
      if (sim_out) sim_out[p] = pr;
      if (raw_out) raw_out[p] = raw;
    }
    if (!BWD) { __syncwarp(); continue; }
    const float graw = dp * sgnn_act_grad(pr, raw, M.final_act, M.yeta);

    // ---- pair layer backward -> dx1, dx2 (length emb_dim) ----
    float* dx1 = my + TL.dx; float* dx2 = dx1 + embp;
    const int D = M.D, K = M.K;
    if (M.pair_kind == SGNN_LAYER_DOT) {
      for (int a = lane; a < D; a += 32) { dx1[a] = graw * x2[a]; dx2[a] = graw * x1[a]; }
    } else {
      const float* U = s_w + M.u_off; const float* V = s_w + M.v_off;
      for (int k = lane; k < K; k += 32) {
        float dt;
        if (M.ntn_mode == SGNN_NTN_REFERENCE) { dt = graw * su; atomicAdd(&s_g[M.u_off + k], graw * st); }
        else { dt = graw * U[k]; atomicAdd(&s_g[M.u_off + k], graw * tb[k]); }
        const float dm = dt * sgnn_act_grad(tb[k], mb[k], M.pair_act, M.yeta);
        dmb[k] = dm;
        if (M.pair_bias) atomicAdd(&s_g[M.b_off + k], dm);
      }
      __syncwarp();
      for (int idx = lane; idx < K * 2 * D; idx += 32) {
        const int k = idx / (2 * D), c = idx - k * 2 * D;
        atomicAdd(&s_g[M.v_off + idx], dmb[k] * (c < D ? x1[c] : x2[c - D]));
      }
      for (int bk = lane; bk < DK; bk += 32) {
        const int b = bk / K, k = bk - b * K;
        const float f = x2[b] * dmb[k];
        float s = 0.f;
        for (int a = 0; a < D; ++a) {
          atomicAdd(&s_g[M.w_off + a * DK + bk], x1[a] * f);
          s = fmaf(s_w[M.w_off + a * DK + bk], x1[a], s);
        }
        vb[bk] = s;  // v[b][k] = sum_a W[a][b][k] x1[a]
      }
      __syncwarp();
      for (int a = lane; a < D; a += 32) {
        float s1 = 0.f, s2 = 0.f;
        for (int k = 0; k < K; ++k) {
          s1 = fmaf(dmb[k], V[k * 2 * D + a] + ub[a * K + k], s1);
          s2 = fmaf(dmb[k], V[k * 2 * D + D + a] + vb[a * K + k], s2);
        }
        dx1[a] = s1; dx2[a] = s2;
      }
    }
    __syncwarp();
    if (dropping && M.pair_drop) {   // gradient w.r.t. the un-dropped representations
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint32_t key = drop_layer_key(dbase, 2 * (p0 + p * pstride) + s, M.pair_ref);
        float* dx = s == 0 ? dx1 : dx2;
        for (int a = lane; a < M.emb_dim; a += 32) dx[a] *= drop_mask(key, 0, a, M.keep, M.inv_keep);
      }
      __syncwarp();
    }

    // ---- pooling + node layers backward for both arms ----
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      float* Gcur = my + TL.gbuf; float* Gnxt = Gcur + NMAX * M.ld; float* T = Gnxt + NMAX * M.ld;
      const float* Hlast = acts[s] + M.L[M.n_node - 1].out_off;
      pool_bwd<NMAX>(M, Gs[s], s_w, s_aux + M.pool_wt_off, Hlast, M.act_ld, s == 0 ? dx1 : dx2,
                     saves[s], Gcur, M.ld, T, s_g, lane);
      graph_bwd<NMAX>(M, Gs[s], s_w, s_aux, acts[s], Gcur, Gnxt, T, s_g, lane, dropping != 0, dbase, 2 * (p0 + p * pstride) + s,
                      nmask[s]);
    }
    __syncwarp();
  }
  if (lane == 0 && loss && loss_acc != 0.f) atomicAdd(loss, loss_acc);
  if (!BWD) return;
  __syncthreads();
  for (int i = threadIdx.x; i < M.n_params; i += blockDim.x) {
    const float v = s_g[i];
    if (v != 0.f) atomicAdd(&grads[i], v);
  }
}

// weight decay gradient + TF-1 Adam (models.py:28-29, 67-74; F14), grads zeroed afterwards
__global__ void adam_kernel(int n, float* __restrict__ params, float* __restrict__ grads,
                            float* __restrict__ m, float* __restrict__ v, float grad_scale, float wd,
                            float lr_t, float b1, float b2, float eps, float* __restrict__ loss,
                            long long* __restrict__ t_dev = nullptr, double lr = 0.0) {
  // t_dev: the step count lives on the device (sgnn_adam_device_step), so that the launch can sit in a replayed CUDA
  // graph: every CTA derives this step's size lr sqrt(1 - b2^t) / (1 - b1^t), t = *t_dev + 1, in the host form's double
  // arithmetic; the LAST CTA to finish advances the count (t_dev[1] counts finished CTAs: all of them have read t by then)
  __shared__ float s_lr;
  if (t_dev) {
    if (threadIdx.x == 0) {
      const double t = (double)(t_dev[0] + 1);
      s_lr = (float)(lr * sqrt(1.0 - pow((double)b2, t)) / (1.0 - pow((double)b1, t)));
    }
    __syncthreads();
    lr_t = s_lr;
  }
  float wdl = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float p = params[i];
    const float g = grads[i] * grad_scale + wd * p;
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
  if (t_dev) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(reinterpret_cast<unsigned long long*>(t_dev + 1), 1ull) == (unsigned long long)gridDim.x - 1ull) {
        t_dev[1] = 0;
        t_dev[0] += 1;
      }
    }
  }
}
