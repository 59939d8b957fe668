// Training / single-pair evaluation kernel, CTA per pair ("team" kernel).
//
// A mini-batch is 128 pairs per GPU (BASELINE configs[2]): with one warp per pair (sgnn_small.cuh,
// the first version) the whole GPU runs 128 warps, each a 42 k-instruction dependent chain -- 166 us
// per step at < 1 warp per SM.  Here a CTA of 2 x TEAM_THREADS threads owns one pair: the two arms (graphs) of the
// pair run on two TEAMS (256 threads each: 128 -> 256 took the 128-pair step from 44 to 36 us, the SimGNN chain from
// 100 to 74 us -- the kernel is latency-bound, not issue-bound) synchronised by named barriers (bar.sync 1 / 2), every loop of the
// forward and backward pass is flattened over its independent outputs (node row x feature column,
// weight row x column, ...) and strided over the team, and the pair layer (NTN / Dot) forward and
// backward use all threads of the CTA.  Same math, same order of floating-point reductions inside one output
// as the warp kernel; parameter gradients accumulate in a per-CTA shared buffer (shared-memory atomics:
// the two teams update the same weights) and are flushed with one global atomic per parameter.
// Reference math: model/Siamese/layers.py:91-118 (GCN), :136-140 (Average), :154-160 (Attention),
// :192-205 (Dense), :223-227 (Padding), :249-252 (Dot), :282-310 (NTN); loss model_mse.py:145-151.
#pragma once
#include "sgnn_small.cuh"

#ifndef SGNN_TEAM_THREADS
#define SGNN_TEAM_THREADS 256
#endif
constexpr int TEAM_THREADS = SGNN_TEAM_THREADS;
constexpr int TEAM_CTA = 2 * TEAM_THREADS;

struct Team {
  int tid;    // thread index inside the team
  int bar;    // named barrier id (1 + arm)
};
__device__ __forceinline__ void team_sync(const Team& t) {   // immediate barrier ids: the CTA reserves 3 barriers, not 16
  if (t.bar == 1) asm volatile("bar.sync 1, %0;" ::"n"(TEAM_THREADS) : "memory");
  else asm volatile("bar.sync 2, %0;" ::"n"(TEAM_THREADS) : "memory");
}

// per-arm shared-memory layout (floats), sized by the batch's max_nodes
struct TeamLayout {
  int nmax, emax;
  int csr, acts, T, G0, G1, emb, save, nmask, dx, arm_total;
  int ub, vb, mb, tb, dmb, scal, hf, hg, hcnt, pair_total;
};

__host__ __device__ inline TeamLayout team_layout(const ModelDev& M, int nmax) {
  TeamLayout t;
  t.nmax = nmax; t.emax = nmax * 8;
  int o = 0;
  t.csr = o; o += round_up4((nmax + 1) + nmax + 2 * t.emax);
  t.acts = o; o += nmax * M.act_ld;
  t.T = o; o += nmax * M.ld;
  t.G0 = o; o += nmax * M.ld;
  t.G1 = o; o += nmax * M.ld;
  t.emb = o; o += round_up4(M.emb_dim);
  t.save = o; o += round_up4(2 * M.node_dim + nmax);
  t.nmask = o; o += round_up4(nmax);
  t.dx = o; o += round_up4(M.emb_dim);
  t.arm_total = round_up4(o);
  const int DK = round_up4(M.D * M.K), Kp = round_up4(M.K);
  o = 0;
  t.ub = o; o += DK;
  t.vb = o; o += DK;
  t.mb = o; o += Kp;
  t.tb = o; o += Kp;
  t.dmb = o; o += Kp;
  t.scal = o; o += 8;
  // SimGNN-style head: activation row (features | layer outputs), pre-activation / gradient row, histogram counters
  t.hf = o; o += round_up4(M.head_floats);
  t.hg = o; o += round_up4(M.head_floats);
  t.hcnt = o; o += M.hist_bins > 0 ? 32 : 0;
  t.pair_total = o;
  return t;
}

__host__ __device__ inline size_t team_smem_floats(const ModelDev& M, int nmax) {
  const TeamLayout t = team_layout(M, nmax);
  const int wfl = round_up4(M.n_params);
  const int wpfl = M.pair_kind == SGNN_LAYER_NTN ? round_up4(M.D * M.D * M.K) : 0;
  return (size_t)2 * wfl + wpfl + round_up4(M.aux_floats) + 2 * (size_t)t.arm_total + t.pair_total;
}
// shared memory of the GARMS variant (arms in the global workspace) and the workspace floats per CTA
__host__ __device__ inline size_t team_smem_floats_garms(const ModelDev& M, int nmax) {
  return team_smem_floats(M, nmax) - 2 * (size_t)team_layout(M, nmax).arm_total;
}
__host__ __device__ inline size_t team_arm_ws_floats(const ModelDev& M, int nmax) {
  return 2 * (size_t)team_layout(M, nmax).arm_total;
}

__device__ __forceinline__ GraphS team_load_graph(const BatchDev& B, int g, float* scratch, const TeamLayout& TL,
                                                  const Team& tm) {
  int* s_rp = reinterpret_cast<int*>(scratch);
  int* s_lab = s_rp + (TL.nmax + 1);
  int* s_col = s_lab + TL.nmax;
  float* s_val = reinterpret_cast<float*>(s_col + TL.emax);
  const int base = B.node_off[g];
  const int n = B.node_off[g + 1] - base;
  const int e0 = B.rowptr[base];
  const int nnz = B.rowptr[base + n] - e0;
  for (int i = tm.tid; i <= n; i += TEAM_THREADS) s_rp[i] = B.rowptr[base + i] - e0;
  for (int i = tm.tid; i < n; i += TEAM_THREADS) s_lab[i] = B.labels[base + i];
  GraphS G;
  G.n = n; G.nnz = nnz; G.rp = s_rp; G.lab = s_lab;
  if (nnz <= TL.emax) {
    for (int e = tm.tid; e < nnz; e += TEAM_THREADS) { s_col[e] = B.col[e0 + e]; s_val[e] = B.val[e0 + e]; }
    G.col = s_col; G.val = s_val;
  } else {
    G.col = B.col + e0; G.val = B.val + e0;
  }
  team_sync(tm);
  return G;
}

// Out[i][j] = act( sum_e A_hat[i,e] * T[col e][j] + b[j] ), flattened over (i, j)
__device__ __forceinline__ void team_aggregate(const GraphS& G, const float* T, int ldt, int dout, const float* bias,
                                               int act, float* Out, int ldo, const Team& tm) {
  for (int idx = tm.tid; idx < G.n * dout; idx += TEAM_THREADS) {
    const int i = idx / dout, j = idx - i * dout;
    float s = 0.f;
    for (int e = G.rp[i]; e < G.rp[i + 1]; ++e) s = fmaf(G.val[e], T[G.col[e] * ldt + j], s);
    Out[i * ldo + j] = sgnn_act(s + (bias ? bias[j] : 0.f), act, 0.f);
  }
}

// Out[i][j] = sum_k H[i][k] * W[k*dout + j] through epi(i, j, v), flattened over (i, j).  A thread carries TWO
// outputs per pass (idx and idx + TEAM_THREADS): the k-chains are serial by definition (fixed summation order),
// so the second chain is what fills the FMA latency.
template <typename Epi>
__device__ __forceinline__ void team_gemm(const float* H, int ldh, int n, int din, const float* W, int dout,
                                          const Team& tm, Epi epi) {
  const int total = n * dout;
  for (int idx = tm.tid; idx < total; idx += 2 * TEAM_THREADS) {
    const int idx2 = idx + TEAM_THREADS;
    const bool two = idx2 < total;
    const int i = idx / dout, j = idx - i * dout;
    const int i2 = two ? idx2 / dout : i, j2 = two ? idx2 - i2 * dout : j;
    const float* h = H + i * ldh;
    const float* h2 = H + i2 * ldh;
    float s = 0.f, s2 = 0.f;
    for (int k = 0; k < din; ++k) {
      s = fmaf(h[k], W[k * dout + j], s);
      s2 = fmaf(h2[k], W[k * dout + j2], s2);
    }
    epi(i, j, s);
    if (two) epi(i2, j2, s2);
  }
}

__device__ __forceinline__ void team_graph_fwd(const ModelDev& M, const GraphS& G, const float* W, float* arm,
                                               const TeamLayout& TL, const Team& tm, bool dropping, uint32_t dbase,
                                               int slot) {
  float* acts = arm + TL.acts;
  float* T = arm + TL.T;
  float* nmask = arm + TL.nmask;
  const int n = G.n, ld = M.ld, ald = M.act_ld;
  float* Hin = nullptr;
  for (int l = 0; l < M.n_node; ++l) {
    const LayerDev& L = M.L[l];
    float* Hout = acts + L.out_off;
    const float* w = W + L.w_off;
    const float* b = L.bias ? W + L.b_off : nullptr;
    const bool drop = dropping && L.drop;
    const uint32_t dkey = drop ? drop_layer_key(dbase, slot, L.ref) : 0u;
    if (drop && l > 0) {   // mask the layer input in place: the saved activation is the masked one
      for (int idx = tm.tid; idx < n * L.din; idx += TEAM_THREADS) {
        const int i = idx / L.din, k = idx - i * L.din;
        Hin[i * ald + k] *= drop_mask(dkey, i, k, M.keep, M.inv_keep);
      }
      team_sync(tm);
    }
    if (L.kind == SGNN_LAYER_GCN && l == 0) {
      if (dropping) {   // sparse_dropout: one draw per node of the one-hot X (layers.py:332-338)
        for (int i = tm.tid; i < n; i += TEAM_THREADS) nmask[i] = drop ? drop_mask(dkey, i, 0, M.keep, M.inv_keep) : 1.f;
        team_sync(tm);
      }
      for (int idx = tm.tid; idx < n * L.dout; idx += TEAM_THREADS) {
        const int i = idx / L.dout, j = idx - i * L.dout;
        float s = 0.f;
        for (int e = G.rp[i]; e < G.rp[i + 1]; ++e) {
          const int c = G.col[e];
          const float v = dropping ? G.val[e] * nmask[c] : G.val[e];
          s = fmaf(v, w[G.lab[c] * L.dout + j], s);
        }
        Hout[i * ald + j] = sgnn_act(s + (b ? b[j] : 0.f), L.act, 0.f);
      }
    } else if (L.kind == SGNN_LAYER_GCN) {
      team_gemm(Hin, ald, n, L.din, w, L.dout, tm, [&](int i, int j, float v) { T[i * ld + j] = v; });
      team_sync(tm);
      team_aggregate(G, T, ld, L.dout, b, L.act, Hout, ald, tm);
    } else {   // Dense
      const int act = L.act;
      team_gemm(Hin, ald, n, L.din, w, L.dout, tm,
                [&](int i, int j, float v) { Hout[i * ald + j] = sgnn_act(v + (b ? b[j] : 0.f), act, 0.f); });
    }
    team_sync(tm);
    Hin = Hout;
  }
  // ---- pooling (layers.py:136-160, 223-227) ----
  const int d = M.node_dim;
  float* emb = arm + TL.emb;
  const float* H = Hin;
  if (M.pool_kind == SGNN_LAYER_PADDING) {
    for (int idx = tm.tid; idx < M.emb_dim; idx += TEAM_THREADS) {
      const int p = idx / d, c = idx - p * d;
      emb[idx] = p < n ? H[p * ald + c] : M.pool_fparam;
    }
  } else if (M.pool_kind == SGNN_LAYER_AVERAGE) {
    for (int j = tm.tid; j < d; j += TEAM_THREADS) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s += H[i * ald + j];
      emb[j] = s / (float)n;
    }
  } else {
    float* mean = arm + TL.save; float* c = mean + d; float* att = c + d;
    const float* wa = W + M.pool_w_off;
    for (int j = tm.tid; j < d; j += TEAM_THREADS) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s += H[i * ald + j];
      mean[j] = s / (float)n;
    }
    team_sync(tm);
    for (int j = tm.tid; j < d; j += TEAM_THREADS) {
      float s = 0.f;
      for (int k = 0; k < d; ++k) s = fmaf(mean[k], wa[k * d + j], s);
      c[j] = tanhf(s);
    }
    team_sync(tm);
    for (int i = tm.tid; i < n; i += TEAM_THREADS) {
      float s = 0.f;
      for (int k = 0; k < d; ++k) s = fmaf(H[i * ald + k], c[k], s);
      att[i] = 1.f / (1.f + expf(-s));
    }
    team_sync(tm);
    for (int j = tm.tid; j < d; j += TEAM_THREADS) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s = fmaf(att[i], H[i * ald + j], s);
      emb[j] = s;
    }
  }
  if (dropping && M.pair_drop) {   // dropout on the pair layer's inputs (layers.py:287-288)
    team_sync(tm);
    const uint32_t key = drop_layer_key(dbase, slot, M.pair_ref);
    for (int a = tm.tid; a < M.emb_dim; a += TEAM_THREADS) emb[a] *= drop_mask(key, 0, a, M.keep, M.inv_keep);
  }
}

// Backward of pooling + node layers of one arm.  demb: dL/d(emb) (already un-dropped).
__device__ __forceinline__ void team_graph_bwd(const ModelDev& M, const GraphS& G, const float* W, const float* s_aux,
                                               float* arm, const TeamLayout& TL, const float* demb, float* s_g,
                                               const Team& tm, bool dropping, uint32_t dbase, int slot) {
  const int n = G.n, ld = M.ld, ald = M.act_ld, d = M.node_dim;
  const float* acts = arm + TL.acts;
  float* Gcur = arm + TL.G0;
  float* Gnxt = arm + TL.G1;
  float* T = arm + TL.T;
  const float* nmask = dropping ? arm + TL.nmask : nullptr;
  const float* H = acts + M.L[M.n_node - 1].out_off;
  // ---- pooling backward -> Gcur [n][d] ----
  if (M.pool_kind == SGNN_LAYER_PADDING) {
    for (int idx = tm.tid; idx < n * d; idx += TEAM_THREADS) {
      const int p = idx / d, c = idx - p * d;
      Gcur[p * ld + c] = demb[idx];
    }
  } else if (M.pool_kind == SGNN_LAYER_AVERAGE) {
    const float inv = 1.f / (float)n;
    for (int idx = tm.tid; idx < n * d; idx += TEAM_THREADS) {
      const int i = idx / d, j = idx - i * d;
      Gcur[i * ld + j] = demb[j] * inv;
    }
  } else {
    const float* mean = arm + TL.save; const float* c = mean + d; const float* att = c + d;
    const float* Wt_att = s_aux + M.pool_wt_off;
    float* dz = T; float* dq = T + TL.nmax; float* dmean = dq + d;   // T is free here (needs nmax + 2 d <= nmax * ld)
    for (int i = tm.tid; i < n; i += TEAM_THREADS) {
      float s = 0.f;
      for (int j = 0; j < d; ++j) s = fmaf(demb[j], H[i * ald + j], s);
      dz[i] = s * att[i] * (1.f - att[i]);
    }
    team_sync(tm);
    for (int k = tm.tid; k < d; k += TEAM_THREADS) {
      float s = 0.f;
      for (int i = 0; i < n; ++i) s = fmaf(dz[i], H[i * ald + k], s);
      dq[k] = s * (1.f - c[k] * c[k]);
    }
    team_sync(tm);
    for (int idx = tm.tid; idx < d * d; idx += TEAM_THREADS) {
      const int k = idx / d, j = idx - k * d;
      atomicAdd(&s_g[M.pool_w_off + idx], mean[k] * dq[j]);
    }
    for (int k = tm.tid; k < d; k += TEAM_THREADS) {
      float s = 0.f;
      for (int j = 0; j < d; ++j) s = fmaf(Wt_att[j * d + k], dq[j], s);
      dmean[k] = s / (float)n;
    }
    team_sync(tm);
    for (int idx = tm.tid; idx < n * d; idx += TEAM_THREADS) {
      const int i = idx / d, j = idx - i * d;
      Gcur[i * ld + j] = att[i] * demb[j] + dz[i] * c[j] + dmean[j];
    }
  }
  team_sync(tm);
  // ---- node layers, last to first ----
  for (int l = M.n_node - 1; l >= 0; --l) {
    const LayerDev& L = M.L[l];
    const float* Hout = acts + L.out_off;
    const int dout = L.dout;
    // If the NEXT node layer dropped its input, the saved Hout is y * mask (mask in {0, 1/keep}) and the incoming
    // gradient is w.r.t. the masked value: dL/dy = g * mask, and y = saved * keep where kept.
    const bool masked = dropping && l + 1 < M.n_node && M.L[l + 1].drop;
    const uint32_t mkey = masked ? drop_layer_key(dbase, slot, M.L[l + 1].ref) : 0u;
    for (int idx = tm.tid; idx < n * dout; idx += TEAM_THREADS) {   // dZ = dOut * act'(out)
      const int i = idx / dout, j = idx - i * dout;
      float y = Hout[i * ald + j];
      float g = Gcur[i * ld + j];
      if (masked) { g *= drop_mask(mkey, i, j, M.keep, M.inv_keep); y *= M.keep; }
      Gcur[i * ld + j] = g * sgnn_act_grad(y, 0.f, L.act, 0.f);
    }
    team_sync(tm);
    if (L.bias)
      for (int j = tm.tid; j < dout; j += TEAM_THREADS) {
        float sb = 0.f;
        for (int i = 0; i < n; ++i) sb += Gcur[i * ld + j];
        atomicAdd(&s_g[L.b_off + j], sb);
      }
    const float* dXW = Gcur;
    if (L.kind == SGNN_LAYER_GCN) {   // A_hat is symmetric: dXW = A_hat . dZ
      team_aggregate(G, Gcur, ld, dout, nullptr, SGNN_ACT_IDENTITY, T, ld, tm);
      team_sync(tm);
      dXW = T;
    }
    if (l == 0) {   // one-hot input: dW[label_i][:] += mask_i * dXW[i][:]
      for (int idx = tm.tid; idx < n * dout; idx += TEAM_THREADS) {
        const int i = idx / dout, j = idx - i * dout;
        const float v = dXW[i * ld + j];
        atomicAdd(&s_g[L.w_off + G.lab[i] * dout + j], nmask ? nmask[i] * v : v);
      }
    } else {
      const float* Hin = acts + M.L[l - 1].out_off;
      const int din = L.din;
      // two outputs per thread and pass (see team_gemm): independent chains hide the FMA latency
      for (int idx = tm.tid; idx < din * dout; idx += 2 * TEAM_THREADS) {   // dW[k][j] = sum_i Hin[i][k] dXW[i][j]
        const int idx2 = idx + TEAM_THREADS;
        const bool two = idx2 < din * dout;
        const int k = idx / dout, j = idx - k * dout;
        const int k2 = two ? idx2 / dout : k, j2 = two ? idx2 - k2 * dout : j;
        float s = 0.f, s2 = 0.f;
        for (int i = 0; i < n; ++i) {
          s = fmaf(Hin[i * ald + k], dXW[i * ld + j], s);
          s2 = fmaf(Hin[i * ald + k2], dXW[i * ld + j2], s2);
        }
        atomicAdd(&s_g[L.w_off + idx], s);
        if (two) atomicAdd(&s_g[L.w_off + idx2], s2);
      }
      const float* Wt = s_aux + L.wt_off;   // [dout][din]
      for (int idx = tm.tid; idx < n * din; idx += 2 * TEAM_THREADS) {      // dHin[i][k] = sum_j dXW[i][j] W[k][j]
        const int idx2 = idx + TEAM_THREADS;
        const bool two = idx2 < n * din;
        const int i = idx / din, k = idx - i * din;
        const int i2 = two ? idx2 / din : i, k2 = two ? idx2 - i2 * din : k;
        float s = 0.f, s2 = 0.f;
        for (int j = 0; j < dout; ++j) {
          s = fmaf(dXW[i * ld + j], Wt[j * din + k], s);
          s2 = fmaf(dXW[i2 * ld + j], Wt[j * din + k2], s2);
        }
        Gnxt[i * ld + k] = s;
        if (two) Gnxt[i2 * ld + k2] = s2;
      }
      float* sw = Gcur; Gcur = Gnxt; Gnxt = sw;
    }
    team_sync(tm);
  }
}

// ---- SimGNN-style head (sgnn.h; build-defined): histogram features + Dense layers on the pair row ----
// All TEAM_CTA threads.  On entry hf[0..K) = U_k * inneract(m_k), hcnt zeroed.  Leaves the raw score in scal[0],
// the layer outputs in hf and the pre-activations in hg (read by the backward pass).
__device__ __forceinline__ void team_head_fwd(const ModelDev& M, const float* W, const float* arms, const TeamLayout& TL,
                                              int n_mine, float* hf, float* hg, int* hcnt, float* scal, bool dropping,
                                              uint32_t dbase, int p) {
  const int tid = threadIdx.x;
  const int K = M.K, bins = M.hist_bins;
  if (bins > 0) {
    // node counts of the two arms (each team knows its own)
    if (tid == 0) scal[3] = __int_as_float(n_mine);
    if (tid == TEAM_THREADS) scal[4] = __int_as_float(n_mine);
    __syncthreads();
    const int n1 = __float_as_int(scal[3]), n2 = __float_as_int(scal[4]);
    const int d = M.node_dim, ald = M.act_ld;
    const float* H1 = arms + TL.acts + M.L[M.n_node - 1].out_off;
    const float* H2 = arms + TL.arm_total + TL.acts + M.L[M.n_node - 1].out_off;
    const int cells = n1 * n2;
    // same arithmetic as pair_histogram_kernel (sgnn_large.cuh): sequential fmaf dot, sigmoid, floor(s * bins)
    for (int c = tid; c < cells; c += TEAM_CTA) {
      const int i = c / n2, j = c - i * n2;
      const float* x = H1 + i * ald;
      const float* y = H2 + j * ald;
      float s = 0.f;
      for (int k = 0; k < d; ++k) s = fmaf(x[k], y[k], s);
      atomicAdd(&hcnt[min(bins - 1, (int)floorf(sgnn_hist_sigmoid(s) * (float)bins))], 1);
    }
    __syncthreads();
    if (tid < bins) hf[K + tid] = cells > 0 ? (float)hcnt[tid] * (1.f / (float)cells) : 0.f;
    __syncthreads();
  }
  int in_off = 0;
  for (int l = 0; l < M.n_head; ++l) {
    const LayerDev& L = M.Hd[l];
    float* in = hf + in_off;
    if (dropping && L.drop) {   // Dense dropout on its input (layers.py:196); the saved input is the masked one
      const uint32_t key = drop_layer_key(dbase, 2 * p, L.ref);
      for (int k = tid; k < L.din; k += TEAM_CTA) in[k] *= drop_mask(key, 0, k, M.keep, M.inv_keep);
      __syncthreads();
    }
    const float* w = W + L.w_off;
    for (int j = tid; j < L.dout; j += TEAM_CTA) {
      float s = 0.f;
      for (int k = 0; k < L.din; ++k) s = fmaf(in[k], w[k * L.dout + j], s);
      s += L.bias ? W[L.b_off + j] : 0.f;
      hg[L.out_off + j] = s;
      hf[L.out_off + j] = sgnn_act(s, L.act, M.yeta);
    }
    __syncthreads();
    in_off = L.out_off;
  }
  if (tid == 0) scal[0] = hf[in_off];
}

// graw = dL/d(raw).  Accumulates the head's parameter gradients into s_g and leaves dL/d(f_k) in hg[0..K).
__device__ __forceinline__ void team_head_bwd(const ModelDev& M, const float* W, float* s_g, const float* hf, float* hg,
                                              float graw, bool dropping, uint32_t dbase, int p) {
  const int tid = threadIdx.x;
  {
    const LayerDev& L = M.Hd[M.n_head - 1];
    if (tid == 0) hg[L.out_off] = graw * sgnn_act_grad(hf[L.out_off], hg[L.out_off], L.act, M.yeta);
    __syncthreads();
  }
  for (int l = M.n_head - 1; l >= 0; --l) {
    const LayerDev& L = M.Hd[l];
    const int in_off = l > 0 ? M.Hd[l - 1].out_off : 0;
    const float* in = hf + in_off;
    const float* dz = hg + L.out_off;
    const float* w = W + L.w_off;
    if (L.bias)
      for (int j = tid; j < L.dout; j += TEAM_CTA) atomicAdd(&s_g[L.b_off + j], dz[j]);
    for (int idx = tid; idx < L.din * L.dout; idx += TEAM_CTA) {
      const int k = idx / L.dout, j = idx - k * L.dout;
      atomicAdd(&s_g[L.w_off + idx], in[k] * dz[j]);
    }
    const bool masked = dropping && L.drop;
    const uint32_t key = masked ? drop_layer_key(dbase, 2 * p, L.ref) : 0u;
    for (int k = tid; k < L.din; k += TEAM_CTA) {
      float s = 0.f;
      for (int j = 0; j < L.dout; ++j) s = fmaf(w[k * L.dout + j], dz[j], s);
      if (masked) s *= drop_mask(key, 0, k, M.keep, M.inv_keep);
      if (l > 0) {
        const LayerDev& Lp = M.Hd[l - 1];
        // the saved output of layer l-1 is masked in place if layer l dropped its input: y = saved * keep where kept
        float y = hf[in_off + k];
        if (masked) y *= M.keep;
        s *= sgnn_act_grad(y, hg[in_off + k], Lp.act, M.yeta);
      }
      hg[in_off + k] = s;
    }
    __syncthreads();
  }
}

// MINB: CTAs per SM the register budget is cut for -- 1 (128 registers) for batches of at most one pair per SM, where
// only the latency of a pair counts; 2 (64 registers, no spills) for larger batches, where two pairs per SM overlap
// their barrier-separated phases
// GARMS: the two arms' buffers (CSR slice, saved activations, gradient rows: ~160 floats per node for GCN32-GCN16) live
// in a GLOBAL workspace slice of this CTA instead of shared memory -- Web-shaped graphs of hundreds of nodes
// (configs[4]) do not fit next to the weights.  Same code, same barriers (bar.sync orders global accesses of the CTA's
// threads as well); the slice stays in L1 / L2.
template <bool BWD, int MINB, bool GARMS = false>
__global__ void __launch_bounds__(TEAM_CTA, MINB)
train_team_kernel(const __grid_constant__ ModelDev M, const BatchDev B, int nmax, const float* __restrict__ params,
                  const int32_t* __restrict__ g1, const int32_t* __restrict__ g2,
                  const float* __restrict__ target_sim, int n_pairs, float inv_global_batch, float tsum, float tsqsum,
                  float* __restrict__ grads, float* __restrict__ loss, float* __restrict__ raw_out,
                  float* __restrict__ sim_out, int dropping, uint32_t dbase, int p0, int pstride,
                  float* __restrict__ arm_ws = nullptr) {
  // local pair p is pair p0 + p * pstride of the GLOBAL batch (data-parallel ranks: the dropout masks are keyed on it)
  extern __shared__ __align__(16) float smem[];
  const TeamLayout TL = team_layout(M, nmax);
  const int wfl = round_up4(M.n_params);
  const int D = M.D, K = M.K, DK = D * K;
  float* s_w = smem;
  float* s_g = s_w + wfl;
  float* s_wp = s_g + wfl;
  const int wpfl = M.pair_kind == SGNN_LAYER_NTN ? round_up4(D * DK) : 0;
  float* s_aux = s_wp + wpfl;
  float* arms = GARMS ? arm_ws + (size_t)blockIdx.x * 2 * TL.arm_total : s_aux + round_up4(M.aux_floats);
  float* pairb = s_aux + round_up4(M.aux_floats) + (GARMS ? 0 : 2 * TL.arm_total);
  stage_weights(M, params, s_w, s_wp);
  if (BWD)
    for (int i = threadIdx.x; i < wfl; i += blockDim.x) s_g[i] = 0.f;
  __syncthreads();
  if (BWD) stage_transposed(M, s_w, s_aux);
  __syncthreads();

  const int arm_id = threadIdx.x / TEAM_THREADS;
  Team tm;
  tm.tid = threadIdx.x - arm_id * TEAM_THREADS;
  tm.bar = 1 + arm_id;
  float* arm = arms + arm_id * TL.arm_total;
  float* ub = pairb + TL.ub; float* vb = pairb + TL.vb; float* mb = pairb + TL.mb;
  float* tb = pairb + TL.tb; float* dmb = pairb + TL.dmb; float* scal = pairb + TL.scal;
  float* hf = pairb + TL.hf; float* hg = pairb + TL.hg; int* hcnt = reinterpret_cast<int*>(pairb + TL.hcnt);
  const float* x1 = arms + TL.emb;
  const float* x2 = arms + TL.arm_total + TL.emb;
  float* dx1 = arms + TL.dx;
  float* dx2 = arms + TL.arm_total + TL.dx;
  const float* V = s_w + M.v_off;
  const float* U = s_w + M.u_off;
  const int tid = threadIdx.x;
  float loss_acc = 0.f;

  for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
    const int gid = arm_id == 0 ? g1[p] : g2[p];
    const int slot = 2 * (p0 + p * pstride) + arm_id;
    const GraphS G = team_load_graph(B, gid, arm + TL.csr, TL, tm);
    team_graph_fwd(M, G, s_w, arm, TL, tm, dropping != 0, dbase, slot);
    __syncthreads();
    // ---- pair layer forward (layers.py:249-252, 282-310), all threads of the CTA ----
    if (M.pair_kind == SGNN_LAYER_DOT) {
      if (tid < 32) {
        float s = 0.f;
        for (int a = tid; a < D; a += 32) s = fmaf(x1[a], x2[a], s);
        s = warp_sum(s);
        if (tid == 0) scal[0] = s;
      }
    } else {
      for (int ak = tid; ak < DK; ak += TEAM_CTA) {
        float s = 0.f;
        for (int b = 0; b < D; ++b) s = fmaf(s_wp[b * DK + ak], x2[b], s);
        ub[ak] = s;
      }
      __syncthreads();
      if (tid < 32) {   // same reduction order as pair_fwd of the warp kernel
        float st = 0.f, su = 0.f, sut = 0.f;
        for (int k = tid; k < K; k += 32) {
          float s = M.pair_bias ? s_w[M.b_off + k] : 0.f;
          for (int a = 0; a < D; ++a) s = fmaf(x1[a], ub[a * K + k], s);
          for (int c = 0; c < D; ++c) {
            s = fmaf(V[k * 2 * D + c], x1[c], s);
            s = fmaf(V[k * 2 * D + D + c], x2[c], s);
          }
          const float t = sgnn_act(s, M.pair_act, M.yeta);
          mb[k] = s; tb[k] = t;
          st += t; su += U[k]; sut = fmaf(U[k], t, sut);
          if (M.n_head > 0) hf[k] = U[k] * t;   // head: the NTN emits its K-vector (sgnn.h)
        }
        st = warp_sum(st); su = warp_sum(su); sut = warp_sum(sut);
        if (tid == 0) {
          scal[0] = M.ntn_mode == SGNN_NTN_REFERENCE ? su * st : sut;   // F5: layers.py:305-308
          scal[1] = st; scal[2] = su;
        }
      } else if (M.hist_bins > 0 && tid >= 32 && tid < 64) {
        hcnt[tid - 32] = 0;
      }
      if (M.n_head > 0) {
        __syncthreads();
        team_head_fwd(M, s_w, arms, TL, G.n, hf, hg, hcnt, scal, dropping != 0, dbase, p0 + p * pstride);
      }
    }
    __syncthreads();
    const float raw = scal[0], st = scal[1], su = scal[2];
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
    if (tid == 0) {
      loss_acc += lc;
      if (sim_out) sim_out[p] = pr;
      if (raw_out) raw_out[p] = raw;
    }
    if (!BWD) { __syncthreads(); continue; }
    const float graw = dp * sgnn_act_grad(pr, raw, M.final_act, M.yeta);

    // ---- pair layer backward -> dx1, dx2 ----
    if (M.pair_kind == SGNN_LAYER_DOT) {
      for (int a = tid; a < D; a += TEAM_CTA) { dx1[a] = graw * x2[a]; dx2[a] = graw * x1[a]; }
    } else {
      if (M.n_head > 0) {   // head backward: leaves dL/d(f_k) in hg[0..K)
        team_head_bwd(M, s_w, s_g, hf, hg, graw, dropping != 0, dbase, p0 + p * pstride);
        __syncthreads();
      }
      for (int k = tid; k < K; k += TEAM_CTA) {
        float dt;
        if (M.n_head > 0) { dt = hg[k] * U[k]; atomicAdd(&s_g[M.u_off + k], hg[k] * tb[k]); }
        else if (M.ntn_mode == SGNN_NTN_REFERENCE) { dt = graw * su; atomicAdd(&s_g[M.u_off + k], graw * st); }
        else { dt = graw * U[k]; atomicAdd(&s_g[M.u_off + k], graw * tb[k]); }
        const float dm = dt * sgnn_act_grad(tb[k], mb[k], M.pair_act, M.yeta);
        dmb[k] = dm;
        if (M.pair_bias) atomicAdd(&s_g[M.b_off + k], dm);
      }
      __syncthreads();
      for (int idx = tid; idx < K * 2 * D; idx += TEAM_CTA) {
        const int k = idx / (2 * D), c = idx - k * 2 * D;
        atomicAdd(&s_g[M.v_off + idx], dmb[k] * (c < D ? x1[c] : x2[c - D]));
      }
      for (int idx = tid; idx < D * DK; idx += TEAM_CTA) {   // dW[a][b][k] = x1[a] x2[b] dm[k]
        const int a = idx / DK, bk = idx - a * DK, b = bk / K, k = bk - b * K;
        atomicAdd(&s_g[M.w_off + idx], x1[a] * (x2[b] * dmb[k]));
      }
      for (int bk = tid; bk < DK; bk += TEAM_CTA) {          // v[b][k] = sum_a W[a][b][k] x1[a]
        float s = 0.f;
        for (int a = 0; a < D; ++a) s = fmaf(s_w[M.w_off + a * DK + bk], x1[a], s);
        vb[bk] = s;
      }
      __syncthreads();
      for (int idx = tid; idx < 2 * D; idx += TEAM_CTA) {
        const int a = idx < D ? idx : idx - D;
        float s = 0.f;
        if (idx < D) {
          for (int k = 0; k < K; ++k) s = fmaf(dmb[k], V[k * 2 * D + a] + ub[a * K + k], s);
          dx1[a] = s;
        } else {
          for (int k = 0; k < K; ++k) s = fmaf(dmb[k], V[k * 2 * D + D + a] + vb[a * K + k], s);
          dx2[a] = s;
        }
      }
    }
    __syncthreads();
    float* dx = arm + TL.dx;
    if (dropping && M.pair_drop) {   // gradient w.r.t. the un-dropped representations
      const uint32_t key = drop_layer_key(dbase, slot, M.pair_ref);
      for (int a = tm.tid; a < M.emb_dim; a += TEAM_THREADS) dx[a] *= drop_mask(key, 0, a, M.keep, M.inv_keep);
      team_sync(tm);
    }
    team_graph_bwd(M, G, s_w, s_aux, arm, TL, dx, s_g, tm, dropping != 0, dbase, slot);
    __syncthreads();
  }
  if (tid == 0 && loss && loss_acc != 0.f) atomicAdd(loss, loss_acc);
  if (!BWD) return;
  __syncthreads();
  for (int i = threadIdx.x; i < M.n_params; i += blockDim.x) {
    const float v = s_g[i];
    if (v != 0.f) atomicAdd(&grads[i], v);
  }
}
