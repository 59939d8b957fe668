// Shared device/host definitions for libsgnn (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/sgnn.h"

#define SGNN_WARP 32

struct LayerDev {
  int kind, din, dout, act, bias, w_off, b_off, aux;
  float fparam;
  int out_off;   // column offset of this layer's output inside a saved-activation row (train)
  int wt_off;    // offset of the transposed copy inside the "aux weights" smem region
  int drop;      // dropout on this layer's INPUT (spec dropout=True and FLAGS.dropout > 0)
  int ref;       // index of the layer in the reference's layer list (dropout stream key)
};

// Flattened description of one model (SgnnConfig after validation).  Passed by value as a
// __grid_constant__ kernel parameter.
struct ModelDev {
  int n_node;                    // node-level layers (GCN / Dense) before the pooling layer
  LayerDev L[SGNN_MAX_LAYERS];
  int pool_kind, pool_aux, pool_w_off, pool_wt_off;
  float pool_fparam;
  int node_dim;                  // width of the last node-level output
  int emb_dim;                   // flattened representation consumed by the pair layer
  int pair_kind, D, K, w_off, v_off, u_off, b_off, pair_bias, pair_act, wp_off;
  int final_act, ntn_mode, loss_mode;
  float yeta;
  int n_params, input_dim;
  int ld;                        // row stride (floats) of per-warp activation buffers
  int act_ld;                    // row stride of the saved-activation rows (train)
  int aux_floats;                // size of the transposed/permuted weight region (train)
  int pair_drop, pair_ref;       // dropout on the pair layer's two inputs (layers.py:287-288)
  float keep, inv_keep;          // keep_prob = 1 - FLAGS.dropout, and 1 / keep_prob
  // SimGNN-style head (sgnn.h): pair-level layers after the NTN
  int n_head;                    // Dense layers after the NTN (0 = the NTN's scalar is the raw score)
  int hist_bins;                 // Histogram layer right after the NTN (0 = none)
  int head_in;                   // K + hist_bins
  int head_floats;               // head_in + sum of the head layers' widths, each rounded up to 4
  LayerDev Hd[SGNN_MAX_HEAD];    // out_off = offset of the layer's output inside the head activation row
};

struct BatchDev {
  int n_graphs;
  const int32_t* node_off;
  const int32_t* labels;
  const int32_t* rowptr;
  const int32_t* col;
  const float* val;
};

__device__ __forceinline__ float sgnn_act(float x, int act, float yeta) {
  switch (act) {
    case SGNN_ACT_RELU: return fmaxf(x, 0.f);
    case SGNN_ACT_SIGMOID: return 1.f / (1.f + expf(-x));
    case SGNN_ACT_TANH: return tanhf(x);
    case SGNN_ACT_SIM_KERNEL: return expf(-yeta * x * x);
    default: return x;
  }
}

// derivative of the activation expressed with its OUTPUT y (and input x for sim_kernel)
__device__ __forceinline__ float sgnn_act_grad(float y, float x, int act, float yeta) {
  switch (act) {
    case SGNN_ACT_RELU: return y > 0.f ? 1.f : 0.f;
    case SGNN_ACT_SIGMOID: return y * (1.f - y);
    case SGNN_ACT_TANH: return 1.f - y * y;
    case SGNN_ACT_SIM_KERNEL: return -2.f * yeta * x * y;
    default: return 1.f;
  }
}

// ---- dropout stream (build-defined counter-based hash; oracle/siamese_oracle.py Dropout) ----
// keep iff floor(keep_prob + u) == 1, kept values scaled by 1/keep_prob (layers.py:332-338, tf.nn.dropout)
__host__ __device__ __forceinline__ uint32_t sgnn_fmix32(uint32_t x) {
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t drop_layer_key(uint32_t base, int slot, int layer) {
  const uint32_t sk = sgnn_fmix32(base ^ ((uint32_t)slot * 0x9E3779B1u + 0x7F4A7C15u));
  return sgnn_fmix32(sk ^ ((uint32_t)(layer + 1) * 0x85EBCA77u));
}
__device__ __forceinline__ float drop_mask(uint32_t layer_key, int row, int col, float keep, float inv_keep) {
  const uint32_t h = sgnn_fmix32(layer_key + (uint32_t)(row * 131 + col) * 0x9E3779B9u);
  const float u = (float)(h >> 8) * 5.9604644775390625e-8f;
  return floorf(keep + u) >= 1.f ? inv_keep : 0.f;
}

// sigmoid of a node-node score in the histogram layer: ONE definition for the three kernels that bin scores
// (pair_histogram_kernel, the training kernel, the all-pairs head kernel), so they agree on every bin index.
// ex2.approx.ftz + rcp.approx.ftz (4 instructions): |error| < 3e-7, far inside the 1e-5 bin-edge band the parity tests exclude.
__device__ __forceinline__ float sgnn_hist_sigmoid(float s) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(s * -1.4426950408889634f));   // exp(-s); +inf for s << 0
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));                    // 1 / inf = 0
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

static inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ __forceinline__ int round_up4(int x) { return (x + 3) & ~3; }
