// All-pairs query x database scoring: the m x n loop of test() (model/Siamese/train.py:57-61)
// as dense work.  The NTN bilinear form factorises per query:
//   m_k(q, g) = x_q W_k x_g^T + V_k [x_q; x_g] + b_k
//             = (x_q W_k + V_k[D:2D]) . x_g + (V_k[0:D] . x_q + b_k)
//             =  P'[q][k][:] . x_g        +  C[q][k]
// so the per-pair cost drops from 2 D^2 K to 2 D K flops and the pair stage becomes the GEMM
// [(Q K) x D] . [D x M] with an activation + reduce-over-k epilogue (layers.py:282-310).
#pragma once
#include "sgnn_common.cuh"

// P[q][k][b] (row stride DEP), C[q][k].  One warp per query.  Dot: K_eff = 1, P = x_q.
__global__ void __launch_bounds__(128)
allpairs_prep_kernel(const __grid_constant__ ModelDev M, const float* __restrict__ params,
                     const float* __restrict__ emb_q, int m, int DEP, float* __restrict__ P,
                     float* __restrict__ C) {
  const int lane = threadIdx.x & 31;
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (q >= m) return;
  const int D = M.emb_dim;
  const float* x = emb_q + (size_t)q * D;
  if (M.pair_kind == SGNN_LAYER_DOT) {
    for (int b = lane; b < DEP; b += 32) P[(size_t)q * DEP + b] = b < D ? x[b] : 0.f;
    if (lane == 0) C[q] = 0.f;
    return;
  }
  const int K = M.K, DK = D * K;
  const float* W = params + M.w_off;
  const float* V = params + M.v_off;
  for (int bk = lane; bk < DK; bk += 32) {
    const int b = bk / K, k = bk - b * K;
    float s = V[k * 2 * D + D + b];
    for (int a = 0; a < D; ++a) s = fmaf(x[a], W[a * DK + bk], s);
    P[((size_t)q * K + k) * DEP + b] = s;
  }
  for (int i = lane; i < K * (DEP - D); i += 32) {  // zero the K padding
    const int k = i / (DEP - D), b = D + i - k * (DEP - D);
    P[((size_t)q * K + k) * DEP + b] = 0.f;
  }
  for (int k = lane; k < K; k += 32) {
    float s = M.pair_bias ? params[M.b_off + k] : 0.f;
    for (int a = 0; a < D; ++a) s = fmaf(V[k * 2 * D + a], x[a], s);
    C[(size_t)q * K + k] = s;
  }
}

// fp32 CUDA-core scorer (parity mode).  Thread = one database graph (embedding in
// registers), CTA loops over a block of queries whose P'/C tiles are staged in shared
// memory and read as warp-uniform broadcasts.
template <int DEP, int QCHUNK>
__global__ void __launch_bounds__(128)
allpairs_fp32_kernel(const __grid_constant__ ModelDev M, const float* __restrict__ params,
                     const float* __restrict__ P, const float* __restrict__ C,
                     const float* __restrict__ emb_db, int m, int n, int q_per_block,
                     float* __restrict__ sim, int64_t ld) {
  extern __shared__ __align__(16) float smem[];
  const bool dot = M.pair_kind == SGNN_LAYER_DOT;
  const int K = dot ? 1 : M.K;
  float* sP = smem;                       // [QCHUNK][K][DEP]
  float* sC = sP + QCHUNK * K * DEP;      // [QCHUNK][K]
  float* sU = sC + QCHUNK * K;            // [K]
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int D = M.emb_dim;
  float x[DEP];
#pragma unroll
  for (int b = 0; b < DEP; ++b) x[b] = (g < n && b < D) ? emb_db[(size_t)g * D + b] : 0.f;
  float su = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) sU[k] = dot ? 1.f : params[M.u_off + k];
  __syncthreads();
  for (int k = 0; k < K; ++k) su += sU[k];
  const bool ref = !dot && M.ntn_mode == SGNN_NTN_REFERENCE;
  const int act = dot ? SGNN_ACT_IDENTITY : M.pair_act;
  const int q_begin = blockIdx.y * q_per_block;
  const int q_end = min(m, q_begin + q_per_block);
  for (int q0 = q_begin; q0 < q_end; q0 += QCHUNK) {
    const int nq = min(QCHUNK, q_end - q0);
    __syncthreads();
    {
      const float4* src = reinterpret_cast<const float4*>(P + (size_t)q0 * K * DEP);
      float4* dst = reinterpret_cast<float4*>(sP);
      for (int i = threadIdx.x; i < nq * K * DEP / 4; i += blockDim.x) dst[i] = src[i];
      for (int i = threadIdx.x; i < nq * K; i += blockDim.x) sC[i] = C[(size_t)q0 * K + i];
    }
    __syncthreads();
    if (g < n) {
      for (int q = 0; q < nq; ++q) {
        float st = 0.f;
        for (int k = 0; k < K; ++k) {
          const float4* p4 = reinterpret_cast<const float4*>(sP + (q * K + k) * DEP);
          float mk = sC[q * K + k];
#pragma unroll
          for (int b = 0; b < DEP / 4; ++b) {
            const float4 p = p4[b];
            mk = fmaf(p.x, x[4 * b + 0], mk);
            mk = fmaf(p.y, x[4 * b + 1], mk);
            mk = fmaf(p.z, x[4 * b + 2], mk);
            mk = fmaf(p.w, x[4 * b + 3], mk);
          }
          const float t = sgnn_act(mk, act, M.yeta);
          st = ref ? st + t : fmaf(sU[k], t, st);
        }
        const float score = ref ? su * st : st;
        sim[(size_t)(q0 + q) * ld + g] = sgnn_act(score, M.final_act, M.yeta);
      }
    }
  }
}
