// Translation unit of the node-per-thread small-graph kernels (sgnn_tile.cuh).
#include "sgnn_large.cuh"
#include "sgnn_tile.cuh"
#include "sgnn_internal.h"
#include <algorithm>

cudaError_t sgnn_launch_embed_tile(const ModelDev& M, const BatchDev& B, int max_nodes, const float* params,
                                   float* emb, int sm_count, int smem_optin, cudaStream_t st) {
  if (max_nodes > 32) return cudaErrorInvalidValue;
  const EmbPlan P = make_emb_plan(M, max_nodes);
  const size_t smem = ((size_t)P.w_floats + (size_t)TILE_WARPS * P.warp_floats) * sizeof(float);
  if (smem > (size_t)smem_optin) return cudaErrorInvalidValue;
  const int n_tiles = (B.n_graphs + P.gpw - 1) / P.gpw;
  const int per_sm = smem * 2 <= (size_t)smem_optin ? 2 : 1;
  int blocks = (n_tiles + TILE_WARPS - 1) / TILE_WARPS;
  if (blocks > sm_count * per_sm) blocks = sm_count * per_sm;
  if ((size_t)M.n_params > (size_t)TILE_WARPS * P.warp_floats) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(embed_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  embed_tile_kernel<<<blocks, TILE_WARPS * 32, smem, st>>>(M, P, B, params, emb);
  return cudaGetLastError();
}

int sgnn_large_row_stride(const ModelDev& M) { return make_emb_plan(M, 32).ldh; }

cudaError_t sgnn_launch_pair_histogram(const float* node_emb, int d, const int32_t* node_off, const int32_t* g1,
                                       const int32_t* g2, int n_pairs, int bins, float* hist, int32_t* counts,
                                       int sm_count, cudaStream_t st) {
  const int blocks = std::min(n_pairs, sm_count * 16);
  pair_histogram_kernel<<<blocks, 128, 0, st>>>(node_emb, d, node_off, g1, g2, n_pairs, bins, hist, counts);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_embed_large(const ModelDev& M, const BatchDev& B, const float* params, float* bufA,
                                    float* bufB, float* emb, float* node_out, int sm_count, int smem_optin,
                                    cudaStream_t st) {
  const EmbPlan P = make_emb_plan(M, 32);
  const int scratch = std::max(M.n_params, (2 + LARGE_THREADS / 32) * M.node_dim);
  const size_t smem = ((size_t)P.w_floats + (size_t)round_up(scratch, 4)) * sizeof(float);
  if (smem > (size_t)smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(embed_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int blocks = std::min(B.n_graphs, sm_count * 8);
  embed_large_kernel<<<blocks, LARGE_THREADS, smem, st>>>(M, P, B, nullptr, B.n_graphs, params, bufA, bufB, emb,
                                                          node_out);
  return cudaGetLastError();
}
