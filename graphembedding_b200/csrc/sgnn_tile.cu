// Translation unit of the node-per-thread small-graph kernels (sgnn_tile.cuh).
#include "sgnn_tile.cuh"
#include "sgnn_internal.h"

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
