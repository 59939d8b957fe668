// Translation unit of the node-per-thread small-graph kernels (sgnn_tile.cuh).
#include "sgnn_tile.cuh"
#include "sgnn_internal.h"

cudaError_t sgnn_launch_embed_tile(const ModelDev& M, const BatchDev& B, int max_nodes, const float* params,
                                   float* emb, int sm_count, int smem_optin, cudaStream_t st) {
  if (max_nodes > 32) return cudaErrorInvalidValue;
  const EmbPlan P = make_emb_plan(M, max_nodes);
  const size_t smem = ((size_t)P.w_floats + (size_t)TILE_WARPS * P.warp_floats) * sizeof(float);
  if (smem > (size_t)smem_optin) return cudaErrorInvalidValue;
  int dmax = 0;
  for (int l = 0; l < M.n_node; ++l) dmax = dmax > M.L[l].dout ? dmax : M.L[l].dout;
  const int n_tiles = (B.n_graphs + P.gpw - 1) / P.gpw;
  const int per_sm = smem * 2 <= (size_t)smem_optin ? 2 : 1;
  int blocks = (n_tiles + TILE_WARPS - 1) / TILE_WARPS;
  if (blocks > sm_count * per_sm) blocks = sm_count * per_sm;
  cudaError_t e;
  if (dmax > 16) {
    e = cudaFuncSetAttribute(embed_tile_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    embed_tile_kernel<32><<<blocks, TILE_WARPS * 32, smem, st>>>(M, P, B, params, emb);
  } else {
    e = cudaFuncSetAttribute(embed_tile_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    embed_tile_kernel<16><<<blocks, TILE_WARPS * 32, smem, st>>>(M, P, B, params, emb);
  }
  return cudaGetLastError();
}
