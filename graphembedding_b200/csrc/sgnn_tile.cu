// Translation unit of the node-per-thread small-graph kernels (sgnn_tile.cuh).
#include "sgnn_large.cuh"
#include "sgnn_tile.cuh"
#include "sgnn_internal.h"
#include <algorithm>

cudaError_t sgnn_launch_embed_tile(const ModelDev& M, const BatchDev& B, int max_nodes, const float* params,
                                   float* emb, float* node_out, void* tc_operand_a, int kch, int parts,
                                   void* tc_operand_b, int ntn_k, int kc, int beside_graphs, int beside_max_nodes,
                                   int sm_count, int smem_optin, cudaStream_t st) {
  if (max_nodes > 32) return cudaErrorInvalidValue;
  const EmbPlan P = make_emb_plan(M, max_nodes);
  // the fused query-side operand stages a tile's pooled rows in one of the warp's activation buffers
  if (tc_operand_b && P.gpw * M.emb_dim > 32 * (P.ldh < P.ldb ? P.ldh : P.ldb)) return cudaErrorInvalidValue;
  const int n_tiles = (B.n_graphs + P.gpw - 1) / P.gpw;
  // Warps (= tiles in flight) per CTA: as many as let two CTAs share an SM (1 KB per CTA is reserved),
  // so that a batch of <= 2 x warps x SMs tiles finishes in one round with every tile resident at once.
  const size_t sm_bytes = (size_t)smem_optin + 1024;
  // query-side operand fusion: + the (D + 1) x C coefficient matrix behind the per-warp buffers
  const size_t wm_floats = tc_operand_b ? (size_t)round_up((M.emb_dim + 1) * kch * 8 * kc, 4) : 0;
  auto bytes_for = [&](int w) { return ((size_t)P.w_floats + (size_t)w * P.warp_floats + wm_floats) * sizeof(float); };
  int warps = TILE_MAX_WARPS;
  while (warps > 4 && 2 * (bytes_for(warps) + 1024) > sm_bytes) --warps;
  if (bytes_for(warps) > (size_t)smem_optin) warps = 1;
  const size_t smem = bytes_for(warps);
  if (smem > (size_t)smem_optin) return cudaErrorInvalidValue;
  const int per_sm = 2 * (smem + 1024) <= sm_bytes ? 2 : 1;
  int blocks = (n_tiles + warps - 1) / warps;
  if (blocks > sm_count * per_sm) blocks = sm_count * per_sm;
  else if (blocks < sm_count) {
    // A small batch that runs BESIDE a large one (the queries next to the database of a ranking step): narrow
    // CTAs on every SM would take registers the large batch's second CTA needs there and push a third of its
    // CTAs into a second wave.  Full-width CTAs in the slots the large batch leaves free do not interfere; the
    // small batch then takes two rounds, still no longer than the large one's single round.
    int free_slots = 0;
    if (beside_graphs > 0 && beside_max_nodes > 0 && beside_max_nodes <= 32) {
      const EmbPlan Pb = make_emb_plan(M, beside_max_nodes);
      const int big_blocks = ((beside_graphs + Pb.gpw - 1) / Pb.gpw + warps - 1) / warps;
      free_slots = sm_count * per_sm - big_blocks;
    }
    if (free_slots >= 8 && blocks <= 4 * free_slots) {
      blocks = std::min(blocks, free_slots);
    } else {   // few tiles, nothing beside them: spread them over all SMs with narrower CTAs
      warps = (n_tiles + sm_count - 1) / sm_count;
      blocks = (n_tiles + warps - 1) / warps;
    }
  }
  cudaError_t e = cudaFuncSetAttribute(embed_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)bytes_for(warps));
  if (e != cudaSuccess) return e;
  embed_tile_kernel<<<blocks, warps * 32, bytes_for(warps), st>>>(M, P, B, params, emb, node_out,
                                                                    TileOperandOut{static_cast<__nv_bfloat16*>(tc_operand_a), kch, parts,
                                                                                   static_cast<__nv_bfloat16*>(tc_operand_b), ntn_k, kc,
                                                                                   kc > 0 ? 128 / kc : 0,
                                                                                   (int)(P.w_floats + (size_t)warps * P.warp_floats)});
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
  const int scratch = (2 + LARGE_THREADS / 32) * M.node_dim;
  const size_t smem = ((size_t)P.w_floats + (size_t)round_up(scratch, 4)) * sizeof(float);
  if (smem > (size_t)smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(embed_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int blocks = std::min(B.n_graphs, sm_count * 8);
  embed_large_kernel<<<blocks, LARGE_THREADS, smem, st>>>(M, P, B, nullptr, B.n_graphs, params, bufA, bufB, emb,
                                                          node_out);
  return cudaGetLastError();
}

#ifdef SGNN_TILE_TRACE
extern "C" int sgnn_debug_tile_trace(unsigned long long* out, int n) {
  cudaDeviceSynchronize();
  return (int)cudaMemcpyFromSymbol(out, g_tile_trace, sizeof(unsigned long long) * (size_t)n);
}
#endif
