// Launch wrappers defined in the kernel translation units (sgnn_rank.cu, ...), called from the C ABI
// in sgnn_api.cu.  Each returns a cudaError_t or, for argument problems, a message through *err.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "sgnn_common.cuh"

cudaError_t sgnn_launch_topk(const float* sim, const int32_t* col_ids, int m, int n, int64_t ld, int k,
                             int32_t* ids, float* vals, cudaStream_t st);
// returns cudaErrorInvalidValue if a row does not fit the shared-memory sorter
// top-k from slice-maximum hints (smax [m][n_slices], 32 columns per slice) written by the tensor-core scorer
cudaError_t sgnn_launch_topk_hint(const float* sim, const unsigned long long* smax, int n_slices, int m, int n, int64_t ld, int k,
                                  int32_t* ids, float* vals, cudaStream_t st);
// push the local top-k into every peer's exchange buffer, then merge the world * k candidates (sgnn_peer.cuh)
cudaError_t sgnn_launch_peer_exchange(const float* vals, const int32_t* ids, int m, int k, int id_offset, int rank,
                                      int world, int64_t slot_elems, void* const* peer_bufs,
                                      unsigned long long* epoch_ctr, unsigned int* done_counter, int32_t* out_ids,
                                      float* out_vals, cudaStream_t st);
cudaError_t sgnn_launch_peer_barrier(int rank, int world, void* const* peer_bufs, unsigned long long* epoch_ctr,
                                     cudaStream_t st);
int64_t sgnn_peer_buffer_bytes(int world, int64_t slot_elems, int n_params);
// push the local gradient blob to every peer, then Adam on the rank-ordered sum of the world slots (sgnn_peer.cuh)
cudaError_t sgnn_launch_peer_adam(int n, int rank, int world, int64_t slot_elems, void* const* peer_bufs,
                                  unsigned long long* epoch_ctr, unsigned int* done_counter, float* params, float* grads,
                                  float* m, float* v, float grad_scale, float wd, float lr_t, float b1, float b2,
                                  float eps, float* loss, int sm_count, cudaStream_t st);
cudaError_t sgnn_launch_sort_rows(const float* sim, int m, int n, int64_t ld, int32_t* ids, int smem_optin,
                                  cudaStream_t st);
cudaError_t sgnn_launch_prec_from_topk(const float* true_sim, int64_t ld, const float* true_vals, const int32_t* pred_ids,
                                       int m, int kmax, const int32_t* ks, int n_ks, double* prec, cudaStream_t st);
cudaError_t sgnn_launch_prec_at_k(const float* true_sim, const float* pred_sim, int m, int n, const int32_t* ks,
                                  int n_ks, double* prec, int smem_optin, cudaStream_t st);
cudaError_t sgnn_launch_kendall(const float* true_sim, const float* pred_sim, int m, int n, long long* counts,
                                double* tau, int smem_optin, cudaStream_t st);

// node-per-thread embedding of small graphs (max_nodes <= 32); cudaErrorInvalidValue if the model
// does not fit this kernel's shared memory (the caller then uses the warp-per-graph kernel)
cudaError_t sgnn_launch_embed_tile(const ModelDev& M, const BatchDev& B, int max_nodes, const float* params,
                                   float* emb, float* node_out, void* tc_operand_a, int kch, int parts,
                                   void* tc_operand_b, int ntn_k, int kc, int beside_graphs, int beside_max_nodes,
                                   int sm_count, int smem_optin, cudaStream_t st);
// beside_graphs / beside_max_nodes (optional, 0 = none): a larger batch that is embedded concurrently -- a small
// batch is then packed into the CTA slots that one leaves free
// tc_operand_a / tc_operand_b (optional): also write the database-side / query-side operand of the tensor-core
// scorer for the embedded graphs (TileOperandOut, sgnn_tile.cuh)

// CTA-per-graph embedding for graphs of any size; bufA / bufB: total_nodes x sgnn_large_row_stride floats each
int sgnn_large_row_stride(const ModelDev& M);
// emb may be NULL (node rows only); node_out may be NULL
cudaError_t sgnn_launch_embed_large(const ModelDev& M, const BatchDev& B, const float* params, float* bufA,
                                    float* bufB, float* emb, float* node_out, int sm_count, int smem_optin,
                                    cudaStream_t st);
cudaError_t sgnn_launch_pair_histogram(const float* node_emb, int d, const int32_t* node_off, const int32_t* g1,
                                       const int32_t* g2, int n_pairs, int bins, float* hist, int32_t* counts,
                                       int sm_count, cudaStream_t st);

// all-pairs scorer for models with a SimGNN-style head (sgnn_head.cu); P / C: per-query NTN tables of
// allpairs_prep_kernel with row stride DEP.  cudaErrorInvalidValue: shape not supported by the kernel.
cudaError_t sgnn_launch_allpairs_head(const ModelDev& M, int DEP, const float* params, const float* Pq, const float* Cq,
                                      const float* emb_db, const float* nodes_q, const int32_t* node_off_q,
                                      int max_nodes_q, const float* nodes_db, const int32_t* node_off_db,
                                      int max_nodes_db, int m, int n, float* sim, int64_t ld, int sm_count,
                                      int smem_optin, cudaStream_t st);
// same head, bin counts of every pair precomputed by the tensor-core histogram (sgnn_hist_tc_geom.h): cnt_in [m][n][16]
cudaError_t sgnn_launch_allpairs_head_counts(const ModelDev& M, int DEP, const float* params, const float* Pq,
                                             const float* Cq, const float* emb_db, const int32_t* node_off_q,
                                             const int32_t* node_off_db, const uint8_t* cnt_in, int m, int n,
                                             float* sim, int64_t ld, int sm_count, int smem_optin, cudaStream_t st);
// register-blocked histogram of listed pairs (sgnn_head.cu); cudaErrorInvalidValue when node_dim > 32
cudaError_t sgnn_launch_pair_histogram_tiled(const float* node_emb, int d, const int32_t* node_off, const int32_t* g1,
                                             const int32_t* g2, int n_pairs, int bins, float* hist, int32_t* counts,
                                             int sm_count, cudaStream_t st);
