// Translation unit of the ranking / metric kernels (sgnn_rank.cuh).
#include "sgnn_rank.cuh"
#include "sgnn_internal.h"
#include <algorithm>

// power of two >= max(n, 16): block_bitonic_u64 works on 16-key blocks
static int pad_pow2(int n) {
  int p = 16;
  while (p < n) p <<= 1;
  return p;
}

cudaError_t sgnn_launch_topk(const float* sim, const int32_t* col_ids, int m, int n, int64_t ld, int k,
                             int32_t* ids, float* vals, cudaStream_t st) {
  if (k <= 256) {
    topk_select_kernel<<<m, 256, 0, st>>>(sim, col_ids, n, ld, k, ids, vals);
  } else {
    const size_t row_bytes = (size_t)n * sizeof(float);
    const int in_smem = row_bytes <= 48 * 1024;
    topk_kernel<<<m, 256, in_smem ? row_bytes : 0, st>>>(sim, col_ids, n, ld, k, in_smem, ids, vals);
  }
  return cudaGetLastError();
}

cudaError_t sgnn_launch_topk_hint(const float* sim, const unsigned long long* smax, int n_slices, int m, int n, int64_t ld, int k,
                                  int32_t* ids, float* vals, cudaStream_t st) {
  const int blocks = (m + 7) / 8;
  if (k <= 32 && n_slices >= 32 && n_slices <= 32 * 10)
    topk_hint_warp_kernel<10><<<blocks, 256, 0, st>>>(sim, smax, n_slices, m, n, ld, k, ids, vals);
  else if (k <= 32 && n_slices >= 32 && n_slices <= 32 * 32)
    topk_hint_warp_kernel<32><<<blocks, 256, 0, st>>>(sim, smax, n_slices, m, n, ld, k, ids, vals);
  else if (n_slices >= 256)
    topk_hint_kernel<<<m, 256, 0, st>>>(sim, smax, n_slices, n, ld, k, ids, vals);
  else
    return cudaErrorInvalidValue;
  return cudaGetLastError();
}

cudaError_t sgnn_launch_sort_rows(const float* sim, int m, int n, int64_t ld, int32_t* ids, int smem_optin,
                                  cudaStream_t st) {
  const int n_pad = pad_pow2(n);
  const size_t bytes = (size_t)n_pad * sizeof(unsigned long long);
  if (bytes > (size_t)smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(sort_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  sort_rows_kernel<<<m, 1024, bytes, st>>>(sim, n, n_pad, ld, ids);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_prec_from_topk(const float* true_sim, int64_t ld, const float* true_vals, const int32_t* pred_ids,
                                       int m, int kmax, const int32_t* ks, int n_ks, double* prec, cudaStream_t st) {
  prec_from_topk_kernel<<<n_ks, 256, 0, st>>>(true_sim, ld, true_vals, pred_ids, m, kmax, ks, n_ks, prec);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_prec_at_k(const float* true_sim, const float* pred_sim, int m, int n, const int32_t* ks,
                                  int n_ks, double* prec, int smem_optin, cudaStream_t st) {
  const int n_pad = pad_pow2(n);
  const size_t bytes = (size_t)n_pad * 8 + (size_t)(2 * n + 2) * 4;
  if (bytes > (size_t)smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(prec_at_k_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(prec, 0, (size_t)n_ks * sizeof(double), st);
  if (e != cudaSuccess) return e;
  prec_at_k_kernel<<<m, 1024, bytes, st>>>(true_sim, pred_sim, n, n_pad, ks, n_ks,
                                          reinterpret_cast<unsigned long long*>(prec));
  prec_finish_kernel<<<(n_ks + 255) / 256, 256, 0, st>>>(ks, n_ks, m, prec);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_kendall(const float* true_sim, const float* pred_sim, int m, int n, long long* counts,
                                double* tau, int smem_optin, cudaStream_t st) {
  const int n_pad = pad_pow2(n);
  const size_t bytes = (size_t)n_pad * 8;
  if (n_pad > 16384 || bytes > (size_t)smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(kendall_tau_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  kendall_tau_kernel<<<m, 1024, bytes, st>>>(true_sim, pred_sim, n, n_pad, counts, tau);
  return cudaGetLastError();
}

#ifdef SGNN_TILE_TRACE
extern "C" int sgnn_debug_topk_trace(int* out, int n) {
  cudaDeviceSynchronize();
  return (int)cudaMemcpyFromSymbol(out, g_topk_dbg, sizeof(int) * (size_t)n);
}
#endif

// ---- peer-memory candidate exchange (sgnn_peer.cuh) ----
#include "sgnn_peer.cuh"

cudaError_t sgnn_launch_peer_exchange(const float* vals, const int32_t* ids, int m, int k, int id_offset, int rank,
                                      int world, int64_t slot_elems, void* const* peer_bufs,
                                      unsigned long long* epoch_ctr, unsigned int* done_counter, int32_t* out_ids,
                                      float* out_vals, cudaStream_t st) {
  if (world < 1 || world > PEER_MAX_WORLD || (int64_t)m * k > slot_elems) return cudaErrorInvalidValue;
  PeerPtrs P;
  for (int r = 0; r < PEER_MAX_WORLD; ++r)
    P.base[r] = r < world ? static_cast<unsigned long long*>(peer_bufs[r]) : nullptr;
  const int count = m * k;
  const int blocks = std::max(1, std::min(64, (count + 255) / 256));
  peer_push_kernel<<<blocks, 256, 0, st>>>(vals, ids, count, id_offset, rank, world, slot_elems, P, epoch_ctr,
                                           done_counter);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const size_t smem = (size_t)8 * world * k * sizeof(unsigned long long);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  peer_merge_kernel<<<(m + 7) / 8, 256, smem, st>>>(P.base[rank], m, k, world, slot_elems, epoch_ctr, out_ids, out_vals);
  return cudaGetLastError();
}

cudaError_t sgnn_launch_peer_barrier(int rank, int world, void* const* peer_bufs, unsigned long long* epoch_ctr,
                                     cudaStream_t st) {
  if (world < 1 || world > PEER_MAX_WORLD) return cudaErrorInvalidValue;
  PeerPtrs P;
  for (int r = 0; r < PEER_MAX_WORLD; ++r)
    P.base[r] = r < world ? static_cast<unsigned long long*>(peer_bufs[r]) : nullptr;
  peer_barrier_kernel<<<1, 32, 0, st>>>(rank, world, P, epoch_ctr + 2);
  return cudaGetLastError();
}

// u64 words of the candidate region = offset of the gradient region
int64_t sgnn_peer_grad_offset(int world, int64_t slot_elems) {
  return (int64_t)(PEER_HEADER_BYTES / 8) + 2 * (int64_t)world * slot_elems;
}

int64_t sgnn_peer_buffer_bytes(int world, int64_t slot_elems, int n_params) {
  const int64_t n_pad = (n_params + 3) / 4 * 4;
  return sgnn_peer_grad_offset(world, slot_elems) * 8 + (int64_t)PEER_HEADER_BYTES + 2 * (int64_t)world * n_pad * 4;
}

cudaError_t sgnn_launch_peer_adam(int n, int rank, int world, int64_t slot_elems, void* const* peer_bufs,
                                  unsigned long long* epoch_ctr, unsigned int* done_counter, float* params, float* grads,
                                  float* m, float* v, float grad_scale, float wd, float lr_t, float b1, float b2,
                                  float eps, float* loss, int sm_count, cudaStream_t st) {
  if (world < 1 || world > PEER_MAX_WORLD) return cudaErrorInvalidValue;
  PeerPtrs P;
  for (int r = 0; r < PEER_MAX_WORLD; ++r)
    P.base[r] = r < world ? static_cast<unsigned long long*>(peer_bufs[r]) : nullptr;
  const int n_pad = (n + 3) / 4 * 4;
  const size_t goff = (size_t)sgnn_peer_grad_offset(world, slot_elems);
  const int blocks = std::max(1, std::min((n + 255) / 256, sm_count));
  peer_grad_push_kernel<<<blocks, 256, 0, st>>>(grads, n, n_pad, rank, world, goff, P, epoch_ctr + 1, done_counter);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  adam_peer_kernel<<<blocks, 256, 0, st>>>(n, n_pad, world, P.base[rank], goff, epoch_ctr + 1, params, grads, m, v, grad_scale,
                                           wd, lr_t, b1, b2, eps, loss);
  return cudaGetLastError();
}
