// Translation unit of the tcgen05 all-pairs scorer (kernels in sgnn_allpairs_tc.cuh).
#include <cstdlib>
#include "sgnn_allpairs_tc.cuh"
#include "sgnn_allpairs_tc_head.cuh"
#include "sgnn_embed_tc.cuh"

namespace {

template <int KT, bool RELU, bool REF, bool SIMK>
cudaError_t launch(const ModelDev& M, const TcGeom& G, int grid, size_t smem, const float* params,
                   const __nv_bfloat16* A, const __nv_bfloat16* B, float* sim, int64_t ld, unsigned long long* smax,
                   cudaStream_t st) {
  auto kern = allpairs_tc_kernel<KT, RELU, REF, SIMK>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kern<<<grid, TC_THREADS, smem, st>>>(M, G, params, A, B, sim, ld, smax);
  return cudaGetLastError();
}

template <int KT>
cudaError_t dispatch(bool relu, bool ref, bool simk, const ModelDev& M, const TcGeom& G, int grid, size_t smem,
                     const float* params, const __nv_bfloat16* A, const __nv_bfloat16* B, float* sim, int64_t ld,
                     unsigned long long* smax, cudaStream_t st) {
#define SGNN_TC3(R, F)                                                              \
  (simk ? launch<KT, R, F, true>(M, G, grid, smem, params, A, B, sim, ld, smax, st)       \
        : launch<KT, R, F, false>(M, G, grid, smem, params, A, B, sim, ld, smax, st))
  return relu ? (ref ? SGNN_TC3(true, true) : SGNN_TC3(true, false))
              : (ref ? SGNN_TC3(false, true) : SGNN_TC3(false, false));
#undef SGNN_TC3
}

}  // namespace

namespace {
cudaError_t prep_launch(const ModelDev& M, const TcGeom& G, bool q_part, const float* params, const float* emb_q,
                        const float* emb_db, __nv_bfloat16* A, __nv_bfloat16* B, cudaStream_t st) {
  const int C = G.KCH * 8 * G.KC;
  const int cpb = (PREP_COLS / G.KC) * G.KC;
  const int col_blocks = (C + cpb - 1) / cpb;
  const int q_blocks = q_part ? col_blocks * ((G.m + PREP_Q - 1) / PREP_Q) : 0;
  const int db_blocks = q_part ? 0 : (G.n_gtiles * 128 * G.KCH + PREP_COLS - 1) / PREP_COLS;
  const size_t prep_smem = ((size_t)(G.D + 1) * PREP_COLS + (size_t)G.D * PREP_Q) * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(tc_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prep_smem);
  if (e != cudaSuccess) return e;
  tc_prep_kernel<<<q_blocks + db_blocks, PREP_COLS, q_part ? prep_smem : 0, st>>>(M, G, q_blocks, col_blocks, params,
                                                                                  emb_q, emb_db, A, B);
  return cudaGetLastError();
}
}  // namespace

cudaError_t sgnn_tc_prep_q(const ModelDev& M, const TcGeom& G, const float* params, const float* emb_q, void* B_ws,
                           float* coef_ws, int mode, cudaStream_t st) {
  static const bool staged = getenv("SGNN_PREP_Q_STAGED") != nullptr;   // the shared-memory version (cross-checks)
  if (staged || !coef_ws) {
    if (mode == 1) return cudaSuccess;
    return prep_launch(M, G, true, params, emb_q, nullptr, nullptr, static_cast<__nv_bfloat16*>(B_ws), st);
  }
  const int C = G.KCH * 8 * G.KC;
  if (mode != 2)
    tc_coef_kernel<<<((G.D + 1) * C + PREPQ_THREADS - 1) / PREPQ_THREADS, PREPQ_THREADS, 0, st>>>(M, G, params, coef_ws);
  if (mode == 1) return cudaGetLastError();
  const dim3 grid((C + PREPQ_THREADS - 1) / PREPQ_THREADS, (G.m + PREPQ_Q - 1) / PREPQ_Q);
  tc_prep_q_kernel<<<grid, PREPQ_THREADS, 0, st>>>(G, coef_ws, emb_q, static_cast<__nv_bfloat16*>(B_ws));
  return cudaGetLastError();
}

cudaError_t sgnn_tc_prep_db(const ModelDev& M, const TcGeom& G, const float* emb_db, void* A_ws, cudaStream_t st) {
  return prep_launch(M, G, false, nullptr, nullptr, emb_db, static_cast<__nv_bfloat16*>(A_ws), nullptr, st);
}

cudaError_t sgnn_tc_score(const ModelDev& M, const TcGeom& G, bool relu, bool ref, bool simk, int grid, size_t smem,
                          const float* params, const void* A_ws, const void* B_ws, float* sim, int64_t ld,
                          unsigned long long* smax, cudaStream_t st) {
  const __nv_bfloat16* A = static_cast<const __nv_bfloat16*>(A_ws);
  const __nv_bfloat16* B = static_cast<const __nv_bfloat16*>(B_ws);
  if (G.K == 10) return dispatch<10>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, smax, st);
  if (G.K == 16) return dispatch<16>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, smax, st);
  return dispatch<0>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, smax, st);
}

// ---- SimGNN-style head on the tensor-core scorer (sgnn_allpairs_tc_head.cuh) ----
bool sgnn_tc_head_supported(const ModelDev& M) {
  if (M.pair_kind != SGNN_LAYER_NTN || M.K != TCH_K || M.n_head != 4) return false;
  if (M.hist_bins != 0 && M.hist_bins != 16) return false;
  const int din0 = TCH_K + M.hist_bins;
  return M.Hd[0].din == din0 && M.Hd[0].dout == TCH_H0 && M.Hd[1].din == TCH_H0 && M.Hd[1].dout == TCH_H1 &&
         M.Hd[2].din == TCH_H1 && M.Hd[2].dout == TCH_H2 && M.Hd[3].din == TCH_H2 && M.Hd[3].dout == 1 &&
         !M.Hd[0].drop && !M.Hd[1].drop && !M.Hd[2].drop && !M.Hd[3].drop;
}
int sgnn_tc_head_smem_bytes(bool hmma) { return TCH_SMEM_BYTES + (hmma ? TCH_HMMA_SMEM : 0); }

// refresh constant slot `slot` from the parameters (stream-ordered: kernel -> staging -> copy into the symbol)
cudaError_t sgnn_tc_head_weights(const ModelDev& M, const float* params, float* staging, int slot, cudaStream_t st) {
  if (slot < 0 || slot >= TCH_SLOTS || !staging) return cudaErrorInvalidValue;
  tch_weights_kernel<<<(TCH_SLOT_FLOATS + 255) / 256, 256, 0, st>>>(M, params, staging);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  return cudaMemcpyToSymbolAsync(c_tch_w, staging, TCH_SLOT_FLOATS * sizeof(float),
                                 (size_t)slot * TCH_SLOT_FLOATS * sizeof(float), cudaMemcpyDeviceToDevice, st);
}
int sgnn_tc_head_slots() { return TCH_SLOTS; }
int sgnn_tc_head_staging_floats() { return TCH_SLOT_FLOATS; }

// slot >= 0: head weights from constant memory (sgnn_tc_head_weights must have run on the stream); < 0: staged in
// shared memory by the kernel itself
cudaError_t sgnn_tc_score_head(const ModelDev& M, const TcGeom& G, int grid, size_t smem, const float* params, int slot,
                               const void* A_ws, const void* B_ws, float* sim, int64_t ld, const uint8_t* cnt,
                               bool cnt_tiled, const int32_t* node_off_q, const int32_t* node_off_db, unsigned long long* smax,
                               cudaStream_t st) {
  if (!sgnn_tc_head_supported(M) || G.KC != TCH_K || G.QH != 8 || slot >= TCH_SLOTS) return cudaErrorInvalidValue;
  const __nv_bfloat16* A = static_cast<const __nv_bfloat16*>(A_ws);
  const __nv_bfloat16* B = static_cast<const __nv_bfloat16*>(B_ws);
  auto launch_it = [&](auto kern) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, TCH_THREADS, smem, st>>>(M, G, params, slot, A, B, sim, ld, cnt, node_off_q, node_off_db, smax);
    return cudaGetLastError();
  };
  const bool hist = M.hist_bins > 0;
  if (cnt_tiled && !hist) return cudaErrorInvalidValue;
  if (cnt_tiled && slot >= 0) {
    // the SimGNN-style chain's activations get the specialised epilogue (FAST of tc_head_epilogue_part)
    const bool fast = M.pair_act == SGNN_ACT_RELU && M.Hd[0].act == SGNN_ACT_RELU && M.Hd[1].act == SGNN_ACT_RELU &&
                      M.Hd[2].act == SGNN_ACT_RELU && M.Hd[3].act == SGNN_ACT_IDENTITY && M.final_act == SGNN_ACT_SIGMOID;
    if (fast) return launch_it(allpairs_tc_head_kernel<true, true, true, true>);
    return launch_it(allpairs_tc_head_kernel<true, true, true>);
  }
  if (cnt_tiled) return launch_it(allpairs_tc_head_kernel<true, false, true>);
  if (slot >= 0) return hist ? launch_it(allpairs_tc_head_kernel<true, true, false>) : launch_it(allpairs_tc_head_kernel<false, true, false>);
  return hist ? launch_it(allpairs_tc_head_kernel<true, false, false>) : launch_it(allpairs_tc_head_kernel<false, false, false>);
}

// ---- graph embedding with the dense GCN layers on the tensor cores (sgnn_embed_tc.cuh) ----
bool sgnn_embed_tc_supported(const ModelDev& M, int max_nodes) { return etc_supported(M, max_nodes); }

cudaError_t sgnn_launch_embed_tc(const ModelDev& M, const BatchDev& B, int max_nodes, int64_t total_rows, const float* params,
                                 float* emb, float* node_out, int32_t* status, int sm_count, int smem_optin, cudaStream_t st) {
  if (!etc_supported(M, max_nodes) || total_rows <= 0) return cudaErrorInvalidValue;
  const EtcPlan P = etc_plan(M, max_nodes, total_rows);
  if (P.total_bytes > smem_optin) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(embed_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P.total_bytes);
  if (e != cudaSuccess) return e;
  const int per_sm = 2 * (P.total_bytes + 1024) <= smem_optin + 1024 ? 2 : 1;
  const int blocks = std::min(P.n_tiles, sm_count * per_sm);
  embed_tc_kernel<<<blocks, ETC_THREADS, P.total_bytes, st>>>(M, P, B, (int)total_rows, params, emb, node_out, status);
  return cudaGetLastError();
}
