// Translation unit of the tcgen05 all-pairs scorer (kernels in sgnn_allpairs_tc.cuh).
#include "sgnn_allpairs_tc.cuh"

namespace {

template <int KT, bool RELU, bool REF, bool SIMK>
cudaError_t launch(const ModelDev& M, const TcGeom& G, int grid, size_t smem, const float* params,
                   const __nv_bfloat16* A, const __nv_bfloat16* B, float* sim, int64_t ld, cudaStream_t st) {
  auto kern = allpairs_tc_kernel<KT, RELU, REF, SIMK>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kern<<<grid, TC_THREADS, smem, st>>>(M, G, params, A, B, sim, ld);
  return cudaGetLastError();
}

template <int KT>
cudaError_t dispatch(bool relu, bool ref, bool simk, const ModelDev& M, const TcGeom& G, int grid, size_t smem,
                     const float* params, const __nv_bfloat16* A, const __nv_bfloat16* B, float* sim, int64_t ld,
                     cudaStream_t st) {
#define SGNN_TC3(R, F)                                                              \
  (simk ? launch<KT, R, F, true>(M, G, grid, smem, params, A, B, sim, ld, st)       \
        : launch<KT, R, F, false>(M, G, grid, smem, params, A, B, sim, ld, st))
  return relu ? (ref ? SGNN_TC3(true, true) : SGNN_TC3(true, false))
              : (ref ? SGNN_TC3(false, true) : SGNN_TC3(false, false));
#undef SGNN_TC3
}

}  // namespace

cudaError_t sgnn_tc_launch(const ModelDev& M, const TcGeom& G, bool relu, bool ref, bool simk, int grid, size_t smem,
                           const float* params, const float* emb_q, const float* emb_db, void* A_ws, void* B_ws,
                           float* sim, int64_t ld, cudaStream_t st) {
  __nv_bfloat16* A = static_cast<__nv_bfloat16*>(A_ws);
  __nv_bfloat16* B = static_cast<__nv_bfloat16*>(B_ws);
  const int C = G.KCH * 8 * G.KC;
  const int col_blocks = (C + (PREP_COLS / G.KC) * G.KC - 1) / ((PREP_COLS / G.KC) * G.KC);
  const int q_blocks = col_blocks * ((G.m + PREP_Q - 1) / PREP_Q);
  const int db_blocks = (G.n_gtiles * 128 * G.KCH + PREP_COLS - 1) / PREP_COLS;
  const size_t prep_smem = ((size_t)(G.D + 1) * PREP_COLS + (size_t)G.D * PREP_Q) * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(tc_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prep_smem);
  if (e != cudaSuccess) return e;
  tc_prep_kernel<<<q_blocks + db_blocks, PREP_COLS, prep_smem, st>>>(M, G, q_blocks, col_blocks, params, emb_q, emb_db,
                                                                      A, B);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (G.K == 10) return dispatch<10>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, st);
  if (G.K == 16) return dispatch<16>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, st);
  return dispatch<0>(relu, ref, simk, M, G, grid, smem, params, A, B, sim, ld, st);
}
