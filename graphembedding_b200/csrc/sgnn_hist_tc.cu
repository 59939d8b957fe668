// Launchers of the tensor-core node-node histogram (sgnn_hist_tc.cuh).
#include <cstdlib>
#include "sgnn_hist_tc.cuh"

bool sgnn_hist_tc_supported(int node_dim, int bins, int max_nodes_q, int max_nodes_db) {
  return node_dim >= 1 && node_dim <= 32 && bins >= 1 && bins <= HT_MAX_BINS && max_nodes_q >= 1 &&
         max_nodes_q <= HT_MAXN && max_nodes_db >= 1 && max_nodes_db <= HT_MAXN_DB &&
         max_nodes_q * max_nodes_db <= 255;   // one byte per bin in the count matrix
}

cudaError_t sgnn_hist_tc_prep_q(const HistTcGeom& G, const float* nodes_q, const int32_t* node_off_q, const HistTcWs& ws,
                                cudaStream_t st) {
  ht_sort_q_kernel<<<1, 1024, 0, st>>>(node_off_q, G.m, G.max_btiles, ws.qmeta, ws.status);
  ht_prep_q_kernel<<<G.max_btiles, 128, 0, st>>>(G, nodes_q, node_off_q, ws.qmeta,
                                                 reinterpret_cast<__nv_bfloat16*>(ws.Bt));
  return cudaGetLastError();
}

cudaError_t sgnn_hist_tc_prep_db(const HistTcGeom& G, const float* nodes_db, const int32_t* node_off_db,
                                 const HistTcWs& ws, cudaStream_t st) {
  ht_prep_db_kernel<<<G.max_atiles, 128, 0, st>>>(G, nodes_db, node_off_db, ws.qmeta, ws.ameta,
                                                  reinterpret_cast<__nv_bfloat16*>(ws.At), ws.status);
  return cudaGetLastError();
}

cudaError_t sgnn_hist_tc_run(const HistTcGeom& G, const HistTcWs& ws, const int32_t* node_off_db, int sm_count,
                             cudaStream_t st) {
  static const bool unpaired = getenv("SGNN_HIST_UNPAIRED") != nullptr;   // one reciprocal per cell (cross-check)
  const size_t smem = (size_t)(HT_GA + HT_STAGES) * ht_tile_bytes(G) + (size_t)4 * HT_SEGB * 128 * 8 +
                      (2 + 2 * HT_STAGES + 2 * HT_BUFS) * 8 + 16;
  auto launch = [&](auto kern) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<sm_count, HT_THREADS, smem, st>>>(G, reinterpret_cast<const __nv_bfloat16*>(ws.At),
                                             reinterpret_cast<const __nv_bfloat16*>(ws.Bt), ws.qmeta, ws.ameta,
                                             node_off_db, ws.cnt);
    return cudaGetLastError();
  };
  return unpaired ? launch(hist_tc_kernel<false>) : launch(hist_tc_kernel<true>);
}
