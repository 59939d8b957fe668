// Layout of the compact wire format of a packed batch (sgnn_pack.cpp builds it on the host, the expansion
// kernel in sgnn_api.cu reads it on the device).  Every section starts on a 16-byte boundary.
#pragma once
#include <stdint.h>

#define SGNN_COMPACT_MAGIC 0x53474e43  /* "SGNC" */

struct SgnnCompactHeader {
  int32_t magic, n_graphs, total_nodes, max_nodes;
  int64_t nnz, total_bytes;
  int32_t n_codes, pad[7];
};   // 64 bytes

struct SgnnCompactLayout {
  int64_t node_off, e_off, labels, row_nnz, col, vcode, dict, total;
};

#if defined(__CUDACC__)
__host__ __device__
#endif
inline SgnnCompactLayout sgnn_compact_layout(int32_t n_graphs, int32_t total_nodes, int64_t nnz, int n_codes) {
  SgnnCompactLayout L;
  int64_t o = 64;
  auto up = [](int64_t x) { return (x + 15) & ~(int64_t)15; };
  L.node_off = o; o = up(o + 4 * ((int64_t)n_graphs + 1));
  L.e_off = o; o = up(o + 4 * ((int64_t)n_graphs + 1));
  L.labels = o; o = up(o + total_nodes);
  L.row_nnz = o; o = up(o + total_nodes);
  L.col = o; o = up(o + nnz);
  L.vcode = o; o = up(o + nnz);
  L.dict = o; o = up(o + 4 * (int64_t)n_codes);
  L.total = o;
  return L;
}
