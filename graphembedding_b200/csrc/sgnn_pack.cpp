// Host-side graph packer: the whole-batch replacement for ModelGraph._preprocess_adj /
// _normalize_adj / _sparse_to_tuple (reference model/Siamese/graphs.py:64-95).
//
// For every graph: A_hat = D^-1/2 (A + I) D^-1/2 with D = rowsum(A + I) and inf -> 0
// (graphs.py:66, 72-76), emitted as (row, col)-sorted CSR.  The reference evaluates
// ((A . D^-1/2)^T . D^-1/2)[i][j] = (A[j][i] * dis[i]) * dis[j] in float64 with
// dis = numpy.power(rowsum, -0.5) and feeds the values through float32 placeholders
// (model_mse.py:14-23).  The same expression is evaluated here in double and rounded once
// to float, so the float32 values are bit-identical to the reference's.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../include/sgnn.h"

extern "C" int sgnn_pack_graphs(int32_t n_graphs, const int32_t* node_counts, const int64_t* edge_off,
                                const int32_t* edges, int32_t* node_off, int32_t* rowptr,
                                int32_t* col, float* val, int64_t* nnz_out, int32_t* max_nodes_out) {
  if (n_graphs < 0 || !node_counts || !edge_off || !node_off || !rowptr || !col || !val) return 1;
  int64_t nodes = 0, nnz = 0;
  int32_t max_n = 0;
  std::vector<std::vector<int32_t>> adj;
  std::vector<double> dis, diag;
  node_off[0] = 0;
  rowptr[0] = 0;
  for (int32_t g = 0; g < n_graphs; ++g) {
    const int32_t n = node_counts[g];
    if (n < 0) return 2;
    max_n = std::max(max_n, n);
    adj.assign(n, std::vector<int32_t>());
    diag.assign(n, 1.0);
    for (int32_t i = 0; i < n; ++i) adj[i].push_back(i);  // + I  (graphs.py:66)
    for (int64_t e = edge_off[g]; e < edge_off[g + 1]; ++e) {
      const int32_t u = edges[2 * e], v = edges[2 * e + 1];
      if (u < 0 || v < 0 || u >= n || v >= n) return 3;
      if (u == v) { diag[u] = 2.0; continue; }  // explicit self loop: (A + I)[u][u] = 2
      adj[u].push_back(v);
      adj[v].push_back(u);
    }
    dis.assign(n, 0.0);
    for (int32_t i = 0; i < n; ++i) {
      std::sort(adj[i].begin(), adj[i].end());
      adj[i].erase(std::unique(adj[i].begin(), adj[i].end()), adj[i].end());
      const double d = static_cast<double>(adj[i].size()) - 1.0 + diag[i];   // graphs.py:72
      const double p = std::pow(d, -0.5);               // graphs.py:73
      dis[i] = std::isinf(p) ? 0.0 : p;                 // graphs.py:74
    }
    for (int32_t i = 0; i < n; ++i) {
      for (int32_t j : adj[i]) {
        col[nnz] = j;
        const double a_ji = (j == i) ? diag[i] : 1.0;
        val[nnz] = static_cast<float>((a_ji * dis[i]) * dis[j]);  // graphs.py:76
        ++nnz;
      }
      rowptr[nodes + i + 1] = static_cast<int32_t>(nnz);
    }
    nodes += n;
    node_off[g + 1] = static_cast<int32_t>(nodes);
  }
  if (nnz_out) *nnz_out = nnz;
  if (max_nodes_out) *max_nodes_out = max_n;
  return 0;
}
