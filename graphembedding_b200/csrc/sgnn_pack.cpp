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

// ---- compact wire format of a packed batch (for the host-buffer entry points) -----------------------
// A ranking step over PCIe is bound by the upload of the batch (287 B per AIDS-sized graph as int32/float
// CSR).  The CSR of small labelled graphs is highly redundant: labels, row lengths and local column indices
// fit a byte, and A_hat takes a handful of distinct values (products of deg^-1/2).  The compact blob carries
//   header | node_off int32[G+1] | e_off int32[G+1] | labels u8[N] | row_nnz u8[N] | col u8[nnz] |
//   vcode u8[nnz] | dict float[n_codes]
// (~80 B per graph) and a device kernel expands it back into EXACTLY the arrays it was built from (values
// come out of the dictionary, so they are bit-identical by construction).  Batches that do not fit the
// byte fields (labels or row lengths > 255, graphs > 256 nodes, > 256 distinct values) return 4 and the
// caller stays on the plain format.
#include <cstring>
#include <map>
#include "sgnn_compact.h"

extern "C" int64_t sgnn_compact_bound(const SgnnGraphBatch* b) {
  if (!b) return -1;
  return sgnn_compact_layout(b->n_graphs, b->total_nodes, b->nnz, 256).total;
}

extern "C" int sgnn_compact_batch(const SgnnGraphBatch* b, uint8_t* out, int64_t cap, int64_t* bytes_out) {
  if (!b || !out) return 1;
  const int32_t G = b->n_graphs, N = b->total_nodes;
  const int64_t nnz = b->nnz;
  if (b->max_nodes > 256) return 4;
  std::map<uint32_t, int> codes;   // by bit pattern: -0.0f / NaN payloads stay distinct
  std::vector<uint8_t> vcode((size_t)nnz);
  for (int64_t e = 0; e < nnz; ++e) {
    uint32_t bits;
    std::memcpy(&bits, &b->val[e], 4);
    auto it = codes.find(bits);
    if (it == codes.end()) {
      if (codes.size() == 256) return 4;
      it = codes.emplace(bits, (int)codes.size()).first;
    }
    vcode[(size_t)e] = (uint8_t)it->second;
  }
  const SgnnCompactLayout L = sgnn_compact_layout(G, N, nnz, (int)codes.size());
  if (L.total > cap) return 5;
  std::memset(out, 0, (size_t)L.total);
  SgnnCompactHeader hd;
  std::memset(&hd, 0, sizeof(hd));
  hd.magic = SGNN_COMPACT_MAGIC; hd.n_graphs = G; hd.total_nodes = N; hd.nnz = nnz; hd.max_nodes = b->max_nodes;
  hd.n_codes = (int32_t)codes.size(); hd.total_bytes = L.total;
  std::memcpy(out, &hd, sizeof(hd));
  int32_t* node_off = reinterpret_cast<int32_t*>(out + L.node_off);
  int32_t* e_off = reinterpret_cast<int32_t*>(out + L.e_off);
  for (int32_t g = 0; g <= G; ++g) {
    node_off[g] = b->node_off[g];
    e_off[g] = b->rowptr[b->node_off[g]];
  }
  for (int32_t i = 0; i < N; ++i) {
    const int32_t len = b->rowptr[i + 1] - b->rowptr[i];
    if (b->labels[i] < 0 || b->labels[i] > 255 || len < 0 || len > 255) return 4;
    out[L.labels + i] = (uint8_t)b->labels[i];
    out[L.row_nnz + i] = (uint8_t)len;
  }
  for (int64_t e = 0; e < nnz; ++e) {
    if (b->col[e] < 0 || b->col[e] > 255) return 4;
    out[L.col + e] = (uint8_t)b->col[e];
    out[L.vcode + e] = vcode[(size_t)e];
  }
  float* dict = reinterpret_cast<float*>(out + L.dict);
  for (const auto& kv : codes) std::memcpy(&dict[kv.second], &kv.first, 4);
  if (bytes_out) *bytes_out = L.total;
  return 0;
}

// Host restatement of the device expansion (tests: blob -> arrays must reproduce the batch bit for bit).
extern "C" int sgnn_expand_compact_host(const uint8_t* blob, int32_t* node_off, int32_t* labels, int32_t* rowptr,
                                        int32_t* col, float* val) {
  if (!blob) return 1;
  SgnnCompactHeader hd;
  std::memcpy(&hd, blob, sizeof(hd));
  if (hd.magic != SGNN_COMPACT_MAGIC) return 2;
  const SgnnCompactLayout L = sgnn_compact_layout(hd.n_graphs, hd.total_nodes, hd.nnz, hd.n_codes);
  const int32_t* no = reinterpret_cast<const int32_t*>(blob + L.node_off);
  const int32_t* eo = reinterpret_cast<const int32_t*>(blob + L.e_off);
  const float* dict = reinterpret_cast<const float*>(blob + L.dict);
  for (int32_t g = 0; g < hd.n_graphs; ++g) {
    node_off[g] = no[g];
    int32_t e = eo[g];
    for (int32_t i = no[g]; i < no[g + 1]; ++i) {
      rowptr[i] = e;
      labels[i] = blob[L.labels + i];
      e += blob[L.row_nnz + i];
    }
  }
  node_off[hd.n_graphs] = no[hd.n_graphs];
  rowptr[hd.total_nodes] = (int32_t)hd.nnz;
  for (int64_t e = 0; e < hd.nnz; ++e) {
    col[e] = blob[L.col + e];
    val[e] = dict[blob[L.vcode + e]];
  }
  return 0;
}
