// Launchers of the tensor-core node-node histogram (sgnn_hist_tc.cuh).
#include <cstdlib>
#include <algorithm>
#include "sgnn_hist_tc.cuh"
#include "sgnn_hist_pairs_tc.cuh"

bool sgnn_hist_tc_supported(int node_dim, int bins, int max_nodes_q, int max_nodes_db) {
  // node rows of at most 16 values: one K step; the 7 operand tiles (3 parts x 4 KB) + counter batches fill shared memory
  return node_dim >= 1 && node_dim <= 16 && bins >= 1 && bins <= HT_MAX_BINS && max_nodes_q >= 1 &&
         max_nodes_q <= HT_MAXN && max_nodes_db >= 1 && max_nodes_db <= HT_MAXN_DB &&
         max_nodes_q * max_nodes_db <= 255;   // one byte per bin in the count matrix
}

// mode 0: bucket sort of the queries + operand tiles; 1: sort only (needs the node offsets, not the node rows: a
// caller may run it beside the query embedding); 2: operand tiles only (the sort has run)
cudaError_t sgnn_hist_tc_prep_q(const HistTcGeom& G, const float* nodes_q, const int32_t* node_off_q, const HistTcWs& ws,
                                cudaStream_t st, int mode) {
  if (mode != 2) ht_sort_q_kernel<<<1, 1024, 0, st>>>(node_off_q, G.m, G.max_btiles, ws.qmeta, ws.status);
  if (mode != 1)
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
  const size_t smem = (size_t)(HT_GA + HT_STAGES) * ht_tile_bytes(G) + (size_t)HT_GROUPS * 2 * HT_SEGB * HT_CSTRIDE * 8 +
                      (2 + 2 * HT_STAGES + 2 * HT_BUFS) * 8 + 16 + 32 * 8;
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

// ---- allocation-free variant (include/sgnn.h: sgnn_hist_all_pairs_static) -------------------------------------------
// The scratch of the histogram (operand tiles of both sides, tile tables, status word) lives in ONE file-scope
// device array instead of the sgnn_reserve workspace: the call allocates nothing, so it can be used where cudaMalloc
// cannot (under stream capture, serving loops, harnesses that forbid allocation).  Every GPU that loads the library
// holds the array until the process exits.  All calls on a device share it and it keeps the previous call's contents:
// everything a call reads it first rewrites (tiles, tables) or clears (status word), and calls that overlap on the GPU
// race -- serialise them on one stream.
#include <cstdio>
#include <mutex>
#include "../../include/sgnn.h"

constexpr size_t HT_STATIC_BYTES = (size_t)SGNN_HIST_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_ht_static_scratch[HT_STATIC_BYTES];

namespace {
constexpr int HT_MAX_DEVICES = 64;
std::mutex g_ht_static_mu;
unsigned char* g_ht_static_ptr[HT_MAX_DEVICES] = {};   // device address of the array, per device (NOT the host symbol)
int g_ht_static_sms[HT_MAX_DEVICES] = {};

int ht_static_fail(char* err, int32_t cap, const char* msg, int rc = 1) {
  if (err && cap > 0) snprintf(err, (size_t)cap, "sgnn_hist_all_pairs_static: %s", msg);
  return rc;
}
inline size_t ht_align256(size_t x) { return (x + 255) & ~(size_t)255; }
}  // namespace

extern "C" int sgnn_hist_all_pairs_static(int32_t node_dim, int32_t bins, const float* nodes_q, const int32_t* node_off_q,
                                          int32_t max_nodes_q, int32_t m, const float* nodes_db,
                                          const int32_t* node_off_db, int32_t max_nodes_db, int32_t n,
                                          uint8_t* counts_dev, void* stream, char* err, int32_t err_cap) {
  if (err && err_cap > 0) err[0] = 0;
  if (!nodes_q || !node_off_q || !nodes_db || !node_off_db || !counts_dev) return ht_static_fail(err, err_cap, "null argument");
  if (m < 0 || n < 0) return ht_static_fail(err, err_cap, "negative size");
  if (m == 0 || n == 0) return 0;
  if (!sgnn_hist_tc_supported(node_dim, bins, max_nodes_q, max_nodes_db))
    return ht_static_fail(err, err_cap, "shape outside the tensor-core histogram (node_dim <= 32, bins <= 16, queries <= 15 "
                                        "nodes, database graphs <= 64 nodes, product <= 255)", 4);
  HistTcGeom G;
  G.m = m; G.n = n; G.d = node_dim; G.ksteps = (node_dim + 15) / 16; G.parts = 3; G.bins = bins;
  G.S = 129 - max_nodes_db;
  G.max_btiles = ht_max_btiles(m);
  G.max_atiles = (int)((int64_t)n * max_nodes_db / G.S + 1);
  // layout of the static scratch: A tiles | B tiles | qmeta | ameta | status
  const size_t tile = (size_t)ht_tile_bytes(G);
  const size_t off_b = ht_align256((size_t)G.max_atiles * tile);
  const size_t off_q = off_b + ht_align256((size_t)G.max_btiles * tile);
  const size_t off_a = off_q + ht_align256((size_t)ht_qmeta_ints(m, G.max_btiles) * 4);
  const size_t off_s = off_a + ht_align256((size_t)(G.max_atiles + 1) * 4);
  const size_t need = off_s + 256;
  if (need > HT_STATIC_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg), "%d queries x %d database graphs of up to %d nodes need %zu bytes of scratch, the static "
             "array holds %zu (SGNN_HIST_STATIC_SCRATCH_BYTES): use sgnn_hist_all_pairs with sgnn_reserve",
             m, n, max_nodes_db, need, HT_STATIC_BYTES);
    return ht_static_fail(err, err_cap, msg, 5);
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return ht_static_fail(err, err_cap, cudaGetErrorString(e));
  if (dev < 0 || dev >= HT_MAX_DEVICES) return ht_static_fail(err, err_cap, "device index out of range");
  unsigned char* base;
  int sms;
  {
    std::lock_guard<std::mutex> lock(g_ht_static_mu);
    if (!g_ht_static_ptr[dev]) {   // once per device, with that device current
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_ht_static_scratch);
      if (e != cudaSuccess || !p) return ht_static_fail(err, err_cap, cudaGetErrorString(e));
      int c = 0;
      e = cudaDeviceGetAttribute(&c, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess || c < 1) return ht_static_fail(err, err_cap, "cudaDeviceGetAttribute failed");
      g_ht_static_sms[dev] = c;
      g_ht_static_ptr[dev] = static_cast<unsigned char*>(p);
    }
    base = g_ht_static_ptr[dev];
    sms = g_ht_static_sms[dev];
  }
  HistTcWs ws;
  ws.At = base; ws.Bt = base + off_b;
  ws.qmeta = reinterpret_cast<int32_t*>(base + off_q);
  ws.ameta = reinterpret_cast<int32_t*>(base + off_a);
  ws.status = reinterpret_cast<int32_t*>(base + off_s);
  ws.cnt = counts_dev;
  ws.parts = G.parts; ws.ksteps = G.ksteps;
  cudaStream_t st = (cudaStream_t)stream;
  e = cudaMemsetAsync(ws.status, 0, 16, st);   // the only word that must start at zero
  if (e == cudaSuccess) e = sgnn_hist_tc_prep_q(G, nodes_q, node_off_q, ws, st);
  if (e == cudaSuccess) e = sgnn_hist_tc_prep_db(G, nodes_db, node_off_db, ws, st);
  if (e == cudaSuccess) e = sgnn_hist_tc_run(G, ws, node_off_db, sms, st);
  if (e != cudaSuccess) return ht_static_fail(err, err_cap, cudaGetErrorString(e));
  return 0;
}

// ---- listed pairs of large graphs (sgnn_hist_pairs_tc.cuh) -----------------------------------------------------------
bool sgnn_hist_pairs_tc_supported(int node_dim, int bins) {
  return node_dim >= 1 && node_dim <= 16 && bins >= 1 && bins <= HT_MAX_BINS;
}

// ws: At / Bt hold max_tiles operand tiles each, tile_off n_graphs + 1 ints, unit_off n_pairs + 1 ints.
// counts [n_pairs][bins] (required), hist [n_pairs][bins] or NULL.
cudaError_t sgnn_hist_pairs_tc_run(int node_dim, int bins, const float* nodes, const int32_t* node_off, int n_graphs,
                                   int64_t max_tiles, const int32_t* g1, const int32_t* g2, int n_pairs,
                                   const HistPairsWs& ws, float* hist, int32_t* counts, int sm_count, cudaStream_t st) {
  if (!sgnn_hist_pairs_tc_supported(node_dim, bins) || !counts) return cudaErrorInvalidValue;
  HpGeom G;
  G.n_graphs = n_graphs; G.n_pairs = n_pairs; G.d = node_dim; G.bins = bins; G.lut_bias = 0x58000000u;
  cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)n_pairs * bins * sizeof(int32_t), st);
  if (e != cudaSuccess) return e;
  hp_scan_kernel<0><<<1, 1024, 0, st>>>(node_off, nullptr, nullptr, n_graphs, ws.tile_off);
  hp_prep_kernel<<<(int)std::min<int64_t>(max_tiles, (int64_t)sm_count * 16), 128, 0, st>>>(
      G, nodes, node_off, ws.tile_off, reinterpret_cast<__nv_bfloat16*>(ws.At), reinterpret_cast<__nv_bfloat16*>(ws.Bt));
  hp_scan_kernel<1><<<1, 1024, 0, st>>>(g1, g2, ws.tile_off, n_pairs, ws.unit_off);
  e = cudaFuncSetAttribute(hist_pairs_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HP_SMEM_BYTES);
  if (e != cudaSuccess) return e;
  hist_pairs_tc_kernel<<<sm_count, HP_THREADS, HP_SMEM_BYTES, st>>>(
      G, reinterpret_cast<const __nv_bfloat16*>(ws.At), reinterpret_cast<const __nv_bfloat16*>(ws.Bt), ws.tile_off,
      ws.unit_off, node_off, g1, g2, counts);
  hp_finalize_kernel<<<(n_pairs + 255) / 256, 256, 0, st>>>(G, node_off, g1, g2, counts, hist);
  return cudaGetLastError();
}
