// Geometry and launch interface of the tensor-core node-node histogram (sgnn_hist_tc.cuh), shared by the API
// translation unit (workspace sizing, stage dispatch) and sgnn_hist_tc.cu (the kernels).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int HT_MAXN = 15;        // nodes of a column-side (query) graph: 4-bit counters hold <= 15 cells per thread
constexpr int HT_MAXN_DB = 64;     // nodes of a lane-side (database) graph: S = 129 - max >= max keeps every tile non-empty
constexpr int HT_MAX_BINS = 16;    // 16 nibbles in a 64-bit register
constexpr int HT_BUFS = 4;         // TMEM accumulator buffers of 128 columns
constexpr int HT_GA = 4;           // database tiles resident per A block
constexpr int HT_STAGES = 3;       // B-tile ring depth
constexpr int HT_SEGB = 16;        // segments (queries) per shared-memory counter batch
constexpr int HT_CSTRIDE = 129;    // counter words per segment row in shared memory (128 lanes + 1: the pair phase reads one lane of 32 segments)
constexpr int HT_EPI_WARPS = 16;
constexpr int HT_GROUP_WARPS = 4;  // warps per epilogue group (4: one warp per TMEM lane quadrant; 8: two that split a tile's segments)
constexpr int HT_GROUPS = HT_EPI_WARPS / HT_GROUP_WARPS;
constexpr int HT_PROD_WARP = 16, HT_MMA_WARP = 17;
constexpr int HT_THREADS = 18 * 32;
constexpr int HT_CNT_BYTES = 16;   // bytes per pair in the count matrix (one byte per bin)

// qmeta (int32): [0] = number of B tiles, [1] = number of A tiles, [4 ..) = bt_n | bt_cnt | bt_first (max_btiles each) | q_sorted (m)
// ameta (int32): a_gfirst[max_atiles + 1]
struct HistTcGeom {
  int m, n;            // queries, database graphs
  int d;               // node row width
  int ksteps;          // MMA K steps of 16 per product (= ceil(d / 16))
  int parts;           // bf16 parts per operand: 1, 2 or 3
  int bins;
  int S;               // database tile stride in node rows (129 - max_nodes_db)
  int max_atiles, max_btiles;
  uint32_t lut_bias = 0x58000000u;   // (0x4B000000 << 3) mod 2^32: see ht_count
  int cnt_m_pad = 0;   // 0: counts as bytes [m][n][16]; > 0 (= round_up(m, 8)): fp16 operand tiles of the head kernel
};

__host__ __device__ inline int ht_tile_bytes(const HistTcGeom& G) { return G.parts * G.ksteps * 4096; }
__host__ __device__ inline int ht_qmeta_ints(int m, int max_btiles) { return 4 + 3 * max_btiles + m; }


// workspace (all device memory, allocated by sgnn_reserve)
struct HistTcWs {
  void* At = nullptr;            // database-side operand tiles
  void* Bt = nullptr;            // query-side operand tiles
  int32_t* qmeta = nullptr;      // ht_qmeta_ints(m, max_btiles)
  int32_t* ameta = nullptr;      // max_atiles + 1
  int32_t* status = nullptr;     // 0 = ok; 1 = a query beyond HT_MAXN nodes; 2 = a database tile beyond 128 rows
  uint8_t* cnt = nullptr;        // [m][n][HT_CNT_BYTES] bin counts of every pair
  int64_t cap_a_tiles = 0, cap_b_tiles = 0, cap_q = 0, cap_cnt = 0;
  int parts = 0, ksteps = 0;
};

// count matrix as fp16 K-major operand tiles [ceil(n / 128)][round_up(m, 8)][2 k chunks][128 graphs][8 bins]: the A
// operand of the head kernel's second MMA (sgnn_allpairs_tc_head.cuh), one 4 KB block per (database tile, query)
inline int64_t ht_cnt_tiled_bytes(int64_t m, int64_t n) { return ((n + 127) / 128) * ((m + 7) / 8 * 8) * 4096; }
inline int ht_max_btiles(int64_t m) { return (int)(m / 8 + 16); }                 // worst case: 15-node queries, 8 per tile
inline int64_t ht_max_atiles(int64_t rows) { return rows / (129 - HT_MAXN_DB) + 2; }

// false: shape outside the tensor-core kernel (the caller keeps the CUDA-core kernel)
bool sgnn_hist_tc_supported(int node_dim, int bins, int max_nodes_q, int max_nodes_db);
cudaError_t sgnn_hist_tc_prep_q(const HistTcGeom& G, const float* nodes_q, const int32_t* node_off_q, const HistTcWs& ws,
                                cudaStream_t st, int mode = 0);
cudaError_t sgnn_hist_tc_prep_db(const HistTcGeom& G, const float* nodes_db, const int32_t* node_off_db,
                                 const HistTcWs& ws, cudaStream_t st);
cudaError_t sgnn_hist_tc_run(const HistTcGeom& G, const HistTcWs& ws, const int32_t* node_off_db, int sm_count,
                             cudaStream_t st);

// ---- listed pairs of large graphs (sgnn_hist_pairs_tc.cuh) ----
struct HistPairsWs {
  void* At = nullptr;            // plain operand tiles of every graph (12 KB each)
  void* Bt = nullptr;            // the same rows scaled by -log2(e)
  int32_t* tile_off = nullptr;   // [n_graphs + 1] first tile of every graph
  int32_t* unit_off = nullptr;   // [n_pairs + 1] first (row tile, column tile) unit of every pair
  int32_t* cnt = nullptr;        // [n_pairs][bins] when the caller wants only the normalised histogram
  int64_t cap_tiles = 0, cap_graphs = 0, cap_pairs = 0, cap_cnt = 0;
};
inline int64_t hp_max_tiles(int64_t n_graphs, int64_t total_nodes) { return total_nodes / 128 + n_graphs; }
bool sgnn_hist_pairs_tc_supported(int node_dim, int bins);
cudaError_t sgnn_hist_pairs_tc_run(int node_dim, int bins, const float* nodes, const int32_t* node_off, int n_graphs,
                                   int64_t max_tiles, const int32_t* g1, const int32_t* g2, int n_pairs,
                                   const HistPairsWs& ws, float* hist, int32_t* counts, int sm_count, cudaStream_t st);
