// Tile geometry of the tensor-core all-pairs scorer, shared by the API translation unit (sizes the
// operand workspaces, plans the grid) and sgnn_tc.cu (the kernels).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "sgnn_common.cuh"

struct TcGeom {
  int m, n, D, K;
  int KC;          // accumulator columns per query: K, or K+1 with the |x| form of relu
  int QH;          // queries per q-block (= floor(128 / KC)): one q-block = one 128-row B tile
  int KCH;         // 16-byte K chunks per part (= round_up(D+1,16)/8)
  int parts;       // 1 = bf16, 2 = bf16x3 (hi + lo)
  int n_gtiles, n_qblocks;
  int GA;          // database tiles kept resident per CTA
  int q_groups, qb_per_group;
  int stages;      // B-tile ring depth; == qb_per_group when all of a CTA's query blocks fit (no recycling)
};

constexpr int TC_NPAD = 128;     // MMA N = accumulator columns of one tile (QH * KC of them used)
constexpr int TC_BUFS = 4;       // TMEM accumulator buffers (4 x 128 columns = all 512)
constexpr int TC_MAX_STAGES = 24; // upper bound of the B-tile ring depth (barrier storage)
constexpr int TC_GA_MAX = 4;
constexpr int TC_EPI_GROUPS = 2;  // epilogue groups (must divide TC_BUFS); a group = 4 TMEM lane quadrants x 2 column parts
constexpr int TC_EPI_WARPS = 8 * TC_EPI_GROUPS;
constexpr int TC_PROD_WARP = TC_EPI_WARPS;        // bulk-copy producer
constexpr int TC_MMA_WARP0 = TC_EPI_WARPS + 1;    // first of TC_BUFS MMA issuer warps
constexpr int TC_THREADS = 32 * (TC_EPI_WARPS + 1 + TC_BUFS);

__host__ __device__ inline int tc_cols_per_query(int K, bool relu) { return relu ? K + 1 : K; }
__host__ __device__ inline int tc_a_bytes(const TcGeom& G) { return G.parts * G.KCH * 128 * 16; }
__host__ __device__ inline int tc_b_bytes(const TcGeom& G) { return G.parts * G.KCH * TC_NPAD * 16; }
constexpr int TC_TAIL_BYTES = (1 + 2 * TC_MAX_STAGES + 2 * TC_BUFS) * 8 + 16 + 4 * 64;   // barriers, TMEM slot, U

#if defined(__CUDACC__)
__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// coefficient of x_q[a] (a == D: the constant term) in column (kk, k) of the query-side operand P' / C
__device__ __forceinline__ float tc_prep_coef(const ModelDev& M, const float* __restrict__ params, int D, int K, int a,
                                              int kk, int k) {   // a == D: the constant term
  if (M.pair_kind == SGNN_LAYER_DOT) return (a < D && a == kk) ? 1.f : 0.f;
  const float* W = params + M.w_off;
  const float* V = params + M.v_off;
  if (kk < D) return a < D ? W[a * D * K + kk * K + k] : V[k * 2 * D + D + kk];
  if (kk == D) return a < D ? V[k * 2 * D + a] : (M.pair_bias ? params[M.b_off + k] : 0.f);
  return 0.f;
}

#endif

// The three stages of the tensor-core scorer (defined in sgnn_tc.cu).  The query-side operand
// (B_ws: the per-query NTN tables, bf16 hi/lo) depends only on the query embeddings and the
// database-side operand (A_ws) only on the database embeddings, so a caller may prepare them on
// different streams; sgnn_tc_score needs both.
// coef_ws: (D + 1) x (KCH * 8 * KC) floats for the query-side coefficient matrix (NULL: staged single-kernel version)
cudaError_t sgnn_tc_prep_q(const ModelDev& M, const TcGeom& G, const float* params, const float* emb_q, void* B_ws,
                           float* coef_ws, int mode, cudaStream_t st);
// mode 0: coefficient matrix + operand; 1: coefficient matrix only (needs no embeddings); 2: operand only
cudaError_t sgnn_tc_prep_db(const ModelDev& M, const TcGeom& G, const float* emb_db, void* A_ws, cudaStream_t st);
cudaError_t sgnn_tc_score(const ModelDev& M, const TcGeom& G, bool relu, bool ref, bool simk, int grid, size_t smem,
                          const float* params, const void* A_ws, const void* B_ws, float* sim, int64_t ld,
                          unsigned long long* smax, cudaStream_t st);
// smax (optional): [m][4 * n_gtiles] largest rank key of every 32-graph slice of every sim row (sgnn_launch_topk_hint)
// SimGNN-style head on the scorer's epilogue (sgnn_allpairs_tc_head.cuh): K = 16, head widths 16 / 8 / 4 / 1, no head
// dropout, 16 bins (cnt = the tensor-core histogram's count matrix) or no histogram; G must be built with KC == K
bool sgnn_tc_head_supported(const ModelDev& M);
int sgnn_tc_head_smem_bytes(bool hmma);   // hmma: the histogram half of layer 0 as a second MMA (counts as fp16 operand tiles)
// head weights in constant memory: TCH slots per device (one per engine); sgnn_tc_head_weights refreshes a slot from
// the parameters on the stream (staging: sgnn_tc_head_staging_floats() device floats owned by the engine)
int sgnn_tc_head_slots();
int sgnn_tc_head_staging_floats();
cudaError_t sgnn_tc_head_weights(const ModelDev& M, const float* params, float* staging, int slot, cudaStream_t st);
cudaError_t sgnn_tc_score_head(const ModelDev& M, const TcGeom& G, int grid, size_t smem, const float* params, int slot,
                               const void* A_ws, const void* B_ws, float* sim, int64_t ld, const uint8_t* cnt,
                               bool cnt_tiled, const int32_t* node_off_q, const int32_t* node_off_db, unsigned long long* smax,
                               cudaStream_t st);
// smax (optional): slice maxima for the hinted top-k, as sgnn_tc_score
// cnt_tiled: cnt holds the fp16 operand tiles of ht_cnt_tiled_bytes (HistTcGeom::cnt_m_pad), not [m][n][16] bytes
inline int tc_n_slices(int n) { return ((n + 127) / 128) * 4; }

// graph embedding with the dense GCN layers as tcgen05 GEMMs over tiles of packed graphs (sgnn_embed_tc.cuh);
// cudaErrorInvalidValue: shape not served (the caller keeps embed_tile_kernel / embed_large_kernel)
struct BatchDev;
bool sgnn_embed_tc_supported(const ModelDev& M, int max_nodes);
cudaError_t sgnn_launch_embed_tc(const ModelDev& M, const BatchDev& B, int max_nodes, int64_t total_rows, const float* params,
                                 float* emb, float* node_out, int32_t* status, int sm_count, int smem_optin, cudaStream_t st);
