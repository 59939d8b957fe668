/*
 * sgnn.h -- C ABI of libsgnn.so: the B200 (sm_100a) Siamese graph-pair similarity path.
 *
 * The reference (kangzf/GraphEmbedding, pure Python/TensorFlow-1) has no FFI or operator
 * registry; the seams this library sits behind are Python-level.
 * Each entry point cites the reference code it replaces (paths relative to the reference
 * root).  Conventions:
 *   - every call returns 0 on success, non-zero on error; sgnn_last_error() gives the text
 *     (no exception crosses the ABI; Python raises RuntimeError from it, like
 *     model/Siamese/layers_factory.py:33 / train.py:88 do);
 *   - plain pointers and sizes only, no torch / numpy types;
 *   - a handle is bound to one GPU; not thread-safe per handle, thread-safe across handles;
 *   - "device" pointers are CUDA device pointers owned by the caller; `stream` is a
 *     cudaStream_t passed as void* (NULL = legacy default stream);
 *   - all device memory the library itself needs is allocated in sgnn_create / sgnn_reserve,
 *     never inside a compute call.
 */
#ifndef SGNN_H_
#define SGNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGNN_MAX_LAYERS 12
#define SGNN_ABI_VERSION 1

typedef struct SgnnHandle_t* SgnnHandle;

/* layer kinds: model/Siamese/layers_factory.py:20-33 (+ Attention, layers.py:143-160) */
enum {
  SGNN_LAYER_GCN = 0,       /* layers.py:52-118  */
  SGNN_LAYER_DENSE = 1,     /* layers.py:163-205 */
  SGNN_LAYER_PADDING = 2,   /* layers.py:208-227 */
  SGNN_LAYER_AVERAGE = 3,   /* layers.py:121-140 */
  SGNN_LAYER_ATTENTION = 4, /* layers.py:143-160 */
  SGNN_LAYER_NTN = 5,       /* layers.py:255-310 */
  SGNN_LAYER_DOT = 6,       /* layers.py:230-252 */
  SGNN_LAYER_HISTOGRAM = 7  /* SimGNN pairwise node-node histogram; build-defined (absent from the reference
                             * snapshot): aux = bins.  Must directly follow the NTN. */
};
/* SimGNN-style head (build-defined): layers AFTER the NTN.  With a head the NTN emits its
 * K-vector f_k = U_k * inneract(m_k) (no sum); an optional Histogram appends `bins` features (node-node
 * similarity histogram of the last node-level outputs, constant for the gradient); Dense layers
 * (layers.py:192-205 on the 1 x (K + bins) pair row) reduce it to the raw score (last output_dim = 1),
 * to which final_act is applied.  At most SGNN_MAX_HEAD Dense layers. */
#define SGNN_MAX_HEAD 6

/* activations: layers_factory.py:101-114 */
enum { SGNN_ACT_RELU = 0, SGNN_ACT_IDENTITY = 1, SGNN_ACT_SIGMOID = 2, SGNN_ACT_TANH = 3,
       SGNN_ACT_SIM_KERNEL = 4 /* exp(-yeta x^2), src/similarity.py:58-60 */ };

enum { SGNN_NTN_REFERENCE = 0 /* (sum U)(sum act(m)), layers.py:305-308 */, SGNN_NTN_PAPER = 1 };
enum { SGNN_LOSS_REFERENCE = 0 /* (B,1)-(B,) broadcast, model_mse.py:145-151 */, SGNN_LOSS_PAIRED = 1 };
enum { SGNN_PREC_FP32 = 0, SGNN_PREC_BF16X3 = 1, SGNN_PREC_BF16 = 2 };

typedef struct {
  int32_t kind;          /* SGNN_LAYER_* */
  int32_t input_dim;     /* GCN/Dense/Attention/NTN */
  int32_t output_dim;    /* GCN/Dense */
  int32_t act;           /* GCN/Dense act, NTN inneract */
  int32_t bias;          /* 0/1 */
  int32_t dropout;       /* 0/1: layer spec `dropout=True` (layers.py:45-49) */
  int32_t sparse_inputs; /* GCN */
  int32_t aux;           /* Padding: max_in_dims; NTN: feature_map_dim; Histogram: bins */
  float fparam;          /* Padding: padding_value */
} SgnnLayerDesc;

/* model/Siamese/config.py flags that reach the device path */
typedef struct {
  int32_t abi_version;   /* SGNN_ABI_VERSION */
  int32_t num_layers;
  SgnnLayerDesc layers[SGNN_MAX_LAYERS];
  int32_t input_dim;     /* one-hot width, graphs.py:113-114 */
  int32_t final_act;     /* config.py:82-83 */
  float yeta;            /* config.py:81 */
  int32_t ntn_mode;      /* SGNN_NTN_* */
  int32_t loss_mode;     /* SGNN_LOSS_* */
  float weight_decay;    /* config.py:104-105 */
  float learning_rate;   /* config.py:107 */
  float beta1, beta2, epsilon; /* tf.train.AdamOptimizer defaults, models.py:28-29 */
  float dropout;         /* config.py:103.  Honoured by sgnn_train_step / sgnn_eval_pairs (the reference
                          * keeps dropout on for train, val and test alike); the batched
                          * embedding / all-pairs entry points are deterministic (no dropout). */
  int32_t precision;     /* SGNN_PREC_* for the all-pairs scorer */
  int32_t device;        /* CUDA ordinal */
} SgnnConfig;

typedef struct {
  int32_t layer;         /* index into SgnnConfig.layers */
  char name[16];         /* weights_0 | bias | weights | weights_W | weights_V | weights_U */
  int32_t ndim;
  int32_t shape[3];
  int64_t offset;        /* in floats, into the flat parameter blob */
} SgnnParamInfo;

/* A batch of graphs as one block-diagonal CSR (what graphs.py:34-95 produces per graph,
 * concatenated).  All arrays live where the call says (host for *_host calls, device
 * otherwise).  Algorithmic bytes per graph: 4N + 4(N+1) + 8 nnz. */
typedef struct {
  int32_t n_graphs;
  int32_t total_nodes;
  int64_t nnz;
  int32_t max_nodes;        /* largest N in the batch */
  const int32_t* node_off;  /* [n_graphs+1] */
  const int32_t* labels;    /* [total_nodes] one-hot column of each node (graphs.py:108-111) */
  const int32_t* rowptr;    /* [total_nodes+1] offsets into col/val */
  const int32_t* col;       /* [nnz] node index local to its graph */
  const float* val;         /* [nnz] A_hat = D^-1/2 (A+I) D^-1/2 (graphs.py:64-76) */
} SgnnGraphBatch;

/* ---- host-side packer (no GPU needed) ------------------------------------------------
 * Replaces ModelGraph._preprocess_adj/_normalize_adj/_sparse_to_tuple (graphs.py:64-95)
 * for a whole list of graphs at once.  edges are undirected pairs of local node indices,
 * edge_off[g]..edge_off[g+1] index pairs of graph g.  Outputs are (row,col)-sorted CSR
 * with float32 values bit-identical to float32(reference float64 value).
 * col/val must have room for total_nodes + 2*num_edges entries. */
int sgnn_pack_graphs(int32_t n_graphs, const int32_t* node_counts, const int64_t* edge_off,
                     const int32_t* edges, int32_t* node_off, int32_t* rowptr, int32_t* col,
                     float* val, int64_t* nnz_out, int32_t* max_nodes_out);

/* ---- lifecycle ----------------------------------------------------------------------- */
/* create_model + Model.__init__ (models_factory.py:5-11, models.py:7-37). */
int sgnn_create(const SgnnConfig* cfg, SgnnHandle* out);
int sgnn_destroy(SgnnHandle h);
/* h may be NULL: returns the last error of a failed sgnn_create on this thread. */
const char* sgnn_last_error(SgnnHandle h);
int sgnn_abi_version(void);

/* layer.vars (layers.py:29, 70-73, 174-177, 269-277) as one flat fp32 blob */
int64_t sgnn_param_count(SgnnHandle h);
int sgnn_param_table(SgnnHandle h, SgnnParamInfo* out, int32_t capacity); /* returns #entries or <0 */
int sgnn_embed_dim(SgnnHandle h);
int sgnn_set_params(SgnnHandle h, const float* host_params);   /* also resets Adam state */
int sgnn_get_params(SgnnHandle h, float* host_params);
/* glorot / zeros (inits.py:11-21), counter-based RNG seeded by `seed` */
int sgnn_init_params(SgnnHandle h, uint64_t seed);
/* device pointers of the handle-owned flat buffers (params, grads, adam m, adam v), each
 * sgnn_param_count floats; for torch.distributed all-reduce on grads. */
int sgnn_device_buffers(SgnnHandle h, float** params, float** grads, float** adam_m, float** adam_v);
/* size the internal workspaces used by the *_host entry points */
int sgnn_reserve(SgnnHandle h, int64_t max_graphs, int64_t max_nodes, int64_t max_nnz,
                 int64_t max_pairs, int64_t max_queries, int64_t max_db);

/* ---- device-resident compute --------------------------------------------------------- */
/* GCN/Dense/pooling stack of every graph in the batch -> emb[n_graphs][embed_dim]
 * (layers.py:91-118, 136-140, 154-160, 192-205, 223-227). */
int sgnn_embed_graphs(SgnnHandle h, const SgnnGraphBatch* batch_dev, float* emb_dev, void* stream);
/* Node-level output of the last GraphConvolution/Dense layer of every graph,
 * node_out[total_nodes][sgnn_node_dim] (the input of the pooling layer, layers.py:136-160).  Needs the
 * node-row workspace: sgnn_reserve(max_nodes >= total_nodes). */
int sgnn_node_dim(SgnnHandle h);
int sgnn_node_embeddings(SgnnHandle h, const SgnnGraphBatch* batch_dev, float* node_out_dev, void* stream);
/* sgnn_embed_graphs and sgnn_node_embeddings in ONE pass: emb_dev [n_graphs][embed_dim] and
 * node_out_dev [total_nodes][sgnn_node_dim].  Graphs of <= 32 nodes need no workspace. */
int sgnn_embed_graphs_nodes(SgnnHandle h, const SgnnGraphBatch* batch_dev, float* emb_dev, float* node_out_dev,
                            void* stream);
/* Pairwise node-node similarity histogram of SimGNN.  NOT in the reference snapshot (only TensorBoard
 * summaries are called "histogram" there, models.py:116 / layers.py:43) -- build-defined:
 *   s_ij = sigmoid(H1_i . H2_j); bin = min(bins-1, floor(s_ij * bins)); hist[b] = count_b / (N1 N2).
 * hist_dev [n_pairs][bins] and counts_dev [n_pairs][bins] (either may be NULL); bins <= 32. */
int sgnn_pair_histogram(SgnnHandle h, const float* node_emb_dev, const int32_t* node_off_dev, const int32_t* g1_dev,
                        const int32_t* g2_dev, int32_t n_pairs, int32_t bins, float* hist_dev, int32_t* counts_dev,
                        void* stream);
/* enable = 1: sgnn_embed_graphs_nodes (and the ranking entry points of models with a Histogram head) run the dense GCN
 * layers H W of layers.py:96-118 as tcgen05 GEMMs over 128-row tiles of packed graphs (csrc/sgnn_embed_tc.cuh: 6-product
 * bf16 split, fp32 in TMEM, A_hat applied outside the MMA through the tile's CSR) whenever the model fits: GCN-only node
 * stacks of widths 16..64 (multiples of 16), graphs of <= 64 nodes, no dropout, Average / Attention / Padding pooling.
 * Same results to fp32 rounding.  OFF by default: measured slower than the CUDA-core tile kernel on ~10-node graphs. */
int sgnn_set_embed_tc(SgnnHandle h, int32_t enable);
/* sgnn_pair_histogram for LARGE graphs on the tensor cores (csrc/sgnn_hist_pairs_tc.cuh; Web-shaped workloads,
 * BASELINE configs[4]): same layer, same outputs (integer counts equal outside the bin-edge band of the 6-product bf16
 * split).  Every graph is cut into tiles of 128 node rows; a work unit is one (row tile, column tile) of one pair, so
 * graphs far below 128 nodes waste most of a tile -- use it from ~50 nodes per graph up (the Python engine routes by
 * that).  n_graphs / total_nodes describe node_off_dev / node_emb_dev.  Needs a tensor-core precision, node width
 * <= 16, bins <= 16 (return code 4 otherwise: use sgnn_pair_histogram).  Build-defined like the layer itself (SURVEY
 * a24); the first call of a size grows an internal workspace (24 KB per 128 node rows), later calls allocate nothing. */
int sgnn_pair_histogram_tc(SgnnHandle h, const float* node_emb_dev, const int32_t* node_off_dev, int32_t n_graphs,
                           int64_t total_nodes, const int32_t* g1_dev, const int32_t* g2_dev, int32_t n_pairs,
                           int32_t bins, float* hist_dev, int32_t* counts_dev, void* stream);
/* NTN/Dot + final_act on listed pairs (layers.py:249-252, 282-310; model_mse.py:39-45).
 * raw_dev / sim_dev may be NULL. */
int sgnn_score_pairs(SgnnHandle h, const float* emb_dev, const int32_t* g1_dev, const int32_t* g2_dev,
                     int32_t n_pairs, float* raw_dev, float* sim_dev, void* stream);
/* one sess.run of test_output / train_val_outputs (train.py:77-102) for n_pairs pairs */
int sgnn_forward_pairs(SgnnHandle h, const SgnnGraphBatch* batch_dev, const int32_t* g1_dev,
                       const int32_t* g2_dev, int32_t n_pairs, float* emb_ws_dev, float* raw_dev,
                       float* sim_dev, void* stream);
/* the m x n loop of test() (train.py:57-61) as one call: sim[i*ld + j] for every
 * (query i, database j). */
int sgnn_score_all_pairs(SgnnHandle h, const float* emb_q_dev, int32_t m, const float* emb_db_dev,
                         int32_t n, float* sim_dev, int64_t ld, void* stream);
/* sgnn_score_all_pairs in separately callable stages (bit flags, any combination, in this order): the query-side
 * operand / tables, the database-side operand, the scorer proper.  Lets a caller prepare the query side while the
 * database is still being uploaded or embedded (what sgnn_rank_host / sgnn_rank_device do internally), and lets
 * bench.py time the scorer kernel alone.  SGNN_STAGE_ALL == sgnn_score_all_pairs. */
enum { SGNN_STAGE_PREP_QUERIES = 1, SGNN_STAGE_PREP_DATABASE = 2, SGNN_STAGE_SCORE = 4, SGNN_STAGE_ALL = 7 };
int sgnn_score_all_pairs_stages(SgnnHandle h, const float* emb_q_dev, int32_t m, const float* emb_db_dev, int32_t n,
                                float* sim_dev, int64_t ld, int32_t stages, void* stream);
/* sgnn_score_all_pairs for models with a Histogram head (SGNN_LAYER_HISTOGRAM): the histogram needs the last
 * node-level rows of both sides (sgnn_embed_graphs_nodes / sgnn_node_embeddings) next to the pooled embeddings.
 * node_off_* [graphs + 1] index rows of nodes_* [total_nodes][sgnn_node_dim]; max_nodes_* = largest graph. */
int sgnn_score_all_pairs_nodes(SgnnHandle h, const float* emb_q_dev, const float* nodes_q_dev,
                               const int32_t* node_off_q_dev, int32_t max_nodes_q, int32_t m, const float* emb_db_dev,
                               const float* nodes_db_dev, const int32_t* node_off_db_dev, int32_t max_nodes_db,
                               int32_t n, float* sim_dev, int64_t ld, void* stream);
/* The Histogram layer of every (query, database) pair as ONE tensor-core GEMM with a binning epilogue
 * (csrc/sgnn_hist_tc.cuh; semantics of sgnn_pair_histogram, build-defined): counts_dev
 * [m][n][16] bytes, byte b = number of node-node cells of pair (i, j) in bin b (bytes >= bins are 0).
 * Needs bins <= 16, node width <= 16, max_nodes_q <= 15, max_nodes_db <= 64, max_nodes_q * max_nodes_db <= 255
 * (return code 4 otherwise: use sgnn_pair_histogram) and an engine created with a tensor-core precision;
 * workspace: sgnn_reserve(max_nodes >= database node rows, max_queries >= m, max_db >= n).  counts_dev may be
 * NULL (the internal count matrix that sgnn_score_all_pairs_nodes reads is filled).  stages: SGNN_STAGE_* flags. */
int sgnn_hist_all_pairs(SgnnHandle h, const float* nodes_q_dev, const int32_t* node_off_q_dev, int32_t max_nodes_q,
                        int32_t m, const float* nodes_db_dev, const int32_t* node_off_db_dev, int32_t max_nodes_db,
                        int32_t n, uint8_t* counts_dev, int32_t stages, void* stream);
/* sgnn_hist_all_pairs WITHOUT any allocation and without a handle.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and the same counts as sgnn_hist_all_pairs on a bf16x3
 * engine; the scratch (operand tiles of both sides, tile tables) lives in a file-scope device array of
 * SGNN_HIST_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  The call runs on the CURRENT
 * device and calls no cudaMalloc / cudaFree.  Problems whose scratch exceeds the array are rejected with return code
 * 5 (at node_dim <= 16 the array serves n * max_nodes_db up to about 23 000 next to 1 000 queries), shapes outside the kernel with 4 (as
 * sgnn_hist_all_pairs); the message goes to err[err_cap] (may be NULL).  All calls on a device SHARE the array,
 * which keeps the previous call's contents (every call rewrites or clears what it reads): calls that overlap on the
 * GPU race -- issue them on one stream.  counts_dev [m][n][16] bytes; rows of queries without nodes are not written. */
#define SGNN_HIST_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int sgnn_hist_all_pairs_static(int32_t node_dim, int32_t bins, const float* nodes_q_dev, const int32_t* node_off_q_dev,
                               int32_t max_nodes_q, int32_t m, const float* nodes_db_dev,
                               const int32_t* node_off_db_dev, int32_t max_nodes_db, int32_t n, uint8_t* counts_dev,
                               void* stream, char* err, int32_t err_cap);
/* descending top-k per row with the reference's tie order (results.py:221-224: stable
 * ascending mergesort reversed => equal scores come out higher index first). */
int sgnn_topk(SgnnHandle h, const float* sim_dev, int32_t m, int32_t n, int64_t ld, int32_t k,
              int32_t* ids_dev, float* vals_dev, void* stream);
/* merge of per-shard top-k candidates (database sharded over GPUs): cand_vals/cand_ids are
 * m x c (c = shards * k), ids are GLOBAL database indices; same tie order as sgnn_topk. */
int sgnn_topk_merge(SgnnHandle h, const float* cand_vals_dev, const int32_t* cand_ids_dev, int32_t m,
                    int32_t c, int32_t k, int32_t* ids_dev, float* vals_dev, void* stream);
/* full descending ranking per row (sort_id_mat, results.py:221-224) */
int sgnn_sort_rows(SgnnHandle h, const float* sim_dev, int32_t m, int32_t n, int64_t ld,
                   int32_t* ids_dev, void* stream);

/* loss + gradient of the MSE part for the listed pairs, ADDED into the handle's grad
 * buffer; loss_dev[0] += mse part (models.py:67-88, model_mse.py:145-151).  targets are
 * similarities exp(-yeta d^2).  global_batch / target_sum / target_sqsum describe the whole
 * (all-rank) batch, needed by the reference-mode loss (F6). */
int sgnn_train_step(SgnnHandle h, const SgnnGraphBatch* batch_dev, const int32_t* g1_dev,
                    const int32_t* g2_dev, const float* target_sim_dev, int32_t n_pairs,
                    int32_t global_batch, double target_sum, double target_sqsum, float* loss_dev,
                    void* stream);
/* forward only on the listed pairs (no full-batch embedding pass): the 'val' and 'test'
 * branches of run_tf (train.py:84-87).  target_sim/loss_dev may be NULL (no loss wanted);
 * raw_dev = pre-final_act score (pred_sim_without_act, models.py:90-91), sim_dev = after. */
int sgnn_eval_pairs(SgnnHandle h, const SgnnGraphBatch* batch_dev, const int32_t* g1_dev,
                    const int32_t* g2_dev, const float* target_sim_dev, int32_t n_pairs,
                    int32_t global_batch, double target_sum, double target_sqsum, float* loss_dev,
                    float* raw_dev, float* sim_dev, void* stream);
/* weight-decay gradient + TF-1 Adam update on every variable (models.py:28-29, 67-74),
 * then zeroes the grad buffer.  grad_scale multiplies the summed gradient first.
 * loss_dev[1] = weight-decay loss term if loss_dev != NULL. */
int sgnn_apply_adam(SgnnHandle h, float grad_scale, float* loss_dev, void* stream);
/* enable = 1: Adam's step count t and the step size lr sqrt(1 - b2^t) / (1 - b1^t) live on the DEVICE (the update
 * kernel derives and advances them itself, same double arithmetic as the host form), so that a sequence of
 * sgnn_train_step + sgnn_apply_adam calls can be captured once into a CUDA graph and replayed: one launch for K optimiser
 * steps instead of 2 K (the 128-pair step of configs[2] is otherwise bound by launch overhead).  enable = 0 copies the
 * count back to the host (synchronises).  The parameters after N steps are bit-identical in both modes. */
int sgnn_adam_device_step(SgnnHandle h, int32_t enable);
int sgnn_zero_grads(SgnnHandle h, void* stream);
/* Dropout masks (layers.py:45-49, 98-100, 196, 287-288, 332-338) come from a counter-based hash of
 * (seed, counter, pair slot, layer, element): each sgnn_train_step / sgnn_eval_pairs call consumes one
 * counter value.  TensorFlow's own random stream cannot be reproduced; the mask SEMANTICS (keep iff
 * floor(keep_prob + u) == 1, scale 1/keep_prob, one draw per node for the sparse first layer) are the
 * reference's. */
int sgnn_set_dropout_stream(SgnnHandle h, uint64_t seed, uint64_t counter);
/* Data-parallel training: local pair p of a call is pair offset + p * stride of the GLOBAL batch (default 0, 1).  The
 * masks are keyed on (seed, counter, global pair index, arm, layer, element), so R ranks that share seed and counter
 * draw exactly the masks one GPU draws for the whole batch -- offset = rank, stride = R for the interleaved split of
 * distributed.pair_slice, offset = rank * pairs_per_rank, stride = 1 for contiguous blocks. */
int sgnn_set_pair_offset(SgnnHandle h, int32_t offset, int32_t stride);

/* ranking metrics on device (src/metrics.py:14-82, results.py:51-109; Kendall tau-b is
 * build-defined, pinned to scipy.stats.kendalltau).  true_sim/pred_sim are m x n. */
int sgnn_precision_at_k(SgnnHandle h, const float* true_sim_dev, const float* pred_sim_dev, int32_t m,
                        int32_t n, const int32_t* ks_dev, int32_t n_ks, double* prec_dev, void* stream);
/* precision@k for SMALL k without sorting the rows: from the true similarities' top-kmax VALUES and the predicted
 * similarities' top-kmax IDS (two sgnn_topk calls).  The tie-inclusive true top-k set (metrics.py:14-35) is
 * { j : true[j] >= k-th largest true value }; prec@k = mean over rows of its overlap with the first k predicted ids,
 * / k.  Same values as sgnn_precision_at_k for every k <= kmax (integer hit counts: exact); entries of ks_dev outside
 * [1, kmax] give NaN.  1 000 x 10 000, k = 10: 0.05 ms instead of 2.4 ms. */
int sgnn_precision_from_topk(SgnnHandle h, const float* true_sim_dev, int64_t ld, const float* true_topk_vals_dev,
                             const int32_t* pred_topk_ids_dev, int32_t m, int32_t kmax, const int32_t* ks_dev,
                             int32_t n_ks, double* prec_dev, void* stream);
int sgnn_kendall_tau(SgnnHandle h, const float* true_sim_dev, const float* pred_sim_dev, int32_t m,
                     int32_t n, int64_t* counts_dev /* [m][5] */, double* tau_dev /* [m] */, void* stream);

/* ---- host-buffer entry points (what a Python caller with numpy arrays uses; H2D / D2H
 * inside).  These are the `e2e` path of bench.py. ---------------------------------------- */
int sgnn_forward_pairs_host(SgnnHandle h, const SgnnGraphBatch* batch_host, const int32_t* g1,
                            const int32_t* g2, int32_t n_pairs, float* raw_host, float* sim_host);
/* The whole ranking step on DEVICE-resident batches (what test() + Result do per model,
 * train.py:99-127): embeds queries and database, scores all pairs into sim_dev [m][ld], and (k > 0,
 * topk_ids_dev != NULL) selects the top-k.  Same result as sgnn_embed_graphs x2 + sgnn_score_all_pairs +
 * sgnn_topk on one stream; the query branch (embedding + query-side scorer tables) runs on an internal
 * stream beside the database branch and joins before the scorer.  Every output is complete in
 * `stream` order. */
int sgnn_rank_device(SgnnHandle h, const SgnnGraphBatch* q_dev, const SgnnGraphBatch* db_dev, int32_t k,
                     float* emb_q_dev, float* emb_db_dev, float* sim_dev, int64_t ld, int32_t* topk_ids_dev,
                     float* topk_vals_dev, void* stream);
/* ---- multi-GPU ranking: top-k candidate exchange over NVLink peer memory ------------------------
 * The database is sharded over the GPUs of one node (one process per GPU); every rank holds a
 * local top-k [m][k] (LOCAL ids + its shard's id_offset) and needs the global one.  Instead of a
 * library all-gather + merge, every rank pushes its candidates into a slot of every peer's exchange
 * buffer (remote stores over NVLink into CUDA-IPC-opened memory), raises an epoch flag, and merges
 * the world * k candidates of each query when all flags of the epoch have arrived.  Tie order as
 * sgnn_topk (results.py:221-224) on GLOBAL ids.
 *   1. every rank: sgnn_peer_create(...)  -> 64-byte IPC handle
 *   2. all-gather the handles on the host (torch.distributed / MPI / files)
 *   3. every rank: sgnn_peer_connect(all_handles [world][64])
 *   4. per step, on EVERY rank, the same number of times: sgnn_topk_exchange(...)
 *   5. host barrier, then sgnn_peer_destroy (also done by sgnn_destroy).
 * Every rank must make the SAME sequence of collective calls (sgnn_topk_exchange, sgnn_rank_sharded,
 * sgnn_apply_adam_allreduce, sgnn_peer_barrier): the kernels wait for the peers' epoch flags on the device.  The wait
 * is bounded: after 30 s without a peer's flag (a rank died, or the call counts differ) the kernel prints a message
 * and traps, so the stream reports a launch failure at the next synchronising call instead of hanging. */
#define SGNN_IPC_HANDLE_BYTES 64
int sgnn_peer_create(SgnnHandle h, int32_t world, int32_t rank, int32_t m_cap, int32_t k_cap,
                     uint8_t* ipc_handle_out);
int sgnn_peer_connect(SgnnHandle h, const uint8_t* all_handles);
int sgnn_topk_exchange(SgnnHandle h, const float* vals_dev, const int32_t* ids_dev, int32_t m, int32_t k,
                       int32_t id_offset, int32_t* out_ids_dev, float* out_vals_dev, void* stream);
int sgnn_peer_destroy(SgnnHandle h);
/* sgnn_rank_device on this rank's database shard + sgnn_topk_exchange, as ONE replayable unit (the exchange
 * epochs live in device memory, so the whole sharded step is a single CUDA graph launch per rank). */
int sgnn_rank_sharded(SgnnHandle h, const SgnnGraphBatch* q_dev, const SgnnGraphBatch* db_shard_dev, int32_t k,
                      int32_t id_offset, float* emb_q_dev, float* emb_db_dev, float* sim_dev, int64_t ld,
                      int32_t* local_ids_dev, float* local_vals_dev, int32_t* out_ids_dev, float* out_vals_dev,
                      void* stream);
/* Stream-ordered barrier between the ranks (one tiny kernel: arrive at every peer, wait for every peer);
 * collective.  Work enqueued after it starts on all GPUs within a few microseconds of each other. */
int sgnn_peer_barrier(SgnnHandle h, void* stream);
/* Data-parallel optimiser step over the same peer buffers: every rank pushes its flat gradient blob
 * (sgnn_train_step output) into every peer's buffer and applies sgnn_apply_adam to the RANK-ORDERED SUM
 * of the world blobs, read in place by the Adam kernel (all ranks apply bit-identical updates; no
 * separate all-reduce).  Collective: every rank calls it once per step. */
int sgnn_apply_adam_allreduce(SgnnHandle h, float grad_scale, float* loss_dev, void* stream);
/* Test hook: the slice-maximum hints the tensor-core scorer left for the top-k kernel during the last
 * sgnn_rank_device / sgnn_rank_host call ([m][4 * ceil(n / 128)] 64-bit rank keys: orderable score << 32 |
 * database id of the largest element of every 32-column slice of sim). */
int sgnn_debug_slice_hints(SgnnHandle h, unsigned long long** hints_dev, int64_t* capacity);
/* embeds queries + database, scores all pairs, returns top-k ids/values per query; the
 * full m x n matrix is copied back only if sim_host != NULL. */
int sgnn_rank_host(SgnnHandle h, const SgnnGraphBatch* queries_host, const SgnnGraphBatch* db_host,
                   int32_t k, int32_t* topk_ids_host, float* topk_vals_host, float* sim_host);
/* ---- compact wire format for the host-buffer entry points ---------------------------------------------
 * Over PCIe a ranking step is bound by the upload of the batch.  sgnn_compact_batch (host, no GPU) re-encodes a
 * packed batch as byte-wide labels / row lengths / column indices plus a dictionary of the distinct A_hat
 * values (~80 B per AIDS-sized graph instead of 287 B); the device expands it back into exactly the arrays it was
 * built from (bit-identical: values come out of the dictionary).  Returns 4 if the batch does not fit the byte
 * fields (labels or row lengths > 255, graphs > 256 nodes, > 256 distinct values) -- stay on sgnn_rank_host then --
 * and 5 if `capacity` < sgnn_compact_bound(batch).  The blob is self-describing and relocatable. */
int64_t sgnn_compact_bound(const SgnnGraphBatch* batch_host);
int sgnn_compact_batch(const SgnnGraphBatch* batch_host, uint8_t* blob_out, int64_t capacity, int64_t* bytes_out);
/* host restatement of the device expansion (tests / tools): blob -> the five arrays of the batch */
int sgnn_expand_compact_host(const uint8_t* blob, int32_t* node_off, int32_t* labels, int32_t* rowptr, int32_t* col,
                             float* val);
/* sgnn_rank_host on two compact blobs (pinned host memory for full PCIe rate) */
int sgnn_rank_host_compact(SgnnHandle h, const uint8_t* queries_blob, const uint8_t* db_blob, int32_t k,
                           int32_t* topk_ids_host, float* topk_vals_host, float* sim_host);
/* one optimizer step: forward + backward + Adam; returns [mse_part, wd_part] */
int sgnn_train_step_host(SgnnHandle h, const SgnnGraphBatch* batch_host, const int32_t* g1,
                         const int32_t* g2, const float* target_sim, int32_t n_pairs, float* loss_host);

#ifdef __cplusplus
}
#endif
#endif /* SGNN_H_ */
