// libsgnn C ABI (include/sgnn.h): model description, buffers, kernel dispatch.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "sgnn_allpairs.cuh"
#include "sgnn_tc_geom.h"
#include "sgnn_hist_tc_geom.h"
#include "sgnn_common.cuh"
#include "sgnn_internal.h"
#include "sgnn_small.cuh"
#include "sgnn_train_team.cuh"
#include "sgnn_compact.h"

namespace {

thread_local std::string g_create_error;

struct HostWs {  // device staging for the *_host entry points (sized by sgnn_reserve)
  int64_t cap_graphs = 0, cap_nodes = 0, cap_nnz = 0, cap_pairs = 0, cap_q = 0, cap_db = 0;
  int32_t *node_off[2] = {nullptr, nullptr}, *labels[2] = {nullptr, nullptr};
  int32_t *rowptr[2] = {nullptr, nullptr}, *col[2] = {nullptr, nullptr};
  float* val[2] = {nullptr, nullptr};
  uint8_t* arena[2] = {nullptr, nullptr};   // one-copy upload target for batches whose host arrays are contiguous
  int64_t arena_cap = 0;
  int32_t *g1 = nullptr, *g2 = nullptr;
  float *target = nullptr, *raw = nullptr, *sim_pairs = nullptr, *emb = nullptr;
  float *emb_q = nullptr, *emb_db = nullptr, *sim = nullptr, *topk_vals = nullptr;
  int32_t* topk_ids = nullptr;
  int64_t cap_topk = 0;
};

}  // namespace

// A ranking step is six dependent-ish kernel launches of 3..30 us each plus two copies: replayed
// as one CUDA graph, the launch gaps and the per-call driver work disappear.  A slot remembers the
// arguments of the last call; the second call with the same arguments captures the stream work,
// later ones replay it.  Any sgnn_reserve (workspaces may move) drops the graphs.
struct GraphSlot {
  uint64_t key = 0;
  bool seen = false;
  uint64_t used = 0;                    // LRU stamp
  cudaGraphExec_t exec = nullptr;
  void drop() {
    if (exec) cudaGraphExecDestroy(exec);
    exec = nullptr; seen = false; key = 0; used = 0;
  }
};

// a few argument sets per entry point (callers often ping-pong between two output buffers)
struct GraphCache {
  GraphSlot slot[4];
  uint64_t clock = 0;
  void drop() { for (auto& s : slot) s.drop(); }
  GraphSlot& find(uint64_t key) {   // the slot holding `key`, else the least recently used one (reset)
    GraphSlot* lru = &slot[0];
    for (auto& s : slot) {
      if (s.seen && s.key == key) { s.used = ++clock; return s; }
      if (s.used < lru->used) lru = &s;
    }
    lru->drop();
    lru->used = ++clock;
    return *lru;
  }
};

struct SgnnHandle_t {
  SgnnConfig cfg;
  // peer-memory top-k exchange (sgnn_peer.cuh): own buffer + the peers' buffers opened through CUDA IPC
  int peer_world = 0, peer_rank = 0;
  int64_t peer_slot_elems = 0;
  unsigned long long* d_peer_epoch = nullptr;   // device words: candidate / gradient / barrier epochs
  void* peer_bufs[16] = {};
  unsigned int* peer_done = nullptr;
  bool use_graphs = true;               // SGNN_NO_GRAPHS=1 or a failed capture turns replay off
  GraphCache g_rank_device, g_rank_host, g_rank_sharded;
  ModelDev M;
  int device = 0, sm_count = 148, smem_optin = 0;
  std::string err;
  std::vector<SgnnParamInfo> table;
  float *d_params = nullptr, *d_grads = nullptr, *d_m = nullptr, *d_v = nullptr, *d_loss = nullptr;
  int64_t adam_t = 0;
  bool embed_tc = false;                      // sgnn_set_embed_tc
  float* d_arm_ws = nullptr;                  // per-CTA arm buffers of the per-pair training kernel for graphs beyond shared memory
  int64_t cap_arm_ws = 0;
  long long* d_adam_t = nullptr;              // sgnn_adam_device_step: the step count (and the step size behind it) on the device
  bool adam_on_device = false;
  uint64_t drop_seed = 0, drop_counter = 0;   // dropout stream: one counter value per train/eval call
  int pair_offset = 0, pair_stride = 1;       // local pair p = pair offset + p * stride of the global batch (sgnn_set_pair_offset)
  float *d_rowsA = nullptr, *d_rowsB = nullptr;   // large-graph path: node-row activations (global workspace)
  int64_t cap_rows = 0;                           // in node rows
  // histogram head: last node-level rows of the query / database side of the ranking entry points
  float* d_hnodes[2] = {nullptr, nullptr};
  int64_t cap_hnodes = 0;                         // in node rows, each side
  struct NodeSide { const float* rows; const int32_t* node_off; int max_nodes; int64_t total_rows; } hside[2] = {};   // total_rows < 0: unknown
  float* d_coef = nullptr;               // tensor-core scorer: query-side coefficient matrix (tc_coef_kernel)
  float *d_P = nullptr, *d_C = nullptr;  // all-pairs query tables
  int64_t cap_P = 0;
  __nv_bfloat16 *d_Atc = nullptr, *d_Btc = nullptr;  // tensor-core operands in UMMA tile order
  int64_t cap_Atc = 0, cap_Btc = 0;                   // in elements
  unsigned long long* d_smax = nullptr;                          // [queries][32-graph slices] slice maxima (top-k hints)
  int64_t cap_smax = 0;
  HistTcWs ht;                            // tensor-core node-node histogram: operand tiles, tile tables, count matrix
  HistPairsWs hp;                         // the same layer on listed pairs of large graphs (grown by the call itself)
  int head_slot = -1;                     // constant-memory slot of the head weights (sgnn_allpairs_tc_head.cuh); -1: none free
  float* d_head_staging = nullptr;        // the block the slot is refreshed from
  int dep = 16;                           // padded embedding width of the all-pairs scorer
  cudaStream_t stream = nullptr;          // used by *_host entry points
  cudaStream_t copy_stream = nullptr;     // host -> device uploads that overlap compute (sgnn_rank_host)
  cudaStream_t aux_stream = nullptr;      // parameter-only work of the ranking step (tc_coef_kernel), beside both branches
  cudaEvent_t chunk_ev[8] = {};
  HostWs ws;
  // launch geometry decided at create
  int embed_wpb[2] = {8, 8}, train_wpb[2] = {4, 4};
  size_t embed_smem[2] = {0, 0}, train_smem[2] = {0, 0}, pair_smem = 0, team_smem32 = 0;
  bool use_team = false;
};

namespace {

// constant-memory slots of the head weights: one per engine and device (a bit mask per device)
std::mutex g_slot_mu;
unsigned g_slot_mask[64] = {};
int acquire_head_slot(int device) {
  std::lock_guard<std::mutex> lock(g_slot_mu);
  if (device < 0 || device >= 64) return -1;
  for (int s = 0; s < sgnn_tc_head_slots(); ++s)
    if (!(g_slot_mask[device] & (1u << s))) { g_slot_mask[device] |= 1u << s; return s; }
  return -1;
}
void release_head_slot(int device, int slot) {
  std::lock_guard<std::mutex> lock(g_slot_mu);
  if (device >= 0 && device < 64 && slot >= 0) g_slot_mask[device] &= ~(1u << slot);
}

int fail(SgnnHandle h, const std::string& msg) {
  if (h) h->err = msg; else g_create_error = msg;
  return 1;
}

#define CUDA_OK(h, expr)                                                                  \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(h, std::string(#expr) + ": " + cudaGetErrorString(e_));                 \
  } while (0)

BatchDev to_dev(const SgnnGraphBatch* b) {
  BatchDev d;
  d.n_graphs = b->n_graphs; d.node_off = b->node_off; d.labels = b->labels;
  d.rowptr = b->rowptr; d.col = b->col; d.val = b->val;
  return d;
}

const char* kind_name(int k) {
  static const char* names[] = {"GraphConvolution", "Dense", "Padding", "Average", "Attention", "NTN", "Dot", "Histogram"};
  return (k >= 0 && k <= 7) ? names[k] : "?";
}

// Validate the layer chain and lay the flat parameter blob out in the reference's variable
// creation order (layers.py:67-73, 146-151, 172-177, 268-277).
int build_model(const SgnnConfig& c, ModelDev& M, std::vector<SgnnParamInfo>& table, std::string& err) {
  memset(&M, 0, sizeof(M));
  if (c.abi_version != SGNN_ABI_VERSION) { err = "SgnnConfig.abi_version mismatch"; return 1; }
  if (c.num_layers < 2 || c.num_layers > SGNN_MAX_LAYERS) { err = "num_layers must be in [2, 12]"; return 1; }
  if (c.input_dim < 1) { err = "input_dim must be >= 1"; return 1; }
  if (!(c.dropout >= 0.f && c.dropout < 1.f)) { err = "dropout must be in [0, 1)"; return 1; }
  int64_t off = 0;
  int cur = c.input_dim, stage = 0, act_cols = 0, dmax = 4, aux = 0, head_cols = 0;
  auto add = [&](int layer, const char* name, int nd, int s0, int s1, int s2) {
    SgnnParamInfo p; memset(&p, 0, sizeof(p));
    p.layer = layer; strncpy(p.name, name, sizeof(p.name) - 1); p.ndim = nd;
    p.shape[0] = s0; p.shape[1] = s1; p.shape[2] = s2; p.offset = off;
    int64_t cnt = (int64_t)s0 * (nd > 1 ? s1 : 1) * (nd > 2 ? s2 : 1);
    table.push_back(p); off += cnt; return (int)p.offset;
  };
  M.pool_kind = -1; M.pair_kind = -1;
  for (int i = 0; i < c.num_layers; ++i) {
    const SgnnLayerDesc& d = c.layers[i];
    if (d.kind == SGNN_LAYER_DENSE && stage == 2) {   // SimGNN-style head: Dense on the pair row (sgnn.h)
      if (M.pair_kind != SGNN_LAYER_NTN) { err = "a Dense head needs an NTN pair layer"; return 1; }
      if (M.n_head == SGNN_MAX_HEAD) { err = "at most 6 Dense layers after the NTN"; return 1; }
      if (d.input_dim != cur) { err = "layer " + std::to_string(i) + ": input_dim does not match the previous width"; return 1; }
      if (d.output_dim < 1 || d.output_dim > 128) { err = "output_dim must be in [1, 128]"; return 1; }
      LayerDev& L = M.Hd[M.n_head];
      L.kind = d.kind; L.din = d.input_dim; L.dout = d.output_dim; L.act = d.act; L.bias = d.bias;
      L.drop = d.dropout && c.dropout > 0.f; L.ref = i;
      L.w_off = add(i, "weights", 2, L.din, L.dout, 1);
      L.b_off = d.bias ? add(i, "bias", 1, L.dout, 1, 1) : 0;
      L.out_off = head_cols; head_cols += round_up(L.dout, 4);
      cur = L.dout; ++M.n_head;
    } else if (d.kind == SGNN_LAYER_HISTOGRAM) {
      if (stage != 2 || M.pair_kind != SGNN_LAYER_NTN || M.n_head != 0 || M.hist_bins != 0) {
        err = "Histogram must directly follow the NTN layer"; return 1;
      }
      if (d.aux < 1 || d.aux > 32) { err = "Histogram: bins must be in [1, 32]"; return 1; }
      M.hist_bins = d.aux; cur = M.K + d.aux; head_cols = round_up(cur, 4);
    } else if (d.kind == SGNN_LAYER_GCN || d.kind == SGNN_LAYER_DENSE) {
      if (stage != 0) { err = std::string(kind_name(d.kind)) + " after the pooling layer is not supported"; return 1; }
      if (i == 0 && d.kind != SGNN_LAYER_GCN) { err = "layer_0 must be GraphConvolution (model_mse.py:97)"; return 1; }
      if (d.input_dim != cur) { err = "layer " + std::to_string(i) + ": input_dim does not match the previous width"; return 1; }
      if (d.output_dim < 1 || d.output_dim > 128) { err = "output_dim must be in [1, 128]"; return 1; }
      LayerDev& L = M.L[M.n_node];
      L.kind = d.kind; L.din = d.input_dim; L.dout = d.output_dim; L.act = d.act; L.bias = d.bias;
      L.drop = d.dropout && c.dropout > 0.f; L.ref = i;
      L.w_off = add(i, d.kind == SGNN_LAYER_GCN ? "weights_0" : "weights", 2, L.din, L.dout, 1);
      L.b_off = d.bias ? add(i, "bias", 1, L.dout, 1, 1) : 0;
      L.out_off = act_cols; act_cols += round_up(L.dout, 4);
      if (M.n_node > 0) { L.wt_off = aux; aux += round_up(L.din * L.dout, 4); }
      dmax = std::max(dmax, L.dout);
      cur = L.dout; ++M.n_node;
    } else if (d.kind == SGNN_LAYER_PADDING || d.kind == SGNN_LAYER_AVERAGE || d.kind == SGNN_LAYER_ATTENTION) {
      if (stage != 0 || M.n_node == 0) { err = "exactly one pooling layer must follow the node layers"; return 1; }
      M.pool_kind = d.kind; M.node_dim = cur; stage = 1;
      if (d.kind == SGNN_LAYER_PADDING) {
        if (d.aux < 1) { err = "Padding: max_in_dims must be >= 1"; return 1; }
        M.pool_aux = d.aux; M.pool_fparam = d.fparam; M.emb_dim = d.aux * cur;
      } else {
        M.emb_dim = cur;
        if (d.kind == SGNN_LAYER_ATTENTION) {
          if (d.input_dim != cur) { err = "Attention: input_dim does not match the node width"; return 1; }
          M.pool_w_off = add(i, "weights", 2, cur, cur, 1);
          M.pool_wt_off = aux; aux += round_up(cur * cur, 4);
        }
      }
    } else if (d.kind == SGNN_LAYER_NTN || d.kind == SGNN_LAYER_DOT) {
      if (stage != 1) { err = "the pair layer (NTN/Dot) must follow a pooling layer"; return 1; }
      if (d.kind == SGNN_LAYER_DOT && i != c.num_layers - 1) { err = "Dot must be the last layer"; return 1; }
      M.pair_kind = d.kind; M.D = M.emb_dim; stage = 2;
      if (d.kind == SGNN_LAYER_NTN) {
        if (d.input_dim != M.emb_dim) { err = "NTN: input_dim must equal the flattened representation width"; return 1; }
        M.K = d.aux; M.pair_act = d.act; M.pair_bias = d.bias;
        M.pair_drop = d.dropout && c.dropout > 0.f; M.pair_ref = i;
        if (M.K < 1 || M.K > 64) { err = "NTN: feature_map_dim must be in [1, 64]"; return 1; }
        if ((int64_t)M.D * M.D * M.K > 32768) { err = "NTN: input_dim^2 * feature_map_dim must be <= 32768"; return 1; }
        M.w_off = add(i, "weights_W", 3, M.D, M.D, M.K);
        M.v_off = add(i, "weights_V", 2, M.K, 2 * M.D, 1);
        M.u_off = add(i, "weights_U", 2, M.K, 1, 1);
        M.b_off = d.bias ? add(i, "bias", 1, M.K, 1, 1) : 0;
        cur = M.K; head_cols = round_up(M.K, 4);
      } else {
        M.K = 1; M.pair_act = SGNN_ACT_IDENTITY;
      }
    } else {
      err = "Unknown layer kind " + std::to_string(d.kind); return 1;
    }
  }
  if (stage != 2) { err = "the layer chain must end with NTN or Dot (or an NTN followed by a Histogram / Dense head)"; return 1; }
  if (M.hist_bins > 0 && M.n_head == 0) { err = "Histogram must be followed by a Dense head"; return 1; }
  if (M.n_head > 0 && M.Hd[M.n_head - 1].dout != 1) { err = "the Dense head must end with output_dim=1"; return 1; }
  M.head_in = M.n_head > 0 ? M.K + M.hist_bins : 0;
  M.head_floats = M.n_head > 0 ? head_cols : 0;
  if (M.emb_dim > 64) { err = "flattened representation wider than 64 is not supported"; return 1; }
  M.final_act = c.final_act; M.yeta = c.yeta; M.ntn_mode = c.ntn_mode; M.loss_mode = c.loss_mode;
  M.keep = 1.0f - c.dropout; M.inv_keep = 1.0f / M.keep;
  M.n_params = (int)off; M.input_dim = c.input_dim;
  M.ld = round_up(dmax, 4) + 4;
  M.act_ld = act_cols + 4;
  M.aux_floats = aux;
  return 0;
}

template <typename Kern>
int set_smem(SgnnHandle h, Kern k, size_t bytes) {
  CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

int plan_launches(SgnnHandle h) {
  const ModelDev& M = h->M;
  const size_t wfl = round_up(M.n_params, 4);
  const size_t budget = (size_t)h->smem_optin;
  const int DK = M.D * M.K;
  const size_t wpfl = M.pair_kind == SGNN_LAYER_NTN ? round_up(M.D * DK, 4) : 0;
  for (int v = 0; v < 2; ++v) {
    const int nmax = v == 0 ? 16 : 32;
    const int csr = v == 0 ? SmallGeom<16>::CSR_WORDS : SmallGeom<32>::CSR_WORDS;
    int wpb = 8;
    size_t need;
    while (true) {
      need = (wfl + (size_t)wpb * embed_warp_floats(M, nmax, csr)) * sizeof(float);
      if (need <= budget || wpb == 1) break;
      wpb >>= 1;
    }
    if (need > budget) return fail(h, "model too large for the small-graph embed kernel's shared memory");
    h->embed_wpb[v] = wpb; h->embed_smem[v] = need;
    const TrainLayout TL = train_layout(M, nmax, csr);
    wpb = 4;
    while (true) {
      need = (2 * wfl + wpfl + round_up(M.aux_floats, 4) + (size_t)wpb * TL.total) * sizeof(float);
      if (need <= budget || wpb == 1) break;
      wpb >>= 1;
    }
    if (need > budget) return fail(h, "model too large for the fused training kernel's shared memory");
    h->train_wpb[v] = wpb; h->train_smem[v] = need;
  }
  {
    const int Dp = round_up(M.D, 4), Kp = round_up(M.K, 4);
    h->pair_smem = (wfl + wpfl + 8 * (size_t)(2 * Dp + round_up(DK, 4) + 2 * Kp)) * sizeof(float);
    if (h->pair_smem > budget) return fail(h, "pair layer too large for shared memory");
  }
  if (set_smem(h, embed_kernel<16>, h->embed_smem[0])) return 1;
  if (set_smem(h, embed_kernel<32>, h->embed_smem[1])) return 1;
  if (set_smem(h, train_pairs_kernel<16, true>, h->train_smem[0])) return 1;
  if (set_smem(h, train_pairs_kernel<32, true>, h->train_smem[1])) return 1;
  if (set_smem(h, train_pairs_kernel<16, false>, h->train_smem[0])) return 1;
  if (set_smem(h, train_pairs_kernel<32, false>, h->train_smem[1])) return 1;
  if (set_smem(h, pair_score_kernel, h->pair_smem)) return 1;
  // CTA-per-pair team kernel (sgnn_train_team.cuh): used whenever it fits; SGNN_TRAIN_KERNEL=warp forces the
  // warp-per-pair kernel (cross-checks)
  h->team_smem32 = team_smem_floats(M, 32) * sizeof(float);
  h->use_team = h->team_smem32 <= budget;
  if (const char* tk = getenv("SGNN_TRAIN_KERNEL")) h->use_team = h->use_team && strcmp(tk, "warp") != 0;
  if (M.n_head > 0 && !h->use_team)
    return fail(h, "a model with a Dense / Histogram head needs the CTA-per-pair training kernel, which does not fit shared memory");
  if (h->use_team) {
    // opt in to all of the SM's shared memory: the per-arm buffers grow with the batch's largest graph, and
    // graphs well beyond 32 nodes still fit for the usual layer widths (launch_pairs checks per call)
    if (set_smem(h, train_team_kernel<true, 1>, (size_t)h->smem_optin)) return 1;
    if (set_smem(h, train_team_kernel<false, 1>, (size_t)h->smem_optin)) return 1;
    if (set_smem(h, train_team_kernel<true, 2>, (size_t)h->smem_optin)) return 1;
    if (set_smem(h, train_team_kernel<false, 2>, (size_t)h->smem_optin)) return 1;
    if (set_smem(h, train_team_kernel<true, 1, true>, (size_t)h->smem_optin)) return 1;
    if (set_smem(h, train_team_kernel<false, 1, true>, (size_t)h->smem_optin)) return 1;
  }
  return 0;
}

uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

template <typename T>
int dev_alloc(SgnnHandle h, T** p, int64_t count) {
  if (*p) { cudaFree(*p); *p = nullptr; }
  if (count <= 0) return 0;
  CUDA_OK(h, cudaMalloc(reinterpret_cast<void**>(p), (size_t)count * sizeof(T)));
  return 0;
}

// the histogram layer of a head model runs on the tensor cores unless the engine was created with the fp32
// (CUDA-core) precision or SGNN_HIST_TC=0
bool hist_tc_enabled(SgnnHandle h) {
  static const bool off = [] { const char* e = getenv("SGNN_HIST_TC"); return e && e[0] == '0'; }();
  return !off && h->M.n_head > 0 && h->M.hist_bins >= 1 && h->M.hist_bins <= HT_MAX_BINS && h->M.node_dim <= 32 &&
         h->cfg.precision != SGNN_PREC_FP32;
}

// NTN + Dense head in the epilogue of the tcgen05 scorer (SGNN_HEAD_TC=0: keep the CUDA-core head kernel)
bool head_tc_enabled(SgnnHandle h) {
  static const bool off = [] { const char* e = getenv("SGNN_HEAD_TC"); return e && e[0] == '0'; }();
  return !off && h->cfg.precision != SGNN_PREC_FP32 && sgnn_tc_head_supported(h->M) &&
         (h->M.hist_bins == 0 || hist_tc_enabled(h));
}

// histogram half of head layer 0 as a second MMA of the head kernel (SGNN_HEAD_HMMA=0: multiplied by the epilogue threads)
bool head_hmma_enabled() {
  static const bool off = [] { const char* e = getenv("SGNN_HEAD_HMMA"); return e && e[0] == '0'; }();
  return !off;
}

// the scorer stage of a head model will run on the tensor-core head kernel (same predicate as score_all_pairs_stages)
bool head_tc_will_run(SgnnHandle h) {
  const ModelDev& M = h->M;
  if (M.n_head == 0 || !head_tc_enabled(h)) return false;
  if (M.hist_bins == 0) return true;
  return hist_tc_enabled(h) && h->hside[0].rows && h->hside[1].rows &&
         sgnn_hist_tc_supported(M.node_dim, M.hist_bins, h->hside[0].max_nodes, h->hside[1].max_nodes);
}

int ensure_query_tables(SgnnHandle h, int64_t m) {
  if (m <= h->cap_P) return 0;
  return fail(h, "query table workspace too small: call sgnn_reserve(max_queries >= m) first");
}


// Stages of sgnn_score_all_pairs: the query-side tables, the database-side operand, the scorer.
// sgnn_rank_device / sgnn_rank_host issue them separately so that the query side overlaps the
// database upload / embedding.
enum { SCORE_PREP_Q = 1, SCORE_PREP_DB = 2, SCORE_MAIN = 4, SCORE_ALL = 7,
       // internal refinements of SCORE_PREP_Q for the tensor-core scorer: the coefficient matrix depends on the
       // parameters only, so the ranking entry points compute it before / beside the query embedding
       SCORE_COEF_ONLY = 8, SCORE_COEF_READY = 16 };

// tcgen05 path of sgnn_score_all_pairs (SGNN_PREC_BF16X3 / SGNN_PREC_BF16)
// head: the SimGNN-style head runs in the scorer's epilogue (sgnn_allpairs_tc_head.cuh): K accumulator columns per
// query (no |x| sum column), cnt = count matrix of the tensor-core histogram (NULL without a Histogram layer)
int score_all_pairs_tc(SgnnHandle h, const float* emb_q, int m, const float* emb_db, int n, float* sim,
                       int64_t ld, cudaStream_t st, int stages, unsigned long long* smax, bool head = false,
                       const uint8_t* cnt = nullptr, const int32_t* node_off_q = nullptr,
                       const int32_t* node_off_db = nullptr, bool cnt_tiled = false) {
  const ModelDev& M = h->M;
  const bool ntn = M.pair_kind == SGNN_LAYER_NTN;
  TcGeom G;
  const bool relu = !head && ntn && M.pair_act == SGNN_ACT_RELU;
  const bool ref = ntn && M.ntn_mode == SGNN_NTN_REFERENCE;
  const bool simk = M.final_act == SGNN_ACT_SIM_KERNEL;
  G.m = m; G.n = n; G.D = M.emb_dim; G.K = ntn ? M.K : 1;
  G.KC = tc_cols_per_query(G.K, relu);
  G.QH = 128 / G.KC;
  G.KCH = round_up(G.D + 1, 16) / 8;
  G.parts = h->cfg.precision == SGNN_PREC_BF16 ? 1 : 2;
  G.n_gtiles = (n + 127) / 128;
  G.n_qblocks = (m + G.QH - 1) / G.QH;
  const int64_t items = (int64_t)G.n_gtiles * G.n_qblocks;
  G.GA = (int)std::max<int64_t>(1, std::min<int64_t>(TC_GA_MAX, items / ((int64_t)h->sm_count * 4)));
  G.GA = std::min(G.GA, G.n_gtiles);
  // head kernel: its units are split evenly over the grid, a unit is GA tiles of ~7 us each for ONE of the two epilogue
  // groups, so the last unit of the longer CTAs is the tail of the kernel: single-tile units (measured 0.8993 -> 0.8947 ms
  // per ranking step; the query block's B tile is then fetched once per tile, 80 MB of L2 reads)
  if (head) G.GA = 1;
  const int g_groups = (G.n_gtiles + G.GA - 1) / G.GA;
  int q_groups = std::max(1, std::min(G.n_qblocks, h->sm_count / g_groups));
  G.qb_per_group = (G.n_qblocks + q_groups - 1) / q_groups;
  q_groups = (G.n_qblocks + G.qb_per_group - 1) / G.qb_per_group;
  G.q_groups = q_groups;
  const int64_t need_a = (int64_t)G.n_gtiles * G.parts * G.KCH * 1024;
  const int64_t need_b = (int64_t)G.n_qblocks * G.parts * G.KCH * (TC_NPAD * 8);
  if (need_a > h->cap_Atc || need_b > h->cap_Btc)
    return fail(h, "tensor-core operand workspace too small: call sgnn_reserve(max_queries, max_db) first");
  const size_t fixed = (size_t)TC_GA_MAX * tc_a_bytes(G) + TC_TAIL_BYTES + 1024 + (head ? sgnn_tc_head_smem_bytes(cnt_tiled) : 0);
  if (fixed + 2 * (size_t)tc_b_bytes(G) > (size_t)h->smem_optin)
    return fail(h, "embedding too wide for the tensor-core all-pairs tiles");
  // the head kernel splits ALL (A block, query block) units evenly over the grid (sgnn_allpairs_tc_head.cuh: TchRange)
  const int64_t head_units = (int64_t)g_groups * G.n_qblocks;
  const int head_grid = (int)std::min<int64_t>(h->sm_count, head_units);
  const int per_cta = head ? (int)((head_units + head_grid - 1) / head_grid) : G.qb_per_group;
  G.stages = (int)std::min<size_t>(std::min(per_cta, TC_MAX_STAGES), ((size_t)h->smem_optin - fixed) / tc_b_bytes(G));
  size_t smem = fixed + (size_t)G.stages * tc_b_bytes(G);
  smem = std::max<size_t>(smem, 120 * 1024);  // one CTA per SM: each CTA owns all 512 TMEM columns
  const int grid = head ? head_grid : g_groups * q_groups;
  cudaError_t e = cudaSuccess;
  if (head && h->head_slot >= 0 &&   // the head weights' constant slot is refreshed wherever the coefficient matrix is (parameters only)
      ((stages & SCORE_COEF_ONLY) || ((stages & SCORE_PREP_Q) && !(stages & SCORE_COEF_READY))))
    e = sgnn_tc_head_weights(M, h->d_params, h->d_head_staging, h->head_slot, st);
  if (e == cudaSuccess && (stages & SCORE_COEF_ONLY)) e = sgnn_tc_prep_q(M, G, h->d_params, emb_q, h->d_Btc, h->d_coef, 1, st);
  if (e == cudaSuccess && (stages & SCORE_PREP_Q))
    e = sgnn_tc_prep_q(M, G, h->d_params, emb_q, h->d_Btc, h->d_coef, (stages & SCORE_COEF_READY) ? 2 : 0, st);
  if (e == cudaSuccess && (stages & SCORE_PREP_DB)) e = sgnn_tc_prep_db(M, G, emb_db, h->d_Atc, st);
  if (e == cudaSuccess && (stages & SCORE_MAIN))
    e = head ? sgnn_tc_score_head(M, G, grid, smem, h->d_params, h->head_slot, h->d_Atc, h->d_Btc, sim, ld, cnt, cnt_tiled, node_off_q, node_off_db, smax, st)
             : sgnn_tc_score(M, G, relu, ref, simk, grid, smem, h->d_params, h->d_Atc, h->d_Btc, sim, ld, smax, st);
  if (e != cudaSuccess) return fail(h, std::string("tensor-core all-pairs launch: ") + cudaGetErrorString(e));
  return 0;
}

int score_all_pairs_stages(SgnnHandle h, const float* emb_q, int32_t m, const float* emb_db, int32_t n,
                           float* sim_dev, int64_t ld, cudaStream_t st, int stages,
                           unsigned long long* smax = nullptr);
// scorer + top-k of the fused ranking entry points: with the tensor-core scorer and a long enough
// row, the scorer emits slice maxima and the top-k kernel reads those instead of the rows
int score_and_topk(SgnnHandle h, const float* emb_q, int m, const float* emb_db, int n, float* sim, int64_t ld,
                   int stages, int k, int32_t* ids, float* vals, cudaStream_t st);

}  // namespace

namespace {
int score_and_topk(SgnnHandle h, const float* emb_q, int m, const float* emb_db, int n, float* sim, int64_t ld,
                   int stages, int k, int32_t* ids, float* vals, cudaStream_t st) {
  const int n_slices = tc_n_slices(n);
  const bool want_topk = k > 0 && ids;
  // hinted kernels: warp per row for k <= 32 and 1024..32768 columns, CTA per row for k <= 256 and >= 8192
  const bool hint = want_topk && h->cfg.precision != SGNN_PREC_FP32 && (h->M.n_head == 0 || head_tc_will_run(h)) &&
                    (int64_t)m * n_slices <= h->cap_smax &&
                    ((k <= 32 && n_slices >= 32 && n_slices <= 1024) || (k <= 256 && n_slices >= 256));
  if (score_all_pairs_stages(h, emb_q, m, emb_db, n, sim, ld, st, stages, hint ? h->d_smax : nullptr)) return 1;
  if (!want_topk) return 0;
  if (!hint) return sgnn_topk(h, sim, m, n, ld, k, ids, vals, st);
  const cudaError_t e = sgnn_launch_topk_hint(sim, h->d_smax, n_slices, m, n, ld, k, ids, vals, st);
  if (e != cudaSuccess) return fail(h, std::string("top-k (hinted): ") + cudaGetErrorString(e));
  return 0;
}

int score_all_pairs_stages(SgnnHandle h, const float* emb_q, int32_t m, const float* emb_db, int32_t n,
                           float* sim_dev, int64_t ld, cudaStream_t st, int stages, unsigned long long* smax) {
  if (m == 0 || n == 0) return 0;
  if (ld < n) return fail(h, "sgnn_score_all_pairs: ld < n");
  if (ensure_query_tables(h, m)) return 1;
  const ModelDev& M = h->M;
  const int K = M.pair_kind == SGNN_LAYER_NTN ? M.K : 1;
  if (M.n_head > 0) {   // SimGNN-style head
    // each stage needs the node rows of ITS side (the ranking entry points prepare the query side before the
    // database has been embedded)
    const bool need_q = M.hist_bins > 0 && (stages & (SCORE_PREP_Q | SCORE_MAIN | SCORE_COEF_ONLY));
    const bool need_db = M.hist_bins > 0 && (stages & (SCORE_PREP_DB | SCORE_MAIN));
    if ((need_q && !h->hside[0].rows) || (need_db && !h->hside[1].rows))
      return fail(h, "the model has a Histogram layer: all-pairs scoring needs the node rows of both sides "
                     "(sgnn_score_all_pairs_nodes, or the sgnn_rank_* entry points)");
    // histogram on the tensor cores (sgnn_hist_tc.cuh) whenever the shapes fit, else the CUDA-core kernel counts.
    // A side's operand is prepared whenever THAT side fits (the other side may not be known yet); the scorer stage
    // decides with both.
    // the rest of the chain (NTN + Dense head) on the tensor-core scorer when the head has the shape that kernel
    // serves; its operands are prepared beside the CUDA-core tables (which side decides is only known at the scorer)
    const bool head_tc = head_tc_enabled(h);
    const bool tc_on = hist_tc_enabled(h);
    const bool tc_q = tc_on && need_q && sgnn_hist_tc_supported(M.node_dim, M.hist_bins, h->hside[0].max_nodes, 1);
    const bool tc_db = tc_on && need_db && sgnn_hist_tc_supported(M.node_dim, M.hist_bins, 1, h->hside[1].max_nodes);
    const bool tc_hist = tc_q && tc_db &&
                         sgnn_hist_tc_supported(M.node_dim, M.hist_bins, h->hside[0].max_nodes, h->hside[1].max_nodes);
    // The database side's NTN operand (tc_prep_kernel) is read by the head kernel only.  When this call also runs the
    // scorer, it goes to the aux stream, beside the histogram's operand kernel and the histogram kernel, and is joined
    // right before the head kernel: one kernel and one launch gap less in front of the 0.54 ms histogram kernel.
    bool db_aside = false;
    if (head_tc) {
      int prep = stages & (SCORE_PREP_Q | SCORE_PREP_DB | SCORE_COEF_ONLY | SCORE_COEF_READY);
      db_aside = (prep & SCORE_PREP_DB) && (stages & SCORE_MAIN) && tc_hist && h->aux_stream && h->aux_stream != st &&
                 h->chunk_ev[6] && h->chunk_ev[7];
      if (db_aside) prep &= ~SCORE_PREP_DB;
      if (prep && score_all_pairs_tc(h, emb_q, m, emb_db, n, sim_dev, ld, st, prep, nullptr, true)) return 1;
      if (db_aside) {
        CUDA_OK(h, cudaEventRecord(h->chunk_ev[6], st));
        CUDA_OK(h, cudaStreamWaitEvent(h->aux_stream, h->chunk_ev[6], 0));
        if (score_all_pairs_tc(h, emb_q, m, emb_db, n, sim_dev, ld, h->aux_stream, SCORE_PREP_DB, nullptr, true)) return 1;
        CUDA_OK(h, cudaEventRecord(h->chunk_ev[7], h->aux_stream));
      }
    }
    HistTcGeom HG;
    if (tc_q || tc_db) {
      const HistTcWs& t = h->ht;
      const int mx_db = tc_db ? h->hside[1].max_nodes : 1;
      HG.m = m; HG.n = n; HG.d = M.node_dim; HG.ksteps = t.ksteps; HG.parts = t.parts; HG.bins = M.hist_bins;
      HG.S = 129 - mx_db;
      HG.max_btiles = (int)std::min<int64_t>(t.cap_b_tiles, ht_max_btiles(m));
      HG.max_atiles = (int)std::min<int64_t>(t.cap_a_tiles, (int64_t)n * mx_db / HG.S + 1);
      if (m > t.cap_q || ht_max_btiles(m) > t.cap_b_tiles || (int64_t)m * n * HT_CNT_BYTES > t.cap_cnt || t.cap_a_tiles == 0 ||
          (tc_db && (h->hside[1].total_rows >= 0 ? h->hside[1].total_rows : (int64_t)n * mx_db) / HG.S + 1 > t.cap_a_tiles))
        return fail(h, "histogram workspace too small: call sgnn_reserve(max_nodes >= total_nodes, max_queries, max_db) first");
    }
    // the per-query P' / C tables feed the CUDA-core head kernels only: with the tensor-core head they are built
    // lazily by the scorer stage, in the rare case that the database side turns out not to fit (22 us off the
    // query branch of the ranking step)
    const bool lazy_tables = head_tc && (M.hist_bins == 0 || tc_q);
    if (stages & SCORE_PREP_Q) {
      if (!lazy_tables) {
        allpairs_prep_kernel<<<(m + 3) / 4, 128, 0, st>>>(M, h->d_params, emb_q, m, h->dep, h->d_P, h->d_C);
        CUDA_OK(h, cudaGetLastError());
      }
      // (with SCORE_COEF_READY the bucket sort has run in the SCORE_COEF_ONLY stage, beside the query embedding)
      if (tc_q) CUDA_OK(h, sgnn_hist_tc_prep_q(HG, h->hside[0].rows, h->hside[0].node_off, h->ht, st,
                                                (stages & SCORE_COEF_READY) ? 2 : 0));
    }
    if ((stages & SCORE_COEF_ONLY) && tc_q)
      CUDA_OK(h, sgnn_hist_tc_prep_q(HG, nullptr, h->hside[0].node_off, h->ht, st, 1));
    if ((stages & SCORE_PREP_DB) && tc_db)
      CUDA_OK(h, sgnn_hist_tc_prep_db(HG, h->hside[1].rows, h->hside[1].node_off, h->ht, st));
    if (!(stages & SCORE_MAIN)) return 0;
    cudaError_t e;
    if (head_tc && (M.hist_bins == 0 || tc_hist)) {
      // the counts leave the histogram kernel as fp16 operand tiles: the head kernel multiplies them with the layer-0
      // rows of the bins on the tensor core (SGNN_HEAD_HMMA=0: bytes, multiplied by the epilogue threads)
      const bool tiled = tc_hist && head_hmma_enabled() && ht_cnt_tiled_bytes(m, n) <= h->ht.cap_cnt;
      if (tiled) HG.cnt_m_pad = (m + 7) / 8 * 8;
      if (tc_hist) CUDA_OK(h, sgnn_hist_tc_run(HG, h->ht, h->hside[1].node_off, h->sm_count, st));
      if (db_aside) CUDA_OK(h, cudaStreamWaitEvent(st, h->chunk_ev[7], 0));   // join: the head kernel reads the NTN operand
      return score_all_pairs_tc(h, emb_q, m, emb_db, n, sim_dev, ld, st, SCORE_MAIN, smax, true,
                                tc_hist ? h->ht.cnt : nullptr, h->hside[0].node_off, h->hside[1].node_off, tiled);
    }
    if (lazy_tables) {   // the tensor-core head was planned but this database does not fit it
      allpairs_prep_kernel<<<(m + 3) / 4, 128, 0, st>>>(M, h->d_params, emb_q, m, h->dep, h->d_P, h->d_C);
      CUDA_OK(h, cudaGetLastError());
    }
    if (tc_hist) {
      CUDA_OK(h, sgnn_hist_tc_run(HG, h->ht, h->hside[1].node_off, h->sm_count, st));
      e = sgnn_launch_allpairs_head_counts(M, h->dep, h->d_params, h->d_P, h->d_C, emb_db, h->hside[0].node_off,
                                           h->hside[1].node_off, h->ht.cnt, m, n, sim_dev, ld, h->sm_count,
                                           h->smem_optin, st);
    } else {
      e = sgnn_launch_allpairs_head(M, h->dep, h->d_params, h->d_P, h->d_C, emb_db, h->hside[0].rows,
                                    h->hside[0].node_off, h->hside[0].max_nodes, h->hside[1].rows,
                                    h->hside[1].node_off, h->hside[1].max_nodes, m, n, sim_dev, ld,
                                    h->sm_count, h->smem_optin, st);
    }
    if (e == cudaErrorInvalidValue)
      return fail(h, "all-pairs head kernel: unsupported shape (head widths and node_dim must be <= 32, embedding <= 32, "
                     "and one graph's node rows must fit shared memory)");
    if (e != cudaSuccess) return fail(h, std::string("all-pairs head kernel: ") + cudaGetErrorString(e));
    return 0;
  }
  if (h->cfg.precision != SGNN_PREC_FP32) return score_all_pairs_tc(h, emb_q, m, emb_db, n, sim_dev, ld, st, stages, smax);
  const int DEP = h->dep;
  if (stages & SCORE_PREP_Q) {
    allpairs_prep_kernel<<<(m + 3) / 4, 128, 0, st>>>(M, h->d_params, emb_q, m, DEP, h->d_P, h->d_C);
    CUDA_OK(h, cudaGetLastError());
  }
  if (!(stages & SCORE_MAIN)) return 0;
  constexpr int QC = 8;
  const int gx = (n + 127) / 128;
  // enough query blocks to fill the machine a few times over
  int gy = std::max(1, std::min((m + QC - 1) / QC, (h->sm_count * 8 + gx - 1) / gx));
  const int qpb = ((m + gy - 1) / gy + QC - 1) / QC * QC;
  gy = (m + qpb - 1) / qpb;
  const size_t smem = ((size_t)QC * K * DEP + QC * K + K) * sizeof(float);
  dim3 grid(gx, gy);
  if (DEP == 16)
    allpairs_fp32_kernel<16, QC><<<grid, 128, smem, st>>>(M, h->d_params, h->d_P, h->d_C, emb_db, m, n, qpb, sim_dev, ld);
  else if (DEP == 32)
    allpairs_fp32_kernel<32, QC><<<grid, 128, smem, st>>>(M, h->d_params, h->d_P, h->d_C, emb_db, m, n, qpb, sim_dev, ld);
  else
    allpairs_fp32_kernel<64, QC><<<grid, 128, smem, st>>>(M, h->d_params, h->d_P, h->d_C, emb_db, m, n, qpb, sim_dev, ld);
  CUDA_OK(h, cudaGetLastError());
  return 0;
}
}  // namespace

namespace {
uint64_t fnv1a(const void* data, size_t bytes) {
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint64_t x = 1469598103934665603ull;
  for (size_t i = 0; i < bytes; ++i) x = (x ^ p[i]) * 1099511628211ull;
  return x ? x : 1;
}

void key_batch(uint64_t* k, const SgnnGraphBatch* b) {
  k[0] = (uint64_t)(uintptr_t)b->node_off; k[1] = (uint64_t)(uintptr_t)b->labels;
  k[2] = (uint64_t)(uintptr_t)b->rowptr;   k[3] = (uint64_t)(uintptr_t)b->col;
  k[4] = (uint64_t)(uintptr_t)b->val;
  k[5] = ((uint64_t)(uint32_t)b->n_graphs << 32) | (uint32_t)b->max_nodes;
  k[6] = ((uint64_t)(uint32_t)b->total_nodes << 32) | (uint32_t)b->nnz;
}

// run `body` (stream work only, starting and ending on `st`) eagerly, or capture / replay it
template <typename F>
int run_cached(SgnnHandle h, GraphCache& cache, uint64_t key, cudaStream_t st, F body) {
  if (!h->use_graphs || st == nullptr || st == cudaStreamLegacy) return body();
  GraphSlot& slot = cache.find(key);
  if (slot.exec && slot.key == key) {
    CUDA_OK(h, cudaGraphLaunch(slot.exec, st));
    return 0;
  }
  if (!(slot.seen && slot.key == key)) {   // first sighting of these arguments: run eagerly
    slot.key = key; slot.seen = true;
    return body();
  }
  if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    h->use_graphs = false;
    return body();
  }
  const int rc = body();
  cudaGraph_t g = nullptr;
  const cudaError_t e = cudaStreamEndCapture(st, &g);
  if (rc != 0 || e != cudaSuccess || g == nullptr) {
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    h->use_graphs = false;
    return rc != 0 ? rc : body();
  }
  const cudaError_t e2 = cudaGraphInstantiate(&slot.exec, g, 0);
  cudaGraphDestroy(g);
  if (e2 != cudaSuccess) {
    cudaGetLastError();
    slot.exec = nullptr;
    h->use_graphs = false;
    return body();
  }
  CUDA_OK(h, cudaGraphLaunch(slot.exec, st));
  return 0;
}
}  // namespace

namespace {
// The Padding pool appends max_in_dims - N rows (layers.py:223-227): tf.pad raises for N > max_in_dims, and so do the
// entry points that take a batch (the kernels would otherwise pool the first max_in_dims nodes only).
int padding_guard(SgnnHandle h, const SgnnGraphBatch* b) {
  if (h->M.pool_kind == SGNN_LAYER_PADDING && b->n_graphs > 0 && b->max_nodes > h->M.pool_aux)
    return fail(h, "Padding: a graph of the batch has " + std::to_string(b->max_nodes) + " nodes, more than max_in_dims=" +
                       std::to_string(h->M.pool_aux) + " (Paddings must be non-negative, layers.py:223-227)");
  return 0;
}

// pooled embeddings AND last node-level rows of a batch in one pass (node_out may be NULL)
int embed_nodes(SgnnHandle h, const SgnnGraphBatch* b, float* emb_dev, float* node_out, cudaStream_t st) {
  if (b->n_graphs == 0) return 0;
  if (padding_guard(h, b)) return 1;
  // opt-in (sgnn_set_embed_tc / SGNN_EMBED_TC=1): the dense GCN layers as tcgen05 GEMMs over tiles of packed graphs.
  // Parity-green, but for graphs of ~10 nodes SLOWER than the CUDA-core tile kernel, hence not the default.
  static const bool etc_env = [] { const char* e = getenv("SGNN_EMBED_TC"); return e && e[0] == '1'; }();
  if ((etc_env || h->embed_tc) && h->cfg.precision != SGNN_PREC_FP32 && sgnn_embed_tc_supported(h->M, b->max_nodes)) {
    const cudaError_t e = sgnn_launch_embed_tc(h->M, to_dev(b), b->max_nodes, b->total_nodes, h->d_params, emb_dev, node_out,
                                               h->ht.status, h->sm_count, h->smem_optin, st);
    if (e == cudaSuccess) return 0;
    if (e != cudaErrorInvalidValue) return fail(h, std::string("embedding (tensor-core kernel): ") + cudaGetErrorString(e));
    cudaGetLastError();
  }
  if (b->max_nodes <= 32) {
    const cudaError_t e = sgnn_launch_embed_tile(h->M, to_dev(b), b->max_nodes, h->d_params, emb_dev, node_out, nullptr,
                                                 0, 0, nullptr, 0, 0, 0, 0, h->sm_count, h->smem_optin, st);
    if (e == cudaSuccess) return 0;
    if (e != cudaErrorInvalidValue) return fail(h, std::string("embedding (tile kernel): ") + cudaGetErrorString(e));
    cudaGetLastError();
  }
  if (b->total_nodes > h->cap_rows)
    return fail(h, "node-row workspace too small: call sgnn_reserve(max_nodes >= total_nodes) first");
  const cudaError_t e = sgnn_launch_embed_large(h->M, to_dev(b), h->d_params, h->d_rowsA, h->d_rowsB, emb_dev, node_out,
                                                h->sm_count, h->smem_optin, st);
  if (e != cudaSuccess) return fail(h, std::string("embedding (large-graph path): ") + cudaGetErrorString(e));
  return 0;
}

// embedding of one side (0 = queries, 1 = database) of a ranking entry point; with a Histogram head the
// side's node rows are kept for the all-pairs head kernel
// *fused (optional): the tile kernel also wrote the side's operand of the tensor-core scorer (database: A tiles,
// queries: B tiles), so the caller skips SCORE_PREP_DB / SCORE_PREP_Q
int embed_side(SgnnHandle h, const SgnnGraphBatch* b, float* emb_dev, int side, cudaStream_t st, bool* fused = nullptr,
               const SgnnGraphBatch* beside = nullptr) {
  if (fused) *fused = false;
  if (padding_guard(h, b)) return 1;
  // Database side: on by default (one kernel less on the critical chain).  Query side: measured neutral (the
  // query embedding gets heavier by what the two small operand kernels cost, 70.9 vs 70.2 us per step), so it stays
  // an opt-in (SGNN_FUSED_QUERY_PREP=1) and the default query branch is embed -> tc_coef -> tc_prep_q.
  static const bool no_fuse = getenv("SGNN_NO_FUSED_PREP") != nullptr;
  static const bool fuse_q = getenv("SGNN_FUSED_QUERY_PREP") != nullptr;
  if (fused && !no_fuse && (side == 1 || fuse_q) && h->M.n_head == 0 && h->cfg.precision != SGNN_PREC_FP32 &&
      b->n_graphs > 0 && b->max_nodes <= 32) {
    const ModelDev& M = h->M;
    const bool ntn = M.pair_kind == SGNN_LAYER_NTN;
    const int K = ntn ? M.K : 1;
    const int kc = tc_cols_per_query(K, ntn && M.pair_act == SGNN_ACT_RELU), qh = 128 / kc;
    const int kch = round_up(M.emb_dim + 1, 16) / 8, parts = h->cfg.precision == SGNN_PREC_BF16 ? 1 : 2;
    const int64_t need = side == 1 ? (int64_t)((b->n_graphs + 127) / 128) * parts * kch * 1024
                                   : (int64_t)((b->n_graphs + qh - 1) / qh) * parts * kch * (TC_NPAD * 8);
    if (need <= (side == 1 ? h->cap_Atc : h->cap_Btc)) {
      const cudaError_t e = sgnn_launch_embed_tile(M, to_dev(b), b->max_nodes, h->d_params, emb_dev, nullptr,
                                                   side == 1 ? h->d_Atc : nullptr, kch, parts,
                                                   side == 0 ? h->d_Btc : nullptr, K, kc, beside ? beside->n_graphs : 0,
                                                   beside ? beside->max_nodes : 0, h->sm_count, h->smem_optin, st);
      if (e == cudaSuccess) { *fused = true; return 0; }
      if (e != cudaErrorInvalidValue) return fail(h, std::string("embedding (tile kernel): ") + cudaGetErrorString(e));
      cudaGetLastError();
    }
  }
  if (h->M.hist_bins == 0) return sgnn_embed_graphs(h, b, emb_dev, st);
  if (b->total_nodes > h->cap_hnodes)
    return fail(h, "node-row workspace of the histogram head too small: call sgnn_reserve(max_nodes >= total_nodes) first");
  if (embed_nodes(h, b, emb_dev, h->d_hnodes[side], st)) return 1;
  h->hside[side].rows = h->d_hnodes[side];
  h->hside[side].node_off = b->node_off;
  h->hside[side].max_nodes = b->max_nodes;
  h->hside[side].total_rows = b->total_nodes;
  return 0;
}
}  // namespace

extern "C" {

int sgnn_abi_version(void) { return SGNN_ABI_VERSION; }

const char* sgnn_last_error(SgnnHandle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int sgnn_create(const SgnnConfig* cfg, SgnnHandle* out) {
  if (!cfg || !out) return fail(nullptr, "sgnn_create: null argument");
  *out = nullptr;
  SgnnHandle h = new SgnnHandle_t();
  h->cfg = *cfg;
  std::string err;
  if (build_model(*cfg, h->M, h->table, err)) { delete h; return fail(nullptr, err); }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    delete h;
    return fail(nullptr, std::string("no CUDA device: the sm_100a path has no CPU fallback (") +
                             cudaGetErrorString(e) + ")");
  }
  h->device = cfg->device;
  if (const char* ng = getenv("SGNN_NO_GRAPHS")) h->use_graphs = !(ng[0] && ng[0] != '0');
  if (cudaSetDevice(h->device) != cudaSuccess) { delete h; return fail(nullptr, "cudaSetDevice failed"); }
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, h->device);
  cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
  const int64_t np = h->M.n_params;
  auto bail = [&](const char* what) {
    g_create_error = std::string(what) + ": " + h->err;
    sgnn_destroy(h);
    return 1;
  };
  if (dev_alloc(h, &h->d_params, np) || dev_alloc(h, &h->d_grads, np) || dev_alloc(h, &h->d_m, np) ||
      dev_alloc(h, &h->d_v, np) || dev_alloc(h, &h->d_loss, 4))
    return bail("allocating parameter buffers");
  cudaMemset(h->d_params, 0, np * sizeof(float));
  cudaMemset(h->d_grads, 0, np * sizeof(float));
  cudaMemset(h->d_m, 0, np * sizeof(float));
  cudaMemset(h->d_v, 0, np * sizeof(float));
  cudaMemset(h->d_loss, 0, 4 * sizeof(float));
  // the side stream carries the (small) query branch of the ranking step next to the (large) database branch: at
  // the highest priority its CTAs are placed first instead of queueing behind the database embedding's
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithPriority(&h->copy_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
      cudaStreamCreateWithPriority(&h->aux_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess)
    return bail("cudaStreamCreate");
  for (auto& ev : h->chunk_ev)
    if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return bail("cudaEventCreate");
  h->dep = h->M.emb_dim <= 16 ? 16 : (h->M.emb_dim <= 32 ? 32 : 64);
  {
    const ModelDev& M = h->M;
    const bool ntn = M.pair_kind == SGNN_LAYER_NTN;
    const int kc = tc_cols_per_query(ntn ? M.K : 1, ntn && M.pair_act == SGNN_ACT_RELU);
    const int64_t cols = (int64_t)(round_up(M.emb_dim + 1, 16) / 8) * 8 * kc;
    if (cudaMalloc(reinterpret_cast<void**>(&h->d_coef), (size_t)((M.emb_dim + 1) * cols) * sizeof(float)) != cudaSuccess)
      h->d_coef = nullptr;   // the staged single-kernel query prep is used instead
  }
  if (plan_launches(h)) return bail("planning launches");
  if (cfg->precision != SGNN_PREC_FP32 && sgnn_tc_head_supported(h->M)) {
    h->head_slot = acquire_head_slot(h->device);
    if (h->head_slot >= 0 &&
        cudaMalloc(reinterpret_cast<void**>(&h->d_head_staging), (size_t)sgnn_tc_head_staging_floats() * sizeof(float)) != cudaSuccess) {
      release_head_slot(h->device, h->head_slot);
      h->head_slot = -1; h->d_head_staging = nullptr;
    }
  }
  *out = h;
  return 0;
}

int sgnn_destroy(SgnnHandle h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->d_params); cudaFree(h->d_grads); cudaFree(h->d_m); cudaFree(h->d_v); cudaFree(h->d_loss);
  cudaFree(h->d_coef); cudaFree(h->d_P); cudaFree(h->d_C); cudaFree(h->d_Atc); cudaFree(h->d_Btc); cudaFree(h->d_smax);
  cudaFree(h->d_rowsA); cudaFree(h->d_rowsB); cudaFree(h->d_hnodes[0]); cudaFree(h->d_hnodes[1]);
  release_head_slot(h->device, h->head_slot);
  h->head_slot = -1;
  cudaFree(h->d_head_staging);
  cudaFree(h->ht.At); cudaFree(h->ht.Bt); cudaFree(h->ht.qmeta); cudaFree(h->ht.ameta); cudaFree(h->ht.status);
  cudaFree(h->ht.cnt);
  cudaFree(h->d_adam_t);
  cudaFree(h->d_arm_ws);
  cudaFree(h->hp.At); cudaFree(h->hp.Bt); cudaFree(h->hp.tile_off); cudaFree(h->hp.unit_off); cudaFree(h->hp.cnt);
  HostWs& w = h->ws;
  for (int s = 0; s < 2; ++s) {
    cudaFree(w.node_off[s]); cudaFree(w.labels[s]); cudaFree(w.rowptr[s]); cudaFree(w.col[s]); cudaFree(w.val[s]);
    cudaFree(w.arena[s]);
  }
  cudaFree(w.g1); cudaFree(w.g2); cudaFree(w.target); cudaFree(w.raw); cudaFree(w.sim_pairs); cudaFree(w.emb);
  cudaFree(w.emb_q); cudaFree(w.emb_db); cudaFree(w.sim); cudaFree(w.topk_vals); cudaFree(w.topk_ids);
  if (h->stream) cudaStreamDestroy(h->stream);
  h->g_rank_device.drop();
  h->g_rank_host.drop();
  h->g_rank_sharded.drop();
  sgnn_peer_destroy(h);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->aux_stream) cudaStreamDestroy(h->aux_stream);
  for (auto& ev : h->chunk_ev) if (ev) cudaEventDestroy(ev);
  delete h;
  return 0;
}

int64_t sgnn_param_count(SgnnHandle h) { return h ? h->M.n_params : -1; }

int sgnn_param_table(SgnnHandle h, SgnnParamInfo* out, int32_t capacity) {
  if (!h) return -1;
  const int n = (int)h->table.size();
  if (out) for (int i = 0; i < n && i < capacity; ++i) out[i] = h->table[i];
  return n;
}

int sgnn_embed_dim(SgnnHandle h) { return h ? h->M.emb_dim : -1; }

int sgnn_set_params(SgnnHandle h, const float* host_params) {
  if (!h || !host_params) return fail(h, "sgnn_set_params: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  const size_t bytes = (size_t)h->M.n_params * sizeof(float);
  CUDA_OK(h, cudaMemcpy(h->d_params, host_params, bytes, cudaMemcpyHostToDevice));
  CUDA_OK(h, cudaMemset(h->d_m, 0, bytes));
  CUDA_OK(h, cudaMemset(h->d_v, 0, bytes));
  CUDA_OK(h, cudaMemset(h->d_grads, 0, bytes));
  h->adam_t = 0;
  if (h->d_adam_t) CUDA_OK(h, cudaMemset(h->d_adam_t, 0, sizeof(long long)));
  return 0;
}

int sgnn_get_params(SgnnHandle h, float* host_params) {
  if (!h || !host_params) return fail(h, "sgnn_get_params: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  CUDA_OK(h, cudaDeviceSynchronize());
  CUDA_OK(h, cudaMemcpy(host_params, h->d_params, (size_t)h->M.n_params * sizeof(float), cudaMemcpyDeviceToHost));
  return 0;
}

int sgnn_init_params(SgnnHandle h, uint64_t seed) {
  if (!h) return 1;
  std::vector<float> p(h->M.n_params, 0.f);
  uint64_t s = seed;
  for (const SgnnParamInfo& t : h->table) {
    if (strcmp(t.name, "bias") == 0) continue;                       // zeros, inits.py:18-21
    const double r = std::sqrt(6.0 / (t.shape[0] + t.shape[1]));      // glorot, inits.py:11-15
    int64_t cnt = (int64_t)t.shape[0] * (t.ndim > 1 ? t.shape[1] : 1) * (t.ndim > 2 ? t.shape[2] : 1);
    for (int64_t i = 0; i < cnt; ++i) {
      const double u = (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0);
      p[t.offset + i] = (float)((2.0 * u - 1.0) * r);
    }
  }
  return sgnn_set_params(h, p.data());
}

int sgnn_device_buffers(SgnnHandle h, float** params, float** grads, float** adam_m, float** adam_v) {
  if (!h) return 1;
  if (params) *params = h->d_params;
  if (grads) *grads = h->d_grads;
  if (adam_m) *adam_m = h->d_m;
  if (adam_v) *adam_v = h->d_v;
  return 0;
}

int sgnn_reserve(SgnnHandle h, int64_t max_graphs, int64_t max_nodes, int64_t max_nnz, int64_t max_pairs,
                 int64_t max_queries, int64_t max_db) {
  if (!h) return 1;
  CUDA_OK(h, cudaSetDevice(h->device));
  CUDA_OK(h, cudaDeviceSynchronize());
  h->g_rank_device.drop();
  h->g_rank_host.drop();
  h->g_rank_sharded.drop();
  HostWs& w = h->ws;
  const int E = h->M.emb_dim;
  if (max_graphs > w.cap_graphs || max_nodes > w.cap_nodes || max_nnz > w.cap_nnz) {
    w.cap_graphs = std::max(w.cap_graphs, max_graphs);
    w.cap_nodes = std::max(w.cap_nodes, max_nodes);
    w.cap_nnz = std::max(w.cap_nnz, max_nnz);
    w.arena_cap = 4 * (w.cap_graphs + 1 + w.cap_nodes + w.cap_nodes + 1 + 2 * w.cap_nnz) + 5 * 256;
    for (int s = 0; s < 2; ++s) {
      if (dev_alloc(h, &w.node_off[s], w.cap_graphs + 1) || dev_alloc(h, &w.labels[s], w.cap_nodes) ||
          dev_alloc(h, &w.rowptr[s], w.cap_nodes + 1) || dev_alloc(h, &w.col[s], w.cap_nnz) ||
          dev_alloc(h, &w.val[s], w.cap_nnz) || dev_alloc(h, &w.arena[s], w.arena_cap))
        return 1;
    }
    if (dev_alloc(h, &w.emb, w.cap_graphs * E) || dev_alloc(h, &w.emb_q, w.cap_graphs * E) ||
        dev_alloc(h, &w.emb_db, w.cap_graphs * E))
      return 1;
  }
  if (max_nodes > h->cap_rows) {   // node-row workspace of the large-graph (N > 32) embedding kernel
    const int64_t fl = max_nodes * (int64_t)sgnn_large_row_stride(h->M);
    if (dev_alloc(h, &h->d_rowsA, fl) || dev_alloc(h, &h->d_rowsB, fl)) return 1;
    h->cap_rows = max_nodes;
  }
  if (h->M.hist_bins > 0 && max_nodes > h->cap_hnodes) {
    const int64_t fl = max_nodes * (int64_t)h->M.node_dim;
    if (dev_alloc(h, &h->d_hnodes[0], fl) || dev_alloc(h, &h->d_hnodes[1], fl)) return 1;
    h->cap_hnodes = max_nodes;
  }
  if (hist_tc_enabled(h)) {   // tensor-core histogram: operand tiles of both sides, tile tables, count matrix
    HistTcWs& t = h->ht;
    t.parts = h->cfg.precision == SGNN_PREC_BF16 ? 1 : 3;
    t.ksteps = (h->M.node_dim + 15) / 16;
    const int64_t tile = (int64_t)t.parts * t.ksteps * 4096;
    const int64_t at = max_nodes > 0 ? ht_max_atiles(max_nodes) : 0, bt = max_queries > 0 ? ht_max_btiles(max_queries) : 0;
    uint8_t* p8 = nullptr;
    if (at > t.cap_a_tiles) {
      p8 = static_cast<uint8_t*>(t.At); if (dev_alloc(h, &p8, at * tile)) return 1; t.At = p8;
      if (dev_alloc(h, &t.ameta, at + 1)) return 1;
      t.cap_a_tiles = at;
    }
    if (bt > t.cap_b_tiles || max_queries > t.cap_q) {
      t.cap_b_tiles = std::max(t.cap_b_tiles, bt);
      t.cap_q = std::max(t.cap_q, max_queries);
      p8 = static_cast<uint8_t*>(t.Bt); if (dev_alloc(h, &p8, t.cap_b_tiles * tile)) return 1; t.Bt = p8;
      if (dev_alloc(h, &t.qmeta, (int64_t)ht_qmeta_ints((int)t.cap_q, (int)t.cap_b_tiles))) return 1;
    }
    if (!t.status) {
      if (dev_alloc(h, &t.status, 4)) return 1;
      CUDA_OK(h, cudaMemset(t.status, 0, 16));
    }
    const int64_t need_cnt = std::max<int64_t>(max_queries * max_db * HT_CNT_BYTES, ht_cnt_tiled_bytes(max_queries, max_db));
    if (need_cnt > t.cap_cnt) {
      if (dev_alloc(h, &t.cnt, need_cnt)) return 1;
      t.cap_cnt = need_cnt;
    }
  }
  if (max_pairs > w.cap_pairs) {
    w.cap_pairs = max_pairs;
    if (dev_alloc(h, &w.g1, max_pairs) || dev_alloc(h, &w.g2, max_pairs) || dev_alloc(h, &w.target, max_pairs) ||
        dev_alloc(h, &w.raw, max_pairs) || dev_alloc(h, &w.sim_pairs, max_pairs))
      return 1;
  }
  if (max_queries > h->cap_P) {
    h->cap_P = max_queries;
    const int K = h->M.pair_kind == SGNN_LAYER_NTN ? h->M.K : 1;
    if (dev_alloc(h, &h->d_P, max_queries * K * 64) || dev_alloc(h, &h->d_C, max_queries * K)) return 1;
  }
  {
    const ModelDev& M = h->M;
    const int K = M.pair_kind == SGNN_LAYER_NTN ? M.K : 1;
    const bool relu = M.pair_kind == SGNN_LAYER_NTN && M.pair_act == SGNN_ACT_RELU;
    const int64_t kch = round_up(M.emb_dim + 1, 16) / 8, qt = 128 / tc_cols_per_query(K, relu);
    const int64_t need_a = ((max_db + 127) / 128) * 2 * kch * 1024;
    const int64_t need_b = ((max_queries + qt - 1) / qt) * 2 * kch * (TC_NPAD * 8);
    if (need_a > h->cap_Atc) {
      if (dev_alloc(h, &h->d_Atc, need_a)) return 1;
      CUDA_OK(h, cudaMemset(h->d_Atc, 0, need_a * sizeof(__nv_bfloat16)));
      h->cap_Atc = need_a;
    }
    if (need_b > h->cap_Btc) {
      if (dev_alloc(h, &h->d_Btc, need_b)) return 1;
      CUDA_OK(h, cudaMemset(h->d_Btc, 0, need_b * sizeof(__nv_bfloat16)));
      h->cap_Btc = need_b;
    }
    const int64_t need_s = max_queries * (int64_t)tc_n_slices((int)max_db);
    if (need_s > h->cap_smax) {
      if (dev_alloc(h, &h->d_smax, need_s)) return 1;
      h->cap_smax = need_s;
    }
  }
  if (max_queries * max_db > w.cap_q * w.cap_db || max_queries > w.cap_q) {
    w.cap_q = std::max(w.cap_q, max_queries);
    w.cap_db = std::max(w.cap_db, max_db);
    if (dev_alloc(h, &w.sim, w.cap_q * w.cap_db)) return 1;
    w.cap_topk = w.cap_q * 64;
    // ids and values of one call lie back to back in the ids allocation (one device -> host copy when the
    // caller's two host arrays are contiguous too)
    if (dev_alloc(h, &w.topk_ids, 2 * w.cap_topk) || dev_alloc(h, &w.topk_vals, w.cap_topk)) return 1;
  }
  return 0;
}

int sgnn_embed_graphs(SgnnHandle h, const SgnnGraphBatch* b, float* emb_dev, void* stream) {
  if (!h || !b || !emb_dev) return fail(h, "sgnn_embed_graphs: null argument");
  if (b->n_graphs == 0) return 0;
  if (padding_guard(h, b)) return 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (b->max_nodes > 32) {   // Web-shaped / large graphs: CTA per graph, rows in the global workspace
    if (b->total_nodes > h->cap_rows)
      return fail(h, "node-row workspace too small for graphs with more than 32 nodes: call sgnn_reserve(max_nodes >= total_nodes) first");
    const cudaError_t e = sgnn_launch_embed_large(h->M, to_dev(b), h->d_params, h->d_rowsA, h->d_rowsB, emb_dev,
                                                  nullptr, h->sm_count, h->smem_optin, st);
    if (e != cudaSuccess) return fail(h, std::string("sgnn_embed_graphs (large-graph path): ") + cudaGetErrorString(e));
    return 0;
  }
  {
    const cudaError_t e = sgnn_launch_embed_tile(h->M, to_dev(b), b->max_nodes, h->d_params, emb_dev, nullptr, nullptr, 0, 0, nullptr, 0, 0, 0, 0, h->sm_count,
                                                 h->smem_optin, st);
    if (e == cudaSuccess) return 0;
    if (e != cudaErrorInvalidValue) return fail(h, std::string("sgnn_embed_graphs: ") + cudaGetErrorString(e));
    cudaGetLastError();   // model too wide for the tile kernel: warp-per-graph kernel below
  }
  const int v = b->max_nodes <= 16 ? 0 : 1;
  const int wpb = h->embed_wpb[v];
  const int blocks = std::min((b->n_graphs + wpb - 1) / wpb, h->sm_count * 4);
  if (v == 0)
    embed_kernel<16><<<blocks, wpb * 32, h->embed_smem[0], st>>>(h->M, to_dev(b), h->d_params, emb_dev);
  else
    embed_kernel<32><<<blocks, wpb * 32, h->embed_smem[1], st>>>(h->M, to_dev(b), h->d_params, emb_dev);
  CUDA_OK(h, cudaGetLastError());
  return 0;
}

int sgnn_node_dim(SgnnHandle h) { return h ? h->M.node_dim : -1; }

int sgnn_node_embeddings(SgnnHandle h, const SgnnGraphBatch* b, float* node_out_dev, void* stream) {
  if (!h || !b || !node_out_dev) return fail(h, "sgnn_node_embeddings: null argument");
  if (b->n_graphs == 0) return 0;
  if (b->total_nodes > h->cap_rows)
    return fail(h, "node-row workspace too small: call sgnn_reserve(max_nodes >= total_nodes) first");
  const cudaError_t e = sgnn_launch_embed_large(h->M, to_dev(b), h->d_params, h->d_rowsA, h->d_rowsB, nullptr,
                                                node_out_dev, h->sm_count, h->smem_optin, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(h, std::string("sgnn_node_embeddings: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_embed_graphs_nodes(SgnnHandle h, const SgnnGraphBatch* b, float* emb_dev, float* node_out_dev, void* stream) {
  if (!h || !b || !emb_dev || !node_out_dev) return fail(h, "sgnn_embed_graphs_nodes: null argument");
  return embed_nodes(h, b, emb_dev, node_out_dev, (cudaStream_t)stream);
}

int sgnn_pair_histogram(SgnnHandle h, const float* node_emb_dev, const int32_t* node_off_dev, const int32_t* g1_dev,
                        const int32_t* g2_dev, int32_t n_pairs, int32_t bins, float* hist_dev, int32_t* counts_dev,
                        void* stream) {
  if (!h || !node_emb_dev || !node_off_dev || !g1_dev || !g2_dev) return fail(h, "sgnn_pair_histogram: null argument");
  if (bins < 1 || bins > 32) return fail(h, "sgnn_pair_histogram: bins must be in [1, 32]");
  if (n_pairs == 0) return 0;
  cudaError_t e = sgnn_launch_pair_histogram_tiled(node_emb_dev, h->M.node_dim, node_off_dev, g1_dev, g2_dev, n_pairs,
                                                   bins, hist_dev, counts_dev, h->sm_count, (cudaStream_t)stream);
  if (e == cudaErrorInvalidValue) {   // node_dim > 32: generic cell-per-lane kernel
    cudaGetLastError();
    e = sgnn_launch_pair_histogram(node_emb_dev, h->M.node_dim, node_off_dev, g1_dev, g2_dev, n_pairs, bins, hist_dev,
                                   counts_dev, h->sm_count, (cudaStream_t)stream);
  }
  if (e != cudaSuccess) return fail(h, std::string("sgnn_pair_histogram: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_pair_histogram_tc(SgnnHandle h, const float* node_emb_dev, const int32_t* node_off_dev, int32_t n_graphs,
                           int64_t total_nodes, const int32_t* g1_dev, const int32_t* g2_dev, int32_t n_pairs,
                           int32_t bins, float* hist_dev, int32_t* counts_dev, void* stream) {
  if (!h || !node_emb_dev || !node_off_dev || !g1_dev || !g2_dev) return fail(h, "sgnn_pair_histogram_tc: null argument");
  if (n_graphs < 1 || total_nodes < 0 || n_pairs < 0) return fail(h, "sgnn_pair_histogram_tc: bad sizes");
  if (h->cfg.precision == SGNN_PREC_FP32 || !sgnn_hist_pairs_tc_supported(h->M.node_dim, bins)) {
    h->err = "sgnn_pair_histogram_tc: needs a tensor-core precision, node rows of <= 16 values and <= 16 bins "
             "(use sgnn_pair_histogram)";
    return 4;
  }
  if (n_pairs == 0) return 0;
  HistPairsWs& w = h->hp;
  const int64_t tiles = hp_max_tiles(n_graphs, total_nodes);
  if (tiles * (int64_t)n_pairs >= (int64_t)1 << 31) return fail(h, "sgnn_pair_histogram_tc: too many (pair, tile) units");
  // workspace grown by the call (the first call of a size allocates; steady-state calls do not)
  if (tiles > w.cap_tiles) {
    uint8_t* p8 = static_cast<uint8_t*>(w.At); if (dev_alloc(h, &p8, tiles * 12288)) return 1; w.At = p8;
    p8 = static_cast<uint8_t*>(w.Bt); if (dev_alloc(h, &p8, tiles * 12288)) return 1; w.Bt = p8;
    w.cap_tiles = tiles;
  }
  if (n_graphs > w.cap_graphs) { if (dev_alloc(h, &w.tile_off, (int64_t)n_graphs + 1)) return 1; w.cap_graphs = n_graphs; }
  if (n_pairs > w.cap_pairs) { if (dev_alloc(h, &w.unit_off, (int64_t)n_pairs + 1)) return 1; w.cap_pairs = n_pairs; }
  int32_t* counts = counts_dev;
  if (!counts) {
    if ((int64_t)n_pairs * bins > w.cap_cnt) { if (dev_alloc(h, &w.cnt, (int64_t)n_pairs * bins)) return 1; w.cap_cnt = (int64_t)n_pairs * bins; }
    counts = w.cnt;
  }
  cudaError_t e = sgnn_hist_pairs_tc_run(h->M.node_dim, bins, node_emb_dev, node_off_dev, n_graphs, tiles, g1_dev, g2_dev,
                                         n_pairs, w, hist_dev, counts, h->sm_count, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(h, std::string("sgnn_pair_histogram_tc: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_score_pairs(SgnnHandle h, const float* emb_dev, const int32_t* g1, const int32_t* g2, int32_t n_pairs,
                     float* raw_dev, float* sim_dev, void* stream) {
  if (!h || !emb_dev || !g1 || !g2) return fail(h, "sgnn_score_pairs: null argument");
  if (h->M.n_head > 0)
    return fail(h, "sgnn_score_pairs: the model has a Dense / Histogram head after the NTN; use sgnn_forward_pairs or sgnn_eval_pairs");
  if (n_pairs == 0) return 0;
  const int blocks = std::min((n_pairs + 7) / 8, h->sm_count * 4);
  pair_score_kernel<<<blocks, 256, h->pair_smem, (cudaStream_t)stream>>>(h->M, h->d_params, emb_dev, g1, g2,
                                                                        n_pairs, raw_dev, sim_dev);
  CUDA_OK(h, cudaGetLastError());
  return 0;
}

static int launch_pairs(SgnnHandle h, bool bwd, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                        const float* target_sim, int32_t n_pairs, int32_t global_batch, double target_sum,
                        double target_sqsum, float* loss, float* raw, float* sim, void* stream, bool no_dropout = false);

int sgnn_forward_pairs(SgnnHandle h, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                       int32_t n_pairs, float* emb_ws_dev, float* raw_dev, float* sim_dev, void* stream) {
  if (h && h->M.n_head > 0) {   // head models: the per-pair kernel holds the node rows the histogram needs
    if (!b || !g1 || !g2) return fail(h, "sgnn_forward_pairs: null argument");
    return launch_pairs(h, false, b, g1, g2, nullptr, n_pairs, n_pairs, 0.0, 0.0, nullptr, raw_dev, sim_dev, stream, true);
  }
  if (sgnn_embed_graphs(h, b, emb_ws_dev, stream)) return 1;
  return sgnn_score_pairs(h, emb_ws_dev, g1, g2, n_pairs, raw_dev, sim_dev, stream);
}

int sgnn_score_all_pairs(SgnnHandle h, const float* emb_q, int32_t m, const float* emb_db, int32_t n,
                         float* sim_dev, int64_t ld, void* stream) {
  if (!h || !emb_q || !emb_db || !sim_dev) return fail(h, "sgnn_score_all_pairs: null argument");
  h->hside[0].rows = h->hside[1].rows = nullptr;
  return score_all_pairs_stages(h, emb_q, m, emb_db, n, sim_dev, ld, (cudaStream_t)stream, SCORE_ALL);
}

int sgnn_score_all_pairs_stages(SgnnHandle h, const float* emb_q, int32_t m, const float* emb_db, int32_t n,
                                float* sim_dev, int64_t ld, int32_t stages, void* stream) {
  if (!h || !emb_q || !emb_db || !sim_dev) return fail(h, "sgnn_score_all_pairs_stages: null argument");
  if (stages < 1 || stages > SGNN_STAGE_ALL) return fail(h, "sgnn_score_all_pairs_stages: bad stage mask");
  static_assert(SGNN_STAGE_PREP_QUERIES == SCORE_PREP_Q && SGNN_STAGE_PREP_DATABASE == SCORE_PREP_DB &&
                SGNN_STAGE_SCORE == SCORE_MAIN, "stage flags");
  return score_all_pairs_stages(h, emb_q, m, emb_db, n, sim_dev, ld, (cudaStream_t)stream, stages);
}

int sgnn_score_all_pairs_nodes(SgnnHandle h, const float* emb_q, const float* nodes_q, const int32_t* node_off_q,
                               int32_t max_nodes_q, int32_t m, const float* emb_db, const float* nodes_db,
                               const int32_t* node_off_db, int32_t max_nodes_db, int32_t n, float* sim_dev, int64_t ld,
                               void* stream) {
  if (!h || !emb_q || !emb_db || !sim_dev || !nodes_q || !node_off_q || !nodes_db || !node_off_db)
    return fail(h, "sgnn_score_all_pairs_nodes: null argument");
  h->hside[0].rows = nodes_q; h->hside[0].node_off = node_off_q; h->hside[0].max_nodes = max_nodes_q;
  h->hside[1].rows = nodes_db; h->hside[1].node_off = node_off_db; h->hside[1].max_nodes = max_nodes_db;
  h->hside[0].total_rows = h->hside[1].total_rows = -1;
  return score_all_pairs_stages(h, emb_q, m, emb_db, n, sim_dev, ld, (cudaStream_t)stream, SCORE_ALL);
}

int sgnn_hist_all_pairs(SgnnHandle h, const float* nodes_q, const int32_t* node_off_q, int32_t max_nodes_q, int32_t m,
                        const float* nodes_db, const int32_t* node_off_db, int32_t max_nodes_db, int32_t n,
                        uint8_t* counts_dev, int32_t stages, void* stream) {
  if (!h || !nodes_q || !node_off_q || !nodes_db || !node_off_db) return fail(h, "sgnn_hist_all_pairs: null argument");
  if (stages < 1 || stages > SGNN_STAGE_ALL) return fail(h, "sgnn_hist_all_pairs: bad stage mask");
  const ModelDev& M = h->M;
  if (!hist_tc_enabled(h)) return fail(h, "sgnn_hist_all_pairs: needs a Histogram model (bins <= 16) and a tensor-core precision");
  if (m == 0 || n == 0) return 0;
  if (!sgnn_hist_tc_supported(M.node_dim, M.hist_bins, max_nodes_q, max_nodes_db)) {
    fail(h, "sgnn_hist_all_pairs: graph sizes outside the tensor-core histogram (queries <= 15 nodes, database <= 64, product <= 255)");
    return 4;
  }
  const HistTcWs& t = h->ht;
  HistTcGeom HG;
  HG.m = m; HG.n = n; HG.d = M.node_dim; HG.ksteps = t.ksteps; HG.parts = t.parts; HG.bins = M.hist_bins;
  HG.S = 129 - max_nodes_db;
  HG.max_btiles = (int)std::min<int64_t>(t.cap_b_tiles, ht_max_btiles(m));
  HG.max_atiles = (int)std::min<int64_t>(t.cap_a_tiles, (int64_t)n * max_nodes_db / HG.S + 1);
  if (m > t.cap_q || ht_max_btiles(m) > t.cap_b_tiles || t.cap_a_tiles == 0 ||
      (!counts_dev && (int64_t)m * n * HT_CNT_BYTES > t.cap_cnt))
    return fail(h, "histogram workspace too small: call sgnn_reserve(max_nodes >= total_nodes, max_queries, max_db) first");
  cudaStream_t st = (cudaStream_t)stream;
  if ((stages & SGNN_STAGE_PREP_DATABASE) && (int64_t)n * max_nodes_db / HG.S + 1 > t.cap_a_tiles) {
    // the size-based bound does not fit the reserved tiles: look at the actual number of node rows (one 4-byte read;
    // callers that cannot synchronise reserve max_nodes >= n * max_nodes_db)
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cs);
    if (cs != cudaStreamCaptureStatusNone)
      return fail(h, "histogram workspace too small for n * max_nodes_db node rows (checked conservatively under stream capture)");
    int32_t total = 0;
    CUDA_OK(h, cudaMemcpyAsync(&total, node_off_db + n, sizeof(total), cudaMemcpyDeviceToHost, st));
    CUDA_OK(h, cudaStreamSynchronize(st));
    if ((int64_t)total / HG.S + 1 > t.cap_a_tiles)
      return fail(h, "histogram workspace too small: call sgnn_reserve(max_nodes >= database node rows) first");
  }
  HistTcWs w = t;
  if (counts_dev) w.cnt = counts_dev;
  if (stages & SGNN_STAGE_PREP_QUERIES) CUDA_OK(h, sgnn_hist_tc_prep_q(HG, nodes_q, node_off_q, w, st));
  if (stages & SGNN_STAGE_PREP_DATABASE) CUDA_OK(h, sgnn_hist_tc_prep_db(HG, nodes_db, node_off_db, w, st));
  if (stages & SGNN_STAGE_SCORE) CUDA_OK(h, sgnn_hist_tc_run(HG, w, node_off_db, h->sm_count, st));
  return 0;
}

#define LAUNCH_OK(h, what, expr)                                                              \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ == cudaErrorInvalidValue) return fail(h, std::string(what) + ": row too long for the shared-memory sorter"); \
    if (e_ != cudaSuccess) return fail(h, std::string(what) + ": " + cudaGetErrorString(e_)); \
  } while (0)

int sgnn_topk(SgnnHandle h, const float* sim_dev, int32_t m, int32_t n, int64_t ld, int32_t k, int32_t* ids_dev,
              float* vals_dev, void* stream) {
  if (!h || !sim_dev || !ids_dev) return fail(h, "sgnn_topk: null argument");
  if (k < 0 || k > n) return fail(h, "Invalid k " + std::to_string(k));  // results.py:68-69
  if (m == 0 || k == 0) return 0;
  LAUNCH_OK(h, "sgnn_topk", sgnn_launch_topk(sim_dev, nullptr, m, n, ld, k, ids_dev, vals_dev, (cudaStream_t)stream));
  return 0;
}

int sgnn_topk_merge(SgnnHandle h, const float* cand_vals_dev, const int32_t* cand_ids_dev, int32_t m, int32_t c,
                    int32_t k, int32_t* ids_dev, float* vals_dev, void* stream) {
  if (!h || !cand_vals_dev || !cand_ids_dev || !ids_dev) return fail(h, "sgnn_topk_merge: null argument");
  if (k < 0 || k > c) return fail(h, "Invalid k " + std::to_string(k));
  if (m == 0 || k == 0) return 0;
  LAUNCH_OK(h, "sgnn_topk_merge",
            sgnn_launch_topk(cand_vals_dev, cand_ids_dev, m, c, c, k, ids_dev, vals_dev, (cudaStream_t)stream));
  return 0;
}

int sgnn_sort_rows(SgnnHandle h, const float* sim_dev, int32_t m, int32_t n, int64_t ld, int32_t* ids_dev,
                   void* stream) {
  if (!h || !sim_dev || !ids_dev) return fail(h, "sgnn_sort_rows: null argument");
  if (m == 0 || n == 0) return 0;
  LAUNCH_OK(h, "sgnn_sort_rows", sgnn_launch_sort_rows(sim_dev, m, n, ld, ids_dev, h->smem_optin, (cudaStream_t)stream));
  return 0;
}

static int launch_pairs(SgnnHandle h, bool bwd, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                        const float* target_sim, int32_t n_pairs, int32_t global_batch, double target_sum,
                        double target_sqsum, float* loss, float* raw, float* sim, void* stream, bool no_dropout) {
  if (n_pairs == 0) return 0;
  if (global_batch < n_pairs) return fail(h, "global_batch < n_pairs");
  if (padding_guard(h, b)) return 1;
  bool garms = false;   // the arms' buffers in a global workspace: graphs beyond what shared memory holds (Web-shaped)
  if (b->max_nodes > 32) {   // beyond the warp-per-pair kernel; the CTA-per-pair kernel takes any size
    const size_t need = h->use_team ? team_smem_floats(h->M, (int)b->max_nodes) * sizeof(float) : 0;
    if (!h->use_team)
      return fail(h, "graphs of up to " + std::to_string(b->max_nodes) + " nodes need the CTA-per-pair training kernel "
                     "(SGNN_TRAIN_KERNEL=warp serves graphs of <= 32 nodes)");
    if (need > (size_t)h->smem_optin) {
      garms = true;
      const size_t need_g = team_smem_floats_garms(h->M, (int)b->max_nodes) * sizeof(float);
      if (need_g > (size_t)h->smem_optin)
        return fail(h, "the model's weights do not fit the per-pair kernel's shared memory (" + std::to_string(need_g / 1024) +
                           " KB needed)");
      const int64_t ws = (int64_t)team_arm_ws_floats(h->M, (int)b->max_nodes) * std::min(n_pairs, h->sm_count);
      if (ws > h->cap_arm_ws) {   // grown by the call (the first call of a size allocates; steady-state calls do not)
        if (dev_alloc(h, &h->d_arm_ws, ws)) return 1;
        h->cap_arm_ws = ws;
      }
    }
  }
  const int v = b->max_nodes <= 16 ? 0 : 1;
  const int wpb = h->train_wpb[v];
  const int blocks = std::min((n_pairs + wpb - 1) / wpb, h->sm_count * 2);
  cudaStream_t st = (cudaStream_t)stream;
  const float invB = 1.0f / (float)global_batch;
  const float s1 = (float)target_sum, s2 = (float)target_sqsum;
  // dropout is on for train, val and test alike (layers capture FLAGS.dropout as a constant)
  const int dropping = h->cfg.dropout > 0.f && !no_dropout;
  uint32_t dbase = 0;
  if (dropping) {
    const uint32_t s32 = (uint32_t)(h->drop_seed ^ (h->drop_seed >> 32));
    dbase = sgnn_fmix32(s32 ^ sgnn_fmix32((uint32_t)h->drop_counter + 0x9E3779B9u));
    h->drop_counter += 1;
  }
  if (h->use_team) {
    const int nmax = std::max(4, (int)b->max_nodes);
    if (garms) {
      const size_t gsmem = team_smem_floats_garms(h->M, nmax) * sizeof(float);
      const int gblocks = std::min(n_pairs, h->sm_count);
      if (bwd)
        train_team_kernel<true, 1, true><<<gblocks, TEAM_CTA, gsmem, st>>>(
            h->M, to_dev(b), nmax, h->d_params, g1, g2, target_sim, n_pairs, invB, s1, s2, h->d_grads, loss, raw, sim,
            dropping, dbase, h->pair_offset, h->pair_stride, h->d_arm_ws);
      else
        train_team_kernel<false, 1, true><<<gblocks, TEAM_CTA, gsmem, st>>>(
            h->M, to_dev(b), nmax, h->d_params, g1, g2, target_sim, n_pairs, invB, s1, s2, h->d_grads, loss, raw, sim,
            dropping, dbase, h->pair_offset, h->pair_stride, h->d_arm_ws);
      CUDA_OK(h, cudaGetLastError());
      return 0;
    }
    const size_t smem = team_smem_floats(h->M, nmax) * sizeof(float);
    // more pairs than SMs: two CTAs per SM (the 64-register variant) when shared memory holds them
    const bool two = n_pairs > h->sm_count && 2 * (smem + 1024) <= (size_t)h->smem_optin + 1024;   // (228 KB per SM, 1 KB reserved per CTA)
    const int tblocks = std::min(n_pairs, h->sm_count * (two ? 2 : 1));
#define SGNN_LAUNCH_TEAM(BW, MB)                                                                                     \
  train_team_kernel<BW, MB><<<tblocks, TEAM_CTA, smem, st>>>(h->M, to_dev(b), nmax, h->d_params, g1, g2, target_sim, \
                                                             n_pairs, invB, s1, s2, h->d_grads, loss, raw, sim,      \
                                                             dropping, dbase, h->pair_offset, h->pair_stride)
    if (bwd) { if (two) SGNN_LAUNCH_TEAM(true, 2); else SGNN_LAUNCH_TEAM(true, 1); }
    else { if (two) SGNN_LAUNCH_TEAM(false, 2); else SGNN_LAUNCH_TEAM(false, 1); }
#undef SGNN_LAUNCH_TEAM
    CUDA_OK(h, cudaGetLastError());
    return 0;
  }
#define SGNN_LAUNCH_PAIRS(NM, BW)                                                                        \
  train_pairs_kernel<NM, BW><<<blocks, wpb * 32, h->train_smem[v], st>>>(                                \
      h->M, to_dev(b), h->d_params, g1, g2, target_sim, n_pairs, invB, s1, s2, h->d_grads, loss, raw, sim,      \
      dropping, dbase, h->pair_offset, h->pair_stride)
  if (v == 0) { if (bwd) SGNN_LAUNCH_PAIRS(16, true); else SGNN_LAUNCH_PAIRS(16, false); }
  else { if (bwd) SGNN_LAUNCH_PAIRS(32, true); else SGNN_LAUNCH_PAIRS(32, false); }
#undef SGNN_LAUNCH_PAIRS
  CUDA_OK(h, cudaGetLastError());
  return 0;
}

int sgnn_train_step(SgnnHandle h, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                    const float* target_sim, int32_t n_pairs, int32_t global_batch, double target_sum,
                    double target_sqsum, float* loss_dev, void* stream) {
  if (!h || !b || !g1 || !g2 || !target_sim) return fail(h, "sgnn_train_step: null argument");
  return launch_pairs(h, true, b, g1, g2, target_sim, n_pairs, global_batch, target_sum, target_sqsum,
                      loss_dev ? loss_dev : h->d_loss, nullptr, nullptr, stream);
}

int sgnn_eval_pairs(SgnnHandle h, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                    const float* target_sim, int32_t n_pairs, int32_t global_batch, double target_sum,
                    double target_sqsum, float* loss_dev, float* raw_dev, float* sim_dev, void* stream) {
  if (!h || !b || !g1 || !g2) return fail(h, "sgnn_eval_pairs: null argument");
  if (!target_sim && loss_dev) return fail(h, "sgnn_eval_pairs: a loss needs targets");
  return launch_pairs(h, false, b, g1, g2, target_sim, n_pairs, global_batch, target_sum, target_sqsum,
                      target_sim ? loss_dev : nullptr, raw_dev, sim_dev, stream);
}

int sgnn_adam_device_step(SgnnHandle h, int32_t enable) {
  if (!h) return 1;
  if (enable && !h->d_adam_t) {
    CUDA_OK(h, cudaMalloc(reinterpret_cast<void**>(&h->d_adam_t), 16));   // [0] step count, [1] finished CTAs of the running update
    CUDA_OK(h, cudaMemset(h->d_adam_t, 0, 16));
  }
  if (enable && !h->adam_on_device) {
    const long long t = h->adam_t;
    CUDA_OK(h, cudaMemcpy(h->d_adam_t, &t, sizeof(t), cudaMemcpyHostToDevice));
  } else if (!enable && h->adam_on_device) {
    long long t = 0;
    CUDA_OK(h, cudaMemcpy(&t, h->d_adam_t, sizeof(t), cudaMemcpyDeviceToHost));   // (synchronises)
    h->adam_t = t;
  }
  h->adam_on_device = enable != 0;
  return 0;
}

int sgnn_apply_adam(SgnnHandle h, float grad_scale, float* loss_dev, void* stream) {
  if (!h) return 1;
  if (h->adam_on_device) {   // step count and step size on the device: nothing in the launch changes from step to step
    const SgnnConfig& c = h->cfg;
    const int n = h->M.n_params;
    const int blocks = std::min((n + 255) / 256, h->sm_count);
    adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(n, h->d_params, h->d_grads, h->d_m, h->d_v, grad_scale,
                                                         c.weight_decay, 0.f, c.beta1, c.beta2, c.epsilon, loss_dev,
                                                         h->d_adam_t, (double)c.learning_rate);
    CUDA_OK(h, cudaGetLastError());
    return 0;
  }
  h->adam_t += 1;
  const SgnnConfig& c = h->cfg;
  const double t = (double)h->adam_t;
  const double lr_t = (double)c.learning_rate * std::sqrt(1.0 - std::pow((double)c.beta2, t)) /
                      (1.0 - std::pow((double)c.beta1, t));
  const int n = h->M.n_params;
  const int blocks = std::min((n + 255) / 256, h->sm_count);
  adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(n, h->d_params, h->d_grads, h->d_m, h->d_v, grad_scale,
                                                       c.weight_decay, (float)lr_t, c.beta1, c.beta2, c.epsilon,
                                                       loss_dev);
  CUDA_OK(h, cudaGetLastError());
  return 0;
}

int sgnn_peer_barrier(SgnnHandle h, void* stream) {
  if (!h) return 1;
  if (h->peer_world == 0) return fail(h, "sgnn_peer_barrier: no peer exchange set up (sgnn_peer_create/connect)");
  for (int r = 0; r < h->peer_world; ++r)
    if (!h->peer_bufs[r]) return fail(h, "sgnn_peer_barrier: peers not connected");
  const cudaError_t e = sgnn_launch_peer_barrier(h->peer_rank, h->peer_world, h->peer_bufs, h->d_peer_epoch,
                                                 (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(h, std::string("sgnn_peer_barrier: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_apply_adam_allreduce(SgnnHandle h, float grad_scale, float* loss_dev, void* stream) {
  if (!h) return 1;
  if (h->peer_world == 0) return fail(h, "sgnn_apply_adam_allreduce: no peer exchange set up (sgnn_peer_create/connect)");
  for (int r = 0; r < h->peer_world; ++r)
    if (!h->peer_bufs[r]) return fail(h, "sgnn_apply_adam_allreduce: peers not connected");
  if (h->adam_on_device)
    return fail(h, "sgnn_apply_adam_allreduce: the Adam step count lives on the device (sgnn_adam_device_step); the peer "
                   "all-reduce form keeps it on the host -- disable the device step count first");
  h->adam_t += 1;
  const SgnnConfig& c = h->cfg;
  const double t = (double)h->adam_t;
  const double lr_t = (double)c.learning_rate * std::sqrt(1.0 - std::pow((double)c.beta2, t)) /
                      (1.0 - std::pow((double)c.beta1, t));
  const cudaError_t e = sgnn_launch_peer_adam(h->M.n_params, h->peer_rank, h->peer_world, h->peer_slot_elems,
                                              h->peer_bufs, h->d_peer_epoch, h->peer_done, h->d_params, h->d_grads,
                                              h->d_m, h->d_v, grad_scale, c.weight_decay, (float)lr_t, c.beta1,
                                              c.beta2, c.epsilon, loss_dev, h->sm_count, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(h, std::string("sgnn_apply_adam_allreduce: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_set_dropout_stream(SgnnHandle h, uint64_t seed, uint64_t counter) {
  if (!h) return 1;
  h->drop_seed = seed; h->drop_counter = counter;
  return 0;
}

int sgnn_set_embed_tc(SgnnHandle h, int32_t enable) {
  if (!h) return 1;
  if (enable && h->cfg.precision == SGNN_PREC_FP32) return fail(h, "sgnn_set_embed_tc: the engine was created with the fp32 (CUDA-core) precision");
  h->embed_tc = enable != 0;
  return 0;
}

int sgnn_set_pair_offset(SgnnHandle h, int32_t offset, int32_t stride) {
  if (!h || offset < 0 || stride < 1) return 1;
  h->pair_offset = offset; h->pair_stride = stride;
  return 0;
}

int sgnn_zero_grads(SgnnHandle h, void* stream) {
  if (!h) return 1;
  CUDA_OK(h, cudaMemsetAsync(h->d_grads, 0, (size_t)h->M.n_params * sizeof(float), (cudaStream_t)stream));
  return 0;
}

int sgnn_precision_at_k(SgnnHandle h, const float* true_sim_dev, const float* pred_sim_dev, int32_t m, int32_t n,
                        const int32_t* ks_dev, int32_t n_ks, double* prec_dev, void* stream) {
  if (!h || !true_sim_dev || !pred_sim_dev || !ks_dev || !prec_dev) return fail(h, "sgnn_precision_at_k: null argument");
  if (m <= 0 || n <= 1 || n_ks <= 0) return fail(h, "sgnn_precision_at_k: empty problem");
  LAUNCH_OK(h, "sgnn_precision_at_k", sgnn_launch_prec_at_k(true_sim_dev, pred_sim_dev, m, n, ks_dev, n_ks, prec_dev,
                                                            h->smem_optin, (cudaStream_t)stream));
  return 0;
}

int sgnn_precision_from_topk(SgnnHandle h, const float* true_sim_dev, int64_t ld, const float* true_topk_vals_dev,
                             const int32_t* pred_topk_ids_dev, int32_t m, int32_t kmax, const int32_t* ks_dev,
                             int32_t n_ks, double* prec_dev, void* stream) {
  if (!h || !true_sim_dev || !true_topk_vals_dev || !pred_topk_ids_dev || !ks_dev || !prec_dev)
    return fail(h, "sgnn_precision_from_topk: null argument");
  if (m <= 0 || kmax <= 0 || n_ks <= 0) return fail(h, "sgnn_precision_from_topk: empty problem");
  LAUNCH_OK(h, "sgnn_precision_from_topk", sgnn_launch_prec_from_topk(true_sim_dev, ld, true_topk_vals_dev, pred_topk_ids_dev,
                                                                      m, kmax, ks_dev, n_ks, prec_dev, (cudaStream_t)stream));
  return 0;
}

int sgnn_kendall_tau(SgnnHandle h, const float* true_sim_dev, const float* pred_sim_dev, int32_t m, int32_t n,
                     int64_t* counts_dev, double* tau_dev, void* stream) {
  if (!h || !true_sim_dev || !pred_sim_dev || !counts_dev) return fail(h, "sgnn_kendall_tau: null argument");
  if (m <= 0 || n <= 1) return fail(h, "sgnn_kendall_tau: empty problem");
  LAUNCH_OK(h, "sgnn_kendall_tau", sgnn_launch_kendall(true_sim_dev, pred_sim_dev, m, n,
                                                       reinterpret_cast<long long*>(counts_dev), tau_dev,
                                                       h->smem_optin, (cudaStream_t)stream));
  return 0;
}

// ---- host-buffer entry points ---------------------------------------------------------
// Host batch -> device.  Every cudaMemcpyAsync costs the calling thread several microseconds, which at
// this problem size is comparable to the kernels, so when the caller's five arrays lie back to back in
// ONE host allocation (PackedBatch.pin() lays them out that way, 16-byte aligned) the whole span goes
// up as a single copy into an arena with the same relative layout; otherwise five copies.
static int upload_batch_on(SgnnHandle h, const SgnnGraphBatch* b, int slot, SgnnGraphBatch* dev, cudaStream_t st) {
  HostWs& w = h->ws;
  if (b->n_graphs > w.cap_graphs || b->total_nodes > w.cap_nodes || b->nnz > w.cap_nnz)
    return fail(h, "host workspace too small: call sgnn_reserve first");
  *dev = *b;
  const uint8_t* p[5] = {(const uint8_t*)b->node_off, (const uint8_t*)b->labels, (const uint8_t*)b->rowptr,
                         (const uint8_t*)b->col, (const uint8_t*)b->val};
  const size_t sz[5] = {(size_t)(b->n_graphs + 1) * 4, (size_t)b->total_nodes * 4, (size_t)(b->total_nodes + 1) * 4,
                        (size_t)b->nnz * 4, (size_t)b->nnz * 4};
  const uint8_t* lo = p[0];
  const uint8_t* hi = p[0] + sz[0];
  size_t total = 0;
  for (int i = 0; i < 5; ++i) {
    lo = p[i] < lo ? p[i] : lo;
    hi = p[i] + sz[i] > hi ? p[i] + sz[i] : hi;
    total += sz[i];
  }
  const size_t span = (size_t)(hi - lo);
  if (span <= total + 5 * 64 && (int64_t)span <= w.arena_cap && ((uintptr_t)lo & 15) == 0) {
    CUDA_OK(h, cudaMemcpyAsync(w.arena[slot], lo, span, cudaMemcpyHostToDevice, st));
    dev->node_off = (const int32_t*)(w.arena[slot] + (p[0] - lo));
    dev->labels = (const int32_t*)(w.arena[slot] + (p[1] - lo));
    dev->rowptr = (const int32_t*)(w.arena[slot] + (p[2] - lo));
    dev->col = (const int32_t*)(w.arena[slot] + (p[3] - lo));
    dev->val = (const float*)(w.arena[slot] + (p[4] - lo));
    return 0;
  }
  CUDA_OK(h, cudaMemcpyAsync(w.node_off[slot], b->node_off, sz[0], cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.labels[slot], b->labels, sz[1], cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.rowptr[slot], b->rowptr, sz[2], cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.col[slot], b->col, sz[3], cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.val[slot], b->val, sz[4], cudaMemcpyHostToDevice, st));
  dev->node_off = w.node_off[slot]; dev->labels = w.labels[slot]; dev->rowptr = w.rowptr[slot];
  dev->col = w.col[slot]; dev->val = w.val[slot];
  return 0;
}

// Expansion of a compact blob (sgnn_compact.h) into the int32 / float CSR arrays the kernels read: one warp
// per graph; row offsets by a warp prefix sum over the row lengths (32 nodes at a time with a carry), then the
// graph's entries are written by all lanes (coalesced), values from the dictionary.
__global__ void __launch_bounds__(256)
expand_compact_kernel(const uint8_t* __restrict__ blob, int32_t* __restrict__ node_off, int32_t* __restrict__ labels,
                      int32_t* __restrict__ rowptr, int32_t* __restrict__ col, float* __restrict__ val) {
  const SgnnCompactHeader* hd = reinterpret_cast<const SgnnCompactHeader*>(blob);
  const int G = hd->n_graphs, N = hd->total_nodes;
  const SgnnCompactLayout L = sgnn_compact_layout(G, N, hd->nnz, hd->n_codes);
  const int32_t* no = reinterpret_cast<const int32_t*>(blob + L.node_off);
  const int32_t* eo = reinterpret_cast<const int32_t*>(blob + L.e_off);
  const float* dict = reinterpret_cast<const float*>(blob + L.dict);
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int g = warp; g < G; g += nwarps) {
    const int n0 = no[g], n1 = no[g + 1], e0 = eo[g], e1 = eo[g + 1];
    if (lane == 0) {
      node_off[g] = n0;
      if (g == G - 1) { node_off[G] = n1; rowptr[N] = e1; }
    }
    int carry = e0;
    for (int i0 = n0; i0 < n1; i0 += 32) {
      const int i = i0 + lane;
      const int len = i < n1 ? blob[L.row_nnz + i] : 0;
      int incl = len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      if (i < n1) { rowptr[i] = carry + incl - len; labels[i] = blob[L.labels + i]; }
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    for (int e = e0 + lane; e < e1; e += 32) { col[e] = blob[L.col + e]; val[e] = dict[blob[L.vcode + e]]; }
  }
}

// compact host blob -> arena (one copy) -> expanded arrays of `slot`
static int upload_compact_on(SgnnHandle h, const uint8_t* blob, int slot, SgnnGraphBatch* dev, cudaStream_t st) {
  HostWs& w = h->ws;
  SgnnCompactHeader hd;
  memcpy(&hd, blob, sizeof(hd));
  if (hd.magic != SGNN_COMPACT_MAGIC) return fail(h, "not a compact batch (sgnn_compact_batch)");
  if (hd.n_graphs > w.cap_graphs || hd.total_nodes > w.cap_nodes || hd.nnz > w.cap_nnz || hd.total_bytes > w.arena_cap)
    return fail(h, "host workspace too small: call sgnn_reserve first");
  CUDA_OK(h, cudaMemcpyAsync(w.arena[slot], blob, (size_t)hd.total_bytes, cudaMemcpyHostToDevice, st));
  const int blocks = std::max(1, std::min((hd.n_graphs + 7) / 8, h->sm_count * 8));
  expand_compact_kernel<<<blocks, 256, 0, st>>>(w.arena[slot], w.node_off[slot], w.labels[slot], w.rowptr[slot],
                                                w.col[slot], w.val[slot]);
  CUDA_OK(h, cudaGetLastError());
  dev->n_graphs = hd.n_graphs; dev->total_nodes = hd.total_nodes; dev->nnz = hd.nnz; dev->max_nodes = hd.max_nodes;
  dev->node_off = w.node_off[slot]; dev->labels = w.labels[slot]; dev->rowptr = w.rowptr[slot];
  dev->col = w.col[slot]; dev->val = w.val[slot];
  return 0;
}

static int upload_batch(SgnnHandle h, const SgnnGraphBatch* b, int slot, SgnnGraphBatch* dev) {
  return upload_batch_on(h, b, slot, dev, h->stream);
}

int sgnn_forward_pairs_host(SgnnHandle h, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                            int32_t n_pairs, float* raw_host, float* sim_host) {
  if (!h || !b) return fail(h, "sgnn_forward_pairs_host: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  HostWs& w = h->ws;
  if (n_pairs > w.cap_pairs) return fail(h, "host workspace too small: call sgnn_reserve first");
  SgnnGraphBatch d;
  if (upload_batch(h, b, 0, &d)) return 1;
  cudaStream_t st = h->stream;
  CUDA_OK(h, cudaMemcpyAsync(w.g1, g1, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.g2, g2, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  if (sgnn_forward_pairs(h, &d, w.g1, w.g2, n_pairs, w.emb, w.raw, w.sim_pairs, st)) return 1;
  if (raw_host) CUDA_OK(h, cudaMemcpyAsync(raw_host, w.raw, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, st));
  if (sim_host) CUDA_OK(h, cudaMemcpyAsync(sim_host, w.sim_pairs, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, st));
  CUDA_OK(h, cudaStreamSynchronize(st));
  return 0;
}

int sgnn_rank_device(SgnnHandle h, const SgnnGraphBatch* q_dev, const SgnnGraphBatch* db_dev, int32_t k,
                     float* emb_q_dev, float* emb_db_dev, float* sim_dev, int64_t ld, int32_t* topk_ids_dev,
                     float* topk_vals_dev, void* stream) {
  if (!h || !q_dev || !db_dev || !emb_q_dev || !emb_db_dev || !sim_dev)
    return fail(h, "sgnn_rank_device: null argument");
  const int m = q_dev->n_graphs, n = db_dev->n_graphs;
  if (m == 0 || n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream, side = h->copy_stream;
  auto body = [&]() -> int {
    // fork: the query branch (embed + query-side scorer tables) runs beside the database branch.  It is enqueued
    // FIRST and on a high-priority stream: the database embedding alone fills every SM, and queued behind it the
    // query branch (22 us of small kernels) would extend the critical path instead of hiding under it.
    CUDA_OK(h, cudaEventRecord(h->chunk_ev[1], st));
    CUDA_OK(h, cudaStreamWaitEvent(side, h->chunk_ev[1], 0));
    // (the query embedding must be the FIRST kernel of the step: its CTAs do not fit beside the database embedding's,
    // the two small operand kernels after it do)
    bool fused_b = false;
    if (embed_side(h, q_dev, emb_q_dev, 0, side, &fused_b, db_dev)) return 1;
    if (!fused_b) {
      // the coefficient matrix needs the parameters only: a third stream takes it off the query chain
      cudaStream_t aux = h->aux_stream;
      CUDA_OK(h, cudaStreamWaitEvent(aux, h->chunk_ev[1], 0));
      if (score_all_pairs_stages(h, emb_q_dev, m, emb_db_dev, n, sim_dev, ld, aux, SCORE_COEF_ONLY)) return 1;
      CUDA_OK(h, cudaEventRecord(h->chunk_ev[5], aux));
      CUDA_OK(h, cudaStreamWaitEvent(side, h->chunk_ev[5], 0));
      if (score_all_pairs_stages(h, emb_q_dev, m, emb_db_dev, n, sim_dev, ld, side, SCORE_PREP_Q | SCORE_COEF_READY)) return 1;
    }
    CUDA_OK(h, cudaEventRecord(h->chunk_ev[2], side));
    bool fused_a = false;
    if (embed_side(h, db_dev, emb_db_dev, 1, st, &fused_a)) return 1;
    CUDA_OK(h, cudaStreamWaitEvent(st, h->chunk_ev[2], 0));   // join (the query branch ended long before the database embedding)
    // database-side operands and the scorer in ONE call: the stage function then takes the NTN operand off this chain
    return score_and_topk(h, emb_q_dev, m, emb_db_dev, n, sim_dev, ld, (fused_a ? 0 : SCORE_PREP_DB) | SCORE_MAIN, k,
                          topk_ids_dev, topk_vals_dev, st);
  };
  uint64_t kb[22] = {};
  key_batch(kb, q_dev);
  key_batch(kb + 7, db_dev);
  kb[14] = (uint64_t)k; kb[15] = (uint64_t)(uintptr_t)emb_q_dev; kb[16] = (uint64_t)(uintptr_t)emb_db_dev;
  kb[17] = (uint64_t)(uintptr_t)sim_dev; kb[18] = (uint64_t)ld; kb[19] = (uint64_t)(uintptr_t)topk_ids_dev;
  kb[20] = (uint64_t)(uintptr_t)topk_vals_dev; kb[21] = (uint64_t)(uintptr_t)st;
  return run_cached(h, h->g_rank_device, fnv1a(kb, sizeof(kb)), st, body);
}

int sgnn_peer_destroy(SgnnHandle h) {
  if (!h) return 1;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < h->peer_world; ++r) {
    if (!h->peer_bufs[r]) continue;
    if (r == h->peer_rank) cudaFree(h->peer_bufs[r]);
    else cudaIpcCloseMemHandle(h->peer_bufs[r]);
    h->peer_bufs[r] = nullptr;
  }
  if (h->peer_done) cudaFree(h->peer_done);
  if (h->d_peer_epoch) cudaFree(h->d_peer_epoch);
  h->peer_done = nullptr; h->d_peer_epoch = nullptr;
  h->g_rank_sharded.drop();
  h->peer_world = 0;
  return 0;
}

int sgnn_peer_create(SgnnHandle h, int32_t world, int32_t rank, int32_t m_cap, int32_t k_cap, uint8_t* ipc_handle_out) {
  if (!h || !ipc_handle_out) return fail(h, "sgnn_peer_create: null argument");
  if (world < 1 || world > 16 || rank < 0 || rank >= world) return fail(h, "sgnn_peer_create: bad world / rank");
  static_assert(sizeof(cudaIpcMemHandle_t) == SGNN_IPC_HANDLE_BYTES, "IPC handle size");
  CUDA_OK(h, cudaSetDevice(h->device));
  sgnn_peer_destroy(h);
  h->peer_world = world; h->peer_rank = rank;
  h->peer_slot_elems = (int64_t)m_cap * k_cap;
  const int64_t bytes = sgnn_peer_buffer_bytes(world, h->peer_slot_elems, h->M.n_params);
  CUDA_OK(h, cudaMalloc(&h->peer_bufs[rank], (size_t)bytes));
  CUDA_OK(h, cudaMemset(h->peer_bufs[rank], 0, (size_t)bytes));
  CUDA_OK(h, cudaMalloc(reinterpret_cast<void**>(&h->peer_done), sizeof(unsigned int)));
  CUDA_OK(h, cudaMemset(h->peer_done, 0, sizeof(unsigned int)));
  CUDA_OK(h, cudaMalloc(reinterpret_cast<void**>(&h->d_peer_epoch), 4 * sizeof(unsigned long long)));
  CUDA_OK(h, cudaMemset(h->d_peer_epoch, 0, 4 * sizeof(unsigned long long)));
  CUDA_OK(h, cudaDeviceSynchronize());
  cudaIpcMemHandle_t hd;
  CUDA_OK(h, cudaIpcGetMemHandle(&hd, h->peer_bufs[rank]));
  memcpy(ipc_handle_out, &hd, sizeof(hd));
  return 0;
}

int sgnn_peer_connect(SgnnHandle h, const uint8_t* all_handles) {
  if (!h || !all_handles) return fail(h, "sgnn_peer_connect: null argument");
  if (h->peer_world == 0) return fail(h, "sgnn_peer_connect: call sgnn_peer_create first");
  CUDA_OK(h, cudaSetDevice(h->device));
  for (int r = 0; r < h->peer_world; ++r) {
    if (r == h->peer_rank) continue;
    cudaIpcMemHandle_t hd;
    memcpy(&hd, all_handles + (size_t)r * SGNN_IPC_HANDLE_BYTES, sizeof(hd));
    CUDA_OK(h, cudaIpcOpenMemHandle(&h->peer_bufs[r], hd, cudaIpcMemLazyEnablePeerAccess));
  }
  return 0;
}

int sgnn_topk_exchange(SgnnHandle h, const float* vals_dev, const int32_t* ids_dev, int32_t m, int32_t k,
                       int32_t id_offset, int32_t* out_ids_dev, float* out_vals_dev, void* stream) {
  if (!h || !vals_dev || !ids_dev || !out_ids_dev) return fail(h, "sgnn_topk_exchange: null argument");
  if (h->peer_world == 0) return fail(h, "sgnn_topk_exchange: no peer exchange set up (sgnn_peer_create/connect)");
  for (int r = 0; r < h->peer_world; ++r)
    if (!h->peer_bufs[r]) return fail(h, "sgnn_topk_exchange: peers not connected");
  if ((int64_t)m * k > h->peer_slot_elems) return fail(h, "sgnn_topk_exchange: m * k exceeds the exchange slot");
  const cudaError_t e = sgnn_launch_peer_exchange(vals_dev, ids_dev, m, k, id_offset, h->peer_rank, h->peer_world,
                                                  h->peer_slot_elems, h->peer_bufs, h->d_peer_epoch, h->peer_done,
                                                  out_ids_dev, out_vals_dev, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(h, std::string("sgnn_topk_exchange: ") + cudaGetErrorString(e));
  return 0;
}

int sgnn_rank_sharded(SgnnHandle h, const SgnnGraphBatch* q_dev, const SgnnGraphBatch* db_dev, int32_t k,
                      int32_t id_offset, float* emb_q_dev, float* emb_db_dev, float* sim_dev, int64_t ld,
                      int32_t* local_ids_dev, float* local_vals_dev, int32_t* out_ids_dev, float* out_vals_dev,
                      void* stream) {
  if (!h || !q_dev || !db_dev || !emb_q_dev || !emb_db_dev || !sim_dev || !local_ids_dev || !local_vals_dev ||
      !out_ids_dev)
    return fail(h, "sgnn_rank_sharded: null argument");
  if (k <= 0) return fail(h, "sgnn_rank_sharded: k must be positive");
  if (h->peer_world == 0) return fail(h, "sgnn_rank_sharded: no peer exchange set up (sgnn_peer_create/connect)");
  const int m = q_dev->n_graphs;
  cudaStream_t st = (cudaStream_t)stream;
  auto body = [&]() -> int {
    const bool saved = h->use_graphs;
    h->use_graphs = false;   // the inner entry points run inline: this call is the unit that is captured
    int rc = sgnn_rank_device(h, q_dev, db_dev, k, emb_q_dev, emb_db_dev, sim_dev, ld, local_ids_dev, local_vals_dev,
                              stream);
    if (!rc) rc = sgnn_topk_exchange(h, local_vals_dev, local_ids_dev, m, k, id_offset, out_ids_dev, out_vals_dev, stream);
    h->use_graphs = saved;
    return rc;
  };
  uint64_t kb[26] = {};
  key_batch(kb, q_dev);
  key_batch(kb + 7, db_dev);
  kb[14] = (uint64_t)k; kb[15] = (uint64_t)(uintptr_t)emb_q_dev; kb[16] = (uint64_t)(uintptr_t)emb_db_dev;
  kb[17] = (uint64_t)(uintptr_t)sim_dev; kb[18] = (uint64_t)ld; kb[19] = (uint64_t)(uintptr_t)local_ids_dev;
  kb[20] = (uint64_t)(uintptr_t)local_vals_dev; kb[21] = (uint64_t)(uintptr_t)st;
  kb[22] = (uint64_t)(uint32_t)id_offset; kb[23] = (uint64_t)(uintptr_t)out_ids_dev;
  kb[24] = (uint64_t)(uintptr_t)out_vals_dev; kb[25] = (uint64_t)(uintptr_t)h->peer_bufs[h->peer_rank];
  return run_cached(h, h->g_rank_sharded, fnv1a(kb, sizeof(kb)), st, body);
}

int sgnn_debug_slice_hints(SgnnHandle h, unsigned long long** hints_dev, int64_t* capacity) {
  if (!h || !hints_dev) return fail(h, "sgnn_debug_slice_hints: null argument");
  *hints_dev = h->d_smax;
  if (capacity) *capacity = h->cap_smax;
  return 0;
}

}  // extern "C"

// shared body of sgnn_rank_host / sgnn_rank_host_compact: `up(slot, dev, stream)` uploads one side
template <typename UpQ, typename UpDb>
static int rank_host_impl(SgnnHandle h, int64_t m, int64_t n, const uint64_t* key_words, int n_key, UpQ up_q, UpDb up_db,
                          int32_t k, int32_t* topk_ids_host, float* topk_vals_host, float* sim_host) {
  HostWs& w = h->ws;
  if (m * n > w.cap_q * w.cap_db || m > w.cap_q || (int64_t)m * k > w.cap_topk)
    return fail(h, "host workspace too small: call sgnn_reserve first");
  cudaStream_t st = h->stream, cs = h->copy_stream;
  const bool want_topk = k > 0 && topk_ids_host;
  auto body = [&]() -> int {
    // the database upload (PCIe) runs on the copy stream while the compute stream uploads and embeds the queries
    SgnnGraphBatch dq, ddb;
    CUDA_OK(h, cudaEventRecord(h->chunk_ev[3], st));
    CUDA_OK(h, cudaStreamWaitEvent(cs, h->chunk_ev[3], 0));
    if (up_db(1, &ddb, cs)) return 1;
    CUDA_OK(h, cudaEventRecord(h->chunk_ev[0], cs));
    if (up_q(0, &dq, st)) return 1;
    bool fused_b = false;
    if (embed_side(h, &dq, w.emb_q, 0, st, &fused_b)) return 1;
    // query-side tables of the scorer are ready before the database lands
    if (!fused_b && score_all_pairs_stages(h, w.emb_q, (int)m, w.emb_db, (int)n, w.sim, n, st, SCORE_PREP_Q)) return 1;
    CUDA_OK(h, cudaStreamWaitEvent(st, h->chunk_ev[0], 0));
    bool fused_a = false;
    if (embed_side(h, &ddb, w.emb_db, 1, st, &fused_a)) return 1;
    if (score_and_topk(h, w.emb_q, (int)m, w.emb_db, (int)n, w.sim, n, (fused_a ? 0 : SCORE_PREP_DB) | SCORE_MAIN,
                       want_topk ? k : 0, w.topk_ids, reinterpret_cast<float*>(w.topk_ids + m * k), st))
      return 1;
    if (want_topk) {
      const size_t bytes = (size_t)m * k * 4;
      if (topk_vals_host == reinterpret_cast<float*>(topk_ids_host + m * k)) {   // contiguous on both sides: one copy
        CUDA_OK(h, cudaMemcpyAsync(topk_ids_host, w.topk_ids, 2 * bytes, cudaMemcpyDeviceToHost, st));
      } else {
        CUDA_OK(h, cudaMemcpyAsync(topk_ids_host, w.topk_ids, bytes, cudaMemcpyDeviceToHost, st));
        if (topk_vals_host)
          CUDA_OK(h, cudaMemcpyAsync(topk_vals_host, w.topk_ids + m * k, bytes, cudaMemcpyDeviceToHost, st));
      }
    }
    if (sim_host) CUDA_OK(h, cudaMemcpyAsync(sim_host, w.sim, (size_t)m * n * 4, cudaMemcpyDeviceToHost, st));
    return 0;
  };
  if (run_cached(h, h->g_rank_host, fnv1a(key_words, sizeof(uint64_t) * (size_t)n_key), st, body)) return 1;
  CUDA_OK(h, cudaStreamSynchronize(st));
  return 0;
}

extern "C" {

int sgnn_rank_host(SgnnHandle h, const SgnnGraphBatch* q, const SgnnGraphBatch* db, int32_t k,
                   int32_t* topk_ids_host, float* topk_vals_host, float* sim_host) {
  if (!h || !q || !db) return fail(h, "sgnn_rank_host: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  uint64_t kb[18] = {};
  key_batch(kb, q);
  key_batch(kb + 7, db);
  kb[14] = (uint64_t)k; kb[15] = (uint64_t)(uintptr_t)topk_ids_host; kb[16] = (uint64_t)(uintptr_t)topk_vals_host;
  kb[17] = (uint64_t)(uintptr_t)sim_host;
  return rank_host_impl(
      h, q->n_graphs, db->n_graphs, kb, 18,
      [&](int slot, SgnnGraphBatch* d, cudaStream_t s) { return upload_batch_on(h, q, slot, d, s); },
      [&](int slot, SgnnGraphBatch* d, cudaStream_t s) { return upload_batch_on(h, db, slot, d, s); }, k, topk_ids_host,
      topk_vals_host, sim_host);
}

int sgnn_rank_host_compact(SgnnHandle h, const uint8_t* q_blob, const uint8_t* db_blob, int32_t k,
                           int32_t* topk_ids_host, float* topk_vals_host, float* sim_host) {
  if (!h || !q_blob || !db_blob) return fail(h, "sgnn_rank_host_compact: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  SgnnCompactHeader hq, hdb;
  memcpy(&hq, q_blob, sizeof(hq));
  memcpy(&hdb, db_blob, sizeof(hdb));
  if (hq.magic != SGNN_COMPACT_MAGIC || hdb.magic != SGNN_COMPACT_MAGIC)
    return fail(h, "sgnn_rank_host_compact: not a compact batch (sgnn_compact_batch)");
  uint64_t kb[12] = {};
  kb[0] = (uint64_t)(uintptr_t)q_blob; kb[1] = (uint64_t)(uintptr_t)db_blob;
  kb[2] = ((uint64_t)(uint32_t)hq.n_graphs << 32) | (uint32_t)hq.max_nodes; kb[3] = (uint64_t)hq.total_bytes;
  kb[4] = ((uint64_t)(uint32_t)hdb.n_graphs << 32) | (uint32_t)hdb.max_nodes; kb[5] = (uint64_t)hdb.total_bytes;
  kb[6] = (uint64_t)k; kb[7] = (uint64_t)(uintptr_t)topk_ids_host; kb[8] = (uint64_t)(uintptr_t)topk_vals_host;
  kb[9] = (uint64_t)(uintptr_t)sim_host; kb[10] = ((uint64_t)(uint32_t)hq.total_nodes << 32) | (uint32_t)hdb.total_nodes;
  kb[11] = 0x636f6d70616374ull;   // distinct from the plain entry point's keys
  return rank_host_impl(
      h, hq.n_graphs, hdb.n_graphs, kb, 12,
      [&](int slot, SgnnGraphBatch* d, cudaStream_t s) { return upload_compact_on(h, q_blob, slot, d, s); },
      [&](int slot, SgnnGraphBatch* d, cudaStream_t s) { return upload_compact_on(h, db_blob, slot, d, s); }, k,
      topk_ids_host, topk_vals_host, sim_host);
}

int sgnn_train_step_host(SgnnHandle h, const SgnnGraphBatch* b, const int32_t* g1, const int32_t* g2,
                         const float* target_sim, int32_t n_pairs, float* loss_host) {
  if (!h || !b) return fail(h, "sgnn_train_step_host: null argument");
  CUDA_OK(h, cudaSetDevice(h->device));
  HostWs& w = h->ws;
  if (n_pairs > w.cap_pairs) return fail(h, "host workspace too small: call sgnn_reserve first");
  double s1 = 0.0, s2 = 0.0;
  for (int i = 0; i < n_pairs; ++i) { s1 += target_sim[i]; s2 += (double)target_sim[i] * target_sim[i]; }
  SgnnGraphBatch d;
  if (upload_batch(h, b, 0, &d)) return 1;
  cudaStream_t st = h->stream;
  CUDA_OK(h, cudaMemcpyAsync(w.g1, g1, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.g2, g2, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemcpyAsync(w.target, target_sim, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemsetAsync(h->d_loss, 0, 4 * sizeof(float), st));
  if (sgnn_train_step(h, &d, w.g1, w.g2, w.target, n_pairs, n_pairs, s1, s2, h->d_loss, st)) return 1;
  if (sgnn_apply_adam(h, 1.0f, h->d_loss, st)) return 1;
  if (loss_host) CUDA_OK(h, cudaMemcpyAsync(loss_host, h->d_loss, 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_OK(h, cudaStreamSynchronize(st));
  return 0;
}

}  // extern "C"
