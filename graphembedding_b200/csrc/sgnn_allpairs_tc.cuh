// All-pairs scorer on the 5th-gen tensor cores (tcgen05 + TMEM), persistent and
// warp-specialised.
//
//   D[g, (q,k)] = sum_b A[g,b] * B[(q,k),b]      A = [x_g, 1, 0..]   B = [P'[q][k][:], C[q][k], 0..]
//
// (see sgnn_allpairs.cuh for P', C: reference layers.py:282-310 factorised per query).  Rows of D
// (TMEM lanes) are database graphs, columns are (query, feature-map) pairs, so the NTN
// "activation + reduce over k" epilogue runs over CONSECUTIVE COLUMNS inside one thread.
//
// Precision: operands are split x = hi + lo (two bf16), and D = Ahi.Bhi + Ahi.Blo + Alo.Bhi is
// accumulated in fp32 in TMEM ("bf16x3", ~2^-17 relative: inside the 1e-4 fp32 parity gate);
// SGNN_PREC_BF16 issues only the hi.hi product.
//
// Shared-memory operand layout (UMMA K-major, no swizzle): [part hi|lo][K chunk of 8][row][8 bf16]
// i.e. 16-byte rows, 8-row core matrices contiguous (SBO = 128 B), K chunks R*16 B apart (LBO).
// The prep kernels write global memory in exactly this order, so a tile is ONE bulk async copy.
//
// Roles (416 threads): warps 0-7 epilogue in two groups of four (group = w/4 takes every other tile,
// warp w reads TMEM lanes 32(w%4)..), warp 8 bulk-copy producer, warp 9 MMA issuer.  TMEM: 4
// accumulator buffers x 128 columns; the two groups are half a tile out of phase, so one group's
// TMEM loads overlap the other's arithmetic on every scheduler.
#pragma once
#include "sgnn_common.cuh"
#include "sgnn_tc.cuh"
#include "sgnn_tc_geom.h"

// Operand preparation, ONE launch: blocks [0, q_blocks) build the B tiles (queries), the rest the A
// tiles (database graphs).
//
// B is a small GEMM: out[q][c] = const[c] + sum_a x_q[a] * Wm[a][c] over the columns c = (kk, k) of
// P' / C (kk < D: P'[q][k][kk] = V[k][D+kk] + sum_a x[a] W[a][kk][k];  kk == D: C[q][k] = b[k] +
// V[k][0:D] . x;  above: 0).  A block stages the coefficient columns of PREP_COLS columns once and
// applies them to PREP_Q queries held in shared memory (x broadcast as float4).  With KC == K + 1
// (relu inner activation) column (kk, K) holds sum_k w_k * column (kk, k) (w = 1 in reference mode,
// U in paper mode): the tensor core then also delivers sum_k w_k m_k, and
//   sum_k w_k relu(m_k) = (sum_k w_k m_k + sum_k w_k |m_k|) / 2
// costs ONE add-with-|x|-modifier per accumulator element in the epilogue instead of max + add.
// A: one thread per (database graph, K chunk).
constexpr int PREP_COLS = 256;   // = block size
constexpr int PREP_Q = 8;

__global__ void __launch_bounds__(PREP_COLS)
tc_prep_kernel(const __grid_constant__ ModelDev M, const TcGeom G, int q_blocks, int col_blocks,
               const float* __restrict__ params, const float* __restrict__ emb_q,
               const float* __restrict__ emb_db, __nv_bfloat16* __restrict__ A, __nv_bfloat16* __restrict__ B) {
  extern __shared__ __align__(16) float prep_smem[];
  if ((int)blockIdx.x >= q_blocks) {
    const int idx = (blockIdx.x - q_blocks) * blockDim.x + threadIdx.x;
    const int total = G.n_gtiles * 128 * G.KCH;
    if (idx >= total) return;
    const int kc = idx % G.KCH, g = idx / G.KCH;
    const int T = g >> 7, r = g & 127;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int b = kc * 8 + j;
      float v = 0.f;
      if (g < G.n) v = b < G.D ? emb_db[(size_t)g * G.D + b] : (b == G.D ? 1.f : 0.f);
      split_bf16(v, hi[j], lo[j]);
    }
    const size_t base = ((size_t)(T * G.parts) * G.KCH + kc) * 1024 + r * 8;
    *reinterpret_cast<uint4*>(A + base) = *reinterpret_cast<const uint4*>(hi);
    if (G.parts == 2)
      *reinterpret_cast<uint4*>(A + base + (size_t)G.KCH * 1024) = *reinterpret_cast<const uint4*>(lo);
    return;
  }
  const int D = G.D, K = G.K, KC = G.KC, DK16 = G.KCH * 8;
  const int C = DK16 * KC;
  const int cpb = (PREP_COLS / KC) * KC;          // columns per block: whole (kk, 0..KC) groups
  const int cb = blockIdx.x % col_blocks, qblk = blockIdx.x / col_blocks;
  const int c = cb * cpb + threadIdx.x;
  const int q0 = qblk * PREP_Q;
  float* Wm = prep_smem;                          // [D + 1][PREP_COLS]
  float* xs = prep_smem + (D + 1) * PREP_COLS;    // [D][PREP_Q] (transposed: queries contiguous)
  const bool on = (int)threadIdx.x < cpb && c < C;
  const int kk = on ? c / KC : 0, k = on ? c - kk * KC : 0;
  const bool ref = M.pair_kind == SGNN_LAYER_DOT || M.ntn_mode == SGNN_NTN_REFERENCE;
  if (on && k < K)
    for (int a = 0; a <= D; ++a) Wm[a * PREP_COLS + threadIdx.x] = tc_prep_coef(M, params, D, K, a, kk, k);
  for (int i = threadIdx.x; i < D * PREP_Q; i += blockDim.x) {
    const int a = i / PREP_Q, qq = i - a * PREP_Q;
    xs[i] = q0 + qq < G.m ? emb_q[(size_t)(q0 + qq) * D + a] : 0.f;
  }
  __syncthreads();
  if (KC > K) {   // the sum columns: one (coefficient row a, column group) per thread, K-term sums in parallel
    const int groups = cpb / KC;
    for (int t = threadIdx.x; t < (D + 1) * groups; t += blockDim.x) {
      const int a = t / groups, col = (t - a * groups) * KC;
      if (cb * cpb + col + K < C) {
        float v = 0.f;
        for (int kq = 0; kq < K; ++kq) v = fmaf(ref ? 1.f : params[M.u_off + kq], Wm[a * PREP_COLS + col + kq], v);
        Wm[a * PREP_COLS + col + K] = v;
      }
    }
  }
  __syncthreads();
  if (!on) return;
  float acc[PREP_Q];
  const float cst = Wm[D * PREP_COLS + threadIdx.x];
#pragma unroll
  for (int qq = 0; qq < PREP_Q; ++qq) acc[qq] = cst;
  for (int a = 0; a < D; ++a) {
    const float w = Wm[a * PREP_COLS + threadIdx.x];
    const float4* x4 = reinterpret_cast<const float4*>(xs + a * PREP_Q);
#pragma unroll
    for (int v4 = 0; v4 < PREP_Q / 4; ++v4) {
      const float4 x = x4[v4];
      acc[4 * v4 + 0] = fmaf(x.x, w, acc[4 * v4 + 0]);
      acc[4 * v4 + 1] = fmaf(x.y, w, acc[4 * v4 + 1]);
      acc[4 * v4 + 2] = fmaf(x.z, w, acc[4 * v4 + 2]);
      acc[4 * v4 + 3] = fmaf(x.w, w, acc[4 * v4 + 3]);
    }
  }
  const size_t lo_off = (size_t)G.KCH * TC_NPAD * 8;
#pragma unroll
  for (int qq = 0; qq < PREP_Q; ++qq) {
    const int q = q0 + qq;
    if (q < G.m) {
      const int qb = q / G.QH, row = (q - qb * G.QH) * KC + k;
      __nv_bfloat16* dst = B + ((size_t)(qb * G.parts) * G.KCH + (kk >> 3)) * (TC_NPAD * 8) + (size_t)row * 8 + (kk & 7);
      __nv_bfloat16 hi, lo;
      split_bf16(acc[qq], hi, lo);
      *dst = hi;
      if (G.parts == 2) dst[lo_off] = lo;
    }
  }
}

// Query-side operand in two small kernels without shared memory, sized to run BESIDE the database embedding
// (whose two resident CTAs leave 4 K registers and ~3 KB of shared memory per SM): on the high-priority side
// stream of the ranking step they start at once instead of waiting for that kernel to drain.  (The staged
// tc_prep_kernel above needs 8 K registers + 11.6 KB per CTA and was serialised behind it: 10 us on the critical
// path of a 70 us step.)
//   tc_coef_kernel:   Wm[a][c], the (D + 1) x C coefficient matrix of the columns c = (kk, k) of P' / C --
//                     depends on the parameters only, so it does not wait for the query embeddings;
//   tc_prep_q_kernel: out[q][c] = Wm[D][c] + sum_a x_q[a] Wm[a][c], thread = one column for PREPQ_Q queries,
//                     coalesced coefficient loads, warp-uniform embedding loads.
// Same arithmetic and summation order as tc_prep_kernel's query branch => identical operand bits.
constexpr int PREPQ_THREADS = 64;
constexpr int PREPQ_Q = 8;

__global__ void __launch_bounds__(PREPQ_THREADS)
tc_coef_kernel(const __grid_constant__ ModelDev M, const TcGeom G, const float* __restrict__ params,
               float* __restrict__ Wm) {
  const int D = G.D, K = G.K, KC = G.KC;
  const int C = G.KCH * 8 * KC;
  const int idx = blockIdx.x * PREPQ_THREADS + threadIdx.x;
  if (idx >= (D + 1) * C) return;
  const int a = idx / C, c = idx - a * C;
  const int kk = c / KC, k = c - kk * KC;
  float v;
  if (k < K) {
    v = tc_prep_coef(M, params, D, K, a, kk, k);
  } else {   // KC == K + 1: the column sum_k w_k (.) of the |x| form of relu
    const bool ref = M.pair_kind == SGNN_LAYER_DOT || M.ntn_mode == SGNN_NTN_REFERENCE;
    v = 0.f;
#pragma unroll 4
    for (int kq = 0; kq < K; ++kq)
      v = fmaf(ref ? 1.f : params[M.u_off + kq], tc_prep_coef(M, params, D, K, a, kk, kq), v);
  }
  Wm[idx] = v;
}

__global__ void __launch_bounds__(PREPQ_THREADS, 16)
tc_prep_q_kernel(const TcGeom G, const float* __restrict__ Wm, const float* __restrict__ emb_q,
                 __nv_bfloat16* __restrict__ B) {
  __shared__ float xs[PREPQ_Q * 64];   // the block's query embeddings (D <= 64): 2 KB, fits the leftover too
  const int D = G.D, KC = G.KC;
  const int C = G.KCH * 8 * KC;
  const int c = blockIdx.x * PREPQ_THREADS + threadIdx.x;
  const int q0 = blockIdx.y * PREPQ_Q;
  const int nq = min(PREPQ_Q, G.m - q0);
  for (int i = threadIdx.x; i < PREPQ_Q * D; i += PREPQ_THREADS)   // PREPQ_Q consecutive rows = one contiguous span
    xs[i] = i < nq * D ? emb_q[(size_t)q0 * D + i] : 0.f;
  __syncthreads();
  if (c >= C) return;
  const int kk = c / KC, k = c - kk * KC;
  float acc[PREPQ_Q];
  const float cst = Wm[(size_t)D * C + c];
#pragma unroll
  for (int qq = 0; qq < PREPQ_Q; ++qq) acc[qq] = cst;
#pragma unroll 5
  for (int a = 0; a < D; ++a) {
    const float w = Wm[(size_t)a * C + c];
#pragma unroll
    for (int qq = 0; qq < PREPQ_Q; ++qq) acc[qq] = fmaf(xs[qq * D + a], w, acc[qq]);
  }
  const size_t lo_off = (size_t)G.KCH * TC_NPAD * 8;
  int qb = q0 / G.QH, r = q0 - qb * G.QH;   // query q = q-block qb, slot r; advanced without further divisions
  __nv_bfloat16* col = B + (size_t)(kk >> 3) * (TC_NPAD * 8) + (kk & 7);
  const size_t qb_stride = (size_t)G.parts * G.KCH * (TC_NPAD * 8);
#pragma unroll
  for (int qq = 0; qq < PREPQ_Q; ++qq) {
    if (qq < nq) {
      __nv_bfloat16* dst = col + (size_t)qb * qb_stride + (size_t)(r * KC + k) * 8;
      __nv_bfloat16 hi, lo;
      split_bf16(acc[qq], hi, lo);
      *dst = hi;
      if (G.parts == 2) dst[lo_off] = lo;
    }
    if (++r == G.QH) { r = 0; ++qb; }
  }
}

struct TcEpi {
  const float* sU;   // [K] in shared memory
  float su, yeta;
  float c2;          // relu + sim_kernel fast path: sim = 2^(c2 * (sum + abs_sum)^2)
  int final_act, pair_act;
  float* sim;
  int64_t ld;
  int m, n;
  unsigned long long* smax;   // optional [m][n_slices]: largest rank key (orderable score << 32 | graph id,
                              // sgnn_rank.cuh) of every 32-graph slice -- hints for topk_hint_kernel
  int n_slices;
#ifdef SGNN_TC_TRACE
  long long* trace; int tile;
#endif
};
#ifdef SGNN_TC_TRACE
#define TC_ESTAMP(slot) do { if (blockIdx.x == 0 && (threadIdx.x & 127) == 0 && (threadIdx.x >> 7) % 2 == 0) E.trace[E.tile * 8 + (slot)] = clock64(); } while (0)
#else
#define TC_ESTAMP(slot) do { } while (0)
#endif

// CREDUX.MAX.F32: one instruction, uniform result.  .NaN: a NaN score (a diverged model) makes the slice maximum NaN, whose
// rank key sorts above every number -- as it does in rank_key() / numpy's order -- so the hinted top-k never skips the
// slice; the callers mark the holders with !(mine < w), which is mine == w for numbers and every lane for a NaN maximum
__device__ __forceinline__ float tc_warp_max(float x) {
  float y;
  asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(y) : "f"(x));
  return y;
}

// same key as rank_key() in sgnn_rank.cuh: descending key order = descending score, ties -> higher id first
__device__ __forceinline__ unsigned long long tc_rank_key(float f, int idx) {
  f += 0.f;
  unsigned int u = __float_as_uint(f);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  if (f != f) u = 0xFFFFFFFFu;
  return ((unsigned long long)u << 32) | (unsigned int)idx;
}

__device__ __forceinline__ float tc_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// score of one (query, database graph) from the reduced accumulator value
template <bool RELU, bool REF, bool SIMK>
__device__ __forceinline__ float tc_finish(const TcEpi& E, float st) {
  if constexpr (RELU && SIMK) {
    return tc_ex2(st * st * E.c2);
  } else {
    float score = RELU ? 0.5f * st : st;
    if (REF) score *= E.su;
    if constexpr (SIMK) return __expf(-E.yeta * score * score);
    return sgnn_act(score, E.final_act, E.yeta);
  }
}

// Epilogue of one tile part: this thread owns TMEM lane (database graph) g and the accumulator
// columns of queries [Q0, Q0 + NQ) of the tile (PART 0: the first ceil(QH/2) queries, PART 1: the
// rest).  KT > 0: K known at compile time; the part is pulled into registers with one wait, the
// warp signals the buffer's MMA issuer at once (8 warps share a buffer), and the arithmetic + stores
// run from registers while the other warps of the scheduler cover the TMEM load latency.
template <int KT, bool RELU, bool REF, bool SIMK, int PART, bool HINT>
__device__ __forceinline__ void tc_epilogue_part(const TcEpi& E, uint32_t taddr, int K, int KCr, int QH, int qrow0,
                                                 int g, uint64_t* t_empty, int lane) {
  const int64_t ld = E.ld;
  if constexpr (KT > 0) {
    constexpr int KC = RELU ? KT + 1 : KT;
    constexpr int QHC = 128 / KC, QA = (QHC + 1) / 2;
    constexpr int Q0 = PART ? QA : 0, NQ = PART ? QHC - QA : QA, NC = NQ * KC;
    const int qmax = g < E.n ? min(NQ, E.m - (qrow0 + Q0)) : 0;   // queries this thread may store
    float* p = E.sim + (size_t)(qrow0 + Q0) * ld + g;
    float v[NC > 0 ? NC : 1];
    tc::tmem_ld_n<NC>(taddr + Q0 * KC, v);
    TC_ESTAMP(5);
    tc::tmem_ld_wait();
    TC_ESTAMP(7);
    tc::tc_fence_before();
    __syncwarp();
    if (lane == 0) tc::mbar_arrive(t_empty);
    float u[KT];
    if constexpr (!REF) {
#pragma unroll
      for (int k = 0; k < KT; ++k) u[k] = E.sU[k];
    }
    static_assert(NQ <= 32, "one lane per query keeps the slice maximum");
    constexpr bool hint = HINT;
    const float off_row = g < E.n ? 0.f : -INFINITY;   // rows past the database never win a slice
    float keep = 0.f;
    unsigned keep_mask = 1u;
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      float st;
      if constexpr (RELU) {
        st = v[q * KC + KT];   // sum_k w_k m_k from the tensor core
#pragma unroll
        for (int k = 0; k < KT; ++k) st = REF ? st + fabsf(v[q * KC + k]) : fmaf(fabsf(v[q * KC + k]), u[k], st);
      } else {
        st = 0.f;
#pragma unroll
        for (int k = 0; k < KT; ++k) {
          const float t = sgnn_act(v[q * KC + k], E.pair_act, E.yeta);
          st = REF ? st + t : fmaf(u[k], t, st);
        }
      }
      const float sv = tc_finish<RELU, REF, SIMK>(E, st);
      if (q < qmax) *p = sv;
      p += ld;
      if (hint) {   // slice maximum and which lanes hold it (ties -> highest graph id, the reference's order)
        const float mine = sv + off_row;
        const float w = tc_warp_max(mine);
        const unsigned who = __ballot_sync(0xffffffffu, !(mine < w));
        if (lane == q) { keep = w; keep_mask = who; }
      }
    }
    if (hint && lane < min(NQ, E.m - (qrow0 + Q0)))
      E.smax[(size_t)(qrow0 + Q0 + lane) * E.n_slices + (g >> 5)] =
          tc_rank_key(keep, (g & ~31) + 31 - __clz(keep_mask | 1u));
  } else {
    const int QA = (QH + 1) / 2;
    const int q0 = PART ? QA : 0, nq = PART ? QH - QA : QA;
    const int qmax = g < E.n ? min(nq, E.m - (qrow0 + q0)) : 0;
    float* p = E.sim + (size_t)(qrow0 + q0) * ld + g;
    const int cols = nq * KCr;
    const uint32_t t0 = taddr + q0 * KCr;
    float st = 0.f;
    int kk = 0, qq = 0;
    constexpr bool hint = HINT;
    const int qrows = min(nq, E.m - (qrow0 + q0));
    float keep = 0.f;
    int keep_lane = 0;
    for (int c0 = 0; c0 < cols; c0 += 8) {
      float v[8];
      tc::tmem_ld8(t0 + c0, v);   // may read past this part's columns, never past the 128-column tile + 7
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (c0 + i < cols) {
          if (kk < K) {
            const float w = REF ? 1.f : E.sU[kk];
            const float t = RELU ? fabsf(v[i]) : sgnn_act(v[i], E.pair_act, E.yeta);
            st = fmaf(w, t, st);
          } else {
            st += v[i];  // the sum column (RELU only: KCr == K + 1)
          }
          if (++kk == KCr) {
            const float sv = tc_finish<RELU, REF, SIMK>(E, st);
            if (qq < qmax) p[(size_t)qq * ld] = sv;
            if (hint) {   // lane (qq mod 32) keeps query qq's slice maximum; flushed every 32 queries
              const float mine = g < E.n ? sv : -INFINITY;
              const float w = tc_warp_max(mine);
              const int top = 31 - __clz(__ballot_sync(0xffffffffu, !(mine < w)));
              if (lane == (qq & 31)) { keep = w; keep_lane = top; }
              if ((qq & 31) == 31 || qq == nq - 1) {
                const int qrow = (qq & ~31) + lane;
                if (qrow <= qq && qrow < qrows)
                  E.smax[(size_t)(qrow0 + q0 + qrow) * E.n_slices + (g >> 5)] = tc_rank_key(keep, (g & ~31) + keep_lane);
              }
            }
            st = 0.f; kk = 0; ++qq;
          }
        }
      }
    }
    tc::tc_fence_before();
    __syncwarp();
    if (lane == 0) tc::mbar_arrive(t_empty);
  }
}

// Roles shared by the scorer kernels: the bulk-copy producer warp and the MMA issuer warps (everything but the epilogue).
struct TcCtaWork { int gt0, ga_count, qb0, nq; };
struct TcSmem {
  uint8_t *sA, *sB;
  uint64_t *a_full, *b_full, *b_empty, *t_full, *t_empty;
  int a_bytes, b_bytes;
};
#ifdef SGNN_TC_TRACE
#define TC_STAMP(tile, slot) do { if (blockIdx.x == 0) trace[(tile) * 8 + (slot)] = clock64(); } while (0)
#else
#define TC_STAMP(tile, slot) do { } while (0)
#endif
__device__ __forceinline__ void tc_producer_mma_roles(const TcGeom& G, const TcCtaWork& Wk, const TcSmem& S,
                                                      const __nv_bfloat16* __restrict__ Ag,
                                                      const __nv_bfloat16* __restrict__ Bg, uint32_t tmem_base, int warp,
                                                      int lane
#ifdef SGNN_TC_TRACE
                                                      , long long* __restrict__ trace
#endif
                                                      ) {
  uint8_t* const sA = S.sA; uint8_t* const sB = S.sB;
  uint64_t* const a_full = S.a_full; uint64_t* const b_full = S.b_full; uint64_t* const b_empty = S.b_empty;
  uint64_t* const t_full = S.t_full; uint64_t* const t_empty = S.t_empty;
  const int a_bytes = S.a_bytes, b_bytes = S.b_bytes;
  const int gt0 = Wk.gt0, ga_count = Wk.ga_count, qb0 = Wk.qb0, nq = Wk.nq;
  if (warp == TC_PROD_WARP) {
    // ===== producer: A tiles once, then the stream of query-block B tiles =====
    if (lane == 0 && nq > 0) {
      tc::mbar_arrive_expect_tx(a_full, ga_count * a_bytes);
      for (int ga = 0; ga < ga_count; ++ga)
        tc::bulk_g2s(sA + ga * a_bytes, reinterpret_cast<const uint8_t*>(Ag) + (size_t)(gt0 + ga) * a_bytes,
                     a_bytes, a_full);
      for (int i = 0; i < nq; ++i) {
        const int s = i % G.stages;
        if (i >= G.stages) tc::mbar_wait(b_empty + s, ((i / G.stages) - 1) & 1);
        tc::mbar_arrive_expect_tx(b_full + s, b_bytes);
        tc::bulk_g2s(sB + s * b_bytes, reinterpret_cast<const uint8_t*>(Bg) + (size_t)(qb0 + i) * b_bytes, b_bytes,
                     b_full + s);
      }
    }
  } else if (warp > TC_PROD_WARP) {
    // ===== MMA issuers: warp 9 + w owns TMEM buffer w and the tiles it = w, w + 4, ... (one thread each).
    // Measured with tools/tc_trace.cu: back-to-back tcgen05.mma into the SAME accumulator serialise at
    // ~250 cycles each (the three bf16x3 products of a tile are such a chain) while chains on different
    // accumulators overlap, and every scalar instruction of a lone issuing thread costs ~8 cycles.  So
    // there is one short chain per buffer in flight, and the loop carries running descriptors instead of
    // recomputing tile coordinates. =====
    if (lane == 0 && nq > 0) {
      const int w = warp - TC_MMA_WARP0;
      const uint32_t idesc = tc::umma_idesc_bf16(128, TC_NPAD);
      const uint32_t a_part = (G.KCH * 128 * 16) >> 4, b_part = (G.KCH * TC_NPAD * 16) >> 4;   // in 16-byte units
      const uint32_t a_ks = (2 * 128 * 16) >> 4, b_ks = (2 * TC_NPAD * 16) >> 4;
      const uint64_t a_desc0 = tc::umma_desc(tc::smem_u32(sA), 128 * 16, 128);
      const uint64_t b_desc0 = tc::umma_desc(tc::smem_u32(sB), TC_NPAD * 16, 128);
      const uint32_t a_tile = (uint32_t)a_bytes >> 4, b_tile = (uint32_t)b_bytes >> 4;
      const int combos = G.parts == 2 ? 3 : 1;
      const int ksteps = G.KCH / 2;
      const bool recycle = nq > G.stages;
      const uint32_t d_addr = tmem_base + w * TC_NPAD;
      const int total = nq * ga_count;
      tc::mbar_wait(a_full, 0);
      // running tile coordinates of it: (i, ga), B stage slot of i, parities
      int i = w / ga_count, ga = w - i * ga_count;
      int waited = 0, released = 0, slot_w = 0, par_w = 0, slot_r = 0;
      int slot_i = i % G.stages;
      uint32_t par_e = 1;   // parity of the previous phase of t_empty[w] (first wait is skipped)
      for (int it = w; it < total; it += TC_BUFS) {
        while (waited <= i) {   // B tiles arrive in order
          tc::mbar_wait(b_full + slot_w, par_w);
          ++waited;
          if (++slot_w == G.stages) { slot_w = 0; par_w ^= 1; }
        }
        TC_STAMP(it, 0);
        if (it >= TC_BUFS) tc::mbar_wait(t_empty + w, par_e);
        par_e ^= 1;
        tc::tc_fence_after();
        TC_STAMP(it, 1);
        const uint64_t ad0 = a_desc0 + (uint64_t)(ga * a_tile);
        const uint64_t bd0 = b_desc0 + (uint64_t)(slot_i * b_tile);
        uint32_t acc = 0;
        for (int c = 0; c < combos; ++c) {   // (hi,hi) (hi,lo) (lo,hi)
          const uint32_t ao = c == 2 ? a_part : 0u, bo = c == 1 ? b_part : 0u;
          for (int ks = 0; ks < ksteps; ++ks) {
            tc::umma_bf16(d_addr, ad0 + ao + (uint64_t)(ks * a_ks), bd0 + bo + (uint64_t)(ks * b_ks), idesc, acc);
            acc = 1;
          }
        }
        tc::umma_commit(t_full + w);
        TC_STAMP(it, 2);
        // advance (i, ga) by TC_BUFS tiles
        ga += TC_BUFS;
        while (ga >= ga_count) {
          ga -= ga_count; ++i;
          if (++slot_i == G.stages) slot_i = 0;
        }
        if (recycle) {   // hand back the B stages all of this warp's MMAs are done with (4 issuers: count 4)
          const int upto = it + TC_BUFS < total ? i : nq;
          for (; released < upto; ++released) {
            tc::umma_commit(b_empty + slot_r);
            if (++slot_r == G.stages) slot_r = 0;
          }
        }
      }
    }
  }
}

template <int KT, bool RELU, bool REF, bool SIMK>
__global__ void __launch_bounds__(TC_THREADS, 1)
allpairs_tc_kernel(const __grid_constant__ ModelDev M, const TcGeom G, const float* __restrict__ params,
                   const __nv_bfloat16* __restrict__ Ag, const __nv_bfloat16* __restrict__ Bg,
                   float* __restrict__ sim, int64_t ld, unsigned long long* __restrict__ smax
#ifdef SGNN_TC_TRACE
                   , long long* __restrict__ trace   // [tile][8] clock64 stamps of CTA 0 (tools/tc_trace.cu)
#endif
                   ) {
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint8_t* smem = tc_smem;
  const int a_bytes = tc_a_bytes(G), b_bytes = tc_b_bytes(G);
  uint8_t* sA = smem;
  uint8_t* sB = sA + TC_GA_MAX * a_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + G.stages * b_bytes);
  uint64_t* a_full = bars;                 // [1]
  uint64_t* b_full = bars + 1;             // [G.stages]
  uint64_t* b_empty = b_full + G.stages;  // [G.stages]
  uint64_t* t_full = b_empty + G.stages;  // [TC_BUFS]
  uint64_t* t_empty = t_full + TC_BUFS;    // [TC_BUFS]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + TC_BUFS);
  float* sU = reinterpret_cast<float*>(tmem_slot + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // this CTA's rectangle of work: GA database tiles x a range of query blocks
  const int gg = blockIdx.x / G.q_groups, qg = blockIdx.x - gg * G.q_groups;
  const int gt0 = gg * G.GA;
  const int ga_count = min(G.GA, G.n_gtiles - gt0);
  const int qb0 = qg * G.qb_per_group;
  const int nq = max(0, min(G.qb_per_group, G.n_qblocks - qb0));

  if (threadIdx.x == 0) {
    tc::mbar_init(a_full, 1);
    for (int s = 0; s < G.stages; ++s) { tc::mbar_init(b_full + s, 1); tc::mbar_init(b_empty + s, TC_BUFS); }
    for (int b = 0; b < TC_BUFS; ++b) { tc::mbar_init(t_full + b, 1); tc::mbar_init(t_empty + b, 8); }
    tc::fence_barrier_init();
  }
  const bool dot = M.pair_kind == SGNN_LAYER_DOT;
  for (int k = threadIdx.x; k < G.K; k += blockDim.x) sU[k] = dot ? 1.f : params[M.u_off + k];
  if (warp == TC_PROD_WARP) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= TC_PROD_WARP) {
    TcCtaWork Wk{gt0, ga_count, qb0, nq};
    TcSmem S{sA, sB, a_full, b_full, b_empty, t_full, t_empty, a_bytes, b_bytes};
    tc_producer_mma_roles(G, Wk, S, Ag, Bg, tmem_base, warp, lane
#ifdef SGNN_TC_TRACE
                          , trace
#endif
                          );
  } else {
    // ===== epilogue warps =====
    TcEpi E;
    E.sU = sU; E.yeta = M.yeta; E.final_act = M.final_act; E.pair_act = dot ? SGNN_ACT_IDENTITY : M.pair_act;
    E.sim = sim; E.ld = ld; E.m = G.m; E.n = G.n;
    E.smax = smax; E.n_slices = G.n_gtiles * 4;
    float su = 0.f;
    for (int k = 0; k < G.K; ++k) su += sU[k];
    E.su = su;
    {
      const float hs = REF ? 0.5f * su : 0.5f;
      E.c2 = -M.yeta * hs * hs * 1.4426950408889634f;
    }
    // group `grp` takes tiles it = grp, grp + 2, ...; tile it lives in TMEM buffer it % 4, so a buffer
    // is only ever read by ONE group (TC_BUFS is a multiple of TC_EPI_GROUPS).  That matters: mbarrier
    // waits are by phase parity, and a group that skipped a buffer's phase (because another group
    // consumed it) can be two phases behind, where the parity wait aliases and returns at once.
    const int lg = warp & 3, part = (warp >> 2) & 1, grp = warp >> 3;
    const int total = nq * ga_count;
    int i = grp / ga_count, ga = grp - i * ga_count;
    for (int it = grp; it < total; it += TC_EPI_GROUPS) {
      const int buf = it & (TC_BUFS - 1);
      if (lg == 0 && lane == 0) TC_STAMP(it, 3);
      tc::mbar_wait(t_full + buf, (it / TC_BUFS) & 1);
      tc::tc_fence_after();
      if (lg == 0 && lane == 0) TC_STAMP(it, 4);
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + buf * TC_NPAD;
      const int g = (gt0 + ga) * 128 + lg * 32 + lane;
#ifdef SGNN_TC_TRACE
      E.trace = trace; E.tile = it;
#endif
      if (smax == nullptr) {
        if (part == 0)
          tc_epilogue_part<KT, RELU, REF, SIMK, 0, false>(E, taddr, G.K, G.KC, G.QH, (qb0 + i) * G.QH, g, t_empty + buf, lane);
        else
          tc_epilogue_part<KT, RELU, REF, SIMK, 1, false>(E, taddr, G.K, G.KC, G.QH, (qb0 + i) * G.QH, g, t_empty + buf, lane);
      } else {
        if (part == 0)
          tc_epilogue_part<KT, RELU, REF, SIMK, 0, true>(E, taddr, G.K, G.KC, G.QH, (qb0 + i) * G.QH, g, t_empty + buf, lane);
        else
          tc_epilogue_part<KT, RELU, REF, SIMK, 1, true>(E, taddr, G.K, G.KC, G.QH, (qb0 + i) * G.QH, g, t_empty + buf, lane);
      }
      if (lg == 0 && lane == 0) TC_STAMP(it, 6);
      ga += TC_EPI_GROUPS;
      while (ga >= ga_count) { ga -= ga_count; ++i; }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == TC_PROD_WARP) tc::tmem_dealloc(tmem_base, 512);
}
