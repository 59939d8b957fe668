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
// Roles (320 threads): warps 0-7 epilogue in two groups of four (group = w/4 takes every other tile,
// warp w reads TMEM lanes 32(w%4)..), warp 8 bulk-copy producer, warp 9 MMA issuer.  TMEM: 4
// accumulator buffers x 128 columns; the two groups are half a tile out of phase, so one group's
// TMEM loads overlap the other's arithmetic on every scheduler.
#pragma once
#include "sgnn_common.cuh"
#include "sgnn_tc.cuh"
#include "sgnn_tc_geom.h"

__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// Operand preparation, ONE launch: blocks [0, q_blocks) build the B tiles (queries), the rest the A
// tiles (database graphs).
//
// B: thread = (query, kk); it walks the K feature maps of its column kk of P' / C.  With
// KC == K + 1 (relu inner activation) row K of the query holds sum_k w_k * row_k (w = 1 in reference
// mode, U in paper mode): the tensor core then also delivers sum_k w_k m_k, and
//   sum_k w_k relu(m_k) = (sum_k w_k m_k + sum_k w_k |m_k|) / 2
// costs ONE add-with-|x|-modifier per accumulator element in the epilogue instead of max + add.
// A: one thread per (database graph, K chunk).
__global__ void __launch_bounds__(128)
tc_prep_kernel(const __grid_constant__ ModelDev M, const TcGeom G, int q_blocks, const float* __restrict__ params,
               const float* __restrict__ emb_q, const float* __restrict__ emb_db, __nv_bfloat16* __restrict__ A,
               __nv_bfloat16* __restrict__ B) {
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
  const int D = G.D, K = G.K, DK16 = G.KCH * 8, DK = D * K;
  const int qpb = blockDim.x / DK16;
  const int ql = threadIdx.x / DK16, kk = threadIdx.x - ql * DK16;
  const int q = blockIdx.x * qpb + ql;
  if (ql >= qpb || q >= G.m) return;
  const int qb = q / G.QH, qq = q - qb * G.QH;
  const int row0 = qq * G.KC;
  const float* x = emb_q + (size_t)q * D;
  const bool dot = M.pair_kind == SGNN_LAYER_DOT;
  const bool ref = M.ntn_mode == SGNN_NTN_REFERENCE;
  const float* W = params + M.w_off;
  const float* V = params + M.v_off;
  __nv_bfloat16* Bq = B + ((size_t)(qb * G.parts) * G.KCH + (kk >> 3)) * (TC_NPAD * 8) + (kk & 7);
  const size_t lo_off = (size_t)G.KCH * TC_NPAD * 8;
  auto put = [&](int row, float v) {
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    Bq[(size_t)row * 8] = hi;
    if (G.parts == 2) Bq[(size_t)row * 8 + lo_off] = lo;
  };
  float sum = 0.f;
  for (int k = 0; k < K; ++k) {
    float v = 0.f;
    if (dot) {
      v = kk < D ? x[kk] : 0.f;
    } else if (kk < D) {          // P'[q][k][kk] = V[k][D + kk] + sum_a x[a] W[a][kk][k]
      v = V[k * 2 * D + D + kk];
      for (int a = 0; a < D; ++a) v = fmaf(x[a], W[a * DK + kk * K + k], v);
    } else if (kk == D) {         // C[q][k] = b[k] + V[k][0:D] . x
      v = M.pair_bias ? params[M.b_off + k] : 0.f;
      for (int a = 0; a < D; ++a) v = fmaf(V[k * 2 * D + a], x[a], v);
    }
    put(row0 + k, v);
    sum = fmaf(dot || ref ? 1.f : params[M.u_off + k], v, sum);
  }
  if (G.KC > K) put(row0 + K, sum);
}

struct TcEpi {
  const float* sU;   // [K] in shared memory
  float su, yeta;
  float c2;          // relu + sim_kernel fast path: sim = 2^(c2 * (sum + abs_sum)^2)
  int final_act, pair_act;
  float* sim;
  int64_t ld;
  int m, n;
};

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

// Epilogue of one accumulator half: this thread owns TMEM lane (database graph) g and walks the
// columns of QH queries x KC accumulator columns.  KT > 0: K known at compile time; the whole half
// is pulled into registers, the TMEM buffer is handed back to the MMA warp at once, and the
// arithmetic + stores run from registers while the next tile's MMAs execute.
template <int KT, bool RELU, bool REF, bool SIMK>
__device__ __forceinline__ void tc_epilogue_half(const TcEpi& E, uint32_t taddr, int K, int KCr, int QH, int qrow0,
                                                 int g, uint64_t* t_empty, int lane) {
  const int qmax = g < E.n ? min(QH, E.m - qrow0) : 0;   // queries this thread may store
  float* p = E.sim + (size_t)qrow0 * E.ld + g;
  const int64_t ld = E.ld;
  if constexpr (KT > 0) {
    constexpr int KC = RELU ? KT + 1 : KT;
    constexpr int QHC = 128 / KC, COLS = QHC * KC, NCH = (COLS + 31) / 32;
    float v[NCH * 32];
#pragma unroll
    for (int c = 0; c < NCH; ++c) tc::tmem_ld32(taddr + c * 32, v + c * 32);
    tc::tmem_ld_wait();
    tc::tc_fence_before();
    __syncwarp();
    if (lane == 0) tc::mbar_arrive(t_empty);
    float u[KT];
    if constexpr (!REF) {
#pragma unroll
      for (int k = 0; k < KT; ++k) u[k] = E.sU[k];
    }
#pragma unroll
    for (int q = 0; q < QHC; ++q) {
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
      const float s = tc_finish<RELU, REF, SIMK>(E, st);
      if (q < qmax) *p = s;
      p += ld;
    }
  } else {
    const int cols = QH * KCr;
    float st = 0.f;
    int kk = 0, qq = 0;
    for (int c0 = 0; c0 < cols; c0 += 8) {
      float v[8];
      tc::tmem_ld8(taddr + c0, v);
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
            const float s = tc_finish<RELU, REF, SIMK>(E, st);
            if (qq < qmax) p[(size_t)qq * ld] = s;
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

template <int KT, bool RELU, bool REF, bool SIMK>
__global__ void __launch_bounds__(TC_THREADS, 1)
allpairs_tc_kernel(const __grid_constant__ ModelDev M, const TcGeom G, const float* __restrict__ params,
                   const __nv_bfloat16* __restrict__ Ag, const __nv_bfloat16* __restrict__ Bg,
                   float* __restrict__ sim, int64_t ld) {
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint8_t* smem = tc_smem;
  const int a_bytes = tc_a_bytes(G), b_bytes = tc_b_bytes(G);
  uint8_t* sA = smem;
  uint8_t* sB = sA + TC_GA_MAX * a_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + TC_STAGES * b_bytes);
  uint64_t* a_full = bars;                 // [1]
  uint64_t* b_full = bars + 1;             // [TC_STAGES]
  uint64_t* b_empty = b_full + TC_STAGES;  // [TC_STAGES]
  uint64_t* t_full = b_empty + TC_STAGES;  // [TC_BUFS]
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
    for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(b_full + s, 1); tc::mbar_init(b_empty + s, 1); }
    for (int b = 0; b < TC_BUFS; ++b) { tc::mbar_init(t_full + b, 1); tc::mbar_init(t_empty + b, 4); }
    tc::fence_barrier_init();
  }
  const bool dot = M.pair_kind == SGNN_LAYER_DOT;
  for (int k = threadIdx.x; k < G.K; k += blockDim.x) sU[k] = dot ? 1.f : params[M.u_off + k];
  if (warp == 8) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ===== producer: A tiles once, then the stream of query-block B tiles =====
    if (lane == 0 && nq > 0) {
      tc::mbar_arrive_expect_tx(a_full, ga_count * a_bytes);
      for (int ga = 0; ga < ga_count; ++ga)
        tc::bulk_g2s(sA + ga * a_bytes, reinterpret_cast<const uint8_t*>(Ag) + (size_t)(gt0 + ga) * a_bytes,
                     a_bytes, a_full);
      for (int i = 0; i < nq; ++i) {
        const int s = i % TC_STAGES;
        if (i >= TC_STAGES) tc::mbar_wait(b_empty + s, ((i / TC_STAGES) - 1) & 1);
        tc::mbar_arrive_expect_tx(b_full + s, b_bytes);
        tc::bulk_g2s(sB + s * b_bytes, reinterpret_cast<const uint8_t*>(Bg) + (size_t)(qb0 + i) * b_bytes, b_bytes,
                     b_full + s);
      }
    }
  } else if (warp == 9) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0 && nq > 0) {
      const uint32_t idesc = tc::umma_idesc_bf16(128, TC_NPAD);
      const uint32_t a_part = G.KCH * 128 * 16, b_part = G.KCH * TC_NPAD * 16;
      tc::mbar_wait(a_full, 0);
      int it = 0;
      for (int i = 0; i < nq; ++i) {
        const int s = i % TC_STAGES;
        tc::mbar_wait(b_full + s, (i / TC_STAGES) & 1);
        tc::tc_fence_after();
        const uint32_t b_addr = tc::smem_u32(sB + s * b_bytes);
        for (int ga = 0; ga < ga_count; ++ga, ++it) {
          const int buf = it & (TC_BUFS - 1);
          if (it >= TC_BUFS) tc::mbar_wait(t_empty + buf, ((it / TC_BUFS) - 1) & 1);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(sA + ga * a_bytes);
          const uint32_t d_addr = tmem_base + buf * TC_NPAD;
          uint32_t acc = 0;
          const int combos = G.parts == 2 ? 3 : 1;
          for (int c = 0; c < combos; ++c) {
            // (hi,hi) (hi,lo) (lo,hi)
            const uint32_t ao = (c == 2) ? a_part : 0, bo = (c == 1) ? b_part : 0;
            for (int ks = 0; ks < G.KCH / 2; ++ks) {
              const uint64_t ad = tc::umma_desc(a_addr + ao + ks * 2 * (128 * 16), 128 * 16, 128);
              const uint64_t bd = tc::umma_desc(b_addr + bo + ks * 2 * (TC_NPAD * 16), TC_NPAD * 16, 128);
              tc::umma_bf16(d_addr, ad, bd, idesc, acc);
              acc = 1;
            }
          }
          tc::umma_commit(t_full + buf);
        }
        tc::umma_commit(b_empty + s);
      }
    }
  } else {
    // ===== epilogue warps =====
    TcEpi E;
    E.sU = sU; E.yeta = M.yeta; E.final_act = M.final_act; E.pair_act = dot ? SGNN_ACT_IDENTITY : M.pair_act;
    E.sim = sim; E.ld = ld; E.m = G.m; E.n = G.n;
    float su = 0.f;
    for (int k = 0; k < G.K; ++k) su += sU[k];
    E.su = su;
    {
      const float hs = REF ? 0.5f * su : 0.5f;
      E.c2 = -M.yeta * hs * hs * 1.4426950408889634f;
    }
    // group `grp` takes tiles it = grp, grp + 2, ...; tile it lives in TMEM buffer it % 4
    const int lg = warp & 3, grp = warp >> 2;
    const int total = nq * ga_count;
    for (int it = grp; it < total; it += 2) {
      const int i = it / ga_count, ga = it - i * ga_count;
      const int buf = it & (TC_BUFS - 1);
      tc::mbar_wait(t_full + buf, (it / TC_BUFS) & 1);
      tc::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + buf * TC_NPAD;
      const int g = (gt0 + ga) * 128 + lg * 32 + lane;
      tc_epilogue_half<KT, RELU, REF, SIMK>(E, taddr, G.K, G.KC, G.QH, (qb0 + i) * G.QH, g, t_empty + buf, lane);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 8) tc::tmem_dealloc(tmem_base, 512);
}
