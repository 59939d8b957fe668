// All-pairs scorer for SimGNN-style chains on the tensor cores: the tcgen05 GEMM of sgnn_allpairs_tc.cuh delivers the
// NTN pre-activations m_k(q, g) (layers.py:282-310, factorised per query) of 128 database graphs x 8 queries per
// tile, and the epilogue -- thread = database graph, two queries at a time -- runs the rest of the chain out of TMEM:
//   f_k = U_k * inneract(m_k);  hist_b = counts[q][g][b] / (N_q N_g)  (bin counts of the tensor-core histogram,
//   sgnn_hist_tc.cuh);  Dense head 32 -> 16 -> 8 -> 4 -> 1 (layers.py:192-205 on the pair row);  final_act.
// The head is BUILD-DEFINED (absent from the reference snapshot; pin = oracle head_layers /
// pair_forward).  Head weights live in shared memory (U folded into the first 16 rows of layer 0) and are read as
// warp-uniform 16-byte broadcasts; all multiply-adds are packed FFMA2 (fma.rn.f32x2), two outputs per instruction.
// A count byte becomes a scaled float with one PRMT + one FMA:  fma(as_float(0x4B000000 | c), s, -2^23 s) = c s
// exactly rounded.  Shapes served: K = 16 feature maps, head widths 16 / 8 / 4 / 1, 16 bins or no histogram;
// everything else stays on the CUDA-core kernel (sgnn_head.cu).
#pragma once
#include <cuda_fp16.h>
#include "sgnn_allpairs_tc.cuh"

constexpr int TCH_K = 16, TCH_H0 = 16, TCH_H1 = 8, TCH_H2 = 4;
// shared-memory weight block (floats)
constexpr int TCH_W0 = 0;                              // [32][16]  rows 0..15: U_k W0[k][:], rows 16..31: W0[16 + b][:]
constexpr int TCH_B0 = TCH_W0 + 32 * TCH_H0;           // [16]
constexpr int TCH_W1 = TCH_B0 + TCH_H0;                // [16][8]
constexpr int TCH_B1 = TCH_W1 + TCH_H0 * TCH_H1;       // [8]
constexpr int TCH_W2 = TCH_B1 + TCH_H1;                // [8][4]
constexpr int TCH_B2 = TCH_W2 + TCH_H1 * TCH_H2;       // [4]
constexpr int TCH_W3 = TCH_B2 + TCH_H2;                // [4]
constexpr int TCH_B3 = TCH_W3 + TCH_H2;                // [1] (+3 pad)
constexpr int TCH_FLOATS = TCH_B3 + 4;
constexpr int TCH_SMEM_BYTES = TCH_FLOATS * 4 + 16;   // + alignment slack
// The same block in CONSTANT memory (one slot per engine of the device): a warp-uniform constant operand reaches the
// FFMA2 through a uniform register (LDCU.128 -> FFMA2 R, R.F32, UR.F32x2, R), per scheduler, while a shared-memory
// broadcast costs 4 cycles of the SM-wide LSU per LDS.128 -- the shared-memory version of this kernel spent 0.39 of
// its 0.43 ms delivering weights.  The slot is refreshed from the parameters by a copy on the stream
// before every scoring call (tch_weights_kernel -> staging -> cudaMemcpyToSymbolAsync), so training between calls is fine.
constexpr int TCH_SLOTS = 8;
constexpr int TCH_SLOT_FLOATS = 768;   // >= TCH_FLOATS, 3 KB
static_assert(TCH_SLOT_FLOATS >= TCH_FLOATS, "slot too small");
__constant__ __align__(16) float c_tch_w[TCH_SLOTS * TCH_SLOT_FLOATS];

struct TcHeadEpi {
  const float* sW;                 // shared-memory weight block
  int slot;                        // constant-memory slot (CONSTW)
  int act_ntn, act0, act1, act2, act3, final_act;
  float yeta;
  float* sim;
  int64_t ld;
  int m, n;
  float hscale;                    // HMMA: inverse of the power-of-two scale of the fp16 weight parts
  const uint8_t* cnt;              // [m][n][16] bin counts (HIST)
  const int32_t* node_off_q;
  const int32_t* node_off_db;
  unsigned long long* smax;        // optional [m][n_slices] slice maxima (rank keys) for the hinted top-k, as TcEpi::smax
  int n_slices;
};

// element i of the head weight block: U folded into the NTN rows of layer 0, missing biases = 0
__device__ __forceinline__ float tch_weight(const ModelDev& M, const float* __restrict__ params, int i) {
  float w = 0.f;
  if (i < TCH_B0) {
    const int r = i / TCH_H0, j = i - r * TCH_H0;
    if (r < M.Hd[0].din) w = params[M.Hd[0].w_off + r * TCH_H0 + j] * (r < TCH_K ? params[M.u_off + r] : 1.f);
  } else if (i < TCH_W1) { if (M.Hd[0].bias) w = params[M.Hd[0].b_off + (i - TCH_B0)]; }
  else if (i < TCH_B1) w = params[M.Hd[1].w_off + (i - TCH_W1)];
  else if (i < TCH_W2) { if (M.Hd[1].bias) w = params[M.Hd[1].b_off + (i - TCH_B1)]; }
  else if (i < TCH_B2) w = params[M.Hd[2].w_off + (i - TCH_W2)];
  else if (i < TCH_W3) { if (M.Hd[2].bias) w = params[M.Hd[2].b_off + (i - TCH_B2)]; }
  else if (i < TCH_B3) w = params[M.Hd[3].w_off + (i - TCH_W3)];
  else if (i == TCH_B3) { if (M.Hd[3].bias) w = params[M.Hd[3].b_off]; }
  return w;
}
__global__ void __launch_bounds__(256)
tch_weights_kernel(const __grid_constant__ ModelDev M, const float* __restrict__ params, float* __restrict__ staging) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < TCH_SLOT_FLOATS) staging[i] = i < TCH_FLOATS ? tch_weight(M, params, i) : 0.f;
}

// Packed pairs of floats live in 64-bit registers (fma.rn.f32x2 operands): with float2 the compiler keeps the halves
// as separate scalars and shuffles them into aligned pairs around every FFMA2 (300 MOVs per pass).
typedef unsigned long long tch_f2;
__device__ __forceinline__ tch_f2 tch_pack(float lo, float hi) {
  tch_f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void tch_unpack(tch_f2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void tch_fma2(tch_f2& d, tch_f2 a, tch_f2 b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
// activations other than relu / identity are out of line: the epilogue has to stay small (instruction fetch)
__device__ __noinline__ float tch_slow_act(float x, int act, float yeta) { return sgnn_act(x, act, yeta); }
__device__ __forceinline__ float tch_act1(float x, int act, float yeta) {
  return act == SGNN_ACT_RELU ? fmaxf(x, 0.f) : (act == SGNN_ACT_IDENTITY ? x : tch_slow_act(x, act, yeta));
}
template <int N>
__device__ __forceinline__ void tch_act(tch_f2* a, int act, float yeta) {
  if (act == SGNN_ACT_IDENTITY) return;
#pragma unroll
  for (int i = 0; i < N; ++i) {
    float lo, hi;
    tch_unpack(a[i], lo, hi);
    a[i] = tch_pack(tch_act1(lo, act, yeta), tch_act1(hi, act, yeta));
  }
}
// the same weight row (NW x 16 bytes = 4 NW columns) applied to two pairs: one shared-memory read per 16 bytes
template <int NW>
__device__ __forceinline__ void tch_row2(tch_f2* accA, tch_f2* accB, float fa, float fb, const float4* __restrict__ w) {
  const tch_f2 a2 = tch_pack(fa, fa), b2 = tch_pack(fb, fb);
  const ulonglong2* w2 = reinterpret_cast<const ulonglong2*>(w);
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const ulonglong2 ww = w2[j];
    tch_fma2(accA[2 * j], a2, ww.x);
    tch_fma2(accA[2 * j + 1], a2, ww.y);
    tch_fma2(accB[2 * j], b2, ww.x);
    tch_fma2(accB[2 * j + 1], b2, ww.y);
  }
}
template <int NW>
__device__ __forceinline__ void tch_bias2(tch_f2* a, tch_f2* b, const float* __restrict__ src) {
  const ulonglong2* w2 = reinterpret_cast<const ulonglong2*>(src);
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const ulonglong2 ww = w2[j];
    a[2 * j] = b[2 * j] = ww.x;
    a[2 * j + 1] = b[2 * j + 1] = ww.y;
  }
}
__device__ __forceinline__ float tch_half(tch_f2 v, int hi) {
  float a, b;
  tch_unpack(v, a, b);
  return hi ? b : a;
}
__device__ __forceinline__ float tch_count(unsigned word, int byte, float s, float ms) {   // count * s, exactly rounded
  return fmaf(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | (unsigned)byte)), s, ms);
}

// layers 1.. of the head on two pairs' layer-0 outputs (weight rows read once for both); final similarities in oa / ob
// FAST: the activations are the SimGNN-style chain's (relu, relu, relu, identity, final sigmoid) at compile time -- no
// run-time dispatch, no out-of-line call sites (they cost register moves around every call), the sigmoid in line
template <bool FAST>
__device__ __forceinline__ void tch_tail2(const TcHeadEpi& E, const float* sW, tch_f2* a0, tch_f2* b0, float& oa, float& ob) {
  tch_act<TCH_H0 / 2>(a0, FAST ? (int)SGNN_ACT_RELU : E.act0, E.yeta);
  tch_act<TCH_H0 / 2>(b0, FAST ? (int)SGNN_ACT_RELU : E.act0, E.yeta);
  tch_f2 a1[TCH_H1 / 2], b1[TCH_H1 / 2];
  tch_bias2<TCH_H1 / 4>(a1, b1, sW + TCH_B1);
#pragma unroll
  for (int k = 0; k < TCH_H0; ++k)
    tch_row2<TCH_H1 / 4>(a1, b1, tch_half(a0[k >> 1], k & 1), tch_half(b0[k >> 1], k & 1),
                         reinterpret_cast<const float4*>(sW + TCH_W1 + k * TCH_H1));
  tch_act<TCH_H1 / 2>(a1, FAST ? (int)SGNN_ACT_RELU : E.act1, E.yeta);
  tch_act<TCH_H1 / 2>(b1, FAST ? (int)SGNN_ACT_RELU : E.act1, E.yeta);
  tch_f2 a2[TCH_H2 / 2], b2[TCH_H2 / 2];
  tch_bias2<TCH_H2 / 4>(a2, b2, sW + TCH_B2);
#pragma unroll
  for (int k = 0; k < TCH_H1; ++k)
    tch_row2<TCH_H2 / 4>(a2, b2, tch_half(a1[k >> 1], k & 1), tch_half(b1[k >> 1], k & 1),
                         reinterpret_cast<const float4*>(sW + TCH_W2 + k * TCH_H2));
  tch_act<TCH_H2 / 2>(a2, FAST ? (int)SGNN_ACT_RELU : E.act2, E.yeta);
  tch_act<TCH_H2 / 2>(b2, FAST ? (int)SGNN_ACT_RELU : E.act2, E.yeta);
  const float4 w3 = *reinterpret_cast<const float4*>(sW + TCH_W3);
  const float c3 = sW[TCH_B3];
  oa = fmaf(tch_half(a2[1], 1), w3.w, fmaf(tch_half(a2[1], 0), w3.z, fmaf(tch_half(a2[0], 1), w3.y, fmaf(tch_half(a2[0], 0), w3.x, c3))));
  ob = fmaf(tch_half(b2[1], 1), w3.w, fmaf(tch_half(b2[1], 0), w3.z, fmaf(tch_half(b2[0], 1), w3.y, fmaf(tch_half(b2[0], 0), w3.x, c3))));
  if (FAST) {   // the expression of sgnn_act(SGNN_ACT_SIGMOID): the same bits as the generic kernel
    oa = 1.f / (1.f + expf(-oa));
    ob = 1.f / (1.f + expf(-ob));
  } else {
    oa = tch_slow_act(tch_act1(oa, E.act3, E.yeta), E.final_act, E.yeta);
    ob = tch_slow_act(tch_act1(ob, E.act3, E.yeta), E.final_act, E.yeta);
  }
}

// Epilogue of one tile part: this thread owns TMEM lane (database graph) g and the accumulator columns of queries
// [qrow0 + 4 part, + 4) of the tile, processed as two passes of two queries.  The code is kept SMALL on purpose (part
// is a run-time value, the k loops are rolled in chunks): 16 epilogue warps at different places of a fully unrolled
// 50 KB epilogue starved on instruction fetch (ncu r2d: stall_no_instruction 5.7 per issue).
// HMMA: the histogram half of layer 0 was done by the tensor core (second accumulator at taddr2: 16 columns per
// query = sum_b count_b W0[16 + b][:]); the epilogue only scales it by 1 / (N_q N_g) into the layer-0 accumulators.
// FAST: the specialised activations of tch_tail2 (measured: 0.923 -> 0.901 ms per ranking step; reading the weights
// from a compile-time constant slot on top of it, which turns half of the LDCU.64 into LDCU.128, gave 0.900, and
// unrolling the k chunks for the other half spilled: 0.912 -- both dropped)
template <bool HIST, bool CONSTW, bool HMMA, bool FAST>
__device__ __forceinline__ void tc_head_epilogue_part(const TcHeadEpi& E, uint32_t taddr, uint32_t taddr2, int qrow0,
                                                      int part, int g, uint64_t* t_empty, int lane) {
  const bool gv = g < E.n;
  int n_g = 1;
  if (HIST && gv) n_g = E.node_off_db[g + 1] - E.node_off_db[g];
#pragma unroll 1
  for (int qp = 0; qp < 2; ++qp) {
    int opaque = 0;
    asm volatile("" : "+r"(opaque));   // per pass: the weight reads must not be hoisted out of the loop (register pressure)
    // (an offset, so that the pointer keeps its address space: LDS / LDCU, not a generic LD)
    const float* sW = CONSTW ? c_tch_w + E.slot * TCH_SLOT_FLOATS : E.sW + opaque;
    const int act_ntn = FAST ? (int)SGNN_ACT_RELU : E.act_ntn;
    const int qa = qrow0 + 4 * part + 2 * qp, qb = qa + 1;
    const bool va = qa < E.m, vb = qb < E.m;
    uint4 ca = make_uint4(0u, 0u, 0u, 0u), cb = ca;
    int nqa = 0, nqb = 0;
    if (HIST && !HMMA) {   // issued first: the loads are in flight while the NTN half runs
      if (gv && va) ca = __ldg(reinterpret_cast<const uint4*>(E.cnt + ((size_t)qa * E.n + g) * 16));
      if (gv && vb) cb = __ldg(reinterpret_cast<const uint4*>(E.cnt + ((size_t)qb * E.n + g) * 16));
    }
    if (HIST) {
      if (va) nqa = E.node_off_q[qa + 1] - E.node_off_q[qa];
      if (vb) nqb = E.node_off_q[qb + 1] - E.node_off_q[qb];
    }
    tch_f2 A[TCH_H0 / 2], B[TCH_H0 / 2];
    tch_bias2<TCH_H0 / 4>(A, B, sW + TCH_B0);
    // the two queries' m_k in chunks of 8 feature maps (16 live accumulator values instead of 32; the other epilogue
    // warps of the scheduler cover the TMEM load latency)
    const uint32_t t0 = taddr + (uint32_t)((4 * part + 2 * qp) * TCH_K);
#pragma unroll 1
    for (int kc = 0; kc < TCH_K; kc += 8) {
      float va8[8], vb8[8];
      tc::tmem_ld8(t0 + kc, va8);
      tc::tmem_ld8(t0 + kc + TCH_K, vb8);
      tc::tmem_ld_wait();
      if (!HMMA && qp == 1 && kc == TCH_K - 8) {   // last TMEM read of this part: hand the buffer back to its MMA issuer
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(t_empty);
      }
      const float4* w = reinterpret_cast<const float4*>(sW + TCH_W0 + kc * TCH_H0);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        tch_row2<TCH_H0 / 4>(A, B, tch_act1(va8[k], act_ntn, E.yeta), tch_act1(vb8[k], act_ntn, E.yeta),
                             w + k * (TCH_H0 / 4));
    }
    if (HIST) {
      const float sa = nqa * n_g > 0 ? 1.f / (float)(nqa * n_g) : 0.f;
      const float sb = nqb * n_g > 0 ? 1.f / (float)(nqb * n_g) : 0.f;
      if (HMMA) {
        float ha[TCH_H0], hb[TCH_H0];
        const uint32_t t2 = taddr2 + (uint32_t)((4 * part + 2 * qp) * TCH_H0);
        tc::tmem_ld16(t2, ha);
        tc::tmem_ld16(t2 + TCH_H0, hb);
        tc::tmem_ld_wait();
        if (qp == 1) {   // last TMEM read of this part (both accumulators of the tile go back together)
          tc::tc_fence_before();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(t_empty);
        }
        const float sah = sa * E.hscale, sbh = sb * E.hscale;
        const tch_f2 sa2 = tch_pack(sah, sah), sb2 = tch_pack(sbh, sbh);
#pragma unroll
        for (int j = 0; j < TCH_H0 / 2; ++j) {
          tch_fma2(A[j], sa2, tch_pack(ha[2 * j], ha[2 * j + 1]));
          tch_fma2(B[j], sb2, tch_pack(hb[2 * j], hb[2 * j + 1]));
        }
      }
      const float msa = -8388608.f * sa, msb = -8388608.f * sb;
#pragma unroll 1
      for (int wd = 0; wd < (HMMA ? 0 : 4); ++wd) {   // four bins per count word; the words rotate through .x
        const float4* w = reinterpret_cast<const float4*>(sW + TCH_W0 + (TCH_K + 4 * wd) * TCH_H0);
#pragma unroll
        for (int b = 0; b < 4; ++b)
          tch_row2<TCH_H0 / 4>(A, B, tch_count(ca.x, b, sa, msa), tch_count(cb.x, b, sb, msb), w + b * (TCH_H0 / 4));
        ca.x = ca.y; ca.y = ca.z; ca.z = ca.w;
        cb.x = cb.y; cb.y = cb.z; cb.z = cb.w;
      }
    }
    float oa, ob;
    tch_tail2<FAST>(E, sW, A, B, oa, ob);
    if (gv && va) E.sim[(size_t)qa * E.ld + g] = oa;
    if (gv && vb) E.sim[(size_t)qb * E.ld + g] = ob;
    if (E.smax) {   // largest rank key of this warp's 32-graph slice, per query (ties -> highest graph id)
      const float ma = gv ? oa : -INFINITY, mb = gv ? ob : -INFINITY;
      const float wa = tc_warp_max(ma), wb = tc_warp_max(mb);
      const unsigned ha = __ballot_sync(0xffffffffu, !(ma < wa)), hb = __ballot_sync(0xffffffffu, !(mb < wb));
      if (lane == 0) {
        const size_t sl = (size_t)(g >> 5);
        if (va) E.smax[(size_t)qa * E.n_slices + sl] = tc_rank_key(wa, (g & ~31) + 31 - __clz(ha | 1u));
        if (vb) E.smax[(size_t)qb * E.n_slices + sl] = tc_rank_key(wb, (g & ~31) + 31 - __clz(hb | 1u));
      }
    }
  }
}

// Producer and MMA roles of the head kernel: ONE issuer warp for all TMEM buffers (a tile's epilogue runs for
// ~20 us, the issue rate does not matter), so the CTA has 18 warps and the epilogue gets 96 registers per thread
// instead of the 80 that the 21-warp scorer leaves (register allocation is per 4 warps).
// HMMA (second MMA per tile): TMEM holds two NTN accumulators (columns 0..255) and two histogram accumulators
// (256..511), one of each per epilogue group.  The bin counts arrive as fp16 K-major operand tiles written by the
// histogram kernel (ht_cnt_tiled_bytes: one 4 KB block per (database tile, query), the 8 queries of a query block
// contiguous), one 32 KB bulk copy per tile into a two-slot ring; the layer-0 rows of the histogram bins are a
// three-way fp16 split staged in shared memory once per CTA.  Per query: D2[128 graphs][16] (+)= counts
// [128][16 bins] x W_part[16 bins][16] for the three parts -- integers <= 255 times fp16 parts, accumulated in fp32.
constexpr int TCH_THREADS = 32 * (TC_EPI_WARPS + 2);
constexpr int TCH_MMA_WARP = TC_EPI_WARPS + 1;
constexpr int TCH_A2_BYTES = 8 * 4096;          // one tile's count blocks (8 queries)
constexpr int TCH_WH_BYTES = 3 * 512;           // three fp16 parts of W0[16 + b][j], K-major 16 x 16
constexpr int TCH_HMMA_SMEM = 2 * TCH_A2_BYTES + TCH_WH_BYTES + 64;
struct TchHmma {
  uint8_t* sA2;          // [2][TCH_A2_BYTES]
  uint8_t* sWh;          // [3][512]
  uint64_t* a2_full;     // [2]
  uint64_t* a2_empty;    // [2]
  const uint8_t* cnt;    // operand tiles in global memory
  int m_pad;
};
__device__ __forceinline__ uint32_t tch_idesc_f16(int M, int N) {   // A fp16, B fp16, D fp32, both K-major
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // (fp16 x bf16 is an illegal instruction)
}
// Work of a CTA: a contiguous range [u0, u1) of units u = (A block ab, query block qb), ab-major -- an even split of ALL
// units over the grid (the 2-D split of the scorer, g-groups x q-groups, left 10 % of this kernel's time to imbalance:
// 79 database tiles in blocks of 4 and 128 query blocks over <= 148 CTAs gave 76 tile-units to most CTAs against a
// mean of 68).  A range that crosses into the next A block reloads the resident database tiles once (a_free).
struct TchRange {
  int u0, u1, n_qb, GA, n_gtiles;
  __device__ __forceinline__ int ab_first() const { return u0 / n_qb; }
  __device__ __forceinline__ int ab_last() const { return (u1 - 1) / n_qb; }
  __device__ __forceinline__ void segment(int ab, int& gt0, int& ga_count, int& qb0, int& nq) const {
    gt0 = ab * GA; ga_count = min(GA, n_gtiles - gt0);
    const int lo = max(u0, ab * n_qb), hi = min(u1, (ab + 1) * n_qb);
    qb0 = lo - ab * n_qb; nq = hi - lo;
  }
};

template <bool HMMA>
__device__ __forceinline__ void tch_producer_mma_roles(const TcGeom& G, const TchRange& R, const TcSmem& S,
                                                       const TchHmma& H, uint64_t* a_free,
                                                       const __nv_bfloat16* __restrict__ Ag,
                                                       const __nv_bfloat16* __restrict__ Bg, uint32_t tmem_base, int warp,
                                                       int lane) {
  constexpr int NB = HMMA ? 2 : TC_BUFS;
  if (lane != 0 || R.u1 <= R.u0) return;
  if (warp == TC_PROD_WARP) {
    int it = 0, qi = 0, seg = 0;
    for (int ab = R.ab_first(); ab <= R.ab_last(); ++ab, ++seg) {
      int gt0, ga_count, qb0, nq;
      R.segment(ab, gt0, ga_count, qb0, nq);
      if (seg > 0) tc::mbar_wait(a_free, (seg - 1) & 1);   // every MMA on the previous A block has completed
      tc::mbar_arrive_expect_tx(S.a_full, ga_count * S.a_bytes);
      for (int ga = 0; ga < ga_count; ++ga)
        tc::bulk_g2s(S.sA + ga * S.a_bytes, reinterpret_cast<const uint8_t*>(Ag) + (size_t)(gt0 + ga) * S.a_bytes,
                     S.a_bytes, S.a_full);
      for (int i = 0; i < nq; ++i, ++qi) {
        const int s = qi % G.stages;
        if (qi >= G.stages) tc::mbar_wait(S.b_empty + s, ((qi / G.stages) - 1) & 1);
        tc::mbar_arrive_expect_tx(S.b_full + s, S.b_bytes);
        tc::bulk_g2s(S.sB + s * S.b_bytes, reinterpret_cast<const uint8_t*>(Bg) + (size_t)(qb0 + i) * S.b_bytes,
                     S.b_bytes, S.b_full + s);
        if (HMMA) {
          for (int ga = 0; ga < ga_count; ++ga, ++it) {
            const int sl = it & 1;
            if (it >= 2) tc::mbar_wait(H.a2_empty + sl, ((it >> 1) - 1) & 1);
            tc::mbar_arrive_expect_tx(H.a2_full + sl, TCH_A2_BYTES);
            tc::bulk_g2s(H.sA2 + sl * TCH_A2_BYTES,
                         H.cnt + ((size_t)(gt0 + ga) * H.m_pad + (size_t)(qb0 + i) * 8) * 4096, TCH_A2_BYTES,
                         H.a2_full + sl);
          }
        }
      }
    }
    return;
  }
  const uint32_t idesc = tc::umma_idesc_bf16(128, TC_NPAD);
  const uint32_t a_part = (G.KCH * 128 * 16) >> 4, b_part = (G.KCH * TC_NPAD * 16) >> 4;   // in 16-byte units
  const uint32_t a_ks = (2 * 128 * 16) >> 4, b_ks = (2 * TC_NPAD * 16) >> 4;
  const uint64_t a_desc0 = tc::umma_desc(tc::smem_u32(S.sA), 128 * 16, 128);
  const uint64_t b_desc0 = tc::umma_desc(tc::smem_u32(S.sB), TC_NPAD * 16, 128);
  const uint32_t a_tile = (uint32_t)S.a_bytes >> 4, b_tile = (uint32_t)S.b_bytes >> 4;
  const int combos = G.parts == 2 ? 3 : 1, ksteps = G.KCH / 2;
  const uint32_t idesc2 = tch_idesc_f16(128, TCH_H0);
  uint64_t a2_desc0 = 0, wh_desc0 = 0;
  if (HMMA) {
    a2_desc0 = tc::umma_desc(tc::smem_u32(H.sA2), 128 * 16, 128);
    wh_desc0 = tc::umma_desc(tc::smem_u32(H.sWh), TCH_H0 * 16, 128);
  }
  int it = 0, qi = 0, seg = 0;
  for (int ab = R.ab_first(); ab <= R.ab_last(); ++ab, ++seg) {
    int gt0, ga_count, qb0, nq;
    R.segment(ab, gt0, ga_count, qb0, nq);
    tc::mbar_wait(S.a_full, seg & 1);
    for (int i = 0; i < nq; ++i, ++qi) {   // tile order: query block outer, database tile inner (as the epilogue walks it)
      const int s = qi % G.stages;
      tc::mbar_wait(S.b_full + s, (qi / G.stages) & 1);
      const uint64_t bd0 = b_desc0 + (uint64_t)((uint32_t)s * b_tile);
      for (int ga = 0; ga < ga_count; ++ga, ++it) {
        const int buf = it & (NB - 1);
        if (it >= NB) tc::mbar_wait(S.t_empty + buf, ((it / NB) - 1) & 1);
        tc::tc_fence_after();
        const uint64_t ad0 = a_desc0 + (uint64_t)((uint32_t)ga * a_tile);
        const uint32_t d_addr = tmem_base + (uint32_t)buf * TC_NPAD;
        uint32_t acc = 0;
        for (int c = 0; c < combos; ++c) {   // (hi,hi) (hi,lo) (lo,hi)
          const uint32_t ao = c == 2 ? a_part : 0u, bo = c == 1 ? b_part : 0u;
          for (int ks = 0; ks < ksteps; ++ks) {
            tc::umma_bf16(d_addr, ad0 + ao + (uint64_t)(ks * a_ks), bd0 + bo + (uint64_t)(ks * b_ks), idesc, acc);
            acc = 1;
          }
        }
        if (HMMA) {
          const int sl = it & 1;
          tc::mbar_wait(H.a2_full + sl, (it >> 1) & 1);
          const uint64_t a2d = a2_desc0 + (uint64_t)((uint32_t)sl * (TCH_A2_BYTES >> 4));
          const uint32_t d2 = tmem_base + 2u * TC_NPAD + (uint32_t)buf * TC_NPAD;
          for (int q = 0; q < 8; ++q)
            for (int sp = 0; sp < 3; ++sp)
              tc::umma_bf16(d2 + (uint32_t)(q * TCH_H0), a2d + (uint64_t)(q * (4096 >> 4)), wh_desc0 + (uint64_t)(sp * (512 >> 4)),
                            idesc2, sp > 0 ? 1u : 0u);
        }
        tc::umma_commit(S.t_full + buf);
        if (HMMA) tc::umma_commit(H.a2_empty + (it & 1));
      }
      tc::umma_commit(S.b_empty + s);   // the stage is free once this block's MMAs have completed
    }
    tc::umma_commit(a_free);
  }
}

template <bool HIST, bool CONSTW, bool HMMA, bool FAST = false>
__global__ void __launch_bounds__(TCH_THREADS, 1)
allpairs_tc_head_kernel(const __grid_constant__ ModelDev M, const TcGeom G, const float* __restrict__ params, int slot,
                        const __nv_bfloat16* __restrict__ Ag, const __nv_bfloat16* __restrict__ Bg,
                        float* __restrict__ sim, int64_t ld, const uint8_t* __restrict__ cnt,
                        const int32_t* __restrict__ node_off_q, const int32_t* __restrict__ node_off_db,
                        unsigned long long* __restrict__ smax) {
  static_assert(!HMMA || HIST, "the second MMA is the histogram half of layer 0");
  constexpr int NB = HMMA ? 2 : TC_BUFS;
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint8_t* smem = tc_smem;
  const int a_bytes = tc_a_bytes(G), b_bytes = tc_b_bytes(G);
  uint8_t* sA = smem;
  uint8_t* sB = sA + TC_GA_MAX * a_bytes;
  uint8_t* sA2 = sB + G.stages * b_bytes;
  uint8_t* sWh = sA2 + (HMMA ? 2 * TCH_A2_BYTES : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sWh + (HMMA ? TCH_WH_BYTES : 0));
  uint64_t* a_full = bars;
  uint64_t* b_full = bars + 1;
  uint64_t* b_empty = b_full + G.stages;
  uint64_t* t_full = b_empty + G.stages;
  uint64_t* t_empty = t_full + TC_BUFS;
  uint64_t* a2_full = t_empty + TC_BUFS;
  uint64_t* a2_empty = a2_full + 2;
  uint64_t* a_free = a2_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_free + 1);
  // TCH_FLOATS, 16-byte aligned (pointer arithmetic, not an integer round trip: the pointer stays a shared-memory pointer)
  float* sW = reinterpret_cast<float*>(tmem_slot + 4) + (((16u - (tc::smem_u32(tmem_slot + 4) & 15u)) & 15u) >> 2);

  // warp index through a shuffle: provably warp-uniform, so that the role branches are uniform control flow and the
  // epilogue's constant operands can go through the uniform datapath (LDCU -> UR operands of FFMA2)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int n_ab = (G.n_gtiles + G.GA - 1) / G.GA;
  const long long total_units = (long long)n_ab * G.n_qblocks;
  TchRange R;
  R.u0 = (int)(total_units * blockIdx.x / gridDim.x); R.u1 = (int)(total_units * (blockIdx.x + 1) / gridDim.x);
  R.n_qb = G.n_qblocks; R.GA = G.GA; R.n_gtiles = G.n_gtiles;

  if (threadIdx.x == 0) {
    tc::mbar_init(a_full, 1);
    tc::mbar_init(a_free, 1);
    for (int s = 0; s < G.stages; ++s) { tc::mbar_init(b_full + s, 1); tc::mbar_init(b_empty + s, 1); }
    for (int b = 0; b < NB; ++b) { tc::mbar_init(t_full + b, 1); tc::mbar_init(t_empty + b, 8); }
    if (HMMA)
      for (int b = 0; b < 2; ++b) { tc::mbar_init(a2_full + b, 1); tc::mbar_init(a2_empty + b, 1); }
    tmem_slot[1] = 0u;
    tc::fence_barrier_init();
  }
  // ---- head weights: U folded into the NTN rows of layer 0, missing biases = 0 ----
  if (!CONSTW)
    for (int i = threadIdx.x; i < TCH_FLOATS; i += blockDim.x) sW[i] = tch_weight(M, params, i);
  if (HMMA) {
    // layer-0 rows of the 16 bins as three fp16 parts, K-major.  The rows are scaled by a power of two that puts
    // the largest |w| in [2^14, 2^15): hi + mid + lo then equals the scaled fp32 weight up to 2^-25 (fp16 subnormal
    // grid), i.e. every weight is exact to 2^-39 of the largest one; the epilogue multiplies the inverse back in.
    uint32_t* s_wmax = tmem_slot + 1;
    float w = 0.f;
    const int b = threadIdx.x >> 4, j = threadIdx.x & 15;
    if (threadIdx.x < 16 * TCH_H0) w = tch_weight(M, params, TCH_W0 + (TCH_K + b) * TCH_H0 + j);
    __syncthreads();   // s_wmax = 0 (thread 0, above)
    if (threadIdx.x < 16 * TCH_H0) atomicMax(s_wmax, __float_as_uint(fabsf(w)));
    __syncthreads();
    const int e = (int)(*s_wmax >> 23) - 127;
    const int S = *s_wmax == 0u ? 0 : max(-100, min(100, 14 - e));
    if (threadIdx.x < 16 * TCH_H0) {
      w *= __uint_as_float((uint32_t)(S + 127) << 23);
      __half* dst = reinterpret_cast<__half*>(sWh + (b >> 3) * 256 + j * 16) + (b & 7);
#pragma unroll
      for (int sp = 0; sp < 3; ++sp) {
        const __half hv = __float2half_rn(w);
        dst[sp * 256] = hv;
        w -= __half2float(hv);
      }
    }
    if (threadIdx.x == 0) *reinterpret_cast<float*>(tmem_slot + 2) = __uint_as_float((uint32_t)(127 - S) << 23);
    tc::fence_proxy_async_smem();
  }
  if (warp == TC_PROD_WARP) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= TC_PROD_WARP) {
    TcSmem S{sA, sB, a_full, b_full, b_empty, t_full, t_empty, a_bytes, b_bytes};
    TchHmma H{sA2, sWh, a2_full, a2_empty, cnt, G.n_qblocks * 8};
    tch_producer_mma_roles<HMMA>(G, R, S, H, a_free, Ag, Bg, tmem_base, warp, lane);
  } else if (R.u1 > R.u0) {
    TcHeadEpi E;
    E.sW = sW; E.slot = slot; E.act_ntn = M.pair_act; E.act0 = M.Hd[0].act; E.act1 = M.Hd[1].act; E.act2 = M.Hd[2].act;
    E.act3 = M.Hd[3].act; E.final_act = M.final_act; E.yeta = M.yeta;
    E.hscale = HMMA ? *reinterpret_cast<const float*>(tmem_slot + 2) : 1.f;
    E.sim = sim; E.ld = ld; E.m = G.m; E.n = G.n; E.cnt = cnt; E.node_off_q = node_off_q; E.node_off_db = node_off_db;
    E.smax = smax; E.n_slices = G.n_gtiles * 4;
    const int lg = warp & 3, part = (warp >> 2) & 1, grp = warp >> 3;
    int it_base = 0;
    for (int ab = R.ab_first(); ab <= R.ab_last(); ++ab) {
      int gt0, ga_count, qb0, nq;
      R.segment(ab, gt0, ga_count, qb0, nq);
      const int total = nq * ga_count;
      for (int lt = (grp - it_base) & (TC_EPI_GROUPS - 1); lt < total; lt += TC_EPI_GROUPS) {
        const int it = it_base + lt;
        const int i = lt / ga_count, ga = lt - i * ga_count;
        const int buf = it & (NB - 1);
        tc::mbar_wait(t_full + buf, (it / NB) & 1);
        tc::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + buf * TC_NPAD;
        const int g = (gt0 + ga) * 128 + lg * 32 + lane;
        tc_head_epilogue_part<HIST, CONSTW, HMMA, FAST>(E, taddr, taddr + 2 * TC_NPAD, (qb0 + i) * G.QH, part, g, t_empty + buf, lane);
      }
      it_base += total;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == TC_PROD_WARP) tc::tmem_dealloc(tmem_base, 512);
}
