// Timeline of the tcgen05 all-pairs kernel on CTA 0: per tile, clock64 stamps of
//   0 MMA thread reaches tile | 1 TMEM buffer free | 2 MMAs + commit issued |
//   3 epilogue (group, lane-quadrant 0) reaches tile | 4 accumulator ready | 6 tile stored
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -DSGNN_TC_TRACE -o tools/tc_trace tools/tc_trace.cu
#include <cstdio>
#include <vector>
#include <algorithm>
#include "../graphembedding_b200/csrc/sgnn_allpairs_tc.cuh"

int main(int argc, char** argv) {
  const int parts = argc > 1 ? atoi(argv[1]) : 2;
  ModelDev M; memset(&M, 0, sizeof(M));
  M.pair_kind = SGNN_LAYER_NTN; M.D = 10; M.K = 10; M.emb_dim = 10; M.pair_act = SGNN_ACT_RELU; M.pair_bias = 1;
  M.final_act = SGNN_ACT_SIM_KERNEL; M.yeta = 0.6f; M.ntn_mode = SGNN_NTN_REFERENCE;
  M.w_off = 0; M.v_off = 1000; M.u_off = 1200; M.b_off = 1210; M.n_params = 1220;
  const int m = 1000, n = 10000;
  TcGeom G; G.m = m; G.n = n; G.D = 10; G.K = 10; G.KC = 11; G.QH = 11; G.KCH = 2; G.parts = parts;
  G.n_gtiles = (n + 127) / 128; G.n_qblocks = (m + G.QH - 1) / G.QH; G.GA = 4;
  const int g_groups = (G.n_gtiles + 3) / 4;
  int q_groups = 148 / g_groups;
  G.qb_per_group = (G.n_qblocks + q_groups - 1) / q_groups;
  q_groups = (G.n_qblocks + G.qb_per_group - 1) / G.qb_per_group; G.q_groups = q_groups;
  float *params, *eq, *ed, *sim; __nv_bfloat16 *A, *B; long long* trace;
  cudaMalloc(&params, 1220 * 4); cudaMalloc(&eq, m * 10 * 4); cudaMalloc(&ed, n * 10 * 4);
  cudaMalloc(&sim, (size_t)m * n * 4);
  std::vector<float> h(n * 10);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (float)((i * 2654435761u) % 1000) / 1000.f - 0.3f;
  cudaMemcpy(params, h.data(), 1220 * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(eq, h.data(), m * 10 * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(ed, h.data(), n * 10 * 4, cudaMemcpyHostToDevice);
  const size_t na = (size_t)G.n_gtiles * 2 * G.KCH * 1024, nb = (size_t)G.n_qblocks * 2 * G.KCH * TC_NPAD * 8;
  cudaMalloc(&A, na * 2); cudaMalloc(&B, nb * 2); cudaMemset(A, 0, na * 2); cudaMemset(B, 0, nb * 2);
  const int ntr = 64 * 8;
  cudaMalloc(&trace, ntr * 8); cudaMemset(trace, 0, ntr * 8);
  {
    const int C = G.KCH * 8 * G.KC, col_blocks = (C + (PREP_COLS / G.KC) * G.KC - 1) / ((PREP_COLS / G.KC) * G.KC);
    const int q_blocks = col_blocks * ((m + PREP_Q - 1) / PREP_Q);
    const int db_blocks = (G.n_gtiles * 128 * G.KCH + PREP_COLS - 1) / PREP_COLS;
    const size_t ps = ((size_t)(G.D + 1) * PREP_COLS + (size_t)G.D * PREP_Q) * sizeof(float);
    cudaFuncSetAttribute(tc_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ps);
    tc_prep_kernel<<<q_blocks + db_blocks, PREP_COLS, ps>>>(M, G, q_blocks, col_blocks, params, eq, ed, A, B);
  }
  G.stages = std::min(G.qb_per_group, TC_MAX_STAGES);
  size_t smem = (size_t)TC_GA_MAX * tc_a_bytes(G) + (size_t)G.stages * tc_b_bytes(G) + TC_TAIL_BYTES + 1024;
  smem = std::max<size_t>(smem, 120 * 1024);
  auto kern = allpairs_tc_kernel<10, true, true, true>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int r = 0; r < 3; ++r) {
#ifdef SGNN_TC_TRACE
    kern<<<g_groups * q_groups, TC_THREADS, smem>>>(M, G, params, A, B, sim, n, trace);
#else
    kern<<<g_groups * q_groups, TC_THREADS, smem>>>(M, G, params, A, B, sim, n);
#endif
  }
  printf("sync: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  printf("bars at smem offset %zu (sA base + ...), a_bytes %d b_bytes %d stages %d\n", (size_t)TC_GA_MAX * tc_a_bytes(G) + (size_t)G.stages * tc_b_bytes(G), tc_a_bytes(G), tc_b_bytes(G), G.stages);
  std::vector<long long> t(ntr);
  cudaMemcpy(t.data(), trace, ntr * 8, cudaMemcpyDeviceToHost);
  const long long t0 = t[0];
  printf("tile  mma_reach buf_free issued | epi_reach acc_ready stored   (cycles since first MMA reach)\n");
  for (int i = 0; i < 52; ++i)
    printf("%3d  [ld_issued %8lld ld_done %8lld] %8lld %8lld %8lld | %8lld %8lld %8lld\n", i, t[i * 8 + 5] - t0, t[i * 8 + 7] - t0, t[i * 8] - t0, t[i * 8 + 1] - t0, t[i * 8 + 2] - t0,
           t[i * 8 + 3] - t0, t[i * 8 + 4] - t0, t[i * 8 + 6] - t0);
  return 0;
}
