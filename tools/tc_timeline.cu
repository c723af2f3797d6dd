// Timeline probe of one Dense tile (diagnostics): clock64 stamps of CTA 0 of dense_tc_kernel.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o tools/tc_timeline tools/tc_timeline.cu
#define KGCNN_TC_TIMELINE 1
#include <cstdio>
#include <vector>
namespace kgcnn { void count_launch(int) {} bool pdl_enabled() { return false; } }
#include "../gcnn_keras_b200/csrc/dense_tc.cu"
#include "../gcnn_keras_b200/csrc/dense_fast.cu"
extern "C" const char* kgcnn_error_string(int) { return ""; }

int main(int argc, char** argv) {
  const int64_t M = argc > 1 ? atoll(argv[1]) : 18432;
  const int K = argc > 2 ? atoi(argv[2]) : 128, U = argc > 3 ? atoi(argv[3]) : 128;
  float *x, *w, *b, *y, *y2;
  cudaMalloc(&x, M * K * 4); cudaMalloc(&w, K * U * 4); cudaMalloc(&b, U * 4); cudaMalloc(&y, M * U * 4); cudaMalloc(&y2, M * U * 4);
  cudaMemset(x, 0, M * K * 4); cudaMemset(w, 0, K * U * 4); cudaMemset(b, 0, U * 4);
  const char* names[3] = {"NN", "NT", "NN+act"};
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 5; ++rep) {
      if (mode == 0) kgcnn_dense_tc(0, x, w, M, K, U, b, y, nullptr, 0, nullptr);
      if (mode == 1 && K == U) kgcnn_dense_tc(1, x, w, M, K, U, nullptr, y, nullptr, 0, nullptr);
      if (mode == 2) kgcnn_dense_act(x, w, M, K, U, b, 1, y, y2, nullptr);
    }
    cudaError_t e = cudaDeviceSynchronize();
    long long t[64];
    cudaMemcpyFromSymbol(t, kgcnn::tc::g_timeline, sizeof(t));
    printf("%s M=%lld (%s): cycles since kernel entry (staging thread 0 / MMA lane 0)\n", names[mode], (long long)M, cudaGetErrorString(e));
    const char* lab[] = {"entry", "loads issued", "after sync (tmem)", "stage0 stored", "publish0", "publish1", "publish2", "publish3",
                         "mma0 done", "mma1 done", "mma2 done", "mma3 done", "drained", "epi tile sync", "stores done", "end sync"};
    for (int i = 0; i < 16; ++i) printf("  %-20s %7lld\n", lab[i], t[i] - t[0]);
    printf("  barrier init done %7lld   tmem alloc done (mma warp) %7lld   pointers ready %7lld  row known %7lld  pa known %7lld\n", t[16] - t[0], t[17] - t[0], t[18] - t[0], t[19] - t[0], t[30] - t[0]);
    printf("  LAST ITERATION (relative to its start): staged+published %lld  mma0 done %lld  mma1 done %lld  drained %lld  prev stores read %lld  bar %lld  tile written+sync %lld  end %lld   | mma warp: full0 seen %lld issued %lld\n",
           t[32] - t[31], t[8] - t[31], t[9] - t[31], t[12] - t[31], t[33] - t[31], t[34] - t[31], t[13] - t[31], t[35] - t[31], t[20] - t[31], t[21] - t[31]);
    printf("  LAST ITERATION 2: alias wait done %lld  chunks 0-2 published %lld  all published %lld  mma0..3 done %lld %lld %lld %lld  drained %lld  tile written+sync %lld  end %lld | mma warp full seen/issued: %lld/%lld %lld/%lld %lld/%lld %lld/%lld\n",
           t[36] - t[31], t[37] - t[31], t[32] - t[31], t[8] - t[31], t[9] - t[31], t[10] - t[31], t[11] - t[31], t[12] - t[31], t[13] - t[31], t[35] - t[31],
           t[20] - t[31], t[21] - t[31], t[22] - t[31], t[23] - t[31], t[24] - t[31], t[25] - t[31], t[26] - t[31], t[27] - t[31]);
    for (int c = 0; c < 4; ++c) printf("  mma warp chunk %d: full seen %7lld  issued %7lld\n", c, t[20 + 2 * c] - t[0], t[21 + 2 * c] - t[0]);
  }
  {   // long-reduction loop (TN): steady-state stamps of chunk 16 of CTA 0
    float* wsb; const size_t wb = kgcnn_dense_tc_workspace_bytes(2, M, K, U); cudaMalloc(&wsb, wb);
    for (int rep = 0; rep < 4; ++rep) kgcnn_dense_tc(2, x, y, M, K, U, nullptr, y2, wsb, wb, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    long long t[64];
    cudaMemcpyFromSymbol(t, kgcnn::tc::g_timeline, sizeof(t));
    printf("TN M=%lld (%s) chunk 16, cycles relative to the start of its step (staging thread 0):\n", (long long)M, cudaGetErrorString(e));
    printf("  stage freed (MMA c-4 done) %lld  staged %lld  loads issued + published %lld  MMA(c-1) done %lld  drained %lld  next step starts %lld\n",
           t[41] - t[40], t[42] - t[40], t[43] - t[40], t[44] - t[40], t[45] - t[40], t[46] - t[40]);
    printf("  loads issued %lld  fence.proxy.async done %lld\n", t[47] - t[40], t[48] - t[40]);
    printf("  MMA warp: full(16) seen %lld  MMAs issued %lld   (MMAs of chunk 15 issued at %lld)\n", t[50] - t[40], t[51] - t[40], t[52] - t[40]);
  }
  return 0;
}
