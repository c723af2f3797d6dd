// Timeline probe of one fused-cfconv tile (diagnostics): clock64 stamps of CTA 0 of cfconv_kernel.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o tools/cf_timeline tools/cf_timeline.cu
#define KGCNN_CF_TIMELINE 1
#include <cstdio>
#include <vector>
namespace kgcnn { void count_launch(int) {} bool pdl_enabled() { return false; } }
#include "../gcnn_keras_b200/csrc/cfconv.cu"
extern "C" const char* kgcnn_error_string(int) { return ""; }

int main(int argc, char** argv) {
  const int64_t E = argc > 1 ? atoll(argv[1]) : 18432;
  const int deg = argc > 2 ? atoi(argv[2]) : 8;
  const int64_t N = E / deg;
  const int R = 25;
  std::vector<int> rowptr(N + 1), rows(E), other(E);
  for (int64_t i = 0; i <= N; ++i) rowptr[i] = (int)(i * deg);
  for (int64_t e = 0; e < E; ++e) { rows[e] = (int)(e / deg); other[e] = (int)((e * 7) % N); }
  std::vector<float> dist(E, 2.0f);
  int *d_rowptr, *d_rows, *d_other;
  float *d_dist, *w1, *w2, *bb, *a, *out, *eo, *ws, *wgbuf;
  cudaMalloc(&d_rowptr, (N + 1) * 4); cudaMalloc(&d_rows, E * 4); cudaMalloc(&d_other, E * 4); cudaMalloc(&d_dist, E * 4);
  cudaMemcpy(d_rowptr, rowptr.data(), (N + 1) * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_rows, rows.data(), E * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_other, other.data(), E * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_dist, dist.data(), E * 4, cudaMemcpyHostToDevice);
  cudaMalloc(&w1, 32 * 128 * 4); cudaMalloc(&w2, 128 * 128 * 4); cudaMalloc(&bb, 128 * 4); cudaMalloc(&a, N * 128 * 4);
  cudaMalloc(&out, N * 128 * 4); cudaMalloc(&eo, E * 4);
  const size_t wsb = kgcnn_cfconv_fused_workspace_bytes(E);
  cudaMalloc(&ws, wsb); cudaMalloc(&wgbuf, (size_t)E * 128 * 4 * 8);
  cudaMemset(w1, 0, 32 * 128 * 4); cudaMemset(w2, 0, 128 * 128 * 4); cudaMemset(bb, 0, 512); cudaMemset(a, 0, N * 128 * 4);
  cudaMemset(eo, 0, E * 4);
  float* wg[8];
  for (int i = 0; i < 8; ++i) wg[i] = wgbuf + (size_t)i * E * 128;
  const char* names[6] = {"mode0", "mode1", "mode2", "mode3", "mode1+WG", "mode3+WG"};
  for (int v = 0; v < 6; ++v) {
    const int mode = v < 4 ? v : (v == 4 ? 1 : 3);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int rc = 0;
    for (int rep = 0; rep < 6; ++rep) {
      if (rep == 5) cudaEventRecord(e0);
      rc = kgcnn_cfconv_fused(mode, d_dist, E, R, 3.125f, 0.f, 0.2f, w1, bb, w2, bb, 128, 1, 0, d_rowptr, d_rows, d_other, nullptr, N,
                              a, a, a, eo, out, eo, v >= 4 ? wg : nullptr, ws, wsb, nullptr);
      if (rep == 5) cudaEventRecord(e1);
    }
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long t[64];
    cudaMemcpyFromSymbol(t, kgcnn::cf::g_timeline, sizeof(t));
    printf("%s E=%lld rc=%d (%s) kernel+fixup %.1f us: cycles since entry (worker thread 0)\n", names[v], (long long)E, rc, cudaGetErrorString(e), ms * 1e3);
    const char* lab[] = {"entry", "W loads issued", "W staged", "after sync", "metadata", "rbf published", "g1 seen", "h0 published",
                         "h1 published", "h2 published", "g2 seen", "epilogue done", "after bar", "g3 seen", "wg done", "end"};
    for (int i = 0; i < 16; ++i) printf("  %-16s %7lld\n", lab[i], t[i] - t[0]);
    printf("  startup: params touched %lld  W loads issued %lld  bias loaded %lld  bias stored %lld\n", t[16] - t[0], t[17] - t[0], t[1] - t[0], t[18] - t[0]);
    printf("  mma warp: rbf seen %lld g1 issued %lld | h0 seen %lld issued %lld | h1 seen %lld issued %lld | h2 seen %lld issued %lld\n",
           t[20] - t[0], t[21] - t[0], t[22] - t[0], t[23] - t[0], t[24] - t[0], t[25] - t[0], t[26] - t[0], t[27] - t[0]);
    for (int i = 0; i < 64; ++i) t[i] = 0;
    cudaMemcpyToSymbol(kgcnn::cf::g_timeline, t, sizeof(t));
  }
  return 0;
}
