// Weight-gradient GEMM of the Dense layers on flat ragged values (kgcnn/layers/modules.py:85 backward):
//     dW[K1,K2] = sum_m X[m,K1]^T G[m,K2]     (+ optional db[K2] = sum_m G[m,K2])
// M is the number of edges / atoms of the batch (1e4 .. 1e6), K1 and K2 are layer widths (1 .. 640): a
// "tall-skinny transposed" product whose output is a handful of 64x64 tiles.  Library GEMMs put one CTA on
// each output tile (4 CTAs on a 148-SM part for 128x128); here the long M axis is split over ~2 CTAs per SM
// and the partial tiles are summed in a fixed order by a second kernel, so the result is deterministic.
// fp32 FMA on CUDA cores: the 1e-5 parity bar rules out single-pass TF32 tensor-core products.
#include <algorithm>

#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kTile = 64;      // output tile edge
constexpr int kStep = 16;      // rows of X / G staged per iteration
constexpr int kThreads = 256;  // 16 x 16 threads, 4 x 4 outputs each

// partial[s, k1, k2] = sum_{m in chunk s} X[m,k1] G[m,k2];  colpart[s, k2] = sum_{m in chunk s} G[m,k2]
__global__ void __launch_bounds__(kThreads)
gemm_tn_partial_kernel(const float* __restrict__ X, const float* __restrict__ G, int64_t M, int K1, int K2,
                       int64_t rows_per_split, float* __restrict__ partial, float* __restrict__ colpart) {
  __shared__ __align__(16) float sa[kStep][kTile];
  __shared__ __align__(16) float sb[kStep][kTile];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int k1_0 = blockIdx.x * kTile, k2_0 = blockIdx.y * kTile;
  const int64_t m_beg = blockIdx.z * rows_per_split;
  const int64_t m_end = min(M, m_beg + rows_per_split);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float csum[4] = {0.f, 0.f, 0.f, 0.f};
  const bool do_col = colpart != nullptr && blockIdx.x == 0 && ty == 0;

  // loader mapping: 256 threads stage 16 rows x 64 columns of each operand (4 floats per thread)
  const int lr = threadIdx.x / 16;         // staged row 0..15
  const int lc = (threadIdx.x % 16) * 4;   // first of 4 columns
  for (int64_t m0 = m_beg; m0 < m_end; m0 += kStep) {
    const int64_t m = m0 + lr;
    float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
    if (m < m_end) {
      const float* xr = X + m * K1 + k1_0 + lc;
      const float* gr = G + m * K2 + k2_0 + lc;
      if (k1_0 + lc + 3 < K1 && (K1 & 3) == 0) va = __ldg(reinterpret_cast<const float4*>(xr));
      else {
        if (k1_0 + lc + 0 < K1) va.x = __ldg(xr + 0);
        if (k1_0 + lc + 1 < K1) va.y = __ldg(xr + 1);
        if (k1_0 + lc + 2 < K1) va.z = __ldg(xr + 2);
        if (k1_0 + lc + 3 < K1) va.w = __ldg(xr + 3);
      }
      if (k2_0 + lc + 3 < K2 && (K2 & 3) == 0) vb = __ldg(reinterpret_cast<const float4*>(gr));
      else {
        if (k2_0 + lc + 0 < K2) vb.x = __ldg(gr + 0);
        if (k2_0 + lc + 1 < K2) vb.y = __ldg(gr + 1);
        if (k2_0 + lc + 2 < K2) vb.z = __ldg(gr + 2);
        if (k2_0 + lc + 3 < K2) vb.w = __ldg(gr + 3);
      }
    }
    *reinterpret_cast<float4*>(&sa[lr][lc]) = va;
    *reinterpret_cast<float4*>(&sb[lr][lc]) = vb;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kStep; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(&sa[r][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&sb[r][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      if (do_col) {
#pragma unroll
        for (int j = 0; j < 4; ++j) csum[j] += bv[j];
      }
    }
    __syncthreads();
  }
  float* out = partial + static_cast<int64_t>(blockIdx.z) * K1 * K2;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k1 = k1_0 + ty * 4 + i;
    if (k1 >= K1) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k2 = k2_0 + tx * 4 + j;
      if (k2 < K2) out[static_cast<int64_t>(k1) * K2 + k2] = acc[i][j];
    }
  }
  if (do_col) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k2 = k2_0 + tx * 4 + j;
      if (k2 < K2) colpart[static_cast<int64_t>(blockIdx.z) * K2 + k2] = csum[j];
    }
  }
}

// out[i] = sum_{s < S} partial[s, i] in ascending s (fixed order -> deterministic)
__global__ void __launch_bounds__(256)
split_reduce_kernel(const float* __restrict__ partial, int S, int64_t n, float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;   // 4 interleaved chains, combined in a fixed order
  int s = 0;
  for (; s + 3 < S; s += 4) {
    s0 += partial[(s + 0) * n + i];
    s1 += partial[(s + 1) * n + i];
    s2 += partial[(s + 2) * n + i];
    s3 += partial[(s + 3) * n + i];
  }
  for (; s < S; ++s) s0 += partial[s * n + i];
  out[i] = (s0 + s1) + (s2 + s3);
}

inline int pick_splits(int64_t M, int K1, int K2) {
  const int tiles = ceil_div(K1, kTile) * ceil_div(K2, kTile);
  int64_t s = std::max<int64_t>(1, (2 * kNumSM) / tiles);           // ~2 CTAs per SM
  s = std::min<int64_t>(s, std::max<int64_t>(1, ceil_div<int64_t>(M, 4 * kStep)));  // >= 64 rows per split
  return static_cast<int>(s);
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

size_t kgcnn_gemm_tn_workspace_bytes(int64_t M, int K1, int K2) {
  if (M <= 0 || K1 <= 0 || K2 <= 0) return 0;
  const int S = pick_splits(M, K1, K2);
  return static_cast<size_t>(S) * (static_cast<size_t>(K1) * K2 + K2) * sizeof(float) + 256;
}

int kgcnn_gemm_tn(const float* X, const float* G, int64_t M, int K1, int K2, float* dW, float* db /*nullable*/,
                  void* workspace, size_t workspace_bytes, kgcnn_stream_t stream) {
  if (M < 0 || K1 <= 0 || K2 <= 0 || !dW) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (M == 0) {
    if (cudaMemsetAsync(dW, 0, sizeof(float) * K1 * K2, st) != cudaSuccess) return KGCNN_E_CUDA;
    if (db && cudaMemsetAsync(db, 0, sizeof(float) * K2, st) != cudaSuccess) return KGCNN_E_CUDA;
    return KGCNN_OK;
  }
  if (!X || !G) return KGCNN_E_INVALID;
  if ((reinterpret_cast<uintptr_t>(X) & 15) || (reinterpret_cast<uintptr_t>(G) & 15)) return KGCNN_E_UNSUPPORTED;
  if (!workspace || workspace_bytes < kgcnn_gemm_tn_workspace_bytes(M, K1, K2)) return KGCNN_E_WORKSPACE;
  const int S = pick_splits(M, K1, K2);
  const int64_t rows_per_split = ceil_div<int64_t>(ceil_div<int64_t>(M, S), kStep) * kStep;
  float* partial = static_cast<float*>(workspace);
  float* colpart = db ? partial + static_cast<size_t>(S) * K1 * K2 : nullptr;
  dim3 grid(ceil_div(K1, kTile), ceil_div(K2, kTile), S);
  gemm_tn_partial_kernel<<<grid, kThreads, 0, st>>>(X, G, M, K1, K2, rows_per_split, partial, colpart);
  const int64_t n = static_cast<int64_t>(K1) * K2;
  split_reduce_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(n, 256)), 256, 0, st>>>(partial, S, n, dW);
  int launches = 2;
  if (db) {
    split_reduce_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(K2, 256)), 256, 0, st>>>(colpart, S, K2, db);
    ++launches;
  }
  return finish_launch(launches);
}

}  // extern "C"
