// Segment softmax over a CSR relation: kgcnn/ops/segment.py:4-24 as used by
// PoolingLocalEdgesAttention (kgcnn/layers/pooling.py:521-533: stable argsort, permute, segment_max, gather,
// exp, segment_sum, gather, divide = 8 TF ops and 5 passes over the attention tensor).  One launch, one warp
// per segment: max, sum of exp and the normalised coefficients, written back in the ORIGINAL edge order.
// Deterministic: lane-strided partial sums combined by a fixed shuffle tree, no atomics.
#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kThreads = 256;

__global__ void segment_softmax_kernel(const float* __restrict__ a, const int32_t* __restrict__ rowptr,
                                       const int32_t* __restrict__ perm, int64_t N, int H, int normalize,
                                       float* __restrict__ out) {
  int64_t row = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (row >= N) return;
  int lo = rowptr[row], hi = rowptr[row + 1];
  if (lo == hi) return;
  for (int h = 0; h < H; ++h) {
    float mx = 0.f;
    if (normalize) {
      mx = -INFINITY;
      for (int p = lo + lane; p < hi; p += 32) {
        int64_t e = perm ? perm[p] : p;
        mx = fmaxf(mx, a[e * H + h]);
      }
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    float s = 0.f;
    for (int p = lo + lane; p < hi; p += 32) {
      int64_t e = perm ? perm[p] : p;
      s += expf(a[e * H + h] - mx);
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    for (int p = lo + lane; p < hi; p += 32) {
      int64_t e = perm ? perm[p] : p;
      out[e * H + h] = expf(a[e * H + h] - mx) / s;
    }
  }
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" int kgcnn_segment_softmax(const float* a, const int32_t* rowptr, const int32_t* perm, int64_t n_out,
                                     int64_t n_elements, int H, int normalize, float* out, kgcnn_stream_t stream) {
  if (n_out < 0 || n_elements < 0 || H <= 0) return KGCNN_E_INVALID;
  if (n_out == 0 || n_elements == 0) return KGCNN_OK;
  if (!a || !rowptr || !out) return KGCNN_E_INVALID;
  unsigned blocks = static_cast<unsigned>(ceil_div<int64_t>(n_out * 32, kThreads));
  segment_softmax_kernel<<<blocks, kThreads, 0, as_stream(stream)>>>(a, rowptr, perm, n_out, H, normalize, out);
  return finish_launch();
}
