// Fused bias + activation (and its derivatives) on the flat values of a ragged tensor:
//     out = g1 * g2 * f^(order)(pre + bias)         order in {0,1,2,3}
// f from kgcnn/ops/activ.py:5-107 (shifted_softplus, softplus2, swish) and the keras built-ins the named
// models use (tanh, relu, sigmoid, linear).  One kernel family is closed under differentiation: the VJP of
// order k is order k+1 with one more multiplier, which is what the force-training double backward needs
// (kgcnn/model/force.py:159-176 differentiates dE/dx once more with respect to the weights).
// Replaces BiasAdd + activation + the Lazy* multiplies around it (kgcnn/layers/modules.py:85).
#include <algorithm>

#include "activation.cuh"
#include "common.cuh"

namespace kgcnn {
namespace {

template <int ACT, bool VEC>
__global__ void __launch_bounds__(256)
bias_act_kernel(const float* __restrict__ pre, const float* __restrict__ bias, const float* __restrict__ g1,
                const float* __restrict__ g2, int order, int64_t total, int U, float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  if (VEC) {
    const int64_t nv = total >> 2;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nv; i += stride) {
      float4 p = reinterpret_cast<const float4*>(pre)[i];
      if (bias) {
        const int c = static_cast<int>((i << 2) % U);
        float4 b = *reinterpret_cast<const float4*>(bias + c);
        p.x += b.x; p.y += b.y; p.z += b.z; p.w += b.w;
      }
      float4 r = make_float4(act_deriv<ACT>(p.x, order), act_deriv<ACT>(p.y, order), act_deriv<ACT>(p.z, order),
                             act_deriv<ACT>(p.w, order));
      if (g1) { float4 m = reinterpret_cast<const float4*>(g1)[i]; r.x *= m.x; r.y *= m.y; r.z *= m.z; r.w *= m.w; }
      if (g2) { float4 m = reinterpret_cast<const float4*>(g2)[i]; r.x *= m.x; r.y *= m.y; r.z *= m.z; r.w *= m.w; }
      reinterpret_cast<float4*>(out)[i] = r;
    }
  } else {
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total; i += stride) {
      float p = pre[i];
      if (bias) p += bias[i % U];
      float r = act_deriv<ACT>(p, order);
      if (g1) r *= g1[i];
      if (g2) r *= g2[i];
      out[i] = r;
    }
  }
}

template <int ACT>
int launch_act(const float* pre, const float* bias, const float* g1, const float* g2, int order, int64_t total,
               int U, float* out, cudaStream_t st) {
  auto al = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = (U % 4 == 0) && al(pre) && al(bias) && al(g1) && al(g2) && al(out);
  const int64_t work = vec ? (total >> 2) : total;
  const unsigned grid = static_cast<unsigned>(std::min<int64_t>(ceil_div<int64_t>(work, 256), kNumSM * 16));
  if (vec) launch_pdl(bias_act_kernel<ACT, true>, dim3(grid), dim3(256), 0, st, pre, bias, g1, g2, order, total, U, out);
  else launch_pdl(bias_act_kernel<ACT, false>, dim3(grid), dim3(256), 0, st, pre, bias, g1, g2, order, total, U, out);
  return finish_launch();
}

// The two VJPs of an order-k node out = g1 * f^(k)(pre + bias) for the cotangent gy, in one pass over (pre, gy, g1):
//     d_pre = gy * g1 * f^(k+1)(pre + bias),   d_g1 = gy * f^(k)(pre + bias)
// (the last sweep of a force-training step evaluates both for every activated Dense).
template <int ACT, bool VEC>
__global__ void __launch_bounds__(256)
bias_act_pair_kernel(const float* __restrict__ pre, const float* __restrict__ bias, const float* __restrict__ gy,
                     const float* __restrict__ g1, int order, int64_t total, int U, float* __restrict__ d_pre,
                     float* __restrict__ d_g1) {
  pdl_trigger();
  pdl_wait();
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  if (VEC) {
    const int64_t nv = total >> 2;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nv; i += stride) {
      float4 p = reinterpret_cast<const float4*>(pre)[i];
      if (bias) {
        const float4 b = *reinterpret_cast<const float4*>(bias + static_cast<int>((i << 2) % U));
        p.x += b.x; p.y += b.y; p.z += b.z; p.w += b.w;
      }
      const float4 y = reinterpret_cast<const float4*>(gy)[i];
      const float4 m = reinterpret_cast<const float4*>(g1)[i];
      reinterpret_cast<float4*>(d_pre)[i] =
          make_float4(y.x * m.x * act_deriv<ACT>(p.x, order + 1), y.y * m.y * act_deriv<ACT>(p.y, order + 1),
                      y.z * m.z * act_deriv<ACT>(p.z, order + 1), y.w * m.w * act_deriv<ACT>(p.w, order + 1));
      reinterpret_cast<float4*>(d_g1)[i] =
          make_float4(y.x * act_deriv<ACT>(p.x, order), y.y * act_deriv<ACT>(p.y, order),
                      y.z * act_deriv<ACT>(p.z, order), y.w * act_deriv<ACT>(p.w, order));
    }
  } else {
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total; i += stride) {
      float p = pre[i];
      if (bias) p += bias[i % U];
      const float y = gy[i];
      d_pre[i] = y * g1[i] * act_deriv<ACT>(p, order + 1);
      d_g1[i] = y * act_deriv<ACT>(p, order);
    }
  }
}

template <int ACT>
int launch_act_pair(const float* pre, const float* bias, const float* gy, const float* g1, int order, int64_t total,
                    int U, float* d_pre, float* d_g1, cudaStream_t st) {
  auto al = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = (U % 4 == 0) && al(pre) && al(bias) && al(gy) && al(g1) && al(d_pre) && al(d_g1);
  const int64_t work = vec ? (total >> 2) : total;
  const unsigned grid = static_cast<unsigned>(std::min<int64_t>(ceil_div<int64_t>(work, 256), kNumSM * 16));
  if (vec) launch_pdl(bias_act_pair_kernel<ACT, true>, dim3(grid), dim3(256), 0, st, pre, bias, gy, g1, order, total, U, d_pre, d_g1);
  else launch_pdl(bias_act_pair_kernel<ACT, false>, dim3(grid), dim3(256), 0, st, pre, bias, gy, g1, order, total, U, d_pre, d_g1);
  return finish_launch();
}

// deterministic column sum: out[u] = sum_m a[m,u]   (bias gradients).  Two fixed-shape stages.
constexpr int kColTile = 32, kRowThreads = 8, kRowsPerBlock = 512;

__global__ void __launch_bounds__(kColTile * kRowThreads)
colsum_kernel(const float* __restrict__ a, int64_t M, int U, int64_t rows_per_block, float* __restrict__ out) {
  __shared__ float sm[kRowThreads][kColTile + 1];
  const int tx = threadIdx.x % kColTile, ty = threadIdx.x / kColTile;
  const int col = blockIdx.x * kColTile + tx;
  const int64_t r0 = blockIdx.y * rows_per_block;
  const int64_t r1 = min(M, r0 + rows_per_block);
  float acc = 0.f;
  if (col < U)
    for (int64_t r = r0 + ty; r < r1; r += kRowThreads) acc += a[r * U + col];
  sm[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && col < U) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kRowThreads; ++k) s += sm[k][tx];
    out[static_cast<int64_t>(blockIdx.y) * U + col] = s;
  }
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_bias_act(const float* pre, const float* bias, const float* g1, const float* g2, int act, int order,
                   int64_t M, int U, float* out, kgcnn_stream_t stream) {
  if (M < 0 || U <= 0 || order < 0 || order > 3) return KGCNN_E_INVALID;
  if (M == 0) return KGCNN_OK;
  if (!pre || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  const int64_t total = M * U;
  switch (act) {
    case KGCNN_ACT_LINEAR: return launch_act<KGCNN_ACT_LINEAR>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_SHIFTED_SOFTPLUS:
      return launch_act<KGCNN_ACT_SHIFTED_SOFTPLUS>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_SOFTPLUS2: return launch_act<KGCNN_ACT_SOFTPLUS2>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_SWISH: return launch_act<KGCNN_ACT_SWISH>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_TANH: return launch_act<KGCNN_ACT_TANH>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_RELU: return launch_act<KGCNN_ACT_RELU>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_SIGMOID: return launch_act<KGCNN_ACT_SIGMOID>(pre, bias, g1, g2, order, total, U, out, st);
    case KGCNN_ACT_LEAKY_RELU:
      return launch_act<KGCNN_ACT_LEAKY_RELU>(pre, bias, g1, g2, order, total, U, out, st);
    default: return KGCNN_E_INVALID;
  }
}

int kgcnn_bias_act_pair(const float* pre, const float* bias, const float* gy, const float* g1, int act, int order,
                        int64_t M, int U, float* d_pre, float* d_g1, kgcnn_stream_t stream) {
  if (M < 0 || U <= 0 || order < 0 || order > 2) return KGCNN_E_INVALID;
  if (M == 0) return KGCNN_OK;
  if (!pre || !gy || !g1 || !d_pre || !d_g1) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  const int64_t total = M * U;
#define KGCNN_PAIR(A) case A: return launch_act_pair<A>(pre, bias, gy, g1, order, total, U, d_pre, d_g1, st);
  switch (act) {
    KGCNN_PAIR(KGCNN_ACT_LINEAR)
    KGCNN_PAIR(KGCNN_ACT_SHIFTED_SOFTPLUS)
    KGCNN_PAIR(KGCNN_ACT_SOFTPLUS2)
    KGCNN_PAIR(KGCNN_ACT_SWISH)
    KGCNN_PAIR(KGCNN_ACT_TANH)
    KGCNN_PAIR(KGCNN_ACT_RELU)
    KGCNN_PAIR(KGCNN_ACT_SIGMOID)
    KGCNN_PAIR(KGCNN_ACT_LEAKY_RELU)
    default: return KGCNN_E_INVALID;
  }
#undef KGCNN_PAIR
}

size_t kgcnn_colsum_workspace_bytes(int64_t M, int U) {
  if (M <= 0 || U <= 0) return 0;
  return static_cast<size_t>(ceil_div<int64_t>(M, kRowsPerBlock)) * U * sizeof(float);
}

int kgcnn_colsum(const float* a, int64_t M, int U, float* out, void* workspace, size_t workspace_bytes,
                 kgcnn_stream_t stream) {
  if (M < 0 || U <= 0 || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (M == 0) return cudaMemsetAsync(out, 0, sizeof(float) * U, st) == cudaSuccess ? KGCNN_OK : KGCNN_E_CUDA;
  if (!a) return KGCNN_E_INVALID;
  const int64_t nblk = ceil_div<int64_t>(M, kRowsPerBlock);
  dim3 block(kColTile * kRowThreads);
  if (nblk == 1) {
    colsum_kernel<<<dim3(ceil_div(U, kColTile), 1), block, 0, st>>>(a, M, U, kRowsPerBlock, out);
    return finish_launch();
  }
  if (!workspace || workspace_bytes < kgcnn_colsum_workspace_bytes(M, U)) return KGCNN_E_WORKSPACE;
  float* partial = static_cast<float*>(workspace);
  colsum_kernel<<<dim3(ceil_div(U, kColTile), static_cast<unsigned>(nblk)), block, 0, st>>>(a, M, U, kRowsPerBlock,
                                                                                             partial);
  colsum_kernel<<<dim3(ceil_div(U, kColTile), 1), block, 0, st>>>(partial, nblk, U, nblk, out);
  return finish_launch(2);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// kgcnn_colsum_static: the column sum for hosts that have no way to hand over a workspace.
//
// kgcnn_colsum needs `ceil(M / 512) * U` floats of scratch between its two stages and expects the caller to bring
// them.  Some callers cannot: a TensorFlow op whose runtime has already reserved nearly all of HBM for its own arena
// (a raw cudaMalloc from inside the op then reports out-of-memory), a plain-C host that was given a fixed device
// budget at start-up, or a serving thread that must not run cudaMalloc / cudaFree because both synchronise the
// whole device.  For them the scratch below is part of the library image: KGCNN_COLSUM_STATIC_MAX_FLOATS floats
// (1 MiB) that every GPU which loads libkgcnn_b200.so keeps until the process exits.
//
// g_colsum_static_partial is a DEVICE symbol.  Host code that spells its name gets the address of a host-side shadow
// object, not device memory, so the launcher never passes the name to a kernel: it asks the runtime once per device
// (with that device current) for the real address and passes that.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_colsum_static_partial[KGCNN_COLSUM_STATIC_MAX_FLOATS];

namespace kgcnn {
namespace {

constexpr int kMaxStaticDevices = 64;
std::mutex g_colsum_static_mu;
float* g_colsum_static_addr[kMaxStaticDevices] = {};  // device address of the array, per device ordinal

// Device address of g_colsum_static_partial on the CURRENT device; resolved by the runtime on the first call per
// device (this may load the module on that device, which is why that call belongs before any stream capture).
int colsum_static_scratch(float** ptr) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return KGCNN_E_CUDA; }
  if (dev < 0 || dev >= kMaxStaticDevices) return KGCNN_E_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_colsum_static_mu);
  if (!g_colsum_static_addr[dev]) {
    float* p = nullptr;
    if (cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_partial) != cudaSuccess || !p) {
      cudaGetLastError();
      return KGCNN_E_CUDA;
    }
    g_colsum_static_addr[dev] = p;
  }
  *ptr = g_colsum_static_addr[dev];
  return KGCNN_OK;
}

}  // namespace
}  // namespace kgcnn

extern "C" int kgcnn_colsum_static(const float* a, int64_t M, int U, float* out, kgcnn_stream_t stream) {
  if (M < 0 || U <= 0 || !out) return KGCNN_E_INVALID;
  if (M > 0 && !a) return KGCNN_E_INVALID;
  const int64_t nblk = ceil_div<int64_t>(M, kRowsPerBlock);
  // the limit is checked before anything touches the device: a too-large problem leaves no trace
  if (nblk > 1 && nblk * U > static_cast<int64_t>(KGCNN_COLSUM_STATIC_MAX_FLOATS)) return KGCNN_E_STATIC_LIMIT;
  float* partial = nullptr;
  const int rc = colsum_static_scratch(&partial);  // every call, so the first call on a device always does the set-up
  if (rc != KGCNN_OK) return rc;
  cudaStream_t st = as_stream(stream);
  if (M == 0) return cudaMemsetAsync(out, 0, sizeof(float) * U, st) == cudaSuccess ? KGCNN_OK : KGCNN_E_CUDA;
  dim3 block(kColTile * kRowThreads);
  if (nblk == 1) {
    colsum_kernel<<<dim3(ceil_div(U, kColTile), 1), block, 0, st>>>(a, M, U, kRowsPerBlock, out);
    return finish_launch();
  }
  // Stage 1 writes partial[0 .. nblk*U) completely before stage 2 (same stream) reads exactly that range, so what an
  // earlier call left in the array is never seen and nothing has to be cleared.
  colsum_kernel<<<dim3(ceil_div(U, kColTile), static_cast<unsigned>(nblk)), block, 0, st>>>(a, M, U, kRowsPerBlock,
                                                                                             partial);
  colsum_kernel<<<dim3(ceil_div(U, kColTile), 1), block, 0, st>>>(partial, nblk, U, nblk, out);
  return finish_launch(2);
}
