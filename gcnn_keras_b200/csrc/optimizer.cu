// Adam with Keras' per-variable clipnorm on ONE flat parameter buffer (force_schnet.py:196-203:
// ks.optimizers.Adam() with optimizer.clipnorm = 1.0).  The per-variable squared norms come from the CSR
// segment kernel (rowptr = parameter offsets); this kernel applies  g <- g * clip / max(||g_var||, clip)  and
// the Keras Adam update  p -= lr*sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps)  in one pass.
// The step counter lives on the device so the whole optimizer can sit inside a captured CUDA graph.
#include "common.cuh"

namespace kgcnn {
namespace {

__device__ __forceinline__ bool guard_tripped(const int32_t* __restrict__ guard,
                                              const int32_t* __restrict__ guard_mask, int n_guard) {
  bool bad = false;
  for (int i = 0; i < n_guard; ++i) bad |= (guard[i] & guard_mask[i]) != 0;      // uniform over the grid
  return bad;
}

__global__ void __launch_bounds__(256)
adam_clipnorm_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                     const int32_t* __restrict__ var_of, const float* __restrict__ sumsq, int64_t n, float lr,
                     const float* __restrict__ lr_device, float b1, float b2, float eps, float clipnorm,
                     const int32_t* __restrict__ step,
                     const int32_t* __restrict__ guard, const int32_t* __restrict__ guard_mask, int n_guard) {
  if (guard_tripped(guard, guard_mask, n_guard)) return;
  const int t = *step + 1;
  const float c1 = 1.f - powf(b1, static_cast<float>(t));
  const float c2 = sqrtf(1.f - powf(b2, static_cast<float>(t)));
  if (lr_device) lr = *lr_device;                    // a learning-rate schedule under a captured graph
  const float lr_t = lr * c2 / c1;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float gi = g[i];
    if (clipnorm > 0.f) {
      const float nrm = sqrtf(sumsq[var_of[i]]);
      gi *= clipnorm / fmaxf(nrm, clipnorm);          // tf.clip_by_norm
    }
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

__global__ void bump_step_kernel(int32_t* step, const int32_t* __restrict__ guard,
                                 const int32_t* __restrict__ guard_mask, int n_guard) {
  if (!guard_tripped(guard, guard_mask, n_guard)) *step += 1;
}

__global__ void guard_fold_kernel(const int32_t* __restrict__ guard, const int32_t* __restrict__ guard_mask, int n_guard,
                                  float* __restrict__ word) {
  *word = guard_tripped(guard, guard_mask, n_guard) ? 1.f : 0.f;
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" int kgcnn_adam_clipnorm(float* params, const float* grads, float* m, float* v, const int32_t* var_of,
                                   const float* sumsq, int64_t n, float lr, const float* lr_device, float beta1,
                                   float beta2, float eps, float clipnorm, int32_t* step, const int32_t* guard,
                                   const int32_t* guard_mask, int n_guard, kgcnn_stream_t stream) {
  if (n < 0 || !step || n_guard < 0 || (n_guard > 0 && (!guard || !guard_mask))) return KGCNN_E_INVALID;
  if (!guard) n_guard = 0;
  if (n == 0) return KGCNN_OK;
  if (!params || !grads || !m || !v || (clipnorm > 0.f && (!var_of || !sumsq))) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(n, 256) < kNumSM * 8 ? ceil_div<int64_t>(n, 256)
                                                                                     : kNumSM * 8);
  adam_clipnorm_kernel<<<grid, 256, 0, st>>>(params, grads, m, v, var_of, sumsq, n, lr, lr_device, beta1, beta2, eps,
                                             clipnorm,
                                             step, guard, guard_mask, n_guard);
  bump_step_kernel<<<1, 1, 0, st>>>(step, guard, guard_mask, n_guard);
  return finish_launch(2);
}

extern "C" int kgcnn_guard_fold(const int32_t* guard, const int32_t* guard_mask, int n_guard, float* word,
                                kgcnn_stream_t stream) {
  if (n_guard < 0 || !word || (n_guard > 0 && (!guard || !guard_mask))) return KGCNN_E_INVALID;
  guard_fold_kernel<<<1, 1, 0, as_stream(stream)>>>(guard, guard_mask, n_guard, word);
  return finish_launch();
}
