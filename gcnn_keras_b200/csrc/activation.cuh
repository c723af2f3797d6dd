// Activation functions of the path and their derivatives up to order 3 (kgcnn/ops/activ.py:5-107 and the keras
// built-ins the named models use).  Shared by the bias/activation kernel family (activation.cu) and the Dense
// epilogue (dense_tc.cu).
#pragma once
#include "common.cuh"

namespace kgcnn {

__device__ __forceinline__ float sigmoidf_acc(float x) {
  // accurate for both signs (no overflow): 1/(1+e^-x) = e^x/(1+e^x)
  if (x >= 0.f) return 1.f / (1.f + expf(-x));
  float e = expf(x);
  return e / (1.f + e);
}

template <int ACT>
__device__ __forceinline__ float act_deriv(float x, int order) {
  if (ACT == KGCNN_ACT_LINEAR) {
    return order == 0 ? x : (order == 1 ? 1.f : 0.f);
  } else if (ACT == KGCNN_ACT_RELU) {
    return order == 0 ? fmaxf(x, 0.f) : (order == 1 ? (x > 0.f ? 1.f : 0.f) : 0.f);
  } else if (ACT == KGCNN_ACT_LEAKY_RELU) {   // kgcnn>leaky_relu, alpha = 0.05 (activ.py:61-80)
    return order == 0 ? (x > 0.f ? x : 0.05f * x) : (order == 1 ? (x > 0.f ? 1.f : 0.05f) : 0.f);
  } else if (ACT == KGCNN_ACT_TANH) {
    float t = tanhf(x);
    float d = 1.f - t * t;
    if (order == 0) return t;
    if (order == 1) return d;
    if (order == 2) return -2.f * t * d;
    return -2.f * d * (1.f - 3.f * t * t);
  } else {
    float s = sigmoidf_acc(x);
    float s1 = s * (1.f - s);
    float s2 = s1 * (1.f - 2.f * s);
    if (ACT == KGCNN_ACT_SHIFTED_SOFTPLUS || ACT == KGCNN_ACT_SOFTPLUS2) {
      // softplus(x) - log 2 (activ.py:15) == relu(x) + log(0.5 exp(-|x|) + 0.5) (activ.py:29)
      if (order == 0) return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))) - 0.69314718055994530942f;
      if (order == 1) return s;
      if (order == 2) return s1;
      return s2;
    }
    if (ACT == KGCNN_ACT_SIGMOID) {
      if (order == 0) return s;
      if (order == 1) return s1;
      if (order == 2) return s2;
      return s2 * (1.f - 2.f * s) - 2.f * s1 * s1;
    }
    // swish x*sigmoid(x) (activ.py:102)
    if (order == 0) return x * s;
    if (order == 1) return s + x * s1;
    if (order == 2) return 2.f * s1 + x * s2;
    float s3 = s2 * (1.f - 2.f * s) - 2.f * s1 * s1;
    return 3.f * s2 + x * s3;
  }
}

// order-0 value with a run-time activation code (uniform branch): the Dense epilogue
__device__ __forceinline__ float act_value(int act, float x) {
  switch (act) {
    case KGCNN_ACT_RELU: return act_deriv<KGCNN_ACT_RELU>(x, 0);
    case KGCNN_ACT_LEAKY_RELU: return act_deriv<KGCNN_ACT_LEAKY_RELU>(x, 0);
    case KGCNN_ACT_TANH: return act_deriv<KGCNN_ACT_TANH>(x, 0);
    case KGCNN_ACT_SHIFTED_SOFTPLUS: return act_deriv<KGCNN_ACT_SHIFTED_SOFTPLUS>(x, 0);
    case KGCNN_ACT_SOFTPLUS2: return act_deriv<KGCNN_ACT_SOFTPLUS2>(x, 0);
    case KGCNN_ACT_SIGMOID: return act_deriv<KGCNN_ACT_SIGMOID>(x, 0);
    case KGCNN_ACT_SWISH: return act_deriv<KGCNN_ACT_SWISH>(x, 0);
    default: return x;
  }
}

}  // namespace kgcnn
