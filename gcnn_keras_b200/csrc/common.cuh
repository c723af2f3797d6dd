// Shared helpers for the sm_100a kernels of libkgcnn_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/kgcnn_b200.h"

namespace kgcnn {

constexpr int kNumSM = 148;  // B200: 2 dies x 74 SMs

void count_launch(int n = 1);

inline int finish_launch(int n = 1) {
  count_launch(n);
  cudaError_t err = cudaPeekAtLastError();
  if (err != cudaSuccess) {
    cudaGetLastError();  // clear so the next call reports its own error
    return KGCNN_E_CUDA;
  }
  return KGCNN_OK;
}

// Function attributes (opt-in dynamic shared memory) are per device: one flag word per call site, one bit per
// device ordinal, set under a lock so a second thread never launches before the attribute is in place.
struct PerDeviceOnce {
  std::mutex mu;
  uint64_t done = 0;
  template <typename Fn>
  void run(Fn&& fn) {
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    std::lock_guard<std::mutex> lock(mu);
    if (done & bit) return;
    fn();
    done |= bit;
  }
};

inline cudaStream_t as_stream(kgcnn_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

template <typename T>
__host__ __device__ inline T ceil_div(T a, T b) {
  return (a + b - 1) / b;
}

// 128-bit streaming load through the read-only path (edge tensors are read exactly once)
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
// 128-bit load that may stay in L1/L2 (gathered node rows are re-used by neighbouring edges)
__device__ __forceinline__ float4 ldg_keep(const float4* p) { return __ldg(p); }

__device__ __forceinline__ void stg_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

}  // namespace kgcnn
