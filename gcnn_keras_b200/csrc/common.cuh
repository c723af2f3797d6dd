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

// Programmatic dependent launch (PDL).  The captured training steps are chains of hundreds of DEPENDENT kernels of
// 2-15 us; with the launch attribute below the next kernel's CTAs are scheduled as soon as every CTA of the current one
// has executed pdl_trigger() (the first instruction of each converted kernel), run their prologue (parameter fetch,
// TMEM allocation, barrier set-up, address arithmetic) under the current kernel's tail and block in pdl_wait() until
// the current grid has completed and its writes are visible.  Rule for every converted kernel: NO global-memory access
// before pdl_wait().  Kernels launched with plain <<<>>> neither wait nor trigger and stay fully serialised on both
// sides, so the two kinds mix freely (a converted kernel launched WITHOUT the attribute returns from pdl_wait()
// at once).  KGCNN_PDL=0 in the environment turns the attribute off.
bool pdl_enabled();
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... Params, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  // Only for grids of at most 16 CTAs per SM: a grid of 1e5 short-lived CTAs (the segment kernels at C5) pays the
  // dependency check in every CTA and has nothing to gain from an early start (measured: C5 layer 9.1 -> 13.3 ms).
  const unsigned long long blocks = static_cast<unsigned long long>(grid.x) * grid.y * grid.z;
  cfg.numAttrs = (pdl_enabled() && blocks <= 16ull * kNumSM) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

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
