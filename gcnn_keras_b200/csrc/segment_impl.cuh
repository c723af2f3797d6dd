// Fused gather * weight -> CSR segment reduce (deterministic, no floating-point atomics).
//   y[i,:] = OP_{p in [rowptr[i], rowptr[i+1])} x[src(p),:] * w[e(p),:]
// replaces tf.gather + LazyMultiply + argsort/gather/segment_*/scatter_nd of
//   kgcnn/layers/gather.py:228, kgcnn/layers/conv/schnet_conv.py:76-78, kgcnn/layers/pooling.py:63-76,
//   kgcnn/layers/pooling.py:152-172 (weighted), :215-218 (per-graph pooling).
//
// Mapping: one warp owns one output row.  The warp is split into G = 32/LPR lane groups; a group owns
// a 128-bit (or 32-bit) column slice per lane and walks every G-th edge of the segment, so a warp keeps
// G*U gathered rows in flight.  Edge ids / source rows for 32 consecutive CSR positions are fetched with
// one coalesced load and broadcast by shuffle.  The in-group order (p ascending) and the cross-group
// combine tree are fixed, so results are bitwise reproducible run to run.
#pragma once
#include "common.cuh"

namespace kgcnn {
namespace seg {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

enum Mode { kSumW0 = 0, kSumW1 = 1, kSumW2 = 2, kMax = 3, kMin = 4 };

template <typename V>
struct Vec;
template <>
struct Vec<float4> {
  static constexpr int kWidth = 4;
  static __device__ __forceinline__ float4 fill(float v) { return make_float4(v, v, v, v); }
  static __device__ __forceinline__ float4 ld_keep(const float4* p) { return __ldg(p); }
  static __device__ __forceinline__ float4 ld_stream(const float4* p) { return ldg_stream(p); }
  static __device__ __forceinline__ void st(float4* p, const float4& v) { *p = v; }
  static __device__ __forceinline__ void fma(float4& a, const float4& x, const float4& w) {
    a.x = fmaf(x.x, w.x, a.x); a.y = fmaf(x.y, w.y, a.y); a.z = fmaf(x.z, w.z, a.z); a.w = fmaf(x.w, w.w, a.w);
  }
  static __device__ __forceinline__ void fmas(float4& a, const float4& x, float w) {
    a.x = fmaf(x.x, w, a.x); a.y = fmaf(x.y, w, a.y); a.z = fmaf(x.z, w, a.z); a.w = fmaf(x.w, w, a.w);
  }
  static __device__ __forceinline__ void add(float4& a, const float4& x) { a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w; }
  static __device__ __forceinline__ void vmax(float4& a, const float4& x) {
    a.x = fmaxf(a.x, x.x); a.y = fmaxf(a.y, x.y); a.z = fmaxf(a.z, x.z); a.w = fmaxf(a.w, x.w);
  }
  static __device__ __forceinline__ void vmin(float4& a, const float4& x) {
    a.x = fminf(a.x, x.x); a.y = fminf(a.y, x.y); a.z = fminf(a.z, x.z); a.w = fminf(a.w, x.w);
  }
  static __device__ __forceinline__ float4 scale(const float4& a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
  static __device__ __forceinline__ float4 shfl_xor(const float4& a, int o) {
    return make_float4(__shfl_xor_sync(0xffffffffu, a.x, o), __shfl_xor_sync(0xffffffffu, a.y, o),
                       __shfl_xor_sync(0xffffffffu, a.z, o), __shfl_xor_sync(0xffffffffu, a.w, o));
  }
};
template <>
struct Vec<float> {
  static constexpr int kWidth = 1;
  static __device__ __forceinline__ float fill(float v) { return v; }
  static __device__ __forceinline__ float ld_keep(const float* p) { return __ldg(p); }
  static __device__ __forceinline__ float ld_stream(const float* p) { return __ldg(p); }
  static __device__ __forceinline__ void st(float* p, const float& v) { *p = v; }
  static __device__ __forceinline__ void fma(float& a, const float& x, const float& w) { a = fmaf(x, w, a); }
  static __device__ __forceinline__ void fmas(float& a, const float& x, float w) { a = fmaf(x, w, a); }
  static __device__ __forceinline__ void add(float& a, const float& x) { a += x; }
  static __device__ __forceinline__ void vmax(float& a, const float& x) { a = fmaxf(a, x); }
  static __device__ __forceinline__ void vmin(float& a, const float& x) { a = fminf(a, x); }
  static __device__ __forceinline__ float scale(const float& a, float s) { return a * s; }
  static __device__ __forceinline__ float shfl_xor(const float& a, int o) { return __shfl_xor_sync(0xffffffffu, a, o); }
};

template <typename V, int LPR, int VPL, int MODE, bool GATHER>
__global__ void __launch_bounds__(kThreads)
gms_kernel(const V* __restrict__ x, const V* __restrict__ wv, const float* __restrict__ ws,
           const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm, const int32_t* __restrict__ gidx,
           int gidx_stride, int gidx_permuted, int64_t n_out, int nvec, int mean, V* __restrict__ y) {
  constexpr int G = 32 / LPR;
  constexpr int U = (VPL >= 3) ? 2 : 4;  // gathered rows in flight per lane group
  using T = Vec<V>;
  const int lane = threadIdx.x & 31;
  const int g = lane / LPR, sub = lane % LPR;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  if (row >= n_out) return;  // warp-uniform
  const int cbase = blockIdx.y * (LPR * VPL) + sub;
  const int beg = __ldg(rowptr + row), end = __ldg(rowptr + row + 1);

  V acc[VPL];
  const float init = (MODE == kMax) ? -INFINITY : (MODE == kMin) ? INFINITY : 0.f;
#pragma unroll
  for (int k = 0; k < VPL; ++k) acc[k] = T::fill(init);

  for (int pb = beg; pb < end; pb += 32) {
    const int p = pb + lane;
    int e_l = 0, s_l = 0;
    if (p < end) {
      e_l = perm ? __ldg(perm + p) : p;
      if (GATHER) s_l = __ldg(gidx + static_cast<int64_t>(gidx_permuted ? p : e_l) * gidx_stride);
      else s_l = e_l;
    }
    const int cnt = min(32, end - pb);
    for (int j0 = 0; j0 < cnt; j0 += G * U) {
      V xv[U][VPL];
      V wvv[U][VPL];
      float wsv[U];
      bool ok[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = j0 + u * G + g;
        ok[u] = j < cnt;
        const int64_t sj = __shfl_sync(0xffffffffu, s_l, j & 31);
        const int64_t ej = __shfl_sync(0xffffffffu, e_l, j & 31);
        if (ok[u]) {
#pragma unroll
          for (int k = 0; k < VPL; ++k) {
            const int c = cbase + k * LPR;
            if (c < nvec) {
              xv[u][k] = GATHER ? T::ld_keep(x + sj * nvec + c) : T::ld_stream(x + sj * nvec + c);
              if (MODE == kSumW2) wvv[u][k] = T::ld_stream(wv + ej * nvec + c);
            }
          }
          if (MODE == kSumW1) wsv[u] = __ldg(ws + ej);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (ok[u]) {
#pragma unroll
          for (int k = 0; k < VPL; ++k) {
            const int c = cbase + k * LPR;
            if (c < nvec) {
              if (MODE == kSumW0) T::add(acc[k], xv[u][k]);
              else if (MODE == kSumW1) T::fmas(acc[k], xv[u][k], wsv[u]);
              else if (MODE == kSumW2) T::fma(acc[k], xv[u][k], wvv[u][k]);
              else if (MODE == kMax) T::vmax(acc[k], xv[u][k]);
              else T::vmin(acc[k], xv[u][k]);
            }
          }
        }
      }
    }
  }
  // fixed combine tree across the G lane groups
#pragma unroll
  for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      V other = T::shfl_xor(acc[k], o);
      if (MODE == kMax) T::vmax(acc[k], other);
      else if (MODE == kMin) T::vmin(acc[k], other);
      else T::add(acc[k], other);
    }
  }
  if (g == 0) {
    const bool empty = end <= beg;
    const float s = (mean && !empty) ? 1.f / static_cast<float>(end - beg) : 1.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int c = cbase + k * LPR;
      if (c < nvec) {
        V out = empty ? T::fill(0.f) : acc[k];
        if (mean) out = T::scale(out, s);
        T::st(y + row * nvec + c, out);
      }
    }
  }
}

template <typename V, int LPR, int VPL>
int launch_shape(int mode, bool gather, const V* x, const V* wv, const float* ws, const int32_t* rowptr,
                 const int32_t* perm, const int32_t* gidx, int gs, int gp, int64_t n_out, int nvec, int mean, V* y,
                 cudaStream_t st) {
  dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(n_out, kWarps)), static_cast<unsigned>(ceil_div(nvec, LPR * VPL)));
#define KGCNN_GMS_LAUNCH(M, GA) \
  gms_kernel<V, LPR, VPL, M, GA><<<grid, kThreads, 0, st>>>(x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y)
  if (mode == kMax) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_GMS_LAUNCH(kMax, false); }
  else if (mode == kMin) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_GMS_LAUNCH(kMin, false); }
  else if (mode == kSumW0) { if (gather) KGCNN_GMS_LAUNCH(kSumW0, true); else KGCNN_GMS_LAUNCH(kSumW0, false); }
  else if (mode == kSumW1) { if (gather) KGCNN_GMS_LAUNCH(kSumW1, true); else KGCNN_GMS_LAUNCH(kSumW1, false); }
  else { if (gather) KGCNN_GMS_LAUNCH(kSumW2, true); else KGCNN_GMS_LAUNCH(kSumW2, false); }
#undef KGCNN_GMS_LAUNCH
  return finish_launch();
}

template <typename V>
int launch(int mode, bool gather, const V* x, const V* wv, const float* ws, const int32_t* rowptr,
           const int32_t* perm, const int32_t* gidx, int gs, int gp, int64_t n_out, int nvec, int mean, V* y,
           cudaStream_t st) {
#define KGCNN_GMS_SHAPE(L, P) \
  return launch_shape<V, L, P>(mode, gather, x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y, st)
  if (nvec <= 1) KGCNN_GMS_SHAPE(1, 1);
  if (nvec <= 2) KGCNN_GMS_SHAPE(2, 1);
  if (nvec <= 4) KGCNN_GMS_SHAPE(4, 1);
  if (nvec <= 8) KGCNN_GMS_SHAPE(8, 1);
  if (nvec <= 16) KGCNN_GMS_SHAPE(16, 1);
  if (nvec <= 32) KGCNN_GMS_SHAPE(32, 1);
  if (nvec <= 64) KGCNN_GMS_SHAPE(32, 2);
  if (nvec <= 96) KGCNN_GMS_SHAPE(32, 3);
  KGCNN_GMS_SHAPE(32, 4);
#undef KGCNN_GMS_SHAPE
}

// defined in segment_v4.cu / segment_v1.cu
int launch_v4(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, cudaStream_t st);
int launch_v1(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, cudaStream_t st);

}  // namespace seg
}  // namespace kgcnn
