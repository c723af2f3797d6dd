// Fused gather * weight -> CSR segment reduce (deterministic, no floating-point atomics).
//   y[i,:] = OP_{p in [rowptr[i], rowptr[i+1])} x[src(p),:] * w[e(p),:]
// replaces tf.gather + LazyMultiply + argsort/gather/segment_*/scatter_nd of
//   kgcnn/layers/gather.py:228, kgcnn/layers/conv/schnet_conv.py:76-78, kgcnn/layers/pooling.py:63-76,
//   kgcnn/layers/pooling.py:152-172 (weighted), :215-218 (per-graph pooling).
//
// Mapping: one warp owns one output row.  The warp is split into G = 32/LPR lane groups; a lane owns VPL
// 128-bit (or 32-bit) column slices and the group walks every G-th edge of the segment with U edges in
// flight, so a warp keeps G*U gathered rows in flight.  The in-group order (p ascending) and the
// cross-group combine tree are fixed, so results are bitwise reproducible run to run.
// r1 ncu (profiles/): with 16 lanes/row + shuffle-broadcast indices the kernel was issue-bound (73 % issue
// slots, 15 % DRAM); 8 lanes/row with 2-4 vectors per lane halves the per-edge instruction count.
#pragma once
#include "common.cuh"

namespace kgcnn {
namespace seg {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

enum Mode { kSumW0 = 0, kSumW1 = 1, kSumW2 = 2, kMax = 3, kMin = 4 };

// optional list of hub rows (segments longer than `thresh`): handled by one CTA each in a second pass
struct LongRows {
  const int32_t* rows = nullptr;   // device [cap]
  const int32_t* count = nullptr;  // device scalar
  int cap = 0;
  int thresh = 0;
};

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

template <typename V, int LPR, int VPL, int MODE, bool GATHER, bool EXACT, bool GPR>
__global__ void __launch_bounds__(kThreads)
gms_kernel(const V* __restrict__ x, const V* __restrict__ wv, const float* __restrict__ ws,
           const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm, const int32_t* __restrict__ gidx,
           int gidx_stride, int gidx_permuted, int64_t n_out, int nvec, int mean, V* __restrict__ y,
           int long_thresh) {
  pdl_trigger();
  pdl_wait();
  constexpr int G = 32 / LPR;
  constexpr int U = (VPL >= 3) ? 2 : 4;  // gathered rows in flight per lane group
  using T = Vec<V>;
  const int lane = threadIdx.x & 31;
  const int g = lane / LPR, sub = lane % LPR;
  // GPR (group per row): every lane group owns its own output row and walks it sequentially -- no
  // cross-group shuffle tree; used for short segments (molecular graphs).  Otherwise the whole warp owns
  // one row and the groups interleave its edges.
  constexpr int PS = GPR ? 1 : G;  // stride of this group's walk over CSR positions
  const int64_t wrow = static_cast<int64_t>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  const int64_t row = GPR ? wrow * G + g : wrow;
  const bool valid = row < n_out;
  if (!GPR && !valid) return;  // warp-uniform
  const int cbase = blockIdx.y * (LPR * VPL) + sub;
  int beg = valid ? __ldg(rowptr + row) : 0, end = valid ? __ldg(rowptr + row + 1) : 0;
  // hub rows (segment longer than long_thresh) are left to the block-per-row pass over the long-row list
  const bool is_long = long_thresh > 0 && (end - beg) > long_thresh;
  if (is_long) end = beg;
  bool cok[VPL];
#pragma unroll
  for (int k = 0; k < VPL; ++k) cok[k] = EXACT || (cbase + k * LPR < nvec);

  V acc[VPL];
  const float init = (MODE == kMax) ? -INFINITY : (MODE == kMin) ? INFINITY : 0.f;
#pragma unroll
  for (int k = 0; k < VPL; ++k) acc[k] = T::fill(init);

  // e = original edge id (for the weight row), s = source row of x
  auto resolve = [&](int p, int& e, int& s) {
    e = perm ? __ldg(perm + p) : p;
    if (GATHER) s = __ldg(gidx + static_cast<int64_t>(gidx_permuted ? p : e) * gidx_stride);
    else s = e;
  };
  auto accumulate = [&](V (&xv)[VPL], V (&wr)[VPL], float wsc) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      if (cok[k]) {
        if (MODE == kSumW0) T::add(acc[k], xv[k]);
        else if (MODE == kSumW1) T::fmas(acc[k], xv[k], wsc);
        else if (MODE == kSumW2) T::fma(acc[k], xv[k], wr[k]);
        else if (MODE == kMax) T::vmax(acc[k], xv[k]);
        else T::vmin(acc[k], xv[k]);
      }
    }
  };

  int p = beg + (GPR ? 0 : g);
  // main loop: U edges per lane group in flight, no predicates on the edge index
  for (; p + (U - 1) * PS < end; p += U * PS) {
    int e[U], s[U];
#pragma unroll
    for (int u = 0; u < U; ++u) resolve(p + u * PS, e[u], s[u]);
    V xv[U][VPL];
    V wr[U][VPL];
    float wsc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const V* xr = x + static_cast<int64_t>(s[u]) * nvec + cbase;
      const V* wrow = wv + static_cast<int64_t>(e[u]) * nvec + cbase;
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        if (cok[k]) {
          xv[u][k] = GATHER ? T::ld_keep(xr + k * LPR) : T::ld_stream(xr + k * LPR);
          if (MODE == kSumW2) wr[u][k] = T::ld_stream(wrow + k * LPR);
        }
      }
      if (MODE == kSumW1) wsc[u] = __ldg(ws + e[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) accumulate(xv[u], wr[u], wsc[u]);
  }
  // tail: one edge per lane group
  for (; p < end; p += PS) {
    int e, s;
    resolve(p, e, s);
    V xv[VPL];
    V wr[VPL];
    float wsc = 0.f;
    const V* xr = x + static_cast<int64_t>(s) * nvec + cbase;
    const V* wrow = wv + static_cast<int64_t>(e) * nvec + cbase;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      if (cok[k]) {
        xv[k] = GATHER ? T::ld_keep(xr + k * LPR) : T::ld_stream(xr + k * LPR);
        if (MODE == kSumW2) wr[k] = T::ld_stream(wrow + k * LPR);
      }
    }
    if (MODE == kSumW1) wsc = __ldg(ws + e);
    accumulate(xv, wr, wsc);
  }
  __syncwarp();
  // fixed combine tree across the G lane groups
#pragma unroll
  for (int o = LPR; o < (GPR ? 0 : 32); o <<= 1) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      V other = T::shfl_xor(acc[k], o);
      if (MODE == kMax) T::vmax(acc[k], other);
      else if (MODE == kMin) T::vmin(acc[k], other);
      else T::add(acc[k], other);
    }
  }
  if ((GPR ? valid : (g == 0)) && !is_long) {
    const bool empty = end <= beg;
    const float sc = (mean && !empty) ? 1.f / static_cast<float>(end - beg) : 1.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      if (cok[k]) {
        V out = empty ? T::fill(0.f) : acc[k];
        if (mean) out = T::scale(out, sc);
        T::st(y + row * nvec + cbase + k * LPR, out);
      }
    }
  }
}

// Block-per-row mapping for outputs with FEW rows (n_out <= 1024: embedding-table gradients, per-type
// reductions): segment lengths are then unbounded (thousands of atoms of one element), so one CTA = 32 lane
// groups walks a row together and the groups are combined through shared memory in a fixed order.
template <typename V, int MODE, bool GATHER, int LPR>
__global__ void __launch_bounds__(kThreads)
gms_block_kernel(const V* __restrict__ x, const V* __restrict__ wv, const float* __restrict__ ws,
                 const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
                 const int32_t* __restrict__ gidx, int gidx_stride, int gidx_permuted, int64_t n_out, int nvec,
                 int mean, V* __restrict__ y, const int32_t* __restrict__ row_map,
                 const int32_t* __restrict__ n_rows_dev) {
  pdl_trigger();
  pdl_wait();
  constexpr int NG = kThreads / LPR;  // lane groups walking the row together (32 .. 256)
  using T = Vec<V>;
  __shared__ V red[NG][LPR];
  const int g = threadIdx.x / LPR, sub = threadIdx.x % LPR;
  if (n_rows_dev && static_cast<int>(blockIdx.x) >= __ldg(n_rows_dev)) return;   // block-uniform
  const int64_t row = row_map ? __ldg(row_map + blockIdx.x) : blockIdx.x;
  const int beg = __ldg(rowptr + row), end = __ldg(rowptr + row + 1);
  const float init = (MODE == kMax) ? -INFINITY : (MODE == kMin) ? INFINITY : 0.f;
  // column slabs of LPR vectors are spread over grid.y: few, very long rows (3 elements own all atoms of an
  // ethanol batch) would otherwise walk the slabs one after the other in a single CTA
  for (int cb = blockIdx.y * LPR; cb < nvec; cb += LPR * gridDim.y) {
    const int c = cb + sub;
    const bool cok = c < nvec;
    V acc = T::fill(init);
    for (int p = beg + g; p < end; p += NG) {
      const int e = perm ? __ldg(perm + p) : p;
      const int s = GATHER ? __ldg(gidx + static_cast<int64_t>(gidx_permuted ? p : e) * gidx_stride) : e;
      if (cok) {
        V xv = GATHER ? T::ld_keep(x + static_cast<int64_t>(s) * nvec + c)
                      : T::ld_stream(x + static_cast<int64_t>(s) * nvec + c);
        if (MODE == kSumW0) T::add(acc, xv);
        else if (MODE == kSumW1) T::fmas(acc, xv, __ldg(ws + e));
        else if (MODE == kSumW2) T::fma(acc, xv, T::ld_stream(wv + static_cast<int64_t>(e) * nvec + c));
        else if (MODE == kMax) T::vmax(acc, xv);
        else T::vmin(acc, xv);
      }
    }
    red[g][sub] = acc;
    __syncthreads();
    if (g == 0 && cok) {
      V tot = red[0][sub];
      for (int k = 1; k < NG; ++k) {   // fixed order over the lane groups
        if (MODE == kMax) T::vmax(tot, red[k][sub]);
        else if (MODE == kMin) T::vmin(tot, red[k][sub]);
        else T::add(tot, red[k][sub]);
      }
      const bool empty = end <= beg;
      if (empty) tot = T::fill(0.f);
      else if (mean) tot = T::scale(tot, 1.f / static_cast<float>(end - beg));
      T::st(y + row * nvec + c, tot);
    }
    __syncthreads();
  }
}

template <typename V>
int launch_block(int mode, bool gather, const V* x, const V* wv, const float* ws, const int32_t* rowptr,
                 const int32_t* perm, const int32_t* gidx, int gs, int gp, int64_t n_out, int nvec, int mean, V* y,
                 const int32_t* row_map, const int32_t* n_rows_dev, int grid_rows, cudaStream_t st) {
  const unsigned grid_x = static_cast<unsigned>(row_map ? grid_rows : n_out);
#define KGCNN_BLK_L(M, GA, L)                                                                                   \
  launch_pdl(gms_block_kernel<V, M, GA, L>, dim3(grid_x, static_cast<unsigned>(ceil_div(nvec, L))), dim3(kThreads), 0, st, x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, \
                                                          y, row_map, n_rows_dev)
#define KGCNN_BLK(M, GA)                                                                                        \
  do {                                                                                                          \
    if (nvec <= 1) KGCNN_BLK_L(M, GA, 1);                                                                       \
    else if (nvec <= 2) KGCNN_BLK_L(M, GA, 2);                                                                  \
    else if (nvec <= 4) KGCNN_BLK_L(M, GA, 4);                                                                  \
    else KGCNN_BLK_L(M, GA, 8);                                                                                 \
  } while (0)
  if (mode == kMax) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_BLK(kMax, false); }
  else if (mode == kMin) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_BLK(kMin, false); }
  else if (mode == kSumW0) { if (gather) KGCNN_BLK(kSumW0, true); else KGCNN_BLK(kSumW0, false); }
  else if (mode == kSumW1) { if (gather) KGCNN_BLK(kSumW1, true); else KGCNN_BLK(kSumW1, false); }
  else { if (gather) KGCNN_BLK(kSumW2, true); else KGCNN_BLK(kSumW2, false); }
#undef KGCNN_BLK
#undef KGCNN_BLK_L
  return finish_launch();
}

template <typename V, int LPR, int VPL>
int launch_shape(int mode, bool gather, const V* x, const V* wv, const float* ws, const int32_t* rowptr,
                 const int32_t* perm, const int32_t* gidx, int gs, int gp, int64_t n_out, int nvec, int mean, V* y,
                 bool short_rows, int long_thresh, cudaStream_t st) {
  constexpr bool kHasGpr = (LPR >= 4 && LPR < 32);
  const bool gpr = kHasGpr && short_rows;
  const int64_t rows_per_block = gpr ? kWarps * (32 / LPR) : kWarps;
  dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(n_out, rows_per_block)),
            static_cast<unsigned>(ceil_div(nvec, LPR * VPL)));
  const bool exact = (nvec == LPR * VPL);
#define KGCNN_GMS_ARGS x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y, long_thresh
#define KGCNN_GMS_LAUNCH(M, GA)                                                                      \
  do {                                                                                               \
    if constexpr (kHasGpr) {                                                                         \
      if (gpr && exact) launch_pdl(gms_kernel<V, LPR, VPL, M, GA, true, true>, grid, dim3(kThreads), 0, st, KGCNN_GMS_ARGS);        \
      else if (gpr) launch_pdl(gms_kernel<V, LPR, VPL, M, GA, false, true>, grid, dim3(kThreads), 0, st, KGCNN_GMS_ARGS);           \
    }                                                                                                \
    if (!gpr && exact) launch_pdl(gms_kernel<V, LPR, VPL, M, GA, true, false>, grid, dim3(kThreads), 0, st, KGCNN_GMS_ARGS);        \
    else if (!gpr) launch_pdl(gms_kernel<V, LPR, VPL, M, GA, false, false>, grid, dim3(kThreads), 0, st, KGCNN_GMS_ARGS);           \
  } while (0)
  if (mode == kMax) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_GMS_LAUNCH(kMax, false); }
  else if (mode == kMin) { if (gather) return KGCNN_E_UNSUPPORTED; KGCNN_GMS_LAUNCH(kMin, false); }
  else if (mode == kSumW0) { if (gather) KGCNN_GMS_LAUNCH(kSumW0, true); else KGCNN_GMS_LAUNCH(kSumW0, false); }
  else if (mode == kSumW1) { if (gather) KGCNN_GMS_LAUNCH(kSumW1, true); else KGCNN_GMS_LAUNCH(kSumW1, false); }
  else { if (gather) KGCNN_GMS_LAUNCH(kSumW2, true); else KGCNN_GMS_LAUNCH(kSumW2, false); }
#undef KGCNN_GMS_LAUNCH
#undef KGCNN_GMS_ARGS
  return finish_launch();
}

template <typename V>
int launch(int mode, bool gather, const V* x, const V* wv, const float* ws, const int32_t* rowptr,
           const int32_t* perm, const int32_t* gidx, int gs, int gp, int64_t n_out, int nvec, int mean, V* y,
           bool short_rows, const LongRows& lr, cudaStream_t st) {
  if (n_out <= 1024)
    return launch_block<V>(mode, gather, x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y, nullptr,
                           nullptr, 0, st);
  const int long_thresh = (lr.rows && lr.cap > 0) ? lr.thresh : 0;
  if (long_thresh > 0) {
    int rc = launch_block<V>(mode, gather, x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y, lr.rows,
                             lr.count, lr.cap, st);
    if (rc != KGCNN_OK) return rc;
  }
#define KGCNN_GMS_SHAPE(L, P)                                                                                    \
  return launch_shape<V, L, P>(mode, gather, x, wv, ws, rowptr, perm, gidx, gs, gp, n_out, nvec, mean, y, short_rows, \
                               long_thresh, st)
  // few lanes per row -> many edges per warp instruction (the kernel is issue-bound, not DRAM-bound, when
  // gathered rows hit in L1/L2); 8 lanes x 128 bit = one full 128 B line per request.
  // Small outputs (SchNet C2: 2304 rows x 128 features = 72 CTAs with 4 vectors per lane) leave half of the 148 SMs
  // idle and serialise 4 x degree dependent loads per lane: split the feature axis over grid.y instead (1 or 2
  // vectors per lane) until the grid covers the machine about twice.
  if (nvec > 8 && nvec <= 32) {
    const int64_t row_blocks = ceil_div<int64_t>(n_out, short_rows ? kWarps * 4 : kWarps);
    if (row_blocks * ceil_div(nvec, 32) < 2 * kNumSM) {
      if (row_blocks * ceil_div(nvec, 16) >= 2 * kNumSM) KGCNN_GMS_SHAPE(8, 2);
      KGCNN_GMS_SHAPE(8, 1);
    }
  }
  if (nvec <= 1) KGCNN_GMS_SHAPE(1, 1);
  if (nvec <= 2) KGCNN_GMS_SHAPE(2, 1);
  if (nvec <= 4) KGCNN_GMS_SHAPE(4, 1);
  if (nvec <= 8) KGCNN_GMS_SHAPE(8, 1);
  if (nvec <= 16) KGCNN_GMS_SHAPE(8, 2);
  if (nvec <= 24) KGCNN_GMS_SHAPE(8, 3);
  if (nvec <= 32) KGCNN_GMS_SHAPE(8, 4);
  if (nvec <= 48) KGCNN_GMS_SHAPE(16, 3);
  if (nvec <= 64) KGCNN_GMS_SHAPE(16, 4);
  if (nvec <= 96) KGCNN_GMS_SHAPE(32, 3);
  KGCNN_GMS_SHAPE(32, 4);
#undef KGCNN_GMS_SHAPE
}

// defined in segment_v4.cu / segment_v1.cu
int launch_v4(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, bool short_rows,
              const LongRows& lr, cudaStream_t st);
int launch_v1(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, bool short_rows,
              const LongRows& lr, cudaStream_t st);

}  // namespace seg
}  // namespace kgcnn
