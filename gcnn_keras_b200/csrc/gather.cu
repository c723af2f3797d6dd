// Row gathers: a2/a3 GatherEmbedding(Selection) (kgcnn/layers/gather.py:83,228) and the two-sided
// gather-product that is the VJP of the fused gather*weight->segment-sum with respect to the weight.
// Layout: a group of LPR lanes owns one row and walks it in 128-bit steps, so every warp issues fully
// coalesced LDG.128/STG.128; each group keeps R rows in flight to cover the index->row dependency.
#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kThreads = 256;
constexpr int kRows = 4;  // rows in flight per lane group

template <typename V>
struct VecOps;
template <>
struct VecOps<float4> {
  static __device__ __forceinline__ float4 load_keep(const float4* p) { return __ldg(p); }
  static __device__ __forceinline__ void store(float4* p, const float4& v) { stg_stream(p, v); }
  static __device__ __forceinline__ float4 mul(const float4& a, const float4& b) {
    return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
  }
  static __device__ __forceinline__ float hsum(const float4& a) { return (a.x + a.y) + (a.z + a.w); }
};
template <>
struct VecOps<float> {
  static __device__ __forceinline__ float load_keep(const float* p) { return __ldg(p); }
  static __device__ __forceinline__ void store(float* p, const float& v) { *p = v; }
  static __device__ __forceinline__ float mul(const float& a, const float& b) { return a * b; }
  static __device__ __forceinline__ float hsum(const float& a) { return a; }
};

// y[e, :] = x[idx[e], :]   (nvec vectors of type V per row)
template <typename V, int LPR>
__global__ void __launch_bounds__(kThreads)
gather_rows_kernel(const V* __restrict__ x, const int32_t* __restrict__ idx, int idx_stride, int64_t E, int nvec,
                   V* __restrict__ y, int64_t y_stride_vec) {
  constexpr int kGroups = kThreads / LPR;
  const int gid = threadIdx.x / LPR, sub = threadIdx.x % LPR;
  const int64_t base = static_cast<int64_t>(blockIdx.x) * (kGroups * kRows);
  int32_t src[kRows];
  int64_t row[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    row[r] = base + r * kGroups + gid;
    src[r] = row[r] < E ? __ldg(idx + row[r] * idx_stride) : -1;
  }
  for (int c = sub; c < nvec; c += LPR) {
    V v[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r)
      if (src[r] >= 0) v[r] = VecOps<V>::load_keep(x + static_cast<int64_t>(src[r]) * nvec + c);
#pragma unroll
    for (int r = 0; r < kRows; ++r)
      if (src[r] >= 0) VecOps<V>::store(y + row[r] * y_stride_vec + c, v[r]);
  }
}

template <int LPR>
__global__ void __launch_bounds__(kThreads)
gather_rows_i64_kernel(const int64_t* __restrict__ x, const int32_t* __restrict__ idx, int idx_stride, int64_t E,
                       int F, int64_t* __restrict__ y, int64_t y_stride) {
  constexpr int kGroups = kThreads / LPR;
  const int gid = threadIdx.x / LPR, sub = threadIdx.x % LPR;
  const int64_t e = static_cast<int64_t>(blockIdx.x) * kGroups + gid;
  if (e >= E) return;
  const int64_t s = __ldg(idx + e * idx_stride);
  for (int c = sub; c < F; c += LPR) y[e * y_stride + c] = __ldg(x + s * F + c);
}

// z[e,:] = a[ia[e],:] * b[ib[e],:]      (REDUCE: z[e] = sum_f ...)
template <typename V, int LPR, bool REDUCE>
__global__ void __launch_bounds__(kThreads)
gather2_mul_kernel(const V* __restrict__ a, const int32_t* __restrict__ ia, int ia_stride, const V* __restrict__ b,
                   const int32_t* __restrict__ ib, int ib_stride, int64_t E, int nvec, float* __restrict__ z) {
  pdl_trigger();
  pdl_wait();
  constexpr int kGroups = kThreads / LPR;
  const int gid = threadIdx.x / LPR, sub = threadIdx.x % LPR;
  const int64_t base = static_cast<int64_t>(blockIdx.x) * (kGroups * kRows);
  int64_t ra[kRows], rb[kRows], row[kRows];
  bool ok[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    row[r] = base + r * kGroups + gid;
    ok[r] = row[r] < E;
    ra[r] = ok[r] ? (ia ? static_cast<int64_t>(__ldg(ia + row[r] * ia_stride)) : row[r]) : 0;
    rb[r] = ok[r] ? (ib ? static_cast<int64_t>(__ldg(ib + row[r] * ib_stride)) : row[r]) : 0;
  }
  float acc[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) acc[r] = 0.f;
  for (int c = sub; c < nvec; c += LPR) {
    V va[kRows], vb[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r)
      if (ok[r]) {
        va[r] = VecOps<V>::load_keep(a + ra[r] * nvec + c);
        vb[r] = VecOps<V>::load_keep(b + rb[r] * nvec + c);
      }
#pragma unroll
    for (int r = 0; r < kRows; ++r)
      if (ok[r]) {
        V p = VecOps<V>::mul(va[r], vb[r]);
        if (REDUCE) acc[r] += VecOps<V>::hsum(p);
        else VecOps<V>::store(reinterpret_cast<V*>(z) + row[r] * nvec + c, p);
      }
  }
  if (REDUCE) {
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      float s = acc[r];
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (sub == 0 && ok[r]) z[row[r]] = s;
    }
  }
}

inline int pick_lpr(int nvec) {
  int l = 1;
  while (l < nvec && l < 32) l <<= 1;
  return l;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename V>
int launch_gather(const V* x, const int32_t* idx, int idx_stride, int64_t E, int nvec, V* y, int64_t ys,
                  cudaStream_t st) {
  int lpr = pick_lpr(nvec);
#define KGCNN_GATHER_CASE(L)                                                                         \
  case L: {                                                                                          \
    unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(E, (kThreads / L) * kRows));             \
    gather_rows_kernel<V, L><<<grid, kThreads, 0, st>>>(x, idx, idx_stride, E, nvec, y, ys);         \
    break;                                                                                           \
  }
  switch (lpr) {
    KGCNN_GATHER_CASE(1)
    KGCNN_GATHER_CASE(2)
    KGCNN_GATHER_CASE(4)
    KGCNN_GATHER_CASE(8)
    KGCNN_GATHER_CASE(16)
    KGCNN_GATHER_CASE(32)
  }
#undef KGCNN_GATHER_CASE
  return finish_launch();
}

template <typename V, bool REDUCE>
int launch_gather2(const V* a, const int32_t* ia, int ias, const V* b, const int32_t* ib, int ibs, int64_t E,
                   int nvec, float* z, cudaStream_t st) {
  int lpr = pick_lpr(nvec);
#define KGCNN_G2_CASE(L)                                                                             \
  case L: {                                                                                          \
    unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(E, (kThreads / L) * kRows));             \
    launch_pdl(gather2_mul_kernel<V, L, REDUCE>, dim3(grid), dim3(kThreads), 0, st, a, ia, ias, b, ib, ibs, E, nvec, z); \
    break;                                                                                           \
  }
  switch (lpr) {
    KGCNN_G2_CASE(1)
    KGCNN_G2_CASE(2)
    KGCNN_G2_CASE(4)
    KGCNN_G2_CASE(8)
    KGCNN_G2_CASE(16)
    KGCNN_G2_CASE(32)
  }
#undef KGCNN_G2_CASE
  return finish_launch();
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_gather_rows(const float* x, int64_t N, int F, const int32_t* idx, int idx_stride, int64_t E, float* y,
                      int64_t y_stride, kgcnn_stream_t stream) {
  if (E < 0 || N < 0 || F <= 0 || idx_stride <= 0 || y_stride < F) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!x || !idx || !y) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (F % 4 == 0 && y_stride % 4 == 0 && aligned16(x) && aligned16(y))
    return launch_gather<float4>(reinterpret_cast<const float4*>(x), idx, idx_stride, E, F / 4,
                                 reinterpret_cast<float4*>(y), y_stride / 4, st);
  return launch_gather<float>(x, idx, idx_stride, E, F, y, y_stride, st);
}

int kgcnn_gather_rows_i64(const int64_t* x, int64_t N, int F, const int32_t* idx, int idx_stride, int64_t E,
                          int64_t* y, int64_t y_stride, kgcnn_stream_t stream) {
  if (E < 0 || N < 0 || F <= 0 || idx_stride <= 0 || y_stride < F) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!x || !idx || !y) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  int lpr = pick_lpr(F);
#define KGCNN_GI_CASE(L)                                                                     \
  case L:                                                                                    \
    gather_rows_i64_kernel<L><<<static_cast<unsigned>(ceil_div<int64_t>(E, kThreads / L)),   \
                               kThreads, 0, st>>>(x, idx, idx_stride, E, F, y, y_stride);    \
    break;
  switch (lpr) {
    KGCNN_GI_CASE(1)
    KGCNN_GI_CASE(2)
    KGCNN_GI_CASE(4)
    KGCNN_GI_CASE(8)
    KGCNN_GI_CASE(16)
    KGCNN_GI_CASE(32)
  }
#undef KGCNN_GI_CASE
  return finish_launch();
}

int kgcnn_gather2_mul(const float* a, const int32_t* ia, int ia_stride, const float* b, const int32_t* ib,
                      int ib_stride, int F, int64_t E, int reduce_f, float* z, kgcnn_stream_t stream) {
  if (E < 0 || F <= 0) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!a || !b || !z) return KGCNN_E_INVALID;
  if ((ia && ia_stride <= 0) || (ib && ib_stride <= 0)) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  bool vec = F % 4 == 0 && aligned16(a) && aligned16(b) && (reduce_f || aligned16(z));
  if (vec) {
    auto a4 = reinterpret_cast<const float4*>(a);
    auto b4 = reinterpret_cast<const float4*>(b);
    return reduce_f ? launch_gather2<float4, true>(a4, ia, ia_stride, b4, ib, ib_stride, E, F / 4, z, st)
                    : launch_gather2<float4, false>(a4, ia, ia_stride, b4, ib, ib_stride, E, F / 4, z, st);
  }
  return reduce_f ? launch_gather2<float, true>(a, ia, ia_stride, b, ib, ib_stride, E, F, z, st)
                  : launch_gather2<float, false>(a, ia, ia_stride, b, ib, ib_stride, E, F, z, st);
}

}  // extern "C"
