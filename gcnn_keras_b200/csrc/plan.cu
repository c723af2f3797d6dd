// Per-batch graph plan: index shift (sample -> batch indexing), sortedness check, CSR row pointer and
// stable permutation.  Replaces the bookkeeping the reference redoes in every layer call:
//   kgcnn/ops/partition.py:97-161   (partition_row_indexing)
//   kgcnn/layers/pooling.py:63-68   (argsort(stable) + permutation gathers)
// Integer work only; results are bit-exact against the oracle.
#include <cstdlib>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <atomic>

#include "common.cuh"

namespace kgcnn {

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("KGCNN_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

namespace {

// last g with splits[g] <= e   (splits ascending, splits[0] == 0, repeated values = empty rows)
__device__ __forceinline__ int row_of(const int64_t* __restrict__ splits, int64_t nrows, int64_t e) {
  int64_t lo = 0, hi = nrows;  // answer in [lo, hi)
  while (hi - lo > 1) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(splits + mid) <= e) lo = mid; else hi = mid;
  }
  return static_cast<int>(lo);
}

__global__ void shift_index_kernel(const int64_t* __restrict__ idx, int64_t E, int K,
                                   const int64_t* __restrict__ node_splits,
                                   const int64_t* __restrict__ edge_splits, int64_t B,
                                   int64_t* __restrict__ out64, int32_t* __restrict__ out32) {
  int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (e >= E) return;
  int g = row_of(edge_splits, B, e);
  int64_t shift = __ldg(node_splits + g);
  for (int k = 0; k < K; ++k) {
    int64_t v = idx[e * K + k] + shift;
    if (out64) out64[e * K + k] = v;
    if (out32) out32[static_cast<int64_t>(k) * E + e] = static_cast<int32_t>(v);  // column-major [K,E]
  }
}

__global__ void value_rowids_kernel(const int64_t* __restrict__ splits, int64_t B, int64_t n,
                                    int32_t* __restrict__ ids) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  ids[i] = row_of(splits, B, i);
}

__global__ void cast_i64_i32_kernel(const int64_t* __restrict__ src, int64_t n, int32_t* __restrict__ dst) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i < n) dst[i] = static_cast<int32_t>(src[i]);
}

// `fixed` (nullable): a copy of the keys clamped into [0, N) for the consumers, so that a batch that violates
// the range contract can never make a later kernel of an already enqueued (captured) step read out of bounds;
// the flags still report the violation and the optimizer skips the update (kgcnn_adam_clipnorm's guard).
__global__ void check_sorted_kernel(const int32_t* __restrict__ key, int stride, int64_t E, int64_t N,
                                    int32_t* __restrict__ flags, int32_t* __restrict__ fixed) {
  int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (e >= E) return;
  int32_t k = key[e * stride];
  if (k < 0 || k >= N) flags[1] = 1;
  if (e > 0 && key[(e - 1) * stride] > k) flags[0] = 1;
  if (fixed && (k < 0 || k >= N)) fixed[e * stride] = k < 0 ? 0 : static_cast<int32_t>(N - 1);
}

// Sorted keys: row pointer from segment boundaries, no atomics, no scan.
__global__ void rowptr_from_sorted_kernel(const int32_t* __restrict__ key, int stride, int64_t E, int64_t N,
                                          int32_t* __restrict__ rowptr) {
  int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (e > E) return;
  if (E == 0) {  // single thread e == 0
    for (int64_t j = 0; j <= N; ++j) rowptr[j] = 0;
    return;
  }
  if (e == E) {  // tail: rows after the last key
    int64_t klast = min(max(static_cast<int64_t>(key[(E - 1) * stride]), int64_t(-1)), N - 1);
    for (int64_t j = klast + 1; j <= N; ++j) rowptr[j] = static_cast<int32_t>(E);
    return;
  }
  int64_t k = min(max(static_cast<int64_t>(key[e * stride]), int64_t(0)), N - 1);
  int64_t kprev = (e == 0) ? -1 : min(max(static_cast<int64_t>(key[(e - 1) * stride]), int64_t(0)), N - 1);
  for (int64_t j = kprev + 1; j <= k; ++j) rowptr[j] = static_cast<int32_t>(e);
}

// unsorted keys: dense copy of the (strided) key column + identity payload for the stable radix sort
__global__ void key_iota_kernel(const int32_t* __restrict__ key, int stride, int64_t E,
                                int32_t* __restrict__ key_dense, int32_t* __restrict__ iota) {
  int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (e >= E) return;
  key_dense[e] = key[e * stride];
  iota[e] = static_cast<int32_t>(e);
}

__global__ void iota_kernel(int32_t* __restrict__ out, int64_t n) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i < n) out[i] = static_cast<int32_t>(i);
}

__global__ void permute_i32_kernel(const int32_t* __restrict__ src, int stride, const int32_t* __restrict__ perm,
                                   int64_t E, int32_t* __restrict__ out) {
  int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (p < E) out[p] = src[static_cast<int64_t>(perm[p]) * stride];
}

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

int key_bits(int64_t N) {
  int bits = 1;
  while ((int64_t(1) << bits) < N && bits < 31) ++bits;
  return bits;
}

size_t cub_temp_bytes(int64_t E, int64_t N) {
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const int32_t*)nullptr, (int32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, static_cast<int>(E), 0,
                                  key_bits(N));
  cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (int32_t*)nullptr, (int32_t*)nullptr,
                                static_cast<int>(N + 1));
  return sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
}

constexpr int kThreads = 256;
inline unsigned blocks_for(int64_t n) { return static_cast<unsigned>(ceil_div<int64_t>(n > 0 ? n : 1, kThreads)); }

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_version(void) { return 100; }

const char* kgcnn_error_string(int code) {
  switch (code) {
    case KGCNN_OK: return "ok";
    case KGCNN_E_INVALID: return "invalid argument";
    case KGCNN_E_UNSUPPORTED: return "unsupported shape";
    case KGCNN_E_CUDA: return "CUDA launch error";
    case KGCNN_E_WORKSPACE: return "workspace too small";
    case KGCNN_E_STATIC_LIMIT: return "problem larger than the fixed scratch array of a *_static entry point";
    default: return "unknown error";
  }
}

long long kgcnn_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
void kgcnn_launch_count_reset(void) { g_launches.store(0, std::memory_order_relaxed); }

int kgcnn_shift_index(const int64_t* idx, int64_t E, int K, const int64_t* node_row_splits,
                      const int64_t* edge_row_splits, int64_t B, int64_t* out_i64, int32_t* out_i32,
                      kgcnn_stream_t stream) {
  if (E < 0 || K <= 0 || B < 0) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!out_i64 && !out_i32) return KGCNN_E_INVALID;
  if (!idx || !node_row_splits || !edge_row_splits || B == 0) return KGCNN_E_INVALID;
  shift_index_kernel<<<blocks_for(E), kThreads, 0, as_stream(stream)>>>(idx, E, K, node_row_splits,
                                                                        edge_row_splits, B, out_i64, out_i32);
  return finish_launch();
}

int kgcnn_value_rowids(const int64_t* row_splits, int64_t B, int64_t n_values, int32_t* ids,
                       kgcnn_stream_t stream) {
  if (B < 0 || n_values < 0) return KGCNN_E_INVALID;
  if (n_values == 0) return KGCNN_OK;
  if (!row_splits || !ids || B == 0) return KGCNN_E_INVALID;
  value_rowids_kernel<<<blocks_for(n_values), kThreads, 0, as_stream(stream)>>>(row_splits, B, n_values, ids);
  return finish_launch();
}

int kgcnn_cast_i64_i32(const int64_t* src, int64_t n, int32_t* dst, kgcnn_stream_t stream) {
  if (n < 0) return KGCNN_E_INVALID;
  if (n == 0) return KGCNN_OK;
  if (!src || !dst) return KGCNN_E_INVALID;
  cast_i64_i32_kernel<<<blocks_for(n), kThreads, 0, as_stream(stream)>>>(src, n, dst);
  return finish_launch();
}

int kgcnn_check_sorted(int32_t* key, int key_stride, int64_t E, int64_t N, int32_t* flags, int clamp_in_place,
                       kgcnn_stream_t stream) {
  if (E < 0 || N < 0 || key_stride <= 0 || !flags) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!key || (clamp_in_place && N == 0)) return KGCNN_E_INVALID;
  check_sorted_kernel<<<blocks_for(E), kThreads, 0, as_stream(stream)>>>(key, key_stride, E, N, flags,
                                                                         clamp_in_place ? key : nullptr);
  return finish_launch();
}

size_t kgcnn_plan_workspace_bytes(int64_t E, int64_t N) {
  if (E < 0 || N < 0) return 0;
  // key_dense[E] + key_sorted[E] + iota[E] + cub temp
  return 3 * align256(static_cast<size_t>(E > 0 ? E : 1) * sizeof(int32_t)) + align256(cub_temp_bytes(E, N)) + 256;
}

int kgcnn_plan_build(const int32_t* key, int key_stride, int64_t E, int64_t N, int assume_sorted,
                     int32_t* rowptr, int32_t* perm, void* workspace, size_t workspace_bytes,
                     kgcnn_stream_t stream) {
  if (E < 0 || N < 0 || key_stride <= 0 || !rowptr) return KGCNN_E_INVALID;
  if (E > 0 && !key) return KGCNN_E_INVALID;
  if (E >= (int64_t(1) << 31) || N >= (int64_t(1) << 31) - 1) return KGCNN_E_UNSUPPORTED;
  cudaStream_t st = as_stream(stream);
  if (assume_sorted) {
    rowptr_from_sorted_kernel<<<blocks_for(E + 1), kThreads, 0, st>>>(key, key_stride, E, N, rowptr);
    int n = 1;
    if (perm && E > 0) {
      iota_kernel<<<blocks_for(E), kThreads, 0, st>>>(perm, E);
      ++n;
    }
    return finish_launch(n);
  }
  if (!perm && E > 0) return KGCNN_E_INVALID;
  if (workspace_bytes < kgcnn_plan_workspace_bytes(E, N) || !workspace) return KGCNN_E_WORKSPACE;
  if (cudaMemsetAsync(rowptr, 0, static_cast<size_t>(N + 1) * sizeof(int32_t), st) != cudaSuccess)
    return KGCNN_E_CUDA;
  if (E == 0) return KGCNN_OK;
  char* ws = static_cast<char*>(workspace);
  size_t seg = align256(static_cast<size_t>(E) * sizeof(int32_t));
  int32_t* key_dense = reinterpret_cast<int32_t*>(ws);
  int32_t* key_sorted = reinterpret_cast<int32_t*>(ws + seg);
  int32_t* iota = reinterpret_cast<int32_t*>(ws + 2 * seg);
  void* cub_ws = ws + 3 * seg;
  size_t cub_bytes = workspace_bytes - 3 * seg;
  key_iota_kernel<<<blocks_for(E), kThreads, 0, st>>>(key, key_stride, E, key_dense, iota);
  // LSD radix sort is stable: equal keys keep their original edge order (tf.argsort(stable=True)).
  if (cub::DeviceRadixSort::SortPairs(cub_ws, cub_bytes, key_dense, key_sorted, iota, perm, static_cast<int>(E), 0,
                                      key_bits(N), st) != cudaSuccess)
    return KGCNN_E_CUDA;
  // row pointer from the segment boundaries of the sorted keys: no histogram atomics (hot receivers made the
  // atomicAdd histogram 10x slower than the sort at 11 M triples), no scan
  rowptr_from_sorted_kernel<<<blocks_for(E + 1), kThreads, 0, st>>>(key_sorted, 1, E, N, rowptr);
  return finish_launch(3);
}

int kgcnn_permute_i32(const int32_t* src, int src_stride, const int32_t* perm, int64_t E, int32_t* out,
                      kgcnn_stream_t stream) {
  if (E < 0 || src_stride <= 0) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!src || !perm || !out) return KGCNN_E_INVALID;
  permute_i32_kernel<<<blocks_for(E), kThreads, 0, as_stream(stream)>>>(src, src_stride, perm, E, out);
  return finish_launch();
}

}  // extern "C"
