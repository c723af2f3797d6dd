// Ragged <-> padded conversion of flat values (a20): RaggedTensor.to_tensor + the ones_like mask of ChangeTensorType
// (kgcnn/layers/casting.py:78-84) in ONE pass, and its inverse, the boolean-mask gather of
// EnergyForceModel._cast_coordinates_pad_to_ragged (kgcnn/model/force.py:204-211).  Pure data movement: elements
// are handled as 32-bit words (float32 / int32: 1 word, int64: 2 words), so results are bit-exact by construction.
// pad and unpad are adjoint (default 0), which closes them under differentiation on the host side.
#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kThreads = 256;

// graph of value n: last b with row_splits[b] <= n (row_splits ascending, B + 1 entries)
__device__ __forceinline__ int64_t graph_of(const int64_t* __restrict__ splits, int64_t B, int64_t n) {
  int64_t lo = 0, hi = B - 1;
  while (lo < hi) {
    const int64_t mid = (lo + hi + 1) >> 1;
    if (__ldg(splits + mid) <= n) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// one thread per VEC words of the padded output; rows are W words wide, W % VEC == 0
template <typename V, int VEC>
__global__ void __launch_bounds__(kThreads)
pad_kernel(const V* __restrict__ values, const int64_t* __restrict__ splits, int64_t B, int64_t max_len, int64_t Wv,
           V fill, V one, V* __restrict__ out, V* __restrict__ mask, uint8_t* __restrict__ mask_bool) {
  const int64_t total = B * max_len * Wv;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(kThreads) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * kThreads) {
    const int64_t c = i % Wv, slot = i / Wv;
    const int64_t j = slot % max_len, b = slot / max_len;
    const int64_t beg = __ldg(splits + b), len = __ldg(splits + b + 1) - beg;
    const bool live = j < len;
    out[i] = live ? __ldg(values + (beg + j) * Wv + c) : fill;
    if (mask) {
      const V zero{};
      mask[i] = live ? one : zero;
    }
    if (mask_bool) {
#pragma unroll
      for (int q = 0; q < VEC; ++q) mask_bool[i * VEC + q] = live ? 1 : 0;
    }
  }
}

template <typename V>
__global__ void __launch_bounds__(kThreads)
unpad_kernel(const V* __restrict__ padded, const int64_t* __restrict__ splits, int64_t B, int64_t max_len,
             int64_t Wv, int64_t n_values, V* __restrict__ values) {
  const int64_t total = n_values * Wv;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(kThreads) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * kThreads) {
    const int64_t c = i % Wv, n = i / Wv;
    const int64_t b = graph_of(splits, B, n);
    const int64_t j = n - __ldg(splits + b);
    values[i] = j < max_len ? __ldg(padded + (b * max_len + j) * Wv + c) : V{};   // rows cut by the pad read as 0
  }
}

inline unsigned grid_for(int64_t items) {
  const int64_t blocks = ceil_div<int64_t>(items, kThreads);
  return static_cast<unsigned>(blocks < 1 ? 1 : (blocks > 8LL * kNumSM ? 8LL * kNumSM : blocks));
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_pad(const void* values, const int64_t* row_splits, int64_t B, int64_t max_len, int64_t words_per_row,
              uint32_t fill_word0, uint32_t fill_word1, int elem_words, void* out, void* mask, int mask_is_bool,
              kgcnn_stream_t stream) {
  if (B < 0 || max_len < 0 || words_per_row <= 0 || (elem_words != 1 && elem_words != 2) ||
      words_per_row % elem_words != 0)
    return KGCNN_E_INVALID;
  if (B == 0 || max_len == 0) return KGCNN_OK;
  if (!row_splits || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  const int64_t W = words_per_row;
  const bool boolmask = mask && mask_is_bool;
  // float mask value 1.0f (ones_like(flat_values)); for 2-word (int64) elements the mask is the integer 1
  const uint32_t one0 = elem_words == 1 ? 0x3f800000u : 1u;
  if (W % 4 == 0 && al16(values) && al16(out) && (!mask || boolmask || al16(mask))) {
    const uint4 fill = elem_words == 1 ? make_uint4(fill_word0, fill_word0, fill_word0, fill_word0)
                                       : make_uint4(fill_word0, fill_word1, fill_word0, fill_word1);
    const uint4 one = elem_words == 1 ? make_uint4(one0, one0, one0, one0) : make_uint4(1u, 0u, 1u, 0u);
    if (boolmask && elem_words == 2) return KGCNN_E_INVALID;      // bool masks are defined on 1-word elements
    pad_kernel<uint4, 4><<<grid_for(B * max_len * (W / 4)), kThreads, 0, st>>>(
        static_cast<const uint4*>(values), row_splits, B, max_len, W / 4, fill, one, static_cast<uint4*>(out),
        boolmask ? nullptr : static_cast<uint4*>(mask), boolmask ? static_cast<uint8_t*>(mask) : nullptr);
  } else if (elem_words == 2) {
    if (boolmask) return KGCNN_E_INVALID;
    pad_kernel<uint2, 2><<<grid_for(B * max_len * (W / 2)), kThreads, 0, st>>>(
        static_cast<const uint2*>(values), row_splits, B, max_len, W / 2, make_uint2(fill_word0, fill_word1),
        make_uint2(1u, 0u), static_cast<uint2*>(out), static_cast<uint2*>(mask), nullptr);
  } else {
    pad_kernel<uint32_t, 1><<<grid_for(B * max_len * W), kThreads, 0, st>>>(
        static_cast<const uint32_t*>(values), row_splits, B, max_len, W, fill_word0, one0,
        static_cast<uint32_t*>(out), boolmask ? nullptr : static_cast<uint32_t*>(mask),
        boolmask ? static_cast<uint8_t*>(mask) : nullptr);
  }
  return finish_launch();
}

int kgcnn_unpad(const void* padded, const int64_t* row_splits, int64_t B, int64_t max_len, int64_t words_per_row,
                int64_t n_values, void* values, kgcnn_stream_t stream) {
  if (B < 0 || max_len < 0 || words_per_row <= 0 || n_values < 0) return KGCNN_E_INVALID;
  if (n_values == 0) return KGCNN_OK;
  if (!padded || !row_splits || !values || B == 0) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  const int64_t W = words_per_row;
  if (W % 4 == 0 && al16(padded) && al16(values))
    unpad_kernel<uint4><<<grid_for(n_values * (W / 4)), kThreads, 0, st>>>(
        static_cast<const uint4*>(padded), row_splits, B, max_len, W / 4, n_values, static_cast<uint4*>(values));
  else
    unpad_kernel<uint32_t><<<grid_for(n_values * W), kThreads, 0, st>>>(
        static_cast<const uint32_t*>(padded), row_splits, B, max_len, W, n_values, static_cast<uint32_t*>(values));
  return finish_launch();
}

}  // extern "C"
