// C-ABI entry of the fused gather*weight->segment-reduce (see segment_impl.cuh for the kernel).
#include "segment_impl.cuh"

using namespace kgcnn;

extern "C" int kgcnn_gather_mul_segment_reduce(const float* x, int F, const float* w, int w_cols,
                                               const int32_t* rowptr, const int32_t* perm, const int32_t* gidx,
                                               int gidx_stride, int gidx_permuted, int64_t n_out, int64_t n_elements,
                                               int op, float* y, const int32_t* long_rows,
                                               const int32_t* n_long, int long_cap, int long_thresh,
                                               kgcnn_stream_t stream) {
  if (F <= 0 || n_out < 0 || n_elements < 0 || op < KGCNN_OP_SUM || op > KGCNN_OP_MIN) return KGCNN_E_INVALID;
  if (w_cols != 0 && w_cols != 1 && w_cols != F) return KGCNN_E_INVALID;
  if ((op == KGCNN_OP_MAX || op == KGCNN_OP_MIN) && (w_cols != 0 || gidx)) return KGCNN_E_UNSUPPORTED;
  if (n_out == 0) return KGCNN_OK;
  if (!x || !rowptr || !y || (w_cols && !w) || (gidx && gidx_stride <= 0)) return KGCNN_E_INVALID;
  int mode;
  if (op == KGCNN_OP_MAX) mode = seg::kMax;
  else if (op == KGCNN_OP_MIN) mode = seg::kMin;
  else if (w_cols == 0) mode = seg::kSumW0;
  else if (w_cols == 1 && F != 1) mode = seg::kSumW1;
  else mode = (F == 1 && w_cols == 1) ? seg::kSumW1 : seg::kSumW2;
  const int mean = (op == KGCNN_OP_MEAN) ? 1 : 0;
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = (F % 4 == 0) && al(x) && al(y) && (mode != seg::kSumW2 || al(w));
  cudaStream_t st = as_stream(stream);
  // the permutation is only needed to address w rows (or x rows when x is the edge tensor itself)
  if (gidx && gidx_permuted && w_cols == 0) perm = nullptr;
  // short segments (molecular graphs): one lane group per row instead of one warp per row
  const bool short_rows = n_elements <= 64 * n_out;
  seg::LongRows lr;
  if (long_rows && n_long && long_cap > 0 && long_thresh > 0) {
    lr.rows = long_rows; lr.count = n_long; lr.cap = long_cap; lr.thresh = long_thresh;
  }
  return vec ? seg::launch_v4(mode, gidx != nullptr, x, w, rowptr, perm, gidx, gidx_stride, gidx_permuted, n_out, F,
                              mean, y, short_rows, lr, st)
             : seg::launch_v1(mode, gidx != nullptr, x, w, rowptr, perm, gidx, gidx_stride, gidx_permuted, n_out, F,
                              mean, y, short_rows, lr, st);
}


namespace {
__global__ void row_stats_kernel(const int32_t* __restrict__ rowptr, int64_t n, int32_t* __restrict__ stats) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  int deg = 0;
  if (i < n) deg = rowptr[i + 1] - rowptr[i];
  for (int o = 16; o > 0; o >>= 1) deg = max(deg, __shfl_xor_sync(0xffffffffu, deg, o));
  if ((threadIdx.x & 31) == 0 && deg > 0) atomicMax(stats, deg);     // integer max: order independent
}
__global__ void long_rows_kernel(const int32_t* __restrict__ rowptr, int64_t n, int thresh, int cap,
                                 int32_t* __restrict__ rows, int32_t* __restrict__ count) {
  int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  if (rowptr[i + 1] - rowptr[i] > thresh) {
    int slot = atomicAdd(count, 1);     // list order is irrelevant: every hub row is reduced by its own CTA
    if (slot < cap) rows[slot] = static_cast<int32_t>(i);
  }
}
}  // namespace

extern "C" int kgcnn_row_stats(const int32_t* rowptr, int64_t n_rows, int32_t* stats, kgcnn_stream_t stream) {
  if (n_rows < 0 || !stats) return KGCNN_E_INVALID;
  if (n_rows == 0) return KGCNN_OK;
  if (!rowptr) return KGCNN_E_INVALID;
  row_stats_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(n_rows, 256)), 256, 0, as_stream(stream)>>>(rowptr, n_rows,
                                                                                                      stats);
  return finish_launch();
}

extern "C" int kgcnn_long_rows_build(const int32_t* rowptr, int64_t n_rows, int thresh, int cap, int32_t* rows,
                                     int32_t* count, kgcnn_stream_t stream) {
  if (n_rows < 0 || thresh <= 0 || cap <= 0 || !rows || !count) return KGCNN_E_INVALID;
  if (n_rows == 0) return KGCNN_OK;
  if (!rowptr) return KGCNN_E_INVALID;
  long_rows_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(n_rows, 256)), 256, 0, as_stream(stream)>>>(
      rowptr, n_rows, thresh, cap, rows, count);
  return finish_launch();
}
