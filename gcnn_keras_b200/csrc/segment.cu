// C-ABI entry of the fused gather*weight->segment-reduce (see segment_impl.cuh for the kernel).
#include "segment_impl.cuh"

using namespace kgcnn;

extern "C" int kgcnn_gather_mul_segment_reduce(const float* x, int F, const float* w, int w_cols,
                                               const int32_t* rowptr, const int32_t* perm, const int32_t* gidx,
                                               int gidx_stride, int gidx_permuted, int64_t n_out, int64_t n_elements,
                                               int op, float* y,
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
  return vec ? seg::launch_v4(mode, gidx != nullptr, x, w, rowptr, perm, gidx, gidx_stride, gidx_permuted, n_out, F,
                              mean, y, short_rows, st)
             : seg::launch_v1(mode, gidx != nullptr, x, w, rowptr, perm, gidx, gidx_stride, gidx_permuted, n_out, F,
                              mean, y, short_rows, st);
}
