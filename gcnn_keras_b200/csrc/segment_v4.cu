// 128-bit instantiations of the fused gather*weight->segment-reduce kernel (F % 4 == 0, 16 B aligned).
#include "segment_impl.cuh"

namespace kgcnn {
namespace seg {

int launch_v4(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, bool short_rows,
              const LongRows& lr, cudaStream_t st) {
  return launch<float4>(mode, gather, reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(w), w,
                        rowptr, perm, gidx, gs, gp, n_out, F / 4, mean, reinterpret_cast<float4*>(y), short_rows, lr, st);
}

}  // namespace seg
}  // namespace kgcnn
