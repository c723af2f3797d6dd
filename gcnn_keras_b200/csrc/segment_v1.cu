// 32-bit instantiations of the fused gather*weight->segment-reduce kernel (any F, any alignment):
// F = 1 (energies, edge weights), 3 (coordinates), 35/50/140/500 (ACSF blocks).
#include "segment_impl.cuh"

namespace kgcnn {
namespace seg {

int launch_v1(int mode, bool gather, const float* x, const float* w, const int32_t* rowptr, const int32_t* perm,
              const int32_t* gidx, int gs, int gp, int64_t n_out, int F, int mean, float* y, bool short_rows,
              const LongRows& lr, cudaStream_t st) {
  return launch<float>(mode, gather, x, w, w, rowptr, perm, gidx, gs, gp, n_out, F, mean, y, short_rows, lr, st);
}

}  // namespace seg
}  // namespace kgcnn
