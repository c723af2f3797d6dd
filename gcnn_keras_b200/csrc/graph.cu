// Batch graph construction on the device: the step in front of the message-passing path.
//   range neighbours : kgcnn/graph/adj.py:466-481 (coordinates_to_distancematrix) + :537-593
//                      (define_adjacency_from_distance), called by SetRange (kgcnn/graph/preprocessor.py:288-313)
//   angle triples    : kgcnn/graph/adj.py:300-385 (get_angle_indices, check_sorted=True order), called by
//                      SetAngle (kgcnn/graph/preprocessor.py:341)
// The reference runs both per molecule on the host (dense N x N distance matrix + argsort; a Python loop over
// every edge).  Here one count pass + scan + emit pass handles the whole ragged batch.  Integer outputs are
// bit-exact: distances are evaluated in the coordinate dtype with round-to-nearest mul/add/sqrt in NumPy's
// order ((dx^2 + dy^2) + dz^2, no FMA contraction), ties in the neighbour ranking are broken by index.
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

__device__ __forceinline__ int row_of64(const int64_t* __restrict__ splits, int64_t nrows, int64_t e) {
  int64_t lo = 0, hi = nrows;
  while (hi - lo > 1) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(splits + mid) <= e) lo = mid; else hi = mid;
  }
  return static_cast<int>(lo);
}

struct CastI64 {
  __host__ __device__ int64_t operator()(const int32_t& v) const { return static_cast<int64_t>(v); }
};

// ---------------------------------------------------------------------------------------------------------
// angle triples
// ---------------------------------------------------------------------------------------------------------
struct AngleFlags {
  int pos_fix, pos_ij, pos_k, allow_multi, allow_reverse, allow_self;
};

// partner edge m of first edge n: the mask of adj.py:346-358
__device__ __forceinline__ bool angle_keep(int m, int n, int m0, int m1, int n0, int n1, const AngleFlags& f) {
  if (m == n) return f.allow_self != 0;
  bool keep = true;
  if (!f.allow_multi) keep = keep && (m0 != n0 || m1 != n1);
  if (!f.allow_reverse) keep = keep && (m0 != n1 || m1 != n0);
  return keep;
}

__global__ void angle_count_kernel(const int32_t* __restrict__ idx, int64_t M, int64_t N,
                                   const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
                                   AngleFlags f, int32_t* __restrict__ count) {
  int64_t n = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (n >= M) return;
  int n0 = idx[n], n1 = idx[M + n];
  int key = f.pos_ij == 0 ? n0 : n1;
  int c = 0;
  if (key >= 0 && key < N) {
    int lo = rowptr[key], hi = rowptr[key + 1];
    for (int p = lo; p < hi; ++p) {
      int m = perm ? perm[p] : p;
      c += angle_keep(m, static_cast<int>(n), idx[m], idx[M + m], n0, n1, f) ? 1 : 0;
    }
  }
  count[n] = c;
}

// one warp per first edge n: partners in ascending label order (the CSR permutation is stable), compacted
// with a ballot so the [A,3] rows of one edge leave as one contiguous run
__global__ void angle_emit_kernel(const int64_t* __restrict__ idx_sample, const int32_t* __restrict__ idx,
                                  int64_t M, int64_t N, const int32_t* __restrict__ rowptr,
                                  const int32_t* __restrict__ perm, AngleFlags f,
                                  const int64_t* __restrict__ offsets, const int64_t* __restrict__ edge_splits,
                                  int64_t B, int64_t* __restrict__ ijk, int64_t* __restrict__ ij_k,
                                  int64_t* __restrict__ angle_splits) {
  int64_t tid = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (angle_splits && tid <= B) angle_splits[tid] = offsets[edge_splits[tid]];
  int64_t n = tid >> 5;
  int lane = threadIdx.x & 31;
  if (n >= M) return;
  int n0 = idx[n], n1 = idx[M + n];
  int key = f.pos_ij == 0 ? n0 : n1;
  if (key < 0 || key >= N) return;
  int lo = rowptr[key], hi = rowptr[key + 1];
  int64_t base = offsets[n];
  if (offsets[n + 1] == base) return;
  int64_t is = idx_sample[2 * n], js = idx_sample[2 * n + 1];
  int64_t e0 = 0;
  if (ij_k) e0 = edge_splits[row_of64(edge_splits, B, n)];
  for (int p = lo; p < hi; p += 32) {
    int q = p + lane;
    bool valid = false;
    int m = 0;
    if (q < hi) {
      m = perm ? perm[q] : q;
      valid = angle_keep(m, static_cast<int>(n), idx[m], idx[M + m], n0, n1, f);
    }
    unsigned mask = __ballot_sync(0xffffffffu, valid);
    if (valid) {
      int64_t o = base + __popc(mask & ((1u << lane) - 1u));
      ijk[3 * o] = is;
      ijk[3 * o + 1] = js;
      ijk[3 * o + 2] = idx_sample[2 * static_cast<int64_t>(m) + f.pos_k];
      if (ij_k) {
        ij_k[2 * o] = n - e0;
        ij_k[2 * o + 1] = m - e0;
      }
    }
    base += __popc(mask);
  }
}

// ---------------------------------------------------------------------------------------------------------
// range neighbours
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float dist_rn(float ax, float ay, float az, float bx, float by, float bz) {
  float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
  return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
}
__device__ __forceinline__ double dist_rn(double ax, double ay, double az, double bx, double by, double bz) {
  double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
  return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
}

struct RangeFlags {
  int use_dist, max_nb, exclusive, self_loops;  // max_nb < 0: no neighbour limit
};

// One warp per centre atom i.  The row d[i, :] of its graph's distance matrix is staged in shared memory;
// atom j is kept if (d_ij < max_distance) and/or (rank of j in the ascending row <= max_neighbours), the two
// terms combined as adj.py:564-587 does (exclusive: and / else: or), diagonal removed unless self_loops.
// EMIT = false: count[i];  EMIT = true: rows [i, j] (sample indexing) in ascending j from offsets[i].
template <typename T, bool EMIT>
__global__ void range_kernel(const T* __restrict__ xyz, const int64_t* __restrict__ node_splits, int64_t B,
                             int64_t N, T max_dist, RangeFlags f, int cap, int32_t* __restrict__ count,
                             const int64_t* __restrict__ offsets, int64_t* __restrict__ edges,
                             T* __restrict__ dist_out, int64_t* __restrict__ edge_splits) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* row = reinterpret_cast<T*>(smem_raw) + static_cast<size_t>(threadIdx.x >> 5) * cap;
  int64_t tid = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (EMIT && edge_splits && tid <= B) edge_splits[tid] = offsets[node_splits[tid]];
  int64_t i = tid >> 5;
  int lane = threadIdx.x & 31;
  if (i >= N) return;
  int g = row_of64(node_splits, B, i);
  int64_t a0 = node_splits[g];
  int ng = static_cast<int>(node_splits[g + 1] - a0);
  T xi = xyz[3 * i], yi = xyz[3 * i + 1], zi = xyz[3 * i + 2];
  for (int j = lane; j < ng; j += 32) {
    const T* pj = xyz + 3 * (a0 + j);
    row[j] = dist_rn(xi, yi, zi, pj[0], pj[1], pj[2]);   // c[i, j] = coord[i] - coord[j]
  }
  __syncwarp();
  // adj.py:574: max_neighbours = min(max_neighbours, num_atoms); the first max_neighbours + 1 of the argsort
  bool use_nb = f.max_nb >= 0;
  int nb = use_nb ? min(f.max_nb, ng) : 0;
  bool rank_needed = use_nb && (nb + 1 < ng);
  int self = static_cast<int>(i - a0);
  int64_t base = EMIT ? offsets[i] : 0;
  int total = 0;
  for (int j0 = 0; j0 < ng; j0 += 32) {
    int j = j0 + lane;
    bool keep = false;
    T dj = 0;
    if (j < ng) {
      dj = row[j];
      bool in_dist = dj < max_dist;
      bool in_nb = true;
      if (rank_needed) {
        int rank = 0;
        for (int k = 0; k < ng; ++k) {
          T dk = row[k];
          rank += (dk < dj || (dk == dj && k < j)) ? 1 : 0;
        }
        in_nb = rank <= nb;
      }
      if (f.exclusive) keep = (!f.use_dist || in_dist) && (!use_nb || in_nb);
      else keep = (f.use_dist && in_dist) || (use_nb && in_nb);
      if (!f.self_loops && j == self) keep = false;
    }
    unsigned mask = __ballot_sync(0xffffffffu, keep);
    if (EMIT && keep) {
      int64_t o = base + __popc(mask & ((1u << lane) - 1u));
      edges[2 * o] = self;
      edges[2 * o + 1] = j;
      if (dist_out) dist_out[o] = dj;
    }
    base += __popc(mask);
    total += __popc(mask);
  }
  if (!EMIT && lane == 0) count[i] = total;
}

inline unsigned blocks_for(int64_t n) { return static_cast<unsigned>(ceil_div<int64_t>(n > 0 ? n : 1, kThreads)); }

template <typename T, bool EMIT>
int launch_range(const void* xyz, const int64_t* node_splits, int64_t B, int64_t N, int max_nodes, double max_distance,
                 RangeFlags f, int32_t* count, const int64_t* offsets, int64_t* edges, void* dist,
                 int64_t* edge_splits, cudaStream_t st) {
  int cap = (max_nodes + 3) & ~3;
  size_t smem = static_cast<size_t>(kWarps) * cap * sizeof(T);
  if (smem > 200 * 1024) return KGCNN_E_UNSUPPORTED;
  auto kern = range_kernel<T, EMIT>;
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return KGCNN_E_CUDA;
  int64_t threads = std::max<int64_t>(N * 32, B + 1);
  kern<<<blocks_for(threads), kThreads, smem, st>>>(static_cast<const T*>(xyz), node_splits, B, N,
                                                    static_cast<T>(max_distance), f, cap, count, offsets, edges,
                                                    static_cast<T*>(dist), edge_splits);
  return finish_launch();
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

size_t kgcnn_offsets_workspace_bytes(int64_t n) {
  if (n < 0) return 0;
  size_t bytes = 0;
  cub::TransformInputIterator<int64_t, CastI64, const int32_t*> it(nullptr, CastI64());
  cub::DeviceScan::InclusiveSum(nullptr, bytes, it, (int64_t*)nullptr, static_cast<int>(n > 0 ? n : 1));
  return bytes + 256;
}

int kgcnn_offsets_from_counts(const int32_t* count, int64_t n, int64_t* offsets, void* workspace,
                              size_t workspace_bytes, kgcnn_stream_t stream) {
  if (n < 0 || !offsets) return KGCNN_E_INVALID;
  if (n >= (int64_t(1) << 31)) return KGCNN_E_UNSUPPORTED;
  cudaStream_t st = as_stream(stream);
  if (cudaMemsetAsync(offsets, 0, sizeof(int64_t), st) != cudaSuccess) return KGCNN_E_CUDA;
  if (n == 0) return KGCNN_OK;
  if (!count) return KGCNN_E_INVALID;
  if (!workspace || workspace_bytes < kgcnn_offsets_workspace_bytes(n)) return KGCNN_E_WORKSPACE;
  cub::TransformInputIterator<int64_t, CastI64, const int32_t*> it(count, CastI64());
  if (cub::DeviceScan::InclusiveSum(workspace, workspace_bytes, it, offsets + 1, static_cast<int>(n), st) !=
      cudaSuccess)
    return KGCNN_E_CUDA;
  return finish_launch();
}

static int angle_flags(int pos_fix, int pos_ij, int pos_k, int allow_multi, int allow_reverse, int allow_self,
                       AngleFlags* f) {
  if ((pos_fix | 1) != 1 || (pos_ij | 1) != 1 || (pos_k | 1) != 1 || pos_k == pos_fix) return KGCNN_E_INVALID;
  // allow_self_edges forces mask[n] = True even when edge n does not match its own key (adj.py:355-356);
  // that only coincides with the CSR list when the fixed column is the key column ('ik', 'kj').
  if (allow_self && pos_fix != pos_ij) return KGCNN_E_UNSUPPORTED;
  *f = AngleFlags{pos_fix, pos_ij, pos_k, allow_multi, allow_reverse, allow_self};
  return KGCNN_OK;
}

int kgcnn_angle_count(const int32_t* idx_cm, int64_t M, int64_t N, const int32_t* rowptr, const int32_t* perm,
                      int pos_fix, int pos_ij, int allow_multi, int allow_reverse, int allow_self, int32_t* count,
                      kgcnn_stream_t stream) {
  AngleFlags f;
  int rc = angle_flags(pos_fix, pos_ij, 1 - pos_fix, allow_multi, allow_reverse, allow_self, &f);
  if (rc != KGCNN_OK) return rc;
  if (M < 0 || N < 0) return KGCNN_E_INVALID;
  if (M == 0) return KGCNN_OK;
  if (M >= (int64_t(1) << 31)) return KGCNN_E_UNSUPPORTED;
  if (!idx_cm || !rowptr || !count) return KGCNN_E_INVALID;
  angle_count_kernel<<<blocks_for(M), kThreads, 0, as_stream(stream)>>>(idx_cm, M, N, rowptr, perm, f, count);
  return finish_launch();
}

int kgcnn_angle_emit(const int64_t* idx_sample, const int32_t* idx_cm, int64_t M, int64_t N, const int32_t* rowptr,
                     const int32_t* perm, int pos_fix, int pos_ij, int allow_multi, int allow_reverse,
                     int allow_self, const int64_t* offsets, const int64_t* edge_row_splits, int64_t B,
                     int64_t* ijk, int64_t* ij_k, int64_t* angle_row_splits, kgcnn_stream_t stream) {
  AngleFlags f;
  int rc = angle_flags(pos_fix, pos_ij, 1 - pos_fix, allow_multi, allow_reverse, allow_self, &f);
  if (rc != KGCNN_OK) return rc;
  if (M < 0 || N < 0 || B < 0 || !offsets || !edge_row_splits) return KGCNN_E_INVALID;
  if (M >= (int64_t(1) << 31)) return KGCNN_E_UNSUPPORTED;
  if (M > 0 && (!idx_sample || !idx_cm || !rowptr)) return KGCNN_E_INVALID;
  int64_t threads = std::max<int64_t>(M * 32, angle_row_splits ? B + 1 : 0);
  if (threads == 0) return KGCNN_OK;
  angle_emit_kernel<<<blocks_for(threads), kThreads, 0, as_stream(stream)>>>(
      idx_sample, idx_cm, M, N, rowptr, perm, f, offsets, edge_row_splits, B, ijk, ij_k, angle_row_splits);
  return finish_launch();
}

int kgcnn_range_count(const void* xyz, int is_f64, const int64_t* node_row_splits, int64_t B, int64_t N,
                      int max_nodes, int use_max_distance, double max_distance, int max_neighbours, int exclusive,
                      int self_loops, int32_t* count, kgcnn_stream_t stream) {
  if (B < 0 || N < 0 || max_nodes < 0) return KGCNN_E_INVALID;
  if (N == 0) return KGCNN_OK;
  if (!xyz || !node_row_splits || !count || B == 0) return KGCNN_E_INVALID;
  RangeFlags f{use_max_distance, max_neighbours, exclusive, self_loops};
  cudaStream_t st = as_stream(stream);
  return is_f64 ? launch_range<double, false>(xyz, node_row_splits, B, N, max_nodes, max_distance, f, count, nullptr,
                                              nullptr, nullptr, nullptr, st)
                : launch_range<float, false>(xyz, node_row_splits, B, N, max_nodes, max_distance, f, count, nullptr,
                                             nullptr, nullptr, nullptr, st);
}

int kgcnn_range_emit(const void* xyz, int is_f64, const int64_t* node_row_splits, int64_t B, int64_t N,
                     int max_nodes, int use_max_distance, double max_distance, int max_neighbours, int exclusive,
                     int self_loops, const int64_t* offsets, int64_t* edges, void* dist, int64_t* edge_row_splits,
                     kgcnn_stream_t stream) {
  if (B < 0 || N < 0 || max_nodes < 0 || !offsets) return KGCNN_E_INVALID;
  if (N > 0 && (!xyz || !node_row_splits || B == 0)) return KGCNN_E_INVALID;
  if (N == 0 && !edge_row_splits) return KGCNN_OK;
  if (!node_row_splits) return KGCNN_E_INVALID;
  RangeFlags f{use_max_distance, max_neighbours, exclusive, self_loops};
  cudaStream_t st = as_stream(stream);
  return is_f64 ? launch_range<double, true>(xyz, node_row_splits, B, N, max_nodes, max_distance, f, nullptr, offsets,
                                             edges, dist, edge_row_splits, st)
                : launch_range<float, true>(xyz, node_row_splits, B, N, max_nodes, max_distance, f, nullptr, offsets,
                                            edges, dist, edge_row_splits, st);
}

}  // extern "C"
