// Dense contractions of the edge/node feature rows on the 5th-generation tensor cores (tcgen05 + TMEM):
//     NN : Y[M,U]  = X[M,K]  W[K,U] (+ bias)        forward of kgcnn/layers/modules.py:85
//     NT : dX[M,K] = G[M,U]  W[K,U]^T               input gradient
//     TN : dW[K,U] = X[M,K]^T G[M,U]                weight gradient (split over M, fixed-order reduce)
// M = edges / atoms of the batch (1e4 .. 1e6), K and U = layer widths (1 .. 640).
//
// Precision: the parity bar is 1e-5 of an fp64 oracle, single-pass TF32 (10-bit mantissa) is ~1e-3.  Every operand
// tile is therefore split while it is staged into shared memory, x = hi + lo with hi = x & 0xffffe000 (exactly
// representable in TF32) and lo = x - hi (exact in fp32), and three MMAs accumulate hi*hi + lo*hi + hi*lo in the
// fp32 TMEM accumulator ("3xTF32"); the dropped lo*lo term is < 2^-20 relative.
//
// Data path: these GEMMs are skinny (K, U <= 640) and memory-bound, and the split needs a register pass anyway, so
// tiles go global -> registers -> (hi, lo) -> shared memory in the canonical no-swizzle K-major UMMA layout (8 x 16 B
// core matrices); operands whose reduction index is the ROW of the global matrix (W in NN, X and G in TN) are
// transposed on the way into shared memory, so every descriptor is K-major; no TMA descriptor is needed.  One elected thread issues tcgen05.mma (kind::tf32, cta_group::1,
// M = 128), tcgen05.commit signals an mbarrier, the four warps read their 32 TMEM lanes back with tcgen05.ld.
#include <algorithm>

#include "common.cuh"

namespace kgcnn {
namespace tc {

constexpr int kThreads = 128;   // 4 warps = 128 TMEM lanes
constexpr int kBM = 128;        // MMA M
constexpr int kKC = 32;         // reduction elements staged per chunk (4 MMA k-steps of 8)
constexpr int kMaxN = 128;      // MMA N per CTA (TMEM columns)

enum { kNN = 0, kNT = 1, kTN = 2 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (sm_100): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo16, uint32_t sbo16) {
  return static_cast<uint64_t>((saddr >> 4) & 0x3FFF) | (static_cast<uint64_t>(lbo16 & 0x3FFF) << 16) |
         (static_cast<uint64_t>(sbo16 & 0x3FFF) << 32) | (1ull << 46);
}

// instruction descriptor kind::tf32: D=F32 [4,6), A=TF32 [7,10), B=TF32 [10,13), a_major 15, b_major 16,
// N>>3 [17,23), M>>4 [24,29)
__device__ __forceinline__ uint32_t make_idesc(int a_mn_major, int b_mn_major, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(a_mn_major) << 15) |
         (static_cast<uint32_t>(b_mn_major) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(kBM >> 4) << 24);
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  int spins = 0;
  while (!done) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && ++spins > (1 << 22)) __trap();   // never hang the device on a lost arrival
  }
}

// K-major operand plane: element (mn, k) lives in float4 slot (k/4)*NRP + mn, component k%4, NRP = NR + 1
// (odd pitch).  Descriptor: LBO = NRP (next 16 B k-chunk), SBO = 8 (next 8 mn), +2*NRP per MMA k-step of 8.
__device__ __forceinline__ float4 load4(const float* __restrict__ g, int64_t row, int64_t row_end, int f, int f_end,
                                        int ld, bool vec_ok) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < row_end) {
    const float* p = g + row * ld + f;
    if (vec_ok && f + 3 < f_end) v = __ldg(reinterpret_cast<const float4*>(p));
    else {
      if (f + 0 < f_end) v.x = __ldg(p + 0);
      if (f + 1 < f_end) v.y = __ldg(p + 1);
      if (f + 2 < f_end) v.z = __ldg(p + 2);
      if (f + 3 < f_end) v.w = __ldg(p + 3);
    }
  }
  return v;
}
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// mn = row of the global matrix, k = feature: rows [row0, row0+NR), features [f0, f0+kKC)
template <int NR>
__device__ __forceinline__ void stage_direct(const float* __restrict__ g, int64_t row0, int64_t row_end, int f0,
                                             int f_end, int ld, bool vec_ok, float4* __restrict__ hi,
                                             float4* __restrict__ lo) {
  constexpr int NRP = NR + 1, NCH = kKC / 4;
  for (int it = threadIdx.x; it < NR * NCH; it += kThreads) {
    const int c = it % NCH, r = it / NCH;                 // consecutive threads: consecutive 16 B of one row
    const float4 v = load4(g, row0 + r, row_end, f0 + 4 * c, f_end, ld, vec_ok);
    const float4 h = make_float4(tf32_hi(v.x), tf32_hi(v.y), tf32_hi(v.z), tf32_hi(v.w));
    hi[c * NRP + r] = h;
    lo[c * NRP + r] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
  }
}

// mn = feature, k = row of the global matrix (transposed on the way in): rows [row0, row0+kKC), features
// [f0, f0+4*nchunks).  Consecutive threads take consecutive rows -> the scalar stores hit 32 different banks.
template <int NR>
__device__ __forceinline__ void stage_transposed(const float* __restrict__ g, int64_t row0, int64_t row_end, int f0,
                                                 int f_end, int ld, int nchunks, bool vec_ok, float4* __restrict__ hi,
                                                 float4* __restrict__ lo) {
  constexpr int NRP = NR + 1;
  float* hs = reinterpret_cast<float*>(hi);
  float* ls = reinterpret_cast<float*>(lo);
  for (int it = threadIdx.x; it < kKC * nchunks; it += kThreads) {
    const int k = it % kKC, j = it / kKC;
    const float4 v = load4(g, row0 + k, row_end, f0 + 4 * j, f_end, ld, vec_ok);
    const float vv[4] = {v.x, v.y, v.z, v.w};
    const int base = ((k >> 2) * NRP + 4 * j) * 4 + (k & 3);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float h = tf32_hi(vv[i]);
      hs[base + 4 * i] = h;
      ls[base + 4 * i] = vv[i] - h;
    }
  }
}

// One CTA = one 128-row output tile (NN / NT) or one (128-feature tile, split of the rows) (TN).
template <int OP>
__global__ void __launch_bounds__(kThreads)
dense_tc_kernel(const float* __restrict__ A, const float* __restrict__ B, int64_t M, int K, int U,
                const float* __restrict__ bias, float* __restrict__ out, int64_t rows_per_split, int vec_a,
                int vec_b) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  constexpr int PLANE = (kKC / 4) * (kBM + 1);     // float4 slots per plane: 8 k-chunks x (128 + 1) mn
  float4* a_hi = reinterpret_cast<float4*>(smem_raw);
  float4* a_lo = a_hi + PLANE;
  float4* b_hi = a_lo + PLANE;
  float4* b_lo = b_hi + PLANE;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // output-tile coordinates
  //   NN: D[rows m0.., cols n0..] = sum_k X[m,k] W[k,n]        reduction over K features of X / rows of W
  //   NT: D[rows m0.., cols n0..] = sum_u G[m,u] W[n,u]        reduction over U; n indexes rows of W (K of them)
  //   TN: D[feat a0.., cols n0..] = sum_rows X[row,a] G[row,n] reduction over the rows of this split
  const int n_total = (OP == kNN) ? U : (OP == kNT ? K : U);
  const int n0 = blockIdx.y * kMaxN;
  const int nt = min(kMaxN, ((n_total - n0 + 15) / 16) * 16);        // MMA N (multiple of 16, zero padded)
  const int64_t m0 = (OP == kTN) ? 0 : static_cast<int64_t>(blockIdx.x) * kBM;
  const int a0 = (OP == kTN) ? blockIdx.x * kBM : 0;
  const int64_t row_beg = (OP == kTN) ? static_cast<int64_t>(blockIdx.z) * rows_per_split : 0;
  const int64_t row_end = (OP == kTN) ? min(M, row_beg + rows_per_split) : 0;
  const int64_t red_len = (OP == kNN) ? K : (OP == kNT ? U : (row_end > row_beg ? row_end - row_beg : 0));

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(static_cast<uint32_t>(kMaxN))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_s;
  const uint32_t bar = smem_u32(&mbar);

  // every operand is K-major with 128 (+1) mn slots per k-chunk: LBO = 129, SBO = 8 (units of 16 B)
  constexpr int NRP = kBM + 1;
  static_assert(kMaxN == kBM, "A and B planes share one pitch");
  constexpr uint32_t kstep = 2 * NRP;                          // descriptor advance per MMA k-step (8 elements)
  const uint32_t idesc = make_idesc(0, 0, nt);

  uint32_t parity = 0;
  bool first = true;
  for (int64_t k0 = 0; k0 < red_len; k0 += kKC) {
    // ---- stage A and B (hi / lo) ----
    if (OP == kNN) {
      stage_direct<kBM>(A, m0, M, static_cast<int>(k0), K, K, vec_a, a_hi, a_lo);                 // X rows, K feats
      stage_transposed<kMaxN>(B, k0, K, n0, U, U, nt / 4, vec_b, b_hi, b_lo);                     // W rows k -> K dim
    } else if (OP == kNT) {
      stage_direct<kBM>(A, m0, M, static_cast<int>(k0), U, U, vec_a, a_hi, a_lo);                 // G rows, U feats
      stage_direct<kMaxN>(B, n0, K, static_cast<int>(k0), U, U, vec_b, b_hi, b_lo);               // W rows n, U feats
    } else {
      stage_transposed<kBM>(A, row_beg + k0, row_end, a0, K, K, kBM / 4, vec_a, a_hi, a_lo);      // X rows -> K dim
      stage_transposed<kMaxN>(B, row_beg + k0, row_end, n0, U, U, nt / 4, vec_b, b_hi, b_lo);     // G rows -> K dim
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> async proxy (MMA)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
      if (lane == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
#pragma unroll
        for (int ks = 0; ks < kKC / 8; ++ks) {
          const uint64_t dah = make_desc(ah + ks * kstep * 16, NRP, 8);
          const uint64_t dal = make_desc(al + ks * kstep * 16, NRP, 8);
          const uint64_t dbh = make_desc(bh + ks * kstep * 16, NRP, 8);
          const uint64_t dbl = make_desc(bl + ks * kstep * 16, NRP, 8);
          mma_tf32(tmem_d, dah, dbh, idesc, (first && ks == 0) ? 0u : 1u);     // hi * hi
          mma_tf32(tmem_d, dal, dbh, idesc, 1u);                                // lo * hi
          mma_tf32(tmem_d, dah, dbl, idesc, 1u);                                // hi * lo
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                     : "memory");
      }
      __syncwarp();
    }
    first = false;
    mbar_wait(bar, parity);          // MMAs of this chunk are done: shared memory may be overwritten
    parity ^= 1;
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  // ---- epilogue: TMEM lane (32*warp + lane) = output row, columns = n ----
  const int d_row = 32 * warp + lane;
  for (int c0 = 0; c0 < nt; c0 += 32) {
    uint32_t v[32];
    if (red_len > 0) {
      const uint32_t taddr = tmem_d + (static_cast<uint32_t>(32 * warp) << 16) + static_cast<uint32_t>(c0);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
            "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
            "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
            "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 0u;
    }
    if (OP == kTN) {
      const int feat = a0 + d_row;
      if (feat < K) {
        float* o = out + (static_cast<int64_t>(blockIdx.z) * K + feat) * U + n0 + c0;
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (n0 + c0 + i < U) o[i] = __uint_as_float(v[i]);
      }
    } else {
      const int64_t row = m0 + d_row;
      if (row < M) {
        float* o = out + row * n_total + n0 + c0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int col = n0 + c0 + i;
          if (col < n_total) o[i] = __uint_as_float(v[i]) + ((OP == kNN && bias) ? __ldg(bias + col) : 0.f);
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                 "r"(static_cast<uint32_t>(kMaxN))
                 : "memory");
  }
}

__global__ void __launch_bounds__(256)
tn_split_reduce_kernel(const float* __restrict__ partial, int S, int64_t n, float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int s = 0;
  for (; s + 3 < S; s += 4) {
    s0 += partial[(s + 0) * n + i];
    s1 += partial[(s + 1) * n + i];
    s2 += partial[(s + 2) * n + i];
    s3 += partial[(s + 3) * n + i];
  }
  for (; s < S; ++s) s0 += partial[s * n + i];
  out[i] = (s0 + s1) + (s2 + s3);
}

constexpr size_t kSmemBytes = 4 * (kKC / 4) * (kBM + 1) * sizeof(float4);   // 4 planes (A hi/lo, B hi/lo) = 64.5 KB

inline int tn_splits(int64_t M, int K, int U) {
  const int tiles = ceil_div(K, kBM) * ceil_div(U, kMaxN);
  int64_t s = std::max<int64_t>(1, (2 * kNumSM) / tiles);
  s = std::min<int64_t>(s, std::max<int64_t>(1, ceil_div<int64_t>(M, 4 * kKC)));
  return static_cast<int>(s);
}

}  // namespace tc
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

size_t kgcnn_dense_tc_workspace_bytes(int op, int64_t M, int K, int U) {
  if (op != tc::kTN || M <= 0 || K <= 0 || U <= 0) return 0;
  return static_cast<size_t>(tc::tn_splits(M, K, U)) * K * U * sizeof(float) + 256;
}

int kgcnn_dense_tc(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias, float* out,
                   void* workspace, size_t workspace_bytes, kgcnn_stream_t stream) {
  if (op < tc::kNN || op > tc::kTN || M < 0 || K <= 0 || U <= 0 || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (M == 0) {
    if (op == tc::kTN && cudaMemsetAsync(out, 0, sizeof(float) * K * U, st) != cudaSuccess) return KGCNN_E_CUDA;
    return KGCNN_OK;
  }
  if (!A || !B) return KGCNN_E_INVALID;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kNN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kNT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kTN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    attr_set = true;
  }
  if (op == tc::kNN) {
    dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(M, tc::kBM)), ceil_div(U, tc::kMaxN));
    tc::dense_tc_kernel<tc::kNN><<<grid, tc::kThreads, tc::kSmemBytes, st>>>(
        A, B, M, K, U, bias, out, 0, (K % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0);
    return finish_launch();
  }
  if (op == tc::kNT) {
    dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(M, tc::kBM)), ceil_div(K, tc::kMaxN));
    tc::dense_tc_kernel<tc::kNT><<<grid, tc::kThreads, tc::kSmemBytes, st>>>(
        A, B, M, K, U, nullptr, out, 0, (U % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0);
    return finish_launch();
  }
  const int S = tc::tn_splits(M, K, U);
  if (!workspace || workspace_bytes < kgcnn_dense_tc_workspace_bytes(op, M, K, U)) return KGCNN_E_WORKSPACE;
  const int64_t rows_per_split = ceil_div<int64_t>(ceil_div<int64_t>(M, S), tc::kKC) * tc::kKC;
  float* partial = static_cast<float*>(workspace);
  dim3 grid(ceil_div(K, tc::kBM), ceil_div(U, tc::kMaxN), S);
  tc::dense_tc_kernel<tc::kTN><<<grid, tc::kThreads, tc::kSmemBytes, st>>>(
      A, B, M, K, U, nullptr, partial, rows_per_split, (K % 4 == 0 && al16(A)) ? 1 : 0,
      (U % 4 == 0 && al16(B)) ? 1 : 0);
  const int64_t n = static_cast<int64_t>(K) * U;
  tc::tn_split_reduce_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(n, 256)), 256, 0, st>>>(partial, S, n, out);
  return finish_launch(2);
}

}  // extern "C"
