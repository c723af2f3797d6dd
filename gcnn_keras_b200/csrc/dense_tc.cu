// Dense contractions of the edge/node feature rows on the 5th-generation tensor cores (tcgen05 + TMEM):
//     NN : Y[M,U]  = X[M,K]  W[K,U] (+ bias)        forward of kgcnn/layers/modules.py:85
//     NT : dX[M,K] = G[M,U]  W[K,U]^T               input gradient
//     TN : dW[K,U] = X[M,K]^T G[M,U]                weight gradient (split over M, fixed-order reduce)
// M = edges / atoms of the batch (1e4 .. 1e6), K and U = layer widths (1 .. 640).
//
// Precision: the parity bar is 1e-5 of an fp64 oracle, single-pass TF32 (10-bit mantissa) is ~1e-3.  Every operand
// tile is therefore split while it is staged into shared memory, x ~ hi + lo with hi = nearest TF32 value and
// lo = nearest TF32 value of x - hi (|x - hi - lo| <= 2^-23 |x|), and three MMAs accumulate lo*hi + hi*lo + hi*hi
// (3xTF32; lo*lo is 2^-22 relative, below the fp32 rounding of the result) in an fp32 TMEM accumulator that is
// drained to fp32 registers every kFlush reduction elements.
//
// Data path: these GEMMs are skinny (K, U <= 640) and sit where tensor time ~ memory time, and the split needs a
// register pass anyway, so tiles go global -> registers -> (hi, lo) -> shared memory in the canonical no-swizzle
// K-major UMMA layout (8 x 16 B core matrices); operands whose reduction index is the ROW of the global matrix (W in
// NN, X and G in TN) are transposed on the way into shared memory, so every descriptor is K-major; no TMA descriptor
// is needed.  The loop over 32-element reduction chunks is a 3-stage software pipeline: while one elected thread's
// tcgen05.mma (kind::tf32, cta_group::1, M = 128) of chunk c run (tcgen05.commit -> one mbarrier per stage), the
// CTA splits and stores chunk c+1, has the global loads of chunk c+2 in flight and drains the previous flush group
// from the second TMEM accumulator with tcgen05.ld.
#include <algorithm>

#include "activation.cuh"
#include "common.cuh"

namespace kgcnn {
namespace tc {

constexpr int kThreads = 512;   // 16 warps stage operands (4 per scheduler: the staging code is issue-latency bound);
                                // warp w reads TMEM lanes 32*(w%4)..+31, columns 32*(w/4)..+31 in the drain / epilogue
constexpr int kCtaThreads = kThreads + 32;   // + warp 8: waits for full stages and issues the MMAs, nothing else
constexpr int kBM = 128;        // MMA M
constexpr int kKC = 32;         // reduction elements staged per chunk (4 MMA k-steps of 8)
constexpr int kLd = 2;          // float4 per thread, operand and register set (= 128*8/512 = 32*32/512)
constexpr int kStages = 3;      // shared-memory stages of (A hi/lo, B hi/lo) planes: 3 x 66 KB
constexpr int kMaxN = 128;      // MMA N per CTA (TMEM columns)
constexpr int kFlush = 64;      // reduction length accumulated inside the tensor core before an fp32 (RN) flush

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
// nearest TF32 value (round half away in magnitude): |x - hi| <= 2^-11 |x|, and x - hi is exact in fp32.  lo is
// rounded the same way (the tensor core would truncate it, a one-sided 2^-22 |x| bias that adds up coherently over
// a reduction); after RN the residual x - hi - lo is <= 2^-23 |x| and unbiased, i.e. fp32-like.
__device__ __forceinline__ float tf32_hi(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

// mn = row of the global matrix, k = feature: rows [row0, row0+128), features [f0, f0+kKC).  Loads are issued as
// one batch of kLd independent float4 per thread (registers), the split + shared-memory stores follow later, so
// the next chunk's global latency overlaps the MMAs of the current one.
// `fast` (CTA-uniform): the whole tile is in bounds and 16-byte aligned -> plain 128-bit loads, no predicates.
__device__ __forceinline__ void load_direct(float4 (&v)[kLd], const float* __restrict__ g, int64_t row0,
                                            int64_t row_end, int f0, int f_end, int ld, bool vec_ok, bool fast) {
  constexpr int NCH = kKC / 4;
#pragma unroll
  for (int i = 0; i < kLd; ++i) {
    const int it = threadIdx.x + i * kThreads;
    const int c = it % NCH, r = it / NCH;                 // consecutive threads: consecutive 16 B of one row
    if (fast) v[i] = __ldg(reinterpret_cast<const float4*>(g + (row0 + r) * ld + f0 + 4 * c));
    else v[i] = load4(g, row0 + r, row_end, f0 + 4 * c, f_end, ld, vec_ok);
  }
}
__device__ __forceinline__ void store_direct(const float4 (&v)[kLd], float4* __restrict__ hi, float4* __restrict__ lo) {
  constexpr int NRP = kBM + 1, NCH = kKC / 4;
#pragma unroll
  for (int i = 0; i < kLd; ++i) {
    const int it = threadIdx.x + i * kThreads;
    const int c = it % NCH, r = it / NCH;
    const float4 h = make_float4(tf32_hi(v[i].x), tf32_hi(v[i].y), tf32_hi(v[i].z), tf32_hi(v[i].w));
    hi[c * NRP + r] = h;
    lo[c * NRP + r] = make_float4(tf32_hi(v[i].x - h.x), tf32_hi(v[i].y - h.y), tf32_hi(v[i].z - h.z),
                                  tf32_hi(v[i].w - h.w));
  }
}

// mn = feature, k = row of the global matrix (transposed on the way in): rows [row0, row0+kKC), features
// [f0, f0+4*nchunks).  Thread -> (row k, 4-feature group j): within a warp 16 rows x 2 adjacent groups, so a load
// instruction reads one full 32-byte sector per row and the scalar stores hit 32 different banks; register i is
// the 16-row block, the warp selects the pair of feature groups.
__device__ __forceinline__ void transposed_coords(int i, int& k, int& j) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  k = 16 * i + (lane & 15);
  j = 2 * warp + (lane >> 4);
}
static_assert(kLd * 16 == kKC && 2 * (kThreads / 32) == kBM / 4, "transposed staging map covers a 32 x 128 chunk");
__device__ __forceinline__ void load_transposed(float4 (&v)[kLd], const float* __restrict__ g, int64_t row0,
                                                int64_t row_end, int f0, int f_end, int ld, int nchunks, bool vec_ok,
                                                bool fast) {
#pragma unroll
  for (int i = 0; i < kLd; ++i) {
    int k, j;
    transposed_coords(i, k, j);
    if (fast) v[i] = __ldg(reinterpret_cast<const float4*>(g + (row0 + k) * ld + f0 + 4 * j));
    else v[i] = (j < nchunks) ? load4(g, row0 + k, row_end, f0 + 4 * j, f_end, ld, vec_ok)
                              : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
__device__ __forceinline__ void store_transposed(const float4 (&v)[kLd], int nchunks, float4* __restrict__ hi,
                                                 float4* __restrict__ lo) {
  constexpr int NRP = kBM + 1;
  float* hs = reinterpret_cast<float*>(hi);
  float* ls = reinterpret_cast<float*>(lo);
#pragma unroll
  for (int i = 0; i < kLd; ++i) {
    int k, j;
    transposed_coords(i, k, j);
    if (j >= nchunks) continue;
    const float vv[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
    const int base = ((k >> 2) * NRP + 4 * j) * 4 + (k & 3);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float h = tf32_hi(vv[q]);
      hs[base + 4 * q] = h;
      ls[base + 4 * q] = tf32_hi(vv[q] - h);
    }
  }
}

// One CTA = one 128-row output tile (NN / NT) or one (128-feature tile, split of the rows) (TN).
// (bx, by, bz) are the tile coordinates (blockIdx of the single-problem kernel; looked up per problem by the grouped
// TN kernel).  colsum_part (TN only, nullable): this split's column sums of G, written by the a0 == 0 tile.
template <int OP>
__device__ __forceinline__ void
dense_tc_body(const float* __restrict__ A, const float* __restrict__ B, int64_t M, int K, int U,
              const float* __restrict__ bias, float* __restrict__ out, int64_t rows_per_split, int vec_a,
              int vec_b, int bx, int by, int bz, float* __restrict__ colsum_part,
              float* __restrict__ act_out = nullptr, int act = 0) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t mbar[kStages];
  __shared__ __align__(8) uint64_t full[kStages];
  __shared__ uint32_t tmem_base_s;
  constexpr int PLANE = (kKC / 4) * (kBM + 1);     // float4 slots per plane: 8 k-chunks x (128 + 1) mn
  float4* planes = reinterpret_cast<float4*>(smem_raw);          // [stage][a_hi, a_lo, b_hi, b_lo][PLANE]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // output-tile coordinates
  //   NN: D[rows m0.., cols n0..] = sum_k X[m,k] W[k,n]        reduction over K features of X / rows of W
  //   NT: D[rows m0.., cols n0..] = sum_u G[m,u] W[n,u]        reduction over U; n indexes rows of W (K of them)
  //   TN: D[feat a0.., cols n0..] = sum_rows X[row,a] G[row,n] reduction over the rows of this split
  const int n_total = (OP == kNN) ? U : (OP == kNT ? K : U);
  const int n0 = by * kMaxN;
  const int nt = min(kMaxN, ((n_total - n0 + 15) / 16) * 16);        // MMA N (multiple of 16, zero padded)
  const int64_t m0 = (OP == kTN) ? 0 : static_cast<int64_t>(bx) * kBM;
  const int a0 = (OP == kTN) ? bx * kBM : 0;
  const int64_t row_beg = (OP == kTN) ? static_cast<int64_t>(bz) * rows_per_split : 0;
  const int64_t row_end = (OP == kTN) ? min(M, row_beg + rows_per_split) : 0;
  const int64_t red_len = (OP == kNN) ? K : (OP == kNT ? U : (row_end > row_beg ? row_end - row_beg : 0));

  if (warp == kThreads / 32) {                                   // two accumulators of kMaxN columns each
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(static_cast<uint32_t>(2 * kMaxN))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
#pragma unroll
    for (int st = 0; st < kStages; ++st) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[st])) : "memory");       // MMAs done
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[st])), "r"(kThreads / 32)
                   : "memory");                                                                        // stage staged
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_s;

  // every operand is K-major with 128 (+1) mn slots per k-chunk: LBO = 129, SBO = 8 (units of 16 B)
  constexpr int NRP = kBM + 1;
  static_assert(kMaxN == kBM, "A and B planes share one pitch");
  constexpr uint32_t kstep = 2 * NRP;                          // descriptor advance per MMA k-step (8 elements)
  const uint32_t idesc = make_idesc(0, 0, nt);

  // two register sets: the global loads of chunk c+2 are issued as soon as chunk c has been stored, so every load
  // has two full pipeline iterations to land (the kernel runs one CTA per SM; registers are not the limit)
  float4 va0[kLd], vb0[kLd], va1[kLd], vb1[kLd];
  // interior tiles (every mn index in range, 16-byte aligned rows) take the unpredicated load path
  const bool in_a = vec_a && ((OP == kTN) ? (a0 + kBM <= K) : (m0 + kBM <= M));
  const bool in_b = vec_b && ((OP == kNT) ? (n0 + kMaxN <= K) : (n0 + kMaxN <= U)) && nt == kMaxN;
  auto load_chunk = [&](int64_t k0, float4 (&va)[kLd], float4 (&vb)[kLd]) {
    const bool in_k = k0 + kKC <= red_len;
    if (OP == kNN) {
      load_direct(va, A, m0, M, static_cast<int>(k0), K, K, vec_a, in_a && in_k);          // X rows, K feats
      load_transposed(vb, B, k0, K, n0, U, U, nt / 4, vec_b, in_b && in_k);                // W rows k -> K dim
    } else if (OP == kNT) {
      load_direct(va, A, m0, M, static_cast<int>(k0), U, U, vec_a, in_a && in_k);          // G rows, U feats
      load_direct(vb, B, n0, K, static_cast<int>(k0), U, U, vec_b, in_b && in_k);          // W rows n, U feats
    } else {
      load_transposed(va, A, row_beg + k0, row_end, a0, K, K, kBM / 4, vec_a, in_a && in_k);   // X rows -> K dim
      load_transposed(vb, B, row_beg + k0, row_end, n0, U, U, nt / 4, vec_b, in_b && in_k);    // G rows -> K dim
    }
  };
  // The tensor core adds into the TMEM accumulator with truncation, which biases long reductions (measured:
  // ~7e-9 relative per reduction element).  Every kFlush elements the accumulator is drained into fp32 registers
  // (round-to-nearest adds) and the next MMAs start a fresh accumulation in the OTHER accumulator, so the drain of
  // group g overlaps the MMAs of group g + 1.
  float racc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) racc[i] = 0.f;
  const int c0 = 32 * (warp >> 2);                       // this warp's 32 accumulator columns (staging warps only)
  auto drain = [&](uint32_t acc_col) {
    if (c0 < nt) {                                       // warp-uniform
      uint32_t v[32];
      const uint32_t taddr = tmem_d + (static_cast<uint32_t>(32 * (warp & 3)) << 16) + acc_col +
                             static_cast<uint32_t>(c0);
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
#pragma unroll
      for (int i = 0; i < 32; ++i) racc[i] += __uint_as_float(v[i]);
    }
  };
  const bool do_colsum = (OP == kTN) && colsum_part != nullptr && a0 == 0;     // CTA-uniform
  float4 csum4[kLd];                                     // bias gradient: this thread's (row lane, 4 features) sums
#pragma unroll
  for (int i = 0; i < kLd; ++i) csum4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  // Warp-specialised software pipeline over chunks c of kKC reduction elements, kStages shared-memory stages:
  //   warps 0-7 (staging):  [wait done(c - kStages)]  split+store(c)  load(c+2)  arrive full(c)  [drain group of c-1]
  //   warp 8   (MMA issue): wait full(c)  tcgen05.mma x 3 x ksteps -> accumulator (group & 1)  commit -> done(c)
  // No block-wide barrier inside the loop; two TMEM accumulators alternate per flush group so that the drain of
  // group g (tcgen05.ld by the staging warps) overlaps the MMAs of group g + 1.
  const int nchunk = static_cast<int>((red_len + kKC - 1) / kKC);
  constexpr int kPerGroup = kFlush / kKC;
  auto wait_chunk = [&](int c) { mbar_wait(smem_u32(&mbar[c % kStages]), static_cast<uint32_t>((c / kStages) & 1)); };
  if (warp == kThreads / 32) {
    // ------------------------------- MMA-issue warp -------------------------------
    const uint64_t desc_hi = (static_cast<uint64_t>(NRP & 0x3FFF) << 16) | (static_cast<uint64_t>(8) << 32) | (1ull << 46);
    for (int c = 0; c < nchunk; ++c) {
      const int st = c % kStages;
      mbar_wait(smem_u32(&full[st]), static_cast<uint32_t>((c / kStages) & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (lane == 0) {
        const uint32_t base = smem_u32(planes + static_cast<size_t>(st) * 4 * PLANE);
        // descriptors differ only in the 14-bit start-address field (16-byte units): + kstep per MMA k-step
        const uint64_t dah0 = desc_hi | ((base >> 4) & 0x3FFF);
        const uint64_t dal0 = dah0 + PLANE, dbh0 = dah0 + 2 * PLANE, dbl0 = dah0 + 3 * PLANE;
        const int64_t left = red_len - static_cast<int64_t>(c) * kKC;
        const int ksteps = (static_cast<int>(left < kKC ? left : kKC) + 7) / 8;           // skip all-zero k-steps
        const bool fresh = (c % kPerGroup) == 0;          // first chunk of a flush group overwrites the accumulator
        const uint32_t acc = tmem_d + static_cast<uint32_t>(((c / kPerGroup) & 1) * kMaxN);
        // small terms first: while they accumulate, the accumulator is ~2^-11 of its final size and so are the
        // truncation errors of the tensor core's adds.  lo*lo (2^-22 relative) is below the fp32 rounding of the
        // result and is left out (3xTF32).
#pragma unroll
        for (int ks = 0; ks < kKC / 8; ++ks) {
          if (ks < ksteps) {
            mma_tf32(acc, dal0 + ks * kstep, dbh0 + ks * kstep, idesc, (fresh && ks == 0) ? 0u : 1u);   // lo * hi
            mma_tf32(acc, dah0 + ks * kstep, dbl0 + ks * kstep, idesc, 1u);                             // hi * lo
          }
        }
#pragma unroll
        for (int ks = 0; ks < kKC / 8; ++ks)
          if (ks < ksteps) mma_tf32(acc, dah0 + ks * kstep, dbh0 + ks * kstep, idesc, 1u);              // hi * hi
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                         smem_u32(&mbar[st]))
                     : "memory");
      }
      __syncwarp();
    }
  } else {
    // ------------------------------- staging warps -------------------------------
    if (nchunk > 0) load_chunk(0, va0, vb0);
    if (nchunk > 1) load_chunk(kKC, va1, vb1);
    auto step = [&](int c, float4 (&va)[kLd], float4 (&vb)[kLd]) {
      const int st = c % kStages;
      float4* a_hi = planes + static_cast<size_t>(st) * 4 * PLANE;
      float4* a_lo = a_hi + PLANE;
      float4* b_hi = a_lo + PLANE;
      float4* b_lo = b_hi + PLANE;
      if (c >= kStages) wait_chunk(c - kStages);           // the MMAs that read this stage are done (long ago)
      // ---- split (hi / lo) and store this chunk ----
      if (OP == kNN) { store_direct(va, a_hi, a_lo); store_transposed(vb, nt / 4, b_hi, b_lo); }
      else if (OP == kNT) { store_direct(va, a_hi, a_lo); store_direct(vb, b_hi, b_lo); }
      else { store_transposed(va, kBM / 4, a_hi, a_lo); store_transposed(vb, nt / 4, b_hi, b_lo); }
      if (OP == kTN && do_colsum) {
#pragma unroll
        for (int i = 0; i < kLd; ++i) {
          csum4[i].x += vb[i].x; csum4[i].y += vb[i].y; csum4[i].z += vb[i].z; csum4[i].w += vb[i].w;
        }
      }
      if (c + 2 < nchunk) load_chunk(static_cast<int64_t>(c + 2) * kKC, va, vb);   // refill this register set
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> async proxy (MMA)
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&full[st])) : "memory");
      // ---- drain the previous flush group while this chunk's MMAs run ----
      if (c > 0 && (c % kPerGroup) == 0) {
        wait_chunk(c - 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        drain(static_cast<uint32_t>((((c - 1) / kPerGroup) & 1) * kMaxN));
      }
    };
    for (int c = 0; c < nchunk; c += 2) {
      step(c, va0, vb0);
      if (c + 1 < nchunk) step(c + 1, va1, vb1);
    }
    if (nchunk > 0) {
      wait_chunk(nchunk - 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain(static_cast<uint32_t>((((nchunk - 1) / kPerGroup) & 1) * kMaxN));
    }
  }

  // ---- epilogue: TMEM lane 32*(warp%4) + lane = output row; warps 0-3 take the even 32-column groups, 4-7 the odd ----
  // The accumulator registers hold one output ROW per lane (TMEM lane = row).  Writing them straight out would make
  // every store instruction touch 32 different rows with 16 bytes each; instead the tile goes through shared memory
  // (the operand stages are free: the last MMAs have completed) and leaves as full 512-byte row segments.
  const int d_row = 32 * (warp & 3) + lane;
  constexpr int OP_PITCH = kMaxN + 4;                              // floats; +4 keeps float4 alignment, spreads banks
  float* otile = reinterpret_cast<float*>(smem_raw);               // [kBM][OP_PITCH] = 67.6 KB
  if (warp < kThreads / 32) {
    if (c0 < nt) {
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(otile + d_row * OP_PITCH + c0 + i) =
            make_float4(racc[i], racc[i + 1], racc[i + 2], racc[i + 3]);
    }
    asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");    // staging warps only
    const int64_t out_rows = (OP == kTN) ? K : M;                  // rows of the output matrix this tile maps to
    const int64_t row_base = (OP == kTN) ? a0 : m0;
    float* obase = (OP == kTN) ? out + static_cast<int64_t>(bz) * K * U : out;
    const bool vec_out = (n_total % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    const int col = n0 + 4 * lane;                                 // this lane's 4 columns of every row
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (OP == kNN && bias) {                                       // scalar loads: the bias may be an unaligned view
      if (col + 0 < n_total) bv.x = __ldg(bias + col + 0);
      if (col + 1 < n_total) bv.y = __ldg(bias + col + 1);
      if (col + 2 < n_total) bv.z = __ldg(bias + col + 2);
      if (col + 3 < n_total) bv.w = __ldg(bias + col + 3);
    }
    for (int r = warp; r < kBM; r += kThreads / 32) {              // one full row segment per warp instruction
      const int64_t row = row_base + r;
      if (row >= out_rows || 4 * lane >= nt) continue;
      float4 v = *reinterpret_cast<const float4*>(otile + r * OP_PITCH + 4 * lane);
      v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
      float* o = obase + row * n_total + col;
      if (vec_out && col + 3 < n_total) *reinterpret_cast<float4*>(o) = v;
      else {
        if (col + 0 < n_total) o[0] = v.x;
        if (col + 1 < n_total) o[1] = v.y;
        if (col + 2 < n_total) o[2] = v.z;
        if (col + 3 < n_total) o[3] = v.w;
      }
      if (OP == kNN && act_out) {          // Dense + activation in one pass: pre (kept for the derivatives) and f(pre)
        const float4 f = make_float4(act_value(act, v.x), act_value(act, v.y), act_value(act, v.z),
                                     act_value(act, v.w));
        float* q = act_out + row * n_total + col;
        if (vec_out && ((reinterpret_cast<uintptr_t>(act_out) & 15) == 0) && col + 3 < n_total)
          *reinterpret_cast<float4*>(q) = f;
        else {
          if (col + 0 < n_total) q[0] = f.x;
          if (col + 1 < n_total) q[1] = f.y;
          if (col + 2 < n_total) q[2] = f.z;
          if (col + 3 < n_total) q[3] = f.w;
        }
      }
    }
  }
  if (OP == kTN && do_colsum && warp < kThreads / 32) {
    // registers i = the two 16-row blocks of a chunk; lanes (l & 15) = rows; (2 warp + (l >> 4)) = 4-feature group:
    // fixed order (i, then xor tree over 16 lanes) -> deterministic
    float4 t = csum4[0];
#pragma unroll
    for (int i = 1; i < kLd; ++i) { t.x += csum4[i].x; t.y += csum4[i].y; t.z += csum4[i].z; t.w += csum4[i].w; }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      t.x += __shfl_xor_sync(0xffffffffu, t.x, o); t.y += __shfl_xor_sync(0xffffffffu, t.y, o);
      t.z += __shfl_xor_sync(0xffffffffu, t.z, o); t.w += __shfl_xor_sync(0xffffffffu, t.w, o);
    }
    const int j = 2 * warp + (lane >> 4);
    if ((lane & 15) == 0 && 4 * j < nt) {
      const int col = n0 + 4 * j;
      float* dst = colsum_part + static_cast<int64_t>(bz) * U;
      if (col + 0 < U) dst[col + 0] = t.x;
      if (col + 1 < U) dst[col + 1] = t.y;
      if (col + 2 < U) dst[col + 2] = t.z;
      if (col + 3 < U) dst[col + 3] = t.w;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kThreads / 32) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                 "r"(static_cast<uint32_t>(2 * kMaxN))
                 : "memory");
  }
}

template <int OP>
__global__ void __launch_bounds__(kCtaThreads)
dense_tc_kernel(const float* __restrict__ A, const float* __restrict__ B, int64_t M, int K, int U,
                const float* __restrict__ bias, float* __restrict__ out, int64_t rows_per_split, int vec_a,
                int vec_b, float* __restrict__ act_out, int act) {
  dense_tc_body<OP>(A, B, M, K, U, bias, out, rows_per_split, vec_a, vec_b, blockIdx.x, blockIdx.y, blockIdx.z,
                    nullptr, act_out, act);
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

// ---------------------------------------------------------------------------------------------------------
// Grouped weight gradients: all (dW, db) += (X^T G, colsum G) products of one backward sweep in TWO launches.
// The problems are independent leaves of the backward graph, so the host defers them and hands over the list;
// CTA -> (problem, feature tile, column tile, row split) through a prefix table passed BY VALUE (baked into a CUDA
// graph node; nothing is read from host memory at replay).  Partials go to the workspace and are summed in a
// fixed order (deterministic); several problems may target the same dW (a weight used by the forward Linear and by
// the force's g W^T): the reduce kernel walks outputs, not problems, so there is no write race.
// ---------------------------------------------------------------------------------------------------------
constexpr int kMaxGroup = 64;

struct TnProblem {          // 64 bytes
  const float* X;           // [M, K]
  const float* G;           // [M, U]
  int64_t M;
  int64_t ws_off;           // float offset of this problem's partials: [splits][K*U] then [splits][U]
  int32_t K, U;
  int32_t splits;
  int32_t rows_per_split;
  int32_t cta_begin;        // first CTA of this problem in the grouped grid
  int32_t tiles_k, tiles_u;
  int32_t want_colsum;
};
struct TnOutput {           // 40 bytes: one (dW, db) pair = a contiguous run of problems
  float* dW;                // [K, U], accumulated into
  float* db;                // [U] or null
  int64_t elem_begin;       // first element of this output in the reduce kernel's index space
  int32_t first, count;     // problems [first, first + count)
  int32_t K, U;
};
struct TnGroup {
  int32_t n_problems, n_outputs, total_ctas, pad;
  int64_t total_elems;
  TnProblem p[kMaxGroup];
  TnOutput o[kMaxGroup];
};

__global__ void __launch_bounds__(kCtaThreads)
tn_grouped_kernel(const __grid_constant__ TnGroup grp, float* __restrict__ ws) {
  __shared__ int which;
  if (threadIdx.x == 0) {
    int lo = 0, hi = grp.n_problems - 1;                  // last problem with cta_begin <= blockIdx.x
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (grp.p[mid].cta_begin <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    which = lo;
  }
  __syncthreads();
  const TnProblem& pr = grp.p[which];
  int t = static_cast<int>(blockIdx.x) - pr.cta_begin;
  const int bx = t % pr.tiles_k; t /= pr.tiles_k;
  const int by = t % pr.tiles_u; t /= pr.tiles_u;
  const int bz = t;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  float* partial = ws + pr.ws_off;
  float* cpart = pr.want_colsum ? partial + static_cast<int64_t>(pr.splits) * pr.K * pr.U : nullptr;
  dense_tc_body<kTN>(pr.X, pr.G, pr.M, pr.K, pr.U, nullptr, partial, pr.rows_per_split,
                     (pr.K % 4 == 0 && al16(pr.X)) ? 1 : 0, (pr.U % 4 == 0 && al16(pr.G)) ? 1 : 0, bx, by, bz, cpart);
}

__global__ void __launch_bounds__(256)
tn_grouped_reduce_kernel(const __grid_constant__ TnGroup grp, const float* __restrict__ ws) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= grp.total_elems) return;
  int lo = 0, hi = grp.n_outputs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (grp.o[mid].elem_begin <= i) lo = mid; else hi = mid - 1;
  }
  const TnOutput& o = grp.o[lo];
  const int64_t e = i - o.elem_begin, kw = static_cast<int64_t>(o.K) * o.U;
  const bool is_w = e < kw;
  float* dst = is_w ? o.dW + e : o.db + (e - kw);
  float acc = *dst;
  for (int q = o.first; q < o.first + o.count; ++q) {     // fixed order: problems, then splits
    const TnProblem& pr = grp.p[q];
    if (!is_w && !pr.want_colsum) continue;
    const float* part = ws + pr.ws_off + (is_w ? e : static_cast<int64_t>(pr.splits) * kw + (e - kw));
    const int64_t stride = is_w ? kw : o.U;
    float s0 = 0.f, s1 = 0.f;
    int sp = 0;
    for (; sp + 1 < pr.splits; sp += 2) { s0 += part[sp * stride]; s1 += part[(sp + 1) * stride]; }
    if (sp < pr.splits) s0 += part[sp * stride];
    acc += s0 + s1;
  }
  *dst = acc;
}

constexpr size_t kSmemBytes = kStages * 4 * (kKC / 4) * (kBM + 1) * sizeof(float4);   // 3 stages x 4 planes x 16.5 KB

inline int tn_splits(int64_t M, int K, int U) {
  const int tiles = ceil_div(K, kBM) * ceil_div(U, kMaxN);
  int64_t s = std::max<int64_t>(1, (2 * kNumSM) / tiles);
  s = std::min<int64_t>(s, std::max<int64_t>(1, ceil_div<int64_t>(M, 4 * kKC)));
  return static_cast<int>(s);
}

// rows per split for a grouped launch: ~2 waves of one resident CTA per SM (the kernel needs ~230 registers per
// thread) over all problems, at least 2 chunks each
inline int64_t grouped_rows_per_split(int n, const int64_t* M, const int32_t* K, const int32_t* U) {
  int64_t chunk_tiles = 0;
  for (int i = 0; i < n; ++i)
    chunk_tiles += ceil_div<int64_t>(std::max<int64_t>(M[i], 1), kKC) * ceil_div(K[i], kBM) * ceil_div(U[i], kMaxN);
  const int64_t target = 2LL * kNumSM;
  // (a search that grows the split until the launch fits two full waves was measured: PaiNN C3 -0.5 %, SchNet C2
  //  +14 % on this kernel because its many small problems then share fewer, longer CTAs -- not kept)
  const int64_t per_cta = std::max<int64_t>(4, ceil_div<int64_t>(chunk_tiles, target));
  return per_cta * kKC;
}

}  // namespace tc
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

size_t kgcnn_dense_tc_workspace_bytes(int op, int64_t M, int K, int U) {
  if (op != tc::kTN || M <= 0 || K <= 0 || U <= 0) return 0;
  return static_cast<size_t>(tc::tn_splits(M, K, U)) * K * U * sizeof(float) + 256;
}

static int dense_tc_launch(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias,
                           float* out, void* workspace, size_t workspace_bytes, kgcnn_stream_t stream,
                           float* act_out, int act) {
  if (op < tc::kNN || op > tc::kTN || M < 0 || K <= 0 || U <= 0 || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (M == 0) {
    if (op == tc::kTN && cudaMemsetAsync(out, 0, sizeof(float) * K * U, st) != cudaSuccess) return KGCNN_E_CUDA;
    return KGCNN_OK;
  }
  if (!A || !B) return KGCNN_E_INVALID;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  static PerDeviceOnce attr;
  attr.run([] {
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kNN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kNT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_tc_kernel<tc::kTN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
  });
  if (op == tc::kNN) {
    dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(M, tc::kBM)), ceil_div(U, tc::kMaxN));
    tc::dense_tc_kernel<tc::kNN><<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
        A, B, M, K, U, bias, out, 0, (K % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0, act_out, act);
    return finish_launch();
  }
  if (op == tc::kNT) {
    dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(M, tc::kBM)), ceil_div(K, tc::kMaxN));
    tc::dense_tc_kernel<tc::kNT><<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
        A, B, M, K, U, nullptr, out, 0, (U % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0, nullptr,
        0);
    return finish_launch();
  }
  const int S = tc::tn_splits(M, K, U);
  if (!workspace || workspace_bytes < kgcnn_dense_tc_workspace_bytes(op, M, K, U)) return KGCNN_E_WORKSPACE;
  const int64_t rows_per_split = ceil_div<int64_t>(ceil_div<int64_t>(M, S), tc::kKC) * tc::kKC;
  float* partial = static_cast<float*>(workspace);
  dim3 grid(ceil_div(K, tc::kBM), ceil_div(U, tc::kMaxN), S);
  tc::dense_tc_kernel<tc::kTN><<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
      A, B, M, K, U, nullptr, partial, rows_per_split, (K % 4 == 0 && al16(A)) ? 1 : 0,
      (U % 4 == 0 && al16(B)) ? 1 : 0, nullptr, 0);
  const int64_t n = static_cast<int64_t>(K) * U;
  tc::tn_split_reduce_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(n, 256)), 256, 0, st>>>(partial, S, n, out);
  return finish_launch(2);
}

int kgcnn_dense_tc(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias, float* out,
                   void* workspace, size_t workspace_bytes, kgcnn_stream_t stream) {
  return dense_tc_launch(op, A, B, M, K, U, bias, out, workspace, workspace_bytes, stream, nullptr, 0);
}

int kgcnn_dense_act(const float* X, const float* W, int64_t M, int K, int U, const float* bias, int act, float* pre,
                    float* y, kgcnn_stream_t stream) {
  if (act < KGCNN_ACT_LINEAR || act > KGCNN_ACT_LEAKY_RELU || !pre || !y) return KGCNN_E_INVALID;
  return dense_tc_launch(tc::kNN, X, W, M, K, U, bias, pre, nullptr, 0, stream, y, act);
}

size_t kgcnn_dense_tn_grouped_workspace_bytes(int n, const int64_t* M, const int32_t* K, const int32_t* U) {
  if (n <= 0 || !M || !K || !U) return 0;
  const int64_t rps = tc::grouped_rows_per_split(n, M, K, U);
  size_t floats = 0;
  for (int i = 0; i < n; ++i) {
    const int64_t splits = std::max<int64_t>(1, ceil_div<int64_t>(M[i], rps));
    floats += static_cast<size_t>(splits) * (static_cast<size_t>(K[i]) * U[i] + U[i]);
    floats = (floats + 63) / 64 * 64;
  }
  return floats * sizeof(float) + 256;
}

int kgcnn_dense_tn_grouped(int n, const float* const* X, const float* const* G, const int64_t* M, const int32_t* K,
                           const int32_t* U, float* const* dW, float* const* db, void* workspace,
                           size_t workspace_bytes, kgcnn_stream_t stream) {
  if (n < 0) return KGCNN_E_INVALID;
  if (n == 0) return KGCNN_OK;
  if (!X || !G || !M || !K || !U || !dW || !db) return KGCNN_E_INVALID;
  if (!workspace || workspace_bytes < kgcnn_dense_tn_grouped_workspace_bytes(n, M, K, U)) return KGCNN_E_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return KGCNN_E_INVALID;
  for (int i = 0; i < n; ++i)
    if (M[i] < 0 || K[i] <= 0 || U[i] <= 0 || !dW[i] || (M[i] > 0 && (!X[i] || !G[i]))) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  static PerDeviceOnce attr;
  attr.run([] {
    cudaFuncSetAttribute(tc::tn_grouped_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
  });
  const int64_t rps = tc::grouped_rows_per_split(n, M, K, U);
  float* ws = static_cast<float*>(workspace);
  int64_t ws_off = 0;
  int launches = 0;
  // Problems are packed into launches greedily.  Within one launch every dW (and its db) belongs to exactly ONE
  // output run of adjacent problems -- the reduce kernel walks outputs, so a second run on the same dW would be a
  // read-modify-write race; such a problem (same dW not adjacent, or a different non-null db) starts the next launch.
  int i = 0;
  while (i < n) {
    tc::TnGroup grp;
    grp.n_problems = 0; grp.n_outputs = 0; grp.total_ctas = 0; grp.pad = 0; grp.total_elems = 0;
    for (; i < n && grp.n_problems < tc::kMaxGroup; ++i) {
      if (M[i] == 0) continue;                                     // contributes nothing
      int hit = -1;
      for (int o = 0; o < grp.n_outputs; ++o)
        if (grp.o[o].dW == dW[i]) hit = o;
      bool merge = false;
      if (hit >= 0) {
        const tc::TnOutput& o = grp.o[hit];
        merge = hit == grp.n_outputs - 1 && o.K == K[i] && o.U == U[i] && (o.db == db[i] || !db[i] || !o.db);
        if (!merge) break;                                         // conflicting destination: next launch
      }
      tc::TnProblem& pr = grp.p[grp.n_problems];
      pr.X = X[i]; pr.G = G[i]; pr.M = M[i]; pr.K = K[i]; pr.U = U[i];
      pr.splits = static_cast<int32_t>(std::max<int64_t>(1, ceil_div<int64_t>(M[i], rps)));
      pr.rows_per_split = static_cast<int32_t>(rps);
      pr.tiles_k = ceil_div(K[i], tc::kBM); pr.tiles_u = ceil_div(U[i], tc::kMaxN);
      pr.cta_begin = grp.total_ctas;
      pr.want_colsum = db[i] ? 1 : 0;
      pr.ws_off = ws_off;
      grp.total_ctas += pr.tiles_k * pr.tiles_u * pr.splits;
      int64_t fl = static_cast<int64_t>(pr.splits) * (static_cast<int64_t>(K[i]) * U[i] + U[i]);
      ws_off += (fl + 63) / 64 * 64;
      if (merge) {
        tc::TnOutput& o = grp.o[grp.n_outputs - 1];
        o.count += 1;
        if (!o.db && db[i]) { o.db = db[i]; grp.total_elems += U[i]; }   // it is the last output: no later shift
      } else {
        tc::TnOutput& o = grp.o[grp.n_outputs++];
        o.dW = dW[i]; o.db = db[i]; o.K = K[i]; o.U = U[i];
        o.first = grp.n_problems; o.count = 1; o.elem_begin = grp.total_elems;
        grp.total_elems += static_cast<int64_t>(K[i]) * U[i] + (db[i] ? U[i] : 0);
      }
      grp.n_problems += 1;
    }
    if (grp.n_problems == 0) continue;
    tc::tn_grouped_kernel<<<static_cast<unsigned>(grp.total_ctas), tc::kCtaThreads, tc::kSmemBytes, st>>>(grp, ws);
    tc::tn_grouped_reduce_kernel<<<static_cast<unsigned>(ceil_div<int64_t>(grp.total_elems, 256)), 256, 0, st>>>(grp, ws);
    launches += 2;
  }
  return launches ? finish_launch(launches) : KGCNN_OK;
}

}  // extern "C"
