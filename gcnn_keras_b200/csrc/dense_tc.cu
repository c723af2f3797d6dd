// Dense contractions of the edge/node feature rows on the 5th-generation tensor cores (tcgen05 + TMEM):
//     NN : Y[M,U]  = X[M,K]  W[K,U] (+ bias)        forward of kgcnn/layers/modules.py:85
//     NT : dX[M,K] = G[M,U]  W[K,U]^T               input gradient
//     TN : dW[K,U] = X[M,K]^T G[M,U]                weight gradient (split over M, fixed-order reduce)
// M = edges / atoms of the batch (1e4 .. 1e6), K and U = layer widths (1 .. 640).
//
// Precision: the parity bar is 1e-5 of an fp64 oracle through six interaction blocks and two backward sweeps, so the
// products must be fp32-grade.  Every operand element is split while it is staged into THREE bf16 pieces,
// x = b1 + b2 + b3 (each the round-to-nearest bf16 of the running remainder; the remainders are exact in fp32, so
// the three pieces carry 24 significant bits: the split is lossless up to 2^-25 |x|), and six kind::f16 MMAs per
// 16 reduction elements accumulate every product of weight >= 2^-16,
//     b3*b1' + b1*b3' + b2*b2'   (2^-16)      b2*b1' + b1*b2'   (2^-8)      b1*b1'   (1),
// small terms first, in an fp32 TMEM accumulator (bf16 x bf16 products are exact in fp32).  The dropped terms are
// 2^-24 relative.  (Round 1 used a 2-piece TF32 split: 22 bits, 3 MMAs per 8 elements = the same tensor time, but a
// representation error of 2^-22 per element that showed up as 1.1-1.6e-5 on PaiNN weight gradients at depth 6,
// profiles/parity_r2.json; the bf16 planes are also 25 % smaller: 6 instead of 8 bytes per element.)  The accumulator
// is drained into fp32 registers after every chunk of 32 reduction elements: the tensor core's accumulate truncates
// (measured ~2^-24 of the accumulator per MMA, one-sided), and with the small terms first only the two b1*b1' MMAs of
// a chunk add at full magnitude (a 64-element flush group measured 4e-7 per Dense, 10 of its 24 MMAs adding small
// terms into a full-size accumulator).
//
// Data path: tiles go global -> registers -> split -> shared memory as 16-byte units of 8 bf16 in the canonical
// NO-SWIZZLE UMMA layouts.  An operand whose reduction index runs along the global rows' feature axis (X in NN, G and
// W in NT) is K-major: unit = 8 reduction elements of one row.  An operand whose reduction index is the global ROW
// (W in NN, X and G in TN) is MN-major: unit = 8 features of one reduction row -- the instruction descriptor's
// a_major / b_major bits tell the tensor core, so nothing is transposed on the way in and every shared-memory store
// is a conflict-free 128-bit store.  kStages x 32 reduction elements are resident, i.e. for K <= 128 every global
// load of a tile is issued up front; one warp issues the MMAs (tcgen05.commit -> one mbarrier per stage), 16 warps
// stage and drain.
#include "dense_tc.cuh"

namespace kgcnn {
namespace tc {

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
  __shared__ float csum_s[kThreads / 32][4][8];
  uint4* planes = reinterpret_cast<uint4*>(smem_raw);            // [stage][a1, a2, a3, b1, b2, b3][kPlaneUnits]
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

  pdl_trigger();
  TL(0);
  if (warp == kThreads / 32) {                                   // four accumulators of kMaxN columns each
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(static_cast<uint32_t>(kTmemCols))
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

  // which layout each operand takes (see the header comment)
  constexpr bool kAMn = (OP == kTN);                 // A: X^T in TN
  constexpr bool kBMn = (OP == kNN) || (OP == kTN);  // B: W in NN, G in TN
  const uint32_t idesc = make_idesc(kAMn ? 1 : 0, kBMn ? 1 : 0, nt);

  // interior tiles (every index in range, 16-byte aligned rows) take the unpredicated load path
  const bool in_a = vec_a && ((OP == kTN) ? (a0 + kBM <= K) : (m0 + kBM <= M));
  const bool in_b = vec_b && ((OP == kNT) ? (n0 + kMaxN <= K) : (n0 + kMaxN <= U)) && nt == kMaxN;
  auto load_chunk = [&](int64_t k0, F8& va, F8& vb) {
    const bool in_k = k0 + kKC <= red_len;
    if (OP == kNN) {
      va = load_kmajor(A, m0, M, static_cast<int>(k0), K, K, vec_a, in_a && in_k);             // X rows, K feats
      vb = load_mnmajor(B, k0, K, n0, U, U, vec_b, in_b && in_k);                              // W rows = reduction
    } else if (OP == kNT) {
      va = load_kmajor(A, m0, M, static_cast<int>(k0), U, U, vec_a, in_a && in_k);             // G rows, U feats
      vb = load_kmajor(B, n0, K, static_cast<int>(k0), U, U, vec_b, in_b && in_k);             // W rows n, U feats
    } else {
      va = load_mnmajor(A, row_beg + k0, row_end, a0, K, K, vec_a, in_a && in_k);              // X rows = reduction
      vb = load_mnmajor(B, row_beg + k0, row_end, n0, U, U, vec_b, in_b && in_k);              // G rows = reduction
    }
  };
  auto load_chunk_a = [&](int64_t k0, F8& va) {
    const bool in_k = k0 + kKC <= red_len;
    if (OP == kTN) va = load_mnmajor(A, row_beg + k0, row_end, a0, K, K, vec_a, in_a && in_k);
    else va = load_kmajor(A, m0, M, static_cast<int>(k0), (OP == kNN) ? K : U, (OP == kNN) ? K : U, vec_a, in_a && in_k);
  };
  auto load_chunk_b = [&](int64_t k0, F8& vb) {
    const bool in_k = k0 + kKC <= red_len;
    if (OP == kNN) vb = load_mnmajor(B, k0, K, n0, U, U, vec_b, in_b && in_k);
    else if (OP == kNT) vb = load_kmajor(B, n0, K, static_cast<int>(k0), U, U, vec_b, in_b && in_k);
    else vb = load_mnmajor(B, row_beg + k0, row_end, n0, U, U, vec_b, in_b && in_k);
  };
  const int nchunk = static_cast<int>((red_len + kKC - 1) / kKC);
  // Short reductions (nchunk <= kStages, i.e. K <= 128: every Dense of SchNet / PaiNN) issue the loads of three A
  // chunks and two B chunks up front (40 registers in flight; the 96-register budget of a 544-thread CTA spills
  // beyond that) and the rest as soon as a register set is free, and drain their per-chunk accumulators once at the
  // end, so no fp32 accumulator registers are live while the loads are in flight; long reductions run two register
  // sets ahead of the stores.
  static_assert(kStages == 4 && kStages * kMaxN == kTmemCols, "four named register sets / accumulators below");
  const bool short_k = nchunk <= kStages;                // CTA-uniform
  F8 va0, vb0, va1, vb1, va2;
  pdl_wait();                                            // first global access below
  if (warp < kThreads / 32) {
    if (nchunk > 0) load_chunk(0, va0, vb0);
    if (nchunk > 1) load_chunk(kKC, va1, vb1);
    if (short_k && nchunk > 2) load_chunk_a(2 * kKC, va2);
  }
  TL(1);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_s;
  TL(2);

  // fp32 (round-to-nearest) running sum of the drained per-chunk accumulators
  float racc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) racc[i] = 0.f;
  const int c0 = 32 * (warp >> 2);                       // this warp's 32 accumulator columns (staging warps only)
  auto drain = [&](uint32_t acc_col) {
    if (c0 < nt) {                                       // warp-uniform
      const uint32_t taddr = tmem_d + (static_cast<uint32_t>(32 * (warp & 3)) << 16) + acc_col +
                             static_cast<uint32_t>(c0);
#pragma unroll
      for (int h = 0; h < 2; ++h) {                      // two 16-column halves: 16 temporaries instead of 32
        uint32_t v[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(taddr + 16u * h)
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 16; ++i) racc[16 * h + i] += __uint_as_float(v[i]);
      }
    }
  };
  const bool do_colsum = (OP == kTN) && colsum_part != nullptr && a0 == 0;     // CTA-uniform
  float csum[8];                                         // bias gradient: this thread's 8 features, its rows k
#pragma unroll
  for (int i = 0; i < 8; ++i) csum[i] = 0.f;
  // Warp-specialised pipeline over chunks c of kKC reduction elements, kStages shared-memory stages:
  //   warps 0-15 (staging): [wait done(c - kStages)]  split+store(c)  load(c + kStages)  arrive full(c)  [drain]
  //   warp 16   (MMA issue): wait full(c)  12 x tcgen05.mma -> accumulator (group & 1)  commit -> done(c)
  // No block-wide barrier inside the loop; the drain of chunk c (tcgen05.ld by the staging warps) overlaps the MMAs
  // of chunk c + 1, which accumulate in another TMEM accumulator.
  static_assert(kFlush == kKC, "one accumulator per chunk");
  auto wait_chunk = [&](int c) { mbar_wait(smem_u32(&mbar[c % kStages]), static_cast<uint32_t>((c / kStages) & 1)); };
  if (warp == kThreads / 32) {
    // ------------------------------- MMA-issue warp -------------------------------
    // descriptors differ only in the 14-bit start-address field (16-byte units)
    const uint64_t da_hi = kAMn ? desc_hi_bits(8, kKC) : desc_hi_bits(kPitchK, 8);
    const uint64_t db_hi = kBMn ? desc_hi_bits(8, kKC) : desc_hi_bits(kPitchK, 8);
    constexpr uint32_t a_step = kAMn ? 16 : 2 * kPitchK;       // start-address advance per MMA k-step (16 elements)
    constexpr uint32_t b_step = kBMn ? 16 : 2 * kPitchK;
    for (int c = 0; c < nchunk; ++c) {
      const int st = c % kStages;
      mbar_wait(smem_u32(&full[st]), static_cast<uint32_t>((c / kStages) & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (c < 4) TL(20 + 2 * c);
      if (c == 16) TL(50);
      if (lane == 0) {
        const uint32_t base = smem_u32(planes + static_cast<size_t>(st) * kStageUnits) >> 4;
        const uint64_t a1 = da_hi | (base & 0x3FFF), a2 = a1 + kPlaneUnits, a3 = a2 + kPlaneUnits;
        const uint64_t b1 = db_hi | ((base + 3 * kPlaneUnits) & 0x3FFF), b2 = b1 + kPlaneUnits, b3 = b2 + kPlaneUnits;
        const int64_t left = red_len - static_cast<int64_t>(c) * kKC;
        const int ksteps = (static_cast<int>(left < kKC ? left : kKC) + 15) / 16;         // skip all-zero k-steps
        constexpr bool fresh = true;                      // every chunk starts a new accumulation (kFlush == kKC)
        const uint32_t acc = tmem_d + static_cast<uint32_t>((short_k ? c : (c & 1)) * kMaxN);
        // small terms first: while they accumulate, the accumulator is 2^-16 .. 2^-8 of its final size and so are
        // the truncation errors of the tensor core's adds
#pragma unroll
        for (int ks = 0; ks < kKC / 16; ++ks) {
          if (ks < ksteps) {
            const uint32_t ao = ks * a_step, bo = ks * b_step;
            mma_bf16(acc, a3 + ao, b1 + bo, idesc, (fresh && ks == 0) ? 0u : 1u);
            mma_bf16(acc, a1 + ao, b3 + bo, idesc, 1u);
            mma_bf16(acc, a2 + ao, b2 + bo, idesc, 1u);
          }
        }
#pragma unroll
        for (int ks = 0; ks < kKC / 16; ++ks) {
          if (ks < ksteps) {
            const uint32_t ao = ks * a_step, bo = ks * b_step;
            mma_bf16(acc, a2 + ao, b1 + bo, idesc, 1u);
            mma_bf16(acc, a1 + ao, b2 + bo, idesc, 1u);
          }
        }
#pragma unroll
        for (int ks = 0; ks < kKC / 16; ++ks)
          if (ks < ksteps) mma_bf16(acc, a1 + ks * a_step, b1 + ks * b_step, idesc, 1u);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                         smem_u32(&mbar[st]))
                     : "memory");
      }
      if (c < 4) TL(21 + 2 * c);
      if (c == 16) TL(51);
      if (c == 15) TL(52);
      __syncwarp();
    }
  } else {
    // ------------------------------- staging warps -------------------------------
    auto stage = [&](int c, const F8& ra, const F8& rb) {         // split + store chunk c
      const int st = c % kStages;
      uint4* pa = planes + static_cast<size_t>(st) * kStageUnits;
      uint4* pb = pa + 3 * kPlaneUnits;
      if (kAMn) store_mnmajor(ra, pa); else store_kmajor(ra, pa);
      if (kBMn) store_mnmajor(rb, pb); else store_kmajor(rb, pb);
      if (OP == kTN && do_colsum) {
        csum[0] += rb.a.x; csum[1] += rb.a.y; csum[2] += rb.a.z; csum[3] += rb.a.w;
        csum[4] += rb.b.x; csum[5] += rb.b.y; csum[6] += rb.b.z; csum[7] += rb.b.w;
      }
    };
    auto publish = [&](int c) {                                    // hand chunk c to the MMA warp
      if (c == 16) TL(47);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> async proxy (MMA)
      if (c == 16) TL(48);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&full[c % kStages])) : "memory");
    };
    if (short_k) {
      if (nchunk > 0) {
        stage(0, va0, vb0);
        TL(3);
        if (nchunk > 2) load_chunk_b(2 * kKC, vb0);
        if (nchunk > 3) load_chunk_a(3 * kKC, va0);
        publish(0);
        TL(4);
      }
      if (nchunk > 1) { stage(1, va1, vb1); if (nchunk > 3) load_chunk_b(3 * kKC, vb1); publish(1); TL(5); }
      if (nchunk > 2) { stage(2, va2, vb0); publish(2); TL(6); }
      if (nchunk > 3) { stage(3, va0, vb1); publish(3); TL(7); }
      for (int c = 0; c < nchunk; ++c) {                   // accumulator c holds chunk c: fp32 (RN) sum in registers
        wait_chunk(c);
        TL(8 + c);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        drain(static_cast<uint32_t>(c * kMaxN));
      }
      TL(12);
    } else {
      auto step = [&](int c, F8& ra, F8& rb) {
        if (c == 16) TL(40);
        if (c >= kStages) wait_chunk(c - kStages);         // the MMAs that read this stage are done
        if (c == 16) TL(41);
        stage(c, ra, rb);
        if (c == 16) TL(42);
        if (c + 2 < nchunk) load_chunk(static_cast<int64_t>(c + 2) * kKC, ra, rb);   // refill this register set
        publish(c);
        if (c == 16) TL(43);
        // ---- drain the previous chunk's accumulator while this chunk's MMAs run ----
        if (c > 0) {
          wait_chunk(c - 1);
          if (c == 16) TL(44);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          drain(static_cast<uint32_t>(((c - 1) & 1) * kMaxN));
          if (c == 16) TL(45);
        }
        if (c == 17) TL(46);
      };
      for (int c = 0; c < nchunk; c += 2) {
        step(c, va0, vb0);
        if (c + 1 < nchunk) step(c + 1, va1, vb1);
      }
      wait_chunk(nchunk - 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain(static_cast<uint32_t>(((nchunk - 1) & 1) * kMaxN));
    }
  }

  // ---- epilogue: TMEM lane 32*(warp%4) + lane = output row; warp w holds columns 32*(w/4)..+31 of its 32 rows ----
  // The accumulator registers hold one output ROW per lane (TMEM lane = row).  Writing them straight out would make
  // every store instruction touch 32 different rows with 16 bytes each; instead the tile (+ bias, and f(pre) beside
  // it for the fused activation) goes through shared memory (the operand stages are free: the last MMAs have
  // completed) and leaves as one bulk asynchronous copy (TMA engine) per 512-byte row.
  const int d_row = 32 * (warp & 3) + lane;
  constexpr int OP_PITCH = kMaxN + 4;                              // floats; +4 keeps 16-byte alignment, spreads banks
  float* otile = reinterpret_cast<float*>(smem_raw);               // [kBM][OP_PITCH] = 67.6 KB
  float* atile = otile + kBM * OP_PITCH;                           // f(pre) tile of the fused activation
  static_assert(kSmemBytesDecl >= 2 * kBM * OP_PITCH * sizeof(float), "two epilogue tiles fit in the stages");
  if (warp < kThreads / 32) {
    const bool with_act = (OP == kNN) && act_out != nullptr;
    if (c0 < nt) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 v = make_float4(racc[i], racc[i + 1], racc[i + 2], racc[i + 3]);
        if (OP == kNN && bias) {                                   // scalar loads: the bias may be an unaligned view
          const int col = n0 + c0 + i;
          if (col + 0 < n_total) v.x += __ldg(bias + col + 0);
          if (col + 1 < n_total) v.y += __ldg(bias + col + 1);
          if (col + 2 < n_total) v.z += __ldg(bias + col + 2);
          if (col + 3 < n_total) v.w += __ldg(bias + col + 3);
        }
        *reinterpret_cast<float4*>(otile + d_row * OP_PITCH + c0 + i) = v;
        if (with_act)       // Dense + activation in one pass: pre (kept for the derivatives) and f(pre)
          *reinterpret_cast<float4*>(atile + d_row * OP_PITCH + c0 + i) =
              make_float4(act_value(act, v.x), act_value(act, v.y), act_value(act, v.z), act_value(act, v.w));
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> async proxy (bulk copy)
    asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");    // staging warps only
    TL(13);
    const int64_t out_rows = (OP == kTN) ? K : M;                  // rows of the output matrix this tile maps to
    const int64_t row_base = (OP == kTN) ? a0 : m0;
    float* obase = (OP == kTN) ? out + static_cast<int64_t>(bz) * K * U : out;
    const int ncols = min(nt, n_total - n0);                       // valid columns of this tile
    const bool bulk_ok = (n_total % 4 == 0) && (ncols % 4 == 0) && ((reinterpret_cast<uintptr_t>(obase) & 15) == 0) &&
                         (!with_act || (reinterpret_cast<uintptr_t>(act_out) & 15) == 0);
    if (bulk_ok) {
      if (threadIdx.x < kBM && row_base + threadIdx.x < out_rows) {
        const int r = threadIdx.x;
        const uint32_t bytes = static_cast<uint32_t>(ncols) * 4u;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                         obase + (row_base + r) * n_total + n0),
                     "r"(smem_u32(otile + r * OP_PITCH)), "r"(bytes)
                     : "memory");
        if (with_act)
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                           act_out + (row_base + r) * n_total + n0),
                       "r"(smem_u32(atile + r * OP_PITCH)), "r"(bytes)
                       : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");     // shared memory is released below
      }
    } else {                                                       // unaligned / ragged widths: per-element stores
      const int col = n0 + 4 * lane;                               // this lane's 4 columns of every row
      for (int r = warp; r < kBM; r += kThreads / 32) {
        const int64_t row = row_base + r;
        if (row >= out_rows || 4 * lane >= nt) continue;
        const float4 v = *reinterpret_cast<const float4*>(otile + r * OP_PITCH + 4 * lane);
        float* o = obase + row * n_total + col;
        if (col + 0 < n_total) o[0] = v.x;
        if (col + 1 < n_total) o[1] = v.y;
        if (col + 2 < n_total) o[2] = v.z;
        if (col + 3 < n_total) o[3] = v.w;
        if (with_act) {
          const float4 f = *reinterpret_cast<const float4*>(atile + r * OP_PITCH + 4 * lane);
          float* q = act_out + row * n_total + col;
          if (col + 0 < n_total) q[0] = f.x;
          if (col + 1 < n_total) q[1] = f.y;
          if (col + 2 < n_total) q[2] = f.z;
          if (col + 3 < n_total) q[3] = f.w;
        }
      }
    }
  }
  if (OP == kTN && do_colsum && warp < kThreads / 32) {
    // this thread summed 8 features (mn8 = 4 (warp / 4) + lane / 8) over its rows k = 8 (warp % 4) + lane % 8 of every
    // chunk: fixed order = xor tree over the 8 row lanes, then the 4 row warps -> deterministic
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float t = csum[i];
      t += __shfl_xor_sync(0xffffffffu, t, 4);
      t += __shfl_xor_sync(0xffffffffu, t, 2);
      t += __shfl_xor_sync(0xffffffffu, t, 1);
      if ((lane & 7) == 0) csum_s[warp][lane >> 3][i] = t;
    }
    asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");
    if (threadIdx.x < kMaxN) {
      const int f = threadIdx.x, mn8 = f >> 3, j = f & 7, w0 = 4 * (mn8 >> 2), sub = mn8 & 3;
      const float t = (csum_s[w0][sub][j] + csum_s[w0 + 1][sub][j]) + (csum_s[w0 + 2][sub][j] + csum_s[w0 + 3][sub][j]);
      if (f < nt && n0 + f < U) colsum_part[static_cast<int64_t>(bz) * U + n0 + f] = t;
    }
  }
  TL(14);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  TL(15);
  if (warp == kThreads / 32) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                 "r"(static_cast<uint32_t>(kTmemCols))
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
constexpr int kMaxGroup = 128;     // problems per launch (the table travels as a 13 KB kernel parameter)

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
  pdl_trigger();
  pdl_wait();
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

// ---------------------------------------------------------------------------------------------------------
// Relational Dense (kgcnn/layers/relational.py:232-234: out[n] = x[n] @ kernel[relation[n]]): the rows arrive
// SORTED by relation and `offs` (device, int32 [R + 1]) holds where each relation's rows begin, so one launch runs
// R independent Dense problems, each against its own [K, U] kernel.  The group sizes are read on the DEVICE: the
// launch shape depends only on (M, R), a CUDA graph replays on batches with another element composition.
// ---------------------------------------------------------------------------------------------------------
constexpr int kRelSplits = 8;     // row splits per relation of the relational weight gradient

// tile -> (relation, tile inside the relation's row range): warp-parallel prefix over the tile counts
__device__ __forceinline__ void rel_find_tile(const int32_t* __restrict__ offs, int R, int t, int* group, int* tile) {
  const int lane = threadIdx.x & 31;
  int base_tiles = 0, found_g = R, found_t = 0;
  for (int g0 = 0; g0 < R && found_g == R; g0 += 32) {
    const int g = g0 + lane;
    const int tiles = (g < R) ? (__ldg(offs + g + 1) - __ldg(offs + g) + kBM - 1) / kBM : 0;
    int incl = tiles;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += v;
    }
    const int excl = base_tiles + incl - tiles;
    const unsigned hit = __ballot_sync(0xffffffffu, g < R && t >= excl && t < excl + tiles);
    if (hit) {
      const int src = __ffs(hit) - 1;
      found_g = g0 + src;
      found_t = t - __shfl_sync(0xffffffffu, excl, src);
    }
    base_tiles += __shfl_sync(0xffffffffu, incl, 31);
  }
  *group = found_g;
  *tile = found_t;
}

template <int OP>
__global__ void __launch_bounds__(kCtaThreads)
dense_rel_kernel(const float* __restrict__ A, const float* __restrict__ W, const int32_t* __restrict__ offs, int R,
                 int K, int U, const float* __restrict__ bias, float* __restrict__ out, int vec_a, int vec_b) {
  __shared__ int s_group, s_tile;
  if (threadIdx.x < 32) {
    int g, t;
    rel_find_tile(offs, R, static_cast<int>(blockIdx.x), &g, &t);
    if (threadIdx.x == 0) { s_group = g; s_tile = t; }
  }
  __syncthreads();
  const int g = s_group;
  if (g >= R) return;                                           // the grid is an upper bound on the tile count
  const int64_t row0 = __ldg(offs + g), cnt = __ldg(offs + g + 1) - row0;
  const int wa = (OP == kNN) ? K : U, wo = (OP == kNN) ? U : K;
  dense_tc_body<OP>(A + row0 * wa, W + static_cast<int64_t>(g) * K * U, cnt, K, U, bias, out + row0 * wo, 0, vec_a,
                    vec_b, s_tile, blockIdx.y, 0, nullptr);
}

// dW[g] = X_g^T G_g: grid (feature tiles, column tiles, R * kRelSplits); partials [R][kRelSplits][K*U]
__global__ void __launch_bounds__(kCtaThreads)
dense_rel_tn_kernel(const float* __restrict__ X, const float* __restrict__ G, const int32_t* __restrict__ offs,
                    int K, int U, float* __restrict__ partial, int vec_a, int vec_b) {
  const int g = blockIdx.z / kRelSplits, s = blockIdx.z % kRelSplits;
  const int64_t row0 = __ldg(offs + g), cnt = __ldg(offs + g + 1) - row0;
  const int64_t rps = ceil_div<int64_t>(ceil_div<int64_t>(cnt, kRelSplits), kKC) * kKC;
  if (rps * s >= cnt) return;                                   // empty split (CTA-uniform): the reduce skips it
  dense_tc_body<kTN>(X + row0 * K, G + row0 * U, cnt, K, U, nullptr,
                     partial + static_cast<int64_t>(g) * kRelSplits * K * U, rps, vec_a, vec_b, blockIdx.x, blockIdx.y,
                     s, nullptr);
}

__global__ void __launch_bounds__(256)
dense_rel_tn_reduce_kernel(const float* __restrict__ partial, const int32_t* __restrict__ offs, int64_t n,
                           float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const int g = blockIdx.y;
  const int64_t cnt = __ldg(offs + g + 1) - __ldg(offs + g);
  const int64_t rps = ceil_div<int64_t>(ceil_div<int64_t>(cnt, kRelSplits), kKC) * kKC;
  float acc = 0.f;
  for (int s = 0; s < kRelSplits && rps * s < cnt; ++s)         // fixed order; relations without rows give 0
    acc += partial[(static_cast<int64_t>(g) * kRelSplits + s) * n + i];
  out[static_cast<int64_t>(g) * n + i] = acc;
}

constexpr size_t kSmemBytes = kSmemBytesDecl;

inline int tn_splits(int64_t M, int K, int U) {
  const int tiles = ceil_div(K, kBM) * ceil_div(U, kMaxN);
  int64_t s = std::max<int64_t>(1, (2 * kNumSM) / tiles);
  s = std::min<int64_t>(s, std::max<int64_t>(1, ceil_div<int64_t>(M, 4 * kKC)));
  return static_cast<int>(s);
}

// rows per split for one grouped launch: the finest split whose CTA count still fits TWO FULL WAVES of one resident
// CTA per SM (a launch of 2.1 waves runs three rounds: ncu of the C2 step, 314 CTAs), at least 4 chunks per CTA.
inline int64_t grouped_rows_per_split(int n, const int64_t* M, const int32_t* K, const int32_t* U) {
  int64_t chunk_tiles = 0;
  for (int i = 0; i < n; ++i)
    chunk_tiles += ceil_div<int64_t>(std::max<int64_t>(M[i], 1), kKC) * ceil_div(K[i], kBM) * ceil_div(U[i], kMaxN);
  const int64_t target = 2LL * kNumSM;
  int64_t per_cta = std::max<int64_t>(4, ceil_div<int64_t>(chunk_tiles, target));
  for (int iter = 0; iter < 4096; ++iter) {
    int64_t ctas = 0;
    for (int i = 0; i < n; ++i) {
      if (M[i] <= 0) continue;
      ctas += ceil_div<int64_t>(ceil_div<int64_t>(M[i], kKC), per_cta) * ceil_div(K[i], kBM) * ceil_div(U[i], kMaxN);
    }
    if (ctas <= target) break;
    per_cta += std::max<int64_t>(1, per_cta / 64);
  }
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
  // every Dense of SchNet / PaiNN: the compact aligned kernel (dense_fast.cu); everything else: the general one
  if (op != tc::kTN && tc::dense_fast_try_launch(op, A, B, M, K, U, bias, out, act_out, act, st)) return finish_launch();
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
    launch_pdl(tc::dense_tc_kernel<tc::kNN>, grid, dim3(tc::kCtaThreads), tc::kSmemBytes, st,
               A, B, M, K, U, bias, out, 0, (K % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0, act_out, act);
    return finish_launch();
  }
  if (op == tc::kNT) {
    dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(M, tc::kBM)), ceil_div(K, tc::kMaxN));
    launch_pdl(tc::dense_tc_kernel<tc::kNT>, grid, dim3(tc::kCtaThreads), tc::kSmemBytes, st,
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
  size_t floats = 0;
  for (int g0 = 0; g0 < n; g0 += tc::kMaxGroup) {            // the launcher packs kMaxGroup problems per launch
    const int cnt = std::min(tc::kMaxGroup, n - g0);
    const int64_t rps = tc::grouped_rows_per_split(cnt, M + g0, K + g0, U + g0);
    for (int i = g0; i < g0 + cnt; ++i) {
      const int64_t splits = std::max<int64_t>(1, ceil_div<int64_t>(M[i], rps));
      floats += static_cast<size_t>(splits) * (static_cast<size_t>(K[i]) * U[i] + U[i]);
      floats = (floats + 63) / 64 * 64;
    }
  }
  // x2: a destination conflict or an empty problem shifts the launch boundaries, and with them the split sizes
  return 2 * floats * sizeof(float) + 256;
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
    // split size of this launch: from the problems it can hold at most (a destination conflict may end it earlier,
    // which only leaves the launch below two waves)
    const int64_t rps = tc::grouped_rows_per_split(std::min(tc::kMaxGroup, n - i), M + i, K + i, U + i);
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
      if (static_cast<size_t>(ws_off) * sizeof(float) > workspace_bytes) return KGCNN_E_WORKSPACE;
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
    launch_pdl(tc::tn_grouped_kernel, dim3(static_cast<unsigned>(grp.total_ctas)), dim3(tc::kCtaThreads), tc::kSmemBytes,
               st, grp, ws);
    launch_pdl(tc::tn_grouped_reduce_kernel, dim3(static_cast<unsigned>(ceil_div<int64_t>(grp.total_elems, 256))),
               dim3(256), 0, st, grp, static_cast<const float*>(ws));
    launches += 2;
  }
  return launches ? finish_launch(launches) : KGCNN_OK;
}


size_t kgcnn_dense_rel_workspace_bytes(int op, int R, int K, int U) {
  if (op != tc::kTN || R <= 0 || K <= 0 || U <= 0) return 0;
  return static_cast<size_t>(R) * tc::kRelSplits * K * U * sizeof(float) + 256;
}

int kgcnn_dense_rel(int op, const float* A, const float* B, const int32_t* offs, int R, int64_t M, int K, int U,
                    const float* bias, float* out, void* workspace, size_t workspace_bytes, kgcnn_stream_t stream) {
  if (op < tc::kNN || op > tc::kTN || R <= 0 || M < 0 || K <= 0 || U <= 0 || !out || !offs) return KGCNN_E_INVALID;
  if (M > (int64_t{1} << 31) - 1) return KGCNN_E_INVALID;       // offsets are int32
  cudaStream_t st = as_stream(stream);
  if (op != tc::kTN && M == 0) return KGCNN_OK;
  if (M > 0 && (!A || !B)) return KGCNN_E_INVALID;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  static PerDeviceOnce attr;
  attr.run([] {
    cudaFuncSetAttribute(tc::dense_rel_kernel<tc::kNN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_rel_kernel<tc::kNT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
    cudaFuncSetAttribute(tc::dense_rel_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(tc::kSmemBytes));
  });
  if (op == tc::kTN) {                                          // A = X [M, K], B = G [M, U], out = dW [R, K, U]
    if (!workspace || workspace_bytes < kgcnn_dense_rel_workspace_bytes(op, R, K, U)) return KGCNN_E_WORKSPACE;
    if (!al16(workspace)) return KGCNN_E_INVALID;
    float* partial = static_cast<float*>(workspace);
    const int64_t n = static_cast<int64_t>(K) * U;
    if (M > 0) {
      dim3 grid(ceil_div(K, tc::kBM), ceil_div(U, tc::kMaxN), static_cast<unsigned>(R) * tc::kRelSplits);
      tc::dense_rel_tn_kernel<<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
          A, B, offs, K, U, partial, (K % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0);
    }
    dim3 rgrid(static_cast<unsigned>(ceil_div<int64_t>(n, 256)), static_cast<unsigned>(R));
    tc::dense_rel_tn_reduce_kernel<<<rgrid, 256, 0, st>>>(partial, offs, n, out);
    return finish_launch(2);
  }
  // sum over relations of ceil(rows / 128) <= M / 128 + R
  const unsigned tiles = static_cast<unsigned>(M / tc::kBM + R);
  if (op == tc::kNN) {                                          // A = X [M, K], B = kernels [R, K, U], out [M, U]
    dim3 grid(tiles, ceil_div(U, tc::kMaxN));
    tc::dense_rel_kernel<tc::kNN><<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
        A, B, offs, R, K, U, bias, out, (K % 4 == 0 && al16(A)) ? 1 : 0, (U % 4 == 0 && al16(B)) ? 1 : 0);
  } else {                                                      // A = G [M, U], B = kernels [R, K, U], out [M, K]
    dim3 grid(tiles, ceil_div(K, tc::kMaxN));
    tc::dense_rel_kernel<tc::kNT><<<grid, tc::kCtaThreads, tc::kSmemBytes, st>>>(
        A, B, offs, R, K, U, nullptr, out, (U % 4 == 0 && al16(A)) ? 1 : 0,
        (U % 4 == 0 && al16(B) && (static_cast<int64_t>(K) * U) % 4 == 0) ? 1 : 0);
  }
  return finish_launch();
}

}  // extern "C"
