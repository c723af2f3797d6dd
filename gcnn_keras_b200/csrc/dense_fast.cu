// Compact NN / NT Dense kernel for the shapes every SchNet / PaiNN layer has: reduction length 32 .. 128 (a multiple of
// 32), 16-byte aligned rows, output width a multiple of 16.  Same arithmetic as dense_tc.cu (3-piece bf16 split, six
// kind::f16 MMAs per 16 reduction elements, one TMEM accumulator per 32-element chunk, fp32 round-to-nearest sum of the
// accumulators in registers), but straight-line code without any predicated edge path: a launch of the general kernel
// runs every instruction exactly once per tile, and its ~10 k SASS instructions of edge handling cost more in
// instruction-cache misses than in issue slots (tools/tc_timeline.cu: 3 900 cycles from entry to the first load).
//   * rows past the end of the matrix are CLAMPED for the loads (their results are never stored);
//   * columns past the layer width are clamped too (the MMA's N covers only the valid ones);
//   * the reduction needs no zero fill because its length is a multiple of the chunk.
// One CTA = one 128-row tile: 16 staging warps (one 8-element unit of A and of B per thread and chunk) + one warp that
// issues the MMAs; all global loads of the A tile are in flight before the first split.  Epilogue: accumulators (+ bias,
// + f(pre) beside it for the fused activation) -> shared memory -> one bulk asynchronous copy per 512-byte row.
// Large M with a short reduction (<= 64: the [E, 20] -> [E, 384] filter Dense of PaiNN on 4e5 edges) runs PERSISTENT:
// one CTA per SM walks over row tiles, so TMEM allocation, barrier set-up, the parameter fetch and the staging of W
// happen once; the next tile's rows are loaded while the current tile drains, and its bulk stores leave while the next
// tile is staged (the output tile then has its own shared memory instead of aliasing the operand stages).
#include "dense_tc.cuh"

namespace kgcnn {
namespace tc {

// persistent walk: W planes (4 chunks) + X planes (4 chunks) or two dense 64 KB output tiles behind the W planes
constexpr size_t kSmemBytesPersistent = static_cast<size_t>(kStages) * 3 * kPlaneUnits * 16 + 2 * kBM * kMaxN * 4;
static_assert(kSmemBytesPersistent >= kSmemBytesDecl && kSmemBytesPersistent + 1024 <= 232448, "fits one SM");

// RAG 0: everything aligned.  RAG 1 (NN): reduction length K arbitrary <= 128, rows of A unaligned (the [E, 20..25]
// radial-basis inputs of the filter networks): element-wise predicated loads of A, rows of W past K read as zero.
// RAG 2 (NT): output width K arbitrary (the gradient with respect to that basis): element-wise stores.
//
// TR (output width a multiple of 128): the MMA computes the TRANSPOSED tile, D^T = W^T X^T -- the operand roles are
// swapped in the instruction (the shared-memory planes and descriptors are the same), so a TMEM lane is an output
// COLUMN and the 32 lanes of a warp write 32 consecutive floats of one output row.  The output tile can then be
// dense in shared memory (pitch 512 B, conflict-free) and a [rows, 128] matrix leaves as ONE bulk copy of up to 64 KB
// instead of 128 copies of one padded row each (measured with tools/tc_timeline.cu: issuing the 128 row copies took
// ~2 400 cycles of a ~6 000-cycle tile).
template <int OP, int ACT, int RAG, bool TR, bool PS>   // ACT < 0: one output; else out = pre, act_out = f_ACT(pre)
                                                          // PS: persistent walk over row tiles (grid.x < tiles)
__global__ void __launch_bounds__(kCtaThreads)
dense_fast_kernel(const float* __restrict__ A, const float* __restrict__ B, int64_t M, int K, int U,
                  const float* __restrict__ bias, float* __restrict__ out, float* __restrict__ act_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t mbar[kStages];
  __shared__ __align__(8) uint64_t full[kStages];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[kMaxN];
  // shared memory: [W planes: chunk][b1, b2, b3] then [X planes: chunk][a1, a2, a3], kPlaneUnits 16-byte units each.
  // The persistent walk keeps the W region for the whole kernel and reuses the X region (and what lies behind it)
  // for the output tile.
  uint4* planes = reinterpret_cast<uint4*>(smem_raw);
  auto b_planes = [&](int c) { return planes + static_cast<size_t>(c) * 3 * kPlaneUnits; };
  auto a_planes = [&](int c) { return planes + static_cast<size_t>(kStages + c) * 3 * kPlaneUnits; };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  TL(0);
  // Touch every kernel parameter first: the constant-bank misses of a freshly launched kernel (~700 cycles per access
  // path, measured with tools/tc_timeline.cu) then overlap instead of queueing behind one another in the prologue.
  asm volatile("" ::"l"(A), "l"(B), "l"(M), "r"(K), "r"(U), "l"(bias), "l"(out), "l"(act_out));
  if (A == nullptr || B == nullptr || M <= 0 || K <= 0 || U <= 0 || out == nullptr) return;   // uniform-path reads
  const int n_total = (OP == kNN) ? U : K;                       // output width
  const int red = (OP == kNN) ? K : U;                           // reduction length = row length of A
  const int n0 = blockIdx.y * kMaxN;
  const int nt = min(kMaxN, (n_total - n0 + 15) / 16 * 16);      // MMA N (multiple of 16; RAG 2: rounded up)
  int64_t m0 = static_cast<int64_t>(blockIdx.x) * kBM;
  const int nchunk = (red + kKC - 1) / kKC;                      // 1 .. 4
  const int64_t ntiles = (M + kBM - 1) / kBM;
  constexpr bool persistent = PS;                                // the launcher guarantees nchunk <= 2 and the smem
  const int my_tiles = PS ? static_cast<int>((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 1;

  if (warp == kThreads / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(static_cast<uint32_t>(kTmemCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    TL(17);
  }
  if (threadIdx.x == 0) {
#pragma unroll
    for (int st = 0; st < kStages; ++st) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[st])) : "memory");
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[st])), "r"(kThreads / 32)
                   : "memory");
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    TL(16);
  }

  // per-thread source pointers of its unit in chunk 0 (16-byte aligned by the launch conditions)
  const float4* pa = nullptr;
  const float4* pb = nullptr;
  int64_t b_step = 0;                                            // float4 per chunk
  if (warp < kThreads / 32) {
    const int t = threadIdx.x;
    const int64_t row = min(m0 + (t >> 2), M - 1);
    pa = reinterpret_cast<const float4*>(A + row * red + 8 * (t & 3));
    if (OP == kNN) {                                             // W[k, u]: reduction along the rows, MN-major
      int k, mn8;
      mn_coords(k, mn8);
      const int f = min(n0 + 8 * mn8, n_total - 8);
      pb = reinterpret_cast<const float4*>(B + static_cast<int64_t>(k) * U + f);
      b_step = static_cast<int64_t>(kKC) * U / 4;
    } else {                                                     // W[n, u]: reduction along the features, K-major
      const int n = min(n0 + (t >> 2), K - 1);
      pb = reinterpret_cast<const float4*>(B + static_cast<int64_t>(n) * U + 8 * (t & 3));
      b_step = kKC / 4;
    }
  }
  auto ld_a = [&](int c) {
    F8 v;
    if (RAG == 1) {                                              // unaligned rows, ragged reduction length
      const float* p = reinterpret_cast<const float*>(pa) + c * kKC;
      const int f = c * kKC + 8 * (threadIdx.x & 3);
      float t[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) t[q] = (f + q < red) ? __ldg(p + q) : 0.f;
      v.a = make_float4(t[0], t[1], t[2], t[3]);
      v.b = make_float4(t[4], t[5], t[6], t[7]);
    } else {
      v.a = __ldg(pa + c * (kKC / 4));
      v.b = __ldg(pa + c * (kKC / 4) + 1);
    }
    return v;
  };
  auto ld_b = [&](int c) {
    F8 v;
    if (RAG == 1 && c * kKC + 8 * (warp & 3) + (lane & 7) >= red) {   // row of W past the reduction length
      v.a = make_float4(0.f, 0.f, 0.f, 0.f);
      v.b = v.a;
    } else {
      v.a = __ldg(pb + c * b_step);
      v.b = __ldg(pb + c * b_step + 1);
    }
    return v;
  };
  F8 va0, va1, va2, vb0, vb1;
  TL(18);
  pdl_wait();                                                    // first global access below
  if (warp < kThreads / 32) {
    va0 = ld_a(0);
    vb0 = ld_b(0);
    if (nchunk > 1) { va1 = ld_a(1); vb1 = ld_b(1); }
    if (nchunk > 2) va2 = ld_a(2);
    if (OP == kNN && threadIdx.x < kMaxN)                        // bias of this column tile -> shared (read in the epilogue)
      bias_s[threadIdx.x] = (bias && threadIdx.x < nt) ? __ldg(bias + n0 + threadIdx.x) : 0.f;
  }
  TL(1);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_s;
  TL(2);

  constexpr bool kBMn = (OP == kNN);
  auto wait_chunk = [&](int c, int it) { mbar_wait(smem_u32(&mbar[c]), static_cast<uint32_t>(it & 1)); };
  float racc[32];
  const int c0 = 32 * (warp >> 2);                       // this warp's 32 accumulator columns (staging warps only)
  if (warp == kThreads / 32) {
    // ------------------------------- MMA-issue warp -------------------------------
    const uint32_t idesc = TR ? make_idesc(kBMn ? 1 : 0, 0, kBM) : make_idesc(0, kBMn ? 1 : 0, nt);
    auto mm = [&](uint32_t acc, uint64_t xd, uint64_t wd, uint32_t flag) {
      if (TR) mma_bf16(acc, wd, xd, idesc, flag); else mma_bf16(acc, xd, wd, idesc, flag);
    };
    const uint64_t da_hi = desc_hi_bits(kPitchK, 8);
    const uint64_t db_hi = kBMn ? desc_hi_bits(8, kKC) : desc_hi_bits(kPitchK, 8);
    constexpr uint32_t a_step = 2 * kPitchK;
    constexpr uint32_t bk_step = kBMn ? 16 : 2 * kPitchK;
    for (int it = 0; it < my_tiles; ++it)
    for (int c = 0; c < nchunk; ++c) {
      mbar_wait(smem_u32(&full[c]), static_cast<uint32_t>(it & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      TL(20 + 2 * c);
      if (lane == 0) {
        const uint32_t base_a = smem_u32(a_planes(c)) >> 4, base_b = smem_u32(b_planes(c)) >> 4;
        const uint64_t a1 = da_hi | (base_a & 0x3FFF), a2 = a1 + kPlaneUnits, a3 = a2 + kPlaneUnits;
        const uint64_t b1 = db_hi | (base_b & 0x3FFF), b2 = b1 + kPlaneUnits, b3 = b2 + kPlaneUnits;
        const uint32_t acc = tmem_d + static_cast<uint32_t>(c * kMaxN);
        // small terms first (see dense_tc.cu)
        mm(acc, a3, b1, 0u);
        mm(acc, a1, b3, 1u);
        mm(acc, a2, b2, 1u);
        mm(acc, a3 + a_step, b1 + bk_step, 1u);
        mm(acc, a1 + a_step, b3 + bk_step, 1u);
        mm(acc, a2 + a_step, b2 + bk_step, 1u);
        mm(acc, a2, b1, 1u);
        mm(acc, a1, b2, 1u);
        mm(acc, a2 + a_step, b1 + bk_step, 1u);
        mm(acc, a1 + a_step, b2 + bk_step, 1u);
        mm(acc, a1, b1, 1u);
        mm(acc, a1 + a_step, b1 + bk_step, 1u);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                         smem_u32(&mbar[c]))
                     : "memory");
      }
      TL(21 + 2 * c);
      __syncwarp();
    }
  } else {
    // ------------------------------- staging warps -------------------------------
    auto stage = [&](int c, const F8& ra, const F8& rb) {
      store_kmajor(ra, a_planes(c));
      if (kBMn) store_mnmajor(rb, b_planes(c)); else store_kmajor(rb, b_planes(c));
    };
    auto publish = [&](int c) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> async proxy (MMA)
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&full[c])) : "memory");
    };
    auto stage_a = [&](int c, const F8& ra) { store_kmajor(ra, a_planes(c)); };
    const int d_row = 32 * (warp & 3) + lane;                      // epilogue: lane = output row, columns c0 .. c0 + 31
    constexpr int OP_PITCH = kMaxN + 4;
    // One-tile launches alias the output tile(s) onto the operand planes from the start of shared memory (free once
    // every MMA has completed).  The persistent walk keeps W staged: its output tile(s) lie behind the X planes of
    // this reduction length when that fits (the bulk stores of a tile then overlap the staging of the next), else ON
    // the X planes (staging waits until the stores have read them).
    constexpr size_t kTileBytes = static_cast<size_t>(kBM) * (TR ? kMaxN : OP_PITCH) * sizeof(float) * (ACT >= 0 ? 2 : 1);
    const size_t x_end = static_cast<size_t>(kStages + nchunk) * 3 * kPlaneUnits * 16;
    const bool alias_x = PS && (x_end + kTileBytes > kSmemBytesPersistent);                  // CTA-uniform
    float* otile = reinterpret_cast<float*>(
        smem_raw + (!PS ? 0 : (alias_x ? static_cast<size_t>(kStages) * 3 * kPlaneUnits * 16 : x_end)));
    float* atile = otile + kBM * (TR ? kMaxN : OP_PITCH);
    auto stage_b = [&](int c, const F8& rb) {
      if (kBMn) store_mnmajor(rb, b_planes(c)); else store_kmajor(rb, b_planes(c));
    };
    if (PS) {                                                      // W is staged once for all tiles
      stage_b(0, vb0);
      if (nchunk > 1) stage_b(1, vb1);
      if (nchunk > 2) { vb0 = ld_b(2); if (nchunk > 3) vb1 = ld_b(3); stage_b(2, vb0); }
      if (nchunk > 3) stage_b(3, vb1);
    }
    for (int it = 0; it < my_tiles; ++it) {
      TL(31);
      if (!PS) {
        stage(0, va0, vb0);
        TL(3);
        if (nchunk > 2) vb0 = ld_b(2);
        if (nchunk > 3) va0 = ld_a(3);
        publish(0);
        TL(4);
        if (nchunk > 1) { stage(1, va1, vb1); if (nchunk > 3) vb1 = ld_b(3); publish(1); TL(5); }
        if (nchunk > 2) { stage(2, va2, vb0); publish(2); TL(6); }
        if (nchunk > 3) { stage(3, va0, vb1); publish(3); TL(7); }
      } else {
        if (it > 0 && alias_x) {                                   // the previous tile's bulk stores have read the X planes
          if (threadIdx.x < kBM) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");
        }
        TL(36);
        stage_a(0, va0);
        if (nchunk > 3) va0 = ld_a(3);                             // pa already points at this tile's rows
        publish(0);
        if (nchunk > 1) { stage_a(1, va1); publish(1); }
        if (nchunk > 2) { stage_a(2, va2); publish(2); }
        TL(37);
        if (nchunk > 3) { stage_a(3, va0); publish(3); }
      }
      TL(32);
      // accumulator c holds chunk c: fp32 (RN) sum in registers
      for (int c = 0; c < nchunk; ++c) {
        wait_chunk(c, it);
        TL(8 + c);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (TR || c0 < nt) {
          const uint32_t taddr = tmem_d + (static_cast<uint32_t>(32 * (warp & 3)) << 16) + static_cast<uint32_t>(c * kMaxN + c0);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
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
            for (int i = 0; i < 16; ++i)
              racc[16 * h + i] = (c == 0) ? __uint_as_float(v[i]) : racc[16 * h + i] + __uint_as_float(v[i]);
          }
        }
      }
      TL(12);
      if (it + 1 < my_tiles) {                                     // next tile's rows: in flight during the epilogue
        const int64_t row = min(m0 + static_cast<int64_t>(gridDim.x) * kBM + (threadIdx.x >> 2), M - 1);
        pa = reinterpret_cast<const float4*>(A + row * red + 8 * (threadIdx.x & 3));
        va0 = ld_a(0);
        if (nchunk > 1) va1 = ld_a(1);
        if (nchunk > 2) va2 = ld_a(2);
      }

      if (it > 0 && !alias_x) {                                    // the previous tile's bulk stores have read otile
        if (threadIdx.x < kBM) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the issuing threads
        TL(33);
        asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");
      }
      TL(34);
      if (TR) {                                                    // lane = output column, registers = 32 rows
        const int col = 32 * (warp & 3) + lane;
        const float bv = (OP == kNN) ? bias_s[col] : 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = racc[i] + bv;
          otile[(c0 + i) * kMaxN + col] = v;
          if (ACT >= 0) atile[(c0 + i) * kMaxN + col] = act_deriv<ACT < 0 ? 0 : ACT>(v, 0);
        }
      } else if (c0 < nt) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float4 v = make_float4(racc[i], racc[i + 1], racc[i + 2], racc[i + 3]);
          if (OP == kNN) {
            const float4 bv = *reinterpret_cast<const float4*>(bias_s + c0 + i);
            v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
          }
          *reinterpret_cast<float4*>(otile + d_row * OP_PITCH + c0 + i) = v;
          if (ACT >= 0)
            *reinterpret_cast<float4*>(atile + d_row * OP_PITCH + c0 + i) =
                make_float4(act_deriv<ACT < 0 ? 0 : ACT>(v.x, 0), act_deriv<ACT < 0 ? 0 : ACT>(v.y, 0),
                            act_deriv<ACT < 0 ? 0 : ACT>(v.z, 0), act_deriv<ACT < 0 ? 0 : ACT>(v.w, 0));
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> async proxy (bulk copy)
      asm volatile("bar.sync 1, %0;" ::"r"(kThreads) : "memory");    // staging warps only
      TL(13);
      if (RAG == 2) {                                               // ragged output width: element-wise stores
        const int ncols = min(nt, n_total - n0);
        for (int r = warp; r < kBM; r += kThreads / 32) {
          if (m0 + r >= M) break;
          for (int cidx = lane; cidx < ncols; cidx += 32)
            out[(m0 + r) * n_total + n0 + cidx] = otile[r * OP_PITCH + cidx];
        }
      } else if (TR && n_total == kMaxN) {                           // the tile is one contiguous range of the output
        if (threadIdx.x == 0) {
          const int64_t rows = min(static_cast<int64_t>(kBM), M - m0);
          const uint32_t bytes = static_cast<uint32_t>(rows) * kMaxN * 4u;
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + m0 * kMaxN),
                       "r"(smem_u32(otile)), "r"(bytes)
                       : "memory");
          if (ACT >= 0)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(act_out + m0 * kMaxN),
                         "r"(smem_u32(atile)), "r"(bytes)
                         : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (it + 1 == my_tiles) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
      } else if (threadIdx.x < kBM) {
        if (m0 + threadIdx.x < M) {
          const int r = threadIdx.x;
          const uint32_t bytes = static_cast<uint32_t>(nt) * 4u;
          const int pitch = TR ? kMaxN : OP_PITCH;
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                           out + (m0 + r) * n_total + n0),
                       "r"(smem_u32(otile + r * pitch)), "r"(bytes)
                       : "memory");
          if (ACT >= 0)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                             act_out + (m0 + r) * n_total + n0),
                         "r"(smem_u32(atile + r * pitch)), "r"(bytes)
                         : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (it + 1 == my_tiles) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      m0 += static_cast<int64_t>(gridDim.x) * kBM;
      TL(35);
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

template <int OP, int ACT, int RAG, bool TR>
static void launch_fast_tr(const float* A, const float* B, int64_t M, int K, int U, const float* bias, float* out,
                        float* act_out, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] {
    cudaFuncSetAttribute(dense_fast_kernel<OP, ACT, RAG, TR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(kSmemBytesDecl));
    cudaFuncSetAttribute(dense_fast_kernel<OP, ACT, RAG, TR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(kSmemBytesPersistent));
  });
  const int n_total = (OP == kNN) ? U : K;
  const int col_tiles = ceil_div(n_total, kMaxN);
  int64_t row_ctas = ceil_div<int64_t>(M, kBM);
  // persistent walk over the row tiles when there are more than two waves of them to amortise the set-up over
  if (row_ctas * col_tiles > 2 * kNumSM && (ACT < 0 || TR)) {     // (two PADDED tiles do not fit behind W)
    dim3 grid(static_cast<unsigned>(std::max(1, kNumSM / col_tiles)), col_tiles);
    launch_pdl(dense_fast_kernel<OP, ACT, RAG, TR, true>, grid, dim3(kCtaThreads), kSmemBytesPersistent, st, A, B, M, K,
               U, bias, out, act_out);
    return;
  }
  dim3 grid(static_cast<unsigned>(row_ctas), col_tiles);
  launch_pdl(dense_fast_kernel<OP, ACT, RAG, TR, false>, grid, dim3(kCtaThreads), kSmemBytesDecl, st, A, B, M, K, U, bias,
             out, act_out);
}

template <int OP, int ACT, int RAG = 0>
static void launch_fast(const float* A, const float* B, int64_t M, int K, int U, const float* bias, float* out,
                        float* act_out, cudaStream_t st) {
  const int n_total = (OP == kNN) ? U : K;
  if (RAG != 2 && n_total % kMaxN == 0) launch_fast_tr<OP, ACT, RAG, (RAG != 2)>(A, B, M, K, U, bias, out, act_out, st);
  else launch_fast_tr<OP, ACT, RAG, false>(A, B, M, K, U, bias, out, act_out, st);
}

bool dense_fast_try_launch(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias,
                           float* out, float* act_out, int act, cudaStream_t st) {
  if (op != kNN && op != kNT) return false;
  const int red = (op == kNN) ? K : U, n_total = (op == kNN) ? U : K;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (M < 1 || red > kStages * kKC || U % 4 != 0 || !al16(B)) return false;
  if (op == kNN && (red % kKC != 0 || !al16(A))) {          // ragged reduction (radial-basis inputs, K = 20 .. 25)
    if (n_total % 16 != 0 || !al16(out) || act_out) return false;
    launch_fast<kNN, -1, 1>(A, B, M, K, U, bias, out, nullptr, st);
    return true;
  }
  if (op == kNT && (n_total % 16 != 0 || !al16(out))) {     // ragged output width (gradient of those inputs)
    if (red % kKC != 0 || !al16(A)) return false;
    launch_fast<kNT, -1, 2>(A, B, M, K, U, nullptr, out, nullptr, st);
    return true;
  }
  if (red % kKC != 0 || n_total % 16 != 0) return false;
  if (!al16(A) || !al16(out) || (bias && !al16(bias)) || (act_out && !al16(act_out))) return false;
  if (op == kNT) {
    launch_fast<kNT, -1>(A, B, M, K, U, nullptr, out, nullptr, st);
    return true;
  }
  if (!act_out) {
    launch_fast<kNN, -1>(A, B, M, K, U, bias, out, nullptr, st);
    return true;
  }
  switch (act) {
    case KGCNN_ACT_SHIFTED_SOFTPLUS: launch_fast<kNN, KGCNN_ACT_SHIFTED_SOFTPLUS>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_SOFTPLUS2: launch_fast<kNN, KGCNN_ACT_SOFTPLUS2>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_SWISH: launch_fast<kNN, KGCNN_ACT_SWISH>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_TANH: launch_fast<kNN, KGCNN_ACT_TANH>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_RELU: launch_fast<kNN, KGCNN_ACT_RELU>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_SIGMOID: launch_fast<kNN, KGCNN_ACT_SIGMOID>(A, B, M, K, U, bias, out, act_out, st); return true;
    case KGCNN_ACT_LEAKY_RELU: launch_fast<kNN, KGCNN_ACT_LEAKY_RELU>(A, B, M, K, U, bias, out, act_out, st); return true;
    default: return false;
  }
}

}  // namespace tc
}  // namespace kgcnn
