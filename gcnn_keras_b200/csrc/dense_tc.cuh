// Shared pieces of the tcgen05 Dense kernels (dense_tc.cu: general shapes, TN, grouped TN; dense_fast.cu: the compact
// aligned NN / NT kernel for reductions of at most 128 elements).  See dense_tc.cu for the design notes.
#pragma once
#include <algorithm>

#include "activation.cuh"
#include "common.cuh"

namespace kgcnn {
namespace tc {

// Timeline probe (tools/tc_timeline.cu defines KGCNN_TC_TIMELINE and includes this file): clock64 stamps of CTA 0.
#ifdef KGCNN_TC_TIMELINE
#ifndef KGCNN_TL_THREAD
#define KGCNN_TL_THREAD 0
#endif
__device__ long long g_timeline[64];
#define TL(slot) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x == KGCNN_TL_THREAD || threadIdx.x == kThreads)) g_timeline[(slot)] = clock64(); } while (0)
#else
#define TL(slot) do { } while (0)
#endif

constexpr int kThreads = 512;   // 16 staging warps: one 8-element unit of A and one of B per thread and chunk;
                                // warp w reads TMEM lanes 32*(w%4)..+31, columns 32*(w/4)..+31 in the drain / epilogue
constexpr int kCtaThreads = kThreads + 32;   // + the MMA-issue warp (544 threads -> 96 registers per thread)
constexpr int kBM = 128;        // MMA M
constexpr int kKC = 32;         // reduction elements staged per chunk (2 MMA k-steps of 16)
constexpr int kStages = 4;      // shared-memory stages of (A b1,b2,b3, B b1,b2,b3) planes: 4 x 48.75 KB
constexpr int kMaxN = 128;      // MMA N per CTA (TMEM columns)
constexpr int kFlush = kKC;     // reduction length accumulated inside the tensor core before an fp32 (RN) flush
constexpr int kTmemCols = 512;  // four accumulators of kMaxN columns (short reductions: one per chunk, drained at the end)
// K-major plane: 16-byte unit (k8, mn) at slot k8 * kPitchK + mn.  130 = 2 mod 8: the 8 lanes of a quarter warp
// (4 k8 columns x 2 rows) land in 8 different 16-byte bank groups.
constexpr int kPitchK = kBM + 2;
// MN-major plane: unit (mn8, k) at slot mn8 * kKC + k: 8 consecutive k are 128 contiguous bytes (one core matrix).
constexpr int kPlaneUnits = (kKC / 8) * kPitchK;                       // 520 units >= 16 * kKC = 512
static_assert(kPlaneUnits >= (kMaxN / 8) * kKC, "plane holds either layout");
constexpr int kStageUnits = 6 * kPlaneUnits;
constexpr size_t kSmemBytesDecl = static_cast<size_t>(kStages) * kStageUnits * 16;   // 4 stages x 6 planes x 8.1 KB

enum { kNN = 0, kNT = 1, kTN = 2 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (sm_100): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46).
// For both majors LBO is the byte offset between core matrices along K, SBO along M/N.
__device__ __forceinline__ uint64_t desc_hi_bits(uint32_t lbo16, uint32_t sbo16) {
  return (static_cast<uint64_t>(lbo16 & 0x3FFF) << 16) | (static_cast<uint64_t>(sbo16 & 0x3FFF) << 32) | (1ull << 46);
}

// instruction descriptor kind::f16: D=F32 [4,6), A=BF16 [7,10), B=BF16 [10,13), a_major 15, b_major 16 (1 = MN-major),
// N>>3 [17,23), M>>4 [24,29)
__device__ __forceinline__ uint32_t make_idesc(int a_mn_major, int b_mn_major, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn_major) << 15) |
         (static_cast<uint32_t>(b_mn_major) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(kBM >> 4) << 24);
}

__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
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

// eight consecutive floats g[row, f .. f+7] (zero beyond row_end / f_end)
struct F8 { float4 a, b; };
__device__ __forceinline__ F8 load8(const float* __restrict__ g, int64_t row, int64_t row_end, int f, int f_end, int ld,
                                    bool vec_ok, bool fast) {
  F8 v;
  if (fast) {
    const float4* p = reinterpret_cast<const float4*>(g + row * ld + f);
    v.a = __ldg(p);
    v.b = __ldg(p + 1);
    return v;
  }
  v.a = make_float4(0.f, 0.f, 0.f, 0.f);
  v.b = v.a;
  if (row < row_end && f < f_end) {
    const float* p = g + row * ld + f;
    if (vec_ok && f + 7 < f_end) {
      v.a = __ldg(reinterpret_cast<const float4*>(p));
      v.b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    } else {
      float t[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) t[q] = (f + q < f_end) ? __ldg(p + q) : 0.f;
      v.a = make_float4(t[0], t[1], t[2], t[3]);
      v.b = make_float4(t[4], t[5], t[6], t[7]);
    }
  }
  return v;
}

// (x0, x1) -> three packed bf16x2 words (x0 in the low half): b1 = rn(x), b2 = rn(x - b1), b3 = rn(x - b1 - b2);
// both remainders are exact in fp32.
__device__ __forceinline__ void split3(float x0, float x1, uint32_t& p1, uint32_t& p2, uint32_t& p3) {
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p1) : "f"(x1), "f"(x0));
  const float r0 = x0 - __uint_as_float(p1 << 16), r1 = x1 - __uint_as_float(p1 & 0xffff0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p2) : "f"(r1), "f"(r0));
  const float s0 = r0 - __uint_as_float(p2 << 16), s1 = r1 - __uint_as_float(p2 & 0xffff0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p3) : "f"(s1), "f"(s0));
}
__device__ __forceinline__ void store_unit(const F8& v, uint4* __restrict__ planes, int slot) {
  uint4 u1, u2, u3;
  split3(v.a.x, v.a.y, u1.x, u2.x, u3.x);
  split3(v.a.z, v.a.w, u1.y, u2.y, u3.y);
  split3(v.b.x, v.b.y, u1.z, u2.z, u3.z);
  split3(v.b.z, v.b.w, u1.w, u2.w, u3.w);
  planes[slot] = u1;
  planes[kPlaneUnits + slot] = u2;
  planes[2 * kPlaneUnits + slot] = u3;
}

// K-major tile: mn = global row (128 of them from row0), k = feature (kKC of them from f0).
// thread -> (k8 = t % 4, mn = t / 4): a warp reads 8 rows x 128 contiguous bytes.
__device__ __forceinline__ F8 load_kmajor(const float* __restrict__ g, int64_t row0, int64_t row_end, int f0, int f_end,
                                          int ld, bool vec_ok, bool fast) {
  const int k8 = threadIdx.x & 3, mn = threadIdx.x >> 2;
  return load8(g, row0 + mn, row_end, f0 + 8 * k8, f_end, ld, vec_ok, fast);
}
__device__ __forceinline__ void store_kmajor(const F8& v, uint4* __restrict__ planes) {
  const int k8 = threadIdx.x & 3, mn = threadIdx.x >> 2;
  store_unit(v, planes, k8 * kPitchK + mn);
}
// MN-major tile: k = global row (kKC of them from row0), mn = feature (128 of them from f0).
// thread -> (k = 8 * (warp % 4) + lane % 8, mn8 = 4 * (warp / 4) + lane / 8): a quarter warp stores one core matrix.
__device__ __forceinline__ void mn_coords(int& k, int& mn8) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  k = 8 * (warp & 3) + (lane & 7);
  mn8 = 4 * (warp >> 2) + (lane >> 3);
}
__device__ __forceinline__ F8 load_mnmajor(const float* __restrict__ g, int64_t row0, int64_t row_end, int f0, int f_end,
                                           int ld, bool vec_ok, bool fast) {
  int k, mn8;
  mn_coords(k, mn8);
  return load8(g, row0 + k, row_end, f0 + 8 * mn8, f_end, ld, vec_ok, fast);
}
__device__ __forceinline__ void store_mnmajor(const F8& v, uint4* __restrict__ planes) {
  int k, mn8;
  mn_coords(k, mn8);
  store_unit(v, planes, mn8 * kKC + k);
}


// dense_fast.cu: returns true if it launched the compact kernel for this problem (aligned operands, reduction length a
// multiple of 32 and at most 128, output width a multiple of 16), false if the general kernel has to take it.
bool dense_fast_try_launch(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias,
                           float* out, float* act_out, int act, cudaStream_t st);

}  // namespace tc
}  // namespace kgcnn
