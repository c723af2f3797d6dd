// SchNet continuous-filter convolution with the filter network INSIDE the gather-scale-scatter
// (kgcnn/layers/conv/schnet_conv.py:73-79: Dense(ssp) -> Dense -> GatherNodesOutgoing -> LazyMultiply ->
// PoolingLocalEdges, fed by GaussBasisLayer, kgcnn/layers/geom.py:567-570).  No [E, units] tensor reaches HBM: per tile
// of 128 edges (CSR order of the output rows) a CTA
//     1. recomputes the Gaussian basis (and its first / second derivative with respect to the distance) from d_e,
//     2. runs  p = rbf W1 + b1  on tcgen05 (3-piece bf16 split, like dense_fast.cu) for every derivative order q,
//     3. applies the activation chain rule in registers:  h = f(p),  h' = f'(p) p',  h'' = f''(p) p'^2 + f'(p) p'',
//        splits the result straight into the MN-major operand planes of the next MMA,
//     4. runs  W^(q) = h^(q) W2 (+ b2)  on tcgen05,
//     5. multiplies with the gathered node rows and reduces over the output rows in edge order.
// The filter depends on ONE scalar per edge, so every derivative of the layer is a forward-mode chain through the
// filter network (phi, phi', phi''): the kernel is closed under the two differentiations force training needs.
//
//     mode 0 (value)      out[i] = sum_e a[o_e] * phi_e                                          y,   i = receiver
//     mode 1 (VJP)        out[i] = sum_e a[o_e] * phi_e        edge[e] = <a[o_e] b[i_e], phi'_e>    gx,  gd:  i = sender
//     mode 2 (JVP)        out[i] = sum_e a[o_e] * phi_e + t_e b[o_e] * phi'_e                     ggy: i = receiver
//     mode 3 (VJP of VJP) out[i] = sum_e t_e a[o_e] * phi'_e   edge[e] = <a[o_e] c[i_e], phi'_e> + t_e <a[o_e] b[i_e], phi''_e>
//
// The MMAs compute the TRANSPOSED tiles (D^T = W^T X^T): a TMEM lane is a feature, a TMEM column an edge, so a warp of
// the epilogue owns 32 features x 32 consecutive edges: the gather of a node row is one coalesced 128-byte load and the
// segment sum is a sequential loop in registers, in edge order (deterministic, no atomics).  Segments that cross a
// 32-edge group leave partial sums in a workspace that a row-parallel fix-up kernel adds in group order; it also
// zeroes the rows without edges.
#include "dense_tc.cuh"

namespace kgcnn {
namespace cf {

using tc::desc_hi_bits;
using tc::F8;
using tc::kBM;
using tc::kKC;
using tc::kPitchK;
using tc::make_idesc;
using tc::mbar_wait;
using tc::mma_bf16;
using tc::smem_u32;
using tc::split3;

// Timeline probe (tools/cf_timeline.cu defines KGCNN_CF_TIMELINE and includes this file): clock64 stamps of CTA 0.
#ifdef KGCNN_CF_TIMELINE
__device__ long long g_timeline[64];
#define CTL(slot) do { if (blockIdx.x == 0 && (threadIdx.x == 0 || threadIdx.x == kWorkers)) g_timeline[(slot)] = clock64(); } while (0)
#else
#define CTL(slot) do { } while (0)
#endif

constexpr int kWorkers = 512;                      // 16 warps: staging, activation chain, epilogue
constexpr int kCtaThreads = kWorkers + 32;         // + the MMA-issue warp
constexpr int kF = 128;                            // units of the filter network (both Dense layers)
// Operand planes (16-byte units of 8 bf16).  "MN-major" unit (mn8, k) at mn8 * K + k holds 8 consecutive mn for one k.
//   W1: (j8, k) at j8 * 32 + k           k < 32 basis functions (rows past R are zero), j = hidden unit
//   W2: (u8, j) at u8 * 128 + j          read MN-major as (M = u, K = j) by the forward GEMM and K-major as
//                                         (M = j, K = u) by the transposed GEMM of the weight-gradient stage: the
//                                         8 x 8 core matrices are the same bytes, only LBO / SBO swap
//   H : (e8, k) at e8 * 128 + k          h^(q) (k = j), later the gathered products A (k = u); N = edge
constexpr int kW1Plane = (kF / 8) * kKC;           // 512 units
constexpr int kBigPlane = (kF / 8) * kF;           // 2048 units = 32 KB
constexpr int kKPlane = (kKC / 8) * kPitchK;       // K-major basis plane: unit (k8, mn) at k8 * kPitchK + mn: 520 units
constexpr int kWUnits = 3 * kW1Plane + 3 * kBigPlane;
constexpr int kHUnits = 3 * kBigPlane;
static_assert(3 * 3 * kKPlane + 16 * 64 <= kHUnits, "basis planes of three derivative orders (+ 1 KB per warp) fit behind the weights");
constexpr size_t kSmemBytes = static_cast<size_t>(kWUnits + kHUnits) * 16;
constexpr int kTmemCols = 512;

struct Params {
  const float* dist;       // [E] distances, original edge order
  const float* W1;         // [R, 128]
  const float* b1;         // [128] or null
  const float* W2;         // [128, 128]
  const float* b2;         // [128] or null
  const int32_t* rows;     // [E] output row of CSR position e (ascending)
  const int32_t* other;    // [E] gathered row of CSR position e
  const int32_t* eperm;    // [E] original edge id of CSR position e, or null (identity)
  const float* a;          // node tensors [*, 128]
  const float* b;
  const float* c;
  const float* tang;       // [E] per-edge scalar, original edge order
  float* out;              // [n_out, 128]
  float* edge_out;         // [E], original edge order
  float* ws;               // [ceil(E / 32)][2][128] partial sums of the segments that cross a 32-edge group
  // weight-gradient operands (WG variants), rows in CSR order: dW2 += h^T A (+ hd^T Bt), dW1 += rbf^T gp (+ rbf1^T gpd)
  float* wg_rbf;           // [E, R]
  float* wg_rbf1;          // [E, R]    mode 3
  float* wg_h;             // [E, 128]
  float* wg_A;             // [E, 128]
  float* wg_gp;            // [E, 128]
  float* wg_hd;            // [E, 128]  mode 3
  float* wg_Bt;            // [E, 128]  mode 3
  float* wg_gpd;           // [E, 128]  mode 3
  int64_t E;
  int R;
  int static_w;            // W1, b1, W2, b2 are variables no kernel of the running chain writes: staged before pdl_wait()
  float gamma, offset, step;
};

__device__ __forceinline__ void store_unit3(const float* v, uint4* base, int slot, int plane_stride) {
  uint4 u1, u2, u3;
  split3(v[0], v[1], u1.x, u2.x, u3.x);
  split3(v[2], v[3], u1.y, u2.y, u3.y);
  split3(v[4], v[5], u1.z, u2.z, u3.z);
  split3(v[6], v[7], u1.w, u2.w, u3.w);
  base[slot] = u1;
  base[plane_stride + slot] = u2;
  base[2 * plane_stride + slot] = u3;
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// sum of v[k] over the 32 lanes for k = 0..7: after three halving exchanges a lane holds the value index
// idx = 4 * bit4 + 2 * bit3 + bit2 of its lane number; two more exchanges complete the sum in every lane.
__device__ __forceinline__ float octet_reduce(const float* v, int lane, int& idx) {
  float w[4], x[2], y;
  const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float keep = h4 ? v[4 + k] : v[k], send = h4 ? v[k] : v[4 + k];
    w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const float keep = h3 ? w[2 + k] : w[k], send = h3 ? w[k] : w[2 + k];
    x[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  {
    const float keep = h2 ? x[1] : x[0], send = h2 ? x[0] : x[1];
    y = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  y += __shfl_xor_sync(0xffffffffu, y, 2);
  y += __shfl_xor_sync(0xffffffffu, y, 1);
  idx = (h4 ? 4 : 0) + (h3 ? 2 : 0) + (h2 ? 1 : 0);
  return y;
}

// shifted softplus f(p) = relu(p) + log1p(exp(-|p|)) - log 2 (kgcnn/ops/activ.py:15,29) and sigma(p) = f'(p) from ONE
// exponential; f'' = sigma (1 - sigma).  ex2 / lg2 / rcp are the hardware approximations (relative 2^-22; lg2 on
// [1, 2] absolute 2^-22): 16 k activations per tile make the accurate library sequences the longest phase of a tile.
__device__ __forceinline__ void ssp_eval(float p, float& f, float& sg) {
#ifdef KGCNN_CF_PRECISE
  const float e = expf(-fabsf(p));
  f = fmaxf(p, 0.f) + log1pf(e) - 0.69314718055994530942f;
  const float r = 1.f / (1.f + e);
#else
  const float e = __expf(-fabsf(p));
  f = fmaxf(p, 0.f) + (__logf(1.f + e) - 0.69314718055994530942f);
  const float r = __fdividef(1.f, 1.f + e);
#endif
  sg = (p >= 0.f) ? r : e * r;
}

template <int MODE, bool WG>
__global__ void __launch_bounds__(kCtaThreads)      // 17 warps are allocated as 20: 96 registers per thread
cfconv_kernel(const Params P) {
  constexpr int ORDER = (MODE == 0) ? 0 : (MODE == 3 ? 2 : 1);   // derivative orders of the filter this mode needs
  constexpr bool kEdgeOut = (MODE == 1 || MODE == 3);
  constexpr int Q0 = (MODE == 3) ? 1 : 0;                        // mode 3 never multiplies with phi itself
  static_assert(!WG || kEdgeOut, "weight-gradient operands come out of the two VJP modes");
  // PIPE (modes 0-2 without the weight-gradient stage): consecutive tiles alternate between the two 256-column halves
  // of TMEM and between two metadata buffers, and the epilogue of tile i runs AFTER the activation chain of tile i + 1
  // has been handed to the tensor core, i.e. under that tile's second MMA (48-96 MMAs) instead of in front of it.
  constexpr bool PIPE = !WG && MODE != 3;
  // TMEM columns (relative to the tile's half when PIPE): p^(q) at 128 q; phi^(q) and the transposed products G
  // behind / over them once they are dead
  constexpr uint32_t kColO0 = (MODE == 0) ? 128u : (PIPE ? 0u : 256u);   // PIPE: over p (in registers by then)
  constexpr uint32_t kColO1 = PIPE ? 128u : 384u;                         // PIPE: over p'

  constexpr uint32_t kColO2 = 256u;                              // over p'' (drained before the MMA is issued)
  constexpr uint32_t kColG0 = (MODE == 3) ? 384u : 256u;         // A W2^T, over a phi^(q) the epilogue has consumed
  constexpr uint32_t kColG1 = 256u;                              // Bt W2^T (mode 3)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar_rbf, bar_g1, bar_h[3], bar_g2[3], bar_a[2], bar_g3[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float b1_s[kF], b2_s[kF];
  __shared__ int row_sb[2][kBM + 2], oth_sb[2][kBM], eid_sb[2][kBM];
  __shared__ float d_sb[2][kBM], t_sb[2][kBM];
  __shared__ float pe_s[4][kBM];

  uint4* w1planes = reinterpret_cast<uint4*>(smem_raw);         // [piece][kW1Plane]
  uint4* w2planes = w1planes + 3 * kW1Plane;                    // [piece][kBigPlane]
  uint4* hplanes = w2planes + 3 * kBigPlane;                    // [piece][kBigPlane]
  uint4* rplanes = hplanes;                                     // [q][piece][kKPlane]   (alias: dead before h is staged)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  CTL(0);
  // touch every parameter line first: the constant-bank misses of a freshly launched kernel then overlap
  asm volatile("" ::"l"(P.dist), "l"(P.rows), "l"(P.a), "l"(P.out), "l"(P.wg_h), "l"(P.E), "r"(P.R));
  const int64_t E = P.E;
  const int64_t ntiles = (E + kBM - 1) / kBM;
  CTL(16);

  if (warp == kWorkers / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(static_cast<uint32_t>(kTmemCols))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    if (lane == 0) {
      auto init = [](uint64_t* b, int n) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n) : "memory");
      };
      init(&bar_rbf, kWorkers / 32);
      init(&bar_g1, 1);
      for (int q = 0; q < 3; ++q) { init(&bar_h[q], kWorkers / 32); init(&bar_g2[q], 1); }
      for (int q = 0; q < 2; ++q) { init(&bar_a[q], kWorkers / 32); init(&bar_g3[q], 1); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  // tile metadata -> shared memory (threads 0 .. 129 of the workers)
  auto load_meta = [&](int64_t e0, int mb) {
    int* row_s = row_sb[mb]; int* oth_s = oth_sb[mb]; int* eid_s = eid_sb[mb];
    float* d_s = d_sb[mb]; float* t_s = t_sb[mb];
    if (threadIdx.x < kBM) {
      const int64_t e = e0 + threadIdx.x;
      int row = -1, oth = 0, eid = 0;
      float d = 0.f, t = 0.f;
      if (e < E) {
        row = __ldg(P.rows + e);
        oth = __ldg(P.other + e);
        eid = P.eperm ? __ldg(P.eperm + e) : static_cast<int>(e);
        d = __ldg(P.dist + eid);
        if (MODE >= 2) t = __ldg(P.tang + eid);
      }
      row_s[1 + threadIdx.x] = row;
      oth_s[threadIdx.x] = oth;
      eid_s[threadIdx.x] = eid;
      d_s[threadIdx.x] = d;
      t_s[threadIdx.x] = t;
    } else if (threadIdx.x == kBM) {
      row_s[0] = (e0 > 0) ? __ldg(P.rows + e0 - 1) : -2;
    } else if (threadIdx.x == kBM + 1) {
      row_s[kBM + 1] = (e0 + kBM < E) ? __ldg(P.rows + e0 + kBM) : -2;
    }
  };
  // Weights that are model variables (written only by the optimizer kernel, which is launched without the PDL
  // attribute and therefore complete before any kernel of this chain started) are staged BEFORE the grid dependency
  // is awaited: the 80 KB weight fetch + split runs under the previous kernel's tail.
  if (!P.static_w) pdl_wait();
  // weights -> operand planes, once per CTA; every global load is in flight before the first split
  if (warp < kWorkers / 32) {
    const int kk = 8 * (warp & 3) + (lane & 7), mn8 = 4 * (warp >> 2) + (lane >> 3);
    float4 lo[5], hi[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const int row = (c == 0) ? kk : 32 * (c - 1) + kk;
      const float* src = ((c == 0) ? P.W1 : P.W2) + static_cast<int64_t>(row) * kF + 8 * mn8;
      const bool live = (c != 0) || (kk < P.R);
      lo[c] = live ? __ldg(reinterpret_cast<const float4*>(src)) : make_float4(0.f, 0.f, 0.f, 0.f);
      hi[c] = live ? __ldg(reinterpret_cast<const float4*>(src) + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    CTL(17);
    float bv1 = 0.f, bv2 = 0.f;
    if (threadIdx.x < kF) {
      if (P.b1) bv1 = __ldg(P.b1 + threadIdx.x);
      if (P.b2) bv2 = __ldg(P.b2 + threadIdx.x);
    }
    CTL(1);
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const float v[8] = {lo[c].x, lo[c].y, lo[c].z, lo[c].w, hi[c].x, hi[c].y, hi[c].z, hi[c].w};
      if (c == 0) store_unit3(v, w1planes, mn8 * kKC + kk, kW1Plane);
      else store_unit3(v, w2planes, mn8 * kF + 32 * (c - 1) + kk, kBigPlane);
    }
    if (threadIdx.x < kF) { b1_s[threadIdx.x] = bv1; b2_s[threadIdx.x] = bv2; }
    CTL(18);
  }
  if (P.static_w) pdl_wait();                                    // everything below reads the chain's tensors
  CTL(2);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  CTL(3);

  if (warp == kWorkers / 32) {
    // ------------------------------------ MMA-issue warp ------------------------------------
    const uint64_t mn_hi = desc_hi_bits(8, kF);                 // MN-major over the 128-deep planes (W2 fwd, h, A)
    const uint64_t w1_hi = desc_hi_bits(8, kKC);                // MN-major W1 (32 deep)
    const uint64_t r_hi = desc_hi_bits(kPitchK, 8);             // K-major basis
    const uint64_t w2t_hi = desc_hi_bits(kF, 8);                // W2 read K-major: (M = j, K = u)
    const uint32_t idesc1 = make_idesc(1, 0, kBM), idesc2 = make_idesc(1, 1, kBM), idesc3 = make_idesc(0, 1, kBM);
    const uint32_t w1base = smem_u32(w1planes) >> 4, w2base = smem_u32(w2planes) >> 4, hbase = smem_u32(hplanes) >> 4;
    constexpr int pw[6] = {0, 2, 1, 0, 1, 0}, ph[6] = {2, 0, 1, 1, 0, 0};   // small products first (dense_tc.cu)
    auto commit = [](uint64_t* bar) {
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                   : "memory");
    };
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t par = static_cast<uint32_t>(it & 1);
      const uint32_t hb = PIPE ? 256u * static_cast<uint32_t>(it & 1) : 0u;     // this tile's half of TMEM
      mbar_wait(smem_u32(&bar_rbf), par);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      CTL(20);
      if (lane == 0) {
#pragma unroll
        for (int q = 0; q <= ORDER; ++q) {
          uint32_t first = 0u;
#pragma unroll
          for (int term = 0; term < 6; ++term) {
            const uint64_t wd = w1_hi | ((w1base + pw[term] * kW1Plane) & 0x3FFF);
            const uint64_t rd = r_hi | ((hbase + (q * 3 + ph[term]) * kKPlane) & 0x3FFF);
            mma_bf16(tmem + hb + 128u * q, wd, rd, idesc1, first);
            first = 1u;
            mma_bf16(tmem + hb + 128u * q, wd + 16, rd + 2 * kPitchK, idesc1, 1u);
          }
        }
        commit(&bar_g1);
      }
      CTL(21);
      __syncwarp();
#pragma unroll
      for (int q = Q0; q <= ORDER; ++q) {
        mbar_wait(smem_u32(&bar_h[q]), par);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        CTL(22 + 2 * q);
        if (lane == 0) {
          const uint32_t acc = tmem + hb + (q == 0 ? kColO0 : (q == 1 ? kColO1 : kColO2));
          uint32_t first = 0u;
#pragma unroll
          for (int term = 0; term < 6; ++term) {
            const uint64_t wd = mn_hi | ((w2base + pw[term] * kBigPlane) & 0x3FFF);
            const uint64_t hd = mn_hi | ((hbase + ph[term] * kBigPlane) & 0x3FFF);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              mma_bf16(acc, wd + 16 * ks, hd + 16 * ks, idesc2, first);
              first = 1u;
            }
          }
          commit(&bar_g2[q]);
        }
        CTL(23 + 2 * q);
        __syncwarp();
      }
      if (WG) {
#pragma unroll
        for (int w = 0; w < (MODE == 3 ? 2 : 1); ++w) {
          mbar_wait(smem_u32(&bar_a[w]), par);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (lane == 0) {
            const uint32_t acc = tmem + (w == 0 ? kColG0 : kColG1);
            uint32_t first = 0u;
#pragma unroll
            for (int term = 0; term < 6; ++term) {
              const uint64_t wd = w2t_hi | ((w2base + pw[term] * kBigPlane) & 0x3FFF);
              const uint64_t ad = mn_hi | ((hbase + ph[term] * kBigPlane) & 0x3FFF);
#pragma unroll
              for (int ks = 0; ks < 8; ++ks) {                  // 16 u per step: two u8 groups of W2, 16 k units of A
                mma_bf16(acc, wd + 2 * kF * ks, ad + 16 * ks, idesc3, first);
                first = 1u;
              }
            }
            commit(&bar_g3[w]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------ worker warps ------------------------------------
    const int cg = warp & 3, eg = warp >> 2;                     // 32-feature group (TMEM lane quarter), 32-edge group
    const int u = 32 * cg + lane;                                // this thread's feature (hidden unit j in the chain)
    const uint32_t trow = tmem + (static_cast<uint32_t>(32 * cg) << 16);
    const int gbase = 32 * eg;                                   // first edge of this warp's group within the tile
    auto publish = [&](uint64_t* bar) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
    };
    auto wait_bar = [&](uint64_t* bar, uint32_t par) {
      mbar_wait(smem_u32(bar), par);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    };
    const float bias1 = b1_s[u], bias2 = b2_s[u];
    // ---- epilogue of one tile: phi^(q) (TMEM half cb) x gathered rows -> segment sums in edge order ----
    auto epilogue = [&](int64_t e0, int nvalid, int mb, uint32_t cb) {
      const int* row_s = row_sb[mb]; const int* oth_s = oth_sb[mb]; const int* eid_s = eid_sb[mb];
      const float* t_s = t_sb[mb];
      float na[8];
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) na[ii] = __ldg(P.a + static_cast<int64_t>(oth_s[gbase + ii]) * kF + u);
      {
        const int64_t group = e0 / 32 + eg;
        const int prev_row = row_s[gbase], next_row = row_s[gbase + 33];
        // bit i: the segment of edge gbase + i ends there (within this group)
        const uint32_t ends = __ballot_sync(0xffffffffu, lane == 31 || row_s[1 + gbase + lane] != row_s[2 + gbase + lane]);
        float acc = 0.f;
        int seg_start = 0;
#pragma unroll
        for (int sb = 0; sb < 4; ++sb) {
          float w0[8], w1[8], w2[8], pe[8], va[8], vb[8], vc[8], A8[8];
          if (MODE != 3) tmem_ld8(trow + cb + kColO0 + gbase + 8 * sb, w0);
          if (ORDER >= 1) tmem_ld8(trow + cb + kColO1 + gbase + 8 * sb, w1);
          if (ORDER >= 2) tmem_ld8(trow + cb + kColO2 + gbase + 8 * sb, w2);
#pragma unroll
          for (int ii = 0; ii < 8; ++ii) {
            const int i = gbase + 8 * sb + ii;
            const int oth = oth_s[i];
            const int row = max(row_s[1 + i], 0);
            va[ii] = na[ii];
            if (MODE == 2) vb[ii] = __ldg(P.b + static_cast<int64_t>(oth) * kF + u);
            if (MODE == 1 || MODE == 3) vb[ii] = __ldg(P.b + static_cast<int64_t>(row) * kF + u);
            if (MODE == 3) vc[ii] = __ldg(P.c + static_cast<int64_t>(row) * kF + u);
          }
          if (sb < 3) {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii)
              na[ii] = __ldg(P.a + static_cast<int64_t>(oth_s[gbase + 8 * (sb + 1) + ii]) * kF + u);
          }
#pragma unroll
          for (int ii = 0; ii < 8; ++ii) {
            const int i = 8 * sb + ii;                             // edge within the group
            const float tt = (MODE >= 2) ? t_s[gbase + i] : 0.f;
            float m;
            if (MODE == 0 || MODE == 1) m = va[ii] * (w0[ii] + bias2);
            else if (MODE == 2) m = va[ii] * (w0[ii] + bias2) + tt * vb[ii] * w1[ii];
            else m = tt * va[ii] * w1[ii];
            if (MODE == 1) { A8[ii] = va[ii] * vb[ii]; pe[ii] = A8[ii] * w1[ii]; }
            if (MODE == 3) { A8[ii] = va[ii] * vc[ii]; pe[ii] = A8[ii] * w1[ii] + tt * (va[ii] * vb[ii]) * w2[ii]; }
            if (WG && gbase + i < nvalid) {
              P.wg_A[(e0 + gbase + i) * kF + u] = A8[ii];
              if (MODE == 3) P.wg_Bt[(e0 + gbase + i) * kF + u] = va[ii] * vb[ii];
            }
            acc += m;
            if ((ends >> i) & 1u) {                                // the segment of `row` ends (within this group)
              const int row = row_s[1 + gbase + i];
              if (row >= 0) {
                const bool open_head = (seg_start == 0) && (row == prev_row);
                const bool open_tail = (i == 31) && (row == next_row);
                if (!open_head && !open_tail) P.out[static_cast<int64_t>(row) * kF + u] = acc;
                else P.ws[(group * 2 + (open_head ? 0 : 1)) * kF + u] = acc;
              }
              acc = 0.f;
              seg_start = i + 1;
            }
          }
          if (kEdgeOut) {
            int idx;
            const float s = octet_reduce(pe, lane, idx);
            if ((lane & 3) == 0) pe_s[cg][gbase + 8 * sb + idx] = s;
          }
          if (WG) store_unit3(A8, hplanes, (4 * eg + sb) * kF + u, kBigPlane);   // A^T: k = u, 8 consecutive edges
        }
      }
      if (WG) publish(&bar_a[0]);
      CTL(11);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"r"(kWorkers) : "memory");
      CTL(12);
      if (kEdgeOut && threadIdx.x < nvalid) {
        const int i = threadIdx.x;
        P.edge_out[eid_s[i]] = ((pe_s[0][i] + pe_s[1][i]) + pe_s[2][i]) + pe_s[3][i];
      }
    };
    int it = 0;
    int64_t prev_e0 = 0;
    int prev_nvalid = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t par = static_cast<uint32_t>(it & 1);
      const int mb = PIPE ? (it & 1) : 0;                          // metadata buffer / TMEM half of this tile
      const uint32_t hb = PIPE ? 256u * static_cast<uint32_t>(it & 1) : 0u;
      const int64_t e0 = tile * kBM;
      const int nvalid = static_cast<int>(min(static_cast<int64_t>(kBM), E - e0));
      const int* row_s = row_sb[mb]; const int* oth_s = oth_sb[mb];
      const float* d_s = d_sb[mb]; const float* t_s = t_sb[mb];
      (void)row_s; (void)oth_s;
      // ---- tile metadata ----
      load_meta(e0, mb);
      // the previous tile's last MMA has completed: the operand planes (which the basis aliases) are free again
      if (PIPE && it > 0) wait_bar(&bar_g2[ORDER], par ^ 1u);
      asm volatile("bar.sync 1, %0;" ::"r"(kWorkers) : "memory");
      CTL(4);

      // ---- Gaussian basis and its derivatives -> K-major planes (thread: 8 basis functions of one edge) ----
      {
        const int k8 = threadIdx.x & 3, mn = threadIdx.x >> 2;
        const float t0 = d_s[mn] - P.offset;
        const float tt = (MODE == 3) ? t_s[mn] : 0.f;
        float g0[8], g1[8], g2[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int k = 8 * k8 + q;
          const float t = t0 - P.step * k;
          const float g = (k < P.R) ? __expf(-P.gamma * t * t) : 0.f;
          g0[q] = g;
          if (ORDER >= 1) g1[q] = -2.f * P.gamma * t * g;
          if (ORDER >= 2) g2[q] = (4.f * P.gamma * P.gamma * t * t - 2.f * P.gamma) * g;
        }
        if (WG) {
          // [E, R] rows of the 8 edges of this warp are one contiguous range: transpose through the unused tail of
          // the H region (behind the basis planes) and write 128-byte lines
          float* rs = reinterpret_cast<float*>(hplanes + 9 * kKPlane) + warp * 256;
          const int el = lane >> 2;
          const int first = 8 * warp;                                // first edge of this warp within the tile
          const int cnt = min(8, nvalid - first) * P.R;              // floats to write (<= 0: nothing)
#pragma unroll
          for (int which = 0; which < (MODE == 3 ? 2 : 1); ++which) {
            const float sc = which ? tt : 1.f;
            const float* g = which ? g1 : g0;
            if (which) __syncwarp();
            *reinterpret_cast<float4*>(rs + el * 32 + 8 * k8) = make_float4(sc * g[0], sc * g[1], sc * g[2], sc * g[3]);
            *reinterpret_cast<float4*>(rs + el * 32 + 8 * k8 + 4) = make_float4(sc * g[4], sc * g[5], sc * g[6], sc * g[7]);
            __syncwarp();
            float* dst = (which ? P.wg_rbf1 : P.wg_rbf) + (e0 + first) * P.R;
            for (int f = lane; f < cnt; f += 32) {
              const int er = f / P.R, kr = f - er * P.R;
              dst[f] = rs[er * 32 + kr];
            }
          }
        }
        {
          const int slot = k8 * kPitchK + mn;
          store_unit3(g0, rplanes, slot, kKPlane);
          if (ORDER >= 1) store_unit3(g1, rplanes + 3 * kKPlane, slot, kKPlane);
          if (ORDER >= 2) store_unit3(g2, rplanes + 6 * kKPlane, slot, kKPlane);
        }
      }
      publish(&bar_rbf);
      CTL(5);

      // ---- activation chain: p^(q) (TMEM) -> h^(q) (planes: k = this thread's hidden unit, edges gbase .. + 31) ----
      wait_bar(&bar_g1, par);
      CTL(6);
      {
        float sg[32];
        // q = 0
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float pre[16], h[16];
          tmem_ld16(trow + hb + 32 * eg + 16 * half, pre);
#pragma unroll
          for (int i = 0; i < 16; ++i) ssp_eval(pre[i] + bias1, h[i], sg[16 * half + i]);
          if (MODE != 3) {
            store_unit3(h, hplanes, (4 * eg + 2 * half) * kF + u, kBigPlane);
            store_unit3(h + 8, hplanes, (4 * eg + 2 * half + 1) * kF + u, kBigPlane);
          }
          if (WG) {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (gbase + 16 * half + i < nvalid) P.wg_h[(e0 + gbase + 16 * half + i) * kF + u] = h[i];
          }
        }
        if (MODE != 3) publish(&bar_h[0]);
        CTL(7);
        if (ORDER >= 1) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            float p1[16], h[16];
            tmem_ld16(trow + hb + 128u + 32 * eg + 16 * half, p1);
#pragma unroll
            for (int i = 0; i < 16; ++i) h[i] = sg[16 * half + i] * p1[i];
            if (MODE != 3 && half == 0) wait_bar(&bar_g2[0], par);   // the previous order's MMAs have read the planes
            store_unit3(h, hplanes, (4 * eg + 2 * half) * kF + u, kBigPlane);
            store_unit3(h + 8, hplanes, (4 * eg + 2 * half + 1) * kF + u, kBigPlane);
            if (WG && MODE == 3) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (gbase + 16 * half + i < nvalid)
                  P.wg_hd[(e0 + gbase + 16 * half + i) * kF + u] = t_s[gbase + 16 * half + i] * h[i];
            }
          }
          publish(&bar_h[1]);
          CTL(8);
        }
        if (ORDER >= 2) {
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            float p1[16], p2[16], h[16];
            tmem_ld16(trow + 128u + 32 * eg + 16 * half, p1);
            tmem_ld16(trow + 256u + 32 * eg + 16 * half, p2);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float s1 = sg[16 * half + i];
              h[i] = s1 * (1.f - s1) * p1[i] * p1[i] + s1 * p2[i];
            }
            if (half == 0) wait_bar(&bar_g2[1], par);
            store_unit3(h, hplanes, (4 * eg + 2 * half) * kF + u, kBigPlane);
            store_unit3(h + 8, hplanes, (4 * eg + 2 * half + 1) * kF + u, kBigPlane);
          }
          publish(&bar_h[2]);
          CTL(9);
        }
      }

      // ---- epilogue: of the PREVIOUS tile when PIPE (it runs under this tile's last MMA, which was handed to the
      //      tensor core just above), of this tile otherwise ----
      if (PIPE && it > 0) epilogue(prev_e0, prev_nvalid, mb ^ 1, hb ^ 256u);
      if (!PIPE) {
        wait_bar(&bar_g2[ORDER], par);
        CTL(10);
        epilogue(e0, nvalid, 0, 0u);
      }

      // ---- weight-gradient operands: gp = (A W2^T) f'(p) (+ second-order terms), rows in CSR order ----
      if (WG) {
        wait_bar(&bar_g3[0], par);
        CTL(13);
        if (MODE == 3) {                                            // Bt (this thread's own stores) -> the A planes
#pragma unroll
          for (int sb = 0; sb < 4; ++sb) {
            float B8[8];
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) {
              const int i = gbase + 8 * sb + ii;
              B8[ii] = (i < nvalid) ? P.wg_Bt[(e0 + i) * kF + u] : 0.f;
            }
            store_unit3(B8, hplanes, (4 * eg + sb) * kF + u, kBigPlane);
          }
          publish(&bar_a[1]);
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float pre[16], gh[16], gp[16];
          tmem_ld16(trow + 32 * eg + 16 * half, pre);
          tmem_ld16(trow + kColG0 + 32 * eg + 16 * half, gh);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float f;
            ssp_eval(pre[i] + bias1, f, pre[i]);                    // pre[i] <- sigma(p)
            gp[i] = gh[i] * pre[i];
          }
          if (MODE == 3) {
            float p1[16], ghd[16];
            tmem_ld16(trow + 128u + 32 * eg + 16 * half, p1);
            if (half == 0) wait_bar(&bar_g3[1], par);
            tmem_ld16(trow + kColG1 + 32 * eg + 16 * half, ghd);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int e = gbase + 16 * half + i;
              const float s1 = pre[i];
              gp[i] += ghd[i] * (t_s[e] * p1[i]) * s1 * (1.f - s1);
              if (e < nvalid) P.wg_gpd[(e0 + e) * kF + u] = ghd[i] * s1;
            }
          }
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (gbase + 16 * half + i < nvalid) P.wg_gp[(e0 + gbase + 16 * half + i) * kF + u] = gp[i];
        }
      }
      CTL(14);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      if (PIPE || kEdgeOut || WG) asm volatile("bar.sync 1, %0;" ::"r"(kWorkers) : "memory");
      prev_e0 = e0;
      prev_nvalid = nvalid;
    }
    if (PIPE && it > 0) {                                           // the last tile's epilogue
      wait_bar(&bar_g2[ORDER], static_cast<uint32_t>((it - 1) & 1));
      CTL(10);
      epilogue(prev_e0, prev_nvalid, (it - 1) & 1, 256u * static_cast<uint32_t>((it - 1) & 1));
    }
  }
  CTL(15);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kWorkers / 32) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(static_cast<uint32_t>(kTmemCols))
                 : "memory");
  }
}

// Rows whose edges cross a 32-edge group: out[r] = last[g0] + first[g0 + 1] + ... + first[g1] (fixed order); rows without
// edges: 0.  One warp per row, 128-bit lanes.
__global__ void __launch_bounds__(256)
cfconv_fixup_kernel(const int32_t* __restrict__ rowptr, int64_t n_out, const float* __restrict__ ws,
                    float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int64_t r = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= n_out) return;
  const int start = __ldg(rowptr + r), end = __ldg(rowptr + r + 1);
  float4* dst = reinterpret_cast<float4*>(out + r * kF) + lane;
  if (start == end) {
    *dst = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const int g0 = start >> 5, g1 = (end - 1) >> 5;
  if (g0 == g1) return;
  const float4* w4 = reinterpret_cast<const float4*>(ws);
  float4 s = __ldg(w4 + (static_cast<int64_t>(g0) * 2 + 1) * (kF / 4) + lane);
  for (int g = g0 + 1; g <= g1; ++g) {
    const float4 v = __ldg(w4 + (static_cast<int64_t>(g) * 2) * (kF / 4) + lane);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  *dst = s;
}

// out[e, k] = scale[e] * d^order/dd^order exp(-gamma (d_e - offset - k step)^2): the operands of the weight-gradient
// GEMMs (which need the basis, unlike the kernels above) without a detour through the coordinates.
__global__ void __launch_bounds__(256)
gauss_basis_kernel(const float* __restrict__ dist, int64_t E, int R, float gamma, float offset, float step, int order,
                   const float* __restrict__ scale, float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= E * R) return;
  const int64_t e = i / R;
  const int k = static_cast<int>(i - e * R);
  const float t = __ldg(dist + e) - offset - step * k;
  const float g = expf(-gamma * t * t);
  float v = g;
  if (order == 1) v = -2.f * gamma * t * g;
  else if (order == 2) v = (4.f * gamma * gamma * t * t - 2.f * gamma) * g;
  if (scale) v *= __ldg(scale + e);
  out[i] = v;
}

template <int MODE, bool WG>
static int launch(const Params& P, const int32_t* rowptr, int64_t n_out, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] {
    cudaFuncSetAttribute(cfconv_kernel<MODE, WG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(kSmemBytes));
  });
  const int64_t ntiles = (P.E + kBM - 1) / kBM;
  int launches = 0;
  if (ntiles > 0) {
    const unsigned grid = static_cast<unsigned>(std::min<int64_t>(ntiles, kNumSM));
    launch_pdl(cfconv_kernel<MODE, WG>, dim3(grid), dim3(kCtaThreads), kSmemBytes, st, P);
    ++launches;
  }
  const int64_t threads = n_out * 32;
  launch_pdl(cfconv_fixup_kernel, dim3(static_cast<unsigned>((threads + 255) / 256)), dim3(256), 0, st, rowptr, n_out,
             static_cast<const float*>(P.ws), P.out);
  return finish_launch(launches + 1);
}

}  // namespace cf
}  // namespace kgcnn

extern "C" {

size_t kgcnn_cfconv_fused_workspace_bytes(int64_t E) {
  return static_cast<size_t>((E + 31) / 32) * 2 * kgcnn::cf::kF * sizeof(float) + 16;
}

int kgcnn_gauss_basis(const float* dist, int64_t E, int R, float gamma, float offset, float step, int order,
                      const float* scale, float* out, kgcnn_stream_t stream) {
  using namespace kgcnn;
  if (E < 0 || R < 1 || order < 0 || order > 2) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!dist || !out) return KGCNN_E_INVALID;
  const int64_t n = E * R;
  cf::gauss_basis_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, as_stream(stream)>>>(
      dist, E, R, gamma, offset, step, order, scale, out);
  return finish_launch();
}

int kgcnn_cfconv_fused(int mode, const float* dist, int64_t E, int R, float gamma, float offset, float step,
                       const float* W1, const float* b1, const float* W2, const float* b2, int units, int act,
                       int static_weights, const int32_t* rowptr, const int32_t* rows, const int32_t* other, const int32_t* eperm,
                       int64_t n_out, const float* a, const float* b, const float* c, const float* tang, float* out,
                       float* edge_out, float* const* wg, void* workspace, size_t workspace_bytes,
                       kgcnn_stream_t stream) {
  using namespace kgcnn;
  if (E < 0 || n_out < 0 || mode < 0 || mode > 3) return KGCNN_E_INVALID;
  if (n_out == 0) return KGCNN_OK;
  if (!rowptr || !out || !W1 || !W2) return KGCNN_E_INVALID;
  if (E > 0 && (!dist || !rows || !other || !a || !workspace)) return KGCNN_E_INVALID;
  if (E > 0 && mode >= 1 && !b) return KGCNN_E_INVALID;
  if (E > 0 && mode >= 2 && !tang) return KGCNN_E_INVALID;
  if (E > 0 && mode == 3 && !c) return KGCNN_E_INVALID;
  if (E > 0 && (mode == 1 || mode == 3) && !edge_out) return KGCNN_E_INVALID;
  if (wg && mode != 1 && mode != 3) return KGCNN_E_INVALID;
  if (units != cf::kF || R < 1 || R > tc::kKC) return KGCNN_E_UNSUPPORTED;
  if (act != KGCNN_ACT_SHIFTED_SOFTPLUS && act != KGCNN_ACT_SOFTPLUS2) return KGCNN_E_UNSUPPORTED;
  if (E >= (1ll << 31) - 256) return KGCNN_E_UNSUPPORTED;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al16(W1) || !al16(W2) || !al16(out) || !al16(workspace)) return KGCNN_E_UNSUPPORTED;
  if (workspace_bytes < kgcnn_cfconv_fused_workspace_bytes(E)) return KGCNN_E_WORKSPACE;
  cf::Params P;
  P.dist = dist; P.W1 = W1; P.b1 = b1; P.W2 = W2; P.b2 = b2;
  P.rows = rows; P.other = other; P.eperm = eperm;
  P.a = a; P.b = b; P.c = c; P.tang = tang;
  P.out = out; P.edge_out = edge_out; P.ws = static_cast<float*>(workspace);
  P.wg_rbf = P.wg_rbf1 = P.wg_h = P.wg_A = P.wg_gp = P.wg_hd = P.wg_Bt = P.wg_gpd = nullptr;
  if (wg) {
    P.wg_rbf = wg[0]; P.wg_h = wg[1]; P.wg_A = wg[2]; P.wg_gp = wg[3];
    if (E > 0 && (!P.wg_rbf || !P.wg_h || !P.wg_A || !P.wg_gp)) return KGCNN_E_INVALID;
    if (mode == 3) {
      P.wg_rbf1 = wg[4]; P.wg_hd = wg[5]; P.wg_Bt = wg[6]; P.wg_gpd = wg[7];
      if (E > 0 && (!P.wg_rbf1 || !P.wg_hd || !P.wg_Bt || !P.wg_gpd)) return KGCNN_E_INVALID;
    }
  }
  P.static_w = static_weights ? 1 : 0;
  P.E = E; P.R = R; P.gamma = gamma; P.offset = offset; P.step = step;
  cudaStream_t st = as_stream(stream);
  switch (mode) {
    case 0: return cf::launch<0, false>(P, rowptr, n_out, st);
    case 1: return wg ? cf::launch<1, true>(P, rowptr, n_out, st) : cf::launch<1, false>(P, rowptr, n_out, st);
    case 2: return cf::launch<2, false>(P, rowptr, n_out, st);
    default: return wg ? cf::launch<3, true>(P, rowptr, n_out, st) : cf::launch<3, false>(P, rowptr, n_out, st);
  }
}

}  // extern "C"
