// Fused PaiNN equivariant message (kgcnn/layers/conv/painn_conv.py:97-115) and all of its derivatives.
//
// The reference computes, per edge e = (receiver i, sender j), with phi = Dense(s) [N,3F], w = Dense(rbf)*env [E,3F],
// v [N,3,F], r_hat [E,3]:
//     sw = phi[j] * w[e]  -> (sw1, sw2, sw3);  ds_i = sum_e sw1;  dv_i[c] = sum_e (sw2 * v[j,c] + sw3 * r[e,c])
// as ~12 TF ops that materialise five [E,3F]-sized tensors (684 MB each at config C3) and sort twice.
//
// Here the message is the multilinear form
//     M(Gs,Gv; Phi; V; W; R) = sum_e sum_f [ Gs[i,f] P1[j,f] W1[e,f]                                   (term 1)
//                                          + sum_c Gv[i,c,f] P2[j,f] V[j,c,f] W2[e,f]                  (term 2)
//                                          + sum_c Gv[i,c,f] P3[j,f] W3[e,f] R[e,c] ]                  (term 3)
// and every tensor that force training needs (forward, backward, double backward) is the partial derivative of M
// with respect to one operand with some operands replaced by cotangents.  Three kernels cover all of them:
//     nodes   : dM/d(Gs,Gv)   = (ds, dv)       one warp per receiver, walks its CSR segment
//     senders : dM/d(Phi,V)                    one warp per sender, walks its CSC segment
//     edges   : dM/d(W,R)                      one warp per edge
// `mask` (bit 0, 1, 2 = term 1, 2, 3) selects the active terms; output blocks of inactive terms are zero.
// No [E,*] intermediate reaches HBM; deterministic (segment order = original edge order, no atomics).
#include "common.cuh"

namespace kgcnn {
namespace painn {

constexpr int kWarps = 8;

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4s(const float* p) { return ldg_stream(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 mul4(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 scl4(float a, float4 b) { return make_float4(a * b.x, a * b.y, a * b.z, a * b.w); }
__device__ __forceinline__ void fma4(float4 a, float4 b, float4& c) {
  c.x = fmaf(a.x, b.x, c.x); c.y = fmaf(a.y, b.y, c.y); c.z = fmaf(a.z, b.z, c.z); c.w = fmaf(a.w, b.w, c.w);
}
__device__ __forceinline__ void fma4s(float a, float4 b, float4& c) {
  c.x = fmaf(a, b.x, c.x); c.y = fmaf(a, b.y, c.y); c.z = fmaf(a, b.z, c.z); c.w = fmaf(a, b.w, c.w);
}
__device__ __forceinline__ float hsum4(float4 a) { return (a.x + a.y) + (a.z + a.w); }
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
constexpr float4 kZero4 = {0.f, 0.f, 0.f, 0.f};

// ---------------------------------------------------------------------------------------------------------
// nodes: ds[i,f] = sum_e P1[j,f] W1[e,f];  dv[i,c,f] = sum_e P2[j,f] W2[e,f] V[j,c,f] + P3[j,f] W3[e,f] R[e,c]
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarps * 32)
nodes_kernel(const float* __restrict__ phi, const float* __restrict__ v, const float* __restrict__ w,
             const float* __restrict__ r, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
             const int32_t* __restrict__ nbr, int nbr_permuted, int64_t N, int F, int mask, float* __restrict__ ds,
             float* __restrict__ dv) {
  const int lane = threadIdx.x & 31;
  const int64_t i = static_cast<int64_t>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  if (i >= N) return;
  const int beg = __ldg(rowptr + i), end = __ldg(rowptr + i + 1);
  const int64_t F3 = 3 * static_cast<int64_t>(F);
  for (int f0 = lane * 4; f0 < F; f0 += 128) {
    float4 as = kZero4, a0 = kZero4, a1 = kZero4, a2 = kZero4;
#pragma unroll 2
    for (int p = beg; p < end; ++p) {
      const int e = perm ? __ldg(perm + p) : p;
      const int j = __ldg(nbr + (nbr_permuted ? p : e));
      const float* wr = w + e * F3 + f0;
      const float* pr = phi + j * F3 + f0;
      if (mask & 1) fma4(ld4(pr), ld4s(wr), as);
      if (mask & 2) {
        const float4 t = mul4(ld4(pr + F), ld4s(wr + F));
        const float* vr = v + j * F3 + f0;
        fma4(t, ld4(vr), a0);
        fma4(t, ld4(vr + F), a1);
        fma4(t, ld4(vr + 2 * F), a2);
      }
      if (mask & 4) {
        const float4 t = mul4(ld4(pr + 2 * F), ld4s(wr + 2 * F));
        const float* rr = r + 3 * static_cast<int64_t>(e);
        fma4s(__ldg(rr), t, a0);
        fma4s(__ldg(rr + 1), t, a1);
        fma4s(__ldg(rr + 2), t, a2);
      }
    }
    if (ds) st4(ds + i * F + f0, as);
    if (dv) {
      st4(dv + i * F3 + f0, a0);
      st4(dv + i * F3 + F + f0, a1);
      st4(dv + i * F3 + 2 * F + f0, a2);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// senders: gphi[j] = ( sum_e W1 Gs[i],  sum_c V[j,c] * A[c],  sum_e W3 sum_c R[e,c] Gv[i,c] ),
//          gvv[j,c] = P2[j] * A[c],     A[c] = sum_e W2[e] Gv[i,c]          (segments of the SENDER relation)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarps * 32)
senders_kernel(const float* __restrict__ gs, const float* __restrict__ gv, const float* __restrict__ phi,
               const float* __restrict__ v, const float* __restrict__ w, const float* __restrict__ r,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm, const int32_t* __restrict__ nbr,
               int nbr_permuted, int64_t N, int F, int mask, float* __restrict__ gphi, float* __restrict__ gvv) {
  const int lane = threadIdx.x & 31;
  const int64_t j = static_cast<int64_t>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  if (j >= N) return;
  const int beg = __ldg(rowptr + j), end = __ldg(rowptr + j + 1);
  const int64_t F3 = 3 * static_cast<int64_t>(F);
  const bool t2 = (mask & 2) != 0, t3 = (mask & 4) != 0 && gphi != nullptr;
  for (int f0 = lane * 4; f0 < F; f0 += 128) {
    float4 c1 = kZero4, a0 = kZero4, a1 = kZero4, a2 = kZero4, b = kZero4;
#pragma unroll 2
    for (int p = beg; p < end; ++p) {
      const int e = perm ? __ldg(perm + p) : p;
      const int i = __ldg(nbr + (nbr_permuted ? p : e));
      const float* wr = w + e * F3 + f0;
      if ((mask & 1) && gphi) fma4(ld4s(wr), ld4(gs + static_cast<int64_t>(i) * F + f0), c1);
      if (t2 || t3) {
        const float* gr = gv + i * F3 + f0;
        const float4 g0 = ld4(gr), g1 = ld4(gr + F), g2 = ld4(gr + 2 * F);
        if (t2) {
          const float4 w2 = ld4s(wr + F);
          fma4(w2, g0, a0);
          fma4(w2, g1, a1);
          fma4(w2, g2, a2);
        }
        if (t3) {
          const float* rr = r + 3 * static_cast<int64_t>(e);
          float4 t = scl4(__ldg(rr), g0);
          fma4s(__ldg(rr + 1), g1, t);
          fma4s(__ldg(rr + 2), g2, t);
          fma4(ld4s(wr + 2 * F), t, b);
        }
      }
    }
    if (gphi) {
      float4 p2 = kZero4;
      if (t2) {
        const float* vr = v + j * F3 + f0;
        p2 = mul4(a0, ld4(vr));
        fma4(a1, ld4(vr + F), p2);
        fma4(a2, ld4(vr + 2 * F), p2);
      }
      st4(gphi + j * F3 + f0, c1);
      st4(gphi + j * F3 + F + f0, p2);
      st4(gphi + j * F3 + 2 * F + f0, b);
    }
    if (gvv) {
      const float4 p2 = t2 ? ld4(phi + j * F3 + F + f0) : kZero4;
      st4(gvv + j * F3 + f0, mul4(p2, a0));
      st4(gvv + j * F3 + F + f0, mul4(p2, a1));
      st4(gvv + j * F3 + 2 * F + f0, mul4(p2, a2));
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// edges: gw[e] = ( P1[j] Gs[i],  P2[j] sum_c Gv[i,c] V[j,c],  P3[j] sum_c R[e,c] Gv[i,c] )     [E,3F]
//        gr[e,c] = sum_f P3[j,f] W3[e,f] Gv[i,c,f]                                              [E,3]
// ---------------------------------------------------------------------------------------------------------
template <bool ACC>
__device__ __forceinline__ void
edges_body(const float* __restrict__ gs, const float* __restrict__ gv, const float* __restrict__ phi,
             const float* __restrict__ v, const float* __restrict__ w, const float* __restrict__ r,
             const int32_t* __restrict__ recv, const int32_t* __restrict__ send, int64_t E, int F, int mask,
             const float* __restrict__ gw_acc, float* __restrict__ gw, float* __restrict__ gr) {
  const int lane = threadIdx.x & 31;
  const int64_t e = static_cast<int64_t>(blockIdx.x) * kWarps + (threadIdx.x >> 5);
  if (e >= E) return;
  const int64_t i = __ldg(recv + e), j = __ldg(send + e);
  const int64_t F3 = 3 * static_cast<int64_t>(F);
  const bool t1 = (mask & 1) && gw, t2 = (mask & 2) && gw, t3 = (mask & 4) != 0;
  float r0 = 0.f, r1 = 0.f, r2 = 0.f;
  if (t3 && gw) { r0 = __ldg(r + 3 * e); r1 = __ldg(r + 3 * e + 1); r2 = __ldg(r + 3 * e + 2); }
  float q0 = 0.f, q1 = 0.f, q2 = 0.f;
  for (int f0 = lane * 4; f0 < F; f0 += 128) {
    const float* pr = phi + j * F3 + f0;
    float4 o1 = kZero4, o2 = kZero4, o3 = kZero4;
    if (t1) o1 = mul4(ld4(pr), ld4(gs + i * F + f0));
    if (t2 || t3) {
      const float* gp = gv + i * F3 + f0;
      const float4 g0 = ld4(gp), g1 = ld4(gp + F), g2 = ld4(gp + 2 * F);
      if (t2) {
        const float* vr = v + j * F3 + f0;
        float4 d = mul4(g0, ld4(vr));
        fma4(g1, ld4(vr + F), d);
        fma4(g2, ld4(vr + 2 * F), d);
        o2 = mul4(ld4(pr + F), d);
      }
      if (t3) {
        const float4 p3 = ld4(pr + 2 * F);
        if (gw) {
          float4 t = scl4(r0, g0);
          fma4s(r1, g1, t);
          fma4s(r2, g2, t);
          o3 = mul4(p3, t);
        }
        if (gr) {
          const float4 q = mul4(p3, ld4s(w + e * F3 + 2 * F + f0));
          q0 += hsum4(mul4(q, g0));
          q1 += hsum4(mul4(q, g1));
          q2 += hsum4(mul4(q, g2));
        }
      }
    }
    if (gw) {
      if (ACC) {             // sum of two substitutions of one backward node: the second call adds onto the first
        const float* ap = gw_acc + e * F3 + f0;
        const float4 a1 = ld4s(ap), a2 = ld4s(ap + F), a3 = ld4s(ap + 2 * F);
        o1.x += a1.x; o1.y += a1.y; o1.z += a1.z; o1.w += a1.w;
        o2.x += a2.x; o2.y += a2.y; o2.z += a2.z; o2.w += a2.w;
        o3.x += a3.x; o3.y += a3.y; o3.z += a3.z; o3.w += a3.w;
      }
      st4(gw + e * F3 + f0, o1);
      st4(gw + e * F3 + F + f0, o2);
      st4(gw + e * F3 + 2 * F + f0, o3);
    }
  }
  if (gr) {
    q0 = warp_sum(q0); q1 = warp_sum(q1); q2 = warp_sum(q2);
    if (lane == 0) { gr[3 * e] = q0; gr[3 * e + 1] = q1; gr[3 * e + 2] = q2; }
  }
}

#define KGCNN_EDGES_PARAMS                                                                                       \
  const float *__restrict__ gs, const float *__restrict__ gv, const float *__restrict__ phi,                     \
      const float *__restrict__ v, const float *__restrict__ w, const float *__restrict__ r,                      \
      const int32_t *__restrict__ recv, const int32_t *__restrict__ send, int64_t E, int F, int mask,             \
      const float *__restrict__ gw_acc, float *__restrict__ gw, float *__restrict__ gr
__global__ void __launch_bounds__(kWarps * 32) edges_kernel(KGCNN_EDGES_PARAMS) {
  edges_body<false>(gs, gv, phi, v, w, r, recv, send, E, F, mask, nullptr, gw, gr);
}
__global__ void __launch_bounds__(kWarps * 32, 3) edges_acc_kernel(KGCNN_EDGES_PARAMS) {   // 3 CTAs / SM like the plain one
  edges_body<true>(gs, gv, phi, v, w, r, recv, send, E, F, mask, gw_acc, gw, gr);
}
#undef KGCNN_EDGES_PARAMS

}  // namespace painn
}  // namespace kgcnn

using namespace kgcnn;

namespace {
inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline bool ok16(const void* p) { return p == nullptr || al16(p); }
}  // namespace

extern "C" {

int kgcnn_painn_msg_nodes(const float* phi, const float* v, const float* w, const float* r, const int32_t* rowptr,
                          const int32_t* perm, const int32_t* nbr, int nbr_permuted, int64_t N, int64_t E, int F,
                          int mask, float* ds, float* dv, kgcnn_stream_t stream) {
  if (N < 0 || E < 0 || F <= 0 || (F & 3) || (mask & ~7)) return KGCNN_E_INVALID;
  if (N == 0 || (!ds && !dv)) return KGCNN_OK;
  if (!rowptr || (E > 0 && (!nbr || !phi || !w)) || ((mask & 2) && !v) || ((mask & 4) && !r)) return KGCNN_E_INVALID;
  if (!ok16(phi) || !ok16(v) || !ok16(w) || !ok16(ds) || !ok16(dv)) return KGCNN_E_INVALID;
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(N, painn::kWarps));
  painn::nodes_kernel<<<grid, painn::kWarps * 32, 0, as_stream(stream)>>>(phi, v, w, r, rowptr, perm, nbr,
                                                                          nbr_permuted, N, F, mask, ds, dv);
  return finish_launch();
}

int kgcnn_painn_msg_senders(const float* gs, const float* gv, const float* phi, const float* v, const float* w,
                            const float* r, const int32_t* rowptr, const int32_t* perm, const int32_t* nbr,
                            int nbr_permuted, int64_t N, int64_t E, int F, int mask, float* gphi, float* gvv,
                            kgcnn_stream_t stream) {
  if (N < 0 || E < 0 || F <= 0 || (F & 3) || (mask & ~7)) return KGCNN_E_INVALID;
  if (N == 0 || (!gphi && !gvv)) return KGCNN_OK;
  if (!rowptr || (E > 0 && (!nbr || !w))) return KGCNN_E_INVALID;
  if ((mask & 1) && gphi && !gs) return KGCNN_E_INVALID;
  if ((mask & 6) && !gv) return KGCNN_E_INVALID;
  if ((mask & 2) && ((gphi && !v) || (gvv && !phi))) return KGCNN_E_INVALID;
  if ((mask & 4) && gphi && !r) return KGCNN_E_INVALID;
  if (!ok16(gs) || !ok16(gv) || !ok16(phi) || !ok16(v) || !ok16(w) || !ok16(gphi) || !ok16(gvv)) return KGCNN_E_INVALID;
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(N, painn::kWarps));
  painn::senders_kernel<<<grid, painn::kWarps * 32, 0, as_stream(stream)>>>(gs, gv, phi, v, w, r, rowptr, perm, nbr,
                                                                            nbr_permuted, N, F, mask, gphi, gvv);
  return finish_launch();
}

int kgcnn_painn_msg_edges(const float* gs, const float* gv, const float* phi, const float* v, const float* w,
                          const float* r, const int32_t* recv, const int32_t* send, int64_t E, int F, int mask,
                          const float* gw_acc, float* gw, float* gr, kgcnn_stream_t stream) {
  if (E < 0 || F <= 0 || (F & 3) || (mask & ~7)) return KGCNN_E_INVALID;
  if (E == 0 || (!gw && !gr)) return KGCNN_OK;
  if (!recv || !send || !phi) return KGCNN_E_INVALID;
  if ((mask & 1) && gw && !gs) return KGCNN_E_INVALID;
  if ((mask & 6) && !gv) return KGCNN_E_INVALID;
  if ((mask & 2) && gw && !v) return KGCNN_E_INVALID;
  if ((mask & 4) && ((gw && !r) || (gr && !w))) return KGCNN_E_INVALID;
  if (!ok16(gs) || !ok16(gv) || !ok16(phi) || !ok16(v) || !ok16(w) || !ok16(gw) || !ok16(gw_acc)) return KGCNN_E_INVALID;
  if (gw_acc && (!gw || gw_acc == gw)) return KGCNN_E_INVALID;
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(E, painn::kWarps));
  if (gw_acc)
    painn::edges_acc_kernel<<<grid, painn::kWarps * 32, 0, as_stream(stream)>>>(gs, gv, phi, v, w, r, recv, send, E,
                                                                                  F, mask, gw_acc, gw, gr);
  else
    painn::edges_kernel<<<grid, painn::kWarps * 32, 0, as_stream(stream)>>>(gs, gv, phi, v, w, r, recv, send, E,
                                                                                   F, mask, nullptr, gw, gr);
  return finish_launch();
}

}  // extern "C"
