// Fused atom-centred symmetry functions (ACSF) of HDNNP: radial G2 and angular G4, each as forward G, VJP
// (dG -> dxyz) and JVP (u -> dG).  Replaces the ~25-op TF chains of
//   kgcnn/layers/conv/acsf_conv.py:199-211 (ACSFG2.call) and :473-494 (ACSFG4.call)
// which materialise the per-bond parameter gather [M,m,3] / [A,m,4] and >= 10 tensors of shape [A,m] (2 GB each at
// batch 256), and end in tensor_scatter_nd_add (atomic on GPU TF, kgcnn/layers/pooling.py:658-663).
// Here nothing of size [pairs|triples, m] ever reaches HBM: one warp owns one centre atom, walks its CSR segment
// of pairs / triples in the original order, lanes run over the m parameter sets, and the (relation, parameter)
// accumulators live in shared memory private to the warp -> deterministic, no atomics.
// The parameters are not trainable in the drivers (param_trainable=False, acsf_conv.py:62,274), so force
// training needs exactly these three maps: G, J^T v and J u.
#include "common.cuh"

namespace kgcnn {
namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxAcc = 1024;   // R*m accumulators per warp in shared memory (G4 of the drivers: 10*50 = 500)
constexpr float kPi = 3.14159265358979323846f;

enum { kFwd = 0, kVjp = 1, kJvp = 2 };

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 ld3(const float* p) { return {__ldg(p), __ldg(p + 1), __ldg(p + 2)}; }
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 axpy(float a, V3 x, V3 y) { return {a * x.x + y.x, a * x.y + y.y, a * x.z + y.z}; }
__device__ __forceinline__ V3 scale(float a, V3 x) { return {a * x.x, a * x.y, a * x.z}; }
__device__ __forceinline__ float norm_ref(V3 v, float eps) {
  // EuclideanNorm: sqrt(relu(sum(v^2)) + eps)   (kgcnn/layers/geom.py:181-187)
  return sqrtf(fmaxf(dot(v, v), 0.f) + eps);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// base^zeta for base >= 0 (pow semantics incl. 0^0 = 1): small integer exponents (the zeta = 1, 2, 4, 8, 16 of
// force_hdnnp2nd.py:52) by branch-free square-and-multiply, exact to fp32 rounding and a third of the instructions
// of exp2f(zeta * log2f(base)); anything else through exp2/log2.
__device__ __forceinline__ float pow_nonneg(float base, float zeta) {
  const int zi = static_cast<int>(zeta);
  if (zeta == static_cast<float>(zi) && zi >= 0 && zi < 32) {
    float b = base, p = (zi & 1) ? b : 1.f;
    b *= b; p *= (zi & 2) ? b : 1.f;
    b *= b; p *= (zi & 4) ? b : 1.f;
    b *= b; p *= (zi & 8) ? b : 1.f;
    b *= b; p *= (zi & 16) ? b : 1.f;
    return p;
  }
  return (base > 0.f) ? exp2f(zeta * log2f(base)) : (zeta == 0.f ? 1.f : 0.f);
}
// f_c(r) = 0.5 (cos(pi clip(r,-Rc,Rc)/Rc) + 1) and its derivative (0 outside the clip)   acsf_conv.py:159-166
__device__ __forceinline__ void cutoff(float r, float rc, float& f, float& df) {
  const float rr = fminf(fmaxf(r, -rc), rc);
  float s, c;
  sincosf(rr * kPi / rc, &s, &c);
  f = 0.5f * (c + 1.f);
  df = (r > -rc && r < rc) ? -0.5f * kPi / rc * s : 0.f;
}

// ---------------------------------------------------------------------------------------------------------
// G2:  G[i, rel(j), k] = sum_j exp(-eta_k (r_ij - mu_k)^2) f_c(r_ij; Rc_k)
// ---------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
acsf_g2_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ rel_atom,
               const int32_t* __restrict__ tset_atom, const int32_t* __restrict__ rowptr,
               const int32_t* __restrict__ perm, const int32_t* __restrict__ send, int send_permuted,
               const float* __restrict__ table, int R, int m, float eps, int64_t N,
               const float* __restrict__ weight, const float* __restrict__ cot, float* __restrict__ out,
               float* __restrict__ edge_out) {
  __shared__ float acc_s[kWarpsPerBlock][kMaxAcc];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = static_cast<int64_t>(blockIdx.x) * kWarpsPerBlock + warp;
  if (i >= N) return;
  float* acc = acc_s[warp];
  const int RM = R * m;
  if (MODE == kVjp) for (int t = lane; t < RM; t += 32) acc[t] = __ldg(cot + i * RM + t);   // dG[i, :, :]
  else for (int t = lane; t < RM; t += 32) acc[t] = 0.f;
  __syncwarp();
  const V3 xi = ld3(xyz + 3 * i);
  V3 ui = {0.f, 0.f, 0.f};
  if (MODE == kJvp) ui = ld3(cot + 3 * i);
  const float* tab_i = table + (tset_atom ? static_cast<int64_t>(max(__ldg(tset_atom + i), 0)) * RM * 3 : 0);
  const bool centre_ok = !tset_atom || __ldg(tset_atom + i) >= 0;
  V3 gi = {0.f, 0.f, 0.f};
  const int beg = __ldg(rowptr + i), end = __ldg(rowptr + i + 1);
  for (int p = beg; p < end; ++p) {
    const int e = perm ? __ldg(perm + p) : p;
    const int j = __ldg(send + (send_permuted ? p : e));
    const int r = __ldg(rel_atom + j);
    V3 gj = {0.f, 0.f, 0.f};
    if (r >= 0 && r < R && centre_ok) {               // unmapped neighbour elements contribute nothing
      const V3 v = sub(xi, ld3(xyz + 3 * j));
      const float d = norm_ref(v, eps);
      const float we = weight ? __ldg(weight + e) : 1.f;     // wACSF: g(Z_j) or an external pair weight
      float ddot = 0.f;
      if (MODE == kJvp) ddot = dot(v, sub(ui, ld3(cot + 3 * j))) / d;
      float a = 0.f;                                   // VJP: sum_k dG * dphi/dd
      for (int k = lane; k < m; k += 32) {
        const float* prm = tab_i + (static_cast<int64_t>(r) * m + k) * 3;
        const float eta = __ldg(prm), mu = __ldg(prm + 1), rc = __ldg(prm + 2);
        float f, df;
        cutoff(d, rc, f, df);
        const float t = d - mu;
        const float g = expf(-eta * t * t);
        if (MODE == kFwd) acc[r * m + k] += we * g * f;
        else {
          const float dphi = we * g * (-2.f * eta * t * f + df);
          if (MODE == kVjp) a += acc[r * m + k] * dphi;
          else acc[r * m + k] += dphi * ddot;
        }
      }
      if (MODE == kVjp) {
        a = warp_sum(a);
        const V3 gv = scale(a / d, v);                 // d/dx_i = +gv, d/dx_j = -gv
        gi = axpy(1.f, gv, gi);
        gj = scale(-1.f, gv);
      }
    }
    if (MODE == kVjp && lane == 0) {
      edge_out[3 * static_cast<int64_t>(e) + 0] = gj.x;
      edge_out[3 * static_cast<int64_t>(e) + 1] = gj.y;
      edge_out[3 * static_cast<int64_t>(e) + 2] = gj.z;
    }
  }
  __syncwarp();
  if (MODE == kVjp) {
    if (lane == 0) { out[3 * i] = gi.x; out[3 * i + 1] = gi.y; out[3 * i + 2] = gi.z; }
  } else {
    for (int t = lane; t < RM; t += 32) out[i * RM + t] = acc[t];
  }
}

// ---------------------------------------------------------------------------------------------------------
// G4:  G[i, pair(Z_j,Z_k), p] = sum_{jk} 2^(1-zeta)(1+lambda cos)^zeta / mult * exp(-eta (rij^2+rik^2+rjk^2)) fc fc fc
//      (acsf_conv.py:419-446, 473-494)
// ---------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
acsf_g4_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ z_atom,
               const int32_t* __restrict__ tset_atom, const int32_t* __restrict__ rev_pair, int zmax,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
               const int32_t* __restrict__ idx_j, const int32_t* __restrict__ idx_k,
               const float* __restrict__ table, int R, int m, float eps, float inv_mult, int uniform_rc, int64_t N,
               const float* __restrict__ mu_tab, const float* __restrict__ weight,
               const float2* __restrict__ fac, const int32_t* __restrict__ fac_idx,
               const float* __restrict__ cot, float* __restrict__ out, float* __restrict__ triple_out,
               int64_t n_triples) {
  __shared__ float acc_s[kWarpsPerBlock][kMaxAcc];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = static_cast<int64_t>(blockIdx.x) * kWarpsPerBlock + warp;
  if (i >= N) return;
  float* acc = acc_s[warp];
  const int RM = R * m;
  if (MODE == kVjp) for (int t = lane; t < RM; t += 32) acc[t] = __ldg(cot + i * RM + t);
  else for (int t = lane; t < RM; t += 32) acc[t] = 0.f;
  __syncwarp();
  const V3 xi = ld3(xyz + 3 * i);
  V3 ui = {0.f, 0.f, 0.f};
  if (MODE == kJvp) ui = ld3(cot + 3 * i);
  const float* tab_i = table + (tset_atom ? static_cast<int64_t>(max(__ldg(tset_atom + i), 0)) * RM * 4 : 0);
  const float* mu_i = mu_tab ? mu_tab + (tset_atom ? static_cast<int64_t>(max(__ldg(tset_atom + i), 0)) * RM : 0)
                             : nullptr;
  const bool centre_ok = !tset_atom || __ldg(tset_atom + i) >= 0;
  const int tset_i = tset_atom ? max(__ldg(tset_atom + i), 0) : 0;
  const float2* fac_i = fac ? fac + static_cast<int64_t>(tset_i) * R * 32 : nullptr;
  const int32_t* idx_i = fac ? fac_idx + static_cast<int64_t>(tset_i) * RM : nullptr;
  V3 gi = {0.f, 0.f, 0.f};
  const int beg = __ldg(rowptr + i), end = __ldg(rowptr + i + 1);
  const float rc0 = uniform_rc ? __ldg(tab_i + 3) : 0.f;           // one cutoff radius for the whole table
  constexpr unsigned kFull = 0xffffffffu;
  // Triples are processed in batches of 32.  Phase A: lane t evaluates the geometry of triple p0 + t ONCE (distances,
  // cosine, cutoffs, tangents) instead of all 32 lanes repeating it.  Phase B: the warp walks the valid triples of
  // the batch in their original order, the owner lane broadcasts its scalars by shuffle and the lanes run over the m
  // parameter sets (accumulation order per (pair type, parameter) = triple order, as before).  Phase C (VJP): every
  // lane assembles the coordinate gradients of its own triple.
  for (int p0 = beg; p0 < end; p0 += 32) {
    const int p = p0 + lane;
    const bool have = p < end;
    int e = 0, j = 0, k = 0, r = -1;
    float rij = 1.f, rik = 1.f, rjk = 1.f, c = 0.f, wt = 0.f;
    float fij = 0.f, dfij = 0.f, fik = 0.f, dfik = 0.f, fjk = 0.f, dfjk = 0.f;
    float t_ij = 0.f, t_ik = 0.f, t_jk = 0.f, t_c = 0.f;             // JVP tangents of (rij, rik, rjk, cos)
    V3 vij = {0.f, 0.f, 0.f}, vik = vij, vjk = vij, dc_ij = vij, dc_ik = vij;
    if (have) {
      e = perm ? __ldg(perm + p) : p;
      j = __ldg(idx_j + e);
      k = __ldg(idx_k + e);
      const int zj = __ldg(z_atom + j), zk = __ldg(z_atom + k);
      if (zj >= 0 && zj < zmax && zk >= 0 && zk < zmax) r = __ldg(rev_pair + zj * zmax + zk);
      if (!(r >= 0 && r < R && centre_ok)) r = -1;
      if (r >= 0) {
        const V3 xj = ld3(xyz + 3 * j), xk = ld3(xyz + 3 * k);
        vij = sub(xi, xj); vik = sub(xi, xk); vjk = sub(xj, xk);
        rij = norm_ref(vij, eps); rik = norm_ref(vik, eps); rjk = norm_ref(vjk, eps);
        const float inv = 1.f / (rij * rik);
        c = dot(vij, vik) / rij / rik;                                   // acsf_conv.py:439
        wt = (weight ? __ldg(weight + e) : 1.f) * inv_mult;              // wACSF: h(Z_j, Z_k) or external
        // d cos / d vij = vik/(rij rik) - c vij/rij^2 ;  d cos / d vik = vij/(rij rik) - c vik/rik^2
        dc_ij = axpy(-c / (rij * rij), vij, scale(inv, vik));
        dc_ik = axpy(-c / (rik * rik), vik, scale(inv, vij));
        if (MODE == kJvp) {
          const V3 uj = ld3(cot + 3 * j), uk = ld3(cot + 3 * k);
          const V3 uij = sub(ui, uj), uik = sub(ui, uk), ujk = sub(uj, uk);
          t_ij = dot(vij, uij) / rij;
          t_ik = dot(vik, uik) / rik;
          t_jk = dot(vjk, ujk) / rjk;
          t_c = dot(dc_ij, uij) + dot(dc_ik, uik);
        }
        if (uniform_rc) {
          cutoff(rij, rc0, fij, dfij);
          cutoff(rik, rc0, fik, dfik);
          cutoff(rjk, rc0, fjk, dfjk);
        }
      }
    }
    float s_ij = 0.f, s_ik = 0.f, s_jk = 0.f, s_c = 0.f;             // VJP: sums over the parameter sets, own triple
    for (unsigned rem = __ballot_sync(kFull, r >= 0); rem; rem &= rem - 1) {
      const int t = __ffs(rem) - 1;
      const int rr = __shfl_sync(kFull, r, t);
      const float cc = __shfl_sync(kFull, c, t), wtt = __shfl_sync(kFull, wt, t);
      const float bij = __shfl_sync(kFull, rij, t), bik = __shfl_sync(kFull, rik, t), bjk = __shfl_sync(kFull, rjk, t);
      float gij = 0.f, gdij = 0.f, gik = 0.f, gdik = 0.f, gjk = 0.f, gdjk = 0.f;
      if (uniform_rc) {
        gij = __shfl_sync(kFull, fij, t); gik = __shfl_sync(kFull, fik, t); gjk = __shfl_sync(kFull, fjk, t);
        if (MODE != kFwd) {
          gdij = __shfl_sync(kFull, dfij, t); gdik = __shfl_sync(kFull, dfik, t); gdjk = __shfl_sync(kFull, dfjk, t);
        }
      }
      float u_ij = 0.f, u_ik = 0.f, u_jk = 0.f, u_c = 0.f;
      if (MODE == kJvp) {
        u_ij = __shfl_sync(kFull, t_ij, t); u_ik = __shfl_sync(kFull, t_ik, t);
        u_jk = __shfl_sync(kFull, t_jk, t); u_c = __shfl_sync(kFull, t_c, t);
      }
      const float r2 = bij * bij + bik * bik + bjk * bjk;
      float a_ij = 0.f, a_ik = 0.f, a_jk = 0.f, a_c = 0.f;
      if (fac_i) {
        // Factorised parameter table (uniform cutoff): the m parameter sets of a pair type are products of few
        // distinct Gaussians exp(-eta sum (r - mu)^2) (slots 0-15) and few distinct angular terms
        // 2^(1-zeta) (1 + lambda cos)^zeta (slots 16-31), e.g. 5 and 10 for the 50 sets of force_hdnnp2nd.py:49-55.
        // Lane s evaluates slot s once per triple; each parameter set then needs two shuffles and a product
        // instead of three transcendental chains.
        const float fff = gij * gik * gjk;
        const float2 key = __ldg(fac_i + rr * 32 + lane);
        float val, dval = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
        if (lane < 16) {                                                 // key = (eta, mu)
          const float sij = bij - key.y, sik = bik - key.y, sjk = bjk - key.y;
          val = expf(-key.x * (mu_i ? sij * sij + sik * sik + sjk * sjk : r2));
          if (MODE != kFwd) {                                            // d(E fff)/d r_ab
            d1 = val * (-2.f * key.x * sij * fff + gdij * gik * gjk);
            d2 = val * (-2.f * key.x * sik * fff + gij * gdik * gjk);
            d3 = val * (-2.f * key.x * sjk * fff + gij * gik * gdjk);
            if (MODE == kJvp) d1 = d1 * u_ij + d2 * u_ik + d3 * u_jk;
          }
          val *= fff;
        } else {                                                         // key = (zeta, lambda)
          const float base = fmaxf(cc * key.y + 1.f, 0.f);
          const float pre = exp2f(1.f - key.x) * wtt;
          const float pw = pow_nonneg(base, key.x);
          val = pre * pw;
          if (MODE != kFwd) {
            const float pw1 = (base > 0.f) ? pw / base : (key.x == 1.f ? 1.f : 0.f);
            dval = pre * key.x * key.y * pw1;                            // dA / dcos
            if (MODE == kJvp) dval *= u_c;
          }
        }
        for (int q0 = 0; q0 < m; q0 += 32) {                             // uniform trip count: shuffles inside
          const int q = q0 + lane;
          const bool on = q < m;
          const int packed = on ? __ldg(idx_i + rr * m + q) : 0x1000;
          const int se = packed & 0xff, sp = packed >> 8;
          const float Ef = __shfl_sync(kFull, val, se);                  // E * fff
          const float P = __shfl_sync(kFull, val, sp);                   // A
          if (MODE == kFwd) {
            if (on) acc[rr * m + q] += P * Ef;
          } else if (MODE == kJvp) {
            const float W = __shfl_sync(kFull, d1, se), dP = __shfl_sync(kFull, dval, sp);
            if (on) acc[rr * m + q] += P * W + dP * Ef;
          } else {
            const float D1 = __shfl_sync(kFull, d1, se), D2 = __shfl_sync(kFull, d2, se);
            const float D3 = __shfl_sync(kFull, d3, se), dP = __shfl_sync(kFull, dval, sp);
            if (on) {
              const float g = acc[rr * m + q], gP = g * P;
              a_ij += gP * D1; a_ik += gP * D2; a_jk += gP * D3; a_c += g * dP * Ef;
            }
          }
        }
      } else
      for (int q = lane; q < m; q += 32) {
        const float4 prm = __ldg(reinterpret_cast<const float4*>(tab_i + (static_cast<int64_t>(rr) * m + q) * 4));
        const float eta = prm.x, zeta = prm.y, lam = prm.z, rc = prm.w;
        if (!uniform_rc) {
          cutoff(bij, rc, gij, gdij);
          cutoff(bik, rc, gik, gdik);
          cutoff(bjk, rc, gjk, gdjk);
        }
        // (1 + lambda cos)^zeta as exp2(zeta log2(base)); base is clamped at 0 (it can only be negative by
        // rounding of cos) -> 0^zeta = 0 like pow
        const float base = fmaxf(cc * lam + 1.f, 0.f);
        const float pre = exp2f(1.f - zeta) * wtt;
        // wACSF shifts the three Gaussians by mu (wacsf_conv.py:277-281); ACSF G4 has mu = 0
        const float mu = mu_i ? __ldg(mu_i + rr * m + q) : 0.f;
        const float sij = bij - mu, sik = bik - mu, sjk = bjk - mu;
        const float pw = (base > 0.f) ? exp2f(zeta * log2f(base)) : (zeta == 0.f ? 1.f : 0.f);
        const float A = pre * pw;
        const float E = expf(-eta * (mu_i ? sij * sij + sik * sik + sjk * sjk : r2));
        const float fff = gij * gik * gjk;
        if (MODE == kFwd) {
          acc[rr * m + q] += A * E * fff;
        } else {
          const float pw1 = (base > 0.f) ? pw / base : (zeta == 1.f ? 1.f : 0.f);   // base^(zeta-1)
          const float dA = pre * zeta * lam * pw1;                      // dA / dcos
          // d rep / d r_ab = A (dE/dr_ab fff + E f'_ab f f)
          const float d_ij = A * E * (-2.f * eta * sij * fff + gdij * gik * gjk);
          const float d_ik = A * E * (-2.f * eta * sik * fff + gij * gdik * gjk);
          const float d_jk = A * E * (-2.f * eta * sjk * fff + gij * gik * gdjk);
          const float d_c = dA * E * fff;
          if (MODE == kVjp) {
            const float g = acc[rr * m + q];
            a_ij += g * d_ij; a_ik += g * d_ik; a_jk += g * d_jk; a_c += g * d_c;
          } else {
            acc[rr * m + q] += d_ij * u_ij + d_ik * u_ik + d_jk * u_jk + d_c * u_c;
          }
        }
      }
      if (MODE == kVjp) {
        a_ij = warp_sum(a_ij); a_ik = warp_sum(a_ik); a_jk = warp_sum(a_jk); a_c = warp_sum(a_c);
        if (lane == t) { s_ij = a_ij; s_ik = a_ik; s_jk = a_jk; s_c = a_c; }
      }
    }
    if (MODE == kVjp && have) {
      V3 gj = {0.f, 0.f, 0.f}, gk = {0.f, 0.f, 0.f};
      if (r >= 0) {
        const V3 g_vij = axpy(s_ij / rij, vij, scale(s_c, dc_ij));
        const V3 g_vik = axpy(s_ik / rik, vik, scale(s_c, dc_ik));
        const V3 g_vjk = scale(s_jk / rjk, vjk);
        gi = axpy(1.f, g_vij, axpy(1.f, g_vik, gi));
        gj = sub(g_vjk, g_vij);
        gk = scale(-1.f, axpy(1.f, g_vik, g_vjk));
      }
      float* o = triple_out + 3 * static_cast<int64_t>(e);            // [2, A, 3]: gradients of atom j, then atom k
      o[0] = gj.x; o[1] = gj.y; o[2] = gj.z;
      o += 3 * n_triples;
      o[0] = gk.x; o[1] = gk.y; o[2] = gk.z;
    }
  }
  if (MODE == kVjp) { gi.x = warp_sum(gi.x); gi.y = warp_sum(gi.y); gi.z = warp_sum(gi.z); }   // fixed tree
  __syncwarp();
  if (MODE == kVjp) {
    if (lane == 0) { out[3 * i] = gi.x; out[3 * i + 1] = gi.y; out[3 * i + 2] = gi.z; }
  } else {
    for (int t = lane; t < RM; t += 32) out[i * RM + t] = acc[t];
  }
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_acsf_g2(const float* xyz, const int32_t* rel_atom, const int32_t* tset_atom, const int32_t* rowptr,
                  const int32_t* perm, const int32_t* send, int send_permuted, const float* table, int R, int m,
                  float eps, int64_t N, const float* weight, int mode, const float* cot, float* out,
                  float* edge_out, kgcnn_stream_t stream) {
  if (N < 0 || R <= 0 || m <= 0 || mode < kFwd || mode > kJvp) return KGCNN_E_INVALID;
  if (R * m > kMaxAcc) return KGCNN_E_UNSUPPORTED;
  if (N == 0) return KGCNN_OK;
  if (!xyz || !rel_atom || !rowptr || !send || !table || !out) return KGCNN_E_INVALID;
  if (mode != kFwd && !cot) return KGCNN_E_INVALID;
  if (mode == kVjp && !edge_out) return KGCNN_E_INVALID;
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(N, kWarpsPerBlock));
  cudaStream_t st = as_stream(stream);
#define KGCNN_G2(M) acsf_g2_kernel<M><<<grid, kWarpsPerBlock * 32, 0, st>>>(xyz, rel_atom, tset_atom, rowptr, perm, \
    send, send_permuted, table, R, m, eps, N, weight, cot, out, edge_out)
  if (mode == kFwd) KGCNN_G2(kFwd); else if (mode == kVjp) KGCNN_G2(kVjp); else KGCNN_G2(kJvp);
#undef KGCNN_G2
  return finish_launch();
}

int kgcnn_acsf_g4(const float* xyz, const int32_t* z_atom, const int32_t* tset_atom, const int32_t* rev_pair,
                  int zmax, const int32_t* rowptr, const int32_t* perm, const int32_t* idx_j, const int32_t* idx_k,
                  const float* table, int R, int m, float eps, float multiplicity, int uniform_rc, int64_t N,
                  const float* mu_table, const float* weight, const float* factor_table,
                  const int32_t* factor_index, int64_t n_triples, int mode, const float* cot, float* out,
                  float* triple_out, kgcnn_stream_t stream) {
  if (N < 0 || R <= 0 || m <= 0 || zmax <= 0 || mode < kFwd || mode > kJvp) return KGCNN_E_INVALID;
  if (R * m > kMaxAcc) return KGCNN_E_UNSUPPORTED;
  if (N == 0) return KGCNN_OK;
  if (!xyz || !z_atom || !rev_pair || !rowptr || !idx_j || !idx_k || !table || !out) return KGCNN_E_INVALID;
  if ((reinterpret_cast<uintptr_t>(table) & 15) != 0) return KGCNN_E_UNSUPPORTED;
  if (mode != kFwd && !cot) return KGCNN_E_INVALID;
  if (mode == kVjp && (!triple_out || n_triples < 0)) return KGCNN_E_INVALID;
  if ((factor_table != nullptr) != (factor_index != nullptr)) return KGCNN_E_INVALID;
  if (factor_table && (!uniform_rc || (reinterpret_cast<uintptr_t>(factor_table) & 7) != 0)) return KGCNN_E_INVALID;
  const float2* fac = reinterpret_cast<const float2*>(factor_table);
  const float inv_mult = multiplicity > 0.f ? 1.f / multiplicity : 1.f;
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(N, kWarpsPerBlock));
  cudaStream_t st = as_stream(stream);
#define KGCNN_G4(M) acsf_g4_kernel<M><<<grid, kWarpsPerBlock * 32, 0, st>>>(xyz, z_atom, tset_atom, rev_pair, zmax, \
    rowptr, perm, idx_j, idx_k, table, R, m, eps, inv_mult, uniform_rc, N, mu_table, weight, fac, factor_index, cot, \
    out, triple_out, n_triples)
  if (mode == kFwd) KGCNN_G4(kFwd); else if (mode == kVjp) KGCNN_G4(kVjp); else KGCNN_G4(kJvp);
#undef KGCNN_G4
  return finish_launch();
}

}  // extern "C"
