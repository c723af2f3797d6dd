// Fused edge geometry (a9-a12): from coordinates and an index pair straight to the per-edge quantities
// the models consume, in one launch instead of the reference's chain of ~15 elementwise TF ops:
//     v = x_i - x_j                                    NodePosition + LazySubtract      geom.py:14-73
//     d = sqrt(relu(sum v^2) + eps)                    NodeDistanceEuclidean            geom.py:181-187, 320-321
//     u = v * divide_no_nan(1, d)                      EdgeDirectionNormalized          geom.py:369-371
//     rbf_k = exp(-gamma (d - offset - k D / bins)^2)  GaussBasisLayer                  geom.py:567-570
//     rbf_n = env_p(d / c) sin(f_n d / c)              BesselBasisLayer                 geom.py:772-785
//     fc = 0.5 (cos(pi clip(d, -c, c) / c) + 1)        CosCutOffEnvelope                geom.py:834-835
// Three maps, like the ACSF kernels (forces are trained on; coordinates are not):
//     mode 0  values;   mode 1  VJP: cotangents of (d, u, rbf, fc) -> g_v[E,3] (scattered +i / -j by the CSR kernel);
//     mode 2  JVP: tangent h[N,3] of the coordinates -> tangents of (d, u, rbf, fc);
//     mode 3  VJP of the VJP with respect to the coordinates (the coordinate-Hessian term): cotangents c of
//             (d, u, rbf, fc) that entered mode 1 and the cotangent h[N,3] of its result ->
//             g_v[E,3] = d/dv sum_out <c_out, D out(v)[eta]>,  eta = h_i - h_j:
//                 scalar outputs g(d):  A2 (u.eta) u + A1 (eta - u (u.eta)) / d,   A1 = sum c g'(d), A2 = sum c g''(d)
//                 direction u = v / d:  (3 (c.u)(u.eta) u - (c.eta) u - (u.eta) c - (c.u) eta) / d^2
// One thread per edge; the [E,R] rows of consecutive edges are contiguous, so the per-k strided accesses of a warp
// stay inside one 32*R*4-byte window that L1 serves after the first touch.
#include "common.cuh"

namespace kgcnn {
namespace {

constexpr float kPiF = 3.14159265358979323846f;
constexpr int kMaxBasis = 64;

enum { kGeoFwd = 0, kGeoVjp = 1, kGeoJvp = 2, kGeoVjpVjp = 3 };
enum { kBasisNone = 0, kBasisGauss = 1, kBasisBessel = 2 };

struct GeoParams {
  int basis, R;
  float eps;            // added under the square root (add_eps), else 0
  int no_nan;           // direction: divide_no_nan
  // gauss
  float gamma, offset, step;      // step = distance / bins
  // bessel
  float inv_cutoff; int p;        // p = envelope_exponent + 1
  // cosine cutoff envelope
  float cos_cutoff;               // <= 0: not requested
};

// Bessel envelope of DimeNet: env(x) = 1/x + a x^(p-1) + b x^p + c x^(p+1) for x < 1, else 0; and d env / dx
__device__ __forceinline__ void bessel_env(float x, int p, float& e, float& de) {
  const float a = -0.5f * (p + 1) * (p + 2), b = static_cast<float>(p) * (p + 2), c = -0.5f * p * (p + 1);
  if (!(x < 1.f)) { e = 0.f; de = 0.f; return; }
  float xp1 = 1.f;                                   // x^(p-1)
  for (int q = 0; q < p - 1; ++q) xp1 *= x;
  const float xp = xp1 * x, xpp = xp * x;
  const float xp2 = (p >= 2) ? xp1 / x : 0.f;        // x^(p-2) (only multiplied when p >= 2)
  e = 1.f / x + a * xp1 + b * xp + c * xpp;
  de = -1.f / (x * x) + a * (p - 1) * xp2 + b * p * xp1 + c * (p + 1) * xp;
}

// second derivative of the envelope (x < 1, else 0)
__device__ __forceinline__ float bessel_env2(float x, int p) {
  if (!(x < 1.f)) return 0.f;
  const float a = -0.5f * (p + 1) * (p + 2), b = static_cast<float>(p) * (p + 2), c = -0.5f * p * (p + 1);
  float xp3 = 1.f;                                   // x^(p-3), built from x^(p-1) / x^2 (p >= 1)
  float xp1 = 1.f;
  for (int q = 0; q < p - 1; ++q) xp1 *= x;
  xp3 = xp1 / (x * x);
  const float xp2 = xp1 / x;
  return 2.f / (x * x * x) + a * (p - 1) * (p - 2) * xp3 + b * p * (p - 1) * xp2 + c * (p + 1) * p * xp1;
}

template <int MODE>
__global__ void __launch_bounds__(256)
edge_geometry_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ idx_i,
                     const int32_t* __restrict__ idx_j, int64_t E, GeoParams prm, const float* __restrict__ freq,
                     // mode 1: cotangents in (nullable each); mode 2: tangent of the coordinates in `h`
                     const float* __restrict__ c_d, const float* __restrict__ c_u, const float* __restrict__ c_rbf,
                     const float* __restrict__ c_fc, const float* __restrict__ h,
                     // mode 0 / 2 outputs (nullable each); mode 1 output g_v
                     float* __restrict__ o_d, float* __restrict__ o_u, float* __restrict__ o_rbf,
                     float* __restrict__ o_fc, float* __restrict__ o_gv) {
  const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (e >= E) return;
  const int i = __ldg(idx_i + e), j = __ldg(idx_j + e);
  const float vx = __ldg(xyz + 3 * i) - __ldg(xyz + 3 * j);
  const float vy = __ldg(xyz + 3 * i + 1) - __ldg(xyz + 3 * j + 1);
  const float vz = __ldg(xyz + 3 * i + 2) - __ldg(xyz + 3 * j + 2);
  const float s = fmaxf(vx * vx + vy * vy + vz * vz, 0.f) + prm.eps;
  const float d = sqrtf(s);
  const float inv = (d == 0.f) ? (prm.no_nan ? 0.f : __int_as_float(0x7f800000)) : 1.f / d;
  const float ux = vx * inv, uy = vy * inv, uz = vz * inv;
  // d d / d v = v / d (0 at d = 0: relu'(0) = 0 in the reference chain)
  const float gdx = (d > 0.f) ? vx / d : 0.f, gdy = (d > 0.f) ? vy / d : 0.f, gdz = (d > 0.f) ? vz / d : 0.f;

  float t_d = 0.f, hx = 0.f, hy = 0.f, hz = 0.f;
  if (MODE == kGeoJvp || MODE == kGeoVjpVjp) {
    hx = __ldg(h + 3 * i) - __ldg(h + 3 * j);
    hy = __ldg(h + 3 * i + 1) - __ldg(h + 3 * j + 1);
    hz = __ldg(h + 3 * i + 2) - __ldg(h + 3 * j + 2);
    t_d = gdx * hx + gdy * hy + gdz * hz;
  }
  float a_d = 0.f;                                    // VJP: total cotangent arriving at d (sum c g'(d))
  float a_dd = 0.f;                                   // mode 3: sum c g''(d)
  constexpr bool kCot = (MODE == kGeoVjp || MODE == kGeoVjpVjp);
  if (kCot && c_d) a_d += __ldg(c_d + e);

  // ---- cosine cutoff envelope ----
  if (prm.cos_cutoff > 0.f && (kCot ? c_fc != nullptr : o_fc != nullptr)) {
    const float rc = prm.cos_cutoff;
    const float dc = fminf(fmaxf(d, -rc), rc);
    float sn, cs;
    sincosf(dc * kPiF / rc, &sn, &cs);
    const bool inside = (d > -rc && d < rc);
    const float dfc = inside ? -0.5f * kPiF / rc * sn : 0.f;
    if (MODE == kGeoFwd) o_fc[e] = 0.5f * (cs + 1.f);
    else if (MODE == kGeoJvp) o_fc[e] = dfc * t_d;
    else {
      const float c = __ldg(c_fc + e);
      a_d += c * dfc;
      if (MODE == kGeoVjpVjp && inside) a_dd += c * (-0.5f * (kPiF / rc) * (kPiF / rc) * cs);
    }
  }
  // ---- radial basis ----
  if (prm.basis != kBasisNone && (kCot ? c_rbf != nullptr : o_rbf != nullptr)) {
    const int R = prm.R;
    if (prm.basis == kBasisGauss) {
      const float t0 = d - prm.offset;
      for (int k = 0; k < R; ++k) {
        const float t = t0 - prm.step * k;
        const float g = expf(-prm.gamma * t * t);
        if (MODE == kGeoFwd) o_rbf[e * R + k] = g;
        else {
          const float dg = -2.f * prm.gamma * t * g;
          if (MODE == kGeoJvp) o_rbf[e * R + k] = dg * t_d;
          else {
            const float c = __ldg(c_rbf + e * R + k);
            a_d += c * dg;
            if (MODE == kGeoVjpVjp) a_dd += c * (4.f * prm.gamma * prm.gamma * t * t - 2.f * prm.gamma) * g;
          }
        }
      }
    } else {
      const float x = d * prm.inv_cutoff;
      float env, denv;
      bessel_env(x, prm.p, env, denv);
      const float ddenv = (MODE == kGeoVjpVjp) ? bessel_env2(x, prm.p) : 0.f;
      for (int k = 0; k < R; ++k) {
        const float f = __ldg(freq + k);
        float sn, cs;
        sincosf(f * x, &sn, &cs);
        if (MODE == kGeoFwd) o_rbf[e * R + k] = env * sn;
        else {
          const float dr = (denv * sn + env * f * cs) * prm.inv_cutoff;        // d rbf / d d
          if (MODE == kGeoJvp) o_rbf[e * R + k] = dr * t_d;
          else {
            const float c = __ldg(c_rbf + e * R + k);
            a_d += c * dr;
            if (MODE == kGeoVjpVjp)
              a_dd += c * (ddenv * sn + 2.f * denv * f * cs - env * f * f * sn) * prm.inv_cutoff * prm.inv_cutoff;
          }
        }
      }
    }
  }
  if (MODE == kGeoFwd) {
    if (o_d) o_d[e] = d;
    if (o_u) { o_u[3 * e] = ux; o_u[3 * e + 1] = uy; o_u[3 * e + 2] = uz; }
  } else if (MODE == kGeoJvp) {
    if (o_d) o_d[e] = t_d;
    if (o_u) {                                         // t_u = (h - (u.h) u) / d
      const float uh = ux * hx + uy * hy + uz * hz;
      o_u[3 * e] = (hx - uh * ux) * inv; o_u[3 * e + 1] = (hy - uh * uy) * inv; o_u[3 * e + 2] = (hz - uh * uz) * inv;
    }
  } else if (MODE == kGeoVjpVjp) {
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (d > 0.f) {
      const float wx = vx / d, wy = vy / d, wz = vz / d;           // v / d (= u unless d == 0)
      const float we = wx * hx + wy * hy + wz * hz;                // u . eta
      const float k1 = a_dd * we - a_d * we / d, k2 = a_d / d;     // A2 (u.eta) u + A1 (eta - u (u.eta)) / d
      gx = k1 * wx + k2 * hx; gy = k1 * wy + k2 * hy; gz = k1 * wz + k2 * hz;
      if (c_u) {
        const float cx = __ldg(c_u + 3 * e), cy = __ldg(c_u + 3 * e + 1), cz = __ldg(c_u + 3 * e + 2);
        const float al = cx * wx + cy * wy + cz * wz, ga = cx * hx + cy * hy + cz * hz, id2 = 1.f / (d * d);
        const float ku = (3.f * al * we - ga) * id2;
        gx += ku * wx - (we * cx + al * hx) * id2;
        gy += ku * wy - (we * cy + al * hy) * id2;
        gz += ku * wz - (we * cz + al * hz) * id2;
      }
    }
    o_gv[3 * e] = gx; o_gv[3 * e + 1] = gy; o_gv[3 * e + 2] = gz;
  } else {
    float gx = a_d * gdx, gy = a_d * gdy, gz = a_d * gdz;
    if (c_u) {                                         // (c_u - (c_u.u) u) / d
      const float cx = __ldg(c_u + 3 * e), cy = __ldg(c_u + 3 * e + 1), cz = __ldg(c_u + 3 * e + 2);
      const float cu = cx * ux + cy * uy + cz * uz;
      gx += (cx - cu * ux) * inv; gy += (cy - cu * uy) * inv; gz += (cz - cu * uz) * inv;
    }
    o_gv[3 * e] = gx; o_gv[3 * e + 1] = gy; o_gv[3 * e + 2] = gz;
  }
}


// ---------------------------------------------------------------------------------------------------------
// Gradient of a scalar with respect to the TRAINABLE Bessel frequencies f_n (kgcnn/layers/geom.py:766-770):
//   tangent_xyz == nullptr :  g_f[n] = sum_e c[e,n] d rbf_n / d f_n              (through the energy)
//   tangent_xyz != nullptr :  g_f[n] = sum_e c[e,n] t_d[e] d2 rbf_n / (d d  d f_n)   (through the forces: c is the
//                              cotangent that entered the coordinate VJP, t_d = (d d / d x) . tangent)
// with  d rbf / d f = env x cos(f x),   d2 rbf / (d d d f) = (env' x cos + env cos - env f x sin) / c,   x = d / c.
// One lane per frequency, one edge per warp and iteration; block partials, then a fixed-order sum: deterministic.
// ---------------------------------------------------------------------------------------------------------
constexpr int kFreqBlocks = 2 * kNumSM;
constexpr int kFreqWarps = 8;

__global__ void __launch_bounds__(kFreqWarps * 32)
bessel_freq_partial_kernel(const float* __restrict__ xyz, const int32_t* __restrict__ idx_i,
                           const int32_t* __restrict__ idx_j, int64_t E, GeoParams prm, const float* __restrict__ freq,
                           const float* __restrict__ cot, const float* __restrict__ h, float* __restrict__ partial) {
  __shared__ float red[kFreqWarps][kMaxBasis];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = prm.R;
  const float f0 = lane < R ? __ldg(freq + lane) : 0.f, f1 = lane + 32 < R ? __ldg(freq + lane + 32) : 0.f;
  float acc0 = 0.f, acc1 = 0.f;
  const int64_t per = (E + gridDim.x - 1) / gridDim.x;
  const int64_t e_beg = blockIdx.x * per, e_end = min(E, e_beg + per);
  for (int64_t e = e_beg + warp; e < e_end; e += kFreqWarps) {
    const int i = __ldg(idx_i + e), j = __ldg(idx_j + e);
    const float vx = __ldg(xyz + 3 * i) - __ldg(xyz + 3 * j);
    const float vy = __ldg(xyz + 3 * i + 1) - __ldg(xyz + 3 * j + 1);
    const float vz = __ldg(xyz + 3 * i + 2) - __ldg(xyz + 3 * j + 2);
    const float d = sqrtf(fmaxf(vx * vx + vy * vy + vz * vz, 0.f) + prm.eps);
    const float x = d * prm.inv_cutoff;
    float env, denv;
    bessel_env(x, prm.p, env, denv);
    float t_d = 1.f;
    if (h) {
      const float hx = __ldg(h + 3 * i) - __ldg(h + 3 * j), hy = __ldg(h + 3 * i + 1) - __ldg(h + 3 * j + 1),
                  hz = __ldg(h + 3 * i + 2) - __ldg(h + 3 * j + 2);
      t_d = (d > 0.f) ? (vx * hx + vy * hy + vz * hz) / d : 0.f;
    }
    auto term = [&](float f) {
      float sn, cs;
      sincosf(f * x, &sn, &cs);
      return h ? t_d * (denv * x * cs + env * cs - env * f * x * sn) * prm.inv_cutoff : env * x * cs;
    };
    if (lane < R) acc0 += __ldg(cot + e * R + lane) * term(f0);
    if (lane + 32 < R) acc1 += __ldg(cot + e * R + lane + 32) * term(f1);
  }
  red[warp][lane] = acc0;
  red[warp][lane + 32] = acc1;
  __syncthreads();
  if (threadIdx.x < kMaxBasis) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kFreqWarps; ++w) s += red[w][threadIdx.x];
    if (threadIdx.x < R) partial[static_cast<int64_t>(blockIdx.x) * R + threadIdx.x] = s;
  }
}

__global__ void bessel_freq_reduce_kernel(const float* __restrict__ partial, int blocks, int R, float* __restrict__ out) {
  const int n = threadIdx.x;
  if (n >= R) return;
  float s0 = 0.f, s1 = 0.f;
  int b = 0;
  for (; b + 1 < blocks; b += 2) { s0 += partial[b * R + n]; s1 += partial[(b + 1) * R + n]; }
  if (b < blocks) s0 += partial[b * R + n];
  out[n] = s0 + s1;
}

}  // namespace
}  // namespace kgcnn

using namespace kgcnn;

extern "C" {

int kgcnn_edge_geometry(const float* xyz, const int32_t* idx_i, const int32_t* idx_j, int64_t E, int mode,
                        int basis, int R, const float* basis_params, const float* frequencies, float eps,
                        int no_nan, float cos_cutoff, const float* cot_d, const float* cot_u, const float* cot_rbf,
                        const float* cot_fc, const float* tangent_xyz, float* out_d, float* out_u, float* out_rbf,
                        float* out_fc, float* out_gv, kgcnn_stream_t stream) {
  if (E < 0 || mode < kGeoFwd || mode > kGeoVjpVjp || basis < kBasisNone || basis > kBasisBessel) return KGCNN_E_INVALID;
  if (basis != kBasisNone && (R <= 0 || R > kMaxBasis || !basis_params)) return KGCNN_E_INVALID;
  if (E == 0) return KGCNN_OK;
  if (!xyz || !idx_i || !idx_j) return KGCNN_E_INVALID;
  if ((mode == kGeoVjp || mode == kGeoVjpVjp) && !out_gv) return KGCNN_E_INVALID;
  if ((mode == kGeoJvp || mode == kGeoVjpVjp) && !tangent_xyz) return KGCNN_E_INVALID;
  if (basis == kBasisBessel && !frequencies) return KGCNN_E_INVALID;
  GeoParams prm;
  prm.basis = basis; prm.R = R; prm.eps = eps; prm.no_nan = no_nan; prm.cos_cutoff = cos_cutoff;
  prm.gamma = prm.offset = prm.step = prm.inv_cutoff = 0.f; prm.p = 1;
  // basis_params is a HOST array of 3 floats: gauss (gamma, offset, distance / bins); bessel (1 / cutoff, p, -)
  if (basis == kBasisGauss) { prm.gamma = basis_params[0]; prm.offset = basis_params[1]; prm.step = basis_params[2]; }
  if (basis == kBasisBessel) {
    prm.inv_cutoff = basis_params[0]; prm.p = static_cast<int>(basis_params[1]);
    if (prm.p < 1) return KGCNN_E_INVALID;
  }
  const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(E, 256));
  cudaStream_t st = as_stream(stream);
#define KGCNN_GEO(M) edge_geometry_kernel<M><<<grid, 256, 0, st>>>(xyz, idx_i, idx_j, E, prm, frequencies, cot_d, \
    cot_u, cot_rbf, cot_fc, tangent_xyz, out_d, out_u, out_rbf, out_fc, out_gv)
  if (mode == kGeoFwd) KGCNN_GEO(kGeoFwd);
  else if (mode == kGeoVjp) KGCNN_GEO(kGeoVjp);
  else if (mode == kGeoJvp) KGCNN_GEO(kGeoJvp);
  else KGCNN_GEO(kGeoVjpVjp);
#undef KGCNN_GEO
  return finish_launch();
}

size_t kgcnn_bessel_frequency_grad_workspace_bytes(int R) {
  return R > 0 ? static_cast<size_t>(kFreqBlocks) * R * sizeof(float) : 0;
}

int kgcnn_bessel_frequency_grad(const float* xyz, const int32_t* idx_i, const int32_t* idx_j, int64_t E, int R,
                                const float* basis_params, const float* frequencies, float eps, const float* cot_rbf,
                                const float* tangent_xyz, float* out, void* workspace, size_t workspace_bytes,
                                kgcnn_stream_t stream) {
  static_assert(kMaxBasis == 64, "two frequencies per lane");
  if (E < 0 || R <= 0 || R > kMaxBasis || !basis_params || !frequencies || !out) return KGCNN_E_INVALID;
  cudaStream_t st = as_stream(stream);
  if (E == 0) return cudaMemsetAsync(out, 0, sizeof(float) * R, st) == cudaSuccess ? KGCNN_OK : KGCNN_E_CUDA;
  if (!xyz || !idx_i || !idx_j || !cot_rbf) return KGCNN_E_INVALID;
  if (!workspace || workspace_bytes < kgcnn_bessel_frequency_grad_workspace_bytes(R)) return KGCNN_E_WORKSPACE;
  GeoParams prm;
  prm.basis = kBasisBessel; prm.R = R; prm.eps = eps; prm.no_nan = 1; prm.cos_cutoff = 0.f;
  prm.gamma = prm.offset = prm.step = 0.f;
  prm.inv_cutoff = basis_params[0]; prm.p = static_cast<int>(basis_params[1]);
  if (prm.p < 1) return KGCNN_E_INVALID;
  const int blocks = static_cast<int>(std::min<int64_t>(kFreqBlocks, ceil_div<int64_t>(E, kFreqWarps)));
  float* partial = static_cast<float*>(workspace);
  bessel_freq_partial_kernel<<<blocks, kFreqWarps * 32, 0, st>>>(xyz, idx_i, idx_j, E, prm, frequencies, cot_rbf,
                                                                 tangent_xyz, partial);
  bessel_freq_reduce_kernel<<<1, kMaxBasis, 0, st>>>(partial, blocks, R, out);
  return finish_launch(2);
}

}  // extern "C"
