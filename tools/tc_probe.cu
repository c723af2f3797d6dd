// Probe: one tcgen05.mma kind::tf32 (M=128, N=32, K=8) on hand-placed operands, for several descriptor variants.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tc_probe tools/tc_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo16, uint32_t sbo16, int version) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(lbo16 & 0x3FFF) << 16) | ((uint64_t)(sbo16 & 0x3FFF) << 32) |
         ((uint64_t)version << 46);
}

// variant: a_mn, b_mn (majors), swap (swap LBO/SBO roles), version bit
__global__ void probe(int a_mn, int b_mn, int swap, int version, int pad, int mode, float* out /*[128][32]*/, int* status) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t mbar;
  __shared__ uint32_t tbase;
  const int N = 32, K = 8;
  float4* a = (float4*)smem;                 // plane A: slots
  float4* b = a + 4096;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // A[m][k] = (m % 7) + 1 if k == (m % 8) else 0  -> picks B row; B[n][k] = n + 10*k + 1
  // K-major operand with NR rows: slot(c, r) = c*NRP + r holds 4 consecutive k (chunk c) of row r
  // MN-major operand: slot(j, k) = j*NRP + k holds 4 consecutive mn (chunk j) at reduction index k
  const int A_NR = a_mn ? K : 128, B_NR = b_mn ? K : N;
  const int A_NRP = A_NR + pad, B_NRP = B_NR + pad;
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) { a[i] = make_float4(0,0,0,0); b[i] = make_float4(0,0,0,0); }
  __syncthreads();
  auto Aval = [](int m, int k) { return (k == (m % 8)) ? float((m % 7) + 1) : 0.f; };
  auto Bval = [](int n, int k) { return float(n + 10 * k + 1); };
  if (threadIdx.x == 0) {
    for (int m = 0; m < 128; ++m) for (int k = 0; k < K; ++k) {
      float v = Aval(m, k);
      if (!a_mn) ((float*)&a[(k / 4) * A_NRP + m])[k % 4] = v;
      else ((float*)&a[(m / 4) * A_NRP + k])[m % 4] = v;
    }
    for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) {
      float v = Bval(n, k);
      if (!b_mn) ((float*)&b[(k / 4) * B_NRP + n])[k % 4] = v;
      else ((float*)&b[(n / 4) * B_NRP + k])[n % 4] = v;
    }
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tbase)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t td = tbase;
  uint32_t a_lbo = a_mn ? 8 : A_NRP, a_sbo = a_mn ? A_NRP : 8;
  uint32_t b_lbo = b_mn ? 8 : B_NRP, b_sbo = b_mn ? B_NRP : 8;
  // swap only the MN-major operands' roles
  if (swap) { if (a_mn) { uint32_t t = a_lbo; a_lbo = a_sbo; a_sbo = t; } if (b_mn) { uint32_t t = b_lbo; b_lbo = b_sbo; b_sbo = t; } }
  if (mode == 1) { if (a_mn) { a_lbo = 1; } if (b_mn) { b_lbo = 1; } }          // LBO = 1 unit
  if (mode == 2) { if (a_mn) { a_lbo = A_NRP; a_sbo = 1; } if (b_mn) { b_lbo = B_NRP; b_sbo = 1; } }
  uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
                   ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (threadIdx.x == 0) {
    uint64_t da = make_desc(smem_u32(a), a_lbo, a_sbo, version), db = make_desc(smem_u32(b), b_lbo, b_sbo, version);
    uint32_t acc = 0;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(td), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
  }
  __syncwarp();
  uint32_t done = 0; int spins = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(smem_u32(&mbar)), "r"(0u) : "memory");
    if (!done && ++spins > (1 << 20)) { if (threadIdx.x == 0) *status = -1; break; }
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[32];
  uint32_t taddr = td + ((uint32_t)(32 * warp) << 16);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 32; ++i) out[(32 * warp + lane) * 32 + i] = __uint_as_float(v[i]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(td), "r"(32u) : "memory");
}

int main() {
  float* out; int* status;
  cudaMalloc(&out, 128 * 32 * 4); cudaMalloc(&status, 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 4096 * 16);
  float h[128 * 32];
  int combos[3][2] = {{0, 1}, {1, 0}, {1, 1}};
  for (int ci = 0; ci < 3; ++ci) for (int pad = 0; pad < 2; ++pad) for (int swap = 0; swap < 2; ++swap) for (int mode = 0; mode < 3; ++mode) {
      int a_mn = combos[ci][0], b_mn = combos[ci][1];
      cudaMemset(out, 0xff, sizeof(h)); cudaMemset(status, 0, 4);
      probe<<<1, 128, 2 * 4096 * 16>>>(a_mn, b_mn, swap, 1, pad, mode, out, status);
      cudaError_t e = cudaDeviceSynchronize();
      int st; cudaMemcpy(&st, status, 4, cudaMemcpyDeviceToHost); cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int m = 0; m < 128; ++m) for (int n = 0; n < 32; ++n) {
        float want = float((m % 7) + 1) * float(n + 10 * (m % 8) + 1);
        if (h[m * 32 + n] != want) ++bad;
      }
      printf("a_mn=%d b_mn=%d pad=%d swap=%d mode=%d : err=%s st=%d mismatches=%d  D[0][0..2]=%g %g %g D[1][0..2]=%g %g %g D[9][5]=%g (want D[0]=1 2 3, D[1]=22 24 26, D[9][5]=48)\n",
             a_mn, b_mn, pad, swap, mode, cudaGetErrorString(e), st, bad, h[0], h[1], h[2], h[32], h[33], h[34], h[9 * 32 + 5]);
      if (e != cudaSuccess) return 1;
    }
  return 0;
}
