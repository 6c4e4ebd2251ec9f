// Probe: which shared-memory address does element (m, k) of an MN-major, non-swizzled TF32 A
// operand of tcgen05.mma come from, as a function of the descriptor's LBO and SBO?
// smem float i holds the value i (exact in TF32 below 2048); B is a K-major selector with
// B[n][k] = (n == k), so D[m][n] = A[m][n] for n < 8 is the float index that was read.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/umma_mn_probe scripts/experiments/umma_mn_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t mkdesc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t lt = 0) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)lt << 61;
  return d;
}

__global__ void probe(float* out, uint32_t lbo, uint32_t sbo, int a_mn, uint32_t lt, uint32_t off, int pattern) {
  __shared__ __align__(1024) float S[8192 + 512];       // A region: S[0..2047]; B selector after it
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  for (int i = t; i < 8192; i += blockDim.x) S[i] = pattern ? (float)(i & 31) : (float)(i >> 2);
  float* Bsel = S + 8192;                                // K-major: [k chunk 2][n 32][4]
  for (int i = t; i < 256; i += blockDim.x) {
    const int c = i & 3, n = (i >> 2) & 31, kc = i >> 7;
    Bsel[i] = (n == kc * 4 + c) ? 1.0f : 0.0f;
  }
  for (int i = 256 + t; i < 512; i += blockDim.x) Bsel[i] = 0.0f;
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(s32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  if (t == 0) {
    uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
    if (a_mn) idesc |= 1u << 15;
    const uint64_t ad = mkdesc(s32(S) + off, lbo, sbo, lt);
    const uint64_t bd = mkdesc(s32(Bsel), 32 * 16, 128);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(ad), "l"(bd),
                 "r"(idesc), "r"(0u) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
  }
  if (warp < 4) {
    for (uint32_t i = 0; i < 4000000u; ++i) {
      uint32_t ok;
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(s32(&bar)), "r"(0u) : "memory");
      if (ok) break;
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(tmem + ((uint32_t)(warp * 32) << 16)) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int k = 0; k < 8; ++k) out[(warp * 32 + lane) * 8 + k] = __uint_as_float(r[k]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem) : "memory");
}

int main() {
  float* d;
  cudaMalloc(&d, 128 * 8 * 4);
  static float h[128 * 8];
  // {LBO, SBO, a_mn, layout type (6 = 32B, 4 = 64B, 2 = 128B, 1 = 128B base 32B), start offset}
  // {LBO, SBO, a_mn, layout type (6 = 32B, 4 = 64B, 2 = 128B, 1 = 128B base 32B), start offset}
  const uint32_t combos[][5] = {{128, 512, 1, 1, 0}, {512, 128, 1, 1, 0}, {128, 1024, 1, 1, 0}, {1024, 128, 1, 1, 0},
                                {1024, 128, 1, 1, 128}, {1024, 128, 1, 1, 256}, {2048, 1024, 1, 1, 0}};
  static float h2[128 * 8];
  for (auto& c : combos) {
    cudaError_t e = cudaSuccess;
    for (int pat = 0; pat < 2; ++pat) {
      cudaMemset(d, 0, sizeof(h));
      probe<<<1, 128>>>(d, c[0], c[1], (int)c[2], c[3], c[4], pat);
      e = cudaDeviceSynchronize();
      cudaMemcpy(pat ? h2 : h, d, sizeof(h), cudaMemcpyDeviceToHost);
    }
    printf("LBO %u SBO %u a_mn %u layout %u start+%u (%s): byte offset read for (m,k) as unit*16 | elem-in-128B:\n", c[0], c[1], c[2], c[3], c[4], cudaGetErrorString(e));
    const int ms[] = {0, 1, 4, 8, 31, 32, 33, 64, 96, 127};
    for (int m : ms) {
      printf("  m=%3d:", m);
      for (int k = 0; k < 8; ++k) printf(" %5.0f|%2.0f", h[m * 8 + k] * 16, h2[m * 8 + k]);
      printf("\n");
    }
    if (e != cudaSuccess) break;
  }
  return 0;
}
