// conv2 forward (10->20 channels, 5x5 valid) as a tcgen05 implicit GEMM with the whole layer
// tail fused into the TMEM epilogue (SURVEY K3+K4+K5, row 8.3; reference: src/model.py:17
// `relu(max_pool2d(conv2_drop(conv2(x)), 2))`).
//
//   GEMM view: M = 64 output positions x 2 images per CTA = 128, N = 20 channels (padded to
//   32), K = 10*5*5 = 250 (padded to 256), TF32 inputs, fp32 accumulation in TMEM.
//
//   1. one elected thread TMA-bulk-copies the two images' [10,12,12] fp32 activations
//      (cp.async.bulk, 11,520 contiguous bytes) into shared memory, completion on an mbarrier
//   2. all 256 threads build the im2col A tile [128 x 256] and the weight B tile [32 x 256]
//      directly in the 128B-swizzled K-major layout the UMMA descriptors expect
//      (8 K-blocks of 32 floats; 16-byte chunk index XOR row%8), then fence.proxy.async
//   3. one thread issues 32 x tcgen05.mma.cta_group::1.kind::tf32 (M=128, N=32, K=8) and commits
//   4. warps 0-3 read the accumulator with tcgen05.ld (lane = GEMM row). Rows are ordered so
//      the four members of each 2x2 pooling window sit in four consecutive lanes: bias,
//      max + first-max arg-max by quad shuffles, Dropout2d channel mask, ReLU, store.
//
// This is the tensor-core statement of the layer. For this 21,840-parameter network the
// fused SIMT step (net_kernels.cu) is faster end to end because it never leaves shared
// memory between layers; both paths are kept and measured (profiles/).
#include "common.cuh"
#include "net_device.cuh"

using namespace net;

namespace {

constexpr int TC_THREADS = 256;
constexpr int KPAD = 256, NPAD = 32, MT = 128;
constexpr int A_BYTES = MT * KPAD * 4;      // 131072
constexpr int B_BYTES = NPAD * KPAD * 4;    // 32768
constexpr int IMG_BYTES = A1_SZ * 4;        // 5760
constexpr int TC_SMEM = A_BYTES + B_BYTES + 2 * IMG_BYTES + 1024 + 64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// Byte offset of element (row r, k) inside a K-major, 128B-swizzled operand made of
// K-blocks of 32 floats; each K-block is `rows` x 128 B, 8-row groups of 1024 B.
__device__ __forceinline__ uint32_t swz_off(int r, int k, int rows) {
  const int kb = k >> 5, c = k & 31;
  return (uint32_t)(kb * rows * 128 + (r >> 3) * 1024 + (r & 7) * 128 +
                    (((c >> 2) ^ (r & 7)) << 4) + ((c & 3) << 2));
}

__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                 // LBO (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;       // SBO: 8-row groups are 1024 B apart
  d |= (uint64_t)1 << 46;                 // sm_100 descriptor version
  d |= (uint64_t)2 << 61;                 // SWIZZLE_128B
  return d;
}

// Row order inside one image's 64 GEMM rows: row = window*4 + j, window = py*4+px,
// j = dy*2+dx, so a pooling window occupies 4 consecutive TMEM lanes.
__device__ __forceinline__ void row_to_yx(int row, int& y, int& x) {
  const int win = row >> 2, j = row & 3;
  y = 2 * (win >> 2) + (j >> 1);
  x = 2 * (win & 3) + (j & 1);
}

__global__ void __launch_bounds__(TC_THREADS, 1)
conv2_fwd_tc_kernel(const float* __restrict__ a1, const float* __restrict__ w,
                    const float* __restrict__ b, float* __restrict__ mask, int mask_mode, float p,
                    unsigned long long seed, unsigned long long step, float* __restrict__ a2,
                    uint8_t* __restrict__ pos2, int B) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* As = smem;
  uint8_t* Bs = As + A_BYTES;
  float* imgs = reinterpret_cast<float*>(Bs + B_BYTES);          // 2 x [10,12,12]
  uint64_t* bar_in = reinterpret_cast<uint64_t*>(Bs + B_BYTES + 2 * IMG_BYTES);
  uint64_t* bar_mma = bar_in + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
  __shared__ float bias_s[NPAD], mask_s[2][NPAD];

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int img0 = blockIdx.x * 2;
  const int nimg = min(2, B - img0);

  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar_in)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar_mma)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(
                     smem_u32(tmem_slot))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (t < NPAD) bias_s[t] = (t < C2) ? b[t] : 0.0f;
  if (t < 2 * NPAD) {
    const int im = t / NPAD, co = t % NPAD;
    float m = 1.0f;
    if (co < C2 && im < nimg) {
      if (mask_mode == 1) {
        m = dropout_mask_value(seed, step, (uint32_t)(img0 + im), 0x20000u + co, p);
        mask[(size_t)(img0 + im) * C2 + co] = m;
      } else if (mask_mode == 2) {
        m = mask[(size_t)(img0 + im) * C2 + co];
      }
    }
    mask_s[im][co] = m;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  // 1. TMA bulk copy of the input activations
  if (t == 0) {
    const uint32_t bytes = (uint32_t)(nimg * IMG_BYTES);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar_in)),
                 "r"(bytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(imgs)),
        "l"(a1 + (size_t)img0 * A1_SZ), "r"(bytes), "r"(smem_u32(bar_in))
        : "memory");
  }
  // weights -> swizzled B tile while the copy is in flight (rows >= 20 and k >= 250 are zero)
  for (int i = t; i < NPAD * KPAD; i += TC_THREADS) {
    const int n = i >> 8, k = i & 255;
    const float v = (n < C2 && k < 250) ? __ldg(w + n * 250 + k) : 0.0f;
    *reinterpret_cast<float*>(Bs + swz_off(n, k, NPAD)) = v;
  }
  {  // wait for the activations
    asm volatile(
        "{\n\t.reg .pred p;\n\tW_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
        "@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(bar_in))
        : "memory");
  }
  // 2. im2col: A[row][k] = a1[img][ci][y+ky][x+kx]; thread <-> (row, 16-wide k slice)
  for (int i = t; i < MT * (KPAD / 4); i += TC_THREADS) {
    const int r = i & (MT - 1), k4 = (i >> 7) << 2;       // 4 consecutive k per iteration
    const int im = r >> 6;
    int y, x;
    row_to_yx(r & 63, y, x);
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k4 + q;
      float val = 0.0f;
      if (k < 250 && im < nimg) {
        const int ci = k / 25, kk = k - ci * 25;
        const int ky = kk / 5, kx = kk - ky * 5;
        val = imgs[im * A1_SZ + ci * 144 + (y + ky) * 12 + (x + kx)];
      }
      v[q] = val;
    }
    *reinterpret_cast<float4*>(As + swz_off(r, k4, MT)) = make_float4(v[0], v[1], v[2], v[3]);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic writes -> tensor core
  __syncthreads();

  // 3. MMA: 32 steps of K = 8 (TF32)
  if (warp == 0 && lane == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) |      // F32 acc, TF32 x TF32
                               ((uint32_t)(NPAD >> 3) << 17) | ((uint32_t)(MT >> 4) << 24);
    const uint32_t a_addr = smem_u32(As), b_addr = smem_u32(Bs);
#pragma unroll 1
    for (int ks = 0; ks < KPAD / 8; ++ks) {
      const int kb = ks >> 2, kin = ks & 3;                    // K-block, 32-byte step inside it
      const uint64_t ad = make_desc_k_sw128(a_addr + kb * (MT * 128) + kin * 32);
      const uint64_t bd = make_desc_k_sw128(b_addr + kb * (NPAD * 128) + kin * 32);
      const uint32_t acc = ks != 0 ? 1u : 0u;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_base),
          "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     smem_u32(bar_mma))
                 : "memory");
  }

  // 4. epilogue: warps 0-3, lane group = warp, TMEM lane = GEMM row
  if (warp < 4) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tW_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t"
        "@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(bar_mma))
        : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t r[32];
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const int row = warp * 32 + lane;          // GEMM row in the tile
    const int im = row >> 6, rin = row & 63;
    const int win = rin >> 2, j = rin & 3;
    const int py = win >> 2, px = win & 3;
#pragma unroll
    for (int co = 0; co < C2; ++co) {
      float v = __uint_as_float(r[co]) + bias_s[co];
      int jj = j;
      // max over the quad with first-max tie breaking (lower j wins on equality)
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oj = __shfl_xor_sync(0xffffffffu, jj, o);
        if (ov > v || (ov == v && oj < jj)) { v = ov; jj = oj; }
      }
      if (j == 0 && im < nimg) {
        const size_t o = (size_t)(img0 + im) * A2_SZ + co * 16 + win;
        a2[o] = fmaxf(v * mask_s[im][co], 0.0f);
        pos2[o] = (uint8_t)((2 * py + (jj >> 1)) * 12 + 2 * px + (jj & 1));
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem_base) : "memory");
  }
}

}  // namespace

extern "C" {

// Same contract as b200_conv2_fwd (include/b200dp.h), computed on the tensor cores in TF32.
int b200_conv2_fwd_tc(const float* a1, const float* w, const float* b, float* mask, int mask_mode,
                      float p, unsigned long long seed, unsigned long long step, float* a2,
                      void* pos2, int B, void* stream) {
  if (B <= 0) return 0;
  B200_REQUIRE(mask_mode == 0 || mask != nullptr, "b200_conv2_fwd_tc: mask buffer required");
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv2_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      TC_SMEM));
    attr_done[dev & 63] = true;
  }
  conv2_fwd_tc_kernel<<<(B + 1) / 2, TC_THREADS, TC_SMEM, (cudaStream_t)stream>>>(
      a1, w, b, mask, mask_mode, p, seed, step, a2, (uint8_t*)pos2, B);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
