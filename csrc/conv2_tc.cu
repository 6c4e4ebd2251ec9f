// conv2 of the MNIST net (10 -> 20 channels, 5x5 valid; src/model.py:17) on the 5th-generation
// tensor cores as an IMPLICIT GEMM WITHOUT im2col: forward, data gradient and weight gradient.
//
// The activations live in a channels-last-by-4 layout ("c4"): [image][chunk 0..3][Y][X][4 floats],
// i.e. the 10 input channels padded to 16 and cut into 4 chunks of 4, so that one pixel's chunk
// is one 16-byte unit. A 5-D TMA tensor map (cp.async.bulk.tensor.5d, SASS UTMALDG) brings TWO
// images into shared memory as [chunk][Y][image][X][4]. In that layout the A operand of filter
// tap (ky, kx) is a plain K-major, non-swizzled UMMA operand that needs no data movement at all:
//     GEMM row r = (2*y + image) * 8 + x  ->  start + x*16 + (r/8)*SBO,   SBO = 12 pixels * 16 B
//     K chunk c                            ->  + c*LBO,                    LBO = one chunk plane
// and only the descriptor's start address moves with the tap: + ((ky*2)*12 + kx) * 16 bytes.
// Forward is 25 taps x 2 K-steps = 50 tcgen05.mma.kind::tf32 (M=128, N=32, K=8) into one TMEM
// accumulator, followed by a TMEM epilogue that applies bias, Dropout2d, 2x2 max-pool with
// arg-max and ReLU (a pooling window's four positions sit in lanes l, l^1, l^16, l^17).
// TF32 operands, fp32 accumulation: the precision torch's cuDNN convolutions use by default.
#include <cuda.h>

#include <mutex>

#include "common.cuh"
#include "net_device.cuh"

using namespace net;

namespace {

constexpr int CT_THREADS = 256;
constexpr int CHUNKS = 4;                              // 16 padded input channels / 4
constexpr int A_CHUNK_BYTES = 12 * 2 * 12 * 16;        // one chunk plane of two images: 4608
constexpr int A_BYTES = CHUNKS * A_CHUNK_BYTES;        // 18432
constexpr int NPAD = 32;                               // 20 output channels padded
constexpr int B_TAP_BYTES = CHUNKS * NPAD * 16;        // 2048
constexpr int B_BYTES = 25 * B_TAP_BYTES;              // 51200
constexpr int STAGES = 4;                              // A-tile ring depth
constexpr int FWD_SMEM = STAGES * A_BYTES + B_BYTES + 256 + 1024;

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major operand without swizzle: 8-row x 16-byte core matrices; LBO = byte distance between
// the two 16-byte K chunks of one MMA, SBO = byte distance between 8-row groups.
__device__ __forceinline__ uint64_t desc_k_noswz(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;                 // sm_100 descriptor version
  return d;                               // layout type 0: no swizzle
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem, uint64_t ad, uint64_t bd, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
      "l"(ad), "l"(bd), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar))
               : "memory");
}
// bounded wait (a wrong descriptor must trap, not hang the GPU)
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t i = 0; i < 4000000u; ++i) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(s32(bar)), "r"(parity) : "memory");
    if (ok) return;
  }
  asm volatile("trap;");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
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
}

constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- layout conversion: [B][10][12][12] fp32 -> c4 [B][4][12][12][4] (channels 10..15 zero) ----
__global__ void a1_to_c4_kernel(const float* __restrict__ a1, float* __restrict__ c4, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one (image, chunk, pixel) per thread
  if (i >= B * CHUNKS * 144) return;
  const int pix = i % 144, ch = (i / 144) % CHUNKS, img = i / (144 * CHUNKS);
  float v[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int ci = ch * 4 + q;
    v[q] = ci < C1 ? a1[((size_t)img * C1 + ci) * 144 + pix] : 0.0f;
  }
  reinterpret_cast<float4*>(c4)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// ---- weights [20][10][5][5] -> B operand [tap 25][chunk 4][co 32][4 floats], zero padded ----------
__global__ void conv2_w_to_operand_kernel(const float* __restrict__ w, float* __restrict__ wb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;       // one float4 of the operand per thread
  if (i >= 25 * CHUNKS * NPAD) return;
  const int co = i % NPAD, ch = (i / NPAD) % CHUNKS, tap = i / (NPAD * CHUNKS);
  float v[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int ci = ch * 4 + q;
    v[q] = (co < C2 && ci < C1) ? __ldg(w + (co * C1 + ci) * 25 + tap) : 0.0f;
  }
  reinterpret_cast<float4*>(wb)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// ---- forward ---------------------------------------------------------------------------------
// Persistent and warp-specialised: a CTA loops over image pairs with a two-deep pipeline.
//   warp 4, one lane : TMA producer  (5-D load of the next pair into A[buf])
//   warp 5, one lane : MMA issuer    (50 tcgen05.mma per pair into TMEM accumulator buf; one commit
//                                      frees A[buf] for the producer, one hands the accumulator over)
//   warps 0-3        : epilogue out of TMEM (bias, Dropout2d, pool + arg-max, ReLU, stores)
// The weights are laid out once per CTA. mask_mode as in b200_conv2_fwd (0 none, 1 Philox + store,
// 2 given).
__device__ __forceinline__ void bar_init(uint64_t* b, uint32_t n) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n));
}
__device__ __forceinline__ void bar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}

__global__ void __launch_bounds__(CT_THREADS, 1)
conv2_fwd_tc2_kernel(const __grid_constant__ CUtensorMap map_a1, const float* __restrict__ w,
                     const float* __restrict__ b, float* __restrict__ mask, int mask_mode, float p,
                     unsigned long long seed, unsigned long long step, float* __restrict__ a2,
                     uint8_t* __restrict__ pos2, int B) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* As = smem;                       // [stage][chunk][Y][img][X][4 floats]
  uint8_t* Bs = As + STAGES * A_BYTES;      // [tap][chunk][co 32][4 floats]
  uint64_t* bars = reinterpret_cast<uint64_t*>(Bs + B_BYTES);
  uint64_t* a_full = bars;                  // [STAGES] TMA landed
  uint64_t* a_free = bars + STAGES;         // [STAGES] the MMAs that read A[stage] are done
  uint64_t* acc_full = bars + 2 * STAGES;   // [2] accumulator buf is complete
  uint64_t* acc_free = acc_full + 2;        // [2] the epilogue has drained accumulator buf
  uint64_t* w_full = acc_free + 2;          // the pre-laid-out weights have landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
  __shared__ float bias_s[NPAD];
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int npairs = (B + 1) / 2;

  if (t == 0) {
    for (int i = 0; i < STAGES; ++i) { bar_init(a_full + i, 1); bar_init(a_free + i, 1); }
    for (int i = 0; i < 2; ++i) { bar_init(acc_full + i, 1); bar_init(acc_free + i, 4); }
    bar_init(w_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(s32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (t < NPAD) bias_s[t] = (t < C2) ? b[t] : 0.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 && lane == 0) {
    // ---------------- TMA producer: the weights (already in operand layout, one 51 KB bulk copy)
    // and then the ring of activation tiles
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(w_full)), "r"((uint32_t)B_BYTES) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(Bs)),
                 "l"(w), "r"((uint32_t)B_BYTES), "r"(s32(w_full)) : "memory");
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int buf = it % STAGES;
      if (it >= STAGES) bar_wait(a_free + buf, ((it / STAGES) - 1) & 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(a_full + buf)), "r"((uint32_t)A_BYTES) : "memory");
      asm volatile(
          "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
          " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(s32(As + buf * A_BYTES)),
          "l"(&map_a1), "r"(s32(a_full + buf)), "r"(0), "r"(0), "r"(2 * pair), "r"(0), "r"(0)
          : "memory");
    }
  } else if (warp == 5 && lane == 0) {
    // ---------------- MMA issuer: only the start addresses move with the tap
    constexpr uint32_t idesc = idesc_tf32(128, NPAD);
    const uint32_t b0 = s32(Bs);
    bar_wait(w_full, 0);
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int buf = it & 1, st = it % STAGES;
      if (it >= 2) bar_wait(acc_free + buf, ((it >> 1) - 1) & 1);
      bar_wait(a_full + st, (it / STAGES) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a0 = s32(As + st * A_BYTES);
      uint32_t acc = 0;
#pragma unroll 1
      for (int tap = 0; tap < 25; ++tap) {
        const int ky = tap / 5, kx = tap - ky * 5;
#pragma unroll
        for (int kc = 0; kc < 2; ++kc) {
          const uint64_t ad = desc_k_noswz(a0 + (2 * kc) * A_CHUNK_BYTES + (ky * 24 + kx) * 16, A_CHUNK_BYTES, 12 * 16);
          const uint64_t bd = desc_k_noswz(b0 + tap * B_TAP_BYTES + (2 * kc) * (NPAD * 16), NPAD * 16, 128);
          mma_tf32(tmem + buf * NPAD, ad, bd, idesc, acc);
          acc = 1;
        }
      }
      mma_commit(a_free + st);
      mma_commit(acc_full + buf);
    }
  } else if (warp < 4) {
    // ---------------- epilogue: lane l of warp w holds GEMM row 32w + l = (g = 4w + l/8, x = l%8),
    // g = 2*y + image. The 2x2 window (py = w, px = x/2) of an image is lanes l, l^1, l^16, l^17.
    const int gl = lane >> 3, x = lane & 7;
    const int im = gl & 1, dy0 = gl >> 1, dx0 = x & 1;
    const int py = warp, px = x >> 1;
    const int j = dy0 * 2 + dx0;                 // scan order inside the window
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int buf = it & 1;
      const int img = 2 * pair + im;
      bar_wait(acc_full + buf, (it >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t r[32];
      tmem_ld32(tmem + buf * NPAD + ((uint32_t)(warp * 32) << 16), r);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) bar_arrive(acc_free + buf);   // the accumulator may be overwritten now
      const bool writer = (j == 0) && img < B;
#pragma unroll
      for (int co = 0; co < C2; ++co) {
        float v = __uint_as_float(r[co]) + bias_s[co];
        int jj = j;
#pragma unroll
        for (int o = 1; o <= 16; o <<= 4) {        // partners l^1, then l^16
          const float ov = __shfl_xor_sync(0xffffffffu, v, o);
          const int oj = __shfl_xor_sync(0xffffffffu, jj, o);
          if (ov > v || (ov == v && oj < jj)) { v = ov; jj = oj; }   // first max in scan order
        }
        if (writer) {
          float m = 1.0f;
          if (mask_mode == 1) {
            m = dropout_mask_value(seed, step, (uint32_t)img, 0x20000u + co, p);
            if (py == 0 && px == 0) mask[(size_t)img * C2 + co] = m;
          } else if (mask_mode == 2) {
            m = __ldg(mask + (size_t)img * C2 + co);
          }
          const size_t o = (size_t)img * A2_SZ + co * 16 + py * 4 + px;
          a2[o] = fmaxf(v * m, 0.0f);
          pos2[o] = (uint8_t)((2 * py + (jj >> 1)) * 12 + 2 * px + (jj & 1));
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// c4 activations [B][4][12][12][4 floats] seen as (e 4, X 12, image B, Y 12, chunk 4): the box
// order IS the shared-memory order [chunk][Y][image][X][4].
int make_c4_map(CUtensorMap* m, const float* c4, int B) {
  EncodeTiledFn enc = encode_fn();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  B200_REQUIRE(((uintptr_t)c4 & 15) == 0, "conv2_tc: c4 activations must be 16-byte aligned");
  const cuuint64_t dims[5] = {4, 12, (cuuint64_t)B, 12, 4};
  const cuuint64_t strides[4] = {16, (cuuint64_t)CHUNKS * 144 * 16, 12 * 16, 144 * 16};   // X, image, Y, chunk
  const cuuint32_t box[5] = {4, 12, 2, 12, 4};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 5, const_cast<float*>(c4), dims, strides,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (c4 activations) failed with code %d", (int)r);
  return 0;
}

}  // namespace

extern "C" {

// [B][10][12][12] fp32 -> c4 layout [B][4][12][12][4] (64 floats... 2304 floats per image).
int b200_a1_to_c4(const float* a1, float* c4, int B, void* stream) {
  if (B <= 0) return 0;
  const int n = B * CHUNKS * 144;
  a1_to_c4_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a1, c4, B);
  B200_LAUNCH_OK();
  return 0;
}

// conv2.weight [20][10][5][5] -> the tensor-core B operand [25 taps][4 chunks][32][4] (12,800
// floats, zero padded). Once per weight update.
int b200_conv2_w_to_operand(const float* w, float* w_operand, void* stream) {
  const int n = 25 * CHUNKS * NPAD;
  conv2_w_to_operand_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, w_operand);
  B200_LAUNCH_OK();
  return 0;
}
int b200_conv2_operand_floats(void) { return B_BYTES / 4; }

// conv2 forward + bias + Dropout2d + 2x2 max-pool + ReLU from c4 activations and the weight
// operand of b200_conv2_w_to_operand; same outputs as b200_conv2_fwd. Ref: src/model.py:17.
int b200_conv2_fwd_tc2(const float* a1_c4, const float* w, const float* b, float* mask, int mask_mode,
                       float p, unsigned long long seed, unsigned long long step, float* a2,
                       void* pos2, int B, void* stream) {
  if (B <= 0) return 0;
  B200_REQUIRE(mask_mode == 0 || mask != nullptr, "b200_conv2_fwd_tc2: mask buffer required");
  B200_REQUIRE(((uintptr_t)w & 15) == 0, "b200_conv2_fwd_tc2: the weight operand must be 16-byte aligned");
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv2_fwd_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr_done[dev & 63] = true;
  }
  CUtensorMap map;
  if (int e = make_c4_map(&map, a1_c4, B)) return e;
  const int npairs = (B + 1) / 2;
  conv2_fwd_tc2_kernel<<<npairs < 148 ? npairs : 148, CT_THREADS, FWD_SMEM, (cudaStream_t)stream>>>(
      map, w, b, mask, mask_mode, p, seed, step, a2, (uint8_t*)pos2, B);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
