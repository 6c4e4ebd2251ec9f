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

// start address + bytes: the field holds address >> 4 in the low 14 bits and a valid shared-memory
// address never carries out of it
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }

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
      const uint64_t ad0 = desc_k_noswz(s32(As + st * A_BYTES), A_CHUNK_BYTES, 12 * 16);
      const uint64_t bd0 = desc_k_noswz(b0, NPAD * 16, 128);
#pragma unroll
      for (int tap = 0; tap < 25; ++tap) {
        const int ky = tap / 5, kx = tap - ky * 5;
#pragma unroll
        for (int kc = 0; kc < 2; ++kc)
          mma_tf32(tmem + buf * NPAD, desc_advance(ad0, (2 * kc) * A_CHUNK_BYTES + (ky * 24 + kx) * 16),
                   desc_advance(bd0, tap * B_TAP_BYTES + (2 * kc) * (NPAD * 16)), idesc, (tap | kc) ? 1u : 0u);
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

// =============================================================================================
// Backward. The gradient before pooling, dZ [image][20][8][8], is nonzero at one position per
// pooling window. It is written into a zero-padded FRAME per image and 4-channel chunk:
//     frame pixel f = 4 + (y + 4) * 12 + x          (pitch 12, four zero rows above and below,
//                                                    x = 8..11 zero, 200 pixels per plane)
// The data gradient reads it with nothing but descriptor arithmetic:
//   da1[q = Y*12 + X][ci] = sum over taps t' = (ky', kx') and co of
//                           frame[q + ky'*12 + kx'][co] * Wflip[t'][co][ci]
//   an implicit GEMM like the forward: A = frame, K-major, rows are consecutive frame pixels
//   (SBO = 128 B); the four zero columns of a row are at once its right margin and the next
//   row's left margin; M = 144 rows per image in two tiles of 128.
// The weight gradient (further down) reduces over pixels and uses its own MN-major layout.
constexpr int FRAME = 200;                              // pixels per chunk plane of the dZ frame
constexpr int DZ_CHUNKS = 6;                            // 20 channels padded to 24
constexpr int DZ_PLANE_BYTES = FRAME * 16;              // 3200
constexpr int DZ_IMG_BYTES = DZ_CHUNKS * DZ_PLANE_BYTES;   // 19200
constexpr int DZ_ORIGIN = 4 + 4 * 12;                   // frame pixel of (y, x) = (0, 0): 52
constexpr int CIPAD = 16;                               // 10 input channels padded
constexpr int WD_TAP_BYTES = DZ_CHUNKS * CIPAD * 16;    // 1536
constexpr int WD_BYTES = 25 * WD_TAP_BYTES;             // 38400
constexpr int DG_STAGES = 3;
// dZ frame [B][6][200][4] from the pooled gradient: one float4 (pixel, chunk) per thread.
__global__ void conv2_dz_frame_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                                      const uint8_t* __restrict__ pos2, const float* __restrict__ mask,
                                      float* __restrict__ frame, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * DZ_CHUNKS * FRAME) return;
  const int f = i % FRAME, ch = (i / FRAME) % DZ_CHUNKS, img = i / (FRAME * DZ_CHUNKS);
  float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  const int q = f - DZ_ORIGIN;                     // y*12 + x of the pre-pool position
  if (q >= 0 && ch < 5) {
    const int y = q / 12, x = q - y * 12;
    if (y < 8 && x < 8) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int co = ch * 4 + c;
        const size_t gi = (size_t)img * A2_SZ + co * 16 + (y >> 1) * 4 + (x >> 1);
        if (pos2[gi] == q && a2[gi] > 0.0f)
          v[c] = da2[gi] * (mask ? mask[(size_t)img * C2 + co] : 1.0f);
      }
    }
  }
  reinterpret_cast<float4*>(frame)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// weights [20][10][5][5] -> data-gradient B operand [tap' 25][chunk 6][ci 16][4 co], flipped taps
__global__ void conv2_w_to_dgrad_operand_kernel(const float* __restrict__ w, float* __restrict__ wd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 25 * DZ_CHUNKS * CIPAD) return;
  const int ci = i % CIPAD, ch = (i / CIPAD) % DZ_CHUNKS, tp = i / (CIPAD * DZ_CHUNKS);
  const int tap = 24 - tp;                         // (4 - ky') * 5 + (4 - kx')
  float v[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int co = ch * 4 + c;
    v[c] = (co < C2 && ci < C1) ? __ldg(w + (co * C1 + ci) * 25 + tap) : 0.0f;
  }
  reinterpret_cast<float4*>(wd)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)),
               "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void bar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
}

// ---- data gradient -----------------------------------------------------------------------------
// Persistent, one image PAIR per pipeline stage: the two frames sit side by side in every chunk
// plane ([chunk 6][image 2][200 pixels][4]), so the GEMM rows run through both images (rows
// 0..143 image 0, 200..343 image 1, the rest unused) and three M tiles of 128 serve two images.
// warp 4 lane 0: bulk copies (the flipped weights once, then 6 planes per image);
// warps 5-7 lane 0: one M tile each, 25 taps x 3 K-steps of tcgen05.mma.kind::tf32 (M=128, N=16,
// K=8) (the issue path of one thread, not the tensor pipe, bounds a stream of small MMAs);
// warps 0-3: TMEM -> da1 [B][10][12][12].
constexpr int DG_PLANE_BYTES = 2 * DZ_PLANE_BYTES;          // 6400: one chunk plane of a pair
constexpr int DG_STAGE_BYTES = DZ_CHUNKS * DG_PLANE_BYTES;  // 38400
constexpr int DG_PAIR_SMEM = DG_STAGES * DG_STAGE_BYTES + WD_BYTES + 256 + 1024;

__global__ void __launch_bounds__(CT_THREADS, 1)
conv2_dgrad_tc_kernel(const float* __restrict__ frame, const float* __restrict__ wd,
                      float* __restrict__ da1, int B) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* As = smem;                              // [stage][chunk 6][image 2][200 pixels][4 floats]
  uint8_t* Bs = As + DG_STAGES * DG_STAGE_BYTES;   // [tap'][chunk 6][ci 16][4 floats]
  uint64_t* bars = reinterpret_cast<uint64_t*>(Bs + WD_BYTES);
  uint64_t* a_full = bars;
  uint64_t* a_free = bars + DG_STAGES;
  uint64_t* acc_full = bars + 2 * DG_STAGES;
  uint64_t* acc_free = acc_full + 2;
  uint64_t* w_full = acc_free + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int npairs = (B + 1) / 2;

  if (t == 0) {
    for (int i = 0; i < DG_STAGES; ++i) { bar_init(a_full + i, 1); bar_init(a_free + i, 3); }
    for (int i = 0; i < 2; ++i) { bar_init(acc_full + i, 3); bar_init(acc_free + i, 4); }
    bar_init(w_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(s32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 && lane == 0) {
    bar_expect(w_full, WD_BYTES);
    bulk_load(Bs, wd, WD_BYTES, w_full);
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int st = it % DG_STAGES;
      if (it >= DG_STAGES) bar_wait(a_free + st, ((it / DG_STAGES) - 1) & 1);
      const int nimg = (2 * pair + 1 < B) ? 2 : 1;
      bar_expect(a_full + st, (uint32_t)(nimg * DZ_IMG_BYTES));
      for (int im = 0; im < nimg; ++im) {
        const float* src = frame + (size_t)(2 * pair + im) * (DZ_IMG_BYTES / 4);
#pragma unroll
        for (int c = 0; c < DZ_CHUNKS; ++c)
          bulk_load(As + st * DG_STAGE_BYTES + c * DG_PLANE_BYTES + im * DZ_PLANE_BYTES,
                    src + c * (DZ_PLANE_BYTES / 4), DZ_PLANE_BYTES, a_full + st);
      }
    }
  } else if (warp >= 5 && lane == 0) {
    // three issuing threads, one per M tile; each commits its own MMAs to the (count 3) barriers
    constexpr uint32_t idesc = idesc_tf32(128, CIPAD);
    const int tile = warp - 5;
    bar_wait(w_full, 0);
    const uint64_t bd0 = desc_k_noswz(s32(Bs), CIPAD * 16, 128);
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int buf = it & 1, st = it % DG_STAGES;
      if (it >= 2) bar_wait(acc_free + buf, ((it >> 1) - 1) & 1);
      bar_wait(a_full + st, (it / DG_STAGES) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint64_t ad0 = desc_k_noswz(s32(As + st * DG_STAGE_BYTES) + tile * (128 * 16), DG_PLANE_BYTES, 128);
      const uint32_t d = tmem + buf * 64 + tile * CIPAD;
#pragma unroll
      for (int tp = 0; tp < 25; ++tp) {
        const int ky = tp / 5, kx = tp - ky * 5;
#pragma unroll
        for (int kc = 0; kc < 3; ++kc)
          mma_tf32(d, desc_advance(ad0, (ky * 12 + kx) * 16 + (2 * kc) * DG_PLANE_BYTES),
                   desc_advance(bd0, tp * WD_TAP_BYTES + (2 * kc) * (CIPAD * 16)), idesc, (tp | kc) ? 1u : 0u);
      }
      mma_commit(a_free + st);
      mma_commit(acc_full + buf);
    }
  } else if (warp < 4) {
    // lane l of warp w holds row 32w + l of each tile: tile 0 in columns 0..15, tile 1 in 16..31,
    // tile 2 in 32..47 of the buffer
    int it = 0;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x, ++it) {
      const int buf = it & 1;
      bar_wait(acc_full + buf, (it >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t r[32], r2[32];
      tmem_ld32(tmem + buf * 64 + ((uint32_t)(warp * 32) << 16), r);
      tmem_ld32(tmem + buf * 64 + 32 + ((uint32_t)(warp * 32) << 16), r2);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) bar_arrive(acc_free + buf);
      const bool two = 2 * pair + 1 < B;
      float* out0 = da1 + (size_t)(2 * pair) * A1_SZ;
      float* out1 = out0 + A1_SZ;
      const int row0 = warp * 32 + lane, row1 = 128 + row0, row2 = 256 + row0;
      // tile 1: rows 128..143 finish image 0, rows 200..255 start image 1; tile 2: rows 256..343
      float* o1 = nullptr;
      if (row1 < 144) o1 = out0 + row1;
      else if (row1 >= FRAME && two) o1 = out1 + (row1 - FRAME);
      float* o2 = (two && row2 - FRAME < 144) ? out1 + (row2 - FRAME) : nullptr;
#pragma unroll
      for (int ci = 0; ci < C1; ++ci) {
        out0[ci * 144 + row0] = __uint_as_float(r[ci]);
        if (o1) o1[ci * 144] = __uint_as_float(r[CIPAD + ci]);
        if (o2) o2[ci * 144] = __uint_as_float(r2[ci]);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
  }
}

// ---- weight (and bias) gradient ----------------------------------------------------------------
// The reduction runs over pixels, so both operands are MN-major. For TF32 the tensor core reads
// MN-major operands in one layout only (measured, scripts/experiments/umma_mn_probe.cu: every
// other layout type returns zeros): layout type 1, "128-byte swizzle with 32-byte atoms":
//     a K row (= pixel) is 128 bytes = 32 channels; 4 K rows per group, groups SBO apart;
//     the next 32 M/N elements are LBO away; 32-byte unit u of K row r sits at u ^ (r & 3),
//     taken from the ABSOLUTE shared-memory address (bits 5-6 ^= bits 7-8).
// Both operands therefore use a "c32" layout, [pixel][32 channels], written pre-swizzled in
// global memory so that plain bulk copies into 512-byte aligned tiles keep the pattern:
//     A = a1  [160 pixels][32]: channels 0..9, channel 10 = 1 (its tap-0 gradient is the bias
//             gradient), rest 0; pixels 144..159 zero
//     B = dZ  [96 pixels][32]:  pitch-12 rows y = 0..7, x = 8..11 zero, channels 20..31 zero
// A's LBO is ONE pixel row: the four 32-channel groups of M = 128 are four consecutive shifts
// j0..j0+3 of the image, so one MMA yields dW[:, :, ky, kx..kx+3] (j0 = ky*12 + kx); ten
// accumulators (5 filter rows x {kx 0..3, kx 4}) cover the filter. Per image: 8 K-steps (one per dZ
// row: its 8 real pixels) x 10 tcgen05.mma (M=128, N=32, K=8). The CTAs' sums meet in fp32 atomics.
constexpr int C32_A1_PIX = 160;
constexpr int C32_A1_BYTES = C32_A1_PIX * 128;          // 20480
constexpr int C32_DZ_PIX = 96;
constexpr int C32_DZ_BYTES = C32_DZ_PIX * 128;          // 12288
constexpr int WG_STAGE_BYTES = C32_A1_BYTES + C32_DZ_BYTES;   // 32768
constexpr int WG_STAGES = 4;
constexpr int WG_SMEM = WG_STAGES * WG_STAGE_BYTES + 256 + 1024;

__device__ __forceinline__ uint64_t desc_mn_sw32atom(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return desc_k_noswz(addr, lbo, sbo) | ((uint64_t)1 << 61);    // layout type 1
}
constexpr uint32_t idesc_tf32_mn(int M, int N) {        // both operands MN-major
  return idesc_tf32(M, N) | (1u << 15) | (1u << 16);
}

// a1 [B][10][12][12] -> swizzled c32 [B][160][32]; one 32-byte unit per thread
__global__ void a1_to_c32_kernel(const float* __restrict__ a1, float* __restrict__ c32, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C32_A1_PIX * 4) return;
  const int u = i & 3, pix = (i >> 2) % C32_A1_PIX, img = i / (4 * C32_A1_PIX);
  float v[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int ci = u * 8 + q;
    v[q] = 0.0f;
    if (pix < 144) v[q] = ci < C1 ? a1[((size_t)img * C1 + ci) * 144 + pix] : (ci == C1 ? 1.0f : 0.0f);
  }
  float4* dst = reinterpret_cast<float4*>(c32 + ((size_t)img * C32_A1_PIX + pix) * 32 + ((u ^ (pix & 3)) * 8));
  dst[0] = make_float4(v[0], v[1], v[2], v[3]);
  dst[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// pooled gradient -> swizzled c32 dZ [B][96][32]; one 32-byte unit per thread
__global__ void conv2_dz_c32_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                                    const uint8_t* __restrict__ pos2, const float* __restrict__ mask,
                                    float* __restrict__ dz, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C32_DZ_PIX * 4) return;
  const int u = i & 3, q = (i >> 2) % C32_DZ_PIX, img = i / (4 * C32_DZ_PIX);
  const int y = q / 12, x = q - y * 12;
  float v[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int co = u * 8 + c;
    v[c] = 0.0f;
    if (x < 8 && co < C2) {
      const size_t gi = (size_t)img * A2_SZ + co * 16 + (y >> 1) * 4 + (x >> 1);
      if (pos2[gi] == q && a2[gi] > 0.0f) v[c] = da2[gi] * (mask ? mask[(size_t)img * C2 + co] : 1.0f);
    }
  }
  float4* dst = reinterpret_cast<float4*>(dz + ((size_t)img * C32_DZ_PIX + q) * 32 + ((u ^ (q & 3)) * 8));
  dst[0] = make_float4(v[0], v[1], v[2], v[3]);
  dst[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// Both dZ layouts of one image from one pass over its pooled gradient: the 320 values are
// scattered into shared-memory images of the frame (data gradient) and of the c32 operand (weight
// gradient), which then go out as coalesced 16-byte stores. One CTA per image.
constexpr int DZ_BOTH_THREADS = 256;
__global__ void __launch_bounds__(DZ_BOTH_THREADS)
conv2_dz_both_kernel(const float* __restrict__ da2, const float* __restrict__ a2,
                     const uint8_t* __restrict__ pos2, const float* __restrict__ mask,
                     float* __restrict__ frame, float* __restrict__ dz32, int B) {
  __shared__ __align__(16) float fs[DZ_IMG_BYTES / 4];     // [chunk 6][200 pixels][4]
  __shared__ __align__(16) float cs[C32_DZ_BYTES / 4];     // [96 pixels][32], swizzled
  const int t = threadIdx.x;
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    for (int i = t; i < DZ_IMG_BYTES / 16; i += DZ_BOTH_THREADS)
      reinterpret_cast<float4*>(fs)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = t; i < C32_DZ_BYTES / 16; i += DZ_BOTH_THREADS)
      reinterpret_cast<float4*>(cs)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    for (int i = t; i < A2_SZ; i += DZ_BOTH_THREADS) {      // one pooled value per thread
      const size_t gi = (size_t)img * A2_SZ + i;
      const int co = i >> 4;
      const float v = (a2[gi] > 0.0f) ? da2[gi] * (mask ? mask[(size_t)img * C2 + co] : 1.0f) : 0.0f;
      const int q = pos2[gi];                               // y*12 + x of the window's arg-max
      fs[((co >> 2) * FRAME + DZ_ORIGIN + q) * 4 + (co & 3)] = v;
      cs[q * 32 + (((co >> 3) ^ (q & 3)) * 8) + (co & 7)] = v;
    }
    __syncthreads();
    float4* fo = reinterpret_cast<float4*>(frame + (size_t)img * (DZ_IMG_BYTES / 4));
    for (int i = t; i < DZ_IMG_BYTES / 16; i += DZ_BOTH_THREADS) fo[i] = reinterpret_cast<const float4*>(fs)[i];
    float4* co4 = reinterpret_cast<float4*>(dz32 + (size_t)img * (C32_DZ_BYTES / 4));
    for (int i = t; i < C32_DZ_BYTES / 16; i += DZ_BOTH_THREADS) co4[i] = reinterpret_cast<const float4*>(cs)[i];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(CT_THREADS, 1)
conv2_wgrad_tc_kernel(const float* __restrict__ a1_c32, const float* __restrict__ dz_c32,
                      float* __restrict__ dw, float* __restrict__ db, int B) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* St = smem;                              // [stage]{a1 c32 20480 B, dZ c32 12288 B}
  uint64_t* bars = reinterpret_cast<uint64_t*>(St + WG_STAGES * WG_STAGE_BYTES);
  uint64_t* s_full = bars;
  uint64_t* s_free = bars + WG_STAGES;
  uint64_t* done = bars + 2 * WG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

  if (t == 0) {
    for (int i = 0; i < WG_STAGES; ++i) { bar_init(s_full + i, 1); bar_init(s_free + i, 2); }
    bar_init(done, 2);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(s32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 && lane == 0) {
    int it = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
      const int st = it % WG_STAGES;
      if (it >= WG_STAGES) bar_wait(s_free + st, ((it / WG_STAGES) - 1) & 1);
      uint8_t* dst = St + st * WG_STAGE_BYTES;
      bar_expect(s_full + st, WG_STAGE_BYTES);
      bulk_load(dst, a1_c32 + (size_t)img * (C32_A1_BYTES / 4), C32_A1_BYTES, s_full + st);
      bulk_load(dst + C32_A1_BYTES, dz_c32 + (size_t)img * (C32_DZ_BYTES / 4), C32_DZ_BYTES, s_full + st);
    }
  } else if ((warp == 5 || warp == 6) && lane == 0) {
    // two issuing threads with five accumulators each (see the data-gradient kernel)
    constexpr uint32_t idesc = idesc_tf32_mn(128, NPAD);
    const int abase = (warp - 5) * 5;
    int it = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
      const int st = it % WG_STAGES;
      bar_wait(s_full + st, (it / WG_STAGES) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a0 = s32(St + st * WG_STAGE_BYTES);
      const uint64_t ad0 = desc_mn_sw32atom(a0, 128, 512);                    // LBO = one pixel
      const uint64_t bd0 = desc_mn_sw32atom(a0 + C32_A1_BYTES, 128, 512);     // N = 32: one group
      // one K-step per dZ row: its 8 real pixels (the row's other 4 are zero and are skipped)
#pragma unroll 2
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t acc = (it > 0 || ks > 0) ? 1u : 0u;
        const uint64_t bd = desc_advance(bd0, ks * (12 * 128));
        const uint64_t adk = desc_advance(ad0, ks * (12 * 128));
#pragma unroll
        for (int i = 0; i < 5; ++i) {
          const int a = abase + i;
          mma_tf32(tmem + a * NPAD, desc_advance(adk, ((a >> 1) * 12 + (a & 1) * 4) * 128), bd, idesc, acc);
        }
      }
      mma_commit(s_free + st);
    }
    mma_commit(done);
  }
  if (warp < 4) {
    bar_wait(done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // lane l of warp w: GEMM row 32w + l = (shift w inside the window of 4, channel l)
#pragma unroll 1
    for (int a = 0; a < 10; ++a) {
      uint32_t r[32];
      tmem_ld32(tmem + a * NPAD + ((uint32_t)(warp * 32) << 16), r);
      const int ky = a >> 1, kx = (a & 1) * 4 + warp, ci = lane;
      if (kx < 5 && ci < C1) {
#pragma unroll
        for (int co = 0; co < C2; ++co)
          atomicAdd(dw + (co * C1 + ci) * 25 + ky * 5 + kx, __uint_as_float(r[co]));
      } else if (ci == C1 && a == 0 && warp == 0) {   // the ones channel at shift 0: bias gradient
#pragma unroll
        for (int co = 0; co < C2; ++co) atomicAdd(db + co, __uint_as_float(r[co]));
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
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

// ---- backward entry points ---------------------------------------------------------------------
int b200_conv2_frame_floats(int B) { return B * (DZ_IMG_BYTES / 4); }
int b200_conv2_dgrad_operand_floats(void) { return WD_BYTES / 4; }

// Pooled gradient da2 [B][20][4][4] (+ a2, pos2, Dropout2d mask or null) -> zero-padded dZ frame
// [B][6][200][4] that both gradient kernels read. Ref: autograd of src/model.py:24.
int b200_conv2_dz_frame(const float* da2, const float* a2, const void* pos2, const float* mask,
                        float* frame, int B, void* stream) {
  if (B <= 0) return 0;
  const int n = B * DZ_CHUNKS * FRAME;
  conv2_dz_frame_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(da2, a2, (const uint8_t*)pos2,
                                                                           mask, frame, B);
  B200_LAUNCH_OK();
  return 0;
}

// conv2.weight -> flipped data-gradient operand (b200_conv2_dgrad_operand_floats() floats).
int b200_conv2_w_to_dgrad_operand(const float* w, float* wd, void* stream) {
  const int n = 25 * DZ_CHUNKS * CIPAD;
  conv2_w_to_dgrad_operand_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, wd);
  B200_LAUNCH_OK();
  return 0;
}

// da1 [B][10][12][12] = conv2 data gradient of the dZ frame (tcgen05, TF32 operands).
int b200_conv2_dgrad_tc(const float* frame, const float* wd, float* da1, int B, void* stream) {
  if (B <= 0) return 0;
  B200_REQUIRE((((uintptr_t)frame | (uintptr_t)wd) & 15) == 0, "b200_conv2_dgrad_tc: operands must be 16-byte aligned");
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv2_dgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_PAIR_SMEM));
    attr_done[dev & 63] = true;
  }
  const int npairs = (B + 1) / 2;
  conv2_dgrad_tc_kernel<<<npairs < 148 ? npairs : 148, CT_THREADS, DG_PAIR_SMEM, (cudaStream_t)stream>>>(frame, wd, da1, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_conv2_c32_a1_floats(int B) { return B * (C32_A1_BYTES / 4); }
int b200_conv2_c32_dz_floats(int B) { return B * (C32_DZ_BYTES / 4); }

// a1 [B][10][12][12] -> the weight gradient's A operand (b200_conv2_c32_a1_floats(B) floats).
int b200_a1_to_c32(const float* a1, float* c32, int B, void* stream) {
  if (B <= 0) return 0;
  const int n = B * C32_A1_PIX * 4;
  a1_to_c32_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a1, c32, B);
  B200_LAUNCH_OK();
  return 0;
}

// pooled gradient -> the weight gradient's B operand (b200_conv2_c32_dz_floats(B) floats).
int b200_conv2_dz_c32(const float* da2, const float* a2, const void* pos2, const float* mask,
                      float* dz, int B, void* stream) {
  if (B <= 0) return 0;
  const int n = B * C32_DZ_PIX * 4;
  conv2_dz_c32_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(da2, a2, (const uint8_t*)pos2,
                                                                         mask, dz, B);
  B200_LAUNCH_OK();
  return 0;
}

// Both dZ layouts in one pass (what b200_conv2_dz_frame + b200_conv2_dz_c32 produce).
int b200_conv2_dz_both(const float* da2, const float* a2, const void* pos2, const float* mask,
                       float* frame, float* dz_c32, int B, void* stream) {
  if (B <= 0) return 0;
  conv2_dz_both_kernel<<<B < 148 * 4 ? B : 148 * 4, DZ_BOTH_THREADS, 0, (cudaStream_t)stream>>>(
      da2, a2, (const uint8_t*)pos2, mask, frame, dz_c32, B);
  B200_LAUNCH_OK();
  return 0;
}

// dw [20][10][5][5] and db [20] (both overwritten) from the two c32 operands (tcgen05, TF32
// operands; the CTAs' partial sums meet in fp32 atomics).
int b200_conv2_wgrad_tc(const float* a1_c32, const float* dz_c32, float* dw, float* db, int B, void* stream) {
  B200_REQUIRE(B > 0, "b200_conv2_wgrad_tc: empty batch");
  B200_REQUIRE((((uintptr_t)dz_c32 | (uintptr_t)a1_c32) & 511) == 0, "b200_conv2_wgrad_tc: operands must be 512-byte aligned");
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv2_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    attr_done[dev & 63] = true;
  }
  B200_CUDA_OK(cudaMemsetAsync(dw, 0, W2_SZ * sizeof(float), (cudaStream_t)stream));
  B200_CUDA_OK(cudaMemsetAsync(db, 0, B2_SZ * sizeof(float), (cudaStream_t)stream));
  conv2_wgrad_tc_kernel<<<B < 148 ? B : 148, CT_THREADS, WG_SMEM, (cudaStream_t)stream>>>(a1_c32, dz_c32, dw, db, B);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
