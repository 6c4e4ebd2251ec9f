// bf16 GEMM on the 5th-generation tensor cores, hand-written for sm_100a (SURVEY row 8.16,
// BASELINE.json config 4: the 784-4096-4096-10 MLP).
//
//   C[M,N] = A * B^T (+ bias) (ReLU) (masked by relu_mask > 0),   fp32 accumulation
//
// One kernel covers the three products of a linear layer y = x W^T + b
// (reference layers: nn.Linear in src/model.py:12-13, widened per BASELINE.json):
//   forward  y  = x  W^T     A = x  [M,K] K-major,        B = W  [N,K] K-major
//   dgrad    dx = dy W       A = dy [M,Kred] K-major,      B = W  [Kred,N] MN-major
//   wgrad    dW = dy^T x     A = dy [Kred,M] MN-major,     B = x  [Kred,N] MN-major
// so no operand is ever transposed in memory: the majorness goes into the TMA box shape,
// the shared-memory matrix descriptor and two bits of the instruction descriptor.
//
// Roles (192 threads per CTA, in all three kernels):
//   warp 0      TMA producer: cp.async.bulk.tensor.2d into a STAGES-deep ring of
//               128B-swizzled A/B tiles, completion on `full` mbarriers (expect_tx)
//   warp 1      TMEM allocator + MMA issuer: one elected lane issues tcgen05.mma.kind::f16
//               (K=16) four times per 64-wide K block; tcgen05.commit releases the smem
//               stage (`empty`) and finally signals `tmem_full`
//   warps 2..5  epilogue: tcgen05.ld 32x32b.x32 (each warp its own 32 TMEM lanes = 32 rows),
//               bias / ReLU / ReLU-mask, bf16 or fp32 stores with edge masking
// Kernels (host-side choose() picks by shape):
//   gemm_bf16_tcgen05_kernel          one 128 x BN tile per CTA (BN 16..256), cta_group::1
//   gemm_bf16_tcgen05_2cta_kernel     one 256 x BN tile per CTA pair, cta_group::2
//   gemm_bf16_tcgen05_persist_kernel  CTA pairs looping over tiles, two TMEM accumulators so
//                                     the epilogue of tile i overlaps the main loop of i+1
// TMA zero-fills out-of-bounds reads, so M, N, K need no padding (only 16-byte row pitch).
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "comm.cuh"

extern "C" int b200_pdl_enabled(void);

namespace {

constexpr int BM = 128;
constexpr int BK = 64;          // bf16 elements per K block = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 192;

// ---------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
// Bounded: a protocol bug traps (CUDA error) instead of hanging the GPU. The wait passes a
// 1 ms suspend-time hint so the four epilogue warps SLEEP on tmem_full during the main loop
// instead of busy-polling (polling stole issue slots from the MMA thread: -13 % throughput).
// The hardware returns from one suspended try_wait after ~0.5 us (ncu: ~380 iterations per
// 200 us wait), so the bound of 2^22 iterations is ~2 s; it is an iteration count, not a
// clock read per poll.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t.reg .u32 n;\n\t"
      "mov.u32 n, 0;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 1000000;\n\t"
      "@p bra DONE_%=;\n\t"
      "add.u32 n, n, 1;\n\t"
      "setp.lt.u32 q, n, 4194304;\n\t"
      "@q bra WAIT_%=;\n\t"
      "mov.u32 %0, 0;\n\t"
      "bra END_%=;\n\t"
      "DONE_%=:\n\t"
      "mov.u32 %0, 1;\n\t"
      "END_%=:\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  if (!ok) __trap();
}
// Wait used by the two single-thread roles (TMA producer, MMA issuer). Same as mbar_wait
// unless built with -DB200_GEMM_SPIN (experiment: plain try_wait polling, no suspend hint).
__device__ __forceinline__ void mbar_wait_role(uint64_t* bar, uint32_t parity) {
#ifdef B200_GEMM_SPIN
  uint32_t done = 0;
  for (uint32_t n = 0; !done; ++n) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (n > (1u << 28)) __trap();
  }
#else
  mbar_wait(bar, parity);
#endif
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout), 128B swizzle.
//   K-major tile  : rows x 128 B; 8-row groups 1024 B apart  -> LBO = 1 (unused), SBO = 1024 B
//   MN-major tile : [K rows][64 MN elems = 128 B]; the next 64 MN elements live `lbo_bytes`
//                   further (one TMA box later); 8-row K groups 1024 B apart -> SBO = 1024 B
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes,
                                                   uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version for sm_100
  d |= (uint64_t)2 << 61;   // SWIZZLE_128B
  return d;
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): bf16 x bf16 -> fp32.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4)                    // c_format  = F32
         | (1u << 7)                  // a_format  = BF16
         | (1u << 10)                 // b_format  = BF16
         | ((uint32_t)a_mn << 15)     // a_major   (1 = MN-major)
         | ((uint32_t)b_mn << 16)     // b_major
         | ((uint32_t)(N >> 3) << 17) // n_dim
         | ((uint32_t)(M >> 4) << 24);// m_dim
}

struct GemmParams {
  int M, N, K;
  void* C;
  int ldc;               // elements
  int c_fp32;            // 1: fp32 output, 0: bf16
  const __nv_bfloat16* bias;       // [N] or null
  int relu;
  const __nv_bfloat16* relu_mask;  // [M, ld_mask] or null: output zeroed where mask <= 0
  int ld_mask;
  // Fused reduce-scatter (wgrad of a data-parallel layer): instead of storing the tile to
  // C, the epilogue stores every 16-byte vector straight into the staging slot of the rank
  // that OWNS that slice of the gradient bucket (two-shot all-reduce, csrc/comm.cu), over
  // NVLink peer memory, while other tiles are still in their main loop.
  int rs_on;
  CommDev rs_comm;
  size_t rs_stage_off;   // arena offset of the bucket's staging [2][world][src_vec]
  size_t rs_src_vec;     // vectors reserved per source rank in one slot
  size_t rs_sl;          // vectors per owner slice
  size_t rs_base_vec;    // vector index of C[0][0] inside the bucket
  const unsigned* rs_seq;  // the bucket channel's sequence counter (slot = (seq+1) & 1)
  // Split-K (persistent kernel only): with ksplit = 2 every output tile is two work units.
  // The unit with the upper half of K stores its raw fp32 accumulators into sk_ws
  // [tile][256][BN] and raises sk_flags[tile]; the unit with the lower half adds them in its
  // epilogue. The flags count 8 producer warps + 8 consumer warps and reset themselves.
  int ksplit;
  float* sk_ws;
  unsigned* sk_flags;
};

template <int BN, bool B_MN> struct Cfg {
  static constexpr int A_BYTES = BM * BK * 2;   // 16 KB
  // an MN-major B tile is made of whole 64-column boxes even when BN < 64
  static constexpr int B_BYTES = ((B_MN && BN < 64) ? 64 : BN) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN >= 256) ? 4 : ((BN >= 128) ? 6 : 8);
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
};

// Second half of a staged bf16 epilogue: the warp's 32 rows sit in shared memory
// (PITCH = BN*2+16 bytes apart); every store instruction now covers whole 16-byte-vector rows
// (512 contiguous bytes for BN = 256), either to C or, with the reduce-scatter epilogue, to the
// staging slot of the rank that owns that slice of the gradient bucket over NVLink.
template <int BN>
__device__ __forceinline__ void store_staged_rows(const GemmParams& p, const uint8_t* sbuf, int lg,
                                                  int lane, int m0, int n0) {
  constexpr int PITCH = BN * 2 + 16;
  constexpr int VPR = BN / 8;                   // 16-byte vectors per tile row
  constexpr int RPI = 32 / VPR;                 // rows written per warp-wide store
  const int cv = lane % VPR;
  const int col = n0 + cv * 8;
  const bool vec_ok = (p.ldc % 8) == 0 && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
  unsigned slot = 0;
  size_t slot_vec = 0;
  if (p.rs_on) {
    slot = (*p.rs_seq + 1u) & 1u;
    slot_vec = (size_t)p.rs_comm.world * p.rs_src_vec;
  }
#pragma unroll 1
  for (int it = 0; it < VPR; ++it) {           // 32 rows / RPI rows per iteration == VPR
    const int rl = it * RPI + lane / VPR;
    const int grow = m0 + lg * 32 + rl;
    if (grow >= p.M || col >= p.N) continue;
    const uint4 val = *reinterpret_cast<const uint4*>(sbuf + rl * PITCH + cv * 16);
    if (p.rs_on) {
      // reduce-scatter epilogue: vector `vec` of the bucket goes to owner vec / sl
      if (col + 8 <= p.N) {
        const size_t vec = p.rs_base_vec + ((size_t)grow * p.ldc + col) / 8;
        const int owner = (int)(vec / p.rs_sl);
        uint4* dst = peer_ptr<uint4>(p.rs_comm, owner, p.rs_stage_off) + slot * slot_vec +
                     (size_t)p.rs_comm.rank * p.rs_src_vec + (vec - (size_t)owner * p.rs_sl);
        *dst = val;
      }
    } else {
      __nv_bfloat16* cptr = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)grow * p.ldc + col;
      if (vec_ok && col + 8 <= p.N) {
        *reinterpret_cast<uint4*>(cptr) = val;
      } else {
        const __nv_bfloat16* sv = reinterpret_cast<const __nv_bfloat16*>(&val);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (col + j < p.N) cptr[j] = sv[j];
      }
    }
  }
}

// ===== epilogue (shared by the 1-CTA and the 2-CTA kernels): warps 2..5 =====
template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint8_t* tiles,
                                              uint64_t* tmem_full, uint32_t tmem_base, int warp,
                                              int lane, int m0, int n0) {
  // Accumulator rows arrive one per lane (tcgen05.ld 32x32b). Bias / ReLU / mask are applied
  // in registers. bf16 results are then staged through shared memory (the tile ring is free
  // once tmem_full has fired) so each warp writes whole 16-byte-vector rows: a store
  // instruction covers 512 contiguous bytes instead of 32 scattered 16-byte pieces. That
  // matters most for the reduce-scatter epilogue, whose stores cross NVLink.
  const int lg = warp & 3;
  mbar_wait(tmem_full, 0);
  tcgen05_fence_after();
#ifdef B200_GEMM_SKIP_EPI
  return;   // experiment: main-loop time only
#endif
  const int row = m0 + lg * 32 + lane;
  const bool row_ok = row < p.M;
  constexpr int CHUNK = BN >= 32 ? 32 : 16;
  constexpr int PITCH = BN * 2 + 16;                       // bytes per staged row
  uint8_t* sbuf = tiles + lg * (32 * PITCH);
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += CHUNK) {
    uint32_t r[CHUNK];
    const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0;
    if constexpr (CHUNK == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r);
    tmem_ld_wait();
    const int col0 = n0 + c0;
    float v[CHUNK];
#pragma unroll
    for (int j = 0; j < CHUNK; ++j) v[j] = __uint_as_float(r[j]);
    if (p.bias) {
#pragma unroll
      for (int j = 0; j < CHUNK; ++j)
        if (col0 + j < p.N) v[j] += __bfloat162float(p.bias[col0 + j]);
    }
    if (p.relu) {
#pragma unroll
      for (int j = 0; j < CHUNK; ++j) v[j] = fmaxf(v[j], 0.0f);
    }
    if (p.relu_mask && row_ok) {
      const __nv_bfloat16* mrow = p.relu_mask + (size_t)row * p.ld_mask + col0;
      if (col0 + CHUNK <= p.N && (p.ld_mask % 8) == 0) {
#pragma unroll
        for (int j = 0; j < CHUNK; j += 8) {
          const uint4 mv = *reinterpret_cast<const uint4*>(mrow + j);
          const uint32_t w[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (__uint_as_float(w[q] << 16) <= 0.0f) v[j + 2 * q] = 0.0f;
            if (__uint_as_float(w[q] & 0xffff0000u) <= 0.0f) v[j + 2 * q + 1] = 0.0f;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < CHUNK; ++j)
          if (col0 + j < p.N && __bfloat162float(mrow[j]) <= 0.0f) v[j] = 0.0f;
      }
    }
    if (p.c_fp32) {
      if (row_ok) {
        float* crow = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
        if (col0 + CHUNK <= p.N && (p.ldc % 4) == 0) {
#pragma unroll
          for (int j = 0; j < CHUNK; j += 4)
            *reinterpret_cast<float4*>(crow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < CHUNK; ++j)
            if (col0 + j < p.N) crow[j] = v[j];
        }
      }
    } else {
      // pack to bf16 and park this lane's row segment in shared memory
#pragma unroll
      for (int j = 0; j < CHUNK; j += 8) {
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(v[j + 2 * q], v[j + 2 * q + 1]);
          w[q] = *reinterpret_cast<const uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(sbuf + lane * PITCH + (c0 + j) * 2) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
  if (!p.c_fp32) {
    __syncwarp();
    store_staged_rows<BN>(p, sbuf, lg, lane, m0, n0);
  }
}

// ===== epilogue of the persistent kernel: registers -> global, no shared memory =====
// During the persistent main loop shared memory is saturated (TMA writes + tensor-core reads),
// so this epilogue never touches it: each lane owns one accumulator row (tcgen05.ld 32x32b),
// applies bias / ReLU / mask and stores 64-byte (bf16) or 128-byte (fp32) row pieces directly.
// `release` is called once, as soon as the last TMEM read of the tile has landed in registers,
// so the MMA warp can start the tile after next in this accumulator buffer.
// `partial`: this lane's row of a split-K partner's fp32 accumulators (BN floats), or null.
template <int BN, typename Release>
__device__ __forceinline__ void epilogue_direct(const GemmParams& p, uint32_t tmem_acc, int warp,
                                                int lane, int m0, int n0, const float* partial,
                                                Release release) {
  const int lg = warp & 3;
  const int row = m0 + lg * 32 + lane;
  const bool row_ok = row < p.M;
  const bool vec16 = (p.ldc % 8) == 0 && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem_acc + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, r);
    tmem_ld_wait();
    if (c0 + 32 >= BN) release();
    const int col0 = n0 + c0;
    if (col0 >= p.N) continue;
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    if (partial) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 q = __ldcg(reinterpret_cast<const float4*>(partial + c0 + j));
        v[j] += q.x; v[j + 1] += q.y; v[j + 2] += q.z; v[j + 3] += q.w;
      }
    }
    const bool full_chunk = col0 + 32 <= p.N;
    if (p.bias) {
      if (full_chunk && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          const uint4 bv = *reinterpret_cast<const uint4*>(p.bias + col0 + j);
          const uint32_t w[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            v[j + 2 * q] += __uint_as_float(w[q] << 16);
            v[j + 2 * q + 1] += __uint_as_float(w[q] & 0xffff0000u);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) v[j] += __bfloat162float(p.bias[col0 + j]);
      }
    }
    if (p.relu) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
    }
    if (!row_ok) continue;
    if (p.relu_mask) {
      const __nv_bfloat16* mrow = p.relu_mask + (size_t)row * p.ld_mask + col0;
      if (full_chunk && (p.ld_mask % 8) == 0 && ((reinterpret_cast<uintptr_t>(p.relu_mask) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          const uint4 mv = *reinterpret_cast<const uint4*>(mrow + j);
          const uint32_t w[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (__uint_as_float(w[q] << 16) <= 0.0f) v[j + 2 * q] = 0.0f;
            if (__uint_as_float(w[q] & 0xffff0000u) <= 0.0f) v[j + 2 * q + 1] = 0.0f;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N && __bfloat162float(mrow[j]) <= 0.0f) v[j] = 0.0f;
      }
    }
    if (p.c_fp32) {
      float* crow = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
      if (full_chunk && (p.ldc % 4) == 0 && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(crow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) crow[j] = v[j];
      }
    } else {
      __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
      if (full_chunk && vec16) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          uint32_t w[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[j + 2 * q], v[j + 2 * q + 1]);
            w[q] = *reinterpret_cast<const uint32_t*>(&h);
          }
          *reinterpret_cast<uint4*>(crow + j) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) crow[j] = __float2bfloat16_rn(v[j]);
      }
    }
  }
}

// Split-K producer epilogue: raw fp32 accumulators of this lane's row -> workspace.
template <int BN, typename Release>
__device__ __forceinline__ void epilogue_store_partial(uint32_t tmem_acc, int warp, int lane,
                                                       float* dst_row, Release release) {
  const int lg = warp & 3;
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem_acc + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, r);
    tmem_ld_wait();
    if (c0 + 32 >= BN) release();
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      __stcg(reinterpret_cast<float4*>(dst_row + c0 + j),
             make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                         __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])));
  }
}

// Reduce-scatter epilogue of the persistent kernel (wgrad of a data-parallel layer). The
// stores cross NVLink, where 512-byte coalescing matters more than shared-memory bandwidth,
// so the bf16 tile rows are staged in a dedicated shared-memory area (the kernel then runs
// with a shorter operand ring) and pushed to the owner ranks row by row.
template <int BN, typename Release>
__device__ __forceinline__ void epilogue_rs_staged(const GemmParams& p, uint8_t* stage,
                                                   uint32_t tmem_acc, int warp, int lane, int m0,
                                                   int n0, Release release) {
  constexpr int PITCH = BN * 2 + 16;
  const int lg = warp & 3;
  uint8_t* sbuf = stage + lg * (32 * PITCH);
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem_acc + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, r);
    tmem_ld_wait();
    if (c0 + 32 >= BN) release();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(r[j + 2 * q]),
                                                       __uint_as_float(r[j + 2 * q + 1]));
        w[q] = *reinterpret_cast<const uint32_t*>(&h);
      }
      *reinterpret_cast<uint4*>(sbuf + lane * PITCH + (c0 + j) * 2) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  __syncwarp();
  store_staged_rows<BN>(p, sbuf, lg, lane, m0, n0);
  __syncwarp();                 // the next tile overwrites the staging rows
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a,
                         const __grid_constant__ CUtensorMap map_b, GemmParams p) {
  using C = Cfg<BN, B_MN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* tiles = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tmem_full = empty + C::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // With the reduce-scatter epilogue consecutive row tiles belong to different owner ranks
  // (owner = bucket vector / slice), so a row-major tile order would send all remote traffic
  // in one burst. Interleave: CTA row y handles tile (y % world) * tiles_per_owner + y / world,
  // which spreads NVLink stores evenly over the kernel's lifetime.
  int tile_y = blockIdx.y;
  if (p.rs_on) {
    const int W = p.rs_comm.world, T = gridDim.y;
    const int per = (T + W - 1) / W;
    const int cand = (tile_y % W) * per + tile_y / W;
    if (T % W == 0) tile_y = cand;
  }
  const int m0 = tile_y * BM, n0 = blockIdx.x * BN;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::TMEM_COLS);
  pdl_trigger();          // the next kernel in the stream may start its own prologue
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the tail
  // of the previous kernel; operands and epilogue inputs may only be read from here on.
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait_role(&empty[s], ph ^ 1);
        uint8_t* a_dst = tiles + s * C::STAGE_BYTES;
        uint8_t* b_dst = a_dst + C::A_BYTES;
        mbar_expect_tx(&full[s], C::STAGE_BYTES);
        const int k0 = kb * BK;
        if (A_MN) {   // global [Kred][M]: two boxes of 64 M-columns x 64 K-rows
          tma_load_2d(a_dst, &map_a, &full[s], m0, k0);
          tma_load_2d(a_dst + BK * 128, &map_a, &full[s], m0 + 64, k0);
        } else {      // global [M][K]: one box of 64 K-columns x 128 rows
          tma_load_2d(a_dst, &map_a, &full[s], k0, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < (BN + 63) / 64; ++j)
            tma_load_2d(b_dst + j * BK * 128, &map_b, &full[s], n0 + 64 * j, k0);
        } else {
          tma_load_2d(b_dst, &map_b, &full[s], k0, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait_role(&full[s], ph);
        tcgen05_fence_after();
        const uint32_t a_addr = smem_u32(tiles + s * C::STAGE_BYTES);
        const uint32_t b_addr = a_addr + C::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // K-major: step 16 elements = 32 B inside the swizzled 128 B row.
          // MN-major: step 16 K-rows = 2048 B.
          const uint64_t adesc = A_MN ? make_smem_desc(a_addr + k * UMMA_K * 128, BK * 128, 1024)
                                      : make_smem_desc(a_addr + k * UMMA_K * 2, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc(b_addr + k * UMMA_K * 128, BK * 128, 1024)
                                      : make_smem_desc(b_addr + k * UMMA_K * 2, 16, 1024);
          umma_bf16(tmem_base, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty[s]);          // frees the smem stage once these MMAs retire
      }
      umma_commit(tmem_full);            // accumulator complete
    }
  } else {
    epilogue_tile<BN>(p, tiles, tmem_full, tmem_base, warp, lane, m0, n0);
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ===========================================================================
// 2-CTA variant: a thread-block cluster of two CTAs (one TPC) computes a 256 x BN tile with
// tcgen05.mma.cta_group::2. Each CTA stages ITS 128 rows of A and ITS half (BN/2 columns) of
// B; the pair's tensor cores read both halves, so every byte of B is fetched from L2 once per
// 256 output rows instead of once per 128: the L2 -> SM operand traffic that caps the 1-CTA
// kernel at ~47 % tensor-pipe activity is cut by a quarter (BN=128) to a half (BN=256).
//   * both CTAs issue their TMA loads with .cta_group::2 and complete them on the LEADER's
//     `full` barrier (address obtained with mapa), which the leader's producer arms with the
//     byte count of both CTAs (the peer never touches the barrier itself)
//   * only the leader's elected thread issues the MMAs; tcgen05.commit ... multicast::cluster
//     releases the stage in both CTAs and finally raises tmem_full in both
//   * each CTA drains its own 128 TMEM lanes with the common epilogue
// ===========================================================================
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
// Arrive on a barrier of another CTA of the cluster (address from mapa). Default semantics
// (release at CTA scope): an explicit .release.cluster here compiles to MEMBAR.ALL.GPU + ERRBAR,
// which serialised the producer loop of the first 2-CTA kernel (ncu: 0.62 PFLOP/s).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* map,
                                                uint32_t leader_bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(map), "r"(leader_bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {   // same offset in both CTAs
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}

template <int BN, bool B_MN> struct Cfg2 {
  static constexpr int A_BYTES = BM * BK * 2;              // this CTA's 128 rows
  static constexpr int B_BYTES = (BN / 2) * BK * 2;        // this CTA's half of the columns
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN >= 256) ? 6 : 8;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256;
  static constexpr int TMEM_COLS = BN;
};

template <int BN, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap map_a,
                              const __grid_constant__ CUtensorMap map_b, GemmParams p) {
  using C = Cfg2<BN, B_MN>;
  static_assert(BN == 128 || BN == 256, "2-CTA tiles are 256x128 or 256x256");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* tiles = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tmem_full = empty + C::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                 // 0 = leader
  const int m0 = (blockIdx.x >> 1) * 256 + (int)rank * BM; // this CTA's rows
  const int n0 = blockIdx.y * BN;
  const int nb0 = n0 + (int)rank * (BN / 2);               // this CTA's half of B
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);     // the leader's producer arms it with the bytes of BOTH CTAs
      mbar_init(&empty[s], 1);    // one multicast commit from the leader's MMA thread
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2sm(tmem_slot, C::TMEM_COLS);
  pdl_trigger();
  tcgen05_fence_before();
  cluster_sync_all();            // both CTAs' barriers and TMEM are ready before anyone signals
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait_role(&empty[s], ph ^ 1);
        uint8_t* a_dst = tiles + s * C::STAGE_BYTES;
        uint8_t* b_dst = a_dst + C::A_BYTES;
        const uint32_t lead_full = mapa_shared(smem_u32(&full[s]), 0);
        if (rank == 0) mbar_expect_tx(&full[s], 2 * C::STAGE_BYTES);
        const int k0 = kb * BK;
        if (A_MN) {
          tma_load_2d_2sm(a_dst, &map_a, lead_full, m0, k0);
          tma_load_2d_2sm(a_dst + BK * 128, &map_a, lead_full, m0 + 64, k0);
        } else {
          tma_load_2d_2sm(a_dst, &map_a, lead_full, k0, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < BN / 2 / 64; ++j)
            tma_load_2d_2sm(b_dst + j * BK * 128, &map_b, lead_full, nb0 + 64 * j, k0);
        } else {
          tma_load_2d_2sm(b_dst, &map_b, lead_full, k0, nb0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(256, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait_role(&full[s], ph);
        tcgen05_fence_after();
        const uint32_t a_addr = smem_u32(tiles + s * C::STAGE_BYTES);
        const uint32_t b_addr = a_addr + C::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          const uint64_t adesc = A_MN ? make_smem_desc(a_addr + k * UMMA_K * 128, BK * 128, 1024)
                                      : make_smem_desc(a_addr + k * UMMA_K * 2, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc(b_addr + k * UMMA_K * 128, BK * 128, 1024)
                                      : make_smem_desc(b_addr + k * UMMA_K * 2, 16, 1024);
          umma_bf16_2sm(tmem_base, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit_2sm(&empty[s]);      // frees the stage in BOTH CTAs
      }
      umma_commit_2sm(tmem_full);        // accumulators complete in both CTAs
    }
  } else {
    epilogue_tile<BN>(p, tiles, tmem_full, tmem_base, warp, lane, m0, n0);
  }
  tcgen05_fence_before();
  cluster_sync_all();            // nobody frees TMEM or exits while the peer still depends on it
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc_2sm(tmem_base, C::TMEM_COLS);
  }
}

// ===========================================================================
// Persistent 2-CTA kernel: one cluster (CTA pair) per TPC loops over 256 x BN output tiles.
// TMEM holds TWO accumulator buffers (2 x BN columns), so the four epilogue warps of each CTA
// drain tile i (tcgen05.ld -> registers -> global) while the MMA thread is already running
// the main loop of tile i+1; the smem ring, its phases and the TMA producer simply keep
// running across tile boundaries. Three pipelines:
//   full/empty[STAGES]   TMA <-> MMA   (as in the one-tile 2-CTA kernel)
//   tmem_full[2]         MMA -> epilogue, multicast commit to both CTAs
//   tmem_empty[2]        epilogue -> MMA, on the LEADER: 4 warps x 2 CTAs arrive (the peer's
//                        warps arrive remotely through mapa)
// Tiles are assigned round-robin (tile = cluster + i * clusters, M fastest), so the ~74 tiles
// in flight share a few column blocks of B and all of A through L2.
// ===========================================================================
template <int BN, bool B_MN, bool RS> struct CfgP {
  static constexpr int A_BYTES = BM * BK * 2;              // this CTA's 128 rows
  static constexpr int B_BYTES = (BN / 2) * BK * 2;        // this CTA's half of the columns
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // 192 KB of operand ring; the reduce-scatter epilogue trades two stages for row staging
  static constexpr int STAGES = (BN >= 256) ? (RS ? 4 : 6) : (RS ? 6 : 8);
  static constexpr int EPI_BYTES = RS ? 4 * 32 * (BN * 2 + 16) : 0;
  static constexpr int SMEM = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
  static constexpr int TMEM_COLS = 2 * BN;                 // double-buffered accumulator
};

template <int BN, bool A_MN, bool B_MN, bool RS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_persist_kernel(const __grid_constant__ CUtensorMap map_a,
                                 const __grid_constant__ CUtensorMap map_b, GemmParams p) {
  using C = CfgP<BN, B_MN, RS>;
  static_assert(BN == 128 || BN == 256, "2-CTA tiles are 256x128 or 256x256");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* tiles = smem;
  uint8_t* epi_stage = smem + C::STAGES * C::STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(epi_stage + C::EPI_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tmem_full = empty + C::STAGES;       // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2], the leader's copy is the one in use
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                 // 0 = leader
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int tiles_m = (p.M + 255) / 256;
  const int num_tiles = tiles_m * ((p.N + BN - 1) / BN);
  const int num_kb = (p.K + BK - 1) / BK;
  // Work units. Without split-K a unit is a tile. With ksplit = 2, units [0, num_tiles) are
  // the PRODUCER halves (upper half of K) and [num_tiles, 2 num_tiles) the consumer halves of
  // the same tiles: every cluster walks its units in ascending order, so whatever a consumer
  // waits for is queued ahead of every consumer on its own cluster, and producers never wait.
  const int ksplit = RS ? 1 : p.ksplit;
  const int num_units = num_tiles * ksplit;
  const int kb_half = (num_kb + ksplit - 1) / ksplit;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);     // armed by the leader's producer with the bytes of both CTAs
      mbar_init(&empty[s], 1);    // one multicast commit from the leader's MMA thread
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 8);   // 4 epilogue warps in each of the two CTAs
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2sm(tmem_slot, C::TMEM_COLS);
  pdl_trigger();
  tcgen05_fence_before();
  cluster_sync_all();            // both CTAs' barriers and TMEM are ready before anyone signals
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ===== TMA producer (both CTAs): the ring keeps rolling across tiles =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = cluster_id; u < num_units; u += num_clusters) {
        const int t = u % num_tiles;
        const bool upper = ksplit == 2 && u < num_tiles;
        const int kb_begin = upper ? kb_half : 0;
        const int kb_end = (ksplit == 2 && !upper) ? kb_half : num_kb;
        const int m0 = (t % tiles_m) * 256 + (int)rank * BM;
        const int nb0 = (t / tiles_m) * BN + (int)rank * (BN / 2);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const uint32_t s = it % C::STAGES;
          const uint32_t ph = (it / C::STAGES) & 1;
          mbar_wait_role(&empty[s], ph ^ 1);
          uint8_t* a_dst = tiles + s * C::STAGE_BYTES;
          uint8_t* b_dst = a_dst + C::A_BYTES;
          const uint32_t lead_full = mapa_shared(smem_u32(&full[s]), 0);
          if (rank == 0) mbar_expect_tx(&full[s], 2 * C::STAGE_BYTES);
          const int k0 = kb * BK;
          if (A_MN) {
            tma_load_2d_2sm(a_dst, &map_a, lead_full, m0, k0);
            tma_load_2d_2sm(a_dst + BK * 128, &map_a, lead_full, m0 + 64, k0);
          } else {
            tma_load_2d_2sm(a_dst, &map_a, lead_full, k0, m0);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 2 / 64; ++j)
              tma_load_2d_2sm(b_dst + j * BK * 128, &map_b, lead_full, nb0 + 64 * j, k0);
          } else {
            tma_load_2d_2sm(b_dst, &map_b, lead_full, k0, nb0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(256, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      uint32_t it = 0, tcount = 0;
      for (int u = cluster_id; u < num_units; u += num_clusters, ++tcount) {
        const bool upper = ksplit == 2 && u < num_tiles;
        const int kb_begin = upper ? kb_half : 0;
        const int kb_end = (ksplit == 2 && !upper) ? kb_half : num_kb;
        const uint32_t buf = tcount & 1, use = tcount >> 1;
        mbar_wait_role(&tmem_empty[buf], (use & 1) ^ 1);    // both CTAs drained this buffer
        tcgen05_fence_after();
        const uint32_t acc = tmem_base + buf * BN;
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const uint32_t s = it % C::STAGES;
          const uint32_t ph = (it / C::STAGES) & 1;
          mbar_wait_role(&full[s], ph);
          tcgen05_fence_after();
          const uint32_t a_addr = smem_u32(tiles + s * C::STAGE_BYTES);
          const uint32_t b_addr = a_addr + C::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = A_MN ? make_smem_desc(a_addr + k * UMMA_K * 128, BK * 128, 1024)
                                        : make_smem_desc(a_addr + k * UMMA_K * 2, 16, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc(b_addr + k * UMMA_K * 128, BK * 128, 1024)
                                        : make_smem_desc(b_addr + k * UMMA_K * 2, 16, 1024);
            umma_bf16_2sm(acc, adesc, bdesc, idesc, (kb != kb_begin || k != 0) ? 1u : 0u);
          }
          umma_commit_2sm(&empty[s]);        // frees the stage in BOTH CTAs
        }
        umma_commit_2sm(&tmem_full[buf]);    // this accumulator is complete in both CTAs
      }
    }
  } else {
    // ===== epilogue warps (both CTAs) =====
    const uint32_t lead_empty0 = mapa_shared(smem_u32(&tmem_empty[0]), 0);
    uint32_t tcount = 0;
    for (int u = cluster_id; u < num_units; u += num_clusters, ++tcount) {
      const int t = u % num_tiles;
      const uint32_t buf = tcount & 1, use = tcount >> 1;
      const int m0 = (t % tiles_m) * 256 + (int)rank * BM;
      const int n0 = (t / tiles_m) * BN;
      mbar_wait(&tmem_full[buf], use & 1);
      tcgen05_fence_after();
      auto release = [&]() {
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(lead_empty0 + buf * 8);
      };
      if constexpr (RS) {
        epilogue_rs_staged<BN>(p, epi_stage, tmem_base + buf * BN, warp, lane, m0, n0, release);
      } else if (ksplit == 2) {
        // this lane's row of the tile's fp32 partial: [tile][256 rows][BN]
        float* prow = p.sk_ws + ((size_t)t * 256 + rank * BM + (warp & 3) * 32 + lane) * BN;
        unsigned* flag = p.sk_flags + t;
        if (u < num_tiles) {
          epilogue_store_partial<BN>(tmem_base + buf * BN, warp, lane, prow, release);
          __threadfence();
          __syncwarp();
          if (lane == 0) red_release_gpu_add_u32(flag, 1u);
        } else {
          if (lane == 0) {        // all 8 producer warps of the tile have published their rows
            uint32_t n = 0;
            while (ld_acquire_gpu_u32(flag) < 8u) {
              __nanosleep(64);
              if (++n > (1u << 25)) __trap();
            }
          }
          __syncwarp();
          epilogue_direct<BN>(p, tmem_base + buf * BN, warp, lane, m0, n0, prow, release);
          __syncwarp();
          if (lane == 0 && atomicAdd(flag, 1u) == 15u) atomicExch(flag, 0u);   // last reader resets
        }
      } else {
        epilogue_direct<BN>(p, tmem_base + buf * BN, warp, lane, m0, n0, nullptr, release);
      }
    }
  }
  tcgen05_fence_before();
  cluster_sync_all();            // nobody frees TMEM or exits while the peer still depends on it
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc_2sm(tmem_base, C::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------
// Host: TMA descriptors through the driver entry point (no link-time libcuda dependency,
// so the library still loads on a machine without a driver)
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
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

// 2-D bf16 tensor [rows][cols] (cols contiguous, `pitch` elements between rows);
// box = box_cols x box_rows, 128B swizzle, zero fill outside.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long pitch,
             int box_cols, int box_rows) {
  EncodeTiledFn enc = get_encode();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  B200_REQUIRE((pitch * 2) % 16 == 0, "gemm: row pitch must be a multiple of 16 bytes (%lld elements)", pitch);
  B200_REQUIRE(((uintptr_t)base & 15) == 0, "gemm: operand base must be 16-byte aligned");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code %d", (int)r);
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, cudaStream_t s) {
  using C = Cfg<BN, B_MN>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN>;
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    attr_done[dev & 63] = true;
  }
  dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM);
  B200_CUDA_OK(b200_launch_pdl(kern, grid, dim3(GEMM_THREADS), (size_t)C::SMEM, s,
                               b200_pdl_enabled() != 0, ma, mb, p));
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
int launch_2cta(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, cudaStream_t s) {
  using C = Cfg2<BN, B_MN>;
  auto kern = gemm_bf16_tcgen05_2cta_kernel<BN, A_MN, B_MN>;
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    attr_done[dev & 63] = true;
  }
  dim3 grid(2 * ((p.M + 255) / 256), (p.N + BN - 1) / BN);   // cluster pairs along x
  B200_CUDA_OK(b200_launch_pdl(kern, grid, dim3(GEMM_THREADS), (size_t)C::SMEM, s,
                               b200_pdl_enabled() != 0, ma, mb, p));
  return 0;
}

template <int BN>
int dispatch_major_2cta(int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                        const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_2cta<BN, false, false>(ma, mb, p, s);
  if (!a_mn && b_mn) return launch_2cta<BN, false, true>(ma, mb, p, s);
  if (a_mn && b_mn) return launch_2cta<BN, true, true>(ma, mb, p, s);
  return launch_2cta<BN, true, false>(ma, mb, p, s);
}

template <int BN, bool A_MN, bool B_MN, bool RS>
int launch_persist(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, cudaStream_t s) {
  using C = CfgP<BN, B_MN, RS>;
  auto kern = gemm_bf16_tcgen05_persist_kernel<BN, A_MN, B_MN, RS>;
  static bool attr_done[64] = {false};
  static int sms[64] = {0};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    B200_CUDA_OK(cudaDeviceGetAttribute(&sms[dev & 63], cudaDevAttrMultiProcessorCount, dev));
    attr_done[dev & 63] = true;
  }
  const int tiles = ((p.M + 255) / 256) * ((p.N + BN - 1) / BN) * (RS ? 1 : p.ksplit);   // work units
  int clusters = sms[dev & 63] / 2;                 // one CTA pair per TPC
  if (clusters > tiles) clusters = tiles;
  // same makespan with fewer pairs when the last wave would be ragged (less L2 contention)
  const int waves = (tiles + clusters - 1) / clusters;
  clusters = (tiles + waves - 1) / waves;
  B200_CUDA_OK(b200_launch_pdl(kern, dim3(2 * clusters), dim3(GEMM_THREADS), (size_t)C::SMEM, s,
                               b200_pdl_enabled() != 0, ma, mb, p));
  return 0;
}

template <int BN>
int dispatch_major_persist(int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                           const GemmParams& p, cudaStream_t s) {
  if (p.rs_on) {   // wgrad is the only product with the reduce-scatter epilogue
    if (a_mn && b_mn) return launch_persist<BN, true, true, true>(ma, mb, p, s);
    B200_REQUIRE(false, "gemm: the persistent reduce-scatter epilogue is built for MN-major operands only");
  }
  if (!a_mn && !b_mn) return launch_persist<BN, false, false, false>(ma, mb, p, s);
  if (!a_mn && b_mn) return launch_persist<BN, false, true, false>(ma, mb, p, s);
  if (a_mn && b_mn) return launch_persist<BN, true, true, false>(ma, mb, p, s);
  return launch_persist<BN, true, false, false>(ma, mb, p, s);
}

// Kernel choice. kind 0: one 128 x BN tile per CTA; 1: one 256 x BN tile per CTA pair;
// 2: persistent CTA pairs with double-buffered TMEM; 3 (forcing only): kind 2 with split-K.
// b200_gemm_set_variant (or B200DP_GEMM_VARIANT=kind[,bn]) pins the kind (and optionally BN)
// for tests and benchmarks; -1 = automatic.
struct Plan { int kind, bn, ksplit; };
int g_force_kind = -2, g_force_bn = 0;
constexpr int SK_MAX_TILES = 37;     // 2 x 37 work units fill the 74 CTA pairs in one wave

Plan choose(int M, int N, int K, bool rs_on, bool both_mn, bool allow_split) {
  if (g_force_kind == -2) {
    g_force_kind = -1;
    if (const char* e = getenv("B200DP_GEMM_VARIANT")) {
      int k = -1, bn = 0;
      if (sscanf(e, "%d,%d", &k, &bn) >= 1) { g_force_kind = k; g_force_bn = bn; }
    }
  }
  const int small_bn = (N <= 16) ? 16 : (N <= 32 ? 32 : (N <= 64 ? 64 : 128));
  // the reduce-scatter epilogue exists in the one-CTA kernel and, for wgrad's MN-major
  // operands, in the persistent kernel
  const bool persist_ok = N >= 128 && M > 128 && (!rs_on || both_mn);
  const bool pair_ok = persist_ok && !rs_on;
  const long long t256 = (long long)((M + 255) / 256) * ((N + 255) / 256);
  const int num_kb = (K + BK - 1) / BK;
  const bool split_ok = allow_split && pair_ok && N >= 256 && t256 <= SK_MAX_TILES && num_kb >= 2;
  if (g_force_kind >= 0) {
    if (g_force_kind == 3 && split_ok) return {2, 256, 2};
    if ((g_force_kind == 1 && pair_ok) || (g_force_kind >= 2 && persist_ok)) {
      int bn = g_force_bn ? g_force_bn : (N >= 256 ? 256 : 128);
      if (bn != 128 && bn != 256) bn = 256;
      if (N < bn) bn = 128;
      return {g_force_kind >= 2 ? 2 : 1, bn, 1};
    }
    int bn = g_force_bn ? g_force_bn : small_bn;
    if (bn == 256 && (N < 256 || rs_on)) bn = small_bn;
    if (bn != 16 && bn != 32 && bn != 64 && bn != 128 && bn != 256) bn = small_bn;
    if (bn > small_bn && bn != 256) bn = small_bn;
    return {0, bn, 1};
  }
  // automatic (thresholds from scripts/gemm_experiment.py on B200)
  // A problem that fits ONE wave of one-tile CTAs (100..148 of them) runs fastest that way: no
  // tile loop, no pair hand-shake (in-kernel times under ncu, profiles/r2/gemm_small_ncu.md:
  // 512x4096x4096 25.3 us against 26.3 for persistent pairs, 4096x784x512 12.7 against 14.3).
  {
    const long long c256 = (long long)((M + 127) / 128) * ((N + 255) / 256);
    const long long c128 = (long long)((M + 127) / 128) * ((N + 127) / 128);
    if (!rs_on && N >= 256 && K <= 1024 && c256 >= 100 && c256 <= 148) return {0, 256, 1};   // measured at K = 512 only
    if (N >= 128 && c128 >= 100 && c128 <= 148) return {0, 128, 1};
  }
  if (persist_ok) {
    if (N >= 256 && t256 >= 56) return {2, 256, 1};
    // Split-K (kind 3) is NOT chosen automatically: at 512 x 4096 x 4096 (32 tiles, 64 K
    // blocks) it measured 0.39 PFLOP/s against 0.53 for 256 x 128 tiles, because the serial
    // tail (producer epilogue -> flag -> consumer epilogue reading the partial tile from L2)
    // costs more than the halved main loop saves.
    const long long t128 = (long long)((M + 255) / 256) * ((N + 127) / 128);
    if (t128 >= 56) return {2, 128, 1};
  }
  if (!rs_on && N >= 256 && (long long)((M + 127) / 128) * ((N + 255) / 256) >= 120) return {0, 256, 1};
  return {0, small_bn, 1};
}

// Split-K workspaces: a small per-device pool (fp32 partial tiles + self-resetting flags),
// one slot per stream that uses split-K, allocated on the first use OUTSIDE stream capture.
// While it does not exist (or every slot is taken) the caller falls back to the unsplit plan,
// so nothing is ever allocated during a capture.
constexpr int SK_SLOTS = 4;
struct SkPool {
  float* ws[SK_SLOTS];
  unsigned* flags[SK_SLOTS];
  cudaStream_t owner[SK_SLOTS];
  int used;
  bool ready;
};
SkPool g_sk[64];
std::mutex g_sk_mu;

bool sk_acquire(cudaStream_t s, float** ws, unsigned** flags) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  std::lock_guard<std::mutex> lock(g_sk_mu);
  SkPool& P = g_sk[dev & 63];
  if (!P.ready) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) {
      cudaGetLastError();
      return false;
    }
    const size_t ws_bytes = (size_t)SK_MAX_TILES * 256 * 256 * sizeof(float);
    bool ok = true;
    for (int i = 0; i < SK_SLOTS && ok; ++i) {
      ok = cudaMalloc((void**)&P.ws[i], ws_bytes) == cudaSuccess &&
           cudaMalloc((void**)&P.flags[i], 256) == cudaSuccess &&
           cudaMemset(P.flags[i], 0, 256) == cudaSuccess;
    }
    ok = ok && cudaDeviceSynchronize() == cudaSuccess;
    if (!ok) {   // e.g. another thread is capturing in global mode: try again at a later call
      cudaGetLastError();
      for (int i = 0; i < SK_SLOTS; ++i) {
        if (P.ws[i]) cudaFree(P.ws[i]);
        if (P.flags[i]) cudaFree(P.flags[i]);
        P.ws[i] = nullptr;
        P.flags[i] = nullptr;
      }
      cudaGetLastError();
      return false;
    }
    P.ready = true;
  }
  int slot = -1;
  for (int i = 0; i < P.used; ++i)
    if (P.owner[i] == s) slot = i;
  if (slot < 0) {
    if (P.used == SK_SLOTS) return false;
    slot = P.used++;
    P.owner[slot] = s;
  }
  *ws = P.ws[slot];
  *flags = P.flags[slot];
  return true;
}

template <int BN>
int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                   const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<BN, false, false>(ma, mb, p, s);
  if (!a_mn && b_mn) return launch<BN, false, true>(ma, mb, p, s);
  if (a_mn && b_mn) return launch<BN, true, true>(ma, mb, p, s);
  return launch<BN, true, false>(ma, mb, p, s);
}

}  // namespace

extern "C" {

// C[M,N] = op(A) * op(B)^T with fp32 accumulation on tcgen05 tensor cores.
//   a_mn = 0: A is [M][K] (lda = K pitch);   a_mn = 1: A is [K][M] (lda = M pitch)
//   b_mn = 0: B is [N][K] (ldb = K pitch);   b_mn = 1: B is [K][N] (ldb = N pitch)
//   C is [M][ldc], bf16 (c_fp32 = 0) or fp32 (c_fp32 = 1)
//   bias: bf16 [N] or null; relu: clamp at 0; relu_mask: bf16 [M][ld_mask] or null, output is
//   zeroed where the mask is <= 0 (ReLU backward fused into dgrad).
// All pitches in elements; operand pitches must be multiples of 8 elements (16 bytes).
int b200_gemm_bf16_rs(const void* A, int a_mn, long long lda, const void* B, int b_mn,
                      long long ldb, void* C, int c_fp32, long long ldc, int M, int N, int K,
                      const void* bias, int relu, const void* relu_mask, long long ld_mask,
                      const void* rs_comm, size_t rs_stage_off, size_t rs_src_bytes,
                      size_t rs_slice_vec, size_t rs_base_vec, const unsigned* rs_seq,
                      void* stream);

// Which kernel a problem would get (host-only, no GPU needed): kind 0 / 1 / 2 and the tile width.
int b200_gemm_plan(int M, int N, int K, int rs_on, int both_mn, int* kind, int* bn, int* ksplit) {
  const Plan p = choose(M, N, K, rs_on != 0, both_mn != 0, true);
  if (kind) *kind = p.kind;
  if (bn) *bn = p.bn;
  if (ksplit) *ksplit = p.ksplit;
  return 0;
}

// Pin the kernel kind / tile width (see choose()); kind = -1 restores the automatic choice.
int b200_gemm_set_variant(int kind, int bn) {
  g_force_kind = kind < 0 ? -1 : kind;
  g_force_bn = kind < 0 ? 0 : bn;
  return 0;
}

int b200_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb,
                   void* C, int c_fp32, long long ldc, int M, int N, int K, const void* bias,
                   int relu, const void* relu_mask, long long ld_mask, void* stream) {
  return b200_gemm_bf16_rs(A, a_mn, lda, B, b_mn, ldb, C, c_fp32, ldc, M, N, K, bias, relu,
                           relu_mask, ld_mask, nullptr, 0, 0, 0, 0, nullptr, stream);
}

// Same product with the reduce-scatter epilogue (see GemmParams::rs_*): C is not written;
// every 16-byte vector of the result lands in the owner rank's staging slot of the bucket's
// two-shot all-reduce channel. Follow it with b200_comm_allreduce(..., push_from = vectors
// produced here) on the same channel.
int b200_gemm_bf16_rs(const void* A, int a_mn, long long lda, const void* B, int b_mn,
                      long long ldb, void* C, int c_fp32, long long ldc, int M, int N, int K,
                      const void* bias, int relu, const void* relu_mask, long long ld_mask,
                      const void* rs_comm, size_t rs_stage_off, size_t rs_src_bytes,
                      size_t rs_slice_vec, size_t rs_base_vec, const unsigned* rs_seq,
                      void* stream) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "b200_gemm_bf16: empty problem");
  cudaStream_t s = (cudaStream_t)stream;
  Plan plan = choose(M, N, K, rs_comm != nullptr, a_mn && b_mn, true);
  float* sk_ws = nullptr;
  unsigned* sk_flags = nullptr;
  if (plan.ksplit == 2 && !sk_acquire(s, &sk_ws, &sk_flags))
    plan = choose(M, N, K, rs_comm != nullptr, a_mn && b_mn, false);
  const int BN = plan.bn;
  const int bn2 = plan.kind ? plan.bn : 0;      // CTA-pair kernels: each CTA loads half of the B tile
  CUtensorMap ma, mb;
  int e;
  if (a_mn) e = make_map(&ma, A, K, M, lda, 64, 64); else e = make_map(&ma, A, M, K, lda, 64, BM);
  if (e) return e;
  // with 2-CTA tiles each CTA loads half of the B tile: K-major boxes are BN/2 rows tall
  if (b_mn) e = make_map(&mb, B, K, N, ldb, 64, 64);
  else e = make_map(&mb, B, N, K, ldb, 64, bn2 ? BN / 2 : BN);
  if (e) return e;
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = (int)ldc; p.c_fp32 = c_fp32;
  p.bias = (const __nv_bfloat16*)bias;
  p.relu = relu;
  p.relu_mask = (const __nv_bfloat16*)relu_mask;
  p.ld_mask = (int)ld_mask;
  p.ksplit = plan.ksplit;
  p.sk_ws = sk_ws;
  p.sk_flags = sk_flags;
  p.rs_on = 0;
  if (rs_comm) {
    B200_REQUIRE(!c_fp32 && ldc == N && N % 8 == 0 && !bias && !relu && !relu_mask,
                 "b200_gemm_bf16: the reduce-scatter epilogue needs a dense bf16 [M,N] output, N %% 8 == 0");
    p.rs_on = 1;
    memcpy(&p.rs_comm, rs_comm, sizeof(CommDev));
    p.rs_stage_off = rs_stage_off;
    p.rs_src_vec = rs_src_bytes / 16;
    p.rs_sl = rs_slice_vec;
    p.rs_base_vec = rs_base_vec;
    p.rs_seq = rs_seq;
  }
  if (plan.kind == 2)
    return BN == 256 ? dispatch_major_persist<256>(a_mn, b_mn, ma, mb, p, s)
                     : dispatch_major_persist<128>(a_mn, b_mn, ma, mb, p, s);
  if (plan.kind == 1)
    return BN == 256 ? dispatch_major_2cta<256>(a_mn, b_mn, ma, mb, p, s)
                     : dispatch_major_2cta<128>(a_mn, b_mn, ma, mb, p, s);
  switch (BN) {
    case 16: return dispatch_major<16>(a_mn, b_mn, ma, mb, p, s);
    case 32: return dispatch_major<32>(a_mn, b_mn, ma, mb, p, s);
    case 64: return dispatch_major<64>(a_mn, b_mn, ma, mb, p, s);
    case 256: return dispatch_major<256>(a_mn, b_mn, ma, mb, p, s);
    default: return dispatch_major<128>(a_mn, b_mn, ma, mb, p, s);
  }
}

}  // extern "C"
