// Persistent training kernel for the MNIST `Net`: many optimizer steps per launch, no kernel
// boundary and no host involvement between steps (reference loop: src/train_dist.py:78-89,
// src/train.py:71-76; DDP's overlapped all-reduce: src/train_dist.py:63,83).
//
// One cooperative launch of <= 148 CTAs (one per SM) that stay resident for `nsteps`
// optimizer steps. The grid is specialised at CTA granularity:
//
//  * WORKER CTAs run forward + loss + backward of their images out of shared memory. With
//    2*B <= workers every image gets a PAIR of CTAs: both run the forward; in the backward
//    each takes half of conv2's output channels (weight gradient), half of conv2's input
//    channels (data gradient) and half of conv1's channels, so the pair never communicates.
//    Otherwise a worker loops over its images and accumulates conv gradients in shared memory.
//    Workers do NOT form the fc weight gradients: they publish per-image activations
//    (a2, dpre, h1, dlogits: 1.7 KB) as soon as fc2's backward has run, about half way
//    through the step.
//  * SERVICE CTAs (the SMs a batch of 64 leaves idle) turn those into updated weights while
//    the workers are still in conv backward:
//      - fc1.weight gradient as a batch GEMM  dW3 = dpre^T a2  over a column slice,
//        fc1.bias / fc2 / loss in one more CTA;
//      - fixed-order sums of the conv2 / conv1 partial rows, each as soon as its producer
//        count is reached (conv2's weight gradient is computed first in the last phase, so its
//        reduction hides behind conv2 dgrad + conv1 wgrad);
//      - multi-GPU: the slice is pushed to every peer as tagged words ({x, seq, y, seq}: the
//        payload validates itself, one NVLink traversal, no fence, no flag) and summed in rank
//        order, so replicas stay bit-identical; 1/world; SGD-momentum; new weights to global
//        memory; a release-increment of the group's "done" counter.
//    Three parameter groups (conv1 | conv2 | fc) are published separately. A worker starts
//    step s+1 as soon as conv1's group is there and picks up conv2 / fc when it reaches
//    them: the fc bucket (76 % of the bytes) and the conv2 bucket are exchanged while
//    backward still runs; only conv1's 1 KB sits on the step-to-step critical path.
//  * Weights and images move global -> shared with cp.async.bulk (UBLKCP) completing on
//    mbarriers; the next image is fetched one item ahead.
//
// Synchronisation is by monotonically increasing counters in global memory
// (red.release.gpu / ld.acquire.gpu); there is no grid-wide barrier. Every wait is bounded:
// on timeout the kernel sets an abort word, reports {where, peer} and drains.
//
// Host-fed mode: inputs arrive in a ring of device slots filled by cudaMemcpyAsync on a
// copy stream; each step's readiness is a 32-bit value written by a stream memory operation
// after that step's copies, which the workers poll; the kernel publishes how many steps it
// has consumed so the copy stream can reuse slots. The step's loss is stored by the kernel
// straight into mapped pinned host memory (4 bytes device -> host per step).
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <ctime>

#include <atomic>
#include <cstring>
#include <new>

#include "comm.cuh"
#include "net_device.cuh"

using namespace net;

namespace {

// ---- layout of the per-step exchange buffers ------------------------------------------
// every segment starts on a float4 so the service CTAs can fetch it with 16-byte loads
constexpr int ACT_STRIDE = 448;  // floats per image: a2[320] dpre[50+2] h1[50+2] dlog[10] loss[1]
constexpr int ACT_A2 = 0, ACT_DPRE = 320, ACT_H1 = 372, ACT_DLOG = 424, ACT_LOSS = 434;
constexpr int ACT_SMALL_F4 = 29; // float4s from ACT_DPRE to the loss: what the small-fc CTA stages
constexpr int P2_ROW = 5020;     // conv2 partial row: W2[5000] b2[20] (flat layout from OFF_W2)
// conv1's 260 gradients / weights travel as self-validating {value, tag} pairs (8 bytes each),
// ordered so that each half (the channels one CTA of an image pair owns, one service CTA
// sums) is a contiguous run of 130 pairs = 65 16-byte words:
//   pair h*130 + i  ->  i < 125 ? W1[h*125 + i] : b1[h*5 + (i - 125)]
constexpr int C1_HALF = 130, C1_PAIRS = 260;
__host__ __device__ constexpr int c1_flat(int half, int i) {
  return i < 125 ? half * 125 + i : OFF_B1 + half * 5 + (i - 125);
}
constexpr int FC_SMALL_LO = OFF_B3;             // b3 | W4 | b4 (+pads): flat [21284, 21848)
constexpr int FC_SMALL_N = FLAT_SZ - OFF_B3;    // 564
static_assert(OFF_W2 % 4 == 0 && P2_ROW % 4 == 0 && OFF_W3 % 4 == 0 && FC_SMALL_LO % 4 == 0 &&
                  FC_SMALL_N % 4 == 0, "groups must be float4 aligned");

// control words, one per 128-byte line
enum { C_ACT = 0, C_P2, C_DFC, C_ABORT, C_STARTED, C_NWORDS };
constexpr int CTL_STRIDE = 32;   // u32 words between counters

struct PersistArgs {
  float* params;
  float* mom;
  float* grad_out;           // optional: the averaged gradient of the LAST step
  const float* hp;           // {lr, momentum, ...}
  const void* x;             // dataset (gather modes) or ring / batch storage (direct modes)
  const long long* labels;
  const int* gidx;
  int x_mode;                // 0 fp32 rows, 1 uint8 rows, 2 uint8 dataset + gidx, 3 fp32 dataset + gidx
  int B;
  int nsteps;
  int update;                // 0: leave params/momentum alone (gradient-only run)
  int n_workers, split, rows, ipc;   // ipc = images per worker per step
  int n_service, n3, kw;             // service CTAs, of which W3 CTAs; W3 columns per CTA
  int idx_stride, n_idx;
  int ring_slots;            // direct modes: slot of local step ls is (ring_pos + ls) % ring_slots
  unsigned ring_pos;
  const unsigned* feed_ready;   // null, or: global step g = feed_base + ls is in its slot once
                                // feed_ready[32 * (g & 1)] > g (two copy streams, one word each)
  unsigned* feed_consumed;
  unsigned feed_base;
  int label_slots;           // labels of local step ls: labels + ((label_pos + ls) % label_slots) * B
  unsigned label_pos;
  int epoch_begin;           // 1: this launch starts an epoch (step_ctr[1] := step_ctr[0])
  float* acts;
  float* part2;
  uint2* part1;              // [rows][260] {value, tag}: conv1 partial rows
  uint2* w1box;              // [260] {value, tag}: conv1's updated weights, service -> workers
  uint2* w2box;              // [5020] {value, tag}: conv2's updated weights, worker slices -> all workers
  unsigned tag_base;         // (launch id << 16): tag of local step ls is tag_base | (ls + 1)
  unsigned* ctl;
  float* loss_ring;
  int ring;
  float* host_loss;          // mapped pinned host memory or null; loss of local step ls -> [ls]
  unsigned long long* step_ctr;   // [0] optimizer steps so far, [1] step at which the epoch began
  unsigned long long seed;
  unsigned sample_base;
  int mask_mode;             // 0 none, 1 Philox, 2 injected (m2buf/m3buf, single step)
  float* m2buf;
  float* m3buf;
  int multi;
  size_t stage_off;
  unsigned* seq_ctr;
  unsigned long long timeout_ns;
  unsigned* err;             // {code, peer, where, seq}
  long long* dbg;            // optional phase stamps [grid][DBG_STEPS][DBG_PH][2] (clock64, globaltimer)
  CommDev comm;
};
constexpr int DBG_STEPS = 32, DBG_PH = 16;
#define PSTAMP(ls_, i_)                                                                   \
  do {                                                                                    \
    if (a.dbg && threadIdx.x == 0 && (ls_) < DBG_STEPS) {                                 \
      long long* d_ = a.dbg + (((size_t)blockIdx.x * DBG_STEPS + (ls_)) * DBG_PH + (i_)) * 2; \
      d_[0] = clock64();                                                                  \
      d_[1] = (long long)globaltimer_ns();                                                \
    }                                                                                     \
  } while (0)

// ---- small PTX wrappers ---------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0xf4240;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded: a copy that never lands (bad pointer would fault first) must not hang the GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  for (unsigned i = 0; i < 20000u; ++i)   // each try suspends for at most 1 ms
    if (mbar_try_wait(bar, parity)) return;
  asm volatile("trap;");
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar` (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ldcg4(const float* p) {
  return __ldcg(reinterpret_cast<const float4*>(p));
}

__device__ __forceinline__ void set_abort(const PersistArgs& a, unsigned where, unsigned peer,
                                          unsigned seq) {
  if (atomicCAS(a.err, 0u, 2u) == 0u) {
    a.err[1] = peer;
    a.err[2] = where;
    a.err[3] = seq;
  }
  atomicExch(a.ctl + C_ABORT * CTL_STRIDE, 1u);
}

// Wait until *p >= target (wrap-safe). One thread calls it. Polls hit L2 (~0.3 us each), so
// the abort word and the clock are looked at every 64 polls.
__device__ __forceinline__ bool wait_ge(const unsigned* p, unsigned target, const PersistArgs& a,
                                        unsigned where, bool sys_scope = false) {
  unsigned v = sys_scope ? ld_acquire_sys_u32(p) : ld_acquire_gpu_u32(p);
  if ((int)(v - target) >= 0) return true;
  const unsigned long long t0 = globaltimer_ns();
  for (;;) {
#pragma unroll 1
    for (int i = 0; i < 64; ++i) {
      v = sys_scope ? ld_acquire_sys_u32(p) : ld_acquire_gpu_u32(p);
      if ((int)(v - target) >= 0) return true;
    }
    if (ld_volatile_u32(a.ctl + C_ABORT * CTL_STRIDE) != 0u) return false;
    // local waits get twice the peer timeout: when a peer is missing, the wait that names it
    // (the gradient exchange) must be the one that fires first
    if (globaltimer_ns() - t0 > 2ull * a.timeout_ns) {
      set_abort(a, where, 0xffffu, target);
      return false;
    }
  }
}

__device__ __forceinline__ void signal_add(unsigned* ctr) {
  // callers: every thread's stores, __threadfence(), __syncthreads(), then ONE thread here
  red_release_gpu_add_u32(ctr, 1u);
}

// ---- worker stages that differ from net_device.cuh ---------------------------------------

// fc1 forward with every operand in shared memory (net::fc1_fwd_img reads the bias with __ldg).
__device__ __forceinline__ void fc1_fwd_s(const float* __restrict__ a2s, const float* __restrict__ W3,
                                          const float* __restrict__ b3, const float* __restrict__ m3s,
                                          float* __restrict__ h1s) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float av[A2_SZ / 32];
#pragma unroll
  for (int i = 0; i < A2_SZ / 32; ++i) av[i] = a2s[lane + 32 * i];
#pragma unroll
  for (int jj = 0; jj < FC1 / (NT / 32); ++jj) {
    const int j = warp + jj * (NT / 32);
    const float* w = W3 + j * A2_SZ;
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < A2_SZ / 32; ++i) s = fmaf(av[i], w[lane + 32 * i], s);
    s = warp_sum(s);
    if (lane == 0) h1s[j] = fmaxf(s + b3[j], 0.0f) * m3s[j];
  }
}

// conv2 weight / bias gradient for output channels [co0, co0 + nco). ACC: add into a
// shared-memory accumulator (several images per CTA) instead of overwriting a global row.
template <bool ACC>
__device__ __forceinline__ void conv2_wgrad_part(const float* __restrict__ g2s,
                                                 const uint8_t* __restrict__ pos2s,
                                                 const float* __restrict__ m2s,
                                                 const float* __restrict__ a1s, float* __restrict__ gW2,
                                                 float* __restrict__ gB2, int co0, int nco,
                                                 bool first) {
  const int t = threadIdx.x;
  if (t < 250) {
    const int ci = t / 25, kk = t - ci * 25;
    const int ky = kk / 5, kx = kk - ky * 5;
    const float* a = a1s + ci * 144 + ky * 12 + kx;
#pragma unroll 1
    for (int co = co0; co < co0 + nco; ++co) {
      float s = 0.0f;
      if (m2s[co] != 0.0f) {   // uniform: Dropout2d removed the whole channel otherwise
        const float4* g4 = reinterpret_cast<const float4*>(g2s + co * 16);
        const uint4 pp = *reinterpret_cast<const uint4*>(pos2s + co * 16);
        const unsigned pw[4] = {pp.x, pp.y, pp.z, pp.w};
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 g = g4[q];
          s0 = fmaf(g.x, a[pw[q] & 0xffu], s0);
          s1 = fmaf(g.y, a[(pw[q] >> 8) & 0xffu], s1);
          s2 = fmaf(g.z, a[(pw[q] >> 16) & 0xffu], s2);
          s3 = fmaf(g.w, a[pw[q] >> 24], s3);
        }
        s = (s0 + s1) + (s2 + s3);
      } else if (ACC && !first) {
        continue;
      }
      if (ACC) gacc(gW2 + co * 250 + t, s, first);
      else gW2[co * 250 + t] = s;
    }
  } else if (t >= 256 && t < 256 + nco) {
    const int co = co0 + (t - 256);
    float s = 0.0f;
#pragma unroll
    for (int p = 0; p < 16; ++p) s += g2s[co * 16 + p];
    if (ACC) gacc(gB2 + co, s, first);
    else gB2[co] = s;
  }
}

// dgrad product T[pos][r] = sum_co D[pos][co] * W2[co][r] for the columns r of input channels
// [ci0, ci0 + nci). With half of the columns (125) the 64 positions are split over two thread
// groups so 250 threads stay busy.
__device__ __forceinline__ void conv2_dgrad_T_part(const float* __restrict__ Ds,
                                                   const int* __restrict__ live, int nlive,
                                                   const float* __restrict__ w2s, float* __restrict__ T,
                                                   int ci0, int nci) {
  const int t = threadIdx.x;
  const int ncol = nci * 25;
  const int groups = (ncol <= 125) ? 2 : 1;
  if (t >= ncol * groups) return;
  const int pg = t / ncol;
  const int r = ci0 * 25 + (t - pg * ncol);
  const int p0 = pg * (64 / groups), p1 = p0 + 64 / groups;
  float w[C2];
#pragma unroll
  for (int sl = 0; sl < C2; ++sl) w[sl] = (sl < nlive) ? w2s[live[sl] * 250 + r] : 0.0f;
  const int nq = (nlive + 3) >> 2;
#pragma unroll 2
  for (int pos = p0; pos < p1; ++pos) {
    const float4* d4 = reinterpret_cast<const float4*>(Ds + pos * C2);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int q = 0; q < C2 / 4; ++q) {
      if (q < nq) {
        const float4 d = d4[q];
        s0 = fmaf(d.x, w[4 * q], s0);
        s1 = fmaf(d.y, w[4 * q + 1], s1);
        s0 = fmaf(d.z, w[4 * q + 2], s0);
        s1 = fmaf(d.w, w[4 * q + 3], s1);
      }
    }
    T[pos * 250 + r] = s0 + s1;
  }
}

// 25-tap gather of T into the gradient of conv1's pooled map, channels [ci0, ci0 + nci).
__device__ __forceinline__ void conv2_dgrad_gather_part(const float* __restrict__ T,
                                                        const float* __restrict__ a1s,
                                                        float* __restrict__ g1s, int ci0, int nci) {
  for (int q = ci0 * 144 + threadIdx.x; q < (ci0 + nci) * 144; q += NT) {
    const int ci = q / 144, r = q - ci * 144;
    const int u = r / 12, v = r - u * 12;
    float s = 0.0f;
    if (a1s[q] > 0.0f) {
      const float* tb = T + (u * 8 + v) * 250 + ci * 25;
#pragma unroll
      for (int ky = 0; ky < 5; ++ky) {
        const bool oky = (unsigned)(u - ky) < 8u;
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
          const bool ok = oky && ((unsigned)(v - kx) < 8u);
          if (ok) s += tb[-ky * 1995 - kx * 249];
        }
      }
    }
    g1s[q] = s;
  }
}

__device__ __forceinline__ void st_pair(uint2* p, float v, unsigned tag) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(tag)
               : "memory");
}

// conv1 weight / bias gradient for the 5 channels of half `h`: 125 weight outputs, each split
// over two adjacent lanes (72 windows each) and combined with one shuffle, + 5 bias sums. The
// results leave as {value, tag} pairs: the service CTA polls the pairs themselves, so there is
// no fence, no flag and no second round trip on this last link of the step.
__device__ __forceinline__ void conv1_wgrad_half(const float* __restrict__ g1s,
                                                 const uint16_t* __restrict__ pos1s,
                                                 const float* __restrict__ xs, uint2* __restrict__ row,
                                                 int h, unsigned tag) {
  const int t = threadIdx.x;
  const int c0 = h * 5;
  if (t < 256) {   // whole warps 0..7: the shuffle below needs every lane
    const int o = t >> 1, half = t & 1;
    float s = 0.0f;
    if (o < 125) {
      const int cl = o / 25, kk = o - cl * 25;
      const int c = c0 + cl;
      const int ky = kk / 5, kx = kk - ky * 5;
      const float* x = xs + ky * IMG + kx;
      const float4* g4 = reinterpret_cast<const float4*>(g1s + c * 144) + half * 18;
      const uint2* p2 = reinterpret_cast<const uint2*>(pos1s + c * 144) + half * 18;
      float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
#pragma unroll 6
      for (int e = 0; e < 18; ++e) {
        const float4 gv = g4[e];
        const uint2 pp = p2[e];
        s0 = fmaf(gv.x, x[pp.x & 0xffffu], s0);
        s1 = fmaf(gv.y, x[pp.x >> 16], s1);
        s2 = fmaf(gv.z, x[pp.y & 0xffffu], s2);
        s3 = fmaf(gv.w, x[pp.y >> 16], s3);
      }
      s = (s0 + s1) + (s2 + s3);
    }
    const float other = __shfl_xor_sync(0xffffffffu, s, 1);
    if (o < 125 && half == 0) st_pair(row + h * C1_HALF + o, s + other, tag);   // lower half first
  } else if (t < 256 + 5) {
    const int c = c0 + (t - 256);
    const float4* g4 = reinterpret_cast<const float4*>(g1s + c * 144);
    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll 6
    for (int e = 0; e < 36; ++e) {
      const float4 g = g4[e];
      s0 += g.x + g.z;
      s1 += g.y + g.w;
    }
    st_pair(row + h * C1_HALF + 125 + (t - 256), s0 + s1, tag);
  }
}

// ---- shared-memory carve of a worker -----------------------------------------------------
struct WorkerSmem {
  float* P;        // mirror of the flat parameter buffer
  float* xs;       // [2][784]
  float* a1s; float* g1s; float* a2s; float* g2s;
  float* h1s; float* m3s; float* dpre; float* m2s; float* logit; float* dlog;
  float* Ds; int* live; int* slot_of;
  float* Ts;
  float* acc;      // conv partial accumulators (flat layout [0, OFF_W3)), several images per CTA
  uint16_t* pos1s; uint8_t* pos2s; uint8_t* xu;   // xu: [2][784] uint8 staging
  unsigned long long* bar;   // [0..2] weight groups, [3..4] image buffers
  int* flags;      // [0] ok word of the last wait
};
constexpr int WS_FLOATS = FLAT_SZ + 2 * IMG_SZ + 2 * A1_SZ + 2 * A2_SZ + 64 * 3 + 32 + 16 + 16 +
                          64 * C2 + 64 + 64 * 250 + OFF_W3;
constexpr int WS_BYTES = WS_FLOATS * 4 + A1_SZ * 2 + A2_SZ + 2 * 784 + 8 * 8 + 64;
static_assert(WS_BYTES <= 227 * 1024, "worker shared memory exceeds one SM");

__device__ __forceinline__ WorkerSmem carve_worker(float* s) {
  WorkerSmem w;
  w.P = s; s += FLAT_SZ;
  w.xs = s; s += 2 * IMG_SZ;
  w.a1s = s; s += A1_SZ;
  w.g1s = s; s += A1_SZ;
  w.a2s = s; s += A2_SZ;
  w.g2s = s; s += A2_SZ;
  w.h1s = s; s += 64;
  w.m3s = s; s += 64;
  w.dpre = s; s += 64;
  w.m2s = s; s += 32;
  w.logit = s; s += 16;
  w.dlog = s; s += 16;
  w.Ds = s; s += 64 * C2;
  w.live = reinterpret_cast<int*>(s); s += 32;
  w.slot_of = reinterpret_cast<int*>(s); s += 32;
  w.Ts = s; s += 64 * 250;
  w.acc = s; s += OFF_W3;
  w.pos1s = reinterpret_cast<uint16_t*>(s);
  uint8_t* b = reinterpret_cast<uint8_t*>(s) + A1_SZ * 2;
  w.pos2s = b; b += A2_SZ;
  w.xu = b; b += 2 * 784;
  w.bar = reinterpret_cast<unsigned long long*>(b); b += 8 * 8;
  w.flags = reinterpret_cast<int*>(b);
  return w;
}
static_assert((WS_FLOATS * 4 + A1_SZ * 2 + A2_SZ) % 16 == 0, "uint8 staging must be 16-byte aligned");
static_assert((WS_FLOATS * 4 + A1_SZ * 2 + A2_SZ + 2 * 784) % 8 == 0, "mbarrier alignment");

__device__ __forceinline__ float normalize_px(unsigned v) {
  return ((float)v / 255.0f - 0.1307f) / 0.3081f;   // ToTensor + Normalize, src/train_dist.py:16-19
}

// =========================================================================================
// Helpers shared by workers and service CTAs: tagged exchange, update, row sums, staging
// =========================================================================================
struct ServiceSmem {
  float4* gs;       // gradient slice in float4 units (<= 1024 items)
  float4* red;      // cross-row-group reduction scratch [NT]
  float* stage;     // activations staging: up to 64 images x 112 floats
  int* flags;
};
constexpr int SV_ITEMS_MAX = 1024;
constexpr int SV_STAGE_FLOATS = 64 * 116;   // max(W3: 64 x (52 + 60), small fc: 64 x 116)
constexpr int SV_BYTES = SV_ITEMS_MAX * 16 + NT * 16 + SV_STAGE_FLOATS * 4 + 64;

// Tagged-word exchange of one float4 with every peer + rank-ordered sum (bit-identical on all
// ranks). `mine` is this rank's value; returns the sum over ranks.
__device__ __forceinline__ void ll_push(const PersistArgs& a, int col4, const float4& v, unsigned seq) {
  constexpr int N4 = FLAT_SZ / 4;
  const unsigned slot = seq & 1u;
  const uint4 w0 = make_uint4(__float_as_uint(v.x), seq, __float_as_uint(v.y), seq);
  const uint4 w1 = make_uint4(__float_as_uint(v.z), seq, __float_as_uint(v.w), seq);
  for (int pp = 1; pp < a.comm.world; ++pp) {
    const int p = (a.comm.rank + pp) % a.comm.world;   // stagger destinations
    uint4* dst = peer_ptr<uint4>(a.comm, p, a.stage_off) +
                 (((size_t)slot * a.comm.world + a.comm.rank) * N4 + col4) * 2;
    st_volatile_u4(dst, w0);
    st_volatile_u4(dst + 1, w1);
  }
}

__device__ __forceinline__ bool ll_gather_sum(const PersistArgs& a, int col4, const float4& mine,
                                              unsigned seq, float4& out) {
  constexpr int N4 = FLAT_SZ / 4;
  const unsigned slot = seq & 1u;
  const int W = a.comm.world, me = a.comm.rank;
  const uint4* base = peer_ptr<uint4>(a.comm, me, a.stage_off) + ((size_t)slot * W * N4 + col4) * 2;
  unsigned long long t0 = 0;
  unsigned polls = 0;
  for (;;) {
    // Every source's two words in one go, UNCONDITIONALLY (ranks beyond the world and this rank
    // itself re-read a valid slot and are ignored): 16 independent loads per round, no local
    // arrays indexed under a predicate, nothing for the compiler to serialise.
    uint4 wa[B200_MAX_WORLD], wb[B200_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < B200_MAX_WORLD; ++r) {
      const int rr = (r < W) ? r : me;
      wa[r] = ld_volatile_u4(base + (size_t)rr * N4 * 2);
      wb[r] = ld_volatile_u4(base + (size_t)rr * N4 * 2 + 1);
    }
    unsigned missing = 0;
#pragma unroll
    for (int r = 0; r < B200_MAX_WORLD; ++r)
      if (r < W && r != me &&
          !(wa[r].y == seq && wa[r].w == seq && wb[r].y == seq && wb[r].w == seq))
        missing |= 1u << r;
    if (!missing) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < B200_MAX_WORLD; ++r)      // rank order: bit-identical on every rank
        if (r < W) {
          const float4 v = (r == me) ? mine
                                     : make_float4(__uint_as_float(wa[r].x), __uint_as_float(wa[r].z),
                                                   __uint_as_float(wb[r].x), __uint_as_float(wb[r].z));
          s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
      out = s;
      return true;
    }
    if ((++polls & 127u) == 0u) {
      if (t0 == 0) t0 = globaltimer_ns();
      bool give_up = ld_volatile_u32(a.ctl + C_ABORT * CTL_STRIDE) != 0u;
      if (!give_up && globaltimer_ns() - t0 > a.timeout_ns) {
        const int peer = __ffs((int)missing) - 1;
        if (atomicCAS(a.comm.err, 0u, 1u) == 0u) {
          a.comm.err[1] = (unsigned)peer; a.comm.err[2] = 12u; a.comm.err[3] = seq;
        }
        atomicExch(a.ctl + C_ABORT * CTL_STRIDE, 1u);
        give_up = true;
      }
      if (give_up) {
        out = make_float4(mine.x * W, mine.y * W, mine.z * W, mine.w * W);   // any finite value: the run is aborted
        return false;
      }
    }
  }
}

// gs[0..nitems) holds this CTA's local gradient slice; item i is flat float4 column col_of(i).
// Exchange (multi-GPU), scale, SGD-momentum, store. All threads call it.
// box != null: the new weights also leave as {value, tag} pairs at box[4 * (col - box_col0) ...]
// (the mailbox the workers read the conv2 group from).
template <typename ColFn>
__device__ __forceinline__ bool exchange_update(const PersistArgs& a, const float4* gs, int nitems,
                                                ColFn col_of, unsigned seq, bool write_grad,
                                                uint2* box = nullptr, int box_col0 = 0,
                                                unsigned tag = 0u, int phase = 0, int t = -1,
                                                int nthreads = NT) {
  // phase 0: push to the peers, then gather + update; 1: push only; 2: gather + update only
  // (the caller pushed earlier). t / nthreads: the threads that take part (default: the CTA).
  if (t < 0) t = threadIdx.x;
  const bool multi = a.multi && a.comm.world > 1;
  if (multi && phase != 2)
    for (int i = t; i < nitems; i += nthreads) ll_push(a, col_of(i), gs[i], seq);
  if (phase == 1) return true;
  const float lr = a.hp[0], mu = a.hp[1];
  const float inv = multi ? 1.0f / (float)a.comm.world : 1.0f;
  bool ok = true;
  for (int i = t; i < nitems; i += nthreads) {
    const int col = col_of(i);
    float4 pv = make_float4(0.f, 0.f, 0.f, 0.f), bv = pv;
    if (a.update || box) pv = ldcg4(a.params + 4 * col);   // operands of the update do not
    if (a.update) bv = ldcg4(a.mom + 4 * col);             // depend on the peers: fetch them first
    float4 g = gs[i];
    if (multi) {
      ok = ll_gather_sum(a, col, g, seq, g) && ok;
      g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
    }
    if (write_grad) reinterpret_cast<float4*>(a.grad_out)[col] = g;
    if (a.update) {
      bv.x = mu * bv.x + g.x; bv.y = mu * bv.y + g.y; bv.z = mu * bv.z + g.z; bv.w = mu * bv.w + g.w;
      pv.x -= lr * bv.x; pv.y -= lr * bv.y; pv.z -= lr * bv.z; pv.w -= lr * bv.w;
      reinterpret_cast<float4*>(a.params)[col] = pv;
      reinterpret_cast<float4*>(a.mom)[col] = bv;
    }
    if (box) {
      uint4* b4 = reinterpret_cast<uint4*>(box + 4 * (size_t)(col - box_col0));
      st_volatile_u4(b4, make_uint4(__float_as_uint(pv.x), tag, __float_as_uint(pv.y), tag));
      st_volatile_u4(b4 + 1, make_uint4(__float_as_uint(pv.z), tag, __float_as_uint(pv.w), tag));
    }
  }
  return ok;
}

// Fixed-order sum over `rows` partial rows of the float4 columns [c0, c0 + nc) into gs[0..nc).
// Thread (rg, c): lanes run along the columns (coalesced), row groups split the rows; a thread's
// loads are issued U at a time so a whole batch costs one L2 round trip.
template <int U>
__device__ __forceinline__ void reduce_rows_slice(const PersistArgs& a, int ls,
                                                  const float* __restrict__ part, int rows,
                                                  int row_floats, int c0, int nc, float4* gs,
                                                  float4* red) {
  const int t = threadIdx.x;
  for (int cb = 0; cb < nc; cb += NT) {            // more columns than threads: several passes
    const int ncp = min(nc - cb, NT);
    const int RG = NT / ncp;                         // >= 1
    const int rg = t / ncp, c = t - rg * ncp;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rg < RG) {
      const float* p = part + 4 * (size_t)(c0 + cb + c);
      for (int r0 = rg; r0 < rows; r0 += U * RG) {
        // unconditional loads (row index clamped) so all 16 are issued before the first use;
        // a predicated load makes the compiler keep each load next to its add, one L2 round
        // trip per row
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldcg4(p + (size_t)min(r0 + u * RG, rows - 1) * row_floats);
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (r0 + u * RG < rows) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
      }
      red[t] = s;
    }
    __syncthreads();
    if (t < ncp) {
      float4 acc = red[t];
      for (int g = 1; g < RG; ++g) {
        const float4 v = red[g * ncp + t];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      gs[cb + t] = acc;
    }
    __syncthreads();
  }
}

// Stage `count` float4s into shared memory with every global load in flight at once (<= 8 per
// thread per batch). addr(i) -> global float4 address, dst(i) -> shared destination.
template <typename AddrFn, typename DstFn>
__device__ __forceinline__ void stage_f4(int count, AddrFn addr, DstFn dst) {
  const int t = threadIdx.x;
  for (int base = 0; base < count; base += 8 * NT) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = ldcg4(addr(min(base + t + u * NT, count - 1)));   // unconditional: all in flight
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = base + t + u * NT;
      if (i < count) *dst(i) = v[u];
    }
  }
}

// =========================================================================================
// WORKER
// =========================================================================================
template <bool SPLIT>
__device__ __forceinline__ void worker_loop(const PersistArgs& a, float* smem) {
  const WorkerSmem w = carve_worker(smem);
  const int t = threadIdx.x;
  const int role = SPLIT ? (int)(blockIdx.x & 1) : 2;
  const int slot = SPLIT ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int nslots = SPLIT ? a.n_workers >> 1 : a.n_workers;
  // the halves of the backward this CTA owns
  const int co0 = SPLIT ? role * 10 : 0, nco = SPLIT ? 10 : C2;
  const int ci0 = SPLIT ? role * 5 : 0, nci = SPLIT ? 5 : C1;
  unsigned* ctl = a.ctl;

  if (t == 0) {
    for (int i = 1; i < 5; ++i) mbar_init(w.bar + i, 1);   // [1] conv2, [2] fc, [3..4] images
    fence_barrier_init();
  }
  const unsigned long long step0 = a.step_ctr[0];
  const unsigned seq0 = a.multi ? *a.seq_ctr : 0u;
  const long long idx_base0 = (a.idx_stride != 0 && !a.epoch_begin)
                                  ? (long long)(step0 - a.step_ctr[1]) * a.idx_stride : 0ll;
  __syncthreads();
  if (t == 0) red_release_gpu_add_u32(ctl + C_STARTED * CTL_STRIDE, 1u);

  const int ipc = a.ipc;
  const int nitems = a.nsteps * ipc;       // (step, image) work items of this CTA, in order
  const size_t row_bytes = (a.x_mode == 0 || a.x_mode == 3) ? IMG_SZ * 4 : IMG_SZ;
  const bool is_u8 = (a.x_mode == 1 || a.x_mode == 2);
  const float inv_B = 1.0f / (float)a.B;
  float* row2 = a.part2 + (size_t)slot * P2_ROW;
  uint2* row1 = a.part1 + (size_t)slot * C1_PAIRS;

  // Issue the bulk copy of item k's image (lane 0 of warp 0 only). Returns false if the
  // item has no image (img >= B) or its data is not there yet and `block` is false.
  int issued = 0;   // items whose image copy has been issued (thread 0's view)
  auto item_img = [&](int k, int& ls, int& img) {
    ls = k / ipc;
    img = slot + (k - ls * ipc) * nslots;
  };
  auto issue_image = [&](int k, bool block) -> bool {
    int ls, img;
    item_img(k, ls, img);
    unsigned long long* bar = w.bar + 3 + (k & 1);
    if (img >= a.B) {            // ragged tail: nothing to fetch, complete the phase empty
      mbar_arrive(bar);
      return true;
    }
    if (a.feed_ready) {
      const unsigned g = a.feed_base + (unsigned)ls;
      const unsigned want = g + 1u;
      const unsigned* word = a.feed_ready + 32u * (g & 1u);
      if (block) {
        if (!wait_ge(word, want, a, 20u, true)) { mbar_arrive(bar); return true; }
      } else if ((int)(ld_acquire_sys_u32(word) - want) < 0) {
        return false;
      }
    }
    size_t r;
    if (a.x_mode >= 2) r = (size_t)__ldg(a.gidx + idx_base0 + (long long)ls * a.idx_stride + img);
    else r = (size_t)((a.ring_pos + (unsigned)ls) % (unsigned)a.ring_slots) * a.B + img;
    const char* src = (const char*)a.x + r * row_bytes;
    void* dst = is_u8 ? (void*)(w.xu + (k & 1) * 784) : (void*)(w.xs + (k & 1) * IMG_SZ);
    fence_proxy_async();
    mbar_expect_tx(bar, (unsigned)row_bytes);
    bulk_g2s(dst, src, (unsigned)row_bytes, bar);
    return true;
  };

  constexpr int FETCH_T = 288;   // lane 0 of warp 9
  constexpr unsigned long long W2_COPY_DELAY_NS = 900ull;
  unsigned long long bubble_end_ns = 0ull;
  bool pending = false;            // a pushed conv2 slice waits for its gather + update
  int pend_c0 = 0, pend_nc = 0;
  unsigned pend_seq = 0u, pend_tag = 0u;
  constexpr int SIGNAL_T = 300;  // a thread of warp 9 that the backward's wide phases leave idle
  for (int k = 0; k < nitems; ++k) {
    int ls, img;
    item_img(k, ls, img);
    const bool first_of_step = (k == ls * ipc);
    const bool last_of_step = (k == ls * ipc + ipc - 1);
    const bool have = img < a.B;
    const unsigned long long step = step0 + (unsigned long long)ls;

    // ---- fetch duties belong to ONE thread of warp 9, the warp conv1 leaves idle: it waits
    // for a parameter group's update, then issues the bulk copy that completes on the
    // group's mbarrier. conv1's group comes first (it is the last one published).
    // (the conv2 and fc groups; conv1's 1 KB comes through the tagged mailbox below)
    auto fetch_fc = [&](bool block) -> bool {
      const unsigned* ctr = ctl + C_DFC * CTL_STRIDE;
      const unsigned want = (unsigned)ls * (unsigned)(a.n3 + 1);
      if (!block && (int)(ld_acquire_gpu_u32(ctr) - want) < 0) return false;
      if (wait_ge(ctr, want, a, 23u)) {
        fence_proxy_async();
        mbar_expect_tx(w.bar + 2, (unsigned)(FLAT_SZ - OFF_W3) * 4u);
        bulk_g2s(w.P + OFF_W3, a.params + OFF_W3, (unsigned)(FLAT_SZ - OFF_W3) * 4u, w.bar + 2);
      } else {
        mbar_arrive(w.bar + 2);   // aborted: complete the phase so nobody hangs on it
      }
      return true;
    };
    // conv2's weights: at the first step of a launch straight from global memory; afterwards
    // the {value, tag} mailbox that every worker filled a slice of at the end of the last step
    // is bulk-copied into the (idle) T buffer and validated word by word after conv1.
    auto fetch_conv2 = [&]() {
      fence_proxy_async();
      if (ls == 0) {
        mbar_expect_tx(w.bar + 1, (unsigned)P2_ROW * 4u);
        bulk_g2s(w.P + OFF_W2, a.params + OFF_W2, (unsigned)P2_ROW * 4u, w.bar + 1);
      } else {
        mbar_expect_tx(w.bar + 1, (unsigned)P2_ROW * 8u);
        bulk_g2s(w.Ts, a.w2box, (unsigned)P2_ROW * 8u, w.bar + 1);
      }
    };
    if (t == FETCH_T && issued <= k) { issue_image(k, true); issued = k + 1; }
    const bool finish_conv2 = first_of_step && pending;
    if (finish_conv2) {
      // Several GPUs: the conv2 slice was only PUSHED to the peers in the bubble. While the
      // other warps wait for conv1's weights (whose exchange is the longer trip), warp 9 waits
      // for the peers' words, sums them, applies SGD-momentum and publishes the slice, then
      // pulls the whole mailbox in: the conv2 exchange overlaps conv1's instead of preceding it.
      if (t >= FETCH_T) {
        const int c0 = pend_c0;
        const bool ok = exchange_update(a, reinterpret_cast<const float4*>(w.Ts), pend_nc,
                                        [=](int i) { return OFF_W2 / 4 + c0 + i; }, pend_seq, false,
                                        a.w2box, OFF_W2 / 4, pend_tag, 2, t - FETCH_T, NT - FETCH_T);
        if (!ok) atomicExch(ctl + C_ABORT * CTL_STRIDE, 1u);
        __syncwarp();
        if (t == FETCH_T) bubble_end_ns = globaltimer_ns();   // published; the copy follows below
      }
      pending = false;
    }
    if (t == FETCH_T && first_of_step && ls > 0 && !finish_conv2) {
      // Every worker published its conv2 slice at the end of its bubble, all within a fraction
      // of a microsecond of each other. Give the last ones time to land, then pull the whole
      // mailbox in with one bulk copy that completes under conv1's forward; a word copied too
      // early is caught by its tag and fetched again.
      while (globaltimer_ns() - bubble_end_ns < W2_COPY_DELAY_NS) { }
      fetch_conv2();
    }

    // dropout masks (Philox keyed by step and sample: replays and the two-kernel path agree)
    PSTAMP(ls, 0);
    int tgt = 0;
    if (have) {
      if (t >= 32 && t < 32 + C2) {
        const int c = t - 32;
        float m = 1.0f;
        if (a.mask_mode == 1) m = dropout_mask_value(a.seed, step, a.sample_base + img, 0x20000u + c, 0.5f);
        else if (a.mask_mode == 2) m = a.m2buf[(size_t)img * C2 + c];
        w.m2s[c] = m;
      } else if (t >= 64 && t < 64 + FC1) {
        const int j = t - 64;
        float m = 1.0f;
        if (a.mask_mode == 1) m = dropout_mask_value(a.seed, step, a.sample_base + img, 0x30000u + j, 0.5f);
        else if (a.mask_mode == 2) m = a.m3buf[(size_t)img * FC1 + j];
        w.m3s[j] = m;
      }
    }
    // ---- image
    mbar_wait(w.bar + 3 + (k & 1), (unsigned)(k >> 1) & 1u);
    if (have && t < 32) {
      // the label is read only now: in host-fed mode the batch is known to have landed once
      // the image copy completed (the fetch thread waited for the step's ready value)
      size_t r;
      if (a.x_mode >= 2) r = (size_t)__ldg(a.gidx + idx_base0 + (long long)ls * a.idx_stride + img);
      else r = (size_t)((a.label_pos + (unsigned)ls) % (unsigned)a.label_slots) * a.B + img;
      tgt = (int)__ldcg(a.labels + r);   // ring slots are rewritten while the kernel runs: not via L1
    }
    float* xs = w.xs + (k & 1) * IMG_SZ;
    if (is_u8 && have) {
      const uchar4* src = reinterpret_cast<const uchar4*>(w.xu + (k & 1) * 784);
      if (t < IMG_SZ / 4) {
        const uchar4 v = src[t];
        reinterpret_cast<float4*>(xs)[t] =
            make_float4(normalize_px(v.x), normalize_px(v.y), normalize_px(v.z), normalize_px(v.w));
      }
    }
    // ---- conv1's weights: at the first step of a launch from global memory, afterwards from
    // the mailbox the two conv1 service CTAs fill with {value, tag} pairs. Each of 130 threads
    // polls its own 16-byte word: the data is the signal, one L2 round trip after it lands.
    if (first_of_step && t < C1_HALF) {
      const int e0 = 2 * t;                       // pairs e0, e0 + 1 (same half: 130 is even)
      const int h = e0 / C1_HALF, i0 = e0 - h * C1_HALF;
      float v0, v1;
      if (ls == 0) {
        v0 = __ldcg(a.params + c1_flat(h, i0));
        v1 = __ldcg(a.params + c1_flat(h, i0 + 1));
      } else {
        const unsigned want = a.tag_base | (unsigned)ls;     // tag of local step ls - 1
        const uint4* src = reinterpret_cast<const uint4*>(a.w1box) + t;
        uint4 v = ld_volatile_u4(src);
        if (!(v.y == want && v.w == want)) {
          const unsigned long long t0 = globaltimer_ns();
          for (unsigned polls = 1;; ++polls) {
            v = ld_volatile_u4(src);
            if (v.y == want && v.w == want) break;
            if ((polls & 255u) == 0u) {
              if (ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE) != 0u) break;
              if (globaltimer_ns() - t0 > 2ull * a.timeout_ns) { set_abort(a, 20u, 0xffffu, want); break; }
            }
          }
        }
        v0 = __uint_as_float(v.x);
        v1 = __uint_as_float(v.z);
      }
      w.P[c1_flat(h, i0)] = v0;
      w.P[c1_flat(h, i0 + 1)] = v1;
    }
    if (t == 0) w.flags[0] = (int)ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE);
    __syncthreads();
    if (w.flags[0] != 0) break;     // uniform: one thread sampled the abort word
    PSTAMP(ls, 1);
    bool fc_issued = !first_of_step;
    if (t == FETCH_T && finish_conv2) {          // under conv1's forward, which warp 9 sits out
      while (globaltimer_ns() - bubble_end_ns < W2_COPY_DELAY_NS) { }
      fetch_conv2();
    }
    if (t == FETCH_T && first_of_step) {
      if (ls == 0) fetch_conv2();                // later steps: issued during the wait for conv1's weights
      fc_issued = fetch_fc(false);               // fc only if it is already published
    }
    if (have) conv1_fwd_img(xs, w.P + OFF_W1, w.P + OFF_B1, w.a1s, w.pos1s);
    if (t == FETCH_T) {
      if (!fc_issued) fc_issued = fetch_fc(false);
      // the other image buffer is free (its item has finished): fetch one item ahead
      if (k + 1 < nitems && issued == k + 1 && issue_image(k + 1, false)) issued = k + 2;
    }
    if (first_of_step) {
      mbar_wait(w.bar + 1, (unsigned)ls & 1u);
      if (ls > 0) {
        // unpack the mailbox copy: a word whose tags are not this step's was copied before its
        // owner had written it; fetch that word again from global memory until it is valid
        const unsigned want = a.tag_base | (unsigned)ls;
        const uint4* st4 = reinterpret_cast<const uint4*>(w.Ts);
        const uint4* box4 = reinterpret_cast<const uint4*>(a.w2box);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = t + u * NT;
          if (i < P2_ROW / 2) {
            uint4 v = st4[i];
            if (!(v.y == want && v.w == want)) {
              const unsigned long long t0 = globaltimer_ns();
              for (unsigned polls = 1;; ++polls) {
                v = ld_volatile_u4(box4 + i);
                if (v.y == want && v.w == want) break;
                if ((polls & 255u) == 0u) {
                  if (ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE) != 0u) break;
                  if (globaltimer_ns() - t0 > 2ull * a.timeout_ns) { set_abort(a, 22u, 0xffffu, want); break; }
                }
              }
            }
            *reinterpret_cast<float2*>(w.P + OFF_W2 + 2 * i) =
                make_float2(__uint_as_float(v.x), __uint_as_float(v.z));
          }
        }
      }
    }
    __syncthreads();
    PSTAMP(ls, 2);
    if (have) conv2_fwd_img(w.a1s, w.P + OFF_W2, w.P + OFF_B2, w.m2s, w.a2s, w.pos2s);
    if (t == FETCH_T && !fc_issued) fetch_fc(true);
    if (first_of_step) mbar_wait(w.bar + 2, (unsigned)ls & 1u);
    __syncthreads();
    PSTAMP(ls, 3);
    if (have) {
      fc1_fwd_s(w.a2s, w.P + OFF_W3, w.P + OFF_B3, w.m3s, w.h1s);
      __syncthreads();
      PSTAMP(ls, 4);
      fc2_fwd_img(w.h1s, w.P + OFF_W4, w.P + OFF_B4, w.logit);
      __syncthreads();
      float loss_img = 0.0f;
      if (t < 32) {
        // log_softmax + NLL + its gradient (src/model.py:22, src/train.py:74), one class per lane
        const float z = (t < NCLS) ? w.logit[t] : -INFINITY;
        const float mx = warp_max(z);
        const float e = (t < NCLS) ? expf(z - mx) : 0.0f;
        const float lse = mx + logf(warp_sum(e));
        const float lp = z - lse;
        if (t < NCLS) w.dlog[t] = (expf(lp) - (t == tgt ? 1.0f : 0.0f)) * inv_B;
        loss_img = -__shfl_sync(0xffffffffu, lp, tgt) * inv_B;
      }
      __syncthreads();
      PSTAMP(ls, 5);
      fc2_bwd_img(w.dlog, w.h1s, w.m3s, w.P + OFF_W4, w.dpre, nullptr, nullptr, true, false);
      __syncthreads();
      // ---- publish this image's fc operands: the service CTAs form dW3, db3, dW4, db4
      if (role != 1) {
        float* act = a.acts + (size_t)img * ACT_STRIDE;
        act[ACT_A2 + t] = w.a2s[t];                                  // NT == 320 == A2_SZ
        if (t < FC1) { act[ACT_DPRE + t] = w.dpre[t]; act[ACT_H1 + t] = w.h1s[t]; }
        if (t < NCLS) act[ACT_DLOG + t] = w.dlog[t];
        if (t == 0) act[ACT_LOSS] = loss_img;
      }
      fc1_bwd_img<true>(w.dpre, w.a2s, w.m2s, w.P + OFF_W3, w.g2s, nullptr, nullptr, true, false);
      __syncthreads();
      PSTAMP(ls, 6);
      build_conv2_dense(w.g2s, w.pos2s, w.m2s, w.Ds, w.live, w.slot_of);
      __syncthreads();
      PSTAMP(ls, 7);
      // Signals are raised by a thread that has no work in the phase that follows (conv2's
      // weight gradient uses threads 0..275): the CTA barrier above ordered every thread's
      // stores before its fence, and the fence's wait for L2 acknowledgements costs nobody.
      if (role != 1 && t == SIGNAL_T) { __threadfence(); signal_add(ctl + C_ACT * CTL_STRIDE); }
      // conv2's weight gradient first: its reduction then hides behind dgrad + conv1 wgrad
      if (SPLIT) {
        conv2_wgrad_part<false>(w.g2s, w.pos2s, w.m2s, w.a1s, row2, row2 + W2_SZ, co0, nco, true);
      } else {
        conv2_wgrad_part<true>(w.g2s, w.pos2s, w.m2s, w.a1s, w.acc + OFF_W2, w.acc + OFF_B2, 0, C2,
                               first_of_step);
      }
      PSTAMP(ls, 8);
      if (SPLIT) __syncthreads();       // every thread's stores of the conv2 partial row are issued
      if (SPLIT && t == SIGNAL_T) { __threadfence(); signal_add(ctl + C_P2 * CTL_STRIDE); }   // idle in the product
      conv2_dgrad_T_part(w.Ds, w.live, w.live[C2], w.P + OFF_W2, w.Ts, ci0, nci);
      __syncthreads();
      PSTAMP(ls, 9);
      conv2_dgrad_gather_part(w.Ts, w.a1s, w.g1s, ci0, nci);
      __syncthreads();
      PSTAMP(ls, 10);
      if (SPLIT) {
        // an idle thread makes sure every CTA's conv2 partial row is there (they were
        // signalled two phases ago, so this returns at once) while the others finish conv1
        if (t == SIGNAL_T)
          w.flags[1] = wait_ge(ctl + C_P2 * CTL_STRIDE, (unsigned)(ls + 1) * (unsigned)a.n_workers, a, 31u) ? 1 : 0;
        conv1_wgrad_half(w.g1s, w.pos1s, xs, row1, role, a.tag_base | (unsigned)(ls + 1));
        __syncthreads();                // xs / g1s are reused by the next item
        PSTAMP(ls, 11);
      } else {
        conv1_wgrad_img(w.g1s, w.pos1s, xs, w.acc + OFF_W1, w.acc + OFF_B1, first_of_step);
        __syncthreads();
      }
    }
    if (!SPLIT && last_of_step) {
      // several images per CTA: one partial row per step, written once
      if (slot < a.B) {
        for (int i = t; i < P2_ROW / 4; i += NT)
          reinterpret_cast<float4*>(row2)[i] = reinterpret_cast<const float4*>(w.acc + OFF_W2)[i];
        if (t < C1_PAIRS) {
          const int h = t / C1_HALF, i = t - h * C1_HALF;
          st_pair(row1 + t, w.acc[c1_flat(h, i)], a.tag_base | (unsigned)(ls + 1));
        }
        __threadfence();
        __syncthreads();
        if (t == 0) {
          signal_add(ctl + C_P2 * CTL_STRIDE);
          w.flags[1] = wait_ge(ctl + C_P2 * CTL_STRIDE, (unsigned)(ls + 1) * (unsigned)a.n_workers, a, 31u) ? 1 : 0;
        }
        __syncthreads();
      }
    }
    if (last_of_step) {
      // ---- the bubble between two steps (conv1's gradients are on their way to the service
      // CTAs and back): every worker sums its slice of the conv2 partial rows, exchanges it with
      // the peers, applies SGD-momentum and publishes the new weights as {value, tag} pairs.
      if (w.flags[1] == 0) break;
      const int total = P2_ROW / 4;                                   // 1255 float4 columns
      const int per = (total + a.n_workers - 1) / a.n_workers;
      const int c0 = (int)blockIdx.x * per, nc = max(0, min(per, total - c0));
      float4* gs = reinterpret_cast<float4*>(w.Ts);
      float4* red = gs + ((per + 3) & ~3);
      reduce_rows_slice<4>(a, ls, a.part2, a.rows, P2_ROW, c0, nc, gs, red);
      // with peers and another step to come, only push now and finish under the next conv1
      const bool defer = a.multi && a.comm.world > 1 && k + 1 < nitems;
      const bool ok = exchange_update(a, gs, nc, [=](int i) { return OFF_W2 / 4 + c0 + i; },
                                      seq0 + (unsigned)ls + 1u, a.grad_out != nullptr && ls == a.nsteps - 1,
                                      a.w2box, OFF_W2 / 4, a.tag_base | (unsigned)(ls + 1), defer ? 1 : 0);
      if (!ok) atomicExch(ctl + C_ABORT * CTL_STRIDE, 1u);
      if (defer) {
        pending = true; pend_c0 = c0; pend_nc = nc;
        pend_seq = seq0 + (unsigned)ls + 1u; pend_tag = a.tag_base | (unsigned)(ls + 1);
      }
      __syncthreads();               // the T buffer is the next step's mailbox landing zone
      if (t == FETCH_T) bubble_end_ns = globaltimer_ns();
      PSTAMP(ls, 12);
    }
  }
}

// =========================================================================================
// SERVICE
// =========================================================================================
// ---- conv1 service: one CTA per half (130 gradients) ---------------------------------------
// Polls the workers' {value, tag} rows, sums them in row order, exchanges with the peers,
// applies SGD-momentum out of registers and answers with {new weight, tag} pairs. Nothing on
// this path is a fence or a flag: it is the only link between two steps' critical paths.
__device__ __forceinline__ void service_conv1(const PersistArgs& a, float* smem, int h,
                                              unsigned long long step0, unsigned seq0) {
  const int t = threadIdx.x;
  float2* red = reinterpret_cast<float2*>(smem);          // [4][65]
  int* flags = reinterpret_cast<int*>(smem + 2 * 4 * 65 + 8);
  unsigned* ctl = a.ctl;
  const int c = t % 65, rg = t / 65;                      // t < 260: uint4 column c, row group rg
  const bool loader = t < 260;
  const int rows = a.rows;
  // element owned by thread i < 130 (registers across steps)
  const int f = c1_flat(h, t < C1_HALF ? t : 0);
  float pv = 0.0f, bv = 0.0f;
  if (t < C1_HALF) { pv = __ldcg(a.params + f); bv = __ldcg(a.mom + f); }
  const bool multi = a.multi && a.comm.world > 1;
  const int W = a.comm.world, me = a.comm.rank;
  for (int ls = 0; ls < a.nsteps; ++ls) {
    PSTAMP(ls, 0);
    const unsigned tag = a.tag_base | (unsigned)(ls + 1);
    const unsigned seq = seq0 + (unsigned)ls + 1u;
    float sx = 0.0f, sy = 0.0f;
    bool ok = true;
    if (loader) {
      const uint4* base = reinterpret_cast<const uint4*>(a.part1) + h * 65 + c;
      for (int r0 = rg; r0 < rows && ok; r0 += 16 * 4) {
        uint4 v[16];
        unsigned pending = 0;
        // first round: every row's word, unconditionally, so the 16 loads overlap
#pragma unroll
        for (int u = 0; u < 16; ++u) v[u] = ld_volatile_u4(base + (size_t)min(r0 + 4 * u, rows - 1) * C1_HALF);
#pragma unroll
        for (int u = 0; u < 16; ++u)
          if (r0 + 4 * u < rows && !(v[u].y == tag && v[u].w == tag)) pending |= 1u << u;
        unsigned long long t0 = 0;
        unsigned polls = 0;
        while (pending) {
#pragma unroll
          for (int u = 0; u < 16; ++u)
            if (pending & (1u << u)) v[u] = ld_volatile_u4(base + (size_t)(r0 + 4 * u) * C1_HALF);
#pragma unroll
          for (int u = 0; u < 16; ++u)
            if ((pending & (1u << u)) && v[u].y == tag && v[u].w == tag) pending &= ~(1u << u);
          if (pending && (++polls & 127u) == 0u) {
            if (t0 == 0) t0 = globaltimer_ns();
            if (ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE) != 0u) { ok = false; break; }
            if (globaltimer_ns() - t0 > 2ull * a.timeout_ns) { set_abort(a, 30u, 0xffffu, tag); ok = false; break; }
          }
        }
#pragma unroll
        for (int u = 0; u < 16; ++u)
          if (r0 + 4 * u < rows) { sx += __uint_as_float(v[u].x); sy += __uint_as_float(v[u].z); }
      }
      red[rg * 65 + c] = make_float2(sx, sy);
    }
    if (t == 0) flags[0] = 1;
    __syncthreads();
    if (!ok) flags[0] = 0;
    __syncthreads();
    if (!flags[0]) break;
    PSTAMP(ls, 1);
    if (t < C1_HALF) {
      const int cc = t >> 1;
      float g = 0.0f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {                       // row-group order: deterministic
        const float2 v = red[q * 65 + cc];
        g += (t & 1) ? v.y : v.x;
      }
      PSTAMP(ls, 2);
      if (multi) {
        // element-granular tagged exchange: 8-byte {value, seq} pairs in the Net staging
        const unsigned slot = seq & 1u;
        const size_t eoff = (size_t)f;
        for (int pp = 1; pp < W; ++pp) {
          const int p = (me + pp) % W;
          uint2* dst = peer_ptr<uint2>(a.comm, p, a.stage_off) + ((size_t)slot * W + me) * FLAT_SZ + eoff;
          st_pair(dst, g, seq);
        }
        const uint2* mine = peer_ptr<uint2>(a.comm, me, a.stage_off) + (size_t)slot * W * FLAT_SZ + eoff;
        float val[B200_MAX_WORLD];
        unsigned pending = ((1u << W) - 1u) & ~(1u << me);
        unsigned long long t0 = 0;
        unsigned polls = 0;
        while (pending) {
          uint2 v[B200_MAX_WORLD];
#pragma unroll
          for (int r = 0; r < B200_MAX_WORLD; ++r)
            if (pending & (1u << r)) {
              asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];"
                           : "=r"(v[r].x), "=r"(v[r].y) : "l"(mine + (size_t)r * FLAT_SZ) : "memory");
            }
#pragma unroll
          for (int r = 0; r < B200_MAX_WORLD; ++r)
            if ((pending & (1u << r)) && v[r].y == seq) { val[r] = __uint_as_float(v[r].x); pending &= ~(1u << r); }
          if (pending && (++polls & 127u) == 0u) {
            if (t0 == 0) t0 = globaltimer_ns();
            if (ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE) != 0u) { ok = false; break; }
            if (globaltimer_ns() - t0 > a.timeout_ns) {
              const int peer = __ffs((int)pending) - 1;
              if (atomicCAS(a.comm.err, 0u, 1u) == 0u) {
                a.comm.err[1] = (unsigned)peer; a.comm.err[2] = 12u; a.comm.err[3] = seq;
              }
              atomicExch(ctl + C_ABORT * CTL_STRIDE, 1u);
              ok = false;
              break;
            }
          }
        }
        float sum = 0.0f;
#pragma unroll
        for (int r = 0; r < B200_MAX_WORLD; ++r)
          if (r < W) sum += (r == me || !ok) ? g : val[r];
        g = sum * (1.0f / (float)W);
      }
      if (a.grad_out && ls == a.nsteps - 1) a.grad_out[f] = g;
      if (a.update) {
        const float lr = a.hp[0], mu = a.hp[1];
        bv = mu * bv + g;
        pv -= lr * bv;
        a.params[f] = pv;
        a.mom[f] = bv;
      }
      st_pair(a.w1box + h * C1_HALF + t, pv, tag);        // the answer the workers poll
    }
    PSTAMP(ls, 3);
    if (h == 0 && t == 0 && a.feed_consumed) {
      // every worker pair has delivered step ls, so its input slot can be refilled (no worker
      // is ever more than one step behind the others; slots are reused half a ring later)
      st_release_sys_u32(a.feed_consumed, a.feed_base + (unsigned)ls + 1u);
    }
    PSTAMP(ls, 4);
    if (!ok) { atomicExch(ctl + C_ABORT * CTL_STRIDE, 1u); break; }
  }
  // ---------------- end of launch: advance the persistent counters exactly once
  if (h == 0 && t == 0) {
    if (wait_ge(ctl + C_STARTED * CTL_STRIDE, gridDim.x, a, 40u)) {
      const bool aborted = ld_volatile_u32(ctl + C_ABORT * CTL_STRIDE) != 0u;
      if (!aborted) {
        a.step_ctr[0] = step0 + (unsigned long long)a.nsteps;
        if (a.epoch_begin) a.step_ctr[1] = step0;
        if (a.multi) *a.seq_ctr = seq0 + (unsigned)a.nsteps;
      }
    }
  }
}

__device__ __forceinline__ void service_loop(const PersistArgs& a, float* smem) {
  ServiceSmem sv;
  sv.gs = reinterpret_cast<float4*>(smem);
  sv.red = sv.gs + SV_ITEMS_MAX;
  sv.stage = reinterpret_cast<float*>(sv.red + NT);
  sv.flags = reinterpret_cast<int*>(sv.stage + SV_STAGE_FLOATS);
  const int t = threadIdx.x;
  const int sidx = (int)blockIdx.x - a.n_workers;
  unsigned* ctl = a.ctl;
  // kind: 0 conv1 half (two CTAs), 1 small fc (+ loss), 3 W3 column slice
  // (conv2's partial rows are summed by the workers themselves between two steps)
  const int kind = sidx < 2 ? 0 : sidx == 2 ? 1 : 3;
  const int part = kind == 0 ? sidx : kind == 3 ? sidx - 3 : 0;
  const unsigned long long step0 = a.step_ctr[0];
  const unsigned seq0 = a.multi ? *a.seq_ctr : 0u;
  __syncthreads();
  if (t == 0) red_release_gpu_add_u32(ctl + C_STARTED * CTL_STRIDE, 1u);
  if (kind == 0) {
    service_conv1(a, smem, part, step0, seq0);
    return;
  }
  const int B = a.B;

  for (int ls = 0; ls < a.nsteps; ++ls) {
    const unsigned seq = seq0 + (unsigned)ls + 1u;
    const bool wg = a.grad_out != nullptr && ls == a.nsteps - 1;
    bool ok = true;
    PSTAMP(ls, 0);
    {
      // ---------------- fc gradients from the published activations
      if (t == 0) sv.flags[0] = wait_ge(ctl + C_ACT * CTL_STRIDE, (unsigned)(ls + 1) * (unsigned)B, a, 33u) ? 1 : 0;
      __syncthreads();
      if (!sv.flags[0]) break;
      PSTAMP(ls, 1);
      if (kind == 3) {
        // dW3[j][k] = sum_b dpre[b][j] * a2[b][k] for k in [k0, k0 + kw): thread item = (j, 4 columns)
        const int k0 = part * a.kw;
        const int kw = max(0, min(a.kw, A2_SZ - k0));
        const int kq = kw >> 2;                       // float4 groups per row
        const int nitems = FC1 * kq;
        float4 acc[3];
#pragma unroll
        for (int u = 0; u < 3; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        float* sd = sv.stage;            // [64][52] dpre (50 + 2 zero pads)
        float* sa = sv.stage + 64 * 52;  // [64][60] a2 columns (kw <= 60)
        const int per = 13 + kq;         // float4s staged per image
        for (int b0 = 0; b0 < B; b0 += 64) {
          const int nb = min(64, B - b0);
          stage_f4(nb * per,
                   [&](int i) {
                     const int b = i / per, q = i - b * per;
                     const float* row = a.acts + (size_t)(b0 + b) * ACT_STRIDE;
                     return q < 13 ? row + ACT_DPRE + 4 * q : row + ACT_A2 + k0 + 4 * (q - 13);
                   },
                   [&](int i) {
                     const int b = i / per, q = i - b * per;
                     return reinterpret_cast<float4*>(q < 13 ? sd + b * 52 + 4 * q : sa + b * 60 + 4 * (q - 13));
                   });
          __syncthreads();
          PSTAMP(ls, 5);
#pragma unroll
          for (int u = 0; u < 3; ++u) {
            const int it = t + u * NT;
            if (it < nitems) {
              const int j = it / kq, q = it - j * kq;
              float4 s = acc[u];
#pragma unroll 8
              for (int b = 0; b < nb; ++b) {
                const float d = sd[b * 52 + j];
                const float4 av = reinterpret_cast<const float4*>(sa + b * 60)[q];
                s.x = fmaf(d, av.x, s.x); s.y = fmaf(d, av.y, s.y);
                s.z = fmaf(d, av.z, s.z); s.w = fmaf(d, av.w, s.w);
              }
              acc[u] = s;
            }
          }
          PSTAMP(ls, 6);
          __syncthreads();
        }
#pragma unroll
        for (int u = 0; u < 3; ++u) {
          const int it = t + u * NT;
          if (it < nitems) sv.gs[it] = acc[u];
        }
        __syncthreads();
        PSTAMP(ls, 2);
        ok = exchange_update(a, sv.gs, nitems,
                             [=](int i) { const int j = i / kq; return (OFF_W3 + j * A2_SZ + k0) / 4 + (i - j * kq); },
                             seq, wg);
      } else {
        // b3 | W4 | b4 (flat [OFF_B3, FLAT_SZ)) and the batch loss
        float* gsf = reinterpret_cast<float*>(sv.gs);
        float accs[2] = {0.0f, 0.0f};
        float lsum = 0.0f;
        float* sd = sv.stage;             // [64][116]: dpre[50+2] h1[50+2] dlog[10] loss[1] (+1)
        constexpr int SD = 4 * ACT_SMALL_F4;   // 116
        constexpr int L_H1 = ACT_H1 - ACT_DPRE, L_DLOG = ACT_DLOG - ACT_DPRE, L_LOSS = ACT_LOSS - ACT_DPRE;
        for (int b0 = 0; b0 < B; b0 += 64) {
          const int nb = min(64, B - b0);
          stage_f4(nb * ACT_SMALL_F4,
                   [&](int i) {
                     const int b = i / ACT_SMALL_F4, q = i - b * ACT_SMALL_F4;
                     return a.acts + (size_t)(b0 + b) * ACT_STRIDE + ACT_DPRE + 4 * q;
                   },
                   [&](int i) { return reinterpret_cast<float4*>(sd) + i; });
          PSTAMP(ls, 6);
          __syncthreads();
          PSTAMP(ls, 5);
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int e = t + u * NT;          // element of the flat slice
            if (e < FC_SMALL_N) {
              const int f = FC_SMALL_LO + e;
              float s = accs[u];
              if (f < OFF_B3 + B3_SZ) {
                const int j = f - OFF_B3;
#pragma unroll 8
                for (int b = 0; b < nb; ++b) s += sd[b * SD + j];
              } else if (f >= OFF_W4 && f < OFF_W4 + W4_SZ) {
                const int c = (f - OFF_W4) / FC1, j = (f - OFF_W4) - c * FC1;
#pragma unroll 8
                for (int b = 0; b < nb; ++b) s = fmaf(sd[b * SD + L_DLOG + c], sd[b * SD + L_H1 + j], s);
              } else if (f >= OFF_B4 && f < OFF_B4 + B4_SZ) {
                const int c = f - OFF_B4;
#pragma unroll 8
                for (int b = 0; b < nb; ++b) s += sd[b * SD + L_DLOG + c];
              }
              accs[u] = s;
            }
          }
          if (t == NT - 1)
            for (int b = 0; b < nb; ++b) lsum += sd[b * SD + L_LOSS];   // image order: deterministic
          __syncthreads();
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int e = t + u * NT;
          if (e < FC_SMALL_N) gsf[e] = accs[u];
        }
        if (t == NT - 1) {
          const unsigned long long step = step0 + (unsigned long long)ls;
          if (a.loss_ring) {
            a.loss_ring[step % (unsigned long long)a.ring] = lsum;
            a.loss_ring[a.ring] = lsum;
          }
          if (a.host_loss) {
            // 4 bytes device -> host per step, written where the caller reads it
            asm volatile("st.global.wt.f32 [%0], %1;" ::"l"(a.host_loss + ls), "f"(lsum) : "memory");
          }
        }
        __syncthreads();
        PSTAMP(ls, 2);
        ok = exchange_update(a, sv.gs, FC_SMALL_N / 4, [=](int i) { return FC_SMALL_LO / 4 + i; }, seq, wg);
      }
    }
    // ---------------- publish: the group's new weights are in global memory
    PSTAMP(ls, 3);
    __threadfence();
    fence_proxy_async();
    __syncthreads();
    if (t == 0) red_release_gpu_add_u32(ctl + C_DFC * CTL_STRIDE, 1u);
    PSTAMP(ls, 4);
    if (!ok) break;
  }
  if (kind == 1 && t == 0 && a.host_loss) __threadfence_system();
}

__global__ void __launch_bounds__(NT, 1)
net_persist_kernel(const __grid_constant__ PersistArgs a) {
  extern __shared__ __align__(128) float smem[];
  if ((int)blockIdx.x < a.n_workers) {
    if (a.split) worker_loop<true>(a, smem);
    else worker_loop<false>(a, smem);
  } else {
    service_loop(a, smem);
  }
}

constexpr int PERSIST_SMEM = (WS_BYTES > SV_BYTES ? WS_BYTES : SV_BYTES) + 128;

struct Plan {
  int grid, n_workers, split, rows, ipc, n_service, n3, kw;
};

Plan make_plan(int B) {
  Plan p;
  const int S = 19;                       // service CTAs: 2 conv1 halves + 1 small fc + 16 W3 slices
  const int maxw = (148 - S) & ~1;        // 128 workers, the rest of the SMs serve
  if (2 * B <= maxw) {
    p.split = 1; p.n_workers = 2 * B; p.rows = B; p.ipc = 1;
  } else {
    p.split = 0; p.n_workers = B < maxw ? B : maxw; p.rows = p.n_workers;
    p.ipc = (B + p.n_workers - 1) / p.n_workers;
  }
  p.n3 = 16;                              // W3 CTAs: 20 columns each, 50 x 5 float4 items = one per thread
  p.kw = 20;
  p.n_service = 3 + p.n3;                 // 2 conv1 halves + small fc + W3 slices
  p.grid = p.n_workers + p.n_service;
  return p;
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" {

struct B200PersistHost {   // mirrored field by field in ops/persist.py (ctypes.Structure)
  float* params; float* mom; float* grad_out; const float* hp;
  const void* x; const long long* labels; const int* gidx;
  int x_mode; int B; int nsteps; int update;
  int idx_stride; int n_idx;
  int ring_slots; unsigned ring_pos;
  const unsigned* feed_ready; unsigned* feed_consumed; unsigned feed_base; int mask_mode;
  float* scratch;            // b200_net_persist_scratch_floats(B) floats
  float* loss_ring; int ring; int multi;
  float* host_loss;
  unsigned long long* step_ctr;
  unsigned long long seed;
  unsigned sample_base; int pad0;
  float* m2buf; float* m3buf;
  const void* comm;          // CommDev blob (host) when multi
  size_t stage_off;
  unsigned* seq_ctr;
  unsigned* err;             // 4 u32 (device), used when not multi
  unsigned long long timeout_ns;
  long long* dbg;            // optional phase stamps (b200_net_persist_dbg_words() int64), or null
  int label_slots; unsigned label_pos;   // direct modes; 0 slots = same ring as the images
  int epoch_begin; int pad1;
};

int b200_net_persist_host_size(void) { return (int)sizeof(B200PersistHost); }

// {grid, workers, split, rows, images per worker, service CTAs, conv2 CTAs, W3 CTAs, W3 columns per CTA}
int b200_net_persist_plan(int B, int* out9) {
  B200_REQUIRE(B > 0, "b200_net_persist_plan: empty batch");
  const Plan p = make_plan(B);
  const int o[9] = {p.grid, p.n_workers, p.split, p.rows, p.ipc, p.n_service, 0, p.n3, p.kw};
  for (int i = 0; i < 9; ++i) out9[i] = o[i];
  return 0;
}

// Scratch floats for a batch of B: activations, conv2 rows, conv1 rows, control words.
long long b200_net_persist_scratch_floats(int B) {
  const Plan p = make_plan(B > 0 ? B : 1);
  return (long long)B * ACT_STRIDE + (long long)p.rows * P2_ROW + (long long)p.rows * C1_PAIRS * 2 +
         C1_PAIRS * 2 + P2_ROW * 2 + C_NWORDS * CTL_STRIDE + 64;
}

int b200_net_persist_smem_bytes(void) { return PERSIST_SMEM; }
// int64 words of the optional phase-stamp buffer: [148][steps][phases][clock64, globaltimer]
long long b200_net_persist_dbg_words(int* steps, int* phases) {
  if (steps) *steps = DBG_STEPS;
  if (phases) *phases = DBG_PH;
  return (long long)148 * DBG_STEPS * DBG_PH * 2;
}

// Run `nsteps` optimizer steps in one cooperative launch on `stream`.
int b200_net_persist_run(const B200PersistHost* h, void* stream) {
  B200_REQUIRE(h && h->B > 0 && h->nsteps > 0, "b200_net_persist_run: bad batch/steps");
  B200_REQUIRE(h->x_mode >= 0 && h->x_mode <= 3, "b200_net_persist_run: bad x_mode");
  B200_REQUIRE(h->x_mode < 2 || h->gidx != nullptr, "b200_net_persist_run: gather modes need indices");
  B200_REQUIRE(h->x_mode < 2 || h->nsteps == 1 || h->idx_stride != 0,
               "b200_net_persist_run: several steps over a dataset need idx_stride");
  B200_REQUIRE(h->x_mode >= 2 || h->ring_slots > 0, "b200_net_persist_run: direct modes need ring_slots");
  B200_REQUIRE(h->mask_mode != 2 || (h->m2buf && h->m3buf && h->nsteps == 1),
               "b200_net_persist_run: injected masks need both buffers and a single step");
  B200_REQUIRE(!h->multi || h->comm, "b200_net_persist_run: multi needs a communicator");
  static bool attr_done[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!attr_done[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(net_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PERSIST_SMEM));
    attr_done[dev & 63] = true;
  }
  const Plan p = make_plan(h->B);
  PersistArgs a;
  memset(&a, 0, sizeof(a));
  a.params = h->params; a.mom = h->mom; a.grad_out = h->grad_out; a.hp = h->hp;
  a.x = h->x; a.labels = h->labels; a.gidx = h->gidx;
  a.x_mode = h->x_mode; a.B = h->B; a.nsteps = h->nsteps; a.update = h->update;
  a.n_workers = p.n_workers; a.split = p.split; a.rows = p.rows; a.ipc = p.ipc;
  a.n_service = p.n_service; a.n3 = p.n3; a.kw = p.kw;
  a.idx_stride = h->idx_stride; a.n_idx = h->n_idx;
  a.ring_slots = h->ring_slots > 0 ? h->ring_slots : 1; a.ring_pos = h->ring_pos;
  a.feed_ready = h->feed_ready; a.feed_consumed = h->feed_consumed; a.feed_base = h->feed_base;
  a.label_slots = h->label_slots > 0 ? h->label_slots : a.ring_slots;
  a.label_pos = h->label_slots > 0 ? h->label_pos : a.ring_pos;
  a.epoch_begin = h->epoch_begin;
  a.acts = h->scratch;
  a.part2 = a.acts + (size_t)h->B * ACT_STRIDE;
  a.part1 = reinterpret_cast<uint2*>(a.part2 + (size_t)p.rows * P2_ROW);
  a.w1box = a.part1 + (size_t)p.rows * C1_PAIRS;
  a.w2box = a.w1box + C1_PAIRS;
  a.ctl = reinterpret_cast<unsigned*>(a.w2box + P2_ROW);
  static std::atomic<unsigned> launch_id{1};
  a.tag_base = (launch_id.fetch_add(1) & 0xffffu) << 16;
  B200_REQUIRE(h->nsteps < 65536, "b200_net_persist_run: at most 65535 steps per launch");
  a.ctl = reinterpret_cast<unsigned*>(((uintptr_t)a.ctl + 127) & ~(uintptr_t)127);
  a.loss_ring = h->loss_ring; a.ring = h->ring > 0 ? h->ring : 1; a.host_loss = h->host_loss;
  a.step_ctr = h->step_ctr; a.seed = h->seed; a.sample_base = h->sample_base;
  a.mask_mode = h->mask_mode; a.m2buf = h->m2buf; a.m3buf = h->m3buf;
  a.multi = h->multi; a.stage_off = h->stage_off; a.seq_ctr = h->seq_ctr;
  a.timeout_ns = h->timeout_ns ? h->timeout_ns : 20000000000ull;
  a.dbg = h->dbg;
  if (h->multi) {
    memcpy(&a.comm, h->comm, sizeof(CommDev));
    a.err = a.comm.err;
    a.timeout_ns = a.comm.timeout_ns;
  } else {
    a.err = h->err;
  }
  B200_REQUIRE(a.err != nullptr, "b200_net_persist_run: error words missing");
  cudaStream_t s = (cudaStream_t)stream;
  B200_CUDA_OK(cudaMemsetAsync(a.ctl, 0, C_NWORDS * CTL_STRIDE * sizeof(unsigned), s));
  void* kargs[] = {(void*)&a};
  B200_CUDA_OK(cudaLaunchCooperativeKernel((const void*)net_persist_kernel, dim3(p.grid), dim3(NT),
                                           kargs, PERSIST_SMEM, s));
  return 0;
}


// -----------------------------------------------------------------------------------------
// Host feeder: `steps` optimizer steps from PINNED host memory through ONE launch of the
// persistent kernel (reference: the DataLoader(pin_memory=True) -> model(data) -> loss.item()
// loop, src/train_dist.py:40-45,78-89). Per step the copy stream carries the batch
// (images, labels: two cudaMemcpyAsync) into a ring slot and then a 32-bit stream write that
// tells the kernel the step is there; the kernel stores each step's loss into the caller's
// pinned buffer itself. The host never waits: when it runs a whole ring ahead of the kernel,
// the COPY STREAM waits (stream wait-value on the kernel's consumed counter).
// -----------------------------------------------------------------------------------------
#define DRVP(name)                                                                          \
  static decltype(&name) fn_##name = nullptr;                                               \
  if (!fn_##name) {                                                                         \
    void* p_ = nullptr;                                                                     \
    cudaDriverEntryPointQueryResult q_;                                                     \
    if (cudaGetDriverEntryPoint(#name, &p_, cudaEnableDefault, &q_) != cudaSuccess ||       \
        q_ != cudaDriverEntryPointSuccess || !p_) {                                         \
      b200_set_error("driver entry point %s is not available", #name);                      \
      return -2;                                                                            \
    }                                                                                       \
    fn_##name = reinterpret_cast<decltype(&name)>(p_);                                      \
  }
#define CUP_OK(expr)                                                                        \
  do {                                                                                      \
    CUresult r_ = (expr);                                                                   \
    if (r_ != CUDA_SUCCESS) {                                                               \
      b200_set_error("%s:%d: %s failed with CUresult %d", __FILE__, __LINE__, #expr, (int)r_); \
      return (int)r_;                                                                       \
    }                                                                                       \
  } while (0)

constexpr int FEED_LABEL_STEPS = 2048;   // label ring (steps): one label copy per launch

struct B200Feeder {
  B200PersistHost tmpl;
  CommDev comm;
  int R;                    // image ring slots (steps)
  size_t x_bytes, y_bytes;  // per step
  char* ringx;
  long long* ringy;         // FEED_LABEL_STEPS steps of labels
  unsigned* flags;          // device: [0] / [32] steps copied by copy stream 0 / 1, [64] steps consumed
  unsigned fed;             // steps fed since creation
  cudaStream_t compute, copy[2];
  cudaEvent_t join_in, join_compute, join_copy[2], labels_done;
  long long launches;
  const float* hloss_lo;    // last pinned loss buffer and its device alias
  float* dloss_lo;
  cudaStream_t last_cs;     // caller stream of the previous run
  bool last_cs_valid;
};

int b200_feeder_create(const B200PersistHost* tmpl, int ring_slots, void** out) {
  B200_REQUIRE(tmpl && out, "b200_feeder_create: null argument");
  B200_REQUIRE(tmpl->x_mode == 0 || tmpl->x_mode == 1, "b200_feeder_create: x_mode must be 0 (fp32) or 1 (uint8)");
  B200_REQUIRE(ring_slots >= 8 && ring_slots % 4 == 0 && FEED_LABEL_STEPS % ring_slots == 0,
               "b200_feeder_create: ring_slots must be a multiple of 4, >= 8, dividing %d", FEED_LABEL_STEPS);
  {
    DRVP(cuStreamWriteValue32)
    DRVP(cuStreamWaitValue32)
  }
  B200Feeder* f = new (std::nothrow) B200Feeder();
  B200_REQUIRE(f != nullptr, "b200_feeder_create: out of host memory");
  memset(f, 0, sizeof(*f));
  f->tmpl = *tmpl;
  if (tmpl->multi) {
    memcpy(&f->comm, tmpl->comm, sizeof(CommDev));
    f->tmpl.comm = &f->comm;
  }
  f->R = ring_slots;
  f->x_bytes = (size_t)tmpl->B * 784 * (tmpl->x_mode == 0 ? 4 : 1);
  f->y_bytes = (size_t)tmpl->B * sizeof(long long);
  B200_CUDA_OK(cudaMalloc((void**)&f->ringx, (size_t)f->R * f->x_bytes));
  B200_CUDA_OK(cudaMalloc((void**)&f->ringy, (size_t)FEED_LABEL_STEPS * f->y_bytes));
  B200_CUDA_OK(cudaMalloc((void**)&f->flags, 96 * sizeof(unsigned)));
  B200_CUDA_OK(cudaMemset(f->ringx, 0, (size_t)f->R * f->x_bytes));
  B200_CUDA_OK(cudaMemset(f->ringy, 0, (size_t)FEED_LABEL_STEPS * f->y_bytes));
  B200_CUDA_OK(cudaMemset(f->flags, 0, 96 * sizeof(unsigned)));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&f->compute, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaEventCreateWithFlags(&f->join_in, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&f->join_compute, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&f->labels_done, cudaEventDisableTiming));
  for (int p = 0; p < 2; ++p) {
    B200_CUDA_OK(cudaStreamCreateWithFlags(&f->copy[p], cudaStreamNonBlocking));
    B200_CUDA_OK(cudaEventCreateWithFlags(&f->join_copy[p], cudaEventDisableTiming));
  }
  B200_CUDA_OK(cudaDeviceSynchronize());
  *out = f;
  return 0;
}

// One launch of at most FEED_LABEL_STEPS steps. Stream operations per step: ONE host->device
// copy (the images) and ONE 32-bit stream write (the step has landed), alternating between two
// copy streams so two copy engines work at once; the labels of the whole launch go up in one
// copy beforehand (they are 8 bytes per sample).
// B200DP_FEED_TRACE=1: host timestamps (us since b200_feeder_run was entered) on stderr
static bool feed_trace_on() {
  static const bool on = [] { const char* e = getenv("B200DP_FEED_TRACE"); return e && atoi(e) != 0; }();
  return on;
}
static double g_feed_t0 = 0.0;
static double feed_now_us() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
#define FEED_TRACE(what) do { if (feed_trace_on()) fprintf(stderr, "[feed] %-28s %7.1f us\n", what, feed_now_us() - g_feed_t0); } while (0)

static int feeder_launch(B200Feeder* f, const char* px, const long long* hy, int steps, float* dloss) {
  DRVP(cuStreamWriteValue32)
  DRVP(cuStreamWaitValue32)
  const unsigned R = (unsigned)f->R, half = R / 2u;
  B200PersistHost a = f->tmpl;
  a.x = f->ringx;
  a.labels = f->ringy;
  a.gidx = nullptr;
  a.nsteps = steps;
  a.update = 1;
  a.grad_out = nullptr;
  a.idx_stride = 0;
  a.ring_slots = f->R;
  a.ring_pos = f->fed % R;
  a.label_slots = FEED_LABEL_STEPS;
  a.label_pos = f->fed % (unsigned)FEED_LABEL_STEPS;
  a.feed_ready = f->flags;
  a.feed_consumed = f->flags + 64;
  a.feed_base = f->fed;
  a.host_loss = dloss;
  a.epoch_begin = 0;
  static const int pre_cfg = [] {
    const char* e = getenv("B200DP_FEED_PRE");
    const int v = e ? atoi(e) : 1;     // measured: 1 copy ahead of the launch is best (0: the launch
    return v < 0 ? 0 : (v > 16 ? 16 : v);   // delays the first copy; 4: the copies delay the launch)
  }();
  const int pre_early = steps < pre_cfg ? steps : pre_cfg;
  if (pre_early == 0) {       // kernel first: its start-up runs while the host enqueues the copies
    if (int e = b200_net_persist_run(&a, f->compute)) return e;
    ++f->launches;
  }
  // labels: one or two contiguous pieces of the label ring
  {
    const unsigned pos = a.label_pos;
    const unsigned first = (unsigned)steps < FEED_LABEL_STEPS - pos ? (unsigned)steps : FEED_LABEL_STEPS - pos;
    B200_CUDA_OK(cudaMemcpyAsync(f->ringy + (size_t)pos * f->tmpl.B, hy, (size_t)first * f->y_bytes,
                                 cudaMemcpyHostToDevice, f->copy[0]));
    if (first < (unsigned)steps)
      B200_CUDA_OK(cudaMemcpyAsync(f->ringy, hy + (size_t)first * f->tmpl.B,
                                   (size_t)(steps - first) * f->y_bytes, cudaMemcpyHostToDevice, f->copy[0]));
    B200_CUDA_OK(cudaEventRecord(f->labels_done, f->copy[0]));
    B200_CUDA_OK(cudaStreamWaitEvent(f->copy[1], f->labels_done, 0));
  }
  FEED_TRACE("labels queued");
  // The first batches go up BEFORE the kernel is launched (the launch costs the host ~15 us,
  // during which the copy engines would otherwise sit idle); the kernel then finds its first
  // inputs in place, and the rest of the copies run ahead of it.
  const int pre = steps < pre_cfg ? steps : pre_cfg;
  for (int i = 0; i < steps; ++i) {
    if (i == pre && pre > 0) {
      FEED_TRACE("first copies queued");
      if (int e = b200_net_persist_run(&a, f->compute)) return e;
      ++f->launches;
      FEED_TRACE("kernel launched");
    }
    const unsigned g = f->fed + (unsigned)i;
    const unsigned p = g & 1u;
    // first step of a half-ring window on this stream: the slots it is about to refill must
    // have been consumed (steps below g - p - R/2)
    if (g >= R && (g - p) % half == 0u)
      CUP_OK(fn_cuStreamWaitValue32((CUstream)f->copy[p], (CUdeviceptr)(f->flags + 64), (g - p) - half,
                                    CU_STREAM_WAIT_VALUE_GEQ));
    const unsigned slot = g % R;
    B200_CUDA_OK(cudaMemcpyAsync(f->ringx + (size_t)slot * f->x_bytes, px + (size_t)i * f->x_bytes,
                                 f->x_bytes, cudaMemcpyHostToDevice, f->copy[p]));
    CUP_OK(fn_cuStreamWriteValue32((CUstream)f->copy[p], (CUdeviceptr)(f->flags + 32u * p), g + 1u,
                                   CU_STREAM_WRITE_VALUE_DEFAULT));
  }
  if (steps <= pre && pre > 0) {
    if (int e = b200_net_persist_run(&a, f->compute)) return e;
    ++f->launches;
  }
  f->fed += (unsigned)steps;
  return 0;
}

// hx: pinned images [steps*B, 784] (fp32 or uint8 per the template), hy: pinned int64 labels,
// hloss: pinned float [steps] or null. Asynchronous; caller_stream waits for the result.
int b200_feeder_run(void* h, const void* hx, const long long* hy, int steps, float* hloss,
                    void* caller_stream) {
  B200Feeder* f = static_cast<B200Feeder*>(h);
  B200_REQUIRE(f && steps > 0, "b200_feeder_run: bad arguments");
  cudaStream_t cs = (cudaStream_t)caller_stream;
  // Order after the caller's stream. When that stream is idle there is nothing to order after
  // (the previous run, which it waited for, has finished too): skip the four calls.
  g_feed_t0 = feed_now_us();
  // (only for the stream the previous run was ordered into: its trailing joins are what tells
  // us that run has finished)
  const cudaError_t busy = (f->last_cs_valid && f->last_cs == cs) ? cudaStreamQuery(cs) : cudaErrorNotReady;
  f->last_cs = cs;
  f->last_cs_valid = true;
  if (busy != cudaSuccess) {
    if (busy != cudaErrorNotReady) B200_CUDA_OK(busy);
    B200_CUDA_OK(cudaEventRecord(f->join_in, cs));
    B200_CUDA_OK(cudaStreamWaitEvent(f->compute, f->join_in, 0));
    for (int p = 0; p < 2; ++p) B200_CUDA_OK(cudaStreamWaitEvent(f->copy[p], f->join_in, 0));
  }
  float* dloss = nullptr;
  if (hloss) {
    // the kernel writes the losses into the caller's pinned buffer through its device alias
    // (looked up once per buffer: callers reuse one pinned loss buffer)
    if (hloss == f->hloss_lo) {
      dloss = f->dloss_lo;
    } else {
      const cudaError_t e = cudaHostGetDevicePointer((void**)&dloss, hloss, 0);
      if (e != cudaSuccess) {
        (void)cudaGetLastError();
        b200_set_error("b200_feeder_run: the loss buffer must be pinned (page-locked) host memory: %s",
                       cudaGetErrorString(e));
        return (int)e;
      }
      f->hloss_lo = hloss; f->dloss_lo = dloss;
    }
  }
  const char* px = static_cast<const char*>(hx);
  for (int done = 0; done < steps;) {
    const int n = steps - done < FEED_LABEL_STEPS ? steps - done : FEED_LABEL_STEPS;
    if (done > 0) {
      // the label ring holds one launch: the next launch's labels may only go up once the
      // previous kernel has finished with its own
      B200_CUDA_OK(cudaEventRecord(f->join_compute, f->compute));
      for (int p = 0; p < 2; ++p) B200_CUDA_OK(cudaStreamWaitEvent(f->copy[p], f->join_compute, 0));
    }
    if (int e = feeder_launch(f, px + (size_t)done * f->x_bytes, hy + (size_t)done * f->tmpl.B, n,
                              dloss ? dloss + done : nullptr))
      return e;
    done += n;
  }
  FEED_TRACE("all copies queued");
  B200_CUDA_OK(cudaEventRecord(f->join_compute, f->compute));
  B200_CUDA_OK(cudaStreamWaitEvent(cs, f->join_compute, 0));
  for (int p = 0; p < 2; ++p) {
    B200_CUDA_OK(cudaEventRecord(f->join_copy[p], f->copy[p]));
    B200_CUDA_OK(cudaStreamWaitEvent(cs, f->join_copy[p], 0));
  }
  FEED_TRACE("joined");
  return 0;
}

long long b200_feeder_launches(void* h) { return h ? static_cast<B200Feeder*>(h)->launches : 0; }

int b200_feeder_destroy(void* h) {
  B200Feeder* f = static_cast<B200Feeder*>(h);
  if (!f) return 0;
  cudaStreamSynchronize(f->compute);
  for (int p = 0; p < 2; ++p) cudaStreamSynchronize(f->copy[p]);
  cudaFree(f->ringx);
  cudaFree(f->ringy);
  cudaFree(f->flags);
  cudaEventDestroy(f->join_in);
  cudaEventDestroy(f->join_compute);
  cudaEventDestroy(f->labels_done);
  for (int p = 0; p < 2; ++p) {
    cudaEventDestroy(f->join_copy[p]);
    cudaStreamDestroy(f->copy[p]);
  }
  cudaStreamDestroy(f->compute);
  delete f;
  return 0;
}

}  // extern "C"
