// Communication runtime + collective kernels over NVSwitch peer memory (SURVEY §5.1,
// rows 8.10-8.14, 8.18). Replaces the reference's Gloo/TCP backend
// (src/train_dist.py:144-146; implicit broadcast at :63, implicit all-reduce at :83;
// send/recv at src/run1.py:13,16).
//
// Host side: symmetric arena allocation and cudaIpc handle export/import.
// Device side, all push-based (posted stores over NVLink, no remote-read round trips):
//   comm_barrier_kernel                 flag barrier across ranks
//   comm_broadcast_kernel               root's buffer -> every rank (DDP init, K16)
//   comm_send_kernel / comm_recv_kernel mailbox + flag point-to-point (K20)
//   net_reduce_allreduce_sgd_kernel     Net's kernel B: partial-row sum + one-shot
//                                       all-reduce + 1/world + SGD-momentum (K17+K15)
//   allreduce_kernel<T, ALGO, UPDATE>   generic bucket all-reduce, one-shot or two-shot
//                                       (reduce-scatter + all-gather), optionally fused
//                                       with the sharded SGD-momentum update
#include <cuda_bf16.h>

#include <cstring>

#include "comm.cuh"
#include "net_device.cuh"

using namespace net;

// ===========================================================================
// Host: symmetric memory
// ===========================================================================
extern "C" int b200_pdl_enabled(void);

extern "C" {

int b200_symm_alloc(size_t bytes, void** dptr, unsigned char* handle64) {
  B200_REQUIRE(bytes > 0 && dptr && handle64, "b200_symm_alloc: bad arguments");
  void* p = nullptr;
  B200_CUDA_OK(cudaMalloc(&p, bytes));
  B200_CUDA_OK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  B200_CUDA_OK(cudaDeviceSynchronize());
  *dptr = p;
  return 0;
}

int b200_symm_open(const unsigned char* handle64, void** peer_ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  B200_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *peer_ptr_out = p;
  return 0;
}

int b200_symm_close(void* peer_ptr_in) {
  B200_CUDA_OK(cudaIpcCloseMemHandle(peer_ptr_in));
  return 0;
}

int b200_symm_free(void* dptr) {
  B200_CUDA_OK(cudaFree(dptr));
  return 0;
}

int b200_device_can_access_peer(int dev, int peer, int* out) {
  B200_CUDA_OK(cudaDeviceCanAccessPeer(out, dev, peer));
  return 0;
}

}  // extern "C"

// ===========================================================================
// Device kernels
// ===========================================================================
namespace {

// ---- barrier ---------------------------------------------------------------
__global__ void comm_barrier_kernel(CommDev c, size_t flag_off, unsigned* seq_ctr) {
  const int p = threadIdx.x;
  const unsigned seq = *seq_ctr + 1u;
  __syncwarp();
  __threadfence_system();
  if (p < c.world) {
    st_release_sys_u32(peer_ptr<unsigned>(c, p, flag_off) + c.rank, seq);
    wait_flag_ge(peer_ptr<unsigned>(c, c.rank, flag_off) + p, seq, c, p, 1u);
  }
  __syncwarp();
  if (p == 0) *seq_ctr = seq;
}

// ---- broadcast (pull from root) ----------------------------------------------
// flags: [0..world) "root ready" (written by root into every rank), [world..2*world)
// "rank done" (written by every rank into root).
__global__ void __launch_bounds__(256)
comm_broadcast_kernel(CommDev c, int root, size_t buf_off, size_t nvec, size_t flag_off,
                      unsigned* seq_ctr, unsigned* arrive_ctr) {
  const unsigned seq = *seq_ctr + 1u;
  unsigned* my_flags = peer_ptr<unsigned>(c, c.rank, flag_off);
  if (c.rank == root) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      if (threadIdx.x < c.world && threadIdx.x != root)
        st_release_sys_u32(peer_ptr<unsigned>(c, threadIdx.x, flag_off) + root, seq);
      // root must not overwrite the buffer before every rank has copied it
      if (threadIdx.x < c.world && threadIdx.x != root)
        wait_flag_ge(my_flags + c.world + threadIdx.x, seq, c, threadIdx.x, 2u);
    }
  } else {
    if (threadIdx.x == 0) wait_flag_ge(my_flags + root, seq, c, root, 3u);
    __syncthreads();
    const uint4* src = peer_ptr<uint4>(c, root, buf_off);
    uint4* dst = peer_ptr<uint4>(c, c.rank, buf_off);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (size_t)gridDim.x * blockDim.x)
      dst[i] = ld_volatile_u4(src + i);
  }
  // last CTA: publish completion, advance the sequence
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(arrive_ctr, 1u);
    if (prev == gridDim.x - 1) {
      *arrive_ctr = 0u;
      if (c.rank != root)
        st_release_sys_u32(peer_ptr<unsigned>(c, root, flag_off) + c.world + c.rank, seq);
      *seq_ctr = seq;
    }
  }
}

// ---- send / recv mailbox -------------------------------------------------------
// Per (dst <- src) pair: a mailbox of `box_bytes` in dst's arena and two counters:
// flags[src] in dst's arena = chunks written so far, acks[dst] in src's arena = chunks
// consumed so far. Counters persist across calls in seq_ctr[peer] (local device memory).
__device__ __forceinline__ void send_body(const CommDev& c, int dst,
                                          const uint4* __restrict__ payload, size_t nvec,
                                          size_t box_off, size_t box_vec, size_t flag_off,
                                          size_t ack_off, unsigned* sent_ctr) {
  unsigned sent = sent_ctr[dst];
  uint4* box = peer_ptr<uint4>(c, dst, box_off) + (size_t)c.rank * box_vec;
  for (size_t done = 0; done < nvec; done += box_vec) {
    if (threadIdx.x == 0)  // wait until the receiver has drained the previous chunk
      wait_flag_ge(peer_ptr<unsigned>(c, c.rank, ack_off) + dst, sent, c, dst, 4u);
    __syncthreads();
    const size_t m = min(box_vec, nvec - done);
    for (size_t i = threadIdx.x; i < m; i += blockDim.x) box[i] = payload[done + i];
    __threadfence_system();
    __syncthreads();
    ++sent;
    if (threadIdx.x == 0)
      st_release_sys_u32(peer_ptr<unsigned>(c, dst, flag_off) + c.rank, sent);
  }
  __syncthreads();
  if (threadIdx.x == 0) sent_ctr[dst] = sent;
}

__device__ __forceinline__ void recv_body(const CommDev& c, int src, uint4* __restrict__ out,
                                          size_t nvec, size_t box_off, size_t box_vec,
                                          size_t flag_off, size_t ack_off, unsigned* recv_ctr) {
  unsigned got = recv_ctr[src];
  const uint4* box = peer_ptr<uint4>(c, c.rank, box_off) + (size_t)src * box_vec;
  for (size_t done = 0; done < nvec; done += box_vec) {
    ++got;
    if (threadIdx.x == 0)
      wait_flag_ge(peer_ptr<unsigned>(c, c.rank, flag_off) + src, got, c, src, 5u);
    __syncthreads();
    const size_t m = min(box_vec, nvec - done);
    for (size_t i = threadIdx.x; i < m; i += blockDim.x) out[done + i] = ld_volatile_u4(box + i);
    __syncthreads();
    if (threadIdx.x == 0)
      st_release_sys_u32(peer_ptr<unsigned>(c, src, ack_off) + c.rank, got);
  }
  __syncthreads();
  if (threadIdx.x == 0) recv_ctr[src] = got;
}

__global__ void __launch_bounds__(1024)
comm_send_kernel(CommDev c, int dst, const uint4* __restrict__ payload, size_t nvec,
                 size_t box_off, size_t box_vec, size_t flag_off, size_t ack_off,
                 unsigned* sent_ctr) {
  send_body(c, dst, payload, nvec, box_off, box_vec, flag_off, ack_off, sent_ctr);
}

__global__ void __launch_bounds__(1024)
comm_recv_kernel(CommDev c, int src, uint4* __restrict__ out, size_t nvec, size_t box_off,
                 size_t box_vec, size_t flag_off, size_t ack_off, unsigned* recv_ctr) {
  recv_body(c, src, out, nvec, box_off, box_vec, flag_off, ack_off, recv_ctr);
}

// A send and a receive that must make progress together (ring shifts, pairwise swaps):
// CTA 0 sends, CTA 1 receives, inside ONE launch, so both are resident by construction.
// Two separate launches on one GPU are not guaranteed to run concurrently, and a
// multi-chunk exchange would then deadlock on the mailbox acks.
__global__ void __launch_bounds__(1024)
comm_sendrecv_kernel(CommDev c, int dst, const uint4* __restrict__ payload, size_t send_vec,
                     int src, uint4* __restrict__ out, size_t recv_vec, size_t box_off,
                     size_t box_vec, size_t flag_off, size_t ack_off, unsigned* sent_ctr,
                     unsigned* recv_ctr) {
  if (blockIdx.x == 0)
    send_body(c, dst, payload, send_vec, box_off, box_vec, flag_off, ack_off, sent_ctr);
  else
    recv_body(c, src, out, recv_vec, box_off, box_vec, flag_off, ack_off, recv_ctr);
}

// ---- Net kernel B, multi-GPU ---------------------------------------------------
// CTA b owns float4 columns [32b, 32b+32) of the flat gradient. It (1) sums the G local
// partial rows, (2) pushes its 512-byte slice into every rank's staging slot
// [seq&1][my rank], (3) raises flag [b][my rank] on every rank, (4) waits for the world's
// flags for slice b, (5) sums the slices in rank order (bit-identical on every rank, so
// replicas never diverge), scales by 1/world and applies SGD-momentum to its own replica.
// No grid-wide or rank-wide barrier: CTA b only ever waits for CTA b of the other ranks.
__global__ void __launch_bounds__(256)
net_reduce_allreduce_sgd_kernel(CommDev c, const float* __restrict__ partial, int G,
                                float* __restrict__ params, float* __restrict__ mom,
                                float* __restrict__ grad_out, const float* __restrict__ hp,
                                const float* __restrict__ loss_partial,
                                float* __restrict__ loss_ring, int ring,
                                unsigned long long* ctl, size_t stage_off, size_t flag_off,
                                unsigned* seq_ctr, unsigned* arrive_ctr) {
  __shared__ float4 sm[8][32];
  constexpr int N4 = FLAT_SZ / 4;
  const int cidx = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cidx;
  const bool valid = col < N4;
  pdl_trigger();
  pdl_wait();
  const unsigned seq = *seq_ctr + 1u;  // every CTA reads it before the last one bumps it
  const unsigned slot = seq & 1u;
  const float4 s = sum_partial_rows_256(partial, G, FLAT_SZ, col, valid, sm);
  __syncthreads();
  if (rg == 0) sm[0][cidx] = s;
  __syncthreads();
  if (rg < c.world && valid) {
    float4* dst = peer_ptr<float4>(c, rg, stage_off) +
                  ((size_t)slot * c.world + c.rank) * N4 + col;
    *dst = sm[0][cidx];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int p = threadIdx.x;
    st_release_sys_u32(peer_ptr<unsigned>(c, p, flag_off) + blockIdx.x * B200_MAX_WORLD + c.rank,
                       seq);
    wait_flag_ge(peer_ptr<unsigned>(c, c.rank, flag_off) + blockIdx.x * B200_MAX_WORLD + p, seq,
                 c, p, 6u);
  }
  __syncthreads();
  if (rg == 0 && valid) {
    const float4* st = peer_ptr<float4>(c, c.rank, stage_off) + (size_t)slot * c.world * N4 + col;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < c.world; ++r) {
      const float4 v = __ldcg(st + (size_t)r * N4);
      g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
    }
    const float inv = 1.0f / (float)c.world;
    g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
    if (grad_out) reinterpret_cast<float4*>(grad_out)[col] = g;
    if (params) {
      const float lr = hp[0], mu = hp[1];
      float4 pv = reinterpret_cast<float4*>(params)[col];
      float4 bv = reinterpret_cast<float4*>(mom)[col];
      bv.x = mu * bv.x + g.x; bv.y = mu * bv.y + g.y; bv.z = mu * bv.z + g.z; bv.w = mu * bv.w + g.w;
      pv.x -= lr * bv.x; pv.y -= lr * bv.y; pv.z -= lr * bv.z; pv.w -= lr * bv.w;
      reinterpret_cast<float4*>(params)[col] = pv;
      reinterpret_cast<float4*>(mom)[col] = bv;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(arrive_ctr, 1u);
    if (prev == gridDim.x - 1) {
      *arrive_ctr = 0u;
      *seq_ctr = seq;
      float l = 0.0f;
      for (int r = 0; r < G; ++r) l += loss_partial[r];
      const unsigned long long step = ctl[0];
      if (loss_ring) {  // this rank's batch loss; slot [ring] always holds the latest
        loss_ring[step % (unsigned long long)ring] = l;
        loss_ring[ring] = l;
      }
      ctl[0] = step + 1ull;
    }
  }
}

// Low-latency variant of the kernel above: every 16-byte store carries two payload floats
// and two copies of the sequence number ({x, seq, y, seq}), so the data validates itself
// (8-byte halves are written atomically). No fence, no separate flag, no second NVLink
// traversal: the receiver polls the payload words until both tags match. This is the
// protocol choice for the 87 KB Net bucket, where the exchange is pure latency.
__global__ void __launch_bounds__(256)
net_reduce_allreduce_sgd_ll_kernel(CommDev c, const float* __restrict__ partial, int G,
                                   float* __restrict__ params, float* __restrict__ mom,
                                   float* __restrict__ grad_out, const float* __restrict__ hp,
                                   const float* __restrict__ loss_partial,
                                   float* __restrict__ loss_ring, int ring,
                                   unsigned long long* ctl, size_t stage_off,
                                   unsigned* seq_ctr, unsigned* arrive_ctr) {
  __shared__ float4 sm[8][32];
  constexpr int N4 = FLAT_SZ / 4;
  const int cidx = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cidx;
  const bool valid = col < N4;
  pdl_trigger();   // the next step's kernel A may start launching
  pdl_wait();      // partial rows come from kernel A; seq_ctr from the previous kernel B
  const unsigned seq = *seq_ctr + 1u;
  const unsigned slot = seq & 1u;
  // this rank's batch loss: one warp of block 0, fixed summation order, off the tail
  if (blockIdx.x == 0 && rg == 7) {
    float l = 0.0f;
    for (int r = cidx; r < G; r += 32) l += loss_partial[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_down_sync(0xffffffffu, l, o);
    if (cidx == 0 && loss_ring) {
      const unsigned long long step = ctl[0];   // bumped only by the last CTA, after we arrive
      loss_ring[step % (unsigned long long)ring] = l;
      loss_ring[ring] = l;
    }
  }
  const float4 s = sum_partial_rows_256(partial, G, FLAT_SZ, col, valid, sm);
  __syncthreads();
  if (rg == 0) sm[0][cidx] = s;
  __syncthreads();
  const float4 mine = sm[0][cidx];
  __syncthreads();
  if (rg < c.world && rg != c.rank && valid) {
    uint4* dst = peer_ptr<uint4>(c, rg, stage_off) +
                 (((size_t)slot * c.world + c.rank) * N4 + col) * 2;
    st_volatile_u4(dst, make_uint4(__float_as_uint(mine.x), seq, __float_as_uint(mine.y), seq));
    st_volatile_u4(dst + 1, make_uint4(__float_as_uint(mine.z), seq, __float_as_uint(mine.w), seq));
  }
  // the update's own operands do not depend on the peers: fetch them while the tags fly
  float4 pv = make_float4(0.f, 0.f, 0.f, 0.f), bv = pv;
  float lr = 0.f, mu = 0.f;
  if (rg == 0 && valid && params) {
    pv = reinterpret_cast<float4*>(params)[col];
    bv = reinterpret_cast<float4*>(mom)[col];
    lr = hp[0];
    mu = hp[1];
  }
  if (rg < c.world && valid) {
    float4 v = mine;
    if (rg != c.rank) {
      const uint4* src = peer_ptr<uint4>(c, c.rank, stage_off) +
                         (((size_t)slot * c.world + rg) * N4 + col) * 2;
      uint4 a = ld_volatile_u4(src), b = ld_volatile_u4(src + 1);
      if (!(a.y == seq && a.w == seq && b.y == seq && b.w == seq)) {
        const unsigned long long t0 = globaltimer_ns();
        unsigned polls = 0;
        for (;;) {
          a = ld_volatile_u4(src);
          b = ld_volatile_u4(src + 1);
          if (a.y == seq && a.w == seq && b.y == seq && b.w == seq) break;
          if ((++polls & 1023u) == 0u && globaltimer_ns() - t0 > c.timeout_ns) {
            if (atomicCAS(c.err, 0u, 1u) == 0u) { c.err[1] = (unsigned)rg; c.err[2] = 6u; c.err[3] = seq; }
            break;
          }
        }
      }
      v = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x),
                      __uint_as_float(b.z));
    }
    sm[rg][cidx] = v;
  }
  __syncthreads();
  if (rg == 0 && valid) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < c.world; ++r) {
      const float4 v = sm[r][cidx];
      g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
    }
    const float inv = 1.0f / (float)c.world;
    g.x *= inv; g.y *= inv; g.z *= inv; g.w *= inv;
    if (grad_out) reinterpret_cast<float4*>(grad_out)[col] = g;
    if (params) {
      bv.x = mu * bv.x + g.x; bv.y = mu * bv.y + g.y; bv.z = mu * bv.z + g.z; bv.w = mu * bv.w + g.w;
      pv.x -= lr * bv.x; pv.y -= lr * bv.y; pv.z -= lr * bv.z; pv.w -= lr * bv.w;
      reinterpret_cast<float4*>(params)[col] = pv;
      reinterpret_cast<float4*>(mom)[col] = bv;
    }
  }
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(arrive_ctr, 1u);
    if (prev == gridDim.x - 1) {
      *arrive_ctr = 0u;
      *seq_ctr = seq;
      ctl[0] = ctl[0] + 1ull;
    }
  }
}

// ---- generic all-reduce --------------------------------------------------------
struct AllreduceArgs {
  const uint4* in;        // local input (gradient bucket), nvec 16-byte vectors
  uint4* out_local;       // ALGO 0: local destination of the reduced vector (may alias in)
  size_t out_off;         // ALGO 1: arena offset of the symmetric destination buffer
  size_t nvec;
  size_t stage_off;       // arena offset of staging [2][world][src_vec]
  size_t src_vec;         // vectors reserved per source rank in one slot
  size_t flag_off;        // arena offset of flags [2][gmax][B200_MAX_WORLD]
  int gmax;
  unsigned* seq_ctr;
  unsigned* arrive_ctr;
  float scale;            // applied to the sum (1/world for the DDP mean)
  // fused SGD (UPDATE = 1): the reduced gradient never reaches memory
  float* master;          // fp32 master weights of the elements this rank updates
  float* mom;             // fp32 momentum of the same elements
  const float* hp;        // {lr, momentum, weight_decay, nesterov}
  size_t state_base;      // element index (in the bucket) that master[0]/mom[0] correspond to
  size_t push_from;       // ALGO 1: vectors below this index were already scattered to their
                          // owners by the producing GEMM's epilogue (b200_gemm_bf16_rs)
};

template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const uint4& u, float* f) {
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y);
    f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
  }
  __device__ static uint4 store(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const uint4& u, float* f) {
    const unsigned w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static uint4 store(const float* f) {
    unsigned w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const unsigned*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// Scale the rank-ordered sum, optionally apply SGD-momentum to the fp32 master weights, and
// pack the vector to publish: the averaged gradient (UPDATE 0) or the new weights.
template <typename T, int UPDATE>
__device__ __forceinline__ uint4 finish_vec(const AllreduceArgs& a, float* acc, size_t v_global) {
  constexpr int N = Vec<T>::N;
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] *= a.scale;
  if (UPDATE) {
    const float lr = a.hp[0], mu = a.hp[1], wd = a.hp[2];
    const size_t e0 = v_global * N - a.state_base;   // multiple of N: 16-byte aligned state
    float4* mw = reinterpret_cast<float4*>(a.master + e0);
    float4* mb = reinterpret_cast<float4*>(a.mom + e0);
#pragma unroll
    for (int q = 0; q < N / 4; ++q) {
      float4 w = mw[q], b = mb[q];
      float* wp = reinterpret_cast<float*>(&w);
      float* bp = reinterpret_cast<float*>(&b);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float g = acc[4 * q + i] + wd * wp[i];
        bp[i] = mu * bp[i] + g;
        wp[i] -= lr * bp[i];
        acc[4 * q + i] = wp[i];
      }
      mw[q] = w;
      mb[q] = b;
    }
  }
  return Vec<T>::store(acc);
}

// Read one tagged (LL) vector: two 16-byte words {x, seq, y, seq} {z, seq, w, seq}; spin until
// all four tags carry this call's sequence number (bounded).
__device__ __forceinline__ uint4 ll_read(const uint4* src, unsigned seq, const CommDev& c,
                                         int peer) {
  uint4 a = ld_volatile_u4(src), b = ld_volatile_u4(src + 1);
  if (!(a.y == seq && a.w == seq && b.y == seq && b.w == seq)) {
    const unsigned long long t0 = globaltimer_ns();
    unsigned polls = 0;
    for (;;) {
      a = ld_volatile_u4(src);
      b = ld_volatile_u4(src + 1);
      if (a.y == seq && a.w == seq && b.y == seq && b.w == seq) break;
      if ((++polls & 1023u) == 0u && globaltimer_ns() - t0 > c.timeout_ns) {
        if (atomicCAS(c.err, 0u, 1u) == 0u) { c.err[1] = (unsigned)peer; c.err[2] = 7u; c.err[3] = seq; }
        break;
      }
    }
  }
  return make_uint4(a.x, a.z, b.x, b.z);
}

constexpr int AR_THREADS = 512;

// ALGO 0 one-shot, low latency: every rank pushes its whole bucket to every rank as tagged
//         words (payload + sequence number in each 8 bytes: no fence, no flag, one NVLink
//         traversal) and every rank reduces everything. For small buckets.
// ALGO 1 two-shot, bandwidth: rank o owns slice o. Everyone pushes slice o to o
//         (reduce-scatter), o reduces (+ SGD on its shard of master weights/momentum) and
//         pushes the result to everyone's destination buffer (all-gather). Optimizer state
//         is sharded across ranks for free. Bulk copies keep 4 x 16 B in flight per thread.
template <typename T, int ALGO, int UPDATE>
__global__ void __launch_bounds__(AR_THREADS) allreduce_kernel(CommDev c, AllreduceArgs a) {
  constexpr int N = Vec<T>::N;
  const unsigned seq = *a.seq_ctr + 1u;
  const unsigned slot = seq & 1u;
  const int b = blockIdx.x, G = gridDim.x, W = c.world;
  const int nt = blockDim.x;
  const size_t slot_vec = (size_t)W * a.src_vec;

  if (ALGO == 0) {
    const size_t per = (a.nvec + G - 1) / G;
    const size_t lo = (size_t)b * per, hi = min(a.nvec, lo + per);
    for (size_t v = lo + threadIdx.x; v < hi; v += nt) {
      const uint4 x = a.in[v];
      const uint4 w0 = make_uint4(x.x, seq, x.y, seq), w1 = make_uint4(x.z, seq, x.w, seq);
      for (int p = 0; p < W; ++p) {
        if (p == c.rank) continue;
        uint4* dst = peer_ptr<uint4>(c, p, a.stage_off) + slot * slot_vec +
                     (size_t)c.rank * a.src_vec + 2 * v;
        st_volatile_u4(dst, w0);
        st_volatile_u4(dst + 1, w1);
      }
    }
    const uint4* st = peer_ptr<uint4>(c, c.rank, a.stage_off) + slot * slot_vec;
    for (size_t v = lo + threadIdx.x; v < hi; v += nt) {
      float acc[N];
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] = 0.0f;
      for (int r = 0; r < W; ++r) {
        const uint4 x = (r == c.rank) ? a.in[v] : ll_read(st + (size_t)r * a.src_vec + 2 * v, seq, c, r);
        float f[N];
        Vec<T>::load(x, f);
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] += f[i];
      }
      a.out_local[v] = finish_vec<T, UPDATE>(a, acc, v);
    }
  } else {
    unsigned* my_flags = peer_ptr<unsigned>(c, c.rank, a.flag_off);
    const size_t fidx1 = ((size_t)0 * a.gmax + b) * B200_MAX_WORLD;
    const size_t fidx2 = ((size_t)1 * a.gmax + b) * B200_MAX_WORLD;
    const size_t sl = (a.nvec + W - 1) / W;       // vectors per owner slice
    const size_t sub = (sl + G - 1) / G;          // vectors per CTA inside a slice
    const size_t slo = (size_t)b * sub, shi = min(sl, slo + sub);
    // reduce-scatter by pushes: my sub-slice of owner o's slice goes to o's staging
    for (int oo = 0; oo < W; ++oo) {
      const int o = (c.rank + oo) % W;            // stagger destinations across ranks
      const size_t base = (size_t)o * sl;
      const size_t lim = min(shi, a.nvec > base ? a.nvec - base : 0);
      uint4* dst = peer_ptr<uint4>(c, o, a.stage_off) + slot * slot_vec + (size_t)c.rank * a.src_vec;
      const uint4* src = a.in + base;
      // vectors with bucket index < push_from are already in place (fused GEMM epilogue)
      const size_t skip = a.push_from > base ? a.push_from - base : 0;
      size_t v0 = slo + threadIdx.x;
      if (v0 < skip) v0 += ((skip - v0 + nt - 1) / nt) * nt;
      for (size_t v = v0; v < lim; v += 4 * (size_t)nt) {
        uint4 x[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (v + (size_t)u * nt < lim) x[u] = src[v + (size_t)u * nt];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (v + (size_t)u * nt < lim) dst[v + (size_t)u * nt] = x[u];
      }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < W) {
      const int p = threadIdx.x;
      st_release_sys_u32(peer_ptr<unsigned>(c, p, a.flag_off) + fidx1 + c.rank, seq);
      wait_flag_ge(my_flags + fidx1 + p, seq, c, p, 8u);
    }
    __syncthreads();
    // reduce my slice (+ SGD on my shard), all-gather by pushes
    const uint4* st = peer_ptr<uint4>(c, c.rank, a.stage_off) + slot * slot_vec;
    const size_t mybase = (size_t)c.rank * sl;
    const size_t mylim = min(shi, a.nvec > mybase ? a.nvec - mybase : 0);
    for (size_t v = slo + threadIdx.x; v < mylim; v += 2 * (size_t)nt) {
      uint4 res[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const size_t vv = v + (size_t)u * nt;
        if (vv < mylim) {
          float acc[N];
#pragma unroll
          for (int i = 0; i < N; ++i) acc[i] = 0.0f;
          for (int r = 0; r < W; ++r) {
            float f[N];
            Vec<T>::load(__ldcg(st + (size_t)r * a.src_vec + vv), f);
#pragma unroll
            for (int i = 0; i < N; ++i) acc[i] += f[i];
          }
          res[u] = finish_vec<T, UPDATE>(a, acc, mybase + vv);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const size_t vv = v + (size_t)u * nt;
        if (vv < mylim)
          for (int pp = 0; pp < W; ++pp) {
            const int p = (c.rank + pp) % W;
            peer_ptr<uint4>(c, p, a.out_off)[mybase + vv] = res[u];
          }
      }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < W) {
      const int p = threadIdx.x;
      st_release_sys_u32(peer_ptr<unsigned>(c, p, a.flag_off) + fidx2 + c.rank, seq);
      wait_flag_ge(my_flags + fidx2 + p, seq, c, p, 9u);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(a.arrive_ctr, 1u);
    if (prev == gridDim.x - 1) {
      *a.arrive_ctr = 0u;
      *a.seq_ctr = seq;
    }
  }
}

}  // namespace

// ===========================================================================
// C ABI: launches
// ===========================================================================
extern "C" {

struct B200CommHost {  // mirrors CommDev; filled by Python through b200_comm_make
  CommDev dev;
};

int b200_comm_struct_size(void) { return (int)sizeof(CommDev); }

// Fill a caller-provided CommDev blob (host memory, b200_comm_struct_size() bytes).
int b200_comm_make(void* blob, int rank, int world, void* const* bases, unsigned* err,
                   unsigned long long timeout_ns) {
  B200_REQUIRE(world >= 1 && world <= B200_MAX_WORLD, "b200_comm_make: world %d not in [1,%d]",
               world, B200_MAX_WORLD);
  B200_REQUIRE(rank >= 0 && rank < world, "b200_comm_make: bad rank %d", rank);
  CommDev c;
  memset(&c, 0, sizeof(c));
  c.rank = rank;
  c.world = world;
  for (int i = 0; i < world; ++i) c.base[i] = (char*)bases[i];
  c.err = err;
  c.timeout_ns = timeout_ns;
  memcpy(blob, &c, sizeof(c));
  return 0;
}

int b200_comm_barrier(const void* comm, size_t flag_off, unsigned* seq_ctr, void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  comm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(c, flag_off, seq_ctr);
  B200_LAUNCH_OK();
  return 0;
}

int b200_comm_broadcast(const void* comm, int root, size_t buf_off, size_t bytes, size_t flag_off,
                        unsigned* seq_ctr, unsigned* arrive_ctr, void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(bytes % 16 == 0 && buf_off % 16 == 0, "b200_comm_broadcast: 16-byte alignment");
  B200_REQUIRE(root >= 0 && root < c.world, "b200_comm_broadcast: bad root");
  const size_t nvec = bytes / 16;
  int blocks = (int)((nvec + 255) / 256);
  if (blocks > 148 * 2) blocks = 148 * 2;
  if (blocks < 1) blocks = 1;
  comm_broadcast_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(c, root, buf_off, nvec, flag_off,
                                                                  seq_ctr, arrive_ctr);
  B200_LAUNCH_OK();
  return 0;
}

int b200_comm_send(const void* comm, int dst, const void* payload, size_t bytes, size_t box_off,
                   size_t box_bytes, size_t flag_off, size_t ack_off, unsigned* sent_ctr,
                   void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(dst >= 0 && dst < c.world && dst != c.rank, "b200_comm_send: bad destination %d", dst);
  B200_REQUIRE(bytes > 0 && bytes % 16 == 0 && box_bytes % 16 == 0, "b200_comm_send: 16-byte granularity");
  comm_send_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(c, dst, (const uint4*)payload, bytes / 16,
                                                         box_off, box_bytes / 16, flag_off, ack_off,
                                                         sent_ctr);
  B200_LAUNCH_OK();
  return 0;
}

int b200_comm_recv(const void* comm, int src, void* out, size_t bytes, size_t box_off,
                   size_t box_bytes, size_t flag_off, size_t ack_off, unsigned* recv_ctr,
                   void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(src >= 0 && src < c.world && src != c.rank, "b200_comm_recv: bad source %d", src);
  B200_REQUIRE(bytes > 0 && bytes % 16 == 0 && box_bytes % 16 == 0, "b200_comm_recv: 16-byte granularity");
  comm_recv_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(c, src, (uint4*)out, bytes / 16, box_off,
                                                         box_bytes / 16, flag_off, ack_off,
                                                         recv_ctr);
  B200_LAUNCH_OK();
  return 0;
}

int b200_comm_sendrecv(const void* comm, int dst, const void* payload, size_t send_bytes, int src,
                       void* out, size_t recv_bytes, size_t box_off, size_t box_bytes,
                       size_t flag_off, size_t ack_off, unsigned* sent_ctr, unsigned* recv_ctr,
                       void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(dst >= 0 && dst < c.world && dst != c.rank, "b200_comm_sendrecv: bad destination");
  B200_REQUIRE(src >= 0 && src < c.world && src != c.rank, "b200_comm_sendrecv: bad source");
  B200_REQUIRE(send_bytes > 0 && recv_bytes > 0 && send_bytes % 16 == 0 && recv_bytes % 16 == 0 &&
                   box_bytes % 16 == 0,
               "b200_comm_sendrecv: 16-byte granularity");
  comm_sendrecv_kernel<<<2, 1024, 0, (cudaStream_t)stream>>>(
      c, dst, (const uint4*)payload, send_bytes / 16, src, (uint4*)out, recv_bytes / 16, box_off,
      box_bytes / 16, flag_off, ack_off, sent_ctr, recv_ctr);
  B200_LAUNCH_OK();
  return 0;
}

// Bytes of staging and flags the Net kernel B needs in the arena.
size_t b200_comm_net_stage_bytes(void) { return (size_t)2 * B200_MAX_WORLD * FLAT_SZ * 4 * 2; }
size_t b200_comm_net_flag_bytes(void) {
  return (size_t)((FLAT_SZ / 4 + 31) / 32) * B200_MAX_WORLD * 4;
}

int b200_comm_net_reduce_allreduce_sgd(const void* comm, const float* partial, int rows,
                                       float* params, float* mom, float* grad_out,
                                       const float* hp, const float* loss_partial,
                                       float* loss_ring, int ring, unsigned long long* ctl,
                                       size_t stage_off, size_t flag_off, unsigned* seq_ctr,
                                       unsigned* arrive_ctr, int protocol, void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(rows > 0 && rows <= 148, "b200_comm_net_reduce_allreduce_sgd: rows out of range");
  const int blocks = (FLAT_SZ / 4 + 31) / 32;
  const bool pdl = b200_pdl_enabled() != 0;
  if (protocol == 0)   // low latency: tagged payload words
    B200_CUDA_OK(b200_launch_pdl(net_reduce_allreduce_sgd_ll_kernel, dim3(blocks), dim3(256), 0,
                                 (cudaStream_t)stream, pdl, c, partial, rows, params, mom, grad_out,
                                 hp, loss_partial, loss_ring, ring, ctl, stage_off, seq_ctr,
                                 arrive_ctr));
  else                 // payload, system fence, then a flag per CTA
    B200_CUDA_OK(b200_launch_pdl(net_reduce_allreduce_sgd_kernel, dim3(blocks), dim3(256), 0,
                                 (cudaStream_t)stream, pdl, c, partial, rows, params, mom, grad_out,
                                 hp, loss_partial, loss_ring, ring, ctl, stage_off, flag_off,
                                 seq_ctr, arrive_ctr));
  return 0;
}

// dtype: 0 fp32, 1 bf16. algo: 0 one-shot, 1 two-shot. update: 0 plain (mean/sum into the
// destination), 1 fused SGD-momentum (destination receives the new weights).
int b200_comm_allreduce(const void* comm, int dtype, int algo, int update, const void* in,
                        void* out_local, size_t out_off, size_t bytes, size_t stage_off,
                        size_t src_bytes, size_t flag_off, int gmax, unsigned* seq_ctr,
                        unsigned* arrive_ctr, float scale, float* master, float* mom,
                        const float* hp, size_t state_base, int blocks, size_t push_from,
                        void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(bytes > 0 && bytes % 16 == 0, "b200_comm_allreduce: size must be a multiple of 16 B");
  B200_REQUIRE(push_from == 0 || algo == 1, "b200_comm_allreduce: push_from needs the two-shot algorithm");
  B200_REQUIRE(src_bytes % 16 == 0 && stage_off % 16 == 0 && out_off % 16 == 0,
               "b200_comm_allreduce: 16-byte alignment");
  AllreduceArgs a;
  a.in = (const uint4*)in;
  a.out_local = (uint4*)out_local;
  a.out_off = out_off;
  a.nvec = bytes / 16;
  a.stage_off = stage_off;
  a.src_vec = src_bytes / 16;
  a.flag_off = flag_off;
  a.gmax = gmax;
  a.seq_ctr = seq_ctr;
  a.arrive_ctr = arrive_ctr;
  a.scale = scale;
  a.master = master;
  a.mom = mom;
  a.hp = hp;
  a.state_base = state_base;
  a.push_from = push_from;
  // one-shot stages tagged words (2x the payload); two-shot stages one slice per source
  const size_t need = (algo == 0) ? 2 * a.nvec : (a.nvec + c.world - 1) / c.world;
  B200_REQUIRE(need <= a.src_vec, "b200_comm_allreduce: %zu B per source exceed the staging (%zu B)",
               need * 16, src_bytes);
  if (blocks <= 0) {
    // one-shot is latency bound: spread it wide (one 16-byte vector per thread, up to 64 CTAs);
    // two-shot moves bulk data: ~32 KB per CTA
    const size_t work = (algo == 0) ? a.nvec : (a.nvec + c.world - 1) / c.world;
    blocks = (algo == 0) ? (int)((work + AR_THREADS - 1) / AR_THREADS) : (int)((work + 2047) / 2048);
    const int cap = (algo == 0) ? 64 : gmax;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
  }
  B200_REQUIRE(blocks <= gmax, "b200_comm_allreduce: %d CTAs exceed the flag table (%d)", blocks, gmax);
  B200_REQUIRE(!update || (master && mom && hp), "b200_comm_allreduce: SGD fusion needs state");
  cudaStream_t s = (cudaStream_t)stream;
#define B200_AR(T, A, U) allreduce_kernel<T, A, U><<<blocks, AR_THREADS, 0, s>>>(c, a)
  if (dtype == 0) {
    if (algo == 0) { if (update) B200_AR(float, 0, 1); else B200_AR(float, 0, 0); }
    else           { if (update) B200_AR(float, 1, 1); else B200_AR(float, 1, 0); }
  } else if (dtype == 1) {
    if (algo == 0) { if (update) B200_AR(__nv_bfloat16, 0, 1); else B200_AR(__nv_bfloat16, 0, 0); }
    else           { if (update) B200_AR(__nv_bfloat16, 1, 1); else B200_AR(__nv_bfloat16, 1, 0); }
  } else {
    B200_REQUIRE(false, "b200_comm_allreduce: unknown dtype %d", dtype);
  }
#undef B200_AR
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
