// sm_100a kernels for the MNIST `Net` (reference: src/model.py, src/train.py:69-104).
//
// Two families, both built from the per-image stages in net_device.cuh:
//
//  * Layered kernels, one launch per layer and direction.
//    They back the autograd path: `Net.forward`, `loss.backward()`.
//  * The fused step kernels (rows 8.5-8.9, 8.15, 8.22): `net_fwdbwd_kernel` runs
//    forward + loss + backward for one image per CTA iteration entirely out of
//    shared memory and emits per-CTA gradient partial rows; `net_reduce_sgd_kernel`
//    sums the rows in fixed order and applies SGD-momentum. Two launches per
//    optimizer step, no zero_grad pass, no host sync; the loss goes to a device ring.
//
// Weight-gradient reductions over the batch are deterministic: every CTA owns a
// private partial row in scratch memory and the rows are summed in row order.
#include "net_device.cuh"

using namespace net;

extern "C" int b200_pdl_enabled(void);

// ---------------------------------------------------------------------------
// Built-in scratch for b200_conv1_bwd_noalloc (see the entry point at the end of this
// file). It lives in the library image: every GPU that loads the library carries these
// bytes until the process exits. Layout: [0, 256) arrival counter, then one 264-float
// partial row per image of the batch.
// ---------------------------------------------------------------------------
#define B200_NOALLOC_MAX_BATCH 128
#define B200_NOALLOC_ROW_FLOATS 264
#define B200_NOALLOC_SCRATCH_BYTES (256 + B200_NOALLOC_MAX_BATCH * B200_NOALLOC_ROW_FLOATS * 4)
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// Optional per-phase timestamps of the fused kernel (debug aid; null in production).
__device__ long long* g_net_phase_dbg = nullptr;
#define B200_PHASE(i)                                                            \
  do {                                                                           \
    if (dbg && threadIdx.x == 0) dbg[blockIdx.x * 16 + (i)] = clock64();         \
  } while (0)

namespace {

__device__ __forceinline__ float normalize_u8(unsigned v) {
  // torchvision ToTensor + Normalize((0.1307,), (0.3081,)) (src/train_dist.py:16-19)
  return ((float)v / 255.0f - 0.1307f) / 0.3081f;
}

// x_mode 0: fp32 [B,784] already normalised; 1: u8 [B,784]; 2: u8 dataset [N,784] gathered
// by idx[img]; 3: fp32 dataset [N,784] gathered by idx[img].
__device__ __forceinline__ void load_image(float* __restrict__ xs, const void* __restrict__ x,
                                           int x_mode, int img, const int* __restrict__ gidx) {
  const int t = threadIdx.x;
  if (x_mode == 0 || x_mode == 3) {
    const size_t row = (x_mode == 3) ? (size_t)gidx[img] : (size_t)img;
    const float4* src = reinterpret_cast<const float4*>((const float*)x + row * IMG_SZ);
    if (t < IMG_SZ / 4) reinterpret_cast<float4*>(xs)[t] = __ldg(src + t);
  } else {
    const size_t row = (x_mode == 2) ? (size_t)gidx[img] : (size_t)img;
    const uchar4* src = reinterpret_cast<const uchar4*>((const uint8_t*)x + row * IMG_SZ);
    if (t < IMG_SZ / 4) {
      const uchar4 v = __ldg(src + t);
      reinterpret_cast<float4*>(xs)[t] =
          make_float4(normalize_u8(v.x), normalize_u8(v.y), normalize_u8(v.z), normalize_u8(v.w));
    }
  }
}

// Last-arriving CTA sums the G partial rows in row order into `out` (n floats).
// Returns true in every thread of the CTA that does the reduction.
__device__ __forceinline__ bool last_block_arrive(unsigned* counter) {
  __shared__ unsigned s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(counter, 1u);
    s_last = (prev == gridDim.x - 1) ? 1u : 0u;
    if (s_last) *counter = 0u;  // self-reset: graph replays need no memset
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0u;
}

__device__ __forceinline__ void reduce_rows(const float* __restrict__ scratch, int G, int stride,
                                            int n, float* __restrict__ out) {
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    float s = 0.0f;
    for (int r = 0; r < G; ++r) s += __ldcg(scratch + (size_t)r * stride + e);
    out[e] = s;
  }
}

// ---------------------------------------------------------------------------
// 8.1 gather + u8 -> fp32 normalise (reference: src/train_dist.py:16-19, DataLoader collate)
// ---------------------------------------------------------------------------
__global__ void gather_normalize_kernel(const uint8_t* __restrict__ data,
                                        const int* __restrict__ idx, float* __restrict__ out,
                                        int B) {
  const int img = blockIdx.x;
  if (img >= B) return;
  const size_t row = idx ? (size_t)idx[img] : (size_t)img;
  const uchar4* src = reinterpret_cast<const uchar4*>(data + row * IMG_SZ);
  float4* dst = reinterpret_cast<float4*>(out + (size_t)img * IMG_SZ);
  for (int t = threadIdx.x; t < IMG_SZ / 4; t += blockDim.x) {
    const uchar4 v = __ldg(src + t);
    dst[t] = make_float4(normalize_u8(v.x), normalize_u8(v.y), normalize_u8(v.z), normalize_u8(v.w));
  }
}

// ---------------------------------------------------------------------------
// 8.2 conv1 + bias + pool + relu, forward / 8.8 backward (wgrad only: the input needs no grad)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(NT)
conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                 const float* __restrict__ b, float* __restrict__ a1,
                 uint16_t* __restrict__ pos1, int B) {
  __shared__ __align__(16) float xs[IMG_SZ];
  __shared__ __align__(16) float a1s[A1_SZ];
  __shared__ float w1s[W1_SZ], b1s[B1_SZ];
  __shared__ uint16_t pos1s[A1_SZ];
  const int t = threadIdx.x;
  if (t < W1_SZ) w1s[t] = w[t];
  if (t < B1_SZ) b1s[t] = b[t];
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    load_image(xs, x, 0, img, nullptr);
    __syncthreads();
    conv1_fwd_img(xs, w1s, b1s, a1s, pos1s);
    __syncthreads();
    for (int i = t; i < A1_SZ; i += NT) {
      a1[(size_t)img * A1_SZ + i] = a1s[i];
      pos1[(size_t)img * A1_SZ + i] = pos1s[i];
    }
  }
}

__global__ void __launch_bounds__(NT)
conv1_bwd_kernel(const float* __restrict__ x, const float* __restrict__ da1,
                 const float* __restrict__ a1, const uint16_t* __restrict__ pos1,
                 float* __restrict__ dw, float* __restrict__ db, float* __restrict__ scratch,
                 unsigned* __restrict__ counter, int B) {
  __shared__ __align__(16) float xs[IMG_SZ];
  __shared__ __align__(16) float g1s[A1_SZ];
  __shared__ uint16_t pos1s[A1_SZ];
  constexpr int ROW = 264;  // 252 (w, padded) + 12 (b, padded)
  float* row = scratch + (size_t)blockIdx.x * ROW;
  const int t = threadIdx.x;
  int it = 0;
  for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
    load_image(xs, x, 0, img, nullptr);
    for (int i = t; i < A1_SZ; i += NT) {
      const size_t gi = (size_t)img * A1_SZ + i;
      g1s[i] = (a1[gi] > 0.0f) ? da1[gi] : 0.0f;
      pos1s[i] = pos1[gi];
    }
    __syncthreads();
    conv1_bwd_img(g1s, pos1s, xs, row, row + 252, it == 0);
    __syncthreads();
  }
  if (last_block_arrive(counter)) {
    const int G = gridDim.x;
    for (int e = t; e < W1_SZ + B1_SZ; e += NT) {
      const int src = (e < W1_SZ) ? e : 252 + (e - W1_SZ);
      float s = 0.0f;
      for (int r = 0; r < G; ++r) s += __ldcg(scratch + (size_t)r * ROW + src);
      if (e < W1_SZ) dw[e] = s; else db[e - W1_SZ] = s;
    }
  }
}

// ---------------------------------------------------------------------------
// 8.3 conv2 + bias + dropout2d + pool + relu, forward / 8.7 backward
// ---------------------------------------------------------------------------
// mask_mode 0: no dropout (mask = 1); 1: generate Philox mask and store it in `mask`;
// 2: use the caller's `mask` (tests inject the oracle's mask this way).
__global__ void __launch_bounds__(NT)
conv2_fwd_kernel(const float* __restrict__ a1, const float* __restrict__ w,
                 const float* __restrict__ b, float* __restrict__ mask, int mask_mode, float p,
                 unsigned long long seed, unsigned long long step, float* __restrict__ a2,
                 uint8_t* __restrict__ pos2, int B) {
  extern __shared__ __align__(16) float smem[];
  float* a1s = smem;              // 1440
  float* w2s = a1s + A1_SZ;       // 5000
  float* a2s = w2s + W2_SZ;       // 320
  float* b2s = a2s + A2_SZ;       // 20
  float* m2s = b2s + 20;          // 20 (+12 pad)
  uint8_t* pos2s = reinterpret_cast<uint8_t*>(m2s + 32);  // 320 B
  const int t = threadIdx.x;
  for (int i = t; i < W2_SZ; i += NT) w2s[i] = w[i];
  if (t < B2_SZ) b2s[t] = b[t];
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    for (int i = t; i < A1_SZ / 4; i += NT)
      reinterpret_cast<float4*>(a1s)[i] =
          __ldg(reinterpret_cast<const float4*>(a1 + (size_t)img * A1_SZ) + i);
    if (t < C2) {
      float m = 1.0f;
      if (mask_mode == 1) {
        m = dropout_mask_value(seed, step, (uint32_t)img, 0x20000u + t, p);
        mask[(size_t)img * C2 + t] = m;
      } else if (mask_mode == 2) {
        m = mask[(size_t)img * C2 + t];
      }
      m2s[t] = m;
    }
    __syncthreads();
    conv2_fwd_img(a1s, w2s, b2s, m2s, a2s, pos2s);
    __syncthreads();
    a2[(size_t)img * A2_SZ + t] = a2s[t];
    pos2[(size_t)img * A2_SZ + t] = pos2s[t];
  }
}

__global__ void __launch_bounds__(NT)
conv2_bwd_kernel(const float* __restrict__ a1, const float* __restrict__ w,
                 const float* __restrict__ da2, const float* __restrict__ a2,
                 const uint8_t* __restrict__ pos2, const float* __restrict__ mask,
                 float* __restrict__ dw, float* __restrict__ db, float* __restrict__ da1,
                 float* __restrict__ scratch, unsigned* __restrict__ counter, int B) {
  extern __shared__ __align__(16) float smem[];
  float* a1s = smem;              // 1440
  float* w2s = a1s + A1_SZ;       // 5000
  float* g1s = w2s + W2_SZ;       // 1440
  float* g2s = g1s + A1_SZ;       // 320
  uint8_t* pos2s = reinterpret_cast<uint8_t*>(g2s + A2_SZ);  // 320 B
  constexpr int ROW = 5020;
  float* row = scratch + (size_t)blockIdx.x * ROW;
  const int t = threadIdx.x;
  for (int i = t; i < W2_SZ; i += NT) w2s[i] = w[i];
  int it = 0;
  for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
    for (int i = t; i < A1_SZ / 4; i += NT)
      reinterpret_cast<float4*>(a1s)[i] =
          __ldg(reinterpret_cast<const float4*>(a1 + (size_t)img * A1_SZ) + i);
    {
      const size_t gi = (size_t)img * A2_SZ + t;
      const float m = mask ? mask[(size_t)img * C2 + (t >> 4)] : 1.0f;
      g2s[t] = (a2[gi] > 0.0f) ? da2[gi] * m : 0.0f;
      pos2s[t] = pos2[gi];
    }
    __syncthreads();
    conv2_bwd_img<false>(g2s, pos2s, a1s, w2s, g1s, row, row + W2_SZ, it == 0, da1 != nullptr);
    __syncthreads();
    if (da1)
      for (int i = t; i < A1_SZ; i += NT) da1[(size_t)img * A1_SZ + i] = g1s[i];
  }
  if (last_block_arrive(counter)) {
    reduce_rows(scratch, gridDim.x, ROW, W2_SZ, dw);
    reduce_rows(scratch + W2_SZ, gridDim.x, ROW, B2_SZ, db);
  }
}

// ---------------------------------------------------------------------------
// 8.4 fc1 + relu + dropout forward / 8.6 backward
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(NT)
fc1_fwd_kernel(const float* __restrict__ a2, const float* __restrict__ W3,
               const float* __restrict__ b3, float* __restrict__ mask, int mask_mode, float p,
               unsigned long long seed, unsigned long long step, float* __restrict__ h1, int B) {
  __shared__ __align__(16) float a2s[A2_SZ];
  __shared__ float m3s[64], h1s[64];
  const int t = threadIdx.x;
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    a2s[t] = a2[(size_t)img * A2_SZ + t];
    if (t < FC1) {
      float m = 1.0f;
      if (mask_mode == 1) {
        m = dropout_mask_value(seed, step, (uint32_t)img, 0x30000u + t, p);
        mask[(size_t)img * FC1 + t] = m;
      } else if (mask_mode == 2) {
        m = mask[(size_t)img * FC1 + t];
      }
      m3s[t] = m;
    }
    __syncthreads();
    fc1_fwd_img<false>(a2s, W3, b3, m3s, h1s);
    __syncthreads();
    if (t < FC1) h1[(size_t)img * FC1 + t] = h1s[t];
  }
}

__global__ void __launch_bounds__(NT)
fc1_bwd_kernel(const float* __restrict__ a2, const float* __restrict__ W3,
               const float* __restrict__ dh1, const float* __restrict__ h1,
               const float* __restrict__ mask, float* __restrict__ dW, float* __restrict__ db,
               float* __restrict__ da2, float* __restrict__ scratch,
               unsigned* __restrict__ counter, int B) {
  __shared__ __align__(16) float a2s[A2_SZ];
  __shared__ float dpre_s[64];
  constexpr int ROW = W3_SZ + 52;
  float* row = scratch + (size_t)blockIdx.x * ROW;
  const int t = threadIdx.x;
  int it = 0;
  for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
    a2s[t] = a2[(size_t)img * A2_SZ + t];
    if (t < FC1) {
      const size_t gi = (size_t)img * FC1 + t;
      const float m = mask ? mask[gi] : 1.0f;
      dpre_s[t] = (h1[gi] > 0.0f) ? dh1[gi] * m : 0.0f;
    }
    __syncthreads();
    // autograd semantics: da2 is the plain gradient w.r.t. fc1's input; conv2's own
    // backward applies its ReLU and channel mask.
    {
      const float a = a2s[t];
      float da = 0.0f;
#pragma unroll 10
      for (int j = 0; j < FC1; ++j) {
        const float d = dpre_s[j];
        gacc(row + j * A2_SZ + t, d * a, it == 0);
        da = fmaf(d, __ldg(W3 + j * A2_SZ + t), da);
      }
      if (da2) da2[(size_t)img * A2_SZ + t] = da;
      if (t < FC1) gacc(row + W3_SZ + t, dpre_s[t], it == 0);
    }
    __syncthreads();
  }
  if (last_block_arrive(counter)) {
    reduce_rows(scratch, gridDim.x, ROW, W3_SZ, dW);
    reduce_rows(scratch + W3_SZ, gridDim.x, ROW, B3_SZ, db);
  }
}

// ---------------------------------------------------------------------------
// fc2 + log_softmax forward/backward, NLL loss forward/backward (8.5 in layered form)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(NT)
fc2_lsm_fwd_kernel(const float* __restrict__ h1, const float* __restrict__ W4,
                   const float* __restrict__ b4, float* __restrict__ logp, int B) {
  __shared__ float w4s[W4_SZ], b4s[16], h1s[64], logit_s[16];
  const int t = threadIdx.x;
  for (int i = t; i < W4_SZ; i += NT) w4s[i] = W4[i];
  if (t < B4_SZ) b4s[t] = b4[t];
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    if (t < FC1) h1s[t] = h1[(size_t)img * FC1 + t];
    __syncthreads();
    fc2_fwd_img(h1s, w4s, b4s, logit_s);
    __syncthreads();
    if (t == 0) {
      float lp[NCLS];
      log_softmax10(logit_s, lp);
#pragma unroll
      for (int c = 0; c < NCLS; ++c) logp[(size_t)img * NCLS + c] = lp[c];
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(NT)
fc2_lsm_bwd_kernel(const float* __restrict__ h1, const float* __restrict__ W4,
                   const float* __restrict__ logp, const float* __restrict__ dlogp,
                   float* __restrict__ dW, float* __restrict__ db, float* __restrict__ dh1,
                   float* __restrict__ scratch, unsigned* __restrict__ counter, int B) {
  __shared__ float w4s[W4_SZ], h1s[64], dlog_s[16];
  constexpr int ROW = 512;
  float* row = scratch + (size_t)blockIdx.x * ROW;
  const int t = threadIdx.x;
  for (int i = t; i < W4_SZ; i += NT) w4s[i] = W4[i];
  int it = 0;
  for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
    if (t < FC1) h1s[t] = h1[(size_t)img * FC1 + t];
    if (t == 0) {
      float s = 0.0f;
#pragma unroll
      for (int c = 0; c < NCLS; ++c) s += dlogp[(size_t)img * NCLS + c];
#pragma unroll
      for (int c = 0; c < NCLS; ++c)
        dlog_s[c] = dlogp[(size_t)img * NCLS + c] - expf(logp[(size_t)img * NCLS + c]) * s;
    }
    __syncthreads();
    // dW4, db4 partials; dh1 is the plain gradient w.r.t. fc2's input here (the
    // dropout/ReLU of fc1 are applied by fc1's own backward in the layered path).
    for (int o = t; o < W4_SZ; o += NT) {
      const int c = o / FC1, j = o - c * FC1;
      gacc(row + o, dlog_s[c] * h1s[j], it == 0);
    }
    if (t < NCLS) gacc(row + W4_SZ + t, dlog_s[t], it == 0);
    if (t >= 64 && t < 64 + FC1) {
      const int j = t - 64;
      float s = 0.0f;
#pragma unroll
      for (int c = 0; c < NCLS; ++c) s = fmaf(dlog_s[c], w4s[c * FC1 + j], s);
      if (dh1) dh1[(size_t)img * FC1 + j] = s;
    }
    __syncthreads();
  }
  if (last_block_arrive(counter)) {
    reduce_rows(scratch, gridDim.x, ROW, W4_SZ, dW);
    reduce_rows(scratch + W4_SZ, gridDim.x, ROW, B4_SZ, db);
  }
}

// reduction: 0 = mean, 1 = sum (F.nll_loss(..., size_average=False), src/train.py:94)
__global__ void nll_fwd_kernel(const float* __restrict__ logp, const long long* __restrict__ target,
                               float* __restrict__ loss, int B, int reduction) {
  __shared__ float red[256];
  float s = 0.0f;
  for (int i = threadIdx.x; i < B; i += 256) s -= logp[(size_t)i * NCLS + (int)target[i]];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = (reduction == 0) ? red[0] / (float)B : red[0];
}

__global__ void nll_bwd_kernel(const float* __restrict__ gloss, const long long* __restrict__ target,
                               float* __restrict__ dlogp, int B, int reduction) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * NCLS) return;
  const int b = i / NCLS, c = i - b * NCLS;
  const float g = *gloss * ((reduction == 0) ? 1.0f / (float)B : 1.0f);
  dlogp[i] = (c == (int)target[b]) ? -g : 0.0f;
}

// ---------------------------------------------------------------------------
// 8.9 SGD with momentum over a flat buffer (reference: optim.SGD, src/train.py:60-61,76)
// hp = {lr, momentum, weight_decay, nesterov(0/1)} lives in device memory so a
// captured CUDA graph picks up a changed learning rate without re-capture.
// ---------------------------------------------------------------------------
__global__ void sgd_momentum_kernel(float* __restrict__ p, const float* __restrict__ g,
                                    float* __restrict__ buf, int n4, const float* __restrict__ hp,
                                    float grad_scale) {
  const float lr = hp[0], mu = hp[1], wd = hp[2];
  const bool nesterov = hp[3] != 0.0f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 bv = reinterpret_cast<float4*>(buf)[i];
#define B200_SGD1(c)                                       \
    {                                                      \
      float gg = gv.c * grad_scale + wd * pv.c;            \
      bv.c = mu * bv.c + gg;                               \
      pv.c -= lr * (nesterov ? gg + mu * bv.c : bv.c);     \
    }
    B200_SGD1(x) B200_SGD1(y) B200_SGD1(z) B200_SGD1(w)
#undef B200_SGD1
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(buf)[i] = bv;
  }
}

// ---------------------------------------------------------------------------
// Fused whole-step kernel A: forward + loss + backward, one image per CTA iteration.
// ---------------------------------------------------------------------------
struct FusedSmem {
  float* xs; float* a1s; float* a2s; float* h1s; float* m3s; float* m2s; float* logit_s;
  float* dlog_s; float* dpre_s; float* g2s; float* g1s; float* w1s; float* b1s; float* w2s;
  float* b2s; float* w4s; float* b4s; uint16_t* pos1s; uint8_t* pos2s;
};
constexpr int FUSED_SMEM_FLOATS = IMG_SZ + A1_SZ + A2_SZ + 64 + 64 + 32 + 16 + 16 + 64 + A2_SZ +
                                  A1_SZ + 252 + 12 + W2_SZ + 20 + W4_SZ + 12 + A1_SZ / 2 +
                                  A2_SZ / 4;
constexpr int FUSED_LIST_OFF = FUSED_SMEM_FLOATS * 4;               // D[64][20] floats
constexpr int FUSED_T_OFF = FUSED_LIST_OFF + 64 * C2 * 4 + 256;      // + live[21], slot_of[20]; then T[64][250]
constexpr int FUSED_W3_OFF = FUSED_T_OFF + 64 * 250 * 4;             // fc1.weight staged in smem
constexpr int FUSED_SMEM_BYTES = FUSED_W3_OFF + W3_SZ * 4;           // ~181 KB: one CTA per SM
constexpr int EVAL_SMEM_BYTES = FUSED_SMEM_FLOATS * 4;
static_assert(FUSED_T_OFF % 16 == 0 && FUSED_W3_OFF % 16 == 0, "smem carve alignment");

__device__ __forceinline__ FusedSmem carve(float* s) {
  FusedSmem f;
  f.xs = s; s += IMG_SZ;
  f.a1s = s; s += A1_SZ;
  f.a2s = s; s += A2_SZ;
  f.g1s = s; s += A1_SZ;
  f.g2s = s; s += A2_SZ;
  f.w2s = s; s += W2_SZ;
  f.w4s = s; s += W4_SZ;
  f.w1s = s; s += 252;
  f.h1s = s; s += 64;
  f.m3s = s; s += 64;
  f.dpre_s = s; s += 64;
  f.m2s = s; s += 32;
  f.b2s = s; s += 20;
  f.logit_s = s; s += 16;
  f.dlog_s = s; s += 16;
  f.b1s = s; s += 12;
  f.b4s = s; s += 12;
  f.pos1s = reinterpret_cast<uint16_t*>(s); s += A1_SZ / 2;
  f.pos2s = reinterpret_cast<uint8_t*>(s);
  return f;
}

__device__ __forceinline__ void load_weights(const FusedSmem& f, const float* __restrict__ P) {
  const int t = threadIdx.x;
  for (int i = t; i < W2_SZ / 4; i += NT)
    reinterpret_cast<float4*>(f.w2s)[i] = __ldg(reinterpret_cast<const float4*>(P + OFF_W2) + i);
  for (int i = t; i < W4_SZ / 4; i += NT)
    reinterpret_cast<float4*>(f.w4s)[i] = __ldg(reinterpret_cast<const float4*>(P + OFF_W4) + i);
  if (t < W1_SZ) f.w1s[t] = P[OFF_W1 + t];
  if (t < B1_SZ) f.b1s[t] = P[OFF_B1 + t];
  if (t < B2_SZ) f.b2s[t] = P[OFF_B2 + t];
  if (t < B4_SZ) f.b4s[t] = P[OFF_B4 + t];
}

// params: flat buffer in the net:: layout. partial: [gridDim.x][FLAT_SZ]. loss_partial: [gridDim.x].
// labels: int64 [B] (x_mode 0/1) or int64 dataset labels gathered by gidx (x_mode 2/3).
// ctl[0] = optimizer step counter (advanced by kernel B), ctl[1] = step at which the current
// epoch started. With idx_stride != 0 the batch's indices are gidx[(step - ctl[1]) * idx_stride ..],
// so a whole epoch over a device-resident dataset replays one captured graph with no host input.
__global__ void __launch_bounds__(NT, 1)
net_fwdbwd_kernel(const float* __restrict__ params, const void* __restrict__ x, int x_mode,
                  const long long* __restrict__ labels, const int* __restrict__ gidx, int B,
                  float inv_B, float* __restrict__ partial, float* __restrict__ loss_partial,
                  int mask_mode, float* __restrict__ m2buf, float* __restrict__ m3buf,
                  unsigned long long seed, const unsigned long long* __restrict__ step_ptr,
                  unsigned sample_base, int idx_stride, int split, int n_idx) {
  extern __shared__ __align__(16) float smem[];
  const FusedSmem f = carve(smem);
  float* Ds = smem + FUSED_SMEM_FLOATS;
  int* live = reinterpret_cast<int*>(Ds + 64 * C2);   // [21]: live channels, [20] = count
  int* slot_of = live + 32;                           // [20]
  // split = 1: two CTAs per image. Both run the (cheap) forward; CTA 2i takes the fc and
  // conv2 weight gradients, CTA 2i+1 takes conv2's data gradient and conv1's gradients.
  // They write disjoint parts of the same partial row and never talk to each other, so a
  // batch of 64 keeps 128 SMs busy instead of 64.
  const int role = split ? (int)(blockIdx.x & 1) : 2;        // 0: fc+conv2-w, 1: dgrad+conv1, 2: all
  const int slot = split ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int nslots = split ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const bool do_w = role != 1, do_d = role != 0;
  long long* dbg = g_net_phase_dbg;
  pdl_trigger();   // kernel B of this step may start launching now
  pdl_wait();      // ... and this kernel may not read parameters before the previous B is done
  B200_PHASE(0);
  float* Ts = reinterpret_cast<float*>(reinterpret_cast<char*>(smem) + FUSED_T_OFF);
  float* w3s = reinterpret_cast<float*>(reinterpret_cast<char*>(smem) + FUSED_W3_OFF);
  const int t = threadIdx.x;
  // All weights stream into shared memory with cp.async, in the order they are needed, so the
  // kernel pays ONE global-memory round trip before conv1 instead of a chain of them:
  //   group 0: conv1 weights/bias      (needed by conv1_fwd)
  //   group 1: conv2 weights/bias      (needed by conv2_fwd)
  //   group 2: fc2 weights/bias + fc1.weight (64 KB)   (needed from fc1_fwd on)
  auto cp16 = [](float* dst, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(
                     (unsigned)__cvta_generic_to_shared(dst)),
                 "l"(src));
  };
  if (t < 63) cp16(f.w1s + 4 * t, params + OFF_W1 + 4 * t);
  if (t >= 64 && t < 67) cp16(f.b1s + 4 * (t - 64), params + OFF_B1 + 4 * (t - 64));
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int i = t; i < W2_SZ / 4; i += NT) cp16(f.w2s + 4 * i, params + OFF_W2 + 4 * i);
  if (t < 5) cp16(f.b2s + 4 * t, params + OFF_B2 + 4 * t);
  asm volatile("cp.async.commit_group;" ::: "memory");
  if (t < W4_SZ / 4) cp16(f.w4s + 4 * t, params + OFF_W4 + 4 * t);
  if (t >= 128 && t < 131) cp16(f.b4s + 4 * (t - 128), params + OFF_B4 + 4 * (t - 128));
  for (int i = t; i < W3_SZ / 4; i += NT) cp16(w3s + 4 * i, params + OFF_W3 + 4 * i);
  asm volatile("cp.async.commit_group;" ::: "memory");
  const unsigned long long step = step_ptr ? step_ptr[0] : 0ull;
  // indices of THIS step start at idx_base; n_idx bounds the prefetch of the next step's rows
  const long long idx_base =
      (idx_stride != 0 && step_ptr) ? (long long)(step - step_ptr[1]) * (long long)idx_stride : 0ll;
  gidx += idx_base;
  float* row = partial + (size_t)slot * FLAT_SZ;
  float loss_acc = 0.0f;
  int it = 0;
  for (int img = slot; img < B; img += nslots, ++it) {
    const bool first = (it == 0);
    load_image(f.xs, x, x_mode, img, gidx);
    // the label is needed only after fc2, but its (dependent) global loads start now
    int tgt = 0;
    if (t < 32) tgt = (int)labels[(x_mode >= 2) ? gidx[img] : img];
    if (t < C2) {
      float m = 1.0f;
      if (mask_mode == 1) {
        m = dropout_mask_value(seed, step, sample_base + img, 0x20000u + t, 0.5f);
        if (m2buf) m2buf[(size_t)img * C2 + t] = m;
      } else if (mask_mode == 2) {
        m = m2buf[(size_t)img * C2 + t];
      }
      f.m2s[t] = m;
    } else if (t >= 32 && t < 32 + FC1) {
      const int j = t - 32;
      float m = 1.0f;
      if (mask_mode == 1) {
        m = dropout_mask_value(seed, step, sample_base + img, 0x30000u + j, 0.5f);
        if (m3buf) m3buf[(size_t)img * FC1 + j] = m;
      } else if (mask_mode == 2) {
        m = m3buf[(size_t)img * FC1 + j];
      }
      f.m3s[j] = m;
    }
    if (first) asm volatile("cp.async.wait_group 2;" ::: "memory");
    __syncthreads();
    B200_PHASE(1);
    conv1_fwd_img(f.xs, f.w1s, f.b1s, f.a1s, f.pos1s);
    if (first) asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    B200_PHASE(2);
    conv2_fwd_img(f.a1s, f.w2s, f.b2s, f.m2s, f.a2s, f.pos2s);
    if (first) asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    B200_PHASE(3);
    fc1_fwd_img<true>(f.a2s, w3s, params + OFF_B3, f.m3s, f.h1s);
    __syncthreads();
    B200_PHASE(4);
    fc2_fwd_img(f.h1s, f.w4s, f.b4s, f.logit_s);
    __syncthreads();
    if (t < 32) {
      // fused log_softmax + NLL + its gradient (src/model.py:22 + src/train.py:74), one class
      // per lane: max and sum by warp shuffles instead of a 20-long serial exp/log chain
      const float z = (t < NCLS) ? f.logit_s[t] : -INFINITY;
      const float mx = warp_max(z);
      const float e = (t < NCLS) ? expf(z - mx) : 0.0f;
      const float lse = mx + logf(warp_sum(e));
      const float lp = z - lse;
      if (t < NCLS) f.dlog_s[t] = (expf(lp) - (t == tgt ? 1.0f : 0.0f)) * inv_B;
      const float lt = __shfl_sync(0xffffffffu, lp, tgt);
      loss_acc -= lt * inv_B;       // every lane of warp 0 tracks it; lane 0 publishes
    }
    __syncthreads();
    B200_PHASE(5);
    fc2_bwd_img(f.dlog_s, f.h1s, f.m3s, f.w4s, f.dpre_s, row + OFF_W4, row + OFF_B4, first, do_w);
    __syncthreads();
    B200_PHASE(6);
    fc1_bwd_img<true>(f.dpre_s, f.a2s, f.m2s, w3s, f.g2s, row + OFF_W3, row + OFF_B3, first, do_w);
    __syncthreads();
    B200_PHASE(7);
    if (do_d) build_conv2_dense(f.g2s, f.pos2s, f.m2s, Ds, live, slot_of);   // uniform per CTA
    __syncthreads();
    B200_PHASE(8);
    if (do_w) conv2_wgrad_img(f.g2s, f.pos2s, f.m2s, f.a1s, row + OFF_W2, row + OFF_B2, first);
    if (do_d) conv2_dgrad_T_img(Ds, live, live[C2], f.w2s, Ts);
    __syncthreads();
    B200_PHASE(9);
    if (do_d) {
      conv2_dgrad_gather_img(Ts, f.a1s, f.g1s);
      __syncthreads();
      B200_PHASE(10);
      conv1_wgrad_img(f.g1s, f.pos1s, f.xs, row + OFF_W1, row + OFF_B1, first);
      __syncthreads();
      B200_PHASE(12);
    }
  }
  // Next step's batch is known already (device-side index stepping): pull its rows from HBM
  // into L2 now, off the critical path of the next launch.
  if (idx_stride != 0 && x_mode >= 2 && role != 1) {
    const int* nidx = gidx + idx_stride;
    const size_t row_bytes = (x_mode == 3) ? IMG_SZ * 4 : IMG_SZ;
    const int lines = (int)((row_bytes + 127) / 128);
    // never read an index beyond the end of the permutation (last step of an epoch)
    const long long left = (long long)n_idx - (idx_base + idx_stride);
    for (int img = slot; img < B && img < left; img += nslots)
      if (t < lines) {
        const char* p = (const char*)x + (size_t)nidx[img] * row_bytes + (size_t)t * 128;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
      }
  }
  B200_PHASE(13);
  if (role != 1) {
    if (t == 0) loss_partial[slot] = loss_acc;
    // zero the 8 pad floats of the row so the flat reduction never reads garbage
    if (t < 2) row[OFF_W1 + W1_SZ + t] = 0.0f;
    if (t >= 2 && t < 4) row[OFF_B1 + B1_SZ + (t - 2)] = 0.0f;
    if (t >= 4 && t < 6) row[OFF_B3 + B3_SZ + (t - 4)] = 0.0f;
    if (t >= 6 && t < 8) row[OFF_B4 + B4_SZ + (t - 6)] = 0.0f;
  }
}

// Fused whole-step kernel B (single GPU): fixed-order sum of the partial rows, then
// SGD-momentum in place (K14+K15: no zero_grad pass, gradients never hit .grad unless asked).
// The multi-GPU version with the all-reduce in the middle lives in comm_kernels.cu.
__global__ void __launch_bounds__(256)
net_reduce_sgd_kernel(const float* __restrict__ partial, int G, float* __restrict__ params,
                      float* __restrict__ mom, float* __restrict__ grad_out,
                      const float* __restrict__ hp, const float* __restrict__ loss_partial,
                      float* __restrict__ loss_ring, int ring, unsigned long long* step_ptr) {
  __shared__ float4 sm[8][32];
  constexpr int N4 = FLAT_SZ / 4;
  const int i = blockIdx.x * 32 + (threadIdx.x & 31);
  pdl_trigger();   // the next step's kernel A may start launching
  pdl_wait();      // partial rows come from kernel A
  const float4 s = sum_partial_rows_256(partial, G, FLAT_SZ, i, i < N4, sm);
  if (threadIdx.x < 32 && i < N4) {
    if (grad_out) reinterpret_cast<float4*>(grad_out)[i] = s;
    if (params) {
      const float lr = hp[0], mu = hp[1];
      float4 pv = reinterpret_cast<float4*>(params)[i];
      float4 bv = reinterpret_cast<float4*>(mom)[i];
      bv.x = mu * bv.x + s.x; bv.y = mu * bv.y + s.y; bv.z = mu * bv.z + s.z; bv.w = mu * bv.w + s.w;
      pv.x -= lr * bv.x; pv.y -= lr * bv.y; pv.z -= lr * bv.z; pv.w -= lr * bv.w;
      reinterpret_cast<float4*>(params)[i] = pv;
      reinterpret_cast<float4*>(mom)[i] = bv;
    }
  }
  if (blockIdx.x == 0 && (threadIdx.x >> 5) == 7) {   // one spare warp: loss + step counter
    const int lane = threadIdx.x & 31;
    float l = 0.0f;
    for (int r = lane; r < G; r += 32) l += loss_partial[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_down_sync(0xffffffffu, l, o);
    if (lane == 0) {
      const unsigned long long step = step_ptr ? *step_ptr : 0ull;
      if (loss_ring) {  // slot [ring] always holds the latest step's loss
        loss_ring[step % (unsigned long long)ring] = l;
        loss_ring[ring] = l;
      }
      if (step_ptr) *step_ptr = step + 1ull;
    }
  }
}

// ---------------------------------------------------------------------------
// 8.15 evaluation: forward (dropout compiled out) + per-image loss / arg-max,
// then a single-block fixed-order reduction to (loss_sum, correct).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 1)
net_eval_kernel(const float* __restrict__ params, const void* __restrict__ x, int x_mode,
                const long long* __restrict__ labels, const int* __restrict__ gidx, int B,
                float* __restrict__ logp_out, float* __restrict__ loss_img,
                int* __restrict__ pred_img) {
  extern __shared__ __align__(16) float smem[];
  const FusedSmem f = carve(smem);
  const int t = threadIdx.x;
  load_weights(f, params);
  if (t < 32) f.m2s[t] = 1.0f;
  if (t < 64) f.m3s[t] = 1.0f;
  for (int img = blockIdx.x; img < B; img += gridDim.x) {
    load_image(f.xs, x, x_mode, img, gidx);
    __syncthreads();
    conv1_fwd_img(f.xs, f.w1s, f.b1s, f.a1s, f.pos1s);
    __syncthreads();
    conv2_fwd_img(f.a1s, f.w2s, f.b2s, f.m2s, f.a2s, f.pos2s);
    __syncthreads();
    fc1_fwd_img<false>(f.a2s, params + OFF_W3, params + OFF_B3, f.m3s, f.h1s);
    __syncthreads();
    fc2_fwd_img(f.h1s, f.w4s, f.b4s, f.logit_s);
    __syncthreads();
    if (t == 0) {
      float lp[NCLS];
      log_softmax10(f.logit_s, lp);
      int best = 0;
#pragma unroll
      for (int c = 1; c < NCLS; ++c) if (lp[c] > lp[best]) best = c;
      if (logp_out)
#pragma unroll
        for (int c = 0; c < NCLS; ++c) logp_out[(size_t)img * NCLS + c] = lp[c];
      if (labels) {
        const int row_lab = (x_mode >= 2) ? gidx[img] : img;
        const int tgt = (int)labels[row_lab];
        float lt = 0.0f;
#pragma unroll
        for (int c = 0; c < NCLS; ++c) if (c == tgt) lt = lp[c];
        loss_img[img] = -lt;
        pred_img[img] = (best == tgt) ? 1 : 0;
      }
    }
    __syncthreads();
  }
}

// out[0] (+)= sum of per-image NLL, out[1] (+)= number correct (as float; exact below 2^24)
__global__ void eval_reduce_kernel(const float* __restrict__ loss_img,
                                   const int* __restrict__ pred_img, int B,
                                   float* __restrict__ out, int accumulate) {
  __shared__ float red[256];
  __shared__ int redc[256];
  float s = 0.0f; int c = 0;
  for (int i = threadIdx.x; i < B; i += 256) { s += loss_img[i]; c += pred_img[i]; }
  red[threadIdx.x] = s; redc[threadIdx.x] = c;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { red[threadIdx.x] += red[threadIdx.x + o]; redc[threadIdx.x] += redc[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (accumulate) { out[0] += red[0]; out[1] += (float)redc[0]; }
    else { out[0] = red[0]; out[1] = (float)redc[0]; }
  }
}

int grid_for(int B, int cap) { return B < cap ? B : cap; }

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

int b200_net_layout(int* offsets9) {
  const int o[9] = {OFF_W1, OFF_B1, OFF_W2, OFF_B2, OFF_W3, OFF_B3, OFF_W4, OFF_B4, FLAT_SZ};
  for (int i = 0; i < 9; ++i) offsets9[i] = o[i];
  return 0;
}

// Floats of scratch each layered backward needs for `rows` CTAs, largest layer (fc1).
int b200_net_scratch_floats(int rows) { return rows * (W3_SZ + 52); }
int b200_net_max_rows(void) { return 148; }

int b200_gather_normalize(const void* data_u8, const int* idx, float* out, int B, void* stream) {
  if (B <= 0) return 0;
  gather_normalize_kernel<<<B, 224, 0, (cudaStream_t)stream>>>((const uint8_t*)data_u8, idx, out, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_conv1_fwd(const float* x, const float* w, const float* b, float* a1, void* pos1, int B,
                   void* stream) {
  if (B <= 0) return 0;
  conv1_fwd_kernel<<<grid_for(B, 148 * 4), NT, 0, (cudaStream_t)stream>>>(x, w, b, a1,
                                                                          (uint16_t*)pos1, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_conv1_bwd(const float* x, const float* da1, const float* a1, const void* pos1, float* dw,
                   float* db, float* scratch, unsigned* counter, int B, void* stream) {
  B200_REQUIRE(B > 0, "b200_conv1_bwd: empty batch");
  conv1_bwd_kernel<<<grid_for(B, 148), NT, 0, (cudaStream_t)stream>>>(
      x, da1, a1, (const uint16_t*)pos1, dw, db, scratch, counter, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_conv2_fwd(const float* a1, const float* w, const float* b, float* mask, int mask_mode,
                   float p, unsigned long long seed, unsigned long long step, float* a2,
                   void* pos2, int B, void* stream) {
  if (B <= 0) return 0;
  B200_REQUIRE(mask_mode == 0 || mask != nullptr, "b200_conv2_fwd: mask buffer required");
  const int smem = (A1_SZ + W2_SZ + A2_SZ + 20 + 32) * 4 + A2_SZ;
  conv2_fwd_kernel<<<grid_for(B, 148 * 4), NT, smem, (cudaStream_t)stream>>>(
      a1, w, b, mask, mask_mode, p, seed, step, a2, (uint8_t*)pos2, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_conv2_bwd(const float* a1, const float* w, const float* da2, const float* a2,
                   const void* pos2, const float* mask, float* dw, float* db, float* da1,
                   float* scratch, unsigned* counter, int B, void* stream) {
  B200_REQUIRE(B > 0, "b200_conv2_bwd: empty batch");
  const int smem = (A1_SZ + W2_SZ + A1_SZ + A2_SZ) * 4 + A2_SZ;
  conv2_bwd_kernel<<<grid_for(B, 148), NT, smem, (cudaStream_t)stream>>>(
      a1, w, da2, a2, (const uint8_t*)pos2, mask, dw, db, da1, scratch, counter, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_fc1_fwd(const float* a2, const float* W, const float* b, float* mask, int mask_mode,
                 float p, unsigned long long seed, unsigned long long step, float* h1, int B,
                 void* stream) {
  if (B <= 0) return 0;
  B200_REQUIRE(mask_mode == 0 || mask != nullptr, "b200_fc1_fwd: mask buffer required");
  fc1_fwd_kernel<<<grid_for(B, 148 * 4), NT, 0, (cudaStream_t)stream>>>(a2, W, b, mask, mask_mode,
                                                                        p, seed, step, h1, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_fc1_bwd(const float* a2, const float* W, const float* dh1, const float* h1,
                 const float* mask, float* dW, float* db, float* da2, float* scratch,
                 unsigned* counter, int B, void* stream) {
  B200_REQUIRE(B > 0, "b200_fc1_bwd: empty batch");
  fc1_bwd_kernel<<<grid_for(B, 148), NT, 0, (cudaStream_t)stream>>>(a2, W, dh1, h1, mask, dW, db,
                                                                    da2, scratch, counter, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_fc2_lsm_fwd(const float* h1, const float* W, const float* b, float* logp, int B,
                     void* stream) {
  if (B <= 0) return 0;
  fc2_lsm_fwd_kernel<<<grid_for(B, 148 * 4), NT, 0, (cudaStream_t)stream>>>(h1, W, b, logp, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_fc2_lsm_bwd(const float* h1, const float* W, const float* logp, const float* dlogp,
                     float* dW, float* db, float* dh1, float* scratch, unsigned* counter, int B,
                     void* stream) {
  B200_REQUIRE(B > 0, "b200_fc2_lsm_bwd: empty batch");
  fc2_lsm_bwd_kernel<<<grid_for(B, 148), NT, 0, (cudaStream_t)stream>>>(
      h1, W, logp, dlogp, dW, db, dh1, scratch, counter, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_nll_fwd(const float* logp, const long long* target, float* loss, int B, int reduction,
                 void* stream) {
  B200_REQUIRE(B > 0, "b200_nll_fwd: empty batch");
  nll_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(logp, target, loss, B, reduction);
  B200_LAUNCH_OK();
  return 0;
}

int b200_nll_bwd(const float* gloss, const long long* target, float* dlogp, int B, int reduction,
                 void* stream) {
  B200_REQUIRE(B > 0, "b200_nll_bwd: empty batch");
  nll_bwd_kernel<<<(B * NCLS + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gloss, target, dlogp, B,
                                                                          reduction);
  B200_LAUNCH_OK();
  return 0;
}

int b200_sgd_momentum(float* p, const float* g, float* buf, long long n, const float* hp,
                      float grad_scale, void* stream) {
  B200_REQUIRE(n % 4 == 0, "b200_sgd_momentum: n must be a multiple of 4 (flat buffers are padded)");
  const int n4 = (int)(n / 4);
  int blocks = (n4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  sgd_momentum_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(p, g, buf, n4, hp, grad_scale);
  B200_LAUNCH_OK();
  return 0;
}

// Debug: per-CTA, per-phase clock64() stamps of the next fused kernel launches go to `buf`
// (device, >= grid*16 int64); pass null to switch off.
int b200_net_set_phase_debug(long long* buf) {
  B200_CUDA_OK(cudaMemcpyToSymbol(g_net_phase_dbg, &buf, sizeof(buf)));
  return 0;
}

// PDL on the fused step's kernels (B200DP_PDL=0 switches it off).
int b200_pdl_enabled(void) {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200DP_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v;
}

// Rows (CTAs) the fused step uses for a batch of B images.
int b200_net_fused_rows(int B) { return grid_for(B, 148); }

static int ensure_fused_attr() {
  static bool done_dev[64] = {false};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  bool& done = done_dev[dev & 63];
  if (!done) {
    B200_CUDA_OK(cudaFuncSetAttribute(net_fwdbwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      FUSED_SMEM_BYTES));
    B200_CUDA_OK(cudaFuncSetAttribute(net_eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      EVAL_SMEM_BYTES));
    done = true;
  }
  return 0;
}

// Kernel A of the fused step. partial: [rows][FLAT_SZ] floats, loss_partial: [rows].
int b200_net_fwdbwd(const float* params, const void* x, int x_mode, const long long* labels,
                    const int* gidx, int B, float* partial, float* loss_partial, int mask_mode,
                    float* m2buf, float* m3buf, unsigned long long seed,
                    const unsigned long long* step_ptr, unsigned sample_base, int idx_stride,
                    int n_idx, void* stream) {
  B200_REQUIRE(B > 0, "b200_net_fwdbwd: empty batch");
  B200_REQUIRE(idx_stride == 0 || n_idx > 0,
               "b200_net_fwdbwd: device-side index stepping needs the length of the index buffer");
  B200_REQUIRE(mask_mode != 2 || (m2buf && m3buf), "b200_net_fwdbwd: mask_mode 2 needs both masks");
  B200_REQUIRE(x_mode < 2 || gidx != nullptr, "b200_net_fwdbwd: gather modes need an index buffer");
  if (int e = ensure_fused_attr()) return e;
  // when the batch leaves at least half of the 148 SMs idle, give every image two CTAs
  const int split = (2 * B <= 148) ? 1 : 0;
  const int grid = split ? 2 * B : grid_for(B, 148);
  B200_CUDA_OK(b200_launch_pdl(net_fwdbwd_kernel, dim3(grid), dim3(NT), FUSED_SMEM_BYTES,
                               (cudaStream_t)stream, b200_pdl_enabled(), params, x, x_mode, labels,
                               gidx, B, 1.0f / (float)B, partial, loss_partial, mask_mode, m2buf,
                               m3buf, seed, step_ptr, sample_base, idx_stride, split, n_idx));
  return 0;
}

// Kernel B of the fused step on one GPU. params/mom may be null to only materialise grad_out.
int b200_net_reduce_sgd(const float* partial, int rows, float* params, float* mom, float* grad_out,
                        const float* hp, const float* loss_partial, float* loss_ring, int ring,
                        unsigned long long* step_ptr, void* stream) {
  B200_REQUIRE(rows > 0 && rows <= 148, "b200_net_reduce_sgd: rows out of range");
  const int blocks = (FLAT_SZ / 4 + 31) / 32;
  B200_CUDA_OK(b200_launch_pdl(net_reduce_sgd_kernel, dim3(blocks), dim3(256), 0,
                               (cudaStream_t)stream, b200_pdl_enabled(), partial, rows, params, mom,
                               grad_out, hp, loss_partial, loss_ring, ring, step_ptr));
  return 0;
}

int b200_net_eval(const float* params, const void* x, int x_mode, const long long* labels,
                  const int* gidx, int B, float* logp_out, float* loss_img, int* pred_img,
                  float* out2, int accumulate, void* stream) {
  if (B <= 0) return 0;
  if (int e = ensure_fused_attr()) return e;
  B200_REQUIRE(!labels || (loss_img && pred_img), "b200_net_eval: labels need loss/pred buffers");
  net_eval_kernel<<<grid_for(B, 148 * 4), NT, EVAL_SMEM_BYTES, (cudaStream_t)stream>>>(
      params, x, x_mode, labels, gidx, B, logp_out, loss_img, pred_img);
  B200_LAUNCH_OK();
  if (labels && out2) {
    eval_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(loss_img, pred_img, B, out2, accumulate);
    B200_LAUNCH_OK();
  }
  return 0;
}

}  // extern "C"


// ===========================================================================
// Entry point that allocates nothing
// ===========================================================================
#include <mutex>

namespace {
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_dev_ptr[64] = {nullptr};  // per device: address of g_noalloc_scratch
}  // namespace

extern "C" {

int b200_conv1_bwd_noalloc_max_batch(void) { return B200_NOALLOC_MAX_BATCH; }
size_t b200_conv1_bwd_noalloc_scratch_bytes(void) { return (size_t)B200_NOALLOC_SCRATCH_BYTES; }

// conv1 weight/bias gradient with NO workspace argument and no allocation of any kind.
//
// Who needs this: code that may not call cudaMalloc (or anything that allocates behind the
// scenes, torch's caching allocator included) at the point where it needs the gradient:
// a host thread whose memory budget was fixed up front, a plugin that is handed its
// buffers and a stream and nothing else, or a latency-critical loop where an allocation
// could synchronise the device. b200_conv1_bwd asks such callers for a scratch buffer and
// a zeroed counter they have no way to obtain; this variant keeps both inside the library
// in a fixed array (g_noalloc_scratch, 135,424 bytes per GPU) instead.
//
// Same inputs, outputs and arithmetic as b200_conv1_bwd; the caller still passes x, da1,
// a1, pos1 and receives dw [10,1,5,5] and db [10].
//
// Limits and rules
//  * B <= 128 images (b200_conv1_bwd_noalloc_max_batch()); larger batches are refused with
//    an error, use b200_conv1_bwd with your own scratch or split the batch.
//  * The FIRST call on each device looks up the array's device address
//    (cudaGetSymbolAddress, with that device current). Make that first call outside any
//    stream capture; later calls only enqueue a 4-byte memset and the kernel.
//  * Every call on a device shares the one array, which still holds the previous call's
//    partial rows. The arrival counter is cleared at the start of each call; the rows need
//    no clearing because each call overwrites every row it later reads. Two calls that
//    overlap on the GPU (different streams, no dependency between them) race on the array
//    and both results are garbage: serialise them.
int b200_conv1_bwd_noalloc(const float* x, const float* da1, const float* a1, const void* pos1,
                           float* dw, float* db, int B, void* stream) {
  B200_REQUIRE(B > 0, "b200_conv1_bwd_noalloc: empty batch");
  B200_REQUIRE(B <= B200_NOALLOC_MAX_BATCH,
               "b200_conv1_bwd_noalloc: batch %d exceeds the built-in scratch (max %d images); "
               "use b200_conv1_bwd with caller-provided scratch or split the batch",
               B, B200_NOALLOC_MAX_BATCH);
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "b200_conv1_bwd_noalloc: device index %d out of range", dev);
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_dev_ptr[dev] == nullptr) {
      // In host code `g_noalloc_scratch` is only a handle for the symbol; the address the
      // kernel can use has to be asked for, once per device, with that device current.
      unsigned char* p = nullptr;
      B200_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
      B200_REQUIRE(p != nullptr, "b200_conv1_bwd_noalloc: symbol address lookup returned null");
      g_noalloc_dev_ptr[dev] = p;
    }
    base = g_noalloc_dev_ptr[dev];
  }
  unsigned* counter = reinterpret_cast<unsigned*>(base);
  float* scratch = reinterpret_cast<float*>(base + 256);
  cudaStream_t s = (cudaStream_t)stream;
  B200_CUDA_OK(cudaMemsetAsync(counter, 0, sizeof(unsigned), s));
  conv1_bwd_kernel<<<grid_for(B, 148), NT, 0, s>>>(x, da1, a1, (const uint16_t*)pos1, dw, db,
                                                   scratch, counter, B);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
