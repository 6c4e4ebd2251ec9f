// Per-image device stages of the MNIST `Net` (reference: src/model.py:9-22).
//
// Every stage works on ONE image whose activations live in shared memory, with
// NT threads of one CTA. The layered kernels (one launch per layer, used by the
// autograd path) and the fused whole-step kernel (forward + loss + backward in
// one launch) are both thin drivers around these functions, so the two paths
// cannot drift apart numerically.
//
// Backward uses the sparsity that max-pool creates: only the arg-max position of
// each 2x2 window carries gradient, so conv wgrad/dgrad iterate over pooled
// windows (16 per conv2 channel, 144 per conv1 channel) instead of over the
// dense 8x8 / 24x24 maps. That is 4x fewer MACs than the dense formulation and
// needs no zero-filled dconv buffer at all.
#pragma once
#include "common.cuh"

namespace net {

constexpr int NT = 320;  // threads per CTA for every Net kernel (10 warps)

constexpr int IMG = 28, IMG_SZ = 784;
constexpr int C1 = 10, P1 = 12, A1_SZ = C1 * P1 * P1;  // 1440
constexpr int C2 = 20, P2 = 4, A2_SZ = C2 * P2 * P2;   // 320
constexpr int FC1 = 50, NCLS = 10;
constexpr int W1_SZ = 250, B1_SZ = 10, W2_SZ = 5000, B2_SZ = 20;
constexpr int W3_SZ = 16000, B3_SZ = 50, W4_SZ = 500, B4_SZ = 10;

// Flat parameter / gradient layout: state_dict order, every tensor start
// aligned to 4 floats (16 B) so the flat buffer can be moved as float4.
constexpr int OFF_W1 = 0;
constexpr int OFF_B1 = 252;
constexpr int OFF_W2 = 264;
constexpr int OFF_B2 = 5264;
constexpr int OFF_W3 = 5284;
constexpr int OFF_B3 = 21284;
constexpr int OFF_W4 = 21336;
constexpr int OFF_B4 = 21836;
constexpr int FLAT_SZ = 21848;  // 87,392 B (21,840 real parameters + 8 pad floats)

// Accumulate one gradient contribution into this CTA's private partial row.
__device__ __forceinline__ void gacc(float* p, float v, bool first) {
  if (first) *p = v; else *p += v;
}

// ---- forward ---------------------------------------------------------------

// conv1 (1->10, 5x5 valid) + bias + 2x2 max-pool + ReLU (src/model.py:16).
// xs[784] -> a1s[10*144] (post-ReLU), pos1s[10*144] = Y*28+X of the window arg-max.
__device__ __forceinline__ void conv1_fwd_img(const float* __restrict__ xs,
                                              const float* __restrict__ w1s,
                                              const float* __restrict__ b1s,
                                              float* __restrict__ a1s,
                                              uint16_t* __restrict__ pos1s) {
  const int t = threadIdx.x;
  if (t < 288) {
    const int half = t / 144, p = t - half * 144;
    const int py = p / 12, px = p - py * 12;
    float patch[36];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c) patch[r * 6 + c] = xs[(2 * py + r) * IMG + 2 * px + c];
#pragma unroll 1
    for (int cc = 0; cc < 5; ++cc) {
      const int ch = half * 5 + cc;
      const float* w = w1s + ch * 25;
      const float b = b1s[ch];
      float s00 = b, s01 = b, s10 = b, s11 = b;
#pragma unroll
      for (int ky = 0; ky < 5; ++ky)
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
          const float wv = w[ky * 5 + kx];
          s00 = fmaf(wv, patch[ky * 6 + kx], s00);
          s01 = fmaf(wv, patch[ky * 6 + kx + 1], s01);
          s10 = fmaf(wv, patch[(ky + 1) * 6 + kx], s10);
          s11 = fmaf(wv, patch[(ky + 1) * 6 + kx + 1], s11);
        }
      // first-max tie-breaking in scan order, like ATen's max_pool2d
      float m = s00; int dy = 0, dx = 0;
      if (s01 > m) { m = s01; dy = 0; dx = 1; }
      if (s10 > m) { m = s10; dy = 1; dx = 0; }
      if (s11 > m) { m = s11; dy = 1; dx = 1; }
      a1s[ch * 144 + p] = fmaxf(m, 0.0f);
      pos1s[ch * 144 + p] = (uint16_t)((2 * py + dy) * IMG + 2 * px + dx);
    }
  }
}

// conv2 (10->20, 5x5 valid) + bias + Dropout2d channel mask + 2x2 max-pool + ReLU
// (src/model.py:17). a1s[10*144] -> a2s[20*16], pos2s[20*16] = Y*12+X.
// m2s[20] holds 0 or 1/(1-p) per channel (all 1 in eval).
__device__ __forceinline__ void conv2_fwd_img(const float* __restrict__ a1s,
                                              const float* __restrict__ w2s,
                                              const float* __restrict__ b2s,
                                              const float* __restrict__ m2s,
                                              float* __restrict__ a2s,
                                              uint8_t* __restrict__ pos2s) {
  const int t = threadIdx.x;  // exactly 320 work items
  const int co = t >> 4, p = t & 15;
  const int py = p >> 2, px = p & 3;
  const float b = b2s[co];
  float s00 = b, s01 = b, s10 = b, s11 = b;
#pragma unroll 2
  for (int ci = 0; ci < C1; ++ci) {
    float patch[36];
    const float* src = a1s + ci * 144 + (2 * py) * 12 + 2 * px;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; c += 2) {
        const float2 v = *reinterpret_cast<const float2*>(src + r * 12 + c);
        patch[r * 6 + c] = v.x;
        patch[r * 6 + c + 1] = v.y;
      }
    const float* w = w2s + (co * C1 + ci) * 25;
#pragma unroll
    for (int ky = 0; ky < 5; ++ky)
#pragma unroll
      for (int kx = 0; kx < 5; ++kx) {
        const float wv = w[ky * 5 + kx];
        s00 = fmaf(wv, patch[ky * 6 + kx], s00);
        s01 = fmaf(wv, patch[ky * 6 + kx + 1], s01);
        s10 = fmaf(wv, patch[(ky + 1) * 6 + kx], s10);
        s11 = fmaf(wv, patch[(ky + 1) * 6 + kx + 1], s11);
      }
  }
  // Dropout2d comes BEFORE the pool in the reference. A positive scale does not
  // move the arg-max, and a dropped channel is all zeros (gradient is zero
  // whatever index is recorded), so pool the raw map and scale afterwards.
  float m = s00; int dy = 0, dx = 0;
  if (s01 > m) { m = s01; dy = 0; dx = 1; }
  if (s10 > m) { m = s10; dy = 1; dx = 0; }
  if (s11 > m) { m = s11; dy = 1; dx = 1; }
  a2s[t] = fmaxf(m * m2s[co], 0.0f);
  pos2s[t] = (uint8_t)((2 * py + dy) * 12 + 2 * px + dx);
}

// fc1 + bias + ReLU + element dropout (src/model.py:19-20).
// a2s[320] -> h1s[50]; W3 is read from global (L2-resident, coalesced).
// W_SMEM: W3 has been staged in shared memory (fused step); otherwise it is global.
template <bool W_SMEM>
__device__ __forceinline__ void fc1_fwd_img(const float* __restrict__ a2s,
                                            const float* __restrict__ W3,
                                            const float* __restrict__ b3,
                                            const float* __restrict__ m3s,
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
    for (int i = 0; i < A2_SZ / 32; ++i)
      s = fmaf(av[i], W_SMEM ? w[lane + 32 * i] : __ldg(w + lane + 32 * i), s);
    s = warp_sum(s);
    if (lane == 0) h1s[j] = fmaxf(s + __ldg(b3 + j), 0.0f) * m3s[j];
  }
}

// fc2 logits (src/model.py:21): h1s[50] -> logit_s[10]; one warp per class.
__device__ __forceinline__ void fc2_fwd_img(const float* __restrict__ h1s,
                                            const float* __restrict__ w4s,
                                            const float* __restrict__ b4s,
                                            float* __restrict__ logit_s) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < NCLS) {
    float s = h1s[lane] * w4s[warp * FC1 + lane];
    if (lane + 32 < FC1) s = fmaf(h1s[lane + 32], w4s[warp * FC1 + lane + 32], s);
    s = warp_sum(s);
    if (lane == 0) logit_s[warp] = s + b4s[warp];
  }
}

// log_softmax over 10 logits by one thread (src/model.py:22). Returns logsumexp.
__device__ __forceinline__ float log_softmax10(const float* __restrict__ logit_s, float* logp) {
  float mx = logit_s[0];
#pragma unroll
  for (int c = 1; c < NCLS; ++c) mx = fmaxf(mx, logit_s[c]);
  float se = 0.0f;
#pragma unroll
  for (int c = 0; c < NCLS; ++c) se += expf(logit_s[c] - mx);
  const float lse = mx + logf(se);
#pragma unroll
  for (int c = 0; c < NCLS; ++c) logp[c] = logit_s[c] - lse;
  return lse;
}

// ---- backward --------------------------------------------------------------

// fc2 backward: dlog_s[10] (gradient w.r.t. logits) -> dW4, db4 partials and
// dpre_s[50] = gradient w.r.t. fc1's pre-activation (through dropout and ReLU).
__device__ __forceinline__ void fc2_bwd_img(const float* __restrict__ dlog_s,
                                            const float* __restrict__ h1s,
                                            const float* __restrict__ m3s,
                                            const float* __restrict__ w4s,
                                            float* __restrict__ dpre_s, float* __restrict__ gW4,
                                            float* __restrict__ gB4, bool first,
                                            bool store = true) {
  const int t = threadIdx.x;
  if (store) {
    for (int o = t; o < W4_SZ; o += NT) {
      const int c = o / FC1, j = o - c * FC1;
      gacc(gW4 + o, dlog_s[c] * h1s[j], first);
    }
    if (t < NCLS) gacc(gB4 + t, dlog_s[t], first);
  }
  if (t >= 64 && t < 64 + FC1) {
    const int j = t - 64;
    float s = 0.0f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) s = fmaf(dlog_s[c], w4s[c * FC1 + j], s);
    // h1 = relu(pre) * m3, so h1 > 0 <=> pre > 0 and the unit was kept.
    dpre_s[j] = (h1s[j] > 0.0f) ? s * m3s[j] : 0.0f;
  }
}

// fc1 backward: dpre_s[50], a2s[320] -> dW3, db3 partials and
// g2s[320] = gradient w.r.t. conv2's pooled pre-ReLU map (ReLU and channel mask applied).
template <bool W_SMEM>
__device__ __forceinline__ void fc1_bwd_img(const float* __restrict__ dpre_s,
                                            const float* __restrict__ a2s,
                                            const float* __restrict__ m2s,
                                            const float* __restrict__ W3,
                                            float* __restrict__ g2s, float* __restrict__ gW3,
                                            float* __restrict__ gB3, bool first,
                                            bool store = true) {
  const int t = threadIdx.x;  // t == k in [0,320)
  const float a = a2s[t];
  float da = 0.0f;
  if (store) {
#pragma unroll 10
    for (int j = 0; j < FC1; ++j) {
      const float d = dpre_s[j];
      gacc(gW3 + j * A2_SZ + t, d * a, first);
      da = fmaf(d, W_SMEM ? W3[j * A2_SZ + t] : __ldg(W3 + j * A2_SZ + t), da);
    }
    if (t < FC1) gacc(gB3 + t, dpre_s[t], first);
  } else {
#pragma unroll 10
    for (int j = 0; j < FC1; ++j)
      da = fmaf(dpre_s[j], W_SMEM ? W3[j * A2_SZ + t] : __ldg(W3 + j * A2_SZ + t), da);
  }
  g2s[t] = (a > 0.0f) ? da * m2s[t >> 4] : 0.0f;
}

// conv2 backward. g2s[20*16] sits at pos2s (arg-max of each pooled window).
// Produces dW2, db2 partials and g1s[10*144] = gradient w.r.t. conv1's pooled
// pre-ReLU map (ReLU applied).
// MASK_BY_A1: fold conv1's ReLU mask (a1 > 0) into g1s and skip the work where it is
// zero (fused step). false gives the plain gradient w.r.t. a1 (autograd path).
template <bool MASK_BY_A1>
__device__ __forceinline__ void conv2_bwd_img(const float* __restrict__ g2s,
                                              const uint8_t* __restrict__ pos2s,
                                              const float* __restrict__ a1s,
                                              const float* __restrict__ w2s,
                                              float* __restrict__ g1s, float* __restrict__ gW2,
                                              float* __restrict__ gB2, bool first,
                                              bool need_dgrad) {
  const int t = threadIdx.x;
  // wgrad: 5000 outputs, 16 pooled windows each
  for (int o = t; o < W2_SZ; o += NT) {
    const int co = o / 250, r = o - co * 250;
    const int ci = r / 25, kk = r - ci * 25;
    const int ky = kk / 5, kx = kk - ky * 5;
    const float* a = a1s + ci * 144 + ky * 12 + kx;
    const float* g = g2s + co * 16;
    const uint8_t* ps = pos2s + co * 16;
    float s = 0.0f;
#pragma unroll
    for (int p = 0; p < 16; ++p) s = fmaf(g[p], a[ps[p]], s);
    gacc(gW2 + o, s, first);
  }
  if (t < C2) {
    float s = 0.0f;
#pragma unroll
    for (int p = 0; p < 16; ++p) s += g2s[t * 16 + p];
    gacc(gB2 + t, s, first);
  }
  if (!need_dgrad) return;
  // dgrad: 1440 outputs; each gathers from the <=3x3 pooled windows whose
  // arg-max can reach it through a 5x5 filter.
  for (int q = t; q < A1_SZ; q += NT) {
    const int ci = q / 144, r = q - ci * 144;
    const int u = r / 12, v = r - u * 12;
    const int pylo = max(0, (u - 4) >> 1), pyhi = min(3, u >> 1);
    const int pxlo = max(0, (v - 4) >> 1), pxhi = min(3, v >> 1);
    float s = 0.0f;
    if (!MASK_BY_A1 || a1s[q] > 0.0f) {
      for (int co = 0; co < C2; ++co) {
        const float* w = w2s + (co * C1 + ci) * 25;
        for (int py = pylo; py <= pyhi; ++py)
          for (int px = pxlo; px <= pxhi; ++px) {
            const int p = co * 16 + py * 4 + px;
            const int pos = pos2s[p];
            const int Y = pos / 12, X = pos - Y * 12;
            const int ky = u - Y, kx = v - X;
            if ((unsigned)ky < 5u && (unsigned)kx < 5u) s = fmaf(g2s[p], w[ky * 5 + kx], s);
          }
      }
    }
    g1s[q] = s;
  }
}

// conv1 weight/bias gradient. g1s[10*144] sits at pos1s inside the 24x24 map.
__device__ __forceinline__ void conv1_bwd_img(const float* __restrict__ g1s,
                                              const uint16_t* __restrict__ pos1s,
                                              const float* __restrict__ xs,
                                              float* __restrict__ gW1, float* __restrict__ gB1,
                                              bool first) {
  const int t = threadIdx.x;
  if (t < W1_SZ) {
    const int c = t / 25, kk = t - c * 25;
    const int ky = kk / 5, kx = kk - ky * 5;
    const float* x = xs + ky * IMG + kx;
    const float* g = g1s + c * 144;
    const uint16_t* ps = pos1s + c * 144;
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
#pragma unroll 4
    for (int p = 0; p < 144; p += 4) {
      s0 = fmaf(g[p], x[ps[p]], s0);
      s1 = fmaf(g[p + 1], x[ps[p + 1]], s1);
      s2 = fmaf(g[p + 2], x[ps[p + 2]], s2);
      s3 = fmaf(g[p + 3], x[ps[p + 3]], s3);
    }
    gacc(gW1 + t, (s0 + s1) + (s2 + s3), first);
  } else if (t >= 256 && t < 256 + C1) {
    const float* g = g1s + (t - 256) * 144;
    float s = 0.0f;
    for (int p = 0; p < 144; ++p) s += g[p];
    gacc(gB1 + (t - 256), s, first);
  }
}


// ---- backward of the fused step -----------------------------------------------
// Measured per phase with clock64 (scripts/phase_timing.py): with one image per CTA the
// backward is LATENCY bound, not instruction bound. Compacted non-zero lists made the loops
// short but chained dependent shared-memory loads (count -> index -> weight), and cost 3x
// more cycles than dense, branch-free, vectorised loops. So the fused step uses:
//   * channel-level sparsity only: a channel dropped by Dropout2d has zero gradient for the
//     whole image, and the test `m2s[co] == 0` is uniform across the CTA;
//   * broadcast 16-byte loads of the 16 window gradients / positions of a channel;
//   * dgrad as a dense [64 pos x 20 ch] x [20 ch x 250] product out of registers (T), then
//     a 25-tap gather with compile-time offsets.

// D[pos][slot] = window gradient of the slot-th LIVE channel if its arg-max sits at conv position
// pos (8x8 frame), else 0. Every (channel, window) owns exactly one distinct entry: no
// conflicts. live[] / *nlive_out list the channels Dropout2d kept (all 20 in eval mode).
__device__ __forceinline__ void build_conv2_dense(const float* __restrict__ g2s,
                                                  const uint8_t* __restrict__ pos2s,
                                                  const float* __restrict__ m2s,
                                                  float* __restrict__ Ds, int* __restrict__ live,
                                                  int* __restrict__ slot_of) {
  const int t = threadIdx.x;  // t == co*16 + p
  for (int i = t; i < 64 * C2 / 4; i += NT)
    reinterpret_cast<float4*>(Ds)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (t < 32) {               // warp 0: compact the live channels
    const bool on = (t < C2) && (m2s[t] != 0.0f);
    const unsigned bal = __ballot_sync(0xffffffffu, on);
    const int sl = __popc(bal & ((1u << t) - 1u));
    if (on) live[sl] = t;
    if (t < C2) slot_of[t] = on ? sl : -1;
    if (t == 0) live[C2] = __popc(bal);   // live[20] = count
  }
  __syncthreads();
  const int sl = slot_of[t >> 4];
  if (sl >= 0) {
    const int pos12 = pos2s[t];
    const int Y = pos12 / 12, X = pos12 - Y * 12;
    Ds[(Y * 8 + X) * C2 + sl] = g2s[t];
  }
}

// conv2 weight / bias gradient: thread <-> one (ci, tap) column, loop over live channels.
__device__ __forceinline__ void conv2_wgrad_img(const float* __restrict__ g2s,
                                                const uint8_t* __restrict__ pos2s,
                                                const float* __restrict__ m2s,
                                                const float* __restrict__ a1s,
                                                float* __restrict__ gW2, float* __restrict__ gB2,
                                                bool first) {
  const int t = threadIdx.x;
  if (t < 250) {
    const int ci = t / 25, kk = t - ci * 25;
    const int ky = kk / 5, kx = kk - ky * 5;
    const float* a = a1s + ci * 144 + ky * 12 + kx;
#pragma unroll 1
    for (int co = 0; co < C2; ++co) {
      float s = 0.0f;
      if (m2s[co] != 0.0f) {   // uniform: the whole channel was dropped otherwise
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
      } else if (!first) {
        continue;              // nothing to add to the running partial
      }
      gacc(gW2 + co * 250 + t, s, first);
    }
  } else if (t >= 256 && t < 256 + C2) {
    const int co = t - 256;
    float s = 0.0f;
#pragma unroll
    for (int p = 0; p < 16; ++p) s += g2s[co * 16 + p];
    gacc(gB2 + co, s, first);
  }
}

// dgrad phase 1: T[pos][r] = sum_co D[pos][co] * W2[co][r], r = ci*25 + tap. D is stored with
// the LIVE channels (not dropped by Dropout2d) compacted to the front, `live[slot]` names the
// channel of a slot and nlive is their count; the 4-channel groups beyond nlive are skipped by a
// uniform branch, so a typical image does 12 instead of 20 FMAs per output.
__device__ __forceinline__ void conv2_dgrad_T_img(const float* __restrict__ Ds,
                                                  const int* __restrict__ live, int nlive,
                                                  const float* __restrict__ w2s,
                                                  float* __restrict__ T) {
  const int t = threadIdx.x;
  if (t >= 250) return;
  float w[C2];
#pragma unroll
  for (int sl = 0; sl < C2; ++sl) w[sl] = (sl < nlive) ? w2s[live[sl] * 250 + t] : 0.0f;
  const int nq = (nlive + 3) >> 2;
#pragma unroll 2
  for (int pos = 0; pos < 64; ++pos) {
    const float4* d4 = reinterpret_cast<const float4*>(Ds + pos * C2);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int q = 0; q < C2 / 4; ++q) {
      if (q < nq) {   // uniform across the CTA
        const float4 d = d4[q];
        s0 = fmaf(d.x, w[4 * q], s0);
        s1 = fmaf(d.y, w[4 * q + 1], s1);
        s0 = fmaf(d.z, w[4 * q + 2], s0);
        s1 = fmaf(d.w, w[4 * q + 3], s1);
      }
    }
    T[pos * 250 + t] = s0 + s1;
  }
}

// dgrad phase 2 (after a barrier): g1s[ci][u][v] = sum over the <=25 filter taps of
// T[(u-ky, v-kx)][ci][tap], with conv1's ReLU mask folded in.
__device__ __forceinline__ void conv2_dgrad_gather_img(const float* __restrict__ T,
                                                       const float* __restrict__ a1s,
                                                       float* __restrict__ g1s) {
  for (int q = threadIdx.x; q < A1_SZ; q += NT) {
    const int ci = q / 144, r = q - ci * 144;
    const int u = r / 12, v = r - u * 12;
    float s = 0.0f;
    if (a1s[q] > 0.0f) {
      // (Y*8+X)*250 + ci*25 + tap with Y=u-ky, X=v-kx is base - ky*1995 - kx*249:
      // every tap is a compile-time offset from one base address.
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

// conv1 weight / bias gradient over all 144 windows of a channel, 4 windows per step with
// 16-byte gradient and 8-byte position broadcasts (zeros contribute nothing).
__device__ __forceinline__ void conv1_wgrad_img(const float* __restrict__ g1s,
                                                const uint16_t* __restrict__ pos1s,
                                                const float* __restrict__ xs,
                                                float* __restrict__ gW1, float* __restrict__ gB1,
                                                bool first) {
  const int t = threadIdx.x;
  if (t < W1_SZ) {
    const int c = t / 25, kk = t - c * 25;
    const int ky = kk / 5, kx = kk - ky * 5;
    const float* x = xs + ky * IMG + kx;
    const float4* g4 = reinterpret_cast<const float4*>(g1s + c * 144);
    const uint2* p2 = reinterpret_cast<const uint2*>(pos1s + c * 144);
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
#pragma unroll 6
    for (int e = 0; e < 36; ++e) {
      const float4 gv = g4[e];
      const uint2 pp = p2[e];
      s0 = fmaf(gv.x, x[pp.x & 0xffffu], s0);
      s1 = fmaf(gv.y, x[pp.x >> 16], s1);
      s2 = fmaf(gv.z, x[pp.y & 0xffffu], s2);
      s3 = fmaf(gv.w, x[pp.y >> 16], s3);
    }
    gacc(gW1 + t, (s0 + s1) + (s2 + s3), first);
  } else if (t >= 256 && t < 256 + C1) {
    const float4* g4 = reinterpret_cast<const float4*>(g1s + (t - 256) * 144);
    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll 6
    for (int e = 0; e < 36; ++e) {
      const float4 g = g4[e];
      s0 += g.x + g.z;
      s1 += g.y + g.w;
    }
    gacc(gB1 + (t - 256), s0 + s1, first);
  }
}

// Fixed-order sum of G gradient partial rows for one float4 column, spread over the 8
// row-groups of a 256-thread CTA (32 columns x 8 groups) so all loads are in flight at
// once. The result is valid in the threads of row-group 0.
__device__ __forceinline__ float4 sum_partial_rows_256(const float* __restrict__ partial, int G,
                                                       int stride, int col4, bool valid,
                                                       float4 (*sm)[32]) {
  const int c = threadIdx.x & 31, rg = threadIdx.x >> 5;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) {
#pragma unroll 4
    for (int r = rg; r < G; r += 8) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(partial + (size_t)r * stride) + col4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  sm[rg][c] = s;
  __syncthreads();
  if (rg == 0) {
#pragma unroll
    for (int k = 1; k < 8; ++k) {
      const float4 v = sm[k][c];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  return s;
}

}  // namespace net
