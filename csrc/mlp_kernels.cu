// Small SIMT kernels around the tcgen05 GEMMs of the wide MLP (BASELINE.json config 4):
// fused softmax + NLL + gradient over padded logits, bias gradient (column sums), and the
// single-GPU SGD-momentum update with fp32 master weights behind bf16 parameters.
// Reference semantics: CrossEntropyLoss / nll_loss on the network output
// (src/train_dist.py:67,82) and optim.SGD (src/train_dist.py:66).
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

// logits fp32 [B][ld] (first ncls columns valid). Writes the mean NLL to *loss and
// dlogits = (softmax - onehot) / B as bf16 [B][ld_d] with zero padding columns.
__global__ void __launch_bounds__(256)
softmax_nll_kernel(const float* __restrict__ logits, int ld, int ncls,
                   const long long* __restrict__ labels, float* __restrict__ loss,
                   __nv_bfloat16* __restrict__ dlogits, int ld_d, int B) {
  __shared__ float red[256];
  float acc = 0.0f;
  const float inv_B = 1.0f / (float)B;
  for (int i = threadIdx.x; i < B; i += 256) {
    const float* row = logits + (size_t)i * ld;
    float mx = row[0];
    for (int c = 1; c < ncls; ++c) mx = fmaxf(mx, row[c]);
    float se = 0.0f;
    for (int c = 0; c < ncls; ++c) se += expf(row[c] - mx);
    const float lse = mx + logf(se);
    const int t = (int)labels[i];
    acc += (lse - row[t]);
    __nv_bfloat16* d = dlogits + (size_t)i * ld_d;
    for (int c = 0; c < ld_d; ++c) {
      float g = 0.0f;
      if (c < ncls) g = (expf(row[c] - lse) - (c == t ? 1.0f : 0.0f)) * inv_B;
      d[c] = __float2bfloat16_rn(g);
    }
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = red[0] * inv_B;
}

// out[n] = sum_b dy[b][n] (bias gradient), fp32 accumulation, bf16 result.
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ dy, long long ld,
                   __nv_bfloat16* __restrict__ out, int B, int N) {
  __shared__ float part[8][33];
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rg = threadIdx.x >> 5;
  float s = 0.0f;
  if (col < N)
    for (int b = rg; b < B; b += 8) s += __bfloat162float(dy[(size_t)b * ld + col]);
  part[rg][threadIdx.x & 31] = s;
  __syncthreads();
  if (rg == 0 && col < N) {
    for (int k = 1; k < 8; ++k) s += part[k][threadIdx.x];
    out[col] = __float2bfloat16_rn(s);
  }
}

// bf16 parameters, bf16 gradients, fp32 master weights and momentum (one GPU).
__global__ void __launch_bounds__(256)
sgd_master_bf16_kernel(__nv_bfloat16* __restrict__ p, const __nv_bfloat16* __restrict__ g,
                       float* __restrict__ master, float* __restrict__ mom, long long n8,
                       const float* __restrict__ hp) {
  const float lr = hp[0], mu = hp[1], wd = hp[2];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    const uint4 gv = reinterpret_cast<const uint4*>(g)[i];
    const unsigned gw[4] = {gv.x, gv.y, gv.z, gv.w};
    float4 m0 = reinterpret_cast<float4*>(master)[2 * i], m1 = reinterpret_cast<float4*>(master)[2 * i + 1];
    float4 b0 = reinterpret_cast<float4*>(mom)[2 * i], b1 = reinterpret_cast<float4*>(mom)[2 * i + 1];
    float w[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    unsigned ow[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float g0 = __uint_as_float(gw[q] << 16) + wd * w[2 * q];
      const float g1 = __uint_as_float(gw[q] & 0xffff0000u) + wd * w[2 * q + 1];
      b[2 * q] = mu * b[2 * q] + g0;
      b[2 * q + 1] = mu * b[2 * q + 1] + g1;
      w[2 * q] -= lr * b[2 * q];
      w[2 * q + 1] -= lr * b[2 * q + 1];
      const __nv_bfloat162 h = __floats2bfloat162_rn(w[2 * q], w[2 * q + 1]);
      ow[q] = *reinterpret_cast<const unsigned*>(&h);
    }
    reinterpret_cast<float4*>(master)[2 * i] = make_float4(w[0], w[1], w[2], w[3]);
    reinterpret_cast<float4*>(master)[2 * i + 1] = make_float4(w[4], w[5], w[6], w[7]);
    reinterpret_cast<float4*>(mom)[2 * i] = make_float4(b[0], b[1], b[2], b[3]);
    reinterpret_cast<float4*>(mom)[2 * i + 1] = make_float4(b[4], b[5], b[6], b[7]);
    reinterpret_cast<uint4*>(p)[i] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
}

}  // namespace

extern "C" {

int b200_softmax_nll_bf16(const float* logits, int ld, int ncls, const long long* labels,
                          float* loss, void* dlogits, int ld_d, int B, void* stream) {
  B200_REQUIRE(B > 0 && ncls > 0 && ncls <= ld && ncls <= ld_d, "b200_softmax_nll_bf16: bad shape");
  softmax_nll_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(logits, ld, ncls, labels, loss,
                                                          (__nv_bfloat16*)dlogits, ld_d, B);
  B200_LAUNCH_OK();
  return 0;
}

int b200_colsum_bf16(const void* dy, long long ld, void* out, int B, int N, void* stream) {
  B200_REQUIRE(B > 0 && N > 0, "b200_colsum_bf16: bad shape");
  colsum_bf16_kernel<<<(N + 31) / 32, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, ld, (__nv_bfloat16*)out, B, N);
  B200_LAUNCH_OK();
  return 0;
}

int b200_sgd_master_bf16(void* p, const void* g, float* master, float* mom, long long n,
                         const float* hp, void* stream) {
  B200_REQUIRE(n > 0 && n % 8 == 0, "b200_sgd_master_bf16: n must be a multiple of 8");
  const long long n8 = n / 8;
  long long blocks = (n8 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  sgd_master_bf16_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)p, (const __nv_bfloat16*)g, master, mom, n8, hp);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
