// NVLink SHARP (NVLS) all-reduce: the switch does the reduction.
//
// Host side: a multicast object spanning all ranks' GPUs (cuMulticastCreate), each rank's
// physical allocation (cuMemCreate) bound into it, and two mappings per rank: a unicast
// pointer to the local memory and a multicast pointer through which
//   multimem.ld_reduce  returns the SUM over every rank's copy of an address (reduced in
//                       the NVSwitch), and
//   multimem.st         writes a value to every rank's copy.
// The POSIX file descriptor of the multicast object travels from rank 0 to the others over
// a Unix socket (Python side). Driver entry points are resolved at run time, so the library
// keeps loading on machines without a driver.
//
// Device side: `nvls_allreduce_kernel`. Rank r owns slice r of the buffer. Per CTA:
//   flag round 1 (everyone's data is in place) -> multimem.ld_reduce of my sub-slice ->
//   scale (+ optional SGD on my shard of master weights / momentum) ->
//   multimem.st of the result into every rank's buffer (plain all-reduce), or peer stores
//   of the new weights into every rank's parameter buffer (fused SGD) -> flag round 2.
// Per-GPU NVLink traffic is ~S instead of 2S(N-1)/N for the two-shot kernel.
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstring>

#include "comm.cuh"

namespace {

#define DRV(name) \
  static decltype(&name) fn_##name = nullptr; \
  if (!fn_##name) { \
    void* p_ = nullptr; cudaDriverEntryPointQueryResult q_; \
    if (cudaGetDriverEntryPoint(#name, &p_, cudaEnableDefault, &q_) != cudaSuccess || \
        q_ != cudaDriverEntryPointSuccess || !p_) { \
      b200_set_error("driver entry point %s is not available", #name); \
      return -2; \
    } \
    fn_##name = reinterpret_cast<decltype(&name)>(p_); \
  }

#define CU_OK(expr) \
  do { \
    CUresult r_ = (expr); \
    if (r_ != CUDA_SUCCESS) { \
      b200_set_error("%s:%d: %s failed with CUresult %d", __FILE__, __LINE__, #expr, (int)r_); \
      return (int)r_; \
    } \
  } while (0)

CUmulticastObjectProp mc_prop(int world, size_t bytes) {
  CUmulticastObjectProp p;
  memset(&p, 0, sizeof(p));
  p.numDevices = (unsigned)world;
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

// ---- device helpers ----------------------------------------------------------------
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const void* mc) {
  uint4 v;  // 8 bf16, accumulated in fp32 inside the switch
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

struct NvlsArgs {
  char* mc_base;          // multicast mapping of the NVLS region
  size_t buf_off;         // byte offset of the buffer inside the region
  size_t nvec;            // 16-byte vectors
  size_t flag_off;        // flags [2][gmax][MAX_WORLD] in the cudaIpc arena
  int gmax;
  unsigned* seq_ctr;
  unsigned* arrive_ctr;
  float scale;
  int bf16;
  // fused SGD: new weights go to out_off (cudaIpc arena) on every rank
  int update;
  size_t out_off;
  float* master;
  float* mom;
  const float* hp;
  size_t state_base;
};

__global__ void __launch_bounds__(512) nvls_allreduce_kernel(CommDev c, NvlsArgs a) {
  const unsigned seq = *a.seq_ctr + 1u;
  const int b = blockIdx.x, G = gridDim.x, W = c.world, nt = blockDim.x;
  unsigned* my_flags = peer_ptr<unsigned>(c, c.rank, a.flag_off);
  const size_t f1 = ((size_t)0 * a.gmax + b) * B200_MAX_WORLD;
  const size_t f2 = ((size_t)1 * a.gmax + b) * B200_MAX_WORLD;
  // round 1: every rank's buffer is complete (stream order on each rank) -> tell everyone
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < W) {
    const int p = threadIdx.x;
    st_release_sys_u32(peer_ptr<unsigned>(c, p, a.flag_off) + f1 + c.rank, seq);
    wait_flag_ge(my_flags + f1 + p, seq, c, p, 10u);
  }
  __syncthreads();
  const size_t sl = (a.nvec + W - 1) / W;
  const size_t sub = (sl + G - 1) / G;
  const size_t mybase = (size_t)c.rank * sl;
  const size_t slo = (size_t)b * sub;
  const size_t lim = min(min(sl, slo + sub), a.nvec > mybase ? a.nvec - mybase : 0);
  char* mc = a.mc_base + a.buf_off;
  const int N = a.bf16 ? 8 : 4;
  // 4 switch reductions in flight per thread: multimem.ld_reduce is a round trip through the
  // NVSwitch, so bandwidth needs depth.
  for (size_t v0 = slo + threadIdx.x; v0 < lim; v0 += 4 * (size_t)nt) {
    uint4 x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t v = v0 + (size_t)u * nt;
      if (v < lim)
        x[u] = a.bf16 ? mm_ld_reduce_bf16(mc + (mybase + v) * 16) : mm_ld_reduce_f32(mc + (mybase + v) * 16);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t v = v0 + (size_t)u * nt;
      if (v >= lim) continue;
      const size_t gv = mybase + v;
      float f[8];
      if (a.bf16) {
        const unsigned w[4] = {x[u].x, x[u].y, x[u].z, x[u].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          f[2 * i] = __uint_as_float(w[i] << 16);
          f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
      } else {
        f[0] = __uint_as_float(x[u].x); f[1] = __uint_as_float(x[u].y);
        f[2] = __uint_as_float(x[u].z); f[3] = __uint_as_float(x[u].w);
        f[4] = f[5] = f[6] = f[7] = 0.0f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] *= a.scale;
      if (a.update) {
        const float lr = a.hp[0], mu = a.hp[1], wd = a.hp[2];
        const size_t e0 = gv * N - a.state_base;      // multiple of N: 16-byte aligned state
        float4* mw = reinterpret_cast<float4*>(a.master + e0);
        float4* mb = reinterpret_cast<float4*>(a.mom + e0);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          if (q >= N / 4) break;
          float4 w = mw[q], m = mb[q];
          float* wp = reinterpret_cast<float*>(&w);
          float* bp = reinterpret_cast<float*>(&m);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float g = f[4 * q + i] + wd * wp[i];
            bp[i] = mu * bp[i] + g;
            wp[i] -= lr * bp[i];
            f[4 * q + i] = wp[i];
          }
          mw[q] = w;
          mb[q] = m;
        }
      }
      uint4 o;
      if (a.bf16) {
        unsigned w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
          w[i] = *reinterpret_cast<const unsigned*>(&h);
        }
        o = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        o = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                       __float_as_uint(f[3]));
      }
      if (a.update) {
        for (int pp = 0; pp < W; ++pp) {
          const int p = (c.rank + pp) % W;
          peer_ptr<uint4>(c, p, a.out_off)[gv] = o;
        }
      } else {
        mm_st(mc + gv * 16, o);   // one store, the switch replicates it to every rank
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < W) {
    const int p = threadIdx.x;
    st_release_sys_u32(peer_ptr<unsigned>(c, p, a.flag_off) + f2 + c.rank, seq);
    wait_flag_ge(my_flags + f2 + p, seq, c, p, 11u);
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

extern "C" {

int b200_nvls_supported(int dev, int* out) {
  DRV(cuDeviceGetAttribute)
  int v = 0;
  CU_OK(fn_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  *out = v;
  return 0;
}

// Round `bytes` up to the multicast granularity for `world` devices.
int b200_nvls_granularity(int world, size_t bytes, size_t* rounded) {
  DRV(cuMulticastGetGranularity)
  CUmulticastObjectProp p = mc_prop(world, bytes);
  size_t g = 0;
  CU_OK(fn_cuMulticastGetGranularity(&g, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  *rounded = (bytes + g - 1) / g * g;
  return 0;
}

// Rank 0: create the multicast object (bytes already rounded) and export it as a POSIX fd.
int b200_nvls_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out) {
  DRV(cuMulticastCreate)
  DRV(cuMemExportToShareableHandle)
  B200_CUDA_OK(cudaFree(0));
  CUmulticastObjectProp p = mc_prop(world, bytes);
  CUmemGenericAllocationHandle h = 0;
  CU_OK(fn_cuMulticastCreate(&h, &p));
  int fd = -1;
  CU_OK(fn_cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc_out = (unsigned long long)h;
  *fd_out = fd;
  return 0;
}

// Other ranks: import the multicast object from the received fd.
int b200_nvls_import(int fd, unsigned long long* mc_out) {
  DRV(cuMemImportFromShareableHandle)
  B200_CUDA_OK(cudaFree(0));
  CUmemGenericAllocationHandle h = 0;
  CU_OK(fn_cuMemImportFromShareableHandle(&h, (void*)(uintptr_t)fd,
                                          CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_out = (unsigned long long)h;
  return 0;
}

int b200_nvls_add_device(unsigned long long mc, int dev) {
  DRV(cuMulticastAddDevice)
  DRV(cuDeviceGet)
  CUdevice d;
  CU_OK(fn_cuDeviceGet(&d, dev));
  CU_OK(fn_cuMulticastAddDevice((CUmemGenericAllocationHandle)mc, d));
  return 0;
}

// Every rank, after ALL ranks have added their device: allocate local memory, bind it into
// the multicast object, map the local (unicast) view and the multicast view.
int b200_nvls_bind_map(unsigned long long mc, int dev, size_t bytes, void** uc_ptr, void** mc_ptr) {
  DRV(cuMemCreate)
  DRV(cuMulticastBindMem)
  DRV(cuMemAddressReserve)
  DRV(cuMemMap)
  DRV(cuMemSetAccess)
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mem = 0;
  CU_OK(fn_cuMemCreate(&mem, bytes, &prop, 0));
  CU_OK(fn_cuMulticastBindMem((CUmemGenericAllocationHandle)mc, 0, mem, 0, bytes, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CUdeviceptr uc = 0, mcva = 0;
  CU_OK(fn_cuMemAddressReserve(&uc, bytes, 0, 0, 0));
  CU_OK(fn_cuMemMap(uc, bytes, 0, mem, 0));
  CU_OK(fn_cuMemSetAccess(uc, bytes, &acc, 1));
  CU_OK(fn_cuMemAddressReserve(&mcva, bytes, 0, 0, 0));
  CU_OK(fn_cuMemMap(mcva, bytes, 0, (CUmemGenericAllocationHandle)mc, 0));
  CU_OK(fn_cuMemSetAccess(mcva, bytes, &acc, 1));
  B200_CUDA_OK(cudaMemset((void*)uc, 0, bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  *uc_ptr = (void*)uc;
  *mc_ptr = (void*)mcva;
  return 0;
}

// In-place all-reduce of [buf_off, buf_off + bytes) of the NVLS region (dtype 0 fp32, 1 bf16),
// optionally fused with sharded SGD (new weights pushed to out_off in the cudaIpc arena).
int b200_nvls_allreduce(const void* comm, void* mc_base, size_t buf_off, size_t bytes, int dtype,
                        size_t flag_off, int gmax, unsigned* seq_ctr, unsigned* arrive_ctr,
                        float scale, int update, size_t out_off, float* master, float* mom,
                        const float* hp, size_t state_base, int blocks, void* stream) {
  const CommDev& c = *(const CommDev*)comm;
  B200_REQUIRE(bytes > 0 && bytes % 16 == 0 && buf_off % 16 == 0,
               "b200_nvls_allreduce: 16-byte granularity");
  B200_REQUIRE(!update || (master && mom && hp), "b200_nvls_allreduce: SGD fusion needs state");
  NvlsArgs a;
  a.mc_base = (char*)mc_base;
  a.buf_off = buf_off;
  a.nvec = bytes / 16;
  a.flag_off = flag_off;
  a.gmax = gmax;
  a.seq_ctr = seq_ctr;
  a.arrive_ctr = arrive_ctr;
  a.scale = scale;
  a.bf16 = dtype == 1;
  a.update = update;
  a.out_off = out_off;
  a.master = master;
  a.mom = mom;
  a.hp = hp;
  a.state_base = state_base;
  if (blocks <= 0) {
    const size_t work = (a.nvec + c.world - 1) / c.world;
    blocks = (int)((work + 2047) / 2048);
    if (blocks > gmax) blocks = gmax;
    if (blocks < 1) blocks = 1;
  }
  B200_REQUIRE(blocks <= gmax, "b200_nvls_allreduce: too many CTAs for the flag table");
  nvls_allreduce_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(c, a);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
