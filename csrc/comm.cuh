// Device-side view of a communicator: one process per GPU, every rank's symmetric arena
// mapped into every other rank's address space (cudaIpc over NVLink 5 / NVSwitch).
//
// The reference's communication backend is Gloo over TCP between VMs
// (src/train_dist.py:144-146). Here "the network" is peer memory: a collective is a kernel
// that stores to / loads from `base[p] + offset` and synchronises through flag words with
// release/acquire semantics at system scope. All flag values are monotonically increasing
// sequence numbers kept in device memory, so the kernels can be captured in CUDA graphs
// and replayed without host-side arguments changing.
#pragma once
#include "common.cuh"

constexpr int B200_MAX_WORLD = 8;

struct CommDev {
  int rank;
  int world;
  char* base[B200_MAX_WORLD];  // arena base of every rank as mapped in THIS process
  unsigned* err;               // local: {code, peer, where, seq}; 0 = healthy
  unsigned long long timeout_ns;
};

#ifdef __CUDACC__

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Wait until *flag (local memory, written by a peer) reaches seq. Bounded: on timeout the
// error word is set and the wait returns false, so a dead peer surfaces as a Python
// exception at the next check instead of a hung GPU.
__device__ __forceinline__ bool wait_flag_ge(const unsigned* flag, unsigned seq, const CommDev& c,
                                             int peer, unsigned where) {
  unsigned v = ld_acquire_sys_u32(flag);
  if ((int)(v - seq) >= 0) return true;
  const unsigned long long t0 = globaltimer_ns();
  for (;;) {
#pragma unroll 1
    for (int i = 0; i < 64; ++i) {
      v = ld_acquire_sys_u32(flag);
      if ((int)(v - seq) >= 0) return true;
    }
    if (globaltimer_ns() - t0 > c.timeout_ns) {
      if (atomicCAS(c.err, 0u, 1u) == 0u) {
        c.err[1] = (unsigned)peer;
        c.err[2] = where;
        c.err[3] = seq;
      }
      return false;
    }
  }
}

template <typename T>
__device__ __forceinline__ T* peer_ptr(const CommDev& c, int p, size_t off) {
  return reinterpret_cast<T*>(c.base[p] + off);
}

#endif  // __CUDACC__
