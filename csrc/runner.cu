// Native step runner: feeds the fused Net step from pinned host memory without any
// Python in the loop.
//
// The reference's loop pulls each batch from a DataLoader, runs ~40 kernels and calls
// loss.item() (src/train_dist.py:78-89). Here one call `b200_runner_run(h, hx, hy, steps,
// hloss, ...)` submits work on three streams so that copies overlap compute and nothing
// synchronises with the host until the caller does. Two submission schemes share the kernels:
//
// Per step (remainders, or when no loss buffer is given), NSLOT staging slots:
//     copy stream   : wait(slot free) -> H2D images -> H2D labels -> record(copied)
//     compute stream: wait(copied)    -> cudaGraphLaunch(one-step graph of that slot)
//                                      -> record(consumed)
//     d2h stream    : wait(consumed)  -> D2H 4-byte loss of that step from the loss ring
//
// Per chunk of CHUNK steps (the steady state), two chunk-sized staging buffers:
//     copy stream   : wait(buffer free) -> per step: H2D images, H2D labels -> record(copied)
//     compute stream: wait(copied) -> cudaGraphLaunch(CHUNK-step graph of that buffer)
//                                   -> record(consumed)
//     d2h stream    : wait(consumed) -> per step: D2H 4-byte loss
// Every step still has its own host->device copies of its inputs and its own device->host
// copy of its loss; what changes is that the 2 x CHUNK kernels of a chunk are ONE graph, so
// they chain with programmatic dependent launch across step boundaries exactly like the
// device-resident path (a one-step graph cannot overlap kernel A of step i+1 with kernel B
// of step i), and the host issues ~3 stream operations per step instead of 9. The copies of
// chunk c+1 overlap the kernels of chunk c.
//
// Measured alternative, not kept: putting the copies INSIDE the chunk graph as memcpy nodes
// (retargeted per chunk with cudaGraphExecMemcpyNodeSetParams1D) was slower than the per-step
// scheme (25.4 vs 24.0 us/step at batch 64): a kernel node with a copy-node dependency loses
// its programmatic edge, and copy nodes cost more than stream copies.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "comm.cuh"

extern "C" {
int b200_net_fwdbwd(const float* params, const void* x, int x_mode, const long long* labels,
                    const int* gidx, int B, float* partial, float* loss_partial, int mask_mode,
                    float* m2buf, float* m3buf, unsigned long long seed,
                    const unsigned long long* step_ptr, unsigned sample_base, int idx_stride,
                    int n_idx, void* stream);
int b200_net_reduce_sgd(const float* partial, int rows, float* params, float* mom, float* grad_out,
                        const float* hp, const float* loss_partial, float* loss_ring, int ring,
                        unsigned long long* step_ptr, void* stream);
int b200_comm_net_reduce_allreduce_sgd(const void* comm, const float* partial, int rows,
                                       float* params, float* mom, float* grad_out,
                                       const float* hp, const float* loss_partial,
                                       float* loss_ring, int ring, unsigned long long* ctl,
                                       size_t stage_off, size_t flag_off, unsigned* seq_ctr,
                                       unsigned* arrive_ctr, int protocol, void* stream);
int b200_net_fused_rows(int B);
}

namespace {

constexpr int NSLOT = 4;
constexpr int CHUNK = 16;

struct RunnerConfig {  // mirrored field by field in ops/runner.py (ctypes.Structure)
  float* params;
  float* mom;
  float* partial;
  float* loss_partial;
  float* hp;
  float* loss_ring;
  unsigned long long* ctl;
  int ring;
  int batch;
  int x_mode;        // 0 fp32 [B,784], 1 uint8 [B,784]
  int mask_mode;     // 0 none, 1 Philox dropout
  unsigned long long seed;
  unsigned sample_base;
  int multi;         // 1: kernel B all-reduces across ranks
  const void* comm;  // CommDev blob (host) when multi
  size_t stage_off;
  size_t flag_off;
  unsigned* seq_ctr;
  unsigned* arrive_ctr;
  int protocol;
  int pad_;
};

struct Runner {
  RunnerConfig cfg;
  CommDev comm;
  size_t x_bytes;
  void* sx[NSLOT];
  long long* sy[NSLOT];
  cudaGraphExec_t exec[NSLOT];
  cudaStream_t compute, copy, d2h;
  cudaEvent_t copied[NSLOT], consumed[NSLOT], join_in, join_compute, join_d2h;
  unsigned long long steps_submitted;
  // chunk path: two staging buffers of CHUNK batches, one CHUNK-step graph per buffer
  int chunk_ok;
  char* cx[2];
  long long* cy[2];
  cudaGraphExec_t chunk_exec[2];
  cudaEvent_t c_copied[2], c_consumed[2];
  unsigned long long chunks_launched;
};

int launch_step(Runner* r, const void* x, const long long* y, cudaStream_t s) {
  const RunnerConfig& c = r->cfg;
  int e = b200_net_fwdbwd(c.params, x, c.x_mode, y, nullptr, c.batch, c.partial,
                          c.loss_partial, c.mask_mode, nullptr, nullptr, c.seed, c.ctl,
                          c.sample_base, 0, 0, s);
  if (e) return e;
  const int rows = b200_net_fused_rows(c.batch);
  if (c.multi)
    return b200_comm_net_reduce_allreduce_sgd(&r->comm, c.partial, rows, c.params, c.mom, nullptr,
                                              c.hp, c.loss_partial, c.loss_ring, c.ring, c.ctl,
                                              c.stage_off, c.flag_off, c.seq_ctr, c.arrive_ctr,
                                              c.protocol, s);
  return b200_net_reduce_sgd(c.partial, rows, c.params, c.mom, nullptr, c.hp, c.loss_partial,
                             c.loss_ring, c.ring, c.ctl, s);
}

// Capture `n` consecutive steps reading batches x + i * x_bytes, y + i * batch.
int capture_steps(Runner* r, const char* x, const long long* y, int n, cudaGraphExec_t* out) {
  cudaGraph_t g = nullptr;
  B200_CUDA_OK(cudaStreamBeginCapture(r->compute, cudaStreamCaptureModeThreadLocal));
  int e = 0;
  for (int i = 0; i < n && !e; ++i)
    e = launch_step(r, x + (size_t)i * r->x_bytes, y + (size_t)i * r->cfg.batch, r->compute);
  const cudaError_t ce = cudaStreamEndCapture(r->compute, &g);
  if (e) {
    if (g) cudaGraphDestroy(g);
    return e;
  }
  B200_CUDA_OK(ce);
  B200_CUDA_OK(cudaGraphInstantiate(out, g, 0));
  B200_CUDA_OK(cudaGraphDestroy(g));
  // upload now: otherwise the first cudaGraphLaunch pays for it inside the caller's loop
  B200_CUDA_OK(cudaGraphUpload(*out, r->compute));
  return 0;
}

}  // namespace

extern "C" {

int b200_runner_config_size(void) { return (int)sizeof(RunnerConfig); }

// Create a runner. The step kernels must have run at least once on this device before
// (function attributes / lazy module loading are not allowed inside stream capture).
int b200_runner_create(const void* cfg_blob, void** out) {
  Runner* r = new (std::nothrow) Runner();
  B200_REQUIRE(r != nullptr, "b200_runner_create: out of host memory");
  memset(r, 0, sizeof(*r));
  memcpy(&r->cfg, cfg_blob, sizeof(RunnerConfig));
  const RunnerConfig& c = r->cfg;
  B200_REQUIRE(c.batch > 0 && (c.x_mode == 0 || c.x_mode == 1), "b200_runner_create: bad config");
  if (c.multi) memcpy(&r->comm, c.comm, sizeof(CommDev));
  r->x_bytes = (size_t)c.batch * 784 * (c.x_mode == 0 ? 4 : 1);
  const size_t y_bytes = (size_t)c.batch * sizeof(long long);
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->compute, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->copy, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->d2h, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_in, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_compute, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_d2h, cudaEventDisableTiming));
  for (int s = 0; s < NSLOT; ++s) {
    B200_CUDA_OK(cudaMalloc(&r->sx[s], r->x_bytes));
    B200_CUDA_OK(cudaMalloc((void**)&r->sy[s], y_bytes));
    B200_CUDA_OK(cudaMemset(r->sx[s], 0, r->x_bytes));
    B200_CUDA_OK(cudaMemset(r->sy[s], 0, y_bytes));
    B200_CUDA_OK(cudaEventCreateWithFlags(&r->copied[s], cudaEventDisableTiming));
    B200_CUDA_OK(cudaEventCreateWithFlags(&r->consumed[s], cudaEventDisableTiming));
  }
  // the chunk path needs every loss of two chunks in flight to have its own ring slot
  const char* env = getenv("B200DP_RUNNER_CHUNK");
  r->chunk_ok = (c.ring >= 4 * CHUNK) && !(env && env[0] == '0');
  if (r->chunk_ok) {
    for (int b = 0; b < 2; ++b) {
      B200_CUDA_OK(cudaMalloc((void**)&r->cx[b], (size_t)CHUNK * r->x_bytes));
      B200_CUDA_OK(cudaMalloc((void**)&r->cy[b], (size_t)CHUNK * y_bytes));
      B200_CUDA_OK(cudaMemset(r->cx[b], 0, (size_t)CHUNK * r->x_bytes));
      B200_CUDA_OK(cudaMemset(r->cy[b], 0, (size_t)CHUNK * y_bytes));
      B200_CUDA_OK(cudaEventCreateWithFlags(&r->c_copied[b], cudaEventDisableTiming));
      B200_CUDA_OK(cudaEventCreateWithFlags(&r->c_consumed[b], cudaEventDisableTiming));
    }
  }
  B200_CUDA_OK(cudaDeviceSynchronize());
  for (int s = 0; s < NSLOT; ++s) {
    const int e = capture_steps(r, static_cast<const char*>(r->sx[s]), r->sy[s], 1, &r->exec[s]);
    if (e) return e;
  }
  if (r->chunk_ok) {
    for (int b = 0; b < 2; ++b) {
      const int e = capture_steps(r, r->cx[b], r->cy[b], CHUNK, &r->chunk_exec[b]);
      if (e) return e;
    }
  }
  *out = r;
  return 0;
}

// Submit `steps` optimizer steps. hx: pinned host images [steps*B, 784] (fp32 or uint8 per
// the config), hy: pinned int64 labels [steps*B], hloss: pinned float [steps] or null.
// Work is ordered after everything already queued on caller_stream, and caller_stream is
// made to wait for its completion; nothing blocks the host except one 8-byte read of the
// step counter at the start (to know which ring slots the losses will land in).
int b200_runner_run(void* h, const void* hx, const long long* hy, int steps, float* hloss,
                    void* caller_stream) {
  Runner* r = static_cast<Runner*>(h);
  const RunnerConfig& c = r->cfg;
  const size_t y_bytes = (size_t)c.batch * sizeof(long long);
  const unsigned long long ring = (unsigned long long)c.ring;
  cudaStream_t cs = (cudaStream_t)caller_stream;
  B200_CUDA_OK(cudaEventRecord(r->join_in, cs));
  B200_CUDA_OK(cudaStreamWaitEvent(r->compute, r->join_in, 0));
  B200_CUDA_OK(cudaStreamWaitEvent(r->copy, r->join_in, 0));
  unsigned long long step0 = 0;
  if (hloss) {
    B200_CUDA_OK(cudaMemcpyAsync(&step0, c.ctl, sizeof(step0), cudaMemcpyDeviceToHost, r->compute));
    B200_CUDA_OK(cudaStreamSynchronize(r->compute));
  }
  const char* px = static_cast<const char*>(hx);
  int i0 = 0;
  // ---- whole chunks: per-step copies into a chunk buffer, one CHUNK-step graph, per-step D2H
  if (hloss && r->chunk_ok) {
    for (; i0 + CHUNK <= steps; i0 += CHUNK) {
      const int b = (int)(r->chunks_launched & 1);
      if (r->chunks_launched >= 2)
        B200_CUDA_OK(cudaStreamWaitEvent(r->copy, r->c_consumed[b], 0));   // buffer free again
      for (int i = 0; i < CHUNK; ++i) {
        B200_CUDA_OK(cudaMemcpyAsync(r->cx[b] + (size_t)i * r->x_bytes,
                                     px + (size_t)(i0 + i) * r->x_bytes, r->x_bytes,
                                     cudaMemcpyHostToDevice, r->copy));
        B200_CUDA_OK(cudaMemcpyAsync(r->cy[b] + (size_t)i * c.batch,
                                     hy + (size_t)(i0 + i) * c.batch, y_bytes,
                                     cudaMemcpyHostToDevice, r->copy));
      }
      B200_CUDA_OK(cudaEventRecord(r->c_copied[b], r->copy));
      B200_CUDA_OK(cudaStreamWaitEvent(r->compute, r->c_copied[b], 0));
      B200_CUDA_OK(cudaGraphLaunch(r->chunk_exec[b], r->compute));
      B200_CUDA_OK(cudaEventRecord(r->c_consumed[b], r->compute));
      B200_CUDA_OK(cudaStreamWaitEvent(r->d2h, r->c_consumed[b], 0));
      for (int i = 0; i < CHUNK; ++i)
        B200_CUDA_OK(cudaMemcpyAsync(hloss + i0 + i, c.loss_ring + ((step0 + i0 + i) % ring),
                                     sizeof(float), cudaMemcpyDeviceToHost, r->d2h));
      ++r->chunks_launched;
    }
  }
  // ---- remaining steps, one at a time through the staging slots
  for (int i = i0; i < steps; ++i) {
    const int s = (int)(r->steps_submitted % NSLOT);
    if (r->steps_submitted >= NSLOT)
      B200_CUDA_OK(cudaStreamWaitEvent(r->copy, r->consumed[s], 0));
    B200_CUDA_OK(cudaMemcpyAsync(r->sx[s], px + (size_t)i * r->x_bytes, r->x_bytes,
                                 cudaMemcpyHostToDevice, r->copy));
    B200_CUDA_OK(cudaMemcpyAsync(r->sy[s], hy + (size_t)i * c.batch, y_bytes,
                                 cudaMemcpyHostToDevice, r->copy));
    B200_CUDA_OK(cudaEventRecord(r->copied[s], r->copy));
    B200_CUDA_OK(cudaStreamWaitEvent(r->compute, r->copied[s], 0));
    B200_CUDA_OK(cudaGraphLaunch(r->exec[s], r->compute));
    B200_CUDA_OK(cudaEventRecord(r->consumed[s], r->compute));
    if (hloss) {
      B200_CUDA_OK(cudaStreamWaitEvent(r->d2h, r->consumed[s], 0));
      B200_CUDA_OK(cudaMemcpyAsync(hloss + i, c.loss_ring + ((step0 + i) % ring), sizeof(float),
                                   cudaMemcpyDeviceToHost, r->d2h));
    }
    ++r->steps_submitted;      // counts per-step submissions only (slot rotation)
  }
  B200_CUDA_OK(cudaEventRecord(r->join_compute, r->compute));
  B200_CUDA_OK(cudaStreamWaitEvent(cs, r->join_compute, 0));
  if (hloss) {
    B200_CUDA_OK(cudaEventRecord(r->join_d2h, r->d2h));
    B200_CUDA_OK(cudaStreamWaitEvent(cs, r->join_d2h, 0));
  }
  return 0;
}

// Number of CHUNK-step graphs launched so far (the other steps went through the per-step path).
long long b200_runner_chunks_launched(void* h) {
  return h ? (long long)static_cast<Runner*>(h)->chunks_launched : 0;
}

int b200_runner_destroy(void* h) {
  Runner* r = static_cast<Runner*>(h);
  if (!r) return 0;
  cudaStreamSynchronize(r->compute);
  cudaStreamSynchronize(r->copy);
  cudaStreamSynchronize(r->d2h);
  for (int s = 0; s < NSLOT; ++s) {
    if (r->exec[s]) cudaGraphExecDestroy(r->exec[s]);
    if (r->sx[s]) cudaFree(r->sx[s]);
    if (r->sy[s]) cudaFree(r->sy[s]);
    if (r->copied[s]) cudaEventDestroy(r->copied[s]);
    if (r->consumed[s]) cudaEventDestroy(r->consumed[s]);
  }
  for (int b = 0; b < 2; ++b) {
    if (r->chunk_exec[b]) cudaGraphExecDestroy(r->chunk_exec[b]);
    if (r->cx[b]) cudaFree(r->cx[b]);
    if (r->cy[b]) cudaFree(r->cy[b]);
    if (r->c_copied[b]) cudaEventDestroy(r->c_copied[b]);
    if (r->c_consumed[b]) cudaEventDestroy(r->c_consumed[b]);
  }
  cudaEventDestroy(r->join_in);
  cudaEventDestroy(r->join_compute);
  cudaEventDestroy(r->join_d2h);
  cudaStreamDestroy(r->compute);
  cudaStreamDestroy(r->copy);
  cudaStreamDestroy(r->d2h);
  delete r;
  return 0;
}

}  // extern "C"
