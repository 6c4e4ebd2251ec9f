// Native step runner: feeds the fused Net step from pinned host memory without any
// Python in the loop.
//
// The reference's loop pulls each batch from a DataLoader, runs ~40 kernels and calls
// loss.item() (src/train_dist.py:78-89). Here one call `b200_runner_run(h, hx, hy, steps,
// hloss, ...)` submits, per step:
//     copy stream   : wait(slot free) -> H2D images -> H2D labels -> record(copied)
//     compute stream: wait(copied)    -> cudaGraphLaunch(step graph of that slot)
//                                      -> record(consumed)
//     d2h stream    : wait(consumed)  -> D2H 4-byte loss of that step from the loss ring
// with NSLOT staging buffers, so the copy of batch i+1 overlaps the kernels of batch i and
// nothing synchronises with the host until the caller does. The step graph (kernel A +
// kernel B, single- or multi-GPU) is captured here, once per slot, from the same launch
// functions the Python path uses.
#include <cstring>
#include <new>

#include "comm.cuh"

extern "C" {
int b200_net_fwdbwd(const float* params, const void* x, int x_mode, const long long* labels,
                    const int* gidx, int B, float* partial, float* loss_partial, int mask_mode,
                    float* m2buf, float* m3buf, unsigned long long seed,
                    const unsigned long long* step_ptr, unsigned sample_base, int idx_stride,
                    void* stream);
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

constexpr int NSLOT = 3;

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
};

int launch_step(Runner* r, int slot, cudaStream_t s) {
  const RunnerConfig& c = r->cfg;
  int e = b200_net_fwdbwd(c.params, r->sx[slot], c.x_mode, r->sy[slot], nullptr, c.batch, c.partial,
                          c.loss_partial, c.mask_mode, nullptr, nullptr, c.seed, c.ctl,
                          c.sample_base, 0, s);
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
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->compute, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->copy, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&r->d2h, cudaStreamNonBlocking));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_in, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_compute, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventCreateWithFlags(&r->join_d2h, cudaEventDisableTiming));
  for (int s = 0; s < NSLOT; ++s) {
    B200_CUDA_OK(cudaMalloc(&r->sx[s], r->x_bytes));
    B200_CUDA_OK(cudaMalloc((void**)&r->sy[s], (size_t)c.batch * sizeof(long long)));
    B200_CUDA_OK(cudaMemset(r->sx[s], 0, r->x_bytes));
    B200_CUDA_OK(cudaMemset(r->sy[s], 0, (size_t)c.batch * sizeof(long long)));
    B200_CUDA_OK(cudaEventCreateWithFlags(&r->copied[s], cudaEventDisableTiming));
    B200_CUDA_OK(cudaEventCreateWithFlags(&r->consumed[s], cudaEventDisableTiming));
  }
  B200_CUDA_OK(cudaDeviceSynchronize());
  for (int s = 0; s < NSLOT; ++s) {
    cudaGraph_t g = nullptr;
    B200_CUDA_OK(cudaStreamBeginCapture(r->compute, cudaStreamCaptureModeThreadLocal));
    const int e = launch_step(r, s, r->compute);
    cudaError_t ce = cudaStreamEndCapture(r->compute, &g);
    if (e) return e;
    B200_CUDA_OK(ce);
    B200_CUDA_OK(cudaGraphInstantiate(&r->exec[s], g, 0));
    B200_CUDA_OK(cudaGraphDestroy(g));
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
  for (int i = 0; i < steps; ++i) {
    const int s = (int)((r->steps_submitted + i) % NSLOT);
    if (r->steps_submitted + i >= NSLOT)
      B200_CUDA_OK(cudaStreamWaitEvent(r->copy, r->consumed[s], 0));
    B200_CUDA_OK(cudaMemcpyAsync(r->sx[s], px + (size_t)i * r->x_bytes, r->x_bytes,
                                 cudaMemcpyHostToDevice, r->copy));
    B200_CUDA_OK(cudaMemcpyAsync(r->sy[s], hy + (size_t)i * c.batch,
                                 (size_t)c.batch * sizeof(long long), cudaMemcpyHostToDevice,
                                 r->copy));
    B200_CUDA_OK(cudaEventRecord(r->copied[s], r->copy));
    B200_CUDA_OK(cudaStreamWaitEvent(r->compute, r->copied[s], 0));
    B200_CUDA_OK(cudaGraphLaunch(r->exec[s], r->compute));
    B200_CUDA_OK(cudaEventRecord(r->consumed[s], r->compute));
    if (hloss) {
      B200_CUDA_OK(cudaStreamWaitEvent(r->d2h, r->consumed[s], 0));
      B200_CUDA_OK(cudaMemcpyAsync(hloss + i, c.loss_ring + ((step0 + i) % (unsigned long long)c.ring),
                                   sizeof(float), cudaMemcpyDeviceToHost, r->d2h));
    }
  }
  r->steps_submitted += (unsigned long long)steps;
  B200_CUDA_OK(cudaEventRecord(r->join_compute, r->compute));
  B200_CUDA_OK(cudaStreamWaitEvent(cs, r->join_compute, 0));
  if (hloss) {
    B200_CUDA_OK(cudaEventRecord(r->join_d2h, r->d2h));
    B200_CUDA_OK(cudaStreamWaitEvent(cs, r->join_d2h, 0));
  }
  return 0;
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
