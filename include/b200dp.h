/* C ABI of libb200dp.so: the native layer of the B200 data-parallel training framework.
 *
 * Conventions
 *  - Every function that can fail returns 0 on success and non-zero on failure;
 *    b200_last_error() then returns a thread-local description.
 *  - Pointers named for device data are CUDA device pointers on the current device.
 *  - `stream` is a cudaStream_t passed as void*; launches are asynchronous on it.
 *  - Batches: B images; images are fp32 [B,1,28,28] normalised, or uint8 as noted.
 *  - "Reference" citations are file:line in the reference repository
 *    (abhishekiitm/CSED_514_Project_Distributed_Training_using_PyTorch).
 */
#ifndef B200DP_H
#define B200DP_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime -------------------------------------------------------------- */
const char* b200_last_error(void);
int b200_version(void);
/* Used by the library's own translation units to record the message b200_last_error returns
 * (printf-style); exported so side-by-side builds of single sources can link against it. */
void b200_set_error(const char* fmt, ...);

/* ---- Net layout / workspace queries ---------------------------------------- */
/* offsets9[0..7]: float offsets of conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b,
 * fc2.w, fc2.b in the flat parameter buffer; offsets9[8]: total floats (21848).
 * Reference: parameter set of src/model.py:9-13. */
int b200_net_layout(int* offsets9);
int b200_net_max_rows(void);            /* max CTAs (= partial rows) a Net kernel uses: 148 */
int b200_net_scratch_floats(int rows);  /* scratch floats the layered backwards need */
int b200_net_fused_rows(int B);         /* partial rows the fused step uses for batch B */

/* ---- layered Net kernels (autograd path) ----------------------------------- */
/* K21: uint8 rows (optionally gathered by idx) -> normalised fp32. Ref: src/train_dist.py:16-19 */
int b200_gather_normalize(const void* data_u8, const int* idx, float* out, int B, void* stream);
/* K1+K2 conv1 + bias + 2x2 max-pool + ReLU. pos1: int16 [B,10,12,12]. Ref: src/model.py:16 */
int b200_conv1_fwd(const float* x, const float* w, const float* b, float* a1, void* pos1, int B,
                   void* stream);
/* K13 conv1 weight/bias gradient. scratch: >= rows*264 floats, counter: zeroed u32. */
int b200_conv1_bwd(const float* x, const float* da1, const float* a1, const void* pos1, float* dw,
                   float* db, float* scratch, unsigned* counter, int B, void* stream);
/* Same gradient as b200_conv1_bwd for callers that cannot allocate device memory: no
 * scratch/counter arguments, the workspace is a fixed 135,424-byte array inside the library
 * (per GPU, held until exit). No counterpart in the reference. B <= 128
 * (b200_conv1_bwd_noalloc_max_batch()), larger batches are refused. The first call on a
 * device resolves the array's address: make it before any stream capture. All calls on a
 * device share the array: calls that overlap on the GPU race, serialise them. */
int b200_conv1_bwd_noalloc(const float* x, const float* da1, const float* a1, const void* pos1,
                           float* dw, float* db, int B, void* stream);
int b200_conv1_bwd_noalloc_max_batch(void);
size_t b200_conv1_bwd_noalloc_scratch_bytes(void);
/* K3+K4+K5 conv2 + bias + Dropout2d + pool + ReLU. mask [B,20]; mask_mode 0 none, 1 generate
 * (Philox, written to mask), 2 use caller's. pos2: uint8 [B,20,4,4]. Ref: src/model.py:17 */
int b200_conv2_fwd(const float* a1, const float* w, const float* b, float* mask, int mask_mode,
                   float p, unsigned long long seed, unsigned long long step, float* a2,
                   void* pos2, int B, void* stream);
/* K12 conv2 backward: dw, db, da1 (da1 may be null). */
int b200_conv2_bwd(const float* a1, const float* w, const float* da2, const float* a2,
                   const void* pos2, const float* mask, float* dw, float* db, float* da1,
                   float* scratch, unsigned* counter, int B, void* stream);
/* K7 fc1 + ReLU + dropout. mask [B,50]. Ref: src/model.py:19-20 */
int b200_fc1_fwd(const float* a2, const float* W, const float* b, float* mask, int mask_mode,
                 float p, unsigned long long seed, unsigned long long step, float* h1, int B,
                 void* stream);
/* K11 fc1 backward. */
int b200_fc1_bwd(const float* a2, const float* W, const float* dh1, const float* h1,
                 const float* mask, float* dW, float* db, float* da2, float* scratch,
                 unsigned* counter, int B, void* stream);
/* K8+K9 fc2 + log_softmax / K10 backward. Ref: src/model.py:21-22 */
int b200_fc2_lsm_fwd(const float* h1, const float* W, const float* b, float* logp, int B,
                     void* stream);
int b200_fc2_lsm_bwd(const float* h1, const float* W, const float* logp, const float* dlogp,
                     float* dW, float* db, float* dh1, float* scratch, unsigned* counter, int B,
                     void* stream);
/* NLL loss over [B,10] log-probabilities; reduction 0 mean, 1 sum. Ref: src/train.py:74,94 */
int b200_nll_fwd(const float* logp, const long long* target, float* loss, int B, int reduction,
                 void* stream);
int b200_nll_bwd(const float* gloss, const long long* target, float* dlogp, int B, int reduction,
                 void* stream);

/* ---- optimizer --------------------------------------------------------------- */
/* K15 SGD-momentum over a flat buffer of n floats (n % 4 == 0).
 * hp (device) = {lr, momentum, weight_decay, nesterov}. Ref: src/train.py:60-61,76 */
int b200_sgd_momentum(float* p, const float* g, float* buf, long long n, const float* hp,
                      float grad_scale, void* stream);

/* ---- fused Net step ------------------------------------------------------------ */
/* Kernel A: forward + log_softmax/NLL + backward, one image per CTA iteration.
 * x_mode 0 fp32 [B,784]; 1 uint8 [B,784]; 2 uint8 dataset gathered by gidx; 3 fp32 dataset
 * gathered by gidx. partial: [rows][21848] floats; loss_partial: [rows].
 * step_ptr (device u64[2]) = {optimizer step, step at epoch start}; idx_stride != 0 makes the
 * batch's indices gidx[(step - epoch_start) * idx_stride ...]; n_idx is then the length of
 * gidx (the next step's rows are prefetched only while they exist). Ref: src/train.py:72-75 */
int b200_net_fwdbwd(const float* params, const void* x, int x_mode, const long long* labels,
                    const int* gidx, int B, float* partial, float* loss_partial, int mask_mode,
                    float* m2buf, float* m3buf, unsigned long long seed,
                    const unsigned long long* step_ptr, unsigned sample_base, int idx_stride,
                    int n_idx, void* stream);
/* Kernel B (1 GPU): fixed-order row sum + SGD-momentum; loss -> loss_ring[step % ring] and
 * loss_ring[ring]; advances step_ptr[0]. Ref: src/train.py:76 */
int b200_net_reduce_sgd(const float* partial, int rows, float* params, float* mom, float* grad_out,
                        const float* hp, const float* loss_partial, float* loss_ring, int ring,
                        unsigned long long* step_ptr, void* stream);
/* K19 evaluation forward + per-image NLL / correct; out2 = {nll_sum, n_correct}.
 * Ref: src/train.py:87-104, src/train_dist.py:92-104 */
int b200_net_eval(const float* params, const void* x, int x_mode, const long long* labels,
                  const int* gidx, int B, float* logp_out, float* loss_img, int* pred_img,
                  float* out2, int accumulate, void* stream);

/* ---- symmetric memory (one process per GPU, cudaIpc over NVLink) ------------------ */
/* Ref: replaces the Gloo/TCP transport set up at src/train_dist.py:144-146 */
int b200_symm_alloc(size_t bytes, void** dptr, unsigned char* handle64);
int b200_symm_open(const unsigned char* handle64, void** peer_ptr_out);
int b200_symm_close(void* peer_ptr_in);
int b200_symm_free(void* dptr);
int b200_device_can_access_peer(int dev, int peer, int* out);

/* ---- communicator + collectives ---------------------------------------------------- */
int b200_comm_struct_size(void);
int b200_comm_make(void* blob, int rank, int world, void* const* bases, unsigned* err,
                   unsigned long long timeout_ns);
int b200_comm_barrier(const void* comm, size_t flag_off, unsigned* seq_ctr, void* stream);
/* K16 DDP-init parameter broadcast. Ref: DDP(model) at src/train_dist.py:63 */
int b200_comm_broadcast(const void* comm, int root, size_t buf_off, size_t bytes, size_t flag_off,
                        unsigned* seq_ctr, unsigned* arrive_ctr, void* stream);
/* K20 point-to-point. Ref: dist.send / dist.recv at src/run1.py:13,16 */
int b200_comm_send(const void* comm, int dst, const void* payload, size_t bytes, size_t box_off,
                   size_t box_bytes, size_t flag_off, size_t ack_off, unsigned* sent_ctr,
                   void* stream);
int b200_comm_recv(const void* comm, int src, void* out, size_t bytes, size_t box_off,
                   size_t box_bytes, size_t flag_off, size_t ack_off, unsigned* recv_ctr,
                   void* stream);
int b200_comm_sendrecv(const void* comm, int dst, const void* payload, size_t send_bytes, int src,
                       void* out, size_t recv_bytes, size_t box_off, size_t box_bytes,
                       size_t flag_off, size_t ack_off, unsigned* sent_ctr, unsigned* recv_ctr,
                       void* stream);
/* K17+K15 Net kernel B across GPUs: row sum + one-shot all-reduce + 1/world + SGD-momentum.
 * Ref: the implicit all-reduce inside backward (src/train_dist.py:83) + optimizer.step (:84) */
size_t b200_comm_net_stage_bytes(void);
size_t b200_comm_net_flag_bytes(void);
int b200_comm_net_reduce_allreduce_sgd(const void* comm, const float* partial, int rows,
                                       float* params, float* mom, float* grad_out,
                                       const float* hp, const float* loss_partial,
                                       float* loss_ring, int ring, unsigned long long* ctl,
                                       size_t stage_off, size_t flag_off, unsigned* seq_ctr,
                                       unsigned* arrive_ctr, int protocol /* 0 LL, 1 flag */,
                                       void* stream);
/* Generic bucket all-reduce. dtype 0 fp32 / 1 bf16; algo 0 one-shot / 1 two-shot;
 * update 1 fuses sharded SGD-momentum (destination receives new weights). */
int b200_comm_allreduce(const void* comm, int dtype, int algo, int update, const void* in,
                        void* out_local, size_t out_off, size_t bytes, size_t stage_off,
                        size_t src_bytes, size_t flag_off, int gmax, unsigned* seq_ctr,
                        unsigned* arrive_ctr, float scale, float* master, float* mom,
                        const float* hp, size_t state_base, int blocks, size_t push_from,
                        void* stream);

/* ---- NVLS: in-switch reduction through an NVLink multicast object --------------------- */
/* Ref: the gradient all-reduce implicit in src/train_dist.py:83; SURVEY §5.1 "NVLS path".
 * Rank 0 creates the multicast object and exports it as a POSIX fd; the other ranks import
 * it (the fd travels over a Unix socket, parallel/comm.py); every rank adds its device, then
 * binds a cuMemCreate allocation and maps both a unicast and a multicast view of it. */
int b200_nvls_supported(int dev, int* out);
int b200_nvls_granularity(int world, size_t bytes, size_t* rounded);
int b200_nvls_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out);
int b200_nvls_import(int fd, unsigned long long* mc_out);
int b200_nvls_add_device(unsigned long long mc, int dev);
int b200_nvls_bind_map(unsigned long long mc, int dev, size_t bytes, void** uc_ptr, void** mc_ptr);
/* All-reduce of `bytes` at buf_off inside the NVLS region: multimem.ld_reduce of this rank's
 * slice, optional fused SGD on the master weights (update = 1), multimem.st of the result. */
int b200_nvls_allreduce(const void* comm, void* mc_base, size_t buf_off, size_t bytes, int dtype,
                        size_t flag_off, int gmax, unsigned* seq_ctr, unsigned* arrive_ctr,
                        float scale, int update, size_t out_off, float* master, float* mom,
                        const float* hp, size_t state_base, int blocks, void* stream);

/* ---- tcgen05 bf16 GEMM ------------------------------------------------------------------ */
/* C[M,N] = op(A) op(B)^T, fp32 accumulation in TMEM. a_mn/b_mn select MN-major operands, so
 * forward, dgrad and wgrad of a linear layer need no transposes. Ref: nn.Linear layers of
 * src/model.py:12-13 widened per BASELINE.json config 4. */
int b200_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb,
                   void* C, int c_fp32, long long ldc, int M, int N, int K, const void* bias,
                   int relu, const void* relu_mask, long long ld_mask, void* stream);
/* Same product with a fused reduce-scatter epilogue: result vectors go straight to the owner
 * ranks' staging of a two-shot all-reduce channel over NVLink peer memory. */
int b200_gemm_bf16_rs(const void* A, int a_mn, long long lda, const void* B, int b_mn,
                      long long ldb, void* C, int c_fp32, long long ldc, int M, int N, int K,
                      const void* bias, int relu, const void* relu_mask, long long ld_mask,
                      const void* rs_comm, size_t rs_stage_off, size_t rs_src_bytes,
                      size_t rs_slice_vec, size_t rs_base_vec, const unsigned* rs_seq,
                      void* stream);
/* Pin the GEMM kernel: kind 0 = one 128 x bn tile per CTA (bn 16..256), 1 = one 256 x bn tile
 * per CTA pair (cta_group::2), 2 = persistent CTA pairs with double-buffered TMEM, 3 = kind 2
 * with every tile split in two along K (where the shape allows); bn 0 = pick.
 * kind -1 restores the automatic choice (also: env B200DP_GEMM_VARIANT=kind[,bn]). */
int b200_gemm_set_variant(int kind, int bn);
/* The kernel kind, tile width and K split b200_gemm_bf16 would use for this shape (host-only). */
int b200_gemm_plan(int M, int N, int K, int rs_on, int both_mn, int* kind, int* bn, int* ksplit);
/* conv2 forward as a TF32 implicit GEMM on tcgen05 (same contract as b200_conv2_fwd). */
int b200_conv2_fwd_tc(const float* a1, const float* w, const float* b, float* mask, int mask_mode,
                      float p, unsigned long long seed, unsigned long long step, float* a2,
                      void* pos2, int B, void* stream);
/* conv2 as an implicit GEMM WITHOUT im2col (csrc/conv2_tc.cu): activations in the c4 layout
 * [B][4 chunks][12][12][4 floats] arrive by one 5-D TMA load per image pair, every filter tap is
 * a descriptor start-address shift, 50 tcgen05.mma.kind::tf32 per pair, pool/ReLU/Dropout2d in the
 * TMEM epilogue. Same outputs as b200_conv2_fwd. Ref: src/model.py:17 */
int b200_a1_to_c4(const float* a1, float* c4, int B, void* stream);
/* conv2.weight -> tensor-core B operand [25 taps][4 chunks][32][4 floats] (b200_conv2_operand_floats()
 * floats); b200_conv2_fwd_tc2 takes THAT as `w_operand`. */
int b200_conv2_w_to_operand(const float* w, float* w_operand, void* stream);
int b200_conv2_operand_floats(void);
int b200_conv2_fwd_tc2(const float* a1_c4, const float* w_operand, const float* b, float* mask, int mask_mode,
                       float p, unsigned long long seed, unsigned long long step, float* a2,
                       void* pos2, int B, void* stream);

/* conv2 backward on tcgen05 (csrc/conv2_tc.cu). b200_conv2_dz_frame scatters the pooled gradient
 * into a zero-padded frame [B][6][200][4] (b200_conv2_frame_floats(B) floats) that
 * the data-gradient kernel reads through UMMA descriptors only, as an implicit GEMM over
 * flipped taps (weights from b200_conv2_w_to_dgrad_operand). Ref: autograd of
 * src/model.py:24 run by src/train_dist.py:83 */
int b200_conv2_frame_floats(int B);
int b200_conv2_dgrad_operand_floats(void);
int b200_conv2_dz_frame(const float* da2, const float* a2, const void* pos2, const float* mask,
                        float* frame, int B, void* stream);
int b200_conv2_w_to_dgrad_operand(const float* w, float* wd, void* stream);
int b200_conv2_dgrad_tc(const float* frame, const float* wd, float* da1, int B, void* stream);
/* weight + bias gradient: both operands MN-major in the swizzled [pixel][32 channels] layout
 * written by b200_a1_to_c32 / b200_conv2_dz_c32 (512-byte aligned buffers) */
int b200_conv2_c32_a1_floats(int B);
int b200_conv2_c32_dz_floats(int B);
int b200_a1_to_c32(const float* a1, float* c32, int B, void* stream);
int b200_conv2_dz_c32(const float* da2, const float* a2, const void* pos2, const float* mask,
                      float* dz, int B, void* stream);
int b200_conv2_dz_both(const float* da2, const float* a2, const void* pos2, const float* mask,
                       float* frame, float* dz_c32, int B, void* stream);   /* frame + c32 in one pass */
int b200_conv2_wgrad_tc(const float* a1_c32, const float* dz_c32, float* dw, float* db, int B, void* stream);
/* MLP helpers: fused softmax + NLL + gradient, bias gradient, master-weight SGD. */
int b200_softmax_nll_bf16(const float* logits, int ld, int ncls, const long long* labels,
                          float* loss, void* dlogits, int ld_d, int B, void* stream);
int b200_colsum_bf16(const void* dy, long long ld, void* out, int B, int N, void* stream);
int b200_sgd_master_bf16(void* p, const void* g, float* master, float* mom, long long n,
                         const float* hp, void* stream);
/* Native step runner (pinned host -> device streaming of the fused step). */
int b200_runner_config_size(void);
int b200_runner_create(const void* cfg_blob, void** out);
int b200_runner_run(void* h, const void* hx, const long long* hy, int steps, float* hloss,
                    void* caller_stream);
long long b200_runner_chunks_launched(void* runner);   /* 16-step graphs launched so far */
int b200_runner_destroy(void* h);
int b200_pdl_enabled(void);
int b200_net_set_phase_debug(long long* buf);

/* ---- persistent step kernel (csrc/net_persist.cu) -------------------------------------- */
/* `nsteps` optimizer steps of the MNIST net in ONE cooperative launch: worker CTAs run
 * forward/backward, service CTAs form the fc gradients, sum the conv partial rows, exchange
 * each bucket with the peers while backward still runs, apply SGD-momentum and publish the new
 * weights. Ref: the loop body of src/train_dist.py:78-89 and DDP's overlapped all-reduce
 * (src/train_dist.py:63,83). The argument block mirrors ops/persist.py:_PersistHost. */
struct B200PersistHost;
int b200_net_persist_host_size(void);
/* out9 = {grid, workers, split, rows, images per worker, service CTAs, conv2 CTAs, W3 CTAs,
 * W3 columns per CTA} for a per-GPU batch of B (host-only). */
int b200_net_persist_plan(int B, int* out9);
long long b200_net_persist_scratch_floats(int B);
int b200_net_persist_smem_bytes(void);
long long b200_net_persist_dbg_words(int* steps, int* phases);   /* phase-stamp buffer (debug) */
int b200_net_persist_run(const struct B200PersistHost* args, void* stream);
/* Host feeder: batches from pinned host memory into a device ring on a copy stream, one 32-bit
 * stream write per step tells the running kernel what has landed; the kernel stores each
 * step's loss into the caller's pinned buffer. Ref: DataLoader(pin_memory=True) + loss.item(),
 * src/train_dist.py:40-45,85. */
int b200_feeder_create(const struct B200PersistHost* tmpl, int ring_slots, void** out);
int b200_feeder_run(void* h, const void* hx, const long long* hy, int steps, float* hloss,
                    void* caller_stream);
long long b200_feeder_launches(void* h);
int b200_feeder_destroy(void* h);

/* ---- bucket planner + reducer core (host only) --------------------------------------- */
/* Ref: torch DDP's bucket rule reached from src/train_dist.py:63 */
int b200_plan_buckets(const long long* bytes, int n, long long first_cap, long long cap,
                      int* bucket_of);
int b200_plan_buckets_by_group(int n, const int* boundaries, int ngroups, int* bucket_of);
void* b200_reducer_create(const int* bucket_of, int n, int nbuckets);
void b200_reducer_destroy(void* h);
void b200_reducer_reset(void* h);
int b200_reducer_mark_ready(void* h, int i);
int b200_reducer_all_done(void* h);
int b200_reducer_pending_buckets(void* h, int* out, int cap);
int b200_reducer_completion_order(void* h, int* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H */
