/* C ABI of libsdml_b200.so -- the hand-written sm_100a kernels of
 * simple_distributed_machine_learning_b200.
 *
 * Conventions
 *   - every entry point returns 0 on success; on failure a non-zero code, with a
 *     human-readable message available from sdml_last_error() (thread-local);
 *   - pointers are device pointers unless the comment says HOST;
 *   - `stream` is a cudaStream_t passed as void*; kernels are enqueued on it and the
 *     call returns without synchronising;
 *   - row-major matrices with an explicit leading dimension in ELEMENTS.
 *
 * "ref:" citations point into /root/reference/simple_distributed.py, the script whose
 * capabilities this framework reproduces.
 */
#ifndef SDML_KERNELS_H_
#define SDML_KERNELS_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---------------------------------------------------------------- */
const char* sdml_last_error(void);
long long sdml_launch_count(void); /* kernels launched by this library so far */
int sdml_abi_version(void);
/* Deadline for every in-kernel wait on another GPU launched after this call (all-gather flags,
 * reduce-scatter / weight-gradient landing slots, refreshed-weight counters, staged DP flags):
 * after timeout_ns (0 = unbounded, the default) the waiting kernel writes {code 11..15, value
 * seen, value wanted, 0} to err_host (4 x uint32 pinned host memory, may be NULL) and carries on;
 * the host is expected to poll err_host.  The reference disables every timeout (ref:36, ref:167). */
int sdml_set_wait_timeout(unsigned long long timeout_ns, void* err_host);

/* ---- bf16 GEMM on tcgen05/TMEM/TMA (csrc/gemm_tcgen05.cu) ---------------------
 * C[M,N] = epilogue(alpha * op(A)[M,K] . op(B)[N,K]^T), fp32 accumulation in TMEM.
 *   a_mn = 0: A stored [M,K] (K contiguous, lda >= K);  a_mn = 1: A stored [K,M].
 *   b_mn = 0: B stored [N,K] (K contiguous, ldb >= K);  b_mn = 1: B stored [K,N].
 * Epilogue, in order: + bias[n] (fp32, may be NULL); ReLU if `relu`; multiply by
 * (aux[m,n] > 0) if aux != NULL (bf16 [M, ld_aux], the ReLU-backward mask);
 * C += result if `accumulate` (requires out_fp32).  Output bf16 or fp32.
 * bn = N-tile (64/128/256; 0 = choose).  lda/ldb multiples of 8, A/B 16 B aligned.
 * Serves ref:63-64,75-77 (fc1/fc2 forward) and their autograd backward (ref:112).
 */
int sdml_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                   long long ldb, long long ldc, int a_mn, int b_mn, const float* bias,
                   const void* aux, long long ld_aux, int relu, int accumulate, int out_fp32,
                   float alpha, int bn, void* stream);

/* ---- fused softmax cross-entropy fwd+bwd (csrc/softmax_ce.cu) ------------------
 * dtype 0 = fp32, 1 = bf16 (logits and dlogits).  targets: int64 [rows].
 * loss_sum (fp32 scalar) += sum_r (logsumexp_r - logit_r[target_r]);
 * dlogits = (softmax - onehot) * grad_scale (skipped if NULL);
 * correct (int32 scalar) += #rows whose arg-max equals the target (skipped if NULL).
 * ref:79 (log_softmax) + ref:111/126 (nll_loss) + ref:127-128 (accuracy).
 */
int sdml_softmax_ce(const void* logits, long long ld, int dtype, const void* targets, int rows,
                    int classes, float* loss_sum, void* dlogits, long long ld_d, float grad_scale,
                    int* correct, void* stream);

/* ---- fused multi-tensor optimizers (csrc/optim.cu) -----------------------------
 * params/grads/moms/m/v/lowp: HOST arrays of n_tensors device pointers (fp32 tensors;
 * lowp = optional bf16 copies of the parameters, refreshed by the update; the array or
 * single entries may be NULL).  sizes: HOST array of element counts.
 * g' = g * grad_scale + weight_decay * p.  zero_grad != 0 clears g afterwards.
 * SGD: buf = first_step ? g' : momentum * buf + g';  p -= lr * buf   (torch.optim.SGD
 * semantics; ref:100-104, ref:113).  Adam: torch.optim.Adam semantics, step >= 1.
 */
int sdml_sgd_multi(int n_tensors, void* const* params, void* const* grads, void* const* moms,
                   void* const* lowp, const long long* sizes, float lr, float momentum,
                   float weight_decay, float grad_scale, int first_step, int zero_grad,
                   void* stream);
int sdml_adam_multi(int n_tensors, void* const* params, void* const* grads, void* const* m,
                    void* const* v, void* const* lowp, const long long* sizes, float lr,
                    float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                    int step, int zero_grad, void* stream);
int sdml_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);
/* out[c] (+)= sum_r x[r,c] for bf16 x [rows, cols] (bias gradient). */
int sdml_colsum_bf16(const void* x, long long ld, int rows, int cols, float* out, int accumulate,
                     void* stream);

/* ---- small fp32 Linear (+bias +ReLU +dropout) (csrc/linear_f32.cu) -------------
 * y[M,N] = dropout_p(relu?(x[M,K] . w[N,K]^T + b)); mask regenerated from (seed, offset +
 * *offset_dev) in the backward pass (offset_dev: optional device int64, for CUDA graphs).  dx / dw+db may be NULL to skip.  ref:75 (fc1+relu+dropout),
 * ref:77 (fc2).
 */
int sdml_linear_f32_fwd(const float* x, const float* w, const float* b, float* y, int M, int N,
                        int K, int relu, float drop_p, long long seed, long long offset,
                        const long long* offset_dev, void* stream);
int sdml_linear_f32_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx,
                        float* dw, float* db, int M, int N, int K, int relu, float drop_p,
                        long long seed, long long offset, const long long* offset_dev,
                        int accumulate, void* stream);

/* ---- data-parallel all-reduce fused with SGD (csrc/dp_allreduce_sgd.cu) ----------
 * One kernel per step and replica: stage the local gradients (scaled by grad_scale) in a
 * buffer the other replicas can read over NVLink, publish an epoch flag to every replica,
 * wait for theirs, sum all replicas' gradients and apply SGD-with-momentum in place
 * (gradients are zeroed).  peer_stage / peer_flags: HOST arrays of `world` device pointers
 * (entry `rank` local, the rest peer-mapped): the staging buffer of THIS epoch's parity
 * (the caller alternates two) and the flag array (`world` uint32, zero at start).
 * local_counter: one zeroed uint32 in local device memory.  epoch: 1, 2, 3, ...
 * Replaces ref:113 (`dist_optim.step`) for DP replicas; the reference has no all-reduce.
 */
int sdml_dp_allreduce_sgd(int n_tensors, void* const* params, void* const* grads,
                          void* const* moms, const long long* sizes, int rank, int world,
                          void* const* peer_stage, void* const* peer_flags, void* local_counter,
                          unsigned int epoch, float lr, float momentum, float weight_decay,
                          float grad_scale, void* stream);

/* The single-replica form of the same kernel for callers that must not allocate device
 * memory (a host runtime that owns the whole GPU heap, a loop that cannot afford
 * cudaMalloc's device-wide stall, a fixed memory plan): staging buffers, flag and counter
 * live in a 512.5 KiB static array inside the library on every GPU that loads it, so this
 * call performs no cudaMalloc / cudaFree.  Tensors are on the current device;
 * p -= lr * (m = momentum * m + grad_scale * g + weight_decay * p), gradients zeroed.
 * At most 65 536 fp32 elements per call (more is an error, nothing is launched).  The first
 * call on a device resolves the array's address there -- make it outside stream capture.
 * All calls on a device share the array: calls whose kernels overlap on one GPU race.
 * One replica only: a static array can be neither IPC-mapped nor peer-addressed.
 * This entry point has no counterpart in the reference.
 */
int sdml_dp_allreduce_sgd_static(int n_tensors, void* const* params, void* const* grads,
                                 void* const* moms, const long long* sizes, float lr,
                                 float momentum, float weight_decay, float grad_scale,
                                 void* stream);

/* Variants of sdml_dp_allreduce_sgd whose epoch lives in device memory (*epoch_dev = number of
 * completed calls; advance it with a stream-ordered increment after the launch), so a captured
 * launch can be replayed from a CUDA graph:
 *   _dev_epoch : staged protocol, staging buffers for odd / even epochs passed separately;
 *   sdml_dp_ll_allreduce_sgd : low-latency protocol -- every replica pushes (value, epoch) pairs
 *     into its slot of every peer's landing buffer (2 * world * total * 8 bytes, zeroed) and polls
 *     its own memory; no flags or fences.  The parameter set is one flat index space: all pushes
 *     of a thread are issued before its first poll (one NVLink trip per call, not per tensor).
 *     The last CTA to finish advances *epoch_dev (exit_counter: device uint32, zero at first), so a
 *     captured launch replays.  timeout_ns != 0 bounds every poll: {7, peer, epoch} is left in err
 *     (4 x uint32, host-visible, may be NULL) and the kernel finishes.  stamps (may be NULL):
 *     64 x 4 uint64 globaltimer probes per call.  The default of the Python front-end.
 * No counterpart in the reference (it has no data parallelism).
 */
int sdml_dp_allreduce_sgd_dev_epoch(int n_tensors, void* const* params, void* const* grads,
                                    void* const* moms, const long long* sizes, int rank, int world,
                                    void* const* peer_stage_odd, void* const* peer_stage_even,
                                    void* const* peer_flags, void* local_counter,
                                    const void* epoch_dev, float lr, float momentum,
                                    float weight_decay, float grad_scale, void* stream);
int sdml_dp_ll_allreduce_sgd(int n_tensors, void* const* params, void* const* grads,
                             void* const* moms, const long long* sizes, int rank, int world,
                             void* const* peer_buf, void* epoch_dev, void* exit_counter, float lr,
                             float momentum, float weight_decay, float grad_scale,
                             unsigned long long timeout_ns, void* err, void* stamps, void* stream);

/* ---- GEMM fused with tensor- / data-parallel collectives (csrc/gemm_tcgen05.cu) ----
 * sdml_gemm_bf16_comm: the CTA-pair tcgen05 GEMM with communication hooks.
 *   ag_world > 1 : A is the local row-stacked [M, K] operand buffer in symmetric memory with only
 *     this rank's rows valid; the other ranks' rows are pulled from ag_src[r] (HOST array of
 *     peer-mapped pointers to their buffers) by extra warps while the GEMM runs, local tiles
 *     first.  ag_flags: HOST array, entry r = rank r's flag array (>= world uint32, zeroed);
 *     ag_count: local chunk counters [world * ag_max_chunks] (zeroed, never reset); epoch 1, 2, ...
 *   rs_world > 1 : rows of the bf16 partial result owned by rank r are TMA-stored into rs_dst[r]
 *     (HOST array: the landing slot reserved for THIS rank on r, [M / rs_world, rs_ld] bf16);
 *     rs_done[r] = r's completion counter.  Follow with sdml_rs_finalize on every owner.
 * sdml_rs_finalize: waits for `expected` on `done` (calls * world * 4 * sdml_gemm_comm_grid()),
 *   then out = mask(relu(sum of `slots` landing slots + bias)).
 * sdml_wgrad_rs: dW = alpha * dz^T . x with the same reduce-scatter epilogue (DP gradient).
 * sdml_rs_finalize_sgd: owner side of the DP reduction: sums the replicas' slots, SGD on its rows
 *   of the fp32 master / momentum, refreshed bf16 rows stored into every replica's weight copy
 *   (wout[q], peer-mapped), +1 on ready[q] for every replica when done.
 * sdml_wait_counter: holds `stream` until *counter >= expected.
 * The reference has neither tensor nor data parallelism: no counterpart (north-star items).
 */
int sdml_gemm_bf16_comm(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                        long long ldb, long long ldc, int b_mn, const float* bias, const void* aux,
                        long long ld_aux, int relu, int out_fp32, float alpha, int ag_world,
                        int ag_rank, void* const* ag_src, void* const* ag_flags, void* ag_count,
                        int ag_max_chunks, int ag_chunk_rows, unsigned int epoch, int rs_world,
                        int rs_rank, void* const* rs_dst, long long rs_ld, void* const* rs_done,
                        void* stream);
int sdml_rs_finalize(const void* landing, int slots, long long slot_stride, long long ld_land,
                     int rows, int cols, const float* bias, const void* aux, long long ld_aux,
                     int relu, void* out, long long ld_out, const void* done, unsigned int expected,
                     void* stream);
int sdml_wgrad_rs(const void* dz, const void* x, int M, int N, int K, long long ld_dz,
                  long long ld_x, float alpha, int rs_world, int rs_rank, void* const* rs_dst,
                  long long rs_ld, void* const* rs_done, void* stream);
int sdml_rs_finalize_sgd(const void* landing, int slots, long long slot_stride, long long ld_land,
                         int rows, int cols, float* master, float* mom, long long ld_w, float lr,
                         float momentum, float weight_decay, void* const* wout, long long ld_wout,
                         int world, const void* done, unsigned int expected, void* const* ready,
                         void* exit_counter, void* stream);
int sdml_wait_counter(const void* counter, unsigned int expected, void* stream);
int sdml_gemm_comm_grid(void); /* CTAs launched by the comm kernels (enters `expected`) */

/* ---- native NCCL point-to-point engine (csrc/p2p_engine.cu) -----------------------
 * One channel = one NCCL communicator (normally 2 ranks: one direction of one stage boundary)
 * with its own high-priority side stream.  Replaces the reference's blocking RPC hop for the
 * activation (ref:47-49, :80) and dist-autograd's gradient RPCs (ref:112).  NCCL is bound at
 * run time from the libnccl already loaded in the process.
 *   sdml_p2p_unique_id   128-byte rendezvous token (create on one rank, distribute out of band)
 *   sdml_p2p_create      collective over the channel's ranks, on the CURRENT device
 *   sdml_p2p_enqueue     n_ops sends / receives as one NCCL group on the side stream; with
 *                        order_after != 0 it starts after everything queued on after_stream
 *                        (the default stream's handle is 0, hence the separate flag); yields a ticket
 *   sdml_p2p_wait_stream `stream` waits on the device for the ticket; the host does not block
 *   sdml_p2p_wait_host   bounded host wait: 0 done, 1 timed out, -1 NCCL / CUDA error
 *   sdml_p2p_destroy     abort != 0: ncclCommAbort (peer is gone) instead of a clean destroy
 */
typedef struct SdmlP2POp {
  int is_send;              /* 1 = send, 0 = receive */
  int peer;                 /* rank inside the channel's communicator */
  void* ptr;
  unsigned long long bytes;
} SdmlP2POp;
int sdml_p2p_nccl_version(void);
int sdml_p2p_unique_id(void* out128);
int sdml_p2p_create(const void* uid128, int nranks, int rank, void** handle);
int sdml_p2p_destroy(void* handle, int abort);
int sdml_p2p_enqueue(void* handle, const SdmlP2POp* ops, int n_ops, void* after_stream,
                     int order_after, long long* ticket_out);
int sdml_p2p_wait_stream(void* handle, long long ticket, void* stream);
int sdml_p2p_wait_host(void* handle, long long ticket, int timeout_ms);

/* ---- native op-list executor (csrc/executor.cu) ------------------------------------
 * One training step of one pipeline rank without returning to the caller's language: replaces
 * the reference's per-batch Python statements (ref:108-113).  `ops` are (kind, micro-batch)
 * pairs from sdml_sched_generate (include/sdml_schedule.h).
 *   sdml_exec_run_hooks        the control loop against caller-supplied hooks (any backend)
 *   sdml_exec_plan_*           GPU binding: every op is a cudaGraphExec_t (kind 0/1/2 = F/B/W of a
 *                              micro-batch, kind 3 = end of step), transfers use sdml_p2p_* channels
 *                              (dir 0 = activations, 1 = gradients) and per-micro-batch buffers
 *   sdml_exec_plan_step        issues every graph launch / receive / wait / send of the step
 */
typedef struct SdmlExecHooks {
  void* user;
  int (*compute)(void* user, int kind, int mb);
  int (*post_recv)(void* user, int dir, int mb, long long* ticket);
  int (*wait_recv)(void* user, int dir, int mb, long long ticket);
  int (*send)(void* user, int dir, int mb, long long* ticket);
  int (*wait_send)(void* user, int dir, int mb, long long ticket);
  int (*end_step)(void* user);
} SdmlExecHooks;
int sdml_exec_run_hooks(const int* ops, int n_ops, int is_first, int is_last, int lookahead,
                        const SdmlExecHooks* hooks);
int sdml_exec_plan_create(const int* ops, int n_ops, int n_mb, int is_first, int is_last,
                          void* compute_stream, void** plan_out);
int sdml_exec_plan_set_graph(void* plan, int kind, int mb, void* graph_exec);
int sdml_exec_plan_set_channel(void* plan, int dir, int is_send, void* p2p_handle);
int sdml_exec_plan_set_buffer(void* plan, int dir, int is_send, int mb, void* ptr,
                              unsigned long long bytes);
int sdml_exec_plan_step(void* plan);
int sdml_exec_plan_destroy(void* plan);

/* ---- MX-fp8 (csrc/gemm_mxfp8.cu) -------------------------------------------------
 * sdml_quant_mxfp8: bf16 [rows, cols] -> E4M3 bytes [rows, cols] + one UE8M0 scale per 32
 *   columns, scales stored as 512-byte atoms in the layout tcgen05.cp copies into TMEM
 *   (sdml_mxfp8_sf_bytes gives the array size; zero it once).  cols % 128 == 0.
 * sdml_quant_mxfp8_dual: one read of bf16 [rows, cols] -> the row-wise copy above (dst_row,
 *   optional) and / or the TRANSPOSED copy dst_col [cols, rows] scaled along the original rows
 *   (optional): the K-major operands of the dgrad / wgrad GEMMs, whose contraction runs over the
 *   other dimension.  rows, cols % 128 == 0.  A row range of a larger transposed operand is
 *   filled by offsetting dst_col / sf_col (see csrc/gemm_mxfp8.cu); groups_col =
 *   ceil(total rows of the transposed operand / 128).
 * sdml_gemm_mxfp8: C[M,N] = epilogue(alpha * deq(A)[M,K] . deq(B)[N,K]^T) on
 *   tcgen05.mma.kind::mxf8f6f4.block_scale; K % 128 == 0, lda / ldb in bytes (multiples of 16);
 *   accumulate (fp32 output only) adds into C.
 * fp8 form of ref:63,75 (fc1 Linear+bias+ReLU) and of its autograd backward for the wide-MLP
 * configs.
 */
long long sdml_mxfp8_sf_bytes(int rows, int cols);
int sdml_quant_mxfp8(const void* src, long long ld_src, int rows, int cols, void* dst,
                     long long ld_dst, void* sf, void* stream);
int sdml_gemm_mxfp8(const void* A, const void* sfa, const void* B, const void* sfb, void* C, int M,
                    int N, int K, long long lda, long long ldb, long long ldc, const float* bias,
                    const void* aux, long long ld_aux, int relu, int out_fp32, int accumulate,
                    float alpha, void* stream);
/* CTA-pair (cta_group::2) kernel, 256 x bn output tile (bn = 192 or 128): half the L2 -> smem
 * operand traffic per FLOP of the single-CTA kernel.  Same contract; scale arrays 16-byte aligned. */
int sdml_gemm_mxfp8_2cta(const void* A, const void* sfa, const void* B, const void* sfb, void* C,
                         int M, int N, int K, long long lda, long long ldb, long long ldc,
                         const float* bias, const void* aux, long long ld_aux, int relu, int out_fp32,
                         int accumulate, float alpha, int bn, void* stream);
/* MX-fp8 GEMM -> reduce-scatter (tensor-parallel row-parallel layer with fp8 operands): rows of the
 * bf16 partial result alpha * A[M,K] . B[N,K]^T owned by rank r are TMA-stored into rs_dst[r] (HOST
 * array: the peer-mapped landing slot reserved for this rank in r's memory, bf16 [M / rs_world, rs_ld]);
 * every epilogue warp then adds 1 to every owner's counter in rs_done (HOST array).  Sum the slots
 * with sdml_rs_finalize (expected count per call: rs_world * 4 * sdml_gemm_comm_grid()).
 * M % (32 * rs_world) == 0, K % 128 == 0, scale arrays 16-byte aligned.  ref:75-77 sharded over GPUs. */
int sdml_gemm_mxfp8_2cta_rs(const void* A, const void* sfa, const void* B, const void* sfb, int M, int N,
                            int K, long long lda, long long ldb, float alpha, int rs_world, int rs_rank,
                            void* const* rs_dst, long long rs_ld, void* const* rs_done, void* stream);

/* Same GEMM (256 x 192 tiles) with FUSED OUTPUT QUANTISATION: the epilogue also writes the MX-fp8
 * copies of the bf16 result that the next GEMMs consume -- q_row: C [M, N] scaled along N,
 * q_col: C^T [N, M] scaled along M (pre-offset into a step-wide transposed buffer like
 * sdml_quant_mxfp8_dual's dst_col) -- bit-identical to sdml_quant_mxfp8_dual(C).  Either may be
 * NULL.  M, N, K % 128 == 0.  (Written in round 1, first hardware run pending.) */
int sdml_gemm_mxfp8_2cta_qout(const void* A, const void* sfa, const void* B, const void* sfb, void* C,
                              int M, int N, int K, long long lda, long long ldb, long long ldc,
                              const float* bias, const void* aux, long long ld_aux, int relu, float alpha,
                              void* q_row, long long ld_qrow, void* sf_qrow, void* q_col,
                              long long ld_qcol, void* sf_qcol, int groups_qcol, void* stream);
int sdml_quant_mxfp8_dual(const void* src, long long ld_src, int rows, int cols, void* dst_row,
                          long long ld_row, void* sf_row, void* dst_col, long long ld_col,
                          void* sf_col, int groups_col, void* stream);
/* Same, with the row-wise copy written into a row range of a larger operand (dst_row / sf_row already
 * point at the range, groups_row = ceil(total_rows / 128)): a tensor-parallel rank quantises only its
 * shard of a gathered operand.  No counterpart in the reference. */
int sdml_quant_mxfp8_dual_at(const void* src, long long ld_src, int rows, int cols, void* dst_row,
                             long long ld_row, void* sf_row, int groups_row, void* dst_col,
                             long long ld_col, void* sf_col, int groups_col, void* stream);

/* ---- the reference model as fused kernels (csrc/lenet_fused.cu) ------------------
 * sdml_lenet_s0_fwd : ref:42-46 (conv1-pool-relu-conv2-Dropout2d-pool-relu-flatten), one CTA
 *   per sample; saves z1 [B,1440], idx1 [B,1440] u8, idx2 [B,320] u8, mask2 [B,20].
 *   zero_this / bump_counter: optional housekeeping done by the launch (clear a float,
 *   advance an int64 device counter by bump_by).
 * sdml_lenet_s1     : ref:75-79 + ref:111 forward AND backward (loss_sum += loss_scale * CE,
 *   dW2/db2/db1 accumulated, dz3 for stage 0, dW1 by a second kernel, optionally on
 *   wgrad_stream); train == 0: loss_sum / correct only (ref:126-128).
 * sdml_lenet_s0_bwd : backward of stage 0, gradients accumulated with atomics.
 */
int sdml_lenet_s0_fwd(const float* x, const float* w1, const float* b1, const float* w2,
                      const float* b2, int B, float* z3, float* z1_save, void* idx1, void* idx2,
                      float* mask2, float drop_p, long long seed, const long long* offset_dev,
                      long long offset_imm, float* zero_this, long long* bump_counter,
                      long long bump_by, void* stream);
int sdml_lenet_s1(const float* z3, const void* targets, const float* w1, const float* b1,
                  const float* w2, const float* b2, int B, int train, float grad_scale,
                  float drop_p, long long seed, const long long* offset_dev, long long offset_imm,
                  float* loss_sum, float loss_scale, int* correct, float* dz3, float* dz4_scratch,
                  float* dw1, float* db1, float* dw2, float* db2, long long* bump_counter,
                  long long bump_by, void* stream, void* wgrad_stream);
int sdml_lenet_s0_bwd(const float* x, const float* w2, const float* z3, const float* dz3,
                      const float* z1_save, const void* idx1, const void* idx2, const float* mask2,
                      int B, float* dw1, float* db1, float* dw2, float* db2, void* stream);

/* ---- symmetric (peer-mapped) buffers (csrc/symm.cu) ---------------------------- */
int sdml_symm_alloc(long long bytes, void** out_ptr);
int sdml_symm_free(void* ptr);
int sdml_symm_export(void* ptr, unsigned char* handle_out /* 64 bytes */);
int sdml_symm_import(const unsigned char* handle, void** out_ptr);
int sdml_symm_unimport(void* ptr);
/* strided device-to-device copy on the copy engine; either side may be peer-mapped */
int sdml_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch, long long width,
                        long long height, void* stream);

/* ---- stage boundary over symmetric memory: push + flag (csrc/boundary.cu) ------
 * Replaces the reference's blocking activation / gradient RPCs (ref:47-49, 71, 80, 112) on GPUs
 * of one NVLink domain: the sender writes the receiver's peer-mapped slot (a GEMM epilogue that
 * stores to the slot, or sdml_bnd_push) and publishes a 32-bit flag; the receiver's stream waits
 * for it.  `cnt` is a device uint32 holding the number of completed rounds (steps), so captured
 * launches replay; comparisons use the signed difference (wrap-safe).  All calls only enqueue.
 *   wait  : stream continues when *flag >= *cnt + add, or after timeout_ns (0 = never), in which
 *           case {code, seen, wanted, 0} goes to err_dev (device, 4 x uint32, zero-initialised; once
 *           non-zero every later wait returns at once) and err_host (pinned host, may be NULL).
 *   signal: *flag_peer = *cnt + add after a system-scope fence (flag_peer may be NULL); then
 *           *cnt += bump.
 *   push  : [ack != NULL: wait *ack >= *cnt]; copy `bytes` (multiple of 16, 16-byte aligned) into
 *           the peer's slot; *flag_peer = *cnt + add.  <= 128 KiB: one single-CTA kernel; larger:
 *           copy engine + signal kernel.
 */
int sdml_bnd_wait(const void* flag, const void* cnt, unsigned int add, unsigned long long timeout_ns,
                  void* err_dev, void* err_host, unsigned int code, void* stream);
int sdml_bnd_signal(void* flag_peer, void* cnt, unsigned int add, unsigned int bump, void* stream);
int sdml_bnd_push(void* dst_peer, const void* src, long long bytes, void* flag_peer, const void* cnt,
                  unsigned int add, const void* ack, unsigned long long timeout_ns, void* err_dev,
                  void* err_host, unsigned int code, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SDML_KERNELS_H_ */
