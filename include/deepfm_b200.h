/* deepfm_b200 C ABI — every entry point of ops/_lib/libdeepfm_b200.so (sm_100a only).
 *
 * Conventions: all functions return 0 on success or a DFM_ERR_* code; dfm_last_error() returns the message of
 * the calling thread's last failure. Pointers are device pointers unless noted. `stream` is a cudaStream_t.
 * Kernels are asynchronous on `stream`. Nothing here allocates device memory unless stated.
 * "Reference" citations: PS = 1-ps-cpu/DeepFM-dist-ps-for-multipleCPU-multiInstance.py,
 *                        HVD = 2-hvd-gpu/DeepFM-hvd-tfrecord-vectorized-map.py  (of the upstream repository).
 */
#ifndef DEEPFM_B200_H_
#define DEEPFM_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define DFM_OK 0
#define DFM_ERR_CUDA 1
#define DFM_ERR_ARG 2
#define DFM_ERR_UNSUPPORTED 3

const char* dfm_last_error(void);
int dfm_abi_version(void);
int dfm_device_is_sm100(void);

/* ---- embedding + FM (replaces embedding_lookup x2 + FM reductions, PS:177-187,203) ---- */
int dfm_emb_fm_fwd(const int* ids, const float* vals, const float* fm_w, const float* fm_v, void* x0_bf16, int ldx,
                   float* S, float* yfm, int B, int F, int K, int N, float* zero_scalars, int n_zero, void* stream);
int dfm_emb_bwd(const int* ids, const float* vals, const float* fm_v, const float* S, const float* dy,
                const void* dx0_bf16, int ldx, float* g_w, float* g_v, int B, int F, int K, int N, void* stream);

/* Deterministic twin of dfm_emb_bwd (csrc/embed_det.cu; TF segment-sums duplicate ids before applying them): sort_ids sorts
 * the batch's lookups by row id into a caller-owned workspace (needs only ids / vals: side stream, step start), emb_bwd_det
 * sums every run of equal ids in a fixed order and gives each touched row ONE plain store (no atomics; g_w / g_v are
 * overwritten, not accumulated). workspace: dfm_emb_bwd_det_workspace_bytes(B, F, K) bytes of device memory. */
long long dfm_emb_bwd_det_workspace_bytes(int B, int F, int K);
int dfm_emb_sort_ids(const int* ids, const float* vals, int B, int F, int K, int N, void* workspace,
                     long long workspace_bytes, void* stream);
int dfm_emb_bwd_det(const float* vals, const float* fm_v, const float* S, const float* dy, const void* dx0_bf16, int ldx,
                    float* g_w, float* g_v, int B, int F, int K, int N, void* workspace, long long workspace_bytes,
                    void* stream);

/* emb_bwd with owner routing for ZeRO-1 data parallelism: every row gradient is added straight into the flat gradient
 * buffer of the rank that reduces it (g_ptrs: `world` peer-mapped buffers; element e belongs to rank e / chunk), so the
 * reduce-scatter of the embedding gradient happens inside this kernel (replaces that share of Horovod's all-reduce, HVD:262). */
int dfm_emb_bwd_routed(const int* ids, const float* vals, const float* fm_v, const float* S, const float* dy,
                       const void* dx0_bf16, int ldx, void* const* g_ptrs, long long gw_off, long long gv_off,
                       long long chunk, int world, int B, int F, int K, int N, void* stream);

/* ---- tcgen05 MLP GEMMs (replaces fully_connected fwd/bwd, PS:212-223) ----
 * mode 0 forward: out[M,N] bf16 = dropout(relu(A[M,K] B[N,K]^T + bias));  mode 1 dgrad: out[M,N] bf16 =
 * (A[M,K] B[K,N]) * (mask != 0) * mask_scale;  mode 2 wgrad: out[M,N] f32 += A[K,M]^T B[K,N].
 * The plan is host memory (holds the TMA descriptors); create once per buffer set, run many times. */
int dfm_gemm_plan_create(void** plan_out, int mode, const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                         void* out, int ldo, const float* bias, int relu, float keep_prob, unsigned long long seed,
                         int layer, int row_offset, const unsigned long long* step_ptr, const void* mask, int ldmask,
                         float mask_scale, int split_k);
int dfm_gemm_plan_info(void* plan, int* bn, int* grid_x, int* grid_y, int* grid_z);
int dfm_gemm_plan_set_row_offset(void* plan, int row_offset);
int dfm_gemm_plan_set_overwrite(void* plan, int on); /* wgrad plans with split_k = 1: out = (plain stores), not += */
int dfm_gemm_plan_set_exclusive(void* plan, int on); /* dgrad plans: no GEMM runs beside it -> whole-SM stage budget */
int dfm_gemm_plan_set_colsum(void* plan, float* colsum); /* dgrad plans: colsum[n] += sum_m out[m,n] (bias gradient) */
int dfm_gemm_plan_set_trace_id(void* plan, int id); /* timeline id (trace build only, see dfm_trace_*) */
int dfm_gemm_plan_set_out_f32(void* plan, int on); /* forward plans: fp32 output (pre-batch-norm activations) */
void dfm_gemm_plan_destroy(void* plan);
int dfm_gemm_run(void* plan, void* stream);

/* ---- output head: logit sum + sigmoid + log-loss + gradients (PS:221-232,244) ---- */
int dfm_logit_loss(const void* x_bf16, int ldx, int H, const float* w_out, const float* b_out, const float* fm_bias,
                   const float* yfm, const float* labels, int B, float inv_batch, float mask_scale, int train,
                   float* prob, float* logit, float* dy, void* dz_bf16, int lddz, float* loss_sum, float* g_b_out,
                   float* g_fm_bias, void* stream);
/* Same, plus (train mode, both pointers non-NULL) g_b_top[j] += sum_b dZ[b,j] and g_w_out[j] += sum_b dy[b] x[b,j]. */
int dfm_logit_loss_cols(const void* x_bf16, int ldx, int H, const float* w_out, const float* b_out, const float* fm_bias,
                        const float* yfm, const float* labels, int B, float inv_batch, float mask_scale, int train,
                        float* prob, float* logit, float* dy, void* dz_bf16, int lddz, float* loss_sum, float* g_b_out,
                        float* g_fm_bias, float* g_b_top, float* g_w_out, void* stream);
/* out[j] += sum_b a[b,j] * (w ? w[b] : 1): bias gradients (w = NULL) and g_w_out = X_top^T dy (w = dy). */
int dfm_colsum_bf16(const void* a_bf16, int lda, int B, int H, const float* w, float* out, void* stream);

/* ---- batch norm after ReLU (+ dropout) for --batch_norm (batch_norm_layer, PS:288-292; applied at PS:215-218) */
int dfm_bn_fwd(const float* z, int ldz, void* out_bf16, int ldo, int B, int H, const float* gamma, const float* beta,
               float* mov_mean, float* mov_var, float* save_mean, float* save_rstd, float decay, int train,
               float keep_prob, unsigned long long seed, int layer, int row_offset, const unsigned long long* step_ptr,
               void* stream);
int dfm_bn_bwd(void* dy_inout_bf16, int ldd, const float* z, int ldz, int B, int H, const float* gamma,
               const float* save_mean, const float* save_rstd, float* g_gamma, float* g_beta, void* stream);

/* ---- streaming AUC histogram (tf.metrics.auc, PS:249-251) ---- */
int dfm_auc_hist(const float* prob, const float* labels, int B, const float* thr, int T, float* hist, void* stream);

/* ---- flat multi-tensor optimizer with TF-1 update rules (PS:260-269); opt: 0 Adam 1 Adagrad 2 Momentum 3 ftrl 4 GD.
 * `done_counter`: 4 bytes of zero-initialised device scratch owned by the caller. */
int dfm_flat_opt_step(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1, float beta2,
                      float eps, float momentum, float grad_scale, long long decay_begin, long long decay_end, float l2,
                      long long shadow_begin, long long shadow_end, void* shadow_bf16, unsigned long long* step_ptr,
                      float* reg_loss, unsigned int* done_counter, void* stream);
/* Same sweep; additionally the last CTA stores `pub_n` floats read from `pub_src` (device) into record (step - 1) %
 * pub_ring_len of `pub_ring`, a device-visible alias of page-locked HOST memory: the step's result (loss sum, L2 term)
 * reaches the host with no D2H copy operation in the stream. pub_ring == NULL disables it. */
int dfm_flat_opt_step_pub(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1,
                          float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                          long long decay_end, float l2, long long shadow_begin, long long shadow_end, void* shadow,
                          unsigned long long* step_ptr, float* reg_loss, unsigned int* done_counter,
                          const float* pub_src, int pub_n, float* pub_ring, int pub_ring_len, void* stream);
/* Fused MLP forward (csrc/mlp_fused.cu): all L <= 4 layers X_{l+1} = dropout(relu(X_l W_l^T + b_l)) in one kernel; a
 * thread-block cluster per 128 batch rows keeps the activations in distributed shared memory between layers. W[l] bf16
 * [H_l, K_l] (ldw[l]), out[l] bf16 [M, H_l] (ldo[l]), keep[l] dropout keep probabilities, layer_ids[l] the Philox stream
 * ids the per-layer GEMM plans would use (identical masks). Widths must be multiples of 16 and <= 448, leading dimensions
 * multiples of 8; otherwise DFM_ERR_ARG and the caller keeps the per-layer plans. Reference: PS:212-226 / HVD:204-218. */
int dfm_mlp_fwd_plan_create(void** plan_out, const void* X, int ldx, int M, int K0, int L, const void* const* W,
                            const int* ldw, const int* H, void* const* out, const int* ldo, const float* const* bias,
                            int relu, const float* keep, unsigned long long seed, const int* layer_ids, int row_offset,
                            const unsigned long long* step_ptr);
int dfm_mlp_fwd_plan_info(void* plan, int* bn, int* cluster, int* m_tiles);
int dfm_mlp_fwd_plan_set_trace_id(void* plan, int id);
int dfm_mlp_fwd_plan_set_debug(void* plan, unsigned long long* dbg); /* development aid: [cta][32] globaltimer stamps */
int dfm_mlp_fwd_run(void* plan, void* stream);
void dfm_mlp_fwd_plan_destroy(void* plan);

/* Flag-wait watchdog: a kernel that waits longer than DFM_FLAG_TIMEOUT_NS (20 s) for a peer's flag records (missing rank,
 * barrier id, expected sequence number, value seen) in 8 words of pinned host memory and traps. dfm_watchdog_record returns
 * that record (allocated on first use, NULL on failure); dfm_watchdog_message formats it into buf and returns 1 if a wait
 * has timed out in this process, else 0 — it works after the trap, when the CUDA context is already dead. */
unsigned int* dfm_watchdog_record(void);
int dfm_watchdog_message(char* buf, int cap);
int dfm_cast_f32_bf16(const float* in, void* out_bf16, long long n, void* stream);

/* ---- split form of the same dense optimizer step (csrc/split_opt.cu) for replicated tables: the whole-table part of the
 * reference's update (dense ApplyAdam on fm_w / fm_v because of the whole-table L2 term, PS:245-246,269) leaves the
 * critical path. touched_bitmap marks the rows a batch looks up; sweep_untouched updates every OTHER row (their gradient is
 * l2*w, nothing from backward is needed) concurrently with forward/backward; tail_opt finishes with the touched rows and the
 * dense ranges d0/d1 (fm_bias block, MLP block), zeroing the gradients it consumes. Results equal dfm_flat_opt_step.
 * union_rowlist (one CTA) ORs `world` per-rank bitmaps of dfm_bitmap_words(N) words (bm_in: host array of device pointers;
 * world == 1: the live bitmap, cleared when clear_inputs != 0) into bm_union and lists the set rows in ascending order
 * (rows: N ints, *row_count entries) for the tail and the exchange. reg_acc: one zero-initialised device float (sum of w^2
 * over the decay range; reset by the finalising tail, which writes *reg_loss = 0.5*l2*sum). */
int dfm_bitmap_words(int N);
int dfm_touched_bitmap(const int* ids, int B, int F, int N, unsigned int* bitmap, void* stream);
int dfm_union_rowlist(const void* const* bm_in, int world, int clear_inputs, int N, unsigned int* bm_union, int* rows,
                      int* row_count, void* stream);
int dfm_union_rowlist_prefix(const void* const* bm_in, int world, int clear_inputs, int N, unsigned int* bm_union, int* rows,
                             int* row_count, int* word_prefix, void* stream); /* + word_prefix[w] = rows listed below 32 w */
int dfm_sweep_untouched(float* w, float* s0, float* s1, int opt, float lr, float beta1, float beta2, float eps,
                        float momentum, float l2, long long fmw_off, long long fmv_off, int N, int K,
                        const unsigned int* bm_union, unsigned long long* step_ptr, float* reg_acc, int grid_blocks,
                        void* stream);
int dfm_tail_opt(float* w, float* g, float* s0, float* s1, int opt, float lr, float beta1, float beta2, float eps,
                 float momentum, float grad_scale, float l2, long long fmw_off, long long fmv_off, int N, int K,
                 long long d0_begin, long long d0_end, long long d1_begin, long long d1_end, long long shadow_begin,
                 long long shadow_end, void* shadow_bf16, const int* rows, const int* row_count,
                 unsigned long long* step_ptr, float* reg_acc, float* reg_loss, unsigned int* done_counter,
                 float* loss_sum, const float* loss_parts, int world, const float* pub_src, int pub_n, float* pub_ring,
                 int pub_ring_len, int do_dense, int do_sparse, int finalize, void* stream);

/* ---- gradient all-reduce fused with the optimizer step over NVLink peer memory (replaces Horovod's NCCL all-reduce
 * HVD:262 + ApplyAdam PS:269). Pointer tables: host arrays of `world` device pointers into symmetric memory; *_mc are
 * NVLS multicast mappings or NULL. Every rank must launch with the same grid_blocks (<= dfm_allreduce_opt_max_blocks()).
 * flags: dfm_allreduce_opt_flag_words() zero-initialised uint32 per rank; reg_parts: 8 floats per rank; seq_ptr: one
 * zero-initialised device word per rank (barrier sequence number, advanced by the kernel). */
int dfm_allreduce_opt_flag_words(void);
int dfm_allreduce_opt_max_blocks(void);
int dfm_allreduce_opt_step(void* const* g_ptrs, void* const* p_ptrs, void* const* sh_ptrs, void* const* flag_ptrs,
                           void* const* reg_ptrs, void* g_mc, void* p_mc, void* sh_mc, float* s0, float* s1, int rank,
                           int world, long long slice_begin, long long slice_end, int opt, float lr, float beta1,
                           float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                           long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                           unsigned long long* step_ptr, float* reg_local, unsigned int* done_counter,
                           unsigned int* seq_ptr, int use_nvls, int grid_blocks, void* stream);

/* Same step; additionally (loss_ptrs: `world` pointers to 8-float inboxes in symmetric memory) every rank's log-loss sum
 * is exchanged at barrier A (reg_extra: optional L2 term of this rank from elsewhere, added to its partial) and the last
 * CTA writes (global loss sum, complete L2 term) to scalars_out[0..1]; [local_begin, local_end): elements whose gradients
 * dfm_emb_bwd_routed already reduced into their owner's buffer (read locally and zeroed, no NVLink load); finalize = 0: an
 * early bucket of the step (launched per bucket so it overlaps backward) with its own flags / seq / scratch words —
 * it neither advances the step nor publishes and into record
 * (step - 1) % pub_ring_len of pub_ring (page-locked HOST memory, 2 floats per record; NULL = no publication). */
int dfm_allreduce_opt_step_pub(void* const* g_ptrs, void* const* p_ptrs, void* const* sh_ptrs, void* const* flag_ptrs,
                               void* const* reg_ptrs, void* g_mc, void* p_mc, void* sh_mc, float* s0, float* s1, int rank,
                               int world, long long slice_begin, long long slice_end, int opt, float lr, float beta1,
                               float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                               long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                               unsigned long long* step_ptr, float* reg_local, unsigned int* done_counter,
                               unsigned int* seq_ptr, int use_nvls, int grid_blocks, void* const* loss_ptrs,
                               const float* loss_local, const float* reg_extra, float* scalars_out, float* pub_ring,
                               int pub_ring_len, long long local_begin, long long local_end, int finalize, void* stream);

/* ---- data-parallel exchange for replicated tables that ships only what differs between ranks (csrc/sparse_allreduce.cu;
 * replaces Horovod's dense NCCL all-reduce, HVD:262). bitmap_exchange: this rank's touched-row bitmap (bm_own, cleared
 * afterwards) into slot [rank] of every rank's [8][words] inbox + flag barrier; then dfm_union_rowlist / dfm_sweep_untouched
 * (above). sparse_allreduce: barrier A (deposits *loss_sum in loss_parts[rank] of every peer), in-place all-reduce of the
 * dense ranges d0 | d1 and of the fm_w / fm_v entries of rows[0 .. *row_count) — each item reduced by one rank
 * (multimem.ld_reduce, or peer loads in rank order) and stored into every rank's buffer — then barrier B. Pointer tables:
 * host arrays of `world` device pointers into symmetric memory; *_mc: NVLS multicast mappings or NULL; flags:
 * dfm_sparse_allreduce_flag_words() zeroed uint32 per rank; seq_*: one zeroed device word each (advanced by the kernels);
 * done_counter: one zeroed scratch word per entry point. */
int dfm_sparse_allreduce_flag_words(void);
int dfm_bitmap_exchange(void* const* bm_ptrs, void* bm_mc, void* const* flag_ptrs, unsigned int* bm_own, int N, int rank,
                        int world, unsigned int* done_counter, unsigned int* seq_c, int use_nvls, void* stream);
int dfm_sparse_allreduce(void* const* g_ptrs, void* g_mc, void* const* flag_ptrs, void* const* loss_ptrs, int rank,
                         int world, long long fmw_off, long long fmv_off, int N, int K, long long d0_begin,
                         long long d0_end, long long d1_begin, long long d1_end, const int* rows, const int* row_count,
                         const float* loss_sum, unsigned int* done_counter, unsigned int* seq_ab, int use_nvls,
                         int grid_blocks, void* stream);

/* ---- state broadcast from one rank into every replica (replaces BroadcastGlobalVariablesHook(0), HVD:372) */
int dfm_bcast_f32(void* const* ptrs, void* mc, long long n, int rank, int world, int use_nvls, void* stream);

/* ---- row-sharded embedding tables with the all-to-all fused into the kernels (no counterpart in the reference beyond
 * the partitioner advice of PDF p32). Pointer tables: host arrays of `world` device pointers into symmetric
 * memory (table shards v/w, gradient shards gv/gw, touched-row bitmaps, barrier flags of dfm_sharded_flag_words() words).
 * Row id lives on rank id % world at local row id / world. grid_blocks must be identical on every rank (<= 1184).
 * do_barrier=1 inside a training step (wait for every owner's row updates), 0 for single-rank evaluation. */
int dfm_sharded_flag_words(void);
int dfm_emb_fm_fwd_a2a(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                       int rank, int world, const int* ids, const float* vals, void* x0_bf16, float* e32, int ldx,
                       float* S, float* yfm, int B, int F, int K, long long N, const unsigned int* seq_ptr,
                       int do_barrier, int grid_blocks, void* stream);
int dfm_emb_bwd_a2a(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                    int rank, int world, const int* ids, const float* vals, const float* e32, const float* S,
                    const float* dy, const void* dx0_bf16, int ldx, int B, int F, int K, long long N, void* stream);
int dfm_sparse_row_opt(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                       int rank, int world, float* v_s0, float* v_s1, float* w_s0, float* w_s1, long long rows_local,
                       int K, int opt, float lr, float beta1, float beta2, float eps, float momentum, float l2,
                       const unsigned long long* step_ptr, int step_offset, unsigned int* seq_ptr,
                       unsigned int* done_counter, float* reg_loss, int grid_blocks, void* stream);

/* push flavour of the fused all-to-all (csrc/sharded_push.cu): requests out, rows pushed back with P2P stores, gradient
 * rows pushed into owner inboxes. cap = B*F; cnt_local = world ints zeroed per step; grid_blocks identical on all ranks. */
int dfm_a2a_push_flag_words(void);
int dfm_a2a_push_forward(void* const* req, void* const* cnt, void* const* rows, void* const* wv, void* const* gin,
                         void* const* gwin, void* const* fl, int rank, int world, long long cap, const int* ids,
                         const float* vals, const float* V, const float* Wt, int* cnt_local, int* pos_of_pair,
                         unsigned int* done_counter, void* x0_bf16, float* e32, int ldx, float* S, float* yfm, int B, int F,
                         int K, long long N, const unsigned int* seq_ptr, int grid_blocks, void* stream);
int dfm_a2a_push_backward(void* const* req, void* const* cnt, void* const* rows, void* const* wv, void* const* gin,
                          void* const* gwin, void* const* fl, int rank, int world, long long cap, const int* ids,
                          const float* vals, const int* pos_of_pair, const float* e32, const float* S, const float* dy,
                          const void* dx0_bf16, int ldx, float* GV, float* GW, void* bitmap, int B, int F, int K,
                          const unsigned int* seq_ptr, int grid_blocks, void* stream);

/* ---- GPU-side tf.train.Example parser (replaces the ParseExample op, HVD:78-83, when the host parser is the bottleneck):
 * blob = serialized records back to back, offsets[n+1]; *err = 0 on entry, 1 + index of the first bad record after. */
int dfm_example_parse(const void* blob, const int* offsets, int n, int F, float* labels, int* ids, float* vals, int* err,
                      void* stream);

/* ---- runtime switches / host helpers */
int dfm_pdl_enabled(void);            /* programmatic dependent launch between the step kernels (default on; DFM_PDL=0) */
void dfm_set_pdl(int on);
long long dfm_set_l2_persist(void* base, unsigned long long bytes); /* persisting-L2 window over the training state;
                                                                         returns bytes set aside, 0 = off/unsupported */
int dfm_host_register(void* ptr, unsigned long long bytes, void** dev_ptr); /* page-lock + map the PS store (HOST pointer) */
/* zero-copy parameter-server pull: out[i,:] = table[ids[i],:], table = device alias of the host store (X3 / row 11) */
int dfm_ps_gather(const float* table_dev, const long long* ids, int n, int K, float* out, void* stream);
int dfm_host_unregister(void* ptr);
/* Pipelined parameter-server worker (parallel/ps.py): the batch's unique rows are a device-resident list with a
 * device-resident count (dfm_touched_bitmap + dfm_union_rowlist_prefix), so nothing waits for a size on the host.
 * remap_ids: inv[p] = position of ids[p] in the list (out-of-range ids -> oob); pull_list: fm_v rows + fm_w scalars of the
 * listed rows out of the mapped host store; push_list: the listed gradient rows, the row ids and the count into page-locked
 * host buffers (device copies zeroed). No reference counterpart beyond the gRPC pulls/pushes of X3. */
int dfm_ps_remap_ids(const int* ids, long long n, int N, const unsigned int* bm_union, const int* word_prefix, int oob,
                     int* inv, void* stream);
int dfm_ps_pull_list(const float* table_v_dev, const float* table_w_dev, const int* rows, const int* count, int max_rows,
                     int K, float* out_v, float* out_w, void* stream);
int dfm_ps_push_list(float* g_v, float* g_w, const int* rows, const int* count, int max_rows, int K, float* host_v,
                     float* host_w, int* host_rows, int* host_count, void* stream);

/* ---- kernel timeline (csrc/common.cuh): %globaltimer stamps per kernel, compiled in only by the trace build
 * (`python -m deepfm_b200.build --trace` -> libdeepfm_b200_trace.so). out: dfm_trace_words() words, 4 per trace id:
 * first CTA entered / first CTA past griddepcontrol.wait / last CTA at its end / phase stamp. No reference counterpart. */
int dfm_trace_enabled(void);
int dfm_trace_words(void);
int dfm_trace_reset(void);
int dfm_trace_read(unsigned long long* out);

/* ---- host-side step driver (csrc/step_driver.cu): one call issues a whole host-fed step - pinned H2D copy into input slot
 * `slot` on copy_stream (after the slot's previous reader), compute_stream waits for it, cudaGraphLaunch of the step's
 * graph, D2H of the result scalars. Asynchronous; host_src must be page-locked. copy_done: 1 = the last copy into the slot
 * has completed (source buffer reusable), 0 = in flight. */
int dfm_driver_create(void** handle, int nslots);
int dfm_driver_destroy(void* handle);
int dfm_driver_step(void* handle, int slot, const void* host_src, void* dev_dst, unsigned long long bytes,
                    void* graph_exec, void* copy_stream, void* compute_stream, const void* dev_scalars,
                    void* host_scalars, unsigned long long scalar_bytes);
int dfm_driver_copy_done(void* handle, int slot);
int dfm_driver_copy_wait(void* handle, int slot);
unsigned long long dfm_driver_steps(void* handle);

/* ---- flat optimizer step that allocates nothing (no counterpart in the reference: TF owns its own allocator).
 * Same update as dfm_flat_opt_step, but the scratch it needs sits in a static device array inside the library
 * (262,400 bytes per GPU, held until exit) instead of being supplied by the caller. For callers that cannot call
 * cudaMalloc: the GPU's memory already belongs to another allocator or a memory cap, or the call site cannot afford
 * the device-wide serialisation of cudaMalloc/cudaFree. Limit: n <= dfm_flat_opt_static_max_elems() (64 Mi);
 * larger n returns DFM_ERR_ARG. The first call on a device does the per-device set-up and should be made before
 * any stream capture begins. All calls on a device share the array; calls that overlap on the GPU race. */
long long dfm_flat_opt_static_max_elems(void);
long long dfm_flat_opt_static_scratch_bytes(void);
int dfm_flat_opt_step_static(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1,
                             float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                             long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                             void* shadow_bf16, unsigned long long* step_ptr, float* reg_loss, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFM_B200_H_ */
