/* C ABI of libdistriflow_b200.so (distriflow_b200/ops/libdistriflow_b200.so).
 *
 * Every function returns 0 on success; on failure a non-zero code, with a message from dfb_last_error().
 * Pointers are device pointers unless stated otherwise.  `stream` is a cudaStream_t passed as void*.
 * dtype codes: 0 = fp32, 1 = bf16, 2 = fp16.  activation codes: 0 = none, 1 = relu, 2 = gelu (tanh form).
 * The reference (/root/reference) has no native code; each entry names the tf.js call site it replaces.
 */
#ifndef DISTRIFLOW_B200_H_
#define DISTRIFLOW_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dfb_last_error(void);
int dfb_version(void);

/* ---- Dense GEMM, tcgen05 + TMEM + TMA (csrc/gemm_sm100.cu) --------------------------------------------
 * D[M,N] = act(A * B^T + bias).  A is [M,K] (a_mn_major = 0, pitch lda) or [K,M] (a_mn_major = 1);
 * B is [N,K] (b_mn_major = 0, pitch ldb) or [K,N] (b_mn_major = 1).  bf16 inputs, fp32 accumulate, bf16 or
 * fp32 output.  block_n in {128, 256}.  splits > 1 = split-K through `workspace` (fp32 scratch of
 * dfb_gemm_workspace_bytes bytes) followed by a reduce kernel.  max_ctas > 0 caps the persistent grid.
 * Replaces predictOnBatch / variableGrads on Dense layers: /root/reference/src/common/models.ts:137-142. */
size_t dfb_gemm_workspace_bytes(int M, int N, int splits);
int dfb_gemm_pick_splits(int M, int N, int K, int block_n, int sms);
int dfb_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K, int lda,
                  int ldb, int ldd, int a_mn_major, int b_mn_major, int out_dtype, int act, int block_n, int splits,
                  void* workspace, size_t workspace_bytes, int max_ctas, void* stream);

/* GEMM with an auxiliary bf16 matrix in the epilogue: aux_mode 1 also stores the pre-activation, aux_mode 2 multiplies by
 * gelu'(aux) (fused GELU backward).  Replaces the activation ops around Dense layers of the tf.js graph. */
int dfb_gemm_bf16_aux(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K, int lda,
                      int ldb, int ldd, int a_mn_major, int b_mn_major, int act, int block_n, void* aux, long aux_ld,
                      int aux_mode, int max_ctas, void* stream);
/* aux_mode 2 (output = accumulator * gelu'(aux)) whose epilogue also emits the column sums of the stored output: the bias
 * gradient of the layer in front of the GELU, with no extra pass over D.  2-CTA kernel; `part` is fp32 scratch. */
size_t dfb_gemm_colsum_part_bytes(int M, int N);
int dfb_gemm_bf16_dgelu_colsum(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb, int ldd,
                               int a_mn_major, int b_mn_major, void* aux, long aux_ld, float* colsum, float* part,
                               size_t part_bytes, int max_ctas, void* stream);

/* All-gather fused into the first consuming GEMM: D = act(A * W + bias) with W[K,N] sharded over ranks; two extra warps
 * per CTA pull the remote shards over NVLink into the local W buffer while the k-loop starts on local rows.
 * Replaces the weight broadcast of /root/reference/src/server/federated_server.ts:80 for that layer (SURVEY K8+K1). */
int dfb_ag_gemm_bf16(const void* A, const void* W_local, void* D, const void* bias, int bias_dtype, int M, int N, int K, int lda,
                     int ldb, int ldd, int out_dtype, int act, int block_n, const long long* pieces, const int* piece_chunk0,
                     int num_pieces, const unsigned* expected, unsigned* ready, int rot_kb, void* stream);

/* Split-K GEMM that allocates nothing: same arguments as dfb_gemm_bf16 minus the workspace, which is a 32 MiB
 * array inside the library (one per GPU, shared by all calls, kept until exit).  For callers that cannot
 * allocate device memory themselves (no-sync step loops, C plug-ins inside a process whose pool owns the GPU).
 * Fails with a clear message when dfb_gemm_workspace_bytes(M, N, splits) exceeds
 * DFB_GEMM_NOALLOC_MAX_WORKSPACE_BYTES.  The first call on a device resolves the array's device address; make
 * it before any stream capture.  Calls that overlap on the GPU race on the array.
 * No counterpart in the reference (it has no native code and no split-K). */
#define DFB_GEMM_NOALLOC_MAX_WORKSPACE_BYTES (32u << 20)
int dfb_gemm_bf16_splitk_noalloc(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K,
                                 int lda, int ldb, int ldd, int a_mn_major, int b_mn_major, int out_dtype, int act,
                                 int block_n, int splits, int max_ctas, void* stream);

/* ---- CUDA-core GEMM for fp32 / tiny-N / unaligned shapes (csrc/gemm_simt.cu) ---------------------------
 * D[m,n] = act(sum_k A[m*sa_m + k*sa_k] * B[n*sb_n + k*sb_k] + bias[n]).  The MNIST heads Dense(10), Dense(5)
 * of /root/reference/experiment/mnist/mnist_server.ts:16-22 and model.json run here. */
int dfb_gemm_simt(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K, long sa_m,
                  long sa_k, long sb_n, long sb_k, long ldd, int a_dtype, int b_dtype, int out_dtype, int act, int ksplit,
                  void* stream);

/* ---- Conv2D, NHWC activations / HWIO filters (csrc/conv_simt.cu, csrc/conv_sm100.cu) ------------------
 * mode 0 forward, 1 data gradient, 2 filter gradient.  Replaces tf.js conv2d and its gradients for the Conv2D
 * layers of /root/reference/experiment/mnist/model.json. */
int dfb_conv2d_simt(int mode, const void* a, const void* b, void* out, const void* bias, int bias_dtype, int N, int H, int W,
                    int C, int OH, int OW, int K, int kh, int kw, int sh, int sw, int pt, int pl, int dtype, int out_is_f32,
                    int act, int ksplit, void* stream);
int dfb_conv2d_tc_supported(int mode, int N, int H, int W, int C, int OH, int OW, int K, int kh, int kw, int sh, int sw);
size_t dfb_conv2d_tc_workspace_bytes(int mode, int N, int H, int W, int C, int OH, int OW, int K, int kh, int kw);
int dfb_conv2d_tc(int mode, const void* a, const void* b, void* out, const void* bias, int bias_dtype, int N, int H, int W,
                  int C, int OH, int OW, int K, int kh, int kw, int sh, int sw, int pt, int pl, int dtype, int out_is_f32,
                  int act, int reserved, void* workspace, size_t workspace_bytes, void* stream);

/* bf16 patch matrix for thin-channel convolutions (ResNet stem): the Dense GEMM then does forward and filter gradient. */
int dfb_col2im_nhwc(const void* dpatches, void* dx, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw,
                    int pt, int pl, int Kpad, void* stream);
int dfb_im2col_nhwc(const void* x, void* out, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw, int pt,
                    int pl, int Kpad, void* stream);

/* ---- Fused softmax-cross-entropy forward + backward (csrc/softmax_ce.cu) ------------------------------
 * Replaces tf.losses.softmaxCrossEntropy(...).mean() + its gradient: /root/reference/src/common/models.ts:139.
 * mode of the dense form: 0 standard, 1 reference-compat (tfjs custom gradient), 2 reference-compat (true gradient). */
int dfb_softmax_ce_index(const void* logits, const void* labels, int label_dtype, float* loss_rows, void* dlogits, int rows,
                         int C, long ld, long ldg, int dtype, float scale, long ignore_index, void* stream);
int dfb_softmax_ce_dense(const void* z, const void* y, float* loss_rows, void* grad, int rows, int C, long ldz, long ldy,
                         long ldg, int dtype, float scale, int mode, void* stream);

/* ---- Element-wise / reduction helpers (csrc/elementwise.cu) -------------------------------------------- */
int dfb_colsum(const void* x, float* out, int M, int N, long ld, int dtype, void* stream);
int dfb_act(const void* a, const void* b, void* out, long n, int dtype, int mode, void* stream);
int dfb_maxpool_nhwc(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw,
                     int pt, int pl, int dtype, void* stream);
int dfb_maxpool_nhwc_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, int OH, int OW, int kh, int kw,
                         int sh, int sw, int pt, int pl, int dtype, void* stream);
/* metrics into a device ring slot (replaces model.evaluate(...).dataSync(): models.ts:106-115) */
/* table_grad[idx[t], :] += dy[t, :] in place (vector reductions): the token embedding's gradient added on top of the tied
 * LM head's weight gradient inside the gradient bucket.  idx int64; D a multiple of 8 (bf16) / 4 (fp32). */
int dfb_embedding_bwd_scatter(const void* dy, const long long* idx, void* table_grad, long n_tok, int D, long V, int dtype,
                              void* stream);
int dfb_accuracy(const void* pred, const void* labels, int label_kind, float* slot, int rows, int C, long ldp, long ldl,
                 int dtype, void* stream);

/* ---- Symmetric heap + collectives over NVLink peer memory (parallel/csrc/comm.cu) ----------------------
 * Transport replacement for socket.io (/root/reference/src/server/abstract_server.ts:81-89,
 * /root/reference/src/client/abstract_client.ts:148-173). */
int dfb_symm_alloc(size_t bytes, void** base_out);
int dfb_symm_free(void* base);
int dfb_symm_get_handle(void* base, void* handle_out64 /* host, 64 bytes */);
int dfb_symm_open_handle(const void* handle64 /* host */, void** peer_ptr_out);
int dfb_symm_close_handle(void* peer_ptr);
int dfb_enable_peer_access(int peer_device);
/* sync path: reduce-scatter + 1/R + optimizer (0 sgd, 1 momentum, 2 adam, 3 none) + all-gather, one launch per bucket
 * (federated_server.ts:92-116).  grads/params/sig are HOST arrays of R device pointers. */
int dfb_sync_rs_opt_ag(void* const* grads, void* const* params, void* const* sig, float* master, float* m, float* v,
                       void* reduced_out, long n, long shard, int rank, int R, unsigned epoch, int push, int grad_dtype,
                       int param_dtype, int opt_kind, float lr, float momentum, float beta1, float beta2, float eps,
                       float weight_decay, int step, int ctas, long nopush_lo, long nopush_hi, void* stream);
/* replayable form for CUDA graphs: barrier generation = *epoch_base + epoch, Adam step = *step_ptr (device memory) */
int dfb_sync_rs_opt_ag_replayable(void* const* grads, void* const* params, void* const* sig, float* master, float* m, float* v,
                                  long n, long shard, int rank, int R, unsigned epoch, int grad_dtype, int param_dtype, int opt_kind,
                                  float lr, float momentum, float beta1, float beta2, float eps, float weight_decay, int ctas,
                                  long nopush_lo, long nopush_hi, const unsigned* epoch_base, const int* step_ptr,
                                  const float* lr_ptr /* nullable: learning rate read on the device */, void* stream);
/* 1 (default): every bucket kernel ends with its own cross-rank barrier; 0: bucket kernels skip it and the step is closed
 * ONCE by dfb_step_close behind its last bucket kernel (federated_server.ts:92-116: mean -> update -> broadcast is one step).
 * epoch_base nullable (device word added to `epoch` for CUDA-graph replay). */
int dfb_sync_close_per_bucket(int on);
int dfb_step_close(void* const* sig, int rank, int R, unsigned epoch, const unsigned* epoch_base, void* stream);
int dfb_allgather_push(void* const* bufs, void* const* sig, long n_bytes, long shard_bytes, int rank, int R, unsigned epoch,
                       int ctas, void* stream);
int dfb_rank_barrier(void* const* sig, int rank, int R, unsigned epoch, void* stream);
/* failure detection (abstract_client.ts:12-13,148-158: connect / upload time-outs): every in-kernel wait for a peer is bounded;
 * a waiter that gives up records 1 + the silent rank in its signal pad, the kernel ends, the host reads the record. */
int dfb_set_barrier_timeout_us(long long us, void* stream);
int dfb_barrier_failures(unsigned* pad, unsigned* host_out2, void* stream);
/* async path (asynchronousSGD_server.ts:65-108): push with staleness guard, in-place apply, pull */
int dfb_async_push(const void* grad, void* const* acc0, void* const* acc1, void* const* sig, const unsigned* seen_version,
                   unsigned* push_state, long n, long shard, int rank, int R, int max_staleness, float scale, int policy_scale,
                   int grad_dtype, int ctas, const unsigned* ticket, void* stream);
int dfb_async_apply(float* acc0, float* acc1, unsigned* ctl_pad, float* master, float* m, float* v, void* param_shard,
                    unsigned* scratch, long shard_elems, int param_dtype, int opt_kind, float lr, float momentum, float beta1,
                    float beta2, float eps, float weight_decay, int step, int ctas,
                    const int* step_ptr /* nullable: Adam step count on the device (graph replay) */,
                    int version_per_apply /* 1: version counts optimizer updates, 0: gradients applied */, void* stream);
int dfb_async_pull(void* const* params, void* const* sig, void* dst, unsigned* seen_version, long n_bytes, long shard_bytes,
                   int rank, int R, int ctas, void* stream);
int dfb_async_read_ctl(const unsigned* ctl_pad, unsigned* host_out9, void* stream);
/* device work queue (dataset.ts:43-67) */
/* `wq`: the queue's control block in the home rank's heap (through any rank's mapping): the slot inside the signal pad
 * (pad + dfb_wq_pad_offset_words(), room for dfb_wq_pad_capacity() batches) or a heap block of dfb_wq_words(batches) words */
int dfb_wq_init(unsigned* wq, unsigned batches, unsigned epochs, void* stream);
int dfb_wq_take(unsigned* wq, unsigned* ticket_out2, void* stream);
int dfb_wq_complete(unsigned* wq, const unsigned* ticket2, const unsigned* accepted /* nullable */, void* stream);
int dfb_wq_words(unsigned batches);
int dfb_wq_pad_offset_words(void);
int dfb_wq_pad_capacity(void);
int dfb_gather_batch(const void* src, void* dst, const unsigned* ticket, long row_bytes, int batch_size, long rows,
                     void* stream);

/* ---- causal self-attention on the packed QKV projection (no counterpart in the reference: its models have no attention;
 * hot path of BASELINE.json config 5).  tcgen05 + TMEM + TMA, head dimension 64, bf16, scale 1/sqrt(D).
 * qkv / dqkv: [B,T,3,H,D]; out / d_out: [B,T,H,D]; lse: [B,H,T] fp32.
 * scratch (dfb_attn_bwd_scratch_bytes): delta [B,H,T] fp32 rounded to 256 B, then the fp32 dQ accumulator [B,T,H,D] and the
 * 256-byte work-item counter of the persistent grid; accumulator and counter must be zero on entry and are zero again on
 * return (allocate the buffer zero-filled once per shape and keep it). */
int dfb_attn_supported(int B, int T, int H, int D);
int dfb_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int D, void* stream);
int dfb_attn_bwd_set_debug(long long* buf);   /* profiling aid: per-CTA cycle counters of the backward kernel (8 int64 per CTA), NULL = off */
size_t dfb_attn_bwd_scratch_bytes(int B, int T, int H, int D);
int dfb_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, void* dqkv,
                 float* dbias /* nullable, fp32 [3*H*D]: receives the column sums of dqkv = bias gradient of the QKV projection */,
                 void* scratch, size_t scratch_bytes, int B, int T, int H, int D, void* stream);

/* ---- normalisation and layout kernels (no counterpart in the reference: its models are Dense / Conv2D / MaxPool) ------------
 * fused residual add + LayerNorm: s = x + r (r nullable), y = LN(s) * gamma + beta; mean / rstd: [rows] fp32 */
int dfb_add_layernorm_fwd(const void* x, const void* r, const void* gamma, const void* beta, void* s_out, void* y, float* mean,
                          float* rstd, int rows, int D, float eps, int dtype, void* stream);
int dfb_layernorm_bwd_staged(int on);   /* 1: rows staged through shared memory by bulk copies; 0 (default): register-only kernel */
int dfb_layernorm_bwd_blocks(int rows);                     /* row blocks of the `part` scratch: 2 * D * blocks floats */
int dfb_add_layernorm_bwd(const void* dy, const void* ds, const void* s, const void* gamma, const float* mean, const float* rstd,
                          void* dx, void* dgamma, void* dbeta,
                          float* dx_colsum /* nullable, fp32 [D]: column sums of dx = bias gradient of the layer that fed the residual */,
                          float* part /* blocks * (dx_colsum ? 3 : 2) * D floats */, int rows, int D, int dtype, void* stream);
/* strided copy of rows of `row_bytes` (multiple of 16): src + sum(i_k * s_k) -> dst + sum(i_k * d_k), strides in bytes */
int dfb_permute_rows(const void* src, void* dst, int n0, int n1, int n2, int n3, long s0, long s1, long s2, long s3, long d0,
                     long d1, long d2, long d3, int row_bytes, void* stream);
/* channels-last BatchNorm fused with the residual add and ReLU of a ResNet block (csrc/batchnorm.cu).
 * x, res (nullable), y: [rows, C] of `dtype`; gamma/beta of `param_dtype`; running stats fp32 (nullable);
 * part: 2 * C * dfb_bn_partial_blocks(rows, C, dtype) floats; ab: 2 * C floats; coef: 3 * C floats. */
int dfb_bn_partial_blocks(long rows, int C, int dtype);
int dfb_bn_fwd_train(const void* x, const void* res, const void* gamma, const void* beta, float* running_mean, float* running_var,
                     void* y, float* save_mean, float* save_invstd, float* part, float* ab, long rows, int C, float eps,
                     float momentum, int relu, int dtype, int param_dtype, void* stream);
int dfb_bn_fwd_eval(const void* x, const void* res, const void* gamma, const void* beta, const float* running_mean,
                    const float* running_var, void* y, float* ab, long rows, int C, float eps, int relu, int dtype,
                    int param_dtype, void* stream);
int dfb_bn_bwd(const void* dy, const void* y, const void* x, const void* gamma, const float* save_mean, const float* save_invstd,
               void* dx, void* dres, void* dgamma, void* dbeta, float* part, float* coef, long rows, int C, int relu, int dtype,
               int param_dtype, void* stream);

/* ---- multicast (NVLS) heap: driver VMM allocation + NVSwitch multicast object (parallel/csrc/mcast.cu).  No counterpart in
 * the reference (socket.io star, abstract_server.ts:81-89).  Descriptors are POSIX fds handed between the rank processes. */
int dfb_mc_supported(int* out);                                            /* 1: device can join a multicast object */
int dfb_mc_round_size(size_t bytes, int ndev, size_t* rounded_out);        /* physical and multicast granularity */
int dfb_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out, void** va_out);      /* zero-filled, mapped */
int dfb_vmm_import(int fd, size_t bytes, unsigned long long* handle_out, void** va_out);           /* peer's allocation */
int dfb_vmm_release(unsigned long long handle, void* va, size_t bytes);
int dfb_mc_create(int ndev, size_t bytes, unsigned long long* mc_out, int* fd_out);                /* rank 0 */
int dfb_mc_import(int fd, unsigned long long* mc_out);                                             /* other ranks */
int dfb_mc_add_device(unsigned long long mc);                               /* every rank, before anybody binds */
int dfb_mc_bind_map(unsigned long long mc, unsigned long long mem, size_t bytes, void** mc_va_out);
int dfb_mc_release(unsigned long long mc, void* mc_va, size_t bytes);
/* sync path through the switch: multimem.ld_reduce of the gradients, multimem.st of the refreshed weights (bf16 only).
 * mc_grads / mc_params: multicast addresses of the bucket; params: unicast bases (ranges that are not pushed);
 * epoch_base / step_ptr / lr_ptr: null for eager launches, device words for CUDA-graph replay. */
int dfb_sync_nvls(const void* mc_grads, void* mc_params, void* const* params, void* const* sig, float* master, float* m, float* v,
                  long n, long shard, int rank, int R, unsigned epoch, int opt_kind, float lr, float momentum, float beta1,
                  float beta2, float eps, float weight_decay, int step, int ctas, long nopush_lo, long nopush_hi,
                  const unsigned* epoch_base, const int* step_ptr, const float* lr_ptr, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DISTRIFLOW_B200_H_ */
