/* C ABI of libtfde_kernels.so -- the hand-written sm_100a kernels of tfde_b200.
 *
 * Conventions
 *  - every function returns a cudaError_t as int (0 == success, 1 == invalid argument);
 *  - tensors are dense, channels-last (NHWC), fp32 unless stated; conv kernels are HWIO, dense kernels
 *    (in, out) -- the Keras variable layout of the reference, which is also the GEMM-native layout;
 *  - `stream` is a cudaStream_t; all launches are asynchronous and CUDA-graph capturable;
 *  - `ws`/`counters`/`partials`: caller-provided scratch. `counters` must be zero on first use and is
 *    left zero by every kernel (self-cleaning tickets); kernels sharing scratch must share a stream.
 *
 * "ref:" comments cite the reference call site whose TensorFlow op the kernel replaces
 * (paths relative to the reference repository root).
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* tfde_stream_t; /* cudaStream_t */

/* ---- GEMM family (csrc/kernels/gemm_fp32.cu) ------------------------------------------------------ */
/* y[M,N] = act(x[M,K] . w[K,N] + bias).  ref: Dense, distributed_with_keras.py:37-38,
 * mnist_keras_distributed.py:103,108 (cuBLAS GEMM + Eigen bias/relu in TF). */
int tfde_dense_fwd(const float* x, const float* w, const float* bias, float* y, int M, int K, int N,
                   int relu, float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
/* dw[K,N] = x^T . (dy * (ymask > 0)), db[N] = column sums (db may be NULL). ref: Dense backward. */
int tfde_dense_wgrad(const float* x, const float* dy, const float* ymask, float* dw, float* db, int M,
                     int K, int N, float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
/* dx[M,K] = (dy * (ymask > 0)) . w^T. ref: Dense backward. */
int tfde_dense_dgrad(const float* dy, const float* ymask, const float* w, float* dx, int M, int K, int N,
                     float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
/* NHWC x HWIO convolution as implicit GEMM. ref: Conv2D, distributed_with_keras.py:34,
 * mnist_keras_distributed.py:83-98 (cuDNN in TF). pt/pl = top/left padding (TF SAME rule). */
int tfde_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int H, int W,
                    int C, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout,
                    int relu, float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
int tfde_conv2d_wgrad(const float* x, const float* dy, const float* ymask, float* dw, float* db, int B,
                      int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl,
                      int Cout, float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
int tfde_conv2d_dgrad(const float* dy, const float* ymask, const float* w, float* dx, int B, int H, int W,
                      int C, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout,
                      float* ws, size_t ws_bytes, unsigned* counters, tfde_stream_t stream);
int tfde_gemm_max_counters(void);
/* Direct convolution for small channel counts (Cin, Cout <= 32, image + filter fit in shared memory): one CTA per
 * image, no GEMM tiles wasted on padding, stride-2 data gradient visits only the taps of its stride phase.
 * ref: the Conv2D(6,3) / Conv2D(12,6,s2) / Conv2D(24,6,s2) layers of mnist_keras_distributed.py:83-98. */
int tfde_smallconv_supported(int B, int H, int W, int Cin, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt,
                             int pl, int Cout, int bwd, int need_dx);
size_t tfde_smallconv_bwd_ws_floats(int B, int Cin, int KH, int KW, int Cout);
int tfde_smallconv_fwd(const float* x, const float* w, const float* bias, float* y, int B, int H, int W, int Cin,
                       int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout, int relu,
                       tfde_stream_t stream);
int tfde_smallconv_bwd(const float* x, const float* dy, const float* w, float* dx, float* dw, float* ws, int B, int H,
                       int W, int Cin, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout,
                       tfde_stream_t stream);
/* tfde_dense_fwd for callers that cannot allocate device memory (a plug-in inside a process whose GPU
 * memory pool belongs to another framework, or a loop where cudaMalloc is forbidden): the split-K scratch
 * is a static __device__ array inside the library (about 2 MiB, held by every GPU that loads the library
 * until exit) instead of caller-provided ws/counters. Shapes whose split-K partial sums exceed 2 MiB are
 * refused with return value 1 and a message on stderr. All calls on a device share the array: calls that
 * may overlap on the GPU race. The first call on each device resolves the array's device address and
 * must happen outside stream capture. Allocates nothing. This entry point has no counterpart in the
 * reference (the reference has no native code at all); it exists for embedders of this library. */
int tfde_dense_fwd_static_scratch(const float* x, const float* w, const float* bias, float* y, int M, int K,
                                  int N, int relu, tfde_stream_t stream);

/* ---- fused first layer of Model A (csrc/kernels/conv_pool.cu) -------------------------------------- */
/* Conv2D(32,3,relu) + MaxPooling2D() in one pass; x may be uint8 (scaled by in_scale on load).
 * ref: distributed_with_keras.py:20-23,34-35. */
int tfde_conv3x3c1_relu_pool_fwd(const void* x, int x_is_u8, float in_scale, const float* w,
                                 const float* bias, float* pooled, uint8_t* idx, int B, int H, int W,
                                 int Cout, tfde_stream_t stream);
size_t tfde_conv3x3c1_relu_pool_bwd_ws_floats(void);
int tfde_conv3x3c1_relu_pool_bwd(const void* x, int x_is_u8, float in_scale, const float* pooled,
                                 const uint8_t* idx, const float* dpooled, float* dw, float* db,
                                 float* partials, unsigned* counter, int B, int H, int W, int Cout,
                                 tfde_stream_t stream);

/* ---- normalisation / pooling / elementwise (csrc/kernels/nn_ops.cu) -------------------------------- */
/* ref: BatchNormalization(scale=False, center=True) + Activation('relu'),
 * mnist_keras_distributed.py:86-88,93-94,99-100,104-105 (cuDNN FusedBatchNormV3 in TF). */
size_t tfde_bn_ws_floats(long M, int C);
int tfde_bn_fwd_train(const float* x, float* y, long M, int C, const float* gamma, const float* beta,
                      float* moving_mean, float* moving_var, float* save_mean, float* save_invstd,
                      float momentum, float eps, int relu, int unbiased_moving, float* partials,
                      unsigned* counters, tfde_stream_t stream);
int tfde_bn_fwd_infer(const float* x, float* y, long M, int C, const float* gamma, const float* beta,
                      const float* moving_mean, const float* moving_var, float eps, int relu,
                      tfde_stream_t stream);
int tfde_bn_bwd(const float* x, const float* y, const float* dy, float* dx, long M, int C,
                const float* gamma, const float* save_mean, const float* save_invstd, int relu,
                float* dbeta, float* dgamma_full, float* partials, unsigned* counters,
                tfde_stream_t stream);
/* ref: MaxPooling2D, distributed_with_keras.py:35 (generic window/stride/padding). */
int tfde_maxpool_fwd(const float* x, float* y, uint8_t* idx, int B, int H, int W, int C, int PH, int PW,
                     int sh, int sw, int pt, int pl, int Ho, int Wo, tfde_stream_t stream);
int tfde_maxpool_bwd(const float* dy, const uint8_t* idx, float* dx, int B, int H, int W, int C, int PH,
                     int PW, int sh, int sw, int pt, int pl, int Ho, int Wo, tfde_stream_t stream);
/* ref: Activation('relu'), mnist_keras_distributed.py:88. */
int tfde_relu_fwd(const float* x, float* y, long n, tfde_stream_t stream);
int tfde_relu_bwd(const float* y, const float* dy, float* dx, long n, tfde_stream_t stream);
int tfde_add(const float* a, const float* b, float* y, long n, int relu, tfde_stream_t stream);
/* ref: image cast + /255, distributed_with_keras.py:20-23, mnist_keras_distributed.py:211. */
int tfde_u8_to_f32(const uint8_t* x, float* y, long n, float scale, tfde_stream_t stream);
/* ref: Dropout(0.5), mnist_keras_distributed.py:106. Philox mask from (seed, step, stream_id, index);
 * seed/step are read from the device hparams block. Same call for forward and backward. */
int tfde_dropout(const float* x, float* y, long n, float rate, const void* hparams, unsigned stream_id,
                 int training, tfde_stream_t stream);
/* ref: Dense(10, activation='softmax'), mnist_keras_distributed.py:108. */
int tfde_softmax_fwd(const float* z, float* p, int rows, int C, tfde_stream_t stream);
int tfde_softmax_bwd(const float* p, const float* dp, float* dz, int rows, int C, tfde_stream_t stream);
int tfde_gap_fwd(const float* x, float* y, int B, int HW, int C, tfde_stream_t stream);
int tfde_gap_bwd(const float* dy, float* dx, int B, int HW, int C, tfde_stream_t stream);

/* ---- loss + single-GPU optimizer (csrc/kernels/loss_optim.cu) -------------------------------------- */
/* Sparse categorical cross-entropy (+gradient +accuracy). out4 = [loss_sum*scale, correct, rows, 0].
 * ref: distributed_with_keras.py:41, tf2_mnist_distributed.py:81-83, mnist_keras_distributed.py:114. */
size_t tfde_xent_ws_floats(void);
int tfde_xent(const float* in, const void* labels, int label_is_i64, float* dgrad, int rows, int C,
              float scale, int from_probs, float* partials, unsigned* counter, float* out4,
              tfde_stream_t stream);
/* opt: 0 SGD, 1 momentum, 2 nesterov, 3 Adam, applied to the flat arena; advances the device-side
 * step. ref: SGD, distributed_with_keras.py:42; GradientDescentOptimizer, mnist_keras_distributed.py:111. */
int tfde_optimizer_apply(float* w, const float* g, float* s1, float* s2, void* w_bf16, long n,
                         int n_extra, float* metrics, float* last_extra, int opt, void* hparams,
                         unsigned* counter, tfde_stream_t stream);
int tfde_hparams_advance(void* hparams, tfde_stream_t stream);

/* ---- peer-to-peer collectives (csrc/kernels/allreduce_p2p.cu) -------------------------------------- */
/* Symmetric memory + CUDA IPC plumbing (set-up only). */
int tfde_set_device(int dev);
int tfde_symm_alloc(size_t bytes, void** out);
int tfde_symm_free(void* p);
int tfde_ipc_get_handle(void* p, void* handle64);
int tfde_ipc_open_handle(const void* handle64, void** out);
int tfde_ipc_close_handle(void* p);
int tfde_ipc_handle_bytes(void);
int tfde_ar_max_ranks(void);
int tfde_ar_max_blocks(void);
size_t tfde_ar_flag_bytes(int world);
size_t tfde_ar_desc_bytes(void);
/* Fused gradient all-reduce (one-shot / two-shot push over NVLink) + scale + optional bf16 wire cast +
 * SGD/momentum/Adam update; `desc` is a TfdeARDesc (layout in allreduce_p2p.cu, mirrored by
 * parallel/comm.py::_ARDesc). ref: the NCCL all-reduce + apply_gradients implied by
 * MultiWorkerMirroredStrategy, distributed_with_keras.py:16,63; MirroredStrategy,
 * mnist_keras_distributed.py:243. */
int tfde_allreduce_fused(const void* desc, tfde_stream_t stream);
/* dst(local) <- src(peer mapped): initial broadcast from the chief. ref: distributed_with_keras.py:58. */
int tfde_p2p_copy(float* dst, const float* src, long n, tfde_stream_t stream);

/* ---- small-batch latency kernels (csrc/kernels/small_batch.cu) -------------------------------------- */
/* Dense forward for small M*N and long K (cluster / DSMEM split-K). ref: Dense(64) over the 5408-wide
 * flattened activation at per-replica batch 64, distributed_with_keras.py:13,37. */
int tfde_dense_skinny_supported(int M, int K, int N);
size_t tfde_dense_skinny_ws_floats(int M, int K, int N);
int tfde_dense_skinny_fwd(const float* x, const float* w, const float* bias, float* y, int M, int K, int N,
                          int relu, float* ws, unsigned* counters, tfde_stream_t stream);
/* Classifier head in one launch: Dense(N<=32) + sparse categorical cross-entropy (+accuracy) + backward of
 * that Dense. ref: Dense(10) + SCCE, distributed_with_keras.py:38,41; Dense(10, softmax) + SCCE on
 * probabilities, mnist_keras_distributed.py:108,114. */
int tfde_head_supported(int K, int N);
size_t tfde_head_ws_floats(int M, int K, int N);
int tfde_head_fwd_bwd(const float* h, const float* w, const float* bias, const void* labels, int label_is_i64,
                      float* out, float* dh, float* dw, float* db, float* out4, int M, int K, int N,
                      int relu_mask, int from_probs, float scale, float* partials, unsigned* counter,
                      tfde_stream_t stream);
/* Backward of Dense(K -> 64) at batch <= 64 in ONE launch: dx [M,K] (may be NULL), dw [K,64], db [64] (may be
 * NULL); ymask = the layer's own output when it had a fused ReLU, else NULL. Replaces two forked 85-block
 * GEMMs on the critical path of the batch-64 step. ref: Dense(64, relu), distributed_with_keras.py:37. */
int tfde_dense_bwd_skinny_supported(int M, int K, int N);
int tfde_dense_bwd_skinny(const float* x, const float* w, const float* dy, const float* ymask, float* dx, float* dw,
                          float* db, int M, int K, int N, tfde_stream_t stream);

/* ---- bf16 tensor-core GEMMs: tcgen05.mma + TMEM + TMA (gemm_tcgen05.cu, gemm_tcgen05_persistent.cu) -- */
/* C[M,N] = act(A[M,K] . B[N,K]^T + bias); A, B bf16 with K contiguous (K % 8 == 0, 16-byte aligned);
 * exactly one of c_f32 / c_bf16 may be NULL. The first call per tile shape raises the kernel's dynamic
 * shared-memory limit and must happen outside stream capture. ref: the cuBLAS GEMMs behind Dense / Conv2D
 * in TensorFlow (distributed_with_keras.py:34-38), BASELINE.json configs 3 and 4. */
int tfde_gemm_bf16_tn(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc, const float* bias,
                      int relu, int M, int N, int K, tfde_stream_t stream);
/* persistent one-CTA-per-SM schedule, double-buffered TMEM accumulator, optional deterministic split-K
 * through the fp32 scratch `ws` (NULL disables splitting). */
int tfde_gemm_bf16_tn_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                                 const float* bias, int relu, int M, int N, int K, float* ws, size_t ws_bytes,
                                 tfde_stream_t stream);
int tfde_cast_bf16(const float* x, void* y, long n, tfde_stream_t stream);
int tfde_cast_transpose_bf16(const float* x, const float* mask, void* y, int R, int C, tfde_stream_t stream);

/* ---- mixed-precision glue (mixed_ops.cu) and convolution operand builders (conv_gather.cu) ------------ */
int tfde_cast_bf16_masked(const float* x, const float* mask, void* y, long n, tfde_stream_t stream);
int tfde_colsum_masked(const float* x, const float* mask, float* out, long R, int C, float* partials,
                       unsigned* counters, tfde_stream_t stream);
/* geometry arguments: B,H,W,C,KH,KW,Ho,Wo,sh,sw,pt,pl,Cout; ld* in elements, multiples of 8.
 * ref: Conv2D forward / filter gradient / data gradient (cuDNN in TF), mnist_keras_distributed.py:83-98. */
int tfde_im2col_bf16(const float* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                     int sh, int sw, int pt, int pl, int Cout, int ldo, tfde_stream_t stream);
int tfde_im2colT_bf16(const float* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                      int sh, int sw, int pt, int pl, int Cout, long ldo, tfde_stream_t stream);
int tfde_dgrad_gather_bf16(const float* dy, const float* mask, void* out, int B, int H, int W, int C, int KH,
                           int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout, int ldo,
                           tfde_stream_t stream);
int tfde_wrot_bf16(const float* w, void* out, int KH, int KW, int Cin, int Cout, int ldo, tfde_stream_t stream);
int tfde_cast_transpose_pad_bf16(const float* x, const float* mask, void* y, long R, int C, long ldy,
                                 tfde_stream_t stream);
int tfde_transpose_bf16(const void* x, void* y, long R, int C, long ldx, long ldy, tfde_stream_t stream);

/* ---- parameter-server push / pull over peer memory (ps_p2p.cu) ---------------------------------------- */
/* Asynchronous SGD on an owner GPU's shard: push adds -lr*g with NVLink reduction atomics (no barrier),
 * pull copies the owner's current values. ref: ParameterServerStrategy, mnist_keras_distributed.py:241-242,
 * tf2_mnist_distributed.py:189 (gRPC RecvTensor in TF). */
int tfde_ps_push(float* w_owner, const float* g, long n, float lr, tfde_stream_t stream);
int tfde_ps_pull(float* w_local, const float* w_owner, long n, tfde_stream_t stream);
/* Whole-model forms used by ParameterServerStrategy when ps and worker tasks share an NVLink domain: `table` is a
 * device array of {owner ptr, local ptr, n, kind} segments (kind 0: trainable -- push adds -lr * local gradient into
 * the owner with reduction atomics, kind 1: non-trainable state -- push stores the local value). push also bumps the
 * owner-side global_step counter (device memory on ps 0) and writes the new value to *step_out.
 * ref: mnist_keras_distributed.py:241-242 (the worker<->ps variable traffic of ParameterServerStrategy). */
size_t tfde_ps_seg_bytes(void);
int tfde_ps_pull_multi(const void* table, int nseg, long max_n, tfde_stream_t stream);
int tfde_ps_push_multi(const void* table, int nseg, long max_n, float lr, unsigned* step_owner, unsigned* step_out,
                       tfde_stream_t stream);

/* ---- the whole training step of the reference's Keras CNN in one cooperative kernel (fused_step.cu) ---------- */
/* Conv2D(32,3,relu) -> MaxPooling2D -> Flatten -> Dense(64,relu) -> Dense(ncls <= 16), sparse CE from logits, SGD /
 * momentum / Adam, per-replica batch <= 64: forward, loss, backward, cross-replica gradient reduction over peer memory
 * (world > 1) and optimizer, `nsteps` steps per launch (step s reads x + s * x_step_bytes and y + s * B).
 * ref: distributed_with_keras.py:33-43 (model, loss, optimizer), :16,63 (the per-step all-reduce under
 * MultiWorkerMirroredStrategy), :13 (batch 64 per worker). */
typedef struct TfdeFusedStepDesc {
  const void* x; const void* y;            /* uint8 or float [nsteps*B,28,28,1]; int64 labels [nsteps*B] */
  int x_is_u8; float in_scale; long x_step_bytes;
  int B, nsteps, ncls, opt;                /* opt: 0 SGD, 1 momentum, 2 nesterov, 3 Adam */
  float* w; float* g; float* s1; float* s2;                      /* flat parameter / gradient / slot arenas */
  long off_ck, off_cb, off_dk, off_db, off_hk, off_hb, n;        /* arena offsets of the 6 tensors; n = trainable floats */
  float loss_scale; int rank; int world; int use_tc;             /* loss_scale = 1 / global batch */
  void* hp; float* metrics; float* tails;  /* hyper-parameter block; running metric sums; per-step [loss,correct,rows,0] */
  float* part; float* hbuf; float* cpart; float* small_block;    /* scratch, sizes from tfde_fused_step_*_floats() */
  void* state; void* status;               /* zero-initialised words: barrier / ticket / epoch state; int[4] status */
  unsigned long long timeout_ns;
  void* rs[8]; void* gsum[8]; void* srecv[8]; void* flags[8];    /* world > 1: peer-mapped receive buffers */
  void* prof;                              /* optional globaltimer stamps [169][16] */
  void* ready; unsigned ready_tag;         /* streamed inputs: uint32 ready[nsteps] (step s usable once >= ready_tag); NULL = resident */
} TfdeFusedStepDesc;
size_t tfde_fused_step_desc_bytes(void);
size_t tfde_fused_step_part_floats(void);
size_t tfde_fused_step_hbuf_floats(void);
size_t tfde_fused_step_cpart_floats(void);
size_t tfde_fused_step_small_floats(void);
size_t tfde_fused_step_state_words(void);
size_t tfde_fused_step_rs_floats(int world);
size_t tfde_fused_step_gsum_floats(void);
size_t tfde_fused_step_srecv_floats(int world);
size_t tfde_fused_step_flag_words(int world);
int tfde_fused_step_blocks(void);
int tfde_fused_step_prof_stamps(void);
int tfde_fused_step_occupancy(int use_tc, size_t dyn_smem);
int tfde_fused_step(const TfdeFusedStepDesc* d, tfde_stream_t stream);
/* End-to-end form (Keras steps_per_execution): per step one cudaMemcpyAsync of that step's pinned images and one of its
 * labels on `copy_stream` into d->x / d->y, an event edge, one launch, `ev_done` recorded behind it.
 * ref: the host->device feed of every batch inside Keras fit, distributed_with_keras.py:63. */
int tfde_fused_exec(const TfdeFusedStepDesc* d, const void* host_x, const void* host_y, size_t x_bytes_per_step,
                    size_t y_bytes_per_step, tfde_stream_t copy_stream, void* ev_copied, void* ev_done,
                    tfde_stream_t stream);
/* Streamed form: the kernel is launched without waiting for the copies; behind every step's images and labels the copy
 * stream writes host_tags[s] (pinned, == d->ready_tag) into d->ready[s], and the kernel takes up step s when it sees it.
 * One launch covers all steps with all copies but the first hidden under compute. */
int tfde_fused_exec_streamed(const TfdeFusedStepDesc* d, const void* host_x, const void* host_y, const void* host_tags,
                             size_t x_bytes_per_step, size_t y_bytes_per_step, tfde_stream_t copy_stream,
                             void* ev_copied, void* ev_done, tfde_stream_t stream);
int tfde_event_create(void** out);
int tfde_event_destroy(void* e);
int tfde_event_synchronize(void* e);
int tfde_event_query(void* e);
/* Self-test of the TF32 tcgen05 path fed from thread-written shared memory (umma_tf32.cuh): act [64][32], W [32][64],
 * dh [64][64] -> D1 = act.W, D2 = act^T.dh, D3 = dh.W^T. ref: the Dense(64) GEMMs of distributed_with_keras.py:37. */
int tfde_umma_tf32_selftest(const float* act, const float* W, const float* dh, float* D1, float* D2, float* D3,
                            int variant, tfde_stream_t stream);
/* tfde_dense_wgrad whose epilogue also stores the tiles into the peers' all-reduce receive slots
 * (`push` is a TfdePush, gemm_fp32.cu): GEMM -> reduce-scatter in one kernel. */
int tfde_dense_wgrad_push(const float* x, const float* dy, const float* ymask, float* dw, float* db, int M,
                          int K, int N, float* ws, size_t ws_bytes, unsigned* counters, const void* push,
                          tfde_stream_t stream);
size_t tfde_push_desc_bytes(void);

/* ---- tcgen05 GEMMs with MN-major operands (gemm_tcgen05_persistent.cu) ------------------------------------ */
/* C[M,N] = A[K,M]^T . B[K,N]: the filter-gradient GEMM with activations [rows, Cin] and output gradients
 * [rows, Cout] as the forward pass left them (both "MN-major": a_major/b_major bits of the instruction
 * descriptor, 64x64 TMA boxes). M % 8 == 0, N % 8 == 0, any K; lda/ldb row strides in elements. Deterministic
 * split-K through `ws`. ref: the dW of every Conv2D/Dense (cuDNN/cuBLAS in TF): distributed_with_keras.py:34-38. */
int tfde_gemm_bf16_nt_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc, int M, int N,
                                 int K, long lda, long ldb, float* ws, size_t ws_bytes, tfde_stream_t stream);
/* C[M,N] = act(A[M,K] . B[kb,N] + bias): A K-major, B MN-major = the weight matrix exactly as Keras stores it
 * (HWIO flattened / (in,out)); kb <= K valid rows of B (padded im2col columns meet zero-filled rows). */
int tfde_gemm_bf16_nn_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                                 const float* bias, int relu, int M, int N, int K, int kb, long ldb, float* ws,
                                 size_t ws_bytes, tfde_stream_t stream);

/* tfde_gemm_bf16_nn_persistent with bf16 output only, whose epilogue also leaves per-32-row partial column sums /
 * sums of squares of the rounded output in stat_sum / stat_sq (fp32 [ceil(M/32)][N]) for the batch-norm that
 * consumes it (tfde_bn_fwd_from_partials_bf16). No split-K; N % 64 == 0. */
int tfde_gemm_bf16_nn_stats_persistent(const void* A, const void* B, void* c_bf16, long ldc, const float* bias, int M,
                                       int N, int K, int kb, long ldb, float* stat_sum, float* stat_sq,
                                       tfde_stream_t stream);

/* ---- implicit-GEMM convolutions (gemm_tcgen05_persistent.cu, CONV modes) ------------------------------------ */
/* The im2col operand is never written to memory: the producer warp of the persistent tcgen05 GEMM fetches
 * 128-pixel x 64-channel tiles of the NHWC bf16 activation with im2col-mode TMA loads (cuTensorMapEncodeIm2col,
 * SASS UTMALDG.4D.IM2COL; padding = the tensor map's zero fill). Channel counts: C % 64 == 0 and Cout % 64 == 0,
 * windows up to 16 x 16, pads = (top, bottom, left, right). bf16 outputs leave through TMA stores (UTMASTG).
 * ref: Conv2D of mnist_keras_distributed.py:83-98 / distributed_with_keras.py:34 (cuDNN in TF). */
int tfde_conv_igemm_supported(int B, int H, int W, int C, int KH, int KW, int Ho, int Wo, int Cout);
/* y[B,Ho,Wo,Cout] = conv(x[B,H,W,C], w16[KH,KW,C,Cout] as stored) + bias; stat_sum/stat_sq: optional fused
 * batch-norm partial statistics as in tfde_gemm_bf16_nn_stats_persistent. */
int tfde_conv_igemm_fwd(const void* x, const void* w16, void* y, const float* bias, int B, int H, int W, int C, int KH,
                        int KW, int Ho, int Wo, int sh, int sw, int pt, int pb, int pl, int pr, int Cout,
                        float* stat_sum, float* stat_sq, float* ws, size_t ws_bytes, tfde_stream_t stream);
/* dx[B,H,W,C] of a STRIDE-1 convolution: the same kernel run over dy with the mirrored padding; the forward
 * pass's bf16 weights are read K-major with the taps mirrored (no rotated copy). */
int tfde_conv_igemm_dgrad(const void* dy, const void* w16, void* dx, int B, int H, int W, int C, int KH, int KW,
                          int Ho, int Wo, int pt, int pl, int Cout, float* ws, size_t ws_bytes, tfde_stream_t stream);
/* dw[KH,KW,C,Cout] (fp32) = im2col(x)^T . dy: 64-pixel x 64-channel im2col boxes as the MN-major A operand,
 * reduced over the B*Ho*Wo output pixels with deterministic split-K through ws. */
int tfde_conv_igemm_wgrad(const void* x, const void* dy, float* dw, int B, int H, int W, int C, int KH, int KW,
                          int Ho, int Wo, int sh, int sw, int pt, int pb, int pl, int pr, int Cout, float* ws,
                          size_t ws_bytes, tfde_stream_t stream);

/* ---- bf16-activation layer kernels (act_bf16.cu) ----------------------------------------------------------- */
/* Keras mixed_bfloat16 semantics: bf16 tensors between layers, fp32 variables / statistics. All tensors are
 * channels-last with C % 8 == 0 (tfde_act16_supported_channels). `ws`/`counters`: per-stream scratch for the
 * two-phase deterministic column reductions. ref: BatchNormalization / Activation / MaxPooling2D / Conv2D,
 * mnist_keras_distributed.py:83-104; ResNet-50 config of BASELINE.json. */
int tfde_act16_supported_channels(int C);
size_t tfde_act16_ws_floats(long M, int C);
int tfde_bn_fwd_train_bf16(const void* x, void* y, long M, int C, const float* gamma, const float* beta,
                           float* moving_mean, float* moving_var, float momentum, float eps, int relu,
                           int unbiased_moving, float* save_mean, float* save_invstd, float* scale, float* shift,
                           float* ws, size_t ws_floats, unsigned* counters, tfde_stream_t stream);
int tfde_bn_apply_bf16(const void* x, void* y, long M, int C, const float* scale, const float* shift, int relu,
                       tfde_stream_t stream);
/* training forward from the partial statistics a GEMM epilogue left behind (psum / psq: fp32 [G][C]) */
int tfde_bn_fwd_from_partials_bf16(const void* x, void* y, long M, int C, const float* psum, const float* psq, int G,
                                   const float* gamma, const float* beta, float* moving_mean, float* moving_var,
                                   float momentum, float eps, int relu, int unbiased_moving, float* save_mean,
                                   float* save_invstd, float* scale, float* shift, float* ws, size_t ws_floats,
                                   unsigned* counters, tfde_stream_t stream);
/* backward; with relu != 0 the mask (y > 0) is recomputed as x*scale+shift > 0 instead of re-reading y */
int tfde_bn_bwd_bf16(const void* x, const void* dy, void* dx, long M, int C, const float* gamma,
                     const float* save_mean, const float* save_invstd, const float* scale, const float* shift,
                     int relu, float* dgamma, float* dbeta, float* ws, size_t ws_floats, unsigned* counters,
                     tfde_stream_t stream);
int tfde_colsum_bf16(const void* x, long M, int C, float* out, float* ws, size_t ws_floats, unsigned* counters,
                     tfde_stream_t stream);
int tfde_add_bf16(const void* a, const void* b, void* y, long n, int relu, tfde_stream_t stream);
int tfde_relu_bwd_bf16(const void* y, const void* dy, const void* addend, void* dx, long n, tfde_stream_t stream);
int tfde_cast_f32_bf16(const float* x, void* y, long n, tfde_stream_t stream);
int tfde_maxpool_fwd_bf16(const void* x, void* y, uint8_t* idx, int Bn, int H, int W, int C, int PH, int PW, int sh,
                          int sw, int pt, int pl, int Ho, int Wo, tfde_stream_t stream);
int tfde_maxpool_bwd_bf16(const void* dy, const uint8_t* idx, void* dx, int Bn, int H, int W, int C, int PH, int PW,
                          int sh, int sw, int pt, int pl, int Ho, int Wo, tfde_stream_t stream);
int tfde_gap_fwd_bf16(const void* x, float* y, int Bn, int HW, int C, tfde_stream_t stream);
int tfde_gap_bwd_bf16(const float* dy, void* dx, int Bn, int HW, int C, tfde_stream_t stream);
/* bf16 -> bf16 operand gathers (one warp per row, 16 bytes per lane) and the compact strided-1x1 data gradient */
int tfde_im2col_b2b(const void* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh,
                    int sw, int pt, int pl, int Cout, int ldo, tfde_stream_t stream);
int tfde_dgrad_gather_b2b(const void* dy, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                          int sh, int sw, int pt, int pl, int Cout, int ldo, tfde_stream_t stream);
int tfde_strided_scatter_b2b(const void* g, void* dx, int Bn, int H, int W, int C, int Ho, int Wo, int sh, int sw,
                             tfde_stream_t stream);

/* ---- NVLS all-reduce (allreduce_p2p.cu, mode 2 of tfde_allreduce_fused) ---------------------------------- */
/* Bytes of the multicast-bound signal pad the NVLS variant needs ([2 phases][256 blocks] counters). The
 * descriptor's mc_g / mc_w / mc_flags are multicast mappings (multimem.ld_reduce / multimem.st / multimem.red),
 * nv_flags this rank's ordinary mapping of the signal pad. ref: the NCCL all-reduce behind
 * MultiWorkerMirroredStrategy, distributed_with_keras.py:16,63. */
size_t tfde_ar_nvls_flag_bytes(void);

/* ---- diagnostics (debug_util.cu) ---------------------------------------------------------------------- */
int tfde_capture_status(tfde_stream_t stream); /* 0 none, 1 capturing, 2 capture invalidated */
int tfde_clear_last_error(void);

#ifdef __cplusplus
}
#endif
