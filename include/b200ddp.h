/*
 * b200ddp — C ABI of the native library (libb200ddp.so).
 *
 * Everything the Python layer calls goes through these entry points (ctypes).  All pointers are
 * device pointers unless a name says "host"; every launch function takes the CUDA stream to launch
 * on (as a void*, i.e. a cudaStream_t) and returns 0 on success or a negative b200 error / a positive
 * cudaError_t.  b200_last_error() gives the text of the last failure on the calling thread.
 *
 * The reference (Neela08/cloud-security-pytorch-sagemaker-distributed-workshop) has no native code of
 * its own; each group below names the external native library call it replaces on the reference's hot
 * path and the reference line that reaches it.
 */
#ifndef B200DDP_H_
#define B200DDP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_MAX_RANKS 8
#define B200_MAX_COMM_BLOCKS 160
#define B200_SIGNAL_BYTES (1u << 20) /* first MiB of every symmetric arena: signal pads + epochs */

#define B200_DTYPE_F32 0
#define B200_DTYPE_BF16 1

const char* b200_last_error(void);
int b200_version(void);

/* ---------------------------------------------------------------------------------------------
 * Symmetric memory (replaces NCCL's P2P/IPC buffer registration; notebooks/2_*.ipynb:400
 * "via P2P/IPC/read").  cuMemCreate + POSIX-fd export, peer import/map, NVSwitch multicast object.
 * ------------------------------------------------------------------------------------------- */
int b200_symm_init(int device);
int b200_symm_granularity(int device, size_t* gran_out, size_t* mc_gran_out);
int b200_symm_alloc(int device, size_t bytes, void** ptr_out, int* fd_out, uint64_t* handle_out);
int b200_symm_import(int device, int fd, size_t bytes, void** ptr_out, uint64_t* handle_out);
int b200_symm_unmap(void* ptr, size_t bytes, uint64_t handle);
int b200_mc_supported(int device);
int b200_mc_create(int ndev, size_t bytes, uint64_t* mc_handle_out, int* fd_out);
int b200_mc_import(int fd, uint64_t* mc_handle_out);
int b200_mc_add_device(uint64_t mc_handle, int device);
int b200_mc_bind_map(int device, uint64_t mc_handle, uint64_t mem_handle, size_t bytes, void** mc_ptr_out);
int b200_mc_release(int device, uint64_t mc_handle, void* mc_ptr, size_t bytes);
int b200_close_fd(int fd);

/* Communicator context: peer base pointers of the symmetric arena (peer[rank] is the local
 * mapping), multicast base (or NULL), host-mapped error flag.  Passed by value to the kernels. */
typedef struct b200_comm {
  int rank;
  int world;
  char* peer[B200_MAX_RANKS];
  char* mc;
  int* err_flag;               /* device pointer to a host-mapped int; non-zero => spin timeout */
  unsigned long long timeout_ns;
} b200_comm;

int b200_comm_alloc_errflag(int** host_ptr_out, int** dev_ptr_out);

/* ---------------------------------------------------------------------------------------------
 * Collectives over the symmetric arena (replace NCCL ring allreduce/broadcast under SMDDP —
 * cifar10-distributed-smddp-gpu.py:148,167; notebooks/2_*.ipynb:350-400).  Offsets are bytes from the
 * arena base and must be 16-byte aligned; `count` is elements of `dtype`.  All fuse `* scale`
 * (gradient averaging) into the reduction.  `algo`: 0 auto, 1 one-shot, 2 two-shot (P2P
 * reduce + push), 3 NVLS (multimem.ld_reduce + multimem.st).  `max_blocks` caps the grid.
 * ------------------------------------------------------------------------------------------- */
#define B200_ALGO_AUTO 0
#define B200_ALGO_ONESHOT 1
#define B200_ALGO_TWOSHOT 2
#define B200_ALGO_NVLS 3
int b200_allreduce(const b200_comm* comm, size_t offset, size_t count, int dtype, float scale,
                   int algo, int max_blocks, void* stream);
/* The one algorithm selector (used by B200_ALGO_AUTO, by the reducer and by the Python layer): one-shot up to
 * oneshot_max_bytes (0 = max(64 KiB, 2 MiB / (world - 1))), then NVLS if a multicast mapping is bound, world > 2 and
 * bytes >= nvls_min_bytes, else two-shot. */
int b200_pick_algo(int world, int has_multicast, size_t bytes, size_t oneshot_max_bytes, size_t nvls_min_bytes);
/* Fused bucket step (the reference's `loss.backward()` all-reduce + `optimizer.step()`, …smddp-gpu.py:167-168, and SMDDP's
 * "balanced fusion buffer", static/images/training/training24.png): rank r reduces shard r of the bucket's fp32 gradients
 * (NVLS ld_reduce or P2P loads), applies SGD with momentum to shard r of the fp32 master weights, writes the new weights
 * (and their bf16 shadow when shadow_offset != 0) into every rank's arena (multimem.st or P2P stores) and zeroes the local
 * gradients -- one kernel.  All offsets are bytes inside the symmetric arena; `momentum_buf` is local (NULL if momentum == 0). */
int b200_allreduce_sgd(const b200_comm* comm, size_t grad_offset, size_t param_offset, size_t shadow_offset,
                       float* momentum_buf, size_t count, float scale, float lr, float momentum, float dampening,
                       float weight_decay, int nesterov, int algo, int max_blocks, void* stream);
/* One-shot all-reduce from the symmetric arena into an arbitrary local buffer (out-of-place). */
int b200_allreduce_oneshot_out(const b200_comm* comm, size_t offset, void* out, size_t count, int dtype,
                               float scale, int max_blocks, void* stream);
int b200_broadcast(const b200_comm* comm, size_t offset, size_t bytes, int root, int max_blocks, void* stream);
int b200_barrier(const b200_comm* comm, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Gradient reducer (replaces torch's C++ DDP Reducer — …native-cpu.py:141, …smddp-gpu.py:148): ready
 * counting per bucket, in-order launch of b200_allreduce on the comm stream behind stream events, join at
 * the end of backward.  Buckets are byte ranges of the symmetric gradient arena.  Stream-ordered only.
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_reducer b200_reducer;
b200_reducer* b200_reducer_create(const b200_comm* comm, int n_params, const int* param_bucket, int n_buckets,
                                  const size_t* bucket_offset_bytes, const size_t* bucket_count, int dtype,
                                  float scale, int max_blocks, size_t oneshot_max_bytes, size_t nvls_min_bytes);
void b200_reducer_destroy(b200_reducer* r);
int b200_reducer_begin(b200_reducer* r);
int b200_reducer_was_ready(const b200_reducer* r, int param);
int b200_reducer_mark_ready(b200_reducer* r, int param, void* producer_stream, void* other_stream, void* comm_stream);
int b200_reducer_finalize(b200_reducer* r, void* main_stream, void* other_stream, void* comm_stream, int sync_grads);
int b200_reducer_bucket_algo(const b200_reducer* r, int bucket);
/* enabled != 0: every bucket launch becomes b200_allreduce_sgd (see there); base offsets are those of element 0 of the flat
 * gradient / parameter / shadow arenas, momentum_base the local momentum buffer laid out like the parameters */
int b200_reducer_set_sgd(b200_reducer* r, int enabled, size_t grad_base, size_t param_base, size_t shadow_base,
                         float* momentum_base, float lr, float momentum, float dampening, float weight_decay, int nesterov);

/* ---------------------------------------------------------------------------------------------
 * Optimizers over flat arenas (replace ATen foreach SGD / Adam — …native-cpu.py:144,156;
 * utils_basic.py:96).  p/g/m/v are fp32; shadow (may be NULL) receives the bf16 copy of p.
 * ------------------------------------------------------------------------------------------- */
int b200_sgd_momentum(float* p, float* g, float* m, void* shadow_bf16, size_t n, float lr, float momentum,
                      float dampening, float weight_decay, int nesterov, float grad_scale, int zero_grad,
                      void* stream);
int b200_adam(float* p, float* g, float* m, float* v, void* shadow_bf16, size_t n, float lr, float beta1,
              float beta2, float eps, float weight_decay, const int* step_dev, float grad_scale,
              int zero_grad, void* stream);
int b200_step_inc(int* step_dev, void* stream);
int b200_cast_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Losses / metrics (replace ATen log_softmax+nll_loss, BCE-with-logits, argmax/eq/sum —
 * …native-cpu.py:143,152,185-187; model_lib/rtNLP_cnn_model.py:64).
 * ------------------------------------------------------------------------------------------- */
int b200_softmax_xent(const void* logits, int dtype, const int64_t* targets, int rows, int cols,
                      float* row_loss, float* loss_mean, void* dlogits, float grad_scale,
                      unsigned int* scratch_counter, void* stream);
int b200_bce_logits(const void* logits, int dtype, const float* targets, int n, float* loss_mean,
                    void* dlogits, float grad_scale, void* stream);
int b200_eval_count(const void* logits, int dtype, const int64_t* targets, int rows, int cols,
                    int* correct_out, float* nll_sum_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Batch norm (+ReLU, +residual), pools, augmentation — NHWC bf16 activations
 * (replace cuDNN BN / ATen relu, add, pools — torchvision resnet18 under …smddp-gpu.py:32).
 * ------------------------------------------------------------------------------------------- */
/* Scratch contract of the BN reductions: `scratch` is global memory beyond the op's inputs/outputs
 * (CTA partial sums + ticket counters).  It must be zero before its first use; every launch leaves
 * the counters at zero again.  Size: b200_bn_scratch_bytes(rows, C). */
size_t b200_bn_scratch_bytes(size_t rows, int C);
/* stat is [4][C] fp32: mean, rstd, scale = gamma*rstd, shift = beta - mean*scale. */
int b200_bn_stats(const void* x_bf16, size_t rows, int C, const float* gamma, const float* beta, float eps,
                  float momentum, float* stat, float* running_mean, float* running_var, void* scratch,
                  size_t scratch_bytes, void* stream);
/* No counterpart in the reference (it has no native code, and no allocation-free calling convention).
 * Same result as b200_bn_stats, for callers that cannot allocate device memory — a process whose own pool
 * already owns the whole GPU, a serving / real-time loop that cannot take cudaMalloc's device-wide
 * synchronisation, a plain-C runtime with no caching allocator to lend a buffer.  The scratch is a fixed
 * B200_BN_NOALLOC_SCRATCH_BYTES array inside the library (held by every GPU that loads it, until exit);
 * problems whose b200_bn_scratch_bytes(rows, C) exceed it are rejected with an error.  The first call on a
 * device looks the array's address up (cudaGetSymbolAddress): make it before any stream capture.  No
 * cudaMalloc/cudaFree, no allocating library call.  Calls share the array and clear its ticket counters on
 * `stream` first; calls that overlap on the GPU race — keep them on one stream or order them with an event. */
#define B200_BN_NOALLOC_SCRATCH_BYTES (1u << 20)
int b200_bn_stats_noalloc(const void* x_bf16, size_t rows, int C, const float* gamma, const float* beta,
                          float eps, float momentum, float* stat, float* running_mean, float* running_var,
                          void* stream);
int b200_bn_eval_prep(const float* gamma, const float* beta, const float* running_mean,
                      const float* running_var, float eps, float* stat, int C, void* stream);
int b200_bn_apply(const void* x_bf16, const void* residual_bf16, void* y_bf16, size_t rows, int C,
                  const float* stat, int relu, void* stream);
/* sums is [2][C]: sum(g), sum(g*xhat) with g = dy * relu_mask; dgamma/dbeta (may be NULL) are written
 * as beta_accum*old + new, i.e. straight into the flat gradient arena. */
int b200_bn_bwd_reduce(const void* dy_bf16, const void* x_bf16, const void* y_bf16, size_t rows, int C,
                       const float* stat, int relu, float* sums, float* dgamma, float* dbeta,
                       float beta_accum, void* scratch, size_t scratch_bytes, void* stream);
int b200_bn_bwd_apply(const void* dy_bf16, const void* x_bf16, const void* y_bf16, void* dx_bf16,
                      void* dres_bf16, size_t rows, int C, const float* stat, const float* sums, int relu,
                      void* stream);
/* One-launch variants for small tensors (rows <= b200_bn_small_rows()): statistics+apply, reduce+apply. */
size_t b200_bn_small_rows(void);
/* b200_bn_bwd_fused: one CTA per 64 channels up to B200DDP_BN_CLUSTER_MIN_ROWS rows, a cluster per 8 channels above */
size_t b200_bn_bwd_fused_max_rows(void);
int b200_bn_fwd_fused(const void* x_bf16, const void* residual_bf16, void* y_bf16, size_t rows, int C,
                      const float* gamma, const float* beta, float eps, float momentum, float* stat,
                      float* running_mean, float* running_var, int relu, void* stream);
int b200_bn_bwd_fused(const void* dy_bf16, const void* x_bf16, const void* y_bf16, void* dx_bf16, void* dres_bf16,
                      size_t rows, int C, const float* stat, int relu, float* dgamma, float* dbeta, float beta_accum,
                      void* stream);
int b200_maxpool_fwd(const void* x, void* y, uint8_t* argmax, int N, int H, int W, int C, int k, int stride,
                     int pad, void* stream);
int b200_maxpool_bwd(const void* dy, const uint8_t* argmax, void* dx, int N, int H, int W, int C, int k,
                     int stride, int pad, void* stream);
int b200_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int b200_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
int b200_relu_bwd(const void* dy, const void* y, void* dx, size_t n, void* stream);
/* out[c] += sum over rows of dy[row][c] for c < c_valid (bias gradients; replaces ATen sum). */
int b200_colsum_accum(const void* dy_bf16, size_t rows, int C, int c_valid, float* out, void* stream);
/* layout 0: out = NHWC bf16 with C_out >= C channels (zero padded); layout 1: NCHW fp32.
 * offsets [N][3] int32 = (crop dy, crop dx in 0..2*pad, flip) or NULL for the identity crop. */
int b200_augment(const uint8_t* img_u8, const int32_t* offsets, void* out, int N, int H, int W, int C,
                 int C_out, int pad, const float* mean, const float* inv_std, int layout, void* stream);
/* x is NCHW fp32; pattern [pattern_channels][p_h][p_w], pattern_channels in {1, C}; mask[N] (or NULL). */
int b200_trojan_stamp(float* x, int N, int C, int H, int W, const float* pattern, int pattern_channels,
                      int p_h, int p_w, int loc_h, int loc_w, float alpha, const uint8_t* mask, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Workshop `Net` megakernel (…native-cpu.py:22-39,146-156): forward + softmax-CE + backward of the LeNet-style
 * CIFAR-10 CNN in two launches, fp32.  params/grads: 10 pointers in state_dict order; grads (may be NULL) are
 * ACCUMULATED (flat arena semantics); krsc != 0: conv weights/grads are channels-last [co][r][s][ci].
 * scratch: b200_net_scratch_bytes(batch) bytes, zero before first use (fc1 operands + a ticket counter).
 * ------------------------------------------------------------------------------------------- */
size_t b200_net_scratch_bytes(int batch);
/* debug: 32 clock64 stamps of CTA 0's phases go to `buf` (device memory; NULL stops) */
void b200_net_set_profile_buffer(long long* buf);
int b200_net_fwd_bwd(const float* x_nchw, const int64_t* targets, int batch, const float* const* params,
                     float* const* grads, int krsc, float* row_loss, float* loss_mean, float* logits,
                     void* scratch, size_t scratch_bytes, void* stream);
/* same, from the raw uint8 NHWC batch: RandomCrop(32,4)+flip (offsets [B][3] = dy, dx, flip; NULL = none) and
   Normalize(mean, 1/std) are applied while the image is staged into shared memory */
int b200_net_fwd_bwd_u8(const uint8_t* img_nhwc, const int32_t* offsets, const float* mean, const float* inv_std,
                        const int64_t* targets, int batch, const float* const* params, float* const* grads, int krsc,
                        float* row_loss, float* loss_mean, float* logits, void* scratch, size_t scratch_bytes,
                        void* stream);

/* ---------------------------------------------------------------------------------------------
 * Tensor-core layers: implicit-GEMM convolution and linear (a 1x1 convolution on [M,1,1,K]) on
 * tcgen05 + TMEM (replace cuDNN conv fwd/dgrad/wgrad and cuBLAS GEMM under nn.Conv2d / nn.Linear —
 * torchvision resnet18 via …smddp-gpu.py:32; model_lib/cifar10_cnn_model.py:11-18).
 * x: NHWC bf16 [N][H][W][C]; w: bf16 [K][R][S][C]; y/dy: NHWC bf16 [N][Ho][Wo][K]; dw: fp32 [K][R][S][C].
 * C and K must be multiples of 8.  b200_conv_wgrad ACCUMULATES into dw (fp32 red.add, split-K).
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_conv_geom {
  int N, H, W, C; /* input */
  int K;          /* output channels */
  int R, S, stride, pad;
} b200_conv_geom;
/* Small problems split K over a thread-block cluster (DSMEM reduction; no scratch).  b200_conv_set_split_k(0)
 * pins fwd/dgrad to one CTA per tile. */
int b200_conv_set_split_k(int enabled);
/* Tuning aid: when non-NULL, CTA (0,0,0) of every conv launch writes 8 clock64() phase stamps there. */
int b200_conv_set_profile_buffer(long long* dev_buf16);   /* 16 clock64 slots, tuning aid */
int b200_conv_fwd(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g, int relu,
                  void* stream);
/* the same convolution without bias/ReLU whose epilogue also accumulates per-channel sum / sum of squares of the output
   into bn_sums[2][Kc] (zero on entry): the statistics of a following training-mode BatchNorm; see b200_bn_finalize */
int b200_conv_fwd_stats(const void* x, const void* w, void* y, const b200_conv_geom* g, float* bn_sums, void* stream);
int b200_bn_finalize(float* sums, size_t rows, int C, const float* gamma, const float* beta, float eps, float momentum,
                     float* stat, float* running_mean, float* running_var, void* stream);
/* bn_finalize + bn_apply in one launch; sums has 2C floats + one ticket word, zero on entry and on exit */
int b200_bn_apply_sums(const void* x_bf16, const void* residual_bf16, void* y_bf16, size_t rows, int C, float* sums,
                       const float* gamma, const float* beta, float eps, float momentum, float* stat,
                       float* running_mean, float* running_var, int relu, void* stream);
/* same launch without the ticket and without clearing `sums`: the caller zeroes all its sum buffers at once before the next
   forward (one pooled buffer, one memset per step) */
int b200_bn_apply_sums_keep(const void* x_bf16, const void* residual_bf16, void* y_bf16, size_t rows, int C, float* sums,
                            const float* gamma, const float* beta, float eps, float momentum, float* stat,
                            float* running_mean, float* running_var, int relu, void* stream);
int b200_conv_dgrad(const void* dy, const void* w, void* dx, const b200_conv_geom* g, void* stream);
/* dx = dgrad(dy, w) + residual: the gradient of the other branch of a forked tensor (ResNet skip) added in the epilogue */
int b200_conv_dgrad_add(const void* dy, const void* w, void* dx, const b200_conv_geom* g, const void* residual, void* stream);
/* dw (fp32, accumulated into) += wgrad(x, dy).  max_splits: 0 = automatic (the reduction over positions is split over
   ~60 % of the SMs), > 0 = cap on the number of splits, < 0 = use -max_splits percent of the SMs (nothing runs alongside) */
int b200_conv_wgrad(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits,
                    void* stream);

/* ---------------------------------------------------------------------------------------------
 * fp32 activation variants.  Every entry point above that reads or writes activations exists a second time with the
 * suffix _f32 and the same argument list: activations, residuals, conv / linear weights and their outputs are then NHWC
 * (or [M][K]) fp32 instead of bf16, convolutions and linear layers run tcgen05.mma kind::tf32 (fp32 operands, the tensor
 * core multiplies their top 19 bits, fp32 accumulation), everything else computes in fp32 as before.  This is the
 * precision of the reference's own training run (fp32 model, cuDNN TF32 convolutions, no autocast:
 * notebooks/code/cifar10-distributed-smddp-gpu.py:145,155-168).  b200_augment writes fp32 NHWC with layout == 2.
 * ------------------------------------------------------------------------------------------- */
int b200_conv_fwd_f32(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g, int relu, void* stream);
int b200_conv_fwd_stats_f32(const void* x, const void* w, void* y, const b200_conv_geom* g, float* bn_sums, void* stream);
int b200_conv_dgrad_f32(const void* dy, const void* w, void* dx, const b200_conv_geom* g, void* stream);
int b200_conv_dgrad_add_f32(const void* dy, const void* w, void* dx, const b200_conv_geom* g, const void* residual, void* stream);
int b200_conv_wgrad_f32(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits, void* stream);
int b200_bn_stats_f32(const void* x, size_t rows, int C, const float* gamma, const float* beta, float eps, float momentum,
                      float* stat, float* running_mean, float* running_var, void* scratch, size_t scratch_bytes, void* stream);
int b200_bn_apply_f32(const void* x, const void* residual, void* y, size_t rows, int C, const float* stat, int relu, void* stream);
int b200_bn_apply_sums_f32(const void* x, const void* residual, void* y, size_t rows, int C, float* sums, const float* gamma,
                           const float* beta, float eps, float momentum, float* stat, float* running_mean, float* running_var,
                           int relu, void* stream);
int b200_bn_apply_sums_keep_f32(const void* x, const void* residual, void* y, size_t rows, int C, float* sums, const float* gamma,
                                const float* beta, float eps, float momentum, float* stat, float* running_mean,
                                float* running_var, int relu, void* stream);
int b200_bn_bwd_reduce_f32(const void* dy, const void* x, const void* y, size_t rows, int C, const float* stat, int relu,
                           float* sums, float* dgamma, float* dbeta, float beta_accum, void* scratch, size_t scratch_bytes,
                           void* stream);
int b200_bn_bwd_apply_f32(const void* dy, const void* x, const void* y, void* dx, void* dres, size_t rows, int C,
                          const float* stat, const float* sums, int relu, void* stream);
int b200_bn_fwd_fused_f32(const void* x, const void* residual, void* y, size_t rows, int C, const float* gamma, const float* beta,
                          float eps, float momentum, float* stat, float* running_mean, float* running_var, int relu, void* stream);
size_t b200_bn_bwd_fused_max_rows_f32(void);
int b200_bn_bwd_fused_f32(const void* dy, const void* x, const void* y, void* dx, void* dres, size_t rows, int C, const float* stat,
                          int relu, float* dgamma, float* dbeta, float beta_accum, void* stream);
int b200_maxpool_fwd_f32(const void* x, void* y, uint8_t* argmax, int N, int H, int W, int C, int k, int stride, int pad, void* stream);
int b200_maxpool_bwd_f32(const void* dy, const uint8_t* argmax, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                         void* stream);
int b200_avgpool_fwd_f32(const void* x, void* y, int N, int HW, int C, void* stream);
int b200_avgpool_bwd_f32(const void* dy, void* dx, int N, int HW, int C, void* stream);
int b200_relu_bwd_f32(const void* dy, const void* y, void* dx, size_t n, void* stream);
int b200_colsum_accum_f32(const void* dy, size_t rows, int C, int c_valid, float* out, void* stream);

/* ---- sequence models of the trojan-detection zoo (csrc/seq.cu), fp32 ---- */
/* C[m][n] = sum_k A[m*sam + k*sak] * B[k*sbk + n*sbn] (+ bias[n]) (+ beta*C) (optional ReLU); transposes are strides */
int b200_sgemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long sam,
               long long sak, long long sbk, long long sbn, int ldc, float beta, int relu, void* stream);
/* one LSTM layer's recurrence (gate order i,f,g,o; H == 100): pre [B][T][4H] = x W_ih^T + b_ih + b_hh */
int b200_lstm_rec_fwd(const float* pre, const float* w_hh, float* hseq, float* gates, float* cseq, int B, int T, int H,
                      void* stream);
int b200_lstm_rec_bwd(const float* dh, const float* gates, const float* cseq, const float* w_hh, float* dgates, int B,
                      int T, int H, void* stream);
/* waveform [B][L] -> (10 log10(mel power) + 50) / 50, [B][1 + L/512][n_mels]; 2048-point Hann STFT, hop 512, reflect */
int b200_mel_frontend(const float* x, const float* window, const float* melfb, float* out, int B, int L, int n_mels,
                      void* stream);
/* gradient w.r.t. the waveform, accumulated into dx [B][L] (zero it first) */
int b200_mel_frontend_bwd(const float* x, const float* window, const float* melfb, const float* dout, float* dx, int B,
                          int L, int n_mels, void* stream);
int b200_gather_rows(const float* table, const int64_t* idx, float* out, long long rows, int dim, long long vocab,
                     void* stream);
/* bias + ReLU + max-over-time of a [B*L][N] conv-as-GEMM output (window length k), and its backward */
int b200_textcnn_pool_fwd(const float* C, const float* bias, float* feat, int32_t* arg, int B, int L, int N, int k, int ldf,
                          int col, void* stream);
int b200_textcnn_pool_bwd(const float* dfeat, const float* feat, const int32_t* arg, float* dC, int B, int L, int N, int ldf,
                          int col, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDP_H_ */
