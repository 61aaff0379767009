/* C ABI of libddpx_kernels.so — hand-written sm_100a kernels (no cuBLAS / cuDNN / NCCL inside).
 *
 * Every entry point takes raw device pointers plus a CUDA stream (as void*) and returns a DDPX_* code;
 * nothing here allocates unless the comment says so. Python reaches these through ctypes
 * (distributed_data_parallel_experiment_b200/_lib.py).
 *
 * Reference parity citations are to /root/reference/multigpu.py (M:) and singlegpu.py (S:).
 */
#ifndef DDPX_KERNELS_H_
#define DDPX_KERNELS_H_

#ifdef __cplusplus
extern "C" {
#endif

enum {
  DDPX_OK = 0,
  DDPX_ERR_ARG = 1,     /* invalid argument combination */
  DDPX_ERR_ALIGN = 2,   /* pointer / leading dimension not 16-byte aligned */
  DDPX_ERR_SHAPE = 3,   /* shape outside what the kernel tiles */
  DDPX_ERR_DRIVER = 4,  /* driver entry point unavailable */
  DDPX_ERR_TMAP = 5,    /* cuTensorMapEncodeTiled rejected the descriptor */
  DDPX_ERR_LAUNCH = 6,  /* kernel launch failed */
  DDPX_ERR_LIMIT = 7    /* size beyond a stated maximum */
};

/* ---- tcgen05 / TMEM / TMA GEMM --------------------------------------------------------------
 * D[M,N] (=|+=) op(A) * op(B), fp32 accumulation in TMEM.
 *   a_mn = 0: A stored [M][K] (K contiguous, ld = lda)   a_mn = 1: A stored [K][M] (M contiguous)
 *   b_mn = 0: B stored [N][K] (K contiguous, ld = ldb)   b_mn = 1: B stored [K][N] (N contiguous)
 *   bf16 = 0: operands fp32 in memory, multiplied as TF32 (what cuDNN/cuBLAS-TF32 do for fp32 tensors)
 *   bf16 = 1: operands bf16.   out_bf16 selects the output type.  bias is per output column (fp32).
 *   act: 0 none, 1 ReLU, 2 GELU(tanh).  split_k > 1 or accumulate != 0 -> fp32 atomics into D.
 * Replaces cuBLAS behind nn.Linear (M:62, S:36-41) and the MLP / GPT-2 GEMMs of BASELINE.json. */
int ddpx_gemm(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
              long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const float* bias, int act,
              int split_k, int accumulate, void* stream);

/* the same with the bias in bf16 (bias_bf16 = 1) or fp32 (0): transformer layers keep their biases in the activation dtype */
int ddpx_gemm_bias(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                   int a_mn, int b_mn, int bf16, int out_bf16, const void* bias, int bias_bf16, int act, int split_k,
                   int accumulate, void* stream);
/* bf16 GEMM with an elementwise pass folded into the store: aux_op 1: D = act(A B + bias) + aux (residual); aux_op 2:
 * D = (A B) * gelu'(aux); out2 (act = 2 only): the pre-activation is stored there as well.  aux / out2: bf16 [M][ldd]. */
int ddpx_gemm_fused(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                    int a_mn, int b_mn, const void* bias, int bias_bf16, int act, const void* aux, int aux_op, void* out2,
                    void* stream);

/* Split-K GEMM, bf16 in and out, with the fix-up inside the kernel (no memset / cast launches): partial tiles are reduced into
 * ws (fp32 [M*N]); the epilogue warp that delivers the last partial sum of a sub-block converts it into D and clears ws.
 * ws and cnt (uint32 [(ceil(M/128)*ceil(N/64)+1)*16]) must be all zero on entry and are left all zero. */
int ddpx_gemm_splitk_bf16(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                          long long ldd, int a_mn, int b_mn, int split_k, float* ws, unsigned int* cnt, void* stream);

/* Batched form: problem (b1,b2), b1 < batch1, b2 < batch2, reads A + b1*a_bs1 + b2*a_bs2 (element strides), same for
 * B and D; D = act(alpha * A*B + bias).  causal = 1 skips output tiles entirely above the diagonal (attention scores),
 * causal = 2 stops the K loop at the diagonal (probabilities x values).  Used for the attention GEMMs of GPT-2. */
int ddpx_gemm_batched(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                      long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const float* bias, int act, float alpha,
                      int split_k, int accumulate, int batch1, int batch2, long long a_bs1, long long a_bs2,
                      long long b_bs1, long long b_bs2, long long d_bs1, long long d_bs2, int causal, void* stream);

/* ---- 3x3 / stride 1 / pad 1 convolution, NHWC fp32, TF32 tensor cores (M:53) ----------------
 * fprop: y[N,H,W,Cout] = conv(x[N,H,W,Cin], wpacked[9][Cout][Cin]) (+bias[Cout]) (+ReLU);
 *        optionally accumulates per-channel sum / sum-of-squares of the pre-activation output into
 *        stat_sum / stat_sq (BatchNorm batch statistics fused into the conv epilogue; pre-zeroed).
 *        The data gradient is the same call with tap-flipped, transposed weights (see ddpx_pack_conv_weights).
 * wgrad: dw[Cout][Cin][3][3] += sum_pixels dy (x) shifted x   (torch layout, fp32 atomics; pre-zeroed or
 *        already holding a gradient to accumulate into). */
int ddpx_conv3x3_fprop(const float* x, const float* wpacked, float* y, int Nimg, int H, int W, int Cin, int Cout,
                       const float* bias, int act, float* stat_sum, float* stat_sq, void* stream);
int ddpx_conv3x3_wgrad(const float* dy, const float* x, float* dw, int Nimg, int H, int W, int Cin, int Cout,
                       int split_k, void* stream);
/* Weight gradient into the tap-major staging layout dw_packed[9][Cout][Cin] (caller zeroes it; 16-byte vector reductions),
 * and the transposition of that layout into torch's [Cout][Cin][3][3]. */
int ddpx_conv3x3_wgrad_packed(const float* dy, const float* x, float* dw_packed, int Nimg, int H, int W, int Cin, int Cout,
                              int split_k, void* stream);
int ddpx_unpack_conv_wgrad(const float* dw_packed, float* dw, int Cout, int Cin, void* stream);
/* same, and the staging buffer is left all-zero: a persistent per-layer staging buffer never needs a memset launch */
int ddpx_unpack_conv_wgrad_rezero(float* dw_packed, float* dw, int Cout, int Cin, void* stream);
/* RGB stem on tensor cores (M:53 conv0, K = 27 padded to 32): patches[P][32] = im2col(x_nchw); forward =
 * ddpx_conv1x1_fprop(patches, wp[Cout][32]) with the same epilogue as the 3x3 kernel; weight gradient =
 * ddpx_gemm(dy^T, patches) with split-K. */
int ddpx_im2col_stem(const float* x_nchw, float* patches, int Nimg, int H, int W, void* stream);
int ddpx_pack_stem_weights(const float* w, float* wp, int Cout, void* stream);
int ddpx_conv1x1_fprop(const float* x, const float* wpacked, float* y, int Nimg, int H, int W, int Cin, int Cout,
                       const float* bias, int act, float* stat_sum, float* stat_sq, void* stream);


/* ---- convolution helpers (csrc/kernels/conv_aux.cu) ---------------------------------------- */
/* torch-layout weights [Cout][Cin][3][3] -> w_fprop [tap][Cout][Cin] and/or w_dgrad [8-tap][Cin][Cout] (either may be NULL) */
int ddpx_pack_conv_weights(const float* w, float* w_fprop, float* w_dgrad, int Cout, int Cin, void* stream);
/* the same for n <= 32 layers in ONE launch; zero[i] / zero_n[i]: floats to clear per layer (BatchNorm accumulators), may
 * be NULL; Cin[i] == 3 marks the RGB stem (wf[i] = [Cout][32] zero-padded rows, wd[i] ignored) */
int ddpx_pack_conv_weights_multi(int n, const void* const* w, void* const* wf, void* const* wd, void* const* zero,
                                 const int* zero_n, const int* Cout, const int* Cin, void* stream);
/* RGB stem (Cin == 3, K = 27): direct conv, NCHW input -> NHWC output, optional bias / ReLU / BN statistics (M:53 conv0) */
int ddpx_conv_stem_fprop(const float* x_nchw, const float* w, const float* bias, float* y_nhwc, int Nimg, int H, int W,
                         int Cout, int act, float* stat_sum, float* stat_sq, void* stream);
int ddpx_conv_stem_wgrad(const float* dy_nhwc, const float* x_nchw, float* dw, int Nimg, int H, int W, int Cout, void* stream);
/* [batch][rows][cols] -> [batch][cols][rows] (NCHW <-> NHWC) */
int ddpx_transpose_batched(const float* in, float* out, int batch, int rows, int cols, void* stream);
int ddpx_relu_fwd(const float* x, float* y, long long n, void* stream);
int ddpx_relu_bwd(const float* y, const float* dy, float* dx, long long n, void* stream);
int ddpx_maxpool2x2_fwd(const float* x, float* y, int Nimg, int H, int W, int C, void* stream);           /* NHWC, M:59 */
int ddpx_maxpool2x2_bwd(const float* x, const float* dy, float* dx, int Nimg, int H, int W, int C, void* stream);
int ddpx_spatial_mean_fwd(const float* x, float* y, int Nimg, int HW, int C, void* stream);               /* M:68 */
int ddpx_spatial_mean_bwd(const float* dy, float* dx, int Nimg, int HW, int C, void* stream);
int ddpx_colsum(const float* x, float* out, long long rows, int cols, int accumulate, void* stream);      /* bias grads */
int ddpx_scale_copy(const float* src, float* dst, long long n, float scale, int accumulate, void* stream);
int ddpx_zero(void* p, long long bytes, void* stream);
/* CUDA-core fp32 GEMM for shapes TMA cannot describe (row pitch not a multiple of 16 B): classifier Linear(512,10) at
 * M:62, toy Linear(20,1).  C[m][n] = act(sum_k A[m*sam+k*sak] * B[n*sbn+k*sbk] + bias[n]) (+C). */
int ddpx_gemm_small(const float* A, const float* B, float* C, int M, int N, int K, long long sam, long long sak,
                    long long sbn, long long sbk, long long ldc, const float* bias, int act, int accumulate, void* stream);

/* ---- BatchNorm2d (+ReLU, +2x2 max-pool) on NHWC fp32 (csrc/kernels/batchnorm.cu; M:54-59) -- */
int ddpx_bn_finalize(const float* sum, const float* sq, float count, float* mean, float* rstd, float* running_mean,
                     float* running_var, long long* num_batches_tracked, float momentum, float eps, int C, void* stream);
int ddpx_bn_eval_stats(const float* running_mean, const float* running_var, float eps, float* mean, float* rstd, int C,
                       void* stream);
int ddpx_bn_stats(const float* x, long long P, int C, float* sum, float* sq, void* stream);
int ddpx_bn_apply(const float* x, float* y, const float* mean, const float* rstd, const float* gamma, const float* beta,
                  int Nimg, int H, int W, int C, int relu, int pool, void* stream);
/* finalize + running-statistics update + normalise (+ReLU, +pool) in ONE launch from the convolution epilogue's raw
 * (sum, sum of squares); training = 0 normalises with the running statistics.  mean_rstd [2][C] is an output. */
int ddpx_bn_apply_stats(const float* x, float* y, const float* sum, const float* sq, float count, float eps, float momentum,
                        const float* gamma, const float* beta, float* running_mean, float* running_var,
                        long long* num_batches_tracked, float* mean_rstd, int training, int Nimg, int H, int W, int C,
                        int relu, int pool, void* stream);
/* one-call training forward: statistics -> finalize -> apply; `workspace` = 2*C floats of scratch (cleared by the call),
 * `mean_rstd` = [2][C] output saved for the backward */
int ddpx_bn_train_fwd(const float* x, float* y, const float* gamma, const float* beta, float* running_mean,
                      float* running_var, long long* num_batches_tracked, float momentum, float eps, int Nimg, int H, int W,
                      int C, int relu, int pool, float* mean_rstd, float* workspace, void* stream);
/* ddpx_bn_train_fwd without the workspace argument, for callers that cannot allocate device memory (they were handed
 * their tensors and own no allocator, their memory was committed up front, or they must not stall the device with
 * cudaMalloc).  The accumulators live in an 8 KiB array inside the library (2 x DDPX_BN_STATIC_MAX_C = 1024 channels of
 * fp32; resident on every GPU that loads the library).  Allocates nothing; C > 1024 returns DDPX_ERR_LIMIT with a
 * message on stderr.  The first call on a device looks the array's address up (make it at start-up, before any stream
 * capture on that device).  All calls share the array: each clears what it uses, and calls that overlap in time on one
 * GPU race.  This entry point has no counterpart in the reference. */
int ddpx_bn_train_fwd_static(const float* x, float* y, const float* gamma, const float* beta, float* running_mean,
                             float* running_var, long long* num_batches_tracked, float momentum, float eps, int Nimg,
                             int H, int W, int C, int relu, int pool, float* mean_rstd, void* stream);
int ddpx_bn_bwd_reduce(const float* x, const float* dout, const float* mean, const float* rstd, const float* gamma,
                       const float* beta, int Nimg, int H, int W, int C, int relu, int pool, float* sum_dz, float* sum_dzx,
                       void* stream);
int ddpx_bn_bwd_apply(const float* x, const float* dout, const float* mean, const float* rstd, const float* gamma,
                      const float* beta, const float* sum_dz, const float* sum_dzx, int Nimg, int H, int W, int C, int relu,
                      int pool, int use_batch_stats, float* dx, void* stream);

/* ---- block-level entry points (csrc/kernels/blocks.cu): every kernel of a conv-BN-ReLU(-pool) group (M:53-59) and
 * direction in one call; all workspaces are the caller's (see the comments in blocks.cu for each argument) ---------- */
int ddpx_conv_bn_fwd(const float* x, const float* wf, const float* gamma, const float* beta, float* running_mean,
                     float* running_var, long long* num_batches_tracked, float momentum, float eps, int Nimg, int H, int W,
                     int Cin, int Cout, int relu, int pool, int training, float* stats, float* y, float* out,
                     float* mean_rstd, float* patches, void* stream);
int ddpx_conv_bn_bwd(const float* y, const float* dout, const float* xin, const float* wd, const float* mean_rstd,
                     const float* gamma, const float* beta, int Nimg, int H, int W, int Cin, int Cout, int relu, int pool,
                     int training, float* sums, float* dconv, float* staging, float* dw, float* dx, void* stream);

/* ---- a whole GPT-2 block per direction in one call (csrc/kernels/blocks_lm.cu; bf16, 64-wide heads).  params[12] =
 * ln_1.{weight,bias}, attn.c_attn.{weight,bias}, attn.c_proj.{weight,bias}, ln_2.{weight,bias}, mlp.c_fc.{weight,bias},
 * mlp.c_proj.{weight,bias}; acts[7] = LN1 out, qkv, attention out, x1, LN2 out, fc pre-activation, fc activation (kept
 * for the backward); stats = fp32 [4*B*T + B*H*Tp]; tmp = bf16 [B*T, C]; backward scratch: wsb = bf16 [B*T, 9C],
 * ws32 = fp32 [4C*C + 13C + B*H*Tp + ddpx_layernorm_bwd_scratch_floats(B*T, C)] (first 4C*C all zero on entry, left all
 * zero), cnt = uint32 [16*(4C*C/8192 + 64)] all
 * zero on entry and exit.  Nothing is allocated.  No counterpart in the reference (a CNN). */
int ddpx_gpt2_block_fwd(const void* x, void* const* params, void* const* acts, float* stats, void* tmp, void* out, int B,
                        int T, int C, int H, float eps, void* stream);
int ddpx_gpt2_block_bwd(const void* x, const void* dout, void* const* params, void* const* acts, const float* stats,
                        void* const* grads, void* dx, void* wsb, float* ws32, unsigned int* cnt, int B, int T, int C, int H,
                        void* stream);

/* ---- fused causal attention, bf16, head dimension 64 (csrc/kernels/attn_sm100.cu, attn_bwd_sm100.cu) ---------------
 * qkv [B][T][3*H*64] (q | k | v, heads contiguous inside each), out / dout [B][T][H*64], lse2 / delta [B][H][Tp] fp32 with
 * Tp = T rounded up to 128 (lse2 = base-2 log-sum-exp of the scaled scores; delta is scratch of the backward).
 * All three kernels are persistent: one launch of at most 148 (forward: 296) CTAs walks every (batch, head, tile) item.
 * No counterpart in the reference (a CNN); serves the GPT-2 configurations of BASELINE.json. */
int ddpx_attn_fwd(const void* qkv, void* out, float* lse2, int B, int T, int H, float scale, void* stream);
int ddpx_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse2, float* delta, void* dqkv, int B,
                  int T, int H, float scale, void* stream);
unsigned int ddpx_attn_read_error_word(void);
unsigned int ddpx_attn_bwd_read_error_word(void);

/* ---- losses and optimizers (csrc/kernels/loss_optim.cu) ------------------------------------ */
/* softmax cross-entropy fwd+bwd fused (M:95): loss_sum[0] = mean NLL, dlogits = (softmax - onehot) * grad_scale / rows */
int ddpx_softmax_ce(const void* logits, const long long* targets, void* dlogits, float* loss_sum, long long rows, int cols,
                    long long ld, int bf16, long long ignore_index, float grad_scale, void* stream);
/* two-kernel form: forward reads the logits once and keeps lse[rows]; backward writes dlogits = (softmax - onehot) * g / rows
 * with the upstream scalar gradient g read from device memory (NULL = 1); columns [cols, ld) get zero */
int ddpx_ce_fwd(const void* logits, const long long* targets, float* lse, float* loss_sum, long long rows, int cols,
                long long ld, int bf16, long long ignore_index, void* stream);
int ddpx_ce_bwd(const void* logits, const long long* targets, const float* lse, const float* g, void* dlogits, long long rows,
                int cols, long long ld, int bf16, long long ignore_index, void* stream);
int ddpx_scale_by_device_scalar(void* p, long long n, const float* g, int bf16, void* stream);
int ddpx_mse(const float* pred, const float* target, float* dpred, float* loss_sum, long long n, void* stream);
/* one launch for all tensors; torch SGD / AdamW semantics (M:97, torch/optim/sgd.py:346-380) */
/* dst[i][:] = src[i][:] * scale for n tensors in one launch per 96 tensors (gradient reducer: gradients -> bucket slots). */
int ddpx_scale_copy_multi(int n, void* const* dst, void* const* src, const long long* numel, float scale, int is_bf16,
                          void* stream);
int ddpx_sgd_multi(int n, void* const* params, void* const* grads, void* const* momentum_bufs, const long long* numel,
                   float lr, float momentum, float dampening, float weight_decay, int nesterov, int first_step,
                   float grad_scale, int is_bf16, void* stream);
int ddpx_adamw_multi(int n, void* const* params, void* const* grads, void* const* exp_avg, void* const* exp_avg_sq,
                     const long long* numel, float lr, float beta1, float beta2, float eps, float weight_decay,
                     long long step, float grad_scale, int is_bf16, void* stream);

/* ---- transformer kernels (csrc/kernels/transformer.cu, attention_aux.cu) -----------------------
 * No counterpart in the reference (a CNN); required by the MLP / GPT-2 configs of BASELINE.json.
 * is_bf16 selects the storage type of the activations/parameters (math is fp32). */
int ddpx_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                       long long rows, int C, float eps, int is_bf16, void* stream);
int ddpx_layernorm_bwd(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd, void* dx,
                       float* dgamma, float* dbeta, long long rows, int C, int is_bf16, void* stream);
/* the same with caller-provided scratch (fp32, ddpx_layernorm_bwd_scratch_floats(rows, C) elements): blocks store partial sums
 * there and a second small kernel adds them up and overwrites dgamma / dbeta (which then need not be zeroed) */
int ddpx_layernorm_bwd_ws(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd, void* dx,
                          float* dgamma, float* dbeta, long long rows, int C, int is_bf16, float* scratch, void* stream);
long long ddpx_layernorm_bwd_scratch_floats(long long rows, int C);
/* ... and with `dres` (optional; shape and dtype of dx, no aliasing) added to dx: the residual branch's gradient */
int ddpx_layernorm_bwd_res(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd, void* dx,
                           float* dgamma, float* dbeta, long long rows, int C, int is_bf16, float* scratch, const void* dres,
                           void* stream);
int ddpx_gelu_fwd(const void* x, void* y, long long n, int is_bf16, void* stream);
int ddpx_gelu_bwd(const void* x, const void* dy, void* dx, long long n, int is_bf16, void* stream);
int ddpx_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long n, void* stream);
int ddpx_add(const void* a, const void* b, void* out, long long n, int is_bf16, void* stream);
int ddpx_cast_scale(const void* src, void* dst, long long n, float scale, int kind, void* stream);
int ddpx_colsum_bf16(const void* x, float* out, long long rows, int cols, void* stream);
/* out[i][c] += column sums of n <= 4 bf16 matrices [rows][cols[i]] in one launch; the fp32 outputs are NOT zeroed here */
int ddpx_colsum_bf16_multi(int n, const void* const* x, float* const* out, long long rows, const int* cols, void* stream);
int ddpx_dropout(const void* x, void* y, long long n, float p, unsigned long long seed, int is_bf16, void* stream);
int ddpx_embedding_fwd(const long long* idx, const void* wte, const void* wpe, void* out, long long BT, int T, int C,
                       int is_bf16, void* stream);
int ddpx_embedding_bwd(const long long* idx, const void* dy, float* dwte, float* dwpe, long long BT, int T, int C,
                       int is_bf16, void* stream);
int ddpx_causal_softmax_fwd(void* s, long long rows, int T, int is_bf16, void* stream);
int ddpx_causal_softmax_bwd(const void* p, void* dp, long long rows, int T, float scale, int is_bf16, void* stream);

/* ---- peer-to-peer NVLink backend (csrc/runtime/p2p.cu) ------------------------------------------
 * Stands where ProcessGroupNCCL stands for the reference (M:32, M:89).
 * ddpx_p2p_malloc / export / import: symmetric buffers through CUDA IPC (these DO allocate / map memory).
 * ddpx_p2p_allreduce_fused: one kernel per gradient bucket = all-reduce over peer memory (+ SGD / AdamW update of the
 *   parameters the bucket covers).  bufs[r] / pads[r] = rank r's bucket / signal pad mapped into this process;
 *   gen = monotonically increasing collective number (identical on all ranks); mode 0 two-shot, 1 one-shot;
 *   opt 0 none, 1 SGD, 2 AdamW; hyper = [lr, momentum, dampening, wd, beta1, beta2, eps, bc1, sqrt(bc2)].
 * ddpx_p2p_broadcast: pull-model broadcast of a chunk of tensors through the source's staging buffer. */
int ddpx_p2p_malloc(void** out, long long bytes);
int ddpx_p2p_free(void* p);
int ddpx_p2p_export(void* p, unsigned char* handle64);
int ddpx_p2p_import(const unsigned char* handle64, void** out);
int ddpx_p2p_unimport(void* p);
int ddpx_p2p_can_access(int dev, int peer);
unsigned int ddpx_p2p_read_error(void* pad);
/* how long a device-side flag wait may last on the current device before the rank traps (seconds; default 600) */
int ddpx_p2p_set_timeout(double seconds);
/* collective sequence numbers shared by every launcher in the process (same order on every rank => same numbers) */
unsigned int ddpx_p2p_next_gen(void);
unsigned int ddpx_p2p_next_bcast_gen(void);
void ddpx_p2p_reset_gens(void);
int ddpx_p2p_allreduce_fused(int rank, int world, void* const* bufs, void* const* pads, unsigned int gen, long long n,
                             int is_bf16, int mode, int opt, const float* hyper, int nesterov, int first_step, int nslots,
                             void* const* params, void* const* state0, void* const* state1, const long long* offsets,
                             const long long* numels, int ctas, void* stream);
/* NVLS flavour: the shard reduction runs in the NVSwitch (multimem.ld_reduce / multimem.st on mc_buf, the multicast
 * mapping of the same symmetric buffer); everything else as the two-shot mode of ddpx_p2p_allreduce_fused. */
int ddpx_p2p_allreduce_nvls(int rank, int world, void* const* bufs, void* mc_buf, void* const* pads, unsigned int gen,
                            long long n, int is_bf16, int opt, const float* hyper, int nesterov, int first_step, int nslots,
                            void* const* params, void* const* state0, void* const* state1, const long long* offsets,
                            const long long* numels, int ctas, void* stream);
/* prev_gen: the generation in which `src` last filled the same staging slot as a source (0 = never); the source waits for
 * every peer's acknowledgement of that generation before it overwrites the slot */
int ddpx_p2p_broadcast(int rank, int world, void* const* pads, int src, unsigned int gen, unsigned int prev_gen,
                       void* staging_of_src, int n, void* const* ptrs, const long long* offs, const long long* bytes,
                       int ctas, void* stream);

/* ---- reducer core (csrc/runtime/reducer_core.cpp): host-side bucket assignment + ready/launch state machine ----
 * Counterpart of c10d::compute_bucket_assignment_by_size and the bookkeeping of c10d::Reducer that torch DDP builds
 * for the reference at M:89. */
/* ---- native reducer (csrc/runtime/reducer.cu): per-gradient / per-bucket work of a synchronised backward without Python:
 * ready counting, strict bucket order, gradient gather (x 1/world), event hand-over to the communication stream, launch of
 * the fused all-reduce(+optimizer) kernel, completion events, per-bucket device timers.  The hot half of c10d::Reducer
 * (M:89); driven by the C++ autograd hooks of csrc/torch_ext/autograd_hooks.cpp. */
void* ddpx_nreducer_create(int n_params, int n_buckets, const int* bucket_of_param, int rank, int world, void* const* pads,
                           void* comm_stream, int ctas);
void ddpx_nreducer_destroy(void* h);
int ddpx_nreducer_set_bucket(void* h, int b, int n, const int* pids, const long long* offsets, const long long* numels,
                             void* const* bufs, void* mc, long long numel, int is_bf16);
int ddpx_nreducer_set_slots(void* h, int b, int n, void* const* params, void* const* state0, void* const* state1);
int ddpx_nreducer_set_step(void* h, int b, int mode, int opt, const float* hyper9, int nesterov, int first_step);
void ddpx_nreducer_arm(void* h, int timing);
/* returns flags: 1 a bucket launched, 2 all buckets launched, 4 the iteration's first bucket launched; < 0 error */
int ddpx_nreducer_mark_ready(void* h, int pid, const void* grad_ptr, void* compute_stream);
int ddpx_nreducer_wait(void* h, void* compute_stream);
int ddpx_nreducer_bucket_times(void* h, float* out_ms, int* out_mode, int max_out);
int ddpx_nreducer_ready_order(void* h, int* out, int max_out);
int ddpx_nreducer_unmarked(void* h, int* out, int max_out);
int ddpx_nreducer_launched(void* h);

/* ---- control-plane store (csrc/runtime/store.cpp): TCP key-value store hosted by rank 0; the bootstrap exchanges CUDA-IPC
 * handles, bucket orders and barriers through it.  Status < 0: -1 missing key / timeout, -2 connection, <= -16 buffer too
 * small (-16 - needed).  Stands where c10d::TCPStore stands for the reference (multigpu.py:30-32). */
void* ddpx_store_server_start(int port, int* bound_port);       /* port 0 = pick one; listens on every interface */
void* ddpx_store_server_start_on(const char* bind_host, int port, int* bound_port);   /* listen on one IPv4 address only */
void ddpx_store_server_stop(void* server);
void* ddpx_store_connect(const char* host, int port, int timeout_ms);
void ddpx_store_close(void* client);
long long ddpx_store_set(void* client, const char* key, const void* data, long long len);
long long ddpx_store_get(void* client, const char* key, void* buf, long long cap, int timeout_ms);
long long ddpx_store_add(void* client, const char* key, long long delta);
long long ddpx_store_delete(void* client, const char* key);
long long ddpx_store_num_keys(void* client);

int ddpx_bucket_assign(int n, const long long* nbytes, const int* keys, int nlimits, const long long* limits,
                       int sort_by_min_position, int* out_bucket_of_position, long long* out_bucket_limit,
                       int* out_num_buckets);
void* ddpx_reducer_create(int n_params, int n_buckets, const int* bucket_of_param);
void ddpx_reducer_destroy(void* h);
void ddpx_reducer_reset(void* h);
int ddpx_reducer_mark_ready(void* h, int param, int* out_buckets, int max_out);
int ddpx_reducer_all_launched(void* h);
int ddpx_reducer_num_marked(void* h);
int ddpx_reducer_ready_order(void* h, int* out, int max_out);
int ddpx_reducer_unmarked(void* h, int* out, int max_out);

/* Last device-side error code written by a bounded wait that expired (0 = none). Synchronises. */
unsigned int ddpx_read_error_word(void);

#ifdef __cplusplus
}
#endif
#endif /* DDPX_KERNELS_H_ */
