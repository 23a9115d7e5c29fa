/*
 * explainn_b200.h -- C ABI of the B200-native ExplaiNN hot path.
 *
 * Every entry point takes raw device pointers, sizes and a CUDA stream, never
 * allocates device memory, never synchronises the stream and returns an int
 * status (0 = ok; otherwise explainn_last_error() holds a thread-local message).
 * Buffers are owned by the caller (PyTorch on the Python side).  No torch types
 * appear here; the Python host code (explainn_b200/_lib.py) binds these with
 * ctypes, and INTEGRATION.md shows the binding a maintainer of the reference
 * would add.
 *
 * The reference (oriolfornes/ExplaiNN) has no FFI: its hot path is a chain of
 * stock torch modules.  Each group below cites the reference lines it replaces
 * (paths relative to the reference checkout).
 *
 * Conventions
 *   B  sequences in the batch        U  cnn_units          k  kernel_size
 *   L  sequence_length               Lc = L-k+1 (conv out) n  = Lc/7 (pool out)
 *   F  n_features                    H  = 100 (hidden width of each unit's MLP)
 *   "planes": the packed one-hot layout, uint32 [B][3][Wp], Wp = ceil(L/32):
 *       plane 0 = low bit of the base code, plane 1 = high bit (A0 C1 G2 T3),
 *       plane 2 = N mask (1 = all-zero one-hot column); bit i%32 of word i/32
 *       is position i; bits at positions >= L have lo=hi=0, N=1.
 */
#ifndef EXPLAINN_B200_H
#define EXPLAINN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* xnn_stream_t; /* cudaStream_t */

#define XNN_OK 0
#define XNN_ERR_ARG 1
#define XNN_ERR_CUDA 2
#define XNN_ERR_UNSUPPORTED 3

#define XNN_HIDDEN 100 /* explainn/architectures/__init__.py:86 (100*cnn_units) */
#define XNN_POOL 7     /* explainn/architectures/__init__.py:81 MaxPool1d(7,7)  */

/* ---- library ---------------------------------------------------------- */
int xnn_version(void);
const char* xnn_last_error(void);
/* number of SMs of the current device (grids are sized in multiples of it) */
int xnn_sm_count(int* out);
/* leave `n` SMs out of every persistent grid launched afterwards (0 = use all): data-parallel training
 * reserves a few SMs so that the NCCL all-reduce of the MLP gradients (issued right after xnn_mlp_bwd) runs
 * under xnn_frontend_bwd instead of behind it.  Process-wide; no counterpart in the reference (single device,
 * explainn/selene/__init__.py:176-180). */
int xnn_set_sm_reserve(int n);
/* largest kernel_size: path 0 = eval forward / interpret / PWM scans on tcgen05 (48), path 1 = training
 * (BatchNorm batch statistics, backward) and `linears[:3]` activations (32).  The reference module has no
 * limit (explainn/architectures/__init__.py:44-45); explainn_b200.ExplaiNN validates against these. */
int xnn_kernel_size_limit(int path);

/* ---- a16: input contract, explainn/sequence/__init__.py:8-29,58-60 ------
 * one_hot_encode gives (4,L) rows A,C,G,T, any other letter an all-zero
 * column; rc_one_hot_encoding reverses both axes. */
/* float one-hot (B,C>=4,L) with element strides -> planes.  Channels 0..3 of
 * each sequence are read (so the reference's (B,4U,L) `repeat` tensor,
 * test.py:148, is accepted as is).  *bad_flag (device int, may be NULL) is set
 * to 1 if a column is neither one-hot nor all-zero. */
int xnn_pack_onehot_f32(const float* x, int64_t batch_stride, int64_t chan_stride,
                        int B, int L, uint32_t* planes, int* bad_flag, xnn_stream_t s);
/* ASCII bases (B,L), row stride in bytes -> planes (upper/lower case ACGT;
 * anything else = N), as sequence/__init__.py:12-27 */
int xnn_pack_ascii(const uint8_t* seqs, int64_t row_stride, int B, int L,
                   uint32_t* planes, xnn_stream_t s);
/* planes -> float one-hot (B,4,L) contiguous */
int xnn_unpack_onehot_f32(const uint32_t* planes, int B, int L, float* x, xnn_stream_t s);
/* reverse complement in the packed domain (sequence/__init__.py:58-60) */
int xnn_revcomp_planes(const uint32_t* planes, int B, int L, uint32_t* out, xnn_stream_t s);

/* ---- a3: conv + BatchNorm + exp + MaxPool1d(7,7) -------------------------
 * explainn/architectures/__init__.py:73-83 (K2-K6) and :109-111 (the repeat,
 * never materialised here). */
/* Train-mode BN1 statistics without a pass over the conv output: integer
 * co-occurrence counts of the one-hot im2col features.
 *   cnt1 int64 [4k]      : s[(c,j)]   = #{(b,p): base(b,p+j)=c}
 *   cnt2 int64 [4k][4k]  : S[(c,j),(c',j')] = #{(b,p): base(p+j)=c & base(p+j')=c'}
 * feature index = c*k + j (same order as the conv weight's (4,k) block).
 * Both buffers are cleared by the call (cudaMemsetAsync on the stream) before the kernel accumulates. */
int xnn_cooc_counts(const uint32_t* planes, int B, int L, int k,
                    long long* cnt1, long long* cnt2, xnn_stream_t s);
/* From counts + weights: batch mean/var of z (biased), the per-unit affine
 * y = scale*z_nobias + shift (conv bias folded), fp64 covariance products for
 * the backward pass, and the running-stat update of nn.BatchNorm1d
 * (momentum 0.1, unbiased var) when running_* != NULL.
 *   cov_w double [U][4k] = Cov . w_u      (saved for xnn_frontend_bwd)
 *   stats double [U][4]  = mean, var, invstd, mean without the conv bias */
int xnn_bn1_train_prepare(const long long* cnt1, const long long* cnt2, long long n_pos,
                          const float* conv_w, const float* conv_b,
                          const float* gamma, const float* beta, float eps, float momentum,
                          int U, int k,
                          float* running_mean, float* running_var,
                          float* scale, float* shift, double* stats, double* cov_w,
                          xnn_stream_t s);
/* Eval-mode fold of a BatchNorm (running stats) and an optional preceding bias
 * into y = scale*x_nobias + shift, for C channels. */
int xnn_bn_fold_eval(const float* gamma, const float* beta, const float* running_mean,
                     const float* running_var, const float* pre_bias /*may be NULL*/,
                     float eps, int C, float* scale, float* shift, xnn_stream_t s);
/* The fused front end.  z = sum_j W[u,base(p+j),j] (N contributes 0),
 * y = scale[u]*z + shift[u], a = exp(y), pool = max over 7 (first max wins).
 *   pool   float [B][U][n]  (required)
 *   zmax   float [B][U][n]  z at the arg-max (NULL in inference)
 *   argmax uint8 [B][U][n]  offset 0..6 of the arg-max (NULL in inference) */
size_t xnn_frontend_work_bytes(int U, int k);
int xnn_frontend_fwd(const uint32_t* planes, int B, int L, int U, int k,
                     const float* conv_w, const float* scale, const float* shift,
                     float* pool, float* zmax, uint8_t* argmax, void* work, xnn_stream_t s);
/* debugging cross-check only: the same contract on the FP32 pipe (shared-memory gather-sum) */
int xnn_frontend_fwd_lut(const uint32_t* planes, int B, int L, int U, int k,
                     const float* conv_w, const float* scale, const float* shift,
                     float* pool, float* zmax, uint8_t* argmax, xnn_stream_t s);
/* Per-position unit activations = linears[:3] (test.py:159-160):
 * act[b,u,p] = exp(scale*z+shift); either or both outputs may be given. */
int xnn_frontend_acts(const uint32_t* planes, int B, int L, int U, int k,
                      const float* conv_w, const float* scale, const float* shift,
                      float* act_f32, uint16_t* act_f16, xnn_stream_t s);

/* ---- a4/a5: per-unit MLP and final Linear ---------------------------------
 * explainn/architectures/__init__.py:84-104,114 (K7-K11). */
/* hpre[b,u,m] = b1[u,m] + sum_q W1[100u+m,q] * pool[b,u,q] */
int xnn_fc1_fwd(const float* pool, const float* w1, const float* b1,
                int B, int U, int n, float* hpre, xnn_stream_t s);
/* Batch statistics of x (B rows, C channels, row-major) for a train-mode
 * BatchNorm: mean/var (biased) -> scale/shift, stats[C][4] (mean,var,invstd,-),
 * running-stat update.  All `work` buffers in this header are caller-owned
 * scratch whose size in BYTES the matching *_work_bytes call returns. */
size_t xnn_bn_work_bytes(int B, int C);
int xnn_bn_train_prepare(const float* x, int B, int C,
                         const float* gamma, const float* beta, float eps, float momentum,
                         float* running_mean, float* running_var,
                         float* scale, float* shift, double* stats, void* work,
                         int phase, long long count, xnn_stream_t s);
/* Synchronised BatchNorm over data-parallel ranks (SURVEY.md 8e option (ii); the reference computes its statistics
 * over the whole single-process batch, architectures/__init__.py:79,90,99): the statistics / backward calls take
 *   phase 0  the whole call (count and out_scale ignored),
 *   phase 1  the partial sums only, left at the start of `work` (xnn_*_partial_bytes) for the host to all-reduce,
 *   phase 2  the rest, with count = values per channel over the GLOBAL batch; parameter gradients that are pure
 *            functions of the all-reduced sums are multiplied by out_scale (1/world) so that the gradient
 *            all-reduce(SUM) that follows restores them. */
size_t xnn_bn_partial_bytes(int B, int C);
/* opre[b,u] = b2[u] + sum_m W2[u,m]*drop(relu(scale2*hpre+shift2)); dropout is
 * a counter-based Bernoulli(1-p)/ (1-p) mask from (seed,b,u,m); p=0 disables.
 * hpre/b1 note: scale2/shift2 are per (u,m) channel (100U). */
int xnn_fc2_fwd(const float* hpre, const float* scale2, const float* shift2,
                const float* w2, const float* b2, int B, int U,
                float drop_p, unsigned long long seed, const uint8_t* mask /*optional injected (B,U,100)*/,
                float* opre, xnn_stream_t s);
/* o = relu(scale3*opre+shift3) (unit outputs, test.py:151);
 * logits = o . Wf^T + bf  (architectures/__init__.py:104,114) */
int xnn_head_fwd(const float* opre, const float* scale3, const float* shift3,
                 const float* wf, const float* bf, int B, int U, int F,
                 float* o, float* logits, xnn_stream_t s);

/* The same MLP on the tensor cores (tcgen05, kind::tf32 with a hi/lo split of both operands: fp32
 * round-off level), without ever writing the (B,100U) hidden activations: each pass recomputes its
 * FC1 tile in TMEM.  Supported for pool widths n <= 96 (xnn_mlp_supported). */
int xnn_mlp_supported(int n);
/* train-mode BatchNorm (linears.7) statistics of hpre = FC1(pool): scale2/shift2, stats2[100U][4],
 * running-stat update -- xnn_fc1_fwd + xnn_bn_train_prepare in one pass over pool. */
size_t xnn_mlp_bn2_work_bytes(int B, int U);
/* Tile images of a training step (optional, xnn_mlp_image_bytes, 128-byte aligned): this first pass over `pool`
 * writes every (unit, 128-row batch tile) operand stage it builds -- three bf16 pieces in the tensor-core operand
 * layout -- and the same for W1; xnn_mlp_fwd and xnn_mlp_bwd of the SAME step, given the same buffer, then fetch
 * their operand stages with one bulk copy (TMA) each instead of re-loading and re-splitting `pool`. */
size_t xnn_mlp_image_bytes(int B, int U, int n);
int xnn_mlp_bn2_train_prepare(const float* pool, const float* w1, const float* b1, int B, int U, int n,
                              const float* gamma, const float* beta, float eps, float momentum,
                              float* running_mean, float* running_var, float* scale, float* shift,
                              double* stats, void* work, void* image, int phase, long long count, xnn_stream_t s);
/* opre[b,u] = b2[u] + sum_m W2[u,m]*drop(relu(scale2*hpre+shift2)) -- xnn_fc1_fwd + xnn_fc2_fwd fused.
 * Dropout: injected byte mask (B,U,100), else Philox keep-bits from (seed,b,u) when drop_p > 0.
 * maskbits uint32 [B][U][4] (NULL in inference): bit m of a (b,u) row = hidden channel m is ACTIVE (kept by
 * dropout and BN2 output > 0).  xnn_mlp_bwd reuses exactly this decision, so the gradient is the one of
 * the forward that was computed even where a recomputed pre-activation rounds to the other side of 0.
 * seed_dev (optional): a device-resident draw counter; the effective seed is seed + *seed_dev * 0x9E3779B97F4A7C15,
 * so a captured CUDA graph draws a new mask on every replay (xnn_counters_advance moves the counter). */
int xnn_mlp_fwd(const float* pool, const float* w1, const float* b1, const float* scale2,
                const float* shift2, const float* w2, const float* b2, int B, int U, int n, float drop_p,
                unsigned long long seed, const unsigned long long* seed_dev, const uint8_t* mask,
                uint32_t* maskbits, float* opre, const void* image, xnn_stream_t s);

/* ---- a6: backward ---------------------------------------------------------
 * autograd of the above (selene/__init__.py:289-292). Gradients are shaped
 * exactly like the reference parameters. */
/* Standalone final Linear (architectures/__init__.py:104; test.py:155 calls it
 * on its own): y = x . W^T + b with x (B,U), W (F,U), b (F). */
int xnn_linear_fwd(const float* x, const float* w, const float* b, int B, int U, int F,
                   float* y, xnn_stream_t s);
/* dy (B,F) -> dx = dy . W (B,U), dW = dy^T . x (F,U), db = column sums (F). */
size_t xnn_linear_bwd_work_bytes(int B, int U, int F);
int xnn_linear_bwd(const float* dy, const float* w, const float* x, int B, int U, int F,
                   float* dx, float* dw, float* db, void* work, xnn_stream_t s);
/* Backward of the whole per-unit MLP on the tensor cores, from dopre = dL/d(opre) (B,U):
 * xnn_fc2_bwd + xnn_bn_bwd (linears.7) + xnn_fc1_bwd fused; hpre is recomputed from pool, neither it
 * nor its gradient is written.  maskbits: the active bits xnn_mlp_fwd wrote (required).
 *   dw1 (100U,n), db1 (100U), dgamma2/dbeta2 (100U), dw2 (U,100), db2 (U),
 *   dy[b,u,q] = pool * sum_m dhpre*W1 (gradient at the arg-max conv output through exp); with `image` given the
 *   factor pool is left out (dy = dL/dpool) and xnn_frontend_bwd applies it (act_scale / act_shift).
 * work: xnn_mlp_bwd_work_bytes, 256-byte aligned. */
size_t xnn_mlp_bwd_work_bytes(int B, int U, int n);
size_t xnn_mlp_bwd_partial_bytes(int B, int U);
int xnn_mlp_bwd(const float* dopre, const float* pool, const float* w1, const float* b1,
                const float* scale2, const float* shift2, const double* stats2, const float* gamma2,
                const float* w2, const uint32_t* maskbits, float drop_p, int B, int U, int n,
                float* dw1, float* db1, float* dgamma2, float* dbeta2, float* dw2, float* db2,
                float* dy, void* work, const void* image, int phase, long long count, float out_scale,
                xnn_stream_t s);
/* Train-mode BatchNorm backward over the batch dimension, C channels:
 *   given g = dL/d(out) (B,C), xpre (B,C), stats (mean,var,invstd), gamma; when
 *   relu_out != NULL the BatchNorm is followed by a ReLU whose output it is and
 *   g is first masked by relu_out > 0:
 *   dgamma, dbeta and dx = gamma*invstd*(g - dbeta/B - xhat*dgamma/B).
 * dx may alias g. work: xnn_bn_work_bytes(B,C) */
int xnn_bn_bwd(const float* g, const float* relu_out, const float* xpre, const double* stats,
               const float* gamma, int B, int C, float* dgamma, float* dbeta, float* dx, void* work,
               int phase, long long count, float out_scale, xnn_stream_t s);
/* FC2 backward + dropout/relu masks: from dopre (B,U) and hpre:
 *   dW2 (U,100), db2 (U), g2[b,u,m] = dopre*W2*mask*(bn2(hpre)>0) (B,U,100). */
size_t xnn_fc2_bwd_work_bytes(int B, int U);
int xnn_fc2_bwd(const float* dopre, const float* hpre, const float* scale2, const float* shift2,
                const float* w2, int B, int U, float drop_p, unsigned long long seed,
                const uint8_t* mask, float* dw2, float* db2, float* g2, void* work,
                xnn_stream_t s);
/* FC1 backward: dhpre (B,U,100), pool (B,U,n) ->
 *   dW1 (100U,n), db1 (100U), dy[b,u,q] = pool*sum_m dhpre*W1 (gradient at the
 *   arg-max conv output through exp). */
size_t xnn_fc1_bwd_work_bytes(int B, int U, int n);
int xnn_fc1_bwd(const float* dhpre, const float* pool, const float* w1,
                int B, int U, int n, float* dw1, float* db1, float* dy, void* work,
                xnn_stream_t s);
/* Front-end backward: sparse scatter of dy at the arg-max windows plus the
 * closed-form dense BatchNorm terms (Cov.w from xnn_bn1_train_prepare).
 *   dconv_w (U,4,k), dconv_b (U) (exactly 0 in train mode), dgamma1, dbeta1 (U) */
size_t xnn_frontend_bwd_work_bytes(int B, int U, int k);
/* act_scale / act_shift (optional, both or neither, (U) each): `dy` then holds dL/dpool and the kernel multiplies
 * every value by pool = exp(act_scale[u] * zmax + act_shift[u]) -- the forward's own expression, bit-identical to
 * the pooled activation it wrote -- before using it (xnn_mlp_bwd on tile images leaves that factor to this call). */
int xnn_frontend_bwd(const uint32_t* planes, int B, int L, int U, int k,
                     const float* dy, const float* zmax, const uint8_t* argmax,
                     const float* gamma, const double* stats, const double* cov_w,
                     const long long* cnt1, long long n_pos,
                     const float* act_scale, const float* act_shift,
                     float* dconv_w, float* dconv_b, float* dgamma, float* dbeta,
                     void* work, xnn_stream_t s);
/* debugging cross-check only: the same contract as a sparse FP64 scatter (no tensor cores) */
int xnn_frontend_bwd_simt(const uint32_t* planes, int B, int L, int U, int k,
                     const float* dy, const float* zmax, const uint8_t* argmax,
                     const float* gamma, const double* stats, const double* cov_w,
                     const long long* cnt1, long long n_pos,
                     float* dconv_w, float* dconv_b, float* dgamma, float* dbeta,
                     void* work, xnn_stream_t s);
/* Fused Adam over a flat fp32 buffer (torch.optim.Adam defaults,
 * architectures/__init__.py:463-464): step is the 1-based step count. */
int xnn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                  long long numel, float lr, float beta1, float beta2, float eps,
                  float weight_decay, long long step, float grad_scale, xnn_stream_t s);

/* The same update over a list of n_tensors parameter tensors in one launch
 * (optimizer.step() of selene/__init__.py:292 walks every parameter).  The five
 * arrays are HOST arrays of device pointers / element counts, read before the call
 * returns; all tensors share step (torch.optim.Adam steps them together).  Results
 * are bit-identical to n_tensors calls of xnn_adam_step. */
/* step_dev (optional): the step count read from DEVICE memory instead (a captured CUDA graph must not bake the
 * host's step into its bias corrections); `step` is then ignored. */
int xnn_adam_step_multi(int n_tensors, float* const* params, const float* const* grads,
                        float* const* exp_avg, float* const* exp_avg_sq, const long long* numels,
                        float lr, float beta1, float beta2, float eps, float weight_decay,
                        long long step, const long long* step_dev, float grad_scale, xnn_stream_t s);

/* ---- step-level pieces (no host scalars, no library kernels inside the training step) ----------------------
 * counters[i] += 1 for i < n: the device-resident Adam step and dropout draw counters. */
int xnn_counters_advance(long long* counters, int n, xnn_stream_t s);
/* *ptrs[i] += 1 for up to 8 counters in separate allocations (ptrs is a HOST array of device pointers): the
 * num_batches_tracked buffers of the three BatchNorm layers (architectures/__init__.py:79,90,99). */
int xnn_counters_advance_list(int n, long long* const* ptrs, xnn_stream_t s);
/* nn.BatchNorm1d's running-statistics update of up to 4 train-mode layers (+ num_batches_tracked += 1 where the
 * pointer is not NULL) in one launch, from the `stats` rows (mean, biased variance, invstd, -) the statistics
 * calls produced: running = (1-momentum)*running + momentum*(mean | var*count/(count-1))
 * (architectures/__init__.py:79,90,99).  All array arguments are HOST arrays of n_layers entries; count = values
 * per channel (B*L' for linears.1, B for linears.7 / .11; the global batch under synchronised statistics). */
int xnn_bn_running_update(int n_layers, const double* const* stats, float* const* running_mean,
                          float* const* running_var, long long* const* num_batches_tracked,
                          const int* channels, const float* momentum, const double* count, xnn_stream_t s);
/* explainn/architectures/__init__.py:446-456: nn.MSELoss() (kind 0) / nn.BCEWithLogitsLoss() (kind 1), 'mean'
 * reduction over the n = B*F elements; deterministic (fp64 partials added in block order).
 * xnn_loss_bwd: dlogits = grad_out[0] * d(loss)/d(logits); grad_out is a DEVICE scalar (NULL = 1). */
size_t xnn_loss_work_bytes(void);
int xnn_loss_fwd(const float* logits, const float* target, long long n, int kind, float* loss, void* work,
                 xnn_stream_t s);
int xnn_loss_bwd(const float* logits, const float* target, const float* grad_out, long long n, int kind,
                 float* dlogits, xnn_stream_t s);

/* ---- a7-a10: interpret ------------------------------------------------------
 * test.py:128-166 (fp16 activations), interpret.py:363-454. */
/* pass 1: per-unit maximum activation, rounded to fp16 like test.py:137, over
 * the sequences listed in idx (int32 [n_idx], NULL = all B):
 *   amax_bits uint32 [U] must be zero-initialised; holds the float bits of the
 *   fp16-rounded maximum (activations are >= 0 so uint order = float order). */
int xnn_act_max(const uint32_t* planes, int B, int L, int U, int k,
                const float* conv_w, const float* scale, const float* shift,
                const int32_t* idx, int n_idx, uint32_t* amax_bits, xnn_stream_t s);
/* pass 2a: per (unit, listed sequence) count of positions whose fp16-rounded
 * activation exceeds thr[u] (thr given as float holding an fp16 value).
 *   counts int32 [U][n_idx] */
int xnn_site_counts(const uint32_t* planes, int B, int L, int U, int k,
                    const float* conv_w, const float* scale, const float* shift,
                    const float* thr, const int32_t* idx, int n_idx,
                    int32_t* counts, xnn_stream_t s);
/* pass 2b: emit sites in the reference order (listed-sequence order, then
 * ascending position), honouring the per-unit cap (interpret.py:414-417):
 *   offsets int64 [U][n_idx] exclusive prefix of counts (+ carry of earlier
 *   strands/ranks); a site with global rank >= cap is dropped.
 *   pfm int32 [U][4][k] accumulated (zero first); N bases are not counted.
 *   sites int32 [U][site_cap][2] = (position in idx list, start) or NULL. */
int xnn_site_emit(const uint32_t* planes, int B, int L, int U, int k,
                  const float* conv_w, const float* scale, const float* shift,
                  const float* thr, const int32_t* idx, int n_idx,
                  const long long* offsets, long long cap,
                  int32_t* pfm, int32_t* sites, long long site_cap, xnn_stream_t s);

/* Passes 1 and 2a on the tensor cores (tcgen05): the Toeplitz GEMM of the conv
 * front end with a per-unit running maximum / site count in the epilogue.  The
 * fp16-rounded activation is monotone in the conv output, so the maximum is
 * taken before the exp and the threshold test is a compare against a per-unit
 * bisected conv value (exp evaluated only within 64 ulps of it).  Same arguments
 * and results as xnn_act_max / xnn_site_counts, plus a 16-byte aligned work
 * buffer of xnn_pwm_scan_tc_work_bytes(U, k) bytes; shapes must satisfy
 * xnn_interpret_tc_supported(L, k). */
int xnn_act_max_tc(const uint32_t* planes, int B, int L, int U, int k,
                   const float* conv_w, const float* scale, const float* shift,
                   const int32_t* idx, int n_idx, uint32_t* amax_bits,
                   void* work, xnn_stream_t s);
int xnn_site_counts_tc(const uint32_t* planes, int B, int L, int U, int k,
                       const float* conv_w, const float* scale, const float* shift,
                       const float* thr, const int32_t* idx, int n_idx,
                       int32_t* counts, void* work, xnn_stream_t s);
/* Pass 2b on the same kernel: identical predicate (so counts and emission
 * always agree), rank = offsets[u][i] + sites at earlier starts; PFM counted
 * with shared-memory atomics.  Arguments as xnn_site_emit plus `work`.
 * xnn_interpret_tc_supported(L, k): 1 when all three *_tc interpret passes fit
 * (use the three of them together, or none: the conv summation order differs
 * from the gather-add passes in the last bit). */
int xnn_interpret_tc_supported(int L, int k);
/* Between passes 2a and 2b (interpret.py:414-417): offsets[u][i] = carry[u] + sum of counts[u][0..i-1] (the rank
 * of the first site of unit u in listed sequence i), then carry[u] += sum of counts[u][:].  counts (U, n) int32
 * from xnn_site_counts[_tc]; carry (U) int64 in/out; offsets (U, n) int64 out. */
int xnn_site_offsets(const int32_t* counts, int U, int n, long long* carry, long long* offsets, xnn_stream_t s);
int xnn_site_emit_tc(const uint32_t* planes, int B, int L, int U, int k,
                     const float* conv_w, const float* scale, const float* shift,
                     const float* thr, const int32_t* idx, int n_idx,
                     const long long* offsets, long long cap,
                     int32_t* pfm, int32_t* sites, long long site_cap,
                     void* work, xnn_stream_t s);

/* ---- a11: PWM module, architectures/__init__.py:119-170 -------------------
 * score+[b,g,p] = sum_j W[g,base(p+j),j]; score- the same on the reverse
 * complement; out[b,g] = max or sum over both strands and all p (fp32).
 * mode: 0 = sum, 1 = max.  argmax int32 [B][G] (NULL ok): index on the
 * concatenated [fwd | rev] axis of torch.max(o, 2)[1]. */
int xnn_pwm_scan(const uint32_t* planes, int B, int L, int G, int w,
                 const float* pwm_w /* (G,4,w) */, int mode,
                 float* out, int32_t* argmax, xnn_stream_t s);

/* The same scan on the tensor cores (tcgen05): the G matrices are the A operand
 * of a Toeplitz GEMM over the bf16 one-hot of x, then of its reverse complement
 * (flip(x), architectures/__init__.py:387-404); weights are split in three bf16
 * pieces so every product is exact and accumulation is fp32.  Same arguments
 * and results as xnn_pwm_scan (scores to fp32 rounding, argmax with the same
 * first-maximum rule) plus a work buffer of xnn_pwm_scan_tc_work_bytes(G, w)
 * bytes, 16-byte aligned.  xnn_pwm_scan_tc_supported: 1 when (L, w) fit this
 * path (w <= 48 and the one-hot buffers fit in shared memory), else 0 -- use
 * xnn_pwm_scan then. */
int xnn_pwm_scan_tc_supported(int L, int w);
size_t xnn_pwm_scan_tc_work_bytes(int G, int w);
int xnn_pwm_scan_tc(const uint32_t* planes, int B, int L, int G, int w,
                    const float* pwm_w /* (G,4,w) */, int mode,
                    float* out, int32_t* argmax, void* work, xnn_stream_t s);
/* xnn_pwm_scan_tc without a work buffer and without any allocation (no cudaMalloc / cudaFree, nothing from an
 * allocating library): for callers under CUDA-graph capture, in tight serving loops or in harnesses that forbid
 * allocation.  NO COUNTERPART IN THE REFERENCE.  The scratch (the
 * bf16 pieces of the G matrices) is a file-scope device array of xnn_pwm_scan_tc_noalloc_max_bytes() = 1 MiB that
 * every GPU loading the library holds until exit; shapes with xnn_pwm_scan_tc_work_bytes(G, w) above it (e.g.
 * more than 896 matrices of width 48, or 1152 of width 35) are rejected with XNN_ERR_ARG and a message naming both
 * sizes.  Same arguments (minus `work`) and bit-identical results to xnn_pwm_scan_tc.  All calls on one device
 * share the array: calls that overlap on the GPU (different streams) race; calls on one stream are safe, each call
 * rewrites every byte it reads. */
size_t xnn_pwm_scan_tc_noalloc_max_bytes(void);
int xnn_pwm_scan_tc_noalloc(const uint32_t* planes, int B, int L, int G, int w,
                            const float* pwm_w /* (G,4,w) */, int mode,
                            float* out, int32_t* argmax, xnn_stream_t s);

/* ---- a13/a14: PWMScan pwm_scoring semantics, pwmscan/pwm_scoring.c ---------
 * Sequences are int8 codes 0..4 (A C G T N) as pwm_scoring.c:627-655, one
 * record per row of `offsets` (int64 [R+1], start of each record in codes).
 * LPM mode (pwm_scoring.c:340-475): lpm double [5][w] (row 4 = N), bg double
 * [5], or [R][5] when bg_per_record (the -q option, :355-372); forward_only as -f.
 *   sum mode : out_sum double [R]
 *   best mode: best double [R], best_pos int32 [R] (first best hit: start for
 *              '+', start+w for '-'), best_strand int8 [R] ('+'/'-'),
 *              n_ties int32 [R] (number of equal later hits, the comma list) */
int xnn_lpm_scan_sum(const int8_t* codes, const long long* offsets, int R, int w,
                     const double* lpm, const double* bg, int bg_per_record, int forward_only,
                     double* out_sum, xnn_stream_t s);
int xnn_lpm_scan_best(const int8_t* codes, const long long* offsets, int R, int w,
                      const double* lpm, const double* bg, int bg_per_record, int forward_only,
                      double* best, int32_t* best_pos, int8_t* best_strand, int32_t* n_ties,
                      int32_t* tie_pos /* [R][tie_cap] later tied hits, may be NULL */, int tie_cap,
                      xnn_stream_t s);
/* integer PWM mode (pwm_scoring.c:477-573): pwm int32 [5][w] (row 4 = INT_MIN,
 * two's-complement wrap preserved); best int32 [R] (INT_MIN and pos -1 when the
 * record is shorter than w), pos int32 [R], strand int8 [R] (0 '+', 1 '-'). */
int xnn_ipwm_scan_best(const int8_t* codes, const long long* offsets, int R, int w,
                       const int32_t* pwm, int forward_only,
                       int32_t* best, int32_t* pos, int8_t* strand, xnn_stream_t s);

/* The same two scans with every record's windows split into xnn_scan_parts(R)
 * contiguous parts, one block per (record, part) -- a chromosome is ONE record --
 * and the parts combined in ascending order by a second kernel (deterministic
 * sums; first-maximum rule and the ascending tie list kept).  work:
 * xnn_scan_work_bytes(R) bytes, 8-byte aligned.  Results as xnn_lpm_scan_sum (sum
 * order differs: 1e-12 relative), xnn_lpm_scan_best and xnn_ipwm_scan_best
 * (bit-exact). */
int xnn_scan_parts(int R);
size_t xnn_scan_work_bytes(int R);
int xnn_lpm_scan_sum_parts(const int8_t* codes, const long long* offsets, int R, int w,
                           const double* lpm, const double* bg, int bg_per_record, int forward_only,
                           double* out_sum, void* work, xnn_stream_t s);
int xnn_lpm_scan_best_parts(const int8_t* codes, const long long* offsets, int R, int w,
                            const double* lpm, const double* bg, int bg_per_record, int forward_only,
                            double* best, int32_t* best_pos, int8_t* best_strand, int32_t* n_ties,
                            int32_t* tie_pos, int tie_cap, void* work, xnn_stream_t s);
int xnn_ipwm_scan_best_parts(const int8_t* codes, const long long* offsets, int R, int w,
                             const int32_t* pwm, int forward_only,
                             int32_t* best, int32_t* pos, int8_t* strand, void* work, xnn_stream_t s);

/* ---- validation metrics on device-resident columns -----
 * Replace the host calls of architectures/__init__.py:458-461 (get_metrics:
 * sklearn roc_auc_score / average_precision_score, scipy pearsonr / spearmanr) as
 * used by selene/__init__.py:359-366 (one flattened column) and
 * pwm-scoring.py:115-120 (one call per PWM column).  Columns are contiguous
 * ([G][N]); the caller sorts them (any stable or unstable sort: ties are handled
 * by value) and passes sorted values plus the permutation (int64 indices into the
 * original order).
 *
 * xnn_metric_auc: sorted_scores DESCENDING; labels uint8 0/1 in ORIGINAL order,
 * column g at labels + g*label_stride (label_stride 0 = one label vector shared
 * by all columns); auroc/aupr double [G] (NaN where a column has one class only /
 * no positive); work = xnn_metric_auc_work_bytes(N, G) bytes.  N < 2^31.
 * xnn_metric_ranks: sorted_vals ASCENDING -> average 1-based ranks (scipy rankdata
 * "average") as double [G][N] in the original order.
 * xnn_metric_pearson_*: r double [G] of column g of x against column g of y
 * (y_stride 0 = shared y); fp64 two-pass; NaN for a constant column. */
/* Every column of vals (G, N) sorted on its own (ascending, or descending), stable, with the int64 permutation:
 * the sort the metric entry points below need (sklearn / scipy sort each column on the host).  One block per
 * column, 8-bit LSD radix sort; work = xnn_sort_columns_work_bytes(N, G) bytes (16 B per element). */
size_t xnn_sort_columns_work_bytes(long long N, int G);
int xnn_sort_columns(const float* vals, long long N, int G, int descending, float* sorted, long long* perm,
                     void* work, xnn_stream_t s);
size_t xnn_metric_auc_work_bytes(long long N, int G);
int xnn_metric_auc(const float* sorted_scores, const long long* perm, const uint8_t* labels,
                   long long label_stride, long long N, int G, double* auroc, double* aupr,
                   void* work, xnn_stream_t s);
int xnn_metric_ranks(const float* sorted_vals, const long long* perm, long long N, int G,
                     double* ranks, xnn_stream_t s);
int xnn_metric_pearson_f32(const float* x, const float* y, long long y_stride, long long N, int G,
                           double* r, xnn_stream_t s);
int xnn_metric_pearson_f64(const double* x, const double* y, long long y_stride, long long N, int G,
                           double* r, xnn_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* EXPLAINN_B200_H */
