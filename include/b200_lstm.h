/*
 * b200_lstm.h -- C ABI of libb200lstm.so: the per-step training loop of the
 * MarkoMiljkovic45/masters-thesis trainer as hand-written sm_100a CUDA kernels.
 *
 * The reference is pure Python over PyTorch/Lightning and has no FFI of its own; each
 * entry point below replaces the LIBRARY CALL the reference makes at the cited
 * file:line (paths relative to the reference checkout).  A binding a maintainer of
 * the reference would add is shown in INTEGRATION.md (ctypes, ~40 lines).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host
 *   - every buffer, including workspace, is owned and allocated by the caller
 *     (query sizes with the *_bytes functions); the library never allocates
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = the
 *     legacy default stream); no call synchronises the device
 *   - return value: 0 on success, negative on error (-1 invalid argument,
 *     -2 unsupported configuration, -1000-cudaError for a CUDA runtime error);
 *     b200_last_error() returns a thread-local message for the last failure
 *   - fp32 everywhere in memory; int64 for row counts
 *
 * Memory layouts (N = n_windows * n_stocks sequences, T = lookback, n = target days,
 * F = input features, H = hidden size, L = layers)
 *   x          [T][N][F]     time-major look-back features [r_s, r_m, r_s*r_m(, r_s^2, r_m^2)]
 *   y_stock    [N][n]        target-window stock returns
 *   y_market   [n_windows][n]
 *   factor     [n_windows][2]  (mean, unbiased variance) of the target-window market return
 *   inv_psi    [N]           1 / unbiased residual variance of the per-window OLS
 *   params / grads / adam m,v : one flat fp32 vector in torch state_dict order
 *        for l in 0..L-1: weight_ih_l (4H x I_l), weight_hh_l (4H x H), bias_ih_l (4H), bias_hh_l (4H)
 *        alpha_head.weight (H), alpha_head.bias (1), beta_head.weight (H), beta_head.bias (1)
 *        with I_0 = F, I_l = H; gate row order i,f,g,o (torch.nn.LSTM)
 *   stash_h, stash_c  [L][T][N][H]  hidden and cell state of every layer and step
 *   hT         = stash_h + ((L-1)*T + (T-1)) * N * H      (final hidden of the top layer)
 *        The tcgen05 engines at hidden 64 (B200_MODE_TC_*) write ONLY hT into stash_h: they keep the hidden states of all
 *        steps inside `workspace` as bf16 hi/lo operand images that later kernels fetch with bulk copies (csrc/lstm_tc.cu);
 *        b200_lstm_hidden_states() returns any layer's states as fp32 for every engine.  stash_c is always fp32.
 *        The wide engine in bf16 mode (hidden 128 / 256 with B200_MODE_TC_FAST, csrc/lstm_wide2.cu) also writes ONLY hT into
 *        stash_h and does not touch stash_c: hidden states, activated gates (bf16 operand images) and the fp32 cell state
 *        live in `workspace`.
 */
#ifndef B200_LSTM_H
#define B200_LSTM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#define B200_OK 0
#define B200_ERR_INVALID (-1)
#define B200_ERR_UNSUPPORTED (-2)

/* LSTM engine selector (`mode` arguments) */
#define B200_MODE_FFMA 0      /* fp32 CUDA-core path: exact-fp32 reference engine */
#define B200_MODE_TC_PRECISE 1 /* tcgen05 bf16x3 split (fp32-equivalent products), fp32 accumulate in TMEM */
#define B200_MODE_TC_FAST 2    /* tcgen05 single bf16 product, fp32 accumulate */

/* loss selector, configs/loss/{mse,nll,combined}.yaml */
#define B200_LOSS_MSE 0
#define B200_LOSS_NLL 1
#define B200_LOSS_COMBINED 2

B200_API int b200_version(void);
B200_API const char* b200_last_error(void);
/* hash of the sources and flags this binary was built from (masters-thesis_b200/build.py source_hash) */
B200_API const char* b200_build_hash(void);
/* 1 when the LSTM engine `mode` is compiled into this library, else 0 */
B200_API int b200_mode_available(int mode);

/* Per-kernel timing for bench.py's roofline leg (no counterpart in the reference).  While enabled,
 * every kernel launch of the LSTM/heads/optimiser paths is bracketed by CUDA events on the launching
 * stream.  b200_timing_report synchronises those events and writes one line per kernel:
 * "<kernel name> <launches> <total milliseconds>\n". */
B200_API int b200_timing_enable(int on);
B200_API int b200_timing_report(char* buf_host, size_t buf_bytes);

/* Tuning knobs for experiments (no counterpart in the reference); process-wide, read at launch time, defaults are
 * the measured best.  The library reads no environment variables.
 *   "tc_prefetch"  bit mask of L2 prefetches in the tcgen05 kernels (csrc/lstm_tc.cu, default 39)
 *   "tc_fwd_eh"    bit 0 / bit 1: two epilogue threads per row in the forward kernel of layers > 0 / layer 0 (default 1)
 *   "tc_wgrad"     products of the weight gradient in layers > 0, precise mode: 3 (default), 2, or 0 = two from 2^22 row-steps on
 *   "tc_dgrad"     products of the data gradient, precise mode: 3 (default) or 2
 *   "wide_fuse"    1 = LSTM cell in the epilogue of the wide engine's recurrent GEMM, fp32-storage path (default 0)
 *   "wide_images"  0 = run bf16 mode of the wide engine on the fp32-storage path of csrc/lstm_wide.cu instead of the
 *                  operand-image engine of csrc/lstm_wide2.cu (default 1); changes b200_lstm_workspace_bytes
 *   "wide_mc"      1 = forward kernel of that engine as two-CTA clusters whose CTAs multicast the activation chunks to each other
 *                  (default 0; used only when the workspace is device memory of the GPU the launch runs on)
 *   "wide_dbg"     timing ablations of that engine (bit mask, results are wrong when set; bit 32 = event trace) */
B200_API int b200_set_option(const char* name, int value);

/* number of fp32 parameters of the model (src/model.py:88-97): 51,074 for F=3,H=64,L=2 */
B200_API int64_t b200_param_count(int n_features, int hidden, int layers);
/* offset (in floats) of layer l's weight_ih inside the flat vector; l == layers gives the heads */
B200_API int64_t b200_param_offset(int n_features, int hidden, int layers, int l);

/* ---------------------------------------------------------------- batch generation
 * Replaces SyntheticLogReturns.generate (src/data.py:49-59, torch.distributions on the host),
 * lookback_target_split / add_quadratic_features / ols_features (src/common.py:81-149) and the
 * DataLoader collate + H2D copy (src/data.py:236-238): window w covers absolute days
 * [w*stride, w*stride + lookback + target) of one Philox4x32-10 addressed series (generator
 * specification: oracle/lstm_oracle.py "generator spec").  window_idx (int64, device) selects
 * arbitrary windows; when NULL, windows first_window .. first_window+n_windows-1 are produced. */
B200_API int b200_gen_batch(uint64_t seed, const int64_t* window_idx, int64_t first_window, int64_t n_windows,
                   int n_stocks, int lookback, int target, int64_t stride, int n_features,
                   float* x, float* y_stock, float* y_market, float* factor, float* inv_psi,
                   void* stream);
/* The same batch from a RESIDENT series instead of the generator: stocks [n_stocks][n_samples], market [n_samples]
 * (device, fp32) are what the reference's prepare_data loads from data_dir/stocks.pt and market.pt
 * (src/data.py:193-196; datamodule=real: 25 portfolios, stride 5).  prediction_task = 0 selects the reconstruction
 * split of lookback_target_split (src/common.py:107-110): the window is `lookback` days long and the target is its
 * last `target` days.  ols_alpha / ols_beta [n_windows*n_stocks] (may be NULL) receive the per-window OLS
 * coefficients that the reference stores in y channels 2, 3 when no true coefficients exist (src/data.py:205-207). */
B200_API int b200_series_batch(const float* stocks, const float* market, int64_t n_samples, const int64_t* window_idx,
                      int64_t first_window, int64_t n_windows, int n_stocks, int lookback, int target,
                      int64_t stride, int n_features, int prediction_task, float* x, float* y_stock,
                      float* y_market, float* factor, float* inv_psi, float* ols_alpha, float* ols_beta,
                      void* stream);
/* A batch in the reference's layout, already on the device (X [B][S][T][F], y [B][S][n][4]: what its DataLoader yields,
 * src/data.py:216, 236-238) -> the kernels' layout (x [T][N][F], y_stock [N][n] = channel 0, y_market [B][n] = channel 1):
 * replaces the flatten(0, 1) / channel slicing of _shared_step (src/model.py:192-196) for host-fed batches. */
B200_API int b200_collate_reference(const float* X, const float* y, int64_t n_windows, int n_stocks, int lookback, int target,
                           int n_features, float* x, float* y_stock, float* y_market, void* stream);
/* days [first_day, first_day + n_days) of the generator's series itself: stocks [n_stocks][n_days], market [n_days]
 * (what SyntheticLogReturns.generate returns, src/data.py:49-59) */
B200_API int b200_gen_series(uint64_t seed, int64_t first_day, int64_t n_days, int n_stocks, float* stocks, float* market,
                    void* stream);
/* per-stock true alpha_i, beta_i of the run (src/data.py:53-54) */
B200_API int b200_gen_coefficients(uint64_t seed, int n_stocks, float* alpha, float* beta, void* stream);
/* raw Philox4x32-10 blocks, for bit-exact checks against the numpy specification:
 * counters [n][4] uint32 -> out [n][4] uint32 */
B200_API int b200_philox4x32(uint64_t seed, const uint32_t* counters, int64_t n, uint32_t* out, void* stream);
/* inter-layer dropout keep-mask exactly as the LSTM kernels regenerate it: out [T][N][H] floats,
 * 0 or 1/(1-p), for the INPUT of layer `layer` (1..L-1).  Replaces aten::bernoulli_ inside
 * nn.LSTM(dropout=p) (src/model.py:92). */
B200_API int b200_dropout_mask(uint64_t seed, int layer, int64_t n_seq, int T, int hidden, float p,
                      float* out, void* stream);

/* ---------------------------------------------------------------- LSTM encoder
 * b200_lstm_fwd replaces self.lstm(lstm_input) (src/model.py:104, nn.LSTM -> oneDNN
 * mkldnn_rnn_layer / cuDNN _cudnn_rnn); b200_lstm_bwd_layer replaces its autograd backward
 * (mkldnn_rnn_layer_backward).  One launch sequence per layer; layers run bottom-up in fwd and
 * must be called top-down (layer = L-1 .. 0) in bwd so that a data-parallel caller can start
 * reducing a layer's gradients while the layers below are still running.
 *
 * training != 0 applies inter-layer dropout p with the Philox mask of b200_dropout_mask.
 * workspace: at least b200_lstm_workspace_bytes(...) bytes, shared by fwd and bwd calls of a step.  In the tcgen05 modes
 * the forward pass leaves state in it that the backward pass of the same step reads (hidden-state operand images, dropout
 * keep bits): pass the SAME pointer and size, untouched in between, with the same dropout_p / dropout_seed / training.
 * A backward launch that finds keep bits of another step (seed, layer or size mismatch) aborts (CUDA launch failure); so does a
 * backward launch of the bf16 wide engine whose workspace was not filled by a forward pass with the same N, T, layers, hidden size,
 * dropout and seed.  The workspace must be memory of the GPU the kernels run on. */
B200_API size_t b200_lstm_workspace_bytes(int mode, int64_t n_seq, int T, int n_features, int hidden, int layers);
B200_API int b200_lstm_fwd(int mode, const float* params, int n_features, int hidden, int layers,
                  const float* x, int64_t n_seq, int T, float* stash_h, float* stash_c,
                  float dropout_p, uint64_t dropout_seed, int training,
                  void* workspace, size_t workspace_bytes, void* stream);
/* dhT: [N][H] gradient of the loss w.r.t. the top layer's final hidden state (from the heads
 * kernels).  grads: flat gradient vector; the call OVERWRITES the slice of layer `layer`. */
B200_API int b200_lstm_bwd_layer(int mode, int layer, const float* params, int n_features, int hidden, int layers,
                        const float* x, int64_t n_seq, int T, const float* stash_h, const float* stash_c,
                        const float* dhT, float* grads,
                        float dropout_p, uint64_t dropout_seed, int training,
                        void* workspace, size_t workspace_bytes, void* stream);
/* all layers, top-down */
B200_API int b200_lstm_bwd(int mode, const float* params, int n_features, int hidden, int layers,
                  const float* x, int64_t n_seq, int T, const float* stash_h, const float* stash_c,
                  const float* dhT, float* grads,
                  float dropout_p, uint64_t dropout_seed, int training,
                  void* workspace, size_t workspace_bytes, void* stream);

/* Hidden states of one layer as fp32 out[T][N][H] after b200_lstm_fwd (no counterpart in the reference: lstm_out of
 * src/model.py:104 for a lower layer): decoded from the workspace images in the tcgen05 modes, copied from stash_h otherwise. */
B200_API int b200_lstm_hidden_states(int mode, int layer, int n_features, int hidden, int layers, int64_t n_seq, int T,
                            const float* stash_h, const void* workspace, size_t workspace_bytes, float* out, void* stream);

/* Unit test hook of the tcgen05 engine (no counterpart in the reference): one GEMM with M = 128 through the
 * same UMMA descriptors, bf16x3 split, tcgen05.mma and TMEM load path as the LSTM kernels (fp32 in/out).
 *   variant 0: D[128][N] = A[128][K] * B[N][K]^T   (forward gate GEMM: both operands K-major)
 *   variant 1: D[128][N] = A[128][K] * B[K][N]     (data gradient: B image read MN-major)
 *   variant 2: D[128][N] = A[K][128]^T * B[K][N]   (weight gradient: A and B images read MN-major)
 * K, N multiples of 16, <= 256. */
B200_API int b200_tc_selftest(const float* A, const float* B, int variant, int K, int N, float* D, void* stream);
/* Microbenchmark hook (no counterpart in the reference): clocks per tcgen05.mma (M = 128, K = 16, bf16, N = n) when `iters` of
 * them are issued back to back, for the interleaved (layout 0) and the 128-byte swizzled (layout 1) operand layout and the
 * three operand orientations of b200_tc_selftest.  clk_per_mma_out: one float on the device. */
B200_API int b200_tc_mma_bench(int layout, int variant, int n, int iters, float* clk_per_mma_out, void* stream);

/* The GEMM of the wide engine (hidden != 64, csrc/lstm_wide.cu) on caller-supplied row-major fp32 device matrices
 * (no counterpart in the reference; what torch's addmm inside nn.LSTM does there): 128 x 256 output tiles, K chunks
 * of 64 double-buffered, operands split to bf16 hi(+lo) while staged, fp32 accumulation in TMEM.
 *   variant 0: C[M][N] = A[M][K] * B[N][K]^T;  variant 1: C = A[M][K] * B[K][N];  variant 2: C = A[K][M]^T * B[K][N]
 * nsplit 2 = three bf16 products (fp32-grade), 1 = plain bf16.  accumulate != 0 adds to C.  N % 64 == 0; K % 64 == 0
 * for variants 0, 1; M % 128 == 0 for variant 2, which splits K over CTAs when a workspace (bytes) is given. */
B200_API int b200_wide_gemm(int nsplit, int variant, const float* A, const float* B, float* C, int64_t M, int N, int64_t K,
                   int accumulate, void* workspace, size_t workspace_bytes, void* stream);

/* Debug hook (no counterpart in the reference): event trace of CTA 0 of the tcgen05 backward kernel.  The buffer
 * holds 4 roles x 8192 records of {tag, clock64}; NULL switches tracing off.  See tools/trace_bwd.py. */
B200_API int b200_debug_trace(void* device_buffer);

/* ---------------------------------------------------------------- heads + objective
 * Replaces alpha_head/beta_head (src/model.py:107-108, addmm), the decoder arithmetic of
 * _shared_step (src/model.py:192-202, 234-249, 286-306), torchmetrics MeanSquaredError,
 * inverse_returns_covariance (src/common.py:50-78) + MultivariateGaussianNLLLoss
 * (src/model.py:54-69: matmul/trace/logdet) and their autograd backward.  The NLL uses the
 * diagonal-plus-rank-1 closed form, O(N*n) instead of O(N^2).
 *
 * Rows are grouped into NLL groups of windows_per_group consecutive windows; a group is one joint
 * Gaussian over windows_per_group*n_stocks rows using the FIRST window's factor (the reference's
 * literal batch semantics, src/model.py:237-247).  Training: windows_per_group = n_windows.
 * Validation/test (batch 1, src/data.py:241-244): windows_per_group = 1.
 *
 * losses_out[4] = { mse (mean over all elements), nll (mean over groups),
 *                   total (mse | nll | nll + mse_weight*mse), mae (mean |pred - target|) }
 * dhT (nullable: forward only) [N][H]; head_grads (nullable) [2H+2] in flat-vector order;
 * alpha_out/beta_out (nullable) [N].  grad_scale multiplies every gradient (1/world_size under
 * data parallelism so that an NCCL sum yields the DDP mean). */
B200_API size_t b200_heads_workspace_bytes(int64_t n_windows, int n_stocks, int n_target, int hidden);
B200_API int b200_heads_loss_fwd_bwd(int loss_kind, float mse_weight, const float* head_params, int hidden,
                            const float* hT, const float* y_stock, const float* y_market,
                            const float* factor, const float* inv_psi,
                            int64_t n_windows, int n_stocks, int n_target, int64_t windows_per_group,
                            float grad_scale, float* alpha_out, float* beta_out, float* losses_out,
                            float* dhT, float* head_grads,
                            void* workspace, size_t workspace_bytes, void* stream);
/* the two names SURVEY.md section 8b lists; thin wrappers over the call above */
B200_API int b200_heads_mse_fwd_bwd(const float* head_params, int hidden, const float* hT,
                           const float* y_stock, const float* y_market,
                           int64_t n_windows, int n_stocks, int n_target, float grad_scale,
                           float* losses_out, float* dhT, float* head_grads,
                           void* workspace, size_t workspace_bytes, void* stream);
B200_API int b200_heads_nll_fwd_bwd(const float* head_params, int hidden, const float* hT,
                           const float* y_stock, const float* factor, const float* inv_psi,
                           int64_t n_windows, int n_stocks, int n_target, int64_t windows_per_group,
                           float grad_scale, float* losses_out, float* dhT, float* head_grads,
                           void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- gradient norm, clip, Adam
 * b200_grad_sqnorm replaces the linalg_vector_norm calls of clip_grad_norm_ (Lightning
 * gradient_clip_val, train.py:171): sqnorm_out[0] = sum g^2 (fp32, deterministic two-level
 * reduction).  workspace: b200_sqnorm_workspace_bytes(count). */
B200_API size_t b200_sqnorm_workspace_bytes(int64_t count);
B200_API int b200_grad_sqnorm(const float* grads, int64_t count, float* sqnorm_out,
                     void* workspace, size_t workspace_bytes, void* stream);
/* No counterpart in the reference.  Same result as b200_grad_sqnorm, for callers that have no way to
 * hand over a workspace: a binding that only holds the gradient pointer (the ctypes stub in
 * INTEGRATION.md owns no allocator), or a training/serving loop that has frozen its memory plan --
 * the framework's caching allocator already holds the device memory, so a fresh cudaMalloc may fail,
 * and it would stall every stream while it runs.  The partial sums go to an 8 KB array compiled into
 * the library (one fp64 slot per 4,096 gradients) instead of caller memory.
 *   - limit: count <= B200_SQNORM_STATIC_MAX (4,194,304 gradients, a 16 MB vector: every model of
 *     configs/model/ including `wide`); larger counts are rejected with B200_ERR_INVALID
 *   - every call looks up the array's device address for the CURRENT device (so it stays right after a
 *     cudaDeviceReset); the first call on a device does the one-time set-up behind that lookup: make it
 *     with the device current and before starting any stream capture.  Later calls only launch kernels.
 *   - nothing needs zeroing between calls (each call overwrites the slots it then reads), but all
 *     calls on a device share the one array: two calls whose kernels can overlap on the GPU
 *     (different streams, no ordering between them) race and return garbage.  Serialise them.
 *   - allocates nothing: no cudaMalloc/cudaFree, no library or torch allocation. */
#define B200_SQNORM_STATIC_MAX 4194304
B200_API int b200_grad_sqnorm_static(const float* grads, int64_t count, float* sqnorm_out, void* stream);

/* clip_grad_norm_ + Adam.step (train.py:171, src/model.py:150-154) in ONE launch for parameter vectors of at most
 * B200_CLIP_ADAM_FUSED_MAX floats (the small and medium reference models): every CTA reduces ||g||^2 over the whole (L2-resident)
 * vector in the same order, then updates its slice with the clip coefficient; sqnorm_out (nullable) receives ||g||^2.
 * clip_val <= 0 disables clipping.  Unlike b200_adam_step it leaves g untouched (every backward call overwrites its slice).
 * Larger vectors: B200_ERR_UNSUPPORTED -> use b200_grad_sqnorm + b200_adam_step. */
#define B200_CLIP_ADAM_FUSED_MAX 131072
B200_API int b200_clip_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t count, float lr,
                        float beta1, float beta2, float eps, float weight_decay, int64_t step, float clip_val,
                        float* sqnorm_out, void* stream);

/* dst[i] += scale * src[i]: accumulates the gradient of one micro-batch into the step's gradient when a batch is
 * processed in several forward/backward passes (one optimiser step per datamodule.batch_size windows,
 * src/data.py:236-238, src/model.py:204-209 -- the reference never splits a batch; this is what lets batches whose
 * activations exceed device memory keep that one-step-per-batch semantics). */
B200_API int b200_grad_accumulate(float* dst, const float* src, int64_t count, float scale, void* stream);

/* Replaces the clip scaling + torch.optim.Adam.step (src/model.py:150-154): L2-coupled weight
 * decay, bias correction, eps outside the sqrt(v_hat).  clip_val <= 0 or sqnorm == NULL disables
 * clipping; otherwise g *= min(1, clip_val / (sqrt(*sqnorm) + 1e-6)).  step is 1-based.
 * zero_grad != 0 clears g after use (Lightning's zero_grad). */
B200_API int b200_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t count,
                   float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                   float clip_val, const float* sqnorm, int zero_grad, void* stream);

/* ---------------------------------------------------------------- data-parallel gradient exchange
 * Replaces the gradient all-reduce of Lightning's implicit DDP (train.py:169: Trainer(strategy='auto') on a multi-GPU box
 * -> torch DistributedDataParallel -> one ncclAllReduce per gradient bucket).  One process per GPU; NCCL is called from
 * inside this library (bound at run time, the host framework's libnccl.so.2 when it is already loaded).
 *   rank 0: b200_dp_unique_id(id_host[128]) -> the caller distributes the 128 bytes to every rank (torchrun's store,
 *           a file, MPI ...) -> every rank, with ITS device current: b200_dp_init(id, world, rank, &comm)
 *   per bucket: b200_dp_allreduce(comm, grads + offset, count, comm_stream): in-place fp32 SUM over the ranks, enqueued
 *           on the given stream (the heads and backward kernels pre-scale gradients by grad_scale = 1/world, so the sum
 *           is the DDP mean).  Buckets are slices of the flat gradient vector; a slice may be reduced as soon as the
 *           b200_lstm_bwd_layer call that writes it has been enqueued and the communication stream waits on an event
 *           recorded after it, so the exchange of the upper layers overlaps the backward kernels of the lower ones.
 *   b200_dp_destroy(comm) at the end.  Errors: -2000 - ncclResult_t. */
B200_API int b200_dp_nccl_version(void);                 /* e.g. 22809; 0 when libnccl cannot be loaded */
B200_API int b200_dp_unique_id(void* id_out_host);       /* 128 bytes */
B200_API int b200_dp_init(const void* id_host, int world_size, int rank, void** comm_out);
B200_API int b200_dp_allreduce(void* comm, float* buf, int64_t count, void* stream);
B200_API int b200_dp_destroy(void* comm);

#ifdef __cplusplus
}
#endif
#endif /* B200_LSTM_H */
