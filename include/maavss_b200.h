/* libmaavss_b200.so -- C ABI of the B200-native MAAVSS training hot path (sm_100a only).
 *
 * Conventions (every entry point):
 *   - returns 0 on success, a negative MAAVSS_ERR_* code otherwise; the message is available
 *     from maavss_last_error() (thread-local);
 *   - all data pointers are DEVICE pointers owned by the caller (contiguous, fp32 unless stated),
 *     the library never allocates on the hot path (plans own their tables; scratch is passed in);
 *   - work is enqueued asynchronously on `stream` (a cudaStream_t passed as void*), no hidden syncs;
 *   - no CPU fallback exists: without a CUDA device every compute call fails.
 * Each function cites the reference interface (file:line under /root/reference) it replaces.
 */
#ifndef MAAVSS_B200_H
#define MAAVSS_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MAAVSS_ABI_VERSION 1

const char* maavss_last_error(void);
int maavss_abi_version(void);
int maavss_device_sm_count(void);
/* Number of kernels this library has launched (or captured into a CUDA graph) in this process so far. */
long long maavss_launch_count(void);

/* ------------------------------------------------------------------ audio front end ---------- */

typedef struct maavss_stft_plan maavss_stft_plan;

/* Twiddle table + window for one (n_fft, hop).  window_host: n_fft floats on the HOST
 * (AV_Dataset.window = torch.hamming_window(fft_len), av_dataset.py:106).  n_fft: power of two in [64, 4096]. */
int maavss_stft_plan_create(maavss_stft_plan** plan, int n_fft, int hop, const float* window_host);
void maavss_stft_plan_destroy(maavss_stft_plan* plan);

/* T_out = n_samples / hop (the 1 + N//hop centre-padded frames minus the dropped last one),
 * F_out = n_fft/2 (trim_stft_end) or n_fft/2 + 1.   av_dataset.py:171-174 */
int maavss_stft_out_dims(int n_samples, int n_fft, int hop, int trim_stft_end, int* T_out, int* F_out);

/* AV_Dataset.stft + the permute(2,1,0) of gen_stft_example (av_dataset.py:157-179, :338):
 * audio (batch, n_samples) -> planes (batch, 2, T_out, F_out) = [real; imag] (or [|z|; angle] when polar). */
int maavss_stft_fwd(const maavss_stft_plan* plan, const float* audio, float* planes, int batch, int n_samples,
                    int trim_stft_end, int normalized, int polar, void* stream);

/* AV_Dataset.istft (av_dataset.py:181-201): planes (batch, 2, n_frames, F_in) -> audio (batch, hop*(n_frames-1)).
 * F_in = n_fft/2 when trim_stft_end (the Nyquist bin is re-padded with zeros), else n_fft/2 + 1. */
int maavss_istft_fwd(const maavss_stft_plan* plan, const float* planes, float* audio, int batch, int n_frames,
                     int trim_stft_end, int normalized, int polar, void* stream);

/* y *= 1 / max(|y| + eps) per clip (normalize_output_fft, av_dataset.py:339-340).  workspace: batch * 4 bytes. */
int maavss_absmax_normalise(float* x, int batch, long long per_clip, float eps, void* workspace, void* stream);

/* x = y + noise * noise_std (AV_Dataset.add_noise, av_dataset.py:217-220; the gaussian draw is the caller's). */
int maavss_add_noise(const float* y, const float* noise, float noise_std, float* x, long long n, void* stream);

/* (outer, 2, plane): [re; im] <-> [|z|; atan2(im, re)]   (av_dataset.py:176-178, :187-191; generator.py:84-95) */
int maavss_magphase_split(const float* in, float* out, long long outer, long long plane, void* stream);
int maavss_magphase_combine(const float* in, float* out, long long outer, long long plane, void* stream);

/* DataGenerator.center_fft_bins / decenter_fft_bins over the last axis of (rows, K)  (generator.py:117-133) */
int maavss_bins_center(const float* in, float* out, long long rows, int K, void* stream);
int maavss_bins_decenter(const float* in, float* out, long long rows, int K, void* stream);

/* DataGenerator.fft / ifft (generator.py:60-81): whole-clip DFT of arbitrary (non power-of-two) length. */
typedef struct maavss_dft_plan maavss_dft_plan;
int maavss_dft_plan_create(maavss_dft_plan** plan, int n);
void maavss_dft_plan_destroy(maavss_dft_plan* plan);
/* audio (batch, n) real -> out (batch, n_keep) complex64 interleaved, bins [0, n_keep); 1/n when norm_forward. */
int maavss_dft_fwd(const maavss_dft_plan* plan, const float* audio, float* out_complex, int batch, int n_keep,
                   int norm_forward, void* stream);
/* in (batch, n/2) complex64, zero-extended to n bins, inverse DFT, real part -> audio (batch, n). */
int maavss_dft_half_inv(const maavss_dft_plan* plan, const float* in_complex, float* audio, int batch,
                        int norm_forward, void* stream);


/* ------------------------------------------------ fusion models: layers, loss, optimiser ------ */
/* All tensors fp32, contiguous, channels-first (N, C, D, H, W) with D = 1 for the 2-D layers.        */

/* One convolution geometry; shared by Conv2d / Conv3d forward (conv_fwd), their input gradient and the
 * ConvTranspose2d forward (conv_dgrad), and the weight gradient (conv_wgrad).
 * nn.Conv2d / nn.BatchNorm2d / nn.Tanh stacks: avse_model.py:427-446,474-502; avse_model_final.py:33-59,75-107;
 * nn.ConvTranspose2d stacks: avse_model.py:449-464,564-592; avse_model_conv.py:448-464,563-594. */
typedef struct maavss_conv_desc {
    int N, Cin, D, H, W;          /* input  (N, Cin, D, H, W) */
    int Cout, KD, KH, KW;         /* weight (Cout, Cin, KD, KH, KW) */
    int SD, SH, SW, PD, PH, PW;   /* stride, zero padding */
    int OD, OH, OW;               /* output (N, Cout, OD, OH, OW); given explicitly so output_padding is expressible */
} maavss_conv_desc;

/* y = conv(x, w) + bias (bias may be NULL). */
int maavss_conv_fwd(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, void* stream);
/* dx = conv_transpose(dy, w) (+ bias per Cin channel when used as the ConvTranspose forward; may be NULL). */
int maavss_conv_dgrad(const maavss_conv_desc* d, const float* dy, const float* w, const float* bias, float* dx, void* stream);
/* dw (+)= sum over batch/positions of dy (x) x.   accumulate != 0 adds into dw (gradient accumulation over num_seq,
 * train.py:158), else dw is overwritten. */
int maavss_conv_wgrad(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate, void* stream);
/* Tensor-core (tcgen05 + TMA, bf16 operands, fp32 accumulate) variant of maavss_conv_fwd (mode 0: act = x, out = y) and
 * maavss_conv_dgrad (mode 1: act = dy, out = dx) for the stride-1 "same" Conv3d / Conv2d layers of the
 * AV_Fusion_Model_Frames visual encoder (nn.Conv3d(k=(3,5,5), bias=False), avse_model_final.py:33-59): channels in
 * {16, 32, 64}, W in {32, 64, 128}, H a multiple of 128 / W.  No bias.  workspace (caller-owned, 1024-byte aligned,
 * maavss_conv_tc_workspace_bytes) holds the channels-last bf16 copy of act and the repacked weights. */
int maavss_conv_tc_supported(const maavss_conv_desc* d, int mode);
long long maavss_conv_tc_workspace_bytes(const maavss_conv_desc* d, int mode);
int maavss_conv_tc(const maavss_conv_desc* d, int mode, const float* act, const float* w, float* out, void* workspace,
                   long long workspace_bytes, void* stream);
/* Tensor-core variant of maavss_conv_wgrad for the same layer envelope (Cout in {16, 32, 64}): both operands are read
 * MN-major from channels-last bf16 copies in the workspace; partial sums meet in dw through atomicAdd. */
int maavss_conv_wgrad_tc_supported(const maavss_conv_desc* d);
long long maavss_conv_wgrad_tc_workspace_bytes(const maavss_conv_desc* d);
int maavss_conv_wgrad_tc(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate, void* workspace,
                         long long workspace_bytes, void* stream);
/* fp32-accurate tensor-core path (3xTF32 on tcgen05, tc_gemm3.cu) for the mid / deep Conv2d layers of the fusion encoders
 * (avse_model.py:427-446,474-502; avse_model_conv.py:427-500): same products as maavss_conv_fwd / _dgrad / _wgrad, <= 5e-6 of
 * torch fp32.  supported(): 2-D layers (D = KD = 1) with Cin >= 16 (or Cin >= 8 and Cout >= 32), Cout a multiple of 4, Cin*KH*KW a multiple of 4,
 * OH*OW a multiple of 32, N*OH*OW >= 128 and an im2col matrix of <= 512 MB.
 *   fwd    writes the im2col matrix col (N*OH*OW, Cin*KH*KW) (caller-owned, maavss_conv_tc3_col_elems floats; keep it for
 *          wgrad) and y (N, Cout, OH, OW) = col w^T + bias
 *   dgrad  dx (N, Cin, H, W) from dy_rows (N*OH*OW, Cout) = dy with the channel axis last (maavss_nchw_to_rows) and w_perm
 *          (Cout, KH*KW, Cin) = the tap-major copy of w (maavss_conv_tc3_permute_weights; weights are constant over a step)
 *   wgrad  dw (Cout, Cin, KH, KW) (+)= dy_rows^T col
 * workspace: maavss_conv_tc3_workspace_bytes(d, mode) with mode 0 fwd / 1 dgrad / 2 wgrad. */
int maavss_conv_tc3_supported(const maavss_conv_desc* d);
long long maavss_conv_tc3_col_elems(const maavss_conv_desc* d);
long long maavss_conv_tc3_workspace_bytes(const maavss_conv_desc* d, int mode);
int maavss_conv_tc3_fwd(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, float* col, void* workspace,
                        long long workspace_bytes, void* stream);
int maavss_conv_tc3_permute_weights(const maavss_conv_desc* d, const float* w, float* w_perm, void* stream);
int maavss_conv_tc3_dgrad(const maavss_conv_desc* d, const float* dy_rows, const float* w_perm, float* dx, void* workspace,
                          long long workspace_bytes, void* stream);
int maavss_conv_tc3_wgrad(const maavss_conv_desc* d, const float* col, const float* dy_rows, float* dw, int accumulate, void* workspace,
                          long long workspace_bytes, void* stream);
/* out[(n * inner + q) * C + c] = in[(n * C + c) * inner + q]   (channels-first -> one row of C channels per position) */
int maavss_nchw_to_rows(const float* in, int N, int C, long long inner, float* out, void* stream);
/* out[c] (+)= sum_{n,i} x[n, c, i]   (bias gradients). workspace: maavss_reduce_workspace_bytes(C). */
int maavss_channel_sum(const float* x, int N, int C, long long inner, float* out, int accumulate, void* workspace, void* stream);
int maavss_reduce_workspace_bytes(int C);
/* Allocation-free variant of maavss_channel_sum.  NO counterpart in the reference: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same kernels and results as maavss_channel_sum, but the partial sums live in a static 512 KB device array inside
 * the library (held by every GPU that loads it until exit) instead of a caller workspace; the call allocates nothing.
 * Limit: C <= maavss_channel_sum_static_max_channels() = 4096, larger C is rejected with MAAVSS_ERR_INVALID.  All calls share
 * the array: calls that may overlap on the GPU (different streams without ordering) race -- serialise them.  The first call
 * on a device resolves the array's address (cudaGetSymbolAddress): make it outside stream capture. */
int maavss_channel_sum_static(const float* x, int N, int C, long long inner, float* out, int accumulate, void* stream);
int maavss_channel_sum_static_max_channels(void);

/* Activations: 0 = identity, 1 = tanh, 2 = leaky_relu(slope), 3 = sigmoid. */
#define MAAVSS_ACT_NONE 0
#define MAAVSS_ACT_TANH 1
#define MAAVSS_ACT_LEAKY 2
#define MAAVSS_ACT_SIGMOID 3
#define MAAVSS_ACT_GELU 4   /* erf GELU; tensor-core GEMM epilogue only */

/* Train-mode BatchNorm (nn.BatchNorm2d / BatchNorm3d, eps 1e-5, momentum 0.1) split so that a data-parallel
 * caller can all-reduce `partial` between the two halves:
 *   bn_reduce   : partial[c][s] = {sum x, sum x^2} (double) over split s of the (n, inner) domain; S = maavss_bn_splits().
 *   bn_finalize : mean / invstd -> stats[c] = {mean, invstd}; running stats and num_batches_tracked updated when
 *                 running_mean != NULL (count = elements per channel over the WHOLE batch, all ranks).
 *                 use_running != 0: eval mode, stats come from running_mean / running_var, partial is ignored.
 *   bn_apply_act: y = act(((x - mean) * invstd) * gamma + beta). */
int maavss_bn_splits(int N, int C, long long inner);
int maavss_bn_reduce(const float* x, int N, int C, long long inner, double* partial, void* stream);
int maavss_bn_finalize(const double* partial, int S, int C, double count, float eps, float momentum, float* running_mean,
                       float* running_var, long long* num_batches_tracked, int use_running, float* stats, void* stream);
int maavss_bn_apply_act(const float* x, const float* stats, const float* gamma, const float* beta, int N, int C,
                        long long inner, int act, float slope, float* y, void* stream);
/* Backward of act(bn(x)):  dz = dy * act'(y);
 *   bn_bwd_reduce  : partial[c][s] = {sum dz, sum dz * xhat}
 *   bn_bwd_finalize: coef[c] = {sum dz / count, sum dz*xhat / count} (zeros in eval mode); dgamma (+)= , dbeta (+)=
 *   bn_bwd_apply   : dx = gamma * invstd * (dz - coef0 - xhat * coef1). */
int maavss_bn_bwd_reduce(const float* dy, const float* y, const float* x, const float* stats, int N, int C, long long inner,
                         int act, float slope, double* partial, void* stream);
int maavss_bn_bwd_finalize(const double* partial, int S, int C, double count, int eval_mode, float* coef, float* dgamma,
                           float* dbeta, int accumulate, void* stream);
int maavss_bn_bwd_apply(const float* dy, const float* y, const float* x, const float* stats, const float* gamma,
                        const float* coef, int N, int C, long long inner, int act, float slope, float* dx, void* stream);

/* Two-launch train-mode BatchNorm (+ activation) over `groups` independent sub-batches of N/groups samples, each with
 * its own batch statistics: the num_seq accumulation windows of one optimizer step (train.py:136-159) batched into one
 * pass (groups = 1 is the plain layer).  reduce -> [data-parallel all-reduce of `partial`] -> apply; the apply kernels
 * re-derive the statistics from the partials themselves (no finalize launch).
 *   partial: (groups, C, S, 2) doubles, S = maavss_bn_splits(N/groups, C, inner);  count = elements per (group, channel)
 *   over ALL ranks;  stats (groups, C, 2) = {mean, invstd} is written by apply and saved for the backward.
 *   Running statistics receive the `groups` momentum updates in window order; num_batches_tracked += groups. */
int maavss_bn_reduce_grouped(const float* x, int N, int C, long long inner, int groups, double* partial, void* stream);
int maavss_bn_apply_grouped(const float* x, const double* partial, int N, int C, long long inner, int groups, double count,
                            const float* gamma, const float* beta, float* running_mean, float* running_var,
                            long long* num_batches_tracked, float momentum, float eps, int act, float slope, float* stats, float* y,
                            void* stream);
int maavss_bn_bwd_reduce_grouped(const float* dy, const float* y, const float* x, const float* stats, int N, int C, long long inner,
                                 int groups, int act, float slope, double* partial, void* stream);
/* partial: sums used for the input gradient (all-reduced when data parallel); partial_local: this rank's sums, used for
 * dgamma / dbeta (+)= (may be the same pointer).  dx may be NULL. */
int maavss_bn_bwd_apply_grouped(const float* dy, const float* y, const float* x, const float* stats, const float* gamma,
                                const double* partial, const double* partial_local, int N, int C, long long inner, int groups,
                                double count, int act, float slope, float* dgamma, float* dbeta, int accumulate, float* dx,
                                void* stream);

/* Elementwise activation and its backward (dz = dy * act'(y), from the OUTPUT y). */
int maavss_act_fwd(const float* x, float* y, long long n, int act, float slope, void* stream);
int maavss_act_bwd(const float* dy, const float* y, float* dz, long long n, int act, float slope, void* stream);

/* nn.MaxPool3d((1, ph, pw)) (stride = kernel, floor mode; avse_model_final.py:36-57) over the (H, W) plane of each of the
 * `planes` = N*C*D leading slices.  idx: one byte per output element = arg-max inside its window (row-major). */
int maavss_maxpool_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, int ph, int pw,
                       void* stream);
int maavss_maxpool_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, int ph, int pw,
                       void* stream);

/* nn.Linear (+ fused activation): y[M, N] = act(x[M, K] w[N, K]^T + bias)   (avse_model.py:550,553,618-627). */
int maavss_linear_fwd(const float* x, const float* w, const float* bias, float* y, int M, int N, int K, int act, float slope,
                      void* stream);
/* dx[M, K] = dz[M, N] w[N, K]   (dx overwritten). */
int maavss_linear_dgrad(const float* dz, const float* w, float* dx, int M, int N, int K, void* stream);
/* dw[N, K] (+)= dz[M, N]^T x[M, K];  dbias[N] (+)= sum_m dz (dbias may be NULL). */
int maavss_linear_wgrad(const float* dz, const float* x, float* dw, float* dbias, int M, int N, int K, int accumulate,
                        void* stream);

/* nn.LSTM(bias=False, batch_first=True, bidirectional=True, num_layers=1)   (avse_model.py:542-543).
 * Gate order i, f, g, o.  gin: (2, B, T, 4H) = x w_ih^T per direction (a linear_fwd each).  One launch per time step
 * handles both directions:  step s uses t = s (dir 0) / T-1-s (dir 1).
 *   gates: (2, B, T, 4H) post-activation gates (saved for backward)      cells: (2, B, T, H)
 *   y: (B, T, 2H)   w_hh: two pointers (H4 x H) */
int maavss_lstm_step_fwd(const float* gin, const float* w_hh_fwd, const float* w_hh_rev, float* gates, float* cells, float* y,
                         int B, int T, int H, int step, void* stream);
/* Backward step s (run s = T-1 .. 0): dgates (2, B, T, 4H) receives the pre-activation gate gradients of time t(s);
 * dcell: (2, B, H) running cell gradient (zero-initialised by the caller). */
int maavss_lstm_step_bwd(const float* dy, const float* gates, const float* cells, const float* w_hh_fwd, const float* w_hh_rev,
                         float* dgates, float* dcell, int B, int T, int H, int step, void* stream);

/* The T recurrent steps of both directions as ONE launch (lstm_persist.cu; H = 256 as the reference builds it,
 * avse_model.py:542): thread-block clusters of 8 CTAs per (direction, 8 batch rows) keep w_hh in registers for the whole
 * sequence and exchange the hidden state / gate gradients through distributed shared memory.  Same arguments and results as
 * running maavss_lstm_step_fwd / _bwd for step = 0..T-1 / T-1..0 (the backward needs no dcell buffer). */
int maavss_bilstm_recurrent_supported(int B, int T, int H);
int maavss_bilstm_recurrent_fwd(const float* gin, const float* w_hh_fwd, const float* w_hh_rev, float* gates, float* cells, float* y, int B,
                                int T, int H, void* stream);
int maavss_bilstm_recurrent_bwd(const float* dy, const float* gates, const float* cells, const float* w_hh_fwd, const float* w_hh_rev,
                                float* dgates, int B, int T, int H, void* stream);

/* The whole layer in one call (SURVEY 8b bilstm_fwd / bilstm_bwd): x (B, T, I) -> y (B, T, 2H).  gin / gates (2, B, T, 4H) and
 * cells (2, B, T, H) are outputs the backward needs (caller-owned).  The input projections and all weight / input gradients
 * run on the 3xTF32 tensor-core GEMM when B*T >= 32, the T recurrent steps are one launch each.
 * bilstm_bwd: dgates (2, B, T, 4H) is scratch / output; dx (B, T, I) and the four dw_* (same shapes as the weights) may be
 * NULL; accumulate != 0 adds into the dw_* (gradient accumulation).  Order on the stream: T steps, dx, then the dw_*.  workspace: maavss_bilstm_workspace_bytes(.., backward). */
long long maavss_bilstm_workspace_bytes(int B, int T, int I, int H, int backward);
int maavss_bilstm_fwd(const float* x, const float* w_ih_fwd, const float* w_hh_fwd, const float* w_ih_rev, const float* w_hh_rev,
                      float* gin, float* gates, float* cells, float* y, int B, int T, int I, int H, void* workspace,
                      long long workspace_bytes, void* stream);
int maavss_bilstm_bwd(const float* dy, const float* x, const float* w_ih_fwd, const float* w_hh_fwd, const float* w_ih_rev,
                      const float* w_hh_rev, const float* gates, const float* cells, const float* y, float* dgates, float* dx,
                      float* dw_ih_fwd, float* dw_hh_fwd, float* dw_ih_rev, float* dw_hh_rev, int accumulate, int B, int T, int I, int H,
                      void* workspace, long long workspace_bytes, void* stream);
/* The weight-gradient half of bilstm_bwd on its own (after a bilstm_bwd call with the four dw_* NULL, dgates kept): lets the
 * caller enqueue it on another stream, off the chain dy -> dx that the rest of the backward pass waits for.  Own workspace
 * of maavss_bilstm_workspace_bytes(.., 1) bytes (must not be shared with a bilstm_bwd still in flight). */
int maavss_bilstm_wgrad(const float* x, const float* y, const float* dgates, float* dw_ih_fwd, float* dw_hh_fwd, float* dw_ih_rev,
                        float* dw_hh_rev, int accumulate, int B, int T, int I, int H, void* workspace, long long workspace_bytes,
                        void* stream);

/* nn.MSELoss (mean): loss[0] = mean((a - b)^2)   (train.py:154-156).  workspace: maavss_reduce_workspace_bytes(8). */
int maavss_mse_fwd(const float* a, const float* b, long long n, float* loss, void* workspace, void* stream);
/* da (+)= gout[0] * 2 (a - b) / n   (gout is a device scalar: no host sync). */
int maavss_mse_bwd(const float* a, const float* b, const float* gout, long long n, float* da, int accumulate, void* stream);

/* torch.optim.Adam step (train.py:95,161) over one flat fp32 arena: 28 B/param of HBM traffic.
 * state: 16 device bytes {int64 step; float step_size; float inv_sqrt_bias_correction2}, zero-initialised by the caller;
 * the step counter is advanced ON THE DEVICE (so a captured CUDA graph of the train step stays correct on replay).
 * grad_scale multiplies g (1/world_size after a sum all-reduce). */
int maavss_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps,
                     float weight_decay, void* state, float grad_scale, void* stream);
/* The same step in two halves, for a train step that updates part of the arena early (train_step.EarlyTail: the LSTM / fc
 * weights while the encoders still back-propagate): adam_prepare advances the device step counter and bias corrections ONCE
 * per optimizer step, adam_apply updates any 16-byte aligned sub-range with them.  zero_grad != 0 also clears g[0..n)
 * (optimizer.zero_grad(), train.py:162, fused into the same pass); max_ctas > 0 caps the grid (an update that shares the
 * GPU with latency-bound kernels). */
int maavss_adam_prepare(void* state, float lr, float beta1, float beta2, void* stream);
int maavss_adam_apply(float* p, float* g, float* m, float* v, long long n, float beta1, float beta2, float eps, float weight_decay,
                      const void* state, float grad_scale, int zero_grad, int max_ctas, void* stream);
/* x[0..n) = value */
int maavss_fill(float* x, long long n, float value, void* stream);

/* ------------------------------------------------------- motion-attention front end ---------- */

/* av_dataset.py:323-328: per clip, out = attn_diff ? cat(0, diff(attn, dim=0)) : attn, then *= 1/max(out)
 * (the SIGNED max, as the reference).  in/out (clips, frames, hw).  workspace: clips * 4 bytes. */
int maavss_attn_diff_normalise(const float* in, float* out, int clips, int frames, long long hw, int diff,
                               void* workspace, void* stream);

/* video_attention.py:56,80-96: cls_rows (frames, heads, cls_stride) = last-block softmax row of the CLS query
 * (entry 0 = CLS itself, then grid_h*grid_w patch tokens); sum over heads, nearest x patch upsample,
 * divide by the per-frame max -> out (frames, grid_h*patch, grid_w*patch). */
int maavss_attn_headsum_upsample(const float* cls_rows, float* out, int frames, int heads, int cls_stride,
                                 int grid_h, int grid_w, int patch, void* stream);

/* ---- attention-map cache: raw uint8 / fp16 frames instead of the per-frame JPEG files the
 * reference writes with torchvision.utils.save_image (av_dataset.py:270-276, save_attn_videos.py:10-19) and reads
 * back through PIL + ToTensor (av_dataset.py:251-269). */
#define MAAVSS_CACHE_U8 0  /* save_image's quantiser: (unsigned char)clamp(x*255 + 0.5, 0, 255); read back as u8 / 255 */
#define MAAVSS_CACHE_F16 1 /* round-to-nearest-even fp16 */
/* maps[0..n) fp32 (the [0,1] per-frame-normalised output of VideoAttention._inference) -> stored[0..n) in `format`.
 * maps 16-byte aligned, stored 8-byte aligned.  5 (u8) / 6 (fp16) bytes of HBM traffic per pixel. */
int maavss_attn_cache_pack(const float* maps, void* stored, long long n, int format, void* stream);
/* stored (clips, frames, hw) in `format` -> out (clips, frames, hw) fp32: dequantise, then av_dataset.py:323-328 per clip
 * (diff ? cat(0, diff(attn, dim=0)) : attn, then *= 1/max, the signed max) with no fp32 intermediate: the stored frames are
 * read twice (max pass, scale pass) -- by one thread-block cluster per clip in a single launch (fp16) or by two launches
 * (uint8, and any frame size that is not a multiple of 4 pixels); s + 4 algorithmic bytes per pixel.
 * workspace: clips * 4 bytes (the two-launch path's per-clip maxima). */
int maavss_attn_cache_load(const void* stored, int format, float* out, int clips, int frames, long long hw, int diff,
                           void* workspace, void* stream);

/* ---- DINO ViT-S/8 last-layer attention maps (video_attention.py:38-103 over dino/vision_transformer.py) ----
 * bf16 operands, fp32 accumulation / residual stream.  All GEMM-shaped work runs on tcgen05 tensor cores fed by TMA.
 *
 * tc_gemm_bf16: D[M, N] = A[M, K] B[N, K]^T (+ bias[N]) (-> GELU) (+ residual[M, N] fp32); A, B bf16 row-major with row
 * pitches lda / ldb (elements, multiples of 8); out fp32 or bf16 with pitch ldc.  nn.Linear of Attention.qkv / proj
 * (dino/vision_transformer.py:75-78), Mlp.fc1 / fc2 (:51-56) and PatchEmbed.proj as a GEMM (:126-130). */
int maavss_tc_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, const float* bias,
                        const float* residual, long long ldr, void* out, long long ldc, int out_bf16, int act, void* stream);
/* Attention.forward core (dino/vision_transformer.py:82-90): qkv (T, N, 3*heads*64) bf16 -> out (T, N, heads*64) bf16,
 * out = softmax(q k^T * scale) v per (frame, head); head_dim is 64. */
int maavss_vit_attention(const void* qkv_bf16, void* out_bf16, int T, int N, int heads, float scale, void* stream);
/* PatchEmbed operand gather: frames (T, C, H, W) fp32 (top-left gh*P x gw*P crop, video_attention.py:43-47) ->
 * patches (T*gh*gw, C*P*P) bf16 with column order (c, ky, kx) = Conv2d weight order. */
int maavss_vit_patchify(const float* frames, void* patches_bf16, int T, int C, int H, int W, int gh, int gw, int P, void* stream);
/* prepare_tokens (dino/vision_transformer.py:196-207): x (T, 1+np, D) = cat(cls, patch_tok (T*np, D)) + pos (1+np, D). */
int maavss_vit_assemble_tokens(const float* patch_tok, const float* cls, const float* pos, float* x, int T, int np, int D,
                               void* stream);
/* nn.LayerNorm(D, eps) over rows of x fp32 -> y bf16 (Block.norm1 / norm2, dino/vision_transformer.py:98,103). */
int maavss_layernorm_bf16(const float* x, const float* gamma, const float* beta, void* y_bf16, long long rows, int D, float eps,
                          void* stream);
int maavss_f32_to_bf16(const float* in, void* out_bf16, long long n, void* stream);
/* Last block with the CLS query only (get_last_selfattention, dino/vision_transformer.py:216-223; video_attention.py:56):
 * q (T rows, pitch ldq) fp32, k (T*N rows, pitch ldk) fp32 -> out (T, heads, N) = softmax_j(q_h . k_jh * scale). */
int maavss_vit_cls_attn_row(const float* q, long long ldq, const float* k, long long ldk, float* out, int T, int heads, int N,
                            float scale, void* stream);

/* Decoded clip frames (frames, H, W, 3) uint8 -> (frames, 3, H, W) fp32 = (x / 255 - mean[c]) / std[c]: the permute, /255. and
 * Normalize of gen_av_example (av_dataset.py:108-112,312-317), bit-identical to the torch expression.  mean / std: 3 HOST
 * floats each.  (RandomResizedCrop is a random augmentation and stays with the caller.) */
int maavss_frames_u8_normalize(const unsigned char* frames_hwc, float* out_chw, long long frames, int H, int W,
                               const float* mean3_host, const float* std3_host, void* stream);

/* ---- the whole extractor as one call: VideoAttention.__load_model / _inference (video_attention.py:38-160).
 * vit_create copies the weights into a plan (GEMM weights as bf16).  All pointers are DEVICE fp32 tensors in dino's
 * state_dict layout: cls_token (dim), pos_embed (1 + pos_grid^2, dim), patch_embed.proj.weight (dim, in_chans*patch*patch),
 * .bias (dim); block_tensors = a HOST array of depth*12 device pointers, per block: norm1.weight, norm1.bias, attn.qkv.weight,
 * attn.qkv.bias, attn.proj.weight, attn.proj.bias, norm2.weight, norm2.bias, mlp.fc1.weight, mlp.fc1.bias, mlp.fc2.weight,
 * mlp.fc2.bias.  head_dim is 64 (vit_tiny / small / base), LayerNorm eps 1e-6 (dino/vision_transformer.py:229-254).
 * vit_prepare evaluates interpolate_pos_encoding (:174-194, bicubic, the +0.1 trick) for one input size, once.
 * vit_attn_maps: frames (T, in_chans, H, W) -> out (T, gh*patch, gw*patch):
 *   post 0: the reference `_inference` output (heads summed, nearest x patch upsample, each frame / its max);
 *   post 1 / 2: followed by av_dataset.py:323-328 per clip of T/clips frames (2 = with the temporal frame difference).
 * workspace: maavss_vit_workspace_bytes(plan, T, H, W, clips), caller-owned. */
typedef struct maavss_vit_plan maavss_vit_plan;
int maavss_vit_create(maavss_vit_plan** plan, int dim, int depth, int patch, int in_chans, int pos_grid, const float* cls_token,
                      const float* pos_embed, const float* patch_w, const float* patch_b, const float* const* block_tensors);
void maavss_vit_destroy(maavss_vit_plan* plan);
int maavss_vit_prepare(maavss_vit_plan* plan, int H, int W);
long long maavss_vit_workspace_bytes(const maavss_vit_plan* plan, int T, int H, int W, int clips);
/* copies the prepared (1 + gh*gw, dim) position embedding of that size to `out` (tests). */
int maavss_vit_pos_embed(const maavss_vit_plan* plan, int H, int W, float* out);
int maavss_vit_attn_maps(const maavss_vit_plan* plan, const float* frames, int T, int H, int W, float* out, int post, int clips,
                         void* workspace, long long workspace_bytes, void* stream);

/* utilities.video_phasegram (utilities.py:206-228; called on every micro-step, train.py:139):
 * frames (B, T, H, W) with clip stride batch_stride (elements) -> out (B, T, R*R): antialiased bilinear resize to (R, R), fft2, fftshift over all four dims,
 * angle, flatten, [cumsum / (2 pi R^2) | (p + pi) / 2 pi], [temporal diff with a zero first frame], [/ max |.| over the
 * whole tensor].  scratch: B*T*R*R floats; max_ws: 4 bytes. */
int maavss_video_phasegram(const float* frames, float* out, float* scratch, void* max_ws, int B, int T, int H, int W, int R,
                           long long batch_stride, int diff, int cumulative, int normalize, void* stream);

/* fp32-accurate GEMM on the tensor cores (3xTF32 on tcgen05.mma.kind::tf32, operands staged by TMA, accumulators in TMEM):
 *     D[P, Q] = A . B^T (+ bias) (-> act) (+ out),  A: P x K, B: Q x K, everything fp32.
 * Used for the GEMM-shaped layers of the fusion models whose parity bar is fp32: nn.Linear fc1 / fc2 / a_fc1 / v_fc1 with the
 * accumulation windows stacked (avse_model.py:550,553,618-627; M = 64 rows) and the Bi-LSTM input projections (:542-543):
 *   forward        y[M, N]  = x w^T      : A = w  (K-major), B = x  (K-major), transposed output, bias per p
 *   data gradient  dx[M, K] = dz w       : A = w  (MN-major: stored [N][K]), B = dz (K-major), transposed output
 *   weight grad.   dw[N, K] += dz^T x    : A = dz (MN-major: stored [M][N]), B = x (MN-major), row-major output, accumulate
 * a_mn_major / b_mn_major: 0 = the operand is stored [rows][K] with pitch lda / ldb, 1 = stored [K][rows].
 * out_transposed: 0 -> out[p * ldc + q], 1 -> out[q * ldc + p].  bias: per p (bias_on_q = 0) or per q, may be NULL.
 * Small products are split along K over the SMs; the partial tiles meet in `workspace` (maavss_gemm_tf32x3_workspace_bytes;
 * 0 when no split is needed) and are reduced in a fixed order.  Pitches must be multiples of 4 floats, pointers 16-byte
 * aligned, and Q, ldc multiples of 4 for a row-major output (maavss_gemm_tf32x3_supported). */
int maavss_gemm_tf32x3_supported(int P, int Q, int K, int a_mn_major, long long lda, int b_mn_major, long long ldb, long long ldc,
                                 int out_transposed);
long long maavss_gemm_tf32x3_workspace_bytes(int P, int Q, int K);
int maavss_gemm_tf32x3(const float* A, int a_mn_major, long long lda, const float* B, int b_mn_major, long long ldb, int P, int Q,
                       int K, float* out, long long ldc, int out_transposed, const float* bias, int bias_on_q, int act, float slope,
                       int accumulate, void* workspace, long long workspace_bytes, void* stream);
/* The same product with the transposed output written in blocks of p_block rows of p (p_block % 32 == 0):
 *   out[(p / p_block) * p_block_stride + q * ldc + p % p_block]
 * -- with p = (sample, position), p_block = positions per sample, ldc = p_block, p_block_stride = channels * p_block this is a
 * convolution output in NCHW straight from the GEMM over im2col rows (conv_tc3.cu). */
int maavss_gemm_tf32x3_blocked(const float* A, int a_mn_major, long long lda, const float* B, int b_mn_major, long long ldb, int P, int Q,
                               int K, float* out, long long ldc, int out_transposed, int p_block, long long p_block_stride,
                               const float* bias, int bias_on_q, int act, float slope, int accumulate, void* workspace,
                               long long workspace_bytes, void* stream);
/* out[c] (+)= sum_r x[r * cols + c]   (bias gradient of a Linear layer from its (rows, cols) output gradient). */
int maavss_colsum(const float* x, int rows, long long cols, float* out, int accumulate, void* stream);

/* The num_seq accumulation windows of one clip batch stacked along the batch axis (x_stft[:, :, hpf*j : hpf*j + span],
 * j = 0..G-1, train.py:140-147): x (B, C, T, F) -> out (G*B, C, span, F), out[j*B + b] = x[b, :, j*hop : j*hop + span]. */
int maavss_stack_windows(const float* x, float* out, int B, int C, int T, int F, int G, int hop, int span, void* stream);

/* The same op split where the reference's own structure lets work be shared: train.py:136-139 calls video_phasegram on
 * the num_seq overlapping windows attn[:, :, j:j+Tw] of one Tc-frame clip (4 windows of 8 over 12 frames), and everything
 * before the temporal diff is per frame.
 *   phasegram_frames       : frames (B, Tc, H, W) -> pflat (B, Tc, R*R), per frame: resize, fft2, spatial fftshift, angle,
 *                            [cumsum / (2 pi R^2) | (p + pi) / 2 pi].  No batch / time roll.  scratch: B*Tc*R*R floats.
 *   phasegram_windows_max  : max_ws[j] = float bits of max |window j| for the G windows j = 0..G-1 (frames j .. j+Tw-1):
 *                            out_j[bd, td] = v(td) - v(td-1) (0 at td = 0) with v(td) = pflat[(bd - batch_roll) mod B][j + (td - Tw/2) mod Tw]
 *                            -- the batch and time rolls of the all-dims fftshift (utilities.py:210) and the diff (:219-222).
 *                            batch_roll = B/2 is the reference on one process; a data-parallel rank whose loader already
 *                            supplies the attention frames of global clip (i - B_global/2) mod B_global passes 0 and
 *                            all-reduces (MAX, as uint32) max_ws between the two calls (utilities.py:226-227 is a whole-batch max).
 *   phasegram_windows_write: out (G, B, Tw, R*R) = out_j * (normalize ? 1 / max_ws[j] : 1): the G windows stacked along the
 *                            batch axis, ready for one batched forward. */
int maavss_phasegram_frames(const float* frames, float* pflat, float* scratch, int B, int Tc, int H, int W, int R,
                            long long batch_stride, int cumulative, void* stream);
int maavss_phasegram_windows_max(const float* pflat, void* max_ws, int B, int Tc, int Tw, int G, int R, int batch_roll, int diff,
                                 void* stream);
int maavss_phasegram_windows_write(const float* pflat, const void* max_ws, float* out, int B, int Tc, int Tw, int G, int R,
                                   int batch_roll, int diff, int normalize, void* stream);

/* ------------------------------------------------------------- sparse optical flow (flow.py) ---------- */
/* flow.py:26 cv2.cvtColor(frame, COLOR_BGR2GRAY) for uint8 frames: (B*3735 + G*19235 + R*9798 + 2^14) >> 15 (OpenCV 4.x). */
int maavss_bgr_to_gray(const unsigned char* bgr, unsigned char* gray, long long pixels, void* stream);
/* flow.py:11-14,28 cv2.goodFeaturesToTrack(gray, maxCorners, qualityLevel, minDistance, blockSize) (Shi-Tomasi: minimum
 * eigenvalue of the block_size x block_size gradient covariance, Sobel 3; threshold at quality * max; 3x3 non-maximum
 * suppression; strongest first with a minimum mutual distance).  gray (H, W) uint8 -> corners_xy (max_corners, 2) = (x, y)
 * floats in OpenCV's order, *count (device int) = how many.  workspace: maavss_flow_features_workspace_bytes(H, W). */
long long maavss_flow_features_workspace_bytes(int H, int W);
int maavss_flow_good_features(const unsigned char* gray, int H, int W, int max_corners, float quality_level, float min_distance,
                              int block_size, float* corners_xy, int* count, void* workspace, long long workspace_bytes, void* stream);
/* flow.py:17-19,37 cv2.calcOpticalFlowPyrLK(prev, next, p0, None, winSize=(win, win), maxLevel, criteria=(EPS|COUNT, max_iter,
 * epsilon)): pyramids by pyrDown, Scharr derivatives, the 14-bit fixed-point bilinear windows and the iteration rules of
 * OpenCV's LKTrackerInvoker.  prev_pts / next_pts (max_points, 2) device floats (x, y); count: optional DEVICE int (points
 * beyond it are skipped; NULL = all max_points); status (max_points) 1 = tracked; err (max_points, may be NULL) = mean |J - I| / 32
 * over the window.  min_eig_threshold: OpenCV's default is 1e-4.  workspace: maavss_flow_lk_workspace_bytes(H, W, max_level). */
long long maavss_flow_lk_workspace_bytes(int H, int W, int max_level);
int maavss_flow_pyr_lk(const unsigned char* prev_gray, const unsigned char* next_gray, int H, int W, const float* prev_pts, int max_points,
                       const int* count, float* next_pts, unsigned char* status, float* err, int win, int max_level, int max_iter,
                       float epsilon, float min_eig_threshold, void* workspace, long long workspace_bytes, void* stream);

/* ------------------------------------------------------------- data parallel ---------- */
/* One NCCL communicator over the GPUs of the box (one process per GPU).  Rank 0 calls comm_unique_id (128 bytes) and hands
 * the bytes to every rank (any side channel: torch.distributed store, MPI, a file); every rank then calls comm_init with its
 * CUDA device current.  NCCL itself is loaded at run time (libnccl.so.2).  The reference is single-process: these are the
 * collectives its loop needs to run N ranks == one process on the concatenated batch:
 *   grad_allreduce     in-place SUM of the flat fp32 gradient arena after the num_seq accumulation windows (train.py:158-161);
 *                      the 1/world average is maavss_adam_step's grad_scale
 *   allreduce_sum_f64  train-mode BatchNorm partial sums between bn_reduce and bn_apply (statistics of the global batch)
 *   allreduce_max_u32  the whole-batch max of utilities.video_phasegram (utilities.py:226-227), float bits as uint32
 * All three enqueue on `stream` (capturable into the train step's CUDA graph). */
/* Small-message sum all-reduce over NVLink peer memory (peer_exchange.cu): the BatchNorm partial sums, 30 exchanges of <= 32 KB
 * per optimizer step on the critical path, as ONE single-CTA kernel per rank and exchange (peer stores + flag, spin on own
 * memory, sum in rank order: bit-identical on every rank) instead of an NCCL all-reduce each.
 *   peer_create     allocates the rank's buffer (channels x 4 slots x 16 ranks x max_doubles doubles) and returns its 64-byte CUDA
 *                   IPC handle; the caller gathers the handles of all ranks (any side channel)
 *   peer_connect    opens the other ranks' buffers: handles = world x 64 bytes, indexed by rank
 *   peer_allreduce_sum_f64   in place, n <= max_doubles, enqueued on `stream` (capturable).  Exchanges of one channel must be
 *                   issued in the same order on every rank; streams that run concurrently use different channels.
 *   peer_status     *error != 0 after a rank waited ~30 s for a peer and gave up (the results are then garbage, not a hang). */
typedef struct maavss_peer maavss_peer;
int maavss_peer_create(maavss_peer** peer, int rank, int world, int channels, int max_doubles, void* handle64);
int maavss_peer_connect(maavss_peer* peer, const void* handles);
int maavss_peer_capacity(const maavss_peer* peer);
int maavss_peer_allreduce_sum_f64(maavss_peer* peer, int channel, double* data, long long n, void* stream);
int maavss_peer_status(maavss_peer* peer, int* error);
int maavss_peer_destroy(maavss_peer* peer);

typedef struct maavss_comm maavss_comm;
int maavss_comm_unique_id(void* id128);
int maavss_comm_init(maavss_comm** comm, int rank, int world, const void* id128);
void maavss_comm_destroy(maavss_comm* comm);
int maavss_grad_allreduce(maavss_comm* comm, float* grad, long long n, void* stream);
int maavss_allreduce_sum_f64(maavss_comm* comm, double* v, long long n, void* stream);
int maavss_allreduce_max_u32(maavss_comm* comm, void* words, long long n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAAVSS_B200_H */
