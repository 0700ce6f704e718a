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

#ifdef __cplusplus
}
#endif
#endif /* MAAVSS_B200_H */
