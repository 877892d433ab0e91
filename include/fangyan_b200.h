/* fangyan_b200 -- C ABI of the B200-native stage 1+2 extraction path.
 *
 * The reference (Autonomof/fangyan-TTS) has no FFI for this path: its boundary is two Python CLIs that call
 * torchaudio / openai-whisper for the front-ends and onnxruntime InferenceSession.run() for the two networks.
 * Every entry point below cites the reference call it replaces (paths relative to /root/reference/CosyVoice).
 *
 * Conventions
 *   - every function returns FY_OK (0) or a negative fy_status; fy_last_error() gives a thread-local message.
 *   - all data pointers are DEVICE pointers unless the parameter name starts with h_ (host).
 *   - utterances are passed PACKED: one contiguous array plus an int32 offsets array of n_utts + 1 entries
 *     (exclusive prefix sums), never padded.  The reference runs every utterance alone at its true length
 *     (tools/extract_embedding.py:25-42), so there is no padding to reproduce.
 *   - launch calls (fy_resample, fy_fbank_cmn, fy_logmel, fy_matcha_mel, fy_campplus_forward, fy_s3tok_encode_fsq,
 *     fy_speaker_*) are asynchronous on the context's stream (fy_set_stream): they neither synchronise nor allocate
 *     nor copy from host memory -- per-call row maps are derived on the device from the offsets arrays -- so a
 *     sequence of them can be captured in a CUDA graph.  h_frame_off arguments are read on the host during the call
 *     (totals for grid sizes and the workspace check) and may be freed as soon as it returns.
 *     The calls that DO allocate and synchronise are the set-up ones: fy_create, fy_load_*, fy_resample_prepare,
 *     fy_profile_read, fy_destroy.
 *   - the caller owns inputs, outputs and the workspace; the library owns weights and lookup tables.
 *   - a fy_ctx is not thread-safe: one per worker / GPU.
 */
#ifndef FANGYAN_B200_H_
#define FANGYAN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fy_ctx fy_ctx;

typedef enum {
  FY_OK = 0,
  FY_ERR_ARG = -1,     /* bad argument / shape / alignment (checked on the host before any launch) */
  FY_ERR_CUDA = -2,    /* a CUDA runtime or driver call failed */
  FY_ERR_STATE = -3,   /* weights not loaded, wrong device, ... */
  FY_ERR_NOMEM = -4    /* workspace too small */
} fy_status;

typedef enum {
  FY_PREC_BF16 = 0,    /* single bf16 tensor-core pass, fp32 accumulate */
  FY_PREC_BF16X3 = 1   /* split-bf16 (hi*hi + hi*lo + lo*hi), fp32-faithful: the parity mode and the default */
} fy_precision;

typedef enum { FY_PCM_S16 = 0, FY_PCM_F32 = 1 } fy_pcm_dtype;

/* One named fp32 weight tensor on the HOST (an ONNX initialiser or a state-dict entry). */
typedef struct {
  const char* name;
  const float* h_data;
  int32_t ndim;
  int64_t shape[4];
} fy_tensor;

/* Hyper-parameters of the S3 tokenizer encoder (UNVERIFIED against the real ONNX graph). */
typedef struct {
  int32_t n_mels, d_model, n_head, n_layer, d_ff, fsmn_kernel, fsq_dim;
  float rope_theta, ln_eps, fsq_scale;
  float mlp_ln_eps;    /* epsilon of the MLP LayerNorm when it differs from the attention one (ln_eps); <= 0: same */
} fy_s3_config;

/* Hyper-parameters of CAMPPlus (UNVERIFIED against the real ONNX graph). */
typedef struct {
  int32_t feat_dim, embed_dim, fcm_channels, init_channels, growth, bn_size, n_blocks;
  int32_t layers[4], kernel[4], dilation[4];
  int32_t seg_len;
  float bn_eps;
} fy_campplus_config;

/* ---- lifetime ------------------------------------------------------------------------------------------- */
int fy_create(int device, int precision, fy_ctx** out);
void fy_destroy(fy_ctx* ctx);
const char* fy_last_error(void);
int fy_set_stream(fy_ctx* ctx, void* cuda_stream);
const char* fy_version(void);

/* ---- weights: replaces onnxruntime.InferenceSession(onnx_path) ---------------------------------------------
 * tools/extract_embedding.py:94, tools/extract_speech_token.py:90, cosyvoice/cli/frontend.py:45-48.
 * BatchNorm folding, bf16 hi/lo splitting and conv-weight re-layout happen here, once. */
int fy_load_campplus(fy_ctx* ctx, const fy_tensor* h_tensors, int n, const fy_campplus_config* cfg);
int fy_load_s3tok(fy_ctx* ctx, const fy_tensor* h_tensors, int n, const fy_s3_config* cfg);

/* ---- A1 / N2: torchaudio.load(path) for RIFF/WAVE files, on the HOST (no device work, no ctx) -------------------
 * tools/extract_embedding.py:26, tools/extract_speech_token.py:28, cosyvoice/utils/file_utils.py:45.
 * fy_wav_scan parses n headers with n_threads plain threads: frames, channels, rate, sample format and the byte
 * offset of the samples (what the host needs to plan batches before decoding anything); a file that cannot be opened
 * or is not RIFF/WAVE gets a negative status in ITS record, the call still returns FY_OK.
 * fy_wav_read_s16 copies n_samples[i] int16 values starting at data_offset[i] of file i into dst[i] -- host memory
 * of the caller, typically slices of a pinned staging buffer so the samples go page cache -> pinned -> GPU with one
 * copy on the host.  Interleaved channels are copied as they are (pass frames * channels).  On failure returns
 * FY_ERR_ARG, names the file in fy_last_error() and stores its index in *bad_index (may be NULL). */
typedef enum {
  FY_WAV_OK = 0,
  FY_WAV_ERR_OPEN = -1,         /* cannot open / read */
  FY_WAV_ERR_FORMAT = -2,       /* not RIFF/WAVE */
  FY_WAV_ERR_LONG_HEADER = -3   /* no fmt + data chunk within the first 4 KiB: needs a full parse */
} fy_wav_status;
typedef struct {
  int64_t frames;        /* samples per channel */
  int64_t data_offset;   /* byte offset of the first sample */
  int32_t channels, sample_rate;
  int32_t format;        /* 1 = integer PCM, 3 = IEEE float (WAVE_FORMAT_EXTENSIBLE resolved to its sub-format) */
  int32_t bits;          /* bits per sample */
  int32_t status;        /* fy_wav_status */
  int32_t reserved;
} fy_wav_info;
int fy_wav_scan(const char* const* paths, int n, fy_wav_info* out, int n_threads);
int fy_wav_read_s16(const char* const* paths, const int64_t* data_offset, const int64_t* n_samples, int16_t* const* dst,
                    int n, int n_threads, int32_t* bad_index);

/* ---- A2: torchaudio.transforms.Resample(orig, 16000) ------------------------------------------------------
 * tools/extract_embedding.py:27-28, tools/extract_speech_token.py:29-30, cosyvoice/utils/file_utils.py:47-49.
 * One launch resamples a packed batch that shares one input rate.  Output length per utterance is
 * ceil(new * n_in / orig) (after dividing both rates by their gcd); out_off must hold those prefix sums and
 * total_out is their last entry (host value).  fy_resample_prepare builds the windowed-sinc filter bank of a rate
 * pair (torchaudio rebuilds it inside every Resample(...) construction) and must have been called once for it. */
int fy_resample_prepare(fy_ctx* ctx, int orig_freq, int new_freq);
int fy_resample(fy_ctx* ctx, const float* in, const int32_t* in_off, float* out, const int32_t* out_off, int n_utts,
                int total_out, int orig_freq, int new_freq);

/* ---- A3 + A4: kaldi.fbank(num_mel_bins=80, dither=0, sample_frequency=16000) then feat - feat.mean(0) ------
 * tools/extract_embedding.py:29-33, cosyvoice/cli/frontend.py:110-114.
 * pcm: packed mono samples (int16, scaled by 1/32768 on the device -- row A1 -- or float32).
 * feat: packed [sum M_b, 80] fp32, M_b = 1 + (N_b - 400) / 160 (0 when N_b < 400); frame_off = prefix sums of M_b.
 * ws: at least fy_fbank_workspace_bytes(n_utts, total_frames) bytes. */
size_t fy_fbank_workspace_bytes(int n_utts, int total_frames);
int fy_fbank_cmn(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                 int n_utts, int total_frames, float* feat, void* ws, size_t ws_bytes);

/* ---- fy_fbank_cmn without any allocation ------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments,
 * kernels and results as fy_fbank_cmn, minus the workspace: the scratch (per-CTA column sums + per-utterance means)
 * is a static __device__ array of FY_FBANK_STATIC_WS_BYTES inside the library, resident on every GPU that loads
 * it until the process exits; fy_create looks its device address up once per device.  The call makes no cudaMalloc /
 * cudaFree / allocating library call and launches three kernels on the context's stream.
 * Limit: fy_fbank_workspace_bytes(n_utts, total_frames) <= FY_FBANK_STATIC_WS_BYTES, i.e.
 * total_frames / 16 + 2 n_utts + 1 <= 13,106 rows (64 utterances of 30 s, or one prompt of any length the path
 * takes); beyond it the call fails with FY_ERR_NOMEM and a message naming both sizes -- use fy_fbank_cmn then.
 * All calls on one GPU share the array and it keeps the previous call's contents (nothing needs clearing: a call
 * writes every element before it reads it), so calls that OVERLAP on the GPU -- two streams, two contexts -- race;
 * serialise them or give the concurrent one a workspace through fy_fbank_cmn. */
#define FY_FBANK_STATIC_WS_BYTES (4u << 20)
int fy_fbank_cmn_static(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                        int n_utts, int total_frames, float* feat);

/* ---- A5: whisper.log_mel_spectrogram(audio, n_mels=128) ---------------------------------------------------
 * tools/extract_speech_token.py:38, cosyvoice/cli/frontend.py:98.
 * mel: packed; utterance b occupies 128 * T_b floats starting at 128 * frame_off[b], laid out [128][T_b] exactly
 * like the reference's [1, 128, T] tensor.  T_b = N_b / 160.  max_frames = largest T_b in the batch (host value). */
size_t fy_logmel_workspace_bytes(int n_utts, int total_frames);
int fy_logmel(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
              int n_utts, int total_frames, int max_frames, float* mel, void* ws, size_t ws_bytes);

/* ---- N4: matcha.utils.audio.mel_spectrogram at 24 kHz --------------------------------
 * The `feat_extractor` of cosyvoice/cli/frontend.py:120-125 (_extract_speech_feat) and of the training loader
 * (cosyvoice/dataset/processor.py:156-181 compute_fbank; conf/cosyvoice3.yaml:140-151): n_fft 1920, hop 480,
 * win 1920, 80 mels, fmin 0, fmax None, center=False (third_party/Matcha-TTS/matcha/utils/audio.py:44-88).
 * pcm: packed mono 24 kHz samples (int16 or float32), every utterance longer than 720 samples (reflect padding).
 * mel: packed time-major [sum T_b, 80] fp32 -- what both callers use after their transpose(0, 1);
 * T_b = 1 + (N_b - 480) / 480; frame_off = prefix sums of T_b. */
int fy_matcha_mel(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                  int n_utts, int total_frames, float* mel);

/* ---- A6: ort_session.run(None, {"input": feat[None]}) on campplus.onnx ------------------------------------
 * tools/extract_embedding.py:36-41, cosyvoice/cli/frontend.py:115-116.
 * feat: packed [sum M_b, 80] fp32 (output of fy_fbank_cmn); emb: [n_utts, 192] fp32. */
size_t fy_campplus_workspace_bytes(fy_ctx* ctx, int n_utts, int total_frames);
int fy_campplus_forward(fy_ctx* ctx, const float* feat, const int32_t* frame_off, const int32_t* h_frame_off,
                        int n_utts, float* emb, void* ws, size_t ws_bytes);

/* Test aid: byte offsets of CAMPPlus activation buffers inside the workspace (see fy_campplus.cu). */
int fy_campplus_layout(fy_ctx* ctx, const int32_t* h_frame_off, int n_utts, int64_t* out, int n_out);

/* ---- A9 + A10: ort_session.run(None, {feat, feat_len}) on speech_tokenizer_v3.onnx ------------------------
 * tools/extract_speech_token.py:40-48, cosyvoice/cli/frontend.py:99-103.
 * mel: packed as written by fy_logmel.  tokens: packed int32, L_b = ((T_b - 1) / 2 + 1 - 1) / 2 + 1 per utterance;
 * tok_off = prefix sums of L_b.  h_frame_off is the host copy of frame_off (needed to plan the launch).
 * fsq_pre (optional, may be NULL): packed [sum L_b, 8] fp32 pre-rounding values, for boundary-distance checks. */
size_t fy_s3tok_workspace_bytes(fy_ctx* ctx, int n_utts, int total_frames);
int fy_s3tok_encode_fsq(fy_ctx* ctx, const float* mel, const int32_t* frame_off, const int32_t* h_frame_off,
                        int n_utts, int32_t* tokens, const int32_t* tok_off, float* fsq_pre, void* ws,
                        size_t ws_bytes);

/* ---- A7: per-speaker mean of utterance embeddings (tools/extract_embedding.py:50-56) ----------------------
 * spk_id[b] in [0, n_spk).  sum: [n_spk, 192] fp32 and count: [n_spk] int32 are ACCUMULATED into (zero them first),
 * so partial sums from several batches / ranks can be combined (all_reduce SUM) before fy_speaker_mean_finish. */
int fy_speaker_sum(fy_ctx* ctx, const float* emb, const int32_t* spk_id, int n_utts, int n_spk, float* sum,
                   int32_t* count);
int fy_speaker_mean_finish(fy_ctx* ctx, const float* sum, const int32_t* count, int n_spk, float* mean);

/* ---- measurement aid: CUDA-event timing per kernel class on the launch stream ----------------------------
 * out[0..2] ms (tcgen05 GEMM, attention, other), out[3..5] algorithmic FLOPs, out[6..8] timed launches,
 * out[9] all kernel launches since the previous read.  fy_profile_read synchronises the stream. */
int fy_profile_enable(fy_ctx* ctx, int on);
int fy_profile_read(fy_ctx* ctx, double* out, int n_out);

/* ---- generic bf16 tensor-core GEMM (exposed for tests and tuning) -----------------------------------------
 * out[M, N] (fp32) = A[M, K] * B[N, K]^T with A, B bf16 row-major. */
int fy_gemm_bf16(fy_ctx* ctx, const void* a_bf16, const void* b_bf16, float* out, int M, int N, int K);

#ifdef __cplusplus
}
#endif
#endif /* FANGYAN_B200_H_ */
