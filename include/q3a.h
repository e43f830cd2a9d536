/* libq3a — B200 (sm_100a) log-mel frontend + Qwen3-ASR / Qwen3-ForcedAligner audio-encoder forward.
 *
 * This is the drop-in boundary for the hot path of qwang-sj/qwen3-asr.cpp.  The reference has no
 * FFI: the path is reached through two C++ headers (reference src/mel_spectrogram.h:42-55 and
 * src/audio_encoder.h:20-41).  Every entry point below names the reference interface it replaces;
 * the C++ classes with the reference's own names and signatures are header-only shims over this
 * C ABI (include/q3a/mel_spectrogram.h, include/q3a/audio_encoder.h).
 *
 * Conventions: plain pointers and sizes only; every function returns Q3A_OK (0) or an error code,
 * and q3a_last_error() returns the message for the calling thread.  There is no CPU fallback: every
 * compute entry point fails with Q3A_ERR_CUDA when no sm_100 device is usable.
 *
 * Layouts (unchanged from the reference):
 *   PCM      float32 mono 16 kHz in [-1,1]                       (mel_spectrogram.h:32-34)
 *   filters  float32 [n_mel=128][n_bins=201] row-major           (mel_spectrogram.h:26-30)
 *   mel      float32 [n_mel=128][n_len] mel-major                (mel_spectrogram.h:18-23)
 *   features float32 [n_tokens][out_dim=1024] row-major          (audio_encoder.cpp:591-595,
 *                                                                 audio_injection.h:59,74)
 */
#ifndef Q3A_H
#define Q3A_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    Q3A_OK = 0,
    Q3A_ERR_ARG = 1,    /* bad argument (null pointer, negative size, mel bins mismatch ...) */
    Q3A_ERR_CUDA = 2,   /* CUDA runtime / driver error, or no usable sm_100 device */
    Q3A_ERR_IO = 3,     /* file could not be opened / read */
    Q3A_ERR_FORMAT = 4, /* not a GGUF v2/v3 file, missing tensor, unsupported tensor type */
    Q3A_ERR_STATE = 5   /* object not loaded */
};

/* MMA operand element type for the encoder's tensor-core kernels (accumulation is always fp32). */
enum { Q3A_OP_BF16 = 0, Q3A_OP_F16 = 1 };

/* Stem / attention behaviour of the two encoders in the reference. */
enum {
    Q3A_MODE_ASR = 0,      /* AudioEncoder::encode: 100-frame chunks at their own length, PE per chunk,
                              full attention over the utterance (audio_encoder.cpp:312-601) */
    Q3A_MODE_ALIGNER = 1,  /* ForcedAligner::encode_audio: chunks zero-padded to 100 frames, valid tokens
                              kept, block-diagonal attention in windows of 104 tokens
                              (forced_aligner.cpp:591-924) */
    Q3A_MODE_NO_CHUNK = 2, /* AudioEncoder::encode_no_chunk: one chunk = whole utterance
                              (audio_encoder.cpp:603-852) */
    Q3A_MODE_CONV_ONLY = 3 /* AudioEncoder::encode_conv_only: stem only, no chunking, no PE, no layers;
                              output is [n_tokens][d_model] (audio_encoder.cpp:854-913) */
};

typedef struct q3a_frontend q3a_frontend; /* resident filterbank + window + FFT tables on one GPU */
typedef struct q3a_encoder q3a_encoder;   /* resident weights + workspace on one GPU */

/* mirrors audio_encoder_hparams + text_decoder_hparams::hidden_size (gguf_loader.h:15-35) */
typedef struct q3a_hparams {
    int32_t n_encoder_layers;
    int32_t d_model;
    int32_t n_attention_heads;
    int32_t ffn_dim;
    int32_t conv_channels;
    int32_t conv_out_dim;
    int32_t n_mel_bins;
    int32_t n_window_infer;
    float layer_norm_eps;
    int32_t out_dim;      /* text hidden size the encoder projects to (1024) */
    int32_t weight_ftype; /* 0 = F32, 1 = F16, 8 = Q8_0 (type of the 2-D Linear weights in the file) */
} q3a_hparams;

/* per-stage device time of the most recent encode call, milliseconds (CUDA events on the compute
 * stream; collected only after q3a_encoder_set_profiling(enc, 1)).  Labels follow the reference's
 * QWEN3_TIMER names (qwen3_asr.cpp:89,104; audio_encoder.cpp:349,574). */
typedef struct q3a_timings {
    float mel_ms;             /* "mel_spectrogram" */
    float conv_chunk_ms;      /* "audio_encoding.conv_chunk": conv1..3 + conv_out (+PE) */
    float transformer_ms;     /* "audio_encoding.transformer": layers + ln_post + proj1/2 */
    float gemm_ms;            /* all tcgen05 GEMM launches (implicit-GEMM convs and Linear layers) */
    float attention_ms;       /* all attention launches */
    float other_ms;           /* conv1, LayerNorm, packing */
    double gemm_flops;        /* algorithmic FLOPs (2*MAC, no padding) of the GEMM launches */
    double attention_flops;   /* 4*hd*H*sum(seg^2) per layer */
    int32_t gemm_launches;
    int32_t attention_launches;
    int32_t total_launches;   /* all kernels of ours launched by the call */
} q3a_timings;

const char* q3a_last_error(void);
int q3a_version(void);
/* number of usable sm_100 devices (0 when there is no driver/GPU) */
int q3a_device_count(void);

/* ---- log-mel frontend ------------------------------------------------------------------ */

/* replaces generate_mel_filters (mel_spectrogram.h:42-43, mel_spectrogram.cpp:361-415).  Host only.
 * out: [n_mels][n_fft/2+1]. */
int q3a_generate_mel_filters(float* out, int n_mels, int n_fft, int sample_rate);
/* n_len the frontend produces for a clip (mel_spectrogram.cpp:517-520): n_samples / 160 */
int q3a_mel_n_len(int n_samples);

/* filters: host [n_mel][n_bins=201] or NULL for the generated bank; n_mel <= 128. */
int q3a_frontend_create(int device, const float* filters, int n_mel, int n_bins, q3a_frontend** out);
void q3a_frontend_free(q3a_frontend* fe);

/* replaces log_mel_spectrogram (mel_spectrogram.h:53-55, mel_spectrogram.cpp:484-628) for a batch of
 * clips in HOST memory: out[b] receives n_mel * n_len[b] floats, mel-major.  n_len may be NULL. */
int q3a_mel_batch(q3a_frontend* fe, const float* const* pcm, const int* n_samples, int n_clips,
                  float* const* out, int* n_len);

/* device-resident variant.  pcm_dev: packed clips, clip b starts at sample pcm_off[b] (host array of
 * n_clips entries, each a multiple of 4).  mel_dev: clip b's [n_mel][n_len_b] block starts at float
 * offset mel_off[b] (host array).  stream: a cudaStream_t (NULL = default stream).
 * finalize = 1: full reference semantics (clamp to max-8, (x+4)/4).  finalize = 0: leaves
 * log10(max(E,1e-10)) in mel_dev and the per-clip maxima in max_dev (device float[n_clips], may be
 * NULL when finalize=1) for a consumer that fuses the clamp (the encoder's conv1). */
int q3a_mel_batch_device(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off, const int* n_samples,
                         int n_clips, float* mel_dev, const int64_t* mel_off, float* max_dev, int finalize,
                         void* stream);

/* Allocation-free variant of q3a_mel_batch_device.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results (bit-identical); the call performs no cudaMalloc / cudaFree
 * and no host allocation: its scratch (clip descriptor table and the per-clip maximum slots) is a static __device__
 * array inside the library (4.5 KB, held by every GPU that loads the library until the process exits), which limits a
 * call to q3a_mel_noalloc_max_clips() = 128 clips; more is rejected with Q3A_ERR_ARG and a message.  All calls on a
 * device share that array and it keeps the last call's contents (each call rewrites the table and clears the maxima
 * itself): calls in one stream are ordered and safe, calls that overlap on one GPU (two streams, two host threads)
 * race.  The first call on a device resolves the array's device address (cudaGetSymbolAddress; not an allocation, but
 * not capturable either: make one call before starting a capture). */
int q3a_mel_batch_device_noalloc(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off, const int* n_samples,
                                 int n_clips, float* mel_dev, const int64_t* mel_off, float* max_dev, int finalize,
                                 void* stream);
int q3a_mel_noalloc_max_clips(void);

/* ---- audio encoder --------------------------------------------------------------------- */

/* replaces AudioEncoder::load_model (audio_encoder.h:25, audio_encoder.cpp:42-83) and the aligner's
 * private loader (forced_aligner.cpp:177-475): reads the audio.encoder.* tensors of a GGUF file
 * (F32 / F16 / Q8_0), takes hyper-parameters from the qwen3-asr.* keys the converter writes
 * (convert_hf_to_gguf.py:441-447) and falls back to tensor shapes / the reference defaults.
 * Supported shapes (both reference presets fit; anything else fails with Q3A_ERR_FORMAT): 128 mel bins,
 * head_dim 64 (d_model = 64 * heads), d_model <= 1024, every dimension a multiple of 8, n_window_infer = 800
 * (104-token windows); layer_norm_eps and n_window_infer are read from "audio.layer_norm_eps" /
 * "audio.n_window_infer" like the reference's loader does (gguf_loader.cpp:76-77), default 1e-5 / 800.
 * Every tensor the reference's graph uses must be present (all biases, ln_post, proj1/2): the reference treats them
 * as optional, this loader does not.  Q8_0 files: the Linear weights are expanded to the 16-bit operand type at load (d * q,
 * rounded once); with Q3A_Q8_FUSED=1 in the environment they stay Q8_0 blocks in device memory and the GEMM's dequant
 * warps expand them tile by tile -- the same bits, half the weight memory, about 1.6x the time. */
int q3a_encoder_load(const char* gguf_path, int device, int operand_dtype, q3a_encoder** out);
void q3a_encoder_free(q3a_encoder* enc);
int q3a_encoder_get_hparams(const q3a_encoder* enc, q3a_hparams* out);
/* tokens the stem emits for n_frames mel frames in a mode (audio_encoder.cpp:304-345) */
int q3a_encoder_n_tokens(int n_frames, int mode);

/* replaces AudioEncoder::encode / encode_no_chunk / encode_conv_only (audio_encoder.h:27-36) and
 * ForcedAligner::encode_audio (forced_aligner.h:241-242) for a batch of utterances in HOST memory.
 * mel[b]: [n_mel][n_frames[b]] mel-major; out[b] receives n_tokens[b] * out_dim floats (d_model for
 * Q3A_MODE_CONV_ONLY).  n_tokens may be NULL. */
int q3a_encoder_encode_batch(q3a_encoder* enc, const float* const* mel, int n_mel, const int* n_frames,
                             int n_utts, int mode, float* const* out, int* n_tokens);

/* device-resident variant: mel_dev packed as produced by q3a_mel_batch_device (finalize=1), or raw
 * log-mel plus max_dev (finalize=0; the clamp and scale are then applied inside conv1's loads).
 * out_dev: [sum n_tokens][out_dim] f32, utterances in order.  tok_off (host, n_utts+1, may be NULL)
 * receives the row offsets.  mel_dev / out_dev must be 16-byte aligned, mel_off entries multiples of 4.
 * Asynchronous on `stream`; the host arrays may be reused as soon as the call returns. */
int q3a_encoder_encode_device(q3a_encoder* enc, const float* mel_dev, const int64_t* mel_off,
                              const float* max_dev, const int* n_frames, int n_utts, int mode,
                              float* out_dev, int64_t* tok_off, void* stream);

/* the whole hot path, host PCM -> host features: Qwen3ASR::transcribe_internal's mel + encode steps
 * (qwen3_asr.cpp:90,105) / ForcedAligner::align's (forced_aligner.cpp:1653,1661) for a batch. */
int q3a_encode_pcm_batch(q3a_frontend* fe, q3a_encoder* enc, const float* const* pcm, const int* n_samples,
                         int n_clips, int mode, float* const* out, int* n_tokens);
/* asynchronous twin: submit enqueues the call and returns a ticket, q3a_encode_wait blocks until that call (and every
 * earlier one) has delivered its features.  Up to two calls overlap: call k+1 uploads and runs its stem under call k's last
 * layers and download.  pcm[i] / out[i] must stay valid until the wait returns; n_tokens is final when submit returns. */
int q3a_encode_pcm_submit(q3a_frontend* fe, q3a_encoder* enc, const float* const* pcm, const int* n_samples,
                          int n_clips, int mode, float* const* out, int* n_tokens, int64_t* ticket);
int q3a_encode_wait(q3a_encoder* enc, int64_t ticket);
/* same from 16-bit PCM as it sits in a WAV file: load_wav's conversion (mel_spectrogram.cpp:195-210: interleaved
 * channels averaged in float, then / 32768) runs on the device, so the upload is 2 bytes per sample.  n_samples[i]
 * counts frames (samples per channel); pcm[i] holds n_samples[i] * n_channels values; n_channels in 1..8.
 * Results are bit-identical to q3a_encode_pcm_batch on load_wav's float output. */
int q3a_encode_pcm16_batch(q3a_frontend* fe, q3a_encoder* enc, const int16_t* const* pcm, const int* n_samples,
                           int n_channels, int n_clips, int mode, float* const* out, int* n_tokens);
/* same with PCM and features resident on the device (packed as above): pcm_dev / out_dev 16-byte aligned, pcm_off
 * entries multiples of 4.  Asynchronous on `stream`.  This is the hand-off a GPU-resident decoder consumes: the rows are
 * the memory text_decoder.cpp:422-456,650-655 copies into inp_audio. */
int q3a_encode_pcm_device(q3a_frontend* fe, q3a_encoder* enc, const float* pcm_dev, const int64_t* pcm_off,
                          const int* n_samples, int n_clips, int mode, float* out_dev, int64_t* tok_off,
                          void* stream);

int q3a_encoder_set_profiling(q3a_encoder* enc, int on);
int q3a_encoder_get_timings(const q3a_encoder* enc, q3a_timings* out);
/* bytes of device memory held by the encoder (weights + workspace) */
int64_t q3a_encoder_device_bytes(const q3a_encoder* enc);

/* ---- forced aligner: NAR decoder prefill (csrc/decoder.cu) ------------------------------------------------------------
 * replaces ForcedAligner::forward_decoder / build_decoder_graph (forced_aligner.cpp:926-1169): one causal forward over the
 * prompt -- token embeddings with the audio features spliced in at [audio_start, audio_start + n_audio) -- through the
 * Qwen3 decoder layers and the classification head: logits [n_tokens][n_classes] f32 row-major, the memory
 * extract_timestamp_classes (forced_aligner.cpp:1279-1308) reads.  Weights: token_embd / blk.N.* / output_norm / output of
 * the same GGUF the aligner's encoder is loaded from (hyper-parameter keys of parse_hparams, :158-168).  head_dim must
 * be 128.  The _device variant takes the audio features where q3a_encode_pcm_device left them (device memory, f32
 * [n_audio][hidden]) and leaves the logits on the device, asynchronously on `stream`; tokens are host memory. */
typedef struct q3a_decoder q3a_decoder;
typedef struct q3a_decoder_hparams {
    int32_t n_layers, hidden, n_heads, n_kv_heads, head_dim, intermediate, vocab, n_classes;
    float rms_eps, rope_theta;
} q3a_decoder_hparams;
int q3a_decoder_load(const char* gguf_path, int device, int operand_dtype, q3a_decoder** out);
void q3a_decoder_free(q3a_decoder* dec);
int q3a_decoder_get_hparams(const q3a_decoder* dec, q3a_decoder_hparams* out);
int q3a_decoder_forward(q3a_decoder* dec, const int32_t* tokens, int n_tokens, const float* audio_embd, int n_audio,
                        int audio_start, float* logits);
int q3a_decoder_forward_device(q3a_decoder* dec, const int32_t* tokens, int n_tokens, const float* audio_dev, int n_audio,
                               int audio_start, float* logits_dev, void* stream);

/* ---- utterance sharding (no collective on the compute path) ------------------------------ */

/* Longest-processing-time assignment of utterances to n_shards replicas by predicted cost
 * a*T + b*sum(seg^2); shard_of[i] receives the shard of utterance i.  Deterministic. Host only. */
int q3a_plan_shards(const int* n_frames, int n_utts, int mode, int n_shards, int* shard_of);

/* ---- in-process multi-GPU executor (csrc/pool.cu) ------------------------------------------------------------
 * The reference is a single process that encodes one utterance at a time (qwen3_asr.cpp:81-149, main.cpp:294-504).  A
 * pool holds one replica (frontend + encoder) and one host worker thread per listed device; q3a_pool_encode_pcm is
 * q3a_encode_pcm_batch for a whole corpus: utterances are assigned to the replicas by q3a_plan_shards, every replica
 * encodes its shard with its own streams, features land in out[i] -- utterance order -- and are bit-identical to what
 * any single replica computes for that utterance.  No collective, no device-to-device traffic.  shard_of (n_clips)
 * and shard_ms (n_devices: wall time of every replica's call) may be NULL.  A device may be listed more than once. */
typedef struct q3a_pool q3a_pool;
int q3a_pool_create(const char* gguf_path, const int* devices, int n_devices, int operand_dtype, q3a_pool** out);
void q3a_pool_free(q3a_pool* pool);
int q3a_pool_n_devices(const q3a_pool* pool);
int q3a_pool_encode_pcm(q3a_pool* pool, const float* const* pcm, const int* n_samples, int n_clips, int mode,
                        float* const* out, int* n_tokens, int* shard_of, float* shard_ms);
/* Host placement helpers for callers of the host-pointer entry points: pinned (portable) staging memory, and binding
 * the calling thread to the CPUs local to a GPU (from /sys/bus/pci/devices/<bus id>/local_cpulist; returns the number
 * of CPUs bound, 0 if nothing was changed).  The pool's workers bind themselves. */
int q3a_host_alloc(size_t bytes, void** out);
void q3a_host_free(void* p);
int q3a_bind_thread_to_device(int device);

/* How a q3a_encode_pcm_batch call over these clips is scheduled: pass_of[i] (a pass bounds the stem workspace;
 * consecutive passes overlap as upload | compute | download), in_part_of[i] (upload part inside the pass: the stem of a
 * part starts as soon as its PCM has landed) and out_part_of[i] (download part: its rows travel back while the next
 * part runs its last layers).  out_width = 1024 (or d_model for Q3A_MODE_CONV_ONLY).  Host only; results never
 * depend on the schedule. */
int q3a_plan_host_call(const int* n_samples, int n_clips, int mode, int out_width, int* pass_of, int* in_part_of,
                       int* out_part_of);

/* ---- unit-test hooks for the tensor-core building blocks (device pointers) ---------------- */

/* D[M][N] (f32) = A[M][K] * W[N][K]^T (+bias[N]) ; A, W are 16-bit (operand_dtype) row-major.
 * K % 8 == 0, N % 8 == 0. */
int q3a_test_gemm(const void* a_dev, const void* w_dev, const float* bias_dev, float* d_dev, int M, int N,
                  int K, int operand_dtype, int block_n, void* stream);
/* out[T][H*64] (16-bit) = softmax(q k^T / 8) v inside each segment; qkv: [T][3*H*64] 16-bit with q, k, v
 * column blocks; cu_seqlens: host int[n_seg+1]. */
int q3a_test_attention(const void* qkv_dev, void* out_dev, int T, int H, const int* cu_seqlens, int n_seg,
                       int operand_dtype, void* stream);
/* same with the kernel generation chosen explicitly: 1 = one query tile per CTA (attention.cu), 2 = persistent CTAs with
 * two ping-ponging query tiles (attention2.cu; correct, measured slower, opt-in with Q3A_ATTN_V2=1), 0 = the encoder's
 * choice (generation 1 unless that switch is set). */
int q3a_test_attention_v(const void* qkv_dev, void* out_dev, int T, int H, const int* cu_seqlens, int n_seg,
                         int operand_dtype, int version, void* stream);

/* Development probe (csrc/probe.cu): cycles that n_mma back-to-back tcgen05.mma (cta_group::1, M = 128, K = 16, bf16) of
 * width N take on each of `grid` CTAs, A from shared memory (a_from_tmem bit 0 clear) or TMEM (bit 0 set), with M = 64
 * instead of 128 when bit 1 is set, accumulating into one TMEM region or alternating between two; cycles_out_dev:
 * device int64[grid]. */
int q3a_probe_mma_rate(int N, int a_from_tmem, int n_mma, int alternate, int k_steps_per_commit, int grid,
                       long long* cycles_out_dev, void* stream);

/* The library's own bounds check (compute-sanitizer's stand-in where that tool is not available).  With Q3A_GUARD=1
 * in the environment when the process first allocates, every device buffer the library owns (weights, activations,
 * staging, descriptor tables) is exact-size and fenced by two 4-KiB pattern bands.  This call synchronises the
 * current device, reads every live band back and reports how many exist and how many were written to;
 * q3a_last_error() then names the damaged buffers.  Q3A_ERR_STATE if the mode is off. */
int q3a_debug_check_guards(int* n_bands, int* n_damaged);
/* Self-test of the detector: writes one byte just past the end of one live guarded buffer, so that the next
 * q3a_debug_check_guards reports exactly one damaged band. */
int q3a_debug_damage_guard(void);

#ifdef __cplusplus
}
#endif
#endif /* Q3A_H */
