// Forced aligner: the non-autoregressive decoder prefill on sm_100a.
//
// Replaces ForcedAligner::build_decoder_graph / forward_decoder (reference forced_aligner.cpp:926-1169): one causal
// forward over the whole prompt -- text tokens with the audio encoder's [n_audio][1024] features spliced in at the
// audio_pad positions (:964-993) -- through 28 Qwen3 layers (RMSNorm, q/k/v, per-head q/k RMSNorm, NeoX RoPE, causal
// grouped-query attention, output projection, RMSNorm, SwiGLU) and the classification head that turns every position
// into 5000 timestamp classes (:1067-1076).  The audio features are taken from device memory, so the encoder's output
// (q3a_encode_pcm_device) feeds it without leaving HBM: the consumer of row f1.
//
// Launch sequence per layer (all weight matmuls on the tcgen05 GEMM of gemm_tc.cuh):
//   QKV GEMM      A = residual stream rounded to 16 bits, RMSNorm folded in (gamma in the weights, rstd in the epilogue)
//   qk_norm_rope  per (token, head): RMSNorm over the 128 dimensions, weight, NeoX rotation -- in place, fp32 arithmetic
//   attention     attention_dec.cu (causal, 16 query heads on 8 key/value heads)
//   O GEMM        + residual, emits the 16-bit copy and the row statistics for the next fold
//   gate|up GEMM  RMSNorm folded in, one GEMM over the concatenated weights
//   swiglu        silu(gate) * up
//   down GEMM     + residual, 16-bit copy + statistics
// then the head GEMM with the final RMSNorm folded in, f32 logits.
#include "gemm.h"
#include "attention.h"
#include "weights.h"

#include <algorithm>

#include <cmath>
#include <cstring>

struct q3a_decoder {
    int device = 0;
    int fmt = q3a::OP_BF16;
    q3a_decoder_hparams hp{};
    q3a::DevBuf arena;
    size_t arena_used = 0;
    const void* tok_embd = nullptr;            // [vocab][hidden] 16-bit
    struct Layer {
        const void *w_qkv = nullptr, *w_o = nullptr, *w_gu = nullptr, *w_down = nullptr;   // 16-bit, norms folded into qkv / gu
        const float *q_norm = nullptr, *k_norm = nullptr;                                  // [head_dim]
    };
    std::vector<Layer> layers;
    const void* w_head = nullptr;              // [n_classes][hidden], output_norm folded in
    const float* b_head = nullptr;             // [n_classes] (zeros when the file has none)
    const float* zeros = nullptr;              // max(N) zeros: bias' and colsum of the RMS folds
    q3a::DevBuf x, x16, stats, qkv, att, gu, act, tok, audio_stage, logits_stage;
    cudaStream_t stream = nullptr;
};

namespace q3a {

// ---- CUDA-core kernels around the GEMMs -------------------------------------------------------------------------
// x[m] = m in [audio_start, audio_start + n_audio) ? audio[m - audio_start] : float(embd[token[m]])   (forced_aligner.cpp:964-993)
// plus what the first folded RMSNorm needs: the 16-bit copy and (sum, sum of squares) in partial 0, zeros in the others.
template <class OpT>
__global__ void __launch_bounds__(256)
dec_embed_kernel(const int* __restrict__ tokens, const OpT* __restrict__ embd, const float* __restrict__ audio, int n_tokens,
                 int n_audio, int audio_start, int hidden, int vocab, float* __restrict__ x, OpT* __restrict__ x16,
                 float2* __restrict__ stats, int n_part) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n_tokens) return;
    const int lane = threadIdx.x & 31;
    const bool is_audio = audio != nullptr && row >= audio_start && row < audio_start + n_audio;
    int t = __ldg(tokens + row);
    t = t < 0 ? 0 : (t >= vocab ? vocab - 1 : t);
    const OpT* er = embd + (size_t)t * hidden;
    const float* ar = audio + (size_t)(row - audio_start) * hidden;
    float sum = 0.0f, sq = 0.0f;
    for (int i = lane * 2; i < hidden; i += 64) {
        float2 v;
        if (is_audio) v = *reinterpret_cast<const float2*>(ar + i);
        else v = OpTraits<OpT>::unpack2(*reinterpret_cast<const uint32_t*>(er + i));
        *reinterpret_cast<float2*>(x + (size_t)row * hidden + i) = v;
        *reinterpret_cast<uint32_t*>(x16 + (size_t)row * hidden + i) = OpTraits<OpT>::pack2(v.x, v.y);
        sum += v.x + v.y; sq = fmaf(v.x, v.x, fmaf(v.y, v.y, sq));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); sq += __shfl_xor_sync(0xffffffffu, sq, o); }
    if (lane < n_part) stats[(size_t)row * n_part + lane] = lane == 0 ? make_float2(sum, sq) : make_float2(0.0f, 0.0f);
}

// In place on the q and k column blocks of qkv [n][(H + 2 Hkv) * 128]: per (token, head) RMSNorm over the head's 128
// dimensions times the per-dimension weight (forced_aligner.cpp:1022-1030), then ggml_rope_ext NEOX with position = token
// index (:1032-1038, :1119-1125): pairs (i, i + 64), angle = pos * theta^(-2i/128).  One warp per (token, head), each lane
// owns dimensions {lane, lane + 32} and their partners {lane + 64, lane + 96}.
template <class OpT>
__global__ void __launch_bounds__(256)
dec_qk_norm_rope_kernel(OpT* __restrict__ qkv, const float* __restrict__ q_norm, const float* __restrict__ k_norm, int n_tokens,
                        int n_heads, int n_kv_heads, float eps, float rope_theta) {
    pdl_trigger();
    pdl_wait();
    const int hq = n_heads + n_kv_heads;                         // heads to treat per token (q heads then k heads)
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (long long)n_tokens * hq) return;
    const int tok = (int)(w / hq), h = (int)(w - (long long)tok * hq);
    const int lane = threadIdx.x & 31;
    const int ld = (n_heads + 2 * n_kv_heads) * 128;
    OpT* p = qkv + (size_t)tok * ld + (size_t)h * 128;           // q heads are columns [0, H*128), k heads follow directly
    const float* wn = h < n_heads ? q_norm : k_norm;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = OpTraits<OpT>::to(p[lane + 32 * j]);
    float sq = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = 1.0f / sqrtf(sq * (1.0f / 128.0f) + eps);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = v[j] * rstd * __ldg(wn + lane + 32 * j);
#pragma unroll
    for (int j = 0; j < 2; ++j) {                                // pair (i, i + 64), i = lane + 32 j
        const int i = lane + 32 * j;
        const float ang = (float)tok * powf(rope_theta, -(float)(2 * i) / 128.0f);
        float sn, cs;
        sincosf(ang, &sn, &cs);
        const float a = v[j], b = v[j + 2];
        p[i] = OpTraits<OpT>::from(a * cs - b * sn);
        p[i + 64] = OpTraits<OpT>::from(a * sn + b * cs);
    }
}

// act[m][i] = silu(gu[m][i]) * gu[m][inter + i]   (forced_aligner.cpp:1052-1059)
template <class OpT>
__global__ void __launch_bounds__(256)
dec_swiglu_kernel(const OpT* __restrict__ gu, OpT* __restrict__ act, long long n_pairs, int inter) {
    pdl_trigger();
    pdl_wait();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_pairs; e += stride) {
        const long long m = e / (inter / 2);
        const int i = (int)(e - m * (inter / 2)) * 2;
        const float2 g = OpTraits<OpT>::unpack2(*reinterpret_cast<const uint32_t*>(gu + m * 2 * inter + i));
        const float2 u = OpTraits<OpT>::unpack2(*reinterpret_cast<const uint32_t*>(gu + m * 2 * inter + inter + i));
        const float a = g.x / (1.0f + __expf(-g.x)) * u.x, b = g.y / (1.0f + __expf(-g.y)) * u.y;
        *reinterpret_cast<uint32_t*>(act + m * inter + i) = OpTraits<OpT>::pack2(a, b);
    }
}

}  // namespace q3a

using namespace q3a;

// ---- loading ---------------------------------------------------------------------------------------------------------
static int dec_load_weights(q3a_decoder* dc, const GgufFile& f) {
    std::string err;
    q3a_decoder_hparams& hp = dc->hp;
    // keys and defaults of ForcedAligner::parse_hparams (forced_aligner.cpp:158-168)
    hp.n_layers = (int)f.get_u32("qwen3-asr.block_count", 28);
    hp.hidden = (int)f.get_u32("qwen3-asr.embedding_length", 1024);
    hp.n_heads = (int)f.get_u32("qwen3-asr.attention.head_count", 16);
    hp.n_kv_heads = (int)f.get_u32("qwen3-asr.attention.head_count_kv", 8);
    hp.intermediate = (int)f.get_u32("qwen3-asr.feed_forward_length", 3072);
    hp.head_dim = (int)f.get_u32("qwen3-asr.attention.key_length", 128);
    hp.rms_eps = f.get_f32("qwen3-asr.attention.layer_norm_rms_epsilon", 1e-6f);
    hp.rope_theta = f.get_f32("qwen3-asr.rope.freq_base", 1000000.0f);
    hp.vocab = (int)f.get_u32("qwen3-asr.vocab_size", 152064);
    hp.n_classes = (int)f.get_u32("qwen3-asr.classify_num", 5000);
    const GgufTensor* t_emb = f.find("token_embd.weight");
    const GgufTensor* t_head = f.find("output.weight");
    if (!t_emb || !t_head || t_emb->ne.size() != 2 || t_head->ne.size() != 2) {
        set_error("GGUF file has no token_embd.weight / output.weight (not a forced-aligner file)");
        return Q3A_ERR_FORMAT;
    }
    hp.hidden = (int)t_emb->ne[0]; hp.vocab = (int)t_emb->ne[1]; hp.n_classes = (int)t_head->ne[1];
    const int d = hp.hidden, dq = hp.n_heads * hp.head_dim, dkv = hp.n_kv_heads * hp.head_dim, inter = hp.intermediate;
    if (hp.head_dim != 128 || hp.n_kv_heads <= 0 || hp.n_heads % hp.n_kv_heads || d % 8 || inter % 8 || hp.n_classes % 8 || d > 8192) {
        set_error("unsupported decoder shape (need head_dim 128, heads a multiple of kv heads, dimensions multiples of 8)");
        return Q3A_ERR_FORMAT;
    }
    ArenaPlan plan;
    std::vector<float> v, v2, v3, g, bias, cs, zero_beta(d, 0.0f);
#define FETCH(name, vec, n) do { if (!fetch(f, (name), vec, err, (n))) { set_error(err); return Q3A_ERR_FORMAT; } } while (0)
    FETCH("token_embd.weight", v, (int64_t)hp.vocab * d); plan.add_op(&dc->tok_embd, v, dc->fmt);
    dc->layers.assign(hp.n_layers, q3a_decoder::Layer{});
    for (int l = 0; l < hp.n_layers; ++l) {
        auto& L = dc->layers[l];
        const std::string B = "blk." + std::to_string(l) + ".";
        FETCH(B + "attn_norm.weight", g, d);
        FETCH(B + "attn_q.weight", v, (int64_t)dq * d);
        FETCH(B + "attn_k.weight", v2, (int64_t)dkv * d);
        FETCH(B + "attn_v.weight", v3, (int64_t)dkv * d);
        v.insert(v.end(), v2.begin(), v2.end()); v.insert(v.end(), v3.begin(), v3.end());
        bias.assign(dq + 2 * dkv, 0.0f);
        fold_layernorm(v, dq + 2 * dkv, d, g, zero_beta, bias, dc->fmt, v2, cs);      // gamma into the weights; beta = 0
        plan.add_op(&L.w_qkv, v, dc->fmt);
        FETCH(B + "attn_q_norm.weight", v, hp.head_dim); plan.add_f32(&L.q_norm, v);
        FETCH(B + "attn_k_norm.weight", v, hp.head_dim); plan.add_f32(&L.k_norm, v);
        FETCH(B + "attn_output.weight", v, (int64_t)d * dq); plan.add_op(&L.w_o, v, dc->fmt);
        FETCH(B + "ffn_norm.weight", g, d);
        FETCH(B + "ffn_gate.weight", v, (int64_t)inter * d);
        FETCH(B + "ffn_up.weight", v2, (int64_t)inter * d);
        v.insert(v.end(), v2.begin(), v2.end());
        bias.assign(2 * inter, 0.0f);
        fold_layernorm(v, 2 * inter, d, g, zero_beta, bias, dc->fmt, v2, cs);
        plan.add_op(&L.w_gu, v, dc->fmt);
        FETCH(B + "ffn_down.weight", v, (int64_t)d * inter); plan.add_op(&L.w_down, v, dc->fmt);
    }
    FETCH("output_norm.weight", g, d);
    FETCH("output.weight", v, (int64_t)hp.n_classes * d);
    bias.assign(hp.n_classes, 0.0f);
    if (f.find("output.bias")) FETCH("output.bias", bias, hp.n_classes);
    fold_layernorm(v, hp.n_classes, d, g, zero_beta, bias, dc->fmt, v2, cs);           // v2 = bias (beta is zero)
    plan.add_op(&dc->w_head, v, dc->fmt);
    plan.add_f32(&dc->b_head, v2);
#undef FETCH
    std::vector<float> zeros((size_t)std::max({dq + 2 * dkv, 2 * inter, hp.n_classes, d}), 0.0f);
    plan.add_f32(&dc->zeros, zeros);
    return upload_arena(plan, dc->arena, &dc->arena_used);
}

extern "C" void q3a_decoder_free(q3a_decoder* dc) {
    if (!dc) return;
    cudaSetDevice(dc->device);
    if (dc->stream) { cudaStreamSynchronize(dc->stream); cudaStreamDestroy(dc->stream); }
    for (DevBuf* b : {&dc->arena, &dc->x, &dc->x16, &dc->stats, &dc->qkv, &dc->att, &dc->gu, &dc->act, &dc->tok, &dc->audio_stage,
                      &dc->logits_stage})
        b->release();
    delete dc;
}

extern "C" int q3a_decoder_load(const char* gguf_path, int device, int operand_dtype, q3a_decoder** out) {
    if (!out) { set_error("decoder_load: out is null"); return Q3A_ERR_ARG; }
    *out = nullptr;
    if (!gguf_path) { set_error("decoder_load: path is null"); return Q3A_ERR_ARG; }
    if (operand_dtype != Q3A_OP_BF16 && operand_dtype != Q3A_OP_F16) { set_error("decoder_load: unknown operand dtype"); return Q3A_ERR_ARG; }
    GgufFile f;
    std::string err;
    if (!f.open(gguf_path, err)) {
        set_error(err);
        return (err.rfind("cannot open", 0) == 0 || err.rfind("cannot stat", 0) == 0 || err.rfind("mmap", 0) == 0) ? Q3A_ERR_IO : Q3A_ERR_FORMAT;
    }
    Q3A_TRY(use_device(device));
    q3a_decoder* dc = new q3a_decoder();
    dc->device = device; dc->fmt = operand_dtype;
    int rc = dec_load_weights(dc, f);
    if (rc == Q3A_OK && cudaStreamCreateWithFlags(&dc->stream, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("cudaStreamCreate failed"); rc = Q3A_ERR_CUDA;
    }
    if (rc != Q3A_OK) { q3a_decoder_free(dc); return rc; }
    *out = dc;
    return Q3A_OK;
}

extern "C" int q3a_decoder_get_hparams(const q3a_decoder* dc, q3a_decoder_hparams* out) {
    if (!dc) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (!out) { set_error("decoder_get_hparams: out is null"); return Q3A_ERR_ARG; }
    *out = dc->hp;
    return Q3A_OK;
}

// ---- forward ---------------------------------------------------------------------------------------------------------
template <class OpT>
static int dec_forward_core(q3a_decoder* dc, const int32_t* tokens_host, int n, const float* audio_dev, int n_audio,
                            int audio_start, float* logits_dev, cudaStream_t st) {
    const q3a_decoder_hparams& hp = dc->hp;
    const int d = hp.hidden, H = hp.n_heads, Hkv = hp.n_kv_heads, inter = hp.intermediate;
    const int dq = H * 128, dqkv = (H + 2 * Hkv) * 128;
    const int n_part = ln_partials(d);
    Q3A_TRY(dc->tok.reserve(sizeof(int32_t) * (size_t)n));
    Q3A_TRY(dc->x.reserve((size_t)n * d * 4));
    Q3A_TRY(dc->x16.reserve((size_t)n * d * 2));
    Q3A_TRY(dc->stats.reserve((size_t)n * n_part * sizeof(float2)));
    Q3A_TRY(dc->qkv.reserve((size_t)n * dqkv * 2));
    Q3A_TRY(dc->att.reserve((size_t)n * dq * 2));
    Q3A_TRY(dc->gu.reserve((size_t)n * 2 * inter * 2));
    Q3A_TRY(dc->act.reserve((size_t)n * inter * 2));
    // pageable source: the runtime stages the tokens before the call returns
    Q3A_CUDA(cudaMemcpyAsync(dc->tok.p, tokens_host, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    float* x = dc->x.as<float>();
    OpT* x16 = dc->x16.as<OpT>();
    float2* stats = dc->stats.as<float2>();
    const bool splice = audio_dev && n_audio > 0 && audio_start >= 0 && audio_start + n_audio <= n;   // forced_aligner.cpp:966
    Q3A_CUDA(launch_pdl(dec_embed_kernel<OpT>, dim3(ceil_div(n, 8)), dim3(256), 0, st, dc->tok.as<int>(),
                        static_cast<const OpT*>(dc->tok_embd), splice ? audio_dev : (const float*)nullptr, n, n_audio, audio_start, d,
                        hp.vocab, x, x16, stats, n_part));
    ++g_launch_counter;
    auto linear = [&](EpiKind kind, const void* A, int K, const void* W, int N, void* out, const float* bias, bool residual) -> int {
        LinearArgs a;
        a.a = A; a.M = n; a.K = K; a.lda = K; a.w = W; a.N = N; a.ldw = K; a.out = out; a.ldc = N;
        a.bias = bias; a.res = residual ? static_cast<const float*>(out) : nullptr;
        a.out16 = x16; a.stats = stats; a.n_part = n_part; a.colsum = dc->zeros;
        a.ln_inv_d = 1.0f / (float)d; a.ln_eps = hp.rms_eps; a.ln_rms = 1;
        return gemm_linear(kind, dc->fmt, a, st);
    };
    for (int l = 0; l < hp.n_layers; ++l) {
        const auto& L = dc->layers[l];
        Q3A_TRY(linear(EPI_LNFOLD_OP, x16, d, L.w_qkv, dqkv, dc->qkv.p, dc->zeros, false));
        {
            const long long warps = (long long)n * (H + Hkv);
            Q3A_CUDA(launch_pdl(dec_qk_norm_rope_kernel<OpT>, dim3((unsigned)ceil_div64(warps, 8)), dim3(256), 0, st, dc->qkv.as<OpT>(),
                                L.q_norm, L.k_norm, n, H, Hkv, hp.rms_eps, hp.rope_theta));
            ++g_launch_counter;
        }
        Q3A_TRY(attention_dec_launch(dc->qkv.p, dc->att.p, n, H, Hkv, dc->fmt, st));
        Q3A_TRY(linear(EPI_BIAS_RES_F32_DUAL, dc->att.p, dq, L.w_o, d, x, nullptr, true));
        Q3A_TRY(linear(EPI_LNFOLD_OP, x16, d, L.w_gu, 2 * inter, dc->gu.p, dc->zeros, false));
        {
            const long long pairs = (long long)n * inter / 2;
            const int blocks = (int)std::min<long long>(ceil_div64(pairs, 256), 148 * 8);
            Q3A_CUDA(launch_pdl(dec_swiglu_kernel<OpT>, dim3(blocks), dim3(256), 0, st, dc->gu.as<OpT>(), dc->act.as<OpT>(), pairs, inter));
            ++g_launch_counter;
        }
        Q3A_TRY(linear(EPI_BIAS_RES_F32_DUAL, dc->act.p, inter, L.w_down, d, x, nullptr, true));
    }
    Q3A_TRY(linear(EPI_LNFOLD_F32, x16, d, dc->w_head, hp.n_classes, logits_dev, dc->b_head, false));
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

extern "C" int q3a_decoder_forward_device(q3a_decoder* dc, const int32_t* tokens, int n_tokens, const float* audio_dev,
                                          int n_audio, int audio_start, float* logits_dev, void* stream) {
    if (!dc) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (n_tokens < 0 || (n_tokens > 0 && (!tokens || !logits_dev))) { set_error("decoder_forward: null argument"); return Q3A_ERR_ARG; }
    if (n_tokens == 0) return Q3A_OK;
    if ((reinterpret_cast<uintptr_t>(logits_dev) & 15u) || (audio_dev && (reinterpret_cast<uintptr_t>(audio_dev) & 7u))) {
        set_error("decoder_forward: logits_dev must be 16-byte aligned, audio_dev 8-byte aligned"); return Q3A_ERR_ARG;
    }
    Q3A_TRY(use_device(dc->device));
    g_launch_counter = 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (dc->fmt == OP_F16) return dec_forward_core<__half>(dc, tokens, n_tokens, audio_dev, n_audio, audio_start, logits_dev, st);
    return dec_forward_core<__nv_bfloat16>(dc, tokens, n_tokens, audio_dev, n_audio, audio_start, logits_dev, st);
}

// host-pointer twin with the reference's signature shape (forced_aligner.cpp:1088-1092): audio_embd and logits on the host
extern "C" int q3a_decoder_forward(q3a_decoder* dc, const int32_t* tokens, int n_tokens, const float* audio_embd, int n_audio,
                                   int audio_start, float* logits) {
    if (!dc) { set_error("Model not loaded"); return Q3A_ERR_STATE; }
    if (n_tokens < 0 || (n_tokens > 0 && (!tokens || !logits))) { set_error("decoder_forward: null argument"); return Q3A_ERR_ARG; }
    if (n_tokens == 0) return Q3A_OK;
    Q3A_TRY(use_device(dc->device));
    const size_t lbytes = sizeof(float) * (size_t)n_tokens * dc->hp.n_classes;
    Q3A_TRY(dc->logits_stage.reserve(lbytes));
    const float* audio_dev = nullptr;
    if (audio_embd && n_audio > 0) {
        const size_t abytes = sizeof(float) * (size_t)n_audio * dc->hp.hidden;
        Q3A_TRY(dc->audio_stage.reserve(abytes));
        Q3A_CUDA(cudaMemcpyAsync(dc->audio_stage.p, audio_embd, abytes, cudaMemcpyHostToDevice, dc->stream));
        audio_dev = dc->audio_stage.as<float>();
    }
    Q3A_TRY(q3a_decoder_forward_device(dc, tokens, n_tokens, audio_dev, n_audio, audio_start, dc->logits_stage.as<float>(), dc->stream));
    Q3A_CUDA(cudaMemcpyAsync(logits, dc->logits_stage.p, lbytes, cudaMemcpyDeviceToHost, dc->stream));
    Q3A_CUDA(cudaStreamSynchronize(dc->stream));
    return Q3A_OK;
}
