// Source-compatible stand-in for the reference's src/audio_encoder.h (+ the hparams structs of
// src/gguf_loader.h:15-35) on the hot path, as a header-only C++17 shim over the libq3a C ABI.
// Qwen3ASR (reference qwen3_asr.cpp:21-43,105) compiles against this header unchanged: same namespace,
// class name, method signatures, bool + get_error() error behaviour, and the same output layout
// (row-major [n_tokens][hidden_size] f32 appended to the caller's std::vector, audio_encoder.cpp:591-595).
//
// Differences, all additive:
//   * encode_aligner(): the windowed / padded-chunk variant that ForcedAligner::encode_audio keeps private
//     (forced_aligner.h:241-242, forced_aligner.cpp:591-924);
//   * encode_batch(): several utterances in one call (the reference encodes one at a time);
//   * set_device() / set_operand_dtype() before load_model().
// There is no CPU fallback: load_model() fails with the library's message when no sm_100 device is usable.
#ifndef Q3A_SHIM_AUDIO_ENCODER_H
#define Q3A_SHIM_AUDIO_ENCODER_H

#include <cstdint>
#include <string>
#include <vector>

#include "../q3a.h"

namespace qwen3_asr {

// gguf_loader.h:15-25
struct audio_encoder_hparams {
    int32_t n_encoder_layers = 18;
    int32_t d_model = 896;
    int32_t n_attention_heads = 14;
    int32_t ffn_dim = 3584;
    int32_t conv_channels = 480;
    int32_t conv_out_dim = 896;
    int32_t n_mel_bins = 128;
    int32_t n_window_infer = 800;
    float layer_norm_eps = 1e-5f;
};

// gguf_loader.h:28-35 (only hidden_size is produced by this path; the rest keep the reference defaults)
struct text_decoder_hparams {
    int32_t hidden_size = 1024;
    int32_t n_decoder_layers = 28;
    int32_t n_attention_heads = 16;
    int32_t n_key_value_heads = 8;
    int32_t intermediate_size = 3072;
    float rms_norm_eps = 1e-6f;
};

class AudioEncoder {
public:
    AudioEncoder() = default;
    ~AudioEncoder() { if (enc_) q3a_encoder_free(enc_); }
    AudioEncoder(const AudioEncoder&) = delete;
    AudioEncoder& operator=(const AudioEncoder&) = delete;

    void set_device(int device) { device_ = device; }
    void set_operand_dtype(int q3a_op) { operand_ = q3a_op; }     // Q3A_OP_BF16 (default) or Q3A_OP_F16

    // audio_encoder.h:25 / audio_encoder.cpp:42-83
    bool load_model(const std::string& model_path) {
        if (enc_) { q3a_encoder_free(enc_); enc_ = nullptr; }
        if (q3a_encoder_load(model_path.c_str(), device_, operand_, &enc_) != Q3A_OK) {
            error_msg_ = std::string("Failed to load model: ") + q3a_last_error();
            return false;
        }
        q3a_hparams hp;
        q3a_encoder_get_hparams(enc_, &hp);
        hparams_.n_encoder_layers = hp.n_encoder_layers;
        hparams_.d_model = hp.d_model;
        hparams_.n_attention_heads = hp.n_attention_heads;
        hparams_.ffn_dim = hp.ffn_dim;
        hparams_.conv_channels = hp.conv_channels;
        hparams_.conv_out_dim = hp.conv_out_dim;
        hparams_.n_mel_bins = hp.n_mel_bins;
        hparams_.n_window_infer = hp.n_window_infer;
        hparams_.layer_norm_eps = hp.layer_norm_eps;
        text_hparams_.hidden_size = hp.out_dim;
        return true;
    }

    // audio_encoder.h:27-28 / audio_encoder.cpp:312-601
    bool encode(const float* mel_data, int n_mel, int n_frames, std::vector<float>& output) {
        return run(mel_data, n_mel, n_frames, Q3A_MODE_ASR, output);
    }
    // audio_encoder.h:30-31 / audio_encoder.cpp:854-913
    bool encode_conv_only(const float* mel_data, int n_mel, int n_frames, std::vector<float>& output) {
        return run(mel_data, n_mel, n_frames, Q3A_MODE_CONV_ONLY, output);
    }
    // audio_encoder.h:34-35 / audio_encoder.cpp:603-852
    bool encode_no_chunk(const float* mel_data, int n_mel, int n_frames, std::vector<float>& output) {
        return run(mel_data, n_mel, n_frames, Q3A_MODE_NO_CHUNK, output);
    }
    // ForcedAligner::encode_audio (forced_aligner.cpp:591-924)
    bool encode_aligner(const float* mel_data, int n_mel, int n_frames, std::vector<float>& output) {
        return run(mel_data, n_mel, n_frames, Q3A_MODE_ALIGNER, output);
    }

    // several utterances at once; outputs[i] receives [n_tokens_i][hidden] like encode()
    bool encode_batch(const std::vector<const float*>& mels, int n_mel, const std::vector<int>& n_frames, int mode,
                      std::vector<std::vector<float>>& outputs) {
        if (!enc_) { error_msg_ = "Model not loaded"; return false; }
        if (mels.size() != n_frames.size()) { error_msg_ = "encode_batch: size mismatch"; return false; }
        const int n = (int)mels.size();
        const int width = mode == Q3A_MODE_CONV_ONLY ? hparams_.d_model : text_hparams_.hidden_size;
        outputs.resize(n);
        std::vector<float*> outs(n);
        for (int i = 0; i < n; ++i) {
            outputs[i].resize((size_t)q3a_encoder_n_tokens(n_frames[i], mode) * width);
            outs[i] = outputs[i].data();
        }
        if (q3a_encoder_encode_batch(enc_, mels.data(), n_mel, n_frames.data(), n, mode, outs.data(), nullptr) !=
            Q3A_OK) {
            error_msg_ = q3a_last_error();
            return false;
        }
        return true;
    }

    const audio_encoder_hparams& get_hparams() const { return hparams_; }
    const text_decoder_hparams& get_text_hparams() const { return text_hparams_; }
    const std::string& get_error() const { return error_msg_; }
    q3a_encoder* handle() const { return enc_; }

private:
    bool run(const float* mel_data, int n_mel, int n_frames, int mode, std::vector<float>& output) {
        if (!enc_) { error_msg_ = "Model not loaded"; return false; }
        const int width = mode == Q3A_MODE_CONV_ONLY ? hparams_.d_model : text_hparams_.hidden_size;
        const int n_tok = q3a_encoder_n_tokens(n_frames, mode);
        output.resize((size_t)(n_tok > 0 ? n_tok : 0) * width);
        float* out = output.data();
        if (q3a_encoder_encode_batch(enc_, &mel_data, n_mel, &n_frames, 1, mode, &out, nullptr) != Q3A_OK) {
            error_msg_ = q3a_last_error();
            return false;
        }
        return true;
    }

    q3a_encoder* enc_ = nullptr;
    int device_ = 0;
    int operand_ = Q3A_OP_BF16;
    audio_encoder_hparams hparams_;
    text_decoder_hparams text_hparams_;
    std::string error_msg_;
};

}  // namespace qwen3_asr

#endif  // Q3A_SHIM_AUDIO_ENCODER_H
