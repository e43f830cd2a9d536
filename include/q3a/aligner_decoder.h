// Source-compatible stand-in for the decoder half of the reference's ForcedAligner (src/forced_aligner.h:244-252,
// src/forced_aligner.cpp:926-1169) as a header-only C++17 shim over the libq3a C ABI: forward_decoder() with the
// reference's argument list (tokens, audio_embd [n_audio][hidden] f32, audio_start_pos) and its output layout (logits
// [n_tokens][classify_num] f32 appended to the caller's std::vector, :1158-1161), bool + get_error() error behaviour.
// forward_decoder_device() is the additive variant that takes the audio features where the encoder left them in HBM.
// There is no CPU fallback.
#ifndef Q3A_SHIM_ALIGNER_DECODER_H
#define Q3A_SHIM_ALIGNER_DECODER_H

#include <cstdint>
#include <string>
#include <vector>

#include "../q3a.h"

namespace qwen3_asr {

class AlignerDecoder {
public:
    AlignerDecoder() = default;
    ~AlignerDecoder() { if (dec_) q3a_decoder_free(dec_); }
    AlignerDecoder(const AlignerDecoder&) = delete;
    AlignerDecoder& operator=(const AlignerDecoder&) = delete;

    void set_device(int device) { device_ = device; }
    void set_operand_dtype(int q3a_op) { operand_ = q3a_op; }

    // the decoder tensors of the aligner GGUF (forced_aligner.cpp:267-314,381-410)
    bool load_model(const std::string& model_path) {
        if (dec_) { q3a_decoder_free(dec_); dec_ = nullptr; }
        if (q3a_decoder_load(model_path.c_str(), device_, operand_, &dec_) != Q3A_OK) {
            error_msg_ = std::string("Failed to load model: ") + q3a_last_error();
            return false;
        }
        q3a_decoder_get_hparams(dec_, &hp_);
        return true;
    }

    // forced_aligner.cpp:1088-1169
    bool forward_decoder(const int32_t* tokens, int32_t n_tokens, const float* audio_embd, int32_t n_audio,
                         int32_t audio_start_pos, std::vector<float>& output) {
        if (!dec_) { error_msg_ = "Model not loaded"; return false; }
        output.resize((size_t)(n_tokens > 0 ? n_tokens : 0) * hp_.n_classes);
        if (q3a_decoder_forward(dec_, tokens, n_tokens, audio_embd, n_audio, audio_start_pos, output.data()) != Q3A_OK) {
            error_msg_ = q3a_last_error();
            return false;
        }
        return true;
    }

    // audio_dev / logits_dev: device memory; stream: a cudaStream_t (nullptr = default stream); asynchronous
    bool forward_decoder_device(const int32_t* tokens, int32_t n_tokens, const float* audio_dev, int32_t n_audio,
                                int32_t audio_start_pos, float* logits_dev, void* stream) {
        if (!dec_) { error_msg_ = "Model not loaded"; return false; }
        if (q3a_decoder_forward_device(dec_, tokens, n_tokens, audio_dev, n_audio, audio_start_pos, logits_dev, stream) != Q3A_OK) {
            error_msg_ = q3a_last_error();
            return false;
        }
        return true;
    }

    const q3a_decoder_hparams& get_hparams() const { return hp_; }
    const std::string& get_error() const { return error_msg_; }

private:
    q3a_decoder* dec_ = nullptr;
    int device_ = 0;
    int operand_ = Q3A_OP_BF16;
    q3a_decoder_hparams hp_{};
    std::string error_msg_;
};

}  // namespace qwen3_asr

#endif  // Q3A_SHIM_ALIGNER_DECODER_H
