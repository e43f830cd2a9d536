// Drives the hot path exactly the way the reference's callers do (Qwen3ASR::transcribe_internal,
// qwen3_asr.cpp:81-113; tests/test_mel.cpp:97; tests/test_encoder.cpp:213-230) but through the shim headers
// over libq3a:   shim_harness <pcm.f32> <model.gguf|-> <mel_out.f32> <feat_out.f32> [mode: asr|aligner]
// Exit codes: 0 ok, 2 mel failed, 3 load failed, 4 encode failed, 1 usage / I/O.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "q3a/aligner_decoder.h"
#include "q3a/audio_encoder.h"
#include "q3a/mel_spectrogram.h"

static bool read_f32(const char* path, std::vector<float>& v) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    v.resize((size_t)n / 4);
    size_t got = std::fread(v.data(), 4, v.size(), f);
    std::fclose(f);
    return got == v.size();
}
static bool write_f32(const char* path, const std::vector<float>& v) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    size_t put = std::fwrite(v.data(), 4, v.size(), f);
    std::fclose(f);
    return put == v.size();
}

int main(int argc, char** argv) {
    if (argc < 5) { std::fprintf(stderr, "usage: %s pcm.f32 model.gguf|- mel_out feat_out [asr|aligner [decoder.gguf logits_out]]\n", argv[0]); return 1; }
    std::vector<float> samples;
    if (!read_f32(argv[1], samples)) { std::fprintf(stderr, "cannot read %s\n", argv[1]); return 1; }

    MelFilters filters;
    generate_mel_filters(filters);                                   // qwen3_asr.cpp:34
    MelSpectrogram mel;
    if (!log_mel_spectrogram(samples.data(), (int)samples.size(), filters, mel, 4)) return 2;   // qwen3_asr.cpp:90
    if (!write_f32(argv[3], mel.data)) return 1;
    std::printf("mel %d x %d\n", mel.n_mel, mel.n_len);
    if (std::strcmp(argv[2], "-") == 0) return 0;

    qwen3_asr::AudioEncoder encoder;
    if (!encoder.load_model(argv[2])) { std::fprintf(stderr, "%s\n", encoder.get_error().c_str()); return 3; }
    std::vector<float> audio_features;
    const bool aligner = argc > 5 && std::strcmp(argv[5], "aligner") == 0;
    const bool ok = aligner ? encoder.encode_aligner(mel.data.data(), mel.n_mel, mel.n_len, audio_features)
                            : encoder.encode(mel.data.data(), mel.n_mel, mel.n_len, audio_features);   // qwen3_asr.cpp:105
    if (!ok) { std::fprintf(stderr, "%s\n", encoder.get_error().c_str()); return 4; }
    const int hidden = encoder.get_text_hparams().hidden_size;
    std::printf("features %d x %d\n", (int)(audio_features.size() / hidden), hidden);   // qwen3_asr.cpp:113
    // error path the reference tests: wrong number of mel bins (audio_encoder.cpp:321-325)
    std::vector<float> dummy;
    if (encoder.encode(mel.data.data(), 64, mel.n_len, dummy)) return 4;
    std::printf("mismatch: %s\n", encoder.get_error().c_str());
    if (!write_f32(argv[4], audio_features)) return 1;
    // optional: decoder.gguf logits_out -- ForcedAligner::align's decoder step (forced_aligner.cpp:1661-1690): a prompt of
    // audio_start + n_audio pads + a few text tokens, the encoder's features spliced in at position 1
    if (argc > 7) {
        qwen3_asr::AlignerDecoder decoder;
        if (!decoder.load_model(argv[6])) { std::fprintf(stderr, "%s\n", decoder.get_error().c_str()); return 5; }
        const int n_audio = (int)(audio_features.size() / hidden);
        if (decoder.get_hparams().hidden != hidden) { std::fprintf(stderr, "decoder width differs from the encoder's\n"); return 5; }
        std::vector<int32_t> tokens((size_t)n_audio + 12);
        for (size_t i = 0; i < tokens.size(); ++i) tokens[i] = (int32_t)((i * 7 + 3) % (size_t)decoder.get_hparams().vocab);
        std::vector<float> logits;
        if (!decoder.forward_decoder(tokens.data(), (int32_t)tokens.size(), audio_features.data(), n_audio, 1, logits)) {
            std::fprintf(stderr, "%s\n", decoder.get_error().c_str()); return 6;
        }
        std::printf("logits %d x %d\n", (int)tokens.size(), decoder.get_hparams().n_classes);
        if (!write_f32(argv[7], logits)) return 1;
    }
    return 0;
}
