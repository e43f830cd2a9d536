// Source-compatible stand-in for the reference's src/mel_spectrogram.h on the hot path, as a
// header-only C++17 shim over the libq3a C ABI (include/q3a.h).  A caller such as
// Qwen3ASR::transcribe_internal (reference qwen3_asr.cpp:90) or ForcedAligner::align
// (forced_aligner.cpp:1649-1653) compiles against this header unchanged and links libq3a.so.
//
// Kept from the reference interface (mel_spectrogram.h:9-30,42-43,53-55):
//   QWEN_* constants, MelSpectrogram, MelFilters, generate_mel_filters(), log_mel_spectrogram().
// Not provided here (host file I/O, out of scope — SURVEY.md §2 row 1): load_wav, *_npy, compare_mel.
//
// There is no CPU fallback: log_mel_spectrogram() returns false (message on stderr, as the
// reference does) when no sm_100 device is usable.
#ifndef Q3A_SHIM_MEL_SPECTROGRAM_H
#define Q3A_SHIM_MEL_SPECTROGRAM_H

#include <cstdint>
#include <cstdio>
#include <mutex>
#include <vector>

#include "../q3a.h"

constexpr int QWEN_SAMPLE_RATE = 16000;
constexpr int QWEN_N_FFT = 400;
constexpr int QWEN_HOP_LENGTH = 160;
constexpr int QWEN_N_MELS = 128;
constexpr int QWEN_CHUNK_SIZE = 30;
constexpr int QWEN_N_SAMPLES = QWEN_SAMPLE_RATE * QWEN_CHUNK_SIZE;
constexpr int QWEN_N_FFT_BINS = QWEN_N_FFT / 2 + 1;

// [n_mel][n_len] mel-major, exactly the memory the reference hands to AudioEncoder::encode
struct MelSpectrogram {
    int32_t n_mel = 0;
    int32_t n_len = 0;
    int32_t n_len_org = 0;
    std::vector<float> data;
};

// [n_mel][n_fft] with n_fft = 201 frequency bins (the reference's field naming)
struct MelFilters {
    int32_t n_mel = 0;
    int32_t n_fft = 0;
    std::vector<float> data;
};

namespace q3a_shim {

// One resident frontend per (device, filterbank contents); created on first use, shared by threads
// (q3a_mel_batch serialises on the frontend's stream, so calls are guarded by the mutex).
struct FrontendCache {
    std::mutex mu;
    q3a_frontend* fe = nullptr;
    std::vector<float> bank;
    int n_mel = 0;
    ~FrontendCache() { if (fe) q3a_frontend_free(fe); }
};
inline FrontendCache& frontend_cache() {
    static FrontendCache c;
    return c;
}
// device the shims use; set before the first call (default 0)
inline int& device_index() {
    static int d = 0;
    return d;
}

}  // namespace q3a_shim

// replaces generate_mel_filters (mel_spectrogram.cpp:361-415)
inline void generate_mel_filters(MelFilters& filters, int n_mels = QWEN_N_MELS, int n_fft = QWEN_N_FFT,
                                 int sample_rate = QWEN_SAMPLE_RATE) {
    filters.n_mel = n_mels;
    filters.n_fft = n_fft / 2 + 1;
    filters.data.assign((size_t)filters.n_mel * filters.n_fft, 0.0f);
    q3a_generate_mel_filters(filters.data.data(), n_mels, n_fft, sample_rate);
}

// replaces log_mel_spectrogram (mel_spectrogram.cpp:484-628).  n_threads is accepted and ignored,
// as in the reference.
inline bool log_mel_spectrogram(const float* samples, int n_samples, const MelFilters& filters, MelSpectrogram& mel,
                                int n_threads = 1) {
    (void)n_threads;
    if (!samples || n_samples <= 0 || filters.n_fft != QWEN_N_FFT_BINS ||
        filters.data.size() != (size_t)filters.n_mel * filters.n_fft) {
        std::fprintf(stderr, "log_mel_spectrogram: bad arguments\n");
        return false;
    }
    auto& c = q3a_shim::frontend_cache();
    std::lock_guard<std::mutex> lock(c.mu);
    if (!c.fe || c.n_mel != filters.n_mel || c.bank != filters.data) {
        if (c.fe) { q3a_frontend_free(c.fe); c.fe = nullptr; }
        if (q3a_frontend_create(q3a_shim::device_index(), filters.data.data(), filters.n_mel, filters.n_fft, &c.fe) !=
            Q3A_OK) {
            std::fprintf(stderr, "log_mel_spectrogram: %s\n", q3a_last_error());
            return false;
        }
        c.bank = filters.data;
        c.n_mel = filters.n_mel;
    }
    mel.n_mel = filters.n_mel;
    mel.n_len = q3a_mel_n_len(n_samples);
    mel.n_len_org = mel.n_len;
    mel.data.resize((size_t)mel.n_mel * mel.n_len);
    if (mel.n_len == 0) return true;            // fewer than one hop of audio: empty spectrogram
    float* out = mel.data.data();
    int n_len = 0;
    if (q3a_mel_batch(c.fe, &samples, &n_samples, 1, &out, &n_len) != Q3A_OK) {
        std::fprintf(stderr, "log_mel_spectrogram: %s\n", q3a_last_error());
        return false;
    }
    return true;
}

#endif  // Q3A_SHIM_MEL_SPECTROGRAM_H
