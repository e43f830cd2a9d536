// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// C-ABI shim around the *reference's own* translation unit
// /root/reference/src/mel_spectrogram.cpp, which compiles stand-alone (libm only).
// The Makefile next to this file compiles that TU where it lies and links it with this
// shim into oracle/_ref/libref_mel.so.  No reference source is copied into the repo.
//
// Wrapped entry points: generate_mel_filters (mel_spectrogram.cpp:361-415) and
// log_mel_spectrogram (mel_spectrogram.cpp:484-628), declared in mel_spectrogram.h:42-55.
#include "mel_spectrogram.h"   // found through -I/root/reference/src
#include <cstring>

extern "C" {

// out: [n_mels * (n_fft/2+1)] floats, row-major [mel][bin]
void ref_generate_mel_filters(float* out, int n_mels, int n_fft, int sample_rate) {
    MelFilters f;
    generate_mel_filters(f, n_mels, n_fft, sample_rate);
    std::memcpy(out, f.data.data(), f.data.size() * sizeof(float));
}

// n_len for a clip of n_samples (what the reference will resize mel.data to)
int ref_mel_n_len(int n_samples) { return n_samples / QWEN_HOP_LENGTH; }

// filters: [n_mel][n_bins] row-major; out: [n_mel][n_len] mel-major.  Returns n_len or -1.
int ref_log_mel(const float* pcm, int n_samples, const float* filters, int n_mel, int n_bins,
                float* out, int out_capacity) {
    MelFilters f;
    f.n_mel = n_mel;
    f.n_fft = n_bins;
    f.data.assign(filters, filters + (size_t)n_mel * n_bins);
    MelSpectrogram mel;
    if (!log_mel_spectrogram(pcm, n_samples, f, mel, 1)) return -1;
    if ((int)mel.data.size() > out_capacity) return -1;
    std::memcpy(out, mel.data.data(), mel.data.size() * sizeof(float));
    return mel.n_len;
}

}  // extern "C"
