// Log-mel frontend object (host side).  Kernels live in mel.cu.
#pragma once
#include "common.cuh"

struct q3a_frontend {
    int device = 0;
    int n_mel = 128;
    int n_bins = 201;
    int n_sms = 148;                // persistent grid of mel_frames_kernel = 2 CTAs per SM
    // resident tables
    float* d_hann = nullptr;        // [400] periodic Hann, float(double value) (mel_spectrogram.cpp:28-58)
    float2* d_twiddle = nullptr;    // [25][9]  W400^(n1*k2)
    int* d_row_bin0 = nullptr;      // [n_mel]   first FFT bin with a non-zero weight
    int* d_row_ptr = nullptr;       // [n_mel+1] offsets into d_weights
    float* d_weights = nullptr;     // compressed filterbank rows
    // per-call staging (host-pointer API) and descriptors
    q3a::DevBuf pcm, mel, desc, maxbuf;
    q3a::PinnedRing ring;
    cudaStream_t stream = nullptr;  // used by the host-pointer API
};

namespace q3a {

struct MelClipDesc {
    long long pcm_off;   // first sample of the clip in the packed PCM buffer
    long long mel_off;   // first float of the clip's [n_mel][n_len] block
    int n_samples;
    int n_len;
    int tile0;           // index of the clip's first frame tile in the grid
    int pad;
};

// enqueue the frontend kernels for a packed batch already resident on the device
int mel_launch(q3a_frontend* fe, const float* pcm_dev, const int64_t* pcm_off, const int* n_samples, int n_clips,
               float* mel_dev, const int64_t* mel_off, float* max_dev, int finalize, cudaStream_t stream);

}  // namespace q3a
