// Conv stem layout + the CUDA-core kernels around the tensor-core GEMMs (stem.cu).
//
// The three 3x3 stride-2 pad-1 convolutions (audio_encoder.cpp:105-128, forced_aligner.cpp:638-657)
// run on "parity-split planes": a conv input with T time steps, F mel rows and C channels is stored as
// four channels-last planes indexed by (time parity, mel parity),
//     plane[(t&1)*2 + (f&1)][chunk*P + 1 + t/2][f/2][c]          (16-bit)
// where every chunk owns P rows and its row 0 is all zero.  With that layout each of the 9 taps of a
// stride-2 conv is a dense box (rows R_out-1 or R_out, cols of-1 or of) of one plane, so the implicit
// GEMM's A tiles are plain TMA boxes and the zero padding is the zero row / TMA out-of-bounds fill.
#pragma once
#include "common.cuh"

namespace q3a {

struct StemChunk {
    long long mel_off;   // float offset of the chunk's first frame in mel row 0 of its utterance
    int row_stride;      // floats between mel rows (the utterance's n_frames)
    int len;             // valid mel frames in this chunk
    int t1_valid;        // conv1 output time steps to compute (the rest are written as zeros)
    int utt;             // utterance index (for the fused clamp's per-clip maximum)
};

// conv1 (1 -> C channels) + bias + GELU, written as the planes conv2 reads.  P2 = rows per chunk.
// clip_max != nullptr: mel holds raw log10 energies and x = (max(x, clip_max[utt]-8)+4)/4 is applied on load
// (mel_spectrogram.cpp:615-625 fused into the consumer).
template <class OpT>
int conv1_launch(const float* mel, const StemChunk* chunks_dev, int n_chunks, int P2, const float* clip_max,
                 const float* w /*[C][9]: df*3+dt*/, const float* bias, int C, OpT* planes, long long plane_stride,
                 cudaStream_t stream);

// Tensor-core conv1 (conv1_tc.cu): same result layout; used whenever the channel count fits (C % 32 == 0, C <= 512).
// w_image: device copy of conv1_tc_weight_image(); R_cap: rows per plane of `planes` (plane stride = R_cap*32*C).
bool conv1_tc_supported(int C);
void conv1_tc_weight_image(const float* w /*[C][9]: df*3+dt*/, int C, std::vector<uint8_t>& image);
template <class OpT>
int conv1_tc_launch(const float* mel, const StemChunk* chunks_dev, int n_chunks, int P2, const float* clip_max,
                    const void* w_image, const float* bias, int C, OpT* planes, int R_cap, int row0, cudaStream_t stream);

// LayerNorm over rows of x [T][d] f32 -> 16-bit operand rows (audio_encoder.cpp:441-448)
template <class OpT>
int layernorm_launch(const float* x, const float* gamma, const float* beta, float eps, int T, int d, OpT* out,
                     cudaStream_t stream);

}  // namespace q3a
