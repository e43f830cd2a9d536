// Audio-encoder object (host side): resident weights, workspace and the launch sequence.
#pragma once
#include "common.cuh"
#include "attention.h"
#include "stem.h"

#include <string>
#include <vector>

struct q3a_encoder {
    int device = 0;
    int fmt = q3a::OP_BF16;          // 16-bit operand type of every tensor-core kernel
    q3a_hparams hp{};
    cudaStream_t stream = nullptr;   // compute stream of the host-pointer entry points
    // host-pointer entry points run as a 3-stage pipeline over passes: upload (copy_in) | compute (stream) |
    // download (copy_out), staging buffers double-buffered by pass parity
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
    static constexpr int kInParts = 4, kOutParts = 4;   // upper bounds; the split tables choose   // finer overlap inside a pass (PartHooks in encoder.cu)
    cudaEvent_t ev_part_in[2][kInParts] = {}, ev_part_out[2][kOutParts] = {};
    long long pass_seq = 0;          // passes enqueued so far: staging buffer / event parity runs across calls
    static constexpr int kTickets = 8;
    cudaEvent_t ev_done[kTickets] = {};   // completion of the last kTickets host-pointer calls (q3a_encode_wait)
    int64_t ticket_seq = 0;

    // ---- weights (one arena) -----------------------------------------------------------------
    q3a::DevBuf arena;
    size_t arena_used = 0;
    const float *conv1_w = nullptr, *conv1_b = nullptr, *conv2_b = nullptr, *conv3_b = nullptr;
    const void *conv2_w = nullptr, *conv3_w = nullptr, *conv_out_w = nullptr;
    const void* conv1_wimg = nullptr;   // smem image of the conv1 weights for conv1_tc_kernel (null: CUDA-core conv1)
    // a Linear weight: [N][K] 16-bit operands, or -- Q8_0 file, fused-dequant path -- p = int8 plane, d = f16 block scales
    struct Wt { const void* p = nullptr; const void* d = nullptr; };
    struct Layer {
        const float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *b_qkv, *b_o, *b_up, *b_down;
        Wt w_qkv, w_o, w_up, w_down;
        // LayerNorm fold (ln_fold): w_qkv / w_up hold gamma o W, b_qkv / b_up hold b + beta . W, s_* the column sums of
        // the folded, rounded weights (epilogues.cuh)
        const float *s_qkv = nullptr, *s_up = nullptr;
    };
    std::vector<Layer> layers;
    const float *lnp_g = nullptr, *lnp_b = nullptr, *b_p1 = nullptr, *b_p2 = nullptr;
    Wt w_p1, w_p2;
    bool q8_fused = false;           // Linear weights stay Q8_0 in device memory, dequantised inside the GEMM (gemm_tc.cuh kWq8)
    const float* s_p1 = nullptr;
    bool ln_fold = false;            // LayerNorm folded into the QKV / FFN-up / proj1 GEMMs (no layernorm_kernel launches)
    int fold_mask = 0;               // which of them: 1 = attn_norm -> QKV, 2 = ffn_norm -> FFN-up, 4 = ln_post -> proj1 (7 = all)
    q3a::DevBuf pe;                  // sinusoidal table [pe_rows][d_model] f32
    int pe_rows = 0;

    // ---- workspace (grows, never shrinks) -----------------------------------------------------
    q3a::DevBuf planes1, planes2, feat, x, h, qkv, att, up, stats, tables, stage_in, stage_out, stage_in2, stage_out2, stage_pcm, stage_mel, stage_max;
    int rcap1 = 0, rcap2 = 0, sigP2 = 0, sigP3 = 0, sigP2_planes2 = 0;   // plane layout the zero rows were prepared for
    q3a::PinnedRing ring;

    // ---- profiling ------------------------------------------------------------------------------
    bool profiling = false;
    std::vector<cudaEvent_t> ev_pool;
    struct Span { int kind; int e0, e1; };
    std::vector<Span> spans;
    int ev_used = 0;
    q3a_timings tm{};
    int launches_last = 0;
};

struct q3a_frontend;

namespace q3a {

// tokens of one stem chunk (audio_encoder.cpp:304-310)
inline int stem_out_len(int n) {
    for (int i = 0; i < 3; ++i) n = (n - 1) / 2 + 1;
    return n;
}
int encoder_n_tokens(int n_frames, int mode);

}  // namespace q3a
