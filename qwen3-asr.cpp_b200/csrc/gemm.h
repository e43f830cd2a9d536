// Host launchers for the tcgen05 GEMM (gemm_tc.cuh) with the fused epilogues of epilogues.cuh.
#pragma once
#include "common.cuh"

namespace q3a {

enum EpiKind : int {
    EPI_BIAS_OP = 0,        // 16-bit out = acc + bias
    EPI_BIAS_GELU_OP = 1,   // 16-bit out = gelu(acc + bias)
    EPI_BIAS_F32 = 2,       // f32 out = acc + bias
    EPI_BIAS_RES_F32 = 3,   // f32 out = acc + bias + res
    EPI_POSEMB_F32 = 4,     // f32 out = acc + pe[pos[m]]
    // LayerNorm folded into the GEMMs (epilogues.cuh): producers of the residual stream also emit its 16-bit copy (out16)
    // and per-row statistics partials (stats), consumers apply mean / rstd in the epilogue
    EPI_BIAS_RES_F32_DUAL = 5,   // f32 out = acc + bias + res ; out16 ; stats
    EPI_POSEMB_F32_DUAL = 6,     // f32 out = acc + pe[pos[m]] ; out16 ; stats
    EPI_LNFOLD_OP = 7,           // 16-bit out = rstd * (acc - mean * colsum) + bias'
    EPI_LNFOLD_GELU_OP = 8,      // 16-bit out = gelu(...)
    EPI_LNFOLD_F32 = 9,          // f32 out = rstd * (acc - mean * colsum) + bias'
};

struct LinearArgs {
    const void* a = nullptr;   // [M][K] 16-bit, leading dimension lda
    int M = 0, K = 0, lda = 0;
    const void* w = nullptr;   // [N][K] 16-bit, leading dimension ldw; or, with w_d set, the int8 plane of Q8_0 weights
    const void* w_d = nullptr; // Q8_0 only: [N][K/32] f16 block scales (fused-dequant path, K % 64 == 0)
    int N = 0, ldw = 0;
    void* out = nullptr;       // [M][ldc], 16-bit or f32 depending on the epilogue
    int ldc = 0;
    const float* bias = nullptr;
    const float* res = nullptr;
    const float* pe = nullptr;
    const int* pos = nullptr;
    int block_n = 0;           // 0 = pick from N
    // LayerNorm fold
    void* out16 = nullptr;     // DUAL: [M][ldc] 16-bit copy of out
    float2* stats = nullptr;   // DUAL: written [M][n_part]; LNFOLD: read
    int n_part = 0;            // partials per row = 2 * ceil(d_model / BLOCK_N of the producer): ln_partials()
    const float* colsum = nullptr;   // LNFOLD: [N]
    float ln_inv_d = 0.0f, ln_eps = 0.0f;
    int ln_rms = 0;            // LNFOLD: RMSNorm instead of LayerNorm (no mean)
};
// statistics partials per row that a DUAL GEMM with N output columns writes
int ln_partials(int N);

int gemm_linear(EpiKind kind, int fmt, const LinearArgs& a, cudaStream_t stream);

struct ConvArgs {
    const void* planes = nullptr;  // input [4][R_cap][Fh][C] 16-bit
    int R_cap = 0, R_used = 0, Fh = 0, C = 0;   // R_cap: rows between planes (stride); R_used: rows this launch covers
    int R_map = 0;                 // rows the tensor map may address from `planes` (0: R_cap); smaller when `planes`
                                   // points into the middle of the arena (a sub-range of chunks)
    const void* w = nullptr;       // [N][9 * cpad] 16-bit, cpad = ceil(C/64)*64, k = tap*cpad + c
    int N = 0;
    const float* bias = nullptr;
    int P = 0;                     // rows per chunk of this conv's row space
    // to-planes epilogue (conv2)
    void* out_planes = nullptr; int P_next = 0; long long out_plane_stride = 0; const int* ot_valid = nullptr;
    // to-features epilogue (conv3)
    void* feat = nullptr; const int* n_tok = nullptr; const int* tok_base = nullptr;
};

int gemm_conv(bool to_feat, int fmt, const ConvArgs& a, cudaStream_t stream);

int pick_block_n(int N);
int sm_count();

// one-time (cached per kernel and device) setup of a gemm_tc_kernel instantiation: opt in to its dynamic shared memory
// and ask how many CTA pairs (clusters of 2) can be resident at once -> persistent grid size
int gemm_prepare_impl(const void* kernel, int smem_bytes, int threads, int* max_pairs);
template <class Kern>
int gemm_prepare(Kern kern, int smem_bytes, int* max_pairs) {
    return gemm_prepare_impl(reinterpret_cast<const void*>(kern), smem_bytes, 384, max_pairs);
}

}  // namespace q3a
