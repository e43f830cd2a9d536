// Stride-2 3x3 convolutions of the stem (conv2, conv3) as implicit GEMM on the tcgen05 kernel.
// Replaces ggml_conv_2d = im2col + mul_mat (reference audio_encoder.cpp:116,123;
// forced_aligner.cpp:645,652) without materialising im2col: the A tiles are TMA boxes over
// parity-split channels-last planes (stem.cuh), zero padding comes from TMA's out-of-bounds fill.
#include "gemm.h"
#include "gemm_tc.cuh"
#include "epilogues.cuh"
#include "tmap.h"

namespace q3a {

template <int BLOCK_N, int kFmt, class Epi>
static int launch_conv(const ConvArgs& a, const Epi& epi, cudaStream_t stream) {
    using S = GemmSmemFor<BLOCK_N, Epi>;
    const int c_blocks = ceil_div(a.C, kBlockK);
    const int cpad = c_blocks * kBlockK;
    CUtensorMap ta, tb;
    {
        uint64_t dims[4] = {(uint64_t)a.C, (uint64_t)a.Fh, (uint64_t)(a.R_map > 0 ? a.R_map : a.R_cap), 4};
        uint64_t strides[3] = {(uint64_t)a.C * 2, (uint64_t)a.Fh * a.C * 2, (uint64_t)a.R_cap * a.Fh * a.C * 2};
        uint32_t box[4] = {64, (uint32_t)a.Fh, (uint32_t)(kBlockM / a.Fh), 1};
        Q3A_TRY(make_tmap_16bit(&ta, a.planes, 4, dims, strides, box));
    }
    Q3A_TRY(make_tmap_2d(&tb, a.w, (uint64_t)a.N, (uint64_t)(9 * cpad), (uint64_t)(9 * cpad), BLOCK_N / 2));
    auto kern = gemm_tc_kernel<BLOCK_N, kFmt, LoaderConv, Epi>;
    int pairs = 0;
    Q3A_TRY(gemm_prepare(kern, S::kTotal, &pairs));
    LoaderConv loader;
    loader.num_k_blocks = 9 * c_blocks;
    loader.c_blocks = c_blocks;
    loader.rows_per_tile = kBlockM / a.Fh;
    loader.last_block_steps = ceil_div(a.C - (c_blocks - 1) * kBlockK, kUmmaK);
    loader.cb_major = feature_enabled("Q3A_CONV_CB_MAJOR", "Q3A_CONV_TAP_MAJOR", kDefaultConvCbMajor) ? 1 : 0;
    const int M = a.R_used * a.Fh;
    const int tiles = ceil_div(M, 2 * kBlockM) * ceil_div(a.N, BLOCK_N);
    const int grid = 2 * (tiles < pairs ? tiles : pairs);
    Q3A_CUDA(launch_pdl(kern, dim3(grid), dim3(kGemmThreads), S::kTotal, stream, ta, tb, ta /*unused: direct epilogue*/, ta, M, a.N, loader, epi, Q8Scales{}));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

template <int BLOCK_N, int kFmt>
static int dispatch_conv(bool to_feat, const ConvArgs& a, cudaStream_t stream) {
    using OpT = typename std::conditional<kFmt == 1, __nv_bfloat16, __half>::type;
    if (to_feat) {
        EpiConvToFeat<OpT> e{(OpT*)a.feat, a.bias, a.n_tok, a.tok_base, a.N, a.Fh, a.P};
        return launch_conv<BLOCK_N, kFmt>(a, e, stream);
    }
    EpiConvToPlanes<OpT> e{(OpT*)a.out_planes, a.bias, a.ot_valid, a.N, a.Fh, a.P, a.P_next, a.out_plane_stride};
    return launch_conv<BLOCK_N, kFmt>(a, e, stream);
}

int gemm_conv(bool to_feat, int fmt, const ConvArgs& a, cudaStream_t stream) {
    if (a.R_used <= 0) return Q3A_OK;
    if (a.C % 8 || a.N % 8 || a.Fh <= 0 || (kBlockM % a.Fh) != 0) {
        set_error("gemm_conv: channels must be multiples of 8 and the mel half-width must divide 128");
        return Q3A_ERR_ARG;
    }
    const bool wide = (a.N % 240 == 0);
    if (fmt == OP_F16) return wide ? dispatch_conv<240, 0>(to_feat, a, stream) : dispatch_conv<128, 0>(to_feat, a, stream);
    return wide ? dispatch_conv<240, 1>(to_feat, a, stream) : dispatch_conv<128, 1>(to_feat, a, stream);
}

}  // namespace q3a
