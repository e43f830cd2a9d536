// Fused epilogues for gemm_tc_kernel: each thread owns one output row and gets 32 consecutive
// fp32 accumulator columns at a time (see the two contracts in gemm_tc.cuh).
#pragma once
#include "common.cuh"
#include "gemm_tc.cuh"

namespace q3a {

__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <class OpT>
__device__ __forceinline__ void store8_op(OpT* p, const float (&v)[8]) {
    uint4 u;
    u.x = OpTraits<OpT>::pack2(v[0], v[1]);
    u.y = OpTraits<OpT>::pack2(v[2], v[3]);
    u.z = OpTraits<OpT>::pack2(v[4], v[5]);
    u.w = OpTraits<OpT>::pack2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void store8_f32(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *(reinterpret_cast<float4*>(p) + 1) = make_float4(v[4], v[5], v[6], v[7]);
}

// one per-column constant per lane (gemm_tc.cuh EpiCols): lane j of the warp holds column n0 + j's value
struct Cols1 { float a; };
struct Cols2 { float a, b; };
__device__ __forceinline__ float col_of(float lane_value, int j) { return __shfl_sync(0xffffffffu, lane_value, j); }

// ---- staged epilogues (Linear layers): the kernel stores v through smem + TMA ---------------------
// out[m][n] = act(acc + bias[n])  -> 16-bit operand for the next GEMM
template <class OpT, bool kGelu>
struct EpiBiasActOp {
    static constexpr bool kStaged = true, kResidual = false;
    static constexpr int kOutBytes = 2;
    const float* bias;
    using Cols = Cols1;
    __device__ __forceinline__ static uint32_t pack2(float a, float b) { return OpTraits<OpT>::pack2(a, b); }
    __device__ __forceinline__ Cols cols(int n0, int nvalid, int lane) const {
        return Cols{bias && lane < nvalid ? __ldg(bias + n0 + lane) : 0.0f};
    }
    __device__ __forceinline__ void apply(int, int, int nvalid, float (&v)[32], const Cols& cv) const {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float x = v[g * 8 + j] + col_of(cv.a, g * 8 + j);
                    v[g * 8 + j] = kGelu ? gelu_tanh(x) : x;
                }
            }
        }
    }
};

// out[m][n] = acc + bias[n] (+ out[m][n])  -> fp32 (residual stream updated in place / final features)
template <bool kRes>
struct EpiBiasF32 {
    static constexpr bool kStaged = true, kResidual = kRes;
    static constexpr int kOutBytes = 4;
    const float* bias;
    int prefetch = 1;        // residual epilogues: L2-prefetch the next tile's residual rows (gemm_tc.cuh)
    using Cols = Cols1;
    __device__ __forceinline__ Cols cols(int n0, int nvalid, int lane) const {
        return Cols{bias && lane < nvalid ? __ldg(bias + n0 + lane) : 0.0f};
    }
    __device__ __forceinline__ void apply(int, int, int nvalid, float (&v)[32], const Cols& cv) const {
        if (!bias) return;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[g * 8 + j] += col_of(cv.a, g * 8 + j);
            }
        }
    }
};

// conv_out: x[m][n] = acc + pe[pos[m]][n]   (sinusoidal table restarts per stem chunk,
// audio_encoder.cpp:400-404 / forced_aligner.cpp:718-735); pe == nullptr -> plain store (conv-only mode)
struct EpiPosEmbF32 {
    static constexpr bool kStaged = true, kResidual = false;
    static constexpr int kOutBytes = 4;
    const float* pe; const int* pos; int ld; int M;
    __device__ __forceinline__ void apply(int m, int n0, int nvalid, float (&v)[32]) const {
        if (!pe || m >= M) return;
        const float* prow = pe + (size_t)__ldg(pos + m) * ld + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float p[8];
                load8(prow + g * 8, p);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[g * 8 + j] += p[j];
            }
        }
    }
};

// ---- LayerNorm folded into the GEMMs (audio_encoder.cpp:441-448,502-508,530-534 without a LayerNorm kernel) -----------
// h = LN(x) * gamma + beta feeds a weight matmul y = h W^T + b.  With W' = gamma o W (folded at load), s[n] = sum_k W'[n][k]
// and c[n] = b[n] + sum_k beta[k] W[n][k]:      y[m][n] = rstd[m] * (sum_k x[m][k] W'[n][k] - mean[m] * s[n]) + c[n],
// so the A operand is the residual stream itself rounded to 16 bits and the statistics enter in the epilogue.  The
// producer of x (conv_out, out-projection, FFN-down) emits that 16-bit copy and per-row partial (sum, sum of squares)
// next to the f32 stream ("dual" epilogues); partials are added here in a fixed order, so results stay bit-identical
// whatever the batch around a row looks like.
//
// producer side: f32 out = acc + bias (+ out), plus the 16-bit copy and the partials (kernel: EpiDual)
template <class OpT, bool kRes>
struct EpiBiasF32Dual {
    static constexpr bool kStaged = true, kResidual = kRes;
    static constexpr int kOutBytes = 4;
    const float* bias;
    float2* stats; int n_part;
    int prefetch = 1;
    __device__ __forceinline__ static uint32_t pack2(float a, float b) { return OpTraits<OpT>::pack2(a, b); }
    using Cols = Cols1;
    __device__ __forceinline__ Cols cols(int n0, int nvalid, int lane) const {
        return Cols{bias && lane < nvalid ? __ldg(bias + n0 + lane) : 0.0f};
    }
    __device__ __forceinline__ void apply(int, int, int nvalid, float (&v)[32], const Cols& cv) const {
        if (!bias) return;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[g * 8 + j] += col_of(cv.a, g * 8 + j);
            }
        }
    }
};
template <class OpT>
struct EpiPosEmbF32Dual {
    static constexpr bool kStaged = true, kResidual = false;
    static constexpr int kOutBytes = 4;
    const float* pe; const int* pos; int ld; int M;
    float2* stats; int n_part;
    __device__ __forceinline__ static uint32_t pack2(float a, float b) { return OpTraits<OpT>::pack2(a, b); }
    __device__ __forceinline__ void apply(int m, int n0, int nvalid, float (&v)[32]) const {
        if (!pe || m >= M) return;
        const float* prow = pe + (size_t)__ldg(pos + m) * ld + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float p[8];
                load8(prow + g * 8, p);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[g * 8 + j] += p[j];
            }
        }
    }
};
template <class OpT, bool kRes> struct EpiDual<EpiBiasF32Dual<OpT, kRes>> { static constexpr bool value = true; };
template <class OpT> struct EpiDual<EpiPosEmbF32Dual<OpT>> { static constexpr bool value = true; };

// consumer side: 16-bit out = act(rstd * (acc - mean * s[n]) + c[n])
struct LnRowState { float mean, rstd; };
__device__ __forceinline__ LnRowState ln_row_state(const float2* stats, int n_part, int M, float inv_d, float eps, int rms, int m) {
    float sum = 0.0f, sq = 0.0f;
    if (m < M) {
        const float2* p = stats + (size_t)m * n_part;
        for (int i = 0; i < n_part; ++i) { const float2 t = __ldg(p + i); sum += t.x; sq += t.y; }
    }
    const float mean = rms ? 0.0f : sum * inv_d;
    const float var = fmaxf(sq * inv_d - mean * mean, 0.0f);
    return LnRowState{mean, 1.0f / sqrtf(var + eps)};
}
template <class OpT, bool kGelu>
struct EpiLnFoldActOp {
    static constexpr bool kStaged = true, kResidual = false;
    static constexpr int kOutBytes = 2;
    using RowState = LnRowState;
    const float* c;          // [N] bias + beta . W
    const float* s;          // [N] column sums of the folded (rounded) weights
    const float2* stats;     // [M][n_part]
    int n_part, M;
    float inv_d, eps;
    int rms;                 // 1: RMSNorm (no mean subtraction; the aligner's decoder, forced_aligner.cpp:1012-1013)
    __device__ __forceinline__ static uint32_t pack2(float a, float b) { return OpTraits<OpT>::pack2(a, b); }
    __device__ __forceinline__ RowState row_state(int m) const { return ln_row_state(stats, n_part, M, inv_d, eps, rms, m); }
    using Cols = Cols2;      // a = c[n], b = s[n]
    __device__ __forceinline__ Cols cols(int n0, int nvalid, int lane) const {
        return lane < nvalid ? Cols{__ldg(c + n0 + lane), __ldg(s + n0 + lane)} : Cols{0.0f, 0.0f};
    }
    __device__ __forceinline__ void apply(int, int, int nvalid, float (&v)[32], const RowState& rs, const Cols& cv) const {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float x = fmaf(rs.rstd, fmaf(-rs.mean, col_of(cv.b, g * 8 + j), v[g * 8 + j]), col_of(cv.a, g * 8 + j));
                    v[g * 8 + j] = kGelu ? gelu_tanh(x) : x;
                }
            }
        }
    }
};

// same fold, f32 output (the decoder's classification head: logits = rmsnorm(x) W^T + b, forced_aligner.cpp:1070-1076)
struct EpiLnFoldF32 {
    static constexpr bool kStaged = true, kResidual = false;
    static constexpr int kOutBytes = 4;
    using RowState = LnRowState;
    const float* c; const float* s; const float2* stats;
    int n_part, M;
    float inv_d, eps;
    int rms;
    __device__ __forceinline__ RowState row_state(int m) const { return ln_row_state(stats, n_part, M, inv_d, eps, rms, m); }
    __device__ __forceinline__ void apply(int, int n0, int nvalid, float (&v)[32], const RowState& rs) const {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float cc[8], ss[8];
                load8(c + n0 + g * 8, cc);
                load8(s + n0 + g * 8, ss);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[g * 8 + j] = fmaf(rs.rstd, fmaf(-rs.mean, ss[j], v[g * 8 + j]), cc[j]);
            }
        }
    }
};

// ---- stem epilogues (layout: stem.cuh) -------------------------------------------------------
// conv2: GELU(acc + bias) -> the parity-split planes conv3 reads.  GEMM row m = R_out*OF + of with
// R_out = chunk*P + 1 + ot (row chunk*P is the zero pad row of the implicit GEMM and is skipped).
// Positions past the chunk's valid length are written as zeros so that a shorter tail chunk sees the
// convolution's zero padding instead of stale data.
template <class OpT>
struct EpiConvToPlanes {
    static constexpr bool kStaged = false, kResidual = false;
    static constexpr int kOutBytes = 4;
    OpT* planes;            // next conv's input: [4][R_cap_next][OF/2][C]
    const float* bias;
    const int* ot_valid;    // per chunk: valid output time steps of this conv
    int C, OF, P, P_next;
    long long plane_stride; // R_cap_next * (OF/2) * C elements
    __device__ __forceinline__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const {
        const int r_out = m / OF, of = m - r_out * OF;
        const int chunk = r_out / P, j = r_out - chunk * P;
        if (j == 0) return;
        const int ot = j - 1;
        const bool valid = ot < __ldg(ot_valid + chunk);
        const int plane = ((ot & 1) << 1) | (of & 1);
        const long long row = (long long)chunk * P_next + 1 + (ot >> 1);
        OpT* dst = planes + plane * plane_stride + (row * (OF >> 1) + (of >> 1)) * C + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float b[8], v[8];
                load8(bias + n0 + g * 8, b);
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) v[jj] = valid ? gelu_tanh(acc[g * 8 + jj] + b[jj]) : 0.0f;
                store8_op<OpT>(dst + g * 8, v);
            }
        }
    }
};

// conv3: GELU(acc + bias) -> token-major feature rows for conv_out.  Feature order inside a row is
// of*C + channel (the conv_out weight columns are permuted to match at load time; the reference's
// order is channel*16 + of, audio_encoder.cpp:139-142).  Only the tokens the mode keeps are written.
template <class OpT>
struct EpiConvToFeat {
    static constexpr bool kStaged = false, kResidual = false;
    static constexpr int kOutBytes = 4;
    OpT* feat;              // [tokens][OF*C]
    const float* bias;
    const int* n_tok;       // per chunk: tokens kept
    const int* tok_base;    // per chunk: first token row
    int C, OF, P;
    __device__ __forceinline__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const {
        const int r_out = m / OF, of = m - r_out * OF;
        const int chunk = r_out / P, j = r_out - chunk * P;
        if (j == 0) return;
        const int ot = j - 1;
        if (ot >= __ldg(n_tok + chunk)) return;
        OpT* dst = feat + ((long long)(__ldg(tok_base + chunk) + ot) * OF + of) * C + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float b[8], v[8];
                load8(bias + n0 + g * 8, b);
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) v[jj] = gelu_tanh(acc[g * 8 + jj] + b[jj]);
                store8_op<OpT>(dst + g * 8, v);
            }
        }
    }
};

}  // namespace q3a
