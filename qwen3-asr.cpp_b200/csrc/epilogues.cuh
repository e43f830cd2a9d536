// Fused epilogues for gemm_tc_kernel: each thread owns one output row and gets 32 consecutive
// fp32 accumulator columns at a time (see the contract in gemm_tc.cuh).
#pragma once
#include "common.cuh"

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

// out[m][n] = act(acc + bias[n])  → 16-bit operand for the next GEMM
template <class OpT, bool kGelu>
struct EpiBiasActOp {
    OpT* out; int ldc; const float* bias;
    __device__ __forceinline__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const {
        OpT* row = out + (size_t)m * ldc + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float b[8], v[8];
                if (bias) load8(bias + n0 + g * 8, b);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float x = acc[g * 8 + j] + (bias ? b[j] : 0.0f);
                    v[j] = kGelu ? gelu_tanh(x) : x;
                }
                store8_op<OpT>(row + g * 8, v);
            }
        }
    }
};

// out[m][n] = acc + bias[n] (+ res[m][n])  → fp32 (residual stream / final features); res may alias out
template <bool kResidual>
struct EpiBiasF32 {
    float* out; int ldc; const float* bias; const float* res;
    __device__ __forceinline__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const {
        float* row = out + (size_t)m * ldc + n0;
        const float* rrow = kResidual ? res + (size_t)m * ldc + n0 : nullptr;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float b[8], r[8], v[8];
                if (bias) load8(bias + n0 + g * 8, b);
                if (kResidual) {
                    float4 a0 = *reinterpret_cast<const float4*>(rrow + g * 8);
                    float4 a1 = *(reinterpret_cast<const float4*>(rrow + g * 8) + 1);
                    r[0] = a0.x; r[1] = a0.y; r[2] = a0.z; r[3] = a0.w; r[4] = a1.x; r[5] = a1.y; r[6] = a1.z; r[7] = a1.w;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    v[j] = acc[g * 8 + j] + (bias ? b[j] : 0.0f) + (kResidual ? r[j] : 0.0f);
                store8_f32(row + g * 8, v);
            }
        }
    }
};

// conv_out: x[m][n] = acc + pe[pos[m]][n]   (sinusoidal table restarts per stem chunk,
// audio_encoder.cpp:400-404 / forced_aligner.cpp:718-735)
struct EpiPosEmbF32 {
    float* out; int ldc; const float* pe; const int* pos;
    __device__ __forceinline__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const {
        float* row = out + (size_t)m * ldc + n0;
        const float* prow = pe + (size_t)__ldg(pos + m) * ldc + n0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (g * 8 < nvalid) {
                float p[8], v[8];
                load8(prow + g * 8, p);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = acc[g * 8 + j] + p[j];
                store8_f32(row + g * 8, v);
            }
        }
    }
};

}  // namespace q3a
