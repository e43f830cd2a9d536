// Shared host/device helpers for libq3a (B200 / sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace q3a {

// ---- error plumbing (C ABI returns int codes; message kept per thread) -------------------
void set_error(const std::string& msg);          // defined in capi.cu
const char* last_error();

#define Q3A_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            ::q3a::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" +     \
                             __FILE__ + ":" + std::to_string(__LINE__) + ")");               \
            return Q3A_ERR_CUDA;                                                             \
        }                                                                                    \
    } while (0)

enum : int {
    Q3A_OK = 0,
    Q3A_ERR_ARG = 1,
    Q3A_ERR_CUDA = 2,
    Q3A_ERR_IO = 3,
    Q3A_ERR_FORMAT = 4,
    Q3A_ERR_STATE = 5,
};

// operand element type fed to the tensor cores (accumulation is always fp32)
enum : int { OP_BF16 = 0, OP_F16 = 1 };

template <typename T> struct OpTraits;
template <> struct OpTraits<__nv_bfloat16> {
    static constexpr int kFmt = 1;   // UMMA F16F32Format::BF16
    __device__ __forceinline__ static __nv_bfloat16 from(float v) { return __float2bfloat16_rn(v); }
    __device__ __forceinline__ static float to(__nv_bfloat16 v) { return __bfloat162float(v); }
    __device__ __forceinline__ static uint32_t pack2(float a, float b) {
        __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&p);
    }
    __device__ __forceinline__ static float2 unpack2(uint32_t u) {
        __nv_bfloat162 p = *reinterpret_cast<__nv_bfloat162*>(&u);
        return __bfloat1622float2(p);
    }
};
template <> struct OpTraits<__half> {
    static constexpr int kFmt = 0;   // UMMA F16F32Format::F16
    __device__ __forceinline__ static __half from(float v) { return __float2half_rn(v); }
    __device__ __forceinline__ static float to(__half v) { return __half2float(v); }
    __device__ __forceinline__ static uint32_t pack2(float a, float b) {
        __half2 p = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&p);
    }
    __device__ __forceinline__ static float2 unpack2(uint32_t u) {
        __half2 p = *reinterpret_cast<__half2*>(&u);
        return __half22float2(p);
    }
};

// ggml_gelu is the tanh form (audio_encoder.cpp:112 → ggml_gelu); MUFU.TANH is accurate to
// ~2^-11 relative, the same order as the f16 table ggml's CPU path reads it through.
__device__ __forceinline__ float gelu_tanh(float x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * x * fmaf(k1 * x, x, 1.0f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    return 0.5f * x * (1.0f + t);
}

__host__ __device__ constexpr int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace q3a
