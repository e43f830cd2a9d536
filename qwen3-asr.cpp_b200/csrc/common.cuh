// Shared host/device helpers for libq3a (B200 / sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/q3a.h"

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

#define Q3A_TRY(expr)                     \
    do {                                  \
        int _rc = (expr);                 \
        if (_rc != Q3A_OK) return _rc;    \
    } while (0)

// operand element type fed to the tensor cores (accumulation is always fp32)
enum : int { OP_BF16 = Q3A_OP_BF16, OP_F16 = Q3A_OP_F16 };

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

// ggml_gelu is the tanh form (audio_encoder.cpp:112 -> ggml_gelu); MUFU.TANH is accurate to
// ~2^-11 relative, the same order as the f16 table ggml's CPU path reads it through.
__device__ __forceinline__ float gelu_tanh(float x) {
    // 5 FP32 instructions + one MUFU (the epilogues of the FFN-up GEMM and of the conv stem are issue-bound on this):
    // u = x (k0 + k0 k1 x^2), y = x/2 + x/2 tanh(u)
    const float k0 = 0.7978845608028654f, k0k1 = 0.7978845608028654f * 0.044715f;
    const float u = x * fmaf(k0k1, x * x, k0);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}

__host__ __device__ constexpr int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ constexpr int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ constexpr int64_t round_up64(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// ---- small RAII helpers (host) ---------------------------------------------------------------
// Growable device buffer: keeps its allocation between calls, grows geometrically.
// With Q3A_GUARD set in the environment every allocation is exact-size and sits between two 4-KiB bands of a
// known byte pattern, registered with guard_check(): the library's own bounds check (q3a_debug_check_guards),
// since an out-of-range store by any kernel lands in a band instead of in the growth slack.
constexpr size_t kGuardBytes = 4096;
bool guard_enabled();
void guard_register(void* base, size_t user_bytes);
void guard_forget(void* base);
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* base = nullptr;            // what cudaMalloc returned (== p unless guarded)
    int reserve(size_t bytes) {
        if (bytes <= cap) return Q3A_OK;
        release();
        if (guard_enabled()) {
            const size_t user = (bytes + 255) & ~size_t(255);
            Q3A_CUDA(cudaMalloc(&base, user + 2 * kGuardBytes));
            Q3A_CUDA(cudaMemset(base, 0xA5, kGuardBytes));
            Q3A_CUDA(cudaMemset(static_cast<char*>(base) + kGuardBytes + user, 0xA5, kGuardBytes));
            p = static_cast<char*>(base) + kGuardBytes;
            cap = user;
            guard_register(base, user);
            return Q3A_OK;
        }
        size_t want = bytes + bytes / 8 + 256;
        Q3A_CUDA(cudaMalloc(&base, want));
        p = base;
        cap = want;
        return Q3A_OK;
    }
    void release() {
        if (base) {
            if (p != base) guard_forget(base);
            cudaFree(base);
        }
        p = base = nullptr;
        cap = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Ring of pinned host staging slots for per-call descriptor tables: a slot is reused only after the
// async copy that read it has completed (event), so back-to-back async calls never race.
struct PinnedRing {
    static constexpr int kSlots = 4;
    void* host[kSlots] = {nullptr, nullptr, nullptr, nullptr};
    size_t cap[kSlots] = {0, 0, 0, 0};
    cudaEvent_t ev[kSlots] = {nullptr, nullptr, nullptr, nullptr};
    bool pending[kSlots] = {false, false, false, false};
    int next = 0;
    // returns a host pointer of at least `bytes`; call mark(stream) after enqueueing the copy
    int acquire(size_t bytes, void** out) {
        int s = next;
        if (pending[s]) { Q3A_CUDA(cudaEventSynchronize(ev[s])); pending[s] = false; }
        if (bytes > cap[s]) {
            if (host[s]) cudaFreeHost(host[s]);
            host[s] = nullptr; cap[s] = 0;
            Q3A_CUDA(cudaMallocHost(&host[s], bytes + 4096));
            cap[s] = bytes + 4096;
        }
        if (!ev[s]) Q3A_CUDA(cudaEventCreateWithFlags(&ev[s], cudaEventDisableTiming));
        *out = host[s];
        return Q3A_OK;
    }
    int mark(cudaStream_t st) {
        int s = next;
        Q3A_CUDA(cudaEventRecord(ev[s], st));
        pending[s] = true;
        next = (next + 1) % kSlots;
        return Q3A_OK;
    }
    void release() {
        for (int s = 0; s < kSlots; ++s) {
            if (pending[s]) cudaEventSynchronize(ev[s]);
            if (host[s]) cudaFreeHost(host[s]);
            if (ev[s]) cudaEventDestroy(ev[s]);
            host[s] = nullptr; ev[s] = nullptr; cap[s] = 0; pending[s] = false;
        }
    }
};

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------
// Every kernel of the path is launched with programmatic stream serialization: its CTAs may become resident
// while the previous kernel of the stream is still draining, run their prologue (barrier init, TMEM
// allocation, descriptor prefetch) and then block in pdl_wait() until the previous grid has completed and
// its writes are visible.  pdl_trigger() lets the next kernel do the same with respect to this one.  A kernel
// must not touch global memory another kernel of the stream writes, nor write global memory, before pdl_wait().
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    static const bool off = getenv("Q3A_NO_PDL") != nullptr;      // debugging aid: plain stream order
    cfg.attrs = at; cfg.numAttrs = off ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---- round-2 code paths and their defaults ------------------------------------------------------------------------
// Each of these replaces a round-1 path that is validated on hardware.  A path becomes the default (constant below set to
// true) once its parity tests have passed on a B200; until then it is opt-in through the "on" variable.  The "off"
// variable always wins, so A/B runs work in both states.  tests/validated_on_b200.json mirrors the constants.
constexpr bool kDefaultAttnV2 = false;       // attention2.cu instead of attention.cu          Q3A_ATTN_V2 / Q3A_ATTN_V1
                                             // (validated r2a: all 76 parity cases green, but 157 us against 133 us per
                                             //  layer on the bench shape -- DESIGN.md 4.4 -- so the first kernel stays default)
constexpr int kDefaultDualEpiWarps = 4;      // epilogue warps of the residual GEMMs with a 16-bit copy   Q3A_DUAL_EPI_WARPS=4|8
                                             // (validated r2s: bit-identical, FFN-down 149.7 -> 134.6 us: a fifth ring stage)
constexpr int kDefaultResEpiWarps = 8;       // the same for the plain residual GEMM                      Q3A_RES_EPI_WARPS=4|8
constexpr bool kDefaultLnFold = true;        // LayerNorm folded into the GEMMs                Q3A_LN_FOLD / Q3A_NO_LN_FOLD
constexpr bool kDefaultQ8Fused = false;      // Q8_0 weights dequantised inside the GEMM       Q3A_Q8_FUSED / Q3A_Q8_DEQUANT_AT_LOAD
                                             // (validated bit-identical, halves the weight arena, but 36.0 against 22.3 ms per
                                             //  64 x 30 s and 3.6 against 1.9 ms for one clip -- DESIGN.md 4.6 -- so opt-in)
constexpr bool kDefaultConvCbMajor = true;   // channel-block-major K order in conv2 / conv3   Q3A_CONV_CB_MAJOR / Q3A_CONV_TAP_MAJOR
inline bool feature_enabled(const char* env_on, const char* env_off, bool dflt) {
    if (getenv(env_off)) return false;
    if (getenv(env_on)) return true;
    return dflt;
}

// make `device` current and check that it is an sm_100 part (there is no other code path)
int use_device(int device);
// opt a kernel in to `bytes` of dynamic shared memory, once per (kernel, current device); thread-safe (several encoders
// on several devices may launch from their own host threads)
int kernel_smem_optin(const void* kernel, int bytes);
// multiprocessor count of the current device (cached per device)
int sm_count_cached();

// counts our kernel launches (reported through q3a_timings::total_launches)
extern thread_local int g_launch_counter;

}  // namespace q3a
