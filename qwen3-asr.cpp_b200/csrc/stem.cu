// CUDA-core kernels of the encoder around the tcgen05 GEMMs: conv1 (K = 9 is too thin for UMMA) and
// LayerNorm.  Layout of the conv planes: stem.h.
#include "stem.h"

namespace q3a {

constexpr int kMelRows = 128;            // n_mel_bins; conv1 output has 64 mel rows

// grid (n_chunks, 2*(P2-1)): one CTA = one conv1 output time step t1 of one chunk = 64 mel positions x C
// channels.  A thread owns 8 consecutive channels (72 weights + 8 biases in registers) and walks the mel
// positions; consecutive threads write consecutive 16-byte groups of a channels-last row.
template <class OpT>
__global__ void __launch_bounds__(256)
conv1_kernel(const float* __restrict__ mel, const StemChunk* __restrict__ chunks, int P2,
             const float* __restrict__ clip_max, const float* __restrict__ w, const float* __restrict__ bias, int C,
             OpT* __restrict__ planes, long long plane_stride) {
    __shared__ float patch[3][kMelRows + 2];          // [dt][f + 1], f = -1 .. 128
    pdl_trigger();
    pdl_wait();
    const StemChunk ck = chunks[blockIdx.x];
    const int t1 = blockIdx.y;
    const int groups = C >> 3;
    const int g = threadIdx.x % groups, p0 = threadIdx.x / groups, ppi = blockDim.x / groups;
    const bool valid = t1 < ck.t1_valid;

    if (valid) {
        float thr = -INFINITY;
        const bool fuse = clip_max != nullptr;
        if (fuse) thr = __ldg(clip_max + ck.utt) - 8.0f;
        for (int i = threadIdx.x; i < 3 * (kMelRows + 2); i += blockDim.x) {
            const int dt = i / (kMelRows + 2), f = i - dt * (kMelRows + 2) - 1;
            const int t = 2 * t1 + dt - 1;
            float v = 0.0f;
            if (t >= 0 && t < ck.len && f >= 0 && f < kMelRows) {
                v = __ldg(mel + ck.mel_off + (long long)f * ck.row_stride + t);
                if (fuse) v = (fmaxf(v, thr) + 4.0f) * 0.25f;
            }
            patch[dt][f + 1] = v;
        }
    }
    float wr[8][9], br[8];
    if (valid && p0 < ppi) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            br[c] = __ldg(bias + g * 8 + c);
#pragma unroll
            for (int k = 0; k < 9; ++k) wr[c][k] = __ldg(w + (g * 8 + c) * 9 + k);
        }
    }
    __syncthreads();
    if (p0 >= ppi) return;

    const long long row = (long long)blockIdx.x * P2 + 1 + (t1 >> 1);
    for (int f1 = p0; f1 < kMelRows / 2; f1 += ppi) {
        float v[8];
        if (valid) {
            float x[9];                                // x[df*3 + dt]
#pragma unroll
            for (int df = 0; df < 3; ++df)
#pragma unroll
                for (int dt = 0; dt < 3; ++dt) x[df * 3 + dt] = patch[dt][2 * f1 + df];   // f = 2*f1 + df - 1
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float acc = br[c];
#pragma unroll
                for (int k = 0; k < 9; ++k) acc = fmaf(wr[c][k], x[k], acc);
                v[c] = gelu_tanh(acc);
            }
        } else {
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = 0.0f;
        }
        const int plane = ((t1 & 1) << 1) | (f1 & 1);
        OpT* dst = planes + plane * plane_stride + ((row * (kMelRows / 4) + (f1 >> 1)) * C + g * 8);
        uint4 u;
        u.x = OpTraits<OpT>::pack2(v[0], v[1]);
        u.y = OpTraits<OpT>::pack2(v[2], v[3]);
        u.z = OpTraits<OpT>::pack2(v[4], v[5]);
        u.w = OpTraits<OpT>::pack2(v[6], v[7]);
        *reinterpret_cast<uint4*>(dst) = u;
    }
}

template <class OpT>
int conv1_launch(const float* mel, const StemChunk* chunks_dev, int n_chunks, int P2, const float* clip_max,
                 const float* w, const float* bias, int C, OpT* planes, long long plane_stride, cudaStream_t stream) {
    if (n_chunks == 0) return Q3A_OK;
    const int groups = C / 8;
    if (C % 8 || groups > 256 || groups < 1) { set_error("conv1: conv_channels must be a multiple of 8, at most 2048"); return Q3A_ERR_ARG; }
    const int ppi = 256 / groups;
    const int threads = ((groups * ppi + 31) / 32) * 32;
    dim3 grid(n_chunks, 2 * (P2 - 1));
    Q3A_CUDA(launch_pdl(conv1_kernel<OpT>, grid, dim3(threads), 0, stream, mel, chunks_dev, P2, clip_max, w, bias, C, planes, plane_stride));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

// ---- LayerNorm ---------------------------------------------------------------------------------
// One warp per row, the row lives in registers (d <= 1024): mean, centred variance, affine, 16-bit store.
// ggml_norm accumulates in double; fp32 two-pass over <= 1024 values is within 1e-6 of it.
template <class OpT>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                 float eps, int T, int d, OpT* __restrict__ out) {
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= T) return;
    const int lane = threadIdx.x & 31;
    const int n4 = d >> 2;                               // float4 per row, <= 256
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * d);
    float4 v[8];
    float sum = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = lane + 32 * i;
        if (idx < n4) { v[i] = xr[idx]; sum += (v[i].x + v[i].y) + (v[i].z + v[i].w); }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum / d;
    float sq = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = lane + 32 * i;
        if (idx < n4) {
            v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
            sq += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = 1.0f / sqrtf(sq / d + eps);
    OpT* orow = out + (size_t)row * d;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = lane + 32 * i;
        if (idx < n4) {
            const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma) + idx);
            const float4 bt = __ldg(reinterpret_cast<const float4*>(beta) + idx);
            uint2 u;
            u.x = OpTraits<OpT>::pack2(v[i].x * rstd * gm.x + bt.x, v[i].y * rstd * gm.y + bt.y);
            u.y = OpTraits<OpT>::pack2(v[i].z * rstd * gm.z + bt.z, v[i].w * rstd * gm.w + bt.w);
            *reinterpret_cast<uint2*>(orow + idx * 4) = u;
        }
    }
}

template <class OpT>
int layernorm_launch(const float* x, const float* gamma, const float* beta, float eps, int T, int d, OpT* out,
                     cudaStream_t stream) {
    if (T == 0) return Q3A_OK;
    if (d % 4 || d > 1024) { set_error("layernorm: d_model must be a multiple of 4, at most 1024"); return Q3A_ERR_ARG; }
    Q3A_CUDA(launch_pdl(layernorm_kernel<OpT>, dim3(ceil_div(T, 8)), dim3(256), 0, stream, x, gamma, beta, eps, T, d, out));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

template int conv1_launch<__nv_bfloat16>(const float*, const StemChunk*, int, int, const float*, const float*,
                                         const float*, int, __nv_bfloat16*, long long, cudaStream_t);
template int conv1_launch<__half>(const float*, const StemChunk*, int, int, const float*, const float*, const float*,
                                  int, __half*, long long, cudaStream_t);
template int layernorm_launch<__nv_bfloat16>(const float*, const float*, const float*, float, int, int, __nv_bfloat16*,
                                             cudaStream_t);
template int layernorm_launch<__half>(const float*, const float*, const float*, float, int, int, __half*, cudaStream_t);

}  // namespace q3a
