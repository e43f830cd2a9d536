// conv1 of the stem on the tensor cores (sm_100a): 3x3 stride-2 pad-1 convolution, 1 -> C channels, + bias +
// tanh-GELU, written as the parity-split planes conv2 reads (layout: stem.h).  Replaces ggml_conv_2d #1 + add +
// gelu (reference audio_encoder.cpp:105-112, forced_aligner.cpp:638-644) and, on the fused path, the
// clamp(max-8) / (x+4)/4 tail of log_mel_spectrogram (mel_spectrogram.cpp:615-625).
//
// The contraction is only 9 taps deep, so the kernel is bound by the epilogue (GELU + pack + store of C values
// per position), not by the MMA; the point of tcgen05 here is to take the 9 FMAs per output off the CUDA cores
// (the CUDA-core version, conv1_kernel in stem.cu, is issue-bound at 4x the HBM-write time).  To keep fp32-level
// accuracy with 16-bit operands, input and weight are split into f16 hi + lo parts and the K axis is laid out as
//     k in [0,9): x_hi*w_hi      k in [9,18): x_lo*w_hi      k in [18,27): x_hi*w_lo      k in [27,32): 0
// (the dropped x_lo*w_lo term is < 2^-22 relative).
//
// One tile = one (chunk, time-row a) = 128 positions: TMEM lane r <-> quadrant q = r/32 = plane (t1&1, f1&1),
// i = r%32: t1 = 2a + (q>>1), f1 = 2i + (q&1); the 32 lanes of a quadrant are 32 consecutive rows of plane q, so
// each epilogue warp stores its 32x32 block with one TMA store.
//   warps 0..3   build the A tile (one position per thread: 9 mel loads, clamp/scale, hi/lo split, swizzled
//                K-major row) into a 3-stage ring
//   warp 4       TMEM allocator + MMA issuer: per tile two N-halves (256 + C-256 columns) = two accumulator
//                buffers, so the epilogue of one half overlaps the MMAs of the next
//   warps 5..20  epilogue (four per TMEM lane quadrant, interleaved over the 32-channel chunks): tcgen05.ld -> bias + GELU -> 16-bit -> smem staging -> TMA store
#include "stem.h"
#include "tc.cuh"
#include "tmap.h"

#include <cstring>

namespace q3a {

constexpr int kC1EpiWarps = 16;                  // 4 per scheduler: the epilogue is what the kernel is made of
constexpr int kC1Threads = (5 + kC1EpiWarps) * 32;
constexpr int kC1Stages = 3;
constexpr int kC1ATile = 128 * 128;             // 128 rows x 128-byte swizzle rows (K = 32 x f16 uses 64 B)
constexpr int kC1EpiBuf = 32 * 64;              // 32 rows x 32 channels x 16 bit
constexpr int kMelRowsC1 = 128;

__host__ __device__ constexpr int c1_smem_bytes(int C) {
    return C * 128 /*W image*/ + kC1Stages * kC1ATile + kC1EpiWarps * 2 * kC1EpiBuf + ((C * 4 + 255) / 256) * 256 /*bias*/ + 256 + 1024;
}

__device__ __forceinline__ float gelu_tanh6(float x) {
    const float x2 = x * x;
    const float w = fmaf(x2, 0.7978845608028654f * 0.044715f, 0.7978845608028654f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(x * w));
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}

template <class OpT>
__global__ void __launch_bounds__(kC1Threads, 1)
conv1_tc_kernel(const float* __restrict__ mel, const StemChunk* __restrict__ chunks, int n_tiles, int tiles_per_chunk,
                int P2, const float* __restrict__ clip_max, const uint4* __restrict__ w_image,
                const float* __restrict__ bias, int C, const __grid_constant__ CUtensorMap tmap_out, int rows_per_plane) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - tc::smem_u32(smem_raw));
    const uint32_t sW = base;
    const uint32_t sA = sW + C * 128;
    const uint32_t sE = sA + kC1Stages * kC1ATile;
    const uint32_t sBias = sE + kC1EpiWarps * 2 * kC1EpiBuf;
    const uint32_t bars = sBias + ((C * 4 + 255) / 256) * 256;
    uint8_t* gA = gen + C * 128;
    uint8_t* gE = gA + kC1Stages * kC1ATile;
    float* gBias = reinterpret_cast<float*>(gE + kC1EpiWarps * 2 * kC1EpiBuf);
    auto a_full = [&](int s) { return bars + 8u * s; };
    auto a_empty = [&](int s) { return bars + 8u * (3 + s); };
    auto tfull = [&](int h) { return bars + 8u * (6 + h); };
    auto tempty = [&](int h) { return bars + 8u * (8 + h); };
    const uint32_t tmem_slot = bars + 80;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + (bars - base) + 80);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N0 = C < 256 ? C : 256, N1 = C - N0;
    const int n_halves = N1 > 0 ? 2 : 1;

    // one-time: weight image and bias into smem, barriers, TMEM (weights are constants: legal before pdl_wait)
    pdl_trigger();
    for (int i = tid; i < C * 8; i += kC1Threads) reinterpret_cast<uint4*>(gen)[i] = __ldg(w_image + i);
    for (int i = tid; i < C; i += kC1Threads) gBias[i] = __ldg(bias + i);
    if (tid == 0) {
        tc::prefetch_tmap(&tmap_out);
        for (int s = 0; s < kC1Stages; ++s) { tc::mbar_init(a_full(s), 4); tc::mbar_init(a_empty(s), 1); }
        for (int h = 0; h < 2; ++h) { tc::mbar_init(tfull(h), 1); tc::mbar_init(tempty(h), kC1EpiWarps); }
        tc::fence_barrier_init();
    }
    if (warp == 4) tc::tmem_alloc<512>(tmem_slot);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();

    if (warp < 4) {
        // ===================== A-tile producers =====================
        const int q = warp, i = lane;
        const bool fuse = clip_max != nullptr;
        int s = 0; uint32_t ph = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int chunk = tile / tiles_per_chunk, a = tile - chunk * tiles_per_chunk;
            const StemChunk ck = chunks[chunk];
            const int t1 = 2 * a + (q >> 1), f1 = 2 * i + (q & 1);
            float x[9];
#pragma unroll
            for (int k = 0; k < 9; ++k) x[k] = 0.0f;
            if (t1 < ck.t1_valid) {
                const float thr = fuse ? __ldg(clip_max + ck.utt) - 8.0f : 0.0f;
#pragma unroll
                for (int df = 0; df < 3; ++df) {
                    const int f = 2 * f1 + df - 1;
#pragma unroll
                    for (int dt = 0; dt < 3; ++dt) {
                        const int t = 2 * t1 + dt - 1;
                        if (f >= 0 && f < kMelRowsC1 && t >= 0 && t < ck.len) {
                            float v = __ldg(mel + ck.mel_off + (long long)f * ck.row_stride + t);
                            if (fuse) v = (fmaxf(v, thr) + 4.0f) * 0.25f;
                            x[df * 3 + dt] = v;
                        }
                    }
                }
            }
            // K layout: [x_hi(9) | x_lo(9) | x_hi(9) | 0(5)] as f16, packed two per word
            unsigned short hb[32];
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const __half hi = __float2half_rn(x[k]);
                const __half lo = __float2half_rn(x[k] - __half2float(hi));
                hb[k] = __half_as_ushort(hi); hb[9 + k] = __half_as_ushort(lo); hb[18 + k] = __half_as_ushort(hi);
            }
#pragma unroll
            for (int k = 27; k < 32; ++k) hb[k] = 0;
            tc::mbar_wait(a_empty(s), ph ^ 1);
            uint8_t* row = gA + s * kC1ATile + (q * 32 + i) * 128;
            const int r7 = (q * 32 + i) & 7;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint4 u;
                u.x = (uint32_t)hb[8 * j + 0] | ((uint32_t)hb[8 * j + 1] << 16);
                u.y = (uint32_t)hb[8 * j + 2] | ((uint32_t)hb[8 * j + 3] << 16);
                u.z = (uint32_t)hb[8 * j + 4] | ((uint32_t)hb[8 * j + 5] << 16);
                u.w = (uint32_t)hb[8 * j + 6] | ((uint32_t)hb[8 * j + 7] << 16);
                *reinterpret_cast<uint4*>(row + ((j ^ r7) << 4)) = u;
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(a_full(s));
            if (++s == kC1Stages) { s = 0; ph ^= 1; }
        }
    } else if (warp == 4) {
        // ===================== MMA issuer =====================
        const uint64_t desc_hi = tc::smem_desc_kmajor_sw128(0);
        const uint32_t idesc0 = tc::make_idesc(0, 128, N0, 0, 0);
        const uint32_t idesc1 = tc::make_idesc(0, 128, N1 > 0 ? N1 : 16, 0, 0);
        int s = 0; uint32_t ph = 0, tph = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            tc::mbar_wait(a_full(s), ph);
            tc::fence_after_sync();
            const uint64_t a_desc = desc_hi | (uint64_t)(((sA + s * kC1ATile) >> 4) & 0x3FFF);
            for (int h = 0; h < n_halves; ++h) {
                tc::mbar_wait(tempty(h), tph ^ 1);
                tc::fence_after_sync();
                if (tc::elect_one()) {
                    const uint64_t b_desc = desc_hi | (uint64_t)(((sW + h * 256 * 128) >> 4) & 0x3FFF);
                    const uint32_t idesc = h ? idesc1 : idesc0;
                    tc::mma_f16_ss(tmem + h * 256, a_desc, b_desc, idesc, 0);
                    tc::mma_f16_ss(tmem + h * 256, a_desc + 2, b_desc + 2, idesc, 1);
                    tc::mma_commit(tfull(h));
                    if (h == n_halves - 1) tc::mma_commit(a_empty(s));
                }
                __syncwarp();
            }
            tph ^= 1;
            if (++s == kC1Stages) { s = 0; ph ^= 1; }
        }
    } else {
        // ===================== epilogue =====================
        const int ew = warp - 5;
        const int q = warp & 3;
        const int hf = ew >> 2;
        const uint32_t my_buf = sE + ew * 2 * kC1EpiBuf;
        uint8_t* my_buf_gen = gE + ew * 2 * kC1EpiBuf;
        uint32_t seq = 0, tph = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int chunk = tile / tiles_per_chunk, a = tile - chunk * tiles_per_chunk;
            const bool valid = (2 * a + (q >> 1)) < __ldg(&chunks[chunk].t1_valid);
            const int out_row = q * rows_per_plane + (chunk * P2 + 1 + a) * 32;
            for (int h = 0; h < n_halves; ++h) {
                const int nch = (h ? N1 : N0) >> 5;
                tc::mbar_wait(tfull(h), tph);
                tc::fence_after_sync();
                const uint32_t t_row = tmem + h * 256 + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
                for (int c = hf; c < nch; c += kC1EpiWarps / 4) {
                    const int n0 = h * 256 + c * 32;
                    uint32_t pk[16];
                    if (valid) {
                        uint32_t r[32];
                        tc::tmem_ld_32x32b_x32(t_row + c * 32, r);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            const float4 b = *reinterpret_cast<const float4*>(gBias + n0 + j);
                            const float v0 = gelu_tanh6(__uint_as_float(r[j + 0]) + b.x);
                            const float v1 = gelu_tanh6(__uint_as_float(r[j + 1]) + b.y);
                            const float v2 = gelu_tanh6(__uint_as_float(r[j + 2]) + b.z);
                            const float v3 = gelu_tanh6(__uint_as_float(r[j + 3]) + b.w);
                            pk[(j >> 1) + 0] = OpTraits<OpT>::pack2(v0, v1);
                            pk[(j >> 1) + 1] = OpTraits<OpT>::pack2(v2, v3);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) pk[j] = 0u;
                    }
                    uint8_t* buf = my_buf_gen + (seq & 1) * kC1EpiBuf;
                    if (lane == 0) tc::bulk_wait_group_read<1>();
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        *reinterpret_cast<uint4*>(buf + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) =
                            make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
                    tc::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        tc::tma_store_2d(&tmap_out, my_buf + (seq & 1) * kC1EpiBuf, n0, out_row);
                        tc::bulk_commit_group();
                    }
                    ++seq;
                }
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(tempty(h));
            }
            tph ^= 1;
        }
        if (lane == 0) tc::bulk_wait_group<0>();
    }
    __syncwarp();
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 4) tc::tmem_dealloc<512>(tmem);
}

// ---- host side -----------------------------------------------------------------------------------------------
bool conv1_tc_supported(int C) { return C >= 32 && C <= 480 + 32 && C % 32 == 0 && (C <= 256 || (C - 256) % 16 == 0); }

// smem image of the weights: row n = [w_hi(9) | w_hi(9) | w_lo(9) | 0(5)] f16, 128-byte rows, 128-byte swizzle
void conv1_tc_weight_image(const float* w /*[C][9]*/, int C, std::vector<uint8_t>& image) {
    image.assign((size_t)C * 128, 0);
    for (int n = 0; n < C; ++n) {
        __half hk[32];
        for (int k = 0; k < 32; ++k) hk[k] = __float2half_rn(0.0f);
        for (int k = 0; k < 9; ++k) {
            const float v = w[n * 9 + k];
            const __half hi = __float2half_rn(v);
            const __half lo = __float2half_rn(v - __half2float(hi));
            hk[k] = hi; hk[9 + k] = hi; hk[18 + k] = lo;
        }
        for (int j = 0; j < 4; ++j)
            std::memcpy(image.data() + (size_t)n * 128 + ((j ^ (n & 7)) << 4), &hk[8 * j], 16);
    }
}

template <class OpT>
int conv1_tc_launch(const float* mel, const StemChunk* chunks_dev, int n_chunks, int P2, const float* clip_max,
                    const void* w_image, const float* bias, int C, OpT* planes, int R_cap, int row0, cudaStream_t stream) {
    if (n_chunks == 0) return Q3A_OK;
    if (!conv1_tc_supported(C)) { set_error("conv1_tc: unsupported channel count"); return Q3A_ERR_ARG; }
    // `chunks_dev` may be a sub-range of the pass's chunks: its first chunk owns plane rows from row0 on, so the
    // map starts there (plane stride unchanged) and stops at the end of the arena
    const int rows_per_plane = R_cap * 32;
    CUtensorMap tm;
    Q3A_TRY(make_tmap_tile32(&tm, planes + (size_t)row0 * 32 * C, 2, (uint64_t)4 * rows_per_plane - (uint64_t)row0 * 32,
                             (uint64_t)C, (uint64_t)C));
    auto kern = conv1_tc_kernel<OpT>;
    const int smem = c1_smem_bytes(C);
    Q3A_TRY(kernel_smem_optin(reinterpret_cast<const void*>(kern), c1_smem_bytes(512)));   // per (kernel, device), thread-safe
    const int tpc = P2 - 1;
    const int n_tiles = n_chunks * tpc;
    const int sms = sm_count_cached();
    const int grid = n_tiles < sms ? n_tiles : sms;
    Q3A_CUDA(launch_pdl(kern, dim3(grid), dim3(kC1Threads), smem, stream, mel, chunks_dev, n_tiles, tpc, P2, clip_max,
                        reinterpret_cast<const uint4*>(w_image), bias, C, tm, rows_per_plane));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

template int conv1_tc_launch<__nv_bfloat16>(const float*, const StemChunk*, int, int, const float*, const void*,
                                            const float*, int, __nv_bfloat16*, int, int, cudaStream_t);
template int conv1_tc_launch<__half>(const float*, const StemChunk*, int, int, const float*, const void*, const float*,
                                     int, __half*, int, int, cudaStream_t);

}  // namespace q3a
