// Variable-length multi-head self-attention on tcgen05 for sm_100a (head_dim = 64).
//
// Replaces the reference's materialised attention: K.Q mul_mat, ggml_soft_max_ext(scale 1/8, mask),
// V.P mul_mat and the permutes around them (audio_encoder.cpp:466-486; block-diagonal variant
// forced_aligner.cpp:737-766,809-821).  Tokens attend only inside their segment: one segment per
// utterance for the ASR encoder (no mask), 104-token windows for the aligner.  Scores are never
// written to HBM.
//
// One CTA = one (128-query tile, head).  Per 128-key block:
//   S[128 x nk] = Q K^T        tcgen05.mma, operands K-major in smem (TMA, 128B swizzle), S in TMEM
//   online softmax             one query row per thread (tcgen05.ld), exp2 with folded scale, fp32
//   P -> smem (16-bit, swizzled K-major), O_blk[128 x 64] = P V   tcgen05.mma, V is the MN-major B operand
//   O = alpha * O + O_blk      fp32 registers
// K/V blocks are double-buffered so the next TMA overlaps the current block's math.
#include "attention.h"
#include "tc.cuh"
#include "tmap.h"

namespace q3a {

constexpr int kHd = 64;
constexpr int kQT = 128;                 // queries per CTA
constexpr int kKT = 128;                 // keys per block
constexpr int kTileBytes = 128 * 64 * 2; // 16 KB: [128 rows][64 x 16-bit]
constexpr int kAttnSmem = 1024 /*align*/ + kTileBytes /*Q*/ + 4 * kTileBytes /*K,V x2*/ + 2 * kTileBytes /*P*/ + 128;
constexpr int kAttnTmemCols = 256;       // S: columns [0,128), O_blk: columns [128,192)

template <class OpT>
__global__ void __launch_bounds__(128, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const AttnTile* __restrict__ tiles,
                 OpT* __restrict__ out, int d_model) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - tc::smem_u32(smem_raw));
    const uint32_t sQ = base;
    auto sK = [&](int b) { return base + kTileBytes * (1 + 2 * b); };
    auto sV = [&](int b) { return base + kTileBytes * (2 + 2 * b); };
    const uint32_t sP = base + 5 * kTileBytes;
    uint8_t* gP = gen + 5 * kTileBytes;
    const uint32_t bars = base + 7 * kTileBytes;
    const uint32_t bar_q = bars, bar_s = bars + 8, bar_o = bars + 16;
    auto bar_kv = [&](int b) { return bars + 24 + 8 * b; };
    const uint32_t tmem_slot = bars + 48;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + 7 * kTileBytes + 48);

    const int tid = threadIdx.x, warp = tid >> 5;
    const AttnTile t = tiles[blockIdx.x];
    const int head = blockIdx.y;
    const int seg_len = t.seg_end - t.seg_start;
    const int nkv = (seg_len + kKT - 1) / kKT;
    const int col_q = head * kHd, col_k = d_model + head * kHd, col_v = 2 * d_model + head * kHd;

    if (tid == 0) {
        tc::prefetch_tmap(&tmap_qkv);
        tc::mbar_init(bar_q, 1); tc::mbar_init(bar_s, 1); tc::mbar_init(bar_o, 1);
        tc::mbar_init(bar_kv(0), 1); tc::mbar_init(bar_kv(1), 1);
        tc::fence_barrier_init();
    }
    if (warp == 0) tc::tmem_alloc<kAttnTmemCols>(tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot_ptr;
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);   // this warp's TMEM lane quadrant

    if (tid == 0) {
        tc::mbar_arrive_expect_tx(bar_q, kTileBytes);
        tc::tma_load_2d(sQ, &tmap_qkv, bar_q, col_q, t.q0);
        tc::mbar_arrive_expect_tx(bar_kv(0), 2 * kTileBytes);
        tc::tma_load_2d(sK(0), &tmap_qkv, bar_kv(0), col_k, t.seg_start);
        tc::tma_load_2d(sV(0), &tmap_qkv, bar_kv(0), col_v, t.seg_start);
    }

    constexpr int kFmt = OpTraits<OpT>::kFmt;
    const float kScale = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)
    float m_run = -INFINITY, l_run = 0.0f;
    float o[kHd];
#pragma unroll
    for (int i = 0; i < kHd; ++i) o[i] = 0.0f;

    for (int j = 0; j < nkv; ++j) {
        const int buf = j & 1;
        const int rem = seg_len - j * kKT;
        const int nk = rem >= kKT ? kKT : ((rem + 15) & ~15);      // keys this block feeds to the MMAs
        if (tid == 0) {
            if (j + 1 < nkv) {                                      // prefetch the next block
                const int nb = buf ^ 1;
                tc::mbar_arrive_expect_tx(bar_kv(nb), 2 * kTileBytes);
                tc::tma_load_2d(sK(nb), &tmap_qkv, bar_kv(nb), col_k, t.seg_start + (j + 1) * kKT);
                tc::tma_load_2d(sV(nb), &tmap_qkv, bar_kv(nb), col_v, t.seg_start + (j + 1) * kKT);
            }
            if (j == 0) tc::mbar_wait(bar_q, 0);
            tc::mbar_wait(bar_kv(buf), (j >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t idesc = tc::make_idesc(kFmt, kQT, nk, 0, 0);
            const uint64_t a = tc::smem_desc_kmajor_sw128(sQ), b = tc::smem_desc_kmajor_sw128(sK(buf));
#pragma unroll
            for (int k = 0; k < kHd / 16; ++k) tc::mma_f16_ss(tmem, a + 2 * k, b + 2 * k, idesc, k != 0);
            tc::mma_commit(bar_s);
        }
        tc::mbar_wait(bar_s, j & 1);
        tc::fence_after_sync();

        // ---- pass 1: row maximum over the block's valid keys ----------------------------------------
        const int n_valid = rem < kKT ? rem : kKT;
        float mx = -INFINITY;
        for (int c = 0; c < nk; c += 32) {
            uint32_t r[32];
            tc::tmem_ld_32x32b_x32(t_row + c, r);
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (c + i < n_valid) mx = fmaxf(mx, __uint_as_float(r[i]));
        }
        const float m_new = fmaxf(m_run, mx * kScale);
        const float alpha = exp2f(m_run - m_new);
        m_run = m_new;
        // ---- pass 2: p = exp2(s*scale - m), row sum, P -> swizzled smem -----------------------------
        float psum = 0.0f;
        for (int c = 0; c < nk; c += 32) {
            uint32_t r[32];
            tc::tmem_ld_32x32b_x32(t_row + c, r);
            tc::tmem_ld_wait();
            float p[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                p[i] = (c + i < n_valid) ? exp2f(fmaf(__uint_as_float(r[i]), kScale, -m_new)) : 0.0f;
                psum += p[i];
            }
            // 64-key sub-tile (c / 64), 16-byte chunks ((c % 64) / 8 + g), XOR-swizzled with row % 8
            uint8_t* prow = gP + (c >> 6) * kTileBytes + tid * 128;
            const int ch0 = (c & 63) >> 3;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                uint4 u;
                u.x = OpTraits<OpT>::pack2(p[g * 8 + 0], p[g * 8 + 1]);
                u.y = OpTraits<OpT>::pack2(p[g * 8 + 2], p[g * 8 + 3]);
                u.z = OpTraits<OpT>::pack2(p[g * 8 + 4], p[g * 8 + 5]);
                u.w = OpTraits<OpT>::pack2(p[g * 8 + 6], p[g * 8 + 7]);
                *reinterpret_cast<uint4*>(prow + (((ch0 + g) ^ (tid & 7)) << 4)) = u;
            }
        }
        l_run = l_run * alpha + psum;
        tc::fence_proxy_async();          // generic-proxy smem writes -> visible to the tensor core
        tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc::fence_after_sync();
            const uint32_t idesc = tc::make_idesc(kFmt, kQT, kHd, 0, 1);
            const uint64_t b0 = tc::smem_desc_mnmajor_sw128(sV(buf));
            const int ks = nk / 16;
            for (int k = 0; k < ks; ++k) {
                const uint64_t a = tc::smem_desc_kmajor_sw128(sP + (k >> 2) * kTileBytes) + 2 * (k & 3);
                tc::mma_f16_ss(tmem + 128, a, b0 + (uint64_t)k * 128, idesc, k != 0);   // 16 key rows = 2048 B
            }
            tc::mma_commit(bar_o);
        }
        tc::mbar_wait(bar_o, j & 1);
        tc::fence_after_sync();
#pragma unroll
        for (int c = 0; c < kHd; c += 32) {
            uint32_t r[32];
            tc::tmem_ld_32x32b_x32(t_row + 128 + c, r);
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[c + i] = fmaf(o[c + i], alpha, __uint_as_float(r[i]));
        }
        tc::fence_before_sync();
    }

    const int row = t.q0 + tid;
    if (row < t.seg_end) {
        const float inv = 1.0f / l_run;
        OpT* dst = out + (size_t)row * d_model + head * kHd;
#pragma unroll
        for (int g = 0; g < kHd / 8; ++g) {
            uint4 u;
            u.x = OpTraits<OpT>::pack2(o[g * 8 + 0] * inv, o[g * 8 + 1] * inv);
            u.y = OpTraits<OpT>::pack2(o[g * 8 + 2] * inv, o[g * 8 + 3] * inv);
            u.z = OpTraits<OpT>::pack2(o[g * 8 + 4] * inv, o[g * 8 + 5] * inv);
            u.w = OpTraits<OpT>::pack2(o[g * 8 + 6] * inv, o[g * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + g * 8) = u;
        }
    }
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<kAttnTmemCols>(tmem);
}

void attention_tiles(const int* cu_seqlens, int n_seg, std::vector<AttnTile>& tiles) {
    tiles.clear();
    for (int s = 0; s < n_seg; ++s)
        for (int q0 = cu_seqlens[s]; q0 < cu_seqlens[s + 1]; q0 += kQT)
            tiles.push_back(AttnTile{q0, cu_seqlens[s], cu_seqlens[s + 1], 0});
}

double attention_flops(const int* cu_seqlens, int n_seg, int n_heads) {
    double f = 0.0;
    for (int s = 0; s < n_seg; ++s) {
        const double len = cu_seqlens[s + 1] - cu_seqlens[s];
        f += 4.0 * kHd * n_heads * len * len;
    }
    return f;
}

int attention_launch(const void* qkv, void* out, int T, int n_heads, const AttnTile* tiles_dev, int n_tiles, int fmt,
                     cudaStream_t stream) {
    if (n_tiles == 0 || T == 0) return Q3A_OK;
    const int d = n_heads * kHd;
    CUtensorMap tm;
    Q3A_TRY(make_tmap_2d(&tm, qkv, (uint64_t)T, (uint64_t)3 * d, (uint64_t)3 * d, 128));
    static bool attr_set = false;
    if (!attr_set) {
        Q3A_CUDA(cudaFuncSetAttribute(attention_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem));
        Q3A_CUDA(cudaFuncSetAttribute(attention_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem));
        attr_set = true;
    }
    dim3 grid(n_tiles, n_heads);
    if (fmt == OP_F16)
        attention_kernel<__half><<<grid, 128, kAttnSmem, stream>>>(tm, tiles_dev, (__half*)out, d);
    else
        attention_kernel<__nv_bfloat16><<<grid, 128, kAttnSmem, stream>>>(tm, tiles_dev, (__nv_bfloat16*)out, d);
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

}  // namespace q3a

extern "C" int q3a_test_attention(const void* qkv_dev, void* out_dev, int T, int H, const int* cu_seqlens, int n_seg,
                                  int operand_dtype, void* stream) {
    using namespace q3a;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return Q3A_ERR_CUDA; }
    Q3A_TRY(use_device(dev));
    if (!qkv_dev || !out_dev || !cu_seqlens || n_seg < 0 || H <= 0) { set_error("test_attention: bad argument"); return Q3A_ERR_ARG; }
    std::vector<AttnTile> tiles;
    attention_tiles(cu_seqlens, n_seg, tiles);
    if (tiles.empty()) return Q3A_OK;
    AttnTile* d_tiles = nullptr;
    Q3A_CUDA(cudaMalloc(&d_tiles, sizeof(AttnTile) * tiles.size()));
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemcpyAsync(d_tiles, tiles.data(), sizeof(AttnTile) * tiles.size(), cudaMemcpyHostToDevice, st);
    int rc = Q3A_OK;
    if (e != cudaSuccess) { set_error(cudaGetErrorString(e)); rc = Q3A_ERR_CUDA; }
    if (rc == Q3A_OK) rc = attention_launch(qkv_dev, out_dev, T, H, d_tiles, (int)tiles.size(), operand_dtype, st);
    cudaStreamSynchronize(st);
    cudaFree(d_tiles);
    return rc;
}
