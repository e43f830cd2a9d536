// Causal grouped-query self-attention on tcgen05 for the forced aligner's NAR decoder prefill (head_dim = 128).
//
// Replaces, per decoder layer, ggml_flash_attn_ext(Q, K, V, causal_mask, 1/sqrt(128)) with GGML_PREC_F32
// (reference forced_aligner.cpp:1036-1042; the mask is k <= q, :1128-1139) for one sequence of n tokens with
// 16 query heads sharing 8 key/value heads (text_attention_heads / text_kv_heads, forced_aligner.h:50-58).
//
// Same scheme as attention.cu (one CTA = one 128-query tile of one head, 64-key blocks, S and O in TMEM, P written
// back over the consumed scores and fed to P.V as the A operand straight from TMEM, lazily rescaled running maximum),
// widened to 128 dimensions as two 64-column swizzle atoms per operand:
//   S   = Q[:, 0:64] K[:, 0:64]^T + Q[:, 64:128] K[:, 64:128]^T       8 MMAs of K = 16 into columns [0,64)
//   O_h += P V[:, 64h:64h+64]   for h = 0, 1                            N = 64 each, columns [64,128) and [128,192)
// Causality: a query tile only walks the key blocks that start at or before its last row, and inside the blocks that
// straddle the diagonal each row masks the keys behind it.  Block 0 always holds key 0 <= every row, so the reference
// maximum of a row is finite from the first block on.
#include "attention.h"
#include "tc.cuh"
#include "tmap.h"

namespace q3a {

namespace {

constexpr int kHdD = 128;                  // head dimension
constexpr int kQTD = 128;                  // queries per CTA
constexpr int kKTD = 64;                   // keys per block
constexpr int kQBytes = 2 * 128 * 64 * 2;  // Q: two [128][64] atoms
constexpr int kKVBytes = 2 * 64 * 64 * 2;  // one K or V block: two [64][64] atoms
constexpr int kDecSmem = kQBytes + 4 * kKVBytes + 128;
constexpr int kDecTmemCols = 256;          // S [0,64), O [64,192)
constexpr float kDecRescale = 8.0f;

}  // namespace

template <class OpT>
__global__ void __launch_bounds__(128, 2)
attention_dec_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                     OpT* __restrict__ out, int n_tokens, int n_heads, int kv_group, int ld_out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = tc::smem_u32(smem_raw);
    if (base & 1023u) { asm volatile("trap;"); }
    const uint32_t sQ = base;                                                   // atoms at +0, +16 KB
    auto sK = [&](int b) { return base + kQBytes + kKVBytes * b; };             // atoms at +0, +8 KB
    auto sV = [&](int b) { return base + kQBytes + kKVBytes * (2 + b); };
    constexpr int kBarOff = kQBytes + 4 * kKVBytes;
    const uint32_t bars = base + kBarOff;
    const uint32_t bar_q = bars, bar_s = bars + 8, bar_o = bars + 16;
    auto bar_k = [&](int b) { return bars + 24 + 8 * b; };
    auto bar_v = [&](int b) { return bars + 40 + 8 * b; };
    const uint32_t tmem_slot = bars + 56;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + kBarOff + 56);

    const int tid = threadIdx.x, warp = tid >> 5;
    const int q0 = blockIdx.x * kQTD, head = blockIdx.y;
    const int q_last = min(q0 + kQTD, n_tokens) - 1;               // last live query row of the tile
    const int nkv = q_last / kKTD + 1;                             // key blocks that start at or before it
    const int d_q = n_heads * kHdD, d_kv = (n_heads / kv_group) * kHdD;
    const int col_q = head * kHdD, col_k = d_q + (head / kv_group) * kHdD, col_v = d_q + d_kv + (head / kv_group) * kHdD;
    constexpr int kFmt = OpTraits<OpT>::kFmt;

    uint32_t tmem = 0;
    auto keys_of = [&](int j) {                                    // keys block j feeds to the MMAs (multiple of 16)
        const int rem = min(n_tokens, q_last + 1) - j * kKTD;
        return rem >= kKTD ? kKTD : ((rem + 15) & ~15);
    };
    auto load_kv = [&](uint32_t dst, uint32_t bar, int col, int row) {          // elected lane: both atoms of a block
        tc::mbar_arrive_expect_tx(bar, kKVBytes);
        tc::tma_load_2d(dst, &tmap_kv, bar, col, row);
        tc::tma_load_2d(dst + kKVBytes / 2, &tmap_kv, bar, col + 64, row);
    };
    auto issue_qk = [&](int j) {                                   // warp 0, converged
        tc::mbar_wait(bar_k(j & 1), (j >> 1) & 1);
        tc::fence_after_sync();
        const uint32_t idesc = tc::make_idesc(kFmt, kQTD, keys_of(j), 0, 0);
        const uint64_t a = tc::smem_desc_kmajor_sw128(sQ), b = tc::smem_desc_kmajor_sw128(sK(j & 1));
        const uint64_t a1 = tc::smem_desc_kmajor_sw128(sQ + kQBytes / 2), b1 = tc::smem_desc_kmajor_sw128(sK(j & 1) + kKVBytes / 2);
        if (tc::elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) tc::mma_f16_ss(tmem, a + 2 * k, b + 2 * k, idesc, k != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) tc::mma_f16_ss(tmem, a1 + 2 * k, b1 + 2 * k, idesc, 1);
            tc::mma_commit(bar_s);
        }
        __syncwarp();
    };

    if (warp == 0) {
        if (tc::elect_one()) {
            pdl_trigger();
            tc::prefetch_tmap(&tmap_q);
            tc::prefetch_tmap(&tmap_kv);
            tc::mbar_init(bar_q, 1); tc::mbar_init(bar_s, 1); tc::mbar_init(bar_o, 1);
            tc::mbar_init(bar_k(0), 1); tc::mbar_init(bar_k(1), 1);
            tc::mbar_init(bar_v(0), 1); tc::mbar_init(bar_v(1), 1);
            tc::fence_barrier_init();
            pdl_wait();
            tc::mbar_arrive_expect_tx(bar_q, kQBytes);
            tc::tma_load_2d(sQ, &tmap_q, bar_q, col_q, q0);
            tc::tma_load_2d(sQ + kQBytes / 2, &tmap_q, bar_q, col_q + 64, q0);
            load_kv(sK(0), bar_k(0), col_k, 0);
            load_kv(sV(0), bar_v(0), col_v, 0);
            if (nkv > 1) load_kv(sK(1), bar_k(1), col_k, kKTD);
        }
        __syncwarp();
        tc::tmem_alloc<kDecTmemCols>(tmem_slot);
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    tmem = *tmem_slot_ptr;
    pdl_wait();
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);
    if (warp == 0) {
        tc::mbar_wait(bar_q, 0);
        issue_qk(0);
    }

    const float kScale = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)
    const bool row_live = (q0 + warp * 32) < n_tokens;                   // warp-uniform
    const int row_pos = q0 + tid;                                        // this thread's query position (= its last key)
    float m_ref = 0.0f, l_run = 0.0f;

    for (int j = 0; j < nkv; ++j) {
        const int nk = keys_of(j);
        // keys of this block the row may see: key index k0 + i <= row_pos and < n_tokens
        const int k0 = j * kKTD;
        const int nv_row = max(0, min(min(kKTD, n_tokens - k0), row_pos - k0 + 1));
        const bool full = (q0 + warp * 32 - k0 + 1) >= kKTD && (n_tokens - k0) >= kKTD;   // warp-uniform: no masking needed
        tc::mbar_wait(bar_s, j & 1);
        tc::fence_after_sync();
        uint32_t s[kKTD];
        if (row_live) {
#pragma unroll
            for (int c = 0; c < kKTD / 32; ++c)
                if (c * 32 < nk) tc::tmem_ld_32x32b_x32(t_row + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
            tc::tmem_ld_wait();
        }
        if (warp == 0 && j + 2 < nkv) {
            if (tc::elect_one()) load_kv(sK(j & 1), bar_k(j & 1), col_k, (j + 2) * kKTD);
            __syncwarp();
        }
        float psum = 0.0f;
        uint32_t* pk = s;
        if (row_live) {
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
            if (full) {
#pragma unroll
                for (int i = 0; i < kKTD; i += 8) {
                    mx0 = fmaxf(mx0, fmaxf(__uint_as_float(s[i + 0]), __uint_as_float(s[i + 1])));
                    mx1 = fmaxf(mx1, fmaxf(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])));
                    mx2 = fmaxf(mx2, fmaxf(__uint_as_float(s[i + 4]), __uint_as_float(s[i + 5])));
                    mx3 = fmaxf(mx3, fmaxf(__uint_as_float(s[i + 6]), __uint_as_float(s[i + 7])));
                }
            } else {
#pragma unroll
                for (int i = 0; i < kKTD; i += 16) {
                    if (i < nk) {
#pragma unroll
                        for (int e = 0; e < 16; e += 4) {
                            if (i + e + 0 < nv_row) mx0 = fmaxf(mx0, __uint_as_float(s[i + e + 0]));
                            if (i + e + 1 < nv_row) mx1 = fmaxf(mx1, __uint_as_float(s[i + e + 1]));
                            if (i + e + 2 < nv_row) mx2 = fmaxf(mx2, __uint_as_float(s[i + e + 2]));
                            if (i + e + 3 < nv_row) mx3 = fmaxf(mx3, __uint_as_float(s[i + e + 3]));
                        }
                    }
                }
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * kScale;   // -inf when the row sees no key of this block
            if (j == 0) {
                m_ref = mx;
            } else {
                const bool grow = mx > m_ref + kDecRescale;
                if (__any_sync(0xffffffffu, grow)) {
                    const float alpha = grow ? tc::ex2_approx(m_ref - mx) : 1.0f;
                    if (grow) m_ref = mx;
                    l_run *= alpha;
                    tc::mbar_wait(bar_o, (j - 1) & 1);
                    tc::fence_after_sync();
#pragma unroll
                    for (int c = 0; c < kHdD; c += 32) {
                        uint32_t r[32];
                        tc::tmem_ld_32x32b_x32(t_row + kKTD + c, r);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tc::tmem_st_32x32b_x32(t_row + kKTD + c, r);
                    }
                    tc::tmem_st_wait();
                }
            }
            const float neg_m = -m_ref;
            if (full) {
                float ps1 = 0.0f;
#pragma unroll
                for (int i = 0; i < kKTD; i += 2) {
                    const float p0 = tc::ex2_approx(fmaf(__uint_as_float(s[i]), kScale, neg_m));
                    const float p1 = tc::ex2_approx(fmaf(__uint_as_float(s[i + 1]), kScale, neg_m));
                    psum += p0; ps1 += p1;
                    pk[i >> 1] = OpTraits<OpT>::pack2(p0, p1);
                }
                psum += ps1;
            } else {
#pragma unroll
                for (int i = 0; i < kKTD; i += 16) {
                    if (i < nk) {
#pragma unroll
                        for (int e = 0; e < 16; e += 2) {
                            float p0 = 0.0f, p1 = 0.0f;
                            if (i + e < nv_row) p0 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e]), kScale, neg_m));
                            if (i + e + 1 < nv_row) p1 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e + 1]), kScale, neg_m));
                            psum += p0 + p1;
                            pk[(i + e) >> 1] = OpTraits<OpT>::pack2(p0, p1);
                        }
                    }
                }
            }
            l_run += psum;
        }
        if (j > 0) { tc::mbar_wait(bar_o, (j - 1) & 1); tc::fence_after_sync(); }
        if (row_live) {
            tc::tmem_st_32x32b_x32(t_row, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
            tc::tmem_st_wait();
        }
        tc::fence_before_sync();
        __syncthreads();
        if (warp == 0) {
            tc::fence_after_sync();
            if (j + 1 < nkv) {
                const int nb = (j + 1) & 1;
                if (tc::elect_one()) load_kv(sV(nb), bar_v(nb), col_v, (j + 1) * kKTD);
                __syncwarp();
            }
            tc::mbar_wait(bar_v(j & 1), (j >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t idesc = tc::make_idesc(kFmt, kQTD, 64, 0, 1);
            const uint64_t b0 = tc::smem_desc_mnmajor_sw128(sV(j & 1)), b1 = tc::smem_desc_mnmajor_sw128(sV(j & 1) + kKVBytes / 2);
            const int ks = nk / 16;
            if (tc::elect_one()) {
#pragma unroll
                for (int k = 0; k < kKTD / 16; ++k)
                    if (k < ks) {
                        tc::mma_f16_ts(tmem + kKTD, tmem + 8 * k, b0 + (uint64_t)k * 128, idesc, (j | k) != 0);
                        tc::mma_f16_ts(tmem + kKTD + 64, tmem + 8 * k, b1 + (uint64_t)k * 128, idesc, (j | k) != 0);
                    }
                tc::mma_commit(bar_o);
            }
            __syncwarp();
            if (j + 1 < nkv) issue_qk(j + 1);
        }
    }

    // ---- epilogue: O / l -> 16-bit rows of 128 dimensions (256 bytes), staged through the idle Q atoms -----------
    tc::mbar_wait(bar_o, (nkv - 1) & 1);
    tc::fence_after_sync();
    if (row_live) {
        const float inv = 1.0f / l_run;
#pragma unroll
        for (int h = 0; h < 2; ++h) {                      // dims 64h .. 64h+63 -> Q atom h, row tid, 16-byte XOR swizzle
            uint32_t r[64];
            tc::tmem_ld_32x32b_x32(t_row + kKTD + 64 * h, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
            tc::tmem_ld_32x32b_x32(t_row + kKTD + 64 * h + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
            tc::tmem_ld_wait();
            uint8_t* stage = smem_raw + h * (kQBytes / 2) + tid * 128;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                uint4 u;
                u.x = OpTraits<OpT>::pack2(__uint_as_float(r[g * 8 + 0]) * inv, __uint_as_float(r[g * 8 + 1]) * inv);
                u.y = OpTraits<OpT>::pack2(__uint_as_float(r[g * 8 + 2]) * inv, __uint_as_float(r[g * 8 + 3]) * inv);
                u.z = OpTraits<OpT>::pack2(__uint_as_float(r[g * 8 + 4]) * inv, __uint_as_float(r[g * 8 + 5]) * inv);
                u.w = OpTraits<OpT>::pack2(__uint_as_float(r[g * 8 + 6]) * inv, __uint_as_float(r[g * 8 + 7]) * inv);
                *reinterpret_cast<uint4*>(stage + ((g ^ (tid & 7)) << 4)) = u;
            }
        }
        __syncwarp();
        const int lane = tid & 31, sub = lane >> 3, g = lane & 7;
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int rr = warp * 32 + i * 4 + sub;
                const uint4 u = *reinterpret_cast<const uint4*>(smem_raw + h * (kQBytes / 2) + rr * 128 + ((g ^ (rr & 7)) << 4));
                if (q0 + rr < n_tokens)
                    *reinterpret_cast<uint4*>(out + (size_t)(q0 + rr) * ld_out + head * kHdD + h * 64 + g * 8) = u;
            }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<kDecTmemCols>(tmem);
}

template <class OpT>
static int attention_dec_launch_t(const void* qkv, void* out, int n_tokens, int n_heads, int n_kv_heads, cudaStream_t stream) {
    const int ld = (n_heads + 2 * n_kv_heads) * kHdD;
    CUtensorMap tm_q, tm_kv;
    Q3A_TRY(make_tmap_2d(&tm_q, qkv, (uint64_t)n_tokens, (uint64_t)ld, (uint64_t)ld, 128));
    Q3A_TRY(make_tmap_2d(&tm_kv, qkv, (uint64_t)n_tokens, (uint64_t)ld, (uint64_t)ld, kKTD));
    auto kern = attention_dec_kernel<OpT>;
    Q3A_TRY(kernel_smem_optin(reinterpret_cast<const void*>(kern), kDecSmem));
    Q3A_CUDA(launch_pdl(kern, dim3((n_tokens + kQTD - 1) / kQTD, n_heads), dim3(128), kDecSmem, stream, tm_q, tm_kv, (OpT*)out,
                        n_tokens, n_heads, n_heads / n_kv_heads, n_heads * kHdD));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

// qkv: [n][(H + 2 Hkv) * 128] 16-bit (q | k | v column blocks, q and k already normalised and rotated);
// out: [n][H * 128] 16-bit
int attention_dec_launch(const void* qkv, void* out, int n_tokens, int n_heads, int n_kv_heads, int fmt, cudaStream_t stream) {
    if (n_tokens <= 0) return Q3A_OK;
    if (n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads) { set_error("attention_dec: heads must be a multiple of kv heads"); return Q3A_ERR_ARG; }
    if (fmt == OP_F16) return attention_dec_launch_t<__half>(qkv, out, n_tokens, n_heads, n_kv_heads, stream);
    return attention_dec_launch_t<__nv_bfloat16>(qkv, out, n_tokens, n_heads, n_kv_heads, stream);
}

}  // namespace q3a
