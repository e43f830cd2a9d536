// Variable-length multi-head self-attention on tcgen05 for sm_100a (head_dim = 64).
//
// Replaces the reference's materialised attention: K.Q mul_mat, ggml_soft_max_ext(scale 1/8, mask),
// V.P mul_mat and the permutes around them (audio_encoder.cpp:466-486; block-diagonal variant
// forced_aligner.cpp:737-766,809-821).  Tokens attend only inside their segment: one segment per
// utterance for the ASR encoder (no mask), 104-token windows for the aligner.  Scores are never
// written to HBM.
//
// One CTA (128 threads, one query row per thread) = one (128-query tile, head); keys go in blocks of KT = 64 (three CTAs
// per SM) for segments below 2048 tokens and KT = 128 (two CTAs per SM) from there on -- chosen per SEGMENT, so that a
// clip's result does not depend on what it is batched with (attention_key_block; Q3A_ATTN_KT=64|128 forces one).
// Per key block j:
//   S_j = Q K_j^T            tcgen05.mma into TMEM columns [0,KT)
//   S_j -> registers         tcgen05.ld.x32 back to back, one wait
//   softmax                  row maximum (four chains), p = exp2(s*scale - m) with a lazily updated reference
//                            maximum m (O in TMEM is rescaled only when the row maximum grows by more than 2^8,
//                            so the common block never reads O back), row sum in fp32, P packed in place
//   P_j -> TMEM              tcgen05.st over the row's own consumed scores, two 16-bit values per column
//   O += P_j V_j             tcgen05.mma with the A operand read from TMEM, accumulating in columns [KT,KT+64);
//                            V is the MN-major B operand.  Q K_{j+1}^T is issued right behind it: the pipe executes
//                            in order, so S_{j+1} overwrites P_j only after P_j V_j has read it, and it completes
//                            while the other CTAs of the SM run their softmax
// One bar.sync per block (P written -> issue).  K and V blocks are double-buffered with separate barriers: K_{j+2} is
// fetched as soon as Q K_j^T has retired, V_{j+1} as soon as P_{j-1} V_{j-1} has.  Warp 0 doubles as the issue warp and
// stays converged (elect.sync around the TMA / MMA instructions only).  Warps whose 32 query rows all lie past the
// end of the segment skip the softmax arithmetic.  The epilogue divides by the row sum, transposes through the idle
// Q tile and stores whole 128-byte rows.  DESIGN.md 4.2 has the measurements behind each of these choices.
#include "attention.h"
#include "tc.cuh"
#include "tmap.h"

#include <cstdlib>

namespace q3a {

constexpr int kHd = 64;
constexpr int kQT = 128;                 // queries per CTA
constexpr int kTileBytes = 128 * 64 * 2; // 16 KB: [128 rows][64 x 16-bit]
// KT = keys per block (128 or 64).  128: fewer blocks, 2 CTAs/SM (227 registers, 80 KB smem, 256 TMEM columns).
// 64: half the per-thread state, so 3 CTAs/SM fit (157 registers, 48 KB, 128 columns) and hide each other's
// barrier / TMEM-load latency better.  (A fourth CTA would need 128 registers: 384 B of spills.)
#ifndef Q3A_ATTN_CTAS
#define Q3A_ATTN_CTAS 3
#endif
template <int KT> struct AttnCfg {
    static constexpr int kKV = KT * 64 * 2;                    // one K or V block
    static constexpr int kSmem = kTileBytes + 4 * kKV + 128;   // Q, 2 x K, 2 x V, barriers (P lives in TMEM)
    static constexpr int kTmemCols = KT == 128 ? 256 : 128;    // S: columns [0,KT), O: columns [KT,KT+64)
    static constexpr int kCtas = KT == 128 ? 2 : Q3A_ATTN_CTAS;
};
constexpr float kRescaleThreshold = 8.0f;   // log2 units: p <= 2^8 between rescales

template <class OpT, int kKT>
__global__ void __launch_bounds__(128, AttnCfg<kKT>::kCtas)
attention_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_kv,
                 const AttnTile* __restrict__ tiles, OpT* __restrict__ out, int d_model) {
    using Cfg = AttnCfg<kKT>;
    constexpr int kKV = Cfg::kKV;
    constexpr int kAttnTmemCols = Cfg::kTmemCols;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = tc::smem_u32(smem_raw);
    if (base & 1023u) { asm volatile("trap;"); }          // the swizzled tiles need 1024-byte alignment
    const uint32_t sQ = base;
    auto sK = [&](int b) { return base + kTileBytes + kKV * b; };
    auto sV = [&](int b) { return base + kTileBytes + kKV * (2 + b); };
    constexpr int kBarOff = kTileBytes + 4 * kKV;
    const uint32_t bars = base + kBarOff;
    const uint32_t bar_q = bars, bar_s = bars + 8, bar_o = bars + 16;
    auto bar_k = [&](int b) { return bars + 24 + 8 * b; };
    auto bar_v = [&](int b) { return bars + 40 + 8 * b; };
    const uint32_t tmem_slot = bars + 56;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + kBarOff + 56);

    const int tid = threadIdx.x, warp = tid >> 5;
    const AttnTile t = tiles[blockIdx.x];                  // (the table is a host upload, not the previous kernel's output)
    if (t.pad != kKT) return;                              // a mixed launch: this tile belongs to the other block size
    const int head = blockIdx.y;
    const int seg_len = t.seg_end - t.seg_start;
    const int nkv = (seg_len + kKT - 1) / kKT;
    const int col_q = head * kHd, col_k = d_model + head * kHd, col_v = 2 * d_model + head * kHd;
    constexpr int kFmt = OpTraits<OpT>::kFmt;

    uint32_t tmem = 0;
    auto keys_of = [&](int j) {                                      // keys block j feeds to the MMAs
        const int rem = seg_len - j * kKT;
        return rem >= kKT ? kKT : ((rem + 15) & ~15);
    };
    // Warp 0 is also the issue warp.  It stays converged: barrier waits and descriptor arithmetic are
    // warp-uniform and only the TMA / tcgen05 instructions themselves sit under elect.sync, so the
    // UTCHMMAs go out back to back (a lone `tid == 0` branch takes ~1000 clk per block to issue).
    auto issue_qk = [&](int j) {                                     // warp 0, converged
        tc::mbar_wait(bar_k(j & 1), (j >> 1) & 1);
        tc::fence_after_sync();
        const uint32_t idesc = tc::make_idesc(kFmt, kQT, keys_of(j), 0, 0);
        const uint64_t a = tc::smem_desc_kmajor_sw128(sQ), b = tc::smem_desc_kmajor_sw128(sK(j & 1));
        if (tc::elect_one()) {
#pragma unroll
            for (int k = 0; k < kHd / 16; ++k) tc::mma_f16_ss(tmem, a + 2 * k, b + 2 * k, idesc, k != 0);
            tc::mma_commit(bar_s);
        }
        __syncwarp();
    };

    if (warp == 0) {
        // barriers, first loads, then the TMEM allocation: the fetch latency runs under it
        if (tc::elect_one()) {
            pdl_trigger();
            tc::prefetch_tmap(&tmap_qkv);
            tc::prefetch_tmap(&tmap_kv);
            tc::mbar_init(bar_q, 1); tc::mbar_init(bar_s, 1); tc::mbar_init(bar_o, 1);
            tc::mbar_init(bar_k(0), 1); tc::mbar_init(bar_k(1), 1);
            tc::mbar_init(bar_v(0), 1); tc::mbar_init(bar_v(1), 1);
            tc::fence_barrier_init();
            pdl_wait();                                    // qkv is the previous kernel's output
            tc::mbar_arrive_expect_tx(bar_q, kTileBytes);
            tc::tma_load_2d(sQ, &tmap_qkv, bar_q, col_q, t.q0);
            tc::mbar_arrive_expect_tx(bar_k(0), kKV);
            tc::tma_load_2d(sK(0), &tmap_kv, bar_k(0), col_k, t.seg_start);
            tc::mbar_arrive_expect_tx(bar_v(0), kKV);
            tc::tma_load_2d(sV(0), &tmap_kv, bar_v(0), col_v, t.seg_start);
            if (nkv > 1) {
                tc::mbar_arrive_expect_tx(bar_k(1), kKV);
                tc::tma_load_2d(sK(1), &tmap_kv, bar_k(1), col_k, t.seg_start + kKT);
            }
        }
        __syncwarp();
        tc::tmem_alloc<kAttnTmemCols>(tmem_slot);
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    tmem = *tmem_slot_ptr;
    pdl_wait();
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);   // this warp's TMEM lane quadrant
    if (warp == 0) {
        tc::mbar_wait(bar_q, 0);
        issue_qk(0);
    }

    const float kScale = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)
    const bool row_live = (t.q0 + warp * 32) < t.seg_end;  // warp-uniform: any valid query row in this warp
    float m_ref = 0.0f, l_run = 0.0f;

    for (int j = 0; j < nkv; ++j) {
        const int rem = seg_len - j * kKT;
        const int nk = keys_of(j);
        const int n_valid = rem < kKT ? rem : kKT;
        // ---- S_j -> registers ---------------------------------------------------------------------
        tc::mbar_wait(bar_s, j & 1);
        tc::fence_after_sync();
        uint32_t s[kKT];
        if (row_live) {
#pragma unroll
            for (int c = 0; c < kKT / 32; ++c)
                if (c * 32 < nk) tc::tmem_ld_32x32b_x32(t_row + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
            tc::tmem_ld_wait();
        }
        if (warp == 0 && j + 2 < nkv) {                    // Q K_j^T has retired: its K buffer is free
            if (tc::elect_one()) {
                tc::mbar_arrive_expect_tx(bar_k(j & 1), kKV);
                tc::tma_load_2d(sK(j & 1), &tmap_kv, bar_k(j & 1), col_k, t.seg_start + (j + 2) * kKT);
            }
            __syncwarp();
        }
        // ---- softmax on registers -------------------------------------------------------------------
        float psum = 0.0f;
        uint32_t* pk = s;                                  // P row, packed 16-bit pairs, overwrites consumed scores in place
        if (row_live) {
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;   // independent chains
            if (n_valid == kKT) {
#pragma unroll
                for (int i = 0; i < kKT; i += 8) {
                    mx0 = fmaxf(mx0, fmaxf(__uint_as_float(s[i + 0]), __uint_as_float(s[i + 1])));
                    mx1 = fmaxf(mx1, fmaxf(__uint_as_float(s[i + 2]), __uint_as_float(s[i + 3])));
                    mx2 = fmaxf(mx2, fmaxf(__uint_as_float(s[i + 4]), __uint_as_float(s[i + 5])));
                    mx3 = fmaxf(mx3, fmaxf(__uint_as_float(s[i + 6]), __uint_as_float(s[i + 7])));
                }
            } else {
#pragma unroll
                for (int i = 0; i < kKT; i += 16) {
                    if (i < nk) {                          // warp-uniform: 16-key groups past the end are skipped
#pragma unroll
                        for (int e = 0; e < 16; e += 4) {
                            if (i + e + 0 < n_valid) mx0 = fmaxf(mx0, __uint_as_float(s[i + e + 0]));
                            if (i + e + 1 < n_valid) mx1 = fmaxf(mx1, __uint_as_float(s[i + e + 1]));
                            if (i + e + 2 < n_valid) mx2 = fmaxf(mx2, __uint_as_float(s[i + e + 2]));
                            if (i + e + 3 < n_valid) mx3 = fmaxf(mx3, __uint_as_float(s[i + e + 3]));
                        }
                    }
                }
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * kScale;
            if (j == 0) {
                m_ref = mx;
            } else {
                const bool grow = mx > m_ref + kRescaleThreshold;
                if (__any_sync(0xffffffffu, grow)) {
                    // rare: bring O (complete up to block j-1) and l to the new reference maximum
                    const float alpha = grow ? tc::ex2_approx(m_ref - mx) : 1.0f;
                    if (grow) m_ref = mx;
                    l_run *= alpha;
                    tc::mbar_wait(bar_o, (j - 1) & 1);
                    tc::fence_after_sync();
#pragma unroll
                    for (int c = 0; c < kHd; c += 32) {
                        uint32_t r[32];
                        tc::tmem_ld_32x32b_x32(t_row + kKT + c, r);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tc::tmem_st_32x32b_x32(t_row + kKT + c, r);
                    }
                    tc::tmem_st_wait();
                }
            }
            const float neg_m = -m_ref;
            if (n_valid == kKT) {
                // two keys per FFMA2 / FADD2 (same roundings as the scalar form: results are bit-identical)
                const uint64_t sc2 = tc::f32x2(kScale, kScale), nm2 = tc::f32x2(neg_m, neg_m);
                uint64_t ps2 = tc::f32x2(0.0f, 0.0f);
#pragma unroll
                for (int i = 0; i < kKT; i += 2) {
                    float x0, x1;
                    tc::f32x2_split(tc::fma_f32x2(tc::f32x2(__uint_as_float(s[i]), __uint_as_float(s[i + 1])), sc2, nm2), x0, x1);
                    const float p0 = tc::ex2_approx(x0), p1 = tc::ex2_approx(x1);
                    ps2 = tc::add_f32x2(ps2, tc::f32x2(p0, p1));
                    pk[i >> 1] = OpTraits<OpT>::pack2(p0, p1);
                }
                float ps0, ps1;
                tc::f32x2_split(ps2, ps0, ps1);
                psum += ps0; psum += ps1;
            } else {
#pragma unroll
                for (int i = 0; i < kKT; i += 16) {
                    if (i < nk) {                          // warp-uniform; P columns >= nk are never read
#pragma unroll
                        for (int e = 0; e < 16; e += 2) {
                            float p0 = 0.0f, p1 = 0.0f;
                            if (i + e < n_valid) p0 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e]), kScale, neg_m));
                            if (i + e + 1 < n_valid) p1 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e + 1]), kScale, neg_m));
                            psum += p0 + p1;
                            pk[(i + e) >> 1] = OpTraits<OpT>::pack2(p0, p1);
                        }
                    }
                }
            }
            l_run += psum;
        }
        // every thread follows the O barrier phase by phase (the epilogue's parity wait relies on it); P V_{j-1} retired
        // long ago, and with it V_{j-1}'s buffer is free
        if (j > 0) { tc::mbar_wait(bar_o, (j - 1) & 1); tc::fence_after_sync(); }
        // ---- P_j -> TMEM, over this row's consumed scores (two 16-bit values per column): the A operand of P V ----
        if (row_live) {
#pragma unroll
            for (int c = 0; c < kKT / 64; ++c)
                if (c * 64 < nk) tc::tmem_st_32x32b_x32(t_row + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[c * 32]));
            tc::tmem_st_wait();
        }
        tc::fence_before_sync();
        __syncthreads();
        if (warp == 0) {
            tc::fence_after_sync();
            if (j + 1 < nkv) {                              // V_{j-1}'s buffer is free: fetch V_{j+1}
                const int nb = (j + 1) & 1;
                if (tc::elect_one()) {
                    tc::mbar_arrive_expect_tx(bar_v(nb), kKV);
                    tc::tma_load_2d(sV(nb), &tmap_kv, bar_v(nb), col_v, t.seg_start + (j + 1) * kKT);
                }
                __syncwarp();
            }
            tc::mbar_wait(bar_v(j & 1), (j >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t idesc = tc::make_idesc(kFmt, kQT, kHd, 0, 1);
            const uint64_t b0 = tc::smem_desc_mnmajor_sw128(sV(j & 1));
            const int ks = nk / 16;
            if (tc::elect_one()) {
#pragma unroll
                for (int k = 0; k < kKT / 16; ++k)          // 16 keys: 8 columns of P, 16 rows of V = 2048 B
                    if (k < ks) tc::mma_f16_ts(tmem + kKT, tmem + 8 * k, b0 + (uint64_t)k * 128, idesc, (j | k) != 0);
                tc::mma_commit(bar_o);
            }
            __syncwarp();
            if (j + 1 < nkv) issue_qk(j + 1);               // behind P V in the pipe: S_{j+1} overwrites P_j only after it was read
        }
    }

    // ---- epilogue: O / l -> 16-bit rows ---------------------------------------------------------------
    tc::mbar_wait(bar_o, (nkv - 1) & 1);
    tc::fence_after_sync();
    // O / l is packed to 16 bits, staged through the (now idle) Q tile with the usual 16-byte XOR swizzle and
    // written out by each warp for its own 32 rows with 8 lanes per row: every store instruction covers four
    // whole 128-byte rows instead of 32 row fragments.
    if (row_live) {
        const float inv = 1.0f / l_run;
        const uint64_t inv2 = tc::f32x2(inv, inv);
        uint32_t r[64];
        tc::tmem_ld_32x32b_x32(t_row + kKT, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        tc::tmem_ld_32x32b_x32(t_row + kKT + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
        tc::tmem_ld_wait();
        uint8_t* stage = smem_raw + tid * 128;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            uint4 u;
            float a[8];
#pragma unroll
            for (int e = 0; e < 8; e += 2)
                tc::f32x2_split(tc::mul_f32x2(tc::f32x2(__uint_as_float(r[g * 8 + e]), __uint_as_float(r[g * 8 + e + 1])), inv2), a[e], a[e + 1]);
            u.x = OpTraits<OpT>::pack2(a[0], a[1]);
            u.y = OpTraits<OpT>::pack2(a[2], a[3]);
            u.z = OpTraits<OpT>::pack2(a[4], a[5]);
            u.w = OpTraits<OpT>::pack2(a[6], a[7]);
            *reinterpret_cast<uint4*>(stage + ((g ^ (tid & 7)) << 4)) = u;
        }
        __syncwarp();
        const int lane = tid & 31, sub = lane >> 3, g = lane & 7;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int rr = warp * 32 + i * 4 + sub;                 // row of the tile
            const uint4 u = *reinterpret_cast<const uint4*>(smem_raw + rr * 128 + ((g ^ (rr & 7)) << 4));
            if (t.q0 + rr < t.seg_end)
                *reinterpret_cast<uint4*>(out + (size_t)(t.q0 + rr) * d_model + head * kHd + g * 8) = u;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<kAttnTmemCols>(tmem);
}

// Keys per block for a segment of this length.  Measured at 24 960 tokens x 14 heads per launch (visit r2f):
// 128 wins from 2048 tokens (347 vs 355 us; 639 vs 662 us at 4096), is level at 1024 (200 vs 201 us) and loses below
// and at 1196 (244 vs 234 us; 152 vs 135 us at 390; 92 vs 78 us on 104-token windows).
int attention_key_block(int seg_len) {
    static const int forced = [] { const char* e = getenv("Q3A_ATTN_KT"); const int v = e ? atoi(e) : 0; return v == 64 || v == 128 ? v : 0; }();
    if (forced) return forced;
    return seg_len >= 2048 ? 128 : 64;
}

int attention_tiles(const int* cu_seqlens, int n_seg, std::vector<AttnTile>& tiles) {
    tiles.clear();
    int mask = 0;
    for (int s = 0; s < n_seg; ++s) {
        const int kt = attention_key_block(cu_seqlens[s + 1] - cu_seqlens[s]);
        if (cu_seqlens[s + 1] > cu_seqlens[s]) mask |= kt == 128 ? 2 : 1;
        for (int q0 = cu_seqlens[s]; q0 < cu_seqlens[s + 1]; q0 += kQT)
            tiles.push_back(AttnTile{q0, cu_seqlens[s], cu_seqlens[s + 1], kt});
    }
    return mask;
}

double attention_flops(const int* cu_seqlens, int n_seg, int n_heads) {
    double f = 0.0;
    for (int s = 0; s < n_seg; ++s) {
        const double len = cu_seqlens[s + 1] - cu_seqlens[s];
        f += 4.0 * kHd * n_heads * len * len;
    }
    return f;
}

template <class OpT, int KT>
static int attention_launch_kt(const void* qkv, void* out, int T, int d, const AttnTile* tiles_dev, int n_tiles, int n_heads,
                               cudaStream_t stream) {
    using Cfg = AttnCfg<KT>;
    CUtensorMap tm_q, tm_kv;
    Q3A_TRY(make_tmap_2d(&tm_q, qkv, (uint64_t)T, (uint64_t)3 * d, (uint64_t)3 * d, 128));
    Q3A_TRY(make_tmap_2d(&tm_kv, qkv, (uint64_t)T, (uint64_t)3 * d, (uint64_t)3 * d, KT));
    auto kern = attention_kernel<OpT, KT>;
    Q3A_TRY(kernel_smem_optin(reinterpret_cast<const void*>(kern), Cfg::kSmem));   // per (kernel, device), thread-safe
    Q3A_CUDA(launch_pdl(kern, dim3(n_tiles, n_heads), dim3(128), Cfg::kSmem, stream, tm_q, tm_kv, tiles_dev, (OpT*)out, d));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

int attention_launch(const void* qkv, void* out, int T, int n_heads, const AttnTile* tiles_dev, int n_tiles, int fmt,
                     int kt_mask, cudaStream_t stream) {
    if (n_tiles == 0 || T == 0) return Q3A_OK;
    const int d = n_heads * kHd;
    // one launch per block size present; in a mixed launch the CTAs of the other size's tiles leave at once
    if (kt_mask & 1) {
        if (fmt == OP_F16) Q3A_TRY((attention_launch_kt<__half, 64>(qkv, out, T, d, tiles_dev, n_tiles, n_heads, stream)));
        else Q3A_TRY((attention_launch_kt<__nv_bfloat16, 64>(qkv, out, T, d, tiles_dev, n_tiles, n_heads, stream)));
    }
    if (kt_mask & 2) {
        if (fmt == OP_F16) Q3A_TRY((attention_launch_kt<__half, 128>(qkv, out, T, d, tiles_dev, n_tiles, n_heads, stream)));
        else Q3A_TRY((attention_launch_kt<__nv_bfloat16, 128>(qkv, out, T, d, tiles_dev, n_tiles, n_heads, stream)));
    }
    return Q3A_OK;
}

int attention_version() {
    return feature_enabled("Q3A_ATTN_V2", "Q3A_ATTN_V1", kDefaultAttnV2) ? 2 : 1;    // read per call: tests switch it
}

}  // namespace q3a

extern "C" int q3a_test_attention_v(const void* qkv_dev, void* out_dev, int T, int H, const int* cu_seqlens, int n_seg,
                                    int operand_dtype, int version, void* stream) {
    using namespace q3a;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return Q3A_ERR_CUDA; }
    Q3A_TRY(use_device(dev));
    if (!qkv_dev || !out_dev || !cu_seqlens || n_seg < 0 || H <= 0) { set_error("test_attention: bad argument"); return Q3A_ERR_ARG; }
    if (version == 0) version = attention_version();
    if (version != 1 && version != 2) { set_error("test_attention: version must be 0 (default), 1 or 2"); return Q3A_ERR_ARG; }
    std::vector<AttnTile> tiles;
    int kt_mask = 0;
    if (version == 1) kt_mask = attention_tiles(cu_seqlens, n_seg, tiles);
    else attention2_items(cu_seqlens, n_seg, tiles);
    if (tiles.empty()) return Q3A_OK;
    static DevBuf d_tiles_buf;                       // test-only entry point: grow-only scratch, never freed
    Q3A_TRY(d_tiles_buf.reserve(sizeof(AttnTile) * tiles.size()));
    AttnTile* d_tiles = d_tiles_buf.as<AttnTile>();
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // pageable source: the runtime stages the table before returning, so `tiles` may go out of scope
    cudaError_t e = cudaMemcpyAsync(d_tiles, tiles.data(), sizeof(AttnTile) * tiles.size(), cudaMemcpyHostToDevice, st);
    int rc = Q3A_OK;
    if (e != cudaSuccess) { set_error(cudaGetErrorString(e)); rc = Q3A_ERR_CUDA; }
    if (rc == Q3A_OK)
        rc = version == 1 ? attention_launch(qkv_dev, out_dev, T, H, d_tiles, (int)tiles.size(), operand_dtype, kt_mask, st)
                          : attention2_launch(qkv_dev, out_dev, T, H, d_tiles, (int)tiles.size(), operand_dtype, st);
    if (!st) cudaStreamSynchronize(st);               // with a caller's stream the call stays asynchronous
    return rc;
}

extern "C" int q3a_test_attention(const void* qkv_dev, void* out_dev, int T, int H, const int* cu_seqlens, int n_seg,
                                  int operand_dtype, void* stream) {
    return q3a_test_attention_v(qkv_dev, out_dev, T, H, cu_seqlens, n_seg, operand_dtype, 0, stream);
}
