// Variable-length multi-head self-attention on tcgen05, second generation: persistent, warp-specialised, two query
// tiles per CTA ping-ponging on the tensor pipe (head_dim = 64).
//
// Replaces the reference's materialised attention -- K.Q mul_mat, ggml_soft_max_ext(scale 1/8, mask), V.P mul_mat and the
// permutes around them (audio_encoder.cpp:466-486; block-diagonal variant forced_aligner.cpp:737-766,809-821).
//
// Why a second kernel: the first one runs ONE softmax chain per CTA -- MMA commit -> mbarrier ->
// tcgen05.ld -> softmax -> tcgen05.st -> bar.sync -> issue -- and every link of that chain is exposed; it reaches 0.18 of
// the tensor peak.  Here a CTA owns up to 256 query rows of one (segment, head) as two 128-row tiles A and B:
//
//   warps 0-3   softmax warpgroup of tile A (one query row per thread, TMEM lane = row)
//   warps 4-7   softmax warpgroup of tile B
//   warp  8     TMA loader: Q_A, Q_B once per work item, K/V blocks through a 3-stage ring shared by both tiles
//   warp  9     MMA issuer: S_X = Q_X K_j^T, O_X += P_X V_j for X = A, B; everything it signals is a tcgen05.commit
//   warps 10-11 idle (setmaxnreg works on whole warpgroups; they give their registers to the softmax warps)
//
// While warpgroup A runs the softmax of block j, the tensor pipe executes P_B V_j and Q_B K_{j+1}^T, and vice versa, so the
// exponentials (MUFU, the real bound at head_dim 64: 16 ex2/clk/SM against 8 k FLOP/clk/SM) never wait for an MMA round
// trip.  K and V are fetched once for both tiles (half the L2 -> SM traffic of one-tile CTAs).  CTAs are persistent (one per
// SM, static round-robin over (item, head) units): the next item's Q and first K/V blocks are in flight, and its first
// Q K^T issued, while the softmax warps still write the previous item's output.
//
// TMEM (512 columns): S_A [0,128), S_B [128,256), O_A [256,320), O_B [320,384).  P overwrites the row's own scores in
// place (two 16-bit values per column) and is the A operand of P V straight from TMEM; Q K_{j+1}^T is issued behind
// P V_j on the same in-order pipe, so S_{j+1} lands on P_j only after it has been read.  The running maximum is updated
// lazily (O rescaled in TMEM only when a row maximum grows by more than 2^8), and that needs no extra barrier: S_j's
// commit also covers P V_{j-1}, and P V_j is not issued before the warpgroup has signalled P_j.
//
// Key blocks: a segment of L keys is cut into ceil(L/128) equal blocks (rounded up to 16 keys, the MMA granule); rows of
// an item: ceil(L/256) items per segment with equal row counts rounded up to whole warps.  390 tokens -> 2 items of
// 224 / 166 rows x 4 blocks of 112 / 112 / 112 / 54 keys: 13 live warps per (clip, head), against 16 warp slots of four
// 128-row tiles before.
#include "attention.h"
#include "tc.cuh"
#include "tmap.h"

#include <cstdlib>

namespace q3a {

namespace {

constexpr int kHd = 64;
constexpr int kQT = 128;                    // rows of one query tile
constexpr int kKT = 128;                    // most keys of one block
constexpr int kTileBytes = 128 * 64 * 2;    // 16 KB: [128 rows][64 x 16-bit], 128-byte swizzle
constexpr int kStages = 3;                  // K/V ring
constexpr int kThreads = 384;
constexpr int kTmemCols = 512;
constexpr int kColS = 0, kColO = 256;       // S_X at kColS + 128 X, O_X at kColO + 64 X
constexpr float kRescaleThreshold = 8.0f;   // log2 units: p <= 2^8 between rescales
// kPoly (template parameter): every kPoly-th pair of exponentials goes to the FMA pipe (ex2_poly) instead of MUFU; 0 = all
// on MUFU.  The kernel is MUFU-bound once the chain is hidden (16 ex2/clk/SM), so moving a quarter to a third of them
// to the 8 x wider FMA pipe raises the ceiling.  Q3A_ATTN2_POLY=<0|2|3|4> picks the instantiation (default below).
constexpr int kDefaultPoly = 0;

struct Smem {
    static constexpr int kQ = 0;                                    // 2 tiles
    static constexpr int kKV = 2 * kTileBytes;                      // kStages x (K, V)
    static constexpr int kOut = kKV + kStages * 2 * kTileBytes;     // 2 staging tiles for the output rows
    static constexpr int kBars = kOut + 2 * kTileBytes;
    static constexpr int kTotal = kBars + 256 + 1024;               // + barriers + alignment slack
};
// barrier slots (8 bytes each)
enum Bar : int {
    B_QFULL = 0,      // [2]
    B_QEMPTY = 2,     // [2]
    B_KFULL = 4,      // [kStages]
    B_VFULL = 7,      // [kStages]
    B_KVEMPTY = 10,   // [kStages]
    B_SFULL = 13,     // [2]
    B_PFULL = 15,     // [2]
    B_OFULL = 17,     // [2]
    B_OFREE = 19,     // [2]
    B_COUNT = 21
};

__device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 216;"); }
__device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 64;"); }

// 2^x for x in about [-126, 10] on the FMA pipe: round-to-nearest split x = i + r, |r| <= 0.5, degree-4 minimax for 2^r
// (relative error 3e-6, far below the 2^-9 / 2^-12 rounding of P), exponent added as an integer.
__device__ __forceinline__ float ex2_poly(float x) {
    x = fmaxf(x, -125.0f);
    const float magic = 12582912.0f;                       // 1.5 * 2^23
    const float t = x + magic;
    const float r = x - (t - magic);
    float p = fmaf(r, 0.00956051f, 0.05591704f);          // least-squares fit of 2^r on [-0.5, 0.5] in relative error
    p = fmaf(p, r, 0.24024981f);
    p = fmaf(p, r, 0.69312197f);
    p = fmaf(p, r, 0.99999919f);
    return __uint_as_float(__float_as_uint(p) + (__float_as_uint(t) << 23));
}

struct Geometry {                            // of one work item, derived identically by every role
    int q0, rows_a, rows_b, seg_start, seg_len, nkv, kb;
    __device__ __forceinline__ Geometry(const AttnTile& t) {
        q0 = t.q0;
        const int rows = t.pad;
        rows_a = rows < kQT ? rows : kQT;
        rows_b = rows - rows_a;
        seg_start = t.seg_start;
        seg_len = t.seg_end - t.seg_start;
        nkv = (seg_len + kKT - 1) / kKT;
        kb = (((seg_len + nkv - 1) / nkv) + 15) & ~15;
    }
    __device__ __forceinline__ int n_valid(int j) const { const int rem = seg_len - j * kb; return rem < kb ? rem : kb; }
    __device__ __forceinline__ int nk(int j) const { return (n_valid(j) + 15) & ~15; }      // keys fed to the MMAs
};

}  // namespace

template <class OpT, int kPoly>
__global__ void __launch_bounds__(kThreads, 1)
attention2_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const AttnTile* __restrict__ items, int n_items, int n_heads,
                  OpT* __restrict__ out, int d_model) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - tc::smem_u32(smem_raw));
    auto sQ = [&](int x) { return base + Smem::kQ + x * kTileBytes; };
    auto sK = [&](int s) { return base + Smem::kKV + s * 2 * kTileBytes; };
    auto sV = [&](int s) { return base + Smem::kKV + s * 2 * kTileBytes + kTileBytes; };
    auto bar = [&](int b) { return base + Smem::kBars + 8u * b; };
    const uint32_t tmem_slot = base + Smem::kBars + 8u * B_COUNT;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem + Smem::kBars + 8 * B_COUNT);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_units = n_items * n_heads;
    constexpr int kFmt = OpTraits<OpT>::kFmt;

    pdl_trigger();
    if (warp == 8 && lane == 0) {
        tc::prefetch_tmap(&tmap_qkv);
        for (int x = 0; x < 2; ++x) {
            tc::mbar_init(bar(B_QFULL + x), 1); tc::mbar_init(bar(B_QEMPTY + x), 1);
            tc::mbar_init(bar(B_SFULL + x), 1); tc::mbar_init(bar(B_PFULL + x), 4);
            tc::mbar_init(bar(B_OFULL + x), 1); tc::mbar_init(bar(B_OFREE + x), 4);
        }
        for (int s = 0; s < kStages; ++s) {
            tc::mbar_init(bar(B_KFULL + s), 1); tc::mbar_init(bar(B_VFULL + s), 1); tc::mbar_init(bar(B_KVEMPTY + s), 1);
        }
        tc::fence_barrier_init();
    }
    if (warp == 9) tc::tmem_alloc<kTmemCols>(tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot_ptr;
    pdl_wait();                                   // qkv is the previous kernel's output; nothing global was touched above

    if (warp >= 8) {
        setmaxnreg_dec();
        if (warp == 8) {
            // ===================== TMA loader ===========================================================
            uint32_t kv_it = 0, q_ph = 0;                  // bit x: parity of tile x's Q barriers (a tile skips items it has no rows in)
            for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
                const int it = u / n_heads, head = u - it * n_heads;
                const Geometry g(items[it]);
                const int col_q = head * kHd, col_k = d_model + head * kHd, col_v = 2 * d_model + head * kHd;
                tc::mbar_wait(bar(B_QEMPTY + 0), (q_ph & 1) ^ 1);
                q_ph ^= 1u;
                if (tc::elect_one()) {
                    tc::mbar_arrive_expect_tx(bar(B_QFULL + 0), kTileBytes);
                    tc::tma_load_2d(sQ(0), &tmap_qkv, bar(B_QFULL + 0), col_q, g.q0);
                }
                __syncwarp();
                if (g.rows_b > 0) {
                    tc::mbar_wait(bar(B_QEMPTY + 1), ((q_ph >> 1) & 1) ^ 1);
                    q_ph ^= 2u;
                    if (tc::elect_one()) {
                        tc::mbar_arrive_expect_tx(bar(B_QFULL + 1), kTileBytes);
                        tc::tma_load_2d(sQ(1), &tmap_qkv, bar(B_QFULL + 1), col_q, g.q0 + kQT);
                    }
                    __syncwarp();
                }
                for (int j = 0; j < g.nkv; ++j, ++kv_it) {
                    const int s = kv_it % kStages;
                    const uint32_t ph = (kv_it / kStages) & 1;
                    tc::mbar_wait(bar(B_KVEMPTY + s), ph ^ 1);
                    if (tc::elect_one()) {
                        const int row = g.seg_start + j * g.kb;
                        tc::mbar_arrive_expect_tx(bar(B_KFULL + s), kTileBytes);
                        tc::tma_load_2d(sK(s), &tmap_qkv, bar(B_KFULL + s), col_k, row);
                        tc::mbar_arrive_expect_tx(bar(B_VFULL + s), kTileBytes);
                        tc::tma_load_2d(sV(s), &tmap_qkv, bar(B_VFULL + s), col_v, row);
                    }
                    __syncwarp();
                }
            }
        } else if (warp == 9) {
            // ===================== MMA issuer (converged warp; the tcgen05 instructions sit under elect.sync) ======
            // Every per-tile barrier (Q, S, P, O) completes one phase per item in which the tile is live, or per block of
            // such an item, so each tile keeps its own counters.
            uint32_t kv_c = 0, item_ph = 0, blk_ph = 0;      // bit x = phase parity of tile x's per-item / per-block barriers
            const uint32_t idesc_pv = tc::make_idesc(kFmt, kQT, kHd, 0, 1);
            for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
                const int it = u / n_heads;
                const Geometry g(items[it]);
                const int ntile = g.rows_b > 0 ? 2 : 1;
                auto issue_qk = [&](int x, int j, int s) {          // S_x = Q_x K_j^T ; K_j sits in ring stage s
                    const uint32_t idesc = tc::make_idesc(kFmt, kQT, g.nk(j), 0, 0);
                    const uint64_t a = tc::smem_desc_kmajor_sw128(sQ(x)), b = tc::smem_desc_kmajor_sw128(sK(s));
                    const uint32_t d = tmem + kColS + x * kKT;
                    const bool last_qk = (j == g.nkv - 1);
                    if (tc::elect_one()) {
#pragma unroll
                        for (int k = 0; k < kHd / 16; ++k) tc::mma_f16_ss(d, a + 2 * k, b + 2 * k, idesc, k != 0);
                        tc::mma_commit(bar(B_SFULL + x));
                        if (last_qk) tc::mma_commit(bar(B_QEMPTY + x));    // Q_x is free once this has retired
                    }
                    __syncwarp();
                };
                {   // block 0
                    const int s0 = kv_c % kStages;
                    tc::mbar_wait(bar(B_KFULL + s0), (kv_c / kStages) & 1);
#pragma unroll 1
                    for (int x = 0; x < ntile; ++x) {
                        tc::mbar_wait(bar(B_QFULL + x), (item_ph >> x) & 1);
                        tc::fence_after_sync();
                        issue_qk(x, 0, s0);
                    }
                }
#pragma unroll 1
                for (int j = 0; j < g.nkv; ++j) {
                    const uint32_t c = kv_c + j;
                    const int s = c % kStages;
                    const uint32_t ph = (c / kStages) & 1;
                    const int sn = (c + 1) % kStages;
                    const uint32_t phn = ((c + 1) / kStages) & 1;
                    const int ks = g.nk(j) / 16;
#pragma unroll 1
                    for (int x = 0; x < ntile; ++x) {
                        tc::mbar_wait(bar(B_PFULL + x), (blk_ph >> x) & 1);        // P_x(j) is in TMEM
                        blk_ph ^= 1u << x;
                        if (j == 0) tc::mbar_wait(bar(B_OFREE + x), ((item_ph >> x) & 1) ^ 1);   // previous item's O_x was read
                        tc::mbar_wait(bar(B_VFULL + s), ph);
                        tc::fence_after_sync();
                        const uint64_t b0 = tc::smem_desc_mnmajor_sw128(sV(s));
                        const uint32_t d = tmem + kColO + x * kHd, a = tmem + kColS + x * kKT;
                        if (tc::elect_one()) {
#pragma unroll
                            for (int k = 0; k < kKT / 16; ++k)       // 16 keys: 8 columns of P, 16 rows of V = 2048 B
                                if (k < ks) tc::mma_f16_ts(d, a + 8 * k, b0 + (uint64_t)k * 128, idesc_pv, (j | k) != 0);
                            if (j == g.nkv - 1) tc::mma_commit(bar(B_OFULL + x));
                            if (x == ntile - 1) tc::mma_commit(bar(B_KVEMPTY + s));  // K_j and V_j have been consumed
                        }
                        __syncwarp();
                        if (j + 1 < g.nkv) {
                            tc::mbar_wait(bar(B_KFULL + sn), phn);
                            tc::fence_after_sync();
                            issue_qk(x, j + 1, sn);      // behind P V on the in-order pipe: S(j+1) lands on P(j) after it was read
                        }
                    }
                }
                kv_c += g.nkv;
                item_ph ^= (ntile == 2) ? 3u : 1u;
            }
            // nobody may leave while a commit of this warp is still in flight: the last one signals the ring stage of the
            // last block, and it retires behind every other MMA and commit of this CTA
            if (kv_c > 0) tc::mbar_wait(bar(B_KVEMPTY + (kv_c - 1) % kStages), ((kv_c - 1) / kStages) & 1);
        }
    } else {
        // ===================== softmax warpgroups ====================================================
        setmaxnreg_inc();
        const int x = warp >> 2;                           // 0: tile A, 1: tile B
        const int wq = warp & 3;                           // TMEM lane quadrant
        const int row_in_tile = wq * 32 + lane;
        const uint32_t t_row = tmem + ((uint32_t)(wq * 32) << 16);
        const uint32_t t_s = t_row + kColS + x * kKT, t_o = t_row + kColO + x * kHd;
        uint8_t* stage = smem + Smem::kOut + x * kTileBytes;
        const float kScale = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)
        uint32_t blk_n = 0, item_n = 0;
        for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
            const int it = u / n_heads, head = u - it * n_heads;
            const Geometry g(items[it]);
            const int rows_x = x == 0 ? g.rows_a : g.rows_b;
            if (rows_x <= 0) continue;                     // this tile does not exist in the item: no barrier is touched
            const bool live = wq * 32 < rows_x;            // warp-uniform: any valid query row in this warp
            float m_ref = 0.0f, l_run = 0.0f;
            for (int j = 0; j < g.nkv; ++j, ++blk_n) {
                const int n_valid = g.n_valid(j), nk = (n_valid + 15) & ~15;
                tc::mbar_wait(bar(B_SFULL + x), blk_n & 1);
                tc::fence_after_sync();
                if (live) {
                    uint32_t s[kKT];
#pragma unroll
                    for (int c = 0; c < kKT / 32; ++c)
                        if (c * 32 < nk) tc::tmem_ld_32x32b_x32(t_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
                    tc::tmem_ld_wait();
                    float psum = 0.0f;
                    uint32_t* pk = s;                      // P row, packed 16-bit pairs, overwrites consumed scores in place
                    float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
                    for (int i = 0; i < kKT; i += 16) {
                        if (i < nk) {                      // warp-uniform: 16-key groups past the end are skipped
                            if (i + 16 <= n_valid) {
#pragma unroll
                                for (int e = 0; e < 16; e += 8) {
                                    mx0 = fmaxf(mx0, fmaxf(__uint_as_float(s[i + e + 0]), __uint_as_float(s[i + e + 1])));
                                    mx1 = fmaxf(mx1, fmaxf(__uint_as_float(s[i + e + 2]), __uint_as_float(s[i + e + 3])));
                                    mx2 = fmaxf(mx2, fmaxf(__uint_as_float(s[i + e + 4]), __uint_as_float(s[i + e + 5])));
                                    mx3 = fmaxf(mx3, fmaxf(__uint_as_float(s[i + e + 6]), __uint_as_float(s[i + e + 7])));
                                }
                            } else {
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
                            // rare: bring O (complete up to block j-1: S_j's commit covers P V_{j-1}) and l to the new maximum
                            const float alpha = grow ? tc::ex2_approx(m_ref - mx) : 1.0f;
                            if (grow) m_ref = mx;
                            l_run *= alpha;
#pragma unroll
                            for (int c = 0; c < kHd; c += 32) {
                                uint32_t r[32];
                                tc::tmem_ld_32x32b_x32(t_o + c, r);
                                tc::tmem_ld_wait();
#pragma unroll
                                for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                                tc::tmem_st_32x32b_x32(t_o + c, r);
                            }
                            tc::tmem_st_wait();
                        }
                    }
                    const float neg_m = -m_ref;
                    float ps1 = 0.0f;
#pragma unroll
                    for (int i = 0; i < kKT; i += 16) {
                        if (i < nk) {                      // warp-uniform; P columns >= nk are never read
                            if (i + 16 <= n_valid) {
#pragma unroll
                                for (int e = 0; e < 16; e += 2) {
                                    const float a0 = fmaf(__uint_as_float(s[i + e]), kScale, neg_m);
                                    const float a1 = fmaf(__uint_as_float(s[i + e + 1]), kScale, neg_m);
                                    float p0, p1;
                                    if (kPoly > 0 && ((i + e) / 2) % (kPoly > 0 ? kPoly : 1) == kPoly - 1) {
                                        p0 = ex2_poly(a0); p1 = ex2_poly(a1);
                                    } else {
                                        p0 = tc::ex2_approx(a0); p1 = tc::ex2_approx(a1);
                                    }
                                    psum += p0; ps1 += p1;
                                    pk[(i + e) >> 1] = OpTraits<OpT>::pack2(p0, p1);
                                }
                            } else {
#pragma unroll
                                for (int e = 0; e < 16; e += 2) {
                                    float p0 = 0.0f, p1 = 0.0f;
                                    if (i + e < n_valid) p0 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e]), kScale, neg_m));
                                    if (i + e + 1 < n_valid) p1 = tc::ex2_approx(fmaf(__uint_as_float(s[i + e + 1]), kScale, neg_m));
                                    psum += p0; ps1 += p1;
                                    pk[(i + e) >> 1] = OpTraits<OpT>::pack2(p0, p1);
                                }
                            }
                        }
                    }
                    l_run += psum + ps1;
                    // P_j -> TMEM over this row's consumed scores: the A operand of P V
#pragma unroll
                    for (int c = 0; c < kKT / 64; ++c)
                        if (c * 64 < nk) tc::tmem_st_32x32b_x32(t_s + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[c * 32]));
                    tc::tmem_st_wait();
                }
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(B_PFULL + x));
            }
            // ---- epilogue: O / l -> 16-bit rows ------------------------------------------------------
            tc::mbar_wait(bar(B_OFULL + x), item_n & 1);
            ++item_n;
            tc::fence_after_sync();
            if (!live) {
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(B_OFREE + x));
            } else {
                uint32_t r[64];
                tc::tmem_ld_32x32b_x32(t_o, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
                tc::tmem_ld_32x32b_x32(t_o + 32, *reinterpret_cast<uint32_t(*)[32]>(&r[32]));
                tc::tmem_ld_wait();
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(B_OFREE + x));  // the issuer may overwrite O_x with the next item
                // packed to 16 bits, staged through this warp's own 4 KB of the tile's staging buffer with the usual
                // 16-byte XOR swizzle, written out with 8 lanes per row: every store covers four whole 128-byte rows
                const float inv = 1.0f / l_run;
                uint8_t* srow = stage + row_in_tile * 128;
#pragma unroll
                for (int gq = 0; gq < 8; ++gq) {
                    uint4 v;
                    v.x = OpTraits<OpT>::pack2(__uint_as_float(r[gq * 8 + 0]) * inv, __uint_as_float(r[gq * 8 + 1]) * inv);
                    v.y = OpTraits<OpT>::pack2(__uint_as_float(r[gq * 8 + 2]) * inv, __uint_as_float(r[gq * 8 + 3]) * inv);
                    v.z = OpTraits<OpT>::pack2(__uint_as_float(r[gq * 8 + 4]) * inv, __uint_as_float(r[gq * 8 + 5]) * inv);
                    v.w = OpTraits<OpT>::pack2(__uint_as_float(r[gq * 8 + 6]) * inv, __uint_as_float(r[gq * 8 + 7]) * inv);
                    *reinterpret_cast<uint4*>(srow + ((gq ^ (row_in_tile & 7)) << 4)) = v;
                }
                __syncwarp();
                const int sub = lane >> 3, gq = lane & 7;
                const int tile_row0 = g.q0 + x * kQT;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = wq * 32 + i * 4 + sub;                 // row of the tile
                    const uint4 v = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((gq ^ (rr & 7)) << 4));
                    if (rr < rows_x)
                        *reinterpret_cast<uint4*>(out + (size_t)(tile_row0 + rr) * d_model + head * kHd + gq * 8) = v;
                }
                __syncwarp();                                             // the staging rows are reused by the next item
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 9) tc::tmem_dealloc<kTmemCols>(tmem);
}

// ---- host side --------------------------------------------------------------------------------------------------
int attention2_item_count(int seg_len) { return (seg_len + 2 * kQT - 1) / (2 * kQT); }

void attention2_items(const int* cu_seqlens, int n_seg, std::vector<AttnTile>& items) {
    items.clear();
    for (int s = 0; s < n_seg; ++s) {
        const int len = cu_seqlens[s + 1] - cu_seqlens[s];
        if (len <= 0) continue;
        const int n = attention2_item_count(len);
        const int rows = (((len + n - 1) / n) + 31) & ~31;                 // whole warps, <= 256
        for (int q = 0; q < len; q += rows)
            items.push_back(AttnTile{cu_seqlens[s] + q, cu_seqlens[s], cu_seqlens[s + 1], std::min(rows, len - q)});
    }
}

template <class OpT, int kPoly>
static int attention2_launch_t(const void* qkv, void* out, int T, int d, const AttnTile* items_dev, int n_items, int n_heads,
                               cudaStream_t stream) {
    CUtensorMap tm;
    Q3A_TRY(make_tmap_2d(&tm, qkv, (uint64_t)T, (uint64_t)3 * d, (uint64_t)3 * d, 128));
    auto kern = attention2_kernel<OpT, kPoly>;
    Q3A_TRY(kernel_smem_optin(reinterpret_cast<const void*>(kern), Smem::kTotal));
    const long long units = (long long)n_items * n_heads;
    const int grid = (int)std::min<long long>(units, sm_count_cached());
    Q3A_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), Smem::kTotal, stream, tm, items_dev, n_items, n_heads, (OpT*)out, d));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

int attention2_launch(const void* qkv, void* out, int T, int n_heads, const AttnTile* items_dev, int n_items, int fmt,
                      cudaStream_t stream) {
    if (n_items == 0 || T == 0) return Q3A_OK;
    const int d = n_heads * kHd;
    const char* pe = getenv("Q3A_ATTN2_POLY");
    const int poly = pe ? atoi(pe) : kDefaultPoly;
#define Q3A_A2(P)                                                                                                          \
    return fmt == OP_F16 ? attention2_launch_t<__half, P>(qkv, out, T, d, items_dev, n_items, n_heads, stream)            \
                         : attention2_launch_t<__nv_bfloat16, P>(qkv, out, T, d, items_dev, n_items, n_heads, stream)
    switch (poly) {
        case 2: Q3A_A2(2);
        case 3: Q3A_A2(3);
        case 4: Q3A_A2(4);
        default: Q3A_A2(0);
    }
#undef Q3A_A2
}

}  // namespace q3a
