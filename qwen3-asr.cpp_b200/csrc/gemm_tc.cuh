// Persistent warp-specialised tcgen05 GEMM for sm_100a on CTA pairs:  D[M,N] = A[M,K] * W[N,K]^T, fp32 accumulate.
//
// Two CTAs on the two SMs of a TPC (cluster of 2) compute one 256 x BLOCK_N tile with
// tcgen05.mma.cta_group::2: each CTA stages its own 128 rows of A and its own BLOCK_N/2 rows of W, the
// tensor cores of both SMs read both halves, so every byte that crosses L2 -> SM feeds twice the math
// of a single-CTA 128 x BLOCK_N tile (the single-CTA version of this kernel measured L2-bound at
// ~44 B/clk/SM, profiles/r1a_*).  Accumulators: 128 lanes x BLOCK_N fp32 columns of TMEM per CTA,
// two buffers, so the epilogue of tile i overlaps the MMAs of tile i+1.
//
//   warp 0       TMA producer (one lane, both CTAs): A/W tiles into a kStages-deep smem ring; completion
//                bytes of both CTAs are signalled on the leader CTA's "full" barrier
//   warp 1       MMA issuer (one lane, leader CTA only); tcgen05.commit multicasts "slot free" and
//                "accumulator ready" to both CTAs
//   warp 2       TMEM allocator / deallocator (both CTAs, cta_group::2)
//   warp 3       idle
//   warps 4..11  epilogue, two warps per TMEM lane quadrant, alternating 32-column chunks:
//                tcgen05.ld -> fused functor -> swizzled smem staging -> TMA store (coalesced, async);
//                the residual stream is brought in by TMA loads one chunk ahead.  The convolution
//                epilogues scatter rows and store straight from registers instead.
//
//   warps 12..15 (kWq8 only) fused Q8_0 dequantisation: the W tile arrives as int8 (TMA, raw ring), these warps expand
//                it -- value = d * q per 32-element block, ggml's Q8_0 definition -- into the swizzled 16-bit
//                operand stage the MMA reads, and signal the "full" barrier next to the A tile's TMA bytes
//
// Both operands are K-major 16-bit (bf16 or f16) staged with 128-byte swizzle, K blocks of 64.  The A-side
// coordinates come from a "loader" policy so the same kernel runs the Linear layers (2-D tensor map) and
// the stride-2 3x3 convolutions of the stem as implicit GEMM (4-D tensor map over parity-split
// channels-last planes, see stem.h).
#pragma once
#include "common.cuh"
#include "tc.cuh"

#include <type_traits>

namespace q3a {

constexpr int kBlockM = 128;                // rows per CTA; a pair covers 256
constexpr int kBlockK = 64;                 // 64 x 16-bit = 128 B = one swizzle span
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 384;           // 4 control warps + 8 epilogue warps
constexpr int kDequantWarps = 4;            // Q8_0 weight path: + 4 warps that expand int8 blocks into the operand stage
constexpr int kEpiWarps = 8;                // default; the kernel takes the count as a template parameter (kEW: 8 or 12)
constexpr int kTmemCols = 512;              // two accumulator buffers of up to 256 columns
constexpr int kSmemLimit = 232448;          // 227 KB opt-in maximum per CTA

// kEpiBufBytes: bytes of one epilogue staging buffer set (32 rows; a dual-output epilogue stages the f32 chunk and its
// 16-bit copy), kEpiBufs: buffer sets per epilogue warp
template <int BLOCK_N, int kEpiBufBytes, int kEpiBufs, bool kWq8 = false, int kEW = kEpiWarps> struct GemmSmem {
    static constexpr int kABytes = kBlockM * kBlockK * 2;              // 16 KB
    static constexpr int kBBytes = (BLOCK_N / 2) * kBlockK * 2;        // this CTA's half of the W tile
    static constexpr int kRawBytes = kWq8 ? (BLOCK_N / 2) * kBlockK : 0;   // the same half as int8 (Q8_0 weights)
    static constexpr int kStageBytes = kABytes + kBBytes + kRawBytes;
    static constexpr int kEpiBytes = kEW * kEpiBufs * kEpiBufBytes;
    static constexpr int kBarrierBytes = 512;
    static constexpr int kAvail = kSmemLimit - 1024 - kBarrierBytes - kEpiBytes;
    static constexpr int kStages = kAvail / kStageBytes > 6 ? 6 : kAvail / kStageBytes;
    static constexpr int kTotal = kStages * kStageBytes + kEpiBytes + kBarrierBytes + 1024;   // +1024: manual alignment
    static_assert(kStageBytes % 1024 == 0, "stage must keep 1024-B swizzle alignment");
    static_assert(kStages >= 3, "need at least three stages");
};
// the shared-memory plan of a (BLOCK_N, epilogue) instantiation, for kernel and host alike
template <class Epi> struct EpiDual { static constexpr bool value = false; };     // specialised by dual-output epilogues
template <int BLOCK_N, class Epi, bool kWq8 = false, int kEW = kEpiWarps>
using GemmSmemFor = GemmSmem<BLOCK_N, Epi::kStaged ? 32 * 32 * Epi::kOutBytes + (EpiDual<Epi>::value ? 32 * 32 * 2 : 0) : 0,
                             Epi::kStaged ? (Epi::kResidual && !EpiDual<Epi>::value ? 3 : 2) : 0, kWq8, kEW>;
// per-row state an epilogue may carry across the column chunks of a tile (LayerNorm-folded epilogues keep mean / rstd)
struct NoRowState {};
template <class Epi, class = void> struct EpiRowState { using type = NoRowState; };
template <class Epi> struct EpiRowState<Epi, decltype((void)sizeof(typename Epi::RowState))> { using type = typename Epi::RowState; };
// per-column constants of a 32-column chunk (bias, fold constants), one column per lane: requested before the accumulator
// load so that their latency runs under it, read back by column with a shuffle.  (Every thread loading all 32 values itself
// exposed four dependent L2 round trips per chunk: 13 % of the stall samples of the FFN-up GEMM, profiles/r2n.)
struct NoCols {};
template <class Epi, class = void> struct EpiCols { using type = NoCols; };
template <class Epi> struct EpiCols<Epi, decltype((void)sizeof(typename Epi::Cols))> { using type = typename Epi::Cols; };

// ---- A-operand coordinate policies (m128 = index of this CTA's 128-row block) -----------------
// The producer and the MMA issuer walk the K blocks of a tile with a small iterator that is advanced
// incrementally: both loops run on a single thread and every integer division in them is latency on the
// critical issue path (the first pair version spent ~1000 clk per K block in the MMA loop, profiles/r1b).
// Linear: A is [M][K] row-major, tensor map dims {K, M}, box {64, 128}.
struct LoaderLinear {
    int num_k_blocks;
    struct It { int kb; };
    __device__ __forceinline__ It begin() const { return It{0}; }
    __device__ __forceinline__ void next(It& it) const { ++it.kb; }
    __device__ __forceinline__ void load_a(uint32_t dst, const void* tmap, uint32_t bar, int m128, const It& it) const {
        tc::tma_load_2d_pair(dst, tmap, bar, it.kb * kBlockK, m128 * kBlockM);
    }
    __device__ __forceinline__ int k_steps(const It&) const { return kBlockK / kUmmaK; }
    __device__ __forceinline__ int k_coord(const It& it) const { return it.kb * kBlockK; }      // W column of this K block
};
// Conv (3x3, stride 2, pad 1) over parity-split planes (layout in stem.h): tensor map dims
// {C, Fh, R, 4 planes}, box {64, Fh, 128/Fh, 1}.  Output row m = R_out * Fh + of where R_out indexes
// (chunk, 1 + ot).  K index = tap * (c_blocks*64) + c, tap = dt*3 + df (dt: time tap, df: mel tap).
//   dt=0 -> odd-time plane, row R_out-1 ; dt=1 -> even-time plane, row R_out ; dt=2 -> odd-time plane, row R_out
//   df=0 -> odd-mel plane, col of-1     ; df=1 -> even-mel plane, col of     ; df=2 -> odd-mel plane, col of
// Out-of-range coordinates (-1, or channels >= C in the last channel block) are zero-filled by TMA,
// which is exactly the convolution's zero padding.
struct LoaderConv {
    int num_k_blocks;       // 9 * c_blocks
    int c_blocks;           // ceil(C / 64)
    int rows_per_tile;      // 128 / Fh
    int last_block_steps;   // k-steps that hold real channels in the last channel block of a tap
    int cb_major;           // 1: channel block outermost, the 9 taps inside it (default); 0: tap outermost
    struct It { int cb, df, dt; };
    __device__ __forceinline__ It begin() const { return It{0, 0, 0}; }
    // K order.  The 9 taps of one channel block read overlapping boxes of the same planes (df = 0 / 2 differ by one mel
    // column, dt = 0 / 2 by one row): walking them back to back re-reads lines that are still in L2, where the tap-major
    // order came back to a channel block 8 K blocks later, after ~150 CTAs had streamed ~90 MB through the 126 MB L2
    // (round 1: 1.48x DRAM read amplification on conv2).  Only the fp32 summation order changes.
    __device__ __forceinline__ void next(It& it) const {
        if (cb_major) { if (++it.df == 3) { it.df = 0; if (++it.dt == 3) { it.dt = 0; ++it.cb; } } }
        else if (++it.cb == c_blocks) { it.cb = 0; if (++it.df == 3) { it.df = 0; ++it.dt; } }
    }
    __device__ __forceinline__ int k_coord(const It& it) const { return ((it.dt * 3 + it.df) * c_blocks + it.cb) * kBlockK; }
    __device__ __forceinline__ void load_a(uint32_t dst, const void* tmap, uint32_t bar, int m128, const It& it) const {
        const int plane = ((it.dt != 1) ? 2 : 0) + ((it.df != 1) ? 1 : 0);
        tc::tma_load_4d_pair(dst, tmap, bar, it.cb * kBlockK, (it.df == 0) ? -1 : 0,
                             m128 * rows_per_tile + ((it.dt == 0) ? -1 : 0), plane);
    }
    __device__ __forceinline__ int k_steps(const It& it) const {
        return (it.cb == c_blocks - 1) ? last_block_steps : kBlockK / kUmmaK;
    }
};

// byte offset of 16-byte chunk j of row r inside a 32-row staging buffer whose rows are kRowBytes long,
// in the layout TMA expects for SWIZZLE_128B (128-B rows) / SWIZZLE_64B (64-B rows)
template <int kRowBytes>
__device__ __forceinline__ uint32_t stage_off(int r, int j) {
    if (kRowBytes == 128) return (uint32_t)(r * 128 + ((j ^ (r & 7)) << 4));
    return (uint32_t)(r * 64 + ((j ^ ((r >> 1) & 3)) << 4));
}

// Q8_0 weights as they sit in device memory (de-interleaved ggml blocks: the int8 plane is read through tmap_b, the
// scales straight from global memory): scale of element (n, k) = d[n * ld_d + k / 32]
struct Q8Scales {
    const __half* d = nullptr;
    int ld_d = 0;               // K / 32
};

// ---- the kernel ------------------------------------------------------------------------------
// Epilogue contracts (epilogues.cuh):
//   staged   struct E { static constexpr bool kStaged = true; using OutT = float | bf16 | f16;
//                       static constexpr bool kResidual;      // out += previous contents of out (f32 only)
//                       __device__ void apply(int m, int n0, int nvalid, float (&v)[32]) const; }
//            v holds the accumulators of row m, columns n0..n0+31 on entry and the values to store on
//            exit (the kernel adds the residual afterwards); tmap_c describes the output matrix.
//            Optional: `using RowState = ...; RowState row_state(int m) const;` -> apply(m, n0, nvalid, v, rs) gets it
//            (loaded once per tile, e.g. the row's LayerNorm statistics).
//            Optional dual output (EpiDual<E>::value, f32 epilogues only): the final f32 chunk is also stored as 16-bit
//            operand type through tmap_c2, and per-row (sum, sum of squares) partials of the final values go to
//            epi.stats[m * epi.n_part + n_blk * 2 + half] -- the next GEMM folds the LayerNorm in (epilogues.cuh).
//   direct   struct E { static constexpr bool kStaged = false;
//                       __device__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const; }
//            called for rows m < M; only the first nvalid (multiple of 8) columns belong to the output.
// kEW epilogue warps: 8 (two per TMEM lane quadrant) or 12 (three).  The 16-bit-output GEMMs with K = 896 are paced by
// their epilogue (FFN-up: 65 % tensor-active with 8 warps); a third warp per quadrant hides more of its latency.
template <int BLOCK_N, int kFmt, class Loader, class Epilogue, bool kWq8 = false, int kEW = kEpiWarps>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 + 32 * kEW + (kWq8 ? 32 * kDequantWarps : 0), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_c2, int M, int N,
               Loader loader, Epilogue epi, Q8Scales q8 = Q8Scales{}) {
    constexpr bool kStaged = Epilogue::kStaged;
    constexpr bool kResidual = Epilogue::kResidual;
    constexpr bool kDual = EpiDual<Epilogue>::value;
    constexpr int kOutBytes = Epilogue::kOutBytes;                     // 4 or 2 (staged only)
    constexpr int kRowBytes = 32 * kOutBytes;                          // staging row: 32 columns
    constexpr int kMainBytes = kStaged ? 32 * kRowBytes : 0;           // the chunk in the output type
    constexpr int kBufBytes = kMainBytes + (kDual ? 32 * 64 : 0);      // + its 16-bit copy (64-byte rows)
    static_assert(!kDual || (kStaged && kOutBytes == 4), "dual output: staged f32 epilogues only");
    // Buffer sets per epilogue warp.  Residual epilogues rotate three (next chunk's residual landing | current chunk |
    // previous chunk being stored); the dual-output ones make do with two 6-KB sets -- the prefetch of chunk i+1 then
    // waits for the store of chunk i-1 to have READ its buffer (lane 0 only, ~100 clk, under the other lanes' tcgen05.ld) --
    // so that the stage ring keeps its depth.
    constexpr int kEpiBufs = kStaged ? (kResidual && !kDual ? 3 : 2) : 0;
    using S = GemmSmemFor<BLOCK_N, Epilogue, kWq8, kEW>;
    constexpr int kStages = S::kStages;
    constexpr int kEpiPerQuad = kEW / 4;                                // epilogue warps sharing a TMEM lane quadrant
    static_assert(kEW == 4 || kEW == 8 || kEW == 12, "4, 8 or 12 epilogue warps");
    static_assert(!EpiDual<Epilogue>::value || kEW <= 8, "the dual-output statistics layout has two partials per column block");
    static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "BLOCK_N must be a multiple of 16, at most 256");
    static_assert(!kStaged || BLOCK_N % 32 == 0, "staged epilogues store whole 32-column chunks");
    static_assert(!kResidual || kOutBytes == 4, "the residual stream is fp32");

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - tc::smem_u32(smem_raw));
    const uint32_t epi_base = smem_base + kStages * S::kStageBytes;
    const uint32_t bar_base = epi_base + S::kEpiBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
    auto res_bar = [&](int w, int b) { return bar_base + 8u * (2 * kStages + 5 + w * 3 + b); };   // 24 barriers
    auto raw_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 5 + 3 * kEW + s); };        // Q8 raw ring
    static_assert(8 * (2 * 6 + 5 + 3 * kEW + 6) <= S::kBarrierBytes, "barrier block too small");
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes + S::kEpiBytes + 8 * (2 * kStages + 4));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();                       // 0 = leader (issues the MMAs)
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    const int m_blocks = ceil_div(M, 2 * kBlockM), n_blocks = ceil_div(N, BLOCK_N);
    const int num_tiles = m_blocks * n_blocks;
    const int num_kb = loader.num_k_blocks;

    pdl_trigger();
    if (warp == 0 && lane == 0) {
        tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
        if (kStaged) tc::prefetch_tmap(&tmap_c);
        if (kDual) tc::prefetch_tmap(&tmap_c2);
    }
    if (warp == 1 && lane == 0) {
        // Q8 path: the dequant warps of both CTAs arrive on the leader's "full" barrier next to the producer's expect_tx
        for (int s = 0; s < kStages; ++s) {
            tc::mbar_init(full_bar(s), kWq8 ? 1 + 2 * kDequantWarps : 1); tc::mbar_init(empty_bar(s), 1);
            if (kWq8) tc::mbar_init(raw_bar(s), 1);
        }
        for (int s = 0; s < 2; ++s) { tc::mbar_init(tfull_bar(s), 1); tc::mbar_init(tempty_bar(s), 2 * kEW); }
        for (int w = 0; w < kEW; ++w)
            for (int b = 0; b < 3; ++b) tc::mbar_init(res_bar(w, b), 1);
        tc::fence_barrier_init();
    }
    if (warp == 2) tc::tmem_alloc_pair<kTmemCols>(tmem_slot);
    tc::fence_before_sync();
    tc::cluster_sync();                       // barriers of both CTAs initialised before any remote signal
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;
    pdl_wait();                               // everything above overlapped the previous kernel's tail

    if (warp == 0) {
        // ===================== TMA producer (both CTAs; converged warp, one elected lane issues) ==========
        int s = 0;
        uint32_t ph = 0;
        const int b_row0 = (int)rank * (BLOCK_N / 2);
        for (int tile = pair; tile < num_tiles; tile += num_pairs) {
            // n fastest: pairs running side by side share the A tile through L2
            const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
            const int m128 = 2 * m_blk + (int)rank;
            const int b_row = n_blk * BLOCK_N + b_row0;
            typename Loader::It li = loader.begin();
            for (int kb = 0; kb < num_kb; ++kb) {
                tc::mbar_wait(empty_bar(s), ph ^ 1);
                if (tc::elect_one()) {
                    const uint32_t a_dst = smem_base + s * S::kStageBytes;
                    const uint32_t bar = tc::mapa(full_bar(s), 0);
                    if constexpr (!kWq8) {
                        if (rank == 0) tc::mbar_arrive_expect_tx(full_bar(s), 2 * S::kStageBytes);
                        loader.load_a(a_dst, &tmap_a, bar, m128, li);
                        tc::tma_load_2d_pair(a_dst + S::kABytes, &tmap_b, bar, loader.k_coord(li), b_row);
                    } else {
                        // A goes to the operand stage as usual; W arrives as int8 in this CTA's raw buffer (own barrier)
                        if (rank == 0) tc::mbar_arrive_expect_tx(full_bar(s), 2 * S::kABytes);
                        loader.load_a(a_dst, &tmap_a, bar, m128, li);
                        tc::mbar_arrive_expect_tx(raw_bar(s), S::kRawBytes);
                        tc::tma_load_2d(a_dst + S::kABytes + S::kBBytes, &tmap_b, raw_bar(s), loader.k_coord(li), b_row);
                    }
                }
                __syncwarp();
                loader.next(li);
                if (++s == kStages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA; the warp stays converged, one elected lane issues) ====
        if (rank == 0) {
            constexpr uint32_t idesc = tc::make_idesc(kFmt, 2 * kBlockM, BLOCK_N, 0, 0);
            const uint64_t desc_hi = tc::smem_desc_kmajor_sw128(0);           // everything but the start address
            int s = 0;
            uint32_t ph = 0, local = 0;
            for (int tile = pair; tile < num_tiles; tile += num_pairs, ++local) {
                const int as = local & 1;
                const uint32_t aph = (local >> 1) & 1;
                tc::mbar_wait(tempty_bar(as), aph ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tmem_base + as * 256;
                typename Loader::It li = loader.begin();
                uint32_t acc = 0;
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int steps = loader.k_steps(li);
                    const uint32_t a_addr = smem_base + s * S::kStageBytes;
                    const uint64_t a_desc = desc_hi | (uint64_t)((a_addr >> 4) & 0x3FFF);
                    const uint64_t b_desc = desc_hi | (uint64_t)(((a_addr + S::kABytes) >> 4) & 0x3FFF);
                    tc::mbar_wait(full_bar(s), ph);
                    tc::fence_after_sync();
                    if (tc::elect_one()) {
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                            // advance 16 elements (32 B) along K inside the swizzle span: +2 in (addr>>4) units
                            if (k < steps) tc::mma_f16_ss_pair(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, acc | k);
                        }
                        tc::mma_commit_pair(empty_bar(s), 3);                    // frees the slot in both CTAs
                    }
                    __syncwarp();
                    acc = 1;
                    loader.next(li);
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
                if (tc::elect_one()) tc::mma_commit_pair(tfull_bar(as), 3);      // accumulators complete
                __syncwarp();
            }
        }
    } else if (kWq8 && warp >= 4 + kEW) {
        // ===================== Q8_0 dequantisation (both CTAs): raw int8 tile -> swizzled 16-bit operand tile =========
        // One W row per thread: 64 int8 (4 x LDS.128, chunk order rotated by row so that a quarter-warp hits eight
        // different 16-byte bank groups) -> value = d * q exactly as ggml defines Q8_0 (one f32 rounding-free product:
        // with x = 1 + (q+128) * 2^-15 built by a byte permute, fma(x, 2^15 d, -(2^15 + 128) d) = q * d exactly) ->
        // rounded once to the operand type, the same value the dequantise-at-load path stores -> 8 x STS.128 into the
        // 128-byte-swizzled K-major layout (conflict-free by construction).
        using OpT = typename std::conditional<kFmt == 1, __nv_bfloat16, __half>::type;
        const int r = (warp - 4 - kEW) * 32 + lane;                  // row of this CTA's half tile
        const bool has_row = r < BLOCK_N / 2;
        const uint32_t full_leader0 = tc::mapa(full_bar(0), 0);
        int s = 0;
        uint32_t ph = 0;
        for (int tile = pair; tile < num_tiles; tile += num_pairs) {
            const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
            const int n = n_blk * BLOCK_N + (int)rank * (BLOCK_N / 2) + r;
            const bool live = has_row && n < N;
            const uint32_t* drow = reinterpret_cast<const uint32_t*>(q8.d + (size_t)(live ? n : 0) * q8.ld_d);
            for (int kb = 0; kb < num_kb; ++kb) {
                const uint32_t dd = live ? __ldg(drow + kb) : 0u;          // the two block scales of this K block (f16 pair)
                tc::mbar_wait(raw_bar(s), ph);
                if (has_row) {
                    const uint8_t* stage = smem_gen + s * S::kStageBytes;
                    const uint8_t* raw = stage + S::kABytes + S::kBBytes + r * 64;
                    uint8_t* dst = const_cast<uint8_t*>(stage) + S::kABytes + r * 128;
                    const __half2 dh = *reinterpret_cast<const __half2*>(&dd);
                    const float d0 = __low2float(dh), d1 = __high2float(dh);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int c = (i + (r >> 1)) & 3;                   // 16 int8 = elements 16c .. 16c+15 of the K block
                        const uint4 qv = *reinterpret_cast<const uint4*>(raw + c * 16);
                        const float dsc = c < 2 ? d0 : d1;
                        const float mul = dsc * 32768.0f, add = -(32768.0f + 128.0f) * dsc;
                        const uint32_t w4[4] = {qv.x ^ 0x80808080u, qv.y ^ 0x80808080u, qv.z ^ 0x80808080u, qv.w ^ 0x80808080u};
                        float f[16];
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            // byte e of the chunk into bits [8,16) of 0x3F800000: 1 + u8 * 2^-15
                            const uint32_t bits = __byte_perm(w4[e >> 2], 0x3F800000u, 0x7604 | ((e & 3) << 4));
                            f[e] = fmaf(__uint_as_float(bits), mul, add);
                        }
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            uint4 u;
                            u.x = OpTraits<OpT>::pack2(f[8 * h + 0], f[8 * h + 1]);
                            u.y = OpTraits<OpT>::pack2(f[8 * h + 2], f[8 * h + 3]);
                            u.z = OpTraits<OpT>::pack2(f[8 * h + 4], f[8 * h + 5]);
                            u.w = OpTraits<OpT>::pack2(f[8 * h + 6], f[8 * h + 7]);
                            const int j = 2 * c + h;                        // 16-byte chunk of the 128-byte operand row
                            *reinterpret_cast<uint4*>(dst + ((j ^ (r & 7)) << 4)) = u;
                        }
                    }
                }
                tc::fence_proxy_async();                                    // generic-proxy writes -> visible to the tensor cores
                __syncwarp();
                if (lane == 0) tc::mbar_arrive_cluster(full_leader0 + 8u * s);
                if (++s == kStages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp >= 4 && warp < 4 + kEW) {
        // ===================== epilogue (both CTAs) =====================
        const int ew = warp - 4;
        const int q = warp & 3;                       // TMEM lane quadrant this warp may read
        const int hf = ew >> 2;                       // which share of the 32-column chunks (0 .. kEpiPerQuad-1)
        const uint32_t tempty_leader0 = tc::mapa(tempty_bar(0), 0), tempty_leader1 = tc::mapa(tempty_bar(1), 0);
        const uint32_t my_bufs = epi_base + ew * kEpiBufs * kBufBytes;
        uint8_t* my_bufs_gen = smem_gen + kStages * S::kStageBytes + ew * kEpiBufs * kBufBytes;
        uint32_t seq = 0;                             // chunks processed by this warp (staging buffer rotation)
        uint32_t rph = 0;                             // phase bits of the residual barriers
        uint32_t local = 0;

        auto chunks_of = [&](int n_blk) { return ceil_div(min(BLOCK_N, N - n_blk * BLOCK_N), 32); };
        // residual prefetch for the chunk with sequence number sq at (tile, chunk c)
        auto issue_res = [&](uint32_t sq, int tile, int c) {
            const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
            const int b = sq % kEpiBufs;
            tc::mbar_arrive_expect_tx(res_bar(ew, b), kMainBytes);
            tc::tma_load_2d(my_bufs + b * kBufBytes, &tmap_c, res_bar(ew, b), n_blk * BLOCK_N + c * 32,
                            (2 * m_blk + (int)rank) * kBlockM + q * 32);
        };
        // first tile at or after t (in this pair's sequence) in which this warp owns a chunk
        auto tile_with_chunk = [&](int t) {
            while (t < num_tiles && hf >= chunks_of(t % n_blocks)) t += num_pairs;
            return t;
        };
        if (kResidual && lane == 0) {
            const int t0 = tile_with_chunk(pair);
            if (t0 < num_tiles) issue_res(0, t0, hf);
        }

        for (int tile = pair; tile < num_tiles; tile += num_pairs, ++local) {
            const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
            const int as = local & 1;
            const uint32_t aph = (local >> 1) & 1;
            const int nch = chunks_of(n_blk);
            if constexpr (kResidual) {
                // The residual stream was written two or three kernels ago and has left L2 since: start this warp's chunks of
                // the pair's NEXT tile on their way from HBM now, a whole tile ahead of the shared-memory loads that use them.
                const int nt = tile + num_pairs;
                if (lane == 0 && nt < num_tiles && epi.prefetch) {
                    const int m_n = nt / n_blocks, n_n = nt - m_n * n_blocks;
                    const int nch_n = chunks_of(n_n);
                    for (int c = hf; c < nch_n; c += kEpiPerQuad)
                        tc::tma_prefetch_2d(&tmap_c, n_n * BLOCK_N + c * 32, (2 * m_n + (int)rank) * kBlockM + q * 32);
                }
            }
            tc::mbar_wait(tfull_bar(as), aph);
            tc::fence_after_sync();
            const int m0 = (2 * m_blk + (int)rank) * kBlockM + q * 32;
            const int m = m0 + lane;
            const uint32_t t_row = tmem_base + as * 256 + ((uint32_t)(q * 32) << 16);
            typename EpiRowState<Epilogue>::type rs{};
            if constexpr (!std::is_same<typename EpiRowState<Epilogue>::type, NoRowState>::value) rs = epi.row_state(m);
            // dual output: this warp's share of the row's statistics.  The two partials of a column block are the sums over
            // its even and its odd 32-column chunks -- with eight warps that is one warp each, with four the one warp keeps
            // both, so the statistics (and everything downstream) are bit-identical for either count.
            float st_sum = 0.0f, st_sq = 0.0f, st_sum1 = 0.0f, st_sq1 = 0.0f;
#pragma unroll 1
            for (int c = hf; c < nch; c += kEpiPerQuad) {
                const int n0 = n_blk * BLOCK_N + c * 32;
                const int nvalid = min(32, min(BLOCK_N - c * 32, N - n0));
                auto prefetch_next_residual = [&]() {
                    if (kResidual && lane == 0) {
                        // next chunk of this warp: same tile, or the first chunk of its next tile
                        int nt = tile, nc = c + kEpiPerQuad;
                        if (nc >= nch) { nt = tile_with_chunk(tile + num_pairs); nc = hf; }
                        if (nt < num_tiles) {
                            tc::bulk_wait_group_read<(kEpiBufs > 2 ? kEpiBufs - 2 : 0)>();   // the store that last used that buffer has read it
                            issue_res(seq + 1, nt, nc);
                        }
                    }
                };
                // three buffer sets: the prefetch goes out first.  Two (dual output): it targets the buffer the previous
                // chunk's store is still reading, so it goes out after the accumulator load and the functor, by which
                // time that read has long finished and lane 0 does not hold up the warp's aligned tcgen05.ld.
                if constexpr (kEpiBufs > 2) prefetch_next_residual();
                typename EpiCols<Epilogue>::type cv{};
                constexpr bool kHasCols = !std::is_same<typename EpiCols<Epilogue>::type, NoCols>::value;
                if constexpr (kHasCols) cv = epi.cols(n0, nvalid, lane);
                uint32_t r[32];
                tc::tmem_ld_32x32b_x32(t_row + c * 32, r);
                tc::tmem_ld_wait();
                if (c + kEpiPerQuad >= nch) {
                    // this warp's last tcgen05.ld of the tile has completed: hand its share of the accumulator buffer back
                    // to the MMA issuer now, before the arithmetic and the stores of the chunk
                    tc::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive_peer_tmem(as ? tempty_leader1 : tempty_leader0);
                }
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                if constexpr (!kStaged) {
                    if (m < M) epi(m, n0, nvalid, v);
                } else {
                    constexpr bool kHasRS = !std::is_same<typename EpiRowState<Epilogue>::type, NoRowState>::value;
                    if constexpr (kHasRS && kHasCols) epi.apply(m, n0, nvalid, v, rs, cv);
                    else if constexpr (kHasRS) epi.apply(m, n0, nvalid, v, rs);
                    else if constexpr (kHasCols) epi.apply(m, n0, nvalid, v, cv);
                    else epi.apply(m, n0, nvalid, v);
                    if constexpr (kEpiBufs <= 2) prefetch_next_residual();
                    const int b = seq % kEpiBufs;
                    uint8_t* buf = my_bufs_gen + b * kBufBytes;
                    if constexpr (kResidual) {
                        tc::mbar_wait(res_bar(ew, b), (rph >> b) & 1);
                        rph ^= 1u << b;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float4* p = reinterpret_cast<float4*>(buf + stage_off<128>(lane, j));
                            float4 x = *p;
                            x.x += v[4 * j + 0]; x.y += v[4 * j + 1]; x.z += v[4 * j + 2]; x.w += v[4 * j + 3];
                            *p = x;
                            if constexpr (kDual) { v[4 * j + 0] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w; }
                        }
                    } else {
                        if (lane == 0) tc::bulk_wait_group_read<kEpiBufs - 1>();   // buffer b's previous store has read it
                        __syncwarp();
                        if constexpr (kOutBytes == 4) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                *reinterpret_cast<float4*>(buf + stage_off<128>(lane, j)) =
                                    make_float4(v[4 * j + 0], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                uint4 u;
                                u.x = Epilogue::pack2(v[8 * j + 0], v[8 * j + 1]);
                                u.y = Epilogue::pack2(v[8 * j + 2], v[8 * j + 3]);
                                u.z = Epilogue::pack2(v[8 * j + 4], v[8 * j + 5]);
                                u.w = Epilogue::pack2(v[8 * j + 6], v[8 * j + 7]);
                                *reinterpret_cast<uint4*>(buf + stage_off<64>(lane, j)) = u;
                            }
                        }
                    }
                    if constexpr (kDual) {
                        // v now holds the final f32 values of the chunk: their 16-bit copy is the next GEMM's A operand, their
                        // sums feed its folded LayerNorm.  Columns past nvalid (a ragged last chunk) do not count.
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (j * 8 < nvalid) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    if (kEW == 4 && (c & 1)) { st_sum1 += v[8 * j + i]; st_sq1 = fmaf(v[8 * j + i], v[8 * j + i], st_sq1); }
                                    else { st_sum += v[8 * j + i]; st_sq = fmaf(v[8 * j + i], v[8 * j + i], st_sq); }
                                }
                            }
                            uint4 u;
                            u.x = Epilogue::pack2(v[8 * j + 0], v[8 * j + 1]);
                            u.y = Epilogue::pack2(v[8 * j + 2], v[8 * j + 3]);
                            u.z = Epilogue::pack2(v[8 * j + 4], v[8 * j + 5]);
                            u.w = Epilogue::pack2(v[8 * j + 6], v[8 * j + 7]);
                            *reinterpret_cast<uint4*>(buf + kMainBytes + stage_off<64>(lane, j)) = u;
                        }
                    }
                    tc::fence_proxy_async();          // generic-proxy smem writes -> visible to the TMA store
                    __syncwarp();
                    if (lane == 0) {
                        tc::tma_store_2d(&tmap_c, my_bufs + b * kBufBytes, n0, m0);
                        if constexpr (kDual) tc::tma_store_2d(&tmap_c2, my_bufs + b * kBufBytes + kMainBytes, n0, m0);
                        tc::bulk_commit_group();
                    }
                    ++seq;
                }
            }
            if constexpr (kDual) {
                // one (sum, sum of squares) partial per (row, column block, warp half), always written (zeros when this
                // warp owned no chunk of the tile) so that the consumer can add all of them in a fixed order
                if (m < M) {
                    epi.stats[(size_t)m * epi.n_part + n_blk * 2 + hf] = make_float2(st_sum, st_sq);
                    if constexpr (kEW == 4) epi.stats[(size_t)m * epi.n_part + n_blk * 2 + 1] = make_float2(st_sum1, st_sq1);
                }
            }
            if (hf >= nch) {                               // a warp that owned no chunk of this tile still owes its arrival
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive_peer_tmem(as ? tempty_leader1 : tempty_leader0);
            }
        }
        if (kStaged && lane == 0) tc::bulk_wait_group<0>();
    }

    __syncwarp();
    tc::fence_before_sync();
    tc::cluster_sync();                       // nobody leaves while the pair's MMAs / remote signals are in flight
    if (warp == 2) tc::tmem_dealloc_pair<kTmemCols>(tmem_base);
}

}  // namespace q3a
