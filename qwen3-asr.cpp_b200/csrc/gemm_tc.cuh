// Persistent warp-specialised tcgen05 GEMM for sm_100a:  D[M,N] = A[M,K] * W[N,K]^T, fp32 accumulate.
//
//   warp 0      : TMA producer (one lane) - A and W tiles into a kStages-deep smem ring
//   warp 1      : MMA issuer  (one lane) - tcgen05.mma, accumulators in TMEM (2 buffers of 256 columns)
//   warp 2      : TMEM allocator / deallocator
//   warp 3      : idle
//   warps 4..7  : epilogue - tcgen05.ld (one accumulator row per thread), fused functor, global store
//
// Both operands are K-major 16-bit (bf16 or f16) staged with 128-byte swizzle; a tile is
// 128 x BLOCK_N with K blocks of 64.  The A-side coordinates come from a "loader" policy so the
// same kernel runs the Linear layers (2-D tensor map) and the stride-2 3x3 convolutions of the
// stem as implicit GEMM (4-D tensor map over parity-split channels-last planes, see stem.cuh).
#pragma once
#include "common.cuh"
#include "tc.cuh"

namespace q3a {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                 // 64 x 16-bit = 128 B = one swizzle span
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 256;
constexpr int kTmemCols = 512;              // two accumulator buffers of up to 256 columns

template <int BLOCK_N> struct GemmSmem {
    static constexpr int kABytes = kBlockM * kBlockK * 2;          // 16 KB
    static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = (200 * 1024) / kStageBytes > 6 ? 6 : (200 * 1024) / kStageBytes;
    static constexpr int kBarrierBytes = 256;
    static constexpr int kTotal = kStages * kStageBytes + kBarrierBytes + 1024;   // +1024: manual alignment
    static_assert(kStageBytes % 1024 == 0, "stage must keep 1024-B swizzle alignment");
    static_assert(kStages >= 2, "need at least double buffering");
};

// ---- A-operand coordinate policies --------------------------------------------------------
// Linear: A is [M][K] row-major, tensor map dims {K, M}, box {64, 128}.
struct LoaderLinear {
    int num_k_blocks;
    __device__ __forceinline__ void load_a(uint32_t dst, const void* tmap, uint32_t bar, int m_blk, int kb) const {
        tc::tma_load_2d(dst, tmap, bar, kb * kBlockK, m_blk * kBlockM);
    }
    __device__ __forceinline__ int k_steps(int) const { return kBlockK / kUmmaK; }
};
// Conv (3x3, stride 2, pad 1) over parity-split planes (layout in stem.cuh): tensor map dims
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
    __device__ __forceinline__ void load_a(uint32_t dst, const void* tmap, uint32_t bar, int m_blk, int kb) const {
        const int tap = kb / c_blocks, cb = kb - tap * c_blocks;
        const int dt = tap / 3, df = tap - dt * 3;
        const int plane = ((dt != 1) ? 2 : 0) + ((df != 1) ? 1 : 0);
        tc::tma_load_4d(dst, tmap, bar, cb * kBlockK, (df == 0) ? -1 : 0,
                        m_blk * rows_per_tile + ((dt == 0) ? -1 : 0), plane);
    }
    __device__ __forceinline__ int k_steps(int kb) const {
        return ((kb % c_blocks) == c_blocks - 1) ? last_block_steps : kBlockK / kUmmaK;
    }
};

// ---- the kernel ------------------------------------------------------------------------------
// Epilogue contract:  struct E { __device__ void operator()(int m, int n0, int nvalid, const float (&acc)[32]) const; }
// is called once per (row m < M, 32-column group starting at n0 < N); only the first `nvalid`
// (a multiple of 8, <= 32) columns of the group belong to the output.
template <int BLOCK_N, int kFmt, class Loader, class Epilogue>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               int M, int N, Loader loader, Epilogue epi) {
    using S = GemmSmem<BLOCK_N>;
    constexpr int kStages = S::kStages;
    static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "BLOCK_N must be a multiple of 16, at most 256");

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * S::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - tc::smem_u32(smem_raw));
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * S::kStageBytes +
                                                                            8 * (2 * kStages + 4));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_blocks = ceil_div(M, kBlockM), n_blocks = ceil_div(N, BLOCK_N);
    const int num_tiles = m_blocks * n_blocks;
    const int num_kb = loader.num_k_blocks;

    if (warp == 0 && lane == 0) {
        tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { tc::mbar_init(full_bar(s), 1); tc::mbar_init(empty_bar(s), 1); }
        for (int s = 0; s < 2; ++s) { tc::mbar_init(tfull_bar(s), 1); tc::mbar_init(tempty_bar(s), 4); }
        tc::fence_barrier_init();
    }
    if (warp == 2) tc::tmem_alloc<kTmemCols>(tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                // n fastest: CTAs running side by side share the A tile through L2
                const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    tc::mbar_wait(empty_bar(s), ph ^ 1);
                    const uint32_t a_dst = smem_base + s * S::kStageBytes;
                    const uint32_t b_dst = a_dst + S::kABytes;
                    tc::mbar_arrive_expect_tx(full_bar(s), S::kStageBytes);
                    loader.load_a(a_dst, &tmap_a, full_bar(s), m_blk, kb);
                    tc::tma_load_2d(b_dst, &tmap_b, full_bar(s), kb * kBlockK, n_blk * BLOCK_N);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = tc::make_idesc(kFmt, kBlockM, BLOCK_N, 0, 0);
        uint32_t it = 0, local = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
            const int as = local & 1;
            const uint32_t aph = (local >> 1) & 1;
            tc::mbar_wait(tempty_bar(as), aph ^ 1);
            tc::fence_after_sync();
            const uint32_t d_tmem = tmem_base + as * 256;
            for (int kb = 0; kb < num_kb; ++kb, ++it) {
                const int s = it % kStages;
                const uint32_t ph = (it / kStages) & 1;
                tc::mbar_wait(full_bar(s), ph);
                tc::fence_after_sync();
                if (lane == 0) {
                    const uint32_t a_addr = smem_base + s * S::kStageBytes;
                    const uint32_t b_addr = a_addr + S::kABytes;
                    const uint64_t a_desc = tc::smem_desc_kmajor_sw128(a_addr);
                    const uint64_t b_desc = tc::smem_desc_kmajor_sw128(b_addr);
                    const int steps = loader.k_steps(kb);
#pragma unroll
                    for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                        // advance 16 elements (32 B) along K inside the swizzle span: +2 in (addr>>4) units
                        if (k < steps) tc::mma_f16_ss(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0);
                    }
                    tc::mma_commit(empty_bar(s));                      // frees the smem slot when the MMAs retire
                    if (kb == num_kb - 1) tc::mma_commit(tfull_bar(as));   // accumulator complete
                }
                __syncwarp();
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;                       // TMEM lane quadrant this warp may read
        uint32_t local = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
            const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
            const int as = local & 1;
            const uint32_t aph = (local >> 1) & 1;
            tc::mbar_wait(tfull_bar(as), aph);
            tc::fence_after_sync();
            const int m = m_blk * kBlockM + q * 32 + lane;
            const uint32_t t_row = tmem_base + as * 256 + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BLOCK_N; c += 32) {
                uint32_t r[32];
                tc::tmem_ld_32x32b_x32(t_row + c, r);
                tc::tmem_ld_wait();
                const int n0 = n_blk * BLOCK_N + c;
                if (m < M && n0 < N) {
                    float acc[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc[j] = __uint_as_float(r[j]);
                    int nvalid = min(32, min(BLOCK_N - c, N - n0));
                    epi(m, n0, nvalid, acc);
                }
            }
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tempty_bar(as));
        }
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 2) tc::tmem_dealloc<kTmemCols>(tmem_base);
}

}  // namespace q3a
