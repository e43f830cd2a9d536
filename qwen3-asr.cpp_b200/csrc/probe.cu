// Development probe: how many cycles one tcgen05.mma (cta_group::1, M = 128, K = 16, 16-bit operands) really occupies
// the tensor pipe, as a function of N, of where A comes from (shared memory or TMEM) and of whether consecutive MMAs
// accumulate into the same TMEM columns.  The attention kernels issue many small MMAs (N = 64 for P.V at head_dim 64);
// ncu showed the pipe 79 % occupied at 21 % of its math rate, and this measures why.  Not on any product path.
#include "common.cuh"
#include "tc.cuh"

namespace q3a {

__global__ void __launch_bounds__(128, 1)
mma_rate_kernel(int N, int a_from_tmem, int n_mma, int alternate, int k_steps_per_commit, long long* __restrict__ cycles) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar = base + 16384 + 32768, slot = bar + 16;
    volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (base - tc::smem_u32(smem_raw)) + 16384 + 32768 + 16);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128)                       // finite operands
        reinterpret_cast<uint32_t*>(smem_raw + (base - tc::smem_u32(smem_raw)))[i] = 0x3C003C00u;
    if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
    if (warp == 0) tc::tmem_alloc<512>(slot);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot_ptr;
    const int M = (a_from_tmem & 2) ? 64 : 128;           // bit 1 of the operand-source argument: the M = 64 shape
    a_from_tmem &= 1;
    if (warp == 0) {
        const uint32_t idesc = tc::make_idesc(1, M, N, 0, 0);
        const uint64_t a = tc::smem_desc_kmajor_sw128(base), b = tc::smem_desc_kmajor_sw128(base + 16384);
        uint32_t ph = 0;
        long long t0 = 0, t1 = 0;
        for (int rep = 0; rep < 2; ++rep) {                                            // rep 0 warms up
            t0 = clock64();
            for (int i = 0; i < n_mma; i += k_steps_per_commit) {
                if (tc::elect_one()) {
                    for (int k = 0; k < k_steps_per_commit; ++k) {
                        const uint32_t d = tmem + (alternate ? (((i + k) & 1) * 256) : 0);
                        if (a_from_tmem) tc::mma_f16_ts(d, tmem + 448 + 8 * (k & 3), b + 2 * (k & 3), idesc, 1);
                        else tc::mma_f16_ss(d, a + 2 * (k & 3), b + 2 * (k & 3), idesc, 1);
                    }
                }
                __syncwarp();
            }
            if (tc::elect_one()) tc::mma_commit(bar);
            __syncwarp();
            tc::mbar_wait(bar, ph);
            ph ^= 1;
            t1 = clock64();
        }
        if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc<512>(tmem);
}

}  // namespace q3a

// cycles_out: device array of `grid` entries (cycles of the timed repetition per CTA)
extern "C" int q3a_probe_mma_rate(int N, int a_from_tmem, int n_mma, int alternate, int k_steps_per_commit, int grid,
                                  long long* cycles_out_dev, void* stream) {
    using namespace q3a;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return Q3A_ERR_CUDA; }
    Q3A_TRY(use_device(dev));
    if (a_from_tmem < 0 || a_from_tmem > 3 || ((a_from_tmem & 2) && N % 8) || N < 16 || N > 256 || (!(a_from_tmem & 2) && N % 16) || n_mma <= 0 || grid <= 0 || k_steps_per_commit <= 0 || !cycles_out_dev || (alternate && N > 192)) {
        set_error("probe_mma_rate: bad argument"); return Q3A_ERR_ARG;
    }
    const int smem = 16384 + 32768 + 64 + 1024;
    Q3A_TRY(kernel_smem_optin(reinterpret_cast<const void*>(mma_rate_kernel), smem));
    mma_rate_kernel<<<grid, 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(N, a_from_tmem, n_mma, alternate, k_steps_per_commit, cycles_out_dev);
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}
