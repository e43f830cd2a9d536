// Linear layers on the tcgen05 GEMM: instantiations and host dispatch.
#include "gemm.h"
#include "gemm_tc.cuh"
#include "epilogues.cuh"
#include "tmap.h"

#include <map>
#include <mutex>

namespace q3a {

int sm_count() { return sm_count_cached(); }

int gemm_prepare_impl(const void* kernel, int smem_bytes, int threads, int* max_pairs) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> cache;
    int dev = 0;
    Q3A_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find({kernel, dev});
    if (it != cache.end()) { *max_pairs = it->second; return Q3A_OK; }
    Q3A_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * sm_count());
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = 2; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
    cfg.attrs = &at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = sm_count() / 2;
    }
    if (n > sm_count() / 2) n = sm_count() / 2;
    cache[{kernel, dev}] = n;
    *max_pairs = n;
    return Q3A_OK;
}

// Q3A_NO_RES_PREFETCH: A/B switch for the residual epilogues' L2 prefetch of the next tile
static int res_prefetch() { return getenv("Q3A_NO_RES_PREFETCH") ? 0 : 1; }

int ln_partials(int N) {
    int bn = pick_block_n(N);
    if (bn == 240) bn = 128;
    return 2 * ceil_div(N, bn);
}

int pick_block_n(int N) {
    if (N % 256 == 0) return 256;
    if (N % 224 == 0) return 224;
    if (N % 240 == 0) return 240;
    if (N >= 1024) return 256;
    return 128;
}

template <int BLOCK_N, int kFmt, class Epi, int kEW>
static int launch_linear_ew(const LinearArgs& a, const Epi& epi, cudaStream_t stream) {
    using S = GemmSmemFor<BLOCK_N, Epi, false, kEW>;
    CUtensorMap ta, tb, tc_;
    Q3A_TRY(make_tmap_2d(&ta, a.a, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)a.lda, kBlockM));
    Q3A_TRY(make_tmap_2d(&tb, a.w, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)a.ldw, BLOCK_N / 2));
    Q3A_TRY(make_tmap_tile32(&tc_, a.out, Epi::kOutBytes, (uint64_t)a.M, (uint64_t)a.N, (uint64_t)a.ldc));
    CUtensorMap tc2 = tc_;
    if (EpiDual<Epi>::value) Q3A_TRY(make_tmap_tile32(&tc2, a.out16, 2, (uint64_t)a.M, (uint64_t)a.N, (uint64_t)a.ldc));
    auto kern = gemm_tc_kernel<BLOCK_N, kFmt, LoaderLinear, Epi, false, kEW>;
    constexpr int kThreads = 128 + 32 * kEW;
    int pairs = 0;
    Q3A_TRY(gemm_prepare_impl(reinterpret_cast<const void*>(kern), S::kTotal, kThreads, &pairs));
    LoaderLinear loader{ceil_div(a.K, kBlockK)};
    const int tiles = ceil_div(a.M, 2 * kBlockM) * ceil_div(a.N, BLOCK_N);
    const int grid = 2 * (tiles < pairs ? tiles : pairs);
    Q3A_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), S::kTotal, stream, ta, tb, tc_, tc2, a.M, a.N, loader, epi, Q8Scales{}));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

// Q3A_EPI_WARPS=8: A/B switch for the 12-warp epilogue of the 16-bit-output GEMMs
static int epi_warps_16bit() { const char* e = getenv("Q3A_EPI_WARPS"); return e && atoi(e) == 8 ? 8 : 12; }

template <int BLOCK_N, int kFmt, class Epi>
static int launch_linear(const LinearArgs& a, const Epi& epi, cudaStream_t stream) {
    if constexpr (Epi::kStaged && Epi::kOutBytes == 2) {
        if (epi_warps_16bit() == 12) return launch_linear_ew<BLOCK_N, kFmt, Epi, 12>(a, epi, stream);
    }
    if constexpr (EpiDual<Epi>::value && Epi::kResidual) {
        // FFN-down / out-projection with the 16-bit copy: K is long, the epilogue has time; four warps halve the staging
        // and buy a fifth ring stage for the operand stream the MMAs were waiting for (Q3A_DUAL_EPI_WARPS=4|8)
        const char* e = getenv("Q3A_DUAL_EPI_WARPS");        // read per launch: the parity test switches it
        if ((e ? atoi(e) : kDefaultDualEpiWarps) == 4) return launch_linear_ew<BLOCK_N, kFmt, Epi, 4>(a, epi, stream);
    }
    if constexpr (!EpiDual<Epi>::value && Epi::kResidual) {
        // the plain residual GEMM (out-projection at fold mask 5): same trade, its own switch (Q3A_RES_EPI_WARPS=4|8)
        const char* e = getenv("Q3A_RES_EPI_WARPS");
        if ((e ? atoi(e) : kDefaultResEpiWarps) == 4) return launch_linear_ew<BLOCK_N, kFmt, Epi, 4>(a, epi, stream);
    }
    return launch_linear_ew<BLOCK_N, kFmt, Epi, 8>(a, epi, stream);
}

// Q8_0 weights, dequantised inside the kernel (gemm_tc.cuh, kWq8): the W tensor map covers the int8 plane
template <int BLOCK_N, int kFmt, class Epi>
static int launch_linear_q8(const LinearArgs& a, const Epi& epi, cudaStream_t stream) {
    using S = GemmSmemFor<BLOCK_N, Epi, true>;
    CUtensorMap ta, tb, tc_;
    Q3A_TRY(make_tmap_2d(&ta, a.a, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)a.lda, kBlockM));
    Q3A_TRY(make_tmap_u8_2d(&tb, a.w, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)a.ldw, BLOCK_N / 2));
    Q3A_TRY(make_tmap_tile32(&tc_, a.out, Epi::kOutBytes, (uint64_t)a.M, (uint64_t)a.N, (uint64_t)a.ldc));
    auto kern = gemm_tc_kernel<BLOCK_N, kFmt, LoaderLinear, Epi, true>;
    int pairs = 0;
    Q3A_TRY(gemm_prepare_impl(reinterpret_cast<const void*>(kern), S::kTotal, kGemmThreads + 32 * kDequantWarps, &pairs));
    LoaderLinear loader{ceil_div(a.K, kBlockK)};
    const int tiles = ceil_div(a.M, 2 * kBlockM) * ceil_div(a.N, BLOCK_N);
    const int grid = 2 * (tiles < pairs ? tiles : pairs);
    Q8Scales q8{static_cast<const __half*>(a.w_d), a.K / 32};
    Q3A_CUDA(launch_pdl(kern, dim3(grid), dim3(kGemmThreads + 32 * kDequantWarps), S::kTotal, stream, ta, tb, tc_, tc_, a.M, a.N,
                        loader, epi, q8));
    ++g_launch_counter;
    Q3A_CUDA(cudaGetLastError());
    return Q3A_OK;
}

template <int BLOCK_N, int kFmt>
static int dispatch_epi_q8(EpiKind kind, const LinearArgs& a, cudaStream_t stream) {
    using OpT = typename std::conditional<kFmt == 1, __nv_bfloat16, __half>::type;
    switch (kind) {
        case EPI_BIAS_OP: return launch_linear_q8<BLOCK_N, kFmt>(a, EpiBiasActOp<OpT, false>{a.bias}, stream);
        case EPI_BIAS_GELU_OP: return launch_linear_q8<BLOCK_N, kFmt>(a, EpiBiasActOp<OpT, true>{a.bias}, stream);
        case EPI_BIAS_F32: return launch_linear_q8<BLOCK_N, kFmt>(a, EpiBiasF32<false>{a.bias}, stream);
        case EPI_BIAS_RES_F32:
            if (a.res != a.out) { set_error("gemm_linear: the residual must be the output matrix (updated in place)"); return Q3A_ERR_ARG; }
            return launch_linear_q8<BLOCK_N, kFmt>(a, EpiBiasF32<true>{a.bias, res_prefetch()}, stream);
        default: break;
    }
    set_error("gemm_linear: this epilogue has no fused-Q8_0 instantiation");
    return Q3A_ERR_ARG;
}

template <int BLOCK_N, int kFmt>
static int dispatch_epi(EpiKind kind, const LinearArgs& a, cudaStream_t stream) {
    if (a.w_d) return dispatch_epi_q8<BLOCK_N, kFmt>(kind, a, stream);
    using OpT = typename std::conditional<kFmt == 1, __nv_bfloat16, __half>::type;
    switch (kind) {
        case EPI_BIAS_OP:
            return launch_linear<BLOCK_N, kFmt>(a, EpiBiasActOp<OpT, false>{a.bias}, stream);
        case EPI_BIAS_GELU_OP:
            return launch_linear<BLOCK_N, kFmt>(a, EpiBiasActOp<OpT, true>{a.bias}, stream);
        case EPI_BIAS_F32:
            return launch_linear<BLOCK_N, kFmt>(a, EpiBiasF32<false>{a.bias}, stream);
        case EPI_BIAS_RES_F32:
            if (a.res != a.out) { set_error("gemm_linear: the residual must be the output matrix (updated in place)"); return Q3A_ERR_ARG; }
            return launch_linear<BLOCK_N, kFmt>(a, EpiBiasF32<true>{a.bias, res_prefetch()}, stream);
        case EPI_POSEMB_F32:
            return launch_linear<BLOCK_N, kFmt>(a, EpiPosEmbF32{a.pe, a.pos, a.ldc, a.M}, stream);
        case EPI_BIAS_RES_F32_DUAL:
        case EPI_POSEMB_F32_DUAL:
            if (!a.out16 || !a.stats || a.n_part != 2 * ceil_div(a.N, BLOCK_N)) {
                set_error("gemm_linear: dual output needs out16, stats and n_part = ln_partials(N)"); return Q3A_ERR_ARG;
            }
            if (kind == EPI_POSEMB_F32_DUAL)
                return launch_linear<BLOCK_N, kFmt>(a, EpiPosEmbF32Dual<OpT>{a.pe, a.pos, a.ldc, a.M, a.stats, a.n_part}, stream);
            if (a.res != a.out) { set_error("gemm_linear: the residual must be the output matrix (updated in place)"); return Q3A_ERR_ARG; }
            return launch_linear<BLOCK_N, kFmt>(a, EpiBiasF32Dual<OpT, true>{a.bias, a.stats, a.n_part, res_prefetch()}, stream);
        case EPI_LNFOLD_OP:
        case EPI_LNFOLD_GELU_OP:
            if (!a.bias || !a.colsum || !a.stats || a.n_part <= 0) { set_error("gemm_linear: LayerNorm fold needs bias', colsum and stats"); return Q3A_ERR_ARG; }
            if (kind == EPI_LNFOLD_OP)
                return launch_linear<BLOCK_N, kFmt>(a, EpiLnFoldActOp<OpT, false>{a.bias, a.colsum, a.stats, a.n_part, a.M, a.ln_inv_d, a.ln_eps, a.ln_rms}, stream);
            return launch_linear<BLOCK_N, kFmt>(a, EpiLnFoldActOp<OpT, true>{a.bias, a.colsum, a.stats, a.n_part, a.M, a.ln_inv_d, a.ln_eps, a.ln_rms}, stream);
        case EPI_LNFOLD_F32:
            if (!a.bias || !a.colsum || !a.stats || a.n_part <= 0) { set_error("gemm_linear: LayerNorm fold needs bias', colsum and stats"); return Q3A_ERR_ARG; }
            return launch_linear<BLOCK_N, kFmt>(a, EpiLnFoldF32{a.bias, a.colsum, a.stats, a.n_part, a.M, a.ln_inv_d, a.ln_eps, a.ln_rms}, stream);
    }
    set_error("gemm_linear: unknown epilogue");
    return Q3A_ERR_ARG;
}

template <int kFmt>
static int dispatch_bn(EpiKind kind, const LinearArgs& a, int bn, cudaStream_t stream) {
    switch (bn) {
        case 128: return dispatch_epi<128, kFmt>(kind, a, stream);
        case 224: return dispatch_epi<224, kFmt>(kind, a, stream);
        case 256: return dispatch_epi<256, kFmt>(kind, a, stream);
    }
    set_error("gemm_linear: unsupported block_n " + std::to_string(bn));
    return Q3A_ERR_ARG;
}

int gemm_linear(EpiKind kind, int fmt, const LinearArgs& a, cudaStream_t stream) {
    if (a.M <= 0) return Q3A_OK;
    if (a.K % 8 || a.N % 8 || a.lda % 8 || a.ldw % 8 || a.ldc % 8 || a.K <= 0 || a.N <= 0) {
        set_error("gemm_linear: K, N and leading dimensions must be positive multiples of 8");
        return Q3A_ERR_ARG;
    }
    if (a.w_d && (a.K % 64 || a.ldw != a.K)) { set_error("gemm_linear: fused Q8_0 weights need K % 64 == 0 and a dense int8 plane"); return Q3A_ERR_ARG; }
    int bn = a.block_n ? a.block_n : pick_block_n(a.N);
    if (bn == 240) bn = 128;
    return fmt == OP_F16 ? dispatch_bn<0>(kind, a, bn, stream) : dispatch_bn<1>(kind, a, bn, stream);
}

}  // namespace q3a

extern "C" int q3a_test_gemm(const void* a_dev, const void* w_dev, const float* bias_dev, float* d_dev, int M, int N,
                             int K, int operand_dtype, int block_n, void* stream) {
    using namespace q3a;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return Q3A_ERR_CUDA; }
    Q3A_TRY(use_device(dev));
    LinearArgs a;
    a.a = a_dev; a.M = M; a.K = K; a.lda = K;
    a.w = w_dev; a.N = N; a.ldw = K;
    a.out = d_dev; a.ldc = N; a.bias = bias_dev; a.block_n = block_n;
    return gemm_linear(EPI_BIAS_F32, operand_dtype, a, reinterpret_cast<cudaStream_t>(stream));
}
