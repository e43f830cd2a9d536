#include "tmap.h"

#include <mutex>

namespace q3a {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_once;

static void resolve() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
        g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    else
        cudaGetLastError();
}

int make_tmap_16bit(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box) {
    std::call_once(g_once, resolve);
    if (!g_encode) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return Q3A_ERR_CUDA; }
    cuuint64_t d[5]; cuuint64_t s[4]; cuuint32_t b[5]; cuuint32_t es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        std::string msg = "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r) + " (rank " +
                          std::to_string(rank) + ", dims";
        for (int i = 0; i < rank; ++i) msg += " " + std::to_string(dims[i]);
        msg += ", box";
        for (int i = 0; i < rank; ++i) msg += " " + std::to_string(box[i]);
        msg += ")";
        set_error(msg);
        return Q3A_ERR_CUDA;
    }
    return Q3A_OK;
}

int make_tmap_tile32(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld) {
    std::call_once(g_once, resolve);
    if (!g_encode) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return Q3A_ERR_CUDA; }
    if (elem_bytes != 2 && elem_bytes != 4) { set_error("make_tmap_tile32: element size must be 2 or 4"); return Q3A_ERR_ARG; }
    cuuint64_t d[2] = {cols, rows};
    cuuint64_t s[1] = {ld * (uint64_t)elem_bytes};
    cuuint32_t b[2] = {32, 32};
    cuuint32_t es[2] = {1, 1};
    CUresult r = g_encode(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                          const_cast<void*>(base), d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          elem_bytes == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (32x32 tile map) failed with code " + std::to_string((int)r) + " rows " +
                  std::to_string(rows) + " cols " + std::to_string(cols) + " ld " + std::to_string(ld));
        return Q3A_ERR_CUDA;
    }
    return Q3A_OK;
}

int make_tmap_u8_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    std::call_once(g_once, resolve);
    if (!g_encode) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return Q3A_ERR_CUDA; }
    cuuint64_t d[2] = {cols, rows};
    cuuint64_t s[1] = {ld};
    cuuint32_t b[2] = {64, box_rows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (int8 plane) failed with code " + std::to_string((int)r) + " rows " + std::to_string(rows) +
                  " cols " + std::to_string(cols));
        return Q3A_ERR_CUDA;
    }
    return Q3A_OK;
}

int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    uint64_t dims[2] = {cols, rows};
    uint64_t strides[1] = {ld * 2};
    uint32_t box[2] = {64, box_rows};
    return make_tmap_16bit(out, base, 2, dims, strides, box);
}

}  // namespace q3a
