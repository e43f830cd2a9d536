// Host-side TMA descriptor (CUtensorMap) construction.  cuTensorMapEncodeTiled is the only driver
// API entry point libq3a needs; it is resolved at run time through cudaGetDriverEntryPoint so the
// library has no link-time dependency on libcuda.so (it must load on a box without a driver).
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace q3a {

// 16-bit elements, 128-byte swizzle, OOB reads return zero.  dims/strides innermost first;
// strides_bytes has rank-1 entries (stride of dims 1..rank-1).
int make_tmap_16bit(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box);

// row-major [rows][cols] 16-bit matrix with leading dimension ld (elements); box = {64, box_rows}
int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);

// row-major [rows][cols] matrix of 2- or 4-byte elements, box = 32 x 32 elements, swizzle = the box's row
// bytes (64 B for 16-bit, 128 B for f32): the GEMM epilogue's staged stores and residual loads
int make_tmap_tile32(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld);

// row-major [rows][cols] int8 matrix (Q8_0 weight plane), box = {64 bytes, box_rows}, no swizzle: the dequant warps read it
int make_tmap_u8_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);

}  // namespace q3a
