// Host-side helper shared by every kernel that feeds tcgen05 through TMA: resolves cuTensorMapEncodeTiled from the
// driver once (no link-time dependency on libcuda) and encodes the one tensor-map shape this library uses -
// a 2-D bf16 tensor [rows x inner] with an arbitrary row pitch, boxes of [box_rows x 64 elements] and the 128-byte
// swizzle that the UMMA shared-memory descriptors in tcgen05.cuh expect.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

#include "common.cuh"

namespace tma_host {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encoder() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

// rows x inner bf16 elements, row pitch `pitch_bytes` (a multiple of 16; may be SMALLER than the row length: overlapping
// rows, see stem_conv.cu), box = [box_rows x 64 elements].  `what` names the caller in error messages.
inline int encode_2d(CUtensorMap* map, const void* base, long long rows, long long inner, long long pitch_bytes, int box_rows,
                     const char* what) {
    EncodeTiledFn enc = encoder();
    DDL_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is not available from the driver", what);
    DDL_REQUIRE(((uintptr_t)base & 15) == 0, "%s: operand pointer must be 16-byte aligned", what);
    DDL_REQUIRE(pitch_bytes % 16 == 0 && pitch_bytes > 0, "%s: operand pitch (%lld bytes) must be a multiple of 16", what, pitch_bytes);
    DDL_REQUIRE(rows >= 1 && inner >= 1 && box_rows >= 1 && box_rows <= 256, "%s: bad tensor-map geometry (rows %lld, inner %lld, box %d)",
                what, rows, inner, box_rows);
    cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)pitch_bytes};
    cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DDL_REQUIRE(r == CUDA_SUCCESS, "%s: cuTensorMapEncodeTiled failed with CUresult %d (rows %lld inner %lld pitch %lld box %d)", what, (int)r,
                rows, inner, pitch_bytes, box_rows);
    return 0;
}

}  // namespace tma_host
