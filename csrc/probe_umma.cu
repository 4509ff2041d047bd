// Hardware probe (not on any training path): does tcgen05.mma accept a 128-byte-swizzled shared-memory operand whose
// descriptor start address is shifted by whole 128-byte rows, i.e. is NOT aligned to the 1024-byte swizzle repeat?
// If it does, ONE halo tile of an activation in shared memory can serve all filter taps of a 3x3 convolution through
// row-shifted descriptors instead of nine TMA loads (csrc/conv_igemm.cu).  Two conventions are tried:
//   variant 0: descriptor built as everywhere else in this library (base-offset field 0);
//   variant 1: base-offset field (bits 49-51) = (start address >> 7) & 7, the rule the PTX ISA gives for unaligned starts.
// mn = 0: A is K-major (rows = M index: shifting rows shifts M);  mn = 1: A is MN-major (rows = K index).
// B is a 64 x 64 identity, so D reproduces the rows of X the MMA really read.  tests/test_gpu_probe.py records the result.
#include <cuda.h>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

namespace {
using namespace tc;

__global__ void __launch_bounds__(128, 1) probe_umma_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_b,
                                                            float* __restrict__ D, int shift, int variant, int mn) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sx = smem;               // 256 rows x 128 B
    uint8_t* sb = smem + 256 * 128;   // 64 rows x 128 B
    uint64_t* bar = reinterpret_cast<uint64_t*>(sb + 64 * 128);
    uint64_t* done = bar + 1;
    uint32_t* tptr = reinterpret_cast<uint32_t*>(done + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(done, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tptr, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tptr;
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, 256 * 128 + 64 * 128);
        tma_load_2d(sx, &tma_x, bar, 0, 0);
        tma_load_2d(sb, &tma_b, bar, 0, 0);
        mbar_wait(bar, 0);
        tc_fence_after();
        const uint32_t a0 = smem_u32(sx) + shift * 128, b0 = smem_u32(sb);
        const uint32_t idesc = make_idesc(128, 64, mn != 0, mn != 0);
        for (int k = 0; k < 4; ++k) {
            uint64_t da, db;
            if (mn) {
                da = make_smem_desc(a0 + k * 2048, 0, 1024);   // second 64-wide M block aliases the first (LBO 0)
                db = make_smem_desc(b0 + k * 2048, 0, 1024);
            } else {
                da = make_smem_desc(a0 + k * 32, 16, 1024);
                db = make_smem_desc(b0 + k * 32, 16, 1024);
            }
            if (variant == 1) da |= (uint64_t)(((a0 + (mn ? k * 2048 : 0)) >> 7) & 7) << 49;
            umma_bf16(tmem, da, db, idesc, k > 0 ? 1u : 0u);
        }
        umma_commit(done);
    }
    mbar_wait(done, 0);
    tc_fence_after();
    uint32_t r[32];
    for (int h = 0; h < 2; ++h) {
        tmem_ld_32x32(tmem + h * 32 + ((uint32_t)(warp * 32) << 16), r);
        tmem_ld_wait();
        for (int e = 0; e < 32; ++e) D[(warp * 32 + lane) * 64 + h * 32 + e] = __uint_as_float(r[e]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, 64);
    }
}
}  // namespace

// X: bf16 [256][64]; B: bf16 [64][64]; D: fp32 [128][64].
// mn = 0: D[i][j] = sum_k X[shift + i][k] B[j][k];  mn = 1: D[m][n] = sum_k X[shift + k][m] B[k][n] (rows m >= 64 repeat).
DDL_API int ddl_probe_umma_shift(const void* X, const void* B, float* D, int shift, int variant, int mn, void* stream_) {
    DDL_REQUIRE(X && B && D && shift >= 0 && shift <= 64, "probe: bad arguments");
    CUtensorMap tx, tb;
    if (int e = tma_host::encode_2d(&tx, X, 256, 64, 128, 256, "probe")) return e;
    if (int e = tma_host::encode_2d(&tb, B, 64, 64, 128, 64, "probe")) return e;
    const int smem = 256 * 128 + 64 * 128 + 64 + 1024;
    static bool attr = false;
    if (!attr) {
        DDL_CHECK(cudaFuncSetAttribute(probe_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    probe_umma_kernel<<<1, 128, smem, (cudaStream_t)stream_>>>(tx, tb, D, shift, variant, mn);
    DDL_LAUNCH_CHECK();
    return 0;
}
