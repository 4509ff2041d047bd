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
// Issue rate of tcgen05.mma (cta_group::1, M = 128, K = 16, both operands in shared memory) as a function of N: one CTA
// issues `iters` back-to-back MMAs on resident tiles and reports the SM cycles until the last one has retired.
__global__ void __launch_bounds__(128, 1) probe_umma_rate_kernel(long long* __restrict__ out, int n, int iters, int kmaj_steps, int mode, int nacc) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* done = reinterpret_cast<uint64_t*>(smem + 128 * 128 + 256 * 128);
    uint32_t* tptr = reinterpret_cast<uint32_t*>(done + 1);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (128 * 128 + 256 * 128) / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (threadIdx.x == 0) {
        mbar_init(done, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tptr;
    if (mode == 0) {
        // issue from one thread inside a divergent region (how the kernels of this library did it at first)
        if (threadIdx.x == 0) {
            const uint32_t a0 = smem_u32(smem), b0 = a0 + 128 * 128;
            const uint32_t idesc = make_idesc(128, n, false, false);
            const long long t0 = clock64();
            for (int i = 0; i < iters; i += 4) {
                const uint32_t d = tmem + ((i >> 2) & (nacc - 1)) * n;  // nacc is a power of two
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(d, make_smem_desc(a0 + k * 32, 16, 1024), make_smem_desc(b0 + k * 32, 16, 1024), idesc, 1u);
            }
            const long long t1 = clock64();
            umma_commit(done);
            mbar_wait(done, 0);
            const long long t2 = clock64();
            out[blockIdx.x * 2] = t1 - t0;
            out[blockIdx.x * 2 + 1] = t2 - t0;
        }
    } else if (mode == 1) {
      if (warp == 0) {
        // the whole warp runs the loop (operands stay warp-uniform); one elected lane issues each instruction
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 128 * 128;
        const uint32_t idesc = make_idesc(128, n, false, false);
        const long long t0 = clock64();
        for (int i = 0; i < iters; i += 4) {
            const uint32_t d = tmem + ((i >> 2) & (nacc - 1)) * n;
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(d, make_smem_desc(a0 + k * 32, 16, 1024), make_smem_desc(b0 + k * 32, 16, 1024), idesc, 1u);
            }
            __syncwarp();
        }
        const long long t1 = clock64();
        if (elect_one()) umma_commit(done);
        __syncwarp();
        mbar_wait(done, 0);
        const long long t2 = clock64();
        if (threadIdx.x == 0) {
            out[blockIdx.x * 2] = t1 - t0;
            out[blockIdx.x * 2 + 1] = t2 - t0;
        }
      }
    } else if (threadIdx.x == 0) {
        // the k-block loop of the GEMM / convolution kernels, piece by piece (one thread, groups of 4 MMAs):
        //   mode 3: + tcgen05.commit to a ring of 8 barriers after every group
        //   mode 4: + tcgen05.fence::after_thread_sync in front of every group
        //   mode 5: + wait (in front of every group) for the commit issued 6 groups earlier (stage reuse)
        //   mode 6: like 5, but the wait is for the commit of the PREVIOUS group (no MMA overlap across groups)
        uint64_t* ring = done + 2;
        for (int r = 0; r < 8; ++r) mbar_init(&ring[r], 1);
        fence_barrier_init();
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 128 * 128;
        const uint32_t idesc = make_idesc(128, n, false, false);
        const int lag = mode == 6 ? 1 : 6;
        const long long t0 = clock64();
        int g = 0;
        for (int i = 0; i < iters; i += 4, ++g) {
            if (mode >= 5 && g >= lag) mbar_wait(&ring[(g - lag) & 7], (uint32_t)(((g - lag) >> 3) & 1));
            if (mode >= 4) tc_fence_after();
            const uint32_t d = tmem + (g & (nacc - 1)) * n;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(d, make_smem_desc(a0 + k * 32, 16, 1024), make_smem_desc(b0 + k * 32, 16, 1024), idesc, 1u);
            umma_commit(&ring[g & 7]);
        }
        const long long t1 = clock64();
        umma_commit(done);
        mbar_wait(done, 0);
        const long long t2 = clock64();
        out[blockIdx.x * 2] = t1 - t0;
        out[blockIdx.x * 2 + 1] = t2 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
}
// How fast can an epilogue drain TMEM, and does a running mainloop slow it down?  `nwarps` warps (4 per TMEM lane quarter) read
// 32 lanes x 32 columns (4 KB per tcgen05.ld) `iters` times each from columns [256, 512) while - if mma_on - one warp keeps
// issuing 128 x 256 x 16 MMAs into columns [0, 256) until the readers are done.  out[cta] = {cycles of the read phase, MMAs
// issued meanwhile}.
__global__ void __launch_bounds__(64 + 16 * 32, 1) probe_tmem_drain_kernel(long long* __restrict__ out, int nwarps, int iters, int mma_on) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* done = reinterpret_cast<uint64_t*>(smem + 128 * 128 + 256 * 128);
    uint32_t* tptr = reinterpret_cast<uint32_t*>(done + 1);
    volatile int* flag = reinterpret_cast<volatile int*>(tptr + 1);  // readers still running
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (128 * 128 + 256 * 128) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (threadIdx.x == 0) {
        mbar_init(done, 1);
        *flag = nwarps;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tptr;
    long long mmas = 0;
    if (warp == 0) {
        if (mma_on) {
            const uint32_t idesc = make_idesc(128, 256, false, false);
            const uint64_t da = make_smem_desc(smem_u32(smem), 16, 1024), db = make_smem_desc(smem_u32(smem) + 128 * 128, 16, 1024);
            while (*flag > 0) {
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem, da + k * 2, db + k * 2, idesc, 1u);
                }
                __syncwarp();
                mmas += 4;
            }
            if (elect_one()) umma_commit(done);
            __syncwarp();
            mbar_wait(done, 0);
        }
    } else if (warp >= 2 && warp < 2 + nwarps) {
        const int quarter = warp & 3;
        uint32_t r[32];
        uint32_t sink = 0;
        __syncwarp();
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            tmem_ld_32x32(tmem + 256 + (uint32_t)((i & 7) * 32) + ((uint32_t)(quarter * 32) << 16), r);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) sink ^= r[e];
        }
        const long long t1 = clock64();
        if (lane == 0) {
            atomicSub(const_cast<int*>(flag), 1);
            if (warp == 2) out[blockIdx.x * 2] = t1 - t0;
            if (sink == 0x12345678u) out[blockIdx.x * 2] = 0;  // keep the loads alive
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0 && lane == 0) out[blockIdx.x * 2 + 1] = mmas;
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, 512);
    }
}
}  // namespace

// out: int64 [grid][2] = {issue cycles, completion cycles} of `iters` MMAs 128 x n x 16 per CTA
DDL_API int ddl_probe_umma_rate(long long* out, int n, int iters, int grid, int mode, int nacc, void* stream_) {
    DDL_REQUIRE(nacc >= 1 && nacc * n <= 512, "probe rate: nacc * n must fit in 512 TMEM columns");
    DDL_REQUIRE(out && n >= 16 && n <= 256 && n % 16 == 0 && iters >= 1 && grid >= 1, "probe rate: bad arguments");
    const int smem = 128 * 128 + 256 * 128 + 256 + 1024;
    static bool attr = false;
    if (!attr) {
        DDL_CHECK(cudaFuncSetAttribute(probe_umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    probe_umma_rate_kernel<<<grid, 128, smem, (cudaStream_t)stream_>>>(out, n, iters, 4, mode, nacc);
    DDL_LAUNCH_CHECK();
    return 0;
}

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

// out: int64 [grid][2] = {cycles the first reader warp needed for `iters` tcgen05.ld of 4 KB each, MMAs issued meanwhile}
DDL_API int ddl_probe_tmem_drain(long long* out, int nwarps, int iters, int mma_on, int grid, void* stream_) {
    DDL_REQUIRE(out && nwarps >= 1 && nwarps <= 16 && iters >= 1 && grid >= 1, "probe drain: bad arguments");
    const int smem = 128 * 128 + 256 * 128 + 256 + 1024;
    static bool attr = false;
    if (!attr) {
        DDL_CHECK(cudaFuncSetAttribute(probe_tmem_drain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr = true;
    }
    probe_tmem_drain_kernel<<<grid, 64 + 16 * 32, smem, (cudaStream_t)stream_>>>(out, nwarps, iters, mma_on);
    DDL_LAUNCH_CHECK();
    return 0;
}
