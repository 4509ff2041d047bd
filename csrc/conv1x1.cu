// 1x1 / stride-1 convolution forward on NHWC bf16 = skinny GEMM  y[M, N] = x[M, K] . w[N, K]^T  with M = batch * H * W
// in the hundreds of thousands and K, N <= 512 (ResNet-50 layer1 / layer2 bottleneck convolutions, SURVEY.md section
// 2.E.2; reference model at notebooks/cv/onnx_experiments.py:19), with the BatchNorm statistics of y produced by the
// epilogue (K9: "conv + BN" without a statistics pass).
//
// These layers are pure streaming problems (0.1 - 0.4 GB in + out against 26 GFLOP): the general 256 x 256-tile GEMM
// of gemm_tcgen05.cu is the wrong shape for them (one k-block per tile, N = 64 fills a quarter of a tile).  This
// kernel is weight-stationary instead:
//   * the whole weight matrix of a CTA's column block (min(N, 256) x K, <= 128 KB) is TMA-loaded into shared memory
//     once; N = 512 is split over blockIdx.y (x is then read twice, the second time from L2);
//   * persistent CTAs walk 128-row tiles of x; warp 0 streams [128 x 64] k-blocks of x through a TMA ring, warp 1
//     issues tcgen05.mma (M = 128, N = 64..256) into 512 / N TMEM accumulator buffers (<= 4), 16 epilogue warps
//     (4 TMEM lane quarters x 4 column groups of 64) drain them;
//   * epilogue per 32 x 64 chunk: tcgen05.ld -> bf16 -> XOR-swizzled staging tile -> (a) per-channel sum / sum of
//     squares accumulated in registers over all tiles of the CTA (lane = channel pair), (b) coalesced 512-byte
//     stores.  Per-CTA partial statistics are written once at the end and folded in fixed order by the BatchNorm
//     finalize kernel (ddl_bn_act_fwd_train_from_partials): deterministic, no atomics.
// Data gradient and weight gradient of these layers stay with the library (benchmarks/conv_bench.py: already at
// 1.05 - 1.4x of their floors).
#include <cuda.h>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

namespace {
using namespace tc;

constexpr int kTileM = 128;
constexpr int kEpiWarps = 16;      // 4 TMEM lane quarters x 4 column groups (one 64-column chunk each)
constexpr int kThreads = 64 + kEpiWarps * 32;  // TMA warp, MMA warp, epilogue warps
constexpr int kMaxStages = 8;
constexpr int kStageBytes = kTileM * 128;  // [128 rows x 64 bf16]

struct C1Params {
    bf16* y;
    float* partials;  // [gridDim.x][2][N]
    int M, N, K;
    int n_cta;        // output columns per CTA = min(N, 256); blockIdx.y selects the column block
    int num_tiles, kblocks, stages;
    int nbuf;         // accumulator buffers of n_cta columns (<= 4)
    int groups;       // n_cta / 64 active column groups
};

// B_MN = false: forward, w is [N][K] (K-major rows).  B_MN = true: data gradient, the same weight tensor is read as
// [K][N] (MN-major): y = dx, x = dy, N = conv input channels, K = conv output channels; no statistics (partials null).
template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1) conv1x1_fwd_kernel(const __grid_constant__ CUtensorMap tma_x,
                                                                   const __grid_constant__ CUtensorMap tma_w, const C1Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int w_bytes = p.n_cta * p.K * 2;
    uint8_t* w_smem = smem;                       // kblocks x [n_cta rows x 128 B]
    uint8_t* a_smem = smem + w_bytes;             // stages x 16 KB
    uint8_t* stg_base = a_smem + p.stages * kStageBytes;  // (4 * groups) x 4 KB
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(stg_base + p.groups * 4 * 4096);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* tmem_full = empty_bar + kMaxStages;  // [4]
    uint64_t* tmem_empty = tmem_full + 4;          // [4]
    uint64_t* w_bar = tmem_empty + 4;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 512;
    const int col_block = blockIdx.y * p.n_cta;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_x);
        tma_prefetch_desc(&tma_w);
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 4; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 4 * p.groups);
        }
        mbar_init(w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================== TMA producer (warp-uniform loop, one elected lane issues) =====================
        if (elect_one()) {
            mbar_expect_tx(w_bar, (uint32_t)w_bytes);
            for (int kb = 0; kb < p.kblocks; ++kb) {
                if (!B_MN) {
                    tma_load_2d(w_smem + (size_t)kb * p.n_cta * 128, &tma_w, w_bar, kb * 64, col_block);
                } else {
                    for (int j = 0; j < p.groups; ++j)  // [64 k-rows x 64 n] boxes, 64-wide n blocks 8 KB apart
                        tma_load_2d(w_smem + (size_t)kb * p.n_cta * 128 + j * 8192, &tma_w, w_bar, col_block + j * 64, kb * 64);
                }
            }
        }
        __syncwarp();
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            for (int kb = 0; kb < p.kblocks; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[stage], kStageBytes);
                    tma_load_2d(a_smem + stage * kStageBytes, &tma_x, &full_bar[stage], kb * 64, tile * kTileM);  // OOB rows: zeros
                }
                __syncwarp();
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue: the whole warp runs the loop, one elected lane issues =====================
        // (descriptors = constant | address >> 4, a k-step only adds to the address field: csrc/conv_igemm.cu)
        const uint32_t idesc = make_idesc(kTileM, p.n_cta, false, B_MN);
        constexpr uint32_t kStepA = 32 >> 4, kStepB = (B_MN ? 2048 : 32) >> 4;
        const uint64_t da0 = make_smem_desc(0, 16, 1024) | (uint64_t)(smem_u32(a_smem) >> 4);
        const uint64_t db0 = (B_MN ? make_smem_desc(0, 8192, 1024) : make_smem_desc(0, 16, 1024)) | (uint64_t)(smem_u32(w_smem) >> 4);
        const uint32_t w_kb16 = (uint32_t)(p.n_cta * 128) >> 4;
        const int kblocks = p.kblocks, stages = p.stages, nbuf = p.nbuf;
        mbar_wait(w_bar, 0);
        tc_fence_after();
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * p.n_cta;
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t da = da0 + (uint32_t)stage * (uint32_t)(kStageBytes >> 4), db = db0 + (uint32_t)kb * w_kb16;
                if (elect_one()) {
                    umma_bf16(d_tmem, da, db, idesc, kb > 0 ? 1u : 0u);
#pragma unroll
                    for (int k = 1; k < 4; ++k) umma_bf16(d_tmem, da + k * kStepA, db + k * kStepB, idesc, 1u);
                    umma_commit(&empty_bar[stage]);
                    if (kb == kblocks - 1) umma_commit(&tmem_full[acc]);
                }
                __syncwarp();
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            if (++acc == nbuf) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;
        const int quarter = warp & 3;  // TMEM lanes 32*quarter .. +31
        const int group = ew >> 2;     // 64-column chunk of the accumulator
        if (group < p.groups) {
            uint8_t* stg = stg_base + ew * 4096;
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;  // channels group*64 + 2*lane, +1
            int acc = 0;
            uint32_t acc_phase = 0;
            const int col0 = col_block + group * 64;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int row_base = tile * kTileM + quarter * 32;
                const bool valid = row_base + lane < p.M;
                mbar_wait(&tmem_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + acc * p.n_cta + group * 64 + ((uint32_t)(quarter * 32) << 16);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + h * 32, r);
                    tmem_ld_wait();
                    if (h == 1) {  // accumulator chunk fully read: hand it back before the slow part
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float f[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) f[e] = valid ? __uint_as_float(r[j * 8 + e]) : 0.f;
                        *reinterpret_cast<uint4*>(stg + lane * 128 + (((h * 4 + j) ^ (lane & 7)) * 16)) = float_to_bf16x8(f);
                    }
                }
                if (++acc == p.nbuf) { acc = 0; acc_phase ^= 1; }
                __syncwarp();
                float a_s0 = 0.f, a_s1 = 0.f, a_q0 = 0.f, a_q1 = 0.f;
                if (!B_MN) {
#pragma unroll 8
                for (int r = 0; r < 32; ++r) {
                    const uint32_t wv = *reinterpret_cast<const uint32_t*>(stg + r * 128 + (((lane >> 2) ^ (r & 7)) * 16) + (lane & 3) * 4);
                    const float v0 = __uint_as_float(wv << 16), v1 = __uint_as_float(wv & 0xffff0000u);
                    a_s0 += v0;
                    a_q0 += v0 * v0;
                    a_s1 += v1;
                    a_q1 += v1 * v1;
                }
                }
                s0 += a_s0;
                s1 += a_s1;
                q0 += a_q0;
                q1 += a_q1;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = i * 4 + (lane >> 3), chunk = lane & 7;
                    const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((chunk ^ (rr & 7)) * 16));
                    if (row_base + rr < p.M) st_global_v4(p.y + (size_t)(row_base + rr) * p.N + col0 + chunk * 8, v);
                }
                __syncwarp();
            }
            if (B_MN) goto done;  // data gradient: no statistics
            // fold the four lane-quarter warps of this column group in fixed order -> per-CTA partials
            float* mine = reinterpret_cast<float*>(stg);  // own staging area, free now
            __syncwarp();
            mine[lane * 4 + 0] = s0;
            mine[lane * 4 + 1] = s1;
            mine[lane * 4 + 2] = q0;
            mine[lane * 4 + 3] = q1;
            asm volatile("bar.sync %0, 128;" ::"r"(1 + group) : "memory");  // the four warps of this group
            if (quarter == 2) {  // first warp of the group (ew = 4 * group)
                float t[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int qw = 0; qw < 4; ++qw) {
                    const float* src = reinterpret_cast<const float*>(stg_base + (group * 4 + qw) * 4096);
#pragma unroll
                    for (int e = 0; e < 4; ++e) t[e] += src[lane * 4 + e];
                }
                float* out = p.partials + (size_t)blockIdx.x * 2 * p.N;
                const int col = col0 + 2 * lane;
                out[col] = t[0];
                out[col + 1] = t[1];
                out[p.N + col] = t[2];
                out[p.N + col + 1] = t[3];
            }
        }
    }
done:
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------ host side
// row-major bf16 [rows, cols]; box [box_rows x 64 cols], 128B swizzle
int make_map(CUtensorMap* map, const void* base, long long rows, long long cols, int box_rows) {
    return tma_host::encode_2d(map, base, rows, cols, cols * 2, box_rows, "conv1x1");
}

bool shape_ok(long long M, int N, int K) {
    auto pow2 = [](int v) { return v == 64 || v == 128 || v == 256 || v == 512; };
    return M >= 1 && M < (1ll << 31) - 256 && pow2(N) && pow2(K) && (long long)N * K <= 65536;
}

}  // namespace

DDL_API int ddl_conv1x1_supported(long long M, int N, int K) { return shape_ok(M, N, K) ? 1 : 0; }
// number of per-CTA partial-statistics rows (= gridDim.x); N > 256 splits the output columns over blockIdx.y and each
// column block gets half of the SMs
DDL_API int ddl_conv1x1_grid(long long M, int N) {
    const long long tiles = (M + kTileM - 1) / kTileM;
    const int col_blocks = N > 256 ? N / 256 : 1;
    long long gx = DDL_SM_COUNT / col_blocks;
    if (gx > tiles) gx = tiles;
    return (int)(gx < 1 ? 1 : gx);
}

static int conv1x1_launch(bool b_mn, const void* x, const void* w, void* y, float* partials, long long M, int N, int K, void* stream_) {
    DDL_REQUIRE(shape_ok(M, N, K), "conv1x1: unsupported shape M=%lld N=%d K=%d (N, K in {64,128,256,512}, N*K <= 65536)", M, N, K);
    DDL_REQUIRE(x && w && y && (partials || b_mn), "conv1x1: null pointer");
    C1Params p;
    p.y = (bf16*)y;
    p.partials = partials;
    p.M = (int)M;
    p.N = N;
    p.K = K;
    p.n_cta = N < 256 ? N : 256;
    const int col_blocks = N / p.n_cta;
    p.num_tiles = (int)((M + kTileM - 1) / kTileM);
    p.kblocks = K / 64;
    p.groups = p.n_cta / 64;
    p.nbuf = 512 / p.n_cta;
    if (p.nbuf > 4) p.nbuf = 4;
    const int fixed = p.n_cta * K * 2 + p.groups * 4 * 4096 + 512 + 1024;
    p.stages = (227 * 1024 - fixed) / kStageBytes;
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    DDL_REQUIRE(p.stages >= 2, "conv1x1: shape N=%d K=%d does not fit in shared memory", N, K);
    const int smem = fixed + p.stages * kStageBytes;
    CUtensorMap tx, tw;
    if (int e = make_map(&tx, x, M, K, kTileM)) return e;
    if (int e = b_mn ? make_map(&tw, w, K, N, 64) : make_map(&tw, w, N, K, p.n_cta)) return e;
    static int attr_smem[2] = {0, 0};
    if (smem > attr_smem[b_mn]) {
        if (b_mn) DDL_CHECK(cudaFuncSetAttribute(conv1x1_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        else DDL_CHECK(cudaFuncSetAttribute(conv1x1_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem[b_mn] = smem;
    }
    dim3 grid(ddl_conv1x1_grid(M, N), col_blocks);
    if (b_mn) conv1x1_fwd_kernel<true><<<grid, kThreads, smem, (cudaStream_t)stream_>>>(tx, tw, p);
    else conv1x1_fwd_kernel<false><<<grid, kThreads, smem, (cudaStream_t)stream_>>>(tx, tw, p);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_conv1x1_fwd(const void* x, const void* w, void* y, float* partials, long long M, int N, int K, void* stream_) {
    return conv1x1_launch(false, x, w, y, partials, M, N, K, stream_);
}

// dx [M, Cin] = dy [M, Cout] . w [Cout][Cin]: the same kernel with the weight read MN-major (supported when
// ddl_conv1x1_supported(M, Cin, Cout))
DDL_API int ddl_conv1x1_dgrad(const void* dy, const void* w, void* dx, long long M, int Cin, int Cout, void* stream_) {
    return conv1x1_launch(true, dy, w, dx, nullptr, M, Cin, Cout, stream_);
}
