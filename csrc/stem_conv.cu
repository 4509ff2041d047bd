// ResNet stem convolution (7x7, stride 2, pad 3, 3 -> 64 channels; reference model: torchvision resnet50 at
// notebooks/cv/onnx_experiments.py:19) as an implicit GEMM on tcgen05 (K9 in SURVEY.md section 2.E.3).
//
// Why its own kernel: with 3 input channels the library has no tensor-core path worth the name (measured on
// B200: 1.47 ms forward + 0.75 ms weight gradient per 256-image step, ~40 TFLOP/s), while the layer only
// moves 0.5 GB - its floor is the write of the 411 MB output.
//
// Formulation.  Space-to-depth turns the 7x7/2 convolution over 3 channels into a 4x4/1 convolution over
// 12 channels (padded to 16):
//     z[n, P, Q, (i*2+j)*3 + c] = xpad[n, 2P+i, 2Q+j, c]            (xpad: 3 pixels of zero padding)
//     y[n, oh, ow, k] = sum_{a,b<4, ch<16} W2[a][k][b*16+ch] * z[n, oh+a, ow+b, ch]
// For fixed (oh, a) the operand row of output pixel ow is z[n, oh+a, ow .. ow+3, :] = 64 contiguous bf16
// = 128 bytes, and the row of pixel ow+1 starts 32 bytes later.  A 2-D TMA tensor map whose row pitch is
// 32 bytes while the row length is 128 bytes (overlapping rows) therefore delivers a [128 pixels x 64]
// K-major im2col tile, already 128B-swizzled, with one cp.async.bulk.tensor - no im2col buffer exists.
//
// Forward: persistent, warp-specialised (TMA / MMA / 4 epilogue warps).  The output is tiled over the
// flattened padded index P = (n*Hp + oh)*Wp + ow so that every tile is one box (rows with ow >= OW or
// oh >= OH are computed and dropped: 5 %).  Four k-steps (a = 0..3) of M=128, N=64, K=64 per tile, the
// packed weights stay in shared memory, four TMEM accumulator buffers decouple MMA and epilogue.  The
// epilogue converts to bf16, stores coalesced NHWC rows and accumulates the BatchNorm statistics
// (per-channel sum / sum of squares of the rounded outputs), so no separate statistics pass reads y.
//
// Weight gradient: dW2[a][b*16+ch][k] = sum_P z[P + a*Wp + b][ch] * dy[P][k] - a GEMM whose reduction runs over
// pixels.  Per output row (n, oh): one box of dy (OW x 64, MN-major B operand) and four boxes of z (MN-major
// A operand, taps a = 0..3 stacked in pairs to M = 128), 7 MMAs of K = 16 per pair for OW = 112; each CTA
// keeps its accumulators in TMEM over all its rows and writes one fp32 partial; a small kernel folds the
// partials in fixed order and scatters them to the [64, 3, 7, 7] gradient.
#include <cuda.h>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

namespace {
using namespace tc;

constexpr int kCout = 64;       // output channels
constexpr int kZC = 16;         // space-to-depth channels (12 used)
constexpr int kTaps = 4;        // vertical taps a (each covers 4 horizontal taps x 16 channels = K 64)
constexpr int kTileM = 128;
constexpr int kFwdStages = 2;   // one whole tile each: 4 vertical taps x 16 KB
constexpr int kFwdThreads = 320;  // TMA warp, MMA warp, 8 epilogue warps
constexpr int kAccBufs = 4;

// ------------------------------------------------------------------------------------------------ prep kernels
// x [N, 3, H, W] (fp32 or bf16, NCHW) -> z [N, Hp, Wp, 16] bf16.  One thread per z pixel.
template <typename T>
__global__ void __launch_bounds__(256) stem_s2d_kernel(const T* __restrict__ x, bf16* __restrict__ z, int N, int H, int W, int Hp,
                                                        int Wp) {
    const size_t total = (size_t)N * Hp * Wp;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int Q = (int)(t % Wp);
        const size_t r = t / Wp;
        const int P = (int)(r % Hp);
        const size_t n = r / Hp;
        float v[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = 0.f;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int row = 2 * P + i - 3;
            if (row < 0 || row >= H) continue;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int col = 2 * Q + j - 3;
                if (col < 0 || col >= W) continue;
#pragma unroll
                for (int c = 0; c < 3; ++c) v[(i * 2 + j) * 3 + c] = (float)x[((n * 3 + c) * H + row) * (size_t)W + col];
            }
        }
        float lo[8], hi[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            lo[e] = v[e];
            hi[e] = v[8 + e];
        }
        st_global_v4(z + t * 16, float_to_bf16x8(lo));
        st_global_v4(z + t * 16 + 8, float_to_bf16x8(hi));
    }
}

// w [64, 3, 7, 7] with element strides (sk, sc, sr, ss) -> W2[a][k][b*16 + ch] bf16 (zero where r or s == 7, ch >= 12)
__global__ void stem_pack_w_kernel(const bf16* __restrict__ w, long long sk, long long sc, long long sr, long long ss,
                                   bf16* __restrict__ w2) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;  // over 4 * 64 * 64
    if (t >= kTaps * kCout * 64) return;
    const int col = t & 63, k = (t >> 6) & 63, a = t >> 12;
    const int b = col >> 4, ch = col & 15;
    bf16 val = __float2bfloat16(0.f);
    if (ch < 12) {
        const int ij = ch / 3, c = ch % 3, i = ij >> 1, j = ij & 1;
        const int r = 2 * a + i, s = 2 * b + j;
        if (r < 7 && s < 7) val = w[k * sk + c * sc + r * sr + s * ss];
    }
    w2[t] = val;
}

// ------------------------------------------------------------------------------------------------ forward
struct StemFwdParams {
    bf16* y;          // [N, OH, OW, 64]
    float* partials;  // [grid][2][64] BatchNorm statistics of this CTA's outputs
    int Hp, Wp, OH, OW;
    long long total_pos;  // N * Hp * Wp
    int num_tiles;
};

constexpr int kFwdWBytes = kTaps * kCout * 128;                         // 32 KB packed weights
constexpr int kFwdTapBytes = kTileM * 128;                              // 16 KB: [128 positions x 64] of one vertical tap
constexpr int kFwdStageBytes = kTaps * kFwdTapBytes;                    // 64 KB
constexpr int kFwdStgOffset = kFwdWBytes + kFwdStages * kFwdStageBytes; // 8 epilogue warps x 2 KB
constexpr int kFwdBarOffset = kFwdStgOffset + 4 * 4096;
constexpr int kFwdSmem = kFwdBarOffset + 512 + 1024;

__global__ void __launch_bounds__(kFwdThreads, 1) stem_fwd_kernel(const __grid_constant__ CUtensorMap tma_z,
                                                                   const __grid_constant__ CUtensorMap tma_w, const StemFwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* w_smem = smem;
    uint8_t* a_smem = smem + kFwdWBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kFwdBarOffset);
    uint64_t* empty_bar = full_bar + kFwdStages;
    uint64_t* tmem_full = empty_bar + kFwdStages;
    uint64_t* tmem_empty = tmem_full + kAccBufs;
    uint64_t* w_bar = tmem_empty + kAccBufs;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = kAccBufs * kCout;  // 256

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_z);
        tma_prefetch_desc(&tma_w);
        for (int s = 0; s < kFwdStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < kAccBufs; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 8);
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
            mbar_expect_tx(w_bar, kFwdWBytes);
            tma_load_2d(w_smem, &tma_w, w_bar, 0, 0);
        }
        __syncwarp();
        // A stage holds all four taps of a tile and is handed back by the SAME commit that publishes the tile's accumulator
        // (the producer waits on tmem_full of the tile that used the stage, two tiles back): one tcgen05.commit per tile.
        // With a commit per tap the kernel ran commit-bound - 5 x ~450 cycles per tile against 16 x 48 cycles of MMAs.
        int it = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const long long p0 = (long long)tile * kTileM;
            const int stage = it & 1;
            if (it >= kFwdStages) {
                const int j = it - kFwdStages;  // (its barrier cannot complete again before this stage is refilled)
                mbar_wait(&tmem_full[j % kAccBufs], (uint32_t)((j / kAccBufs) & 1));
            }
            if (elect_one()) {
                mbar_expect_tx(&full_bar[stage], kFwdStageBytes);
                // rows past the end of z are zero-filled by TMA (they only feed dropped outputs)
#pragma unroll
                for (int a = 0; a < kTaps; ++a)
                    tma_load_2d(a_smem + stage * kFwdStageBytes + a * kFwdTapBytes, &tma_z, &full_bar[stage], 0, (int)(p0 + (long long)a * p.Wp));
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===================== MMA issue: the whole warp runs the loop, one elected lane issues =====================
        constexpr uint32_t idesc = make_idesc(kTileM, kCout, false, false);
        const uint64_t d_const = make_smem_desc(0, 16, 1024);
        const uint64_t da0 = d_const | (uint64_t)(smem_u32(a_smem) >> 4), db0 = d_const | (uint64_t)(smem_u32(w_smem) >> 4);
        mbar_wait(w_bar, 0);
        tc_fence_after();
        int it = 0, acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int stage = it & 1;
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            mbar_wait(&full_bar[stage], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * kCout;
            if (elect_one()) {
#pragma unroll
                for (int a = 0; a < kTaps; ++a) {
                    const uint64_t da = da0 + (uint32_t)(stage * (kFwdStageBytes >> 4) + a * (kFwdTapBytes >> 4));
                    const uint64_t db = db0 + (uint32_t)(a * ((kCout * 128) >> 4));
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(d_tmem, da + k * 2, db + k * 2, idesc, (a | k) ? 1u : 0u);
                }
                umma_commit(&tmem_full[acc]);
            }
            __syncwarp();
            if (++acc == kAccBufs) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ===================== epilogue warps (2..9): TMEM lane quarter x 32-column half =====================
        // (with four warps the epilogue - ~1500 cycles per tile per warp - was what bounded the kernel, not the 768 cycles
        // of MMAs; offsets are 32-bit, checked on the host)
        const int ew = warp - 2, quarter = warp & 3, h = ew >> 2;
        uint8_t* stg = smem + kFwdStgOffset + ew * 2048;
        float cs[8], cq[8];  // columns h * 32 + t_ch * 8 + q
#pragma unroll
        for (int q = 0; q < 8; ++q) cs[q] = cq[q] = 0.f;
        const int t_row = lane >> 2, t_ch = lane & 3;
        const int img = p.Hp * p.Wp, total = (int)p.total_pos;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            // my row: flattened padded position -> element offset of the output row (or -1: dropped)
            const int pos = tile * kTileM + quarter * 32 + lane;
            int out_off = -1;
            if (pos < total) {
                const int n = pos / img, rem = pos - n * img;
                const int oh = rem / p.Wp, ow = rem - oh * p.Wp;
                if (oh < p.OH && ow < p.OW) out_off = ((n * p.OH + oh) * p.OW + ow) * kCout;
            }
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + acc * kCout + h * 32 + ((uint32_t)(quarter * 32) << 16), r);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            if (++acc == kAccBufs) { acc = 0; acc_phase ^= 1; }
            // bf16 row into the warp's staging tile (16-byte chunks XOR-swizzled: conflict-free both ways)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
                *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = float_to_bf16x8(f);
            }
            __syncwarp();
            // transposed domain: 4 lanes = one 64-byte row piece; coalesced stores, statistics of the stored values
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int rr = it * 8 + t_row;
                const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((t_ch ^ ((rr >> 1) & 3)) << 4));
                const int off = __shfl_sync(0xffffffffu, out_off, rr);
                if (off >= 0) {
                    st_global_v4(p.y + off + h * 32 + t_ch * 8, v);
                    float f[8];
                    bf16x8_to_float(v, f);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        cs[q] += f[q];
                        cq[q] = fmaf(f[q], f[q], cq[q]);
                    }
                }
            }
            __syncwarp();
        }
        // fold the 8 row groups of the warp, then - in fixed order - the four lane-quarter warps of each column half
#pragma unroll
        for (int q = 0; q < 8; ++q) {
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                cs[q] += __shfl_xor_sync(0xffffffffu, cs[q], o);
                cq[q] += __shfl_xor_sync(0xffffffffu, cq[q], o);
            }
        }
        float* mine = reinterpret_cast<float*>(stg);  // own staging area, free now: [32 columns][2]
        if (t_row == 0) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                mine[(t_ch * 8 + q) * 2] = cs[q];
                mine[(t_ch * 8 + q) * 2 + 1] = cq[q];
            }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight epilogue warps only
        if (ew == 0) {
            // lane -> columns lane (half 0) and 32 + lane (half 1)
            float* out = p.partials + (size_t)blockIdx.x * 2 * kCout;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                float ts = 0.f, tq = 0.f;
#pragma unroll
                for (int qw = 0; qw < 4; ++qw) {
                    const float* src = reinterpret_cast<const float*>(smem + kFwdStgOffset + (hh * 4 + qw) * 2048);
                    ts += src[lane * 2];
                    tq += src[lane * 2 + 1];
                }
                out[hh * 32 + lane] = ts;
                out[kCout + hh * 32 + lane] = tq;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------ weight gradient
struct StemWgradParams {
    float* partials;  // [grid][2 pairs][128 m][64 k] fp32
    int Hp, Wp, OH, OW;
    int num_rows;     // N * OH
    int stages;
};

constexpr int kWgThreads = 192;

// grid = persistent CTAs; each loops over output rows (n, oh).  Shared memory per stage: dy box + 4 z boxes of OW x 128 B.
__global__ void __launch_bounds__(kWgThreads, 1) stem_wgrad_kernel(const __grid_constant__ CUtensorMap tma_z,
                                                                    const __grid_constant__ CUtensorMap tma_dy,
                                                                    const StemWgradParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int box_bytes = p.OW * 128;
    const int stage_bytes = 5 * box_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.stages;
    uint64_t* done_bar = empty_bar + p.stages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(done_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 128;  // two accumulators of 64 columns

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_z);
        tma_prefetch_desc(&tma_dy);
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    const bool has_work = (int)blockIdx.x < p.num_rows;

    if (warp == 0) {
        // producer: warp-uniform loop, one elected lane issues
        int stage = 0;
        uint32_t phase = 0;
        for (int row = blockIdx.x; row < p.num_rows; row += gridDim.x) {
            const int n = row / p.OH, oh = row - n * p.OH;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* st = smem + (size_t)stage * stage_bytes;
            if (elect_one()) {
                mbar_expect_tx(&full_bar[stage], stage_bytes);
                tma_load_2d(st, &tma_dy, &full_bar[stage], 0, row * p.OW);
#pragma unroll
                for (int a = 0; a < kTaps; ++a) tma_load_2d(st + (1 + a) * box_bytes, &tma_z, &full_bar[stage], 0, (n * p.Hp + oh + a) * p.Wp);
            }
            __syncwarp();
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // A: z boxes, MN-major, M = 128 = taps (a, a+1) x 64; B: dy box, MN-major, N = 64; K = pixels of the row.  The
        // whole warp runs the loop, one elected lane issues; a k-step of 16 pixels adds 2048 B to both descriptors.
        constexpr uint32_t idesc = make_idesc(128, kCout, true, true);
        const uint64_t d_const = make_smem_desc(0, box_bytes, 1024);
        const uint32_t smem_base = smem_u32(smem);
        int stage = 0;
        uint32_t phase = 0;
        bool first = true;
        const int ksteps = p.OW / 16;
        for (int row = blockIdx.x; row < p.num_rows; row += gridDim.x) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t sdy = smem_base + (uint32_t)stage * (uint32_t)stage_bytes;
            const uint64_t db0 = d_const | (uint64_t)(sdy >> 4);
            if (elect_one()) {
#pragma unroll
                for (int pair = 0; pair < 2; ++pair) {
                    const uint64_t da0 = d_const | (uint64_t)((sdy + (1 + 2 * pair) * box_bytes) >> 4);
                    for (int k = 0; k < ksteps; ++k) umma_bf16(tmem_base + pair * kCout, da0 + k * 128, db0 + k * 128, idesc, (!first || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
            first = false;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(done_bar);
        __syncwarp();
    } else {
        // epilogue: TMEM -> this CTA's fp32 partial (rows m = 32*quarter + lane)
        const int quarter = warp & 3;
        float* out = p.partials + (size_t)blockIdx.x * 2 * 128 * kCout;
        if (has_work) {
            mbar_wait(done_bar, 0);
            tc_fence_after();
        }
#pragma unroll
        for (int pair = 0; pair < 2; ++pair)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t r[32];
                if (has_work) {
                    tmem_ld_32x32(tmem_base + pair * kCout + h * 32 + ((uint32_t)(quarter * 32) << 16), r);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) r[e] = 0u;
                }
                float* dst = out + ((size_t)pair * 128 + quarter * 32 + lane) * kCout + h * 32;
#pragma unroll
                for (int e = 0; e < 32; e += 4)
                    *reinterpret_cast<float4*>(dst + e) = make_float4(__uint_as_float(r[e]), __uint_as_float(r[e + 1]),
                                                                      __uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]));
            }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// dW[k, c, r, s] (element strides sk, sc, sr, ss) = sum over CTAs of partial[cta][a/2][(a%2)*64 + b*16 + ch][k]
__global__ void stem_wgrad_finalize_kernel(const float* __restrict__ partials, int nparts, bf16* __restrict__ dw, long long sk,
                                           long long sc, long long sr, long long ss) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;  // over 3*7*7*64, k fastest
    if (t >= 3 * 49 * kCout) return;
    const int k = t & 63, crs = t >> 6;
    const int s = crs % 7, r = (crs / 7) % 7, c = crs / 49;
    const int a = r >> 1, i = r & 1, b = s >> 1, j = s & 1;
    const int m = (a & 1) * 64 + b * 16 + (i * 2 + j) * 3 + c;
    const float* src = partials + ((size_t)(a >> 1) * 128 + m) * kCout + k;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int q = 0;
    for (; q + 3 < nparts; q += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) acc[u] += src[(size_t)(q + u) * 2 * 128 * kCout];
    }
    for (; q < nparts; ++q) acc[0] += src[(size_t)q * 2 * 128 * kCout];
    dw[k * sk + c * sc + r * sr + s * ss] = __float2bfloat16((acc[0] + acc[1]) + (acc[2] + acc[3]));
}

// ------------------------------------------------------------------------------------------------ host side
// [rows x 64] bf16 with row pitch `pitch_bytes` (32 for the overlapping im2col rows of z), box [box_rows x 64], 128B swizzle
int make_rows_map(CUtensorMap* map, const void* base, long long rows, long long pitch_bytes, int box_rows) {
    return tma_host::encode_2d(map, base, rows, 64, pitch_bytes, box_rows, "stem conv");
}

int check_geom(int N, int H, int W) {
    DDL_REQUIRE(N >= 1 && H >= 8 && W >= 8 && H % 2 == 0 && W % 2 == 0, "stem conv: need even H, W >= 8 (got %d x %d)", H, W);
    DDL_REQUIRE((long long)N * (H / 2 + 3) * (W / 2 + 3) + 4ll * (W / 2 + 3) < 0x7fffffffll, "stem conv: batch too large for 32-bit TMA coordinates");
    return 0;
}

int grid_1d(size_t total) {
    long long b = (long long)((total + 255) / 256);
    const long long cap = DDL_SM_COUNT * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

DDL_API size_t ddl_stem_z_bytes(int N, int H, int W) { return (size_t)N * (H / 2 + 3) * (W / 2 + 3) * kZC * sizeof(bf16); }
DDL_API int ddl_stem_fwd_grid(int N, int H, int W) {
    const long long tiles = ((long long)N * (H / 2 + 3) * (W / 2 + 3) + kTileM - 1) / kTileM;
    return (int)(tiles < DDL_SM_COUNT ? tiles : DDL_SM_COUNT);
}
DDL_API int ddl_stem_wgrad_grid(int N, int H, int W) {
    const long long rows = (long long)N * (H / 2);
    return (int)(rows < DDL_SM_COUNT ? rows : DDL_SM_COUNT);
}
DDL_API size_t ddl_stem_wgrad_scratch_bytes(int N, int H, int W) {
    return (size_t)ddl_stem_wgrad_grid(N, H, W) * 2 * 128 * kCout * sizeof(float);
}

DDL_API int ddl_stem_s2d(const void* x, int x_dtype, void* z, int N, int H, int W, void* stream_) {
    if (int e = check_geom(N, H, W)) return e;
    DDL_REQUIRE(x && z, "stem s2d: null pointer");
    DDL_REQUIRE(x_dtype == DDL_F32 || x_dtype == DDL_BF16, "stem s2d: input must be float32 or bfloat16");
    const int Hp = H / 2 + 3, Wp = W / 2 + 3;
    const int grid = grid_1d((size_t)N * Hp * Wp);
    if (x_dtype == DDL_F32) stem_s2d_kernel<float><<<grid, 256, 0, (cudaStream_t)stream_>>>((const float*)x, (bf16*)z, N, H, W, Hp, Wp);
    else stem_s2d_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream_>>>((const bf16*)x, (bf16*)z, N, H, W, Hp, Wp);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_stem_conv_fwd(const void* z, const void* w, long long sk, long long sc, long long sr, long long ss, void* w2, void* y,
                              float* partials, int N, int H, int W, void* stream_) {
    if (int e = check_geom(N, H, W)) return e;
    DDL_REQUIRE(z && w && w2 && y && partials, "stem fwd: null pointer");
    cudaStream_t s = (cudaStream_t)stream_;
    stem_pack_w_kernel<<<(kTaps * kCout * 64 + 255) / 256, 256, 0, s>>>((const bf16*)w, sk, sc, sr, ss, (bf16*)w2);
    DDL_LAUNCH_CHECK();
    StemFwdParams p;
    p.y = (bf16*)y;
    p.partials = partials;
    p.Hp = H / 2 + 3;
    p.Wp = W / 2 + 3;
    p.OH = H / 2;
    p.OW = W / 2;
    p.total_pos = (long long)N * p.Hp * p.Wp;
    DDL_REQUIRE(p.total_pos < (1ll << 31) - kTileM && (long long)N * p.OH * p.OW * kCout < (1ll << 31), "stem conv: batch too large for 32-bit offsets");
    p.num_tiles = (int)((p.total_pos + kTileM - 1) / kTileM);
    CUtensorMap tz, tw;
    if (int e = make_rows_map(&tz, z, p.total_pos - 3, kZC * 2, kTileM)) return e;
    if (int e = make_rows_map(&tw, w2, kTaps * kCout, 128, kTaps * kCout)) return e;
    static bool attr_set = false;
    if (!attr_set) {
        DDL_CHECK(cudaFuncSetAttribute(stem_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        attr_set = true;
    }
    stem_fwd_kernel<<<ddl_stem_fwd_grid(N, H, W), kFwdThreads, kFwdSmem, s>>>(tz, tw, p);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_stem_conv_wgrad(const void* z, const void* dy, void* dw, long long sk, long long sc, long long sr, long long ss,
                                void* scratch, size_t scratch_bytes, int N, int H, int W, void* stream_) {
    if (int e = check_geom(N, H, W)) return e;
    DDL_REQUIRE(z && dy && dw && scratch, "stem wgrad: null pointer");
    DDL_REQUIRE(scratch_bytes >= ddl_stem_wgrad_scratch_bytes(N, H, W), "stem wgrad: scratch too small");
    StemWgradParams p;
    p.partials = (float*)scratch;
    p.Hp = H / 2 + 3;
    p.Wp = W / 2 + 3;
    p.OH = H / 2;
    p.OW = W / 2;
    p.num_rows = N * p.OH;
    DDL_REQUIRE(p.OW % 16 == 0 && p.OW <= 256, "stem wgrad: output width %d must be a multiple of 16 and <= 256", p.OW);
    const int stage_bytes = 5 * p.OW * 128;
    p.stages = (220 * 1024) / stage_bytes;
    if (p.stages > 4) p.stages = 4;
    DDL_REQUIRE(p.stages >= 1, "stem wgrad: output width %d does not fit in shared memory", p.OW);
    const int smem = p.stages * stage_bytes + 512 + 1024;
    CUtensorMap tz, tdy;
    if (int e = make_rows_map(&tz, z, (long long)N * p.Hp * p.Wp - 3, kZC * 2, p.OW)) return e;
    if (int e = make_rows_map(&tdy, dy, (long long)N * p.OH * p.OW, 128, p.OW)) return e;
    static int attr_smem = 0;
    if (smem > attr_smem) {
        DDL_CHECK(cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem = smem;
    }
    cudaStream_t s = (cudaStream_t)stream_;
    const int grid = ddl_stem_wgrad_grid(N, H, W);
    stem_wgrad_kernel<<<grid, kWgThreads, smem, s>>>(tz, tdy, p);
    DDL_LAUNCH_CHECK();
    stem_wgrad_finalize_kernel<<<(3 * 49 * kCout + 255) / 256, 256, 0, s>>>(p.partials, grid, (bf16*)dw, sk, sc, sr, ss);
    DDL_LAUNCH_CHECK();
    return 0;
}
