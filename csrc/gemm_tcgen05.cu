// bf16 GEMM on the 5th-generation tensor cores: tcgen05.mma with TMEM accumulators, TMA
// operand loads, mbarrier pipelines, warp-specialised persistent CTAs (K8, SURVEY.md §2.E.3;
// BASELINE.json:5 "tcgen05/TMEM GEMM").  Written directly in PTX; descriptor layouts follow
// the canonical UMMA layouts (K-major / MN-major, 128-byte swizzle).
//
//      C[M, N] = epilogue( A[M, K] . B[N, K]^T )          fp32 accumulation in TMEM
//
// Operand storage (row-major in global memory, bf16):
//      A: [M, K] (K-major)   or  [K, M] (MN-major, `a_mn`)
//      B: [N, K] (K-major)   or  [K, N] (MN-major, `b_mn`)
// which covers the three GEMMs of a linear layer without any transpose kernel:
//      forward  y  = x W^T        A = x  [M,K]  K-major      B = W  [N,K] K-major
//      dgrad    dx = dy W         A = dy [M,N]  K-major      B = W  stored [K=N, N=Kin]  MN-major
//      wgrad    dW = dy^T x       A = dy stored [K=tok, M=N] MN-major,  B = x stored [K=tok, N=Kin] MN-major
//
// CTA = 10 warps: warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2-9 epilogue
// (two warps per TMEM lane quarter, each taking half of the tile's columns).  Tile
// 128 x BN (BN = 128 | 256) x 64, UMMA 128 x BN x 16, `kStages` smem stages, two TMEM
// accumulator stages so the epilogue of tile i overlaps the mainloop of tile i+1.
//
// Epilogues (fused, per BASELINE.json:5): bias add, exact (erf) GELU with the pre-activation
// kept for backward, dGELU multiply, residual / accumulate add, bf16 store; split-K with
// fp32 `red.global.add` into a caller-zeroed buffer for the skinny wgrad shapes.
#include <cuda.h>

#include <mutex>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;
constexpr int kThreads = (2 + kEpiWarps) * 32;

using namespace tc;

// ---------------------------------------------------------------------------- epilogue math
// GELU(x) = x * Phi(x), Phi from Abramowitz-Stegun 7.1.26 (|erf error| < 1.5e-7, far below bf16 resolution).
// With z = |x|/sqrt2, t = 1/(1 + p z), q = (a1 t + ... + a5 t^5) exp(-z^2):  erf(z) = 1 - q, hence
//     gelu(x)  = max(x, 0) - 0.5 |x| q          (12 instructions, two of them MUFU)
//     gelu'(x) = Phi(x) + x phi(x),  Phi(x) = 0.5 + sign(x) (0.5 - 0.5 q),  phi(x) = exp(-x^2/2)/sqrt(2 pi)
// and exp(-z^2) = exp(-x^2/2) is the same exponential in both.
__device__ __forceinline__ float gelu_q(float x, float& e) {
    const float ax = fabsf(x);
    const float t = __fdividef(1.f, fmaf(0.3275911f * 0.70710678f, ax, 1.f));
    float pl = fmaf(1.061405429f, t, -1.453152027f);
    pl = fmaf(pl, t, 1.421413741f);
    pl = fmaf(pl, t, -0.284496736f);
    pl = fmaf(pl, t, 0.254829592f);
    e = exp2f(x * x * (-0.5f * 1.4426950408889634f));
    return pl * t * e;
}
__device__ __forceinline__ float gelu_f(float x) {
    float e;
    const float q = gelu_q(x, e);
    return fmaf(-0.5f * fabsf(x), q, fmaxf(x, 0.f));
}
__device__ __forceinline__ float dgelu_f(float x) {
    float e;
    const float q = gelu_q(x, e);
    const float cdf = 0.5f + copysignf(fmaf(-0.5f, q, 0.5f), x);
    return fmaf(x * 0.39894228f, e, cdf);
}

struct GemmParams {
    int M, N, K;
    bf16* C;           // [M, ldc]
    long long ldc;
    const bf16* bias;  // [N] or null
    const bf16* residual;  // [M, ldc] or null: added to the result (also used as "accumulate into grad")
    bf16* aux;         // EPI_GELU: pre-activation out [M, ldc];  EPI_DGELU: pre-activation in
    float* acc_f32;    // split-K: fp32 [M, N] accumulation target (caller-zeroed), ld = N
    float* col_stats;  // optional [2][N] fp32 (caller-zeroed): per-column sum and sum of squares of the stored output
    int epilogue;      // DDL_EPI_*
    int split_k;
    int num_m_blocks, num_n_blocks, k_blocks_total;
};

// Epilogue of one 128-row x BN-column accumulator tile, executed by one epilogue warp (32 rows x BN/2
// columns).  TMEM gives each thread one ROW; global memory wants one 128-byte line per 8 lanes.  The
// warp stages each 32 x 64 bf16 chunk in a private 4 KB swizzled smem buffer and re-reads it
// transposed: lane l <-> (row it*4 + l/8, 16-byte chunk l%8), so every global access (output, aux,
// residual) is 4 full lines per instruction instead of 32 partial ones.
template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint8_t* stg, uint32_t tmem_base, int acc_stage, int m_blk,
                                              int n_blk, int quarter, int half, int lane, uint64_t* full_bar, uint32_t full_phase) {
    constexpr int kColsPerWarp = BN / 2;
    const int t_sr = lane >> 3, t_ch = lane & 7;
    auto stg_row = [&](int r, int c) -> uint4* { return reinterpret_cast<uint4*>(stg + r * 128 + ((c ^ (r & 7)) << 4)); };
    const int row_base = m_blk * BM + quarter * 32;   // first row of this warp's 32
    const int row = row_base + lane;
    // the one tensor this epilogue READS per element (dGELU: saved pre-activation, else: residual)
    const bf16* ext_ptr = p.epilogue == DDL_EPI_DGELU ? p.aux : p.residual;
    const bool has_ext = ext_ptr != nullptr && p.split_k == 1;
#pragma unroll 1
    for (int c = 0; c < kColsPerWarp; c += 64) {
        const int col0 = n_blk * BN + half * kColsPerWarp + c;
        // Issue every external load of this chunk FIRST, in the coalesced (transposed) mapping, so that the
        // 8 independent 16-byte requests per lane are in flight while the accumulator is fetched and the
        // math runs (for the first chunk: while the tile's last MMAs are still executing).
        uint4 ext[8];
        if (has_ext) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int rr = it * 4 + t_sr;
                ext[it] = make_uint4(0, 0, 0, 0);
                if (row_base + rr < p.M && col0 + t_ch * 8 < p.N)
                    ext[it] = ld_stream_v4(ext_ptr + (size_t)(row_base + rr) * p.ldc + col0 + t_ch * 8);
            }
        }
        if (c == 0) {
            mbar_wait(full_bar, full_phase);
            tc_fence_after();
        }
        uint32_t r0[32], r1[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc_stage * BN + half * kColsPerWarp + c);
        tmem_ld_32x32(taddr, r0);
        tmem_ld_32x32(taddr + 32, r1);
        tmem_ld_wait();
        if (col0 >= p.N) continue;
        if (p.split_k > 1) {
            if (row < p.M) {
                float* dst = p.acc_f32 + (size_t)row * p.N + col0;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    if (col0 + j < p.N)
                        red_add_f32x4(dst + j, __uint_as_float(r0[j]), __uint_as_float(r0[j + 1]), __uint_as_float(r0[j + 2]), __uint_as_float(r0[j + 3]));
                    if (col0 + 32 + j < p.N)
                        red_add_f32x4(dst + 32 + j, __uint_as_float(r1[j]), __uint_as_float(r1[j + 1]), __uint_as_float(r1[j + 2]), __uint_as_float(r1[j + 3]));
                }
            }
            continue;
        }
        // row domain (thread = row): + bias, ReLU; round to bf16 and stage
#pragma unroll
        for (int g = 0; g < 8; ++g) {  // 8 column groups of 8
            float w[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) w[q] = __uint_as_float(g < 4 ? r0[g * 8 + q] : r1[(g - 4) * 8 + q]);
            if (p.bias && col0 + g * 8 < p.N) {
                float bsv[8];
                bf16x8_to_float(ld_global_v4(p.bias + col0 + g * 8), bsv);
#pragma unroll
                for (int q = 0; q < 8; ++q) w[q] += bsv[q];
            }
            if (p.epilogue == DDL_EPI_RELU) {
#pragma unroll
                for (int q = 0; q < 8; ++q) w[q] = fmaxf(w[q], 0.f);
            }
            *stg_row(lane, g) = float_to_bf16x8(w);
        }
        __syncwarp();
        // transposed domain (8 lanes = one 128-byte line): activation / dGELU / residual, coalesced stores
        float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, cq[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + t_sr;
            uint4 pv = *stg_row(rr, t_ch);
            if (row_base + rr < p.M && col0 + t_ch * 8 < p.N) {
                const size_t off = (size_t)(row_base + rr) * p.ldc + col0 + t_ch * 8;
                if (p.epilogue == DDL_EPI_GELU) {
                    if (p.aux) st_global_v4(p.aux + off, pv);  // backward sees exactly this rounded pre-activation
                    float f[8];
                    bf16x8_to_float(pv, f);
#pragma unroll
                    for (int q = 0; q < 8; ++q) f[q] = gelu_f(f[q]);
                    if (has_ext) {
                        float rs[8];
                        bf16x8_to_float(ext[it], rs);
#pragma unroll
                        for (int q = 0; q < 8; ++q) f[q] += rs[q];
                    }
                    pv = float_to_bf16x8(f);
                } else if (p.epilogue == DDL_EPI_DGELU) {
                    float f[8], pre[8];
                    bf16x8_to_float(pv, f);
                    bf16x8_to_float(ext[it], pre);
#pragma unroll
                    for (int q = 0; q < 8; ++q) f[q] *= dgelu_f(pre[q]);
                    pv = float_to_bf16x8(f);
                } else if (has_ext) {
                    float f[8], rs[8];
                    bf16x8_to_float(pv, f);
                    bf16x8_to_float(ext[it], rs);
#pragma unroll
                    for (int q = 0; q < 8; ++q) f[q] += rs[q];
                    pv = float_to_bf16x8(f);
                }
                st_global_v4(p.C + off, pv);
                if (p.col_stats) {
                    float f[8];
                    bf16x8_to_float(pv, f);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        cs[q] += f[q];
                        cq[q] += f[q] * f[q];
                    }
                }
            }
        }
        if (p.col_stats) {
            // BatchNorm statistics of a 1x1 convolution come out of its GEMM for free: fold the 4 row groups
            // of the warp, then one red.add per column from lanes 0..7 (K9: conv + BN-statistics epilogue)
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                cs[q] += __shfl_xor_sync(0xffffffffu, cs[q], 8);
                cs[q] += __shfl_xor_sync(0xffffffffu, cs[q], 16);
                cq[q] += __shfl_xor_sync(0xffffffffu, cq[q], 8);
                cq[q] += __shfl_xor_sync(0xffffffffu, cq[q], 16);
            }
            if (t_sr == 0 && col0 + t_ch * 8 < p.N) {
                float* d0 = p.col_stats + col0 + t_ch * 8;
                red_add_f32x4(d0, cs[0], cs[1], cs[2], cs[3]);
                red_add_f32x4(d0 + 4, cs[4], cs[5], cs[6], cs[7]);
                red_add_f32x4(d0 + p.N, cq[0], cq[1], cq[2], cq[3]);
                red_add_f32x4(d0 + p.N + 4, cq[4], cq[5], cq[6], cq[7]);
            }
        }
        __syncwarp();  // buffer free for the next chunk
    }
}

template <int BN, bool A_MN, bool B_MN, int kStages>
struct SmemLayout {
    static constexpr int kABytes = BM * BK * 2;
    static constexpr int kBBytes = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStageOffset = kStages * kStageBytes;          // 8 epilogue warps x 4 KB staging
    static constexpr int kBarOffset = kStageOffset + kEpiWarps * 4096;
    static constexpr int kTotal = kBarOffset + 256 + 1024;  // barriers + alignment slack
};

template <int BN, bool A_MN, bool B_MN, int kStages, int kCluster>
__global__ void __launch_bounds__(kThreads, 1) gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                                const __grid_constant__ CUtensorMap tma_b,
                                                                const __grid_constant__ CUtensorMap tma_b_half, const GemmParams p) {
    using L = SmemLayout<BN, A_MN, B_MN, kStages>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full = empty_bar + kStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;         // [2]
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 2 * BN;  // two accumulator stages (256 or 512 columns)

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kCluster);  // every CTA that receives the multicast B must release the stage
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], kEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    if (kCluster > 1) cluster_sync_all();  // peers' barriers exist before any multicast can land
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    // kCluster == 2: the two CTAs of a cluster take vertically adjacent tiles (m, n) / (m+1, n) and
    // share the B tile: each loads half of it and multicasts that half to both, which cuts the
    // L2 -> SM operand traffic per CTA from A + B to A + B/2 (the mainloop is L2-bandwidth bound).
    const int cta_rank = kCluster > 1 ? (int)cluster_ctarank() : 0;
    const int sched_id = blockIdx.x / kCluster, sched_n = gridDim.x / kCluster;
    const int m_groups = (p.num_m_blocks + kCluster - 1) / kCluster;
    const int tiles_mn = m_groups * p.num_n_blocks;
    const int num_tiles = tiles_mn * p.split_k;
    const int kb_per_split = (p.k_blocks_total + p.split_k - 1) / p.split_k;

    if (warp == 0) {
        // ===================== TMA producer (warp-uniform loop, one elected lane issues) =====================
        {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
                const int ks = tile / tiles_mn, mn = tile % tiles_mn;
                const int m_blk = (mn / p.num_n_blocks) * kCluster + cta_rank, n_blk = mn % p.num_n_blocks;
                const int kb0 = ks * kb_per_split;
                const int kb1 = min(kb0 + kb_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::kStageBytes;
                    uint8_t* sb = sa + L::kABytes;
                    if (elect_one()) {
                    mbar_expect_tx(&full_bar[stage], L::kStageBytes);
                    if (A_MN) {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tma_a, &full_bar[stage], m_blk * BM + j * 64, kb * BK);
                    } else {
                        tma_load_2d(sa, &tma_a, &full_bar[stage], kb * BK, m_blk * BM);
                    }
                    if (kCluster == 1) {
                        if (B_MN) {
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tma_b, &full_bar[stage], n_blk * BN + j * 64, kb * BK);
                        } else {
                            tma_load_2d(sb, &tma_b, &full_bar[stage], kb * BK, n_blk * BN);
                        }
                    } else {
                        // my half of B, delivered to both CTAs (same offsets, both full barriers)
                        constexpr int kHalf = BN / 2;
                        if (B_MN) {
#pragma unroll
                            for (int j = 0; j < kHalf / 64; ++j) {
                                const int jj = cta_rank * (kHalf / 64) + j;
                                tma_load_2d_mcast(sb + jj * (BK * 128), &tma_b, &full_bar[stage], n_blk * BN + jj * 64, kb * BK, 0x3);
                            }
                        } else {
                            tma_load_2d_mcast(sb + cta_rank * (kHalf * 128), &tma_b_half, &full_bar[stage], kb * BK, n_blk * BN + cta_rank * kHalf, 0x3);
                        }
                    }
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue: the whole warp runs the loop, one elected lane issues =====================
        // (stage descriptors = constant | address >> 4, a k-step adds 32 B (K-major) or 2048 B (MN-major) to the address)
        {
            constexpr uint32_t idesc = make_idesc(BM, BN, A_MN, B_MN);
            const uint64_t da_const = A_MN ? make_smem_desc(0, BK * 128, 1024) : make_smem_desc(0, 16, 1024);
            const uint64_t db_const = B_MN ? make_smem_desc(0, BK * 128, 1024) : make_smem_desc(0, 16, 1024);
            constexpr uint32_t kStepA = (A_MN ? 2048 : 32) >> 4, kStepB = (B_MN ? 2048 : 32) >> 4;
            const uint32_t smem_base = smem_u32(smem);
            int stage = 0;
            uint32_t phase = 0;
            int acc_stage = 0;
            uint32_t acc_phase = 0;
            for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
                const int ks = tile / tiles_mn;
                const int kb0 = ks * kb_per_split;
                const int kb1 = min(kb0 + kb_per_split, p.k_blocks_total);
                mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc_stage * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    // (the address field is 14 bits of (address >> 4); in a cluster the shared-window address carries the CTA rank above it)
                    const uint32_t sa = (smem_base + (uint32_t)stage * (uint32_t)L::kStageBytes) & 0x3FFFFu;
                    const uint64_t da = da_const | (uint64_t)(sa >> 4), db = db_const | (uint64_t)((sa + L::kABytes) >> 4);
                    if (elect_one()) {
                        umma_bf16(d_tmem, da, db, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
                        for (int k = 1; k < BK / UMMA_K; ++k) umma_bf16(d_tmem, da + k * kStepA, db + k * kStepB, idesc, 1u);
                        // frees the smem stage once these MMAs retire (in both CTAs when B is shared)
                        if (kCluster > 1) umma_commit_mcast(&empty_bar[stage], 0x3);
                        else umma_commit(&empty_bar[stage]);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit(&tmem_full[acc_stage]);  // accumulator complete -> epilogue
                __syncwarp();
                if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;              // 0..7
        const int quarter = warp & 3;         // TMEM lane quarter this warp may access
        const int half = ew >> 2;             // which half of the tile's columns
        uint8_t* stg = smem + L::kStageOffset + ew * 4096;
        int acc_stage = 0;
        uint32_t acc_phase = 0;
        for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
            const int mn = tile % tiles_mn;
            const int m_blk = (mn / p.num_n_blocks) * kCluster + cta_rank, n_blk = mn % p.num_n_blocks;
            epilogue_tile<BN>(p, stg, tmem_base, acc_stage, m_blk, n_blk, quarter, half, lane, &tmem_full[acc_stage], acc_phase);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);
            if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (kCluster > 1) cluster_sync_all();  // nobody leaves while a peer may still multicast into its smem
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------
// cta_group::2 variant: a CTA PAIR (cluster of 2 on one TPC) computes one 256 x 256 tile.
// Each CTA stages its own 128 rows of A and its own half (128 rows) of B per k-block (32 KB per
// stage instead of 48 KB -> 6 stages, and half the shared-memory read traffic per MMA); the leader
// issues tcgen05.mma.cta_group::2 (M = 256, N = 256) which reads both CTAs' shared memory and
// accumulates into both CTAs' tensor memory; each CTA's epilogue drains its own 128 rows.
//   full barrier   : leader's; its producer arms 64 KB, both CTAs' TMA complete on it
//   empty barriers : one per CTA, released together by the leader's multicast commit
//   tmem_full      : one per CTA (multicast commit); tmem_empty: leader's, 16 warp arrivals
template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreads, 1) gemm2sm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                                   const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
    constexpr int BN = 256, kHalfN = 128, kStages = 6;
    constexpr int kABytes = BM * BK * 2, kBBytes = kHalfN * BK * 2, kStageBytes = kABytes + kBBytes;  // 32 KB
    constexpr int kStageOffset = kStages * kStageBytes, kBarOffset = kStageOffset + kEpiWarps * 4096;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBarOffset);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full = empty_bar + kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta_rank = (int)cluster_ctarank();
    const bool leader = cta_rank == 0;
    constexpr uint32_t kTmemCols = 512;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 2 * kEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_2sm(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int sched_id = blockIdx.x >> 1, sched_n = gridDim.x >> 1;
    const int m_pairs = (p.num_m_blocks + 1) / 2;  // 256-row blocks
    const int tiles_mn = m_pairs * p.num_n_blocks;
    const int num_tiles = tiles_mn * p.split_k;
    const int kb_per_split = (p.k_blocks_total + p.split_k - 1) / p.split_k;

    if (warp == 0) {
        // producer: warp-uniform loop, one elected lane issues
        {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
                const int ks = tile / tiles_mn, mn = tile % tiles_mn;
                const int m_blk = (mn / p.num_n_blocks) * 2 + cta_rank, n_blk = mn % p.num_n_blocks;
                const int kb0 = ks * kb_per_split;
                const int kb1 = min(kb0 + kb_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * kStageBytes;
                    uint8_t* sb = sa + kABytes;
                    if (elect_one()) {
                    if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
                    if (A_MN) {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j) tma_load_2d_2sm(sa + j * (BK * 128), &tma_a, &full_bar[stage], m_blk * BM + j * 64, kb * BK);
                    } else {
                        tma_load_2d_2sm(sa, &tma_a, &full_bar[stage], kb * BK, m_blk * BM);
                    }
                    const int n0 = n_blk * BN + cta_rank * kHalfN;
                    if (B_MN) {
#pragma unroll
                        for (int j = 0; j < kHalfN / 64; ++j) tma_load_2d_2sm(sb + j * (BK * 128), &tma_b, &full_bar[stage], n0 + j * 64, kb * BK);
                    } else {
                        tma_load_2d_2sm(sb, &tma_b, &full_bar[stage], kb * BK, n0);
                    }
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // The whole warp runs the loop (stage addresses and descriptors stay warp-uniform: they are computed in uniform
        // registers and feed UTCHMMA without an ELECT / R2UR round trip per instruction); one elected lane issues.  A stage's
        // descriptors are a constant plus the 14-bit address field, a k-step adds 32 B (K-major) or 2048 B (MN-major).
        if (leader) {
            constexpr uint32_t idesc = make_idesc(2 * BM, BN, A_MN, B_MN);
            const uint64_t da_const = A_MN ? make_smem_desc(0, BK * 128, 1024) : make_smem_desc(0, 16, 1024);
            const uint64_t db_const = B_MN ? make_smem_desc(0, BK * 128, 1024) : make_smem_desc(0, 16, 1024);
            constexpr uint32_t kStepA = (A_MN ? 2048 : 32) >> 4, kStepB = (B_MN ? 2048 : 32) >> 4;
            const uint32_t smem_base = smem_u32(smem);
            int stage = 0;
            uint32_t phase = 0;
            int acc_stage = 0;
            uint32_t acc_phase = 0;
            for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
                const int ks = tile / tiles_mn;
                const int kb0 = ks * kb_per_split;
                const int kb1 = min(kb0 + kb_per_split, p.k_blocks_total);
                mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc_stage * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + (uint32_t)stage * (uint32_t)kStageBytes;
                    const uint64_t da = da_const | (uint64_t)(sa >> 4), db = db_const | (uint64_t)((sa + kABytes) >> 4);
                    if (elect_one()) {
                        umma_bf16_2sm(d_tmem, da, db, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
                        for (int k = 1; k < BK / UMMA_K; ++k) umma_bf16_2sm(d_tmem, da + k * kStepA, db + k * kStepB, idesc, 1u);
                        umma_commit_2sm(&empty_bar[stage]);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit_2sm(&tmem_full[acc_stage]);
                __syncwarp();
                if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
            }
        }
    } else {
        const int ew = warp - 2;
        const int quarter = warp & 3;
        const int half = ew >> 2;
        uint8_t* stg = smem + kStageOffset + ew * 4096;
        int acc_stage = 0;
        uint32_t acc_phase = 0;
        for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
            const int mn = tile % tiles_mn;
            const int m_blk = (mn / p.num_n_blocks) * 2 + cta_rank, n_blk = mn % p.num_n_blocks;
            epilogue_tile<BN>(p, stg, tmem_base, acc_stage, m_blk, n_blk, quarter, half, lane, &tmem_full[acc_stage], acc_phase);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&tmem_empty[acc_stage]);
            if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2sm(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------
// K6: all-gather fused into the consuming GEMM (BASELINE.json:5, FSDP mode).
// The weight W [N, K] is sharded by rows over the ranks (rank r owns rows [r N/W, (r+1) N/W) in its
// symmetric heap).  ONE kernel per rank does both jobs:
//   * two copier warps in every CTA pull the shards out of the peers' memory over NVLink into the local
//     full-W buffer (own shard first) and publish a per-shard ready counter (red.release.gpu);
//   * the TMA producer walks the output column blocks shard by shard, starting with the LOCAL shard, and
//     only waits (ld.acquire.gpu + fence.proxy.async) for a shard the first time it needs it -
//     so the math on the shards that already arrived overlaps the transfer of the ones still in flight.
// A cross-rank flag barrier (same protocol as csrc/collectives.cu) brackets the copies: peers' shards are
// final before anyone reads them and nobody's shard is overwritten while a peer still reads it.
struct GatherParams {
    const char* shard[DDL_MAX_WORLD];  // peer-mapped shard bases ([N/W, K] bf16, contiguous)
    char* full;                         // local [N, K] bf16
    uint32_t* flags;                    // local [DDL_MAX_WORLD + 1], zero before the first call (self-resetting)
    uint32_t* pad[DDL_MAX_WORLD];       // signal pads [blocks][DDL_MAX_WORLD]
    uint32_t* epoch;                    // local per-block barrier epochs
    int rank, world, blocks_per_shard;
    size_t shard_bytes;
};

__device__ __forceinline__ void gather_rank_barrier(const GatherParams& g, uint32_t epoch, int lane) {
    if (lane < g.world) {
        uint32_t* dst = g.pad[lane] + (size_t)blockIdx.x * DDL_MAX_WORLD + g.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
        const uint32_t* mine = g.pad[g.rank] + (size_t)blockIdx.x * DDL_MAX_WORLD + lane;
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        } while ((int32_t)(v - epoch) < 0);
    }
    __syncwarp();
}

constexpr int kGatherThreads = kThreads + 64;

__global__ void __launch_bounds__(kGatherThreads, 1) gemm_allgather_kernel(const __grid_constant__ CUtensorMap tma_a,
                                                                           const __grid_constant__ CUtensorMap tma_b, const GemmParams p,
                                                                           const GatherParams g) {
    constexpr int BN = 128, kStages = 6;
    using L = SmemLayout<BN, false, false, kStages>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full = empty_bar + kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 2 * BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], kEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int num_tiles = p.num_m_blocks * p.num_n_blocks;
    const uint32_t copiers_total = gridDim.x * 2;  // copier warps in the grid
    // column blocks in shard order, local shard first; all row blocks of one column block are consecutive
    auto tile_coords = [&](int tile, int& m_blk, int& n_blk) {
        const int n_idx = tile / p.num_m_blocks;
        m_blk = tile % p.num_m_blocks;
        n_blk = (n_idx + g.rank * g.blocks_per_shard) % p.num_n_blocks;
    };

    if (warp == 0) {
        {   // warp-uniform loop (every lane polls the shard flag), one elected lane issues the loads
            int stage = 0;
            uint32_t phase = 0, ready = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int m_blk, n_blk;
                tile_coords(tile, m_blk, n_blk);
                const int sh = n_blk / g.blocks_per_shard;
                if (!((ready >> sh) & 1u)) {
                    uint32_t v;
                    do {
                        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(g.flags + sh) : "memory");
                    } while (v < copiers_total);
                    __syncwarp();
                    fence_proxy_async_all();  // copiers wrote with ordinary stores; TMA reads through the async proxy
                    ready |= 1u << sh;
                }
                for (int kb = 0; kb < p.k_blocks_total; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::kStageBytes;
                    uint8_t* sb = sa + L::kABytes;
                    if (elect_one()) {
                        mbar_expect_tx(&full_bar[stage], L::kStageBytes);
                        tma_load_2d(sa, &tma_a, &full_bar[stage], kb * BK, m_blk * BM);
                        tma_load_2d(sb, &tma_b, &full_bar[stage], kb * BK, n_blk * BN);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        {   // warp-uniform loop, one elected lane issues
            constexpr uint32_t idesc = make_idesc(BM, BN, false, false);
            const uint64_t d_const = make_smem_desc(0, 16, 1024);
            const uint32_t smem_base = smem_u32(smem);
            int stage = 0, acc_stage = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc_stage * BN;
                for (int kb = 0; kb < p.k_blocks_total; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + (uint32_t)stage * (uint32_t)L::kStageBytes;
                    const uint64_t da = d_const | (uint64_t)(sa >> 4), db = d_const | (uint64_t)((sa + L::kABytes) >> 4);
                    if (elect_one()) {
                        umma_bf16(d_tmem, da, db, idesc, kb > 0 ? 1u : 0u);
#pragma unroll
                        for (int k = 1; k < BK / UMMA_K; ++k) umma_bf16(d_tmem, da + k * 2, db + k * 2, idesc, 1u);
                        umma_commit(&empty_bar[stage]);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit(&tmem_full[acc_stage]);
                __syncwarp();
                if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp < 2 + kEpiWarps) {
        const int ew = warp - 2, quarter = warp & 3, half = ew >> 2;
        uint8_t* stg = smem + L::kStageOffset + ew * 4096;
        int acc_stage = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            int m_blk, n_blk;
            tile_coords(tile, m_blk, n_blk);
            epilogue_tile<BN>(p, stg, tmem_base, acc_stage, m_blk, n_blk, quarter, half, lane, &tmem_full[acc_stage], acc_phase);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);
            if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
        }
    } else {
        // ===================== copier warps: the all-gather =====================
        const int cw = warp - (2 + kEpiWarps);  // 0 or 1
        uint32_t epoch = g.epoch[blockIdx.x];
        if (cw == 0) gather_rank_barrier(g, epoch + 1, lane);  // every peer's shard is final
        asm volatile("bar.sync 1, 64;" ::: "memory");           // the two copier warps
        const size_t n16 = g.shard_bytes >> 4;
        const size_t tid = (size_t)blockIdx.x * 64 + cw * 32 + lane, nthreads = (size_t)gridDim.x * 64;
        for (int r = 0; r < g.world; ++r) {
            int s = g.rank + r;
            if (s >= g.world) s -= g.world;
            const char* src = g.shard[s];
            char* dst = g.full + (size_t)s * g.shard_bytes;
            for (size_t v = tid; v < n16; v += nthreads * 4) {
                uint4 t[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (v + u * nthreads < n16) t[u] = ld_weak_noalloc_v4(src + (v + u * nthreads) * 16);
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (v + u * nthreads < n16) st_global_v4(dst + (v + u * nthreads) * 16, t[u]);
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(g.flags + s) : "memory");
        }
        asm volatile("bar.sync 1, 64;" ::: "memory");
        if (cw == 0) {
            gather_rank_barrier(g, epoch + 2, lane);  // peers may now overwrite their shards
            if (lane == 0) g.epoch[blockIdx.x] = epoch + 2;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        // last CTA out resets the ready counters for the next launch (CUDA-graph replayable)
        const uint32_t old = atomicAdd(g.flags + DDL_MAX_WORLD, 1u);
        if (old == gridDim.x - 1) {
            for (int i = 0; i <= DDL_MAX_WORLD; ++i) g.flags[i] = 0;
            __threadfence();
        }
    }
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// fp32 split-K accumulator -> bf16.  flags bit 0: accumulate into the existing bf16 output; bit 1: write zeros back to the
// accumulator after reading it, so that a cached accumulator is ready for the next split-K GEMM without a memset kernel
__global__ void f32_to_bf16_kernel(float* __restrict__ in, bf16* __restrict__ out, size_t n8, int flags) {
    const bool accumulate = flags & 1, rezero = flags & 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        float f[8];
        *reinterpret_cast<float4*>(&f[0]) = *reinterpret_cast<const float4*>(in + i * 8);
        *reinterpret_cast<float4*>(&f[4]) = *reinterpret_cast<const float4*>(in + i * 8 + 4);
        if (rezero) {
            *reinterpret_cast<float4*>(in + i * 8) = make_float4(0.f, 0.f, 0.f, 0.f);
            *reinterpret_cast<float4*>(in + i * 8 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (accumulate) {
            float g[8];
            bf16x8_to_float(ld_global_v4(out + i * 8), g);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += g[j];
        }
        st_global_v4(out + i * 8, float_to_bf16x8(f));
    }
}

// out = dy * gelu'(pre), 8 elements per thread (the stand-alone form of the DDL_EPI_DGELU epilogue)
__global__ void dgelu_mul_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ pre, bf16* __restrict__ out, size_t n8) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        float g[8], p[8];
        bf16x8_to_float(ld_stream_v4(dy + i * 8), g);
        bf16x8_to_float(ld_stream_v4(pre + i * 8), p);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] *= dgelu_f(p[j]);
        st_global_v4(out + i * 8, float_to_bf16x8(g));
    }
}

// column sums of a bf16 [rows, cols] matrix (bias gradient): out[c] = sum_r x[r, c]
// block = 32 column-vectors (512 contiguous bytes per warp) x 8 rows; grid.y splits the rows
__global__ void __launch_bounds__(256) colsum_partial_kernel(const bf16* __restrict__ x, int rows, int cols, long long ld,
                                                             float* __restrict__ partials) {
    __shared__ float sh[8][32 * 8 + 1];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int vc = blockIdx.x * 32 + tx;
    const bool ok = vc * 8 < cols;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ok) {
        for (int r = blockIdx.y * 8 + ty; r < rows; r += gridDim.y * 8 * 2) {
            uint4 u0 = ld_stream_v4(x + (size_t)r * ld + vc * 8), u1;
            const int r1 = r + gridDim.y * 8;
            const bool has1 = r1 < rows;
            if (has1) u1 = ld_stream_v4(x + (size_t)r1 * ld + vc * 8);
            float f[8];
            bf16x8_to_float(u0, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) s[j] += f[j];
            if (has1) {
                bf16x8_to_float(u1, f);
#pragma unroll
                for (int j = 0; j < 8; ++j) s[j] += f[j];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[ty][tx * 8 + j] = s[j];
    __syncthreads();
    const int col = threadIdx.x;  // 0..255 = (tx, j)
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += sh[w][col];
    const int c = blockIdx.x * 256 + col;
    if (c < cols) partials[(size_t)blockIdx.y * cols + c] = a;
}
// grid = ceil(cols / 32), block = 256: lane = column (coalesced 128-byte rows), 8 warps fold `nparts` rows two at a time
__global__ void __launch_bounds__(256) colsum_final_kernel(const float* __restrict__ partials, int nparts, int cols, bf16* __restrict__ out) {
    __shared__ float sh[8][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float a0 = 0.f, a1 = 0.f;
    if (c < cols) {
        int i = w;
        for (; i + 8 < nparts; i += 16) {
            a0 += partials[(size_t)i * cols + c];
            a1 += partials[(size_t)(i + 8) * cols + c];
        }
        if (i < nparts) a0 += partials[(size_t)i * cols + c];
    }
    sh[w][lane] = a0 + a1;
    __syncthreads();
    if (w != 0 || c >= cols) return;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += sh[k][lane];
    out[c] = __float2bfloat16(acc);
}

// ---------------------------------------------------------------------------- host side
// 2-D bf16 row-major tensor [rows, cols] with pitch `ld` elements; box = [box_rows, 64 cols], 128B swizzle
int make_map(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
    DDL_REQUIRE((ld * 2) % 16 == 0, "gemm: operand pitch (%lld elements) must be a multiple of 8", ld);
    return tma_host::encode_2d(map, ptr, rows, cols, ld * 2, box_rows, "gemm");
}

template <int BN, bool A_MN, bool B_MN, int kCluster>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tbh, const GemmParams& p, int grid, cudaStream_t stream) {
    constexpr int kStages = BN == 256 ? 4 : 6;
    using L = SmemLayout<BN, A_MN, B_MN, kStages>;
    auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, kStages, kCluster>;
    static bool attr_set = false;
    if (!attr_set) {
        DDL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = L::kTotal;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DDL_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, tbh, p));
    return 0;
}

template <bool A_MN, bool B_MN>
int launch_2sm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t stream) {
    constexpr int kSmem = 6 * (BM * BK * 2 + 128 * BK * 2) + kEpiWarps * 4096 + 256 + 1024;
    auto kern = gemm2sm_bf16_kernel<A_MN, B_MN>;
    static bool attr_set = false;
    if (!attr_set) {
        DDL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DDL_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
    return 0;
}

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
// hung the GPU.  The CTA-pair kernel (cp.async.bulk.tensor ... cta_group::2) and the cluster kernel (multicast::cluster)
// are therefore only chosen when both operands are allocations of the current device; anything else (a peer mapping, an
// IPC import, host memory) takes the one-CTA kernel, whose loads are plain unicast TMA.
bool operand_is_local(const void* ptr) {
    cudaPointerAttributes attr;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeDevice && attr.device == dev;
}

int g_sm_count = 0;
int g_two_sm = 1;  // use the cta_group::2 kernel for BN = 256 problems with >= 2 row blocks (measured fastest)
int g_force_cluster = 0;  // 0 = auto, 1 or 2 = forced (benchmarks)

}  // namespace

DDL_API int ddl_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb, void* C, long long ldc, int M,
                          int N, int K, const void* bias, const void* residual, void* aux, int epilogue, int split_k, float* acc_f32,
                          int block_n, void* stream_) {
    return ddl_gemm_bf16_ex(A, a_mn, lda, B, b_mn, ldb, C, ldc, M, N, K, bias, residual, aux, epilogue, split_k, acc_f32, nullptr, block_n,
                            stream_);
}

DDL_API int ddl_gemm_bf16_ex(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb, void* C, long long ldc, int M,
                             int N, int K, const void* bias, const void* residual, void* aux, int epilogue, int split_k, float* acc_f32,
                             float* col_stats, int block_n, void* stream_) {
    DDL_REQUIRE(col_stats == nullptr || split_k == 1, "gemm: column statistics need split_k == 1");
    DDL_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
    DDL_REQUIRE(N % 8 == 0 && K % 8 == 0, "gemm: N (%d) and K (%d) must be multiples of 8", N, K);
    DDL_REQUIRE(!a_mn || M % 8 == 0, "gemm: MN-major A needs M %% 8 == 0");
    DDL_REQUIRE(ldc % 8 == 0, "gemm: ldc must be a multiple of 8");
    DDL_REQUIRE(split_k >= 1, "gemm: split_k must be >= 1");
    DDL_REQUIRE(split_k == 1 || (acc_f32 != nullptr && epilogue == DDL_EPI_NONE && !bias && !residual),
                "gemm: split-K writes raw fp32 partial sums (no epilogue) into acc_f32");
    DDL_REQUIRE(split_k > 1 || C != nullptr, "gemm: null output");
    DDL_REQUIRE(epilogue != DDL_EPI_DGELU || aux != nullptr, "gemm: dGELU epilogue needs the pre-activation in aux");
    if (g_sm_count == 0) {
        int dev = 0;
        DDL_CHECK(cudaGetDevice(&dev));
        DDL_CHECK(cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const int BN = block_n == 128 ? 128 : (block_n == 256 ? 256 : (N % 256 == 0 || N > 1024 ? 256 : 128));
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.C = (bf16*)C; p.ldc = ldc;
    p.bias = (const bf16*)bias; p.residual = (const bf16*)residual; p.aux = (bf16*)aux;
    p.acc_f32 = acc_f32; p.epilogue = epilogue; p.split_k = split_k;
    p.col_stats = col_stats;
    p.num_m_blocks = (M + BM - 1) / BM;
    p.num_n_blocks = (N + BN - 1) / BN;
    p.k_blocks_total = (K + BK - 1) / BK;
    DDL_REQUIRE(split_k <= p.k_blocks_total, "gemm: split_k %d exceeds the number of K blocks %d", split_k, p.k_blocks_total);
    {
        // no empty K slices: an empty slice would publish an accumulator nobody wrote
        const int per = (p.k_blocks_total + split_k - 1) / split_k;
        p.split_k = (p.k_blocks_total + per - 1) / per;
    }
    CUtensorMap ta, tb;
    // K-major operand [R=M|N, K]: box 64 (K) x BM|BN rows.  MN-major operand stored [K, M|N]: box 64 (MN) x 64 (K) rows.
    if (int e = a_mn ? make_map(&ta, A, K, M, lda, BK) : make_map(&ta, A, M, K, lda, BM)) return e;
    if (int e = b_mn ? make_map(&tb, B, K, N, ldb, BK) : make_map(&tb, B, N, K, ldb, BN)) return e;
    const bool local_ops = operand_is_local(A) && operand_is_local(B);
    if (g_two_sm && BN == 256 && p.num_m_blocks >= 2 && local_ops) {
        // CTA pairs: 256 x 256 tiles, half-height B boxes (K-major) / 64 x 64 boxes (MN-major)
        CUtensorMap tb2 = tb;
        if (!b_mn) {
            if (int e = make_map(&tb2, B, N, K, ldb, 128)) return e;
        }
        const int m_pairs = (p.num_m_blocks + 1) / 2;
        const int pair_tiles = m_pairs * p.num_n_blocks * p.split_k;
        const int grid2 = pair_tiles * 2 < g_sm_count ? pair_tiles * 2 : (g_sm_count / 2) * 2;
        cudaStream_t s2 = (cudaStream_t)stream_;
        if (!a_mn && !b_mn) return launch_2sm<false, false>(ta, tb2, p, grid2, s2);
        if (!a_mn && b_mn) return launch_2sm<false, true>(ta, tb2, p, grid2, s2);
        if (a_mn && b_mn) return launch_2sm<true, true>(ta, tb2, p, grid2, s2);
        return launch_2sm<true, false>(ta, tb2, p, grid2, s2);
    }
    // B shared by a 2-CTA cluster (multicast) whenever there are at least two row blocks
    const int cluster = (g_force_cluster == 2 && p.num_m_blocks >= 2 && local_ops) ? 2 : 1;  // multicast sharing measured no faster than 1 CTA
    CUtensorMap tbh = tb;
    if (cluster == 2 && !b_mn) {
        if (int e = make_map(&tbh, B, N, K, ldb, BN / 2)) return e;  // half-height box for the K-major B tile
    }
    const int m_groups = (p.num_m_blocks + cluster - 1) / cluster;
    const int tiles = m_groups * p.num_n_blocks * p.split_k;  // per cluster
    int grid = tiles * cluster < g_sm_count ? tiles * cluster : (g_sm_count / cluster) * cluster;
    cudaStream_t s = (cudaStream_t)stream_;
#define GO(BN_, AM, BMN)                                                         \
    do {                                                                         \
        if (cluster == 2) return launch<BN_, AM, BMN, 2>(ta, tb, tbh, p, grid, s); \
        return launch<BN_, AM, BMN, 1>(ta, tb, tbh, p, grid, s);                   \
    } while (0)
    if (BN == 256) {
        if (!a_mn && !b_mn) GO(256, false, false);
        if (!a_mn && b_mn) GO(256, false, true);
        if (a_mn && b_mn) GO(256, true, true);
        GO(256, true, false);
    } else {
        if (!a_mn && !b_mn) GO(128, false, false);
        if (!a_mn && b_mn) GO(128, false, true);
        if (a_mn && b_mn) GO(128, true, true);
        GO(128, true, false);
    }
#undef GO
}

DDL_API int ddl_allgather_gemm_bf16(const ddl_comm_t* comm, void* const* shard_ptrs, void* full_w, void* flags, const void* A, long long lda,
                                    void* C, long long ldc, int M, int N, int K, const void* bias, const void* residual, void* aux,
                                    int epilogue, void* stream_) {
    DDL_REQUIRE(comm && shard_ptrs && full_w && flags && A && C, "allgather_gemm: null pointer");
    DDL_REQUIRE(comm->world >= 1 && comm->world <= DDL_MAX_WORLD, "allgather_gemm: bad world size");
    DDL_REQUIRE(N % (comm->world * 128) == 0, "allgather_gemm: N (%d) must be a multiple of world * 128", N);
    DDL_REQUIRE(K % 8 == 0 && ldc % 8 == 0, "allgather_gemm: K and ldc must be multiples of 8");
    if (g_sm_count == 0) {
        int dev = 0;
        DDL_CHECK(cudaGetDevice(&dev));
        DDL_CHECK(cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.C = (bf16*)C; p.ldc = ldc;
    p.bias = (const bf16*)bias; p.residual = (const bf16*)residual; p.aux = (bf16*)aux;
    p.acc_f32 = nullptr; p.col_stats = nullptr; p.epilogue = epilogue; p.split_k = 1;
    p.num_m_blocks = (M + BM - 1) / BM;
    p.num_n_blocks = N / 128;
    p.k_blocks_total = (K + BK - 1) / BK;
    GatherParams g;
    for (int i = 0; i < DDL_MAX_WORLD; ++i) { g.shard[i] = nullptr; g.pad[i] = nullptr; }
    for (int i = 0; i < comm->world; ++i) {
        DDL_REQUIRE(shard_ptrs[i] && comm->signal_pads[i], "allgather_gemm: peer pointer %d is null", i);
        g.shard[i] = (const char*)shard_ptrs[i];
        g.pad[i] = (uint32_t*)comm->signal_pads[i];
    }
    g.full = (char*)full_w;
    g.flags = (uint32_t*)flags;
    g.epoch = (uint32_t*)comm->epoch;
    g.rank = comm->rank; g.world = comm->world;
    g.blocks_per_shard = (N / comm->world) / 128;
    g.shard_bytes = (size_t)(N / comm->world) * K * 2;
    CUtensorMap ta, tb;
    if (int e = make_map(&ta, A, M, K, lda, BM)) return e;
    if (int e = make_map(&tb, full_w, N, K, K, 128)) return e;
    using L = SmemLayout<128, false, false, 6>;
    static bool attr_set = false;
    if (!attr_set) {
        DDL_CHECK(cudaFuncSetAttribute(gemm_allgather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    // every CTA carries two copier warps, so the whole grid must be resident: one CTA per SM
    const int tiles = p.num_m_blocks * p.num_n_blocks;
    int grid = tiles < g_sm_count ? tiles : g_sm_count;
    if (grid > DDL_COMM_MAX_BLOCKS) grid = DDL_COMM_MAX_BLOCKS;
    gemm_allgather_kernel<<<grid, kGatherThreads, L::kTotal, (cudaStream_t)stream_>>>(ta, tb, p, g);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_gemm_set_cluster(int cluster) {
    // 0 = auto (CTA pairs with tcgen05.mma.cta_group::2 on 256 x 256 tiles when the problem has >= 2 row
    //     blocks and BN = 256, else one CTA per tile), 1 = always one CTA per tile,
    // 2 = 2-CTA cluster sharing B by TMA multicast (cta_group::1 MMAs), 3 = same as 0
    DDL_REQUIRE(cluster >= 0 && cluster <= 3, "cluster mode must be 0..3");
    g_two_sm = (cluster == 0 || cluster == 3);
    g_force_cluster = cluster;
    return 0;
}

DDL_API int ddl_f32_to_bf16(const float* in, void* out, size_t numel, int flags, void* stream_) {
    DDL_REQUIRE(numel % 8 == 0, "ddl_f32_to_bf16: numel must be a multiple of 8");
    if (numel == 0) return 0;
    int blocks = ddl_ceil_div((long long)(numel / 8), 256);
    if (blocks > DDL_SM_COUNT * 8) blocks = DDL_SM_COUNT * 8;
    f32_to_bf16_kernel<<<blocks, 256, 0, (cudaStream_t)stream_>>>(const_cast<float*>(in), (bf16*)out, numel / 8, flags);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_dgelu_mul(const void* dy, const void* pre, void* out, size_t numel, void* stream_) {
    DDL_REQUIRE(numel % 8 == 0, "ddl_dgelu_mul: numel must be a multiple of 8");
    if (numel == 0) return 0;
    int blocks = ddl_ceil_div((long long)(numel / 8), 256);
    if (blocks > DDL_SM_COUNT * 8) blocks = DDL_SM_COUNT * 8;
    dgelu_mul_kernel<<<blocks, 256, 0, (cudaStream_t)stream_>>>((const bf16*)dy, (const bf16*)pre, (bf16*)out, numel / 8);
    DDL_LAUNCH_CHECK();
    return 0;
}

static int colsum_parts(int rows, int cols) {
    const int col_tiles = (cols / 8 + 31) / 32;
    int parts = (DDL_SM_COUNT * 6 + col_tiles - 1) / col_tiles;  // ~6 blocks per SM in total
    const int max_parts = (rows + 15) / 16;
    if (parts > max_parts) parts = max_parts;
    return parts < 1 ? 1 : parts;
}

DDL_API size_t ddl_colsum_scratch_bytes(int rows, int cols) { return (size_t)colsum_parts(rows, cols) * cols * sizeof(float); }

DDL_API int ddl_colsum_bf16(const void* x, int rows, int cols, long long ld, void* out, void* scratch, size_t scratch_bytes, void* stream_) {
    DDL_REQUIRE(cols % 8 == 0 && ld % 8 == 0, "ddl_colsum_bf16: cols and ld must be multiples of 8");
    DDL_REQUIRE(scratch_bytes >= ddl_colsum_scratch_bytes(rows, cols), "ddl_colsum_bf16: scratch too small");
    const int parts = colsum_parts(rows, cols);
    dim3 grid((cols / 8 + 31) / 32, parts);
    cudaStream_t s = (cudaStream_t)stream_;
    colsum_partial_kernel<<<grid, 256, 0, s>>>((const bf16*)x, rows, cols, ld, (float*)scratch);
    DDL_LAUNCH_CHECK();
    colsum_final_kernel<<<ddl_ceil_div(cols, 32), 256, 0, s>>>((const float*)scratch, parts, cols, (bf16*)out);
    DDL_LAUNCH_CHECK();
    return 0;
}
