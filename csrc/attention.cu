// Self-attention for BERT (head_dim 64, sequence 128..512, right-padding mask, dropout) on the
// 5th-generation tensor cores (K11; BASELINE.json:5 "softmax/attention").
//
// Layout: the QKV projection leaves one packed matrix  qkv[B*S, 3H]  (row = token; columns
// q|k|v, each H = heads*64 wide).  Every operand tile this kernel touches is a [128 rows x 64
// columns] bf16 box of that matrix, fetched by ONE tensor map with the 128-byte swizzle, so
// the same shared-memory tile serves as a K-major operand (reduction over head_dim) and as an
// MN-major operand (reduction over sequence positions).  Output ctx[B*S, H]; backward writes
// the packed gradient dqkv[B*S, 3H] directly (no concatenation pass).
//
// Forward, one CTA per (batch, head, 128-query tile), 128 threads (thread = query row):
//   TMA   : Q tile, K tiles, V tiles (only the tiles below the sequence's valid length)
//   MMA 1 : S = Q K^T                      tcgen05.mma SS, fp32 accumulators in TMEM [0, Skv)
//   soft  : row max / exp2 / row sum on TMEM rows (tcgen05.ld), dropout, P -> bf16 written back
//           over the S columns it came from (tcgen05.st)
//   MMA 2 : O = P V                        tcgen05.mma TS: A = P from TMEM, B = V tiles (MN-major)
//   epil. : O / rowsum -> bf16 ctx, LSE saved for backward
//
// Backward, one CTA per (batch, head, 128-key tile), looping over query tiles; everything is
// computed transposed (rows = keys) so that P^T and dS^T come out of the softmax threads as
// K-major A operands, and the very same dS^T tile read MN-major is the A operand of dQ:
//   S^T = K Q^T, dP^T = V dO^T  ->  P^T, dS^T (bf16, swizzled smem)  ->
//   dV += P^T dO, dK += dS^T Q (TMEM accumulators across query tiles), dQ = dS K (per tile;
//   written straight to dqkv when the sequence has one key tile, else red.add into fp32).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

using namespace tc;

namespace {

constexpr int D = 64;          // head dim
constexpr int TILE = 128;      // query / key tile
constexpr int kTileBytes = TILE * D * 2;  // 16 KB
constexpr float kLog2e = 1.4426950408889634f;

struct AttnParams {
    int B, S, heads, H;        // H = heads * 64
    const int* seqlens;        // [B] valid lengths (right padding) or null
    bf16* ctx;                 // [B*S, H]
    float* lse;                // [B, heads, S]  (natural log of the softmax denominator, incl. max)
    float scale;               // 1/sqrt(64)
    uint32_t drop_thr;         // byte threshold (p = thr/256), 0 = no dropout
    float drop_scale;          // 1/(1-p)
    uint64_t seed;
    const uint64_t* seed_dev;
    // backward
    const float* dvec;         // [B, heads, S] rowsum(dO * O)
    bf16* dqkv;                // [B*S, 3H]
    float* dq_acc;             // [B*S, H] fp32, only when S > 128
};

// 32-bit finaliser (two multiplies): cheap enough to run once per 4 probabilities in forward and once
// per probability in backward (where a thread owns one key and walks the queries)
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
// One 32-bit hash covers the 4 probabilities (q, 4*kv4 .. 4*kv4+3) of head `bh`: byte i decides element i.
// Forward (thread = query, walks keys) uses all four bytes; backward (thread = key, walks queries) one of them.
__device__ __forceinline__ uint32_t drop_hash(uint32_t seed32, uint32_t bh, uint32_t q, uint32_t kv4, uint32_t S) {
    const uint32_t idx = (bh * S + q) * (S >> 2) + kv4;
    return mix32(idx * 0x9E3779B1u + seed32);
}
__device__ __forceinline__ uint32_t drop_keep4(uint32_t seed32, uint32_t bh, uint32_t q, uint32_t kv4, uint32_t S, uint32_t thr) {
    const uint32_t h = drop_hash(seed32, bh, q, kv4, S);
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) m |= (((h >> (8 * i)) & 0xffu) >= thr ? 1u : 0u) << i;
    return m;
}

__device__ __forceinline__ uint64_t resolve_seed(uint64_t seed, const uint64_t* seed_dev, uint32_t thr) {
    if (thr && seed_dev) seed ^= (*seed_dev + 0x632BE59BD9B4E019ull) * 0x9E3779B97F4A7C15ull;
    return seed;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <int kMaxKvTiles>
__global__ void __launch_bounds__(kMaxKvTiles == 1 ? 128 : 256, kMaxKvTiles == 1 ? 4 : 1) attn_fwd_kernel(const __grid_constant__ CUtensorMap tma_qkv,
                                                                                  const AttnParams p) {
    constexpr int kHalves = kMaxKvTiles == 1 ? 1 : 2;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kTileBytes;
    uint8_t* sV = sK + kMaxKvTiles * kTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kMaxKvTiles * kTileBytes);  // [0] load, [1] S ready, [2] O ready
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 4);
    float* sRed = reinterpret_cast<float*>(bars + 8);  // [2][128] partner exchange (row max, row sum)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tile = blockIdx.x, hd = blockIdx.y, b = blockIdx.z;
    const int kv_len = p.seqlens ? min(max(p.seqlens[b], 1), p.S) : p.S;
    const int n_kv = (kv_len + TILE - 1) / TILE;
    // One key tile: by the time O is produced, S (128 fp32 columns) has been consumed and P (bf16 pairs) sits in
    // columns [0, 64), so O takes columns [64, 128): 128 TMEM columns per CTA instead of 256 -> four CTAs per SM
    // (their serial TMA -> MMA -> softmax -> MMA chains overlap each other).
    constexpr uint32_t kTmemCols = kMaxKvTiles == 1 ? 128 : 512;
    const uint32_t o_col = kMaxKvTiles == 1 ? 64u : ((kMaxKvTiles * TILE < 512) ? (uint32_t)(kMaxKvTiles * TILE) : 256u);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tma_qkv);
        for (int i = 0; i < 3; ++i) mbar_init(&bars[i], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int row0 = b * p.S + q_tile * TILE;
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bars[0], (uint32_t)((1 + 2 * n_kv) * kTileBytes));
        tma_load_2d(sQ, &tma_qkv, &bars[0], hd * D, row0);
        for (int j = 0; j < n_kv; ++j) {
            tma_load_2d(sK + j * kTileBytes, &tma_qkv, &bars[0], p.H + hd * D, b * p.S + j * TILE);
            tma_load_2d(sV + j * kTileBytes, &tma_qkv, &bars[0], 2 * p.H + hd * D, b * p.S + j * TILE);
        }
        // ---- MMA 1: S = Q K^T
        mbar_wait(&bars[0], 0);
        tc_fence_after();
        constexpr uint32_t idesc_s = make_idesc(TILE, TILE, false, false);
        for (int j = 0; j < n_kv; ++j) {
#pragma unroll
            for (int k = 0; k < D / 16; ++k) {
                const uint64_t da = make_smem_desc(smem_u32(sQ) + k * 32, 16, 1024);
                const uint64_t db = make_smem_desc(smem_u32(sK + j * kTileBytes) + k * 32, 16, 1024);
                umma_bf16(tmem_base + j * TILE, da, db, idesc_s, k > 0 ? 1u : 0u);
            }
        }
        umma_commit(&bars[1]);
    }

    // ---- softmax: thread = (query row, column half).  One key tile: 128 threads, a thread owns its row.
    // More key tiles: 256 threads - warps w and w+4 share a TMEM lane quarter and take alternate 32-column
    // chunks, which halves the exp / hash work per thread; row max and row sum are combined through smem.
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float sl2 = p.scale * kLog2e;
    const uint64_t seed64 = resolve_seed(p.seed, p.seed_dev, p.drop_thr);
    const uint32_t seed = mix32((uint32_t)seed64) ^ (uint32_t)(seed64 >> 32);
    auto pair_sync = [&]() {  // the two warps of one lane quarter
        if (kHalves == 2) asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory");
    };
    mbar_wait(&bars[1], 0);
    __syncwarp();
    tc_fence_after();
    const int n_chunks = n_kv * (TILE / 32);
    float mx = -INFINITY;
    for (int c = half; c < n_chunks; c += kHalves) {
        uint32_t v[32];
        tmem_ld_32x32(lane_addr + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (c * 32 + j < kv_len) mx = fmaxf(mx, __uint_as_float(v[j]));
    }
    if (kHalves == 2) {
        sRed[half * TILE + r] = mx;
        pair_sync();
        mx = fmaxf(mx, sRed[(half ^ 1) * TILE + r]);
        pair_sync();
    }
    const float mxs = mx * sl2;
    float sum = 0.f;
    const uint32_t bh = (uint32_t)(b * p.heads + hd);
    const uint32_t q_idx = (uint32_t)(q_tile * TILE + r);
    for (int c = half; c < n_chunks; c += kHalves) {
        uint32_t v[32];
        tmem_ld_32x32(lane_addr + c * 32, v);
        tmem_ld_wait();
        // P (bf16 pairs) of chunk c overwrites S columns [16c, 16c+16), i.e. part of chunk c/2.  With one thread
        // per row that chunk was consumed earlier; with two, both partners have loaded this step's chunks before
        // either stores (and chunk c/2 belongs to an earlier step for c >= 2).
        pair_sync();
        uint32_t packed[16];
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
            float e[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int col = c * 32 + j4 * 4 + t;
                e[t] = col < kv_len ? exp2f(__uint_as_float(v[j4 * 4 + t]) * sl2 - mxs) : 0.f;
                sum += e[t];
            }
            if (p.drop_thr) {
                const uint32_t keep = drop_keep4(seed, bh, q_idx, (uint32_t)(c * 8 + j4), (uint32_t)p.S, p.drop_thr);
#pragma unroll
                for (int t = 0; t < 4; ++t) e[t] = ((keep >> t) & 1u) ? e[t] : 0.f;
            }
            packed[j4 * 2] = pack_bf16x2(e[0], e[1]);
            packed[j4 * 2 + 1] = pack_bf16x2(e[2], e[3]);
        }
        tmem_st_32x16(lane_addr + c * 16, packed);
    }
    if (kHalves == 2) {
        sRed[half * TILE + r] = sum;
        pair_sync();
        sum += sRed[(half ^ 1) * TILE + r];
    }
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();

    if (threadIdx.x == 0) {
        tc_fence_after();
        // ---- MMA 2: O = P V   (A from TMEM, B = V tile read MN-major: N = head_dim contiguous, K = key rows)
        constexpr uint32_t idesc_o = make_idesc(TILE, D, false, true);
        for (int j = 0; j < n_kv; ++j) {
#pragma unroll
            for (int k = 0; k < TILE / 16; ++k) {
                const uint64_t db = make_smem_desc(smem_u32(sV + j * kTileBytes) + k * 2048, 8192, 1024);
                umma_bf16_ts(tmem_base + o_col, tmem_base + j * (TILE / 2) + k * 8, db, idesc_o, (j > 0 || k > 0) ? 1u : 0u);
            }
        }
        umma_commit(&bars[2]);
    }
    mbar_wait(&bars[2], 0);
    __syncwarp();
    tc_fence_after();
    {
        const float inv = (p.drop_thr ? p.drop_scale : 1.f) / sum;
        bf16* out = p.ctx + (size_t)(row0 + r) * p.H + hd * D;
#pragma unroll
        for (int c = half * (2 / kHalves); c < (half + 1) * (2 / kHalves); ++c) {
            uint32_t v[32];
            tmem_ld_32x32(lane_addr + o_col + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                float o[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) o[t] = __uint_as_float(v[j + t]) * inv;
                st_global_v4(out + c * 32 + j, float_to_bf16x8(o));
            }
        }
        if (p.lse && half == 0) p.lse[((size_t)b * p.heads + hd) * p.S + q_tile * TILE + r] = mx * p.scale + __logf(sum);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------
// forward, streaming variant for S > 128: one CTA per (batch, head, 128-query tile), 256 threads, key / value tiles
// streamed through a two-stage TMA ring with an online softmax, so that a CTA needs 80 KB of shared memory and 256
// TMEM columns (S [0,128) fp32, P [128,192) bf16 pairs, O [192,256) fp32) and TWO CTAs fit per SM - their serial
// TMA -> MMA -> softmax -> MMA chains overlap each other (the resident variant above holds all of S in 512 columns
// and runs one CTA per SM).  The running maximum is only raised when a tile exceeds it by more than 8 in the log2
// domain (P then stays <= 256, exact in the final O / l), so the O accumulator is rarely read back and rescaled.
__global__ void __launch_bounds__(256, 2) attn_fwd_stream_kernel(const __grid_constant__ CUtensorMap tma_qkv, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sKV = sQ + kTileBytes;  // [2 stages][K tile, V tile]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + 4 * kTileBytes);  // [0] Q  [1],[2] K/V stage  [3] S ready  [4] PV done
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 8);
    float* sRed = reinterpret_cast<float*>(bars + 10);  // [2][128] partner exchange

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tile = blockIdx.x, hd = blockIdx.y, b = blockIdx.z;
    const int kv_len = p.seqlens ? min(max(p.seqlens[b], 1), p.S) : p.S;
    const int n_kv = (kv_len + TILE - 1) / TILE;
    constexpr uint32_t kTmemCols = 256, cS = 0, cP = 128, cO = 192;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tma_qkv);
        for (int i = 0; i < 5; ++i) mbar_init(&bars[i], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int row0 = b * p.S + q_tile * TILE;
    auto load_kv = [&](int j) {  // thread 0
        uint8_t* st = sKV + (j & 1) * 2 * kTileBytes;
        mbar_expect_tx(&bars[1 + (j & 1)], 2 * kTileBytes);
        tma_load_2d(st, &tma_qkv, &bars[1 + (j & 1)], p.H + hd * D, b * p.S + j * TILE);
        tma_load_2d(st + kTileBytes, &tma_qkv, &bars[1 + (j & 1)], 2 * p.H + hd * D, b * p.S + j * TILE);
    };
    auto issue_scores = [&](int j) {  // thread 0: S = Q K_j^T
        mbar_wait(&bars[1 + (j & 1)], (uint32_t)((j >> 1) & 1));
        tc_fence_after();
        constexpr uint32_t idesc_s = make_idesc(TILE, TILE, false, false);
        const uint32_t k_s = smem_u32(sKV + (j & 1) * 2 * kTileBytes);
#pragma unroll
        for (int k = 0; k < D / 16; ++k)
            umma_bf16(tmem_base + cS, make_smem_desc(smem_u32(sQ) + k * 32, 16, 1024), make_smem_desc(k_s + k * 32, 16, 1024), idesc_s,
                      k > 0 ? 1u : 0u);
        umma_commit(&bars[3]);
    };
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bars[0], kTileBytes);
        tma_load_2d(sQ, &tma_qkv, &bars[0], hd * D, row0);
        load_kv(0);
        if (n_kv > 1) load_kv(1);
        mbar_wait(&bars[0], 0);
        issue_scores(0);
    }

    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const float sl2 = p.scale * kLog2e;
    const uint64_t seed64 = resolve_seed(p.seed, p.seed_dev, p.drop_thr);
    const uint32_t seed = mix32((uint32_t)seed64) ^ (uint32_t)(seed64 >> 32);
    const uint32_t bh = (uint32_t)(b * p.heads + hd);
    const uint32_t q_idx = (uint32_t)(q_tile * TILE + r);
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory"); };

    float m_used = -INFINITY;  // running maximum (log2 domain, i.e. already multiplied by scale * log2 e)
    float l_part = 0.f;        // this thread's share of the row sum, relative to m_used
    for (int j = 0; j < n_kv; ++j) {
        mbar_wait(&bars[3], (uint32_t)(j & 1));
        __syncwarp();
        tc_fence_after();
        // ---- tile maximum over this thread's two 32-column chunks
        float mt = -INFINITY;
#pragma unroll
        for (int ci = 0; ci < 2; ++ci) {
            const int c = half + 2 * ci;
            uint32_t v[32];
            tmem_ld_32x32(lane_addr + cS + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int t = 0; t < 32; ++t)
                if (j * TILE + c * 32 + t < kv_len) mt = fmaxf(mt, __uint_as_float(v[t]));
        }
        sRed[half * TILE + r] = mt;
        pair_sync();
        mt = fmaxf(mt, sRed[(half ^ 1) * TILE + r]) * sl2;
        pair_sync();
        // raise the running maximum only when this tile exceeds it by more than 8 (P stays <= 2^8)
        const bool raise = mt > m_used + 8.f;
        const float m_new = raise ? mt : m_used;
        const float alpha = raise ? exp2f(m_used - m_new) : 1.f;  // exp2f(-inf) = 0 on the first tile
        m_used = m_new;
        l_part *= alpha;
        if (j > 0) {
            mbar_wait(&bars[4], (uint32_t)((j - 1) & 1));  // P columns free, O complete up to tile j-1
            __syncwarp();
            tc_fence_after();
            if (__any_sync(0xffffffffu, raise)) {
                uint32_t o[32];
                tmem_ld_32x32(lane_addr + cO + half * 32, o);
                tmem_ld_wait();
#pragma unroll
                for (int t = 0; t < 32; ++t) o[t] = __float_as_uint(__uint_as_float(o[t]) * alpha);
                // 32 columns as two 16-column stores
                uint32_t lo[16], hi[16];
#pragma unroll
                for (int t = 0; t < 16; ++t) {
                    lo[t] = o[t];
                    hi[t] = o[16 + t];
                }
                tmem_st_32x16(lane_addr + cO + half * 32, lo);
                tmem_st_32x16(lane_addr + cO + half * 32 + 16, hi);
            }
        }
        // ---- P = exp2(s - m), dropout, bf16 pairs into the P columns
#pragma unroll
        for (int ci = 0; ci < 2; ++ci) {
            const int c = half + 2 * ci;
            uint32_t v[32];
            tmem_ld_32x32(lane_addr + cS + c * 32, v);
            tmem_ld_wait();
            uint32_t packed[16];
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
                float e[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int col = j * TILE + c * 32 + j4 * 4 + t;
                    e[t] = col < kv_len ? exp2f(__uint_as_float(v[j4 * 4 + t]) * sl2 - m_used) : 0.f;
                    l_part += e[t];
                }
                if (p.drop_thr) {
                    const uint32_t keep = drop_keep4(seed, bh, q_idx, (uint32_t)(j * (TILE / 4) + c * 8 + j4), (uint32_t)p.S, p.drop_thr);
#pragma unroll
                    for (int t = 0; t < 4; ++t) e[t] = ((keep >> t) & 1u) ? e[t] : 0.f;
                }
                packed[j4 * 2] = pack_bf16x2(e[0], e[1]);
                packed[j4 * 2 + 1] = pack_bf16x2(e[2], e[3]);
            }
            tmem_st_32x16(lane_addr + cP + c * 16, packed);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncthreads();  // S consumed, P (and the rescaled O) written by everybody
        if (threadIdx.x == 0) {
            tc_fence_after();
            // O (+)= P V_j : A = P from TMEM, B = V tile read MN-major
            constexpr uint32_t idesc_o = make_idesc(TILE, D, false, true);
            const uint32_t v_s = smem_u32(sKV + (j & 1) * 2 * kTileBytes + kTileBytes);
#pragma unroll
            for (int k = 0; k < TILE / 16; ++k)
                umma_bf16_ts(tmem_base + cO, tmem_base + cP + k * 8, make_smem_desc(v_s + k * 2048, 8192, 1024), idesc_o,
                             (j > 0 || k > 0) ? 1u : 0u);
            umma_commit(&bars[4]);
            if (j + 1 < n_kv) issue_scores(j + 1);  // runs behind PV_j
            if (j + 2 < n_kv) {
                mbar_wait(&bars[4], (uint32_t)(j & 1));  // stage j&1 (K_j, V_j) no longer read
                load_kv(j + 2);
            }
        }
    }
    // ---- epilogue
    sRed[half * TILE + r] = l_part;
    pair_sync();
    const float l = l_part + sRed[(half ^ 1) * TILE + r];
    mbar_wait(&bars[4], (uint32_t)((n_kv - 1) & 1));
    __syncwarp();
    tc_fence_after();
    {
        const float inv = (p.drop_thr ? p.drop_scale : 1.f) / l;
        bf16* out = p.ctx + (size_t)(row0 + r) * p.H + hd * D;
        uint32_t v[32];
        tmem_ld_32x32(lane_addr + cO + half * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int t0 = 0; t0 < 32; t0 += 8) {
            float o[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) o[t] = __uint_as_float(v[t0 + t]) * inv;
            st_global_v4(out + half * 32 + t0, float_to_bf16x8(o));
        }
        // natural-log LSE: m_used is in the log2 domain of the scaled scores
        if (p.lse && half == 0) p.lse[((size_t)b * p.heads + hd) * p.S + q_tile * TILE + r] = m_used * (1.f / kLog2e) + __logf(l);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// D[b, h, s] = sum_d dO[b, s, h, d] * O[b, s, h, d]
__global__ void attn_dvec_kernel(const bf16* __restrict__ dO, const bf16* __restrict__ O, float* __restrict__ dvec, int B, int S, int heads) {
    const size_t total = (size_t)B * S * heads;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int hd = (int)(t % heads);
    const size_t tok = t / heads;
    const int s = (int)(tok % S), b = (int)(tok / S);
    const bf16* a = dO + tok * heads * D + hd * D;
    const bf16* o = O + tok * heads * D + hd * D;
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < D; j += 8) {
        float x[8], y[8];
        bf16x8_to_float(ld_global_v4(a + j), x);
        bf16x8_to_float(ld_global_v4(o + j), y);
#pragma unroll
        for (int q = 0; q < 8; ++q) acc += x[q] * y[q];
    }
    dvec[((size_t)b * heads + hd) * S + s] = acc;
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
// write 8 bf16 (one 16-byte chunk) of row `row`, logical chunk `chunk` (0..15 across 128 columns)
// into a [128 x 128] bf16 tile stored as two [128 x 64] K-major 128B-swizzled blocks
__device__ __forceinline__ void st_swizzled_chunk(uint8_t* tile, int row, int chunk, uint4 v) {
    const int blk = chunk >> 3, cc = chunk & 7;
    uint8_t* dst = tile + blk * kTileBytes + row * 128 + ((cc ^ (row & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = v;
}

// kSingle (S == 128): one query tile, nothing accumulates across iterations, so the gradient
// accumulators reuse consumed S^T / dP^T columns (256 TMEM columns, 96 KB smem -> two CTAs per SM).
// Otherwise: dV / dK accumulate over the query tiles in their own columns, and the Q / dO tiles of
// the next iteration are prefetched by TMA while this one is in its softmax phase.
template <bool kSingle>
__global__ void __launch_bounds__(kSingle ? 128 : 256, kSingle ? 2 : 1) attn_bwd_kernel(const __grid_constant__ CUtensorMap tma_qkv,
                                                                        const __grid_constant__ CUtensorMap tma_do, const AttnParams p) {
    constexpr int kBufs = kSingle ? 1 : 2;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sK = smem;                          // [128 kv x 64]
    uint8_t* sV = sK + kTileBytes;
    uint8_t* sQ = sV + kTileBytes;               // [kBufs][128 q x 64]
    uint8_t* sdO = sQ + kBufs * kTileBytes;      // [kBufs]
    uint8_t* sdSt = sdO + kBufs * kTileBytes;    // dS^T [128 kv x 128 q], two swizzled 64-column blocks (A of dQ, read MN-major)
    float* sLse = reinterpret_cast<float*>(sdSt + 2 * kTileBytes);  // [128]
    float* sDv = sLse + TILE;                                       // [128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDv + TILE);
    // bars: [0] K/V  [1],[2] Q/dO buffer 0/1  [3] S^T,dP^T ready  [4] dV done (P^T columns free)  [5] all gradients of the tile ready
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 8);

    // thread -> (key row, column half).  kSingle: 128 threads, one per row.  Otherwise 256 threads: warps w and
    // w+4 share a TMEM lane quarter and split the query columns, which halves the exp/hash work per thread.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quarter = warp & 3, half = warp >> 2;
    constexpr int kHalves = kSingle ? 1 : 2;
    constexpr int kChunksPerThread = (TILE / 32) / kHalves;
    const int kv_tile = blockIdx.x, hd = blockIdx.y, b = blockIdx.z;
    const int kv_len = p.seqlens ? min(max(p.seqlens[b], 1), p.S) : p.S;
    const int n_q = p.S / TILE;
    // TMEM columns.  S^T fp32 [0,128) is overwritten in place by P^T bf16 [0,64); dP^T fp32 [128,256) by dS^T bf16 [128,192).
    constexpr uint32_t kTmemCols = kSingle ? 256 : 512;
    // With 512 columns (not kSingle) the bf16 copies get their own columns [384,448) / [448,512): two warps work on
    // one lane quarter concurrently there, and in-place aliasing would let one overwrite what the other has not read.
    // Multi-tile variant: dS^T is NOT copied to TMEM; dK reads it from the swizzled smem tile (K-major A operand), which
    // leaves [448,512) for dQ.  S^T / dP^T are then dead as soon as the softmax threads have read them, so the MMAs of the
    // NEXT query tile are issued right behind this tile's gradient MMAs and run while dQ is being written out.
    constexpr uint32_t cST = 0, cDP = 128, cDQ = kSingle ? 0 : 448;
    constexpr uint32_t cDV = kSingle ? 64 : 256, cDK = kSingle ? 192 : 320;
    constexpr uint32_t cPT = kSingle ? cST : 384, cDST = cDP;  // cDST only used by kSingle

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tma_qkv);
        tma_prefetch_desc(&tma_do);
        for (int i = 0; i < 6; ++i) mbar_init(&bars[i], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);

    const int kv_row0 = b * p.S + kv_tile * TILE;
    const int r = quarter * 32 + lane;              // this thread's key row inside the tile
    const int kv_idx = kv_tile * TILE + r;
    const bool kv_valid = kv_idx < kv_len;
    const bool tile_empty = kv_tile * TILE >= kv_len;  // fully padded key tile: gradients are zero
    const float sl2 = p.scale * kLog2e;
    const uint64_t seed64 = resolve_seed(p.seed, p.seed_dev, p.drop_thr);
    const uint32_t seed = mix32((uint32_t)seed64) ^ (uint32_t)(seed64 >> 32);
    const uint32_t bh = (uint32_t)(b * p.heads + hd);

    if (!tile_empty) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(&bars[0], 2 * kTileBytes);
            tma_load_2d(sK, &tma_qkv, &bars[0], p.H + hd * D, kv_row0);
            tma_load_2d(sV, &tma_qkv, &bars[0], 2 * p.H + hd * D, kv_row0);
            mbar_expect_tx(&bars[1], 2 * kTileBytes);
            tma_load_2d(sQ, &tma_qkv, &bars[1], hd * D, b * p.S);
            tma_load_2d(sdO, &tma_do, &bars[1], hd * D, b * p.S);
        }
        uint32_t ph = 0;            // phase of the once-per-iteration barriers [3], [4], [5]
        uint32_t qph[2] = {0, 0};   // phase of each Q/dO buffer's barrier
        // S^T = K Q^T and dP^T = V dO^T of the query tile in buffer `bf` (thread 0 only)
        auto issue_scores = [&](int bf, uint32_t phase) {
            mbar_wait(&bars[1 + bf], phase);
            tc_fence_after();
            constexpr uint32_t idesc = make_idesc(TILE, TILE, false, false);
            const uint32_t q_s = smem_u32(sQ + bf * kTileBytes), do_s = smem_u32(sdO + bf * kTileBytes);
#pragma unroll
            for (int k = 0; k < D / 16; ++k)
                umma_bf16(tmem_base + cST, make_smem_desc(smem_u32(sK) + k * 32, 16, 1024), make_smem_desc(q_s + k * 32, 16, 1024), idesc,
                          k > 0 ? 1u : 0u);
#pragma unroll
            for (int k = 0; k < D / 16; ++k)
                umma_bf16(tmem_base + cDP, make_smem_desc(smem_u32(sV) + k * 32, 16, 1024), make_smem_desc(do_s + k * 32, 16, 1024), idesc,
                          k > 0 ? 1u : 0u);
            umma_commit(&bars[3]);
        };
        for (int qi = 0; qi < n_q; ++qi) {
            const int buf = kSingle ? 0 : (qi & 1);
            const int q_row0 = b * p.S + qi * TILE;
            uint8_t* cQ = sQ + buf * kTileBytes;
            uint8_t* cdO = sdO + buf * kTileBytes;
            if (!kSingle && threadIdx.x == 0 && qi + 1 < n_q) {
                // the other buffer was last read by iteration qi-1's MMAs, which have completed
                const int nb = buf ^ 1;
                mbar_expect_tx(&bars[1 + nb], 2 * kTileBytes);
                tma_load_2d(sQ + nb * kTileBytes, &tma_qkv, &bars[1 + nb], hd * D, q_row0 + TILE);
                tma_load_2d(sdO + nb * kTileBytes, &tma_do, &bars[1 + nb], hd * D, q_row0 + TILE);
            }
            if (half == 0) {
                sLse[r] = p.lse[((size_t)b * p.heads + hd) * p.S + qi * TILE + r] * kLog2e;
                sDv[r] = p.dvec[((size_t)b * p.heads + hd) * p.S + qi * TILE + r];
            }
            __syncthreads();
            if (threadIdx.x == 0 && (kSingle || qi == 0)) {
                mbar_wait(&bars[0], 0);
                issue_scores(buf, qph[buf]);
            }
            qph[buf] ^= 1;
            mbar_wait(&bars[3], ph);
            __syncwarp();
            tc_fence_after();
            // ---- softmax backward on row `r` (a key); columns = the queries of this tile
#pragma unroll 1
            for (int c = half * kChunksPerThread; c < (half + 1) * kChunksPerThread; ++c) {
                uint32_t s[32], dp[32];
                tmem_ld_32x32(lane_addr + cST + c * 32, s);
                tmem_ld_32x32(lane_addr + cDP + c * 32, dp);
                tmem_ld_wait();
                uint32_t ppk[16], dpk[16];
                // dropout: the 4 lanes that own keys 4k .. 4k+3 need the SAME 32-bit hash per query (one byte each), so each
                // computes it for 8 of the chunk's 32 queries and the group exchanges them by shuffle (a shuffle costs one
                // instruction, the hash eleven)
                uint32_t hq[8];
                if (p.drop_thr) {
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        hq[u] = drop_hash(seed, bh, (uint32_t)(qi * TILE + c * 32 + u * 4 + (lane & 3)), (uint32_t)(kv_idx >> 2), (uint32_t)p.S);
                }
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    float pt[8], ds[8], lse8[8], dv8[8];
                    const int qc0 = c * 32 + g * 8;
                    *reinterpret_cast<float4*>(&lse8[0]) = *reinterpret_cast<const float4*>(&sLse[qc0]);
                    *reinterpret_cast<float4*>(&lse8[4]) = *reinterpret_cast<const float4*>(&sLse[qc0 + 4]);
                    *reinterpret_cast<float4*>(&dv8[0]) = *reinterpret_cast<const float4*>(&sDv[qc0]);
                    *reinterpret_cast<float4*>(&dv8[4]) = *reinterpret_cast<const float4*>(&sDv[qc0 + 4]);
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        const float pv = kv_valid ? exp2f(__uint_as_float(s[g * 8 + t]) * sl2 - lse8[t]) : 0.f;
                        float dpv = __uint_as_float(dp[g * 8 + t]);
                        float pdrop = pv;
                        if (p.drop_thr) {
                            const uint32_t h = __shfl_sync(0xffffffffu, hq[(g * 8 + t) >> 2], (lane & ~3) | (t & 3));
                            const bool kept = ((h >> (8 * (kv_idx & 3))) & 0xffu) >= p.drop_thr;
                            pdrop = kept ? pv * p.drop_scale : 0.f;
                            dpv = kept ? dpv * p.drop_scale : 0.f;
                        }
                        pt[t] = pdrop;
                        ds[t] = pv * (dpv - dv8[t]) * p.scale;
                    }
                    const uint4 dsv = float_to_bf16x8(ds);
                    const uint4 ptv = float_to_bf16x8(pt);
                    st_swizzled_chunk(sdSt, r, c * 4 + g, dsv);
                    ppk[g * 4 + 0] = ptv.x; ppk[g * 4 + 1] = ptv.y; ppk[g * 4 + 2] = ptv.z; ppk[g * 4 + 3] = ptv.w;
                    dpk[g * 4 + 0] = dsv.x; dpk[g * 4 + 1] = dsv.y; dpk[g * 4 + 2] = dsv.z; dpk[g * 4 + 3] = dsv.w;
                }
                // kSingle: bf16 pairs go back over the fp32 columns they came from (16c + 15 < 32c for c >= 1)
                tmem_st_32x16(lane_addr + cPT + c * 16, ppk);
                if (kSingle) tmem_st_32x16(lane_addr + cDST + c * 16, dpk);
            }
            tmem_st_wait();
            fence_proxy_async();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
            tc_fence_before();
            __syncthreads();
            if (threadIdx.x == 0) {
                tc_fence_after();
                constexpr uint32_t idesc_ts = make_idesc(TILE, D, false, true);   // A from TMEM (K-major), B MN-major
                constexpr uint32_t idesc_mm = make_idesc(TILE, D, true, true);    // A, B both MN-major from smem
                const uint32_t acc0 = (!kSingle && qi > 0) ? 1u : 0u;
                // dV += P^T dO : A = P^T in TMEM [kv lanes x q], B = dO tile read MN-major (N = d, K = q rows)
#pragma unroll
                for (int k = 0; k < TILE / 16; ++k)
                    umma_bf16_ts(tmem_base + cDV, tmem_base + cPT + k * 8, make_smem_desc(smem_u32(cdO) + k * 2048, 8192, 1024), idesc_ts,
                                 (k > 0) ? 1u : acc0);
                umma_commit(&bars[4]);
                // dK += dS^T Q : A = dS^T (kSingle: in TMEM; else: the swizzled smem tile read K-major, two 64-query blocks),
                //                B = Q tile read MN-major
                if (kSingle) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)
                        umma_bf16_ts(tmem_base + cDK, tmem_base + cDST + k * 8, make_smem_desc(smem_u32(cQ) + k * 2048, 8192, 1024), idesc_ts,
                                     (k > 0) ? 1u : acc0);
                } else {
                    constexpr uint32_t idesc_kmn = make_idesc(TILE, D, false, true);  // A K-major smem, B MN-major smem
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)
                        umma_bf16(tmem_base + cDK, make_smem_desc(smem_u32(sdSt) + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                                  make_smem_desc(smem_u32(cQ) + k * 2048, 8192, 1024), idesc_kmn, (k > 0) ? 1u : acc0);
                }
                // dQ = dS K : accumulates into the columns P^T occupied, so wait until the dV MMAs have consumed them.
                //   A = dS^T smem tile read MN-major (M = q contiguous, 64-wide blocks kTileBytes apart, K = kv rows)
                //   B = K tile read MN-major (N = d, K = kv rows)
                if (kSingle) {
                    mbar_wait(&bars[4], ph);
                    tc_fence_after();
                }
#pragma unroll
                for (int k = 0; k < TILE / 16; ++k)
                    umma_bf16(tmem_base + cDQ, make_smem_desc(smem_u32(sdSt) + k * 2048, kTileBytes, 1024),
                              make_smem_desc(smem_u32(sK) + k * 2048, 8192, 1024), idesc_mm, k > 0 ? 1u : 0u);
                umma_commit(&bars[5]);
                // software pipeline: the next query tile's S^T / dP^T run behind these MMAs while the threads write dQ out
                if (!kSingle && qi + 1 < n_q) issue_scores(buf ^ 1, qph[buf ^ 1]);
            }
            mbar_wait(&bars[5], ph);
            __syncwarp();
            tc_fence_after();
            // dQ rows = queries (TMEM lane = query index inside the tile)
            {
                const size_t tok = (size_t)q_row0 + r;
#pragma unroll
                for (int c = half * (2 / kHalves); c < (half + 1) * (2 / kHalves); ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_addr + cDQ + c * 32, v);
                    tmem_ld_wait();
                    if (kSingle) {
                        bf16* dst = p.dqkv + tok * 3 * p.H + hd * D + c * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            float o[8];
#pragma unroll
                            for (int t = 0; t < 8; ++t) o[t] = __uint_as_float(v[j + t]);
                            st_global_v4(dst + j, float_to_bf16x8(o));
                        }
                    } else {
                        float* dst = p.dq_acc + tok * p.H + hd * D + c * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            red_add_f32x4(dst + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                          __uint_as_float(v[j + 3]));
                    }
                }
            }
            tc_fence_before();
            __syncthreads();  // S^T / dP^T / dQ columns and the dS^T tile are free again
            ph ^= 1;
        }
    }
    // ---- dK, dV rows = keys
    {
        const size_t tok = (size_t)kv_row0 + r;
        bf16* dk = p.dqkv + tok * 3 * p.H + p.H + hd * D;
        bf16* dv = p.dqkv + tok * 3 * p.H + 2 * p.H + hd * D;
        if (tile_empty) {
            const uint4 z = make_uint4(0, 0, 0, 0);
            for (int j = half * (D / kHalves); j < (half + 1) * (D / kHalves); j += 8) {
                st_global_v4(dk + j, z);
                st_global_v4(dv + j, z);
            }
        } else {
            tc_fence_after();
#pragma unroll
            for (int c = half * (2 / kHalves); c < (half + 1) * (2 / kHalves); ++c) {
                uint32_t a[32], g[32];
                tmem_ld_32x32(lane_addr + cDK + c * 32, a);
                tmem_ld_32x32(lane_addr + cDV + c * 32, g);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    float o[8], w[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        o[t] = __uint_as_float(a[j + t]);
                        w[t] = __uint_as_float(g[j + t]);
                    }
                    st_global_v4(dk + c * 32 + j, float_to_bf16x8(o));
                    st_global_v4(dv + c * 32 + j, float_to_bf16x8(w));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// dq_acc [tokens, H] fp32 -> q part of dqkv [tokens, 3H] bf16
__global__ void attn_dq_convert_kernel(const float* __restrict__ acc, bf16* __restrict__ dqkv, size_t tokens, int H) {
    const int vc = H >> 3;
    const size_t total = tokens * vc;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t tok = t / vc;
        const int c8 = (int)(t % vc);
        float f[8];
        *reinterpret_cast<float4*>(&f[0]) = *reinterpret_cast<const float4*>(acc + tok * H + c8 * 8);
        *reinterpret_cast<float4*>(&f[4]) = *reinterpret_cast<const float4*>(acc + tok * H + c8 * 8 + 4);
        st_global_v4(dqkv + tok * 3 * H + c8 * 8, float_to_bf16x8(f));
    }
}

// ---------------------------------------------------------------------------- host
int make_tile_map(CUtensorMap* map, const void* ptr, long long rows, long long cols) {
    DDL_REQUIRE(cols % 8 == 0, "attention: operand pitch must be a multiple of 8 elements");
    return tma_host::encode_2d(map, ptr, rows, cols, cols * 2, 128, "attention");
}

int fill_params(AttnParams* p, int B, int S, int heads, const int* seqlens, float dropout_p, unsigned long long seed,
                const unsigned long long* seed_dev) {
    DDL_REQUIRE(S % TILE == 0 && S >= TILE && S <= 512, "attention: sequence length %d must be a multiple of 128 in [128, 512]", S);
    DDL_REQUIRE(B >= 1 && heads >= 1, "attention: bad shape");
    p->B = B; p->S = S; p->heads = heads; p->H = heads * D;
    p->seqlens = seqlens;
    p->scale = 0.125f;
    int thr = (int)(dropout_p * 256.f + 0.5f);
    thr = thr < 0 ? 0 : (thr > 255 ? 255 : thr);
    p->drop_thr = (uint32_t)thr;
    p->drop_scale = 1.f / (1.f - (float)thr / 256.f);
    p->seed = seed;
    p->seed_dev = (const uint64_t*)seed_dev;
    p->ctx = nullptr; p->lse = nullptr; p->dvec = nullptr; p->dqkv = nullptr; p->dq_acc = nullptr;
    return 0;
}

// S > 128: streaming kernel (two CTAs per SM) unless DDL_ATTN_FWD=resident asks for the all-of-S-in-TMEM variant
bool attn_fwd_streaming() {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("DDL_ATTN_FWD");
        mode = (e && strcmp(e, "resident") == 0) ? 0 : 1;
    }
    return mode == 1;
}

template <int KV>
int launch_fwd(const CUtensorMap& map, const AttnParams& p, cudaStream_t s) {
    constexpr int smem = (1 + 2 * KV) * kTileBytes + 64 + 2 * TILE * 4 + 1024;
    static bool set = false;
    if (!set) {
        DDL_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<KV>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        set = true;
    }
    dim3 grid(p.S / TILE, p.heads, p.B);
    attn_fwd_kernel<KV><<<grid, KV == 1 ? 128 : 256, smem, s>>>(map, p);
    DDL_LAUNCH_CHECK();
    return 0;
}

}  // namespace

DDL_API int ddl_attention_fwd(const void* qkv, const int* seqlens, void* ctx, float* lse, int B, int S, int heads, float dropout_p,
                              unsigned long long seed, const unsigned long long* seed_dev, void* stream_) {
    AttnParams p;
    if (int e = fill_params(&p, B, S, heads, seqlens, dropout_p, seed, seed_dev)) return e;
    DDL_REQUIRE(qkv && ctx && lse, "attention fwd: null pointer");
    p.ctx = (bf16*)ctx;
    p.lse = lse;
    CUtensorMap map;
    if (int e = make_tile_map(&map, qkv, (long long)B * S, 3LL * p.H)) return e;
    cudaStream_t s = (cudaStream_t)stream_;
    if (S > TILE && attn_fwd_streaming()) {
        constexpr int smem = 5 * kTileBytes + 128 + 2 * TILE * 4 + 1024;
        static bool set = false;
        if (!set) {
            DDL_CHECK(cudaFuncSetAttribute(attn_fwd_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            set = true;
        }
        dim3 grid(p.S / TILE, p.heads, p.B);
        attn_fwd_stream_kernel<<<grid, 256, smem, s>>>(map, p);
        DDL_LAUNCH_CHECK();
        return 0;
    }
    switch (S / TILE) {
        case 1: return launch_fwd<1>(map, p, s);
        case 2: return launch_fwd<2>(map, p, s);
        case 3: return launch_fwd<3>(map, p, s);
        default: return launch_fwd<4>(map, p, s);
    }
}

DDL_API size_t ddl_attention_bwd_scratch_bytes(int B, int S, int heads) {
    size_t n = (size_t)B * heads * S * sizeof(float);                       // D vector
    if (S > TILE) n += (size_t)B * S * heads * D * sizeof(float);           // fp32 dQ accumulator
    return n;
}

DDL_API int ddl_attention_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse, const int* seqlens, void* dqkv,
                              void* scratch, size_t scratch_bytes, int B, int S, int heads, float dropout_p, unsigned long long seed,
                              const unsigned long long* seed_dev, void* stream_) {
    AttnParams p;
    if (int e = fill_params(&p, B, S, heads, seqlens, dropout_p, seed, seed_dev)) return e;
    DDL_REQUIRE(qkv && ctx && dctx && lse && dqkv && scratch, "attention bwd: null pointer");
    DDL_REQUIRE(scratch_bytes >= ddl_attention_bwd_scratch_bytes(B, S, heads), "attention bwd: scratch too small");
    cudaStream_t s = (cudaStream_t)stream_;
    float* dvec = (float*)scratch;
    float* dq_acc = S > TILE ? dvec + (size_t)B * heads * S : nullptr;
    p.lse = const_cast<float*>(lse);
    p.dvec = dvec;
    p.dqkv = (bf16*)dqkv;
    p.dq_acc = dq_acc;
    const size_t rows = (size_t)B * S * heads;
    attn_dvec_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, s>>>((const bf16*)dctx, (const bf16*)ctx, dvec, B, S, heads);
    DDL_LAUNCH_CHECK();
    if (dq_acc) DDL_CHECK(cudaMemsetAsync(dq_acc, 0, (size_t)B * S * p.H * sizeof(float), s));
    CUtensorMap mq, md;
    if (int e = make_tile_map(&mq, qkv, (long long)B * S, 3LL * p.H)) return e;
    if (int e = make_tile_map(&md, dctx, (long long)B * S, (long long)p.H)) return e;
    dim3 grid(S / TILE, heads, B);
    if (S == TILE) {
        constexpr int smem = 6 * kTileBytes + 2 * TILE * 4 + 128 + 1024;
        static bool set1 = false;
        if (!set1) {
            DDL_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            set1 = true;
        }
        attn_bwd_kernel<true><<<grid, 128, smem, s>>>(mq, md, p);
    } else {
        constexpr int smem = 8 * kTileBytes + 2 * TILE * 4 + 128 + 1024;
        static bool set2 = false;
        if (!set2) {
            DDL_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            set2 = true;
        }
        attn_bwd_kernel<false><<<grid, 256, smem, s>>>(mq, md, p);
    }
    DDL_LAUNCH_CHECK();
    if (dq_acc) {
        const size_t total = (size_t)B * S * (p.H / 8);
        int blocks = (int)((total + 255) / 256);
        if (blocks > DDL_SM_COUNT * 8) blocks = DDL_SM_COUNT * 8;
        attn_dq_convert_kernel<<<blocks, 256, 0, s>>>(dq_acc, (bf16*)dqkv, (size_t)B * S, p.H);
        DDL_LAUNCH_CHECK();
    }
    return 0;
}
