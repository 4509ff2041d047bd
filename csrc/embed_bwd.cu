// Gradients of the three BERT embedding tables (K14 "embedding gather" backward; model at BASELINE.json:9-10, HF
// BertEmbeddings) from dz [B*S, H] bf16, the gradient of word[ids] + pos[s] + type[tt]:
//   word  [V, H]: scatter-add by token id - fp32 red.global.add into a caller-zeroed accumulator (token ids repeat),
//                 converted to bf16 by ddl_f32_to_bf16 afterwards;
//   pos   [P, H]: dpos[s] = sum_b dz[b, s]   - deterministic, written as bf16 directly (rows s >= S are zero);
//   type  [T, H]: dtype[t] = sum_{rows with tt = t} dz - the 16 k rows fall onto T = 2 table rows, so atomics would
//                 serialise (the ATen index_add this replaces took 0.12 ms per call for exactly that reason); instead
//                 every block sums its slab of rows per type in registers and a second kernel folds the blocks in
//                 fixed order.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

constexpr int kMaxTypes = 4;
constexpr int kTypeBlocks = 296;

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// one warp per token row; lanes stride over the H/8 vectors
__global__ void __launch_bounds__(256) embed_bwd_word_kernel(const bf16* __restrict__ dz, const long long* __restrict__ ids,
                                                             float* __restrict__ dword, int rows, int H, int V) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nv = H >> 3;
    for (long long row = (long long)blockIdx.x * 8 + warp; row < rows; row += (long long)gridDim.x * 8) {
        const long long id = ids[row];
        if (id < 0 || id >= V) continue;
        float* dst = dword + (size_t)id * H;
        for (int c = lane; c < nv; c += 32) {
            float f[8];
            bf16x8_to_float(ld_stream_v4(dz + (size_t)row * H + c * 8), f);
            red_add_v4(dst + c * 8, f[0], f[1], f[2], f[3]);
            red_add_v4(dst + c * 8 + 4, f[4], f[5], f[6], f[7]);
        }
    }
}

// thread per (position s, vector c): sum over the batch; positions >= S get zeros
__global__ void __launch_bounds__(256) embed_bwd_pos_kernel(const bf16* __restrict__ dz, bf16* __restrict__ dpos, int B, int S, int H,
                                                            int P) {
    const int nv = H >> 3;
    const size_t total = (size_t)P * nv;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % nv), s = (int)(t / nv);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (s < S) {
            for (int b = 0; b < B; ++b) {
                float f[8];
                bf16x8_to_float(ld_stream_v4(dz + ((size_t)b * S + s) * H + c * 8), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += f[j];
            }
        }
        st_global_v4(dpos + (size_t)s * H + c * 8, float_to_bf16x8(acc));
    }
}

// block = H/8 threads (one vector each), a contiguous slab of rows per block -> partial [block][T][H]
__global__ void __launch_bounds__(256) embed_bwd_type_partial_kernel(const bf16* __restrict__ dz, const long long* __restrict__ tts,
                                                                     float* __restrict__ partial, int rows, int H, int T) {
    const int c = threadIdx.x;
    const int nv = H >> 3;
    const int per = (rows + gridDim.x - 1) / gridDim.x;
    const int r0 = blockIdx.x * per, r1 = min(rows, r0 + per);
    float acc[kMaxTypes][8];
#pragma unroll
    for (int t = 0; t < kMaxTypes; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
    if (c < nv) {
        for (int r = r0; r < r1; ++r) {
            const int tt = (int)tts[r];
            float f[8];
            bf16x8_to_float(ld_stream_v4(dz + (size_t)r * H + c * 8), f);
#pragma unroll
            for (int t = 0; t < kMaxTypes; ++t) {
                const float m = tt == t ? 1.f : 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[t][j] += m * f[j];
            }
        }
        for (int t = 0; t < T && t < kMaxTypes; ++t) {
            float* dst = partial + ((size_t)blockIdx.x * T + t) * H + c * 8;
            *reinterpret_cast<float4*>(dst) = make_float4(acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[t][4], acc[t][5], acc[t][6], acc[t][7]);
        }
    }
}

// grid = ceil(T*H / 32), block = 256: lane = column, 8 warps fold the blocks in fixed order
__global__ void __launch_bounds__(256) embed_bwd_type_final_kernel(const float* __restrict__ partial, int nblocks, int cols,
                                                                   bf16* __restrict__ out) {
    __shared__ float sh[8][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float a0 = 0.f, a1 = 0.f;
    if (c < cols) {
        int i = w;
        for (; i + 8 < nblocks; i += 16) {
            a0 += partial[(size_t)i * cols + c];
            a1 += partial[(size_t)(i + 8) * cols + c];
        }
        if (i < nblocks) a0 += partial[(size_t)i * cols + c];
    }
    sh[w][lane] = a0 + a1;
    __syncthreads();
    if (w != 0 || c >= cols) return;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += sh[k][lane];
    out[c] = __float2bfloat16(acc);
}

}  // namespace

DDL_API size_t ddl_embed_bwd_scratch_bytes(int H, int T) { return (size_t)kTypeBlocks * T * H * sizeof(float); }

// dword_f32 [V, H] must be zero on entry (accumulated into); dpos [P, H] and dtype [T, H] bf16 are overwritten.
DDL_API int ddl_embed_bwd(const void* dz, const long long* ids, const long long* token_types, float* dword_f32, void* dpos, void* dtype,
                          void* scratch, size_t scratch_bytes, int B, int S, int H, int V, int P, int T, void* stream_) {
    DDL_REQUIRE(H % 8 == 0 && H >= 8 && H <= 2048, "embed_bwd: hidden size %d must be a multiple of 8 and <= 2048", H);
    DDL_REQUIRE(T >= 1 && T <= kMaxTypes, "embed_bwd: at most %d token types (got %d)", kMaxTypes, T);
    DDL_REQUIRE(S <= P, "embed_bwd: sequence length %d exceeds the position table (%d)", S, P);
    DDL_REQUIRE(dz && ids && token_types && dword_f32 && dpos && dtype && scratch, "embed_bwd: null pointer");
    DDL_REQUIRE(scratch_bytes >= ddl_embed_bwd_scratch_bytes(H, T), "embed_bwd: scratch too small");
    const int rows = B * S;
    if (rows == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream_;
    int grid = (rows + 7) / 8;
    if (grid > DDL_SM_COUNT * 8) grid = DDL_SM_COUNT * 8;
    embed_bwd_word_kernel<<<grid, 256, 0, s>>>((const bf16*)dz, ids, dword_f32, rows, H, V);
    DDL_LAUNCH_CHECK();
    const size_t ptotal = (size_t)P * (H / 8);
    int pgrid = (int)((ptotal + 255) / 256);
    if (pgrid > DDL_SM_COUNT * 8) pgrid = DDL_SM_COUNT * 8;
    embed_bwd_pos_kernel<<<pgrid, 256, 0, s>>>((const bf16*)dz, (bf16*)dpos, B, S, H, P);
    DDL_LAUNCH_CHECK();
    int tblocks = kTypeBlocks < rows ? kTypeBlocks : rows;
    embed_bwd_type_partial_kernel<<<tblocks, 256, 0, s>>>((const bf16*)dz, token_types, (float*)scratch, rows, H, T);
    DDL_LAUNCH_CHECK();
    embed_bwd_type_final_kernel<<<ddl_ceil_div(T * H, 32), 256, 0, s>>>((const float*)scratch, tblocks, T * H, (bf16*)dtype);
    DDL_LAUNCH_CHECK();
    return 0;
}
