// Pooling on NHWC bf16 activations (K14): the 3x3/stride-2/pad-1 max pool of
// the ResNet stem (112^2 -> 56^2, reference model at notebooks/cv/onnx_experiments.py:19) and
// the global average pool before the classifier (7^2 -> 1), forward and backward.
//
// Max pool (odd image sizes; even sizes: see the *_even_* kernels below): one thread per output pixel x 8 channels (one
// 16-byte vector): 9 vector loads, one vector store, plus one byte per element recording which of the 9 taps won.  Backward
// is a gather: one thread per INPUT pixel x 8 channels looks at the (at most 4) windows that
// contain it and sums the gradients of those whose recorded tap points here - no atomics,
// every dx element written exactly once.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) maxpool3x3s2_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y,
                                                                unsigned char* __restrict__ idx, int N, int H, int W, int C, int OH,
                                                                int OW) {
    const int vc = C >> 3;
    const size_t total = (size_t)N * OH * OW * vc;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(t % vc);
        size_t r = t / vc;
        const int ow = (int)(r % OW);
        r /= OW;
        const int oh = (int)(r % OH);
        const int n = (int)(r / OH);
        float best[8];
        int arg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            best[j] = -INFINITY;
            arg[j] = 0;
        }
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int ih = oh * 2 - 1 + kh;
            if (ih < 0 || ih >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int iw = ow * 2 - 1 + kw;
                if (iw < 0 || iw >= W) continue;
                float f[8];
                bf16x8_to_float(ld_global_v4(x + (((size_t)n * H + ih) * W + iw) * C + c8 * 8), f);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (f[j] > best[j]) {
                        best[j] = f[j];
                        arg[j] = kh * 3 + kw;
                    }
            }
        }
        const size_t o = (((size_t)n * OH + oh) * OW + ow) * C + c8 * 8;
        st_global_v4(y + o, float_to_bf16x8(best));
        uint2 packed;
        packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
        packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
        *reinterpret_cast<uint2*>(idx + o) = packed;
    }
}

// dy2 (nullable): a second gradient of the pooled tensor (the other branch of a residual fork), summed on the fly
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ dy2,
                                                                const unsigned char* __restrict__ idx, bf16* __restrict__ dx, int N, int H,
                                                                int W, int C, int OH, int OW) {
    const int vc = C >> 3;
    const size_t total = (size_t)N * H * W * vc;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(t % vc);
        size_t r = t / vc;
        const int iw = (int)(r % W);
        r /= W;
        const int ih = (int)(r % H);
        const int n = (int)(r / H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // windows (oh, ow) with 2*oh - 1 <= ih <= 2*oh + 1
        const int oh0 = ih >> 1, oh1 = (ih + 1) >> 1;
        const int ow0 = iw >> 1, ow1 = (iw + 1) >> 1;
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const int oh = a == 0 ? oh0 : oh1;
            if (a == 1 && oh1 == oh0) continue;
            if (oh >= OH) continue;
            const int kh = ih - (oh * 2 - 1);
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int ow = b == 0 ? ow0 : ow1;
                if (b == 1 && ow1 == ow0) continue;
                if (ow >= OW) continue;
                const int kw = iw - (ow * 2 - 1);
                const int code = kh * 3 + kw;
                const size_t o = (((size_t)n * OH + oh) * OW + ow) * C + c8 * 8;
                const uint2 packed = *reinterpret_cast<const uint2*>(idx + o);
                float g[8];
                bf16x8_to_float(ld_global_v4(dy + o), g);
                if (dy2) {
                    float g2[8];
                    bf16x8_to_float(ld_global_v4(dy2 + o), g2);
#pragma unroll
                    for (int j = 0; j < 8; ++j) g[j] += g2[j];
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int aj = (int)(((j < 4 ? packed.x : packed.y) >> (8 * (j & 3))) & 0xffu);
                    if (aj == code) acc[j] += g[j];
                }
            }
        }
        st_global_v4(dx + t * 8, float_to_bf16x8(acc));
    }
}

// ---- even H and W (the ResNet stem: 112^2 -> 56^2): the same two passes with a fraction of the instructions ---------
// The generic kernels above spend ~600 instructions per 16-byte vector on 64-bit index divisions and per-element float
// compares and ran at 2.5 - 4x their memory time.  Here a block owns one output row (forward) or one pair of input rows
// (backward) of one image, so that no division is left; the maximum is taken on packed bf16 pairs (HMNMX2) and the
// winning tap is found by comparing the taps with it, last to first (the lowest tap index wins a tie, as in ATen); the
// tap codes are stored as 4 bits per element (one 32-bit word per 8 channels).  The backward pass handles a 2 x 2 patch
// of input pixels per thread: the four windows that reach it are loaded once (the generic kernel loads nine per patch).
__device__ __forceinline__ uint32_t hmax2_u32(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
// 0xffff in each half where a == b
__device__ __forceinline__ uint32_t heq2_mask(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("set.eq.u32.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

__global__ void __launch_bounds__(256) maxpool3x3s2_even_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y,
                                                                     uint32_t* __restrict__ idx, int H, int W, int C, int OH, int OW) {
    const int vc = C >> 3, oh = blockIdx.x, n = blockIdx.y;
    const bf16* xin = x + (size_t)n * H * W * C;
    const size_t orow = ((size_t)n * OH + oh) * OW;
    for (int t = threadIdx.x; t < OW * vc; t += blockDim.x) {
        const int ow = t / vc, c8 = t - ow * vc;
        uint4 v[9];
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int ih = oh * 2 - 1 + kh;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int iw = ow * 2 - 1 + kw;
                // out of the image: -inf in every element (0xff80 = bf16 -inf); never wins, never ties with a finite value
                v[kh * 3 + kw] = (ih >= 0 && iw >= 0) ? ld_global_v4(xin + ((size_t)ih * W + iw) * C + c8 * 8)
                                                      : make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);
            }
        }
        uint32_t m[4], arg[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t* w0 = reinterpret_cast<const uint32_t*>(&v[0]);
            uint32_t mm = w0[q];
#pragma unroll
            for (int tap = 1; tap < 9; ++tap) mm = hmax2_u32(mm, reinterpret_cast<const uint32_t*>(&v[tap])[q]);
            uint32_t a = 0;
#pragma unroll
            for (int tap = 8; tap >= 0; --tap) {
                const uint32_t eq = heq2_mask(reinterpret_cast<const uint32_t*>(&v[tap])[q], mm);
                a = (a & ~eq) | ((uint32_t)tap * 0x00010001u & eq);
            }
            m[q] = mm;
            arg[q] = a;
        }
        const size_t o = orow + ow;
        st_global_v4(y + o * C + c8 * 8, make_uint4(m[0], m[1], m[2], m[3]));
        // element 2q -> bits [8q, 8q+4), element 2q+1 -> bits [8q+4, 8q+8)
        uint32_t packed = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) packed |= ((arg[q] & 0xfu) | ((arg[q] >> 12) & 0xf0u)) << (8 * q);
        idx[o * vc + c8] = packed;
    }
}

__global__ void __launch_bounds__(256) maxpool3x3s2_even_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ dy2,
                                                                     const uint32_t* __restrict__ idx, bf16* __restrict__ dx, int H, int W, int C,
                                                                     int OH, int OW) {
    const int vc = C >> 3, a = blockIdx.x, n = blockIdx.y;  // input rows 2a, 2a + 1
    bf16* dxn = dx + (size_t)n * H * W * C;
    for (int t = threadIdx.x; t < OW * vc; t += blockDim.x) {
        const int b = t / vc, c8 = t - b * vc;  // input columns 2b, 2b + 1
        // all loads of the item first (with them behind per-window branches the kernel ran latency-bound)
        uint32_t cdw[4];
        uint4 gv[4], g2v[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int oh = a + (w >> 1), ow = b + (w & 1);
            const bool ok = oh < OH && ow < OW;
            const size_t o = ((size_t)n * OH + (ok ? oh : a)) * OW + (ok ? ow : b);
            const uint32_t c = idx[o * vc + c8];
            cdw[w] = ok ? c : 0xffffffffu;  // 15 in every nibble matches no tap
            gv[w] = ld_global_v4(dy + o * C + c8 * 8);
            if (dy2) g2v[w] = ld_global_v4(dy2 + o * C + c8 * 8);
        }
        float acc[4][8];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[i][e] = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int di = w >> 1, dj = w & 1;
            const uint32_t codes = cdw[w];
            float g[8];
            bf16x8_to_float(gv[w], g);
            if (dy2) {
                float g2[8];
                bf16x8_to_float(g2v[w], g2);
#pragma unroll
                for (int e = 0; e < 8; ++e) g[e] += g2[e];
            }
            // pixel (i, j) of the patch is tap (i + 1 - 2 di, j + 1 - 2 dj) of this window when that lies in [0, 2]
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                if (di == 1 && i == 0) continue;
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    if (dj == 1 && j == 0) continue;
                    // bit 4e + 3 of m is set where nibble e of the code word equals this tap (exact: no carry crosses a nibble)
                    const uint32_t t_ = codes ^ ((uint32_t)((i + 1 - 2 * di) * 3 + (j + 1 - 2 * dj)) * 0x11111111u);
                    const uint32_t m = ~(((t_ & 0x77777777u) + 0x77777777u) | t_) & 0x88888888u;
#pragma unroll
                    for (int e = 0; e < 8; ++e)
                        if (m & (8u << (4 * e))) acc[i * 2 + j][e] += g[e];
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
                st_global_v4(dxn + ((size_t)(2 * a + i) * W + 2 * b + j) * C + c8 * 8, float_to_bf16x8(acc[i * 2 + j]));
    }
}

// global average pool: y[n, c] = mean_hw x[n, hw, c]
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int N, int HW, int C) {
    const int vc = C >> 3;
    const size_t total = (size_t)N * vc;
    const float inv = 1.f / (float)HW;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(t % vc);
        const int n = (int)(t / vc);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int p = 0; p < HW; ++p) {
            float f[8];
            bf16x8_to_float(ld_stream_v4(x + ((size_t)n * HW + p) * C + c8 * 8), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] *= inv;
        st_global_v4(y + t * 8, float_to_bf16x8(acc));
    }
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const bf16* __restrict__ dy, bf16* __restrict__ dx, int N, int HW, int C) {
    const int vc = C >> 3;
    const size_t total = (size_t)N * HW * vc;
    const float inv = 1.f / (float)HW;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(t % vc);
        const int n = (int)(t / ((size_t)HW * vc));
        float f[8];
        bf16x8_to_float(ld_global_v4(dy + ((size_t)n * vc + c8) * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] *= inv;
        st_global_v4(dx + t * 8, float_to_bf16x8(f));
    }
}

// Network input: [N, C, H*W] planes (fp32 or bf16, C <= 8) -> [N, H*W, 8] bf16 pixels with the channels
// C..7 zero.  One thread per pixel: C coalesced plane reads, one 16-byte store.  Eight channels make the stem
// convolution an ordinary tensor-core implicit GEMM for the library (a 3-channel NHWC input is not).
template <typename T>
__global__ void __launch_bounds__(256) nchw_to_nhwc8_kernel(const T* __restrict__ x, bf16* __restrict__ y, int N, int C, size_t HW) {
    const size_t total = (size_t)N * HW;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t n = t / HW, px = t - n * HW;
        float v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) v[c] = c < C ? (float)x[(n * C + c) * HW + px] : 0.f;
        st_global_v4(y + t * 8, float_to_bf16x8(v));
    }
}

// even image sizes take the row-per-block kernels (their index format differs: forward and backward decide alike)
bool maxpool_even(int N, int H, int W) { return H % 2 == 0 && W % 2 == 0 && N <= 65535 && H >= 2 && W >= 2; }

int grid_for(size_t total) {
    long long b = (long long)((total + 255) / 256);
    const long long cap = DDL_SM_COUNT * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

DDL_API int ddl_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream_) {
    DDL_REQUIRE(C % 8 == 0, "maxpool: channels must be a multiple of 8");
    DDL_REQUIRE(x && y && idx, "maxpool: null pointer");
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    if (maxpool_even(N, H, W)) {  // idx: 4-bit tap codes, C / 2 bytes per output pixel (the buffer holds C)
        DDL_REQUIRE(((uintptr_t)idx & 3) == 0, "maxpool: index buffer must be 4-byte aligned");
        maxpool3x3s2_even_fwd_kernel<<<dim3(OH, N), 256, 0, (cudaStream_t)stream_>>>((const bf16*)x, (bf16*)y, (uint32_t*)idx, H, W, C, OH, OW);
        DDL_LAUNCH_CHECK();
        return 0;
    }
    const size_t total = (size_t)N * OH * OW * (C / 8);
    maxpool3x3s2_fwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream_>>>((const bf16*)x, (bf16*)y, (unsigned char*)idx, N, H, W, C,
                                                                               OH, OW);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_maxpool3x3s2_bwd_ex(const void* dy, const void* dy2, const void* idx, void* dx, int N, int H, int W, int C,
                                    void* stream_) {
    DDL_REQUIRE(C % 8 == 0, "maxpool: channels must be a multiple of 8");
    DDL_REQUIRE(dy && dx && idx, "maxpool: null pointer");
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    if (maxpool_even(N, H, W)) {
        maxpool3x3s2_even_bwd_kernel<<<dim3(H / 2, N), 256, 0, (cudaStream_t)stream_>>>((const bf16*)dy, (const bf16*)dy2, (const uint32_t*)idx,
                                                                                        (bf16*)dx, H, W, C, OH, OW);
        DDL_LAUNCH_CHECK();
        return 0;
    }
    const size_t total = (size_t)N * H * W * (C / 8);
    maxpool3x3s2_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream_>>>((const bf16*)dy, (const bf16*)dy2, (const unsigned char*)idx,
                                                                               (bf16*)dx, N, H, W, C, OH, OW);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, void* stream_) {
    return ddl_maxpool3x3s2_bwd_ex(dy, nullptr, idx, dx, N, H, W, C, stream_);
}

DDL_API int ddl_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream_) {
    DDL_REQUIRE(C % 8 == 0, "avgpool: channels must be a multiple of 8");
    avgpool_fwd_kernel<<<grid_for((size_t)N * (C / 8)), 256, 0, (cudaStream_t)stream_>>>((const bf16*)x, (bf16*)y, N, HW, C);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream_) {
    DDL_REQUIRE(C % 8 == 0, "avgpool: channels must be a multiple of 8");
    avgpool_bwd_kernel<<<grid_for((size_t)N * HW * (C / 8)), 256, 0, (cudaStream_t)stream_>>>((const bf16*)dy, (bf16*)dx, N, HW, C);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_nchw_to_nhwc8_bf16(const void* x, int x_dtype, void* y, int N, int C, long long HW, void* stream_) {
    DDL_REQUIRE(x && y, "nchw_to_nhwc8: null pointer");
    DDL_REQUIRE(C >= 1 && C <= 8 && N >= 1 && HW >= 1, "nchw_to_nhwc8: need 1 <= C <= 8 (got %d)", C);
    DDL_REQUIRE(x_dtype == DDL_F32 || x_dtype == DDL_BF16, "nchw_to_nhwc8: input must be float32 or bfloat16");
    const int grid = grid_for((size_t)N * HW);
    if (x_dtype == DDL_F32) nchw_to_nhwc8_kernel<float><<<grid, 256, 0, (cudaStream_t)stream_>>>((const float*)x, (bf16*)y, N, C, (size_t)HW);
    else nchw_to_nhwc8_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream_>>>((const bf16*)x, (bf16*)y, N, C, (size_t)HW);
    DDL_LAUNCH_CHECK();
    return 0;
}
