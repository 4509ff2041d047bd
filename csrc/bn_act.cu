// BatchNorm (training) + residual add + ReLU on NHWC bf16 activations, forward and backward
// (the BN/ReLU half of K9 "conv-BN-ReLU"; workload: the 53 BatchNorm2d of
// the reference's ResNet-50, notebooks/cv/onnx_experiments.py:19).
//
// The activation is a [M = N*H*W, C] matrix with C contiguous.  A thread owns one
// 8-channel (16-byte) column group for its whole life: the grid's total thread count is a
// multiple of C/8, so striding over the flat vector index never changes a thread's
// channels and per-channel constants stay in registers.
//
//   forward : stats pass   (read x)                      -> per-block partial sum / sumsq
//             finalize     (C threads)                   -> mean, invstd, running stats, scale/shift
//             apply pass   (read x [+res], write y)      -> y = relu(x*scale + shift [+ res])
//   backward: reduce pass  (read dy, y, x [write g])     -> partial sum(g), sum(g*xhat)
//             finalize                                   -> dgamma, dbeta, coefficients
//             apply pass   (read g|dy,y and x, write dx) -> dx
//
// versus stock eager (BN kernel, ReLU kernel, add kernel, each a full read+write), the
// forward moves 2R+1W (+1R residual) instead of up to 5R+3W.  All cross-block reductions
// go through a scratch buffer folded in fixed order: deterministic, no atomics.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

constexpr int kBnMaxThreads = 256;
constexpr int kBnGridCap = DDL_SM_COUNT * 2;
constexpr int kUnroll = 4;

struct BnGeom {
    int vec_cols;  // C / 8
    int threads;   // multiple of vec_cols, <= 256
    int grid;
    size_t nvec;   // M * C / 8
};

int make_geom(long long M, int C, BnGeom* g) {
    DDL_REQUIRE(C % 8 == 0 && C >= 8 && C <= 2048, "bn: channels %d must be a multiple of 8 and <= 2048", C);
    DDL_REQUIRE(M >= 1, "bn: empty activation");
    g->vec_cols = C / 8;
    g->threads = (kBnMaxThreads / g->vec_cols) * g->vec_cols;
    g->nvec = (size_t)M * g->vec_cols;
    long long want = (long long)((g->nvec + (size_t)g->threads * kUnroll - 1) / ((size_t)g->threads * kUnroll));
    g->grid = (int)(want < 1 ? 1 : (want > kBnGridCap ? kBnGridCap : want));
    return 0;
}

// block-level fold of per-thread [2][8] accumulators over the threads that share a column
// group; result for column group vc lands in partial[block][which][vc*8 + j]
__device__ __forceinline__ void fold_to_partials(const float (&a)[8], const float (&b)[8], int vec_cols, int C,
                                                 float* __restrict__ partial_block, float* sh /*[threads][17]*/) {
    float* mine = sh + threadIdx.x * 17;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mine[j] = a[j];
        mine[8 + j] = b[j];
    }
    __syncthreads();
    // thread t < vec_cols*16 handles (vc = t / 16, slot = t % 16)
    for (int t = threadIdx.x; t < vec_cols * 16; t += blockDim.x) {
        const int vc = t >> 4, slot = t & 15;
        float acc = 0.f;
        for (int r = vc; r < (int)blockDim.x; r += vec_cols) acc += sh[r * 17 + slot];
        const int which = slot >> 3, j = slot & 7;
        partial_block[which * C + vc * 8 + j] = acc;
    }
}

__global__ void __launch_bounds__(kBnMaxThreads) bn_stats_kernel(const bf16* __restrict__ x, size_t nvec, int vec_cols, int C,
                                                                 float* __restrict__ partials) {
    extern __shared__ float sh[];
    float s[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * kUnroll) {
        uint4 u[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) u[k] = ld_stream_v4(x + (v0 + k * stride) * 8);
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                float f[8];
                bf16x8_to_float(u[k], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    s[j] += f[j];
                    q[j] += f[j] * f[j];
                }
            }
    }
    fold_to_partials(s, q, vec_cols, C, partials + (size_t)blockIdx.x * 2 * C, sh);
}

// Fold the per-block partials of 32 channels with kFoldWarps warps in parallel (fixed order ->
// deterministic): warp w sums partial rows w, w+kFoldWarps, ... four at a time with independent
// accumulators (the loads are L2 hits whose latency is the whole cost of this kernel, so they
// must all be in flight together); shared memory folds the warps.
constexpr int kFoldWarps = 32, kFoldThreads = kFoldWarps * 32;
__device__ __forceinline__ void fold_partials32(const float* __restrict__ partials, int nblocks, int C, int c, bool ok, float& s,
                                                float& q) {
    __shared__ float sh[kFoldWarps][2][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (ok) {
        int i = w;
        for (; i + 3 * kFoldWarps < nblocks; i += 4 * kFoldWarps) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float* row = partials + (size_t)(i + u * kFoldWarps) * 2 * C;
                a[u] += row[c];
                b[u] += row[C + c];
            }
        }
        for (; i < nblocks; i += kFoldWarps) {
            a[0] += partials[(size_t)i * 2 * C + c];
            b[0] += partials[(size_t)i * 2 * C + C + c];
        }
    }
    sh[w][0][lane] = (a[0] + a[1]) + (a[2] + a[3]);
    sh[w][1][lane] = (b[0] + b[1]) + (b[2] + b[3]);
    __syncthreads();
    s = 0.f;
    q = 0.f;
    if (w == 0) {
#pragma unroll
        for (int k = 0; k < kFoldWarps; ++k) {
            s += sh[k][0][lane];
            q += sh[k][1][lane];
        }
    }
}

// mean/var from partials; running-stat update (unbiased var, like torch); scale/shift
// grid = ceil(C / 32), block = kFoldThreads
__global__ void __launch_bounds__(kFoldThreads) bn_fwd_finalize_kernel(const float* __restrict__ partials, int nblocks, long long M, int C,
                                       const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                                       float* __restrict__ running_mean, float* __restrict__ running_var, float momentum, float eps,
                                       float* __restrict__ mean_out, float* __restrict__ invstd_out, float* __restrict__ scale_out,
                                       float* __restrict__ shift_out) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    float s, q;
    fold_partials32(partials, nblocks, C, c, c < C, s, q);
    if (c >= C || threadIdx.x >= 32) return;
    const float inv_m = 1.f / (float)M;
    const float mean = s * inv_m;
    float var = q * inv_m - mean * mean;
    var = var < 0.f ? 0.f : var;
    const float invstd = rsqrtf(var + eps);
    mean_out[c] = mean;
    invstd_out[c] = invstd;
    const float g = gamma ? __bfloat162float(gamma[c]) : 1.f, b = beta ? __bfloat162float(beta[c]) : 0.f;
    scale_out[c] = g * invstd;
    shift_out[c] = b - mean * g * invstd;
    if (running_mean) {
        const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}

// scale/shift from running statistics (eval mode)
__global__ void bn_eval_coeff_kernel(int C, const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                                     const float* __restrict__ running_mean, const float* __restrict__ running_var, float eps,
                                     float* __restrict__ scale_out, float* __restrict__ shift_out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float invstd = rsqrtf(running_var[c] + eps);
    const float g = gamma ? __bfloat162float(gamma[c]) : 1.f, b = beta ? __bfloat162float(beta[c]) : 0.f;
    scale_out[c] = g * invstd;
    shift_out[c] = b - running_mean[c] * g * invstd;
}

template <bool kRelu, bool kRes>
__global__ void __launch_bounds__(kBnMaxThreads) bn_apply_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res,
                                                                 const float* __restrict__ scale, const float* __restrict__ shift,
                                                                 bf16* __restrict__ y, unsigned char* __restrict__ mask, size_t nvec,
                                                                 int vec_cols) {
    const int vc = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) % vec_cols);
    float sc[8], sf[8];
    *reinterpret_cast<float4*>(&sc[0]) = *reinterpret_cast<const float4*>(scale + vc * 8);
    *reinterpret_cast<float4*>(&sc[4]) = *reinterpret_cast<const float4*>(scale + vc * 8 + 4);
    *reinterpret_cast<float4*>(&sf[0]) = *reinterpret_cast<const float4*>(shift + vc * 8);
    *reinterpret_cast<float4*>(&sf[4]) = *reinterpret_cast<const float4*>(shift + vc * 8 + 4);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * kUnroll) {
        uint4 u[kUnroll], r[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                u[k] = ld_stream_v4(x + (v0 + k * stride) * 8);
                if (kRes) r[k] = ld_stream_v4(res + (v0 + k * stride) * 8);
            }
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                float f[8];
                bf16x8_to_float(u[k], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = f[j] * sc[j] + sf[j];
                if (kRes) {
                    float g[8];
                    bf16x8_to_float(r[k], g);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] += g[j];
                }
                if (kRelu) {
                    // one byte per vector: bit j = "channel j passed the ReLU"; backward reads this instead of y
                    unsigned m = 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        m |= (f[j] > 0.f ? 1u : 0u) << j;
                        f[j] = fmaxf(f[j], 0.f);
                    }
                    if (mask) mask[v0 + k * stride] = (unsigned char)m;
                }
                st_global_v4(y + (v0 + k * stride) * 8, float_to_bf16x8(f));
            }
    }
}

// backward pass 1: g = (dy [+ dy2]) * relu_mask; partial sum(g), sum(g * (x - mean)); optionally write g.
// The ReLU mask is the byte-per-vector bitmask written by the forward (kMask) or, for callers that did not
// keep one, recomputed from the saved output y.  dy2 (nullable) is a second gradient of the same tensor -
// the other branch of a residual fork - summed here instead of by a separate elementwise kernel.
template <bool kRelu, bool kWriteG, bool kMask>
__global__ void __launch_bounds__(kBnMaxThreads) bn_bwd_reduce_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ dy2,
                                                                      const bf16* __restrict__ y, const unsigned char* __restrict__ mask,
                                                                      const bf16* __restrict__ x, const float* __restrict__ mean,
                                                                      bf16* __restrict__ g_out, size_t nvec, int vec_cols, int C,
                                                                      float* __restrict__ partials) {
    extern __shared__ float sh[];
    const int vc = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) % vec_cols);
    float mu[8], s[8], q[8];
    *reinterpret_cast<float4*>(&mu[0]) = *reinterpret_cast<const float4*>(mean + vc * 8);
    *reinterpret_cast<float4*>(&mu[4]) = *reinterpret_cast<const float4*>(mean + vc * 8 + 4);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const bool add2 = dy2 != nullptr;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * kUnroll) {
        uint4 ud[kUnroll], ue[kUnroll], uy[kUnroll], ux[kUnroll];
        unsigned mk[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                ud[k] = ld_stream_v4(dy + (v0 + k * stride) * 8);
                if (add2) ue[k] = ld_stream_v4(dy2 + (v0 + k * stride) * 8);
                if (kRelu && !kMask) uy[k] = ld_stream_v4(y + (v0 + k * stride) * 8);
                if (kRelu && kMask) mk[k] = mask[v0 + k * stride];
                ux[k] = ld_stream_v4(x + (v0 + k * stride) * 8);
            }
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                float d[8], xx[8];
                bf16x8_to_float(ud[k], d);
                bf16x8_to_float(ux[k], xx);
                if (add2) {
                    float e[8];
                    bf16x8_to_float(ue[k], e);
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] += e[j];
                }
                if (kRelu && !kMask) {
                    float yy[8];
                    bf16x8_to_float(uy[k], yy);
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] = yy[j] > 0.f ? d[j] : 0.f;
                }
                if (kRelu && kMask) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] = (mk[k] >> j) & 1u ? d[j] : 0.f;
                }
                if (kWriteG) st_global_v4(g_out + (v0 + k * stride) * 8, float_to_bf16x8(d));
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    s[j] += d[j];
                    q[j] += d[j] * (xx[j] - mu[j]);
                }
            }
    }
    fold_to_partials(s, q, vec_cols, C, partials + (size_t)blockIdx.x * 2 * C, sh);
}

// dgamma, dbeta and the coefficients of pass 2:  dx = a * (g - b - (x - mean) * c)
// grid = ceil(C / 32), block = kFoldThreads
__global__ void __launch_bounds__(kFoldThreads) bn_bwd_finalize_kernel(const float* __restrict__ partials, int nblocks, long long M, int C,
                                       const bf16* __restrict__ gamma, const float* __restrict__ invstd, void* __restrict__ dgamma,
                                       void* __restrict__ dbeta, int grad_dtype, float* __restrict__ coef /*[3][C]*/) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    float sg, sgx;
    fold_partials32(partials, nblocks, C, c, c < C, sg, sgx);
    if (c >= C || threadIdx.x >= 32) return;
    const float is = invstd[c];
    const float g = gamma ? __bfloat162float(gamma[c]) : 1.f;
    const float dgam = sgx * is, dbet = sg;
    if (dgamma) {
        if (grad_dtype == DDL_F32) ((float*)dgamma)[c] = dgam; else ((bf16*)dgamma)[c] = __float2bfloat16(dgam);
    }
    if (dbeta) {
        if (grad_dtype == DDL_F32) ((float*)dbeta)[c] = dbet; else ((bf16*)dbeta)[c] = __float2bfloat16(dbet);
    }
    const float inv_m = 1.f / (float)M;
    coef[c] = g * is;
    coef[C + c] = sg * inv_m;
    coef[2 * C + c] = sgx * inv_m * is * is;
}

// backward pass 2.  kFromG: g was materialised by pass 1 (residual blocks); else apply the ReLU mask
// (bitmask when kMask, recomputed from y otherwise) to dy again.
template <bool kRelu, bool kFromG, bool kMask>
__global__ void __launch_bounds__(kBnMaxThreads) bn_bwd_apply_kernel(const bf16* __restrict__ dy_or_g, const bf16* __restrict__ y,
                                                                     const unsigned char* __restrict__ mask, const bf16* __restrict__ x,
                                                                     const float* __restrict__ mean, const float* __restrict__ coef,
                                                                     bf16* __restrict__ dx, size_t nvec, int vec_cols, int C) {
    const int vc = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) % vec_cols);
    float mu[8], a[8], b[8], cc[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        *reinterpret_cast<float4*>(&mu[4 * h]) = *reinterpret_cast<const float4*>(mean + vc * 8 + 4 * h);
        *reinterpret_cast<float4*>(&a[4 * h]) = *reinterpret_cast<const float4*>(coef + vc * 8 + 4 * h);
        *reinterpret_cast<float4*>(&b[4 * h]) = *reinterpret_cast<const float4*>(coef + C + vc * 8 + 4 * h);
        *reinterpret_cast<float4*>(&cc[4 * h]) = *reinterpret_cast<const float4*>(coef + 2 * C + vc * 8 + 4 * h);
    }
    constexpr bool kNeedMask = kRelu && !kFromG;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * kUnroll) {
        uint4 ud[kUnroll], uy[kUnroll], ux[kUnroll];
        unsigned mk[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                ud[k] = ld_stream_v4(dy_or_g + (v0 + k * stride) * 8);
                if (kNeedMask && !kMask) uy[k] = ld_stream_v4(y + (v0 + k * stride) * 8);
                if (kNeedMask && kMask) mk[k] = mask[v0 + k * stride];
                ux[k] = ld_stream_v4(x + (v0 + k * stride) * 8);
            }
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                float d[8], xx[8], o[8];
                bf16x8_to_float(ud[k], d);
                bf16x8_to_float(ux[k], xx);
                if (kNeedMask && !kMask) {
                    float yy[8];
                    bf16x8_to_float(uy[k], yy);
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] = yy[j] > 0.f ? d[j] : 0.f;
                }
                if (kNeedMask && kMask) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] = (mk[k] >> j) & 1u ? d[j] : 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = a[j] * (d[j] - b[j] - (xx[j] - mu[j]) * cc[j]);
                st_global_v4(dx + (v0 + k * stride) * 8, float_to_bf16x8(o));
            }
    }
}

// ---- stem: BatchNorm + ReLU + 3x3/2 max pool in one forward and two backward kernels --------------------------------------
// relu(maxpool(bn(x))) = maxpool(relu(bn(x))): the normalised 112 x 112 activation (411 MB at batch 256) is never written or
// read; the pooled tensor and a 4-bit code per pooled element (winning tap 0..8, or 15 = "pooled value is 0: no gradient")
// carry everything the backward needs.  Even H and W only (every window starts at an even pixel - 1).
//   forward : block = one output row of one image; z = x * scale + shift in fp32 for the 9 taps, lowest tap wins a tie
//   reduce  : elementwise over the pooled tensors: g = (dy + dy2) where y > 0; partial sum(g), sum(g * (x - mean)) per channel in
//             the layout of bn_bwd_reduce_kernel, with (x - mean) recovered from y (see the kernel)
//   apply   : thread = 2 x 2 input patch x 8 channels; the gradient of a pixel is the sum of g over the (at most four) windows
//             whose winning tap is that pixel; dx = a * (g - b - (x - mean) * c) with the coefficients of bn_bwd_finalize_kernel

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint32_t bf16x2_min(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("min.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
// 0xffff in each half where a == b
__device__ __forceinline__ uint32_t bf16x2_eq_mask(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("set.eq.u32.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// z = x * scale + shift is monotonic in x per channel, so the window's winner is the largest x where scale > 0 and the
// smallest x where scale < 0: maximum AND minimum are taken on packed bf16 pairs, the winner's tap is found by comparing the
// taps with it, last to first (lowest tap wins a tie, as ATen does on equal values), and z is evaluated once per element.
__global__ void __launch_bounds__(256) bn_relu_pool_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ scale,
                                                                const float* __restrict__ shift, bf16* __restrict__ y,
                                                                uint32_t* __restrict__ codes, int H, int W, int C, int OH, int OW) {
    const int vc = C >> 3, oh = blockIdx.x, n = blockIdx.y;
    const bf16* xin = x + (size_t)n * H * W * C;
    const size_t orow = ((size_t)n * OH + oh) * OW;
    const int c8 = threadIdx.x % vc;  // blockDim.x is a multiple of vc: the same for every item of this thread
    float sc[8], sf[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        *reinterpret_cast<float4*>(&sc[4 * h]) = *reinterpret_cast<const float4*>(scale + c8 * 8 + 4 * h);
        *reinterpret_cast<float4*>(&sf[4 * h]) = *reinterpret_cast<const float4*>(shift + c8 * 8 + 4 * h);
    }
    uint32_t use_max[4];  // 0xffff in the halves whose channel has scale >= 0
#pragma unroll
    for (int q = 0; q < 4; ++q) use_max[q] = (sc[2 * q] >= 0.f ? 0x0000ffffu : 0u) | (sc[2 * q + 1] >= 0.f ? 0xffff0000u : 0u);
    for (int t = threadIdx.x; t < OW * vc; t += blockDim.x) {
        const int ow = t / vc;
        const bool top = oh == 0, left = ow == 0;  // taps of row 0 / column 0 lie outside the image there
        uint4 v[9];
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                // an outside tap is replaced by the centre tap: a duplicate changes neither the maximum nor the minimum
                const int ih = oh * 2 - 1 + ((kh == 0 && top) ? 1 : kh), iw = ow * 2 - 1 + ((kw == 0 && left) ? 1 : kw);
                v[kh * 3 + kw] = ld_global_v4(xin + ((size_t)ih * W + iw) * C + c8 * 8);
            }
        float best[8];
        uint32_t packed = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t mx = reinterpret_cast<const uint32_t*>(&v[0])[q], mn = mx;
#pragma unroll
            for (int tap = 1; tap < 9; ++tap) {
                const uint32_t w = reinterpret_cast<const uint32_t*>(&v[tap])[q];
                mx = bf16x2_max(mx, w);
                mn = bf16x2_min(mn, w);
            }
            const uint32_t win = (mx & use_max[q]) | (mn & ~use_max[q]);
            uint32_t a = 0;
#pragma unroll
            for (int tap = 8; tap >= 0; --tap) {
                if ((tap < 3 && top) || (tap % 3 == 0 && left)) continue;  // not a tap of this window
                const uint32_t eq = bf16x2_eq_mask(reinterpret_cast<const uint32_t*>(&v[tap])[q], win);
                a = (a & ~eq) | ((uint32_t)tap * 0x00010001u & eq);
            }
            const float z0 = fmaf(__uint_as_float(win << 16), sc[2 * q], sf[2 * q]);
            const float z1 = fmaf(__uint_as_float(win & 0xffff0000u), sc[2 * q + 1], sf[2 * q + 1]);
            best[2 * q] = z0 > 0.f ? z0 : 0.f;
            best[2 * q + 1] = z1 > 0.f ? z1 : 0.f;
            packed |= ((z0 > 0.f ? (a & 0xfu) : 15u) | ((z1 > 0.f ? ((a >> 16) & 0xfu) : 15u) << 4)) << (8 * q);
        }
        const size_t o = orow + ow;
        st_global_v4(y + o * C + c8 * 8, float_to_bf16x8(best));
        codes[o * vc + c8] = packed;
    }
}

// The winning tap's x never has to be fetched: where a gradient passes, the pooled value IS the normalised winner,
// y = (x - mean) * invstd * gamma + beta, so (x - mean) = (y - beta) / (gamma * invstd) (y carries bf16 rounding, 2^-9
// relative - noise for a sum over ~10^5 pixels per channel; a channel with gamma == 0 contributes 0 to sum(g * (x - mean)),
// its dx is 0 anyway).  The pass is then elementwise over the POOLED tensors only.
__global__ void __launch_bounds__(kBnMaxThreads) bn_relu_pool_bwd_reduce_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ dy2,
                                                                                 const bf16* __restrict__ y, const bf16* __restrict__ gamma,
                                                                                 const bf16* __restrict__ beta, const float* __restrict__ invstd,
                                                                                 float* __restrict__ partials, size_t nvec, int vec_cols,
                                                                                 int C) {
    extern __shared__ float sh[];
    const int vc = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) % vec_cols);
    float bt[8], rs[8], s[8], q[8];
    {
        float gm[8], is[8];
        bf16x8_to_float(ld_global_v4(gamma + vc * 8), gm);
        bf16x8_to_float(ld_global_v4(beta + vc * 8), bt);
#pragma unroll
        for (int h = 0; h < 2; ++h) *reinterpret_cast<float4*>(&is[4 * h]) = *reinterpret_cast<const float4*>(invstd + vc * 8 + 4 * h);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float d = gm[j] * is[j];
            rs[j] = d != 0.f ? 1.f / d : 0.f;
            s[j] = q[j] = 0.f;
        }
    }
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += stride * kUnroll) {
        uint4 ud[kUnroll], u2[kUnroll], uy[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                ud[k] = ld_stream_v4(dy + (v0 + k * stride) * 8);
                if (dy2) u2[k] = ld_stream_v4(dy2 + (v0 + k * stride) * 8);
                uy[k] = ld_stream_v4(y + (v0 + k * stride) * 8);
            }
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
            if (v0 + k * stride < nvec) {
                float g[8], yy[8];
                bf16x8_to_float(ud[k], g);
                bf16x8_to_float(uy[k], yy);
                if (dy2) {
                    float g2[8];
                    bf16x8_to_float(u2[k], g2);
#pragma unroll
                    for (int j = 0; j < 8; ++j) g[j] += g2[j];
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float gj = yy[j] > 0.f ? g[j] : 0.f;
                    s[j] += gj;
                    q[j] = fmaf(gj, yy[j] - bt[j], q[j]);
                }
            }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] *= rs[j];
    fold_to_partials(s, q, vec_cols, C, partials + (size_t)blockIdx.x * 2 * C, sh);
}

__global__ void __launch_bounds__(256, 2) bn_relu_pool_bwd_apply_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ dy2,
                                                                      const uint32_t* __restrict__ codes, const bf16* __restrict__ x,
                                                                      const float* __restrict__ mean, const float* __restrict__ coef,
                                                                      bf16* __restrict__ dx, int H, int W, int C, int OH, int OW) {
    const int vc = C >> 3, a_ = blockIdx.x, n = blockIdx.y;  // input rows 2 a_, 2 a_ + 1
    const int c8 = threadIdx.x % vc;
    // dx = a * (g - b - (x - mean) * c) = ka * g + kx * x + k0   (three coefficient vectors in registers instead of four)
    float ka[8], kx[8], k0[8];
    {
        float mu[8], kb[8], kc[8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            *reinterpret_cast<float4*>(&mu[4 * h]) = *reinterpret_cast<const float4*>(mean + c8 * 8 + 4 * h);
            *reinterpret_cast<float4*>(&ka[4 * h]) = *reinterpret_cast<const float4*>(coef + c8 * 8 + 4 * h);
            *reinterpret_cast<float4*>(&kb[4 * h]) = *reinterpret_cast<const float4*>(coef + C + c8 * 8 + 4 * h);
            *reinterpret_cast<float4*>(&kc[4 * h]) = *reinterpret_cast<const float4*>(coef + 2 * C + c8 * 8 + 4 * h);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            kx[e] = -ka[e] * kc[e];
            k0[e] = ka[e] * (kc[e] * mu[e] - kb[e]);
        }
    }
    const bf16* xin = x + (size_t)n * H * W * C;
    bf16* dxn = dx + (size_t)n * H * W * C;
    for (int t = threadIdx.x; t < OW * vc; t += blockDim.x) {
        const int b = t / vc;  // input columns 2b, 2b + 1
        uint4 xv[4];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) xv[i * 2 + j] = ld_stream_v4(xin + ((size_t)(2 * a_ + i) * W + 2 * b + j) * C + c8 * 8);
        // all loads of the item first (the four windows' codes and gradients are independent of each other and of x):
        // with the loads behind per-window branches this kernel ran latency-bound at 2.4x its memory time
        uint32_t cdw[4];
        uint4 gv[4], g2v[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int oh = a_ + (w >> 1), ow = b + (w & 1);
            const bool ok = oh < OH && ow < OW;
            const size_t o = ((size_t)n * OH + (ok ? oh : a_)) * OW + (ok ? ow : b);
            const uint32_t c = codes[o * vc + c8];
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
            const uint32_t cd = cdw[w];
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
                    const uint32_t t_ = cd ^ ((uint32_t)((i + 1 - 2 * di) * 3 + (j + 1 - 2 * dj)) * 0x11111111u);
                    const uint32_t m = ~(((t_ & 0x77777777u) + 0x77777777u) | t_) & 0x88888888u;
#pragma unroll
                    for (int e = 0; e < 8; ++e)
                        if (m & (8u << (4 * e))) acc[i * 2 + j][e] += g[e];
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float xx[8], o8[8];
            bf16x8_to_float(xv[i], xx);
#pragma unroll
            for (int e = 0; e < 8; ++e) o8[e] = fmaf(ka[e], acc[i][e], fmaf(kx[e], xx[e], k0[e]));
            st_global_v4(dxn + ((size_t)(2 * a_ + (i >> 1)) * W + 2 * b + (i & 1)) * C + c8 * 8, float_to_bf16x8(o8));
        }
    }
}

bool bn_pool_ok(int N, int H, int W, int C) {
    return N >= 1 && N <= 65535 && H >= 2 && W >= 2 && H % 2 == 0 && W % 2 == 0 && C % 8 == 0 && C >= 8 && C <= 2048 && 256 % (C / 8) == 0;
}

size_t smem_bytes(const BnGeom& g) { return (size_t)g.threads * 17 * sizeof(float); }

}  // namespace

// scratch layout: [grid][2][C] partials, then [3][C] coefficients
DDL_API size_t ddl_bn_scratch_bytes(long long M, int C) {
    BnGeom g;
    if (make_geom(M, C, &g)) return 0;
    return ((size_t)g.grid * 2 * C + 3 * (size_t)C) * sizeof(float);
}

static int bn_apply_impl(const void* x, const void* residual, const float* scale, const float* shift, void* y, void* relu_mask,
                         long long M, int C, int relu, void* stream_) {
    BnGeom g;
    if (int e = make_geom(M, C, &g)) return e;
    cudaStream_t s = (cudaStream_t)stream_;
    const bf16 *xp = (const bf16*)x, *rp = (const bf16*)residual;
    bf16* yp = (bf16*)y;
    unsigned char* mp = (unsigned char*)relu_mask;
    if (relu && rp) bn_apply_kernel<true, true><<<g.grid, g.threads, 0, s>>>(xp, rp, scale, shift, yp, mp, g.nvec, g.vec_cols);
    else if (relu) bn_apply_kernel<true, false><<<g.grid, g.threads, 0, s>>>(xp, rp, scale, shift, yp, mp, g.nvec, g.vec_cols);
    else if (rp) bn_apply_kernel<false, true><<<g.grid, g.threads, 0, s>>>(xp, rp, scale, shift, yp, nullptr, g.nvec, g.vec_cols);
    else bn_apply_kernel<false, false><<<g.grid, g.threads, 0, s>>>(xp, rp, scale, shift, yp, nullptr, g.nvec, g.vec_cols);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_bn_apply(const void* x, const void* residual, const float* scale, const float* shift, void* y, long long M, int C,
                         int relu, void* stream_) {
    return bn_apply_impl(x, residual, scale, shift, y, nullptr, M, C, relu, stream_);
}

DDL_API int ddl_bn_act_fwd_train_mask(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                      float* running_var, void* y, void* relu_mask, float* mean, float* invstd, float* scale, float* shift,
                                      void* scratch, size_t scratch_bytes, long long M, int C, float eps, float momentum, int relu,
                                      void* stream_) {
    BnGeom g;
    if (int e = make_geom(M, C, &g)) return e;
    DDL_REQUIRE(x && y && mean && invstd && scale && shift && scratch, "bn fwd: null pointer");
    DDL_REQUIRE(scratch_bytes >= ddl_bn_scratch_bytes(M, C), "bn fwd: scratch too small");
    cudaStream_t s = (cudaStream_t)stream_;
    float* partials = (float*)scratch;
    bn_stats_kernel<<<g.grid, g.threads, smem_bytes(g), s>>>((const bf16*)x, g.nvec, g.vec_cols, C, partials);
    DDL_LAUNCH_CHECK();
    bn_fwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(partials, g.grid, M, C, (const bf16*)gamma, (const bf16*)beta, running_mean,
                                                               running_var, momentum, eps, mean, invstd, scale, shift);
    DDL_LAUNCH_CHECK();
    return bn_apply_impl(x, residual, scale, shift, y, relu ? relu_mask : nullptr, M, C, relu, stream_);
}

DDL_API int ddl_bn_act_fwd_train(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                 float* running_var, void* y, float* mean, float* invstd, float* scale, float* shift, void* scratch,
                                 size_t scratch_bytes, long long M, int C, float eps, float momentum, int relu, void* stream_) {
    return ddl_bn_act_fwd_train_mask(x, residual, gamma, beta, running_mean, running_var, y, nullptr, mean, invstd, scale, shift, scratch,
                                     scratch_bytes, M, C, eps, momentum, relu, stream_);
}

// statistics already available as per-block partials [nblocks][2][C] (e.g. written by the stem convolution's epilogue)
DDL_API int ddl_bn_act_fwd_train_from_partials(const void* x, const void* residual, const void* gamma, const void* beta,
                                               float* running_mean, float* running_var, void* y, void* relu_mask, float* mean,
                                               float* invstd, float* scale, float* shift, const float* partials, int nblocks,
                                               long long M, int C, float eps, float momentum, int relu, void* stream_) {
    BnGeom g;
    if (int e = make_geom(M, C, &g)) return e;
    DDL_REQUIRE(x && y && mean && invstd && scale && shift && partials && nblocks >= 1, "bn fwd (partials): null pointer");
    cudaStream_t s = (cudaStream_t)stream_;
    bn_fwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(partials, nblocks, M, C, (const bf16*)gamma, (const bf16*)beta,
                                                                        running_mean, running_var, momentum, eps, mean, invstd, scale, shift);
    DDL_LAUNCH_CHECK();
    return bn_apply_impl(x, residual, scale, shift, y, relu ? relu_mask : nullptr, M, C, relu, stream_);
}

DDL_API int ddl_bn_act_fwd_train_from_sums(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                           float* running_var, void* y, float* mean, float* invstd, float* scale, float* shift,
                                           const float* sums, long long M, int C, float eps, float momentum, int relu, void* stream_) {
    BnGeom g;
    if (int e = make_geom(M, C, &g)) return e;
    DDL_REQUIRE(x && y && mean && invstd && scale && shift && sums, "bn fwd (from sums): null pointer");
    cudaStream_t s = (cudaStream_t)stream_;
    // `sums` has the layout of ONE partial block: [2][C]
    bn_fwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(sums, 1, M, C, (const bf16*)gamma, (const bf16*)beta, running_mean, running_var,
                                                              momentum, eps, mean, invstd, scale, shift);
    DDL_LAUNCH_CHECK();
    return ddl_bn_apply(x, residual, scale, shift, y, M, C, relu, stream_);
}

DDL_API int ddl_bn_act_fwd_eval(const void* x, const void* residual, const void* gamma, const void* beta, const float* running_mean,
                                const float* running_var, void* y, float* scale, float* shift, long long M, int C, float eps, int relu,
                                void* stream_) {
    DDL_REQUIRE(x && y && running_mean && running_var && scale && shift, "bn eval: null pointer");
    bn_eval_coeff_kernel<<<ddl_ceil_div(C, 128), 128, 0, (cudaStream_t)stream_>>>(C, (const bf16*)gamma, (const bf16*)beta, running_mean,
                                                                                 running_var, eps, scale, shift);
    DDL_LAUNCH_CHECK();
    return ddl_bn_apply(x, residual, scale, shift, y, M, C, relu, stream_);
}

DDL_API int ddl_bn_act_bwd_ex(const void* dy, const void* dy2, const void* y, const void* relu_mask, const void* x, const void* gamma,
                              const float* mean, const float* invstd, void* dx, void* dres, void* dgamma, void* dbeta,
                              int param_grad_dtype, void* scratch, size_t scratch_bytes, long long M, int C, int relu, void* stream_) {
    BnGeom g;
    if (int e = make_geom(M, C, &g)) return e;
    DDL_REQUIRE(dy && x && mean && invstd && dx && scratch, "bn bwd: null pointer");
    DDL_REQUIRE(!relu || y || relu_mask, "bn bwd: relu needs the saved output or the relu bitmask of the forward");
    DDL_REQUIRE(scratch_bytes >= ddl_bn_scratch_bytes(M, C), "bn bwd: scratch too small");
    // g must be materialised when it is itself an output (gradient of the residual branch)
    const bool write_g = dres != nullptr && relu;
    DDL_REQUIRE(!dy2 || write_g, "bn bwd: a second incoming gradient is only fused for relu + residual layers");
    cudaStream_t s = (cudaStream_t)stream_;
    float* partials = (float*)scratch;
    float* coef = partials + (size_t)g.grid * 2 * C;
    const bf16 *dyp = (const bf16*)dy, *dy2p = (const bf16*)dy2, *yp = (const bf16*)y, *xp = (const bf16*)x;
    const unsigned char* mp = (const unsigned char*)relu_mask;
    const size_t sm = smem_bytes(g);
#define BN_REDUCE(R, W, K, GOUT) \
    bn_bwd_reduce_kernel<R, W, K><<<g.grid, g.threads, sm, s>>>(dyp, dy2p, yp, mp, xp, mean, GOUT, g.nvec, g.vec_cols, C, partials)
    if (write_g && mp) BN_REDUCE(true, true, true, (bf16*)dres);
    else if (write_g) BN_REDUCE(true, true, false, (bf16*)dres);
    else if (relu && mp) BN_REDUCE(true, false, true, nullptr);
    else if (relu) BN_REDUCE(true, false, false, nullptr);
    else BN_REDUCE(false, false, false, nullptr);
#undef BN_REDUCE
    DDL_LAUNCH_CHECK();
    bn_bwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(partials, g.grid, M, C, (const bf16*)gamma, invstd, dgamma, dbeta,
                                                               param_grad_dtype, coef);
    DDL_LAUNCH_CHECK();
#define BN_APPLY(R, F, K, SRC) \
    bn_bwd_apply_kernel<R, F, K><<<g.grid, g.threads, 0, s>>>(SRC, yp, mp, xp, mean, coef, (bf16*)dx, g.nvec, g.vec_cols, C)
    if (write_g) BN_APPLY(true, true, false, (const bf16*)dres);
    else if (relu && mp) BN_APPLY(true, false, true, dyp);
    else if (relu) BN_APPLY(true, false, false, dyp);
    else BN_APPLY(false, false, false, dyp);
#undef BN_APPLY
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_bn_act_bwd(const void* dy, const void* y, const void* x, const void* gamma, const float* mean, const float* invstd,
                           void* dx, void* dres, void* dgamma, void* dbeta, int param_grad_dtype, void* scratch, size_t scratch_bytes,
                           long long M, int C, int relu, void* stream_) {
    return ddl_bn_act_bwd_ex(dy, nullptr, y, nullptr, x, gamma, mean, invstd, dx, dres, dgamma, dbeta, param_grad_dtype, scratch,
                             scratch_bytes, M, C, relu, stream_);
}

// ---- stem: BatchNorm + ReLU + max pool (see the kernels above) ----------------------------------------------------------
DDL_API int ddl_bn_relu_maxpool_supported(int N, int H, int W, int C) { return bn_pool_ok(N, H, W, C) ? 1 : 0; }

// scratch of the backward: the partials + coefficients of a BatchNorm backward over the pooled tensor
DDL_API size_t ddl_bn_relu_maxpool_bwd_scratch_bytes(int N, int H, int W, int C) {
    if (!bn_pool_ok(N, H, W, C)) return 0;
    return ddl_bn_scratch_bytes((long long)N * (H / 2) * (W / 2), C);
}

// x [N, H, W, C] bf16 = the convolution output; statistics as per-CTA partials [nblocks][2][C] (as for
// ddl_bn_act_fwd_train_from_partials); y [N, H/2, W/2, C] bf16 = maxpool3x3s2(relu(bn(x))); codes: N * H/2 * W/2 * C/2 bytes.
DDL_API int ddl_bn_relu_maxpool_fwd_train(const void* x, const void* gamma, const void* beta, float* running_mean, float* running_var,
                                          void* y, void* codes, float* mean, float* invstd, float* scale, float* shift,
                                          const float* partials, int nblocks, int N, int H, int W, int C, float eps, float momentum,
                                          void* stream_) {
    DDL_REQUIRE(bn_pool_ok(N, H, W, C), "bn + relu + maxpool: unsupported shape (N %d, %d x %d, C %d)", N, H, W, C);
    DDL_REQUIRE(x && y && codes && mean && invstd && scale && shift && partials && nblocks >= 1, "bn + relu + maxpool fwd: null pointer");
    DDL_REQUIRE(((uintptr_t)codes & 3) == 0, "bn + relu + maxpool: code buffer must be 4-byte aligned");
    cudaStream_t s = (cudaStream_t)stream_;
    const long long M = (long long)N * H * W;
    bn_fwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(partials, nblocks, M, C, (const bf16*)gamma, (const bf16*)beta,
                                                                        running_mean, running_var, momentum, eps, mean, invstd, scale, shift);
    DDL_LAUNCH_CHECK();
    bn_relu_pool_fwd_kernel<<<dim3(H / 2, N), 256, 0, s>>>((const bf16*)x, scale, shift, (bf16*)y, (uint32_t*)codes, H, W, C, H / 2, W / 2);
    DDL_LAUNCH_CHECK();
    return 0;
}

// dy (+ optional dy2) and y (the forward's output) [N, H/2, W/2, C] -> dx [N, H, W, C], dgamma, dbeta
DDL_API int ddl_bn_relu_maxpool_bwd(const void* dy, const void* dy2, const void* y, const void* codes, const void* x, const void* gamma,
                                    const void* beta, const float* mean, const float* invstd, void* dx, void* dgamma, void* dbeta,
                                    int param_grad_dtype, void* scratch, size_t scratch_bytes, int N, int H, int W, int C, void* stream_) {
    DDL_REQUIRE(bn_pool_ok(N, H, W, C), "bn + relu + maxpool: unsupported shape (N %d, %d x %d, C %d)", N, H, W, C);
    DDL_REQUIRE(dy && y && codes && x && gamma && beta && mean && invstd && dx && scratch, "bn + relu + maxpool bwd: null pointer");
    DDL_REQUIRE(scratch_bytes >= ddl_bn_relu_maxpool_bwd_scratch_bytes(N, H, W, C), "bn + relu + maxpool bwd: scratch too small");
    cudaStream_t s = (cudaStream_t)stream_;
    const int OH = H / 2, OW = W / 2;
    BnGeom g;
    if (int e = make_geom((long long)N * OH * OW, C, &g)) return e;
    float* partials = (float*)scratch;
    float* coef = partials + (size_t)g.grid * 2 * C;
    bn_relu_pool_bwd_reduce_kernel<<<g.grid, g.threads, smem_bytes(g), s>>>((const bf16*)dy, (const bf16*)dy2, (const bf16*)y, (const bf16*)gamma,
                                                                            (const bf16*)beta, invstd, partials, g.nvec, g.vec_cols, C);
    DDL_LAUNCH_CHECK();
    bn_bwd_finalize_kernel<<<ddl_ceil_div(C, 32), kFoldThreads, 0, s>>>(partials, g.grid, (long long)N * H * W, C, (const bf16*)gamma, invstd,
                                                                        dgamma, dbeta, param_grad_dtype, coef);
    DDL_LAUNCH_CHECK();
    bn_relu_pool_bwd_apply_kernel<<<dim3(H / 2, N), 256, 0, s>>>((const bf16*)dy, (const bf16*)dy2, (const uint32_t*)codes, (const bf16*)x, mean,
                                                                 coef, (bf16*)dx, H, W, C, OH, OW);
    DDL_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant of ddl_bn_act_fwd_train.
//
// Who needs it: a host program that has handed the whole GPU to its own pre-sized arena (a
// serving process, an embedded runtime with a fixed memory budget) has nothing left to
// cudaMalloc from, and a latency-critical loop cannot afford the device-wide
// synchronisation that cudaMalloc / cudaFree imply.  Such callers can provide inputs and
// outputs but no scratch buffer.  For them the cross-block partial sums live in the
// file-scope array below, which is part of the loaded module: every GPU that loads the
// library holds DDL_BN_NOALLOC_MAX_BYTES (2 MiB) of it until the process exits.
//
// Rules for the caller:
//  * The first call on each device resolves the array's device address with
//    cudaGetSymbolAddress (the name `g_bn_noalloc_scratch` is only a host-side stand-in in
//    host code, not an address a kernel may be given).  That set-up call must be made with
//    the device current and should happen before any stream capture begins.
//  * Sizes whose scratch need (ddl_bn_scratch_bytes(M, C)) exceeds the array are rejected
//    with cudaErrorInvalidValue and a message in ddl_last_error().
//  * All calls on a device share the one array, and it keeps whatever the last call left
//    in it.  Nothing has to start at zero: every word a call reads (per-block partial sums,
//    then the folded coefficients) is written earlier in the same call, so stale contents
//    are harmless.  Two calls that overlap on the GPU (different streams, or different host
//    threads) race on the array and corrupt each other's statistics: serialise them.
//  * No cudaMalloc / cudaFree / allocating library routine is called on this path.
#define DDL_BN_NOALLOC_MAX_FLOATS (DDL_BN_NOALLOC_MAX_BYTES / sizeof(float))

static __device__ __align__(256) float g_bn_noalloc_scratch[DDL_BN_NOALLOC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
void* g_bn_noalloc_ptr[kMaxDevices] = {};  // per-device address of g_bn_noalloc_scratch
}  // namespace

DDL_API int ddl_bn_act_fwd_train_noalloc(const void* x, const void* residual, const void* gamma, const void* beta,
                                         float* running_mean, float* running_var, void* y, float* mean, float* invstd, float* scale,
                                         float* shift, long long M, int C, float eps, float momentum, int relu, void* stream_) {
    const size_t need = ddl_bn_scratch_bytes(M, C);
    DDL_REQUIRE(need != 0, "bn noalloc: unsupported shape M=%lld C=%d (C must be a multiple of 8, <= 2048; M >= 1)", M, C);
    DDL_REQUIRE(need <= (size_t)DDL_BN_NOALLOC_MAX_BYTES,
                "bn noalloc: M=%lld C=%d needs %zu bytes of scratch, more than the built-in %zu; use ddl_bn_act_fwd_train with a "
                "caller-provided scratch buffer",
                M, C, need, (size_t)DDL_BN_NOALLOC_MAX_BYTES);
    int dev = -1;
    DDL_CHECK(cudaGetDevice(&dev));
    DDL_REQUIRE(dev >= 0 && dev < kMaxDevices, "bn noalloc: device ordinal %d out of range", dev);
    void* ptr = __atomic_load_n(&g_bn_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (ptr == nullptr) {
        // once per device, with that device current; idempotent, so a race between two host
        // threads only resolves the same address twice
        DDL_CHECK(cudaGetSymbolAddress((void**)&ptr, g_bn_noalloc_scratch));
        DDL_REQUIRE(ptr != nullptr, "bn noalloc: cudaGetSymbolAddress returned a null address");
        __atomic_store_n(&g_bn_noalloc_ptr[dev], ptr, __ATOMIC_RELEASE);
    }
    return ddl_bn_act_fwd_train(x, residual, gamma, beta, running_mean, running_var, y, mean, invstd, scale, shift, ptr,
                                (size_t)DDL_BN_NOALLOC_MAX_BYTES, M, C, eps, momentum, relu, stream_);
}
