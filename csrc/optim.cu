// Fused AdamW over flat buffers (K13 in SURVEY.md §2.E.3; BASELINE.json:5 "fused AdamW").
//
// One launch updates a whole gradient bucket: reads the (already averaged) gradient in
// bf16 or fp32, updates fp32 master weights and both moments, and writes the bf16 compute
// copy of the parameters - 28 bytes of traffic per element, no per-tensor launches.
// Hyper-parameters live in device memory ({lr, beta1, beta2, eps, weight_decay, step}) so
// that a captured CUDA graph sees new values (LR schedule, step count) on every replay.
// The reference has no optimizer.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

template <typename TG>
__device__ __forceinline__ void load_grad8(const TG* g, size_t i8, float (&f)[8]);
template <>
__device__ __forceinline__ void load_grad8<bf16>(const bf16* g, size_t i8, float (&f)[8]) {
    bf16x8_to_float(ld_stream_v4(g + i8 * 8), f);
}
template <>
__device__ __forceinline__ void load_grad8<float>(const float* g, size_t i8, float (&f)[8]) {
    const uint4 a = ld_stream_v4(g + i8 * 8), b = ld_stream_v4(g + i8 * 8 + 4);
    f[0] = __uint_as_float(a.x); f[1] = __uint_as_float(a.y); f[2] = __uint_as_float(a.z); f[3] = __uint_as_float(a.w);
    f[4] = __uint_as_float(b.x); f[5] = __uint_as_float(b.y); f[6] = __uint_as_float(b.z); f[7] = __uint_as_float(b.w);
}

template <typename TG, typename TP>
__global__ void __launch_bounds__(256) adamw_kernel(const TG* __restrict__ grad, float* __restrict__ master, float* __restrict__ m,
                                                    float* __restrict__ v, TP* __restrict__ param_out, size_t n8,
                                                    const float* __restrict__ hyper, float grad_scale) {
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5];
    const float inv_bc1 = 1.f / (1.f - powf(b1, step));
    const float inv_sqrt_bc2 = rsqrtf(1.f - powf(b2, step));
    const float decay = 1.f - lr * wd;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        float g[8], w[8], mm[8], vv[8];
        load_grad8<TG>(grad, i, g);
        *reinterpret_cast<float4*>(&w[0]) = *reinterpret_cast<const float4*>(master + i * 8);
        *reinterpret_cast<float4*>(&w[4]) = *reinterpret_cast<const float4*>(master + i * 8 + 4);
        *reinterpret_cast<float4*>(&mm[0]) = *reinterpret_cast<const float4*>(m + i * 8);
        *reinterpret_cast<float4*>(&mm[4]) = *reinterpret_cast<const float4*>(m + i * 8 + 4);
        *reinterpret_cast<float4*>(&vv[0]) = *reinterpret_cast<const float4*>(v + i * 8);
        *reinterpret_cast<float4*>(&vv[4]) = *reinterpret_cast<const float4*>(v + i * 8 + 4);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gr = g[j] * grad_scale;
            mm[j] = b1 * mm[j] + (1.f - b1) * gr;
            vv[j] = b2 * vv[j] + (1.f - b2) * gr * gr;
            const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
            w[j] = w[j] * decay - lr * (mm[j] * inv_bc1) / denom;
        }
        *reinterpret_cast<float4*>(master + i * 8) = *reinterpret_cast<float4*>(&w[0]);
        *reinterpret_cast<float4*>(master + i * 8 + 4) = *reinterpret_cast<float4*>(&w[4]);
        *reinterpret_cast<float4*>(m + i * 8) = *reinterpret_cast<float4*>(&mm[0]);
        *reinterpret_cast<float4*>(m + i * 8 + 4) = *reinterpret_cast<float4*>(&mm[4]);
        *reinterpret_cast<float4*>(v + i * 8) = *reinterpret_cast<float4*>(&vv[0]);
        *reinterpret_cast<float4*>(v + i * 8 + 4) = *reinterpret_cast<float4*>(&vv[4]);
        if (param_out != nullptr) {
            if constexpr (sizeof(TP) == 2) {
                *reinterpret_cast<uint4*>(param_out + i * 8) = float_to_bf16x8(w);
            } else {
                *reinterpret_cast<float4*>(param_out + i * 8) = *reinterpret_cast<float4*>(&w[0]);
                *reinterpret_cast<float4*>(param_out + i * 8 + 4) = *reinterpret_cast<float4*>(&w[4]);
            }
        }
    }
}

__global__ void hyper_tick_kernel(float* hyper) { hyper[5] += 1.f; }

}  // namespace

DDL_API int ddl_adamw(const void* grad, int grad_dtype, float* master, float* m, float* v, void* param_out, int param_dtype,
                      size_t numel, const float* hyper, float grad_scale, void* stream_) {
    DDL_REQUIRE(numel % 8 == 0, "ddl_adamw: numel %zu must be a multiple of 8 (pad the flat buffer)", numel);
    DDL_REQUIRE(grad && master && m && v && hyper, "ddl_adamw: null pointer");
    DDL_REQUIRE((((uintptr_t)grad | (uintptr_t)master | (uintptr_t)m | (uintptr_t)v | (uintptr_t)param_out) & 15) == 0,
                "ddl_adamw: buffers must be 16-byte aligned");
    const size_t n8 = numel / 8;
    if (n8 == 0) return 0;
    int blocks = ddl_ceil_div((long long)n8, 256);
    const int cap = DDL_SM_COUNT * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t s = (cudaStream_t)stream_;
    if (grad_dtype == DDL_BF16 && param_dtype == DDL_BF16)
        adamw_kernel<bf16, bf16><<<blocks, 256, 0, s>>>((const bf16*)grad, master, m, v, (bf16*)param_out, n8, hyper, grad_scale);
    else if (grad_dtype == DDL_F32 && param_dtype == DDL_BF16)
        adamw_kernel<float, bf16><<<blocks, 256, 0, s>>>((const float*)grad, master, m, v, (bf16*)param_out, n8, hyper, grad_scale);
    else if (grad_dtype == DDL_F32 && param_dtype == DDL_F32)
        adamw_kernel<float, float><<<blocks, 256, 0, s>>>((const float*)grad, master, m, v, (float*)param_out, n8, hyper, grad_scale);
    else if (grad_dtype == DDL_BF16 && param_dtype == DDL_F32)
        adamw_kernel<bf16, float><<<blocks, 256, 0, s>>>((const bf16*)grad, master, m, v, (float*)param_out, n8, hyper, grad_scale);
    else
        DDL_REQUIRE(false, "ddl_adamw: bad dtypes");
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_adamw_tick(float* hyper, void* stream_) {
    DDL_REQUIRE(hyper != nullptr, "ddl_adamw_tick: null");
    hyper_tick_kernel<<<1, 1, 0, (cudaStream_t)stream_>>>(hyper);
    DDL_LAUNCH_CHECK();
    return 0;
}
