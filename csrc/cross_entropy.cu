// Fused cross-entropy: log-softmax + NLL loss + gradient in ONE pass over the logits
// (K12 in SURVEY.md §2.E.3).  1000 classes for ResNet-50, 10 for the CIFAR-shape config,
// 2-few for BERT classification.  One block per row; the row is read once, kept in
// registers/shared memory, and d(logits) = (softmax - onehot) * grad_scale is written in
// the logits' dtype.  Row losses go to a [B] fp32 buffer that a one-block kernel averages
// in a fixed order (deterministic).  The reference computes a softmax only for display
// (notebooks/cv/onnx_experiments.py:85-88, :174); it has no loss.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

constexpr int kCeThreads = 128;
constexpr int kMaxPerThread = 16;  // classes <= 128 * 16 * ... handled by the strided loop below

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f(float v);
template <>
__device__ __forceinline__ float from_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16(v); }

__device__ __forceinline__ float block_reduce(float v, bool is_max, float* sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = is_max ? warp_max(v) : warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    float r = sh[0];
#pragma unroll
    for (int w = 1; w < kCeThreads / 32; ++w) r = is_max ? fmaxf(r, sh[w]) : r + sh[w];
    return r;
}

template <typename T>
__global__ void __launch_bounds__(kCeThreads) ce_fwd_bwd_kernel(const T* __restrict__ logits, const long long* __restrict__ labels,
                                                                float* __restrict__ row_loss, T* __restrict__ dlogits, int C,
                                                                float grad_scale, long long ignore_index) {
    __shared__ float sh[kCeThreads / 32];
    const long long row = blockIdx.x;
    const T* x = logits + row * C;
    const long long label = labels[row];
    float vals[kMaxPerThread];
    float mx = -INFINITY;
    int cnt = 0;
    for (int c = threadIdx.x; c < C && cnt < kMaxPerThread; c += kCeThreads, ++cnt) {
        vals[cnt] = to_f<T>(x[c]);
        mx = fmaxf(mx, vals[cnt]);
    }
    // rows wider than kCeThreads * kMaxPerThread: stream the tail from memory
    for (int c = threadIdx.x + kCeThreads * kMaxPerThread; c < C; c += kCeThreads) mx = fmaxf(mx, to_f<T>(x[c]));
    mx = block_reduce(mx, true, sh);
    float se = 0.f;
    cnt = 0;
    for (int c = threadIdx.x; c < C && cnt < kMaxPerThread; c += kCeThreads, ++cnt) {
        vals[cnt] = __expf(vals[cnt] - mx);
        se += vals[cnt];
    }
    for (int c = threadIdx.x + kCeThreads * kMaxPerThread; c < C; c += kCeThreads) se += __expf(to_f<T>(x[c]) - mx);
    se = block_reduce(se, false, sh);
    const float inv = 1.f / se;
    const bool ignored = label == ignore_index;
    if (threadIdx.x == 0) {
        float l = 0.f;
        if (!ignored) l = -(to_f<T>(x[label]) - mx - __logf(se));
        row_loss[row] = l;
    }
    if (dlogits) {
        T* d = dlogits + row * C;
        const float gs = ignored ? 0.f : grad_scale;
        cnt = 0;
        for (int c = threadIdx.x; c < C && cnt < kMaxPerThread; c += kCeThreads, ++cnt)
            d[c] = from_f<T>((vals[cnt] * inv - (c == label ? 1.f : 0.f)) * gs);
        for (int c = threadIdx.x + kCeThreads * kMaxPerThread; c < C; c += kCeThreads)
            d[c] = from_f<T>((__expf(to_f<T>(x[c]) - mx) * inv - (c == label ? 1.f : 0.f)) * gs);
    }
}

__global__ void mean_kernel(const float* __restrict__ v, int n, float scale, float* __restrict__ out) {
    __shared__ float sh[256 / 32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += 256) acc += v[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += sh[w];
        *out = t * scale;
    }
}

}  // namespace

DDL_API int ddl_cross_entropy(const void* logits, int dtype, const long long* labels, float* row_loss, float* mean_loss,
                              void* dlogits, int rows, int classes, float grad_scale, long long ignore_index, void* stream_) {
    DDL_REQUIRE(logits && labels && row_loss, "cross_entropy: null pointer");
    DDL_REQUIRE(classes >= 1, "cross_entropy: classes must be >= 1");
    if (rows == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream_;
    if (dtype == DDL_BF16)
        ce_fwd_bwd_kernel<bf16><<<rows, kCeThreads, 0, s>>>((const bf16*)logits, labels, row_loss, (bf16*)dlogits, classes, grad_scale, ignore_index);
    else if (dtype == DDL_F32)
        ce_fwd_bwd_kernel<float><<<rows, kCeThreads, 0, s>>>((const float*)logits, labels, row_loss, (float*)dlogits, classes, grad_scale, ignore_index);
    else
        DDL_REQUIRE(false, "cross_entropy: bad dtype");
    DDL_LAUNCH_CHECK();
    if (mean_loss) {
        mean_kernel<<<1, 256, 0, s>>>(row_loss, rows, 1.f / (float)rows, mean_loss);
        DDL_LAUNCH_CHECK();
    }
    return 0;
}
