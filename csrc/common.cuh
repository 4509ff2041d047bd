// Shared device/host helpers for every translation unit of libb200ddl.
// sm_100a only (no multi-arch dispatch).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DDL_API extern "C" __attribute__((visibility("default")))

#define DDL_MAX_WORLD 8
#define DDL_SM_COUNT 148

// ----------------------------------------------------------------------------- errors
void ddl_set_error(const char* fmt, ...);  // runtime.cu

#define DDL_CHECK(expr)                                                                  \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            ddl_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                              \
        }                                                                                \
    } while (0)

#define DDL_REQUIRE(cond, ...)                                                           \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            ddl_set_error(__VA_ARGS__);                                                  \
            return (int)cudaErrorInvalidValue;                                           \
        }                                                                                \
    } while (0)

#define DDL_LAUNCH_CHECK() DDL_CHECK(cudaGetLastError())

// ----------------------------------------------------------------------------- vectors
typedef __nv_bfloat16 bf16;

struct alignas(16) Vec16 {
    uint4 u;
};

__device__ __forceinline__ uint4 ld_global_v4(const void* p) {
    return *reinterpret_cast<const uint4*>(p);
}
// streaming (read-once) 16 B load that does not allocate in L1
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// weak (coalescing) 16 B load that skips L1 allocation: bulk peer-memory reads.  A warp's 32 x 16 B
// become 128-byte NVLink requests; the strong .sys form below issues one request per lane.
__device__ __forceinline__ uint4 ld_weak_noalloc_v4(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
// volatile-ish 16 B load for peer memory written during this kernel's lifetime
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
    uint4 r;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_global_v4(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
__device__ __forceinline__ void st_stream_v4(void* p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ void bf16x8_to_float(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(h[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 float_to_bf16x8(const float (&f)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return u;
}

// ----------------------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ----------------------------------------------------------------------------- RNG
// Stateless counter-based hash (squares-style) so that forward and backward regenerate the
// same dropout mask from (seed, element index) without storing it.
__device__ __forceinline__ uint32_t hash_u32(uint64_t seed, uint64_t idx) {
    uint64_t x = (idx + 1) * 0x9E3779B97F4A7C15ull ^ seed;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return (uint32_t)x;
}
// keep-probability test on 8 consecutive elements: idx8 = index of the 8-element group
__device__ __forceinline__ uint32_t keep_mask8(uint64_t seed, uint64_t idx8, uint32_t drop_threshold_u8) {
    // two 32-bit hashes -> 8 bytes -> 8 independent Bernoulli draws with resolution 1/256
    uint32_t a = hash_u32(seed, idx8 * 2), b = hash_u32(seed, idx8 * 2 + 1);
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m |= (((a >> (8 * i)) & 0xffu) >= drop_threshold_u8 ? 1u : 0u) << i;
        m |= (((b >> (8 * i)) & 0xffu) >= drop_threshold_u8 ? 1u : 0u) << (4 + i);
    }
    return m;
}

static inline int ddl_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
