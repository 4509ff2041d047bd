// Peer-memory collectives for one NVSwitch domain (K1-K5, K7 in SURVEY.md §2.E.3).
//
// Every rank maps every peer's symmetric heap (runtime.cu / parallel/symm.py), so a
// collective is ONE kernel per rank that issues plain ld/st (or multimem.*) on peer
// addresses; NVLink 5 / NVSwitch moves the bytes.  No NCCL on this path (BASELINE.json:5).
//
// Synchronisation: per-block flag words in each rank's signal pad.  Block b of rank r
// announces "I reached barrier #e" by storing e into pad[peer][b][r] with st.release.sys and
// waits until pad[r][b][peer] >= e with ld.acquire.sys.  Epochs only grow (kept in device
// memory between launches), so nothing is ever reset and a captured CUDA graph can replay
// the kernels.  Blocks are independent of one another: no co-residency requirement, so
// the comm kernel can share the GPU with backward-pass kernels.
//
// Reduction is fused with the 1/world scale and the output cast (fp32|bf16 -> fp32|bf16),
// accumulating in fp32.
//
// The reference has no collective call site at all; the requirement comes
// from BASELINE.json:5.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

constexpr int kThreads = 512;

struct PeerPtrs {
    char* p[DDL_MAX_WORLD];
};

struct Ctx {
    uint32_t* pad[DDL_MAX_WORLD];  // signal pad of every rank (peer mapped), [blocks][DDL_MAX_WORLD]
    uint32_t* epoch;               // local: per-block barrier epoch, [blocks]
    int rank;
    int world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acquire_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// All threads of the block call this.  `epoch` is a per-thread copy of the block's epoch.
__device__ __forceinline__ void block_barrier(const Ctx& c, uint32_t& epoch, bool fence_peer_writes) {
    if (fence_peer_writes) __threadfence_system();
    __syncthreads();
    epoch += 1;
    if ((int)threadIdx.x < c.world) {
        const int peer = threadIdx.x;
        st_release_sys(c.pad[peer] + (size_t)blockIdx.x * DDL_MAX_WORLD + c.rank, epoch);
        const uint32_t* mine = c.pad[c.rank] + (size_t)blockIdx.x * DDL_MAX_WORLD + peer;
        // (polling with a relaxed load + one fence.acq_rel.sys afterwards measured 4 us SLOWER per
        // collective on B200 than acquire loads: the sys-scope fence is the expensive part)
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
        }
    }
    __syncthreads();
}

__device__ __forceinline__ uint32_t epoch_begin(const Ctx& c) { return c.epoch[blockIdx.x]; }
__device__ __forceinline__ void epoch_end(const Ctx& c, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = epoch;
}

// ---- 8-element granules ---------------------------------------------------------------
template <typename T>
struct G8;  // load / store 8 consecutive elements as floats
template <>
struct G8<float> {
    static constexpr int kBytes = 32;
    __device__ static void load(const char* base, size_t g, float (&f)[8]) {
        const uint4 a = ld_weak_noalloc_v4(base + g * 32), b = ld_weak_noalloc_v4(base + g * 32 + 16);
        f[0] = __uint_as_float(a.x); f[1] = __uint_as_float(a.y); f[2] = __uint_as_float(a.z); f[3] = __uint_as_float(a.w);
        f[4] = __uint_as_float(b.x); f[5] = __uint_as_float(b.y); f[6] = __uint_as_float(b.z); f[7] = __uint_as_float(b.w);
    }
    __device__ static void store(char* base, size_t g, const float (&f)[8]) {
        uint4 a, b;
        a.x = __float_as_uint(f[0]); a.y = __float_as_uint(f[1]); a.z = __float_as_uint(f[2]); a.w = __float_as_uint(f[3]);
        b.x = __float_as_uint(f[4]); b.y = __float_as_uint(f[5]); b.z = __float_as_uint(f[6]); b.w = __float_as_uint(f[7]);
        st_global_v4(base + g * 32, a);
        st_global_v4(base + g * 32 + 16, b);
    }
};
template <>
struct G8<bf16> {
    static constexpr int kBytes = 16;
    __device__ static void load(const char* base, size_t g, float (&f)[8]) {
        bf16x8_to_float(ld_weak_noalloc_v4(base + g * 16), f);
    }
    __device__ static void store(char* base, size_t g, const float (&f)[8]) {
        st_global_v4(base + g * 16, float_to_bf16x8(f));
    }
};

// ---- K1: one-shot all-reduce ----------------------------------------------------------
// Every rank reads the same granules from all peers, sums, and (after a barrier that proves
// every peer has finished reading) writes the result.  Latency-optimal: 2 barriers, one pass.
template <typename Tin, typename Tout, int UNR>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(Ctx c, PeerPtrs in, char* out, size_t n8, float scale) {
    uint32_t epoch = epoch_begin(c);
    float acc[UNR][8];
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    block_barrier(c, epoch, false);
#pragma unroll
    for (int k = 0; k < UNR; ++k) {
        const size_t g = g0 + k * stride;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
        if (g < n8) {
            // rank order 0..W-1 on every rank: all replicas get bit-identical sums (a staggered
            // order would let fp32 rounding differ between ranks and the model replicas drift)
            for (int peer = 0; peer < c.world; ++peer) {
                float f[8];
                G8<Tin>::load(in.p[peer], g, f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[k][j] += f[j];
            }
        }
    }
    block_barrier(c, epoch, false);  // all peers have read my input: in-place write is safe
#pragma unroll
    for (int k = 0; k < UNR; ++k) {
        const size_t g = g0 + k * stride;
        if (g < n8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[k][j] *= scale;
            G8<Tout>::store(out, g, acc[k]);
        }
    }
    epoch_end(c, epoch);
}

// ---- K2: two-shot all-reduce ----------------------------------------------------------
// Rank r owns slice r: pull-reduce it from all peers (inbound links), then push the
// result to all peers (outbound links).  Both NVLink directions are busy at once; bytes
// per GPU per direction = (W-1)/W * S, the bandwidth-optimal volume without switch help.
// A peer load takes ~2 us round trip, so bandwidth comes from bytes in flight: every thread
// first issues ALL loads of its UNR granules from ALL W ranks (UNR*W = 16 independent 16-byte
// requests for bf16), and only then starts adding.  W is a compile-time constant so that the
// load batch is fully unrolled (W = 0 selects the generic run-time loop).
template <typename T>
__device__ __forceinline__ void g8_load_raw(const char* base, size_t g, uint4 (&raw)[G8<T>::kBytes / 16]) {
#pragma unroll
    for (int i = 0; i < G8<T>::kBytes / 16; ++i) raw[i] = ld_weak_noalloc_v4(base + g * G8<T>::kBytes + i * 16);
}
template <typename T>
__device__ __forceinline__ void g8_accumulate(const uint4 (&raw)[G8<T>::kBytes / 16], float (&acc)[8]) {
    if constexpr (sizeof(T) == 2) {
        float f[8];
        bf16x8_to_float(raw[0], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
    } else {
        acc[0] += __uint_as_float(raw[0].x); acc[1] += __uint_as_float(raw[0].y); acc[2] += __uint_as_float(raw[0].z); acc[3] += __uint_as_float(raw[0].w);
        acc[4] += __uint_as_float(raw[1].x); acc[5] += __uint_as_float(raw[1].y); acc[6] += __uint_as_float(raw[1].z); acc[7] += __uint_as_float(raw[1].w);
    }
}

template <typename Tin, typename Tout, int W, int UNR>
__global__ void __launch_bounds__(kThreads) allreduce_twoshot_kernel(Ctx c, PeerPtrs in, PeerPtrs out, size_t n8, float scale) {
    uint32_t epoch = epoch_begin(c);
    const int world = W > 0 ? W : c.world;
    const size_t slice = n8 / world;  // host guarantees divisibility
    const size_t base = slice * c.rank;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    constexpr int kVec = G8<Tin>::kBytes / 16;
    block_barrier(c, epoch, false);
    for (size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g0 < slice; g0 += stride * UNR) {
        if constexpr (W > 0) {
            uint4 raw[UNR][W][kVec];
#pragma unroll
            for (int k = 0; k < UNR; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
#pragma unroll
                    for (int r = 0; r < W; ++r) {
                        int peer = c.rank + r;
                        if (peer >= W) peer -= W;
                        g8_load_raw<Tin>(in.p[peer], base + g, raw[k][r]);
                    }
                }
            }
            // reduce + push granule by granule: the raw registers die as they are consumed
#pragma unroll
            for (int k = 0; k < UNR; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
                    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r = 0; r < W; ++r) g8_accumulate<Tin>(raw[k][r], acc);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] *= scale;
#pragma unroll
                    for (int r = 0; r < W; ++r) {
                        int peer = c.rank + r;
                        if (peer >= W) peer -= W;
                        G8<Tout>::store(out.p[peer], base + g, acc);
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < UNR; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
                    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    for (int r = 0; r < world; ++r) {
                        int peer = c.rank + r;
                        if (peer >= world) peer -= world;
                        float f[8];
                        G8<Tin>::load(in.p[peer], base + g, f);
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[j] += f[j];
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] *= scale;
                    for (int r = 0; r < world; ++r) {
                        int peer = c.rank + r;
                        if (peer >= world) peer -= world;
                        G8<Tout>::store(out.p[peer], base + g, acc);
                    }
                }
            }
        }
    }
    block_barrier(c, epoch, true);  // my pushes are visible before any peer leaves its kernel
    epoch_end(c, epoch);
}

// ---- K3: NVLS (multimem) all-reduce -----------------------------------------------------
// The switch performs the reduction (multimem.ld_reduce on the multicast address returns
// the sum over all ranks) and the broadcast (multimem.st).  bf16 in/out, fp32 accumulate.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint4 multimem_ld_reduce_f32x4(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

template <bool kBf16, int UNR>
__global__ void __launch_bounds__(kThreads) allreduce_nvls_kernel(Ctx c, char* mc_in, char* mc_out, size_t n16, float scale) {
    // n16 = number of 16-byte vectors in the whole message
    uint32_t epoch = epoch_begin(c);
    const size_t slice = n16 / c.world;
    const size_t base = slice * c.rank;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    block_barrier(c, epoch, false);
    for (size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g0 < slice; g0 += stride * UNR) {
        uint4 v[UNR];
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            const size_t g = g0 + k * stride;
            if (g < slice) v[k] = kBf16 ? multimem_ld_reduce_bf16x8(mc_in + (base + g) * 16) : multimem_ld_reduce_f32x4(mc_in + (base + g) * 16);
        }
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            const size_t g = g0 + k * stride;
            if (g < slice) {
                if (kBf16) {
                    // The switch returns the fp32 sum rounded to bf16 (multimem.ld_reduce ... .acc::f32.bf16x2); the scale
                    // below is 1 / world = a power of two for 2, 4 and 8 ranks, so this second rounding is exact there
                    // (one rounding in total, like the peer-load kernels); other world sizes pay one extra bf16 rounding.
                    float f[8];
                    bf16x8_to_float(v[k], f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] *= scale;
                    multimem_st_bf16x8(mc_out + (base + g) * 16, float_to_bf16x8(f));
                } else {
                    uint4 o;
                    o.x = __float_as_uint(__uint_as_float(v[k].x) * scale);
                    o.y = __float_as_uint(__uint_as_float(v[k].y) * scale);
                    o.z = __float_as_uint(__uint_as_float(v[k].z) * scale);
                    o.w = __float_as_uint(__uint_as_float(v[k].w) * scale);
                    multimem_st_f32x4(mc_out + (base + g) * 16, o);
                }
            }
        }
    }
    block_barrier(c, epoch, true);
    epoch_end(c, epoch);
}

// ---- hybrid all-reduce: switch reduction and peer loads / stores in one launch -----------------
// The NVLS kernel alone moves S bytes per direction per GPU but keeps the links only about half busy (the multimem
// reduce / broadcast rate of the switch is the limit: 16 MB take ~39 us of data phase at 8 GPUs); the two-shot kernel
// moves 1.75 S per direction through plain loads and stores.  Running both on disjoint parts of ONE message at the same
// time fills the links: CTAs [0, nb_a) reduce granules [0, n_a) through the multicast address, CTAs [nb_a, gridDim)
// two-shot granules [n_a, n_a + n_b).  Every CTA does its own pairwise flag barriers (as in the other kernels), so the two
// parts need no synchronisation with each other.  bf16 in place; every replica gets bit-identical results (each granule is
// reduced once - by the switch or by its owner rank - and broadcast).
template <int W, int UNR_A, int UNR_B>
__global__ void __launch_bounds__(kThreads) allreduce_hybrid_kernel(Ctx c, char* mc_in, char* mc_out, PeerPtrs in, PeerPtrs out, size_t n_a,
                                                                    size_t n_b, int nb_a, float scale) {
    uint32_t epoch = epoch_begin(c);
    block_barrier(c, epoch, false);
    if ((int)blockIdx.x < nb_a) {
        const size_t slice = n_a / W, base = slice * c.rank;
        const size_t stride = (size_t)nb_a * blockDim.x;
        for (size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g0 < slice; g0 += stride * UNR_A) {
            uint4 v[UNR_A];
#pragma unroll
            for (int k = 0; k < UNR_A; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) v[k] = multimem_ld_reduce_bf16x8(mc_in + (base + g) * 16);
            }
#pragma unroll
            for (int k = 0; k < UNR_A; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
                    float f[8];
                    bf16x8_to_float(v[k], f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] *= scale;
                    multimem_st_bf16x8(mc_out + (base + g) * 16, float_to_bf16x8(f));
                }
            }
        }
    } else {
        const int vb = blockIdx.x - nb_a, vn = gridDim.x - nb_a;
        const size_t slice = n_b / W, base = n_a + slice * c.rank;
        const size_t stride = (size_t)vn * blockDim.x;
        for (size_t g0 = (size_t)vb * blockDim.x + threadIdx.x; g0 < slice; g0 += stride * UNR_B) {
            uint4 raw[UNR_B][W][1];
#pragma unroll
            for (int k = 0; k < UNR_B; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
#pragma unroll
                    for (int r = 0; r < W; ++r) {
                        int peer = c.rank + r;
                        if (peer >= W) peer -= W;
                        g8_load_raw<bf16>(in.p[peer], base + g, raw[k][r]);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < UNR_B; ++k) {
                const size_t g = g0 + k * stride;
                if (g < slice) {
                    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r = 0; r < W; ++r) g8_accumulate<bf16>(raw[k][r], acc);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] *= scale;
#pragma unroll
                    for (int r = 0; r < W; ++r) {
                        int peer = c.rank + r;
                        if (peer >= W) peer -= W;
                        G8<bf16>::store(out.p[peer], base + g, acc);
                    }
                }
            }
        }
    }
    block_barrier(c, epoch, true);
    epoch_end(c, epoch);
}

// ---- all-gather / reduce-scatter / broadcast ------------------------------------------
// all-gather: out[r * shard + i] = peer r's shard[i]   (pull; 16-byte vectors)
__global__ void __launch_bounds__(kThreads) allgather_kernel(Ctx c, PeerPtrs shard, char* out, size_t shard16) {
    uint32_t epoch = epoch_begin(c);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    block_barrier(c, epoch, false);
    for (int r = 0; r < c.world; ++r) {
        int peer = c.rank + r;
        if (peer >= c.world) peer -= c.world;
        const char* src = shard.p[peer];
        char* dst = out + (size_t)peer * shard16 * 16;
        for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < shard16; g += stride * 4) {
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (g + k * stride < shard16) v[k] = ld_weak_noalloc_v4(src + (g + k * stride) * 16);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (g + k * stride < shard16) st_global_v4(dst + (g + k * stride) * 16, v[k]);
        }
    }
    block_barrier(c, epoch, false);  // peers are done reading my shard
    epoch_end(c, epoch);
}

// reduce-scatter: out_shard[i] = scale * sum_r full_r[rank * shard + i]
template <typename Tin, typename Tout, int UNR>
__global__ void __launch_bounds__(kThreads) reducescatter_kernel(Ctx c, PeerPtrs full, char* out_shard, size_t shard8, float scale) {
    uint32_t epoch = epoch_begin(c);
    const size_t base = shard8 * c.rank;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    block_barrier(c, epoch, false);
    for (size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g0 < shard8; g0 += stride * UNR) {
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            const size_t g = g0 + k * stride;
            if (g < shard8) {
                float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int r = 0; r < c.world; ++r) {
                    int peer = c.rank + r;
                    if (peer >= c.world) peer -= c.world;
                    float f[8];
                    G8<Tin>::load(full.p[peer], base + g, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] += f[j];
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] *= scale;
                G8<Tout>::store(out_shard, g, acc);
            }
        }
    }
    block_barrier(c, epoch, false);
    epoch_end(c, epoch);
}

// K7: reduce-scatter fused with the sharded AdamW update.  The reduced gradient never
// touches memory: it is consumed in registers by the optimizer, which updates the fp32
// master / moments of this rank's shard and writes the bf16 parameter shard.
// hyper = {lr, beta1, beta2, eps, weight_decay, step} in device memory (graph friendly).
template <typename Tin, int W>
__global__ void __launch_bounds__(kThreads) reducescatter_adamw_kernel(Ctx c, PeerPtrs full, size_t shard8, float scale,
                                                                       float* __restrict__ master, float* __restrict__ m,
                                                                       float* __restrict__ v, bf16* __restrict__ param_out,
                                                                       const float* __restrict__ hyper) {
    uint32_t epoch = epoch_begin(c);
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5];
    const float bc1 = 1.f - powf(b1, step), bc2 = 1.f - powf(b2, step);
    const float inv_bc1 = 1.f / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
    const int world = W > 0 ? W : c.world;
    const size_t base = shard8 * c.rank;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    constexpr int kVec = G8<Tin>::kBytes / 16;
    block_barrier(c, epoch, false);
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < shard8; g += stride) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // issue every load of this granule first: W peer gradient vectors + 6 optimizer-state vectors
        float w[8], mm[8], vv[8];
        if constexpr (W > 0) {
            uint4 raw[W][kVec];
#pragma unroll
            for (int r = 0; r < W; ++r) {
                int peer = c.rank + r;
                if (peer >= W) peer -= W;
                g8_load_raw<Tin>(full.p[peer], base + g, raw[r]);
            }
            *reinterpret_cast<float4*>(&w[0]) = *reinterpret_cast<const float4*>(master + g * 8);
            *reinterpret_cast<float4*>(&w[4]) = *reinterpret_cast<const float4*>(master + g * 8 + 4);
            *reinterpret_cast<float4*>(&mm[0]) = *reinterpret_cast<const float4*>(m + g * 8);
            *reinterpret_cast<float4*>(&mm[4]) = *reinterpret_cast<const float4*>(m + g * 8 + 4);
            *reinterpret_cast<float4*>(&vv[0]) = *reinterpret_cast<const float4*>(v + g * 8);
            *reinterpret_cast<float4*>(&vv[4]) = *reinterpret_cast<const float4*>(v + g * 8 + 4);
#pragma unroll
            for (int r = 0; r < W; ++r) g8_accumulate<Tin>(raw[r], acc);
        } else {
            for (int r = 0; r < world; ++r) {
                int peer = c.rank + r;
                if (peer >= world) peer -= world;
                float f[8];
                G8<Tin>::load(full.p[peer], base + g, f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += f[j];
            }
            *reinterpret_cast<float4*>(&w[0]) = *reinterpret_cast<const float4*>(master + g * 8);
            *reinterpret_cast<float4*>(&w[4]) = *reinterpret_cast<const float4*>(master + g * 8 + 4);
            *reinterpret_cast<float4*>(&mm[0]) = *reinterpret_cast<const float4*>(m + g * 8);
            *reinterpret_cast<float4*>(&mm[4]) = *reinterpret_cast<const float4*>(m + g * 8 + 4);
            *reinterpret_cast<float4*>(&vv[0]) = *reinterpret_cast<const float4*>(v + g * 8);
            *reinterpret_cast<float4*>(&vv[4]) = *reinterpret_cast<const float4*>(v + g * 8 + 4);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gr = acc[j] * scale;
            mm[j] = b1 * mm[j] + (1.f - b1) * gr;
            vv[j] = b2 * vv[j] + (1.f - b2) * gr * gr;
            const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
            w[j] = w[j] * (1.f - lr * wd) - lr * (mm[j] * inv_bc1) / denom;
        }
        *reinterpret_cast<float4*>(master + g * 8) = *reinterpret_cast<float4*>(&w[0]);
        *reinterpret_cast<float4*>(master + g * 8 + 4) = *reinterpret_cast<float4*>(&w[4]);
        *reinterpret_cast<float4*>(m + g * 8) = *reinterpret_cast<float4*>(&mm[0]);
        *reinterpret_cast<float4*>(m + g * 8 + 4) = *reinterpret_cast<float4*>(&mm[4]);
        *reinterpret_cast<float4*>(v + g * 8) = *reinterpret_cast<float4*>(&vv[0]);
        *reinterpret_cast<float4*>(v + g * 8 + 4) = *reinterpret_cast<float4*>(&vv[4]);
        if (param_out) *reinterpret_cast<uint4*>(param_out + g * 8) = float_to_bf16x8(w);
    }
    block_barrier(c, epoch, false);
    epoch_end(c, epoch);
}

// broadcast (K5): everyone pulls root's buffer
__global__ void __launch_bounds__(kThreads) broadcast_kernel(Ctx c, PeerPtrs buf, int root, size_t n16) {
    uint32_t epoch = epoch_begin(c);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    block_barrier(c, epoch, false);
    if (c.rank != root) {
        const char* src = buf.p[root];
        char* dst = buf.p[c.rank];
        for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n16; g += stride)
            st_global_v4(dst + g * 16, ld_weak_noalloc_v4(src + g * 16));
    }
    block_barrier(c, epoch, false);
    epoch_end(c, epoch);
}

__global__ void barrier_kernel(Ctx c) {
    uint32_t epoch = epoch_begin(c);
    block_barrier(c, epoch, true);
    epoch_end(c, epoch);
}

// ---- host side ------------------------------------------------------------------------
int make_ctx(const ddl_comm_t* comm, Ctx* c) {
    DDL_REQUIRE(comm != nullptr, "null comm");
    DDL_REQUIRE(comm->world >= 1 && comm->world <= DDL_MAX_WORLD, "world size %d not in [1,%d]", comm->world, DDL_MAX_WORLD);
    DDL_REQUIRE(comm->rank >= 0 && comm->rank < comm->world, "bad rank %d", comm->rank);
    for (int i = 0; i < comm->world; ++i) {
        DDL_REQUIRE(comm->signal_pads[i] != nullptr, "signal pad %d is null", i);
        c->pad[i] = (uint32_t*)comm->signal_pads[i];
    }
    DDL_REQUIRE(comm->epoch != nullptr, "epoch buffer is null");
    c->epoch = (uint32_t*)comm->epoch;
    c->rank = comm->rank;
    c->world = comm->world;
    return 0;
}

int fill_ptrs(PeerPtrs* pp, void* const* ptrs, int world, const char* what) {
    for (int i = 0; i < DDL_MAX_WORLD; ++i) pp->p[i] = nullptr;
    for (int i = 0; i < world; ++i) {
        DDL_REQUIRE(ptrs && ptrs[i], "%s: peer pointer %d is null", what, i);
        DDL_REQUIRE(((uintptr_t)ptrs[i] & 15) == 0, "%s: peer pointer %d not 16-byte aligned", what, i);
        pp->p[i] = (char*)ptrs[i];
    }
    return 0;
}

int clamp_blocks(int num_blocks, size_t work_items, int per_block) {
    if (num_blocks <= 0) num_blocks = 32;
    if (num_blocks > DDL_COMM_MAX_BLOCKS) num_blocks = DDL_COMM_MAX_BLOCKS;
    long long need = (long long)((work_items + per_block - 1) / per_block);
    if (need < 1) need = 1;
    if (need < num_blocks) num_blocks = (int)need;
    return num_blocks;
}

constexpr int kOneShotUnroll = 4;

}  // namespace

DDL_API size_t ddl_comm_signal_pad_bytes(void) { return (size_t)DDL_COMM_MAX_BLOCKS * DDL_MAX_WORLD * sizeof(uint32_t); }
DDL_API size_t ddl_comm_epoch_bytes(void) { return (size_t)DDL_COMM_MAX_BLOCKS * sizeof(uint32_t); }

DDL_API size_t ddl_allreduce_oneshot_max_bytes(int dtype, int num_blocks) {
    if (num_blocks <= 0) num_blocks = 32;
    if (num_blocks > DDL_COMM_MAX_BLOCKS) num_blocks = DDL_COMM_MAX_BLOCKS;
    const size_t esz = dtype == DDL_BF16 ? 2 : 4;
    return (size_t)num_blocks * kThreads * kOneShotUnroll * 8 * esz;
}

DDL_API int ddl_allreduce(const ddl_comm_t* comm, void* const* in_ptrs, void* const* out_ptrs, size_t numel, int in_dtype,
                          int out_dtype, float scale, int algo, void* mc_in, void* mc_out, int num_blocks, void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    cudaStream_t stream = (cudaStream_t)stream_;
    DDL_REQUIRE(numel % (8 * (size_t)c.world) == 0, "ddl_allreduce: numel %zu must be a multiple of 8*world", numel);
    DDL_REQUIRE(in_dtype == DDL_F32 || in_dtype == DDL_BF16, "bad in_dtype");
    DDL_REQUIRE(out_dtype == DDL_F32 || out_dtype == DDL_BF16, "bad out_dtype");
    PeerPtrs in, out;
    if (int e = fill_ptrs(&in, in_ptrs, c.world, "in")) return e;
    if (int e = fill_ptrs(&out, out_ptrs, c.world, "out")) return e;
    const size_t n8 = numel / 8;
    const size_t in_bytes = numel * (in_dtype == DDL_BF16 ? 2 : 4);
    if (algo == DDL_ALGO_AUTO) {
        if (in_bytes <= ddl_allreduce_oneshot_max_bytes(in_dtype, num_blocks) && in_bytes <= (256u << 10))
            algo = DDL_ALGO_ONESHOT;
        else if (mc_in && mc_out && in_dtype == out_dtype)
            algo = DDL_ALGO_NVLS;
        else
            algo = DDL_ALGO_TWOSHOT;
    }
    if (algo == DDL_ALGO_ONESHOT) {
        const int nb = clamp_blocks(num_blocks, n8, kThreads);
        DDL_REQUIRE(n8 <= (size_t)nb * kThreads * kOneShotUnroll, "one-shot all-reduce: %zu elements exceed the single-wave limit; use two-shot", numel);
        char* o = out.p[c.rank];
#define ONE(TI, TO) allreduce_oneshot_kernel<TI, TO, kOneShotUnroll><<<nb, kThreads, 0, stream>>>(c, in, o, n8, scale)
        if (in_dtype == DDL_BF16 && out_dtype == DDL_BF16) ONE(bf16, bf16);
        else if (in_dtype == DDL_F32 && out_dtype == DDL_BF16) ONE(float, bf16);
        else if (in_dtype == DDL_F32 && out_dtype == DDL_F32) ONE(float, float);
        else ONE(bf16, float);
#undef ONE
    } else if (algo == DDL_ALGO_TWOSHOT) {
        DDL_REQUIRE(in_dtype == out_dtype || in_ptrs[c.rank] != out_ptrs[c.rank], "two-shot with a dtype change cannot be in place");
        const int nb = clamp_blocks(num_blocks, n8 / c.world, kThreads * 4);
#define TWO_W(TI, TO, W_, U_) allreduce_twoshot_kernel<TI, TO, W_, U_><<<nb, kThreads, 0, stream>>>(c, in, out, n8, scale)
#define TWO(TI, TO)                                   \
    do {                                              \
        constexpr int kV = sizeof(TI) == 2 ? 1 : 2;   \
        if (c.world == 2) TWO_W(TI, TO, 2, 8 / kV);   \
        else if (c.world == 4) TWO_W(TI, TO, 4, 4 / kV); \
        else if (c.world == 8) TWO_W(TI, TO, 8, 2 / kV); \
        else TWO_W(TI, TO, 0, 2);                     \
    } while (0)
        if (in_dtype == DDL_BF16 && out_dtype == DDL_BF16) TWO(bf16, bf16);
        else if (in_dtype == DDL_F32 && out_dtype == DDL_BF16) TWO(float, bf16);
        else if (in_dtype == DDL_F32 && out_dtype == DDL_F32) TWO(float, float);
        else TWO(bf16, float);
#undef TWO_W
#undef TWO
    } else if (algo == DDL_ALGO_NVLS) {
        DDL_REQUIRE(mc_in && mc_out, "NVLS all-reduce needs multicast addresses (not exposed on this host?)");
        DDL_REQUIRE(in_dtype == out_dtype, "NVLS all-reduce keeps the dtype");
        const size_t n16 = in_bytes / 16;
        DDL_REQUIRE(n16 % c.world == 0, "NVLS all-reduce: size not divisible");
        const int nb = clamp_blocks(num_blocks, n16 / c.world, kThreads * 8);
        if (in_dtype == DDL_BF16) allreduce_nvls_kernel<true, 8><<<nb, kThreads, 0, stream>>>(c, (char*)mc_in, (char*)mc_out, n16, scale);
        else allreduce_nvls_kernel<false, 8><<<nb, kThreads, 0, stream>>>(c, (char*)mc_in, (char*)mc_out, n16, scale);
    } else {
        DDL_REQUIRE(false, "unknown all-reduce algorithm %d", algo);
    }
    DDL_LAUNCH_CHECK();
    return 0;
}

// bf16, in place (in == out buffers).  nvls_permille of the message goes through the switch with nvls_blocks CTAs, the
// rest through two-shot peer loads / stores with p2p_blocks CTAs.  World size 2, 4 or 8.
DDL_API int ddl_allreduce_hybrid(const ddl_comm_t* comm, void* const* buf_ptrs, size_t numel, float scale, void* mc_buf, int nvls_permille,
                                 int nvls_blocks, int p2p_blocks, void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    DDL_REQUIRE(mc_buf != nullptr, "hybrid all-reduce needs the multicast address of the buffer");
    DDL_REQUIRE(c.world == 2 || c.world == 4 || c.world == 8, "hybrid all-reduce: world size must be 2, 4 or 8");
    DDL_REQUIRE(numel % (8 * (size_t)c.world) == 0, "hybrid all-reduce: numel %zu must be a multiple of 8*world", numel);
    DDL_REQUIRE(nvls_permille >= 0 && nvls_permille <= 1000 && nvls_blocks >= 1 && p2p_blocks >= 1, "hybrid all-reduce: bad split");
    DDL_REQUIRE(nvls_blocks + p2p_blocks <= DDL_COMM_MAX_BLOCKS, "hybrid all-reduce: too many blocks");
    PeerPtrs io;
    if (int e = fill_ptrs(&io, buf_ptrs, c.world, "buf")) return e;
    const size_t n8 = numel / 8, unit = (size_t)c.world;
    size_t n_a = (size_t)((double)n8 * nvls_permille / 1000.0) / unit * unit;
    if (n_a > n8) n_a = n8;
    const size_t n_b = n8 - n_a;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int nb = nvls_blocks + p2p_blocks;
#define HYB(W_, UB_) allreduce_hybrid_kernel<W_, 8, UB_><<<nb, kThreads, 0, stream>>>(c, (char*)mc_buf, (char*)mc_buf, io, io, n_a, n_b, nvls_blocks, scale)
    if (c.world == 2) HYB(2, 8);
    else if (c.world == 4) HYB(4, 4);
    else HYB(8, 2);
#undef HYB
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_allgather(const ddl_comm_t* comm, void* const* shard_ptrs, void* out_full, size_t shard_bytes, int num_blocks,
                          void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    DDL_REQUIRE(shard_bytes % 16 == 0, "ddl_allgather: shard bytes must be a multiple of 16");
    DDL_REQUIRE(((uintptr_t)out_full & 15) == 0, "ddl_allgather: output not 16-byte aligned");
    PeerPtrs sp;
    if (int e = fill_ptrs(&sp, shard_ptrs, c.world, "shard")) return e;
    const size_t n16 = shard_bytes / 16;
    const int nb = clamp_blocks(num_blocks, n16, kThreads * 4);
    allgather_kernel<<<nb, kThreads, 0, (cudaStream_t)stream_>>>(c, sp, (char*)out_full, n16);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_reducescatter(const ddl_comm_t* comm, void* const* full_ptrs, void* out_shard, size_t shard_numel, int in_dtype,
                              int out_dtype, float scale, int num_blocks, void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    DDL_REQUIRE(shard_numel % 8 == 0, "ddl_reducescatter: shard numel must be a multiple of 8");
    PeerPtrs fp;
    if (int e = fill_ptrs(&fp, full_ptrs, c.world, "full")) return e;
    const size_t s8 = shard_numel / 8;
    const int nb = clamp_blocks(num_blocks, s8, kThreads * 2);
    cudaStream_t stream = (cudaStream_t)stream_;
#define RS(TI, TO) reducescatter_kernel<TI, TO, 2><<<nb, kThreads, 0, stream>>>(c, fp, (char*)out_shard, s8, scale)
    if (in_dtype == DDL_BF16 && out_dtype == DDL_BF16) RS(bf16, bf16);
    else if (in_dtype == DDL_F32 && out_dtype == DDL_BF16) RS(float, bf16);
    else if (in_dtype == DDL_F32 && out_dtype == DDL_F32) RS(float, float);
    else if (in_dtype == DDL_BF16 && out_dtype == DDL_F32) RS(bf16, float);
    else DDL_REQUIRE(false, "bad dtypes");
#undef RS
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_reducescatter_adamw(const ddl_comm_t* comm, void* const* full_ptrs, size_t shard_numel, int in_dtype, float scale,
                                    float* master, float* m, float* v, void* param_out_bf16, const float* hyper, int num_blocks,
                                    void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    DDL_REQUIRE(shard_numel % 8 == 0, "ddl_reducescatter_adamw: shard numel must be a multiple of 8");
    DDL_REQUIRE(master && m && v && hyper, "ddl_reducescatter_adamw: null optimizer state");
    PeerPtrs fp;
    if (int e = fill_ptrs(&fp, full_ptrs, c.world, "full")) return e;
    const size_t s8 = shard_numel / 8;
    const int nb = clamp_blocks(num_blocks, s8, kThreads);
    cudaStream_t stream = (cudaStream_t)stream_;
#define RSA(TI, W_) reducescatter_adamw_kernel<TI, W_><<<nb, kThreads, 0, stream>>>(c, fp, s8, scale, master, m, v, (bf16*)param_out_bf16, hyper)
#define RSA_W(TI)                       \
    do {                                \
        if (c.world == 1) RSA(TI, 1);   \
        else if (c.world == 2) RSA(TI, 2); \
        else if (c.world == 4) RSA(TI, 4); \
        else if (c.world == 8) RSA(TI, 8); \
        else RSA(TI, 0);                \
    } while (0)
    if (in_dtype == DDL_BF16) RSA_W(bf16);
    else if (in_dtype == DDL_F32) RSA_W(float);
    else DDL_REQUIRE(false, "bad dtype");
#undef RSA_W
#undef RSA
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_broadcast(const ddl_comm_t* comm, void* const* buf_ptrs, int root, size_t bytes, int num_blocks, void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    DDL_REQUIRE(bytes % 16 == 0, "ddl_broadcast: bytes must be a multiple of 16");
    DDL_REQUIRE(root >= 0 && root < c.world, "ddl_broadcast: bad root");
    PeerPtrs bp;
    if (int e = fill_ptrs(&bp, buf_ptrs, c.world, "buf")) return e;
    const int nb = clamp_blocks(num_blocks, bytes / 16, kThreads * 4);
    broadcast_kernel<<<nb, kThreads, 0, (cudaStream_t)stream_>>>(c, bp, root, bytes / 16);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_barrier(const ddl_comm_t* comm, void* stream_) {
    Ctx c;
    if (int e = make_ctx(comm, &c)) return e;
    barrier_kernel<<<1, 32, 0, (cudaStream_t)stream_>>>(c);
    DDL_LAUNCH_CHECK();
    return 0;
}
