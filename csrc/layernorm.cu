// Fused (dropout +) residual-add + LayerNorm, forward and backward (K10),
// plus the fused embedding-gather + LayerNorm used at the bottom of BERT (part of K14).
//
// bf16 in / bf16 out, fp32 statistics.  One warp owns one row (hidden 768 / 1024 -> 3 / 4
// 16-byte vectors per lane, the whole row lives in registers, two-pass variance).
// Backward writes dz and accumulates dgamma / dbeta in registers across all rows a warp
// visits, then reduces per block into a [blocks, 2, H] fp32 scratch that a second kernel
// folds in a fixed order (deterministic, no atomics).
// Dropout masks come from a stateless hash of (seed, element index): nothing is stored and
// backward regenerates the same mask.  The reference has no such op.
#include "../include/b200ddl.h"
#include "common.cuh"

namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kLnThreads = kWarpsPerBlock * 32;

struct DropCfg {
    uint64_t seed;               // per-call-site salt (host value)
    const uint64_t* seed_dev;    // optional device word mixed in at run time (changes per graph replay)
    uint32_t thr;                // drop iff byte < thr   (p = thr / 256)
    float scale;                 // 1 / (1 - thr/256)
};

__device__ __forceinline__ void resolve_seed(DropCfg& d) {
    if (d.thr && d.seed_dev) d.seed ^= (*d.seed_dev + 0x632BE59BD9B4E019ull) * 0x9E3779B97F4A7C15ull;
}

__device__ __forceinline__ void apply_dropout8(float (&f)[8], const DropCfg& d, uint64_t idx8) {
    const uint32_t keep = keep_mask8(d.seed, idx8, d.thr);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = ((keep >> j) & 1u) ? f[j] * d.scale : 0.f;
}

// z = dropout(x) + residual ; y = LN(z) * gamma + beta
template <int VPL>
__global__ void __launch_bounds__(kLnThreads, (VPL <= 4 ? 2 : 1)) add_ln_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ residual,
                                                                const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                                                                bf16* __restrict__ z_out, bf16* __restrict__ y,
                                                                float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                                int rows, int H, float eps, DropCfg drop) {
    resolve_seed(drop);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nv = H >> 3;
    const float inv_h = 1.f / (float)H;
    // software pipeline across rows: the loads of the next row are in flight while this one is reduced and stored
    const long long stride = (long long)gridDim.x * kWarpsPerBlock;
    long long row = (long long)blockIdx.x * kWarpsPerBlock + warp;
    uint4 xv[VPL], rv[VPL];
    auto load_row = [&](long long r, uint4 (&a)[VPL], uint4 (&b)[VPL]) {
        if (r < rows) {
#pragma unroll
            for (int k = 0; k < VPL; ++k) {
                const int c = lane + 32 * k;
                if (c < nv) {
                    const size_t off = (size_t)r * H + c * 8;
                    a[k] = ld_stream_v4(x + off);
                    if (residual) b[k] = ld_stream_v4(residual + off);
                }
            }
        }
    };
    load_row(row, xv, rv);
    for (; row < rows; row += stride) {
        uint4 nxv[VPL], nrv[VPL];
        load_row(row + stride, nxv, nrv);
        float v[VPL][8];
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            const int c = lane + 32 * k;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[k][j] = 0.f;
            if (c < nv) {
                const size_t off = (size_t)row * H + c * 8;
                bf16x8_to_float(xv[k], v[k]);
                if (drop.thr) apply_dropout8(v[k], drop, ((size_t)row * nv + c));
                if (residual) {
                    float r[8];
                    bf16x8_to_float(rv[k], r);
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[k][j] += r[j];
                }
                if (z_out) {
                    // round once so that backward sees exactly what forward normalised
                    const uint4 q = float_to_bf16x8(v[k]);
                    st_global_v4(z_out + off, q);
                    bf16x8_to_float(q, v[k]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) s += v[k][j];
            }
        }
        const float mean = warp_sum(s) * inv_h;
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            if (lane + 32 * k < nv) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float d = v[k][j] - mean;
                    q += d * d;
                }
            }
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_h + eps);
        if (lane == 0) {
            mean_out[row] = mean;
            rstd_out[row] = rstd;
        }
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            const int c = lane + 32 * k;
            if (c < nv) {
                float g[8], b[8], o[8];
                bf16x8_to_float(ld_global_v4(gamma + c * 8), g);
                bf16x8_to_float(ld_global_v4(beta + c * 8), b);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = (v[k][j] - mean) * rstd * g[j] + b[j];
                st_global_v4(y + (size_t)row * H + c * 8, float_to_bf16x8(o));
            }
        }
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            xv[k] = nxv[k];
            rv[k] = nrv[k];
        }
    }
}

// dz = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma
// dx = dz * keep * scale (only when dropout is active; else dx aliases dz on the host side)
//
// A warp owns one row at a time, and a row is a chain load -> reduce -> store, so the loads of the NEXT row are issued
// before the current one is processed (software pipeline across the row loop; without it each row paid a full memory
// latency: 7 400 cycles per row measured).  To have registers for that, the dgamma / dbeta column sums of a warp live
// in shared memory ([warp][which][k][half][lane] float4: conflict-free 512-byte rows) instead of 2 * VPL * 8 registers.
// COLSUM: a third column sum, of the gradient that leaves through the first input (dx when dropout is active, else dz):
// that is the bias gradient of the linear layer in front of this LayerNorm, which then needs no column-sum pass of its own.
template <int VPL, bool COLSUM>
__global__ void __launch_bounds__(kLnThreads, (VPL <= 4 ? 2 : 1)) ln_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z,
                                                               const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                                                               const bf16* __restrict__ gamma, bf16* __restrict__ dz,
                                                               bf16* __restrict__ dx, float* __restrict__ partials, int rows, int H,
                                                               DropCfg drop) {
    resolve_seed(drop);
    extern __shared__ float4 ln_acc[];  // [kWarpsPerBlock][kAcc][VPL][2][32]
    constexpr int kAcc = COLSUM ? 3 : 2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nv = H >> 3;
    const float inv_h = 1.f / (float)H;
    float4* acc_g = ln_acc + (size_t)(warp * kAcc + 0) * VPL * 64;  // dgamma: [k][half][lane]
    float4* acc_b = ln_acc + (size_t)(warp * kAcc + 1) * VPL * 64;  // dbeta
    float4* acc_x = ln_acc + (size_t)(warp * kAcc + (COLSUM ? 2 : 0)) * VPL * 64;  // column sum of dx (COLSUM only)
#pragma unroll
    for (int k = 0; k < VPL; ++k)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            acc_g[(k * 2 + h) * 32 + lane] = make_float4(0.f, 0.f, 0.f, 0.f);
            acc_b[(k * 2 + h) * 32 + lane] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (COLSUM) acc_x[(k * 2 + h) * 32 + lane] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    uint4 gam[VPL];
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
        const int c = lane + 32 * k;
        gam[k] = make_uint4(0, 0, 0, 0);
        if (c < nv) gam[k] = ld_global_v4(gamma + c * 8);
    }
    const long long stride = (long long)gridDim.x * kWarpsPerBlock;
    long long row = (long long)blockIdx.x * kWarpsPerBlock + warp;
    uint4 dyv[VPL], zv[VPL];
    float mean = 0.f, rstd = 0.f;
    auto load_row = [&](long long r, uint4 (&a)[VPL], uint4 (&b)[VPL], float& mu, float& rs) {
        if (r < rows) {
            mu = mean_in[r];
            rs = rstd_in[r];
#pragma unroll
            for (int k = 0; k < VPL; ++k) {
                const int c = lane + 32 * k;
                if (c < nv) {
                    const size_t off = (size_t)r * H + c * 8;
                    a[k] = ld_stream_v4(dy + off);
                    b[k] = ld_stream_v4(z + off);
                }
            }
        }
    };
    load_row(row, dyv, zv, mean, rstd);
    for (; row < rows; row += stride) {
        uint4 ndy[VPL], nz[VPL];
        float nmean = 0.f, nrstd = 0.f;
        load_row(row + stride, ndy, nz, nmean, nrstd);  // in flight while this row is processed
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            if (lane + 32 * k < nv) {
                float d[8], zz[8], gm[8], dg[8];
                bf16x8_to_float(dyv[k], d);
                bf16x8_to_float(zv[k], zz);
                bf16x8_to_float(gam[k], gm);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (zz[j] - mean) * rstd;
                    const float g = d[j] * gm[j];
                    s1 += g;
                    s2 += g * xh;
                    dg[j] = d[j] * xh;
                }
                // only this lane touches these four float4 slots: plain read-modify-write
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float4 a = acc_g[(k * 2 + h) * 32 + lane], b = acc_b[(k * 2 + h) * 32 + lane];
                    a.x += dg[4 * h + 0]; a.y += dg[4 * h + 1]; a.z += dg[4 * h + 2]; a.w += dg[4 * h + 3];
                    b.x += d[4 * h + 0];  b.y += d[4 * h + 1];  b.z += d[4 * h + 2];  b.w += d[4 * h + 3];
                    acc_g[(k * 2 + h) * 32 + lane] = a;
                    acc_b[(k * 2 + h) * 32 + lane] = b;
                }
            }
        }
        const float c1 = warp_sum(s1) * inv_h, c2 = warp_sum(s2) * inv_h;
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            const int c = lane + 32 * k;
            if (c < nv) {
                const size_t off = (size_t)row * H + c * 8;
                float d[8], zz[8], gm[8], o[8];
                bf16x8_to_float(dyv[k], d);
                bf16x8_to_float(zv[k], zz);
                bf16x8_to_float(gam[k], gm);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = rstd * (d[j] * gm[j] - c1 - (zz[j] - mean) * rstd * c2);
                uint4 ov = float_to_bf16x8(o);
                st_global_v4(dz + off, ov);
                if (dx != nullptr && drop.thr) {
                    apply_dropout8(o, drop, ((size_t)row * nv + c));
                    ov = float_to_bf16x8(o);
                    st_global_v4(dx + off, ov);
                }
                if (COLSUM) {  // of the values as stored (bf16-rounded), like a column-sum pass over the tensor would see them
                    float r[8];
                    bf16x8_to_float(ov, r);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float4 a = acc_x[(k * 2 + h) * 32 + lane];
                        a.x += r[4 * h + 0]; a.y += r[4 * h + 1]; a.z += r[4 * h + 2]; a.w += r[4 * h + 3];
                        acc_x[(k * 2 + h) * 32 + lane] = a;
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            dyv[k] = ndy[k];
            zv[k] = nz[k];
        }
        mean = nmean;
        rstd = nrstd;
    }
    // block fold of the 8 warps' column sums -> this block's partial row [kAcc][H]
    __syncthreads();
    float* out = partials + (size_t)blockIdx.x * kAcc * H;
    const float* accf = reinterpret_cast<const float*>(ln_acc);
    for (int idx = threadIdx.x; idx < kAcc * H; idx += blockDim.x) {
        const int which = idx / H, col = idx - which * H;
        const int c = col >> 3, e8 = col & 7;       // vector index, element inside the vector
        const int k = c >> 5, ln = c & 31, h = e8 >> 2, e = e8 & 3;
        float a = 0.f;
#pragma unroll
        for (int w = 0; w < kWarpsPerBlock; ++w)
            a += accf[((((size_t)(w * kAcc + which) * VPL + k) * 2 + h) * 32 + ln) * 4 + e];
        out[idx] = a;
    }
}

// grid = ceil(2H / 32), block = 1024: 32 warps fold the per-block partial rows in parallel, four rows at a time with
// independent accumulators (the loads are L2 hits whose latency is the whole cost of this kernel); fixed order
constexpr int kLnFoldWarps = 32;
__global__ void __launch_bounds__(kLnFoldWarps * 32) ln_bwd_finalize_kernel(const float* __restrict__ partials, int nblocks, int H,
                                                                            bf16* __restrict__ dgamma, bf16* __restrict__ dbeta,
                                                                            float* __restrict__ dgamma_f32, float* __restrict__ dbeta_f32,
                                                                            int nacc, bf16* __restrict__ dxsum) {
    __shared__ float sh[kLnFoldWarps][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;
    const int tot = nacc * H;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (i < tot) {
        int b = w;
        for (; b + 3 * kLnFoldWarps < nblocks; b += 4 * kLnFoldWarps) {
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] += partials[(size_t)(b + u * kLnFoldWarps) * tot + i];
        }
        for (; b < nblocks; b += kLnFoldWarps) a[0] += partials[(size_t)b * tot + i];
    }
    sh[w][lane] = (a[0] + a[1]) + (a[2] + a[3]);
    __syncthreads();
    if (w != 0 || i >= tot) return;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < kLnFoldWarps; ++k) acc += sh[k][lane];
    if (i < H) {
        if (dgamma) dgamma[i] = __float2bfloat16(acc);
        if (dgamma_f32) dgamma_f32[i] = acc;
    } else if (i < 2 * H) {
        if (dbeta) dbeta[i - H] = __float2bfloat16(acc);
        if (dbeta_f32) dbeta_f32[i - H] = acc;
    } else if (dxsum) {
        dxsum[i - 2 * H] = __float2bfloat16(acc);
    }
}

// BERT embeddings: y = LN(word[ids] + pos[s] + type[tt]) (+ dropout on the output)
template <int VPL>
__global__ void __launch_bounds__(kLnThreads) embed_ln_fwd_kernel(const long long* __restrict__ ids, const long long* __restrict__ tts,
                                                                  const bf16* __restrict__ word, const bf16* __restrict__ pos,
                                                                  const bf16* __restrict__ type, const bf16* __restrict__ gamma,
                                                                  const bf16* __restrict__ beta, bf16* __restrict__ z_out,
                                                                  bf16* __restrict__ y, float* __restrict__ mean_out,
                                                                  float* __restrict__ rstd_out, int rows, int S, int H, float eps,
                                                                  DropCfg drop) {
    resolve_seed(drop);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nv = H >> 3;
    const float inv_h = 1.f / (float)H;
    for (long long row = (long long)blockIdx.x * kWarpsPerBlock + warp; row < rows; row += (long long)gridDim.x * kWarpsPerBlock) {
        const long long id = ids[row], tt = tts ? tts[row] : 0;
        const int s = (int)(row % S);
        float v[VPL][8];
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            const int c = lane + 32 * k;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[k][j] = 0.f;
            if (c < nv) {
                float a[8], b[8], t[8];
                bf16x8_to_float(ld_global_v4(word + (size_t)id * H + c * 8), a);
                bf16x8_to_float(ld_global_v4(pos + (size_t)s * H + c * 8), b);
                bf16x8_to_float(ld_global_v4(type + (size_t)tt * H + c * 8), t);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[k][j] = a[j] + b[j] + t[j];
                const uint4 q = float_to_bf16x8(v[k]);
                st_global_v4(z_out + (size_t)row * H + c * 8, q);
                bf16x8_to_float(q, v[k]);
#pragma unroll
                for (int j = 0; j < 8; ++j) sum += v[k][j];
            }
        }
        const float mean = warp_sum(sum) * inv_h;
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < VPL; ++k)
            if (lane + 32 * k < nv) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float d = v[k][j] - mean;
                    q += d * d;
                }
            }
        const float rstd = rsqrtf(warp_sum(q) * inv_h + eps);
        if (lane == 0) {
            mean_out[row] = mean;
            rstd_out[row] = rstd;
        }
#pragma unroll
        for (int k = 0; k < VPL; ++k) {
            const int c = lane + 32 * k;
            if (c < nv) {
                float g[8], b[8], o[8];
                bf16x8_to_float(ld_global_v4(gamma + c * 8), g);
                bf16x8_to_float(ld_global_v4(beta + c * 8), b);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = (v[k][j] - mean) * rstd * g[j] + b[j];
                if (drop.thr) apply_dropout8(o, drop, ((size_t)row * nv + c));
                st_global_v4(y + (size_t)row * H + c * 8, float_to_bf16x8(o));
            }
        }
    }
}

// elementwise dropout backward / forward on bf16 (used for the embedding output mask)
__global__ void dropout_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, size_t n8, DropCfg drop) {
    resolve_seed(drop);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        float f[8];
        bf16x8_to_float(ld_stream_v4(in + i * 8), f);
        apply_dropout8(f, drop, i);
        st_global_v4(out + i * 8, float_to_bf16x8(f));
    }
}

DropCfg make_drop(float p, unsigned long long seed, const unsigned long long* seed_dev = nullptr) {
    DropCfg d;
    d.seed = seed;
    d.seed_dev = (const uint64_t*)seed_dev;
    int thr = (int)(p * 256.f + 0.5f);
    if (thr < 0) thr = 0;
    if (thr > 255) thr = 255;
    d.thr = (uint32_t)thr;
    d.scale = 1.f / (1.f - (float)thr / 256.f);
    return d;
}

int ln_grid(int rows, int per_sm = 2) {
    int blocks = ddl_ceil_div(rows, kWarpsPerBlock);
    const int cap = DDL_SM_COUNT * per_sm;
    return blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
}

#define LN_DISPATCH(H, CALL)                                  \
    do {                                                      \
        const int _nv = (H) / 8;                              \
        if (_nv <= 32) { constexpr int VPL = 1; CALL; }       \
        else if (_nv <= 64) { constexpr int VPL = 2; CALL; }  \
        else if (_nv <= 96) { constexpr int VPL = 3; CALL; }  \
        else if (_nv <= 128) { constexpr int VPL = 4; CALL; } \
        else if (_nv <= 192) { constexpr int VPL = 6; CALL; } \
        else { constexpr int VPL = 8; CALL; }                 \
    } while (0)

}  // namespace

DDL_API float ddl_dropout_effective_p(float p) {
    const DropCfg d = make_drop(p, 0);
    return (float)d.thr / 256.f;
}

DDL_API int ddl_add_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* z_out, void* y,
                                  float* mean, float* rstd, int rows, int H, float eps, float dropout_p, unsigned long long seed,
                                  const unsigned long long* seed_dev, void* stream_) {
    DDL_REQUIRE(H % 8 == 0 && H >= 8 && H <= 2048, "layernorm: hidden size %d must be a multiple of 8 and <= 2048", H);
    DDL_REQUIRE(x && gamma && beta && y && mean && rstd, "layernorm fwd: null pointer");
    if (rows == 0) return 0;
    const DropCfg d = make_drop(dropout_p, seed, seed_dev);
    const int grid = ln_grid(rows);
    cudaStream_t s = (cudaStream_t)stream_;
    LN_DISPATCH(H, (add_ln_fwd_kernel<VPL><<<grid, kLnThreads, 0, s>>>((const bf16*)x, (const bf16*)residual, (const bf16*)gamma,
                                                                      (const bf16*)beta, (bf16*)z_out, (bf16*)y, mean, rstd, rows, H, eps, d)));
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API size_t ddl_layernorm_bwd_scratch_bytes(int rows, int H) { return (size_t)ln_grid(rows) * 2 * H * sizeof(float); }
DDL_API size_t ddl_layernorm_bwd_colsum_scratch_bytes(int rows, int H) { return (size_t)ln_grid(rows) * 3 * H * sizeof(float); }

template <bool COLSUM>
static int ln_bwd_impl(const void* dy, const void* z, const float* mean, const float* rstd, const void* gamma, void* dz, void* dx, void* dgamma,
                       void* dbeta, void* dxsum, int param_grad_dtype, void* scratch, size_t scratch_bytes, int rows, int H, float dropout_p,
                       unsigned long long seed, const unsigned long long* seed_dev, void* stream_) {
    constexpr int kAcc = COLSUM ? 3 : 2;
    DDL_REQUIRE(H % 8 == 0 && H >= 8 && H <= 2048, "layernorm: hidden size %d must be a multiple of 8 and <= 2048", H);
    DDL_REQUIRE(dy && z && mean && rstd && gamma && dz && scratch && (!COLSUM || dxsum), "layernorm bwd: null pointer");
    const size_t need = (size_t)ln_grid(rows) * kAcc * H * sizeof(float);
    DDL_REQUIRE(scratch_bytes >= need, "layernorm bwd: scratch too small (%zu < %zu)", scratch_bytes, need);
    if (rows == 0) return 0;
    const DropCfg d = make_drop(dropout_p, seed, seed_dev);
    const int grid = ln_grid(rows);
    cudaStream_t s = (cudaStream_t)stream_;
    LN_DISPATCH(H, ({
                    const int smem = kWarpsPerBlock * kAcc * VPL * 64 * (int)sizeof(float4);
                    static bool attr_set = false;  // one per (VPL, COLSUM) instantiation
                    if (!attr_set) {
                        DDL_CHECK(cudaFuncSetAttribute(ln_bwd_kernel<VPL, COLSUM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                        attr_set = true;
                    }
                    ln_bwd_kernel<VPL, COLSUM><<<grid, kLnThreads, smem, s>>>((const bf16*)dy, (const bf16*)z, mean, rstd, (const bf16*)gamma,
                                                                              (bf16*)dz, (bf16*)dx, (float*)scratch, rows, H, d);
                }));
    DDL_LAUNCH_CHECK();
    const bool f32 = param_grad_dtype == DDL_F32;
    ln_bwd_finalize_kernel<<<ddl_ceil_div(kAcc * H, 32), kLnFoldWarps * 32, 0, s>>>(
        (const float*)scratch, grid, H, f32 ? nullptr : (bf16*)dgamma, f32 ? nullptr : (bf16*)dbeta, f32 ? (float*)dgamma : nullptr,
        f32 ? (float*)dbeta : nullptr, kAcc, (bf16*)dxsum);
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const void* gamma, void* dz,
                              void* dx, void* dgamma, void* dbeta, int param_grad_dtype, void* scratch, size_t scratch_bytes, int rows,
                              int H, float dropout_p, unsigned long long seed, const unsigned long long* seed_dev, void* stream_) {
    return ln_bwd_impl<false>(dy, z, mean, rstd, gamma, dz, dx, dgamma, dbeta, nullptr, param_grad_dtype, scratch, scratch_bytes, rows, H,
                              dropout_p, seed, seed_dev, stream_);
}

// as ddl_layernorm_bwd, plus dxsum [H] bf16 = column sums of the gradient of the first input (dx if dropout is active, else dz) =
// the bias gradient of a linear layer that produced that input.  scratch: ddl_layernorm_bwd_colsum_scratch_bytes(rows, H).
DDL_API int ddl_layernorm_bwd_colsum(const void* dy, const void* z, const float* mean, const float* rstd, const void* gamma, void* dz,
                                     void* dx, void* dgamma, void* dbeta, void* dxsum, int param_grad_dtype, void* scratch,
                                     size_t scratch_bytes, int rows, int H, float dropout_p, unsigned long long seed,
                                     const unsigned long long* seed_dev, void* stream_) {
    return ln_bwd_impl<true>(dy, z, mean, rstd, gamma, dz, dx, dgamma, dbeta, dxsum, param_grad_dtype, scratch, scratch_bytes, rows, H,
                             dropout_p, seed, seed_dev, stream_);
}

DDL_API int ddl_embed_layernorm_fwd(const long long* ids, const long long* token_types, const void* word, const void* pos,
                                    const void* type, const void* gamma, const void* beta, void* z_out, void* y, float* mean,
                                    float* rstd, int rows, int S, int H, float eps, float dropout_p, unsigned long long seed,
                                    const unsigned long long* seed_dev, void* stream_) {
    DDL_REQUIRE(H % 8 == 0 && H >= 8 && H <= 2048, "embed_layernorm: hidden size %d must be a multiple of 8 and <= 2048", H);
    DDL_REQUIRE(ids && word && pos && type && gamma && beta && z_out && y && mean && rstd, "embed_layernorm: null pointer");
    if (rows == 0) return 0;
    const DropCfg d = make_drop(dropout_p, seed, seed_dev);
    const int grid = ln_grid(rows);
    cudaStream_t s = (cudaStream_t)stream_;
    LN_DISPATCH(H, (embed_ln_fwd_kernel<VPL><<<grid, kLnThreads, 0, s>>>(ids, token_types, (const bf16*)word, (const bf16*)pos, (const bf16*)type,
                                                                        (const bf16*)gamma, (const bf16*)beta, (bf16*)z_out, (bf16*)y, mean,
                                                                        rstd, rows, S, H, eps, d)));
    DDL_LAUNCH_CHECK();
    return 0;
}

DDL_API int ddl_dropout(const void* in, void* out, size_t numel, float dropout_p, unsigned long long seed,
                        const unsigned long long* seed_dev, void* stream_) {
    DDL_REQUIRE(numel % 8 == 0, "ddl_dropout: numel must be a multiple of 8");
    if (numel == 0) return 0;
    const DropCfg d = make_drop(dropout_p, seed, seed_dev);
    int blocks = ddl_ceil_div((long long)(numel / 8), 256);
    if (blocks > DDL_SM_COUNT * 8) blocks = DDL_SM_COUNT * 8;
    dropout_kernel<<<blocks, 256, 0, (cudaStream_t)stream_>>>((const bf16*)in, (bf16*)out, numel / 8, d);
    DDL_LAUNCH_CHECK();
    return 0;
}
