// Convolutions on NHWC bf16 as implicit GEMMs on tcgen05: forward, data gradient and weight gradient for the 1x1 and
// 3x3 layers (stride 1 and 2) of the reference's ResNet-50 (notebooks/cv/onnx_experiments.py:19; shapes in SURVEY.md
// section 2.E.2; K9 in section 2.E.3, BASELINE.json:5 "conv-BN-ReLU epilogues ... no cuBLAS/cuDNN on these paths").
//
// No im2col buffer exists anywhere.  The activation tensor is described to TMA as a 4-D tensor [C, W, H, N].  By default the
// operand tile of a filter tap is ONE im2col-mode load (cuTensorMapEncodeIm2col): 128 consecutive output positions x 64
// channels, displaced by the tap offset, zero-filled outside the image - that IS the padding - with the convolution stride as
// traversal stride; tiles are dense across image rows and images.  (ddl_conv_set_im2col(0) goes back to tiled boxes
// [64 channels, bw, bh, nb] with four "parity" maps for stride 2.)  3x3 / stride-1 layers on wide images take the halo-tile
// kernel of conv_halo.cuh; 256-column blocks with K <= 256 keep their weight panel in shared memory (DataParams::stay).
//
//   data kernel   out[pix, n] = sum_taps  A_tap[pix, :] . W_tap[n, :]           (forward and data gradient)
//       forward : A = x boxes,  W_tap = w[n][tap][:]  (K-major rows of the channels-last weight),  BatchNorm statistics
//                 (per-channel sum / sum of squares of the stored bf16 output) accumulated by the epilogue warps in
//                 registers and written as per-CTA partials for ddl_bn_act_fwd_train_from_partials (deterministic).
//       dgrad s1: A = dy boxes with mirrored tap offsets, W_tap = w[:][tap][n] read MN-major from the same tensor.
//       dgrad s2: the input pixels split into 4 parity classes; each class is a dense convolution over dy with the
//                 subset of taps that reaches it, stored with pixel stride 2.
//   weight kernel dw[(tap, c), k] = sum_pix  x_tap[pix, c] . dy[pix, k]         (reduction over pixels)
//       both operands MN-major straight from the NHWC tensors, up to 512 TMEM columns of accumulators per CTA alive for
//       the CTA's whole pixel range, fp32 partials per pixel split, folded in fixed order (and transposed to the weight
//       layout) by a second kernel.
//
// Structure of all kernels as in gemm_tcgen05.cu / conv1x1.cu: warp 0 = TMA producer, warp 1 = MMA issue (both run their loops
// warp-uniform, elect.sync picks the issuing lane), remaining warps = epilogue; mbarrier ring between producer and MMA,
// tcgen05.commit releases stages.
#include <cuda.h>

#include <algorithm>

#include "../include/b200ddl.h"
#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.h"

namespace {
using namespace tc;

struct Tap {
    int8_t dh, dw;  // offset of the box in (rows, columns) of the tensor map it reads
    int8_t map;     // which of the (up to 4) input tensor maps
    int8_t wtap;    // r * S + s: which filter tap's weights
};

struct alignas(64) ConvMaps {
    CUtensorMap in[4];
    CUtensorMap w;  // data kernel: weights; weight kernel: dy
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

// im2col-mode TMA load: `pixels` (fixed in the tensor map) consecutive output positions starting at the base pixel
// (w, h, n) - walking along W, then H, then N inside the map's bounding box - each displaced by the filter-tap offset
// (off_w, off_h); positions outside the image are zero-filled.  One instruction delivers a dense [pixels x 64 channels]
// operand tile that may span image rows and images.
__device__ __forceinline__ void tma_load_im2col(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c, int w, int h, int n, int off_w,
                                                int off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"((uint16_t)off_w), "h"((uint16_t)off_h)
        : "memory");
}

// mbarrier wait that adds the cycles spent waiting to *acc when timing is on
__device__ __forceinline__ void mbar_wait_timed(uint64_t* bar, uint32_t parity, bool timing, long long& acc) {
    if (!timing) {
        mbar_wait(bar, parity);
        return;
    }
    const long long t0 = clock64();
    mbar_wait(bar, parity);
    acc += clock64() - t0;
}

// x / d through a host-computed multiplier M = ceil(2^40 / d): exact while x * d < 2^40 (checked where M is computed)
__device__ __forceinline__ int fast_div(int x, unsigned long long M) { return (int)(((unsigned long long)(unsigned)x * M) >> 40); }

// ------------------------------------------------------------------------------------------------ data kernel
constexpr int kTileM = 128;
constexpr int kEpiWarps = 16;  // 4 TMEM lane quarters x 4 column groups of 64
constexpr int kThreads = 64 + kEpiWarps * 32;
constexpr int kStageA = kTileM * 128;  // [128 pixels x 64 channels] bf16
constexpr int kStgBytes = 2048;        // per epilogue warp: 32 rows x 32 columns bf16
constexpr int kMaxStages = 8;

struct DataParams {
    bf16* out;
    float* partials;  // [m_groups][2][Cn] or null (data gradient)
    int ncls;         // output parity classes (1, or up to 4 for the stride-2 data gradient)
    int cls_ntaps[4], cls_oph[4], cls_opw[4];
    Tap taps[4][9];
    int N, OHt, OWt;   // grid of output pixels of one class
    int OHf, OWf, os;  // full output extent and the pixel stride of a class inside it
    int Cn;            // output channels
    int bw, bh, nb;    // box: columns, rows, images (bw * bh * nb <= 128)
    int tiles_w, tiles_h, tiles_per_cls, num_tiles;
    int num_n_blocks, bn, groups;  // bn output channels per CTA, groups = bn / 64
    int cblocks;                   // reduction channels / 64
    int w_tap_stride;              // elements between consecutive taps inside one weight row (= conv input channels)
    int stages;
    int group;  // ring slots are released `group` at a time (one tcgen05.commit per group, see PairParams::bgroup)
    uint32_t a_bytes;
    // im2col mode: tiles are 128 consecutive positions of the class grid [N][OHt][OWt] (dense, spanning rows and images);
    // taps[].dw / dh are offsets in input pixels, the base pixel of output (a, b) is (b * cstride + low_w, a * cstride + low_h)
    int im2col, cstride, low_w, low_h;
    long long cls_pix;  // N * OHt * OWt
    // stay: all weight tiles of the CTA's column block (ntaps * cblocks * bn * 128 bytes <= 128 KB, one class) are loaded once
    // and stay in shared memory; the ring then carries activation tiles only (launch_data decides)
    int stay;
    uint32_t panel_bytes;
    int ablate;  // timing experiments only (ddl_conv_set_debug): 1 = no epilogue staging / stores, 2 = no activation loads
    int dense;       // output position of MMA row r of tile t is t * 128 + r (set by launch_data)
    unsigned long long mul_ow, mul_oh, mul_tpc;  // fast_div multipliers of OWt, OHt, tiles_per_cls (0 = use the plain division)
    long long* dbg;  // optional [gridDim.x][8] cycle counters (ddl_conv_set_debug): MMA warp total / wait accumulator / wait
                     // operands, first epilogue warp total / wait accumulator
};

template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1) conv_data_kernel(const __grid_constant__ ConvMaps maps, const DataParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int stage_b = p.bn * 128;
    const int stage_bytes = p.stay ? kStageA : kStageA + stage_b;
    uint8_t* ring = smem + (p.stay ? p.panel_bytes : 0u);  // (the resident weight panel, if any, comes first)
    uint8_t* stg_base = ring + (size_t)p.stages * stage_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(stg_base + kEpiWarps * kStgBytes);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* tmem_full = empty_bar + kMaxStages;  // [2]
    uint64_t* tmem_empty = tmem_full + 2;           // [2]
    uint64_t* w_bar = tmem_empty + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 512;
    // a CTA keeps ONE column block for its whole life, so that the BatchNorm sums can stay in registers
    const int n_blk = blockIdx.x % p.num_n_blocks, m_group = blockIdx.x / p.num_n_blocks;
    const int m_groups = gridDim.x / p.num_n_blocks;
    const int tiles_hw = p.tiles_h * p.tiles_w;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&maps.in[i]);
        tma_prefetch_desc(&maps.w);
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 4 * p.groups);
        }
        mbar_init(w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================== TMA producer =====================
        // (the whole warp runs the loop and one elected lane issues, as in the MMA warp below: inside an `if (lane == 0)`
        // region every TMA instruction is wrapped in an ELECT / R2UR.BROADCAST loop and the B_MN layers, with up to five
        // loads per k-block, ran producer-bound)
        {
            auto load_b = [&](uint8_t* sb, uint64_t* bar, int wcol, int cb) {
                if (!B_MN) {
                    tma_load_2d(sb, &maps.w, bar, wcol + cb * 64, n_blk * p.bn);
                } else {
                    for (int j = 0; j < p.groups; ++j) tma_load_2d(sb + j * 8192, &maps.w, bar, wcol + n_blk * p.bn + j * 64, cb * 64);
                }
            };
            if (p.stay) {
                // the weight panel of this column block, once: tile (tap ti, channel block cb) at index ti * cblocks + cb
                if (elect_one()) {
                    mbar_expect_tx(w_bar, p.panel_bytes);
                    for (int ti = 0; ti < p.cls_ntaps[0]; ++ti)
                        for (int cb = 0; cb < p.cblocks; ++cb)
                            load_b(smem + (size_t)(ti * p.cblocks + cb) * stage_b, w_bar, p.taps[0][ti].wtap * p.w_tap_stride, cb);
                }
                __syncwarp();
            }
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx_bytes = p.a_bytes + (p.stay ? 0u : (uint32_t)stage_b);
            for (int tile = m_group; tile < p.num_tiles; tile += m_groups) {
                // (divisions through the host-computed multipliers where they exist; none at all for flat dense tiles)
                const int cls = p.ncls == 1 ? 0 : (p.mul_tpc ? fast_div(tile, p.mul_tpc) : tile / p.tiles_per_cls);
                const int t = tile - cls * p.tiles_per_cls;
                int n0 = 0, oh0 = 0, ow0 = t * kTileM;  // flat dense tiles: one box row of 128 pixels per tile
                int iw = 0, ih = 0, in_ = 0;             // im2col mode: base pixel of the tile's first position
                if (p.im2col) {
                    const int m0 = t * kTileM;  // < 2^31 (checked on the host)
                    const int t2 = p.mul_ow ? fast_div(m0, p.mul_ow) : m0 / p.OWt, b0 = m0 - t2 * p.OWt;
                    in_ = p.mul_oh ? fast_div(t2, p.mul_oh) : t2 / p.OHt;
                    iw = b0 * p.cstride + p.low_w;
                    ih = (t2 - in_ * p.OHt) * p.cstride + p.low_h;
                } else if (!p.dense) {
                    const int ng = t / tiles_hw, rem = t - ng * tiles_hw;
                    const int hb = rem / p.tiles_w, wb = rem - hb * p.tiles_w;
                    n0 = ng * p.nb; oh0 = hb * p.bh; ow0 = wb * p.bw;
                }
                const int ntaps = p.cls_ntaps[cls];
                for (int ti = 0; ti < ntaps; ++ti) {
                    const Tap tp = p.taps[cls][ti];
                    const int wcol = tp.wtap * p.w_tap_stride;
                    for (int cb = 0; cb < p.cblocks; ++cb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = ring + (size_t)stage * stage_bytes;
                        if (elect_one()) {
                            if (p.ablate & 2) {
                                mbar_arrive(&full_bar[stage]);
                            } else {
                                mbar_expect_tx(&full_bar[stage], tx_bytes);
                                if (p.im2col) tma_load_im2col(sa, &maps.in[0], &full_bar[stage], cb * 64, iw, ih, in_, tp.dw - p.low_w, tp.dh - p.low_h);
                                else tma_load_4d(sa, &maps.in[tp.map], &full_bar[stage], cb * 64, ow0 + tp.dw, oh0 + tp.dh, n0);
                                if (!p.stay) load_b(sa + kStageA, &full_bar[stage], wcol, cb);
                            }
                        }
                        __syncwarp();
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // The whole warp runs the loop so that stage addresses and descriptors stay warp-uniform (uniform registers feed
        // UTCHMMA directly); one elected lane issues.  The descriptors of a stage differ from a per-kernel constant only in
        // the 14-bit address field, and a k-step only adds to it: +32 B (K-major) or +2048 B (MN-major B).
        {
            const uint32_t idesc = make_idesc(kTileM, p.bn, false, B_MN);
            const uint64_t da_const = make_smem_desc(0, 16, 1024);
            const uint64_t db_const = B_MN ? make_smem_desc(0, 8192, 1024) : make_smem_desc(0, 16, 1024);
            constexpr uint32_t kStepA = 32 >> 4, kStepB = (B_MN ? 2048 : 32) >> 4;
            const uint32_t ring_base = smem_u32(ring), panel_base = smem_u32(smem);
            const bool stay = p.stay != 0;
            if (stay) {
                mbar_wait(w_bar, 0);
                tc_fence_after();
            }
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            const bool timing = p.dbg != nullptr;
            long long w_te = 0, w_full = 0;
            const long long t_start = timing ? clock64() : 0;
            for (int tile = m_group; tile < p.num_tiles; tile += m_groups) {
                const int cls = tile / p.tiles_per_cls;
                const int nkb = p.cls_ntaps[cls] * p.cblocks;
                mbar_wait_timed(&tmem_empty[acc], acc_phase ^ 1, timing, w_te);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * p.bn;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait_timed(&full_bar[stage], phase, timing, w_full);
                    tc_fence_after();
                    const uint32_t sa = ring_base + (uint32_t)stage * (uint32_t)stage_bytes;
                    const uint32_t sb = stay ? panel_base + (uint32_t)kb * (uint32_t)stage_b : sa + kStageA;
                    const uint64_t da = da_const | (uint64_t)(sa >> 4), db = db_const | (uint64_t)(sb >> 4);
                    if (elect_one()) {
                        umma_bf16(d_tmem, da, db, idesc, kb > 0 ? 1u : 0u);
#pragma unroll
                        for (int k = 1; k < 4; ++k) umma_bf16(d_tmem, da + k * kStepA, db + k * kStepB, idesc, 1u);
                        umma_commit(&empty_bar[stage]);
                    }
                    __syncwarp();
                    if (++stage == p.stages) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit(&tmem_full[acc]);
                __syncwarp();
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            if (timing && lane == 0) {
                long long* d = p.dbg + (size_t)blockIdx.x * 8;
                d[0] = clock64() - t_start;
                d[1] = w_te;
                d[2] = w_full;
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;
        const int quarter = warp & 3;  // TMEM lanes 32 * quarter .. + 31
        const int group = ew >> 2;     // 64-column chunk of the accumulator
        if (group < p.groups) {
            uint8_t* stg = stg_base + ew * kStgBytes;
            const bool stats = p.partials != nullptr;
            const bool timing = p.dbg != nullptr && ew == 0;
            long long w_tf = 0;
            const long long t_start = timing ? clock64() : 0;
            float cs[16], cq[16];  // columns col0 + h * 32 + t_ch * 8 + q  at index h * 8 + q
#pragma unroll
            for (int i = 0; i < 16; ++i) cs[i] = cq[i] = 0.f;
            const int col0 = n_blk * p.bn + group * 64;
            const int t_row = lane >> 2, t_ch = lane & 3;
            const int rows_img = p.bw * p.bh, rows_box = rows_img * p.nb;
            int acc = 0;
            uint32_t acc_phase = 0;
            // dense: MMA row r of tile t IS output position t * 128 + r (1x1 / stride-1 layers and every stride-1 class in im2col
            // mode).  No division per tile then: a chain of ~13 integer divisions in front of every accumulator (the tiled-box
            // decode) cost ~2000 cycles per tile and made the epilogue, not the mainloop, the bottleneck of the short-K layers.
            const bool dense = p.dense != 0;
            for (int tile = m_group; tile < p.num_tiles; tile += m_groups) {
                int cls = 0, t = tile, n0 = 0, oh0 = 0, ow0 = 0, oph = 0, opw = 0;
                if (!dense) {
                    cls = p.mul_tpc ? fast_div(tile, p.mul_tpc) : tile / p.tiles_per_cls;
                    t = tile - cls * p.tiles_per_cls;
                    const int ng = t / tiles_hw, rem = t - ng * tiles_hw;
                    const int hb = rem / p.tiles_w, wb = rem - hb * p.tiles_w;
                    n0 = ng * p.nb; oh0 = hb * p.bh; ow0 = wb * p.bw;
                    oph = p.cls_oph[cls]; opw = p.cls_opw[cls];
                }
                // element offset of the 16-byte piece this lane stores for each of its 4 rows per 32-column half
                long long off[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = quarter * 32 + it * 8 + t_row;
                    off[it] = -1;
                    if (dense) {
                        const int m = t * kTileM + r;
                        if (m < (int)p.cls_pix) off[it] = (long long)m * p.Cn + col0 + t_ch * 8;
                    } else if (p.im2col) {
                        const int m = t * kTileM + r;
                        if (m < (int)p.cls_pix) {
                            if (p.os == 1) {  // dense output: position m IS the pixel index
                                off[it] = (long long)m * p.Cn + col0 + t_ch * 8;
                            } else {
                                const int t2 = p.mul_ow ? fast_div(m, p.mul_ow) : m / p.OWt, b = m - t2 * p.OWt;
                                const int nn = p.mul_oh ? fast_div(t2, p.mul_oh) : t2 / p.OHt, a = t2 - nn * p.OHt;
                                off[it] = (((long long)nn * p.OHf + a * p.os + oph) * p.OWf + b * p.os + opw) * p.Cn + col0 + t_ch * 8;
                            }
                        }
                    } else if (r < rows_box) {
                        const int j = r / rows_img, rr = r - j * rows_img;
                        const int a = rr / p.bw, b = rr - a * p.bw;
                        if (n0 + j < p.N && oh0 + a < p.OHt && ow0 + b < p.OWt)
                            off[it] = (((long long)(n0 + j) * p.OHf + (oh0 + a) * p.os + oph) * p.OWf + (ow0 + b) * p.os + opw) * p.Cn + col0 + t_ch * 8;
                    }
                }
                mbar_wait_timed(&tmem_full[acc], acc_phase, timing, w_tf);
                tc_fence_after();
                const uint32_t taddr = tmem_base + acc * p.bn + group * 64 + ((uint32_t)(quarter * 32) << 16);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + h * 32, r);
                    tmem_ld_wait();
                    if (h == 1) {  // accumulator chunk fully read: hand it back before the stores
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
                    }
                    if (p.ablate & 1) continue;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float f[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
                        *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = float_to_bf16x8(f);
                    }
                    __syncwarp();
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int rr = it * 8 + t_row;
                        const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((t_ch ^ ((rr >> 1) & 3)) << 4));
                        if (off[it] >= 0) {
                            st_global_v4(p.out + off[it] + h * 32, v);
                            if (stats) {
                                float f[8];
                                bf16x8_to_float(v, f);
#pragma unroll
                                for (int q = 0; q < 8; ++q) {
                                    cs[h * 8 + q] += f[q];
                                    cq[h * 8 + q] = fmaf(f[q], f[q], cq[h * 8 + q]);
                                }
                            }
                        }
                    }
                    __syncwarp();
                }
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            if (timing && lane == 0) {
                long long* d = p.dbg + (size_t)blockIdx.x * 8;
                d[3] = clock64() - t_start;
                d[4] = w_tf;
            }
            if (stats) {
                // fold the 8 row groups of the warp (lanes with equal t_ch), then the 4 lane-quarter warps of this
                // column group in fixed order -> this CTA's partial statistics
#pragma unroll
                for (int i = 0; i < 16; ++i) {
#pragma unroll
                    for (int o = 4; o < 32; o <<= 1) {
                        cs[i] += __shfl_xor_sync(0xffffffffu, cs[i], o);
                        cq[i] += __shfl_xor_sync(0xffffffffu, cq[i], o);
                    }
                }
                float* mine = reinterpret_cast<float*>(stg);  // own staging area, free now: [64 columns][2]
                if (t_row == 0) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int c = (i >> 3) * 32 + t_ch * 8 + (i & 7);
                        mine[c * 2] = cs[i];
                        mine[c * 2 + 1] = cq[i];
                    }
                }
                asm volatile("bar.sync %0, 128;" ::"r"(1 + group) : "memory");  // the four warps of this group
                if (quarter == 2) {  // ew = 4 * group
                    float t4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int qw = 0; qw < 4; ++qw) {
                        const float* src = reinterpret_cast<const float*>(stg_base + (group * 4 + qw) * kStgBytes);
#pragma unroll
                        for (int e = 0; e < 4; ++e) t4[e] += src[lane * 4 + e];
                    }
                    float* outp = p.partials + (size_t)m_group * 2 * p.Cn;
                    const int col = col0 + 2 * lane;
                    outp[col] = t4[0];
                    outp[p.Cn + col] = t4[1];
                    outp[col + 1] = t4[2];
                    outp[p.Cn + col + 1] = t4[3];
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------------------------------------ pair data kernel
// Same contraction as conv_data_kernel, for layers with >= 128 output channels, restructured around what the profile of
// that kernel showed (profiles/ncu_prof_conv_v1.md: tensor pipe busy but fed at ~60 % - shared memory is shared between
// the MMA operand reads and the TMA writes of operands that are re-loaded far too often):
//   * a CTA PAIR (cluster of 2, tcgen05.mma.cta_group::2) owns two pixel tiles x bn channels: M = 256, each CTA stages
//     its own pixel tile and only HALF of the weight tile per k-block (half the weight traffic and half the B reads per SM);
//   * 3x3 / stride-1 layers (forward and data gradient): the activation tile is loaded ONCE per 64-channel block as a
//     halo box [64, W + 1, th + 2, nb] and the nine taps are the same shared-memory tile read through descriptors shifted
//     by (dh + 1) * (W + 1) + (dw + 1) rows (the UMMA swizzle is a function of the absolute address: csrc/probe_umma.cu).
//     The box starts at column -1, so with a row pitch of W + 1 the zero column of row h + 1 doubles as the right-hand
//     padding of row h.  MMA row m = a * (W + 1) + b; the rows with b = W are computed and dropped.
//   * A and B move through separate rings (A slots are reused by nine k-blocks in halo mode).
// The tensor maps of this kernel cover activations and weights of the launching GPU only (its TMA loads are the
// cta_group::2 form; ddl_conv_fwd / ddl_conv_dgrad check the pointers and otherwise use conv_data_kernel).
constexpr int kPairMaxStages = 8;

struct PairParams {
    bf16* out;
    float* partials;  // [2 * pair groups][2][Cn] or null
    int ncls;
    int cls_ntaps[4], cls_oph[4], cls_opw[4];
    Tap taps[4][9];
    int N, OHt, OWt, OHf, OWf, os, Cn;
    int bw, bh, nb;      // the TMA box (halo mode: including the halo)
    int tiles_w, tiles_h, tiles_per_cls, pair_tiles_per_cls, num_pair_tiles;
    int th;              // output rows per tile (halo mode: bh - 2, direct: bh)
    int pitch, rows_img; // row-index geometry of the accumulator: m = j * rows_img + a * pitch + b
    int num_n_blocks, bn, groups;
    int cblocks, w_tap_stride;
    int halo;
    int a_slot_bytes, a_stages, b_stages;
    int bgroup;          // B (and, in direct mode, A) slots are handed back to the producer `bgroup` at a time: a
                         // tcgen05.commit costs ~450 cycles of tensor-pipe time (csrc/probe_umma.cu), four MMAs of a
                         // 64-deep k-block take 256 - 512, so one commit per k-block would halve the MMA rate
    uint32_t a_bytes;
    int im2col, cstride, low_w, low_h;  // im2col mode (direct mode only), as in DataParams
    long long cls_pix;
    long long* dbg;  // optional [gridDim.x][8] cycle counters (ddl_conv_set_debug): where each role of a CTA waits
    int ablate;      // development only: bit 0 = epilogue drops its staging / stores, bit 1 = producer issues no TMA loads
};

__device__ __forceinline__ void tma_load_im2col_2sm(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c, int w, int h, int n, int off_w,
                                                    int off_h) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c), "r"(w), "r"(h), "r"(n), "h"((uint16_t)off_w), "h"((uint16_t)off_h)
        : "memory");
}

__device__ __forceinline__ void tma_load_4d_2sm(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1) conv_pair_kernel(const __grid_constant__ ConvMaps maps, const PairParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int b_half = p.bn >> 1;
    const int b_slot = b_half * 128;
    uint8_t* a_ring = smem;
    uint8_t* b_ring = a_ring + (size_t)p.a_stages * p.a_slot_bytes;
    uint8_t* stg_base = b_ring + (size_t)p.b_stages * b_slot;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(stg_base + kEpiWarps * kStgBytes);
    uint64_t* a_empty = a_full + kPairMaxStages;
    uint64_t* b_full = a_empty + kPairMaxStages;
    uint64_t* b_empty = b_full + kPairMaxStages;
    uint64_t* tmem_full = b_empty + kPairMaxStages;  // [2]
    uint64_t* tmem_empty = tmem_full + 2;             // [2]
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta_rank = (int)cluster_ctarank();
    const bool leader = cta_rank == 0;
    constexpr uint32_t kTmemCols = 512;
    const int cluster_id = blockIdx.x >> 1;
    const int n_blk = cluster_id % p.num_n_blocks, pgroup = cluster_id / p.num_n_blocks;
    const int pgroups = (gridDim.x >> 1) / p.num_n_blocks;
    const int tiles_hw = p.tiles_h * p.tiles_w;

    if (p.halo) {
        // the row right behind each halo box is read by the last valid pixel's (+1, +1) tap: it must be zero
        const int box_rows = p.bw * p.bh * p.nb;
        for (int s = warp; s < p.a_stages; s += kThreads / 32)
            if (lane < 8) reinterpret_cast<uint4*>(a_ring + (size_t)s * p.a_slot_bytes + (size_t)box_rows * 128)[lane] = make_uint4(0, 0, 0, 0);
        fence_proxy_async();
    }
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&maps.in[i]);
        tma_prefetch_desc(&maps.w);
        for (int s = 0; s < kPairMaxStages; ++s) {
            mbar_init(&a_full[s], 1);
            mbar_init(&a_empty[s], 1);
            mbar_init(&b_full[s], 1);
            mbar_init(&b_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 2 * 4 * p.groups);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_2sm(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    // pair tile q -> class and this CTA's tile inside the class; a tile index past the end decodes to an image index
    // >= N: its loads are zero-filled and nothing is stored
    auto decode = [&](int q, int& cls, int& n0, int& oh0, int& ow0) {
        cls = q / p.pair_tiles_per_cls;
        const int t = (q - cls * p.pair_tiles_per_cls) * 2 + cta_rank;
        if (p.im2col) {  // tile t = positions [128 t, 128 t + 128) of the class grid; n0 / oh0 / ow0 = its first position
            const int m0 = t * kTileM, t2 = m0 / p.OWt;
            ow0 = m0 % p.OWt;
            oh0 = t2 % p.OHt;
            n0 = t2 / p.OHt;
            return;
        }
        const int ng = t / tiles_hw, rem = t - ng * tiles_hw;
        const int hb = rem / p.tiles_w, wb = rem - hb * p.tiles_w;
        n0 = ng * p.nb;
        oh0 = hb * p.th;
        ow0 = wb * (p.bw - p.halo);
    };

    if (warp == 0) {
        // ===================== TMA producer (both CTAs; all bytes complete on the leader's barriers) =====================
        if (lane == 0) {
            int as = 0, bs = 0;
            uint32_t aph = 0, bph = 0;
            const bool timing = p.dbg != nullptr;
            long long w_ae = 0, w_be = 0;
            const long long t_start = timing ? clock64() : 0;
            for (int q = pgroup; q < p.num_pair_tiles; q += pgroups) {
                int cls, n0, oh0, ow0;
                decode(q, cls, n0, oh0, ow0);
                const int ntaps = p.cls_ntaps[cls];
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    for (int ti = 0; ti < ntaps; ++ti) {
                        const Tap tp = p.taps[cls][ti];
                        if (!p.halo || ti == 0) {
                            // direct mode: A slot i and B slot i are released together (b_empty), no separate wait
                            if (p.halo) mbar_wait_timed(&a_empty[as], aph ^ 1, timing, w_ae);
                            else if (bs % p.bgroup == 0) mbar_wait_timed(&b_empty[bs / p.bgroup], bph ^ 1, timing, w_be);
                            uint8_t* sa = a_ring + (size_t)as * p.a_slot_bytes;
                            if (p.ablate & 2) {
                                if (leader) mbar_arrive(&a_full[as]);
                            } else {
                                if (leader) mbar_expect_tx(&a_full[as], 2 * p.a_bytes);
                                if (p.halo) tma_load_4d_2sm(sa, &maps.in[0], &a_full[as], cb * 64, ow0 - 1, oh0 - 1, n0);
                                else if (p.im2col)
                                    tma_load_im2col_2sm(sa, &maps.in[0], &a_full[as], cb * 64, ow0 * p.cstride + p.low_w, oh0 * p.cstride + p.low_h, n0,
                                                        tp.dw - p.low_w, tp.dh - p.low_h);
                                else tma_load_4d_2sm(sa, &maps.in[tp.map], &a_full[as], cb * 64, ow0 + tp.dw, oh0 + tp.dh, n0);
                            }
                            if (++as == p.a_stages) { as = 0; aph ^= 1; }
                        }
                        if (p.halo && bs % p.bgroup == 0) mbar_wait_timed(&b_empty[bs / p.bgroup], bph ^ 1, timing, w_be);
                        uint8_t* sb = b_ring + (size_t)bs * b_slot;
                        const int nb0 = n_blk * p.bn + cta_rank * b_half;
                        if (p.ablate & 2) {
                            if (leader) mbar_arrive(&b_full[bs]);
                        } else if (leader) {
                            mbar_expect_tx(&b_full[bs], 2 * (uint32_t)b_slot);
                        }
                        if (p.ablate & 2) {
                        } else if (!B_MN) {
                            tma_load_2d_2sm(sb, &maps.w, &b_full[bs], tp.wtap * p.w_tap_stride + cb * 64, nb0);
                        } else {
                            for (int j = 0; j < (b_half >> 6); ++j)
                                tma_load_2d_2sm(sb + j * 8192, &maps.w, &b_full[bs], tp.wtap * p.w_tap_stride + nb0 + j * 64, cb * 64);
                        }
                        if (++bs == p.b_stages) { bs = 0; bph ^= 1; }
                    }
                }
            }
            if (timing) {
                long long* d = p.dbg + (size_t)blockIdx.x * 8;
                d[0] = clock64() - t_start;  // producer: total, waiting for a free A slot, waiting for a free B slot
                d[1] = w_ae;
                d[2] = w_be;
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader only) =====================
        if (lane == 0 && leader) {
            const uint32_t idesc = make_idesc(256, p.bn, false, B_MN);
            int as = 0, bs = 0, acc = 0;
            uint32_t aph = 0, bph = 0, acc_phase = 0;
            const bool timing = p.dbg != nullptr;
            long long w_af = 0, w_bf = 0, w_te = 0, w_issue = 0, w_commit = 0;
            const long long t_start = timing ? clock64() : 0;
            for (int q = pgroup; q < p.num_pair_tiles; q += pgroups) {
                const int cls = q / p.pair_tiles_per_cls;
                const int ntaps = p.cls_ntaps[cls];
                mbar_wait_timed(&tmem_empty[acc], acc_phase ^ 1, timing, w_te);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * p.bn;
                bool first = true;
                for (int cb = 0; cb < p.cblocks; ++cb) {
                    for (int ti = 0; ti < ntaps; ++ti) {
                        if (!p.halo || ti == 0) {
                            mbar_wait_timed(&a_full[as], aph, timing, w_af);
                            tc_fence_after();
                        }
                        mbar_wait_timed(&b_full[bs], bph, timing, w_bf);
                        tc_fence_after();
                        uint32_t sa = smem_u32(a_ring + (size_t)as * p.a_slot_bytes);
                        if (p.halo) {
                            const Tap tp = p.taps[cls][ti];
                            sa += (uint32_t)(((int)tp.dh + 1) * p.pitch + (int)tp.dw + 1) * 128u;
                        }
                        const uint32_t sb = smem_u32(b_ring + (size_t)bs * b_slot);
                        const long long ti0 = timing ? clock64() : 0;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const uint64_t da = make_smem_desc(sa + k * 32, 16, 1024);
                            const uint64_t db = B_MN ? make_smem_desc(sb + k * 2048, 8192, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
                            umma_bf16_2sm(d_tmem, da, db, idesc, (!first || k > 0) ? 1u : 0u);
                        }
                        const long long ti1 = timing ? clock64() : 0;
                        first = false;
                        if (bs % p.bgroup == p.bgroup - 1) umma_commit_2sm(&b_empty[bs / p.bgroup]);
                        if (timing) {
                            w_issue += ti1 - ti0;
                            w_commit += clock64() - ti1;
                        }
                        if (++bs == p.b_stages) { bs = 0; bph ^= 1; }
                        if (!p.halo || ti == ntaps - 1) {
                            if (p.halo) umma_commit_2sm(&a_empty[as]);
                            if (++as == p.a_stages) { as = 0; aph ^= 1; }
                        }
                    }
                }
                umma_commit_2sm(&tmem_full[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            if (timing) {
                long long* d = p.dbg + (size_t)blockIdx.x * 8;
                d[3] = clock64() - t_start;  // MMA issuer: total, waiting for A data, for B data, for a free accumulator
                d[4] = w_af;
                d[5] = w_bf;
                d[6] = w_te;
                d[7] = w_issue;
                d[0] = w_commit;  // (overwrites the leader's producer total)
            }
        }
    } else {
        // ===================== epilogue warps (each CTA drains its own 128 rows) =====================
        const int ew = warp - 2;
        const int quarter = warp & 3;
        const int group = ew >> 2;
        if (group < p.groups) {
            uint8_t* stg = stg_base + ew * kStgBytes;
            const bool stats = p.partials != nullptr;
            float cs[16], cq[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) cs[i] = cq[i] = 0.f;
            const int col0 = n_blk * p.bn + group * 64;
            const int t_row = lane >> 2, t_ch = lane & 3;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int q = pgroup; q < p.num_pair_tiles; q += pgroups) {
                int cls, n0, oh0, ow0;
                decode(q, cls, n0, oh0, ow0);
                const int oph = p.cls_oph[cls], opw = p.cls_opw[cls];
                long long off[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = quarter * 32 + it * 8 + t_row;
                    off[it] = -1;
                    if (p.im2col) {
                        const int m = (n0 * p.OHt + oh0) * p.OWt + ow0 + r;
                        if (m < (int)p.cls_pix) {
                            if (p.os == 1) {
                                off[it] = (long long)m * p.Cn + col0 + t_ch * 8;
                            } else {
                                const int b = m % p.OWt, t2 = m / p.OWt;
                                const int a = t2 % p.OHt, nn = t2 / p.OHt;
                                off[it] = (((long long)nn * p.OHf + a * p.os + oph) * p.OWf + b * p.os + opw) * p.Cn + col0 + t_ch * 8;
                            }
                        }
                        continue;
                    }
                    const int j = r / p.rows_img, rr = r - j * p.rows_img;
                    const int a = rr / p.pitch, b = rr - a * p.pitch;
                    if (j < p.nb && a < p.th && b < p.bw - p.halo && n0 + j < p.N && oh0 + a < p.OHt && ow0 + b < p.OWt)
                        off[it] = (((long long)(n0 + j) * p.OHf + (oh0 + a) * p.os + oph) * p.OWf + (ow0 + b) * p.os + opw) * p.Cn + col0 + t_ch * 8;
                }
                mbar_wait(&tmem_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + acc * p.bn + group * 64 + ((uint32_t)(quarter * 32) << 16);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + h * 32, r);
                    tmem_ld_wait();
                    if (h == 1) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_leader(&tmem_empty[acc]);
                    }
                    if (p.ablate & 1) continue;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float f[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
                        *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = float_to_bf16x8(f);
                    }
                    __syncwarp();
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int rr = it * 8 + t_row;
                        const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((t_ch ^ ((rr >> 1) & 3)) << 4));
                        if (off[it] >= 0) {
                            st_global_v4(p.out + off[it] + h * 32, v);
                            if (stats) {
                                float f[8];
                                bf16x8_to_float(v, f);
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    cs[h * 8 + e] += f[e];
                                    cq[h * 8 + e] = fmaf(f[e], f[e], cq[h * 8 + e]);
                                }
                            }
                        }
                    }
                    __syncwarp();
                }
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            if (stats) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
#pragma unroll
                    for (int o = 4; o < 32; o <<= 1) {
                        cs[i] += __shfl_xor_sync(0xffffffffu, cs[i], o);
                        cq[i] += __shfl_xor_sync(0xffffffffu, cq[i], o);
                    }
                }
                float* mine = reinterpret_cast<float*>(stg);
                if (t_row == 0) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int c = (i >> 3) * 32 + t_ch * 8 + (i & 7);
                        mine[c * 2] = cs[i];
                        mine[c * 2 + 1] = cq[i];
                    }
                }
                asm volatile("bar.sync %0, 128;" ::"r"(1 + group) : "memory");
                if (quarter == 2) {
                    float t4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int qw = 0; qw < 4; ++qw) {
                        const float* src = reinterpret_cast<const float*>(stg_base + (group * 4 + qw) * kStgBytes);
#pragma unroll
                        for (int e = 0; e < 4; ++e) t4[e] += src[lane * 4 + e];
                    }
                    float* outp = p.partials + (size_t)(pgroup * 2 + cta_rank) * 2 * p.Cn;
                    const int col = col0 + 2 * lane;
                    outp[col] = t4[0];
                    outp[p.Cn + col] = t4[1];
                    outp[col + 1] = t4[2];
                    outp[p.Cn + col + 1] = t4[3];
                }
            }
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

#include "conv_halo.cuh"

// ------------------------------------------------------------------------------------------------ weight gradient
// D[m = (tap, c)][n = k] = sum over pixels of x_tap[pix][c] * dy[pix][k].  Both operands are MN-major: a pixel's
// 64-channel slice is one 128-byte row, the reduction index (pixel) walks the rows.  The (tap, c) axis is cut into
// 64-channel "units"; an MMA takes two consecutive units as its M = 128 (two 64-wide blocks one slot apart, LBO =
// slot size) and bn output channels as N.  A CTA owns `G` such unit pairs x one column block (G * bn <= 512 TMEM
// columns) and one slice of the pixel range; it writes its fp32 accumulators as a partial that the fold kernel sums in
// fixed order.  Slots hold ceil16(pixels) rows; the rows past the box are zeroed once and never written again, so the
// K padding multiplies zeros.
constexpr int kWgThreads = 192;  // TMA, MMA, 4 epilogue warps
constexpr int kWgMaxUnits = 8;    // 64-channel units (two per accumulator pair) a CTA may own

struct WgradParams {
    float* partials;  // [ksplit][Mtot][Cout]
    int ntaps;
    Tap taps[9];
    int N, OHt, OWt, bw, bh, nb, tiles_w, tiles_h, num_chunks;
    int cblocks, U, G, mgroups;
    int Cout, bn, nblk;
    int ksplit, cps;
    int slot_bytes, ksteps, stages, pix;
    long long Mtot;
    // im2col mode: chunk = `pix` consecutive output positions (dense, spanning rows and images); x taps through the im2col
    // map maps.in[0], dy through the flattened map maps.w; nothing needs zero-filling (TMA fills what lies past the end)
    int im2col, cstride, low_w, low_h;
    unsigned long long mul_ow, mul_oh;  // fast_div multipliers of OWt, OHt (0 = plain division)
};

__global__ void __launch_bounds__(kWgThreads, 1) conv_wgrad_kernel(const __grid_constant__ ConvMaps maps, const WgradParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int dy_slots = p.bn >> 6;
    const int nslots = dy_slots + 2 * p.G;
    const int stage_bytes = nslots * p.slot_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* done_bar = empty_bar + kMaxStages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(done_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 512;

    // zero every slot once when a chunk is not a whole number of 16-pixel k-steps: the rows behind each box (K padding)
    // must read as zeros
    if (p.pix * 128 != p.slot_bytes) {
        const int n16 = p.stages * stage_bytes / 16;
        uint4* s4 = reinterpret_cast<uint4*>(smem);
        for (int i = threadIdx.x; i < n16; i += kWgThreads) s4[i] = make_uint4(0, 0, 0, 0);
        fence_proxy_async();
    }
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&maps.in[i]);
        tma_prefetch_desc(&maps.w);
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(done_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int nb_ = blockIdx.x % p.nblk, t_ = blockIdx.x / p.nblk;
    const int mg = t_ % p.mgroups, ks = t_ / p.mgroups;
    const int u0 = mg * 2 * p.G;
    const int nu = min(2 * p.G, p.U - u0);  // 64-channel units of this CTA
    const int npairs = (nu + 1) >> 1;
    const int c0 = ks * p.cps, c1 = min(c0 + p.cps, p.num_chunks);
    const int tiles_hw = p.tiles_h * p.tiles_w;

    if (warp == 0) {
        // producer: warp-uniform loop, one elected lane issues the (up to 4 + 8) loads of a chunk; the unit -> (tap, channel
        // block) decode is hoisted out of the chunk loop
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t tx_bytes = (uint32_t)(dy_slots + nu) * (uint32_t)p.pix * 128u;
        int u_c[kWgMaxUnits], u_dw[kWgMaxUnits], u_dh[kWgMaxUnits], u_map[kWgMaxUnits];
#pragma unroll
        for (int i = 0; i < kWgMaxUnits; ++i) {
            const int u = u0 + (i < nu ? i : 0);
            const int ti = u / p.cblocks;
            const Tap tp = p.taps[ti];
            u_c[i] = (u - ti * p.cblocks) * 64;
            u_dw[i] = tp.dw; u_dh[i] = tp.dh; u_map[i] = tp.map;
        }
        for (int chunk = c0; chunk < c1; ++chunk) {
            // (a chain of plain integer divisions here is ~1000 cycles of dependent latency per chunk - as much as the chunk's MMAs)
            int n0 = 0, oh0 = 0, ow0 = 0, iw = 0, ih = 0, in_ = 0;
            const int m0 = chunk * p.pix;
            if (p.im2col) {
                const int t2 = p.mul_ow ? fast_div(m0, p.mul_ow) : m0 / p.OWt;
                in_ = p.mul_oh ? fast_div(t2, p.mul_oh) : t2 / p.OHt;
                iw = (m0 - t2 * p.OWt) * p.cstride + p.low_w;
                ih = (t2 - in_ * p.OHt) * p.cstride + p.low_h;
            } else {
                const int ng = chunk / tiles_hw, rem = chunk - ng * tiles_hw;
                const int hb = rem / p.tiles_w, wb = rem - hb * p.tiles_w;
                n0 = ng * p.nb; oh0 = hb * p.bh; ow0 = wb * p.bw;
            }
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* st = smem + (size_t)stage * stage_bytes;
            if (elect_one()) {
                mbar_expect_tx(&full_bar[stage], tx_bytes);
                if (p.im2col) {
                    for (int j = 0; j < dy_slots; ++j) tma_load_4d(st + j * p.slot_bytes, &maps.w, &full_bar[stage], nb_ * p.bn + j * 64, m0, 0, 0);
#pragma unroll
                    for (int i = 0; i < kWgMaxUnits; ++i)
                        if (i < nu)
                            tma_load_im2col(st + (dy_slots + i) * p.slot_bytes, &maps.in[0], &full_bar[stage], u_c[i], iw, ih, in_, u_dw[i] - p.low_w,
                                            u_dh[i] - p.low_h);
                } else {
                    for (int j = 0; j < dy_slots; ++j) tma_load_4d(st + j * p.slot_bytes, &maps.w, &full_bar[stage], nb_ * p.bn + j * 64, ow0, oh0, n0);
#pragma unroll
                    for (int i = 0; i < kWgMaxUnits; ++i)
                        if (i < nu)
                            tma_load_4d(st + (dy_slots + i) * p.slot_bytes, &maps.in[u_map[i]], &full_bar[stage], u_c[i], ow0 + u_dw[i], oh0 + u_dh[i], n0);
                }
            }
            __syncwarp();
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // whole warp runs the loop (uniform addresses / descriptors), one elected lane issues; a k-step of 16 pixels is
        // +2048 B in both MN-major operands, the second 64-wide block of either operand lies one slot further (LBO)
        const uint32_t idesc = make_idesc(128, p.bn, true, true);
        const uint64_t d_const = make_smem_desc(0, (uint32_t)p.slot_bytes, 1024);
        const uint32_t smem_base = smem_u32(smem);
        int stage = 0;
        uint32_t phase = 0;
        for (int chunk = c0; chunk < c1; ++chunk) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t sdy = smem_base + (uint32_t)stage * (uint32_t)stage_bytes;
            const uint64_t db0 = d_const | (uint64_t)(sdy >> 4);
            if (elect_one()) {
                for (int g = 0; g < npairs; ++g) {
                    const uint64_t da0 = d_const | (uint64_t)((sdy + (uint32_t)(dy_slots + 2 * g) * (uint32_t)p.slot_bytes) >> 4);
                    const uint32_t d_tmem = tmem_base + g * p.bn;
                    if (p.ksteps == 8) {
                        umma_bf16(d_tmem, da0, db0, idesc, chunk > c0 ? 1u : 0u);
#pragma unroll
                        for (int k = 1; k < 8; ++k) umma_bf16(d_tmem, da0 + k * 128, db0 + k * 128, idesc, 1u);
                    } else {
                        for (int k = 0; k < p.ksteps; ++k) umma_bf16(d_tmem, da0 + k * 128, db0 + k * 128, idesc, (chunk > c0 || k > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(done_bar);
        __syncwarp();
    } else {
        // epilogue: TMEM lane = row m of the pair (lanes 0..63 first unit, 64..127 second unit) -> fp32 partial rows
        const int quarter = warp & 3;
        mbar_wait(done_bar, 0);
        tc_fence_after();
        for (int g = 0; g < npairs; ++g) {
            const int ui = 2 * g + (quarter >> 1);
            if (ui >= nu) continue;  // warp-uniform
            const long long m = (long long)(u0 + ui) * 64 + (quarter & 1) * 32 + lane;
            float* dst = p.partials + ((size_t)ks * p.Mtot + m) * p.Cout + nb_ * p.bn;
            for (int cc = 0; cc < p.bn; cc += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + g * p.bn + cc + ((uint32_t)(quarter * 32) << 16), r);
                tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; e += 4)
                    *reinterpret_cast<float4*>(dst + cc + e) =
                        make_float4(__uint_as_float(r[e]), __uint_as_float(r[e + 1]), __uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// dw[k][r][s][c] (element strides sk, sr, ss, sc) = sum over splits of partial[split][(r*S + s) * C + c][k], folded in a
// fixed order.  A block owns (8 >> lanes_log2) rows m x 32 columns k; the 8 warps are (row, split lane): each thread sums
// every lanes-th split of one element (4 independent accumulators, 128-byte coalesced reads along k), the lanes are then
// folded through shared memory.  Few splits (big weight tensors): lanes = 1, 8 rows per block, 16-byte runs along c in
// the output.  Many splits (small weight tensors, up to 148): lanes = 8, one row per block, so that the grid stays large.
__global__ void __launch_bounds__(256) conv_wgrad_fold_kernel(const float* __restrict__ partials, int ksplit, long long Mtot, int Cout,
                                                              bf16* __restrict__ dw, int C, int S, long long sk, long long sr, long long ss,
                                                              long long sc, int lanes_log2) {
    __shared__ float sm[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int lanes = 1 << lanes_log2, rows = 8 >> lanes_log2;
    const int row = ty >> lanes_log2, lane = ty & (lanes - 1);
    const long long m = (long long)blockIdx.x * rows + row;
    const int k0 = blockIdx.y * 32;
    const size_t stride = (size_t)Mtot * Cout;
    const float* src = partials + (size_t)m * Cout + k0 + tx;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int q = lane;
    for (; q + 3 * lanes < ksplit; q += 4 * lanes) {
        a0 += src[(size_t)q * stride];
        a1 += src[(size_t)(q + lanes) * stride];
        a2 += src[(size_t)(q + 2 * lanes) * stride];
        a3 += src[(size_t)(q + 3 * lanes) * stride];
    }
    for (; q < ksplit; q += lanes) a0 += src[(size_t)q * stride];
    sm[ty][tx] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    if ((int)threadIdx.x < rows * 32) {
        const int r = threadIdx.x % rows, kk = threadIdx.x / rows;  // m fastest: neighbouring threads write neighbouring c
        float acc = 0.f;
        for (int l = 0; l < lanes; ++l) acc += sm[r * lanes + l][kk];
        const long long mm = (long long)blockIdx.x * rows + r;
        const int tap = (int)(mm / C), c = (int)(mm - (long long)tap * C);
        const int fr = tap / S, fs = tap - fr * S;
        dw[(long long)(k0 + kk) * sk + fr * sr + fs * ss + c * sc] = __float2bfloat16(acc);
    }
}

// ------------------------------------------------------------------------------------------------ host side
int make_map4(CUtensorMap* map, const void* base, const cuuint64_t (&dims)[4], const cuuint64_t (&strides)[3], const cuuint32_t (&box)[4],
              const char* what) {
    tma_host::EncodeTiledFn enc = tma_host::encoder();
    DDL_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeTiled is not available from the driver", what);
    DDL_REQUIRE(((uintptr_t)base & 15) == 0, "%s: tensor must be 16-byte aligned", what);
    cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DDL_REQUIRE(r == CUDA_SUCCESS, "%s: cuTensorMapEncodeTiled (4-D) failed with CUresult %d (dims %llu %llu %llu %llu, box %u %u %u %u)", what,
                (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2], (unsigned long long)dims[3],
                box[0], box[1], box[2], box[3]);
    return 0;
}

// im2col-mode tensor map of an NHWC activation [N, H, W, C]: `pixels` positions x 64 channels per load; base positions run
// over [low, dim + up) in steps of `stride` in W and H (see tma_load_im2col)
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const int*,
                                   const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeIm2colFn im2col_encoder() {
    static EncodeIm2colFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        void* q = nullptr;
        cudaDriverEntryPointQueryResult r;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &q, cudaEnableDefault, &r) == cudaSuccess && r == cudaDriverEntryPointSuccess)
            fn = (EncodeIm2colFn)q;
        tried = true;
    }
    return fn;
}

int make_im2col_map(CUtensorMap* map, const void* base, int N, int H, int W, int C, int low_w, int low_h, int up_w, int up_h, int stride,
                    int pixels, const char* what) {
    EncodeIm2colFn enc = im2col_encoder();
    DDL_REQUIRE(enc != nullptr, "%s: cuTensorMapEncodeIm2col is not available from the driver", what);
    DDL_REQUIRE(((uintptr_t)base & 15) == 0, "%s: tensor must be 16-byte aligned", what);
    const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    const cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    const int lower[2] = {low_w, low_h}, upper[2] = {up_w, up_h};
    const cuuint32_t estr[4] = {1u, (cuuint32_t)stride, (cuuint32_t)stride, 1u};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, str, lower, upper, 64u, (cuuint32_t)pixels, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DDL_REQUIRE(r == CUDA_SUCCESS, "%s: cuTensorMapEncodeIm2col failed with CUresult %d (N %d H %d W %d C %d, corners [%d %d] [%d %d], stride %d)",
                what, (int)r, N, H, W, C, low_w, low_h, up_w, up_h, stride);
    return 0;
}

// how one operand tile (box) walks the pixel grid [N][GH][GW]
struct BoxGeo {
    int N, GH, GW;  // grid the boxes tile (flattened 1x1 layers: N = GH = 1, GW = all pixels)
    int bw, bh, nb, tiles_w, tiles_h, ngroups;
    int pix() const { return bw * bh * nb; }
    int tiles() const { return ngroups * tiles_h * tiles_w; }
};

BoxGeo choose_box(int N, int GH, int GW, int max_rows) {
    BoxGeo g;
    g.N = N; g.GH = GH; g.GW = GW;
    g.bw = GW < max_rows ? GW : max_rows;
    g.tiles_w = (GW + g.bw - 1) / g.bw;
    int bh = max_rows / g.bw;
    if (bh < 1) bh = 1;
    if (bh > GH) bh = GH;
    g.tiles_h = (GH + bh - 1) / bh;
    g.bh = (GH + g.tiles_h - 1) / g.tiles_h;  // same number of tiles, evenly filled
    g.nb = 1;
    if (g.bh == GH && g.bw == GW) {
        g.nb = max_rows / (g.bw * g.bh);
        if (g.nb > N) g.nb = N;
        if (g.nb < 1) g.nb = 1;
    }
    g.ngroups = (N + g.nb - 1) / g.nb;
    return g;
}

struct ConvShape {
    int N, H, W, C, Cout, R, S, stride, pad, OH, OW;
    bool flat;  // 1x1 / stride 1: no spatial structure
};

bool conv_ok(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad) {
    if (!(N >= 1 && H >= 1 && W >= 1 && R == S && ((R == 1 && pad == 0) || (R == 3 && pad == 1)) && (stride == 1 || stride == 2))) return false;
    if (C % 64 != 0 || Cout % 64 != 0 || C < 64 || Cout < 64) return false;
    if (!(Cout <= 256 || Cout % 256 == 0) || !(C <= 256 || C % 256 == 0)) return false;
    if (stride == 2 && (H % 2 != 0 || W % 2 != 0)) return false;
    if ((long long)N * H * W >= (1ll << 31) - 256) return false;
    return true;
}

ConvShape make_shape(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad) {
    ConvShape s;
    s.N = N; s.H = H; s.W = W; s.C = C; s.Cout = Cout; s.R = R; s.S = S; s.stride = stride; s.pad = pad;
    s.OH = (H + 2 * pad - R) / stride + 1;
    s.OW = (W + 2 * pad - S) / stride + 1;
    s.flat = (R == 1 && stride == 1);
    return s;
}

inline int floor_div2(int e) { return e >= 0 ? e / 2 : -((-e + 1) / 2); }

// tensor maps of an NHWC activation [N, H, W, C] read with boxes [64, bw, bh, nb]: one map (stride 1), one flattened map
// (1x1 layers), or four parity maps (stride 2: map ph * 2 + pw sees the pixels (2i + ph, 2j + pw))
int make_act_maps(CUtensorMap* out4, const void* base, int N, int H, int W, int C, int stride, bool flat, const BoxGeo& g, const char* what) {
    const cuuint32_t box[4] = {64u, (cuuint32_t)g.bw, (cuuint32_t)g.bh, (cuuint32_t)g.nb};
    if (flat) {
        const cuuint64_t P = (cuuint64_t)N * H * W;
        const cuuint64_t dims[4] = {(cuuint64_t)C, P, 1, 1};
        const cuuint64_t str[3] = {(cuuint64_t)C * 2, P * C * 2, P * C * 2};
        if (int e = make_map4(&out4[0], base, dims, str, box, what)) return e;
        for (int i = 1; i < 4; ++i) out4[i] = out4[0];
        return 0;
    }
    if (stride == 1) {
        const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
        const cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
        if (int e = make_map4(&out4[0], base, dims, str, box, what)) return e;
        for (int i = 1; i < 4; ++i) out4[i] = out4[0];
        return 0;
    }
    for (int ph = 0; ph < 2; ++ph)
        for (int pw = 0; pw < 2; ++pw) {
            const char* b = (const char*)base + ((size_t)ph * W + pw) * C * 2;
            const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)(W / 2), (cuuint64_t)(H / 2), (cuuint64_t)N};
            const cuuint64_t str[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 2};
            if (int e = make_map4(&out4[ph * 2 + pw], b, dims, str, box, what)) return e;
        }
    return 0;
}

// taps of the forward convolution (also the x-side taps of the weight gradient): input pixel = stride * o + r - pad
int forward_taps(const ConvShape& s, Tap* taps) {
    int n = 0;
    for (int r = 0; r < s.R; ++r)
        for (int q = 0; q < s.S; ++q) {
            Tap t;
            if (s.stride == 1) {
                t.dh = (int8_t)(r - s.pad); t.dw = (int8_t)(q - s.pad); t.map = 0;
            } else {
                const int eh = r - s.pad, ew = q - s.pad;
                const int dh = floor_div2(eh), dw = floor_div2(ew);
                t.dh = (int8_t)dh; t.dw = (int8_t)dw;
                t.map = (int8_t)((eh - 2 * dh) * 2 + (ew - 2 * dw));
            }
            t.wtap = (int8_t)(r * s.S + q);
            taps[n++] = t;
        }
    return n;
}

int g_im2col = 1;  // activation tiles through im2col-mode TMA (dense 128-pixel tiles) instead of tiled boxes
int g_ring_group = 1;  // ring slots released per tcgen05.commit (1 = every k-block; larger values measured slower)
int g_wgrad_rows = 0;  // 0 = auto, 64 / 128 = forced pixel-chunk size of the weight gradient (benchmarks)

int g_stay = 1;  // resident weight panel where it fits (ddl_conv_set_stay)
long long* g_conv_dbg = nullptr;  // ddl_conv_set_debug
int g_conv_ablate = 0;

int launch_data(bool b_mn, const ConvMaps& maps, DataParams& p, cudaStream_t stream) {
    p.groups = p.bn / 64;
    const int fixed = kEpiWarps * kStgBytes + 512 + 1024;
    // Resident weights: one class, 256-column blocks, the whole [ntaps * cblocks] x bn panel within 128 KB (K <= 256) and four
    // activation stages beside it.  Measured gain is small (256 -> 1024 at 14 x 14: 49.2 -> 47.1 us forward; narrower column
    // blocks got slower): what bounded these short-K layers was the index arithmetic of the epilogue (see `dense` below and
    // benchmarks/conv_ablate.py), not the weight re-loads.
    const long long panel = (long long)p.cls_ntaps[0] * p.cblocks * p.bn * 128;
    p.stay = 0;
    p.panel_bytes = 0;
    if (g_stay && p.ncls == 1 && p.bn == 256 && panel <= 128 * 1024 && (227 * 1024 - fixed - panel) / kStageA >= 4) {
        p.stay = 1;
        p.panel_bytes = (uint32_t)panel;
    }
    const int stage_bytes = p.stay ? kStageA : kStageA + p.bn * 128;
    p.ablate = g_conv_ablate;
    p.dbg = g_conv_dbg;
    auto magic = [](long long xmax, int d) -> unsigned long long {  // 0 when the exactness bound x * d < 2^40 does not hold
        if (d <= 0 || xmax * (long long)d >= (1ll << 40) || xmax / d >= (1ll << 23)) return 0ull;  // (second bound: no 64-bit overflow)
        return ((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d;
    };
    const long long max_pos = (long long)p.N * p.OHt * p.OWt + kTileM;
    p.mul_ow = magic(max_pos, p.OWt);
    p.mul_oh = magic(max_pos, p.OHt);
    p.mul_tpc = magic((long long)p.num_tiles + 1, p.tiles_per_cls);
    // dense tiles: one class whose positions are stored with pixel stride 1 and that is tiled 128 positions at a time, either by
    // im2col-mode loads or as a flat 1x1 layer (one box row of 128 pixels per tile)
    p.dense = 0;
    if (p.ncls == 1 && p.os == 1 && (p.im2col || (p.N == 1 && p.OHt == 1 && p.bw == kTileM && p.bh == 1 && p.nb == 1))) {
        p.dense = 1;
        if (!p.im2col) p.cls_pix = p.OWt;
    }
    p.stages = (227 * 1024 - fixed - (int)p.panel_bytes) / stage_bytes;
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    p.group = 1;
    DDL_REQUIRE(p.stages >= 2, "conv: tile does not fit in shared memory");
    const int smem = fixed + (int)p.panel_bytes + p.stages * stage_bytes;
    static int attr_smem[2] = {0, 0};
    if (smem > attr_smem[b_mn ? 1 : 0]) {
        if (b_mn) DDL_CHECK(cudaFuncSetAttribute(conv_data_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        else DDL_CHECK(cudaFuncSetAttribute(conv_data_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem[b_mn ? 1 : 0] = smem;
    }
    int m_groups = DDL_SM_COUNT / p.num_n_blocks;
    if (m_groups > p.num_tiles) m_groups = p.num_tiles;
    if (m_groups < 1) m_groups = 1;
    const int grid = m_groups * p.num_n_blocks;
    if (b_mn) conv_data_kernel<true><<<grid, kThreads, smem, stream>>>(maps, p);
    else conv_data_kernel<false><<<grid, kThreads, smem, stream>>>(maps, p);
    DDL_LAUNCH_CHECK();
    return 0;
}

void fill_geo(DataParams& p, const BoxGeo& g) {
    p.N = g.N; p.OHt = g.GH; p.OWt = g.GW;
    p.bw = g.bw; p.bh = g.bh; p.nb = g.nb;
    p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h;
    p.tiles_per_cls = g.tiles();
    p.a_bytes = (uint32_t)g.pix() * 128u;
}

int fwd_m_groups(const ConvShape& s) {
    const BoxGeo g = s.flat ? choose_box(1, 1, s.N * s.OH * s.OW, kTileM) : choose_box(s.N, s.OH, s.OW, kTileM);
    const int bn = s.Cout <= 256 ? s.Cout : 256;
    int m_groups = DDL_SM_COUNT / (s.Cout / bn);
    const long long tiles = (g_im2col && !s.flat) ? ((long long)s.N * s.OH * s.OW + kTileM - 1) / kTileM : g.tiles();
    if (m_groups > tiles) m_groups = (int)tiles;
    return m_groups < 1 ? 1 : m_groups;
}

// forward taps as offsets in input pixels (im2col mode): input pixel = stride * o + (r - pad)
int im2col_fwd_taps(const ConvShape& s, Tap* taps) {
    int n = 0;
    for (int r = 0; r < s.R; ++r)
        for (int q = 0; q < s.S; ++q) {
            Tap t;
            t.dh = (int8_t)(r - s.pad); t.dw = (int8_t)(q - s.pad); t.map = 0; t.wtap = (int8_t)(r * s.S + q);
            taps[n++] = t;
        }
    return n;
}

// Turn a filled DataParams (classes, taps in INPUT-pixel offsets, class grid N x OHt x OWt) into im2col mode over the
// activation [aN, aH, aW, aC] read with traversal stride `cstride`, and build its tensor map.
int use_im2col(DataParams& p, CUtensorMap* map, const void* act, int aN, int aH, int aW, int aC, int cstride, int pixels, const char* what) {
    int low_w = 127, low_h = 127;
    for (int c = 0; c < p.ncls; ++c)
        for (int i = 0; i < p.cls_ntaps[c]; ++i) {
            low_w = p.taps[c][i].dw < low_w ? p.taps[c][i].dw : low_w;
            low_h = p.taps[c][i].dh < low_h ? p.taps[c][i].dh : low_h;
        }
    p.im2col = 1;
    p.cstride = cstride;
    p.low_w = low_w; p.low_h = low_h;
    p.cls_pix = (long long)p.N * p.OHt * p.OWt;
    p.tiles_per_cls = (int)((p.cls_pix + pixels - 1) / pixels);
    p.tiles_w = p.tiles_per_cls; p.tiles_h = 1;  // (unused by the im2col decode; kept non-zero)
    p.num_tiles = p.ncls * p.tiles_per_cls;
    p.a_bytes = (uint32_t)pixels * 128u;
    // the base positions must number exactly OWt x OHt per image
    const int up_w = (p.OWt - 1) * cstride + 1 + low_w - aW, up_h = (p.OHt - 1) * cstride + 1 + low_h - aH;
    return make_im2col_map(map, act, aN, aH, aW, aC, low_w, low_h, up_w, up_h, cstride, pixels, what);
}

// ---- pair kernel launch plan -----------------------------------------------------------------------------
int g_pair_mode = 0;  // 0 = never use the CTA-pair kernel, 1 = pair kernel (halo tiles where possible), 2 = pair kernel, no halo

bool ptr_is_local(const void* ptr) {
    cudaPointerAttributes attr;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeDevice && attr.device == dev;
}

struct PairPlan {
    bool halo, im2col;
    BoxGeo g;       // the TMA box and how it tiles the grid (halo mode: box includes the halo; g.tiles() counts output tiles)
    int th, pitch, rows_img;
    int a_slot, a_stages, b_stages, bgroup;
    int pair_tiles_per_cls, pgroups;
};

// d: classes / taps / grid / bn already filled in.  halo_ok: 3x3 stride-1 geometry (all taps are shifts of one box).
bool plan_pair(const DataParams& d, bool halo_ok, bool flat, PairPlan& pl) {
    if (g_pair_mode == 0 || d.bn % 128 != 0) return false;
    pl.im2col = false;
    const int b_slot = (d.bn / 2) * 128;
    const int avail = 227 * 1024 - kEpiWarps * kStgBytes - 1024 - 1024;
    pl.halo = false;
    if (halo_ok && g_pair_mode == 1 && d.ncls == 1 && d.OWt + 1 <= 128) {
        const int pitch = d.OWt + 1;
        int th = 128 / pitch;
        if (th > d.OHt) th = d.OHt;
        const int tiles_h = (d.OHt + th - 1) / th;
        th = (d.OHt + tiles_h - 1) / tiles_h;
        const int rows_img = (th + 2) * pitch;
        int nb = 1;
        if (th == d.OHt) {
            nb = (128 - th * pitch) / rows_img + 1;
            if (nb > d.N) nb = d.N;
            if (nb < 1) nb = 1;
        }
        BoxGeo g;
        g.N = d.N; g.GH = d.OHt; g.GW = d.OWt;
        g.bw = pitch; g.bh = th + 2; g.nb = nb;
        g.tiles_w = 1; g.tiles_h = tiles_h; g.ngroups = (d.N + nb - 1) / nb;
        const int box_rows = g.pix();
        int slot_rows = box_rows + 1;
        if (slot_rows < 2 * pitch + 2 + 128) slot_rows = 2 * pitch + 2 + 128;
        const int a_slot = (slot_rows * 128 + 1023) / 1024 * 1024;
        int b_stages = (avail - 2 * a_slot) / b_slot;
        if (b_stages > kPairMaxStages) b_stages = kPairMaxStages;
        const int grp = g_ring_group;
        b_stages = b_stages / grp * grp;
        if (b_stages >= 3 && g.bh <= 256 && g.nb <= 256) {
            pl.halo = true;
            pl.g = g; pl.th = th; pl.pitch = pitch; pl.rows_img = rows_img;
            pl.a_slot = a_slot; pl.a_stages = 2; pl.b_stages = b_stages; pl.bgroup = grp;
        }
    }
    if (!pl.halo) {
        pl.g = choose_box(d.N, d.OHt, d.OWt, kTileM);
        pl.th = pl.g.bh; pl.pitch = pl.g.bw; pl.rows_img = pl.g.bw * pl.g.bh;
        pl.a_slot = kStageA;
        int st = avail / (kStageA + b_slot);
        if (st > kPairMaxStages) st = kPairMaxStages;
        const int grp = g_ring_group;
        st = st / grp * grp;
        if (st < 2) return false;
        pl.a_stages = pl.b_stages = st;
        pl.bgroup = grp;
    }
    int tiles_per_cls = pl.g.tiles();
    if (!pl.halo && g_im2col && !flat) {
        pl.im2col = true;  // dense 128-position tiles of the class grid
        tiles_per_cls = (int)(((long long)d.N * d.OHt * d.OWt + kTileM - 1) / kTileM);
    }
    pl.pair_tiles_per_cls = (tiles_per_cls + 1) / 2;
    int pgroups = (DDL_SM_COUNT / 2) / d.num_n_blocks;
    if (pgroups > pl.pair_tiles_per_cls * d.ncls) pgroups = pl.pair_tiles_per_cls * d.ncls;
    pl.pgroups = pgroups < 1 ? 1 : pgroups;
    return true;
}

// act: the activation tensor the boxes read ([aN, aH, aW, aC] NHWC, map stride act_stride); w: [w_rows][w_inner] weights
int launch_pair(bool b_mn, const DataParams& d0, const PairPlan& pl, const void* act, int aN, int aH, int aW, int aC, int act_stride, bool flat,
                const void* w, long long w_rows, long long w_inner, cudaStream_t stream) {
    DataParams d = d0;
    ConvMaps maps;
    if (pl.im2col) {
        if (int e = use_im2col(d, &maps.in[0], act, aN, aH, aW, aC, act_stride, kTileM, "conv (pair, im2col)")) return e;
        for (int i = 1; i < 4; ++i) maps.in[i] = maps.in[0];
    }
    PairParams p = {};
    p.out = d.out; p.partials = d.partials;
    p.ncls = d.ncls;
    for (int i = 0; i < 4; ++i) {
        p.cls_ntaps[i] = d.cls_ntaps[i]; p.cls_oph[i] = d.cls_oph[i]; p.cls_opw[i] = d.cls_opw[i];
        for (int j = 0; j < 9; ++j) p.taps[i][j] = d.taps[i][j];
    }
    p.N = d.N; p.OHt = d.OHt; p.OWt = d.OWt; p.OHf = d.OHf; p.OWf = d.OWf; p.os = d.os; p.Cn = d.Cn;
    p.bw = pl.g.bw; p.bh = pl.g.bh; p.nb = pl.g.nb;
    p.tiles_w = pl.g.tiles_w; p.tiles_h = pl.g.tiles_h;
    p.tiles_per_cls = pl.im2col ? d.tiles_per_cls : pl.g.tiles();
    p.pair_tiles_per_cls = pl.pair_tiles_per_cls;
    p.im2col = pl.im2col ? 1 : 0;
    p.cstride = d.cstride; p.low_w = d.low_w; p.low_h = d.low_h; p.cls_pix = d.cls_pix;
    p.num_pair_tiles = pl.pair_tiles_per_cls * d.ncls;
    p.th = pl.th; p.pitch = pl.pitch; p.rows_img = pl.rows_img;
    p.num_n_blocks = d.num_n_blocks; p.bn = d.bn; p.groups = d.bn / 64;
    p.cblocks = d.cblocks; p.w_tap_stride = d.w_tap_stride;
    p.halo = pl.halo ? 1 : 0;
    p.a_slot_bytes = pl.a_slot; p.a_stages = pl.a_stages; p.b_stages = pl.b_stages;
    p.bgroup = pl.bgroup;
    p.a_bytes = pl.im2col ? (uint32_t)kTileM * 128u : (uint32_t)pl.g.pix() * 128u;
    p.dbg = g_conv_dbg;
    p.ablate = g_conv_ablate;
    if (!pl.im2col)
        if (int e = make_act_maps(maps.in, act, aN, aH, aW, aC, act_stride, flat, pl.g, "conv (pair) activations")) return e;
    if (int e = tma_host::encode_2d(&maps.w, w, w_rows, w_inner, w_inner * 2, b_mn ? 64 : d.bn / 2, "conv (pair) weights")) return e;
    const int smem = pl.a_stages * pl.a_slot + pl.b_stages * (d.bn / 2) * 128 + kEpiWarps * kStgBytes + 1024 + 1024;
    static int attr_smem[2] = {0, 0};
    if (smem > attr_smem[b_mn ? 1 : 0]) {
        if (b_mn) DDL_CHECK(cudaFuncSetAttribute(conv_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        else DDL_CHECK(cudaFuncSetAttribute(conv_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem[b_mn ? 1 : 0] = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pl.pgroups * d.num_n_blocks);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (b_mn) DDL_CHECK(cudaLaunchKernelEx(&cfg, conv_pair_kernel<true>, maps, p));
    else DDL_CHECK(cudaLaunchKernelEx(&cfg, conv_pair_kernel<false>, maps, p));
    return 0;
}

// ---- halo kernel launch plan (conv_halo.cuh) -------------------------------------------------------------
int g_halo = 1;        // 3x3 / stride-1 / pad-1 layers on halo tiles; 0 = im2col kernel
int g_halo_mt = 0;     // pixel tiles per weight tile: 0 = auto, 1 / 2 forced (benchmarks)
int g_halo_min_w = 28; // narrower images fill too few of the 128 MMA rows per tile

struct HaloPlan {
    int th, pitch, tiles_h, num_tiles, mt, bn, nblk, cblocks, a_stages, b_stages, a_slot, box_rows, m_groups, smem;
};

// Cred = reduction channels (forward: C, data gradient: Cout), Cn = output channels of the contraction
bool plan_halo(int N, int H, int W, int Cred, int Cn, int R, int stride, int pad, HaloPlan& h) {
    if (!g_halo || R != 3 || stride != 1 || pad != 1) return false;
    if (W < g_halo_min_w || W + 1 > 128 || Cn > 256 || (long long)N * H > (1 << 28)) return false;
    h.pitch = W + 1;
    int th = 128 / h.pitch;
    if (th > H) th = H;
    h.tiles_h = (H + th - 1) / th;
    h.th = (H + h.tiles_h - 1) / h.tiles_h;
    h.num_tiles = N * h.tiles_h;
    h.bn = Cn;
    h.nblk = 1;
    h.cblocks = Cred / 64;
    h.box_rows = (h.th + 2) * h.pitch;
    int slot_rows = h.box_rows + 1;
    if (slot_rows < 2 * h.pitch + 2 + 128) slot_rows = 2 * h.pitch + 2 + 128;
    h.a_slot = (slot_rows * 128 + 1023) / 1024 * 1024;
    const int b_slot = h.bn * 128, b_all = 9 * h.cblocks * b_slot;
    for (int mt = 2; mt >= 1; --mt) {
        if (g_halo_mt && mt != g_halo_mt) continue;
        if (2 * mt * h.bn > 512) continue;
        const int fixed = 4 * (h.bn / 64) * mt * kStgBytes + 512 + 1024;
        const int avail = 227 * 1024 - fixed;
        const bool stay = b_all + 2 * mt * h.a_slot <= avail;
        int a_stages, b_stages;
        if (stay) {
            b_stages = 0;
            a_stages = (avail - b_all) / (mt * h.a_slot);
            if (a_stages > 4) a_stages = 4;
        } else {
            a_stages = 2;
            b_stages = (avail - 2 * mt * h.a_slot) / b_slot;
            if (b_stages > kHaloMaxStages) b_stages = kHaloMaxStages;
            if (b_stages < 3) continue;
        }
        h.mt = mt; h.a_stages = a_stages; h.b_stages = b_stages;
        h.smem = fixed + a_stages * mt * h.a_slot + (b_stages ? b_stages : 9 * h.cblocks) * b_slot;
        const int units = (h.num_tiles + mt - 1) / mt;
        h.m_groups = DDL_SM_COUNT < units ? DDL_SM_COUNT : units;
        return true;
    }
    return false;
}

// act [N, H, W, Cact] NHWC; w [w_rows][w_inner]; out [N, H, W, Cn]
int launch_halo(bool b_mn, const HaloPlan& h, const void* act, int N, int H, int W, int Cact, const void* w, long long w_rows, long long w_inner,
                int w_tap_stride, void* out, float* partials, int Cn, cudaStream_t stream) {
    ConvMaps maps;
    {
        const cuuint64_t dims[4] = {(cuuint64_t)Cact, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
        const cuuint64_t str[3] = {(cuuint64_t)Cact * 2, (cuuint64_t)W * Cact * 2, (cuuint64_t)H * W * Cact * 2};
        const cuuint32_t box[4] = {64u, (cuuint32_t)h.pitch, (cuuint32_t)(h.th + 2), 1u};
        if (int e = make_map4(&maps.in[0], act, dims, str, box, "conv (halo) activations")) return e;
        for (int i = 1; i < 4; ++i) maps.in[i] = maps.in[0];
    }
    if (int e = tma_host::encode_2d(&maps.w, w, w_rows, w_inner, w_inner * 2, b_mn ? 64 : h.bn, "conv (halo) weights")) return e;
    HaloParams p = {};
    p.out = (bf16*)out; p.partials = partials;
    p.N = N; p.H = H; p.W = W; p.Cn = Cn;
    p.th = h.th; p.pitch = h.pitch; p.tiles_h = h.tiles_h; p.num_tiles = h.num_tiles;
    p.mt = h.mt; p.num_units = (h.num_tiles + h.mt - 1) / h.mt;
    p.num_n_blocks = h.nblk; p.bn = h.bn; p.groups = h.bn / 64;
    p.cblocks = h.cblocks; p.w_tap_stride = w_tap_stride;
    p.flip = b_mn ? 1 : 0;
    p.a_stages = h.a_stages; p.b_stages = h.b_stages;
    p.ablate = g_conv_ablate;
    p.a_slot = (uint32_t)h.a_slot; p.box_rows = (uint32_t)h.box_rows; p.a_box_bytes = (uint32_t)h.box_rows * 128u;
    const bool stay = h.b_stages == 0;
    void (*kern)(const ConvMaps, const HaloParams) = nullptr;
    if (b_mn) kern = h.mt == 2 ? (stay ? conv_halo_kernel<true, 2, true> : conv_halo_kernel<true, 2, false>)
                               : (stay ? conv_halo_kernel<true, 1, true> : conv_halo_kernel<true, 1, false>);
    else kern = h.mt == 2 ? (stay ? conv_halo_kernel<false, 2, true> : conv_halo_kernel<false, 2, false>)
                          : (stay ? conv_halo_kernel<false, 1, true> : conv_halo_kernel<false, 1, false>);
    static int attr_smem[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int variant = (b_mn ? 4 : 0) + (h.mt == 2 ? 2 : 0) + (stay ? 1 : 0);
    if (h.smem > attr_smem[variant]) {
        DDL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, h.smem));
        attr_smem[variant] = h.smem;
    }
    kern<<<h.m_groups * h.nblk, kThreads, h.smem, stream>>>(maps, p);
    DDL_LAUNCH_CHECK();
    return 0;
}

// classes / taps / grid of the forward convolution (shared by the launch and the partial-row query)
void fill_fwd(DataParams& p, const ConvShape& s, const BoxGeo& g) {
    p.ncls = 1;
    p.cls_ntaps[0] = (g_im2col && !s.flat) ? im2col_fwd_taps(s, p.taps[0]) : forward_taps(s, p.taps[0]);
    fill_geo(p, g);
    p.OHf = g.GH; p.OWf = g.GW; p.os = 1;
    p.Cn = s.Cout;
    p.num_tiles = p.tiles_per_cls;
    p.bn = s.Cout <= 256 ? s.Cout : 256;
    p.num_n_blocks = s.Cout / p.bn;
    p.cblocks = s.C / 64;
    p.w_tap_stride = s.C;
}

// rows of BatchNorm partials the forward writes (the pair kernel when the shape allows it, else the one-CTA kernel)
int fwd_partial_rows(const ConvShape& s) {
    const BoxGeo g = s.flat ? choose_box(1, 1, s.N * s.OH * s.OW, kTileM) : choose_box(s.N, s.OH, s.OW, kTileM);
    HaloPlan hp;
    if (plan_halo(s.N, s.OH, s.OW, s.C, s.Cout, s.R, s.stride, s.pad, hp)) return hp.m_groups;
    DataParams p = {};
    fill_fwd(p, s, g);
    PairPlan pl;
    if (plan_pair(p, s.R == 3 && s.stride == 1, s.flat, pl)) return 2 * pl.pgroups;
    return fwd_m_groups(s);
}

// weight-gradient launch plan (shared by the scratch-size query and the launch)
struct WgradPlan {
    bool im2col;
    BoxGeo g;
    int bn, nblk, U, G, mgroups, ksteps, slot_bytes, stages, ksplit, cps;
};

bool plan_wgrad(const ConvShape& s, WgradPlan& w) {
    w.bn = s.Cout <= 256 ? s.Cout : 256;
    w.nblk = s.Cout / w.bn;
    w.U = s.R * s.S * (s.C / 64);
    int Gmax = 512 / w.bn;
    if (Gmax > kWgMaxUnits / 2) Gmax = kWgMaxUnits / 2;
    if (Gmax > (w.U + 1) / 2) Gmax = (w.U + 1) / 2;
    const int avail = 227 * 1024 - 1024 - 512;
    // 128-pixel chunks first (with as many accumulator pairs as still leave two stages), 64-pixel chunks only when even one
    // pair does not fit: the cost per chunk (barrier round trip, commit) outweighs both a shallower prefetch and fewer
    // accumulators per CTA (measured, benchmarks/conv_bench.py --wgrad-rows)
    int G = 0;
    const int rows_opts[2] = {128, 64};
    for (int o = 0; o < 2 && G == 0; ++o) {
        const int rows = g_wgrad_rows ? g_wgrad_rows : rows_opts[o];
        // im2col mode chunks the flattened output positions exactly like a 1x1 layer does
        const BoxGeo g = (s.flat || g_im2col) ? choose_box(1, 1, s.N * s.OH * s.OW, rows) : choose_box(s.N, s.OH, s.OW, rows);
        const int ksteps = (g.pix() + 15) / 16;
        const int slot = ksteps * 2048;
        for (int gg = Gmax; gg >= 1; --gg) {
            int stages = avail / ((w.bn / 64 + 2 * gg) * slot);
            if (stages > 4) stages = 4;
            if (stages >= 2) {
                G = gg;
                w.g = g; w.ksteps = ksteps; w.slot_bytes = slot; w.stages = stages;
                break;
            }
        }
        if (g_wgrad_rows) break;
    }
    if (G == 0) return false;
    w.G = G;
    w.im2col = g_im2col && !s.flat;
    w.mgroups = (w.U + 2 * G - 1) / (2 * G);
    const int items = w.nblk * w.mgroups;
    const int chunks = w.g.tiles();
    int ksplit = DDL_SM_COUNT / items;
    if (ksplit < 1) ksplit = 1;
    if (ksplit > chunks) ksplit = chunks;
    w.cps = (chunks + ksplit - 1) / ksplit;
    w.ksplit = (chunks + w.cps - 1) / w.cps;
    return true;
}

}  // namespace

DDL_API int ddl_conv_supported(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad) {
    return conv_ok(N, H, W, C, Cout, R, S, stride, pad) ? 1 : 0;
}

DDL_API int ddl_conv_fwd_groups(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad) {
    if (!conv_ok(N, H, W, C, Cout, R, S, stride, pad)) return 0;
    return fwd_partial_rows(make_shape(N, H, W, C, Cout, R, S, stride, pad));
}

DDL_API int ddl_conv_set_im2col(int on) {
    g_im2col = on ? 1 : 0;
    return 0;
}

DDL_API int ddl_conv_set_tuning(int wgrad_rows, int pair_mode) {
    DDL_REQUIRE(wgrad_rows == 0 || wgrad_rows == 64 || wgrad_rows == 128, "conv tuning: wgrad_rows must be 0, 64 or 128");
    DDL_REQUIRE(pair_mode >= 0 && pair_mode <= 2, "conv tuning: pair_mode must be 0 (one-CTA kernel), 1 (CTA pairs + halo tiles) or 2 (CTA pairs)");
    g_wgrad_rows = wgrad_rows;
    g_pair_mode = pair_mode;
    return 0;
}

DDL_API int ddl_conv_set_stay(int on) {
    g_stay = on ? 1 : 0;
    return 0;
}

DDL_API int ddl_conv_set_halo(int on, int mt, int min_w) {
    DDL_REQUIRE(mt >= 0 && mt <= 2, "conv tuning: halo mt must be 0 (auto), 1 or 2");
    g_halo = on ? 1 : 0;
    g_halo_mt = mt;
    g_halo_min_w = min_w > 0 ? min_w : 28;
    return 0;
}

// development aid: device buffer of [296][8] int64 that the next pair-kernel launches fill with per-CTA wait-cycle counters
// (producer: total, wait A slot, wait B slot; MMA issuer: total, wait A data, wait B data, wait accumulator); null = off
DDL_API int ddl_conv_set_debug(void* buf, int ablate) {
    g_conv_dbg = (long long*)buf;
    g_conv_ablate = ablate;  // bit 0: epilogue drops staging / stores, bit 1: no TMA loads (results are garbage: timing only)
    return 0;
}

DDL_API int ddl_conv_fwd(const void* x, const void* w, void* y, float* partials, int N, int H, int W, int C, int Cout, int R, int S, int stride,
                         int pad, void* stream_) {
    DDL_REQUIRE(conv_ok(N, H, W, C, Cout, R, S, stride, pad), "conv fwd: unsupported shape (N %d, %dx%d, C %d -> %d, k %d, stride %d, pad %d)", N,
                H, W, C, Cout, R, stride, pad);
    DDL_REQUIRE(x && w && y, "conv fwd: null pointer");
    const ConvShape s = make_shape(N, H, W, C, Cout, R, S, stride, pad);
    {
        HaloPlan hp;
        if (plan_halo(N, s.OH, s.OW, C, Cout, R, stride, pad, hp))
            return launch_halo(false, hp, x, N, H, W, C, w, Cout, (long long)R * S * C, C, y, partials, Cout, (cudaStream_t)stream_);
    }
    const BoxGeo g = s.flat ? choose_box(1, 1, N * s.OH * s.OW, kTileM) : choose_box(N, s.OH, s.OW, kTileM);
    DataParams p = {};
    p.out = (bf16*)y;
    p.partials = partials;
    fill_fwd(p, s, g);
    cudaStream_t stream = (cudaStream_t)stream_;
    PairPlan pl;
    const bool pair_shape = plan_pair(p, R == 3 && stride == 1, s.flat, pl);
    if (pair_shape && ptr_is_local(x) && ptr_is_local(w))
        return launch_pair(false, p, pl, x, N, H, W, C, stride, s.flat, w, Cout, (long long)R * S * C, stream);
    ConvMaps maps;
    if (g_im2col && !s.flat) {
        if (int e = use_im2col(p, &maps.in[0], x, N, H, W, C, stride, kTileM, "conv fwd (im2col)")) return e;
        for (int i = 1; i < 4; ++i) maps.in[i] = maps.in[0];
    } else {
        if (int e = make_act_maps(maps.in, x, N, H, W, C, stride, s.flat, g, "conv fwd")) return e;
    }
    if (int e = tma_host::encode_2d(&maps.w, w, Cout, (long long)R * S * C, (long long)R * S * C * 2, p.bn, "conv fwd weights")) return e;
    if (int e = launch_data(false, maps, p, stream)) return e;
    if (pair_shape && partials) {
        // the caller sized the partials for the pair kernel (ddl_conv_fwd_groups): the rows this kernel does not write are zeros
        const int rows = fwd_m_groups(s), want = 2 * pl.pgroups;
        if (want > rows) DDL_CHECK(cudaMemsetAsync(partials + (size_t)rows * 2 * Cout, 0, (size_t)(want - rows) * 2 * Cout * sizeof(float), stream));
    }
    return 0;
}

// dx [N, H, W, C] from dy [N, OH, OW, Cout] and w [Cout][R][S][C]
DDL_API int ddl_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad,
                           void* stream_) {
    DDL_REQUIRE(conv_ok(N, H, W, C, Cout, R, S, stride, pad), "conv dgrad: unsupported shape (N %d, %dx%d, C %d -> %d, k %d, stride %d, pad %d)",
                N, H, W, C, Cout, R, stride, pad);
    DDL_REQUIRE(dy && w && dx, "conv dgrad: null pointer");
    const ConvShape s = make_shape(N, H, W, C, Cout, R, S, stride, pad);
    cudaStream_t stream = (cudaStream_t)stream_;
    {
        HaloPlan hp;
        if (plan_halo(N, H, W, Cout, C, R, stride, pad, hp))
            return launch_halo(true, hp, dy, N, s.OH, s.OW, Cout, w, Cout, (long long)R * S * C, C, dx, nullptr, C, stream);
    }
    // boxes walk the dy grid; a class of the stride-2 case covers H/2 x W/2 = OH x OW input pixels
    const BoxGeo g = s.flat ? choose_box(1, 1, N * s.OH * s.OW, kTileM) : choose_box(N, s.OH, s.OW, kTileM);
    DataParams p = {};
    p.out = (bf16*)dx;
    p.partials = nullptr;
    fill_geo(p, g);
    p.Cn = C;
    p.bn = C <= 256 ? C : 256;
    p.num_n_blocks = C / p.bn;
    p.cblocks = Cout / 64;
    p.w_tap_stride = C;
    if (stride == 1) {
        p.ncls = 1;
        int n = 0;
        for (int r = 0; r < R; ++r)
            for (int q = 0; q < S; ++q) {
                Tap t;
                t.dh = (int8_t)(pad - r); t.dw = (int8_t)(pad - q); t.map = 0; t.wtap = (int8_t)(r * S + q);
                p.taps[0][n++] = t;
            }
        p.cls_ntaps[0] = n;
        p.OHf = g.GH; p.OWf = g.GW; p.os = 1;
    } else {
        // input pixel (2a + ph, 2b + pw) receives dy[(2a + ph + pad - r) / 2, ...] for the taps r of matching parity
        struct Cls { int ph, pw, n; Tap taps[9]; } cls[4];
        int nc = 0;
        bool uncovered = false;
        for (int ph = 0; ph < 2; ++ph)
            for (int pw = 0; pw < 2; ++pw) {
                Cls c;
                c.ph = ph; c.pw = pw; c.n = 0;
                for (int r = 0; r < R; ++r)
                    for (int q = 0; q < S; ++q) {
                        const int eh = ph + pad - r, ew = pw + pad - q;
                        if ((eh & 1) || (ew & 1)) continue;
                        Tap t;
                        t.dh = (int8_t)(eh / 2); t.dw = (int8_t)(ew / 2); t.map = 0; t.wtap = (int8_t)(r * S + q);
                        c.taps[c.n++] = t;
                    }
                if (c.n == 0) uncovered = true;
                else cls[nc++] = c;
            }
        // pixels that no tap reaches (1x1 / stride 2: three of four) get zeros
        if (uncovered) DDL_CHECK(cudaMemsetAsync(dx, 0, (size_t)N * H * W * C * 2, stream));
        std::sort(cls, cls + nc, [](const Cls& a, const Cls& b) { return a.n > b.n; });  // heavy classes first
        p.ncls = nc;
        for (int i = 0; i < nc; ++i) {
            p.cls_ntaps[i] = cls[i].n;
            p.cls_oph[i] = cls[i].ph;
            p.cls_opw[i] = cls[i].pw;
            for (int j = 0; j < cls[i].n; ++j) p.taps[i][j] = cls[i].taps[j];
        }
        p.OHf = H; p.OWf = W; p.os = 2;
    }
    p.num_tiles = p.ncls * p.tiles_per_cls;
    PairPlan pl;
    if (plan_pair(p, R == 3 && stride == 1, s.flat, pl) && ptr_is_local(dy) && ptr_is_local(w))
        return launch_pair(true, p, pl, dy, N, s.OH, s.OW, Cout, 1, s.flat, w, Cout, (long long)R * S * C, stream);
    ConvMaps maps;
    if (g_im2col && !s.flat) {
        // the taps are already offsets in dy pixels; every class reads the same map
        if (int e = use_im2col(p, &maps.in[0], dy, N, s.OH, s.OW, Cout, 1, kTileM, "conv dgrad (im2col)")) return e;
        for (int i = 1; i < 4; ++i) maps.in[i] = maps.in[0];
    } else if (int e = make_act_maps(maps.in, dy, N, s.OH, s.OW, Cout, 1, s.flat, g, "conv dgrad")) return e;
    if (int e = tma_host::encode_2d(&maps.w, w, Cout, (long long)R * S * C, (long long)R * S * C * 2, 64, "conv dgrad weights")) return e;
    return launch_data(true, maps, p, stream);
}

DDL_API size_t ddl_conv_wgrad_scratch_bytes(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad) {
    if (!conv_ok(N, H, W, C, Cout, R, S, stride, pad)) return 0;
    WgradPlan w;
    if (!plan_wgrad(make_shape(N, H, W, C, Cout, R, S, stride, pad), w)) return 0;
    return (size_t)w.ksplit * (size_t)w.U * 64 * Cout * sizeof(float);
}

// dw [Cout, C, R, S] addressed with element strides (sk, sc, sr, ss) from x [N, H, W, C] and dy [N, OH, OW, Cout]
DDL_API int ddl_conv_wgrad(const void* x, const void* dy, void* dw, long long sk, long long sc, long long sr, long long ss, void* scratch,
                           size_t scratch_bytes, int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad, void* stream_) {
    DDL_REQUIRE(conv_ok(N, H, W, C, Cout, R, S, stride, pad), "conv wgrad: unsupported shape (N %d, %dx%d, C %d -> %d, k %d, stride %d, pad %d)",
                N, H, W, C, Cout, R, stride, pad);
    DDL_REQUIRE(x && dy && dw && scratch, "conv wgrad: null pointer");
    const ConvShape s = make_shape(N, H, W, C, Cout, R, S, stride, pad);
    WgradPlan w;
    DDL_REQUIRE(plan_wgrad(s, w), "conv wgrad: no launch plan fits in shared memory");
    DDL_REQUIRE(scratch_bytes >= ddl_conv_wgrad_scratch_bytes(N, H, W, C, Cout, R, S, stride, pad), "conv wgrad: scratch too small");
    WgradParams p = {};
    p.partials = (float*)scratch;
    p.ntaps = w.im2col ? im2col_fwd_taps(s, p.taps) : forward_taps(s, p.taps);
    p.N = w.g.N; p.OHt = w.g.GH; p.OWt = w.g.GW;
    p.bw = w.g.bw; p.bh = w.g.bh; p.nb = w.g.nb;
    p.tiles_w = w.g.tiles_w; p.tiles_h = w.g.tiles_h;
    p.num_chunks = w.g.tiles();
    p.cblocks = C / 64; p.U = w.U; p.G = w.G; p.mgroups = w.mgroups;
    p.Cout = Cout; p.bn = w.bn; p.nblk = w.nblk;
    p.ksplit = w.ksplit; p.cps = w.cps;
    p.slot_bytes = w.slot_bytes; p.ksteps = w.ksteps; p.stages = w.stages; p.pix = w.g.pix();
    p.Mtot = (long long)w.U * 64;
    ConvMaps maps;
    if (w.im2col) {
        p.im2col = 1;
        p.cstride = stride; p.low_w = -pad; p.low_h = -pad;
        p.OHt = s.OH; p.OWt = s.OW;
        {
            const long long xmax = (long long)N * s.OH * s.OW + 256;
            auto magic = [&](int d) -> unsigned long long {
                if (d <= 0 || xmax * (long long)d >= (1ll << 40) || xmax / d >= (1ll << 23)) return 0ull;
                return ((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d;
            };
            p.mul_ow = magic(s.OW);
            p.mul_oh = magic(s.OH);
        }
        const int up_w = (s.OW - 1) * stride + 1 - pad - W, up_h = (s.OH - 1) * stride + 1 - pad - H;
        if (int e = make_im2col_map(&maps.in[0], x, N, H, W, C, -pad, -pad, up_w, up_h, stride, w.g.pix(), "conv wgrad x (im2col)")) return e;
        for (int i = 1; i < 4; ++i) maps.in[i] = maps.in[0];
    } else {
        if (int e = make_act_maps(maps.in, x, N, H, W, C, stride, s.flat, w.g, "conv wgrad x")) return e;
    }
    {
        CUtensorMap tmp[4];
        if (int e = make_act_maps(tmp, dy, N, s.OH, s.OW, Cout, 1, s.flat || w.im2col, w.g, "conv wgrad dy")) return e;
        maps.w = tmp[0];
    }
    const int smem = p.stages * (p.bn / 64 + 2 * p.G) * p.slot_bytes + 512 + 1024;
    static int attr_smem = 0;
    if (smem > attr_smem) {
        DDL_CHECK(cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_smem = smem;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const int grid = p.nblk * p.mgroups * p.ksplit;
    conv_wgrad_kernel<<<grid, kWgThreads, smem, stream>>>(maps, p);
    DDL_LAUNCH_CHECK();
    int lanes_log2 = 0;
    while (lanes_log2 < 3 && (4 << lanes_log2) < p.ksplit) ++lanes_log2;  // about 4 splits per lane before adding lanes
    dim3 fgrid((unsigned)(p.Mtot / (8 >> lanes_log2)), (unsigned)(Cout / 32));
    conv_wgrad_fold_kernel<<<fgrid, 256, 0, stream>>>(p.partials, p.ksplit, p.Mtot, Cout, (bf16*)dw, C, S, sk, sr, ss, sc, lanes_log2);
    DDL_LAUNCH_CHECK();
    return 0;
}
