// 3x3 / stride-1 / pad-1 convolutions (forward and data gradient) on halo tiles.  Included by conv_igemm.cu inside its
// anonymous namespace (it uses that file's ConvMaps, tma_load_4d and epilogue constants).
//
// What the im2col kernel (conv_data_kernel) pays for on these layers, per 128-pixel tile: nine activation tiles (the same
// pixels nine times over through L2 and shared memory), nine weight tiles, one tcgen05.commit per 64-deep k-block (~450
// cycles of tensor-pipe time each, csrc/probe_umma.cu, against 128 - 256 cycles of MMAs) and one trip of the MMA warp through
// its per-k-block loop: the 64-channel layers of ResNet-50's first stage ran at 2.7x the library's time.  Here
//   * the activation tile is loaded ONCE per 64-channel block as a box [64, W + 1, th + 2, 1] that starts at column -1 and
//     row oh0 - 1; TMA zero-fills everything outside the image - that is the padding.  With a row pitch of W + 1 the zero
//     column of image row h + 1 doubles as the right-hand padding of row h.  MMA row m = a * (W + 1) + b, and tap (r, s)
//     is the SAME shared-memory tile read through a descriptor that starts (r * (W + 1) + s) rows later (the 128-byte
//     swizzle is a function of the absolute shared-memory address, so a row-shifted start needs no base offset).  Rows
//     with b = W or a >= th are computed and dropped (87.5 % of the MMA rows are pixels at W = 56 and W = 28);
//   * the weights stay in shared memory for the CTA's whole life when all nine taps fit (64 -> 64 channels: 72 KB), else
//     they move through their own ring and every weight tile is used for `mt` = 2 pixel tiles (two accumulators), which
//     halves the weight traffic per pixel and doubles the MMA work per commit;
//   * commits: one per activation slot (9 k-blocks), one per accumulator, one per weight slot in ring mode.
// Roles as in conv_data_kernel: warp 0 = TMA producer, warp 1 = MMA issue (warp-uniform loop, one elected lane), 16
// epilogue warps (4 TMEM lane quarters x up to 4 (pixel tile, 64-column group) pairs) with the BatchNorm statistics of
// the stored bf16 values as per-CTA partials.
#pragma once

struct HaloParams {
    bf16* out;
    float* partials;  // [m_groups][2][Cn] or null (data gradient)
    int N, H, W, Cn;  // images, output (= input) extent, output channels
    int th, pitch, tiles_h, num_tiles, num_units;  // a unit = `mt` consecutive tiles, one CTA iteration
    int mt;  // = the kernel's MT
    int num_n_blocks, bn, groups;
    int cblocks, w_tap_stride;
    int flip;                // data gradient: shift t = (r', s') multiplies the weights of tap 8 - t
    int a_stages, b_stages;  // b_stages == 0: the 9 * cblocks weight tiles are loaded once and stay
    uint32_t a_slot, a_box_bytes, box_rows;
    int ablate;  // timing experiments only (ddl_conv_set_debug): 1 = no epilogue stores, 2 = no activation loads
};

constexpr int kHaloMaxStages = 8;

template <bool B_MN, int MT, bool STAY>
__global__ void __launch_bounds__(kThreads, 1) conv_halo_kernel(const __grid_constant__ ConvMaps maps, const HaloParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t b_slot = (uint32_t)p.bn * 128u;
    const int nb_slots = p.b_stages ? p.b_stages : 9 * p.cblocks;
    const int n_act = 4 * p.groups * MT;  // epilogue warps with work
    uint8_t* a_ring = smem;
    uint8_t* b_ring = a_ring + (size_t)p.a_stages * MT * p.a_slot;
    uint8_t* stg_base = b_ring + (size_t)nb_slots * b_slot;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(stg_base + (size_t)n_act * kStgBytes);
    uint64_t* a_empty = a_full + kHaloMaxStages;
    uint64_t* b_full = a_empty + kHaloMaxStages;
    uint64_t* b_empty = b_full + kHaloMaxStages;
    uint64_t* tmem_full = b_empty + kHaloMaxStages;  // [2]
    uint64_t* tmem_empty = tmem_full + 2;             // [2]
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 512;
    const int n_blk = blockIdx.x % p.num_n_blocks, m_group = blockIdx.x / p.num_n_blocks;
    const int m_groups = gridDim.x / p.num_n_blocks;

    // Everything behind the box inside a slot is zeroed once and never written again: the row right behind the box is read
    // by the last pixel's (+1, +1) tap and must be zero; the rest only feeds MMA rows that are dropped.
    {
        const uint32_t tail0 = p.box_rows * 128u, tail = p.a_slot - tail0;
        for (int s = 0; s < p.a_stages * MT; ++s) {
            uint8_t* base = a_ring + (size_t)s * p.a_slot + tail0;
            for (uint32_t i = threadIdx.x * 16u; i < tail; i += kThreads * 16u) *reinterpret_cast<uint4*>(base + i) = make_uint4(0, 0, 0, 0);
        }
        fence_proxy_async();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.in[0]);
        tma_prefetch_desc(&maps.w);
        for (int s = 0; s < kHaloMaxStages; ++s) {
            mbar_init(&a_full[s], 1);
            mbar_init(&a_empty[s], 1);
            mbar_init(&b_full[s], 1);
            mbar_init(&b_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], n_act);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================== TMA producer (warp-uniform loop, one elected lane issues) =====================
        auto load_b = [&](uint8_t* sb, uint64_t* bar, int tap, int cb) {
            const int wt = p.flip ? 8 - tap : tap;
            if (!B_MN) {
                tma_load_2d(sb, &maps.w, bar, wt * p.w_tap_stride + cb * 64, n_blk * p.bn);
            } else {
                for (int j = 0; j < p.groups; ++j) tma_load_2d(sb + j * 8192, &maps.w, bar, wt * p.w_tap_stride + n_blk * p.bn + j * 64, cb * 64);
            }
        };
        if (STAY) {
            if (elect_one()) {
                mbar_expect_tx(&b_full[0], (uint32_t)nb_slots * b_slot);
                for (int cb = 0; cb < p.cblocks; ++cb)
                    for (int tap = 0; tap < 9; ++tap) load_b(b_ring + (size_t)(cb * 9 + tap) * b_slot, &b_full[0], tap, cb);
            }
            __syncwarp();
        }
        int as = 0, bs = 0;
        uint32_t aph = 0, bph = 0;
        const bool no_loads = (p.ablate & 2) != 0;
        for (int unit = m_group; unit < p.num_units; unit += m_groups) {
            // a tile past the end decodes to an image index >= N: zero-filled, nothing is stored for it
            int tn[MT], th0[MT];
#pragma unroll
            for (int t = 0; t < MT; ++t) {
                const int tile = unit * MT + t;
                tn[t] = tile / p.tiles_h;
                th0[t] = (tile - tn[t] * p.tiles_h) * p.th - 1;
            }
            for (int cb = 0; cb < p.cblocks; ++cb) {
                mbar_wait(&a_empty[as], aph ^ 1);
                if (elect_one()) {
                    if (no_loads) {
                        mbar_arrive(&a_full[as]);
                    } else {
                        mbar_expect_tx(&a_full[as], (uint32_t)MT * p.a_box_bytes);
#pragma unroll
                        for (int t = 0; t < MT; ++t)
                            tma_load_4d(a_ring + (size_t)(as * MT + t) * p.a_slot, &maps.in[0], &a_full[as], cb * 64, -1, th0[t], tn[t]);
                    }
                }
                __syncwarp();
                if (++as == p.a_stages) { as = 0; aph ^= 1; }
                if (!STAY) {
                    for (int tap = 0; tap < 9; ++tap) {
                        mbar_wait(&b_empty[bs], bph ^ 1);
                        if (elect_one()) {
                            mbar_expect_tx(&b_full[bs], b_slot);
                            load_b(b_ring + (size_t)bs * b_slot, &b_full[bs], tap, cb);
                        }
                        __syncwarp();
                        if (++bs == p.b_stages) { bs = 0; bph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue: the whole warp runs the loop, one elected lane issues =====================
        // This warp's own instruction stream is what bounds the small-N layers (one warp retires a dependent instruction
        // every few cycles; 4 MMAs of 128 x 64 x 16 are 192 cycles of tensor-pipe time): the nine taps are straight-line
        // code, a descriptor is (constant | address >> 4) and every step only adds to it - +2 per 16-deep k-step, +8 per
        // tile row (one tap column), +8 * pitch per tap row, + slot size per tile / weight tile.
        const uint32_t idesc = make_idesc(kTileM, p.bn, false, B_MN);
        constexpr uint32_t kStepA = 32 >> 4, kStepB = (B_MN ? 2048 : 32) >> 4;
        const uint32_t row16 = (uint32_t)p.pitch * 8u, a_slot16 = p.a_slot >> 4, b_slot16 = b_slot >> 4;
        const uint64_t da0 = make_smem_desc(0, 16, 1024) | (uint64_t)(smem_u32(a_ring) >> 4);
        const uint64_t db0 = (B_MN ? make_smem_desc(0, 8192, 1024) : make_smem_desc(0, 16, 1024)) | (uint64_t)(smem_u32(b_ring) >> 4);
        const uint32_t bn = (uint32_t)p.bn;
        const int a_stages = p.a_stages, b_stages = p.b_stages, cblocks = p.cblocks;
        int as = 0, bs = 0, acc = 0;
        uint32_t aph = 0, bph = 0, acc_phase = 0;
        if (STAY) {
            mbar_wait(&b_full[0], 0);
            tc_fence_after();
        }
        for (int unit = m_group; unit < p.num_units; unit += m_groups) {
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d0 = tmem_base + (uint32_t)acc * (MT * bn);
            for (int cb = 0; cb < cblocks; ++cb) {
                mbar_wait(&a_full[as], aph);
                tc_fence_after();
                const uint64_t da_s = da0 + (uint32_t)(as * MT) * a_slot16;
                if (STAY) {
                    const uint64_t db_c = db0 + (uint32_t)(cb * 9) * b_slot16;
                    if (elect_one()) {
#pragma unroll
                        for (int j = 0; j < 9; ++j) {
#pragma unroll
                            for (int t = 0; t < MT; ++t) {
                                const uint64_t da = da_s + (uint32_t)t * a_slot16 + (uint32_t)(j / 3) * row16 + (uint32_t)(j % 3) * 8u;
                                const uint64_t db = db_c + (uint32_t)j * b_slot16;
                                const uint32_t d = d0 + (uint32_t)t * bn;
                                umma_bf16(d, da, db, idesc, j == 0 ? (cb != 0 ? 1u : 0u) : 1u);
#pragma unroll
                                for (int k = 1; k < 4; ++k) umma_bf16(d, da + k * kStepA, db + k * kStepB, idesc, 1u);
                            }
                        }
                        umma_commit(&a_empty[as]);
                    }
                    __syncwarp();
                } else {
#pragma unroll
                    for (int j = 0; j < 9; ++j) {
                        mbar_wait(&b_full[bs], bph);
                        tc_fence_after();
                        const uint64_t db = db0 + (uint32_t)bs * b_slot16;
                        if (elect_one()) {
#pragma unroll
                            for (int t = 0; t < MT; ++t) {
                                const uint64_t da = da_s + (uint32_t)t * a_slot16 + (uint32_t)(j / 3) * row16 + (uint32_t)(j % 3) * 8u;
                                const uint32_t d = d0 + (uint32_t)t * bn;
                                umma_bf16(d, da, db, idesc, j == 0 ? (cb != 0 ? 1u : 0u) : 1u);
#pragma unroll
                                for (int k = 1; k < 4; ++k) umma_bf16(d, da + k * kStepA, db + k * kStepB, idesc, 1u);
                            }
                            umma_commit(&b_empty[bs]);
                            if (j == 8) umma_commit(&a_empty[as]);
                        }
                        __syncwarp();
                        if (++bs == b_stages) { bs = 0; bph ^= 1; }
                    }
                }
                if (++as == a_stages) { as = 0; aph ^= 1; }
            }
            if (elect_one()) umma_commit(&tmem_full[acc]);
            __syncwarp();
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;
        const int quarter = warp & 3;  // TMEM lanes 32 * quarter .. + 31
        const int vg = ew >> 2;        // (pixel tile of the unit, 64-column group)
        if (vg < p.groups * MT) {
            const int ts = vg / p.groups, group = vg - ts * p.groups;
            uint8_t* stg = stg_base + ew * kStgBytes;
            const bool stats = p.partials != nullptr;
            float cs[16], cq[16];  // columns col0 + h * 32 + t_ch * 8 + q  at index h * 8 + q
#pragma unroll
            for (int i = 0; i < 16; ++i) cs[i] = cq[i] = 0.f;
            const int col0 = n_blk * p.bn + group * 64;
            const int t_row = lane >> 2, t_ch = lane & 3;
            // MMA row -> pixel of the tile: the same for every tile
            int row_a[4], rel[4];
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int r = quarter * 32 + it * 8 + t_row;
                const int a = r / p.pitch, b = r - a * p.pitch;
                row_a[it] = (a < p.th && b < p.W) ? a : (1 << 20);
                rel[it] = (a * p.W + b) * p.Cn + col0 + t_ch * 8;
            }
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int unit = m_group; unit < p.num_units; unit += m_groups) {
                const int tile = unit * MT + ts;
                const int n = tile / p.tiles_h, oh0 = (tile - n * p.tiles_h) * p.th;
                const long long base = ((long long)n * p.H + oh0) * p.W * p.Cn;
                long long off[4];
#pragma unroll
                for (int it = 0; it < 4; ++it) off[it] = (tile < p.num_tiles && oh0 + row_a[it] < p.H) ? base + rel[it] : -1;
                mbar_wait(&tmem_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (uint32_t)((acc * MT + ts) * p.bn + group * 64) + ((uint32_t)(quarter * 32) << 16);
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
            if (stats) {
                // fold the 8 row groups of the warp (lanes with equal t_ch), then - in fixed order - the warps that hold
                // the same 64 columns (4 lane quarters x mt pixel tiles) -> this CTA's partial statistics
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
                asm volatile("bar.sync 1, %0;" ::"r"(n_act * 32) : "memory");  // all epilogue warps with work
                if (quarter == 2 && ts == 0) {
                    float t4[4] = {0.f, 0.f, 0.f, 0.f};
                    for (int tt = 0; tt < MT; ++tt) {
#pragma unroll
                        for (int qw = 0; qw < 4; ++qw) {
                            const float* src = reinterpret_cast<const float*>(stg_base + (size_t)(((tt * p.groups + group) * 4 + qw)) * kStgBytes);
#pragma unroll
                            for (int e = 0; e < 4; ++e) t4[e] += src[lane * 4 + e];
                        }
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
