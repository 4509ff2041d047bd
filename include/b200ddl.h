/* libb200ddl - C ABI of the hand-written sm_100a kernels and the native runtime.
 *
 * Every entry point returns 0 on success or a non-zero cudaError_t value; the message is
 * available from ddl_last_error() (thread local).  Pointers are device pointers unless the
 * name says otherwise; `stream` is a cudaStream_t passed as void*.  Nothing here allocates
 * device memory except ddl_symm_alloc: scratch buffers are passed in by the caller, which
 * keeps every kernel launch capturable in a CUDA graph.
 *
 * "reference" in the comments below = /root/reference/notebooks/cv/onnx_experiments.py, the
 * only code file of the upstream repository.  It contains no native
 * code and no distributed code, so most entry points cite BASELINE.json instead.
 */
#ifndef B200DDL_H
#define B200DDL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { DDL_F32 = 0, DDL_BF16 = 1 };
enum { DDL_ALGO_AUTO = 0, DDL_ALGO_ONESHOT = 1, DDL_ALGO_TWOSHOT = 2, DDL_ALGO_NVLS = 3 };
enum { DDL_ACT_NONE = 0, DDL_ACT_GELU = 1, DDL_ACT_RELU = 2 };
enum { DDL_EPI_NONE = 0, DDL_EPI_GELU = 1, DDL_EPI_DGELU = 2, DDL_EPI_RELU = 3 };

#define DDL_COMM_MAX_BLOCKS 320

/* ---- runtime (csrc/runtime.cu) ------------------------------------------------------ */
const char* ddl_last_error(void);
int ddl_version(void);
int ddl_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem);
/* symmetric heap slab: cudaMalloc + zero fill; handle export/import over CUDA IPC
 * (target layer T1; BASELINE.json:5 "load and store directly from peer GPUs") */
int ddl_symm_alloc(size_t bytes, void** ptr);
int ddl_symm_free(void* ptr);
int ddl_ipc_export(void* ptr, unsigned char* handle64);
int ddl_ipc_open(const unsigned char* handle64, void** ptr);
int ddl_ipc_close(void* ptr);
int ddl_can_access_peer(int device, int peer, int* ok);
int ddl_stream_sync(void* stream);

/* ---- peer-memory collectives (csrc/collectives.cu) ---------------------------------- */
typedef struct {
    void* signal_pads[8]; /* peer-mapped address of every rank's signal pad (one channel) */
    void* epoch;          /* local device memory, ddl_comm_epoch_bytes(), zero-initialised */
    int rank;
    int world;
} ddl_comm_t;

size_t ddl_comm_signal_pad_bytes(void);
size_t ddl_comm_epoch_bytes(void);
size_t ddl_allreduce_oneshot_max_bytes(int dtype, int num_blocks);
/* All-reduce (mean when scale = 1/world) fused with the output cast.  in_ptrs[r] / out_ptrs[r]
 * are this message's address inside rank r's heap as mapped into THIS process.
 * K1 one-shot, K2 two-shot, K3 NVLS multimem (needs multicast addresses mc_in/mc_out).
 * BASELINE.json:5; the reference has no collective. */
int ddl_allreduce(const ddl_comm_t* comm, void* const* in_ptrs, void* const* out_ptrs, size_t numel, int in_dtype, int out_dtype,
                  float scale, int algo, void* mc_in, void* mc_out, int num_blocks, void* stream);
/* Hybrid all-reduce (bf16, in place): nvls_permille / 1000 of the message is reduced by the switch (multimem, nvls_blocks
 * CTAs), the rest by two-shot peer loads / stores (p2p_blocks CTAs), in ONE kernel - the two paths use different link
 * resources and overlap.  buf_ptrs[r] = the buffer in rank r's heap, mc_buf its multicast address.  BASELINE.json:5
 * (">= 80 % of the NVLink roofline for buckets >= 16 MB"); no counterpart in the reference. */
int ddl_allreduce_hybrid(const ddl_comm_t* comm, void* const* buf_ptrs, size_t numel, float scale, void* mc_buf, int nvls_permille,
                         int nvls_blocks, int p2p_blocks, void* stream);
int ddl_allgather(const ddl_comm_t* comm, void* const* shard_ptrs, void* out_full, size_t shard_bytes, int num_blocks, void* stream);
int ddl_reducescatter(const ddl_comm_t* comm, void* const* full_ptrs, void* out_shard, size_t shard_numel, int in_dtype,
                      int out_dtype, float scale, int num_blocks, void* stream);
/* K7: reduce-scatter fused with the sharded AdamW update (BASELINE.json:5, FSDP mode).
 * hyper = device float[8]: {lr, beta1, beta2, eps, weight_decay, step, -, -}. */
int ddl_reducescatter_adamw(const ddl_comm_t* comm, void* const* full_ptrs, size_t shard_numel, int in_dtype, float scale,
                            float* master, float* m, float* v, void* param_out_bf16, const float* hyper, int num_blocks,
                            void* stream);
int ddl_broadcast(const ddl_comm_t* comm, void* const* buf_ptrs, int root, size_t bytes, int num_blocks, void* stream);
int ddl_barrier(const ddl_comm_t* comm, void* stream);

/* ---- fused AdamW (csrc/optim.cu), K13 ---------------------------------------------- */
int ddl_adamw(const void* grad, int grad_dtype, float* master, float* m, float* v, void* param_out, int param_dtype, size_t numel,
              const float* hyper, float grad_scale, void* stream);
int ddl_adamw_tick(float* hyper, void* stream); /* hyper[5] += 1 (step counter), graph friendly */

/* ---- LayerNorm family (csrc/layernorm.cu), K10 ------------------------------------- */
float ddl_dropout_effective_p(float p); /* p rounded to a multiple of 1/256 */
/* z = dropout(x) + residual ; y = LN(z)*gamma + beta.  residual, z_out may be NULL. */
int ddl_add_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* z_out, void* y,
                          float* mean, float* rstd, int rows, int H, float eps, float dropout_p, unsigned long long seed,
                          const unsigned long long* seed_dev, void* stream);
size_t ddl_layernorm_bwd_scratch_bytes(int rows, int H);
/* dz (grad of z) always; dx = dropout-masked dz only when dropout_p > 0 and dx != NULL.
 * dgamma/dbeta in param_grad_dtype.  scratch: ddl_layernorm_bwd_scratch_bytes(rows, H). */
int ddl_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const void* gamma, void* dz, void* dx,
                      void* dgamma, void* dbeta, int param_grad_dtype, void* scratch, size_t scratch_bytes, int rows, int H,
                      float dropout_p, unsigned long long seed, const unsigned long long* seed_dev, void* stream);
/* as ddl_layernorm_bwd, plus dxsum [H] bf16 = column sums of the gradient of the first input (dx when dropout is active, else
 * dz): the bias gradient of the linear layer that produced that input.  scratch: ddl_layernorm_bwd_colsum_scratch_bytes. */
size_t ddl_layernorm_bwd_colsum_scratch_bytes(int rows, int H);
int ddl_layernorm_bwd_colsum(const void* dy, const void* z, const float* mean, const float* rstd, const void* gamma, void* dz, void* dx,
                             void* dgamma, void* dbeta, void* dxsum, int param_grad_dtype, void* scratch, size_t scratch_bytes, int rows,
                             int H, float dropout_p, unsigned long long seed, const unsigned long long* seed_dev, void* stream);
int ddl_embed_layernorm_fwd(const long long* ids, const long long* token_types, const void* word, const void* pos,
                            const void* type, const void* gamma, const void* beta, void* z_out, void* y, float* mean, float* rstd,
                            int rows, int S, int H, float eps, float dropout_p, unsigned long long seed,
                            const unsigned long long* seed_dev, void* stream);
/* Gradients of the BERT embedding tables from dz [B*S, H] bf16 (csrc/embed_bwd.cu): dword_f32 [V, H] fp32 is accumulated
 * into (zero it first; convert with ddl_f32_to_bf16), dpos [P, H] and dtype [T, H] bf16 are overwritten (T <= 4). */
size_t ddl_embed_bwd_scratch_bytes(int H, int T);
int ddl_embed_bwd(const void* dz, const long long* ids, const long long* token_types, float* dword_f32, void* dpos, void* dtype,
                  void* scratch, size_t scratch_bytes, int B, int S, int H, int V, int P, int T, void* stream);
int ddl_dropout(const void* in, void* out, size_t numel, float dropout_p, unsigned long long seed,
                const unsigned long long* seed_dev, void* stream);

/* ---- cross-entropy (csrc/cross_entropy.cu), K12 ------------------------------------ */
/* row_loss[rows] fp32; mean_loss (optional) = mean(row_loss); dlogits (optional) =
 * (softmax - onehot) * grad_scale in the logits' dtype.  The reference only computes a
 * display softmax (reference:85-88, :174). */
int ddl_cross_entropy(const void* logits, int dtype, const long long* labels, float* row_loss, float* mean_loss, void* dlogits,
                      int rows, int classes, float grad_scale, long long ignore_index, void* stream);

/* ---- BatchNorm + residual + ReLU, NHWC bf16 (csrc/bn_act.cu), K9 epilogue half ------ */
size_t ddl_bn_scratch_bytes(long long M, int C);
int ddl_bn_apply(const void* x, const void* residual, const float* scale, const float* shift, void* y, long long M, int C, int relu,
                 void* stream);
int ddl_bn_act_fwd_train(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                         float* running_var, void* y, float* mean, float* invstd, float* scale, float* shift, void* scratch,
                         size_t scratch_bytes, long long M, int C, float eps, float momentum, int relu, void* stream);
/* same, also writing relu_mask (nullable; M*C/8 bytes, bit j of byte v = channel j of 8-channel vector v passed the
 * ReLU): backward then reads 1 bit instead of 16 per element to rebuild the mask (ddl_bn_act_bwd_ex). */
int ddl_bn_act_fwd_train_mask(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                              float* running_var, void* y, void* relu_mask, float* mean, float* invstd, float* scale, float* shift,
                              void* scratch, size_t scratch_bytes, long long M, int C, float eps, float momentum, int relu,
                              void* stream);
int ddl_bn_act_fwd_eval(const void* x, const void* residual, const void* gamma, const void* beta, const float* running_mean,
                        const float* running_var, void* y, float* scale, float* shift, long long M, int C, float eps, int relu,
                        void* stream);
/* Training-mode BN + residual + ReLU when the per-channel sum / sum of squares already exist (sums = fp32 [2][C],
 * e.g. from ddl_gemm_bf16_ex): skips the statistics pass over x. */
int ddl_bn_act_fwd_train_from_sums(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                   float* running_var, void* y, float* mean, float* invstd, float* scale, float* shift,
                                   const float* sums, long long M, int C, float eps, float momentum, int relu, void* stream);
/* same with per-block partial statistics [nblocks][2][C] (written by ddl_stem_conv_fwd) and the optional relu_mask */
int ddl_bn_act_fwd_train_from_partials(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                       float* running_var, void* y, void* relu_mask, float* mean, float* invstd, float* scale,
                                       float* shift, const float* partials, int nblocks, long long M, int C, float eps,
                                       float momentum, int relu, void* stream);
/* Allocation-free variant of ddl_bn_act_fwd_train: same arguments minus the scratch buffer.
 * No counterpart in the reference (it has no native code at all).  It exists for callers
 * that cannot allocate device memory - a process whose own arena already owns the whole
 * GPU, or a latency-critical loop that must not hit cudaMalloc's device-wide sync.  The
 * scratch is a 2 MiB array inside the library image (held by every GPU that loads the
 * library until exit); shapes needing more (ddl_bn_scratch_bytes(M, C) >
 * DDL_BN_NOALLOC_MAX_BYTES; every M fits when C <= 512, wider C only for small M -
 * ask ddl_bn_scratch_bytes) are rejected.  The first call on a device looks up the array's
 * address: make that call with the device current and before any stream capture.  All
 * calls share the array; calls that overlap on the GPU race - serialise them.  Nothing in
 * it needs zeroing between calls.  The looked-up address is cached per device for the life
 * of the process: a host program that calls cudaDeviceReset() invalidates it (and every
 * other handle of this library) and must not call the entry point on that device again. */
#define DDL_BN_NOALLOC_MAX_BYTES (2u << 20)
int ddl_bn_act_fwd_train_noalloc(const void* x, const void* residual, const void* gamma, const void* beta, float* running_mean,
                                 float* running_var, void* y, float* mean, float* invstd, float* scale, float* shift, long long M,
                                 int C, float eps, float momentum, int relu, void* stream);
/* dres (optional) receives the gradient of the residual branch (= ReLU-masked dy). */
int ddl_bn_act_bwd(const void* dy, const void* y, const void* x, const void* gamma, const float* mean, const float* invstd, void* dx,
                   void* dres, void* dgamma, void* dbeta, int param_grad_dtype, void* scratch, size_t scratch_bytes, long long M,
                   int C, int relu, void* stream);
/* Extended backward: the ReLU mask comes from relu_mask (ddl_bn_act_fwd_train_mask) when given, else from y; dy2
 * (nullable, relu + residual layers only) is a second incoming gradient of the same tensor - the other branch of the
 * residual fork of the next block (reference model: torchvision ResNet-50, reference:19) - added inside the
 * reduction pass instead of by a separate elementwise kernel. */
int ddl_bn_act_bwd_ex(const void* dy, const void* dy2, const void* y, const void* relu_mask, const void* x, const void* gamma,
                      const float* mean, const float* invstd, void* dx, void* dres, void* dgamma, void* dbeta,
                      int param_grad_dtype, void* scratch, size_t scratch_bytes, long long M, int C, int relu, void* stream);
/* Stem: BatchNorm (batch statistics from per-CTA partials, as ddl_bn_act_fwd_train_from_partials) + ReLU + 3x3/2 max pool in
 * ONE pass over x [N, H, W, C] (even H, W; 256 % (C / 8) == 0): y [N, H/2, W/2, C]; codes = N * H/2 * W/2 * C/2 bytes (4 bits per
 * pooled element: winning tap, or 15 where the pooled value is 0).  Backward: dx, dgamma, dbeta from dy (+ optional dy2), y, the
 * codes and x; scratch from ddl_bn_relu_maxpool_bwd_scratch_bytes. */
int ddl_bn_relu_maxpool_supported(int N, int H, int W, int C);
size_t ddl_bn_relu_maxpool_bwd_scratch_bytes(int N, int H, int W, int C);
int ddl_bn_relu_maxpool_fwd_train(const void* x, const void* gamma, const void* beta, float* running_mean, float* running_var, void* y,
                                  void* codes, float* mean, float* invstd, float* scale, float* shift, const float* partials, int nblocks,
                                  int N, int H, int W, int C, float eps, float momentum, void* stream);
int ddl_bn_relu_maxpool_bwd(const void* dy, const void* dy2, const void* y, const void* codes, const void* x, const void* gamma,
                            const void* beta, const float* mean, const float* invstd, void* dx, void* dgamma, void* dbeta,
                            int param_grad_dtype, void* scratch, size_t scratch_bytes, int N, int H, int W, int C, void* stream);

/* ---- self-attention on tcgen05 (csrc/attention.cu), K11 ---------------------------------- */
/* qkv: packed projections [B*S, 3*heads*64] bf16 (q | k | v); seqlens[B] = valid (right-padded)
 * lengths or NULL; ctx [B*S, heads*64]; lse [B, heads, S] fp32 saved for backward.  head_dim is 64,
 * S a multiple of 128 up to 512 (the BASELINE.json:9-10 configs: 128 and 512).  Dropout acts on the
 * attention probabilities; its mask is regenerated in backward from (seed, *seed_dev).
 * Backward writes the packed gradient dqkv [B*S, 3*heads*64]; scratch from
 * ddl_attention_bwd_scratch_bytes.  The reference has no attention ("Long-context"). */
int ddl_attention_fwd(const void* qkv, const int* seqlens, void* ctx, float* lse, int B, int S, int heads, float dropout_p,
                      unsigned long long seed, const unsigned long long* seed_dev, void* stream);
size_t ddl_attention_bwd_scratch_bytes(int B, int S, int heads);
int ddl_attention_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse, const int* seqlens, void* dqkv,
                      void* scratch, size_t scratch_bytes, int B, int S, int heads, float dropout_p, unsigned long long seed,
                      const unsigned long long* seed_dev, void* stream);

/* ---- pooling on NHWC bf16 (csrc/pool.cu), K14 -------------------------------------------- */
/* 3x3 / stride 2 / pad 1 max pool (ResNet stem; model at reference:19).  idx: one byte per output
 * element (winning tap 0..8), consumed by the gather-style backward. */
int ddl_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream);
int ddl_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, void* stream);
/* same with a second incoming gradient dy2 (nullable) summed on the fly: the two branches of a residual fork */
int ddl_maxpool3x3s2_bwd_ex(const void* dy, const void* dy2, const void* idx, void* dx, int N, int H, int W, int C, void* stream);
/* global average pool [N, HW, C] -> [N, C] and its backward */
int ddl_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int ddl_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
/* network input [N, C, HW] (x_dtype DDL_F32 / DDL_BF16, C <= 8) -> [N, HW, 8] bf16, channels C..7 zero: the stem
 * convolution of the reference's ResNet-50 (reference:19) then has 8 input channels and runs on tensor cores */
int ddl_nchw_to_nhwc8_bf16(const void* x, int x_dtype, void* y, int N, int C, long long HW, void* stream);

/* ---- ResNet stem convolution on tcgen05 (csrc/stem_conv.cu), K9 ---------------------------
 * 7x7 / stride 2 / pad 3, 3 -> 64 channels (torchvision resnet50 conv1, reference:19) as a 4x4 convolution over the
 * space-to-depth input z [N, H/2+3, W/2+3, 16] bf16; the im2col tiles are fetched by TMA through a tensor map with
 * overlapping rows.  H, W even.  w / dw: bf16 [64, 3, 7, 7] addressed with element strides (sk, sc, sr, ss), so both
 * contiguous and channels-last weights work.  w2: 32 KB scratch for the packed weights.  y: [N, H/2, W/2, 64] bf16 NHWC.
 * partials: fp32 [ddl_stem_fwd_grid][2][64] per-CTA BatchNorm statistics of y (feed ddl_bn_act_fwd_train_from_partials). */
size_t ddl_stem_z_bytes(int N, int H, int W);
int ddl_stem_fwd_grid(int N, int H, int W);
int ddl_stem_wgrad_grid(int N, int H, int W);
size_t ddl_stem_wgrad_scratch_bytes(int N, int H, int W);
int ddl_stem_s2d(const void* x, int x_dtype, void* z, int N, int H, int W, void* stream);
int ddl_stem_conv_fwd(const void* z, const void* w, long long sk, long long sc, long long sr, long long ss, void* w2, void* y,
                      float* partials, int N, int H, int W, void* stream);
/* weight gradient from z and dy [N, H/2, W/2, 64]; W/2 must be a multiple of 16 (the reduction runs over whole output rows) */
int ddl_stem_conv_wgrad(const void* z, const void* dy, void* dw, long long sk, long long sc, long long sr, long long ss,
                        void* scratch, size_t scratch_bytes, int N, int H, int W, void* stream);

/* ---- 1x1 / stride-1 convolution forward, weight-stationary (csrc/conv1x1.cu), K9 --------
 * y[M, N] = x[M, K] . w[N, K]^T on NHWC bf16 (M = batch*H*W; N, K in {64,128,256,512}, N*K <= 65536: the layer1 / layer2
 * bottleneck convolutions of the reference's ResNet-50, reference:19) with the BatchNorm statistics of y written as
 * per-CTA partials fp32 [ddl_conv1x1_grid(M, N)][2][N] for ddl_bn_act_fwd_train_from_partials. */
int ddl_conv1x1_supported(long long M, int N, int K);
int ddl_conv1x1_grid(long long M, int N);
int ddl_conv1x1_fwd(const void* x, const void* w, void* y, float* partials, long long M, int N, int K, void* stream);
/* data gradient of the same layers, same kernel with w [Cout][Cin] read MN-major: dx [M, Cin] = dy [M, Cout] . w
 * (supported when ddl_conv1x1_supported(M, Cin, Cout)) */
int ddl_conv1x1_dgrad(const void* dy, const void* w, void* dx, long long M, int Cin, int Cout, void* stream);

/* ---- 1x1 / 3x3 convolutions as implicit GEMMs on tcgen05 (csrc/conv_igemm.cu), K9 -----------------------------
 * NHWC bf16 activations, channels-last weights [Cout][R][S][C]; R = S in {1 (pad 0), 3 (pad 1)}, stride 1 or 2 (even H, W),
 * C and Cout multiples of 64 (<= 256 or multiples of 256): every convolution of the reference's ResNet-50 (reference:19,
 * SURVEY.md 2.E.2) except the 3-channel stem (csrc/stem_conv.cu).  The operand tiles are TMA boxes of the 4-D activation
 * tensor shifted by the filter tap (zero fill = padding); nothing is unfolded in memory.
 *   ddl_conv_fwd   : y [N, OH, OW, Cout]; partials (nullable) = fp32 [ddl_conv_fwd_groups(...)][2][Cout] per-CTA BatchNorm
 *                    statistics of y for ddl_bn_act_fwd_train_from_partials.
 *   ddl_conv_dgrad : dx [N, H, W, C] from dy [N, OH, OW, Cout] (every element of dx is written).
 *   ddl_conv_wgrad : dw addressed with element strides (sk, sc, sr, ss) of a [Cout, C, R, S] tensor; scratch of
 *                    ddl_conv_wgrad_scratch_bytes(...) holds the fp32 partials of the pixel splits (fixed-order fold).
 * The reference only ever runs these layers forward through library kernels (reference:32, :174). */
int ddl_conv_supported(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad);
int ddl_conv_fwd_groups(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad);
int ddl_conv_fwd(const void* x, const void* w, void* y, float* partials, int N, int H, int W, int C, int Cout, int R, int S, int stride,
                 int pad, void* stream);
int ddl_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad,
                   void* stream);
size_t ddl_conv_wgrad_scratch_bytes(int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad);
int ddl_conv_wgrad(const void* x, const void* dy, void* dw, long long sk, long long sc, long long sr, long long ss, void* scratch,
                   size_t scratch_bytes, int N, int H, int W, int C, int Cout, int R, int S, int stride, int pad, void* stream);
/* benchmarks: pixel-chunk size of the weight gradient (0 = auto, 64, 128); forward / data-gradient kernel for layers with
 * >= 128 output channels: 0 = one CTA per tile, 1 = CTA pairs (cta_group::2) with halo tiles for 3x3 / stride 1 (default),
 * 2 = CTA pairs without halo tiles */
int ddl_conv_set_tuning(int wgrad_rows, int pair_mode);
/* activation tiles through im2col-mode TMA (1, default: dense 128-pixel tiles) or through tiled boxes (0; comparison runs) */
int ddl_conv_set_im2col(int on);
/* 3x3 / stride-1 / pad-1 layers (forward and data gradient) on halo tiles (csrc/conv_halo.cuh): on = 1 (default) / 0; mt =
 * pixel tiles per weight tile (0 = auto, 1, 2); min_w = narrowest image that takes this kernel (0 = default 28) */
int ddl_conv_set_halo(int on, int mt, int min_w);
/* weight panel of a CTA's column block resident in shared memory where it fits (1, default) or streamed with every tile (0) */
int ddl_conv_set_stay(int on);
/* development aid: device buffer [296][8] int64 that following forward / data-gradient launches of the CTA-pair kernel fill
 * with per-CTA wait-cycle counters (which barrier each role waits on); NULL switches it off.  ablate (timing experiments
 * only, results are garbage): bit 0 = epilogue drops its staging and stores, bit 1 = no TMA loads are issued */
int ddl_conv_set_debug(void* buf, int ablate);

/* ---- hardware probes (csrc/probe_umma.cu; benchmarks/probe_umma*.py, probe_tmem_drain.py; not on any training path) ---- */
int ddl_probe_umma_shift(const void* X, const void* B, float* D, int shift, int variant, int mn, void* stream);
int ddl_probe_umma_rate(long long* out, int n, int iters, int grid, int mode, int nacc, void* stream);
/* cycles for `iters` tcgen05.ld (32 lanes x 32 columns) per warp with `nwarps` reader warps, with or without a concurrent stream
 * of 128 x 256 x 16 MMAs on other TMEM columns; out: int64 [grid][2] = {cycles, MMAs issued meanwhile} */
int ddl_probe_tmem_drain(long long* out, int nwarps, int iters, int mma_on, int grid, void* stream);

/* ---- tcgen05 / TMEM / TMA GEMM (csrc/gemm_tcgen05.cu), K8 ------------------------------ */
/* C[M,N] = epilogue(A . B^T), bf16 operands, fp32 accumulation in tensor memory.
 * A is [M,K] row-major (a_mn = 0) or stored as [K,M] (a_mn = 1); B is [N,K] (b_mn = 0) or stored
 * as [K,N] (b_mn = 1); lda/ldb/ldc are row pitches in elements (multiples of 8).
 * Epilogue, in order: + bias[N]; DDL_EPI_GELU (exact GELU; pre-activation written to aux when
 * given) | DDL_EPI_DGELU (multiply by GELU'(aux)) | DDL_EPI_RELU; + residual[M,ldc]; bf16 store.
 * split_k > 1: raw fp32 partial sums are red.add-ed into acc_f32[M,N] (caller zeroes it), no
 * epilogue; finish with ddl_f32_to_bf16.  block_n: 0 = auto, 128 or 256.
 * BASELINE.json:5 ("tcgen05/TMEM GEMM ... TMA-fed tiles"); the reference has no GEMM of its own
 * (its only matrix product is the library-run fc layer of ResNet-50, reference:19). */
int ddl_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb, void* C, long long ldc, int M, int N,
                  int K, const void* bias, const void* residual, void* aux, int epilogue, int split_k, float* acc_f32, int block_n,
                  void* stream);
/* Same, plus col_stats: optional fp32 [2][N] (caller-zeroed) that receives the per-column sum and sum of squares
 * of the stored output - the BatchNorm statistics of a 1x1 convolution for free (K9, conv + BN epilogue). */
int ddl_gemm_bf16_ex(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb, void* C, long long ldc, int M, int N,
                     int K, const void* bias, const void* residual, void* aux, int epilogue, int split_k, float* acc_f32,
                     float* col_stats, int block_n, void* stream);
/* K6 - all-gather fused into the consuming GEMM (BASELINE.json:5, FSDP): C = epilogue(A . W^T) where W [N, K] is
 * sharded by rows over the ranks (shard_ptrs[r] = rank r's [N/world, K] shard as mapped into this process).  One
 * kernel per rank: copier warps pull the shards over NVLink into full_w [N, K] (local scratch) and publish
 * per-shard ready counters; the TMA producer consumes the column blocks shard by shard, local shard first, so math
 * on arrived shards overlaps the transfer of the rest.  flags: local uint32[9], zero before the first call.
 * N must be a multiple of world * 128.  No counterpart in the reference. */
int ddl_allgather_gemm_bf16(const ddl_comm_t* comm, void* const* shard_ptrs, void* full_w, void* flags, const void* A, long long lda,
                            void* C, long long ldc, int M, int N, int K, const void* bias, const void* residual, void* aux, int epilogue,
                            void* stream);
/* launch mode: 0 = auto (CTA pairs, tcgen05.mma.cta_group::2, 256x256 tiles, when eligible), 1 = one CTA per
 * tile, 2 = 2-CTA cluster sharing B by TMA multicast, 3 = same as 0 */
int ddl_gemm_set_cluster(int cluster);
/* flags: bit 0 = add to the existing bf16 output, bit 1 = write zeros back to `in` after reading it (a cached split-K
 * accumulator is then ready for the next GEMM without a memset) */
int ddl_f32_to_bf16(const float* in, void* out, size_t numel, int flags, void* stream);
/* out = dy * GELU'(pre) elementwise on bf16 (numel a multiple of 8): the stand-alone form of the DDL_EPI_DGELU epilogue */
int ddl_dgelu_mul(const void* dy, const void* pre, void* out, size_t numel, void* stream);
size_t ddl_colsum_scratch_bytes(int rows, int cols);
/* out[c] = sum_r x[r,c] (bias gradients), deterministic two-stage fold */
int ddl_colsum_bf16(const void* x, int rows, int cols, long long ld, void* out, void* scratch, size_t scratch_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DDL_H */
