/* pdt_b200.h — C ABI of libpdt_b200.so (hand-written sm_100a kernels).
 *
 * Conventions
 *   - every function returns 0 on success, else a cudaError_t value (or a
 *     negative PDT_E* code), and only ENQUEUES work on `stream`;
 *   - pointers are device pointers unless stated otherwise;
 *   - activations are NHWC ("channels_last"), dtype codes PDT_BF16/FP16/FP32;
 *   - nothing here allocates device memory: callers own every buffer
 *     (torch tensors on the Python side).
 * Each declaration cites what it replaces in the reference call graph
 * (/root/reference/*.py and the torch internals those reach).
 */
#ifndef PDT_B200_H_
#define PDT_B200_H_

#include <stddef.h>
#include <stdint.h>
#include <cuda_runtime_api.h>

/* Bumped whenever an entry point is added or changes its signature; the Python binding refuses a library
 * that reports a different number (a stale build of an older checkout). */
#define PDT_ABI_VERSION 16

#ifdef __cplusplus
extern "C" {
#endif

#define PDT_MAX_RANKS 8
#define PDT_BF16 0
#define PDT_FP16 1
#define PDT_FP32 2

#define PDT_E_UNSUPPORTED (-2)   /* shape/dtype outside what the kernel was written for */
#define PDT_E_TOO_LARGE (-3)     /* size beyond a documented fixed limit */

/* ---- fabric pad layout (one pad per rank, peer-mapped) ------------------- */
#define PDT_COMM_CHANNELS 4            /* 0: gradient stream, 1: main stream (metrics, broadcasts), 2: SyncBatchNorm exchanges, 3: spare */
#define PDT_COMM_MAX_BLOCKS 256
#define PDT_SCRATCH_FLOATS 8192        /* per (channel, source rank) */
#define PDT_PAD_INF_OFFSET_WORDS (PDT_COMM_CHANNELS * PDT_COMM_MAX_BLOCKS * PDT_MAX_RANKS)
#define PDT_PAD_SCRATCH_OFFSET_WORDS (PDT_PAD_INF_OFFSET_WORDS + PDT_COMM_MAX_BLOCKS * PDT_MAX_RANKS)
#define PDT_PAD_BYTES ((size_t)4 * (PDT_PAD_SCRATCH_OFFSET_WORDS + PDT_COMM_CHANNELS * PDT_MAX_RANKS * PDT_SCRATCH_FLOATS))

#ifndef PDT_PEERS_DEFINED
#define PDT_PEERS_DEFINED
typedef struct PdtPeers { void* p[PDT_MAX_RANKS]; } PdtPeers;   /* same VA range seen from this rank, one per rank */
#endif

typedef struct PdtCommDesc {
  int rank, world;            /* world: any value in 1..PDT_MAX_RANKS (the reference ships a 3-GPU launch, start.sh:2-4) */
  void* pad[PDT_MAX_RANKS];   /* PDT_PAD_BYTES each, zero-initialised once */
  void* pad_mc;               /* the pad through an NVLS multicast mapping (all ranks at once), or NULL */
} PdtCommDesc;

/* Up to PDT_MAX_SEGMENTS ranges [off, off+len) of a flat fp32 arena, in elements (host struct). */
#define PDT_MAX_SEGMENTS 8
typedef struct PdtSegments {
  int n;
  size_t off[PDT_MAX_SEGMENTS], len[PDT_MAX_SEGMENTS];
} PdtSegments;

size_t pdt_comm_pad_bytes(void);

/* amp_state: 8 x 32-bit words on the device:
 *   [0] loss scale (f32)  [1] growth tracker (i32)  [2] found_inf, this rank (i32)
 *   [3] skipped steps (i32)  [4] found_inf consensus (i32)
 * hyper: 4 floats on the device: lr, momentum, weight_decay, amp_enabled(0/1).   */

/* Gradient bucket all-reduce, first half (c10d Reducer + ncclAllReduce at
 * distributed.py:262) fused with 1/world, AMP unscale and inf check
 * (_amp_foreach_non_finite_check_and_unscale_, distributed_syncBN_amp.py:276). */
/* grads_mc: the gradient arena through its NVLS multicast mapping, or NULL.  With it the W-way sum of a
 * 16-byte slot is ONE multimem.ld_reduce through the switch instead of W peer loads.  seg_len % (4 * world) == 0. */
int pdt_reduce_scatter_f32(const PdtCommDesc* comm, const PdtPeers* grads, const void* grads_mc, size_t seg_off,
                           size_t seg_len, float* amp_state, int use_flags, int blocks, cudaStream_t stream);

/* Second half: inf consensus + momentum SGD (torch/optim/sgd.py:343-380) on the
 * owned shard + store of fp32 weights and low-precision shadow to every rank
 * (optimizer.step() at distributed.py:263; autocast weight casts).
 * All segments of the arena are served by ONE launch.  params_mc / shadow_mc: multicast mappings of the two
 * arenas (or NULL): each updated 16 bytes then leave this GPU once (multimem.st) instead of once per rank. */
int pdt_sgd_allgather(const PdtCommDesc* comm, const PdtPeers* grads, const PdtPeers* params, const PdtPeers* shadow,
                      void* params_mc, void* shadow_mc, int shadow_dtype, float* momentum, const PdtSegments* segments,
                      const float* hyper, float* amp_state, int use_flags, int blocks, cudaStream_t stream);

/* In-place two-shot all-reduce (dist.all_reduce, distributed.py:80). */
int pdt_allreduce_f32(const PdtCommDesc* comm, const PdtPeers* bufs, size_t off, size_t len, float post_scale,
                      int blocks, cudaStream_t stream);

/* One-shot all-reduce of <= PDT_SCRATCH_FLOATS floats: reduce_mean(loss),
 * reduce_mean(acc1) and the barrier before them (distributed.py:253-255). */
int pdt_allreduce_small_f32(const PdtCommDesc* comm, const float* in, float* out, int n, float post_scale,
                            int channel, cudaStream_t stream);
/* The same exchange for n <= 1024 metrics, followed by AverageMeter.update on the device
 * (losses.update / top1.update, distributed.py:257-258): meter[0..n) += out * weight, meter[n] += weight,
 * meter[n+1 .. 2n+1) = out.  comm == NULL or world == 1: no exchange, just the meter update. */
int pdt_metrics_update(const PdtCommDesc* comm, const float* in, float* out, int n, float post_scale, int channel,
                       float* meter, float weight, cudaStream_t stream);

/* Pull [off, off+len) of the root's arena into every rank's arena (DDP
 * constructor _sync_module_states and per-forward _sync_buffers,
 * torch/nn/parallel/distributed.py:879-890, 2214-2257). */
int pdt_bcast_pull(const PdtCommDesc* comm, const PdtPeers* bufs, size_t off_bytes, size_t len_bytes, int root,
                   int channel, int blocks, cudaStream_t stream);

/* dist.barrier() as a device-side rendezvous (distributed.py:253). */
int pdt_barrier(const PdtCommDesc* comm, int channel, cudaStream_t stream);

/* GradScaler.update() (_amp_update_scale_, distributed_syncBN_amp.py:277). */
int pdt_scale_update(float* amp_state, float growth, float backoff, int interval, cudaStream_t stream);

/* Rank emulation for tests (no reference counterpart): the reduce-scatter, SGD/all-gather,
 * small all-reduce and broadcast bodies run for ALL `world` ranks inside one cooperative launch on one GPU
 * (rank = blockIdx.y), every rank's arena / pad / state a separate buffer of that GPU.  P2P data plane. */
typedef struct PdtEmuRanks {
  void* pad[PDT_MAX_RANKS];
  float* amp_state[PDT_MAX_RANKS];
  float* momentum[PDT_MAX_RANKS];
  float* io[PDT_MAX_RANKS];
} PdtEmuRanks;
int pdt_emu_step(const PdtEmuRanks* ranks, int world, const PdtPeers* grads, const PdtPeers* params, const PdtPeers* shadow,
                 const PdtSegments* segments, const float* hyper, int blocks, cudaStream_t stream);
int pdt_emu_small(const PdtEmuRanks* ranks, int world, int n, float post_scale, const PdtPeers* bufs, size_t len_bytes,
                  int root, cudaStream_t stream);

/* foreach momentum SGD on one flat arena (torch/optim/sgd.py:383-477). */
int pdt_sgd_flat(float* p, const float* g, float* m, void* shadow, int shadow_dtype, size_t n, const float* hyper,
                 float* amp_state, cudaStream_t stream);

/* ---- BatchNorm / SyncBatchNorm on NHWC activations ------------------------
 * x, y, dy, dx, res: [M][C] in `dtype`; statistics and affine parameters fp32.
 * scratch: pdt_bn_scratch_bytes(C) bytes, the first 256 bytes (ticket words of the
 * two-level cross-CTA fold) zero before the first call (self-resetting
 * afterwards); do not share one scratch between kernels that may overlap.  comm == NULL or world == 1: plain BatchNorm (cuDNN/ATen
 * batch_norm, 20 layers of torchvision resnet18).  world > 1: SyncBatchNorm
 * with the cross-rank exchange inside the kernel (replaces
 * torch/nn/modules/_functions.py:39-209; distributed_syncBN_amp.py:143-145). */
#define PDT_BN_MAX_BLOCKS 592
size_t pdt_bn_scratch_bytes(int C);
int pdt_bn_fwd_stats(const void* x, long M, int C, int dtype, void* scratch, const PdtCommDesc* comm,
                     float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                     float* save_invstd, cudaStream_t stream);
/* Allocation-free BatchNorm statistics (no counterpart in the reference: it exists for callers of THIS
 * library that cannot allocate the scratch buffer pdt_bn_fwd_stats needs — code inside latency-critical
 * or graph-replayed loops where cudaMalloc is off limits because it synchronises the device, plugins
 * that must not grow a host framework's memory pool, C callers without an allocator).  Same result as
 * pdt_bn_fwd_stats with comm == NULL; the scratch is a fixed array inside the library.
 *   limit   : C <= PDT_BN_STATIC_MAX_C, else PDT_E_TOO_LARGE; each GPU that loads the library keeps
 *             PDT_BN_STATIC_SCRATCH_BYTES resident until exit;
 *   sharing : all calls on a device use the same array; calls that may overlap on the GPU race;
 *   set-up  : the first call on each device looks the array's address up (safe from several host threads at
 *             once: the cached pointer is published atomically) — do it before any stream capture.
 * Calls no cudaMalloc / cudaFree / allocating routine. */
#define PDT_BN_STATIC_MAX_C 512
#define PDT_BN_STATIC_SCRATCH_BYTES (256 + PDT_BN_MAX_BLOCKS * 2 * PDT_BN_STATIC_MAX_C * 4)
int pdt_bn_fwd_stats_noalloc(const void* x, long M, int C, int dtype, float* running_mean, float* running_var,
                             float momentum, float eps, float* save_mean, float* save_invstd, cudaStream_t stream);
int pdt_bn_fwd_apply(const void* x, const void* res, void* y, long M, int C, int dtype, const float* mean,
                     const float* stat2, const float* gamma, const float* beta, int relu, int eval, float eps,
                     cudaStream_t stream);
/* y may be NULL when relu != 0 and no residual entered the ReLU: the mask is then recomputed from x. */
int pdt_bn_bwd_reduce(const void* dy, const void* y, const void* x, long M, int C, int dtype, const float* mean,
                      const float* invstd, const float* gamma, const float* beta, int relu, void* scratch,
                      const PdtCommDesc* comm, float* dgamma, float* dbeta, float* out_k, cudaStream_t stream);
int pdt_bn_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, long M, int C, int dtype,
                     const float* mean, const float* invstd, const float* gamma, const float* beta, const float* k,
                     int relu, cudaStream_t stream);

/* BatchNorm + ReLU + MaxPool fused (ResNet stem tail: batch_norm + relu_ + max_pool2d_with_indices and
 * their backward kernels).  The normalised tensor is never written: forward pools relu(bn(x)) straight
 * from the conv output x and also saves, per pooled element, the arg-max tap (`idx`) and the raw conv
 * value at the arg-max (`xmax`).  Backward: the per-channel reductions are taken over the POOLED tensors
 * (dpool masked by ypool > 0, times xmax - mean: every pooled gradient lands on exactly one input pixel),
 * then the dx pass gathers the pooled gradient through `idx` and recomputes the ReLU mask from x.
 * Statistics come from pdt_bn_fwd_stats on x. */
int pdt_bn_pool_fwd(const void* x, void* y, int8_t* idx, void* xmax, int dtype, int N, int H, int W, int C, int ks, int st,
                    int pd, const float* mean, const float* invstd, const float* gamma, const float* beta,
                    cudaStream_t stream);
int pdt_bn_pool_bwd_reduce(const void* dpool, const void* ypool, const void* xmax, int dtype, int N, int H, int W, int C,
                           int ks, int st, int pd, const float* mean, const float* invstd, const float* gamma,
                           const float* beta, void* scratch, const PdtCommDesc* comm, float* dgamma, float* dbeta,
                           float* out_k, cudaStream_t stream);
int pdt_bn_pool_bwd_apply(const void* dpool, const int8_t* idx, const void* x, void* dx, int dtype, int N, int H, int W,
                          int C, int ks, int st, int pd, const float* mean, const float* invstd, const float* gamma,
                          const float* beta, const float* k, cudaStream_t stream);

/* ---- loss / pooling / layout ----------------------------------------------
 * pdt_ce_fwd: nn.CrossEntropyLoss (mean) + utils.accuracy top-1 in one pass
 * (distributed.py:247,250; utils.py:105-111).  out = {loss, hit fraction};
 * dlogits_unit = d loss / d logits in fp32.  scratch: pdt_ce_scratch_bytes(B),
 * first word zero before the first call. */
size_t pdt_ce_scratch_bytes(int B);
int pdt_ce_fwd(const void* logits, int dtype, const long long* target, int B, int K, float* dlogits_unit, void* scratch,
               float* out, cudaStream_t stream);
/* out = in * (*scale) cast to dtype (loss-scaled dlogits; GradScaler.scale, distributed_syncBN_amp.py:275). */
int pdt_scale_cast(const float* in, const float* scale, void* out, int dtype, size_t n, cudaStream_t stream);
/* nn.MaxPool2d / nn.AdaptiveAvgPool2d(1) of torchvision ResNet (ATen max_pool2d_with_indices, mean). */
int pdt_maxpool_fwd(const void* x, void* y, int8_t* idx, int dtype, int N, int H, int W, int C, int ks, int st, int pd,
                    cudaStream_t stream);
int pdt_maxpool_bwd(const void* dy, const int8_t* idx, void* dx, int dtype, int N, int H, int W, int C, int ks, int st,
                    int pd, cudaStream_t stream);
int pdt_avgpool_fwd(const void* x, void* y, int dtype, int N, int HW, int C, cudaStream_t stream);
int pdt_avgpool_bwd(const void* dy, void* dx, int dtype, int N, int HW, int C, cudaStream_t stream);
/* autocast input cast + channels_last conversion with channel padding (autocast_mode.cpp casts). */
int pdt_nchw_to_nhwc(const float* x, void* y, int dtype, int N, int Cin, int H, int W, int Cpad, cudaStream_t stream);

/* Destination-side pull of `nbytes` from a peer-mapped pointer (DataParallel scatter / gather /
 * buffer broadcast; torch.cuda.comm.scatter/gather/broadcast_coalesced, dataparallel.py:119). */
int pdt_peer_copy(const void* src, void* dst, size_t nbytes, cudaStream_t stream);
int pdt_abi_version(void);

/* ---- implicit-GEMM convolution on tcgen05 / TMEM / TMA --------------------
 * x: [N][H][W][C], w: [K][R][S][C] (KRSC), y: [N][P][Q][K], all bf16, fp16 (kind::f16) or fp32
 * (kind::tf32: operands read as TF32, fp32 accumulate, like cuDNN with allow_tf32); K % 8 == 0,
 * channel rows 16-byte aligned, 16-byte aligned pointers.  Forward convolution of nn.Conv2d (cuDNN
 * cudnnConvolutionForward; every conv of torchvision resnet18, distributed.py:137,246) and,
 * with dy as x and the flipped/transposed filter as w (pad' = R-1-pad), the data gradient of
 * stride-1 convolutions (cudnnConvolutionBackwardData, distributed.py:262). */
int pdt_conv_fprop(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, int R, int S,
                   int stride, int pad, cudaStream_t stream);
/* y = conv(x, w) + addend, addend laid out like y (same residual-gradient use as pdt_conv3x3_halo_add). */
int pdt_conv_fprop_add(const void* x, const void* w, void* y, const void* addend, int dtype, int N, int H, int W, int C, int K,
                       int R, int S, int stride, int pad, cudaStream_t stream);
/* Convolution + BatchNorm batch statistics in one kernel: the epilogue also produces the per-channel
 * mean / invstd of y (and updates the running statistics), so BatchNorm needs no statistics pass of its
 * own (cudnnConvolutionForward + batch_norm statistics kernel).  Dense NHWC output, K <= 512.
 * scratch: pdt_bn_scratch_bytes(K) bytes, first 256 bytes zero before the first call.
 * comm != NULL with world > 1: SyncBatchNorm — the CTA that finalises the statistics exchanges the raw moments
 * with its peers over NVLink before it writes mean / invstd (the all_gather of torch's sync_batch_norm,
 * torch/nn/modules/_functions.py:74-76, inside the convolution kernel).  Same for the two kernels below. */
int pdt_conv_fprop_bnstats(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, int R, int S,
                           int stride, int pad, void* scratch, const PdtCommDesc* comm, float* running_mean,
                           float* running_var, float momentum, float eps, float* save_mean, float* save_invstd,
                           cudaStream_t stream);
/* 3x3 / stride 1 / pad 1 convolution with input-patch reuse for the large feature maps (W <= 56, K in {64, 128},
 * C % 64 == 0): one tiled TMA load per (TR+2) x (W+2) patch serves all nine taps through row-shifted shared-memory
 * descriptors; the 64x64 filter bank stays resident in shared memory.  scratch != NULL also yields the BatchNorm
 * batch statistics of y (as pdt_conv_fprop_bnstats).  Same cuDNN calls replaced as pdt_conv_fprop. */
int pdt_conv3x3_halo_supported(int dtype, int N, int H, int W, int C, int K);
/* pdt_conv3x3_halo with an optional addend [N][H][W][K] summed into the output in the epilogue (y = conv + addend;
 * requires scratch == NULL).  As the data gradient of a residual block's first convolution it takes the shortcut
 * gradient, replacing the ATen add that autograd issues per block in the reference's backward (distributed.py:262). */
int pdt_conv3x3_halo_add(const void* x, const void* w, void* y, const void* addend, int dtype, int N, int H, int W, int C, int K,
                         void* scratch, const PdtCommDesc* comm, float* running_mean, float* running_var, float momentum,
                         float eps, float* save_mean, float* save_invstd, cudaStream_t stream);
int pdt_conv3x3_halo(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int K, void* scratch,
                     float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                     float* save_invstd, cudaStream_t stream);
/* 3x3 / stride 1 / pad 1 weight gradient for 64 -> 64 channels with the same patch reuse (one x patch + one
 * dy tile per work item, nine row-shifted views, two taps per M=128 instruction); dw [64][3][3][64] fp32 is
 * accumulated into.  pdt_conv_wgrad routes eligible shapes here (PDT_WGRAD_HALO=0 disables). */
int pdt_conv3x3_wgrad_halo_supported(int dtype, int N, int H, int W, int C, int K);
int pdt_conv3x3_wgrad_halo(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int K,
                           cudaStream_t stream);
/* Stem: the 4x4 / stride 1 convolution over the space-to-depth image (pdt_stem_s2d), 16 channels per pixel,
 * 64 output channels: one linear TMA load per 128-pixel tile, sixteen row-shifted K = 16 views (32-byte
 * swizzle), resident filter, optional BatchNorm statistics in the epilogue (scratch != NULL).
 * xs [N][HP][WP][16], wp [64][4][4][16], y [N][HP-3][WP-3][64]. */
int pdt_stem_conv_supported(int dtype, int N, int HP, int WP, int K);
int pdt_stem_conv(const void* xs, const void* wp, void* y, int dtype, int N, int HP, int WP, int K, void* scratch,
                  const PdtCommDesc* comm, float* running_mean, float* running_var, float momentum, float eps,
                  float* save_mean, float* save_invstd, cudaStream_t stream);
/* Stem weight gradient with the same patch reuse (16-bit, 64 output channels, output rows of 16..128 pixels in
 * 16-pixel steps): dwp [64][4][4][16] fp32 += sum over output pixels (zero it first; accumulates with red.add,
 * so the deterministic mode keeps pdt_conv_wgrad_ex).  xs [N][HP][WP][16], dy [N][HP-3][WP-3][64]. */
int pdt_stem_wgrad_supported(int dtype, int N, int HP, int WP, int K);
int pdt_stem_wgrad(const void* xs, const void* dy, float* dwp, int dtype, int N, int HP, int WP, int K, cudaStream_t stream);
/* Weight gradient (cudnnConvolutionBackwardFilter): dw [K][R][S][C] fp32 += sum over pixels; split-K over
 * the pixel dimension with red.global.add, so zero dw first.  C % 64 == 0, K % 64 == 0. */
int pdt_conv_wgrad(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int K, int R, int S,
                   int stride, int pad, cudaStream_t stream);
/* Deterministic weight gradients on the CURRENT device (the reference's --seed turns on cudnn.deterministic,
 * distributed.py:91-99): with a workspace registered, split-K partial tiles are stored into slabs of it and
 * folded in a fixed order instead of red.global.add, and the atomics-based patch-reuse kernel is bypassed.
 * ws == NULL restores the default.  Launches sharing the workspace must be stream-ordered. */
int pdt_set_wgrad_workspace(void* ws, size_t bytes);
/* General forms: fp32 bias [K] in the epilogue (nn.Linear as a 1x1 convolution, cuBLASLt GEMM + bias);
 * separate lower/upper padding with an explicit output size; placement of output pixel (p,q) at
 * (p*os + oa, q*os + ob) of an OH x OW image (one parity class of a strided data gradient, cuDNN strided
 * dgrad); explicit byte strides / channels-per-pixel of the activation as TMA sees it (space-to-depth stem). */
int pdt_conv_fprop_ex(const void* x, const void* w, void* y, const float* bias, int dtype, int N, int H, int W, int C,
                      int Cx, long long x_sw, long long x_sh, long long x_sn, int K, int R, int S, int stride,
                      int pad_lo_h, int pad_lo_w, int pad_hi_h, int pad_hi_w, int P, int Q, int OH, int OW, int os,
                      int oa, int ob, int dil_w, const void* addend /* laid out like the OH x OW output, or NULL */,
                      cudaStream_t stream);
/* The general form with the BatchNorm-statistics epilogue of pdt_conv_fprop_bnstats (dense P x Q output, no bias, no
 * addend, K <= 512): the fp32 stem uses it. */
int pdt_conv_fprop_ex_bnstats(const void* x, const void* w, void* y, int dtype, int N, int H, int W, int C, int Cx, long long x_sw,
                              long long x_sh, long long x_sn, int K, int R, int S, int stride, int pad_lo_h, int pad_lo_w,
                              int pad_hi_h, int pad_hi_w, int P, int Q, int dil_w, void* scratch, const PdtCommDesc* comm,
                              float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                              float* save_invstd, cudaStream_t stream);
int pdt_conv_wgrad_ex(const void* x, const void* dy, float* dw, int dtype, int N, int H, int W, int C, int Cx,
                      long long x_sw, long long x_sh, long long x_sn, int K, int R, int S, int stride, int pad_lo_h,
                      int pad_lo_w, int pad_hi_h, int pad_hi_w, int P, int Q, int dil_w, cudaStream_t stream);
/* Stem input transform: NCHW fp32 -> 2x2 space-to-depth NHWC(16) low precision, zero padded by `pad`
 * (autocast input cast + layout change for the 7x7 stride-2 stem of torchvision resnet). */
int pdt_stem_s2d(const float* x, void* y, int dtype, int N, int H, int W, int pad, cudaStream_t stream);
/* The same from raw uint8 NCHW pixels, with ToTensor + Normalize fused: value = x * scale[c] + bias[c]
 * (scale = 1 / (255 * std), bias = -mean / std; host arrays of 3 floats).  The loader then ships one byte per
 * value over PCIe instead of four. */
int pdt_stem_s2d_u8(const unsigned char* x, void* y, int dtype, int N, int H, int W, int pad, const float* scale,
                    const float* bias, cudaStream_t stream);
/* Filter of the data-gradient convolution, 16-bit elements: out[c][i][j][k] = w[k][rs[i]][ss[j]][c], w = [K][R][S][C],
 * out = [C][Th][Tw][K] (the flip / index_select / transpose / contiguous chain of the ATen formulation in one
 * launch).  rs / ss are host arrays of Th / Tw tap indices (<= 8 each). */
int pdt_dgrad_weight(const void* w, void* out, int K, int C, int R, int S, int Th, int Tw, const int* rs, const int* ss,
                     cudaStream_t stream);
/* The same for many filters in one launch.  `jobs_dev` is a DEVICE array of njobs descriptors; tile_begin is
 * the running sum of ceil(K/t) * ceil(C/t) * Th * Tw over the preceding jobs with t = pdt_dgrad_tile() (64),
 * total_tiles the grand total. */
typedef struct PdtDgradJob {
  const void* src;
  void* dst;
  int K, C, R, S, Th, Tw;
  int rs[8], ss[8];
  int tile_begin;
  int pad_;
} PdtDgradJob;
int pdt_dgrad_weight_batched(const PdtDgradJob* jobs_dev, int njobs, int total_tiles, cudaStream_t stream);
int pdt_dgrad_tile(void);
/* Stem filter [K][3][7][7] (any element strides) -> [K][4][4][16], the filter of the 4x4 / 1 convolution over the
 * space-to-depth image, and the inverse for its fp32 gradient (writes every element of dw). */
int pdt_stem_weight_s2d(const void* w, void* wp, int dtype, int K, long long sk, long long sc, long long sr, long long ss,
                        cudaStream_t stream);
int pdt_stem_weight_grad_unpack(const float* dwp, float* dw, int K, long long sk, long long sc, long long sr, long long ss,
                                cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PDT_B200_H_ */
