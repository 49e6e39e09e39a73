/* dps_b200 native library -- compute kernels (C ABI).  See dps_b200.h for conventions.
 *
 * These replace the reference model's library calls (ATen/oneDNN on CPU, cuDNN in the single-process
 * baseline): reference src/parameter_server/server.py:21-76 (ResNet18 / BasicBlock) and the
 * nn.CrossEntropyLoss of src/workers/worker.py:133,342.  Activations are NHWC bf16.
 */
#ifndef DPS_B200_COMPUTE_H
#define DPS_B200_COMPUTE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- implicit-GEMM convolution on tcgen05/TMEM/TMA (reference: nn.Conv2d, server.py:24,26,32,48) ----
 * kind 0 = fprop : x[N][H][W][Cin]                 -> out[N][H/s][W/s][Cout] (+ per-channel sum/sumsq in stats)
 * kind 1 = dgrad : x = dY[N][H/s][W/s][Cout]       -> out = dX[N][H][W][Cin] (+ addend, may alias out)
 * w is [Cout][w_pitch] bf16 with w_cols valid columns ordered (r, s, ci).  H, W are the conv INPUT dims.
 * Plans own their TMA descriptors; buffers must stay at the same addresses for the plan's lifetime. */
int dps_conv_plan_create(int kind, const void* x, const void* w, void* out, const void* addend, long long* stats,
                         int N, int H, int W, int Cin, int Cout, int R, int stride, int w_cols, int w_pitch,
                         void** out_plan);
/* wgrad: dw[Cout][dw_pitch] fp32 += dY^T * im2col(x)   (atomic accumulation: the caller zeroes dw) */
int dps_wgrad_plan_create(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R,
                          int stride, int valid_cols, int dw_pitch, int nsplit, void** out_plan);
/* dgrad plans only: also perform the next BatchNorm's backward reduction in the epilogue (writes dz = dy*(y>0),
 * accumulates [2C] int64 sums x 2^40) / take the addend at half resolution for parity (0,0) (stride-2 only). */
int dps_conv_plan_fuse_bn_bwd(void* plan, const void* y, const void* x, const float* mean, const float* rstd,
                              long long* sums);
/* after dps_conv_plan_fuse_bn_bwd, for a BN WITHOUT residual input: recompute the ReLU mask from x (needs that BN's
 * gamma / beta, fp32 [C]) instead of reading its forward output y */
int dps_conv_plan_fuse_bn_mask_from_x(void* plan, const float* gamma, const float* beta);
int dps_conv_plan_addend_compact(void* plan, int on);
int dps_conv_plan_run(void* plan, void* stream);
int dps_conv_plan_destroy(void* plan);
int dps_conv_plan_info(void* plan, int* grid_xyz, int* block_n, int* smem_bytes);

/* ---- fused BatchNorm(+residual)+ReLU (reference: nn.BatchNorm2d + torch.relu + `out += shortcut(x)`,
 *      server.py:25-40,49,67-68).  x, residual, y: [P][C] bf16 (NHWC flattened); stats = [2C] 64-bit fixed-point
 *      sum (x2^32) / sumsq (x2^24) produced by the conv epilogue.  training == 0 normalises with the running statistics. ---- */
int dps_bn_forward(const void* x, const long long* stats, const float* gamma, const float* beta, const void* residual,
                   void* y, float* save_mean, float* save_rstd, float* running_mean, float* running_var,
                   long long* num_batches, long long P, int C, int relu, int training, float eps, float momentum,
                   void* stream);
/* sums: [2C] int64 scratch (fixed point x 2^40), zero on entry.   dz (masked gradient) may alias dy.  dgamma/dbeta are accumulated. */
int dps_bn_backward(const void* dy, const void* y, const void* x, const float* mean, const float* rstd,
                    const float* gamma, void* dz, long long* sums, void* dx, float* dgamma, float* dbeta, long long P,
                    int C, int relu, void* stream);

/* Apply pass only: dz and the [2C] int64 sums were already produced by a dgrad plan with dps_conv_plan_fuse_bn_bwd. */
int dps_bn_backward_apply(const void* dz, const void* x, const float* mean, const float* rstd, const float* gamma,
                          const long long* sums, void* dx, float* dgamma, float* dbeta, long long P, int C,
                          void* stream);

/* No counterpart in the reference.  dps_bn_backward for callers that cannot allocate the [2*C] scratch
 * (memory fully owned by a fixed-budget pool, loops that cannot take cudaMalloc's device-wide sync, hooks that
 * only receive their inputs and outputs): the scratch is a built-in device array, max C = 512 channels
 * (8 KiB per GPU, held until exit).  It is cleared on `stream` at the start of every call; calls that overlap
 * on the GPU race on it.  The first call on a device resolves the array's address -- make it before any
 * stream capture.  The array that is used belongs to the device that is CURRENT when the call is made, so the
 * device that owns `stream` must be the current device (a stream of another GPU would be handed the wrong
 * GPU's array).  Larger C is rejected with an error.  Same outputs as dps_bn_backward. */
int dps_bn_backward_noalloc(const void* dy, const void* y, const void* x, const float* mean, const float* rstd,
                            const float* gamma, void* dz, void* dx, float* dgamma, float* dbeta, long long P, int C,
                            int relu, void* stream);

/* ---- fused head: avg-pool + Linear + softmax cross-entropy, forward and backward (reference:
 *      server.py:56-57,73-75; worker.py:133,341-345).  act: [N][HW][F] bf16; w: [K][F] bf16; out[0] = mean loss,
 *      out[1] = number of correct top-1 predictions.  dlogits scratch is [N][128]; dw/db are accumulated.
 *      One launch of <= 128 co-resident CTAs with a grid barrier inside: launches on one device must not overlap.
 *      F must be a multiple of 128 (<= 512), K <= 128. ---- */
int dps_head_forward_backward(const void* act, const void* w, const float* bias, const long long* labels,
                              float* pooled, float* dlogits, float* loss_n, int* correct_n, void* d_act, float* dw,
                              float* db, float* out, int N, int HW, int F, int K, int training, void* stream);

/* ---- stem input staging: fp32 NCHW images -> bf16 im2col rows [N*H*W][64] (27 valid columns, K order (r, s, c)).
 *      zero / zero_words: optional buffer of 64-bit words cleared by the same launch (the step's BN scratch). ---- */
int dps_stem_im2col(const float* img, void* col, int N, int H, int W, void* zero, int zero_words, void* stream);

/* ---- device-resident input pipeline (reference: src/workers/worker.py:146-158 transforms, :181-187 DataLoader):
 *      out[b] = Normalize(ToTensor(Flip(Crop(ZeroPad(src[idx[b]], pad))))) as fp32 NCHW, out_labels[b] = src_labels[idx[b]].
 *      src is the uint8 NHWC dataset in device memory.  train != 0: crop offsets (dy, dx in [0, 2*pad]) and the flip bit
 *      come from splitmix64(seed * 0x100000001B3 + counter0 + b): dy = z % (2*pad+1), dx = (z >> 16) % (2*pad+1),
 *      flip = (z >> 40) & 1.  train == 0: centre crop, no flip.  Rows b >= n_valid are zero-filled with label -1. ---- */
int dps_augment_batch(const unsigned char* src, const long long* src_labels, const long long* idx, float* out,
                      long long* out_labels, int B, int S, int n_valid, unsigned long long seed,
                      unsigned long long counter0, int train, int pad, const float* mean3, const float* std3,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPS_B200_COMPUTE_H */
