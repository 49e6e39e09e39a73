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
int dps_conv_plan_create(int kind, const void* x, const void* w, void* out, const void* addend, float* stats,
                         int N, int H, int W, int Cin, int Cout, int R, int stride, int w_cols, int w_pitch,
                         void** out_plan);
/* wgrad: dw[Cout][dw_pitch] fp32 += dY^T * im2col(x)   (atomic accumulation: the caller zeroes dw) */
int dps_wgrad_plan_create(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R,
                          int stride, int valid_cols, int dw_pitch, int nsplit, void** out_plan);
int dps_conv_plan_run(void* plan, void* stream);
int dps_conv_plan_destroy(void* plan);
int dps_conv_plan_info(void* plan, int* grid_xyz, int* block_n, int* smem_bytes);

#ifdef __cplusplus
}
#endif
#endif /* DPS_B200_COMPUTE_H */
