// Fused classification head: global average pool + Linear(512, classes) + softmax cross-entropy,
// forward AND backward, in two launches.
//
// Replaces AdaptiveAvgPool2d + view + nn.Linear + nn.CrossEntropyLoss and their backward
// (reference: src/parameter_server/server.py:56-57,73-75; src/workers/worker.py:133,341-345).
//
//  head_sample_kernel (one CTA per sample): pool -> logits -> log-softmax -> loss_n, correct_n,
//      dlogits = (p - onehot)/N, dpooled = dlogits * W, and the broadcast d_act = dpooled / HW (bf16).
//  head_param_kernel (one CTA per class): dW[j] += dlogits[:, j]^T * pooled, db[j] += sum dlogits[:, j];
//      CTA 0 also reduces the mean loss and the correct count.
#include "common.cuh"
#include "../include/dps_b200_compute.h"

namespace dps {

constexpr int kHeadThreads = 512;
constexpr int kMaxFeat = 512;
constexpr int kMaxClasses = 128;

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

__global__ void __launch_bounds__(kHeadThreads) head_sample_kernel(
    const __nv_bfloat16* __restrict__ act, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
    const long long* __restrict__ labels, float* __restrict__ pooled_out, float* __restrict__ dlogits_out,
    float* __restrict__ loss_n, int* __restrict__ correct_n, __nv_bfloat16* __restrict__ d_act, int N, int HW,
    int F, int K, int training) {
  __shared__ float s_pool[kMaxFeat];
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_logit[kMaxClasses];
  __shared__ float s_red[2];
  const int n = blockIdx.x;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const __nv_bfloat16* a = act + (size_t)n * HW * F;
  const float inv_hw = 1.f / (float)HW;
  for (int c = tid; c < F; c += blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(a[(size_t)p * F + c]);
    s *= inv_hw;
    s_pool[c] = s;
    pooled_out[(size_t)n * F + c] = s;
  }
  __syncthreads();
  for (int j = warp; j < K; j += kHeadThreads / 32) {
    const __nv_bfloat16* wj = w + (size_t)j * F;
    float s = 0.f;
    for (int c = lane * 2; c < F; c += 64) {
      const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(wj + c));
      s += f.x * s_pool[c] + f.y * s_pool[c + 1];
    }
    s = warp_sum(s);
    if (lane == 0) s_logit[j] = s + bias[j];
  }
  __syncthreads();
  const int label = (int)labels[n];
  if (warp == 0) {
    float m = -INFINITY;
    for (int j = lane; j < K; j += 32) m = fmaxf(m, s_logit[j]);
    m = warp_max(m);
    float se = 0.f;
    for (int j = lane; j < K; j += 32) se += __expf(s_logit[j] - m);
    se = warp_sum(se);
    // first index attaining the max (torch.max semantics of the reference eval loop, worker.py:324)
    int arg = K;
    for (int j = lane; j < K; j += 32)
      if (s_logit[j] == m) arg = min(arg, j);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) arg = min(arg, __shfl_xor_sync(0xffffffffu, arg, o));
    if (lane == 0) {
      s_red[0] = m;
      s_red[1] = se;
      const bool valid = label >= 0 && label < K;        // label < 0 marks a padding row (eval tail batch)
      loss_n[n] = valid ? (m + __logf(se)) - s_logit[label] : 0.f;
      correct_n[n] = (valid && arg == label) ? 1 : 0;
    }
  }
  __syncthreads();
  if (!training) return;
  const float m = s_red[0], inv_se = 1.f / s_red[1], inv_n = 1.f / (float)N;
  for (int j = tid; j < K; j += blockDim.x) {
    const float pj = __expf(s_logit[j] - m) * inv_se;
    const float d = (pj - (j == label ? 1.f : 0.f)) * inv_n;
    s_logit[j] = d;
    dlogits_out[(size_t)n * kMaxClasses + j] = d;
  }
  __syncthreads();
  for (int c = tid; c < F; c += blockDim.x) {
    float s = 0.f;
    for (int j = 0; j < K; ++j) s += s_logit[j] * __bfloat162float(w[(size_t)j * F + c]);
    s_pool[c] = s * inv_hw;                       // d_act value for every spatial position
  }
  __syncthreads();
  __nv_bfloat16* da = d_act + (size_t)n * HW * F;
  for (int i = tid; i < HW * F / 2; i += blockDim.x) {
    const int c = (i * 2) % F;
    *reinterpret_cast<uint32_t*>(da + (size_t)i * 2) = pack_bf16x2(s_pool[c], s_pool[c + 1]);
  }
}

// grid = (K classes, F / 128 feature chunks), 128 threads: one (class, feature) pair per thread.
__global__ void __launch_bounds__(128) head_param_kernel(
    const float* __restrict__ pooled, const float* __restrict__ dlogits, const float* __restrict__ loss_n,
    const int* __restrict__ correct_n, float* __restrict__ dw, float* __restrict__ db, float* __restrict__ out,
    int N, int F, int K, int training) {
  __shared__ float s_dl[1024];
  pdl_launch_dependents();
  pdl_wait();
  const int j = blockIdx.x;
  const int tid = threadIdx.x;
  if (training && j < K) {
    for (int n0 = 0; n0 < N; n0 += 1024) {
      const int cnt = min(1024, N - n0);
      __syncthreads();
      for (int n = tid; n < cnt; n += blockDim.x) s_dl[n] = dlogits[(size_t)(n0 + n) * kMaxClasses + j];
      __syncthreads();
      const int c = blockIdx.y * 128 + tid;
      if (c < F) {
        float acc = 0.f;
        int n = 0;
        for (; n + 8 <= cnt; n += 8) {
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = pooled[(size_t)(n0 + n + u) * F + c];
#pragma unroll
          for (int u = 0; u < 8; ++u) acc += s_dl[n + u] * v[u];
        }
        for (; n < cnt; ++n) acc += s_dl[n] * pooled[(size_t)(n0 + n) * F + c];
        dw[(size_t)j * F + c] += acc;
      }
      if (blockIdx.y == 0 && tid < 32) {
        float sdl = 0.f;
        for (int n = tid; n < cnt; n += 32) sdl += s_dl[n];
        sdl = warp_sum(sdl);
        if (tid == 0) db[j] += sdl;
      }
    }
  }
  if (j == 0 && blockIdx.y == 0 && tid >= 32 && tid < 64) {
    const int lane = tid - 32;
    float sl = 0.f;
    int c = 0;
    for (int n = lane; n < N; n += 32) { sl += loss_n[n]; c += correct_n[n]; }
    sl = warp_sum(sl);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) { out[0] = sl / (float)N; out[1] = (float)c; }
  }
}

// Input staging for the stem: fp32 NCHW images -> bf16 im2col rows [N*H*W][64]
// (27 valid columns ordered (r, s, c) + zero padding), so the 3-channel stem conv runs as a K=64 GEMM on
// the same tcgen05 kernel as every other conv (reference stem: server.py:48).
__global__ void __launch_bounds__(256) stem_im2col_kernel(const float* __restrict__ img, __nv_bfloat16* __restrict__ col,
                                                          int N, int H, int W) {
  pdl_launch_dependents();
  pdl_wait();
  // only columns 0..31 are written (27 valid + 5 zeros); columns 32..63 of the buffer are zero from allocation
  const long long total = (long long)N * H * W * 4;        // 4 vectors of 8 columns per pixel
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int vec = (int)(i & 3);
    long long pix = i >> 2;
    const int x = (int)(pix % W);
    const int y = (int)((pix / W) % H);
    const int n = (int)(pix / ((long long)W * H));
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = vec * 8 + e;
      float val = 0.f;
      if (k < 27) {
        const int c = k % 3, s = (k / 3) % 3, r = k / 9;
        const int iy = y + r - 1, ix = x + s - 1;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) val = __ldg(img + (((size_t)n * 3 + c) * H + iy) * W + ix);
      }
      v[e] = val;
    }
    st_v4(col + pix * 64 + vec * 8, make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                  pack_bf16x2(v[6], v[7])));
  }
}

}  // namespace dps

using namespace dps;

DPS_API int dps_head_forward_backward(const void* act, const void* w, const float* bias, const long long* labels,
                                      float* pooled, float* dlogits, float* loss_n, int* correct_n, void* d_act,
                                      float* dw, float* db, float* out, int N, int HW, int F, int K, int training,
                                      void* stream) {
  if (F > kMaxFeat || F % 64 || K > kMaxClasses) DPS_FAIL(1, "head: unsupported F=%d K=%d", F, K);
  cudaStream_t st = (cudaStream_t)stream;
  DPS_CHECK(launch_pdl(head_sample_kernel, dim3(N), dim3(kHeadThreads), 0, st, (const __nv_bfloat16*)act, (const __nv_bfloat16*)w, bias, labels,
                                                 pooled, dlogits, loss_n, correct_n, (__nv_bfloat16*)d_act, N, HW, F,
                                                 K, training));
  DPS_CHECK(cudaGetLastError());
  DPS_CHECK(launch_pdl(head_param_kernel, dim3(training ? K : 1, training ? (F + 127) / 128 : 1), dim3(128), 0, st,
      pooled, dlogits, loss_n, correct_n, dw, db, out, N, F, K, training));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_stem_im2col(const float* img, void* col, int N, int H, int W, void* stream) {
  const long long total = (long long)N * H * W * 4;
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  // first kernel of a step: its predecessor is a memset / H2D copy node, so it is launched without the
  // programmatic-dependency attribute (griddepcontrol.wait in the kernel is then a no-op)
  stem_im2col_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(img, (__nv_bfloat16*)col, N, H, W);
  DPS_CHECK(cudaGetLastError());
  return 0;
}
