// Fused classification head: global average pool + Linear(512, classes) + softmax cross-entropy,
// forward AND backward, in ONE launch (a grid barrier separates the per-sample part from the parameter gradients).
//
// Replaces AdaptiveAvgPool2d + view + nn.Linear + nn.CrossEntropyLoss and their backward
// (reference: src/parameter_server/server.py:56-57,73-75; src/workers/worker.py:133,341-345).
//
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

// Grid-wide barrier of the head kernel (all CTAs are co-resident: grid <= 128 CTAs of one per SM).  Sense-reversing,
// so it is correct for any grid size from launch to launch; launches of the head on ONE device must not overlap.
__device__ unsigned int g_head_bar_count = 0;
__device__ volatile unsigned int g_head_bar_gen = 0;

__device__ __forceinline__ void head_grid_barrier(unsigned int nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int gen = g_head_bar_gen;
    __threadfence();
    if (atomicAdd(&g_head_bar_count, 1u) == nblocks - 1) {
      g_head_bar_count = 0;
      __threadfence();
      g_head_bar_gen = gen + 1;
    } else {
      const long long t0 = clock64();
      while (g_head_bar_gen == gen) {
        __nanosleep(64);
        if (clock64() - t0 > 4000000000LL) { printf("dps: head grid barrier timeout (block %d)\n", blockIdx.x); __trap(); }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// ONE launch, grid = min(N, 128) CTAs x 512 threads, the whole classifier matrix resident in shared memory.
//  phase 1 (per sample, CTAs stride over samples): 16-byte pooled loads -> logits (warp per class, weights from
//      shared memory) -> log-softmax, loss_n, correct_n -> dlogits -> dpooled = dlogits * W (shared memory) ->
//      d_act broadcast (16-byte stores); pooled / dlogits also go to global memory for phase 2.
//  grid barrier
//  phase 2 (work item = 128 features x 4 classes, batch split over 4 thread slices): dW += dlogits^T * pooled,
//      db += sum dlogits; CTA 0 reduces the mean loss and the correct count.
__global__ void __launch_bounds__(kHeadThreads, 1) head_fused_kernel(
    const __nv_bfloat16* __restrict__ act, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
    const long long* __restrict__ labels, float* __restrict__ pooled_out, float* __restrict__ dlogits_out,
    float* __restrict__ loss_n, int* __restrict__ correct_n, __nv_bfloat16* __restrict__ d_act,
    float* __restrict__ dw, float* __restrict__ db, float* __restrict__ out, int N, int HW, int F, int K,
    int training) {
  extern __shared__ __align__(16) uint8_t head_smem[];
  __nv_bfloat16* s_w = reinterpret_cast<__nv_bfloat16*>(head_smem);                  // [K][F]
  float* s_pool = reinterpret_cast<float*>(head_smem + (size_t)kMaxClasses * kMaxFeat * 2);   // [F]
  float* s_part = s_pool + kMaxFeat;                                                 // [8][F] pooling partials / phase-2 scratch
  float* s_logit = s_part + 8 * kMaxFeat;                                            // [kMaxClasses]
  float* s_red = s_logit + kMaxClasses;                                              // [2]
  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // classifier weights do not depend on the previous kernel: start copying them before the dependency wait
  const int wvecs = K * F / 8;
  for (int i = tid; i < wvecs; i += kHeadThreads)
    *reinterpret_cast<uint4*>(s_w + (size_t)i * 8) = ld_nc_v4(w + (size_t)i * 8);
  pdl_wait();
  const float inv_hw = 1.f / (float)HW, inv_n = 1.f / (float)N;
  const int cgs = F / 8;                       // 8-channel groups (<= 64)
  const int psub = kHeadThreads / cgs;         // position subsets (>= 8)
  for (int n = blockIdx.x; n < N; n += gridDim.x) {
    __syncthreads();
    // ---- average pool: thread = (8 channels, subset of positions), 16-byte loads
    {
      const int cg = tid % cgs, ps = tid / cgs;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const __nv_bfloat16* a = act + (size_t)n * HW * F + cg * 8;
      for (int p = ps; p < HW; p += psub) {
        const uint4 v = ld_nc_v4(a + (size_t)p * F);
        float2 f;
        f = unpack_bf16x2(v.x); acc[0] += f.x; acc[1] += f.y;
        f = unpack_bf16x2(v.y); acc[2] += f.x; acc[3] += f.y;
        f = unpack_bf16x2(v.z); acc[4] += f.x; acc[5] += f.y;
        f = unpack_bf16x2(v.w); acc[6] += f.x; acc[7] += f.y;
      }
      if (ps < 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) s_part[ps * kMaxFeat + cg * 8 + i] = acc[i];
      }
      __syncthreads();
      // subsets 8.. (only when F < 512) fold into the first eight in a second round
      if (ps >= 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) atomicAdd(&s_part[(ps & 7) * kMaxFeat + cg * 8 + i], acc[i]);
      }
      __syncthreads();
      for (int c = tid; c < F; c += kHeadThreads) {
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) sum += s_part[k * kMaxFeat + c];
        sum *= inv_hw;
        s_pool[c] = sum;
        pooled_out[(size_t)n * F + c] = sum;
      }
    }
    __syncthreads();
    // ---- logits: one warp per class, weights from shared memory (lane reads 2 channels per step)
    for (int j = warp; j < K; j += kHeadThreads / 32) {
      const __nv_bfloat16* wj = s_w + (size_t)j * F;
      float sum = 0.f;
      for (int c = lane * 2; c < F; c += 64) {
        const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(wj + c));
        sum += f.x * s_pool[c] + f.y * s_pool[c + 1];
      }
      sum = warp_sum(sum);
      if (lane == 0) s_logit[j] = sum + bias[j];
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
    if (training) {
      const float m = s_red[0], inv_se = 1.f / s_red[1];
      for (int j = tid; j < K; j += kHeadThreads) {
        const float pj = __expf(s_logit[j] - m) * inv_se;
        const float d = (pj - (j == label ? 1.f : 0.f)) * inv_n;
        s_logit[j] = d;
        dlogits_out[(size_t)n * kMaxClasses + j] = d;
      }
      __syncthreads();
      // ---- dpooled = dlogits * W: thread = channel pair, weights from shared memory
      for (int c2 = tid; c2 < F / 2; c2 += kHeadThreads) {
        float s0 = 0.f, s1 = 0.f;
#pragma unroll 4
        for (int j = 0; j < K; ++j) {
          const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(s_w + (size_t)j * F + c2 * 2));
          const float d = s_logit[j];
          s0 += d * f.x; s1 += d * f.y;
        }
        s_pool[c2 * 2] = s0 * inv_hw;                  // d_act value for every spatial position
        s_pool[c2 * 2 + 1] = s1 * inv_hw;
      }
      __syncthreads();
      __nv_bfloat16* da = d_act + (size_t)n * HW * F;
      const int cg = tid % cgs;
      const uint4 v = make_uint4(pack_bf16x2(s_pool[cg * 8 + 0], s_pool[cg * 8 + 1]), pack_bf16x2(s_pool[cg * 8 + 2], s_pool[cg * 8 + 3]),
                                 pack_bf16x2(s_pool[cg * 8 + 4], s_pool[cg * 8 + 5]), pack_bf16x2(s_pool[cg * 8 + 6], s_pool[cg * 8 + 7]));
      for (int p = tid / cgs; p < HW; p += psub) st_v4(da + (size_t)p * F + cg * 8, v);
    }
  }
  // ---- everything of phase 1 is visible to every CTA from here on
  head_grid_barrier(gridDim.x);
  if (blockIdx.x == 0 && warp == 0) {
    float sl = 0.f;
    int c = 0;
    for (int n = lane; n < N; n += 32) { sl += __ldcg(loss_n + n); c += __ldcg(correct_n + n); }
    sl = warp_sum(sl);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) { out[0] = sl / (float)N; out[1] = (float)c; }
  }
  if (!training) return;
  // ---- phase 2: dW / db.  work item = (feature chunk of 128, class group of 4); thread = (feature, batch slice)
  float* s_dl = s_part;                                  // [256][4] dlogits of the current sample chunk
  float* s_acc = s_part + 1024;                          // [4 slices][4 classes][128 features]
  const int nfc = F / 128, ncg = (K + 3) / 4;
  const int f = tid & 127, q = tid >> 7;
  for (int item = blockIdx.x; item < nfc * ncg; item += gridDim.x) {
    const int fc = item % nfc, cg = item / nfc;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    float dbs = 0.f;
    for (int n0 = 0; n0 < N; n0 += 256) {
      const int cnt = min(256, N - n0);
      __syncthreads();
      for (int i = tid; i < cnt * 4; i += kHeadThreads) {
        const int j = cg * 4 + (i & 3);
        s_dl[i] = j < K ? __ldcg(dlogits_out + (size_t)(n0 + (i >> 2)) * kMaxClasses + j) : 0.f;
      }
      __syncthreads();
      const float* pp = pooled_out + (size_t)n0 * F + fc * 128 + f;
#pragma unroll 8
      for (int n = q; n < cnt; n += 4) {
        const float pv = __ldcg(pp + (size_t)n * F);
        const float4 d = *reinterpret_cast<const float4*>(s_dl + n * 4);
        acc[0] += pv * d.x; acc[1] += pv * d.y; acc[2] += pv * d.z; acc[3] += pv * d.w;
      }
      if (fc == 0 && tid < 4)
        for (int n = 0; n < cnt; ++n) dbs += s_dl[n * 4 + tid];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s_acc[(q * 4 + k) * 128 + f] = acc[k];
    __syncthreads();
    {
      const int j = cg * 4 + q;                          // thread (q, f) finishes class q of the group
      if (j < K) {
        const float sum = s_acc[(0 * 4 + q) * 128 + f] + s_acc[(1 * 4 + q) * 128 + f] + s_acc[(2 * 4 + q) * 128 + f] +
                          s_acc[(3 * 4 + q) * 128 + f];
        dw[(size_t)j * F + fc * 128 + f] += sum;
      }
    }
    if (fc == 0 && tid < 4 && cg * 4 + tid < K) db[cg * 4 + tid] += dbs;
  }
}

// Input staging for the stem: fp32 NCHW images -> bf16 im2col rows [N*H*W][64]
// (27 valid columns ordered (r, s, c) + zero padding), so the 3-channel stem conv runs as a K=64 GEMM on
// the same tcgen05 kernel as every other conv (reference stem: server.py:48).
// The kernel also clears `zero` (the step's BatchNorm statistics scratch), which saves a memset node per step.
//
// W == 32 (the CIFAR shape): one warp per image row.  A lane loads the 3 channels x 3 rows at its own x (nine
// fully coalesced 128-byte warp loads), gets the x-1 / x+1 neighbours from the adjacent lanes by shuffle (lanes 0 and
// 31 are the image border: zero), and writes its 64 bytes.  Other widths use the generic one-vector-per-thread path.
__global__ void __launch_bounds__(256) stem_im2col_kernel(const float* __restrict__ img, __nv_bfloat16* __restrict__ col,
                                                          int N, int H, int W, unsigned long long* __restrict__ zero,
                                                          int zero_words) {
  pdl_launch_dependents();
  pdl_wait();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < zero_words; i += gridDim.x * blockDim.x) zero[i] = 0ull;
  if (W == 32) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < N * H; row += warps) {
      const int n = row / H, y = row - n * H;
      float v[3][3][3];                                   // [r][s][c]
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int iy = y + r - 1;
        const bool in = iy >= 0 && iy < H;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float mid = in ? __ldg(img + (((size_t)n * 3 + c) * H + iy) * 32 + lane) : 0.f;
          const float left = __shfl_up_sync(0xffffffffu, mid, 1), right = __shfl_down_sync(0xffffffffu, mid, 1);
          v[r][0][c] = lane == 0 ? 0.f : left;
          v[r][1][c] = mid;
          v[r][2][c] = lane == 31 ? 0.f : right;
        }
      }
      const float* f = &v[0][0][0];                       // k = (r*3 + s)*3 + c, exactly the K order of the weights
      __nv_bfloat16* dst = col + ((size_t)row * 32 + lane) * 64;
      st_v4(dst, make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7])));
      st_v4(dst + 8, make_uint4(pack_bf16x2(f[8], f[9]), pack_bf16x2(f[10], f[11]), pack_bf16x2(f[12], f[13]), pack_bf16x2(f[14], f[15])));
      st_v4(dst + 16, make_uint4(pack_bf16x2(f[16], f[17]), pack_bf16x2(f[18], f[19]), pack_bf16x2(f[20], f[21]), pack_bf16x2(f[22], f[23])));
      st_v4(dst + 24, make_uint4(pack_bf16x2(f[24], f[25]), pack_bf16x2(f[26], 0.f), 0u, 0u));
    }
    return;
  }
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

constexpr int kHeadSmem = kMaxClasses * kMaxFeat * 2 + (kMaxFeat + 8 * kMaxFeat + kMaxClasses + 8) * 4;

DPS_API int dps_head_forward_backward(const void* act, const void* w, const float* bias, const long long* labels,
                                      float* pooled, float* dlogits, float* loss_n, int* correct_n, void* d_act,
                                      float* dw, float* db, float* out, int N, int HW, int F, int K, int training,
                                      void* stream) {
  if (F > kMaxFeat || F % 128 || K > kMaxClasses || K < 1 || N < 1 || HW < 1)
    DPS_FAIL(1, "head: unsupported N=%d HW=%d F=%d K=%d", N, HW, F, K);
  cudaStream_t st = (cudaStream_t)stream;
  static bool configured = false;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(head_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kHeadSmem));
    configured = true;
  }
  // every CTA must be resident for the grid barrier: one CTA per SM (149 KiB of shared memory), at most 128 of them
  const int grid = N < 128 ? N : 128;
  DPS_CHECK(launch_pdl(head_fused_kernel, dim3(grid), dim3(kHeadThreads), kHeadSmem, st, (const __nv_bfloat16*)act,
                       (const __nv_bfloat16*)w, bias, labels, pooled, dlogits, loss_n, correct_n, (__nv_bfloat16*)d_act,
                       dw, db, out, N, HW, F, K, training));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

// zero / zero_words: optional buffer of 64-bit words cleared by the same launch (NULL / 0: nothing)
DPS_API int dps_stem_im2col(const float* img, void* col, int N, int H, int W, void* zero, int zero_words,
                            void* stream) {
  const long long total = W == 32 ? (long long)N * H * 32 : (long long)N * H * W * 4;
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  // first kernel of a step: its predecessor is an H2D copy node, so it is launched without the
  // programmatic-dependency attribute (griddepcontrol.wait in the kernel is then a no-op)
  stem_im2col_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(img, (__nv_bfloat16*)col, N, H, W,
                                                             (unsigned long long*)zero, zero_words);
  DPS_CHECK(cudaGetLastError());
  return 0;
}
