// Fused BatchNorm(+residual)+ReLU forward and backward on NHWC bf16 activations.
//
// Replaces nn.BatchNorm2d + torch.relu + the residual `out += shortcut(x)` of the reference blocks
// (reference: src/parameter_server/server.py:25-40,49,67-68) and their autograd backward.
//
// Forward: the per-channel batch sums (sum, sum of squares) arrive from the producing conv's epilogue
// (conv_tc.cu), so the forward is ONE bandwidth-bound pass: normalise + affine (+ residual) (+ ReLU),
// bf16 out, plus saved mean / rstd and the running-statistics update (momentum 0.1, unbiased variance,
// num_batches_tracked) done by CTA 0.
// Backward: two passes.  reduce: dz = dy * (y > 0); per-channel sum(dz), sum(dz * xhat) (registers ->
// shared -> one global atomic per channel per CTA), writes dz.  apply: dx = gamma*rstd*(dz - mean(dz)
// - xhat*mean(dz*xhat)), and CTA 0 accumulates dgamma / dbeta into the flat gradient arena.
#include "common.cuh"
#include "../include/dps_b200_compute.h"

namespace dps {

constexpr int kBnThreads = 256;
constexpr int kMaxC = 512;
constexpr double kGradSumScale = 1099511627776.0;   // 2^40 (keep in sync with conv_tc.cu)

struct V8 {
  float v[8];
};
__device__ __forceinline__ V8 unpack8(uint4 u) {
  V8 r;
  float2 f;
  f = unpack_bf16x2(u.x); r.v[0] = f.x; r.v[1] = f.y;
  f = unpack_bf16x2(u.y); r.v[2] = f.x; r.v[3] = f.y;
  f = unpack_bf16x2(u.z); r.v[4] = f.x; r.v[5] = f.y;
  f = unpack_bf16x2(u.w); r.v[6] = f.x; r.v[7] = f.y;
  return r;
}
__device__ __forceinline__ uint4 pack8(const V8& r) {
  return make_uint4(pack_bf16x2(r.v[0], r.v[1]), pack_bf16x2(r.v[2], r.v[3]), pack_bf16x2(r.v[4], r.v[5]),
                    pack_bf16x2(r.v[6], r.v[7]));
}

__global__ void __launch_bounds__(kBnThreads, 3) bn_fwd_kernel(
    const __nv_bfloat16* __restrict__ x, const long long* __restrict__ stats, const float* __restrict__ gamma,
    const float* __restrict__ beta, const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ y,
    float* __restrict__ save_mean, float* __restrict__ save_rstd, float* __restrict__ running_mean,
    float* __restrict__ running_var, long long* __restrict__ num_batches, long long P, int C, int relu,
    int training, float eps, float momentum, double inv_sum, double inv_sq) {
  // inv_sum = 1 / (2^32 P), inv_sq = 1 / (2^24 P), computed on the host: the FP64 pipe of this part is vestigial, and
  // two double divisions per channel in every CTA's prologue were a measurable share of the small layers' ~4 us
  __shared__ float s_scale[kMaxC], s_shift[kMaxC];
  pdl_launch_dependents();
  pdl_wait();
  // The first batch of activation loads is issued BEFORE the per-channel prologue (statistics -> scale / shift), so
  // the two dependent L2 round trips of a small layer overlap instead of adding up.
  constexpr int kUnroll = 4;
  const int cg = C / 8;
  const long long nvec = P * cg;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long v0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  uint4 xa[kUnroll], ra[kUnroll];
#pragma unroll
  for (int j = 0; j < kUnroll; ++j) {
    const long long vv = v0 + j * stride;
    if (vv < nvec) {
      xa[j] = ld_nc_v4(x + vv * 8);
      if (residual != nullptr) ra[j] = ld_nc_v4(residual + vv * 8);
    }
  }
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float mean, var;
    if (training) {
      // 64-bit fixed point from the conv epilogue: sum x 2^32, sum of squares x 2^24
      // fp32 from here on (the fixed-point SUMS are exact; E[x^2] - mean^2 in fp32 loses ~6e-8 * (1 + mean^2 / var)
      // relative precision, far below bf16 activations) -- no FP64 instruction in any CTA's prologue
      mean = __ll2float_rn(stats[c]) * (float)inv_sum;
      var = fmaxf(__ll2float_rn(stats[C + c]) * (float)inv_sq - mean * mean, 0.f);
    } else {
      mean = running_mean[c];
      var = running_var[c];
    }
    const float rstd = rsqrtf(var + eps);
    const float sc = gamma[c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = beta[c] - mean * sc;
    if (training && (c % gridDim.x) == blockIdx.x) {    // spread over the grid: no single CTA is delayed by it
      save_mean[c] = mean;
      save_rstd[c] = rstd;
      const float unbiased = var * (float)P / (float)max(1LL, P - 1);
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
      if (c == 0) *num_batches += 1;     // channel 0 is owned by CTA 0 only
    }
  }
  __syncthreads();
  // blockDim (256) is a multiple of C/8, so a thread always lands on the same 8 channels: keep their
  // scale/shift in registers and issue kUnroll independent 16-byte loads per tensor before using any of them.
  const int c0 = (int)(v0 % cg) * 8;
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { sc[i] = s_scale[c0 + i]; sh[i] = s_shift[c0 + i]; }
  for (long long v = v0; v < nvec; v += kUnroll * stride) {
    if (v != v0) {
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) {
        const long long vv = v + j * stride;
        if (vv < nvec) {
          xa[j] = ld_nc_v4(x + vv * 8);
          if (residual != nullptr) ra[j] = ld_nc_v4(residual + vv * 8);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const long long vv = v + j * stride;
      if (vv >= nvec) break;
      V8 a = unpack8(xa[j]);
      V8 r = {};
      if (residual != nullptr) r = unpack8(ra[j]);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float o = a.v[i] * sc[i] + sh[i] + r.v[i];
        a.v[i] = relu ? fmaxf(o, 0.f) : o;
      }
      st_v4(y + vv * 8, pack8(a));
    }
  }
}

__global__ void __launch_bounds__(kBnThreads) bn_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
    const float* __restrict__ mean, const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dz,
    long long* __restrict__ sums, long long P, int C, int relu) {
  // Cross-thread reduction in a fixed order (no shared-memory atomics), then ONE 64-bit fixed-point (x 2^40)
  // integer atomic per channel per CTA: integer adds are exact and order-independent, so the activation gradients
  // of a step are bit-reproducible run to run (bf16 re-rounding would otherwise amplify fp32 summation jitter).
  __shared__ float s_part[kBnThreads * 16];            // [row slot][2C] partials
  pdl_launch_dependents();
  pdl_wait();
  const int cg = C / 8;                               // channel groups; 256 % cg == 0
  const int my_cg = threadIdx.x % cg;
  const int rows_per_pass = blockDim.x / cg;
  const int c0 = my_cg * 8;
  float mu[8], rs[8], a_dz[8], a_dzx[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { mu[i] = mean[c0 + i]; rs[i] = rstd[c0 + i]; a_dz[i] = 0.f; a_dzx[i] = 0.f; }
  constexpr int kUnroll = 4;
  const long long rstride = (long long)gridDim.x * rows_per_pass;
  for (long long row = (long long)blockIdx.x * rows_per_pass + threadIdx.x / cg; row < P; row += kUnroll * rstride) {
    uint4 ga[kUnroll], xa[kUnroll], ya[kUnroll];
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const long long rr = row + j * rstride;
      if (rr < P) {
        const long long off = rr * C + c0;
        ga[j] = ld_nc_v4(dy + off);
        xa[j] = ld_nc_v4(x + off);
        if (relu) ya[j] = ld_nc_v4(y + off);
      }
    }
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const long long rr = row + j * rstride;
      if (rr >= P) break;
      const long long off = rr * C + c0;
      V8 g = unpack8(ga[j]);
      V8 xv = unpack8(xa[j]);
      if (relu) {
        V8 yv = unpack8(ya[j]);
#pragma unroll
        for (int i = 0; i < 8; ++i) g.v[i] = yv.v[i] > 0.f ? g.v[i] : 0.f;
        st_v4(dz + off, pack8(g));
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a_dz[i] += g.v[i];
        a_dzx[i] += g.v[i] * (xv.v[i] - mu[i]) * rs[i];
      }
    }
  }
  const int slot = threadIdx.x / cg;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    s_part[slot * 2 * C + c0 + i] = a_dz[i];
    s_part[slot * 2 * C + C + c0 + i] = a_dzx[i];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    double acc = 0.0;
    for (int k = 0; k < rows_per_pass; ++k) acc += (double)s_part[k * 2 * C + c];
    atomicAdd(reinterpret_cast<unsigned long long*>(sums + c),
              (unsigned long long)__double2ll_rn(acc * kGradSumScale));
  }
}

__global__ void __launch_bounds__(kBnThreads, 3) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dz, const __nv_bfloat16* __restrict__ x, const float* __restrict__ mean,
    const float* __restrict__ rstd, const float* __restrict__ gamma, const long long* __restrict__ sums,
    __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, long long P, int C) {
  __shared__ float s_a[kMaxC], s_b[kMaxC], s_c[kMaxC];   // dx = s_a*dz + s_b*x + s_c
  pdl_launch_dependents();
  pdl_wait();
  // first batch of loads before the per-channel prologue (see bn_fwd_kernel)
  constexpr int kUnroll = 4;
  const int cg = C / 8;
  const long long nvec = P * cg;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long v0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  uint4 ga[kUnroll], xa[kUnroll];
#pragma unroll
  for (int j = 0; j < kUnroll; ++j) {
    const long long vv = v0 + j * stride;
    if (vv < nvec) { ga[j] = ld_nc_v4(dz + vv * 8); xa[j] = ld_nc_v4(x + vv * 8); }
  }
  const float invP = 1.f / (float)P;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float sdz = __ll2float_rn(sums[c]) * (float)(1.0 / kGradSumScale);
    const float sdzx = __ll2float_rn(sums[C + c]) * (float)(1.0 / kGradSumScale);
    const float gr = gamma[c] * rstd[c];
    // dx = gr * (dz - sdz/P - xhat * sdzx/P),  xhat = (x - mean) * rstd
    const float k = gr * rstd[c] * sdzx * invP;
    s_a[c] = gr;
    s_b[c] = -k;
    s_c[c] = -gr * sdz * invP + k * mean[c];
    if ((c % gridDim.x) == blockIdx.x) {
      dgamma[c] += sdzx;
      dbeta[c] += sdz;
    }
  }
  __syncthreads();
  const int c0 = (int)(v0 % cg) * 8;
  float ka[8], kb[8], kc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { ka[i] = s_a[c0 + i]; kb[i] = s_b[c0 + i]; kc[i] = s_c[c0 + i]; }
  for (long long v = v0; v < nvec; v += kUnroll * stride) {
    if (v != v0) {
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) {
        const long long vv = v + j * stride;
        if (vv < nvec) { ga[j] = ld_nc_v4(dz + vv * 8); xa[j] = ld_nc_v4(x + vv * 8); }
      }
    }
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      const long long vv = v + j * stride;
      if (vv >= nvec) break;
      V8 g = unpack8(ga[j]);
      V8 xv = unpack8(xa[j]);
#pragma unroll
      for (int i = 0; i < 8; ++i) g.v[i] = ka[i] * g.v[i] + kb[i] * xv.v[i] + kc[i];
      st_v4(dx + vv * 8, pack8(g));
    }
  }
}

}  // namespace dps

using namespace dps;

static int bn_grid(long long nvec) {
  // at least 4 vectors (one unrolled batch) per thread: small layers use few CTAs, every CTA repeats the prologue
  long long want = (nvec + 4 * kBnThreads - 1) / (4 * kBnThreads);
  static const int per_sm = [] { const char* e = getenv("DPS_BN_CTAS_PER_SM"); return e ? atoi(e) : 3; }();
  long long cap = 148 * per_sm;  // default: exactly one wave at 3 CTAs / SM (77-80 registers), no partial last wave
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

// Reduction grid: every thread gets a full unrolled batch of four rows before another CTA is added.  Each CTA ends with
// 2*C 64-bit atomics on the same 2*C addresses, so a small layer cut into 512 CTAs of four rows each (the first version)
// spent most of its 12 us in those atomics.
static int bn_reduce_grid(long long P, int C) {
  const long long rows_per_cta = (long long)(kBnThreads / (C / 8)) * 4;
  const long long want = (P + rows_per_cta - 1) / rows_per_cta;
  return (int)(want < 1 ? 1 : (want > 148 * 4 ? 148 * 4 : want));
}

DPS_API int dps_bn_forward(const void* x, const long long* stats, const float* gamma, const float* beta,
                           const void* residual, void* y, float* save_mean, float* save_rstd, float* running_mean,
                           float* running_var, long long* num_batches, long long P, int C, int relu, int training,
                           float eps, float momentum, void* stream) {
  if (C % 8 || C > kMaxC || kBnThreads % (C / 8)) DPS_FAIL(1, "bn: unsupported channel count %d", C);
  DPS_CHECK(launch_pdl(bn_fwd_kernel, dim3(bn_grid(P * (C / 8))), dim3(kBnThreads), 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)x, stats, gamma, beta, (const __nv_bfloat16*)residual, (__nv_bfloat16*)y, save_mean,
      save_rstd, running_mean, running_var, num_batches, P, C, relu, training, eps, momentum,
      1.0 / (4294967296.0 * (double)P), 1.0 / (16777216.0 * (double)P)));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

// sums must be zero on entry ([2*C] int64, fixed point x 2^40).  dz may alias dy.  When relu == 0, dz is not written and the
// apply pass should be given dy.
DPS_API int dps_bn_backward(const void* dy, const void* y, const void* x, const float* mean, const float* rstd,
                            const float* gamma, void* dz, long long* sums, void* dx, float* dgamma, float* dbeta,
                            long long P, int C, int relu, void* stream) {
  if (C % 8 || C > kMaxC || kBnThreads % (C / 8)) DPS_FAIL(1, "bn: unsupported channel count %d", C);
  const int g1 = bn_reduce_grid(P, C);
  DPS_CHECK(launch_pdl(bn_bwd_reduce_kernel, dim3(g1), dim3(kBnThreads), 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)x, mean, rstd, (__nv_bfloat16*)dz,
      sums, P, C, relu));
  DPS_CHECK(cudaGetLastError());
  DPS_CHECK(launch_pdl(bn_bwd_apply_kernel, dim3(bn_grid(P * (C / 8))), dim3(kBnThreads), 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)(relu ? dz : dy), (const __nv_bfloat16*)x, mean, rstd, gamma, sums, (__nv_bfloat16*)dx,
      dgamma, dbeta, P, C));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_bn_backward_apply(const void* dz, const void* x, const float* mean, const float* rstd,
                                  const float* gamma, const long long* sums, void* dx, float* dgamma, float* dbeta,
                                  long long P, int C, void* stream) {
  if (C % 8 || C > kMaxC || kBnThreads % (C / 8)) DPS_FAIL(1, "bn: unsupported channel count %d", C);
  DPS_CHECK(launch_pdl(bn_bwd_apply_kernel, dim3(bn_grid(P * (C / 8))), dim3(kBnThreads), 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)dz, (const __nv_bfloat16*)x, mean, rstd, gamma, sums, (__nv_bfloat16*)dx, dgamma, dbeta,
      P, C));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------
// BatchNorm backward for callers that must not allocate.
//
// dps_bn_backward() needs a [2*C] 64-bit reduction scratch from its caller.  Some callers have nowhere to
// get one: a serving or training process whose device memory was handed in full to a pool with a hard
// budget at start-up (a later cudaMalloc fails or is forbidden by policy), a latency-critical loop where the
// implicit device-wide synchronisation of cudaMalloc/cudaFree is not acceptable, or a hook / plug-in callback
// that only receives its input and output buffers.  For them the scratch lives inside the library image:
// one file-scope __device__ array of kNoAllocMaxC = 512 channels (2 * 512 64-bit words = 8 KiB), resident on every
// GPU that loads the library until the process exits.
//
// Rules that follow from sharing one array:
//   * the array keeps whatever the previous call left in it, so every call first clears the 2*C entries it
//     is about to accumulate into (a memset on `stream`, ordered before the kernels);
//   * two calls that can overlap on the GPU (different streams, or different host threads feeding different
//     streams) race on the array -- serialise them on one stream or with events;
//   * C above 512 is rejected with an error; nothing is truncated.
//
// Set-up: in host code `g_bn_noalloc_scratch` is only a stand-in for the symbol, not a device address, so
// its address is resolved with cudaGetSymbolAddress the first time the entry point runs on a device (with
// that device current), checked, cached per device and then passed to the kernels as an ordinary pointer.
// Make that first call on each device before starting any stream capture.  The per-device cache is indexed by the
// CURRENT device (cudaGetDevice): the caller must make the device that owns `stream` current before the call.
// ------------------------------------------------------------------------------------------
namespace dps {
constexpr int kNoAllocMaxC = 512;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) long long g_bn_noalloc_scratch[2 * kNoAllocMaxC];
}  // namespace dps

#include <mutex>

static int bn_noalloc_scratch_ptr(long long** out) {
  static std::mutex mu;
  static long long* resolved[kNoAllocMaxDevices] = {nullptr};
  int dev = -1;
  DPS_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) DPS_FAIL(1, "bn_backward_noalloc: device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (resolved[dev] == nullptr) {
    void* p = nullptr;
    // once per device, with that device current: the symbol has a different address on every GPU
    DPS_CHECK(cudaGetSymbolAddress(&p, g_bn_noalloc_scratch));
    if (p == nullptr) DPS_FAIL(1, "bn_backward_noalloc: cudaGetSymbolAddress returned a null address");
    resolved[dev] = static_cast<long long*>(p);
  }
  *out = resolved[dev];
  return 0;
}

// Same contract and results as dps_bn_backward, minus the `sums` argument.  Calls no cudaMalloc / cudaFree and no
// allocating library routine; the caller still passes every input and output buffer.
DPS_API int dps_bn_backward_noalloc(const void* dy, const void* y, const void* x, const float* mean,
                                    const float* rstd, const float* gamma, void* dz, void* dx, float* dgamma,
                                    float* dbeta, long long P, int C, int relu, void* stream) {
  if (C > kNoAllocMaxC)
    DPS_FAIL(1, "bn_backward_noalloc: C=%d exceeds the built-in scratch (max %d channels); use dps_bn_backward "
                "with a caller-provided [2*C] int64 buffer", C, kNoAllocMaxC);
  if (C <= 0 || C % 8 || kBnThreads % (C / 8)) DPS_FAIL(1, "bn: unsupported channel count %d", C);
  if (P <= 0) DPS_FAIL(1, "bn_backward_noalloc: P must be positive");
  long long* sums = nullptr;
  if (int e = bn_noalloc_scratch_ptr(&sums)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  // the shared array still holds the previous call's sums: clear exactly what this call accumulates into
  DPS_CHECK(cudaMemsetAsync(sums, 0, sizeof(long long) * 2 * (size_t)C, st));
  const int g1 = bn_reduce_grid(P, C);
  // plain launch: it follows a memset, not a kernel, so there is no programmatic edge to give it (and a captured graph
  // gets an ordinary memset -> kernel dependency); the apply kernel behind it still overlaps its launch with this one
  bn_bwd_reduce_kernel<<<dim3(g1), dim3(kBnThreads), 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,
                                                             (const __nv_bfloat16*)x, mean, rstd, (__nv_bfloat16*)dz, sums,
                                                             P, C, relu);
  DPS_CHECK(cudaGetLastError());
  DPS_CHECK(launch_pdl(bn_bwd_apply_kernel, dim3(bn_grid(P * (C / 8))), dim3(kBnThreads), 0, st, (const __nv_bfloat16*)(relu ? dz : dy),
                                                                  (const __nv_bfloat16*)x, mean, rstd, gamma, sums,
                                                                  (__nv_bfloat16*)dx, dgamma, dbeta, P, C));
  DPS_CHECK(cudaGetLastError());
  return 0;
}
