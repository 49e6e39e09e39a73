// Implicit-GEMM convolutions on the 5th-gen tensor cores: tcgen05.mma with TMEM accumulators,
// operands staged by TMA (cp.async.bulk.tensor) into 128-byte-swizzled shared memory.
//
// Replaces the reference's nn.Conv2d calls -- ATen/oneDNN on CPU, cuDNN in the single-process
// baseline (reference: src/parameter_server/server.py:24,26,32,48 and autograd's backward of them).
// Activations are NHWC bf16, weights K-R-S-C bf16, accumulation fp32.
//
// Two kernels cover all twenty convs of the CIFAR-style ResNet-18 in all three passes:
//
//  tapconv_kernel  (fprop and dgrad)   D[128 pixels, BLOCK_N channels] = sum over "taps" of A_tap * B_tap
//      A_tap : a 128-pixel x 64-channel box of the input, fetched by ONE 5-D tiled TMA at the tap's
//              spatial offset.  Padding is TMA out-of-bounds zero fill (coordinates may be negative), so
//              there is no im2col buffer and no boundary code.  Stride-2 convs read a "parity view"
//              tensor map (dims: [pw*C+c, w/2, ph, h/2, n]) so every tap is again a dense box.
//      B_tap : the weight slice of that tap.  fprop reads W as [Cout][(r,s,ci)] K-major; dgrad reads
//              the SAME tensor MN-major ([co rows][ci contiguous]) -- no transposed weight copy.
//      stride-2 dgrad is decomposed into the 4 output-parity classes (1/2/2/4 taps), no zero-stuffing.
//      epilogue: tcgen05.ld -> bf16 -> smem staging -> coalesced stores; optional fused residual-gradient
//              add (`addend`), optional fused per-channel sum / sum-of-squares for the following BatchNorm.
//
//  wgrad_kernel    dW[co, tap, ci] += sum over pixels X_tap[pix, ci] * dY[pix, co]
//      both operands are pixel-major in memory == MN-major UMMA operands straight from TMA boxes; the activation
//      tap is the M operand, so accumulator rows are ci and a warp's registers reduce into dW with coalesced fp32
//      red.global.add (no transpose); up to 4 taps x 128 co of fp32 accumulators live in TMEM (512 columns);
//      split over pixel ranges across CTAs.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM lane quadrant = warp_id % 4).
#include "common.cuh"
#include "../include/dps_b200.h"
#include "../include/dps_b200_compute.h"

#include <mutex>
#include <vector>

namespace dps {

constexpr int kMinStages = 3;      // 2 CTAs / SM when the grid has several waves
constexpr int kMaxStages = 6;      // 1 CTA / SM (grid <= one wave): deeper pipeline hides the L2->smem latency
constexpr int kCtrlBytes = 128;    // barriers + TMEM slot, placed right after the stage buffers
constexpr int kTileM = 128;         // pixels per CTA tile == UMMA M
constexpr int kKBlock = 64;         // reduction elements per pipeline stage (one 128-byte swizzle row)
constexpr int kABytes = kTileM * kKBlock * 2;   // 16 KiB
constexpr int kThreads = 192;
constexpr double kStatSumScale = 4294967296.0;   // 2^32  (keep in sync with bn.cu)
constexpr double kStatSqScale = 16777216.0;      // 2^24
constexpr double kGradSumScale = 1099511627776.0;  // 2^40: BN-backward sums (keep in sync with bn.cu)

struct Tap {
  int16_t c_off;   // added to the A channel coordinate (parity view: pw*C)
  int16_t db;      // A coordinate offset along w
  int16_t ph;      // A coordinate along the parity dim
  int16_t da;      // A coordinate offset along h
  int32_t k_off;   // B column offset of this tap
};

struct ConvParams {
  int BW, BH, BN;                 // pixel box of a tile (BW*BH*BN == 128), all powers of two
  int bw_sh, bhw_sh;              // log2(BW), log2(BW*BH)
  int tiles_w, tiles_h;           // tiles along w and h (tiles along n = gridDim.x / (tiles_w*tiles_h))
  int red_blocks;                 // 64-wide reduction blocks per tap
  int ngroups;                    // parity groups (gridDim.z)
  int nstages;                    // pipeline depth (kMinStages or kMaxStages)
  int tiles_total;                // halo kernel: number of 16x8-pixel tiles (persistent loop)
  int ntaps[4];
  Tap taps[4][9];
  int oph[4], opw[4];             // output parity offsets per group
  int os;                         // output pixel stride (1, or 2 for stride-2 dgrad)
  int out_H, out_W, out_C;        // output tensor dims (pixels are written at (h*os+oph, w*os+opw))
  __nv_bfloat16* out;
  const __nv_bfloat16* addend;    // optional, same shape as out
  // optional fusion of the NEXT BatchNorm's backward reduction into this (dgrad) epilogue: the tile just computed
  // is dy of that BN; with its forward output y (ReLU mask) and input x the epilogue stores dz = dy * (y > 0)
  // and accumulates sum(dz), sum(dz * xhat) per channel (64-bit fixed point x 2^40) -- bn_bwd_reduce disappears.
  const __nv_bfloat16* fuse_y;
  const __nv_bfloat16* fuse_x;
  const float* fuse_mean;
  const float* fuse_rstd;
  const float* fuse_gamma;        // optional (with fuse_beta): that BN has NO residual input, so its ReLU mask is recomputed
  const float* fuse_beta;         // from x -- y = relu(x*sc + sh) -- and y is not read at all (a quarter of the epilogue's loads)
  long long* fuse_sums;
  int addend_compact;             // addend is indexed at tile resolution (stride-2 dgrad + 1x1 shortcut gradient)
  long long* stats;               // optional [2*out_C]: per-channel sum (x 2^32) and sum of squares (x 2^24) in
                                  // 64-bit fixed point: integer atomics are exact and order-independent, so
                                  // BN statistics are bit-identical run to run (and full-rate in L2)
};

template <int BLOCK_N>
struct ConvSmem {
  static constexpr int kBBytes = BLOCK_N * kKBlock * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kPitch = BLOCK_N * 2 + 16;                 // staging row pitch (bank-conflict free)
  static constexpr int total(int nst) { return 1024 + nst * kStageBytes + kCtrlBytes; }   // 1024 = alignment slack
  static_assert(2 * kTileM * kPitch + 4096 <= kMinStages * kStageBytes, "staging must fit in the pipeline buffers");
};

// ---- explicit shared-space accessors (the staging tile is addressed with 32-bit shared addresses so the
// compiler emits LDS/STS instead of generic loads) ----
__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_v4(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}
__device__ __forceinline__ void red_fixed(long long* dst, float v, float scale) {
  // exact: power-of-two scaling in fp32, then float -> int64 (no FP64 pipe involved)
  atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)__float2ll_rn(v * scale));
}

// Epilogue shared by the conv kernels, executed by EPI_WARPS warps (threads et0 .. et0+32*EPI_WARPS, named barrier BAR;
// persistent kernels run two such groups side by side, one per TMEM accumulator):
// TMEM accumulator -> bf16 staging tile -> (BN statistics | addend + fused BN-backward reduction) -> coalesced stores.
// Warp w may only read TMEM lanes of quadrant (w & 3); with 8 warps each quadrant is split in two column halves.
template <int BLOCK_N, int EPI_WARPS, int MAX_GROUP = 4, bool MASK_X = true>
__device__ __forceinline__ void conv_epilogue(const ConvParams& p, int grp, int n0, int h0, int w0, int c_out0,
                                              uint32_t stage, uint32_t tmem_acc, uint64_t* tmem_full,
                                              uint32_t full_parity, uint64_t* tmem_empty, int et0 = 64, int bar_id = 1) {
  constexpr int T = EPI_WARPS * 32;
  constexpr int kPitch = ConvSmem<BLOCK_N>::kPitch;
  constexpr int kChunks = BLOCK_N / 8;                  // 16-byte chunks per row
  constexpr int kRowsPerPass = T / kChunks;
  constexpr int kPasses = kTileM / kRowsPerPass;
  constexpr int kGroup = kPasses < MAX_GROUP ? kPasses : MAX_GROUP;   // rows prefetched together (12 registers each)
  constexpr int kColsPerWarp = BLOCK_N / (EPI_WARPS / 4);
  constexpr int kPairs = BLOCK_N / 2;                   // reductions work on bf16 pairs
  constexpr int kParts = T / kPairs;                    // row partitions of the column reductions
  constexpr int kRowsPerPart = kTileM / kParts;
  const int et = threadIdx.x - et0;          // index inside this epilogue group
  const int ew = et >> 5, lane = et & 31;
  const int q = (threadIdx.x >> 5) & 3;                 // TMEM lane quadrant of this warp
  const int row = q * 32 + lane;                        // tile row == pixel index in the box
  const int col0 = (ew >> 2) * kColsPerWarp;
  const uint32_t stage_x = stage + kTileM * kPitch;     // second staging tile (x of the fused BN)
  const uint32_t part_s = stage + 2 * kTileM * kPitch;  // [kParts][4][kPairs] fp32 partials
  const int chunk = et % kChunks, row0 = et / kChunks;
  const bool fuse = p.fuse_y != nullptr;
  const bool mask_from_x = MASK_X && fuse && p.fuse_gamma != nullptr;   // (compiled out of the register-tight halo kernel)
  const bool use_addend = p.addend != nullptr && (!p.addend_compact || (p.oph[grp] == 0 && p.opw[grp] == 0));
  // box dims are powers of two: shifts instead of divisions
  const int bw_sh = p.bw_sh, bhw_sh = p.bhw_sh, bw_m = (1 << bw_sh) - 1, bhw_m = (1 << bhw_sh) - 1;
  auto out_off = [&](int rr) -> size_t {
    const int n = n0 + (rr >> bhw_sh), rem = rr & bhw_m;
    const int oh = (h0 + (rem >> bw_sh)) * p.os + p.oph[grp], ow = (w0 + (rem & bw_m)) * p.os + p.opw[grp];
    return (((size_t)n * p.out_H + oh) * p.out_W + ow) * p.out_C + c_out0 + chunk * 8;
  };
  auto addend_off = [&](int rr) -> size_t {
    if (!p.addend_compact) return out_off(rr);
    const int n = n0 + (rr >> bhw_sh), rem = rr & bhw_m;
    return (((size_t)n * (p.out_H / p.os) + h0 + (rem >> bw_sh)) * (p.out_W / p.os) + w0 + (rem & bw_m)) * p.out_C +
           c_out0 + chunk * 8;
  };
  // Operands that do not depend on the accumulator (residual-gradient addend, y / x of the fused BN) are fetched
  // in groups of kGroup rows, all loads of a group in flight at once; the first group is issued BEFORE waiting
  // for the MMAs so its latency hides under the mainloop.
  uint4 aa[kGroup], ya[kGroup], xa[kGroup];
  auto issue = [&](int g) {
#pragma unroll
    for (int j = 0; j < kGroup; ++j) {
      const int rr = row0 + (g * kGroup + j) * kRowsPerPass;
      if (use_addend) aa[j] = ld_nc_v4(p.addend + addend_off(rr));
      if (fuse) {
        const size_t off = out_off(rr);
        if (!mask_from_x) ya[j] = ld_nc_v4(p.fuse_y + off);
        xa[j] = ld_nc_v4(p.fuse_x + off);
      }
    }
  };
  if (use_addend || fuse) issue(0);
  // scale / shift of the fused BN for this thread's 8 channels, exactly as bn_fwd_kernel forms them
  float msc[8], msh[8];
  if (mask_from_x) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = c_out0 + chunk * 8 + i;
      msc[i] = p.fuse_gamma[c] * p.fuse_rstd[c];
      msh[i] = p.fuse_beta[c] - p.fuse_mean[c] * msc[i];
    }
  }
  mbar_wait(tmem_full, full_parity);
  tc_fence_after();
  if constexpr (kColsPerWarp >= 32) {
#pragma unroll
    for (int c = 0; c < kColsPerWarp / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + ((uint32_t)(q * 32) << 16) + col0 + c * 32, r);
      tmem_ld_wait();
      const uint32_t dst = stage + row * kPitch + (col0 + c * 32) * 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sts_v4(dst + j * 16,
               make_uint4(pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]))));
      }
    }
  } else {
    static_assert(kColsPerWarp == 16 || kColsPerWarp >= 32, "unsupported epilogue split");
    uint32_t r[16];
    tmem_ld_32x16(tmem_acc + ((uint32_t)(q * 32) << 16) + col0, r);
    tmem_ld_wait();
    const uint32_t dst = stage + row * kPitch + col0 * 2;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      sts_v4(dst + j * 16,
             make_uint4(pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]))));
    }
  }
  tc_fence_before();
  if (tmem_empty != nullptr) {             // persistent kernels: the accumulator may be overwritten from here on
    __syncwarp();
    if (lane == 0) mbar_arrive(tmem_empty);
  }
  asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(T) : "memory");      // epilogue warps only

  const int pair = et % kPairs, part = et / kPairs;
  // deterministic two-level column reduction: per-thread partial over its rows -> shared -> fixed-order sum ->
  // ONE 64-bit fixed-point atomic per channel and tile
  auto reduce_store = [&](float a0, float a1, float b0, float b1, long long* dst, int cstride, float sc0, float sc1) {
    sts_f32(part_s + ((part * 4 + 0) * kPairs + pair) * 4, a0);
    sts_f32(part_s + ((part * 4 + 1) * kPairs + pair) * 4, a1);
    sts_f32(part_s + ((part * 4 + 2) * kPairs + pair) * 4, b0);
    sts_f32(part_s + ((part * 4 + 3) * kPairs + pair) * 4, b1);
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(T) : "memory");
    for (int idx = et; idx < 2 * BLOCK_N; idx += T) {
      const int which = idx / BLOCK_N, ch = idx % BLOCK_N;
      const int k = which * 2 + (ch & 1), pr = ch >> 1;
      float acc = 0.f;
#pragma unroll
      for (int pp = 0; pp < kParts; ++pp) acc += lds_f32(part_s + ((pp * 4 + k) * kPairs + pr) * 4);
      red_fixed(dst + which * cstride + c_out0 + ch, acc, which ? sc1 : sc0);
    }
  };

  if (p.stats != nullptr) {
    // fused BatchNorm statistics over the bf16 values that are actually stored
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    const uint32_t base = stage + (part * kRowsPerPart) * kPitch + pair * 4;
#pragma unroll 8
    for (int rr = 0; rr < kRowsPerPart; ++rr) {
      const float2 v = unpack_bf16x2(lds_u32(base + rr * kPitch));
      s0 += v.x; s1 += v.y;
      q0 += v.x * v.x; q1 += v.y * v.y;
    }
    reduce_store(s0, s1, q0, q1, p.stats, p.out_C, (float)kStatSumScale, (float)kStatSqScale);
  }

#pragma unroll 1
  for (int g = 0; g < kPasses / kGroup; ++g) {
    if (g > 0) issue(g);
#pragma unroll
    for (int j = 0; j < kGroup; ++j) {
      const int rr = row0 + (g * kGroup + j) * kRowsPerPass;
      const size_t off = out_off(rr);
      uint4 v = lds_v4(stage + rr * kPitch + chunk * 16);
      if (use_addend) {
        const uint4 a = aa[j];
        float2 x, y;
        x = unpack_bf16x2(v.x); y = unpack_bf16x2(a.x); v.x = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.y); y = unpack_bf16x2(a.y); v.y = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.z); y = unpack_bf16x2(a.z); v.z = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.w); y = unpack_bf16x2(a.w); v.w = pack_bf16x2(x.x + y.x, x.y + y.y);
      }
      if (fuse) {
        if (mask_from_x) {
          // y = relu(x * sc + sh) > 0  <=>  x * sc + sh > 0: the forward output is not read
          auto maskx = [&](uint32_t gg, uint32_t xx, int i) -> uint32_t {
            const float2 xf = unpack_bf16x2(xx);
            const uint32_t lo = fmaf(xf.x, msc[i], msh[i]) > 0.f ? 0x0000FFFFu : 0u;
            const uint32_t hi = fmaf(xf.y, msc[i + 1], msh[i + 1]) > 0.f ? 0xFFFF0000u : 0u;
            return gg & (lo | hi);
          };
          v.x = maskx(v.x, xa[j].x, 0); v.y = maskx(v.y, xa[j].y, 2); v.z = maskx(v.z, xa[j].z, 4); v.w = maskx(v.w, xa[j].w, 6);
        } else {
          const uint4 yv = ya[j];
          // ReLU mask per bf16 lane: y > 0  <=>  sign bit clear and magnitude non-zero
          auto mask2 = [](uint32_t gg, uint32_t yy) -> uint32_t {
            const uint32_t lo = ((yy & 0x8000u) == 0u && (yy & 0x7FFFu) != 0u) ? 0x0000FFFFu : 0u;
            const uint32_t hi = ((yy & 0x80000000u) == 0u && (yy & 0x7FFF0000u) != 0u) ? 0xFFFF0000u : 0u;
            return gg & (lo | hi);
          };
          v.x = mask2(v.x, yv.x); v.y = mask2(v.y, yv.y); v.z = mask2(v.z, yv.z); v.w = mask2(v.w, yv.w);
        }
        sts_v4(stage + rr * kPitch + chunk * 16, v);       // dz, for the reduction below
        sts_v4(stage_x + rr * kPitch + chunk * 16, xa[j]);
      }
      st_v4(p.out + off, v);
    }
  }
  if (fuse) {
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(T) : "memory");
    const float mu0 = p.fuse_mean[c_out0 + 2 * pair], mu1 = p.fuse_mean[c_out0 + 2 * pair + 1];
    const float rs0 = p.fuse_rstd[c_out0 + 2 * pair], rs1 = p.fuse_rstd[c_out0 + 2 * pair + 1];
    float d0 = 0.f, d1 = 0.f, x0 = 0.f, x1 = 0.f;
    const uint32_t off = (part * kRowsPerPart) * kPitch + pair * 4;
#pragma unroll 8
    for (int rr = 0; rr < kRowsPerPart; ++rr) {
      const float2 g = unpack_bf16x2(lds_u32(stage + off + rr * kPitch));
      const float2 xv = unpack_bf16x2(lds_u32(stage_x + off + rr * kPitch));
      d0 += g.x; d1 += g.y;
      x0 += g.x * (xv.x - mu0) * rs0;
      x1 += g.y * (xv.y - mu1) * rs1;
    }
    reduce_store(d0, d1, x0, x1, p.fuse_sums, p.out_C, (float)kGradSumScale, (float)kGradSumScale);
  }
  if (tmem_empty != nullptr) asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(T) : "memory");   // staging is reused by the next tile
}

// Register-reduction variant of the epilogue: ONE staging tile instead of two.  The column reductions (BN statistics,
// or the fused BN-backward sums) ride along in the store loop as per-thread accumulators for the thread's 8 channels,
// are combined by a butterfly over the lanes that hold the same channels, then per warp in shared memory, then in
// fixed order across warps.  The fused path accumulates sum(dz*x) and centres / scales it once at the end
// (sum dz*xhat = rstd * (sum dz*x - mean * sum dz)).  Needs kTileM*kPitch + 8 KiB of staging: used where shared memory
// is the constraint (halo2conv_kernel keeps 144 KiB of weights resident); measured ~2 us slower per launch than
// conv_epilogue on the epilogue-bound layer-1 kernel, identical results.
template <int BLOCK_N, int EPI_WARPS>
__device__ __forceinline__ void conv_epilogue_rr(const ConvParams& p, int grp, int n0, int h0, int w0, int c_out0,
                                              uint32_t stage, uint32_t tmem_acc, uint64_t* tmem_full,
                                              uint32_t full_parity, uint64_t* tmem_empty) {
  constexpr int T = EPI_WARPS * 32;
  constexpr int kPitch = ConvSmem<BLOCK_N>::kPitch;
  constexpr int kChunks = BLOCK_N / 8;                  // 16-byte chunks per row
  constexpr int kRowsPerPass = T / kChunks;
  constexpr int kPasses = kTileM / kRowsPerPass;
  constexpr int kGroup = kPasses < 4 ? kPasses : 4;
  constexpr int kColsPerWarp = BLOCK_N / (EPI_WARPS / 4);
  const int et = threadIdx.x - 64;
  const int ew = et >> 5, lane = et & 31;
  const int q = (threadIdx.x >> 5) & 3;                 // TMEM lane quadrant of this warp
  const int row = q * 32 + lane;                        // tile row == pixel index in the box
  const int col0 = (ew >> 2) * kColsPerWarp;
  const uint32_t part_s = stage + kTileM * kPitch;      // [EPI_WARPS][2][BLOCK_N] fp32 per-warp partial sums
  const int chunk = et % kChunks, row0 = et / kChunks;  // chunk == lane % kChunks (kChunks divides 32)
  const bool fuse = p.fuse_y != nullptr;
  const bool stats = p.stats != nullptr;
  const bool use_addend = p.addend != nullptr && (!p.addend_compact || (p.oph[grp] == 0 && p.opw[grp] == 0));
  // box dims are powers of two: shifts instead of divisions
  const int bw_sh = p.bw_sh, bhw_sh = p.bhw_sh, bw_m = (1 << bw_sh) - 1, bhw_m = (1 << bhw_sh) - 1;
  auto out_off = [&](int rr) -> size_t {
    const int n = n0 + (rr >> bhw_sh), rem = rr & bhw_m;
    const int oh = (h0 + (rem >> bw_sh)) * p.os + p.oph[grp], ow = (w0 + (rem & bw_m)) * p.os + p.opw[grp];
    return (((size_t)n * p.out_H + oh) * p.out_W + ow) * p.out_C + c_out0 + chunk * 8;
  };
  auto addend_off = [&](int rr) -> size_t {
    if (!p.addend_compact) return out_off(rr);
    const int n = n0 + (rr >> bhw_sh), rem = rr & bhw_m;
    return (((size_t)n * (p.out_H / p.os) + h0 + (rem >> bw_sh)) * (p.out_W / p.os) + w0 + (rem & bw_m)) * p.out_C +
           c_out0 + chunk * 8;
  };
  // Operands that do not depend on the accumulator (residual-gradient addend, y / x of the fused BN) are fetched
  // in groups of kGroup rows, all loads of a group in flight at once; the first group is issued BEFORE waiting
  // for the MMAs so its latency hides under the mainloop.
  uint4 aa[kGroup], ya[kGroup], xa[kGroup];
  auto issue = [&](int g) {
#pragma unroll
    for (int j = 0; j < kGroup; ++j) {
      const int rr = row0 + (g * kGroup + j) * kRowsPerPass;
      if (use_addend) aa[j] = ld_nc_v4(p.addend + addend_off(rr));
      if (fuse) {
        const size_t off = out_off(rr);
        ya[j] = ld_nc_v4(p.fuse_y + off);
        xa[j] = ld_nc_v4(p.fuse_x + off);
      }
    }
  };
  if (use_addend || fuse) issue(0);
  // per-thread column accumulators for its 8 channels: (sum, sum of squares) or (sum dz, sum dz*xhat)
  // (the fused path accumulates sum dz*x and centres / scales it once at the end: sum dz*xhat = rstd*(sum dz*x - mean*sum dz))
  float acc0[8], acc1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { acc0[i] = 0.f; acc1[i] = 0.f; }
  mbar_wait(tmem_full, full_parity);
  tc_fence_after();
  if constexpr (kColsPerWarp >= 32) {
#pragma unroll
    for (int c = 0; c < kColsPerWarp / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + ((uint32_t)(q * 32) << 16) + col0 + c * 32, r);
      tmem_ld_wait();
      const uint32_t dst = stage + row * kPitch + (col0 + c * 32) * 2;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sts_v4(dst + j * 16,
               make_uint4(pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5])),
                          pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]))));
      }
    }
  } else {
    static_assert(kColsPerWarp == 16 || kColsPerWarp >= 32, "unsupported epilogue split");
    uint32_t r[16];
    tmem_ld_32x16(tmem_acc + ((uint32_t)(q * 32) << 16) + col0, r);
    tmem_ld_wait();
    const uint32_t dst = stage + row * kPitch + col0 * 2;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      sts_v4(dst + j * 16,
             make_uint4(pack_bf16x2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5])),
                        pack_bf16x2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]))));
    }
  }
  tc_fence_before();
  if (tmem_empty != nullptr) {             // persistent kernels: the accumulator may be overwritten from here on
    __syncwarp();
    if (lane == 0) mbar_arrive(tmem_empty);
  }
  asm volatile("bar.sync 1, %0;" ::"n"(T) : "memory");      // epilogue warps only

  // One pass over the tile: coalesced 16-byte stores, and the column reductions ride along in registers
  // (BN statistics for fprop; masked gradient + BN-backward sums for dgrad).
#pragma unroll 1
  for (int g = 0; g < kPasses / kGroup; ++g) {
    if (g > 0) issue(g);
#pragma unroll
    for (int j = 0; j < kGroup; ++j) {
      const int rr = row0 + (g * kGroup + j) * kRowsPerPass;
      const size_t off = out_off(rr);
      uint4 v = lds_v4(stage + rr * kPitch + chunk * 16);
      if (use_addend) {
        const uint4 a = aa[j];
        float2 x, y;
        x = unpack_bf16x2(v.x); y = unpack_bf16x2(a.x); v.x = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.y); y = unpack_bf16x2(a.y); v.y = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.z); y = unpack_bf16x2(a.z); v.z = pack_bf16x2(x.x + y.x, x.y + y.y);
        x = unpack_bf16x2(v.w); y = unpack_bf16x2(a.w); v.w = pack_bf16x2(x.x + y.x, x.y + y.y);
      }
      if (fuse) {
        const uint4 yv = ya[j];
        // ReLU mask per bf16 lane: y > 0  <=>  sign bit clear and magnitude non-zero
        auto mask2 = [](uint32_t gg, uint32_t yy) -> uint32_t {
          const uint32_t lo = ((yy & 0x8000u) == 0u && (yy & 0x7FFFu) != 0u) ? 0x0000FFFFu : 0u;
          const uint32_t hi = ((yy & 0x80000000u) == 0u && (yy & 0x7FFF0000u) != 0u) ? 0xFFFF0000u : 0u;
          return gg & (lo | hi);
        };
        v.x = mask2(v.x, yv.x); v.y = mask2(v.y, yv.y); v.z = mask2(v.z, yv.z); v.w = mask2(v.w, yv.w);
      }
      st_v4(p.out + off, v);
      if (stats || fuse) {
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
        const uint32_t xx[4] = {xa[j].x, xa[j].y, xa[j].z, xa[j].w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const float2 f = unpack_bf16x2(vv[h]);
          acc0[2 * h] += f.x; acc0[2 * h + 1] += f.y;
          if (fuse) {
            const float2 xv = unpack_bf16x2(xx[h]);
            acc1[2 * h] += f.x * xv.x;
            acc1[2 * h + 1] += f.y * xv.y;
          } else {
            acc1[2 * h] += f.x * f.x; acc1[2 * h + 1] += f.y * f.y;
          }
        }
      }
    }
  }
  if (stats || fuse) {
    // lanes l, l + kChunks, ... of a warp hold the same 8 channels (different rows): butterfly over those lanes,
    // then one partial per warp in shared memory, then a fixed-order sum over the warps -> ONE 64-bit fixed-point
    // atomic per channel and tile (deterministic).
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int o = kChunks; o < 32; o <<= 1) {
        acc0[i] += __shfl_xor_sync(0xffffffffu, acc0[i], o);
        acc1[i] += __shfl_xor_sync(0xffffffffu, acc1[i], o);
      }
    }
    if (lane < kChunks) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        sts_f32(part_s + ((ew * 2 + 0) * BLOCK_N + chunk * 8 + i) * 4, acc0[i]);
        sts_f32(part_s + ((ew * 2 + 1) * BLOCK_N + chunk * 8 + i) * 4, acc1[i]);
      }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(T) : "memory");
    long long* dst = stats ? p.stats : p.fuse_sums;
    const float sc0 = stats ? (float)kStatSumScale : (float)kGradSumScale;
    const float sc1 = stats ? (float)kStatSqScale : (float)kGradSumScale;
    for (int idx = et; idx < 2 * BLOCK_N; idx += T) {
      const int which = idx / BLOCK_N, ch = idx % BLOCK_N;
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < EPI_WARPS; ++w) acc += lds_f32(part_s + ((w * 2 + which) * BLOCK_N + ch) * 4);
      if (fuse && which) {
        float sdz = 0.f;
#pragma unroll
        for (int w = 0; w < EPI_WARPS; ++w) sdz += lds_f32(part_s + ((w * 2) * BLOCK_N + ch) * 4);
        acc = (acc - p.fuse_mean[c_out0 + ch] * sdz) * p.fuse_rstd[c_out0 + ch];
      }
      red_fixed(dst + which * p.out_C + c_out0 + ch, acc, which ? sc1 : sc0);
    }
  }
  if (tmem_empty != nullptr) asm volatile("bar.sync 1, %0;" ::"n"(T) : "memory");   // staging is reused by the next tile
}

// CL > 1: the CL CTAs of a cluster compute CL adjacent N tiles of the SAME pixel tile, so they need the same
// activation box.  Each CTA fetches 1/CL of its rows and TMA-multicasts them into all CL CTAs (L2->SM activation
// traffic / CL); the weight tile stays private.  A stage may be refilled only after ALL CL consumers released it:
// every MMA thread multicasts its tcgen05.commit to the `empty` barrier of all CTAs (barrier count = CL).
// EPI = epilogue warps (4, or 8 for one-wave grids: one CTA per SM, so the wider epilogue costs no occupancy and halves
// the serial tail of the kernel -- TMEM drain, statistics, stores -- that nothing overlaps when a CTA has ONE tile).
template <int BLOCK_N, bool B_MN, int CL, int EPI = 4>
__global__ void __launch_bounds__(64 + 32 * EPI) tapconv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const __grid_constant__ ConvParams p) {
  using S = ConvSmem<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nst = p.nstages;
  uint8_t* stages = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + nst * S::kStageBytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tmem_full = empty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int grp = blockIdx.z;
  const int ntaps = p.ntaps[grp];
  const int nkb = ntaps * p.red_blocks;

  // tile -> first pixel
  int t = blockIdx.x;
  const int tw = t % p.tiles_w; t /= p.tiles_w;
  const int th = t % p.tiles_h; t /= p.tiles_h;
  const int w0 = tw * p.BW, h0 = th * p.BH, n0 = t * p.BN;
  const int c_out0 = blockIdx.y * BLOCK_N;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < nst; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], CL); }
      mbar_init(tmem_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, BLOCK_N);        // BLOCK_N fp32 columns (64 or 128: power of two >= 32)
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();          // peers' barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();            // prologue above overlapped the previous kernel's tail; dependent data from here on
  constexpr uint16_t kClusterMask = (uint16_t)((1u << CL) - 1u);

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The issue loops of the producer and of the MMA thread are single-thread critical paths (one iteration per
    // 64-deep reduction block): no integer divisions and no parameter-space loads inside them -- stage / phase are
    // running counters and the tap is fetched once per tap.  (Measured with bench_tools/exp/ingest_probe.cu: one
    // thread sustains a TMA every ~160 cycles and an SM ingests ~100 B/clk; the division-based loops of the first
    // version needed 800-1100 cycles per block and capped every tap-loop conv at ~31 B/clk per SM.)
    if (elect_one()) {
      int s = 0;
      uint32_t ph = 0;
      const int red_blocks = p.red_blocks;
      // cluster variant: my 128/CL rows of the box (tile order: w fastest, then h, then n)
      const int q = (CL > 1) ? (int)cluster_ctarank() : 0;
      const int base = (q * (kTileM / CL)) >> p.bw_sh;              // first (h, n) line of my slice
      const int bh_sh = p.bhw_sh - p.bw_sh;
      const int hq = base & ((1 << bh_sh) - 1), nq = base >> bh_sh;
      for (int t = 0; t < ntaps; ++t) {
        const Tap tp = p.taps[grp][t];
        const int cw = w0 + tp.db, chh = h0 + tp.da + hq, cn = n0 + nq, cph = tp.ph;
        int ca = tp.c_off, ckb = tp.k_off;
        for (int cb = 0; cb < red_blocks; ++cb, ca += kKBlock, ckb += kKBlock) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* a = stages + s * S::kStageBytes;
          uint8_t* b = a + kABytes;
          mbar_expect_tx(&full[s], S::kStageBytes);
          if (CL == 1) tma_load_5d(a, &tmA, &full[s], ca, cw, cph, chh, cn);
          else         tma_load_5d_mc(a + q * (kTileM / CL) * 128, &tmA, &full[s], ca, cw, cph, chh, cn, kClusterMask);
          if (!B_MN) {
            tma_load_2d(b, &tmB, &full[s], ckb, c_out0);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_N / 64; ++j)
              tma_load_2d(b + j * 8192, &tmB, &full[s], tp.k_off + c_out0 + j * 64, cb * kKBlock);
          }
          if (++s == nst) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kTileM, BLOCK_N, 0, B_MN ? 1 : 0);
      const uint32_t stage0 = smem_u32(stages);
      // descriptors of stage 0; a stage advances the 14-bit start-address field by kStageBytes / 16
      const uint64_t adesc0 = make_smem_desc(stage0, 16, 1024);
      const uint64_t bdesc0 = B_MN ? make_smem_desc(stage0 + kABytes, 8192, 1024) : make_smem_desc(stage0 + kABytes, 16, 1024);
      int s = 0;
      uint32_t ph = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint64_t adesc = adesc0 + (uint64_t)((s * S::kStageBytes) >> 4);
        const uint64_t bdesc = bdesc0 + (uint64_t)((s * S::kStageBytes) >> 4);
#pragma unroll
        for (int k = 0; k < kKBlock / 16; ++k) {
          // advance 16 reduction elements: K-major +32 B inside the swizzle row; MN-major +16 rows of 128 B
          const uint64_t ad = adesc + (uint64_t)(k * 2);
          const uint64_t bd = bdesc + (uint64_t)(B_MN ? k * 128 : k * 2);
          umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        if (CL == 1) umma_commit(&empty[s]);            // smem slot free once these MMAs have read it
        else umma_commit_mc(&empty[s], kClusterMask);   // ... in every CTA of the cluster
        if (++s == nst) { s = 0; ph ^= 1u; }
      }
      umma_commit(tmem_full);              // accumulator complete
    }
  } else {
    // ===================== epilogue (4 warps) =====================
    conv_epilogue<BLOCK_N, EPI>(p, grp, n0, h0, w0, c_out0, smem_u32(stages), tmem_base, tmem_full, 0, nullptr);
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();          // nobody leaves while peers may still multicast into / signal this CTA
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
}

// ------------------------------------------------------------------------------------------
// pairconv_kernel: tapconv on CTA PAIRS (tcgen05 cta_group::2), 256 pixels x 128 channels per pair.
//
// The two CTAs of a cluster sit on the two SMs of a TPC.  Each loads the activation box of ITS 128-pixel tile and
// only HALF of the 128-wide weight tile (64 rows / columns); one thread of the leader CTA issues
// tcgen05.mma.cta_group::2 (M = 256), which reads A and the two B halves from both CTAs' shared memory and writes
// rows 0..127 / 128..255 of the accumulator into the TMEM of CTA 0 / CTA 1.  Per SM and reduction block that is
// 16 KB + 8 KB instead of 16 KB + 16 KB: L2->SM traffic -- the bound of these layers (profiles/launches_v5) --
// drops by a quarter against 128x128 tiles and by half against 128x64 tiles.
// Protocol: both producers wait on their own `empty` barrier and signal bytes on the LEADER's `full` barrier
// (peer bit cleared); the leader arms it with the bytes of both CTAs; tcgen05.commit.cta_group::2 multicasts the
// release of a stage (and the accumulator-ready signal) to both CTAs.
// ------------------------------------------------------------------------------------------
constexpr int kPairN = 128;
constexpr int kPairStage = kABytes + 64 * kKBlock * 2;      // 16 KiB activations + 8 KiB half weight tile

template <bool B_MN, int EPI = 4>
__global__ void __launch_bounds__(64 + 32 * EPI) pairconv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                            const __grid_constant__ CUtensorMap tmB,
                                                            const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nst = p.nstages;
  uint8_t* stages = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + nst * kPairStage);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tmem_full = empty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int grp = blockIdx.z;
  const int nkb = p.ntaps[grp] * p.red_blocks;
  const int rank = (int)cluster_ctarank();               // 0 = leader (issues the MMAs)

  int t = blockIdx.x;                                    // consecutive pixel tiles form a pair
  const int tw = t % p.tiles_w; t /= p.tiles_w;
  const int th = t % p.tiles_h; t /= p.tiles_h;
  const int w0 = tw * p.BW, h0 = th * p.BH, n0 = t * p.BN;
  const int c_out0 = blockIdx.y * kPairN;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < nst; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      mbar_init(tmem_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_2sm(tmem_slot, kPairN);                   // same warp id in both CTAs, same destination address
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                    // both CTAs' barriers and TMEM are ready
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      int s = 0;
      uint32_t ph = 0;
      const int ntaps = p.ntaps[grp], red_blocks = p.red_blocks;
      for (int t = 0; t < ntaps; ++t) {                       // division-free issue loop (see tapconv_kernel)
        const Tap tp = p.taps[grp][t];
        const int cw = w0 + tp.db, chh = h0 + tp.da, cph = tp.ph;
        for (int cb = 0; cb < red_blocks; ++cb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* a = stages + s * kPairStage;
          uint8_t* b = a + kABytes;
          if (rank == 0) mbar_expect_tx(&full[s], 2 * kPairStage);        // bytes of BOTH CTAs land on my barrier
          tma_load_5d_2sm(a, &tmA, &full[s], tp.c_off + cb * kKBlock, cw, cph, chh, n0);
          if (!B_MN) tma_load_2d_2sm(b, &tmB, &full[s], tp.k_off + cb * kKBlock, c_out0 + rank * 64);
          else       tma_load_2d_2sm(b, &tmB, &full[s], tp.k_off + c_out0 + rank * 64, cb * kKBlock);
          if (++s == nst) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(256, kPairN, 0, B_MN ? 1 : 0);
      const uint32_t stage0 = smem_u32(stages);
      const uint64_t adesc0 = make_smem_desc(stage0, 16, 1024);
      const uint64_t bdesc0 = B_MN ? make_smem_desc(stage0 + kABytes, 8192, 1024) : make_smem_desc(stage0 + kABytes, 16, 1024);
      int s = 0;
      uint32_t ph = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint64_t adesc = adesc0 + (uint64_t)((s * kPairStage) >> 4);
        const uint64_t bdesc = bdesc0 + (uint64_t)((s * kPairStage) >> 4);
#pragma unroll
        for (int k = 0; k < kKBlock / 16; ++k)
          umma_bf16_2sm(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(B_MN ? k * 128 : k * 2), idesc,
                        (kb | k) != 0 ? 1u : 0u);
        umma_commit_2sm(&empty[s], 3);       // stage free in both CTAs once these MMAs have read it
        if (++s == nst) { s = 0; ph ^= 1u; }
      }
      umma_commit_2sm(tmem_full, 3);         // accumulators of both CTAs complete
    }
  } else {
    conv_epilogue<kPairN, EPI>(p, grp, n0, h0, w0, c_out0, smem_u32(stages), tmem_base, tmem_full, 0, nullptr);
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                    // the peer may still be reading / signalling
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, kPairN);
  }
}

// ------------------------------------------------------------------------------------------
// haloconv_kernel: 3x3 stride-1 convolutions with 64 reduction channels and 64 output channels (layer1 fprop
// and dgrad), persistent, weights resident.
//
// tapconv re-reads the activation box once per filter tap and the whole weight matrix once per pixel tile: for
// layer1 that is 216 KB of L2->SM traffic per 128x64 output tile, which bounds the kernel (profiles/
// ncu_tapconv_layer1_fprop.md).  Here each CTA keeps the 9 x 64 x 64 weights (72 KB) in shared memory for its whole
// life and loads ONE halo tile of (16+2) x (8+2) pixels x 64 channels per output tile (23 KB, TMA zero-fills the
// padding).  Each filter tap is then just a *shifted window* of that halo tile: the UMMA A descriptor starts at
// ((1+dh)*10 + (1+dw)) * 128 B with a stride-byte-offset of 10 * 128 B between 8-row groups -- one image row of the
// 16x8 tile is one 8-row group.  (The 128-byte swizzle is a function of absolute shared-memory address bits, so
// windows that are not 1024-byte aligned read correctly with base_offset = 0; probed in bench_tools/exp/exp_halo.cu.)
// Two 64-column TMEM accumulators and TWO epilogue groups of 8 warps (own staging tile, own named barrier): group g
// drains accumulator g, so the epilogues of tiles i and i+1 are in flight together and overlap the MMAs of tile i+2.
// The per-tile epilogue is a chain of latencies (TMEM load, staging round trip, three block barriers, global
// loads / atomics), ~2 us per tile; one group left the tensor pipe idle two thirds of the time.
// ------------------------------------------------------------------------------------------
constexpr int kHaloW = 8, kHaloH = 16;
constexpr int kHaloRowsW = kHaloW + 2;
constexpr int kHaloBytes = (kHaloH + 2) * kHaloRowsW * 128;      // 23040
constexpr int kHaloSlot = 23552;                                  // rounded up to 1024
constexpr int kHaloSlots = 2;
constexpr int kHaloWBytes = 9 * 8192;
constexpr int kHaloEpiWarps = 16;                                         // halo2conv_kernel: one group of 16
constexpr int kHaloGroupWarps = 8;                                        // haloconv_kernel: two groups of 8
constexpr int kHaloThreads = 64 + 32 * kHaloEpiWarps;
constexpr int kHaloStaging = 2 * kTileM * ConvSmem<64>::kPitch + 8192;   // dz + x tiles + reduction partials
constexpr int kHaloSmem = 1024 + kHaloWBytes + kHaloSlots * kHaloSlot + 2 * kHaloStaging + 256;
static_assert(kHaloSmem <= 227 * 1024, "haloconv_kernel shared memory");

template <bool B_MN>
__global__ void __launch_bounds__(kHaloThreads, 1) haloconv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem;
  uint8_t* sA = sW + kHaloWBytes;
  uint8_t* staging = sA + kHaloSlots * kHaloSlot;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(staging + 2 * kHaloStaging);     // 9 barriers: one per tap
  uint64_t* fullA = wfull + 9;
  uint64_t* emptyA = fullA + kHaloSlots;
  uint64_t* tfull = emptyA + kHaloSlots;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int t = 0; t < 9; ++t) mbar_init(&wfull[t], 1);
      for (int s = 0; s < kHaloSlots; ++s) { mbar_init(&fullA[s], 1); mbar_init(&emptyA[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], kHaloGroupWarps); }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 128);            // two 64-column accumulators
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  const int tiles_per_img = p.tiles_w * p.tiles_h;
  if (warp == 0) {
    if (elect_one()) {
      // weights: one barrier per tap, so the first MMAs start after 8 KiB instead of 72 KiB; the first halo tile is
      // requested right after tap 0
      for (int t = 0; t < 9; ++t) {
        mbar_expect_tx(&wfull[t], 8192);
        tma_load_2d(sW + t * 8192, &tmB, &wfull[t], t * 64, 0);     // fprop: [64 co rows][64 k]; dgrad: [64 co(k) rows][64 ci]
        if (t == 0 && (int)blockIdx.x < p.tiles_total) {
          const int tt = blockIdx.x;
          const int n = tt / tiles_per_img, r = tt % tiles_per_img;
          mbar_expect_tx(&fullA[0], kHaloBytes);
          tma_load_5d(sA, &tmA, &fullA[0], 0, (r % p.tiles_w) * kHaloW - 1, 0, (r / p.tiles_w) * kHaloH - 1, n);
        }
      }
      int i = 1;
      for (int t = blockIdx.x + gridDim.x; t < p.tiles_total; t += gridDim.x, ++i) {
        const int s = i % kHaloSlots, ph = (i / kHaloSlots) & 1;
        const int n = t / tiles_per_img, r = t % tiles_per_img;
        const int h0 = (r / p.tiles_w) * kHaloH, w0 = (r % p.tiles_w) * kHaloW;
        mbar_wait(&emptyA[s], ph ^ 1);
        mbar_expect_tx(&fullA[s], kHaloBytes);
        tma_load_5d(sA + s * kHaloSlot, &tmA, &fullA[s], 0, w0 - 1, 0, h0 - 1, n);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kTileM, 64, 0, B_MN ? 1 : 0);
      const uint32_t w_base = smem_u32(sW);
      int i = 0;
      for (int t = blockIdx.x; t < p.tiles_total; t += gridDim.x, ++i) {
        const int s = i % kHaloSlots, ph = (i / kHaloSlots) & 1;
        const int acc = i & 1;
        mbar_wait(&tempty[acc], ((i >> 1) & 1) ^ 1);          // epilogue group `acc` has drained this accumulator
        mbar_wait(&fullA[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sA + s * kHaloSlot);
#pragma unroll
        for (int tp = 0; tp < 9; ++tp) {
          // tap tp = (r, s): fprop reads the window shifted by (r-1, s-1), dgrad by (1-r, 1-s) -- compile-time offsets
          const int r3 = tp / 3, s3 = tp % 3;
          const int wrow = B_MN ? (2 - r3) * kHaloRowsW + (2 - s3) : r3 * kHaloRowsW + s3;
          if (i == 0) { mbar_wait(&wfull[tp], 0); tc_fence_after(); }
          const uint64_t adesc = make_smem_desc(a_base + wrow * 128, 16, kHaloRowsW * 128);
          const uint32_t b_addr = w_base + tp * 8192;
          const uint64_t bdesc = B_MN ? make_smem_desc(b_addr, 8192, 1024) : make_smem_desc(b_addr, 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + acc * 64, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(B_MN ? k * 128 : k * 2), idesc,
                      (tp | k) != 0 ? 1u : 0u);
        }
        umma_commit(&emptyA[s]);
        umma_commit(&tfull[acc]);
      }
    }
  } else {
    // two epilogue groups of 8 warps; group g serves accumulator g == tiles with (i & 1) == g
    const int g = (warp - 2) / kHaloGroupWarps;
    const uint32_t my_staging = smem_u32(staging) + g * kHaloStaging;
    int i = g;
    for (int t = blockIdx.x + g * gridDim.x; t < p.tiles_total; t += 2 * gridDim.x, i += 2) {
      const int n = t / tiles_per_img, r = t % tiles_per_img;
      const int h0 = (r / p.tiles_w) * kHaloH, w0 = (r % p.tiles_w) * kHaloW;
      conv_epilogue<64, kHaloGroupWarps, 2, false>(p, 0, n, h0, w0, 0, my_staging, tmem_base + g * 64, &tfull[g], (i >> 1) & 1,
                                         &tempty[g], 64 + g * 32 * kHaloGroupWarps, 1 + g);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}
// ------------------------------------------------------------------------------------------
// halo2conv_kernel: the halo scheme for 3x3 stride-1 convolutions with 128 reduction channels (layer2: 16x16
// images, 128 -> 128 channels; fprop and dgrad).  The output channels are cut into 64-wide slices (gridDim.y); a CTA
// serves ONE slice for its whole life and keeps that slice's weights resident: 9 taps x 2 reduction blocks x 8 KiB =
// 144 KiB.  Per 128-pixel tile it loads two halo tiles (one per 64-channel reduction block, 23 KiB each) into a
// two-slot ring -- the MMAs of block 1 overlap the load of the next tile's block 0 -- and issues 9 x 2 x 4 MMAs into
// one of two TMEM accumulators.  L2->SM traffic per 128x128 outputs: 2 x 46 KiB instead of 576 KiB for the tap loop
// (these convs are bound by the per-SM ingest rate, profiles/conv_bound_analysis.md).
// Shared memory: 144 KiB weights + 46 KiB ring + 26 KiB staging -> the register-reduction epilogue (one staging tile).
// ------------------------------------------------------------------------------------------
constexpr int kH2RB = 2;                                          // reduction blocks (128 channels)
constexpr int kH2Slots = 2;
constexpr int kH2WBytes = 9 * kH2RB * 8192;                       // 147456
constexpr int kH2Staging = kTileM * ConvSmem<64>::kPitch + 8192;  // one output tile + per-warp partials
constexpr int kH2Smem = 1024 + kH2WBytes + kH2Slots * kHaloSlot + kH2Staging + 256;
static_assert(kH2Smem <= 227 * 1024, "halo2conv_kernel shared memory");

template <bool B_MN>
__global__ void __launch_bounds__(kHaloThreads, 1) halo2conv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem;
  uint8_t* sA = sW + kH2WBytes;
  uint8_t* staging = sA + kH2Slots * kHaloSlot;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(staging + kH2Staging);
  uint64_t* fullA = wfull + 1;
  uint64_t* emptyA = fullA + kH2Slots;
  uint64_t* tfull = emptyA + kH2Slots;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int c_out0 = blockIdx.y * 64;                  // this CTA's 64-wide slice of the GEMM N dimension
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (elect_one()) {
      mbar_init(wfull, 1);
      for (int s = 0; s < kH2Slots; ++s) { mbar_init(&fullA[s], 1); mbar_init(&emptyA[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], kHaloEpiWarps); }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 128);            // two 64-column accumulators
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  const int tiles_per_img = p.tiles_w * p.tiles_h;
  if (warp == 0) {
    if (elect_one()) {
      // resident weights: chunk (tap, cb) = [64 rows][64 cols] bf16, 128-byte swizzled
      mbar_expect_tx(wfull, kH2WBytes);
      for (int t = 0; t < 9; ++t)
        for (int cb = 0; cb < kH2RB; ++cb) {
          uint8_t* dst = sW + (t * kH2RB + cb) * 8192;
          if (!B_MN) tma_load_2d(dst, &tmB, wfull, p.taps[0][t].k_off + cb * kKBlock, c_out0);   // rows = co slice, cols = k
          else       tma_load_2d(dst, &tmB, wfull, p.taps[0][t].k_off + c_out0, cb * kKBlock);   // rows = k (co), cols = ci slice
        }
      int i = 0;
      for (int t = blockIdx.x; t < p.tiles_total; t += gridDim.x, ++i) {
        const int n = t / tiles_per_img, r = t % tiles_per_img;
        const int h0 = (r / p.tiles_w) * kHaloH, w0 = (r % p.tiles_w) * kHaloW;
        for (int cb = 0; cb < kH2RB; ++cb) {
          const int j = i * kH2RB + cb;
          const int s = j % kH2Slots, ph = (j / kH2Slots) & 1;
          mbar_wait(&emptyA[s], ph ^ 1);
          mbar_expect_tx(&fullA[s], kHaloBytes);
          tma_load_5d(sA + s * kHaloSlot, &tmA, &fullA[s], cb * kKBlock, w0 - 1, 0, h0 - 1, n);
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kTileM, 64, 0, B_MN ? 1 : 0);
      mbar_wait(wfull, 0);
      int i = 0;
      for (int t = blockIdx.x; t < p.tiles_total; t += gridDim.x, ++i) {
        const int acc = i & 1;
        mbar_wait(&tempty[acc], ((i >> 1) & 1) ^ 1);          // epilogue has drained this accumulator
        for (int cb = 0; cb < kH2RB; ++cb) {
          const int j = i * kH2RB + cb;
          const int s = j % kH2Slots, ph = (j / kH2Slots) & 1;
          mbar_wait(&fullA[s], ph);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + s * kHaloSlot);
#pragma unroll
          for (int tp = 0; tp < 9; ++tp) {
            const Tap tap = p.taps[0][tp];
            const uint32_t a_addr = a_base + ((1 + tap.da) * kHaloRowsW + (1 + tap.db)) * 128;
            const uint64_t adesc = make_smem_desc(a_addr, 16, kHaloRowsW * 128);
            const uint32_t b_addr = smem_u32(sW + (tp * kH2RB + cb) * 8192);
            const uint64_t bdesc = B_MN ? make_smem_desc(b_addr, 8192, 1024) : make_smem_desc(b_addr, 16, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(tmem_base + acc * 64, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(B_MN ? k * 128 : k * 2),
                        idesc, (cb | tp | k) != 0 ? 1u : 0u);
          }
          umma_commit(&emptyA[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else {
    int i = 0;
    for (int t = blockIdx.x; t < p.tiles_total; t += gridDim.x, ++i) {
      const int n = t / tiles_per_img, r = t % tiles_per_img;
      const int h0 = (r / p.tiles_w) * kHaloH, w0 = (r % p.tiles_w) * kHaloW;
      const int acc = i & 1;
      conv_epilogue_rr<64, kHaloEpiWarps>(p, 0, n, h0, w0, c_out0, smem_u32(staging), tmem_base + acc * 64, &tfull[acc],
                                          (i >> 1) & 1, &tempty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// ------------------------------------------------------------------------------------------
// wgrad
// ------------------------------------------------------------------------------------------
struct WgradParams {
  int BW, BH, BN;                 // pixel box of one reduction block (BW*BH*BN == 64)
  int tiles_w, tiles_h, tiles_n;  // reduction blocks along w, h, n
  int nsplit;                     // CTAs along the pixel range (gridDim.x)
  int co_tiles;                   // ceil(Cout/128)
  int ci_chunks;                  // Cin / NB
  int co_alias;                   // 1 when Cout == 64: rows 64..127 alias rows 0..63 (LBO = 0)
  int ntap_groups;
  int ntaps[4];
  Tap taps[4][9];                 // A(X) coordinate offsets + dW column offset (tap*Cin)
  int nstages, stage_bytes;       // pipeline depth / bytes per stage (dY tile + the group's tap tiles)
  int Cout, dw_pitch;             // dW row pitch in floats (9*Cin, or 32 for the stem)
  int valid_cols;                 // columns of the ci chunk that exist in dW (stem: 27)
  float* dw;
};

template <int NB>
struct WgradSmem {
  static constexpr int kDyBytes = 64 * 128 * 2;                   // [64 pix][128 co] as two 64-wide atoms
  static constexpr int kXBytes = 64 * NB * 2;                     // one tap: [64 pix][NB ci]
  static constexpr int kMaxTaps = 512 / NB;                       // TMEM columns
  static constexpr int kBudget = 200 * 1024;                      // stage buffers per CTA
  static constexpr int total(int nst, int stage_bytes) { return 1024 + nst * stage_bytes + kCtrlBytes; }
};

constexpr int kWgradThreads = 64 + 32 * 8;     // TMA warp, MMA warp, EIGHT reduction warps (one CTA per SM: 512 TMEM columns)

template <int NB>
__global__ void __launch_bounds__(kWgradThreads) wgrad_kernel(const __grid_constant__ CUtensorMap tmDy,
                                                         const __grid_constant__ CUtensorMap tmX,
                                                         const __grid_constant__ WgradParams p) {
  using S = WgradSmem<NB>;
  const int kWS = p.nstages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stages = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)kWS * p.stage_bytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tmem_full = empty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  // blockIdx.y enumerates (co_tile, ci_chunk, tap_group)
  int y = blockIdx.y;
  const int tg = y % p.ntap_groups; y /= p.ntap_groups;
  const int cic = y % p.ci_chunks;  y /= p.ci_chunks;
  const int cot = y;
  const int ntaps = p.ntaps[tg];
  const int total_blocks = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (total_blocks + p.nsplit - 1) / p.nsplit;
  const int kb0 = blockIdx.x * per;
  const int kb1 = min(total_blocks, kb0 + per);
  const int nkb = max(0, kb1 - kb0);
  const uint32_t dy_atom_bytes = 64 * 128;                         // one [64 pix][64 co] atom
  const uint32_t stage_tx = (p.co_alias ? dy_atom_bytes : 2 * dy_atom_bytes) + ntaps * S::kXBytes;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmDy);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < kWS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      mbar_init(tmem_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();            // prologue above overlapped the previous kernel's tail; dependent data from here on

  if (warp == 0) {
    if (elect_one()) {
      // division-free issue loop: (tw, th, tn) and (stage, phase) are running counters; the taps of this group live
      // in registers (at most 512 / NB of them)
      int t = kb0;
      int tw = t % p.tiles_w; t /= p.tiles_w;
      int th = t % p.tiles_h; t /= p.tiles_h;
      int tn = t;
      const int tiles_w = p.tiles_w, tiles_h = p.tiles_h, BW = p.BW, BH = p.BH, BN = p.BN;
      const int stage_bytes = p.stage_bytes, co_alias = p.co_alias;
      int tc[S::kMaxTaps], tdb[S::kMaxTaps], tph[S::kMaxTaps], tda[S::kMaxTaps];
#pragma unroll
      for (int tp = 0; tp < S::kMaxTaps; ++tp) {
        const Tap tap = p.taps[tg][tp < ntaps ? tp : 0];
        tc[tp] = tap.c_off + cic * NB; tdb[tp] = tap.db; tph[tp] = tap.ph; tda[tp] = tap.da;
      }
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&empty[s], ph ^ 1);
        const int w0 = tw * BW, h0 = th * BH, n0 = tn * BN;
        uint8_t* dy = stages + (size_t)s * stage_bytes;
        uint8_t* x = dy + S::kDyBytes;
        mbar_expect_tx(&full[s], stage_tx);
        tma_load_5d(dy, &tmDy, &full[s], cot * 128, w0, 0, h0, n0);
        if (!co_alias) tma_load_5d(dy + dy_atom_bytes, &tmDy, &full[s], cot * 128 + 64, w0, 0, h0, n0);
#pragma unroll
        for (int tp = 0; tp < S::kMaxTaps; ++tp) {
          if (tp < ntaps) {
#pragma unroll
            for (int j = 0; j < NB / 64; ++j)
              tma_load_5d(x + tp * S::kXBytes + j * 8192, &tmX, &full[s], tc[tp] + j * 64, w0 + tdb[tp], tph[tp],
                          h0 + tda[tp], n0);
          }
        }
        if (++s == kWS) { s = 0; ph ^= 1u; }
        if (++tw == tiles_w) { tw = 0; if (++th == tiles_h) { th = 0; ++tn; } }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // D[(ci)][co] = X_tap^T * dY: the activation tap is the M operand (rows = ci; with 64 input channels rows
      // 64..127 alias rows 0..63 through LBO = 0), dY the N operand.  Rows of the accumulator are then ci, so a warp's
      // 32 lanes hold 32 consecutive ci of one co and the split-K reduction is a coalesced fp32 red.global.add straight
      // from registers (no shared-memory transpose, no block barriers in the epilogue) -- the scheme of wgrad_halo_kernel.
      const int nco = p.co_alias ? 64 : 128;                           // accumulator columns per tap
      const uint32_t idesc = p.co_alias ? make_idesc_bf16(128, 64, 1, 1) : make_idesc_bf16(128, 128, 1, 1);
      const uint32_t stage0 = smem_u32(stages);
      const uint32_t stage_bytes = (uint32_t)p.stage_bytes;
      const uint64_t adesc0 = make_smem_desc(stage0 + S::kDyBytes, NB == 128 ? 8192 : 0, 1024);     // X tap
      const uint64_t bdesc0 = make_smem_desc(stage0, dy_atom_bytes, 1024);                          // dY
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint64_t soff = (uint64_t)((s * stage_bytes) >> 4);
        const uint64_t bdesc = bdesc0 + soff;
        for (int tp = 0; tp < ntaps; ++tp) {
          const uint64_t adesc = adesc0 + soff + (uint64_t)((tp * S::kXBytes) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)          // 64 pixels = 4 x UMMA_K(16); +16 rows of 128 B per step
            umma_bf16(tmem_base + tp * nco, adesc + (uint64_t)(k * 128), bdesc + (uint64_t)(k * 128), idesc,
                      (i | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty[s]);
        if (++s == kWS) { s = 0; ph ^= 1u; }
      }
      umma_commit(tmem_full);
    }
  } else {
    const int q = warp & 3;
    const int lane = lane_id();
    if (nkb > 0 && q * 32 < NB) {                          // NB == 64: accumulator rows 64..127 are aliases
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      const int nco = p.co_alias ? 64 : 128;
      const int ci = q * 32 + lane;
      const bool row_ok = cic * NB + ci < p.valid_cols;    // stem: only 27 of the 64 padded columns exist in dW
      const int half = (warp - 2) >> 2;                    // warps 2..5 take the even 32-column chunks, 6..9 the odd ones
      for (int tp = 0; tp < ntaps; ++tp) {
        float* dst = p.dw + (size_t)(cot * 128) * p.dw_pitch + p.taps[tg][tp].k_off + cic * NB + ci;
        for (int c = half; c < nco / 32; c += 2) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + tp * nco + c * 32, r);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(dst + (size_t)(c * 32 + j) * p.dw_pitch, __uint_as_float(r[j]));
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------
// wgrad_halo_kernel: weight gradient of 3x3 stride-1 convolutions (H, W multiples of 8).
//
// wgrad_kernel re-reads a shifted activation box for every filter tap: 48-64 KB of L2->SM traffic per 64-pixel
// reduction block, which bounds it.  Here one reduction block is an 8x8-pixel patch of one image:
//   B operand = dY patch   [64 pix][64 co]  MN-major (dense TMA box),
//   A operand = TWO filter taps at once: the activation HALO patch (10x10 pixels x 64 ci, loaded ONCE for all nine
//               taps) is addressed as an MN-major M=128 operand whose two 64-wide atoms are the shifted windows of
//               taps t and t+1 -- the descriptor's leading-byte-offset is simply the distance between the two
//               windows, and the stride-byte-offset (10*128 B) steps from one image row of the patch to the next.
// D[(tap pair, ci)][co] accumulates in five 64-column TMEM regions; 20.8 KB of traffic feed 9 taps x 64 x 64 x 64
// MACs (4.7x less than before).  The accumulator rows are (tap, ci), so a warp's 32 lanes hold 32 consecutive ci of
// one co: the split-K reduction is a coalesced fp32 red.global.add straight from registers, no transpose.
// ------------------------------------------------------------------------------------------
struct WgradHaloParams {
  int tiles_w, tiles_h, tiles_n;  // 8x8 patches along w, h; images
  int nsplit;                     // CTAs along the patch range (gridDim.x)
  int ci_chunks;                  // Cin / 64   (gridDim.y = (Cout/64) * ci_chunks)
  int Cin, dw_pitch;
  int nstages;
  float* dw;
};
constexpr int kWhDyBytes = 64 * 128;                   // [64 pix][64 co]
constexpr int kWhXBytes = 100 * 128;                   // 10 x 10 halo pixels x 64 ci
constexpr int kWhStage = kWhDyBytes + 13312;           // X slot rounded up to 1024
constexpr int kWhMaxStages = 8;

__global__ void __launch_bounds__(kWgradThreads) wgrad_halo_kernel(const __grid_constant__ CUtensorMap tmDy,
                                                              const __grid_constant__ CUtensorMap tmX,
                                                              const __grid_constant__ WgradHaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nst = p.nstages;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + nst * kWhStage);
  uint64_t* empty = full + kWhMaxStages;
  uint64_t* tmem_full = empty + kWhMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int cot = blockIdx.y / p.ci_chunks, cic = blockIdx.y % p.ci_chunks;
  const int total = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (total + p.nsplit - 1) / p.nsplit;
  const int kb0 = blockIdx.x * per, kb1 = min(total, kb0 + per);
  const int nkb = max(0, kb1 - kb0);

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmDy);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < nst; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      mbar_init(tmem_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      int t = kb0;                                            // division-free issue loop (see tapconv_kernel)
      int tw = t % p.tiles_w; t /= p.tiles_w;
      int th = t % p.tiles_h; t /= p.tiles_h;
      int tn = t;
      const int tiles_w = p.tiles_w, tiles_h = p.tiles_h;
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&empty[s], ph ^ 1);
        uint8_t* dy = smem + s * kWhStage;
        mbar_expect_tx(&full[s], kWhDyBytes + kWhXBytes);
        tma_load_5d(dy, &tmDy, &full[s], cot * 64, tw * 8, 0, th * 8, tn);
        tma_load_5d(dy + kWhDyBytes, &tmX, &full[s], cic * 64, tw * 8 - 1, 0, th * 8 - 1, tn);
        if (++s == nst) { s = 0; ph ^= 1u; }
        if (++tw == tiles_w) { tw = 0; if (++th == tiles_h) { th = 0; ++tn; } }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);      // both operands MN-major
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t dy_addr = smem_u32(smem) + s * kWhStage;
        const uint32_t x_addr = dy_addr + kWhDyBytes;
        const uint64_t bdesc = make_smem_desc(dy_addr, 8192, 1024);
#pragma unroll
        for (int pr = 0; pr < 5; ++pr) {
          const int ta = 2 * pr, tb = (2 * pr + 1 < 9) ? 2 * pr + 1 : 8;
          const int offa = (ta / 3) * 10 + ta % 3, offb = (tb / 3) * 10 + tb % 3;   // window origin in halo rows
          const uint64_t adesc = make_smem_desc(x_addr + offa * 128, (offb - offa) * 128, 10 * 128);
#pragma unroll
          for (int k = 0; k < 4; ++k)          // 16 pixels = 2 patch rows: A +2*1280 B, B +16*128 B
            umma_bf16(tmem_base + pr * 64, adesc + (uint64_t)(k * 160), bdesc + (uint64_t)(k * 128), idesc,
                      (i | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty[s]);
        if (++s == nst) { s = 0; ph ^= 1u; }
      }
      umma_commit(tmem_full);
    }
  } else if (nkb > 0) {
    const int q = warp & 3;
    const int lane = lane_id();
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int ci = (q & 1) * 32 + lane;                       // rows 0..63 -> first tap of the pair, 64..127 -> second
#pragma unroll 1
    for (int pr = 0; pr < 5; ++pr) {
      const int tap = 2 * pr + (q >> 1);
      if (tap > 8) continue;                                  // second half of the padded fifth pair
      float* dst = p.dw + (size_t)(cot * 64) * p.dw_pitch + tap * p.Cin + cic * 64 + ci;
      {
        const int c = (warp - 2) >> 2;                        // eight reduction warps: 2..5 -> columns 0..31, 6..9 -> 32..63
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + pr * 64 + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) atomicAdd(dst + (size_t)(c * 32 + j) * p.dw_pitch, __uint_as_float(r[j]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------
// host side: tensor maps + plans
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 5-D activation map over an NHWC bf16 tensor [N][H][W][C].
//   parity == 0: dims (C, W, 1, H, N)
//   parity == 1: dims (2C, W/2, 2, H/2, N)   (x[n, 2a+ph, 2b+pw, c] at coordinate (pw*C+c, b, ph, a, n))
static int make_act_map(CUtensorMap* m, const void* base, int N, int H, int W, int C, int parity,
                        int box_c, int bw, int bh, int bn) {
  EncodeTiledFn enc = get_encode();
  if (!enc) DPS_FAIL(3, "cuTensorMapEncodeTiled not available");
  cuuint64_t dims[5], strides[4];
  const cuuint64_t e = 2;
  if (!parity) {
    dims[0] = (cuuint64_t)C; dims[1] = W; dims[2] = 1; dims[3] = H; dims[4] = N;
    strides[0] = (cuuint64_t)C * e;            // w
    strides[1] = (cuuint64_t)W * C * e;        // (ph) -- size 1, value irrelevant but must be valid
    strides[2] = (cuuint64_t)W * C * e;        // h
    strides[3] = (cuuint64_t)H * W * C * e;    // n
  } else {
    dims[0] = (cuuint64_t)2 * C; dims[1] = W / 2; dims[2] = 2; dims[3] = H / 2; dims[4] = N;
    strides[0] = (cuuint64_t)2 * C * e;        // b
    strides[1] = (cuuint64_t)W * C * e;        // ph
    strides[2] = (cuuint64_t)2 * W * C * e;    // a
    strides[3] = (cuuint64_t)H * W * C * e;    // n
  }
  cuuint32_t box[5] = {(cuuint32_t)box_c, (cuuint32_t)bw, 1, (cuuint32_t)bh, (cuuint32_t)bn};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) DPS_FAIL(4, "cuTensorMapEncodeTiled(act) failed: %d (N%d H%d W%d C%d par%d box %d,%d,%d,%d)",
                                  (int)r, N, H, W, C, parity, box_c, bw, bh, bn);
  return 0;
}

// 2-D weight map [rows][cols] bf16 with row pitch `pitch` elements.
static int make_w_map(CUtensorMap* m, const void* base, int rows, int cols, int pitch, int box_cols, int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) DPS_FAIL(3, "cuTensorMapEncodeTiled not available");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) DPS_FAIL(4, "cuTensorMapEncodeTiled(weight) failed: %d (rows %d cols %d pitch %d)", (int)r,
                                  rows, cols, pitch);
  return 0;
}

// choose a pixel box (bw, bh, bn) with bw*bh*bn == pixels, bw | W, bh | H, bn | N, preferring wide rows
static bool pick_box(int pixels, int N, int H, int W, int* bw, int* bh, int* bn) {
  for (int w = pixels; w >= 1; w >>= 1) {
    if (w > W || W % w) continue;
    for (int h = pixels / w; h >= 1; h >>= 1) {
      if (h > H || H % h) continue;
      const int n = pixels / (w * h);
      if (w * h * n != pixels || n > N || N % n) continue;
      *bw = w; *bh = h; *bn = n;
      return true;
    }
  }
  return false;
}

struct ConvPlan {
  int kind;                       // 0 fprop, 1 dgrad, 2 wgrad
  int block_n;
  bool b_mn;
  bool halo;                      // persistent halo kernel (layer1-type convs)
  bool halo2;                     // halo2conv_kernel (128 reduction channels, 64-wide resident weight slices)
  int cluster;                    // CTAs per cluster along N (activation multicast); 1 = none
  bool pair;                      // pairconv_kernel (cta_group::2)
  bool epi8;                      // one-wave grid: 8 epilogue warps
  CUtensorMap tmA, tmB;
  ConvParams cp;
  WgradParams wp;
  WgradHaloParams whp;
  bool whalo;                     // wgrad_halo_kernel
  dim3 grid;
  size_t smem;
};

template <int BN_, bool MN_, int CL_ = 1, int EPI_ = 4>
static int launch_tapconv(const ConvPlan& pl, cudaStream_t st) {
  auto kern = tapconv_kernel<BN_, MN_, CL_, EPI_>;
  static bool configured = false;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   ConvSmem<BN_>::total(kMaxStages)));
    configured = true;
  }
  DPS_CHECK(launch_pdl_cluster(kern, pl.grid, dim3(64 + 32 * EPI_), ConvSmem<BN_>::total(pl.cp.nstages), st,
                               dim3(1, CL_, 1), pl.tmA, pl.tmB, pl.cp));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

template <bool MN_, int EPI_ = 4>
static int launch_pair(const ConvPlan& pl, cudaStream_t st) {
  auto kern = pairconv_kernel<MN_, EPI_>;
  static bool configured = false;
  const int smem_max = 1024 + kMaxStages * kPairStage + kCtrlBytes;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    configured = true;
  }
  DPS_CHECK(launch_pdl_cluster(kern, pl.grid, dim3(64 + 32 * EPI_), 1024 + pl.cp.nstages * kPairStage + kCtrlBytes, st,
                               dim3(2, 1, 1), pl.tmA, pl.tmB, pl.cp));
  return 0;
}

template <bool MN_>
static int launch_halo(const ConvPlan& pl, cudaStream_t st) {
  auto kern = haloconv_kernel<MN_>;
  static bool configured = false;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kHaloSmem));
    configured = true;
  }
  DPS_CHECK(launch_pdl(kern, pl.grid, dim3(kHaloThreads), kHaloSmem, st, pl.tmA, pl.tmB, pl.cp));
  return 0;
}

template <bool MN_>
static int launch_halo2(const ConvPlan& pl, cudaStream_t st) {
  auto kern = halo2conv_kernel<MN_>;
  static bool configured = false;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kH2Smem));
    configured = true;
  }
  DPS_CHECK(launch_pdl(kern, pl.grid, dim3(kHaloThreads), kH2Smem, st, pl.tmA, pl.tmB, pl.cp));
  return 0;
}

static int launch_wgrad_halo(const ConvPlan& pl, cudaStream_t st) {
  static bool configured = false;
  const int smem = 1024 + kWhMaxStages * kWhStage + 256;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  DPS_CHECK(launch_pdl(wgrad_halo_kernel, pl.grid, dim3(kWgradThreads), 1024 + pl.whp.nstages * kWhStage + 256, st,
                       pl.tmA, pl.tmB, pl.whp));
  return 0;
}

template <int NB_>
static int launch_wgrad(const ConvPlan& pl, cudaStream_t st) {
  auto kern = wgrad_kernel<NB_>;
  static bool configured = false;
  if (!configured) {
    DPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   WgradSmem<NB_>::total(1, WgradSmem<NB_>::kBudget)));
    configured = true;
  }
  DPS_CHECK(launch_pdl(kern, pl.grid, dim3(kWgradThreads), WgradSmem<NB_>::total(pl.wp.nstages, pl.wp.stage_bytes), st,
                       pl.tmA, pl.tmB, pl.wp));
  DPS_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace dps

using namespace dps;

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
// fprop:  x [N][H][W][Cin] -> out [N][H/stride][W/stride][Cout];  w [Cout][R*R*Cin] (row pitch w_pitch)
// dgrad:  x is dY [N][H/stride][W/stride][Cout] -> out dX [N][H][W][Cin] (H, W are ALWAYS the conv input dims)
DPS_API int dps_conv_plan_create(int kind, const void* x, const void* w, void* out, const void* addend,
                                 long long* stats, int N, int H, int W, int Cin, int Cout, int R, int stride,
                                 int w_cols, int w_pitch, void** out_plan) {
  if (kind != 0 && kind != 1) DPS_FAIL(1, "kind must be 0 (fprop) or 1 (dgrad)");
  if (R != 1 && R != 3) DPS_FAIL(1, "R must be 1 or 3");
  if (stride != 1 && stride != 2) DPS_FAIL(1, "stride must be 1 or 2");
  if (H % stride || W % stride) DPS_FAIL(1, "H, W must be divisible by stride");
  const int pad = R / 2;
  const int Ho = H / stride, Wo = W / stride;
  const int red = (kind == 0) ? Cin : Cout;        // reduction channels
  const int ngemm = (kind == 0) ? Cout : Cin;      // GEMM N
  if (red % 64) DPS_FAIL(1, "reduction channels (%d) must be a multiple of 64", red);
  if (ngemm % 64) DPS_FAIL(1, "output channels (%d) must be a multiple of 64", ngemm);
  ConvPlan* pl = new ConvPlan();
  memset(&pl->cp, 0, sizeof(pl->cp));
  pl->kind = kind;
  pl->block_n = (ngemm % 128 == 0) ? 128 : 64;
  {
    // small layers (few pixel tiles): prefer 64-wide N tiles so the grid covers more of the 148 SMs
    const long long tiles = (long long)N * Ho * Wo / kTileM;
    const int groups = (kind == 1 && stride == 2 && R == 3) ? 4 : 1;
    if (pl->block_n == 128 && tiles * (ngemm / 128) * groups < 100) pl->block_n = 64;
    // CTA pairs (cta_group::2): 256 x 128 pair tiles when N is a multiple of 128 and the pixel tiles pair up
    pl->pair = false;
    {
      const char* ep = getenv("DPS_PAIR");
      // measured: neutral for the 256/512-wide layers, slightly slower than plain 128x128 tiles at N = 128
      if (!(ep && ep[0] == '0') && ngemm % 128 == 0 && ngemm >= 256 && tiles % 2 == 0) pl->pair = true;
    }
    // wide layers: 64-wide N tiles in clusters of 4 (or 2) that share the activation box by TMA multicast
    pl->cluster = 1;
    const char* e = getenv("DPS_CLUSTER");
    if (pl->pair) {
      pl->block_n = 128;
    } else if (!(e && e[0] == '0') && ngemm >= 256) {
      pl->block_n = 64;
      const int ny = ngemm / 64;
      pl->cluster = (ny % 4 == 0) ? 4 : (ny % 2 == 0 ? 2 : 1);
    }
  }
  pl->b_mn = (kind == 1);
  ConvParams& cp = pl->cp;
  cp.red_blocks = red / 64;
  // pixel tiles run over the conv OUTPUT grid for fprop, and over the dY-resolution grid for dgrad
  // (stride 1: same as dX; stride 2: one tile of (a, b) per output-parity class)
  const int TH = Ho, TW = Wo;
  {
    const char* e = getenv("DPS_HALO");
    pl->halo = !(e && e[0] == '0') && R == 3 && stride == 1 && red == 64 && ngemm == 64 && H % kHaloH == 0 &&
               W % kHaloW == 0;
  }
  {
    // 128 reduction channels: resident 64-wide weight slices (DPS_HALO2=1 enables; validated shapes: 16x16 images)
    const char* e = getenv("DPS_HALO2");
    pl->halo2 = (e && e[0] == '1') && !pl->halo && R == 3 && stride == 1 && red == 64 * kH2RB && ngemm <= 148 * 64 &&
                H % kHaloH == 0 && W % kHaloW == 0;
  }
  const bool any_halo = pl->halo || pl->halo2;
  if (any_halo) {
    cp.BW = kHaloW; cp.BH = kHaloH; cp.BN = 1;
  } else if (!pick_box(kTileM, N, TH, TW, &cp.BW, &cp.BH, &cp.BN)) {
    delete pl;
    DPS_FAIL(1, "no 128-pixel box divides N=%d H=%d W=%d", N, TH, TW);
  }
  if (any_halo) { pl->cluster = 1; pl->pair = false; }
  if (pl->cluster > 1) {
    // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
    // memory hung the GPU: the cluster variant (activation box multicast) is only chosen for a buffer that lives on
    // the current device; anything else (a peer mapping) takes the plain per-CTA loads.
    cudaPointerAttributes at;
    int cur = -1;
    const bool local = cudaGetDevice(&cur) == cudaSuccess && cudaPointerGetAttributes(&at, x) == cudaSuccess &&
                       at.type == cudaMemoryTypeDevice && at.device == cur;
    if (!local) { cudaGetLastError(); pl->cluster = 1; }
  }
  if (pl->halo2) pl->block_n = 64;
  while (pl->cluster > 1 && (kTileM / pl->cluster) % cp.BW != 0) pl->cluster /= 2;   // a slice is whole box lines
  // each CTA of a cluster loads 128/cluster rows of the box: sub-box = BW x sub_bh x sub_bn
  const int lines = kTileM / pl->cluster / cp.BW;
  const int sub_bh = lines < cp.BH ? lines : cp.BH;
  const int sub_bn = lines / sub_bh;
  for (cp.bw_sh = 0; (1 << cp.bw_sh) < cp.BW; ++cp.bw_sh) {}
  for (cp.bhw_sh = 0; (1 << cp.bhw_sh) < cp.BW * cp.BH; ++cp.bhw_sh) {}
  cp.tiles_w = TW / cp.BW;
  cp.tiles_h = TH / cp.BH;
  const int tiles_n = N / cp.BN;
  cp.out = reinterpret_cast<__nv_bfloat16*>(out);
  cp.addend = reinterpret_cast<const __nv_bfloat16*>(addend);
  cp.stats = stats;
  int rc = 0;
  if (kind == 0) {
    cp.out_H = Ho; cp.out_W = Wo; cp.out_C = Cout; cp.os = 1;
    cp.ngroups = 1;
    int nt = 0;
    for (int r = 0; r < R; ++r)
      for (int s = 0; s < R; ++s) {
        Tap t;
        if (stride == 1) {
          t.c_off = 0; t.ph = 0; t.da = (int16_t)(r - pad); t.db = (int16_t)(s - pad);
        } else {
          // ih = 2*oh + r - pad  ->  parity ph = (r - pad) & 1, a = oh + floor((r - pad) / 2)
          const int dr = r - pad, ds = s - pad;
          const int ph = ((dr % 2) + 2) % 2, pw = ((ds % 2) + 2) % 2;
          t.ph = (int16_t)ph; t.da = (int16_t)((dr - ph) / 2);
          t.c_off = (int16_t)(pw * Cin); t.db = (int16_t)((ds - pw) / 2);
        }
        t.k_off = (r * R + s) * Cin;
        cp.taps[0][nt++] = t;
      }
    cp.ntaps[0] = nt;
    rc = make_act_map(&pl->tmA, x, N, H, W, Cin, stride == 2, 64, cp.BW + (any_halo ? 2 : 0), sub_bh + (any_halo ? 2 : 0),
                      sub_bn);
    if (!rc) rc = make_w_map(&pl->tmB, w, Cout, w_cols, w_pitch, 64, pl->pair ? 64 : pl->block_n);  // pair: half tile per CTA
  } else {
    cp.out_H = H; cp.out_W = W; cp.out_C = Cin; cp.os = stride;
    if (stride == 1) {
      cp.ngroups = 1;
      int nt = 0;
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < R; ++s) {
          Tap t;
          t.c_off = 0; t.ph = 0; t.da = (int16_t)(pad - r); t.db = (int16_t)(pad - s);
          t.k_off = (r * R + s) * Cin;
          cp.taps[0][nt++] = t;
        }
      cp.ntaps[0] = nt;
    } else {
      // dX[n, 2a+ph, 2b+pw] = sum over (r, s) with (ph + pad - r), (pw + pad - s) even of
      //                       dY[n, a + (ph+pad-r)/2, b + (pw+pad-s)/2] * W[:, r, s, :]
      int g = 0;
      for (int ph = 0; ph < 2; ++ph)
        for (int pw = 0; pw < 2; ++pw) {
          int nt = 0;
          for (int r = 0; r < R; ++r)
            for (int s = 0; s < R; ++s) {
              const int dr = ph + pad - r, ds = pw + pad - s;
              if ((dr & 1) || (ds & 1)) continue;
              Tap t;
              t.c_off = 0; t.ph = 0; t.da = (int16_t)(dr / 2); t.db = (int16_t)(ds / 2);
              t.k_off = (r * R + s) * Cin;
              cp.taps[g][nt++] = t;
            }
          if (nt == 0) continue;               // 1x1 stride-2: only parity (0,0) receives gradient
          cp.ntaps[g] = nt; cp.oph[g] = ph; cp.opw[g] = pw;
          ++g;
        }
      cp.ngroups = g;
    }
    rc = make_act_map(&pl->tmA, x, N, Ho, Wo, Cout, 0, 64, cp.BW + (any_halo ? 2 : 0), sub_bh + (any_halo ? 2 : 0),
                      sub_bn);
    if (!rc) rc = make_w_map(&pl->tmB, w, Cout, w_cols, w_pitch, 64, 64);
  }
  if (rc) { delete pl; return rc; }
  pl->grid = dim3(cp.tiles_w * cp.tiles_h * tiles_n, ngemm / pl->block_n, cp.ngroups);
  static const int one_wave_stages = [] { const char* e = getenv("DPS_STAGES1"); const int v = e ? atoi(e) : kMaxStages;
                                         return v < kMinStages ? kMinStages : (v > kMaxStages ? kMaxStages : v); }();
  cp.nstages = ((long long)pl->grid.x * pl->grid.y * pl->grid.z <= 148) ? one_wave_stages : kMinStages;
  if (pl->pair) {
    const long long ctas = (long long)pl->grid.x * pl->grid.y * pl->grid.z;
    cp.nstages = ctas <= 148 ? (one_wave_stages < 4 ? 4 : one_wave_stages) : 4;          // staging needs >= 3 stages of 24 KiB
  }
  {
    const char* e = getenv("DPS_EPI8");
    pl->epi8 = !(e && e[0] == '0') && cp.nstages == kMaxStages && pl->cluster == 1;
  }
  if (pl->halo) {
    cp.tiles_total = cp.tiles_w * cp.tiles_h * tiles_n;
    pl->grid = dim3(cp.tiles_total < 148 ? cp.tiles_total : 148, 1, 1);      // persistent: one CTA per SM
  }
  if (pl->halo2) {
    cp.tiles_total = cp.tiles_w * cp.tiles_h * tiles_n;
    const int slices = ngemm / 64, per = 148 / slices;                         // one CTA per SM, `per` CTAs per slice
    pl->grid = dim3(cp.tiles_total < per ? cp.tiles_total : per, slices, 1);
  }
  *out_plan = pl;
  return 0;
}

// wgrad: dw[Cout][dw_pitch] (fp32, accumulated with atomics: caller zeroes) += dY^T * im2col(X)
DPS_API int dps_wgrad_plan_create(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout,
                                  int R, int stride, int valid_cols, int dw_pitch, int nsplit, void** out_plan) {
  if (R != 1 && R != 3) DPS_FAIL(1, "R must be 1 or 3");
  if (Cin % 64 || Cout % 64) DPS_FAIL(1, "channels must be multiples of 64");
  const int pad = R / 2;
  const int Ho = H / stride, Wo = W / stride;
  ConvPlan* pl = new ConvPlan();
  memset(&pl->wp, 0, sizeof(pl->wp));
  pl->kind = 2;
  pl->whalo = false;
  {
    const char* e = getenv("DPS_WHALO");
    if (!(e && e[0] == '0') && R == 3 && stride == 1 && H % 8 == 0 && W % 8 == 0 && valid_cols == Cin &&
        dw_pitch == 9 * Cin) {
      WgradHaloParams& hp = pl->whp;
      hp.tiles_w = W / 8; hp.tiles_h = H / 8; hp.tiles_n = N;
      hp.ci_chunks = Cin / 64; hp.Cin = Cin; hp.dw_pitch = dw_pitch; hp.dw = dw;
      static const int wh_stages = [] { const char* e = getenv("DPS_WH_STAGES"); const int v = e ? atoi(e) : kWhMaxStages;
                                        return v < 2 ? 2 : (v > kWhMaxStages ? kWhMaxStages : v); }();
      hp.nstages = wh_stages;
      const int ny = (Cout / 64) * hp.ci_chunks;
      const int total = hp.tiles_w * hp.tiles_h * hp.tiles_n;
      int ns = nsplit > 0 ? nsplit : (148 / ny > 0 ? 148 / ny : 1);
      if (ns > total) ns = total;
      hp.nsplit = ns;
      int rc = make_act_map(&pl->tmA, dy, N, H, W, Cout, 0, 64, 8, 8, 1);
      if (!rc) rc = make_act_map(&pl->tmB, x, N, H, W, Cin, 0, 64, 10, 10, 1);
      if (rc) { delete pl; return rc; }
      pl->whalo = true;
      pl->block_n = 64;
      pl->grid = dim3(ns, ny, 1);
      *out_plan = pl;
      return 0;
    }
  }
  WgradParams& wp = pl->wp;
  const int NB = (Cin % 128 == 0) ? 128 : 64;
  pl->block_n = NB;
  if (!pick_box(64, N, Ho, Wo, &wp.BW, &wp.BH, &wp.BN)) {
    delete pl;
    DPS_FAIL(1, "no 64-pixel box divides N=%d H=%d W=%d", N, Ho, Wo);
  }
  wp.tiles_w = Wo / wp.BW; wp.tiles_h = Ho / wp.BH; wp.tiles_n = N / wp.BN;
  wp.co_alias = (Cout == 64);
  wp.co_tiles = (Cout + 127) / 128;
  if (!wp.co_alias && Cout % 128) { delete pl; DPS_FAIL(1, "Cout must be 64 or a multiple of 128"); }
  wp.ci_chunks = Cin / NB;
  wp.Cout = Cout; wp.dw_pitch = dw_pitch; wp.dw = dw;
  wp.valid_cols = valid_cols;
  const int nco = (Cout == 64) ? 64 : 128;                    // accumulator columns per tap (TMEM: 512 in all)
  const int max_taps = (512 / nco) < (512 / NB) ? (512 / nco) : (512 / NB);
  const int total_taps = R * R;
  // balanced tap groups
  const int ngroups = (total_taps + max_taps - 1) / max_taps;
  wp.ntap_groups = ngroups;
  int idx = 0;
  for (int g = 0; g < ngroups; ++g) {
    const int cnt = total_taps / ngroups + (g < total_taps % ngroups ? 1 : 0);
    wp.ntaps[g] = cnt;
    for (int i = 0; i < cnt; ++i, ++idx) {
      const int r = idx / R, s = idx % R;
      Tap t;
      if (stride == 1) {
        t.c_off = 0; t.ph = 0; t.da = (int16_t)(r - pad); t.db = (int16_t)(s - pad);
      } else {
        const int dr = r - pad, ds = s - pad;
        const int ph = ((dr % 2) + 2) % 2, pw = ((ds % 2) + 2) % 2;
        t.ph = (int16_t)ph; t.da = (int16_t)((dr - ph) / 2);
        t.c_off = (int16_t)(pw * Cin); t.db = (int16_t)((ds - pw) / 2);
      }
      t.k_off = idx * Cin;
      wp.taps[g][i] = t;
    }
  }
  const int total_blocks = wp.tiles_w * wp.tiles_h * wp.tiles_n;
  const int ny = wp.co_tiles * wp.ci_chunks * ngroups;
  // one CTA per SM (512 TMEM columns each): keep the whole grid inside ONE wave of 148 CTAs
  if (nsplit <= 0) nsplit = 148 / ny > 0 ? 148 / ny : 1;
  if (nsplit > total_blocks) nsplit = total_blocks;
  wp.nsplit = nsplit;
  {
    int max_group = 0;
    for (int g = 0; g < ngroups; ++g) max_group = wp.ntaps[g] > max_group ? wp.ntaps[g] : max_group;
    wp.stage_bytes = 64 * 128 * 2 + max_group * 64 * NB * 2;
    int nst = (200 * 1024) / wp.stage_bytes;
    wp.nstages = nst > kMaxStages ? kMaxStages : (nst < 2 ? 2 : nst);
  }
  int rc = make_act_map(&pl->tmA, dy, N, Ho, Wo, Cout, 0, 64, wp.BW, wp.BH, wp.BN);
  if (!rc) rc = make_act_map(&pl->tmB, x, N, H, W, Cin, stride == 2, 64, wp.BW, wp.BH, wp.BN);
  if (rc) { delete pl; return rc; }
  pl->grid = dim3(nsplit, ny, 1);
  *out_plan = pl;
  return 0;
}

// Fuse the following BatchNorm's backward reduction into a dgrad plan (see ConvParams::fuse_*).  sums: [2*C] int64,
// zero on entry, fixed point x 2^40.  y / x: forward output and input of that BN, same shape as the plan's output.
DPS_API int dps_conv_plan_fuse_bn_bwd(void* plan, const void* y, const void* x, const float* mean, const float* rstd,
                                      long long* sums) {
  ConvPlan* pl = reinterpret_cast<ConvPlan*>(plan);
  if (pl->kind != 1) DPS_FAIL(1, "fuse_bn_bwd: only dgrad plans");
  pl->cp.fuse_y = reinterpret_cast<const __nv_bfloat16*>(y);
  pl->cp.fuse_x = reinterpret_cast<const __nv_bfloat16*>(x);
  pl->cp.fuse_mean = mean;
  pl->cp.fuse_rstd = rstd;
  pl->cp.fuse_sums = sums;
  pl->cp.fuse_gamma = pl->cp.fuse_beta = nullptr;
  return 0;
}

// The fused BatchNorm has no residual input and is followed by ReLU: its mask can be recomputed from x
// (y = relu(x * gamma*rstd + beta - mean*gamma*rstd)), so the epilogue stops reading y.
DPS_API int dps_conv_plan_fuse_bn_mask_from_x(void* plan, const float* gamma, const float* beta) {
  ConvPlan* pl = reinterpret_cast<ConvPlan*>(plan);
  if (pl->kind != 1 || pl->cp.fuse_y == nullptr) DPS_FAIL(1, "fuse_bn_mask_from_x: call dps_conv_plan_fuse_bn_bwd first");
  if (pl->halo2 || pl->halo) return 0;             // those epilogues keep reading y (register budget of the persistent kernels)
  pl->cp.fuse_gamma = gamma;
  pl->cp.fuse_beta = beta;
  return 0;
}

// Stride-2 dgrad only: `addend` is [N][H/2][W/2][C] and is added to the output pixels of parity (0,0)
// (the gradient coming back through a 1x1 stride-2 shortcut convolution).
DPS_API int dps_conv_plan_addend_compact(void* plan, int on) {
  ConvPlan* pl = reinterpret_cast<ConvPlan*>(plan);
  if (pl->kind != 1 || pl->cp.os != 2) DPS_FAIL(1, "addend_compact: only stride-2 dgrad plans");
  pl->cp.addend_compact = on;
  return 0;
}

DPS_API int dps_conv_plan_run(void* plan, void* stream) {
  ConvPlan* pl = reinterpret_cast<ConvPlan*>(plan);
  cudaStream_t st = (cudaStream_t)stream;
  if (pl->kind == 2 && pl->whalo) return launch_wgrad_halo(*pl, st);
  if (pl->kind == 2) return pl->block_n == 128 ? launch_wgrad<128>(*pl, st) : launch_wgrad<64>(*pl, st);
  if (pl->halo) return pl->b_mn ? launch_halo<true>(*pl, st) : launch_halo<false>(*pl, st);
  if (pl->halo2) return pl->b_mn ? launch_halo2<true>(*pl, st) : launch_halo2<false>(*pl, st);
  if (pl->pair && pl->epi8) return pl->b_mn ? launch_pair<true, 8>(*pl, st) : launch_pair<false, 8>(*pl, st);
  if (pl->pair) return pl->b_mn ? launch_pair<true>(*pl, st) : launch_pair<false>(*pl, st);
  if (pl->block_n == 64 && pl->cluster == 4)
    return pl->b_mn ? launch_tapconv<64, true, 4>(*pl, st) : launch_tapconv<64, false, 4>(*pl, st);
  if (pl->block_n == 64 && pl->cluster == 2)
    return pl->b_mn ? launch_tapconv<64, true, 2>(*pl, st) : launch_tapconv<64, false, 2>(*pl, st);
  if (pl->epi8 && pl->cluster == 1) {
    if (pl->b_mn) return pl->block_n == 128 ? launch_tapconv<128, true, 1, 8>(*pl, st) : launch_tapconv<64, true, 1, 8>(*pl, st);
    return pl->block_n == 128 ? launch_tapconv<128, false, 1, 8>(*pl, st) : launch_tapconv<64, false, 1, 8>(*pl, st);
  }
  if (pl->b_mn) return pl->block_n == 128 ? launch_tapconv<128, true>(*pl, st) : launch_tapconv<64, true>(*pl, st);
  return pl->block_n == 128 ? launch_tapconv<128, false>(*pl, st) : launch_tapconv<64, false>(*pl, st);
}

DPS_API int dps_conv_plan_destroy(void* plan) {
  delete reinterpret_cast<ConvPlan*>(plan);
  return 0;
}

DPS_API int dps_conv_plan_info(void* plan, int* grid_xyz, int* block_n, int* smem_bytes) {
  ConvPlan* pl = reinterpret_cast<ConvPlan*>(plan);
  grid_xyz[0] = pl->grid.x; grid_xyz[1] = pl->grid.y; grid_xyz[2] = pl->grid.z;
  *block_n = pl->block_n;
  *smem_bytes = (pl->kind == 2 && pl->whalo) ? 1024 + pl->whp.nstages * kWhStage + 256
                : pl->kind == 2 ? WgradSmem<64>::total(pl->wp.nstages, pl->wp.stage_bytes)
                : pl->halo   ? kHaloSmem
                : pl->halo2  ? kH2Smem
                              : (pl->block_n == 128 ? ConvSmem<128>::total(pl->cp.nstages)
                                                    : ConvSmem<64>::total(pl->cp.nstages));
  return 0;
}
