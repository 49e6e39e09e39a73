// Hardware probe (test-only entry point): can a tcgen05 K-major SWIZZLE_128B A operand be a *shifted window* of a
// halo tile in shared memory?  One TMA box of (BH+2) x (BW+2) pixels x 64 channels is loaded once; each of the 9
// filter taps is then addressed by moving the descriptor start by ((1+dh)*(BW+2) + (1+dw)) * 128 B with a
// stride-byte-offset of (BW+2)*128 B between 8-row groups (BW == 8, so one image row == one 8-row group).
// variant 0: base_offset field = 0;  variant 1: base_offset = (start_address >> 7) & 7.
#include "common.cuh"
// build (not part of libdps_b200.so): nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -shared -Xcompiler -fPIC \
//   -I csrc -I include -o bench_tools/bin/libexp_halo.so bench_tools/exp/exp_halo.cu
thread_local char g_dps_err[512];

namespace dps {

constexpr int kHB = 8, kHH = 16;                       // tile: 16 rows x 8 cols = 128 pixels
constexpr int kHaloRows = (kHH + 2) * (kHB + 2);       // 180 rows of 128 B
constexpr int kHaloBytes = kHaloRows * 128;            // 23040
constexpr int kHaloPad = 23552;                        // rounded to 1024

__global__ void __launch_bounds__(128) exp_halo_kernel(const __grid_constant__ CUtensorMap tmA,
                                                       const __grid_constant__ CUtensorMap tmB, float* out, int h0, int w0,
                                                       int n, int variant) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kHaloPad;                       // 9 taps x 8 KiB
  uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 9 * 8192);
  uint64_t* done = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    if (elect_one()) { mbar_init(bar, 1); mbar_init(done, 1); fence_barrier_init(); }
    __syncwarp();
    tmem_alloc(slot, 64);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar, kHaloBytes + 9 * 8192);
    tma_load_5d(sA, &tmA, bar, 0, w0 - 1, 0, h0 - 1, n);
    for (int t = 0; t < 9; ++t) tma_load_2d(sB + t * 8192, &tmB, bar, t * 64, 0);
    mbar_wait(bar, 0);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
    for (int t = 0; t < 9; ++t) {
      const int dh = t / 3 - 1, dw = t % 3 - 1;
      const uint32_t a_addr = smem_u32(sA) + ((1 + dh) * (kHB + 2) + (1 + dw)) * 128;
      uint64_t adesc = make_smem_desc(a_addr, 16, (kHB + 2) * 128);
      if (variant == 1) adesc |= (uint64_t)((a_addr >> 7) & 7) << 49;
      const uint64_t bdesc = make_smem_desc(smem_u32(sB + t * 8192), 16, 1024);
      for (int k = 0; k < 4; ++k) umma_bf16(tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (t | k) != 0);
    }
    umma_commit(done);
  }
  mbar_wait(done, 0);
  tc_fence_after();
  const int row = warp * 32 + lane_id();
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, r);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[row * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 64); }
}

}  // namespace dps
using namespace dps;

typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                          const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                          CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// x: [N][H][W][64] bf16, w: [64][9*64] bf16 (K-R-S-C), out: [128][64] fp32 = the 16x8-pixel tile at (n, h0, w0)
DPS_API int dps_exp_halo_tile(const void* x, const void* w, float* out, int N, int H, int W, int n, int h0, int w0,
                              int variant, void* stream) {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  DPS_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncFn enc = reinterpret_cast<EncFn>(fp);
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[5] = {64, (cuuint64_t)W, 1, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t st[4] = {128, (cuuint64_t)W * 128, (cuuint64_t)W * 128, (cuuint64_t)H * W * 128};
    cuuint32_t box[5] = {64, kHB + 2, 1, kHH + 2, 1}, es[5] = {1, 1, 1, 1, 1};
    if (enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(x), dims, st, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) DPS_FAIL(4, "encode A failed");
  }
  {
    cuuint64_t dims[2] = {576, 64};
    cuuint64_t st[1] = {576 * 2};
    cuuint32_t box[2] = {64, 64}, es[2] = {1, 1};
    if (enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w), dims, st, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) DPS_FAIL(4, "encode B failed");
  }
  const int smem = 1024 + kHaloPad + 9 * 8192 + 64;
  DPS_CHECK(cudaFuncSetAttribute(exp_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  exp_halo_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(tmA, tmB, out, h0, w0, n, variant);
  DPS_CHECK(cudaGetLastError());
  return 0;
}
