// Hardware probe (not part of the product library): how fast can ONE SM ingest L2-resident data into shared
// memory with TMA, as a function of the box size, the number of issuing threads and the CTAs per SM?
// `nthr` elected threads (one per warp) each own a ring of S (power of two) stages and issue `iters` loads, waiting
// only when they lap their ring; nothing consumes the data.  The issue loop has no divisions and no parameter loads.
//   variant 0: 2-D tiled TMA, SWIZZLE_128B, box 64 x ROWS (ROWS*128 bytes), row pitch 9216 B   (GEMM K-major operand)
//   variant 2: 5-D tiled TMA like the layer-4 conv activations  (C=512, W=4, H=4, box 64 x 4 x 1 x 4 x 8 = 16 KiB)
//   variant 3: 5-D tiled TMA like the layer-2 conv activations  (C=128, W=16, H=16, box 64 x 16 x 1 x 8 x 1 = 16 KiB)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc -I include -o bench_tools/bin/ingest_probe bench_tools/exp/ingest_probe.cu -lcuda
#include "common.cuh"
#include <vector>
#include <cstring>
thread_local char g_dps_err[512];
using namespace dps;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn enc() {
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  return (EncodeTiledFn)p;
}

template <int VARIANT, int LOG_S, int NTHR, int NBOX = 1>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm, int iters, int stage_bytes,
                                             int rows, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int S = 1 << LOG_S;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + NTHR * S * stage_bytes);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NTHR * S; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tm);
  }
  __syncthreads();
  if (warp < NTHR && elect_one()) {
    const long long t0 = clock64();
    uint8_t* base = smem + warp * S * stage_bytes;
    uint64_t* bar = full + warp * S;
    int c0 = ((blockIdx.x * 5 + warp * 3) & 63) * 64, r0 = 0;
    for (int i = 0; i < iters; ++i) {
      const int s = i & (S - 1);
      if (i >= S) mbar_wait(&bar[s], ((i >> LOG_S) - 1) & 1);
      mbar_expect_tx(&bar[s], stage_bytes);
      uint8_t* dst = base + s * stage_bytes;
      if (VARIANT == 0) {
#pragma unroll
        for (int b = 0; b < NBOX; ++b) tma_load_2d(dst + b * (stage_bytes / NBOX), &tm, &bar[s], c0, r0 + b * (rows / NBOX));
      }
      if (VARIANT == 2) tma_load_5d(dst, &tm, &bar[s], c0 & 511, -1 + (i & 1), 0, -1 + ((i >> 1) & 1), r0 & 127);
      if (VARIANT == 3) tma_load_5d(dst, &tm, &bar[s], c0 & 127, -1 + (i & 1), 0, (i & 8), r0 & 127);
      c0 += 64;
      if (c0 >= 4608) { c0 = 0; r0 = (r0 + rows) & 2047; }
    }
    for (int i = iters - S; i < iters; ++i) mbar_wait(&bar[i & (S - 1)], (i >> LOG_S) & 1);
    const long long t1 = clock64();
    if (warp == 0) cycles[blockIdx.x] = t1 - t0;
  }
}

template <int VARIANT, int LOG_S, int NTHR, int NBOX = 1>
static void run(const CUtensorMap& tm, int stage_bytes, int rows, long long* cyc) {
  auto k = probe<VARIANT, LOG_S, NTHR, NBOX>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int S = 1 << LOG_S;
  const size_t smem = 1024 + NTHR * S * stage_bytes + 512;
  if (smem > 220 * 1024) return;
  for (int grid : {64, 148, 296}) {
    if (grid == 296 && smem > 110 * 1024) continue;
    const int iters = (32 << 20) / stage_bytes / NTHR;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<<<grid, 128, smem>>>(tm, iters, stage_bytes, rows, cyc);
    cudaEventRecord(a);
    k<<<grid, 128, smem>>>(tm, iters, stage_bytes, rows, cyc);
    cudaEventRecord(b);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("variant %d: %s\n", VARIANT, cudaGetErrorString(err)); exit(1); }
    float ms; cudaEventElapsedTime(&ms, a, b);
    std::vector<long long> h(grid); cudaMemcpy(h.data(), cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    long long mx = 0; for (auto v : h) mx = v > mx ? v : mx;
    const double per_cta = (double)iters * NTHR * stage_bytes;
    const int per_sm = grid > 148 ? 2 : 1;
    printf("v%d nbox %d stage %2d KB  thr %d  S %2d  grid %3d : %7.1f us  %6.1f B/clk/SM  %5.0f cyc/TMA  %8.0f GB/s\n", VARIANT, NBOX,
           stage_bytes / 1024, NTHR, S, grid, ms * 1e3, per_cta * per_sm / (double)mx, (double)mx / iters,
           per_cta * grid / (ms * 1e-3) / 1e9);
  }
}

int main() {
  const size_t bytes = 64u << 20;
  uint8_t* src; cudaMalloc(&src, bytes); cudaMemset(src, 1, bytes);
  long long* cyc; cudaMalloc(&cyc, 296 * sizeof(long long));
  EncodeTiledFn e = enc();
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  for (int nbox : {1, 2, 4, 8}) {
    const int rows = 128;                       // 16 KiB stage split into nbox boxes of rows/nbox rows
    CUtensorMap tm; memset(&tm, 0, sizeof(tm));
    cuuint64_t dims[2] = {4608, 4096}; cuuint64_t st[1] = {4608 * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(rows / nbox)};
    CUresult r = e(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, src, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    if (nbox == 1) run<0, 2, 1, 1>(tm, rows * 128, rows, cyc);
    if (nbox == 2) run<0, 2, 1, 2>(tm, rows * 128, rows, cyc);
    if (nbox == 4) run<0, 2, 1, 4>(tm, rows * 128, rows, cyc);
    if (nbox == 8) run<0, 2, 1, 8>(tm, rows * 128, rows, cyc);
  }
  return 0;
}
