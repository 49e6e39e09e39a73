// Device-resident input pipeline: gather + RandomCrop(pad 4) + RandomHorizontalFlip + ToTensor + Normalize in one
// launch, reading the uint8 dataset that lives in HBM and writing the engine's fp32 NCHW input batch.
//
// Replaces torchvision's transform stack inside DataLoader worker processes
// (reference: src/workers/worker.py:146-158 transforms, :181-187 DataLoader(num_workers=2)).  CIFAR-100 is 150 MB as
// uint8: on a 180 GB part it is resident for the whole job, so a training step needs NO host work and NO host->device
// copy for its inputs -- with 8 ranks on one host the CPU transform stack, not the GPU, was the bound (measured:
// 84 k img/s with one rank, 31 k img/s per rank with eight).
//
// Randomness is counter-based (splitmix64 of seed and the global sample counter), so a batch is a pure function of
// (seed, counter0, indices): reproducible, checkable against a host restatement, identical under graph replay or
// eager launch.
#include "common.cuh"
#include "../include/dps_b200_compute.h"

namespace dps {

__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  unsigned long long z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

struct AugNorm {
  float scale[3];   // 1 / (255 * std)
  float shift[3];   // -mean / std
};

// One thread per output pixel (all three channels): reads 3 bytes, writes 3 floats (coalesced along w per channel).
__global__ void __launch_bounds__(256) augment_batch_kernel(const unsigned char* __restrict__ src,        // [n,S,S,3]
                                                            const long long* __restrict__ src_labels,     // [n]
                                                            const long long* __restrict__ idx,            // [B]
                                                            float* __restrict__ out,                      // [B,3,S,S]
                                                            long long* __restrict__ out_labels,           // [B]
                                                            int B, int S, int n_valid, unsigned long long seed,
                                                            unsigned long long counter0, int train, int pad,
                                                            AugNorm nm) {
  const int per_img = S * S;
  const long long total = (long long)B * per_img;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(t / per_img), pix = (int)(t % per_img);
    const int h = pix / S, w = pix % S;
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;                 // rows past n_valid: zeros (+ label -1), never counted
    if (b < n_valid) {
      const long long s = idx[b];
      int dy = pad, dx = pad, flip = 0;
      if (train) {
        const unsigned long long z = splitmix64(seed * 0x100000001B3ull + counter0 + (unsigned long long)b);
        dy = (int)(z % (unsigned long long)(2 * pad + 1));
        dx = (int)((z >> 16) % (unsigned long long)(2 * pad + 1));
        flip = (int)((z >> 40) & 1ull);
      }
      const int wc = flip ? (S - 1 - w) : w;            // crop first, then mirror the cropped image
      const int sh = h + dy - pad, sw = wc + dx - pad;
      float p0 = 0.f, p1 = 0.f, p2 = 0.f;               // zero padding is applied BEFORE Normalize: pad -> -mean/std
      if (sh >= 0 && sh < S && sw >= 0 && sw < S) {
        const unsigned char* px = src + ((size_t)s * per_img + (size_t)sh * S + sw) * 3;
        p0 = (float)px[0]; p1 = (float)px[1]; p2 = (float)px[2];
      }
      v0 = fmaf(p0, nm.scale[0], nm.shift[0]);
      v1 = fmaf(p1, nm.scale[1], nm.shift[1]);
      v2 = fmaf(p2, nm.scale[2], nm.shift[2]);
      if (pix == 0) out_labels[b] = src_labels[s];
    } else if (pix == 0) {
      out_labels[b] = -1;
    }
    float* o = out + (size_t)b * 3 * per_img + pix;
    o[0] = v0; o[per_img] = v1; o[2 * per_img] = v2;
  }
}

}  // namespace dps

DPS_API int dps_augment_batch(const unsigned char* src, const long long* src_labels, const long long* idx, float* out,
                              long long* out_labels, int B, int S, int n_valid, unsigned long long seed,
                              unsigned long long counter0, int train, int pad, const float* mean3, const float* std3,
                              void* stream) {
  if (B <= 0 || S <= 0 || n_valid < 0 || n_valid > B || pad < 0 || pad > 16)
    DPS_FAIL(1, "augment_batch: bad arguments B=%d S=%d n_valid=%d pad=%d", B, S, n_valid, pad);
  dps::AugNorm nm;
  for (int c = 0; c < 3; ++c) {
    nm.scale[c] = 1.0f / (255.0f * std3[c]);
    nm.shift[c] = -mean3[c] / std3[c];
  }
  const long long total = (long long)B * S * S;
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  dps::augment_batch_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, src_labels, idx, out, out_labels, B, S, n_valid,
                                                                    seed, counter0, train, pad, nm);
  DPS_CHECK(cudaGetLastError());
  return 0;
}
