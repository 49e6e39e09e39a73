// Fabric: symmetric arenas shared between the per-GPU processes of one box through CUDA IPC.
//
// Replaces the transport role of the reference's gRPC runtime on the box (reference:
// src/parameter_server/server.py:370-403 server, src/workers/worker.py:199-231 channel+register):
// rank == worker id == shard id == GPU index; each process allocates one arena, exports a 64-byte IPC
// handle, the handles travel once through torch.distributed (bootstrap only) and every peer maps every
// arena.  After that all traffic is ld/st/red on peer addresses inside the PS kernels (ps_kernels.cu).
#include "common.cuh"
#include "../include/dps_b200.h"

DPS_API int dps_device_count(int* n) {
  DPS_CHECK(cudaGetDeviceCount(n));
  return 0;
}

DPS_API int dps_arena_alloc(int device, uint64_t bytes, void** out_ptr) {
  DPS_CHECK(cudaSetDevice(device));
  void* p = nullptr;
  DPS_CHECK(cudaMalloc(&p, bytes));
  DPS_CHECK(cudaMemset(p, 0, bytes));
  DPS_CHECK(cudaDeviceSynchronize());
  *out_ptr = p;
  return 0;
}

DPS_API int dps_arena_free(int device, void* ptr) {
  DPS_CHECK(cudaSetDevice(device));
  DPS_CHECK(cudaFree(ptr));
  return 0;
}

DPS_API int dps_ipc_export(void* ptr, unsigned char out_handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  DPS_CHECK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(out_handle, &h, 64);
  return 0;
}

DPS_API int dps_ipc_open(int device, const unsigned char handle[64], void** out_ptr) {
  DPS_CHECK(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  DPS_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = p;
  return 0;
}

DPS_API int dps_ipc_close(void* ptr) {
  DPS_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

DPS_API int dps_enable_peer(int device, int peer) {
  if (device == peer) return 0;
  DPS_CHECK(cudaSetDevice(device));
  int can = 0;
  DPS_CHECK(cudaDeviceCanAccessPeer(&can, device, peer));
  if (!can) DPS_FAIL(2, "device %d cannot access peer %d", device, peer);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  DPS_CHECK(e);
  return 0;
}
