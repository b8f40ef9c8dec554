// P2P substrate: symmetric buffers shared between the per-GPU processes of one box, plus the small
// NVLink kernels that are not fused into compute (epoch broadcast, weight push, peer bandwidth probe).
//
// The reference moves every byte through TCP sockets relayed by the server (4 hops per frame,
// server.py:56,89,60,73).  Here a buffer is cudaMalloc'ed by its owner, exported with a CUDA IPC
// handle and mapped by the peers; kernels then load/store it directly over NVLink/NVSwitch
// (csrc/preproc.cu pulls frames, csrc/detect.cu pushes detection rows).
#include <string.h>

#include "common.cuh"

namespace dvc {

// K14: membership epoch word written into every peer's copy (release at system scope so kernels that
// poll it with ld.acquire.sys observe the change).  dst_ptrs is a device array of n peer addresses.
__global__ void epoch_broadcast_kernel(uint32_t** dst_ptrs, int n, uint32_t value) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) st_release_sys_u32(dst_ptrs[i], value);
}

// K13: push one source buffer (packed weights) into n peer destinations with 16-byte stores; every
// source vector is read once and stored n times (software multicast).
__global__ void __launch_bounds__(256) push_multicast_kernel(const uint4* __restrict__ src, uint4** dst_ptrs, int n, long long nvec) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
    const uint4 v = __ldg(src + i);
    for (int k = 0; k < n; ++k) dst_ptrs[k][i] = v;
  }
}

// Peer read bandwidth probe: sums a (possibly remote) buffer with 16-byte loads.
__global__ void __launch_bounds__(256) peer_read_kernel(const uint4* __restrict__ src, long long nvec, unsigned* sink) {
  unsigned acc = 0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
    const uint4 v = ldg_nc_v4(src + i);
    acc += v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0xdeadbeef) *sink = acc;
}

}  // namespace dvc

using namespace dvc;

extern "C" int dvc_p2p_alloc(size_t bytes, void** out) {
  DVC_REQUIRE(out && bytes > 0, "p2p_alloc: bad arguments");
  DVC_CHECK_CUDA(cudaMalloc(out, bytes));
  DVC_CHECK_CUDA(cudaMemset(*out, 0, bytes));
  return DVC_OK;
}
extern "C" int dvc_p2p_free(void* p) {
  DVC_CHECK_CUDA(cudaFree(p));
  return DVC_OK;
}
extern "C" int dvc_p2p_get_handle(void* p, void* h_handle64) {
  DVC_REQUIRE(p && h_handle64, "p2p_get_handle: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  DVC_CHECK_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(h_handle64), p));
  return DVC_OK;
}
extern "C" int dvc_p2p_open(const void* h_handle64, void** out) {
  DVC_REQUIRE(h_handle64 && out, "p2p_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, h_handle64, sizeof(h));
  DVC_CHECK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return DVC_OK;
}
extern "C" int dvc_p2p_close(void* p) {
  DVC_CHECK_CUDA(cudaIpcCloseMemHandle(p));
  return DVC_OK;
}
/* Same-process peers (one process driving several GPUs): enable direct access from the current device. */
extern "C" int dvc_p2p_enable_peer(int peer_device) {
  int dev = 0, can = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev == peer_device) return DVC_OK;
  DVC_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  DVC_REQUIRE(can, "p2p: device %d cannot access device %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return DVC_OK; }
  DVC_CHECK_CUDA(e);
  return DVC_OK;
}
extern "C" int dvc_epoch_broadcast(void** d_dst_ptrs, int n, uint32_t value, void* stream) {
  DVC_REQUIRE(d_dst_ptrs && n > 0, "epoch_broadcast: bad arguments");
  epoch_broadcast_kernel<<<ceil_div(n, 32), 32, 0, (cudaStream_t)stream>>>(reinterpret_cast<uint32_t**>(d_dst_ptrs), n, value);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
extern "C" int dvc_push_multicast(const void* src, void** d_dst_ptrs, int n, size_t bytes, void* stream) {
  DVC_REQUIRE(src && d_dst_ptrs && n > 0 && bytes % 16 == 0, "push_multicast: bytes must be a multiple of 16");
  const long long nvec = (long long)(bytes / 16);
  int grid = (int)ceil_div64(nvec, 256);
  if (grid > 148 * 8) grid = 148 * 8;
  push_multicast_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4**>(d_dst_ptrs), n, nvec);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
extern "C" int dvc_peer_read(const void* src, size_t bytes, void* d_sink, void* stream) {
  DVC_REQUIRE(src && d_sink && bytes % 16 == 0, "peer_read: bytes must be a multiple of 16");
  peer_read_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(src), (long long)(bytes / 16), reinterpret_cast<unsigned*>(d_sink));
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
