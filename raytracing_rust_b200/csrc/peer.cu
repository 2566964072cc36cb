// peer.cu — the multi-GPU exchange step over peer memory (NVLink 5 / NVSwitch).
//
// The reference's pixel loop keeps one accumulator per pixel and finishes with
//   col /= ns; c = clamp(sqrt(c), 0, 1); (255.99 * c) as i32          tests/test.rs:71-78
// With samples sharded over G GPUs (rtb_render_opts.sample_begin/stride) every GPU holds a partial
// float3[W*H] sum.  k_reduce_resolve is the exchange step and the epilogue in ONE kernel on the
// root GPU: each thread pulls its pixels' partial sums from all G buffers (its own plus the peers'
// buffers mapped through CUDA IPC, i.e. P2P loads over NVLink), adds them in rank order (so the
// result does not depend on timing), and writes the quantised image — no intermediate reduced
// buffer, no separate collective launch.  raytracing_rust_b200/distributed.py drives it and
// keeps the NCCL reduce + rtb_resolve_device path beside it.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>

#include "rtb200.h"
#include "scene.hpp"

namespace rtb {

struct PeerBufs {
  const float* p[RTB_MAX_PEER_BUFFERS];
  int n;
};

__device__ __forceinline__ uint8_t quantise(float sum, float inv_ns) {
  float c = sqrtf(sum * inv_ns);
  c = c > 1.0f ? 1.0f : (c < 0.0f ? 0.0f : c);  // NaN passes through, then `as i32` -> 0
  const float q = 255.99f * c;
  const int v = (q != q) ? 0 : (int)q;
  return (uint8_t)min(max(v, 0), 255);
}

// 16 bytes per thread per buffer: with G = 8 each thread keeps eight independent 128-bit P2P
// loads in flight before the first add.
template <bool kVec>
__global__ void __launch_bounds__(256) k_reduce_resolve(PeerBufs bufs, size_t n_values, float inv_ns, float* __restrict__ total,
                                                        uint8_t* __restrict__ rgb8) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  if (kVec) {
    const size_t n4 = n_values >> 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      float4 v[RTB_MAX_PEER_BUFFERS];
#pragma unroll
      for (int g = 0; g < RTB_MAX_PEER_BUFFERS; ++g)
        if (g < bufs.n) v[g] = __ldcs(reinterpret_cast<const float4*>(bufs.p[g]) + i);  // streamed once
      float4 s = v[0];
#pragma unroll
      for (int g = 1; g < RTB_MAX_PEER_BUFFERS; ++g)
        if (g < bufs.n) s.x += v[g].x, s.y += v[g].y, s.z += v[g].z, s.w += v[g].w;
      if (total) reinterpret_cast<float4*>(total)[i] = s;
      uchar4 q;
      q.x = quantise(s.x, inv_ns), q.y = quantise(s.y, inv_ns), q.z = quantise(s.z, inv_ns), q.w = quantise(s.w, inv_ns);
      reinterpret_cast<uchar4*>(rgb8)[i] = q;
    }
  } else {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_values; i += stride) {
      float s = bufs.p[0][i];
      for (int g = 1; g < bufs.n; ++g) s += bufs.p[g][i];
      if (total) total[i] = s;
      rgb8[i] = quantise(s, inv_ns);
    }
  }
}

}  // namespace rtb

using namespace rtb;

#define PEER_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

extern "C" {

int rtb_device_alloc(size_t bytes, int device, void** d_ptr) {
  if (!d_ptr || bytes == 0) return fail(RTB_ERR_INVALID, "device_alloc: bad argument");
  PEER_TRY(cudaSetDevice(device));
  PEER_TRY(cudaMalloc(d_ptr, bytes));  // a whole allocation: its base pointer can be exported over IPC
  PEER_TRY(cudaMemset(*d_ptr, 0, bytes));
  return RTB_OK;
}

int rtb_device_free(void* d_ptr) {
  if (d_ptr) PEER_TRY(cudaFree(d_ptr));
  return RTB_OK;
}

int rtb_peer_export(const void* d_ptr, void* handle) {
  if (!d_ptr || !handle) return fail(RTB_ERR_INVALID, "peer_export: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == RTB_IPC_HANDLE_BYTES, "IPC handle size");
  PEER_TRY(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle), const_cast<void*>(d_ptr)));
  return RTB_OK;
}

int rtb_peer_open(const void* handle, int device, void** d_ptr) {
  if (!handle || !d_ptr) return fail(RTB_ERR_INVALID, "peer_open: bad argument");
  PEER_TRY(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  PEER_TRY(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RTB_OK;
}

int rtb_peer_close(void* d_ptr) {
  if (d_ptr) PEER_TRY(cudaIpcCloseMemHandle(d_ptr));
  return RTB_OK;
}

int rtb_reduce_resolve_device(const float* const* d_sums, int n_bufs, int nx, int ny, int ns, float* d_total, uint8_t* d_rgb8,
                              void* cuda_stream) {
  if (!d_sums || n_bufs < 1 || n_bufs > RTB_MAX_PEER_BUFFERS || nx <= 0 || ny <= 0 || ns <= 0 || !d_rgb8)
    return fail(RTB_ERR_INVALID, "reduce_resolve_device: bad argument");
  PeerBufs b;
  b.n = n_bufs;
  bool vec = ((size_t)3 * nx * ny) % 4 == 0 && ((uintptr_t)d_rgb8 & 3) == 0 && (!d_total || ((uintptr_t)d_total & 15) == 0);
  for (int g = 0; g < RTB_MAX_PEER_BUFFERS; ++g) {
    b.p[g] = g < n_bufs ? d_sums[g] : nullptr;
    if (g < n_bufs && !d_sums[g]) return fail(RTB_ERR_INVALID, "reduce_resolve_device: null buffer");
    if (g < n_bufs && ((uintptr_t)d_sums[g] & 15)) vec = false;
  }
  const size_t n = (size_t)3 * nx * ny;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const size_t work = vec ? n / 4 : n;
  const unsigned blocks = (unsigned)std::min<size_t>((work + 255) / 256, (size_t)sms * 8);  // grid-stride, 8 CTAs per SM
  if (vec)
    k_reduce_resolve<true><<<blocks, 256, 0, (cudaStream_t)cuda_stream>>>(b, n, 1.0f / (float)ns, d_total, d_rgb8);
  else
    k_reduce_resolve<false><<<blocks, 256, 0, (cudaStream_t)cuda_stream>>>(b, n, 1.0f / (float)ns, d_total, d_rgb8);
  PEER_TRY(cudaGetLastError());
  return RTB_OK;
}

}  // extern "C"
