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

// Row-band sharding: quantise the 4-row bands band_begin, band_begin + band_stride, ... of a full-size accumulation
// buffer.  compact == 0: each row goes to its own place in a full-size rgb8 image (this GPU's or, through a peer-mapped
// pointer, the root's: the bands of the G shards tile the image without overlap).  compact != 0: the shard's rows are
// packed consecutively (for a gather).  One thread per value.
__global__ void __launch_bounds__(256) k_resolve_bands(const float* __restrict__ sum, int nx, int ny, float inv_ns, int band_begin,
                                                       int band_stride, int compact, size_t n_values, uint8_t* __restrict__ rgb8) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // index into the shard's packed rows
  if (idx >= n_values) return;
  const size_t row_len = (size_t)nx * 3;
  const size_t lrow = idx / row_len, col = idx - lrow * row_len;
  const size_t band = lrow / RTB_BAND_ROWS, r = lrow - band * RTB_BAND_ROWS;
  const size_t row = ((size_t)band_begin + band * band_stride) * RTB_BAND_ROWS + r;
  if (row >= (size_t)ny) return;
  const size_t src = row * row_len + col;
  rgb8[compact ? idx : src] = quantise(sum[src], inv_ns);
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
  if (nx <= 0 || ny <= 0) return fail(RTB_ERR_INVALID, "reduce_resolve_device: bad argument");
  return rtb_reduce_resolve_range_device(d_sums, n_bufs, 0, (size_t)3 * nx * ny, ns, d_total, d_rgb8, cuda_stream);
}

int rtb_reduce_resolve_range_device(const float* const* d_sums, int n_bufs, size_t first_value, size_t n_values, int ns,
                                    float* d_total, uint8_t* d_rgb8, void* cuda_stream) {
  if (!d_sums || n_bufs < 1 || n_bufs > RTB_MAX_PEER_BUFFERS || ns <= 0 || !d_rgb8)
    return fail(RTB_ERR_INVALID, "reduce_resolve_device: bad argument");
  if (n_values == 0) return RTB_OK;
  d_rgb8 += first_value;
  if (d_total) d_total += first_value;
  PeerBufs b;
  b.n = n_bufs;
  bool vec = n_values % 4 == 0 && ((uintptr_t)d_rgb8 & 3) == 0 && (!d_total || ((uintptr_t)d_total & 15) == 0);
  for (int g = 0; g < RTB_MAX_PEER_BUFFERS; ++g) {
    b.p[g] = g < n_bufs && d_sums[g] ? d_sums[g] + first_value : nullptr;
    if (g < n_bufs && !d_sums[g]) return fail(RTB_ERR_INVALID, "reduce_resolve_device: null buffer");
    if (g < n_bufs && ((uintptr_t)b.p[g] & 15)) vec = false;
  }
  const size_t n = n_values;
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

int rtb_resolve_bands_device(const float* d_sum_rgb, int nx, int ny, int ns, int band_begin, int band_stride, uint8_t* d_rgb8,
                             int compact, void* cuda_stream) {
  if (!d_sum_rgb || !d_rgb8 || nx <= 0 || ny <= 0 || ns <= 0 || band_begin < 0 || band_stride < 1)
    return fail(RTB_ERR_INVALID, "resolve_bands_device: bad argument");
  const int bands_all = (ny + RTB_BAND_ROWS - 1) / RTB_BAND_ROWS;
  const int my_bands = band_begin < bands_all ? (bands_all - band_begin + band_stride - 1) / band_stride : 0;
  const size_t n = (size_t)my_bands * RTB_BAND_ROWS * nx * 3;
  if (n == 0) return RTB_OK;
  k_resolve_bands<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(d_sum_rgb, nx, ny, 1.0f / (float)ns, band_begin,
                                                                                     band_stride, compact, n, d_rgb8);
  PEER_TRY(cudaGetLastError());
  return RTB_OK;
}

}  // extern "C"
