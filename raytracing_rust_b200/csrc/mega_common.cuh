// mega_common.cuh — pieces shared by the persistent path-tracing kernels (render.cu: k_render_mega,
// dual.cu: k_render_dual): launch parameters, lane-state codes, the accumulation-tile flush.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_types.h"
#include "philox.h"

namespace rtb {

struct MegaParams {
  int nx, ny;
  int ns;             // total samples per pixel of the image
  int sample_first;   // first sample index this launch renders
  int sample_step;    // distance between consecutive samples of this launch (GPU sharding stride)
  int sample_count;   // samples per pixel in this launch
  int slices;         // work items per pixel tile (each takes every slices-th sample of the launch)
  int n_items;        // tiles * slices; warps draw items from a global counter until none is left
  int band_begin, band_stride;  // row-band sharding: tile row r of this launch is tile row band_begin + r * band_stride of the image
  int max_depth;
  float t_min;
  int background;
  int k_node, k_leaf, k_shade;  // lanes a phase needs before the warp's scheduler runs it
  uint64_t seed;
  PhiloxKeys keys;  // philox_keys(seed)
  // Scenes whose (at most two) constant media are all bounded by one static sphere carry the media in the kernel
  // parameters: centre.xyz, radius | -(1/density), material.  n_const_media = -1: use the tables (sample_media).
  int n_const_media;
  float4 cmedia_sphere[2];
  float cmedia_neg_inv_density[2];
  int cmedia_mat[2];
};

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Lane states of the megakernel's per-warp scheduler, stored in `cur`:
//   cur >= 0                inner BVH node to test next
//   CUR_DONE < cur < 0      leaf code ~(first << 3 | count): primitives still to test
//   CUR_SHADE               traversal finished, waiting for the shade phase
//   CUR_NEW                 needs a new camera path
//   CUR_DONE                pool exhausted
#define CUR_SHADE (-2147483647 - 1)
#define CUR_NEW (-2147483647)
#define CUR_DONE (-2147483645)
#define CUR_WAIT (-2147483646)  // needs a path but the warp's current item is used up: waits for the next item

// Adds one warp's finished accumulation tile to the image (one RED.global per non-zero
// pixel-channel) and clears it.  All 32 lanes call it together.
__device__ __forceinline__ void flush_tile(float* tile_acc, int4 info, int nx, int lane, float* __restrict__ accum) {
  if (info.x >= 0) {
    const int x0 = info.x, y0 = info.y, w_t = info.z & 255, n_pix = info.z >> 8;
    for (int q = lane; q < n_pix * 3; q += 32) {
      const float v = tile_acc[q];
      if (v != 0.0f) {
        const int pl = q / 3, c = q - pl * 3;
        const int row = y0 + pl / w_t, i = x0 + pl % w_t;
        atomicAdd(accum + 3ull * ((size_t)row * nx + i) + c, v);
        tile_acc[q] = 0.0f;
      }
    }
  }
  __syncwarp();
}

}  // namespace rtb
