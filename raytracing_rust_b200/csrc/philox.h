// philox.h — Philox4x32-10 counter-based RNG, shared by the kernels and the host-side
// primary-ray generator so both see the same draws.
//
// The reference draws rand::thread_rng() everywhere (src/util.rs:5,16, src/camera.rs:61,
// src/material.rs:118, src/medium.rs:29, tests/test.rs:56); only the distributions are
// contractual.  Here every path (pixel, sample) owns the stream
//     key = (seed_lo, seed_hi),  counter = (block, pixel, sample, 0),  block = 0,1,2,...
// which makes an image independent of how samples are sharded over GPUs.
// U01 = (u32 >> 8) * 2^-24 in [0,1).
#pragma once
#include <stdint.h>

#include "device_types.h"

namespace rtb {

struct u4 {
  uint32_t x, y, z, w;
};

RTB_HD inline uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

RTB_HD inline u4 philox4x32_10(u4 c, uint32_t k0, uint32_t k1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = mulhi32(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = mulhi32(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = u4{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

RTB_HD inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// Draw layout of a camera sample (raygen): block 0 = (jitter_u, jitter_v, lens_r, lens_phi),
// block 1 = (time, -, -, -); bounce draws start at block 2.
#define RTB_RNG_FIRST_BOUNCE_BLOCK 2u

}  // namespace rtb
