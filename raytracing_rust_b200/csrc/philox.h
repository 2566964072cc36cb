// philox.h — Philox4x32-10 counter-based RNG, shared by the kernels and the host-side
// primary-ray generator so both see the same draws.
//
// The reference draws rand::thread_rng() everywhere (src/util.rs:5,16, src/camera.rs:61,
// src/material.rs:118, src/medium.rs:29, tests/test.rs:56); only the distributions are
// contractual.  Here every draw is a pure function of
//     key = (seed_lo, seed_hi),  counter = (block, pixel, sample, purpose)
// which makes an image independent of how samples are sharded over GPUs and needs no RNG state
// in the path.  One 128-bit Philox output is cut into FIVE 24-bit uniforms (the top 24 bits of
// each word, plus the four low bytes of x,y,z): enough for one whole camera sample or one
// whole bounce.  U(0,1) = (k + 0.5) * 2^-24, strictly inside (0,1).
#pragma once
#include <stdint.h>

#include "device_types.h"

namespace rtb {

struct u4 {
  uint32_t x, y, z, w;
};

RTB_HD inline u4 philox4x32_10(u4 c, uint32_t k0, uint32_t k1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 10; ++i) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c.x;  // one IMAD.WIDE.U32 each
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c.z;
    c = u4{(uint32_t)(p1 >> 32) ^ c.y ^ k0, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c.w ^ k1, (uint32_t)p0};
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

struct Rand5 {
  float a, b, c, d, e;
};

RTB_HD inline float u24_to_unit(uint32_t k) { return ((float)k + 0.5f) * (1.0f / 16777216.0f); }

RTB_HD inline Rand5 rand5(uint64_t seed, uint32_t block, uint32_t pixel, uint32_t sample, uint32_t purpose) {
  const u4 r = philox4x32_10(u4{block, pixel, sample, purpose}, (uint32_t)seed, (uint32_t)(seed >> 32));
  Rand5 o;
  o.a = u24_to_unit(r.x >> 8);
  o.b = u24_to_unit(r.y >> 8);
  o.c = u24_to_unit(r.z >> 8);
  o.d = u24_to_unit(r.w >> 8);
  o.e = u24_to_unit(((r.x & 255u) << 16) | ((r.y & 255u) << 8) | (r.z & 255u));
  return o;
}

// Counter layout.
//   purpose 0, block 0      camera sample: (jitter_u, jitter_v, lens_r, lens_phi, time)
//   purpose 1, block depth  bounce `depth`: (ball0 | schlick, ball1, ball2, medium0, medium1)
//   purpose 2+g, block depth  free-flight draws of media 2+5g .. 6+5g on that segment
#define RTB_RNG_CAMERA 0u
#define RTB_RNG_BOUNCE 1u
#define RTB_RNG_MEDIA 2u

}  // namespace rtb
