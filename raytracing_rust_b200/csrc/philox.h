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
// whole bounce.  U = fl((k + 0.5) * 2^-24) in (0, 1]: never 0 (no log(0), no zero-length ball sample); the single
// value k = 2^24 - 1 rounds to exactly 1.0f.
#pragma once
#include <stdint.h>

#include "device_types.h"

namespace rtb {

struct u4 {
  uint32_t x, y, z, w;
};

// 32 x 32 -> (hi, lo).  On the device the two halves are asked for separately (one IMAD.WIDE.U32 after ptxas merges
// them); the 64-bit product form made nvcc add a zero to every high word (+2 instructions per round).
RTB_HD inline void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#if defined(__CUDA_ARCH__)
  hi = __umulhi(a, b), lo = a * b;
#else
  const uint64_t p = (uint64_t)a * b;
  hi = (uint32_t)(p >> 32), lo = (uint32_t)p;
#endif
}

RTB_HD inline u4 philox4x32_10(u4 c, uint32_t k0, uint32_t k1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 10; ++i) {
    uint32_t h0, l0, h1, l1;
    philox_mulhilo(0xD2511F53u, c.x, h0, l0);
    philox_mulhilo(0xCD9E8D57u, c.z, h1, l1);
    c = u4{h1 ^ c.y ^ k0, l1, h0 ^ c.w ^ k1, l0};
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

// The ten round keys of a seed, made once on the host: passed to a kernel by value they sit in the constant bank and
// enter each round's LOP3 as an operand, so a Philox call is 4 instructions per round (2 IMAD.WIDE + 2 LOP3).
struct PhiloxKeys {
  uint32_t k[20];
};
RTB_HD inline PhiloxKeys philox_keys(uint64_t seed) {
  PhiloxKeys K;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int i = 0; i < 10; ++i) {
    K.k[2 * i] = k0, K.k[2 * i + 1] = k1;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return K;
}
RTB_HD inline u4 philox4x32_10(u4 c, const PhiloxKeys& K) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int i = 0; i < 10; ++i) {
    uint32_t h0, l0, h1, l1;
    philox_mulhilo(0xD2511F53u, c.x, h0, l0);
    philox_mulhilo(0xCD9E8D57u, c.z, h1, l1);
    c = u4{h1 ^ c.y ^ K.k[2 * i], l1, h0 ^ c.w ^ K.k[2 * i + 1], l0};
  }
  return c;
}

struct Rand5 {
  float a, b, c, d, e;
};

// (k + 0.5) * 2^-24 as one FMA: k * 2^-24 is exact and the single rounding of the sum is the rounding of k + 0.5 scaled
// by a power of two, so the value is the same as the two-step form on host and device alike
RTB_HD inline float u24_to_unit(uint32_t k) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn((float)k, 1.0f / 16777216.0f, 0.5f / 16777216.0f);
#else
  return ((float)k + 0.5f) * (1.0f / 16777216.0f);
#endif
}

RTB_HD inline Rand5 rand5_from(const u4 r) {
  Rand5 o;
  o.a = u24_to_unit(r.x >> 8);
  o.b = u24_to_unit(r.y >> 8);
  o.c = u24_to_unit(r.z >> 8);
  o.d = u24_to_unit(r.w >> 8);
  o.e = u24_to_unit(((r.x & 255u) << 16) | ((r.y & 255u) << 8) | (r.z & 255u));
  return o;
}
RTB_HD inline Rand5 rand5(uint64_t seed, uint32_t block, uint32_t pixel, uint32_t sample, uint32_t purpose) {
  return rand5_from(philox4x32_10(u4{block, pixel, sample, purpose}, (uint32_t)seed, (uint32_t)(seed >> 32)));
}
RTB_HD inline Rand5 rand5(const PhiloxKeys& K, uint32_t block, uint32_t pixel, uint32_t sample, uint32_t purpose) {
  return rand5_from(philox4x32_10(u4{block, pixel, sample, purpose}, K));
}

// Counter layout.
//   purpose 0, block 0      camera sample: (jitter_u, jitter_v, lens_r, lens_phi, time)
//   purpose 1, block depth  bounce `depth`: (ball0 | schlick, ball1, ball2, medium0, medium1)
//   purpose 2+g, block depth  free-flight draws of media 2+5g .. 6+5g on that segment
#define RTB_RNG_CAMERA 0u
#define RTB_RNG_BOUNCE 1u
#define RTB_RNG_MEDIA 2u

}  // namespace rtb
