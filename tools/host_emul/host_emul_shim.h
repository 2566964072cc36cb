// host_emul_shim.h — DEVELOPER TOOL, not product code.  Lets g++ compile csrc/trace.cuh (the kernels' device functions)
// for the host, so traversal / shading LOGIC can be debugged and counted in a container that has no GPU.  SFU
// approximations become libm calls, read-only loads become plain loads.  librtb200.so never contains this.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

#define RTB_DEV inline
struct float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
template <class T> inline T __ldg(const T* p) { return *p; }
inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
inline int __float_as_int(float f) { int v; std::memcpy(&v, &f, 4); return v; }
inline unsigned __float_as_uint(float f) { unsigned v; std::memcpy(&v, &f, 4); return v; }
inline float __uint_as_float(unsigned v) { float f; std::memcpy(&f, &v, 4); return f; }
inline float __log2f(float x) { return std::log2(x); }
inline float __logf(float x) { return std::log(x); }
inline float __sinf(float x) { return std::sin(x); }
inline void __sincosf(float x, float* s, float* c) { *s = std::sin(x), *c = std::cos(x); }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz((unsigned)v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
#include <math.h>  // the C names (floorf, fabsf, ...) in the global namespace
#define CUDART_INF_F (__builtin_inff())
#define CUDART_PI_F 3.141592654f
