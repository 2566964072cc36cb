// trace.cuh — device-side building blocks shared by the megakernel and the wavefront
// kernels: RNG draws, ray/primitive intersection, BVH traversal, constant media, textures and
// material scattering.  f32 throughout; each routine cites the f64 reference code whose result
// it reproduces.
#pragma once
#if defined(RTB_HOST_EMUL)
// developer build only (tools/host_emul): the same source compiled for the host by g++, to debug kernel LOGIC in a
// container without a GPU.  Never part of librtb200.so.
#include "host_emul_shim.h"
#else
#include <cuda_runtime.h>
#include <math_constants.h>
#define RTB_DEV __device__ __forceinline__
#endif


#include "device_types.h"
#include "philox.h"
#include "rtb_scene_format.h"

namespace rtb {

// ---- self-checks (-DRTB_DEBUG_CHECKS=1 builds only; compute-sanitizer is not available on the GPU pool) --------------
// Every index the kernels form — traversal stack slots, node / primitive / material / texture / texel indices, shared
// accumulation tiles, output pixels — is compared with its bound and violations are COUNTED per class in a device array
// the host reads back (rtb_debug_check_failures).  Release builds compile the checks away.
enum RtbCheck { RTB_CHK_STACK = 0, RTB_CHK_NODE = 1, RTB_CHK_PRIM = 2, RTB_CHK_MAT = 3, RTB_CHK_TEX = 4, RTB_CHK_TEXEL = 5,
                RTB_CHK_TILE = 6, RTB_CHK_PIXEL = 7, RTB_CHK_QUEUE = 8, RTB_CHK_NAN = 9, RTB_CHK_COUNT = 16 };
#if defined(RTB_DEBUG_CHECKS) && !defined(RTB_HOST_EMUL)
// one counter array per translation unit (no relocatable device code in this build): each .cu reports its own kernels
static __device__ unsigned int g_rtb_check_fail[RTB_CHK_COUNT];
#define RTB_CHECK(cond, cls)                            \
  do {                                                  \
    if (!(cond)) atomicAdd(&g_rtb_check_fail[cls], 1u); \
  } while (0)
// each translation unit defines its reader once (render.cu, wavefront.cu, debug.cu); rtb_debug_check_failures adds them up
#define RTB_DEFINE_CHECK_READER(fn)                                                                  \
  int fn(unsigned int* acc) {                                                                        \
    unsigned int h[RTB_CHK_COUNT];                                                                   \
    if (cudaMemcpyFromSymbol(h, g_rtb_check_fail, sizeof h) != cudaSuccess) return -1;               \
    for (int i = 0; i < RTB_CHK_COUNT; ++i) acc[i] += h[i], h[i] = 0;                                \
    return cudaMemcpyToSymbol(g_rtb_check_fail, h, sizeof h) == cudaSuccess ? 0 : -1;                \
  }
#else
#define RTB_CHECK(cond, cls) ((void)0)
#define RTB_DEFINE_CHECK_READER(fn) \
  int fn(unsigned int*) { return 0; }
#endif

// ---------------------------------------------------------------------------------------------
struct V3f {
  float x, y, z;
};
RTB_DEV V3f mk(float x, float y, float z) { return V3f{x, y, z}; }
RTB_DEV V3f operator+(V3f a, V3f b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
RTB_DEV V3f operator-(V3f a, V3f b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
RTB_DEV V3f operator-(V3f a) { return mk(-a.x, -a.y, -a.z); }
RTB_DEV V3f operator*(float s, V3f a) { return mk(s * a.x, s * a.y, s * a.z); }
RTB_DEV V3f operator*(V3f a, V3f b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
RTB_DEV float dot(V3f a, V3f b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
RTB_DEV V3f madd(float t, V3f d, V3f o) {  // o + t*d
  return mk(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
}
RTB_DEV float comp(V3f v, int axis) { return axis == 0 ? v.x : (axis == 1 ? v.y : v.z); }
RTB_DEV void set_comp(V3f& v, int axis, float s) {
  if (axis == 0) v.x = s;
  else if (axis == 1) v.y = s;
  else v.z = s;
}

RTB_DEV float4 ldg4(const f4* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
RTB_DEV int4 ldg4i(const i4* p) { return __ldg(reinterpret_cast<const int4*>(p)); }
// one 256-bit read-only load (sm_100a: ld.global.nc.v8.f32 -> LDG.E.256.CONSTANT); p is 32-byte aligned
struct alignas(32) Ld256 {
  float v[8];
};
RTB_DEV Ld256 ld256(const Ld256* p) {
  Ld256 r;
#if defined(RTB_HOST_EMUL)
  return *p;
#else
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
      : "l"(p));
  return r;
#endif
}

// Single-instruction SFU forms (MUFU.RCP / RSQ / SQRT / EX2 / LG2, <= 2 ulp).  The f64 reference
// is matched statistically and through closest-hit IDs, not bit-wise, so the IEEE-rounded
// division / sqrt sequences (and their slow-path subroutines) buy nothing here.
RTB_DEV float rcp_fast(float x) {
#if defined(RTB_HOST_EMUL)
  return 1.0f / x;
#else
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}
RTB_DEV float sqrt_fast(float x) {
#if defined(RTB_HOST_EMUL)
  return sqrtf(x);
#else
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}
RTB_DEV float rsqrt_fast(float x) {
#if defined(RTB_HOST_EMUL)
  return 1.0f / sqrtf(x);
#else
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}
RTB_DEV float cbrt_fast01(float x) {  // x in (0,1): 2^(log2(x)/3)
  return exp2f(__log2f(x) * (1.0f / 3.0f));
}
// sin(x) for |x| up to ~1e5: two-constant Cody-Waite reduction to [-pi, pi], then MUFU.SIN
RTB_DEV float sin_reduced(float x) {
  float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(-k, 6.2831854820251465f, x);  // 2*pi rounded to f32
  r = fmaf(-k, -1.7484555e-7f, r);              // 2*pi - fl(2*pi)
  return __sinf(r);
}

// ---- RNG: one Philox4x32-10 call = four U(0,1) draws ------------------------------------------
// counter = (block, pixel, sample, purpose), key = seed.  purpose 0: block 0 = camera
// (jitter_u, jitter_v, lens_r, lens_phi), block 1 = (time,-,-,-); purpose 1: block = depth,
// the scatter draws of that bounce (ball x3, schlick); purpose 2 + g: block = depth, the
// free-flight draws of media 4g .. 4g+3 on that segment.
// (Rand5 / rand5 live in philox.h: five 24-bit uniforms per Philox call, in (0,1] and never 0,
// so there is no zero-length ball sample and no log(0).)

// uniform in the unit ball == the distribution of the rejection loop at src/util.rs:4-13,
// sampled directly from 3 draws (no divergence)
RTB_DEV V3f sample_unit_ball(float r0, float r1, float r2) {
  float z = 1.0f - 2.0f * r0;
  float phi = (2.0f * r1 - 1.0f) * CUDART_PI_F;
  float r = cbrt_fast01(r2);
  float s = sqrt_fast(fmaxf(0.0f, 1.0f - z * z)) * r;
  float sn, cs;
  __sincosf(phi, &sn, &cs);
  return mk(s * cs, s * sn, r * z);
}

// ---- rays -----------------------------------------------------------------------------------
struct RayF {
  V3f o, d;
  float time;
};

struct Hit {
  float t;
  int prim;  // index into prims (or bprims for boundary queries); -1 = none
  int face;  // BOX: face 0..5 in the order of src/cube.rs:21-74
  int id;    // primitive ID of the hit (first ID + face), for the tie rule
};

RTB_DEV V3f xf_point(const DXform& x, V3f p) {  // local = M p + b
  return mk(fmaf(x.m[0], p.x, fmaf(x.m[1], p.y, fmaf(x.m[2], p.z, x.b[0]))),
            fmaf(x.m[3], p.x, fmaf(x.m[4], p.y, fmaf(x.m[5], p.z, x.b[1]))),
            fmaf(x.m[6], p.x, fmaf(x.m[7], p.y, fmaf(x.m[8], p.z, x.b[2]))));
}
RTB_DEV V3f xf_vec(const DXform& x, V3f v) {  // M v
  return mk(fmaf(x.m[0], v.x, fmaf(x.m[1], v.y, x.m[2] * v.z)), fmaf(x.m[3], v.x, fmaf(x.m[4], v.y, x.m[5] * v.z)),
            fmaf(x.m[6], v.x, fmaf(x.m[7], v.y, x.m[8] * v.z)));
}
RTB_DEV V3f xf_vec_t(const DXform& x, V3f v) {  // M^T v
  return mk(fmaf(x.m[0], v.x, fmaf(x.m[3], v.y, x.m[6] * v.z)), fmaf(x.m[1], v.x, fmaf(x.m[4], v.y, x.m[7] * v.z)),
            fmaf(x.m[2], v.x, fmaf(x.m[5], v.y, x.m[8] * v.z)));
}
RTB_DEV V3f xf_point_inv(const DXform& x, V3f p) {  // world = M^T (p - b)
  return xf_vec_t(x, mk(p.x - x.b[0], p.y - x.b[1], p.z - x.b[2]));
}

RTB_DEV DXform load_xform(const DXform* xs, int i) {
  DXform x;
  const float4* p = reinterpret_cast<const float4*>(xs + i);
  float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  x.m[0] = a.x, x.m[1] = a.y, x.m[2] = a.z, x.m[3] = a.w;
  x.m[4] = b.x, x.m[5] = b.y, x.m[6] = b.z, x.m[7] = b.w;
  x.m[8] = c.x;
  x.b[0] = c.y, x.b[1] = c.z, x.b[2] = c.w;
  return x;
}

RTB_DEV V3f sphere_center(float4 a, float4 b, int4 c, float time) {
  // MovingSphere::center, src/sphere.rs:115-118
  V3f ctr = mk(a.x, a.y, a.z);
  if (RTB_PRIM_TYPE(c.x) == DP_MOVING_SPHERE) {
    float f = (time - b.w) * __int_as_float(c.w);
    ctr = madd(f, mk(b.x, b.y, b.z), ctr);
  }
  return ctr;
}

struct Counters {
  unsigned long long node_tests, prim_tests, medium_tests;
};

// One primitive against one ray over (t_min, t_max).  `self` means the ray origin lies ON this
// primitive (it is the scattering surface): in exact arithmetic that root is t = 0 < t_min, so
// it is excluded analytically instead of trusting f32 round-off against t_min = 1e-3.
//   Sphere: src/sphere.rs:37-77 (strict t_min < t < t_max, near root first), with the
//           closest-approach form t = tc -/+ h, h^2 = (r^2 - |oc + tc d|^2)/a, which does not
//           cancel for the r = 1000 / r = 5000 spheres the way b^2 - a c does in f32.
//   Rect:   src/rect.rs:39-69 (all comparisons inclusive).
//   Box:    the six Rects of src/cube.rs:21-74 scanned by HittableList::hit == nearest of
//           {slab entry, slab exit} inside [t_min, t_max].
// Plane distances are always (k - o) * rcp(d), in Rect and Box alike, so coincident planes
// (a cube standing on a floor rect) produce bit-identical t and the ID tie rule can apply.
// kClass: what the caller knows about the record at compile time (the flat scan of render.cu walks its primitives class by
// class, so the type / plane / transform tests and the axis selects fold away; the arithmetic is the same either way).
enum PrimClass { PC_ANY = -1, PC_SPHERE = 0, PC_RECT_YZ = 1, PC_RECT_ZX = 2, PC_RECT_XY = 3 };  // rect classes: no transform
template <bool kStats, int kClass>
RTB_DEV bool intersect_prim_abc(const float4 A, const float4 B, const int4 C, const DXform* __restrict__ xforms,
                                const RayF& ray, float inv_a, float t_min, float t_max, bool self,
                                int self_face, float* t_out, int* face_out, int* id_out, Counters* cnt) {
  if (kStats) cnt->prim_tests++;
  const bool k_rect = kClass >= PC_RECT_YZ && kClass <= PC_RECT_XY;
  const int type = kClass == PC_SPHERE ? (int)DP_SPHERE : (k_rect ? (int)DP_RECT : RTB_PRIM_TYPE(C.x));
  *id_out = C.z;
  *face_out = 0;
  if (type <= DP_MOVING_SPHERE) {
    V3f ctr = kClass == PC_SPHERE ? mk(A.x, A.y, A.z) : sphere_center(A, B, C, ray.time);
    V3f oc = ray.o - ctr;
    float b = dot(oc, ray.d);
    float tc = -b * inv_a;
    float t;
    if (self) {
      t = 2.0f * tc;  // roots are 0 and -2b/a
      *t_out = t;
      return t < t_max && t > t_min;
    }
    V3f l = madd(tc, ray.d, oc);
    float disc = fmaf(A.w, A.w, -dot(l, l));
    if (disc > 0.0f) {
      float h = sqrt_fast(disc * inv_a);
      t = tc - h;
      if (!(t < t_max && t > t_min)) t = tc + h;
      *t_out = t;
      return t < t_max && t > t_min;
    }
    return false;
  }
  // planar primitives live in the local space of their instance transform
  V3f o = ray.o, d = ray.d;
  const int xi = k_rect ? 0 : RTB_PRIM_XFORM(C.x);
  if (xi != 0) {
    DXform x = load_xform(xforms, xi);
    o = xf_point(x, ray.o);
    d = xf_vec(x, ray.d);
  }
  if (type == DP_RECT) {
    if (self) return false;  // a ray leaving a plane never meets it again
    const int plane = k_rect ? kClass - PC_RECT_YZ : RTB_PRIM_PLANE(C.x);
    const int k_axis = plane, a_axis = plane == 2 ? 0 : plane + 1, b_axis = plane == 0 ? 2 : plane - 1;
    float t = (B.x - comp(o, k_axis)) * rcp_fast(comp(d, k_axis));
    if (t < t_min || t > t_max) return false;
    float x = fmaf(t, comp(d, a_axis), comp(o, a_axis));
    float y = fmaf(t, comp(d, b_axis), comp(o, b_axis));
    if (x < A.x || x > A.z || y < A.y || y > A.w) return false;
    *t_out = t;
    return true;
  }
  // DP_BOX
  float ix = rcp_fast(d.x), iy = rcp_fast(d.y), iz = rcp_fast(d.z);
  float tx0 = (A.x - o.x) * ix, tx1 = (B.x - o.x) * ix;
  float ty0 = (A.y - o.y) * iy, ty1 = (B.y - o.y) * iy;
  float tz0 = (A.z - o.z) * iz, tz1 = (B.z - o.z) * iz;
  float nx = fminf(tx0, tx1), fx = fmaxf(tx0, tx1);
  float ny = fminf(ty0, ty1), fy = fmaxf(ty0, ty1);
  float nz = fminf(tz0, tz1), fz = fmaxf(tz0, tz1);
  if (self) {
    // origin lies on face `self_face` (axis 2 - face/2; odd = min plane)
    const int ax = 2 - (self_face >> 1);
    const bool is_min = self_face & 1;
    const float dk = comp(d, ax);
    const bool entering = is_min ? dk > 0.0f : dk < 0.0f;  // that plane is the near plane of its slab
    const float nn = entering ? -CUDART_INF_F : (ax == 0 ? nx : (ax == 1 ? ny : nz));
    const float ff = entering ? (ax == 0 ? fx : (ax == 1 ? fy : fz)) : 0.0f;
    if (ax == 0) nx = nn, fx = ff;
    else if (ax == 1) ny = nn, fy = ff;
    else nz = nn, fz = ff;
  }
  float t_enter = fmaxf(fmaxf(nx, ny), nz);
  float t_exit = fminf(fminf(fx, fy), fz);
  if (!(t_enter <= t_exit)) return false;
  // candidates in list-scan terms: the entry face, then the exit face
  const bool use_enter = t_enter >= t_min && t_enter <= t_max;
  const bool use_exit = t_exit >= t_min && t_exit <= t_max;
  if (!use_enter && !use_exit) return false;
  int ax;
  bool is_min;
  if (use_enter) {
    ax = (nx >= ny && nx >= nz) ? 0 : (ny >= nz ? 1 : 2);
    is_min = comp(d, ax) > 0.0f;  // entering through the min plane when d > 0
    *t_out = t_enter;
  } else {
    ax = (fx <= fy && fx <= fz) ? 0 : (fy <= fz ? 1 : 2);
    is_min = !(comp(d, ax) > 0.0f);  // leaving through the max plane when d > 0
    *t_out = t_exit;
  }
  const int face = ((2 - ax) << 1) | (is_min ? 1 : 0);
  *face_out = face;
  *id_out = C.z + face;
  return true;
}

// the record fetched from the primitive table (3 x LDG.128)
template <bool kStats>
RTB_DEV bool intersect_prim(const DPrim* __restrict__ prims, int pi, const DXform* __restrict__ xforms,
                                               const RayF& ray, float inv_a, float t_min, float t_max, bool self,
                                               int self_face, float* t_out, int* face_out, int* id_out, Counters* cnt) {
  const float4 A = ldg4(&prims[pi].a);
  const float4 B = ldg4(&prims[pi].b);
  const int4 C = ldg4i(&prims[pi].c);
  return intersect_prim_abc<kStats, PC_ANY>(A, B, C, xforms, ray, inv_a, t_min, t_max, self, self_face, t_out, face_out, id_out, cnt);
}

// ---- BVH node step -----------------------------------------------------------------------------
struct RayInv {  // per-ray constants of the slab test
  float idx, idy, idz, oodx, oody, oodz;
};
RTB_DEV RayInv make_ray_inv(const RayF& ray) {
  auto safe_inv = [](float v) { return rcp_fast(fabsf(v) > 1e-20f ? v : copysignf(1e-20f, v)); };
  RayInv r;
  r.idx = safe_inv(ray.d.x), r.idy = safe_inv(ray.d.y), r.idz = safe_inv(ray.d.z);
  r.oodx = ray.o.x * r.idx, r.oody = ray.o.y * r.idy, r.oodz = ray.o.z * r.idz;
  return r;
}

// Tests both children of inner node `cur` against (t_min, t_best).  Returns the number of
// children hit (0, 1, 2); *next = the child to visit next (the nearer one), *push = the other
// (valid when 2 is returned).
RTB_DEV int node_step(const DNode* __restrict__ nodes, int cur, const RayInv& ri, float t_min, float t_best,
                                         int* next, int* push) {
  // two 256-bit loads (LDG.E.256.CONSTANT): (n0, n1) and (n2, n3)
  const Ld256 h0v = ld256(reinterpret_cast<const Ld256*>(nodes + cur));
  const Ld256 h1v = ld256(reinterpret_cast<const Ld256*>(nodes + cur) + 1);
  const float4 n0 = make_float4(h0v.v[0], h0v.v[1], h0v.v[2], h0v.v[3]), n1 = make_float4(h0v.v[4], h0v.v[5], h0v.v[6], h0v.v[7]);
  const float4 n2 = make_float4(h1v.v[0], h1v.v[1], h1v.v[2], h1v.v[3]);
  const int4 n3 = make_int4(__float_as_int(h1v.v[4]), __float_as_int(h1v.v[5]), __float_as_int(h1v.v[6]), __float_as_int(h1v.v[7]));
  float c0lox = fmaf(n0.x, ri.idx, -ri.oodx), c0hix = fmaf(n0.y, ri.idx, -ri.oodx);
  float c0loy = fmaf(n0.z, ri.idy, -ri.oody), c0hiy = fmaf(n0.w, ri.idy, -ri.oody);
  float c0loz = fmaf(n2.x, ri.idz, -ri.oodz), c0hiz = fmaf(n2.y, ri.idz, -ri.oodz);
  float c1lox = fmaf(n1.x, ri.idx, -ri.oodx), c1hix = fmaf(n1.y, ri.idx, -ri.oodx);
  float c1loy = fmaf(n1.z, ri.idy, -ri.oody), c1hiy = fmaf(n1.w, ri.idy, -ri.oody);
  float c1loz = fmaf(n2.z, ri.idz, -ri.oodz), c1hiz = fmaf(n2.w, ri.idz, -ri.oodz);
  float tn0 = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
  float tf0 = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), t_best));
  float tn1 = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
  float tf1 = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), t_best));
  const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
  const bool swap = tn1 < tn0;
  const int c0 = n3.x, c1 = n3.y;
  if (h0 && h1) {
    *next = swap ? c1 : c0;
    *push = swap ? c0 : c1;
    return 2;
  }
  *next = h0 ? c0 : c1;
  return (h0 || h1) ? 1 : 0;
}

// One step of the FOUR-wide tree (DNode4): tests the four child boxes of inner node `cur` against (t_min, t_best), pushes
// every hit child but the nearest on the stack (slot order: a full distance sort buys < 0.5 % fewer steps on the shipped
// scenes, profiles/r2_experiments.txt) and returns the nearest; with no hit it pops.  The stack's bottom entry ends the
// traversal (render.cu), so a pop needs no emptiness test.  Ties (a ray origin inside several boxes gives each t_min)
// go to the lowest slot, so exactly one child is "nearest".
RTB_DEV int node_step4(const DNode4* __restrict__ nodes, int cur, const RayInv& ri, float t_min, float t_best, int*& top) {
  const Ld256* nd = reinterpret_cast<const Ld256*>(nodes + cur);
  const Ld256 X = ld256(nd), Y = ld256(nd + 1), Z = ld256(nd + 2);
  const int4 C = __ldg(reinterpret_cast<const int4*>(nd + 3));
  float k[4];
  bool h[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float ax = fmaf(X.v[i], ri.idx, -ri.oodx), bx = fmaf(X.v[4 + i], ri.idx, -ri.oodx);
    const float ay = fmaf(Y.v[i], ri.idy, -ri.oody), by = fmaf(Y.v[4 + i], ri.idy, -ri.oody);
    const float az = fmaf(Z.v[i], ri.idz, -ri.oodz), bz = fmaf(Z.v[4 + i], ri.idz, -ri.oodz);
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_best));
    h[i] = tn <= tf;
    k[i] = h[i] ? tn : CUDART_INF_F;
  }
  // nearest child: a two-level tournament, ties to the lower slot (so exactly one slot is "nearest")
  const bool p01 = k[1] < k[0], p23 = k[3] < k[2];
  const bool p = fminf(k[2], k[3]) < fminf(k[0], k[1]);
  int next = p ? (p23 ? C.w : C.z) : (p01 ? C.y : C.x);
  // every other hit child, in slot order
  if (h[0] && (p01 || p)) *top++ = C.x;
  if (h[1] && (!p01 || p)) *top++ = C.y;
  if (h[2] && (p23 || !p)) *top++ = C.z;
  if (h[3] && (!p23 || !p)) *top++ = C.w;
  if (!(h[0] || h[1] || h[2] || h[3])) next = *--top;
  return next;
}

#define RTB_STACK_DEPTH 64

// Closest surface hit over (t_min, +inf): the result of world.hit(ray, 0.001, f64::MAX)
// (src/color.rs:7) restricted to surfaces.  Stack-based traversal of the two-child BVH,
// near child first, far child pushed, the query interval shrinking with every accepted hit.
// Exact ties: Rect accepts t == t_max (src/rect.rs:46) so in the reference's list scan a LATER
// planar primitive replaces an earlier one at equal t (Q23; e.g. cornell_box's two floors and
// the cube bottoms resting on them); later == higher primitive ID, which is what wins here.
#define RTB_TRAVERSAL_END (-2147483647 - 1)  // stack bottom: ends a traversal (== CUR_SHADE of the megakernel)

// The same query over the four-wide tree (DNode4), the form the megakernel's NODE / LEAF phases run.
template <bool kStats>
RTB_DEV Hit trace_closest4(const SceneView& sv, const RayF& ray, float t_min, int self_prim, int self_face, Counters* cnt,
                           float t_max) {
  Hit best;
  best.t = t_max, best.prim = -1, best.face = 0, best.id = -1;
  const float inv_a = rcp_fast(dot(ray.d, ray.d));
  const RayInv ri = make_ray_inv(ray);
  int stack[RTB_STACK_DEPTH + 1];
  stack[0] = RTB_TRAVERSAL_END;
  int* top = stack + 1;
  int cur = sv.root4;
  for (;;) {
    while (cur >= 0) {
      if (kStats) cnt->node_tests += 2;
      RTB_CHECK(cur < sv.n_nodes4, RTB_CHK_NODE);
      cur = node_step4(sv.nodes4, cur, ri, t_min, best.t, top);
      RTB_CHECK(top >= stack && top <= stack + RTB_STACK_DEPTH, RTB_CHK_STACK);
    }
    if (cur == RTB_TRAVERSAL_END) return best;
    const unsigned code = ~(unsigned)cur;
    const int first = (int)(code >> 3), count = (int)(code & 7u);
    RTB_CHECK(first >= 0 && first + count <= sv.n_prims, RTB_CHK_PRIM);
    for (int i = 0; i < count; ++i) {
      const int pi = first + i;
      float t;
      int face, id;
      if (intersect_prim<kStats>(sv.prims, pi, sv.xforms, ray, inv_a, t_min, best.t, pi == self_prim, self_face, &t, &face, &id, cnt) &&
          (t < best.t || id > best.id)) {
        best.t = t, best.prim = pi, best.face = face, best.id = id;
      }
    }
    cur = *--top;
  }
}

template <bool kStats>
RTB_DEV Hit trace_closest(const SceneView& sv, const RayF& ray, float t_min, int self_prim,
                                             int self_face, Counters* cnt, float t_max) {
  // t_max: only surfaces nearer than this are of interest (a ConstantMedium that already scattered the ray there)
  if (sv.nodes4) return trace_closest4<kStats>(sv, ray, t_min, self_prim, self_face, cnt, t_max);
  Hit best;
  best.t = t_max, best.prim = -1, best.face = 0, best.id = -1;
  const float inv_a = rcp_fast(dot(ray.d, ray.d));
  const RayInv ri = make_ray_inv(ray);
  int stack[RTB_STACK_DEPTH];
  int sp = 0;
  int cur = sv.root;
  for (;;) {
    while (cur >= 0) {
      if (kStats) cnt->node_tests++;
      RTB_CHECK(cur < sv.n_nodes, RTB_CHK_NODE);
      int next, push;
      int nh = node_step(sv.nodes, cur, ri, t_min, best.t, &next, &push);
      RTB_CHECK(sp >= 0 && sp < RTB_STACK_DEPTH, RTB_CHK_STACK);
      if (nh == 2) stack[sp++] = push;
      if (nh == 0) {
        if (sp == 0) return best;
        next = stack[--sp];
      }
      cur = next;
    }
    const unsigned code = ~(unsigned)cur;
    const int first = (int)(code >> 3), count = (int)(code & 7u);
    for (int i = 0; i < count; ++i) {
      const int pi = first + i;
      float t;
      int face, id;
      if (intersect_prim<kStats>(sv.prims, pi, sv.xforms, ray, inv_a, t_min, best.t, pi == self_prim, self_face, &t,
                                 &face, &id, cnt) &&
          (t < best.t || id > best.id)) {
        best.t = t, best.prim = pi, best.face = face, best.id = id;
      }
    }
    if (sp == 0) return best;
    cur = stack[--sp];
  }
}

template <bool kStats>
RTB_DEV Hit trace_closest(const SceneView& sv, const RayF& ray, float t_min, int self_prim, int self_face, Counters* cnt) {
  return trace_closest<kStats>(sv, ray, t_min, self_prim, self_face, cnt, CUDART_INF_F);
}

// ---- surface point ------------------------------------------------------------------------------
struct Surface {
  V3f p, n;  // world-space hit point and shading normal (FlipNormals applied)
  float u, v;
  int mat;
};

RTB_DEV void sphere_uv(V3f n, float* u, float* v) {  // src/sphere.rs:9-15 (Q5: FRAC_2_PI)
  float phi = atan2f(n.z, n.x);
  float theta = asinf(fminf(1.0f, fmaxf(-1.0f, n.y)));
  *u = 1.0f - (phi + CUDART_PI_F) * (0.5f / CUDART_PI_F);
  *v = (theta + 0.63661977236758134308f) * (1.0f / CUDART_PI_F);
}

// Fills the HitRecord fields (src/hittable.rs:9-16) of an accepted hit.  u,v are only computed
// when the material's texture chain can read them (ImageTexture); the reference computes
// atan2/asin on every sphere hit (src/sphere.rs:50,64).
RTB_DEV Surface finalize_hit(const SceneView& sv, const DPrim* __restrict__ prims, const RayF& ray,
                                                const Hit& h, int4* mat_record = nullptr, bool force_uv = false) {
  Surface s;
  RTB_CHECK(h.prim >= 0 && h.prim < sv.n_prims, RTB_CHK_PRIM);
  const float4 A = ldg4(&prims[h.prim].a);
  const float4 B = ldg4(&prims[h.prim].b);
  const int4 C = ldg4i(&prims[h.prim].c);
  const int type = RTB_PRIM_TYPE(C.x);
  const int xi = RTB_PRIM_XFORM(C.x);
  s.mat = C.y;
  RTB_CHECK(C.y >= 0 && C.y < sv.n_mats, RTB_CHK_MAT);
  s.u = 0.0f, s.v = 0.0f;
  // the material record (type, texture, parameter, needs_uv) in one load, handed on to the caller
  const int4 mrec = __ldg(reinterpret_cast<const int4*>(sv.mats + C.y));
  if (mat_record) *mat_record = mrec;
  const bool need_uv = force_uv || mrec.w != 0;  // force_uv: the field-level parity hook (debug.cu) wants every record complete
  if (type <= DP_MOVING_SPHERE) {
    V3f ctr = sphere_center(A, B, C, ray.time);
    s.p = madd(h.t, ray.d, ray.o);
    s.n = rcp_fast(A.w) * (s.p - ctr);
    if (need_uv) {
      V3f nl = s.n;
      if (xi != 0) nl = xf_vec(load_xform(sv.xforms, xi), s.n);  // uv is taken in the sphere's own frame
      sphere_uv(nl, &s.u, &s.v);
    }
  } else {
    V3f o = ray.o, d = ray.d;
    DXform x;
    if (xi != 0) {
      x = load_xform(sv.xforms, xi);
      o = xf_point(x, ray.o);
      d = xf_vec(x, ray.d);
    }
    V3f pl = madd(h.t, d, o);
    int k_axis;
    bool is_min = false;
    if (type == DP_RECT) {
      k_axis = RTB_PRIM_PLANE(C.x);
      set_comp(pl, k_axis, B.x);  // the hit point lies on the plane exactly
    } else {
      k_axis = 2 - (h.face >> 1);
      is_min = h.face & 1;
      set_comp(pl, k_axis, is_min ? comp(mk(A.x, A.y, A.z), k_axis) : comp(mk(B.x, B.y, B.z), k_axis));
    }
    if (need_uv) {  // src/rect.rs:54-55 with the face's (x0,y0,x1,y1) in its (a,b) axes
      const int aa = k_axis == 2 ? 0 : k_axis + 1, bb = k_axis == 0 ? 2 : k_axis - 1;
      float a0, b0, a1, b1;
      if (type == DP_RECT) {
        a0 = A.x, b0 = A.y, a1 = A.z, b1 = A.w;
      } else {
        a0 = comp(mk(A.x, A.y, A.z), aa), b0 = comp(mk(A.x, A.y, A.z), bb);
        a1 = comp(mk(B.x, B.y, B.z), aa), b1 = comp(mk(B.x, B.y, B.z), bb);
      }
      s.u = (comp(pl, aa) - a0) / (a1 - a0);
      s.v = (comp(pl, bb) - b0) / (b1 - b0);
    }
    V3f nl = mk(k_axis == 0 ? 1.0f : 0.0f, k_axis == 1 ? 1.0f : 0.0f, k_axis == 2 ? 1.0f : 0.0f);  // always +e_k (Q9)
    if (xi != 0) {
      s.p = xf_point_inv(x, pl);
      s.n = xf_vec_t(x, nl);
    } else {
      s.p = pl;
      s.n = nl;
    }
  }
  if (RTB_PRIM_FLIP(C.x)) s.n = -s.n;
  return s;
}

// Slab interval [t_enter, t_exit] of the ray's LINE against a BOX primitive (the six Rects of src/cube.rs:21-74 over
// (-MAX, MAX)): the two boundary hits ConstantMedium::hit asks for (src/medium.rs:30-31) in one pass.
RTB_DEV bool box_interval(const DPrim* __restrict__ prims, int pi, const int4 C, const DXform* __restrict__ xforms, const RayF& ray,
                          float* t_enter, float* t_exit) {
  const float4 A = ldg4(&prims[pi].a);
  const float4 B = ldg4(&prims[pi].b);
  V3f o = ray.o, d = ray.d;
  const int xi = RTB_PRIM_XFORM(C.x);
  if (xi != 0) {
    DXform x = load_xform(xforms, xi);
    o = xf_point(x, ray.o);
    d = xf_vec(x, ray.d);
  }
  const float ix = rcp_fast(d.x), iy = rcp_fast(d.y), iz = rcp_fast(d.z);
  const float tx0 = (A.x - o.x) * ix, tx1 = (B.x - o.x) * ix;
  const float ty0 = (A.y - o.y) * iy, ty1 = (B.y - o.y) * iy;
  const float tz0 = (A.z - o.z) * iz, tz1 = (B.z - o.z) * iz;
  *t_enter = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fminf(tz0, tz1));
  *t_exit = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fmaxf(tz0, tz1));
  return *t_enter <= *t_exit;
}

// The reference restarts the boundary query at hit1.t + 0.0001 (src/medium.rs:31).  In f32 that increment is below one
// ulp once |t| > 1024 and the second query would find the first surface again (a medium far from the ray origin
// vanished: tests/field_checks.py build_far_medium), so the restart point is at least four ulps beyond hit1.t.
RTB_DEV float medium_restart(float t1) { return t1 + fmaxf(0.0001f, fabsf(t1) * 4.8e-7f); }

// ---- constant media: src/medium.rs:28-56 -------------------------------------------------------
// Runs every medium against the ray after the surface query.  Per medium (reference order):
// hit1 = boundary.hit(-MAX, MAX); hit2 = boundary.hit(hit1.t + 1e-4, MAX) — each the closest
// boundary primitive in list order; clamp to [t_min, closest_so_far]; exponential free flight.
// Independent exponentials clipped to the running closest hit give the same distribution as
// the reference's list scan (Q13).  Returns the index of the medium that scattered, or -1.
template <bool kStats>
RTB_DEV int sample_media(const SceneView& sv, const RayF& ray, float t_min, float* t_closest,
                                            float xi0, float xi1, uint64_t seed, uint32_t pixel, uint32_t sample,
                                            uint32_t depth, Counters* cnt) {
  // xi0, xi1: the free-flight draws of media 0 and 1 (slots d, e of the bounce block); media
  // 2.. take theirs from extra Philox blocks (purpose RTB_RNG_MEDIA + g), five at a time.
  int which = -1;
  const float a = dot(ray.d, ray.d);
  const float inv_a = rcp_fast(a);
  const float inv_len = rsqrt_fast(a);
  const float len = a * inv_len;
  Rand5 extra;
  extra.a = extra.b = extra.c = extra.d = extra.e = 0.5f;
  for (int mi = 0; mi < sv.n_media; ++mi) {
    if (kStats) cnt->medium_tests++;
    float xi = mi == 0 ? xi0 : xi1;
    if (mi >= 2) {
      const int q = (mi - 2) % 5;
      if (q == 0) extra = rand5(seed, depth, pixel, sample, RTB_RNG_MEDIA + (uint32_t)((mi - 2) / 5));
      xi = q == 0 ? extra.a : (q == 1 ? extra.b : (q == 2 ? extra.c : (q == 3 ? extra.d : extra.e)));
    }
    const int4 raw = __ldg(reinterpret_cast<const int4*>(sv.media + mi));
    const int first = raw.x, count = raw.y;
    float t1 = 0.0f, t2 = 0.0f;
    bool ok = true;
    const int4 C0 = ldg4i(&sv.bprims[first].c);
    if (count == 1 && RTB_PRIM_TYPE(C0.x) <= DP_MOVING_SPHERE) {
      // a sphere boundary: hit1 is the near root over (-MAX, MAX), hit2 the far root if it
      // lies beyond hit1.t + 1e-4 (src/sphere.rs:45-62 called twice by src/medium.rs:30-31)
      if (kStats) cnt->prim_tests += 2;
      const float4 A = ldg4(&sv.bprims[first].a);
      const float4 B = ldg4(&sv.bprims[first].b);
      V3f oc = ray.o - sphere_center(A, B, C0, ray.time);
      float tc = -dot(oc, ray.d) * inv_a;
      V3f l = madd(tc, ray.d, oc);
      float disc = fmaf(A.w, A.w, -dot(l, l));
      float h = sqrt_fast(fmaxf(disc, 0.0f) * inv_a);
      t1 = tc - h, t2 = tc + h;
      ok = disc > 0.0f && t2 > medium_restart(t1);
    } else if (count == 1 && RTB_PRIM_TYPE(C0.x) == DP_BOX) {
      // a Cube boundary (cornell_smoke): entry and exit face of one slab pass; the second query accepts
      // t >= hit1.t + 1e-4 (Rect bounds are inclusive, src/rect.rs:46)
      if (kStats) cnt->prim_tests += 2;
      ok = box_interval(sv.bprims, first, C0, sv.xforms, ray, &t1, &t2) && t2 >= medium_restart(t1);
    } else {
#pragma unroll 1
    for (int pass = 0; pass < 2 && ok; ++pass) {
      const float lo = pass == 0 ? -CUDART_INF_F : medium_restart(t1);
      float best = CUDART_INF_F;
      bool any = false;
      for (int i = 0; i < count; ++i) {
        float t;
        int face, id;
        if (intersect_prim<kStats>(sv.bprims, first + i, sv.xforms, ray, inv_a, lo, best, false, 0, &t, &face, &id, cnt)) {
          best = t;
          any = true;
        }
      }
      ok = any;
      if (pass == 0) t1 = best; else t2 = best;
    }
    }
    if (!ok) continue;
    if (t1 < t_min) t1 = t_min;
    if (t2 > *t_closest) t2 = *t_closest;
    if (t1 < t2) {
      float distance_inside = (t2 - t1) * len;
      float hit_distance = __int_as_float(raw.w) * __logf(xi);  // -(1/density) * ln(xi)
      if (hit_distance < distance_inside) {
        *t_closest = fmaf(hit_distance, inv_len, t1);
        which = mi;
      }
    }
  }
  return which;
}

// ---- textures: src/texture.rs, src/perlin.rs --------------------------------------------------
RTB_DEV float perlin_noise(const DPerlin* __restrict__ P, V3f p) {  // src/perlin.rs:76-97
  float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  float u = p.x - fx, v = p.y - fy, w = p.z - fz;
  // `floor(x) as usize`: negatives (and NaN) saturate to 0 (Q18); large values wrap with & 255
  // exactly like the usize arithmetic as long as they fit 2^31, which scene coordinates do
  int i = fx > 0.0f ? (int)fminf(fx, 2147483520.0f) : 0;
  int j = fy > 0.0f ? (int)fminf(fy, 2147483520.0f) : 0;
  int k = fz > 0.0f ? (int)fminf(fz, 2147483520.0f) : 0;
  float uu = u * u * (3.0f - 2.0f * u);
  float vv = v * v * (3.0f - 2.0f * v);
  float ww = w * w * (3.0f - 2.0f * w);
  unsigned px0 = __ldg(&P->perm_x[i & 255]), px1 = __ldg(&P->perm_x[(i + 1) & 255]);
  unsigned py0 = __ldg(&P->perm_y[j & 255]), py1 = __ldg(&P->perm_y[(j + 1) & 255]);
  unsigned pz0 = __ldg(&P->perm_z[k & 255]), pz1 = __ldg(&P->perm_z[(k + 1) & 255]);
  float accum = 0.0f;
#pragma unroll
  for (int di = 0; di < 2; ++di)
#pragma unroll
    for (int dj = 0; dj < 2; ++dj)
#pragma unroll
      for (int dk = 0; dk < 2; ++dk) {
        unsigned idx = (di ? px1 : px0) ^ (dj ? py1 : py0) ^ (dk ? pz1 : pz0);
        float4 c = ldg4(&P->ranvec[idx]);
        float wgt = (di ? uu : 1.0f - uu) * (dj ? vv : 1.0f - vv) * (dk ? ww : 1.0f - ww);
        accum = fmaf(wgt, fmaf(c.x, u - (float)di, fmaf(c.y, v - (float)dj, c.z * (w - (float)dk))), accum);
      }
  return accum;
}

RTB_DEV float perlin_turb(const DPerlin* __restrict__ P, V3f p) {  // src/perlin.rs:99-109, depth 7
  float accum = 0.0f, weight = 1.0f;
#pragma unroll 1
  for (int d = 0; d < 7; ++d) {
    accum = fmaf(weight, perlin_noise(P, p), accum);
    weight *= 0.5f;
    p = 2.0f * p;
  }
  return fabsf(accum);
}

struct TexCounters {
  unsigned long long n[4];
};

// sign(sin(10 x)) == parity of floor(10 x / pi): the checker only needs the sign of the
// product of three sines (src/texture.rs:41-42)
RTB_DEV int sin10_negative(float x) { return ((int)floorf(x * 3.1830988618379067f)) & 1; }

template <bool kStats>
RTB_DEV V3f texture_value(const SceneView& sv, int tex, float u, float v, V3f p, TexCounters* tc) {
  for (;;) {
    // both halves of the 32-byte record at once: the colour load does not wait for the type test
    RTB_CHECK(tex >= 0 && tex < sv.n_texs, RTB_CHK_TEX);
    const int4 hd = __ldg(reinterpret_cast<const int4*>(sv.texs + tex));
    const float4 col = __ldg(reinterpret_cast<const float4*>(sv.texs + tex) + 1);
    const int type = hd.x;
    if (kStats) tc->n[type & 3]++;
    if (type == RTB_TEX_CHECKER) {  // src/texture.rs:40-47: product of three sines < 0 -> `odd`
      const int neg = sin10_negative(p.x) ^ sin10_negative(p.y) ^ sin10_negative(p.z);
      tex = neg ? hd.y : hd.z;
      continue;
    }
    if (type == RTB_TEX_SOLID) return mk(col.x, col.y, col.z);
    if (type == RTB_TEX_NOISE) {  // src/texture.rs:66-70
      float t = perlin_turb(sv.perlin + hd.y, p);
      float g = 0.5f * (1.0f + sin_reduced(fmaf(col.w, p.x, 5.0f * t)));
      return mk(g, g, g);
    }
    // RTB_TEX_IMAGE, src/texture.rs:87-107: saturating casts, clamp high, nearest texel
    const int nx = hd.y, ny = hd.z;
    float fi = u * (float)nx, fj = (1.0f - v) * (float)ny;
    int i = fi > 0.0f ? (int)fminf(fi, 2147483520.0f) : 0;
    int j = fj > 0.0f ? (int)fminf(fj, 2147483520.0f) : 0;
    if (i > nx - 1) i = nx - 1;
    if (j > ny - 1) j = ny - 1;
    RTB_CHECK(i >= 0 && j >= 0 && hd.w >= 0 && hd.w + i + nx * j < sv.n_texels, RTB_CHK_TEXEL);
    uint32_t texel = __ldg(&sv.texels[hd.w + i + nx * j]);
    const float k = 1.0f / 255.0f;
    return mk((float)(texel & 255u) * k, (float)((texel >> 8) & 255u) * k, (float)((texel >> 16) & 255u) * k);
  }
}

// ---- materials: src/material.rs ------------------------------------------------------------------
RTB_DEV V3f reflect(V3f v, V3f n) { return v - (2.0f * dot(v, n)) * n; }  // :9-11

// Material::scatter.  Returns false when the path ends (DiffuseLight, absorbed Metal).
// On success `ray` becomes the scattered ray (origin = hit.p, same time) and *atten the albedo.
// rnd: the four draws of this bounce (ball x3, schlick).
template <bool kStats>
RTB_DEV bool scatter(const SceneView& sv, int mat_type, int mat_tex, float mat_param, const Surface& s,
                                        RayF& ray, const Rand5& rnd, V3f* atten, TexCounters* tc) {
  if (mat_type == RTB_MAT_DIFFUSE_LIGHT) return false;  // :144-146
  if (mat_type == RTB_MAT_DIELECTRIC) {                 // :106-126
    *atten = mk(1.0f, 1.0f, 1.0f);
    const float ref_idx = mat_param;
    const float ddn = dot(ray.d, s.n);
    const float inv_len = rsqrt_fast(dot(ray.d, ray.d));
    V3f outward_normal;
    float ni_over_nt, cosine;
    if (ddn > 0.0f) {
      cosine = ref_idx * ddn * inv_len;
      outward_normal = -s.n;
      ni_over_nt = ref_idx;
    } else {
      cosine = -ddn * inv_len;
      outward_normal = s.n;
      ni_over_nt = rcp_fast(ref_idx);
    }
    // refract(), :13-23
    V3f uv = inv_len * ray.d;
    float dt = dot(uv, outward_normal);
    float discriminant = 1.0f - ni_over_nt * ni_over_nt * (1.0f - dt * dt);
    V3f dir = reflect(ray.d, s.n);  // un-normalised d, unflipped hit.normal (Q14)
    if (discriminant > 0.0f) {
      float r0 = (1.0f - ref_idx) * rcp_fast(1.0f + ref_idx);  // schlick(), :25-28
      r0 = r0 * r0;
      float m = 1.0f - cosine;
      float m2 = m * m;
      float reflect_prob = r0 + (1.0f - r0) * (m2 * m2 * m);
      if (rnd.a >= reflect_prob) dir = ni_over_nt * (uv - dt * outward_normal) - sqrt_fast(discriminant) * outward_normal;
    }
    ray.o = s.p, ray.d = dir;
    return true;
  }
  // Lambertian :49-53, Metal :76-86, Isotropic :165-168 share the ball sample and the albedo
  V3f ball = sample_unit_ball(rnd.a, rnd.b, rnd.c);
  V3f dir;
  if (mat_type == RTB_MAT_LAMBERTIAN) {
    dir = s.n + ball;  // target - p = normal + ball
  } else if (mat_type == RTB_MAT_METAL) {
    float inv_len = rsqrt_fast(dot(ray.d, ray.d));
    dir = reflect(inv_len * ray.d, s.n);
    if (mat_param > 0.0f) dir = madd(mat_param, ball, dir);
    if (!(dot(dir, s.n) > 0.0f)) return false;
  } else {
    dir = ball;
  }
  *atten = texture_value<kStats>(sv, mat_tex, s.u, s.v, s.p, tc);
  ray.o = s.p, ray.d = dir;
  return true;
}

// Material::scatter without the albedo lookup (the caller evaluates the material's texture once
// for both `emitted` and the attenuation): returns false when the path ends, else `ray` becomes
// the scattered ray.  Same arithmetic as scatter() above.
RTB_DEV bool scatter_dir(int mat_type, float mat_param, const Surface& s, RayF& ray, const Rand5& rnd) {
  if (mat_type == RTB_MAT_DIFFUSE_LIGHT) return false;  // src/material.rs:144-146
  if (mat_type == RTB_MAT_DIELECTRIC) {                 // :106-126
    const float ref_idx = mat_param;
    const float ddn = dot(ray.d, s.n);
    const float inv_len = rsqrt_fast(dot(ray.d, ray.d));
    V3f outward_normal;
    float ni_over_nt, cosine;
    if (ddn > 0.0f) {
      cosine = ref_idx * ddn * inv_len;
      outward_normal = -s.n;
      ni_over_nt = ref_idx;
    } else {
      cosine = -ddn * inv_len;
      outward_normal = s.n;
      ni_over_nt = rcp_fast(ref_idx);
    }
    V3f uv = inv_len * ray.d;  // refract(), :13-23
    float dt = dot(uv, outward_normal);
    float discriminant = 1.0f - ni_over_nt * ni_over_nt * (1.0f - dt * dt);
    V3f dir = reflect(ray.d, s.n);  // un-normalised d, unflipped hit.normal (Q14)
    if (discriminant > 0.0f) {
      float r0 = (1.0f - ref_idx) * rcp_fast(1.0f + ref_idx);  // schlick(), :25-28
      r0 = r0 * r0;
      float m = 1.0f - cosine;
      float m2 = m * m;
      float reflect_prob = r0 + (1.0f - r0) * (m2 * m2 * m);
      if (rnd.a >= reflect_prob) dir = ni_over_nt * (uv - dt * outward_normal) - sqrt_fast(discriminant) * outward_normal;
    }
    ray.o = s.p, ray.d = dir;
    return true;
  }
  V3f ball = sample_unit_ball(rnd.a, rnd.b, rnd.c);  // Lambertian :49-53, Metal :76-86, Isotropic :165-168
  V3f dir;
  if (mat_type == RTB_MAT_LAMBERTIAN) {
    dir = s.n + ball;
  } else if (mat_type == RTB_MAT_METAL) {
    float inv_len = rsqrt_fast(dot(ray.d, ray.d));
    dir = reflect(inv_len * ray.d, s.n);
    if (mat_param > 0.0f) dir = madd(mat_param, ball, dir);
    if (!(dot(dir, s.n) > 0.0f)) return false;
  } else {
    dir = ball;
  }
  ray.o = s.p, ray.d = dir;
  return true;
}

// ---- camera: src/camera.rs:53-67 ------------------------------------------------------------------
// Draws: block 0 = (jitter_u, jitter_v, lens_r, lens_phi), block 1 = (time).  The lens sample is
// the uniform unit disk (r = sqrt(xi), phi = 2 pi xi) — same distribution as the rejection loop
// at src/util.rs:15-24; skipped iff lens_radius == 0 (src/camera.rs:54).  The time draw always
// happens, as upstream (src/camera.rs:61).
template <class Seed>  // uint64_t seed, or the PhiloxKeys made from it
RTB_DEV RayF camera_ray(const DCamera& cam, int i, int j, int nx, int ny, const Seed& seed, uint32_t pixel,
                                           uint32_t sample) {
  const Rand5 r5 = rand5(seed, 0u, pixel, sample, RTB_RNG_CAMERA);
  const float ju = r5.a, jv = r5.b, lr = r5.c, lp = r5.d, xt = r5.e;
  float s = ((float)i + ju) * rcp_fast((float)nx);  // tests/test.rs:66-67
  float t = ((float)j + jv) * rcp_fast((float)ny);
  V3f origin = mk(cam.origin[0], cam.origin[1], cam.origin[2]);
  if (cam.lens_radius != 0.0f) {
    float r = cam.lens_radius * sqrt_fast(lr);
    float sn, cs;
    __sincosf((2.0f * lp - 1.0f) * CUDART_PI_F, &sn, &cs);
    float rx = r * cs, ry = r * sn;
    origin = mk(fmaf(cam.u[0], rx, fmaf(cam.v[0], ry, origin.x)), fmaf(cam.u[1], rx, fmaf(cam.v[1], ry, origin.y)),
                fmaf(cam.u[2], rx, fmaf(cam.v[2], ry, origin.z)));
  }
  RayF ray;
  ray.o = origin;
  ray.d = mk(cam.llc[0] + s * cam.horizontal[0] + t * cam.vertical[0] - origin.x,
             cam.llc[1] + s * cam.horizontal[1] + t * cam.vertical[1] - origin.y,
             cam.llc[2] + s * cam.horizontal[2] + t * cam.vertical[2] - origin.z);
  ray.time = cam.time0 + xt * (cam.time1 - cam.time0);
  return ray;
}

RTB_DEV V3f background_color(int background, V3f d) {
  if (background == 1) {  // the gradient commented out at src/color.rs:18-20 (lit variant only)
    float inv = rsqrt_fast(dot(d, d));
    float t = 0.5f * (d.y * inv + 1.0f);
    return mk((1.0f - t) + t * 0.5f, (1.0f - t) + t * 0.7f, (1.0f - t) + t * 1.0f);
  }
  return mk(0.0f, 0.0f, 0.0f);  // src/color.rs:21
}

}  // namespace rtb
