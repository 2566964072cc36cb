// device_types.h — flat, GPU-resident scene tables (SoA-of-records) produced by flatten.cpp
// from the rtb_scene object graph and consumed by the sm_100a kernels in render.cu.
//
// The reference walks a tree of boxed trait objects per ray (src/hittable.rs:37-47,
// src/bvh.rs:70-89).  Here the tree is compiled away at upload time:
//   * every leaf shape becomes one 48-byte DPrim (3 x 128-bit loads) carrying its composed
//     instance transform index, its FlipNormals parity and its material;
//   * Rotate/Traslate chains (src/rotate.rs:85-113, src/traslate.rs:18-24) compose into one
//     rigid transform  local = M * world + b ; spheres are pre-transformed to world space;
//   * a Cube (six Rects, src/cube.rs:21-74) is one BOX primitive solved by a slab test;
//   * all surfaces share one SAH BVH of 64-byte two-child nodes (4 x 128-bit loads);
//   * every ConstantMedium (src/medium.rs:28-56) is a DMedium with its own boundary
//     primitive range, evaluated after the surface query.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RTB_HD __host__ __device__
#else
#define RTB_HD
#endif

namespace rtb {

struct f4 {
  float x, y, z, w;
};
struct i4 {
  int x, y, z, w;
};

enum DPrimType { DP_SPHERE = 0, DP_MOVING_SPHERE = 1, DP_RECT = 2, DP_BOX = 3 };

// flags word: bits 0-2 type, bits 3-4 rect plane (RtbPlane), bit 5 FlipNormals parity,
// bits 16-31 transform index (0 = identity)
#define RTB_PRIM_TYPE(f) ((f)&7)
#define RTB_PRIM_PLANE(f) (((f) >> 3) & 3)
#define RTB_PRIM_FLIP(f) (((f) >> 5) & 1)
#define RTB_PRIM_XFORM(f) ((int)((unsigned)(f) >> 16))

struct alignas(16) DPrim {
  // SPHERE: a = (cx,cy,cz,r) in WORLD space
  // MOVING_SPHERE: a = (c0.xyz, r), b = (c1-c0 .xyz, time0), c.w = bits of 1/(time1-time0)
  // RECT: a = (x0,y0,x1,y1), b.x = k            (local space of its transform)
  // BOX:  a = (pmin.xyz, -), b = (pmax.xyz, -)  (local space of its transform)
  f4 a, b;
  i4 c;  // (flags, material, first primitive ID, aux)
};

// 64-byte BVH node holding the boxes of both children, fetched as two 256-bit loads per visit
// (LDG.E.256 on sm_100a: half the L1 wavefronts of four 128-bit loads, and the L1 data pipe is the
// busiest unit of the traversal, profiles/).
struct alignas(32) DNode {
  f4 n0;  // child0: lo.x hi.x lo.y hi.y
  f4 n1;  // child1: lo.x hi.x lo.y hi.y
  f4 n2;  // child0 lo.z hi.z, child1 lo.z hi.z
  i4 n3;  // child0, child1 (>= 0 inner node index; < 0 leaf: ~v = first_prim << 3 | count, count 0..7), -, -
};

// Four-wide node (the two-child SAH tree collapsed by surface area, flatten.cpp): 128 bytes = three 256-bit loads of
// plane coordinates (x, y, z: the four children's lo then hi) + one 128-bit load of child references.  About half the
// traversal steps of the two-child tree for the same number of box tests (profiles/r2_experiments.txt).
struct alignas(32) DNode4 {
  float lox[4], hix[4];
  float loy[4], hiy[4];
  float loz[4], hiz[4];
  int child[4];  // >= 0: DNode4 index; < 0: leaf code as in DNode; unused slots: an empty leaf (~0) behind a box no ray meets
  int pad[4];
};

struct alignas(16) DXform {
  float m[9];  // row-major M:  local = M * world + b
  float b[3];
};

struct alignas(16) DMaterial {
  int type;  // RtbMatType
  int tex;
  float param;  // metal: fuzz (already clamped, src/material.rs:70); dielectric: ref_idx
  int needs_uv;  // the texture chain can reach an ImageTexture, the only reader of (u,v)
};

struct alignas(16) DTexture {
  int type;  // RtbTexType
  int a, b;  // checker: odd, even; noise: perlin table; image: nx, ny
  int off;   // image: first texel (RGBA8 units) in the texel pool
  float r, g, bl, scale;
};

struct alignas(16) DPerlin {
  f4 ranvec[256];
  uint8_t perm_x[256], perm_y[256], perm_z[256];
};

struct alignas(16) DMedium {
  int first, count;  // boundary primitive range in `bprims`
  int mat;           // Isotropic material (src/medium.rs:22)
  float neg_inv_density;
};

// What the kernels see (passed by value as a kernel parameter).
struct SceneView {
  const DNode* nodes;
  const DNode4* nodes4;  // the same tree four-wide (nullptr when only the two-child tree exists: device-built LBVH)
  const DPrim* prims;    // surfaces, BVH leaf order
  const DPrim* bprims;   // medium boundaries
  const DMedium* media;
  const DXform* xforms;
  const DMaterial* mats;
  const DTexture* texs;
  const DPerlin* perlin;
  const uint32_t* texels;  // RGBA8
  int n_nodes, n_prims, n_media, root;  // root: node index (always an inner node)
  int root4;                            // root of nodes4
  int n_nodes4, n_mats, n_texs, n_texels;  // table sizes (the RTB_DEBUG_CHECKS build compares every index with them)
  int flat;  // != 0: the megakernel scans `prims` instead of walking a tree (scenes of a handful of primitives)
};

// The primitives of a flat-scan scene (k_render_mega kTree = 2) as a kernel parameter: read through the constant bank, grouped
// by class so that each group's loop tests one known kind of primitive (trace.cuh PrimClass): [0, end[0]) static spheres,
// [end[0], end[1]) / [end[1], end[2]) / [end[2], end[3]) untransformed YZ / ZX / XY rects, [end[3], n) everything else.
#define RTB_FLAT_MAX_PRIMS 16
struct FlatTables {
  int n;
  int end[4];
  int index[RTB_FLAT_MAX_PRIMS];  // position of each record in SceneView::prims (what Hit::prim / self_prim name)
  DPrim prims[RTB_FLAT_MAX_PRIMS];
};

struct DCamera {  // src/camera.rs:8-18, f32
  float origin[3], llc[3], horizontal[3], vertical[3], u[3], v[3];
  float time0, time1, lens_radius;
};

}  // namespace rtb
