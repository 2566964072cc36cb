// flatten.hpp — object graph -> flat device tables (host side, no CUDA).
#pragma once
#include <string>
#include <vector>

#include "device_types.h"
#include "scene.hpp"

namespace rtb {

struct FlatScene {
  std::vector<DNode> nodes;
  std::vector<DNode4> nodes4;  // the same tree collapsed to four children per node (empty when the device builds the BVH)
  int root4 = 0;
  std::vector<DPrim> prims;   // BVH leaf order
  std::vector<DPrim> bprims;  // medium boundaries, grouped per medium
  std::vector<DMedium> media;
  std::vector<DXform> xforms;  // [0] = identity
  std::vector<DMaterial> mats;
  std::vector<DTexture> texs;
  std::vector<DPerlin> perlin;
  std::vector<uint32_t> texels;
  std::vector<float> prim_boxes;  // n x (lo.xyz, hi.xyz), only when the BVH is left to the device builder
  int root = -1;
  int n_prim_ids = 0;
  int bvh_depth = 0;
  double time_lo = 0.0, time_hi = 1.0;  // shutter range the moving-sphere bounds cover
  bool has_moving = false;
  // >= 4 of the six faces of the scene's bounding box are covered by one thin primitive each (a room: cornell_box).  Rays
  // inside such a scene always hit something, which moves the point where scanning all primitives beats walking a tree.
  bool enclosed = false;
};

// Returns 0 or an RtbStatus error (message via set_error).
// host_bvh = false leaves `nodes` empty (root = -1) and fills `prim_boxes` for the device builder.
int flatten_scene(const rtb_scene& s, int world, double time_lo, double time_hi, FlatScene* out, bool host_bvh = true);

}  // namespace rtb
