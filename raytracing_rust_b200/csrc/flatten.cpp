// flatten.cpp — compiles the reference-shaped object graph (Box<dyn Hittable> tree,
// src/hittable.rs:18-21) into the flat device tables of device_types.h and builds the
// SAH BVH the kernels traverse.  Host-only; runs once per scene upload.
//
// What replaces what:
//   HittableList::hit / BVHNode::hit (src/hittable.rs:37-47, src/bvh.rs:70-89) -> one BVH over
//     every surface primitive; closest hit is independent of the tree shape, so the median
//     split on a random axis (src/bvh.rs:17-66) is replaced by a full-sweep SAH build.
//   Rotate / Traslate wrappers -> composed rigid transform per primitive.
//   FlipNormals wrappers -> one parity bit per primitive.
//   ConstantMedium -> DMedium + boundary primitive range.
// Primitive IDs follow the depth-first insertion order of the graph (a Cube owns six, in the
// order of src/cube.rs:21-74), exactly as the oracle numbers them.
#include "flatten.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <cstdio>
#include <limits>

namespace rtb {
namespace {

struct Xf {  // local = M * world + b   (double on the host)
  double m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double b[3] = {0, 0, 0};
  bool identity = true;
};

struct Box3 {
  double lo[3] = {1e300, 1e300, 1e300};
  double hi[3] = {-1e300, -1e300, -1e300};
  void grow(const double p[3]) {
    for (int i = 0; i < 3; ++i) lo[i] = std::min(lo[i], p[i]), hi[i] = std::max(hi[i], p[i]);
  }
  void grow(const Box3& o) {
    for (int i = 0; i < 3; ++i) lo[i] = std::min(lo[i], o.lo[i]), hi[i] = std::max(hi[i], o.hi[i]);
  }
  double area() const {
    double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0) return 0.0;
    return 2.0 * (dx * dy + dy * dz + dz * dx);
  }
};

void to_world(const Xf& x, const double pl[3], double pw[3]) {  // world = M^T (local - b)
  double q[3] = {pl[0] - x.b[0], pl[1] - x.b[1], pl[2] - x.b[2]};
  for (int i = 0; i < 3; ++i) pw[i] = x.m[0 * 3 + i] * q[0] + x.m[1 * 3 + i] * q[1] + x.m[2 * 3 + i] * q[2];
}

struct Walker {
  const rtb_scene& s;
  FlatScene& out;
  double time_lo, time_hi;
  std::vector<Box3> prim_boxes;  // world-space bounds of out.prims (pre-reorder)
  std::string err;

  int add_xform(const Xf& x) {
    if (x.identity) return 0;
    DXform d;
    for (int i = 0; i < 9; ++i) d.m[i] = (float)x.m[i];
    for (int i = 0; i < 3; ++i) d.b[i] = (float)x.b[i];
    // dedupe against the most recent entry (instances arrive consecutively)
    if (out.xforms.size() > 1 && std::memcmp(&out.xforms.back(), &d, sizeof d) == 0) return (int)out.xforms.size() - 1;
    out.xforms.push_back(d);
    if (out.xforms.size() > 65535) err = "too many distinct instance transforms";
    return (int)out.xforms.size() - 1;
  }

  void emit(const DPrim& p, const Box3& box, bool boundary) {
    if (boundary) {
      out.bprims.push_back(p);
    } else {
      out.prims.push_back(p);
      prim_boxes.push_back(box);
    }
  }

  // returns false on error
  bool walk(int h, const Xf& x, int flip, bool boundary) {
    const RtbObjRec& o = s.obj[h];
    switch (o.type) {
      case RTB_OBJ_SPHERE: {
        double cw[3];
        to_world(x, o.v, cw);
        DPrim p{};
        p.a = {(float)cw[0], (float)cw[1], (float)cw[2], (float)o.v[3]};
        p.c.x = DP_SPHERE | (flip << 5) | (add_xform(x) << 16);
        p.c.y = o.mat;
        p.c.z = out.n_prim_ids++;
        Box3 b;
        double r = std::fabs(o.v[3]);
        for (int i = 0; i < 3; ++i) b.lo[i] = cw[i] - r, b.hi[i] = cw[i] + r;
        emit(p, b, boundary);
        return true;
      }
      case RTB_OBJ_MOVING_SPHERE: {
        double c0[3], c1[3];
        to_world(x, o.v, c0);
        to_world(x, o.v + 3, c1);
        double t0 = o.v[6], t1 = o.v[7], r = std::fabs(o.v[8]);
        DPrim p{};
        p.a = {(float)c0[0], (float)c0[1], (float)c0[2], (float)o.v[8]};
        p.b = {(float)(c1[0] - c0[0]), (float)(c1[1] - c0[1]), (float)(c1[2] - c0[2]), (float)t0};
        float inv_dt = (float)(1.0 / (t1 - t0));
        std::memcpy(&p.c.w, &inv_dt, 4);
        p.c.x = DP_MOVING_SPHERE | (flip << 5) | (add_xform(x) << 16);
        p.c.y = o.mat;
        p.c.z = out.n_prim_ids++;
        Box3 b;  // centre is linear in time: bound it at both ends of the shutter range
        for (double tm : {time_lo, time_hi}) {
          double f = (tm - t0) / (t1 - t0);
          for (int i = 0; i < 3; ++i) {
            double c = c0[i] + f * (c1[i] - c0[i]);
            b.lo[i] = std::min(b.lo[i], c - r), b.hi[i] = std::max(b.hi[i], c + r);
          }
        }
        out.has_moving = true;
        emit(p, b, boundary);
        return true;
      }
      case RTB_OBJ_RECT: {
        DPrim p{};
        p.a = {(float)o.v[0], (float)o.v[1], (float)o.v[2], (float)o.v[3]};
        p.b = {(float)o.v[4], 0, 0, 0};
        p.c.x = DP_RECT | (o.sel << 3) | (flip << 5) | (add_xform(x) << 16);
        p.c.y = o.mat;
        p.c.z = out.n_prim_ids++;
        int k_axis, a_axis, b_axis;  // src/rect.rs:40-44
        switch (o.sel) {
          case RTB_PLANE_YZ: k_axis = 0, a_axis = 1, b_axis = 2; break;
          case RTB_PLANE_ZX: k_axis = 1, a_axis = 2, b_axis = 0; break;
          default: k_axis = 2, a_axis = 0, b_axis = 1; break;
        }
        Box3 b;  // true bounds (the reference's Rect::bounding_box ignores `plane`, Q8)
        for (int i = 0; i < 4; ++i) {
          double pl[3], pw[3];
          pl[k_axis] = o.v[4];
          pl[a_axis] = (i & 1) ? o.v[2] : o.v[0];
          pl[b_axis] = (i & 2) ? o.v[3] : o.v[1];
          to_world(x, pl, pw);
          b.grow(pw);
        }
        emit(p, b, boundary);
        return true;
      }
      case RTB_OBJ_CUBE: {
        DPrim p{};
        p.a = {(float)o.v[0], (float)o.v[1], (float)o.v[2], 0};
        p.b = {(float)o.v[3], (float)o.v[4], (float)o.v[5], 0};
        p.c.x = DP_BOX | (flip << 5) | (add_xform(x) << 16);
        p.c.y = o.mat;
        p.c.z = out.n_prim_ids;
        out.n_prim_ids += 6;
        Box3 b;
        for (int i = 0; i < 8; ++i) {
          double pl[3] = {(i & 1) ? o.v[3] : o.v[0], (i & 2) ? o.v[4] : o.v[1], (i & 4) ? o.v[5] : o.v[2]};
          double pw[3];
          to_world(x, pl, pw);
          b.grow(pw);
        }
        emit(p, b, boundary);
        return true;
      }
      case RTB_OBJ_FLIP_NORMALS: return walk(o.child, x, flip ^ 1, boundary);
      case RTB_OBJ_TRANSLATE: {  // src/traslate.rs:19: child ray origin = origin - offset
        Xf c = x;
        for (int i = 0; i < 3; ++i) c.b[i] -= o.v[i];
        c.identity = false;
        return walk(o.child, c, flip, boundary);
      }
      case RTB_OBJ_ROTATE: {  // src/rotate.rs:85-96
        int a_axis, b_axis;
        switch (o.sel) {
          case RTB_AXIS_X: a_axis = 1, b_axis = 2; break;
          case RTB_AXIS_Y: a_axis = 2, b_axis = 0; break;
          default: a_axis = 0, b_axis = 1; break;
        }
        double radians = (M_PI / 180.0) * o.v[0];
        double sn = std::sin(radians), cs = std::cos(radians);
        double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        R[a_axis * 3 + a_axis] = cs, R[a_axis * 3 + b_axis] = sn;
        R[b_axis * 3 + a_axis] = -sn, R[b_axis * 3 + b_axis] = cs;
        Xf c;
        c.identity = false;
        for (int i = 0; i < 3; ++i) {
          for (int j = 0; j < 3; ++j) {
            double acc = 0;
            for (int k = 0; k < 3; ++k) acc += R[i * 3 + k] * x.m[k * 3 + j];
            c.m[i * 3 + j] = acc;
          }
          c.b[i] = R[i * 3 + 0] * x.b[0] + R[i * 3 + 1] * x.b[1] + R[i * 3 + 2] * x.b[2];
        }
        return walk(o.child, c, flip, boundary);
      }
      case RTB_OBJ_CONSTANT_MEDIUM: {
        if (boundary) {
          err = "ConstantMedium nested inside a medium boundary is not supported";
          return false;
        }
        DMedium m{};
        m.first = (int)out.bprims.size();
        if (!walk(o.child, x, 0, true)) return false;
        m.count = (int)out.bprims.size() - m.first;
        DMaterial iso{};  // Isotropic::new(texture), src/medium.rs:22
        iso.type = RTB_MAT_ISOTROPIC, iso.tex = o.tex;  // medium hits carry u = v = 0 (src/medium.rs:46-47)
        out.mats.push_back(iso);
        m.mat = (int)out.mats.size() - 1;
        m.neg_inv_density = (float)(-1.0 / o.v[0]);
        out.media.push_back(m);
        return true;
      }
      case RTB_OBJ_LIST:
      case RTB_OBJ_BVH:
        for (int k = 0; k < o.list_n; ++k)
          if (!walk(s.list_pool[o.list_off + k], x, flip, boundary)) return false;
        return true;
    }
    err = "unknown object type";
    return false;
  }
};

}  // namespace
static int tex_needs_uv(const rtb_scene& s, int t) {
  const RtbTexRec& r = s.tex[t];
  if (r.type == RTB_TEX_IMAGE) return 1;
  if (r.type == RTB_TEX_CHECKER) return tex_needs_uv(s, r.a) | tex_needs_uv(s, r.b);
  return 0;
}
namespace {

// ---- SAH BVH ------------------------------------------------------------------------------------
struct BuildNode {
  Box3 box;
  int left = -1, right = -1;  // BuildNode indices
  int first = 0, count = 0;   // leaf: range in the ordered primitive list
};

// Full-sweep SAH over the three axes on presorted lists: every axis list is sorted once by
// (centroid key, primitive index) and stably partitioned at each split, so a node costs O(count)
// and the whole build O(n log n) (1 400 primitives: 0.3 ms; the per-node re-sorting form took 4 ms).
struct Builder {
  const std::vector<Box3>& boxes;
  std::vector<int> order;  // primitive indices in leaf order (axis-0 list at the end of the build)
  std::vector<BuildNode> nodes;
  int max_leaf = 1;  // single-primitive leaves measured best (profiles/r1_bvh_sweep.txt)
  double c_trav = 1.0;  // cost of a node visit in units of one primitive test
  int depth_max = 0;
  std::vector<int> sorted[3], scratch;
  std::vector<double> right_area;
  std::vector<uint8_t> goes_left;

  void prepare() {
    const int n = (int)boxes.size();
    std::vector<double> key(n);
    for (int axis = 0; axis < 3; ++axis) {
      for (int i = 0; i < n; ++i) key[i] = boxes[i].lo[axis] + boxes[i].hi[axis];
      sorted[axis].resize(n);
      for (int i = 0; i < n; ++i) sorted[axis][i] = i;
      std::sort(sorted[axis].begin(), sorted[axis].end(), [&](int a, int b) { return key[a] < key[b] || (key[a] == key[b] && a < b); });
    }
    scratch.resize(n), right_area.resize(n), goes_left.assign(n, 0);
  }

  int build(int first, int count, int depth) {
    depth_max = std::max(depth_max, depth);
    BuildNode n;
    for (int i = 0; i < count; ++i) n.box.grow(boxes[sorted[0][first + i]]);
    int idx = (int)nodes.size();
    nodes.push_back(n);
    if (count == 1) {
      nodes[idx].first = first, nodes[idx].count = count;
      return idx;
    }
    double best = std::numeric_limits<double>::infinity();
    int best_axis = -1, best_split = -1;
    for (int axis = 0; axis < 3; ++axis) {
      const int* list = sorted[axis].data() + first;
      Box3 acc;
      for (int i = count - 1; i > 0; --i) {
        acc.grow(boxes[list[i]]);
        right_area[i] = acc.area();
      }
      Box3 left;
      for (int i = 1; i < count; ++i) {
        left.grow(boxes[list[i - 1]]);
        double cost = left.area() * i + right_area[i] * (count - i);
        if (cost < best) best = cost, best_axis = axis, best_split = i;
      }
    }
    double parent_area = std::max(n.box.area(), 1e-300);
    double split_cost = c_trav + best / parent_area;  // traversal step + expected leaf tests
    if (count <= max_leaf && split_cost >= (double)count) {
      nodes[idx].first = first, nodes[idx].count = count;
      return idx;
    }
    // the first best_split entries of the winning axis go left; the other two lists are partitioned stably
    for (int i = 0; i < count; ++i) goes_left[sorted[best_axis][first + i]] = i < best_split;
    for (int axis = 0; axis < 3; ++axis) {
      if (axis == best_axis) continue;
      int* list = sorted[axis].data() + first;
      int nl = 0, nr = 0;
      for (int i = 0; i < count; ++i) {
        if (goes_left[list[i]]) list[nl++] = list[i];
        else scratch[nr++] = list[i];
      }
      std::copy(scratch.begin(), scratch.begin() + nr, list + nl);
    }
    int l = build(first, best_split, depth + 1);
    int r = build(first + best_split, count - best_split, depth + 1);
    nodes[idx].left = l, nodes[idx].right = r;
    return idx;
  }
};

// f32 box, rounded outward and padded so f32 slab arithmetic stays conservative
void pad_box(const Box3& b, float lo[3], float hi[3]) {
  for (int i = 0; i < 3; ++i) {
    double m = std::max(1.0, std::max(std::fabs(b.lo[i]), std::fabs(b.hi[i])));
    double eps = 4e-6 * m + 1e-5 * (b.hi[i] - b.lo[i]);
    lo[i] = std::nextafterf((float)(b.lo[i] - eps), -INFINITY);
    hi[i] = std::nextafterf((float)(b.hi[i] + eps), INFINITY);
  }
}

}  // namespace

namespace {
// a four-wide node whose only child is one leaf (or nothing): scenes with zero or one surface primitive
DNode4 lone_leaf_node4(const float* lo, const float* hi, int leaf) {
  const float far_away = 1e30f;
  DNode4 dn{};
  for (int k = 0; k < 4; ++k) {
    dn.lox[k] = dn.hix[k] = dn.loy[k] = dn.hiy[k] = dn.loz[k] = dn.hiz[k] = far_away;
    dn.child[k] = ~0;
  }
  if (lo) {
    dn.lox[0] = lo[0], dn.hix[0] = hi[0], dn.loy[0] = lo[1], dn.hiy[0] = hi[1], dn.loz[0] = lo[2], dn.hiz[0] = hi[2];
    dn.child[0] = leaf;
  }
  return dn;
}
}  // namespace

int flatten_scene(const rtb_scene& s, int world, double time_lo, double time_hi, FlatScene* out, bool host_bvh) {
  if (world < 0 || (size_t)world >= s.obj.size()) return fail(RTB_ERR_INVALID, "flatten: bad world handle");
  *out = FlatScene();
  out->time_lo = time_lo, out->time_hi = time_hi;
  out->xforms.push_back(DXform{{1, 0, 0, 0, 1, 0, 0, 0, 1}, {0, 0, 0}});

  // tables that map one-to-one
  for (const RtbTexRec& t : s.tex) {
    DTexture d{};
    d.type = t.type, d.a = t.a, d.b = t.b;
    d.r = (float)t.v[0], d.g = (float)t.v[1], d.bl = (float)t.v[2];
    d.scale = (float)t.v[0];
    if (t.type == RTB_TEX_IMAGE) {
      d.off = (int)out->texels.size();
      const uint8_t* src = s.image_pool.data() + t.image_off;
      size_t n = (size_t)t.a * t.b;
      out->texels.resize((size_t)d.off + n);
      uint32_t* dst = out->texels.data() + d.off;
      for (size_t i = 0; i < n; ++i) dst[i] = (uint32_t)src[3 * i] | ((uint32_t)src[3 * i + 1] << 8) | ((uint32_t)src[3 * i + 2] << 16);
    }
    out->texs.push_back(d);
  }
  for (const RtbMatRec& m : s.mat) {
    DMaterial d{};
    d.type = m.type, d.tex = m.tex;
    d.param = (float)(m.type == RTB_MAT_METAL ? (m.param < 1.0 ? m.param : 1.0) : m.param);  // src/material.rs:70
    d.needs_uv = m.tex >= 0 ? tex_needs_uv(s, m.tex) : 0;
    out->mats.push_back(d);
  }
  for (const RtbPerlinTable& p : s.perlin) {
    DPerlin d;
    for (int i = 0; i < 256; ++i) d.ranvec[i] = {(float)p.ranvec[i][0], (float)p.ranvec[i][1], (float)p.ranvec[i][2], 0.f};
    std::memcpy(d.perm_x, p.perm_x, 256);
    std::memcpy(d.perm_y, p.perm_y, 256);
    std::memcpy(d.perm_z, p.perm_z, 256);
    out->perlin.push_back(d);
  }

  const bool timing = std::getenv("RTB200_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!timing) return;
    auto t = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[rtb200 timing] flatten: %s %.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count());
    t_prev = t;
  };
  lap("texture / material / perlin tables");
  Walker w{s, *out, time_lo, time_hi, {}, {}};
  if (!w.walk(world, Xf(), 0, false) || !w.err.empty())
    return fail(RTB_ERR_INVALID, "flatten: " + (w.err.empty() ? std::string("walk failed") : w.err));

  int n = (int)out->prims.size();
  {  // FlatScene::enclosed: count the faces of the world box that a thin primitive (a wall) covers to >= 80 %
    Box3 world_box;
    for (const Box3& b : w.prim_boxes) world_box.grow(b);
    int covered = 0;
    for (int a = 0; a < 3 && n > 0; ++a) {
      const int u = (a + 1) % 3, v = (a + 2) % 3;
      const double ext = world_box.hi[a] - world_box.lo[a];
      const double face = (world_box.hi[u] - world_box.lo[u]) * (world_box.hi[v] - world_box.lo[v]);
      if (!(ext > 0.0) || !(face > 0.0)) continue;
      for (int side = 0; side < 2; ++side) {
        const double plane = side ? world_box.hi[a] : world_box.lo[a];
        for (const Box3& b : w.prim_boxes) {
          const bool thin = b.hi[a] - b.lo[a] <= 1e-3 * ext;
          const bool there = std::fabs((side ? b.hi[a] : b.lo[a]) - plane) <= 1e-3 * ext;
          if (thin && there && (b.hi[u] - b.lo[u]) * (b.hi[v] - b.lo[v]) >= 0.8 * face) {
            ++covered;
            break;
          }
        }
      }
    }
    out->enclosed = covered >= 4;
  }
  if (n == 0) {  // only media, or an empty list: nothing to traverse
    DNode dn{};  // two empty leaves
    dn.n0 = {0, 0, 0, 0}, dn.n1 = dn.n0, dn.n2 = dn.n0;
    dn.n3 = {~0, ~0, 0, 0};
    out->nodes.push_back(dn);
    out->root = 0;
    out->nodes4.push_back(lone_leaf_node4(nullptr, nullptr, ~0));
    return RTB_OK;
  }
  if (!host_bvh && n >= 2) {
    // the BVH is built on the device (bvh_build.cu): hand over the primitives in insertion order and
    // their conservatively padded f32 boxes (lo.xyz, hi.xyz)
    out->prim_boxes.resize((size_t)n * 6);
    for (int i = 0; i < n; ++i) pad_box(w.prim_boxes[i], &out->prim_boxes[(size_t)i * 6], &out->prim_boxes[(size_t)i * 6 + 3]);
    out->root = -1;
    return RTB_OK;
  }
  Builder b{w.prim_boxes, {}, {}, 1, 1.0, 0, {}, {}, {}, {}};
  if (const char* e = std::getenv("RTB200_MAX_LEAF")) b.max_leaf = std::max(1, std::min(7, std::atoi(e)));
  if (const char* e = std::getenv("RTB200_CTRAV")) b.c_trav = std::atof(e);
  b.prepare();
  lap("graph walk");
  int root = b.build(0, n, 0);
  b.order = b.sorted[0];
  lap("SAH build");
  out->bvh_depth = b.depth_max;
  if (b.depth_max > 60) return fail(RTB_ERR_INVALID, "flatten: BVH too deep for the traversal stack");

  std::vector<DPrim> ordered(n);
  for (int i = 0; i < n; ++i) ordered[i] = out->prims[b.order[i]];
  out->prims.swap(ordered);

  // emit 2-child nodes in DFS order; leaves are encoded in the parent's child slot
  auto leaf_code = [](const BuildNode& bn) { return ~((bn.first << 3) | bn.count); };
  if (b.nodes[root].left < 0) {
    // single leaf: wrap it in one node whose second child is an empty leaf (count 0)
    DNode dn{};
    float lo[3], hi[3];
    pad_box(b.nodes[root].box, lo, hi);
    dn.n0 = {lo[0], hi[0], lo[1], hi[1]};
    dn.n1 = dn.n0;
    dn.n2 = {lo[2], hi[2], lo[2], hi[2]};
    dn.n3 = {leaf_code(b.nodes[root]), ~0, 0, 0};
    out->nodes.push_back(dn);
    out->root = 0;
    out->nodes4.push_back(lone_leaf_node4(lo, hi, leaf_code(b.nodes[root])));
    return RTB_OK;
  }
  std::vector<int> dev_index(b.nodes.size(), -1);
  int next = 0;
  {  // device indices in DFS pre-order from the root (the creation order, unless the tree was re-linked afterwards)
    std::vector<int> st{root};
    while (!st.empty()) {
      const int i = st.back();
      st.pop_back();
      if (b.nodes[i].left < 0) continue;
      dev_index[i] = next++;
      st.push_back(b.nodes[i].right);
      st.push_back(b.nodes[i].left);
    }
  }
  out->nodes.resize(next);
  for (size_t i = 0; i < b.nodes.size(); ++i) {
    const BuildNode& bn = b.nodes[i];
    if (bn.left < 0 || dev_index[i] < 0) continue;
    const BuildNode& L = b.nodes[bn.left];
    const BuildNode& R = b.nodes[bn.right];
    float l0[3], h0[3], l1[3], h1[3];
    pad_box(L.box, l0, h0);
    pad_box(R.box, l1, h1);
    DNode dn{};
    dn.n0 = {l0[0], h0[0], l0[1], h0[1]};
    dn.n1 = {l1[0], h1[0], l1[1], h1[1]};
    dn.n2 = {l0[2], h0[2], l1[2], h1[2]};
    dn.n3 = {L.left >= 0 ? dev_index[bn.left] : leaf_code(L), R.left >= 0 ? dev_index[bn.right] : leaf_code(R), 0, 0};
    out->nodes[dev_index[i]] = dn;
  }
  out->root = dev_index[root];

  // ---- the same tree, four-wide: a node adopts its grandchildren, largest surface area first, until it has four children
  // (or only leaves).  Children keep the two-child tree's boxes (already padded the same way), leaves keep their codes.
  {
    std::vector<DNode4>& n4 = out->nodes4;
    n4.clear();
    const float far_away = 1e30f;  // unused slots: a point no ray meets (and an empty leaf behind it if one ever does)
    struct Todo { int bn, slot, depth; };
    std::vector<Todo> todo;  // (BuildNode index, DNode4 index, depth of that node)
    n4.push_back(DNode4{});
    todo.push_back({root, 0, 1});
    int depth4 = 0;
    while (!todo.empty()) {
      const int bn_i = todo.back().bn, slot_i = todo.back().slot, depth_i = todo.back().depth;
      todo.pop_back();
      depth4 = std::max(depth4, depth_i);
      const BuildNode& bn = b.nodes[bn_i];
      std::vector<int> kids = {bn.left, bn.right};
      while (kids.size() < 4) {
        int best = -1;
        double best_area = -1.0;
        for (size_t k = 0; k < kids.size(); ++k)
          if (b.nodes[kids[k]].left >= 0 && b.nodes[kids[k]].box.area() > best_area) best_area = b.nodes[kids[k]].box.area(), best = (int)k;
        if (best < 0) break;
        const BuildNode& g = b.nodes[kids[best]];
        kids[best] = g.left;
        kids.push_back(g.right);
      }
      DNode4 dn{};
      for (int k = 0; k < 4; ++k) {
        if (k < (int)kids.size()) {
          const BuildNode& c = b.nodes[kids[k]];
          float lo[3], hi[3];
          pad_box(c.box, lo, hi);
          dn.lox[k] = lo[0], dn.hix[k] = hi[0], dn.loy[k] = lo[1], dn.hiy[k] = hi[1], dn.loz[k] = lo[2], dn.hiz[k] = hi[2];
          if (c.left >= 0) {
            dn.child[k] = (int)n4.size();
            n4.push_back(DNode4{});
            todo.push_back({kids[k], dn.child[k], depth_i + 1});
          } else {
            dn.child[k] = leaf_code(c);
          }
        } else {
          dn.lox[k] = dn.hix[k] = dn.loy[k] = dn.hiy[k] = dn.loz[k] = dn.hiz[k] = far_away;
          dn.child[k] = ~0;
        }
      }
      n4[slot_i] = dn;
    }
    out->root4 = 0;
    // a step pushes up to three children: the traversal stack (RTB_STACK_DEPTH = 64 entries) must hold 3 per level
    if (3 * depth4 + 1 > 60) n4.clear();
  }
  return RTB_OK;
}

}  // namespace rtb
