// wide_exp.cpp — DEVELOPER TOOL (part of libemul, see emul.cpp): BVH layout experiments on the host.  Collapses the
// two-child SAH tree of flatten.cpp into W-wide nodes and counts traversal steps / box tests / stack pushes for the ray
// set of a real render, so a layout can be judged before any kernel is written for it.
#define RTB_HOST_EMUL 1
#include <vector>

#include "flatten.hpp"
#include "rtb200.h"
#include "trace.cuh"

using namespace rtb;

struct EmuSceneView {  // mirrors the head of EmuScene in emul.cpp
  rtb_scene* s;
  FlatScene flat;
  SceneView view;
};

namespace {
struct WChild {
  float lo[3], hi[3];
  int ref;  // >= 0: wide node index; < 0: leaf code of the two-child tree
};
struct WNode {
  WChild c[8];
  int n;
};
struct Wide {
  std::vector<WNode> nodes;
  int width;
};

float area(const WChild& b) {
  float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  return 2.f * (dx * dy + dy * dz + dz * dx);
}
void children_of(const DNode& n, WChild out[2]) {
  out[0] = WChild{{n.n0.x, n.n0.z, n.n2.x}, {n.n0.y, n.n0.w, n.n2.y}, n.n3.x};
  out[1] = WChild{{n.n1.x, n.n1.z, n.n2.z}, {n.n1.y, n.n1.w, n.n2.w}, n.n3.y};
}
int collapse(const std::vector<DNode>& bvh2, int node2, Wide& w) {
  const int idx = (int)w.nodes.size();
  w.nodes.push_back(WNode{});
  std::vector<WChild> kids(2);
  children_of(bvh2[node2], kids.data());
  while ((int)kids.size() < w.width) {
    int best = -1;
    float best_a = -1.f;
    for (size_t i = 0; i < kids.size(); ++i)
      if (kids[i].ref >= 0 && area(kids[i]) > best_a) best_a = area(kids[i]), best = (int)i;
    if (best < 0) break;
    WChild two[2];
    children_of(bvh2[kids[best].ref], two);
    kids[best] = two[0];
    kids.push_back(two[1]);
  }
  WNode wn{};
  wn.n = (int)kids.size();
  for (int i = 0; i < wn.n; ++i) {
    wn.c[i] = kids[i];
    if (kids[i].ref >= 0) wn.c[i].ref = collapse(bvh2, kids[i].ref, w);
  }
  w.nodes[idx] = wn;
  return idx;
}
}  // namespace

static uint16_t* g_per_ray_steps = nullptr;  // optional: node steps of each ray (emu_wide_steps)

extern "C" {

// rays: n x {o, d, time, self_prim, self_face} (9 floats; the last two are ints stored as floats).
// mode 0: hit children fully sorted by entry distance; 1: nearest first, the rest pushed in slot order.
// out: [0] node steps, [1] box tests, [2] pushes, [3] leaf visits (= primitive tests), [4] hits, [5] max stack depth,
//      [6 + k] steps whose hit count was k (k = 0..8)
void emu_wide_stats(void* scene, const float* rays, size_t n, int width, int mode, uint64_t* out) {
  EmuSceneView* e = (EmuSceneView*)scene;
  const SceneView& sv = e->view;
  Wide w;
  w.width = width;
  collapse(e->flat.nodes, sv.root, w);
  for (int k = 0; k < 16; ++k) out[k] = 0;
  out[15] = w.nodes.size();
  Counters cnt = {0, 0, 0};
  for (size_t i = 0; i < n; ++i) {
    const float* r = rays + 9 * i;
    RayF ray;
    ray.o = mk(r[0], r[1], r[2]), ray.d = mk(r[3], r[4], r[5]), ray.time = r[6];
    const int self_prim = (int)r[7], self_face = (int)r[8];
    const float t_min = 1e-3f;
    const float inv_a = rcp_fast(dot(ray.d, ray.d));
    const RayInv ri = make_ray_inv(ray);
    Hit best;
    best.t = CUDART_INF_F, best.prim = -1, best.face = 0, best.id = -1;
    int stack[256];
    int sp = 0;
    int cur = 0;
    const uint64_t steps_before = out[0];
    for (;;) {
      while (cur >= 0) {
        out[0]++;
        const WNode& wn = w.nodes[cur];
        float tn[8];
        int ref[8], nh = 0;
        for (int k = 0; k < wn.n; ++k) {
          out[1]++;
          const WChild& c = wn.c[k];
          float lx = fmaf(c.lo[0], ri.idx, -ri.oodx), hx = fmaf(c.hi[0], ri.idx, -ri.oodx);
          float ly = fmaf(c.lo[1], ri.idy, -ri.oody), hy = fmaf(c.hi[1], ri.idy, -ri.oody);
          float lz = fmaf(c.lo[2], ri.idz, -ri.oodz), hz = fmaf(c.hi[2], ri.idz, -ri.oodz);
          float a = fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fmaxf(fminf(lz, hz), t_min));
          float b = fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fminf(fmaxf(lz, hz), best.t));
          if (a <= b) tn[nh] = a, ref[nh] = c.ref, nh++;
        }
        out[6 + nh]++;
        if (nh == 0) {
          if (sp == 0) goto done;
          cur = stack[--sp];
          continue;
        }
        // nearest to the front
        int m = 0;
        for (int k = 1; k < nh; ++k)
          if (tn[k] < tn[m]) m = k;
        std::swap(tn[0], tn[m]), std::swap(ref[0], ref[m]);
        if (mode == 0)  // the rest farthest-first onto the stack
          for (int a = 1; a < nh; ++a)
            for (int b = a + 1; b < nh; ++b)
              if (tn[b] > tn[a]) std::swap(tn[a], tn[b]), std::swap(ref[a], ref[b]);
        for (int k = 1; k < nh; ++k) stack[sp++] = ref[k], out[2]++;
        if ((uint64_t)sp > out[5]) out[5] = sp;
        cur = ref[0];
      }
      {
        const unsigned code = ~(unsigned)cur;
        const int first = (int)(code >> 3), count = (int)(code & 7u);
        for (int q = 0; q < count; ++q) {
          const int pi = first + q;
          float t;
          int face, id;
          out[3]++;
          if (intersect_prim<false>(sv.prims, pi, sv.xforms, ray, inv_a, t_min, best.t, pi == self_prim, self_face, &t, &face, &id, &cnt) &&
              (t < best.t || id > best.id))
            best.t = t, best.prim = pi, best.face = face, best.id = id;
        }
      }
      if (sp == 0) break;
      cur = stack[--sp];
    }
  done:
    if (best.prim >= 0) out[4]++;
    if (g_per_ray_steps) g_per_ray_steps[i] = (uint16_t)std::min<uint64_t>(out[0] - steps_before, 65535);
  }
}

// same, and also the node steps each ray took (steps[n])
void emu_wide_steps(void* scene, const float* rays, size_t n, int width, int mode, uint64_t* out, uint16_t* steps) {
  g_per_ray_steps = steps;
  emu_wide_stats(scene, rays, n, width, mode, out);
  g_per_ray_steps = nullptr;
}

}  // extern "C"
