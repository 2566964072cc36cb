// emul.cpp — DEVELOPER TOOL, not product code and not the oracle.  Compiles the kernels' own device functions
// (csrc/trace.cuh, f32) for the host and drives them with plain loops, so that in a container without a GPU one can
//   * check kernel LOGIC (traversal, self-hit exclusion, hit records, scatter, media) against the f64 oracle,
//   * count traversal steps / box tests of a BVH layout before spending GPU time on it.
// SFU approximations are libm calls here, so numbers differ from the device in the last bits.  Nothing in
// raytracing_rust_b200/ links, imports or executes this; the product fails loudly without a CUDA device.
#define RTB_HOST_EMUL 1
#include <atomic>
#include <cstdlib>
#include <thread>
#include <vector>

#include "flatten.hpp"
#include "rtb200.h"
#include "trace.cuh"

namespace rtb {
void destroy_device_scene(DeviceScene*) {}  // scene.cpp's handle destructor; no device here
}
using namespace rtb;

struct EmuScene {
  rtb_scene* s = nullptr;
  FlatScene flat;
  SceneView view{};
  std::vector<int> id_table;
};

static void make_view(EmuScene* e) {
  const FlatScene& f = e->flat;
  SceneView& v = e->view;
  v.nodes = f.nodes.data(), v.prims = f.prims.data(), v.bprims = f.bprims.data(), v.media = f.media.data();
  v.xforms = f.xforms.data(), v.mats = f.mats.data(), v.texs = f.texs.data(), v.perlin = f.perlin.data();
  v.texels = f.texels.data();
  v.n_nodes = (int)f.nodes.size(), v.n_prims = (int)f.prims.size(), v.n_media = (int)f.media.size(), v.root = f.root;
  v.nodes4 = f.nodes4.empty() ? nullptr : f.nodes4.data(), v.root4 = f.root4;
  e->id_table.assign(std::max(1, f.n_prim_ids), 0);
  for (int p = 0; p < v.n_prims; ++p) {
    const int n = RTB_PRIM_TYPE(f.prims[p].c.x) == DP_BOX ? 6 : 1;
    for (int k = 0; k < n; ++k) e->id_table[f.prims[p].c.z + k] = p;
  }
}

extern "C" {

void* emu_scene_load(const void* bytes, size_t len, int world) {
  rtb_scene* s = rtb_scene_deserialize(bytes, len);
  if (!s) return nullptr;
  EmuScene* e = new EmuScene();
  e->s = s;
  if (flatten_scene(*s, world, 0.0, 1.0, &e->flat, true) != RTB_OK) {
    rtb_scene_destroy(s);
    delete e;
    return nullptr;
  }
  make_view(e);
  return e;
}
// width 2: the two-child tree; 4: the four-wide tree (default when flatten built it)
void emu_set_bvh_width(void* p, int width) {
  EmuScene* e = (EmuScene*)p;
  e->view.nodes4 = (width == 4 && !e->flat.nodes4.empty()) ? e->flat.nodes4.data() : nullptr;
}
// what the flattener found out about the scene: [0] primitives, [1] FlatScene::enclosed
void emu_scene_facts(void* p, int* out) {
  EmuScene* e = (EmuScene*)p;
  out[0] = (int)e->flat.prims.size(), out[1] = e->flat.enclosed ? 1 : 0;
}
void emu_scene_free(void* p) {
  EmuScene* e = (EmuScene*)p;
  if (!e) return;
  rtb_scene_destroy(e->s);
  delete e;
}

// same contract as rtb_closest_hit_ex; counters (optional): [0] node steps, [1] primitive tests
void emu_closest_hit_ex(void* p, const double* rays, size_t n, double t_min, const int32_t* self_id, int32_t* out_prim,
                        float* out_rec, uint64_t* counters) {
  EmuScene* e = (EmuScene*)p;
  const SceneView& sv = e->view;
  Counters cnt = {0, 0, 0};
  for (size_t i = 0; i < n; ++i) {
    const double* r = rays + 7 * i;
    RayF ray;
    ray.o = mk((float)r[0], (float)r[1], (float)r[2]), ray.d = mk((float)r[3], (float)r[4], (float)r[5]), ray.time = (float)r[6];
    int self_prim = -1, self_face = 0;
    if (self_id && self_id[i] >= 0) {
      self_prim = e->id_table[self_id[i]];
      self_face = self_id[i] - sv.prims[self_prim].c.z;
    }
    const Hit h = trace_closest<true>(sv, ray, (float)t_min, self_prim, self_face, &cnt);
    int id = -1;
    float rec[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (h.prim >= 0) {
      const int4 C = ldg4i(&sv.prims[h.prim].c);
      id = C.z + (RTB_PRIM_TYPE(C.x) == DP_BOX ? h.face : 0);
      const Surface s = finalize_hit(sv, sv.prims, ray, h, nullptr, true);
      rec[0] = h.t, rec[1] = s.u, rec[2] = s.v, rec[3] = s.p.x, rec[4] = s.p.y, rec[5] = s.p.z;
      rec[6] = s.n.x, rec[7] = s.n.y, rec[8] = s.n.z;
    }
    out_prim[i] = id;
    if (out_rec) std::memcpy(out_rec + 9 * i, rec, sizeof rec);
  }
  if (counters) counters[0] = cnt.node_tests, counters[1] = cnt.prim_tests;
}

void emu_camera_rays(const RtbCamera* c, int nx, int ny, int sidx, uint64_t seed, float* out);
void emu_scatter(void* p, int mat, const double* in, size_t n, uint64_t seed, float* out) {
  EmuScene* e = (EmuScene*)p;
  const SceneView& sv = e->view;
  for (size_t idx = 0; idx < n; ++idx) {
    const double* q = in + 16 * idx;
    RayF ray;
    ray.o = mk((float)q[0], (float)q[1], (float)q[2]), ray.d = mk((float)q[3], (float)q[4], (float)q[5]), ray.time = (float)q[6];
    Surface s;
    s.u = (float)q[8], s.v = (float)q[9], s.p = mk((float)q[10], (float)q[11], (float)q[12]);
    s.n = mk((float)q[13], (float)q[14], (float)q[15]), s.mat = mat;
    const int4 m = __ldg(reinterpret_cast<const int4*>(sv.mats + mat));
    const Rand5 rnd = rand5(seed, 0u, (uint32_t)idx, 0u, RTB_RNG_BOUNCE);
    TexCounters tc;
    V3f tex_rgb = mk(1.f, 1.f, 1.f);
    if (m.x != RTB_MAT_DIELECTRIC) tex_rgb = texture_value<false>(sv, m.y, s.u, s.v, s.p, &tc);
    const bool ok = scatter_dir(m.x, __int_as_float(m.z), s, ray, rnd);
    float* o = out + 10 * idx;
    o[0] = ok ? 1.f : 0.f;
    o[1] = ok ? ray.o.x : 0.f, o[2] = ok ? ray.o.y : 0.f, o[3] = ok ? ray.o.z : 0.f;
    o[4] = ok ? ray.d.x : 0.f, o[5] = ok ? ray.d.y : 0.f, o[6] = ok ? ray.d.z : 0.f;
    o[7] = ok ? tex_rgb.x : 0.f, o[8] = ok ? tex_rgb.y : 0.f, o[9] = ok ? tex_rgb.z : 0.f;
  }
}

void emu_sample(int kind, uint64_t seed, size_t n, float* out) {
  for (size_t idx = 0; idx < n; ++idx) {
    V3f p;
    if (kind == 0) {
      const Rand5 r = rand5(seed, 0u, (uint32_t)idx, 0u, RTB_RNG_BOUNCE);
      p = sample_unit_ball(r.a, r.b, r.c);
    } else {
      const Rand5 r = rand5(seed, 0u, (uint32_t)idx, 0u, RTB_RNG_CAMERA);
      const float rr = sqrt_fast(r.c);
      float sn, cs;
      __sincosf((2.0f * r.d - 1.0f) * CUDART_PI_F, &sn, &cs);
      p = mk(rr * cs, rr * sn, 0.f);
    }
    out[3 * idx] = p.x, out[3 * idx + 1] = p.y, out[3 * idx + 2] = p.z;
  }
}

static void host_camera(const RtbCamera& c, DCamera* out) {  // render.cu make_device_camera (src/camera.rs:21-51)
  auto sub = [](const double* a, const double* b, double* r) { for (int i = 0; i < 3; ++i) r[i] = a[i] - b[i]; };
  auto nrm = [](double* a) { double n = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); for (int i = 0; i < 3; ++i) a[i] /= n; };
  auto crs = [](const double* a, const double* b, double* r) {
    r[0] = a[1] * b[2] - a[2] * b[1], r[1] = a[2] * b[0] - a[0] * b[2], r[2] = a[0] * b[1] - a[1] * b[0];
  };
  double half_height = c.focus_dist * std::tan(c.vertical_fov * M_PI / 180.0 / 2.0), half_width = c.aspect * half_height;
  double w[3], u[3], v[3];
  sub(c.look_from, c.look_at, w), nrm(w), crs(c.view_up, w, u), nrm(u), crs(w, u, v);
  for (int i = 0; i < 3; ++i) {
    out->origin[i] = (float)c.look_from[i];
    out->llc[i] = (float)(c.look_from[i] - half_width * u[i] - half_height * v[i] - c.focus_dist * w[i]);
    out->horizontal[i] = (float)(2.0 * half_width * u[i]), out->vertical[i] = (float)(2.0 * half_height * v[i]);
    out->u[i] = (float)u[i], out->v[i] = (float)v[i];
  }
  out->time0 = (float)c.time0, out->time1 = (float)c.time1, out->lens_radius = (float)(c.aperture / 2.0);
}

void emu_camera_rays(const RtbCamera* c, int nx, int ny, int sidx, uint64_t seed, float* out) {
  DCamera cam;
  host_camera(*c, &cam);
  const PhiloxKeys keys = philox_keys(seed);
  for (int pixel = 0; pixel < nx * ny; ++pixel) {
    const int row = pixel / nx, i = pixel - row * nx;
    const RayF ray = camera_ray(cam, i, ny - 1 - row, nx, ny, keys, (uint32_t)pixel, (uint32_t)sidx);
    float* o = out + 7 * (size_t)pixel;
    o[0] = ray.o.x, o[1] = ray.o.y, o[2] = ray.o.z, o[3] = ray.d.x, o[4] = ray.d.y, o[5] = ray.d.z, o[6] = ray.time;
  }
}

// The SHADE logic of render.cu's k_render_mega as a plain per-path loop over the same device functions (for radiance
// checks of the f32 path without a GPU).  out_sum: 3*nx*ny floats (SUM of samples); stats: [0] segments, [1] node steps,
// [2] prim tests.
void emu_render(void* p, const RtbCamera* c, int nx, int ny, int ns, int max_depth, double t_min_d, int background, uint64_t seed,
                int nthreads, float* out_sum, uint64_t* stats, float* seg_dump, size_t seg_cap) {
  EmuScene* e = (EmuScene*)p;
  const SceneView& sv = e->view;
  DCamera cam;
  host_camera(*c, &cam);
  const PhiloxKeys keys = philox_keys(seed);
  const float t_min = (float)t_min_d;
  const bool media_first = std::getenv("EMU_MEDIA_FIRST") != nullptr;
  std::atomic<int> next_row{0};
  std::atomic<uint64_t> a_seg{0}, a_node{0}, a_prim{0};
  std::atomic<size_t> dump_n{0};  // seg_dump (optional): every segment's {o, d, time, self_prim, self_face}, 9 floats
  auto worker = [&]() {
    Counters cnt = {0, 0, 0};
    TexCounters tc = {{0, 0, 0, 0}};
    uint64_t segs = 0;
    for (;;) {
      const int row = next_row.fetch_add(1);
      if (row >= ny) break;
      for (int i = 0; i < nx; ++i) {
        const uint32_t pixel = (uint32_t)(row * nx + i);
        double acc[3] = {0, 0, 0};
        for (int smp = 0; smp < ns; ++smp) {
          RayF ray = camera_ray(cam, i, ny - 1 - row, nx, ny, keys, pixel, (uint32_t)smp);
          V3f thr = mk(1.f, 1.f, 1.f);
          int self_prim = -1, self_face = 0;
          for (int depth = 0;; ) {
            if (seg_dump) {
              const size_t k = dump_n.fetch_add(1);
              if (k < seg_cap) {
                float* q = seg_dump + 9 * k;
                q[0] = ray.o.x, q[1] = ray.o.y, q[2] = ray.o.z, q[3] = ray.d.x, q[4] = ray.d.y, q[5] = ray.d.z, q[6] = ray.time;
                q[7] = (float)self_prim, q[8] = (float)self_face;
              }
            }
            const Rand5 rnd = rand5(keys, (uint32_t)depth, pixel, (uint32_t)smp, RTB_RNG_BOUNCE);
            int medium = -1;
            Hit best;
            if (media_first) {
              // the media's free-flight distances first (against the unclipped boundary intervals); the surface query then only
              // looks nearer than the scatter point: same draws, same outcome, fewer node tests
              float t_med = CUDART_INF_F;
              if (sv.n_media > 0)
                medium = sample_media<true>(sv, ray, t_min, &t_med, rnd.d, rnd.e, seed, pixel, (uint32_t)smp, (uint32_t)depth, &cnt);
              best = trace_closest<true>(sv, ray, t_min, self_prim, self_face, &cnt, t_med);
              if (best.prim >= 0) medium = -1;
            } else {
              best = trace_closest<true>(sv, ray, t_min, self_prim, self_face, &cnt);
              if (sv.n_media > 0)
                medium = sample_media<true>(sv, ray, t_min, &best.t, rnd.d, rnd.e, seed, pixel, (uint32_t)smp, (uint32_t)depth, &cnt);
            }
            ++segs;
            bool ended = true;
            V3f rad = mk(0.f, 0.f, 0.f);
            if (medium >= 0 || best.prim >= 0) {
              Surface s;
              int4 m;
              if (medium >= 0) {
                s.p = madd(best.t, ray.d, ray.o), s.n = mk(1.f, 0.f, 0.f), s.u = 0.f, s.v = 0.f;
                s.mat = sv.media[medium].mat;
                m = __ldg(reinterpret_cast<const int4*>(sv.mats + s.mat));
                self_prim = -1;
              } else {
                s = finalize_hit(sv, sv.prims, ray, best, &m);
                self_prim = best.prim, self_face = best.face;
              }
              V3f tex_rgb = mk(1.f, 1.f, 1.f);
              if (m.x != RTB_MAT_DIELECTRIC && (m.x == RTB_MAT_DIFFUSE_LIGHT || depth < max_depth))
                tex_rgb = texture_value<true>(sv, m.y, s.u, s.v, s.p, &tc);
              if (m.x == RTB_MAT_DIFFUSE_LIGHT) rad = thr * tex_rgb;
              if (depth < max_depth && scatter_dir(m.x, __int_as_float(m.z), s, ray, rnd)) {
                thr = thr * tex_rgb;
                ++depth;
                ended = false;
              }
            } else {
              rad = thr * background_color(background, ray.d);
            }
            acc[0] += rad.x, acc[1] += rad.y, acc[2] += rad.z;
            if (ended) break;
          }
        }
        float* o = out_sum + 3 * (size_t)pixel;
        o[0] = (float)acc[0], o[1] = (float)acc[1], o[2] = (float)acc[2];
      }
    }
    a_seg += segs, a_node += cnt.node_tests, a_prim += cnt.prim_tests;
  };
  std::vector<std::thread> th;
  for (int t = 1; t < std::max(1, nthreads); ++t) th.emplace_back(worker);
  worker();
  for (auto& t : th) t.join();
  if (stats) stats[0] = a_seg, stats[1] = a_node, stats[2] = a_prim;
}

}  // extern "C"
