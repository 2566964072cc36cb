// rtb200.hpp — C++17 host API that keeps the shape of the raytracing_rust crate
// (module-by-module: sphere::Sphere, rect::Rect, material::Lambertian, ... and the
// create_image driver of tests/test.rs:55) on top of the C ABI in rtb200.h.
//
// The classes are pure descriptions, like the Rust value types: objects own their materials,
// materials own their textures, copying is `.clone()`.  Nothing is intersected on the host:
// create_image() lowers the object graph into an rtb_scene (one C-ABI builder call per
// reference constructor call) and renders it on the GPU.  Header-only; link librtb200.so.
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "rtb200.h"

namespace rtb200 {

struct Vector3 {  // nalgebra::Vector3<f64> as far as scene construction needs it
  double x = 0, y = 0, z = 0;
  Vector3() = default;
  Vector3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
  Vector3 operator+(const Vector3& o) const { return {x + o.x, y + o.y, z + o.z}; }
  Vector3 operator-(const Vector3& o) const { return {x - o.x, y - o.y, z - o.z}; }
  double magnitude() const;
  const double* data() const { return &x; }
};
inline double Vector3::magnitude() const { return __builtin_sqrt(x * x + y * y + z * z); }

// rand::thread_rng() stand-in: a per-thread seeded stream (the reference is unseeded, F5).
class ThreadRng {
 public:
  void seed(uint64_t s) { s_ = s; }
  uint64_t next_u64() {
    uint64_t z = (s_ += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double gen_f64() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }

 private:
  uint64_t s_ = 0x5EED;
};
inline ThreadRng& thread_rng() {
  static thread_local ThreadRng r;
  return r;
}

// Lowering context: memoises shared Perlin tables so clones of one NoiseTexture map to one handle.
struct Lowering {
  rtb_scene* scene;
  std::map<const void*, int> memo;
};
inline int check(int h, const char* what) {
  if (h < 0) throw std::runtime_error(std::string(what) + ": " + rtb_last_error());
  return h;
}

// ---- texture.rs -------------------------------------------------------------------------
struct Texture {
  virtual ~Texture() = default;
  virtual int lower(Lowering&) const = 0;
};
using TexturePtr = std::shared_ptr<const Texture>;

struct SolidTexture : Texture {  // src/texture.rs:9-26
  double r, g, b;
  SolidTexture(double r_, double g_, double b_) : r(r_), g(g_), b(b_) {}
  int lower(Lowering& L) const override { return check(rtb_tex_solid(L.scene, r, g, b), "SolidTexture"); }
};

struct CheckerTexture : Texture {  // src/texture.rs:28-49 — (odd, even)
  TexturePtr odd, even;
  template <class T, class U>
  CheckerTexture(T odd_, U even_)
      : odd(std::make_shared<T>(std::move(odd_))), even(std::make_shared<U>(std::move(even_))) {}
  int lower(Lowering& L) const override {
    int a = odd->lower(L), b = even->lower(L);
    return check(rtb_tex_checker(L.scene, a, b), "CheckerTexture");
  }
};

struct NoiseTexture : Texture {  // src/texture.rs:51-72; tables drawn at construction (Perlin::new)
  double scale;
  std::shared_ptr<const uint64_t> table_seed;  // shared by clones, like the cloned Vec tables
  explicit NoiseTexture(double scale_)
      : scale(scale_), table_seed(std::make_shared<uint64_t>(thread_rng().next_u64())) {}
  int lower(Lowering& L) const override {
    auto it = L.memo.find(table_seed.get());
    if (it != L.memo.end()) return it->second;  // a clone: same tables, same scale
    int h = check(rtb_tex_noise(L.scene, scale, *table_seed), "NoiseTexture");
    L.memo[table_seed.get()] = h;
    return h;
  }
};

// image::open(path).expect(..).to_rgb8() of tests/test.rs:201-205: the decoded pixels of a JPEG file
struct RgbImage {
  std::vector<uint8_t> data;
  uint32_t nx = 0, ny = 0;
  std::vector<uint8_t> into_raw() && { return std::move(data); }
};
inline RgbImage open_image(const std::string& path) {
  uint8_t* p = nullptr;
  RgbImage im;
  check(rtb_jpeg_decode_file(path.c_str(), &p, &im.nx, &im.ny), "image::open");
  im.data.assign(p, p + (size_t)3 * im.nx * im.ny);
  rtb_image_free(p);
  return im;
}

struct ImageTexture : Texture {  // src/texture.rs:74-108
  std::shared_ptr<const std::vector<uint8_t>> data;
  uint32_t nx, ny;
  ImageTexture(std::vector<uint8_t> d, uint32_t nx_, uint32_t ny_)
      : data(std::make_shared<std::vector<uint8_t>>(std::move(d))), nx(nx_), ny(ny_) {}
  int lower(Lowering& L) const override {
    auto it = L.memo.find(data.get());
    if (it != L.memo.end()) return it->second;
    int h = check(rtb_tex_image(L.scene, data->data(), nx, ny), "ImageTexture");
    L.memo[data.get()] = h;
    return h;
  }
};

// ---- material.rs ------------------------------------------------------------------------
struct Material {
  virtual ~Material() = default;
  virtual int lower(Lowering&) const = 0;
};
using MaterialPtr = std::shared_ptr<const Material>;

struct Lambertian : Material {  // src/material.rs:36-58
  TexturePtr albedo;
  template <class T>
  explicit Lambertian(T t) : albedo(std::make_shared<T>(std::move(t))) {}
  int lower(Lowering& L) const override {
    return check(rtb_mat_lambertian(L.scene, albedo->lower(L)), "Lambertian");
  }
};
struct Metal : Material {  // src/material.rs:61-92 (fuzz clamp :70 happens where it is used)
  TexturePtr albedo;
  double fuzz;
  template <class T>
  Metal(T t, double fuzz_) : albedo(std::make_shared<T>(std::move(t))), fuzz(fuzz_) {}
  int lower(Lowering& L) const override {
    return check(rtb_mat_metal(L.scene, albedo->lower(L), fuzz), "Metal");
  }
};
struct Dielectric : Material {  // src/material.rs:95-131
  double ref_idx;
  explicit Dielectric(double r) : ref_idx(r) {}
  int lower(Lowering& L) const override { return check(rtb_mat_dielectric(L.scene, ref_idx), "Dielectric"); }
};
struct DiffuseLight : Material {  // src/material.rs:133-151
  TexturePtr emit;
  template <class T>
  explicit DiffuseLight(T t) : emit(std::make_shared<T>(std::move(t))) {}
  int lower(Lowering& L) const override {
    return check(rtb_mat_diffuse_light(L.scene, emit->lower(L)), "DiffuseLight");
  }
};
struct Isotropic : Material {  // src/material.rs:154-173
  TexturePtr albedo;
  template <class T>
  explicit Isotropic(T t) : albedo(std::make_shared<T>(std::move(t))) {}
  int lower(Lowering& L) const override {
    return check(rtb_mat_isotropic(L.scene, albedo->lower(L)), "Isotropic");
  }
};

// ---- hittable.rs and the primitives -----------------------------------------------------
struct Hittable {
  virtual ~Hittable() = default;
  virtual int lower(Lowering&) const = 0;
};
using HittablePtr = std::shared_ptr<const Hittable>;  // Box<dyn Hittable> / Rc<dyn Hittable>

template <class H>
HittablePtr boxed(H h) {
  return std::make_shared<H>(std::move(h));
}

struct Sphere : Hittable {  // src/sphere.rs:20-85
  Vector3 center;
  double radius;
  MaterialPtr material;
  template <class M>
  Sphere(Vector3 c, double r, M m) : center(c), radius(r), material(std::make_shared<M>(std::move(m))) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_sphere(L.scene, center.data(), radius, material->lower(L)), "Sphere");
  }
};

struct MovingSphere : Hittable {  // src/sphere.rs:87-175
  Vector3 center0, center1;
  double time0, time1, radius;
  MaterialPtr material;
  template <class M>
  MovingSphere(Vector3 c0, Vector3 c1, double t0, double t1, double r, M m)
      : center0(c0), center1(c1), time0(t0), time1(t1), radius(r),
        material(std::make_shared<M>(std::move(m))) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_moving_sphere(L.scene, center0.data(), center1.data(), time0, time1, radius,
                                       material->lower(L)),
                 "MovingSphere");
  }
};

enum class Plane { YZ = RTB_PLANE_YZ, ZX = RTB_PLANE_ZX, XY = RTB_PLANE_XY };  // src/rect.rs:8-12
struct Rect : Hittable {  // src/rect.rs:14-76
  Plane plane;
  double x0, y0, x1, y1, k;
  MaterialPtr material;
  template <class M>
  Rect(Plane p, double x0_, double y0_, double x1_, double y1_, double k_, M m)
      : plane(p), x0(x0_), y0(y0_), x1(x1_), y1(y1_), k(k_), material(std::make_shared<M>(std::move(m))) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_rect(L.scene, (int)plane, x0, y0, x1, y1, k, material->lower(L)), "Rect");
  }
};

struct Cube : Hittable {  // src/cube.rs:8-94
  Vector3 p_min, p_max;
  MaterialPtr material;
  template <class M>
  Cube(Vector3 a, Vector3 b, M m) : p_min(a), p_max(b), material(std::make_shared<M>(std::move(m))) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_cube(L.scene, p_min.data(), p_max.data(), material->lower(L)), "Cube");
  }
};

struct FlipNormals : Hittable {  // src/hittable.rs:67-88
  HittablePtr hittable;
  template <class H>
  explicit FlipNormals(H h) : hittable(boxed(std::move(h))) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_flip_normals(L.scene, hittable->lower(L)), "FlipNormals");
  }
};

enum class Axis { X = RTB_AXIS_X, Y = RTB_AXIS_Y, Z = RTB_AXIS_Z };  // src/rotate.rs:7-11
struct Rotate : Hittable {  // src/rotate.rs:21-118
  Axis axis;
  HittablePtr hittable;
  double angle;
  template <class H>
  Rotate(Axis a, H h, double angle_deg) : axis(a), hittable(boxed(std::move(h))), angle(angle_deg) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_rotate(L.scene, (int)axis, hittable->lower(L), angle), "Rotate");
  }
};

struct Traslate : Hittable {  // src/traslate.rs:6-33 (sic)
  HittablePtr hitable;
  Vector3 offset;
  template <class H>
  Traslate(H h, Vector3 off) : hitable(boxed(std::move(h))), offset(off) {}
  int lower(Lowering& L) const override {
    return check(rtb_obj_translate(L.scene, hitable->lower(L), offset.data()), "Traslate");
  }
};

struct ConstantMedium : Hittable {  // src/medium.rs:11-61
  HittablePtr boundary;
  double density;
  TexturePtr texture;
  template <class H, class T>
  ConstantMedium(H b, double d, T t)
      : boundary(boxed(std::move(b))), density(d), texture(std::make_shared<T>(std::move(t))) {}
  int lower(Lowering& L) const override {
    int b = boundary->lower(L);
    return check(rtb_obj_constant_medium(L.scene, b, density, texture->lower(L)), "ConstantMedium");
  }
};

struct HittableList : Hittable {  // src/hittable.rs:23-65
  std::vector<HittablePtr> list;
  template <class H>
  void push(H h) { list.push_back(boxed(std::move(h))); }
  int lower(Lowering& L) const override {
    std::vector<int> hs;
    for (auto& h : list) hs.push_back(h->lower(L));
    return check(rtb_obj_list(L.scene, hs.data(), (int)hs.size()), "HittableList");
  }
};

struct BVHNode : Hittable {  // src/bvh.rs:9-94 — BVHNode::new(&mut [Rc<dyn Hittable>], time0, time1)
  std::vector<HittablePtr> hittable;
  double time0, time1;
  uint64_t seed;  // split-axis stream of the oracle's reference-style build (src/bvh.rs:40)
  BVHNode(const std::vector<HittablePtr>& hs, double t0, double t1)
      : hittable(hs), time0(t0), time1(t1), seed(thread_rng().next_u64()) {}
  int lower(Lowering& L) const override {
    std::vector<int> hs;
    for (auto& h : hittable) hs.push_back(h->lower(L));
    return check(rtb_obj_bvh(L.scene, hs.data(), (int)hs.size(), time0, time1, seed), "BVHNode");
  }
};

// ---- camera.rs ----------------------------------------------------------------------------
struct Camera {  // Camera::new argument order, src/camera.rs:21-31
  RtbCamera c;
  Camera(Vector3 look_from, Vector3 look_at, Vector3 view_up, double vertical_fov, double aspect,
         double aperture, double focus_dist, double time0, double time1) {
    for (int i = 0; i < 3; ++i) {
      c.look_from[i] = look_from.data()[i];
      c.look_at[i] = look_at.data()[i];
      c.view_up[i] = view_up.data()[i];
    }
    c.vertical_fov = vertical_fov, c.aspect = aspect, c.aperture = aperture;
    c.focus_dist = focus_dist, c.time0 = time0, c.time1 = time1;
  }
};

// set_camera, tests/test.rs:30-53 (note focus_dist precedes aperture here, as upstream)
inline Camera set_camera(size_t nx, size_t ny, Vector3 look_from, Vector3 look_at, Vector3 view_up,
                         double vertical_fov, double focus_dist, double aperture, double time0,
                         double time1) {
  return Camera(look_from, look_at, view_up, vertical_fov, (double)nx / (double)ny, aperture, focus_dist,
                time0, time1);
}

// RAII scene lowered from a world
struct LoweredScene {
  rtb_scene* scene = nullptr;
  int world = -1;
  explicit LoweredScene(const Hittable& w) {
    scene = rtb_scene_create();
    Lowering L{scene, {}};
    try {
      world = w.lower(L);
    } catch (...) {
      rtb_scene_destroy(scene);
      throw;
    }
  }
  ~LoweredScene() { rtb_scene_destroy(scene); }
  LoweredScene(const LoweredScene&) = delete;
  LoweredScene& operator=(const LoweredScene&) = delete;
};

// create_image(ny, nx, ns, cam, world) -> String, tests/test.rs:55-85 — same argument order
// (ny before nx), same P3 text out; the pixel loop runs on the GPU.
// n_gpus: 1 = the device in opts (default 0); 0 = every GPU of the node, k > 1 = the first k (rtb_render_multi).
inline std::string create_image(size_t ny, size_t nx, size_t ns, const Camera& cam, const Hittable& world,
                                const rtb_render_opts* opts = nullptr, int n_gpus = 1) {
  LoweredScene ls(world);
  rtb_render_opts o;
  if (opts) o = *opts; else rtb_render_opts_default(&o);
  std::vector<uint8_t> rgb8(3 * nx * ny);
  const int rc = n_gpus == 1
                     ? rtb_render(ls.scene, ls.world, &cam.c, (int)nx, (int)ny, (int)ns, &o, nullptr, rgb8.data(), nullptr)
                     : rtb_render_multi(ls.scene, ls.world, &cam.c, (int)nx, (int)ny, (int)ns, &o, nullptr, n_gpus, nullptr,
                                        rgb8.data(), nullptr);
  if (rc != 0)
    throw std::runtime_error(std::string("create_image: ") + rtb_last_error());
  size_t n = rtb_ppm_p3(rgb8.data(), (int)nx, (int)ny, nullptr, 0);
  std::string out(n, '\0');
  rtb_ppm_p3(rgb8.data(), (int)nx, (int)ny, out.data(), n);
  return out;
}

}  // namespace rtb200
