// oracle.cpp — CPU f64 restatement of the reference renderer (DrStiev/raytracing_rust).
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / `--impl reference` legs may load liboracle.so.  The product
// (librtb200.so) never links, loads or calls it, and has no CPU fallback.
//
// It follows the reference line by line, keeping its control flow (virtual dispatch over the
// object graph, list order, BVH both-children rule, the medium's two-call trick in f64) so
// that its closest-hit answers, ties included, are the reference's.  Each function cites the
// reference file:line it restates.  The Rust reference itself cannot be built here (no
// cargo/rustc, crates not vendored), so this restatement is the pin.
//
// PARITY PIN: the reference's only known-answer vectors are output/final_scene.ppm and
// output/cornell_smoke.ppm (800x800, 100 spp, both all "0 0 0"); tests/test_oracle.py checks
// the oracle reproduces them byte for byte from the as-written scenes.  Everything else the
// reference leaves unpinned (its RNG is unseeded), and is checked here through analytic
// properties (furnace, Beer-Lambert, Schlick) in tests/test_oracle.py.
//
// Randomness: the reference draws rand::thread_rng() (unseeded ChaCha12).  Only the
// distributions are contractual; here every pixel owns a seeded splitmix64 stream and
// gen::<f64>() = (u64 >> 11) * 2^-53 as in rand 0.8.5.
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../include/rtb_scene_format.h"

namespace {

// ---- nalgebra::Vector3<f64> -------------------------------------------------------------
struct V3 {
  double x, y, z;
  V3() : x(0), y(0), z(0) {}
  V3(double a, double b, double c) : x(a), y(b), z(c) {}
  double& operator[](int i) { return (&x)[i]; }
  double operator[](int i) const { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a / norm(a); }
inline V3 mul(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // zip_map(|l,r| l*r)

const double F64_MAX = DBL_MAX;

// Rust `as usize` on f64: saturating, NaN -> 0 (src/perlin.rs:83-85, src/texture.rs:91-92)
inline uint64_t as_usize(double v) {
  if (!(v > 0.0)) return 0;  // negatives and NaN
  if (v >= 18446744073709551615.0) return UINT64_MAX;
  return (uint64_t)v;
}
// Rust `as i32` on f64: saturating, NaN -> 0 (tests/test.rs:76-78)
inline int32_t as_i32(double v) {
  if (v != v) return 0;
  if (v <= -2147483648.0) return INT32_MIN;
  if (v >= 2147483647.0) return INT32_MAX;
  return (int32_t)v;
}

// ---- rand stand-in ------------------------------------------------------------------------
struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next_u64() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double gen() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
  uint64_t gen_range(uint64_t n) { return next_u64() % n; }
};

// ---- src/ray.rs:3-30 -----------------------------------------------------------------------
struct Ray {
  V3 origin, direction;
  double time;
  V3 pointing_at(double t) const { return origin + t * direction; }
};

// ---- src/util.rs:4-24 ---------------------------------------------------------------------
V3 random_in_unit_sphere(Rng& rng) {
  V3 unit(1.0, 1.0, 1.0);
  for (;;) {
    double a = rng.gen(), b = rng.gen(), c = rng.gen();
    V3 p = 2.0 * V3(a, b, c) - unit;
    if (dot(p, p) < 1.0) return p;
  }
}
V3 random_in_unit_disk(Rng& rng) {
  V3 unit(1.0, 1.0, 0.0);
  for (;;) {
    double a = rng.gen(), b = rng.gen();
    V3 p = 2.0 * V3(a, b, 0.0) - unit;
    if (dot(p, p) < 1.0) return p;
  }
}

// ---- src/perlin.rs:38-110 -----------------------------------------------------------------
struct Perlin {
  const RtbPerlinTable* t;
  static double interpolation(const V3 c[2][2][2], double u, double v, double w) {  // :38-56
    double uu = u * u * (3.0 - 2.0 * u);
    double vv = v * v * (3.0 - 2.0 * v);
    double ww = w * w * (3.0 - 2.0 * w);
    double accum = 0.0;
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 2; ++j)
        for (int k = 0; k < 2; ++k) {
          V3 weight(u - i, v - j, w - k);
          accum += (i * uu + (1 - i) * (1.0 - uu)) * (j * vv + (1 - j) * (1.0 - vv)) *
                   (k * ww + (1 - k) * (1.0 - ww)) * dot(c[i][j][k], weight);
        }
    return accum;
  }
  double noise(V3 p) const {  // :76-97
    double u = p.x - std::floor(p.x);
    double v = p.y - std::floor(p.y);
    double w = p.z - std::floor(p.z);
    uint64_t i = as_usize(std::floor(p.x));  // negative -> 0 (Q18)
    uint64_t j = as_usize(std::floor(p.y));
    uint64_t k = as_usize(std::floor(p.z));
    V3 c[2][2][2];
    for (int di = 0; di < 2; ++di)
      for (int dj = 0; dj < 2; ++dj)
        for (int dk = 0; dk < 2; ++dk) {
          unsigned idx = t->perm_x[(i + di) & 255] ^ t->perm_y[(j + dj) & 255] ^ t->perm_z[(k + dk) & 255];
          c[di][dj][dk] = V3(t->ranvec[idx][0], t->ranvec[idx][1], t->ranvec[idx][2]);
        }
    return interpolation(c, u, v, w);
  }
  double turb(V3 p, int depth) const {  // :99-109
    double accum = 0.0;
    V3 temp_p = p;
    double weight = 1.0;
    for (int d = 0; d < depth; ++d) {
      accum += weight * noise(temp_p);
      weight *= 0.5;
      temp_p = temp_p * 2.0;
    }
    return std::fabs(accum);
  }
};

// ---- src/texture.rs -------------------------------------------------------------------------
struct Texture {
  virtual ~Texture() = default;
  virtual V3 value(double u, double v, const V3& p) const = 0;
};
struct SolidTexture : Texture {  // :9-26
  V3 color;
  V3 value(double, double, const V3&) const override { return color; }
};
struct CheckerTexture : Texture {  // :28-49
  const Texture *odd, *even;
  V3 value(double u, double v, const V3& p) const override {
    double s = std::sin(10.0 * p.x) * std::sin(10.0 * p.y) * std::sin(10.0 * p.z);
    if (s < 0.0) return odd->value(u, v, p);
    return even->value(u, v, p);
  }
};
struct NoiseTexture : Texture {  // :51-72
  Perlin noise;
  double scale;
  V3 value(double, double, const V3& p) const override {
    return V3(1.0, 1.0, 1.0) * 0.5 * (1.0 + std::sin(scale * p.x + 5.0 * noise.turb(p, 7)));
  }
};
struct ImageTexture : Texture {  // :74-108
  const uint8_t* data;
  uint32_t nx_, ny_;
  V3 value(double u, double v, const V3&) const override {
    uint64_t nx = nx_, ny = ny_;
    uint64_t i = as_usize(u * (double)nx);
    uint64_t j = as_usize((1.0 - v) * (double)ny);
    if (i > nx - 1) i = nx - 1;
    if (j > ny - 1) j = ny - 1;
    uint64_t idx = 3 * i + 3 * nx * j;
    return V3(data[idx] / 255.0, data[idx + 1] / 255.0, data[idx + 2] / 255.0);
  }
};

// ---- src/hittable.rs:9-16 --------------------------------------------------------------------
struct Material;
struct HitRecord {
  double t, u, v;
  V3 p, normal;
  const Material* material;
  int32_t prim;  // oracle bookkeeping: insertion-order primitive ID (not in the reference)
};

// ---- src/material.rs ---------------------------------------------------------------------------
V3 reflect(const V3& v, const V3& n) { return v - 2.0 * dot(v, n) * n; }  // :9-11
bool refract(const V3& v, const V3& n, double ni_over_nt, V3* out) {      // :13-23
  V3 uv = normalize(v);
  double dt = dot(uv, n);
  double discriminant = 1.0 - ni_over_nt * ni_over_nt * (1.0 - dt * dt);
  if (discriminant > 0.0) {
    *out = ni_over_nt * (uv - n * dt) - n * std::sqrt(discriminant);
    return true;
  }
  return false;
}
double schlick(double cosine, double ref_idx) {  // :25-28
  double r0 = (1.0 - ref_idx) / (1.0 + ref_idx);
  r0 = r0 * r0;
  double m = 1.0 - cosine;
  return r0 + (1.0 - r0) * (m * m * m * m * m);
}

struct Material {
  int type = 0;  // RtbMatType, for counters only
  virtual ~Material() = default;
  virtual bool scatter(const Ray& ray, const HitRecord& hit, Rng& rng, Ray* scattered, V3* attenuation) const = 0;
  virtual V3 emitted(double, double, const V3&) const { return V3(); }
};
struct Lambertian : Material {  // :36-58
  const Texture* albedo;
  bool scatter(const Ray& ray, const HitRecord& hit, Rng& rng, Ray* s, V3* a) const override {
    V3 target = hit.p + hit.normal + random_in_unit_sphere(rng);
    *s = Ray{hit.p, target - hit.p, ray.time};
    *a = albedo->value(hit.u, hit.v, hit.p);
    return true;
  }
};
struct Metal : Material {  // :61-92
  const Texture* albedo;
  double fuzz;
  bool scatter(const Ray& ray, const HitRecord& hit, Rng& rng, Ray* s, V3* a) const override {
    V3 reflected = reflect(normalize(ray.direction), hit.normal);
    if (fuzz > 0.0) reflected = reflected + fuzz * random_in_unit_sphere(rng);
    if (dot(reflected, hit.normal) > 0.0) {
      *s = Ray{hit.p, reflected, ray.time};
      *a = albedo->value(hit.u, hit.v, hit.p);
      return true;
    }
    return false;
  }
};
struct Dielectric : Material {  // :95-131
  double ref_idx;
  bool scatter(const Ray& ray, const HitRecord& hit, Rng& rng, Ray* s, V3* a) const override {
    *a = V3(1.0, 1.0, 1.0);
    V3 outward_normal;
    double ni_over_nt, cosine;
    if (dot(ray.direction, hit.normal) > 0.0) {
      cosine = ref_idx * dot(ray.direction, hit.normal) / norm(ray.direction);
      outward_normal = -hit.normal;
      ni_over_nt = ref_idx;
    } else {
      cosine = -dot(ray.direction, hit.normal) / norm(ray.direction);
      outward_normal = hit.normal;
      ni_over_nt = 1.0 / ref_idx;
    }
    V3 refracted;
    if (refract(ray.direction, outward_normal, ni_over_nt, &refracted)) {
      double reflect_prob = schlick(cosine, ref_idx);
      if (rng.gen() >= reflect_prob) {
        *s = Ray{hit.p, refracted, ray.time};
        return true;
      }
    }
    V3 reflected = reflect(ray.direction, hit.normal);
    *s = Ray{hit.p, reflected, ray.time};
    return true;
  }
};
struct DiffuseLight : Material {  // :133-151
  const Texture* emit;
  bool scatter(const Ray&, const HitRecord&, Rng&, Ray*, V3*) const override { return false; }
  V3 emitted(double u, double v, const V3& p) const override { return emit->value(u, v, p); }
};
struct Isotropic : Material {  // :154-173
  const Texture* albedo;
  bool scatter(const Ray& ray, const HitRecord& hit, Rng& rng, Ray* s, V3* a) const override {
    *s = Ray{hit.p, random_in_unit_sphere(rng), ray.time};
    *a = albedo->value(hit.u, hit.v, hit.p);
    return true;
  }
};

// ---- src/aabb.rs ----------------------------------------------------------------------------
struct AABB {
  V3 min, max;
  bool hit(const Ray& ray, double t_min, double t_max) const {  // :31-44
    for (int a = 0; a < 3; ++a) {
      double inv_d = 1.0 / ray.direction[a];
      double t0 = (min[a] - ray.origin[a]) * inv_d;
      double t1 = (max[a] - ray.origin[a]) * inv_d;
      if (inv_d < 0.0) std::swap(t0, t1);
      t_min = std::fmax(t_min, t0);  // f64::max: NaN-ignoring, like fmax
      t_max = std::fmin(t_max, t1);
      if (t_max <= t_min) return false;
    }
    return true;
  }
};
AABB surrounding_box(const AABB& a, const AABB& b) {  // :6-18
  return AABB{V3(std::fmin(a.min.x, b.min.x), std::fmin(a.min.y, b.min.y), std::fmin(a.min.z, b.min.z)),
              V3(std::fmax(a.max.x, b.max.x), std::fmax(a.max.y, b.max.y), std::fmax(a.max.z, b.max.z))};
}

// Per-query context: the rng (thread_rng in the reference) and oracle-only switches/counters.
struct Ctx {
  Rng* rng;
  bool media_enabled;
  uint64_t aabb_tests = 0, prim_tests = 0;
};

// ---- src/hittable.rs:18-21 -------------------------------------------------------------------
struct Hittable {
  virtual ~Hittable() = default;
  virtual bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const = 0;
  virtual bool bounding_box(double t0, double t1, AABB* out) const = 0;
};

void get_sphere_uv(const V3& p, double* u, double* v) {  // src/sphere.rs:9-15
  double phi = std::atan2(p.z, p.x);
  double theta = std::asin(p.y);
  *u = 1.0 - (phi + M_PI) / (2.0 * M_PI);
  *v = (theta + M_2_PI) / M_PI;  // FRAC_2_PI = 2/pi, as written (Q5)
}

struct Sphere : Hittable {  // src/sphere.rs:20-85
  V3 center;
  double radius;
  const Material* material;
  int32_t prim;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    ++cx.prim_tests;
    V3 oc = ray.origin - center;
    double a = dot(ray.direction, ray.direction);
    double b = dot(oc, ray.direction);
    double c = dot(oc, oc) - radius * radius;
    double discriminant = b * b - a * c;
    if (discriminant > 0.0) {
      double sq = std::sqrt(discriminant);
      double t = (-b - sq) / a;
      if (t < t_max && t > t_min) return fill(ray, t, center, out);
      t = (-b + sq) / a;
      if (t < t_max && t > t_min) return fill(ray, t, center, out);
    }
    return false;
  }
  bool fill(const Ray& ray, double t, const V3& c, HitRecord* out) const {
    V3 p = ray.pointing_at(t);
    V3 normal = (p - c) / radius;
    double u, v;
    get_sphere_uv(normal, &u, &v);
    *out = HitRecord{t, u, v, p, normal, material, prim};
    return true;
  }
  bool bounding_box(double, double, AABB* out) const override {  // :79-84
    V3 r(radius, radius, radius);
    *out = AABB{center - r, center + r};
    return true;
  }
};

struct MovingSphere : Hittable {  // src/sphere.rs:87-175
  V3 center0, center1;
  double time0, time1, radius;
  const Material* material;
  int32_t prim;
  V3 center(double time) const {  // :115-118
    return center0 + ((time - time0) / (time1 - time0)) * (center1 - center0);
  }
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    ++cx.prim_tests;
    V3 ctr = center(ray.time);
    V3 oc = ray.origin - ctr;
    double a = dot(ray.direction, ray.direction);
    double b = dot(oc, ray.direction);
    double c = dot(oc, oc) - radius * radius;
    double discriminant = b * b - a * c;
    if (discriminant > 0.0) {
      double sq = std::sqrt(discriminant);
      double t = (-b - sq) / a;
      if (t < t_max && t > t_min) return fill(ray, t, ctr, out);
      t = (-b + sq) / a;
      if (t < t_max && t > t_min) return fill(ray, t, ctr, out);
    }
    return false;
  }
  bool fill(const Ray& ray, double t, const V3& c, HitRecord* out) const {
    V3 p = ray.pointing_at(t);
    V3 normal = (p - c) / radius;
    double u, v;
    get_sphere_uv(normal, &u, &v);
    *out = HitRecord{t, u, v, p, normal, material, prim};
    return true;
  }
  bool bounding_box(double t0, double t1, AABB* out) const override {  // :165-174
    V3 r(radius, radius, radius);
    AABB a0{center(t0) - r, center(t0) + r};
    AABB a1{center(t1) - r, center(t1) + r};
    *out = surrounding_box(a0, a1);
    return true;
  }
};

struct Rect : Hittable {  // src/rect.rs:14-76
  int plane;
  double x0, y0, x1, y1, k;
  const Material* material;
  int32_t prim;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    ++cx.prim_tests;
    int k_axis, a_axis, b_axis;
    switch (plane) {
      case RTB_PLANE_YZ: k_axis = 0, a_axis = 1, b_axis = 2; break;
      case RTB_PLANE_ZX: k_axis = 1, a_axis = 2, b_axis = 0; break;
      default: k_axis = 2, a_axis = 0, b_axis = 1; break;
    }
    double t = (k - ray.origin[k_axis]) / ray.direction[k_axis];
    if (t < t_min || t > t_max) return false;
    double x = ray.origin[a_axis] + t * ray.direction[a_axis];
    double y = ray.origin[b_axis] + t * ray.direction[b_axis];
    if (x < x0 || x > x1 || y < y0 || y > y1) return false;
    double u = (x - x0) / (x1 - x0);
    double v = (y - y0) / (y1 - y0);
    V3 p = ray.pointing_at(t);
    V3 normal;
    normal[k_axis] = 1.0;
    *out = HitRecord{t, u, v, p, normal, material, prim};
    return true;
  }
  bool bounding_box(double, double, AABB* out) const override {  // :71-75, ignores `plane` (Q8)
    *out = AABB{V3(x0, y0, k - 0.0001), V3(x1, y1, k + 0.0001)};
    return true;
  }
};

struct HittableList : Hittable {  // src/hittable.rs:23-65
  std::vector<const Hittable*> list;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    double closest_so_far = t_max;
    bool hit_anything = false;
    HitRecord rec;
    for (const Hittable* h : list) {
      if (h->hit(ray, t_min, closest_so_far, cx, &rec)) {
        closest_so_far = rec.t;
        *out = rec;
        hit_anything = true;
      }
    }
    return hit_anything;
  }
  bool bounding_box(double t0, double t1, AABB* out) const override {
    if (list.empty()) return false;
    AABB acc;
    if (!list[0]->bounding_box(t0, t1, &acc)) return false;
    for (size_t i = 1; i < list.size(); ++i) {
      AABB b;
      if (!list[i]->bounding_box(t0, t1, &b)) return false;
      acc = surrounding_box(acc, b);
    }
    *out = acc;
    return true;
  }
};

struct Cube : Hittable {  // src/cube.rs:8-94
  V3 p_min, p_max;
  HittableList sides;
  std::vector<std::unique_ptr<Rect>> rects;
  Cube(V3 a, V3 b, const Material* m, int32_t prim0) : p_min(a), p_max(b) {
    auto push = [&](int plane, double x0, double y0, double x1, double y1, double k) {
      auto r = std::make_unique<Rect>();
      r->plane = plane, r->x0 = x0, r->y0 = y0, r->x1 = x1, r->y1 = y1, r->k = k;
      r->material = m;
      r->prim = prim0 + (int32_t)rects.size();
      sides.list.push_back(r.get());
      rects.push_back(std::move(r));
    };
    push(RTB_PLANE_XY, a.x, a.y, b.x, b.y, b.z);  // :21-29
    push(RTB_PLANE_XY, a.x, a.y, b.x, b.y, a.z);  // :30-38
    push(RTB_PLANE_ZX, a.z, a.x, b.z, b.x, b.y);  // :39-47
    push(RTB_PLANE_ZX, a.z, a.x, b.z, b.x, a.y);  // :48-56
    push(RTB_PLANE_YZ, a.y, a.z, b.y, b.z, b.x);  // :57-65
    push(RTB_PLANE_YZ, a.y, a.z, b.y, b.z, a.x);  // :66-74
  }
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    return sides.hit(ray, t_min, t_max, cx, out);
  }
  bool bounding_box(double, double, AABB* out) const override {
    *out = AABB{p_min, p_max};
    return true;
  }
};

struct FlipNormals : Hittable {  // src/hittable.rs:67-88
  const Hittable* hittable;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    if (!hittable->hit(ray, t_min, t_max, cx, out)) return false;
    out->normal = -out->normal;
    return true;
  }
  bool bounding_box(double t0, double t1, AABB* out) const override { return hittable->bounding_box(t0, t1, out); }
};

void get_axis(int axis, int* r, int* a, int* b) {  // src/rotate.rs:13-19
  switch (axis) {
    case RTB_AXIS_X: *r = 0, *a = 1, *b = 2; break;
    case RTB_AXIS_Y: *r = 1, *a = 2, *b = 0; break;
    default: *r = 2, *a = 0, *b = 1; break;
  }
}

struct Rotate : Hittable {  // src/rotate.rs:21-118
  int axis;
  double sin_theta, cos_theta;
  const Hittable* hittable;
  bool has_bbox;
  AABB bbox;
  Rotate(int axis_, const Hittable* h, double angle) : axis(axis_), hittable(h) {  // :30-81
    int r_axis, a_axis, b_axis;
    get_axis(axis, &r_axis, &a_axis, &b_axis);
    double radians = (M_PI / 180.0) * angle;
    sin_theta = std::sin(radians);
    cos_theta = std::cos(radians);
    AABB bb;
    has_bbox = h->bounding_box(0.0, 1.0, &bb);
    if (has_bbox) {
      // as written: max starts at f64::MAX and min at f64::MIN, so no corner ever updates
      // them and the box is (-MAX, +MAX)^3 (Q10)
      V3 max(F64_MAX, F64_MAX, F64_MAX);
      V3 min(-F64_MAX, -F64_MAX, -F64_MAX);
      for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j)
          for (int k = 0; k < 2; ++k) {
            double r = k * bb.max[r_axis] + (1 - k) * bb.min[r_axis];
            double a = i * bb.max[a_axis] + (1 - i) * bb.min[a_axis];
            double b = j * bb.max[b_axis] + (1 - j) * bb.min[b_axis];
            double new_a = cos_theta * a - sin_theta * b;
            double new_b = sin_theta * a + cos_theta * b;
            if (new_a < min[a_axis]) min[a_axis] = new_a;
            if (new_b < min[b_axis]) min[b_axis] = new_b;
            if (r < min[r_axis]) min[r_axis] = r;
            if (new_a > max[a_axis]) max[a_axis] = new_a;
            if (new_b > max[b_axis]) max[b_axis] = new_b;
            if (r > max[r_axis]) max[r_axis] = r;
          }
      bbox = AABB{min, max};
    }
  }
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {  // :85-113
    int r_axis, a_axis, b_axis;
    get_axis(axis, &r_axis, &a_axis, &b_axis);
    V3 origin = ray.origin, direction = ray.direction;
    origin[a_axis] = cos_theta * ray.origin[a_axis] + sin_theta * ray.origin[b_axis];
    origin[b_axis] = -sin_theta * ray.origin[a_axis] + cos_theta * ray.origin[b_axis];
    direction[a_axis] = cos_theta * ray.direction[a_axis] + sin_theta * ray.direction[b_axis];
    direction[b_axis] = -sin_theta * ray.direction[a_axis] + cos_theta * ray.direction[b_axis];
    Ray rotated{origin, direction, ray.time};
    if (!hittable->hit(rotated, t_min, t_max, cx, out)) return false;
    V3 p = out->p, normal = out->normal;
    p[a_axis] = cos_theta * out->p[a_axis] - sin_theta * out->p[b_axis];
    p[b_axis] = sin_theta * out->p[a_axis] + cos_theta * out->p[b_axis];
    normal[a_axis] = cos_theta * out->normal[a_axis] - sin_theta * out->normal[b_axis];
    normal[b_axis] = sin_theta * out->normal[a_axis] + cos_theta * out->normal[b_axis];
    out->p = p;
    out->normal = normal;
    return true;
  }
  bool bounding_box(double, double, AABB* out) const override {
    if (has_bbox) *out = bbox;
    return has_bbox;
  }
};

struct Traslate : Hittable {  // src/traslate.rs:6-33
  const Hittable* hitable;
  V3 offset;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    Ray moved{ray.origin - offset, ray.direction, ray.time};
    if (!hitable->hit(moved, t_min, t_max, cx, out)) return false;
    out->p = out->p + offset;
    return true;
  }
  bool bounding_box(double t0, double t1, AABB* out) const override {
    if (!hitable->bounding_box(t0, t1, out)) return false;
    out->min = out->min + offset;
    out->max = out->max + offset;
    return true;
  }
};

struct ConstantMedium : Hittable {  // src/medium.rs:11-61
  const Hittable* boundary;
  double density;
  Isotropic phase_function;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {
    if (!cx.media_enabled) return false;  // oracle-only switch for the closest-hit ID test
    HitRecord hit1, hit2;
    if (boundary->hit(ray, -F64_MAX, F64_MAX, cx, &hit1)) {
      if (boundary->hit(ray, hit1.t + 0.0001, F64_MAX, cx, &hit2)) {
        if (hit1.t < t_min) hit1.t = t_min;
        if (hit2.t > t_max) hit2.t = t_max;
        if (hit1.t < hit2.t) {
          double distance_inside_boundary = (hit2.t - hit1.t) * norm(ray.direction);
          double hit_distance = -(1.0 / density) * std::log(cx.rng->gen());
          if (hit_distance < distance_inside_boundary) {
            double t = hit1.t + hit_distance / norm(ray.direction);
            *out = HitRecord{t, 0.0, 0.0, ray.pointing_at(t), V3(1.0, 0.0, 0.0), &phase_function, -2};
            return true;
          }
        }
      }
    }
    return false;
  }
  bool bounding_box(double t0, double t1, AABB* out) const override { return boundary->bounding_box(t0, t1, out); }
};

struct BVHNode : Hittable {  // src/bvh.rs:9-94
  const Hittable *left, *right;
  AABB bbox;
  bool hit(const Ray& ray, double t_min, double t_max, Ctx& cx, HitRecord* out) const override {  // :70-89
    ++cx.aabb_tests;
    if (bbox.hit(ray, t_min, t_max)) {
      HitRecord l, r;
      bool hl = left->hit(ray, t_min, t_max, cx, &l);
      bool hr = right->hit(ray, t_min, t_max, cx, &r);
      if (hl && hr) {
        *out = (l.t < r.t) ? l : r;
        return true;
      }
      if (hl) { *out = l; return true; }
      if (hr) { *out = r; return true; }
    }
    return false;
  }
  bool bounding_box(double, double, AABB* out) const override {
    *out = bbox;
    return true;
  }
};

// ---- scene loaded from .rtscene bytes ------------------------------------------------------
struct Scene {
  std::vector<uint8_t> bytes;  // keeps perlin tables / texels alive
  std::vector<std::unique_ptr<Texture>> tex;
  std::vector<std::unique_ptr<Material>> mat;
  std::vector<std::unique_ptr<Hittable>> owned;  // every instantiated hittable
  const Hittable* world = nullptr;
  int32_t n_prims = 0;
  std::string error;
  // bvh shape summary for tests
  uint64_t bvh_nodes = 0;
};

struct Loader {
  Scene& sc;
  const RtbTexRec* texrec;
  const RtbMatRec* matrec;
  const RtbObjRec* objrec;
  const int32_t* list_pool;
  RtbSceneHeader h;

  template <class T>
  T* own(std::unique_ptr<T> p) {
    T* raw = p.get();
    sc.owned.push_back(std::move(p));
    return raw;
  }

  // BVHNode::new, src/bvh.rs:17-66: random axis, in-place sort on bbox.min[axis] with the
  // never-Equal comparator, median split, dup leaf on len==1.
  const Hittable* bvh_new(const Hittable** hs, size_t len, double time0, double time1, Rng& rng) {
    int axis = (int)rng.gen_range(3);
    bool no_bbox = false;
    std::stable_sort(hs, hs + len, [&](const Hittable* a, const Hittable* b) {
      AABB ba, bb;
      if (!a->bounding_box(time0, time1, &ba) || !b->bounding_box(time0, time1, &bb)) {
        no_bbox = true;
        return false;
      }
      return ba.min[axis] - bb.min[axis] < 0.0;
    });
    if (no_bbox) { sc.error = "No bounding box in BVHNode"; return nullptr; }
    auto node = std::make_unique<BVHNode>();
    if (len == 1) {
      node->left = hs[0], node->right = hs[0];
    } else if (len == 2) {
      node->left = hs[0], node->right = hs[1];
    } else {
      node->left = bvh_new(hs, len / 2, time0, time1, rng);
      node->right = bvh_new(hs + len / 2, len - len / 2, time0, time1, rng);
      if (!node->left || !node->right) return nullptr;
    }
    AABB lb, rb;
    if (!node->left->bounding_box(time0, time1, &lb) || !node->right->bounding_box(time0, time1, &rb)) {
      sc.error = "No bounding box in BVHNode";
      return nullptr;
    }
    node->bbox = surrounding_box(lb, rb);
    ++sc.bvh_nodes;
    return own(std::move(node));
  }

  // Instantiate object `h` (a fresh instance per reference: the graph is a tree of values).
  // Primitive IDs are handed out depth-first in insertion order.
  const Hittable* build(int hnd) {
    if (hnd < 0 || (uint32_t)hnd >= h.n_obj) { sc.error = "bad object handle"; return nullptr; }
    const RtbObjRec& o = objrec[hnd];
    switch (o.type) {
      case RTB_OBJ_SPHERE: {
        auto s = std::make_unique<Sphere>();
        s->center = V3(o.v[0], o.v[1], o.v[2]), s->radius = o.v[3];
        s->material = sc.mat[o.mat].get(), s->prim = sc.n_prims++;
        return own(std::move(s));
      }
      case RTB_OBJ_MOVING_SPHERE: {
        auto s = std::make_unique<MovingSphere>();
        s->center0 = V3(o.v[0], o.v[1], o.v[2]), s->center1 = V3(o.v[3], o.v[4], o.v[5]);
        s->time0 = o.v[6], s->time1 = o.v[7], s->radius = o.v[8];
        s->material = sc.mat[o.mat].get(), s->prim = sc.n_prims++;
        return own(std::move(s));
      }
      case RTB_OBJ_RECT: {
        auto r = std::make_unique<Rect>();
        r->plane = o.sel, r->x0 = o.v[0], r->y0 = o.v[1], r->x1 = o.v[2], r->y1 = o.v[3], r->k = o.v[4];
        r->material = sc.mat[o.mat].get(), r->prim = sc.n_prims++;
        return own(std::move(r));
      }
      case RTB_OBJ_CUBE: {
        auto c = std::make_unique<Cube>(V3(o.v[0], o.v[1], o.v[2]), V3(o.v[3], o.v[4], o.v[5]),
                                        sc.mat[o.mat].get(), sc.n_prims);
        sc.n_prims += 6;
        return own(std::move(c));
      }
      case RTB_OBJ_FLIP_NORMALS: {
        const Hittable* c = build(o.child);
        if (!c) return nullptr;
        auto f = std::make_unique<FlipNormals>();
        f->hittable = c;
        return own(std::move(f));
      }
      case RTB_OBJ_ROTATE: {
        const Hittable* c = build(o.child);
        if (!c) return nullptr;
        return own(std::make_unique<Rotate>(o.sel, c, o.v[0]));
      }
      case RTB_OBJ_TRANSLATE: {
        const Hittable* c = build(o.child);
        if (!c) return nullptr;
        auto t = std::make_unique<Traslate>();
        t->hitable = c, t->offset = V3(o.v[0], o.v[1], o.v[2]);
        return own(std::move(t));
      }
      case RTB_OBJ_CONSTANT_MEDIUM: {
        const Hittable* c = build(o.child);
        if (!c) return nullptr;
        auto m = std::make_unique<ConstantMedium>();
        m->boundary = c, m->density = o.v[0];
        m->phase_function.type = RTB_MAT_ISOTROPIC;
        m->phase_function.albedo = sc.tex[o.tex].get();
        return own(std::move(m));
      }
      case RTB_OBJ_LIST: {
        auto l = std::make_unique<HittableList>();
        for (int k = 0; k < o.list_n; ++k) {
          const Hittable* c = build(list_pool[o.list_off + k]);
          if (!c) return nullptr;
          l->list.push_back(c);
        }
        return own(std::move(l));
      }
      case RTB_OBJ_BVH: {
        std::vector<const Hittable*> hs;
        for (int k = 0; k < o.list_n; ++k) {
          const Hittable* c = build(list_pool[o.list_off + k]);
          if (!c) return nullptr;
          hs.push_back(c);
        }
        if (hs.empty()) { sc.error = "No bounding box in BVHNode"; return nullptr; }
        Rng rng(o.seed);
        return bvh_new(hs.data(), hs.size(), o.v[0], o.v[1], rng);
      }
    }
    sc.error = "unknown object type";
    return nullptr;
  }
};

// ---- src/camera.rs --------------------------------------------------------------------------
struct Camera {
  V3 origin, lower_left_corner, horizontal, vertical, u, v;
  double time0, time1, lens_radius;
  explicit Camera(const RtbCamera& c) {  // :21-51
    V3 look_from(c.look_from[0], c.look_from[1], c.look_from[2]);
    V3 look_at(c.look_at[0], c.look_at[1], c.look_at[2]);
    V3 view_up(c.view_up[0], c.view_up[1], c.view_up[2]);
    double theta = c.vertical_fov * M_PI / 180.0;
    double half_height = c.focus_dist * std::tan(theta / 2.0);
    double half_width = c.aspect * half_height;
    V3 w = normalize(look_from - look_at);
    u = normalize(cross(view_up, w));
    v = cross(w, u);
    origin = look_from;
    lower_left_corner = look_from - half_width * u - half_height * v - c.focus_dist * w;
    horizontal = 2.0 * half_width * u;
    vertical = 2.0 * half_height * v;
    time0 = c.time0, time1 = c.time1;
    lens_radius = c.aperture / 2.0;
  }
  // :53-67 with the two random inputs made explicit (rd = unit-disk sample, xi_time in [0,1))
  Ray get_ray_with(double s, double t, V3 rd_unit, double xi_time) const {
    V3 org = origin;
    if (lens_radius != 0.0) {
      V3 rd = lens_radius * rd_unit;
      V3 offset = u * rd.x + v * rd.y;
      org = origin + offset;
    }
    double time = time0 + xi_time * (time1 - time0);
    return Ray{org, lower_left_corner + s * horizontal + t * vertical - org, time};
  }
  Ray get_ray(double s, double t, Rng& rng) const {
    V3 rd;
    if (lens_radius != 0.0) rd = random_in_unit_disk(rng);
    double xi = rng.gen();
    return get_ray_with(s, t, rd, xi);
  }
};

// ---- src/color.rs:6-23 -------------------------------------------------------------------------
struct RenderParams {
  int max_depth;
  double t_min;
  int background;  // 0 black (as written), 1 sky (the commented-out gradient :18-20)
};

V3 color(const Ray& ray, const Hittable* world, int depth, Ctx& cx, const RenderParams& rp, uint64_t* segments) {
  ++*segments;
  HitRecord hit;
  if (world->hit(ray, rp.t_min, F64_MAX, cx, &hit)) {
    V3 emitted = hit.material->emitted(hit.u, hit.v, hit.p);
    if (depth < rp.max_depth) {
      Ray scattered;
      V3 attenuation;
      if (hit.material->scatter(ray, hit, *cx.rng, &scattered, &attenuation))
        return emitted + mul(attenuation, color(scattered, world, depth + 1, cx, rp, segments));
    }
    return emitted;
  }
  if (rp.background == 1) {
    V3 unit_direction = normalize(ray.direction);
    double t = 0.5 * (unit_direction.y + 1.0);
    return (1.0 - t) * V3(1.0, 1.0, 1.0) + t * V3(0.5, 0.7, 1.0);
  }
  return V3(0.0, 0.0, 0.0);
}

inline uint64_t pixel_seed(uint64_t seed, uint64_t pixel) {
  Rng r(seed ^ (pixel * 0xD1342543DE82EF95ull + 0x632BE59BD9B4E019ull));
  r.next_u64();
  return r.next_u64();
}

}  // namespace

// ================================ C interface for the tests ===================================
extern "C" {

struct orc_scene;  // opaque = Scene

const char* orc_about() { return "oracle: CPU f64 restatement of raytracing_rust (test infrastructure)"; }

void* orc_scene_load(const void* buf, size_t len, int world, char* err, size_t errcap) {
  auto fail = [&](const std::string& m) -> void* {
    if (err && errcap) { std::strncpy(err, m.c_str(), errcap - 1); err[errcap - 1] = 0; }
    return nullptr;
  };
  if (!buf || len < sizeof(RtbSceneHeader)) return fail("short buffer");
  auto sc = std::make_unique<Scene>();
  sc->bytes.assign((const uint8_t*)buf, (const uint8_t*)buf + len);
  Loader L{*sc, nullptr, nullptr, nullptr, nullptr, {}};
  std::memcpy(&L.h, sc->bytes.data(), sizeof L.h);
  if (L.h.magic != RTB_SCENE_MAGIC || L.h.version != RTB_SCENE_VERSION) return fail("bad magic");
  const uint8_t* p = sc->bytes.data() + sizeof L.h;
  L.texrec = (const RtbTexRec*)p; p += (size_t)L.h.n_tex * sizeof(RtbTexRec);
  L.matrec = (const RtbMatRec*)p; p += (size_t)L.h.n_mat * sizeof(RtbMatRec);
  L.objrec = (const RtbObjRec*)p; p += (size_t)L.h.n_obj * sizeof(RtbObjRec);
  L.list_pool = (const int32_t*)p; p += (size_t)L.h.n_list * sizeof(int32_t);
  const RtbPerlinTable* perlin = (const RtbPerlinTable*)p; p += (size_t)L.h.n_perlin * sizeof(RtbPerlinTable);
  const uint8_t* images = p; p += L.h.n_image_bytes;
  if ((size_t)(p - sc->bytes.data()) > len) return fail("truncated");
  // textures (children precede parents by construction)
  for (uint32_t i = 0; i < L.h.n_tex; ++i) {
    const RtbTexRec& t = L.texrec[i];
    switch (t.type) {
      case RTB_TEX_SOLID: {
        auto s = std::make_unique<SolidTexture>();
        s->color = V3(t.v[0], t.v[1], t.v[2]);
        sc->tex.push_back(std::move(s));
        break;
      }
      case RTB_TEX_CHECKER: {
        if (t.a < 0 || (uint32_t)t.a >= i || t.b < 0 || (uint32_t)t.b >= i) return fail("bad checker handle");
        auto c = std::make_unique<CheckerTexture>();
        c->odd = sc->tex[t.a].get(), c->even = sc->tex[t.b].get();
        sc->tex.push_back(std::move(c));
        break;
      }
      case RTB_TEX_NOISE: {
        if (t.a < 0 || (uint32_t)t.a >= L.h.n_perlin) return fail("bad perlin index");
        auto n = std::make_unique<NoiseTexture>();
        n->noise.t = &perlin[t.a], n->scale = t.v[0];
        sc->tex.push_back(std::move(n));
        break;
      }
      case RTB_TEX_IMAGE: {
        auto im = std::make_unique<ImageTexture>();
        im->data = images + t.image_off, im->nx_ = (uint32_t)t.a, im->ny_ = (uint32_t)t.b;
        sc->tex.push_back(std::move(im));
        break;
      }
      default: return fail("bad texture type");
    }
  }
  for (uint32_t i = 0; i < L.h.n_mat; ++i) {
    const RtbMatRec& m = L.matrec[i];
    const Texture* tx = (m.tex >= 0 && (uint32_t)m.tex < L.h.n_tex) ? sc->tex[m.tex].get() : nullptr;
    if (m.type != RTB_MAT_DIELECTRIC && !tx) return fail("bad material texture");
    std::unique_ptr<Material> out;
    switch (m.type) {
      case RTB_MAT_LAMBERTIAN: { auto x = std::make_unique<Lambertian>(); x->albedo = tx; out = std::move(x); break; }
      case RTB_MAT_METAL: {
        auto x = std::make_unique<Metal>();
        x->albedo = tx;
        x->fuzz = m.param < 1.0 ? m.param : 1.0;  // Metal::new, src/material.rs:70
        out = std::move(x);
        break;
      }
      case RTB_MAT_DIELECTRIC: { auto x = std::make_unique<Dielectric>(); x->ref_idx = m.param; out = std::move(x); break; }
      case RTB_MAT_DIFFUSE_LIGHT: { auto x = std::make_unique<DiffuseLight>(); x->emit = tx; out = std::move(x); break; }
      case RTB_MAT_ISOTROPIC: { auto x = std::make_unique<Isotropic>(); x->albedo = tx; out = std::move(x); break; }
      default: return fail("bad material type");
    }
    out->type = m.type;
    sc->mat.push_back(std::move(out));
  }
  sc->world = L.build(world);
  if (!sc->world) return fail(sc->error.empty() ? "build failed" : sc->error);
  return sc.release();
}

void orc_scene_free(void* s) { delete (Scene*)s; }
int orc_scene_prim_count(void* s) { return ((Scene*)s)->n_prims; }
uint64_t orc_scene_bvh_nodes(void* s) { return ((Scene*)s)->bvh_nodes; }

// create_image, tests/test.rs:55-85.  Rows are distributed over `nthreads` threads; each
// pixel owns a seeded stream, so the image does not depend on the thread count.
// out_mean_rgb: 3*nx*ny doubles, row 0 = top (j = ny-1), the per-pixel mean BEFORE gamma;
// out_rgb8: 3*nx*ny bytes, the reference quantisation.  Either may be NULL.
// Returns the number of ray segments (world.hit queries issued by color()).
uint64_t orc_render(void* scene, const RtbCamera* camera, int nx, int ny, int ns, int max_depth, double t_min,
                    int background, uint64_t seed, int nthreads, double* out_mean_rgb, uint8_t* out_rgb8) {
  Scene* sc = (Scene*)scene;
  Camera cam(*camera);
  RenderParams rp{max_depth, t_min, background};
  if (nthreads < 1) nthreads = 1;
  std::atomic<int> next_row{0};
  std::atomic<uint64_t> total_segments{0};
  auto worker = [&]() {
    uint64_t segs = 0;
    for (;;) {
      int row = next_row.fetch_add(1);  // row 0 = top = j = ny-1  (tests/test.rs:62)
      if (row >= ny) break;
      int j = ny - 1 - row;
      for (int i = 0; i < nx; ++i) {
        Rng rng(pixel_seed(seed, (uint64_t)row * nx + i));
        Ctx cx{&rng, true};
        V3 col(0.0, 0.0, 0.0);
        for (int s = 0; s < ns; ++s) {
          double u = ((double)i + rng.gen()) / (double)nx;
          double v = ((double)j + rng.gen()) / (double)ny;
          Ray ray = cam.get_ray(u, v, rng);
          col = col + color(ray, sc->world, 0, cx, rp, &segs);
        }
        col = col / (double)ns;
        size_t o = 3 * ((size_t)row * nx + i);
        if (out_mean_rgb) out_mean_rgb[o] = col.x, out_mean_rgb[o + 1] = col.y, out_mean_rgb[o + 2] = col.z;
        if (out_rgb8) {
          for (int c = 0; c < 3; ++c) {
            double x = std::sqrt(col[c]);
            x = x > 1.0 ? 1.0 : (x < 0.0 ? 0.0 : x);  // nalgebra::clamp (NaN passes through)
            int32_t q = as_i32(255.99 * x);
            out_rgb8[o + c] = (uint8_t)q;
          }
        }
      }
    }
    total_segments += segs;
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; ++t) th.emplace_back(worker);
  worker();
  for (auto& t : th) t.join();
  return total_segments.load();
}

// world.hit(ray, t_min, f64::MAX) (src/color.rs:7) for a batch, media disabled.
// rays: n x {ox,oy,oz,dx,dy,dz,time}.  out_prim: primitive ID or -1.  out_rec (optional):
// n x {t,u,v,px,py,pz,nx,ny,nz}.
void orc_closest_hit(void* scene, const double* rays, size_t n, double t_min, int media_enabled, uint64_t seed,
                     int32_t* out_prim, double* out_t, double* out_rec, uint64_t* counters) {
  Scene* sc = (Scene*)scene;
  Rng rng(seed);
  Ctx cx{&rng, media_enabled != 0};
  for (size_t i = 0; i < n; ++i) {
    const double* r = rays + 7 * i;
    Ray ray{V3(r[0], r[1], r[2]), V3(r[3], r[4], r[5]), r[6]};
    HitRecord hit;
    if (sc->world->hit(ray, t_min, F64_MAX, cx, &hit)) {
      out_prim[i] = hit.prim;
      if (out_t) out_t[i] = hit.t;
      if (out_rec) {
        double* o = out_rec + 9 * i;
        o[0] = hit.t, o[1] = hit.u, o[2] = hit.v;
        o[3] = hit.p.x, o[4] = hit.p.y, o[5] = hit.p.z;
        o[6] = hit.normal.x, o[7] = hit.normal.y, o[8] = hit.normal.z;
      }
    } else {
      out_prim[i] = -1;
      if (out_t) out_t[i] = 0.0;
      if (out_rec) std::memset(out_rec + 9 * i, 0, 9 * sizeof(double));
    }
  }
  if (counters) counters[0] = cx.aabb_tests, counters[1] = cx.prim_tests;
}

// Texture::value for a batch: in n x {u,v,px,py,pz}, out n x {r,g,b}
int orc_texture_value(void* scene, int tex, const double* uvp, size_t n, double* out) {
  Scene* sc = (Scene*)scene;
  if (tex < 0 || (size_t)tex >= sc->tex.size()) return -1;
  for (size_t i = 0; i < n; ++i) {
    const double* q = uvp + 5 * i;
    V3 c = sc->tex[tex]->value(q[0], q[1], V3(q[2], q[3], q[4]));
    out[3 * i] = c.x, out[3 * i + 1] = c.y, out[3 * i + 2] = c.z;
  }
  return 0;
}

// Material::scatter for a batch with per-item seeded rng.  in: n x {ray o[3], d[3], time,
// hit t,u,v,p[3],normal[3]} (16 doubles); out: n x {scattered?, o[3], d[3], atten[3]} (10).
int orc_scatter(void* scene, int mat, const double* in, size_t n, uint64_t seed, double* out) {
  Scene* sc = (Scene*)scene;
  if (mat < 0 || (size_t)mat >= sc->mat.size()) return -1;
  for (size_t i = 0; i < n; ++i) {
    const double* q = in + 16 * i;
    Ray ray{V3(q[0], q[1], q[2]), V3(q[3], q[4], q[5]), q[6]};
    HitRecord hit{q[7], q[8], q[9], V3(q[10], q[11], q[12]), V3(q[13], q[14], q[15]), sc->mat[mat].get(), 0};
    Rng rng(pixel_seed(seed, i));
    Ray s{};
    V3 a;
    bool ok = sc->mat[mat]->scatter(ray, hit, rng, &s, &a);
    double* o = out + 10 * i;
    o[0] = ok ? 1.0 : 0.0;
    o[1] = s.origin.x, o[2] = s.origin.y, o[3] = s.origin.z;
    o[4] = s.direction.x, o[5] = s.direction.y, o[6] = s.direction.z;
    o[7] = a.x, o[8] = a.y, o[9] = a.z;
  }
  return 0;
}

// Camera::get_ray with explicit random inputs: in n x {s,t,rdx,rdy,xi_time}, out n x 7
void orc_camera_rays(const RtbCamera* camera, const double* in, size_t n, double* out) {
  Camera cam(*camera);
  for (size_t i = 0; i < n; ++i) {
    const double* q = in + 5 * i;
    Ray r = cam.get_ray_with(q[0], q[1], V3(q[2], q[3], 0.0), q[4]);
    double* o = out + 7 * i;
    o[0] = r.origin.x, o[1] = r.origin.y, o[2] = r.origin.z;
    o[3] = r.direction.x, o[4] = r.direction.y, o[5] = r.direction.z, o[6] = r.time;
  }
}

// Samplers for distribution tests: kind 0 = random_in_unit_sphere, 1 = random_in_unit_disk
void orc_sample(int kind, uint64_t seed, size_t n, double* out) {
  Rng rng(seed);
  for (size_t i = 0; i < n; ++i) {
    V3 p = kind == 0 ? random_in_unit_sphere(rng) : random_in_unit_disk(rng);
    out[3 * i] = p.x, out[3 * i + 1] = p.y, out[3 * i + 2] = p.z;
  }
}

}  // extern "C"
