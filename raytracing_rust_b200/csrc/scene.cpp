// scene.cpp — scene-graph builders and (de)serialisation behind the C ABI (include/rtb200.h).
// Each builder records the reference constructor call it stands in for; nothing is evaluated
// here.  The CUDA flattener (flatten.cpp) and the CPU oracle both consume these records.
#include "scene.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <exception>
#include <vector>

namespace rtb {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

// src/perlin.rs:12-36 with the draws taken from a seeded stream, in Perlin::new's field
// order (ran_vec, perm_x, perm_y, perm_z; src/perlin.rs:67-74).
static void perlin_generate_perm(SceneRng& rng, uint8_t* p) {
  int tmp[256];
  for (int i = 0; i < 256; ++i) tmp[i] = i;
  for (int i = 255; i >= 0; --i) {  // permute(): src/perlin.rs:4-10
    int target = (int)rng.gen_range((uint64_t)i + 1);
    int t = tmp[i];
    tmp[i] = tmp[target];
    tmp[target] = t;
  }
  for (int i = 0; i < 256; ++i) p[i] = (uint8_t)tmp[i];
}

void make_perlin_table(uint64_t seed, RtbPerlinTable* t) {
  SceneRng rng(seed);
  for (int i = 0; i < 256; ++i) {
    double x = -1.0 + 2.0 * rng.gen_f64();
    double y = -1.0 + 2.0 * rng.gen_f64();
    double z = -1.0 + 2.0 * rng.gen_f64();
    double n = std::sqrt(x * x + y * y + z * z);
    t->ranvec[i][0] = x / n;
    t->ranvec[i][1] = y / n;
    t->ranvec[i][2] = z / n;
  }
  perlin_generate_perm(rng, t->perm_x);
  perlin_generate_perm(rng, t->perm_y);
  perlin_generate_perm(rng, t->perm_z);
}

}  // namespace rtb

using rtb::fail;

extern "C" {

const char* rtb_last_error(void) { return rtb::g_last_error.c_str(); }
const char* rtb_version(void) { return "rtb200 0.1 (sm_100a)"; }

rtb_scene* rtb_scene_create(void) { return new rtb_scene(); }
void rtb_scene_destroy(rtb_scene* s) { delete s; }

#define CHECK_SCENE(s) \
  if (!(s)) return fail(RTB_ERR_INVALID, "null scene")

int rtb_tex_solid(rtb_scene* s, double r, double g, double b) {
  CHECK_SCENE(s);
  RtbTexRec t{};
  t.type = RTB_TEX_SOLID;
  t.a = t.b = -1;
  t.v[0] = r, t.v[1] = g, t.v[2] = b;
  s->tex.push_back(t);
  return (int)s->tex.size() - 1;
}

int rtb_tex_checker(rtb_scene* s, int odd, int even) {
  CHECK_SCENE(s);
  if (!s->tex_ok(odd) || !s->tex_ok(even)) return fail(RTB_ERR_INVALID, "checker: bad texture handle");
  RtbTexRec t{};
  t.type = RTB_TEX_CHECKER;
  t.a = odd, t.b = even;
  s->tex.push_back(t);
  return (int)s->tex.size() - 1;
}

int rtb_tex_noise_tables(rtb_scene* s, double scale, const double* ranvec, const uint8_t* px,
                         const uint8_t* py, const uint8_t* pz) {
  CHECK_SCENE(s);
  if (!ranvec || !px || !py || !pz) return fail(RTB_ERR_INVALID, "noise: null table");
  RtbPerlinTable tab;
  std::memcpy(tab.ranvec, ranvec, sizeof(tab.ranvec));
  std::memcpy(tab.perm_x, px, 256);
  std::memcpy(tab.perm_y, py, 256);
  std::memcpy(tab.perm_z, pz, 256);
  s->perlin.push_back(tab);
  RtbTexRec t{};
  t.type = RTB_TEX_NOISE;
  t.a = (int)s->perlin.size() - 1;
  t.b = -1;
  t.v[0] = scale;
  s->tex.push_back(t);
  return (int)s->tex.size() - 1;
}

int rtb_tex_noise(rtb_scene* s, double scale, uint64_t seed) {
  CHECK_SCENE(s);
  RtbPerlinTable tab;
  rtb::make_perlin_table(seed, &tab);
  return rtb_tex_noise_tables(s, scale, &tab.ranvec[0][0], tab.perm_x, tab.perm_y, tab.perm_z);
}

int rtb_tex_image(rtb_scene* s, const uint8_t* rgb8, uint32_t nx, uint32_t ny) {
  CHECK_SCENE(s);
  if (!rgb8 || nx == 0 || ny == 0) return fail(RTB_ERR_INVALID, "image: empty image");
  RtbTexRec t{};
  t.type = RTB_TEX_IMAGE;
  t.a = (int)nx, t.b = (int)ny;
  t.image_off = s->image_pool.size();
  s->image_pool.insert(s->image_pool.end(), rgb8, rgb8 + (size_t)nx * ny * 3);
  s->tex.push_back(t);
  return (int)s->tex.size() - 1;
}

static int push_mat(rtb_scene* s, int type, int tex, double param) {
  RtbMatRec m{};
  m.type = type, m.tex = tex, m.param = param;
  s->mat.push_back(m);
  return (int)s->mat.size() - 1;
}

int rtb_mat_lambertian(rtb_scene* s, int tex) {
  CHECK_SCENE(s);
  if (!s->tex_ok(tex)) return fail(RTB_ERR_INVALID, "lambertian: bad texture handle");
  return push_mat(s, RTB_MAT_LAMBERTIAN, tex, 0.0);
}
int rtb_mat_metal(rtb_scene* s, int tex, double fuzz) {
  CHECK_SCENE(s);
  if (!s->tex_ok(tex)) return fail(RTB_ERR_INVALID, "metal: bad texture handle");
  return push_mat(s, RTB_MAT_METAL, tex, fuzz);
}
int rtb_mat_dielectric(rtb_scene* s, double ref_idx) {
  CHECK_SCENE(s);
  return push_mat(s, RTB_MAT_DIELECTRIC, -1, ref_idx);
}
int rtb_mat_diffuse_light(rtb_scene* s, int tex) {
  CHECK_SCENE(s);
  if (!s->tex_ok(tex)) return fail(RTB_ERR_INVALID, "diffuse_light: bad texture handle");
  return push_mat(s, RTB_MAT_DIFFUSE_LIGHT, tex, 0.0);
}
int rtb_mat_isotropic(rtb_scene* s, int tex) {
  CHECK_SCENE(s);
  if (!s->tex_ok(tex)) return fail(RTB_ERR_INVALID, "isotropic: bad texture handle");
  return push_mat(s, RTB_MAT_ISOTROPIC, tex, 0.0);
}

static RtbObjRec blank_obj(int type) {
  RtbObjRec o{};
  o.type = type;
  o.mat = o.child = o.tex = -1;
  o.sel = 0;
  o.list_off = o.list_n = 0;
  return o;
}
static int push_obj(rtb_scene* s, const RtbObjRec& o) {
  s->obj.push_back(o);
  return (int)s->obj.size() - 1;
}

int rtb_obj_sphere(rtb_scene* s, const double c[3], double radius, int mat) {
  CHECK_SCENE(s);
  if (!c || !s->mat_ok(mat)) return fail(RTB_ERR_INVALID, "sphere: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_SPHERE);
  o.mat = mat;
  o.v[0] = c[0], o.v[1] = c[1], o.v[2] = c[2], o.v[3] = radius;
  return push_obj(s, o);
}

int rtb_obj_moving_sphere(rtb_scene* s, const double c0[3], const double c1[3], double t0, double t1,
                          double radius, int mat) {
  CHECK_SCENE(s);
  if (!c0 || !c1 || !s->mat_ok(mat)) return fail(RTB_ERR_INVALID, "moving_sphere: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_MOVING_SPHERE);
  o.mat = mat;
  for (int i = 0; i < 3; ++i) o.v[i] = c0[i], o.v[3 + i] = c1[i];
  o.v[6] = t0, o.v[7] = t1, o.v[8] = radius;
  return push_obj(s, o);
}

int rtb_obj_rect(rtb_scene* s, int plane, double x0, double y0, double x1, double y1, double k,
                 int mat) {
  CHECK_SCENE(s);
  if (plane < 0 || plane > 2 || !s->mat_ok(mat)) return fail(RTB_ERR_INVALID, "rect: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_RECT);
  o.mat = mat;
  o.sel = plane;
  o.v[0] = x0, o.v[1] = y0, o.v[2] = x1, o.v[3] = y1, o.v[4] = k;
  return push_obj(s, o);
}

int rtb_obj_cube(rtb_scene* s, const double p_min[3], const double p_max[3], int mat) {
  CHECK_SCENE(s);
  if (!p_min || !p_max || !s->mat_ok(mat)) return fail(RTB_ERR_INVALID, "cube: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_CUBE);
  o.mat = mat;
  for (int i = 0; i < 3; ++i) o.v[i] = p_min[i], o.v[3 + i] = p_max[i];
  return push_obj(s, o);
}

int rtb_obj_flip_normals(rtb_scene* s, int obj) {
  CHECK_SCENE(s);
  if (!s->obj_ok(obj)) return fail(RTB_ERR_INVALID, "flip_normals: bad object handle");
  RtbObjRec o = blank_obj(RTB_OBJ_FLIP_NORMALS);
  o.child = obj;
  return push_obj(s, o);
}

int rtb_obj_rotate(rtb_scene* s, int axis, int obj, double angle_deg) {
  CHECK_SCENE(s);
  if (axis < 0 || axis > 2 || !s->obj_ok(obj)) return fail(RTB_ERR_INVALID, "rotate: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_ROTATE);
  o.child = obj;
  o.sel = axis;
  o.v[0] = angle_deg;
  return push_obj(s, o);
}

int rtb_obj_translate(rtb_scene* s, int obj, const double off[3]) {
  CHECK_SCENE(s);
  if (!off || !s->obj_ok(obj)) return fail(RTB_ERR_INVALID, "translate: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_TRANSLATE);
  o.child = obj;
  o.v[0] = off[0], o.v[1] = off[1], o.v[2] = off[2];
  return push_obj(s, o);
}

int rtb_obj_constant_medium(rtb_scene* s, int boundary, double density, int tex) {
  CHECK_SCENE(s);
  if (!s->obj_ok(boundary) || !s->tex_ok(tex))
    return fail(RTB_ERR_INVALID, "constant_medium: bad argument");
  RtbObjRec o = blank_obj(RTB_OBJ_CONSTANT_MEDIUM);
  o.child = boundary;
  o.tex = tex;
  o.v[0] = density;
  return push_obj(s, o);
}

static int push_list(rtb_scene* s, int type, const int* objs, int n) {
  if (n < 0 || (n > 0 && !objs)) return fail(RTB_ERR_INVALID, "list: bad argument");
  for (int i = 0; i < n; ++i)
    if (!s->obj_ok(objs[i])) return fail(RTB_ERR_INVALID, "list: bad object handle");
  RtbObjRec o = blank_obj(type);
  o.list_off = (int)s->list_pool.size();
  o.list_n = n;
  s->list_pool.insert(s->list_pool.end(), objs, objs + n);
  return push_obj(s, o);
}

int rtb_obj_list(rtb_scene* s, const int* objs, int n) {
  CHECK_SCENE(s);
  return push_list(s, RTB_OBJ_LIST, objs, n);
}

int rtb_obj_bvh(rtb_scene* s, const int* objs, int n, double t0, double t1, uint64_t seed) {
  CHECK_SCENE(s);
  // BVHNode::new on an empty slice indexes hittable[0] -> panic (src/bvh.rs:43-51).
  if (n <= 0) return fail(RTB_ERR_NO_BBOX, "bvh: empty object slice");
  int h = push_list(s, RTB_OBJ_BVH, objs, n);
  if (h < 0) return h;
  s->obj[h].v[0] = t0;
  s->obj[h].v[1] = t1;
  s->obj[h].seed = seed;
  return h;
}

// ---- serialisation ----------------------------------------------------------------------
size_t rtb_scene_serialized_size(const rtb_scene* s) {
  if (!s) return 0;
  return sizeof(RtbSceneHeader) + s->tex.size() * sizeof(RtbTexRec) + s->mat.size() * sizeof(RtbMatRec) +
         s->obj.size() * sizeof(RtbObjRec) + s->list_pool.size() * sizeof(int32_t) +
         s->perlin.size() * sizeof(RtbPerlinTable) + s->image_pool.size();
}

int rtb_scene_serialize(const rtb_scene* s, void* buf, size_t cap) {
  CHECK_SCENE(s);
  size_t need = rtb_scene_serialized_size(s);
  if (!buf || cap < need) return fail(RTB_ERR_INVALID, "serialize: buffer too small");
  RtbSceneHeader h{};
  h.magic = RTB_SCENE_MAGIC, h.version = RTB_SCENE_VERSION;
  h.n_tex = (uint32_t)s->tex.size(), h.n_mat = (uint32_t)s->mat.size();
  h.n_obj = (uint32_t)s->obj.size(), h.n_list = (uint32_t)s->list_pool.size();
  h.n_perlin = (uint32_t)s->perlin.size();
  h.n_image_bytes = s->image_pool.size();
  uint8_t* p = (uint8_t*)buf;
  auto put = [&p](const void* src, size_t n) {
    if (n) std::memcpy(p, src, n);
    p += n;
  };
  put(&h, sizeof h);
  put(s->tex.data(), s->tex.size() * sizeof(RtbTexRec));
  put(s->mat.data(), s->mat.size() * sizeof(RtbMatRec));
  put(s->obj.data(), s->obj.size() * sizeof(RtbObjRec));
  put(s->list_pool.data(), s->list_pool.size() * sizeof(int32_t));
  put(s->perlin.data(), s->perlin.size() * sizeof(RtbPerlinTable));
  put(s->image_pool.data(), s->image_pool.size());
  return RTB_OK;
}

rtb_scene* rtb_scene_deserialize(const void* buf, size_t len) {
  if (!buf || len < sizeof(RtbSceneHeader)) {
    rtb::set_error("deserialize: short buffer");
    return nullptr;
  }
  RtbSceneHeader h;
  std::memcpy(&h, buf, sizeof h);
  if (h.magic != RTB_SCENE_MAGIC || h.version != RTB_SCENE_VERSION) {
    rtb::set_error("deserialize: bad magic/version");
    return nullptr;
  }
  // every count is bounded by the bytes that are left before it is multiplied or added: a header cannot make the sum wrap
  size_t left = len - sizeof h;
  auto take = [&left](uint64_t count, size_t each) {
    if (count > left / each) return false;
    left -= (size_t)count * each;
    return true;
  };
  if (!take(h.n_tex, sizeof(RtbTexRec)) || !take(h.n_mat, sizeof(RtbMatRec)) || !take(h.n_obj, sizeof(RtbObjRec)) ||
      !take(h.n_list, sizeof(int32_t)) || !take(h.n_perlin, sizeof(RtbPerlinTable)) || !take(h.n_image_bytes, 1)) {
    rtb::set_error("deserialize: truncated");
    return nullptr;
  }
  rtb_scene* s = nullptr;
  try {
    s = new rtb_scene();
    s->tex.resize(h.n_tex), s->mat.resize(h.n_mat), s->obj.resize(h.n_obj);
    s->list_pool.resize(h.n_list), s->perlin.resize(h.n_perlin), s->image_pool.resize((size_t)h.n_image_bytes);
  } catch (const std::exception&) {  // nothing may unwind through the extern "C" boundary
    delete s;
    rtb::set_error("deserialize: out of memory");
    return nullptr;
  }
  const uint8_t* p = (const uint8_t*)buf + sizeof h;
  auto get = [&p](void* dst, size_t n) {
    if (n) std::memcpy(dst, p, n);
    p += n;
  };
  get(s->tex.data(), s->tex.size() * sizeof(RtbTexRec));
  get(s->mat.data(), s->mat.size() * sizeof(RtbMatRec));
  get(s->obj.data(), s->obj.size() * sizeof(RtbObjRec));
  get(s->list_pool.data(), s->list_pool.size() * sizeof(int32_t));
  get(s->perlin.data(), s->perlin.size() * sizeof(RtbPerlinTable));
  get(s->image_pool.data(), s->image_pool.size());
  // validate handles so later walks cannot index out of range
  bool ok = true;
  for (size_t i = 0; i < s->tex.size(); ++i) {
    const auto& t = s->tex[i];
    switch (t.type) {
      case RTB_TEX_SOLID: break;
      case RTB_TEX_CHECKER:  // children precede parents, as the builders create them: no cycles (a self-referencing checker
        // would recurse without end in flatten's tex_needs_uv and spin in the device's texture_value loop)
        ok &= t.a >= 0 && (size_t)t.a < i && t.b >= 0 && (size_t)t.b < i;
        break;
      case RTB_TEX_NOISE: ok &= t.a >= 0 && (size_t)t.a < s->perlin.size(); break;
      case RTB_TEX_IMAGE: {
        const uint64_t pool = s->image_pool.size();
        ok &= t.a > 0 && t.b > 0 && t.image_off <= pool && (uint64_t)t.a * (uint64_t)t.b <= (pool - std::min<uint64_t>(t.image_off, pool)) / 3;
        break;
      }
      default: ok = false;  // an unknown type would fall through to the image branch on the device
    }
  }
  for (auto& m : s->mat) {
    ok &= m.type >= RTB_MAT_LAMBERTIAN && m.type <= RTB_MAT_ISOTROPIC;
    if (m.type != RTB_MAT_DIELECTRIC) ok &= s->tex_ok(m.tex);
  }
  for (size_t i = 0; i < s->obj.size(); ++i) {
    auto& o = s->obj[i];
    switch (o.type) {
      case RTB_OBJ_RECT:
        ok &= o.sel >= RTB_PLANE_YZ && o.sel <= RTB_PLANE_XY;
        ok &= s->mat_ok(o.mat);
        break;
      case RTB_OBJ_SPHERE: case RTB_OBJ_MOVING_SPHERE: case RTB_OBJ_CUBE:
        ok &= s->mat_ok(o.mat);
        break;
      case RTB_OBJ_ROTATE:
        ok &= o.sel >= RTB_AXIS_X && o.sel <= RTB_AXIS_Z;
        ok &= o.child >= 0 && (size_t)o.child < i;
        break;
      case RTB_OBJ_FLIP_NORMALS: case RTB_OBJ_TRANSLATE:
        ok &= o.child >= 0 && (size_t)o.child < i;  // children precede parents: no cycles
        break;
      case RTB_OBJ_CONSTANT_MEDIUM:
        ok &= o.child >= 0 && (size_t)o.child < i && s->tex_ok(o.tex);
        break;
      case RTB_OBJ_LIST: case RTB_OBJ_BVH:
        ok &= o.list_off >= 0 && o.list_n >= 0 && (size_t)o.list_off + o.list_n <= s->list_pool.size();
        if (ok)
          for (int k = 0; k < o.list_n; ++k) {
            int c = s->list_pool[o.list_off + k];
            ok &= c >= 0 && (size_t)c < i;
          }
        break;
      default: ok = false;
    }
  }
  if (!ok) {
    delete s;
    rtb::set_error("deserialize: invalid handle in scene");
    return nullptr;
  }
  return s;
}

static int count_prims(const rtb_scene* s, int h) {
  const RtbObjRec& o = s->obj[h];
  switch (o.type) {
    case RTB_OBJ_SPHERE: case RTB_OBJ_MOVING_SPHERE: case RTB_OBJ_RECT: return 1;
    case RTB_OBJ_CUBE: return 6;
    case RTB_OBJ_FLIP_NORMALS: case RTB_OBJ_ROTATE: case RTB_OBJ_TRANSLATE:
    case RTB_OBJ_CONSTANT_MEDIUM:
      return count_prims(s, o.child);
    default: {
      int n = 0;
      for (int k = 0; k < o.list_n; ++k) n += count_prims(s, s->list_pool[o.list_off + k]);
      return n;
    }
  }
}

int rtb_scene_prim_count(rtb_scene* s, int world) {
  CHECK_SCENE(s);
  if (!s->obj_ok(world)) return fail(RTB_ERR_INVALID, "prim_count: bad world handle");
  return count_prims(s, world);
}

size_t rtb_ppm_p3(const uint8_t* rgb8, int nx, int ny, char* out, size_t cap) {
  // tests/test.rs:59 header, :79 one "r g b\n" line per pixel, rows top to bottom.
  std::string s = "P3\n" + std::to_string(nx) + " " + std::to_string(ny) + "\n255\n";
  s.reserve(s.size() + (size_t)nx * ny * 12);
  char line[16];
  static const char* digits = "0123456789";
  for (size_t i = 0; i < (size_t)nx * ny; ++i) {
    int n = 0;
    for (int c = 0; c < 3; ++c) {
      unsigned v = rgb8[3 * i + c];
      if (v >= 100) line[n++] = digits[v / 100];
      if (v >= 10) line[n++] = digits[(v / 10) % 10];
      line[n++] = digits[v % 10];
      line[n++] = c == 2 ? '\n' : ' ';
    }
    s.append(line, n);
  }
  if (out && cap) {
    size_t n = s.size() < cap ? s.size() : cap;
    std::memcpy(out, s.data(), n);
    if (n < cap) out[n] = 0;
  }
  return s.size();
}

int rtb_scene_set_bvh_builder(rtb_scene* s, int builder) {
  CHECK_SCENE(s);
  if (builder < RTB_BVH_AUTO || builder > RTB_BVH_DEVICE_LBVH) return fail(RTB_ERR_INVALID, "set_bvh_builder: bad builder");
  if (s->bvh_builder != builder && s->dev) {  // rebuild at the next upload
    rtb::destroy_device_scene(s->dev);
    s->dev = nullptr;
  }
  s->bvh_builder = builder;
  return RTB_OK;
}

size_t rtb_ppm_p6(const uint8_t* rgb8, int nx, int ny, char* out, size_t cap) {
  // binary sibling of the P3 text above: same header fields, raw bytes, rows top to bottom
  std::string hdr = "P6\n" + std::to_string(nx) + " " + std::to_string(ny) + "\n255\n";
  const size_t body = (size_t)3 * nx * ny, total = hdr.size() + body;
  if (out && cap) {
    size_t n = hdr.size() < cap ? hdr.size() : cap;
    std::memcpy(out, hdr.data(), n);
    if (cap > hdr.size()) std::memcpy(out + hdr.size(), rgb8, std::min(body, cap - hdr.size()));
  }
  return total;
}

int rtb_image_write_ppm(const char* path, const uint8_t* rgb8, int nx, int ny, int binary) {
  // what the #[test] wrappers do with create_image's String (tests/test.rs:536,557: File::create + write!)
  if (!path || !rgb8 || nx <= 0 || ny <= 0) return fail(RTB_ERR_INVALID, "image_write_ppm: bad argument");
  const size_t n = binary ? rtb_ppm_p6(rgb8, nx, ny, nullptr, 0) : rtb_ppm_p3(rgb8, nx, ny, nullptr, 0);
  std::vector<char> buf(n + 1);
  if (binary) rtb_ppm_p6(rgb8, nx, ny, buf.data(), buf.size()); else rtb_ppm_p3(rgb8, nx, ny, buf.data(), buf.size());
  FILE* f = std::fopen(path, "wb");
  if (!f) return fail(RTB_ERR_INVALID, std::string("image_write_ppm: cannot create ") + path);
  const size_t w = std::fwrite(buf.data(), 1, n, f);
  if (std::fclose(f) != 0 || w != n) return fail(RTB_ERR_INVALID, std::string("image_write_ppm: short write to ") + path);
  return RTB_OK;
}

int rtb_scene_save(const rtb_scene* s, const char* path) {
  CHECK_SCENE(s);
  if (!path) return fail(RTB_ERR_INVALID, "scene_save: null path");
  std::vector<uint8_t> buf(rtb_scene_serialized_size(s));
  int rc = rtb_scene_serialize(s, buf.data(), buf.size());
  if (rc != RTB_OK) return rc;
  FILE* f = std::fopen(path, "wb");
  if (!f) return fail(RTB_ERR_INVALID, std::string("scene_save: cannot create ") + path);
  const size_t w = std::fwrite(buf.data(), 1, buf.size(), f);
  if (std::fclose(f) != 0 || w != buf.size()) return fail(RTB_ERR_INVALID, std::string("scene_save: short write to ") + path);
  return RTB_OK;
}

rtb_scene* rtb_scene_load(const char* path) {
  FILE* f = path ? std::fopen(path, "rb") : nullptr;
  if (!f) {
    rtb::set_error(std::string("scene_load: cannot open ") + (path ? path : "(null)"));
    return nullptr;
  }
  std::vector<uint8_t> buf;
  uint8_t chunk[1 << 16];
  size_t n;
  while ((n = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + n);
  std::fclose(f);
  return rtb_scene_deserialize(buf.data(), buf.size());
}

}  // extern "C"
