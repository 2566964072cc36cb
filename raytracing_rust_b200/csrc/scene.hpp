// scene.hpp — host-side scene graph behind the rtb_scene handle (internal).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "rtb200.h"

namespace rtb {

// Seeded stand-in for rand::thread_rng() in scene construction (tests/test.rs:90,420,
// src/perlin.rs:5,13): splitmix64; gen::<f64>() = (u64 >> 11) * 2^-53 as in rand 0.8.5.
struct SceneRng {
  uint64_t s;
  explicit SceneRng(uint64_t seed) : s(seed) {}
  uint64_t next_u64() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double gen_f64() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
  uint64_t gen_range(uint64_t n) { return next_u64() % n; }
};

struct DeviceScene;  // flatten.hpp / render.cu
void destroy_device_scene(DeviceScene*);

void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

}  // namespace rtb

struct rtb_scene {
  std::vector<RtbTexRec> tex;
  std::vector<RtbMatRec> mat;
  std::vector<RtbObjRec> obj;
  std::vector<int32_t> list_pool;
  std::vector<RtbPerlinTable> perlin;
  std::vector<uint8_t> image_pool;
  rtb::DeviceScene* dev = nullptr;  // uploaded flattened scene (rtb_scene_upload)
  int dev_world = -1;
  int bvh_builder = 0;  // RtbBvhBuilder
  // rtb_render_multi: copies of this scene for the other devices (one device copy per handle) and the
  // serialised bytes they were made from
  std::vector<rtb_scene*> replicas;
  std::vector<uint8_t> replica_blob;
  ~rtb_scene() {
    for (rtb_scene* r : replicas) delete r;
    if (dev) rtb::destroy_device_scene(dev);
  }
  bool tex_ok(int h) const { return h >= 0 && (size_t)h < tex.size(); }
  bool mat_ok(int h) const { return h >= 0 && (size_t)h < mat.size(); }
  bool obj_ok(int h) const { return h >= 0 && (size_t)h < obj.size(); }
};
