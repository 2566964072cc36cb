// scenes_named.cpp — the eight scene constructors of the reference's driver
// (tests/test.rs:89-523) and the cameras its #[test]s use (tests/test.rs:543-830), written
// against the crate-shaped C++ API (include/rtb200.hpp).  thread_rng() is the seeded stand-in,
// drawn in the same order as upstream.  Two labelled lit variants are added for non-vacuous
// image parity; they are NOT reference scenes.
#include <cstring>

#include "rtb200.hpp"
#include "scene.hpp"

using namespace rtb200;

namespace {

struct EarthImage {
  const uint8_t* rgb8;
  uint32_t nx, ny;
};

CheckerTexture make_checker() {  // tests/test.rs:93-96, :166-169
  return CheckerTexture(SolidTexture(0.2, 0.3, 0.1), SolidTexture(0.9, 0.9, 0.9));
}

HittablePtr random_scene() {  // tests/test.rs:89-163
  auto& rng = thread_rng();
  Vector3 origin(4.0, 0.2, 0.0);
  std::vector<HittablePtr> world;
  world.push_back(boxed(Sphere(Vector3(0.0, -1000.0, 0.0), 1000.0, Lambertian(make_checker()))));
  for (int a = -10; a < 10; ++a) {
    for (int b = -10; b < 10; ++b) {
      double choose_material = rng.gen_f64();
      double cx = a + 0.9 * rng.gen_f64();
      double cz = b + 0.9 * rng.gen_f64();
      Vector3 center(cx, 0.2, cz);
      if ((center - origin).magnitude() > 0.9) {
        if (choose_material < 0.8) {  // diffuse, moving up by 0.5*xi over the shutter
          double dy = 0.5 * rng.gen_f64();
          double r = rng.gen_f64() * rng.gen_f64();
          double g = rng.gen_f64() * rng.gen_f64();
          double bl = rng.gen_f64() * rng.gen_f64();
          world.push_back(boxed(MovingSphere(center, center + Vector3(0.0, dy, 0.0), 0.0, 1.0, 0.2,
                                             Lambertian(SolidTexture(r, g, bl)))));
        } else if (choose_material < 0.95) {  // metal
          double r = 0.5 * (1.0 + rng.gen_f64());
          double g = 0.5 * (1.0 + rng.gen_f64());
          double bl = 0.5 * (1.0 + rng.gen_f64());
          double fuzz = 0.5 * rng.gen_f64();
          world.push_back(boxed(Sphere(center, 0.2, Metal(SolidTexture(r, g, bl), fuzz))));
        } else {  // glass
          world.push_back(boxed(Sphere(center, 0.2, Dielectric(1.5))));
        }
      }
    }
  }
  world.push_back(boxed(Sphere(Vector3(0.0, 1.0, 0.0), 1.0, Dielectric(1.5))));
  world.push_back(boxed(Sphere(Vector3(-4.0, 1.0, 0.0), 1.0, Lambertian(SolidTexture(0.4, 0.2, 0.1)))));
  world.push_back(boxed(Sphere(Vector3(4.0, 1.0, 0.0), 1.0, Metal(SolidTexture(0.7, 0.6, 0.5), 0.0))));
  return boxed(BVHNode(world, 0.0, 1.0));
}

HittablePtr two_spheres() {  // tests/test.rs:165-182
  CheckerTexture checker = make_checker();
  HittableList world;
  world.push(Sphere(Vector3(0.0, -10.0, 0.0), 10.0, Lambertian(checker)));
  world.push(Sphere(Vector3(0.0, 10.0, 0.0), 10.0, Lambertian(checker)));
  return boxed(std::move(world));
}

HittablePtr two_perlin_sphere() {  // tests/test.rs:184-198
  NoiseTexture noise(4.0);
  HittableList world;
  world.push(Sphere(Vector3(0.0, -1000.0, 0.0), 1000.0, Lambertian(noise)));
  world.push(Sphere(Vector3(0.0, 2.0, 0.0), 2.0, Lambertian(noise)));
  return boxed(std::move(world));
}

ImageTexture earth_texture(const EarthImage& e) {  // tests/test.rs:201-206, :489-494
  if (!e.rgb8) throw std::runtime_error("image not found");  // .expect("image not found")
  return ImageTexture(std::vector<uint8_t>(e.rgb8, e.rgb8 + (size_t)e.nx * e.ny * 3), e.nx, e.ny);
}

HittablePtr earth(const EarthImage& e) {  // tests/test.rs:200-209 — a bare sphere, not a list
  return boxed(Sphere(Vector3(0.0, 0.0, 0.0), 2.0, Lambertian(earth_texture(e))));
}

HittablePtr simple_light() {  // tests/test.rs:211-240
  NoiseTexture noise(4.0);
  HittableList world;
  world.push(Sphere(Vector3(0.0, -1000.0, 0.0), 1000.0, Lambertian(noise)));
  world.push(Sphere(Vector3(0.0, 2.0, 0.0), 2.0, Lambertian(noise)));
  world.push(Sphere(Vector3(0.0, 7.0, 0.0), 2.0, DiffuseLight(SolidTexture(4.0, 4.0, 4.0))));
  world.push(Rect(Plane::XY, 3.0, 1.0, 5.0, 3.0, -2.0, DiffuseLight(SolidTexture(4.0, 4.0, 4.0))));
  return boxed(std::move(world));
}

HittablePtr cornell_box() {  // tests/test.rs:242-323 — two floors at y=0, no ceiling (Q24)
  Lambertian red(SolidTexture(0.65, 0.05, 0.05));
  Lambertian white(SolidTexture(0.73, 0.73, 0.73));
  Lambertian green(SolidTexture(0.12, 0.45, 0.15));
  DiffuseLight light(SolidTexture(15.0, 15.0, 15.0));
  HittableList world;
  world.push(FlipNormals(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 555.0, green)));
  world.push(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 0.0, red));
  world.push(Rect(Plane::ZX, 227.0, 213.0, 332.0, 343.0, 554.0, light));
  world.push(FlipNormals(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white)));
  world.push(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white));
  world.push(FlipNormals(Rect(Plane::XY, 0.0, 0.0, 555.0, 555.0, 555.0, white)));
  world.push(Traslate(Rotate(Axis::Y, Cube(Vector3(0, 0, 0), Vector3(165, 165, 165), white), -18.0),
                      Vector3(130.0, 0.0, 65.0)));
  world.push(Traslate(Rotate(Axis::Y, Cube(Vector3(0, 0, 0), Vector3(165, 330, 165), white), 15.0),
                      Vector3(265.0, 0.0, 295.0)));
  return boxed(std::move(world));
}

// tests/test.rs:325-417 as written (book == false): floor flipped, ceiling unflipped, the
// sixth wall at z = 0 between camera and box => renders black (Q24).
// book == true is the labelled variant `cornell_smoke_book`: back wall at z = 555, floor
// unflipped at y = 0, ceiling flipped at y = 555.
HittablePtr cornell_smoke(bool book) {
  Lambertian red(SolidTexture(0.65, 0.05, 0.05));
  Lambertian white(SolidTexture(0.73, 0.73, 0.73));
  Lambertian green(SolidTexture(0.12, 0.45, 0.15));
  DiffuseLight light(SolidTexture(7.0, 7.0, 7.0));
  HittableList world;
  world.push(FlipNormals(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 555.0, green)));
  world.push(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 0.0, red));
  world.push(Rect(Plane::ZX, 127.0, 113.0, 432.0, 443.0, 554.0, light));
  if (!book) {
    world.push(FlipNormals(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white)));
    world.push(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 555.0, white));
    world.push(FlipNormals(Rect(Plane::XY, 0.0, 0.0, 555.0, 555.0, 0.0, white)));
  } else {
    world.push(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white));
    world.push(FlipNormals(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 555.0, white)));
    world.push(FlipNormals(Rect(Plane::XY, 0.0, 0.0, 555.0, 555.0, 555.0, white)));
  }
  Traslate box1(Rotate(Axis::Y, Cube(Vector3(0, 0, 0), Vector3(165, 165, 165), white), -18.0),
                Vector3(130.0, 0.0, 65.0));
  Traslate box2(Rotate(Axis::Y, Cube(Vector3(0, 0, 0), Vector3(165, 330, 165), white), 15.0),
                Vector3(265.0, 0.0, 295.0));
  world.push(ConstantMedium(box1, 0.01, SolidTexture(1.0, 1.0, 1.0)));
  world.push(ConstantMedium(box2, 0.01, SolidTexture(0.0, 0.0, 0.0)));
  return boxed(std::move(world));
}

// tests/test.rs:419-523.  lit == false is the scene as written: the light Rect has
// x0 = 147 > x1 = 123 and can never be hit (src/rect.rs:51) => renders black (Q24).
// lit == true is the labelled variant `final_scene_lit` with the bounds ordered.
HittablePtr final_scene(const EarthImage& e, bool lit) {
  auto& rng = thread_rng();
  Lambertian white(SolidTexture(0.73, 0.73, 0.73));
  Lambertian ground(SolidTexture(0.48, 0.83, 0.53));
  HittableList world;
  std::vector<HittablePtr> box_list1;
  const int nb = 20;
  for (int i = 0; i < nb; ++i) {
    for (int j = 0; j < 20; ++j) {
      double w = 100.0;
      double x0 = -1000.0 + i * w, z0 = -1000.0 + j * w, y0 = 0.0;
      double x1 = x0 + w, y1 = 100.0 * (rng.gen_f64() + 0.01), z1 = z0 + w;
      box_list1.push_back(boxed(Cube(Vector3(x0, y0, z0), Vector3(x1, y1, z1), ground)));
    }
  }
  world.push(BVHNode(box_list1, 0.0, 1.0));
  DiffuseLight light(SolidTexture(7.0, 7.0, 7.0));
  if (!lit)
    world.push(Rect(Plane::ZX, 147.0, 412.0, 123.0, 423.0, 554.0, light));
  else
    world.push(Rect(Plane::ZX, 147.0, 123.0, 412.0, 423.0, 554.0, light));
  Vector3 center(400.0, 400.0, 200.0);
  world.push(MovingSphere(center, center + Vector3(30.0, 0.0, 0.0), 0.0, 1.0, 50.0,
                          Lambertian(SolidTexture(0.7, 0.3, 0.1))));
  world.push(Sphere(Vector3(260.0, 150.0, 45.0), 50.0, Dielectric(1.5)));
  world.push(Sphere(Vector3(0.0, 150.0, 145.0), 50.0, Metal(SolidTexture(0.8, 0.8, 0.9), 10.0)));
  Sphere boundary(Vector3(360.0, 150.0, 145.0), 70.0, Dielectric(1.5));
  world.push(boundary);
  world.push(ConstantMedium(boundary, 0.2, SolidTexture(0.2, 0.4, 0.9)));
  Sphere fog(Vector3(0.0, 0.0, 0.0), 5000.0, Dielectric(1.5));
  world.push(ConstantMedium(fog, 0.0001, SolidTexture(1.0, 1.0, 1.0)));
  world.push(Sphere(Vector3(400.0, 200.0, 400.0), 100.0, Lambertian(earth_texture(e))));
  world.push(Sphere(Vector3(220.0, 280.0, 300.0), 80.0, Lambertian(NoiseTexture(0.1))));
  std::vector<HittablePtr> box_list2;
  const int ns = 1000;
  for (int k = 0; k < ns; ++k) {
    double x = 165.0 * rng.gen_f64();
    double y = 165.0 * rng.gen_f64();
    double z = 165.0 * rng.gen_f64();
    box_list2.push_back(boxed(Sphere(Vector3(x, y, z), 10.0, white)));
  }
  world.push(Traslate(Rotate(Axis::Y, BVHNode(box_list2, 0.0, 0.1), 15.0), Vector3(-100.0, 270.0, 395.0)));
  return boxed(std::move(world));
}

// Cameras of the #[test]s, tests/test.rs:543-554 etc.  All: view_up (0,1,0), focus 10,
// aperture 0.1, shutter [0,1].
Camera camera_for(const std::string& name, int nx, int ny) {
  Vector3 up(0.0, 1.0, 0.0);
  if (name == "simple_light")  // :702-713
    return set_camera(nx, ny, Vector3(13, 3, 3), Vector3(0, 0, 0), up, 50.0, 10.0, 0.1, 0.0, 1.0);
  if (name == "cornell_box" || name == "cornell_smoke" || name == "cornell_smoke_book")  // :741-752, :780-791
    return set_camera(nx, ny, Vector3(278, 278, -800), Vector3(278, 278, 0), up, 40.0, 10.0, 0.1, 0.0, 1.0);
  if (name == "final_scene" || name == "final_scene_lit")  // :819-830
    return set_camera(nx, ny, Vector3(478, 278, -600), Vector3(278, 278, 0), up, 40.0, 10.0, 0.1, 0.0, 1.0);
  // random_scene, two_spheres, two_perlin_sphere, earth: :543-554, :582-593, :622-633, :662-673
  return set_camera(nx, ny, Vector3(13, 2, 3), Vector3(0, 0, 0), up, 20.0, 10.0, 0.1, 0.0, 1.0);
}

}  // namespace

extern "C" int rtb_scene_named(rtb_scene* s, const char* name_c, uint64_t seed, const uint8_t* earth_rgb8,
                               uint32_t earth_nx, uint32_t earth_ny, int nx, int ny, RtbCamera* camera) {
  if (!s || !name_c || nx <= 0 || ny <= 0) return rtb::fail(RTB_ERR_INVALID, "scene_named: bad argument");
  std::string name(name_c);
  EarthImage e{earth_rgb8, earth_nx, earth_ny};
  try {
    thread_rng().seed(seed);
    HittablePtr world;
    if (name == "random_scene") world = random_scene();
    else if (name == "two_spheres") world = two_spheres();
    else if (name == "two_perlin_sphere") world = two_perlin_sphere();
    else if (name == "earth") world = earth(e);
    else if (name == "simple_light") world = simple_light();
    else if (name == "cornell_box") world = cornell_box();
    else if (name == "cornell_smoke") world = cornell_smoke(false);
    else if (name == "cornell_smoke_book") world = cornell_smoke(true);
    else if (name == "final_scene") world = final_scene(e, false);
    else if (name == "final_scene_lit") world = final_scene(e, true);
    else return rtb::fail(RTB_ERR_INVALID, "scene_named: unknown scene '" + name + "'");
    Lowering L{s, {}};
    int h = world->lower(L);
    if (camera) *camera = camera_for(name, nx, ny).c;
    return h;
  } catch (const std::exception& ex) {
    return rtb::fail(RTB_ERR_IO, std::string("scene_named: ") + ex.what());
  }
}
