// test_scenes.cpp — the reference's #[test] bodies (tests/test.rs:527-838) written against the C++ host API
// (include/rtb200.hpp): build the world with the crate's constructors, set_camera, create_image(ny, nx, ns, ..),
// write the P3 text.  Each scene prints "<name> <bytes of P3 text> <FNV-1a of the text>" so the pytest driver
// (tests/test_gpu_parity.py) can compare with the image the C ABI's named scene renders.
//   usage: test_scenes <out_dir> <nx> <ny> <ns> [n_gpus]
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>

#include "rtb200.hpp"

using namespace rtb200;

static uint64_t fnv1a(const std::string& s) {
  uint64_t h = 1469598103934665603ull;
  for (unsigned char c : s) h = (h ^ c) * 1099511628211ull;
  return h;
}

static void emit(const std::string& dir, const char* name, const std::string& ppm) {
  std::ofstream(dir + "/" + name + ".ppm", std::ios::binary) << ppm;  // fs::write(format!("output/{}.ppm", ..), tests/test.rs:561
  std::printf("%s %zu %016llx\n", name, ppm.size(), (unsigned long long)fnv1a(ppm));
}

int main(int argc, char** argv) {
  if (argc < 5) return 2;
  const std::string dir = argv[1];
  const size_t nx = std::atoi(argv[2]), ny = std::atoi(argv[3]), ns = std::atoi(argv[4]);
  const int n_gpus = argc > 5 ? std::atoi(argv[5]) : 1;
  rtb_render_opts o;
  rtb_render_opts_default(&o);
  o.seed = 7;
  const Vector3 up(0.0, 1.0, 0.0);
  try {
    {  // test_two_spheres, tests/test.rs:575-611 (scene :165-182), with the sky the reference left commented out
      CheckerTexture checker(SolidTexture(0.2, 0.3, 0.1), SolidTexture(0.9, 0.9, 0.9));
      HittableList world;
      world.push(Sphere(Vector3(0.0, -10.0, 0.0), 10.0, Lambertian(checker)));
      world.push(Sphere(Vector3(0.0, 10.0, 0.0), 10.0, Lambertian(checker)));
      Camera cam = set_camera(nx, ny, Vector3(13, 2, 3), Vector3(0, 0, 0), up, 20.0, 10.0, 0.1, 0.0, 1.0);
      rtb_render_opts sky = o;
      sky.background = RTB_BG_SKY;
      emit(dir, "two_spheres", create_image(ny, nx, ns, cam, world, &sky, n_gpus));
    }
    {  // test_cornell_box, tests/test.rs:735-768 (scene :242-323): walls as written, a light, two rotated boxes
      Lambertian red(SolidTexture(0.65, 0.05, 0.05)), white(SolidTexture(0.73, 0.73, 0.73)), green(SolidTexture(0.12, 0.45, 0.15));
      DiffuseLight light(SolidTexture(15.0, 15.0, 15.0));
      HittableList world;
      world.push(FlipNormals(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 555.0, green)));
      world.push(Rect(Plane::YZ, 0.0, 0.0, 555.0, 555.0, 0.0, red));
      world.push(Rect(Plane::ZX, 227.0, 213.0, 332.0, 343.0, 554.0, light));
      world.push(FlipNormals(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white)));  // as written: a second floor, no ceiling
      world.push(Rect(Plane::ZX, 0.0, 0.0, 555.0, 555.0, 0.0, white));
      world.push(FlipNormals(Rect(Plane::XY, 0.0, 0.0, 555.0, 555.0, 555.0, white)));
      world.push(Traslate(Rotate(Axis::Y, Cube(Vector3(0.0, 0.0, 0.0), Vector3(165.0, 165.0, 165.0), white), -18.0),
                          Vector3(130.0, 0.0, 65.0)));
      world.push(Traslate(Rotate(Axis::Y, Cube(Vector3(0.0, 0.0, 0.0), Vector3(165.0, 330.0, 165.0), white), 15.0),
                          Vector3(265.0, 0.0, 295.0)));
      Camera cam = set_camera(nx, ny, Vector3(278, 278, -800), Vector3(278, 278, 0), up, 40.0, 10.0, 0.1, 0.0, 1.0);
      emit(dir, "cornell_box", create_image(ny, nx, ns, cam, world, &o, n_gpus));
    }
    if (argc > 6) {  // test_earth, tests/test.rs:613-649 (scene :200-209): a bare Sphere with a JPEG ImageTexture, sky on
      RgbImage image = open_image(argv[6]);  // image::open("texture/earthmap.jpg").expect(..).to_rgb8()
      const uint32_t inx = image.nx, iny = image.ny;
      ImageTexture texture(std::move(image).into_raw(), inx, iny);
      Sphere earth(Vector3(0.0, 0.0, 0.0), 2.0, Lambertian(texture));
      Camera cam = set_camera(nx, ny, Vector3(13, 2, 3), Vector3(0, 0, 0), up, 20.0, 10.0, 0.0, 0.0, 1.0);
      rtb_render_opts sky = o;
      sky.background = RTB_BG_SKY;
      emit(dir, "earth", create_image(ny, nx, ns, cam, earth, &sky, n_gpus));
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  return 0;
}
