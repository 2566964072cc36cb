// debug.cu — field-level parity hooks: the device functions of trace.cuh run one stage at a time over a batch, so the
// tests can compare them number against number with the CPU oracle instead of only through rendered images.
//
//   rtb_closest_hit_ex      world.hit(ray, t_min, MAX) (src/color.rs:7) with the scattering surface named per ray
//                           (the kernels' analytic self-hit exclusion, trace.cuh intersect_prim `self`) and the whole
//                           HitRecord {t,u,v,p,normal} (src/hittable.rs:9-16) of the answer
//   rtb_debug_camera_rays   Camera::get_ray (src/camera.rs:53-67) as the render kernels evaluate it (device f32)
//   rtb_debug_camera_draws  the Philox draws behind those rays in the oracle's input layout (host, f64)
//   rtb_debug_scatter       Material::scatter (src/material.rs:49-168) + the albedo lookup, device f32
//   rtb_debug_sample        random_in_unit_sphere / _disk (src/util.rs:4-24) as sampled on the device
// Test hooks only: nothing here is on the render path.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <mutex>
#include <string>
#include <vector>

#include "render_internal.hpp"
#include "trace.cuh"

namespace rtb {

RTB_DEFINE_CHECK_READER(read_checks_debug)


#define CUDA_TRY(expr)                                                                                              \
  do {                                                                                                              \
    cudaError_t e_ = (expr);                                                                                        \
    if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));           \
  } while (0)

// primitive ID -> index into prims (a BOX owns six consecutive IDs, src/cube.rs:21-74)
__global__ void k_id_table(const DPrim* __restrict__ prims, int n_prims, int* __restrict__ table, int n_ids) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_prims) return;
  const int4 C = ldg4i(&prims[p].c);
  const int n = RTB_PRIM_TYPE(C.x) == DP_BOX ? 6 : 1;
  for (int f = 0; f < n; ++f)
    if (C.z + f < n_ids) table[C.z + f] = p;
}

__global__ void __launch_bounds__(256) k_closest_hit_ex(SceneView sv, const float* __restrict__ rays, size_t n, float t_min,
                                                        const int* __restrict__ self_id, const int* __restrict__ id_table,
                                                        int* __restrict__ out_prim, float* __restrict__ out_rec) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const float* r = rays + 7 * idx;
  RayF ray;
  ray.o = mk(r[0], r[1], r[2]);
  ray.d = mk(r[3], r[4], r[5]);
  ray.time = r[6];
  int self_prim = -1, self_face = 0;
  if (self_id && self_id[idx] >= 0) {
    self_prim = id_table[self_id[idx]];
    self_face = self_id[idx] - ldg4i(&sv.prims[self_prim].c).z;
  }
  Counters cnt;
  const Hit h = trace_closest<false>(sv, ray, t_min, self_prim, self_face, &cnt);
  int id = -1;
  float rec[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (h.prim >= 0) {
    const int4 C = ldg4i(&sv.prims[h.prim].c);
    id = C.z + (RTB_PRIM_TYPE(C.x) == DP_BOX ? h.face : 0);
    const Surface s = finalize_hit(sv, sv.prims, ray, h, nullptr, true);
    rec[0] = h.t, rec[1] = s.u, rec[2] = s.v;
    rec[3] = s.p.x, rec[4] = s.p.y, rec[5] = s.p.z;
    rec[6] = s.n.x, rec[7] = s.n.y, rec[8] = s.n.z;
  }
  out_prim[idx] = id;
  if (out_rec)
    for (int q = 0; q < 9; ++q) out_rec[9 * idx + q] = rec[q];
}

__global__ void k_camera_rays(DCamera cam, int nx, int ny, int sidx, uint64_t seed, float* __restrict__ out) {
  const int pixel = blockIdx.x * blockDim.x + threadIdx.x;
  if (pixel >= nx * ny) return;
  const int row = pixel / nx, i = pixel - row * nx;
  const PhiloxKeys keys = philox_keys(seed);
  const RayF ray = camera_ray(cam, i, ny - 1 - row, nx, ny, keys, (uint32_t)pixel, (uint32_t)sidx);
  float* o = out + 7 * (size_t)pixel;
  o[0] = ray.o.x, o[1] = ray.o.y, o[2] = ray.o.z, o[3] = ray.d.x, o[4] = ray.d.y, o[5] = ray.d.z, o[6] = ray.time;
}

// in: n x {ray o[3], d[3], time, hit t,u,v,p[3],normal[3]} (16 floats); out: n x {scattered?, o[3], d[3], atten[3]}.
// The draws of item i are those of bounce 0 of (pixel i, sample 0): the block the render kernels use.
__global__ void k_scatter(SceneView sv, int mat, const float* __restrict__ in, size_t n, uint64_t seed, float* __restrict__ out) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const float* q = in + 16 * idx;
  RayF ray;
  ray.o = mk(q[0], q[1], q[2]), ray.d = mk(q[3], q[4], q[5]), ray.time = q[6];
  Surface s;
  s.u = q[8], s.v = q[9], s.p = mk(q[10], q[11], q[12]), s.n = mk(q[13], q[14], q[15]), s.mat = mat;
  const int4 m = __ldg(reinterpret_cast<const int4*>(sv.mats + mat));
  const Rand5 rnd = rand5(seed, 0u, (uint32_t)idx, 0u, RTB_RNG_BOUNCE);
  TexCounters tc;
  // the render kernels' order (render.cu SHADE): one texture lookup, then scatter_dir
  V3f tex_rgb = mk(1.f, 1.f, 1.f);
  if (m.x != RTB_MAT_DIELECTRIC) tex_rgb = texture_value<false>(sv, m.y, s.u, s.v, s.p, &tc);
  const bool ok = scatter_dir(m.x, __int_as_float(m.z), s, ray, rnd);
  float* o = out + 10 * idx;
  o[0] = ok ? 1.f : 0.f;
  o[1] = ok ? ray.o.x : 0.f, o[2] = ok ? ray.o.y : 0.f, o[3] = ok ? ray.o.z : 0.f;
  o[4] = ok ? ray.d.x : 0.f, o[5] = ok ? ray.d.y : 0.f, o[6] = ok ? ray.d.z : 0.f;
  o[7] = ok ? tex_rgb.x : 0.f, o[8] = ok ? tex_rgb.y : 0.f, o[9] = ok ? tex_rgb.z : 0.f;
}

// kind 0: the unit-ball sample of the scatter code; kind 1: the unit-disk sample of the lens (camera_ray's r, phi form)
__global__ void k_sample(int kind, uint64_t seed, size_t n, float* __restrict__ out) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
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

namespace {
struct DevBuf {  // freed on every return path
  void* p = nullptr;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, std::max<size_t>(bytes, 16)); }
  template <class T>
  T* as() const {
    return (T*)p;
  }
};
}  // namespace

}  // namespace rtb

using namespace rtb;

extern "C" {

int rtb_closest_hit_ex(rtb_scene* s, int world, const double* rays, size_t n, double t_min, const int32_t* self_id, int device,
                       int32_t* out_prim, float* out_rec) {
  if (!s || (n && (!rays || !out_prim))) return fail(RTB_ERR_INVALID, "closest_hit_ex: bad argument");
  double tlo = 0.0, thi = 1.0;
  for (size_t i = 0; i < n; ++i) tlo = std::min(tlo, rays[7 * i + 6]), thi = std::max(thi, rays[7 * i + 6]);
  int rc = ensure_uploaded(s, world, device, tlo, thi);
  if (rc != RTB_OK) return rc;
  if (n == 0) return RTB_OK;
  const DeviceScene* d = s->dev;
  const int n_ids = d->flat.n_prim_ids;
  if (self_id)
    for (size_t i = 0; i < n; ++i)
      if (self_id[i] >= n_ids) return fail(RTB_ERR_INVALID, "closest_hit_ex: self_id out of range");
  CUDA_TRY(cudaSetDevice(device));
  std::vector<float> h((size_t)7 * n);
  for (size_t i = 0; i < 7 * n; ++i) h[i] = (float)rays[i];
  DevBuf d_rays, d_self, d_table, d_prim, d_rec;
  CUDA_TRY(d_rays.alloc(h.size() * sizeof(float)));
  CUDA_TRY(d_prim.alloc(n * sizeof(int)));
  CUDA_TRY(cudaMemcpy(d_rays.p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  if (out_rec) CUDA_TRY(d_rec.alloc(9 * n * sizeof(float)));
  if (self_id) {
    CUDA_TRY(d_self.alloc(n * sizeof(int)));
    CUDA_TRY(cudaMemcpy(d_self.p, self_id, n * sizeof(int), cudaMemcpyHostToDevice));
    CUDA_TRY(d_table.alloc((size_t)std::max(1, n_ids) * sizeof(int)));
    CUDA_TRY(cudaMemset(d_table.p, 0, (size_t)std::max(1, n_ids) * sizeof(int)));
    if (d->view.n_prims > 0)
      k_id_table<<<(d->view.n_prims + 255) / 256, 256>>>(d->view.prims, d->view.n_prims, d_table.as<int>(), n_ids);
  }
  k_closest_hit_ex<<<(unsigned)((n + 255) / 256), 256>>>(view_with_width(d, 0), d_rays.as<float>(), n, (float)t_min, d_self.as<int>(),
                                                         d_table.as<int>(), d_prim.as<int>(), d_rec.as<float>());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out_prim, d_prim.p, n * sizeof(int), cudaMemcpyDeviceToHost));
  if (out_rec) CUDA_TRY(cudaMemcpy(out_rec, d_rec.p, 9 * n * sizeof(float), cudaMemcpyDeviceToHost));
  return RTB_OK;
}

int rtb_debug_camera_rays(const RtbCamera* c, int nx, int ny, int sidx, uint64_t seed, int device, float* out) {
  if (!c || !out || nx <= 0 || ny <= 0) return fail(RTB_ERR_INVALID, "debug_camera_rays: bad argument");
  CUDA_TRY(cudaSetDevice(device));
  DCamera cam;
  make_device_camera(*c, &cam);
  const size_t n = (size_t)nx * ny;
  DevBuf d_out;
  CUDA_TRY(d_out.alloc(7 * n * sizeof(float)));
  k_camera_rays<<<(unsigned)((n + 255) / 256), 256>>>(cam, nx, ny, sidx, seed, d_out.as<float>());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out, d_out.p, 7 * n * sizeof(float), cudaMemcpyDeviceToHost));
  return RTB_OK;
}

// The draws of sample `sidx` of every pixel in the oracle's explicit-input layout (orc_camera_rays):
// n x {s, t, rd.x, rd.y, xi_time}, rd = the unit-disk point (before the lens-radius scale).  Host code, f64.
int rtb_debug_camera_draws(int nx, int ny, int sidx, uint64_t seed, double* out) {
  if (!out || nx <= 0 || ny <= 0) return fail(RTB_ERR_INVALID, "debug_camera_draws: bad argument");
  for (int row = 0; row < ny; ++row) {
    const int j = ny - 1 - row;
    for (int i = 0; i < nx; ++i) {
      const uint32_t pixel = (uint32_t)(row * nx + i);
      const Rand5 r5 = rand5(seed, 0u, pixel, (uint32_t)sidx, RTB_RNG_CAMERA);
      double* o = out + 5 * (size_t)pixel;
      o[0] = ((double)i + (double)r5.a) / (double)nx;  // tests/test.rs:66-67
      o[1] = ((double)j + (double)r5.b) / (double)ny;
      const double r = std::sqrt((double)r5.c), phi = (2.0 * (double)r5.d - 1.0) * M_PI;
      o[2] = r * std::cos(phi), o[3] = r * std::sin(phi);
      o[4] = (double)r5.e;
    }
  }
  return RTB_OK;
}

int rtb_debug_scatter(rtb_scene* s, int world, int mat, const double* in, size_t n, uint64_t seed, int device, float* out) {
  if (!s || !in || !out || !s->mat_ok(mat)) return fail(RTB_ERR_INVALID, "debug_scatter: bad argument");
  int rc = ensure_uploaded(s, world, device, 0.0, 1.0);
  if (rc != RTB_OK) return rc;
  if (n == 0) return RTB_OK;
  CUDA_TRY(cudaSetDevice(device));
  std::vector<float> h((size_t)16 * n);
  for (size_t i = 0; i < 16 * n; ++i) h[i] = (float)in[i];
  DevBuf d_in, d_out;
  CUDA_TRY(d_in.alloc(h.size() * sizeof(float)));
  CUDA_TRY(d_out.alloc(10 * n * sizeof(float)));
  CUDA_TRY(cudaMemcpy(d_in.p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  k_scatter<<<(unsigned)((n + 255) / 256), 256>>>(s->dev->view, mat, d_in.as<float>(), n, seed, d_out.as<float>());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out, d_out.p, 10 * n * sizeof(float), cudaMemcpyDeviceToHost));
  return RTB_OK;
}

int rtb_debug_sample(int kind, uint64_t seed, size_t n, int device, float* out) {
  if (!out || kind < 0 || kind > 1) return fail(RTB_ERR_INVALID, "debug_sample: bad argument");
  if (n == 0) return RTB_OK;
  CUDA_TRY(cudaSetDevice(device));
  DevBuf d_out;
  CUDA_TRY(d_out.alloc(3 * n * sizeof(float)));
  k_sample<<<(unsigned)((n + 255) / 256), 256>>>(kind, seed, n, d_out.as<float>());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out, d_out.p, 3 * n * sizeof(float), cudaMemcpyDeviceToHost));
  return RTB_OK;
}

}  // extern "C"

// ---- closest-hit query that allocates nothing (rtb_closest_hit_noalloc) ------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  rtb_closest_hit_ex needs one piece of device
// scratch — the table "primitive ID -> index into prims" that turns a ray's self_id into the primitive it leaves — and
// allocates it (plus staging for its host buffers) per call.  Here the caller passes DEVICE inputs and outputs and the
// table lives in a file-scope array every GPU that loads the library holds until exit: RTB_NOALLOC_MAX_PRIM_IDS ints =
// 256 KB.  In host code `g_noalloc_id_table` is only a stand-in for the symbol: its device address is fetched once per
// device with cudaGetSymbolAddress (that device current) and it is that pointer the kernels receive.
namespace rtb {
static __device__ __align__(256) int g_noalloc_id_table[RTB_NOALLOC_MAX_PRIM_IDS];

namespace {
const int kNoallocMaxDevices = 64;
int* g_noalloc_table_ptr[kNoallocMaxDevices];  // device address of the array on each device, nullptr = not fetched yet
std::mutex g_noalloc_mutex;

// once per device, with that device current
cudaError_t noalloc_table_for_current_device(int device, int** out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mutex);
  if (!g_noalloc_table_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_id_table);
    if (e != cudaSuccess) return e;
    g_noalloc_table_ptr[device] = (int*)p;
  }
  *out = g_noalloc_table_ptr[device];
  return cudaSuccess;
}
}  // namespace
}  // namespace rtb

extern "C" int rtb_closest_hit_noalloc(rtb_scene* s, const float* d_rays, size_t n, double t_min, const int32_t* d_self_id,
                                       int32_t* d_out_prim, float* d_out_rec, void* cuda_stream) {
  if (!s || (n && (!d_rays || !d_out_prim))) return fail(RTB_ERR_INVALID, "closest_hit_noalloc: bad argument");
  if (!s->dev)  // uploading allocates: it is the caller's set-up step (rtb_scene_upload), not this call's
    return fail(RTB_ERR_INVALID, "closest_hit_noalloc: scene not uploaded (call rtb_scene_upload first; this entry point allocates nothing)");
  const DeviceScene* d = s->dev;
  const int n_ids = d->flat.n_prim_ids;
  if (n_ids > RTB_NOALLOC_MAX_PRIM_IDS)
    return fail(RTB_ERR_INVALID, "closest_hit_noalloc: the scene has " + std::to_string(n_ids) + " primitive IDs, the static scratch table holds " +
                                     std::to_string(RTB_NOALLOC_MAX_PRIM_IDS) + " (RTB_NOALLOC_MAX_PRIM_IDS); use rtb_closest_hit_ex");
  if (d->device < 0 || d->device >= kNoallocMaxDevices) return fail(RTB_ERR_INVALID, "closest_hit_noalloc: device ordinal out of range");
  if (n == 0) return RTB_OK;
  CUDA_TRY(cudaSetDevice(d->device));
  int* table = nullptr;
  CUDA_TRY(noalloc_table_for_current_device(d->device, &table));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  if (d_self_id) {
    // the array is shared by every call and keeps the last call's contents: start from zeros, then fill this scene's IDs
    CUDA_TRY(cudaMemsetAsync(table, 0, (size_t)std::max(1, n_ids) * sizeof(int), stream));
    if (d->view.n_prims > 0)
      k_id_table<<<(d->view.n_prims + 255) / 256, 256, 0, stream>>>(d->view.prims, d->view.n_prims, table, n_ids);
  }
  k_closest_hit_ex<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(view_with_width(d, 0), d_rays, n, (float)t_min, d_self_id, table,
                                                                     d_out_prim, d_out_rec);
  CUDA_TRY(cudaGetLastError());
  return RTB_OK;
}

