// multi.cu — rtb_render_multi: create_image on several GPUs of one node from ONE host call.
//
// The reference's create_image (tests/test.rs:55-85) is one thread and one accumulator per pixel.
// bench.py and raytracing_rust_b200/distributed.py scale it the way the rest of this repo expects
// (one process per GPU, samples sharded by rank, one exchange step); a Rust caller of the crate has
// a single process, so this entry point does the same thing inside it:
//   * one host thread per device renders its shard into that device's accumulation buffer: by default the sample
//     slice  s = i, i + G, i + 2G, ...  of every pixel (counter-based draws: the union of the slices is exactly the
//     single-GPU sample set: same samples whatever G is, sums equal up to float addition order, i.e. <= 1 LSB after
//     quantisation); with opts->multi_shard = RTB_SHARD_BANDS every sample of the 4-row bands i, i + G, ... instead
//     (the "image tiles" split: disjoint pixels, nothing to add up),
//   * the exchange is sharded over the devices (peer.cu): device g sums + quantises 1/G of the pixels from all G
//     buffers through peer access (NVLink loads) and stores its rgb8 slice into the first device's image; band shards
//     just quantise their own rows into it.  Devices without full peer connectivity fall back to staged copies and one
//     kernel on the first device.  The image goes to the host once.
// Scene replicas for devices 1..G-1 are made from the caller's scene by serialisation and cached in
// the scene handle until it changes.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "render_internal.hpp"
#include "scene.hpp"

using namespace rtb;

namespace {

struct Shard {
  int device = 0;
  rtb_scene* scene = nullptr;
  float* d_sum = nullptr;
  rtb_stats stats{};
  int rc = RTB_OK;
  std::string err;
};

void run_shard(Shard* sh, int world, const RtbCamera* cam, int nx, int ny, int ns, rtb_render_opts o) {
  auto bail = [&](int rc) {
    sh->rc = rc;
    sh->err = rtb_last_error();  // thread-local in the worker: hand it to the caller's thread
  };
  o.device = sh->device;
  int rc = ensure_uploaded(sh->scene, world, sh->device, std::min(cam->time0, cam->time1), std::max(cam->time0, cam->time1));
  if (rc != RTB_OK) return bail(rc);
  const size_t bytes = (size_t)3 * nx * ny * sizeof(float);
  cudaError_t e = cudaSetDevice(sh->device);
  if (e == cudaSuccess) e = pool_alloc(sh->device, bytes, (void**)&sh->d_sum);  // kept between calls: no 100 MB cudaMalloc / cudaFree per image
  if (e == cudaSuccess) e = cudaMemsetAsync(sh->d_sum, 0, bytes, 0);
  if (e != cudaSuccess) {
    fail(RTB_ERR_CUDA, std::string("render_multi: ") + cudaGetErrorString(e));
    return bail(RTB_ERR_CUDA);
  }
  rc = rtb_render_device(sh->scene, cam, nx, ny, ns, &o, sh->d_sum, nullptr, &sh->stats);
  if (rc != RTB_OK) return bail(rc);
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    fail(RTB_ERR_CUDA, std::string("render_multi: ") + cudaGetErrorString(e));
    return bail(RTB_ERR_CUDA);
  }
}

}  // namespace

extern "C" int rtb_render_multi(rtb_scene* s, int world, const RtbCamera* camera, int nx, int ny, int ns,
                                const rtb_render_opts* opts, const int* devices, int n_devices, float* out_sum_rgb,
                                uint8_t* out_rgb8, rtb_stats* stats) {
  if (!s || !camera || nx <= 0 || ny <= 0 || ns <= 0) return fail(RTB_ERR_INVALID, "render_multi: bad argument");
  rtb_render_opts o;
  if (opts) o = *opts; else rtb_render_opts_default(&o);
  if (o.sample_begin != 0 || o.sample_stride > 1)
    return fail(RTB_ERR_INVALID, "render_multi: the call shards the samples itself (sample_begin/stride must be 0/1)");
  const int have = rtb_device_count();
  if (have <= 0) return fail(RTB_ERR_NO_DEVICE, "render_multi: no CUDA device");
  std::vector<int> devs;
  if (devices && n_devices > 0) devs.assign(devices, devices + n_devices);
  else for (int i = 0; i < (n_devices > 0 ? std::min(n_devices, have) : have); ++i) devs.push_back(i);
  if ((int)devs.size() > RTB_MAX_PEER_BUFFERS) return fail(RTB_ERR_INVALID, "render_multi: more than RTB_MAX_PEER_BUFFERS devices");
  for (size_t i = 0; i < devs.size(); ++i) {
    if (devs[i] < 0 || devs[i] >= have) return fail(RTB_ERR_INVALID, "render_multi: bad device ordinal");
    for (size_t j = 0; j < i; ++j)
      if (devs[j] == devs[i]) return fail(RTB_ERR_INVALID, "render_multi: device listed twice");
  }
  const bool bands = o.multi_shard == RTB_SHARD_BANDS;
  if (o.multi_shard != RTB_SHARD_SAMPLES && !bands) return fail(RTB_ERR_INVALID, "render_multi: bad multi_shard");
  if (o.band_stride > 1 || o.band_begin != 0) return fail(RTB_ERR_INVALID, "render_multi: the call shards the rows itself (band_begin/stride must be 0)");
  // never more shards than samples (or than 4-row bands)
  const int G = (int)std::min<size_t>(devs.size(), bands ? (size_t)((ny + RTB_BAND_ROWS - 1) / RTB_BAND_ROWS) : (size_t)ns);
  // "the current world" (world < 0) is a property of the caller's handle: resolve it once, the replicas have none
  if (world < 0) world = s->dev_world;
  if (!s->obj_ok(world)) return fail(RTB_ERR_INVALID, "render_multi: bad world handle");

  // replicas of the scene for devices 1..G-1 (a scene handle holds one device copy), rebuilt when the scene changed
  if (G > 1) {
    std::vector<uint8_t> blob(rtb_scene_serialized_size(s));
    int rc = rtb_scene_serialize(s, blob.data(), blob.size());
    if (rc != RTB_OK) return rc;
    if (blob != s->replica_blob || (int)s->replicas.size() < G - 1) {
      for (rtb_scene* r : s->replicas) delete r;
      s->replicas.clear();
      for (int i = 1; i < G; ++i) {
        rtb_scene* r = rtb_scene_deserialize(blob.data(), blob.size());
        if (!r) {  // leave no half-built replica set behind a stale blob
          for (rtb_scene* q : s->replicas) delete q;
          s->replicas.clear();
          s->replica_blob.clear();
          return RTB_ERR_INVALID;
        }
        r->bvh_builder = s->bvh_builder;
        s->replicas.push_back(r);
      }
      s->replica_blob.swap(blob);
    }
  }

  std::vector<Shard> shards(G);
  for (int i = 0; i < G; ++i) {
    shards[i].device = devs[i];
    shards[i].scene = i == 0 ? s : s->replicas[i - 1];
  }
  {
    std::vector<std::thread> th;
    for (int i = 0; i < G; ++i) {
      rtb_render_opts oi = o;
      if (bands) oi.band_begin = i, oi.band_stride = G;  // every sample of the 4-row bands i, i+G, ...
      else oi.sample_begin = i, oi.sample_stride = G;    // samples i, i+G, ... of every pixel
      th.emplace_back(run_shard, &shards[i], world, camera, nx, ny, ns, oi);
    }
    for (auto& t : th) t.join();
  }
  auto cleanup = [&]() {
    for (Shard& sh : shards)
      if (sh.d_sum) {
        cudaSetDevice(sh.device);
        cudaDeviceSynchronize();
        pool_free(sh.device, sh.d_sum, (size_t)3 * nx * ny * sizeof(float));
      }
  };
  for (Shard& sh : shards)
    if (sh.rc != RTB_OK) {
      cleanup();
      return fail(sh.rc, "render_multi: device " + std::to_string(sh.device) + ": " + sh.err);
    }

  // ---- exchange step ---------------------------------------------------------------------------------------------
  // With peer access between all the devices the exchange is SHARDED: device g sums + quantises 1/G of the pixels,
  // pulling that slice of every peer's partial sums over NVLink, and stores the rgb8 slice (and the f32 total, if
  // asked for) straight into the first device's buffers — G * (G-1)/G slices in flight on G GPUs instead of the
  // first device pulling G-1 whole images.  Row bands need no arithmetic: device g quantises its own bands into the
  // first device's image.  Without full peer access the first device does everything (staged copies).
  const size_t n = (size_t)3 * nx * ny;
  const int root = devs[0];
  std::vector<void*> staging;
  uint8_t* d_rgb8 = nullptr;
  float* d_total = nullptr;
  int rc = RTB_OK;
  std::string err;
  auto cuda_ok = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    rc = RTB_ERR_CUDA, err = std::string("render_multi: ") + what + ": " + cudaGetErrorString(e);
    return false;
  };
  bool all_p2p = G > 1;
  for (int i = 0; i < G && all_p2p; ++i)
    for (int j = 0; j < G && all_p2p; ++j) {
      if (i == j) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devs[i], devs[j]);
      if (!can) all_p2p = false;
    }
  if (all_p2p)
    for (int i = 0; i < G && all_p2p; ++i) {
      if (cudaSetDevice(devs[i]) != cudaSuccess) { all_p2p = false; break; }
      for (int j = 0; j < G; ++j) {
        if (i == j) continue;
        cudaError_t e = cudaDeviceEnablePeerAccess(devs[j], 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError(), e = cudaSuccess;
        if (e != cudaSuccess) { cudaGetLastError(); all_p2p = false; break; }
      }
    }
  std::vector<const float*> bufs(G);
  for (int i = 0; i < G; ++i) bufs[i] = shards[i].d_sum;
  if (cuda_ok(cudaSetDevice(root), "cudaSetDevice")) {
    cuda_ok(pool_alloc(root, n, (void**)&d_rgb8), "cudaMalloc(rgb8)");
    if (rc == RTB_OK && out_sum_rgb && G > 1 && !bands) cuda_ok(pool_alloc(root, n * sizeof(float), (void**)&d_total), "cudaMalloc(total)");
    if (rc == RTB_OK && G == 1) {
      rc = bands ? rtb_resolve_bands_device(bufs[0], nx, ny, ns, 0, 1, d_rgb8, 0, nullptr)
                 : rtb_reduce_resolve_device(bufs.data(), 1, nx, ny, ns, nullptr, d_rgb8, nullptr);
      if (rc != RTB_OK) err = rtb_last_error();
    } else if (rc == RTB_OK && all_p2p) {
      const size_t per = ((n + G - 1) / G + 3) & ~(size_t)3;  // slice length, a multiple of 4 values (128-bit loads)
      for (int g = 0; g < G && rc == RTB_OK; ++g) {
        if (!cuda_ok(cudaSetDevice(devs[g]), "cudaSetDevice")) break;
        if (bands) {
          rc = rtb_resolve_bands_device(bufs[g], nx, ny, ns, g, G, d_rgb8, 0, nullptr);
        } else {
          const size_t first = std::min(n, (size_t)g * per), count = std::min(n - first, per);
          rc = rtb_reduce_resolve_range_device(bufs.data(), G, first, count, ns, d_total, d_rgb8, nullptr);
        }
        if (rc != RTB_OK) err = rtb_last_error();
      }
      for (int g = 0; g < G; ++g)  // every slice must have landed before the first device reads the image
        if (cudaSetDevice(devs[g]) == cudaSuccess && rc == RTB_OK) cuda_ok(cudaDeviceSynchronize(), "cudaDeviceSynchronize(slice)");
      cudaSetDevice(root);
    } else if (rc == RTB_OK) {
      // no full peer connectivity: bring every partial sum to the first device, which does the whole exchange
      for (int i = 1; i < G && rc == RTB_OK; ++i) {
        void* st = nullptr;
        if (!cuda_ok(cudaMalloc(&st, n * sizeof(float)), "cudaMalloc(staging)")) break;
        staging.push_back(st);
        if (!cuda_ok(cudaMemcpyPeer(st, root, shards[i].d_sum, devs[i], n * sizeof(float)), "cudaMemcpyPeer")) break;
        bufs[i] = (const float*)st;
      }
      // (band shards are disjoint and zero elsewhere, so their sum is the image: the same kernel serves both modes)
      if (rc == RTB_OK) {
        rc = rtb_reduce_resolve_device(bufs.data(), G, nx, ny, ns, d_total, d_rgb8, nullptr);
        if (rc != RTB_OK) err = rtb_last_error();
      }
    }
    if (rc == RTB_OK && out_rgb8) cuda_ok(cudaMemcpy(out_rgb8, d_rgb8, n, cudaMemcpyDeviceToHost), "cudaMemcpy(rgb8)");
    if (rc == RTB_OK && out_sum_rgb) {
      if (bands && G > 1) {  // each device's bands straight to their place in the caller's buffer
        const size_t chunk = (size_t)RTB_BAND_ROWS * nx * 3;  // floats per full band
        const int bands_all = (ny + RTB_BAND_ROWS - 1) / RTB_BAND_ROWS;
        for (int g = 0; g < G && rc == RTB_OK; ++g) {
          if (!cuda_ok(cudaSetDevice(devs[g]), "cudaSetDevice")) break;
          const float* src = shards[g].d_sum;
          for (int b = g; b < bands_all && rc == RTB_OK; b += G) {
            const size_t off = (size_t)b * chunk, len = std::min(chunk, n - off);
            cuda_ok(cudaMemcpyAsync(out_sum_rgb + off, src + off, len * sizeof(float), cudaMemcpyDeviceToHost, 0), "cudaMemcpy(sum band)");
          }
          if (rc == RTB_OK) cuda_ok(cudaDeviceSynchronize(), "cudaDeviceSynchronize(sum bands)");
        }
        cudaSetDevice(root);
      } else {
        cuda_ok(cudaMemcpy(out_sum_rgb, G > 1 ? d_total : shards[0].d_sum, n * sizeof(float), cudaMemcpyDeviceToHost), "cudaMemcpy(sum)");
      }
    }
    if (rc == RTB_OK) cuda_ok(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    for (void* p : staging) cudaFree(p);
    if (d_rgb8) pool_free(root, d_rgb8, n);
    if (d_total) pool_free(root, d_total, n * sizeof(float));
  }
  cleanup();
  if (rc != RTB_OK) return fail(rc, err);

  if (stats) {
    std::memset(stats, 0, sizeof *stats);
    for (const Shard& sh : shards) {
      stats->paths += sh.stats.paths, stats->segments += sh.stats.segments;
      stats->node_tests += sh.stats.node_tests, stats->prim_tests += sh.stats.prim_tests, stats->medium_tests += sh.stats.medium_tests;
      for (int k = 0; k < 5; ++k) stats->shade[k] += sh.stats.shade[k];
      for (int k = 0; k < 4; ++k) stats->tex_evals[k] += sh.stats.tex_evals[k];
      stats->kernel_launches += sh.stats.kernel_launches;
      stats->device_ms = std::max(stats->device_ms, sh.stats.device_ms);  // shards run side by side
    }
    stats->kernel_launches += 1;  // the reduce + resolve kernel
    stats->reserved[0] = (uint64_t)G;
  }
  return RTB_OK;
}
