// wavefront.cu — the wavefront form of the hot path: ray-gen / extend / material-sorted shade
// queues over a pool of path slots resident in HBM (north_star's pipeline).
//
// Same arithmetic as the megakernel (trace.cuh), different scheduling: instead of a warp
// juggling its 32 paths through phases, every phase is its own kernel over a compacted queue of
// slot indices, so each kernel starts with full warps and each shade kernel runs ONE material:
//   k_wf_extend        world.hit(ray, t_min, MAX) + constant media (src/color.rs:7) for every
//                      slot in the extend queue; appends the slot to the queue of the material
//                      it hit (or the miss queue)
//   k_wf_shade<M>      hit record + emitted + Material::scatter (src/color.rs:8-16) for material
//                      M in {lambertian, metal, dielectric, diffuse_light, isotropic, miss};
//                      survivors go to the next extend queue, finished paths add their
//                      radiance to the accumulation buffer and free their slot
//   k_wf_generate      free slots take the next (pixel, sample) and a camera ray (raygen)
//   k_wf_advance       one thread: rolls the device-side counters between iterations
// Path state is SoA in 16-byte records (4 x 128-bit per slot, 64 B); queues are int32 arrays with
// warp-aggregated atomic appends.  Counter-based RNG (philox.h) means no RNG state is stored.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>

#include "render_internal.hpp"
#include "scene.hpp"
#include "trace.cuh"

namespace rtb {

RTB_DEFINE_CHECK_READER(read_checks_wavefront)


enum { WQ_LAMBERTIAN = 0, WQ_METAL, WQ_DIELECTRIC, WQ_LIGHT, WQ_ISOTROPIC, WQ_MISS, WQ_COUNT };

struct WfCounters {
  int n_ext[2];           // extend queue sizes (ping-pong)
  int n_new;              // free slots waiting for a path
  int n_shade[WQ_COUNT];  // per-material queue sizes
  int pad;
  unsigned long long next_path, total_paths;
  unsigned long long segments;
};

struct WfView {
  float4* ray_o;   // o.xyz, time
  float4* ray_d;   // d.xyz, hit t
  float4* thr;     // throughput.xyz, -
  int4* meta;      // pixel, sample, depth | self_face << 8, self_prim ; (hit prim/face live in `hit`)
  int2* hit;       // prim (or -1 miss, <= -2: medium -2-m), face
  int* q_ext[2];
  int* q_new;
  int* q_shade[WQ_COUNT];
  WfCounters* cnt;
  int capacity;
};

struct WfParams {
  int nx, ny;
  int sample_first, sample_step;  // sample index of path-sample k: sample_first + k * sample_step
  int max_depth;
  float t_min;
  int background;
  uint64_t seed;
};

// append `slot` to a queue with one atomic per warp
__device__ __forceinline__ void queue_push(int* __restrict__ q, int* __restrict__ count, int slot, bool pred, int capacity) {
  const unsigned m = __ballot_sync(0xffffffffu, pred);  // callers keep whole warps converged here
  if (!pred) return;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(count, __popc(m));
  base = __shfl_sync(m, base, leader);
  RTB_CHECK(base >= 0 && base + __popc(m) <= capacity && slot >= 0 && slot < capacity, RTB_CHK_QUEUE);
  (void)capacity;
  q[base + __popc(m & ((1u << lane) - 1u))] = slot;
}

__global__ void __launch_bounds__(256) k_wf_generate(WfView wv, DCamera cam, WfParams P, int parity) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_new = wv.cnt->n_new;
  const unsigned long long id = wv.cnt->next_path + (unsigned long long)i;
  const bool live = i < n_new && id < wv.cnt->total_paths;
  const int slot = i < n_new ? wv.q_new[i] : 0;
  if (live) {
    // path id -> (sample k, pixel in 8x4-tile order) so that neighbouring slots start coherent
    const unsigned npix = (unsigned)P.nx * (unsigned)P.ny;
    const unsigned k = (unsigned)(id / npix), p = (unsigned)(id % npix);
    const unsigned tiles_x = (unsigned)(P.nx + 7) >> 3;
    // tile-major order over full tiles is only a coherence hint; fall back to row-major when ragged
    unsigned px, row;
    if ((P.nx & 7) == 0 && (P.ny & 3) == 0) {
      const unsigned tile = p >> 5, in = p & 31u;
      px = (tile % tiles_x) * 8u + (in & 7u);
      row = (tile / tiles_x) * 4u + (in >> 3);
    } else {
      px = p % (unsigned)P.nx, row = p / (unsigned)P.nx;
    }
    const uint32_t pixel = row * (unsigned)P.nx + px;
    const uint32_t sample = (uint32_t)(P.sample_first + (int)k * P.sample_step);
    RayF ray = camera_ray(cam, (int)px, P.ny - 1 - (int)row, P.nx, P.ny, P.seed, pixel, sample);
    wv.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
    wv.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
    wv.thr[slot] = make_float4(1.f, 1.f, 1.f, 0.f);
    wv.meta[slot] = make_int4((int)pixel, (int)sample, 0, -1);
  }
  queue_push(wv.q_ext[parity], &wv.cnt->n_ext[parity], slot, live, wv.capacity);
}

// rolls the counters after an iteration: `consumed` extend queue emptied, shade queues emptied,
// the paths handed out by k_wf_generate accounted for
__global__ void k_wf_advance(WfView wv, int consumed_parity) {
  WfCounters* c = wv.cnt;
  c->next_path = min(c->total_paths, c->next_path + (unsigned long long)c->n_new);
  c->n_new = 0;
  c->n_ext[consumed_parity] = 0;
  for (int q = 0; q < WQ_COUNT; ++q) c->n_shade[q] = 0;
}

__global__ void __launch_bounds__(256) k_wf_extend(SceneView sv, WfView wv, WfParams P, int parity) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = wv.cnt->n_ext[parity];
  const bool on = i < n;
  int slot = 0, which = WQ_MISS;
  if (on) {
    slot = wv.q_ext[parity][i];
    const float4 o4 = wv.ray_o[slot], d4 = wv.ray_d[slot];
    const int4 meta = wv.meta[slot];
    RayF ray;
    ray.o = mk(o4.x, o4.y, o4.z), ray.d = mk(d4.x, d4.y, d4.z), ray.time = o4.w;
    const int depth = meta.z & 255, self_face = (meta.z >> 8) & 7;
    Counters cnt;
    Hit h = trace_closest<false>(sv, ray, P.t_min, meta.w, self_face, &cnt);
    int medium = -1;
    if (sv.n_media > 0) {
      const Rand5 rnd = rand5(P.seed, (uint32_t)depth, (uint32_t)meta.x, (uint32_t)meta.y, RTB_RNG_BOUNCE);
      medium = sample_media<false>(sv, ray, P.t_min, &h.t, rnd.d, rnd.e, P.seed, (uint32_t)meta.x, (uint32_t)meta.y,
                                   (uint32_t)depth, &cnt);
    }
    int mat = -1;
    if (medium >= 0) {
      mat = __ldg(&sv.media[medium].mat);
      h.prim = -2 - medium;
    } else if (h.prim >= 0) {
      mat = __ldg(&sv.prims[h.prim].c.y);
    }
    if (mat >= 0) which = __ldg(&sv.mats[mat].type);
    wv.ray_d[slot].w = h.t;
    wv.hit[slot] = make_int2(h.prim, h.face);
  }
  // one vote + one atomic per (warp, queue)
#pragma unroll
  for (int q = 0; q < WQ_COUNT; ++q) queue_push(wv.q_shade[q], &wv.cnt->n_shade[q], slot, on && which == q, wv.capacity);
  const unsigned m_on = __ballot_sync(0xffffffffu, on);
  if (m_on && (threadIdx.x & 31) == 0) atomicAdd(&wv.cnt->segments, (unsigned long long)__popc(m_on));
}

// Material-sorted shading: every slot in queue kQ hit material type kQ (or missed).
template <int kQ>
__global__ void __launch_bounds__(256) k_wf_shade(SceneView sv, WfView wv, WfParams P, int next_parity,
                                                  float* __restrict__ accum) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = wv.cnt->n_shade[kQ];
  const bool on = i < n;
  bool survive = false;
  int slot = 0;
  if (on) {
    slot = wv.q_shade[kQ][i];
    const float4 o4 = wv.ray_o[slot], d4 = wv.ray_d[slot], t4 = wv.thr[slot];
    int4 meta = wv.meta[slot];
    RayF ray;
    ray.o = mk(o4.x, o4.y, o4.z), ray.d = mk(d4.x, d4.y, d4.z), ray.time = o4.w;
    V3f thr = mk(t4.x, t4.y, t4.z);
    V3f rad = mk(0.f, 0.f, 0.f);
    int depth = meta.z & 255;
    if (kQ == WQ_MISS) {
      rad = thr * background_color(P.background, ray.d);  // src/color.rs:17-22
    } else {
      const int2 hp = wv.hit[slot];
      Hit h;
      h.t = d4.w, h.prim = hp.x, h.face = hp.y, h.id = 0;
      Surface s;
      int self_prim = -1, self_face = 0;
      if (h.prim <= -2) {  // constant-medium event, src/medium.rs:44-51
        s.p = madd(h.t, ray.d, ray.o);
        s.n = mk(1.f, 0.f, 0.f);
        s.u = 0.f, s.v = 0.f;
        s.mat = __ldg(&sv.media[-2 - h.prim].mat);
      } else {
        s = finalize_hit(sv, sv.prims, ray, h);
        self_prim = h.prim, self_face = h.face;
      }
      const int4 m = __ldg(reinterpret_cast<const int4*>(sv.mats + s.mat));
      TexCounters tc;
      if (kQ == WQ_LIGHT) rad = thr * texture_value<false>(sv, m.y, s.u, s.v, s.p, &tc);  // emitted, src/color.rs:8
      if (kQ != WQ_LIGHT && depth < P.max_depth) {  // src/color.rs:9-12
        const Rand5 rnd = rand5(P.seed, (uint32_t)depth, (uint32_t)meta.x, (uint32_t)meta.y, RTB_RNG_BOUNCE);
        V3f atten;
        if (scatter<false>(sv, kQ, m.y, __int_as_float(m.z), s, ray, rnd, &atten, &tc)) {
          thr = thr * atten;
          ++depth;
          survive = true;
          wv.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
          wv.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
          wv.thr[slot] = make_float4(thr.x, thr.y, thr.z, 0.f);
          meta.z = depth | (self_face << 8);
          meta.w = self_prim;
          wv.meta[slot] = meta;
        }
      }
    }
    if (rad.x != 0.f || rad.y != 0.f || rad.z != 0.f) {
      float* a = accum + 3ull * (unsigned)meta.x;
      atomicAdd(a, rad.x);
      atomicAdd(a + 1, rad.y);
      atomicAdd(a + 2, rad.z);
    }
  }
  queue_push(wv.q_ext[next_parity], &wv.cnt->n_ext[next_parity], slot, on && survive, wv.capacity);
  queue_push(wv.q_new, &wv.cnt->n_new, slot, on && !survive, wv.capacity);
}

__global__ void k_wf_init(WfView wv, unsigned long long total_paths) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < wv.capacity) wv.q_new[i] = i;
  if (i == 0) {
    WfCounters c = {};
    c.n_new = wv.capacity;
    c.total_paths = total_paths;
    *wv.cnt = c;
  }
}

// ================================ host side =====================================================
struct WavefrontState {
  WfView view{};
  std::vector<void*> allocs;
  int capacity = 0;
};

void destroy_wavefront_state(WavefrontState* w) {
  if (!w) return;
  for (void* p : w->allocs) cudaFree(p);
  delete w;
}

#define WF_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

static int ensure_state(DeviceScene* d, int capacity) {
  if (d->wf && d->wf->capacity >= capacity) return RTB_OK;
  if (d->wf) destroy_wavefront_state(d->wf), d->wf = nullptr;
  WavefrontState* w = new WavefrontState();
  w->capacity = capacity;
  auto alloc = [&](size_t bytes, void** out) -> cudaError_t {
    cudaError_t e = cudaMalloc(out, bytes);
    if (e == cudaSuccess) w->allocs.push_back(*out);
    return e;
  };
  WfView& v = w->view;
  v.capacity = capacity;
  const size_t n = (size_t)capacity;
  d->wf = w;  // owned from here on (freed with the scene even if an allocation below fails)
  WF_TRY(alloc(n * sizeof(float4), (void**)&v.ray_o));
  WF_TRY(alloc(n * sizeof(float4), (void**)&v.ray_d));
  WF_TRY(alloc(n * sizeof(float4), (void**)&v.thr));
  WF_TRY(alloc(n * sizeof(int4), (void**)&v.meta));
  WF_TRY(alloc(n * sizeof(int2), (void**)&v.hit));
  WF_TRY(alloc(n * sizeof(int), (void**)&v.q_ext[0]));
  WF_TRY(alloc(n * sizeof(int), (void**)&v.q_ext[1]));
  WF_TRY(alloc(n * sizeof(int), (void**)&v.q_new));
  for (int q = 0; q < WQ_COUNT; ++q) WF_TRY(alloc(n * sizeof(int), (void**)&v.q_shade[q]));
  WF_TRY(alloc(sizeof(WfCounters), (void**)&v.cnt));
  return RTB_OK;
}

int launch_wavefront(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                     cudaStream_t stream, uint64_t* launches) {
  const int stride = std::max(1, o.sample_stride), begin = std::max(0, o.sample_begin);
  const long long my_samples = begin < ns ? (ns - begin + stride - 1) / stride : 0;
  const unsigned long long total = (unsigned long long)my_samples * (unsigned long long)nx * ny;
  if (total == 0) return RTB_OK;
  // slots in flight: enough to fill the machine many times over, bounded by the job and by memory
  const long long want = o.reserved[4] > 0 ? (long long)o.reserved[4] : (1ll << 20);  // 64 MB of path state: stays in the 126 MB L2 (profiles/r1_wavefront.txt)
  const int capacity = (int)std::max<long long>(256, std::min<long long>(want, (long long)std::min<unsigned long long>(total, 1ull << 30)));
  int rc = ensure_state(d, capacity);
  if (rc != RTB_OK) return rc;
  WfView wv = d->wf->view;
  wv.capacity = capacity;
  WfParams P;
  P.nx = nx, P.ny = ny, P.sample_first = begin, P.sample_step = stride;
  P.max_depth = o.max_depth, P.t_min = (float)o.t_min, P.background = o.background, P.seed = o.seed;
  const int block = 256, grid = (capacity + block - 1) / block;
  k_wf_init<<<grid, block, 0, stream>>>(wv, total);
  ++*launches;
  int parity = 0;  // generate fills q_ext[parity]; extend consumes it; shade fills q_ext[parity ^ 1]
  k_wf_generate<<<grid, block, 0, stream>>>(wv, cam, P, parity);
  k_wf_advance<<<1, 1, 0, stream>>>(wv, parity ^ 1);
  *launches += 2;
  WfCounters h{};
  const int check_every = 4;
  for (long long it = 0;; ++it) {
    k_wf_extend<<<grid, block, 0, stream>>>(d->view, wv, P, parity);
    k_wf_shade<WQ_LAMBERTIAN><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_shade<WQ_METAL><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_shade<WQ_DIELECTRIC><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_shade<WQ_LIGHT><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_shade<WQ_ISOTROPIC><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_shade<WQ_MISS><<<grid, block, 0, stream>>>(d->view, wv, P, parity ^ 1, d_sum);
    k_wf_generate<<<grid, block, 0, stream>>>(wv, cam, P, parity ^ 1);
    k_wf_advance<<<1, 1, 0, stream>>>(wv, parity);
    *launches += 9;
    parity ^= 1;
    if ((it % check_every) == check_every - 1) {
      WF_TRY(cudaMemcpyAsync(&h, wv.cnt, sizeof h, cudaMemcpyDeviceToHost, stream));
      WF_TRY(cudaStreamSynchronize(stream));
      if (h.n_ext[parity] == 0 && h.next_path >= h.total_paths) break;
    }
  }
  WF_TRY(cudaGetLastError());
  // fold the segment count into the scene's stats block (slot 1, like the megakernel)
  WF_TRY(cudaMemcpyAsync(d->d_stats + 1, &wv.cnt->segments, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, stream));
  return RTB_OK;
}

}  // namespace rtb
