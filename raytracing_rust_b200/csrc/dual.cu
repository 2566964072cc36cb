// dual.cu — k_render_dual: the megakernel's per-warp phase scheduler with TWO paths per lane.
//
// Same hot path as render.cu's k_render_mega (create_image tests/test.rs:55-85, color
// src/color.rs:6-23) and the same device code for every reference formula (trace.cuh).  In the
// megakernel a lane owns one path, so whenever the warp runs a phase (NODE / LEAF / SHADE) the
// lanes whose path is in another phase idle: 15.8 of 32 lanes are active on final_scene (ncu,
// profiles/).  Here a lane owns two path slots.  A phase runs for every lane that has EITHER slot
// in that phase, so a lane only idles when both of its paths wait for something else.
//
// Where the state lives
//   shared memory   everything a path is: ray, 1/d products, closest hit so far, throughput,
//                   pixel / sample / depth — s_slot[word][slot][thread], conflict-free because a
//                   lane only touches its own column.  LEAF and SHADE work straight from there.
//   registers       the scheduling state of both slots (cur, sp — the `active` slot's pair and the
//                   `other` slot's pair, exchanged by a role swap, never indexed) and a cached copy
//                   of one slot's NODE operands (ri, t_best), reloaded only when the lane starts
//                   walking the other slot.  The NODE loop itself touches no shared memory.
//   local memory    the two traversal stacks.
// Work distribution, accumulation tiles, generations and draws are the megakernel's: persistent
// warps draw (8x4 tile x sample slice) items from a device counter; images are bit-identical to
// k_render_mega's up to the order of the float adds into a pixel.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>

#include "render_internal.hpp"
#include "scene.hpp"
#include "trace.cuh"
#include "mega_common.cuh"

namespace rtb {

enum DualWord {
  W_OX = 0, W_OY, W_OZ, W_DX, W_DY, W_DZ, W_TIME, W_INVA,
  W_RI0, W_RI1, W_RI2, W_RI3, W_RI4, W_RI5,
  W_BT, W_BPRIM, W_BID, W_FACES,  // W_FACES: best.face | self_face << 4
  W_SELF,
  W_THRX, W_THRY, W_THRZ, W_PIXEL, W_SAMPLE, W_DEPTH, W_PIXLOCAL,
  W_COUNT
};
constexpr int kDualThreads = 256;
constexpr int kDualStack = RTB_STACK_DEPTH;
constexpr size_t kDualSlotBytes = (size_t)W_COUNT * 2 * kDualThreads * sizeof(float);

__device__ __forceinline__ bool st_leaf(int c) { return c < 0 && c > CUR_DONE; }
__device__ __forceinline__ bool st_shade(int c) { return c == CUR_SHADE || c == CUR_NEW; }
__device__ __forceinline__ bool st_alive(int c) { return c != CUR_DONE && c != CUR_WAIT && c != CUR_NEW; }

template <bool kStats, int kMinBlocks>
__global__ void __launch_bounds__(kDualThreads, kMinBlocks)
    k_render_dual(SceneView sv, DCamera cam, MegaParams P, float* __restrict__ accum, unsigned long long* __restrict__ stats,
                  unsigned int* __restrict__ work_counter) {
  extern __shared__ float s_slot[];  // [W_COUNT][2][256]
  __shared__ float s_acc[8][2][96];
  __shared__ int s_next[8];
  __shared__ int4 s_item[8][2];
  __shared__ int s_gen[8];
  __shared__ int s_pool[8];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int tiles_x = (P.nx + 7) >> 3;
#define SL_F(w, sl) s_slot[((w) * 2 + (sl)) * kDualThreads + threadIdx.x]
#define SL_I(w, sl) reinterpret_cast<int*>(s_slot)[((w) * 2 + (sl)) * kDualThreads + threadIdx.x]

  for (int q = lane; q < 192; q += 32) (&s_acc[wib][0][0])[q] = 0.0f;
  if (lane == 0) {
    s_next[wib] = 0, s_gen[wib] = 0, s_pool[wib] = 0;
    s_item[wib][0] = make_int4(-1, 0, 0, 0), s_item[wib][1] = make_int4(-1, 0, 0, 0);
  }
  SL_I(W_PIXLOCAL, 0) = 0, SL_I(W_PIXLOCAL, 1) = 0;
  __syncwarp();

  Counters cnt = {0, 0, 0};
  TexCounters tc = {{0, 0, 0, 0}};
  unsigned long long n_segments = 0;
  unsigned long long n_shade[5] = {0, 0, 0, 0, 0};

  // scheduling state: (cur, sp) of the active slot `a`, (ocur, osp) of the other one
  int cur = CUR_WAIT, ocur = CUR_WAIT, sp = 0, osp = 0, a = 0;
  int loaded = -1;  // slot whose NODE operands sit in (ri, bt); -1: none
  RayInv ri = {1.f, 1.f, 1.f, 0.f, 0.f, 0.f};
  float bt = CUDART_INF_F;
  int stack[2 * kDualStack];
  bool items_left = true;
#define ROLE_SWAP()                \
  do {                             \
    int t_ = cur;                  \
    cur = ocur, ocur = t_;         \
    t_ = sp, sp = osp, osp = t_;   \
    a ^= 1;                        \
  } while (0)

  for (;;) {
    unsigned m_node = __ballot_sync(0xffffffffu, (cur & ocur) >= 0);
    while (__popc(m_node) >= P.k_node) {
      // ------------------------------------------------------------------ NODE phase
      if (cur < 0 && ocur >= 0) ROLE_SWAP();
      if (cur >= 0) {
        if (loaded != a) {
          ri.idx = SL_F(W_RI0, a), ri.idy = SL_F(W_RI1, a), ri.idz = SL_F(W_RI2, a);
          ri.oodx = SL_F(W_RI3, a), ri.oody = SL_F(W_RI4, a), ri.oodz = SL_F(W_RI5, a);
          bt = SL_F(W_BT, a);
          loaded = a;
        }
        if (kStats) cnt.node_tests++;
        int next, push;
        const int nh = node_step(sv.nodes, cur, ri, P.t_min, bt, &next, &push);
        if (nh == 2) stack[a * kDualStack + sp++] = push;
        if (nh == 0) next = sp > 0 ? stack[a * kDualStack + --sp] : CUR_SHADE;
        cur = next;
      }
      m_node = __ballot_sync(0xffffffffu, (cur & ocur) >= 0);
    }
    const unsigned m_wait = __ballot_sync(0xffffffffu, cur == CUR_WAIT || ocur == CUR_WAIT);
    if (m_wait != 0u) {
      // the current item is used up; the previous item's paths (other tile) must have drained
      // before that tile can be flushed and reused for the next item
      const int g = s_gen[wib];
      const bool old_a = st_alive(cur) && (SL_I(W_PIXLOCAL, a) >> 5) != g;
      const bool old_o = st_alive(ocur) && (SL_I(W_PIXLOCAL, a ^ 1) >> 5) != g;
      const unsigned m_old = __ballot_sync(0xffffffffu, old_a || old_o);
      if (m_old == 0u) {
        const int g2 = g ^ 1;
        flush_tile(s_acc[wib][g2], s_item[wib][g2], P.nx, lane, accum);
        int item = P.n_items;
        if (items_left) {
          if (lane == 0) item = (int)atomicAdd(work_counter, 1u);
          item = __shfl_sync(0xffffffffu, item, 0);
        }
        items_left = item < P.n_items;
        __syncwarp();
        if (lane == 0) {
          int4 info = make_int4(-1, 0, 0, 0);
          if (items_left) {
            const int tile = item / P.slices, slice = item - tile * P.slices;
            const int x0 = (tile % tiles_x) * 8, y0 = (tile / tiles_x) * 4;
            const int w_t = min(8, P.nx - x0), n_pix = w_t * min(4, P.ny - y0);
            const int my_count = slice < P.sample_count ? (P.sample_count - slice + P.slices - 1) / P.slices : 0;
            info = make_int4(x0, y0, w_t | (n_pix << 8), slice);
            s_gen[wib] = g2, s_pool[wib] = n_pix * my_count, s_next[wib] = 0;
          }
          s_item[wib][g2] = info;
        }
        __syncwarp();
        if (cur == CUR_WAIT) cur = items_left ? CUR_NEW : CUR_DONE;
        if (ocur == CUR_WAIT) ocur = items_left ? CUR_NEW : CUR_DONE;
      }
    }
    const unsigned m_leaf = __ballot_sync(0xffffffffu, st_leaf(cur) || st_leaf(ocur));
    const unsigned m_shade = __ballot_sync(0xffffffffu, st_shade(cur) || st_shade(ocur));
    if ((m_node | m_leaf | m_shade) == 0u) break;
    const int n_node = __popc(m_node), n_leaf = __popc(m_leaf), n_shade_l = __popc(m_shade);
    int phase;  // 0 node, 1 leaf, 2 shade
    if (n_leaf >= P.k_leaf) phase = 1;
    else if (n_shade_l >= P.k_shade) phase = 2;
    else phase = (n_node >= n_leaf && n_node >= n_shade_l) ? 0 : (n_leaf >= n_shade_l ? 1 : 2);

    if (phase == 2) {
      // ------------------------------------------------------------------ SHADE phase
      if (!st_shade(cur) && st_shade(ocur)) ROLE_SWAP();
      if (st_shade(cur)) {
        bool need_new = cur == CUR_NEW;
        V3f thr = mk(SL_F(W_THRX, a), SL_F(W_THRY, a), SL_F(W_THRZ, a));
        uint32_t pixel = (uint32_t)SL_I(W_PIXEL, a), sample = (uint32_t)SL_I(W_SAMPLE, a);
        int depth = SL_I(W_DEPTH, a), pix_local = SL_I(W_PIXLOCAL, a);
        RayF ray;
        int self_prim = -1, self_face = 0;
        if (!need_new) {
          ray.o = mk(SL_F(W_OX, a), SL_F(W_OY, a), SL_F(W_OZ, a));
          ray.d = mk(SL_F(W_DX, a), SL_F(W_DY, a), SL_F(W_DZ, a));
          ray.time = SL_F(W_TIME, a);
          Hit best;
          best.t = SL_F(W_BT, a), best.prim = SL_I(W_BPRIM, a), best.id = SL_I(W_BID, a);
          best.face = SL_I(W_FACES, a) & 15;
          n_segments++;
          const Rand5 rnd = rand5(P.keys, (uint32_t)depth, pixel, sample, RTB_RNG_BOUNCE);
          int medium = -1;
          if (sv.n_media > 0)
            medium = sample_media<kStats>(sv, ray, P.t_min, &best.t, rnd.d, rnd.e, P.seed, pixel, sample, (uint32_t)depth,
                                          &cnt);
          bool ended = true;
          V3f rad = mk(0.f, 0.f, 0.f);
          if (medium >= 0 || best.prim >= 0) {
            Surface s;
            if (medium >= 0) {  // src/medium.rs:44-51
              s.p = madd(best.t, ray.d, ray.o);
              s.n = mk(1.f, 0.f, 0.f);
              s.u = 0.f, s.v = 0.f;
              s.mat = __ldg(&sv.media[medium].mat);
              self_prim = -1;
            } else {
              s = finalize_hit(sv, sv.prims, ray, best);
              self_prim = best.prim, self_face = best.face;
            }
            const int4 m = __ldg(reinterpret_cast<const int4*>(sv.mats + s.mat));
            if (kStats) n_shade[m.x]++;
            V3f tex_rgb = mk(1.f, 1.f, 1.f);
            if (m.x != RTB_MAT_DIELECTRIC && (m.x == RTB_MAT_DIFFUSE_LIGHT || depth < P.max_depth))
              tex_rgb = texture_value<kStats>(sv, m.y, s.u, s.v, s.p, &tc);
            if (m.x == RTB_MAT_DIFFUSE_LIGHT) rad = thr * tex_rgb;
            if (depth < P.max_depth) {  // src/color.rs:9
              if (scatter_dir(m.x, __int_as_float(m.z), s, ray, rnd)) {
                thr = thr * tex_rgb;
                ++depth;
                ended = false;
              }
            }
          } else {
            rad = thr * background_color(P.background, ray.d);  // miss, src/color.rs:17-22
          }
          if (rad.x != 0.f || rad.y != 0.f || rad.z != 0.f) {
            float* acc = &s_acc[wib][pix_local >> 5][(pix_local & 31) * 3];
            atomicAdd(acc, rad.x);
            atomicAdd(acc + 1, rad.y);
            atomicAdd(acc + 2, rad.z);
          }
          need_new = ended;
        }
        // lanes whose path ended take the next paths of the warp's current item, in lane order
        const unsigned m_need = __ballot_sync(m_shade, need_new);
        if (need_new) {
          const int base = s_next[wib], pool = s_pool[wib], g = s_gen[wib];
          const int id = base + __popc(m_need & lt_mask);
          __syncwarp(m_need);
          if ((m_need & lt_mask) == 0u) s_next[wib] = min(pool, base + __popc(m_need));
          if (id < pool) {
            const int4 info = s_item[wib][g];
            const int w_t = info.z & 255, n_pix = info.z >> 8;
            int pl, ks, lx, ly;
            if (info.z == (8 | (32 << 8))) {  // full 8x4 tile
              pl = id & 31, ks = id >> 5, lx = pl & 7, ly = pl >> 3;
            } else {
              pl = id % n_pix, ks = id / n_pix, lx = pl % w_t, ly = pl / w_t;
            }
            pix_local = pl | (g << 5);
            const int k = info.w + ks * P.slices;
            sample = (uint32_t)(P.sample_first + k * P.sample_step);
            const int i = info.x + lx, row = info.y + ly;  // row 0 = top of the image = j = ny-1 (tests/test.rs:62)
            pixel = (uint32_t)(row * P.nx + i);
            ray = camera_ray(cam, i, P.ny - 1 - row, P.nx, P.ny, P.keys, pixel, sample);
            thr = mk(1.f, 1.f, 1.f);
            depth = 0, self_prim = -1, self_face = 0;
            SL_I(W_PIXEL, a) = (int)pixel, SL_I(W_SAMPLE, a) = (int)sample, SL_I(W_PIXLOCAL, a) = pix_local;
          } else {
            cur = CUR_WAIT;
          }
        }
        if (cur != CUR_WAIT) {  // start the next segment: world.hit(ray, t_min, MAX), src/color.rs:7
          SL_F(W_THRX, a) = thr.x, SL_F(W_THRY, a) = thr.y, SL_F(W_THRZ, a) = thr.z;
          SL_I(W_DEPTH, a) = depth;
          SL_F(W_OX, a) = ray.o.x, SL_F(W_OY, a) = ray.o.y, SL_F(W_OZ, a) = ray.o.z;
          SL_F(W_DX, a) = ray.d.x, SL_F(W_DY, a) = ray.d.y, SL_F(W_DZ, a) = ray.d.z;
          SL_F(W_TIME, a) = ray.time;
          ri = make_ray_inv(ray);
          SL_F(W_RI0, a) = ri.idx, SL_F(W_RI1, a) = ri.idy, SL_F(W_RI2, a) = ri.idz;
          SL_F(W_RI3, a) = ri.oodx, SL_F(W_RI4, a) = ri.oody, SL_F(W_RI5, a) = ri.oodz;
          SL_F(W_INVA, a) = rcp_fast(dot(ray.d, ray.d));
          bt = CUDART_INF_F;
          loaded = a;
          SL_F(W_BT, a) = bt, SL_I(W_BPRIM, a) = -1, SL_I(W_BID, a) = -1, SL_I(W_FACES, a) = self_face << 4;
          SL_I(W_SELF, a) = self_prim;
          sp = 0;
          cur = sv.root;
        }
      }
    } else if (phase == 0) {
      if (cur < 0 && ocur >= 0) ROLE_SWAP();
      if (cur >= 0) {  // below-threshold NODE step
        if (loaded != a) {
          ri.idx = SL_F(W_RI0, a), ri.idy = SL_F(W_RI1, a), ri.idz = SL_F(W_RI2, a);
          ri.oodx = SL_F(W_RI3, a), ri.oody = SL_F(W_RI4, a), ri.oodz = SL_F(W_RI5, a);
          bt = SL_F(W_BT, a);
          loaded = a;
        }
        if (kStats) cnt.node_tests++;
        int next, push;
        const int nh = node_step(sv.nodes, cur, ri, P.t_min, bt, &next, &push);
        if (nh == 2) stack[a * kDualStack + sp++] = push;
        if (nh == 0) next = sp > 0 ? stack[a * kDualStack + --sp] : CUR_SHADE;
        cur = next;
      }
    } else {
      // ------------------------------------------------------------------ LEAF phase
      if (!st_leaf(cur) && st_leaf(ocur)) ROLE_SWAP();
      if (st_leaf(cur)) {
        const unsigned code = ~(unsigned)cur;
        const int first = (int)(code >> 3), count = (int)(code & 7u);
        RayF ray;
        ray.o = mk(SL_F(W_OX, a), SL_F(W_OY, a), SL_F(W_OZ, a));
        ray.d = mk(SL_F(W_DX, a), SL_F(W_DY, a), SL_F(W_DZ, a));
        ray.time = SL_F(W_TIME, a);
        const float inv_a = SL_F(W_INVA, a);
        float best_t = SL_F(W_BT, a);
        int best_id = SL_I(W_BID, a);
        const int faces = SL_I(W_FACES, a);
        const int self_prim = SL_I(W_SELF, a), self_face = faces >> 4;
        for (int q = 0; q < count; ++q) {
          const int pi = first + q;
          float t;
          int face, id;
          if (intersect_prim<kStats>(sv.prims, pi, sv.xforms, ray, inv_a, P.t_min, best_t, pi == self_prim, self_face, &t,
                                     &face, &id, &cnt) &&
              (t < best_t || id > best_id)) {
            best_t = t, best_id = id;
            SL_F(W_BT, a) = t, SL_I(W_BPRIM, a) = pi, SL_I(W_BID, a) = id, SL_I(W_FACES, a) = face | (self_face << 4);
            if (loaded == a) bt = t;
          }
        }
        cur = sp > 0 ? stack[a * kDualStack + --sp] : CUR_SHADE;
      }
    }
  }

  __syncwarp();
  flush_tile(s_acc[wib][0], s_item[wib][0], P.nx, lane, accum);
  flush_tile(s_acc[wib][1], s_item[wib][1], P.nx, lane, accum);
  unsigned long long v = warp_sum(n_segments);
  if (lane == 0 && v) atomicAdd(stats + 1, v);
  if (kStats) {
    unsigned long long vals[14] = {0,          0,          cnt.node_tests, cnt.prim_tests, cnt.medium_tests,
                                   n_shade[0], n_shade[1], n_shade[2],     n_shade[3],     n_shade[4],
                                   tc.n[0],    tc.n[1],    tc.n[2],        tc.n[3]};
#pragma unroll
    for (int q = 2; q < 14; ++q) {
      unsigned long long w = warp_sum(vals[q]);
      if (lane == 0 && w) atomicAdd(stats + q, w);
    }
  }
#undef SL_F
#undef SL_I
#undef ROLE_SWAP
}

#define DUAL_TRY(expr)                                                                      \
  do {                                                                                      \
    cudaError_t e_ = (expr);                                                                \
    if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

template <bool kStats, int kMinBlocks>
static int launch_dual_variant(DeviceScene* d, const DCamera& cam, const MegaParams& P, long long blocks, int carveout,
                               float* d_sum, cudaStream_t stream) {
  auto kern = k_render_dual<kStats, kMinBlocks>;
  DUAL_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDualSlotBytes));
  DUAL_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carveout));
  DUAL_TRY(cudaMemsetAsync(d->d_work, 0, sizeof(unsigned int), stream));
  kern<<<(unsigned)blocks, kDualThreads, kDualSlotBytes, stream>>>(d->view, cam, P, d_sum, d->d_stats, d->d_work);
  DUAL_TRY(cudaGetLastError());
  return RTB_OK;
}

// opts.reserved[0..2]: lanes a phase needs (defaults 12 / 12 / 12 of the 32 lanes, each holding two paths);
// reserved[3]: resident blocks per SM (2: <= 128 registers, 3: <= 80); reserved[5]: shared-memory carveout override (%)
int launch_dual(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                cudaStream_t stream, uint64_t* launches) {
  const int stride = std::max(1, o.sample_stride);
  const int begin = std::max(0, o.sample_begin);
  const int my_samples = begin < ns ? (ns - begin + stride - 1) / stride : 0;
  const long long tiles = (long long)((nx + 7) / 8) * ((ny + 3) / 4);
  const long long want_warps = (long long)d->prop.multiProcessorCount * 64;
  const int chunk_max = 64;
  for (int done = 0; done < my_samples; done += chunk_max) {
    MegaParams P;
    P.nx = nx, P.ny = ny, P.ns = ns;
    P.sample_count = std::min(chunk_max, my_samples - done);
    P.sample_first = begin + done * stride;
    P.sample_step = stride;
    int slices = (int)std::min<long long>(std::max<long long>(1, want_warps / std::max<long long>(1, tiles)), P.sample_count);
    P.slices = std::max(1, slices);
    P.max_depth = o.max_depth, P.t_min = (float)o.t_min, P.background = o.background, P.seed = o.seed;
    P.keys = philox_keys(o.seed);
    P.k_node = o.reserved[0] > 0 ? o.reserved[0] : 12;
    P.k_leaf = o.reserved[1] > 0 ? o.reserved[1] : 12;
    P.k_shade = o.reserved[2] > 0 ? o.reserved[2] : 12;
    const int min_blocks = o.reserved[3] == 3 ? 3 : 2;
    const long long items = tiles * P.slices;
    if (items > 0x7fffffffLL) return fail(RTB_ERR_INVALID, "render: image too large for one launch");
    P.n_items = (int)items;
    const long long blocks = std::min<long long>((items + 7) / 8, (long long)d->prop.multiProcessorCount * min_blocks);
    const int per_block = (int)kDualSlotBytes + 7 * 1024 + 1024;
    const int carveout = o.reserved[5] > 0 ? o.reserved[5] : std::min(100, (min_blocks * per_block * 100) / (228 * 1024) + 3);
    int rc;
    if (o.collect_stats) rc = launch_dual_variant<true, 2>(d, cam, P, blocks, carveout, d_sum, stream);
    else if (min_blocks == 3) rc = launch_dual_variant<false, 3>(d, cam, P, blocks, carveout, d_sum, stream);
    else rc = launch_dual_variant<false, 2>(d, cam, P, blocks, carveout, d_sum, stream);
    if (rc != RTB_OK) return rc;
    ++*launches;
  }
  return RTB_OK;
}

}  // namespace rtb
