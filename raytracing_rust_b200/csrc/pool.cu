// pool.cu — the SM-resident wavefront: a pool of paths in shared memory, regrouped by phase.
//
// Same hot path as render.cu's megakernel (create_image tests/test.rs:55-85, color
// src/color.rs:6-23) and the same device code for every reference formula (trace.cuh); what
// differs is who runs which path when.  The megakernel binds one path to one lane for the path's
// whole life, so a warp always holds a mix of lanes walking BVH nodes, lanes testing a leaf
// primitive and lanes waiting to shade, and each phase's code runs with the lanes it happens to
// have (ncu: 14 / 8 / 20 of 32).  Here a CTA (one per SM) owns kRows x 32 path slots whose state
// lives in shared memory, slot (row r, column c) = [field][r][c], and every warp is a worker:
//   1. it looks at the three per-column queue masks (bit r of Q[phase][c] = slot (r, c) waits for
//      that phase), picks the phase most of its 32 columns have work for,
//   2. lane c claims one waiting slot of column c (atomicAnd on the mask bit), loads only the
//      fields that phase needs, runs the phase's tight code — NODE: BVH node steps until it
//      reaches a leaf or the end of the traversal; LEAF: the leaf's primitive test; SHADE: media,
//      hit record, emitted, scatter, and regeneration of finished paths from a device counter,
//   3. stores what changed and sets the slot's bit in the queue of its next phase.
// Because a lane only ever touches slots of its own column, every shared-memory access is
// conflict-free, and because ~kRows/16 slots per lane wait at any time, each phase runs with
// nearly full warps.  The traversal stack is per slot in shared memory (kStack levels; scenes whose
// BVH is deeper use the megakernel).  Radiance goes straight to the image with RED.global.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>

#include "render_internal.hpp"
#include "scene.hpp"
#include "trace.cuh"

namespace rtb {

struct PoolParams {
  int nx, ny;
  int sample_first, sample_step, sample_count;  // sample index of path-sample k: sample_first + k * sample_step
  int tiles_x;
  unsigned total_ids;  // tiles * 32 * sample_count path ids; id -> (tile, sample k, pixel of the 8x4 tile)
  int max_depth;
  float t_min;
  int background;
  int k_leaf;  // hybrid worker: a leaf test runs when at least this many lanes hold a leaf
  int k_node;  // a NODE run goes on while at least this many of the warp's lanes are still at inner nodes
  uint64_t seed;
};

// slot fields (float and int planes)
enum { PF_OX = 0, PF_OY, PF_OZ, PF_DX, PF_DY, PF_DZ, PF_TIME, PF_IDX, PF_IDY, PF_IDZ, PF_BEST_T, PF_THR_R, PF_THR_G, PF_THR_B, PF_COUNT };
enum { PI_BEST_PF = 0, PI_BEST_ID, PI_SELF, PI_PIXEL, PI_SAMPLE, PI_DEPTH, PI_CUR, PI_SP, PI_COUNT };

#ifndef RTB_POOL_HYBRID
#define RTB_POOL_HYBRID 1
#endif
#define POOL_COMPILER_FENCE() asm volatile("" ::: "memory")
#define PCUR_SHADE (-2147483647 - 1)
#define PCUR_NEW (-2147483647)

template <int kRows, int kStack>
struct PoolLayout {
  static constexpr size_t kPlane = (size_t)kRows * 32 * 4;
  static constexpr size_t kOffI = PF_COUNT * kPlane;
  static constexpr size_t kOffStack = kOffI + PI_COUNT * kPlane;
  static constexpr size_t kOffQ = kOffStack + kStack * kPlane;
  static constexpr size_t kOffLive = kOffQ + 3 * 32 * 4;
  static constexpr size_t kBytes = kOffLive + 16;
};

__device__ __forceinline__ unsigned long long pool_warp_sum(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <bool kStats, int kRows, int kStack>
__global__ void __launch_bounds__(RTB_POOL_THREADS, 1) k_render_pool(SceneView sv, DCamera cam, PoolParams P, float* __restrict__ accum,
                                                        unsigned long long* __restrict__ stats,
                                                        unsigned int* __restrict__ work_counter) {
  using L = PoolLayout<kRows, kStack>;
  extern __shared__ __align__(16) unsigned char pool_smem[];
  float(*F)[kRows][32] = reinterpret_cast<float(*)[kRows][32]>(pool_smem);
  int(*I)[kRows][32] = reinterpret_cast<int(*)[kRows][32]>(pool_smem + L::kOffI);
  int(*STK)[kRows][32] = reinterpret_cast<int(*)[kRows][32]>(pool_smem + L::kOffStack);
  unsigned(*Q)[32] = reinterpret_cast<unsigned(*)[32]>(pool_smem + L::kOffQ);
  int* s_live = reinterpret_cast<int*>(pool_smem + L::kOffLive);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;

  // every slot starts without a path, queued for SHADE (which regenerates)
  for (int r = wib; r < kRows; r += RTB_POOL_THREADS / 32) I[PI_CUR][r][lane] = PCUR_NEW;
  if (wib == 0) {
    Q[0][lane] = 0u, Q[1][lane] = 0u;
    Q[2][lane] = kRows >= 32 ? 0xffffffffu : ((1u << kRows) - 1u);
    if (lane == 0) *s_live = kRows * 32;
  }
  __syncthreads();

  Counters cnt = {0, 0, 0};
  TexCounters tc = {{0, 0, 0, 0}};
  unsigned long long n_segments = 0;
  unsigned long long n_shade[5] = {0, 0, 0, 0, 0};
  unsigned iter = (unsigned)wib * 7u;

  for (;;) {
    const unsigned mN = reinterpret_cast<volatile unsigned*>(Q[0])[lane];
    const unsigned mL = reinterpret_cast<volatile unsigned*>(Q[1])[lane];
    const unsigned mS = reinterpret_cast<volatile unsigned*>(Q[2])[lane];
    const int nN = __popc(__ballot_sync(0xffffffffu, mN != 0u));
    const int nL = __popc(__ballot_sync(0xffffffffu, mL != 0u));
    const int nS = __popc(__ballot_sync(0xffffffffu, mS != 0u));
    if ((nN | nL | nS) == 0) {
      if (*reinterpret_cast<volatile int*>(s_live) <= 0) break;
      __nanosleep(64);
      continue;
    }
    const int ph = (nN >= nL && nN >= nS) ? 0 : (nL >= nS ? 1 : 2);
    // claim one waiting slot of this lane's column (rotating start so no row is favoured); a lane
    // that loses the race for its bit to another warp sits this run out
    const unsigned m = ph == 0 ? mN : (ph == 1 ? mL : mS);
    const unsigned rot = (iter++) & 31u;
    const bool cand = m != 0u;  // this lane has a candidate slot: row rr of its column
    const unsigned mr = __funnelshift_r(m, m, rot);
    const int rr = cand ? (int)((__ffs((int)mr) - 1 + rot) & 31u) : 0;
    unsigned claimed = 0u;
    if (cand) claimed = atomicAnd(&Q[ph][lane], ~(1u << rr));
    // Shared-memory operations of one thread reach the SM's shared memory in program order, so the
    // state loads below are issued right behind the claim without waiting for its result, and a
    // publisher's state stores land before the atomicOr that sets the slot's bit; only the compiler
    // has to be kept from reordering them.
    POOL_COMPILER_FENCE();
#if RTB_POOL_HYBRID
    if (ph <= 1) {
      // ------------------------------------------------------------------ TRAVERSE worker
      // Claimed from the NODE or the LEAF queue, so it starts with all its lanes in one phase; it
      // then keeps alternating node steps and leaf tests for as long as at least k_node / k_leaf of
      // its lanes want one — a slot only goes back through a queue when the warp's lanes have
      // drifted apart (or its traversal is done).
      RayF ray;
      ray.o = mk(0.f, 0.f, 0.f), ray.d = mk(0.f, 0.f, 1.f), ray.time = 0.f;
      RayInv ri;
      ri.idx = ri.idy = ri.idz = 1.f, ri.oodx = ri.oody = ri.oodz = 0.f;
      float best_t = 0.f;
      int best_pf = -1, best_id = -1, self = -1, cur = PCUR_SHADE, sp = 0;
      if (cand) {
        ray.o = mk(F[PF_OX][rr][lane], F[PF_OY][rr][lane], F[PF_OZ][rr][lane]);
        ray.d = mk(F[PF_DX][rr][lane], F[PF_DY][rr][lane], F[PF_DZ][rr][lane]);
        ray.time = F[PF_TIME][rr][lane];
        ri.idx = F[PF_IDX][rr][lane], ri.idy = F[PF_IDY][rr][lane], ri.idz = F[PF_IDZ][rr][lane];
        ri.oodx = ray.o.x * ri.idx, ri.oody = ray.o.y * ri.idy, ri.oodz = ray.o.z * ri.idz;
        best_t = F[PF_BEST_T][rr][lane];
        best_pf = I[PI_BEST_PF][rr][lane], best_id = I[PI_BEST_ID][rr][lane], self = I[PI_SELF][rr][lane];
        cur = I[PI_CUR][rr][lane], sp = I[PI_SP][rr][lane];
      }
      const bool act = cand && ((claimed >> rr) & 1u);
      if (!act) cur = PCUR_SHADE;
      bool first = true;
      for (;;) {
        const int n_node = __popc(__ballot_sync(0xffffffffu, cur >= 0));
        if (n_node >= P.k_node || (first && ph == 0 && n_node > 0)) {
          first = false;
          if (cur >= 0) {
            if (kStats) cnt.node_tests++;
            int next, push;
            const int nh = node_step(sv.nodes, cur, ri, P.t_min, best_t, &next, &push);
            if (nh == 2) STK[sp++][rr][lane] = push;
            if (nh == 0) next = sp > 0 ? STK[--sp][rr][lane] : PCUR_SHADE;
            cur = next;
          }
          continue;
        }
        const bool at_leaf = cur < 0 && cur != PCUR_SHADE;
        const int n_leaf = __popc(__ballot_sync(0xffffffffu, at_leaf));
        if (n_leaf >= P.k_leaf || (first && n_leaf > 0)) {
          first = false;
          if (at_leaf) {
            const float inv_a = rcp_fast(dot(ray.d, ray.d));
            const unsigned code = ~(unsigned)cur;
            const int lfirst = (int)(code >> 3), count = (int)(code & 7u);
            for (int q = 0; q < count; ++q) {
              const int pi = lfirst + q;
              float t;
              int face, id;
              if (intersect_prim<kStats>(sv.prims, pi, sv.xforms, ray, inv_a, P.t_min, best_t, pi == (self >> 3), self & 7, &t,
                                         &face, &id, &cnt) &&
                  (t < best_t || id > best_id)) {
                best_t = t, best_pf = (pi << 3) | face, best_id = id;
              }
            }
            cur = sp > 0 ? STK[--sp][rr][lane] : PCUR_SHADE;
          }
          continue;
        }
        break;
      }
      if (act) {
        F[PF_BEST_T][rr][lane] = best_t;
        I[PI_BEST_PF][rr][lane] = best_pf, I[PI_BEST_ID][rr][lane] = best_id;
        I[PI_CUR][rr][lane] = cur, I[PI_SP][rr][lane] = sp;
        POOL_COMPILER_FENCE();
        atomicOr(&Q[cur >= 0 ? 0 : (cur == PCUR_SHADE ? 2 : 1)][lane], 1u << rr);
      }
#else
    if (ph == 0) {
      // ------------------------------------------------------------------ NODE worker
      RayInv ri;
      ri.idx = ri.idy = ri.idz = 1.f, ri.oodx = ri.oody = ri.oodz = 0.f;
      float best_t = 0.f;
      int cur = PCUR_SHADE, sp = 0;
      if (cand) {
        ri.idx = F[PF_IDX][rr][lane], ri.idy = F[PF_IDY][rr][lane], ri.idz = F[PF_IDZ][rr][lane];
        ri.oodx = F[PF_OX][rr][lane] * ri.idx, ri.oody = F[PF_OY][rr][lane] * ri.idy, ri.oodz = F[PF_OZ][rr][lane] * ri.idz;
        best_t = F[PF_BEST_T][rr][lane];
        cur = I[PI_CUR][rr][lane], sp = I[PI_SP][rr][lane];
      }
      const bool act = cand && ((claimed >> rr) & 1u);
      if (!act) cur = PCUR_SHADE;
      do {
        if (cur >= 0) {
          if (kStats) cnt.node_tests++;
          int next, push;
          const int nh = node_step(sv.nodes, cur, ri, P.t_min, best_t, &next, &push);
          if (nh == 2) STK[sp++][rr][lane] = push;
          if (nh == 0) next = sp > 0 ? STK[--sp][rr][lane] : PCUR_SHADE;
          cur = next;
        }
      } while (__popc(__ballot_sync(0xffffffffu, cur >= 0)) >= P.k_node);
      if (act) {
        I[PI_CUR][rr][lane] = cur, I[PI_SP][rr][lane] = sp;
        POOL_COMPILER_FENCE();
        atomicOr(&Q[cur >= 0 ? 0 : (cur == PCUR_SHADE ? 2 : 1)][lane], 1u << rr);
      }
    } else if (ph == 1) {
      // ------------------------------------------------------------------ LEAF worker
      const bool act = cand && ((claimed >> rr) & 1u);
      if (act) {
        RayF ray;
        ray.o = mk(F[PF_OX][rr][lane], F[PF_OY][rr][lane], F[PF_OZ][rr][lane]);
        ray.d = mk(F[PF_DX][rr][lane], F[PF_DY][rr][lane], F[PF_DZ][rr][lane]);
        ray.time = F[PF_TIME][rr][lane];
        float best_t = F[PF_BEST_T][rr][lane];
        int best_pf = I[PI_BEST_PF][rr][lane], best_id = I[PI_BEST_ID][rr][lane];
        const int self = I[PI_SELF][rr][lane];
        const int self_prim = self >> 3, self_face = self & 7;  // self = -1 -> prim -1
        int cur = I[PI_CUR][rr][lane], sp = I[PI_SP][rr][lane];
        const float inv_a = rcp_fast(dot(ray.d, ray.d));
        const unsigned code = ~(unsigned)cur;
        const int first = (int)(code >> 3), count = (int)(code & 7u);
        for (int q = 0; q < count; ++q) {
          const int pi = first + q;
          float t;
          int face, id;
          if (intersect_prim<kStats>(sv.prims, pi, sv.xforms, ray, inv_a, P.t_min, best_t, pi == self_prim, self_face, &t, &face,
                                     &id, &cnt) &&
              (t < best_t || id > best_id)) {
            best_t = t, best_pf = (pi << 3) | face, best_id = id;
          }
        }
        cur = sp > 0 ? STK[--sp][rr][lane] : PCUR_SHADE;
        F[PF_BEST_T][rr][lane] = best_t;
        I[PI_BEST_PF][rr][lane] = best_pf, I[PI_BEST_ID][rr][lane] = best_id;
        I[PI_CUR][rr][lane] = cur, I[PI_SP][rr][lane] = sp;
        POOL_COMPILER_FENCE();
        atomicOr(&Q[cur >= 0 ? 0 : (cur == PCUR_SHADE ? 2 : 1)][lane], 1u << rr);
      }
#endif
    } else {
      // ------------------------------------------------------------------ SHADE worker
      const bool act = cand && ((claimed >> rr) & 1u);
      RayF ray;
      ray.o = mk(0.f, 0.f, 0.f), ray.d = mk(0.f, 0.f, 1.f), ray.time = 0.f;
      V3f thr = mk(1.f, 1.f, 1.f);
      uint32_t pixel = 0, sample = 0;
      int depth = 0, self = -1;
      bool need_new = false;
      if (act) {
        need_new = I[PI_CUR][rr][lane] == PCUR_NEW;
        if (!need_new) {
          ray.o = mk(F[PF_OX][rr][lane], F[PF_OY][rr][lane], F[PF_OZ][rr][lane]);
          ray.d = mk(F[PF_DX][rr][lane], F[PF_DY][rr][lane], F[PF_DZ][rr][lane]);
          ray.time = F[PF_TIME][rr][lane];
          thr = mk(F[PF_THR_R][rr][lane], F[PF_THR_G][rr][lane], F[PF_THR_B][rr][lane]);
          pixel = (uint32_t)I[PI_PIXEL][rr][lane], sample = (uint32_t)I[PI_SAMPLE][rr][lane];
          depth = I[PI_DEPTH][rr][lane];
          Hit best;
          best.t = F[PF_BEST_T][rr][lane];
          const int pf = I[PI_BEST_PF][rr][lane];
          best.prim = pf >> 3, best.face = pf & 7, best.id = 0;
          // the segment's closest surface hit is in `best`; media may pre-empt it
          n_segments++;
          const Rand5 rnd = rand5(P.seed, (uint32_t)depth, pixel, sample, RTB_RNG_BOUNCE);
          int medium = -1;
          if (sv.n_media > 0)
            medium = sample_media<kStats>(sv, ray, P.t_min, &best.t, rnd.d, rnd.e, P.seed, pixel, sample, (uint32_t)depth, &cnt);
          bool ended = true;
          V3f rad = mk(0.f, 0.f, 0.f);
          if (medium >= 0 || best.prim >= 0) {
            Surface s;
            if (medium >= 0) {  // src/medium.rs:44-51
              s.p = madd(best.t, ray.d, ray.o);
              s.n = mk(1.f, 0.f, 0.f);
              s.u = 0.f, s.v = 0.f;
              s.mat = __ldg(&sv.media[medium].mat);
              self = -1;
            } else {
              s = finalize_hit(sv, sv.prims, ray, best);
              self = pf;
            }
            const int4 m4 = __ldg(reinterpret_cast<const int4*>(sv.mats + s.mat));
            if (kStats) n_shade[m4.x]++;
            V3f tex_rgb = mk(1.f, 1.f, 1.f);
            if (m4.x != RTB_MAT_DIELECTRIC && (m4.x == RTB_MAT_DIFFUSE_LIGHT || depth < P.max_depth))
              tex_rgb = texture_value<kStats>(sv, m4.y, s.u, s.v, s.p, &tc);
            if (m4.x == RTB_MAT_DIFFUSE_LIGHT) rad = thr * tex_rgb;  // emitted, src/color.rs:8
            if (depth < P.max_depth && scatter_dir(m4.x, __int_as_float(m4.z), s, ray, rnd)) {  // src/color.rs:9-12
              thr = thr * tex_rgb;
              ++depth;
              ended = false;
            }
          } else {
            rad = thr * background_color(P.background, ray.d);  // miss, src/color.rs:17-22
          }
          if (rad.x != 0.f || rad.y != 0.f || rad.z != 0.f) {
            float* a = accum + 3ull * pixel;
            atomicAdd(a, rad.x), atomicAdd(a + 1, rad.y), atomicAdd(a + 2, rad.z);
          }
          need_new = ended;
        }
      }
      // finished (or empty) slots take the next path ids: one device atomic per warp
      const unsigned m_need = __ballot_sync(0xffffffffu, need_new);
      bool dead = false;
      if (m_need != 0u) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(work_counter, (unsigned)__popc(m_need));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (need_new) {
          const unsigned id = base + (unsigned)__popc(m_need & lt_mask);
          if (id < P.total_ids && base <= id) {
            const unsigned per_tile = 32u * (unsigned)P.sample_count;
            const unsigned tile = id / per_tile, rem = id - tile * per_tile;
            const int k = (int)(rem >> 5), p = (int)(rem & 31u);
            const int i = (int)(tile % (unsigned)P.tiles_x) * 8 + (p & 7), row = (int)(tile / (unsigned)P.tiles_x) * 4 + (p >> 3);
            if (i < P.nx && row < P.ny) {
              sample = (uint32_t)(P.sample_first + k * P.sample_step);
              pixel = (uint32_t)(row * P.nx + i);  // row 0 = top of the image = j = ny-1 (tests/test.rs:62)
              ray = camera_ray(cam, i, P.ny - 1 - row, P.nx, P.ny, P.seed, pixel, sample);
              thr = mk(1.f, 1.f, 1.f);
              depth = 0, self = -1;
              need_new = false;
            }
          } else {
            dead = true;  // no path ids left: the slot retires
          }
        }
      }
      if (act) {
        if (dead) {
          atomicSub(s_live, 1);
        } else if (need_new) {
          // an id outside a ragged edge tile: stay queued for SHADE and draw again
          I[PI_CUR][rr][lane] = PCUR_NEW;
          POOL_COMPILER_FENCE();
          atomicOr(&Q[2][lane], 1u << rr);
        } else {
          // start the next segment: world.hit(ray, t_min, MAX), src/color.rs:7
          const RayInv ri = make_ray_inv(ray);
          F[PF_OX][rr][lane] = ray.o.x, F[PF_OY][rr][lane] = ray.o.y, F[PF_OZ][rr][lane] = ray.o.z;
          F[PF_DX][rr][lane] = ray.d.x, F[PF_DY][rr][lane] = ray.d.y, F[PF_DZ][rr][lane] = ray.d.z;
          F[PF_TIME][rr][lane] = ray.time;
          F[PF_IDX][rr][lane] = ri.idx, F[PF_IDY][rr][lane] = ri.idy, F[PF_IDZ][rr][lane] = ri.idz;
          F[PF_BEST_T][rr][lane] = CUDART_INF_F;
          F[PF_THR_R][rr][lane] = thr.x, F[PF_THR_G][rr][lane] = thr.y, F[PF_THR_B][rr][lane] = thr.z;
          I[PI_BEST_PF][rr][lane] = -1, I[PI_BEST_ID][rr][lane] = -1, I[PI_SELF][rr][lane] = self;
          I[PI_PIXEL][rr][lane] = (int)pixel, I[PI_SAMPLE][rr][lane] = (int)sample, I[PI_DEPTH][rr][lane] = depth;
          I[PI_CUR][rr][lane] = sv.root, I[PI_SP][rr][lane] = 0;
          POOL_COMPILER_FENCE();
          atomicOr(&Q[0][lane], 1u << rr);
        }
      }
    }
  }

  unsigned long long v = pool_warp_sum(n_segments);
  if (lane == 0 && v) atomicAdd(stats + 1, v);
  if (kStats) {
    unsigned long long vals[14] = {0,          0,          cnt.node_tests, cnt.prim_tests, cnt.medium_tests,
                                   n_shade[0], n_shade[1], n_shade[2],     n_shade[3],     n_shade[4],
                                   tc.n[0],    tc.n[1],    tc.n[2],        tc.n[3]};
#pragma unroll
    for (int q = 2; q < 14; ++q) {
      unsigned long long w = pool_warp_sum(vals[q]);
      if (lane == 0 && w) atomicAdd(stats + q, w);
    }
  }
}

// ================================ host ==========================================================
#define POOL_TRY(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t e_ = (expr);                                                                             \
    if (e_ != cudaSuccess) return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

template <bool kStats, int kRows, int kStack>
static int launch_pool_variant(DeviceScene* d, const DCamera& cam, const PoolParams& P, float* d_sum, cudaStream_t stream) {
  using L = PoolLayout<kRows, kStack>;
  auto kern = k_render_pool<kStats, kRows, kStack>;
  static bool configured = false;
  if (!configured) {
    POOL_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::kBytes));
    configured = true;
  }
  POOL_TRY(cudaMemsetAsync(d->d_work, 0, sizeof(unsigned int), stream));
  kern<<<d->prop.multiProcessorCount, RTB_POOL_THREADS, L::kBytes, stream>>>(d->view, cam, P, d_sum, d->d_stats, d->d_work);
  POOL_TRY(cudaGetLastError());
  return RTB_OK;
}

bool pool_supports(const DeviceScene* d) { return d->flat.bvh_depth < RTB_POOL_STACK; }

int launch_pool(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                cudaStream_t stream, uint64_t* launches) {
  const int stride = std::max(1, o.sample_stride);
  const int begin = std::max(0, o.sample_begin);
  const int my_samples = begin < ns ? (ns - begin + stride - 1) / stride : 0;
  const long long tiles_x = (nx + 7) / 8, tiles = tiles_x * ((ny + 3) / 4);
  // path ids of one launch fit 31 bits
  const int chunk_max = (int)std::max<long long>(1, std::min<long long>(64, (1LL << 31) / (tiles * 32) - 1));
  for (int done = 0; done < my_samples; done += chunk_max) {
    PoolParams P;
    P.nx = nx, P.ny = ny;
    P.sample_count = std::min(chunk_max, my_samples - done);
    P.sample_first = begin + done * stride;
    P.sample_step = stride;
    P.tiles_x = (int)tiles_x;
    P.total_ids = (unsigned)(tiles * 32 * P.sample_count);
    P.max_depth = o.max_depth, P.t_min = (float)o.t_min, P.background = o.background, P.seed = o.seed;
    P.k_node = std::max(1, o.reserved[0] > 0 ? o.reserved[0] : 8);
    P.k_leaf = std::max(1, o.reserved[1] > 0 ? o.reserved[1] : 8);
    int rc;
    if (o.collect_stats)
      rc = launch_pool_variant<true, RTB_POOL_ROWS, RTB_POOL_STACK>(d, cam, P, d_sum, stream);
    else
      rc = launch_pool_variant<false, RTB_POOL_ROWS, RTB_POOL_STACK>(d, cam, P, d_sum, stream);
    if (rc != RTB_OK) return rc;
    ++*launches;
  }
  return RTB_OK;
}

}  // namespace rtb
