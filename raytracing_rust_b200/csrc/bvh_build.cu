// bvh_build.cu — the BVH built on the GPU.
//
// Replaces BVHNode::new (src/bvh.rs:17-66: random axis, sort by bbox.min[axis], median split,
// recursion on the host) for large scenes.  Closest-hit answers do not depend on the tree, so
// the device builder is free to use a construction that parallelises: a linear BVH over
// Morton-sorted primitive centroids (Karras 2012, "Maximizing parallelism in the construction of
// BVHs, octrees and k-d trees"):
//   k_bounds      scene bounds of the primitive centroids (block reduction + ordered-int atomics)
//   k_morton      30-bit Morton code per primitive
//   cub radix sort (code, primitive index)
//   k_hierarchy   one thread per internal node: its key range, split and children
//   k_refit       one thread per leaf walks up; the second child to arrive at a node unions the
//                 child boxes (atomic arrival counter), so every box is computed exactly once
//   k_emit        writes the traversal format (DNode: both child boxes in the parent, leaf =
//                 ~(primitive << 3 | 1)) and gathers the primitives into leaf order
//   k_depth       tree depth (for the traversal stacks)
// Moving spheres enter with the union of their boxes over the shutter interval
// (src/sphere.rs:165-174), exactly as in the host builder: flatten.cpp computes the padded f32
// primitive boxes for both.
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <string>
#include <vector>

#include "render_internal.hpp"
#include "scene.hpp"

namespace rtb {

namespace {

__device__ __forceinline__ unsigned f2ord(float f) {  // order-preserving float -> uint
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// bounds[0..2] = min of centroids (ordered uint), bounds[3..5] = max
__global__ void __launch_bounds__(256) k_bounds(const float* __restrict__ boxes, int n, unsigned* __restrict__ bounds) {
  __shared__ unsigned s_lo[3][8], s_hi[3][8];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  if (i < n) {
    const float* b = boxes + 6 * (size_t)i;
#pragma unroll
    for (int a = 0; a < 3; ++a) lo[a] = hi[a] = f2ord(0.5f * (b[a] + b[3 + a]));
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = min(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = max(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
    if ((threadIdx.x & 31) == 0) s_lo[a][threadIdx.x >> 5] = lo[a], s_hi[a][threadIdx.x >> 5] = hi[a];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    unsigned l = 0xffffffffu, h = 0u;
    for (int w = 0; w < 8; ++w) l = min(l, s_lo[threadIdx.x][w]), h = max(h, s_hi[threadIdx.x][w]);
    atomicMin(bounds + threadIdx.x, l);
    atomicMax(bounds + 3 + threadIdx.x, h);
  }
}

__device__ __forceinline__ unsigned expand10(unsigned v) {  // 10 bits -> every third bit
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

__global__ void __launch_bounds__(256) k_morton(const float* __restrict__ boxes, int n, const unsigned* __restrict__ bounds,
                                                unsigned* __restrict__ keys, unsigned* __restrict__ idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* b = boxes + 6 * (size_t)i;
  unsigned q[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float lo = ord2f(bounds[a]), hi = ord2f(bounds[3 + a]);
    const float ext = fmaxf(hi - lo, 1e-30f);
    const float t = (0.5f * (b[a] + b[3 + a]) - lo) / ext;
    q[a] = (unsigned)fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f);
  }
  keys[i] = (expand10(q[0]) << 2) | (expand10(q[1]) << 1) | expand10(q[2]);
  idx[i] = (unsigned)i;
}

// length of the common prefix of keys i and j (ties broken by the position, Karras §4)
__device__ __forceinline__ int delta(const unsigned* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const unsigned a = keys[i], b = keys[j];
  return a != b ? __clz(a ^ b) : 32 + __clz((unsigned)i ^ (unsigned)j);
}

// internal node i in [0, n-2]; children: index >= 0 internal, ~leaf for sorted leaf position
__global__ void __launch_bounds__(256) k_hierarchy(const unsigned* __restrict__ keys, int n, int2* __restrict__ children,
                                                   int* __restrict__ parent_int, int* __restrict__ parent_leaf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    if (t == 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  int2 c;
  if (lo == gamma) {
    c.x = ~gamma;
    parent_leaf[gamma] = i;
  } else {
    c.x = gamma;
    parent_int[gamma] = i;
  }
  if (hi == gamma + 1) {
    c.y = ~(gamma + 1);
    parent_leaf[gamma + 1] = i;
  } else {
    c.y = gamma + 1;
    parent_int[gamma + 1] = i;
  }
  children[i] = c;
  if (i == 0) parent_int[0] = -1;
}

struct Box6 {
  float lo[3], hi[3];
};
__device__ __forceinline__ Box6 load_box(const float* __restrict__ p) {
  Box6 b;
  b.lo[0] = p[0], b.lo[1] = p[1], b.lo[2] = p[2], b.hi[0] = p[3], b.hi[1] = p[4], b.hi[2] = p[5];
  return b;
}

// node_boxes: (n-1) internal boxes; leaf boxes are read through the sorted index
__global__ void __launch_bounds__(256) k_refit(const float* __restrict__ prim_boxes, const unsigned* __restrict__ idx, int n,
                                               const int2* __restrict__ children, const int* __restrict__ parent_int,
                                               const int* __restrict__ parent_leaf, float* __restrict__ node_boxes,
                                               int* __restrict__ arrived) {
  const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
  if (leaf >= n) return;
  int node = parent_leaf[leaf];
  while (node >= 0) {
    if (atomicAdd(arrived + node, 1) == 0) return;  // the sibling subtree is not finished: it will do this node
    __threadfence();
    const int2 c = children[node];
    const Box6 a = load_box(c.x < 0 ? prim_boxes + 6 * (size_t)idx[~c.x] : node_boxes + 6 * (size_t)c.x);
    const Box6 b = load_box(c.y < 0 ? prim_boxes + 6 * (size_t)idx[~c.y] : node_boxes + 6 * (size_t)c.y);
    float* o = node_boxes + 6 * (size_t)node;
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = fminf(a.lo[k], b.lo[k]), o[3 + k] = fmaxf(a.hi[k], b.hi[k]);
    __threadfence();
    node = parent_int[node];
  }
}

__global__ void __launch_bounds__(256) k_emit(const float* __restrict__ prim_boxes, const unsigned* __restrict__ idx, int n,
                                              const int2* __restrict__ children, const float* __restrict__ node_boxes,
                                              const DPrim* __restrict__ prims_in, DNode* __restrict__ nodes,
                                              DPrim* __restrict__ prims_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) prims_out[i] = prims_in[idx[i]];
  if (i >= n - 1) return;
  const int2 c = children[i];
  const Box6 a = load_box(c.x < 0 ? prim_boxes + 6 * (size_t)idx[~c.x] : node_boxes + 6 * (size_t)c.x);
  const Box6 b = load_box(c.y < 0 ? prim_boxes + 6 * (size_t)idx[~c.y] : node_boxes + 6 * (size_t)c.y);
  DNode dn;
  dn.n0 = {a.lo[0], a.hi[0], a.lo[1], a.hi[1]};
  dn.n1 = {b.lo[0], b.hi[0], b.lo[1], b.hi[1]};
  dn.n2 = {a.lo[2], a.hi[2], b.lo[2], b.hi[2]};
  dn.n3 = {c.x < 0 ? ~(((~c.x) << 3) | 1) : c.x, c.y < 0 ? ~(((~c.y) << 3) | 1) : c.y, 0, 0};
  nodes[i] = dn;
}

__global__ void __launch_bounds__(256) k_depth(int n, const int* __restrict__ parent_int, const int* __restrict__ parent_leaf,
                                               int* __restrict__ depth) {
  const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
  if (leaf >= n) return;
  int d = 0;
  for (int node = parent_leaf[leaf]; node >= 0; node = parent_int[node]) ++d;
  atomicMax(depth, d);
}

}  // namespace

#define BVH_TRY(expr)                                                                                     \
  do {                                                                                                    \
    cudaError_t e_ = (expr);                                                                              \
    if (e_ != cudaSuccess) {                                                                              \
      for (void* p_ : tmp) cudaFree(p_);                                                                  \
      return fail(RTB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));                      \
    }                                                                                                     \
  } while (0)

// Builds the BVH of d->flat.prims (insertion order, n >= 2) on the current device; on success
// d->view.nodes / d->view.prims / d->view.root are set, d->flat.bvh_depth holds the depth and the
// allocations are owned by d->allocs.
int build_bvh_device(DeviceScene* d) {
  FlatScene& f = d->flat;
  const int n = (int)f.prims.size();
  if (n < 2 || f.prim_boxes.size() != (size_t)n * 6) return fail(RTB_ERR_INVALID, "build_bvh_device: needs >= 2 primitives with boxes");
  std::vector<void*> tmp;
  auto talloc = [&tmp](void** p, size_t bytes) {
    cudaError_t e = cudaMalloc(p, std::max<size_t>(bytes, 16));
    if (e == cudaSuccess) tmp.push_back(*p);
    return e;
  };
  float *d_boxes = nullptr, *d_node_boxes = nullptr;
  unsigned *d_bounds = nullptr, *d_keys = nullptr, *d_idx = nullptr, *d_keys2 = nullptr, *d_idx2 = nullptr;
  int2* d_children = nullptr;
  int *d_parent_int = nullptr, *d_parent_leaf = nullptr, *d_arrived = nullptr, *d_depth = nullptr;
  DPrim* d_prims_in = nullptr;
  void* d_sort_tmp = nullptr;
  BVH_TRY(talloc((void**)&d_boxes, sizeof(float) * 6 * n));
  BVH_TRY(talloc((void**)&d_node_boxes, sizeof(float) * 6 * (n - 1)));
  BVH_TRY(talloc((void**)&d_bounds, sizeof(unsigned) * 8));
  BVH_TRY(talloc((void**)&d_keys, sizeof(unsigned) * n));
  BVH_TRY(talloc((void**)&d_idx, sizeof(unsigned) * n));
  BVH_TRY(talloc((void**)&d_keys2, sizeof(unsigned) * n));
  BVH_TRY(talloc((void**)&d_idx2, sizeof(unsigned) * n));
  BVH_TRY(talloc((void**)&d_children, sizeof(int2) * (n - 1)));
  BVH_TRY(talloc((void**)&d_parent_int, sizeof(int) * (n - 1)));
  BVH_TRY(talloc((void**)&d_parent_leaf, sizeof(int) * n));
  BVH_TRY(talloc((void**)&d_arrived, sizeof(int) * (n - 1)));
  BVH_TRY(talloc((void**)&d_depth, sizeof(int) * 4));
  BVH_TRY(talloc((void**)&d_prims_in, sizeof(DPrim) * n));
  BVH_TRY(cudaMemcpy(d_boxes, f.prim_boxes.data(), sizeof(float) * 6 * n, cudaMemcpyHostToDevice));
  BVH_TRY(cudaMemcpy(d_prims_in, f.prims.data(), sizeof(DPrim) * n, cudaMemcpyHostToDevice));
  const unsigned init_bounds[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
  BVH_TRY(cudaMemcpy(d_bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice));
  BVH_TRY(cudaMemset(d_arrived, 0, sizeof(int) * (n - 1)));
  BVH_TRY(cudaMemset(d_depth, 0, sizeof(int) * 4));
  // the outputs live as long as the scene
  DNode* d_nodes = nullptr;
  DPrim* d_prims = nullptr;
  BVH_TRY(cudaMalloc((void**)&d_nodes, sizeof(DNode) * (n - 1)));
  d->allocs.push_back(d_nodes);
  BVH_TRY(cudaMalloc((void**)&d_prims, sizeof(DPrim) * n));
  d->allocs.push_back(d_prims);

  const unsigned gb = (unsigned)((n + 255) / 256);
  k_bounds<<<gb, 256>>>(d_boxes, n, d_bounds);
  k_morton<<<gb, 256>>>(d_boxes, n, d_bounds, d_keys, d_idx);
  size_t sort_bytes = 0;
  BVH_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, d_keys, d_keys2, d_idx, d_idx2, n, 0, 30));
  BVH_TRY(talloc(&d_sort_tmp, sort_bytes));
  BVH_TRY(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_bytes, d_keys, d_keys2, d_idx, d_idx2, n, 0, 30));
  k_hierarchy<<<gb, 256>>>(d_keys2, n, d_children, d_parent_int, d_parent_leaf);
  k_refit<<<gb, 256>>>(d_boxes, d_idx2, n, d_children, d_parent_int, d_parent_leaf, d_node_boxes, d_arrived);
  k_emit<<<gb, 256>>>(d_boxes, d_idx2, n, d_children, d_node_boxes, d_prims_in, d_nodes, d_prims);
  k_depth<<<gb, 256>>>(n, d_parent_int, d_parent_leaf, d_depth);
  BVH_TRY(cudaGetLastError());
  int depth = 0;
  BVH_TRY(cudaMemcpy(&depth, d_depth, sizeof(int), cudaMemcpyDeviceToHost));
  for (void* p : tmp) cudaFree(p);
  tmp.clear();
  if (depth > 60) return fail(RTB_ERR_INVALID, "build_bvh_device: BVH too deep for the traversal stack");
  f.bvh_depth = depth;
  f.root = 0;
  d->view.nodes = d_nodes;
  d->view.prims = d_prims;
  d->view.n_nodes = n - 1;
  d->n_nodes_device = n - 1;
  return RTB_OK;
}

}  // namespace rtb
