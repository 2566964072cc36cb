// render_internal.hpp — host-side state shared by render.cu and wavefront.cu (internal).
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "device_types.h"
#include "flatten.hpp"
#include "rtb200.h"

namespace rtb {

struct WavefrontState;  // wavefront.cu
void destroy_wavefront_state(WavefrontState*);

struct DeviceScene {
  int device = 0;
  FlatScene flat;   // host copy of what was uploaded (bounds, counts)
  SceneView view{};  // device pointers
  std::vector<void*> allocs;
  unsigned long long* d_stats = nullptr;  // 32 counters
  unsigned int* d_work = nullptr;         // work-item counter of the running megakernel launch
  cudaDeviceProp prop{};
  WavefrontState* wf = nullptr;
  int n_nodes_device = 0;  // BVH nodes resident on the device (host- or device-built)
};

// bvh_build.cu: LBVH over d->flat.prims on the device; sets view.nodes / view.prims / flat.root / flat.bvh_depth
int build_bvh_device(DeviceScene* d);

int ensure_uploaded(rtb_scene* s, int world, int device, double tlo, double thi);
void make_device_camera(const RtbCamera& c, DCamera* out);
// pool.cu: path pool in shared memory (RTB_PIPE_POOL); needs bvh depth < RTB_POOL_STACK
#ifndef RTB_POOL_ROWS
#define RTB_POOL_ROWS 32
#endif
#ifndef RTB_POOL_THREADS
#define RTB_POOL_THREADS 768
#endif
#ifndef RTB_POOL_STACK
#define RTB_POOL_STACK 16
#endif
bool pool_supports(const DeviceScene* d);
int launch_pool(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                cudaStream_t stream, uint64_t* launches);
// dual.cu: the megakernel's scheduler with two paths per lane (RTB_PIPE_DUAL)
int launch_dual(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                cudaStream_t stream, uint64_t* launches);
int launch_wavefront(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                     cudaStream_t stream, uint64_t* launches);

}  // namespace rtb
