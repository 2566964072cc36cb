// render_internal.hpp — host-side state shared by render.cu and wavefront.cu (internal).
#pragma once
#include <cuda_runtime.h>

#include <utility>
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
  std::vector<void*> allocs;                      // plain cudaMalloc'ed (device BVH builder)
  std::vector<std::pair<void*, size_t>> pooled;  // from the per-device allocation pool (render.cu pool_alloc)
  unsigned long long* d_stats = nullptr;  // 32 counters
  unsigned int* d_work = nullptr;         // work-item counter of the running megakernel launch
  cudaDeviceProp prop{};
  WavefrontState* wf = nullptr;
  int n_nodes_device = 0;  // BVH nodes resident on the device (host- or device-built)
  FlatTables flat_tables{};  // the primitives by class when the scene fits the flat scan (<= RTB_FLAT_MAX_PRIMS), else n = 0
  const DNode4* d_nodes4 = nullptr;  // the four-wide form of the tree (host builder only); view.nodes4 = this or nullptr
};

// The scene tables as a kernel should see them: width 0 = the upload-time choice (d->view), 1 = no tree (the megakernel's flat
// scan; only scenes that fit FlatTables), 2 = the two-child tree, 4 = the four-wide tree when it exists.  RTB200_BVH_WIDTH=1|2|4
// in the environment overrides width 0 (experiments).
SceneView view_with_width(const DeviceScene* d, int width);

// bvh_build.cu: LBVH over d->flat.prims on the device; sets view.nodes / view.prims / flat.root / flat.bvh_depth
int build_bvh_device(DeviceScene* d);

// per-translation-unit readers of the RTB_DEBUG_CHECKS counters (trace.cuh RTB_DEFINE_CHECK_READER): add into acc[16], clear
int read_checks_render(unsigned int* acc);
int read_checks_wavefront(unsigned int* acc);
int read_checks_debug(unsigned int* acc);

// per-device pool of device allocations keyed by size (render.cu); `device` must be the current device
cudaError_t pool_alloc(int device, size_t bytes, void** out);
void pool_free(int device, void* p, size_t bytes);

int ensure_uploaded(rtb_scene* s, int world, int device, double tlo, double thi);
void make_device_camera(const RtbCamera& c, DCamera* out);
int launch_wavefront(DeviceScene* d, const DCamera& cam, int nx, int ny, int ns, const rtb_render_opts& o, float* d_sum,
                     cudaStream_t stream, uint64_t* launches);

}  // namespace rtb
