/* rtb200.h — C ABI of librtb200.so, the B200 (sm_100a) path tracer that replaces the
 * per-pixel Monte-Carlo hot path of DrStiev/raytracing_rust:
 *     create_image (tests/test.rs:55-85)  ->  color (src/color.rs:6-23)
 *     -> Hittable::hit / Material::scatter,emitted / Texture::value.
 *
 * The reference has no FFI; its seam is the trait trio plus the `::new` constructors.
 * Each entry point below names the reference item it stands in for.  A Rust `-sys` crate
 *  binds exactly these symbols.
 *
 * Conventions: every call returns 0 on success, non-zero on error (rtb_last_error() has the
 * text); builder calls return a handle >= 0 or a negative error.  A scene and its handles are
 * owned by one rtb_scene and freed together.  Not thread-safe per scene.  Render calls block
 * until the image is in the caller's host buffers.  There is NO CPU fallback: render and
 * closest-hit calls fail with RTB_ERR_NO_DEVICE when no CUDA device is usable.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>
#include "rtb_scene_format.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rtb_scene rtb_scene;

enum RtbStatus {
  RTB_OK = 0,
  RTB_ERR_INVALID = -1,   /* bad handle / argument */
  RTB_ERR_NO_DEVICE = -2, /* no CUDA device / driver */
  RTB_ERR_CUDA = -3,      /* a CUDA call failed */
  RTB_ERR_NO_BBOX = -4,   /* reference: panic!("No bounding box in BVHNode") src/bvh.rs:30,58 */
  RTB_ERR_IO = -5
};

enum RtbBackground {
  RTB_BG_BLACK = 0, /* src/color.rs:21 — the reference as written */
  RTB_BG_SKY = 1    /* the gradient commented out at src/color.rs:18-20 (lit variant) */
};

enum RtbPipeline {
  RTB_PIPE_AUTO = 0,
  RTB_PIPE_MEGAKERNEL = 1, /* one persistent kernel, path regeneration */
  RTB_PIPE_WAVEFRONT = 2,  /* raygen / extend / per-material shade queues, path state in HBM */
  RTB_PIPE_POOL = 3,       /* removed (round-1 experiment: per-SM path pool in shared memory; slower, profiles/r1_pool.txt) */
  RTB_PIPE_DUAL = 4        /* removed (round-1 experiment: two paths per lane; slower, profiles/r1_dual.txt) */
};

const char* rtb_last_error(void);
const char* rtb_version(void);
/* number of usable CUDA devices (0 when there is none; never fails) */
int rtb_device_count(void);

/* ---- scene lifetime ------------------------------------------------------------------- */
rtb_scene* rtb_scene_create(void);
void rtb_scene_destroy(rtb_scene*);

/* ---- textures: src/texture.rs --------------------------------------------------------- */
int rtb_tex_solid(rtb_scene*, double r, double g, double b);              /* SolidTexture::new :14 */
int rtb_tex_checker(rtb_scene*, int odd, int even);                       /* CheckerTexture::new :34 */
/* NoiseTexture::new :57.  The reference draws fresh Perlin tables from thread_rng
 * (src/perlin.rs:12-36); here the tables come from `seed` (same construction, reproducible).
 * Cloning a NoiseTexture shares its tables: pass the same handle twice. */
int rtb_tex_noise(rtb_scene*, double scale, uint64_t seed);
/* same, with caller-provided tables (ranvec[256*3], perm_{x,y,z}[256]) */
int rtb_tex_noise_tables(rtb_scene*, double scale, const double* ranvec, const uint8_t* perm_x,
                         const uint8_t* perm_y, const uint8_t* perm_z);
int rtb_tex_image(rtb_scene*, const uint8_t* rgb8, uint32_t nx, uint32_t ny); /* ImageTexture::new :81 */

/* ---- materials: src/material.rs ------------------------------------------------------- */
int rtb_mat_lambertian(rtb_scene*, int albedo_tex);          /* Lambertian::new :41 */
int rtb_mat_metal(rtb_scene*, int albedo_tex, double fuzz);  /* Metal::new :67 */
int rtb_mat_dielectric(rtb_scene*, double ref_idx);          /* Dielectric::new :100 */
int rtb_mat_diffuse_light(rtb_scene*, int emit_tex);         /* DiffuseLight::new :138 */
int rtb_mat_isotropic(rtb_scene*, int albedo_tex);           /* Isotropic::new :159 */

/* ---- hittables ------------------------------------------------------------------------ */
int rtb_obj_sphere(rtb_scene*, const double center[3], double radius, int mat); /* src/sphere.rs:27 */
int rtb_obj_moving_sphere(rtb_scene*, const double c0[3], const double c1[3], double t0, double t1,
                          double radius, int mat);                              /* src/sphere.rs:97 */
int rtb_obj_rect(rtb_scene*, int plane /*RtbPlane*/, double x0, double y0, double x1, double y1,
                 double k, int mat);                                            /* src/rect.rs:25 */
int rtb_obj_cube(rtb_scene*, const double p_min[3], const double p_max[3], int mat); /* src/cube.rs:15 */
int rtb_obj_flip_normals(rtb_scene*, int obj);                                  /* src/hittable.rs:72 */
int rtb_obj_rotate(rtb_scene*, int axis /*RtbAxis*/, int obj, double angle_deg); /* src/rotate.rs:30 */
int rtb_obj_translate(rtb_scene*, int obj, const double offset[3]);             /* src/traslate.rs:12 */
int rtb_obj_constant_medium(rtb_scene*, int boundary, double density, int tex); /* src/medium.rs:18 */
int rtb_obj_list(rtb_scene*, const int* objs, int n);         /* HittableList push order, src/hittable.rs:23-34 */
/* BVHNode::new(&mut [Rc<dyn Hittable>], time0, time1), src/bvh.rs:17.  Fails with
 * RTB_ERR_NO_BBOX if n == 0.  `seed` drives the oracle's split-axis stream; the GPU builds its
 * own SAH tree over the TRUE bounds of the same primitives, so a BVHNode answers exactly like the
 * HittableList of the same children.  That equals the reference wherever the reference's bounding
 * boxes are right — every shipped scene — and differs, deliberately, in two corners where upstream
 * bounding-box bugs leak into its results:
 *   * Rect::bounding_box ignores `plane` (src/rect.rs:71-75, Q8): a ZX / YZ Rect placed DIRECTLY in
 *     a BVHNode is culled upstream by a box in the wrong plane; here it is hit (XY Rects and Cubes,
 *     which override the box, agree).  tests/test_gpu_fields.py pins this.
 *   * a one-element BVHNode stores its child in both slots (src/bvh.rs:43-45): a ConstantMedium
 *     there draws two free-flight distances upstream and keeps the nearer; here it draws one. */
int rtb_obj_bvh(rtb_scene*, const int* objs, int n, double time0, double time1, uint64_t seed);

/* ---- named scenes: tests/test.rs:89-523 ---------------------------------------------------
 * name in {"random_scene","two_spheres","two_perlin_sphere","earth","simple_light",
 *          "cornell_box","cornell_smoke","final_scene"} as written in the reference, plus the
 * labelled lit variants {"final_scene_lit","cornell_smoke_book"}.  `seed` replaces thread_rng
 * in the constructors; `earth_rgb8` (1024x512 decoded earthmap.jpg) is needed by earth and
 * final_scene*, else may be NULL.  Returns the world handle; *camera gets the camera the
 * matching #[test] uses (tests/test.rs:543-830) for the given nx, ny. */
int rtb_scene_named(rtb_scene*, const char* name, uint64_t seed, const uint8_t* earth_rgb8,
                    uint32_t earth_nx, uint32_t earth_ny, int nx, int ny, RtbCamera* camera);

/* ---- serialisation (.rtscene, include/rtb_scene_format.h) ----------------------------- */
size_t rtb_scene_serialized_size(const rtb_scene*);
int rtb_scene_serialize(const rtb_scene*, void* buf, size_t cap);
rtb_scene* rtb_scene_deserialize(const void* buf, size_t len);
/* the same bytes as a .rtscene file (the format host, oracle and GPU library share) */
int rtb_scene_save(const rtb_scene*, const char* path);
rtb_scene* rtb_scene_load(const char* path);
/* number of leaf-primitive IDs under `world` (depth-first insertion order; a Cube owns 6,
 * in the order of src/cube.rs:21-74) */
int rtb_scene_prim_count(rtb_scene*, int world);

/* Who builds the BVH the kernels traverse (BVHNode::new, src/bvh.rs:17-66, is a host recursion
 * upstream; closest-hit answers do not depend on the tree).  HOST_SAH: full-sweep SAH on the host
 * (best trees).  DEVICE_LBVH: Morton-sorted linear BVH built by CUDA kernels (csrc/bvh_build.cu),
 * for large scenes.  AUTO: the device builder from RTB_BVH_AUTO_DEVICE_MIN primitive IDs on.
 * Takes effect at the next upload. */
typedef enum RtbBvhBuilder { RTB_BVH_AUTO = 0, RTB_BVH_HOST_SAH = 1, RTB_BVH_DEVICE_LBVH = 2 } RtbBvhBuilder;
#define RTB_BVH_AUTO_DEVICE_MIN 100000
int rtb_scene_set_bvh_builder(rtb_scene*, int builder);

/* ---- rendering ---------------------------------------------------------------------------- */
typedef struct rtb_render_opts {
  int max_depth;       /* 50  (src/color.rs:9) */
  double t_min;        /* 1e-3 (src/color.rs:7) */
  int background;      /* RtbBackground */
  uint64_t seed;       /* render seed: Philox4x32-10 key; counter = (block, pixel, sample) */
  int sample_begin;    /* this call renders samples s = sample_begin + i*sample_stride < ns */
  int sample_stride;   /* (rank, world_size) for sample sharding across GPUs; 0,1 for all */
  int pipeline;        /* RtbPipeline */
  int device;          /* CUDA device ordinal */
  int collect_stats;   /* count segments / node tests / primitive tests (slower, untimed runs) */
  int band_begin;      /* row-band sharding (the "image tiles" split of the pixel loop, tests/test.rs:62-63): this call */
  int band_stride;     /* renders the 4-row bands b = band_begin + i*band_stride of the image; 0,1 (or 0,0) for all rows */
  int multi_shard;     /* rtb_render_multi only: RtbMultiShard — how the call splits the work over its devices */
  int reserved[7];     /* 0 = defaults.  Tuning / experiment knobs of the megakernel, no effect on the result beyond float
                        * summation order: [0..2] lanes a NODE / LEAF / SHADE round waits for, [3] occupancy point (2..4 blocks
                        * of 256 threads per SM, 5..7 of 128), [4] wavefront pipeline: path slots, [5] shared-memory carveout %,
                        * [6] closest-hit form: 1 flat scan (scenes of <= 16 primitives), 2 two-child BVH, 4 four-wide BVH */
} rtb_render_opts;
/* rtb_render_multi: RTB_SHARD_SAMPLES — device g renders samples g, g+G, ... of every pixel, one sum-reduce closes the
 * image (best when there are many samples per pixel); RTB_SHARD_BANDS — device g renders every sample of the 4-row bands
 * g, g+G, ...: disjoint pixels, no arithmetic in the exchange, only rgb8 rows move (best at low samples per pixel). */
typedef enum RtbMultiShard { RTB_SHARD_SAMPLES = 0, RTB_SHARD_BANDS = 1 } RtbMultiShard;
#define RTB_BAND_ROWS 4

typedef struct rtb_stats {
  uint64_t paths, segments;
  uint64_t node_tests, prim_tests, medium_tests;
  uint64_t shade[5];        /* per RtbMatType */
  uint64_t tex_evals[4];    /* per RtbTexType */
  uint64_t kernel_launches; /* launches of this library's kernels in the call */
  double device_ms;         /* CUDA-event time of the render kernels */
  uint64_t reserved[8];
} rtb_stats;

void rtb_render_opts_default(rtb_render_opts*);

/* create_image(ny, nx, ns, cam, world) -> String, tests/test.rs:55-85 (argument order is the
 * caller's business here: nx is the width).  Host buffers; either output may be NULL.
 *   out_sum_rgb : 3*nx*ny floats, row 0 = top of the image (j = ny-1), SUM of the radiance
 *                 samples this call rendered (divide by the sample count for the mean)
 *   out_rgb8    : 3*nx*ny bytes, the reference quantisation tests/test.rs:71-78 of the mean
 *                 (only meaningful when the call renders all ns samples) */
int rtb_render(rtb_scene*, int world, const RtbCamera*, int nx, int ny, int ns,
               const rtb_render_opts*, float* out_sum_rgb, uint8_t* out_rgb8, rtb_stats*);
/* rtb_render keeps its device accumulation / rgb8 buffers per device between calls (grow-only; a 4K image:
 * 124 MB).  This frees the ones not in use.  Always RTB_OK. */
int rtb_release_workspace(void);

/* create_image in chunks of `chunk_spp` samples with real progress reporting (in place of the
 * reference's dead src/progressbar.rs and the start/end lines of src/logger.rs): after every chunk
 * `cb(user, samples_done, samples_total, rgb8_or_NULL, mpaths_per_s, eta_seconds)` is called; every
 * `preview_every`-th chunk (0 = never) rgb8 is the host image so far, 3*nx*ny bytes valid during
 * the call.  A non-zero return stops the render; the outputs then hold the mean of the
 * *samples_done samples rendered.  Single-GPU call (sample_begin/stride must be 0/1). */
typedef int (*rtb_progress_fn)(void* user, int samples_done, int samples_total, const uint8_t* rgb8,
                               double mpaths_per_s, double eta_seconds);
int rtb_render_progressive(rtb_scene*, int world, const RtbCamera*, int nx, int ny, int ns,
                           const rtb_render_opts*, int chunk_spp, int preview_every, rtb_progress_fn cb,
                           void* user, float* out_sum_rgb, uint8_t* out_rgb8, int* samples_done);

/* Device-resident variant for multi-GPU and benchmarking: the scene is uploaded once
 * (rtb_scene_upload), samples are ADDED into d_sum_rgb (device pointer, 3*nx*ny floats,
 * caller-zeroed) on `cuda_stream` (a cudaStream_t cast to void*, NULL = default stream)
 * without synchronising. */
int rtb_scene_upload(rtb_scene*, int world, int device);
/* bytes of flattened scene tables resident on the device after rtb_scene_upload (0 if none) */
size_t rtb_scene_device_bytes(const rtb_scene*);
int rtb_render_device(rtb_scene*, const RtbCamera*, int nx, int ny, int ns, const rtb_render_opts*,
                      float* d_sum_rgb, void* cuda_stream, rtb_stats*);
/* tests/test.rs:71-78: /ns, sqrt, clamp[0,1], (255.99*c) as i32 — on device */
int rtb_resolve_device(const float* d_sum_rgb, int nx, int ny, int ns, uint8_t* d_rgb8,
                       void* cuda_stream);

/* ---- multi-GPU exchange step over peer memory (csrc/peer.cu) ------------------------------
 * The reference has no counterpart (single thread, tests/test.rs:64-70 keeps one accumulator per
 * pixel); with samples sharded over G GPUs these calls combine the G partial sums.
 *   rtb_device_alloc / _free : a zeroed accumulation buffer that is a whole CUDA allocation, so
 *                              its base pointer can be exported to the other ranks' processes
 *   rtb_peer_export          : cudaIpcGetMemHandle -> 64 opaque bytes to hand to the root rank
 *   rtb_peer_open / _close   : map a peer's buffer into this process (P2P over NVLink)
 *   rtb_reduce_resolve_device: ONE kernel on the root: sum of n_bufs buffers (own + peer-mapped,
 *                              added in array order) -> tests/test.rs:71-78 quantisation into
 *                              d_rgb8, and the f32 total into d_total if it is not NULL */
#define RTB_IPC_HANDLE_BYTES 64
#define RTB_MAX_PEER_BUFFERS 16
int rtb_device_alloc(size_t bytes, int device, void** d_ptr);
int rtb_device_free(void* d_ptr);
int rtb_peer_export(const void* d_ptr, void* handle64);
int rtb_peer_open(const void* handle64, int device, void** d_ptr);
int rtb_peer_close(void* d_ptr);
int rtb_reduce_resolve_device(const float* const* d_sums, int n_bufs, int nx, int ny, int ns,
                              float* d_total, uint8_t* d_rgb8, void* cuda_stream);
/* The SHARDED form of that exchange: the same kernel over the value range [first_value, first_value + n_values) of the
 * full-size buffers (all pointers are BASE pointers; first_value a multiple of 4 keeps the 128-bit path).  With G ranks
 * each calling it for its own 1/G of the image — d_sums = every rank's buffer (peer-mapped), d_rgb8 = the root's rgb8
 * buffer (peer-mapped) — every GPU pulls (G-1)/G of ITS slice over NVLink instead of the root pulling G-1 whole images,
 * and the quantised slice lands on the root directly. */
int rtb_reduce_resolve_range_device(const float* const* d_sums, int n_bufs, size_t first_value, size_t n_values, int ns,
                                    float* d_total, uint8_t* d_rgb8, void* cuda_stream);
/* Row-band sharding (rtb_render_opts.band_begin / band_stride): quantise this shard's 4-row bands of d_sum_rgb (a
 * full-size accumulation buffer).  compact == 0: into their place in a full-size rgb8 image — the caller's own or the
 * root's peer-mapped one (the shards' bands tile the image, no arithmetic in the exchange); compact != 0: packed
 * consecutively, band i of the shard at row 4i (for a gather). */
int rtb_resolve_bands_device(const float* d_sum_rgb, int nx, int ny, int ns, int band_begin, int band_stride,
                             uint8_t* d_rgb8, int compact, void* cuda_stream);

/* create_image on several GPUs of this node from one call (csrc/multi.cu): one host thread per device renders its
 * shard — the sample slice s = i, i+G, ... of every pixel (opts->multi_shard = RTB_SHARD_SAMPLES, default) or every
 * sample of the 4-row bands i, i+G, ... (RTB_SHARD_BANDS) — then the exchange runs sharded over the devices: device g
 * sums + quantises 1/G of the pixels from all G partial images over peer access (band shards: quantises its own rows)
 * and stores the result into devices[0]'s image, which goes to the host once.
 * devices == NULL: ordinals 0 .. n_devices-1 (n_devices <= 0: every visible device).  The image is the one rtb_render
 * produces for the same seed, whatever G is (the draws are counter-based: same sample set; sums equal up to the order
 * of the float adds, <= 1 LSB after quantisation).  opts->device, sample_begin/stride and band_begin/stride are ignored /
 * must be 0,1.  stats (optional): sums over the devices, device_ms = the slowest shard, reserved[0] = G. */
int rtb_render_multi(rtb_scene*, int world, const RtbCamera*, int nx, int ny, int ns, const rtb_render_opts*,
                     const int* devices, int n_devices, float* out_sum_rgb, uint8_t* out_rgb8, rtb_stats*);

/* Philox4x32-10 (the renderer's counter-based generator, csrc/philox.h) on the host, for known-answer tests. */
int rtb_philox4x32_10(const uint32_t counter[4], const uint32_t key[2], uint32_t out[4]);

/* Issue-rate calibration for the roofline denominator (csrc/calib.cu): lane-instructions per second
 * (in 10^12) the whole device sustains on FFMA chains alone and on FFMA + LOP3 chains interleaved
 * 1:1 (both the FMA and the ALU pipe busy), and on the 1 FFMA : 2 FMNMX mix of the BVH slab test.
 * ~0.1 s of GPU time. */
int rtb_issue_peak(int device, double* fma_tlane_per_s, double* mixed_tlane_per_s, double* slab_tlane_per_s);

/* Closest-hit query for the primitive-ID parity test: Hittable::hit(ray, t_min, f64::MAX)
 * (src/color.rs:7) with ConstantMedium objects skipped.  rays = n x {ox,oy,oz,dx,dy,dz,time}
 * (host, double); out_prim[i] = primitive ID or -1; out_t[i] = t (undefined on a miss). */
int rtb_closest_hit(rtb_scene*, int world, const double* rays, size_t n, double t_min, int device,
                    int32_t* out_prim, float* out_t);
/* Camera::get_ray (src/camera.rs:53-67) for the fixed parity batch: ray i is sample `s` of
 * pixel i of the nx*ny image, drawn from the same Philox stream the renderer uses; host
 * output n x 7 doubles.  Pure host code (no device). */
int rtb_primary_rays(const RtbCamera*, int nx, int ny, int s, uint64_t seed, double* out_rays);

/* Texture::value (src/texture.rs:4-6) over a batch, evaluated by the device texture code —
 * a test hook for Perlin / checker / image parity.  uvp = n x {u,v,px,py,pz} (host, double);
 * out = n x {r,g,b} (host, float). */
int rtb_debug_texture_value(rtb_scene*, int world, int tex, const double* uvp, size_t n, int device,
                            float* out);

/* ---- field-level parity hooks (csrc/debug.cu): one stage of the device path at a time ------------------------
 * rtb_closest_hit_ex: rtb_closest_hit plus (a) self_id (optional, n entries): the primitive ID of the surface ray i
 * leaves (-1: none).  The kernels exclude that surface's own root analytically where the f64 reference relies on
 * t_min = 0.001 (src/color.rs:7); this is the query every scattered ray of a render issues.  (b) out_rec (optional):
 * n x {t, u, v, p.xyz, normal.xyz} = the HitRecord of src/hittable.rs:9-16 (u,v always computed, as upstream). */
int rtb_closest_hit_ex(rtb_scene*, int world, const double* rays, size_t n, double t_min, const int32_t* self_id,
                       int device, int32_t* out_prim, float* out_rec);
/* Camera::get_ray (src/camera.rs:53-67) evaluated by the device code of the render kernels (f32) for sample `s` of
 * every pixel: out = nx*ny x {o.xyz, d.xyz, time} floats.  rtb_primary_rays is the same batch in host f64. */
int rtb_debug_camera_rays(const RtbCamera*, int nx, int ny, int s, uint64_t seed, int device, float* out_rays);
/* The random inputs behind that batch, in the layout an explicit-input get_ray wants: nx*ny x {s, t, rd.x, rd.y,
 * xi_time} with rd the unit-disk point of src/util.rs:15-24 (before the lens-radius scale).  Host code. */
int rtb_debug_camera_draws(int nx, int ny, int s, uint64_t seed, double* out);
/* Material::scatter (src/material.rs:49-168) + the albedo lookup as the SHADE phase runs them.  in = n x {ray o.xyz,
 * d.xyz, time, hit t, u, v, p.xyz, normal.xyz} (16 doubles); out = n x {scattered?, o.xyz, d.xyz, attenuation.rgb}
 * (10 floats, zeros when the path ends).  Item i uses the bounce-0 draws of (pixel i, sample 0) under `seed`. */
int rtb_debug_scatter(rtb_scene*, int world, int mat, const double* in, size_t n, uint64_t seed, int device, float* out);
/* The device samplers that replace the rejection loops of src/util.rs:4-24: kind 0 = uniform in the unit ball (scatter),
 * kind 1 = uniform in the unit disk (lens).  out = n x {x, y, z} floats. */
int rtb_debug_sample(int kind, uint64_t seed, size_t n, int device, float* out);

/* rtb_closest_hit_ex for callers that cannot allocate (no counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc).  The call makes
 * NO allocation of any kind: inputs and outputs are DEVICE pointers of the caller — d_rays n x {o.xyz, d.xyz, time} floats,
 * d_self_id n ints or NULL, d_out_prim n ints, d_out_rec n x 9 floats or NULL — and the one piece of scratch the query needs
 * (the primitive-ID -> primitive table behind self_id) lives in a static device array of RTB_NOALLOC_MAX_PRIM_IDS entries
 * (256 KB, held by every GPU that loads the library).  The scene must already be on its device (rtb_scene_upload, which does
 * allocate, is the caller's set-up step; moving spheres are bounded for the shutter [0, 1]); scenes with more primitive IDs
 * than the table holds are refused with RTB_ERR_INVALID.  Work is enqueued on `cuda_stream` (NULL = default stream) without
 * synchronising.  Every call shares the one table: it is cleared and refilled per call, and two calls that overlap on the
 * GPU (different streams, or different scenes from different threads) race on it — serialise them.
 * Answers are those of rtb_closest_hit_ex on the same rays, bit for bit. */
#define RTB_NOALLOC_MAX_PRIM_IDS 65536
int rtb_closest_hit_noalloc(rtb_scene*, const float* d_rays, size_t n, double t_min, const int32_t* d_self_id,
                            int32_t* d_out_prim, float* d_out_rec, void* cuda_stream);

/* Self-checking builds (make NVEXTRA=-DRTB_DEBUG_CHECKS=1): every index the kernels form (traversal stack, node /
 * primitive / material / texture / texel, shared tiles, output pixels, wavefront queues) is bounds-checked on the device and
 * violations are counted per class; this reads and clears the 16 counters (all zero = clean).  The stand-in for
 * compute-sanitizer, which the GPU pool does not offer.  Release builds: RTB_ERR_INVALID. */
int rtb_debug_check_failures(int device, uint32_t out[16]);

/* P3 text exactly as tests/test.rs:59,79: "P3\n{nx} {ny}\n255\n" + "{r} {g} {b}\n" per pixel.
 * Returns the byte count (excluding NUL); writes at most cap bytes. */
size_t rtb_ppm_p3(const uint8_t* rgb8, int nx, int ny, char* out, size_t cap);
/* Binary PPM (P6) of the same image: "P6\n{nx} {ny}\n255\n" + 3*nx*ny raw bytes. */
size_t rtb_ppm_p6(const uint8_t* rgb8, int nx, int ny, char* out, size_t cap);
/* File::create(path) + write!(file, "{}", create_image(..)) of the #[test] wrappers
 * (tests/test.rs:536,557): binary == 0 writes the reference's P3 text, else P6. */
int rtb_image_write_ppm(const char* path, const uint8_t* rgb8, int nx, int ny, int binary);
/* The same image as an 8-bit RGB PNG (zlib stored blocks; any PNG reader opens it). */
int rtb_image_write_png(const char* path, const uint8_t* rgb8, int nx, int ny);

/* ---- JPEG input: image::open(path).to_rgb8() of tests/test.rs:201-205,489-493 ---------------
 * (upstream: crates image 0.25.1 / zune-jpeg 0.4.11, Cargo.lock:376-394,1363-1364).  Baseline and
 * extended-sequential Huffman JPEG, 8 bit, grey or 3 components with 1x1 / 2x1 / 2x2 chroma
 * sampling, restart intervals; integer slow IDCT, triangle chroma interpolation and fixed-point
 * YCbCr->RGB, i.e. byte-identical to the libjpeg(-turbo) default decode (tests/test_image_io.py).
 * *rgb8_out is malloc'ed (3*nx*ny bytes, row 0 = top) and released with rtb_image_free.
 * Progressive / arithmetic / 12-bit files fail with RTB_ERR_INVALID; unreadable files RTB_ERR_IO. */
int rtb_jpeg_decode(const uint8_t* data, size_t len, uint8_t** rgb8_out, uint32_t* nx, uint32_t* ny);
int rtb_jpeg_decode_file(const char* path, uint8_t** rgb8_out, uint32_t* nx, uint32_t* ny);
void rtb_image_free(uint8_t* rgb8);
/* ImageTexture::new(image::open(path).to_rgb8().into_raw(), nx, ny): decode + rtb_tex_image. */
int rtb_tex_image_file(rtb_scene*, const char* path);

#ifdef __cplusplus
}
#endif
#endif
