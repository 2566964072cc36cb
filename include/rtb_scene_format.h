/* rtb_scene_format.h — the flat scene-graph records shared by the host scene builder,
 * the CUDA flattener and the CPU oracle ("same bytes in => same scene on both sides").
 *
 * A scene is three tables of fixed-size records (textures, materials, objects) plus three
 * blob pools (object-handle lists, Perlin tables, image texels).  Handles are indices into
 * the tables.  The object table is a graph that mirrors the reference crate's object
 * nesting one-to-one (raytracing_rust: src/hittable.rs:18-21 `trait Hittable`,
 * src/material.rs:30-33 `trait Material`, src/texture.rs:4-6 `trait Texture`).
 *
 * On-disk / in-memory serialisation (.rtscene), little endian:
 *   RtbSceneHeader | RtbTexRec[n_tex] | RtbMatRec[n_mat] | RtbObjRec[n_obj]
 *   | int32 list_pool[n_list] | RtbPerlinTable[n_perlin] | uint8 image_pool[n_image_bytes]
 */
#ifndef RTB_SCENE_FORMAT_H
#define RTB_SCENE_FORMAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_SCENE_MAGIC 0x4E435352u /* "RSCN" */
#define RTB_SCENE_VERSION 1u

/* src/texture.rs:9,28,51,74 */
enum RtbTexType { RTB_TEX_SOLID = 0, RTB_TEX_CHECKER = 1, RTB_TEX_NOISE = 2, RTB_TEX_IMAGE = 3 };
/* src/material.rs:36,61,95,133,154 */
enum RtbMatType {
  RTB_MAT_LAMBERTIAN = 0,
  RTB_MAT_METAL = 1,
  RTB_MAT_DIELECTRIC = 2,
  RTB_MAT_DIFFUSE_LIGHT = 3,
  RTB_MAT_ISOTROPIC = 4
};
/* src/sphere.rs:20,87  src/rect.rs:14  src/cube.rs:8  src/hittable.rs:24,67
 * src/rotate.rs:21  src/traslate.rs:6  src/medium.rs:11  src/bvh.rs:9 */
enum RtbObjType {
  RTB_OBJ_SPHERE = 0,
  RTB_OBJ_MOVING_SPHERE = 1,
  RTB_OBJ_RECT = 2,
  RTB_OBJ_CUBE = 3,
  RTB_OBJ_FLIP_NORMALS = 4,
  RTB_OBJ_ROTATE = 5,
  RTB_OBJ_TRANSLATE = 6,
  RTB_OBJ_CONSTANT_MEDIUM = 7,
  RTB_OBJ_LIST = 8,
  RTB_OBJ_BVH = 9
};
/* src/rect.rs:8-12 — (k,a,b) axes: YZ (0,1,2), ZX (1,2,0), XY (2,0,1) */
enum RtbPlane { RTB_PLANE_YZ = 0, RTB_PLANE_ZX = 1, RTB_PLANE_XY = 2 };
/* src/rotate.rs:7-19 — (r,a,b) axes: X (0,1,2), Y (1,2,0), Z (2,0,1) */
enum RtbAxis { RTB_AXIS_X = 0, RTB_AXIS_Y = 1, RTB_AXIS_Z = 2 };

typedef struct RtbSceneHeader {
  uint32_t magic, version;
  uint32_t n_tex, n_mat, n_obj, n_list, n_perlin;
  uint32_t reserved;
  uint64_t n_image_bytes;
} RtbSceneHeader;

typedef struct RtbTexRec {
  int32_t type;   /* RtbTexType */
  int32_t a, b;   /* CHECKER: a = `odd` texture, b = `even` texture (src/texture.rs:34);
                     NOISE: a = Perlin table index; IMAGE: nx = a, ny = b */
  int32_t pad;
  uint64_t image_off; /* IMAGE: byte offset of RGB8 texels in the image pool */
  double v[3];    /* SOLID: r,g,b; NOISE: v[0] = scale */
} RtbTexRec;

typedef struct RtbMatRec {
  int32_t type;  /* RtbMatType */
  int32_t tex;   /* albedo / emit texture handle; -1 for DIELECTRIC */
  double param;  /* METAL: fuzz as passed by the caller (clamped to <= 1 on use,
                    src/material.rs:70); DIELECTRIC: ref_idx */
} RtbMatRec;

typedef struct RtbObjRec {
  int32_t type;     /* RtbObjType */
  int32_t mat;      /* SPHERE / MOVING_SPHERE / RECT / CUBE: material handle */
  int32_t child;    /* FLIP / ROTATE / TRANSLATE: wrapped object; MEDIUM: boundary object */
  int32_t sel;      /* RECT: RtbPlane; ROTATE: RtbAxis */
  int32_t list_off, list_n; /* LIST / BVH: slice of the list pool (insertion order) */
  int32_t tex;      /* MEDIUM: phase-function albedo texture */
  int32_t pad;
  uint64_t seed;    /* BVH: seed of the split-axis stream (the reference draws thread_rng,
                       src/bvh.rs:40; here it is reproducible) */
  /* SPHERE: cx,cy,cz,r                      (src/sphere.rs:27)
   * MOVING_SPHERE: c0[3], c1[3], t0, t1, r  (src/sphere.rs:97-104)
   * RECT: x0,y0,x1,y1,k                     (src/rect.rs:25)
   * CUBE: pmin[3], pmax[3]                  (src/cube.rs:15-19)
   * ROTATE: angle in degrees                (src/rotate.rs:30)
   * TRANSLATE: offset[3]                    (src/traslate.rs:12)
   * MEDIUM: density                         (src/medium.rs:18)
   * BVH: time0, time1                       (src/bvh.rs:17) */
  double v[9];
} RtbObjRec;

/* src/perlin.rs:59-64 — 256 unit vectors + three permutations of 0..255 */
typedef struct RtbPerlinTable {
  double ranvec[256][3];
  uint8_t perm_x[256], perm_y[256], perm_z[256];
} RtbPerlinTable;

/* src/camera.rs:21-31 — constructor arguments, in the constructor's order */
typedef struct RtbCamera {
  double look_from[3], look_at[3], view_up[3];
  double vertical_fov, aspect, aperture, focus_dist, time0, time1;
} RtbCamera;

#ifdef __cplusplus
}
#endif
#endif
