/* b2rt.h -- C ABI of the B200-native ray-intersection / photon-mapping backend.
 *
 * This is the drop-in boundary underneath the reference's pbrt-facing entry points
 * (cuda_render/cudaapi.h:8-19, class CudaRender cuda_render/cudarender.h:22-91).  The C++ shim
 * that implements those four functions against pbrt headers calls ONLY the functions below.
 * Plain pointers and sizes; no pbrt, torch, OptiX or C++ types.  Every call returns 0 on success
 * or a negative B2RT_E_* code; b2rt_last_error() gives the message.  Single-threaded contract, like
 * the reference (one static CudaRender*, cuda_render/cudaapi.cpp:3).
 *
 * There is NO CPU fallback: if no CUDA device is usable every call fails with B2RT_E_CUDA.
 *
 * Primitive ids are global and follow submission order: the triangles of the first mesh are
 * 0..ntris-1, the next shape continues from there (same rule pbrt uses for shapeId,
 * pbrt-v2/src/core/shape.cpp:45).
 */
#ifndef B2RT_H
#define B2RT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2rt_ctx b2rt_ctx;

enum {
    B2RT_OK = 0,
    B2RT_E_ARG = -1,      /* bad argument */
    B2RT_E_CUDA = -2,     /* CUDA runtime error (message has the cudaError string) */
    B2RT_E_STATE = -3,    /* call out of order (e.g. trace before build) */
    B2RT_E_OVERFLOW = -4  /* a device-side capacity was exceeded (stack, photon buffer) */
};

enum { B2RT_HOST = 0, B2RT_DEVICE = 1 };  /* where a caller-supplied buffer lives */

/* 32-byte ray: {ox,oy,oz,tmin, dx,dy,dz,tmax}.  Same fields as optix::Ray as the reference fills it
 * (cuda_render/simple_render/simplerender.cu:22-28) minus the ray type. */
typedef struct { float ox, oy, oz, tmin, dx, dy, dz, tmax; } b2rt_ray;
/* 16-byte hit: t = +INF and prim = -1 on a miss; (b1,b2) are pbrt's barycentrics
 * (pbrt-v2/src/shapes/trianglemesh.cpp:164-172); 0 for spheres/disks (their attributes are
 * recomputed from the hit point by the shade stage). */
typedef struct { float t; int32_t prim; float b1, b2; } b2rt_hit;

/* materials / lights: the subset cuda_render translates (util/material/cudamaterial.cpp:8-21,
 * util/light/cudalight.cpp:16-71) */
enum { B2RT_MAT_MATTE = 0, B2RT_MAT_MIRROR = 1, B2RT_MAT_GLASS = 2 };
enum { B2RT_LIGHT_POINT = 0, B2RT_LIGHT_DISK_AREA = 1 };

typedef struct {
    int32_t n_prims, n_tris, n_wide_nodes, n_leaf_slots;
    int32_t max_depth, wide_width;
    float   bounds[6];          /* scene AABB lo.xyz hi.xyz */
    float   sah_cost;           /* SAH cost of the wide tree (node cost 1, prim cost 1) */
    float   build_ms;           /* device time of the last b2rt_build_accel */
} b2rt_accel_info;

/* ---- lifetime ------------------------------------------------------------------------------ */
/* replaces CudaRenderInit (cuda_render/cudaapi.cpp:4-7, cudarender.cpp:17-36) */
int  b2rt_create(int device, uint64_t seed, b2rt_ctx **out);
void b2rt_destroy(b2rt_ctx *ctx);
const char *b2rt_last_error(b2rt_ctx *ctx);   /* ctx may be NULL: last create error */
int  b2rt_device_count(void);

/* ---- scene ingestion: replaces CudaRender::createCudaShape (cudarender.cpp:136-196) ----------- */
/* world-space vertices (pbrt pre-transforms them, shapes/trianglemesh.cpp:69-71).  N, uv may be NULL.
 * *first_prim receives the global id of triangle 0. */
int b2rt_add_mesh(b2rt_ctx *ctx, int32_t nverts, const float *P, int32_t ntris, const int32_t *idx,
                  const float *N, const float *uv, int32_t material_id, int32_t arealight_id,
                  int32_t *first_prim);
/* o2w, w2o: row-major 3x4 object->world and its inverse exactly as pbrt holds them
 * (cuda_render/util/shape/cudasphere.cpp:15-40); w2o may be NULL (then it is computed here). */
int b2rt_add_sphere(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float radius, int32_t material_id,
                    int32_t arealight_id, int32_t *prim);
/* disk at object z = height (cuda_render/util/shape/cudadisk.cpp:15-45) */
int b2rt_add_disk(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float height, float radius,
                  float inner_radius, int32_t material_id, int32_t arealight_id, int32_t *prim);
int b2rt_add_material(b2rt_ctx *ctx, int32_t type, const float rgb_kd_or_kr[3], const float rgb_kt[3],
                      float ior, int32_t *material_id);
/* point: p = position, rgb = I.  disk area: rgb = L, prim = the disk primitive, nsamples. */
int b2rt_add_light(b2rt_ctx *ctx, int32_t type, const float p[3], const float rgb[3], int32_t prim,
                   int32_t nsamples, int32_t *light_id);
int b2rt_clear_scene(b2rt_ctx *ctx);

/* ---- acceleration structure: replaces OptiX "Sbvh"/"Bvh" (cudarender.cpp:44,50) ---------------- */
int b2rt_build_accel(b2rt_ctx *ctx);
int b2rt_accel_info_get(b2rt_ctx *ctx, b2rt_accel_info *info);
/* parity/determinism hooks.  Any pointer may be NULL.  Sizes from b2rt_accel_info.
 *   morton[n_prims]      sorted 32-bit keys          prim_order[n_prims]  sorted prim ids
 *   bin_parent[2n-1]     Karras parent index (internal i at i, leaf j at n-1+j; root parent -1)
 *   bin_child[2*(n-1)]   left,right of each internal (>=n-1 means leaf id + n-1)
 *   bin_box[6*(2n-1)]    refit AABBs
 *   wide_nodes[32*n_wide_nodes] raw 128-byte nodes (as uint32)   leaf_prims[n_leaf_slots] */
int b2rt_accel_download(b2rt_ctx *ctx, uint32_t *morton, int32_t *prim_order, int32_t *bin_parent,
                        int32_t *bin_child, float *bin_box, uint32_t *wide_nodes, int32_t *leaf_prims);

/* ---- rays: replaces rtTrace call sites (simplerender.cu:27,68; raytracing.cu:24,76,102) --------- */
/* where = B2RT_HOST: caller memory, copied in/out inside the call.  B2RT_DEVICE: device pointers,
 * asynchronous on the context stream. */
int b2rt_trace_closest(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, b2rt_hit *hits, int where);
int b2rt_trace_any(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, uint8_t *occluded, int where);

/* generic device sort exposed for tests/bench of the hand-written radix sort (K2):
 * stable LSD sort of (key32,value32) pairs on `key_bits` low bits; device pointers only. */
int b2rt_sort_pairs(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, int key_bits);

/* ---- timing / stream ----------------------------------------------------------------------- */
int b2rt_sync(b2rt_ctx *ctx);
int b2rt_timer_start(b2rt_ctx *ctx);            /* cudaEventRecord on the context stream */
int b2rt_timer_stop(b2rt_ctx *ctx, float *ms);  /* record + synchronize + elapsed */
int b2rt_launch_count(b2rt_ctx *ctx, int64_t *n, int reset); /* kernels launched by this ctx */
void *b2rt_stream(b2rt_ctx *ctx);               /* the cudaStream_t all work is issued on */

#ifdef __cplusplus
}
#endif
#endif
