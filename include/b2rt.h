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
enum { B2RT_LIGHT_POINT = 0, B2RT_LIGHT_DISK_AREA = 1, B2RT_LIGHT_SPOT = 2, B2RT_LIGHT_DISTANT = 3 };

typedef struct {
    int32_t n_prims, n_tris, n_wide_nodes, n_leaf_slots;
    int32_t max_depth, wide_width;
    float   bounds[6];          /* scene AABB lo.xyz hi.xyz */
    float   sah_cost;           /* SAH cost of the wide tree (node cost 1, prim cost 1) */
    float   build_ms;           /* device time of the last b2rt_build_accel: bounds, Morton, sort, topology, refit,
                                   wide collapse (scene already resident) */
    float   upload_ms;          /* host->device copy of the flattened scene + buffer allocation before it */
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
/* The same with orientation flags.  N (3 floats per vertex, WORLD space: the caller applies the mesh's object-to-world
 * normal transform) and uv (2 per vertex) are pbrt's "N" and "uv" / "st" mesh parameters: a hit then carries the
 * interpolated shading normal and the (dpdu, dpdv) frame of the uv parametrisation exactly as Triangle::Intersect +
 * Triangle::GetShadingGeometry + BSDF::BSDF form them (pbrt-v2/src/shapes/trianglemesh.cpp:181-201, 311-389; reference
 * device code cuda_render/util/shape/cudatrianglemesh.cu:25-78). */
enum { B2RT_MESH_REVERSE = 1 };   /* pbrt ReverseOrientation ^ TransformSwapsHandedness: flips the normals (core/diffgeom.cpp:52-54) */
int b2rt_add_mesh_ex(b2rt_ctx *ctx, int32_t nverts, const float *P, int32_t ntris, const int32_t *idx,
                     const float *N, const float *uv, int32_t material_id, int32_t arealight_id, int32_t flags,
                     int32_t *first_prim);
/* o2w, w2o: row-major 3x4 object->world and its inverse exactly as pbrt holds them
 * (cuda_render/util/shape/cudasphere.cpp:15-40); w2o may be NULL (then it is computed here). */
int b2rt_add_sphere(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float radius, int32_t material_id,
                    int32_t arealight_id, int32_t *prim);
/* partial sphere: pbrt's "zmin" / "zmax" / "phimax" (shapes/sphere.cpp:40-49, 84-106; the reference's device program ignores
 * all three, cudasphere.cu:27-72): the surface with object-space z outside [zmin, zmax] or azimuth beyond phimax_deg is cut
 * away, a ray that meets the cut-away part first continues to the far intersection.  Values are clamped as pbrt does.  The
 * azimuth test uses the side of the cut's direction instead of atan2f (not reproducible between math libraries): equal to
 * pbrt's except for points within rounding of the cut.  b2rt_add_disk_ex: the same "phimax" for a disk (shapes/disk.cpp:
 * 73-76); not for area lights. */
int b2rt_add_sphere_ex(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float radius, float zmin, float zmax,
                       float phimax_deg, int32_t material_id, int32_t arealight_id, int32_t *prim);
int b2rt_add_disk_ex(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float height, float radius, float inner_radius,
                     float phimax_deg, int32_t material_id, int32_t arealight_id, int32_t *prim);
/* disk at object z = height (cuda_render/util/shape/cudadisk.cpp:15-45) */
int b2rt_add_disk(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float height, float radius,
                  float inner_radius, int32_t material_id, int32_t arealight_id, int32_t *prim);
int b2rt_add_material(b2rt_ctx *ctx, int32_t type, const float rgb_kd_or_kr[3], const float rgb_kt[3],
                      float ior, int32_t *material_id);
/* point: p = position, rgb = I.  disk area: rgb = L, prim = the disk primitive, nsamples. */
int b2rt_add_light(b2rt_ctx *ctx, int32_t type, const float p[3], const float rgb[3], int32_t prim,
                   int32_t nsamples, int32_t *light_id);
/* pbrt SpotLight (lights/spot.cpp:40-121; the reference's cudalight.cpp:61-71 only warns): l2w / w2l = the light's
 * LightToWorld and its inverse (row-major 3x4; the light looks along its +z), rgb = I, cos_total_width / cos_falloff_start =
 * the two cosines pbrt derives from "coneangle" and "conedeltaangle".  pbrt DistantLight (lights/distant.cpp:40-105):
 * dir_to_light = Normalize(LightToWorld(from - to)), rgb = L; its emission disk follows the scene bounds at build time. */
int b2rt_add_spot_light(b2rt_ctx *ctx, const float l2w[12], const float w2l[12], const float rgb[3], float cos_total_width,
                        float cos_falloff_start, int32_t *light_id);
int b2rt_add_distant_light(b2rt_ctx *ctx, const float dir_to_light[3], const float rgb[3], int32_t *light_id);
int b2rt_clear_scene(b2rt_ctx *ctx);

/* ---- object instancing without flattening: replaces CudaRender::assembleObject / objectInstance, which hang every pbrt
 * ObjectInstance under an OptiX Transform node over the object's own Group (cuda_render/cudarender.cpp:38-103).  Here: a
 * two-level BVH.  The shapes added between b2rt_object_begin and b2rt_object_end form an OBJECT with a tree of its own,
 * built once in the space the shapes are given in (pbrt: the CTM inside ObjectBegin / ObjectEnd); primitive ids returned by
 * the b2rt_add_* calls in between are local to the object, area lights are refused (as pbrt does, core/api.cpp:1009-1011),
 * material ids refer to the context's material table.  b2rt_add_instance places the object under the instance-to-world
 * transform o2w (w2o = its inverse as pbrt holds it; NULL: computed here): nothing is copied, rays are taken into the
 * instance's space as pbrt's TransformedPrimitive::Intersect does (core/primitive.cpp:95-120: origin as a point,
 * direction as a vector, not renormalised, so t carries over) and the hit attributes come back to world space the same way.
 * Hit ids: the items of the top-level tree (primitives and instances, in submission order) take the ids [0, n_items); the
 * primitives of instance k then take [first_prim_k, first_prim_k + n_prims of its object), in instance order after
 * n_items -- b2rt_instance_info after b2rt_build_accel.  Objects do not nest (pbrt: core/api.cpp:1107-1110, 1143-1146). */
int b2rt_object_begin(b2rt_ctx *ctx, int32_t *object_id);
int b2rt_object_end(b2rt_ctx *ctx);
int b2rt_add_instance(b2rt_ctx *ctx, int32_t object_id, const float o2w[12], const float *w2o, int32_t *instance_id);
int b2rt_instance_info(b2rt_ctx *ctx, int32_t instance_id, int32_t *first_prim, int32_t *n_prims, int32_t *object_id);

/* ---- binary scene cache (the reference has none: every run goes through pbrt's text parser).
 * b2rt_scene_save writes everything staged so far -- meshes with their attributes, spheres / disks, objects and instances,
 * materials, lights -- to one file; b2rt_scene_load replaces the staged scene by the file's (then b2rt_build_accel as
 * usual).  The trees are not stored: the build is deterministic and takes milliseconds.  A file written by a library
 * build with other struct layouts is refused. */
int b2rt_scene_save(b2rt_ctx *ctx, const char *path);
int b2rt_scene_load(b2rt_ctx *ctx, const char *path);
/* what is staged: counts = {top-level items (primitives + instances), lights, materials, objects, instances} */
int b2rt_scene_counts(b2rt_ctx *ctx, int32_t counts[5]);

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


/* parity hook for the hit attributes (K6): closest hit of every ray + what the shade stage derives from it -- replaces the
 * attribute half of trianglemesh_intersect / sphere_intersect / disk_intersect (cuda_render/util/shape/cudatrianglemesh.cu:
 * 25-78).  prim = -1 on a miss (the other fields are then 0).  ng = geometric normal dg.nn, n / sn / tn = the BSDF frame
 * (nn, sn, tn of core/reflection.cpp:601-609) built from the shading geometry. */
typedef struct { float p[3]; float t; float ng[3]; float u; float n[3]; float v; float sn[3]; float ray_eps; float tn[3];
                 int32_t prim; } b2rt_surface;     /* 80 bytes */
int b2rt_surface_at(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, b2rt_surface *out, int where);

/* ---- photon pass: replaces photontracing_camera / photontracing_closest_hit (cuda_render/photon_mapping/
 * photontracing.cu:80-185) and the host-side CreatePhotonMap kd-tree build (photonmappingrenderer.cpp:150-180) -- */
enum { B2RT_MAP_CAUSTIC = 0, B2RT_MAP_INDIRECT = 1,
       B2RT_MAP_DIRECT = 2,      /* final gathering only: photons of a path's FIRST diffuse hit (photonmap.cpp:468-472) */
       B2RT_MAP_RADIANCE = 3 };  /* final gathering only: radiance photons; wi_oct holds the surface normal, alpha the
                                    reflectance rho_r until b2rt_radiance_compute turns it into the outgoing radiance Lo */
/* 32-byte photon (the reference's CudaPhoton is 40 B, photon_mapping/photonmapping.h:32-40).
 * wi_oct: incident direction, octahedral 2 x snorm16 (low half x, high half y); path = emitting path index. */
typedef struct { float p[3]; uint32_t wi_oct; float alpha[3]; uint32_t path; } b2rt_photon;
typedef struct {
    float   origin[3];        /* lower corner of the grid = min photon position per axis */
    float   cell_size, inv_cell;
    int32_t dims[3];
    int32_t n_cells;
    int64_t n_photons;
    float   build_ms;
} b2rt_grid_info;

/* Trace paths [first_path, first_path+n_paths) from the lights (counter-based RNG: the photons of a path depend only
 * on (seed, path index)), append the deposits in path order to the caustic / indirect maps.  Deposit, bounce and
 * Russian-roulette rules follow pbrt's PhotonShootingTask (integrators/photonmap.cpp:416-514), finalgather off.
 * flags: bit0 = collect caustic photons, bit1 = collect indirect photons.  n_stored[2] (may be NULL) receives the
 * number of photons this call appended to each map. */
int b2rt_photons_trace(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int32_t max_depth, int32_t flags,
                       int64_t n_stored[2]);
/* pbrt's stopping rule (photonmap.cpp:521-576): shoot batches of `batch_paths` until each map holds at least the
 * wanted number of photons; a map stops collecting (and stops counting paths) once it is full.  pbrt's give-up rule
 * (:182-184, :526-537) too: after 500 000 paths a wanted map that holds fewer than 4 photons ends the shooting, ALL maps
 * are erased and the call still returns B2RT_OK -- n_stored_out is {0, 0} and b2rt_last_error() has the message; the
 * frame then renders without photon maps, which is what pbrt does (e.g. a diffuse-only scene with causticphotons > 0). */
int b2rt_photons_shoot(b2rt_ctx *ctx, int64_t n_caustic_wanted, int64_t n_indirect_wanted, int32_t max_depth,
                       int64_t batch_paths, int64_t n_paths_out[2], int64_t n_stored_out[2]);
int b2rt_photons_clear(b2rt_ctx *ctx);
int b2rt_photons_count(b2rt_ctx *ctx, int32_t map, int64_t *n_photons, int64_t *n_paths);
/* copy the photons of a map (path order) to the host / replace them from the host (checkpoint, tests, and the
 * receive side of the multi-GPU allgather).  where = B2RT_HOST or B2RT_DEVICE. */
int b2rt_photons_download(b2rt_ctx *ctx, int32_t map, b2rt_photon *out, int64_t capacity, int where);
int b2rt_photons_upload(b2rt_ctx *ctx, int32_t map, const b2rt_photon *in, int64_t n, int64_t n_paths, int where);
void *b2rt_photons_device_ptr(b2rt_ctx *ctx, int32_t map);   /* path-order array on the device (NCCL send buffer) */

/* hash-grid build (cell-key sort + prefix-sum cell ranges).  cell_size <= 0: chosen from the photon density so that
 * an occupied cell holds about `target_per_cell` photons (<= 0: 8).  Bit-exact definition in DESIGN.md. */
int b2rt_photon_grid_build(b2rt_ctx *ctx, int32_t map, float cell_size, float target_per_cell);
int b2rt_photon_grid_info(b2rt_ctx *ctx, int32_t map, b2rt_grid_info *info);
/* parity hooks: sorted cell keys[n], sorted->path-order index[n], cell_start[n_cells+1], sorted photons[n] */
int b2rt_photon_grid_download(b2rt_ctx *ctx, int32_t map, uint32_t *keys, uint32_t *order, uint32_t *cell_start,
                              b2rt_photon *sorted);

/* ---- radiance estimate: replaces KdTreeLookup / photonGatheringCamera (photon_mapping/gathering.cu:25-126) with
 * pbrt's estimator LPhoton (integrators/photonmap.cpp:214-260): k nearest photons within sqrt(max_dist2), Simpson
 * kernel, k/(nPaths*maxDist2) normalisation.  k <= 0 selects the fixed-radius mode (all photons within the radius).
 * query: 48 bytes {p[3], n[3] (faceforwarded shading normal Nf), kd[3] (rho of the diffuse lobe), 3 pad}. */
typedef struct { float p[3]; float n[3]; float kd[3]; float pad[3]; } b2rt_query;
int b2rt_gather(b2rt_ctx *ctx, int32_t map, int64_t nqueries, const b2rt_query *queries, int32_t k, float max_dist2,
                float *out_rgb, float *out_radius2, int32_t *out_found, int where);
/* parity hook: the selected photon set per query (indices into the SORTED photon array, ascending (d2, index));
 * idx_out[nqueries*k], -1 padded */
int b2rt_gather_sets(b2rt_ctx *ctx, int32_t map, int64_t nqueries, const b2rt_query *queries, int32_t k,
                     float max_dist2, int32_t *idx_out, int where);

/* ---- final gathering: replaces nothing in cuda_render (its finalGatheringCamera, photon_mapping/gathering.cu:129-146,
 * only averages the k-nearest estimate); follows pbrt's PhotonIntegrator with finalgather true, the integrator's
 * default (pbrt-v2/src/integrators/photonmap.cpp:468-490 radiance-photon creation, :585-621 ComputeRadianceTask,
 * :263-283 EPhoton, :156-172 RadiancePhotonProcess, :645-758 the gather itself).  SURVEY.md 8(f) rank 3. ---------- */
/* Re-trace paths [first_path, first_path + n_paths) (same Philox streams, hence the same paths as b2rt_photons_trace)
 * and append (a) the direct photon of each path to B2RT_MAP_DIRECT and (b) a radiance photon at every deposit
 * (direct, indirect or caustic) that draws u < 0.125 from the path's own stream to B2RT_MAP_RADIANCE, both in path
 * order.  n_stored[0] = direct photons, n_stored[1] = radiance photons appended (may be NULL). */
int b2rt_photons_trace_fg(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int32_t max_depth, int64_t n_stored[2]);
/* pbrt only creates a radiance photon (and a direct photon) where a photon was actually DEPOSITED (depositedPhoton,
 * photonmap.cpp:455-490): a caustic-type hit (all-specular path, > 1 intersection) counts while the caustic map is still
 * collecting, i.e. for paths < caustic_paths; a direct / indirect hit for paths < indirect_paths.  The two numbers are
 * n_paths_out of b2rt_photons_shoot; pass them before b2rt_photons_trace_fg over [0, max of the two).  Negative = no limit
 * (the default: every matte hit of the traced paths counts). */
int b2rt_fg_deposit_limits(b2rt_ctx *ctx, int64_t caustic_paths, int64_t indirect_paths);
/* EPhoton: irradiance at (p, n) from the k nearest photons of `map` within sqrt(max_dist2):
 * sum of alpha over the photons with Dot(n, wi) > 0, divided by (nPaths * md2 * pi).  out_E[3 * nqueries]. */
int b2rt_irradiance(b2rt_ctx *ctx, int32_t map, int64_t nqueries, const b2rt_query *queries, int32_t k, float max_dist2,
                    float *out_E, int where);
/* ComputeRadianceTask: Lo = (INV_PI * rho_r) * (E_direct + E_indirect + E_caustic) for every radiance photon, from the
 * maps whose grid is built, then the hash grid over the radiance photons.  Call once after the maps are complete. */
int b2rt_radiance_compute(b2rt_ctx *ctx, int32_t nused, float max_dist2);
/* RadiancePhotonProcess with an unbounded search radius: the nearest radiance photon (ties: lower sorted index) whose
 * normal satisfies Dot(rp.n, n) > 0.  out_rgb[3 * nqueries] = its Lo (0 if none), out_idx[nqueries] = its index in the
 * SORTED radiance-photon array or -1; either may be NULL.  Only p and n of the queries are read. */
int b2rt_radiance_lookup(b2rt_ctx *ctx, int64_t nqueries, const b2rt_query *queries, float *out_rgb, int32_t *out_idx,
                         int where);
/* b2rt_render_frame's indirect term from now on: enable = 0, LPhoton on the indirect map (finalgather false);
 * enable = 1, one-bounce final gathering with gather_samples rays per shading point (pbrt finalgathersamples, default
 * 32: half BSDF-sampled, half sampled in cones of gather_angle_deg degrees, pbrt gatherangle default 10, around the
 * directions of the 50 nearest indirect photons, combined with the power heuristic). */
int b2rt_final_gather_set(b2rt_ctx *ctx, int32_t enable, int32_t gather_samples, float gather_angle_deg);

/* parity hook: the gather of b2rt_render_frame for caller-supplied shading points.  kd = rho * path throughput; the
 * queries' pad carries what the camera pass leaves there: pad[0] = ray epsilon of the hit, pad[1] = bits of the pixel id,
 * pad[2] = bits of (sample number | specular-tree node << 22) -- the Philox counter words of the point's gather rays.
 * Needs the indirect map's grid and b2rt_radiance_compute.  out_rgb[3 * nqueries]. */
int b2rt_final_gather(b2rt_ctx *ctx, int64_t nqueries, const b2rt_query *queries, float max_dist2, float *out_rgb, int where);

/* ---- camera pass: replaces the four launches of PhotonMappingRenderer::render / SimpleRenderer::render
 * (cuda_render/photon_mapping/photonmappingrenderer.cpp:31-45,108-139,228-277; simple_render/simplerender.cpp:18-103)
 * including the CPU camera-ray loop (util/camera/pbrtcamera.cpp:57-122) and the CPU film loop. ------------------- */
/* pinhole camera: ray origin o, direction Normalize(A + x*B + y*C) for continuous raster coordinates (x, y) --
 * exactly the rays pbrt's PerspectiveCamera::GenerateRay produces with lensradius 0 (cameras/perspective.cpp:81-114) */
typedef struct { float o[3]; float A[3]; float B[3]; float C[3]; int32_t xres, yres; } b2rt_camera;
typedef struct {
    int32_t spp;                 /* samples per pixel (box filter, radius 0.5) */
    int32_t max_specular_depth;  /* pbrt maxspeculardepth (default 5) */
    int32_t nused;               /* photons per estimate (pbrt nused, default 50); <= 0: fixed radius */
    float   max_dist2;           /* pbrt maxdist squared */
    int32_t use_caustic, use_indirect, direct_lighting;   /* which terms of PhotonIntegrator::Li to evaluate */
    int32_t x0, x1, y0, y1;      /* crop window in pixels [x0,x1) x [y0,y1): the image tile this call renders */
} b2rt_frame_params;
/* out_rgb: (x1-x0)*(y1-y0)*3 floats, row-major within the crop window.  Photon maps must have been shot and their
 * grids built if use_caustic / use_indirect are set (maps that are empty or unbuilt contribute nothing). */
int b2rt_render_frame(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, int where);

/* parity hook for the camera pass proper (K14 camera rays, K6 attributes, K7 direct lighting with MIS, the specular tree):
 * per camera sample of the crop window (pixel-major, spp samples each) the radiance render_li_kernel computes before the
 * photon-map terms are added -- Le + UniformSampleAllLights + both specular branches (pbrt PhotonIntegrator::Li,
 * integrators/photonmap.cpp:624-638, core/integrator.cpp:47-258; reference: raytracing_closest_hit + directLight,
 * cuda_render/photon_mapping/raytracing.cu:49-128) -- the number of diffuse leaves of its specular tree and, if
 * out_queries != NULL, the leaves themselves (*slots_out per sample) as the gather / final-gather stage receives them.
 * Host buffers.  Bit-exact against oracle/oracle_render.h port_camera_samples. */
int b2rt_camera_samples(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, uint32_t *out_nleaves,
                        b2rt_query *out_queries, int32_t *slots_out);

/* ---- image-plane samples: replaces the CPU sample loop PbrtCamera::preLaunch (cuda_render/util/camera/pbrtcamera.cpp:
 * 57-122), which pulls every CameraSample out of pbrt's Sampler and uploads 72 bytes per ray.  The pattern of
 * b2rt_render_frame's samples inside each pixel: pbrt "random" (the default), "stratified" (xsamples * ysamples = spp
 * jittered strata, core/montecarlo.cpp:158-168) or "lowdiscrepancy" (a scrambled (0,2)-sequence per pixel,
 * core/montecarlo.h:272-295,317-325).  Randomness is Philox(seed, pixel, sample): tiling-independent. ---- */
enum { B2RT_SAMPLER_RANDOM = 0, B2RT_SAMPLER_STRATIFIED = 1, B2RT_SAMPLER_LOWDISCREPANCY = 2 };
int b2rt_sampler_set(b2rt_ctx *ctx, int32_t type, int32_t xsamples, int32_t ysamples);
/* parity hook: the continuous raster positions (imageX, imageY) of the samples of window = {x0, y0, x1, y1} (cells may
 * lie up to 64 pixels outside the xres x yres image, as a filtered frame's halo does); xy[cells * spp * 2], cell-major */
int b2rt_sample_positions(b2rt_ctx *ctx, int32_t xres, int32_t yres, const int32_t window[4], int32_t spp, float *xy,
                          int where);

/* ---- film: replaces the reference's CPU film loop, camera->film->AddSample for every launch sample
 * (cuda_render/photon_mapping/photonmappingrenderer.cpp:247-273, simple_render/simplerender.cpp:84-100), i.e. pbrt's
 * ImageFilm::AddSample + the normalisation in ImageFilm::WriteImage (pbrt-v2/src/film/image.cpp:77-137,178-213) with
 * the pixel filters of pbrt-v2/src/filters: same 16x16 weight table, same extent arithmetic, XYZ accumulation. ---- */
enum { B2RT_FILTER_BOX = 0, B2RT_FILTER_TRIANGLE = 1, B2RT_FILTER_GAUSSIAN = 2, B2RT_FILTER_MITCHELL = 3,
       B2RT_FILTER_SINC = 4 };
/* p0, p1: gaussian alpha | mitchell B, C | sinc tau (pbrt's parameter names); widths in pixels, (0, 16] */
typedef struct { int32_t type; float xwidth, ywidth, p0, p1; } b2rt_filter;
/* the filter b2rt_render_frame reconstructs with from now on (NULL: pbrt's default, box 0.5 x 0.5).  With a support
 * wider than half a pixel the frame also traces the sample cells around the crop window that reach into it
 * (ImageFilm::GetSampleExtent, film/image.cpp:155-166), so image bands stay bit-identical to the whole frame. */
int b2rt_film_set_filter(b2rt_ctx *ctx, const b2rt_filter *filter);
/* stand-alone reconstruction: cells = {cx0, cy0, cx1, cy1} sample cells (cell (cx, cy) holds spp samples whose raster
 * position xy lies in [cx, cx+1) x [cy, cy+1)), samples cell-major: xy[2*s], rgb[3*s], s = cell * spp + k;
 * pixels = {x0, y0, x1, y1} = the film's crop window; out_rgb (x1-x0)*(y1-y0)*3 row-major.  A pixel sums its samples
 * in (cell row, cell column, k) order: deterministic, and equal to pbrt's film fed in that order bit for bit. */
int b2rt_film_reconstruct(b2rt_ctx *ctx, const b2rt_filter *filter, const int32_t cells[4], int32_t spp, const float *xy,
                          const float *rgb, const int32_t pixels[4], float *out_rgb, int where);

/* ---- several GPUs of one box behind the same single-threaded boundary (the reference has no multi-GPU
 * path, one implicit device: cuda_render/cudarender.cpp:17-36).  One process, one context per device, one worker thread
 * per device inside the library.  Results are bit-identical to the single-GPU calls for any GPU count:
 *   - photon passes: the path range is cut into one contiguous slice per GPU; the compaction kernel of each GPU stores
 *     its photons at their final position in the photon arrays of EVERY GPU (peer stores over NVLink: compute and exchange
 *     in one kernel).  B2RT_MGPU_EXCHANGE=nccl selects a grouped ncclBroadcast per slice instead (libnccl.so.2 is dlopen'ed),
 *     =copy one cudaMemcpyPeerAsync per slice (also the fallback without peer access).
 *   - grids, radiance photons: replicated (deterministic, same bytes everywhere).
 *   - camera pass: 64 x 64 image tiles, tile t on GPU t mod n, each GPU renders its tiles in one pass and stores the
 *     pixels straight into the frame on GPU 0; pixel filters wider than a pixel use one row band per GPU instead.
 * Scene, sampler, filter and final-gather settings are per context: call b2rt_add_* / b2rt_add_material / ... on
 * b2rt_multi_ctx(m, i) for every i (same calls, same order), or the b2rt_multi_*_set helpers. */
typedef struct b2rt_multi b2rt_multi;
typedef struct {
    float   trace_ms, reserve_ms, exchange_ms;   /* wall clock of the three steps of the sharded photon passes (summed) */
    float   grid_ms, render_ms;
    int64_t exchange_bytes;                      /* bytes of photons every GPU holds after the passes (all slices) */
    int32_t passes;
} b2rt_multi_stats;
int  b2rt_multi_create(int ngpus, const int *devices /* NULL: 0 .. ngpus-1 */, uint64_t seed, b2rt_multi **out);
/* grow the group: the new contexts receive a copy of context 0's staged scene, materials, lights and settings, so the
 * scene is ingested once even when the GPU count is learnt late; call b2rt_multi_build_accel afterwards */
int  b2rt_multi_resize(b2rt_multi *m, int ngpus, const int *devices);
void b2rt_multi_destroy(b2rt_multi *m);
const char *b2rt_multi_last_error(b2rt_multi *m);
int  b2rt_multi_count(b2rt_multi *m);
b2rt_ctx *b2rt_multi_ctx(b2rt_multi *m, int i);
/* peer_access: 1 if every pair of the devices has peer access; exchange_mode: 0 fused peer stores, 1 NCCL, 2 peer copies,
 * 3 push kernel (B2RT_MGPU_EXCHANGE=push: local compaction, then one coalesced copy kernel stores the slice into the peers) */
int  b2rt_multi_info(b2rt_multi *m, int32_t *peer_access, int32_t *exchange_mode);
int  b2rt_multi_stats_get(b2rt_multi *m, b2rt_multi_stats *out, int reset);
int  b2rt_multi_build_accel(b2rt_multi *m);                     /* every device builds the same deterministic BVH */
int  b2rt_multi_photons_clear(b2rt_multi *m);
/* as b2rt_photons_trace / _shoot / _trace_fg; afterwards EVERY context holds the complete maps */
int  b2rt_multi_photons_trace(b2rt_multi *m, int64_t first_path, int64_t n_paths, int32_t max_depth, int32_t flags,
                              int64_t n_stored[2]);
int  b2rt_multi_photons_shoot(b2rt_multi *m, int64_t n_caustic_wanted, int64_t n_indirect_wanted, int32_t max_depth,
                              int64_t batch_paths, int64_t n_paths_out[2], int64_t n_stored_out[2]);
int  b2rt_multi_photons_trace_fg(b2rt_multi *m, int64_t first_path, int64_t n_paths, int32_t max_depth, int64_t n_stored[2]);
int  b2rt_multi_photon_grid_build(b2rt_multi *m, int32_t map, float cell_size, float target_per_cell);
int  b2rt_multi_radiance_compute(b2rt_multi *m, int32_t nused, float max_dist2);
int  b2rt_multi_fg_deposit_limits(b2rt_multi *m, int64_t caustic_paths, int64_t indirect_paths);
int  b2rt_multi_final_gather_set(b2rt_multi *m, int32_t enable, int32_t gather_samples, float gather_angle_deg);
int  b2rt_multi_sampler_set(b2rt_multi *m, int32_t type, int32_t xsamples, int32_t ysamples);
int  b2rt_multi_film_set_filter(b2rt_multi *m, const b2rt_filter *filter);
/* as b2rt_render_frame; where = B2RT_HOST, or B2RT_DEVICE for a buffer on the FIRST device of the group */
int  b2rt_multi_render_frame(b2rt_multi *m, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, int where);

/* traversal work counters: wide nodes fetched and primitive records tested by b2rt_trace_* since the last
 * reset -- L2 -> SM bytes per ray = (nodes * 128 + prims * 48) / rays.  Only in the -DB2RT_STATS build (libb2rt_stats.so,
 * `make -C cuda_raytrace_b200/csrc stats`; tools/trace_stats.py); the regular library returns B2RT_E_STATE. */
int b2rt_trace_stats(b2rt_ctx *ctx, uint64_t *nodes, uint64_t *prims, int reset);

/* generic device sort exposed for tests/bench of the hand-written radix sort (K2):
 * stable LSD sort of (key32,value32) pairs on `key_bits` low bits; device pointers only. */
int b2rt_sort_pairs(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, int key_bits);
/* The same sort for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  No counterpart in the reference.  Allocates nothing: the ping-pong
 * arrays, histograms / look-back states and scan temporaries live in a static device array inside the library (9.1 MB that
 * every GPU that loads libb2rt.so holds until the process exits), so n is limited to B2RT_SORT_NOALLOC_MAX pairs -- larger
 * n returns B2RT_E_ARG with a message.  Same kernels and schedule as b2rt_sort_pairs, hence the same bytes; device pointers
 * only; asynchronous on the context stream.  All calls on one device share the array: calls that overlap on the GPU
 * (two contexts / streams of one device) race, and the array keeps the previous call's contents (the call clears what it
 * needs at zero itself).  n = 0 does the once-per-device set-up only (cudaGetSymbolAddress of the array): call it once
 * before a graph capture. */
#define B2RT_SORT_NOALLOC_MAX (1 << 20)
int b2rt_sort_pairs_noalloc(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, int key_bits);

/* ---- pinned host buffers ---------------------------------------------------------------------- */
/* Page-locked host memory for the B2RT_HOST calls (rays in, hits / images out): with pageable memory the driver
 * stages every copy and the chunked upload / trace / download pipeline of b2rt_trace_* cannot overlap.  The reference
 * maps its output buffers instead (photonmappingrenderer.cpp:247-251, bOutput->map()).  Returns NULL on failure. */
void *b2rt_host_alloc(b2rt_ctx *ctx, int64_t bytes);
void  b2rt_host_free(b2rt_ctx *ctx, void *p);

/* ---- per-stage device times and NVTX ranges (the reference has no timing anywhere on its GPU path) -----
 * Every stage entry point below opens an NVTX range named after itself (visible in Nsight Systems / Compute; free when no
 * tool is attached).  With timings enabled (b2rt_timings_enable, or B2RT_TIMINGS=1 in the environment at b2rt_create) it
 * also brackets the stage with CUDA events on the context stream and SYNCHRONISES at its end -- a measuring mode, not for
 * production pipelines -- and b2rt_get_timings returns the accumulated device milliseconds and call counts per stage. */
enum { B2RT_STAGE_BUILD = 0,        /* b2rt_build_accel */
       B2RT_STAGE_TRACE_CLOSEST,    /* b2rt_trace_closest */
       B2RT_STAGE_TRACE_ANY,        /* b2rt_trace_any */
       B2RT_STAGE_PHOTONS,          /* b2rt_photons_trace / _shoot / _trace_fg */
       B2RT_STAGE_GRID,             /* b2rt_photon_grid_build, b2rt_radiance_compute */
       B2RT_STAGE_GATHER,           /* b2rt_gather / _gather_sets / _irradiance / _radiance_lookup / _final_gather */
       B2RT_STAGE_RENDER,           /* b2rt_render_frame, b2rt_camera_samples, b2rt_surface_at */
       B2RT_STAGE_FILM,             /* b2rt_film_reconstruct */
       B2RT_STAGE_COUNT };
typedef struct { double ms[B2RT_STAGE_COUNT]; int64_t calls[B2RT_STAGE_COUNT]; } b2rt_timings;
int b2rt_timings_enable(b2rt_ctx *ctx, int on);
int b2rt_get_timings(b2rt_ctx *ctx, b2rt_timings *out, int reset);

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
