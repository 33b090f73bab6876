// Internal declarations shared by the b2rt translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "b2rt.h"
#include <nvtx3/nvToolsExt.h>

#define B2RT_SM_COUNT_DEFAULT 148

struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    // owns its allocation: `delete ctx` (b2rt_destroy, with the context's device current) frees every buffer, including
    // the ones added later -- an explicit release list in b2rt_destroy went stale once (the final-gather buffers leaked)
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : p(o.p), cap(o.cap) { o.p = nullptr; o.cap = 0; }
    DevBuf &operator=(DevBuf &&o) noexcept {
        if (this != &o) { release(); p = o.p; cap = o.cap; o.p = nullptr; o.cap = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    // grow-only device buffer; returns cudaError
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *)p; }
};

// 128-byte 4-wide BVH node, SoA inside the node so each field is one 16-byte load.
//   lox hix loy hiy loz hiz : 4 children each
//   child[4]  : >=0 internal node index; <0: leaf = 0x80000000 | (first_slot << 3) | (count-1);
//               B2RT_CHILD_EMPTY for an unused lane (its box is inverted so it is never entered)
//   meta[4]   : meta[0] = number of used children, meta[1] = binary-node id it came from
struct __align__(128) WideNode {
    float lox[4], hix[4], loy[4], hiy[4], loz[4], hiz[4];
    int32_t child[4];
    int32_t meta[4];
};
static_assert(sizeof(WideNode) == 128, "wide node must be one cache line");
#define B2RT_CHILD_EMPTY 0x7fffffff
#define B2RT_LEAF_MAX 4          // prims per leaf (pbrt default maxnodeprims 4, accelerators/bvh.cpp:494)
#define B2RT_PRIM_TRI 0
#define B2RT_PRIM_SPHERE 1
#define B2RT_PRIM_DISK 2
#define B2RT_PRIM_INSTANCE 3     // top-level tree only: one record per object instance (two-level BVH), see InstanceRec

// Leaf-ordered primitive record, 48 B = 3 x float4:
//   a = {p1.xyz, bits(prim id)}   b = {e1.xyz, bits(type)}   c = {e2.xyz, bits(analytic index)}
// e1 = p2 - p1, e2 = p3 - p1 are the SAME binary32 values pbrt computes per test
// (pbrt-v2/src/shapes/trianglemesh.cpp:155-156), hoisted to build time.
struct PrimRec { float4 a, b, c; };

// analytic primitives (spheres / disks), world<->object as 3x4 row-major
struct AnalyticPrim {
    float o2w[12], w2o[12];
    float radius, height, inner, pad;
};

// One instance of an object (pbrt ObjectInstance -> TransformedPrimitive, core/primitive.cpp:95-120; the reference's OptiX
// Transform node over a Group, cuda_render/cudarender.cpp:88-103).  The object's own BVH (built once, in the object's space)
// is shared by all its instances; nothing is flattened.  An instance occupies the global primitive ids
// [prim_base, prim_base + n_prims): prim_base + the primitive's id inside the object.
// Its top-level leaf record: a = {lo32(ptr), hi32(ptr), 0, bits(prim_base)}, b.w = B2RT_PRIM_INSTANCE, c.w = instance index;
// ptr = device address of this record, so the traversal kernels need no extra argument.
struct DScene;
struct __align__(16) InstanceRec {
    float o2w[12], w2o[12];          // instance-to-world and its inverse, row-major 3x4, exactly as pbrt holds them
    const WideNode *nodes;           // the object's tree / records / analytic table (device pointers)
    const PrimRec *recs;
    const AnalyticPrim *analytic;
    const DScene *object;            // the object's scene view for the shade stage (attributes, per-prim materials)
    int32_t prim_base, n_prims, object_id, pad;
    float box[6];                    // the object's AABB in its own space (lo.xyz, hi.xyz)
};
static_assert(sizeof(InstanceRec) % 16 == 0, "instance_ray reads the transform rows as float4");

struct Material { int32_t type; float kd[3]; float kt[3]; float ior; };
// analytic/area/n are derived at build time for disk area lights: analytic = index into the analytic table,
// area = Disk::Area() (shapes/disk.cpp:132-135), n = Normalize(ObjectToWorld(Normal(0,0,1))) (disk.cpp:154)
// spot: p = position, l2w / w2l = rotation parts of LightToWorld / WorldToLight, the two cone cosines.  distant: n = direction
// TO the light, wc / wr = bounding sphere of the scene box (BBox::BoundingSphere), filled in when the tables are uploaded.
struct Light { int32_t type; float p[3]; float rgb[3]; int32_t prim; int32_t nsamples; int32_t analytic; float area; float n[3];
               float l2w[9], w2l[9]; float cos_total, cos_falloff; float wc[3], wr; };

// ---- device view of the scene -------------------------------------------------------------------------------------
struct DScene {
    const WideNode *nodes;
    const PrimRec *recs;            // leaf-slot order
    const AnalyticPrim *analytic;
    const int32_t *prim_material;   // per global prim id, -1 = none (default grey matte 0.5, cudamaterial.cpp:20)
    const int32_t *prim_light;      // per global prim id, -1 = not an area light
    const Material *materials;
    const Light *lights;
    const float *light_cdf;         // n_lights+1 entries, cdf[0]=0, cdf[n]=1; then n_lights pdfs
    int32_t n_lights, n_materials;
    // per-vertex shading attributes (K6).  tri_flags == NULL: no mesh of the scene has any (the common case; make_surface
    // then never touches the rest).  prim_kind[prim] = triangle index; tri_flags[tri]: bit 0 normals, bit 1 uv, bit 2 flip.
    const uint8_t *tri_flags;
    const int32_t *prim_kind, *tri_idx;
    const float *verts, *vnormal, *vuv;
    // object instances (two-level BVH): NULL / 0 without instances.  A hit with HitD::inst >= 0 lies in instances[inst];
    // its attributes come from instances[inst].object (the object's own scene view), transformed to world space.
    const InstanceRec *instances;
    int32_t n_instances;
};

// one photon map (caustic or indirect): photons in path order as traced, then the hash-grid built over them
struct PhotonMap {
    DevBuf photons;          // b2rt_photon[n], path order (bit-identical for any GPU count)
    int64_t n = 0;
    int64_t n_paths = 0;     // paths emitted while this map was still collecting (pbrt nCausticPaths / nIndirectPaths)
    // grid
    bool grid_built = false;
    uint64_t grid_epoch = 0; // bumped by every grid build (tables derived from the grid check it)
    DevBuf sorted;           // b2rt_photon[n] in (cell key, path-order index) order
    DevBuf cell_start;       // uint32[n_cells+1]
    DevBuf keys, order;      // uint32[n] sorted cell keys / original indices (kept for parity downloads)
    DevBuf pos4;             // float4[n] {x, y, z, bits(sorted index)}: what the gather's distance tests read
    DevBuf coarse;           // uint32[cdims.x*cdims.y*cdims.z] photons per coarse cell (4^3 fine cells)
    int32_t cdims[3] = {0, 0, 0};
    b2rt_grid_info grid{};
};

struct b2rt_ctx {
    int device = 0;
    int sm_count = B2RT_SM_COUNT_DEFAULT;
    uint64_t seed = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evb0 = nullptr, evb1 = nullptr, evb_mid = nullptr;
    // host-buffer calls: upload / download streams and the events that chain the chunks
    cudaStream_t stream_in = nullptr, stream_out = nullptr;
    cudaEvent_t ev_in = nullptr, ev_k = nullptr, ev_out = nullptr;
    std::string err;
    int64_t launches = 0;
    // kernel variants (environment overrides for A/B measurements: B2RT_TRACE_VARIANT, B2RT_TRACE_REFILL,
    // B2RT_GATHER_VARIANT); results are identical for every setting
    int trace_variant = 2;       // 0: one thread per ray, 1: persistent grid + ray pool, 2: + postponed leaves
    int trace_refill = 0;        // persistent: refill idle lanes when fewer than this many lanes are active
    int pt_blocks_per_sm[4] = {0, 0, 0, 0};
    int trace_sort = -1;         // variant 2 child order: 2 = near to far, 1 = nearest first only, 0 = slot order, < 0 = 2 for closest / 0 for any hit
    int trace_planes = 1;        // variant 2: near / far slab planes picked by the ray's direction signs (load address) instead
                                 // of min / max per child: 1 = closest-hit rays only (default), 2 = any-hit rays too, 0 = never
                                 // (B2RT_TRACE_PLANES)
    int trace_pool_flip = 0;     // which of the two ray-pool counters the next speculative-kernel launch uses
    int trace_chunk = 32;        // persistent: consecutive rays a warp takes from the pool per atomicAdd
    int64_t host_chunk = 1 << 18; // rays per chunk of the host-buffer pipeline (B2RT_HOST_CHUNK)
    int sort_variant = 1;        // K2: 1 = one-sweep (one histogram pass + one look-back scatter launch per digit) for inputs
                                 // below 2^22 pairs, the three-launch pass (histogram, scan, scatter) above; 2 = one-sweep
                                 // always, 0 = never; B2RT_SORT_VARIANT
    int sort_rank = 1;           // K2 scatter: how lanes find their digit peers: 1 = eight ballots (default: 1.3-1.4 x faster on
                                 // B200, r06h), 0 = match.any (B2RT_SORT_RANK)
    int photon_variant = 1;      // 0: one thread per path, 1: resident grid with path regeneration
    int photon_blocks_per_sm = 0;
    int collapse_blocks_per_sm = 0;   // occupancy of the cooperative wide-collapse kernel (build.cu)
    int gather_blocks_per_sm[4] = {0, 0, 0, 0};
    float gather_target = 1.6f;  // first gather radius: sized to hold gather_target * k photons at the local density
    int gather_tcap = 1024;      // gather variant 2: positions in the shared tile (1024 -> 7 blocks per SM, 2048 -> 5)
    int gather_qt = 8;           // gather variant 2: consecutive queries that share one staged tile (8, 16 or 32)
    int gather_variant = 2;      // 0: row-by-row scan with direct loads, 1: coarse table + TMA-staged tiles per warp,
                                 // 2: one TMA-staged tile shared by a block's consecutive queries (default)

    // host-side staging of the scene (flattened)
    std::vector<float> h_verts;      // xyz
    std::vector<int32_t> h_idx;      // 3 per tri, global vertex indices
    std::vector<int32_t> h_tri_prim; // global prim id of each triangle
    // per-vertex shading attributes (world-space normals, uv) of the meshes that have them; sized like h_verts once any
    // mesh brought one, empty otherwise.  h_tri_flags[tri]: bit 0 normals, bit 1 uv, bit 2 reversed orientation
    std::vector<float> h_vnormal, h_vuv;
    std::vector<uint8_t> h_tri_flags;
    bool any_attr = false;
    std::vector<AnalyticPrim> h_analytic;
    std::vector<int32_t> h_analytic_prim;  // global prim id
    std::vector<int32_t> h_analytic_type;
    std::vector<int32_t> h_prim_material, h_prim_light;
    std::vector<Material> h_materials;
    std::vector<Light> h_lights;
    int32_t n_prims = 0;             // items of THIS context's tree: primitives and (top level only) one item per instance
    // objects and instances (two-level BVH).  An object is a child context (own staging, own tree, same stream) that the
    // b2rt_add_* calls between b2rt_object_begin / _end fill; instances live in the parent only.
    std::vector<b2rt_ctx *> objects;
    b2rt_ctx *cur_object = nullptr;  // != nullptr between b2rt_object_begin and b2rt_object_end
    b2rt_ctx *parent = nullptr;      // set in an object's context
    struct HostInstance { float o2w[12], w2o[12]; int32_t object; int32_t item; };
    std::vector<HostInstance> h_instances;
    std::vector<int32_t> h_instance_base;   // global id of each instance's first primitive (assigned by the build)
    int64_t n_virtual_prims = 0;     // n_prims + the primitives of all instances = the size of the global id space
    DevBuf d_instances, d_objects;   // InstanceRec[n_instances], DScene[n_objects]

    // device scene + accel
    bool built = false;
    DevBuf d_verts, d_idx, d_analytic, d_prim_kind;   // prim_kind[n_prims]: tri index or ~analytic index
    DevBuf d_vnormal, d_vuv, d_tri_flags;             // uploaded only when any_attr
    DevBuf d_leaf_box;      // float[6*n] per-prim AABB (prim id order)
    DevBuf d_scene_box;     // uint32[12]: centroid bounds + scene bounds as ordered ints
    DevBuf d_keys, d_vals, d_keys_alt, d_vals_alt, d_hist;  // sort
    DevBuf d_bin_parent, d_bin_child, d_bin_box, d_bin_flag, d_bin_cost, d_bin_count, d_bin_info;
    DevBuf d_wide_tmp, d_wide_flag, d_wide_index, d_queue_a, d_queue_b, d_counters;
    DevBuf d_wide_nodes, d_prim_recs, d_leaf_prims;
    DevBuf d_rays, d_hits, d_occ;  // staging for host-pointer calls
    DevBuf d_scan_tmp;
    DevBuf d_prim_material, d_prim_light, d_materials, d_lights, d_light_cdf;
    DevBuf d_slots, d_cnt_c, d_cnt_i, d_grid_box;
    DevBuf d_queries, d_q_rgb, d_q_r2, d_q_found, d_q_idx;
    DevBuf d_samples, d_qslots, d_qcnt, d_q_rgb2, d_frame;
    DevBuf d_film_xy, d_film_rgb;       // per-sample raster positions / radiance of a frame reconstructed with a pixel filter
    b2rt_filter film_filter = {B2RT_FILTER_BOX, 0.5f, 0.5f, 0.f, 0.f};   // b2rt_film_set_filter
    int sampler_type = B2RT_SAMPLER_RANDOM, sampler_nx = 1, sampler_ny = 1;   // b2rt_sampler_set
    int film_variant = 1;               // 1: shared-memory tiles (film_tile_kernel), 0: one thread per pixel (B2RT_FILM_VARIANT)
    size_t film_smem_opt_in = 0;
    PhotonMap maps[4];       // B2RT_MAP_CAUSTIC, B2RT_MAP_INDIRECT, B2RT_MAP_DIRECT, B2RT_MAP_RADIANCE
    // final gathering (finalgather.cu)
    int fg_enable = 0, fg_samples = 32;                 // b2rt_final_gather_set
    int fg_variant = 3;                                 // 3 (default): trace / cell-sorted lookup / combine as separate launches;
                                                        // 0: the fused warp-per-shading-point kernel (B2RT_FG_VARIANT, A/B)
    float fg_cos_angle = 0.98480775301220802f;          // cos(10 degrees)
    int64_t fg_lim_c = INT64_MAX, fg_lim_i = INT64_MAX; // b2rt_fg_deposit_limits
    bool radiance_computed = false;                     // maps[3].alpha holds Lo (true) or rho_r (false)
    DevBuf d_fg_queries, d_fg_E, d_fg_Etmp, d_fg_idx, d_fg_r2, d_fg_found, d_fg_slots, d_fg_cnt, d_fg_cnt2;
    DevBuf d_fg_dirs, d_fg_pend, d_fg_pq, d_fg_pidx, d_fg_flag, d_fg_coarse2, d_fg_nbox1, d_fg_nbox2;
    DevBuf d_fg_rpos, d_fg_rnrm;                        // packed {p, oct normal} / decoded normals of the sorted radiance photons
    uint64_t fg_tables_epoch = 0;                       // grid_epoch of maps[3] the tables above were built for (0 = none)
    DevBuf d_fg_rayrec, d_fg_rkeys, d_fg_rvals, d_fg_rkeys2, d_fg_rvals2, d_fg_ri;   // split gather pipeline, per gather ray
    b2rt_accel_info info{};
    // b2rt_get_timings: accumulated device time per stage while timings_on (StageScope)
    bool timings_on = false;
    int stage_depth = 0;                                // entry points call each other (shoot -> trace): only the outermost counts
    cudaEvent_t ev_s0 = nullptr, ev_s1 = nullptr;
    b2rt_timings timings{};
};

// NVTX range + (optionally) CUDA-event timing of one stage entry point
struct StageScope {
    b2rt_ctx *ctx; int stage; bool timed;
    StageScope(b2rt_ctx *c, int s, const char *name) : ctx(c), stage(s), timed(false) {
        nvtxRangePushA(name);
        if (!ctx) return;
        if (ctx->stage_depth++ == 0 && ctx->timings_on && ctx->ev_s0 && cudaSetDevice(ctx->device) == cudaSuccess)
            timed = cudaEventRecord(ctx->ev_s0, ctx->stream) == cudaSuccess;
    }
    ~StageScope() {
        if (ctx) {
            --ctx->stage_depth;
            if (timed && cudaEventRecord(ctx->ev_s1, ctx->stream) == cudaSuccess && cudaEventSynchronize(ctx->ev_s1) == cudaSuccess) {
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, ctx->ev_s0, ctx->ev_s1) == cudaSuccess) { ctx->timings.ms[stage] += ms; ctx->timings.calls[stage]++; }
            }
        }
        nvtxRangePop();
    }
};
#define B2RT_STAGE(ctx, stage) StageScope stage_scope__((ctx), (stage), __func__)

#define B2RT_CUDA(ctx, call)                                                              \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);             \
            return B2RT_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

#define B2RT_LAUNCH_CHECK(ctx)                                                            \
    do {                                                                                  \
        (ctx)->launches++;                                                                \
        cudaError_t e__ = cudaGetLastError();                                             \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string("kernel launch: ") + cudaGetErrorString(e__);        \
            return B2RT_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

// where a GPU's freshly traced photons go: its slice of the two maps' arrays on each of the n GPUs of the box (n = 1: the
// local arrays).  a / b = caustic / indirect (photon pass), direct / radiance (final-gathering pass).
#define B2RT_MAX_GPUS 16
struct PhotonDst { b2rt_photon *a[B2RT_MAX_GPUS]; b2rt_photon *b[B2RT_MAX_GPUS]; int n; };

// ---- stage entry points (host functions, each enqueues on ctx->stream) -----------------------
int b2rt_sort_pairs_impl(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt,
                         uint32_t *vals_alt, int key_bits, bool *result_in_alt);
int b2rt_build_impl(b2rt_ctx *ctx);
int b2rt_copy_staged_scene(const b2rt_ctx *src, b2rt_ctx *dst);
int b2rt_trace_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_hit *d_hits, uint8_t *d_occ,
                    bool any_hit);
struct DScene;
int b2rt_upload_tables(b2rt_ctx *ctx);
int b2rt_photons_trace_impl(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int flags,
                            int64_t n_stored[2]);
int b2rt_photons_trace_begin(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int flags, uint32_t tot[2]);
int b2rt_photons_trace_commit(b2rt_ctx *ctx, int64_t n_paths, int max_depth, const PhotonDst &dst);
int b2rt_photons_trace_fg_begin(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, uint32_t tot[2]);
int b2rt_photons_trace_fg_commit(b2rt_ctx *ctx, int64_t n_paths, int max_depth, const PhotonDst &dst);
int b2rt_photons_reserve(b2rt_ctx *ctx, DevBuf &b, int64_t keep, int64_t want);
int b2rt_photons_push(b2rt_ctx *ctx, const b2rt_photon *src, int64_t n_photons, const PhotonDst &dst);   // dst.a[0 .. n): peers
int b2rt_photon_grid_build_impl(b2rt_ctx *ctx, int map, float cell_size, float target_per_cell);
int b2rt_render_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *d_out);
int b2rt_render_pixels_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, const uint32_t *d_pixmap,
                            int64_t n_map, float *d_out);
int b2rt_camera_samples_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, int slots, float *d_rgb,
                             uint32_t *d_nq, b2rt_query *d_q);
int b2rt_camera_slots(b2rt_ctx *ctx, int max_specular_depth);
int b2rt_surface_at_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_surface *d_out);
int b2rt_sample_positions_impl(b2rt_ctx *ctx, int xres, int yres, const int32_t window[4], int spp, float *d_xy);
bool b2rt_filter_valid(const b2rt_filter *f);
void b2rt_filter_halo(const b2rt_filter *f, int *hx, int *hy);
int b2rt_film_impl(b2rt_ctx *ctx, const b2rt_filter *f, const int32_t cells[4], int32_t spp, const float *d_xy,
                   const float *d_rgb, const int32_t pixels[4], float *d_out);
int b2rt_gather_impl(b2rt_ctx *ctx, int map, int64_t nq, const b2rt_query *d_queries, int k, float max_dist2,
                     float *d_rgb, float *d_r2, int32_t *d_found, int32_t *d_idx);
void b2rt_fill_dscene(b2rt_ctx *ctx, DScene *sc);
// final gathering (finalgather.cu)
int b2rt_photons_trace_fg_impl(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int64_t n_stored[2]);
int b2rt_irradiance_impl(b2rt_ctx *ctx, int map, int64_t nq, const b2rt_query *d_queries, int k, float max_dist2,
                         float *d_E, bool accumulate);
int b2rt_radiance_compute_impl(b2rt_ctx *ctx, int nused, float max_dist2);
int b2rt_radiance_lookup_impl(b2rt_ctx *ctx, int64_t nq, const b2rt_query *d_queries, float *d_rgb, int32_t *d_idx);
int b2rt_final_gather_impl(b2rt_ctx *ctx, int64_t nq, const b2rt_query *d_queries, float max_dist2, float *d_rgb);
// exclusive scan of n uint32 in place (single pass, small n helper used by build)
int b2rt_exclusive_scan_u32(b2rt_ctx *ctx, uint32_t *d, int64_t n, uint32_t *d_total);
