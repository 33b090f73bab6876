/* TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatements used as the parity oracle for the GPU path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 * Build: make -C oracle port   (gcc -O2 -std=c11 -ffp-contract=off)                              */
#ifndef ORACLE_PORT_H
#define ORACLE_PORT_H
#include <stdint.h>
#define PORT_ANALYTIC_STRIDE 32

/* (1) pbrt primitive tests + the exhaustive loop of renderers/aggregatetest.cpp:93-102.
 * rays: 8 floats {o,tmin,d,tmax}.  prim kinds: kind[p] >= 0 -> triangle index, < 0 -> ~analytic index.
 * analytic record = PORT_ANALYTIC_STRIDE (32) floats {o2w[12], w2o[12], radius, height, inner, type(1 sphere,2 disk),
 * phimax in radians (0 = full turn), 3 unused}; for a sphere height / inner hold pbrt's zmin / zmax (zmin >= zmax: none).
 * Winner = (min t, then min prim id).  Outputs t (INF on miss), prim (-1), b1, b2.           */
void port_intersect_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                          const float *analytic, int64_t nrays, const float *rays, float *t_out,
                          int32_t *prim_out, float *b1_out, float *b2_out, int nthreads);
void port_occluded_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                         const float *analytic, int64_t nrays, const float *rays, uint8_t *occ, int nthreads);

/* (2) restatement of OUR deterministic LBVH builder (csrc/build.cu) for bit-exact topology checks.
 * Inputs: per-prim AABBs box[6n].  Outputs (caller-allocated):
 *   morton[n], order[n], parent[2n-1], child[2(n-1)], bin_box[6(2n-1)], collapse[n-1],
 *   wide nodes as raw 32-uint32 records (capacity n), returns number of wide nodes.        */
int port_lbvh_build(int n, const float *box, uint32_t *morton, int32_t *order, int32_t *parent,
                    int32_t *child, float *bin_box, uint8_t *collapse, uint32_t *wide_nodes);
void port_prim_boxes(int nprims, const int32_t *kind, const float *verts, const int32_t *idx, float *box);

/* (3) stable LSD radix sort restatement (std stable sort semantics) */
void port_sort_pairs(int64_t n, uint32_t *keys, uint32_t *vals, int key_bits);

/* (4) photon pass / hash grid / radiance estimate restatements (oracle_photon.c) */
typedef struct {
    int32_t nprims; const int32_t *kind; const float *verts; const int32_t *idx; const float *analytic; /* as in (1) */
    const int32_t *prim_material;   /* per prim, -1 = default grey matte 0.5 */
    int32_t nmaterials; const float *materials;   /* 8 floats each: type(0 matte,1 mirror,2 glass), kd/kr[3], kt[3], ior */
    int32_t nlights; const float *lights;         /* 8 floats each: type(0 point,1 disk area), p[3], rgb[3], prim */
    /* per-vertex shading attributes (may all be NULL): world-space normals 3 / vertex, uv 2 / vertex, and per TRIANGLE
     * flags: bit 0 the mesh has normals, bit 1 uv, bit 2 reversed orientation (same meaning as csrc/b2rt_shade.cuh) */
    const float *vnormal; const float *vuv; const uint8_t *tri_flags;
    /* camera pass only (may be NULL): area-light index per prim (-1 none), light samples per light */
    const int32_t *prim_light; const int32_t *light_nsamples;
    /* spot / distant lights (types 2 / 3; may be NULL): 24 floats per light {l2w[9], w2l[9], cos_total, cos_falloff,
     * dir_to_light[3], pad}; world_bounds[6] = the scene box the distant light's emission disk is derived from */
    const float *light_ext; const float *world_bounds;
} port_scene;
/* per-sample oracle of the camera pass (oracle_render.h): see there */
void port_camera_samples(const port_scene *sc, uint64_t seed, const float cam[12], int64_t n, const float *xy,
                         const uint32_t *gpix, const uint32_t *ks, int sampler_type, int max_specular_depth,
                         int direct_lighting, int want_queries, int slots, float *out_rgb, uint32_t *out_nq, float *out_q);
/* hit attributes (K6): closest hit of each ray + pbrt's Triangle::Intersect tail, Triangle::GetShadingGeometry and
 * BSDF::BSDF (shapes/trianglemesh.cpp:181-201, 311-389; core/reflection.cpp:601-609).  out: 20 floats per ray laid out as
 * b2rt_surface {p, t, ng, u, n, v, sn, ray_eps, tn, bits(prim)}; prim = -1 and zeros on a miss. */
void port_surface_at(const port_scene *sc, int64_t nrays, const float *rays, float *out);
typedef struct { float p[3]; uint32_t wi_oct; float alpha[3]; uint32_t path; } port_photon;   /* == b2rt_photon */
void port_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);
void port_sincos_quarter(float q, float *s, float *c);
/* traces paths [first, first+n); appends to out_c / out_i (capacities cap_c / cap_i); returns 0, or -1 on overflow */
int port_photon_trace(const port_scene *sc, uint64_t seed, int64_t first_path, int64_t n_paths, int max_depth,
                      int flags, port_photon *out_c, int64_t cap_c, int64_t *n_c, port_photon *out_i, int64_t cap_i,
                      int64_t *n_i);
/* grid over n photons: fills origin[3], dims[3], *cell, *inv, keys[n] (sorted), order[n], cell_start (caller allocates
 * after a first call with cell_start == NULL to learn dims) */
void port_grid_build(int64_t n, const port_photon *ph, float cell_size, float target_per_cell, float origin[3],
                     int32_t dims[3], float *cell, float *inv, uint32_t *keys, uint32_t *order, uint32_t *cell_start);
/* brute-force k nearest (by (d2, index)) within max_dist2 over the SORTED photon array + pbrt's LPhoton for a diffuse
 * surface.  query = 12 floats {p, n, kd, pad}.  k <= 0: fixed radius.  idx_out may be NULL (k entries per query). */
void port_gather(int64_t n, const port_photon *sorted, int64_t n_paths, int64_t nq, const float *queries, int k,
                 float max_dist2, float *rgb, float *r2_out, int32_t *found, int32_t *idx_out);
#endif
