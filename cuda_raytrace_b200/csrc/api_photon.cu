// extern "C" surface of the photon pass, the hash grid and the radiance estimate (include/b2rt.h).
// Host-side only: argument checks, staging of host buffers, buffer management.
#include "b2rt_internal.cuh"
#include <string.h>
#include <math.h>

static int need_built(b2rt_ctx *ctx, const char *who) {
    if (!ctx->built) { ctx->err = std::string(who) + ": b2rt_build_accel has not been called"; return B2RT_E_STATE; }
    return B2RT_OK;
}
static int check_map(b2rt_ctx *ctx, int32_t map, const char *who) {
    if (map < B2RT_MAP_CAUSTIC || map > B2RT_MAP_RADIANCE) { ctx->err = std::string(who) + ": bad map id"; return B2RT_E_ARG; }
    return B2RT_OK;
}
static int status_check(b2rt_ctx *ctx) {
    uint32_t s = 0;
    B2RT_CUDA(ctx, cudaMemcpyAsync(&s, ctx->d_counters.as<uint32_t>() + 8, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (s) {
        cudaMemsetAsync(ctx->d_counters.as<uint32_t>() + 8, 0, 4, ctx->stream);
        ctx->err = "traversal stack overflow on " + std::to_string(s) + " rays";
        return B2RT_E_OVERFLOW;
    }
    return B2RT_OK;
}

extern "C" {

int b2rt_photons_clear(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    for (PhotonMap &m : ctx->maps) { m.n = 0; m.n_paths = 0; m.grid_built = false; }
    ctx->radiance_computed = false;
    return B2RT_OK;
}

int b2rt_photons_trace(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int32_t max_depth, int32_t flags,
                       int64_t n_stored[2]) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_PHOTONS);
    if (first_path < 0 || n_paths < 0 || max_depth < 1 || max_depth > 64 || !(flags & 3)) {
        ctx->err = "photons_trace: bad arguments"; return B2RT_E_ARG;
    }
    int rc = need_built(ctx, "photons_trace");
    if (rc) return rc;
    if (ctx->h_lights.empty()) { ctx->err = "photons_trace: the scene has no lights"; return B2RT_E_STATE; }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    rc = b2rt_photons_trace_impl(ctx, first_path, n_paths, max_depth, flags, n_stored);
    if (rc) return rc;
    return status_check(ctx);
}

int b2rt_photons_shoot(b2rt_ctx *ctx, int64_t n_caustic_wanted, int64_t n_indirect_wanted, int32_t max_depth,
                       int64_t batch_paths, int64_t n_paths_out[2], int64_t n_stored_out[2]) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_PHOTONS);
    if (n_caustic_wanted < 0 || n_indirect_wanted < 0 || batch_paths <= 0) { ctx->err = "photons_shoot: bad arguments"; return B2RT_E_ARG; }
    int rc = b2rt_photons_clear(ctx);
    if (rc) return rc;
    int64_t first = 0;
    bool gave_up = false;
    bool c_done = n_caustic_wanted == 0, i_done = n_indirect_wanted == 0;
    while (!(c_done && i_done)) {
        int flags = (c_done ? 0 : 1) | (i_done ? 0 : 2);
        int64_t st[2];
        rc = b2rt_photons_trace(ctx, first, batch_paths, max_depth, flags, st);
        if (rc) return rc;
        first += batch_paths;
        if (ctx->maps[0].n >= n_caustic_wanted) c_done = true;
        if (ctx->maps[1].n >= n_indirect_wanted) i_done = true;
        // pbrt gives up when nothing is being stored (photonmap.cpp:182-184, 526-537): after 500 000 paths a map that is
        // still wanted and holds fewer than blockSize / 1024 = 4 photons (pbrt passes its block size, 4096, as `shot`,
        // not the paths shot so far).  Like pbrt this is not fatal: it reports, ERASES every map and renders on
        // without them -- a diffuse-only scene with the default causticphotons = 20000 ends up here in pbrt as well.
        if (first > 500000 && ((!c_done && ctx->maps[0].n < 4) || (!i_done && ctx->maps[1].n < 4))) {
            rc = b2rt_photons_clear(ctx);
            if (rc) return rc;
            ctx->err = "photons_shoot: unable to store enough photons, giving up (photon maps erased, as pbrt does)";
            gave_up = true;
            break;
        }
    }
    for (int m = 0; m < 2; ++m) {
        if (n_paths_out) n_paths_out[m] = ctx->maps[m].n_paths;
        if (n_stored_out) n_stored_out[m] = ctx->maps[m].n;
    }
    (void)gave_up;       // reported through the counts (0 photons although some were wanted) and b2rt_last_error
    return B2RT_OK;
}

int b2rt_photons_count(b2rt_ctx *ctx, int32_t map, int64_t *n_photons, int64_t *n_paths) {
    if (!ctx) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "photons_count");
    if (rc) return rc;
    if (n_photons) *n_photons = ctx->maps[map].n;
    if (n_paths) *n_paths = ctx->maps[map].n_paths;
    return B2RT_OK;
}

int b2rt_photons_download(b2rt_ctx *ctx, int32_t map, b2rt_photon *out, int64_t capacity, int where) {
    if (!ctx) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "photons_download");
    if (rc) return rc;
    PhotonMap &m = ctx->maps[map];
    if (!out || capacity < m.n) { ctx->err = "photons_download: buffer too small"; return B2RT_E_ARG; }
    if (m.n == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    B2RT_CUDA(ctx, cudaMemcpyAsync(out, m.photons.p, (size_t)m.n * sizeof(b2rt_photon),
                                   where == B2RT_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B2RT_OK;
}

int b2rt_photons_upload(b2rt_ctx *ctx, int32_t map, const b2rt_photon *in, int64_t n, int64_t n_paths, int where) {
    if (!ctx) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "photons_upload");
    if (rc) return rc;
    if (n < 0 || (n && !in) || n >= (1ll << 32)) { ctx->err = "photons_upload: bad arguments"; return B2RT_E_ARG; }
    PhotonMap &m = ctx->maps[map];
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    if ((const void *)in != m.photons.p) {
        B2RT_CUDA(ctx, m.photons.reserve((size_t)(n ? n : 1) * sizeof(b2rt_photon)));
        if (n)
            B2RT_CUDA(ctx, cudaMemcpyAsync(m.photons.p, in, (size_t)n * sizeof(b2rt_photon),
                                           where == B2RT_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                           ctx->stream));
        B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    m.n = n; m.n_paths = n_paths; m.grid_built = false;
    if (map == B2RT_MAP_RADIANCE) ctx->radiance_computed = false;    // alpha of uploaded radiance photons is rho_r
    return B2RT_OK;
}

void *b2rt_photons_device_ptr(b2rt_ctx *ctx, int32_t map) {
    if (!ctx || map < 0 || map > B2RT_MAP_RADIANCE) return nullptr;
    return ctx->maps[map].photons.p;
}

int b2rt_photon_grid_build(b2rt_ctx *ctx, int32_t map, float cell_size, float target_per_cell) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_GRID);
    int rc = check_map(ctx, map, "photon_grid_build");
    if (rc) return rc;
    if (ctx->maps[map].n >= (1ll << 32)) { ctx->err = "photon_grid_build: more than 2^32 photons"; return B2RT_E_ARG; }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    return b2rt_photon_grid_build_impl(ctx, map, cell_size, target_per_cell);
}

int b2rt_photon_grid_info(b2rt_ctx *ctx, int32_t map, b2rt_grid_info *info) {
    if (!ctx || !info) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "photon_grid_info");
    if (rc) return rc;
    if (!ctx->maps[map].grid_built) { ctx->err = "photon_grid_info: grid not built"; return B2RT_E_STATE; }
    *info = ctx->maps[map].grid;
    return B2RT_OK;
}

int b2rt_photon_grid_download(b2rt_ctx *ctx, int32_t map, uint32_t *keys, uint32_t *order, uint32_t *cell_start,
                              b2rt_photon *sorted) {
    if (!ctx) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "photon_grid_download");
    if (rc) return rc;
    PhotonMap &m = ctx->maps[map];
    if (!m.grid_built) { ctx->err = "photon_grid_download: grid not built"; return B2RT_E_STATE; }
    cudaStream_t st = ctx->stream;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    auto dn = [&](void *dst, const DevBuf &src, size_t bytes) -> cudaError_t {
        if (!dst || !bytes) return cudaSuccess;
        return cudaMemcpyAsync(dst, src.p, bytes, cudaMemcpyDeviceToHost, st);
    };
    B2RT_CUDA(ctx, dn(keys, m.keys, (size_t)m.n * 4));
    B2RT_CUDA(ctx, dn(order, m.order, (size_t)m.n * 4));
    B2RT_CUDA(ctx, dn(cell_start, m.cell_start, (size_t)(m.grid.n_cells + 1) * 4));
    B2RT_CUDA(ctx, dn(sorted, m.sorted, (size_t)m.n * sizeof(b2rt_photon)));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

static int gather_common(b2rt_ctx *ctx, int32_t map, int64_t nq, const b2rt_query *queries, int32_t k, float max_dist2,
                         float *out_rgb, float *out_r2, int32_t *out_found, int32_t *idx_out, int where) {
    if (!ctx) return B2RT_E_ARG;
    int rc = check_map(ctx, map, "gather");
    if (rc) return rc;
    if (nq < 0 || (nq && !queries) || !(max_dist2 > 0.f) || k > 224 || (idx_out && k <= 0)) {
        ctx->err = "gather: bad arguments (k <= 224, max_dist2 > 0)"; return B2RT_E_ARG;
    }
    PhotonMap &m = ctx->maps[map];
    if (!m.grid_built) { ctx->err = "gather: b2rt_photon_grid_build has not been called for this map"; return B2RT_E_STATE; }
    if (nq == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    if (where == B2RT_DEVICE) return b2rt_gather_impl(ctx, map, nq, queries, k, max_dist2, out_rgb, out_r2, out_found, idx_out);
    if (where != B2RT_HOST) { ctx->err = "gather: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
    cudaStream_t st = ctx->stream;
    B2RT_CUDA(ctx, ctx->d_queries.reserve((size_t)nq * sizeof(b2rt_query)));
    B2RT_CUDA(ctx, ctx->d_q_rgb.reserve((size_t)nq * 12));
    B2RT_CUDA(ctx, ctx->d_q_r2.reserve((size_t)nq * 4));
    B2RT_CUDA(ctx, ctx->d_q_found.reserve((size_t)nq * 4));
    if (idx_out) B2RT_CUDA(ctx, ctx->d_q_idx.reserve((size_t)nq * k * 4));
    B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_queries.p, queries, (size_t)nq * sizeof(b2rt_query), cudaMemcpyHostToDevice, st));
    rc = b2rt_gather_impl(ctx, map, nq, ctx->d_queries.as<b2rt_query>(), k, max_dist2, ctx->d_q_rgb.as<float>(),
                          ctx->d_q_r2.as<float>(), ctx->d_q_found.as<int32_t>(),
                          idx_out ? ctx->d_q_idx.as<int32_t>() : nullptr);
    if (rc) return rc;
    if (out_rgb) B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, ctx->d_q_rgb.p, (size_t)nq * 12, cudaMemcpyDeviceToHost, st));
    if (out_r2) B2RT_CUDA(ctx, cudaMemcpyAsync(out_r2, ctx->d_q_r2.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    if (out_found) B2RT_CUDA(ctx, cudaMemcpyAsync(out_found, ctx->d_q_found.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    if (idx_out) B2RT_CUDA(ctx, cudaMemcpyAsync(idx_out, ctx->d_q_idx.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, st));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

int b2rt_gather(b2rt_ctx *ctx, int32_t map, int64_t nqueries, const b2rt_query *queries, int32_t k, float max_dist2,
                float *out_rgb, float *out_radius2, int32_t *out_found, int where) {
    B2RT_STAGE(ctx, B2RT_STAGE_GATHER);
    return gather_common(ctx, map, nqueries, queries, k, max_dist2, out_rgb, out_radius2, out_found, nullptr, where);
}

int b2rt_gather_sets(b2rt_ctx *ctx, int32_t map, int64_t nqueries, const b2rt_query *queries, int32_t k,
                     float max_dist2, int32_t *idx_out, int where) {
    if (ctx && !idx_out) { ctx->err = "gather_sets: idx_out is NULL"; return B2RT_E_ARG; }
    B2RT_STAGE(ctx, B2RT_STAGE_GATHER);
    return gather_common(ctx, map, nqueries, queries, k, max_dist2, nullptr, nullptr, nullptr, idx_out, where);
}

// ---- final gathering (finalgather.cu) ----------------------------------------------------------------------------------
int b2rt_photons_trace_fg(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int32_t max_depth, int64_t n_stored[2]) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_PHOTONS);
    if (first_path < 0 || n_paths < 0 || max_depth < 1 || max_depth > 64) { ctx->err = "photons_trace_fg: bad arguments"; return B2RT_E_ARG; }
    int rc = need_built(ctx, "photons_trace_fg");
    if (rc) return rc;
    if (ctx->h_lights.empty()) { ctx->err = "photons_trace_fg: the scene has no lights"; return B2RT_E_STATE; }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    rc = b2rt_photons_trace_fg_impl(ctx, first_path, n_paths, max_depth, n_stored);
    if (rc) return rc;
    return status_check(ctx);
}

int b2rt_irradiance(b2rt_ctx *ctx, int32_t map, int64_t nq, const b2rt_query *queries, int32_t k, float max_dist2,
                    float *out_E, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_GATHER);
    int rc = check_map(ctx, map, "irradiance");
    if (rc) return rc;
    if (nq < 0 || (nq && (!queries || !out_E)) || !(max_dist2 > 0.f) || k < 1 || k > 224) {
        ctx->err = "irradiance: bad arguments (1 <= k <= 224, max_dist2 > 0)"; return B2RT_E_ARG;
    }
    PhotonMap &m = ctx->maps[map];
    if (!m.grid_built) { ctx->err = "irradiance: b2rt_photon_grid_build has not been called for this map"; return B2RT_E_STATE; }
    if (nq == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    if (where == B2RT_DEVICE) return b2rt_irradiance_impl(ctx, map, nq, queries, k, max_dist2, out_E, false);
    if (where != B2RT_HOST) { ctx->err = "irradiance: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
    cudaStream_t st = ctx->stream;
    B2RT_CUDA(ctx, ctx->d_queries.reserve((size_t)nq * sizeof(b2rt_query)));
    B2RT_CUDA(ctx, ctx->d_q_rgb.reserve((size_t)nq * 12));
    B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_queries.p, queries, (size_t)nq * sizeof(b2rt_query), cudaMemcpyHostToDevice, st));
    rc = b2rt_irradiance_impl(ctx, map, nq, ctx->d_queries.as<b2rt_query>(), k, max_dist2, ctx->d_q_rgb.as<float>(), false);
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaMemcpyAsync(out_E, ctx->d_q_rgb.p, (size_t)nq * 12, cudaMemcpyDeviceToHost, st));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

int b2rt_radiance_compute(b2rt_ctx *ctx, int32_t nused, float max_dist2) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_GRID);
    if (nused < 1 || nused > 224 || !(max_dist2 > 0.f)) { ctx->err = "radiance_compute: bad arguments (1 <= nused <= 224)"; return B2RT_E_ARG; }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = b2rt_radiance_compute_impl(ctx, nused, max_dist2);
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B2RT_OK;
}

int b2rt_radiance_lookup(b2rt_ctx *ctx, int64_t nq, const b2rt_query *queries, float *out_rgb, int32_t *out_idx, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_GATHER);
    if (nq < 0 || (nq && !queries)) { ctx->err = "radiance_lookup: bad arguments"; return B2RT_E_ARG; }
    PhotonMap &m = ctx->maps[B2RT_MAP_RADIANCE];
    if (!m.grid_built) { ctx->err = "radiance_lookup: the radiance-photon grid is not built (b2rt_radiance_compute)"; return B2RT_E_STATE; }
    if (nq == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    if (where == B2RT_DEVICE) return b2rt_radiance_lookup_impl(ctx, nq, queries, out_rgb, out_idx);
    if (where != B2RT_HOST) { ctx->err = "radiance_lookup: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
    cudaStream_t st = ctx->stream;
    B2RT_CUDA(ctx, ctx->d_queries.reserve((size_t)nq * sizeof(b2rt_query)));
    B2RT_CUDA(ctx, ctx->d_q_rgb.reserve((size_t)nq * 12));
    B2RT_CUDA(ctx, ctx->d_q_found.reserve((size_t)nq * 4));
    B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_queries.p, queries, (size_t)nq * sizeof(b2rt_query), cudaMemcpyHostToDevice, st));
    int rc = b2rt_radiance_lookup_impl(ctx, nq, ctx->d_queries.as<b2rt_query>(), ctx->d_q_rgb.as<float>(), ctx->d_q_found.as<int32_t>());
    if (rc) return rc;
    if (out_rgb) B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, ctx->d_q_rgb.p, (size_t)nq * 12, cudaMemcpyDeviceToHost, st));
    if (out_idx) B2RT_CUDA(ctx, cudaMemcpyAsync(out_idx, ctx->d_q_found.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

int b2rt_final_gather(b2rt_ctx *ctx, int64_t nq, const b2rt_query *queries, float max_dist2, float *out_rgb, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_GATHER);
    if (nq < 0 || (nq && (!queries || !out_rgb)) || !(max_dist2 > 0.f)) { ctx->err = "final_gather: bad arguments"; return B2RT_E_ARG; }
    int rc = need_built(ctx, "final_gather");
    if (rc) return rc;
    if (!ctx->maps[B2RT_MAP_INDIRECT].grid_built || !ctx->radiance_computed || !ctx->maps[B2RT_MAP_RADIANCE].grid_built) {
        ctx->err = "final_gather: needs the indirect map's grid and b2rt_radiance_compute"; return B2RT_E_STATE;
    }
    if (nq == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const b2rt_query *dq = queries;
    float *d_rgb = out_rgb;
    if (where != B2RT_DEVICE) {
        if (where != B2RT_HOST) { ctx->err = "final_gather: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
        B2RT_CUDA(ctx, ctx->d_queries.reserve((size_t)nq * sizeof(b2rt_query)));
        B2RT_CUDA(ctx, ctx->d_q_rgb.reserve((size_t)nq * 12));
        B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_queries.p, queries, (size_t)nq * sizeof(b2rt_query), cudaMemcpyHostToDevice, st));
        dq = ctx->d_queries.as<b2rt_query>(); d_rgb = ctx->d_q_rgb.as<float>();
    }
    rc = b2rt_final_gather_impl(ctx, nq, dq, max_dist2, d_rgb);
    if (rc) return rc;
    if (where != B2RT_DEVICE) B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, d_rgb, (size_t)nq * 12, cudaMemcpyDeviceToHost, st));
    return status_check(ctx);
}

int b2rt_fg_deposit_limits(b2rt_ctx *ctx, int64_t caustic_paths, int64_t indirect_paths) {
    if (!ctx) return B2RT_E_ARG;
    ctx->fg_lim_c = caustic_paths < 0 ? INT64_MAX : caustic_paths;
    ctx->fg_lim_i = indirect_paths < 0 ? INT64_MAX : indirect_paths;
    return B2RT_OK;
}

int b2rt_final_gather_set(b2rt_ctx *ctx, int32_t enable, int32_t gather_samples, float gather_angle_deg) {
    if (!ctx) return B2RT_E_ARG;
    if (enable && (gather_samples < 1 || gather_samples > 4096 || !(gather_angle_deg > 0.f) || !(gather_angle_deg < 90.f))) {
        ctx->err = "final_gather_set: gather_samples in [1, 4096], gather_angle in (0, 90) degrees"; return B2RT_E_ARG;
    }
    ctx->fg_enable = enable ? 1 : 0;
    if (enable) {
        ctx->fg_samples = gather_samples;
        // cos(Radians(ga)) (photonmap.cpp:296): float degrees -> float radians -> cos in double, rounded once
        const float rad = (3.14159265358979323846f / 180.f) * gather_angle_deg;
        ctx->fg_cos_angle = (float)cos((double)rad);
    }
    return B2RT_OK;
}

int b2rt_surface_at(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, b2rt_surface *out, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_RENDER);
    if (nrays < 0 || (nrays && (!rays || !out))) { ctx->err = "surface_at: bad arguments"; return B2RT_E_ARG; }
    int rc = need_built(ctx, "surface_at");
    if (rc || nrays == 0) return rc;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const b2rt_ray *dr = rays;
    b2rt_surface *d_out = out;
    if (where != B2RT_DEVICE) {
        if (where != B2RT_HOST) { ctx->err = "surface_at: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
        B2RT_CUDA(ctx, ctx->d_rays.reserve((size_t)nrays * sizeof(b2rt_ray)));
        B2RT_CUDA(ctx, ctx->d_samples.reserve((size_t)nrays * sizeof(b2rt_surface)));
        B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_rays.p, rays, (size_t)nrays * sizeof(b2rt_ray), cudaMemcpyHostToDevice, ctx->stream));
        dr = ctx->d_rays.as<b2rt_ray>(); d_out = ctx->d_samples.as<b2rt_surface>();
    }
    rc = b2rt_surface_at_impl(ctx, nrays, dr, d_out);
    if (rc) return rc;
    if (where != B2RT_DEVICE)
        B2RT_CUDA(ctx, cudaMemcpyAsync(out, d_out, (size_t)nrays * sizeof(b2rt_surface), cudaMemcpyDeviceToHost, ctx->stream));
    return status_check(ctx);
}

int b2rt_camera_samples(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, uint32_t *out_nleaves,
                        b2rt_query *out_queries, int32_t *slots_out) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_RENDER);
    if (!cam || !fp || !out_rgb || !out_nleaves || fp->spp < 1 || fp->x0 < 0 || fp->y0 < 0 || fp->x1 > cam->xres ||
        fp->y1 > cam->yres || fp->x0 >= fp->x1 || fp->y0 >= fp->y1) { ctx->err = "camera_samples: bad arguments"; return B2RT_E_ARG; }
    int rc = need_built(ctx, "camera_samples");
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const int slots = b2rt_camera_slots(ctx, fp->max_specular_depth);
    if (slots_out) *slots_out = slots;
    const int64_t n = (int64_t)(fp->x1 - fp->x0) * (fp->y1 - fp->y0) * fp->spp;
    B2RT_CUDA(ctx, ctx->d_samples.reserve((size_t)n * 12));
    B2RT_CUDA(ctx, ctx->d_qcnt.reserve((size_t)n * 4));
    if (out_queries) B2RT_CUDA(ctx, ctx->d_qslots.reserve((size_t)n * slots * sizeof(b2rt_query)));
    rc = b2rt_camera_samples_impl(ctx, cam, fp, slots, ctx->d_samples.as<float>(), ctx->d_qcnt.as<uint32_t>(),
                                  out_queries ? ctx->d_qslots.as<b2rt_query>() : nullptr);
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, ctx->d_samples.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    B2RT_CUDA(ctx, cudaMemcpyAsync(out_nleaves, ctx->d_qcnt.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_queries)
        B2RT_CUDA(ctx, cudaMemcpyAsync(out_queries, ctx->d_qslots.p, (size_t)n * slots * sizeof(b2rt_query), cudaMemcpyDeviceToHost, ctx->stream));
    return status_check(ctx);
}

int b2rt_render_frame(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_RENDER);
    if (!cam || !fp || !out_rgb || fp->spp < 1 || fp->x0 < 0 || fp->y0 < 0 || fp->x1 > cam->xres || fp->y1 > cam->yres ||
        fp->x0 >= fp->x1 || fp->y0 >= fp->y1 || fp->nused > 224 || !(fp->max_dist2 > 0.f)) {
        ctx->err = "render_frame: bad arguments"; return B2RT_E_ARG;
    }
    int rc = need_built(ctx, "render_frame");
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t bytes = (size_t)(fp->x1 - fp->x0) * (fp->y1 - fp->y0) * 12;
    float *d_out = out_rgb;
    if (where != B2RT_DEVICE) {
        B2RT_CUDA(ctx, ctx->d_frame.reserve(bytes));
        d_out = ctx->d_frame.as<float>();
    }
    rc = b2rt_render_impl(ctx, cam, fp, d_out);
    if (rc) return rc;
    if (where != B2RT_DEVICE)
        B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, d_out, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    rc = status_check(ctx);
    if (rc) return rc;
    uint32_t dropped = 0;
    B2RT_CUDA(ctx, cudaMemcpyAsync(&dropped, ctx->d_counters.as<uint32_t>() + 9, 4, cudaMemcpyDeviceToHost, ctx->stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (dropped) {
        cudaMemsetAsync(ctx->d_counters.as<uint32_t>() + 9, 0, 4, ctx->stream);
        ctx->err = "render_frame: " + std::to_string(dropped) + " samples had more diffuse leaves than slots";
        return B2RT_E_OVERFLOW;
    }
    return B2RT_OK;
}

}  // extern "C"
