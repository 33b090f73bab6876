// The extern "C" surface declared in include/b2rt.h.  Host-side only: argument checks, scene
// staging, buffer management; every compute step is a kernel in sort.cu / build.cu / trace.cu.
// There is deliberately no CPU path: without a usable CUDA device b2rt_create fails.
#include "b2rt_internal.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

static std::string g_create_error;

static int invert34(const float m[12], float out[12]) {
    // general 3x4 affine inverse in double
    double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (det == 0.0) return -1;
    double id = 1.0 / det;
    double r[9] = {(e * i - f * h) * id, (c * h - b * i) * id, (b * f - c * e) * id,
                   (f * g - d * i) * id, (a * i - c * g) * id, (c * d - a * f) * id,
                   (d * h - e * g) * id, (b * g - a * h) * id, (a * e - b * d) * id};
    for (int row = 0; row < 3; ++row) {
        for (int col = 0; col < 3; ++col) out[4 * row + col] = (float)r[3 * row + col];
        out[4 * row + 3] = (float)-(r[3 * row] * m[3] + r[3 * row + 1] * m[7] + r[3 * row + 2] * m[11]);
    }
    return 0;
}

extern "C" {

int b2rt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int b2rt_create(int device, uint64_t seed, b2rt_ctx **out) {
    if (!out) return B2RT_E_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_error = std::string("b2rt_create: no CUDA device (") + cudaGetErrorString(e) +
                         "); this library has no CPU fallback";
        cudaGetLastError();
        return B2RT_E_CUDA;
    }
    if (device < 0 || device >= n) { g_create_error = "b2rt_create: bad device index"; return B2RT_E_ARG; }
    b2rt_ctx *ctx = new b2rt_ctx();
    ctx->device = device; ctx->seed = seed;
#define CK(call) if ((e = (call)) != cudaSuccess) { g_create_error = std::string(#call) + ": " + cudaGetErrorString(e); delete ctx; return B2RT_E_CUDA; }
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        g_create_error = "b2rt_create: device is not sm_100-class (kernels are built for sm_100a only)";
        delete ctx; return B2RT_E_CUDA;
    }
    ctx->sm_count = prop.multiProcessorCount;
    if (const char *v = getenv("B2RT_TRACE_VARIANT")) ctx->trace_variant = atoi(v);
    if (const char *v = getenv("B2RT_TRACE_REFILL")) ctx->trace_refill = atoi(v);
    if (const char *v = getenv("B2RT_TRACE_CHUNK")) ctx->trace_chunk = atoi(v);
    if (const char *v = getenv("B2RT_GATHER_VARIANT")) ctx->gather_variant = atoi(v);
    if (const char *v = getenv("B2RT_SORT_VARIANT")) ctx->sort_variant = atoi(v);
    if (const char *v = getenv("B2RT_SORT_RANK")) ctx->sort_rank = atoi(v);
    if (const char *v = getenv("B2RT_PHOTON_VARIANT")) ctx->photon_variant = atoi(v);
    if (const char *v = getenv("B2RT_FG_VARIANT")) ctx->fg_variant = atoi(v);
    if (const char *v = getenv("B2RT_TRACE_SORT")) ctx->trace_sort = atoi(v);
    if (const char *v = getenv("B2RT_TRACE_PLANES")) ctx->trace_planes = atoi(v);
    if (const char *v = getenv("B2RT_FILM_VARIANT")) ctx->film_variant = atoi(v);
    if (const char *v = getenv("B2RT_GATHER_QT")) ctx->gather_qt = atoi(v);
    if (const char *v = getenv("B2RT_GATHER_TCAP")) ctx->gather_tcap = atoi(v);
    if (const char *v = getenv("B2RT_GATHER_TARGET")) ctx->gather_target = fminf(fmaxf((float)atof(v), 1.05f), 3.4f);
    if (const char *v = getenv("B2RT_HOST_CHUNK")) { long c = atol(v); if (c >= 1024) ctx->host_chunk = c; }
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&ctx->ev0)); CK(cudaEventCreate(&ctx->ev1));
    CK(cudaEventCreate(&ctx->evb0)); CK(cudaEventCreate(&ctx->evb1)); CK(cudaEventCreate(&ctx->evb_mid));
    CK(cudaStreamCreateWithFlags(&ctx->stream_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&ctx->stream_out, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_k, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_out, cudaEventDisableTiming));
    CK(cudaEventCreate(&ctx->ev_s0)); CK(cudaEventCreate(&ctx->ev_s1));
    if (const char *v = getenv("B2RT_TIMINGS")) ctx->timings_on = atoi(v) != 0;
    CK(ctx->d_counters.reserve(256));
    CK(cudaMemsetAsync(ctx->d_counters.p, 0, 256, ctx->stream));
#undef CK
    *out = ctx;
    return B2RT_OK;
}

// ---- objects (two-level BVH): child contexts that share the parent's device and stream --------------------------------
static void destroy_object(b2rt_ctx *c) {
    if (!c) return;
    cudaEventDestroy(c->evb0); cudaEventDestroy(c->evb1); cudaEventDestroy(c->evb_mid);
    delete c;
}
static b2rt_ctx *new_object(b2rt_ctx *ctx) {
    b2rt_ctx *c = new b2rt_ctx();
    c->device = ctx->device; c->seed = ctx->seed; c->sm_count = ctx->sm_count; c->stream = ctx->stream; c->parent = ctx;
    c->sort_variant = ctx->sort_variant; c->sort_rank = ctx->sort_rank;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaEventCreate(&c->evb0) != cudaSuccess || cudaEventCreate(&c->evb1) != cudaSuccess ||
        cudaEventCreate(&c->evb_mid) != cudaSuccess || c->d_counters.reserve(256) != cudaSuccess ||
        cudaMemsetAsync(c->d_counters.p, 0, 256, ctx->stream) != cudaSuccess) {
        ctx->err = std::string("object_begin: ") + cudaGetErrorString(cudaGetLastError());
        destroy_object(c);
        return nullptr;
    }
    return c;
}

// everything b2rt_add_* / b2rt_object_* / b2rt_add_instance staged on the host side of src -> dst (b2rt_multi_resize)
static void copy_staging(const b2rt_ctx *c0, b2rt_ctx *c) {
    c->h_verts = c0->h_verts; c->h_idx = c0->h_idx; c->h_tri_prim = c0->h_tri_prim; c->h_analytic = c0->h_analytic;
    c->h_analytic_prim = c0->h_analytic_prim; c->h_analytic_type = c0->h_analytic_type;
    c->h_prim_material = c0->h_prim_material; c->h_prim_light = c0->h_prim_light;
    c->h_materials = c0->h_materials; c->h_lights = c0->h_lights;
    c->h_vnormal = c0->h_vnormal; c->h_vuv = c0->h_vuv; c->h_tri_flags = c0->h_tri_flags; c->any_attr = c0->any_attr;
    c->n_prims = c0->n_prims;
    c->built = false;
}
}  // extern "C"
int b2rt_copy_staged_scene(const b2rt_ctx *src, b2rt_ctx *dst) {
    copy_staging(src, dst);
    for (b2rt_ctx *c : dst->objects) destroy_object(c);
    dst->objects.clear(); dst->cur_object = nullptr;
    for (const b2rt_ctx *o : src->objects) {
        b2rt_ctx *c = new_object(dst);
        if (!c) return B2RT_E_CUDA;
        copy_staging(o, c);
        dst->objects.push_back(c);
    }
    dst->h_instances = src->h_instances;
    return B2RT_OK;
}
extern "C" {

// ---- binary scene cache: the staged scene -> one file and back --------------------------------------
// Layout: "B2RTSCN2", then the struct sizes (a cache written by another build of the library is refused), then one block
// per context -- the top level first, then every object -- of length-prefixed arrays, then the instances.
}  // extern "C"
namespace {
const char SCENE_MAGIC[8] = {'B', '2', 'R', 'T', 'S', 'C', 'N', '2'};
template <typename T> bool put_vec(FILE *f, const std::vector<T> &v) {
    const uint64_t n = v.size();
    return fwrite(&n, 8, 1, f) == 1 && (n == 0 || fwrite(v.data(), sizeof(T), n, f) == n);
}
template <typename T> bool get_vec(FILE *f, std::vector<T> &v, uint64_t limit) {
    uint64_t n = 0;
    if (fread(&n, 8, 1, f) != 1 || n > limit) return false;
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}
bool put_staging(FILE *f, const b2rt_ctx *c) {
    const int32_t hdr[2] = {c->n_prims, c->any_attr ? 1 : 0};
    return fwrite(hdr, 4, 2, f) == 2 && put_vec(f, c->h_verts) && put_vec(f, c->h_idx) && put_vec(f, c->h_tri_prim) &&
           put_vec(f, c->h_vnormal) && put_vec(f, c->h_vuv) && put_vec(f, c->h_tri_flags) && put_vec(f, c->h_analytic) &&
           put_vec(f, c->h_analytic_prim) && put_vec(f, c->h_analytic_type) && put_vec(f, c->h_prim_material) &&
           put_vec(f, c->h_prim_light) && put_vec(f, c->h_materials) && put_vec(f, c->h_lights);
}
bool get_staging(FILE *f, b2rt_ctx *c) {
    int32_t hdr[2];
    const uint64_t L = 1ull << 33;
    if (fread(hdr, 4, 2, f) != 2 || hdr[0] < 0) return false;
    bool ok = get_vec(f, c->h_verts, L) && get_vec(f, c->h_idx, L) && get_vec(f, c->h_tri_prim, L) && get_vec(f, c->h_vnormal, L) &&
              get_vec(f, c->h_vuv, L) && get_vec(f, c->h_tri_flags, L) && get_vec(f, c->h_analytic, L) &&
              get_vec(f, c->h_analytic_prim, L) && get_vec(f, c->h_analytic_type, L) && get_vec(f, c->h_prim_material, L) &&
              get_vec(f, c->h_prim_light, L) && get_vec(f, c->h_materials, L) && get_vec(f, c->h_lights, L);
    if (!ok) return false;
    c->n_prims = hdr[0]; c->any_attr = hdr[1] != 0; c->built = false;
    // consistency: what the builder indexes must be in range
    if (c->h_idx.size() != 3 * c->h_tri_prim.size() || c->h_prim_material.size() != (size_t)c->n_prims ||
        c->h_prim_light.size() != (size_t)c->n_prims || c->h_analytic.size() != c->h_analytic_prim.size() ||
        c->h_analytic.size() != c->h_analytic_type.size() || c->h_verts.size() % 3 != 0)
        return false;
    const int64_t nv = (int64_t)c->h_verts.size() / 3;
    for (int32_t v : c->h_idx) if (v < 0 || v >= nv) return false;
    for (int32_t q : c->h_tri_prim) if (q < 0 || q >= c->n_prims) return false;
    for (int32_t q : c->h_analytic_prim) if (q < 0 || q >= c->n_prims) return false;
    if (c->any_attr && (c->h_vnormal.size() != c->h_verts.size() || c->h_vuv.size() != 2 * (size_t)nv || c->h_tri_flags.size() != c->h_tri_prim.size()))
        return false;
    return true;
}
}  // namespace
extern "C" {

int b2rt_scene_counts(b2rt_ctx *ctx, int32_t counts[5]) {
    if (!ctx || !counts) return B2RT_E_ARG;
    counts[0] = ctx->n_prims; counts[1] = (int32_t)ctx->h_lights.size(); counts[2] = (int32_t)ctx->h_materials.size();
    counts[3] = (int32_t)ctx->objects.size(); counts[4] = (int32_t)ctx->h_instances.size();
    return B2RT_OK;
}

int b2rt_scene_save(b2rt_ctx *ctx, const char *path) {
    if (!ctx || !path) return B2RT_E_ARG;
    if (ctx->cur_object) { ctx->err = "scene_save: b2rt_object_end is missing"; return B2RT_E_STATE; }
    FILE *f = fopen(path, "wb");
    if (!f) { ctx->err = std::string("scene_save: cannot open ") + path; return B2RT_E_ARG; }
    const uint32_t sizes[4] = {(uint32_t)sizeof(AnalyticPrim), (uint32_t)sizeof(Material), (uint32_t)sizeof(Light), (uint32_t)sizeof(b2rt_ctx::HostInstance)};
    const uint64_t nobj = ctx->objects.size();
    bool ok = fwrite(SCENE_MAGIC, 1, 8, f) == 8 && fwrite(sizes, 4, 4, f) == 4 && put_staging(f, ctx) && fwrite(&nobj, 8, 1, f) == 1;
    for (size_t o = 0; ok && o < ctx->objects.size(); ++o) ok = put_staging(f, ctx->objects[o]);
    ok = ok && put_vec(f, ctx->h_instances);
    ok = (fclose(f) == 0) && ok;
    if (!ok) { ctx->err = std::string("scene_save: write to ") + path + " failed"; return B2RT_E_ARG; }
    return B2RT_OK;
}

int b2rt_scene_load(b2rt_ctx *ctx, const char *path) {
    if (!ctx || !path) return B2RT_E_ARG;
    if (ctx->parent) { ctx->err = "scene_load: not on an object"; return B2RT_E_STATE; }
    FILE *f = fopen(path, "rb");
    if (!f) { ctx->err = std::string("scene_load: cannot open ") + path; return B2RT_E_ARG; }
    char magic[8]; uint32_t sizes[4]; uint64_t nobj = 0;
    const uint32_t want[4] = {(uint32_t)sizeof(AnalyticPrim), (uint32_t)sizeof(Material), (uint32_t)sizeof(Light), (uint32_t)sizeof(b2rt_ctx::HostInstance)};
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, SCENE_MAGIC, 8) == 0 && fread(sizes, 4, 4, f) == 4 && memcmp(sizes, want, 16) == 0;
    if (!ok) { fclose(f); ctx->err = std::string("scene_load: ") + path + " is not a scene cache of this library build"; return B2RT_E_ARG; }
    b2rt_clear_scene(ctx);
    ok = get_staging(f, ctx) && fread(&nobj, 8, 1, f) == 1 && nobj < (1u << 24);
    for (uint64_t o = 0; ok && o < nobj; ++o) {
        b2rt_ctx *c = new_object(ctx);
        if (!c) { fclose(f); return B2RT_E_CUDA; }
        ctx->objects.push_back(c);
        ok = get_staging(f, c);
    }
    ok = ok && get_vec(f, ctx->h_instances, 1ull << 28);
    fclose(f);
    for (size_t k = 0; ok && k < ctx->h_instances.size(); ++k) {
        const b2rt_ctx::HostInstance &H = ctx->h_instances[k];
        ok = H.object >= 0 && (size_t)H.object < ctx->objects.size() && H.item >= 0 && H.item < ctx->n_prims && ctx->objects[H.object]->n_prims > 0;
    }
    if (!ok) { b2rt_clear_scene(ctx); ctx->err = std::string("scene_load: ") + path + " is truncated or inconsistent"; return B2RT_E_ARG; }
    return B2RT_OK;
}

int b2rt_object_begin(b2rt_ctx *ctx, int32_t *object_id) {
    if (!ctx) return B2RT_E_ARG;
    if (ctx->parent || ctx->cur_object) { ctx->err = "object_begin: objects do not nest (pbrt: core/api.cpp:1107-1110)"; return B2RT_E_STATE; }
    b2rt_ctx *c = new_object(ctx);
    if (!c) return B2RT_E_CUDA;
    if (object_id) *object_id = (int32_t)ctx->objects.size();
    ctx->objects.push_back(c);
    ctx->cur_object = c;
    return B2RT_OK;
}

int b2rt_object_end(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    if (!ctx->cur_object) { ctx->err = "object_end: no object is open"; return B2RT_E_STATE; }
    ctx->cur_object = nullptr;
    return B2RT_OK;
}

int b2rt_add_instance(b2rt_ctx *ctx, int32_t object_id, const float o2w[12], const float *w2o, int32_t *instance_id) {
    if (!ctx) return B2RT_E_ARG;
    if (ctx->cur_object) { ctx->err = "add_instance: not inside an object definition (pbrt: core/api.cpp:1143-1146)"; return B2RT_E_STATE; }
    if (object_id < 0 || object_id >= (int32_t)ctx->objects.size() || !o2w) { ctx->err = "add_instance: bad arguments"; return B2RT_E_ARG; }
    if (ctx->objects[object_id]->n_prims == 0) { ctx->err = "add_instance: the object has no primitives"; return B2RT_E_ARG; }
    if ((int64_t)ctx->n_prims + 1 >= (1ll << 28)) { ctx->err = "add_instance: more than 2^28 items"; return B2RT_E_ARG; }
    b2rt_ctx::HostInstance H;
    memcpy(H.o2w, o2w, sizeof(H.o2w));
    if (w2o) memcpy(H.w2o, w2o, sizeof(H.w2o));
    else if (invert34(o2w, H.w2o)) { ctx->err = "add_instance: singular transform"; return B2RT_E_ARG; }
    H.object = object_id;
    H.item = ctx->n_prims++;
    ctx->h_prim_material.push_back(-1);
    ctx->h_prim_light.push_back(-1);
    if (instance_id) *instance_id = (int32_t)ctx->h_instances.size();
    ctx->h_instances.push_back(H);
    ctx->built = false;
    return B2RT_OK;
}

int b2rt_instance_info(b2rt_ctx *ctx, int32_t instance_id, int32_t *first_prim, int32_t *n_prims, int32_t *object_id) {
    if (!ctx) return B2RT_E_ARG;
    if (!ctx->built) { ctx->err = "instance_info: no acceleration structure built"; return B2RT_E_STATE; }
    if (instance_id < 0 || instance_id >= (int32_t)ctx->h_instances.size()) { ctx->err = "instance_info: bad instance id"; return B2RT_E_ARG; }
    const int o = ctx->h_instances[instance_id].object;
    if (first_prim) *first_prim = ctx->h_instance_base[instance_id];
    if (n_prims) *n_prims = ctx->objects[o]->n_prims;
    if (object_id) *object_id = o;
    return B2RT_OK;
}

void b2rt_destroy(b2rt_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->stream_in);
    cudaStreamSynchronize(ctx->stream_out);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1);
    cudaEventDestroy(ctx->evb0); cudaEventDestroy(ctx->evb1); cudaEventDestroy(ctx->evb_mid);
    cudaEventDestroy(ctx->ev_in); cudaEventDestroy(ctx->ev_k); cudaEventDestroy(ctx->ev_out);
    cudaEventDestroy(ctx->ev_s0); cudaEventDestroy(ctx->ev_s1);
    cudaStreamDestroy(ctx->stream_in); cudaStreamDestroy(ctx->stream_out);
    cudaStreamDestroy(ctx->stream);
    for (b2rt_ctx *c : ctx->objects) destroy_object(c);
    delete ctx;      // ~DevBuf frees every device buffer of the context and of its photon maps
}

const char *b2rt_last_error(b2rt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int b2rt_clear_scene(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    ctx->h_verts.clear(); ctx->h_idx.clear(); ctx->h_tri_prim.clear(); ctx->h_analytic.clear();
    ctx->h_analytic_prim.clear(); ctx->h_analytic_type.clear(); ctx->h_prim_material.clear();
    ctx->h_prim_light.clear(); ctx->h_materials.clear(); ctx->h_lights.clear();
    ctx->h_vnormal.clear(); ctx->h_vuv.clear(); ctx->h_tri_flags.clear(); ctx->any_attr = false;
    ctx->n_prims = 0; ctx->built = false;
    for (b2rt_ctx *c : ctx->objects) destroy_object(c);
    ctx->objects.clear(); ctx->cur_object = nullptr; ctx->h_instances.clear(); ctx->h_instance_base.clear(); ctx->n_virtual_prims = 0;
    return B2RT_OK;
}

int b2rt_add_mesh_ex(b2rt_ctx *ctx, int32_t nverts, const float *P, int32_t ntris, const int32_t *idx,
                     const float *N, const float *uv, int32_t material_id, int32_t arealight_id, int32_t flags,
                     int32_t *first_prim) {
    if (!ctx) return B2RT_E_ARG;
    if (ctx->cur_object) {              // inside b2rt_object_begin / _end: the mesh belongs to the object, ids are local to it
        if (arealight_id >= 0) { ctx->err = "add_mesh: area lights are not supported inside objects (pbrt: core/api.cpp:1009-1011)"; return B2RT_E_ARG; }
        int rc = b2rt_add_mesh_ex(ctx->cur_object, nverts, P, ntris, idx, N, uv, material_id, -1, flags, first_prim);
        if (rc) ctx->err = ctx->cur_object->err;
        return rc;
    }
    if (nverts < 0 || ntris < 0 || (nverts && !P) || (ntris && !idx)) { ctx->err = "add_mesh: bad sizes/pointers"; return B2RT_E_ARG; }
    if ((int64_t)ctx->n_prims + ntris >= (1ll << 28)) { ctx->err = "add_mesh: more than 2^28 primitives"; return B2RT_E_ARG; }
    if (flags & ~B2RT_MESH_REVERSE) { ctx->err = "add_mesh: unknown flag bits"; return B2RT_E_ARG; }
    for (int64_t k = 0; k < 3ll * ntris; ++k)
        if (idx[k] < 0 || idx[k] >= nverts) { ctx->err = "add_mesh: vertex index out of range"; return B2RT_E_ARG; }
    int32_t vbase = (int32_t)(ctx->h_verts.size() / 3);
    ctx->h_verts.insert(ctx->h_verts.end(), P, P + 3ll * nverts);
    // per-vertex normals / uv (K6, reference: cuda_render/util/shape/cudatrianglemesh.cpp:24-44 uploads them the same way)
    const uint8_t fl = (uint8_t)((N ? 1 : 0) | (uv ? 2 : 0) | ((flags & B2RT_MESH_REVERSE) ? 4 : 0));
    if (fl) ctx->any_attr = true;
    if (ctx->any_attr) {
        ctx->h_vnormal.resize(3 * (size_t)vbase, 0.f); ctx->h_vuv.resize(2 * (size_t)vbase, 0.f);
        if (N) ctx->h_vnormal.insert(ctx->h_vnormal.end(), N, N + 3ll * nverts);
        if (uv) ctx->h_vuv.insert(ctx->h_vuv.end(), uv, uv + 2ll * nverts);
        ctx->h_vnormal.resize(3 * (size_t)(vbase + nverts), 0.f); ctx->h_vuv.resize(2 * (size_t)(vbase + nverts), 0.f);
        ctx->h_tri_flags.resize(ctx->h_tri_prim.size(), 0);
    }
    size_t ib = ctx->h_idx.size();
    ctx->h_idx.resize(ib + 3ll * ntris);
    for (int64_t k = 0; k < 3ll * ntris; ++k) ctx->h_idx[ib + k] = idx[k] + vbase;
    if (first_prim) *first_prim = ctx->n_prims;
    for (int32_t t = 0; t < ntris; ++t) {
        ctx->h_tri_prim.push_back(ctx->n_prims++);
        ctx->h_prim_material.push_back(material_id);
        ctx->h_prim_light.push_back(arealight_id);
    }
    if (ctx->any_attr) ctx->h_tri_flags.resize(ctx->h_tri_prim.size(), fl);
    ctx->built = false;
    return B2RT_OK;
}

int b2rt_add_mesh(b2rt_ctx *ctx, int32_t nverts, const float *P, int32_t ntris, const int32_t *idx,
                  const float *N, const float *uv, int32_t material_id, int32_t arealight_id,
                  int32_t *first_prim) {
    return b2rt_add_mesh_ex(ctx, nverts, P, ntris, idx, N, uv, material_id, arealight_id, 0, first_prim);
}

// phimax in degrees as pbrt's parameter -> the record's value: radians after pbrt's clamp (Radians(Clamp(pm, 0, 360)),
// sphere.cpp:48, disk.cpp:46), 0 for the full turn
static float phimax_field(float phimax_deg) {
    const float pm = fminf(fmaxf(phimax_deg, 0.f), 360.f);
    return pm >= 360.f ? 0.f : ((float)M_PI / 180.f) * pm;
}

static int add_analytic(b2rt_ctx *ctx, int type, const float o2w[12], const float *w2o, float radius,
                        float height, float inner, int32_t material_id, int32_t arealight_id, int32_t *prim, float phimax = 0.f) {
    if (!ctx) return B2RT_E_ARG;
    if (ctx->cur_object) {
        if (arealight_id >= 0) { ctx->err = "add_sphere/disk: area lights are not supported inside objects (pbrt: core/api.cpp:1009-1011)"; return B2RT_E_ARG; }
        int rc = add_analytic(ctx->cur_object, type, o2w, w2o, radius, height, inner, material_id, -1, prim, phimax);
        if (rc) ctx->err = ctx->cur_object->err;
        return rc;
    }
    if (!o2w || !(radius > 0.f)) { ctx->err = "add_sphere/disk: bad arguments"; return B2RT_E_ARG; }
    AnalyticPrim a;
    memcpy(a.o2w, o2w, sizeof(a.o2w));
    if (w2o) memcpy(a.w2o, w2o, sizeof(a.w2o));
    else if (invert34(o2w, a.w2o)) { ctx->err = "add_sphere/disk: singular transform"; return B2RT_E_ARG; }
    a.radius = radius; a.height = height; a.inner = inner; a.pad = phimax;
    ctx->h_analytic.push_back(a);
    ctx->h_analytic_type.push_back(type);
    if (prim) *prim = ctx->n_prims;
    ctx->h_analytic_prim.push_back(ctx->n_prims++);
    ctx->h_prim_material.push_back(material_id);
    ctx->h_prim_light.push_back(arealight_id);
    ctx->built = false;
    return B2RT_OK;
}

int b2rt_add_sphere(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float radius, int32_t material_id,
                    int32_t arealight_id, int32_t *prim) {
    return add_analytic(ctx, B2RT_PRIM_SPHERE, o2w, w2o, radius, 0.f, 0.f, material_id, arealight_id, prim);
}

int b2rt_add_sphere_ex(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float radius, float zmin, float zmax,
                       float phimax_deg, int32_t material_id, int32_t arealight_id, int32_t *prim) {
    if (!ctx) return B2RT_E_ARG;
    if (!(radius > 0.f) || !(zmin == zmin) || !(zmax == zmax) || !(phimax_deg > 0.f)) { ctx->err = "add_sphere_ex: bad arguments"; return B2RT_E_ARG; }
    // Sphere::Sphere (shapes/sphere.cpp:40-49): zmin = Clamp(min(z0, z1), -radius, radius), zmax likewise
    float z0 = fminf(fmaxf(fminf(zmin, zmax), -radius), radius), z1 = fminf(fmaxf(fmaxf(zmin, zmax), -radius), radius);
    if (!(z0 < z1)) { ctx->err = "add_sphere_ex: zmin and zmax leave no surface"; return B2RT_E_ARG; }
    if (z0 <= -radius && z1 >= radius) z0 = z1 = 0.f;      // no z clipping (the record's "zmin >= zmax")
    return add_analytic(ctx, B2RT_PRIM_SPHERE, o2w, w2o, radius, z0, z1, material_id, arealight_id, prim, phimax_field(phimax_deg));
}

int b2rt_add_disk_ex(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float height, float radius, float inner_radius,
                     float phimax_deg, int32_t material_id, int32_t arealight_id, int32_t *prim) {
    if (!ctx) return B2RT_E_ARG;
    if (!(phimax_deg > 0.f)) { ctx->err = "add_disk_ex: bad arguments"; return B2RT_E_ARG; }
    if (arealight_id >= 0 && phimax_field(phimax_deg) > 0.f) { ctx->err = "add_disk_ex: an area light needs the full disk"; return B2RT_E_ARG; }
    return add_analytic(ctx, B2RT_PRIM_DISK, o2w, w2o, radius, height, inner_radius, material_id, arealight_id, prim, phimax_field(phimax_deg));
}

int b2rt_add_disk(b2rt_ctx *ctx, const float o2w[12], const float *w2o, float height, float radius,
                  float inner_radius, int32_t material_id, int32_t arealight_id, int32_t *prim) {
    return add_analytic(ctx, B2RT_PRIM_DISK, o2w, w2o, radius, height, inner_radius, material_id, arealight_id,
                        prim);
}

int b2rt_add_material(b2rt_ctx *ctx, int32_t type, const float rgb[3], const float rgb_kt[3], float ior,
                      int32_t *material_id) {
    if (!ctx) return B2RT_E_ARG;
    if (type < B2RT_MAT_MATTE || type > B2RT_MAT_GLASS || !rgb) { ctx->err = "add_material: bad arguments"; return B2RT_E_ARG; }
    Material m;
    m.type = type; memcpy(m.kd, rgb, 12);
    if (rgb_kt) memcpy(m.kt, rgb_kt, 12); else m.kt[0] = m.kt[1] = m.kt[2] = 1.f;
    m.ior = ior;
    if (material_id) *material_id = (int32_t)ctx->h_materials.size();
    ctx->h_materials.push_back(m);
    return B2RT_OK;
}

int b2rt_add_light(b2rt_ctx *ctx, int32_t type, const float p[3], const float rgb[3], int32_t prim,
                   int32_t nsamples, int32_t *light_id) {
    if (!ctx) return B2RT_E_ARG;
    if (type < B2RT_LIGHT_POINT || type > B2RT_LIGHT_DISK_AREA || !rgb) { ctx->err = "add_light: bad arguments"; return B2RT_E_ARG; }
    Light l;
    memset(&l, 0, sizeof(l));
    l.type = type;
    if (p) memcpy(l.p, p, 12); else l.p[0] = l.p[1] = l.p[2] = 0.f;
    memcpy(l.rgb, rgb, 12);
    l.prim = prim; l.nsamples = nsamples > 0 ? nsamples : 1;
    if (light_id) *light_id = (int32_t)ctx->h_lights.size();
    ctx->h_lights.push_back(l);
    return B2RT_OK;
}

int b2rt_add_spot_light(b2rt_ctx *ctx, const float l2w[12], const float w2l[12], const float rgb[3], float cos_total_width,
                        float cos_falloff_start, int32_t *light_id) {
    if (!ctx) return B2RT_E_ARG;
    if (!l2w || !w2l || !rgb || !(cos_total_width < cos_falloff_start || cos_total_width == cos_falloff_start)) {
        ctx->err = "add_spot_light: bad arguments"; return B2RT_E_ARG;
    }
    Light l;
    memset(&l, 0, sizeof(l));
    l.type = B2RT_LIGHT_SPOT;
    l.p[0] = l2w[3]; l.p[1] = l2w[7]; l.p[2] = l2w[11];                     // LightToWorld(Point(0, 0, 0))
    memcpy(l.rgb, rgb, 12);
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { l.l2w[3 * r + c] = l2w[4 * r + c]; l.w2l[3 * r + c] = w2l[4 * r + c]; }
    l.cos_total = cos_total_width; l.cos_falloff = cos_falloff_start;
    l.prim = -1; l.nsamples = 1; l.analytic = -1; l.n[2] = 1.f;
    if (light_id) *light_id = (int32_t)ctx->h_lights.size();
    ctx->h_lights.push_back(l);
    return B2RT_OK;
}

int b2rt_add_distant_light(b2rt_ctx *ctx, const float dir_to_light[3], const float rgb[3], int32_t *light_id) {
    if (!ctx) return B2RT_E_ARG;
    if (!dir_to_light || !rgb) { ctx->err = "add_distant_light: bad arguments"; return B2RT_E_ARG; }
    Light l;
    memset(&l, 0, sizeof(l));
    l.type = B2RT_LIGHT_DISTANT;
    memcpy(l.rgb, rgb, 12);
    memcpy(l.n, dir_to_light, 12);
    l.prim = -1; l.nsamples = 1; l.analytic = -1;
    if (light_id) *light_id = (int32_t)ctx->h_lights.size();
    ctx->h_lights.push_back(l);
    return B2RT_OK;
}

int b2rt_timings_enable(b2rt_ctx *ctx, int on) {
    if (!ctx) return B2RT_E_ARG;
    ctx->timings_on = on != 0;
    return B2RT_OK;
}

int b2rt_get_timings(b2rt_ctx *ctx, b2rt_timings *out, int reset) {
    if (!ctx || !out) return B2RT_E_ARG;
    *out = ctx->timings;
    if (reset) ctx->timings = b2rt_timings{};
    return B2RT_OK;
}

int b2rt_build_accel(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_BUILD);
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = b2rt_build_impl(ctx);
    if (rc) return rc;
    return b2rt_upload_tables(ctx);          // after the build: distant lights need the scene bounds
}

int b2rt_accel_info_get(b2rt_ctx *ctx, b2rt_accel_info *info) {
    if (!ctx || !info) return B2RT_E_ARG;
    if (!ctx->built) { ctx->err = "accel_info: no acceleration structure built"; return B2RT_E_STATE; }
    *info = ctx->info;
    return B2RT_OK;
}

int b2rt_accel_download(b2rt_ctx *ctx, uint32_t *morton, int32_t *prim_order, int32_t *bin_parent,
                        int32_t *bin_child, float *bin_box, uint32_t *wide_nodes, int32_t *leaf_prims) {
    if (!ctx) return B2RT_E_ARG;
    if (!ctx->built) { ctx->err = "accel_download: no acceleration structure built"; return B2RT_E_STATE; }
    const int64_t n = ctx->n_prims;
    cudaStream_t st = ctx->stream;
    auto dn = [&](void *dst, const DevBuf &src, size_t bytes) -> cudaError_t {
        if (!dst || !bytes) return cudaSuccess;
        return cudaMemcpyAsync(dst, src.p, bytes, cudaMemcpyDeviceToHost, st);
    };
    B2RT_CUDA(ctx, dn(morton, ctx->d_keys, n * 4));
    B2RT_CUDA(ctx, dn(prim_order, ctx->d_vals, n * 4));
    if (n >= 1) {
        B2RT_CUDA(ctx, dn(bin_parent, ctx->d_bin_parent, (2 * n - 1) * 4));
        B2RT_CUDA(ctx, dn(bin_child, ctx->d_bin_child, 2 * (n - 1) * 4));
        B2RT_CUDA(ctx, dn(bin_box, ctx->d_bin_box, (2 * n - 1) * 24));
    }
    B2RT_CUDA(ctx, dn(wide_nodes, ctx->d_wide_nodes, (size_t)ctx->info.n_wide_nodes * 128));
    B2RT_CUDA(ctx, dn(leaf_prims, ctx->d_leaf_prims, n * 4));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

static int check_status(b2rt_ctx *ctx) {
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

static int trace_common(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, b2rt_hit *hits, uint8_t *occ,
                        int where, bool any) {
    if (!ctx) return B2RT_E_ARG;
    if (nrays < 0 || (nrays && (!rays || (!hits && !occ)))) { ctx->err = "trace: bad arguments"; return B2RT_E_ARG; }
    if (!ctx->built) { ctx->err = "trace: b2rt_build_accel has not been called"; return B2RT_E_STATE; }
    if (nrays == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    if (where == B2RT_DEVICE) return b2rt_trace_impl(ctx, nrays, rays, hits, occ, any);
    if (where != B2RT_HOST) { ctx->err = "trace: where must be B2RT_HOST or B2RT_DEVICE"; return B2RT_E_ARG; }
    cudaStream_t st = ctx->stream;
    B2RT_CUDA(ctx, ctx->d_rays.reserve((size_t)nrays * sizeof(b2rt_ray)));
    if (any) B2RT_CUDA(ctx, ctx->d_occ.reserve((size_t)nrays));
    else B2RT_CUDA(ctx, ctx->d_hits.reserve((size_t)nrays * sizeof(b2rt_hit)));
    // Chunked pipeline over three streams: the upload of chunk i+1, the traversal of chunk i and the download of
    // chunk i-1 overlap (PCIe is full duplex; with pinned caller memory the copies are asynchronous).
    // The loop body is a lambda so that EVERY way out of it -- a failing CUDA call, a failing launch -- passes the join
    // below: copies that still read the caller's `rays` or write its `hits` / `occ` must have drained before the
    // caller sees the error and frees or reuses those buffers.
    const int64_t CH = ctx->host_chunk;
    auto pipeline = [&]() -> int {
        B2RT_CUDA(ctx, cudaEventRecord(ctx->ev_k, st));                 // everything queued so far on the main stream
        B2RT_CUDA(ctx, cudaStreamWaitEvent(ctx->stream_in, ctx->ev_k, 0));
        for (int64_t c0 = 0; c0 < nrays; c0 += CH) {
            const int64_t cn = nrays - c0 < CH ? nrays - c0 : CH;
            b2rt_ray *dr = ctx->d_rays.as<b2rt_ray>() + c0;
            B2RT_CUDA(ctx, cudaMemcpyAsync(dr, rays + c0, (size_t)cn * sizeof(b2rt_ray), cudaMemcpyHostToDevice, ctx->stream_in));
            B2RT_CUDA(ctx, cudaEventRecord(ctx->ev_in, ctx->stream_in));
            B2RT_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_in, 0));
            int rc = any ? b2rt_trace_impl(ctx, cn, dr, nullptr, ctx->d_occ.as<uint8_t>() + c0, true)
                         : b2rt_trace_impl(ctx, cn, dr, ctx->d_hits.as<b2rt_hit>() + c0, nullptr, false);
            if (rc) return rc;
            B2RT_CUDA(ctx, cudaEventRecord(ctx->ev_k, st));
            B2RT_CUDA(ctx, cudaStreamWaitEvent(ctx->stream_out, ctx->ev_k, 0));
            if (any)
                B2RT_CUDA(ctx, cudaMemcpyAsync(occ + c0, ctx->d_occ.as<uint8_t>() + c0, (size_t)cn, cudaMemcpyDeviceToHost, ctx->stream_out));
            else
                B2RT_CUDA(ctx, cudaMemcpyAsync(hits + c0, ctx->d_hits.as<b2rt_hit>() + c0, (size_t)cn * sizeof(b2rt_hit),
                                               cudaMemcpyDeviceToHost, ctx->stream_out));
        }
        B2RT_CUDA(ctx, cudaEventRecord(ctx->ev_out, ctx->stream_out));
        B2RT_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_out, 0));
        return B2RT_OK;
    };
    const int prc = pipeline();
    if (prc) {
        const std::string msg = ctx->err;
        cudaStreamSynchronize(ctx->stream_in); cudaStreamSynchronize(st); cudaStreamSynchronize(ctx->stream_out);
        ctx->err = msg;
        return prc;
    }
    return check_status(ctx);
}

int b2rt_trace_closest(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, b2rt_hit *hits, int where) {
    B2RT_STAGE(ctx, B2RT_STAGE_TRACE_CLOSEST);
    return trace_common(ctx, nrays, rays, hits, nullptr, where, false);
}
int b2rt_trace_any(b2rt_ctx *ctx, int64_t nrays, const b2rt_ray *rays, uint8_t *occluded, int where) {
    B2RT_STAGE(ctx, B2RT_STAGE_TRACE_ANY);
    return trace_common(ctx, nrays, rays, nullptr, occluded, where, true);
}

// -DB2RT_STATS build only (make -C cuda_raytrace_b200/csrc stats -> libb2rt_stats.so): wide nodes fetched and primitive
// records tested by the traversal kernel since the last reset; SURVEY 8(d) asks for the L2 -> SM bytes per ray they imply
// (nodes * 128 + prims * 48).  The regular build returns B2RT_E_STATE.
int b2rt_trace_stats(b2rt_ctx *ctx, uint64_t *nodes, uint64_t *prims, int reset) {
    if (!ctx) return B2RT_E_ARG;
#ifdef B2RT_STATS
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned long long v[2] = {0, 0};
    B2RT_CUDA(ctx, cudaMemcpyAsync(v, ctx->d_counters.as<uint32_t>() + 24, 16, cudaMemcpyDeviceToHost, ctx->stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (reset) B2RT_CUDA(ctx, cudaMemsetAsync(ctx->d_counters.as<uint32_t>() + 24, 0, 16, ctx->stream));
    if (nodes) *nodes = v[0];
    if (prims) *prims = v[1];
    return B2RT_OK;
#else
    (void)nodes; (void)prims; (void)reset;
    ctx->err = "trace_stats: this library was built without -DB2RT_STATS (make -C cuda_raytrace_b200/csrc stats)";
    return B2RT_E_STATE;
#endif
}

int b2rt_sort_pairs(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, int key_bits) {
    if (!ctx) return B2RT_E_ARG;
    if (n < 0 || key_bits < 1 || key_bits > 32 || (n && (!keys || !vals))) { ctx->err = "sort_pairs: bad arguments"; return B2RT_E_ARG; }
    if (n <= 1) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    B2RT_CUDA(ctx, ctx->d_keys_alt.reserve((size_t)n * 4));
    B2RT_CUDA(ctx, ctx->d_vals_alt.reserve((size_t)n * 4));
    bool in_alt = false;
    int rc = b2rt_sort_pairs_impl(ctx, n, keys, vals, ctx->d_keys_alt.as<uint32_t>(), ctx->d_vals_alt.as<uint32_t>(),
                                  key_bits, &in_alt);
    if (rc) return rc;
    if (in_alt) {
        B2RT_CUDA(ctx, cudaMemcpyAsync(keys, ctx->d_keys_alt.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        B2RT_CUDA(ctx, cudaMemcpyAsync(vals, ctx->d_vals_alt.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return B2RT_OK;
}

void *b2rt_host_alloc(b2rt_ctx *ctx, int64_t bytes) {
    if (!ctx || bytes <= 0) return nullptr;
    void *p = nullptr;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocPortable) != cudaSuccess) {
        ctx->err = std::string("host_alloc: ") + cudaGetErrorString(cudaGetLastError());
        return nullptr;
    }
    return p;
}

void b2rt_host_free(b2rt_ctx *ctx, void *p) {
    (void)ctx;
    if (p) cudaFreeHost(p);
}

int b2rt_sync(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return check_status(ctx);
}

int b2rt_timer_start(b2rt_ctx *ctx) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));        // a caller that drives several contexts may have another device current
    B2RT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return B2RT_OK;
}
int b2rt_timer_stop(b2rt_ctx *ctx, float *ms) {
    if (!ctx || !ms) return B2RT_E_ARG;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    B2RT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    B2RT_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    B2RT_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return B2RT_OK;
}
int b2rt_launch_count(b2rt_ctx *ctx, int64_t *n, int reset) {
    if (!ctx) return B2RT_E_ARG;
    if (n) *n = ctx->launches;
    if (reset) ctx->launches = 0;
    return B2RT_OK;
}
void *b2rt_stream(b2rt_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

}  // extern "C"
