// K8 + K9: photon emission and tracing, deterministic compaction; K10 + K11: photon hash-grid build.
//
// Replaces photontracing_camera / photontracing_closest_hit (cuda_render/photon_mapping/photontracing.cu:80-185),
// the cuRAND buffer they read (photonmappingrenderer.cpp:182-221) and the host-side kd-tree build
// CreatePhotonMap (photonmappingrenderer.cpp:150-180).  Path logic follows pbrt's PhotonShootingTask::Run
// (pbrt-v2/src/integrators/photonmap.cpp:416-514) with finalgather off: caustic deposit = first diffuse hit of an
// all-specular path with >1 intersections, indirect deposit = any later diffuse hit, Russian roulette on the
// luminance ratio, maxphotondepth.
//
// One thread follows one path from emission to termination with the traversal inlined (no wavefront queues: the
// bounce state lives in registers).  Randomness is Philox(key = seed, counter = (path, bounce)),
// so a path's photons depend only on (seed, path index).  Deposits go to per-path slots, then an exclusive scan
// over the per-path counts places them in path order: the photon array is byte-identical for any launch shape and
// any split of the path range over GPUs.
#include "b2rt_shade.cuh"

namespace {

constexpr int PHOTON_THREADS = 128;

// octahedral encoding of a unit vector into 2 x snorm16; deterministic (+,*,/,floor only)
__device__ __forceinline__ uint32_t oct_encode(f3 v) {
    float l1 = fabsf(v.x) + fabsf(v.y) + fabsf(v.z);
    float x = v.x / l1, y = v.y / l1;
    if (v.z < 0.f) {
        float ox = (1.f - fabsf(y)) * (x >= 0.f ? 1.f : -1.f);
        float oy = (1.f - fabsf(x)) * (y >= 0.f ? 1.f : -1.f);
        x = ox; y = oy;
    }
    int qx = (int)floorf(fminf(fmaxf(x, -1.f), 1.f) * 32767.f + 0.5f);
    int qy = (int)floorf(fminf(fmaxf(y, -1.f), 1.f) * 32767.f + 0.5f);
    return ((uint32_t)qx & 0xffffu) | ((uint32_t)qy << 16);
}

// K8+K9.  slots[(path-first)*max_depth + j]: j-th deposit of the path; cnt_c / cnt_i: how many of them are caustic /
// indirect (a path deposits at most one caustic photon and it is always its first deposit, because specularPath
// can only go from true to false).
template <bool INST>
__global__ void __launch_bounds__(PHOTON_THREADS)
photon_trace_kernel(DScene sc, uint32_t seed_lo, uint32_t seed_hi, int64_t first_path, int64_t n_paths,
                    int max_depth, int flags, b2rt_photon *__restrict__ slots, uint32_t *__restrict__ cnt_c,
                    uint32_t *__restrict__ cnt_i, uint32_t *status) {
    int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_paths) return;
    const uint64_t path = (uint64_t)(first_path + li);
    const uint32_t p_lo = (uint32_t)path, p_hi = (uint32_t)(path >> 32);
    const bool want_c = flags & 1, want_i = flags & 2;
    uint32_t nc = 0, ni = 0;
    b2rt_photon *my = slots + li * max_depth;

    // emission: 6 dims -> blocks 0 and 1 of bounce 0   (u[0] light, u[1..2] position, u[3] component (unused: one
    // shape per light), u[4..5] direction)
    u4 r0 = philox4x32_10(p_lo, p_hi, 0u, 0u, seed_lo, seed_hi);
    u4 r1 = philox4x32_10(p_lo, p_hi, 0u, 1u, seed_lo, seed_hi);
    f3 o, d, alpha;
    float tmin;
    bool alive = emit_photon(sc, u01(r0.x), u01(r0.y), u01(r0.z), u01(r1.x), u01(r1.y), o, d, tmin, alpha);
    bool specular_path = true;
    int n_isect = 0;
    bool overflow = false;
    while (alive) {
        HitD h;
        if (!traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, o.x, o.y, o.z, d.x, d.y, d.z, tmin, INFINITY, h, overflow))
            break;
        ++n_isect;
        Surface s = make_surface<INST>(sc, h, o, d);
        Mat m = load_material(sc, s.material);
        f3 wo = -d;
        if (m.type == B2RT_MAT_MATTE) {     // hasNonSpecular
            bool caustic = specular_path && n_isect > 1;
            bool dep = caustic ? want_c : (n_isect > 1 && want_i);
            if (dep) {
                b2rt_photon ph;
                ph.p[0] = s.p.x; ph.p[1] = s.p.y; ph.p[2] = s.p.z; ph.wi_oct = oct_encode(wo);
                ph.alpha[0] = alpha.x; ph.alpha[1] = alpha.y; ph.alpha[2] = alpha.z; ph.path = p_lo;
                float4 *dst = reinterpret_cast<float4 *>(my + nc + ni);
                dst[0] = make_float4(ph.p[0], ph.p[1], ph.p[2], __uint_as_float(ph.wi_oct));
                dst[1] = make_float4(ph.alpha[0], ph.alpha[1], ph.alpha[2], __uint_as_float(ph.path));
                if (caustic) ++nc; else ++ni;
            }
        }
        if (n_isect >= max_depth) break;
        u4 rb = philox4x32_10(p_lo, p_hi, (uint32_t)n_isect, 0u, seed_lo, seed_hi);
        BsdfSample bs;
        if (!sample_bsdf(s, m, wo, u01(rb.x), u01(rb.y), u01(rb.z), bs)) break;
        if (is_black(bs.f) || bs.pdf == 0.f) break;
        // anew = alpha * fr * AbsDot(wi, nn) / pdf
        f3 anew = mk3(alpha.x * bs.f.x * bs.absdot / bs.pdf, alpha.y * bs.f.y * bs.absdot / bs.pdf,
                      alpha.z * bs.f.z * bs.absdot / bs.pdf);
        float cont = fminf(1.f, lum(anew) / lum(alpha));
        if (u01(rb.w) > cont) break;
        alpha = mk3(anew.x / cont, anew.y / cont, anew.z / cont);
        specular_path = specular_path && bs.specular;
        if (!want_i && !specular_path) break;      // indirectDone && !specularPath
        o = s.p; d = bs.wi; tmin = s.ray_eps;
    }
    cnt_c[li] = nc;
    cnt_i[li] = ni;
    if (overflow) atomicAdd(status, 1u);
}

// Path-regeneration variant of K8: a resident grid, every lane always owns a live path.  Each loop iteration is one
// bounce; lanes whose path ended take the next path index from a global pool (ranks from __ballot_sync + __popc, one
// atomicAdd per warp) and emit there, so the traversal and shading of an iteration run with (nearly) all 32 lanes
// instead of the lanes that happen to survive to bounce b.  A path's randomness and its slot row depend only on its
// index, so the output is byte-identical to photon_trace_kernel.
// resident blocks per SM the kernel is compiled for: 8 (64 registers, 44 B of spills) instead of ptxas' own 80 registers --
// Cfg3 shoot 10.5 -> 10.1 ms; 5 / 7 / 9 / 10 / 12 blocks: 11.1 / 10.4 / 10.1 / 10.4 / 10.6 ms (r07t, r07u)
#ifndef B2RT_PHOTON_BLOCKS
#define B2RT_PHOTON_BLOCKS 8
#endif
template <bool INST>
__global__ void __launch_bounds__(PHOTON_THREADS, B2RT_PHOTON_BLOCKS)
photon_trace_regen_kernel(DScene sc, uint32_t seed_lo, uint32_t seed_hi, int64_t first_path, int64_t n_paths,
                          int max_depth, int flags, b2rt_photon *__restrict__ slots, uint32_t *__restrict__ cnt_c,
                          uint32_t *__restrict__ cnt_i, unsigned long long *pool, uint32_t *status) {
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const bool want_c = flags & 1, want_i = flags & 2;
    bool alive = false, exhausted = false, overflow = false;
    int64_t li = 0;
    uint32_t p_lo = 0, p_hi = 0, nc = 0, ni = 0;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f), alpha = mk3(0.f, 0.f, 0.f);
    float tmin = 0.f;
    bool specular_path = true;
    int n_isect = 0;
    for (;;) {
        const uint32_t idle = __ballot_sync(0xffffffffu, !alive);
        if (idle && !exhausted) {
            const int want = __popc(idle);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(pool, (unsigned long long)want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (!alive) {
                const int64_t i = (int64_t)base + __popc(idle & lt);
                if (i < n_paths) {
                    li = i;
                    const uint64_t path = (uint64_t)(first_path + li);
                    p_lo = (uint32_t)path; p_hi = (uint32_t)(path >> 32);
                    nc = 0; ni = 0; specular_path = true; n_isect = 0;
                    u4 r0 = philox4x32_10(p_lo, p_hi, 0u, 0u, seed_lo, seed_hi);
                    u4 r1 = philox4x32_10(p_lo, p_hi, 0u, 1u, seed_lo, seed_hi);
                    alive = emit_photon(sc, u01(r0.x), u01(r0.y), u01(r0.z), u01(r1.x), u01(r1.y), o, d, tmin, alpha);
                    if (!alive) { cnt_c[li] = 0; cnt_i[li] = 0; }
                }
            }
            if ((int64_t)base + want >= n_paths) exhausted = true;
        }
        if (!__any_sync(0xffffffffu, alive)) {
            if (exhausted) break;
            continue;
        }
        if (alive) {
            // one bounce
            bool cont_path = false;
            HitD h;
            if (traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, o.x, o.y, o.z, d.x, d.y, d.z, tmin, INFINITY, h, overflow)) {
                ++n_isect;
                Surface s = make_surface<INST>(sc, h, o, d);
                Mat m = load_material(sc, s.material);
                f3 wo = -d;
                if (m.type == B2RT_MAT_MATTE) {
                    bool caustic = specular_path && n_isect > 1;
                    bool dep = caustic ? want_c : (n_isect > 1 && want_i);
                    if (dep) {
                        float4 *dst = reinterpret_cast<float4 *>(slots + li * max_depth + nc + ni);
                        dst[0] = make_float4(s.p.x, s.p.y, s.p.z, __uint_as_float(oct_encode(wo)));
                        dst[1] = make_float4(alpha.x, alpha.y, alpha.z, __uint_as_float(p_lo));
                        if (caustic) ++nc; else ++ni;
                    }
                }
                if (n_isect < max_depth) {
                    u4 rb = philox4x32_10(p_lo, p_hi, (uint32_t)n_isect, 0u, seed_lo, seed_hi);
                    BsdfSample bs;
                    if (sample_bsdf(s, m, wo, u01(rb.x), u01(rb.y), u01(rb.z), bs) && !is_black(bs.f) && bs.pdf != 0.f) {
                        f3 anew = mk3(alpha.x * bs.f.x * bs.absdot / bs.pdf, alpha.y * bs.f.y * bs.absdot / bs.pdf,
                                      alpha.z * bs.f.z * bs.absdot / bs.pdf);
                        float cont = fminf(1.f, lum(anew) / lum(alpha));
                        if (!(u01(rb.w) > cont)) {
                            alpha = mk3(anew.x / cont, anew.y / cont, anew.z / cont);
                            specular_path = specular_path && bs.specular;
                            if (want_i || specular_path) {
                                o = s.p; d = bs.wi; tmin = s.ray_eps;
                                cont_path = true;
                            }
                        }
                    }
                }
            }
            if (!cont_path) {
                cnt_c[li] = nc; cnt_i[li] = ni;
                alive = false;
            }
        }
    }
    if (overflow) atomicAdd(status, 1u);
}

__global__ void photon_pool_reset_kernel(unsigned long long *pool) { *pool = 0ull; }

// place the deposits in path order: out_c[off_c[path] + 0..nc), out_i[off_i[path] + 0..ni).
// The compaction IS the exchange step of the multi-GPU photon pass: dst holds the position of this GPU's slice in the
// photon arrays of EVERY GPU of the box (dst.n of them; peer pointers, the stores travel over NVLink / NVSwitch), so the
// photon maps are assembled by the kernel that produces them -- no send buffer, no collective afterwards, no copies.
__global__ void __launch_bounds__(256)
photon_compact_kernel(const b2rt_photon *__restrict__ slots, const uint32_t *__restrict__ off_c,
                      const uint32_t *__restrict__ off_i, const uint32_t *__restrict__ tot, int64_t n_paths,
                      int max_depth, const PhotonDst dst) {
    int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_paths) return;
    uint32_t oc = off_c[li], oi = off_i[li];
    uint32_t nc = ((li + 1 < n_paths) ? off_c[li + 1] : tot[0]) - oc;
    uint32_t ni = ((li + 1 < n_paths) ? off_i[li + 1] : tot[1]) - oi;
    const float4 *src = reinterpret_cast<const float4 *>(slots + li * max_depth);
    for (uint32_t j = 0; j < nc; ++j) {
        const float4 a = src[2 * j], b = src[2 * j + 1];
        for (int g = 0; g < dst.n; ++g) {
            float4 *d = reinterpret_cast<float4 *>(dst.a[g] + oc + j);
            d[0] = a; d[1] = b;
        }
    }
    for (uint32_t j = 0; j < ni; ++j) {
        const float4 a = src[2 * (nc + j)], b = src[2 * (nc + j) + 1];
        for (int g = 0; g < dst.n; ++g) {
            float4 *d = reinterpret_cast<float4 *>(dst.b[g] + oi + j);
            d[0] = a; d[1] = b;
        }
    }
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------------------------
void b2rt_fill_dscene(b2rt_ctx *ctx, DScene *sc) {
    sc->nodes = ctx->d_wide_nodes.as<WideNode>();
    sc->recs = ctx->d_prim_recs.as<PrimRec>();
    sc->analytic = ctx->d_analytic.as<AnalyticPrim>();
    sc->prim_material = ctx->d_prim_material.as<int32_t>();
    sc->prim_light = ctx->d_prim_light.as<int32_t>();
    sc->materials = ctx->d_materials.as<Material>();
    sc->lights = ctx->d_lights.as<Light>();
    sc->light_cdf = ctx->d_light_cdf.as<float>();
    sc->n_lights = (int32_t)ctx->h_lights.size();
    sc->n_materials = (int32_t)ctx->h_materials.size();
    sc->tri_flags = ctx->any_attr ? ctx->d_tri_flags.as<uint8_t>() : nullptr;
    sc->prim_kind = ctx->d_prim_kind.as<int32_t>();
    sc->tri_idx = ctx->d_idx.as<int32_t>();
    sc->verts = ctx->d_verts.as<float>();
    sc->vnormal = ctx->d_vnormal.as<float>();
    sc->vuv = ctx->d_vuv.as<float>();
    sc->instances = ctx->h_instances.empty() ? nullptr : ctx->d_instances.as<InstanceRec>();
    sc->n_instances = (int32_t)ctx->h_instances.size();
}

// material / light tables -> device (called from b2rt_build_accel).  Light power CDF as pbrt's
// ComputeLightSamplingCDF + Distribution1D (core/integrator.cpp:261-268, core/montecarlo.h).
int b2rt_upload_tables(b2rt_ctx *ctx) {
    cudaStream_t st = ctx->stream;
    auto up = [&](DevBuf &b, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = b.reserve(bytes ? bytes : 16);
        if (e != cudaSuccess || !bytes) return e;
        return cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st);
    };
    const size_t nl = ctx->h_lights.size();
    std::vector<float> power(nl), cdf(2 * nl + 2, 0.f);
    for (size_t i = 0; i < nl; ++i) {
        Light &L = ctx->h_lights[i];
        L.analytic = -1; L.area = 0.f;
        if (L.type != B2RT_LIGHT_DISTANT) { L.n[0] = L.n[1] = 0.f; L.n[2] = 1.f; }
        float P[3];
        if (L.type == B2RT_LIGHT_POINT) {
            float f = (float)(4.f * M_PI);
            for (int c = 0; c < 3; ++c) P[c] = f * L.rgb[c];
        } else if (L.type == B2RT_LIGHT_SPOT) {                  // SpotLight::Power (lights/spot.cpp:72-75)
            float f = (float)M_PI, g = 1.f - .5f * (L.cos_falloff + L.cos_total);
            for (int c = 0; c < 3; ++c) P[c] = L.rgb[c] * 2.f * f * g;
        } else if (L.type == B2RT_LIGHT_DISTANT) {               // DistantLight::Power over BBox::BoundingSphere of the scene
            const float *b = ctx->info.bounds;
            float d2 = 0.f;
            for (int a = 0; a < 3; ++a) { L.wc[a] = .5f * b[a] + .5f * b[3 + a]; float d = L.wc[a] - b[3 + a]; d2 += d * d; }
            L.wr = sqrtf(d2);
            float f = (float)M_PI;
            for (int c = 0; c < 3; ++c) P[c] = L.rgb[c] * f * L.wr * L.wr;
        } else {
            int a = -1;
            for (size_t k = 0; k < ctx->h_analytic_prim.size(); ++k)
                if (ctx->h_analytic_prim[k] == L.prim && ctx->h_analytic_type[k] == B2RT_PRIM_DISK) a = (int)k;
            if (a < 0) { ctx->err = "light " + std::to_string(i) + ": area light primitive is not a disk"; return B2RT_E_ARG; }
            const AnalyticPrim &D = ctx->h_analytic[a];
            float phimax = ((float)M_PI / 180.f) * 360.f;
            L.analytic = a;
            L.area = phimax * 0.5f * (D.radius * D.radius - D.inner * D.inner);
            float nx = D.w2o[8], ny = D.w2o[9], nz = D.w2o[10];   // w2o^T (0,0,1)
            float inv = 1.f / sqrtf(nx * nx + ny * ny + nz * nz);
            L.n[0] = nx * inv; L.n[1] = ny * inv; L.n[2] = nz * inv;
            float pi = (float)M_PI;
            for (int c = 0; c < 3; ++c) P[c] = L.rgb[c] * L.area * pi;
        }
        power[i] = 0.212671f * P[0] + 0.715160f * P[1] + 0.072169f * P[2];
    }
    if (nl) {
        for (size_t i = 1; i <= nl; ++i) cdf[i] = cdf[i - 1] + power[i - 1] / (float)nl;
        float func_int = cdf[nl];
        for (size_t i = 1; i <= nl; ++i) cdf[i] = func_int > 0.f ? cdf[i] / func_int : (float)i / (float)nl;
        for (size_t i = 0; i < nl; ++i)
            cdf[nl + 1 + i] = func_int > 0.f ? power[i] / (func_int * (float)nl) : 1.f / (float)nl;
    }
    B2RT_CUDA(ctx, up(ctx->d_prim_material, ctx->h_prim_material.data(), ctx->h_prim_material.size() * 4));
    B2RT_CUDA(ctx, up(ctx->d_prim_light, ctx->h_prim_light.data(), ctx->h_prim_light.size() * 4));
    B2RT_CUDA(ctx, up(ctx->d_materials, ctx->h_materials.data(), ctx->h_materials.size() * sizeof(Material)));
    B2RT_CUDA(ctx, up(ctx->d_lights, ctx->h_lights.data(), nl * sizeof(Light)));
    B2RT_CUDA(ctx, up(ctx->d_light_cdf, cdf.data(), cdf.size() * sizeof(float)));
    // scene views of the objects for the shade stage of instanced hits: their own geometry / attribute / per-primitive
    // material arrays, this context's material table (InstanceRec::object points into d_objects, reserved by the build)
    std::vector<DScene> osc(ctx->objects.size());
    for (size_t o = 0; o < ctx->objects.size(); ++o) {
        b2rt_fill_dscene(ctx->objects[o], &osc[o]);
        osc[o].materials = ctx->d_materials.as<Material>(); osc[o].n_materials = (int32_t)ctx->h_materials.size();
        osc[o].prim_light = nullptr; osc[o].lights = nullptr; osc[o].light_cdf = nullptr; osc[o].n_lights = 0;
    }
    if (!osc.empty()) B2RT_CUDA(ctx, up(ctx->d_objects, osc.data(), osc.size() * sizeof(DScene)));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}

// grow a photon array keeping its first `keep` photons
int b2rt_photons_reserve(b2rt_ctx *ctx, DevBuf &b, int64_t keep, int64_t want) {
    size_t need = (size_t)want * sizeof(b2rt_photon);
    if (need <= b.cap) return B2RT_OK;
    size_t ncap = need + need / 2 + 4096;
    void *np = nullptr;
    B2RT_CUDA(ctx, cudaMalloc(&np, ncap));
    if (keep > 0)
        B2RT_CUDA(ctx, cudaMemcpyAsync(np, b.p, (size_t)keep * sizeof(b2rt_photon), cudaMemcpyDeviceToDevice, ctx->stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (b.p) cudaFree(b.p);
    b.p = np; b.cap = ncap;
    return B2RT_OK;
}

// B2RT_MGPU_EXCHANGE=push: the slice was compacted into this GPU's own array; a grid-stride copy stores it into the same
// place of the peers' arrays, 16 bytes per thread and destination: whole 512-byte lines per warp over NVLink where the
// compaction kernel's peer stores are 32-byte pieces at path-dependent offsets.
__global__ void __launch_bounds__(256)
photon_push_kernel(const float4 *__restrict__ src, int64_t n16, const PhotonDst dst) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = src[i];
        for (int g = 0; g < dst.n; ++g) reinterpret_cast<float4 *>(dst.a[g])[i] = v;
    }
}
int b2rt_photons_push(b2rt_ctx *ctx, const b2rt_photon *src, int64_t n_photons, const PhotonDst &dst) {
    if (n_photons <= 0 || dst.n <= 0) return B2RT_OK;
    const int64_t n16 = n_photons * 2;
    const int64_t blocks = std::min<int64_t>((n16 + 255) / 256, (int64_t)ctx->sm_count * 8);
    photon_push_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(reinterpret_cast<const float4 *>(src), n16, dst);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

// Photon pass in two steps so that several GPUs can meet in the middle (multi.cu): `begin` traces the paths into per-path
// slots, scans the per-path counts and returns how many photons each map gains; `commit` compacts them to wherever the
// caller says -- the local arrays (single GPU) or this GPU's slice of every GPU's arrays.
int b2rt_photons_trace_begin(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int flags, uint32_t tot[2]) {
    cudaStream_t st = ctx->stream;
    tot[0] = tot[1] = 0;
    if (n_paths == 0) return B2RT_OK;
    if (n_paths > (1ll << 30)) { ctx->err = "photons_trace: at most 2^30 paths per call"; return B2RT_E_ARG; }
    DScene sc;
    b2rt_fill_dscene(ctx, &sc);
    B2RT_CUDA(ctx, ctx->d_slots.reserve((size_t)n_paths * max_depth * sizeof(b2rt_photon)));
    B2RT_CUDA(ctx, ctx->d_cnt_c.reserve((size_t)n_paths * 4));
    B2RT_CUDA(ctx, ctx->d_cnt_i.reserve((size_t)n_paths * 4));
    uint32_t *cnt_c = ctx->d_cnt_c.as<uint32_t>(), *cnt_i = ctx->d_cnt_i.as<uint32_t>();
    uint32_t *counters = ctx->d_counters.as<uint32_t>();
    unsigned grid = (unsigned)((n_paths + PHOTON_THREADS - 1) / PHOTON_THREADS);
    if (ctx->photon_variant == 1) {
        if (ctx->photon_blocks_per_sm == 0) {
            B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->photon_blocks_per_sm, photon_trace_regen_kernel<false>,
                                                                         PHOTON_THREADS, 0));
            if (ctx->photon_blocks_per_sm < 1) ctx->photon_blocks_per_sm = 1;
        }
        unsigned long long *pool = reinterpret_cast<unsigned long long *>(counters + 18);
        photon_pool_reset_kernel<<<1, 1, 0, st>>>(pool);
        B2RT_LAUNCH_CHECK(ctx);
        unsigned pgrid = (unsigned)std::min<int64_t>(grid, (int64_t)ctx->sm_count * ctx->photon_blocks_per_sm);
        (sc.instances ? photon_trace_regen_kernel<true> : photon_trace_regen_kernel<false>)<<<pgrid, PHOTON_THREADS, 0, st>>>(sc, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32),
                                                                   first_path, n_paths, max_depth, flags,
                                                                   ctx->d_slots.as<b2rt_photon>(), cnt_c, cnt_i, pool,
                                                                   counters + 8);
    } else {
        (sc.instances ? photon_trace_kernel<true> : photon_trace_kernel<false>)<<<grid, PHOTON_THREADS, 0, st>>>(sc, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32), first_path,
                                                             n_paths, max_depth, flags, ctx->d_slots.as<b2rt_photon>(),
                                                             cnt_c, cnt_i, counters + 8);
    }
    B2RT_LAUNCH_CHECK(ctx);
    int rc = b2rt_exclusive_scan_u32(ctx, cnt_c, n_paths, counters + 10);
    if (rc) return rc;
    rc = b2rt_exclusive_scan_u32(ctx, cnt_i, n_paths, counters + 11);
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaMemcpyAsync(tot, counters + 10, 8, cudaMemcpyDeviceToHost, st));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}
int b2rt_photons_trace_commit(b2rt_ctx *ctx, int64_t n_paths, int max_depth, const PhotonDst &dst) {
    if (n_paths == 0) return B2RT_OK;
    photon_compact_kernel<<<(unsigned)((n_paths + 255) / 256), 256, 0, ctx->stream>>>(
        ctx->d_slots.as<b2rt_photon>(), ctx->d_cnt_c.as<uint32_t>(), ctx->d_cnt_i.as<uint32_t>(),
        ctx->d_counters.as<uint32_t>() + 10, n_paths, max_depth, dst);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

int b2rt_photons_trace_impl(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int flags,
                            int64_t n_stored[2]) {
    if (n_stored) n_stored[0] = n_stored[1] = 0;
    if (n_paths == 0) return B2RT_OK;
    uint32_t tot[2];
    int rc = b2rt_photons_trace_begin(ctx, first_path, n_paths, max_depth, flags, tot);
    if (rc) return rc;
    PhotonMap &mc = ctx->maps[B2RT_MAP_CAUSTIC], &mi = ctx->maps[B2RT_MAP_INDIRECT];
    rc = b2rt_photons_reserve(ctx, mc.photons, mc.n, mc.n + tot[0]);
    if (rc) return rc;
    rc = b2rt_photons_reserve(ctx, mi.photons, mi.n, mi.n + tot[1]);
    if (rc) return rc;
    PhotonDst dst;
    dst.n = 1; dst.a[0] = mc.photons.as<b2rt_photon>() + mc.n; dst.b[0] = mi.photons.as<b2rt_photon>() + mi.n;
    rc = b2rt_photons_trace_commit(ctx, n_paths, max_depth, dst);
    if (rc) return rc;
    mc.n += tot[0]; mi.n += tot[1];
    if (flags & 1) mc.n_paths += n_paths;
    if (flags & 2) mi.n_paths += n_paths;
    mc.grid_built = mi.grid_built = false;
    if (n_stored) { n_stored[0] = tot[0]; n_stored[1] = tot[1]; }
    return B2RT_OK;
}

// ---- K10 + K11: hash grid -----------------------------------------------------------------------------------------
// Bit-exact definition (restated in oracle/oracle_port.c: port_grid_*):
//   lo[a], hi[a]  = min / max photon coordinate per axis
//   cell c, inv = 1.f / c ; dims[a] = clamp((int)floorf((hi[a]-lo[a])*inv) + 1, 1, 1024)
//   ci[a] = clamp((int)floorf((p[a]-lo[a])*inv), 0, dims[a]-1) ; key = (ci.z*dims.y + ci.y)*dims.x + ci.x
//   order = stable sort by key of the path-order indices ; cell_start[k] = #photons with key < k
namespace {

__device__ __forceinline__ uint32_t g_f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void grid_box_init_kernel(uint32_t *box) {
    if (threadIdx.x < 6) box[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;
}

__global__ void __launch_bounds__(256)
grid_bounds_kernel(const b2rt_photon *__restrict__ ph, int64_t n, uint32_t *__restrict__ box) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float4 a = __ldg(reinterpret_cast<const float4 *>(ph + i));
        lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
        hi[0] = fmaxf(hi[0], a.x); hi[1] = fmaxf(hi[1], a.y); hi[2] = fmaxf(hi[2], a.z);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int a = 0; a < 3; ++a)
            if (lo[a] <= hi[a]) { atomicMin(&box[a], g_f2ord(lo[a])); atomicMax(&box[3 + a], g_f2ord(hi[a])); }
}

struct GridParams { float lo[3]; float inv; int dims[3]; };

__device__ __forceinline__ uint32_t grid_key(const GridParams &g, float x, float y, float z) {
    int cx = min(max((int)floorf((x - g.lo[0]) * g.inv), 0), g.dims[0] - 1);
    int cy = min(max((int)floorf((y - g.lo[1]) * g.inv), 0), g.dims[1] - 1);
    int cz = min(max((int)floorf((z - g.lo[2]) * g.inv), 0), g.dims[2] - 1);
    return (uint32_t)((cz * g.dims[1] + cy) * g.dims[0] + cx);
}

__global__ void __launch_bounds__(256)
grid_keys_kernel(const b2rt_photon *__restrict__ ph, int64_t n, GridParams g, uint32_t *__restrict__ keys,
                 uint32_t *__restrict__ vals, uint32_t *__restrict__ cell_count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = __ldg(reinterpret_cast<const float4 *>(ph + i));
    uint32_t k = grid_key(g, a.x, a.y, a.z);
    keys[i] = k; vals[i] = (uint32_t)i;
    atomicAdd(&cell_count[k], 1u);
}

__global__ void __launch_bounds__(256)
grid_reorder_kernel(const b2rt_photon *__restrict__ ph, const uint32_t *__restrict__ order, int64_t n,
                    b2rt_photon *__restrict__ sorted, float4 *__restrict__ pos4) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *src = reinterpret_cast<const float4 *>(ph + order[i]);
    float4 a = __ldg(src), b = __ldg(src + 1);
    float4 *dst = reinterpret_cast<float4 *>(sorted + i);
    dst[0] = a; dst[1] = b;
    pos4[i] = make_float4(a.x, a.y, a.z, __uint_as_float((uint32_t)i));
}

// photons per coarse cell (4^3 fine cells): 16 fine rows, each an x-run of <= 4 cells = one cell_start difference
__global__ void __launch_bounds__(256)
grid_coarse_kernel(const uint32_t *__restrict__ cell_start, int dx, int dy, int dz, int cx, int cy, int cz,
                   uint32_t *__restrict__ coarse) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cx * cy * cz) return;
    const int X = i % cx, Y = (i / cx) % cy, Z = i / (cx * cy);
    const int xa = X * 4, xb = min(xa + 4, dx);
    uint32_t cnt = 0;
    for (int zs = 0; zs < 4; ++zs)
        for (int ys = 0; ys < 4; ++ys) {
            const int z = Z * 4 + zs, y = Y * 4 + ys;
            if (z < dz && y < dy) {
                const int64_t rb = ((int64_t)z * dy + y) * dx;
                cnt += cell_start[rb + xb] - cell_start[rb + xa];
            }
        }
    coarse[i] = cnt;
}

}  // namespace

int b2rt_photon_grid_build_impl(b2rt_ctx *ctx, int map, float cell_size, float target_per_cell) {
    PhotonMap &m = ctx->maps[map];
    cudaStream_t st = ctx->stream;
    const int64_t n = m.n;
    m.grid_built = false;
    m.grid = b2rt_grid_info{};
    m.grid.n_photons = n;
    B2RT_CUDA(ctx, cudaEventRecord(ctx->evb0, st));
    const b2rt_photon *ph = m.photons.as<b2rt_photon>();
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (n > 0) {
        B2RT_CUDA(ctx, ctx->d_grid_box.reserve(32));
        uint32_t *box = ctx->d_grid_box.as<uint32_t>();
        grid_box_init_kernel<<<1, 32, 0, st>>>(box);
        B2RT_LAUNCH_CHECK(ctx);
        int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
        grid_bounds_kernel<<<blocks, 256, 0, st>>>(ph, n, box);
        B2RT_LAUNCH_CHECK(ctx);
        uint32_t hb[6];
        B2RT_CUDA(ctx, cudaMemcpyAsync(hb, box, sizeof(hb), cudaMemcpyDeviceToHost, st));
        B2RT_CUDA(ctx, cudaStreamSynchronize(st));
        for (int a = 0; a < 6; ++a) {
            uint32_t o = hb[a], b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
            memcpy(a < 3 ? &lo[a] : &hi[a - 3], &b, 4);
        }
    }
    float ext[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
    float c = cell_size;
    if (!(c > 0.f)) {
        float target = target_per_cell > 0.f ? target_per_cell : 8.f;
        float A = 2.f * (ext[0] * ext[1] + ext[1] * ext[2] + ext[2] * ext[0]);
        float mx = fmaxf(ext[0], fmaxf(ext[1], ext[2]));
        c = (A > 0.f && n > 0) ? sqrtf(A * target / (float)n) : (mx > 0.f ? mx / 64.f : 1.f);
    }
    int dims[3];
    float inv;
    for (;;) {
        inv = 1.f / c;
        int64_t prod = 1;
        for (int a = 0; a < 3; ++a) {
            int d = (int)floorf(ext[a] * inv) + 1;
            dims[a] = d < 1 ? 1 : (d > 1024 ? 1024 : d);
            prod *= dims[a];
        }
        if (prod <= (1ll << 26) && (int)floorf(fmaxf(ext[0], fmaxf(ext[1], ext[2])) * inv) + 1 <= 1024) break;
        c = c * 2.f;
    }
    const int n_cells = dims[0] * dims[1] * dims[2];
    m.grid.cell_size = c; m.grid.inv_cell = inv; m.grid.n_cells = n_cells;
    for (int a = 0; a < 3; ++a) { m.grid.origin[a] = lo[a]; m.grid.dims[a] = dims[a]; }

    B2RT_CUDA(ctx, m.cell_start.reserve((size_t)(n_cells + 1) * 4));
    B2RT_CUDA(ctx, cudaMemsetAsync(m.cell_start.p, 0, (size_t)(n_cells + 1) * 4, st));
    const int64_t nn = n > 0 ? n : 1;
    B2RT_CUDA(ctx, m.keys.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, m.order.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, m.sorted.reserve((size_t)nn * sizeof(b2rt_photon)));
    B2RT_CUDA(ctx, m.pos4.reserve((size_t)nn * sizeof(float4)));
    for (int a = 0; a < 3; ++a) m.cdims[a] = (dims[a] + 3) / 4;
    const int n_coarse = m.cdims[0] * m.cdims[1] * m.cdims[2];
    B2RT_CUDA(ctx, m.coarse.reserve((size_t)n_coarse * 4));
    if (n > 0) {
        B2RT_CUDA(ctx, ctx->d_keys_alt.reserve((size_t)n * 4));
        B2RT_CUDA(ctx, ctx->d_vals_alt.reserve((size_t)n * 4));
        GridParams g;
        for (int a = 0; a < 3; ++a) { g.lo[a] = lo[a]; g.dims[a] = dims[a]; }
        g.inv = inv;
        unsigned grid = (unsigned)((n + 255) / 256);
        grid_keys_kernel<<<grid, 256, 0, st>>>(ph, n, g, m.keys.as<uint32_t>(), m.order.as<uint32_t>(),
                                               m.cell_start.as<uint32_t>());
        B2RT_LAUNCH_CHECK(ctx);
        int bits = 1;
        while ((1ll << bits) < n_cells) ++bits;
        bool in_alt = false;
        int rc = b2rt_sort_pairs_impl(ctx, n, m.keys.as<uint32_t>(), m.order.as<uint32_t>(),
                                      ctx->d_keys_alt.as<uint32_t>(), ctx->d_vals_alt.as<uint32_t>(), bits, &in_alt);
        if (rc) return rc;
        if (in_alt) { std::swap(m.keys, ctx->d_keys_alt); std::swap(m.order, ctx->d_vals_alt); }
        rc = b2rt_exclusive_scan_u32(ctx, m.cell_start.as<uint32_t>(), n_cells + 1, nullptr);
        if (rc) return rc;
        grid_reorder_kernel<<<grid, 256, 0, st>>>(ph, m.order.as<uint32_t>(), n, m.sorted.as<b2rt_photon>(),
                                                  m.pos4.as<float4>());
        B2RT_LAUNCH_CHECK(ctx);
        grid_coarse_kernel<<<(n_coarse + 255) / 256, 256, 0, st>>>(m.cell_start.as<uint32_t>(), dims[0], dims[1], dims[2],
                                                                   m.cdims[0], m.cdims[1], m.cdims[2],
                                                                   m.coarse.as<uint32_t>());
        B2RT_LAUNCH_CHECK(ctx);
    } else {
        B2RT_CUDA(ctx, cudaMemsetAsync(m.coarse.p, 0, (size_t)n_coarse * 4, st));
    }
    B2RT_CUDA(ctx, cudaEventRecord(ctx->evb1, st));
    B2RT_CUDA(ctx, cudaEventSynchronize(ctx->evb1));
    B2RT_CUDA(ctx, cudaEventElapsedTime(&m.grid.build_ms, ctx->evb0, ctx->evb1));
    m.grid_built = true;
    ++m.grid_epoch;
    return B2RT_OK;
}
