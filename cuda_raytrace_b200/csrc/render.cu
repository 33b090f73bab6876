// K14 camera rays, K6/K7 shading with direct lighting and the specular tree, K13 film: the camera pass.
//
// Replaces raytracing_camera / raytracing_closest_hit / directLight / shadow_any_hit
// (cuda_render/photon_mapping/raytracing.cu:19-147), simple_camera / simple_cloest_hit
// (simple_render/simplerender.cu:22-78), finalGatheringCamera (photon_mapping/gathering.cu:129-146), the CPU camera
// ray loop PbrtCamera::preLaunch (util/camera/pbrtcamera.cpp:57-122) and the CPU film loop
// (photonmappingrenderer.cpp:247-273).  Radiance follows pbrt's PhotonIntegrator::Li with finalgather off
// (pbrt-v2/src/integrators/photonmap.cpp:624-781): Le + UniformSampleAllLights with MIS
// (core/integrator.cpp:47-174) + LPhoton(caustic) + LPhoton(indirect) + SpecularReflect + SpecularTransmit
// (core/integrator.cpp:177-258, both branches followed, depth < maxspeculardepth).
//
// One thread per camera sample walks the specular tree depth-first with an explicit stack; shadow rays and MIS rays
// are traced inline.  Every diffuse leaf leaves a gather query (throughput folded into its kd) in the sample's slot
// row; a scan over the per-sample leaf counts compacts them, K12 runs once per photon map over the compact array,
// and the film kernel adds everything up per pixel in a fixed order.  Pixel jitter and light samples are
// Philox(seed, pixel, sample), so an image tile is bit-identical no matter which GPU renders it.
#include "b2rt_shade.cuh"
#include <algorithm>

namespace {

constexpr int RENDER_THREADS = 128;
// Resident blocks per SM the camera-pass kernel is compiled for.  Left to itself ptxas gives it 128 registers (4 blocks) and
// still spills 258 bytes; capped at 80 registers (6 blocks) it spills more (354 B) and is faster all the same -- the kernel
// lives on latency hiding: Cfg4 frame (8 M triangles, 4K, 16 spp) 37.3 -> 33.2 ms, Cfg3 camera pass 16.9 -> 16.7 ms; 3 / 5 / 7 /
// 8 blocks: 40.0 / 34.3 / 33.6 / 33.8 ms (profiles/r07q_render_blocks_ab.txt).
#ifndef B2RT_RENDER_BLOCKS
#define B2RT_RENDER_BLOCKS 6
#endif
constexpr int MAX_TREE_DEPTH = 8;

// ---- K14 image-plane sample patterns: where inside its pixel sample k lies ------------------------------------------
struct SamplerD { int type, nx, ny; };
constexpr float ONE_MINUS_EPS = 0.9999999403953552f;          // core/montecarlo.h:48
// (0,2)-sequence, pbrt Sample02 = VanDerCorput + Sobol2 with one random scramble each (core/montecarlo.h:272-295)
__host__ __device__ __forceinline__ float ld_vdc(uint32_t n, uint32_t scramble) {
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
    n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
    n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
    n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
    n ^= scramble;
    return fminf((float)((n >> 8) & 0xffffffu) / (float)(1 << 24), ONE_MINUS_EPS);
}
__host__ __device__ __forceinline__ float ld_sobol2(uint32_t n, uint32_t scramble) {
    for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
        if (n & 1u) scramble ^= v;
    return fminf((float)((scramble >> 8) & 0xffffffu) / (float)(1 << 24), ONE_MINUS_EPS);
}
// offset of sample k inside pixel `gpix`.  random: two Philox uniforms (pbrt "random").  stratified: stratum
// (k % nx, k / nx) with a Philox jitter, StratifiedSample2D (core/montecarlo.cpp:158-168; pbrt "stratified",
// samplers/stratified.cpp:79-82).  lowdiscrepancy: point k of a (0,2)-sequence whose two scrambles are the pixel's own
// Philox words - LDShuffleScrambled2D without its shuffle, which only permutes the order inside the pixel
// (core/montecarlo.h:317-325, pbrt "lowdiscrepancy", samplers/lowdiscrepancy.cpp:67-79).
__device__ __forceinline__ float2 pixel_offset(const SamplerD &sm, uint32_t gpix, uint32_t k, uint32_t k0, uint32_t k1) {
    if (sm.type == B2RT_SAMPLER_LOWDISCREPANCY) {
        const u4 sc = philox4x32_10(gpix, 0xffffffffu, 1u, 0u, k0, k1);
        return make_float2(ld_vdc(k, sc.x), ld_sobol2(k, sc.y));
    }
    const u4 rj = philox4x32_10(gpix, k, 0u, 0u, k0, k1);
    if (sm.type == B2RT_SAMPLER_STRATIFIED) {
        const float dx = 1.f / (float)sm.nx, dy = 1.f / (float)sm.ny;
        const int sx = (int)(k % (uint32_t)sm.nx), sy = (int)(k / (uint32_t)sm.nx);
        return make_float2(fminf(((float)sx + u01(rj.x)) * dx, ONE_MINUS_EPS), fminf(((float)sy + u01(rj.y)) * dy, ONE_MINUS_EPS));
    }
    return make_float2(u01(rj.x), u01(rj.y));
}
__device__ __forceinline__ uint32_t global_pixel_id(int px, int py, int xres, int yres) {
    uint32_t gpix = (uint32_t)py * (uint32_t)xres + (uint32_t)px;             // tiling-independent
    // sample cells around the image (a pixel filter wider than the pixel reaches them, ImageFilm::GetSampleExtent):
    // ids after the image's own, unique for cells up to 64 pixels outside
    if (px < 0 || py < 0 || px >= xres || py >= yres)
        gpix = (uint32_t)xres * (uint32_t)yres + (uint32_t)(py + 64) * (uint32_t)(xres + 128) + (uint32_t)(px + 64);
    return gpix;
}

__device__ __forceinline__ float power_heuristic(float f, float g) { return (f * f) / (f * f + g * g); }

// Shape::Pdf(p, wi) for the disk light (core/shape.cpp:95-108); also returns the hit distance
__device__ __forceinline__ float disk_light_pdf(const AnalyticPrim &D, const LightD &L, f3 p, f3 wi) {
    RayD r;
    r.ox = p.x; r.oy = p.y; r.oz = p.z; r.dx = wi.x; r.dy = wi.y; r.dz = wi.z; r.tmin = 1e-3f;
    float t;
    if (!disk_test(r, D, INFINITY, t)) return 0.f;
    f3 dv = wi * t;
    float d2 = dv.x * dv.x + dv.y * dv.y + dv.z * dv.z;
    float pdf = d2 / (fabsf(dot3(L.n, -wi)) * L.area);
    return isinf(pdf) ? 0.f : pdf;
}

// UniformSampleAllLights + EstimateDirect for a Lambertian surface (flags = BSDF_ALL & ~BSDF_SPECULAR)
template <bool INST>
__device__ f3 direct_lighting(const DScene &sc, const Surface &s, const Mat &m, f3 wo, uint32_t c0, uint32_t c1,
                              uint32_t node, uint32_t k0, uint32_t k1, bool &overflow, const SamplerD &sm) {
    f3 L = mk3(0.f, 0.f, 0.f);
    const f3 f_lambert = m.kd * B2_INV_PI;
    // With pbrt's stratified / low-discrepancy samplers the light and BSDF samples of the camera ray's own hit are not
    // independent either: every light gets its ns 2D samples per pixel sample out of ONE pattern over the pixel's
    // spp * ns samples (LDShuffleScrambled2D(ns, spp), samplers/lowdiscrepancy.cpp:67-79, core/sampler.cpp:96-106).
    // Here: points [k * ns, (k + 1) * ns) of a (0,2)-sequence whose scrambles are Philox words of (pixel, light).
    // Deeper nodes of the specular tree keep independent uniforms (pbrt re-uses the same sample values there).
    const bool pattern = sm.type != B2RT_SAMPLER_RANDOM && node == 1u;
    for (int li = 0; li < sc.n_lights; ++li) {
        LightD Lt = load_light(sc, li);
        const bool delta = Lt.type != B2RT_LIGHT_DISK_AREA;           // point, spot, distant: one sample, no MIS
        const int ns = delta ? 1 : max(Lt.nsamples, 1);
        f3 Ld = mk3(0.f, 0.f, 0.f);
        u4 scr = {0u, 0u, 0u, 0u};
        if (pattern && !delta) scr = philox4x32_10(c0, 0xfffffffeu, (uint32_t)li, 0u, k0, k1);
        for (int j = 0; j < ns; ++j) {
            u4 r = philox4x32_10(c0, c1, node, 0x100u + (uint32_t)(li * 64 + j), k0, k1);
            float u_l1 = u01(r.x), u_l2 = u01(r.y), u_b1 = u01(r.z), u_b2 = u01(r.w);
            if (pattern) {
                const uint32_t n = c1 * (uint32_t)ns + (uint32_t)j;
                u_l1 = ld_vdc(n, scr.x); u_l2 = ld_sobol2(n, scr.y); u_b1 = ld_vdc(n, scr.z); u_b2 = ld_sobol2(n, scr.w);
            }
            HitD h;
            if (Lt.type == B2RT_LIGHT_DISTANT) {
                // DistantLight::Sample_L (lights/distant.cpp:48-55): wi = lightDir, Li = L, visibility ray to infinity
                const f3 wi = Lt.n;
                if (!(dot3(wi, s.ng) * dot3(wo, s.ng) > 0.f)) continue;
                if (traverse<true, INST>(sc.nodes, sc.recs, sc.analytic, s.p.x, s.p.y, s.p.z, wi.x, wi.y, wi.z, s.ray_eps, INFINITY, h,
                                   overflow))
                    continue;
                const float w = fabsf(dot3(wi, s.n));
                Ld = Ld + mk3(f_lambert.x * Lt.rgb.x * w, f_lambert.y * Lt.rgb.y * w, f_lambert.z * Lt.rgb.z * w);
                continue;
            }
            if (delta) {
                // PointLight / SpotLight::Sample_L (lights/point.cpp:50-57, lights/spot.cpp:50-57): delta light, no MIS
                f3 dl = Lt.p - s.p;
                float d2 = dl.x * dl.x + dl.y * dl.y + dl.z * dl.z;
                f3 wi = normalize3(dl);
                bool same = dot3(wi, s.ng) * dot3(wo, s.ng) > 0.f;      // BSDF::f: sides by the geometric normal
                if (!same) continue;
                f3 I = Lt.rgb;
                if (Lt.type == B2RT_LIGHT_SPOT) {
                    const float fo = spot_falloff(sc.lights[li], -wi);
                    if (fo == 0.f) continue;
                    I = I * fo;
                }
                // visibility segment p -> lightPos: direction (lightPos - p), [rayEps, 1]
                if (traverse<true, INST>(sc.nodes, sc.recs, sc.analytic, s.p.x, s.p.y, s.p.z, dl.x, dl.y, dl.z, s.ray_eps, 1.f, h,
                                   overflow))
                    continue;
                float w = fabsf(dot3(wi, s.n));     // / lightPdf (= 1)
                Ld = Ld + mk3(f_lambert.x * (I.x / d2) * w, f_lambert.y * (I.y / d2) * w, f_lambert.z * (I.z / d2) * w);
                continue;
            }
            const AnalyticPrim &D = sc.analytic[Lt.analytic];
            // ---- light sampling (lights/diffuse.cpp:69-81) ----
            {
                f3 ps = disk_sample_point(D, u_l1, u_l2);
                f3 dl = ps - s.p;
                f3 wi = normalize3(dl);
                float lpdf = disk_light_pdf(D, Lt, s.p, wi);
                bool emits = dot3(Lt.n, -wi) > 0.f;
                bool same = dot3(wi, s.ng) * dot3(wo, s.ng) > 0.f;      // BSDF::f: sides by the geometric normal
                if (lpdf > 0.f && emits && same) {
                    if (!traverse<true, INST>(sc.nodes, sc.recs, sc.analytic, s.p.x, s.p.y, s.p.z, dl.x, dl.y, dl.z, s.ray_eps,
                                        1.f - 1e-3f, h, overflow)) {
                        f3 wil = to_local(s, wi), wol = to_local(s, wo);
                        float bpdf = (wol.z * wil.z > 0.f) ? fabsf(wil.z) * B2_INV_PI : 0.f;
                        float w = fabsf(dot3(wi, s.n)) * power_heuristic(lpdf, bpdf) / lpdf;
                        Ld = Ld + mk3(f_lambert.x * Lt.rgb.x * w, f_lambert.y * Lt.rgb.y * w, f_lambert.z * Lt.rgb.z * w);
                    }
                }
            }
            // ---- BSDF sampling ----
            {
                BsdfSample bs;
                if (!sample_bsdf(s, m, wo, u_b1, u_b2, 0.f, bs) || is_black(bs.f) || !(bs.pdf > 0.f)) continue;
                float lpdf = disk_light_pdf(D, Lt, s.p, bs.wi);
                if (lpdf == 0.f) continue;
                float weight = power_heuristic(bs.pdf, lpdf);
                if (traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, s.p.x, s.p.y, s.p.z, bs.wi.x, bs.wi.y, bs.wi.z, s.ray_eps,
                                    INFINITY, h, overflow) &&
                    h.prim == Lt.prim && dot3(Lt.n, -bs.wi) > 0.f) {
                    float w = bs.absdot * weight / bs.pdf;
                    Ld = Ld + mk3(bs.f.x * Lt.rgb.x * w, bs.f.y * Lt.rgb.y * w, bs.f.z * Lt.rgb.z * w);
                }
            }
        }
        float inv = 1.f / (float)ns;
        L = L + Ld * inv;
    }
    return L;
}

struct TreeNode { f3 o, d; float tmin; f3 T; int depth; uint32_t id; };


// K14 + K6 + K7.  sample s (chunk-local) -> pixel (crop-window row-major) and sample number k.
template <bool INST>
__global__ void __launch_bounds__(RENDER_THREADS, B2RT_RENDER_BLOCKS)
render_li_kernel(DScene sc, b2rt_camera cam, b2rt_frame_params fp, uint32_t seed_lo, uint32_t seed_hi, int64_t first_sample,
                 int64_t n_samples, int slots, float *__restrict__ sample_rgb, b2rt_query *__restrict__ qslots,
                 uint32_t *__restrict__ qcnt, uint32_t *status, float2 *__restrict__ sample_xy, SamplerD sm,
                 const uint32_t *__restrict__ pixmap) {
    int64_t ls = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ls >= n_samples) return;
    const int64_t s = first_sample + ls;
    const int cw = fp.x1 - fp.x0;
    const int64_t pix = s / fp.spp;
    const int k = (int)(s % fp.spp);
    int px = fp.x0 + (int)(pix % cw), py = fp.y0 + (int)(pix / cw);
    if (pixmap) {               // this GPU's pixels of a frame shared by several (image tiles): x | y << 16
        const uint32_t pm = __ldg(pixmap + pix);
        px = (int)(pm & 0xffffu); py = (int)(pm >> 16);
    }
    const uint32_t gpix = global_pixel_id(px, py, cam.xres, cam.yres);
    const uint32_t k0 = seed_lo ^ 0x43414d52u, k1 = seed_hi;                     // 'CAMR' stream
    const float2 off = pixel_offset(sm, gpix, (uint32_t)k, k0, k1);
    const float fx = (float)px + off.x, fy = (float)py + off.y;
    if (sample_xy) sample_xy[ls] = make_float2(fx, fy);
    f3 d0 = mk3(cam.A[0] + fx * cam.B[0] + fy * cam.C[0], cam.A[1] + fx * cam.B[1] + fy * cam.C[1],
                cam.A[2] + fx * cam.B[2] + fy * cam.C[2]);

    TreeNode stack[MAX_TREE_DEPTH + 1];
    int sp = 0;
    stack[sp].o = mk3(cam.o[0], cam.o[1], cam.o[2]); stack[sp].d = normalize3(d0); stack[sp].tmin = 0.f;
    stack[sp].T = mk3(1.f, 1.f, 1.f); stack[sp].depth = 0; stack[sp].id = 1u; ++sp;
    f3 L = mk3(0.f, 0.f, 0.f);
    uint32_t nq = 0;
    bool overflow = false, dropped = false;
    b2rt_query *myq = qslots + ls * slots;
    while (sp > 0) {
        TreeNode nd = stack[--sp];
        HitD h;
        if (!traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, nd.o.x, nd.o.y, nd.o.z, nd.d.x, nd.d.y, nd.d.z, nd.tmin, INFINITY, h,
                             overflow))
            continue;
        Surface sf = make_surface<INST>(sc, h, nd.o, nd.d);
        Mat m = load_material(sc, sf.material);
        f3 wo = -nd.d;
        if (sf.light >= 0 && sf.light < sc.n_lights && dot3(sf.ng, wo) > 0.f) {     // isect.Le(wo): DiffuseAreaLight::L
            const Light &Lg = sc.lights[sf.light];
            L = L + nd.T * mk3(Lg.rgb[0], Lg.rgb[1], Lg.rgb[2]);
        }
        if (m.type == B2RT_MAT_MATTE) {
            if (fp.direct_lighting)
                L = L + nd.T * direct_lighting<INST>(sc, sf, m, wo, gpix, (uint32_t)k, nd.id, k0, k1, overflow, sm);
            if (fp.use_caustic || fp.use_indirect) {
                if ((int)nq < slots) {
                    f3 nf = dot3(sf.n, wo) < 0.f ? -sf.n : sf.n;              // Faceforward(n, wo)
                    f3 kd = m.kd * nd.T;
                    float4 *dst = reinterpret_cast<float4 *>(myq + nq);
                    dst[0] = make_float4(sf.p.x, sf.p.y, sf.p.z, nf.x);
                    dst[1] = make_float4(nf.y, nf.z, kd.x, kd.y);
                    // pad: what the final-gather stage needs to continue from this leaf (ray epsilon, and the Philox
                    // counter words: pixel id, sample number | specular-tree node); K12 never reads it
                    dst[2] = make_float4(kd.z, sf.ray_eps, __uint_as_float(gpix),
                                         __uint_as_float((uint32_t)k | (nd.id << 22)));
                    ++nq;
                } else dropped = true;
            }
            continue;
        }
        if (nd.depth + 1 >= fp.max_specular_depth) continue;
        // SpecularReflect / SpecularTransmit: single-lobe sampling, pdf 1, weight f*|cos|/pdf = F*Kr or (1-F)*Kt
        f3 wol = to_local(sf, wo);
        float F = (m.type == B2RT_MAT_GLASS) ? fresnel_dielectric(wol.z, 1.f, m.ior) : 1.f;
        if (m.type == B2RT_MAT_GLASS && !is_black(m.kt)) {
            f3 wil;
            if (refract_local(wol, m.ior, wil) && wil.z != 0.f && sp < MAX_TREE_DEPTH) {
                float ac = fabsf(wil.z);
                f3 w = mk3((1.f - F) * m.kt.x / ac * ac, (1.f - F) * m.kt.y / ac * ac, (1.f - F) * m.kt.z / ac * ac);
                if (!is_black(w)) {
                    stack[sp].o = sf.p; stack[sp].d = to_world(sf, wil); stack[sp].tmin = sf.ray_eps;
                    stack[sp].T = nd.T * w; stack[sp].depth = nd.depth + 1; stack[sp].id = nd.id * 2u + 1u; ++sp;
                }
            }
        }
        if (!is_black(m.kd) && wol.z != 0.f && sp < MAX_TREE_DEPTH) {
            f3 wil = mk3(-wol.x, -wol.y, wol.z);
            float ac = fabsf(wil.z);
            f3 w = mk3(F * m.kd.x / ac * ac, F * m.kd.y / ac * ac, F * m.kd.z / ac * ac);
            if (!is_black(w)) {
                stack[sp].o = sf.p; stack[sp].d = to_world(sf, wil); stack[sp].tmin = sf.ray_eps;
                stack[sp].T = nd.T * w; stack[sp].depth = nd.depth + 1; stack[sp].id = nd.id * 2u; ++sp;
            }
        }
    }
    sample_rgb[3 * ls + 0] = L.x; sample_rgb[3 * ls + 1] = L.y; sample_rgb[3 * ls + 2] = L.z;
    qcnt[ls] = nq;
    if (overflow) atomicAdd(status, 1u);
    if (dropped) atomicAdd(status + 1, 1u);
}

}  // namespace

namespace {

__global__ void __launch_bounds__(256)
query_compact_kernel(const b2rt_query *__restrict__ qslots, const uint32_t *__restrict__ qoff, const uint32_t *__restrict__ qtot,
                     int64_t n_samples, int slots, b2rt_query *__restrict__ out) {
    int64_t ls = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ls >= n_samples) return;
    uint32_t o = qoff[ls];
    uint32_t n = ((ls + 1 < n_samples) ? qoff[ls + 1] : qtot[0]) - o;
    const float4 *src = reinterpret_cast<const float4 *>(qslots + ls * slots);
    float4 *dst = reinterpret_cast<float4 *>(out + o);
    for (uint32_t j = 0; j < 3 * n; ++j) dst[j] = src[j];
}

// K13 (box filter, radius 0.5: a sample only reaches its own pixel, film/image.cpp:77-137 with filters/box.cpp):
// pixel = mean over its spp samples of (Le + direct + sum over the sample's leaves of caustic + indirect estimates)
__global__ void __launch_bounds__(256)
film_resolve_kernel(const float *__restrict__ sample_rgb, const uint32_t *__restrict__ qoff, const uint32_t *__restrict__ qtot,
                    const float *__restrict__ rgb_c, const float *__restrict__ rgb_i, int64_t n_samples, int spp,
                    int64_t first_pixel, int64_t n_pixels, float *__restrict__ out_rgb, const uint32_t *__restrict__ pixmap,
                    int xres) {
    int64_t lp = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= n_pixels) return;
    f3 acc = mk3(0.f, 0.f, 0.f);
    for (int k = 0; k < spp; ++k) {
        int64_t ls = lp * spp + k;
        f3 L = mk3(sample_rgb[3 * ls], sample_rgb[3 * ls + 1], sample_rgb[3 * ls + 2]);
        uint32_t o = qoff[ls];
        uint32_t n = ((ls + 1 < n_samples) ? qoff[ls + 1] : qtot[0]) - o;
        for (uint32_t j = o; j < o + n; ++j) {
            if (rgb_c) L = L + mk3(rgb_c[3 * j], rgb_c[3 * j + 1], rgb_c[3 * j + 2]);
            if (rgb_i) L = L + mk3(rgb_i[3 * j], rgb_i[3 * j + 1], rgb_i[3 * j + 2]);
        }
        acc = acc + L;
    }
    float inv = 1.f / (float)spp;
    int64_t at = first_pixel + lp;
    if (pixmap) {               // tile pixel: straight to its place in the WHOLE image, which may live on another GPU
        const uint32_t pm = __ldg(pixmap + at);      // (the composite on GPU 0 is this store, over NVLink)
        at = (int64_t)(pm >> 16) * xres + (int64_t)(pm & 0xffffu);
    }
    float *o = out_rgb + 3 * at;
    o[0] = acc.x * inv; o[1] = acc.y * inv; o[2] = acc.z * inv;
}

// per-sample radiance (camera-pass terms + the photon-map estimates of the sample's leaves) for the filtered film
__global__ void __launch_bounds__(256)
sample_resolve_kernel(const float *__restrict__ sample_rgb, const uint32_t *__restrict__ qoff, const uint32_t *__restrict__ qtot,
                      const float *__restrict__ rgb_c, const float *__restrict__ rgb_i, int64_t n_samples,
                      float *__restrict__ out_rgb) {
    int64_t ls = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ls >= n_samples) return;
    f3 L = mk3(sample_rgb[3 * ls], sample_rgb[3 * ls + 1], sample_rgb[3 * ls + 2]);
    uint32_t o = qoff[ls];
    uint32_t n = ((ls + 1 < n_samples) ? qoff[ls + 1] : qtot[0]) - o;
    for (uint32_t j = o; j < o + n; ++j) {
        if (rgb_c) L = L + mk3(rgb_c[3 * j], rgb_c[3 * j + 1], rgb_c[3 * j + 2]);
        if (rgb_i) L = L + mk3(rgb_i[3 * j], rgb_i[3 * j + 1], rgb_i[3 * j + 2]);
    }
    out_rgb[3 * ls] = L.x; out_rgb[3 * ls + 1] = L.y; out_rgb[3 * ls + 2] = L.z;
}

// parity hook: the raster positions of the camera samples of a crop window, sample-major as the camera pass walks them
__global__ void __launch_bounds__(256)
sample_positions_kernel(int xres, int yres, int x0, int y0, int cw, int spp, uint32_t seed_lo, uint32_t seed_hi, int64_t n,
                        SamplerD sm, float2 *__restrict__ out) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int64_t pix = s / spp;
    const int k = (int)(s % spp);
    const int px = x0 + (int)(pix % cw), py = y0 + (int)(pix / cw);
    const float2 off = pixel_offset(sm, global_pixel_id(px, py, xres, yres), (uint32_t)k, seed_lo ^ 0x43414d52u, seed_hi);
    out[s] = make_float2((float)px + off.x, (float)py + off.y);
}

}  // namespace

// parity hook for K6: closest hit + the surface the shade stage derives from it, one thread per ray
template <bool INST>
__global__ void __launch_bounds__(128)
surface_at_kernel(DScene sc, const b2rt_ray *__restrict__ rays, int64_t n, b2rt_surface *__restrict__ out, uint32_t *status) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
    const float4 a = __ldg(rp), b = __ldg(rp + 1);
    HitD h;
    bool overflow = false;
    float4 *o = reinterpret_cast<float4 *>(out + i);
    if (!traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w, h, overflow)) {
        for (int k = 0; k < 4; ++k) o[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        o[4] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
    } else {
        const Surface s = make_surface<INST>(sc, h, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z));
        o[0] = make_float4(s.p.x, s.p.y, s.p.z, h.t);
        o[1] = make_float4(s.ng.x, s.ng.y, s.ng.z, s.u);
        o[2] = make_float4(s.n.x, s.n.y, s.n.z, s.v);
        o[3] = make_float4(s.sn.x, s.sn.y, s.sn.z, s.ray_eps);
        o[4] = make_float4(s.tn.x, s.tn.y, s.tn.z, __int_as_float(h.prim));
    }
    if (overflow) atomicAdd(status, 1u);
}
int b2rt_surface_at_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_surface *d_out) {
    if (n <= 0) return B2RT_OK;
    DScene sc;
    b2rt_fill_dscene(ctx, &sc);
    (sc.instances ? surface_at_kernel<true> : surface_at_kernel<false>)<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(sc, d_rays, n, d_out, ctx->d_counters.as<uint32_t>() + 8);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

int b2rt_sample_positions_impl(b2rt_ctx *ctx, int xres, int yres, const int32_t window[4], int spp, float *d_xy) {
    SamplerD sm = {ctx->sampler_type, ctx->sampler_nx, ctx->sampler_ny};
    if (sm.type == B2RT_SAMPLER_STRATIFIED && (int64_t)sm.nx * sm.ny != spp) {
        ctx->err = "sample_positions: stratified sampler needs xsamples * ysamples == spp"; return B2RT_E_ARG;
    }
    const int cw = window[2] - window[0];
    const int64_t n = (int64_t)cw * (window[3] - window[1]) * spp;
    sample_positions_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
        xres, yres, window[0], window[1], cw, spp, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32), n, sm,
        reinterpret_cast<float2 *>(d_xy));
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

// parity hook for K6 / K7 / K14: what render_li_kernel leaves per camera sample, before any photon-map term is added.
// d_rgb[3 n], d_nq[n], d_q[n * slots] (NULL: no queries wanted), n = crop-window pixels * spp, sample-major per pixel.
int b2rt_camera_samples_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, int slots, float *d_rgb,
                             uint32_t *d_nq, b2rt_query *d_q) {
    DScene sc;
    b2rt_fill_dscene(ctx, &sc);
    b2rt_frame_params p = *fp;
    p.use_caustic = 0; p.use_indirect = d_q ? 1 : 0;
    SamplerD sm = {ctx->sampler_type, ctx->sampler_nx, ctx->sampler_ny};
    if (sm.type == B2RT_SAMPLER_STRATIFIED && (int64_t)sm.nx * sm.ny != fp->spp) {
        ctx->err = "camera_samples: stratified sampler needs xsamples * ysamples == spp"; return B2RT_E_ARG;
    }
    const int64_t n = (int64_t)(p.x1 - p.x0) * (p.y1 - p.y0) * fp->spp;
    const int64_t chunk = 1 << 20;
    for (int64_t s0 = 0; s0 < n; s0 += chunk) {
        const int64_t ns = std::min(chunk, n - s0);
        (sc.instances ? render_li_kernel<true> : render_li_kernel<false>)<<<(unsigned)((ns + RENDER_THREADS - 1) / RENDER_THREADS), RENDER_THREADS, 0, ctx->stream>>>(
            sc, *cam, p, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32), s0, ns, slots, d_rgb + 3 * s0,
            d_q ? d_q + s0 * slots : nullptr, d_nq + s0, ctx->d_counters.as<uint32_t>() + 8, nullptr, sm, nullptr);
        B2RT_LAUNCH_CHECK(ctx);
    }
    return B2RT_OK;
}
int b2rt_camera_slots(b2rt_ctx *ctx, int max_specular_depth) {
    for (const Material &m : ctx->h_materials)
        if (m.type != B2RT_MAT_MATTE) { int d = max_specular_depth < 1 ? 1 : max_specular_depth; return 1 << (d - 1 > 5 ? 5 : d - 1); }
    return 1;
}

// d_out: device buffer of (x1-x0)*(y1-y0)*3 floats, crop-window row-major
int b2rt_render_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, float *d_out) {
    return b2rt_render_pixels_impl(ctx, cam, fp, nullptr, 0, d_out);
}
// The same for an arbitrary pixel list (d_pixmap[n_map], x | y << 16, on this device): the multi-GPU frame's image tiles.
// d_out is then the WHOLE xres * yres * 3 image (possibly a peer pointer) and every pixel is stored at its own place.
// Box filter of radius <= 0.5 only (a wider filter needs the halo cells of every tile; multi.cu uses row bands then).
int b2rt_render_pixels_impl(b2rt_ctx *ctx, const b2rt_camera *cam, const b2rt_frame_params *fp, const uint32_t *d_pixmap,
                            int64_t n_map, float *d_out) {
    cudaStream_t st = ctx->stream;
    DScene sc;
    b2rt_fill_dscene(ctx, &sc);
    const bool any_spec = [&] {
        for (const Material &m : ctx->h_materials) if (m.type != B2RT_MAT_MATTE) return true;
        return false;
    }();
    int slots = 1;
    if (any_spec) { int d = fp->max_specular_depth < 1 ? 1 : fp->max_specular_depth; slots = 1 << (d - 1 > 5 ? 5 : d - 1); }
    const bool use_c = fp->use_caustic && ctx->maps[0].grid_built && ctx->maps[0].n > 0;
    // final gathering with no radiance photon at all: nothing to look up, the indirect term is zero
    const bool use_i = fp->use_indirect && ctx->maps[1].grid_built && ctx->maps[1].n > 0 &&
                       !(ctx->fg_enable && ctx->radiance_computed && !ctx->maps[B2RT_MAP_RADIANCE].grid_built);
    const bool use_fg = use_i && ctx->fg_enable;
    if (use_fg && !ctx->radiance_computed) {
        ctx->err = "render_frame: final gathering is on but b2rt_radiance_compute has not been called"; return B2RT_E_STATE;
    }
    b2rt_frame_params p = *fp;
    p.use_caustic = use_c; p.use_indirect = use_i;
    // pixel filter wider than the pixel: trace the sample cells around the crop window that reach into it, keep every
    // sample's raster position and radiance, reconstruct at the end (film.cu).  The box filter of radius 0.5 (pbrt's
    // default) keeps the direct per-pixel mean below.
    const b2rt_filter &flt = ctx->film_filter;
    const bool filtered = !(flt.type == B2RT_FILTER_BOX && flt.xwidth <= 0.5f && flt.ywidth <= 0.5f);
    if (filtered && d_pixmap) { ctx->err = "render_pixels: pixel lists need the box filter"; return B2RT_E_STATE; }
    if (filtered) {
        int hx, hy;
        b2rt_filter_halo(&flt, &hx, &hy);
        if (hx > 64 || hy > 64) { ctx->err = "render_frame: filter too wide"; return B2RT_E_ARG; }
        p.x0 -= hx; p.x1 += hx; p.y0 -= hy; p.y1 += hy;
    }
    SamplerD sm = {ctx->sampler_type, ctx->sampler_nx, ctx->sampler_ny};
    if (sm.type == B2RT_SAMPLER_STRATIFIED && (int64_t)sm.nx * sm.ny != fp->spp) {
        ctx->err = "render_frame: stratified sampler needs xsamples * ysamples == spp"; return B2RT_E_ARG;
    }
    const int64_t n_cells = d_pixmap ? n_map : (int64_t)(p.x1 - p.x0) * (p.y1 - p.y0);
    if (n_cells <= 0) return B2RT_OK;
    if (filtered) {
        B2RT_CUDA(ctx, ctx->d_film_xy.reserve((size_t)n_cells * fp->spp * 8));
        B2RT_CUDA(ctx, ctx->d_film_rgb.reserve((size_t)n_cells * fp->spp * 12));
    }
    // chunk = whole pixels, at most ~2M samples
    int64_t chunk_pixels = (1ll << 21) / fp->spp;
    if (chunk_pixels < 1) chunk_pixels = 1;
    if (chunk_pixels > n_cells) chunk_pixels = n_cells;
    const int64_t cs = chunk_pixels * fp->spp;
    B2RT_CUDA(ctx, ctx->d_samples.reserve((size_t)cs * 12));
    B2RT_CUDA(ctx, ctx->d_qslots.reserve((size_t)cs * slots * sizeof(b2rt_query)));
    B2RT_CUDA(ctx, ctx->d_qcnt.reserve((size_t)cs * 4));
    B2RT_CUDA(ctx, ctx->d_queries.reserve((size_t)cs * slots * sizeof(b2rt_query)));
    B2RT_CUDA(ctx, ctx->d_q_rgb.reserve((size_t)cs * slots * 12));
    B2RT_CUDA(ctx, ctx->d_q_rgb2.reserve((size_t)cs * slots * 12));
    uint32_t *counters = ctx->d_counters.as<uint32_t>();
    for (int64_t p0 = 0; p0 < n_cells; p0 += chunk_pixels) {
        const int64_t np = (p0 + chunk_pixels <= n_cells) ? chunk_pixels : n_cells - p0;
        const int64_t ns = np * fp->spp;
        uint32_t *qcnt = ctx->d_qcnt.as<uint32_t>();
        (sc.instances ? render_li_kernel<true> : render_li_kernel<false>)<<<(unsigned)((ns + RENDER_THREADS - 1) / RENDER_THREADS), RENDER_THREADS, 0, st>>>(
            sc, *cam, p, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32), p0 * fp->spp, ns, slots,
            ctx->d_samples.as<float>(), ctx->d_qslots.as<b2rt_query>(), qcnt, counters + 8,
            filtered ? ctx->d_film_xy.as<float2>() + p0 * fp->spp : nullptr, sm, d_pixmap);
        B2RT_LAUNCH_CHECK(ctx);
        int rc = b2rt_exclusive_scan_u32(ctx, qcnt, ns, counters + 12);
        if (rc) return rc;
        const float *rgb_c = nullptr, *rgb_i = nullptr;
        if (use_c || use_i) {
            uint32_t tot = 0;
            B2RT_CUDA(ctx, cudaMemcpyAsync(&tot, counters + 12, 4, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaStreamSynchronize(st));
            query_compact_kernel<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(ctx->d_qslots.as<b2rt_query>(), qcnt,
                                                                               counters + 12, ns, slots,
                                                                               ctx->d_queries.as<b2rt_query>());
            B2RT_LAUNCH_CHECK(ctx);
            if (use_c) {
                rc = b2rt_gather_impl(ctx, 0, tot, ctx->d_queries.as<b2rt_query>(), fp->nused, fp->max_dist2,
                                      ctx->d_q_rgb.as<float>(), nullptr, nullptr, nullptr);
                if (rc) return rc;
                rgb_c = ctx->d_q_rgb.as<float>();
            }
            if (use_i) {
                if (use_fg)       // one-bounce final gathering instead of the direct estimate (photonmap.cpp:645-758)
                    rc = b2rt_final_gather_impl(ctx, tot, ctx->d_queries.as<b2rt_query>(), fp->max_dist2, ctx->d_q_rgb2.as<float>());
                else
                    rc = b2rt_gather_impl(ctx, 1, tot, ctx->d_queries.as<b2rt_query>(), fp->nused, fp->max_dist2,
                                          ctx->d_q_rgb2.as<float>(), nullptr, nullptr, nullptr);
                if (rc) return rc;
                rgb_i = ctx->d_q_rgb2.as<float>();
            }
        }
        if (filtered)
            sample_resolve_kernel<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(
                ctx->d_samples.as<float>(), qcnt, counters + 12, rgb_c, rgb_i, ns, ctx->d_film_rgb.as<float>() + 3 * p0 * fp->spp);
        else
            film_resolve_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(ctx->d_samples.as<float>(), qcnt, counters + 12,
                                                                              rgb_c, rgb_i, ns, fp->spp, p0, np, d_out, d_pixmap,
                                                                              cam->xres);
        B2RT_LAUNCH_CHECK(ctx);
    }
    if (filtered) {
        const int32_t cells[4] = {p.x0, p.y0, p.x1, p.y1}, pixels[4] = {fp->x0, fp->y0, fp->x1, fp->y1};
        return b2rt_film_impl(ctx, &flt, cells, fp->spp, ctx->d_film_xy.as<float>(), ctx->d_film_rgb.as<float>(), pixels, d_out);
    }
    return B2RT_OK;
}
