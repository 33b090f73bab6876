// Hit attributes (K6), BSDFs and light sampling shared by the photon pass (photon.cu) and the camera pass
// (render.cu).  Behaviour follows pbrt-v2 (the CPU oracle), NOT the reference's simplified device code where the
// two differ (SURVEY.md 8(a) "behavioural gaps"): Fresnel glass with both lobes, mirror honours Kr, Russian
// roulette, per-hit ray epsilon.
//   reference device code being replaced: cuda_render/util/material/cudamaterial.cu.h:17-173,
//   util/light/cudalight.cu.h:18-138, util/shape/cudatrianglemesh.cu:25-78 (attributes)
//   pbrt code being followed: core/reflection.cpp:60-67,115-175,311-319,522-580; lights/point.cpp:50-88;
//   lights/diffuse.cpp:69-101; shapes/disk.cpp:148-157; core/shape.cpp:95-108
#pragma once
#include "b2rt_device.cuh"

struct Surface {
    f3 p;        // world hit point  o + d*t
    f3 n;        // geometric normal (= shading normal: per-vertex normals are not used, like the reference's
                 // fallback when a mesh has none, cudatrianglemesh.cu:60-75)
    f3 sn, tn;   // BSDF frame: sn = Normalize(dpdu) with (dpdu,dpdv) = CoordinateSystem(n), tn = Cross(n, sn)
    float ray_eps;
    int32_t material, light;
};

// K6: from (slot, t) and the ray.  Triangle: n = Normalize(Cross(e2, e1)) which is what pbrt gets for meshes with
// default/degenerate uv (shapes/trianglemesh.cpp:181-201: CoordinateSystem(Normalize(Cross(e2,e1)),..));
// rayEpsilon = 1e-3*t (trianglemesh.cpp:224).  Sphere / disk: 5e-4*t (sphere.cpp:154, disk.cpp:100).
__device__ __forceinline__ Surface make_surface(const DScene &sc, const HitD &h, f3 o, f3 d) {
    Surface s;
    s.p = o + d * h.t;
    const float4 *pp = reinterpret_cast<const float4 *>(sc.recs + h.slot);
    float4 B = __ldg(pp + 1), C = __ldg(pp + 2);
    int type = __float_as_int(B.w);
    if (type == B2RT_PRIM_TRI) {
        s.n = normalize3(cross3(mk3(C.x, C.y, C.z), mk3(B.x, B.y, B.z)));
        s.ray_eps = 1e-3f * h.t;
    } else {
        const AnalyticPrim &A = sc.analytic[__float_as_int(C.w)];
        f3 no;
        if (type == B2RT_PRIM_SPHERE) xf_point(A.w2o, s.p.x, s.p.y, s.p.z, no.x, no.y, no.z);
        else no = mk3(0.f, 0.f, 1.f);
        s.n = normalize3(xf_normal(A.w2o, no));
        s.ray_eps = 5e-4f * h.t;
    }
    f3 dpdv;
    coordinate_system(s.n, s.sn, dpdv);
    s.sn = normalize3(s.sn);
    s.tn = cross3(s.n, s.sn);
    s.material = sc.prim_material ? sc.prim_material[h.prim] : -1;
    s.light = sc.prim_light ? sc.prim_light[h.prim] : -1;
    return s;
}

__device__ __forceinline__ f3 to_local(const Surface &s, f3 v) { return mk3(dot3(v, s.sn), dot3(v, s.tn), dot3(v, s.n)); }
__device__ __forceinline__ f3 to_world(const Surface &s, f3 v) {
    return mk3(s.sn.x * v.x + s.tn.x * v.y + s.n.x * v.z, s.sn.y * v.x + s.tn.y * v.y + s.n.y * v.z,
               s.sn.z * v.x + s.tn.z * v.y + s.n.z * v.z);
}

struct Mat { int type; f3 kd, kt; float ior; };
__device__ __forceinline__ Mat load_material(const DScene &sc, int id) {
    Mat m;
    if (id < 0 || id >= sc.n_materials) {      // grey matte 0.5 (cuda_render/util/material/cudamaterial.cpp:20)
        m.type = B2RT_MAT_MATTE; m.kd = mk3(.5f, .5f, .5f); m.kt = mk3(0.f, 0.f, 0.f); m.ior = 1.5f;
        return m;
    }
    const Material &M = sc.materials[id];
    m.type = M.type; m.kd = mk3(M.kd[0], M.kd[1], M.kd[2]); m.kt = mk3(M.kt[0], M.kt[1], M.kt[2]); m.ior = M.ior;
    return m;
}

// pbrt FresnelDielectric::Evaluate (core/reflection.cpp:115-137) + FrDiel (:60-67); scalar eta (RGB equal)
__device__ __forceinline__ float fresnel_dielectric(float cosi, float eta_i, float eta_t) {
    cosi = fminf(fmaxf(cosi, -1.f), 1.f);
    bool entering = cosi > 0.f;
    float ei = eta_i, et = eta_t;
    if (!entering) { float tmp = ei; ei = et; et = tmp; }
    float sint = ei / et * sqrtf(fmaxf(0.f, 1.f - cosi * cosi));
    if (sint >= 1.f) return 1.f;
    float cost = sqrtf(fmaxf(0.f, 1.f - sint * sint));
    float ac = fabsf(cosi);
    float Rparl = ((et * ac) - (ei * cost)) / ((et * ac) + (ei * cost));
    float Rperp = ((ei * ac) - (et * cost)) / ((ei * ac) + (et * cost));
    return (Rparl * Rparl + Rperp * Rperp) / 2.f;
}

// mirror lobe in the local frame: wi = (-wo.x,-wo.y,wo.z), f*|cos|/pdf = F*R  (pdf 1; reflection.cpp:138-144)
// glass transmission lobe (reflection.cpp:147-169); returns false on total internal reflection
__device__ __forceinline__ bool refract_local(f3 wo, float ior, f3 &wi) {
    bool entering = wo.z > 0.f;
    float ei = 1.f, et = ior;
    if (!entering) { ei = ior; et = 1.f; }
    float sini2 = fmaxf(0.f, 1.f - wo.z * wo.z);
    float eta = ei / et;
    float sint2 = eta * eta * sini2;
    if (sint2 >= 1.f) return false;
    float cost = sqrtf(fmaxf(0.f, 1.f - sint2));
    if (entering) cost = -cost;
    wi = mk3(eta * -wo.x, eta * -wo.y, cost);
    return true;
}

// BSDF::Sample_f with BSDF_ALL (core/reflection.cpp:522-580) specialised to the three materials the reference
// translates.  Returns weight = f*|cos(wi)|/pdf as pbrt would form it inside `alpha * fr * AbsDot / pdf`:
// the caller multiplies alpha by `f`, then by `absdot`, then divides by `pdf` (same order as photonmap.cpp:499-501).
struct BsdfSample { f3 wi; f3 f; float pdf; float absdot; bool specular; };
__device__ __forceinline__ bool sample_bsdf(const Surface &s, const Mat &m, f3 woW, float u1, float u2, float ucomp,
                                            BsdfSample &o) {
    f3 wo = to_local(s, woW);
    if (m.type == B2RT_MAT_MATTE) {
        f3 wi = cosine_sample_hemisphere(u1, u2);
        if (wo.z < 0.f) wi.z *= -1.f;
        o.pdf = (wo.z * wi.z > 0.f) ? fabsf(wi.z) * B2_INV_PI : 0.f;
        if (o.pdf == 0.f) return false;
        o.wi = to_world(s, wi);
        // f re-evaluated for non-specular lobes: reflection only when wi, wo on the same side of ng
        bool same = dot3(o.wi, s.n) * dot3(woW, s.n) > 0.f;
        o.f = same ? m.kd * B2_INV_PI : mk3(0.f, 0.f, 0.f);
        o.specular = false;
    } else {
        bool has_r = !is_black(m.kd), has_t = (m.type == B2RT_MAT_GLASS) && !is_black(m.kt);
        int ncomp = (has_r ? 1 : 0) + (has_t ? 1 : 0);
        if (ncomp == 0) return false;
        int which = min((int)floorf(ucomp * (float)ncomp), ncomp - 1);
        bool reflect = has_r && (which == 0);
        f3 wi;
        float F = (m.type == B2RT_MAT_GLASS) ? fresnel_dielectric(wo.z, 1.f, m.ior) : 1.f;
        if (reflect) {
            wi = mk3(-wo.x, -wo.y, wo.z);
            float ac = fabsf(wi.z);
            o.f = mk3(F * m.kd.x / ac, F * m.kd.y / ac, F * m.kd.z / ac);
        } else {
            if (!refract_local(wo, m.ior, wi)) return false;
            float ac = fabsf(wi.z);
            o.f = mk3((1.f - F) * m.kt.x / ac, (1.f - F) * m.kt.y / ac, (1.f - F) * m.kt.z / ac);
        }
        o.pdf = 1.f;
        if (ncomp > 1) o.pdf = o.pdf / (float)ncomp;
        o.wi = to_world(s, wi);
        o.specular = true;
    }
    o.absdot = fabsf(dot3(o.wi, s.n));
    return true;
}

// ---- lights ------------------------------------------------------------------------------------------------------
struct LightD { int type; f3 p, rgb; int prim; int nsamples; int analytic; float area; f3 n; };
__device__ __forceinline__ LightD load_light(const DScene &sc, int i) {
    const Light &L = sc.lights[i];
    LightD l;
    l.type = L.type; l.p = mk3(L.p[0], L.p[1], L.p[2]); l.rgb = mk3(L.rgb[0], L.rgb[1], L.rgb[2]);
    l.prim = L.prim; l.nsamples = L.nsamples; l.analytic = L.analytic; l.area = L.area;
    l.n = mk3(L.n[0], L.n[1], L.n[2]);
    return l;
}

// Distribution1D::SampleDiscrete (core/montecarlo.h) over light power luminance (core/integrator.cpp:261-268)
__device__ __forceinline__ int sample_light(const DScene &sc, float u, float &pdf) {
    int n = sc.n_lights, i = 0;
    while (i + 1 < n && u >= sc.light_cdf[i + 1]) ++i;
    pdf = sc.light_cdf[n + 1 + i];
    return i;
}

// point on the disk light for (u1,u2): Disk::Sample (shapes/disk.cpp:148-157)
__device__ __forceinline__ f3 disk_sample_point(const AnalyticPrim &D, float u1, float u2) {
    float x, y;
    concentric_sample_disk(u1, u2, x, y);
    x *= D.radius; y *= D.radius;
    f3 p;
    xf_point(D.o2w, x, y, D.height, p.x, p.y, p.z);
    return p;
}

// K8: emission.  Light::Sample_L(scene, ls, u1, u2, ...) for point (lights/point.cpp:80-88) and diffuse disk area
// lights (lights/diffuse.cpp:89-101).  Returns false when the path carries no energy.
// alpha = AbsDot(Nl, d) * Le / (pdf * lightPdf)   (integrators/photonmap.cpp:435)
__device__ __forceinline__ bool emit_photon(const DScene &sc, float u0, float u1, float u2, float u4, float u5, f3 &o,
                                            f3 &d, float &tmin, f3 &alpha) {
    float light_pdf;
    int li = sample_light(sc, u0, light_pdf);
    LightD L = load_light(sc, li);
    f3 Nl, Le;
    float pdf;
    if (L.type == B2RT_LIGHT_POINT) {
        o = L.p;
        d = uniform_sample_sphere(u1, u2);
        Nl = d; tmin = 0.f;
        pdf = 1.f / (4.f * B2_PI);
        Le = L.rgb;
    } else {
        const AnalyticPrim &D = sc.analytic[L.analytic];
        o = disk_sample_point(D, u1, u2);
        Nl = L.n;
        d = uniform_sample_sphere(u4, u5);
        if (dot3(d, Nl) < 0.f) d = d * -1.f;
        tmin = 1e-3f;
        pdf = (1.f / L.area) * B2_INV_TWOPI;
        Le = (dot3(Nl, d) > 0.f) ? L.rgb : mk3(0.f, 0.f, 0.f);
    }
    if (pdf == 0.f || is_black(Le)) return false;
    float ad = fabsf(dot3(Nl, d));
    float den = pdf * light_pdf;
    alpha = mk3(ad * Le.x / den, ad * Le.y / den, ad * Le.z / den);
    return !is_black(alpha);
}
