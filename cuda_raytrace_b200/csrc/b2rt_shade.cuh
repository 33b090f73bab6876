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
    f3 n;        // SHADING normal, BSDF::nn = dgShading.nn: the z axis of the BSDF frame, the cosine factors and the
                 // Faceforward of the photon lookups (core/reflection.cpp:601-609, integrators/photonmap.cpp:624-638)
    f3 ng;       // GEOMETRIC normal, dg.nn: which side wi / wo are on in BSDF::f (reflection.cpp:616), isect.Le, the
                 // normals of radiance photons and gather hits (photonmap.cpp:481, :672).  Equal to n unless the mesh has
                 // per-vertex normals.
    f3 sn, tn;   // BSDF frame: sn = Normalize(dgShading.dpdu), tn = Cross(nn, sn)
    float u, v;  // surface parametrisation (triangles with attributes only; 0 otherwise)
    float ray_eps;
    int32_t material, light;
};

// Triangle with per-vertex uv and / or normals, or one whose orientation is reversed: pbrt's Triangle::Intersect tail
// (shapes/trianglemesh.cpp:181-201, the uv determinant and its CoordinateSystem fallback) and
// Triangle::GetShadingGeometry (:311-389) followed by BSDF::BSDF, operation for operation (the reference's device code
// for the same job: cuda_render/util/shape/cudatrianglemesh.cu:25-78).  Normals arrive in WORLD space (the caller applies
// the mesh's object-to-world normal transform; pbrt applies it after the interpolation, the same value up to rounding
// and the same bits for the identity).  flags: bit 0 the mesh has normals, bit 1 it has uv, bit 2 ReverseOrientation ^
// TransformSwapsHandedness (core/diffgeom.cpp:52-54).  dndu / dndv only feed texture filtering and are not formed.
// I != NULL: the triangle belongs to an object instance.  pbrt's TransformedPrimitive::Intersect takes dg.nn (as a normal),
// dpdu and dpdv (as vectors) to world space BEFORE the shading geometry is formed, and GetShadingGeometry receives the whole
// object-to-world chain for the interpolated normal (core/primitive.cpp:104-117, shapes/trianglemesh.cpp:338): the
// Gram-Schmidt step below then runs in world space, which is not the same as transforming its object-space result when the
// instance transform is not rigid.
static __device__ __noinline__ void tri_surface_attr(const DScene &sc, const HitD &h, int tri, int flags, Surface &s,
                                                     const InstanceRec *I = nullptr) {
    const int i0 = sc.tri_idx[3 * tri], i1 = sc.tri_idx[3 * tri + 1], i2 = sc.tri_idx[3 * tri + 2];
    const f3 p1 = mk3(sc.verts[3 * i0], sc.verts[3 * i0 + 1], sc.verts[3 * i0 + 2]);
    const f3 p2 = mk3(sc.verts[3 * i1], sc.verts[3 * i1 + 1], sc.verts[3 * i1 + 2]);
    const f3 p3 = mk3(sc.verts[3 * i2], sc.verts[3 * i2 + 1], sc.verts[3 * i2 + 2]);
    float uv[3][2] = {{0.f, 0.f}, {1.f, 0.f}, {1.f, 1.f}};                  // Triangle::GetUVs defaults (trianglemesh.h:98-100)
    if (flags & 2) {
        uv[0][0] = sc.vuv[2 * i0]; uv[0][1] = sc.vuv[2 * i0 + 1];
        uv[1][0] = sc.vuv[2 * i1]; uv[1][1] = sc.vuv[2 * i1 + 1];
        uv[2][0] = sc.vuv[2 * i2]; uv[2][1] = sc.vuv[2 * i2 + 1];
    }
    const float du1 = uv[0][0] - uv[2][0], du2 = uv[1][0] - uv[2][0], dv1 = uv[0][1] - uv[2][1], dv2 = uv[1][1] - uv[2][1];
    const f3 dp1 = p1 - p3, dp2 = p2 - p3;
    const float determinant = du1 * dv2 - dv1 * du2;
    f3 dpdu, dpdv;
    if (determinant == 0.f) {
        const f3 e1 = p2 - p1, e2 = p3 - p1;
        coordinate_system(normalize3(cross3(e2, e1)), dpdu, dpdv);
    } else {
        const float invdet = 1.f / determinant;
        dpdu = (dp1 * dv2 - dp2 * dv1) * invdet;
        dpdv = (dp1 * -du2 + dp2 * du1) * invdet;
    }
    const float b0 = 1.f - h.b1 - h.b2;
    s.u = b0 * uv[0][0] + h.b1 * uv[1][0] + h.b2 * uv[2][0];
    s.v = b0 * uv[0][1] + h.b1 * uv[1][1] + h.b2 * uv[2][1];
    s.ng = normalize3(cross3(dpdu, dpdv));
    if (flags & 4) s.ng = s.ng * -1.f;
    if (I) {
        s.ng = normalize3(xf_normal(I->w2o, s.ng));
        xf_vec(I->o2w, dpdu.x, dpdu.y, dpdu.z, dpdu.x, dpdu.y, dpdu.z);
        xf_vec(I->o2w, dpdv.x, dpdv.y, dpdv.z, dpdv.x, dpdv.y, dpdv.z);
    }
    f3 sdpdu = dpdu;                        // dgShading.dpdu
    s.n = s.ng;
    if (flags & 1) {
        float b[3];
        const float A00 = uv[1][0] - uv[0][0], A01 = uv[2][0] - uv[0][0], A10 = uv[1][1] - uv[0][1], A11 = uv[2][1] - uv[0][1];
        const float C0 = s.u - uv[0][0], C1 = s.v - uv[0][1];
        const float det = A00 * A11 - A01 * A10;                               // SolveLinearSystem2x2 (core/transform.cpp:39-49)
        bool ok = !(fabsf(det) < 1e-10f);
        if (ok) {
            b[1] = (A11 * C0 - A01 * C1) / det;
            b[2] = (A00 * C1 - A10 * C0) / det;
            ok = !(isnan(b[1]) || isnan(b[2]));
        }
        if (!ok) b[0] = b[1] = b[2] = 1.f / 3.f;
        else b[0] = 1.f - b[1] - b[2];
        const f3 n0 = mk3(sc.vnormal[3 * i0], sc.vnormal[3 * i0 + 1], sc.vnormal[3 * i0 + 2]);
        const f3 n1 = mk3(sc.vnormal[3 * i1], sc.vnormal[3 * i1 + 1], sc.vnormal[3 * i1 + 2]);
        const f3 n2 = mk3(sc.vnormal[3 * i2], sc.vnormal[3 * i2 + 1], sc.vnormal[3 * i2 + 2]);
        f3 nsum = n0 * b[0] + n1 * b[1] + n2 * b[2];
        if (I) nsum = xf_normal(I->w2o, nsum);
        const f3 ns = normalize3(nsum);
        f3 ss = normalize3(dpdu);
        f3 ts = cross3(ss, ns);
        if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) {
            ts = normalize3(ts);
            ss = cross3(ts, ns);
        } else {
            coordinate_system(ns, ss, ts);
        }
        sdpdu = ss;
        s.n = normalize3(cross3(ss, ts));      // DifferentialGeometry ctor of dgShading (core/diffgeom.cpp:46)
        if (flags & 4) s.n = s.n * -1.f;
    }
    s.sn = normalize3(sdpdu);
    s.tn = cross3(s.n, s.sn);
}

// K6: from (slot, t) and the ray.  Triangle without attributes: pbrt's default uv (0,0), (1,0), (1,1) have determinant 1, so
// Triangle::Intersect forms dpdu = p2 - p1 and dpdv = p3 - p2 (shapes/trianglemesh.cpp:181-201, up to the rounding of its
// dp1 / dp2 sums) and dg.nn = Normalize(Cross(dpdu, dpdv)) = the direction of Cross(e1, e2): the winding decides the side,
// which matters for glass meshes (entering / leaving, core/reflection.cpp:147-151).  Here from the record's edges: n =
// Normalize(Cross(e1, e2)), sn = Normalize(e1).  (Until round 2 this was Cross(e2, e1) + CoordinateSystem -- pbrt's
// degenerate-uv fallback, the opposite side; found by the instancing tests against pbrt's own attributes.)
// rayEpsilon = 1e-3*t (trianglemesh.cpp:224).  Sphere / disk: 5e-4*t (sphere.cpp:154, disk.cpp:100).
// I != NULL (h.prim then local to the object, sc = the object's scene view, o / d the ray in the instance's space): the
// attributes are formed in the object's space exactly as for a top-level primitive and then taken to world space the way
// TransformedPrimitive::Intersect does -- p as a point, the normal by the inverse transpose and renormalised, dpdu as a
// vector (core/primitive.cpp:104-117); rayEpsilon stays the shape's.
__device__ __forceinline__ Surface make_surface_impl(const DScene &sc, const HitD &h, f3 o, f3 d, const InstanceRec *I) {
    Surface s;
    s.p = o + d * h.t;
    s.u = 0.f; s.v = 0.f;
    const float4 *pp = reinterpret_cast<const float4 *>(sc.recs + h.slot);
    float4 B = __ldg(pp + 1), C = __ldg(pp + 2);
    int type = __float_as_int(B.w);
    s.material = sc.prim_material ? sc.prim_material[h.prim] : -1;
    s.light = (!I && sc.prim_light) ? sc.prim_light[h.prim] : -1;     // pbrt: no area lights inside instances (core/api.cpp:1009-1011)
    if (I) xf_point(I->o2w, s.p.x, s.p.y, s.p.z, s.p.x, s.p.y, s.p.z);
    if (type == B2RT_PRIM_TRI) {
        s.ray_eps = 1e-3f * h.t;
        if (sc.tri_flags) {                                  // some mesh of the scene carries attributes
            const int tri = sc.prim_kind[h.prim];
            const int fl = sc.tri_flags[tri];
            if (fl) { tri_surface_attr(sc, h, tri, fl, s, I); return s; }
        }
        s.n = normalize3(cross3(mk3(B.x, B.y, B.z), mk3(C.x, C.y, C.z)));
        s.sn = mk3(B.x, B.y, B.z);
    } else {
        const AnalyticPrim &A = sc.analytic[__float_as_int(C.w)];
        f3 no;
        if (type == B2RT_PRIM_SPHERE) {
            const f3 po = I ? o + d * h.t : s.p;             // the hit point in the space the sphere's transform starts from
            xf_point(A.w2o, po.x, po.y, po.z, no.x, no.y, no.z);
        } else no = mk3(0.f, 0.f, 1.f);
        s.n = normalize3(xf_normal(A.w2o, no));
        s.ray_eps = 5e-4f * h.t;
        f3 dpdv;
        coordinate_system(s.n, s.sn, dpdv);
    }
    if (I) {
        s.n = normalize3(xf_normal(I->w2o, s.n));
        xf_vec(I->o2w, s.sn.x, s.sn.y, s.sn.z, s.sn.x, s.sn.y, s.sn.z);
    }
    s.ng = s.n;
    s.sn = normalize3(s.sn);
    s.tn = cross3(s.n, s.sn);
    return s;
}

// hit inside an object instance: the ray goes into the instance's space as the traversal took it there (instance_ray)
static __device__ __noinline__ Surface make_surface_instanced(const DScene &sc, const HitD &h, f3 o, f3 d) {
    const InstanceRec *I = sc.instances + h.inst;
    RayD r;
    instance_ray(I, o.x, o.y, o.z, d.x, d.y, d.z, 0.f, r);
    HitD lh = h;
    lh.prim = h.prim - I->prim_base;
    lh.inst = -1;
    const DScene *os = I->object;
    DScene osc = *os;
    osc.materials = sc.materials; osc.n_materials = sc.n_materials;
    return make_surface_impl(osc, lh, mk3(r.ox, r.oy, r.oz), mk3(r.dx, r.dy, r.dz), I);
}

template <bool INST>
__device__ __forceinline__ Surface make_surface(const DScene &sc, const HitD &h, f3 o, f3 d) {
    if (INST && h.inst >= 0) return make_surface_instanced(sc, h, o, d);
    return make_surface_impl(sc, h, o, d, nullptr);
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
        bool same = dot3(o.wi, s.ng) * dot3(woW, s.ng) > 0.f;
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
// SpotLight::Falloff (lights/spot.cpp:60-70): w in world space, pointing away from the light
__device__ __forceinline__ float spot_falloff(const Light &L, f3 w) {
    f3 wl = normalize3(mk3(L.w2l[0] * w.x + L.w2l[1] * w.y + L.w2l[2] * w.z, L.w2l[3] * w.x + L.w2l[4] * w.y + L.w2l[5] * w.z,
                           L.w2l[6] * w.x + L.w2l[7] * w.y + L.w2l[8] * w.z));
    const float costheta = wl.z;
    if (costheta < L.cos_total) return 0.f;
    if (costheta > L.cos_falloff) return 1.f;
    const float delta = (costheta - L.cos_total) / (L.cos_falloff - L.cos_total);
    return delta * delta * delta * delta;
}
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
    } else if (L.type == B2RT_LIGHT_SPOT) {
        // SpotLight::Sample_L (lights/spot.cpp:112-121): UniformSampleCone in light space (core/montecarlo.cpp:405-410)
        const Light &G = sc.lights[li];
        const float costheta = (1.f - u1) + u1 * G.cos_total;
        const float sintheta = sqrtf(1.f - costheta * costheta);
        float sn, cs;
        b2_sincos_quarter(4.f * u2, sn, cs);
        const f3 v = mk3(cs * sintheta, sn * sintheta, costheta);
        o = L.p;
        d = mk3(G.l2w[0] * v.x + G.l2w[1] * v.y + G.l2w[2] * v.z, G.l2w[3] * v.x + G.l2w[4] * v.y + G.l2w[5] * v.z,
                G.l2w[6] * v.x + G.l2w[7] * v.y + G.l2w[8] * v.z);
        Nl = d; tmin = 0.f;
        pdf = (float)(1.0 / ((double)(2.f) * 3.14159265358979323846 * (double)(1.f - G.cos_total)));      // UniformConePdf
        Le = L.rgb * spot_falloff(G, d);
    } else if (L.type == B2RT_LIGHT_DISTANT) {
        // DistantLight::Sample_L (lights/distant.cpp:84-105): a disk of the scene's bounding radius facing the light
        const Light &G = sc.lights[li];
        f3 v1, v2;
        coordinate_system(L.n, v1, v2);
        float d1, d2;
        concentric_sample_disk(u1, u2, d1, d2);
        const f3 wc = mk3(G.wc[0], G.wc[1], G.wc[2]);
        const f3 pdisk = wc + (v1 * d1 + v2 * d2) * G.wr;
        o = pdisk + L.n * G.wr;
        d = -L.n;
        Nl = d; tmin = 0.f;
        pdf = (float)(1.0 / (3.14159265358979323846 * (double)G.wr * (double)G.wr));
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
