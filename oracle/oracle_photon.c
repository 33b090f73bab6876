/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the photon pass, the photon hash grid and the radiance
 * estimate.  Never linked into or called from the product path (cuda_raytrace_b200/).
 *
 * What it restates, and from where:
 *   - path logic, deposit rules, Russian roulette: pbrt-v2/src/integrators/photonmap.cpp:416-514 (finalgather off)
 *   - emission: lights/point.cpp:80-88, lights/diffuse.cpp:89-101, shapes/disk.cpp:148-157, core/integrator.cpp:261-268
 *   - BSDFs: core/reflection.cpp:60-67 (FrDiel), 115-137 (FresnelDielectric), 138-169 (specular lobes),
 *            311-319 (cosine lobe), 522-580 (BSDF::Sample_f)
 *   - sampling: core/montecarlo.cpp:283-290, 306-348; core/montecarlo.h:128-133; core/geometry.h:508-518
 *   - estimator: integrators/photonmap.cpp:187-260 (PhotonProcess heap + LPhoton), core/kdtree.h:151-184
 * Deliberate differences from pbrt, shared with the GPU path and stated in DESIGN.md: random numbers come from
 * Philox-4x32-10 keyed by (seed, path, bounce) instead of Halton + MT19937; sinf/cosf are a fixed polynomial
 * (b2_sincos_quarter) so that CPU and GPU agree bit for bit; the photon container is a hash grid, not a kd-tree
 * (the kNN set is the same up to distance ties); shading frame = CoordinateSystem(n) for every primitive.
 * Parity with pbrt itself is pinned statistically (tests/test_oracle_photon_cpu.py: flux, counts, estimator vs
 * pbrt's own KdTree + LPhoton through oracle/_ref) -- photon-by-photon parity with pbrt is impossible by design.
 *
 * Compile with -ffp-contract=off.                                                                              */
#define _GNU_SOURCE
#include "oracle_port.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define O_PI 3.14159265358979323846f
#define O_INV_PI 0.31830988618379067154f
#define O_INV_TWOPI 0.15915494309189533577f

typedef struct { float x, y, z; } v3;
static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vscale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline int black(v3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }
static inline float lumin(v3 c) { return 0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z; }
static inline v3 vcross(v3 a, v3 b) {   /* pbrt Cross: double products (core/geometry.h:477-484) */
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return V((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
static inline v3 vnormalize(v3 v) {
    float inv = 1.f / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    return V(v.x * inv, v.y * inv, v.z * inv);
}
static void coord_system(v3 v1, v3 *v2, v3 *v3o) {
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        *v2 = V(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        *v2 = V(0.f, v1.z * invLen, -v1.y * invLen);
    }
    *v3o = vcross(v1, *v2);
}

/* Philox-4x32-10 (Salmon, Moraes, Dror, Shaw 2011), the published algorithm */
static void philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
void port_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    philox(c0, c1, c2, c3, k0, k1, out);
}
static inline float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }

void port_sincos_quarter(float q, float *s, float *c) {
    float kf = floorf(q + 0.5f);
    int k = (int)kf;
    float f = q - kf;
    float x = f * 1.57079632679489661923f;
    float z = x * x;
    float ps = ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z + -1.6666654611e-1f) * z * x + x;
    float pc = ((2.443315711809948e-5f * z + -1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z +
               (1.f - 0.5f * z);
    switch (k & 3) {
        case 0: *s = ps; *c = pc; break;
        case 1: *s = pc; *c = -ps; break;
        case 2: *s = -ps; *c = -pc; break;
        default: *s = -pc; *c = ps; break;
    }
}
static void concentric_disk(float u1, float u2, float *dx, float *dy) {
    float r, theta;
    float sx = 2.f * u1 - 1.f, sy = 2.f * u2 - 1.f;
    if (sx == 0.f && sy == 0.f) { *dx = 0.f; *dy = 0.f; return; }
    if (sx >= -sy) {
        if (sx > sy) { r = sx; theta = (sy > 0.f) ? sy / r : 8.f + sy / r; }
        else { r = sy; theta = 2.f - sx / r; }
    } else {
        if (sx <= sy) { r = -sx; theta = 4.f - sy / r; }
        else { r = -sy; theta = 6.f + sx / r; }
    }
    float s, c;
    port_sincos_quarter(theta * 0.5f, &s, &c);
    *dx = r * c; *dy = r * s;
}
static v3 cosine_hemisphere(float u1, float u2) {
    float x, y;
    concentric_disk(u1, u2, &x, &y);
    return V(x, y, sqrtf(fmaxf(0.f, 1.f - x * x - y * y)));
}
static v3 uniform_sphere(float u1, float u2) {
    float z = 1.f - 2.f * u1;
    float r = sqrtf(fmaxf(0.f, 1.f - z * z));
    float s, c;
    port_sincos_quarter(4.f * u2, &s, &c);
    return V(r * c, r * s, z);
}

/* ---- scene access ---------------------------------------------------------------------------------------------- */
typedef struct { int type; v3 kd, kt; float ior; } omat;
static omat get_material(const port_scene *sc, int prim) {
    omat m;
    int id = sc->prim_material ? sc->prim_material[prim] : -1;
    if (id < 0 || id >= sc->nmaterials) { m.type = 0; m.kd = V(.5f, .5f, .5f); m.kt = V(0, 0, 0); m.ior = 1.5f; return m; }
    const float *M = sc->materials + 8 * id;
    m.type = (int)M[0]; m.kd = V(M[1], M[2], M[3]); m.kt = V(M[4], M[5], M[6]); m.ior = M[7];
    return m;
}
static void xfp(const float *m, v3 p, v3 *o) {
    o->x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    o->y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    o->z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
}
static v3 xfn(const float *w2o, v3 n) {   /* Normal transform = transpose of the inverse */
    return V(w2o[0] * n.x + w2o[4] * n.y + w2o[8] * n.z, w2o[1] * n.x + w2o[5] * n.y + w2o[9] * n.z,
             w2o[2] * n.x + w2o[6] * n.y + w2o[10] * n.z);
}

typedef struct { v3 p, n, ng, sn, tn; float ray_eps, u, v, t; } osurf;   /* n = shading normal (BSDF nn), ng = geometric dg.nn */
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
/* triangle with per-vertex normals / uv / reversed orientation: pbrt shapes/trianglemesh.cpp:181-201 (dpdu, dpdv from the uv
 * determinant or CoordinateSystem), :311-389 (GetShadingGeometry), core/diffgeom.cpp:46-54, core/reflection.cpp:601-609 */
static void tri_attr(const port_scene *sc, int tri, int fl, float b1, float b2, osurf *s) {
    const int32_t *vi = sc->idx + 3 * (int64_t)tri;
    const float *q1 = sc->verts + 3 * (int64_t)vi[0], *q2 = sc->verts + 3 * (int64_t)vi[1], *q3 = sc->verts + 3 * (int64_t)vi[2];
    const v3 p1 = V(q1[0], q1[1], q1[2]), p2 = V(q2[0], q2[1], q2[2]), p3 = V(q3[0], q3[1], q3[2]);
    float uv[3][2] = {{0.f, 0.f}, {1.f, 0.f}, {1.f, 1.f}};
    if (fl & 2)
        for (int j = 0; j < 3; ++j) { uv[j][0] = sc->vuv[2 * (int64_t)vi[j]]; uv[j][1] = sc->vuv[2 * (int64_t)vi[j] + 1]; }
    const float du1 = uv[0][0] - uv[2][0], du2 = uv[1][0] - uv[2][0], dv1 = uv[0][1] - uv[2][1], dv2 = uv[1][1] - uv[2][1];
    const v3 dp1 = vsub(p1, p3), dp2 = vsub(p2, p3);
    const float determinant = du1 * dv2 - dv1 * du2;
    v3 dpdu, dpdv;
    if (determinant == 0.f) {
        coord_system(vnormalize(vcross(vsub(p3, p1), vsub(p2, p1))), &dpdu, &dpdv);
    } else {
        const float invdet = 1.f / determinant;
        dpdu = vscale(vsub(vscale(dp1, dv2), vscale(dp2, dv1)), invdet);
        dpdv = vscale(vadd(vscale(dp1, -du2), vscale(dp2, du1)), invdet);
    }
    const float b0 = 1.f - b1 - b2;
    s->u = b0 * uv[0][0] + b1 * uv[1][0] + b2 * uv[2][0];
    s->v = b0 * uv[0][1] + b1 * uv[1][1] + b2 * uv[2][1];
    s->ng = vnormalize(vcross(dpdu, dpdv));
    if (fl & 4) s->ng = vscale(s->ng, -1.f);
    v3 sdpdu = dpdu;
    s->n = s->ng;
    if (fl & 1) {
        float b[3];
        const float A00 = uv[1][0] - uv[0][0], A01 = uv[2][0] - uv[0][0], A10 = uv[1][1] - uv[0][1], A11 = uv[2][1] - uv[0][1];
        const float C0 = s->u - uv[0][0], C1 = s->v - uv[0][1];
        const float det = A00 * A11 - A01 * A10;
        int ok = !(fabsf(det) < 1e-10f);
        if (ok) {
            b[1] = (A11 * C0 - A01 * C1) / det;
            b[2] = (A00 * C1 - A10 * C0) / det;
            ok = !(isnan(b[1]) || isnan(b[2]));
        }
        if (!ok) b[0] = b[1] = b[2] = 1.f / 3.f;
        else b[0] = 1.f - b[1] - b[2];
        const float *m0 = sc->vnormal + 3 * (int64_t)vi[0], *m1 = sc->vnormal + 3 * (int64_t)vi[1], *m2 = sc->vnormal + 3 * (int64_t)vi[2];
        const v3 ns = vnormalize(vadd(vadd(vscale(V(m0[0], m0[1], m0[2]), b[0]), vscale(V(m1[0], m1[1], m1[2]), b[1])),
                                      vscale(V(m2[0], m2[1], m2[2]), b[2])));
        v3 ss = vnormalize(dpdu);
        v3 ts = vcross(ss, ns);
        if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) {
            ts = vnormalize(ts);
            ss = vcross(ts, ns);
        } else {
            coord_system(ns, &ss, &ts);
        }
        sdpdu = ss;
        s->n = vnormalize(vcross(ss, ts));
        if (fl & 4) s->n = vscale(s->n, -1.f);
    }
    s->sn = vnormalize(sdpdu);
    s->tn = vcross(s->n, s->sn);
}
/* exhaustive closest hit (min t, min prim id) + hit attributes */
static int closest_range(const port_scene *sc, v3 o, v3 d, float tmin, float tmax, osurf *s, int *prim_out) {
    float ray[8] = {o.x, o.y, o.z, tmin, d.x, d.y, d.z, tmax};
    float t, b1, b2; int32_t prim;
    port_intersect_brute(sc->nprims, sc->kind, sc->verts, sc->idx, sc->analytic, 1, ray, &t, &prim, &b1, &b2, 1);
    if (prim < 0) return 0;
    *prim_out = prim;
    s->p = vadd(o, vscale(d, t));
    s->t = t; s->u = 0.f; s->v = 0.f;
    int k = sc->kind[prim];
    if (k >= 0) {
        s->ray_eps = 1e-3f * t;
        const int fl = sc->tri_flags ? sc->tri_flags[k] : 0;
        if (fl) { tri_attr(sc, k, fl, b1, b2, s); return 1; }
        const int32_t *vi = sc->idx + 3 * (int64_t)k;
        const float *p1 = sc->verts + 3 * (int64_t)vi[0], *p2 = sc->verts + 3 * (int64_t)vi[1], *p3 = sc->verts + 3 * (int64_t)vi[2];
        v3 e1 = V(p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]), e2 = V(p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2]);
        s->n = vnormalize(vcross(e1, e2));      /* pbrt: Cross(dpdu, dpdv) with the default uv = the side of Cross(e1, e2) */
        s->sn = e1;                             /* dpdu = p2 - p1 */
    } else {
        const float *A = sc->analytic + PORT_ANALYTIC_STRIDE * (int64_t)(~k);
        v3 no;
        if ((int)A[27] == 1) xfp(A + 12, s->p, &no); else no = V(0.f, 0.f, 1.f);
        s->n = vnormalize(xfn(A + 12, no));
        s->ray_eps = 5e-4f * t;
        v3 dpdv;
        coord_system(s->n, &s->sn, &dpdv);
    }
    s->ng = s->n;
    s->sn = vnormalize(s->sn);
    s->tn = vcross(s->n, s->sn);
    return 1;
}
static int closest(const port_scene *sc, v3 o, v3 d, float tmin, osurf *s, int *prim_out) {
    return closest_range(sc, o, d, tmin, INFINITY, s, prim_out);
}
void port_surface_at(const port_scene *sc, int64_t nrays, const float *rays, float *out) {
    for (int64_t i = 0; i < nrays; ++i) {
        const float *r = rays + 8 * i;
        float *o = out + 20 * i;
        osurf s; int prim = -1;
        memset(o, 0, 80);
        if (!closest_range(sc, V(r[0], r[1], r[2]), V(r[4], r[5], r[6]), r[3], r[7], &s, &prim)) {
            int32_t m1 = -1; memcpy(o + 19, &m1, 4);
            continue;
        }
        o[0] = s.p.x; o[1] = s.p.y; o[2] = s.p.z; o[3] = s.t;
        o[4] = s.ng.x; o[5] = s.ng.y; o[6] = s.ng.z; o[7] = s.u;
        o[8] = s.n.x; o[9] = s.n.y; o[10] = s.n.z; o[11] = s.v;
        o[12] = s.sn.x; o[13] = s.sn.y; o[14] = s.sn.z; o[15] = s.ray_eps;
        o[16] = s.tn.x; o[17] = s.tn.y; o[18] = s.tn.z;
        int32_t pi = prim; memcpy(o + 19, &pi, 4);
    }
}
static v3 to_local(const osurf *s, v3 v) { return V(vdot(v, s->sn), vdot(v, s->tn), vdot(v, s->n)); }
static v3 to_world(const osurf *s, v3 v) {
    return V(s->sn.x * v.x + s->tn.x * v.y + s->n.x * v.z, s->sn.y * v.x + s->tn.y * v.y + s->n.y * v.z,
             s->sn.z * v.x + s->tn.z * v.y + s->n.z * v.z);
}

static float fresnel_dielectric(float cosi, float eta_i, float eta_t) {
    cosi = fminf(fmaxf(cosi, -1.f), 1.f);
    float ei = eta_i, et = eta_t;
    if (!(cosi > 0.f)) { float tmp = ei; ei = et; et = tmp; }
    float sint = ei / et * sqrtf(fmaxf(0.f, 1.f - cosi * cosi));
    if (sint >= 1.f) return 1.f;
    float cost = sqrtf(fmaxf(0.f, 1.f - sint * sint));
    float ac = fabsf(cosi);
    float Rparl = ((et * ac) - (ei * cost)) / ((et * ac) + (ei * cost));
    float Rperp = ((ei * ac) - (et * cost)) / ((ei * ac) + (et * cost));
    return (Rparl * Rparl + Rperp * Rperp) / 2.f;
}

typedef struct { v3 wi, f; float pdf, absdot; int specular; } osample;
static int sample_bsdf(const osurf *s, const omat *m, v3 woW, float u1, float u2, float ucomp, osample *o) {
    v3 wo = to_local(s, woW);
    if (m->type == 0) {
        v3 wi = cosine_hemisphere(u1, u2);
        if (wo.z < 0.f) wi.z *= -1.f;
        o->pdf = (wo.z * wi.z > 0.f) ? fabsf(wi.z) * O_INV_PI : 0.f;
        if (o->pdf == 0.f) return 0;
        o->wi = to_world(s, wi);
        int same = vdot(o->wi, s->ng) * vdot(woW, s->ng) > 0.f;
        o->f = same ? vscale(m->kd, O_INV_PI) : V(0, 0, 0);
        o->specular = 0;
    } else {
        int has_r = !black(m->kd), has_t = (m->type == 2) && !black(m->kt);
        int ncomp = has_r + has_t;
        if (ncomp == 0) return 0;
        int which = (int)floorf(ucomp * (float)ncomp);
        if (which > ncomp - 1) which = ncomp - 1;
        int reflect = has_r && which == 0;
        float F = (m->type == 2) ? fresnel_dielectric(wo.z, 1.f, m->ior) : 1.f;
        v3 wi;
        if (reflect) {
            wi = V(-wo.x, -wo.y, wo.z);
            float ac = fabsf(wi.z);
            o->f = V(F * m->kd.x / ac, F * m->kd.y / ac, F * m->kd.z / ac);
        } else {
            int entering = wo.z > 0.f;
            float ei = 1.f, et = m->ior;
            if (!entering) { ei = m->ior; et = 1.f; }
            float sini2 = fmaxf(0.f, 1.f - wo.z * wo.z);
            float eta = ei / et;
            float sint2 = eta * eta * sini2;
            if (sint2 >= 1.f) return 0;
            float cost = sqrtf(fmaxf(0.f, 1.f - sint2));
            if (entering) cost = -cost;
            wi = V(eta * -wo.x, eta * -wo.y, cost);
            float ac = fabsf(wi.z);
            o->f = V((1.f - F) * m->kt.x / ac, (1.f - F) * m->kt.y / ac, (1.f - F) * m->kt.z / ac);
        }
        o->pdf = 1.f;
        if (ncomp > 1) o->pdf = o->pdf / (float)ncomp;
        o->wi = to_world(s, wi);
        o->specular = 1;
    }
    o->absdot = fabsf(vdot(o->wi, s->n));
    return 1;
}

static uint32_t oct_encode(v3 v) {
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
static v3 oct_decode(uint32_t w) {
    float x = (float)(int)(int16_t)(w & 0xffffu) / 32767.f, y = (float)(int)(int16_t)(w >> 16) / 32767.f;
    float z = 1.f - fabsf(x) - fabsf(y);
    if (z < 0.f) {
        float ox = (1.f - fabsf(y)) * (x >= 0.f ? 1.f : -1.f);
        float oy = (1.f - fabsf(x)) * (y >= 0.f ? 1.f : -1.f);
        x = ox; y = oy;
    }
    return vnormalize(V(x, y, z));
}

/* ---- lights + emission ------------------------------------------------------------------------------------------ */
typedef struct { int type; v3 p, rgb; int analytic; float area; v3 n; const float *ext; v3 wc; float wr; } olight;
/* SpotLight::Falloff (lights/spot.cpp:60-70) */
static float spot_falloff(const olight *L, v3 w) {
    const float *m = L->ext + 9;
    v3 wl = vnormalize(V(m[0] * w.x + m[1] * w.y + m[2] * w.z, m[3] * w.x + m[4] * w.y + m[5] * w.z, m[6] * w.x + m[7] * w.y + m[8] * w.z));
    float costheta = wl.z, ct = L->ext[18], cf = L->ext[19];
    if (costheta < ct) return 0.f;
    if (costheta > cf) return 1.f;
    float delta = (costheta - ct) / (cf - ct);
    return delta * delta * delta * delta;
}
static void derive_lights(const port_scene *sc, olight *L, float *cdf /* 2n+2 */) {
    int n = sc->nlights;
    float power[64];
    for (int i = 0; i < n; ++i) {
        const float *l = sc->lights + 8 * i;
        L[i].type = (int)l[0]; L[i].p = V(l[1], l[2], l[3]); L[i].rgb = V(l[4], l[5], l[6]);
        L[i].analytic = -1; L[i].area = 0.f; L[i].n = V(0, 0, 1);
        L[i].ext = sc->light_ext ? sc->light_ext + 24 * i : NULL; L[i].wc = V(0, 0, 0); L[i].wr = 0.f;
        v3 P;
        if (L[i].type == 0) {
            float f = (float)(4.f * M_PI);                       /* PointLight::Power, lights/point.cpp:60-62 */
            P = V(f * L[i].rgb.x, f * L[i].rgb.y, f * L[i].rgb.z);
        } else if (L[i].type == 2) {                             /* SpotLight::Power, lights/spot.cpp:72-75 */
            float f = (float)M_PI, g = 1.f - .5f * (L[i].ext[19] + L[i].ext[18]);
            P = V(L[i].rgb.x * 2.f * f * g, L[i].rgb.y * 2.f * f * g, L[i].rgb.z * 2.f * f * g);
        } else if (L[i].type == 3) {                             /* DistantLight::Power over BBox::BoundingSphere */
            const float *b = sc->world_bounds;
            float c[3], d2 = 0.f;
            for (int a = 0; a < 3; ++a) { c[a] = .5f * b[a] + .5f * b[3 + a]; float d = c[a] - b[3 + a]; d2 += d * d; }
            L[i].wc = V(c[0], c[1], c[2]); L[i].wr = sqrtf(d2);
            L[i].n = V(L[i].ext[20], L[i].ext[21], L[i].ext[22]);
            float f = (float)M_PI;
            P = V(L[i].rgb.x * f * L[i].wr * L[i].wr, L[i].rgb.y * f * L[i].wr * L[i].wr, L[i].rgb.z * f * L[i].wr * L[i].wr);
        } else {
            int prim = (int)l[7];
            int a = ~sc->kind[prim];
            const float *D = sc->analytic + PORT_ANALYTIC_STRIDE * (int64_t)a;
            float phimax = ((float)M_PI / 180.f) * 360.f;
            L[i].analytic = a;
            L[i].area = phimax * 0.5f * (D[24] * D[24] - D[26] * D[26]);       /* Disk::Area */
            L[i].n = vnormalize(V(D[12 + 8], D[12 + 9], D[12 + 10]));   /* w2o^T (0,0,1) */
            float pi = (float)M_PI;                              /* DiffuseAreaLight::Power = Lemit * area * M_PI */
            P = V(L[i].rgb.x * L[i].area * pi, L[i].rgb.y * L[i].area * pi, L[i].rgb.z * L[i].area * pi);
        }
        power[i] = lumin(P);
    }
    cdf[0] = 0.f;
    for (int i = 1; i <= n; ++i) cdf[i] = cdf[i - 1] + power[i - 1] / (float)n;
    float fi = cdf[n];
    for (int i = 1; i <= n; ++i) cdf[i] = fi > 0.f ? cdf[i] / fi : (float)i / (float)n;
    for (int i = 0; i < n; ++i) cdf[n + 1 + i] = fi > 0.f ? power[i] / (fi * (float)n) : 1.f / (float)n;
}

/* Light::Sample_L for photon emission: point (lights/point.cpp:80-88), diffuse disk (lights/diffuse.cpp:89-101), spot
 * (lights/spot.cpp:112-121), distant (lights/distant.cpp:84-105).  r0 / r1: the path's first two Philox blocks. */
static void o_emit(const port_scene *sc, const olight *Ls, const uint32_t r0[4], const uint32_t r1[4], v3 *o, v3 *d, v3 *Nl,
                   v3 *Le, float *tmin, float *pdf) {
    if (Ls->type == 0) {
        *o = Ls->p; *d = uniform_sphere(u01(r0[1]), u01(r0[2])); *Nl = *d; *tmin = 0.f;
        *pdf = 1.f / (4.f * O_PI); *Le = Ls->rgb;
    } else if (Ls->type == 2) {
        const float u1 = u01(r0[1]), u2 = u01(r0[2]), ct = Ls->ext[18];
        const float costheta = (1.f - u1) + u1 * ct;
        const float sintheta = sqrtf(1.f - costheta * costheta);
        float sn, cs;
        port_sincos_quarter(4.f * u2, &sn, &cs);
        const v3 v = V(cs * sintheta, sn * sintheta, costheta);
        const float *m = Ls->ext;
        *o = Ls->p;
        *d = V(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z);
        *Nl = *d; *tmin = 0.f;
        *pdf = (float)(1.0 / ((double)(2.f) * 3.14159265358979323846 * (double)(1.f - ct)));
        *Le = vscale(Ls->rgb, spot_falloff(Ls, *d));
    } else if (Ls->type == 3) {
        v3 v1, v2;
        coord_system(Ls->n, &v1, &v2);
        float d1, d2;
        concentric_disk(u01(r0[1]), u01(r0[2]), &d1, &d2);
        const v3 pdisk = vadd(Ls->wc, vscale(vadd(vscale(v1, d1), vscale(v2, d2)), Ls->wr));
        *o = vadd(pdisk, vscale(Ls->n, Ls->wr));
        *d = vneg(Ls->n);
        *Nl = *d; *tmin = 0.f;
        *pdf = (float)(1.0 / (3.14159265358979323846 * (double)Ls->wr * (double)Ls->wr));
        *Le = Ls->rgb;
    } else {
        const float *D = sc->analytic + PORT_ANALYTIC_STRIDE * (int64_t)Ls->analytic;
        float x, y;
        concentric_disk(u01(r0[1]), u01(r0[2]), &x, &y);
        x *= D[24]; y *= D[24];
        xfp(D, V(x, y, D[25]), o);
        *Nl = Ls->n;
        *d = uniform_sphere(u01(r1[0]), u01(r1[1]));
        if (vdot(*d, *Nl) < 0.f) *d = vscale(*d, -1.f);
        *tmin = 1e-3f;
        *pdf = (1.f / Ls->area) * O_INV_TWOPI;
        *Le = (vdot(*Nl, *d) > 0.f) ? Ls->rgb : V(0, 0, 0);
    }
}

int port_photon_trace(const port_scene *sc, uint64_t seed, int64_t first_path, int64_t n_paths, int max_depth,
                      int flags, port_photon *out_c, int64_t cap_c, int64_t *n_c, port_photon *out_i, int64_t cap_i,
                      int64_t *n_i) {
    olight L[64]; float cdf[130];
    if (sc->nlights < 1 || sc->nlights > 64) return -2;
    derive_lights(sc, L, cdf);
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    const int want_c = flags & 1, want_i = flags & 2;
    int64_t nc = *n_c, ni = *n_i;
    for (int64_t li = 0; li < n_paths; ++li) {
        uint64_t path = (uint64_t)(first_path + li);
        uint32_t p_lo = (uint32_t)path, p_hi = (uint32_t)(path >> 32);
        uint32_t r0[4], r1[4];
        philox(p_lo, p_hi, 0u, 0u, k0, k1, r0);
        philox(p_lo, p_hi, 0u, 1u, k0, k1, r1);
        /* light choice (Distribution1D::SampleDiscrete) */
        float u0 = u01(r0[0]);
        int lsel = 0;
        while (lsel + 1 < sc->nlights && u0 >= cdf[lsel + 1]) ++lsel;
        float light_pdf = cdf[sc->nlights + 1 + lsel];
        const olight *Ls = &L[lsel];
        v3 o, d, Nl, Le; float tmin, pdf;
        o_emit(sc, Ls, r0, r1, &o, &d, &Nl, &Le, &tmin, &pdf);
        if (pdf == 0.f || black(Le)) continue;
        float ad = fabsf(vdot(Nl, d)), den = pdf * light_pdf;
        v3 alpha = V(ad * Le.x / den, ad * Le.y / den, ad * Le.z / den);
        if (black(alpha)) continue;
        int specular_path = 1, n_isect = 0;
        for (;;) {
            osurf s; int prim;
            if (!closest(sc, o, d, tmin, &s, &prim)) break;
            ++n_isect;
            omat m = get_material(sc, prim);
            v3 wo = vneg(d);
            if (m.type == 0) {
                int caustic = specular_path && n_isect > 1;
                int dep = caustic ? want_c : (n_isect > 1 && want_i);
                if (dep) {
                    port_photon ph;
                    ph.p[0] = s.p.x; ph.p[1] = s.p.y; ph.p[2] = s.p.z; ph.wi_oct = oct_encode(wo);
                    ph.alpha[0] = alpha.x; ph.alpha[1] = alpha.y; ph.alpha[2] = alpha.z; ph.path = p_lo;
                    if (caustic) { if (nc >= cap_c) return -1; out_c[nc++] = ph; }
                    else { if (ni >= cap_i) return -1; out_i[ni++] = ph; }
                }
            }
            if (n_isect >= max_depth) break;
            uint32_t rb[4];
            philox(p_lo, p_hi, (uint32_t)n_isect, 0u, k0, k1, rb);
            osample bs;
            if (!sample_bsdf(&s, &m, wo, u01(rb[0]), u01(rb[1]), u01(rb[2]), &bs)) break;
            if (black(bs.f) || bs.pdf == 0.f) break;
            v3 anew = V(alpha.x * bs.f.x * bs.absdot / bs.pdf, alpha.y * bs.f.y * bs.absdot / bs.pdf,
                        alpha.z * bs.f.z * bs.absdot / bs.pdf);
            float cont = fminf(1.f, lumin(anew) / lumin(alpha));
            if (u01(rb[3]) > cont) break;
            alpha = V(anew.x / cont, anew.y / cont, anew.z / cont);
            specular_path = specular_path && bs.specular;
            if (!want_i && !specular_path) break;
            o = s.p; d = bs.wi; tmin = s.ray_eps;
        }
    }
    *n_c = nc; *n_i = ni;
    return 0;
}

/* ---- hash grid (definition in csrc/photon.cu header comment / DESIGN.md) ------------------------------------------ */
void port_grid_build(int64_t n, const port_photon *ph, float cell_size, float target_per_cell, float origin[3],
                     int32_t dims[3], float *cell, float *inv_out, uint32_t *keys, uint32_t *order,
                     uint32_t *cell_start) {
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (n > 0) {
        for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
        for (int64_t i = 0; i < n; ++i)
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], ph[i].p[a]); hi[a] = fmaxf(hi[a], ph[i].p[a]); }
    }
    float ext[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
    float c = cell_size;
    if (!(c > 0.f)) {
        float target = target_per_cell > 0.f ? target_per_cell : 8.f;
        float A = 2.f * (ext[0] * ext[1] + ext[1] * ext[2] + ext[2] * ext[0]);
        float mx = fmaxf(ext[0], fmaxf(ext[1], ext[2]));
        c = (A > 0.f && n > 0) ? sqrtf(A * target / (float)n) : (mx > 0.f ? mx / 64.f : 1.f);
    }
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
    for (int a = 0; a < 3; ++a) origin[a] = lo[a];
    *cell = c; *inv_out = inv;
    if (!cell_start) return;
    int64_t ncell = (int64_t)dims[0] * dims[1] * dims[2];
    memset(cell_start, 0, (size_t)(ncell + 1) * 4);
    uint32_t *k0 = (uint32_t *)malloc((size_t)(n ? n : 1) * 4);
    for (int64_t i = 0; i < n; ++i) {
        int ci[3];
        for (int a = 0; a < 3; ++a) {
            int v = (int)floorf((ph[i].p[a] - lo[a]) * inv);
            ci[a] = v < 0 ? 0 : (v > dims[a] - 1 ? dims[a] - 1 : v);
        }
        k0[i] = (uint32_t)((ci[2] * dims[1] + ci[1]) * dims[0] + ci[0]);
        cell_start[k0[i] + 1]++;
    }
    for (int64_t k = 0; k < ncell; ++k) cell_start[k + 1] += cell_start[k];   /* cell_start[k] = #keys < k */
    uint32_t *cur = (uint32_t *)malloc((size_t)(ncell + 1) * 4);
    memcpy(cur, cell_start, (size_t)(ncell + 1) * 4);
    for (int64_t i = 0; i < n; ++i) {          /* counting sort = stable sort by key of the path-order indices */
        uint32_t pos = cur[k0[i]]++;
        keys[pos] = k0[i]; order[pos] = (uint32_t)i;
    }
    free(cur); free(k0);
}

/* ---- radiance estimate: brute-force k nearest by (d2, index) + LPhoton (photonmap.cpp:207-260), diffuse surface;
 * the sum runs over the selected set in ascending photon index ---- */
static int cmp_u64(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}
static int cmp_idx(const void *a, const void *b) {
    uint32_t x = (uint32_t)*(const uint64_t *)a, y = (uint32_t)*(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}
void port_gather(int64_t n, const port_photon *sorted, int64_t n_paths, int64_t nq, const float *queries, int k,
                 float max_dist2, float *rgb, float *r2_out, int32_t *found, int32_t *idx_out) {
    uint64_t *keys = (uint64_t *)malloc((size_t)(n ? n : 1) * 8);
    const float npf = (float)n_paths;
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        v3 nf = V(Q[3], Q[4], Q[5]), kd = V(Q[6], Q[7], Q[8]);
        int64_t m = 0;
        for (int64_t i = 0; i < n; ++i) {
            float dx = sorted[i].p[0] - Q[0], dy = sorted[i].p[1] - Q[1], dz = sorted[i].p[2] - Q[2];
            float d2 = dx * dx + dy * dy + dz * dz;
            if (d2 < max_dist2) { uint32_t b; memcpy(&b, &d2, 4); keys[m++] = ((uint64_t)b << 32) | (uint32_t)i; }
        }
        v3 acc = V(0, 0, 0);
        float md2 = max_dist2;
        int64_t cnt = m;
        if (k > 0) {
            qsort(keys, (size_t)m, 8, cmp_u64);
            if (cnt > k) cnt = k;
            if (cnt == k) { uint32_t b = (uint32_t)(keys[k - 1] >> 32); memcpy(&md2, &b, 4); }
            if (idx_out) for (int j = 0; j < k; ++j) idx_out[q * k + j] = j < cnt ? (int32_t)(uint32_t)keys[j] : -1;
            /* summation order of the device kernels: the selected set in ascending photon index (= the order the
             * grid cells store them, so a kernel that selects by threshold needs no sort), 32 strided partial sums,
             * then an xor butterfly */
            qsort(keys, (size_t)cnt, 8, cmp_idx);
            v3 part[32];
            for (int l = 0; l < 32; ++l) {
                part[l] = V(0, 0, 0);
                for (int64_t j = l; j < cnt; j += 32) {
                    uint32_t i = (uint32_t)keys[j], b = (uint32_t)(keys[j] >> 32);
                    float d2; memcpy(&d2, &b, 4);
                    v3 wi = oct_decode(sorted[i].wi_oct);
                    if (vdot(nf, wi) > 0.f) {
                        float sk = 1.f - d2 / md2;
                        float w = 3.f * O_INV_PI * sk * sk / (npf * md2);
                        part[l] = vadd(part[l], V(w * sorted[i].alpha[0], w * sorted[i].alpha[1], w * sorted[i].alpha[2]));
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1) {
                v3 nx[32];
                for (int l = 0; l < 32; ++l) nx[l] = vadd(part[l], part[l ^ o]);
                memcpy(part, nx, sizeof(part));
            }
            acc = part[0];
        } else {
            double ax = 0, ay = 0, az = 0;    /* fixed radius: order-free reference sum in double */
            for (int64_t j = 0; j < m; ++j) {
                uint32_t i = (uint32_t)keys[j], b = (uint32_t)(keys[j] >> 32);
                float d2; memcpy(&d2, &b, 4);
                v3 wi = oct_decode(sorted[i].wi_oct);
                if (vdot(nf, wi) > 0.f) {
                    float sk = 1.f - d2 / md2;
                    float w = 3.f * O_INV_PI * sk * sk / (npf * md2);
                    ax += w * sorted[i].alpha[0]; ay += w * sorted[i].alpha[1]; az += w * sorted[i].alpha[2];
                }
            }
            acc = V((float)ax, (float)ay, (float)az);
        }
        if (rgb) { rgb[3 * q] = acc.x * kd.x * O_INV_PI; rgb[3 * q + 1] = acc.y * kd.y * O_INV_PI; rgb[3 * q + 2] = acc.z * kd.z * O_INV_PI; }
        if (r2_out) r2_out[q] = md2;
        if (found) found[q] = (int32_t)cnt;
    }
    free(keys);
}

/* final gathering (finalgather true): direct / radiance photons, EPhoton, radiance-photon lookup, the gather */
#include "oracle_fg.h"
#include "oracle_render.h"
