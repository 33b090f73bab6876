// Device-side building blocks shared by trace.cu, photon.cu, gather.cu and render.cu (sm_100a only).
//
// Every float expression in this file is written as explicit single operations and the library is compiled
// with -fmad=false, so each result is the correctly rounded IEEE binary32 value of exactly the expression
// spelled here; oracle/oracle_port.c restates the same expressions (gcc -ffp-contract=off), which is what makes
// the GPU results comparable bit for bit.  Transcendentals are avoided: the only sin/cos used on the hot path is
// b2_sincos_quarter() below, a fixed polynomial evaluated with + and * only.
#pragma once
#include "b2rt_internal.cuh"
#include <math.h>

#define B2_PI 3.14159265358979323846f
#define B2_INV_PI 0.31830988618379067154f
#define B2_INV_TWOPI 0.15915494309189533577f

struct f3 { float x, y, z; };
__host__ __device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
// pbrt Dot(): (x*x' + y*y') + z*z'   (core/geometry.h)
__host__ __device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ bool is_black(f3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }
// pbrt RGBSpectrum::y()  (core/spectrum.h:425-426)
__host__ __device__ __forceinline__ float lum(f3 c) { return 0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z; }

// pbrt Cross(): products and difference in double, rounded once to float (core/geometry.h:477-484).
// a*b is exact in double, so fma(a,b,-(c*d)) == (a*b)-(c*d) evaluated in double.
__device__ __forceinline__ float cross_comp(double a, double b, double c, double d) {
    return (float)__fma_rn(a, b, -__dmul_rn(c, d));
}
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
    const double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return mk3(cross_comp(ay, bz, az, by), cross_comp(az, bx, ax, bz), cross_comp(ax, by, ay, bx));
}
// pbrt Normalize(): v / Length(v), Length = sqrtf(x*x+y*y+z*z); operator/ multiplies by 1.f/f
__device__ __forceinline__ f3 normalize3(f3 v) {
    float inv = 1.f / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    return mk3(v.x * inv, v.y * inv, v.z * inv);
}
// pbrt CoordinateSystem (core/geometry.h:508-518)
__device__ __forceinline__ void coordinate_system(f3 v1, f3 &v2, f3 &v3) {
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        v2 = mk3(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        v2 = mk3(0.f, v1.z * invLen, -v1.y * invLen);
    }
    v3 = cross3(v1, v2);
}

// ---- counter-based RNG: Philox-4x32-10 (Salmon et al. 2011) ---------------------------------------------------
// key = (seed_lo, seed_hi ^ stream), counter = (index_lo, index_hi, bounce, block).  Replaces the reference's
// cuRAND MTGP32 buffer (cuda_render/util/random/cudarandom.cpp:15-28) and its Halton table
// (photon_mapping/photontracing.cu:19-43): no state, no buffers, reproducible per photon for any GPU count.
struct u4 { uint32_t x, y, z, w; };
__host__ __device__ __forceinline__ u4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                     uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    u4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}
// 24 random bits -> [0,1): exact in binary32, never 1
__host__ __device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }

// ---- deterministic sin/cos ---------------------------------------------------------------------------------------
// angle = q quarter turns, q in [0,4].  k = nearest integer, f = q-k in [-1/2,1/2] (exact), x = f*(pi/2);
// Cephes single-precision minimax polynomials on [-pi/4,pi/4], Horner with separate * and +.
__host__ __device__ __forceinline__ void b2_sincos_quarter(float q, float &s, float &c) {
    float kf = floorf(q + 0.5f);
    int k = (int)kf;
    float f = q - kf;
    float x = f * 1.57079632679489661923f;
    float z = x * x;
    float ps = ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z + -1.6666654611e-1f) * z * x + x;
    float pc = ((2.443315711809948e-5f * z + -1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z +
               (1.f - 0.5f * z);
    switch (k & 3) {
        case 0: s = ps; c = pc; break;
        case 1: s = pc; c = -ps; break;
        case 2: s = -ps; c = -pc; break;
        default: s = -pc; c = ps; break;
    }
}

// pbrt ConcentricSampleDisk (core/montecarlo.cpp:306-348) with cosf/sinf(theta*pi/4) -> b2_sincos_quarter(theta/2)
__host__ __device__ __forceinline__ void concentric_sample_disk(float u1, float u2, float &dx, float &dy) {
    float r, theta;
    float sx = 2.f * u1 - 1.f, sy = 2.f * u2 - 1.f;
    if (sx == 0.f && sy == 0.f) { dx = 0.f; dy = 0.f; return; }
    if (sx >= -sy) {
        if (sx > sy) { r = sx; theta = (sy > 0.f) ? sy / r : 8.f + sy / r; }
        else { r = sy; theta = 2.f - sx / r; }
    } else {
        if (sx <= sy) { r = -sx; theta = 4.f - sy / r; }
        else { r = -sy; theta = 6.f + sx / r; }
    }
    float s, c;
    b2_sincos_quarter(theta * 0.5f, s, c);
    dx = r * c; dy = r * s;
}
// pbrt CosineSampleHemisphere (core/montecarlo.h:128-133)
__host__ __device__ __forceinline__ f3 cosine_sample_hemisphere(float u1, float u2) {
    float x, y;
    concentric_sample_disk(u1, u2, x, y);
    return mk3(x, y, sqrtf(fmaxf(0.f, 1.f - x * x - y * y)));
}
// pbrt UniformSampleSphere (core/montecarlo.cpp:283-290), phi = 2*pi*u2 = 4*u2 quarter turns
__host__ __device__ __forceinline__ f3 uniform_sample_sphere(float u1, float u2) {
    float z = 1.f - 2.f * u1;
    float r = sqrtf(fmaxf(0.f, 1.f - z * z));
    float s, c;
    b2_sincos_quarter(4.f * u2, s, c);
    return mk3(r * c, r * s, z);
}

struct RayD {           // per-ray constants
    float ox, oy, oz, dx, dy, dz, tmin;
    double ddx, ddy, ddz;   // direction widened once (pbrt's Cross works in double)
};

// slot: leaf-record slot of the hit primitive (in the object's records for an instanced hit); inst: instance index or -1
struct HitD { float t; int32_t prim; float b1, b2; int32_t slot; int32_t inst; };

// returns true and writes t,b1,b2 if pbrt's Triangle::Intersect (shapes/trianglemesh.cpp:146-179) would accept
// on [tmin, tmax]
__device__ __forceinline__ bool tri_test(const RayD &r, const float4 &A, const float4 &B, const float4 &C,
                                         float tmax, float &t_out, float &b1_out, float &b2_out) {
    const double e1x = B.x, e1y = B.y, e1z = B.z, e2x = C.x, e2y = C.y, e2z = C.z;
    // s1 = Cross(ray.d, e2)
    float s1x = cross_comp(r.ddy, e2z, r.ddz, e2y);
    float s1y = cross_comp(r.ddz, e2x, r.ddx, e2z);
    float s1z = cross_comp(r.ddx, e2y, r.ddy, e2x);
    float divisor = s1x * B.x + s1y * B.y + s1z * B.z;
    if (divisor == 0.f) return false;
    float inv = 1.f / divisor;
    float dvx = r.ox - A.x, dvy = r.oy - A.y, dvz = r.oz - A.z;
    float b1 = (dvx * s1x + dvy * s1y + dvz * s1z) * inv;
    if (b1 < 0.f || b1 > 1.f) return false;
    // s2 = Cross(d, e1)
    const double ddvx = dvx, ddvy = dvy, ddvz = dvz;
    float s2x = cross_comp(ddvy, e1z, ddvz, e1y);
    float s2y = cross_comp(ddvz, e1x, ddvx, e1z);
    float s2z = cross_comp(ddvx, e1y, ddvy, e1x);
    float b2 = (r.dx * s2x + r.dy * s2y + r.dz * s2z) * inv;
    if (b2 < 0.f || b1 + b2 > 1.f) return false;
    float t = (C.x * s2x + C.y * s2y + C.z * s2z) * inv;
    if (t < r.tmin || t > tmax) return false;
    t_out = t; b1_out = b1; b2_out = b2;
    return true;
}

__device__ __forceinline__ void xf_point(const float *m, float x, float y, float z, float &ox, float &oy,
                                         float &oz) {
    ox = m[0] * x + m[1] * y + m[2] * z + m[3];
    oy = m[4] * x + m[5] * y + m[6] * z + m[7];
    oz = m[8] * x + m[9] * y + m[10] * z + m[11];
}
__device__ __forceinline__ void xf_vec(const float *m, float x, float y, float z, float &ox, float &oy,
                                       float &oz) {
    ox = m[0] * x + m[1] * y + m[2] * z;
    oy = m[4] * x + m[5] * y + m[6] * z;
    oz = m[8] * x + m[9] * y + m[10] * z;
}
// pbrt Transform::operator()(Normal): transpose of the inverse (core/transform.h), n' = w2o^T n
__device__ __forceinline__ f3 xf_normal(const float *w2o, f3 n) {
    return mk3(w2o[0] * n.x + w2o[4] * n.y + w2o[8] * n.z, w2o[1] * n.x + w2o[5] * n.y + w2o[9] * n.z,
               w2o[2] * n.x + w2o[6] * n.y + w2o[10] * n.z);
}

// phi > phimax of pbrt's partial quadrics (phi = atan2f(y, x) lifted to [0, 2 pi), shapes/sphere.cpp:88-90, disk.cpp:73-75)
// WITHOUT atan2f, which is not reproducible between math libraries: the side of the direction (cos phimax, sin phimax)
// the point lies on, from the deterministic sin / cos polynomial above.  Equal to pbrt's test except for points within
// rounding of the cut.
__device__ __forceinline__ bool phi_clipped(float x, float y, float phimax) {
    float s, c;
    b2_sincos_quarter(phimax * 0.63661977236758134308f, s, c);       // radians -> quarter turns
    const float cr = c * y - s * x;                                  // |r| sin(phi - phimax)
    return phimax <= B2_PI ? !(y >= 0.f && cr <= 0.f) : (y < 0.f && cr > 0.f);
}

// pbrt Sphere::Intersect arithmetic (shapes/sphere.cpp:58-106) with zmin / zmax clipping; the reference's device program
// ignores zmin / zmax / phiMax (cuda_render/util/shape/cudasphere.cu:27-72)
static __device__ __noinline__ bool sphere_test(const RayD &r, const AnalyticPrim &S, float tmax, float &t_out) {
    float ox, oy, oz, dx, dy, dz;
    xf_point(S.w2o, r.ox, r.oy, r.oz, ox, oy, oz);
    xf_vec(S.w2o, r.dx, r.dy, r.dz, dx, dy, dz);
    float A = dx * dx + dy * dy + dz * dz;
    float B = 2 * (dx * ox + dy * oy + dz * oz);
    float C = ox * ox + oy * oy + oz * oz - S.radius * S.radius;
    float discrim = B * B - 4.f * A * C;
    if (discrim < 0.f) return false;
    float root = sqrtf(discrim);
    float q = (B < 0) ? -.5f * (B - root) : -.5f * (B + root);
    float t0 = q / A, t1 = C / q;
    if (t0 > t1) { float s = t0; t0 = t1; t1 = s; }
    if (t0 > tmax || t1 < r.tmin) return false;
    float thit = t0;
    if (t0 < r.tmin) {
        thit = t1;
        if (thit > tmax) return false;
    }
    // partial sphere: pbrt's zmin / zmax / phimax clipping (shapes/sphere.cpp:84-106), which the reference's device
    // program ignores.  For a sphere the record's height / inner fields hold zmin / zmax (already clamped to [-radius,
    // radius] as pbrt's constructor does; zmin >= zmax = no z clipping) and pad holds phimax in radians (0 = full turn).
    const float zmin = S.height, zmax = S.inner, phimax = S.pad;
    if (zmin < zmax || phimax > 0.f) {
        auto clipped = [&](float t) {
            float px = ox + dx * t;
            const float py = oy + dy * t, pz = oz + dz * t;
            if (px == 0.f && py == 0.f) px = 1e-5f * S.radius;
            bool c = (zmin < zmax) && ((zmin > -S.radius && pz < zmin) || (zmax < S.radius && pz > zmax));
            if (phimax > 0.f) c = c || phi_clipped(px, py, phimax);
            return c;
        };
        if (clipped(thit)) {
            if (thit == t1) return false;
            if (t1 > tmax) return false;
            thit = t1;
            if (clipped(thit)) return false;
        }
    }
    t_out = thit;
    return true;
}

// pbrt Disk::Intersect (shapes/disk.cpp:56-76)
static __device__ __noinline__ bool disk_test(const RayD &r, const AnalyticPrim &D, float tmax, float &t_out) {
    float ox, oy, oz, dx, dy, dz;
    xf_point(D.w2o, r.ox, r.oy, r.oz, ox, oy, oz);
    xf_vec(D.w2o, r.dx, r.dy, r.dz, dx, dy, dz);
    if ((double)fabsf(dz) < 1e-7) return false;
    float thit = (D.height - oz) / dz;
    if (thit < r.tmin || thit > tmax) return false;
    float px = ox + dx * thit, py = oy + dy * thit;
    float dist2 = px * px + py * py;
    if (dist2 > D.radius * D.radius || dist2 < D.inner * D.inner) return false;
    if (D.pad > 0.f && phi_clipped(px, py, D.pad)) return false;    // pbrt's phimax (shapes/disk.cpp:73-76); pad = radians, 0 = full
    t_out = thit;
    return true;
}

constexpr int B2RT_STACK_SIZE = 64;

// Reciprocal direction for the slab test, clamped to +-FLT_MAX.  With 1/0 = inf a ray that runs exactly in the plane of a
// box face gives (plane - origin) * inf = 0 * inf = NaN, fminf(NaN, +inf) = +inf, and the box - which the ray grazes -
// would be rejected: the node test must only ever err on the accepting side (found by
// tests/test_gpu_parity.py::test_axis_parallel_and_degenerate_rays_match_oracle: rays straight down through mesh
// vertices lost their hits).  With a finite factor the product is 0 and the slab is [0, big].
__device__ __forceinline__ float slab_rcp(float d) {
    return fminf(fmaxf(1.f / d, -3.402823466e38f), 3.402823466e38f);
}

// Same operations and roundings as tri_test, evaluated without early exits: one predicate at the end.  Every
// comparison is false on NaN, so a zero divisor (inv = inf, products NaN or inf) is rejected exactly as in tri_test.
__device__ __forceinline__ bool tri_test_nobranch(const RayD &r, const float4 &A, const float4 &B, const float4 &C,
                                                  float tmax, float &t_out, float &b1_out, float &b2_out) {
    const double e1x = B.x, e1y = B.y, e1z = B.z, e2x = C.x, e2y = C.y, e2z = C.z;
    const float s1x = cross_comp(r.ddy, e2z, r.ddz, e2y);
    const float s1y = cross_comp(r.ddz, e2x, r.ddx, e2z);
    const float s1z = cross_comp(r.ddx, e2y, r.ddy, e2x);
    const float divisor = s1x * B.x + s1y * B.y + s1z * B.z;
    const float inv = 1.f / divisor;
    const float dvx = r.ox - A.x, dvy = r.oy - A.y, dvz = r.oz - A.z;
    const float b1 = (dvx * s1x + dvy * s1y + dvz * s1z) * inv;
    const double ddvx = dvx, ddvy = dvy, ddvz = dvz;
    const float s2x = cross_comp(ddvy, e1z, ddvz, e1y);
    const float s2y = cross_comp(ddvz, e1x, ddvx, e1z);
    const float s2z = cross_comp(ddvx, e1y, ddvy, e1x);
    const float b2 = (r.dx * s2x + r.dy * s2y + r.dz * s2z) * inv;
    const float t = (C.x * s2x + C.y * s2y + C.z * s2z) * inv;
    const bool ok = (divisor != 0.f) & (b1 >= 0.f) & (b1 <= 1.f) & (b2 >= 0.f) & (b1 + b2 <= 1.f) & (t >= r.tmin) & (t <= tmax);
    t_out = t; b1_out = b1; b2_out = b2;
    return ok;
}

// ---- object instances: the second level of the two-level BVH ---------------------------------------------------------
// Replaces the OptiX Transform node the reference hangs every ObjectInstance under (cuda_render/cudarender.cpp:88-103) and
// follows pbrt's TransformedPrimitive::Intersect (core/primitive.cpp:95-103): the ray is taken into the instance's space
// by Transform::operator()(Ray) -- origin as a point, direction as a vector, NOT renormalised, so t is the same parameter
// in both spaces and [tmin, tmax] carry over -- and the object's own tree is searched there with the very same primitive
// arithmetic.  t therefore equals what pbrt computes for the instanced shape bit for bit.
__device__ __forceinline__ const InstanceRec *instance_of(const float4 &A) {
    return reinterpret_cast<const InstanceRec *>(((uint64_t)__float_as_uint(A.y) << 32) | (uint64_t)__float_as_uint(A.x));
}
__device__ __forceinline__ void instance_ray(const InstanceRec *__restrict__ I, float ox, float oy, float oz, float dx, float dy,
                                             float dz, float tmin, RayD &r) {
    float m[12];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 row = __ldg(reinterpret_cast<const float4 *>(I->w2o) + k);
        m[4 * k] = row.x; m[4 * k + 1] = row.y; m[4 * k + 2] = row.z; m[4 * k + 3] = row.w;
    }
    xf_point(m, ox, oy, oz, r.ox, r.oy, r.oz);
    xf_vec(m, dx, dy, dz, r.dx, r.dy, r.dz);
    r.tmin = tmin;
    r.ddx = r.dx; r.ddy = r.dy; r.ddz = r.dz;
}

// Closest (or any) accepted primitive of one instance on [r.tmin, tmax].  Plain if-if loop with a private stack: it is
// entered from divergent code (one lane of a warp may be here while its neighbours test triangles), so it does not
// cooperate across the warp.  Same conservative node test, same winner rule (min t, then min id) as the top level; the id
// returned is local to the object.
template <bool ANY>
static __device__ __noinline__ bool instance_test(const RayD &rw, const InstanceRec *__restrict__ I, float tmax, float &t_out,
                                                  float &b1_out, float &b2_out, int &prim_out, int &slot_out, bool &overflow) {
    RayD r;
    instance_ray(I, rw.ox, rw.oy, rw.oz, rw.dx, rw.dy, rw.dz, rw.tmin, r);
    const WideNode *__restrict__ nodes = I->nodes;
    const PrimRec *__restrict__ recs = I->recs;
    const AnalyticPrim *__restrict__ analytic = I->analytic;
    const float idx = slab_rcp(r.dx), idy = slab_rcp(r.dy), idz = slab_rcp(r.dz);
    float best_t = tmax, best_b1 = 0.f, best_b2 = 0.f;
    int best_prim = 0x7fffffff, best_slot = -1;
    bool found = false;
    int32_t stack[B2RT_STACK_SIZE];
    int sp = 0;
    int32_t cur = 0;
    for (;;) {
        if (cur >= 0) {
            const float4 *np = reinterpret_cast<const float4 *>(nodes + cur);
            const float4 lox = __ldg(np), hix = __ldg(np + 1), loy = __ldg(np + 2), hiy = __ldg(np + 3), loz = __ldg(np + 4),
                         hiz = __ldg(np + 5);
            const int4 ch = __ldg(reinterpret_cast<const int4 *>(np + 6));
            float tn[4]; int32_t cr[4];
            const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, hx[4] = {hix.x, hix.y, hix.z, hix.w};
            const float ly[4] = {loy.x, loy.y, loy.z, loy.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w};
            const float lz[4] = {loz.x, loz.y, loz.z, loz.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
            cr[0] = ch.x; cr[1] = ch.y; cr[2] = ch.z; cr[3] = ch.w;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float t0x = (lx[k] - r.ox) * idx, t1x = (hx[k] - r.ox) * idx;
                float t0y = (ly[k] - r.oy) * idy, t1y = (hy[k] - r.oy) * idy;
                float t0z = (lz[k] - r.oz) * idz, t1z = (hz[k] - r.oz) * idz;
                float tnear = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), r.tmin));
                float tfar = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), best_t));
                float tn_p = __fmaf_rn(fabsf(tnear), -2e-6f, tnear);
                float tf_p = __fmaf_rn(fabsf(tfar), 2e-6f, tfar);
                bool h = (tn_p <= tf_p) && (cr[k] != B2RT_CHILD_EMPTY);
                tn[k] = h ? tnear : INFINITY;
            }
#define B2RT_CSWAP(a, b)                                                                  \
    if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; int32_t cc = cr[a]; cr[a] = cr[b]; cr[b] = cc; }
            B2RT_CSWAP(0, 1) B2RT_CSWAP(2, 3) B2RT_CSWAP(0, 2) B2RT_CSWAP(1, 3) B2RT_CSWAP(1, 2)
#undef B2RT_CSWAP
            const int nhit = (tn[0] < INFINITY) + (tn[1] < INFINITY) + (tn[2] < INFINITY) + (tn[3] < INFINITY);
            if (nhit == 0) {
                if (sp == 0) break;
                cur = stack[--sp];
                continue;
            }
            if (sp + 3 > B2RT_STACK_SIZE) { overflow = true; break; }
            if (nhit > 3) stack[sp++] = cr[3];
            if (nhit > 2) stack[sp++] = cr[2];
            if (nhit > 1) stack[sp++] = cr[1];
            cur = cr[0];
        } else {
            const uint32_t ref = (uint32_t)cur & 0x7fffffffu;
            const uint32_t first = ref >> 3, cnt = (ref & 7u) + 1u;
            for (uint32_t s = first; s < first + cnt; ++s) {
                const float4 *pp = reinterpret_cast<const float4 *>(recs + s);
                const float4 A = __ldg(pp), B = __ldg(pp + 1), C = __ldg(pp + 2);
                const int prim = __float_as_int(A.w);
                const int type = __float_as_int(B.w);
                float t, b1 = 0.f, b2 = 0.f;
                bool h;
                if (type == B2RT_PRIM_TRI) h = tri_test_nobranch(r, A, B, C, best_t, t, b1, b2);
                else if (type == B2RT_PRIM_SPHERE) h = sphere_test(r, analytic[__float_as_int(C.w)], best_t, t);
                else h = disk_test(r, analytic[__float_as_int(C.w)], best_t, t);
                if (h) {
                    if (ANY) { found = true; break; }
                    if (t < best_t || prim < best_prim) {
                        best_t = t; best_prim = prim; best_b1 = b1; best_b2 = b2; best_slot = (int)s; found = true;
                    }
                }
            }
            if ((ANY && found) || sp == 0) break;
            cur = stack[--sp];
        }
    }
    t_out = best_t; b1_out = best_b1; b2_out = best_b2; prim_out = best_prim; slot_out = best_slot;
    return found;
}

// Wide-BVH traversal of one ray, inlined into the photon and camera kernels.  Winner among all accepted primitives is
// (min t, then min primitive id): order-independent, so the answer does not depend on tree shape or visiting order.
// Node boxes are tested conservatively (padded slab test) so the tree never hides a primitive that the exhaustive loop
// would accept.  Speculative while-while loop: a lane that reaches a leaf parks it and keeps descending until every
// lane that is still in the node loop has one, then the parked leaves are tested together (the if-if form ran the
// primitive tests of these kernels with 2.5-6 of 32 threads per instruction).
// Returns true on a hit (ANY: any accepted primitive).  `overflow` is set if the stack would have overflowed.
// INST: the scene has object instances (instance records in the top-level tree).  A separate instantiation so that kernels
// of scenes without instances keep their registers: with the nested call compiled in, photon_trace_kernel went from 80 to 128
// registers and trace_spec_kernel started to spill (the Cfg2 step lost 19 %, run r06f).
template <bool ANY, bool INST>
__device__ __forceinline__ bool traverse(const WideNode *__restrict__ nodes, const PrimRec *__restrict__ recs,
                                         const AnalyticPrim *__restrict__ analytic, float ox, float oy, float oz,
                                         float dx, float dy, float dz, float tmin, float tmax, HitD &hit,
                                         bool &overflow) {
    RayD r;
    r.ox = ox; r.oy = oy; r.oz = oz; r.tmin = tmin;
    r.dx = dx; r.dy = dy; r.dz = dz;
    r.ddx = dx; r.ddy = dy; r.ddz = dz;
    const float idx = slab_rcp(dx), idy = slab_rcp(dy), idz = slab_rcp(dz);
    constexpr int32_t NONE = B2RT_CHILD_EMPTY;      // "nothing": never a valid node index or leaf reference

    float best_t = tmax;
    int best_prim = 0x7fffffff, best_slot = -1, best_inst = -1;
    float best_b1 = 0.f, best_b2 = 0.f;
    bool found = false;

    int32_t stack[B2RT_STACK_SIZE];
    int sp = 0;
    int32_t cur = 0, leaf = NONE;  // root

    while (cur != NONE || leaf != NONE) {
        // ---- node loop -----------------------------------------------------------------------------
        while (cur >= 0 && cur != NONE) {
            const float4 *np = reinterpret_cast<const float4 *>(nodes + cur);
            float4 lox = __ldg(np), hix = __ldg(np + 1), loy = __ldg(np + 2), hiy = __ldg(np + 3),
                   loz = __ldg(np + 4), hiz = __ldg(np + 5);
            int4 ch = __ldg(reinterpret_cast<const int4 *>(np + 6));
            float tn[4]; int32_t cr[4];
            const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, hx[4] = {hix.x, hix.y, hix.z, hix.w};
            const float ly[4] = {loy.x, loy.y, loy.z, loy.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w};
            const float lz[4] = {loz.x, loz.y, loz.z, loz.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
            cr[0] = ch.x; cr[1] = ch.y; cr[2] = ch.z; cr[3] = ch.w;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float t0x = (lx[k] - r.ox) * idx, t1x = (hx[k] - r.ox) * idx;
                float t0y = (ly[k] - r.oy) * idy, t1y = (hy[k] - r.oy) * idy;
                float t0z = (lz[k] - r.oz) * idz, t1z = (hz[k] - r.oz) * idz;
                float tnear = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), r.tmin));
                float tfar = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), best_t));
                // conservative padding: widen the slab interval by a few ulps on both ends
                float tn_p = __fmaf_rn(fabsf(tnear), -2e-6f, tnear);
                float tf_p = __fmaf_rn(fabsf(tfar), 2e-6f, tfar);
                bool h = (tn_p <= tf_p) && (cr[k] != B2RT_CHILD_EMPTY);
                tn[k] = h ? tnear : INFINITY;
            }
            // sort the 4 (tn, child) pairs ascending by tn with a 5-comparator network
#define B2RT_CSWAP(a, b)                                                                  \
    if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; int32_t cc = cr[a]; cr[a] = cr[b]; cr[b] = cc; }
            if (!ANY) { B2RT_CSWAP(0, 1) B2RT_CSWAP(2, 3) B2RT_CSWAP(0, 2) B2RT_CSWAP(1, 3) B2RT_CSWAP(1, 2) }
#undef B2RT_CSWAP
            int nhit = (tn[0] < INFINITY) + (tn[1] < INFINITY) + (tn[2] < INFINITY) + (tn[3] < INFINITY);
            if (ANY) {                          // any hit: the answer does not depend on the order, skip the network
                if (sp + 3 > B2RT_STACK_SIZE) { overflow = true; sp = 0; cur = NONE; leaf = NONE; break; }
                int32_t nxt = NONE;
#pragma unroll
                for (int k = 3; k >= 0; --k)
                    if (tn[k] < INFINITY) { if (nxt != NONE) stack[sp++] = nxt; nxt = cr[k]; }
                cur = nxt != NONE ? nxt : (sp > 0 ? stack[--sp] : NONE);
            } else if (nhit == 0) {
                cur = sp > 0 ? stack[--sp] : NONE;
            } else {
                if (sp + 3 > B2RT_STACK_SIZE) { overflow = true; sp = 0; cur = NONE; leaf = NONE; break; }
                if (nhit > 3) stack[sp++] = cr[3];
                if (nhit > 2) stack[sp++] = cr[2];
                if (nhit > 1) stack[sp++] = cr[1];
                cur = cr[0];
            }
            if (cur < 0 && leaf == NONE) {      // park this leaf, continue with the next entry
                leaf = cur;
                cur = sp > 0 ? stack[--sp] : NONE;
            }
            if (!__any_sync(__activemask(), leaf == NONE)) break;
        }
        // ---- primitive loop: the parked leaf, then `cur` while it is a leaf as well ---------------------
        while (leaf != NONE) {
            uint32_t ref = (uint32_t)leaf & 0x7fffffffu;
            uint32_t first = ref >> 3, cnt = (ref & 7u) + 1u;
            for (uint32_t s = first; s < first + cnt; ++s) {
                const float4 *pp = reinterpret_cast<const float4 *>(recs + s);
                float4 A = __ldg(pp), B = __ldg(pp + 1), C = __ldg(pp + 2);
                int prim = __float_as_int(A.w);
                int type = __float_as_int(B.w);
                float t, b1 = 0.f, b2 = 0.f;
                bool h;
                int hslot = (int)s, hinst = -1;
                if (type == B2RT_PRIM_TRI) h = tri_test_nobranch(r, A, B, C, best_t, t, b1, b2);
                else if (type == B2RT_PRIM_SPHERE) h = sphere_test(r, analytic[__float_as_int(C.w)], best_t, t);
                else if (!INST || type == B2RT_PRIM_DISK) h = disk_test(r, analytic[__float_as_int(C.w)], best_t, t);
                else {                                      // object instance: search its tree in its own space
                    int lprim;
                    h = instance_test<ANY>(r, instance_of(A), best_t, t, b1, b2, lprim, hslot, overflow);
                    prim += lprim;                          // A.w of an instance record = its first global id
                    hinst = __float_as_int(C.w);
                }
                if (h) {
                    if (ANY) { found = true; break; }
                    if (t < best_t || prim < best_prim) {   // t <= best_t is guaranteed by the test
                        best_t = t; best_prim = prim; best_b1 = b1; best_b2 = b2; best_slot = hslot; best_inst = hinst; found = true;
                    }
                }
            }
            if (ANY && found) { cur = NONE; leaf = NONE; break; }
            leaf = NONE;
            if (cur < 0) {                      // the next entry is a leaf too
                leaf = cur;
                cur = sp > 0 ? stack[--sp] : NONE;
            }
        }
    }
    hit.t = found ? best_t : INFINITY;
    hit.prim = found ? best_prim : -1;
    hit.b1 = best_b1; hit.b2 = best_b2; hit.slot = best_slot; hit.inst = best_inst;
    return found;
}
