// K5: wide-BVH traversal for closest-hit and any-hit rays.
//
// Replaces OptiX's closed "Bvh" traversal behind every rtTrace in the reference
// (cuda_render/simple_render/simplerender.cu:27,68; photon_mapping/raytracing.cu:24,76,102;
// photontracing.cu:102,131,184) and the per-candidate intersection programs
// (util/shape/cudatrianglemesh.cu:12-81, cudasphere.cu:27-72, cudadisk.cu:18-50).
//
// Parity contract: the primitive tests reproduce the arithmetic of the reference's CPU path
// bit for bit -- Triangle::Intersect (pbrt-v2/src/shapes/trianglemesh.cpp:146-179), including
// pbrt's Cross() which forms its products in double (core/geometry.h:477-484), Sphere::Intersect
// (shapes/sphere.cpp:58-83) and Disk::Intersect (shapes/disk.cpp:56-72).  This file is compiled
// with -fmad=false so no float expression is contracted.  Winner among all accepted primitives is
// (min t, then min primitive id): order-independent, so the answer does not depend on tree shape.
// Node boxes are tested conservatively (padded slab test) so the tree never hides a primitive that
// the exhaustive loop (renderers/aggregatetest.cpp:93-102) would accept.
#include "b2rt_internal.cuh"
#include <math.h>

namespace {

constexpr int TRACE_THREADS = 128;
constexpr int STACK_SIZE = 96;

struct RayD {           // per-ray constants
    float ox, oy, oz, dx, dy, dz, tmin;
    double ddx, ddy, ddz;   // direction widened once (pbrt's Cross works in double)
};

// pbrt Cross(): double products, double subtract, round to float.  a*b is exact in double, so
// fma(a,b,-(c*d)) rounds exactly like (a*b)-(c*d) evaluated in double.
__device__ __forceinline__ float cross_comp(double a, double b, double c, double d) {
    return (float)__fma_rn(a, b, -__dmul_rn(c, d));
}

// returns true and writes t,b1,b2 if pbrt's Triangle::Intersect would accept on [tmin, tmax]
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

// full sphere (the reference ignores zmin/zmax/phiMax too, cudasphere.cu:27-72)
__device__ bool sphere_test(const RayD &r, const AnalyticPrim &S, float tmax, float &t_out) {
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
    t_out = thit;
    return true;
}

__device__ bool disk_test(const RayD &r, const AnalyticPrim &D, float tmax, float &t_out) {
    float ox, oy, oz, dx, dy, dz;
    xf_point(D.w2o, r.ox, r.oy, r.oz, ox, oy, oz);
    xf_vec(D.w2o, r.dx, r.dy, r.dz, dx, dy, dz);
    if ((double)fabsf(dz) < 1e-7) return false;
    float thit = (D.height - oz) / dz;
    if (thit < r.tmin || thit > tmax) return false;
    float px = ox + dx * thit, py = oy + dy * thit;
    float dist2 = px * px + py * py;
    if (dist2 > D.radius * D.radius || dist2 < D.inner * D.inner) return false;
    t_out = thit;
    return true;
}

template <bool ANY>
__global__ void __launch_bounds__(TRACE_THREADS)
trace_kernel(const WideNode *__restrict__ nodes, const PrimRec *__restrict__ recs,
             const AnalyticPrim *__restrict__ analytic, const b2rt_ray *__restrict__ rays,
             b2rt_hit *__restrict__ hits, uint8_t *__restrict__ occ, int64_t nrays, uint32_t *status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrays) return;
    const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
    float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    RayD r;
    r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.tmin = r0.w;
    r.dx = r1.x; r.dy = r1.y; r.dz = r1.z;
    r.ddx = r1.x; r.ddy = r1.y; r.ddz = r1.z;
    const float idx = 1.f / r.dx, idy = 1.f / r.dy, idz = 1.f / r.dz;

    float best_t = r1.w;
    int best_prim = 0x7fffffff;
    float best_b1 = 0.f, best_b2 = 0.f;
    bool found = false;

    int32_t stack[STACK_SIZE];
    int sp = 0;
    int32_t cur = 0;  // root
    bool overflow = false;

    while (true) {
        if (cur >= 0) {
            // ---- internal node: test 4 child boxes ------------------------------------------------
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
                float tn_p = tnear - fabsf(tnear) * 2e-6f;
                float tf_p = tfar + fabsf(tfar) * 2e-6f;
                bool hit = (tn_p <= tf_p) && (cr[k] != B2RT_CHILD_EMPTY);
                tn[k] = hit ? tnear : INFINITY;
            }
            // sort the 4 (tn, child) pairs ascending by tn with a 5-comparator network
#define B2RT_CSWAP(a, b)                                                                  \
    if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; int32_t cc = cr[a]; cr[a] = cr[b]; cr[b] = cc; }
            B2RT_CSWAP(0, 1) B2RT_CSWAP(2, 3) B2RT_CSWAP(0, 2) B2RT_CSWAP(1, 3) B2RT_CSWAP(1, 2)
#undef B2RT_CSWAP
            // push far-to-near, continue with the nearest
            int nhit = (tn[0] < INFINITY) + (tn[1] < INFINITY) + (tn[2] < INFINITY) + (tn[3] < INFINITY);
            if (nhit == 0) {
                if (sp == 0) break;
                cur = stack[--sp];
                continue;
            }
            if (sp + 3 > STACK_SIZE) { overflow = true; break; }
            if (nhit > 3) stack[sp++] = cr[3];
            if (nhit > 2) stack[sp++] = cr[2];
            if (nhit > 1) stack[sp++] = cr[1];
            cur = cr[0];
            // leaf refs are negative, but B2RT_CHILD_EMPTY is positive and never pushed
        } else {
            // ---- leaf: 1..8 primitive records -----------------------------------------------------
            uint32_t ref = (uint32_t)cur & 0x7fffffffu;
            uint32_t first = ref >> 3, cnt = (ref & 7u) + 1u;
            for (uint32_t s = first; s < first + cnt; ++s) {
                const float4 *pp = reinterpret_cast<const float4 *>(recs + s);
                float4 A = __ldg(pp), B = __ldg(pp + 1), C = __ldg(pp + 2);
                int prim = __float_as_int(A.w);
                int type = __float_as_int(B.w);
                float t, b1 = 0.f, b2 = 0.f;
                bool h;
                if (type == B2RT_PRIM_TRI) h = tri_test(r, A, B, C, best_t, t, b1, b2);
                else if (type == B2RT_PRIM_SPHERE) h = sphere_test(r, analytic[__float_as_int(C.w)], best_t, t);
                else h = disk_test(r, analytic[__float_as_int(C.w)], best_t, t);
                if (h) {
                    if (ANY) { found = true; break; }
                    if (t < best_t || prim < best_prim) {   // t <= best_t is guaranteed by the test
                        best_t = t; best_prim = prim; best_b1 = b1; best_b2 = b2; found = true;
                    }
                }
            }
            if (ANY && found) break;
            if (sp == 0) break;
            cur = stack[--sp];
        }
    }
    if (overflow) atomicAdd(status, 1u);
    if (ANY) {
        occ[i] = found ? 1 : 0;
    } else {
        float4 h;
        h.x = found ? best_t : INFINITY;
        h.y = __int_as_float(found ? best_prim : -1);
        h.z = best_b1; h.w = best_b2;
        *reinterpret_cast<float4 *>(hits + i) = h;
    }
}

}  // namespace

int b2rt_trace_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_hit *d_hits, uint8_t *d_occ,
                    bool any_hit) {
    if (n <= 0) return B2RT_OK;
    const WideNode *nodes = ctx->d_wide_nodes.as<WideNode>();
    const PrimRec *recs = ctx->d_prim_recs.as<PrimRec>();
    const AnalyticPrim *an = ctx->d_analytic.as<AnalyticPrim>();
    uint32_t *status = ctx->d_counters.as<uint32_t>() + 8;
    int64_t grid = (n + TRACE_THREADS - 1) / TRACE_THREADS;
    if (any_hit)
        trace_kernel<true><<<(unsigned)grid, TRACE_THREADS, 0, ctx->stream>>>(nodes, recs, an, d_rays, nullptr,
                                                                              d_occ, n, status);
    else
        trace_kernel<false><<<(unsigned)grid, TRACE_THREADS, 0, ctx->stream>>>(nodes, recs, an, d_rays, d_hits,
                                                                               nullptr, n, status);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}
