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
#include "b2rt_device.cuh"

namespace {

constexpr int TRACE_THREADS = 128;

template <bool ANY>
__global__ void __launch_bounds__(TRACE_THREADS)
trace_kernel(const WideNode *__restrict__ nodes, const PrimRec *__restrict__ recs,
             const AnalyticPrim *__restrict__ analytic, const b2rt_ray *__restrict__ rays,
             b2rt_hit *__restrict__ hits, uint8_t *__restrict__ occ, int64_t nrays, uint32_t *status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrays) return;
    const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
    float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    HitD h;
    bool overflow = false;
    bool found = traverse<ANY>(nodes, recs, analytic, r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r0.w, r1.w, h, overflow);
    if (overflow) atomicAdd(status, 1u);
    if (ANY) {
        occ[i] = found ? 1 : 0;
    } else {
        float4 o;
        o.x = h.t; o.y = __int_as_float(h.prim); o.z = h.b1; o.w = h.b2;
        *reinterpret_cast<float4 *>(hits + i) = o;
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
