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
#include <type_traits>

namespace {

constexpr int TRACE_THREADS = 128;

template <bool ANY, bool INST>
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
    bool found = traverse<ANY, INST>(nodes, recs, analytic, r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r0.w, r1.w, h, overflow);
    if (overflow) atomicAdd(status, 1u);
    if (ANY) {
        occ[i] = found ? 1 : 0;
    } else {
        float4 o;
        o.x = h.t; o.y = __int_as_float(h.prim); o.z = h.b1; o.w = h.b2;
        *reinterpret_cast<float4 *>(hits + i) = o;
    }
}


// ---- persistent variant -----------------------------------------------------------------------------------------
// One resident grid (blocks = SMs x occupancy) pulls rays from a global pool instead of one thread per ray:
//   * a warp fetches 32 consecutive rays at a time (caller order, so a warp stays coherent for camera rays);
//   * when fewer than `refill` lanes of the warp are still traversing, the idle lanes are refilled from the pool
//     -- ranks from __ballot_sync + __popc, one atomicAdd per warp -- so long rays do not hold 31 idle lanes;
//   * the traversal stack lives in shared memory, one column per thread (bank = lane: conflict-free); only entries
//     deeper than PT_SSTACK spill to a local array;
//   * while-while loop: descend through internal nodes without leaving the node branch, then test the leaf.
// Results are identical to trace_kernel by construction: the winner is (min t, then min prim id), independent of
// the visiting order.
constexpr int PT_THREADS = 128;
#ifndef B2RT_TRACE_BLOCKS_CLOSEST
#define B2RT_TRACE_BLOCKS_CLOSEST 7      // resident blocks per SM the speculative kernel is compiled for (72 / 64 registers);
                                         // re-checked r08e: 6 / 7 -> 6711, 8 / 9 -> 6511, 9 / 10 -> 6177 against 6843 Mrays/s
#endif
#ifndef B2RT_TRACE_BLOCKS_ANY
#define B2RT_TRACE_BLOCKS_ANY 8
#endif
constexpr int PT_SSTACK = 12;

template <bool ANY, bool INST>
__global__ void __launch_bounds__(PT_THREADS)
trace_persistent_kernel(const WideNode *__restrict__ nodes, const PrimRec *__restrict__ recs,
                        const AnalyticPrim *__restrict__ analytic, const b2rt_ray *__restrict__ rays,
                        b2rt_hit *__restrict__ hits, uint8_t *__restrict__ occ, int64_t nrays,
                        unsigned long long *pool, int refill, int chunk, uint32_t *status) {
    __shared__ int32_t s_stack[PT_SSTACK][PT_THREADS];
    int32_t l_stack[B2RT_STACK_SIZE - PT_SSTACK];
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt = (1u << lane) - 1u;

    RayD r;
    float idx = 0.f, idy = 0.f, idz = 0.f;
    float best_t = 0.f, best_b1 = 0.f, best_b2 = 0.f;
    int best_prim = 0x7fffffff;
    bool found = false, active = false, exhausted = false;
    int sp = 0;
    int32_t cur = 0;
    int64_t my = 0, w_next = 0, w_end = 0;      // warp-uniform: the warp's current chunk of the ray pool

#define PT_PUSH(v) do { if (sp < PT_SSTACK) s_stack[sp][tid] = (v); else l_stack[sp - PT_SSTACK] = (v); ++sp; } while (0)
#define PT_POP() (--sp, (sp < PT_SSTACK) ? s_stack[sp][tid] : l_stack[sp - PT_SSTACK])

    for (;;) {
        // ---- refill idle lanes: the warp owns a chunk [w_next, w_end) of consecutive rays (one atomicAdd on the
        // global pool per chunk), idle lanes take the next rays of the chunk in lane order ---------------------------
        uint32_t idle = __ballot_sync(0xffffffffu, !active);
        while (idle && !exhausted) {
            if (w_next >= w_end) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(pool, (unsigned long long)chunk);
                base = __shfl_sync(0xffffffffu, base, 0);
                if ((int64_t)base >= nrays) { exhausted = true; break; }
                w_next = (int64_t)base;
                w_end = (int64_t)base + chunk < nrays ? (int64_t)base + chunk : nrays;
            }
            const int rank = __popc(idle & lt);
            const int64_t i = w_next + rank;
            if (!active && i < w_end) {
                const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
                const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
                r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.tmin = r0.w;
                r.dx = r1.x; r.dy = r1.y; r.dz = r1.z;
                r.ddx = r1.x; r.ddy = r1.y; r.ddz = r1.z;
                idx = slab_rcp(r1.x); idy = slab_rcp(r1.y); idz = slab_rcp(r1.z);
                best_t = r1.w; best_prim = 0x7fffffff; best_b1 = 0.f; best_b2 = 0.f;
                found = false; sp = 0; cur = 0; my = i; active = true;
            }
            const int64_t avail = w_end - w_next;
            const int want = __popc(idle);
            w_next += want < avail ? want : avail;
            idle = __ballot_sync(0xffffffffu, !active);
        }
        if (!__any_sync(0xffffffffu, active)) break;

        // ---- traverse until this ray ends or the warp has thinned out ------------------------------------------
        while (active) {
            bool done = false;
            // internal nodes
            while (cur >= 0) {
                const float4 *np = reinterpret_cast<const float4 *>(nodes + cur);
                const float4 lox = __ldg(np), hix = __ldg(np + 1), loy = __ldg(np + 2), hiy = __ldg(np + 3),
                             loz = __ldg(np + 4), hiz = __ldg(np + 5);
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
                    float tn_p = tnear - fabsf(tnear) * 2e-6f;
                    float tf_p = tfar + fabsf(tfar) * 2e-6f;
                    bool h = (tn_p <= tf_p) && (cr[k] != B2RT_CHILD_EMPTY);
                    tn[k] = h ? tnear : INFINITY;
                }
#define B2RT_CSWAP(a, b)                                                                  \
    if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; int32_t cc = cr[a]; cr[a] = cr[b]; cr[b] = cc; }
                B2RT_CSWAP(0, 1) B2RT_CSWAP(2, 3) B2RT_CSWAP(0, 2) B2RT_CSWAP(1, 3) B2RT_CSWAP(1, 2)
#undef B2RT_CSWAP
                const int nhit = (tn[0] < INFINITY) + (tn[1] < INFINITY) + (tn[2] < INFINITY) + (tn[3] < INFINITY);
                if (nhit == 0) {
                    if (sp == 0) { done = true; break; }
                    cur = PT_POP();
                    continue;
                }
                if (sp + 3 > B2RT_STACK_SIZE) { atomicAdd(status, 1u); done = true; break; }
                if (nhit > 3) PT_PUSH(cr[3]);
                if (nhit > 2) PT_PUSH(cr[2]);
                if (nhit > 1) PT_PUSH(cr[1]);
                cur = cr[0];
            }
            if (!done) {
                // leaf: 1..8 primitive records
                const uint32_t ref = (uint32_t)cur & 0x7fffffffu;
                const uint32_t first = ref >> 3, cnt = (ref & 7u) + 1u;
                for (uint32_t s = first; s < first + cnt; ++s) {
                    const float4 *pp = reinterpret_cast<const float4 *>(recs + s);
                    const float4 A = __ldg(pp), B = __ldg(pp + 1), C = __ldg(pp + 2);
                    int prim = __float_as_int(A.w);
                    const int type = __float_as_int(B.w);
                    float t, b1 = 0.f, b2 = 0.f;
                    bool h;
                    if (type == B2RT_PRIM_TRI) h = tri_test(r, A, B, C, best_t, t, b1, b2);
                    else if (type == B2RT_PRIM_SPHERE) h = sphere_test(r, analytic[__float_as_int(C.w)], best_t, t);
                    else if (!INST || type == B2RT_PRIM_DISK) h = disk_test(r, analytic[__float_as_int(C.w)], best_t, t);
                    else {                                  // object instance (two-level BVH): b2rt_device.cuh instance_test
                        int lprim, lslot;
                        bool iovf = false;
                        h = instance_test<ANY>(r, instance_of(A), best_t, t, b1, b2, lprim, lslot, iovf);
                        prim += lprim;
                        if (iovf) atomicAdd(status, 1u);
                    }
                    if (h) {
                        if (ANY) { found = true; break; }
                        if (t < best_t || prim < best_prim) {
                            best_t = t; best_prim = prim; best_b1 = b1; best_b2 = b2; found = true;
                        }
                    }
                }
                if ((ANY && found) || sp == 0) done = true;
                else cur = PT_POP();
            }
            if (done) {
                if (ANY) {
                    occ[my] = found ? 1 : 0;
                } else {
                    float4 o;
                    o.x = found ? best_t : INFINITY; o.y = __int_as_float(found ? best_prim : -1);
                    o.z = best_b1; o.w = best_b2;
                    *reinterpret_cast<float4 *>(hits + my) = o;
                }
                active = false;
                break;
            }
            if (!exhausted && __popc(__activemask()) < refill) break;
        }
        __syncwarp();
    }
#undef PT_PUSH
#undef PT_POP
}

// ---- speculative variant (B2RT_TRACE_VARIANT=2) ------------------------------------------------------------------------
// Same pool / shared-memory stack as above, plus:
//   * postponed leaves (Aila & Laine's speculative while-while): a lane that reaches a leaf parks it and keeps
//     descending; the warp leaves the node loop only when every lane has a parked leaf or nothing left to visit,
//     so both the node loop and the primitive loop run with more lanes;
//   * a branch-free triangle test (same operations and roundings as tri_test, one predicate at the end).
// Box tests only prune; accepted primitives and their (t, b1, b2) are computed by the same expressions, so the
// results are bit-identical to the other variants.
// SORT: 2 = children visited in full near-to-far order, 1 = nearest first, the rest in slot order, 0 = slot order
// PL (B2RT_TRACE_PLANES, default on): the near and far plane of every slab are picked by the SIGN of the ray direction --
// as the ADDRESS of the 16-byte load, lo or hi array of the node -- instead of min / max per child: the same entry and exit
// distances (rounded subtraction and multiplication are monotonic, so (lo - o) * id <= (hi - o) * id for id >= 0 and the
// reverse for id < 0, with the clamped reciprocal also for d = +-0) with 6 instead of 12 FMNMX per child.  ncu r04i: the
// ALU pipe (min / max, selects, compares) is the busiest pipe of this kernel, 48 / 54 % of its peak at 18 / 21.6 active
// threads per instruction.
template <bool ANY, int SORT, bool PL, bool INST>
__global__ void __launch_bounds__(PT_THREADS, ANY ? B2RT_TRACE_BLOCKS_ANY : B2RT_TRACE_BLOCKS_CLOSEST)
trace_spec_kernel(const WideNode *__restrict__ nodes, const PrimRec *__restrict__ recs,
                  const AnalyticPrim *__restrict__ analytic, const b2rt_ray *__restrict__ rays,
                  b2rt_hit *__restrict__ hits, uint8_t *__restrict__ occ, int64_t nrays,
                  unsigned long long *pool, int refill, int chunk, uint32_t *status, unsigned long long *next_pool) {
    __shared__ int32_t s_stack[PT_SSTACK][PT_THREADS];
    int32_t l_stack[B2RT_STACK_SIZE - PT_SSTACK];
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt = (1u << lane) - 1u;
    // the ray-pool counter of the NEXT launch (the two counters alternate): zeroing it here saves the separate reset
    // launch in front of every trace call
    if (blockIdx.x == 0 && tid == 0) *next_pool = 0ull;
    constexpr int32_t NONE = B2RT_CHILD_EMPTY;       // "nothing": never a valid node index or leaf reference

    RayD r;
    float idx = 0.f, idy = 0.f, idz = 0.f;
#ifdef B2RT_STATS
    uint32_t st_nodes = 0, st_prims = 0;    // -DB2RT_STATS: wide nodes fetched / primitive records tested by this lane
#endif
    int sg_ = 0;                            // PL: bit a set = direction component a is negative (near plane = hi array)
    float best_t = 0.f, best_b1 = 0.f, best_b2 = 0.f;
    int best_prim = 0x7fffffff;
    bool found = false, active = false, exhausted = false;
    int sp = 0;
    int32_t cur = NONE, leaf = NONE;
    int64_t my = 0, w_next = 0, w_end = 0;

#define PT_PUSH(v) do { if (sp < PT_SSTACK) s_stack[sp][tid] = (v); else l_stack[sp - PT_SSTACK] = (v); ++sp; } while (0)
#define PT_POP() (--sp, (sp < PT_SSTACK) ? s_stack[sp][tid] : l_stack[sp - PT_SSTACK])

    for (;;) {
        uint32_t idle = __ballot_sync(0xffffffffu, !active);
        while (idle && !exhausted) {
            if (w_next >= w_end) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(pool, (unsigned long long)chunk);
                base = __shfl_sync(0xffffffffu, base, 0);
                if ((int64_t)base >= nrays) { exhausted = true; break; }
                w_next = (int64_t)base;
                w_end = (int64_t)base + chunk < nrays ? (int64_t)base + chunk : nrays;
            }
            const int64_t i = w_next + __popc(idle & lt);
            if (!active && i < w_end) {
                const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
                const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
                r.ox = r0.x; r.oy = r0.y; r.oz = r0.z; r.tmin = r0.w;
                r.dx = r1.x; r.dy = r1.y; r.dz = r1.z;
                r.ddx = r1.x; r.ddy = r1.y; r.ddz = r1.z;
                idx = slab_rcp(r1.x); idy = slab_rcp(r1.y); idz = slab_rcp(r1.z);
                if (PL) sg_ = (signbit(idx) ? 1 : 0) | (signbit(idy) ? 2 : 0) | (signbit(idz) ? 4 : 0);
                best_t = r1.w; best_prim = 0x7fffffff; best_b1 = 0.f; best_b2 = 0.f;
                found = false; sp = 0; cur = 0; leaf = NONE; my = i; active = true;
            }
            const int64_t avail = w_end - w_next;
            const int want = __popc(idle);
            w_next += want < avail ? want : avail;
            idle = __ballot_sync(0xffffffffu, !active);
        }
        if (!__any_sync(0xffffffffu, active)) break;

        while (active) {
            // ---- node loop: descend; park the first leaf met and keep going until the whole warp has one ---------
            while (cur >= 0 && cur != NONE) {
                const float4 *np = reinterpret_cast<const float4 *>(nodes + cur);
#ifdef B2RT_STATS
                ++st_nodes;
#endif
                // PL: lox / loy / loz hold the NEAR planes, hix / hiy / hiz the FAR ones (picked by the load address)
                const int nx_ = sg_ & 1, ny_ = 2 + ((sg_ >> 1) & 1), nz_ = 4 + ((sg_ >> 2) & 1);
                const float4 lox = __ldg(np + (PL ? nx_ : 0)), hix = __ldg(np + (PL ? (nx_ ^ 1) : 1)),
                             loy = __ldg(np + (PL ? ny_ : 2)), hiy = __ldg(np + (PL ? (ny_ ^ 1) : 3)),
                             loz = __ldg(np + (PL ? nz_ : 4)), hiz = __ldg(np + (PL ? (nz_ ^ 1) : 5));
                const int4 ch = __ldg(reinterpret_cast<const int4 *>(np + 6));
                float tn[4]; int32_t cr[4];
                const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, hx[4] = {hix.x, hix.y, hix.z, hix.w};
                const float ly[4] = {loy.x, loy.y, loy.z, loy.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w};
                const float lz[4] = {loz.x, loz.y, loz.z, loz.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
                cr[0] = ch.x; cr[1] = ch.y; cr[2] = ch.z; cr[3] = ch.w;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float t0x = (lx[k] - r.ox) * idx, t1x = (hx[k] - r.ox) * idx;
                    const float t0y = (ly[k] - r.oy) * idy, t1y = (hy[k] - r.oy) * idy;
                    const float t0z = (lz[k] - r.oz) * idz, t1z = (hz[k] - r.oz) * idz;
                    const float tnear = PL ? fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, r.tmin))
                                           : fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), r.tmin));
                    const float tfar = PL ? fminf(fminf(t1x, t1y), fminf(t1z, best_t))
                                          : fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), best_t));
                    const float tn_p = __fmaf_rn(fabsf(tnear), -2e-6f, tnear);      // conservative pad, one FFMA each
                    const float tf_p = __fmaf_rn(fabsf(tfar), 2e-6f, tfar);
                    const bool h = (tn_p <= tf_p) && (cr[k] != B2RT_CHILD_EMPTY);
                    tn[k] = h ? tnear : INFINITY;
                }
#define B2RT_CSWAP(a, b)                                                                  \
    if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; int32_t cc = cr[a]; cr[a] = cr[b]; cr[b] = cc; }
                if (SORT >= 1) { B2RT_CSWAP(0, 1) B2RT_CSWAP(2, 3) B2RT_CSWAP(0, 2) }
                if (SORT == 2) { B2RT_CSWAP(1, 3) B2RT_CSWAP(1, 2) }
#undef B2RT_CSWAP
                const int nhit = (tn[0] < INFINITY) + (tn[1] < INFINITY) + (tn[2] < INFINITY) + (tn[3] < INFINITY);
                if (SORT < 2) {
                    if (sp + 3 > B2RT_STACK_SIZE) { atomicAdd(status, 1u); sp = 0; cur = NONE; leaf = NONE; break; }
                    int32_t nxt = NONE;
#pragma unroll
                    for (int k = 3; k >= 0; --k)
                        if (tn[k] < INFINITY) { if (nxt != NONE) PT_PUSH(nxt); nxt = cr[k]; }
                    cur = nxt != NONE ? nxt : (sp > 0 ? PT_POP() : NONE);
                } else if (nhit == 0) {
                    cur = sp > 0 ? PT_POP() : NONE;
                } else {
                    if (sp + 3 > B2RT_STACK_SIZE) { atomicAdd(status, 1u); sp = 0; cur = NONE; leaf = NONE; break; }
                    if (nhit > 3) PT_PUSH(cr[3]);
                    if (nhit > 2) PT_PUSH(cr[2]);
                    if (nhit > 1) PT_PUSH(cr[1]);
                    cur = cr[0];
                }
                if (cur < 0 && leaf == NONE) {          // park this leaf, continue with the next entry
                    leaf = cur;
                    cur = sp > 0 ? PT_POP() : NONE;
                }
                if (!__any_sync(__activemask(), leaf == NONE)) break;
            }
            // ---- primitive loop: the parked leaf, then `cur` while it is a leaf as well ------------------------------
            while (leaf != NONE) {
                const uint32_t ref = (uint32_t)leaf & 0x7fffffffu;
                const uint32_t first = ref >> 3, cnt = (ref & 7u) + 1u;
                for (uint32_t s = first; s < first + cnt; ++s) {
                    const float4 *pp = reinterpret_cast<const float4 *>(recs + s);
                    const float4 A = __ldg(pp), B = __ldg(pp + 1), C = __ldg(pp + 2);
#ifdef B2RT_STATS
                    ++st_prims;
#endif
                    int prim = __float_as_int(A.w);
                    const int type = __float_as_int(B.w);
                    float t, b1 = 0.f, b2 = 0.f;
                    bool h;
                    if (type == B2RT_PRIM_TRI) h = tri_test_nobranch(r, A, B, C, best_t, t, b1, b2);
                    else if (type == B2RT_PRIM_SPHERE) h = sphere_test(r, analytic[__float_as_int(C.w)], best_t, t);
                    else if (!INST || type == B2RT_PRIM_DISK) h = disk_test(r, analytic[__float_as_int(C.w)], best_t, t);
                    else {                                  // object instance (two-level BVH): b2rt_device.cuh instance_test
                        int lprim, lslot;
                        bool iovf = false;
                        h = instance_test<ANY>(r, instance_of(A), best_t, t, b1, b2, lprim, lslot, iovf);
                        prim += lprim;
                        if (iovf) atomicAdd(status, 1u);
                    }
                    if (h) {
                        if (ANY) { found = true; break; }
                        if (t < best_t || prim < best_prim) {
                            best_t = t; best_prim = prim; best_b1 = b1; best_b2 = b2; found = true;
                        }
                    }
                }
                if (ANY && found) { cur = NONE; sp = 0; leaf = NONE; break; }
                leaf = NONE;
                if (cur < 0) {                          // the next entry is a leaf too
                    leaf = cur;
                    cur = sp > 0 ? PT_POP() : NONE;
                }
            }
            if (cur == NONE) {                          // nothing left: leaf == NONE here
                if (ANY) {
                    occ[my] = found ? 1 : 0;
                } else {
                    float4 o;
                    o.x = found ? best_t : INFINITY; o.y = __int_as_float(found ? best_prim : -1);
                    o.z = best_b1; o.w = best_b2;
                    *reinterpret_cast<float4 *>(hits + my) = o;
                }
                active = false;
                break;
            }
            if (!exhausted && __popc(__activemask()) < refill) break;
        }
        __syncwarp();
    }
#ifdef B2RT_STATS
    // one pair of 64-bit adds per warp: d_counters words 24-27
    for (int o = 16; o > 0; o >>= 1) { st_nodes += __shfl_xor_sync(0xffffffffu, st_nodes, o); st_prims += __shfl_xor_sync(0xffffffffu, st_prims, o); }
    if (lane == 0) {
        atomicAdd(reinterpret_cast<unsigned long long *>(status + 16), (unsigned long long)st_nodes);
        atomicAdd(reinterpret_cast<unsigned long long *>(status + 18), (unsigned long long)st_prims);
    }
#endif
#undef PT_PUSH
#undef PT_POP
}

__global__ void pool_reset_kernel(unsigned long long *pool) { *pool = 0ull; }

}  // namespace
int b2rt_trace_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_hit *d_hits, uint8_t *d_occ,
                    bool any_hit) {
    if (n <= 0) return B2RT_OK;
    const WideNode *nodes = ctx->d_wide_nodes.as<WideNode>();
    const PrimRec *recs = ctx->d_prim_recs.as<PrimRec>();
    const AnalyticPrim *an = ctx->d_analytic.as<AnalyticPrim>();
    uint32_t *status = ctx->d_counters.as<uint32_t>() + 8;
    const bool inst = !ctx->h_instances.empty();
    if (ctx->trace_variant == 1 || ctx->trace_variant == 2) {
        const bool spec = ctx->trace_variant == 2;
        // persistent grid: every SM holds as many blocks as the occupancy calculator allows
        int &per_sm = ctx->pt_blocks_per_sm[(spec ? 2 : 0) + (any_hit ? 1 : 0)];
        if (per_sm == 0 && spec) {
            if (any_hit) B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_spec_kernel<true, 0, true, false>, PT_THREADS, 0));
            else B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_spec_kernel<false, 2, true, false>, PT_THREADS, 0));
            if (per_sm < 1) per_sm = 1;
        }
        if (per_sm == 0) {
            if (any_hit) B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_persistent_kernel<true, false>, PT_THREADS, 0));
            else B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_persistent_kernel<false, false>, PT_THREADS, 0));
            if (per_sm < 1) per_sm = 1;
        }
        int64_t blocks = (int64_t)ctx->sm_count * per_sm;
        int64_t need = (n + PT_THREADS - 1) / PT_THREADS;
        if (blocks > need) blocks = need;
        unsigned long long *pool = reinterpret_cast<unsigned long long *>(ctx->d_counters.as<uint32_t>() + 16);
        const int chunk = ctx->trace_chunk >= 32 ? (ctx->trace_chunk & ~31) : 32;
        unsigned long long *next_pool = pool;
        if (spec) {                     // two counters (uint32 slots 16-17 and 20-21, zero at creation) used in turn
            if (ctx->trace_pool_flip) pool = reinterpret_cast<unsigned long long *>(ctx->d_counters.as<uint32_t>() + 20);
            else next_pool = reinterpret_cast<unsigned long long *>(ctx->d_counters.as<uint32_t>() + 20);
            ctx->trace_pool_flip ^= 1;
        } else {
            pool_reset_kernel<<<1, 1, 0, ctx->stream>>>(pool);
            B2RT_LAUNCH_CHECK(ctx);
        }
        if (spec) {
            // default (trace_sort < 0): closest-hit rays visit children near to far; any-hit rays take them in slot order
            // (no sorting network: +9 % / +5 % on the 1 M / 8 M-triangle shadow rays, run r02b; for closest-hit rays slot
            // order costs 25-35 % and nearest-first-only is +2 % / -5 %)
            const int so = ctx->trace_sort < 0 ? (any_hit ? 0 : 2) : ctx->trace_sort;
            auto pick = [&](auto planes) {
                constexpr bool P = decltype(planes)::value;
                if (so == 2) return any_hit ? trace_spec_kernel<true, 2, P, false> : trace_spec_kernel<false, 2, P, false>;
                if (so == 1) return any_hit ? trace_spec_kernel<true, 1, P, false> : trace_spec_kernel<false, 1, P, false>;
                return any_hit ? trace_spec_kernel<true, 0, P, false> : trace_spec_kernel<false, 0, P, false>;
            };
            // r04k A/B (identical bytes): closest-hit rays 6426 -> 6570 Mrays/s, incoherent 3141 -> 3191, any-hit rays 7533 ->
            // 7455 -- so the sign-selected planes are the default for closest hit only (1), 2 = both, 0 = neither
            const bool planes = ctx->trace_planes == 2 || (ctx->trace_planes == 1 && !any_hit);
            auto kern = planes ? pick(std::true_type()) : pick(std::false_type());
            // scenes with object instances: the instantiation that knows the instance record (one per ray type, default child
            // order and plane selection) -- kept apart so that the instance call costs the other scenes no register or spill
            if (inst) kern = any_hit ? trace_spec_kernel<true, 0, false, true> : trace_spec_kernel<false, 2, true, true>;
            kern<<<(unsigned)blocks, PT_THREADS, 0, ctx->stream>>>(nodes, recs, an, d_rays, any_hit ? nullptr : d_hits,
                                                                   any_hit ? d_occ : nullptr, n, pool, ctx->trace_refill,
                                                                   chunk, status, next_pool);
        }
        else {
            auto kern = any_hit ? (inst ? trace_persistent_kernel<true, true> : trace_persistent_kernel<true, false>)
                                : (inst ? trace_persistent_kernel<false, true> : trace_persistent_kernel<false, false>);
            kern<<<(unsigned)blocks, PT_THREADS, 0, ctx->stream>>>(nodes, recs, an, d_rays, any_hit ? nullptr : d_hits,
                                                                   any_hit ? d_occ : nullptr, n, pool, ctx->trace_refill, chunk, status);
        }
        B2RT_LAUNCH_CHECK(ctx);
        return B2RT_OK;
    }
    int64_t grid = (n + TRACE_THREADS - 1) / TRACE_THREADS;
    auto kern = any_hit ? (inst ? trace_kernel<true, true> : trace_kernel<true, false>) : (inst ? trace_kernel<false, true> : trace_kernel<false, false>);
    kern<<<(unsigned)grid, TRACE_THREADS, 0, ctx->stream>>>(nodes, recs, an, d_rays, any_hit ? nullptr : d_hits, any_hit ? d_occ : nullptr, n, status);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}
