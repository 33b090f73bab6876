// K1 + K3 + K4: deterministic LBVH builder (bounds + Morton, Karras topology + refit + SAH leaf
// collapse, greedy wide collapse to 4-wide 128-byte nodes).  K2 (the sort) lives in sort.cu.
//
// Replaces the reference's request for OptiX's closed "Sbvh" builder (cuda_render/cudarender.cpp:44,50)
// fed by trianglemesh_bounds / sphere_bounds / disk_bounds (cuda_render/util/shape/cudatrianglemesh.cu:83-99,
// cudasphere.cu:74-79, cudadisk.cu:87-96).  CPU analogue in the oracle: BVHAccel
// (pbrt-v2/src/accelerators/bvh.cpp:153-380) -- a different (SAH-recursive) tree, so topology parity
// is against oracle/oracle_port.c (same arithmetic, -ffp-contract=off there, -fmad=false here) and
// hit parity is against pbrt itself.
//
// Every float result below is order-independent (min/max) or evaluated in one fixed order, and every
// index comes from integer prefix sums, so the output bytes are identical across runs and GPUs.
#include "b2rt_internal.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
#include <math.h>
#include <stdlib.h>

namespace {

__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(b);
}

__global__ void init_scene_box_kernel(uint32_t *box) {
    int i = threadIdx.x;
    if (i < 12) box[i] = ((i % 6) < 3) ? 0xffffffffu : 0u;  // mins start at +max, maxes at 0
}

__device__ __forceinline__ void xform_point(const float *m, float x, float y, float z, float &ox, float &oy,
                                            float &oz) {
    ox = m[0] * x + m[1] * y + m[2] * z + m[3];
    oy = m[4] * x + m[5] * y + m[6] * z + m[7];
    oz = m[8] * x + m[9] * y + m[10] * z + m[11];
}

// K1a: per-prim AABB -> leaf_box[6*p]; block-reduced centroid and scene bounds -> ordered-int atomics.
__global__ void __launch_bounds__(256)
prim_bounds_kernel(const float *__restrict__ verts, const int32_t *__restrict__ idx,
                   const int32_t *__restrict__ prim_kind, const AnalyticPrim *__restrict__ analytic,
                   const int32_t *__restrict__ analytic_type, const InstanceRec *__restrict__ instances, int n,
                   float *__restrict__ leaf_box, uint32_t *__restrict__ scene_box) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool ok = p < n;
    if (ok) {
        int kind = prim_kind[p];
        if (kind >= 0) {
            const int32_t *t = idx + 3 * (int64_t)kind;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float *v = verts + 3 * (int64_t)t[k];
#pragma unroll
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], v[a]); hi[a] = fmaxf(hi[a], v[a]); }
            }
        } else if ((~kind) & 0x40000000) {
            // object instance: the object's box through the instance transform (8 corners), padded like the analytic boxes
            const InstanceRec &I = instances[(~kind) & 0x3fffffff];
            for (int c = 0; c < 8; ++c) {
                float x, y, z;
                xform_point(I.o2w, I.box[(c & 1) ? 3 : 0], I.box[(c & 2) ? 4 : 1], I.box[(c & 4) ? 5 : 2], x, y, z);
                float q[3] = {x, y, z};
#pragma unroll
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], q[a]); hi[a] = fmaxf(hi[a], q[a]); }
            }
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                float pad = 1e-5f * fmaxf(fmaxf(fabsf(lo[a]), fabsf(hi[a])), hi[a] - lo[a]);
                lo[a] -= pad; hi[a] += pad;
            }
        } else {
            const AnalyticPrim &A = analytic[~kind];
            float r = A.radius;
            float z0 = analytic_type[~kind] == B2RT_PRIM_SPHERE ? -r : A.height;
            float z1 = analytic_type[~kind] == B2RT_PRIM_SPHERE ? r : A.height;
            for (int c = 0; c < 8; ++c) {
                float x, y, z;
                xform_point(A.o2w, (c & 1) ? r : -r, (c & 2) ? r : -r, (c & 4) ? z1 : z0, x, y, z);
                float q[3] = {x, y, z};
#pragma unroll
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], q[a]); hi[a] = fmaxf(hi[a], q[a]); }
            }
            // pad analytic boxes by a few ulps of their extent: the corner transform is rounded
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                float pad = 1e-5f * fmaxf(fmaxf(fabsf(lo[a]), fabsf(hi[a])), hi[a] - lo[a]);
                lo[a] -= pad; hi[a] += pad;
            }
        }
        float *b = leaf_box + 6 * (int64_t)p;
        b[0] = lo[0]; b[1] = lo[1]; b[2] = lo[2]; b[3] = hi[0]; b[4] = hi[1]; b[5] = hi[2];
    }
    // centroid (same expression as the Morton kernel)
    float c[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) c[a] = 0.5f * (lo[a] + hi[a]);
    float cmin[3], cmax[3], smin[3], smax[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        cmin[a] = ok ? c[a] : INFINITY; cmax[a] = ok ? c[a] : -INFINITY;
        smin[a] = lo[a]; smax[a] = hi[a];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cmin[a] = fminf(cmin[a], __shfl_xor_sync(0xffffffffu, cmin[a], o));
            cmax[a] = fmaxf(cmax[a], __shfl_xor_sync(0xffffffffu, cmax[a], o));
            smin[a] = fminf(smin[a], __shfl_xor_sync(0xffffffffu, smin[a], o));
            smax[a] = fmaxf(smax[a], __shfl_xor_sync(0xffffffffu, smax[a], o));
        }
    }
    // warp results -> shared-memory atomics -> 12 global atomics per BLOCK (min / max are order-independent, so the
    // bounds stay bit-exact; one global atomic per warp and value serialised 375 k atomics on 12 addresses at 1 M prims)
    __shared__ uint32_t s_box[12];
    if (threadIdx.x < 12) s_box[threadIdx.x] = (threadIdx.x % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (cmin[a] <= cmax[a]) {
                atomicMin(&s_box[a], f2ord(cmin[a]));
                atomicMax(&s_box[3 + a], f2ord(cmax[a]));
                atomicMin(&s_box[6 + a], f2ord(smin[a]));
                atomicMax(&s_box[9 + a], f2ord(smax[a]));
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const uint32_t v = s_box[threadIdx.x];
        if ((threadIdx.x % 6) < 3) { if (v != 0xffffffffu) atomicMin(&scene_box[threadIdx.x], v); }
        else if (v != 0u) atomicMax(&scene_box[threadIdx.x], v);
    }
}

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// K1b: 30-bit Morton code of the quantised centroid; value = prim id (so a stable sort breaks ties by id)
__global__ void __launch_bounds__(256)
morton_kernel(const float *__restrict__ leaf_box, const uint32_t *__restrict__ scene_box, int n,
              uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float *b = leaf_box + 6 * (int64_t)p;
    uint32_t q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float cmin = ord2f(scene_box[a]), cmax = ord2f(scene_box[3 + a]);
        float ext = cmax - cmin;
        float scale = ext > 0.f ? 1024.f / ext : 0.f;
        float c = 0.5f * (b[a] + b[3 + a]);
        float f = (c - cmin) * scale;
        int qi = (int)f;
        q[a] = (uint32_t)min(max(qi, 0), 1023);
    }
    keys[p] = (expand_bits10(q[0]) << 2) | (expand_bits10(q[1]) << 1) | expand_bits10(q[2]);
    vals[p] = (uint32_t)p;
}

// ---- K3: Karras 2012 topology -----------------------------------------------------------------
__device__ __forceinline__ int delta(const uint32_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(a ^ b);
}

// node numbering: internal i in [0,n-1) -> i ; leaf j -> (n-1)+j
__global__ void __launch_bounds__(256)
karras_topology_kernel(const uint32_t *__restrict__ keys, int n, int32_t *__restrict__ parent,
                       int32_t *__restrict__ child, int2 *__restrict__ range) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    int gamma = i + s * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    int left = (first == gamma) ? (n - 1) + gamma : gamma;
    int right = (last == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    child[2 * i] = left; child[2 * i + 1] = right;
    parent[left] = i; parent[right] = i;
    range[i] = make_int2(first, last);
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ float box_area(const float *b) {
    float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return 2.f * (dx * dy + dy * dz + dz * dx);
}

// bottom-up refit + SAH leaf-collapse decision.  flag[i]: arrival counter.  collapse[i]=1 when the
// whole subtree of internal i is cheaper as one leaf (count <= B2RT_LEAF_MAX).
// The kernel is bound by the dependent round trips to L2 on the way to the root (1 M prims: ~40 levels), not by bytes: the
// climbing thread keeps its own box and cost in registers and reads only its sibling's; the arrival counter, the parent,
// the children and the range of a node are requested together (one round trip instead of four in a row).  Min / max and
// cl + cr are commutative, so taking (mine, sibling) instead of (left, right) gives the same bits.
__global__ void __launch_bounds__(256)
refit_kernel(const float *__restrict__ leaf_box, const uint32_t *__restrict__ vals, int n,
             const int32_t *__restrict__ parent, const int32_t *__restrict__ child,
             const int2 *__restrict__ range, float *box, float *cost, uint32_t *flag,
             uint8_t *__restrict__ collapse, int4 *__restrict__ info) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int node = (n - 1) + j;
    float b[6], c_me;
    {
        const float *s = leaf_box + 6 * (int64_t)vals[j];
        float *d = box + 6 * (int64_t)node;
#pragma unroll
        for (int a = 0; a < 6; ++a) { b[a] = s[a]; d[a] = b[a]; }
        c_me = box_area(b);  // one prim, prim cost 1
        cost[node] = c_me;
    }
    int cur = parent[node];
    while (cur >= 0) {
        const int2 ch = *reinterpret_cast<const int2 *>(child + 2 * (int64_t)cur);
        const int2 rg = range[cur];
        const int up = parent[cur];
        // release: my box / cost are visible before I announce myself; acquire: the sibling's are visible to me once I have
        // seen its announcement.  One acq_rel atomic instead of __threadfence() (a sequentially consistent MEMBAR.SC) on
        // both sides of a relaxed one: the two fences were 55 % of the kernel's stall samples (ncu r07i: membar 39.8 warps
        // per issue).
        uint32_t seen;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(seen) : "l"(flag + cur) : "memory");
        if (seen == 0u) return;                            // first arriver leaves; second owns the node
        const int sib = ch.x == node ? ch.y : ch.x;
        const volatile float *bs = box + 6 * (int64_t)sib;
        float sb[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) sb[a] = bs[a];
        const float c_sib = ((const volatile float *)cost)[sib];
#pragma unroll
        for (int a = 0; a < 3; ++a) { b[a] = fminf(b[a], sb[a]); b[3 + a] = fmaxf(b[3 + a], sb[3 + a]); }
        float *d = box + 6 * (int64_t)cur;
#pragma unroll
        for (int a = 0; a < 6; ++a) d[a] = b[a];
        float area = box_area(b);
        float c_int = area + (c_me + c_sib);  // node cost 1
        int cnt = rg.y - rg.x + 1;
        float c_leaf = area * (float)cnt;
        bool col = (cnt <= B2RT_LEAF_MAX) && (c_leaf <= c_int);
        collapse[cur] = col ? 1 : 0;
        c_me = col ? c_leaf : c_int;
        cost[cur] = c_me;
        // everything the wide collapse needs to know about this node in one 16-byte record
        info[cur] = make_int4(ch.x, ch.y, __float_as_int(area), col ? 0 : 1);
        node = cur;
        cur = up;
    }
}

// ---- K4: wide collapse ---------------------------------------------------------------------------
struct WideTmp { int32_t c[4]; };  // binary node ids of the (up to 4) children, -1 = unused

__device__ __forceinline__ bool is_openable(int node, int n, const uint8_t *collapse) {
    return node < n - 1 && !collapse[node];
}

// The collapse without the host: ONE cooperative launch walks the levels itself, grid.sync() between them (one thread per
// wide root of the level's queue).  The level-synchronous host loop of round 1 (one launch per level) cost one launch, one 4-byte download and one stream synchronisation per level (15 levels
// for the 1 M-triangle scene).  A first device-side version (r04c) used a single work queue with a ticket, a tail and a
// pending counter: three same-address atomics per wide node, 568 us for 266 k nodes -- slower than the host loop.  Here the
// only shared counter is the next level's queue tail, bumped once per warp (ballot + prefix over the lanes' child counts).
// The queue order is scheduling-dependent but never reaches the output: a root's children are a function of the binary
// tree alone and node indices come from the prefix sum over wide_flag.
// ctl[0 .. 2] = queue sizes of levels L, L + 1, L + 2 in rotation (the third one is zeroed while level L runs, so one
// grid.sync per level is enough), ctl[5] = number of levels
__global__ void __launch_bounds__(256)
wide_collapse_coop_kernel(int n, const int4 *__restrict__ info, WideTmp *__restrict__ tmp,
                          uint32_t *__restrict__ wide_flag, int32_t *qa, int32_t *qb, uint32_t *ctl) {
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int64_t gsize = (int64_t)gridDim.x * blockDim.x;
    int cur = 0, levels = 0;
    for (;;) {
        const int c0 = levels % 3, c1 = (levels + 1) % 3, c2 = (levels + 2) % 3;
        const uint32_t nq = ((volatile uint32_t *)ctl)[c0];
        if (nq == 0) break;
        ++levels;
        if (blockIdx.x == 0 && threadIdx.x == 0) ((volatile uint32_t *)ctl)[c2] = 0u;     // the level after next starts empty
        const int32_t *qin = cur ? qb : qa;
        int32_t *qout = cur ? qa : qb;
        // whole warps iterate together (the ballot below needs every lane)
        for (int64_t w = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); w < (int64_t)nq; w += gsize) {
            const int64_t qi = w + lane;
            int open[4], no = 0;
            bool live = false;
            if (qi < (int64_t)nq) {
                const int root = qin[qi];
                // node record {left, right, bits(area), openable} written by the refit: the chain of dependent loads per
                // level is queue -> record of the root -> records of its two children -> records of the two children of
                // the first node opened (was: queue -> children -> collapse flags -> boxes -> children -> flags -> boxes ...)
                const int4 ri = __ldcg(info + root);
                if (ri.w) {                                 // the queue may hold collapsed nodes: they are pushed unseen
                    live = true;
                    int c[4], kl[4], kr[4], op[4]; float ar[4];
                    auto fetch = [&](int k) {
                        op[k] = 0; kl[k] = -1; kr[k] = -1; ar[k] = 0.f;
                        if (c[k] < n - 1) {
                            const int4 t = __ldcg(info + c[k]);
                            kl[k] = t.x; kr[k] = t.y; ar[k] = __int_as_float(t.z); op[k] = t.w;
                        }
                    };
                    int m = 2;
                    c[0] = ri.x; c[1] = ri.y; c[2] = -1; c[3] = -1;
                    fetch(0); fetch(1);
                    op[2] = op[3] = 0; kl[2] = kl[3] = kr[2] = kr[3] = -1; ar[2] = ar[3] = 0.f;
                    while (m < 4) {
                        int best = -1; float best_area = -1.f;
                        for (int k = 0; k < m; ++k)
                            if (op[k] && ar[k] > best_area) { best_area = ar[k]; best = k; }   // first max wins: deterministic
                        if (best < 0) break;
                        const int l = kl[best], r = kr[best];
                        // keep order: left child takes the slot, right child is inserted right after it
                        for (int k = m; k > best + 1; --k) { c[k] = c[k - 1]; kl[k] = kl[k - 1]; kr[k] = kr[k - 1]; op[k] = op[k - 1]; ar[k] = ar[k - 1]; }
                        c[best] = l; c[best + 1] = r;
                        ++m;
                        if (m < 4) { fetch(best); fetch(best + 1); }
                        else { op[best] = l < n - 1; op[best + 1] = r < n - 1; }     // not looked at: the next level filters them
                    }
                    WideTmp t;
#pragma unroll
                    for (int k = 0; k < 4; ++k) t.c[k] = c[k];
                    tmp[root] = t;
                    wide_flag[root] = 1u;
                    for (int k = 0; k < m; ++k)
                        if (op[k]) open[no++] = c[k];
                }
            }
            if (__any_sync(0xffffffffu, live) && lane == 0) atomicMax(&ctl[6], (uint32_t)levels);
            // one atomic per warp: exclusive prefix of the lanes' child counts
            int inc = no;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            const int total = __shfl_sync(0xffffffffu, inc, 31);
            uint32_t at = 0;
            if (lane == 31 && total > 0) at = atomicAdd(&ctl[c1], (uint32_t)total);
            at = __shfl_sync(0xffffffffu, at, 31) + (uint32_t)(inc - no);
            for (int k = 0; k < no; ++k) qout[at + k] = open[k];
        }
        cur ^= 1;
        grid.sync();
    }
    grid.sync();
    if (blockIdx.x == 0 && threadIdx.x == 0) ctl[5] = ((volatile uint32_t *)ctl)[6];      // levels that opened a node
}
__global__ void wide_collapse_init_kernel(int32_t *qa, uint32_t *ctl, const uint8_t *collapse) {
    // queue A = {root 0} unless the whole tree collapsed into one leaf (then there is nothing to do)
    const bool any = !collapse[0];
    ctl[0] = any ? 1u : 0u; ctl[1] = 0u; ctl[2] = 0u; ctl[5] = 0u; ctl[6] = 0u;
    if (any) qa[0] = 0;
}

// Absolute margin added to every child box stored in a wide node: 2^-21 of the largest root-box coordinate (~4 ulps at
// scene scale).  A slab test cannot follow a ray that runs exactly IN the plane shared by two sibling boxes (a ray
// through a mesh vertex, an axis-parallel ray along a wall): the interval of one of the two boxes ends at exactly 0 and
// no relative padding of t widens it, so triangles touching the plane from that side were never tested although
// pbrt's triangle test accepts the boundary hit (tests/test_gpu_parity.py::
// test_axis_parallel_and_degenerate_rays_match_oracle).  With the margin the ray is strictly inside both boxes.
// One float subtract / add per plane; oracle/oracle_port.c (port_lbvh_build) does the same, so node bytes stay bit-exact.
__device__ __forceinline__ float node_box_pad(const float *root) {
    float m = fmaxf(fmaxf(fmaxf(fabsf(root[0]), fabsf(root[1])), fmaxf(fabsf(root[2]), fabsf(root[3]))),
                    fmaxf(fabsf(root[4]), fabsf(root[5])));
    return m * 4.76837158203125e-07f;
}

__global__ void __launch_bounds__(128)
wide_emit_kernel(int n, const WideTmp *__restrict__ tmp, const uint32_t *__restrict__ wide_flag,
                 const uint32_t *__restrict__ wide_index, const float *__restrict__ box,
                 const int2 *__restrict__ range, const uint8_t *__restrict__ collapse,
                 WideNode *__restrict__ nodes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !wide_flag[i]) return;
    WideTmp t = tmp[i];
    WideNode w;
    int used = 0;
    const float pad = node_box_pad(box);        // box[0..5] = root
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int c = t.c[k];
        if (c < 0) {
            w.lox[k] = w.loy[k] = w.loz[k] = INFINITY;
            w.hix[k] = w.hiy[k] = w.hiz[k] = -INFINITY;
            w.child[k] = B2RT_CHILD_EMPTY;
            continue;
        }
        ++used;
        const float *b = box + 6 * (int64_t)c;
        w.lox[k] = b[0] - pad; w.loy[k] = b[1] - pad; w.loz[k] = b[2] - pad;
        w.hix[k] = b[3] + pad; w.hiy[k] = b[4] + pad; w.hiz[k] = b[5] + pad;
        if (c >= n - 1) {
            w.child[k] = (int32_t)(0x80000000u | ((uint32_t)(c - (n - 1)) << 3));          // 1 prim
        } else if (collapse[c]) {
            int2 rg = range[c];
            w.child[k] = (int32_t)(0x80000000u | ((uint32_t)rg.x << 3) | (uint32_t)(rg.y - rg.x));
        } else {
            w.child[k] = (int32_t)wide_index[c];
        }
    }
    w.meta[0] = used; w.meta[1] = i; w.meta[2] = 0; w.meta[3] = 0;
    nodes[wide_index[i]] = w;
}

// n == 1 (or a root that collapsed to a single leaf): one node, one leaf child covering everything
__global__ void wide_single_kernel(int n, const float *__restrict__ box_root, WideNode *__restrict__ nodes) {
    if (threadIdx.x != 0) return;
    WideNode w;
    for (int k = 0; k < 4; ++k) {
        w.lox[k] = w.loy[k] = w.loz[k] = INFINITY;
        w.hix[k] = w.hiy[k] = w.hiz[k] = -INFINITY;
        w.child[k] = B2RT_CHILD_EMPTY; w.meta[k] = 0;
    }
    if (n > 0) {
        const float pad = node_box_pad(box_root);
        w.lox[0] = box_root[0] - pad; w.loy[0] = box_root[1] - pad; w.loz[0] = box_root[2] - pad;
        w.hix[0] = box_root[3] + pad; w.hiy[0] = box_root[4] + pad; w.hiz[0] = box_root[5] + pad;
        w.child[0] = (int32_t)(0x80000000u | (uint32_t)(n - 1));
        w.meta[0] = 1;
    }
    nodes[0] = w;
}

// leaf-ordered primitive records + leaf_prims (prim id per slot)
__global__ void __launch_bounds__(256)
prim_records_kernel(const float *__restrict__ verts, const int32_t *__restrict__ idx,
                    const int32_t *__restrict__ prim_kind, const int32_t *__restrict__ analytic_type,
                    const InstanceRec *__restrict__ instances, const uint32_t *__restrict__ vals, int n,
                    PrimRec *__restrict__ recs, int32_t *__restrict__ leaf_prims) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int p = (int)vals[j];
    leaf_prims[j] = p;
    int kind = prim_kind[p];
    PrimRec r;
    if (kind >= 0) {
        const int32_t *t = idx + 3 * (int64_t)kind;
        const float *p1 = verts + 3 * (int64_t)t[0], *p2 = verts + 3 * (int64_t)t[1],
                    *p3 = verts + 3 * (int64_t)t[2];
        r.a = make_float4(p1[0], p1[1], p1[2], __int_as_float(p));
        r.b = make_float4(p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2], __int_as_float(B2RT_PRIM_TRI));
        r.c = make_float4(p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2], __int_as_float(0));
    } else if ((~kind) & 0x40000000) {
        const int k = (~kind) & 0x3fffffff;
        const uint64_t ptr = reinterpret_cast<uint64_t>(instances + k);     // the record carries the instance's address
        r.a = make_float4(__uint_as_float((uint32_t)ptr), __uint_as_float((uint32_t)(ptr >> 32)), 0.f, __int_as_float(instances[k].prim_base));
        r.b = make_float4(0, 0, 0, __int_as_float(B2RT_PRIM_INSTANCE));
        r.c = make_float4(0, 0, 0, __int_as_float(k));
    } else {
        r.a = make_float4(0, 0, 0, __int_as_float(p));
        r.b = make_float4(0, 0, 0, __int_as_float(analytic_type[~kind]));
        r.c = make_float4(0, 0, 0, __int_as_float(~kind));
    }
    recs[j] = r;
}

}  // namespace

int b2rt_build_impl(b2rt_ctx *ctx) {
    const int n = ctx->n_prims;
    cudaStream_t st = ctx->stream;
    ctx->built = false;
    B2RT_CUDA(ctx, cudaEventRecord(ctx->evb0, st));

    // ---- upload flattened scene -----------------------------------------------------------------
    std::vector<int32_t> kind(n);
    for (size_t t = 0; t < ctx->h_tri_prim.size(); ++t) kind[ctx->h_tri_prim[t]] = (int32_t)t;
    for (size_t a = 0; a < ctx->h_analytic_prim.size(); ++a) kind[ctx->h_analytic_prim[a]] = ~(int32_t)a;
    // two-level BVH: every object (child context: same device, same stream) builds its own tree in its own space first;
    // an instance is one item of this tree whose leaf record points at the InstanceRec uploaded below
    const size_t n_inst = ctx->h_instances.size();
    std::vector<InstanceRec> irec(n_inst);
    ctx->h_instance_base.assign(n_inst, 0);
    ctx->n_virtual_prims = n;
    if (n_inst || !ctx->objects.empty()) {
        if (ctx->cur_object) { ctx->err = "build_accel: b2rt_object_end is missing"; return B2RT_E_STATE; }
        std::vector<char> used(ctx->objects.size(), 0);
        for (size_t k = 0; k < n_inst; ++k) used[ctx->h_instances[k].object] = 1;
        for (size_t o = 0; o < ctx->objects.size(); ++o) {
            b2rt_ctx *c = ctx->objects[o];
            if (!used[o] || c->built) continue;
            c->sort_variant = ctx->sort_variant;
            int rc = b2rt_build_impl(c);
            if (!rc) rc = b2rt_upload_tables(c);
            if (rc) { ctx->err = "object " + std::to_string(o) + ": " + c->err; return rc; }
        }
        B2RT_CUDA(ctx, ctx->d_objects.reserve(std::max<size_t>(ctx->objects.size(), 1) * sizeof(DScene)));
        B2RT_CUDA(ctx, ctx->d_instances.reserve(std::max<size_t>(n_inst, 1) * sizeof(InstanceRec)));
        for (size_t k = 0; k < n_inst; ++k) {
            const b2rt_ctx::HostInstance &H = ctx->h_instances[k];
            const b2rt_ctx *c = ctx->objects[H.object];
            InstanceRec &I = irec[k];
            memcpy(I.o2w, H.o2w, sizeof(I.o2w)); memcpy(I.w2o, H.w2o, sizeof(I.w2o));
            I.nodes = c->d_wide_nodes.as<WideNode>(); I.recs = c->d_prim_recs.as<PrimRec>(); I.analytic = c->d_analytic.as<AnalyticPrim>();
            I.object = ctx->d_objects.as<DScene>() + H.object;
            I.prim_base = (int32_t)ctx->n_virtual_prims; I.n_prims = c->n_prims; I.object_id = H.object; I.pad = 0;
            memcpy(I.box, c->info.bounds, sizeof(I.box));
            ctx->h_instance_base[k] = I.prim_base;
            ctx->n_virtual_prims += c->n_prims;
            if (ctx->n_virtual_prims >= (1ll << 31)) { ctx->err = "build_accel: more than 2^31 instanced primitives"; return B2RT_E_ARG; }
            kind[H.item] = ~(int32_t)(0x40000000 | (int32_t)k);
        }
        if (n_inst)
            B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_instances.p, irec.data(), n_inst * sizeof(InstanceRec), cudaMemcpyHostToDevice, st));
    }
    const InstanceRec *d_inst = ctx->d_instances.as<InstanceRec>();
    auto up = [&](DevBuf &b, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = b.reserve(bytes ? bytes : 16);
        if (e != cudaSuccess || !bytes) return e;
        return cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st);
    };
    DevBuf &d_atype = ctx->d_queue_b;  // analytic types live in their own small buffer below
    (void)d_atype;
    B2RT_CUDA(ctx, up(ctx->d_verts, ctx->h_verts.data(), ctx->h_verts.size() * sizeof(float)));
    B2RT_CUDA(ctx, up(ctx->d_idx, ctx->h_idx.data(), ctx->h_idx.size() * sizeof(int32_t)));
    B2RT_CUDA(ctx, up(ctx->d_prim_kind, kind.data(), kind.size() * sizeof(int32_t)));
    if (ctx->any_attr) {
        B2RT_CUDA(ctx, up(ctx->d_vnormal, ctx->h_vnormal.data(), ctx->h_vnormal.size() * sizeof(float)));
        B2RT_CUDA(ctx, up(ctx->d_vuv, ctx->h_vuv.data(), ctx->h_vuv.size() * sizeof(float)));
        B2RT_CUDA(ctx, up(ctx->d_tri_flags, ctx->h_tri_flags.data(), ctx->h_tri_flags.size()));
    }
    // analytic prims + their types packed in one buffer: [AnalyticPrim x m][int32 type x m]
    size_t m = ctx->h_analytic.size();
    std::vector<uint8_t> apack(m * sizeof(AnalyticPrim) + m * sizeof(int32_t) + 16);
    if (m) {
        memcpy(apack.data(), ctx->h_analytic.data(), m * sizeof(AnalyticPrim));
        memcpy(apack.data() + m * sizeof(AnalyticPrim), ctx->h_analytic_type.data(), m * sizeof(int32_t));
    }
    B2RT_CUDA(ctx, up(ctx->d_analytic, apack.data(), apack.size()));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));  // host vectors above go out of scope
    const AnalyticPrim *d_an = ctx->d_analytic.as<AnalyticPrim>();
    const int32_t *d_an_type = (const int32_t *)((const uint8_t *)ctx->d_analytic.p + m * sizeof(AnalyticPrim));

    const int nn = n > 0 ? n : 1;
    B2RT_CUDA(ctx, ctx->d_leaf_box.reserve((size_t)nn * 6 * sizeof(float)));
    B2RT_CUDA(ctx, ctx->d_scene_box.reserve(16 * sizeof(uint32_t)));
    B2RT_CUDA(ctx, ctx->d_keys.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_vals.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_keys_alt.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_vals_alt.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_bin_parent.reserve((size_t)(2 * nn) * 4));
    B2RT_CUDA(ctx, ctx->d_bin_child.reserve((size_t)(2 * nn) * 4));
    B2RT_CUDA(ctx, ctx->d_bin_count.reserve((size_t)nn * sizeof(int2)));   // ranges
    B2RT_CUDA(ctx, ctx->d_bin_box.reserve((size_t)(2 * nn) * 6 * sizeof(float)));
    B2RT_CUDA(ctx, ctx->d_bin_cost.reserve((size_t)(2 * nn) * sizeof(float)));
    B2RT_CUDA(ctx, ctx->d_bin_info.reserve((size_t)nn * sizeof(int4)));      // {left, right, area, openable} per internal node
    B2RT_CUDA(ctx, ctx->d_bin_flag.reserve((size_t)nn * 5));               // uint32 flag + uint8 collapse
    B2RT_CUDA(ctx, ctx->d_wide_tmp.reserve((size_t)nn * sizeof(WideTmp)));
    B2RT_CUDA(ctx, ctx->d_wide_flag.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_wide_index.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_queue_a.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_queue_b.reserve((size_t)nn * 4));
    B2RT_CUDA(ctx, ctx->d_counters.reserve(256));
    B2RT_CUDA(ctx, ctx->d_prim_recs.reserve((size_t)nn * sizeof(PrimRec)));
    B2RT_CUDA(ctx, ctx->d_leaf_prims.reserve((size_t)nn * 4));

    B2RT_CUDA(ctx, cudaEventRecord(ctx->evb_mid, st));      // scene resident, buffers reserved: the build proper starts here
    b2rt_accel_info &info = ctx->info;
    info = b2rt_accel_info{};
    info.n_prims = n; info.n_tris = (int)ctx->h_tri_prim.size(); info.wide_width = 4; info.n_leaf_slots = n;

    uint32_t *scene_box = ctx->d_scene_box.as<uint32_t>();
    float *leaf_box = ctx->d_leaf_box.as<float>();
    float *bin_box = ctx->d_bin_box.as<float>();
    int n_wide = 1, depth = 1;

    if (n == 0) {
        B2RT_CUDA(ctx, ctx->d_wide_nodes.reserve(sizeof(WideNode)));
        wide_single_kernel<<<1, 32, 0, st>>>(0, nullptr, ctx->d_wide_nodes.as<WideNode>());
        B2RT_LAUNCH_CHECK(ctx);
    } else {
        const int grid = (n + 255) / 256;
        uint32_t sb[12];
        bool have_sb = false;
        init_scene_box_kernel<<<1, 32, 0, st>>>(scene_box);
        B2RT_LAUNCH_CHECK(ctx);
        prim_bounds_kernel<<<grid, 256, 0, st>>>(ctx->d_verts.as<float>(), ctx->d_idx.as<int32_t>(),
                                                 ctx->d_prim_kind.as<int32_t>(), d_an, d_an_type, d_inst, n, leaf_box,
                                                 scene_box);
        B2RT_LAUNCH_CHECK(ctx);
        morton_kernel<<<grid, 256, 0, st>>>(leaf_box, scene_box, n, ctx->d_keys.as<uint32_t>(),
                                            ctx->d_vals.as<uint32_t>());
        B2RT_LAUNCH_CHECK(ctx);
        bool in_alt = false;
        int rc = b2rt_sort_pairs_impl(ctx, n, ctx->d_keys.as<uint32_t>(), ctx->d_vals.as<uint32_t>(),
                                      ctx->d_keys_alt.as<uint32_t>(), ctx->d_vals_alt.as<uint32_t>(), 30, &in_alt);
        if (rc) return rc;
        if (in_alt) { std::swap(ctx->d_keys, ctx->d_keys_alt); std::swap(ctx->d_vals, ctx->d_vals_alt); }
        uint32_t *keys = ctx->d_keys.as<uint32_t>(), *vals = ctx->d_vals.as<uint32_t>();

        prim_records_kernel<<<grid, 256, 0, st>>>(ctx->d_verts.as<float>(), ctx->d_idx.as<int32_t>(),
                                                  ctx->d_prim_kind.as<int32_t>(), d_an_type, d_inst, vals, n,
                                                  ctx->d_prim_recs.as<PrimRec>(), ctx->d_leaf_prims.as<int32_t>());
        B2RT_LAUNCH_CHECK(ctx);

        int32_t *parent = ctx->d_bin_parent.as<int32_t>(), *child = ctx->d_bin_child.as<int32_t>();
        int2 *range = ctx->d_bin_count.as<int2>();
        uint32_t *flag = ctx->d_bin_flag.as<uint32_t>();
        uint8_t *collapse = (uint8_t *)(flag + nn);
        float *cost = ctx->d_bin_cost.as<float>();
        B2RT_CUDA(ctx, cudaMemsetAsync(flag, 0, (size_t)nn * 5, st));
        if (n == 1) {
            B2RT_CUDA(ctx, cudaMemcpyAsync(bin_box, leaf_box, 6 * sizeof(float), cudaMemcpyDeviceToDevice, st));
            int32_t m1 = -1;
            B2RT_CUDA(ctx, cudaMemcpyAsync(parent, &m1, 4, cudaMemcpyHostToDevice, st));
            B2RT_CUDA(ctx, ctx->d_wide_nodes.reserve(sizeof(WideNode)));
            wide_single_kernel<<<1, 32, 0, st>>>(1, bin_box, ctx->d_wide_nodes.as<WideNode>());
            B2RT_LAUNCH_CHECK(ctx);
        } else {
            karras_topology_kernel<<<(n - 1 + 255) / 256, 256, 0, st>>>(keys, n, parent, child, range);
            B2RT_LAUNCH_CHECK(ctx);
            refit_kernel<<<grid, 256, 0, st>>>(leaf_box, vals, n, parent, child, range, bin_box, cost, flag,
                                               collapse, ctx->d_bin_info.as<int4>());
            B2RT_LAUNCH_CHECK(ctx);
            // greedy collapse to 4-wide nodes: one persistent launch over a device-side queue, the node order from a
            // prefix sum over wide_flag; then ONE host read for everything the host needs (root collapsed?, node count,
            // depth, SAH cost, root box) -- the build used to synchronise ~17 times
            uint32_t *wide_flag = ctx->d_wide_flag.as<uint32_t>();
            uint32_t *wide_index = ctx->d_wide_index.as<uint32_t>();
            uint32_t *counter = ctx->d_counters.as<uint32_t>();
            int32_t *qa = ctx->d_queue_a.as<int32_t>(), *qb = ctx->d_queue_b.as<int32_t>();
            uint32_t *ctl = counter + 40;
            B2RT_CUDA(ctx, cudaMemsetAsync(wide_flag, 0, (size_t)nn * 4, st));
            wide_collapse_init_kernel<<<1, 1, 0, st>>>(qa, ctl, collapse);
            B2RT_LAUNCH_CHECK(ctx);
            {
                if (ctx->collapse_blocks_per_sm == 0) {
                    B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->collapse_blocks_per_sm, wide_collapse_coop_kernel, 256, 0));
                    if (ctx->collapse_blocks_per_sm < 1) ctx->collapse_blocks_per_sm = 1;
                }
                // every block must be resident (grid.sync); more than 2 per SM only adds to the barrier
                int per_sm = std::min(ctx->collapse_blocks_per_sm, 2);
                if (const char *v = getenv("B2RT_COLLAPSE_BLOCKS")) per_sm = std::max(1, std::min(ctx->collapse_blocks_per_sm, atoi(v)));
                int64_t blocks = (int64_t)ctx->sm_count * per_sm;
                const int64_t want = ((int64_t)(n - 1) + 255) / 256;
                if (blocks > want) blocks = want;
                WideTmp *wtmp = ctx->d_wide_tmp.as<WideTmp>();
                int nn_ = n;
                const int4 *info_ = ctx->d_bin_info.as<int4>();
                void *args[] = {&nn_, &info_, &wtmp, &wide_flag, &qa, &qb, &ctl};
                B2RT_CUDA(ctx, cudaLaunchCooperativeKernel((const void *)wide_collapse_coop_kernel, dim3((unsigned)blocks), dim3(256), args, 0, st));
                B2RT_LAUNCH_CHECK(ctx);
            }
            B2RT_CUDA(ctx, cudaMemcpyAsync(wide_index, wide_flag, (size_t)(n - 1) * 4, cudaMemcpyDeviceToDevice, st));
            rc = b2rt_exclusive_scan_u32(ctx, wide_index, n - 1, counter + 4);
            if (rc) return rc;
            uint8_t root_collapsed = 0;
            uint32_t total = 0, deepest = 0;
            float c0 = 0.f, rootbox[6];
            B2RT_CUDA(ctx, cudaMemcpyAsync(&root_collapsed, collapse, 1, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaMemcpyAsync(&total, counter + 4, 4, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaMemcpyAsync(&deepest, ctl + 5, 4, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaMemcpyAsync(&c0, cost, 4, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaMemcpyAsync(rootbox, bin_box, 24, cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaMemcpyAsync(sb, scene_box, sizeof(sb), cudaMemcpyDeviceToHost, st));
            have_sb = true;
            B2RT_CUDA(ctx, cudaStreamSynchronize(st));
            if (root_collapsed) {
                B2RT_CUDA(ctx, ctx->d_wide_nodes.reserve(sizeof(WideNode)));
                wide_single_kernel<<<1, 32, 0, st>>>(n, bin_box, ctx->d_wide_nodes.as<WideNode>());
                B2RT_LAUNCH_CHECK(ctx);
            } else {
                n_wide = (int)total;
                depth = (int)deepest;
                B2RT_CUDA(ctx, ctx->d_wide_nodes.reserve((size_t)n_wide * sizeof(WideNode)));
                wide_emit_kernel<<<(n - 1 + 127) / 128, 128, 0, st>>>(
                    n, ctx->d_wide_tmp.as<WideTmp>(), wide_flag, wide_index, bin_box, range, collapse,
                    ctx->d_wide_nodes.as<WideNode>());
                B2RT_LAUNCH_CHECK(ctx);
            }
            float dx = rootbox[3] - rootbox[0], dy = rootbox[4] - rootbox[1], dz = rootbox[5] - rootbox[2];
            float ra = 2.f * (dx * dy + dy * dz + dz * dx);
            info.sah_cost = ra > 0.f ? c0 / ra : 0.f;
        }
        if (!have_sb) {
            B2RT_CUDA(ctx, cudaMemcpyAsync(sb, scene_box, sizeof(sb), cudaMemcpyDeviceToHost, st));
            B2RT_CUDA(ctx, cudaStreamSynchronize(st));
        }
        for (int a = 0; a < 6; ++a) {
            uint32_t o = sb[6 + a];
            uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
            memcpy(&info.bounds[a], &b, 4);
        }
    }
    info.n_wide_nodes = n_wide;
    info.max_depth = depth;
    B2RT_CUDA(ctx, cudaEventRecord(ctx->evb1, st));
    B2RT_CUDA(ctx, cudaEventSynchronize(ctx->evb1));
    B2RT_CUDA(ctx, cudaEventElapsedTime(&info.upload_ms, ctx->evb0, ctx->evb_mid));
    B2RT_CUDA(ctx, cudaEventElapsedTime(&info.build_ms, ctx->evb_mid, ctx->evb1));
    ctx->built = true;
    return B2RT_OK;
}
