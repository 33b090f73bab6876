// Internal declarations shared by the b2rt translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "b2rt.h"

#define B2RT_SM_COUNT_DEFAULT 148

struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    // grow-only device buffer; returns cudaError
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T *as() const { return (T *)p; }
};

// 128-byte 4-wide BVH node, SoA inside the node so each field is one 16-byte load.
//   lox hix loy hiy loz hiz : 4 children each
//   child[4]  : >=0 internal node index; <0: leaf = 0x80000000 | (first_slot << 3) | (count-1);
//               B2RT_CHILD_EMPTY for an unused lane (its box is inverted so it is never entered)
//   meta[4]   : meta[0] = number of used children, meta[1] = binary-node id it came from
struct __align__(128) WideNode {
    float lox[4], hix[4], loy[4], hiy[4], loz[4], hiz[4];
    int32_t child[4];
    int32_t meta[4];
};
static_assert(sizeof(WideNode) == 128, "wide node must be one cache line");
#define B2RT_CHILD_EMPTY 0x7fffffff
#define B2RT_LEAF_MAX 4          // prims per leaf (pbrt default maxnodeprims 4, accelerators/bvh.cpp:494)
#define B2RT_PRIM_TRI 0
#define B2RT_PRIM_SPHERE 1
#define B2RT_PRIM_DISK 2

// Leaf-ordered primitive record, 48 B = 3 x float4:
//   a = {p1.xyz, bits(prim id)}   b = {e1.xyz, bits(type)}   c = {e2.xyz, bits(analytic index)}
// e1 = p2 - p1, e2 = p3 - p1 are the SAME binary32 values pbrt computes per test
// (pbrt-v2/src/shapes/trianglemesh.cpp:155-156), hoisted to build time.
struct PrimRec { float4 a, b, c; };

// analytic primitives (spheres / disks), world<->object as 3x4 row-major
struct AnalyticPrim {
    float o2w[12], w2o[12];
    float radius, height, inner, pad;
};

struct Material { int32_t type; float kd[3]; float kt[3]; float ior; };
struct Light { int32_t type; float p[3]; float rgb[3]; int32_t prim; int32_t nsamples; };

struct b2rt_ctx {
    int device = 0;
    int sm_count = B2RT_SM_COUNT_DEFAULT;
    uint64_t seed = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evb0 = nullptr, evb1 = nullptr;
    std::string err;
    int64_t launches = 0;

    // host-side staging of the scene (flattened)
    std::vector<float> h_verts;      // xyz
    std::vector<int32_t> h_idx;      // 3 per tri, global vertex indices
    std::vector<int32_t> h_tri_prim; // global prim id of each triangle
    std::vector<AnalyticPrim> h_analytic;
    std::vector<int32_t> h_analytic_prim;  // global prim id
    std::vector<int32_t> h_analytic_type;
    std::vector<int32_t> h_prim_material, h_prim_light;
    std::vector<Material> h_materials;
    std::vector<Light> h_lights;
    int32_t n_prims = 0;

    // device scene + accel
    bool built = false;
    DevBuf d_verts, d_idx, d_analytic, d_prim_kind;   // prim_kind[n_prims]: tri index or ~analytic index
    DevBuf d_leaf_box;      // float[6*n] per-prim AABB (prim id order)
    DevBuf d_scene_box;     // uint32[12]: centroid bounds + scene bounds as ordered ints
    DevBuf d_keys, d_vals, d_keys_alt, d_vals_alt, d_hist;  // sort
    DevBuf d_bin_parent, d_bin_child, d_bin_box, d_bin_flag, d_bin_cost, d_bin_count;
    DevBuf d_wide_tmp, d_wide_flag, d_wide_index, d_queue_a, d_queue_b, d_counters;
    DevBuf d_wide_nodes, d_prim_recs, d_leaf_prims;
    DevBuf d_rays, d_hits, d_occ;  // staging for host-pointer calls
    DevBuf d_scan_tmp;
    b2rt_accel_info info{};
};

#define B2RT_CUDA(ctx, call)                                                              \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);             \
            return B2RT_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

#define B2RT_LAUNCH_CHECK(ctx)                                                            \
    do {                                                                                  \
        (ctx)->launches++;                                                                \
        cudaError_t e__ = cudaGetLastError();                                             \
        if (e__ != cudaSuccess) {                                                         \
            (ctx)->err = std::string("kernel launch: ") + cudaGetErrorString(e__);        \
            return B2RT_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

// ---- stage entry points (host functions, each enqueues on ctx->stream) -----------------------
int b2rt_sort_pairs_impl(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt,
                         uint32_t *vals_alt, int key_bits, bool *result_in_alt);
int b2rt_build_impl(b2rt_ctx *ctx);
int b2rt_trace_impl(b2rt_ctx *ctx, int64_t n, const b2rt_ray *d_rays, b2rt_hit *d_hits, uint8_t *d_occ,
                    bool any_hit);
// exclusive scan of n uint32 in place (single pass, small n helper used by build)
int b2rt_exclusive_scan_u32(b2rt_ctx *ctx, uint32_t *d, int64_t n, uint32_t *d_total);
