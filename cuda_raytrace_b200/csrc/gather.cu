// K12: radiance estimate over the photon hash grid -- k-nearest and fixed-radius modes.
//
// Replaces KdTreeLookup / processPhoton / photonGatheringCamera (cuda_render/photon_mapping/gathering.cu:11-126),
// which walk a CPU-built kd-tree per thread with a private stack.  The estimator is pbrt's LPhoton
// (pbrt-v2/src/integrators/photonmap.cpp:207-260) for a diffuse surface, not the reference's unweighted PPM sum:
//   maxDist2 = d2 of the k-th nearest photon if >= k lie within the search radius, else the radius itself
//   L = INV_PI * Kd * sum_{i in set, Dot(Nf, wi_i) > 0}  (3*INV_PI*(1 - d2_i/maxDist2)^2 / (nPaths*maxDist2)) * alpha_i
// The set is the k smallest (d2, index-in-sorted-array) pairs with d2 < radius2 -- exactly what pbrt's heap keeps
// (photonmap.cpp:187-204) up to ties in d2.
//
// One warp per query.  Shells of cells around the query's cell are scanned outwards; within a shell each row of
// x-adjacent cells is one contiguous range of the sorted photon array, read with coalesced 16-byte loads, lanes
// testing 32 photons at a time.  Candidates (64-bit key = d2 bits << 32 | index) are appended to a per-warp
// shared-memory buffer by ballot compaction; when it fills, a warp-cooperative bitonic sort keeps the k smallest and
// tightens the bound.  The scan stops after the first shell whose outer boundary is farther than the bound.
#include "b2rt_shade.cuh"

namespace {

#ifndef B2RT_GATHER_BLOCKS
#define B2RT_GATHER_BLOCKS 7
#endif
constexpr int GATHER_WARPS = 4;
constexpr int GATHER_CAP = 256;          // candidate buffer entries per warp (k <= GATHER_CAP - 32)
constexpr int GATHER_PER_LANE = GATHER_CAP / 32;
// candidate slot of element e: XOR swizzle inside aligned groups of 8 so that the sort's lane-strided 64-bit
// accesses (element lane*PL + j) spread over all banks; consecutive elements still map to distinct slots
#define CI(e) ((e) ^ (((e) >> 3) & 7))

struct GatherGrid {
    float lo[3]; float cell, inv, slack;
    int dims[3];
};

__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int m) {
    uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m), hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}

// ascending bitonic sort of 32*PL keys; element e lives in lane e / PL, register e % PL.
// The stages that stay inside a lane (stride < PL) are unrolled; the cross-lane stages run as ROLLED loops over
// (size, lane mask) so the whole network is a few hundred instructions and stays resident in the instruction cache
// (the fully unrolled 256-key network is ~37 KB of straight-line code: ncu showed the kernel stalled on
// instruction fetch, "no_instruction", for 10 warps per issue).
// `up` differs between lanes: the swap decision is formed with logic ops so it compiles to predicates + selects, not to
// a divergent branch (ncu: the branchy form ran with 13 of 32 threads per instruction)
__device__ __forceinline__ void key_ce(uint64_t &a, uint64_t &b, bool up) {
    // one 64-bit comparison: ascending swaps when b < a, descending when !(b < a); the latter also swaps equal keys,
    // which changes nothing
    const bool sw = (b < a) == up;
    const uint64_t lo = sw ? b : a, hi = sw ? a : b;
    a = lo; b = hi;
}
template <int PL>
__device__ __forceinline__ void warp_sort(uint64_t (&v)[PL], int lane) {
#pragma unroll
    for (int size = 2; size <= PL; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
            for (int j = 0; j < PL; ++j)
                if ((j & stride) == 0) key_ce(v[j], v[j + stride], ((lane * PL + j) & size) == 0);
        }
    }
#pragma unroll 1
    for (int size = 2 * PL; size <= 32 * PL; size <<= 1) {
        const bool up = ((lane * PL) & size) == 0;
#pragma unroll 1
        for (int lm = size / (2 * PL); lm > 0; lm >>= 1) {
            const bool take_min = (up == ((lane & lm) == 0));
#pragma unroll
            for (int j = 0; j < PL; ++j) {
                const uint64_t o = shfl_xor_u64(v[j], lm);
                v[j] = ((o < v[j]) == take_min) ? o : v[j];      // min or max of the pair with one comparison
            }
        }
#pragma unroll
        for (int stride = PL >> 1; stride > 0; stride >>= 1) {
#pragma unroll
            for (int j = 0; j < PL; ++j)
                if ((j & stride) == 0) key_ce(v[j], v[j + stride], up);
        }
    }
}

template <int PL>
__device__ __noinline__ void warp_sort_buffer(uint64_t *cand, int count, int lane) {
    uint64_t v[PL];
#pragma unroll
    for (int j = 0; j < PL; ++j) {
        int e = lane * PL + j;
        v[j] = e < count ? cand[CI(e)] : 0xffffffffffffffffull;
    }
    warp_sort<PL>(v, lane);
#pragma unroll
    for (int j = 0; j < PL; ++j) cand[CI(lane * PL + j)] = v[j];
    __syncwarp();
}

// keep the k smallest of cand[0..count): sorts the buffer (network sized to the next of 64 / 128 / 256 entries),
// returns min(count, k); cand[0..min(count,k)) is ascending afterwards
__device__ __forceinline__ int warp_select(uint64_t *cand, int count, int k, int lane) {
    if (count <= 64) warp_sort_buffer<2>(cand, count, lane);
    else if (count <= 128) warp_sort_buffer<4>(cand, count, lane);
    else warp_sort_buffer<GATHER_PER_LANE>(cand, count, lane);
    return count < k ? count : k;
}

// Threshold selection: keep the k smallest keys of cand[0..count), count > k, IN BUFFER ORDER (no sort).
// The d2 word T of the k-th smallest key is bracketed in the bit domain: a probe t costs one compare per held key and
// one warp-wide REDUX; as soon as a probe has exactly k - 1 (or k) keys below it, T is the smallest key >= t (the
// largest < t).  Probes alternate between interpolation on the counts (photons on a surface: the count grows linearly
// in d2) and the midpoint of the bracket (bounds the worst case at 64 probes); typically 3-6 probes.  The first version
// bisected bit by bit: ~27 dependent probes, slower than the sort network it replaced.  Ties in d2 at the threshold
// are resolved by the index word (rare), and the survivors are ballot-compacted to the front of the buffer (exactly k
// of them).  Returns the largest kept key.  Element e lives in lane e % 32, register e / 32 (conflict-free with the
// slot swizzle).
template <int PL>
__device__ __noinline__ uint64_t warp_keep_k(uint64_t *cand, int count, int k, int lane) {
    const uint32_t FULL = 0xffffffffu, lt = (1u << lane) - 1u;
    uint32_t d[PL], dmax = 0u;
#pragma unroll
    for (int j = 0; j < PL; ++j) {
        const int e = j * 32 + lane;
        d[j] = FULL;
        if (e < count) { d[j] = (uint32_t)(cand[CI(e)] >> 32); dmax = max(dmax, d[j]); }
    }
    // invariant: #(d < lo) = clo <= k - 1 and #(d < hi) = chi >= k, so lo <= T < hi
    uint32_t lo = 0u, hi = __reduce_max_sync(FULL, dmax) + 1u, clo = 0u, chi = (uint32_t)count, T;
    for (int it = 0;; ++it) {
        if (hi - lo <= 1u) { T = lo; break; }
        uint32_t t = lo + ((hi - lo) >> 1);
        if (!(it & 1)) {
            const float flo = __uint_as_float(lo), fhi = __uint_as_float(hi);
            const float f = flo + (fhi - flo) * (((float)(k - (int)clo) - 0.5f) / (float)(chi - clo));
            t = min(max(__float_as_uint(f), lo + 1u), hi - 1u);
        }
        uint32_t c = 0;
#pragma unroll
        for (int j = 0; j < PL; ++j) c += d[j] < t ? 1u : 0u;
        c = __reduce_add_sync(FULL, c);
        if ((int)c == k - 1) {            // T = the smallest key at or above the probe
            uint32_t m = FULL;
#pragma unroll
            for (int j = 0; j < PL; ++j) m = min(m, d[j] >= t ? d[j] : FULL);
            T = __reduce_min_sync(FULL, m);
            break;
        }
        if ((int)c == k) {                // T = the largest key below the probe
            uint32_t m = 0u;
#pragma unroll
            for (int j = 0; j < PL; ++j) m = max(m, d[j] < t ? d[j] : 0u);
            T = __reduce_max_sync(FULL, m);
            break;
        }
        if ((int)c < k) { lo = t; clo = c; } else { hi = t; chi = c; }
    }
    uint32_t c2 = 0, mx = 0;
#pragma unroll
    for (int j = 0; j < PL; ++j) {
        c2 += (d[j] < T ? 1u : 0u) + (d[j] == T ? 0x10000u : 0u);
        if (d[j] == T) mx = max(mx, (uint32_t)cand[CI(j * 32 + lane)]);
    }
    c2 = __reduce_add_sync(FULL, c2);
    const int need = k - (int)(c2 & 0xffffu), ce = (int)(c2 >> 16);      // 1 <= need <= ce
    uint32_t I;
    if (ce == need) {
        I = __reduce_max_sync(FULL, mx);
    } else {                              // more keys tie at the threshold than fit: the `need` smallest indices
        I = 0u;
        for (uint32_t bit = 0x80000000u; bit; bit >>= 1) {
            const uint32_t t = I | bit;
            uint32_t c = 0;
#pragma unroll
            for (int j = 0; j < PL; ++j)
                if (d[j] == T) c += (uint32_t)cand[CI(j * 32 + lane)] < t ? 1u : 0u;
            c = __reduce_add_sync(FULL, c);
            if ((int)c < need) I = t;
        }
    }
    const uint64_t kth = ((uint64_t)T << 32) | I;
    int out = 0;
#pragma unroll
    for (int j = 0; j < PL; ++j) {        // writes land at or before the slots this round has already read
        const int e = j * 32 + lane;
        const uint64_t key = e < count ? cand[CI(e)] : 0xffffffffffffffffull;
        const bool keep = key <= kth;
        const uint32_t m = __ballot_sync(FULL, keep);
        if (keep) cand[CI(out + __popc(m & lt))] = key;
        out += __popc(m);
    }
    __syncwarp();
    return kth;
}
__device__ __forceinline__ uint64_t warp_keep(uint64_t *cand, int count, int k, int lane) {
    if (count <= 64) return warp_keep_k<2>(cand, count, k, lane);
    if (count <= 128) return warp_keep_k<4>(cand, count, k, lane);
    return warp_keep_k<GATHER_PER_LANE>(cand, count, k, lane);
}
// reorder cand[0..count) by ascending photon index (the summation order of the estimator); used where the candidates
// were not collected in cell order
__device__ __forceinline__ void warp_sort_by_index(uint64_t *cand, int count, int lane) {
    for (int e = lane; e < count; e += 32) { const uint64_t v = cand[CI(e)]; cand[CI(e)] = (v << 32) | (v >> 32); }
    __syncwarp();
    warp_select(cand, count, count, lane);
    for (int e = lane; e < count; e += 32) { const uint64_t v = cand[CI(e)]; cand[CI(e)] = (v << 32) | (v >> 32); }
    __syncwarp();
}

__device__ __forceinline__ f3 oct_decode(uint32_t w) {
    float x = (float)(int)(int16_t)(w & 0xffffu) / 32767.f, y = (float)(int)(int16_t)(w >> 16) / 32767.f;
    float z = 1.f - fabsf(x) - fabsf(y);
    if (z < 0.f) {
        float ox = (1.f - fabsf(y)) * (x >= 0.f ? 1.f : -1.f);
        float oy = (1.f - fabsf(x)) * (y >= 0.f ? 1.f : -1.f);
        x = ox; y = oy;
    }
    return normalize3(mk3(x, y, z));
}

// KNN = true: k nearest within radius;  false: every photon within the radius (no candidate buffer)
template <bool KNN>
__global__ void __launch_bounds__(GATHER_WARPS * 32)
gather_kernel(const b2rt_photon *__restrict__ photons, const uint32_t *__restrict__ cell_start, GatherGrid g,
              const b2rt_query *__restrict__ queries, int64_t nq, int k, float max_dist2, float n_paths,
              float *__restrict__ out_rgb, float *__restrict__ out_r2, int32_t *__restrict__ out_found,
              int32_t *__restrict__ idx_out) {
    __shared__ uint64_t s_cand[GATHER_WARPS][GATHER_CAP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint64_t *cand = s_cand[warp];
    const uint32_t lt = (1u << lane) - 1u;
    for (int64_t q = (int64_t)blockIdx.x * GATHER_WARPS + warp; q < nq; q += (int64_t)gridDim.x * GATHER_WARPS) {
        const float4 *qp = reinterpret_cast<const float4 *>(queries + q);
        float4 q0 = __ldg(qp), q1 = __ldg(qp + 1), q2 = __ldg(qp + 2);
        const float px = q0.x, py = q0.y, pz = q0.z;
        const f3 nf = mk3(q0.w, q1.x, q1.y), kd = mk3(q1.z, q1.w, q2.x);
        int c[3];
        c[0] = min(max((int)floorf((px - g.lo[0]) * g.inv), 0), g.dims[0] - 1);
        c[1] = min(max((int)floorf((py - g.lo[1]) * g.inv), 0), g.dims[1] - 1);
        c[2] = min(max((int)floorf((pz - g.lo[2]) * g.inv), 0), g.dims[2] - 1);
        const float qa[3] = {px, py, pz};
        uint64_t bound = (uint64_t)__float_as_uint(max_dist2) << 32;   // accept keys < bound  <=>  d2 < radius2
        int count = 0;
        f3 acc = mk3(0.f, 0.f, 0.f);       // fixed-radius mode accumulates directly
        int found_fr = 0;
        const int smax = max(max(g.dims[0], g.dims[1]), g.dims[2]);
        for (int s = 0; s < smax; ++s) {
            const int z0 = max(c[2] - s, 0), z1 = min(c[2] + s, g.dims[2] - 1);
            const int y0 = max(c[1] - s, 0), y1 = min(c[1] + s, g.dims[1] - 1);
            const int x0 = c[0] - s, x1 = c[0] + s;
            for (int z = z0; z <= z1; ++z) {
                for (int y = y0; y <= y1; ++y) {
                    const bool face = (abs(z - c[2]) == s) || (abs(y - c[1]) == s);
                    const int row = (z * g.dims[1] + y) * g.dims[0];
                    // face rows: the whole x run; interior rows: the two end cells only
                    for (int part = 0; part < (face ? 1 : 2); ++part) {
                        int xa, xb;
                        if (face) { xa = max(x0, 0); xb = min(x1, g.dims[0] - 1); }
                        else if (part == 0) { xa = xb = x0; if (x0 < 0) continue; }
                        else { xa = xb = x1; if (x1 > g.dims[0] - 1 || s == 0) continue; }
                        const uint32_t b = __ldg(cell_start + row + xa), e = __ldg(cell_start + row + xb + 1);
                        for (uint32_t i0 = b; i0 < e; i0 += 32) {
                            const uint32_t i = i0 + lane;
                            bool pass = false;
                            uint64_t key = 0;
                            float4 pa;
                            if (i < e) {
                                pa = __ldg(reinterpret_cast<const float4 *>(photons + i));
                                float dx = pa.x - px, dy = pa.y - py, dz = pa.z - pz;
                                float d2 = dx * dx + dy * dy + dz * dz;
                                key = ((uint64_t)__float_as_uint(d2) << 32) | i;
                                pass = key < bound;
                            }
                            if (KNN) {
                                uint32_t m = __ballot_sync(0xffffffffu, pass);
                                if (m) {
                                    if (pass) cand[CI(count + __popc(m & lt))] = key;
                                    count += __popc(m);
                                    __syncwarp();
                                    if (count > GATHER_CAP - 32) {
                                        count = warp_select(cand, count, k, lane);
                                        if (count == k) bound = cand[CI(k - 1)];       // only keys below the current k-th
                                    }
                                }
                            } else if (pass) {
                                float d2 = __uint_as_float((uint32_t)(key >> 32));
                                f3 wi = oct_decode(__float_as_uint(pa.w));
                                if (dot3(nf, wi) > 0.f) {
                                    float4 pb = __ldg(reinterpret_cast<const float4 *>(photons + i) + 1);
                                    float sk = 1.f - d2 / max_dist2;
                                    float w = 3.f * B2_INV_PI * sk * sk / (n_paths * max_dist2);
                                    acc = acc + mk3(w * pb.x, w * pb.y, w * pb.z);
                                }
                                ++found_fr;
                            }
                        }
                    }
                }
            }
            // distance from the query to the outside of the scanned block; sides already at the grid edge do not count
            float dmin = INFINITY;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                if (c[a] - s > 0) dmin = fminf(dmin, qa[a] - (g.lo[a] + (float)(c[a] - s) * g.cell));
                if (c[a] + s < g.dims[a] - 1) dmin = fminf(dmin, (g.lo[a] + (float)(c[a] + s + 1) * g.cell) - qa[a]);
            }
            if (dmin == INFINITY) break;                 // whole grid scanned
            dmin -= g.slack;
            float r2 = __uint_as_float((uint32_t)(bound >> 32));
            if (KNN && count > k) {                      // tighten before deciding
                count = warp_select(cand, count, k, lane);
                if (count == k) bound = cand[CI(k - 1)];
                r2 = __uint_as_float((uint32_t)(bound >> 32));
            }
            if (dmin > 0.f && dmin * dmin >= r2) break;
        }
        if (KNN) {
            count = warp_select(cand, count, k, lane);
            const float md2 = (count == k) ? __uint_as_float((uint32_t)(cand[CI(k - 1)] >> 32)) : max_dist2;
            if (idx_out)
                for (int j = lane; j < k; j += 32) idx_out[q * k + j] = j < count ? (int32_t)(uint32_t)cand[CI(j)] : -1;
            __syncwarp();
            warp_sort_by_index(cand, count, lane);
            for (int j = lane; j < count; j += 32) {
                uint64_t key = cand[CI(j)];
                uint32_t i = (uint32_t)key;
                float d2 = __uint_as_float((uint32_t)(key >> 32));
                const float4 *pp = reinterpret_cast<const float4 *>(photons + i);
                float4 pa = __ldg(pp), pb = __ldg(pp + 1);
                f3 wi = oct_decode(__float_as_uint(pa.w));
                if (dot3(nf, wi) > 0.f) {
                    float sk = 1.f - d2 / md2;
                    float w = 3.f * B2_INV_PI * sk * sk / (n_paths * md2);
                    acc = acc + mk3(w * pb.x, w * pb.y, w * pb.z);
                }
            }
            if (lane == 0) {
                if (out_r2) out_r2[q] = md2;
                if (out_found) out_found[q] = count;
            }
            __syncwarp();
        } else {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) found_fr += __shfl_xor_sync(0xffffffffu, found_fr, o);
            if (lane == 0) {
                if (out_r2) out_r2[q] = max_dist2;
                if (out_found) out_found[q] = found_fr;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc.x = acc.x + __shfl_xor_sync(0xffffffffu, acc.x, o);
            acc.y = acc.y + __shfl_xor_sync(0xffffffffu, acc.y, o);
            acc.z = acc.z + __shfl_xor_sync(0xffffffffu, acc.z, o);
        }
        if (lane == 0 && out_rgb) {
            out_rgb[3 * q + 0] = acc.x * kd.x * B2_INV_PI;
            out_rgb[3 * q + 1] = acc.y * kd.y * B2_INV_PI;
            out_rgb[3 * q + 2] = acc.z * kd.z * B2_INV_PI;
        }
    }
}


// ---- tiled variant (default) ---------------------------------------------------------------------------------------
// Two-level scan with TMA-staged photon tiles.  The fine grid keeps its bit-exact definition; on top of it sits a
// coarse table (G2_C^3 fine cells per coarse cell, photon count each) built with the grid.
//   * shells of COARSE cells around the query; the 32 lanes look up 32 coarse cells at once (count > 0 and box
//     closer than the current bound), so empty space costs one 4-byte load per 64 fine cells;
//   * a surviving coarse cell is 16 fine rows (z, y) of 4 x-adjacent cells = 16 contiguous runs of the sorted
//     position array; 16 lanes fetch the run bounds at once (two coarse cells per round);
//   * the runs are copied into the warp's shared-memory tile by the TMA: each lane issues one
//     cp.async.bulk.shared.global for its run at the lane's prefix offset, all completing on the warp's mbarrier.
//     The tile comes out dense, so the distance tests run on full 32-lane batches with conflict-free LDS.128;
//   * positions are stored as {x, y, z, bits(index)}: the candidate key (d2 << 32 | index) needs no index
//     arithmetic; alpha / wi are only touched for the k selected photons.
// The selected set is the k smallest (d2, index) keys below the radius -- the same definition as the simple
// kernel and the brute-force oracle, so results are bit-identical whatever the scan order.
constexpr int G2_C = 4, G2_SHIFT = 2;
constexpr int G2_STAGE = 256;            // staged positions per warp (4 KB: six blocks per SM fit)

struct Gather2Grid {
    float lo[3]; float cell, inv, slack;
    int dims[3]; int cdims[3];
};

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// squared distance from q to the box [lo, hi] per axis; sides flagged open extend to infinity (edge cells also hold
// the photons that were clamped into them)
__device__ __forceinline__ float axis_gap(float q, float lo, float hi, bool open_lo, bool open_hi, float slack) {
    float d = 0.f;
    if (!open_lo) d = fmaxf(d, lo - q - slack);
    if (!open_hi) d = fmaxf(d, q - hi - slack);
    return d;
}

struct GatherArgs {
    const float4 *pos4; const b2rt_photon *photons; const uint32_t *cell_start; const uint32_t *coarse;
    Gather2Grid g;
    const b2rt_query *queries; int64_t nq; int k; float max_dist2, n_paths, target;
    float *out_rgb, *out_r2; int32_t *out_found, *idx_out;
};

// final selection, estimator sum and output of one query; cand[0..count) holds every candidate below the bound
// `ordered`: the buffer is in ascending photon index (candidates collected in cell order)
template <bool KNN>
__device__ __forceinline__ void gather_finish(const GatherArgs &A, int64_t q, f3 nf, f3 kd, uint64_t *cand, int count,
                                              f3 acc, int found_fr, int lane, bool ordered) {
    const b2rt_photon *__restrict__ photons = A.photons;
    const int k = A.k;
    const float max_dist2 = A.max_dist2, n_paths = A.n_paths;
    float *__restrict__ out_rgb = A.out_rgb, *__restrict__ out_r2 = A.out_r2;
    int32_t *__restrict__ out_found = A.out_found, *__restrict__ idx_out = A.idx_out;
    {
        if (KNN) {
            float md2 = max_dist2;
            if (idx_out) {                               // parity hook: the set ascending by (d2, index)
                if (count > 0) count = warp_select(cand, count, k, lane);
                if (count == k) md2 = __uint_as_float((uint32_t)(cand[CI(k - 1)] >> 32));
                for (int j = lane; j < k; j += 32) idx_out[q * k + j] = j < count ? (int32_t)(uint32_t)cand[CI(j)] : -1;
                __syncwarp();
                ordered = false;
            } else if (count > k) {
                md2 = __uint_as_float((uint32_t)(warp_keep(cand, count, k, lane) >> 32));
                count = k;
            } else if (count == k) {
                uint32_t mx = 0u;
                for (int j = lane; j < count; j += 32) mx = max(mx, (uint32_t)(cand[CI(j)] >> 32));
                md2 = __uint_as_float(__reduce_max_sync(0xffffffffu, mx));
            }
            if (!ordered && count > 1) warp_sort_by_index(cand, count, lane);
            for (int j = lane; j < count; j += 32) {
                const uint64_t key = cand[CI(j)];
                const uint32_t i = (uint32_t)key;
                const float d2 = __uint_as_float((uint32_t)(key >> 32));
                const float4 *pp = reinterpret_cast<const float4 *>(photons + i);
                const float4 pa = __ldg(pp), pb = __ldg(pp + 1);
                const f3 wi = oct_decode(__float_as_uint(pa.w));
                if (dot3(nf, wi) > 0.f) {
                    const float sk = 1.f - d2 / md2;
                    const float wgt = 3.f * B2_INV_PI * sk * sk / (n_paths * md2);
                    acc = acc + mk3(wgt * pb.x, wgt * pb.y, wgt * pb.z);
                }
            }
            if (lane == 0) {
                if (out_r2) out_r2[q] = md2;
                if (out_found) out_found[q] = count;
            }
            __syncwarp();
        } else {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) found_fr += __shfl_xor_sync(0xffffffffu, found_fr, o);
            if (lane == 0) {
                if (out_r2) out_r2[q] = max_dist2;
                if (out_found) out_found[q] = found_fr;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc.x = acc.x + __shfl_xor_sync(0xffffffffu, acc.x, o);
            acc.y = acc.y + __shfl_xor_sync(0xffffffffu, acc.y, o);
            acc.z = acc.z + __shfl_xor_sync(0xffffffffu, acc.z, o);
        }
        if (lane == 0 && out_rgb) {
            out_rgb[3 * q + 0] = acc.x * kd.x * B2_INV_PI;
            out_rgb[3 * q + 1] = acc.y * kd.y * B2_INV_PI;
            out_rgb[3 * q + 2] = acc.z * kd.z * B2_INV_PI;
        }
    }
}

// one query, one warp: coarse shells, TMA-staged rows, candidate buffer (the scan described above)
template <bool KNN>
__device__ __forceinline__ void gather_query_warp(const GatherArgs &A, int64_t q, float r2_first, uint64_t *cand,
                                                  const float4 *stage, uint32_t stage_a, uint32_t bar, uint32_t &phase,
                                                  int lane) {
    const float4 *__restrict__ pos4 = A.pos4;
    const b2rt_photon *__restrict__ photons = A.photons;
    const uint32_t *__restrict__ cell_start = A.cell_start, *__restrict__ coarse = A.coarse;
    const Gather2Grid &g = A.g;
    const b2rt_query *__restrict__ queries = A.queries;
    const int k = A.k;
    const float max_dist2 = A.max_dist2, n_paths = A.n_paths;
    const uint32_t lt = (1u << lane) - 1u;
    {
        const float4 *qp = reinterpret_cast<const float4 *>(queries + q);
        const float4 q0 = __ldg(qp), q1 = __ldg(qp + 1), q2 = __ldg(qp + 2);
        const float qa[3] = {q0.x, q0.y, q0.z};
        const f3 nf = mk3(q0.w, q1.x, q1.y), kd = mk3(q1.z, q1.w, q2.x);
        int C[3], cf[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            cf[a] = min(max((int)floorf((qa[a] - g.lo[a]) * g.inv), 0), g.dims[a] - 1);
            C[a] = cf[a] >> G2_SHIFT;
        }
        int smax = 0;
#pragma unroll
        for (int a = 0; a < 3; ++a) smax = max(smax, max(C[a], g.cdims[a] - 1 - C[a]));
        // First search radius from the local photon density (performance only: the result is defined by the k smallest
        // keys below max_dist2 and does not depend on it).  N8 = photons in the 2x2x2 coarse cells nearest to the query,
        // taken as lying on a surface patch of (8 cells)^2; radius expected to hold 1.6 k photons (<= 128 candidates: the small sort network).  If fewer than k are
        // found the search is repeated with 4x the squared radius.
        float r2_try = max_dist2;
        if (KNN) {
            uint32_t n8 = 0;
            if (lane < 8) {
                const int o[3] = {lane & 1, (lane >> 1) & 1, lane >> 2};
                int P[3];
                bool in = true;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    P[a] = C[a] - ((cf[a] & (G2_C - 1)) < G2_C / 2 ? 1 : 0) + o[a];
                    in = in && P[a] >= 0 && P[a] < g.cdims[a];
                }
                if (in) n8 = __ldg(coarse + ((int64_t)P[2] * g.cdims[1] + P[1]) * g.cdims[0] + P[0]);
            }
            n8 = __reduce_add_sync(0xffffffffu, n8);
            if (n8 > 0) {
                const float est = (A.target * 64.f * B2_INV_PI) * (float)k * g.cell * g.cell / (float)n8;
                r2_try = fminf(est, max_dist2);
            }
        }
        if (KNN && r2_first > 0.f) r2_try = fminf(r2_first, max_dist2);     // caller already knows a radius that was too small
        uint64_t bound;
        int count;
        f3 acc = mk3(0.f, 0.f, 0.f);
        int found_fr = 0;
      for (;;) {
        bound = (uint64_t)__float_as_uint(r2_try) << 32;
        count = 0;
        // The first round covers the query's coarse cell and its 26 neighbours at once (shells 0 and 1: where almost every
        // query ends), decoded with constant divisors; from S = 2 on, one shell surface per iteration.
        for (int S = 1; S <= smax || S == 1; ++S) {
            const int w = 2 * S + 1, n_face = w * w;
            const int ntask = S == 1 ? 27 : 2 * n_face + (2 * S - 1) * 8 * S;
            for (int t0 = 0; t0 < ntask; t0 += 32) {
                const int t = t0 + lane;
                bool ok = t < ntask;
                int ox = 0, oy = 0, oz = 0;
                if (S == 1) {
                    ox = t % 3 - 1; oy = (t / 3) % 3 - 1; oz = t / 9 - 1;
                } else if (ok) {
                    if (t < 2 * n_face) {
                        const int u = t < n_face ? t : t - n_face;
                        oz = t < n_face ? -S : S;
                        oy = u / w - S; ox = u % w - S;
                    } else {
                        const int tt = t - 2 * n_face, per = 8 * S;
                        const int layer = tt / per, r = tt % per, side = r / (2 * S), o = r % (2 * S);
                        oz = layer - (S - 1);
                        if (side == 0) { ox = -S + o; oy = -S; }
                        else if (side == 1) { ox = S; oy = -S + o; }
                        else if (side == 2) { ox = S - o; oy = S; }
                        else { ox = -S; oy = S - o; }
                    }
                }
                const int X = C[0] + ox, Y = C[1] + oy, Z = C[2] + oz;
                ok = ok && X >= 0 && Y >= 0 && Z >= 0 && X < g.cdims[0] && Y < g.cdims[1] && Z < g.cdims[2];
                if (ok) {
                    const float r2 = __uint_as_float((uint32_t)(bound >> 32));
                    const int P[3] = {X, Y, Z};
                    float d2 = 0.f;
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int f0 = P[a] << G2_SHIFT, f1 = min(f0 + G2_C, g.dims[a]);
                        const float gap = axis_gap(qa[a], g.lo[a] + (float)f0 * g.cell, g.lo[a] + (float)f1 * g.cell,
                                                   f0 == 0, f1 == g.dims[a], g.slack);
                        d2 += gap * gap;
                    }
                    ok = d2 < r2 && __ldg(coarse + ((int64_t)Z * g.cdims[1] + Y) * g.cdims[0] + X) != 0u;
                }
                uint32_t mask = __ballot_sync(0xffffffffu, ok);
                const int packed = X | (Y << 10) | (Z << 20);
                while (mask) {
                    // two coarse cells per round: lanes 0-15 take the rows of the first, lanes 16-31 of the second
                    const int l0 = __ffs(mask) - 1;
                    mask &= mask - 1;
                    int l1 = -1;
                    if (mask) { l1 = __ffs(mask) - 1; mask &= mask - 1; }
                    const int src = lane < 16 ? l0 : l1;
                    const int pk = __shfl_sync(0xffffffffu, packed, src < 0 ? 0 : src);
                    const int cx = pk & 1023, cy = (pk >> 10) & 1023, cz = (pk >> 20) & 1023;
                    const int row = lane & 15;
                    const int z = (cz << G2_SHIFT) + (row >> 2), y = (cy << G2_SHIFT) + (row & 3);
                    const int xa = cx << G2_SHIFT, xb = min(xa + G2_C, g.dims[0]);     // cells [xa, xb)
                    bool valid = src >= 0 && z < g.dims[2] && y < g.dims[1];
                    if (valid) {
                        const float r2 = __uint_as_float((uint32_t)(bound >> 32));
                        float gx = axis_gap(qa[0], g.lo[0] + (float)xa * g.cell, g.lo[0] + (float)xb * g.cell, xa == 0,
                                            xb == g.dims[0], g.slack);
                        float gy = axis_gap(qa[1], g.lo[1] + (float)y * g.cell, g.lo[1] + (float)(y + 1) * g.cell, y == 0,
                                            y + 1 == g.dims[1], g.slack);
                        float gz = axis_gap(qa[2], g.lo[2] + (float)z * g.cell, g.lo[2] + (float)(z + 1) * g.cell, z == 0,
                                            z + 1 == g.dims[2], g.slack);
                        valid = gx * gx + gy * gy + gz * gz < r2;
                    }
                    uint32_t b = 0, e = 0;
                    if (valid) {
                        const int64_t rb = ((int64_t)z * g.dims[1] + y) * g.dims[0];
                        b = __ldg(cell_start + rb + xa);
                        e = __ldg(cell_start + rb + xb);
                    }
                    // ---- stage the runs [b, e) of all lanes into the warp's tile, G2_STAGE positions at a time ----
                    while (true) {
                        const int n = (int)(e - b);
                        int inc = n;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int up = __shfl_up_sync(0xffffffffu, inc, o);
                            if (lane >= o) inc += up;
                        }
                        const int all = __shfl_sync(0xffffffffu, inc, 31);
                        if (all == 0) break;
                        const int off = inc - n;
                        const int take = off < G2_STAGE ? min(n, G2_STAGE - off) : 0;
                        const int total = min(all, G2_STAGE);
                        fence_proxy_async_smem();          // earlier generic reads of the tile before the async writes
                        __syncwarp();
                        if (lane == 0) mbar_expect_tx(bar, (uint32_t)total * 16u);
                        __syncwarp();
                        if (take > 0) tma_bulk_g2s(stage_a + (uint32_t)off * 16u, pos4 + b, (uint32_t)take * 16u, bar);
                        b += (uint32_t)take;
                        mbar_wait(bar, phase);
                        phase ^= 1u;
                        for (int i0 = 0; i0 < total; i0 += 32) {
                            const int i = i0 + lane;
                            bool pass = false;
                            uint64_t key = 0;
                            if (i < total) {
                                const float4 pa = stage[i];
                                const float dx = pa.x - qa[0], dy = pa.y - qa[1], dz = pa.z - qa[2];
                                const float d2 = dx * dx + dy * dy + dz * dz;
                                key = ((uint64_t)__float_as_uint(d2) << 32) | __float_as_uint(pa.w);
                                pass = key < bound;
                            }
                            if (KNN) {
                                const uint32_t m = __ballot_sync(0xffffffffu, pass);
                                if (m) {
                                    if (pass) cand[CI(count + __popc(m & lt))] = key;
                                    count += __popc(m);
                                    __syncwarp();
                                    if (count > GATHER_CAP - 32) { bound = warp_keep(cand, count, k, lane); count = k; }
                                }
                            } else if (pass) {
                                const float d2 = __uint_as_float((uint32_t)(key >> 32));
                                const float4 *pp = reinterpret_cast<const float4 *>(photons + (uint32_t)key);
                                const float4 pa = __ldg(pp);
                                const f3 wi = oct_decode(__float_as_uint(pa.w));
                                if (dot3(nf, wi) > 0.f) {
                                    const float4 pb = __ldg(pp + 1);
                                    const float sk = 1.f - d2 / max_dist2;
                                    const float wgt = 3.f * B2_INV_PI * sk * sk / (n_paths * max_dist2);
                                    acc = acc + mk3(wgt * pb.x, wgt * pb.y, wgt * pb.z);
                                }
                                ++found_fr;
                            }
                        }
                        __syncwarp();
                    }
                }
            }
            // ---- after the shell: tighten, then decide whether anything outside the scanned block can still matter ----
            // The first radius already comes from the local density, so the bound is only re-tightened between shells
            // when the buffer is more than half full (keeps later sorts on the small networks); otherwise the single
            // sort at the end is the only one a query pays for.
            if (KNN && count > k && count > GATHER_CAP / 2) { bound = warp_keep(cand, count, k, lane); count = k; }
            float dmin = INFINITY;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int f0 = (C[a] - S) << G2_SHIFT, f1 = (C[a] + S + 1) << G2_SHIFT;
                if (f0 > 0) dmin = fminf(dmin, qa[a] - (g.lo[a] + (float)f0 * g.cell));
                if (f1 < g.dims[a]) dmin = fminf(dmin, (g.lo[a] + (float)f1 * g.cell) - qa[a]);
            }
            if (dmin == INFINITY) break;
            dmin -= g.slack;
            const float r2 = __uint_as_float((uint32_t)(bound >> 32));
            if (dmin > 0.f && dmin * dmin >= r2) break;
        }
        if (!KNN || count >= k || r2_try >= max_dist2) break;
        r2_try = fminf(r2_try * 4.f, max_dist2);
      }

        gather_finish<KNN>(A, q, nf, kd, cand, count, acc, found_fr, lane, false);   // shells are not in cell order
    }
}

template <bool KNN>
__global__ void __launch_bounds__(GATHER_WARPS * 32, 6)
gather_tiled_kernel(const GatherArgs A) {
    __shared__ __align__(128) float4 s_stage[GATHER_WARPS][G2_STAGE];
    __shared__ uint64_t s_cand[GATHER_WARPS][GATHER_CAP];
    __shared__ __align__(8) uint64_t s_bar[GATHER_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bar = smem_addr(&s_bar[warp]);
    if (lane == 0) mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    uint32_t phase = 0;
    for (int64_t q = (int64_t)blockIdx.x * GATHER_WARPS + warp; q < A.nq; q += (int64_t)gridDim.x * GATHER_WARPS)
        gather_query_warp<KNN>(A, q, -1.f, s_cand[warp], s_stage[warp], smem_addr(s_stage[warp]), bar, phase, lane);
}

// ---- block-shared tile variant (default for k-nearest) ---------------------------------------------------------------
// Queries arrive in pixel order, and neighbouring pixels select almost the same photons (a pixel is ~2 mm on the
// Cornell walls, the k = 64 radius ~2 cm).  A block therefore takes GB_QT consecutive queries at once:
//   A. warp 0, one lane per query: first radius from the coarse-cell density, the query's fine-cell box for that radius,
//      warp-reduced to the union box of the tile;
//   B. all threads: one (z, y) row of the union box each = one contiguous run of the position array; block prefix sum;
//   C. every thread issues cp.async.bulk for its runs into ONE shared tile (<= GB_TCAP positions), one block mbarrier;
//   D. each warp scans the shared tile for its 8 queries (distance test, ballot-compacted candidates, one sort);
//   E. queries the tile cannot answer - fewer than k photons inside the first radius, a union box with too many rows
//      or photons (sparse caustic regions, incoherent tiles, caustic hot spots) - fall back to gather_query_warp with
//      a private slice of the tile as staging buffer.
// Rows, staging and prefix sums are paid once per 32 queries instead of once per query (they were ~40 % of the
// per-warp kernel's instructions).  The result set of a query is still the k smallest keys below the final bound.
constexpr int GB_ROWS = 2 * GATHER_WARPS * 32;

// GB_TCAP = positions in the shared tile, GB_MINB = resident blocks per SM the register budget is capped for
template <int GB_QT, int GB_TCAP, int GB_MINB>
__global__ void __launch_bounds__(GATHER_WARPS * 32, GB_MINB)
gather_block_kernel(const GatherArgs A) {
    static_assert(GATHER_WARPS * G2_STAGE <= GB_TCAP, "fallback staging slices must fit in the shared tile");
    __shared__ __align__(128) float4 s_tile[GB_TCAP];
    __shared__ uint64_t s_cand[GATHER_WARPS][GATHER_CAP];
    __shared__ float4 s_q[GB_QT];
    __shared__ int s_box[6];
    __shared__ uint32_t s_wsum[GATHER_WARPS];
    __shared__ __align__(8) uint64_t s_bar[GATHER_WARPS + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Gather2Grid &g = A.g;
    const uint32_t lt = (1u << lane) - 1u;
    uint64_t *cand = s_cand[warp];
    const uint32_t wbar = smem_addr(&s_bar[warp]), bbar = smem_addr(&s_bar[GATHER_WARPS]), tile_a = smem_addr(s_tile);
    if (lane == 0) mbar_init(wbar, 1);
    if (tid == 0) mbar_init(bbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    uint32_t wphase = 0, bphase = 0;
    const int64_t ntiles = (A.nq + GB_QT - 1) / GB_QT;

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        fence_proxy_async_smem();       // generic reads of the previous tile before this tile's async writes
        __syncthreads();
        const int64_t q0 = tile * GB_QT;
        const int nqt = (int)(A.nq - q0 < GB_QT ? A.nq - q0 : GB_QT);
        // ---- A: per-query radius and cell box, union over the tile --------------------------------------------------
        if (warp == 0) {
            int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {-1, -1, -1};
            if (lane < nqt) {
                const float4 a0 = __ldg(reinterpret_cast<const float4 *>(A.queries + q0 + lane));
                const float pa[3] = {a0.x, a0.y, a0.z};
                int C[3], cf[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    cf[a] = min(max((int)floorf((pa[a] - g.lo[a]) * g.inv), 0), g.dims[a] - 1);
                    C[a] = cf[a] >> G2_SHIFT;
                }
                uint32_t n8 = 0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int o[3] = {c & 1, (c >> 1) & 1, c >> 2};
                    int P[3];
                    bool in = true;
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        P[a] = C[a] - ((cf[a] & (G2_C - 1)) < G2_C / 2 ? 1 : 0) + o[a];
                        in = in && P[a] >= 0 && P[a] < g.cdims[a];
                    }
                    if (in) n8 += __ldg(A.coarse + ((int64_t)P[2] * g.cdims[1] + P[1]) * g.cdims[0] + P[0]);
                }
                float r2_try = A.max_dist2;
                if (n8 > 0) r2_try = fminf((A.target * 64.f * B2_INV_PI) * (float)A.k * g.cell * g.cell / (float)n8, A.max_dist2);
                const float r = sqrtf(r2_try) * 1.0001f + g.slack;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    lo[a] = min(max((int)floorf((pa[a] - r - g.lo[a]) * g.inv), 0), g.dims[a] - 1);
                    hi[a] = min(max((int)floorf((pa[a] + r - g.lo[a]) * g.inv), 0), g.dims[a] - 1);
                }
                s_q[lane] = make_float4(pa[0], pa[1], pa[2], r2_try);
            }
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                lo[a] = __reduce_min_sync(0xffffffffu, lo[a]);
                hi[a] = __reduce_max_sync(0xffffffffu, hi[a]);
            }
            if (lane == 0) { s_box[0] = lo[0]; s_box[1] = lo[1]; s_box[2] = lo[2]; s_box[3] = hi[0]; s_box[4] = hi[1]; s_box[5] = hi[2]; }
        }
        __syncthreads();
        const int x0 = s_box[0], y0 = s_box[1], z0 = s_box[2], x1 = s_box[3], y1 = s_box[4], z1 = s_box[5];
        const int ny = y1 - y0 + 1, nz = z1 - z0 + 1;
        const int R = ny * nz;
        bool tile_ok = R <= GB_ROWS;
        // ---- B: one row of the union box per thread (two if R > 128), block prefix sum ------------------------------
        uint32_t rb_[2] = {0, 0}, rn_[2] = {0, 0}, off = 0, T = 0;
        if (tile_ok) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int row = tid + j * GATHER_WARPS * 32;
                if (row < R) {
                    const int z = z0 + row / ny, y = y0 + row % ny;
                    const int64_t rb = ((int64_t)z * g.dims[1] + y) * g.dims[0];
                    const uint32_t b = __ldg(A.cell_start + rb + x0), e = __ldg(A.cell_start + rb + x1 + 1);
                    rb_[j] = b; rn_[j] = e - b;
                }
            }
            const uint32_t mine = rn_[0] + rn_[1];
            uint32_t inc = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += up;
            }
            if (lane == 31) s_wsum[warp] = inc;
            __syncthreads();
            uint32_t wbase = 0;
#pragma unroll
            for (int w = 0; w < GATHER_WARPS; ++w) { if (w < warp) wbase += s_wsum[w]; T += s_wsum[w]; }
            off = wbase + inc - mine;
            tile_ok = T <= GB_TCAP;
        }
        // ---- C: stage every run into the shared tile ---------------------------------------------------------------
        if (tile_ok && T > 0) {
            if (tid == 0) mbar_expect_tx(bbar, T * 16u);
            if (rn_[0]) tma_bulk_g2s(tile_a + off * 16u, A.pos4 + rb_[0], rn_[0] * 16u, bbar);
            if (rn_[1]) tma_bulk_g2s(tile_a + (off + rn_[0]) * 16u, A.pos4 + rb_[1], rn_[1] * 16u, bbar);
            mbar_wait(bbar, bphase);
            bphase ^= 1u;
        }
        // ---- D: each warp scans the shared tile for its queries ------------------------------------------------------
        uint32_t failed = 0;
        for (int j = 0; j < GB_QT / GATHER_WARPS; ++j) {
            const int qi = warp + j * GATHER_WARPS;
            if (qi >= nqt) break;
            if (!tile_ok) { failed |= 1u << j; continue; }
            const float4 qv = s_q[qi];
            uint64_t bound = (uint64_t)__float_as_uint(qv.w) << 32;
            int count = 0;
            for (uint32_t i0 = 0; i0 < T; i0 += 32) {
                const uint32_t i = i0 + lane;
                bool pass = false;
                uint64_t key = 0;
                if (i < T) {
                    const float4 pa = s_tile[i];
                    const float dx = pa.x - qv.x, dy = pa.y - qv.y, dz = pa.z - qv.z;
                    const float d2 = dx * dx + dy * dy + dz * dz;
                    key = ((uint64_t)__float_as_uint(d2) << 32) | __float_as_uint(pa.w);
                    pass = key < bound;
                }
                const uint32_t m = __ballot_sync(0xffffffffu, pass);
                if (m) {
                    if (pass) cand[CI(count + __popc(m & lt))] = key;
                    count += __popc(m);
                    __syncwarp();
                    if (count > GATHER_CAP - 32) { bound = warp_keep(cand, count, A.k, lane); count = A.k; }   // keeps cell order
                }
            }
            if (count >= A.k || qv.w >= A.max_dist2) {
                const float4 *qp = reinterpret_cast<const float4 *>(A.queries + q0 + qi);
                const float4 a0 = __ldg(qp), a1 = __ldg(qp + 1), a2 = __ldg(qp + 2);
                gather_finish<true>(A, q0 + qi, mk3(a0.w, a1.x, a1.y), mk3(a1.z, a1.w, a2.x), cand, count, mk3(0.f, 0.f, 0.f), 0, lane,
                                    R <= GATHER_WARPS * 32);     // one row per thread: the tile is in ascending photon index
            } else {
                failed |= 1u << j;
            }
        }
        // ---- E: fall-backs, after every warp is done with the shared tile ---------------------------------------------
        fence_proxy_async_smem();       // this thread's reads of the shared tile before the fall-backs' async writes
        if (__syncthreads_or(failed != 0u)) {
            for (int j = 0; j < GB_QT / GATHER_WARPS; ++j) {
                if (!((failed >> j) & 1u)) continue;
                const int qi = warp + j * GATHER_WARPS;
                const float r2_next = tile_ok ? s_q[qi].w * 4.f : -1.f;
                gather_query_warp<true>(A, q0 + qi, r2_next, cand, s_tile + warp * G2_STAGE,
                                        tile_a + (uint32_t)(warp * G2_STAGE) * 16u, wbar, wphase, lane);
            }
        }
    }
}


// (A warp-specialised producer / consumer version of the same tiles -- one producer warp TMA-staging tile t + 1 into a
// double-buffered shared tile while four consumer warps scan tile t, hand-over by full / empty mbarriers -- was built in round
// 1 as B2RT_GATHER_VARIANT=3: identical bytes, but 19.1 against 15.9 ms for the camera pass, because the
// per-query work wants many resident warps.  Removed in round 2; the code is in the history at commit 4f8443c.)


}  // namespace

int b2rt_gather_impl(b2rt_ctx *ctx, int map, int64_t nq, const b2rt_query *d_queries, int k, float max_dist2,
                     float *d_rgb, float *d_r2, int32_t *d_found, int32_t *d_idx) {
    PhotonMap &m = ctx->maps[map];
    if (nq <= 0) return B2RT_OK;
    GatherGrid g;
    for (int a = 0; a < 3; ++a) { g.lo[a] = m.grid.origin[a]; g.dims[a] = m.grid.dims[a]; }
    g.cell = m.grid.cell_size; g.inv = m.grid.inv_cell;
    float mx = 0.f;
    for (int a = 0; a < 3; ++a) {
        float hi = m.grid.origin[a] + (float)m.grid.dims[a] * m.grid.cell_size;
        mx = fmaxf(mx, fmaxf(fabsf(m.grid.origin[a]), fabsf(hi)));
    }
    g.slack = 1e-5f * mx;
    int64_t blocks = (nq + GATHER_WARPS - 1) / GATHER_WARPS;
    int64_t cap = (int64_t)ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    float npaths = (float)m.n_paths;    // pbrt: int nPaths * float maxDist2
    if (ctx->gather_variant >= 1) {
        Gather2Grid g2;
        for (int a = 0; a < 3; ++a) { g2.lo[a] = g.lo[a]; g2.dims[a] = g.dims[a]; g2.cdims[a] = m.cdims[a]; }
        g2.cell = g.cell; g2.inv = g.inv; g2.slack = g.slack;
        int &per_sm = ctx->gather_blocks_per_sm[k > 0 ? 1 : 0];      // resident grid, persistent over the queries
        if (per_sm == 0) {
            if (k > 0) B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gather_tiled_kernel<true>, GATHER_WARPS * 32, 0));
            else B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gather_tiled_kernel<false>, GATHER_WARPS * 32, 0));
            if (per_sm < 1) per_sm = 1;
        }
        cap = (int64_t)ctx->sm_count * per_sm;
        blocks = (nq + GATHER_WARPS - 1) / GATHER_WARPS;
        if (blocks > cap) blocks = cap;
        GatherArgs A;
        A.pos4 = m.pos4.as<float4>(); A.photons = m.sorted.as<b2rt_photon>(); A.cell_start = m.cell_start.as<uint32_t>();
        A.coarse = m.coarse.as<uint32_t>(); A.g = g2; A.queries = d_queries; A.nq = nq; A.k = k; A.max_dist2 = max_dist2;
        A.n_paths = npaths; A.target = ctx->gather_target; A.out_rgb = d_rgb; A.out_r2 = d_r2; A.out_found = d_found; A.idx_out = d_idx;
        if (k > 0 && ctx->gather_variant >= 2) {
            const int qt = ctx->gather_qt == 32 ? 32 : (ctx->gather_qt == 16 ? 16 : 8);
            // default: 1024-position tile, 7 blocks per SM (r01z sweep: camera pass 15.9 ms against 17.2 ms with 2048 / 5)
            void (*kern)(const GatherArgs) = gather_block_kernel<8, 1024, B2RT_GATHER_BLOCKS>;
            if (qt == 32) kern = gather_block_kernel<32, 2048, 5>;
            else if (qt == 16) kern = gather_block_kernel<16, 2048, 5>;
            else if (ctx->gather_tcap == 2048) kern = gather_block_kernel<8, 2048, 5>;
            int &bsm = ctx->gather_blocks_per_sm[2];
            if (bsm == 0) {
                B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bsm, kern, GATHER_WARPS * 32, 0));
                if (bsm < 1) bsm = 1;
            }
            int64_t nb = (nq + qt - 1) / qt;
            if (nb > (int64_t)ctx->sm_count * bsm) nb = (int64_t)ctx->sm_count * bsm;
            kern<<<(unsigned)nb, GATHER_WARPS * 32, 0, ctx->stream>>>(A);
        } else if (k > 0) gather_tiled_kernel<true><<<(unsigned)blocks, GATHER_WARPS * 32, 0, ctx->stream>>>(A);
        else gather_tiled_kernel<false><<<(unsigned)blocks, GATHER_WARPS * 32, 0, ctx->stream>>>(A);
        B2RT_LAUNCH_CHECK(ctx);
        return B2RT_OK;
    }
    if (k > 0)
        gather_kernel<true><<<(unsigned)blocks, GATHER_WARPS * 32, 0, ctx->stream>>>(
            m.sorted.as<b2rt_photon>(), m.cell_start.as<uint32_t>(), g, d_queries, nq, k, max_dist2, npaths, d_rgb,
            d_r2, d_found, d_idx);
    else
        gather_kernel<false><<<(unsigned)blocks, GATHER_WARPS * 32, 0, ctx->stream>>>(
            m.sorted.as<b2rt_photon>(), m.cell_start.as<uint32_t>(), g, d_queries, nq, k, max_dist2, npaths, d_rgb,
            d_r2, d_found, d_idx);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}
