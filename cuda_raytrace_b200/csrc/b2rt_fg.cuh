// Final gathering: the pieces that are pure arithmetic over arrays (no BVH traversal), written __host__ __device__ so
// tests/hostemu can run the very same code on the CPU against the oracle before it ever reaches a GPU.
//   pbrt code being followed: integrators/photonmap.cpp:156-172 (RadiancePhotonProcess), :263-283 (EPhoton),
//   :696-704,:741-752 (photon-direction pdf + power heuristic), core/montecarlo.cpp:400-420 (uniform cone)
#pragma once
#include "b2rt_device.cuh"

#ifndef FG_PHASE
#define FG_PHASE(x)
#endif
#ifndef FG_STAT
#define FG_STAT(counter, amount)          // tests/hostemu defines it to count the work of a lookup on the host
#endif
// pstride / nrm: how the radiance photons are laid out for the lookup.  pstride = 2, nrm = NULL: the grid-sorted b2rt_photon
// array itself (float4 pairs, the normal octahedral in the first one's w).  pstride = 1, nrm != NULL: the packed copies
// fg_pack_kernel makes for the camera pass -- 16 bytes per distance test instead of a 32-byte record, and the normals
// already decoded (fg_oct_decode + normalise were 18 % of the gather kernel's instructions, ncu r03g); same values, same bits.
struct FgGrid { float lo[3]; float cell, inv, slack; int dims[3]; int cdims[3]; int c2dims[3]; int pstride; const float4 *nrm; };
__host__ __device__ __forceinline__ f3 fg_photon_normal(const FgGrid &g, uint32_t i, float w);

__host__ __device__ __forceinline__ f3 fg_normalize(f3 v) {      // pbrt Normalize (same arithmetic as normalize3)
    float inv = 1.f / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    return mk3(v.x * inv, v.y * inv, v.z * inv);
}
// octahedral 2 x snorm16 <-> unit vector (same code as the photon records' wi, photon.cu / gather.cu)
__host__ __device__ __forceinline__ uint32_t fg_oct_encode(f3 v) {
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
__host__ __device__ __forceinline__ f3 fg_oct_decode(uint32_t w) {
    float x = (float)(int)(int16_t)(w & 0xffffu) / 32767.f, y = (float)(int)(int16_t)(w >> 16) / 32767.f;
    float z = 1.f - fabsf(x) - fabsf(y);
    if (z < 0.f) {
        float ox = (1.f - fabsf(y)) * (x >= 0.f ? 1.f : -1.f);
        float oy = (1.f - fabsf(x)) * (y >= 0.f ? 1.f : -1.f);
        x = ox; y = oy;
    }
    return fg_normalize(mk3(x, y, z));
}
__host__ __device__ __forceinline__ uint32_t fg_f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}

// Nearest radiance photon whose normal faces n (Dot(rp.n, n) > 0), unbounded radius; ties go to the lower sorted index.
// photons: the grid-sorted b2rt_photon array viewed as float4 pairs ({p.xyz, bits(oct normal)}, {Lo.rgb, -}).
//
// Phase 0 walks Chebyshev shells of FINE cells around the query's (clamped) cell, r = 0 .. FG_FINE_SHELLS.  After shell r
// every unseen photon lies outside the box of cells [c - r, c + r]; faces of that box at or beyond the grid's edge have
// nothing behind them.  The walk stops once the best squared distance is below (distance to the nearest remaining face -
// slack)^2.  Most lookups end here (a gather ray that hits a wall finds a photon of that wall next to it).
// Phase 1 serves the rest -- hits on specular objects, on surfaces no photon reached: the nearest facing photon may be
// tens of cells away and a fine-cell walk would spend its time in empty space (r03a: 43 s per 960 x 540 frame).  It walks
// shells of LEVEL-2 cells (16^3 fine cells) and descends only where the occupancy tables (coarse1: photons per 4^3 fine
// cells, coarse2: per 16^3) are non-zero, the cell's normal bounds (nbox1 / nbox2, below) admit a photon that faces n and
// the cell's box is not farther than the best photon so far: level-2 cell ->
// its 4^3 level-1 cells -> their 16 fine rows (one contiguous photon range each).  Same stopping rule with the level-2
// cell size.  Both phases only ever skip photons that cannot win, so the result is the exact global nearest: identical to
// a brute-force scan (tests/test_oracle_fg_cpu.py runs this very code on the host against one).
#define FG_FINE_SHELLS 2
#ifndef FG_REFINE_PASSES
#define FG_REFINE_PASSES 1
#endif
__host__ __device__ __forceinline__ f3 fg_photon_normal(const FgGrid &g, uint32_t i, float w) {
    if (g.nrm) { const float4 N = g.nrm[i]; return mk3(N.x, N.y, N.z); }
    return fg_oct_decode(fg_f2u(w));
}
__host__ __device__ __forceinline__ void fg_scan_range(const FgGrid &g, const float4 *__restrict__ photons, uint32_t i0, uint32_t i1,
                                                       f3 p, f3 n, int &best, float &best_d2) {
    FG_STAT(0, i1 - i0);
    for (uint32_t i = i0; i < i1; ++i) {
        const float4 A = photons[(size_t)g.pstride * i];
        const float dx = A.x - p.x, dy = A.y - p.y, dz = A.z - p.z;
        const float d2 = dx * dx + dy * dy + dz * dz;
        if (d2 < best_d2 || (d2 == best_d2 && (int)i < best)) {
            if (dot3(fg_photon_normal(g, i, A.w), n) > 0.f) { best = (int)i; best_d2 = d2; }
        }
    }
}
// lower bound of the squared distance from p to any photon binned in fine cells [i0, i1) per axis.  Faces on the grid's
// edge are open (edge cells also hold the photons clamped into them); the slack covers the rounding of the binning.
__host__ __device__ __forceinline__ float fg_box_d2(const FgGrid &g, const float *pa, int x0, int x1, int y0, int y1, int z0,
                                                    int z1) {
    const int i0[3] = {x0, y0, z0}, i1[3] = {x1, y1, z1};
    float d2 = 0.f;
    for (int a = 0; a < 3; ++a) {
        float d = 0.f;
        if (i0[a] > 0) d = fmaxf(d, (g.lo[a] + (float)i0[a] * g.cell) - pa[a] - g.slack);
        if (i1[a] < g.dims[a]) d = fmaxf(d, pa[a] - (g.lo[a] + (float)i1[a] * g.cell) - g.slack);
        d2 += d * d;
    }
    return d2;
}
// Cell bounds: per level-1 / level-2 cell 12 floats, the componentwise min and max of its photons' normals (lo.xyz,
// hi.xyz) and then of their positions.  No photon of the cell can face n when sum_a max(n_a * lo_a, n_a * hi_a) <= 0: every product and the two
// additions are the ones dot3() performs on a photon of the cell with each factor replaced by its bound, and rounded
// multiplication and addition are monotonic, so the photon's own dot product is <= 0 too and the scan would reject it.
// Without this a query whose neighbourhood only holds photons facing away (a gather ray that hits a mirror ball next to
// a wall) scans every one of them without ever tightening the search (r03b: 27 s per 960 x 540 frame).
__host__ __device__ __forceinline__ bool fg_may_face(const float *__restrict__ nb, f3 n) {
    const float b = fmaxf(n.x * nb[0], n.x * nb[3]) + fmaxf(n.y * nb[1], n.y * nb[4]) + fmaxf(n.z * nb[2], n.z * nb[5]);
    return b > 0.f;
}
// squared distance from p to the bounding box of a cell's photon POSITIONS (pb = lo.xyz, hi.xyz): a lower bound of what
// fg_scan_range computes for any photon of the cell (same operations on bounds, monotonic rounding).  Much tighter than
// the cell's own box when the photons lie on a surface that crosses it: with the grid box, a query above a floor opens
// every floor cell within sqrt(2 h cell) of the foot point, a hundred cells, however good its current best is.
__host__ __device__ __forceinline__ float fg_pbox_d2(const float *__restrict__ pb, f3 p) {
    const float dx = fmaxf(fmaxf(pb[0] - p.x, p.x - pb[3]), 0.f), dy = fmaxf(fmaxf(pb[1] - p.y, p.y - pb[4]), 0.f),
                dz = fmaxf(fmaxf(pb[2] - p.z, p.z - pb[5]), 0.f);
    return dx * dx + dy * dy + dz * dz;
}
__host__ __device__ inline void fg_nbox_level1(const FgGrid &g, const float4 *__restrict__ photons,
                                               const uint32_t *__restrict__ cell_start, int X, int Y, int Z, float *out) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float plo[3] = {INFINITY, INFINITY, INFINITY}, phi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const int fx0 = 4 * X, fx1 = min(fx0 + 4, g.dims[0]);
    for (int z = 4 * Z; z < min(4 * Z + 4, g.dims[2]); ++z)
        for (int y = 4 * Y; y < min(4 * Y + 4, g.dims[1]); ++y) {
            const size_t row = ((size_t)z * g.dims[1] + y) * g.dims[0];
            for (uint32_t i = cell_start[row + fx0]; i < cell_start[row + fx1]; ++i) {
                const float4 A = photons[(size_t)g.pstride * i];
                const f3 m = fg_photon_normal(g, i, A.w);
                lo[0] = fminf(lo[0], m.x); lo[1] = fminf(lo[1], m.y); lo[2] = fminf(lo[2], m.z);
                hi[0] = fmaxf(hi[0], m.x); hi[1] = fmaxf(hi[1], m.y); hi[2] = fmaxf(hi[2], m.z);
                plo[0] = fminf(plo[0], A.x); plo[1] = fminf(plo[1], A.y); plo[2] = fminf(plo[2], A.z);
                phi[0] = fmaxf(phi[0], A.x); phi[1] = fmaxf(phi[1], A.y); phi[2] = fmaxf(phi[2], A.z);
            }
        }
    for (int a = 0; a < 3; ++a) { out[a] = lo[a]; out[3 + a] = hi[a]; out[6 + a] = plo[a]; out[9 + a] = phi[a]; }
}
__host__ __device__ inline void fg_nbox_level2(const FgGrid &g, const float *__restrict__ nbox1,
                                               const uint32_t *__restrict__ coarse1, int X, int Y, int Z, float *out,
                                               uint32_t *count) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float plo[3] = {INFINITY, INFINITY, INFINITY}, phi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t cnt = 0;
    for (int z = 4 * Z; z < min(4 * Z + 4, g.cdims[2]); ++z)
        for (int y = 4 * Y; y < min(4 * Y + 4, g.cdims[1]); ++y)
            for (int x = 4 * X; x < min(4 * X + 4, g.cdims[0]); ++x) {
                const size_t i = ((size_t)z * g.cdims[1] + y) * g.cdims[0] + x;
                if (coarse1[i] == 0u) continue;
                cnt += coarse1[i];
                for (int a = 0; a < 3; ++a) {
                    lo[a] = fminf(lo[a], nbox1[12 * i + a]); hi[a] = fmaxf(hi[a], nbox1[12 * i + 3 + a]);
                    plo[a] = fminf(plo[a], nbox1[12 * i + 6 + a]); phi[a] = fmaxf(phi[a], nbox1[12 * i + 9 + a]);
                }
            }
    for (int a = 0; a < 3; ++a) { out[a] = lo[a]; out[3 + a] = hi[a]; out[6 + a] = plo[a]; out[9 + a] = phi[a]; }
    *count = cnt;
}
// min over the warp by (d2, index): every lane ends with the winner.  NL = 1 (one thread owns the lookup): nothing to do.
template <int NL>
__host__ __device__ __forceinline__ void fg_lanes_min(int &best, float &best_d2) {
#ifdef __CUDA_ARCH__
    if (NL > 1) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float od2 = __shfl_xor_sync(0xffffffffu, best_d2, o);
            const int ob = __shfl_xor_sync(0xffffffffu, best, o);
            if (ob >= 0 && (od2 < best_d2 || (od2 == best_d2 && (best < 0 || ob < best)))) { best_d2 = od2; best = ob; }
        }
    }
#endif
}
// lower bound for the photons of fine cells [x0, x1) x {y} x {z} inside a level-1 cell with position bounds pb: per axis a
// photon lies in the grid interval AND in the cell's position interval, so its distance is at least the larger of the
// two (a floor's photons all have y = 0: the row's grid box alone would put them up to a cell nearer than they are)
__host__ __device__ __forceinline__ float fg_row_d2(const FgGrid &g, const float *pa, const float *__restrict__ pb, int x0, int x1,
                                                    int y, int z) {
    const int i0[3] = {x0, y, z}, i1[3] = {x1, y + 1, z + 1};
    float d2 = 0.f;
    for (int a = 0; a < 3; ++a) {
        float d = fmaxf(fmaxf(pb[a] - pa[a], pa[a] - pb[3 + a]), 0.f);
        if (i0[a] > 0) d = fmaxf(d, (g.lo[a] + (float)i0[a] * g.cell) - pa[a] - g.slack);
        if (i1[a] < g.dims[a]) d = fmaxf(d, pa[a] - (g.lo[a] + (float)i1[a] * g.cell) - g.slack);
        d2 += d * d;
    }
    return d2;
}
// one level-2 cell (X, Y, Z): descend through the occupancy tables.  NL lanes share the cell: lane l takes the level-1
// children l, l + NL, ... with a private best and the lanes agree on the minimum before they leave (all decisions up to
// the children loop depend only on values every lane holds, so a warp stays converged around the shuffles).
template <int NL>
__host__ __device__ inline void fg_visit_level2(const FgGrid &g, const float4 *__restrict__ photons,
                                                const uint32_t *__restrict__ cell_start, const uint32_t *__restrict__ coarse1,
                                                const uint32_t *__restrict__ coarse2, const float *__restrict__ nbox1,
                                                const float *__restrict__ nbox2, const float *pa, f3 p, f3 n, int X, int Y,
                                                int Z, bool sample, int lane, int &best, float &best_d2) {
    const size_t i2 = ((size_t)Z * g.c2dims[1] + Y) * g.c2dims[0] + X;
    FG_STAT(1, 1);
    if (coarse2[i2] == 0u || !fg_may_face(nbox2 + 12 * i2, n)) return;
    if (fg_pbox_d2(nbox2 + 12 * i2 + 6, p) > best_d2) return;
    // One thread per lookup: open the admissible child whose position box is nearest first.  The nearest photon is most
    // likely in it, and with best = its distance the tight boxes of the other children prune them without a distance test
    // (raster order would scan every child that lies before it).  Lanes sharing a lookup take their children as they come.
    int kfirst = -1;
    if (NL == 1) {
        float dmin = best_d2;
        for (int k = 0; k < 64; ++k) {
            const int x1 = 4 * X + (k & 3), y1 = 4 * Y + ((k >> 2) & 3), z1 = 4 * Z + (k >> 4);
            if (x1 >= g.cdims[0] || y1 >= g.cdims[1] || z1 >= g.cdims[2]) continue;
            const size_t i1c = ((size_t)z1 * g.cdims[1] + y1) * g.cdims[0] + x1;
            FG_STAT(2, 1);
            if (coarse1[i1c] == 0u || !fg_may_face(nbox1 + 12 * i1c, n)) continue;
            const float d = fg_pbox_d2(nbox1 + 12 * i1c + 6, p);
            if (d <= dmin) { if (d < dmin || kfirst < 0) { dmin = d; kfirst = k; } }
        }
        if (kfirst < 0) return;
    }
    for (int kk = (NL == 1 ? -1 : lane); kk < 64; kk += NL) {
                const int k = kk < 0 ? kfirst : kk;
                if (kk >= 0 && k == kfirst) continue;
                const int x1 = 4 * X + (k & 3), y1 = 4 * Y + ((k >> 2) & 3), z1 = 4 * Z + (k >> 4);
                if (x1 >= g.cdims[0] || y1 >= g.cdims[1] || z1 >= g.cdims[2]) continue;
                const size_t i1c = ((size_t)z1 * g.cdims[1] + y1) * g.cdims[0] + x1;
                if (NL > 1) FG_STAT(2, 1);
                if (coarse1[i1c] == 0u || !fg_may_face(nbox1 + 12 * i1c, n)) continue;
                if (fg_pbox_d2(nbox1 + 12 * i1c + 6, p) > best_d2) continue;
                const int fx0 = 4 * x1, fx1 = min(fx0 + 4, g.dims[0]);
                const int fy1 = min(4 * y1 + 4, g.dims[1]), fz1 = min(4 * z1 + 4, g.dims[2]);
                for (int z = 4 * z1; z < fz1; ++z)
                    for (int y = 4 * y1; y < fy1; ++y) {
                        const size_t row = ((size_t)z * g.dims[1] + y) * g.dims[0];
                        const uint32_t i0 = cell_start[row + fx0], i1 = cell_start[row + fx1];
                        if (i0 == i1) continue;
                        if (fg_row_d2(g, pa, nbox1 + 12 * i1c + 6, fx0, fx1, y, z) > best_d2) continue;
                        FG_STAT(3, sample ? 1 : 0);
                        fg_scan_range(g, photons, i0, sample ? i0 + 1u : i1, p, n, best, best_d2);
                    }
    }
    fg_lanes_min<NL>(best, best_d2);
}
// phase 0; returns true when the lookup is finished (best / best_d2 final), false when phase 1 has to continue from them
// NL = 1: one thread owns the lookup.  NL = 32: the lanes of a warp share it (same p, n on all): they stride over the
// photons of every row range with a private best and agree on the minimum after each shell, so the stopping rule sees the
// same values on every lane and the warp stays converged.
template <int NL>
__host__ __device__ inline bool fg_radiance_phase0(const FgGrid &g, const float4 *__restrict__ photons,
                                                   const uint32_t *__restrict__ cell_start, f3 p, f3 n, int lane, int &best,
                                                   float &best_d2) {
    const float pa[3] = {p.x, p.y, p.z};
    int c[3];
    for (int a = 0; a < 3; ++a) {
        float f = floorf((pa[a] - g.lo[a]) * g.inv);
        f = fminf(fmaxf(f, 0.f), (float)(g.dims[a] - 1));
        c[a] = (int)f;
    }
    int rmax = 0;
    for (int a = 0; a < 3; ++a) { rmax = max(rmax, c[a]); rmax = max(rmax, g.dims[a] - 1 - c[a]); }
    best = -1;
    best_d2 = INFINITY;
    FG_PHASE(0);
    // ---- phase 0: fine shells -----------------------------------------------------------------------------------------
    for (int r = 0; r <= min(rmax, FG_FINE_SHELLS); ++r) {
        const int z0 = max(c[2] - r, 0), z1 = min(c[2] + r, g.dims[2] - 1);
        const int y0 = max(c[1] - r, 0), y1 = min(c[1] + r, g.dims[1] - 1);
        const int x0 = max(c[0] - r, 0), x1 = min(c[0] + r, g.dims[0] - 1);
        for (int z = z0; z <= z1; ++z) {
            const bool zface = (z == c[2] - r) || (z == c[2] + r);
            for (int y = y0; y <= y1; ++y) {
                const bool full = zface || (y == c[1] - r) || (y == c[1] + r);
                const size_t row = ((size_t)z * g.dims[1] + y) * g.dims[0];
                // full row of the shell: cells x0..x1 are one contiguous photon range; otherwise only the two end cells
                for (int part = 0; part < (full ? 1 : 2); ++part) {
                    int xa, xb;
                    if (full) { xa = x0; xb = x1; }
                    else if (part == 0) { xa = xb = c[0] - r; if (xa < 0) continue; }
                    else { xa = xb = c[0] + r; if (xa >= g.dims[0]) continue; }
                    const uint32_t i0 = cell_start[row + xa], i1 = cell_start[row + xb + 1];
                    if (NL == 1) fg_scan_range(g, photons, i0, i1, p, n, best, best_d2);
                    else
                        for (uint32_t i = i0 + (uint32_t)lane; i < i1; i += NL) fg_scan_range(g, photons, i, i + 1u, p, n, best, best_d2);
                }
            }
        }
        fg_lanes_min<NL>(best, best_d2);
        float bound = INFINITY;
        for (int a = 0; a < 3; ++a) {
            if (c[a] - r > 0) bound = fminf(bound, pa[a] - (g.lo[a] + (float)(c[a] - r) * g.cell));
            if (c[a] + r + 1 < g.dims[a]) bound = fminf(bound, (g.lo[a] + (float)(c[a] + r + 1) * g.cell) - pa[a]);
        }
        if (bound == INFINITY) return true;               // the whole grid has been scanned
        bound -= g.slack;
        if (best >= 0 && bound > 0.f && best_d2 < bound * bound) return true;
    }
    return false;
}
// ---- phase 1: level-2 shells, descent through the occupancy tables; NL lanes work on ONE lookup (same p, n, best on all) ---
template <int NL>
__host__ __device__ inline void fg_radiance_phase1(const FgGrid &g, const float4 *__restrict__ photons,
                                                   const uint32_t *__restrict__ cell_start, const uint32_t *__restrict__ coarse1,
                                                   const uint32_t *__restrict__ coarse2, const float *__restrict__ nbox1,
                                                   const float *__restrict__ nbox2, f3 p, f3 n, int lane, int &best,
                                                   float &best_d2) {
    const float pa[3] = {p.x, p.y, p.z};
    int C[3];
    for (int a = 0; a < 3; ++a) {
        float f = floorf((pa[a] - g.lo[a]) * g.inv);
        f = fminf(fmaxf(f, 0.f), (float)(g.dims[a] - 1));
        C[a] = (int)f >> 4;
    }
    int Rmax = 0;
    for (int a = 0; a < 3; ++a) { Rmax = max(Rmax, C[a]); Rmax = max(Rmax, g.c2dims[a] - 1 - C[a]); }
    const float cell2 = 16.f * g.cell;
    // Two sweeps.  The first only SAMPLES the admissible cells (the first photon of every fine row): a facing photon near
    // the true nearest turns up after a small fraction of the distance tests, and its distance is a valid upper bound.
    // The second, exact sweep then only opens cells that intersect the ball of that radius.  (One exact sweep in raster
    // order spends most of its tests on cells it would have skipped had it seen the near ones first: tools/fg_lookup_work.py.)
    for (int sweep = 0; sweep < 2; ++sweep) {
    const bool sample = sweep == 0;
    FG_PHASE(sample ? 1 : 3);
    for (int R = 0; R <= Rmax; ++R) {
        const int Z0 = max(C[2] - R, 0), Z1 = min(C[2] + R, g.c2dims[2] - 1);
        const int Y0 = max(C[1] - R, 0), Y1 = min(C[1] + R, g.c2dims[1] - 1);
        const int X0 = max(C[0] - R, 0), X1 = min(C[0] + R, g.c2dims[0] - 1);
        // mode 0 finds the admissible cell of the shell whose position box is nearest and opens it first (see
        // fg_visit_level2), mode 1 opens the others in raster order
        int fX = -1, fY = -1, fZ = -1;
        for (int mode = 0; mode < 2; ++mode) {
            float dmin = best_d2;
            for (int Z = Z0; Z <= Z1; ++Z) {
                const bool zface = (Z == C[2] - R) || (Z == C[2] + R);
                for (int Y = Y0; Y <= Y1; ++Y) {
                    const bool full = zface || (Y == C[1] - R) || (Y == C[1] + R);
                    for (int X = full ? X0 : max(C[0] - R, X0); X <= X1; ++X) {
                        if (!full && X != C[0] - R && X != C[0] + R) {        // interior of the row: jump to its far end
                            if (C[0] + R > X1 || X > C[0] + R) break;
                            X = C[0] + R;
                        }
                        if (mode == 0) {
                            const size_t i2 = ((size_t)Z * g.c2dims[1] + Y) * g.c2dims[0] + X;
                            if (coarse2[i2] == 0u || !fg_may_face(nbox2 + 12 * i2, n)) continue;
                            const float d = fg_pbox_d2(nbox2 + 12 * i2 + 6, p);
                            if (d <= dmin && (d < dmin || fX < 0)) { dmin = d; fX = X; fY = Y; fZ = Z; }
                        } else if (X != fX || Y != fY || Z != fZ) {
                            fg_visit_level2<NL>(g, photons, cell_start, coarse1, coarse2, nbox1, nbox2, pa, p, n, X, Y, Z, sample, lane, best, best_d2);
                        }
                    }
                }
            }
            if (mode == 0 && fX >= 0)
                fg_visit_level2<NL>(g, photons, cell_start, coarse1, coarse2, nbox1, nbox2, pa, p, n, fX, fY, fZ, sample, lane, best, best_d2);
        }
        float bound = INFINITY;
        for (int a = 0; a < 3; ++a) {
            if (C[a] - R > 0) bound = fminf(bound, pa[a] - (g.lo[a] + (float)(C[a] - R) * cell2));
            if (C[a] + R + 1 < g.c2dims[a]) bound = fminf(bound, (g.lo[a] + (float)(C[a] + R + 1) * cell2) - pa[a]);
        }
        if (bound == INFINITY) break;
        bound -= g.slack;
        if (best >= 0 && bound > 0.f && best_d2 < bound * bound) break;
    }
    // between the sweeps: all photons of the 3 x 3 x 3 fine cells around the sampled winner.  The exact sweep has to open
    // every admissible cell that reaches into the ball of the current best, i.e. a layer around the true nearest as thick
    // as the sample's error; this brings the error from a couple of cells down to the local photon spacing.
    for (int pass = 0, prev = -1; sample && best >= 0 && best != prev && pass < FG_REFINE_PASSES; ++pass) {
        prev = best;
        FG_PHASE(2);
        const float4 B = photons[(size_t)g.pstride * best];
        const float pb[3] = {B.x, B.y, B.z};
        int cb[3];
        for (int a = 0; a < 3; ++a) {
            float f = floorf((pb[a] - g.lo[a]) * g.inv);
            f = fminf(fmaxf(f, 0.f), (float)(g.dims[a] - 1));
            cb[a] = (int)f;
        }
        const int xa = max(cb[0] - 1, 0), xb = min(cb[0] + 1, g.dims[0] - 1);
        for (int k = lane; k < 9; k += NL) {
            const int y = cb[1] - 1 + k % 3, z = cb[2] - 1 + k / 3;
            if (y < 0 || z < 0 || y >= g.dims[1] || z >= g.dims[2]) continue;
            const size_t row = ((size_t)z * g.dims[1] + y) * g.dims[0];
            fg_scan_range(g, photons, cell_start[row + xa], cell_start[row + xb + 1], p, n, best, best_d2);
        }
        fg_lanes_min<NL>(best, best_d2);
    }
    }
}
// the whole lookup by one thread
__host__ __device__ inline int fg_radiance_nearest(const FgGrid &g, const float4 *__restrict__ photons,
                                                  const uint32_t *__restrict__ cell_start, const uint32_t *__restrict__ coarse1,
                                                  const uint32_t *__restrict__ coarse2, const float *__restrict__ nbox1,
                                                  const float *__restrict__ nbox2, f3 p, f3 n, float &best_d2) {
    int best;
    if (!fg_radiance_phase0<1>(g, photons, cell_start, p, n, 0, best, best_d2))
        fg_radiance_phase1<1>(g, photons, cell_start, coarse1, coarse2, nbox1, nbox2, p, n, 0, best, best_d2);
    return best;
}

// pbrt PowerHeuristic(nf, fPdf, ng, gPdf) (core/montecarlo.h:264-267)
__host__ __device__ __forceinline__ float fg_power_heuristic(int nf, float fPdf, int ng, float gPdf) {
    float f = (float)nf * fPdf, g = (float)ng * gPdf;
    return (f * f) / (f * f + g * g);
}
// UniformConePdf (core/montecarlo.cpp:400-402): the 2*pi is a double product in pbrt, rounded once
__host__ __device__ __forceinline__ float fg_cone_pdf(float cos_max) {
    return (float)(1.0 / ((double)(2.f) * 3.14159265358979323846 * (double)(1.f - cos_max)));
}
// pdf of choosing direction wi by "pick one of nd photon directions, sample its cone" (photonmap.cpp:696-703)
__host__ __device__ __forceinline__ float fg_photon_pdf(const f3 *dirs, int nd, f3 wi, float cos_angle, float cone_pdf) {
    float pdf = 0.f;
    const float thr = .999f * cos_angle;
    for (int j = 0; j < nd; ++j)
        if (dot3(dirs[j], wi) > thr) pdf += cone_pdf;
    return nd > 0 ? pdf / (float)nd : 0.f;
}
// pbrt CoordinateSystem + Cross with the double products of core/geometry.h, host/device neutral
__host__ __device__ __forceinline__ f3 fg_cross(f3 a, f3 b) {
    const double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return mk3((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
__host__ __device__ __forceinline__ void fg_coordinate_system(f3 v1, f3 &v2, f3 &v3) {
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        v2 = mk3(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        v2 = mk3(0.f, v1.z * invLen, -v1.y * invLen);
    }
    v3 = fg_cross(v1, v2);
}
// UniformSampleCone(u1, u2, costhetamax, x, y, z) (core/montecarlo.cpp:413-420); phi = 2*pi*u2 = 4*u2 quarter turns
__host__ __device__ __forceinline__ f3 fg_sample_cone(float u1, float u2, float cos_max, f3 x, f3 y, f3 z) {
    float costheta = (1.f - u1) * cos_max + u1 * 1.f;
    float sintheta = sqrtf(fmaxf(0.f, 1.f - costheta * costheta));
    float s, c;
    b2_sincos_quarter(4.f * u2, s, c);
    float a = c * sintheta, b = s * sintheta;
    return mk3(a * x.x + b * y.x + costheta * z.x, a * x.y + b * y.y + costheta * z.y, a * x.z + b * y.z + costheta * z.z);
}
