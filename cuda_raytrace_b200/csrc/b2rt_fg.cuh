// Final gathering: the pieces that are pure arithmetic over arrays (no BVH traversal), written __host__ __device__ so
// tests/hostemu can run the very same code on the CPU against the oracle before it ever reaches a GPU.
//   pbrt code being followed: integrators/photonmap.cpp:156-172 (RadiancePhotonProcess), :263-283 (EPhoton),
//   :696-704,:741-752 (photon-direction pdf + power heuristic), core/montecarlo.cpp:400-420 (uniform cone)
#pragma once
#include "b2rt_device.cuh"

struct FgGrid { float lo[3]; float cell, inv, slack; int dims[3]; };

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
// The search walks Chebyshev shells of cells around the query's (clamped) cell.  After shell r every unseen photon lies
// outside the box of cells [c - r, c + r]; faces of that box at or beyond the grid's edge have nothing behind them.  The
// walk stops once the best squared distance is below (distance to the nearest remaining face - slack)^2, so the result
// is the exact global nearest: identical to a brute-force scan.
__host__ __device__ inline int fg_radiance_nearest(const FgGrid &g, const float4 *__restrict__ photons,
                                                  const uint32_t *__restrict__ cell_start, f3 p, f3 n, float &best_d2) {
    const float pa[3] = {p.x, p.y, p.z};
    int c[3];
    for (int a = 0; a < 3; ++a) {
        float f = floorf((pa[a] - g.lo[a]) * g.inv);
        f = fminf(fmaxf(f, 0.f), (float)(g.dims[a] - 1));
        c[a] = (int)f;
    }
    int rmax = 0;
    for (int a = 0; a < 3; ++a) { rmax = max(rmax, c[a]); rmax = max(rmax, g.dims[a] - 1 - c[a]); }
    int best = -1;
    best_d2 = INFINITY;
    for (int r = 0; r <= rmax; ++r) {
        const int z0 = max(c[2] - r, 0), z1 = min(c[2] + r, g.dims[2] - 1);
        const int y0 = max(c[1] - r, 0), y1 = min(c[1] + r, g.dims[1] - 1);
        const int x0 = max(c[0] - r, 0), x1 = min(c[0] + r, g.dims[0] - 1);
        for (int z = z0; z <= z1; ++z) {
            const bool zface = (z == c[2] - r) || (z == c[2] + r);
            for (int y = y0; y <= y1; ++y) {
                const bool full = zface || (y == c[1] - r) || (y == c[1] + r);
                const uint32_t row = ((uint32_t)z * (uint32_t)g.dims[1] + (uint32_t)y) * (uint32_t)g.dims[0];
                // full row of the shell: cells x0..x1 are one contiguous photon range; otherwise only the two end cells
                for (int part = 0; part < (full ? 1 : 2); ++part) {
                    int xa, xb;
                    if (full) { xa = x0; xb = x1; }
                    else if (part == 0) { xa = xb = c[0] - r; if (xa < 0) continue; }
                    else { xa = xb = c[0] + r; if (r == 0 || xa >= g.dims[0]) continue; }
                    const uint32_t i0 = cell_start[row + (uint32_t)xa], i1 = cell_start[row + (uint32_t)xb + 1u];
                    for (uint32_t i = i0; i < i1; ++i) {
                        const float4 A = photons[2 * (size_t)i];
                        const float dx = A.x - p.x, dy = A.y - p.y, dz = A.z - p.z;
                        const float d2 = dx * dx + dy * dy + dz * dz;
                        if (d2 < best_d2 || (d2 == best_d2 && (int)i < best)) {
                            if (dot3(fg_oct_decode(fg_f2u(A.w)), n) > 0.f) { best = (int)i; best_d2 = d2; }
                        }
                    }
                }
            }
        }
        float bound = INFINITY;
        for (int a = 0; a < 3; ++a) {
            if (c[a] - r > 0) bound = fminf(bound, pa[a] - (g.lo[a] + (float)(c[a] - r) * g.cell));
            if (c[a] + r + 1 < g.dims[a]) bound = fminf(bound, (g.lo[a] + (float)(c[a] + r + 1) * g.cell) - pa[a]);
        }
        if (bound == INFINITY) break;                     // the whole grid has been scanned
        bound -= g.slack;
        if (best >= 0 && bound > 0.f && best_d2 < bound * bound) break;
    }
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
