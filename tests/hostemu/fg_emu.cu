// TEST INFRASTRUCTURE ONLY.  Compiles the __host__ __device__ arithmetic of csrc/b2rt_fg.cuh (the radiance-photon shell
// search and the MIS helpers the final-gather kernel calls) for the HOST, so the CPU test suite can run the very code the
// GPU runs against the C oracle without a GPU.  Never loaded by the product path.
// work counters of the lookup: [0] photons distance-tested, [1] level-2 cells looked at, [2] level-1 cells looked at
static unsigned long long g_fg_stat[4], g_fg_phase_tests[4];
static int g_fg_phase;
#ifdef __CUDA_ARCH__
#define FG_STAT(counter, amount)
#else
#define FG_STAT(counter, amount) (g_fg_stat[counter] += (unsigned long long)(amount), (counter) == 0 ? (g_fg_phase_tests[g_fg_phase] += (unsigned long long)(amount)) : 0ull)
#define FG_PHASE(x) (g_fg_phase = (x))
#endif
#ifndef FG_PHASE
#define FG_PHASE(x)
#endif
#include "../../cuda_raytrace_b200/csrc/b2rt_fg.cuh"

extern "C" {
// grid: lo[3], cell, inv, slack, dims[3];  photons: b2rt_photon[n] in grid order;  queries: 12 floats each
// coarse1: photons per 4^3 fine cells (the grid's own coarse table), dims (d + 3) / 4, row-major x fastest.  The level-2
// counts and both levels' normal bounds are built here with the functions the device kernels call.
// packed != 0: run the lookup over the packed position array + decoded normals (the layout the camera pass uses)
// instead of the sorted photon records; the answer must not change.
void emu_radiance_nearest_layout(const float *lo, float cell, float inv, float slack, const int *dims, const void *photons_in,
                                 const uint32_t *cell_start, const uint32_t *coarse1, long long nq, const float *queries,
                                 int *idx_out, float *d2_out, long long n_photons, int packed) {
    FgGrid g;
    g.pstride = 2; g.nrm = nullptr;
    const void *photons = photons_in;
    float4 *pos = nullptr, *nrm = nullptr;
    if (packed) {
        pos = new float4[n_photons > 0 ? n_photons : 1]; nrm = new float4[n_photons > 0 ? n_photons : 1];
        for (long long i = 0; i < n_photons; ++i) {
            pos[i] = ((const float4 *)photons_in)[2 * i];
            const f3 m = fg_oct_decode(fg_f2u(pos[i].w));
            nrm[i] = make_float4(m.x, m.y, m.z, 0.f);
        }
        g.pstride = 1; g.nrm = nrm; photons = pos;
    }
    for (int a = 0; a < 3; ++a) {
        g.lo[a] = lo[a]; g.dims[a] = dims[a]; g.cdims[a] = (dims[a] + 3) / 4; g.c2dims[a] = (g.cdims[a] + 3) / 4;
    }
    g.cell = cell; g.inv = inv; g.slack = slack;
    const int n1 = g.cdims[0] * g.cdims[1] * g.cdims[2], n2 = g.c2dims[0] * g.c2dims[1] * g.c2dims[2];
    float *nbox1 = new float[12 * (size_t)n1], *nbox2 = new float[12 * (size_t)n2];
    uint32_t *coarse2 = new uint32_t[n2];
    for (int i = 0; i < n1; ++i)
        fg_nbox_level1(g, (const float4 *)photons, cell_start, i % g.cdims[0], (i / g.cdims[0]) % g.cdims[1],
                       i / (g.cdims[0] * g.cdims[1]), nbox1 + 12 * (size_t)i);
    for (int i = 0; i < n2; ++i)
        fg_nbox_level2(g, nbox1, coarse1, i % g.c2dims[0], (i / g.c2dims[0]) % g.c2dims[1], i / (g.c2dims[0] * g.c2dims[1]),
                       nbox2 + 12 * (size_t)i, coarse2 + i);
    for (long long q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        float d2;
        idx_out[q] = fg_radiance_nearest(g, (const float4 *)photons, cell_start, coarse1, coarse2, nbox1, nbox2,
                                         mk3(Q[0], Q[1], Q[2]), mk3(Q[3], Q[4], Q[5]), d2);
        if (d2_out) d2_out[q] = d2;
    }
    delete[] nbox1; delete[] nbox2; delete[] coarse2;
    delete[] pos; delete[] nrm;
}
void emu_radiance_nearest(const float *lo, float cell, float inv, float slack, const int *dims, const void *photons,
                          const uint32_t *cell_start, const uint32_t *coarse1, long long nq, const float *queries,
                          int *idx_out, float *d2_out) {
    emu_radiance_nearest_layout(lo, cell, inv, slack, dims, photons, cell_start, coarse1, nq, queries, idx_out, d2_out, 0, 0);
}
// distance tests per phase of the lookup: fine shells, sampling sweep, refinement around the sampled winner, exact sweep
void emu_phase_tests(unsigned long long *out, int reset) {
    for (int i = 0; i < 4; ++i) { out[i] = g_fg_phase_tests[i]; if (reset) g_fg_phase_tests[i] = 0; }
}
void emu_stat(unsigned long long *out, int reset) {
    for (int i = 0; i < 4; ++i) { out[i] = g_fg_stat[i]; if (reset) g_fg_stat[i] = 0; }
}
float emu_cone_pdf(float cos_max) { return fg_cone_pdf(cos_max); }
float emu_power_heuristic(int nf, float f, int ng, float g) { return fg_power_heuristic(nf, f, ng, g); }
void emu_sample_cone(float u1, float u2, float cos_max, const float *z, float *out) {
    f3 zz = mk3(z[0], z[1], z[2]), vx, vy;
    fg_coordinate_system(zz, vx, vy);
    f3 w = fg_sample_cone(u1, u2, cos_max, vx, vy, zz);
    out[0] = w.x; out[1] = w.y; out[2] = w.z;
}
uint32_t emu_oct_encode(const float *v) { return fg_oct_encode(mk3(v[0], v[1], v[2])); }
void emu_oct_decode(uint32_t w, float *o) { f3 v = fg_oct_decode(w); o[0] = v.x; o[1] = v.y; o[2] = v.z; }
}
