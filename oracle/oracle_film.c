/* TEST INFRASTRUCTURE - CPU restatement of pbrt-v2's image film, used only by tests/, __graft_entry__.smoke() and
 * bench.py's CPU baseline as the checker of the CUDA film kernel (K13).  Never linked into the product library.
 *
 * Follows (reference file:line):
 *   filter weight table 16x16          pbrt-v2/src/film/image.cpp:56-69
 *   ImageFilm::AddSample               pbrt-v2/src/film/image.cpp:77-137
 *   ImageFilm::WriteImage normalisation pbrt-v2/src/film/image.cpp:178-213 (splat term is zero on this path)
 *   filters                            pbrt-v2/src/filters/box.cpp:39, triangle.cpp:39-42, gaussian.cpp:39 + gaussian.h:47-58,
 *                                      mitchell.cpp:39 + mitchell.h:50-59, sinc.cpp:39 + sinc.h:49-57
 *   RGB <-> XYZ                        pbrt-v2/src/core/spectrum.h:72-87 (RGBSpectrum::ToXYZ / FromXYZ)
 * Pin: oracle/_ref's own ImageFilm fed with the same samples (tests/test_oracle_film_cpu.py, live and golden).
 *
 * Samples arrive cell-major: cell (cx, cy) of the window [cx0, cx1) x [cy0, cy1) holds spp samples whose continuous
 * raster position lies inside the cell - what a per-pixel sampler produces.  pbrt adds samples in arrival order; the
 * port (and the CUDA kernel) add them per pixel in (cell row-major, sample) order, so the two differ from pbrt by
 * rounding only and agree with each other bit for bit. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define FT 16

enum { F_BOX = 0, F_TRIANGLE = 1, F_GAUSSIAN = 2, F_MITCHELL = 3, F_SINC = 4 };

static float gauss1(float d, float alpha, float expv) {
    float v = (float)(expf(-alpha * d * d) - expv);
    return v > 0.f ? v : 0.f;
}
static float mitchell1(float x, float B, float C) {
    x = fabsf(2.f * x);
    if (x > 1.f)
        return ((-B - 6 * C) * x * x * x + (6 * B + 30 * C) * x * x + (-12 * B - 48 * C) * x + (8 * B + 24 * C)) * (1.f / 6.f);
    return ((12 - 9 * B - 6 * C) * x * x * x + (-18 + 12 * B + 6 * C) * x * x + (6 - 2 * B)) * (1.f / 6.f);
}
static float sinc1(float x, float tau) {
    x = fabsf(x);
    if (x < 1e-5) return 1.f;
    if (x > 1.) return 0.f;
    x *= 3.14159265358979323846f;      /* pbrt's M_PI is a float literal (core/pbrt.h) */
    float sinc = sinf(x) / x;
    float lanczos = sinf(x * tau) / (x * tau);
    return sinc * lanczos;
}
static float filter_eval(int type, float xw, float yw, float p0, float p1, float x, float y) {
    switch (type) {
    case F_TRIANGLE: { float a = xw - fabsf(x), b = yw - fabsf(y); return (a > 0.f ? a : 0.f) * (b > 0.f ? b : 0.f); }
    case F_GAUSSIAN: return gauss1(x, p0, expf(-p0 * xw * xw)) * gauss1(y, p0, expf(-p0 * yw * yw));
    case F_MITCHELL: return mitchell1(x * (1.f / xw), p0, p1) * mitchell1(y * (1.f / yw), p0, p1);
    case F_SINC: return sinc1(x * (1.f / xw), p0) * sinc1(y * (1.f / yw), p0);
    default: return 1.f;
    }
}
void port_filter_table(int type, float xw, float yw, float p0, float p1, float *table) {
    for (int y = 0; y < FT; ++y) {
        float fy = ((float)y + .5f) * yw / FT;
        for (int x = 0; x < FT; ++x) {
            float fx = ((float)x + .5f) * xw / FT;
            table[y * FT + x] = filter_eval(type, xw, yw, p0, p1, fx, fy);
        }
    }
}

static int ceil2int(float v) { return (int)ceilf(v); }
static int floor2int(float v) { return (int)floorf(v); }

/* xy[2*s], rgb[3*s] for s = (cell row-major in the window) * spp + k;  out: (x1-x0)*(y1-y0)*3, pixel window row-major.
 * The pixel window plays the role of the film's crop window (AddSample clamps to it). */
void port_film(int type, float xw, float yw, float p0, float p1, int cx0, int cy0, int cx1, int cy1, int spp,
               const float *xy, const float *rgb, int x0, int y0, int x1, int y1, float *out) {
    float table[FT * FT];
    port_filter_table(type, xw, yw, p0, p1, table);
    const float invx = 1.f / xw, invy = 1.f / yw;
    const int W = x1 - x0, H = y1 - y0, cw = cx1 - cx0;
    float *acc = (float *)calloc((size_t)W * H * 4, sizeof(float));
    const int64_t ns = (int64_t)cw * (cy1 - cy0) * spp;
    for (int64_t s = 0; s < ns; ++s) {
        const float dx = xy[2 * s] - 0.5f, dy = xy[2 * s + 1] - 0.5f;
        int sx0 = ceil2int(dx - xw), sx1 = floor2int(dx + xw), sy0 = ceil2int(dy - yw), sy1 = floor2int(dy + yw);
        if (sx0 < x0) sx0 = x0;
        if (sx1 > x1 - 1) sx1 = x1 - 1;
        if (sy0 < y0) sy0 = y0;
        if (sy1 > y1 - 1) sy1 = y1 - 1;
        if (sx1 - sx0 < 0 || sy1 - sy0 < 0) continue;
        const float r = rgb[3 * s], g = rgb[3 * s + 1], b = rgb[3 * s + 2];
        float xyz[3];
        xyz[0] = 0.412453f * r + 0.357580f * g + 0.180423f * b;
        xyz[1] = 0.212671f * r + 0.715160f * g + 0.072169f * b;
        xyz[2] = 0.019334f * r + 0.119193f * g + 0.950227f * b;
        for (int y = sy0; y <= sy1; ++y) {
            int iy = floor2int(fabsf(((float)y - dy) * invy * FT));
            if (iy > FT - 1) iy = FT - 1;
            for (int x = sx0; x <= sx1; ++x) {
                int ix = floor2int(fabsf(((float)x - dx) * invx * FT));
                if (ix > FT - 1) ix = FT - 1;
                const float w = table[iy * FT + ix];
                float *p = acc + 4 * ((int64_t)(y - y0) * W + (x - x0));
                p[0] += w * xyz[0]; p[1] += w * xyz[1]; p[2] += w * xyz[2]; p[3] += w;
            }
        }
    }
    for (int64_t i = 0; i < (int64_t)W * H; ++i) {
        const float *p = acc + 4 * i;
        float c[3];
        c[0] = 3.240479f * p[0] - 1.537150f * p[1] - 0.498535f * p[2];
        c[1] = -0.969256f * p[0] + 1.875991f * p[1] + 0.041556f * p[2];
        c[2] = 0.055648f * p[0] - 0.204043f * p[1] + 1.057311f * p[2];
        if (p[3] != 0.f) {
            const float inv = 1.f / p[3];
            for (int a = 0; a < 3; ++a) { c[a] = c[a] * inv; if (!(c[a] > 0.f)) c[a] = 0.f; }
        }
        out[3 * i] = c[0]; out[3 * i + 1] = c[1]; out[3 * i + 2] = c[2];
    }
    free(acc);
}

/* ---- image-plane sample patterns (K14): restatement of the per-pixel offsets of csrc/render.cu pixel_offset ----------
 *   random          two Philox uniforms
 *   stratified      StratifiedSample2D, pbrt-v2/src/core/montecarlo.cpp:158-168 (jitter = Philox uniforms)
 *   lowdiscrepancy  Sample02 = VanDerCorput + Sobol2, pbrt-v2/src/core/montecarlo.h:272-295, scrambles = Philox words
 * Pin: pbrt's own VanDerCorput / Sobol2 / StratifiedSample2D through oracle/ref_film.cpp (tests/test_oracle_film_cpu.py). */
void port_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);
static const float ONE_MINUS_EPS = 0.9999999403953552f;

float port_vdc(uint32_t n, uint32_t scramble) {
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
    n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
    n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
    n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
    n ^= scramble;
    float v = (float)((n >> 8) & 0xffffffu) / (float)(1 << 24);
    return v < ONE_MINUS_EPS ? v : ONE_MINUS_EPS;
}
float port_sobol2(uint32_t n, uint32_t scramble) {
    for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
        if (n & 1u) scramble ^= v;
    float r = (float)((scramble >> 8) & 0xffffffu) / (float)(1 << 24);
    return r < ONE_MINUS_EPS ? r : ONE_MINUS_EPS;
}
static float u01f(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }

/* type 0 random, 1 stratified (nx * ny == spp), 2 lowdiscrepancy; window = sample cells, out xy[cells * spp * 2] */
void port_sample_positions(int type, int nx, int ny, uint64_t seed, int xres, int yres, int x0, int y0, int x1, int y1,
                           int spp, float *out) {
    const uint32_t k0 = (uint32_t)seed ^ 0x43414d52u, k1 = (uint32_t)(seed >> 32);
    int64_t s = 0;
    for (int py = y0; py < y1; ++py)
        for (int px = x0; px < x1; ++px) {
            uint32_t gpix = (uint32_t)py * (uint32_t)xres + (uint32_t)px;
            if (px < 0 || py < 0 || px >= xres || py >= yres)
                gpix = (uint32_t)xres * (uint32_t)yres + (uint32_t)(py + 64) * (uint32_t)(xres + 128) + (uint32_t)(px + 64);
            for (int k = 0; k < spp; ++k, ++s) {
                uint32_t r[4];
                float ox, oy;
                if (type == 2) {
                    port_philox(gpix, 0xffffffffu, 1u, 0u, k0, k1, r);
                    ox = port_vdc((uint32_t)k, r[0]); oy = port_sobol2((uint32_t)k, r[1]);
                } else {
                    port_philox(gpix, (uint32_t)k, 0u, 0u, k0, k1, r);
                    if (type == 1) {
                        const float dx = 1.f / (float)nx, dy = 1.f / (float)ny;
                        const int sx = k % nx, sy = k / nx;
                        ox = ((float)sx + u01f(r[0])) * dx; oy = ((float)sy + u01f(r[1])) * dy;
                        if (!(ox < ONE_MINUS_EPS)) ox = ONE_MINUS_EPS;
                        if (!(oy < ONE_MINUS_EPS)) oy = ONE_MINUS_EPS;
                    } else { ox = u01f(r[0]); oy = u01f(r[1]); }
                }
                out[2 * s] = (float)px + ox; out[2 * s + 1] = (float)py + oy;
            }
        }
}
