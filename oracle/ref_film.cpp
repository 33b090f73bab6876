// TEST INFRASTRUCTURE - C hooks around the reference's OWN film (pbrt-v2 ImageFilm + its pixel filters), compiled from
// /root/reference into oracle/_ref/libpbrt_oracle.so.  Used to pin oracle_film.c and to generate tests/golden/film.npz.
// The image leaves pbrt the only way ImageFilm offers: WriteImage() to a .pfm file (core/imageio.cpp:404-441).
#include "pbrt.h"
#include "film/image.h"
#include "filters/box.h"
#include "filters/triangle.h"
#include "filters/gaussian.h"
#include "filters/mitchell.h"
#include "filters/sinc.h"
#include "sampler.h"
#include "spectrum.h"

extern "C" {

// type: 0 box, 1 triangle, 2 gaussian (p0 = alpha), 3 mitchell (p0 = B, p1 = C), 4 lanczos sinc (p0 = tau).
// crop = {x0/xres, x1/xres, y0/yres, y1/yres}; samples: n x {imageX, imageY, r, g, b}; writes pfm_path.
int ref_film(int type, float xw, float yw, float p0, float p1, int xres, int yres, const float crop[4], long long n,
             const float *samples, const char *pfm_path) {
    Filter *f;
    switch (type) {
    case 1: f = new TriangleFilter(xw, yw); break;
    case 2: f = new GaussianFilter(xw, yw, p0); break;
    case 3: f = new MitchellFilter(p0, p1, xw, yw); break;
    case 4: f = new LanczosSincFilter(xw, yw, p0); break;
    default: f = new BoxFilter(xw, yw); break;
    }
    ImageFilm film(xres, yres, f, crop, pfm_path, false);
    for (long long i = 0; i < n; ++i) {
        CameraSample cs;
        cs.imageX = samples[5 * i]; cs.imageY = samples[5 * i + 1];
        cs.lensU = cs.lensV = 0.5f; cs.time = 0.f;
        float rgb[3] = {samples[5 * i + 2], samples[5 * i + 3], samples[5 * i + 4]};
        film.AddSample(cs, Spectrum::FromRGB(rgb));
    }
    film.WriteImage(1.f);
    return 0;
}

}

// pbrt's own (0,2)-sequence point and stratified strata centres (core/montecarlo.h:272-295, montecarlo.cpp:158-168)
#include "montecarlo.h"
#include "rng.h"
extern "C" {
void ref_sample02(unsigned n, unsigned s0, unsigned s1, float out[2]) {
    uint32_t sc[2] = {s0, s1};
    Sample02(n, sc, out);
}
void ref_stratified_centres(int nx, int ny, float *out) {
    RNG rng(1);
    StratifiedSample2D(out, nx, ny, rng, false);
}
}
