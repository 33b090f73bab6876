// K13 film reconstruction with pbrt's pixel filters.
//
// Replaces the reference's CPU film loop - camera->film->AddSample for every launch sample
// (cuda_render/photon_mapping/photonmappingrenderer.cpp:247-273, simple_render/simplerender.cpp:84-100) - i.e. pbrt's
// ImageFilm::AddSample + the normalisation of ImageFilm::WriteImage (pbrt-v2/src/film/image.cpp:77-137, 178-213) with
// the 16 x 16 filter weight table of the ImageFilm constructor (image.cpp:56-69) and the box / triangle / gaussian /
// mitchell / lanczos-sinc filters (pbrt-v2/src/filters/*.cpp).
//
// pbrt scatters: every sample adds itself to the pixels inside the filter support, in sample arrival order, with
// atomics when tasks overlap.  Here a pixel GATHERS: one thread per pixel walks the sample cells within the support
// in (row, column, sample) order and tests each sample with pbrt's own extent arithmetic (Ceil2Int / Floor2Int of
// imageX -+ width), so the sum order is fixed, no atomics are needed and the image is bit-identical from run to run
// and for every image tiling.  Contract: samples arrive cell-major (spp per cell, row-major cells) and lie inside
// their cell - what a per-pixel sampler produces.
#include "b2rt_internal.cuh"

#include <cmath>

namespace {

constexpr int FT = 16;       // FILTER_TABLE_SIZE, image.cpp:57

// ---- filter functions, evaluated on the host for the table only -----------------------------------------------------
float gauss1(float d, float alpha, float expv) {             // filters/gaussian.h:56-58
    float v = (float)(expf(-alpha * d * d) - expv);
    return v > 0.f ? v : 0.f;
}
float mitchell1(float x, float B, float C) {                 // filters/mitchell.h:50-59
    x = fabsf(2.f * x);
    if (x > 1.f)
        return ((-B - 6 * C) * x * x * x + (6 * B + 30 * C) * x * x + (-12 * B - 48 * C) * x + (8 * B + 24 * C)) * (1.f / 6.f);
    return ((12 - 9 * B - 6 * C) * x * x * x + (-18 + 12 * B + 6 * C) * x * x + (6 - 2 * B)) * (1.f / 6.f);
}
float sinc1(float x, float tau) {                            // filters/sinc.h:49-57 (pbrt's M_PI is a float literal)
    x = fabsf(x);
    if (x < 1e-5) return 1.f;
    if (x > 1.) return 0.f;
    x *= 3.14159265358979323846f;
    float sinc = sinf(x) / x;
    float lanczos = sinf(x * tau) / (x * tau);
    return sinc * lanczos;
}
float filter_eval(const b2rt_filter &f, float x, float y) {
    switch (f.type) {
    case B2RT_FILTER_TRIANGLE: {
        float a = f.xwidth - fabsf(x), b = f.ywidth - fabsf(y);
        return (a > 0.f ? a : 0.f) * (b > 0.f ? b : 0.f);
    }
    case B2RT_FILTER_GAUSSIAN:
        return gauss1(x, f.p0, expf(-f.p0 * f.xwidth * f.xwidth)) * gauss1(y, f.p0, expf(-f.p0 * f.ywidth * f.ywidth));
    case B2RT_FILTER_MITCHELL: return mitchell1(x * (1.f / f.xwidth), f.p0, f.p1) * mitchell1(y * (1.f / f.ywidth), f.p0, f.p1);
    case B2RT_FILTER_SINC: return sinc1(x * (1.f / f.xwidth), f.p0) * sinc1(y * (1.f / f.ywidth), f.p0);
    default: return 1.f;
    }
}

struct FilmTable { float w[FT * FT]; };

struct FilmArgs {
    float xw, yw, invx, invy;
    int hx, hy;                      // sample cells beyond a pixel that can reach it
    int cx0, cy0, cx1, cy1, spp;     // sample cells
    int x0, y0, x1, y1;              // pixel window (the film's crop window)
};

__global__ void __launch_bounds__(128)
film_filter_kernel(const FilmTable tab, const FilmArgs a, const float2 *__restrict__ xy, const float *__restrict__ rgb,
                   float *__restrict__ out) {
    __shared__ float s_tab[FT * FT];
    for (int i = threadIdx.x; i < FT * FT; i += blockDim.x) s_tab[i] = tab.w[i];
    __syncthreads();
    const int W = a.x1 - a.x0;
    const int64_t lp = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= (int64_t)W * (a.y1 - a.y0)) return;
    const int x = a.x0 + (int)(lp % W), y = a.y0 + (int)(lp / W);
    const int ca = max(a.cx0, x - a.hx), cb = min(a.cx1 - 1, x + a.hx);
    const int ra = max(a.cy0, y - a.hy), rb = min(a.cy1 - 1, y + a.hy);
    const int cw = a.cx1 - a.cx0;
    float X = 0.f, Y = 0.f, Z = 0.f, wsum = 0.f;
    for (int cy = ra; cy <= rb; ++cy) {
        for (int cx = ca; cx <= cb; ++cx) {
            const int64_t s0 = ((int64_t)(cy - a.cy0) * cw + (cx - a.cx0)) * a.spp;
            for (int k = 0; k < a.spp; ++k) {
                const float2 p = __ldg(xy + s0 + k);
                const float dx = p.x - 0.5f, dy = p.y - 0.5f;                // image.cpp:80-85
                if (x < (int)ceilf(dx - a.xw) || x > (int)floorf(dx + a.xw) || y < (int)ceilf(dy - a.yw) ||
                    y > (int)floorf(dy + a.yw))
                    continue;
                int ix = (int)floorf(fabsf(((float)x - dx) * a.invx * FT));   // image.cpp:102-113
                int iy = (int)floorf(fabsf(((float)y - dy) * a.invy * FT));
                ix = min(ix, FT - 1); iy = min(iy, FT - 1);
                const float w = s_tab[iy * FT + ix];
                const float r = __ldg(rgb + 3 * (s0 + k)), g = __ldg(rgb + 3 * (s0 + k) + 1), b = __ldg(rgb + 3 * (s0 + k) + 2);
                const float sx = 0.412453f * r + 0.357580f * g + 0.180423f * b;      // Spectrum::ToXYZ, core/spectrum.h:58-62
                const float sy = 0.212671f * r + 0.715160f * g + 0.072169f * b;
                const float sz = 0.019334f * r + 0.119193f * g + 0.950227f * b;
                X += w * sx; Y += w * sy; Z += w * sz; wsum += w;
            }
        }
    }
    float c[3];                                                               // image.cpp:186-196, core/spectrum.h:51-55
    c[0] = 3.240479f * X - 1.537150f * Y - 0.498535f * Z;
    c[1] = -0.969256f * X + 1.875991f * Y + 0.041556f * Z;
    c[2] = 0.055648f * X - 0.204043f * Y + 1.057311f * Z;
    if (wsum != 0.f) {
        const float inv = 1.f / wsum;
#pragma unroll
        for (int j = 0; j < 3; ++j) { c[j] = c[j] * inv; if (!(c[j] > 0.f)) c[j] = 0.f; }
    }
    out[3 * lp] = c[0]; out[3 * lp + 1] = c[1]; out[3 * lp + 2] = c[2];
}

// Tiled version (default): a block owns a 32 x 8 pixel tile and walks the cell rows that reach it from top to bottom.
// Each row's samples are staged ONCE in shared memory, already turned into what every pixel needs from them - the
// shifted raster position and the XYZ radiance - laid out [sample k][cell] with an odd row pitch, so the staging stores
// and the per-pixel reads (32 lanes = 32 consecutive cells, same k) are both conflict-free.  A pixel still sums its
// samples in (cell row, cell column, k) order with the expressions of film_filter_kernel: same bits.
// The one-thread-per-pixel kernel read every sample from global memory (2h+1)^2 times with one dependent load chain
// per thread: 29 ms for a 3840 x 2160 x 16 spp Mitchell frame, 60 warps stalled on loads per issue (run r02e).
constexpr int FILM_TW = 32, FILM_TH = 8;

__global__ void __launch_bounds__(FILM_TW * FILM_TH)
film_tile_kernel(const FilmTable tab, const FilmArgs a, const float2 *__restrict__ xy, const float *__restrict__ rgb,
                 float *__restrict__ out, int pitch) {
    extern __shared__ float s_dyn[];
    __shared__ float s_tab[FT * FT];
    const int tid = threadIdx.y * FILM_TW + threadIdx.x;
    for (int i = tid; i < FT * FT; i += FILM_TW * FILM_TH) s_tab[i] = tab.w[i];
    const int W = a.x1 - a.x0, cw = a.cx1 - a.cx0;
    const int tx0 = a.x0 + blockIdx.x * FILM_TW, ty0 = a.y0 + blockIdx.y * FILM_TH;
    const int x = tx0 + threadIdx.x, y = ty0 + threadIdx.y;
    const bool live = x < a.x1 && y < a.y1;
    const int col0 = max(a.cx0, tx0 - a.hx), col1 = min(a.cx1 - 1, min(tx0 + FILM_TW - 1, a.x1 - 1) + a.hx);    // staged cell columns
    const int ncol = col1 - col0 + 1;
    const int row0 = max(a.cy0, ty0 - a.hy), row1 = min(a.cy1 - 1, min(ty0 + FILM_TH - 1, a.y1 - 1) + a.hy);
    const int ca = max(col0, x - a.hx), cb = min(col1, x + a.hx);
    const int plane = pitch * a.spp;            // floats per staged quantity
    float *s_dx = s_dyn, *s_dy = s_dyn + plane, *s_x = s_dyn + 2 * plane, *s_y = s_dyn + 3 * plane, *s_z = s_dyn + 4 * plane;
    const float fx = (float)x, fy = (float)y;
    float X = 0.f, Y = 0.f, Z = 0.f, wsum = 0.f;
    for (int cy = row0; cy <= row1; ++cy) {
        __syncthreads();                        // the previous row is consumed (first pass: s_tab is written)
        const int64_t rowbase = ((int64_t)(cy - a.cy0) * cw + (col0 - a.cx0)) * a.spp;
        const int nrow = ncol * a.spp;
        for (int i = tid; i < nrow; i += FILM_TW * FILM_TH) {
            const int cell = i / a.spp, k = i - cell * a.spp;
            const float2 p = __ldg(xy + rowbase + i);
            const float r = __ldg(rgb + 3 * (rowbase + i)), g = __ldg(rgb + 3 * (rowbase + i) + 1),
                        b = __ldg(rgb + 3 * (rowbase + i) + 2);
            const int o = k * pitch + cell;
            s_dx[o] = p.x - 0.5f; s_dy[o] = p.y - 0.5f;
            s_x[o] = 0.412453f * r + 0.357580f * g + 0.180423f * b;
            s_y[o] = 0.212671f * r + 0.715160f * g + 0.072169f * b;
            s_z[o] = 0.019334f * r + 0.119193f * g + 0.950227f * b;
        }
        __syncthreads();
        if (!live || cy < y - a.hy || cy > y + a.hy) continue;
        for (int cx = ca; cx <= cb; ++cx) {
            const int c = cx - col0;
            for (int k = 0; k < a.spp; ++k) {
                const int o = k * pitch + c;
                const float dx = s_dx[o], dy = s_dy[o];
                // x >= Ceil2Int(v) <=> x >= v and x <= Floor2Int(v) <=> x <= v for an integer x
                if (fx < dx - a.xw || fx > dx + a.xw || fy < dy - a.yw || fy > dy + a.yw) continue;
                int ix = (int)floorf(fabsf((fx - dx) * a.invx * FT));
                int iy = (int)floorf(fabsf((fy - dy) * a.invy * FT));
                ix = min(ix, FT - 1); iy = min(iy, FT - 1);
                const float w = s_tab[iy * FT + ix];
                X += w * s_x[o]; Y += w * s_y[o]; Z += w * s_z[o]; wsum += w;
            }
        }
    }
    if (!live) return;
    float c[3];
    c[0] = 3.240479f * X - 1.537150f * Y - 0.498535f * Z;
    c[1] = -0.969256f * X + 1.875991f * Y + 0.041556f * Z;
    c[2] = 0.055648f * X - 0.204043f * Y + 1.057311f * Z;
    if (wsum != 0.f) {
        const float inv = 1.f / wsum;
#pragma unroll
        for (int j = 0; j < 3; ++j) { c[j] = c[j] * inv; if (!(c[j] > 0.f)) c[j] = 0.f; }
    }
    const int64_t lp = (int64_t)(y - a.y0) * W + (x - a.x0);
    out[3 * lp] = c[0]; out[3 * lp + 1] = c[1]; out[3 * lp + 2] = c[2];
}

}  // namespace

bool b2rt_filter_valid(const b2rt_filter *f) {
    return f && f->type >= B2RT_FILTER_BOX && f->type <= B2RT_FILTER_SINC && f->xwidth > 0.f && f->ywidth > 0.f &&
           f->xwidth <= 16.f && f->ywidth <= 16.f;
}
void b2rt_filter_halo(const b2rt_filter *f, int *hx, int *hy) {
    *hx = (int)floorf(f->xwidth + 0.5f);
    *hy = (int)floorf(f->ywidth + 0.5f);
}

// all pointers are device pointers
int b2rt_film_impl(b2rt_ctx *ctx, const b2rt_filter *f, const int32_t cells[4], int32_t spp, const float *d_xy,
                   const float *d_rgb, const int32_t pixels[4], float *d_out) {
    FilmTable tab;
    for (int y = 0; y < FT; ++y) {
        const float fy = ((float)y + .5f) * f->ywidth / FT;
        for (int x = 0; x < FT; ++x) {
            const float fx = ((float)x + .5f) * f->xwidth / FT;
            tab.w[y * FT + x] = filter_eval(*f, fx, fy);
        }
    }
    FilmArgs a;
    a.xw = f->xwidth; a.yw = f->ywidth; a.invx = 1.f / f->xwidth; a.invy = 1.f / f->ywidth;
    b2rt_filter_halo(f, &a.hx, &a.hy);
    a.cx0 = cells[0]; a.cy0 = cells[1]; a.cx1 = cells[2]; a.cy1 = cells[3]; a.spp = spp;
    a.x0 = pixels[0]; a.y0 = pixels[1]; a.x1 = pixels[2]; a.y1 = pixels[3];
    const int64_t np = (int64_t)(a.x1 - a.x0) * (a.y1 - a.y0);
    // tiled kernel when one staged cell row fits in shared memory (5 floats per sample), else one thread per pixel
    const int pitch = (FILM_TW + 2 * a.hx) | 1;
    const size_t smem = (size_t)pitch * spp * 5 * sizeof(float);
    if (ctx->film_variant != 0 && smem <= 160 * 1024) {
        if (smem > 48 * 1024 && smem > ctx->film_smem_opt_in) {
            B2RT_CUDA(ctx, cudaFuncSetAttribute(film_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            ctx->film_smem_opt_in = smem;
        }
        dim3 grid((unsigned)((a.x1 - a.x0 + FILM_TW - 1) / FILM_TW), (unsigned)((a.y1 - a.y0 + FILM_TH - 1) / FILM_TH));
        film_tile_kernel<<<grid, dim3(FILM_TW, FILM_TH), smem, ctx->stream>>>(tab, a, reinterpret_cast<const float2 *>(d_xy),
                                                                              d_rgb, d_out, pitch);
    } else {
        film_filter_kernel<<<(unsigned)((np + 127) / 128), 128, 0, ctx->stream>>>(tab, a, reinterpret_cast<const float2 *>(d_xy),
                                                                                 d_rgb, d_out);
    }
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

extern "C" {

int b2rt_film_set_filter(b2rt_ctx *ctx, const b2rt_filter *f) {
    if (!ctx) return B2RT_E_ARG;
    if (!f) { ctx->film_filter = b2rt_filter{B2RT_FILTER_BOX, 0.5f, 0.5f, 0.f, 0.f}; return B2RT_OK; }
    if (!b2rt_filter_valid(f)) { ctx->err = "film_set_filter: bad filter"; return B2RT_E_ARG; }
    ctx->film_filter = *f;
    return B2RT_OK;
}

int b2rt_sampler_set(b2rt_ctx *ctx, int32_t type, int32_t xsamples, int32_t ysamples) {
    if (!ctx) return B2RT_E_ARG;
    if (type < B2RT_SAMPLER_RANDOM || type > B2RT_SAMPLER_LOWDISCREPANCY ||
        (type == B2RT_SAMPLER_STRATIFIED && (xsamples < 1 || ysamples < 1))) {
        ctx->err = "sampler_set: bad arguments"; return B2RT_E_ARG;
    }
    ctx->sampler_type = type;
    ctx->sampler_nx = type == B2RT_SAMPLER_STRATIFIED ? xsamples : 1;
    ctx->sampler_ny = type == B2RT_SAMPLER_STRATIFIED ? ysamples : 1;
    return B2RT_OK;
}

int b2rt_sample_positions(b2rt_ctx *ctx, int32_t xres, int32_t yres, const int32_t window[4], int32_t spp, float *xy,
                          int where) {
    if (!ctx) return B2RT_E_ARG;
    if (!window || !xy || spp < 1 || xres < 1 || yres < 1 || window[0] >= window[2] || window[1] >= window[3] ||
        window[0] < -64 || window[1] < -64 || window[2] > xres + 64 || window[3] > yres + 64) {
        ctx->err = "sample_positions: bad arguments"; return B2RT_E_ARG;
    }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)(window[2] - window[0]) * (window[3] - window[1]) * spp;
    float *d_xy = xy;
    if (where != B2RT_DEVICE) {
        B2RT_CUDA(ctx, ctx->d_film_xy.reserve((size_t)n * 8));
        d_xy = ctx->d_film_xy.as<float>();
    }
    int rc = b2rt_sample_positions_impl(ctx, xres, yres, window, spp, d_xy);
    if (rc) return rc;
    if (where != B2RT_DEVICE) {
        B2RT_CUDA(ctx, cudaMemcpyAsync(xy, d_xy, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
        B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return B2RT_OK;
}

int b2rt_film_reconstruct(b2rt_ctx *ctx, const b2rt_filter *f, const int32_t cells[4], int32_t spp, const float *xy,
                          const float *rgb, const int32_t pixels[4], float *out_rgb, int where) {
    if (!ctx) return B2RT_E_ARG;
    B2RT_STAGE(ctx, B2RT_STAGE_FILM);
    if (!b2rt_filter_valid(f) || !cells || !pixels || spp < 1 || !xy || !rgb || !out_rgb || cells[0] >= cells[2] ||
        cells[1] >= cells[3] || pixels[0] >= pixels[2] || pixels[1] >= pixels[3]) {
        ctx->err = "film_reconstruct: bad arguments"; return B2RT_E_ARG;
    }
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t ns = (int64_t)(cells[2] - cells[0]) * (cells[3] - cells[1]) * spp;
    const int64_t np = (int64_t)(pixels[2] - pixels[0]) * (pixels[3] - pixels[1]);
    const float *d_xy = xy, *d_rgb = rgb;
    float *d_out = out_rgb;
    if (where != B2RT_DEVICE) {
        B2RT_CUDA(ctx, ctx->d_film_xy.reserve((size_t)ns * 8));
        B2RT_CUDA(ctx, ctx->d_film_rgb.reserve((size_t)ns * 12));
        B2RT_CUDA(ctx, ctx->d_frame.reserve((size_t)np * 12));
        B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_film_xy.p, xy, (size_t)ns * 8, cudaMemcpyHostToDevice, ctx->stream));
        B2RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_film_rgb.p, rgb, (size_t)ns * 12, cudaMemcpyHostToDevice, ctx->stream));
        d_xy = ctx->d_film_xy.as<float>(); d_rgb = ctx->d_film_rgb.as<float>(); d_out = ctx->d_frame.as<float>();
    }
    int rc = b2rt_film_impl(ctx, f, cells, spp, d_xy, d_rgb, pixels, d_out);
    if (rc) return rc;
    if (where != B2RT_DEVICE) {
        B2RT_CUDA(ctx, cudaMemcpyAsync(out_rgb, d_out, (size_t)np * 12, cudaMemcpyDeviceToHost, ctx->stream));
        B2RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return B2RT_OK;
}

}
