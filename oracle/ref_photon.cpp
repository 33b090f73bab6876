// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// extern "C" face over the UNMODIFIED reference CPU photon mapper (pbrt-v2 PhotonIntegrator), compiled from
// /root/reference/pbrt-v2/src where it lies.  integrators/photonmap.cpp keeps Photon, PhotonShootingTask and
// LPhoton file-local, so that one source file is compiled a second time INSIDE namespace refpm (by #include from
// its original location -- nothing is copied), with `private` opened so the harness can read the maps.
//
//   ref_pm_shoot     PhotonIntegrator::Preprocess           integrators/photonmap.cpp:339-402 (tasks :405-582)
//   ref_pm_lphoton   KdTree<Photon>::Lookup + LPhoton        core/kdtree.h:151-184, photonmap.cpp:187-260
// KdTree's node data is reached through the fork's own `friend class PhotonMappingRenderer` (core/kdtree.h:84).
#include "pbrt.h"
#include "stdafx.h"
#include "integrator.h"
#include "kdtree.h"
#include "scene.h"
#include "montecarlo.h"
#include "sampler.h"
#include "progressreporter.h"
#include "intersection.h"
#include "paramset.h"
#include "camera.h"
#include "renderer.h"
#include "reflection.h"
#include "light.h"
#include "primitive.h"
#include "parallel.h"
#include "shapes/trianglemesh.h"
#include "shapes/sphere.h"
#include "shapes/disk.h"
#include "accelerators/bvh.h"
#include "materials/matte.h"
#include "materials/mirror.h"
#include "materials/glass.h"
#include "textures/constant.h"
#include "lights/point.h"
#include "lights/diffuse.h"
#include <atomic>
#include <thread>
#include <chrono>
#include <algorithm>
#include <vector>

namespace refpm {
#define private public
#include "integrators/photonmap.cpp"
#undef private
}  // namespace refpm

// the fork declares this class a friend of KdTree (core/kdtree.h:84): use it to read the flattened tree
class PhotonMappingRenderer {
public:
    template <typename T> static uint32_t count(const KdTree<T> *t) { return t ? t->nNodes : 0; }
    template <typename T> static const T *data(const KdTree<T> *t) { return t->nodeData; }
};

namespace {
struct StubRenderer : public Renderer {
    void Render(const Scene *) {}
    Spectrum Li(const Scene *, const RayDifferential &, const Sample *, RNG &, MemoryArena &, Intersection *,
                Spectrum *) const { return 0.f; }
    Spectrum Transmittance(const Scene *, const RayDifferential &, const Sample *, RNG &, MemoryArena &) const {
        return 1.f;
    }
};

struct PmScene {
    std::vector<Reference<Material> > materials;
    std::vector<Reference<Primitive> > prims;
    std::vector<Light *> lights;
    Reference<Material> dflt;
    Scene *scene;
    PmScene() : scene(NULL) {
        dflt = new MatteMaterial(new ConstantTexture<Spectrum>(Spectrum(0.5f)), new ConstantTexture<float>(0.f), NULL);
    }
    Reference<Material> mat(int id) { return (id >= 0 && id < (int)materials.size()) ? materials[id] : dflt; }
};
Matrix4x4 m12(const float m[12]) {
    return Matrix4x4(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], 0, 0, 0, 1);
}
Spectrum rgb(const float *c) { float v[3] = {c[0], c[1], c[2]}; return Spectrum::FromRGB(v); }
}  // namespace

extern "C" {

void *ref_pm_create() { return new PmScene(); }

// type 0 matte (Kd), 1 mirror (Kr), 2 glass (Kr, Kt, index) -- the materials cuda_render translates
int ref_pm_add_material(void *h, int type, const float kd[3], const float kt[3], float ior) {
    PmScene *s = (PmScene *)h;
    Reference<Texture<Spectrum> > a = new ConstantTexture<Spectrum>(rgb(kd));
    if (type == 0) s->materials.push_back(new MatteMaterial(a, new ConstantTexture<float>(0.f), NULL));
    else if (type == 1) s->materials.push_back(new MirrorMaterial(a, NULL));
    else s->materials.push_back(new GlassMaterial(a, new ConstantTexture<Spectrum>(rgb(kt)),
                                                  new ConstantTexture<float>(ior), NULL));
    return (int)s->materials.size() - 1;
}

void ref_pm_add_mesh(void *h, int nverts, const float *P, int ntris, const int *idx, int material) {
    PmScene *s = (PmScene *)h;
    Transform *I = new Transform();
    Reference<Shape> mesh = new TriangleMesh(I, I, false, ntris, nverts, idx, (const Point *)P, NULL, NULL, NULL,
                                             Reference<Texture<float> >());
    s->prims.push_back(new GeometricPrimitive(mesh, s->mat(material), NULL));
}

void ref_pm_add_sphere(void *h, const float o2w[12], const float w2o[12], float radius, int material) {
    PmScene *s = (PmScene *)h;
    Transform *t = new Transform(m12(o2w), m12(w2o)), *ti = new Transform(m12(w2o), m12(o2w));
    s->prims.push_back(new GeometricPrimitive(new Sphere(t, ti, false, radius, -radius, radius, 360.f),
                                              s->mat(material), NULL));
}

// L == NULL: plain disk; else a DiffuseAreaLight with emitted radiance L and nsamples
void ref_pm_add_disk(void *h, const float o2w[12], const float w2o[12], float height, float radius, float inner,
                     int material, const float *L, int nsamples) {
    PmScene *s = (PmScene *)h;
    Transform *t = new Transform(m12(o2w), m12(w2o)), *ti = new Transform(m12(w2o), m12(o2w));
    Reference<Shape> d = new Disk(t, ti, false, height, radius, inner, 360.f);
    AreaLight *al = NULL;
    if (L) {
        al = new DiffuseAreaLight(*t, rgb(L), nsamples, d);
        s->lights.push_back(al);
    }
    s->prims.push_back(new GeometricPrimitive(d, s->mat(material), al));
}

void ref_pm_add_point_light(void *h, const float p[3], const float I[3]) {
    PmScene *s = (PmScene *)h;
    s->lights.push_back(new PointLight(Translate(Vector(p[0], p[1], p[2])), rgb(I)));
}

void ref_pm_build(void *h) {
    PmScene *s = (PmScene *)h;
    ParamSet ps;
    Primitive *accel = CreateBVHAccelerator(s->prims, ps);
    s->scene = new Scene(accel, s->lights, NULL);
}

// Runs pbrt's own photon shooting.  out_*: 9 floats per photon {p, alpha(rgb), wi}.  Returns 0 or -1 (capacity).
// npaths[2] = {nCausticPaths, nIndirectPaths}, nstored[2].  ncores <= 0: all cores.
int ref_pm_shoot(void *h, int n_caustic, int n_indirect, int max_depth, int ncores, float *out_c, int64_t cap_c,
                 float *out_i, int64_t cap_i, int64_t npaths[2], int64_t nstored[2]) {
    PmScene *s = (PmScene *)h;
    PbrtOptions.nCores = ncores > 0 ? ncores : 0;
    PbrtOptions.quiet = true;
    refpm::PhotonIntegrator pi(n_caustic, n_indirect, 50, 5, max_depth, 0.1f, false, 32, 10.f);
    StubRenderer r;
    pi.Preprocess(s->scene, NULL, &r);
    TasksCleanup();
    const KdTree<refpm::Photon> *maps[2] = {pi.causticMap, pi.indirectMap};
    float *outs[2] = {out_c, out_i};
    int64_t caps[2] = {cap_c, cap_i};
    npaths[0] = pi.nCausticPaths; npaths[1] = pi.nIndirectPaths;
    for (int m = 0; m < 2; ++m) {
        uint32_t n = PhotonMappingRenderer::count(maps[m]);
        nstored[m] = n;
        if ((int64_t)n > caps[m]) return -1;
        for (uint32_t i = 0; i < n; ++i) {
            const refpm::Photon &p = PhotonMappingRenderer::data(maps[m])[i];
            float c[3];
            p.alpha.ToRGB(c);
            float *o = outs[m] + 9 * (int64_t)i;
            o[0] = p.p.x; o[1] = p.p.y; o[2] = p.p.z; o[3] = c[0]; o[4] = c[1]; o[5] = c[2];
            o[6] = p.wi.x; o[7] = p.wi.y; o[8] = p.wi.z;
        }
    }
    return 0;
}

// pbrt's estimator on caller-supplied photons: KdTree<Photon> build + Lookup + LPhoton for a Lambertian surface.
// photons: 9 floats each {p, alpha, wi}; queries: 12 floats {p, n (shading normal), kd, pad}; wo = n.
void ref_pm_lphoton(int64_t n, const float *photons, int n_paths, int64_t nq, const float *queries, int k,
                    float max_dist2, float *rgb_out) {
    std::vector<refpm::Photon> v((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const float *p = photons + 9 * i;
        v[i] = refpm::Photon(Point(p[0], p[1], p[2]), rgb(p + 3), Vector(p[6], p[7], p[8]));
    }
    KdTree<refpm::Photon> *map = n ? new KdTree<refpm::Photon>(v) : NULL;
    std::vector<refpm::ClosePhoton> buf((size_t)(k > 0 ? k : 1));
    MemoryArena arena;
    RNG rng(1);
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        Normal nn(Q[3], Q[4], Q[5]);
        Vector dpdu, dpdv;
        CoordinateSystem(Vector(nn), &dpdu, &dpdv);
        DifferentialGeometry dg(Point(Q[0], Q[1], Q[2]), dpdu, dpdv, Normal(0, 0, 0), Normal(0, 0, 0), 0.f, 0.f, NULL);
        dg.nn = nn;
        BSDF *bsdf = BSDF_ALLOC(arena, BSDF)(dg, nn);
        bsdf->Add(BSDF_ALLOC(arena, Lambertian)(rgb(Q + 6)));
        Intersection isect;
        isect.dg = dg;
        Spectrum L = refpm::LPhoton(map, n_paths, k, &buf[0], bsdf, rng, isect, Vector(nn), max_dist2);
        L.ToRGB(rgb_out + 3 * q);
        arena.FreeAll();
    }
    delete map;
}

// The same estimator as a TIMED, multi-threaded loop (the CPU baseline of the gather stage: bench.py): kd-tree build and
// lookups timed apart; ranges of 1024 queries handed out by an atomic cursor, one arena / lookup buffer per thread.
void ref_pm_lphoton_timed(int64_t n, const float *photons, int n_paths, int64_t nq, const float *queries, int k,
                          float max_dist2, int nthreads, float *rgb_out, double *build_seconds, double *lookup_seconds) {
    std::vector<refpm::Photon> v((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const float *p = photons + 9 * i;
        v[i] = refpm::Photon(Point(p[0], p[1], p[2]), rgb(p + 3), Vector(p[6], p[7], p[8]));
    }
    auto t0 = std::chrono::steady_clock::now();
    KdTree<refpm::Photon> *map = n ? new KdTree<refpm::Photon>(v) : NULL;
    auto t1 = std::chrono::steady_clock::now();
    std::atomic<int64_t> cursor(0);
    auto work = [&]() {
        std::vector<refpm::ClosePhoton> buf((size_t)(k > 0 ? k : 1));
        MemoryArena arena;
        RNG rng(1);
        for (;;) {
            const int64_t lo = cursor.fetch_add(1024);
            if (lo >= nq) break;
            for (int64_t q = lo; q < std::min(nq, lo + 1024); ++q) {
                const float *Q = queries + 12 * q;
                Normal nn(Q[3], Q[4], Q[5]);
                Vector dpdu, dpdv;
                CoordinateSystem(Vector(nn), &dpdu, &dpdv);
                DifferentialGeometry dg(Point(Q[0], Q[1], Q[2]), dpdu, dpdv, Normal(0, 0, 0), Normal(0, 0, 0), 0.f, 0.f, NULL);
                dg.nn = nn;
                BSDF *bsdf = BSDF_ALLOC(arena, BSDF)(dg, nn);
                bsdf->Add(BSDF_ALLOC(arena, Lambertian)(rgb(Q + 6)));
                Intersection isect;
                isect.dg = dg;
                Spectrum L = refpm::LPhoton(map, n_paths, k, &buf[0], bsdf, rng, isect, Vector(nn), max_dist2);
                L.ToRGB(rgb_out + 3 * q);
                arena.FreeAll();
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < std::max(1, nthreads); ++t) th.emplace_back(work);
    for (auto &t : th) t.join();
    auto t2 = std::chrono::steady_clock::now();
    delete map;
    if (build_seconds) *build_seconds = std::chrono::duration<double>(t1 - t0).count();
    if (lookup_seconds) *lookup_seconds = std::chrono::duration<double>(t2 - t1).count();
}

// pbrt's irradiance estimate on caller-supplied photons: KdTree<Photon> build + Lookup + EPhoton (photonmap.cpp:263-283).
// photons: 9 floats each {p, alpha, wi}; queries: 12 floats {p, n, -, -}
void ref_pm_ephoton(int64_t n, const float *photons, int n_paths, int64_t nq, const float *queries, int k,
                    float max_dist2, float *E_out) {
    std::vector<refpm::Photon> v((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const float *p = photons + 9 * i;
        v[i] = refpm::Photon(Point(p[0], p[1], p[2]), rgb(p + 3), Vector(p[6], p[7], p[8]));
    }
    KdTree<refpm::Photon> *map = n ? new KdTree<refpm::Photon>(v) : NULL;
    std::vector<refpm::ClosePhoton> buf((size_t)(k > 0 ? k : 1));
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        Spectrum E = refpm::EPhoton(map, n_paths, k, &buf[0], max_dist2, Point(Q[0], Q[1], Q[2]), Normal(Q[3], Q[4], Q[5]));
        E.ToRGB(E_out + 3 * q);
    }
    delete map;
}

// pbrt's radiance-photon lookup: KdTree<RadiancePhoton> + RadiancePhotonProcess with md2 = INFINITY, exactly as the
// final gather calls it (photonmap.cpp:156-172, 688-693).  photons: 9 floats each {p, n, Lo}; out: Lo of the photon found
void ref_pm_radiance_lookup(int64_t n, const float *photons, int64_t nq, const float *queries, float *rgb_out) {
    std::vector<refpm::RadiancePhoton> v((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const float *p = photons + 9 * i;
        v[i] = refpm::RadiancePhoton(Point(p[0], p[1], p[2]), Normal(p[3], p[4], p[5]));
        v[i].Lo = rgb(p + 6);
    }
    KdTree<refpm::RadiancePhoton> *map = n ? new KdTree<refpm::RadiancePhoton>(v) : NULL;
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        Normal nn(Q[3], Q[4], Q[5]);
        refpm::RadiancePhotonProcess proc(nn);
        float md2 = INFINITY;
        if (map) map->Lookup(Point(Q[0], Q[1], Q[2]), proc, md2);
        Spectrum L = proc.photon ? proc.photon->Lo : Spectrum(0.f);
        L.ToRGB(rgb_out + 3 * q);
    }
    delete map;
}

}  // extern "C"
