// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// A thin extern "C" face over the UNMODIFIED reference CPU renderer (pbrt-v2, compiled from
// /root/reference/pbrt-v2/src by oracle/Makefile into oracle/_ref/libpbrt.a).  It lets the
// Python tests and bench.py's cpu_baseline / --impl reference arm push the same flat arrays the
// GPU library receives through pbrt's own BVHAccel / Triangle / Sphere / Disk code.
//
// What it exercises in the reference (all paths relative to /root/reference):
//   scene build      pbrt-v2/src/shapes/trianglemesh.cpp:42-72, 119-124 (Refine),
//                    pbrt-v2/src/accelerators/bvh.cpp:153-202 (SAH build, maxPrims 4 as api.cpp:494)
//   closest hit      pbrt-v2/src/accelerators/bvh.cpp:388-439 -> shapes/trianglemesh.cpp:146-227
//   any hit          pbrt-v2/src/accelerators/bvh.cpp:442-488 -> shapes/trianglemesh.cpp:230-300
//   brute force      the loop of pbrt-v2/src/renderers/aggregatetest.cpp:93-102
//
// Primitive ids: pbrt numbers shapes by creation order (core/shape.cpp:45); we refine every mesh
// ourselves, in submission order, and remember shapeId -> global primitive index, so ids agree
// with b2rt's "order of submission" numbering.
#include "pbrt.h"
#include "shapes/trianglemesh.h"
#include "shapes/sphere.h"
#include "shapes/disk.h"
#include "accelerators/bvh.h"
#include "materials/matte.h"
#include "textures/constant.h"
#include "primitive.h"
#include "intersection.h"
#include "paramset.h"
#include "reflection.h"
#include "memory.h"
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#include <unordered_map>
#include <chrono>

namespace {
struct RefScene {
    std::vector<Transform *> xforms;                 // kept alive for the shapes
    std::vector<Reference<Shape> > shapes;           // intersectable shapes, global prim order
    std::vector<Reference<Primitive> > prims;
    std::unordered_map<uint32_t, int> id2prim;       // pbrt shapeId -> global prim index
    Reference<Material> mtl;
    Primitive *accel;
    double build_seconds;
    RefScene() : accel(NULL), build_seconds(0), cur_object(-1) {
        Reference<Texture<Spectrum> > kd = new ConstantTexture<Spectrum>(Spectrum(0.5f));
        Reference<Texture<float> > sig = new ConstantTexture<float>(0.f);
        mtl = new MatteMaterial(kd, sig, NULL);
    }
    // object instancing (pbrt ObjectBegin / ObjectEnd / ObjectInstance, core/api.cpp:1098-1178): while cur_object >= 0 the
    // shapes go into that object's own primitive list with ids local to it
    struct Object { std::vector<Reference<Primitive> > prims; int nshapes; bool aggregated; Object() : nshapes(0), aggregated(false) {} };
    std::vector<Object> objects;
    int cur_object;
    struct Instance { int object; uint32_t primitiveId; };
    std::vector<Instance> instances;
    std::unordered_map<uint32_t, int> obj_id2prim;   // pbrt shapeId of an object's shape -> id local to the object
    std::unordered_map<uint32_t, int> inst_of;       // primitiveId of a TransformedPrimitive -> instance index
    void push(const Reference<Shape> &s) {
        if (cur_object >= 0) {
            Object &o = objects[cur_object];
            obj_id2prim[s->shapeId] = o.nshapes++;
            o.prims.push_back(new GeometricPrimitive(s, mtl, NULL));
            return;
        }
        id2prim[s->shapeId] = (int)shapes.size();
        shapes.push_back(s);
        prims.push_back(new GeometricPrimitive(s, mtl, NULL));
    }
};

Matrix4x4 mat12(const float m[12]) {
    return Matrix4x4(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], 0, 0, 0, 1);
}
// o2w and w2o are both caller-supplied so the GPU library and pbrt see the very same floats
void make_xforms(RefScene *s, const float o2w[12], const float w2o[12], Transform **t, Transform **ti) {
    *t = new Transform(mat12(o2w), mat12(w2o));
    *ti = new Transform(mat12(w2o), mat12(o2w));
    s->xforms.push_back(*t); s->xforms.push_back(*ti);
}

// Work distribution of the timed loops.  With as many threads as pbrt's own pool has, the ranges go through pbrt's task
// system itself (EnqueueTasks / WaitForAllTasks, core/parallel.cpp:800-826, the way SamplerRenderer::Render hands out
// image tiles): a worker takes the next task when it finishes one.  Otherwise the same dynamic scheme with std::thread
// and an atomic cursor.  Ranges are 4096 rays: a 1080p frame where half the rays miss is badly balanced by one static
// range per thread (the first version: 25.5 Mrays/s on 16 threads where 32 gave 46.9).
template <typename F> struct RangeTask : public Task {
    RangeTask(F *f, int64_t lo, int64_t hi) : f(f), lo(lo), hi(hi) {}
    void Run() { (*f)(lo, hi); }
    F *f; int64_t lo, hi;
};
template <typename F> void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n <= 0) { f(0, n); return; }
    const int64_t grain = 4096;
    if (nthreads == NumSystemCores() && PbrtOptions.nCores != 1) {
        vector<Task *> tasks;
        for (int64_t lo = 0; lo < n; lo += grain) tasks.push_back(new RangeTask<F>(&f, lo, std::min(n, lo + grain)));
        EnqueueTasks(tasks);
        WaitForAllTasks();
        for (size_t i = 0; i < tasks.size(); ++i) delete tasks[i];
        return;
    }
    std::atomic<int64_t> cursor(0);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&]() {
            for (;;) {
                const int64_t lo = cursor.fetch_add(grain);
                if (lo >= n) break;
                f(lo, std::min(n, lo + grain));
            }
        });
    for (auto &t : th) t.join();
}
}  // namespace

extern "C" {

void *ref_scene_create() { return new RefScene(); }

// world-space vertices (identity transform), same as what the GPU library is given
int ref_scene_add_mesh(void *h, int nverts, const float *P, int ntris, const int *idx) {
    RefScene *s = (RefScene *)h;
    Transform *I = new Transform();
    s->xforms.push_back(I);
    Reference<Shape> mesh = new TriangleMesh(I, I, false, ntris, nverts, idx, (const Point *)P,
                                             NULL, NULL, NULL, Reference<Texture<float> >());
    std::vector<Reference<Shape> > tris;
    mesh->Refine(tris);
    int first = (int)s->shapes.size();
    for (size_t i = 0; i < tris.size(); ++i) s->push(tris[i]);
    return first;
}

// the same with pbrt's "N" and "uv" mesh parameters and ReverseOrientation (identity object-to-world: N is world space)
int ref_scene_add_mesh_attr(void *h, int nverts, const float *P, int ntris, const int *idx, const float *N,
                            const float *uv, int reverse) {
    RefScene *s = (RefScene *)h;
    Transform *I = new Transform();
    s->xforms.push_back(I);
    Reference<Shape> mesh = new TriangleMesh(I, I, reverse != 0, ntris, nverts, idx, (const Point *)P,
                                             (const Normal *)N, NULL, uv, Reference<Texture<float> >());
    std::vector<Reference<Shape> > tris;
    mesh->Refine(tris);
    int first = (int)s->shapes.size();
    for (size_t i = 0; i < tris.size(); ++i) s->push(tris[i]);
    return first;
}

int ref_scene_add_sphere(void *h, const float o2w[12], const float w2o[12], float radius) {
    RefScene *s = (RefScene *)h;
    Transform *t, *ti;
    make_xforms(s, o2w, w2o, &t, &ti);
    int first = (int)s->shapes.size();
    s->push(new Sphere(t, ti, false, radius, -radius, radius, 360.f));
    return first;
}

// pbrt's partial sphere ("zmin" / "zmax" / "phimax") and partial disk ("phimax")
int ref_scene_add_disk_ex(void *h, const float o2w[12], const float w2o[12], float height, float radius, float inner, float phimax) {
    RefScene *s = (RefScene *)h;
    Transform *t, *ti;
    make_xforms(s, o2w, w2o, &t, &ti);
    int first = (int)s->shapes.size();
    s->push(new Disk(t, ti, false, height, radius, inner, phimax));
    return first;
}
int ref_scene_add_sphere_ex(void *h, const float o2w[12], const float w2o[12], float radius, float zmin, float zmax, float phimax) {
    RefScene *s = (RefScene *)h;
    Transform *t, *ti;
    make_xforms(s, o2w, w2o, &t, &ti);
    int first = (int)s->shapes.size();
    s->push(new Sphere(t, ti, false, radius, zmin, zmax, phimax));
    return first;
}

int ref_scene_add_disk(void *h, const float o2w[12], const float w2o[12], float height, float radius,
                       float inner) {
    RefScene *s = (RefScene *)h;
    Transform *t, *ti;
    make_xforms(s, o2w, w2o, &t, &ti);
    int first = (int)s->shapes.size();
    s->push(new Disk(t, ti, false, height, radius, inner, 360.f));
    return first;
}

// BVHAccel exactly as RenderOptions::MakeScene makes it by default ("sah", maxnodeprims 4)
double ref_scene_build(void *h) {
    RefScene *s = (RefScene *)h;
    auto t0 = std::chrono::steady_clock::now();
    ParamSet ps;
    s->accel = CreateBVHAccelerator(s->prims, ps);
    s->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return s->build_seconds;
}

int ref_scene_nprims(void *h) { return (int)((RefScene *)h)->shapes.size(); }

// ---- object instances, exactly as pbrtObjectInstance makes them (core/api.cpp:1140-1172) -----------------------------
int ref_scene_object_begin(void *h) {
    RefScene *s = (RefScene *)h;
    s->objects.push_back(RefScene::Object());
    s->cur_object = (int)s->objects.size() - 1;
    return s->cur_object;
}
void ref_scene_object_end(void *h) { ((RefScene *)h)->cur_object = -1; }
// o2w = instance-to-world (pbrt's CTM at ObjectInstance), w2o its inverse.  The instance takes ONE slot in the scene's
// primitive list; returns the instance index.  Hits inside it are reported as (instance, id local to the object).
int ref_scene_add_instance(void *h, int object, const float o2w[12], const float w2o[12]) {
    RefScene *s = (RefScene *)h;
    RefScene::Object &o = s->objects[object];
    if (o.prims.empty()) return -1;
    if (!o.aggregated && (o.prims.size() > 1 || !o.prims[0]->CanIntersect())) {
        Reference<Primitive> accel = CreateBVHAccelerator(o.prims, ParamSet());
        o.prims.erase(o.prims.begin(), o.prims.end());
        o.prims.push_back(accel);
    }
    o.aggregated = true;
    Transform *t, *ti;
    make_xforms(s, o2w, w2o, &t, &ti);
    AnimatedTransform w2i(ti, 0.f, ti, 1.f);
    Reference<Primitive> prim = new TransformedPrimitive(o.prims[0], w2i);
    RefScene::Instance I; I.object = object; I.primitiveId = prim->primitiveId;
    s->inst_of[prim->primitiveId] = (int)s->instances.size();
    s->instances.push_back(I);
    s->prims.push_back(prim);
    s->shapes.push_back(Reference<Shape>());     // keeps "item" numbering: one slot per instance, never reported
    return (int)s->instances.size() - 1;
}
// closest hit with the instance resolved: inst_out = instance index or -1, prim_out = id local to the object for an
// instanced hit, global item id otherwise
double ref_intersect_inst(void *h, int64_t nrays, const float *rays, float *t_out, int *prim_out, int *inst_out, int nthreads) {
    RefScene *s = (RefScene *)h;
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(nrays, nthreads, [=](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const float *r = rays + 8 * i;
            Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7]);
            Intersection isect;
            t_out[i] = INFINITY; prim_out[i] = -1; inst_out[i] = -1;
            if (s->accel->Intersect(ray, &isect)) {
                t_out[i] = ray.maxt;
                std::unordered_map<uint32_t, int>::const_iterator it = s->inst_of.find(isect.primitiveId);
                if (it != s->inst_of.end()) { inst_out[i] = it->second; prim_out[i] = s->obj_id2prim.find(isect.shapeId)->second; }
                else prim_out[i] = s->id2prim.find(isect.shapeId)->second;
            }
        }
    });
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void ref_scene_destroy(void *h) {
    RefScene *s = (RefScene *)h;
    if (!s) return;
    delete s->accel;   // shapes are ref-counted; transforms leak on purpose (test infra)
    delete s;
}

// rays: 8 floats each {ox,oy,oz,tmin, dx,dy,dz,tmax}.  Outputs: t (INF on miss), prim (-1), u, v.
// Returns seconds spent inside the traversal loop.
double ref_intersect(void *h, int64_t nrays, const float *rays, float *t_out, int *prim_out,
                     float *uv_out, int nthreads) {
    RefScene *s = (RefScene *)h;
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(nrays, nthreads, [=](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const float *r = rays + 8 * i;
            Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7]);
            Intersection isect;
            if (s->accel->Intersect(ray, &isect)) {
                t_out[i] = ray.maxt;
                prim_out[i] = s->id2prim.find(isect.shapeId)->second;
                if (uv_out) { uv_out[2 * i] = isect.dg.u; uv_out[2 * i + 1] = isect.dg.v; }
            } else {
                t_out[i] = INFINITY; prim_out[i] = -1;
                if (uv_out) { uv_out[2 * i] = 0; uv_out[2 * i + 1] = 0; }
            }
        }
    });
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Hit attributes as pbrt forms them: Intersection::GetBSDF runs Shape::GetShadingGeometry and MatteMaterial::GetBSDF,
// i.e. BSDF(dgShading, dgGeom.nn) (core/reflection.cpp:601-609).  The BSDF's frame is private; LocalToWorld of the unit
// axes returns sn, tn, nn (x * 1 + y * 0 + z * 0: exact up to the sign of a zero).  out: 20 floats per ray = b2rt_surface.
void ref_surface_inst(void *h, int64_t nrays, const float *rays, float *out, int *inst_out, int nthreads);
void ref_surface(void *h, int64_t nrays, const float *rays, float *out, int nthreads) {
    ref_surface_inst(h, nrays, rays, out, NULL, nthreads);
}
// the same for scenes with object instances: inst_out (may be NULL) = instance index or -1; the prim field of an instanced
// hit is the id local to its object.  TransformedPrimitive::Intersect has taken dg to world space and composed the
// transforms GetShadingGeometry uses (core/primitive.cpp:104-117).
void ref_surface_inst(void *h, int64_t nrays, const float *rays, float *out, int *inst_out, int nthreads) {
    RefScene *s = (RefScene *)h;
    parallel_for(nrays, nthreads, [=](int64_t lo, int64_t hi) {
        MemoryArena arena;
        for (int64_t i = lo; i < hi; ++i) {
            const float *r = rays + 8 * i;
            float *o = out + 20 * i;
            for (int k = 0; k < 20; ++k) o[k] = 0.f;
            Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7]);
            Intersection isect;
            int prim = -1;
            if (inst_out) inst_out[i] = -1;
            if (s->accel->Intersect(ray, &isect)) {
                std::unordered_map<uint32_t, int>::const_iterator it = s->inst_of.find(isect.primitiveId);
                if (it != s->inst_of.end()) { if (inst_out) inst_out[i] = it->second; prim = s->obj_id2prim.find(isect.shapeId)->second; }
                else prim = s->id2prim.find(isect.shapeId)->second;
                BSDF *bsdf = isect.GetBSDF(RayDifferential(ray), arena);
                const Vector sn = bsdf->LocalToWorld(Vector(1, 0, 0)), tn = bsdf->LocalToWorld(Vector(0, 1, 0)),
                             nn = bsdf->LocalToWorld(Vector(0, 0, 1));
                o[0] = isect.dg.p.x; o[1] = isect.dg.p.y; o[2] = isect.dg.p.z; o[3] = ray.maxt;
                o[4] = isect.dg.nn.x; o[5] = isect.dg.nn.y; o[6] = isect.dg.nn.z; o[7] = isect.dg.u;
                o[8] = nn.x; o[9] = nn.y; o[10] = nn.z; o[11] = isect.dg.v;
                o[12] = sn.x; o[13] = sn.y; o[14] = sn.z; o[15] = isect.rayEpsilon;
                o[16] = tn.x; o[17] = tn.y; o[18] = tn.z;
                arena.FreeAll();
            }
            memcpy(o + 19, &prim, 4);
        }
    });
}

double ref_intersectp(void *h, int64_t nrays, const float *rays, unsigned char *occ, int nthreads) {
    RefScene *s = (RefScene *)h;
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(nrays, nthreads, [=](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const float *r = rays + 8 * i;
            Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7]);
            occ[i] = s->accel->IntersectP(ray) ? 1 : 0;
        }
    });
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Exhaustive loop over every shape with pbrt's own Shape::Intersect, each against the ORIGINAL
// [tmin,tmax] (no shrinking), winner = (min t, then min prim index).  nties counts rays whose
// winning t is shared by >1 primitive (where pbrt's BVH answer depends on traversal order).
double ref_intersect_brute(void *h, int64_t nrays, const float *rays, float *t_out, int *prim_out,
                           int64_t *nties, int nthreads) {
    RefScene *s = (RefScene *)h;
    std::atomic<int64_t> ties(0);
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(nrays, nthreads, [&](int64_t lo, int64_t hi) {
        int64_t myties = 0;
        for (int64_t i = lo; i < hi; ++i) {
            const float *r = rays + 8 * i;
            Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7]);
            float best = INFINITY; int bp = -1; int cnt = 0;
            for (size_t k = 0; k < s->shapes.size(); ++k) {
                float t, eps; DifferentialGeometry dg;
                if (s->shapes[k] && s->shapes[k]->Intersect(ray, &t, &eps, &dg)) {
                    if (t < best) { best = t; bp = (int)k; cnt = 1; }
                    else if (t == best) ++cnt;
                }
            }
            t_out[i] = best; prim_out[i] = bp;
            if (cnt > 1) ++myties;
        }
        ties += myties;
    });
    if (nties) *nties = ties.load();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int ref_num_cores() { return NumSystemCores(); }

}  // extern "C"
