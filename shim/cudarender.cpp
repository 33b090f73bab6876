// Implementation of the pbrt-facing boundary on top of the b2rt C ABI.  See cudarender.h / cuda_render/cudaapi.h.
//
// What each piece replaces in the reference (paths relative to /root/reference):
//   CudaRenderInit / CreateCudaShape / CudaObjectInstance / CreateCudaRenderer   cuda_render/cudaapi.cpp:3-26
//   CudaRender::createCudaShape (+ material cache, light index)                  cuda_render/cudarender.cpp:136-196
//   shape conversion        cuda_render/util/shape/{cudatrianglemesh,cudasphere,cudadisk}.cpp
//   material conversion     cuda_render/util/material/cudamaterial.cpp:8-75
//   light conversion        cuda_render/util/light/cudalight.cpp:16-133
//   camera rays             cuda_render/util/camera/pbrtcamera.cpp:57-122 (CPU loop -> three probe rays)
//   film                    cuda_render/photon_mapping/photonmappingrenderer.cpp:247-283
#include "cudarender.h"
#include "cuda_render/cudaapi.h"
#include "scene.h"
#include "camera.h"
#include "film.h"
#include "sampler.h"
#include "light.h"
#include "material.h"
#include "reflection.h"
#include "intersection.h"
#include "shapes/trianglemesh.h"
#include "shapes/sphere.h"
#include "shapes/disk.h"
#include "lights/point.h"
#include "lights/diffuse.h"
#include "film/image.h"
#include "samplers/stratified.h"
#include "samplers/lowdiscrepancy.h"
#include "filters/box.h"
#include "cameras/perspective.h"
#include "lights/spot.h"
#include "lights/distant.h"
#include "filters/triangle.h"
#include "filters/gaussian.h"
#include "filters/mitchell.h"
#include "filters/sinc.h"
#include <string.h>

// ---- friend accessors: the fork's pbrt headers befriend these class names (shapes/trianglemesh.h:69,
// ---- shapes/sphere.h:62, shapes/disk.h:45, lights/point.h:62, lights/diffuse.h:66) ----------------------------------
class CudaTriangleMesh {
public:
    static int ntris(const TriangleMesh *m) { return m->ntris; }
    static int nverts(const TriangleMesh *m) { return m->nverts; }
    static const int *indices(const TriangleMesh *m) { return m->vertexIndex; }
    static const Point *points(const TriangleMesh *m) { return m->p; }     // already in world space (trianglemesh.cpp:69-71)
    static const Normal *normals(const TriangleMesh *m) { return m->n; }   // OBJECT space, NULL if the mesh has none
    static const float *uvs(const TriangleMesh *m) { return m->uvs; }
    static bool hasTangents(const TriangleMesh *m) { return m->s != NULL; }
    static bool hasAlpha(const TriangleMesh *m) { return m->alphaTexture.GetPtr() != NULL; }
};
class CudaSphere {
public:
    static float radius(const Sphere *s) { return s->radius; }
    static bool clipped(const Sphere *s) { return s->zmin > -s->radius || s->zmax < s->radius; }
    static float zmin(const Sphere *s) { return s->zmin; }
    static float zmax(const Sphere *s) { return s->zmax; }
    static float phiMaxDeg(const Sphere *s) { return s->phiMax >= 2.f * (float)M_PI ? 360.f : s->phiMax * (180.f / (float)M_PI); }
};
class CudaDisk {
public:
    static float height(const Disk *d) { return d->height; }
    static float radius(const Disk *d) { return d->radius; }
    static float inner(const Disk *d) { return d->innerRadius; }
    static float phiMaxDeg(const Disk *d) { return d->phiMax >= 2.f * (float)M_PI ? 360.f : d->phiMax * (180.f / (float)M_PI); }
};
class CudaLight {
public:
    static Point pos(const PointLight *l) { return l->lightPos; }
    static Spectrum intensity(const PointLight *l) { return l->Intensity; }
    static Spectrum emitted(const DiffuseAreaLight *l) { return l->Lemit; }
};

static CudaRender *gCudaRender = NULL;     // one renderer per process, like the reference (cudaapi.cpp:3)

void CudaRenderInit() { gCudaRender = new CudaRender(); }

void CreateCudaShape(const std::string &name, Reference<Shape> &shape, std::vector<Reference<Primitive> > *currentInstance,
                     const Material *material, int lightIndex) {
    gCudaRender->createCudaShape(name, shape, currentInstance, material, lightIndex);
}

void CudaObjectInstance(std::vector<Reference<Primitive> > *key, const Transform &transform) {
    gCudaRender->objectInstance(key, transform);
}

Renderer *CreateCudaRenderer(Sampler *sampler, Camera *camera, const ParamSet &params, const std::string &rendername) {
    gCudaRender->createRenderer(sampler, camera, params, rendername);
    return gCudaRender;      // pbrt deletes it after Render (core/api.cpp:1199)
}

static void m34(const Matrix4x4 &m, float out[12]) {
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) out[4 * r + c] = m.m[r][c];
}
static void rgb3(const Spectrum &s, float out[3]) { s.ToRGB(out); }

CudaRender::CudaRender() : multi_(NULL), ctx_(NULL), renderer_(NULL), sampler_(NULL), camera_(NULL) {
    int device = 0;
    if (const char *e = getenv("B2RT_DEVICE")) device = atoi(e);
    uint64_t seed = 0;
    if (const char *e = getenv("B2RT_SEED")) seed = strtoull(e, NULL, 0);
    // a group of ONE GPU while the scene is parsed; Render() grows it when the Renderer line asks for more
    int rc = b2rt_multi_create(1, &device, seed, &multi_);
    if (rc) Severe("CudaRender: %s", b2rt_last_error(NULL));      // no CPU fallback: abort like cudashape.cpp:32
    ctx_ = b2rt_multi_ctx(multi_, 0);
}

CudaRender::~CudaRender() {
    delete renderer_;
    delete sampler_;
    delete camera_;
    b2rt_multi_destroy(multi_);
    if (gCudaRender == this) gCudaRender = NULL;
}

void CudaRender::check(int rc, const char *what) {
    if (rc) Severe("CudaRender: %s failed: %s", what, b2rt_last_error(ctx_));
}
void CudaRender::checkMulti(int rc, const char *what) {
    if (rc) Severe("CudaRender: %s failed: %s", what, b2rt_multi_last_error(multi_));
}

// "integer gpus" (SURVEY 8(e)): how many GPUs of the box the frame is spread over
static int gpusParam(const ParamSet &params) {
    int def = 1;
    if (const char *e = getenv("B2RT_GPUS")) def = atoi(e);
    int g = params.FindOneInt("gpus", def);
    // B2RT_GPUS_ON_ONE_DEVICE: test hook, the group is made of several contexts on the first device (a one-GPU box still
    // runs the sharded photon pass, the tile lists and the composite)
    const int have = getenv("B2RT_GPUS_ON_ONE_DEVICE") ? 16 : b2rt_device_count();
    if (g <= 0 || g > have) {
        if (g > have) Warning("CudaRender: \"gpus\" %d but %d visible, using %d", g, have, have);
        g = have;
    }
    return g < 1 ? 1 : g;
}

void CudaRender::createRenderer(Sampler *sampler, Camera *camera, const ParamSet &params, const std::string &name) {
    sampler_ = sampler; camera_ = camera;
    delete renderer_;
    if (name == "simple") renderer_ = new SimpleRenderer(params);          // cudarender.cpp:126-134
    else renderer_ = new PhotonMappingRenderer(params);
}

// Material -> (type, Kd|Kr, Kt, ior) through the public GetBSDF interface: evaluate the BSDF the material builds on a
// dummy hit (textures are evaluated once at a default DifferentialGeometry like cudamaterial.cpp:35-42, so constants
// only) and read the lobes back.  Unlike the reference this recovers glass Kr/Kt/index, which GlassMaterial keeps
// private without a friend (materials/glass.h:44-60).
int CudaRender::materialId(const Material *material) {
    std::map<const Material *, int>::iterator it = materialCache_.find(material);
    if (it != materialCache_.end()) return it->second;
    int id = -1;
    if (material) {
        MemoryArena arena;
        Vector dpdu(1, 0, 0), dpdv(0, 1, 0);
        DifferentialGeometry dg(Point(0, 0, 0), dpdu, dpdv, Normal(0, 0, 0), Normal(0, 0, 0), 0.f, 0.f, NULL);
        BSDF *bsdf = material->GetBSDF(dg, dg, arena);
        Vector n(dg.nn), wi;
        float kd[3] = {0.5f, 0.5f, 0.5f}, kt[3] = {0, 0, 0}, pdf = 0.f, ior = bsdf->eta;
        int type = -1;
        if (bsdf->NumComponents(BxDFType(BSDF_DIFFUSE | BSDF_REFLECTION)) > 0) {
            rgb3(bsdf->f(n, n, BxDFType(BSDF_DIFFUSE | BSDF_REFLECTION)) * M_PI, kd);
            type = B2RT_MAT_MATTE;
        } else if (bsdf->NumComponents(BxDFType(BSDF_SPECULAR | BSDF_TRANSMISSION)) > 0) {
            // glass at normal incidence: reflection lobe = F*Kr/cos, transmission lobe = (1-F)*Kt/cos, cos = 1
            float F = ((ior - 1.f) / (ior + 1.f)) * ((ior - 1.f) / (ior + 1.f));
            Spectrum fr = bsdf->Sample_f(n, &wi, BSDFSample(0.5f, 0.5f, 0.5f), &pdf, BxDFType(BSDF_SPECULAR | BSDF_REFLECTION));
            rgb3(pdf > 0.f && F > 0.f ? fr / F : Spectrum(0.f), kd);
            Spectrum ft = bsdf->Sample_f(n, &wi, BSDFSample(0.5f, 0.5f, 0.5f), &pdf, BxDFType(BSDF_SPECULAR | BSDF_TRANSMISSION));
            rgb3(pdf > 0.f ? ft / (1.f - F) : Spectrum(0.f), kt);
            type = B2RT_MAT_GLASS;
        } else if (bsdf->NumComponents(BxDFType(BSDF_SPECULAR | BSDF_REFLECTION)) > 0) {
            Spectrum fr = bsdf->Sample_f(n, &wi, BSDFSample(0.5f, 0.5f, 0.5f), &pdf, BxDFType(BSDF_SPECULAR | BSDF_REFLECTION));
            rgb3(fr, kd);
            type = B2RT_MAT_MIRROR;
        }
        if (type < 0) {
            Warning("CudaRender: material not supported, using grey matte");        // cudamaterial.cpp:20
            type = B2RT_MAT_MATTE;
        }
        // lobes that are not translated (the glossy lobe of plastic / uber / substrate, Oren-Nayar roughness ...): say so,
        // as the reference does for whole materials it cannot translate (cudamaterial.cpp:20)
        const int translated = type == B2RT_MAT_GLASS ? 2 : 1;
        if (bsdf->NumComponents() > translated)
            Warning("CudaRender: material has %d BxDF lobes, the CUDA renderer keeps %d (%s); the others are dropped",
                    bsdf->NumComponents(), translated,
                    type == B2RT_MAT_MATTE ? "Lambertian" : type == B2RT_MAT_GLASS ? "specular reflection + transmission"
                                                                                   : "specular reflection");
        check(b2rt_add_material(ctx_, type, kd, kt, ior, &id), "b2rt_add_material");
    }
    materialCache_[material] = id;
    return id;
}

// hand one flattened shape (optionally moved by an instance transform) to the GPU library
int CudaRender::submit(const ShapeRec &r, const Transform *extra, int lightIndex) {
    int prim = -1;
    const int al = lightIndex;      // b2rt area-light id == pbrt light index (lights are added in scene->lights order)
    if (r.kind == 0) {
        std::vector<float> P = r.P, N = r.N;
        if (extra) {
            for (size_t i = 0; i + 2 < P.size(); i += 3) {
                Point q = (*extra)(Point(P[i], P[i + 1], P[i + 2]));
                P[i] = q.x; P[i + 1] = q.y; P[i + 2] = q.z;
            }
            for (size_t i = 0; i + 2 < N.size(); i += 3) {           // TransformedPrimitive: normals by the instance transform
                Normal q = (*extra)(Normal(N[i], N[i + 1], N[i + 2]));
                N[i] = q.x; N[i + 1] = q.y; N[i + 2] = q.z;
            }
        }
        check(b2rt_add_mesh_ex(ctx_, (int)(P.size() / 3), P.data(), (int)(r.idx.size() / 3), r.idx.data(),
                               N.empty() ? NULL : N.data(), r.uv.empty() ? NULL : r.uv.data(), r.material, al,
                               r.reverse ? B2RT_MESH_REVERSE : 0, &prim), "b2rt_add_mesh_ex");
    } else {
        float o2w[12], w2o[12];
        memcpy(o2w, r.o2w, sizeof(o2w)); memcpy(w2o, r.w2o, sizeof(w2o));
        if (extra) {
            Matrix4x4 a(r.o2w[0], r.o2w[1], r.o2w[2], r.o2w[3], r.o2w[4], r.o2w[5], r.o2w[6], r.o2w[7], r.o2w[8], r.o2w[9],
                        r.o2w[10], r.o2w[11], 0, 0, 0, 1);
            Transform t = (*extra) * Transform(a);
            m34(t.GetMatrix(), o2w); m34(t.GetInverseMatrix(), w2o);
        }
        if (r.kind == 1 && r.clipped)
            check(b2rt_add_sphere_ex(ctx_, o2w, w2o, r.radius, r.height, r.inner, r.phimax, r.material, al, &prim), "b2rt_add_sphere_ex");
        else if (r.kind == 1) check(b2rt_add_sphere(ctx_, o2w, w2o, r.radius, r.material, al, &prim), "b2rt_add_sphere");
        else if (r.clipped) check(b2rt_add_disk_ex(ctx_, o2w, w2o, r.height, r.radius, r.inner, r.phimax, r.material, al, &prim), "b2rt_add_disk_ex");
        else check(b2rt_add_disk(ctx_, o2w, w2o, r.height, r.radius, r.inner, r.material, al, &prim), "b2rt_add_disk");
    }
    return prim;
}

// by shape NAME, like CudaShape::CreateCudaShape (util/shape/cudashape.cpp:13-28); anything else warns and is skipped
// (cudarender.cpp:140-144)
void CudaRender::createCudaShape(const std::string &name, Reference<Shape> &shape,
                                 std::vector<Reference<Primitive> > *currentInstance, const Material *material,
                                 int lightIndex) {
    ShapeRec r;
    r.radius = r.height = r.inner = 0.f;
    r.reverse = false; r.clipped = false; r.phimax = 360.f;
    if (name == "trianglemesh") {
        const TriangleMesh *m = static_cast<const TriangleMesh *>(shape.GetPtr());
        r.kind = 0;
        const Point *p = CudaTriangleMesh::points(m);
        r.P.resize(3 * (size_t)CudaTriangleMesh::nverts(m));
        for (int i = 0; i < CudaTriangleMesh::nverts(m); ++i) { r.P[3 * i] = p[i].x; r.P[3 * i + 1] = p[i].y; r.P[3 * i + 2] = p[i].z; }
        r.idx.assign(CudaTriangleMesh::indices(m), CudaTriangleMesh::indices(m) + 3 * (size_t)CudaTriangleMesh::ntris(m));
        // per-vertex attributes (reference: cudatrianglemesh.cpp:24-44 uploads "N" and "uv" the same way).  pbrt keeps N in
        // OBJECT space and transforms the interpolated normal (trianglemesh.cpp:338-341); the library takes world space.
        const int nv = CudaTriangleMesh::nverts(m);
        if (const Normal *n = CudaTriangleMesh::normals(m)) {
            r.N.resize(3 * (size_t)nv);
            for (int i = 0; i < nv; ++i) {
                Normal w = (*shape->ObjectToWorld)(n[i]);
                r.N[3 * i] = w.x; r.N[3 * i + 1] = w.y; r.N[3 * i + 2] = w.z;
            }
        }
        if (const float *uv = CudaTriangleMesh::uvs(m)) r.uv.assign(uv, uv + 2 * (size_t)nv);
        r.reverse = shape->ReverseOrientation ^ shape->TransformSwapsHandedness;       // core/diffgeom.cpp:52-54
        if (CudaTriangleMesh::hasTangents(m)) Warning("CudaRender: per-vertex tangents \"S\" of a triangle mesh are ignored");
        if (CudaTriangleMesh::hasAlpha(m)) Warning("CudaRender: alpha textures are not supported (cudatrianglemesh.cu:66), mesh is opaque");
    } else if (name == "sphere") {
        const Sphere *s = static_cast<const Sphere *>(shape.GetPtr());
        r.kind = 1; r.radius = CudaSphere::radius(s);
        r.phimax = CudaSphere::phiMaxDeg(s);               // zmin / zmax / phimax are honoured (b2rt_add_sphere_ex)
        r.clipped = CudaSphere::clipped(s) || r.phimax < 360.f;
        r.height = CudaSphere::zmin(s); r.inner = CudaSphere::zmax(s);
    } else if (name == "disk") {
        const Disk *d = static_cast<const Disk *>(shape.GetPtr());
        r.kind = 2; r.radius = CudaDisk::radius(d); r.height = CudaDisk::height(d); r.inner = CudaDisk::inner(d);
        r.phimax = CudaDisk::phiMaxDeg(d);
        if (r.phimax < 360.f && lightIndex >= 0) {
            Warning("CudaRender: \"phimax\" of a disk that is an area light is ignored (full disk)");
            r.phimax = 360.f;
        }
        r.clipped = r.phimax < 360.f;
    } else {
        Warning("CudaRender: shape \"%s\" is not supported by the CUDA renderer, skipped", name.c_str());
        return;
    }
    if (r.kind != 0) {
        m34(shape->ObjectToWorld->GetMatrix(), r.o2w);
        m34(shape->WorldToObject->GetMatrix(), r.w2o);
    }
    r.material = materialId(material);
    if (currentInstance) { instances_[currentInstance].push_back(r); return; }      // inside ObjectBegin/End: deferred
    int prim = submit(r, NULL, lightIndex);
    if (lightIndex >= 0) {
        if (r.kind == 2) lightPrims_[lightIndex] = prim;
        else Warning("CudaRender: only disks can be area lights (cudalight.cpp:35-56); emission of light %d ignored", lightIndex);
    }
}

// Instances are NOT flattened: the shapes of an ObjectBegin block become one b2rt object (its own tree, built once, in the
// block's space) the first time the block is instantiated, and every ObjectInstance adds a transform over it -- the
// two-level BVH that stands in for the reference's OptiX Transform node over the object's Group (cudarender.cpp:38-103).
// B2RT_FLATTEN_INSTANCES=1 re-submits the shapes under the instance transform instead (the round-1 behaviour, for A/B).
void CudaRender::objectInstance(std::vector<Reference<Primitive> > *key, const Transform &transform) {
    std::map<std::vector<Reference<Primitive> > *, std::vector<ShapeRec> >::iterator it = instances_.find(key);
    if (it == instances_.end() || it->second.empty()) return;
    const char *flat = getenv("B2RT_FLATTEN_INSTANCES");
    if (flat && atoi(flat)) {
        for (size_t i = 0; i < it->second.size(); ++i) submit(it->second[i], &transform, -1);
        return;
    }
    std::map<std::vector<Reference<Primitive> > *, int>::iterator ob = objectIds_.find(key);
    int object = -1;
    if (ob == objectIds_.end()) {
        check(b2rt_object_begin(ctx_, &object), "b2rt_object_begin");
        for (size_t i = 0; i < it->second.size(); ++i) submit(it->second[i], NULL, -1);
        check(b2rt_object_end(ctx_), "b2rt_object_end");
        objectIds_[key] = object;
    } else object = ob->second;
    float o2w[12], w2o[12];
    m34(transform.GetMatrix(), o2w);
    m34(transform.GetInverseMatrix(), w2o);
    check(b2rt_add_instance(ctx_, object, o2w, w2o, NULL), "b2rt_add_instance");
}

void CudaRenderer::cacheParams(const ParamSet &params) {
    saveScene = params.FindOneString("savescene", "");
    loadScene = params.FindOneString("loadscene", "");
}

void CudaRender::Render(const Scene *scene) {
    const bool fromCache = !renderer_->loadScene.empty();
    if (fromCache) {
        Info("CudaRender: scene from the cache %s (shapes and lights of the scene file are ignored)", renderer_->loadScene.c_str());
        check(b2rt_scene_load(ctx_, renderer_->loadScene.c_str()), "b2rt_scene_load");
    }
    // lights, in scene->lights order so that ids agree with the lightIndex api.cpp passed for area-light shapes
    for (size_t i = 0; !fromCache && i < scene->lights.size(); ++i) {
        const Light *l = scene->lights[i];
        float rgb[3], p[3] = {0, 0, 0};
        int id = -1;
        if (const PointLight *pl = dynamic_cast<const PointLight *>(l)) {
            Point q = CudaLight::pos(pl);
            p[0] = q.x; p[1] = q.y; p[2] = q.z;
            rgb3(CudaLight::intensity(pl), rgb);
            check(b2rt_add_light(ctx_, B2RT_LIGHT_POINT, p, rgb, -1, 1, &id), "b2rt_add_light");
        } else if (const DiffuseAreaLight *al = dynamic_cast<const DiffuseAreaLight *>(l)) {
            std::map<int, int>::const_iterator it = lightPrims_.find((int)i);
            rgb3(CudaLight::emitted(al), rgb);
            if (it == lightPrims_.end()) {
                Warning("CudaRender: area light %d is not on a disk, treated as black", (int)i);
                rgb[0] = rgb[1] = rgb[2] = 0.f;
                check(b2rt_add_light(ctx_, B2RT_LIGHT_POINT, p, rgb, -1, 1, &id), "b2rt_add_light");
            } else {
                check(b2rt_add_light(ctx_, B2RT_LIGHT_DISK_AREA, NULL, rgb, it->second, l->nSamples, &id), "b2rt_add_light");
            }
        } else if (const SpotLight *sl = dynamic_cast<const SpotLight *>(l)) {
            // SpotLight keeps everything private without a friend line (lights/spot.h:57-61; the reference only warns for
            // it, cudalight.cpp:61-71); this file is compiled with -fno-access-control
            float l2w[12], w2l[12];
            m34(sl->LightToWorld.GetMatrix(), l2w);
            m34(sl->WorldToLight.GetMatrix(), w2l);
            rgb3(sl->Intensity, rgb);
            check(b2rt_add_spot_light(ctx_, l2w, w2l, rgb, sl->cosTotalWidth, sl->cosFalloffStart, &id), "b2rt_add_spot_light");
        } else if (const DistantLight *dl = dynamic_cast<const DistantLight *>(l)) {
            float d[3] = {dl->lightDir.x, dl->lightDir.y, dl->lightDir.z};
            rgb3(dl->L, rgb);
            check(b2rt_add_distant_light(ctx_, d, rgb, &id), "b2rt_add_distant_light");
        } else {
            Warning("CudaRender: light %d has a type the CUDA renderer does not translate, treated as black", (int)i);
            rgb[0] = rgb[1] = rgb[2] = 0.f;
            check(b2rt_add_light(ctx_, B2RT_LIGHT_POINT, p, rgb, -1, 1, &id), "b2rt_add_light");
        }
    }
    if (!renderer_->saveScene.empty()) {
        check(b2rt_scene_save(ctx_, renderer_->saveScene.c_str()), "b2rt_scene_save");
        Info("CudaRender: scene cache written to %s", renderer_->saveScene.c_str());
    }
    if (renderer_->gpus > 1) {
        // the other GPUs get a copy of the scene, lights and materials staged in the first context
        int first = 0;
        if (const char *e = getenv("B2RT_DEVICE")) first = atoi(e);
        std::vector<int> devs;
        const int have = b2rt_device_count();
        for (int i = 0; i < renderer_->gpus; ++i) devs.push_back(getenv("B2RT_GPUS_ON_ONE_DEVICE") ? first : (first + i) % have);
        checkMulti(b2rt_multi_resize(multi_, renderer_->gpus, devs.data()), "b2rt_multi_resize");
        Info("CudaRender: %d GPUs", renderer_->gpus);
    }
    checkMulti(b2rt_multi_build_accel(multi_), "b2rt_multi_build_accel");
    renderer_->render(scene, this, camera_);
}

// ---- camera + film helpers -----------------------------------------------------------------------------------------
// Recover the pinhole model d(x,y) ~ A + x*B + y*C from three rays of the pbrt camera itself (public GenerateRay), so
// any projective pbrt camera with lensradius 0 works without friend access (cameras/perspective.cpp:81-114).
static b2rt_camera probeCamera(const Camera *camera) {
    const int W = camera->film->xResolution, H = camera->film->yResolution;
    // what the library's camera is: a pinhole (b2rt_camera).  Say so when the scene's camera is something else instead of
    // silently rendering a different picture (the reference uploads pbrt's own rays, pbrtcamera.cpp:57-122, so it follows
    // any camera; depth of field there costs a CPU ray loop per frame).
    if (const ProjectiveCamera *pc = dynamic_cast<const ProjectiveCamera *>(camera)) {
        if (pc->lensRadius > 0.f)
            Warning("CudaRender: \"lensradius\" %g ignored, the CUDA renderer's camera is a pinhole (no depth of field)", pc->lensRadius);
        if (!dynamic_cast<const PerspectiveCamera *>(camera))
            Warning("CudaRender: only the perspective camera is translated; this projective camera is rendered as a pinhole "
                    "through its centre ray");
    } else {
        Warning("CudaRender: camera type not translated, rendered as a pinhole fitted to its corner rays");
    }
    struct { float x, y; } pts[4] = {{W * .5f, H * .5f}, {0, 0}, {(float)W, 0}, {0, (float)H}};
    Vector d[4]; Point o;
    for (int i = 0; i < 4; ++i) {
        CameraSample cs;
        cs.imageX = pts[i].x; cs.imageY = pts[i].y; cs.lensU = cs.lensV = 0.5f; cs.time = 0.f;
        Ray r;
        camera->GenerateRay(cs, &r);
        d[i] = Normalize(r.d); o = r.o;
    }
    Vector u[4];
    for (int i = 1; i < 4; ++i) u[i] = d[i] / Dot(d[i], d[0]);      // un-normalise onto the plane dot(., forward) = 1
    b2rt_camera c;
    c.o[0] = o.x; c.o[1] = o.y; c.o[2] = o.z;
    Vector B = (u[2] - u[1]) / (float)W, C = (u[3] - u[1]) / (float)H;
    c.A[0] = u[1].x; c.A[1] = u[1].y; c.A[2] = u[1].z;
    c.B[0] = B.x; c.B[1] = B.y; c.B[2] = B.z;
    c.C[0] = C.x; c.C[1] = C.y; c.C[2] = C.z;
    c.xres = W; c.yres = H;
    return c;
}

// The scene's `PixelFilter`: ImageFilm keeps it in a private member and the filter classes keep their shape parameters
// private as well (film/image.h:66, filters/gaussian.h:52, mitchell.h:61, sinc.h:59); this file is compiled with
// -fno-access-control (shim/Makefile) so the shim can read them without touching the reference's headers.
// Hands the filter to the library; returns the ImageFilm, or NULL for a film type the shim does not know.
static ImageFilm *useFilmFilter(b2rt_multi *ctx, Camera *camera) {
    ImageFilm *film = dynamic_cast<ImageFilm *>(camera->film);
    if (!film || !film->filter) { b2rt_multi_film_set_filter(ctx, NULL); return NULL; }
    Filter *f = film->filter;
    b2rt_filter bf;
    bf.type = B2RT_FILTER_BOX; bf.xwidth = f->xWidth; bf.ywidth = f->yWidth; bf.p0 = bf.p1 = 0.f;
    if (dynamic_cast<TriangleFilter *>(f)) bf.type = B2RT_FILTER_TRIANGLE;
    else if (GaussianFilter *g = dynamic_cast<GaussianFilter *>(f)) { bf.type = B2RT_FILTER_GAUSSIAN; bf.p0 = g->alpha; }
    else if (MitchellFilter *m = dynamic_cast<MitchellFilter *>(f)) { bf.type = B2RT_FILTER_MITCHELL; bf.p0 = m->B; bf.p1 = m->C; }
    else if (LanczosSincFilter *l = dynamic_cast<LanczosSincFilter *>(f)) { bf.type = B2RT_FILTER_SINC; bf.p0 = l->tau; }
    else if (!dynamic_cast<BoxFilter *>(f)) Warning("CudaRender: unknown pixel filter, reconstructing with a box of its widths");
    if (b2rt_multi_film_set_filter(ctx, &bf) != B2RT_OK) { b2rt_multi_film_set_filter(ctx, NULL); return NULL; }
    return film;
}

// Reconstructed pixels -> the film's pixel array (Lxyz with weight 1: WriteImage's normalisation, film/image.cpp:
// 178-213, then returns exactly these values) -> WriteImage.  The reference pushes every sample through
// Film::AddSample on the CPU instead (photonmappingrenderer.cpp:251-283); here the filter already ran on the GPU
// (csrc/film.cu), so adding the pixels as samples would filter twice.  Unknown film types get the old hand-off, one
// AddSample per pixel centre.
static void writeFilm(Camera *camera, ImageFilm *film, const std::vector<float> &rgb) {
    const int W = camera->film->xResolution, H = camera->film->yResolution;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const float *c = &rgb[3 * ((size_t)y * W + x)];
            float v[3] = {c[0], c[1], c[2]};
            for (int k = 0; k < 3; ++k) if (!(v[k] >= 0.f) || isinf(v[k])) v[k] = 0.f;      // NaN / negative / inf guard
            if (film) {
                const int fx = x - film->xPixelStart, fy = y - film->yPixelStart;
                if (fx < 0 || fy < 0 || fx >= film->xPixelCount || fy >= film->yPixelCount) continue;   // outside the crop window
                ImageFilm::Pixel &px = (*film->pixels)(fx, fy);
                RGBToXYZ(v, px.Lxyz);
                px.weightSum = 1.f;
            } else {
                CameraSample cs;
                cs.imageX = x + 0.5f; cs.imageY = y + 0.5f; cs.lensU = cs.lensV = 0.5f; cs.time = 0.f;
                camera->film->AddSample(cs, Spectrum::FromRGB(v));
            }
        }
    camera->film->WriteImage();
}

// The scene's `Sampler`: its pattern inside the pixel is reproduced on the GPU (b2rt_sampler_set) instead of pulling
// every CameraSample through Sampler::GetMoreSamples on the CPU as the reference does (pbrtcamera.cpp:57-122).
// stratified's strata counts are private members (samplers/stratified.h:56; this file has -fno-access-control).
static void useSampler(b2rt_multi *ctx, Sampler *sampler) {
    int rc = B2RT_OK;
    if (StratifiedSampler *st = dynamic_cast<StratifiedSampler *>(sampler))
        rc = b2rt_multi_sampler_set(ctx, B2RT_SAMPLER_STRATIFIED, st->xPixelSamples, st->yPixelSamples);
    else if (dynamic_cast<LDSampler *>(sampler)) rc = b2rt_multi_sampler_set(ctx, B2RT_SAMPLER_LOWDISCREPANCY, 1, 1);
    else rc = b2rt_multi_sampler_set(ctx, B2RT_SAMPLER_RANDOM, 1, 1);       // "random" and the samplers the GPU path has no pattern for
    if (rc != B2RT_OK) b2rt_multi_sampler_set(ctx, B2RT_SAMPLER_RANDOM, 1, 1);
}

static b2rt_frame_params frameParams(const b2rt_camera &cam, Sampler *sampler) {
    b2rt_frame_params fp;
    memset(&fp, 0, sizeof(fp));
    fp.spp = sampler ? sampler->samplesPerPixel : 1;
    fp.max_specular_depth = 5; fp.nused = 50; fp.max_dist2 = 0.01f;
    fp.direct_lighting = 1;
    fp.x0 = 0; fp.x1 = cam.xres; fp.y0 = 0; fp.y1 = cam.yres;
    return fp;
}

SimpleRenderer::SimpleRenderer(const ParamSet &params) { gpus = gpusParam(params); cacheParams(params); }

void SimpleRenderer::render(const Scene *, CudaRender *cr, Camera *camera) {
    b2rt_camera cam = probeCamera(camera);
    b2rt_frame_params fp = frameParams(cam, cr->sampler());
    std::vector<float> img((size_t)cam.xres * cam.yres * 3);
    ImageFilm *film = useFilmFilter(cr->multi(), camera);
    useSampler(cr->multi(), cr->sampler());
    cr->checkMulti(b2rt_multi_render_frame(cr->multi(), &cam, &fp, img.data(), B2RT_HOST), "b2rt_multi_render_frame");
    writeFilm(camera, film, img);
}

PhotonMappingRenderer::PhotonMappingRenderer(const ParamSet &params) {
    // pbrt hands the hook the parameters of the `Renderer` line only (RenderOptions::MakeRenderer, core/api.cpp:1245-1253):
    // a scene that configures `SurfaceIntegrator "photonmap" "integer indirectphotons" [...]` and is run with -c gets the
    // defaults below unless the same parameters are repeated on a Renderer line.  Not silently.
    if (params.ints.empty() && params.floats.empty() && params.bools.empty())
        Warning("CudaRender: no photon-mapping parameters on the Renderer line; SurfaceIntegrator parameters are not visible "
                "to the CUDA renderer (core/api.cpp:1249), using pbrt's photonmap defaults (causticphotons 20000, "
                "indirectphotons 100000, nused 50, maxdist 0.1, finalgather true).  Repeat them as "
                "Renderer \"cuda\" \"integer indirectphotons\" [...] to set them.");
    nCaustic = params.FindOneInt("causticphotons", 20000);          // names and defaults: photonmap.cpp:784-801
    nIndirect = params.FindOneInt("indirectphotons", 100000);
    nUsed = params.FindOneInt("nused", 50);
    maxSpecularDepth = params.FindOneInt("maxspeculardepth", 5);
    maxPhotonDepth = params.FindOneInt("maxphotondepth", 5);
    maxDist = params.FindOneFloat("maxdist", .1f);
    finalGather = params.FindOneBool("finalgather", true);          // pbrt's default is final gathering ON (photonmap.cpp:793)
    gatherSamples = params.FindOneInt("finalgathersamples", 32);
    if (PbrtOptions.quickRender) gatherSamples = max(1, gatherSamples / 4);   // photonmap.cpp:795
    gatherAngle = params.FindOneFloat("gatherangle", 10.f);
    gpus = gpusParam(params);
    cacheParams(params);
}

void PhotonMappingRenderer::render(const Scene *scene, CudaRender *cr, Camera *camera) {
    b2rt_multi *mg = cr->multi();
    int32_t counts[5] = {0, 0, 0, 0, 0};
    cr->check(b2rt_scene_counts(b2rt_multi_ctx(mg, 0), counts), "b2rt_scene_counts");     // the lights may come from a scene cache
    if (counts[1] > 0 && (nCaustic > 0 || nIndirect > 0)) {
        int64_t paths[2], stored[2];
        int64_t batch = 262144;           // the same for every GPU count: the stopping rule counts whole batches, and the
                                          // image must not depend on "gpus"
        cr->checkMulti(b2rt_multi_photons_shoot(mg, nCaustic, nIndirect, maxPhotonDepth, batch, paths, stored), "b2rt_multi_photons_shoot");
        if (stored[0] == 0 && stored[1] == 0)     // pbrt's give-up (photonmap.cpp:526-537): report, render without photon maps
            Error("Unable to store enough photons.  Giving up.\n");
        for (int m = 0; m < 2; ++m) cr->checkMulti(b2rt_multi_photon_grid_build(mg, m, 0.f, 0.f), "b2rt_multi_photon_grid_build");
        Info("CudaRender: %lld caustic photons from %lld paths, %lld indirect photons from %lld paths",
             (long long)stored[0], (long long)paths[0], (long long)stored[1], (long long)paths[1]);
        const bool fg = finalGather && nIndirect > 0 && stored[1] > 0;
        if (fg) {
            // PhotonIntegrator::Preprocess with finalGather (photonmap.cpp:468-490, 383-399): direct photons and radiance
            // photons of the paths that filled the indirect map, then ComputeRadianceTask + the radiance-photon map
            int64_t st[2];
            // pbrt's depositedPhoton rule: direct / indirect hits count while the indirect map collected (paths[1]), caustic
            // hits while the caustic map did (paths[0]); it keeps making radiance photons until BOTH maps are full
            cr->checkMulti(b2rt_multi_fg_deposit_limits(mg, paths[0], paths[1]), "b2rt_multi_fg_deposit_limits");
            const int64_t fgPaths = paths[0] > paths[1] ? paths[0] : paths[1];
            cr->checkMulti(b2rt_multi_photons_trace_fg(mg, 0, fgPaths, maxPhotonDepth, st), "b2rt_multi_photons_trace_fg");
            cr->checkMulti(b2rt_multi_photon_grid_build(mg, B2RT_MAP_DIRECT, 0.f, 0.f), "b2rt_multi_photon_grid_build");
            cr->checkMulti(b2rt_multi_radiance_compute(mg, nUsed, maxDist * maxDist), "b2rt_multi_radiance_compute");
            Info("CudaRender: final gathering, %lld direct photons, %lld radiance photons", (long long)st[0], (long long)st[1]);
        }
        cr->checkMulti(b2rt_multi_final_gather_set(mg, fg ? 1 : 0, gatherSamples, gatherAngle), "b2rt_multi_final_gather_set");
    }
    b2rt_camera cam = probeCamera(camera);
    b2rt_frame_params fp = frameParams(cam, cr->sampler());
    fp.max_specular_depth = maxSpecularDepth; fp.nused = nUsed; fp.max_dist2 = maxDist * maxDist;
    fp.use_caustic = nCaustic > 0; fp.use_indirect = nIndirect > 0;
    std::vector<float> img((size_t)cam.xres * cam.yres * 3);
    ImageFilm *film = useFilmFilter(mg, camera);
    useSampler(mg, cr->sampler());
    cr->checkMulti(b2rt_multi_render_frame(mg, &cam, &fp, img.data(), B2RT_HOST), "b2rt_multi_render_frame");
    writeFilm(camera, film, img);
}
