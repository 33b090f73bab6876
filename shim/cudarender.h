// CudaRender: the pbrt Renderer plugin of the B200 backend (class surface of the reference's
// cuda_render/cudarender.h:14-91: `CudaRender : public Renderer` owning the GPU scene and one CudaRenderer
// sub-renderer picked by name -- "simple" or photon mapping, cudarender.cpp:126-134).
// Host C++ compiled against pbrt headers; talks to the GPU only through the C ABI of include/b2rt.h.
#ifndef B2RT_SHIM_CUDARENDER_H
#define B2RT_SHIM_CUDARENDER_H
#include "pbrt.h"
#include "renderer.h"
#include "paramset.h"
#include "transform.h"
#include "b2rt.h"
#include <map>
#include <string>
#include <vector>

class CudaRender;

// sub-renderer interface (reference: cudarender.h:14-19; the CudaCamera argument is gone because camera rays are
// generated on the device, so the pbrt Camera itself is handed over)
class CudaRenderer {
public:
    CudaRenderer() : gpus(1) {}
    virtual ~CudaRenderer() {}
    int gpus;            // "integer gpus" of the Renderer line (default: $B2RT_GPUS, else 1; <= 0: every visible GPU)
    // binary scene cache: "string savescene" writes what the parse staged (shapes, instances, materials,
    // lights) to a file; "string loadscene" renders THAT scene instead of the parsed one, so a stub file with camera, film,
    // sampler and the Renderer line starts without pbrt's text parser touching a single triangle
    std::string saveScene, loadScene;
    void cacheParams(const ParamSet &params);
    virtual void render(const Scene *scene, CudaRender *cudarender, Camera *camera) = 0;
};

class CudaRender : public Renderer {
public:
    CudaRender();
    ~CudaRender();
    // Renderer interface (pbrt-v2/src/core/renderer.h:43-54); Li / Transmittance are stubs as in the reference
    // (cudarender.h:27-32): radiance is computed on the GPU for whole frames, not per ray
    void Render(const Scene *scene);
    Spectrum Li(const Scene *, const RayDifferential &, const Sample *, RNG &, MemoryArena &, Intersection *isect = NULL,
                Spectrum *T = NULL) const { return Spectrum(0.f); }
    Spectrum Transmittance(const Scene *, const RayDifferential &, const Sample *, RNG &, MemoryArena &) const {
        return Spectrum(1.f);
    }
    // called through cudaapi.h
    void createRenderer(Sampler *sampler, Camera *camera, const ParamSet &params, const std::string &name);
    void createCudaShape(const std::string &name, Reference<Shape> &shape,
                         std::vector<Reference<Primitive> > *currentInstance, const Material *material, int lightIndex);
    void objectInstance(std::vector<Reference<Primitive> > *key, const Transform &transform);

    b2rt_ctx *ctx() { return ctx_; }           // context of the first GPU: scene ingestion goes here
    b2rt_multi *multi() { return multi_; }     // the GPU group every pass runs on (one GPU unless "integer gpus" says more)
    void checkMulti(int rc, const char *what);
    Sampler *sampler() { return sampler_; }
    // area-light prims: pbrt light index (position in scene->lights) -> b2rt primitive id of its disk
    const std::map<int, int> &lightPrims() const { return lightPrims_; }
    void check(int rc, const char *what);

private:
    struct ShapeRec {                 // one flattened shape, kept so that object instances can be replayed
        int kind;                     // 0 mesh, 1 sphere, 2 disk
        std::vector<float> P; std::vector<int> idx;
        std::vector<float> N, uv;     // per-vertex world-space normals / uv of a mesh (empty: none)
        bool reverse;                 // ReverseOrientation ^ TransformSwapsHandedness
        bool clipped;                 // sphere with zmin / zmax (kept in height / inner) or phimax; disk with phimax
        float phimax;                 // degrees
        float o2w[12], w2o[12], radius, height, inner;
        int material;
    };
    int materialId(const Material *m);
    int submit(const ShapeRec &r, const Transform *extra, int lightIndex);
    b2rt_multi *multi_;
    b2rt_ctx *ctx_;
    CudaRenderer *renderer_;
    Sampler *sampler_;
    Camera *camera_;
    std::map<const Material *, int> materialCache_;
    std::map<std::vector<Reference<Primitive> > *, std::vector<ShapeRec> > instances_;
    std::map<std::vector<Reference<Primitive> > *, int> objectIds_;     // ObjectBegin block -> b2rt object, made on first use
    std::map<int, int> lightPrims_;
};

// direct lighting only (reference: simple_render/simplerender.{h,cpp})
class SimpleRenderer : public CudaRenderer {
public:
    explicit SimpleRenderer(const ParamSet &params);
    void render(const Scene *scene, CudaRender *cudarender, Camera *camera);
};

// photon mapping (reference: photon_mapping/photonmappingrenderer.{h,cpp}); parameters use pbrt's PhotonIntegrator
// names (integrators/photonmap.cpp:784-801)
class PhotonMappingRenderer : public CudaRenderer {
public:
    explicit PhotonMappingRenderer(const ParamSet &params);
    void render(const Scene *scene, CudaRender *cudarender, Camera *camera);
private:
    int nCaustic, nIndirect, nUsed, maxSpecularDepth, maxPhotonDepth, gatherSamples;
    float maxDist, gatherAngle;
    bool finalGather;
};

#endif
