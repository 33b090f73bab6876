/* The pbrt-facing entry points of the CUDA renderer -- the drop-in boundary.
 *
 * pbrt-v2's core/api.cpp includes this file as "cuda_render/cudaapi.h" when built with -DHAS_CUDA_RENDER
 * (pbrt-v2/src/core/api.cpp:128-130) and calls the four functions below from pbrtInit (:691-695), pbrtShape
 * (:1020-1030), pbrtObjectInstance (:1173-1177) and RenderOptions::MakeRenderer (:1245-1253).  Names, argument
 * types and meaning are those of the reference's cuda_render/cudaapi.h:8-19; the implementation behind them
 * (shim/cudarender.cpp) flattens pbrt's objects into plain arrays and calls only the C ABI in include/b2rt.h.
 * Like the original, this header must not pull in any CUDA header: it is seen by pbrt's own translation units.
 */
#ifndef B2RT_SHIM_CUDAAPI_H
#define B2RT_SHIM_CUDAAPI_H

Renderer *CreateCudaRenderer(Sampler *sampler, Camera *camera, const ParamSet &params, const std::string &rendername);

void CudaRenderInit();

void CreateCudaShape(const std::string &name, Reference<Shape> &shape,
                     std::vector<Reference<Primitive> > *currentInstance, const Material *material, int lightIndex);

void CudaObjectInstance(std::vector<Reference<Primitive> > *key, const Transform &transform);

#endif
