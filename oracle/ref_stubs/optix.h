// TEST INFRASTRUCTURE ONLY — stand-in: the OptiX SDK is not on this image.  The MLVQ headers only reach it through
// Optix7.hpp's error-check macros, and RayTracer.hpp only names its handle types as members of host classes the host
// drivers never instantiate; opaque handles are enough for both.
#pragma once
typedef int OptixResult;
#define OPTIX_SUCCESS 0
typedef unsigned long long OptixTraversableHandle;
typedef struct OptixDenoiser_t* OptixDenoiser;
typedef struct OptixDeviceContext_t* OptixDeviceContext;
typedef struct OptixModule_t* OptixModule;
typedef struct OptixPipeline_t* OptixPipeline;
typedef struct OptixProgramGroup_t* OptixProgramGroup;
struct OptixModuleCompileOptions { int unused; };
struct OptixPipelineCompileOptions { int unused; };
struct OptixPipelineLinkOptions { int unused; };
struct OptixShaderBindingTable { int unused; };
