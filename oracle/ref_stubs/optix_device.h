// TEST INFRASTRUCTURE ONLY — stand-in for the OptiX device header (the SDK is not on this image).  Declares the device
// intrinsics the reference's programs call; the host drivers define them: ref_env_driver.cpp only needs tex2D (to record
// SampleCubeMap's addressing), ref_ie_driver.cpp implements all of them as a small host emulation of the OptiX execution
// model (optixTrace = the oracle's BVH traversal + calls into the reference's any-hit / closest-hit / miss programs).
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <utility>
#include <optix.h>

typedef unsigned int OptixVisibilityMask;
enum OptixPrimitiveType {
  OPTIX_PRIMITIVE_TYPE_CUSTOM = 0x2500,
  OPTIX_PRIMITIVE_TYPE_ROUND_QUADRATIC_BSPLINE = 0x2501,
  OPTIX_PRIMITIVE_TYPE_ROUND_CUBIC_BSPLINE = 0x2502,
  OPTIX_PRIMITIVE_TYPE_ROUND_LINEAR = 0x2503,
  OPTIX_PRIMITIVE_TYPE_TRIANGLE = 0x2531
};
enum OptixRayFlags {
  OPTIX_RAY_FLAG_NONE = 0u,
  OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1u << 0,
  OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT = 1u << 2,
  OPTIX_RAY_FLAG_DISABLE_CLOSESTHIT = 1u << 3
};
#define OPTIX_SBT_RECORD_ALIGNMENT 16ull
#define OPTIX_SBT_RECORD_HEADER_SIZE ((size_t)32)

unsigned int optixGetPayload_0();
unsigned int optixGetPayload_1();
uint3 optixGetLaunchIndex();
float3 optixGetWorldRayOrigin();
float3 optixGetWorldRayDirection();
float optixGetRayTmax();
unsigned int optixGetPrimitiveIndex();
OptixPrimitiveType optixGetPrimitiveType();
float2 optixGetTriangleBarycentrics();
float optixGetCurveParameter();
float3 optixTransformPointFromWorldToObjectSpace(float3 p);
CUdeviceptr optixGetSbtDataPointer();
void optixIgnoreIntersection();
void optixTerminateRay();
void optixTrace(OptixTraversableHandle handle, float3 origin, float3 direction, float tmin, float tmax, float time,
                OptixVisibilityMask mask, unsigned int flags, unsigned int sbt_offset, unsigned int sbt_stride,
                unsigned int miss_index, unsigned int& p0, unsigned int& p1);
template <class T>
T tex2D(cudaTextureObject_t tex, float u, float v);
