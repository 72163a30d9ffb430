// TEST INFRASTRUCTURE ONLY — stand-in for the OptiX device header: RayTracerUtilities.cuh names the payload getters
// inside a template the host driver never instantiates, and tex2D inside SampleCubeMap, which the driver does
// instantiate — ref_env_driver.cpp specialises it to record (face handle, u, v).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <math.h>
#include <utility>
static inline unsigned int optixGetPayload_0() { return 0; }
static inline unsigned int optixGetPayload_1() { return 0; }
template <class T>
T tex2D(cudaTextureObject_t tex, float u, float v);
