// TEST INFRASTRUCTURE ONLY — empty stand-in: the OptiX SDK is not on this image and the MLVQ headers only reach it
// through Optix7.hpp's error-check macros, which the host driver never expands.
#pragma once
typedef int OptixResult;
#define OPTIX_SUCCESS 0
