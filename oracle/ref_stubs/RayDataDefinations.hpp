// TEST INFRASTRUCTURE ONLY — shadows RayTracer/RayDataDefinations.hpp (launch parameters, SBT records and curve
// intersectors built on OptiX device intrinsics) for the host compile of RayTracerUtilities.cuh / Environment.cuh /
// BSDF.cuh, none of whose functions use anything it declares.
#pragma once
