// ref_btf_driver.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Runs the reference's OWN compressed-BTF evaluator on the host: the MLVQ headers are compiled UNMODIFIED from where
// they lie (/root/reference/include/RayTracerFacility/MLVQ/{BTFBase,PDF6D,PDF4D,PDF3D,PDF2D,PDF1D,IndexAB,VectorColor,
// CIELab,SharedCoordinates}.cuh) against the stand-in headers of oracle/ref_stubs (glm subset, empty optix.h, export
// macro) with `__device__` taken from the CUDA toolkit's host_defines.h (an ignored attribute under g++).
// The output (oracle/_ref/libbtf_ref.so) validates the restatement orc::btf_value_deg (oracle_btf.h) and generates
// tests/golden/btf_ref.npz.  The table set is filled the way CompressedBTF::ImportFromFolder does it
// (src/RayTracerFacility/MLVQ/CompressedBTF.cpp:152-191, 320-385), through the reference's own Init() methods.
#include <cstdlib>
#include <cstring>

#include "ref_btf_fill.h"

using namespace RayTracerFacility;

extern "C" __attribute__((visibility("default"))) int ref_btf_value_deg(const rtf_btf_desc* d, const float* beta_angles,
                                                                         unsigned long long n, const float* q6, float* rgb) {
  BTFBase b;
  ref_fill_btf(b, d, beta_angles);
  for (unsigned long long i = 0; i < n; i++) {
    const float* q = q6 + 6 * i;
    glm::vec3 out(1.0f);  // SurfaceCompressedBTF::GetValue presets 1 (RayDataDefinations.hpp:334)
    b.GetValueDeg(glm::vec2(q[0], q[1]), q[2], q[3], q[4], q[5], out, getenv("REF_BTF_PRINT") != nullptr);
    rgb[3 * i] = out.x, rgb[3 * i + 1] = out.y, rgb[3 * i + 2] = out.z;
  }
  return 0;
}
