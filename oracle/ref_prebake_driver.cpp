// ref_prebake_driver.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Host driver around the REFERENCE's own geometry pre-kernels, CopyVerticesInstancedKernel and CopySkinnedVerticesKernel
// (src/RayTracerFacility/RayTracer.cu:1148-1175 and :1194-1249).  RayTracer.cu as a whole needs the OptiX SDK and CUDA
// device code, but these two kernels are self-contained glm arithmetic: the Makefile cuts exactly those line ranges out of
// the reference file at build time into oracle/_ref/prebake_extract.inc (git-ignored, never committed) and this file
// includes them unmodified, with `__global__` defined away and threadIdx / blockIdx / blockDim supplied below, over the
// glm stand-in of ref_stubs/ (operation orders stated there).  One "thread" per call, the loop plays the grid.
// Output: oracle/_ref/libprebake_ref.so, used by tests/golden/make_golden.py (prebake_ref.npz) and re-run by
// tests/test_ref_programs.py where it exists.
#include <cstdint>
#include <cstring>

#include <glm/glm.hpp>
#include <Vertex.hpp>

namespace {
struct Dim3 {
  unsigned x = 0, y = 0, z = 0;
};
Dim3 threadIdx, blockIdx, blockDim;
}  // namespace
#define __global__ static
#include "_ref/prebake_extract.inc"
#undef __global__

static_assert(sizeof(UniEngine::Vertex) == 80, "Vertex layout");
static_assert(sizeof(UniEngine::SkinnedVertex) == 144, "SkinnedVertex layout");
static_assert(sizeof(glm::mat4) == 64 && sizeof(glm::vec4) == 16 && sizeof(glm::vec3) == 12, "glm stand-in layout");

extern "C" __attribute__((visibility("default"))) void ref_copy_vertices_instanced(int n_matrices, int n_vertices,
                                                                                    const float* matrices, const float* colors,
                                                                                    const void* vertices, float* out_positions,
                                                                                    void* out_vertices) {
  blockDim.x = 1, threadIdx.x = 0;
  for (blockIdx.x = 0; blockIdx.x < (unsigned)(n_matrices * n_vertices) + 3u; blockIdx.x++)  // + 3: the bounds test is the kernel's
    CopyVerticesInstancedKernel(n_matrices, n_vertices, (glm::mat4*)matrices, (glm::vec4*)colors, (UniEngine::Vertex*)vertices,
                                (glm::vec3*)out_positions, (UniEngine::Vertex*)out_vertices);
}
extern "C" __attribute__((visibility("default"))) void ref_copy_skinned_vertices(int n, const void* skinned_vertices,
                                                                                  const float* bone_matrices, float* out_positions,
                                                                                  void* out_vertices) {
  blockDim.x = 1, threadIdx.x = 0;
  for (blockIdx.x = 0; blockIdx.x < (unsigned)n + 3u; blockIdx.x++)
    CopySkinnedVerticesKernel(n, (UniEngine::SkinnedVertex*)skinned_vertices, (glm::mat4*)bone_matrices, (glm::vec3*)out_positions,
                              (UniEngine::Vertex*)out_vertices);
}
