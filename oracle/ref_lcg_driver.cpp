// ref_lcg_driver.cpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Exposes the reference's OWN random number generator, compiled verbatim from where it lies:
//   /root/reference/include/RayTracerFacility/RayTracer/LinearCongruenceGenerator.hpp:6-35
// (the only file on the hot path that builds without OptiX / glm / UniEngine; `__device__` is defined away
// on the command line, see the `ref` target of oracle/Makefile).  The output goes to oracle/_ref/ and is used
// to (1) generate tests/golden/lcg_kat.json and (2) validate the restatement orc::LCG in oracle_math.h.
#include <RayTracerFacility/RayTracer/LinearCongruenceGenerator.hpp>

extern "C" __attribute__((visibility("default"))) void ref_lcg(unsigned v0, unsigned v1, int n, float* out,
                                                               unsigned* state_after_init) {
  RayTracerFacility::LinearCongruenceGenerator<16> g;
  g.Init(v0, v1);
  if (state_after_init) *state_after_init = g.state;
  for (int i = 0; i < n; i++) out[i] = g();
}
