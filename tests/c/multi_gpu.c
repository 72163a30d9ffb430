/* A C host (no Python) driving n GPUs through the C ABI: rtf_create_multi -> replicated scene -> sharded illumination
 * (both modes), LiDAR grid, camera, raw closest hit — each compared with ONE device doing the whole job through the plain
 * single-GPU entry points.  usage: multi_gpu <n_devices>; prints one line per check and exits non-zero on a mismatch.
 * The reference's host is C++ on one device (CUDAModule.cpp:25 cudaSetDevice(0)); this is what its multi-GPU caller
 * would look like. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rtf/rtf.h"

#define CHECK(x)                                                              \
  do {                                                                        \
    int rc_ = (x);                                                            \
    if (rc_ < 0) {                                                            \
      fprintf(stderr, "%s failed (%d): %s\n", #x, rc_, rtf_last_error());     \
      return 2;                                                               \
    }                                                                         \
  } while (0)

static unsigned lcg(unsigned* s) { return *s = *s * 1664525u + 1013904223u; }
static float rnd(unsigned* s) { return (float)(lcg(s) >> 8) / 16777216.0f; }

int main(int argc, char** argv) {
  const int n_dev = argc > 1 ? atoi(argv[1]) : 1;
  const int n_tri = argc > 2 ? atoi(argv[2]) : 20000;
  int devices[64];
  for (int k = 0; k < n_dev; k++) devices[k] = k;
  /* scene: a canopy of random small triangles over a ground quad */
  const int nv = 3 * n_tri + 4;
  rtf_vertex* v = (rtf_vertex*)calloc((size_t)nv, sizeof(rtf_vertex));
  uint32_t* t = (uint32_t*)malloc(sizeof(uint32_t) * 3 * (size_t)(n_tri + 2));
  unsigned seed = 12345u;
  for (int i = 0; i < n_tri; i++) {
    const float cx = 8.0f * rnd(&seed) - 4.0f, cy = 0.2f + 1.5f * rnd(&seed), cz = 8.0f * rnd(&seed) - 4.0f;
    for (int k = 0; k < 3; k++) {
      rtf_vertex* p = &v[3 * i + k];
      p->position[0] = cx + 0.1f * (rnd(&seed) - 0.5f), p->position[1] = cy + 0.1f * (rnd(&seed) - 0.5f);
      p->position[2] = cz + 0.1f * (rnd(&seed) - 0.5f);
      p->normal[1] = 1.0f;
      p->color[0] = p->color[1] = p->color[2] = p->color[3] = 1.0f;
      t[3 * i + k] = (uint32_t)(3 * i + k);
    }
  }
  const float g[4][2] = {{-6, -6}, {6, -6}, {6, 6}, {-6, 6}};
  for (int k = 0; k < 4; k++) {
    rtf_vertex* p = &v[3 * n_tri + k];
    p->position[0] = g[k][0], p->position[2] = g[k][1], p->normal[1] = 1.0f;
    p->color[0] = p->color[1] = p->color[2] = p->color[3] = 1.0f;
  }
  const uint32_t q[6] = {0, 2, 1, 0, 3, 2};
  for (int k = 0; k < 6; k++) t[3 * n_tri + k] = (uint32_t)(3 * n_tri) + q[k];
  rtf_material mat;
  memset(&mat, 0, sizeof(mat));
  mat.albedo_color[0] = 0.3f, mat.albedo_color[1] = 0.6f, mat.albedo_color[2] = 0.2f, mat.roughness = 1.0f;
  const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};

  rtf_multi* m = NULL;
  CHECK(rtf_create_multi(devices, n_dev, &m));
  CHECK(rtf_multi_set_geometry(m, 1, v, (size_t)nv, sizeof(rtf_vertex), t, (size_t)n_tri + 2, 1));
  CHECK(rtf_multi_set_material(m, 1, &mat));
  CHECK(rtf_multi_set_instance(m, 1, 1, 1, ident, 77));
  CHECK(rtf_multi_commit(m, 63));
  rtf_ctx* one = NULL; /* the single-device reference: its own context on device 0 */
  CHECK(rtf_create(0, &one));
  CHECK(rtf_set_geometry(one, 1, v, (size_t)nv, sizeof(rtf_vertex), t, (size_t)n_tri + 2, 1));
  CHECK(rtf_set_material(one, 1, &mat));
  CHECK(rtf_set_instance(one, 1, 1, 1, ident, 77));
  CHECK(rtf_commit(one, 63));
  uint64_t nt = 0;
  CHECK(rtf_triangle_count(one, &nt));

  rtf_env env;
  memset(&env, 0, sizeof(env));
  env.color[0] = env.color[1] = env.color[2] = 1.0f;
  env.skylight_intensity = 1.0f, env.gamma = 1.0f, env.lighting_type = RTF_ENV_SCENE;
  rtf_ray_props rp;
  rp.bounces = 3, rp.samples = 16;
  int bad = 0;

  /* illumination */
  float* e1 = (float*)malloc(24 * nt);
  float* es = (float*)malloc(24 * nt);
  float* et = (float*)malloc(24 * nt);
  CHECK(rtf_estimate_illumination_mesh(one, &env, &rp, 7u, 1e-3f, 1, 1, 0, nt, 0, rp.samples, 1, e1));
  CHECK(rtf_multi_estimate_illumination_mesh(m, &env, &rp, 7u, 1e-3f, 1, 1, RTF_SHARD_SAMPLES, es));
  CHECK(rtf_multi_estimate_illumination_mesh(m, &env, &rp, 7u, 1e-3f, 1, 1, RTF_SHARD_TRIANGLES, et));
  double worst = 0.0, mean = 0.0;
  for (uint64_t i = 0; i < nt; i++)
    for (int k = 0; k < 3; k++) {
      const double a = e1[6 * i + k], b = es[6 * i + k];
      const double rel = fabs(a - b) / (fabs(a) > 1e-30 ? fabs(a) : 1e-30);
      if (rel > worst) worst = rel;
      mean += a;
    }
  const int tri_same = memcmp(e1, et, 24 * nt) == 0;
  printf("illumination: %llu triangles, mean energy %.6f, sample-sharded max rel err %.2e (bar 1e-5), triangle-sharded bit-identical: %d\n",
         (unsigned long long)nt, mean / (3.0 * (double)nt), worst, tri_same);
  if (!(worst <= 1e-5) || !tri_same || !(mean > 0.0)) bad |= 1;

  /* LiDAR grid */
  rtf_scan_grid grid;
  memset(&grid, 0, sizeof(grid));
  grid.center[1] = 10.0f, grid.left[0] = 1.0f, grid.up[2] = 1.0f, grid.direction[1] = -1.0f;
  grid.size[0] = grid.size[1] = 10.0f, grid.distance[0] = grid.distance[1] = 0.01f;
  uint64_t nr = 0;
  CHECK(rtf_scan_grid_ray_count(&grid, &nr));
  float* t1 = (float*)malloc(4 * nr);
  float* tm = (float*)malloc(4 * nr);
  uint32_t* p1 = (uint32_t*)malloc(4 * nr);
  uint32_t* pm = (uint32_t*)malloc(4 * nr);
  CHECK(rtf_scan_grid_trace(one, &grid, 0, nr, t1, p1, NULL));
  CHECK(rtf_multi_scan_grid_trace(m, &grid, tm, pm, NULL));
  uint64_t hits = 0;
  for (uint64_t i = 0; i < nr; i++) hits += p1[i] != 0xFFFFFFFFu;
  const int scan_same = memcmp(t1, tm, 4 * nr) == 0 && memcmp(p1, pm, 4 * nr) == 0;
  printf("lidar: %llu rays, %llu hits, bit-identical: %d\n", (unsigned long long)nr, (unsigned long long)hits, scan_same);
  if (!scan_same || hits == 0) bad |= 2;

  /* camera */
  const float pos[3] = {0.0f, 1.5f, 6.0f}, rot[4] = {0.9950042f, -0.0998334f, 0.0f, 0.0f};
  rtf_camera* c1 = NULL;
  rtf_multi_camera* cm = NULL;
  rtf_camera_props *pp1 = NULL, *ppm = NULL;
  CHECK(rtf_camera_create(one, &c1));
  CHECK(rtf_camera_resize(c1, 320, 200));
  CHECK(rtf_camera_get_props(c1, &pp1));
  pp1->accumulate = 0;
  CHECK(rtf_camera_set(c1, pos, rot));
  CHECK(rtf_multi_camera_create(m, &cm));
  CHECK(rtf_multi_camera_resize(cm, 320, 200));
  CHECK(rtf_multi_camera_get_props(cm, &ppm));
  ppm->accumulate = 0;
  CHECK(rtf_multi_camera_set(cm, pos, rot));
  rp.samples = 4;
  CHECK(rtf_render_camera(one, &env, c1, &rp));
  CHECK(rtf_multi_render_camera(m, &env, cm, &rp, 7));
  float* f1 = (float*)malloc(16u * 320u * 200u);
  float* fm = (float*)malloc(16u * 320u * 200u);
  int cam_same = 1;
  double img = 0.0;
  for (int which = 0; which < 3; which++) {
    CHECK(rtf_camera_download(c1, which, f1));
    CHECK(rtf_multi_camera_download(cm, which, fm));
    cam_same = cam_same && memcmp(f1, fm, 16u * 320u * 200u) == 0;
    if (which == 0)
      for (int i = 0; i < 320 * 200; i++) img += f1[4 * i];
  }
  printf("camera: 320x200x4spp, mean red %.6f, colour/normal/albedo bit-identical: %d\n", img / (320.0 * 200.0), cam_same);
  if (!cam_same || !(img > 0.0)) bad |= 4;

  /* raw closest hit with host buffers */
  const size_t nray = 1u << 18;
  float* rays = (float*)malloc(32 * nray);
  for (size_t i = 0; i < nray; i++) {
    float* r = rays + 8 * i;
    r[0] = 10.0f * rnd(&seed) - 5.0f, r[1] = 3.0f * rnd(&seed), r[2] = 10.0f * rnd(&seed) - 5.0f, r[3] = 0.0f;
    float d[3] = {rnd(&seed) - 0.5f, rnd(&seed) - 0.5f, rnd(&seed) - 0.5f};
    const float l = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) + 1e-9f;
    r[4] = d[0] / l, r[5] = d[1] / l, r[6] = d[2] / l, r[7] = 1e20f;
  }
  void* h1 = malloc(16 * nray);
  void* hm = malloc(16 * nray);
  CHECK(rtf_trace_closest(one, rays, nray, h1));
  CHECK(rtf_multi_trace_closest(m, rays, nray, hm));
  const int hit_same = memcmp(h1, hm, 16 * nray) == 0;
  printf("closest hit: %zu rays, bit-identical: %d\n", nray, hit_same);
  if (!hit_same) bad |= 8;

  rtf_camera_destroy(c1);
  rtf_multi_camera_destroy(cm);
  rtf_destroy(one);
  rtf_multi_destroy(m);
  printf("%s: %d device(s) through rtf_create_multi\n", bad ? "FAILED" : "OK", n_dev);
  return bad ? 1 : 0;
}
