/* Plain-C client of the C ABI (include/rt_b200.h): builds the README example scene (README.md:41-55 of the reference)
 * exactly as a foreign caller would — no Python, no torch — renders it and prints the mean radiance.
 * usage: readme_scene <width> <spp>      exit code = RtStatus of the first failing call (7 = no sm_100 device). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "rt_b200.h"

#define CHECK(call)                                                                  \
  do {                                                                               \
    int32_t st_ = (call);                                                            \
    if (st_ != RT_OK) {                                                              \
      fprintf(stderr, "%s -> status %d: %s\n", #call, st_, rt_last_error());        \
      return st_;                                                                    \
    }                                                                                \
  } while (0)

static uint32_t add_tex(RtScene* s, int kind, int i0, int i1, double r0, double g0, double b0, double r1, double g1, double b1, int32_t* st) {
  RtTextureDesc t;
  memset(&t, 0, sizeof t);
  t.kind = kind; t.i0 = i0; t.i1 = i1;
  t.c0[0] = r0; t.c0[1] = g0; t.c0[2] = b0; t.c1[0] = r1; t.c1[1] = g1; t.c1[2] = b1;
  uint32_t id = 0;
  *st = rt_scene_add_texture(s, &t, &id);
  return id;
}
static uint32_t add_mat(RtScene* s, int kind, uint32_t tex, int32_t* st) {
  RtMaterialDesc m;
  memset(&m, 0, sizeof m);
  m.kind = kind; m.texture = tex;
  uint32_t id = 0;
  *st = rt_scene_add_material(s, &m, &id);
  return id;
}
static uint32_t add_node(RtScene* s, const RtNodeDesc* d, int32_t* st) {
  uint32_t id = 0;
  *st = rt_scene_add_node(s, d, &id);
  return id;
}
static uint32_t with_material(RtScene* s, uint32_t mat, uint32_t child, int32_t* st) {
  RtNodeDesc d;
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_WITH_MATERIAL; d.material = mat; d.n_children = 1; d.children = &child;
  return add_node(s, &d, st);
}

int main(int argc, char** argv) {
  int width = argc > 1 ? atoi(argv[1]) : 64, spp = argc > 2 ? atoi(argv[2]) : 16;
  int32_t st = RT_OK;
  RtScene* s = NULL;
  CHECK(rt_scene_create(&s));
  uint32_t checker = add_tex(s, RT_TEX_CHECKER, 20, 10, 0.2, 0.2, 0.2, 0.8, 0.8, 0.8, &st); CHECK(st);
  uint32_t blue = add_tex(s, RT_TEX_CONSTANT, 0, 0, 0.0, 0.2, 0.5, 0, 0, 0, &st); CHECK(st);
  uint32_t grey = add_tex(s, RT_TEX_CONSTANT, 0, 0, 0.8, 0.8, 0.8, 0, 0, 0, &st); CHECK(st);
  uint32_t m0 = add_mat(s, RT_MAT_LAMBERTIAN, checker, &st); CHECK(st);
  uint32_t m1 = add_mat(s, RT_MAT_LAMBERTIAN, blue, &st); CHECK(st);
  uint32_t m2 = add_mat(s, RT_MAT_MIRROR, grey, &st); CHECK(st);
  RtNodeDesc d;
  memset(&d, 0, sizeof d); d.kind = RT_NODE_SPHERE; d.s[0] = 1.0;                        /* sphere (0,0,0) 1 */
  uint32_t g0 = with_material(s, m0, add_node(s, &d, &st), &st); CHECK(st);
  memset(&d, 0, sizeof d); d.kind = RT_NODE_SPHERE; d.v[0][1] = -1000.0; d.s[0] = 999.0;  /* sphere (0,-1000,0) 999 */
  uint32_t g1 = with_material(s, m1, add_node(s, &d, &st), &st); CHECK(st);
  memset(&d, 0, sizeof d); d.kind = RT_NODE_PARALLELOGRAM;                                /* parallelogram q u v */
  d.v[0][0] = -3.25; d.v[0][1] = -1; d.v[0][2] = -0.75; d.v[1][0] = 1.25; d.v[1][2] = -1.25; d.v[2][1] = 2;
  uint32_t g2 = with_material(s, m2, add_node(s, &d, &st), &st); CHECK(st);
  uint32_t kids[3] = {g0, g1, g2};
  memset(&d, 0, sizeof d); d.kind = RT_NODE_GROUP; d.n_children = 3; d.children = kids;
  uint32_t root = add_node(s, &d, &st); CHECK(st);
  CHECK(rt_scene_commit(s, root));
  RtSceneInfo info;
  CHECK(rt_scene_info(s, &info));
  printf("committed: %u prims, %u bvh nodes, %llu device bytes\n", info.n_prims, info.n_bvh_nodes, (unsigned long long)info.device_bytes);

  RtCamera cam;
  memset(&cam, 0, sizeof cam);
  cam.center[0] = -0.75; cam.center[2] = 2; cam.look_at[2] = -1; cam.up[1] = 1;
  cam.vfov = 3.14159265358979323846 / 2; cam.aspect_ratio = 16.0 / 9.0; cam.focus_dist = 10;
  cam.bg_kind = RT_BG_CONSTANT; cam.bg_c0[0] = cam.bg_c0[1] = cam.bg_c0[2] = 1; cam.bg_c1[0] = cam.bg_c1[1] = cam.bg_c1[2] = 1;
  cam.image_width = width; cam.samples_per_pixel = spp; cam.max_recursion_depth = 10;
  int32_t height = 0;
  CHECK(rt_image_height(&cam, &height));
  float* img = (float*)malloc((size_t)width * height * 3 * sizeof(float));
  RtStats stats;
  st = rt_render(s, 0, &cam, 100, 0, (uint64_t)spp, img, NULL, &stats);
  if (st != RT_OK) { fprintf(stderr, "rt_render -> status %d: %s\n", st, rt_last_error()); rt_scene_destroy(s); free(img); return st; }
  double sum = 0;
  for (size_t k = 0; k < (size_t)width * height * 3; k++) sum += img[k];
  printf("rendered %dx%d at %d spp: mean %.6f, rays/sample %.4f, %.3f ms on the device, %llu launches\n", width, height, spp,
         sum / ((double)width * height * 3), (double)stats.rays / (double)stats.samples, stats.ms_total, (unsigned long long)stats.launches);
  free(img);
  CHECK(rt_scene_destroy(s));
  return 0;
}
