/* Plain-C client of the multi-GPU entry points (include/rt_b200.h: rt_context_create / rt_render_multi): builds the
 * Cornell box (test/Main.hs:188-218 of the reference: six walls + light, two rotated cuboids, one redirect target) as a
 * foreign caller would - no Python, no torch, one process - renders it once on device 0 with rt_render and once on ALL
 * visible GPUs with rt_render_multi, and compares the two images (same sample set: equal up to fp32 summation order).
 * usage: cornell_multi <width> <spp> [n_devices]     exit code = RtStatus of the first failing call, 100 = images differ. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "rt_b200.h"

#define CHECK(call)                                                                  \
  do {                                                                               \
    int32_t st_ = (call);                                                            \
    if (st_ != RT_OK) {                                                              \
      fprintf(stderr, "%s -> status %d: %s\n", #call, st_, rt_last_error());        \
      return st_;                                                                    \
    }                                                                                \
  } while (0)

static int32_t g_st = RT_OK;
static uint32_t lambertian(RtScene* s, int kind, double r, double g, double b) {
  RtTextureDesc t;
  memset(&t, 0, sizeof t);
  t.kind = RT_TEX_CONSTANT; t.c0[0] = r; t.c0[1] = g; t.c0[2] = b;
  uint32_t tex = 0, mat = 0;
  if (g_st == RT_OK) g_st = rt_scene_add_texture(s, &t, &tex);
  RtMaterialDesc m;
  memset(&m, 0, sizeof m);
  m.kind = kind; m.texture = tex;
  if (g_st == RT_OK) g_st = rt_scene_add_material(s, &m, &mat);
  return mat;
}
static uint32_t node(RtScene* s, const RtNodeDesc* d) {
  uint32_t id = 0;
  if (g_st == RT_OK) g_st = rt_scene_add_node(s, d, &id);
  return id;
}
static uint32_t with_material(RtScene* s, uint32_t mat, uint32_t child) {
  RtNodeDesc d;
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_WITH_MATERIAL; d.material = mat; d.n_children = 1; d.children = &child;
  return node(s, &d);
}
static uint32_t pgram(RtScene* s, uint32_t mat, double qx, double qy, double qz, double ux, double uy, double uz, double vx, double vy, double vz) {
  RtNodeDesc d;
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_PARALLELOGRAM;
  d.v[0][0] = qx; d.v[0][1] = qy; d.v[0][2] = qz; d.v[1][0] = ux; d.v[1][1] = uy; d.v[1][2] = uz; d.v[2][0] = vx; d.v[2][1] = vy; d.v[2][2] = vz;
  return with_material(s, mat, node(s, &d));
}
/* transform (translate t . rotateY deg) (mat <$ cuboid (fromCorners 0 (sx, sy, sz))) */
static uint32_t box(RtScene* s, uint32_t mat, double sx, double sy, double sz, double deg, double tx, double ty, double tz) {
  RtNodeDesc d;
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_CUBOID; d.v[1][0] = sx; d.v[1][1] = sy; d.v[1][2] = sz;
  uint32_t c = with_material(s, mat, node(s, &d));
  const double a = deg * 3.14159265358979323846 / 180.0, co = cos(a), si = sin(a);
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_TRANSFORM; d.n_children = 1; d.children = &c;
  const double m[16] = {co, 0, si, tx, 0, 1, 0, ty, -si, 0, co, tz, 0, 0, 0, 1};   /* rotateY, Geometry.hs:412-417, then translate */
  memcpy(d.m, m, sizeof m);
  return node(s, &d);
}

int main(int argc, char** argv) {
  const int width = argc > 1 ? atoi(argv[1]) : 96, spp = argc > 2 ? atoi(argv[2]) : 32, want_dev = argc > 3 ? atoi(argv[3]) : 0;
  RtScene* s = NULL;
  CHECK(rt_scene_create(&s));
  const uint32_t red = lambertian(s, RT_MAT_LAMBERTIAN, 0.65, 0.05, 0.05), white = lambertian(s, RT_MAT_LAMBERTIAN, 0.73, 0.73, 0.73),
                 green = lambertian(s, RT_MAT_LAMBERTIAN, 0.12, 0.45, 0.15), light = lambertian(s, RT_MAT_LIGHT, 15, 15, 15);
  uint32_t kids[8];
  kids[0] = pgram(s, green, 555, 0, 0, 0, 555, 0, 0, 0, 555);
  kids[1] = pgram(s, red, 0, 0, 0, 0, 555, 0, 0, 0, 555);
  kids[2] = pgram(s, light, 343, 554, 332, -130, 0, 0, 0, 0, -105);
  kids[3] = pgram(s, white, 0, 0, 0, 555, 0, 0, 0, 0, 555);
  kids[4] = pgram(s, white, 555, 555, 555, -555, 0, 0, 0, 0, -555);
  kids[5] = pgram(s, white, 0, 0, 555, 555, 0, 0, 0, 555, 0);
  kids[6] = box(s, white, 165, 330, 165, 15, 265, 0, 295);
  kids[7] = box(s, white, 165, 165, 165, -18, 130, 0, 65);
  RtNodeDesc d;
  memset(&d, 0, sizeof d);
  d.kind = RT_NODE_GROUP; d.n_children = 8; d.children = kids;
  const uint32_t root = node(s, &d);
  CHECK(g_st);
  CHECK(rt_scene_commit(s, root));
  RtSceneInfo info;
  CHECK(rt_scene_info(s, &info));
  printf("committed: %u prims\n", info.n_prims);

  RtRedirectTarget tg = {0.25, {343, 554, 332}, {-130, 0, 0}, {0, 0, -105}};
  RtCamera cam;
  memset(&cam, 0, sizeof cam);
  cam.center[0] = 278; cam.center[1] = 278; cam.center[2] = -800; cam.look_at[0] = 278; cam.look_at[1] = 278; cam.up[1] = 1;
  cam.vfov = 40.0 * 3.14159265358979323846 / 180.0; cam.aspect_ratio = 1.0; cam.focus_dist = 10;
  cam.bg_kind = RT_BG_CONSTANT;
  cam.image_width = width; cam.samples_per_pixel = spp; cam.max_recursion_depth = 50;
  cam.n_redirect_targets = 1; cam.redirect_targets = &tg;
  int32_t height = 0;
  CHECK(rt_image_height(&cam, &height));
  const size_t n = (size_t)width * height * 3;

  RtContext* ctx = NULL;
  int32_t devs[16];
  for (int k = 0; k < 16; k++) devs[k] = k;
  int32_t st = rt_context_create(want_dev > 0 ? devs : NULL, want_dev > 0 ? want_dev : 0, &ctx);
  if (st != RT_OK) { fprintf(stderr, "rt_context_create -> status %d: %s\n", st, rt_last_error()); rt_scene_destroy(s); return st; }
  float *one = (float*)malloc(n * sizeof(float)), *all = NULL, *sq1 = (float*)malloc(n * sizeof(float)), *sqn = (float*)malloc(n * sizeof(float));
  CHECK(rt_host_alloc(n * sizeof(float), (void**)&all));      /* page-locked: every GPU DMA-copies its slice straight into it */
  RtStats st1, stn;
  CHECK(rt_render(s, 0, &cam, 234, 0, (uint64_t)spp, one, sq1, &st1));
  CHECK(rt_render_multi(ctx, s, &cam, 234, all, NULL, NULL));            /* first call allocates the context's image buffers */
  CHECK(rt_render_multi(ctx, s, &cam, 234, all, sqn, &stn));
  RtContextInfo ci;
  CHECK(rt_context_info(ctx, &ci));
  double sum1 = 0, sumn = 0, worst = 0, worst_sq = 0;
  size_t bad = 0;
  for (size_t k = 0; k < n; k++) {
    sum1 += one[k]; sumn += all[k];
    const double diff = fabs((double)one[k] - (double)all[k]), tol = 2e-4 * fabs((double)one[k]) + 2e-5;
    if (diff > tol) bad++;
    if (diff > worst) worst = diff;
    const double dsq = fabs((double)sq1[k] - (double)sqn[k]) / (fabs((double)sq1[k]) + 1e-3);
    if (dsq > worst_sq) worst_sq = dsq;
  }
  printf("multi: %d devices, reduce mode %d%s%s; %dx%d at %d spp: mean one GPU %.6f, all GPUs %.6f, max |diff| %.3g, sumsq rel diff %.3g, "
         "pixels beyond tolerance %zu; rays %llu vs %llu; device ms %.3f vs %.3f\n", ci.n_devices, ci.reduce_mode,
         ci.reduce_note[0] ? " - " : "", ci.reduce_note, width, height, spp, sum1 / (double)n, sumn / (double)n, worst, worst_sq, bad,
         (unsigned long long)st1.rays, (unsigned long long)stn.rays, st1.ms_total, stn.ms_total);
  free(one); free(sq1); free(sqn);
  CHECK(rt_host_free(all));
  CHECK(rt_context_destroy(ctx));
  CHECK(rt_scene_destroy(s));
  if (bad != 0 || st1.rays != stn.rays || st1.samples != stn.samples || worst_sq > 1e-3) return 100;
  return 0;
}
