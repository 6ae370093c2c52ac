// rt_shade.cuh — hit reconstruction, textures, Perlin noise, materials and the scatter step
// of the integrator (reference: Ray.hs:174-224, Material.hs, Texture.hs, Noise.hs).
#pragma once
#include "rt_trace.cuh"

namespace rtb {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;

struct Surf {          // HitRecord (Core.hs:155-162) + what `<$` attached
  f3 p, n;             // world point, unit normal facing the ray
  f3 uvn;              // sphere: LOCAL outward normal (sphereUV input); planar: hit point in the primitive's space
  bool front, sphere_uv;
  uint32_t material, geom;
  uint32_t prim, tf;   // primitive index / type bits: hr_uv is LAZY as in the reference (Geometry.hs:90) — the planar
                       // coordinates are only computed when a texture reads them (surf_uv)
};

// ---- hit record from (ray, t, primitive) ----------------------------------------------------------------
__device__ __forceinline__ Surf reconstruct(const DScene& sc, const RayT& ray, const Hit& hit) {
  Surf s;
  s.p = fma3(hit.t, ray.d, ray.o);
  RT_ASSERT(hit.prim != kNone && (hit.prim >= kMediumBase ? hit.prim - kMediumBase < sc.n_media : hit.prim < sc.n_prims));
  if (hit.prim >= kMediumBase) {       // constantMedium hit (Geometry.hs:321-327)
    s.n = -ray.d;
    s.front = true;
    s.sphere_uv = false;
    s.uvn = mk3(0.f, 0.f, 0.f);
    s.material = sc.media[hit.prim - kMediumBase].material;
    s.geom = hit.prim;   // kMediumBase + k: the next segment starts at a scatter point INSIDE medium k (no surface geometry has such an id)
    s.prim = kNone;
    s.tf = 0;
    return s;
  }
  const float4* P = sc.prims + 4 * (size_t)hit.prim;
  const float4 m4 = __ldg(P + 3), p0 = __ldg(P);
  const uint32_t tf = __float_as_uint(m4.x);
  s.geom = __float_as_uint(m4.y);
  s.material = __float_as_uint(m4.w);
  s.prim = hit.prim;
  s.tf = tf;
  f3 o = ray.o, d = ray.d;
  const DXform* X = nullptr;
  if (tf & (PRIM_FLAG_MOVING | PRIM_FLAG_UVXFORM)) X = sc.xforms + __float_as_uint(m4.z);
  if (tf & PRIM_FLAG_MOVING) {
    o = mat3(X->Ainv, ray.o) + fma3(ray.time, ld3(X->binv1), ld3(X->binv0));
    d = mat3(X->Ainv, ray.d);
  }
  if ((tf & PRIM_TYPE_MASK) == PRIM_SPHERE) {
    f3 pl = fma3(hit.t, d, o);
    f3 on;
    if (tf & PRIM_FLAG_BIG) {
      const float4 p1 = __ldg(P + 1);
      double r = (double)p0.w + (double)p1.w;
      on.x = (float)((((double)pl.x - (double)p0.x) - (double)p1.x) / r);
      on.y = (float)((((double)pl.y - (double)p0.y) - (double)p1.y) / r);
      on.z = (float)((((double)pl.z - (double)p0.z) - (double)p1.z) / r);
    } else {
      on = __frcp_rn(p0.w) * (pl - mk3(p0.x, p0.y, p0.z));   // outward normal, Geometry.hs:83
    }
    f3 onw = on;                                             // world-space outward normal
    if (tf & PRIM_FLAG_MOVING) { s.uvn = on; onw = mat3(X->A, on); }
    else if (tf & PRIM_FLAG_UVXFORM) s.uvn = mat3(X->Ainv, on);
    else s.uvn = on;
    s.front = hit.front;
    s.n = s.front ? onw : -onw;
    s.sphere_uv = true;
  } else {
    const float4 p1 = __ldg(P + 1);
    f3 n = mk3(p0.w, p1.x, p1.y);
    if (tf & PRIM_FLAG_MOVING) n = mat3(X->A, n);
    s.front = hit.front;
    s.n = s.front ? n : -n;
    s.sphere_uv = false;
    s.uvn = fma3(hit.t, d, o);                               // local-space hit point, for surf_uv
  }
  return s;
}

__device__ __forceinline__ void surf_uv(const DScene& sc, const Surf& s, float& u, float& v) {
  if (s.sphere_uv) {   // sphereUV, Geometry.hs:100-104
    u = atan2f(s.uvn.x, s.uvn.z) * (0.5f * kInvPi) + 0.5f;
    v = acosf(fminf(1.f, fmaxf(-1.f, -s.uvn.y))) * kInvPi;
  } else if (s.prim == kNone) {   // constantMedium hit: uv = (0, 0) (Geometry.hs:326)
    u = 0.f;
    v = 0.f;
  } else {             // planar coordinates a, b (Geometry.hs:130-131), then the shape's own uv map
    const float4* P = sc.prims + 4 * (size_t)s.prim;
    const float4 p0 = __ldg(P), p1 = __ldg(P + 1), p2 = __ldg(P + 2);
    const f3 pr = s.uvn - mk3(p0.x, p0.y, p0.z);
    const float a = dot3(mk3(p1.z, p1.w, p2.x), pr), b = dot3(mk3(p2.y, p2.z, p2.w), pr);
    if ((s.tf & PRIM_TYPE_MASK) == PRIM_TRI) {               // Geometry.hs:175
      const float4 u0 = __ldg(sc.tri_uv + 2 * (size_t)s.prim), u1 = __ldg(sc.tri_uv + 2 * (size_t)s.prim + 1);
      const float c0 = 1.f - a - b;
      u = c0 * u0.x + a * u0.z + b * u1.x;
      v = c0 * u0.y + a * u0.w + b * u1.y;
    } else {
      u = a;
      v = b;
    }
  }
}

// ---- Perlin noise (Noise.hs:15-50); tables staged in shared memory by the shading kernel ---------------
struct NoiseTab {
  const float* grad;     // 256*3
  const uint8_t* perm;   // 3*256
};
__device__ __forceinline__ float smoothstep_(float x) { return x * x * (3.f - 2.f * x); }
__device__ float perlin_noise(const NoiseTab& T, f3 p) {
  float flx = floorf(p.x), fly = floorf(p.y), flz = floorf(p.z);
  int ix = (int)flx, iy = (int)fly, iz = (int)flz;
  float fx = p.x - flx, fy = p.y - fly, fz = p.z - flz;
  float wx[2] = {smoothstep_(1.f - fx), smoothstep_(fx)};
  float wy[2] = {smoothstep_(1.f - fy), smoothstep_(fy)};
  float wz[2] = {smoothstep_(1.f - fz), smoothstep_(fz)};
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
      for (int k = 0; k < 2; k++) {
        int h = T.perm[(ix + i) & 255] ^ T.perm[256 + ((iy + j) & 255)] ^ T.perm[512 + ((iz + k) & 255)];
        const float* g = T.grad + 3 * h;
        float dp = g[0] * (fx - (float)i) + g[1] * (fy - (float)j) + g[2] * (fz - (float)k);
        acc += wx[i] * wy[j] * wz[k] * dp;
      }
  return acc;
}
__device__ float fractal_noise(const NoiseTab& T, int depth, f3 p) {
  float acc = 0.f, coef = 1.f;
  for (int o = 0; o < depth; o++) {
    acc += coef * perlin_noise(T, p);
    coef *= 0.5f;
    p = 2.f * p;
  }
  return acc;
}

// ---- textures (Texture.hs:18-79) ------------------------------------------------------------------------------
__device__ __forceinline__ int floor_mod(int a, int m) {
  int r = a % m;
  return r < 0 ? r + m : r;
}
__device__ f3 eval_texture(const DScene& sc, const NoiseTab& T, uint32_t id, const Surf& s) {
  RT_ASSERT(id < sc.n_textures);
  const DTexture& tx = sc.textures[id];
  switch (tx.kind) {
    case RT_TEX_CHECKER: {
      float u, v;
      surf_uv(sc, s, u, v);
      int i = (int)floorf(u * (float)tx.i0), j = (int)floorf(v * (float)tx.i1);
      return ((i + j) & 1) == 0 ? ld3(tx.c0) : ld3(tx.c1);
    }
    case RT_TEX_IMAGE: {
      float u, v;
      surf_uv(sc, s, u, v);
      int i = floor_mod((int)floorf(u * (float)tx.w), tx.w);
      int j = floor_mod((int)floorf((1.f - v) * (float)tx.h), tx.h);
      RT_ASSERT(i >= 0 && i < tx.w && j >= 0 && j < tx.h);
      float4 t = __ldg(sc.texels + tx.image_offset + (size_t)j * tx.w + i);
      return mk3(t.x, t.y, t.z);
    }
    case RT_TEX_NOISE: {
      float n = fractal_noise(T, tx.i0, fma3(tx.freq, s.p, ld3(tx.shift))) * 0.625f + 0.5f;
      f3 c0 = ld3(tx.c0);
      return fma3(n, ld3(tx.c1) - c0, c0);
    }
    case RT_TEX_MARBLE: {
      float sin_arg = tx.freq * dot3(ld3(tx.dir), s.p);
      float n = 10.f * fabsf(fractal_noise(T, 7, fma3(0.25f * tx.freq, s.p, ld3(tx.shift))));
      float g = 0.5f + 0.5f * sinf(sin_arg + n);
      return mk3(g, g, g);
    }
    default: return ld3(tx.c0);   // RT_TEX_CONSTANT
  }
}

__device__ __forceinline__ f3 background(const DCamera& cam, f3 d) {
  if (cam.bg_kind == RT_BG_CONSTANT) return ld3(cam.bg0);
  float a = 0.5f * (d.y + 1.f);             // sky / grayFade, test/Main.hs:19-28
  return (1.f - a) * ld3(cam.bg0) + a * ld3(cam.bg1);
}

__device__ __forceinline__ f3 reflect3(f3 n, f3 v) { return fma3(-2.f * dot3(n, v), n, v); }   // Core.hs:49-51

// solid-angle pdf of redirect target k for direction dir from p (Ray.hs:199-202)
__device__ __forceinline__ float target_pdf(const DTarget& tg, f3 p, f3 dir) {
  f3 n = ld3(tg.n);
  float denom = dot3(n, dir);
  if (!(fabsf(denom) > 1e-8f)) return 0.f;
  f3 qo = ld3(tg.origin) - p;
  float t = __fdividef(dot3(n, qo), denom);      // 2-ulp fast division: these feed a Monte Carlo weight, not a hit distance
  if (!(t > 0.f)) return 0.f;
  f3 pr = fma3(t, dir, -qo);
  float a = dot3(ld3(tg.wa), pr), b = dot3(ld3(tg.wb), pr);
  if (!(0.f <= a && a <= 1.f && 0.f <= b && b <= 1.f)) return 0.f;
  return __fdividef(t * t, fabsf(dot3(ld3(tg.crs), dir)));
}

// One bounce of rayColor (Ray.hs:178-224) after the closest-hit query.  Returns true when the path
// continues (ray / thr / origin_geom updated); `contrib` is the radiance this path adds to its pixel
// (non-zero only when the path ends on an emitter or the background: no library material both emits
// and scatters).
// Material-supplied direction generator (the generalisation the reference's author sketches in Material.hs:36-38: "allowing
// the material to specify a unit vector generator (and pdf) to use in the case of no redirection ... One issue is that it
// would require converting a single vector into an orthonormal basis").  RtOptions.material_sampling = 1 (OFF by default: it
// changes which samples are drawn, not what they converge to): `anisotropic g` then draws the new direction from its own
// Henyey-Greenstein phase function about the incoming direction instead of uniformly on the sphere; the mixture pdf with the
// redirect targets and the weight follow Ray.hs:187-224 with that generator's pdf in place of 1 / 4 pi.  Without redirect
// targets the weight is exactly the albedo (the estimator's variance from the phase function is gone).
__device__ __forceinline__ void onb_from(f3 n, f3& t, f3& b) {   // branch-light orthonormal basis around unit n
  const float sg = copysignf(1.f, n.z), a = -__frcp_rn(sg + n.z), c = n.x * n.y * a;
  t = mk3(1.f + sg * n.x * n.x * a, sg * c, -sg * n.x);
  b = mk3(c, sg + n.y * n.y * a, -n.y);
}
__device__ __forceinline__ float hg_phase_4pi(float g, float mu) {   // Material.hs:124-129: phase function times 4 pi
  const float b = 1.f + g * g - 2.f * g * mu;
  return __fdividef(1.f - g * g, b * sqrtf(b));
}
__device__ __forceinline__ f3 sample_hg(f3 fwd, float g, float u1, float u2) {
  const float q = __fdividef(1.f - g * g, 1.f - g + 2.f * g * u1);
  const float mu = fminf(1.f, fmaxf(-1.f, __fdividef(1.f + g * g - q * q, 2.f * g)));
  const float r = sqrtf(fmaxf(0.f, 1.f - mu * mu));
  float sn, cs;
  sincospif(2.f * u2, &sn, &cs);
  f3 t, b;
  onb_from(fwd, t, b);
  return fma3(r * cs, t, fma3(r * sn, b, mu * fwd));
}

template <bool material_sampling = false>
__device__ __forceinline__ bool shade_bounce(const DScene& sc, const DCamera& cam, const NoiseTab& T, RayT& ray,
                                             const Hit& hit, f3& thr, uint32_t& origin_geom, uint4 rnd, f3& contrib) {
  contrib = mk3(0.f, 0.f, 0.f);
  if (hit.prim == kNone) {   // Ray.hs:179
    contrib = thr * background(cam, ray.d);
    return false;
  }
  const Surf s = reconstruct(sc, ray, hit);
  RT_ASSERT(s.material < sc.n_materials);
  const DMaterial mat = sc.materials[s.material];
  const f3 in_dir = ray.d;
  f3 dir;
  switch (mat.kind) {
    case RT_MAT_LIGHT:                     // Material.hs:41-42
      contrib = thr * eval_texture(sc, T, mat.texture, s);
      return false;
    case RT_MAT_PITCH_BLACK: return false; // Material.hs:46-47
    case RT_MAT_MIRROR:                    // Material.hs:64-67
      thr = thr * eval_texture(sc, T, mat.texture, s);
      dir = reflect3(s.n, in_dir);
      break;
    case RT_MAT_METAL: {                   // Material.hs:72-78
      f3 u = unit_vector(u01(rnd.x), u01(rnd.y));
      f3 d2 = fma3(mat.param, u, reflect3(s.n, in_dir));
      if (!(dot3(d2, s.n) > 0.f)) return false;
      thr = thr * eval_texture(sc, T, mat.texture, s);
      dir = normalize3(d2);
      break;
    }
    case RT_MAT_DIELECTRIC: {              // Material.hs:81-106
      float ratio = s.front ? __frcp_rn(mat.param) : mat.param;
      float cos_t = fminf(1.f, -dot3(s.n, in_dir));
      float sin_t = sqrtf(fmaxf(0.f, 1.f - cos_t * cos_t));
      bool cannot = ratio * sin_t > 1.f;
      float r0 = (1.f - ratio) / (1.f + ratio);
      r0 *= r0;
      float om = 1.f - cos_t, om2 = om * om;
      float refl = r0 + (1.f - r0) * om2 * om2 * om;   // Schlick
      if (cannot || u01(rnd.x) < refl) dir = reflect3(s.n, in_dir);
      else {
        f3 perp = ratio * fma3(cos_t, s.n, in_dir);
        float para = -sqrtf(fabsf(1.f - dot3(perp, perp)));
        dir = fma3(para, s.n, perp);
      }
      break;
    }
    case RT_MAT_TRANSPARENT:               // Material.hs:109-112
      thr = thr * eval_texture(sc, T, mat.texture, s);
      dir = in_dir;
      break;
    default: {                             // HemisphereF / SphereF (Ray.hs:187-224)
      const bool hemi = mat.kind == RT_MAT_LAMBERTIAN || mat.kind == RT_MAT_LOMMEL_SEELIGER;
      const float r = u01(rnd.x);
      int choice = -1;
      for (int k = 0; k < cam.n_targets; k++)
        if (r < cam.targets[k].cumulative) { choice = k; break; }   // Ray.hs:150-151
      // the material's own generator (Material.hs:36-38), used in the case of no redirection
      const bool own = material_sampling && mat.kind == RT_MAT_ANISOTROPIC && fabsf(mat.param) > 1e-3f && fabsf(mat.param) < 0.999f;
      if (choice < 0) {
        if (own) dir = sample_hg(in_dir, mat.param, u01(rnd.y), u01(rnd.z));
        else {
          f3 uu = unit_vector(u01(rnd.y), u01(rnd.z));
          dir = hemi ? s.n + uu : uu;
        }
      } else {
        const DTarget& tg = cam.targets[choice];
        f3 lp = fma3(u01(rnd.y), ld3(tg.U), fma3(u01(rnd.z), ld3(tg.V), ld3(tg.origin)));
        dir = lp - s.p;
      }
      float l2 = dot3(dir, dir);
      if (!(l2 > 1e-12f)) return false;
      dir = rsqrtf(l2) * dir;
      float pdf1 = hemi ? dot3(dir, s.n) * kInvPi : 0.25f * kInvPi;
      if (hemi && !(pdf1 > 0.f)) return false;                      // Ray.hs:197-198
      const float gen = own ? pdf1 * hg_phase_4pi(mat.param, dot3(in_dir, dir)) : pdf1;   // pdf of the no-redirection generator
      float pdf = cam.rem_prob * gen;
      for (int k = 0; k < cam.n_targets; k++) {
        // The target the direction was SAMPLED from is hit by construction (Ray.hs:199-202 finds it again with a
        // parallelogram test): its pdf is t^2 / |crs . dir| with t = |lp - p|.  Re-deriving a, b in fp32 would reject a light
        // point within a few ulp of an edge (1e-6 of the samples) and, with target probabilities summing to 1, leave
        // pdf = 0 -> an infinite weight that poisons the pixel.
        const float pk = k == choice ? __fdividef(l2, fabsf(dot3(ld3(cam.targets[k].crs), dir))) : target_pdf(cam.targets[k], s.p, dir);
        pdf = fmaf(cam.targets[k].p, pk, pdf);
      }
      if (cam.n_targets && !(pdf > 0.f)) return false;              // direction no strategy can produce: no contribution
      f3 f = eval_texture(sc, T, mat.texture, s);
      float w = cam.n_targets ? __fdividef(pdf1, pdf) : 1.f;
      if (mat.kind == RT_MAT_LOMMEL_SEELIGER) {                     // Material.hs:56-61
        float mu0 = -dot3(in_dir, s.n), mu1 = dot3(dir, s.n);
        w *= __fdividef(0.25f, mu0 + mu1);
      } else if (mat.kind == RT_MAT_ANISOTROPIC) {                  // Material.hs:124-129
        // own generator, no redirection: phase / (generator pdf x 4 pi) = 1 exactly - the weight is the albedo
        if (!(own && cam.n_targets == 0)) w *= hg_phase_4pi(mat.param, dot3(in_dir, dir));
      }
      thr = w * (thr * f);
      break;
    }
  }
  ray.o = s.p;
  ray.d = dir;
  origin_geom = s.geom;
  if (sc.n_media == 1u && sc.media[0].closed != 0u && s.geom >= sc.media[0].geom_lo && s.geom <= sc.media[0].geom_hi) {
    // leaving the closed boundary of the only medium: s.n faces the incoming ray, so the outward normal is s.n on a
    // front-side hit and -s.n otherwise; a new direction against it goes into the solid
    const float side = dot3(dir, s.n);
    if (s.front ? side < 0.f : side > 0.f) origin_geom |= kInsideFlag;
  }
  return true;
}

}  // namespace rtb
