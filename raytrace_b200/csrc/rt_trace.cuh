// rt_trace.cuh — device-side math, Philox RNG, primitive tests and BVH traversal (sm_100a).
//
// Reference semantics followed (src/Graphics/Ray/*.hs):
//   * open hit intervals tmin < t < tmax (Core.hs:83-84); closed, conservative box test
//     (Core.hs:95-106,147-152) — here on padded fp32 boxes, NaN never culls;
//   * sphere: nearer root if inside the interval, else the farther one (Geometry.hs:63-80);
//   * plane shapes: |n.d| > 1e-8, planar coordinates a, b, inclusive shape tests
//     (Geometry.hs:123-132,150,174);
//   * exact ties: the primitive earlier in the reference's evaluation order wins
//     (Geometry.hs:340-345,356-361) = lower primitive index here.
// Deliberate fp32 measures: a ray never re-hits the planar primitive it
// starts on and only sees the far root of the sphere it starts on (exact in real arithmetic,
// and what the reference's fp64 + tmin = 1e-4 yields); spheres with r >= 64 use an fp64
// discriminant.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_device.h"

namespace rtb {

// ---- small vector helpers -----------------------------------------------------------------------
struct f3 {
  float x, y, z;
};
__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 ld3(const float* p) { return f3{p[0], p[1], p[2]}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator-(f3 a) { return f3{-a.x, -a.y, -a.z}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return f3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return f3{a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ f3 fma3(float s, f3 a, f3 b) { return f3{fmaf(s, a.x, b.x), fmaf(s, a.y, b.y), fmaf(s, a.z, b.z)}; }
__device__ __forceinline__ f3 mat3(const float* M, f3 v) {
  return f3{fmaf(M[0], v.x, fmaf(M[1], v.y, M[2] * v.z)), fmaf(M[3], v.x, fmaf(M[4], v.y, M[5] * v.z)),
            fmaf(M[6], v.x, fmaf(M[7], v.y, M[8] * v.z))};
}
// linear's `normalize` (leaves near-zero / already-unit vectors alone)
__device__ __forceinline__ f3 normalize3(f3 v) {
  float l = dot3(v, v);
  if (fabsf(l) <= 1e-12f) return v;
  return rsqrtf(l) * v;
}

// ---- Philox4x32-10 counter RNG -------------------------------------------------------------------
// key = render seed; counter = (pixel, sample lo, bounce<<8 | block, sample hi).  The sample set
// therefore depends only on (seed, pixel, global sample index), never on pool size or GPU count.
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }  // [0,1)

enum : uint32_t { RNG_RAYGEN = 0xF0u, RNG_MEDIA = 0x10u, RNG_SHADE = 0x40u };
__device__ __forceinline__ uint4 rng4(uint2 key, uint32_t pixel, uint64_t sample, uint32_t bounce, uint32_t block) {
  return philox4x32(make_uint4(pixel, (uint32_t)sample, (bounce << 8) | block, (uint32_t)(sample >> 32)), key);
}

// uniform direction on the unit sphere (same distribution as the rejection sampler, Core.hs:54-60)
__device__ __forceinline__ f3 unit_vector(float u1, float u2) {
  float z = 1.f - 2.f * u1;
  float r = sqrtf(fmaxf(0.f, 1.f - z * z));
  float s, c;
  sincospif(2.f * u2, &s, &c);
  return f3{r * c, r * s, z};
}

// ---- ray ---------------------------------------------------------------------------------------------
struct RayT {
  f3 o, d;
  float time;
};

struct Hit {
  float t;
  uint32_t prim;   // primitive index, kNone = miss
  bool front;
};

// fast reciprocal (MUFU.RCP, <= 1 ulp): operands here are never denormal (|n.d| > 1e-8, q != 0)
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// 32-byte read-only global load (sm_100: LDG.E.256): BVH node pairs and primitive records are fetched by divergent
// lanes, where every load instruction costs the L1 one tag lookup per lane — half as many instructions, half as many
// lookups (ncu r01g pawn: l1tex at 67 % of peak, 40 % of stall samples waiting on these loads).  p must be 32 B aligned.
struct F8 { float4 a, b; };
__device__ __forceinline__ F8 ldg256(const float4* p) {
#ifndef RT_NO_LDG256
  F8 r;
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
      : "l"(p));
  return r;
#else
  return F8{__ldg(p), __ldg(p + 1)};
#endif
}

// Where a primitive record comes from: global memory (two 32-byte loads: {r0, r1} and {r2, meta}) ...
struct GlobalRec {
  F8 lo, hi;
  static constexpr bool kBoxes = true;    // aggregated cuboids are answered at their head face
  __device__ __forceinline__ float4 meta() const { return hi.b; }
  __device__ __forceinline__ float4 r0() const { return lo.a; }
  __device__ __forceinline__ float4 r1() const { return lo.b; }
  __device__ __forceinline__ float4 r2() const { return hi.a; }
};
// ... or the kernel parameter block (constant bank: the values become instruction operands, no loads at all) —
// used by the flat-list kernel, where every lane of a warp tests the same primitive at the same time.
struct ParamRec {
  const float4* R;      // 4 consecutive float4 in param space
  static constexpr bool kBoxes = false;   // the flat-list kernels keep boxes in a list of their own
  __device__ __forceinline__ float4 meta() const { return R[3]; }
  __device__ __forceinline__ float4 r0() const { return R[0]; }
  __device__ __forceinline__ float4 r1() const { return R[1]; }
  __device__ __forceinline__ float4 r2() const { return R[2]; }
};

// Planar coordinates a, b are computed in fp32 from p = o + t d (rounding ~ 3e-7 at Cornell's scale): with the exact bounds
// 0 <= a, b <= 1 (Geometry.hs:150,174) a ray through the shared edge of two parallelograms / triangles can fall OUTSIDE both
// and see the background through a closed wall (28 pixel centres on Cornell's wall / floor edge, one on the pawn).  The
// bounds are therefore widened by kEdgeEps: on the edge both primitives are hit and the closer one - or, on an exact tie,
// the one the reference evaluates first - wins, which is what the fp64 reference does on one side or the other.
constexpr float kEdgeEps = 1e-6f;

// Exact tie between two primitives (same t to the last bit: coplanar duplicates, shared edges): the reference keeps the
// one it evaluates first (Geometry.hs:340-345,356-361).  Device indices follow the device tree, so the evaluation order
// is looked up — on ties only.
__device__ __forceinline__ bool evaluated_before(const uint32_t* __restrict__ ref_order, uint32_t a, uint32_t b) {
  return b == kNone || __ldg(ref_order + a) < __ldg(ref_order + b);
}

// Closest hit of an aggregated cuboid = closest hit over its six parallelogram faces (Geometry.hs:154-166 + :117-144),
// by the slab method: the boundary of a convex solid is crossed at most twice, at t_in (from outside: front = True) and
// t_out (from inside).  Per face the reference requires |n.d| > 1e-8 (a parallel ray misses both faces of that axis; the
// closed interior test 0 <= a, b <= 1 of the other faces then decides by |coordinate| <= half extent), t in (tmin, best]
// and that the face is not the one the ray starts on (skip_a) / must not report again (skip_b).  The closest face passing
// all of that is t_in's face if it passes, else t_out's.  b0..b3 = box record (rt_device.h), `first` / `geom0` = primitive
// index / geometry id of face 0; faces +z -z -x +x +y -y = 0..5, axes tested z, x, y so that an exact edge tie goes to the
// lower face index like the reference's scan.  strict: caller scans in index order (ties lose to earlier primitives).
__device__ __forceinline__ void intersect_box(const uint32_t* __restrict__ ref_order, const float4 b0, const float4 b1, const float4 b2,
                                              const float4 b3, uint32_t first, uint32_t geom0, f3 o, f3 d, float tmin, uint32_t skip_a,
                                              uint32_t skip_b, bool strict, Hit& best) {
  const f3 oc = mk3(b0.x, b0.y, b0.z) - o;
  // one axis: entry / exit distance of its slab and the face the ray enters through (the exit face is that ^ 1).
  // A parallel ray (|D| <= 1e-8, Geometry.hs:125) gets a huge finite 1/D: inside the slab the two distances become
  // -huge / +huge (no constraint), outside both take the same sign (the box test below fails) — no branch, no NaN.
  auto axis = [&](f3 e, float h, uint32_t face_minus, float& tn, float& tf, uint32_t& near_face) {
    const float D = dot3(e, d), P = dot3(e, oc);       // local coordinate of the ray: -P + t D
    const float inv = fabsf(D) > 1e-8f ? rcp_approx(D) : copysignf(1e30f, D);
    const float tm = (P - h) * inv, tp = (P + h) * inv; // planes at -h and +h
    tn = fminf(tm, tp);
    tf = fmaxf(tm, tp);
    near_face = face_minus ^ (D > 0.f ? 0u : 1u);      // moving up the axis: enters through the -h face
  };
  float nz, fz, nx, fx, ny, fy;
  uint32_t az, ax, ay;
  axis(mk3(b3.x, b3.y, b3.z), b2.w, 1u, nz, fz, az);   // z: faces 1 (-z) / 0 (+z)
  axis(mk3(b1.x, b1.y, b1.z), b0.w, 2u, nx, fx, ax);   // x: faces 2 (-x) / 3 (+x)
  axis(mk3(b2.x, b2.y, b2.z), b1.w, 5u, ny, fy, ay);   // y: faces 5 (-y) / 4 (+y)
  const float t_in = fmaxf(fmaxf(nz, nx), ny), t_out = fminf(fminf(fz, fx), fy);
  const uint32_t f_in = t_in == nz ? az : (t_in == nx ? ax : ay);
  const uint32_t f_out = (t_out == fz ? az : (t_out == fx ? ax : ay)) ^ 1u;
  const bool crosses = t_in <= t_out;
  const uint32_t g_in = geom0 + f_in, g_out = geom0 + f_out;
  const bool ok_in = crosses & (t_in > tmin) & (g_in != skip_a) & (g_in != skip_b);
  const bool ok_out = crosses & (t_out > tmin) & (g_out != skip_a) & (g_out != skip_b);
  const float t = ok_in ? t_in : t_out;
  const uint32_t idx = first + (ok_in ? f_in : f_out);
  if ((ok_in | ok_out) && (t < best.t || (!strict && t == best.t && evaluated_before(ref_order, idx, best.prim)))) {
    best.t = t;
    best.prim = idx;
    best.front = ok_in;
  }
}

// BVH kernels: the box record comes from global memory.  (Out of line — __noinline__ — was tried to keep the slab test's
// registers out of the traversal loop: the by-reference result then lives on the stack, 80 registers + spills; inline: 72.)
__device__ __forceinline__ void intersect_box_global(const uint32_t* __restrict__ sc_ref_order, const float4* __restrict__ B, uint32_t first, uint32_t geom0, f3 o, f3 d, float tmin,
                                                     uint32_t skip_a, uint32_t skip_b, Hit& best) {
  const F8 b01 = ldg256(B), b23 = ldg256(B + 2);
  intersect_box(sc_ref_order, b01.a, b01.b, b23.a, b23.b, first, geom0, o, d, tmin, skip_a, skip_b, false, best);
}

// one primitive; `skip_a` = geometry the ray ORIGIN lies on, `skip_b` = geometry of a hit at tmin
// that must not be reported again (second boundary query of constantMedium, Geometry.hs:311).
// F = scene feature bits (TF_*): code for absent primitive kinds / instances is compiled out.
// kScan: the caller tests primitives in increasing index order with skip_b == kNone (flat-list kernel): an exact tie
// then always loses to the earlier primitive, and the skip_b comparison disappears.
template <uint32_t F, class Rec, bool kScan = false>
__device__ __forceinline__ void intersect_rec(const DScene& sc, const Rec rec, uint32_t i, const RayT& ray, float tmin,
                                              uint32_t skip_a, uint32_t skip_b, Hit& best) {
  const float4 m4 = rec.meta();
  const uint32_t tf = __float_as_uint(m4.x), geom = __float_as_uint(m4.y);
  const float4 p0 = rec.r0();
  f3 o = ray.o, d = ray.d;
  if constexpr ((F & TF_MOVING) != 0) {
    if (tf & PRIM_FLAG_MOVING) {  // transform / moving instance kept local (Geometry.hs:388-389,452-454)
      RT_ASSERT(__float_as_uint(m4.z) < sc.n_xforms);
      const DXform& X = sc.xforms[__float_as_uint(m4.z)];
      f3 b = fma3(ray.time, ld3(X.binv1), ld3(X.binv0));
      o = mat3(X.Ainv, ray.o) + b;
      d = mat3(X.Ainv, ray.d);
    }
  }
  float t;
  bool front;
  const bool is_sphere = (F & TF_PLANAR) == 0 || ((F & TF_SPHERE) != 0 && (tf & PRIM_TYPE_MASK) == PRIM_SPHERE);
  if (is_sphere) {
    if constexpr ((F & TF_SPHERE) != 0) {
      const bool on_a = geom == skip_a, on_b = !kScan && geom == skip_b;
      if (tf & PRIM_FLAG_BIG) {  // fp64 up to the discriminant, stable roots in fp32
        const float4 p1 = rec.r1();
        double cx = (double)p0.x + (double)p1.x, cy = (double)p0.y + (double)p1.y, cz = (double)p0.z + (double)p1.z;
        double r = (double)p0.w + (double)p1.w;
        double ox = cx - (double)o.x, oy = cy - (double)o.y, oz = cz - (double)o.z;
        double hd = (double)d.x * ox + (double)d.y * oy + (double)d.z * oz;
        double c = ox * ox + oy * oy + oz * oz - r * r;
        // |d|^2 = 1 + O(1e-7) after the fp32 normalise; at |oc| ~ 1e3 the term (|d|^2 - 1) c is NOT negligible
        // against the discriminant, so the quadratic keeps its leading coefficient
        double dd = (double)d.x * (double)d.x + (double)d.y * (double)d.y + (double)d.z * (double)d.z;
        double disc = hd * hd - dd * c;
        if (!(disc >= 0.0)) return;
        float sq = sqrtf((float)disc);
        // q = h + sign(h) sqrt(disc): roots are q/|d|^2 and c/q, neither suffers cancellation
        double q = hd >= 0.0 ? hd + (double)sq : hd - (double)sq;
        float rq = (float)(q * (2.0 - dd)), rc = q != 0.0 ? __fdividef((float)c, (float)q) : 0.f;
        float near_t = fminf(rq, rc), far_t = fmaxf(rq, rc);
        if (on_a || on_b) {
          if (on_a && !(hd > 0.0)) return;
          t = far_t;
          if (!(t > tmin)) return;
        } else {
          t = near_t;
          if (!(t > tmin && t <= best.t)) { t = far_t; if (!(t > tmin)) return; }
        }
        front = (double)t <= hd;
      } else {
        f3 oc = mk3(p0.x, p0.y, p0.z) - o;
        float h = dot3(d, oc);
        f3 l = fma3(-h, d, oc);                 // component of oc perpendicular to the ray
        float disc = fmaf(p0.w, p0.w, -dot3(l, l));
        if (!(disc >= 0.f)) return;
        float sq = sqrtf(disc);
        if (on_a || on_b) {
          if (on_a && !(h > 0.f)) return;
          t = h + sq;
          if (!(t > tmin)) return;
        } else {
          t = h - sq;
          if (!(t > tmin && t <= best.t)) { t = h + sq; if (!(t > tmin)) return; }
        }
        front = t <= h;                          // d.(p-c) = t - h  (Geometry.hs:83-84, `<=`)
      }
      if (p0.w < 0.f) front = !front;
    } else return;
  } else {
    if constexpr ((F & TF_PLANAR) != 0) {
      if constexpr (Rec::kBoxes && (F & TF_BOX) != 0) {
        if (tf & (PRIM_FLAG_BOXHEAD | PRIM_FLAG_BOXTAIL)) {   // aggregated cuboid: one slab test at the head face
          if (tf & PRIM_FLAG_BOXHEAD) intersect_box_global(sc.ref_order, sc.tri_uv + 2 * (size_t)i, i, geom, o, d, tmin, skip_a, skip_b, best);
          return;
        }
      }
      const float4 p1 = rec.r1();
      f3 q = mk3(p0.x, p0.y, p0.z), n = mk3(p0.w, p1.x, p1.y);
      float denom = dot3(n, d);
      f3 qo = q - o;
      t = dot3(n, qo) * rcp_approx(denom);
#ifndef RT_SCAN_BRANCHY
      if constexpr (kScan) {
        // Linear scan over a constant-bank list: every lane tests every primitive, and in a warp of 32 incoherent rays
        // some lane nearly always passes the distance test, so an early exit saves no issue slots — it only adds
        // BSSY/BSYNC pairs and branch-resolve stalls.  Straight-line code, one predicated update at the end.
        // (t is inf/NaN for a parallel ray: every comparison below is then false.)
        const float4 p2 = rec.r2();
        const f3 pr = fma3(t, d, -qo);             // p - q
        const float a = dot3(mk3(p1.z, p1.w, p2.x), pr), b = dot3(mk3(p2.y, p2.z, p2.w), pr);
        const bool pgram = (tf & PRIM_TYPE_MASK) == PRIM_PGRAM;           // uniform (constant bank)
        const float lim = pgram ? fmaxf(a, b) : a + b;                    // a, b <= 1  |  a + b <= 1
        const bool ok = (a >= -kEdgeEps) & (b >= -kEdgeEps) & (lim <= 1.f + kEdgeEps) & (fabsf(denom) > 1e-8f) & (t > tmin) & (t < best.t) & (geom != skip_a);
        if (ok) {
          best.t = t;
          best.prim = i;
          best.front = denom < 0.f;
        }
        return;
      }
#endif
      // one exit for: parallel ray (Geometry.hs:125), outside the open interval, beyond the closest hit so far,
      // the primitive the ray starts on / must not report again
      if (!(fabsf(denom) > 1e-8f && t > tmin && (kScan ? t < best.t : t <= best.t)) || geom == skip_a || (!kScan && geom == skip_b)) return;
      const float4 p2 = rec.r2();
      f3 pr = fma3(t, d, -qo);                   // p - q
      float a = dot3(mk3(p1.z, p1.w, p2.x), pr), b = dot3(mk3(p2.y, p2.z, p2.w), pr);
      if ((tf & PRIM_TYPE_MASK) == PRIM_PGRAM) { if (!(-kEdgeEps <= a && a <= 1.f + kEdgeEps && -kEdgeEps <= b && b <= 1.f + kEdgeEps)) return; }
      else { if (!(a >= -kEdgeEps && b >= -kEdgeEps && a + b <= 1.f + kEdgeEps)) return; }
      front = denom < 0.f;
    } else return;
  }
  if (t < best.t || (!kScan && t == best.t && evaluated_before(sc.ref_order, i, best.prim))) {
    best.t = t;
    best.prim = i;
    best.front = front;
  }
}

// Tests primitive i and returns how many consecutive primitives that answered (6 for the head face of an aggregated
// cuboid, else 1): leaf loops advance by it.
template <uint32_t F>
__device__ __forceinline__ uint32_t intersect_prim(const DScene& sc, uint32_t i, const RayT& ray, float tmin,
                                                   uint32_t skip_a, uint32_t skip_b, Hit& best) {
  RT_ASSERT(i < sc.n_prims);
  const float4* P = sc.prims + 4 * (size_t)i;
  const GlobalRec rec{ldg256(P), ldg256(P + 2)};
  intersect_rec<F>(sc, rec, i, ray, tmin, skip_a, skip_b, best);
  if constexpr ((F & TF_BOX) != 0) return (__float_as_uint(rec.hi.b.x) & PRIM_FLAG_BOXHEAD) ? 6u : 1u;
  return 1u;
}

// TF_MOVLIST: the scene's moving / instanced primitives live outside the tree; every closest-hit query over the SURFACES
// scans them once.
template <uint32_t F>
__device__ __forceinline__ void scan_moving(const DScene& sc, const RayT& ray, float tmin, uint32_t skip_a, Hit& best) {
  for (uint32_t k = 0; k < sc.moving_count;) k += intersect_prim<(F & ~TF_MOVLIST) | TF_MOVING>(sc, sc.moving_first + k, ray, tmin, skip_a, kNone, best);
}

// slab test of one 32-byte node against the clipped interval (tmin, tmax]
__device__ __forceinline__ bool box_hit(const float4 lo, const float4 hi, f3 inv, f3 oinv, float tmin, float tmax, float& tn) {
  float x0 = fmaf(lo.x, inv.x, -oinv.x), x1 = fmaf(hi.x, inv.x, -oinv.x);
  float y0 = fmaf(lo.y, inv.y, -oinv.y), y1 = fmaf(hi.y, inv.y, -oinv.y);
  float z0 = fmaf(lo.z, inv.z, -oinv.z), z1 = fmaf(hi.z, inv.z, -oinv.z);
  tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
  float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
  return tn <= tf;   // fminf/fmaxf drop NaN operands: a NaN slab never culls (Appendix A.8)
}

// Closest hit below node reference `root` ("while-while" traversal: every lane first descends through inner
// nodes until it holds a leaf, then the warp tests primitives together — keeps the expensive primitive code
// SIMD-dense).  `stack` is this thread's column of the shared-memory short stack (stride words apart).
// `best` carries (tmax, kNone) in and the result out.
template <uint32_t F>
__device__ __forceinline__ void closest_hit(const DScene& sc, uint32_t root, const RayT& ray, float tmin,
                                            uint32_t skip_a, uint32_t skip_b, uint32_t* stack, uint32_t stride, Hit& best) {
  f3 inv;
  inv.x = fabsf(ray.d.x) > 1e-30f ? __frcp_rn(ray.d.x) : copysignf(1e30f, ray.d.x);
  inv.y = fabsf(ray.d.y) > 1e-30f ? __frcp_rn(ray.d.y) : copysignf(1e30f, ray.d.y);
  inv.z = fabsf(ray.d.z) > 1e-30f ? __frcp_rn(ray.d.z) : copysignf(1e30f, ray.d.z);
  const f3 oinv = ray.o * inv;
  uint32_t cur = root;          // kNone (leaf bit set, all ones) = nothing left
  int sp = 0;
  while (cur != kNone) {
    while (!(cur & kLeafBit)) {
      RT_ASSERT(cur + 1u < sc.n_nodes && (cur & 1u) == 0u);
      const float4* N = sc.nodes + 2 * (size_t)cur;
      const F8 c0 = ldg256(N), c1 = ldg256(N + 2);
      const float4 l0 = c0.a, h0 = c0.b, l1 = c1.a, h1 = c1.b;
      float tn0, tn1;
      const bool hit0 = box_hit(l0, h0, inv, oinv, tmin, best.t, tn0);
      const bool hit1 = box_hit(l1, h1, inv, oinv, tmin, best.t, tn1);
      const uint32_t r0 = __float_as_uint(l0.w), r1 = __float_as_uint(l1.w);
      if (hit0 && hit1) {
        const bool first0 = tn0 <= tn1;     // near child first; results do not depend on the order
        RT_ASSERT((uint32_t)sp < sc.stack_entries);
        stack[sp * stride] = first0 ? r1 : r0;
        sp++;
        cur = first0 ? r0 : r1;
      } else if (hit0) cur = r0;
      else if (hit1) cur = r1;
      else if (sp > 0) { sp--; cur = stack[sp * stride]; }
      else cur = kNone;
    }
    if (cur == kNone) break;
    const uint32_t first = (cur & ~kLeafBit) >> kLeafShift, count = (cur & kLeafCountMask) + 1u;
    for (uint32_t k = 0; k < count;) k += intersect_prim<F>(sc, first + k, ray, tmin, skip_a, skip_b, best);
    if (sp > 0) { sp--; cur = stack[sp * stride]; }
    else cur = kNone;
  }
}

// A segment that starts at a scatter point of the scene's ONLY medium, whose boundary is part of the surfaces
// (`dielectric <$ mesh` + `constantMedium rho mesh`: subsurface scattering) and a closed surface (DMedium::closed: the
// first boundary hit from an interior point is then always a back face), is inside that medium from tmin on, so the
// reference's outcome (Geometry.hs:313-327) is: the closest surface hit if it comes before tmin + d, else a scatter event
// at tmin + d, with d = -ln(rand)/density drawn from the counter RNG - the same number whether it is drawn before or
// after the traversal.  Drawing it first turns the unbounded surface query of an interior ray (which crosses the whole
// mesh to find where it would leave) into one bounded by the free-flight distance: a mean free path of 0.2 in a pawn four
// units tall culls nearly every node.  Returns the bound (inf: not applicable / rand = 0).
__device__ __forceinline__ bool starts_inside_only_medium(const DScene& sc, uint32_t origin_geom) {
  const bool scattered = origin_geom == kMediumBase;                                         // starts at a scatter point
  const bool inwards = (origin_geom & kInsideFlag) != 0u && origin_geom < kMediumBase;       // leaves the boundary into the solid
  return (scattered || inwards) && sc.n_media == 1u && sc.media[0].closed != 0u;             // closed implies shared
}
__device__ __forceinline__ uint32_t strip_inside_flag(uint32_t origin_geom) {
  return origin_geom < kMediumBase ? origin_geom & ~kInsideFlag : origin_geom;
}
__device__ __forceinline__ float free_flight_bound(const DScene& sc, float tmin, uint2 key, uint32_t pixel, uint64_t sample, uint32_t bounce) {
  const uint4 rnd = rng4(key, pixel, sample, bounce, RNG_MEDIA);
  return tmin + sc.media[0].neg_inv_density * logf(u01(rnd.x));   // rand = 0 -> +inf -> unbounded, no scatter (as in the reference)
}

// ---- full closest-hit query of a path segment: surfaces, then every medium (Geometry.hs:298-330) ---
template <uint32_t F>
__device__ __forceinline__ Hit trace_segment(const DScene& sc, const RayT& ray, uint32_t origin_geom, bool with_media,
                                             uint2 key, uint32_t pixel, uint64_t sample, uint32_t bounce,
                                             uint32_t* stack, uint32_t stride) {
  const float tmin = 1e-4f;  // Ray.hs:178
  const float inf = __int_as_float(0x7f800000);
  Hit best{inf, kNone, false};
  if constexpr ((F & TF_MEDIA) != 0) {
    if (with_media && starts_inside_only_medium(sc, origin_geom)) best.t = free_flight_bound(sc, tmin, key, pixel, sample, bounce);
    origin_geom = strip_inside_flag(origin_geom);
  }
  // the bounded query looks at (0, bound]: a scatter point lies on no surface, and a boundary crossing inside (0, tmin] -
  // which the reference's open interval skips, so that it sees the ray as already outside - must be noticed
  if constexpr ((F & TF_MOVLIST) != 0) scan_moving<F>(sc, ray, best.t < inf ? 0.f : tmin, origin_geom, best);
  closest_hit<F>(sc, sc.surface_root, ray, best.t < inf ? 0.f : tmin, origin_geom, kNone, stack, stride, best);
  if constexpr ((F & TF_MEDIA) != 0) {
    if (best.prim == kNone && best.t < inf) {   // nothing before the free-flight distance: the scatter event
      best.prim = kMediumBase;
      best.front = true;
      return best;
    }
    if (best.prim != kNone && !(best.t > tmin)) {   // crossing within tmin: an ordinary segment after all (~1e-4 of them)
      best = Hit{inf, kNone, false};
      if constexpr ((F & TF_MOVLIST) != 0) scan_moving<F>(sc, ray, tmin, origin_geom, best);
      closest_hit<F>(sc, sc.surface_root, ray, tmin, origin_geom, kNone, stack, stride, best);
    }
    if (with_media) {
      uint4 rnd = make_uint4(0, 0, 0, 0);
      for (uint32_t m = 0; m < sc.n_media; m++) {
        if ((m & 3u) == 0) rnd = rng4(key, pixel, sample, bounce, RNG_MEDIA + (m >> 2));
        const uint32_t rbits = (m & 3u) == 0 ? rnd.x : ((m & 3u) == 1 ? rnd.y : ((m & 3u) == 2 ? rnd.z : rnd.w));
        const DMedium& md = sc.media[m];
        // boundary query 1 on (tmin, inf); if it is a front-side hit before the current closest hit, query 2 on
        // (t1, inf) finds where the ray leaves again (Geometry.hs:307-313).  One code copy serves both.
        float t1 = tmin, t2 = 0.f, qmin = tmin;
        uint32_t skip_b = kNone;
        bool have_segment = false;
        for (int pass = 0; pass < 2; pass++) {
          Hit h{inf, kNone, false};
          bool answered = false;
          if (pass == 0 && md.shared && (best.prim == kNone || best.prim < kMediumBase)) {
            // boundary ⊆ surfaces (DMedium::shared): a surface miss is a boundary miss, and a surface hit ON the
            // boundary is the closest boundary hit
            if (best.prim == kNone) break;
            const uint32_t g = __float_as_uint(__ldg(sc.prims + 4 * (size_t)best.prim + 3).y);
            if (g >= md.geom_lo && g <= md.geom_hi) { h = best; answered = true; }
          }
          if (!answered) closest_hit<F>(sc, md.root, ray, qmin, origin_geom, skip_b, stack, stride, h);
          if (h.prim == kNone) break;
          if (pass == 0) {
            if (!h.front) { t1 = tmin; t2 = fminf(best.t, h.t); have_segment = true; break; }
            if (!(h.t < best.t)) break;
            t1 = qmin = h.t;
            skip_b = __float_as_uint(__ldg(sc.prims + 4 * (size_t)h.prim + 3).y);
          } else {
            t2 = fminf(best.t, h.t);
            have_segment = true;
          }
        }
        if (!have_segment) continue;
        const float hit_dist = md.neg_inv_density * logf(u01(rbits));   // rand = 0 → +inf → no hit
        if (hit_dist < t2 - t1) {
          best.t = t1 + hit_dist;
          best.prim = kMediumBase + m;
          best.front = true;
        }
      }
    }
  }
  return best;
}

}  // namespace rtb
