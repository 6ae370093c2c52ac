// rt_device.h — flattened scene records shared by the host flattener (rt_flatten.cpp) and the
// CUDA kernels (rt_kernels.cu).  Everything here is what lives in HBM after rt_scene_upload.
//
// Layout summary (DESIGN.md §"Data layout in HBM"):
//   nodes   : 32-byte BVH nodes, 2 x float4 {lo.xyz, a} {hi.xyz, b}; siblings adjacent (one 64 B
//             fetch tests both children).  a = node reference: inner → index of the pair holding its children;
//             leaf → kLeafBit | first primitive << 5 | (count - 1), count <= 32.  b = primitive count (leaf) or 0.
//             Roots are bare references (no node, no box test of their own, as in the reference).
//   prims   : 64-byte records, 4 x float4, stored in the leaf order of the device tree (a leaf = a contiguous range,
//             in the reference's evaluation order inside the leaf).  The reference's evaluation order (group: list
//             order, Geometry.hs:340-345; bvhNode: left first, :356-361) is kept in `ref_order`: on an exact tie
//             the primitive that the reference evaluates first wins, whatever the shape of the device tree.
//             sphere : {c.xyz, r}        {-}                {-}                  {meta}
//             planar : {q.xyz, n.x} {n.y, n.z, wa.xy} {wa.z, wb.xyz} {meta}
//                      with wa = v x nS, wb = nS x u (planar coords a = wa.(p-q), b = wb.(p-q);
//                      Geometry.hs:130-131 rewritten by the scalar triple product)
//             meta   : {type|flags, geom id, xform id, material id} as uint32 bit patterns
//   tri_uv  : 32 bytes of auxiliary data per primitive: a triangle's {uv0, uv1} {uv2, -, -}; for an AGGREGATED CUBOID
//             (below) the 80-byte box record {c.xyz, h0} {e0.xyz, h1} {e1.xyz, h2} {e2.xyz, -} (centre, unit axes = images
//             of the local x, y, z axes, half extents) starts in the head face's slot and runs over the next two
//             (parallelogram faces have no uv data of their own)
// Aggregated cuboids: the reference's `cuboid` is a group of six parallelograms (Geometry.hs:154-166).  All six face
// records stay in `prims` (shading reads the face that was hit), but traversal answers the closest hit of the whole
// solid with ONE slab test at the head face (PRIM_FLAG_BOXHEAD, ~60 instructions instead of 6 x 33) and treats the
// other five (PRIM_FLAG_BOXTAIL) as already answered.  The six faces are consecutive in index order (+z -z -x +x +y -y),
// their geometry ids are consecutive too, and the flattener never splits them across leaves.
#pragma once
#include <stdint.h>
#ifdef RT_CHECKED
#include <assert.h>   // (outside the namespace: other headers use assert() too)
#endif

namespace rtb {

constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kLeafShift = 5u, kLeafCountMask = 31u, kLeafCap = 32u;   // leaf ref = kLeafBit | first << 5 | (count - 1)
constexpr uint32_t kMediumBase = 0xFFFF0000u;  // hit.prim >= kMediumBase: medium (prim - base)
// origin-geometry word of a queued path (Queue::d.w): the geometry id of the surface the ray starts on (< 2^26), kNone,
// or kMediumBase + k after a scatter event in medium k.  kInsideFlag on a surface id: the surface is the closed boundary
// of the scene's only medium and the ray leaves it INWARDS (refracted into the solid, or reflected back from inside), so
// the segment starts inside the medium like one that starts at a scatter point (rt_trace.cuh::starts_inside_only_medium).
// Traversal strips the flag before the id is compared with anything.
constexpr uint32_t kInsideFlag = 0x80000000u;
constexpr int kStackDepth = 32;                // shared-memory short stack entries per thread
constexpr int kMaxTargets = 8;                 // cs_redirectTargets entries supported on device
constexpr int kMaxMedia = 16;

// primitive type / flag bits (meta.x)
constexpr uint32_t PRIM_SPHERE = 0u;
constexpr uint32_t PRIM_PGRAM = 1u;
constexpr uint32_t PRIM_TRI = 2u;
constexpr uint32_t PRIM_TYPE_MASK = 3u;
constexpr uint32_t PRIM_FLAG_BIG = 4u;      // large-radius sphere: fp64 discriminant
constexpr uint32_t PRIM_FLAG_MOVING = 8u;   // stays in local space; ray is mapped by xform(time)
constexpr uint32_t PRIM_FLAG_UVXFORM = 16u; // sphere under a rotation: uv needs the local normal
constexpr uint32_t PRIM_FLAG_BOXHEAD = 32u; // face 0 of an aggregated cuboid: traversal tests the whole box here
constexpr uint32_t PRIM_FLAG_BOXTAIL = 64u; // faces 1..5 of an aggregated cuboid: answered by the head

// scene feature bits: which code the traversal kernel must contain (template specialisation)
constexpr uint32_t TF_SPHERE = 1u, TF_PLANAR = 2u, TF_MOVING = 4u, TF_MEDIA = 8u;
constexpr uint32_t TF_BOX = 16u;   // some cuboids are aggregated (PRIM_FLAG_BOXHEAD / BOXTAIL occur)
// The scene's few moving / instanced primitives are kept OUT of the tree, in the short list [moving_first, moving_first +
// moving_count) that every segment scans once: the thousands of static primitives then run the traversal code without the
// per-primitive transform branch (demo2: one moving sphere used to put all 3407 primitives on the catch-all kernel).
// TF_MOVING (moving primitives inside the tree, transform branch in every primitive test) remains the fallback.
constexpr uint32_t TF_MOVLIST = 32u;
constexpr uint32_t kMaxMovingList = 32u;

struct DNode {  // 32 bytes
  float lo[3];
  uint32_t a;
  float hi[3];
  uint32_t b;
};

// One input record of the device-side tree build (rt_build.cuh): an item is what a leaf may not split - one primitive, or the
// six faces of an aggregated cuboid.  Items are listed in the reference's evaluation order; `prov` is the first of the item's
// `n` consecutive PROVISIONAL storage slots (the order the host packed the records in), the box is rounded outward from fp64.
struct BuildItem {  // 32 bytes = two float4
  float lo[3];
  uint32_t prov;
  float hi[3];
  uint32_t n;
};

struct DXform {  // world->local: o' = Ainv*o + binv0 + time*binv1 ; local->world normal: A*n
  float Ainv[9];
  float binv0[3], binv1[3];
  float A[9];
};

struct DMaterial {  // Material.hs
  int32_t kind;     // RtMaterialKind
  uint32_t texture;
  float param;
  int32_t _pad;
};

struct DTexture {   // Texture.hs
  int32_t kind;     // RtTextureKind
  int32_t i0, i1;   // checker n_u n_v / noise layers
  int32_t w, h;     // image size
  uint32_t image_offset;  // first texel (float4) in the image pool
  float freq;
  float c0[3], c1[3], shift[3], dir[3];
};

struct DMedium {    // constantMedium, Geometry.hs:298-330
  float neg_inv_density;
  uint32_t root;        // node reference of the boundary tree
  uint32_t material;
  // Boundary primitives whose geometry ids form the dense range [geom_lo, geom_hi] AND which are all also part of the
  // surface tree (e.g. `dielectric <$ mesh` + `constantMedium rho mesh`): the surface closest hit then answers the first
  // boundary query (closest over a subset = closest over the set when it lies in the subset).  shared = 0 otherwise.
  uint32_t geom_lo, geom_hi, shared;
  // shared boundary that is a CLOSED, outward-oriented surface (a watertight triangle mesh: every directed edge once and
  // its reverse once, positive volume; or one sphere): from a point inside, every ray leaves through a back face, so a
  // path that has just scattered in this medium is known to be inside it (rt_trace.cuh::starts_inside_only_medium).
  // The Stanford bunny has holes -> 0; the pawn is watertight -> 1.
  uint32_t closed;
};

struct DTarget {    // RedirectTarget, Ray.hs:112-118 (+ its parallelogram record for the pdf probe)
  float p, cumulative;
  float origin[3], U[3], V[3], crs[3];
  float n[3], wa[3], wb[3];
};

struct DCamera {    // Ray.hs:123-155, evaluated in double on the host, stored fp32
  int32_t W, H, spp, max_depth;
  // top_left is stored RELATIVE to the camera centre (|.| ~ focus distance), so that pixel offsets are not
  // rounded at world-coordinate magnitude (Cornell: 278 → 3e-6 direction error otherwise)
  float center[3], top_left[3], pixel_u[3], pixel_v[3], defocus_u[3], defocus_v[3];
  int32_t has_defocus;
  int32_t bg_kind;
  float bg0[3], bg1[3];
  int32_t n_targets;
  float rem_prob;
  DTarget targets[kMaxTargets];
};

struct DScene {     // device pointers of one uploaded scene
  const float4* nodes;
  const float4* prims;
  const float4* tri_uv;
  const DXform* xforms;
  const DMaterial* materials;
  const DTexture* textures;
  const float4* texels;
  const uint32_t* canonical;   // prim index -> canonical primitive id (probe output)
  const uint32_t* ref_order;   // prim index -> position in the reference's evaluation order (exact ties only)
  const float* noise_grad;     // 256*3
  const uint8_t* noise_perm;   // 3*256
  DMedium media[kMaxMedia];
  uint32_t n_media;
  uint32_t surface_root;       // node reference or kNone
  uint32_t has_noise;
  uint32_t n_prims;
  // conservative world-space box of everything a ray can meet (all surface primitives and medium boundaries, swept over
  // the shutter, rounded outward); has_world_box = 0 when it is not finite.  Camera rays that miss it end on the
  // background inside k_raygen and never enter the queue (rt_kernels.cu::k_raygen).
  float world_lo[3], world_hi[3];
  uint32_t has_world_box;
  uint32_t moving_first, moving_count;   // TF_MOVLIST: the moving primitives, outside the tree (scanned once per segment)
  // array sizes, read only by the RT_CHECKED build's assertions (RT_ASSERT below)
  uint32_t n_nodes, n_materials, n_textures, n_xforms, stack_entries;
};

// -DRT_CHECKED: every index the kernels compute (queue slot, node, primitive, stack entry, pixel, material, texture) is
// asserted against the size of the array it goes into.  compute-sanitizer is not available on the GPU pool this was developed
// on, so the checked build run through the whole GPU test suite stands in for a memcheck pass (profiles/r02u_checked_build.txt).
#ifdef RT_CHECKED
#define RT_ASSERT(c) assert(c)
#else
#define RT_ASSERT(c) ((void)0)
#endif

}  // namespace rtb
