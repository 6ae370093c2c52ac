// rt_flatten.cpp — description tree → flattened device scene (host side, double precision).
//
// Replaces, for the GPU path, what the reference does implicitly by nesting closures
// (src/Graphics/Ray/Geometry.hs): every combinator of the tree is resolved HERE, once:
//   * `transform` / `moving` chains (Geometry.hs:382-391, 449-456) collapse into one affine map
//     o' = Ainv*o + binv0 + time*binv1 per leaf; static chains are baked into world-space
//     primitives, moving ones keep the primitive local and flag it;
//   * `<$` (Geometry.hs:44-47): the OUTERMOST material wins (fmap (const m) applied last);
//   * `constantMedium` (Geometry.hs:298-330) nodes are hoisted out of the surface tree: closest
//     hit is order independent, so "surfaces first, then each medium on (tmin, t_best)" gives
//     the same distribution as the reference's left-to-right `group` fold;
//   * `bvhTree` (Geometry.hs:369-377) is built with the reference's policy — axis = longestDim
//     of the joined box with argMax's tie rule (Core.hs:43-46,143-144), STABLE sort on the bbox
//     midpoint (Core.hs:87-88), left gets n div 2 — on the reference's own bounding boxes;
//   * `group` (Geometry.hs:336-347), a linear scan in the reference, becomes a BVH built by the
//     same policy (or one multi-primitive leaf when small): identical results, fewer tests;
//   * primitives are numbered in the reference's evaluation order so that "lower index wins an
//     exact tie" equals the reference's tie rules (earlier child / left child wins).
#include <algorithm>
#include <array>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <unordered_map>
#include <unordered_set>

#include "rt_host.h"

namespace rtb {
namespace {

constexpr double kInf = std::numeric_limits<double>::infinity();

struct D3 {
  double x = 0, y = 0, z = 0;
};
inline D3 operator+(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline D3 operator-(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline D3 operator*(double s, D3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 mk(const double* p) { return {p[0], p[1], p[2]}; }
inline double get(D3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

struct BoxD {
  double lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
  void add(D3 p) {
    for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], get(p, k)); hi[k] = std::max(hi[k], get(p, k)); }
  }
  void join(const BoxD& b) {
    for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
  }
  void pad(double e) {
    for (int k = 0; k < 3; k++) { lo[k] -= e; hi[k] += e; }
  }
  bool empty() const { return lo[0] > hi[0]; }
};
inline int longest_dim(const BoxD& b) {  // Core.hs:143-144 + argMax Core.hs:43-46
  double x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
  if (x > y) return x > z ? 0 : 2;
  return y > z ? 1 : 2;
}

// affine map p -> M*p + c0 + time*c1
struct Affine {
  double M[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double c0[3] = {0, 0, 0}, c1[3] = {0, 0, 0};
  D3 lin(D3 v) const { return {M[0] * v.x + M[1] * v.y + M[2] * v.z, M[3] * v.x + M[4] * v.y + M[5] * v.z, M[6] * v.x + M[7] * v.y + M[8] * v.z}; }
  D3 pt(D3 p, double time) const {
    D3 r = lin(p);
    return {r.x + c0[0] + time * c1[0], r.y + c0[1] + time * c1[1], r.z + c0[2] + time * c1[2]};
  }
};
// (outer ∘ inner)(p) = outer(inner(p)); motion terms stay linear in time because the linear
// parts do not depend on time.
Affine compose(const Affine& outer, const Affine& inner) {
  Affine r;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += outer.M[i * 3 + k] * inner.M[k * 3 + j];
      r.M[i * 3 + j] = s;
    }
  D3 a0 = outer.lin(mk(inner.c0)), a1 = outer.lin(mk(inner.c1));
  r.c0[0] = a0.x + outer.c0[0]; r.c0[1] = a0.y + outer.c0[1]; r.c0[2] = a0.z + outer.c0[2];
  r.c1[0] = a1.x + outer.c1[0]; r.c1[1] = a1.y + outer.c1[1]; r.c1[2] = a1.z + outer.c1[2];
  return r;
}

struct XfState {          // accumulated transform of the current subtree
  Affine fwd;             // local -> world
  Affine inv;             // world -> local
  bool identity = true, moving = false;
  std::vector<uint32_t> chain;  // description node ids of the transform/moving nodes above
};

struct HPrim {            // one primitive instance (double precision, final space)
  uint32_t type = PRIM_SPHERE, flags = 0;
  D3 c;                   // sphere centre / planar q
  double r = 0;
  D3 n, wa, wb;           // planar
  double uv[3][2] = {{0, 0}, {1, 0}, {0, 1}};
  uint32_t geom = kNone, xform = 0, material = kNone, canonical = kNone;
  BoxD box;               // world-space box (swept over the shutter for moving instances)
  uint32_t ref_index = kNone;   // position in the reference's evaluation order (decides exact ties)
  uint32_t index = kNone;       // device (storage) index: leaf order of the device tree
  D3 corner[3];           // triangles: the three vertices as given (local space) - watertightness test of medium boundaries
  bool has_corners = false;
  D3 box_c, box_e[3];     // PRIM_FLAG_BOXHEAD: centre and unit axes of the aggregated cuboid (same space as c / n)
  double box_h[3] = {0, 0, 0};   // ... and its half extents
};

// reference-shaped tree: exactly the structure the reference would evaluate
struct RNode {
  enum Kind { LEAF, GROUP, BVH } kind = LEAF;
  BoxD ref_box;                     // the reference's bounding box of this sub-geometry
  uint32_t prim = kNone;            // LEAF: index into Builder::prims
  std::vector<int> children;        // GROUP: list order; BVH: {left, right}
};

// device-shaped tree
struct HNode {
  BoxD box;
  int left = -1, right = -1;        // inner
  uint32_t first = 0, count = 0;    // leaf: final primitive indices [first, first+count)
};

struct FlattenError {
  int32_t status;
  std::string msg;
};

struct Builder {
  Scene& sc;
  std::vector<HPrim> prims;
  std::vector<RNode> rnodes;
  std::vector<HNode> hnodes;
  std::map<std::vector<uint32_t>, uint32_t> xform_ids;
  std::vector<XfState> xforms;      // index = xform id (0 = identity)
  // Ids are handed out in blocks of one description node's sub-shapes (1; 6 cuboid faces; a mesh's triangles), at the
  // node's first encounter: the same numbering as one id per (node, sub) in walk order, with one map lookup per node
  // instead of one per triangle.
  std::map<uint32_t, uint32_t> canon_base;                          // desc node -> canonical id of its sub-shape 0
  std::map<std::pair<uint32_t, uint32_t>, uint32_t> geom_base;      // (desc node, xform) -> geometry id of sub-shape 0
  uint32_t next_canonical = 0, next_geom = 0;
  std::vector<uint32_t> memo_chain;
  uint32_t memo_chain_id = kNone;
  uint32_t memo_node = kNone, memo_xf = kNone, memo_canon = 0, memo_geom = 0;   // last (node, xform) looked up
  struct MeshMemo { uint32_t p0, p1; int r0, r1, root; };          // prims / rnodes ranges of a walked mesh, and its tree's root
  std::map<std::pair<uint32_t, uint32_t>, MeshMemo> mesh_memo;      // (desc node, xform id)
  struct MediumRec { double density; int rtree; uint32_t material; };
  std::vector<MediumRec> media;
  std::vector<char> mat_used, tex_used;
  int depth_guard = 0;
  double ms_ref_tree = 0;           // time inside the meshes' reference-policy trees (trace_commit)

  explicit Builder(Scene& s) : sc(s) {
    XfState id;
    xforms.push_back(id);
    xform_ids[{}] = 0;
  }

  [[noreturn]] void fail(int32_t st, const std::string& m) { throw FlattenError{st, m}; }

  uint32_t sub_shapes(uint32_t node) const {
    const HostNode& n = sc.nodes[node];
    if (n.kind == RT_NODE_CUBOID) return 6;
    if (n.kind == RT_NODE_MESH) return (uint32_t)(n.faces.size() / 6);
    return 1;
  }
  void assign_ids(HPrim& p, uint32_t node, uint32_t sub) {
    if (node != memo_node || p.xform != memo_xf) {
      const uint32_t k = sub_shapes(node);
      auto c = canon_base.find(node);
      if (c == canon_base.end()) { c = canon_base.emplace(node, next_canonical).first; next_canonical += k; }
      auto g = geom_base.find({node, p.xform});
      if (g == geom_base.end()) { g = geom_base.emplace(std::make_pair(node, p.xform), next_geom).first; next_geom += k; }
      memo_node = node; memo_xf = p.xform; memo_canon = c->second; memo_geom = g->second;
    }
    p.canonical = memo_canon + sub;
    p.geom = memo_geom + sub;
  }
  uint32_t xform_id(const XfState& x) {
    if (x.identity) return 0;
    if (memo_chain_id != kNone && x.chain == memo_chain) return memo_chain_id;   // consecutive faces of one instance
    auto it = xform_ids.find(x.chain);
    uint32_t id;
    if (it != xform_ids.end()) id = it->second;
    else {
      id = (uint32_t)xforms.size();
      xforms.push_back(x);
      xform_ids[x.chain] = id;
    }
    memo_chain = x.chain;
    memo_chain_id = id;
    return id;
  }

  int new_rnode(RNode::Kind k) {
    rnodes.emplace_back();
    rnodes.back().kind = k;
    return (int)rnodes.size() - 1;
  }

  // ---- primitives ---------------------------------------------------------------------------
  void use_material(uint32_t m) {
    if (m == kNone) return;
    if (m >= sc.materials.size()) fail(RT_ERR_INVALID_ARG, "material id out of range");
    const RtMaterialDesc& d = sc.materials[m];
    if (d.kind == RT_MAT_USER)
      fail(RT_ERR_USER_DEFINED, "scene uses a user-defined Material closure; only the library's own "
                                "materials (Material.hs:41-129) can be flattened — no CPU fallback");
    if (d.kind < RT_MAT_LIGHT || d.kind > RT_MAT_USER) fail(RT_ERR_INVALID_ARG, "unknown material kind");
    mat_used[m] = 1;
    if (d.kind != RT_MAT_PITCH_BLACK && d.kind != RT_MAT_DIELECTRIC) {
      if (d.texture >= sc.textures.size()) fail(RT_ERR_INVALID_ARG, "texture id out of range");
      int tk = sc.textures[d.texture].d.kind;
      if (tk == RT_TEX_USER)
        fail(RT_ERR_USER_DEFINED, "scene uses a user-defined Texture closure (solidTexture / uvTexture / "
                                  "Texture); only Texture.hs:18-79's constructors can be flattened — no CPU fallback");
      if (tk < RT_TEX_CONSTANT || tk > RT_TEX_USER) fail(RT_ERR_INVALID_ARG, "unknown texture kind");
      tex_used[d.texture] = 1;
    }
  }

  int leaf_for(HPrim&& p, const BoxD& ref_box) {
    prims.push_back(std::move(p));
    int r = new_rnode(RNode::LEAF);
    rnodes[r].prim = (uint32_t)prims.size() - 1;
    rnodes[r].ref_box = ref_box;
    return r;
  }

  void finish_prim(HPrim& p, const XfState& xf, uint32_t node, uint32_t sub, uint32_t material, const BoxD& local_box) {
    p.xform = xform_id(xf);
    assign_ids(p, node, sub);
    p.material = material;
    // world box: corners of the local box mapped forward (at both shutter ends when moving)
    BoxD wb;
    for (int i = 0; i < 8; i++) {
      D3 c{(i & 1) ? local_box.hi[0] : local_box.lo[0], (i & 2) ? local_box.hi[1] : local_box.lo[1], (i & 4) ? local_box.hi[2] : local_box.lo[2]};
      wb.add(xf.fwd.pt(c, 0));
      if (xf.moving) wb.add(xf.fwd.pt(c, 1));
    }
    p.box = wb;
    if (xf.moving) p.flags |= PRIM_FLAG_MOVING;
  }

  int make_sphere(uint32_t node, uint32_t sub, D3 c, double r, const XfState& xf, uint32_t material) {
    HPrim p;
    p.type = PRIM_SPHERE;
    BoxD ref;  // Geometry.hs:60-61
    ref.add({c.x - r, c.y - r, c.z - r});
    ref.add({c.x + r, c.y + r, c.z + r});
    BoxD local = ref;
    if (xf.moving) { p.c = c; }
    else { p.c = xf.fwd.pt(c, 0); }
    p.r = r;
    finish_prim(p, xf, node, sub, material, local);
    if (!xf.moving) {  // exact world box of a sphere under a Euclidean map
      BoxD wb;
      wb.add({p.c.x - std::fabs(r), p.c.y - std::fabs(r), p.c.z - std::fabs(r)});
      wb.add({p.c.x + std::fabs(r), p.c.y + std::fabs(r), p.c.z + std::fabs(r)});
      p.box = wb;
      // uv (sphereUV, Geometry.hs:100-104) is a function of the LOCAL outward normal
      bool rot = false;
      for (int i = 0; i < 9; i++) if (std::fabs(xf.fwd.M[i] - ((i % 4 == 0) ? 1.0 : 0.0)) > 1e-15) rot = true;
      if (rot) p.flags |= PRIM_FLAG_UVXFORM;
    } else p.flags |= PRIM_FLAG_UVXFORM;
    // fp32 cannot hold |oc|^2 - r^2 for large radii next to small scene detail → fp64 discriminant
    if (std::fabs(r) >= 64.0) p.flags |= PRIM_FLAG_BIG;
    return leaf_for(std::move(p), ref);
  }

  // planeShape (Geometry.hs:117-144) for the two library shapes
  // `corners` (triangles): the three vertices as given, so that the bounding box equals the reference's
  // `boxHull [p0, p1, p2]` bit for bit (p0 + (p1 - p0) need not round back to p1).
  int make_planar(uint32_t node, uint32_t sub, bool tri, D3 q, D3 u, D3 v, const double uv[3][2], const XfState& xf, uint32_t material,
                  const D3* corners = nullptr) {
    HPrim p;
    p.type = tri ? PRIM_TRI : PRIM_PGRAM;
    D3 cp = cross(u, v);
    double ncp = std::sqrt(dot(cp, cp));
    D3 n{0, 0, 0}, wa{0, 0, 0}, wb{0, 0, 0};
    if (ncp > 0 && std::isfinite(ncp)) {
      n = (1.0 / ncp) * cp;
      D3 nS = (1.0 / ncp) * n;
      wa = cross(v, nS);   // a = nS . ((p-q) x v) = (p-q) . (v x nS)
      wb = cross(nS, u);   // b = nS . (u x (p-q)) = (p-q) . (nS x u)
    }  // degenerate shape: normal is NaN in the reference → `abs denom > 1e-8` is False → never hit
    BoxD hull;
    if (corners) { hull.add(corners[0]); hull.add(corners[1]); hull.add(corners[2]); }
    else { hull.add(q); hull.add(q + u); hull.add(q + v); if (!tri) hull.add(q + u + v); }
    BoxD ref = hull;
    ref.pad(0.0001);       // Geometry.hs:144
    if (xf.moving) { p.c = q; p.n = n; p.wa = wa; p.wb = wb; }
    else { p.c = xf.fwd.pt(q, 0); p.n = xf.fwd.lin(n); p.wa = xf.fwd.lin(wa); p.wb = xf.fwd.lin(wb); }
    if (uv) std::memcpy(p.uv, uv, sizeof p.uv);
    if (corners && tri) { p.corner[0] = corners[0]; p.corner[1] = corners[1]; p.corner[2] = corners[2]; p.has_corners = true; }
    finish_prim(p, xf, node, sub, material, hull);
    return leaf_for(std::move(p), ref);
  }

  // stable sort of node ids by box centre along axis d (sortOn, Geometry.hs:375): keys are gathered once into a
  // contiguous array instead of being recomputed through two indirections per comparison
  template <class BoxOf>
  void sort_by_centre(std::vector<int>& ids, BoxOf box_of, int d) {
    sort_keys.resize(ids.size());
    for (size_t i = 0; i < ids.size(); i++) {
      const BoxD& b = box_of(ids[i]);
      sort_keys[i] = {(b.lo[d] + b.hi[d]) / 2, ids[i]};
    }
    std::stable_sort(sort_keys.begin(), sort_keys.end(),
                     [](const std::pair<double, int>& a, const std::pair<double, int>& b) { return a.first < b.first; });
    for (size_t i = 0; i < ids.size(); i++) ids[i] = sort_keys[i].second;
  }
  std::vector<std::pair<double, int>> sort_keys;

  // ---- reference-policy BVH over RNodes (Geometry.hs:369-377) ------------------------------------
  int bvh_tree(std::vector<int> obs) {
    if (obs.size() == 1) return obs[0];
    BoxD all;
    for (int o : obs) all.join(rnodes[o].ref_box);
    int d = longest_dim(all);
    sort_by_centre(obs, [&](int o) -> const BoxD& { return rnodes[o].ref_box; }, d);
    size_t n = obs.size() / 2;
    std::vector<int> left(obs.begin(), obs.begin() + n), right(obs.begin() + n, obs.end());
    int l = bvh_tree(std::move(left));
    int r = bvh_tree(std::move(right));
    int node = new_rnode(RNode::BVH);
    rnodes[node].children = {l, r};
    rnodes[node].ref_box = rnodes[l].ref_box;
    rnodes[node].ref_box.join(rnodes[r].ref_box);
    return node;
  }

  // ---- description tree walk ------------------------------------------------------------------------
  // Returns the RNode of the SURFACE geometry of this subtree (-1 when it has none, e.g. a medium) and
  // the reference's bounding box of the whole subtree in *ref_box.
  int walk(uint32_t id, const XfState& xf, uint32_t material, BoxD* ref_box) {
    if (id >= sc.nodes.size()) fail(RT_ERR_INVALID_ARG, "node id out of range");
    if (++depth_guard > 4096) fail(RT_ERR_INVALID_ARG, "description tree too deep (cycle?)");
    const HostNode& n = sc.nodes[id];
    int result = -1;
    BoxD rb;
    switch (n.kind) {
      case RT_NODE_SPHERE: {
        result = make_sphere(id, 0, mk(n.v[0]), n.s[0], xf, material);
        rb = rnodes[result].ref_box;
        break;
      }
      case RT_NODE_PARALLELOGRAM: {
        result = make_planar(id, 0, false, mk(n.v[0]), mk(n.v[1]), mk(n.v[2]), nullptr, xf, material);
        rb = rnodes[result].ref_box;
        break;
      }
      case RT_NODE_TRIANGLE: {
        D3 pc[3] = {mk(n.v[0]), mk(n.v[1]), mk(n.v[2])};
        result = make_planar(id, 0, true, pc[0], pc[1] - pc[0], pc[2] - pc[0], n.uv, xf, material, pc);
        rb = rnodes[result].ref_box;
        break;
      }
      case RT_NODE_CUBOID: {  // Geometry.hs:154-166: face order front, back, left, right, top, bottom
        double xmin = n.v[0][0], ymin = n.v[0][1], zmin = n.v[0][2], xmax = n.v[1][0], ymax = n.v[1][1], zmax = n.v[1][2];
        D3 dx{xmax - xmin, 0, 0}, dy{0, ymax - ymin, 0}, dz{0, 0, zmax - zmin};
        D3 mdx = -1.0 * dx, mdz = -1.0 * dz;
        struct F { D3 q, u, v; } faces[6] = {
            {{xmin, ymin, zmax}, dx, dy}, {{xmax, ymin, zmin}, mdx, dy}, {{xmin, ymin, zmin}, dz, dy},
            {{xmax, ymin, zmax}, mdz, dy}, {{xmin, ymax, zmax}, dx, mdz}, {{xmin, ymin, zmin}, dx, dz}};
        int g = new_rnode(RNode::GROUP);
        std::vector<int> ch;
        for (uint32_t f = 0; f < 6; f++) {
          int leaf = make_planar(id, f, false, faces[f].q, faces[f].u, faces[f].v, nullptr, xf, material);
          ch.push_back(leaf);
          rb.join(rnodes[leaf].ref_box);
        }
        // A proper solid (positive extents) whose six faces will share one leaf is traversed as ONE slab test
        // (rt_device.h, "aggregated cuboids"): same closest face, same t up to rounding.
        if (aggregate_boxes && kLeafMax >= 6 && xmax > xmin && ymax > ymin && zmax > zmin && std::isfinite(xmax - xmin) &&
            std::isfinite(ymax - ymin) && std::isfinite(zmax - zmin)) {
          HPrim& head = prims[rnodes[ch[0]].prim];
          const D3 c_local{(xmin + xmax) / 2, (ymin + ymax) / 2, (zmin + zmax) / 2};
          const D3 ax[3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
          head.box_c = xf.moving ? c_local : xf.fwd.pt(c_local, 0);
          for (int k = 0; k < 3; k++) head.box_e[k] = xf.moving ? ax[k] : xf.fwd.lin(ax[k]);
          head.box_h[0] = (xmax - xmin) / 2; head.box_h[1] = (ymax - ymin) / 2; head.box_h[2] = (zmax - zmin) / 2;
          head.flags |= PRIM_FLAG_BOXHEAD;
          for (int f = 1; f < 6; f++) prims[rnodes[ch[f]].prim].flags |= PRIM_FLAG_BOXTAIL;
        }
        rnodes[g].children = ch;
        rnodes[g].ref_box = rb;
        result = g;
        break;
      }
      case RT_NODE_MESH: {  // triangleMesh, Geometry.hs:288-294
        size_t nf = n.faces.size() / 6;
        if (nf == 0) fail(RT_ERR_EMPTY_BVH, "bvhTree: empty list");
        // The same mesh under the same transform again (`dielectric <$ mesh` + `constantMedium rho mesh`): its triangles and
        // its reference-policy tree (a sort at every level) are those of the first visit - clone them, new material only.
        const uint32_t xf_id = xform_id(xf);
        auto seen = mesh_memo.find({id, xf_id});
        if (seen != mesh_memo.end()) {
          const MeshMemo mm = seen->second;
          const uint32_t dp = (uint32_t)prims.size() - mm.p0;
          const int dr = (int)rnodes.size() - mm.r0;
          prims.reserve(prims.size() + (mm.p1 - mm.p0));
          for (uint32_t i = mm.p0; i < mm.p1; i++) { prims.push_back(prims[i]); prims.back().material = material; }
          rnodes.reserve(rnodes.size() + (size_t)(mm.r1 - mm.r0));
          for (int r = mm.r0; r < mm.r1; r++) {
            rnodes.push_back(rnodes[r]);
            RNode& c = rnodes.back();
            if (c.kind == RNode::LEAF) c.prim += dp;
            else for (int& ch : c.children) ch += dr;
          }
          result = mm.root + dr;
          rb = rnodes[result].ref_box;
          break;
        }
        const uint32_t memo_p0 = (uint32_t)prims.size();
        const int memo_r0 = (int)rnodes.size();
        static const double def_uv[3][2] = {{0, 0}, {1, 0}, {0, 1}};
        std::vector<int> tris;
        tris.reserve(nf);
        prims.reserve(prims.size() + nf);
        rnodes.reserve(rnodes.size() + 2 * nf);
        size_t nv = n.verts.size() / 3, nt = n.uvs.size() / 2;
        for (size_t f = 0; f < nf; f++) {
          const int32_t* fi = &n.faces[6 * f];
          double uv[3][2];
          D3 p[3];
          for (int k = 0; k < 3; k++) {
            int32_t vi = fi[2 * k], ti = fi[2 * k + 1];
            if (vi < 0 || (size_t)vi >= nv) fail(RT_ERR_INVALID_ARG, "mesh vertex index out of range");
            if (ti >= 0 && (size_t)ti >= nt) fail(RT_ERR_INVALID_ARG, "mesh uv index out of range");
            p[k] = mk(&n.verts[3 * vi]);
            uv[k][0] = ti >= 0 ? n.uvs[2 * ti] : def_uv[k][0];
            uv[k][1] = ti >= 0 ? n.uvs[2 * ti + 1] : def_uv[k][1];
          }
          tris.push_back(make_planar(id, (uint32_t)f, true, p[0], p[1] - p[0], p[2] - p[0], uv, xf, material, p));
        }
        const auto t_tree = std::chrono::steady_clock::now();
        result = bvh_tree(std::move(tris));
        ms_ref_tree += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_tree).count();
        rb = rnodes[result].ref_box;
        mesh_memo[{id, xf_id}] = MeshMemo{memo_p0, (uint32_t)prims.size(), memo_r0, (int)rnodes.size(), result};
        break;
      }
      case RT_NODE_GROUP:
      case RT_NODE_BVH_TREE:
      case RT_NODE_BVH_NODE: {
        if (n.kind == RT_NODE_BVH_NODE && n.children.size() != 2) fail(RT_ERR_INVALID_ARG, "bvhNode needs exactly two children");
        if (n.children.empty()) {
          if (n.kind == RT_NODE_BVH_TREE) fail(RT_ERR_EMPTY_BVH, "bvhTree: empty list");
          fail(RT_ERR_INVALID_ARG, "group: empty list (boxJoin of an empty list)");
        }
        std::vector<int> ch;
        for (uint32_t c : n.children) {
          BoxD cb;
          int r = walk(c, xf, material, &cb);
          rb.join(cb);
          if (r >= 0) { rnodes[r].ref_box = cb; ch.push_back(r); }
        }
        if (ch.empty()) { result = -1; break; }
        if (n.kind == RT_NODE_GROUP) {
          if (ch.size() == 1) { result = ch[0]; break; }
          int g = new_rnode(RNode::GROUP);
          rnodes[g].children = ch;
          result = g;
        } else if (n.kind == RT_NODE_BVH_TREE) {
          result = bvh_tree(ch);
        } else {
          if (ch.size() == 1) { result = ch[0]; break; }
          int b = new_rnode(RNode::BVH);
          rnodes[b].children = ch;
          result = b;
        }
        break;
      }
      case RT_NODE_TRANSFORM: {  // Geometry.hs:382-391
        if (n.children.size() != 1) fail(RT_ERR_INVALID_ARG, "transform needs one child");
        const double* m = n.m;
        if (std::fabs(m[12]) > 1e-12 || std::fabs(m[13]) > 1e-12 || std::fabs(m[14]) > 1e-12 || std::fabs(m[15] - 1) > 1e-12)
          fail(RT_ERR_NON_EUCLIDEAN, "transform: bottom row of the matrix must be 0 0 0 1");
        Affine f;
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) f.M[r * 3 + c] = m[r * 4 + c]; f.c0[r] = m[r * 4 + 3]; }
        for (int i = 0; i < 3; i++)
          for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += f.M[k * 3 + i] * f.M[k * 3 + j];
            if (std::fabs(s - (i == j ? 1.0 : 0.0)) > 1e-9)
              fail(RT_ERR_NON_EUCLIDEAN, "transform: matrix is not Euclidean (rotation/reflection + translation); "
                                         "use transformVertices for scaling (Geometry.hs:379-381,434)");
          }
        Affine g;  // inverse of a Euclidean map: M^T, -M^T c
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) g.M[i * 3 + j] = f.M[j * 3 + i];
        D3 t = g.lin(mk(f.c0));
        g.c0[0] = -t.x; g.c0[1] = -t.y; g.c0[2] = -t.z;
        XfState x2;
        x2.fwd = compose(xf.fwd, f);
        x2.inv = compose(g, xf.inv);
        x2.identity = false;
        x2.moving = xf.moving;
        x2.chain = xf.chain;
        x2.chain.push_back(id);
        BoxD cb;
        result = walk(n.children[0], x2, material, &cb);
        for (int i = 0; i < 8; i++) {  // boxHull of the mapped corners (Geometry.hs:386-387)
          D3 c{(i & 4) ? cb.hi[0] : cb.lo[0], (i & 2) ? cb.hi[1] : cb.lo[1], (i & 1) ? cb.hi[2] : cb.lo[2]};
          rb.add(f.pt(c, 0));
        }
        break;
      }
      case RT_NODE_MOVING: {  // Geometry.hs:449-456: shift = (1-time) v0 + time v1
        if (n.children.size() != 1) fail(RT_ERR_INVALID_ARG, "moving needs one child");
        Affine f, g;
        for (int k = 0; k < 3; k++) { f.c0[k] = n.v[0][k]; f.c1[k] = n.v[1][k] - n.v[0][k]; g.c0[k] = -f.c0[k]; g.c1[k] = -f.c1[k]; }
        XfState x2;
        x2.fwd = compose(xf.fwd, f);
        x2.inv = compose(g, xf.inv);
        x2.identity = false;
        x2.moving = true;
        x2.chain = xf.chain;
        x2.chain.push_back(id);
        BoxD cb;
        result = walk(n.children[0], x2, material, &cb);
        for (int k = 0; k < 3; k++) {
          rb.lo[k] = std::min(cb.lo[k] + n.v[0][k], cb.lo[k] + n.v[1][k]);
          rb.hi[k] = std::max(cb.hi[k] + n.v[0][k], cb.hi[k] + n.v[1][k]);
        }
        break;
      }
      case RT_NODE_MEDIUM: {  // Geometry.hs:298-330 — hoisted
        if (n.children.size() != 1) fail(RT_ERR_INVALID_ARG, "constantMedium needs one child");
        if (material == kNone) fail(RT_ERR_TYPE, "constantMedium without a material (expected `material <$ constantMedium ...`)");
        if (!(n.s[0] > 0)) fail(RT_ERR_INVALID_ARG, "constantMedium: density must be positive");
        BoxD cb;
        int b = walk(n.children[0], xf, kNone, &cb);
        if (b < 0) fail(RT_ERR_INVALID_ARG, "constantMedium: boundary has no surface");
        rnodes[b].ref_box = cb;
        media.push_back({n.s[0], b, material});
        rb = cb;
        result = -1;
        break;
      }
      case RT_NODE_WITH_MATERIAL: {  // outermost `<$` wins
        if (n.children.size() != 1) fail(RT_ERR_INVALID_ARG, "<$ needs one child");
        uint32_t m = material;
        if (m == kNone) { m = n.material; use_material(m); }
        result = walk(n.children[0], xf, m, &rb);
        break;
      }
      case RT_NODE_PURE: {
        if (n.children.size() != 1) fail(RT_ERR_INVALID_ARG, "pureGeometry needs one child");
        result = walk(n.children[0], xf, material, &rb);
        break;
      }
      case RT_NODE_USER:
        fail(RT_ERR_USER_DEFINED, "scene uses a user-defined shape (planeShape with a custom predicate or a raw "
                                  "Geometry closure); only the library's own shapes can be flattened — no CPU fallback");
      default: fail(RT_ERR_INVALID_ARG, "unknown node kind");
    }
    --depth_guard;
    *ref_box = rb;
    return result;
  }

  // ---- numbering in the reference's evaluation order ---------------------------------------------------
  uint32_t next_index = 0;
  void number(int r, bool surface) {
    RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) {
      HPrim& p = prims[n.prim];
      if (surface && p.material == kNone)
        fail(RT_ERR_TYPE, "raytrace: world is a Geometry m () — attach materials with `<$` (Ray.hs:121)");
      p.ref_index = next_index++;
      if (!sah) p.index = p.ref_index;   // reference-shaped device tree: storage order = evaluation order
      return;
    }
    for (int c : n.children) number(c, surface);
  }

  // ---- device-shaped tree --------------------------------------------------------------------------------
  // A `group` subtree made only of groups and primitives with at most kLeafMax primitives becomes ONE leaf (a linear
  // scan, exactly what the reference does for `group`): every lane of a warp then runs the same instruction stream,
  // which beats a divergent BVH walk for small lists.  RT_B200_LEAF overrides (<= 32 by the 5-bit count field).
  uint32_t kLeafMax = 24;
  // A bvhTree subtree with at most kCollapse primitives becomes one leaf (the reference's tree, truncated: same split
  // policy above, a short scan below).  RT_B200_COLLAPSE overrides.
  uint32_t kCollapse = 4;
  bool aggregate_boxes = true;      // RT_B200_NO_BOXES=1: every cuboid face is tested on its own (A/B runs)
  // number of primitives below a subtree made only of GROUPs and LEAFs; 0 if it contains a BVH node
  uint32_t flat_count(const RNode& n) const {
    if (n.kind == RNode::LEAF) return 1;
    if (n.kind != RNode::GROUP) return 0;
    uint32_t tot = 0;
    for (int c : n.children) {
      uint32_t k = flat_count(rnodes[c]);
      if (k == 0) return 0;
      tot += k;
      if (tot > kLeafCap) return kLeafCap + 1;
    }
    return tot;
  }
  void count_collect(const RNode& n, uint32_t& k, uint32_t& first, BoxD& box) const {
    if (k > kLeafCap) return;
    if (n.kind == RNode::LEAF) { k++; first = std::min(first, prims[n.prim].index); box.join(prims[n.prim].box); return; }
    for (int c : n.children) count_collect(rnodes[c], k, first, box);
  }
  void flat_collect(const RNode& n, uint32_t& first, BoxD& box) const {
    if (n.kind == RNode::LEAF) { first = std::min(first, prims[n.prim].index); box.join(prims[n.prim].box); return; }
    for (int c : n.children) flat_collect(rnodes[c], first, box);
  }
  int median_build(std::vector<int> hs) {  // the reference's split policy on device boxes
    if (hs.size() == 1) return hs[0];
    BoxD all;
    for (int h : hs) all.join(hnodes[h].box);
    int d = longest_dim(all);
    sort_by_centre(hs, [&](int h) -> const BoxD& { return hnodes[h].box; }, d);
    size_t n = hs.size() / 2;
    std::vector<int> l(hs.begin(), hs.begin() + n), r(hs.begin() + n, hs.end());
    int li = median_build(std::move(l)), ri = median_build(std::move(r));
    hnodes.emplace_back();
    int me = (int)hnodes.size() - 1;
    hnodes[me].left = li; hnodes[me].right = ri;
    hnodes[me].box = hnodes[li].box;
    hnodes[me].box.join(hnodes[ri].box);
    return me;
  }
  // ---- device tree, default policy: binned SAH over ALL primitives below a root ------------------------------------
  // The reference's tree fixes only the EVALUATION ORDER of the primitives (ref_index: who wins an exact tie); closest
  // hit over a set does not depend on how the set is searched, so the device tree is free to be a better one (SURVEY
  // §8f4).  The reference's policy — object median along the longest axis, Geometry.hs:369-377 — ages badly with mixed
  // sizes (demo1's ground sphere of radius 1000 sits deep inside a tree of 0.2-radius spheres and inflates every
  // ancestor box), and a `group` of bvhTrees is several trees searched one after the other.  Here: one tree per root,
  // top-down, 16 centroid bins x 3 axes, surface-area cost; leaves of <= kCollapse primitives (an aggregated cuboid is
  // one item); leaves store their primitives in evaluation order, storage indices follow the leaf order.
  // RT_B200_BVH=reference keeps the reference-shaped tree (A/B runs, tests).
  bool sah = true;
  double sah_node_cost = 1.0;   // RT_B200_SAH_NODE_COST overrides (sweep: profiles/r01g_sweep_sah.txt)
  uint32_t next_storage = 0;
  struct Item { uint32_t prim, n; BoxD box; double c[3]; };
  static double half_area(const BoxD& b) {
    if (b.empty()) return 0.0;
    double x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
    return x * y + y * z + z * x;
  }
  void collect_items(int r, std::vector<Item>& out) {   // DFS = evaluation order
    const RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) {
      const HPrim& p = prims[n.prim];
      if ((p.flags & PRIM_FLAG_BOXTAIL) && !out.empty() && out.back().prim + out.back().n == n.prim) {   // faces 1..5 travel with their head
        out.back().n++;
        out.back().box.join(p.box);
        return;
      }
      out.push_back(Item{n.prim, 1u, p.box, {0, 0, 0}});
      return;
    }
    for (int c : n.children) collect_items(c, out);
  }
  // Moving / instanced primitives of the surface tree kept out of it (TF_MOVLIST, rt_device.h): decided by
  // plan_moving_list() before the surface tree is built
  bool split_moving = false;
  uint32_t moving_first = 0, moving_count = 0;
  uint32_t count_moving(int r) const {
    const RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) return (prims[n.prim].flags & PRIM_FLAG_MOVING) ? 1u : 0u;
    uint32_t t = 0;
    for (int c : n.children) t += count_moving(c);
    return t;
  }
  uint32_t count_leaves(int r) const {
    const RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) return 1u;
    uint32_t t = 0;
    for (int c : n.children) t += count_leaves(c);
    return t;
  }
  void plan_moving_list(int surf, bool enabled) {
    split_moving = false;
    if (!enabled || !sah || surf < 0) return;
    const uint32_t nm = count_moving(surf), nall = count_leaves(surf);
    if (nm == 0 || nm > kMaxMovingList || nall - nm <= kLeafMax) return;      // nothing to gain for a short list of statics
    for (const MediumRec& m : media) if (count_moving(m.rtree)) return;        // a moving medium boundary: keep the fallback
    split_moving = true;
  }
  int sah_leaf(std::vector<Item>& items, size_t lo, size_t hi) {
    std::sort(items.begin() + lo, items.begin() + hi, [&](const Item& a, const Item& b) { return prims[a.prim].ref_index < prims[b.prim].ref_index; });
    hnodes.emplace_back();
    const int me = (int)hnodes.size() - 1;
    HNode& h = hnodes[me];
    h.first = next_storage;
    for (size_t i = lo; i < hi; i++) {
      // an item's primitives are consecutive entries of `prims` (make_planar appends the six faces in order)
      for (uint32_t k = 0; k < items[i].n; k++) prims[items[i].prim + k].index = next_storage++;
      h.box.join(items[i].box);
    }
    h.count = next_storage - h.first;
    return me;
  }
  int sah_node(std::vector<Item>& items, size_t lo, size_t hi, int depth) {
    uint32_t n_prims = 0;
    BoxD cb;   // centroid bounds
    for (size_t i = lo; i < hi; i++) { n_prims += items[i].n; cb.add({items[i].c[0], items[i].c[1], items[i].c[2]}); }
    if (hi - lo == 1) return sah_leaf(items, lo, hi);
    auto item_cost = [](const Item& it) { return it.n == 6 ? 2.0 : 1.0; };   // slab test ~ two planar tests
    const double kNodeCost = sah_node_cost;   // one node step (two slab tests + stack work) in units of a primitive test
    size_t mid = lo;
    bool split_found = false;
    // The traversal stack holds kStackDepth (32) entries: SAH may split as unevenly as it likes only while a balanced
    // median split of what is left would still fit (depth + ceil(log2 n) stays below 30).
    int log2n = 0;
    while (((size_t)1 << log2n) < hi - lo) log2n++;
    if (depth + 1 + log2n < 30) {
      constexpr int kBins = 16;
      BoxD all;
      double leaf_cost = 0;
      for (size_t i = lo; i < hi; i++) { all.join(items[i].box); leaf_cost += item_cost(items[i]); }
      double best = std::numeric_limits<double>::infinity();
      int best_axis = -1, best_bin = 0;
      for (int ax = 0; ax < 3; ax++) {
        const double ext = cb.hi[ax] - cb.lo[ax];
        if (!(ext > 0) || !std::isfinite(ext)) continue;
        BoxD bb[kBins];
        double bc[kBins] = {0};
        for (size_t i = lo; i < hi; i++) {
          int k = std::min(kBins - 1, (int)((items[i].c[ax] - cb.lo[ax]) / ext * kBins));
          bb[k].join(items[i].box);
          bc[k] += item_cost(items[i]);
        }
        BoxD r;
        double rc = 0, ra[kBins], rn[kBins];
        for (int k = kBins - 1; k > 0; k--) { r.join(bb[k]); rc += bc[k]; ra[k] = half_area(r); rn[k] = rc; }
        BoxD l;
        double lc = 0;
        for (int k = 0; k < kBins - 1; k++) {
          l.join(bb[k]);
          lc += bc[k];
          if (lc == 0 || rn[k + 1] == 0) continue;
          const double c = half_area(l) * lc + ra[k + 1] * rn[k + 1];
          if (c < best) { best = c; best_axis = ax; best_bin = k; }
        }
      }
      const double area = half_area(all);
      const bool must_split = n_prims > kCollapse;
      if (best_axis >= 0 && area > 0 && (must_split || kNodeCost + best / area < leaf_cost)) {
        const double ext = cb.hi[best_axis] - cb.lo[best_axis];
        auto it = std::partition(items.begin() + lo, items.begin() + hi, [&](const Item& x) {
          int k = std::min(kBins - 1, (int)((x.c[best_axis] - cb.lo[best_axis]) / ext * kBins));
          return k <= best_bin;
        });
        mid = (size_t)(it - items.begin());
        split_found = mid > lo && mid < hi;
      } else if (!must_split) {
        return sah_leaf(items, lo, hi);
      }
    }
    if (!split_found) {   // coincident centroids, or too deep for the traversal stack: balanced object-median split
      if (n_prims <= kCollapse) return sah_leaf(items, lo, hi);
      const int ax = longest_dim(cb.empty() ? BoxD() : cb);
      mid = lo + (hi - lo) / 2;
      std::nth_element(items.begin() + lo, items.begin() + mid, items.begin() + hi,
                       [&](const Item& a, const Item& b) { return a.c[ax] < b.c[ax]; });
    }
    const int l = sah_node(items, lo, mid, depth + 1);
    const int r = sah_node(items, mid, hi, depth + 1);
    hnodes.emplace_back();
    const int me = (int)hnodes.size() - 1;
    hnodes[me].left = l;
    hnodes[me].right = r;
    hnodes[me].box = hnodes[l].box;
    hnodes[me].box.join(hnodes[r].box);
    return me;
  }
  int to_device_sah(int r, bool is_surface = false) {
    std::vector<Item> items;
    collect_items(r, items);
    std::vector<Item> movers;
    if (is_surface && split_moving) {
      std::vector<Item> keep;
      for (const Item& it : items) ((prims[it.prim].flags & PRIM_FLAG_MOVING) ? movers : keep).push_back(it);
      items.swap(keep);
    }
    struct AfterTree {     // the moving list is stored right behind the tree's primitives, in evaluation order
      Builder& b; std::vector<Item>& mv;
      ~AfterTree() {
        if (mv.empty()) return;
        std::sort(mv.begin(), mv.end(), [&](const Item& x, const Item& y) { return b.prims[x.prim].ref_index < b.prims[y.prim].ref_index; });
        b.moving_first = b.next_storage;
        for (const Item& it : mv) for (uint32_t k = 0; k < it.n; k++) b.prims[it.prim + k].index = b.next_storage++;
        b.moving_count = b.next_storage - b.moving_first;
      }
    } after{*this, movers};
    uint32_t n_prims = 0;
    for (Item& it : items) {
      n_prims += it.n;
      for (int k = 0; k < 3; k++) it.c[k] = (it.box.lo[k] + it.box.hi[k]) / 2;
    }
    if (n_prims <= kLeafMax) return sah_leaf(items, 0, items.size());   // short list: one leaf, scanned (flat kernels)
    if (is_surface && defer_to_device(items)) return kDeferred;
    return sah_node(items, 0, items.size(), 0);
  }

  // ---- surface tree left to the device (RtOptions.gpu_build, rt_build.cuh) ---------------------------------------------
  // The host keeps what only it can do - the reference-policy pass that fixes the evaluation order - and hands the
  // binned-SAH build (same bins, same cost model as sah_node above) to the GPU: the items travel in evaluation order with
  // outward-rounded fp32 boxes, the primitive records are packed in that same order (provisional slots), and the first
  // upload builds the tree, reads nodes and leaf order back and permutes the records once.
  int gpu_build = -1;                          // -1: trees of >= kGpuBuildMin items; 0: never; 1: every tree with an inner node
  static constexpr int kDeferred = -2;
  static constexpr size_t kGpuBuildMin = 1024, kGpuBuildMax = 1u << 20;
  DeferredTree deferred;
  bool defer_to_device(const std::vector<Item>& items) {
    if (gpu_build == 0 || items.size() < 2 || items.size() > kGpuBuildMax) return false;
    if (gpu_build < 0 && items.size() < kGpuBuildMin) return false;
    for (const Item& it : items) {
      if (it.n != 1 && !(it.n == 6 && (prims[it.prim].flags & PRIM_FLAG_BOXHEAD))) return false;
      for (int k = 0; k < 3; k++)
        if (!std::isfinite(it.box.lo[k]) || !std::isfinite(it.box.hi[k]) || std::fabs(it.box.lo[k]) > 1e30 || std::fabs(it.box.hi[k]) > 1e30) return false;
    }
    deferred.pending = true;
    deferred.first = next_storage;
    deferred.collapse = kCollapse;
    deferred.node_cost = (float)sah_node_cost;
    deferred.items.resize(items.size());
    for (size_t i = 0; i < items.size(); i++) {
      BuildItem& bi = deferred.items[i];
      for (int k = 0; k < 3; k++) { bi.lo[k] = round_down(items[i].box.lo[k]); bi.hi[k] = round_up(items[i].box.hi[k]); }
      bi.prov = next_storage;
      bi.n = items[i].n;
      for (uint32_t k = 0; k < items[i].n; k++) prims[items[i].prim + k].index = next_storage++;
    }
    deferred.n_prims = next_storage - deferred.first;
    return true;
  }

  // storage slots for a subtree's primitives without a tree over them (evaluation order)
  void assign_storage_linear(int r) {
    std::vector<Item> items;
    collect_items(r, items);
    std::sort(items.begin(), items.end(), [&](const Item& x, const Item& y) { return prims[x.prim].ref_index < prims[y.prim].ref_index; });
    for (const Item& it : items) for (uint32_t k = 0; k < it.n; k++) prims[it.prim + k].index = next_storage++;
  }

  int to_device(int r, bool is_surface = false) {
    if (sah) return to_device_sah(r, is_surface);
    const RNode n = rnodes[r];  // copy: hnodes/rnodes may grow
    if (n.kind == RNode::LEAF) {
      hnodes.emplace_back();
      int me = (int)hnodes.size() - 1;
      hnodes[me].first = prims[n.prim].index; hnodes[me].count = 1;
      hnodes[me].box = prims[n.prim].box;
      return me;
    }
    if (n.kind == RNode::GROUP) {
      const uint32_t k = flat_count(n);
      if (k >= 1 && k <= kLeafMax) {   // e.g. a cuboid (6 parallelograms) or the whole Cornell box (18): one leaf;
        hnodes.emplace_back();         // indices are contiguous because numbering is a DFS in evaluation order
        int me = (int)hnodes.size() - 1;
        uint32_t first = kNone;
        BoxD box;
        flat_collect(n, first, box);
        hnodes[me].first = first;
        hnodes[me].count = k;
        hnodes[me].box = box;
        return me;
      }
    }
    if (n.kind == RNode::BVH && kCollapse > 1) {
      uint32_t k = 0, first = kNone;
      BoxD box;
      count_collect(n, k, first, box);
      if (k <= kCollapse) {
        hnodes.emplace_back();
        int me = (int)hnodes.size() - 1;
        hnodes[me].first = first; hnodes[me].count = k; hnodes[me].box = box;
        return me;
      }
    }
    std::vector<int> hs;
    for (int c : n.children) hs.push_back(to_device(c));
    if (n.kind == RNode::BVH) {
      hnodes.emplace_back();
      int me = (int)hnodes.size() - 1;
      hnodes[me].left = hs[0]; hnodes[me].right = hs[1];
      hnodes[me].box = hnodes[hs[0]].box;
      hnodes[me].box.join(hnodes[hs[1]].box);
      return me;
    }
    return median_build(std::move(hs));  // GROUP → BVH (results identical to the linear scan)
  }

  void collect_prims(int r, std::vector<uint32_t>& out) const {
    const RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) { out.push_back(n.prim); return; }
    for (int c : n.children) collect_prims(c, out);
  }
  // Is the boundary under RNode r a closed, outward-oriented surface?  Conservative: true only for one static sphere of
  // positive radius, or a set of static triangles under ONE transform in which every directed edge (keyed by the exact
  // vertex coordinates, so duplicated obj vertices still match) occurs once and its reverse once, with positive volume.
  bool boundary_is_closed(int r) const {
    std::vector<uint32_t> ps;
    collect_prims(r, ps);
    if (ps.empty()) return false;
    if (ps.size() == 1) {
      const HPrim& p = prims[ps[0]];
      return p.type == PRIM_SPHERE && p.r > 0 && !(p.flags & PRIM_FLAG_MOVING);
    }
    // vertices are identified by their exact coordinates, directed edges by the pair of vertex ids (hash maps: O(n))
    const uint32_t xf0 = prims[ps[0]].xform;
    std::vector<std::array<double, 3>> pts;
    pts.reserve(3 * ps.size());
    double vol6 = 0;
    for (uint32_t pi : ps) {
      const HPrim& p = prims[pi];
      if (p.type != PRIM_TRI || !p.has_corners || p.xform != xf0 || (p.flags & PRIM_FLAG_MOVING)) return false;
      for (int k = 0; k < 3; k++) {
        if (!(std::isfinite(p.corner[k].x) && std::isfinite(p.corner[k].y) && std::isfinite(p.corner[k].z))) return false;
        pts.push_back({p.corner[k].x + 0.0, p.corner[k].y + 0.0, p.corner[k].z + 0.0});   // + 0.0: -0 and +0 are one vertex
      }
      vol6 += dot(p.corner[0], cross(p.corner[1], p.corner[2]));
    }
    // vertex ids by exact coordinates (hash of the three bit patterns), directed edges as id pairs packed into 64 bits
    struct KeyHash {
      size_t operator()(const std::array<uint64_t, 3>& k) const {
        uint64_t h = k[0] * 0x9e3779b97f4a7c15ULL;
        h ^= (k[1] + 0x9e3779b97f4a7c15ULL + (h << 6) + (h >> 2)) * 0xbf58476d1ce4e5b9ULL;
        h ^= (k[2] + 0x94d049bb133111ebULL + (h << 6) + (h >> 2)) * 0xff51afd7ed558ccdULL;
        return (size_t)(h ^ (h >> 31));
      }
    };
    std::unordered_map<std::array<uint64_t, 3>, uint32_t, KeyHash> ids;
    ids.reserve(pts.size());
    std::unordered_set<uint64_t> edges;
    edges.reserve(pts.size() * 2);
    for (size_t t = 0; t < pts.size(); t += 3) {
      uint64_t id[3];
      for (int k = 0; k < 3; k++) {
        std::array<uint64_t, 3> key;
        std::memcpy(key.data(), pts[t + k].data(), 24);
        id[k] = ids.emplace(key, (uint32_t)ids.size()).first->second;
      }
      if (id[0] == id[1] || id[1] == id[2] || id[0] == id[2]) return false;   // degenerate triangle
      for (int k = 0; k < 3; k++)
        if (!edges.insert(id[k] << 32 | id[(k + 1) % 3]).second) return false;     // a directed edge used twice
    }
    for (uint64_t e : edges)                                                        // ... and its reverse exactly once
      if (!edges.count((e << 32) | (e >> 32))) return false;
    // orientation in WORLD space: a reflecting instance transform (det < 0) turns the surface inside out
    const double* M = xforms[xf0].fwd.M;
    const double det = M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
    return std::isfinite(vol6) && std::isfinite(det) && vol6 * det > 0;
  }

  void collect_geoms(int r, std::vector<uint32_t>& out) const {
    const RNode& n = rnodes[r];
    if (n.kind == RNode::LEAF) { out.push_back(prims[n.prim].geom); return; }
    for (int c : n.children) collect_geoms(c, out);
  }

  static float round_down(double x) { float f = (float)x; return (double)f > x ? std::nextafterf(f, -INFINITY) : f; }
  static float round_up(double x) { float f = (float)x; return (double)f < x ? std::nextafterf(f, INFINITY) : f; }
  void write_node(DNode& d, const HNode& h) {
    double m = 0;
    for (int k = 0; k < 3; k++) m = std::max(m, std::max(std::fabs(h.box.lo[k]), std::fabs(h.box.hi[k])));
    double pad = 1e-5 * m + 1e-6;  // slab test runs in fp32: keep boxes conservative
    for (int k = 0; k < 3; k++) { d.lo[k] = round_down(h.box.lo[k] - pad); d.hi[k] = round_up(h.box.hi[k] + pad); }
  }
  // Node reference: leaf → kLeafBit | first << 5 | (count - 1); inner → index of the node PAIR holding its two
  // children (siblings adjacent: one 64-byte fetch tests both).  A root has no node of its own: like the
  // reference, the top-level geometry's own box is never tested.
  uint32_t emit_ref(FlatScene& fs, int hn, uint32_t depth, uint32_t& max_depth) {
    const HNode h = hnodes[hn];
    max_depth = std::max(max_depth, depth);
    if (h.left < 0) return kLeafBit | (h.first << kLeafShift) | (h.count - 1);
    uint32_t at = (uint32_t)fs.nodes.size();
    fs.nodes.emplace_back();
    fs.nodes.emplace_back();
    int ids[2] = {h.left, h.right};
    for (int k = 0; k < 2; k++) {
      DNode d;
      write_node(d, hnodes[ids[k]]);
      d.a = emit_ref(fs, ids[k], depth + 1, max_depth);
      d.b = hnodes[ids[k]].left < 0 ? hnodes[ids[k]].count : 0;
      fs.nodes[at + k] = d;
    }
    return at;
  }

  void pack_prims(FlatScene& fs) {
    size_t n = prims.size();
    fs.prims.assign(n * 16, 0.f);
    fs.tri_uv.assign(n * 8, 0.f);
    fs.canonical.assign(n, kNone);
    fs.ref_order.assign(n, kNone);
    for (const HPrim& p : prims) {
      if (p.index == kNone) continue;
      float* o = &fs.prims[(size_t)p.index * 16];
      if (p.type == PRIM_SPHERE) {
        o[0] = (float)p.c.x; o[1] = (float)p.c.y; o[2] = (float)p.c.z; o[3] = (float)p.r;
        // fp64 copy for the large-radius path: hi parts in o[0..3], lo parts in o[4..7]
        o[4] = (float)(p.c.x - (double)o[0]); o[5] = (float)(p.c.y - (double)o[1]); o[6] = (float)(p.c.z - (double)o[2]);
        o[7] = (float)(p.r - (double)o[3]);
        fs.n_spheres++;
      } else {
        o[0] = (float)p.c.x; o[1] = (float)p.c.y; o[2] = (float)p.c.z;
        o[3] = (float)p.n.x; o[4] = (float)p.n.y; o[5] = (float)p.n.z;
        o[6] = (float)p.wa.x; o[7] = (float)p.wa.y; o[8] = (float)p.wa.z;
        o[9] = (float)p.wb.x; o[10] = (float)p.wb.y; o[11] = (float)p.wb.z;
        fs.n_planars++;
      }
      uint32_t meta[4] = {p.type | p.flags, p.geom, p.xform, p.material};
      std::memcpy(o + 12, meta, 16);
      float* u = &fs.tri_uv[(size_t)p.index * 8];
      if (p.flags & PRIM_FLAG_BOXHEAD) {   // box record over the aux slots of faces 0..1 (parallelograms have no uv data)
        const float rec[16] = {(float)p.box_c.x, (float)p.box_c.y, (float)p.box_c.z, (float)p.box_h[0],
                               (float)p.box_e[0].x, (float)p.box_e[0].y, (float)p.box_e[0].z, (float)p.box_h[1],
                               (float)p.box_e[1].x, (float)p.box_e[1].y, (float)p.box_e[1].z, (float)p.box_h[2],
                               (float)p.box_e[2].x, (float)p.box_e[2].y, (float)p.box_e[2].z, 0.f};
        std::memcpy(u, rec, sizeof rec);
      } else if (p.type == PRIM_TRI) {
        for (int k = 0; k < 3; k++) { u[2 * k] = (float)p.uv[k][0]; u[2 * k + 1] = (float)p.uv[k][1]; }
      }
      fs.canonical[p.index] = p.canonical;
      fs.ref_order[p.index] = p.ref_index;
    }
  }
};

void fill_d3(float* o, const double* v) { o[0] = (float)v[0]; o[1] = (float)v[1]; o[2] = (float)v[2]; }

}  // namespace

uint64_t FlatScene::bytes() const {
  return nodes.size() * sizeof(DNode) + prims.size() * 4 + tri_uv.size() * 4 + canonical.size() * 4 +
         xforms.size() * sizeof(DXform) + materials.size() * sizeof(DMaterial) + textures.size() * sizeof(DTexture) +
         texels.size() * 4 + noise_grad.size() * 4 + noise_perm.size();
}

// ---- default Perlin tables (Noise.hs:52-92) ---------------------------------------------------------------
// The three permutation tables are the reference's literals.  The gradient table is what
// `evalState (replicateM 256 randomUnitVector) (mkStdGen 666)` yields under a restatement of
// SplitMix (the `random` package is not part of the reference tree: PARITY UNPINNED — a Haskell
// front end passes its own table through rt_scene_set_noise_tables).
namespace {
const uint8_t kPX[256] = {
    179, 60, 35, 16, 220, 94, 67, 236, 106, 112, 65, 166, 83, 101, 246, 140, 219, 186, 113, 88, 153, 70, 34, 63, 157, 210, 212, 188, 54, 74, 23, 161,
    28, 137, 126, 107, 183, 58, 134, 127, 211, 225, 17, 123, 150, 243, 160, 68, 75, 239, 173, 221, 89, 109, 61, 72, 159, 80, 154, 18, 214, 144, 197, 24,
    105, 32, 84, 226, 136, 29, 139, 97, 230, 167, 165, 238, 27, 14, 50, 193, 46, 253, 240, 111, 69, 196, 130, 102, 104, 118, 204, 12, 169, 202, 142, 25,
    245, 215, 149, 138, 185, 48, 223, 247, 47, 98, 143, 26, 87, 251, 103, 52, 234, 232, 218, 205, 92, 228, 162, 85, 122, 191, 242, 164, 129, 192, 255, 231,
    147, 91, 178, 213, 176, 36, 120, 155, 241, 222, 177, 20, 152, 141, 51, 171, 250, 95, 71, 119, 254, 172, 53, 146, 135, 124, 125, 163, 235, 99, 7, 22,
    100, 229, 93, 174, 3, 189, 116, 66, 217, 158, 237, 55, 151, 0, 148, 45, 86, 64, 216, 43, 252, 121, 200, 115, 39, 184, 82, 56, 9, 181, 62, 2,
    81, 209, 44, 79, 19, 110, 41, 10, 194, 15, 132, 224, 249, 96, 233, 117, 49, 203, 5, 37, 11, 59, 168, 114, 90, 131, 31, 145, 40, 206, 13, 187,
    133, 207, 4, 199, 170, 78, 30, 182, 248, 21, 6, 227, 57, 180, 73, 42, 128, 175, 108, 33, 244, 201, 198, 77, 195, 8, 38, 190, 76, 156, 208, 1};
const uint8_t kPY[256] = {
    252, 123, 131, 151, 243, 143, 12, 247, 196, 179, 99, 0, 178, 109, 71, 160, 93, 205, 127, 38, 142, 117, 152, 124, 166, 95, 200, 121, 15, 17, 10, 190,
    116, 158, 173, 75, 248, 191, 197, 58, 70, 184, 226, 146, 6, 239, 165, 113, 218, 34, 83, 77, 74, 5, 176, 85, 112, 147, 59, 66, 14, 31, 2, 21,
    198, 108, 255, 11, 36, 156, 96, 73, 1, 189, 126, 50, 220, 16, 249, 23, 139, 135, 141, 92, 159, 4, 119, 174, 171, 253, 86, 227, 251, 172, 7, 149,
    207, 212, 224, 44, 187, 91, 175, 84, 28, 211, 180, 195, 52, 98, 244, 125, 138, 54, 210, 201, 209, 219, 133, 240, 9, 202, 199, 42, 51, 48, 104, 154,
    88, 53, 64, 105, 242, 63, 223, 222, 67, 238, 32, 134, 72, 62, 101, 150, 94, 161, 19, 236, 215, 97, 206, 22, 188, 230, 170, 18, 13, 162, 129, 90,
    246, 130, 35, 46, 43, 213, 29, 76, 241, 61, 30, 136, 235, 87, 114, 68, 183, 177, 250, 132, 3, 122, 110, 145, 81, 78, 232, 69, 60, 80, 216, 65,
    164, 237, 157, 203, 25, 221, 254, 181, 8, 182, 214, 24, 40, 102, 37, 106, 228, 39, 229, 163, 111, 186, 245, 55, 41, 225, 234, 49, 45, 231, 107, 103,
    168, 153, 47, 56, 118, 120, 137, 155, 148, 82, 33, 204, 79, 169, 144, 20, 27, 128, 192, 217, 100, 208, 115, 140, 233, 89, 193, 185, 167, 57, 26, 194};
const uint8_t kPZ[256] = {
    153, 90, 163, 138, 20, 136, 79, 100, 93, 38, 185, 31, 193, 43, 161, 2, 30, 37, 87, 6, 127, 207, 96, 51, 27, 227, 203, 215, 155, 190, 106, 94,
    65, 183, 114, 71, 74, 219, 245, 39, 140, 216, 195, 191, 3, 214, 13, 23, 168, 179, 218, 133, 102, 53, 194, 124, 166, 108, 116, 246, 35, 109, 220, 121,
    205, 110, 83, 242, 252, 231, 25, 128, 61, 57, 187, 222, 189, 228, 148, 101, 239, 162, 48, 150, 55, 174, 178, 42, 200, 160, 58, 206, 11, 92, 204, 67,
    8, 113, 34, 172, 181, 177, 66, 16, 243, 159, 197, 135, 249, 241, 28, 147, 158, 139, 176, 88, 29, 149, 254, 226, 192, 201, 202, 186, 69, 45, 72, 199,
    107, 99, 209, 10, 230, 217, 247, 89, 50, 129, 15, 64, 248, 167, 180, 146, 210, 169, 4, 81, 97, 238, 5, 237, 125, 236, 95, 9, 104, 221, 32, 86,
    165, 134, 224, 182, 198, 234, 253, 59, 164, 52, 60, 208, 103, 22, 46, 188, 54, 12, 1, 112, 144, 137, 47, 156, 98, 78, 18, 82, 175, 225, 19, 184,
    244, 49, 120, 157, 130, 70, 80, 251, 212, 250, 119, 118, 196, 223, 105, 68, 84, 21, 145, 173, 56, 235, 91, 17, 126, 76, 131, 117, 152, 111, 33, 36,
    233, 141, 122, 85, 7, 123, 44, 62, 115, 26, 0, 40, 229, 211, 213, 63, 143, 77, 14, 24, 142, 154, 73, 132, 171, 41, 170, 240, 75, 151, 232, 255};

// SplitMix64 as a small generator object
class SplitMix {
 public:
  explicit SplitMix(uint64_t n) : state_(finalize(n)), gamma_(odd_gamma(n + kGolden)) {}
  uint64_t word() { state_ += gamma_; return finalize(state_); }
  double unit() { return (double)(word() >> 11) * 0x1.0p-53; }
  double range(double lo, double hi) { return lo + (hi - lo) * unit(); }
 private:
  static constexpr uint64_t kGolden = 0x9e3779b97f4a7c15ULL;
  static uint64_t finalize(uint64_t z) {
    z ^= z >> 33; z *= 0xff51afd7ed558ccdULL;
    z ^= z >> 33; z *= 0xc4ceb9fe1a85ec53ULL;
    return z ^ (z >> 33);
  }
  static uint64_t odd_gamma(uint64_t z) {
    z ^= z >> 30; z *= 0xbf58476d1ce4e5b9ULL;
    z ^= z >> 27; z *= 0x94d049bb133111ebULL;
    z = (z ^ (z >> 31)) | 1ULL;
    return __builtin_popcountll(z ^ (z >> 1)) < 24 ? z ^ 0xaaaaaaaaaaaaaaaaULL : z;
  }
  uint64_t state_, gamma_;
};
}  // namespace

void default_noise_tables(int32_t perm[3][256], double grad[256][3]) {
  for (int i = 0; i < 256; i++) { perm[0][i] = kPX[i]; perm[1][i] = kPY[i]; perm[2][i] = kPZ[i]; }
  SplitMix g(666);
  for (int i = 0; i < 256; i++) {
    for (;;) {  // randomUnitVector, Core.hs:54-60
      double x = g.range(-1, 1), y = g.range(-1, 1), z = g.range(-1, 1);
      double q = x * x + y * y + z * z;
      if (1e-8 <= q && q <= 1) {
        double s = std::sqrt(q);
        grad[i][0] = x / s; grad[i][1] = y / s; grad[i][2] = z / s;
        break;
      }
    }
  }
}

int32_t flatten_scene(Scene& sc, uint32_t root) {
  try {
    Builder b(sc);
    // RT_B200_TRACE_COMMIT=1: phase times of the host scene pipeline on stderr (tuning aid)
    const RtOptions& O = sc.opts;
    const bool trace = O.trace_commit > 0;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
      if (!trace) return;
      auto now = std::chrono::steady_clock::now();
      std::fprintf(stderr, "[rt_b200 commit] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
      t_last = now;
    };
    if (O.leaf_max >= 0) b.kLeafMax = std::min<uint32_t>(kLeafCap, std::max<uint32_t>(1, (uint32_t)O.leaf_max));
    if (O.bvh_reference >= 0) b.sah = O.bvh_reference == 0;
    if (O.sah_node_cost >= 0) b.sah_node_cost = O.sah_node_cost;
    if (O.aggregate_boxes >= 0) b.aggregate_boxes = O.aggregate_boxes != 0;
    if (O.collapse >= 0) b.kCollapse = std::min<uint32_t>(kLeafCap, std::max<uint32_t>(1, (uint32_t)O.collapse));
    b.mat_used.assign(sc.materials.size(), 0);
    b.tex_used.assign(sc.textures.size(), 0);
    XfState id;
    BoxD rb;
    int surf = b.walk(root, id, kNone, &rb);
    lap("walk (description tree)");
    if (trace) std::fprintf(stderr, "[rt_b200 commit]   of which mesh reference trees %8.3f ms\n", b.ms_ref_tree);
    if (surf >= 0) b.number(surf, true);
    for (auto& m : b.media) b.number(m.rtree, false);
    lap("number");
    if (b.media.size() > (size_t)kMaxMedia) b.fail(RT_ERR_UNSUPPORTED, "more than 16 constantMedium volumes in one scene");

    FlatScene fs;
    uint32_t max_depth = 0;
    // measured (profiles/r02c_sweep.txt, demo2): scanning the moving list once per segment costs more than the per-primitive
    // transform branch it removes (45.2 vs 43.8 ms per 100 spp) - opt-in only
    b.plan_moving_list(surf, O.split_moving > 0);
    if (O.gpu_build >= 0) b.gpu_build = O.gpu_build;
    bool deferred = false;
    if (surf >= 0) {
      int h = b.to_device(surf, true);
      deferred = h == Builder::kDeferred;
      if (!deferred) fs.surface_root = b.emit_ref(fs, h, 0, max_depth);
    }
    std::vector<uint32_t> share_root;     // media that walk the surface tree (their root is known once the surface root is)
    fs.moving_first = b.moving_first;
    fs.moving_count = b.moving_count;
    lap("surface tree to device");
    std::vector<uint32_t> surf_geoms;
    if (surf >= 0) b.collect_geoms(surf, surf_geoms);
    std::sort(surf_geoms.begin(), surf_geoms.end());
    surf_geoms.erase(std::unique(surf_geoms.begin(), surf_geoms.end()), surf_geoms.end());
    for (size_t i = 0; i < b.media.size(); i++) {
      DMedium& dm = fs.media[i];
      std::vector<uint32_t> g;
      b.collect_geoms(b.media[i].rtree, g);
      std::sort(g.begin(), g.end());
      g.erase(std::unique(g.begin(), g.end()), g.end());
      // The boundary IS the surface geometry (`dielectric <$ mesh` + `constantMedium rho mesh`: pawn, bunny): the same
      // primitives under the same transforms, so the boundary queries walk the SURFACE tree - no second SAH build over the
      // same triangles (bunny commit 35 -> 20 ms) and one tree in the GPU's caches instead of two.  A boundary answer is only
      // read for t, the front flag and the geometry id, all of which the surface primitive carries identically.
      const bool same_tree = surf >= 0 && O.shared_boundary != 0 && !b.split_moving && !g.empty() && g == surf_geoms &&
                             b.count_leaves(b.media[i].rtree) == b.count_leaves(surf);
      if (same_tree) {
        if (b.sah) b.assign_storage_linear(b.media[i].rtree);   // the boundary's own records keep their slots (host bookkeeping)
        dm.root = fs.surface_root;
        share_root.push_back((uint32_t)i);
        fs.n_shared_trees++;
      } else {
        int h = b.to_device(b.media[i].rtree);
        dm.root = b.emit_ref(fs, h, 0, max_depth);
      }
      dm.neg_inv_density = (float)(-(1.0 / b.media[i].density));
      dm.material = b.media[i].material;
      dm.geom_lo = dm.geom_hi = dm.shared = dm.closed = 0;
      // shared-boundary shortcut: dense geometry-id range, every id also present in the surface tree
      if (!g.empty() && g.back() - g.front() + 1 == g.size() && O.shared_boundary != 0 &&
          std::includes(surf_geoms.begin(), surf_geoms.end(), g.begin(), g.end())) {
        dm.geom_lo = g.front(); dm.geom_hi = g.back(); dm.shared = 1;
        if (O.interior_bound != 0 && b.boundary_is_closed(b.media[i].rtree)) dm.closed = 1;
      }
    }
    if (deferred) {   // the device writes its node pairs BEHIND the host-built ones (the media trees above): the array only grows
      b.deferred.node_base = (uint32_t)fs.nodes.size();
      fs.surface_root = b.deferred.node_base;
      for (uint32_t i : share_root) fs.media[i].root = fs.surface_root;
      fs.deferred = std::move(b.deferred);
    }
    fs.n_media = (uint32_t)b.media.size();
    fs.max_depth = max_depth;
    {   // world box: every primitive instance (surface or medium boundary), padded like the node boxes
      BoxD wb;
      for (const HPrim& p : b.prims) wb.join(p.box);
      bool finite = !b.prims.empty();
      for (int k = 0; k < 3; k++) finite = finite && std::isfinite(wb.lo[k]) && std::isfinite(wb.hi[k]) && wb.lo[k] <= wb.hi[k];
      if (finite && O.world_box != 0) {
        double mx = 0;
        for (int k = 0; k < 3; k++) mx = std::max(mx, std::max(std::fabs(wb.lo[k]), std::fabs(wb.hi[k])));
        const double pad = 1e-5 * mx + 1e-6;
        for (int k = 0; k < 3; k++) { fs.world_lo[k] = Builder::round_down(wb.lo[k] - pad); fs.world_hi[k] = Builder::round_up(wb.hi[k] + pad); }
        fs.has_world_box = 1;
      }
    }
    if (max_depth + 1 > (uint32_t)kStackDepth)
      b.fail(RT_ERR_UNSUPPORTED, "flattened BVH is deeper than the shared-memory traversal stack (32 levels)");
    lap("media trees to device");
    b.pack_prims(fs);
    lap("pack primitives");
    if (fs.canonical.size() >= (1u << 26)) b.fail(RT_ERR_UNSUPPORTED, "more than 2^26 primitives");
    for (const HPrim& p : b.prims) {
      if (p.index == kNone) continue;
      fs.features |= (p.type == PRIM_SPHERE) ? TF_SPHERE : TF_PLANAR;
      if (p.flags & PRIM_FLAG_MOVING) fs.features |= b.split_moving ? TF_MOVLIST : TF_MOVING;
      if (p.flags & PRIM_FLAG_BOXHEAD) fs.features |= TF_BOX;
    }
    if (!b.media.empty()) fs.features |= TF_MEDIA;

    fs.xforms.resize(b.xforms.size());
    for (size_t i = 0; i < b.xforms.size(); i++) {
      DXform& d = fs.xforms[i];
      const XfState& x = b.xforms[i];
      for (int k = 0; k < 9; k++) { d.Ainv[k] = (float)x.inv.M[k]; d.A[k] = (float)x.fwd.M[k]; }
      fill_d3(d.binv0, x.inv.c0);
      fill_d3(d.binv1, x.inv.c1);
    }
    fs.materials.resize(sc.materials.size());
    for (size_t i = 0; i < sc.materials.size(); i++) {
      fs.materials[i].kind = sc.materials[i].kind;
      fs.materials[i].texture = sc.materials[i].texture;
      fs.materials[i].param = (float)sc.materials[i].param;
      fs.materials[i]._pad = 0;
    }
    fs.textures.resize(sc.textures.size());
    for (size_t i = 0; i < sc.textures.size(); i++) {
      const HostTexture& t = sc.textures[i];
      DTexture& d = fs.textures[i];
      std::memset(&d, 0, sizeof d);
      d.kind = t.d.kind; d.i0 = t.d.i0; d.i1 = t.d.i1; d.freq = (float)t.d.freq;
      fill_d3(d.c0, t.d.c0); fill_d3(d.c1, t.d.c1); fill_d3(d.shift, t.d.shift); fill_d3(d.dir, t.d.dir);
      if (!b.tex_used[i]) continue;
      if (t.d.kind == RT_TEX_IMAGE) {
        if (t.d.width <= 0 || t.d.height <= 0 || t.image.empty()) b.fail(RT_ERR_INVALID_ARG, "imageTexture: empty image");
        d.w = t.d.width; d.h = t.d.height;
        d.image_offset = (uint32_t)(fs.texels.size() / 4);
        fs.texels.insert(fs.texels.end(), t.image.begin(), t.image.end());      // already float4 texels
      }
      if (t.d.kind == RT_TEX_NOISE || t.d.kind == RT_TEX_MARBLE) fs.has_noise = 1;
    }
    fs.noise_grad.resize(256 * 3);
    fs.noise_perm.resize(3 * 256);
    for (int i = 0; i < 256; i++) {
      for (int k = 0; k < 3; k++) {
        fs.noise_grad[3 * i + k] = (float)sc.grad[i][k];
        fs.noise_perm[k * 256 + i] = (uint8_t)(sc.perm[k][i] & 255);
      }
    }
    lap("materials, textures, noise");
    sc.flat = std::move(fs);
    sc.committed = true;
    return RT_OK;
  } catch (const FlattenError& e) {
    set_error(e.msg);
    return e.status;
  } catch (const std::exception& e) {
    set_error(std::string("internal error: ") + e.what());
    return RT_ERR_INVALID_ARG;
  }
}

}  // namespace rtb
