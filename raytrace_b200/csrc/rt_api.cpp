// rt_api.cpp — host-only half of the C ABI (include/rt_b200.h): scene description, commit,
// inspection, .obj loader.  The CUDA half (upload / render / probe / tonemap) is rt_kernels.cu.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>

#include "rt_host.h"

namespace rtb {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

void options_init(RtOptions& o) {
  std::memset(&o, 0xFF, sizeof o);   // every integer field = -1
  o.sah_node_cost = -1.0;
}
Scene::Scene() { default_noise_tables(perm, grad); options_init(opts); }
Scene::~Scene() {
  for (auto& kv : devices) destroy_device_scene(kv.second);
}

// ---- .obj loader: the rules of parseObj (Geometry.hs:207-285) ------------------------------------
struct Mesh {
  std::vector<double> verts, uvs;
  std::vector<int32_t> faces;
};

namespace {
struct ObjError {
  std::string msg;
};
inline bool hs_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v'; }
std::vector<std::string> words(const std::string& s) {  // Haskell `words`
  std::vector<std::string> out;
  size_t i = 0;
  while (i < s.size()) {
    while (i < s.size() && hs_space(s[i])) i++;
    size_t j = i;
    while (j < s.size() && !hs_space(s[j])) j++;
    if (j > i) out.push_back(s.substr(i, j - i));
    i = j;
  }
  return out;
}
// `readMaybe :: String -> Maybe Double`: [-] digits [. digits] [e [+-] digits]
bool read_double(const std::string& t, double* out) {
  size_t i = 0, n = t.size();
  if (i < n && t[i] == '-') i++;
  size_t d0 = i;
  while (i < n && t[i] >= '0' && t[i] <= '9') i++;
  if (i == d0) return false;
  if (i < n && t[i] == '.') {
    i++;
    size_t f0 = i;
    while (i < n && t[i] >= '0' && t[i] <= '9') i++;
    if (i == f0) return false;
  }
  if (i < n && (t[i] == 'e' || t[i] == 'E')) {
    i++;
    if (i < n && (t[i] == '+' || t[i] == '-')) i++;
    size_t e0 = i;
    while (i < n && t[i] >= '0' && t[i] <= '9') i++;
    if (i == e0) return false;
  }
  if (i != n) return false;
  *out = std::strtod(t.c_str(), nullptr);
  return true;
}
std::string with_line(int k, const std::string& err) { return "line " + std::to_string(k) + ": " + err; }

bool extract_int(const std::string& s, size_t& pos, long long* out) {  // extractInt / extractNat
  bool neg = false;
  if (pos < s.size() && s[pos] == '-') { neg = true; pos++; }
  size_t d0 = pos;
  long long v = 0;
  while (pos < s.size() && s[pos] >= '0' && s[pos] <= '9') { v = v * 10 + (s[pos] - '0'); pos++; }
  if (pos == d0) return false;
  *out = neg ? -v : v;
  return true;
}
long long process_ix(long long len, long long i) {  // processIx
  if (1 <= i && i <= len) return i - 1;
  if (-len <= i && i <= -1) return i + len;
  throw ObjError{"index out of bounds: " + std::to_string(i)};
}
void get_indices(long long nv, long long nt, const std::string& tok, int32_t* vi, int32_t* ti) {
  size_t pos = 0;
  long long i;
  if (!extract_int(tok, pos, &i)) throw ObjError{"expected number"};
  *vi = (int32_t)process_ix(nv, i);
  std::string rest = tok.substr(pos);
  if (rest.empty()) { *ti = -1; return; }
  if (rest.size() >= 2 && rest[0] == '/' && rest[1] == '/') { *ti = -1; return; }
  if (rest[0] == '/') {
    size_t p2 = 1;
    long long j;
    if (!extract_int(rest, p2, &j)) throw ObjError{"expected number"};
    *ti = (int32_t)process_ix(nt, j);
    return;
  }
  throw ObjError{std::string("unexpected character '") + rest[0] + "'"};
}
}  // namespace

static int32_t parse_obj(const char* text, size_t len, Mesh& m) {
  struct Line { int k; std::string rest; };
  std::vector<Line> vl, vtl, fl;
  size_t pos = 0;
  int k = 0;
  while (pos < len) {  // `lines`, then `takeWhile (/= '#')`
    size_t e = pos;
    while (e < len && text[e] != '\n') e++;
    k++;
    std::string line(text + pos, e - pos);
    size_t h = line.find('#');
    if (h != std::string::npos) line.resize(h);
    if (line.size() >= 2 && line[0] == 'v' && line[1] == ' ') vl.push_back({k, line.substr(2)});
    else if (line.size() >= 3 && line[0] == 'v' && line[1] == 't' && line[2] == ' ') vtl.push_back({k, line.substr(3)});
    else if (line.size() >= 2 && line[0] == 'f' && line[1] == ' ') fl.push_back({k, line.substr(2)});
    pos = e + 1;
  }
  for (auto& l : vl) {
    auto w = words(l.rest);
    double x, y, z;
    if (w.size() < 3 || !read_double(w[0], &x) || !read_double(w[1], &y) || !read_double(w[2], &z)) {
      set_error(with_line(l.k, "invalid 'v' statement"));
      return RT_ERR_PARSE;
    }
    m.verts.insert(m.verts.end(), {x, y, z});
  }
  for (auto& l : vtl) {
    auto w = words(l.rest);
    double u, v = 0;
    bool ok = (w.size() == 1 && read_double(w[0], &u)) ||
              (w.size() >= 2 && read_double(w[0], &u) && read_double(w[1], &v));
    if (!ok) {
      set_error(with_line(l.k, "invalid 'vt' statement"));
      return RT_ERR_PARSE;
    }
    if (w.size() == 1) v = 0;
    m.uvs.insert(m.uvs.end(), {u, v});
  }
  long long nv = (long long)m.verts.size() / 3, nt = (long long)m.uvs.size() / 2;
  for (auto& l : fl) {
    auto w = words(l.rest);
    std::vector<int32_t> vi(w.size()), ti(w.size());
    try {
      for (size_t i = 0; i < w.size(); i++) get_indices(nv, nt, w[i], &vi[i], &ti[i]);
    } catch (const ObjError& e) {
      set_error(with_line(l.k, e.msg));
      return RT_ERR_PARSE;
    }
    if (w.size() < 3) {
      set_error(with_line(l.k, "invalid 'f' statement (fewer than 3 vertices)"));
      return RT_ERR_PARSE;
    }
    for (size_t i = 1; i + 1 < w.size(); i++)  // fan triangulation (Geometry.hs:248,251-255)
      m.faces.insert(m.faces.end(), {vi[0], ti[0], vi[i], ti[i], vi[i + 1], ti[i + 1]});
  }
  return RT_OK;
}

}  // namespace rtb

using namespace rtb;

extern "C" {

uint32_t rt_version(void) { return RT_B200_VERSION; }
const char* rt_last_error(void) { return g_error.c_str(); }

int32_t rt_scene_create(RtScene** out) {
  if (!out) { set_error("rt_scene_create: null out pointer"); return RT_ERR_INVALID_ARG; }
  *out = reinterpret_cast<RtScene*>(new (std::nothrow) Scene());
  if (!*out) { set_error("out of memory"); return RT_ERR_INVALID_ARG; }
  return RT_OK;
}
int32_t rt_scene_destroy(RtScene* s) {
  delete reinterpret_cast<Scene*>(s);
  return RT_OK;
}

int32_t rt_scene_add_texture(RtScene* sp, const RtTextureDesc* d, uint32_t* id) {
  if (!sp || !d || !id) { set_error("rt_scene_add_texture: null argument"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  if (d->kind < RT_TEX_CONSTANT || d->kind > RT_TEX_USER) { set_error("unknown texture kind"); return RT_ERR_INVALID_ARG; }
  HostTexture t;
  t.d = *d;
  t.d.image = nullptr;
  if (d->kind == RT_TEX_IMAGE) {
    if (!d->image || d->width <= 0 || d->height <= 0) { set_error("imageTexture: null or empty image"); return RT_ERR_INVALID_ARG; }
    // stored as float4 texels right away (the device format): ONE pass over the image, commit only memcpy's it
    const size_t n = (size_t)d->width * d->height;
    t.image.resize(n * 4);
    const float* src = d->image;
    float* dst = t.image.data();
    for (size_t k = 0; k < n; k++, dst += 4, src += 3) { dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = 0.f; }
  }
  *id = (uint32_t)s.textures.size();
  s.textures.push_back(std::move(t));
  s.committed = false;
  return RT_OK;
}

int32_t rt_scene_add_material(RtScene* sp, const RtMaterialDesc* d, uint32_t* id) {
  if (!sp || !d || !id) { set_error("rt_scene_add_material: null argument"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  if (d->kind < RT_MAT_LIGHT || d->kind > RT_MAT_USER) { set_error("unknown material kind"); return RT_ERR_INVALID_ARG; }
  *id = (uint32_t)s.materials.size();
  s.materials.push_back(*d);
  s.committed = false;
  return RT_OK;
}

int32_t rt_scene_add_node(RtScene* sp, const RtNodeDesc* d, uint32_t* id) {
  if (!sp || !d || !id) { set_error("rt_scene_add_node: null argument"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  if (d->kind < RT_NODE_SPHERE || d->kind > RT_NODE_USER) { set_error("unknown node kind"); return RT_ERR_INVALID_ARG; }
  HostNode n;
  n.kind = d->kind;
  n.material = d->material;
  if (d->n_children && !d->children) { set_error("rt_scene_add_node: null children"); return RT_ERR_INVALID_ARG; }
  for (uint32_t i = 0; i < d->n_children; i++) {
    if (d->children[i] >= s.nodes.size()) { set_error("rt_scene_add_node: child id out of range (children must be added first)"); return RT_ERR_INVALID_ARG; }
    n.children.push_back(d->children[i]);
  }
  std::memcpy(n.v, d->v, sizeof n.v);
  std::memcpy(n.s, d->s, sizeof n.s);
  std::memcpy(n.uv, d->uv, sizeof n.uv);
  std::memcpy(n.m, d->m, sizeof n.m);
  if (d->kind == RT_NODE_MESH) {
    if ((d->n_verts && !d->verts) || (d->n_faces && !d->faces) || (d->n_uvs && !d->uvs)) { set_error("mesh: null array"); return RT_ERR_INVALID_ARG; }
    n.verts.assign(d->verts, d->verts + (size_t)d->n_verts * 3);
    if (d->n_uvs) n.uvs.assign(d->uvs, d->uvs + (size_t)d->n_uvs * 2);
    n.faces.assign(d->faces, d->faces + (size_t)d->n_faces * 6);
  }
  if (d->kind == RT_NODE_WITH_MATERIAL && d->material >= s.materials.size()) { set_error("<$: material id out of range"); return RT_ERR_INVALID_ARG; }
  *id = (uint32_t)s.nodes.size();
  s.nodes.push_back(std::move(n));
  s.committed = false;
  return RT_OK;
}

int32_t rt_scene_add_batch(RtScene* sp, const RtTextureDesc* tex, uint32_t n_tex, const RtMaterialDesc* mat, uint32_t n_mat,
                           const RtNodeDesc* nodes, uint32_t n_nodes, uint32_t* first_ids) {
  if (!sp || !first_ids || (n_tex && !tex) || (n_mat && !mat) || (n_nodes && !nodes)) { set_error("rt_scene_add_batch: null argument"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  first_ids[0] = (uint32_t)s.textures.size();
  first_ids[1] = (uint32_t)s.materials.size();
  first_ids[2] = (uint32_t)s.nodes.size();
  s.textures.reserve(s.textures.size() + n_tex);
  s.materials.reserve(s.materials.size() + n_mat);
  s.nodes.reserve(s.nodes.size() + n_nodes);
  uint32_t id = 0;
  for (uint32_t k = 0; k < n_tex; k++) { const int32_t st = rt_scene_add_texture(sp, tex + k, &id); if (st != RT_OK) return st; }
  for (uint32_t k = 0; k < n_mat; k++) { const int32_t st = rt_scene_add_material(sp, mat + k, &id); if (st != RT_OK) return st; }
  for (uint32_t k = 0; k < n_nodes; k++) { const int32_t st = rt_scene_add_node(sp, nodes + k, &id); if (st != RT_OK) return st; }
  return RT_OK;
}

int32_t rt_scene_set_noise_tables(RtScene* sp, const int32_t* px, const int32_t* py, const int32_t* pz, const double* g) {
  if (!sp) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  const int32_t* src[3] = {px, py, pz};
  for (int k = 0; k < 3; k++)
    if (src[k]) for (int i = 0; i < 256; i++) s.perm[k][i] = src[k][i];
  if (g) for (int i = 0; i < 256; i++) for (int k = 0; k < 3; k++) s.grad[i][k] = g[3 * i + k];
  s.committed = false;
  return RT_OK;
}

int32_t rt_scene_commit(RtScene* sp, uint32_t root) {
  if (!sp) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
  Scene& s = *reinterpret_cast<Scene*>(sp);
  {
    std::lock_guard<std::mutex> lock(s.devices_mu);   // get_device_scene takes the same lock
    for (auto& kv : s.devices) destroy_device_scene(kv.second);
    s.devices.clear();
  }
  return flatten_scene(s, root);
}

int32_t rt_options_init(RtOptions* o) {
  if (!o) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  options_init(*o);
  return RT_OK;
}
int32_t rt_scene_set_options(RtScene* sp, const RtOptions* o) {
  if (!sp || !o) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  reinterpret_cast<Scene*>(sp)->opts = *o;
  return RT_OK;
}
int32_t rt_scene_get_options(const RtScene* sp, RtOptions* o) {
  if (!sp || !o) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  *o = reinterpret_cast<const Scene*>(sp)->opts;
  return RT_OK;
}

int32_t rt_scene_info(const RtScene* sp, RtSceneInfo* info) {
  if (!sp || !info) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  const Scene& s = *reinterpret_cast<const Scene*>(sp);
  if (!s.committed) { set_error("scene not committed"); return RT_ERR_NOT_COMMITTED; }
  const FlatScene& f = s.flat;
  std::memset(info, 0, sizeof *info);
  info->n_prims = (uint32_t)f.canonical.size();
  info->n_spheres = f.n_spheres;
  info->n_planars = f.n_planars;
  info->n_bvh_nodes = (uint32_t)f.nodes.size();
  info->n_media = f.n_media;
  info->n_materials = (uint32_t)f.materials.size();
  info->n_textures = (uint32_t)f.textures.size();
  info->n_xforms = (uint32_t)f.xforms.size();
  info->surface_root = f.surface_root;
  info->max_depth = f.max_depth;
  info->device_bytes = f.bytes();
  info->n_closed_media = 0;
  for (uint32_t k = 0; k < f.n_media; k++) info->n_closed_media += f.media[k].closed ? 1u : 0u;
  info->n_moving_list = f.moving_count;
  info->n_shared_trees = f.n_shared_trees;
  info->gpu_build = f.deferred.pending ? 1u : f.deferred.built ? 2u : 0u;
  return RT_OK;
}

int32_t rt_scene_get_bvh(const RtScene* sp, float* nodes, uint32_t* canon) {
  if (!sp) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
  const Scene& s = *reinterpret_cast<const Scene*>(sp);
  if (!s.committed) { set_error("scene not committed"); return RT_ERR_NOT_COMMITTED; }
  if (s.flat.deferred.pending) {
    set_error("the surface tree of this scene is built on the device by the first upload (RtOptions.gpu_build): nothing to read yet");
    return RT_ERR_NOT_COMMITTED;
  }
  if (nodes) std::memcpy(nodes, s.flat.nodes.data(), s.flat.nodes.size() * sizeof(DNode));
  if (canon) std::memcpy(canon, s.flat.canonical.data(), s.flat.canonical.size() * 4);
  return RT_OK;
}

int32_t rt_scene_get_eval_order(const RtScene* sp, uint32_t* out) {
  if (!sp || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  const Scene& s = *reinterpret_cast<const Scene*>(sp);
  if (!s.committed) { set_error("scene not committed"); return RT_ERR_NOT_COMMITTED; }
  for (size_t i = 0; i < s.flat.ref_order.size(); i++) out[s.flat.ref_order[i]] = s.flat.canonical[i];
  return RT_OK;
}

int32_t rt_image_height(const RtCamera* cam, int32_t* height) {
  if (!cam || !height) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  if (!(cam->aspect_ratio > 0) || cam->image_width <= 0) { set_error("bad image width / aspect ratio"); return RT_ERR_INVALID_ARG; }
  *height = (int32_t)std::nearbyint((double)cam->image_width / cam->aspect_ratio);  // round-half-even, Ray.hs:123
  return RT_OK;
}

int32_t rt_set_pool_size(RtScene* sp, uint32_t n) {
  if (!sp) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
  reinterpret_cast<Scene*>(sp)->opts.pool = n ? (int64_t)n : -1;
  return RT_OK;
}

int32_t rt_obj_parse(const char* text, size_t len, RtMesh** out) {
  if (!text || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  auto* m = new (std::nothrow) Mesh();
  if (!m) { set_error("out of memory"); return RT_ERR_INVALID_ARG; }
  int32_t st = parse_obj(text, len, *m);
  if (st != RT_OK) { delete m; *out = nullptr; return st; }
  *out = reinterpret_cast<RtMesh*>(m);
  return RT_OK;
}
int32_t rt_mesh_sizes(const RtMesh* mp, uint32_t* nv, uint32_t* nt, uint32_t* nf) {
  if (!mp) { set_error("null mesh"); return RT_ERR_INVALID_ARG; }
  const Mesh& m = *reinterpret_cast<const Mesh*>(mp);
  if (nv) *nv = (uint32_t)(m.verts.size() / 3);
  if (nt) *nt = (uint32_t)(m.uvs.size() / 2);
  if (nf) *nf = (uint32_t)(m.faces.size() / 6);
  return RT_OK;
}
int32_t rt_mesh_data(const RtMesh* mp, const double** v, const double** t, const int32_t** f) {
  if (!mp) { set_error("null mesh"); return RT_ERR_INVALID_ARG; }
  const Mesh& m = *reinterpret_cast<const Mesh*>(mp);
  if (v) *v = m.verts.data();
  if (t) *t = m.uvs.data();
  if (f) *f = m.faces.data();
  return RT_OK;
}
int32_t rt_mesh_destroy(RtMesh* mp) {
  delete reinterpret_cast<Mesh*>(mp);
  return RT_OK;
}

}  // extern "C"
