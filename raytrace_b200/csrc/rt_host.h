// rt_host.h — host-side scene object behind the C ABI (include/rt_b200.h).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_device.h"

namespace rtb {

void set_error(const std::string& msg);  // thread-local last-error string

struct HostTexture {
  RtTextureDesc d;
  std::vector<float> image;  // w*h float4 texels {r, g, b, 0}, linear (converted at add time)
};

struct HostNode {  // one node of the description tree, copied at add time
  int32_t kind = 0;
  uint32_t material = 0;
  std::vector<uint32_t> children;
  double v[3][3] = {};
  double s[2] = {};
  double uv[3][2] = {};
  double m[16] = {};
  std::vector<double> verts, uvs;
  std::vector<int32_t> faces;
};

// Surface tree left to the device (RtOptions.gpu_build; rt_build.cuh): rt_scene_commit stops after the reference-policy pass
// (evaluation order) and packs the primitive records in that order; the FIRST upload builds the binned-SAH tree on the GPU,
// reads nodes + leaf order back, permutes the host arrays once, and every later upload (other devices) is a plain copy.
struct DeferredTree {
  bool pending = false;        // commit deferred the tree and no device has built it yet
  bool built = false;          // the tree in FlatScene::nodes came from the device
  std::vector<BuildItem> items;
  uint32_t first = 0;          // storage slots [first, first + n_prims) belong to the tree
  uint32_t n_prims = 0;
  uint32_t node_base = 0;      // DNode index of the root pair (= FlatScene::surface_root)
  uint32_t collapse = 4;
  float node_cost = 1.f;
  uint32_t levels = 0, launches = 0;
  double build_ms = 0, device_ms = 0;   // whole step (upload of items .. host arrays permuted) / kernels only
};

// flattened, device-ready scene (host copy)
struct FlatScene {
  std::vector<DNode> nodes;
  std::vector<float> prims;      // 16 floats per primitive
  std::vector<float> tri_uv;     // 8 floats per primitive
  std::vector<uint32_t> canonical;
  std::vector<uint32_t> ref_order;   // device primitive index -> position in the reference's evaluation order
  std::vector<DXform> xforms;
  std::vector<DMaterial> materials;
  std::vector<DTexture> textures;
  std::vector<float> texels;     // 4 floats per texel
  std::vector<float> noise_grad; // 256*3
  std::vector<uint8_t> noise_perm; // 3*256
  DMedium media[kMaxMedia];
  uint32_t n_media = 0;
  uint32_t surface_root = kNone;
  uint32_t has_noise = 0;
  uint32_t n_spheres = 0, n_planars = 0;
  uint32_t max_depth = 0;
  uint32_t features = 0;       // TF_* bits: selects the kernel specialisation
  float world_lo[3] = {0, 0, 0}, world_hi[3] = {0, 0, 0};   // box of all primitives (outward rounded)
  uint32_t has_world_box = 0;
  uint32_t moving_first = 0, moving_count = 0;   // TF_MOVLIST
  DeferredTree deferred;
  uint32_t n_shared_trees = 0;   // media whose boundary queries walk the surface tree (boundary geometry == surface geometry)
  uint64_t bytes() const;
};

struct DeviceScene;   // rt_kernels.cu: uploaded buffers + render workspace of one device

struct Scene {
  std::vector<HostTexture> textures;
  std::vector<RtMaterialDesc> materials;
  std::vector<HostNode> nodes;
  int32_t perm[3][256];
  double grad[256][3];
  bool committed = false;
  FlatScene flat;
  RtOptions opts;          // tuning / A-B options (negative field = default); the library reads no environment variables
  std::map<int, DeviceScene*> devices;
  std::mutex build_mu;     // serialises the one device-side tree build of a deferred scene (get_device_scene)
  std::mutex devices_mu;   // uploads may race: renders of one committed scene can come from several OS threads
  Scene();
  ~Scene();
};

// rt_flatten.cpp
int32_t flatten_scene(Scene& scene, uint32_t root);
void default_noise_tables(int32_t perm[3][256], double grad[256][3]);
void options_init(RtOptions& o);

// rt_kernels.cu
void destroy_device_scene(DeviceScene* d);

}  // namespace rtb
