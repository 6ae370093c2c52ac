/*
 * rt_b200.h — C ABI of the B200-native backend for the `raytrace` path-tracing loop.
 *
 * This is the drop-in boundary for ONE hot path of UnaryPlus/raytrace:
 *     raytrace :: ToRandom m => CameraSettings -> Geometry m Material -> StdGen -> A.Matrix D Color
 *     (reference: src/Graphics/Ray.hs:121-238)
 * The reference has no FFI at all (raytrace.cabal:30-49); its "plugin interface" is the
 * combinator vocabulary exported by Graphics.Ray.* (Ray.hs:7-20, Geometry.hs:5-16,
 * Material.hs:3-6, Texture.hs:2-5).  A front end (Haskell `foreign import ccall`, the Python
 * mirror in raytrace_b200/, or plain C++) reifies those combinators as a DESCRIPTION TREE and
 * hands it over with the calls below; the library validates it, flattens it into device
 * buffers, builds the BVH with the reference's split policy (Geometry.hs:369-377) and renders
 * with hand-written sm_100a CUDA kernels.
 *
 * Conventions
 *   - plain C, `extern "C"`, POD structs, caller-owned pointers (the library copies what it
 *     needs during the call), no C++ exceptions or aborts across the boundary;
 *   - every entry point returns an RtStatus; rt_last_error() gives a thread-local message;
 *   - scenes with user-defined Haskell closures (Material / Texture / planeShape predicate /
 *     background that cannot be classified) are REJECTED with RT_ERR_USER_DEFINED — there is
 *     no CPU fallback anywhere in this library;
 *   - all inputs are double (the reference's numeric type, Core.hs:29-31); the device
 *     computes in fp32 (fp64 only for the large-radius sphere discriminant).
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_VERSION 0x00020001u

typedef enum RtStatus {
  RT_OK = 0,
  RT_ERR_INVALID_ARG = 1,     /* null pointer, bad id, bad enum                            */
  RT_ERR_EMPTY_BVH = 2,       /* `bvhTree []` — reference: error "bvhTree: empty list" (Geometry.hs:371) */
  RT_ERR_USER_DEFINED = 3,    /* user-defined material/texture/shape/background: explicit rejection */
  RT_ERR_NON_EUCLIDEAN = 4,   /* `transform` with a non-Euclidean matrix (Geometry.hs:379-381) */
  RT_ERR_TYPE = 5,            /* description tree ill-typed (e.g. root without material)    */
  RT_ERR_CUDA = 6,            /* CUDA runtime failure                                       */
  RT_ERR_NO_DEVICE = 7,       /* no sm_100 device visible                                   */
  RT_ERR_NOT_COMMITTED = 8,   /* render before commit                                       */
  RT_ERR_PARSE = 9,           /* .obj parse error (message mirrors Geometry.hs:230-285)     */
  RT_ERR_UNSUPPORTED = 10     /* valid reference construct this build cannot flatten yet    */
} RtStatus;

/* ---- textures: Graphics.Ray.Texture (Texture.hs:15-79) ------------------------------- */
typedef enum RtTextureKind {
  RT_TEX_CONSTANT = 1,  /* constantTexture c0                       Texture.hs:18-19 */
  RT_TEX_CHECKER = 2,   /* checkerTexture i0 i1 c0 c1               Texture.hs:45-53 */
  RT_TEX_IMAGE = 3,     /* imageTexture (linear RGB, rows top→down) Texture.hs:31-41 */
  RT_TEX_NOISE = 4,     /* noiseTexture i0 freq shift c0 c1         Texture.hs:56-67 */
  RT_TEX_MARBLE = 5,    /* marbleTexture dir freq shift             Texture.hs:70-79 */
  RT_TEX_USER = 6       /* solidTexture / uvTexture / Texture with a user closure → rejected */
} RtTextureKind;

typedef struct RtTextureDesc {
  int32_t kind;
  int32_t i0, i1;       /* checker: n_u, n_v; noise: number of layers k */
  int32_t width, height;/* image */
  double freq;          /* noise / marble */
  double c0[3], c1[3];  /* colours */
  double shift[3];      /* noise / marble */
  double dir[3];        /* marble stripe direction */
  const float* image;   /* width*height*3 linear-light floats, row 0 = top row (readImage, Ray.hs:241-245) */
} RtTextureDesc;

/* ---- materials: Graphics.Ray.Material (Material.hs:41-129) ---------------------------- */
typedef enum RtMaterialKind {
  RT_MAT_LIGHT = 1,           /* lightSource tex           Material.hs:41-42  */
  RT_MAT_PITCH_BLACK = 2,     /* pitchBlack                Material.hs:46-47  */
  RT_MAT_LAMBERTIAN = 3,      /* lambertian tex            Material.hs:51-53  */
  RT_MAT_LOMMEL_SEELIGER = 4, /* lommelSeeliger tex        Material.hs:56-61  */
  RT_MAT_MIRROR = 5,          /* mirror tex                Material.hs:64-67  */
  RT_MAT_METAL = 6,           /* metal param tex           Material.hs:72-78  */
  RT_MAT_DIELECTRIC = 7,      /* dielectric param          Material.hs:89-106 */
  RT_MAT_TRANSPARENT = 8,     /* transparent tex           Material.hs:109-112 */
  RT_MAT_ISOTROPIC = 9,       /* isotropic tex             Material.hs:116-118 */
  RT_MAT_ANISOTROPIC = 10,    /* anisotropic param tex     Material.hs:124-129 */
  RT_MAT_USER = 11            /* `Material` built from a user closure → rejected */
} RtMaterialKind;

typedef struct RtMaterialDesc {
  int32_t kind;
  uint32_t texture;     /* id from rt_scene_add_texture (ignored by pitchBlack / dielectric) */
  double param;         /* metal: fuzz; dielectric: ior; anisotropic: g */
} RtMaterialDesc;

/* ---- geometry description tree: Graphics.Ray.Geometry (Geometry.hs:42-456) ------------ */
typedef enum RtNodeKind {
  RT_NODE_SPHERE = 1,         /* sphere v[0] s[0]                          Geometry.hs:58-94   */
  RT_NODE_PARALLELOGRAM = 2,  /* parallelogram q=v[0] u=v[1] v=v[2]        Geometry.hs:147-151 */
  RT_NODE_TRIANGLE = 3,       /* triangle (v[0],uv[0]) (v[1],uv[1]) (v[2],uv[2])  Geometry.hs:169-176 */
  RT_NODE_CUBOID = 4,         /* cuboid box: lo=v[0], hi=v[1]              Geometry.hs:154-166 */
  RT_NODE_MESH = 5,           /* triangleMesh (Mesh verts uvs faces)       Geometry.hs:179-184,288-294 */
  RT_NODE_GROUP = 6,          /* group children                            Geometry.hs:336-347 */
  RT_NODE_BVH_NODE = 7,       /* bvhNode children[0] children[1]           Geometry.hs:352-363 */
  RT_NODE_BVH_TREE = 8,       /* bvhTree children                          Geometry.hs:369-377 */
  RT_NODE_TRANSFORM = 9,      /* transform m children[0]                   Geometry.hs:382-391 */
  RT_NODE_MOVING = 10,        /* moving v[0] v[1] children[0]              Geometry.hs:449-456 */
  RT_NODE_MEDIUM = 11,        /* constantMedium s[0] children[0]           Geometry.hs:298-330 */
  RT_NODE_WITH_MATERIAL = 12, /* material <$ children[0]                   Geometry.hs:44-47   */
  RT_NODE_PURE = 13,          /* pureGeometry children[0]                  Geometry.hs:50-51   */
  RT_NODE_USER = 14           /* planeShape with user predicate / raw Geometry closure → rejected */
} RtNodeKind;

typedef struct RtNodeDesc {
  int32_t kind;
  uint32_t material;          /* WITH_MATERIAL: id from rt_scene_add_material */
  uint32_t n_children;
  uint32_t n_verts, n_uvs, n_faces;   /* MESH */
  const uint32_t* children;   /* node ids, in the reference's list order */
  const double* verts;        /* MESH: n_verts*3 */
  const double* uvs;          /* MESH: n_uvs*2 (may be NULL when n_uvs == 0) */
  const int32_t* faces;       /* MESH: n_faces*6 = (v0,t0,v1,t1,v2,t2), 0-based, t = -1 for Nothing */
  double v[3][3];             /* vector slots, see RtNodeKind */
  double s[2];                /* scalar slots: sphere radius / medium density */
  double uv[3][2];            /* TRIANGLE texture coordinates */
  double m[16];               /* TRANSFORM: M44 row-major, bottom row 0 0 0 1 */
} RtNodeDesc;

/* ---- camera: CameraSettings (Ray.hs:40-98), field for field --------------------------- */
typedef enum RtBackgroundKind {
  RT_BG_CONSTANT = 1,   /* const c0                                  (default, Ray.hs:94)   */
  RT_BG_YGRADIENT = 2   /* (1-a)*c0 + a*c1, a = 0.5*(dir.y+1)        (sky / grayFade, test/Main.hs:19-28) */
} RtBackgroundKind;

typedef struct RtRedirectTarget {   /* one element of cs_redirectTargets (Ray.hs:64) */
  double p;
  double q[3], s0[3], s1[3];
} RtRedirectTarget;

typedef struct RtCamera {
  double center[3];             /* cs_center  */
  double look_at[3];            /* cs_lookAt  */
  double up[3];                 /* cs_up      */
  double vfov;                  /* cs_vfov (radians) */
  double aspect_ratio;          /* cs_aspectRatio */
  double defocus_angle;         /* cs_defocusAngle */
  double focus_dist;            /* cs_focusDist */
  double bg_c0[3], bg_c1[3];    /* cs_background, classified by the front end */
  int32_t bg_kind;
  int32_t image_width;          /* cs_imageWidth */
  int32_t samples_per_pixel;    /* cs_samplesPerPixel */
  int32_t max_recursion_depth;  /* cs_maxRecursionDepth */
  int32_t n_redirect_targets;   /* length cs_redirectTargets */
  int32_t _pad;
  const RtRedirectTarget* redirect_targets;
} RtCamera;

typedef struct RtStats {
  uint64_t samples;        /* camera paths started                                               */
  uint64_t rays;           /* path segments that reached a closest-hit query (Ray.hs:178)         */
  uint64_t launches;       /* kernels launched by this call                                       */
  uint32_t iterations;     /* wavefront iterations                                                */
  uint32_t pool;           /* live-ray pool size used                                             */
  double ms_total;         /* device time of the render (CUDA events), copies excluded            */
  double ms_traverse;      /* device time spent in the traversal/closest-hit kernel               */
  double ms_shade;         /* device time spent in the shading/scatter/compaction kernel          */
  double ms_raygen;        /* device time spent in camera-ray generation                          */
  double ms_copy;          /* host<->device copies done by this call                              */
  double ms_finish;        /* device time of the tail kernel (k_finish); ms_raygen is what is left of ms_total */
  uint64_t rays_finish;    /* segments traced by the tail kernel (included in rays)               */
  uint64_t paths_finish;   /* paths the tail kernel carried to their end                          */
  /* Work EXECUTED by the GPU traversal (k_traverse_p), counted on the device when RtOptions.count_work = 1 (an
   * instrumented variant of the kernel; zeros otherwise).  These - not the reference's test counts - are what the
   * traversal roofline is computed from: bytes = 64 per node pair, 64 per primitive record, 64 per box record. */
  uint64_t node_pairs;     /* 64-byte child-pair fetches (two slab tests each)                    */
  uint64_t sphere_tests;   /* sphere records fetched and tested                                   */
  uint64_t planar_tests;   /* parallelogram / triangle records fetched and tested                 */
  uint64_t box_tests;      /* aggregated-cuboid slab tests (one 64-byte box record each)          */
  uint64_t rays_counted;   /* path segments covered by the four counters above                    */
} RtStats;

typedef struct RtSceneInfo {
  uint32_t n_prims, n_spheres, n_planars;
  uint32_t n_bvh_nodes;       /* 32-byte nodes uploaded */
  uint32_t n_media, n_materials, n_textures, n_xforms;
  uint32_t surface_root;
  uint32_t max_depth;         /* deepest root→leaf path in the device tree */
  uint64_t device_bytes;      /* scene bytes resident in HBM */
  uint32_t n_closed_media;    /* media whose boundary is a shared, closed, outward-oriented surface (interior queries
                               * are then bounded by the free-flight distance) */
  uint32_t n_moving_list;     /* moving / instanced primitives kept outside the tree in a short list scanned once per segment
                               * (RtOptions.split_moving): the static primitives then run the kernel without the transform branch */
  uint32_t n_shared_trees;    /* media whose boundary geometry IS the surface geometry: their boundary queries walk the surface
                               * tree, no second tree is built or uploaded (pawn, bunny) */
  uint32_t gpu_build;         /* surface tree: 0 = built by the host at commit; 1 = left to the device, not built yet (the first
                               * upload / render builds it: n_bvh_nodes and max_depth then cover the media trees only);
                               * 2 = built on the device (RtOptions.gpu_build) */
} RtSceneInfo;

typedef struct RtScene RtScene;

/*
 * Tuning / A-B options of one scene.  Every field: negative = library default (rt_options_init sets them all to -1).
 * Commit-time fields are read by rt_scene_commit, render-time fields by every render of the scene.  Every combination
 * renders the SAME samples (equal-seed images equal up to fp32 summation order; tests/test_gpu_edge_cases.py) - these are
 * performance knobs and parity switches, not quality settings.  The library reads no environment variables.
 */
typedef struct RtOptions {
  /* ---- commit time ---- */
  int32_t bvh_reference;     /* 1: device tree = the reference's object-median tree (Geometry.hs:369-377); 0: binned SAH (default) */
  int32_t leaf_max;          /* a `group` of at most this many primitives is one scanned leaf (default 24, <= 32) */
  int32_t collapse;          /* subtrees of at most this many primitives become one leaf (default 4) */
  int32_t aggregate_boxes;   /* 1 (default): a cuboid's six faces are answered by one slab test; 0: six planar tests */
  int32_t shared_boundary;   /* 1 (default): a medium boundary that is also a surface is answered from the surface hit */
  int32_t interior_bound;    /* 1 (default): interior segments of a closed shared boundary are bounded by the free-flight distance */
  int32_t world_box;         /* 1 (default): keep the world box (camera rays that miss it end in k_raygen) */
  int32_t split_moving;      /* 1: moving instances live in a short list outside the tree so that static primitives run the kernel without the per-primitive transform branch (default 0: measured slower on demo2) */
  int32_t trace_commit;      /* 1: print the phases of the host scene pipeline to stderr */
  int32_t gpu_build;         /* surface tree built by the GPU at the first upload (binned SAH, same cost model as the host builder):
                              * default: trees of >= 1024 items; 0: never (host builder); 1: every tree with an inner node */
  double sah_node_cost;      /* cost of a node step in primitive tests (default 1.0) */
  /* ---- render time ---- */
  int64_t pool;              /* live-path pool size (default 2^25; rounded up to 1024) */
  int64_t finish_max;        /* tail kernel takes over at <= this many live paths (0 = never, >= 2^32-1 = as soon as all paths started) */
  int32_t finish_refill;     /* tail kernel: idle lanes per warp that trigger a claim (default 4) */
  int32_t tail_batch;        /* iterations per host round trip once all paths have started (default 3) */
  int32_t node_burst;        /* k_traverse_p: node steps per vote (default 4) */
  int32_t refill;            /* k_traverse_p: idle lanes per warp that trigger a fetch (default 24) */
  int32_t trav_blocks;       /* k_traverse_p: blocks per SM in its persistent grid (default: what the occupancy query says is resident, 8 or 9) */
  int32_t tile;              /* 1: camera paths start in 8x4 pixel tiles (default for BVH scenes) */
  int32_t primary_cull;      /* 1/0: force camera-ray culling against the world box on/off (default: on when the lens is outside it) */
  int32_t shade_sort;        /* 1: k_shade handles paths grouped by material class (default: on when the scene has Perlin or image textures) */
  int32_t timing;            /* 0: no per-kernel events even when stats are requested (default 1) */
  int32_t count_work;        /* 1: run the instrumented traversal kernel and fill RtStats.node_pairs ... (default 0) */
  int32_t debug_sync;        /* 1: synchronise after every launch and name a faulting kernel on stderr */
  int32_t raygen_block;      /* culling k_raygen: consecutive samples of one pixel owned by a thread (default 8; 1 = one atomic triple per culled sample) */
  int32_t material_sampling; /* 1: a material draws the direction of the no-redirection case from its own generator - the generalisation
                              * sketched in Material.hs:36-38; today `anisotropic g` samples its Henyey-Greenstein phase function.  Same
                              * expectation, other samples, less variance.  Default 0: the reference's estimator, sample for sample */
  int32_t smem_carveout;     /* shared-memory carve-out asked for the traversal kernels: default = the smallest supported size that holds
                              * the resident blocks' stacks (the rest of the 256 KB is L1 for node pairs and primitive records; pawn
                              * traversal -4 %); 0: the attribute is never touched (the driver's choice: 132 KB for 63 KB used);
                              * > 0: that percentage of the SM's maximum as the hint */
  int32_t reserved[5];
} RtOptions;

/* library */
uint32_t rt_version(void);
const char* rt_last_error(void);
int32_t rt_device_count(int32_t* n);                 /* sm_100 devices visible */

/* scene building (description tree) */
int32_t rt_scene_create(RtScene** out);
int32_t rt_scene_destroy(RtScene* scene);
int32_t rt_scene_add_texture(RtScene* scene, const RtTextureDesc* desc, uint32_t* id);
int32_t rt_scene_add_material(RtScene* scene, const RtMaterialDesc* desc, uint32_t* id);
int32_t rt_scene_add_node(RtScene* scene, const RtNodeDesc* desc, uint32_t* id);
/* A whole description in ONE call: n_textures + n_materials + n_nodes records, added in array order exactly as the three
 * calls above would (textures first, then materials, then nodes; ids are consecutive, children before parents, so a front
 * end that marshals an immutable world once - the constructors of Geometry.hs / Material.hs / Texture.hs evaluated into
 * these records - can hand the same arrays to every scene it creates).  first_ids[3]: id of textures[0], materials[0],
 * nodes[0].  Stops at the first invalid record (the scene keeps what was added before it). */
int32_t rt_scene_add_batch(RtScene* scene, const RtTextureDesc* textures, uint32_t n_textures, const RtMaterialDesc* materials,
                           uint32_t n_materials, const RtNodeDesc* nodes, uint32_t n_nodes, uint32_t* first_ids /*3*/);
/* Perlin tables of Noise.hs:52-92; the three permutation tables default to the reference's
 * literals, the gradient table to a SplitMix restatement of `mkStdGen 666` (parity unpinned). */
int32_t rt_scene_set_noise_tables(RtScene* scene, const int32_t* perm_x, const int32_t* perm_y,
                                  const int32_t* perm_z, const double* gradients /*256*3*/);

/* validate + flatten + build BVH (host only; no GPU needed) */
int32_t rt_scene_commit(RtScene* scene, uint32_t root);
int32_t rt_scene_info(const RtScene* scene, RtSceneInfo* info);
/* read back the flattened host-side buffers (tests: BVH shape equality with the oracle) */
int32_t rt_scene_get_bvh(const RtScene* scene, float* nodes /*n_bvh_nodes*8, may be NULL*/,
                         uint32_t* prim_canonical_id /*n_prims, may be NULL*/);
/* canonical primitive ids in the REFERENCE's evaluation order (group: list order, Geometry.hs:340-345; bvhNode / bvhTree:
 * left subtree first, :356-361,369-377) — the order that decides exact ties.  The device tree is free to store the
 * primitives in another order (its own leaf order); this is the order the oracle's traversal must reproduce. */
int32_t rt_scene_get_eval_order(const RtScene* scene, uint32_t* canonical_id_out /*n_prims*/);

/* upload to a device (one scene may be resident on several devices) */
int32_t rt_scene_upload(RtScene* scene, int32_t device);

/* camera helper: imageHeight = round-half-even (W / aspect)  (Ray.hs:123) */
int32_t rt_image_height(const RtCamera* cam, int32_t* height);

/*
 * Render global sample indices [first_sample, first_sample + n_samples) of every pixel.
 * rgb_out: HOST buffer, H*W*3 floats, row-major, row 0 = top (index (j :. i), Ray.hs:238),
 *          linear RGB, unclamped, already divided by cam->samples_per_pixel.
 * rgb_sumsq_out: optional HOST buffer, same shape: Σ x² per channel (not divided).
 */
int32_t rt_render(RtScene* scene, int32_t device, const RtCamera* cam, uint64_t seed,
                  uint64_t first_sample, uint64_t n_samples, float* rgb_out,
                  float* rgb_sumsq_out, RtStats* stats);

/*
 * Same, accumulating UN-normalised sums into DEVICE buffers owned by the caller (e.g. a torch
 * tensor), on the caller's stream; used by the one-process-per-GPU sample-sharded path where
 * an NCCL reduce follows.  d_sum / d_sumsq must be zero-initialised by the caller.
 * cuda_stream is a cudaStream_t (NULL = the default stream).
 * Completion is ordered on the stream: operations the caller queues on it afterwards see the finished sums.  With
 * stats == NULL nothing is read back, and when the launch sequence is known up front (flat scene whose sample range fits
 * the path pool - one GPU's share of a small render) the call returns without synchronising the host.
 */
int32_t rt_render_device(RtScene* scene, int32_t device, const RtCamera* cam, uint64_t seed,
                         uint64_t first_sample, uint64_t n_samples, float* d_sum,
                         float* d_sumsq /*nullable*/, void* cuda_stream, RtStats* stats);

/* Deterministic pixel-centre primary rays (no lens, given shutter time): canonical primitive
 * id (0xFFFFFFFF = miss, 0xFFFFFFFE = medium) and hit distance per pixel.  HOST buffers H*W. */
int32_t rt_probe_primary(RtScene* scene, int32_t device, const RtCamera* cam, double time,
                         uint32_t* prim_id_out, float* t_out);

/* Deterministic shading inputs of the same pixel-centre primary hit: 10 floats per pixel = world normal facing the ray
 * (HitRecord hr_normal, Core.hs:158), uv (hr_uv), front flag (hr_frontSide), the material's texture colour at the hit
 * (Texture.hs:15: tex point uv; zero for pitchBlack / dielectric) and the material kind; zeros on a miss or a medium.
 * HOST buffer H*W*10. */
int32_t rt_probe_surface(RtScene* scene, int32_t device, const RtCamera* cam, double time, float* out);

/* The one collective of the sample-sharded path (one reduce of the fp32 image to the root), done over NVLink
 * peer memory by the root's own kernel instead of ncclReduce.  Every rank keeps its un-normalised sum buffer in symmetric
 * memory (mapped by all peers; torch.distributed._symmetric_memory or cuMem + cuMulticast in a C host); after a barrier
 * across ranks the ROOT calls this on its stream:
 *   multicast_ptr != NULL: one pass of multimem.ld_reduce over the NVSwitch multicast mapping - the switch sums the G
 *                          buffers in the fabric (NVLS) and the root ingests the image once;
 *   else:                  the root sums peer_ptrs[0 .. n_peers) (its own buffer included) with peer loads.
 * d_out[k] = scale * sum_r buffer_r[k] (scale = 1/spp, Ray.hs:232), d_out a root-local device buffer that is NOT one
 * of the summed buffers.  n_floats must be a multiple of 4 and all pointers 16-byte aligned (pad the image). */
int32_t rt_reduce_image(int32_t device, const void* multicast_ptr, const void* const* peer_ptrs, int32_t n_peers,
                        float* d_out, uint64_t n_floats, float scale, void* cuda_stream);

/*
 * Multi-GPU rendering inside the library: the reference's `raytrace` is ONE call that uses the whole machine
 * (Ray.hs:121, parallel array :238).  A context owns a set of GPUs in one process; rt_render_multi splits the samples per
 * pixel across them (the sample set does not depend on the GPU count), renders every share on its own device
 * (one host thread and stream per device, the committed scene uploaded once per device) and assembles the image with a
 * reduce-scatter over NVLink: every device reduces ITS 1/G slice of the image through the NVSwitch multicast mapping
 * (multimem.ld_reduce; cuMemCreate + cuMulticast*, resolved at run time), normalises it by 1/spp (Ray.hs:232) and copies
 * that slice straight into the caller's host image.  Falls back to peer loads, then to a gather on the first device,
 * when the fabric has no multicast / no peer access (RtContextInfo.reduce_mode says which) - always on the GPUs.
 */
typedef struct RtContext RtContext;
typedef struct RtContextInfo {
  int32_t n_devices;
  int32_t reduce_mode;      /* 0 = single device (no reduce), 1 = NVSwitch multicast, 2 = peer loads, 3 = gather on the first device;
                               decided at the first render */
  int32_t devices[16];
  char reduce_note[160];    /* why multicast was not used, if it was not */
} RtContextInfo;
/* devices == NULL / n_devices == 0: every visible sm_100 device.  A device may be listed more than once (each entry is a
 * share with its own host thread and stream; the reduce then uses peer loads) - how a one-GPU box exercises the G > 1 path. */
int32_t rt_context_create(const int32_t* devices, int32_t n_devices, RtContext** out);
int32_t rt_context_destroy(RtContext* ctx);
int32_t rt_context_info(const RtContext* ctx, RtContextInfo* info);
/* Same result as rt_render(scene, dev, cam, seed, 0, cam->samples_per_pixel, ...) up to fp32 summation order.
 * stats (nullable): sums over the devices; times are the maximum over the devices. */
int32_t rt_render_multi(RtContext* ctx, RtScene* scene, const RtCamera* cam, uint64_t seed, float* rgb_out,
                        float* rgb_sumsq_out /*nullable*/, RtStats* stats /*nullable*/);

/* Optional page-locked host buffers for rt_render's outputs.  rt_render accepts ANY host pointer; when rgb_out (and
 * rgb_sumsq_out) point into page-locked memory - from rt_host_alloc, cudaMallocHost or cudaHostRegister - the image is
 * DMA-copied straight into them, otherwise it goes through the library's pinned staging buffer plus one host memcpy
 * (0.5 ms for a 600x600 image).  Freed buffers are kept on a small free list and handed out again (cudaMallocHost costs
 * more than the copy it saves).  The Haskell binding wraps the pointer in a ForeignPtr whose finaliser is rt_host_free,
 * which is what the massiv array of the result (Ray.hs:238) is built over. */
int32_t rt_host_alloc(size_t bytes, void** out);
int32_t rt_host_free(void* p);

/* tuning knobs */
int32_t rt_options_init(RtOptions* opts);                              /* every field = -1 (library default) */
int32_t rt_scene_set_options(RtScene* scene, const RtOptions* opts);   /* commit-time fields: call before rt_scene_commit */
int32_t rt_scene_get_options(const RtScene* scene, RtOptions* opts);
int32_t rt_set_pool_size(RtScene* scene, uint32_t n_paths);            /* = options.pool (0 = default) */

/* image output step after the path (writeImage / writeImageSqrt, Ray.hs:248-260):
 * linear RGB float → 8-bit with the sRGB OETF (mode 0) or sqrt gamma (mode 1), clamped. */
int32_t rt_tonemap_u8(int32_t device, const float* rgb_host, int32_t n_pixels, int32_t mode,
                      uint8_t* rgb8_host);

/*
 * Allocation-free variant of rt_tonemap_u8.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernel arithmetic and same result as rt_tonemap_u8, but the staging memory is a file-scope
 * `static __device__` array inside the library instead of two cudaMalloc'ed buffers: the call performs no cudaMalloc,
 * cudaFree, stream / event creation or any other allocating routine.
 *   - limit: n_pixels <= RT_TONEMAP_STATIC_MAX_PIXELS (2 Mi pixels - a 1920 x 1080 frame fits); more is rejected with
 *     RT_ERR_INVALID_ARG and a message.  Every GPU that loads the library holds the array (30 MiB) until the process exits.
 *   - rgb / rgb8_out may be host OR device pointers (the copies use cudaMemcpyDefault); the work is queued on cuda_stream
 *     (a cudaStream_t, NULL = the default stream) and the call waits for it unless that stream is being captured.
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress, which loads the module on that
 *     device): make that call once outside any capture.
 *   - every call on a device SHARES the one array, and it keeps the last call's contents (nothing in it needs to start at
 *     zero: a call overwrites what it reads).  Calls that overlap on the same GPU - two streams, two threads - race on it:
 *     serialise them.
 */
#define RT_TONEMAP_STATIC_MAX_PIXELS 2097152
int32_t rt_tonemap_u8_static(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode,
                             uint8_t* rgb8_out, void* cuda_stream);

/* .obj loader with the rules of parseObj (Geometry.hs:207-285). Buffers are owned by the
 * returned handle. */
typedef struct RtMesh RtMesh;
int32_t rt_obj_parse(const char* text, size_t len, RtMesh** out);
int32_t rt_mesh_sizes(const RtMesh* mesh, uint32_t* n_verts, uint32_t* n_uvs, uint32_t* n_faces);
int32_t rt_mesh_data(const RtMesh* mesh, const double** verts, const double** uvs,
                     const int32_t** faces);
int32_t rt_mesh_destroy(RtMesh* mesh);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
