// rt_kernels.cu — the wavefront path tracer (sm_100a) and the CUDA half of the C ABI.
//
// One render = a loop of three kernels over a pool of P live paths:
//   k_raygen   : refill the free tail of the current queue with camera rays (Ray.hs:157-172,229),
//                Philox counter RNG keyed by (seed; pixel, global sample, bounce, block)
//   k_traverse : closest hit per live ray — BVH traversal with __ldg float4 node-pair fetches and a
//                shared-memory short stack; media boundary queries + free-flight sampling
//                (three variants: flat constant-bank list for leaf-only scenes, persistent vote-scheduled warps for BVH
//                scenes, static while-while as the A/B baseline)
//   k_shade    : material / texture / redirect-mixture evaluation, radiance accumulation, and block-aggregated
//                ballot/popc stream compaction of the surviving rays into the next queue (queue tiles prefetched by TMA
//                bulk copies); the last block to finish prepares the next iteration (no host round trip, no extra launch)
//   k_finish   : tail kernel — once few paths are left, one thread carries one path to its end
// The host reads the control block back only once per batch of iterations.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "rt_host.h"
#include "rt_shade.cuh"
#include "rt_build.cuh"

namespace rtb {

#define RT_CUDA(expr)                                                                            \
  do {                                                                                           \
    cudaError_t e_ = (expr);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      set_error(std::string(#expr) + ": " + cudaGetErrorString(e_));                             \
      return RT_ERR_CUDA;                                                                        \
    }                                                                                            \
  } while (0)

// Every ABI entry that selects a device restores the caller's current device on the way out (a torch process must not find
// torch.cuda.current_device() changed behind its back).
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); } }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
struct DevBuf {                   // cudaMalloc'ed scratch freed on every exit path
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
};

// No C++ exception crosses the C ABI: an entry point's body runs inside this guard (std::bad_alloc from a vector, ...).
template <class F>
static int32_t abi_guard(const char* what, F&& body) {
  try {
    return body();
  } catch (const std::exception& e) {
    set_error(std::string(what) + ": " + e.what());
    return RT_ERR_CUDA;
  } catch (...) {
    set_error(std::string(what) + ": unknown C++ exception");
    return RT_ERR_CUDA;
  }
}

constexpr int kTraverseBlock = 128;
constexpr int kShadeBlock = 256;
#ifndef RT_SHADE_MINB
#define RT_SHADE_MINB 4   // resident k_shade blocks per SM the register allocation must allow (4 → 64 regs)
#endif
constexpr int kRaygenBlock = 256;

struct Ctl {                      // device-resident control block of one render
  unsigned long long next_item;   // next (sample, pixel) work item to start
  unsigned long long total_items;
  unsigned long long rays;        // path segments traversed so far
  unsigned long long fill_item0;  // this iteration: first work item k_raygen turns into a camera ray
  unsigned int n_cur, n_next;     // live paths entering the iteration / compacted survivors leaving it
  unsigned int n_live, fill_lo;   // this iteration: queue slots [0, n_live) are live, [fill_lo, n_live) freshly generated
  unsigned int iterations, pool;
  unsigned int trav_next;         // next queue slot the persistent traversal warps fetch
  unsigned int blocks_done;       // k_shade blocks finished (the last one prepares the next iteration)
  unsigned int paths_finish;      // paths the tail kernel took over
  unsigned int n_gen, gen_done;   // k_raygen with primary-ray culling: camera paths queued so far / blocks finished
  unsigned int culled;            // ... camera rays of this iteration that missed the world box (ended on the background)
  unsigned long long rays_finish; // segments traced by the tail kernel (also counted in rays)
  unsigned long long work[5];     // instrumented traversal (RtOptions.count_work): node pairs, sphere / planar / box tests, rays
};

struct Queue {                    // SoA live-ray queue (56 B per path)
  float4* o;                      // origin.xyz, shutter time
  float4* d;                      // direction.xyz, origin geometry id (bits)
  float4* thr;                    // throughput.rgb, pixel index (bits)
  uint2* meta;                    // global sample index (lo 32), bounce
};

struct RenderArgs {
  DCamera cam;
  uint2 key;
  unsigned long long first_sample;
  uint32_t n_pixels;
  float* sum;
  float* sumsq;
  float2* hits;                   // per queue slot: t, prim (bits)
  // per queue slot: the free-flight bound of an interior segment (rt_trace.cuh::starts_inside_only_medium), drawn by the kernel
  // that queued the path - k_shade, which waits on HBM and has issue slots to spare - so that the latency-bound traversal
  // fetches 4 bytes instead of running Philox + log behind two more dependent queue loads.  Other slots: unspecified.
  float* bound;
  int refill;                     // persistent traversal: idle lanes per warp that trigger a fetch
  int node_burst;                 // persistent traversal: node steps per vote
  uint32_t finish_max;            // k_finish takes over when at most this many paths are left (0 = never)
  int finish_refill;              // k_finish: idle lanes per warp that trigger a claim of new paths
  uint32_t tile;                  // camera paths start in 8x4 pixel tiles (1) or in row-major pixel order (0)
  uint32_t cull;                  // k_raygen: camera rays that miss the world box end there (never queued)
  float world_lo[3], world_hi[3];
  // work-item order: with kblock = K > 1, K consecutive items are K consecutive samples of ONE pixel (a k_raygen thread owns
  // them and sums the background of its culled camera rays in registers: one atomic triple per pixel instead of one per
  // sample); kblock = 1: an item is (sample pass, pixel).  A bijection onto (pixel, sample) either way - the RNG is keyed by
  // (pixel, global sample), so the sample set does not depend on it.
  uint32_t shade_sort;                  // k_shade: deal the paths of a tile to the threads by material class
  uint32_t kblock, ktail;               // K; n_samples % K
  unsigned long long kfull_items;       // (n_samples / K) * K * n_pixels: items before the ragged last block of samples
};

// Closes iteration k and prepares iteration k+1 (single thread: the last k_shade block, or k_begin for k = 0).
__device__ void prepare_iteration(volatile Ctl* ctl, bool first) {
  unsigned int n_cur = 0;
  if (!first) {
    n_cur = ctl->n_next;
    ctl->rays += ctl->n_live;
    if (ctl->n_live) ctl->iterations += 1;   // launches past the end of the render are empty and do not count
  }
  const unsigned long long rem = ctl->total_items - ctl->next_item;
  const unsigned int room = ctl->pool - n_cur;
  const unsigned int fill = rem < room ? (unsigned int)rem : room;
  ctl->n_cur = n_cur;
  ctl->n_next = 0;
  ctl->trav_next = 0;
  ctl->fill_lo = n_cur;
  ctl->n_live = n_cur + fill;
  ctl->fill_item0 = ctl->next_item;
  ctl->next_item += fill;
}

// ---- camera rays -----------------------------------------------------------------------------------
__device__ __forceinline__ RayT camera_ray(const DCamera& cam, uint32_t pixel, float time, float lu, float lv, float px, float py) {
  uint32_t j = pixel / (uint32_t)cam.W, i = pixel - j * (uint32_t)cam.W;
  RayT r;
  r.time = time;
  r.o = ld3(cam.center);
  // target - centre, built from small quantities only
  f3 rel = fma3((float)i + px, ld3(cam.pixel_u), fma3((float)j + py, ld3(cam.pixel_v), ld3(cam.top_left)));
  if (cam.has_defocus) {   // uniform point on the lens disk (Core.hs:63-68; polar map, same distribution)
    float rad = sqrtf(lu), s, c;
    sincospif(2.f * lv, &s, &c);
    f3 lens = fma3(rad * c, ld3(cam.defocus_u), (rad * s) * ld3(cam.defocus_v));
    r.o = r.o + lens;
    rel = rel - lens;
  }
  r.d = normalize3(rel);
  return r;
}

// Position p of a sample pass → pixel, walking the image in 8x4 tiles (bands of four rows, eight columns at a time;
// the last band / last tile of a band may be smaller): the 32 camera rays of a warp then span a compact patch instead
// of a 32x1 strip, so they visit the same BVH nodes for longer.  A bijection on [0, W*H) for every W, H; the sample
// set is unchanged because the RNG is keyed by the pixel, not by p.
__device__ __forceinline__ uint32_t tiled_pixel(uint32_t p, uint32_t W, uint32_t H) {
  const uint32_t band = p / (4u * W), k = p - band * 4u * W;
  const uint32_t bh = min(4u, H - 4u * band);
  const uint32_t t = k / (8u * bh), r = k - t * 8u * bh;
  const uint32_t tw = min(8u, W - 8u * t);
  const uint32_t jj = r / tw, ii = r - jj * tw;
  return (4u * band + jj) * W + 8u * t + ii;
}

struct NewPath { RayT ray; uint32_t pixel; uint32_t sample_lo; };
// camera path of (pixel, local sample index) (Ray.hs:157-172,229)
__device__ __forceinline__ NewPath make_camera_path(const RenderArgs& A, uint32_t pixel, unsigned long long s_local) {
  const unsigned long long sample = A.first_sample + s_local;
  const uint4 r0 = rng4(A.key, pixel, sample, 0, RNG_RAYGEN);
  float lu = 0.f, lv = 0.f;
  if (A.cam.has_defocus) {
    const uint4 r1 = rng4(A.key, pixel, sample, 0, RNG_RAYGEN + 1);
    lu = u01(r1.x);
    lv = u01(r1.y);
  }
  return NewPath{camera_ray(A.cam, pixel, u01(r0.x), lu, lv, u01(r0.y), u01(r0.z)), pixel, (uint32_t)sample};
}
// Work item -> (position in the pass, local sample index).  kblock <= 1: item = sample pass * n_pixels + position.  kblock = K:
// K consecutive items are K consecutive samples of ONE position (a bijection either way; RenderArgs::kblock).
struct ItemCursor {
  uint32_t pos;                 // position in the sample pass (pixel before the 8x4 tile mapping)
  uint32_t j, kcur;             // sample within the current block of kcur samples
  unsigned long long blk_s0;    // first local sample index of the current block
  __device__ __forceinline__ unsigned long long s_local() const { return blk_s0 + j; }
};
__device__ __forceinline__ ItemCursor decode_item(const RenderArgs& A, unsigned long long item) {
  ItemCursor c;
  if (A.kblock <= 1u) {
    const unsigned long long pass = item / A.n_pixels;
    c.pos = (uint32_t)(item - pass * A.n_pixels);
    c.j = 0u; c.kcur = 1u; c.blk_s0 = pass;
  } else if (item < A.kfull_items) {
    const uint32_t per_blk = A.kblock * A.n_pixels;                   // < 2^32: checked on the host
    const unsigned long long blk = item / per_blk;
    const uint32_t rem = (uint32_t)(item - blk * per_blk);
    c.pos = rem / A.kblock;
    c.j = rem - c.pos * A.kblock; c.kcur = A.kblock; c.blk_s0 = blk * A.kblock;
  } else {                                                           // last block of samples: ktail = n_samples % K per pixel
    const uint32_t idx = (uint32_t)(item - A.kfull_items);
    c.pos = idx / A.ktail;
    c.j = idx - c.pos * A.ktail; c.kcur = A.ktail; c.blk_s0 = A.kfull_items / A.n_pixels;
  }
  return c;
}
// the next work item, without divisions; returns true when the position (hence the pixel) changed
__device__ __forceinline__ bool advance_item(const RenderArgs& A, ItemCursor& c) {
  if (++c.j < c.kcur) return false;
  c.j = 0u;
  if (++c.pos == A.n_pixels) {          // next block of samples
    c.pos = 0u;
    c.blk_s0 += c.kcur;
    if (A.kblock > 1u && c.blk_s0 * A.n_pixels >= A.kfull_items) c.kcur = A.ktail;
  }
  return true;
}
__device__ __forceinline__ uint32_t item_pixel(const RenderArgs& A, const ItemCursor& c) {
  return A.tile ? tiled_pixel(c.pos, (uint32_t)A.cam.W, (uint32_t)A.cam.H) : c.pos;
}
// work item -> camera path
__device__ __forceinline__ NewPath new_camera_path(const RenderArgs& A, unsigned long long item) {
  const ItemCursor c = decode_item(A, item);
  return make_camera_path(A, item_pixel(A, c), c.s_local());
}

// Camera rays of the refilled queue tail.  With A.cull (the scene has a finite world box) a camera ray that misses the box
// cannot hit anything: its path ends on the background right here (Ray.hs:179: miss -> cs_background) and is never
// queued - no 56-byte record, no traversal fetch, no hit record, no shading pass for the large part of the frame an
// object in front of a backdrop leaves empty (pawn, bunny: ~80 % of the camera rays).  Survivors are compacted into
// [fill_lo, fill_lo + n_gen) (warp ballot ranks, one atomic per 256-item tile); the last block publishes n_live.
__global__ void __launch_bounds__(kRaygenBlock) k_raygen(Ctl* __restrict__ ctl, Queue q, const __grid_constant__ RenderArgs A) {
  const uint32_t lo = ctl->fill_lo, n = ctl->n_live;
  const unsigned long long item0 = ctl->fill_item0;
  if (!A.cull) {
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; lo + t < n; t += gridDim.x * blockDim.x) {
      const NewPath np = new_camera_path(A, item0 + t);
      const uint32_t slot = lo + t;
      RT_ASSERT(slot < ctl->pool);
      q.o[slot] = make_float4(np.ray.o.x, np.ray.o.y, np.ray.o.z, np.ray.time);
      q.d[slot] = make_float4(np.ray.d.x, np.ray.d.y, np.ray.d.z, __uint_as_float(kNone));
      q.thr[slot] = make_float4(1.f, 1.f, 1.f, __uint_as_float(np.pixel));
      q.meta[slot] = make_uint2(np.sample_lo, 0u);
    }
    return;
  }
  __shared__ uint32_t s_cnt[kRaygenBlock / 32];
  __shared__ uint32_t s_base;
  // nothing to generate: n_live == fill_lo, or k_finish has just ended the render (n_live = 0 with a stale fill_lo)
  const uint32_t fill = n > lo ? n - lo : 0u, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const float4 wlo = make_float4(A.world_lo[0], A.world_lo[1], A.world_lo[2], 0.f), whi = make_float4(A.world_hi[0], A.world_hi[1], A.world_hi[2], 0.f);
  unsigned int culled = 0;
  // A thread owns K = A.kblock consecutive work items (K samples of one pixel, unless its run straddles a pixel boundary
  // because this iteration's first item is not a multiple of K): the background radiance of its culled rays is summed in
  // registers and flushed with ONE atomic triple when the pixel changes.  The block compacts the survivors of a whole chunk at
  // once (queue order: thread after thread = pixel after pixel inside a 256-pixel tile, a pixel's samples side by side).
  const uint32_t K = max(A.kblock, 1u);
  const uint32_t chunks = (fill + K - 1u) / K;
  uint32_t acc_pixel = kNone;
  f3 acc = mk3(0.f, 0.f, 0.f), acc2 = mk3(0.f, 0.f, 0.f);
  auto flush = [&]() {
    if (acc_pixel != kNone && (acc.x != 0.f || acc.y != 0.f || acc.z != 0.f)) {
      float* sp = A.sum + 3 * (size_t)acc_pixel;
      atomicAdd(sp, acc.x);
      atomicAdd(sp + 1, acc.y);
      atomicAdd(sp + 2, acc.z);
      if (A.sumsq) {
        float* s2 = A.sumsq + 3 * (size_t)acc_pixel;
        atomicAdd(s2, acc2.x);
        atomicAdd(s2 + 1, acc2.y);
        atomicAdd(s2 + 2, acc2.z);
      }
    }
    acc = mk3(0.f, 0.f, 0.f);
    acc2 = mk3(0.f, 0.f, 0.f);
  };
  for (uint32_t cbase = blockIdx.x * blockDim.x; cbase < chunks; cbase += gridDim.x * blockDim.x) {   // uniform per block
    const uint32_t c = cbase + threadIdx.x;
    // index arithmetic once per chunk (a 64-bit division and the tile mapping's four divisions), then increments
    ItemCursor cur{};
    uint32_t cur_pixel = 0u;
    if (c < chunks) { cur = decode_item(A, item0 + (unsigned long long)c * K); cur_pixel = item_pixel(A, cur); }
    const ItemCursor cur0 = cur;
    const uint32_t pixel0 = cur_pixel;
    // Pass 1: the K camera rays of the chunk against the world box.  Misses end here (background summed in registers, one
    // atomic triple per pixel); hits only leave a bit.  No barrier, no atomic inside the loop (the first version compacted
    // the survivors of every one of the K rounds across the block: three barriers and one slot atomic per sample).
    unsigned long long keepmask = 0ull;
    for (uint32_t k = 0; k < K; k++) {
      const uint32_t t = c * K + k;
      if (c < chunks && t < fill) {
        if (k != 0u && advance_item(A, cur)) cur_pixel = item_pixel(A, cur);
        const NewPath np = make_camera_path(A, cur_pixel, cur.s_local());
        f3 inv;
        inv.x = fabsf(np.ray.d.x) > 1e-30f ? __frcp_rn(np.ray.d.x) : copysignf(1e30f, np.ray.d.x);
        inv.y = fabsf(np.ray.d.y) > 1e-30f ? __frcp_rn(np.ray.d.y) : copysignf(1e30f, np.ray.d.y);
        inv.z = fabsf(np.ray.d.z) > 1e-30f ? __frcp_rn(np.ray.d.z) : copysignf(1e30f, np.ray.d.z);
        float tn;
        const bool keep = box_hit(wlo, whi, inv, np.ray.o * inv, 1e-4f, __int_as_float(0x7f800000), tn);
        RT_ASSERT(np.pixel < A.n_pixels);
        if (keep) keepmask |= 1ull << k;
        else {   // miss: emitted = background, path ends (throughput is still 1)
          const f3 cbg = background(A.cam, np.ray.d);
          culled++;
          if (np.pixel != acc_pixel) { flush(); acc_pixel = np.pixel; }
          acc = acc + cbg;
          acc2 = acc2 + cbg * cbg;
        }
      }
    }
    flush();
    acc_pixel = kNone;
    // ONE compaction per chunk: exclusive rank of this thread's survivors in the block (warp scan + warp totals), one slot atomic
    const uint32_t mine = (uint32_t)__popcll(keepmask);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if ((int)lane >= o) incl += up;
    }
    if (lane == 31u) s_cnt[warp] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kRaygenBlock / 32; w++) {
      const uint32_t cw = s_cnt[w];
      before += (w < (int)warp) ? cw : 0u;
      total += cw;
    }
    if (threadIdx.x == 0 && total) s_base = atomicAdd(&ctl->n_gen, total);
    __syncthreads();
    // Pass 2: the survivors again (same counters -> the same rays), straight into their slots.  A thread's survivors are
    // consecutive samples of one pixel and sit side by side in the queue.
    if (keepmask) {
      uint32_t slot = lo + s_base + before + (incl - mine);
      cur = cur0;
      cur_pixel = pixel0;
      for (uint32_t k = 0; k < K; k++) {
        if (k != 0u && advance_item(A, cur)) cur_pixel = item_pixel(A, cur);
        if ((keepmask >> k) & 1ull) {
          const NewPath np = make_camera_path(A, cur_pixel, cur.s_local());
          RT_ASSERT(slot < ctl->pool);
          q.o[slot] = make_float4(np.ray.o.x, np.ray.o.y, np.ray.o.z, np.ray.time);
          q.d[slot] = make_float4(np.ray.d.x, np.ray.d.y, np.ray.d.z, __uint_as_float(kNone));
          q.thr[slot] = make_float4(1.f, 1.f, 1.f, __uint_as_float(np.pixel));
          q.meta[slot] = make_uint2(np.sample_lo, 0u);
          slot++;
        }
      }
    }
    __syncthreads();   // s_cnt / s_base are reused by the next chunk
  }
  for (int o = 16; o > 0; o >>= 1) culled += __shfl_down_sync(0xFFFFFFFFu, culled, o);
  if (lane == 0 && culled) atomicAdd(&ctl->culled, culled);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl->gen_done, 1u) == gridDim.x - 1) {   // last block: publish what this iteration really holds
      __threadfence();
      volatile Ctl* c = ctl;
      c->gen_done = 0;
      if (fill) c->n_live = lo + c->n_gen;
      c->n_gen = 0;
      c->rays += c->culled;          // a culled camera ray is a path segment like any other (Ray.hs:178)
      c->culled = 0;
    }
  }
}

// ---- traversal / closest hit ---------------------------------------------------------------------------
// F = scene feature bits (TF_*): one specialisation per combination that occurs (sphere-only, planar-only,
// mixed; with / without media; a catch-all with moving instances), chosen at launch from FlatScene::features.

// ---- persistent traversal with dynamic ray fetch ----------------------------------------------------------------
// Per-ray work varies by two orders of magnitude (a background ray leaves after the root pair; a ray through a mesh
// walks ~40 nodes, and a medium adds boundary queries), so a static ray→lane map leaves most lanes idle (ncu, round 1:
// 8–13 of 32 threads active per instruction).  Here every lane is a small state machine {query → next query →
// result}; a warp whose idle-lane count reaches kRefill grabs that many new queue slots with ONE atomic
// (ballot/popc ranks), so lanes are replaced as they finish instead of waiting for the slowest ray of the warp.
// Query sequence of one path segment (Geometry.hs:298-330 hoisted, see rt_trace.cuh::trace_segment — same results):
//   surfaces on (tmin, inf) → for each medium: boundary on (tmin, inf) [+ boundary on (t1, inf)] + free-flight sample.

constexpr int kPrimBurst = 8;   // primitives tested per PRIMITIVE step

// Resident-block floor of k_traverse_p (the kernel is bound by memory latency, so warps in flight matter): 9 blocks = 56
// registers = 36 warps / SM where that fits without spills - every specialisation WITHOUT moving instances (pawn / bunny <10>,
// demo1 <1>: 0 bytes spilled; traversal -3 %) - and 8 blocks = 64 registers for the catch-all variants with the per-primitive
// transform (<15>, <31>, <59>: 66 - 120 bytes of spills at 56 registers, demo2 +4 %).  profiles/r02p_sweep_trav_minb.txt; round 1
// found 10 blocks = 48 registers far worse (profiles/r01d_rejected_experiments.txt).
#ifdef RT_TRAV_MINB
constexpr int trav_minb(uint32_t) { return RT_TRAV_MINB; }
#else
constexpr int trav_minb(uint32_t F) { return (F & (TF_MOVING | TF_MOVLIST)) ? 8 : 9; }
#endif
// kCount: instrumented variant (RtOptions.count_work) - counts the node pairs and primitive records this traversal really
// fetches (per lane, warp-reduced once at the end) so that the roofline is computed from EXECUTED work.
template <uint32_t F, bool kCount>
__global__ void __launch_bounds__(kTraverseBlock, trav_minb(F)) k_traverse_p(Ctl* __restrict__ ctl, Queue q, const __grid_constant__ DScene sc,
                                                               const __grid_constant__ RenderArgs A) {
  extern __shared__ uint32_t s_stack[];
  uint32_t w_pairs = 0, w_sph = 0, w_pl = 0, w_box = 0, w_rays = 0;
  auto count_prim = [&](uint32_t i) {   // classify primitive i the way intersect_rec does
    if constexpr (kCount) {
      const uint32_t tf = __float_as_uint(__ldg(sc.prims + 4 * (size_t)i + 3).x);
      if (tf & PRIM_FLAG_BOXHEAD) w_box++;
      else if (tf & PRIM_FLAG_BOXTAIL) {}
      else if ((tf & PRIM_TYPE_MASK) == PRIM_SPHERE) w_sph++;
      else w_pl++;
    }
  };   // (scene BVH depth + 1) x kTraverseBlock words: a deeper static stack would only eat L1
  uint32_t* const stack = s_stack + threadIdx.x;
  constexpr uint32_t stride = kTraverseBlock;
  constexpr bool kMedia = (F & TF_MEDIA) != 0;
  constexpr uint32_t kFull = 0xFFFFFFFFu;
  const uint32_t n = ctl->n_live;
  const uint32_t lane = threadIdx.x & 31u;
  const float tmin0 = 1e-4f;   // Ray.hs:178
  const float inf = __int_as_float(0x7f800000);

  bool active = false, exhausted = false;
  uint32_t slot = 0, cur = kNone, origin_geom = kNone, skip_b = kNone;
  uint32_t pc = 0, pe = 0;        // pending primitive range of the current leaf
  int sp = 0;
  RayT ray{};
  f3 inv{}, oinv{};
  float tmin = tmin0;
  Hit h{inf, kNone, false};      // running result of the current query
  Hit best{inf, kNone, false};   // media scenes: closest event of the segment so far
  uint32_t m = 0;                // media scenes: 0 = surface query, k+1 = medium k
  int pass = 0;
  float t1 = 0.f;

  for (;;) {
    // Every lane is in one of three states, each served by different code:
    //   N  cur is an inner node        → NODE step (fetch the 64-byte child pair, two slab tests, push/pop)
    //   P  pc < pe (rest of a leaf)    → PRIMITIVE step (up to kPrimBurst primitive tests)
    //   F  query finished              → FINISH step (media bookkeeping, result write) followed by the ray fetch
    // Running them "as the lanes come" serialises the three codes at 7–12 of 32 active threads (ncu, round 1), so the
    // warp votes each iteration and runs only ONE kind: F when enough lanes wait for it (or nothing else is left),
    // otherwise the majority of N and P; the other lanes keep their state and join later.  Closest hit is order
    // independent, so results do not change.
    if (active && pc == pe && (cur & kLeafBit) && cur != kNone) {   // leaf reference → primitive range; next node from the stack
      pc = (cur & ~kLeafBit) >> kLeafShift;
      pe = pc + (cur & kLeafCountMask) + 1u;
      RT_ASSERT(pe <= sc.n_prims);
      if (sp > 0) { sp--; cur = stack[sp * stride]; }
      else cur = kNone;
    }
    const bool want_p = active && pc < pe;
    const bool want_n = active && !want_p && cur != kNone;
    const bool want_f = active && !want_p && !want_n;
    const int n_p = __popc(__ballot_sync(kFull, want_p)), n_n = __popc(__ballot_sync(kFull, want_n));
    const int n_f = __popc(__ballot_sync(kFull, want_f));

    if (n_n + n_p != 0 && n_f < A.refill) {
      if (n_n >= n_p) {
        // up to A.node_burst node steps per vote: the vote costs ~20 instructions for every lane, a step ~45 for the
        // lanes that take it; lanes that reach a leaf early sit out the rest of the burst
#pragma unroll 1
        for (int burst = 0; burst < A.node_burst && !(cur & kLeafBit) && want_n; burst++) {
          RT_ASSERT(cur + 1u < sc.n_nodes && (cur & 1u) == 0u);
          const float4* N = sc.nodes + 2 * (size_t)cur;
          const F8 c0 = ldg256(N), c1 = ldg256(N + 2);
          const float4 l0 = c0.a, h0 = c0.b, l1 = c1.a, h1 = c1.b;
          if constexpr (kCount) w_pairs++;
          float tn0, tn1;
          const bool hit0 = box_hit(l0, h0, inv, oinv, tmin, h.t, tn0);
          const bool hit1 = box_hit(l1, h1, inv, oinv, tmin, h.t, tn1);
          const uint32_t r0 = __float_as_uint(l0.w), r1 = __float_as_uint(l1.w);
          if (hit0 && hit1) {
            const bool first0 = tn0 <= tn1;
            RT_ASSERT((uint32_t)sp < sc.stack_entries);
            stack[sp * stride] = first0 ? r1 : r0;
            sp++;
            cur = first0 ? r0 : r1;
          } else if (hit0) cur = r0;
          else if (hit1) cur = r1;
          else if (sp > 0) { sp--; cur = stack[sp * stride]; }
          else cur = kNone;
        }
      } else if (want_p) {
        const uint32_t stop = min(pe, pc + (uint32_t)kPrimBurst);
        while (pc < stop) { count_prim(pc); pc += intersect_prim<F>(sc, pc, ray, tmin, origin_geom, skip_b, h); }   // a cuboid head answers 6 (never past pe)
      }
      continue;
    }

    // ---- FINISH step: next query of the segment, or its result --------------------------------------------------
    if (want_f) {
      if constexpr (!kMedia) {
        RT_ASSERT(slot < n);
        __stcs(A.hits + slot, make_float2(h.front ? h.t : -h.t, __uint_as_float(h.prim)));
        active = false;
      } else {
        bool have_h = true;          // h holds a boundary / surface answer to evaluate
        for (;;) {
          if (have_h) {
            float s1 = 0.f, s2 = -1.f;   // free-flight segment (s1, s2); empty when s2 < s1
            if (m == 0) {
              if (tmin == 0.f && h.prim != kNone && !(h.t > tmin0)) {
                // bounded interior query (fetch, below) met the boundary within tmin of the scatter point: the reference
                // skips that crossing (open interval (1e-4, inf)) and sees the ray as already outside - redo the segment
                // as an ordinary one (once per ~1e4 interior segments)
                tmin = tmin0;
                h = Hit{inf, kNone, false};
                if constexpr ((F & TF_MOVLIST) != 0) scan_moving<F>(sc, ray, tmin, origin_geom, h);
                cur = sc.surface_root;
                break;
              }
              const bool bounded = tmin == 0.f;
              tmin = tmin0;
              best = h;
              // the bounded query found nothing before the free-flight distance: the scatter event
              if (h.prim == kNone && h.t < inf) { best.prim = kMediumBase; best.front = true; m = sc.n_media; }
              // ... or a surface before it: the free-flight sample of the only medium IS the bound, so no scatter event comes
              // first (what the evaluation below would find out with a second Philox draw)
              else if (bounded && h.prim != kNone) m = sc.n_media;
            }
            else if (pass == 0) {
              if (h.prim != kNone) {
                if (!h.front) { s1 = tmin0; s2 = fminf(best.t, h.t); }           // we are inside (Geometry.hs:313)
                else if (h.t < best.t) {                                        // entry before the closest event
                  t1 = h.t;
                  tmin = h.t;
                  skip_b = __float_as_uint(__ldg(sc.prims + 4 * (size_t)h.prim + 3).y);
                  pass = 1;
                  h = Hit{inf, kNone, false};
                  const uint32_t root = sc.media[m - 1].root;
                  if (root & kLeafBit) {    // boundary is a short primitive list: answer the query right here
                    const uint32_t first = (root & ~kLeafBit) >> kLeafShift, count = (root & kLeafCountMask) + 1u;
                    for (uint32_t k = 0; k < count;) { count_prim(first + k); k += intersect_prim<F>(sc, first + k, ray, tmin, origin_geom, skip_b, h); }
                    continue;
                  }
                  cur = root;
                  break;                                                        // → second boundary query on (t1, inf)
                }
              }
            } else if (h.prim != kNone) { s1 = t1; s2 = fminf(best.t, h.t); }
            if (s2 >= s1) {
              const DMedium& md = sc.media[m - 1];
              const uint32_t pixel = __float_as_uint(__ldcs(q.thr + slot).w);
              const uint2 meta = __ldcs(q.meta + slot);
              const unsigned long long sample = (A.first_sample & 0xFFFFFFFF00000000ull) | meta.x;
              const uint32_t k = m - 1;
              const uint4 rnd = rng4(A.key, pixel, sample, meta.y, RNG_MEDIA + (k >> 2));
              const uint32_t rbits = (k & 3u) == 0 ? rnd.x : ((k & 3u) == 1 ? rnd.y : ((k & 3u) == 2 ? rnd.z : rnd.w));
              const float hit_dist = md.neg_inv_density * logf(u01(rbits));      // rand = 0 → +inf → no hit
              if (hit_dist < s2 - s1) { best.t = s1 + hit_dist; best.prim = kMediumBase + k; best.front = true; }
            }
            m++;
          }
          if (m > sc.n_media) {           // all media done: the segment's closest event
            __stcs(A.hits + slot, make_float2(best.front ? best.t : -best.t, __uint_as_float(best.prim)));
            active = false;
            break;
          }
          const DMedium& md = sc.media[m - 1];   // set up this medium's first boundary query
          pass = 0;
          have_h = true;
          if (md.shared && (best.prim == kNone || best.prim < kMediumBase)) {
            // boundary ⊆ surfaces: the surface answer is the boundary answer when it lies on the boundary, and a
            // surface miss is a boundary miss
            if (best.prim == kNone) { h = best; continue; }
            const uint32_t g = __float_as_uint(__ldg(sc.prims + 4 * (size_t)best.prim + 3).y);
            if (g >= md.geom_lo && g <= md.geom_hi) { h = best; continue; }
          }
          tmin = tmin0;
          skip_b = kNone;
          h = Hit{inf, kNone, false};
          if (md.root & kLeafBit) {
            const uint32_t first = (md.root & ~kLeafBit) >> kLeafShift, count = (md.root & kLeafCountMask) + 1u;
            for (uint32_t k = 0; k < count;) { count_prim(first + k); k += intersect_prim<F>(sc, first + k, ray, tmin, origin_geom, skip_b, h); }
            continue;
          }
          cur = md.root;
          break;
        }
        sp = 0;
      }
    }
    // ---- ray fetch: one atomic per warp hands a queue slot to every lane without a ray ----------------------------
    if (!exhausted) {
      const uint32_t idle = __ballot_sync(kFull, !active);
      if (idle) {
        const int want = __popc(idle), leader = __ffs(idle) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&ctl->trav_next, (unsigned int)want);
        base = __shfl_sync(kFull, base, leader);
        exhausted = base + (uint32_t)want >= n;
        if (!active) {
          const uint32_t i = base + __popc(idle & ((1u << lane) - 1u));
          if (i < n) {
            const float4 o4 = __ldcs(q.o + i), d4 = __ldcs(q.d + i);   // streaming: queue records must not evict the BVH from L1
            float bound_i = inf;
            if constexpr (kMedia) {                                    // (issued with the two loads above; meaningful for interior segments only)
              if (sc.n_media == 1u && sc.media[0].closed != 0u) bound_i = __ldcs(A.bound + i);
            }
            ray = RayT{mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), o4.w};
            origin_geom = __float_as_uint(d4.w);
            inv.x = fabsf(ray.d.x) > 1e-30f ? __frcp_rn(ray.d.x) : copysignf(1e30f, ray.d.x);
            inv.y = fabsf(ray.d.y) > 1e-30f ? __frcp_rn(ray.d.y) : copysignf(1e30f, ray.d.y);
            inv.z = fabsf(ray.d.z) > 1e-30f ? __frcp_rn(ray.d.z) : copysignf(1e30f, ray.d.z);
            oinv = ray.o * inv;
            slot = i;
            active = true;
            if constexpr (kCount) w_rays++;
            cur = sc.surface_root;
            sp = 0;
            tmin = tmin0;
            skip_b = kNone;
            h = Hit{inf, kNone, false};
            m = 0;
            if constexpr (kMedia) {
              // interior segment of a subsurface-scattering mesh: the free-flight distance bounds the surface query
              // (rt_trace.cuh::starts_inside_only_medium)
              const bool inside = starts_inside_only_medium(sc, origin_geom);
              origin_geom = strip_inside_flag(origin_geom);
              if (inside) {
                // free_flight_bound of this segment, drawn by the kernel that queued it (RenderArgs::bound)
                h.t = bound_i;
                // the origin is a scatter point or a surface the ray never re-hits (skip_a): look at (0, bound] so that a
                // boundary crossing inside (0, tmin0] is noticed
                if (h.t < inf) tmin = 0.f;
              }
            }
            if constexpr ((F & TF_MOVLIST) != 0) {   // the moving primitives, outside the tree: once per segment
              if constexpr (kCount) for (uint32_t k = 0; k < sc.moving_count; k++) count_prim(sc.moving_first + k);
              scan_moving<F>(sc, ray, tmin, origin_geom, h);
            }
          }
        }
      }
    }
    if (__ballot_sync(kFull, active) == 0u) break;
  }
  if constexpr (kCount) {
    for (int o = 16; o > 0; o >>= 1) {
      w_pairs += __shfl_down_sync(kFull, w_pairs, o); w_sph += __shfl_down_sync(kFull, w_sph, o);
      w_pl += __shfl_down_sync(kFull, w_pl, o); w_box += __shfl_down_sync(kFull, w_box, o); w_rays += __shfl_down_sync(kFull, w_rays, o);
    }
    if (lane == 0) {
      atomicAdd(&ctl->work[0], (unsigned long long)w_pairs); atomicAdd(&ctl->work[1], (unsigned long long)w_sph);
      atomicAdd(&ctl->work[2], (unsigned long long)w_pl); atomicAdd(&ctl->work[3], (unsigned long long)w_box);
      atomicAdd(&ctl->work[4], (unsigned long long)w_rays);
    }
  }
}


// ---- flat-list traversal ------------------------------------------------------------------------------------------
// Scenes whose whole surface geometry is one short list (the reference's `group` of a few primitives: README scene,
// Cornell box) need no tree and no stack: every lane tests the same primitive at the same time, so the list travels in
// the KERNEL PARAMETER block and, with the loop unrolled, each record field is an immediate constant-bank operand of
// the FFMA that uses it — zero loads, zero divergence between lanes.
// Aggregated cuboids (rt_device.h) travel in a list of their own: their six faces are left out of `rec`, and one slab
// test per box runs after the face scan (with the full tie rule, so the result is that of a scan in index order).
constexpr int kFlatBoxes = 4;
struct FlatList {
  uint32_t n, n_box, _pad[2];
  uint32_t index[kLeafCap];         // primitive index of rec[k] (increasing)
  float4 rec[kLeafCap][4];
  float4 box[kFlatBoxes][4];        // box records
  uint32_t box_first[kFlatBoxes];   // primitive index / geometry id of face 0
  uint32_t box_geom[kFlatBoxes];
};

// closest hit over the whole list: unrolled, every record field a constant-bank operand
template <uint32_t F>
__device__ __forceinline__ Hit scan_flat(const DScene& sc, const FlatList& L, const RayT& ray, uint32_t origin_geom) {
  Hit h{__int_as_float(0x7f800000), kNone, false};
#pragma unroll
  for (int k = 0; k < (int)kLeafCap; k++) {
    if (k >= (int)L.n) break;
    intersect_rec<F, ParamRec, true>(sc, ParamRec{L.rec[k]}, L.index[k], ray, 1e-4f, origin_geom, kNone, h);
  }
  if constexpr ((F & TF_PLANAR) != 0) {
#pragma unroll
    for (int k = 0; k < kFlatBoxes; k++) {
      if (k >= (int)L.n_box) break;
      intersect_box(sc.ref_order, L.box[k][0], L.box[k][1], L.box[k][2], L.box[k][3], L.box_first[k], L.box_geom[k], ray.o, ray.d, 1e-4f,
                    origin_geom, kNone, false, h);
    }
  }
  return h;
}
using TraverseFn = void (*)(Ctl*, Queue, const DScene, const RenderArgs);
template <template <uint32_t> class K, uint32_t B>   // B = TF_BOX or 0
static TraverseFn pick_planar_variant(uint32_t features) {
  if (features & TF_MOVING) return K<TF_SPHERE | TF_PLANAR | TF_MOVING | TF_MEDIA | B>::fn();
  switch (features & (TF_SPHERE | TF_PLANAR | TF_MEDIA)) {
    case TF_PLANAR: return K<TF_PLANAR | B>::fn();
    case TF_SPHERE | TF_PLANAR: return K<TF_SPHERE | TF_PLANAR | B>::fn();
    case TF_PLANAR | TF_MEDIA: return K<TF_PLANAR | TF_MEDIA | B>::fn();
    case TF_SPHERE | TF_PLANAR | TF_MEDIA: return K<TF_SPHERE | TF_PLANAR | TF_MEDIA | B>::fn();
    default: return K<TF_SPHERE | TF_PLANAR | TF_MOVING | TF_MEDIA | B>::fn();
  }
}
template <template <uint32_t> class K>
static TraverseFn pick_by_features(uint32_t features) {
  if (features & TF_MOVLIST) return K<TF_SPHERE | TF_PLANAR | TF_MEDIA | TF_BOX | TF_MOVLIST>::fn();   // static tree + short moving list
  if (features & TF_BOX) return pick_planar_variant<K, TF_BOX>(features);   // aggregated cuboids present (implies planar)
  if (!(features & TF_MOVING)) {
    if ((features & (TF_SPHERE | TF_PLANAR | TF_MEDIA)) == TF_SPHERE) return K<TF_SPHERE>::fn();
    if ((features & (TF_SPHERE | TF_PLANAR | TF_MEDIA)) == (TF_SPHERE | TF_MEDIA)) return K<TF_SPHERE | TF_MEDIA>::fn();
  }
  return pick_planar_variant<K, 0u>(features);
}
template <uint32_t F> struct PersistentK { static TraverseFn fn() { return k_traverse_p<F, false>; } };
template <uint32_t F> struct CountedK { static TraverseFn fn() { return k_traverse_p<F, true>; } };   // RtOptions.count_work

// ---- shading, accumulation, compaction -------------------------------------------------------------------
// ---- TMA (bulk async copy) + mbarrier helpers: the shading kernel prefetches its next tile of queue records into
// shared memory while it shades the current one (cp.async.bulk → UBLKCP in SASS) ---------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
  }
}

// Shading of ONE path segment whose closest hit is known (Ray.hs:176-198): material / texture / redirect mixture,
// radiance into the accumulators, depth limit.  Updates ray / thr / origin_geom / meta for the next segment and returns
// whether the path continues.  Shared by k_shade, k_trace_regions and k_finish: same RNG counters, same results.
// HOOK: materials draw the no-redirection direction from their own generator (RtOptions.material_sampling) - a compile-time
// switch, so that the default kernels carry none of its code (as a run-time flag it cost demo2's k_shade 7 %).
template <bool HOOK = false>
__device__ __forceinline__ bool shade_path(const DScene& sc, const RenderArgs& A, const NoiseTab& T, RayT& ray, const Hit& hit, f3& thr,
                                           uint32_t& origin_geom, uint32_t pixel, uint2& meta) {
  const unsigned long long sample = (A.first_sample & 0xFFFFFFFF00000000ull) | meta.x;
  const uint4 rnd = rng4(A.key, pixel, sample, meta.y, RNG_SHADE);
  f3 contrib;
  RT_ASSERT(pixel < A.n_pixels);
  bool alive = shade_bounce<HOOK>(sc, A.cam, T, ray, hit, thr, origin_geom, rnd, contrib);
  // a non-finite weight (pdf underflow at a grazing light direction) must not reach the accumulators: one NaN would own
  // the pixel for the whole render
  if (!(fabsf(thr.x + thr.y + thr.z) < 3.0e38f)) { alive = false; thr = mk3(0.f, 0.f, 0.f); }
  if (!(fabsf(contrib.x + contrib.y + contrib.z) < 3.0e38f)) contrib = mk3(0.f, 0.f, 0.f);
  if (contrib.x != 0.f || contrib.y != 0.f || contrib.z != 0.f) {
    float* s = A.sum + 3 * (size_t)pixel;
    atomicAdd(s, contrib.x);
    atomicAdd(s + 1, contrib.y);
    atomicAdd(s + 2, contrib.z);
    if (A.sumsq) {
      float* s2 = A.sumsq + 3 * (size_t)pixel;
      atomicAdd(s2, contrib.x * contrib.x);
      atomicAdd(s2 + 1, contrib.y * contrib.y);
      atomicAdd(s2 + 2, contrib.z * contrib.z);
    }
  }
  meta.y += 1;
  if (meta.y >= (uint32_t)A.cam.max_depth) alive = false;             // rayColor 0 = zero (Ray.hs:176)
  if (thr.x == 0.f && thr.y == 0.f && thr.z == 0.f) alive = false;     // nothing downstream can contribute
  return alive;
}

// One block-iteration = kShadeBlock consecutive queue slots.  Slot reservation is aggregated per BLOCK (warp
// ballot/popc ranks → shared-memory warp counts → ONE global atomic): the per-warp atomic of round 1 spent half of the
// kernel waiting on a single contended address (ncu: 49 % of stall samples on the SHFL after it).
// Survivors are compacted into the next queue; the atomic hands out dense slots.  (Replacing finished paths by new camera
// paths in their own slot — no compaction — was measured and dropped: it mixes fresh camera rays with bounced rays inside
// warps, traversal +15 %, shading +20 %; profiles/r01b.)
//
// FF != 0 (flat scenes, see k_trace_flat below): the closest-hit query runs HERE, over the constant-bank primitive list,
// instead of being read from `hits`, and with GEN the camera rays of the refilled tail are generated here too — the whole
// iteration is one launch, and new paths never touch the queue before their first bounce.
template <uint32_t FF, bool GEN, bool BOUND = false, bool HOOK = false>   // BOUND: survivors that start inside the closed medium get their free-flight bound (RenderArgs::bound)
__device__ __forceinline__ void shade_tiles(Ctl* __restrict__ ctl, const Queue& q, const Queue& qn, const DScene& sc, const RenderArgs& A,
                                            const FlatList* L) {
  __shared__ float s_grad[256 * 3];
  __shared__ uint8_t s_perm[3 * 256];
  __shared__ uint32_t s_cnt[2][kShadeBlock / 32];
  __shared__ uint32_t s_base[2];
  __shared__ uint32_t s_bin[FF ? 1 : 64];        // material-coherent shading: class histogram and permutation of the tile
  __shared__ uint8_t s_order[FF ? 1 : kShadeBlock];
  // double-buffered tile of queue records, filled by TMA one iteration ahead (16 KB per stage)
  __shared__ __align__(128) float4 s_o[2][kShadeBlock];
  __shared__ __align__(128) float4 s_d[2][kShadeBlock];
  __shared__ __align__(128) float4 s_t[2][kShadeBlock];
  __shared__ __align__(128) uint2 s_m[2][kShadeBlock];
  __shared__ __align__(128) float2 s_h[2][FF ? 1 : kShadeBlock];
  __shared__ __align__(8) uint64_t s_bar[2];
  constexpr uint32_t kTileBytes = kShadeBlock * (16 + 16 + 16 + 8 + (FF ? 0 : 8));
  auto issue_tile = [&](int st, uint32_t base) {   // one thread: arm the barrier, launch the bulk copies
    mbar_expect_tx(&s_bar[st], kTileBytes);
    bulk_g2s(s_o[st], q.o + base, kShadeBlock * 16, &s_bar[st]);
    bulk_g2s(s_d[st], q.d + base, kShadeBlock * 16, &s_bar[st]);
    bulk_g2s(s_t[st], q.thr + base, kShadeBlock * 16, &s_bar[st]);
    bulk_g2s(s_m[st], q.meta + base, kShadeBlock * 8, &s_bar[st]);
    if (!FF) bulk_g2s(s_h[st], A.hits + base, kShadeBlock * 8, &s_bar[st]);
  };
  if (threadIdx.x == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (sc.has_noise) {      // Perlin tables staged in shared memory (Noise.hs:52-92)
    for (int k = threadIdx.x; k < 768; k += blockDim.x) {
      s_grad[k] = sc.noise_grad[k];
      s_perm[k] = sc.noise_perm[k];
    }
    __syncthreads();
  }
  const NoiseTab T{s_grad, s_perm};
  const uint32_t n = ctl->n_live;
  // queue slots below n_stored hold paths written by the previous iteration; with GEN the refilled tail
  // [fill_lo, n_live) is generated here (its tiles are a suffix of every block's tile sequence: no loads for them)
  const uint32_t n_stored = GEN ? ctl->fill_lo : n;
  const unsigned long long item0 = GEN ? ctl->fill_item0 : 0ull;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  int par = 0;
  const uint32_t stride = gridDim.x * kShadeBlock;
  if (threadIdx.x == 0 && blockIdx.x * kShadeBlock < n_stored && blockIdx.x * kShadeBlock < n) issue_tile(0, blockIdx.x * kShadeBlock);
  uint32_t tile = 0;
  for (uint32_t base = blockIdx.x * kShadeBlock; base < n; base += stride, par ^= 1, tile++) {
    const uint32_t i = base + threadIdx.x;
    // prefetch the NEXT tile (tiles never run past the allocation: the pool is a multiple of 1024 slots); the stage it
    // lands in was last read one iteration ago, before that iteration's __syncthreads
    if (threadIdx.x == 0 && base + stride < n_stored && base + stride < n) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue_tile(par ^ 1, base + stride);
    }
    if (base < n_stored) mbar_wait(&s_bar[par], (tile >> 1) & 1u);
    bool alive = false;
    RayT ray{};
    f3 thr = mk3(0.f, 0.f, 0.f);
    uint32_t origin_geom = kNone, pixel = 0;
    uint2 meta = make_uint2(0, 0);
    // Material-coherent shading (the reference author's own note, Material.hs:36-38): the 256 paths of the tile
    // are re-dealt to the threads by MATERIAL CLASS (miss / material kind / texture family - a counting sort through shared
    // memory, where the tile already sits), so that a warp shades one kind of surface instead of serialising ten.  Only the
    // thread -> path assignment changes: same paths, same RNG counters, same results.
    uint32_t src = threadIdx.x;         // which path of the tile this thread shades
    if constexpr (FF == 0) {
      if (A.shade_sort) {
        uint32_t key = 63u;               // slots past the end of the queue sort last
        if (i < n) {
          const uint32_t prim = __float_as_uint(s_h[par][threadIdx.x].y);
          key = 0u;                       // miss: background
          if (prim != kNone) {
            const uint32_t m = prim >= kMediumBase ? sc.media[prim - kMediumBase].material : __float_as_uint(__ldg(sc.prims + 4 * (size_t)prim + 3).w);
            const DMaterial mat = sc.materials[m];
            key = (uint32_t)mat.kind;     // 1 .. 10
            if (mat.kind != RT_MAT_PITCH_BLACK && mat.kind != RT_MAT_DIELECTRIC) {
              const int tk = sc.textures[mat.texture].kind;
              if (tk == RT_TEX_NOISE || tk == RT_TEX_MARBLE) key += 16u;          // Perlin octaves: by far the longest shader
              else if (tk == RT_TEX_IMAGE || tk == RT_TEX_CHECKER) key += 32u;     // needs uv (atan2 / acos on spheres)
            }
          }
        }
        if (threadIdx.x < 64) s_bin[threadIdx.x] = 0u;
        __syncthreads();
        const uint32_t rank = atomicAdd(&s_bin[key], 1u);
        __syncthreads();
        if (warp == 0) {                  // exclusive prefix over the 64 bins: two bins per lane, one shuffle scan
          const uint32_t v0 = s_bin[2 * lane], v1 = s_bin[2 * lane + 1];
          uint32_t incl = v0 + v1;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((int)lane >= o) incl += up;
          }
          s_bin[2 * lane] = incl - v0 - v1;
          s_bin[2 * lane + 1] = incl - v1;
        }
        __syncthreads();
        s_order[s_bin[key] + rank] = (uint8_t)threadIdx.x;
        __syncthreads();
        src = s_order[threadIdx.x];
      }
    }
    const uint32_t i_src = base + src;
    if (i_src < n) {
      if (GEN && i_src >= n_stored) {   // Ray.hs:229-231 for work item fill_item0 + (slot - fill_lo), as k_raygen numbers them
        const NewPath np = new_camera_path(A, item0 + (i_src - n_stored));
        ray = np.ray;
        thr = mk3(1.f, 1.f, 1.f);
        pixel = np.pixel;
        meta = make_uint2(np.sample_lo, 0u);
      } else {
        const float4 o4 = s_o[par][src], d4 = s_d[par][src], t4 = s_t[par][src];
        meta = s_m[par][src];
        ray = RayT{mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), o4.w};
        thr = mk3(t4.x, t4.y, t4.z);
        pixel = __float_as_uint(t4.w);
        origin_geom = __float_as_uint(d4.w);
      }
      Hit hit;
      if (FF) {   // closest hit over the constant-bank list
        hit = scan_flat<FF ? FF : TF_SPHERE>(sc, *L, ray, origin_geom);
      } else {
        const float2 h2 = s_h[par][src];
        hit = Hit{fabsf(h2.x), __float_as_uint(h2.y), !(h2.x < 0.f)};
      }
      alive = shade_path<HOOK>(sc, A, T, ray, hit, thr, origin_geom, pixel, meta);
    }
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, alive);
    if (lane == 0) s_cnt[par][warp] = __popc(mask);
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kShadeBlock / 32; w++) {
      const uint32_t c = s_cnt[par][w];
      before += (w < (int)warp) ? c : 0u;
      total += c;
    }
    if (threadIdx.x == 0 && total) s_base[par] = atomicAdd(&ctl->n_next, total);
    __syncthreads();
    if (alive) {
      const uint32_t slot = s_base[par] + before + __popc(mask & ((1u << lane) - 1u));
      RT_ASSERT(slot < ctl->pool);
      qn.o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
      qn.d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(origin_geom));
      qn.thr[slot] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(pixel));
      qn.meta[slot] = meta;
      if constexpr (BOUND) {     // the next segment starts inside the closed medium: its free-flight bound, for the traversal's fetch
        if (starts_inside_only_medium(sc, origin_geom))
          A.bound[slot] = free_flight_bound(sc, 1e-4f, A.key, pixel, (A.first_sample & 0xFFFFFFFF00000000ull) | meta.x, meta.y);
      }
    }
  }
  // the last block to finish closes the iteration and prepares the next (no separate launch, no host round trip)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl->blocks_done, 1u) == gridDim.x - 1) {
      __threadfence();
      ctl->blocks_done = 0;
      prepare_iteration(ctl, false);
    }
  }
}

// BOUND = the scene's only medium has a closed boundary shared with the surfaces (interior segments are bounded queries)
template <bool BOUND, bool HOOK = false>
__global__ void __launch_bounds__(kShadeBlock, RT_SHADE_MINB) k_shade(Ctl* __restrict__ ctl, Queue q, Queue qn, const __grid_constant__ DScene sc,
                                                                      const __grid_constant__ RenderArgs A) {
  shade_tiles<0, false, BOUND, HOOK>(ctl, q, qn, sc, A, nullptr);
}

// Flat scenes (no BVH, no media, no moving instances): raygen + closest hit + shading + compaction of one wavefront
// iteration in ONE launch.  The traversal part is issue-bound (constant-bank FFMAs), the shading part waits on HBM and
// atomics: in one kernel the warp schedulers overlap the two, and the `hits` round trip (16 B per ray), the second read
// of origin/direction (32 B) and the first write + read of every new camera path (112 B) disappear.
#ifndef RT_TRACE_FLAT_MINB
#define RT_TRACE_FLAT_MINB 4
#endif
template <uint32_t F>
__global__ void __launch_bounds__(kShadeBlock, RT_TRACE_FLAT_MINB) k_trace_flat(Ctl* __restrict__ ctl, Queue q, Queue qn, const __grid_constant__ DScene sc,
                                                                                const __grid_constant__ RenderArgs A,
                                                                                const __grid_constant__ FlatList L) {
  shade_tiles<F, true>(ctl, q, qn, sc, A, &L);
}

using TraceFlatFn = void (*)(Ctl*, Queue, Queue, const DScene, const RenderArgs, const FlatList);
static TraceFlatFn pick_trace_flat(uint32_t features) {
  if (features & (TF_MOVING | TF_MEDIA | TF_MOVLIST)) return nullptr;
  switch (features & (TF_SPHERE | TF_PLANAR)) {
    case TF_SPHERE: return k_trace_flat<TF_SPHERE>;
    case TF_PLANAR: return k_trace_flat<TF_PLANAR>;
    case TF_SPHERE | TF_PLANAR: return k_trace_flat<TF_SPHERE | TF_PLANAR>;
    default: return nullptr;
  }
}

// Initialises the control block ON the device (no host-to-device copy whose pinned source a later, still queued render
// of the same workspace could overwrite) and prepares iteration 0.
__global__ void k_begin(Ctl* ctl, unsigned long long total_items, unsigned int pool) {
  Ctl c{};
  c.total_items = total_items;
  c.pool = pool;
  *ctl = c;
  prepare_iteration(ctl, true);
}

// ---- tail kernel ------------------------------------------------------------------------------------------------
// Once every camera path has been started and only a few long paths survive (depth up to 50: dozens of nearly empty
// iterations, three launches each), one launch finishes them: each thread carries ONE path through its remaining
// segments (closest hit + shading in a loop, no queue traffic).  Same device functions, same RNG counters → the same
// samples as the wavefront iterations it replaces.  No-op unless the work items are exhausted and n_live <= A.finish_max.
// FF != 0: flat scene, closest hit by the constant-bank list scan of k_trace_flat (no BVH, no stack); else trace_segment<F>.
template <uint32_t F, uint32_t FF>
__device__ __forceinline__ void finish_paths(Ctl* __restrict__ ctl, const Queue& q, const DScene& sc, const RenderArgs& A, const FlatList* L) {
  extern __shared__ uint32_t s_stack[];
  __shared__ float s_grad[256 * 3];
  __shared__ uint8_t s_perm[3 * 256];
  const uint32_t n = ctl->n_live;
  // not yet: work items left, too many paths, or the prepared iteration still has camera rays to generate in
  // [fill_lo, n_live) (those slots hold no paths until k_raygen / k_trace_flat has run)
  if (n == 0 || n > A.finish_max || ctl->next_item < ctl->total_items || ctl->fill_lo != n) return;
  if (sc.has_noise) {
    for (int k = threadIdx.x; k < 768; k += blockDim.x) { s_grad[k] = sc.noise_grad[k]; s_perm[k] = sc.noise_perm[k]; }
    __syncthreads();
  }
  const NoiseTab T{s_grad, s_perm};
  uint32_t* stack = s_stack + threadIdx.x;
  unsigned int segments = 0;
  // Every lane is a path carrier: when its path ends it takes the next unclaimed queue slot instead of waiting for the
  // longest path of the warp (claims are warp-aggregated: one atomic on ctl->trav_next, which prepare_iteration left at 0).
  constexpr uint32_t kFull = 0xFFFFFFFFu;
  const uint32_t lane = threadIdx.x & 31u;
  bool have = false, more = true;   // more (warp-uniform): the queue may still hold unclaimed paths
  RayT ray{};
  f3 thr = mk3(0.f, 0.f, 0.f);
  uint2 meta = make_uint2(0u, 0u);
  uint32_t pixel = 0, origin_geom = kNone;
  unsigned long long sample = 0;
  for (;;) {
    if (more) {
      const unsigned int need = __ballot_sync(kFull, !have);
      if (__popc(need) >= A.finish_refill || need == kFull) {
        const int leader = __ffs(need) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&ctl->trav_next, (unsigned int)__popc(need));
        base = __shfl_sync(kFull, base, leader);
        if (!have) {
          const uint32_t i = base + __popc(need & ((1u << lane) - 1u));
          if (i < n) {
            const float4 o4 = q.o[i], d4 = q.d[i], t4 = q.thr[i];
            meta = q.meta[i];
            ray = RayT{mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), o4.w};
            thr = mk3(t4.x, t4.y, t4.z);
            pixel = __float_as_uint(t4.w);
            origin_geom = __float_as_uint(d4.w);
            sample = (A.first_sample & 0xFFFFFFFF00000000ull) | meta.x;
            have = true;
          }
        }
        more = base + (uint32_t)__popc(need) < n;
      }
    }
    if (!__any_sync(kFull, have)) break;
    if (have) {
      Hit hit;
      if (FF) hit = scan_flat<FF ? FF : TF_SPHERE>(sc, *L, ray, origin_geom);
      else hit = trace_segment<F>(sc, ray, origin_geom, true, A.key, pixel, sample, meta.y, stack, kTraverseBlock);
      segments++;
      have = shade_path(sc, A, T, ray, hit, thr, origin_geom, pixel, meta);
    }
  }
  // bookkeeping: segments traced, and (last block) the render is complete
  for (int o = 16; o > 0; o >>= 1) segments += __shfl_down_sync(0xFFFFFFFFu, segments, o);
  if ((threadIdx.x & 31) == 0 && segments) { atomicAdd(&ctl->rays, (unsigned long long)segments); atomicAdd(&ctl->rays_finish, (unsigned long long)segments); }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl->blocks_done, 1u) == gridDim.x - 1) {
      ctl->blocks_done = 0;
      ctl->paths_finish += n;
      ctl->n_live = 0;
      ctl->n_cur = 0;
      ctl->fill_lo = 0;          // a launch queued behind the end of the render must see an EMPTY iteration (no stale tail range)
      ctl->fill_item0 = ctl->next_item;
      ctl->iterations += 1;
    }
  }
}
template <uint32_t F>
__global__ void __launch_bounds__(kTraverseBlock) k_finish(Ctl* __restrict__ ctl, Queue q, const __grid_constant__ DScene sc,
                                                           const __grid_constant__ RenderArgs A) {
  finish_paths<F, 0>(ctl, q, sc, A, nullptr);
}
template <uint32_t F> struct FinishK { static TraverseFn fn() { return k_finish<F>; } };

template <uint32_t FF>
__global__ void __launch_bounds__(kTraverseBlock, 8) k_finish_flat(Ctl* __restrict__ ctl, Queue q, const __grid_constant__ DScene sc,
                                                                   const __grid_constant__ RenderArgs A, const __grid_constant__ FlatList L) {
  finish_paths<FF, FF>(ctl, q, sc, A, &L);
}
using FinishFlatFn = void (*)(Ctl*, Queue, const DScene, const RenderArgs, const FlatList);
static FinishFlatFn pick_finish_flat(uint32_t features) {
  switch (features & (TF_SPHERE | TF_PLANAR)) {
    case TF_SPHERE: return k_finish_flat<TF_SPHERE>;
    case TF_PLANAR: return k_finish_flat<TF_PLANAR>;
    case TF_SPHERE | TF_PLANAR: return k_finish_flat<TF_SPHERE | TF_PLANAR>;
    default: return nullptr;
  }
}

// ---- deterministic primary-ray probe (parity gate: primitive id + t) --------------------------------------
__global__ void __launch_bounds__(kTraverseBlock) k_probe(const __grid_constant__ DScene sc, const __grid_constant__ DCamera cam, float time,
                                                          uint32_t* __restrict__ ids, float* __restrict__ ts) {
  extern __shared__ uint32_t s_stack[];   // (scene BVH depth + 1) x kTraverseBlock words: a deeper static stack would only eat L1
  const uint32_t n = (uint32_t)cam.W * (uint32_t)cam.H;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    DCamera const& c = cam;
    uint32_t j = p / (uint32_t)c.W, i = p - j * (uint32_t)c.W;
    RayT ray;
    ray.time = time;
    ray.o = ld3(c.center);
    ray.d = normalize3(fma3((float)i + 0.5f, ld3(c.pixel_u), fma3((float)j + 0.5f, ld3(c.pixel_v), ld3(c.top_left))));
    Hit h = trace_segment<TF_SPHERE | TF_PLANAR | TF_MOVING | TF_BOX | TF_MOVLIST>(sc, ray, kNone, false, make_uint2(0, 0), 0, 0, 0, s_stack + threadIdx.x, kTraverseBlock);
    ids[p] = h.prim == kNone ? kNone : sc.canonical[h.prim];
    ts[p] = h.t;
  }
}

// ---- deterministic surface probe: what the shader sees at the pixel-centre primary hit ------------------------------
__global__ void __launch_bounds__(kTraverseBlock) k_probe_surface(const __grid_constant__ DScene sc, const __grid_constant__ DCamera cam,
                                                                  float time, float* __restrict__ out) {
  extern __shared__ uint32_t s_stack[];
  __shared__ float s_grad[256 * 3];
  __shared__ uint8_t s_perm[3 * 256];
  for (int k = threadIdx.x; k < 768; k += blockDim.x) { s_grad[k] = sc.noise_grad[k]; s_perm[k] = sc.noise_perm[k]; }
  __syncthreads();
  const NoiseTab T{s_grad, s_perm};
  const uint32_t n = (uint32_t)cam.W * (uint32_t)cam.H;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    uint32_t j = p / (uint32_t)cam.W, i = p - j * (uint32_t)cam.W;
    RayT ray;
    ray.time = time;
    ray.o = ld3(cam.center);
    ray.d = normalize3(fma3((float)i + 0.5f, ld3(cam.pixel_u), fma3((float)j + 0.5f, ld3(cam.pixel_v), ld3(cam.top_left))));
    Hit h = trace_segment<TF_SPHERE | TF_PLANAR | TF_MOVING | TF_BOX | TF_MOVLIST>(sc, ray, kNone, false, make_uint2(0, 0), 0, 0, 0, s_stack + threadIdx.x, kTraverseBlock);
    float* o = out + 10 * (size_t)p;
    for (int k = 0; k < 10; k++) o[k] = 0.f;
    if (h.prim == kNone) continue;
    const Surf s = reconstruct(sc, ray, h);
    const DMaterial mat = sc.materials[s.material];
    float u, v;
    surf_uv(sc, s, u, v);
    o[0] = s.n.x; o[1] = s.n.y; o[2] = s.n.z; o[3] = u; o[4] = v; o[5] = s.front ? 1.f : 0.f;
    if (mat.kind != RT_MAT_PITCH_BLACK && mat.kind != RT_MAT_DIELECTRIC) {
      const f3 c = eval_texture(sc, T, mat.texture, s);
      o[6] = c.x; o[7] = c.y; o[8] = c.z;
    }
    o[9] = (float)mat.kind;
  }
}

// ---- image output step (writeImage / writeImageSqrt, Ray.hs:248-260) -----------------------------------------
__global__ void k_tonemap(const float* __restrict__ rgb, int n, int mode, uint8_t* __restrict__ out) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    float x = rgb[k];
    x = fminf(1.f, fmaxf(0.f, x));     // NaN → 0
    float y = mode == 1 ? sqrtf(x) : (x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 1.f / 2.4f) - 0.055f);
    // the reference's encoder (massiv-io / Color `toWord8`) quantises as floor(256 y) clamped to 255: pinned against
    // pawn_demo.png's background gradient (99.7 % of 208k pixels exact; round(255 y) matches only 81 %)
    out[k] = (uint8_t)min(255, (int)floorf(y * 256.f));
  }
}

// ---- image reduction over NVLink peer memory (sample-sharded renders) -------------------------------------
// Every rank's un-normalised sum buffer lives in symmetric memory.  With NVSwitch multicast the root reads the image ONCE
// through the multicast address: multimem.ld_reduce makes the switch fetch the element from every GPU and add them in the
// fabric (NVLS), so the root ingests W*H*3 floats instead of (G-1) times that, and the 1/spp normalisation (Ray.hs:232)
// is fused into the same pass.  Without multicast the root sums the peers' buffers with plain peer loads.
__device__ __forceinline__ float4 multimem_sum4(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__global__ void __launch_bounds__(256) k_reduce_multimem(const float4* mc, float4* __restrict__ out, size_t n4, float scale) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += (size_t)gridDim.x * blockDim.x) {
    const float4 v = multimem_sum4(mc + k);
    out[k] = make_float4(v.x * scale, v.y * scale, v.z * scale, v.w * scale);
  }
}
constexpr int kMaxPeers = 16;
struct PeerPtrs { const float4* p[kMaxPeers]; };
__global__ void __launch_bounds__(256) k_reduce_peers(const __grid_constant__ PeerPtrs peers, int n_peers, float4* __restrict__ out, size_t n4, float scale) {
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += (size_t)gridDim.x * blockDim.x) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int r = 0; r < n_peers; r++) {
      const float4 v = __ldcv(peers.p[r] + k);     // volatile-cached load: peer memory written by another GPU
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    out[k] = make_float4(a.x * scale, a.y * scale, a.z * scale, a.w * scale);
  }
}

__global__ void k_scale(const float* __restrict__ in, float* __restrict__ out, int n, float s) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) out[k] = in[k] * s;
}

// =================================================================================================================
// host side
// =================================================================================================================
struct DeviceScene {              // one committed scene resident on one device: a single arena allocation
  int device = 0;
  int sm_count = 148;
  DScene d{};
  void* arena = nullptr;
  bool has_flat = false;      // surface geometry is a single leaf: k_traverse_flat applies
  FlatList flat{};
};

// Render workspace (queues, hit buffer, control block, accumulators, events, stream).  Workspaces belong to the
// DEVICE, not to a scene: a render borrows one from the per-device free list and returns it, so building a new scene
// does not re-allocate half a gigabyte, and concurrent renders (GHC safe calls from several OS threads) never share one.
struct Workspace {
  int device = 0;
  uint32_t pool = 0;        // capacity of the queues (only ever grows: cudaFree synchronises the whole device)
  uint32_t pool_used = 0;   // logical pool of the current render (<= pool)
  Queue q[2]{};
  float2* hits = nullptr;
  float* bound = nullptr;
  Ctl* ctl = nullptr;
  Ctl* h_ctl = nullptr;   // pinned, two snapshot slots
  cudaEvent_t snap_ev[2] = {nullptr, nullptr};
  cudaEvent_t copy_ev[2] = {nullptr, nullptr};   // brackets the image read-back of rt_render (RtStats.ms_copy)
  float* sum = nullptr;
  float* sumsq = nullptr;
  float* out = nullptr;
  float* h_out = nullptr;     // pinned staging for the image read-back (2 x accum_elems: image, sum of squares)
  size_t accum_elems = 0;
  std::vector<void*> allocs;
  std::vector<cudaEvent_t> events;
  cudaStream_t stream = nullptr;
  cudaEvent_t last_done = nullptr;   // recorded when a render returned with work still queued (single-shot renders)
  bool pending = false;              // ... the next user of the workspace orders its stream behind it
};

static std::mutex g_ws_mutex;
static std::map<int, std::vector<Workspace*>> g_ws_free;

static void destroy_workspace(Workspace* w) {
  if (!w) return;
  for (void* p : w->allocs) cudaFree(p);
  for (auto e : w->events) cudaEventDestroy(e);
  if (w->h_ctl) cudaFreeHost(w->h_ctl);
  if (w->h_out) cudaFreeHost(w->h_out);
  if (w->last_done) cudaEventDestroy(w->last_done);
  for (cudaEvent_t ev : w->snap_ev) if (ev) cudaEventDestroy(ev);
  for (cudaEvent_t ev : w->copy_ev) if (ev) cudaEventDestroy(ev);
  if (w->stream) cudaStreamDestroy(w->stream);
  delete w;
}

static int32_t acquire_workspace(int device, Workspace** out) {
  {
    std::lock_guard<std::mutex> lock(g_ws_mutex);
    auto& v = g_ws_free[device];
    if (!v.empty()) { *out = v.back(); v.pop_back(); return RT_OK; }
  }
  auto* w = new Workspace();
  w->device = device;
  cudaError_t e = cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&w->h_ctl), 2 * sizeof(Ctl));
  for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&w->snap_ev[k], cudaEventDisableTiming);
  for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaEventCreate(&w->copy_ev[k]);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&w->last_done, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    set_error(std::string("workspace creation failed: ") + cudaGetErrorString(e));
    destroy_workspace(w);
    return RT_ERR_CUDA;
  }
  *out = w;
  return RT_OK;
}

static void release_workspace(Workspace* w) {
  std::lock_guard<std::mutex> lock(g_ws_mutex);
  auto& v = g_ws_free[w->device];
  if (v.size() < 4) v.push_back(w);
  else destroy_workspace(w);
}

struct WorkspaceLease {           // RAII: returns the workspace on every exit path
  Workspace* w = nullptr;
  ~WorkspaceLease() { if (w) release_workspace(w); }
};

void destroy_device_scene(DeviceScene* d) {
  if (!d) return;
  DeviceGuard guard;
  cudaSetDevice(d->device);
  if (d->arena) cudaFree(d->arena);
  delete d;
}

static int32_t check_device(int device) {
  static std::mutex mu;
  static std::map<int, int> major_of;     // cudaGetDeviceProperties costs milliseconds: ask once per device
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    set_error("no CUDA device visible: raytrace_b200 renders only on an sm_100 GPU (there is no CPU fallback)");
    return RT_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= n) { set_error("device index out of range"); return RT_ERR_INVALID_ARG; }
  int major = 0;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = major_of.find(device);
    if (it == major_of.end()) {
      RT_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
      major_of[device] = major;
    } else major = it->second;
  }
  if (major < 10) {
    set_error("device " + std::to_string(device) + " is not sm_100 (kernels are built for sm_100a only)");
    return RT_ERR_NO_DEVICE;
  }
  return RT_OK;
}

// ---- device-side tree build (rt_build.cuh) ---------------------------------------------------------------------------
// Scratch of the builder: one grow-only allocation per device, kept between builds (a cudaMalloc + cudaFree pair costs more
// than the whole build of a small mesh); builds on one device are serialised by g_build_mu.
struct BuildScratch {
  void* mem = nullptr;
  size_t bytes = 0;
  size_t clean_bins = 0;       // the first `clean_bins` Bin records hold the identity (every build leaves them so)
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  uint32_t max_grid = 0;       // co-resident blocks of k_build_tree on this device
};
static std::mutex g_build_mu;
static std::map<int, BuildScratch> g_build_scratch;

static int32_t build_deferred_tree(Scene& s, int device) {
  using namespace gpubuild;
  const auto t0 = std::chrono::steady_clock::now();
  FlatScene& f = s.flat;
  DeferredTree& D = f.deferred;
  const uint32_t n = (uint32_t)D.items.size();
  std::lock_guard<std::mutex> lock(g_build_mu);
  RT_CUDA(cudaSetDevice(device));
  BuildScratch& S = g_build_scratch[device];
  if (!S.stream) {
    RT_CUDA(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking));
    for (auto& e : S.ev) RT_CUDA(cudaEventCreate(&e));
  }
  BuildParams P{};
  P.n_items = n; P.first = D.first; P.collapse = D.collapse; P.node_cost = D.node_cost;
  P.bank_slots = n / 2u + 1u; P.node_base = D.node_base;
  const size_t n_bins = (size_t)2 * P.bank_slots * (3 * kBins);
  // layout (256-byte aligned sections): bins first, so that a kept scratch keeps its clean bins
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) / 256 * 256; return at; };
  const size_t o_bins = take(n_bins * sizeof(Bin)), o_items = take((size_t)n * sizeof(BuildItem)), o_idx0 = take((size_t)n * 4),
               o_idx1 = take((size_t)n * 4), o_nod0 = take((size_t)n * 4), o_nod1 = take((size_t)n * 4), o_off = take((size_t)n * 4),
               o_perm = take((size_t)D.n_prims * 4), o_nodes = take((size_t)2 * n * sizeof(BNode)),
               o_dn = take((size_t)2 * n * sizeof(DNode)), o_ctl = take(sizeof(BuildCtl));
  if (off > S.bytes) {
    if (S.mem) cudaFree(S.mem);
    S.mem = nullptr; S.bytes = 0; S.clean_bins = 0;
    cudaError_t e = cudaMalloc(&S.mem, off);
    if (e != cudaSuccess) { set_error(std::string("device tree build: scratch allocation failed: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    S.bytes = off;
  }
  unsigned char* base = static_cast<unsigned char*>(S.mem);
  Bin* bins = reinterpret_cast<Bin*>(base + o_bins);
  float4* items = reinterpret_cast<float4*>(base + o_items);
  uint32_t* idx[2] = {reinterpret_cast<uint32_t*>(base + o_idx0), reinterpret_cast<uint32_t*>(base + o_idx1)};
  uint32_t* nod[2] = {reinterpret_cast<uint32_t*>(base + o_nod0), reinterpret_cast<uint32_t*>(base + o_nod1)};
  uint32_t* offs = reinterpret_cast<uint32_t*>(base + o_off);
  uint32_t* perm = reinterpret_cast<uint32_t*>(base + o_perm);
  BNode* nodes = reinterpret_cast<BNode*>(base + o_nodes);
  DNode* dnodes = reinterpret_cast<DNode*>(base + o_dn);
  BuildCtl* ctl = reinterpret_cast<BuildCtl*>(base + o_ctl);
  cudaStream_t st = S.stream;
  const size_t had_clean = S.clean_bins;
  S.clean_bins = 0;                      // until this build has run to its end
  // root node and control block are prepared here; everything else happens inside ONE cooperative launch
  BNode root{};
  root.lo = 0; root.hi = n; root.slot = kNone; root.depth = 0; root.kind = kPending; root.bslot = 0; root.child = kNone;
  for (int k = 0; k < 3; k++) { root.bx[k] = kEncMax; root.bx[3 + k] = kEncMin; root.cb[k] = kEncMax; root.cb[3 + k] = kEncMin; }
  BuildCtl h{};
  h.bslots[0] = 1;                                                 // the root owns slot 0 of bank 0
  h.level_begin[0] = 0; h.level_end[0] = 1;
  RT_CUDA(cudaMemcpyAsync(items, D.items.data(), (size_t)n * sizeof(BuildItem), cudaMemcpyHostToDevice, st));
  RT_CUDA(cudaMemcpyAsync(nodes, &root, sizeof root, cudaMemcpyHostToDevice, st));
  RT_CUDA(cudaMemcpyAsync(ctl, &h, sizeof h, cudaMemcpyHostToDevice, st));
  RT_CUDA(cudaEventRecord(S.ev[0], st));
  uint32_t launches = 0;
  if (had_clean < n_bins) { k_build_fill_bins<<<592, 256, 0, st>>>(bins, n_bins); launches++; }
  const bool unit_items = D.n_prims == n;       // a mesh: every item is one primitive, storage offset = position
  BuildArgs A{};
  A.items = items; A.idx[0] = idx[0]; A.idx[1] = idx[1]; A.node_of[0] = nod[0]; A.node_of[1] = nod[1];
  A.off = unit_items ? nullptr : offs; A.perm = perm; A.nodes = nodes; A.bins = bins; A.ctl = ctl; A.dnodes = dnodes; A.P = P;
  if (!S.max_grid) {
    int per_sm = 0, sms = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_build_tree, kBlock, 0));
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    S.max_grid = (uint32_t)std::max(1, per_sm * sms);
  }
  // a persistent grid: one block per 256 items while they are all co-resident (a barrier among few blocks is cheaper), else
  // every slot of the machine and the blocks loop
  // (never fewer than one block per SM: the split phase works a warp per node, and the middle levels of a small tree have
  // hundreds of nodes)
  int sm_count = 148;
  cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
  const uint32_t grid = std::min<uint32_t>(S.max_grid, std::max<uint32_t>((n + kBlock - 1) / kBlock, (uint32_t)sm_count));
  void* kargs[] = {&A};
  RT_CUDA(cudaLaunchCooperativeKernel((const void*)k_build_tree, dim3(grid), dim3(kBlock), kargs, 0, st));
  launches++;
  RT_CUDA(cudaEventRecord(S.ev[1], st));
  // read-back: small trees fetch the worst-case node count together with the control block (one wait); large ones ask first
  const uint32_t max_dn = 2u * (n - 1u);
  const bool two_step = n > 65536u;
  std::vector<DNode> got(two_step ? 0 : max_dn);
  std::vector<uint32_t> hperm(D.n_prims);
  if (!two_step) RT_CUDA(cudaMemcpyAsync(got.data(), dnodes, (size_t)max_dn * sizeof(DNode), cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaMemcpyAsync(hperm.data(), perm, (size_t)D.n_prims * 4, cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaMemcpyAsync(&h, ctl, sizeof h, cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaStreamSynchronize(st));
  RT_CUDA(cudaGetLastError());
  const uint32_t n_dn = 2u * h.n_inner;
  if (n_dn > max_dn || (n_dn & 1u)) { set_error("device tree build: bad node count (internal error)"); return RT_ERR_CUDA; }
  got.resize(n_dn);
  if (two_step) {
    RT_CUDA(cudaMemcpyAsync(got.data(), dnodes, (size_t)n_dn * sizeof(DNode), cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
  }
  S.clean_bins = n_bins;                  // k_build_split put the identity back into every bin it consumed
  float dev_ms = 0.f;
  cudaEventElapsedTime(&dev_ms, S.ev[0], S.ev[1]);
  if (h.overflow) { set_error("device tree build: a leaf holds more than 32 primitives (internal error)"); return RT_ERR_CUDA; }
  if (h.max_depth + 1u > (uint32_t)kStackDepth) {
    set_error("flattened BVH is deeper than the shared-memory traversal stack (32 levels)");
    return RT_ERR_UNSUPPORTED;
  }
  // host arrays: node pairs behind the host-built ones; primitive records from provisional (evaluation) order to leaf order
  if (f.nodes.size() != D.node_base) { set_error("device tree build: node base moved (internal error)"); return RT_ERR_CUDA; }
  f.nodes.insert(f.nodes.end(), got.begin(), got.end());
  {
    const size_t np = D.n_prims, first = D.first;
    std::vector<float> prims(np * 16), uv(np * 8);
    std::vector<uint32_t> canon(np), ref(np);
    for (size_t j = 0; j < np; j++) {
      const size_t src = hperm[j];
      if (src < first || src >= first + np) { set_error("device tree build: bad permutation (internal error)"); return RT_ERR_CUDA; }
      std::memcpy(&prims[j * 16], &f.prims[src * 16], 64);
      std::memcpy(&uv[j * 8], &f.tri_uv[src * 8], 32);
      canon[j] = f.canonical[src];
      ref[j] = f.ref_order[src];
    }
    std::memcpy(&f.prims[first * 16], prims.data(), np * 64);
    std::memcpy(&f.tri_uv[first * 8], uv.data(), np * 32);
    std::memcpy(&f.canonical[first], canon.data(), np * 4);
    std::memcpy(&f.ref_order[first], ref.data(), np * 4);
  }
  f.max_depth = std::max(f.max_depth, h.max_depth);
  D.levels = h.levels; D.launches = launches; D.device_ms = dev_ms;
  D.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  D.built = true;                         // (before `pending` drops: a deferred scene never shows both flags clear)
  D.pending = false;
  std::vector<BuildItem>().swap(D.items);
  if (s.opts.trace_commit > 0)
    std::fprintf(stderr, "[rt_b200 upload] device tree build: %u items, %u node pairs, depth %u, %u levels, %u launch(es), grid %u: %.3f ms on the device, "
                 "%.3f ms with copies and the host permutation\n", n, n_dn / 2u, h.max_depth, h.levels, launches, grid, dev_ms, D.build_ms);
  return RT_OK;
}

static int32_t get_device_scene(Scene& s, int device, DeviceScene** out) {
  if (!s.committed) { set_error("scene not committed"); return RT_ERR_NOT_COMMITTED; }
  if (s.flat.deferred.pending || s.flat.deferred.built) {   // a scene whose surface tree was left to the device: built by the first upload
    int32_t st0 = check_device(device);
    if (st0 != RT_OK) return st0;
    std::lock_guard<std::mutex> lock(s.build_mu);
    if (s.flat.deferred.pending) {
      st0 = build_deferred_tree(s, device);
      if (st0 != RT_OK) return st0;
    }
  }
  {
    std::lock_guard<std::mutex> lock(s.devices_mu);
    auto it = s.devices.find(device);
    if (it != s.devices.end()) { *out = it->second; RT_CUDA(cudaSetDevice(device)); return RT_OK; }
  }
  // The upload itself runs WITHOUT the lock: rt_render_multi uploads a fresh scene to all its devices from one thread per
  // device, and every GPU has its own PCIe link (8 x 1.1 ms one after the other for demo2's 8.8 MB otherwise).  Two threads
  // that race for the SAME device both upload; the loser's copy is dropped below.
  int32_t st = check_device(device);
  if (st != RT_OK) return st;
  RT_CUDA(cudaSetDevice(device));
  auto* ds = new DeviceScene();
  ds->device = device;
  cudaDeviceGetAttribute(&ds->sm_count, cudaDevAttrMultiProcessorCount, device);
  const FlatScene& f = s.flat;
  DScene& d = ds->d;
  // one arena, one staging buffer, one H2D copy (256-byte aligned sections)
  struct Sec { const void* src; size_t bytes; size_t off; };
  Sec secs[11] = {
      {f.nodes.data(), f.nodes.size() * sizeof(DNode), 0},   {f.prims.data(), f.prims.size() * sizeof(float), 0},
      {f.tri_uv.data(), f.tri_uv.size() * sizeof(float), 0}, {f.xforms.data(), f.xforms.size() * sizeof(DXform), 0},
      {f.materials.data(), f.materials.size() * sizeof(DMaterial), 0}, {f.textures.data(), f.textures.size() * sizeof(DTexture), 0},
      {f.texels.data(), f.texels.size() * sizeof(float), 0}, {f.canonical.data(), f.canonical.size() * sizeof(uint32_t), 0},
      {f.noise_grad.data(), f.noise_grad.size() * sizeof(float), 0}, {f.noise_perm.data(), f.noise_perm.size(), 0},
      {f.ref_order.data(), f.ref_order.size() * sizeof(uint32_t), 0}};
  // arena layout: the small sections first (they travel in ONE staged copy), large ones (image texels: 8 MB for demo2) behind
  // them - those are copied straight from their vectors, with no zero-filled staging image and no extra host pass
  constexpr size_t kDirect = 256u << 10;
  size_t total_bytes = 0;
  for (Sec& x : secs) if (x.bytes < kDirect) { x.off = total_bytes; total_bytes += (x.bytes + 255) / 256 * 256 + 256; }
  const size_t staged_end = total_bytes;
  for (Sec& x : secs) if (x.bytes >= kDirect) { x.off = total_bytes; total_bytes += (x.bytes + 255) / 256 * 256 + 256; }
  std::vector<unsigned char> staging(staged_end, 0);
  for (const Sec& x : secs) if (x.bytes && x.bytes < kDirect) std::memcpy(staging.data() + x.off, x.src, x.bytes);
  cudaError_t e = cudaMalloc(&ds->arena, total_bytes);
  if (e == cudaSuccess && staged_end) e = cudaMemcpy(ds->arena, staging.data(), staged_end, cudaMemcpyHostToDevice);
  for (const Sec& x : secs)
    if (e == cudaSuccess && x.bytes >= kDirect) e = cudaMemcpy(static_cast<unsigned char*>(ds->arena) + x.off, x.src, x.bytes, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error(std::string("scene upload failed: ") + cudaGetErrorString(e));
    destroy_device_scene(ds);
    return RT_ERR_CUDA;
  }
  const unsigned char* base = static_cast<const unsigned char*>(ds->arena);
  d.nodes = reinterpret_cast<const float4*>(base + secs[0].off);
  d.prims = reinterpret_cast<const float4*>(base + secs[1].off);
  d.tri_uv = reinterpret_cast<const float4*>(base + secs[2].off);
  d.xforms = reinterpret_cast<const DXform*>(base + secs[3].off);
  d.materials = reinterpret_cast<const DMaterial*>(base + secs[4].off);
  d.textures = reinterpret_cast<const DTexture*>(base + secs[5].off);
  d.texels = reinterpret_cast<const float4*>(base + secs[6].off);
  d.canonical = reinterpret_cast<const uint32_t*>(base + secs[7].off);
  d.noise_grad = reinterpret_cast<const float*>(base + secs[8].off);
  d.noise_perm = reinterpret_cast<const uint8_t*>(base + secs[9].off);
  d.ref_order = reinterpret_cast<const uint32_t*>(base + secs[10].off);
  if (f.surface_root != kNone && (f.surface_root & kLeafBit)) {
    const uint32_t first = (f.surface_root & ~kLeafBit) >> kLeafShift, count = (f.surface_root & kLeafCountMask) + 1u;
    FlatList& L = ds->flat;
    bool fits = true;
    for (uint32_t i = first; i < first + count; i++) {
      const float* rec = f.prims.data() + (size_t)i * 16;
      uint32_t meta[4];
      std::memcpy(meta, rec + 12, 16);
      if (meta[0] & PRIM_FLAG_BOXTAIL) continue;
      if (meta[0] & PRIM_FLAG_BOXHEAD) {
        if (L.n_box == (uint32_t)kFlatBoxes) { fits = false; break; }   // (cannot happen with leaves of <= 24 primitives)
        std::memcpy(L.box[L.n_box], f.tri_uv.data() + (size_t)i * 8, 64);
        L.box_first[L.n_box] = i;
        L.box_geom[L.n_box] = meta[1];
        L.n_box++;
        continue;
      }
      std::memcpy(L.rec[L.n], rec, 64);
      L.index[L.n] = i;
      L.n++;
    }
    ds->has_flat = fits;
  }
  std::memcpy(d.media, f.media, sizeof d.media);
  d.n_media = f.n_media;
  d.surface_root = f.surface_root;
  d.has_noise = f.has_noise;
  for (int k = 0; k < 3; k++) { d.world_lo[k] = f.world_lo[k]; d.world_hi[k] = f.world_hi[k]; }
  d.has_world_box = f.has_world_box;
  d.n_prims = (uint32_t)f.canonical.size();
  d.moving_first = f.moving_first;
  d.moving_count = f.moving_count;
  d.n_nodes = (uint32_t)f.nodes.size();
  d.n_materials = (uint32_t)f.materials.size();
  d.n_textures = (uint32_t)f.textures.size();
  d.n_xforms = (uint32_t)f.xforms.size();
  d.stack_entries = f.max_depth + 1;
  {
    std::lock_guard<std::mutex> lock(s.devices_mu);
    auto it = s.devices.find(device);
    if (it != s.devices.end()) {          // another thread uploaded this device meanwhile: keep its copy
      destroy_device_scene(ds);
      RT_CUDA(cudaSetDevice(device));
      *out = it->second;
      return RT_OK;
    }
    s.devices[device] = ds;
  }
  *out = ds;
  return RT_OK;
}

// Camera of Ray.hs:123-155 evaluated in double, stored fp32
static int32_t make_camera(const RtCamera& c, DCamera* out) {
  if (c.image_width <= 0 || !(c.aspect_ratio > 0) || c.samples_per_pixel <= 0) { set_error("bad camera settings"); return RT_ERR_INVALID_ARG; }
  if (c.n_redirect_targets < 0 || c.n_redirect_targets > kMaxTargets) { set_error("more than 8 cs_redirectTargets"); return RT_ERR_UNSUPPORTED; }
  if (c.n_redirect_targets && !c.redirect_targets) { set_error("null redirect target array"); return RT_ERR_INVALID_ARG; }
  if (c.bg_kind != RT_BG_CONSTANT && c.bg_kind != RT_BG_YGRADIENT) {
    set_error("cs_background is a user-defined closure (neither constant nor affine in dir.y): rejected, no CPU fallback");
    return RT_ERR_USER_DEFINED;
  }
  DCamera k{};
  k.W = c.image_width;
  k.H = (int)std::nearbyint((double)c.image_width / c.aspect_ratio);
  if (k.H <= 0) { set_error("image height is zero"); return RT_ERR_INVALID_ARG; }
  k.spp = c.samples_per_pixel;
  k.max_depth = c.max_recursion_depth;
  auto sub = [](const double* a, const double* b, double* r) { for (int i = 0; i < 3; i++) r[i] = a[i] - b[i]; };
  auto crs = [](const double* a, const double* b, double* r) {
    r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0];
  };
  auto nrm = [](double* a) {
    double l = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
    if (std::fabs(l) <= 1e-12 || std::fabs(1 - l) <= 1e-12) return;   // linear's normalize
    double s = 1 / std::sqrt(l);
    for (int i = 0; i < 3; i++) a[i] *= s;
  };
  double vh = c.focus_dist * std::tan(c.vfov / 2) * 2;
  double vw = vh * (double)k.W / (double)k.H;
  double w[3], u[3], v[3];
  sub(c.center, c.look_at, w);
  nrm(w);
  crs(c.up, w, u);
  nrm(u);
  crs(w, u, v);
  double rad = c.focus_dist * std::tan(c.defocus_angle / 2);
  for (int i = 0; i < 3; i++) {
    double across = vw * u[i], down = -(vh * v[i]);
    k.center[i] = (float)c.center[i];
    k.top_left[i] = (float)(-w[i] * c.focus_dist - across / 2 - down / 2);   // relative to the centre
    k.pixel_u[i] = (float)(across / k.W);
    k.pixel_v[i] = (float)(down / k.H);
    k.defocus_u[i] = (float)(u[i] * rad);
    k.defocus_v[i] = (float)(v[i] * rad);
    k.bg0[i] = (float)c.bg_c0[i];
    k.bg1[i] = (float)c.bg_c1[i];
  }
  k.has_defocus = rad != 0.0;
  k.bg_kind = c.bg_kind;
  k.n_targets = c.n_redirect_targets;
  double acc = 0;
  for (int t = 0; t < k.n_targets; t++) {
    const RtRedirectTarget& r = c.redirect_targets[t];
    DTarget& d = k.targets[t];
    acc += r.p;
    d.p = (float)r.p;
    d.cumulative = (float)acc;
    double cp[3];
    crs(r.s0, r.s1, cp);
    double ncp = std::sqrt(cp[0] * cp[0] + cp[1] * cp[1] + cp[2] * cp[2]);
    double n[3] = {0, 0, 0}, nS[3] = {0, 0, 0}, wa[3] = {0, 0, 0}, wb[3] = {0, 0, 0};
    if (ncp > 0) {
      for (int i = 0; i < 3; i++) { n[i] = cp[i] / ncp; nS[i] = n[i] / ncp; }
      crs(r.s1, nS, wa);
      crs(nS, r.s0, wb);
    }
    for (int i = 0; i < 3; i++) {
      d.origin[i] = (float)r.q[i]; d.U[i] = (float)r.s0[i]; d.V[i] = (float)r.s1[i]; d.crs[i] = (float)cp[i];
      d.n[i] = (float)n[i]; d.wa[i] = (float)wa[i]; d.wb[i] = (float)wb[i];
    }
  }
  k.rem_prob = (float)(1.0 - acc);
  *out = k;
  return RT_OK;
}

static int32_t ensure_workspace(Workspace& ds, uint32_t pool, size_t accum_elems) {
  ds.pool_used = pool;
  if (ds.pool < pool) {
    auto drop = [&](void* p) {
      if (!p) return;
      cudaFree(p);
      ds.allocs.erase(std::remove(ds.allocs.begin(), ds.allocs.end(), p), ds.allocs.end());
    };
    for (int k = 0; k < 2; k++) { drop(ds.q[k].o); drop(ds.q[k].d); drop(ds.q[k].thr); drop(ds.q[k].meta); ds.q[k] = Queue{}; }
    drop(ds.hits); drop(ds.bound); drop(ds.ctl);
    ds.hits = nullptr; ds.bound = nullptr; ds.ctl = nullptr; ds.pool = 0;
    auto grab = [&](void** p, size_t bytes) -> cudaError_t {
      cudaError_t e = cudaMalloc(p, bytes);
      if (e == cudaSuccess) ds.allocs.push_back(*p);
      return e;
    };
    for (int k = 0; k < 2; k++) {
      RT_CUDA(grab((void**)&ds.q[k].o, (size_t)pool * 16));
      RT_CUDA(grab((void**)&ds.q[k].d, (size_t)pool * 16));
      RT_CUDA(grab((void**)&ds.q[k].thr, (size_t)pool * 16));
      RT_CUDA(grab((void**)&ds.q[k].meta, (size_t)pool * 8));
    }
    RT_CUDA(grab((void**)&ds.hits, (size_t)pool * 8));
    RT_CUDA(grab((void**)&ds.bound, (size_t)pool * 4));
    RT_CUDA(grab((void**)&ds.ctl, sizeof(Ctl)));
    ds.pool = pool;
  }
  if (ds.accum_elems < accum_elems) {
    for (float* old : {ds.sum, ds.sumsq, ds.out}) {
      if (!old) continue;
      cudaFree(old);
      ds.allocs.erase(std::remove(ds.allocs.begin(), ds.allocs.end(), (void*)old), ds.allocs.end());
    }
    ds.sum = ds.sumsq = ds.out = nullptr;
    ds.accum_elems = 0;
    void* p[3] = {nullptr, nullptr, nullptr};
    for (int k = 0; k < 3; k++) {
      RT_CUDA(cudaMalloc(&p[k], accum_elems * sizeof(float)));
      ds.allocs.push_back(p[k]);
    }
    ds.sum = (float*)p[0]; ds.sumsq = (float*)p[1]; ds.out = (float*)p[2];
    if (ds.h_out) cudaFreeHost(ds.h_out);   // the staging buffer is re-created on demand (pageable destinations only)
    ds.h_out = nullptr;
    ds.accum_elems = accum_elems;
  }
  return RT_OK;
}

static uint32_t choose_pool(const Scene& s, unsigned long long total_items) {
  uint32_t pool = s.opts.pool > 0 ? (uint32_t)std::min<int64_t>(s.opts.pool, 1ll << 30) : 0u;
  if (pool == 0) pool = 1u << 25;   // fewer, larger iterations: launch gaps and kernel tails amortise (profiles/r02x_sweep_pool.txt); 3.8 GB
  pool = (std::max<uint32_t>(pool, 1024) + 1023u) / 1024u * 1024u;   // whole 256-slot tiles for the TMA prefetch in k_shade
  if (total_items < pool) pool = (uint32_t)std::max<unsigned long long>(1024, (total_items + 1023) / 1024 * 1024);
  return pool;
}

// the wavefront loop, accumulating UN-normalised sums into d_sum (+ d_sumsq)
static int32_t run_wavefront_range(Scene& s, DeviceScene& dev, Workspace& ds, const DCamera& cam, uint64_t seed, uint64_t first_sample,
                             uint64_t n_samples, float* d_sum, float* d_sumsq, cudaStream_t stream, RtStats* stats) {
  const uint32_t n_pixels = (uint32_t)cam.W * (uint32_t)cam.H;
  const unsigned long long total = (unsigned long long)n_pixels * n_samples;
  if (cam.max_depth <= 0) {   // rayColor 0 _ _ = zero (Ray.hs:176): every path is black before its first segment
    if (stats) { stats->samples = total; stats->pool = ds.pool_used; }
    return RT_OK;
  }
  RenderArgs A{};
  A.cam = cam;
  A.key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  A.first_sample = first_sample;
  A.n_pixels = n_pixels;
  A.sum = d_sum;
  A.sumsq = d_sumsq;
  A.hits = ds.hits;
  A.bound = ds.bound;
  A.refill = 24;
  A.node_burst = 4;
  // k_finish takes over the tail.  Flat scenes: with the constant-bank scan a lane that carries its path to the end
  // (and claims the next one when it dies) is faster per ray than a wavefront iteration once no camera rays are left to
  // generate, so it takes over as soon as every path has started, whatever their number.  BVH scenes: worth it only when
  // the depth limit leaves dozens of near-empty iterations to replace (demo1 / demo2, depth 50); at depth 20 with
  // subsurface media (pawn, bunny) the iterations are cheaper.  Sweeps: profiles/r01h_sweep_tail.txt
  const RtOptions& O = s.opts;
  // "flat scene" = the fused one-launch kernel applies: the surfaces are one short list AND nothing needs the generic
  // traversal (no media, no moving instances); a short list with media / movers runs the BVH kernels on a leaf root
  // RtOptions.material_sampling (off by default: parity with the reference's estimator) lives in k_shade<., true> only: the
  // fused flat-scene kernel and the tail kernel stand aside for it
  const bool hook = O.material_sampling > 0;
  const TraceFlatFn fused = dev.has_flat && !hook ? pick_trace_flat(s.flat.features) : nullptr;
  const bool flat_scene = fused != nullptr;
  A.finish_max = flat_scene ? 0xFFFFFFFFu : (cam.max_depth > 24 ? 1u << 20 : 0u);
  if (O.finish_max >= 0) A.finish_max = (uint32_t)std::min<int64_t>(O.finish_max, 0xFFFFFFFFll);
  if (hook) A.finish_max = 0u;
  A.finish_refill = 4;
  if (O.finish_refill >= 0) A.finish_refill = std::min(32, std::max(1, (int)O.finish_refill));
  A.tile = flat_scene ? 0u : 1u;   // coherence only matters to the BVH kernels; the flat scan is the same for every ray
  if (O.tile >= 0) A.tile = O.tile ? 1u : 0u;
  // Camera rays that miss the world box end inside k_raygen.  Worth the test (and the compaction it brings) only when the
  // lens lies OUTSIDE the box - from inside it every ray passes (demo1 / demo2: +1 ms for nothing).  Flat scenes: the fused
  // kernel generates its own camera rays.
  A.cull = 0u;
  if (dev.d.has_world_box && !flat_scene) {
    float lens = 0.f;
    if (cam.has_defocus)
      lens = std::max(std::sqrt(cam.defocus_u[0] * cam.defocus_u[0] + cam.defocus_u[1] * cam.defocus_u[1] + cam.defocus_u[2] * cam.defocus_u[2]),
                      std::sqrt(cam.defocus_v[0] * cam.defocus_v[0] + cam.defocus_v[1] * cam.defocus_v[1] + cam.defocus_v[2] * cam.defocus_v[2]));
    for (int k = 0; k < 3; k++)
      if (cam.center[k] + lens < dev.d.world_lo[k] || cam.center[k] - lens > dev.d.world_hi[k]) A.cull = 1u;
  }
  if (O.primary_cull >= 0) A.cull = (O.primary_cull && dev.d.has_world_box && !flat_scene) ? 1u : 0u;
  for (int k = 0; k < 3; k++) { A.world_lo[k] = dev.d.world_lo[k]; A.world_hi[k] = dev.d.world_hi[k]; }
  {   // Material-coherent shading pays only when some shader is MUCH longer than the others (Perlin octaves, image-texture
      // uv): demo2 47.0 -> 45.2 ms per 100 spp.  With plain constant textures the sort costs more than the divergence it
      // removes (demo1 22.3 -> 23.3, pawn 23.2 -> 24.1; profiles/r02c_sweep.txt), so it stays off there.
    uint32_t texk = 0;
    for (const DTexture& t : s.flat.textures) texk |= 1u << (t.kind & 31);
    A.shade_sort = (texk & ((1u << RT_TEX_NOISE) | (1u << RT_TEX_MARBLE) | (1u << RT_TEX_IMAGE))) != 0 ? 1u : 0u;
    if (O.shade_sort >= 0) A.shade_sort = O.shade_sort ? 1u : 0u;
  }
  // culling iterations: a k_raygen thread owns K consecutive samples of a pixel (one background atomic triple per K samples)
  A.kblock = 1u;
  if (A.cull) {
    A.kblock = 8u;
    if (O.raygen_block >= 0) A.kblock = (uint32_t)std::min(64, std::max(1, (int)O.raygen_block));
    while (A.kblock > 1u && (unsigned long long)A.kblock * n_pixels >= 0xFFFFFFFFull) A.kblock /= 2u;
  }
  A.ktail = (uint32_t)(n_samples % A.kblock);
  A.kfull_items = (n_samples / A.kblock) * A.kblock * (unsigned long long)n_pixels;
  int tail_batch = 3;                // iterations (each followed by a k_finish check) per host round trip once all paths have started
  if (O.tail_batch >= 0) tail_batch = std::min(64, std::max(1, (int)O.tail_batch));
  if (O.node_burst >= 0) A.node_burst = std::min(8, std::max(1, (int)O.node_burst));
  if (O.refill >= 0) A.refill = std::min(32, std::max(1, (int)O.refill));
  const bool timing = stats != nullptr && O.timing != 0;
  const size_t kMaxTimed = 4096;            // iterations with per-kernel events
  if (timing && ds.events.size() < kMaxTimed * 4 + 2) {
    size_t want = kMaxTimed * 4 + 2;
    while (ds.events.size() < want) {
      cudaEvent_t e;
      RT_CUDA(cudaEventCreate(&e));
      ds.events.push_back(e);
    }
  }
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  if (timing) { ev_begin = ds.events[kMaxTimed * 4]; ev_end = ds.events[kMaxTimed * 4 + 1]; RT_CUDA(cudaEventRecord(ev_begin, stream)); }

  const bool count_work = stats != nullptr && O.count_work > 0;
  const TraverseFn traverse = count_work ? pick_by_features<CountedK>(s.flat.features) : pick_by_features<PersistentK>(s.flat.features);
  const TraverseFn finish = pick_by_features<FinishK>(s.flat.features);
  const FinishFlatFn finish_flat = fused ? pick_finish_flat(s.flat.features) : nullptr;
  const size_t stack_bytes = (size_t)(s.flat.max_depth + 1) * kTraverseBlock * sizeof(uint32_t);
  // L1 / shared-memory split of the traversal kernels (RtOptions.smem_carveout).  Left alone, the driver gives k_traverse_p a
  // 132 KB shared-memory carve-out for the 63 KB its nine resident blocks use on pawn (ncu: launch__shared_mem_config_size
  // 135 KB), i.e. 124 KB of L1 for a kernel whose node pairs and primitive records are fetched by divergent lanes at L1
  // latency or ten times that.  Default: ask for the smallest supported carve-out that holds the resident blocks' stacks
  // (the hint sits halfway between that size and the next smaller one: the driver's rounding at the edges is not what the
  // percentages suggest - 28 % came back as 100 KB).  pawn (64 KB): L1 hit rate 72 -> 80 %+, traversal -4 %.  The attribute
  // belongs to the FUNCTION, not the scene: it is set again whenever the size a render needs differs from the one in force
  // (a 64 KB carve-out left behind by pawn cost bunny, same specialisation, two resident blocks and 5 %).
  // 0: never touch the attribute; > 0: that percentage as the hint.  profiles/r02zc_sweep_smem_carveout.txt
  if (O.smem_carveout != 0) {
    int max_smem = 233472;
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev.device);
    // resident blocks as the REGISTERS allow (no dynamic shared memory in the query: the answer must not depend on the
    // carve-out a previous scene left on the function - the query below, with the stacks, runs after the attribute is set)
    int by_regs = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&by_regs, traverse, kTraverseBlock, 0) != cudaSuccess || by_regs <= 0) { cudaGetLastError(); by_regs = 9; }
    if (O.trav_blocks >= 0) by_regs = std::max(1, (int)O.trav_blocks);
    const size_t need = (size_t)by_regs * (stack_bytes + 1024);
    size_t prev = 0, hint = (size_t)max_smem;
    for (size_t kb : {32u, 64u, 100u, 132u, 164u, 196u, 228u}) {      // (never below 32 KB: the register-limited query above must stay unconstrained)
      if (kb * 1024 >= need) { hint = (prev + kb * 1024) / 2; break; }
      prev = kb * 1024;
    }
    const int pct = O.smem_carveout > 0 ? std::min(100, (int)O.smem_carveout) : (int)std::min<size_t>(100, hint * 100 / (size_t)max_smem);
    static std::mutex carve_mu;
    static std::map<std::pair<const void*, int>, int> carve_set;      // (function, device) -> hint in force + 1
    std::lock_guard<std::mutex> lock(carve_mu);
    for (const void* fn : {(const void*)traverse, (const void*)finish}) {
      int& have = carve_set[{fn, dev.device}];
      if (have != pct + 1) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, pct) == cudaSuccess) have = pct + 1;
        else cudaGetLastError();
      }
    }
  }
  int g_trav = dev.sm_count * 8, g_shade = dev.sm_count * 8, g_gen = dev.sm_count * 4;
  {   // the persistent traversal grid = exactly the blocks that are resident at once (8 or 9 per SM, by specialisation)
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, traverse, kTraverseBlock, stack_bytes) == cudaSuccess && per_sm > 0)
      g_trav = dev.sm_count * per_sm;
    else cudaGetLastError();
  }
  if (O.trav_blocks >= 0) g_trav = dev.sm_count * std::max(1, (int)O.trav_blocks);
  uint64_t launches = 0;
  size_t timed = 0;
  std::vector<char> has_fin(timing ? kMaxTimed : 0, 0);   // iteration k was followed by a (timed) k_finish launch
  uint32_t iter = 0;
  int cur = 0;
  // The loop is device driven: the last k_shade block prepares the next iteration, the host only launches batches
  // and reads the 80-byte control block back between them.  Batch length follows the measured consumption of work
  // items (iterations past the end are empty launches, so the estimate stays a little short).
  k_begin<<<1, 1, 0, stream>>>(ds.ctl, total, ds.pool_used);
  launches++;
  // tail mode: every camera path has started (known up front when the whole range fits the pool, e.g. one GPU's share
  // of a small render).  From then on k_finish is checked after EVERY iteration - a ~2 us no-op until few enough paths
  // are left - so the takeover never waits for a host round trip, and launches queued past it are empty.
  bool tail = total <= ds.pool_used;
  // flat scenes, unbounded tail kernel: once every path has started, ONE more iteration + k_finish_flat completes the
  // render, so the launch sequence is known without reading the control block back
  const bool finish_all = finish_flat != nullptr && A.finish_max == 0xFFFFFFFFu;
  const bool single_shot = tail && finish_all;   // the whole sample range fits the pool: k_begin, one iteration, tail
  const unsigned long long tail_chunk = finish_all ? 1ull : (A.finish_max ? (unsigned long long)tail_batch : 8ull);

  // RT_B200_DEBUG_SYNC=1: synchronise after every launch and name the kernel that faulted (stderr)
  const bool debug_sync = O.debug_sync > 0;
  auto dbg = [&](const char* what) {
    if (!debug_sync) return;
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) std::fprintf(stderr, "[rt_b200] %s (iteration %u): %s\n", what, iter, cudaGetErrorString(e));
  };
  const bool shade_bound = dev.d.n_media == 1u && dev.d.media[0].closed != 0u;   // k_shade draws the free-flight bound of interior segments
  auto launch_iteration = [&](bool fin_check) {
    Queue& q = ds.q[cur];
    Queue& qn = ds.q[cur ^ 1];
    const bool t_this = timing && timed < kMaxTimed;
    if (fused) {   // one launch per iteration; reported as shade time (ms_traverse = ms_raygen = 0)
      if (t_this) { cudaEventRecord(ds.events[timed * 4 + 0], stream); cudaEventRecord(ds.events[timed * 4 + 1], stream); }
      fused<<<g_shade, kShadeBlock, 0, stream>>>(ds.ctl, q, qn, dev.d, A, dev.flat);
      launches += 1;
    } else {
      k_raygen<<<g_gen, kRaygenBlock, 0, stream>>>(ds.ctl, q, A);
      dbg("k_raygen");
      if (t_this) cudaEventRecord(ds.events[timed * 4 + 0], stream);
      traverse<<<g_trav, kTraverseBlock, stack_bytes, stream>>>(ds.ctl, q, dev.d, A);
      dbg("traverse");
      if (t_this) cudaEventRecord(ds.events[timed * 4 + 1], stream);
      if (hook) {
        if (shade_bound) k_shade<true, true><<<g_shade, kShadeBlock, 0, stream>>>(ds.ctl, q, qn, dev.d, A);
        else k_shade<false, true><<<g_shade, kShadeBlock, 0, stream>>>(ds.ctl, q, qn, dev.d, A);
      } else if (shade_bound) k_shade<true><<<g_shade, kShadeBlock, 0, stream>>>(ds.ctl, q, qn, dev.d, A);
      else k_shade<false><<<g_shade, kShadeBlock, 0, stream>>>(ds.ctl, q, qn, dev.d, A);
      dbg("k_shade");
      launches += 3;
    }
    if (t_this) { cudaEventRecord(ds.events[timed * 4 + 2], stream); timed++; }
    cur ^= 1;
    iter++;
    if (A.finish_max && fin_check) {   // device-side decision: no-op until the work items are exhausted and few paths are left
      if (finish_flat) finish_flat<<<dev.sm_count * 8, kTraverseBlock, 0, stream>>>(ds.ctl, ds.q[cur], dev.d, A, dev.flat);
      else finish<<<dev.sm_count * 8, kTraverseBlock, stack_bytes, stream>>>(ds.ctl, ds.q[cur], dev.d, A);
      launches++;
      dbg("k_finish");
      if (t_this) { cudaEventRecord(ds.events[(timed - 1) * 4 + 3], stream); has_fin[timed - 1] = 1; }
    }
  };
  // A chunk = some iterations, then a snapshot of the control block into a pinned slot + an event; the host waits for
  // the snapshot, plans the next chunk from the measured consumption of work items (iterations past the end are empty
  // launches, so the estimate stays a little short) and launches it.  (Keeping one more chunk queued behind the one
  // waited for was measured in round 1 and dropped: the ~5 round trips of a render cost less than the staler plans.)
  struct Flight { int slot; unsigned long long iters; };
  int slot = 0;
  auto launch_chunk = [&](unsigned long long n) -> Flight {
    for (unsigned long long b2 = 0; b2 < n; b2++) launch_iteration(tail || b2 + 1 == n);
    cudaMemcpyAsync(&ds.h_ctl[slot], ds.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, stream);
    cudaEventRecord(ds.snap_ev[slot], stream);
    Flight f{slot, (unsigned long long)iter};
    slot ^= 1;
    return f;
  };
  unsigned long long chunk_next = tail ? tail_chunk : 2ull;
  unsigned long long prev_started = 0, prev_iters = 0;
  Flight cur_f = launch_chunk(tail ? tail_chunk : 4ull);
  if (single_shot && !stats) {
    // nothing to read back: return with the work queued (stream-ordered completion, like any device-buffer API);
    // whoever borrows this workspace next orders its stream behind last_done
    RT_CUDA(cudaEventRecord(ds.last_done, stream));
    ds.pending = true;
    RT_CUDA(cudaGetLastError());
    return RT_OK;
  }
  const Ctl* final_ctl = nullptr;
  for (;;) {
    RT_CUDA(cudaEventSynchronize(ds.snap_ev[cur_f.slot]));
    const Ctl& c = ds.h_ctl[cur_f.slot];
    if (c.n_live == 0) { final_ctl = &c; break; }   // the render was complete at this snapshot
    const unsigned long long started = c.next_item;
    const double per_iter = cur_f.iters > prev_iters ? (double)(started - prev_started) / (double)(cur_f.iters - prev_iters) : 0.0;
    prev_started = started;
    prev_iters = cur_f.iters;
    if (started >= total) { tail = true; chunk_next = tail_chunk; }
    else if (per_iter >= 1.0) chunk_next = (unsigned long long)std::min(4096.0, std::max(2.0, 0.9 * (double)(total - started) / per_iter));
    else chunk_next = 2;
    cur_f = launch_chunk(chunk_next);
    if (iter > 100000000u) { set_error("wavefront loop did not terminate"); return RT_ERR_CUDA; }
  }
  if (timing) { RT_CUDA(cudaEventRecord(ev_end, stream)); RT_CUDA(cudaEventSynchronize(ev_end)); }
  RT_CUDA(cudaGetLastError());
  if (stats) {
    stats->samples = total;
    stats->rays = final_ctl->rays;
    stats->rays_finish = final_ctl->rays_finish;
    stats->paths_finish = final_ctl->paths_finish;
    stats->launches = launches;
    stats->iterations = final_ctl->iterations;
    stats->pool = ds.pool_used;
    stats->node_pairs = final_ctl->work[0]; stats->sphere_tests = final_ctl->work[1]; stats->planar_tests = final_ctl->work[2];
    stats->box_tests = final_ctl->work[3]; stats->rays_counted = final_ctl->work[4];
    if (timing) {
      float ms = 0;
      cudaEventElapsedTime(&ms, ev_begin, ev_end);
      stats->ms_total = ms;
      double tt = 0, tsd = 0, tfin = 0;
      for (size_t k = 0; k < timed; k++) {
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, ds.events[k * 4 + 0], ds.events[k * 4 + 1]);
        cudaEventElapsedTime(&b, ds.events[k * 4 + 1], ds.events[k * 4 + 2]);
        if (has_fin[k]) cudaEventElapsedTime(&c, ds.events[k * 4 + 2], ds.events[k * 4 + 3]);
        tt += a;
        tsd += b;
        tfin += c;
      }
      // iterations beyond the timed window (very long renders) are extrapolated by iteration count
      double scale = timed ? (double)iter / (double)timed : 0.0;
      if (timed == kMaxTimed && iter > timed) { tt *= scale; tsd *= scale; }
      stats->ms_traverse = tt;
      stats->ms_shade = tsd;
      stats->ms_finish = tfin;
      stats->ms_raygen = std::max(0.0, (double)ms - tt - tsd - tfin);
    }
  }
  return RT_OK;
}

// Queue records carry the low 32 bits of the global sample index (the high bits come from first_sample), so a range
// that crosses a multiple of 2^32 is rendered as two consecutive ranges into the same accumulators.
static int32_t run_wavefront(Scene& s, DeviceScene& dev, Workspace& ds, const DCamera& cam, uint64_t seed, uint64_t first_sample,
                             uint64_t n_samples, float* d_sum, float* d_sumsq, cudaStream_t stream, RtStats* stats) {
  RtStats total{};
  while (n_samples) {
    const uint64_t room = 0x100000000ull - (first_sample & 0xFFFFFFFFull);
    const uint64_t n = std::min<uint64_t>(n_samples, room);
    RtStats part{};
    int32_t st = run_wavefront_range(s, dev, ds, cam, seed, first_sample, n, d_sum, d_sumsq, stream, stats ? &part : nullptr);
    if (st != RT_OK) return st;
    total.samples += part.samples; total.rays += part.rays; total.launches += part.launches; total.iterations += part.iterations;
    total.pool = part.pool; total.ms_total += part.ms_total; total.ms_traverse += part.ms_traverse; total.ms_shade += part.ms_shade;
    total.ms_raygen += part.ms_raygen; total.ms_finish += part.ms_finish; total.rays_finish += part.rays_finish; total.paths_finish += part.paths_finish;
    total.node_pairs += part.node_pairs; total.sphere_tests += part.sphere_tests; total.planar_tests += part.planar_tests;
    total.box_tests += part.box_tests; total.rays_counted += part.rays_counted;
    first_sample += n;
    n_samples -= n;
  }
  if (stats) *stats = total;
  return RT_OK;
}

}  // namespace rtb

using namespace rtb;

extern "C" {

int32_t rt_device_count(int32_t* n) {
  if (!n) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  int c = 0, good = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) c = 0;
  for (int i = 0; i < c; i++) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major >= 10) good++;
  }
  *n = good;
  return RT_OK;
}

static int32_t rt_scene_upload_impl(RtScene* sp, int32_t device);
int32_t rt_scene_upload(RtScene* sp, int32_t device) {
  return abi_guard("rt_scene_upload", [&]() { return rt_scene_upload_impl(sp, device); });
}
static int32_t rt_scene_upload_impl(RtScene* sp, int32_t device) {
  if (!sp) { set_error("null scene"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  DeviceScene* ds = nullptr;
  return get_device_scene(*reinterpret_cast<Scene*>(sp), device, &ds);
}

static int32_t rt_render_device_impl(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats);
int32_t rt_render_device(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats) {
  return abi_guard("rt_render_device", [&]() { return rt_render_device_impl(sp, device, rc, seed, first_sample, n_samples, d_sum, d_sumsq, cuda_stream, stats); });
}
static int32_t rt_render_device_impl(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* d_sum, float* d_sumsq, void* cuda_stream, RtStats* stats) {
  if (!sp || !rc || !d_sum) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  Scene& s = *reinterpret_cast<Scene*>(sp);
  DeviceScene* ds = nullptr;
  int32_t st = get_device_scene(s, device, &ds);
  if (st != RT_OK) return st;
  DCamera cam;
  st = make_camera(*rc, &cam);
  if (st != RT_OK) return st;
  if (stats) std::memset(stats, 0, sizeof *stats);
  if (n_samples == 0) return RT_OK;
  const unsigned long long total = (unsigned long long)cam.W * cam.H * n_samples;
  WorkspaceLease lease;
  st = acquire_workspace(device, &lease.w);
  if (st != RT_OK) return st;
  st = ensure_workspace(*lease.w, choose_pool(s, total), 0);
  if (st != RT_OK) return st;
  // the caller's stream, NULL = the default stream (CUDA convention): what the caller queues next is ordered behind the render
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  if (lease.w->pending) { RT_CUDA(cudaStreamWaitEvent(stream, lease.w->last_done, 0)); lease.w->pending = false; }
  return run_wavefront(s, *ds, *lease.w, cam, seed, first_sample, n_samples, d_sum, d_sumsq, stream, stats);
}

// ---- page-locked host buffers for the caller's image (recycled: cudaMallocHost / cudaFreeHost cost ~0.3 ms each) ----
static std::mutex g_host_mu;
static std::map<void*, size_t> g_host_live;                  // handed out: pointer → capacity
static std::multimap<size_t, void*> g_host_free;             // capacity → buffer, at most kHostFreeMax bytes kept
static size_t g_host_free_bytes = 0;
constexpr size_t kHostFreeMax = 256u << 20;

int32_t rt_host_alloc(size_t bytes, void** out) {
  if (!out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  *out = nullptr;
  if (bytes == 0) bytes = 1;
  const size_t cap = (bytes + 4095) / 4096 * 4096;
  {
    std::lock_guard<std::mutex> lock(g_host_mu);
    auto it = g_host_free.lower_bound(cap);
    if (it != g_host_free.end() && it->first <= cap + cap / 4) {   // close enough in size: reuse
      *out = it->second;
      g_host_live[it->second] = it->first;
      g_host_free_bytes -= it->first;
      g_host_free.erase(it);
      return RT_OK;
    }
  }
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    set_error("no CUDA device visible: page-locked buffers need the CUDA driver (there is no CPU fallback)");
    return RT_ERR_NO_DEVICE;
  }
  void* p = nullptr;
  cudaError_t e = cudaHostAlloc(&p, cap, cudaHostAllocPortable);
  if (e != cudaSuccess) { set_error(std::string("cudaHostAlloc failed: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
  std::lock_guard<std::mutex> lock(g_host_mu);
  g_host_live[p] = cap;
  *out = p;
  return RT_OK;
}

int32_t rt_host_free(void* p) {
  if (!p) return RT_OK;
  size_t cap = 0;
  bool keep = false;
  {
    std::lock_guard<std::mutex> lock(g_host_mu);
    auto it = g_host_live.find(p);
    if (it == g_host_live.end()) { set_error("rt_host_free: not a buffer from rt_host_alloc"); return RT_ERR_INVALID_ARG; }
    cap = it->second;
    g_host_live.erase(it);
    if (g_host_free_bytes + cap <= kHostFreeMax) { g_host_free.emplace(cap, p); g_host_free_bytes += cap; keep = true; }
  }
  if (!keep) cudaFreeHost(p);
  return RT_OK;
}

// true when [p, p + bytes) is page-locked memory the copy engines can write directly
static bool is_pinned_host(const void* p, size_t bytes) {
  cudaPointerAttributes a0{}, a1{};
  if (cudaPointerGetAttributes(&a0, p) != cudaSuccess) { cudaGetLastError(); return false; }
  if (cudaPointerGetAttributes(&a1, static_cast<const char*>(p) + bytes - 1) != cudaSuccess) { cudaGetLastError(); return false; }
  return a0.type == cudaMemoryTypeHost && a1.type == cudaMemoryTypeHost;
}

static int32_t rt_render_impl(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* rgb_out, float* rgb_sumsq_out, RtStats* stats);
int32_t rt_render(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* rgb_out, float* rgb_sumsq_out, RtStats* stats) {
  return abi_guard("rt_render", [&]() { return rt_render_impl(sp, device, rc, seed, first_sample, n_samples, rgb_out, rgb_sumsq_out, stats); });
}
static int32_t rt_render_impl(RtScene* sp, int32_t device, const RtCamera* rc, uint64_t seed, uint64_t first_sample, uint64_t n_samples, float* rgb_out, float* rgb_sumsq_out, RtStats* stats) {
  if (!sp || !rc || !rgb_out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  Scene& s = *reinterpret_cast<Scene*>(sp);
  DeviceScene* ds = nullptr;
  int32_t st = get_device_scene(s, device, &ds);
  if (st != RT_OK) return st;
  DCamera cam;
  st = make_camera(*rc, &cam);
  if (st != RT_OK) return st;
  if (stats) std::memset(stats, 0, sizeof *stats);
  const size_t elems = (size_t)cam.W * cam.H * 3;
  const unsigned long long total = (unsigned long long)cam.W * cam.H * n_samples;
  WorkspaceLease lease;
  st = acquire_workspace(device, &lease.w);
  if (st != RT_OK) return st;
  Workspace& ws = *lease.w;
  st = ensure_workspace(ws, choose_pool(s, std::max<unsigned long long>(total, 1)), elems);
  if (st != RT_OK) return st;
  cudaStream_t stream = ws.stream;
  if (ws.pending) { RT_CUDA(cudaStreamWaitEvent(stream, ws.last_done, 0)); ws.pending = false; }
  RT_CUDA(cudaMemsetAsync(ws.sum, 0, elems * sizeof(float), stream));
  if (rgb_sumsq_out) RT_CUDA(cudaMemsetAsync(ws.sumsq, 0, elems * sizeof(float), stream));
  if (n_samples) {
    st = run_wavefront(s, *ds, ws, cam, seed, first_sample, n_samples, ws.sum, rgb_sumsq_out ? ws.sumsq : nullptr, stream, stats);
    if (st != RT_OK) return st;
  }
  const cudaEvent_t c0 = ws.copy_ev[0], c1 = ws.copy_ev[1];
  k_scale<<<ds->sm_count * 4, 256, 0, stream>>>(ws.sum, ws.out, (int)elems, 1.0f / (float)cam.spp);   // Ray.hs:232
  RT_CUDA(cudaEventRecord(c0, stream));
  // page-locked caller buffers (rt_host_alloc, …) are written by the copy engine directly; pageable ones go
  // device → pinned staging (full PCIe rate, asynchronous) → one host memcpy
  const bool direct = is_pinned_host(rgb_out, elems * sizeof(float)) && (!rgb_sumsq_out || is_pinned_host(rgb_sumsq_out, elems * sizeof(float)));
  if (!direct && !ws.h_out) RT_CUDA(cudaMallocHost(reinterpret_cast<void**>(&ws.h_out), 2 * ws.accum_elems * sizeof(float)));
  float* h_rgb = direct ? rgb_out : ws.h_out;
  float* h_sq = direct ? rgb_sumsq_out : ws.h_out + ws.accum_elems;
  RT_CUDA(cudaMemcpyAsync(h_rgb, ws.out, elems * sizeof(float), cudaMemcpyDeviceToHost, stream));
  if (rgb_sumsq_out) RT_CUDA(cudaMemcpyAsync(h_sq, ws.sumsq, elems * sizeof(float), cudaMemcpyDeviceToHost, stream));
  RT_CUDA(cudaEventRecord(c1, stream));
  RT_CUDA(cudaStreamSynchronize(stream));
  if (!direct) {
    std::memcpy(rgb_out, ws.h_out, elems * sizeof(float));
    if (rgb_sumsq_out) std::memcpy(rgb_sumsq_out, ws.h_out + ws.accum_elems, elems * sizeof(float));
  }
  if (stats) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c0, c1);
    stats->ms_copy = ms;
    stats->launches += 1;
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

static int32_t rt_probe_primary_impl(RtScene* sp, int32_t device, const RtCamera* rc, double time, uint32_t* ids, float* ts);
int32_t rt_probe_primary(RtScene* sp, int32_t device, const RtCamera* rc, double time, uint32_t* ids, float* ts) {
  return abi_guard("rt_probe_primary", [&]() { return rt_probe_primary_impl(sp, device, rc, time, ids, ts); });
}
static int32_t rt_probe_primary_impl(RtScene* sp, int32_t device, const RtCamera* rc, double time, uint32_t* ids, float* ts) {
  if (!sp || !rc || !ids || !ts) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  Scene& s = *reinterpret_cast<Scene*>(sp);
  DeviceScene* ds = nullptr;
  int32_t st = get_device_scene(s, device, &ds);
  if (st != RT_OK) return st;
  DCamera cam;
  st = make_camera(*rc, &cam);
  if (st != RT_OK) return st;
  WorkspaceLease lease;
  st = acquire_workspace(device, &lease.w);
  if (st != RT_OK) return st;
  cudaStream_t stream = lease.w->stream;
  const size_t n = (size_t)cam.W * cam.H;
  DevBuf b_ids, b_ts;
  RT_CUDA(cudaMalloc(&b_ids.p, n * 4));
  RT_CUDA(cudaMalloc(&b_ts.p, n * 4));
  uint32_t* d_ids = static_cast<uint32_t*>(b_ids.p);
  float* d_ts = static_cast<float*>(b_ts.p);
  k_probe<<<ds->sm_count * 8, kTraverseBlock, (size_t)(s.flat.max_depth + 1) * kTraverseBlock * sizeof(uint32_t), stream>>>(ds->d, cam, (float)time, d_ids, d_ts);
  cudaError_t e = cudaMemcpyAsync(ids, d_ids, n * 4, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ts, d_ts, n * 4, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  RT_CUDA(e);
  return RT_OK;
}

static int32_t rt_probe_surface_impl(RtScene* sp, int32_t device, const RtCamera* rc, double time, float* out);
int32_t rt_probe_surface(RtScene* sp, int32_t device, const RtCamera* rc, double time, float* out) {
  return abi_guard("rt_probe_surface", [&]() { return rt_probe_surface_impl(sp, device, rc, time, out); });
}
static int32_t rt_probe_surface_impl(RtScene* sp, int32_t device, const RtCamera* rc, double time, float* out) {
  if (!sp || !rc || !out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  Scene& s = *reinterpret_cast<Scene*>(sp);
  DeviceScene* ds = nullptr;
  int32_t st = get_device_scene(s, device, &ds);
  if (st != RT_OK) return st;
  DCamera cam;
  st = make_camera(*rc, &cam);
  if (st != RT_OK) return st;
  WorkspaceLease lease;
  st = acquire_workspace(device, &lease.w);
  if (st != RT_OK) return st;
  cudaStream_t stream = lease.w->stream;
  const size_t n = (size_t)cam.W * cam.H * 10;
  DevBuf b_out;
  RT_CUDA(cudaMalloc(&b_out.p, n * 4));
  float* d_out = static_cast<float*>(b_out.p);
  k_probe_surface<<<ds->sm_count * 8, kTraverseBlock, (size_t)(s.flat.max_depth + 1) * kTraverseBlock * sizeof(uint32_t), stream>>>(ds->d, cam, (float)time, d_out);
  cudaError_t e = cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  RT_CUDA(e);
  return RT_OK;
}

int32_t rt_reduce_image(int32_t device, const void* multicast_ptr, const void* const* peer_ptrs, int32_t n_peers,
                        float* d_out, uint64_t n_floats, float scale, void* cuda_stream) {
  if (!d_out || (!multicast_ptr && (!peer_ptrs || n_peers <= 0))) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  if (n_peers > kMaxPeers) { set_error("rt_reduce_image: at most 16 peers"); return RT_ERR_INVALID_ARG; }
  if (n_floats % 4 != 0 || (reinterpret_cast<uintptr_t>(d_out) & 15u) || (reinterpret_cast<uintptr_t>(multicast_ptr) & 15u)) {
    set_error("rt_reduce_image: buffers must be 16-byte aligned and hold a multiple of 4 floats (pad the image)");
    return RT_ERR_INVALID_ARG;
  }
  int32_t st = check_device(device);
  if (st != RT_OK) return st;
  DeviceGuard guard;
  RT_CUDA(cudaSetDevice(device));
  if (n_floats == 0) return RT_OK;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  const size_t n4 = n_floats / 4;
  int sm = 0;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
  const int grid = (int)std::min<size_t>((n4 + 255) / 256, (size_t)sm * 8);
  if (multicast_ptr) {
    k_reduce_multimem<<<grid, 256, 0, stream>>>(static_cast<const float4*>(multicast_ptr), reinterpret_cast<float4*>(d_out), n4, scale);
  } else {
    PeerPtrs pp{};
    for (int r = 0; r < n_peers; r++) {
      if (!peer_ptrs[r] || (reinterpret_cast<uintptr_t>(peer_ptrs[r]) & 15u)) { set_error("rt_reduce_image: bad peer pointer"); return RT_ERR_INVALID_ARG; }
      pp.p[r] = static_cast<const float4*>(peer_ptrs[r]);
    }
    k_reduce_peers<<<grid, 256, 0, stream>>>(pp, n_peers, reinterpret_cast<float4*>(d_out), n4, scale);
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

static int32_t rt_tonemap_u8_impl(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out);
int32_t rt_tonemap_u8(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out) {
  return abi_guard("rt_tonemap_u8", [&]() { return rt_tonemap_u8_impl(device, rgb, n_pixels, mode, out); });
}
static int32_t rt_tonemap_u8_impl(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out) {
  if (!rgb || !out || n_pixels < 0) { set_error("bad argument"); return RT_ERR_INVALID_ARG; }
  int32_t st = check_device(device);
  if (st != RT_OK) return st;
  DeviceGuard guard;
  RT_CUDA(cudaSetDevice(device));
  const size_t n = (size_t)n_pixels * 3;
  if (n == 0) return RT_OK;
  WorkspaceLease lease;            // the workspace's own stream: no legacy default stream, no implicit device-wide sync
  st = acquire_workspace(device, &lease.w);
  if (st != RT_OK) return st;
  cudaStream_t stream = lease.w->stream;
  if (lease.w->pending) { RT_CUDA(cudaStreamWaitEvent(stream, lease.w->last_done, 0)); lease.w->pending = false; }
  DevBuf b_in, b_out;
  RT_CUDA(cudaMalloc(&b_in.p, n * 4));
  RT_CUDA(cudaMalloc(&b_out.p, n));
  int sm = 148;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
  const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)sm * 8);
  RT_CUDA(cudaMemcpyAsync(b_in.p, rgb, n * 4, cudaMemcpyHostToDevice, stream));
  k_tonemap<<<grid, 256, 0, stream>>>(static_cast<const float*>(b_in.p), (int)n, mode, static_cast<uint8_t*>(b_out.p));
  RT_CUDA(cudaMemcpyAsync(out, b_out.p, n, cudaMemcpyDeviceToHost, stream));
  RT_CUDA(cudaStreamSynchronize(stream));
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

// ---- allocation-free tone mapping (no counterpart in the reference) ------
// rt_tonemap_u8 stages the image through two cudaMalloc'ed buffers per call.  Some callers cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it), so this variant stages through a FILE-SCOPE
// static device array instead: every GPU that loads the library holds kStaticTonemapBytes (30 MiB) until exit.
// Layout: [0, 12 * kStaticTonemapMaxPixels) the linear float image, then the 8-bit result.  In HOST code the array's name is
// only a host-side stand-in, not a device address: the address a kernel may use comes from cudaGetSymbolAddress, asked
// once per device with that device current (static_scratch_for, below).
constexpr size_t kStaticTonemapMaxPixels = RT_TONEMAP_STATIC_MAX_PIXELS;            // 2 Mi pixels: covers 1920 x 1080
constexpr size_t kStaticTonemapOutOffset = kStaticTonemapMaxPixels * 3 * sizeof(float);   // a multiple of 256
constexpr size_t kStaticTonemapBytes = kStaticTonemapOutOffset + kStaticTonemapMaxPixels * 3;
static __device__ __align__(256) unsigned char g_tonemap_scratch[kStaticTonemapBytes];

}  // extern "C"

namespace rtb {

// k_tonemap on the static scratch: reads n floats at `scratch`, writes n bytes at `scratch + kStaticTonemapOutOffset`
__global__ void __launch_bounds__(256) k_tonemap_static(unsigned char* __restrict__ scratch, int n, int mode) {
  const float* __restrict__ rgb = reinterpret_cast<const float*>(scratch);
  uint8_t* __restrict__ out = scratch + kStaticTonemapOutOffset;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    float x = rgb[k];
    x = fminf(1.f, fmaxf(0.f, x));     // NaN -> 0
    const float y = mode == 1 ? sqrtf(x) : (x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 1.f / 2.4f) - 0.055f);
    out[k] = (uint8_t)min(255, (int)floorf(y * 256.f));      // the reference encoder's quantiser (see k_tonemap)
  }
}

// device address of g_tonemap_scratch on `device` (which must be current): resolved once per device, kept in a fixed table
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[kStaticMaxDevices] = {};
static int g_static_sms[kStaticMaxDevices] = {};
static int32_t static_scratch_for(int device, unsigned char** ptr, int* sms) {
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    RT_CUDA(cudaGetSymbolAddress(&p, g_tonemap_scratch));
    int major = 0, sm = 0;
    RT_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    RT_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device));
    if (major < 10) { set_error("device " + std::to_string(device) + " is not sm_100 (kernels are built for sm_100a only)"); return RT_ERR_NO_DEVICE; }
    g_static_sms[device] = sm;
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *ptr = g_static_ptr[device];
  *sms = g_static_sms[device];
  return RT_OK;
}

}  // namespace rtb

extern "C" {

static int32_t rt_tonemap_u8_static_impl(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out, void* cuda_stream);
int32_t rt_tonemap_u8_static(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out, void* cuda_stream) {
  return abi_guard("rt_tonemap_u8_static", [&]() { return rt_tonemap_u8_static_impl(device, rgb, n_pixels, mode, out, cuda_stream); });
}
static int32_t rt_tonemap_u8_static_impl(int32_t device, const float* rgb, int32_t n_pixels, int32_t mode, uint8_t* out, void* cuda_stream) {
  if (!rgb || !out || n_pixels < 0) { set_error("rt_tonemap_u8_static: bad argument"); return RT_ERR_INVALID_ARG; }
  if ((size_t)n_pixels > kStaticTonemapMaxPixels) {
    set_error("rt_tonemap_u8_static: " + std::to_string(n_pixels) + " pixels exceed the static scratch array (at most " +
              std::to_string(kStaticTonemapMaxPixels) + " pixels per call); use rt_tonemap_u8, which allocates, or split the image");
    return RT_ERR_INVALID_ARG;
  }
  int visible = 0;
  if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
    cudaGetLastError();
    set_error("no CUDA device visible: raytrace_b200 runs only on an sm_100 GPU (there is no CPU fallback)");
    return RT_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= visible || device >= kStaticMaxDevices) { set_error("device index out of range"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  RT_CUDA(cudaSetDevice(device));
  unsigned char* scratch = nullptr;
  int sms = 0;
  int32_t st = static_scratch_for(device, &scratch, &sms);
  if (st != RT_OK) return st;
  const size_t n = (size_t)n_pixels * 3;
  if (n == 0) return RT_OK;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  // nothing in the array has to start at zero: [0, 4n) and the n result bytes are overwritten before they are read, and
  // nothing beyond them is touched, so what an earlier (larger) call left behind is never seen
  RT_CUDA(cudaMemcpyAsync(scratch, rgb, n * sizeof(float), cudaMemcpyDefault, stream));
  const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)sms * 8);
  k_tonemap_static<<<grid, 256, 0, stream>>>(scratch, (int)n, mode);
  RT_CUDA(cudaGetLastError());
  RT_CUDA(cudaMemcpyAsync(out, scratch + kStaticTonemapOutOffset, n, cudaMemcpyDefault, stream));
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
  if (cap == cudaStreamCaptureStatusNone) RT_CUDA(cudaStreamSynchronize(stream));    // a capturing stream must not be synchronised
  return RT_OK;
}

}  // extern "C"

#include "rt_multi.cuh"   // rt_context_* / rt_render_multi (uses the static helpers above)
