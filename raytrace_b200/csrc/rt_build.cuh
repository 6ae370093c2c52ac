// rt_build.cuh — device-side binned-SAH build of the surface tree.
//
// The reference builds its BVH on the host, recursively, by object median along the longest axis (Geometry.hs:369-377,
// Core.hs:43-46,143-144).  Of that tree the device needs one thing only: the order in which its leaves are evaluated (who wins
// an exact tie) - the host's reference-policy pass keeps that (rt_flatten.cpp::number).  The tree the GPU WALKS is a binned-SAH
// tree (16 centroid bins x 3 axes, surface-area cost, the cost model of rt_flatten.cpp::sah_node), and this file builds it on
// the GPU in ONE cooperative launch: a persistent grid (at most one block per 256 items, at most what is co-resident on the 148
// SMs) walks the levels, grid-wide barriers between the three phases of a level, and stops on the device when a level produced
// no node - no host round trip, no launch per level (a 1200-triangle pawn spent 0.8 ms in 39 launches before).
//
//   bin phase      one thread per item: centroid -> bin on each axis, box / centroid bounds / counts accumulated per bin.
//                  A block's 256 consecutive positions mostly belong to ONE node in the upper levels: its 48 bins live in
//                  shared memory and are flushed with at most 48 x 14 global atomics per block (the root level of a 1 M
//                  triangle mesh would otherwise send 42 M atomics to 48 addresses).
//   split phase    one thread per node of the level: sweeps the 3 x 15 split candidates, decides leaf / bin split / position
//                  split (coincident centroids, or too deep for the 32-entry traversal stack), allocates the node PAIR
//                  (rt_device.h: siblings adjacent, one 64-byte fetch tests both) and the two child build nodes, whose boxes
//                  and centroid bounds come straight from the bins; writes its own padded box into its parent's pair.
//   part phase     one thread per position: moves the item to its child's segment (warp-aggregated slot atomics).
//   leaf phase     one thread per leaf: primitives of a leaf in evaluation order (= item order), leaf reference
//                  kLeafBit | first << 5 | (count - 1), and the permutation provisional slot -> storage slot.
//
// HBM-bound integer / compare work: nothing here is GEMM-shaped.  Per level the algorithmic traffic is one 32-byte item
// record + 8 bytes of index state per item in the bin phase and again in the part phase (80 B / item / level); a 16-level bunny
// build moves ~13 MB, far below what the launch latencies cost (DESIGN.md §6c has the measured split).
#pragma once
#include <cooperative_groups.h>

namespace rtb {
namespace gpubuild {

constexpr int kBins = 16;
constexpr uint32_t kPending = 0u, kBinSplit = 1u, kPosSplit = 2u, kLeaf = 3u;
constexpr int kMaxLevels = 40;           // the depth rule below keeps every tree under kStackDepth (32) levels
constexpr int kBlock = 256;

struct BNode {          // one node under construction, 128 bytes
  uint32_t lo, hi;      // its items are positions [lo, hi) of the index array
  uint32_t slot;        // DNode index of its record inside its parent's pair (kNone: the root has no record, rt_device.h)
  uint32_t depth;
  int bx[6];            // box of its items: lo xyz, hi xyz (order-preserving int encoding of the floats)
  int cb[6];            // bounds of its items' centroids
  uint32_t n_prims, cost;
  uint32_t kind, axis, bin, nl;   // decision of k_build_split; nl = items going left
  uint32_t child;       // build-node id of the left child (right = child + 1)
  uint32_t bslot;       // its 48 bins: slot in the bank of its level's parity
  uint32_t fill[2];     // k_build_part: next free position of the left / right segment
  uint32_t pad_[6];
};
static_assert(sizeof(BNode) == 128, "BNode layout");

struct Bin {            // 64 bytes
  int bx[6];
  int cb[6];
  uint32_t cnt, cost;   // items; cost = 1 per primitive item, 2 per aggregated cuboid (a slab test ~ two planar tests)
  uint32_t pad_[2];
};
static_assert(sizeof(Bin) == 64, "Bin layout");

struct BuildCtl {
  uint32_t n_inner;       // inner nodes so far: build nodes = 1 + 2 n_inner, node pairs = n_inner
  uint32_t unused_;
  uint32_t max_depth;
  uint32_t bslots[2];     // bin slots handed out in each bank
  uint32_t n_leaves;
  uint32_t overflow;      // a leaf of more than kLeafCap primitives (cannot happen with collapse <= 32; checked by the host)
  uint32_t levels;        // levels the build took
  uint32_t level_begin[kMaxLevels + 1], level_end[kMaxLevels + 1];
};

struct BuildParams {
  uint32_t n_items;
  uint32_t first;         // storage slot of the tree's first primitive
  uint32_t collapse;
  float node_cost;
  uint32_t bank_slots;    // bin slots per bank
  uint32_t node_base;     // DNode index of dnodes[0] (the device tree is stored behind the host-built media trees)
};

__device__ __forceinline__ int enc(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__device__ __forceinline__ float dec(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }
constexpr int kEncMax = 0x7FFFFFFF, kEncMin = (int)0x80000000;

// the ONE binning formula (k_build_bin and k_build_part must agree bit for bit; sub, div, mul: nothing to contract)
__device__ __forceinline__ int bin_of(float c, float lo, float ext) {
  const int k = (int)((c - lo) / ext * (float)kBins);
  return min(kBins - 1, max(k, 0));
}
__device__ __forceinline__ float centre(float lo, float hi) { return 0.5f * (lo + hi); }

__device__ __forceinline__ void bin_identity(Bin& b) {
#pragma unroll
  for (int k = 0; k < 3; k++) { b.bx[k] = kEncMax; b.bx[3 + k] = kEncMin; b.cb[k] = kEncMax; b.cb[3 + k] = kEncMin; }
  b.cnt = 0; b.cost = 0; b.pad_[0] = 0; b.pad_[1] = 0;
}


__global__ void k_build_fill_bins(Bin* bins, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) bin_identity(bins[i]);
}

// accumulate one item into 12 bounds (+ 2 counters) in shared or global memory
__device__ __forceinline__ void accumulate_bounds(int* bx, int* cb, const float4 a, const float4 b) {
  atomicMin(&bx[0], enc(a.x)); atomicMin(&bx[1], enc(a.y)); atomicMin(&bx[2], enc(a.z));
  atomicMax(&bx[3], enc(b.x)); atomicMax(&bx[4], enc(b.y)); atomicMax(&bx[5], enc(b.z));
  const int cx = enc(centre(a.x, b.x)), cy = enc(centre(a.y, b.y)), cz = enc(centre(a.z, b.z));
  atomicMin(&cb[0], cx); atomicMin(&cb[1], cy); atomicMin(&cb[2], cz);
  atomicMax(&cb[3], cx); atomicMax(&cb[4], cy); atomicMax(&cb[5], cz);
}
__device__ __forceinline__ void accumulate(int* bx, int* cb, uint32_t* cnt, uint32_t* cost, const float4 a, const float4 b, uint32_t n) {
  accumulate_bounds(bx, cb, a, b);
  atomicAdd(cnt, 1u);
  atomicAdd(cost, n == 6u ? 2u : 1u);
}

struct BuildArgs {
  const float4* items;     // BuildItem records as two float4 (read-only for the whole kernel: __ldg)
  uint32_t* idx[2];        // position -> item, double buffered (the part phase writes the other one)
  uint32_t* node_of[2];    // position -> build node (kNone once its leaf is final)
  uint32_t* off;           // position -> primitive offset (scenes with six-primitive items), else nullptr
  uint32_t* perm;          // out: storage slot (relative to P.first) -> provisional slot
  BNode* nodes;
  Bin* bins;               // two banks of P.bank_slots x 48
  BuildCtl* ctl;
  DNode* dnodes;           // out: node pairs, dnodes[0] = DNode index P.node_base
  BuildParams P;
};

// ---- phase 0: identity permutation, root bounds ------------------------------------------------------------------------
__device__ void phase_init(const BuildArgs& A, Bin& sb) {
  const uint32_t n = A.P.n_items;
  for (uint32_t p0 = blockIdx.x * kBlock; p0 < n; p0 += gridDim.x * kBlock) {
    if (threadIdx.x == 0) bin_identity(sb);
    __syncthreads();
    const uint32_t p = p0 + threadIdx.x;
    if (p < n) {
      A.idx[0][p] = p;
      A.node_of[0][p] = 0;
      const float4 a = __ldg(A.items + 2 * p), b = __ldg(A.items + 2 * p + 1);
      accumulate(sb.bx, sb.cb, &sb.cnt, &sb.cost, a, b, __float_as_uint(b.w));
    }
    __syncthreads();
    if (threadIdx.x == 0 && sb.cnt) {
      BNode& r = A.nodes[0];
      for (int k = 0; k < 3; k++) {
        atomicMin(&r.bx[k], sb.bx[k]); atomicMax(&r.bx[3 + k], sb.bx[3 + k]);
        atomicMin(&r.cb[k], sb.cb[k]); atomicMax(&r.cb[3 + k], sb.cb[3 + k]);
      }
      atomicAdd(&r.cost, sb.cost);
      atomicAdd(&r.n_prims, sb.cnt + 5u * (sb.cost - sb.cnt));     // items are 1 primitive (cost 1) or 6 (cost 2)
    }
    __syncthreads();
  }
}

// ---- bin phase ---------------------------------------------------------------------------------------------------------
__device__ void phase_bin(const BuildArgs& A, int cur, uint32_t level, Bin* sb, uint32_t& s_primary) {
  const BuildParams& P = A.P;
  const uint32_t* idx = A.idx[cur];
  const uint32_t* node_of = A.node_of[cur];
  Bin* bank = A.bins + (size_t)(level & 1u) * P.bank_slots * (3 * kBins);
  for (uint32_t p0 = blockIdx.x * kBlock; p0 < P.n_items; p0 += gridDim.x * kBlock) {
    const uint32_t p = p0 + threadIdx.x;
    if (threadIdx.x == 0) s_primary = node_of[p0];
    if (threadIdx.x < 3 * kBins) bin_identity(sb[threadIdx.x]);
    __syncthreads();
    const uint32_t primary = s_primary;
    if (p < P.n_items) {
      const uint32_t nd = node_of[p];
      if (nd != kNone) {
        const BNode& N = A.nodes[nd];
        if (N.hi - N.lo >= 2u) {
          const uint32_t it = idx[p];
          const float4 a = __ldg(A.items + 2 * it), b = __ldg(A.items + 2 * it + 1);
          const float c[3] = {centre(a.x, b.x), centre(a.y, b.y), centre(a.z, b.z)};
          Bin* dst = nd == primary ? sb : bank + (size_t)N.bslot * (3 * kBins);
#pragma unroll
          for (int ax = 0; ax < 3; ax++) {
            const float lo = dec(N.cb[ax]), ext = dec(N.cb[3 + ax]) - lo;
            if (!(ext > 0.f) || !isfinite(ext)) continue;
            Bin& B = dst[ax * kBins + bin_of(c[ax], lo, ext)];
            accumulate(B.bx, B.cb, &B.cnt, &B.cost, a, b, __float_as_uint(b.w));
          }
        }
      }
    }
    __syncthreads();
    if (threadIdx.x < 3 * kBins && primary != kNone) {
      const Bin& S = sb[threadIdx.x];
      if (S.cnt) {
        Bin& B = bank[(size_t)A.nodes[primary].bslot * (3 * kBins) + threadIdx.x];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          atomicMin(&B.bx[k], S.bx[k]); atomicMax(&B.bx[3 + k], S.bx[3 + k]);
          atomicMin(&B.cb[k], S.cb[k]); atomicMax(&B.cb[3 + k], S.cb[3 + k]);
        }
        atomicAdd(&B.cnt, S.cnt);
        atomicAdd(&B.cost, S.cost);
      }
    }
    __syncthreads();
  }
}

struct Agg {            // running union while sweeping bins
  float lo[3], hi[3];
  uint32_t cnt, cost;
  __device__ void clear() { for (int k = 0; k < 3; k++) { lo[k] = INFINITY; hi[k] = -INFINITY; } cnt = 0; cost = 0; }
  __device__ void join(const Bin& b) {
    if (!b.cnt) return;
    for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], dec(b.bx[k])); hi[k] = fmaxf(hi[k], dec(b.bx[3 + k])); }
    cnt += b.cnt; cost += b.cost;
  }
  __device__ float half_area() const {
    if (!cnt) return 0.f;
    const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    return x * y + y * z + z * x;
  }
};

// node box -> DNode record: padded like rt_flatten.cpp::write_node (the slab test runs in fp32), rounded outward
__device__ __forceinline__ void write_box(DNode& d, const int* bx) {
  float lo[3], hi[3], m = 0.f;
  for (int k = 0; k < 3; k++) { lo[k] = dec(bx[k]); hi[k] = dec(bx[3 + k]); m = fmaxf(m, fmaxf(fabsf(lo[k]), fabsf(hi[k]))); }
  const float pad = __fmaf_ru(1e-5f, m, 1e-6f);
  for (int k = 0; k < 3; k++) { d.lo[k] = __fsub_rd(lo[k], pad); d.hi[k] = __fadd_ru(hi[k], pad); }
}

__device__ __forceinline__ void init_child(BNode& C, uint32_t lo, uint32_t hi, uint32_t slot, uint32_t depth) {
  C.lo = lo; C.hi = hi; C.slot = slot; C.depth = depth;
  for (int k = 0; k < 3; k++) { C.bx[k] = kEncMax; C.bx[3 + k] = kEncMin; C.cb[k] = kEncMax; C.cb[3 + k] = kEncMin; }
  C.n_prims = 0; C.cost = 0; C.kind = kPending; C.axis = 0; C.bin = 0; C.nl = 0; C.child = kNone; C.bslot = kNone;
  C.fill[0] = 0; C.fill[1] = 0;
}


// ---- split phase -------------------------------------------------------------------------------------------------------
// Two variants of the same decision.  Wide levels (more nodes than twice the warps of the grid): one THREAD per node, the bins
// swept sequentially - thousands of nodes hide each other's latency.  Narrow levels (the top of every tree; all of a small
// tree): one WARP per node, lane k holding bin k, the 15 candidate splits of an axis evaluated at once from a prefix and a
// suffix scan - a lone thread sweeping 48 bins out of L2 took ~50 us per level, which was the whole build time of a pawn.
struct Part {            // union of some bins (ints: the order-preserving encoding, so min / max are exact)
  int bx[6], cb[6];
  uint32_t cnt, cost;
};
__device__ __forceinline__ void part_identity(Part& p) {
#pragma unroll
  for (int k = 0; k < 3; k++) { p.bx[k] = kEncMax; p.bx[3 + k] = kEncMin; p.cb[k] = kEncMax; p.cb[3 + k] = kEncMin; }
  p.cnt = 0; p.cost = 0;
}
__device__ __forceinline__ float enc_half_area(const int* bx, uint32_t cnt) {
  if (!cnt) return 0.f;
  const float x = dec(bx[3]) - dec(bx[0]), y = dec(bx[4]) - dec(bx[1]), z = dec(bx[5]) - dec(bx[2]);
  return x * y + y * z + z * x;
}

// decision shared by both variants (rt_flatten.cpp::sah_node): kind and, for a position split, nl
__device__ __forceinline__ uint32_t decide(const BNode& N, const BuildParams& P, bool sah_allowed, bool have, float best, uint32_t& nl) {
  const uint32_t n_items = N.hi - N.lo;
  if (n_items == 1u) return kLeaf;
  uint32_t kind = kPending;
  if (sah_allowed) {
    const float area = enc_half_area(N.bx, 1u);
    const bool must_split = N.n_prims > P.collapse || N.slot == kNone;     // (the root is a pair: FlatScene::surface_root)
    if (have && area > 0.f && (must_split || P.node_cost + best / area < (float)N.cost)) kind = kBinSplit;
    else if (!must_split) kind = kLeaf;
  }
  if (kind == kPending) {    // coincident centroids, or too deep: halve the segment as it stands
    if (N.n_prims <= P.collapse && N.slot != kNone) kind = kLeaf;
    else { kind = kPosSplit; nl = n_items / 2u; }
  }
  return kind;
}

// allocate the pair and the two children of an inner node (one thread)
__device__ __forceinline__ void make_children(const BuildArgs& A, uint32_t level, BNode& N, uint32_t kind, uint32_t axis, uint32_t bin, uint32_t nl,
                                              const Part* L, const Part* R) {
  BuildCtl* ctl = A.ctl;
  // Two independent atomics (their round trips overlap): the inner node's rank gives both its pair (node_base + 2 rank) and its
  // children's build-node ids (1 + 2 rank, 2 + 2 rank); one add hands out the bin slots of the children that need any.
  const uint32_t want = (nl >= 2u ? 1u : 0u) + ((N.hi - N.lo) - nl >= 2u ? 1u : 0u);
  const uint32_t rank = atomicAdd(&ctl->n_inner, 1u);
  uint32_t slot = want ? atomicAdd(&ctl->bslots[(level + 1u) & 1u], want) : 0u;
  const uint32_t pair = A.P.node_base + 2u * rank, child = 1u + 2u * rank;
  if (N.slot != kNone) { A.dnodes[N.slot - A.P.node_base].a = pair; A.dnodes[N.slot - A.P.node_base].b = 0u; }
  BNode C[2];
  init_child(C[0], N.lo, N.lo + nl, pair, N.depth + 1u);
  init_child(C[1], N.lo + nl, N.hi, pair + 1u, N.depth + 1u);
  if (L && R) {
    const Part* src[2] = {L, R};
    for (int s = 0; s < 2; s++) {
      for (int q = 0; q < 6; q++) { C[s].bx[q] = src[s]->bx[q]; C[s].cb[q] = src[s]->cb[q]; }
      C[s].cost = src[s]->cost;
      C[s].n_prims = src[s]->cnt + 5u * (src[s]->cost - src[s]->cnt);
    }
  }                          // (position split: the part phase accumulates the children's bounds)
  for (int s = 0; s < 2; s++)
    if (C[s].hi - C[s].lo >= 2u) C[s].bslot = slot++;
  A.nodes[child] = C[0];
  A.nodes[child + 1u] = C[1];
  N.kind = kind; N.axis = axis; N.bin = bin; N.nl = nl; N.child = child;
  N.fill[0] = 0; N.fill[1] = 0;
}

__device__ void phase_split_thread(const BuildArgs& A, uint32_t level, uint32_t begin, uint32_t end) {
  const BuildParams& P = A.P;
  Bin* bank = A.bins + (size_t)(level & 1u) * P.bank_slots * (3 * kBins);
  for (uint32_t nd = begin + blockIdx.x * kBlock + threadIdx.x; nd < end; nd += gridDim.x * kBlock) {
    BNode& N = A.nodes[nd];
    const uint32_t n_items = N.hi - N.lo;
    atomicMax(&A.ctl->max_depth, N.depth);
    if (N.slot != kNone) write_box(A.dnodes[N.slot - P.node_base], N.bx);
    Bin* B = n_items >= 2u ? bank + (size_t)N.bslot * (3 * kBins) : nullptr;
    uint32_t log2n = 0;
    while ((1u << log2n) < n_items) log2n++;
    // The traversal stack holds kStackDepth entries: SAH may split as unevenly as it likes only while a balanced split of
    // what is left would still fit (the rule of rt_flatten.cpp::sah_node).
    const bool sah_allowed = n_items >= 2u && N.depth + 1u + log2n < 30u;
    float best = INFINITY;
    uint32_t best_axis = 0, best_bin = 0;
    bool have = false;
    if (sah_allowed) {
      for (int ax = 0; ax < 3; ax++) {
        const float ext = dec(N.cb[3 + ax]) - dec(N.cb[ax]);
        if (!(ext > 0.f) || !isfinite(ext)) continue;
        const Bin* X = B + ax * kBins;
        float ra[kBins], rn[kBins];
        Agg r;
        r.clear();
        for (int k = kBins - 1; k > 0; k--) { r.join(X[k]); ra[k] = r.half_area(); rn[k] = (float)r.cost; }
        Agg l;
        l.clear();
        for (int k = 0; k < kBins - 1; k++) {
          l.join(X[k]);
          if (l.cost == 0u || rn[k + 1] == 0.f) continue;
          const float c = l.half_area() * (float)l.cost + ra[k + 1] * rn[k + 1];
          if (c < best) { best = c; best_axis = (uint32_t)ax; best_bin = (uint32_t)k; have = true; }
        }
      }
    }
    uint32_t nl = 0;
    const uint32_t kind = decide(N, P, sah_allowed, have, best, nl);
    if (kind == kLeaf) {
      N.kind = kLeaf;
      atomicAdd(&A.ctl->n_leaves, 1u);
    } else if (kind == kPosSplit) {
      make_children(A, level, N, kind, 0u, 0u, nl, nullptr, nullptr);
    } else {
      Part L, R;
      part_identity(L);
      part_identity(R);
      const Bin* X = B + best_axis * kBins;
      for (uint32_t k = 0; k < (uint32_t)kBins; k++) {
        const Bin& b = X[k];
        if (!b.cnt) continue;
        Part& C = k <= best_bin ? L : R;
        for (int q = 0; q < 3; q++) {
          C.bx[q] = min(C.bx[q], b.bx[q]); C.bx[3 + q] = max(C.bx[3 + q], b.bx[3 + q]);
          C.cb[q] = min(C.cb[q], b.cb[q]); C.cb[3 + q] = max(C.cb[3 + q], b.cb[3 + q]);
        }
        C.cnt += b.cnt; C.cost += b.cost;
      }
      make_children(A, level, N, kind, best_axis, best_bin, L.cnt, &L, &R);
    }
    if (B) for (int k = 0; k < 3 * kBins; k++) if (B[k].cnt) bin_identity(B[k]);     // the bank is clean again for level + 2
  }
}

__device__ void phase_split_warp(const BuildArgs& A, uint32_t level, uint32_t begin, uint32_t end) {
  const BuildParams& P = A.P;
  constexpr unsigned kAll = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * kBlock + threadIdx.x) >> 5, n_warps = gridDim.x * (kBlock / 32);
  Bin* bank = A.bins + (size_t)(level & 1u) * P.bank_slots * (3 * kBins);
  for (uint32_t nd = begin + warp; nd < end; nd += n_warps) {
    BNode& N = A.nodes[nd];
    const uint32_t n_items = N.hi - N.lo;
    if (lane == 0) {
      atomicMax(&A.ctl->max_depth, N.depth);
      if (N.slot != kNone) write_box(A.dnodes[N.slot - P.node_base], N.bx);
    }
    uint32_t log2n = 0;
    while ((1u << log2n) < n_items) log2n++;
    const bool sah_allowed = n_items >= 2u && N.depth + 1u + log2n < 30u;
    float best = INFINITY;
    uint32_t best_key = 0xFFFFFFFFu;
    Part mine[3];                      // lane k < 16: bin k of each axis (the three loads are in flight together)
#pragma unroll
    for (int ax = 0; ax < 3; ax++) part_identity(mine[ax]);
    if (n_items >= 2u && lane < kBins) {
      Bin* b0 = bank + (size_t)N.bslot * (3 * kBins) + lane;
      int4 w[3][4];
#pragma unroll
      for (int ax = 0; ax < 3; ax++) {
        const int4* src = reinterpret_cast<const int4*>(b0 + ax * kBins);
#pragma unroll
        for (int q = 0; q < 4; q++) w[ax][q] = src[q];
      }
#pragma unroll
      for (int ax = 0; ax < 3; ax++) {
        Part& m = mine[ax];
        m.bx[0] = w[ax][0].x; m.bx[1] = w[ax][0].y; m.bx[2] = w[ax][0].z; m.bx[3] = w[ax][0].w;
        m.bx[4] = w[ax][1].x; m.bx[5] = w[ax][1].y; m.cb[0] = w[ax][1].z; m.cb[1] = w[ax][1].w;
        m.cb[2] = w[ax][2].x; m.cb[3] = w[ax][2].y; m.cb[4] = w[ax][2].z; m.cb[5] = w[ax][2].w;
        m.cnt = (uint32_t)w[ax][3].x; m.cost = (uint32_t)w[ax][3].y;
        if (m.cnt) bin_identity(b0[ax * kBins]);       // the bank is clean again for level + 2
      }
    }
#pragma unroll
    for (int ax = 0; ax < 3; ax++) {
      const float ext = dec(N.cb[3 + ax]) - dec(N.cb[ax]);
      if (!sah_allowed || !(ext > 0.f) || !isfinite(ext)) continue;
      // inclusive prefix (left part up to bin k) and suffix (right part from bin k) of box and cost
      int l[6], r[6];
      uint32_t lc = mine[ax].cost, rc = mine[ax].cost, ln = mine[ax].cnt, rn = mine[ax].cnt;
#pragma unroll
      for (int q = 0; q < 6; q++) { l[q] = mine[ax].bx[q]; r[q] = mine[ax].bx[q]; }
#pragma unroll
      for (int d = 1; d < kBins; d <<= 1) {
        int t[6], u[6];
#pragma unroll
        for (int q = 0; q < 6; q++) { t[q] = __shfl_up_sync(kAll, l[q], d); u[q] = __shfl_down_sync(kAll, r[q], d); }
        const uint32_t tc = __shfl_up_sync(kAll, lc, d), tn = __shfl_up_sync(kAll, ln, d);
        const uint32_t uc = __shfl_down_sync(kAll, rc, d), un = __shfl_down_sync(kAll, rn, d);
        if (lane >= d) {
#pragma unroll
          for (int q = 0; q < 3; q++) { l[q] = min(l[q], t[q]); l[3 + q] = max(l[3 + q], t[3 + q]); }
          lc += tc; ln += tn;
        }
        if (lane + d < 32) {           // lanes >= 16 hold the identity: the suffix of lane k < 16 covers bins k..15
#pragma unroll
          for (int q = 0; q < 3; q++) { r[q] = min(r[q], u[q]); r[3 + q] = max(r[3 + q], u[3 + q]); }
          rc += uc; rn += un;
        }
      }
      // candidate k: bins 0..k left, k+1..15 right
      const float la = enc_half_area(l, ln);
      const float ra = __shfl_down_sync(kAll, enc_half_area(r, rn), 1);
      const uint32_t rcost = __shfl_down_sync(kAll, rc, 1);
      if (lane < kBins - 1 && lc != 0u && rcost != 0u) {
        const float c = la * (float)lc + ra * (float)rcost;
        if (c < best) { best = c; best_key = (uint32_t)(ax * kBins + lane); }
      }
    }
    // first minimum in (axis, bin) order, as the sequential sweep finds it
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      const float ob = __shfl_xor_sync(kAll, best, d);
      const uint32_t ok = __shfl_xor_sync(kAll, best_key, d);
      if (ob < best || (ob == best && ok < best_key)) { best = ob; best_key = ok; }
    }
    const bool have = best_key != 0xFFFFFFFFu;
    uint32_t nl = 0;
    const uint32_t kind = decide(N, P, sah_allowed, have, best, nl);
    if (kind == kLeaf) {
      if (lane == 0) { N.kind = kLeaf; atomicAdd(&A.ctl->n_leaves, 1u); }
      continue;
    }
    if (kind == kPosSplit) {
      if (lane == 0) make_children(A, level, N, kind, 0u, 0u, nl, nullptr, nullptr);
      continue;
    }
    const uint32_t axis = best_key / kBins, bin = best_key % kBins;
    Part m = axis == 0u ? mine[0] : axis == 1u ? mine[1] : mine[2];
    Part L = m, R = m;
    if (lane > (int)bin) part_identity(L);
    if (lane <= (int)bin || lane >= kBins) part_identity(R);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
      for (int q = 0; q < 3; q++) {
        L.bx[q] = min(L.bx[q], __shfl_xor_sync(kAll, L.bx[q], d)); L.bx[3 + q] = max(L.bx[3 + q], __shfl_xor_sync(kAll, L.bx[3 + q], d));
        L.cb[q] = min(L.cb[q], __shfl_xor_sync(kAll, L.cb[q], d)); L.cb[3 + q] = max(L.cb[3 + q], __shfl_xor_sync(kAll, L.cb[3 + q], d));
        R.bx[q] = min(R.bx[q], __shfl_xor_sync(kAll, R.bx[q], d)); R.bx[3 + q] = max(R.bx[3 + q], __shfl_xor_sync(kAll, R.bx[3 + q], d));
        R.cb[q] = min(R.cb[q], __shfl_xor_sync(kAll, R.cb[q], d)); R.cb[3 + q] = max(R.cb[3 + q], __shfl_xor_sync(kAll, R.cb[3 + q], d));
      }
      L.cnt += __shfl_xor_sync(kAll, L.cnt, d); L.cost += __shfl_xor_sync(kAll, L.cost, d);
      R.cnt += __shfl_xor_sync(kAll, R.cnt, d); R.cost += __shfl_xor_sync(kAll, R.cost, d);
    }
    if (lane == 0) make_children(A, level, N, kind, axis, bin, L.cnt, &L, &R);
  }
}

__device__ __forceinline__ void phase_split(const BuildArgs& A, uint32_t level, uint32_t begin, uint32_t end) {
  if (end - begin > 2u * gridDim.x * (kBlock / 32)) phase_split_thread(A, level, begin, end);
  else phase_split_warp(A, level, begin, end);
}

// ---- part phase --------------------------------------------------------------------------------------------------------
__device__ void phase_part(const BuildArgs& A, int cur, uint32_t level) {
  const BuildParams& P = A.P;
  const uint32_t* idx = A.idx[cur];
  const uint32_t* node_of = A.node_of[cur];
  uint32_t* idx_out = A.idx[cur ^ 1];
  uint32_t* node_out = A.node_of[cur ^ 1];
  for (uint32_t p0 = blockIdx.x * kBlock; p0 < P.n_items; p0 += gridDim.x * kBlock) {
    const uint32_t p = p0 + threadIdx.x;
    bool move = false;
    uint32_t key = 0, it = 0, nd = kNone, side = 0;
    if (p < P.n_items) {
      nd = node_of[p];
      it = idx[p];
      if (nd != kNone) {
        const BNode& N = A.nodes[nd];
        if (N.kind == kLeaf) nd = kNone;
        else if (N.kind == kBinSplit) {
          const float4 a = __ldg(A.items + 2 * it), b = __ldg(A.items + 2 * it + 1);
          const float c = N.axis == 0 ? centre(a.x, b.x) : N.axis == 1 ? centre(a.y, b.y) : centre(a.z, b.z);
          const float lo = dec(N.cb[N.axis]), ext = dec(N.cb[3 + N.axis]) - lo;
          side = (uint32_t)bin_of(c, lo, ext) > N.bin ? 1u : 0u;
          move = true;
          key = nd * 2u + side;
        } else {                // position split: the item stays where it is; its child learns its bounds here
          side = (p - N.lo) >= N.nl ? 1u : 0u;
          BNode& C = A.nodes[N.child + side];
          const float4 a = __ldg(A.items + 2 * it), b = __ldg(A.items + 2 * it + 1);
          const uint32_t n = __float_as_uint(b.w);
          accumulate_bounds(C.bx, C.cb, a, b);
          atomicAdd(&C.cost, n == 6u ? 2u : 1u);
          atomicAdd(&C.n_prims, n);
          idx_out[p] = it;
          node_out[p] = N.child + side;
        }
      }
      if (nd == kNone) { idx_out[p] = it; node_out[p] = kNone; }       // finished (leaf) positions carry through
    }
    // one slot atomic per (node, side) group of the warp
    const unsigned movers = __ballot_sync(0xFFFFFFFFu, move);
    if (move) {
      const unsigned peers = __match_any_sync(movers, key);
      const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
      const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
      BNode& N = A.nodes[nd];
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(&N.fill[side], (uint32_t)__popc(peers));
      base = __shfl_sync(peers, base, leader);
      const uint32_t dest = (side ? N.lo + N.nl : N.lo) + base + rank;
      idx_out[dest] = it;
      node_out[dest] = N.child + side;
    }
  }
}

// ---- exclusive prefix of the items' primitive counts in position order (block 0; scenes with aggregated cuboids) --------
__device__ void phase_scan(const BuildArgs& A, int cur, uint32_t* s_warp, uint32_t& s_carry) {
  const uint32_t n = A.P.n_items;
  const uint32_t* idx = A.idx[cur];
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n; base += kBlock) {
    const uint32_t p = base + threadIdx.x;
    const uint32_t v = p < n ? __float_as_uint(__ldg(A.items + 2 * idx[p] + 1).w) : 0u;
    uint32_t x = v;
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d); if (lane >= d) x += y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    uint32_t before_warps = 0;
    for (int w = 0; w < warp; w++) before_warps += s_warp[w];
    const uint32_t before = s_carry + before_warps + x - v;
    if (p < n) A.off[p] = before;
    __syncthreads();
    if (threadIdx.x == kBlock - 1) s_carry = before + v;
    __syncthreads();
  }
}

// ---- leaf phase ----------------------------------------------------------------------------------------------------------
__device__ void phase_leaf(const BuildArgs& A, int cur, uint32_t n_nodes) {
  const BuildParams& P = A.P;
  uint32_t* idx = A.idx[cur];
  for (uint32_t nd = blockIdx.x * kBlock + threadIdx.x; nd < n_nodes; nd += gridDim.x * kBlock) {
    const BNode& N = A.nodes[nd];
    if (N.kind != kLeaf) continue;
    // evaluation order inside the leaf = item order (the items were listed in the reference's evaluation order)
    for (uint32_t i = N.lo + 1u; i < N.hi; i++) {
      const uint32_t v = idx[i];
      uint32_t j = i;
      while (j > N.lo && idx[j - 1u] > v) { idx[j] = idx[j - 1u]; j--; }
      idx[j] = v;
    }
    const uint32_t rel = A.off ? A.off[N.lo] : N.lo;
    uint32_t count = 0;
    for (uint32_t i = N.lo; i < N.hi; i++) {
      const uint32_t it = idx[i];
      const uint32_t prov = __float_as_uint(__ldg(A.items + 2 * it).w), n = __float_as_uint(__ldg(A.items + 2 * it + 1).w);
      for (uint32_t k = 0; k < n; k++) A.perm[rel + count + k] = prov + k;
      count += n;
    }
    if (count > kLeafCap || count == 0u) { atomicAdd(&A.ctl->overflow, 1u); continue; }
    if (N.slot != kNone) {
      A.dnodes[N.slot - P.node_base].a = kLeafBit | ((P.first + rel) << kLeafShift) | (count - 1u);
      A.dnodes[N.slot - P.node_base].b = count;
    }
  }
}

// The whole build: ONE cooperative launch (the host has put the root node and the control block in place).
__global__ void __launch_bounds__(kBlock) k_build_tree(BuildArgs A) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ Bin sb[3 * kBins];
  __shared__ uint32_t s_word[kBlock / 32 + 2];
  volatile BuildCtl* vctl = A.ctl;
  phase_init(A, sb[0]);
  grid.sync();
  int cur = 0;
  uint32_t level = 0;
  for (; level < (uint32_t)kMaxLevels; level++) {
    const uint32_t begin = vctl->level_begin[level], end = vctl->level_end[level];
    if (begin == end) break;                                   // the previous level produced no node: the tree is closed
    phase_bin(A, cur, level, sb, s_word[0]);
    grid.sync();
    phase_split(A, level, begin, end);
    grid.sync();
    if (blockIdx.x == 0 && threadIdx.x == 0) {                 // publish the next level's node range (read after the next barrier)
      vctl->level_begin[level + 1] = end;
      vctl->level_end[level + 1] = 1u + 2u * vctl->n_inner;
      vctl->bslots[level & 1u] = 0u;
    }
    phase_part(A, cur, level);
    cur ^= 1;
    grid.sync();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) vctl->levels = level;
  if (A.off) {
    if (blockIdx.x == 0) phase_scan(A, cur, s_word + 1, s_word[0]);
    grid.sync();
  }
  phase_leaf(A, cur, 1u + 2u * vctl->n_inner);
}

}  // namespace gpubuild
}  // namespace rtb
