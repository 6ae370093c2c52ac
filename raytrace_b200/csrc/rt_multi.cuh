// rt_multi.cuh — multi-GPU rendering INSIDE the library (included at the end of rt_kernels.cu).
//
// The reference's `raytrace` is one call that uses the whole machine (Ray.hs:121, parallel array :238).  Here one
// RtContext owns a set of GPUs in ONE process: rt_render_multi splits the samples per pixel across them (global sample
// ranges, SURVEY 8e: the Philox counter carries the global sample index, so the sample set does not depend on the GPU
// count), renders every share on its own device from its own host thread and stream, and then assembles the image with a
// REDUCE-SCATTER over NVLink written here, not with ncclReduce:
//   * every device accumulates into a buffer of its own that is bound to an NVSwitch multicast object (cuMemCreate +
//     cuMulticastCreate / AddDevice / BindMem, driver entry points resolved at run time - the library has no link-time
//     dependency on libcuda, so it still loads on a machine without a driver);
//   * once all shares are done (cross-device stream waits on per-device events: no host barrier, no flags), EVERY device
//     reduces its own 1/G slice of the image with multimem.ld_reduce (the switch adds the G copies in the fabric),
//     normalises it by 1/spp (Ray.hs:232) and DMA-copies that slice straight into the caller's host image - so the
//     read-back shrinks with G as well as the render;
//   * without multicast support the slices are summed with plain peer loads (cudaDeviceEnablePeerAccess), and without
//     peer access the shares are gathered on the first device with cudaMemcpyPeerAsync and summed there.
// The committed scene is uploaded once per device and stays resident (Scene::devices).
#pragma once
#include <cuda.h>

#include <thread>

namespace rtb {

// ---- driver entry points (resolved through the runtime: no -lcuda) ---------------------------------------------------
struct DriverApi {
  bool ok = false;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
};

static const DriverApi& driver_api() {
  static DriverApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    bool all = true;
    auto get = [&](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
      if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !*fn) {
        cudaGetLastError();
        *fn = nullptr;
        all = false;
      }
    };
    get("cuMemGetAllocationGranularity", (void**)&api.MemGetAllocationGranularity);
    get("cuMemCreate", (void**)&api.MemCreate);
    get("cuMemRelease", (void**)&api.MemRelease);
    get("cuMemAddressReserve", (void**)&api.MemAddressReserve);
    get("cuMemAddressFree", (void**)&api.MemAddressFree);
    get("cuMemMap", (void**)&api.MemMap);
    get("cuMemUnmap", (void**)&api.MemUnmap);
    get("cuMemSetAccess", (void**)&api.MemSetAccess);
    get("cuMulticastCreate", (void**)&api.MulticastCreate);
    get("cuMulticastAddDevice", (void**)&api.MulticastAddDevice);
    get("cuMulticastBindMem", (void**)&api.MulticastBindMem);
    get("cuMulticastUnbind", (void**)&api.MulticastUnbind);
    get("cuMulticastGetGranularity", (void**)&api.MulticastGetGranularity);
    get("cuDeviceGet", (void**)&api.DeviceGet);
    get("cuDeviceGetAttribute", (void**)&api.DeviceGetAttribute);
    api.ok = all;
  });
  return api;
}

enum ReduceMode : int32_t { kReduceNone = 0, kReduceMulticast = 1, kReducePeer = 2, kReduceGather = 3 };

struct CtxDevice {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;          // this device's share has been rendered into its accumulator
  cudaEvent_t t0 = nullptr, t1 = nullptr;   // device time of the share (RtStats.ms_total = max over devices)
  float* acc = nullptr;                // un-normalised sums [sum | sumsq], 2 * n_pad floats (this device's own mapping)
  float* mc = nullptr;                 // multicast mapping of all devices' accumulators, as seen from this device
  float* out = nullptr;                // this device's normalised slice (cudaMalloc, 2 * slice floats)
  float* gather = nullptr;             // kReduceGather, first device only: G accumulators side by side
  CUmemGenericAllocationHandle mem = 0;
  CUdeviceptr acc_va = 0, mc_va = 0;
  bool acc_mapped = false, mc_mapped = false, mc_bound = false;
};

struct Context {
  std::vector<CtxDevice> dev;
  std::mutex mu;                       // one render at a time per context
  ReduceMode mode = kReduceNone;
  std::string mode_why;
  size_t n_pad = 0;                    // capacity: floats per accumulator half (multiple of 4 * G)
  size_t map_bytes = 0;                // mapped size of one accumulator (granularity aligned)
  CUmemGenericAllocationHandle mc_handle = 0;
  bool mc_created = false;
  float* h_stage = nullptr;            // pinned staging for pageable destinations
  size_t h_stage_floats = 0;
};

static void context_free_buffers(Context& c) {
  const DriverApi& D = driver_api();
  for (CtxDevice& d : c.dev) {
    cudaSetDevice(d.device);
    cudaStreamSynchronize(d.stream);
    if (d.mc_mapped) { D.MemUnmap(d.mc_va, c.map_bytes); d.mc_mapped = false; }
    if (d.mc_va) { D.MemAddressFree(d.mc_va, c.map_bytes); d.mc_va = 0; }
    if (d.mc_bound) {
      CUdevice cd;
      if (D.DeviceGet(&cd, d.device) == CUDA_SUCCESS) D.MulticastUnbind(c.mc_handle, cd, 0, c.map_bytes);
      d.mc_bound = false;
    }
    if (d.acc_mapped) { D.MemUnmap(d.acc_va, c.map_bytes); d.acc_mapped = false; }
    if (d.acc_va) { D.MemAddressFree(d.acc_va, c.map_bytes); d.acc_va = 0; }
    if (d.mem) { D.MemRelease(d.mem); d.mem = 0; }
    else if (d.acc) cudaFree(d.acc);
    d.acc = d.mc = nullptr;
    if (d.out) { cudaFree(d.out); d.out = nullptr; }
    if (d.gather) { cudaFree(d.gather); d.gather = nullptr; }
  }
  if (c.mc_created) { D.MemRelease(c.mc_handle); c.mc_created = false; c.mc_handle = 0; }
  c.n_pad = 0;
  c.map_bytes = 0;
}

// accumulators bound to one multicast object; false (with everything released) when any step fails
static bool context_alloc_multicast(Context& c, size_t bytes, std::string& why) {
  const DriverApi& D = driver_api();
  if (!D.ok) { why = "driver entry points for cuMem / cuMulticast not available"; return false; }
  const int G = (int)c.dev.size();
  for (int a = 0; a < G; a++)
    for (int b = a + 1; b < G; b++)
      if (c.dev[a].device == c.dev[b].device) { why = "a device is listed more than once"; return false; }
  std::vector<CUdevice> cu(G);
  for (int g = 0; g < G; g++) {
    if (D.DeviceGet(&cu[g], c.dev[g].device) != CUDA_SUCCESS) { why = "cuDeviceGet failed"; return false; }
    int sup = 0;
    if (D.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cu[g]) != CUDA_SUCCESS || !sup) {
      why = "device " + std::to_string(c.dev[g].device) + " has no multicast support";
      return false;
    }
  }
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)G;
  mp.handleTypes = 0;          // CU_MEM_HANDLE_TYPE_NONE: one process, nothing is exported
  mp.flags = 0;
  mp.size = bytes;
  size_t gran = 0;
  if (D.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) { why = "cuMulticastGetGranularity failed"; return false; }
  for (int g = 0; g < G; g++) {      // the physical allocations must satisfy their own granularity too
    CUmemAllocationProp ap{};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = c.dev[g].device;
    size_t ag = 0;
    if (D.MemGetAllocationGranularity(&ag, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && ag > gran) gran = ag;
  }
  const size_t size = (bytes + gran - 1) / gran * gran;
  mp.size = size;
  c.map_bytes = size;
  auto fail = [&](const std::string& w, CUresult r) { why = w + " (CUresult " + std::to_string((int)r) + ")"; context_free_buffers(c); return false; };
  CUresult r = D.MulticastCreate(&c.mc_handle, &mp);
  if (r != CUDA_SUCCESS) return fail("cuMulticastCreate", r);
  c.mc_created = true;
  for (int g = 0; g < G; g++) {
    r = D.MulticastAddDevice(c.mc_handle, cu[g]);
    if (r != CUDA_SUCCESS) return fail("cuMulticastAddDevice", r);
  }
  for (int g = 0; g < G; g++) {
    CtxDevice& d = c.dev[g];
    cudaSetDevice(d.device);
    CUmemAllocationProp ap{};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d.device;
    r = D.MemCreate(&d.mem, size, &ap, 0);
    if (r != CUDA_SUCCESS) return fail("cuMemCreate", r);
    r = D.MemAddressReserve(&d.acc_va, size, gran, 0, 0);
    if (r != CUDA_SUCCESS) return fail("cuMemAddressReserve", r);
    r = D.MemMap(d.acc_va, size, 0, d.mem, 0);
    if (r != CUDA_SUCCESS) return fail("cuMemMap", r);
    d.acc_mapped = true;
    CUmemAccessDesc ad{};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = d.device;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = D.MemSetAccess(d.acc_va, size, &ad, 1);
    if (r != CUDA_SUCCESS) return fail("cuMemSetAccess", r);
    r = D.MulticastBindMem(c.mc_handle, 0, d.mem, 0, size, 0);
    if (r != CUDA_SUCCESS) return fail("cuMulticastBindMem", r);
    d.mc_bound = true;
    d.acc = reinterpret_cast<float*>(d.acc_va);
  }
  for (int g = 0; g < G; g++) {      // every device maps the multicast object into its own address range
    CtxDevice& d = c.dev[g];
    cudaSetDevice(d.device);
    r = D.MemAddressReserve(&d.mc_va, size, gran, 0, 0);
    if (r != CUDA_SUCCESS) return fail("cuMemAddressReserve (multicast)", r);
    r = D.MemMap(d.mc_va, size, 0, c.mc_handle, 0);
    if (r != CUDA_SUCCESS) return fail("cuMemMap (multicast)", r);
    d.mc_mapped = true;
    CUmemAccessDesc ad{};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = d.device;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = D.MemSetAccess(d.mc_va, size, &ad, 1);
    if (r != CUDA_SUCCESS) return fail("cuMemSetAccess (multicast)", r);
    d.mc = reinterpret_cast<float*>(d.mc_va);
  }
  return true;
}

// (re)allocates the context's image buffers for accumulators of n_pad floats per half
static int32_t context_ensure(Context& c, size_t n_pad) {
  if (c.n_pad >= n_pad) return RT_OK;
  context_free_buffers(c);
  const int G = (int)c.dev.size();
  const size_t bytes = 2 * n_pad * sizeof(float);
  c.mode = kReduceNone;
  c.mode_why.clear();
  if (G > 1) {
    std::string why;
    if (context_alloc_multicast(c, bytes, why)) c.mode = kReduceMulticast;
    else c.mode_why = why;
  }
  if (c.mode != kReduceMulticast) {
    bool peers = G > 1;
    for (int a = 0; a < G && peers; a++)
      for (int b = 0; b < G && peers; b++) {
        if (c.dev[a].device == c.dev[b].device) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, c.dev[a].device, c.dev[b].device) != cudaSuccess || !can) peers = false;
      }
    if (peers) {
      for (int a = 0; a < G; a++) {
        cudaSetDevice(c.dev[a].device);
        for (int b = 0; b < G; b++) {
          if (c.dev[a].device == c.dev[b].device) continue;
          cudaError_t e = cudaDeviceEnablePeerAccess(c.dev[b].device, 0);
          if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) peers = false;
          cudaGetLastError();
        }
      }
    }
    c.mode = G == 1 ? kReduceNone : (peers ? kReducePeer : kReduceGather);
    for (CtxDevice& d : c.dev) {
      RT_CUDA(cudaSetDevice(d.device));
      RT_CUDA(cudaMalloc((void**)&d.acc, bytes));
    }
    if (c.mode == kReduceGather) {
      RT_CUDA(cudaSetDevice(c.dev[0].device));
      RT_CUDA(cudaMalloc((void**)&c.dev[0].gather, bytes * G));
    }
  }
  const size_t slice = n_pad / G;
  for (int g = 0; g < G; g++) {
    CtxDevice& d = c.dev[g];
    RT_CUDA(cudaSetDevice(d.device));
    const size_t out_floats = (c.mode == kReduceGather) ? (g == 0 ? 2 * n_pad : 0) : 2 * slice;
    if (out_floats) RT_CUDA(cudaMalloc((void**)&d.out, out_floats * sizeof(float)));
  }
  c.n_pad = n_pad;
  return RT_OK;
}

// slice [off, off + n) of the image: out[k] = scale * sum over devices of acc_g[off + k]
static void launch_slice_reduce(Context& c, int g, size_t base_floats, size_t off, size_t n, float* out, float scale) {
  CtxDevice& d = c.dev[g];
  const size_t n4 = n / 4;
  if (n4 == 0) return;
  const int grid = (int)std::min<size_t>((n4 + 255) / 256, (size_t)d.sm_count * 8);
  if (c.mode == kReduceMulticast) {
    k_reduce_multimem<<<grid, 256, 0, d.stream>>>(reinterpret_cast<const float4*>(d.mc + base_floats + off), reinterpret_cast<float4*>(out), n4, scale);
  } else {
    PeerPtrs pp{};
    const int G = (int)c.dev.size();
    for (int r = 0; r < G; r++)
      pp.p[r] = reinterpret_cast<const float4*>((c.mode == kReduceGather ? c.dev[0].gather + (size_t)r * 2 * c.n_pad : c.dev[r].acc) + base_floats + off);
    k_reduce_peers<<<grid, 256, 0, d.stream>>>(pp, G, reinterpret_cast<float4*>(out), n4, scale);
  }
}

}  // namespace rtb

using namespace rtb;

extern "C" {

static int32_t context_create_impl(const int32_t* devices, int32_t n_devices, RtContext** out);
int32_t rt_context_create(const int32_t* devices, int32_t n_devices, RtContext** out) {
  try {
    return context_create_impl(devices, n_devices, out);
  } catch (const std::exception& e) {
    set_error(std::string("rt_context_create: ") + e.what());
    if (out) *out = nullptr;
    return RT_ERR_CUDA;
  }
}
static int32_t context_create_impl(const int32_t* devices, int32_t n_devices, RtContext** out) {
  if (!out || n_devices < 0 || (n_devices > 0 && !devices)) { set_error("rt_context_create: bad argument"); return RT_ERR_INVALID_ARG; }
  *out = nullptr;
  DeviceGuard guard;
  int visible = 0;
  if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
    cudaGetLastError();
    set_error("no CUDA device visible: raytrace_b200 renders only on sm_100 GPUs (there is no CPU fallback)");
    return RT_ERR_NO_DEVICE;
  }
  std::vector<int> ids;
  if (n_devices == 0) {           // every visible sm_100 device
    for (int i = 0; i < visible; i++) if (check_device(i) == RT_OK) ids.push_back(i);
    if (ids.empty()) { set_error("no sm_100 device visible"); return RT_ERR_NO_DEVICE; }
  } else {
    for (int i = 0; i < n_devices; i++) {
      int32_t st = check_device(devices[i]);
      if (st != RT_OK) return st;
      ids.push_back(devices[i]);      // a device may be listed more than once: every entry is a share with its own stream
                                      // (no multicast then - peer-load reduce; lets a one-GPU box run the G > 1 path)
    }
  }
  if ((int)ids.size() > kMaxPeers) { set_error("rt_context_create: at most 16 devices"); return RT_ERR_INVALID_ARG; }
  auto* c = new (std::nothrow) Context();
  if (!c) { set_error("out of memory"); return RT_ERR_INVALID_ARG; }
  struct Owner {          // a throwing allocation below must not leak the half-built context
    Context* c;
    ~Owner() { if (c) rt_context_destroy(reinterpret_cast<RtContext*>(c)); }
  } owner{c};
  c->dev.resize(ids.size());
  cudaError_t e = cudaSuccess;
  for (size_t g = 0; g < ids.size() && e == cudaSuccess; g++) {
    CtxDevice& d = c->dev[g];
    d.device = ids[g];
    e = cudaSetDevice(d.device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, d.device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d.done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreate(&d.t0);
    if (e == cudaSuccess) e = cudaEventCreate(&d.t1);
  }
  if (e != cudaSuccess) {
    set_error(std::string("rt_context_create: ") + cudaGetErrorString(e));
    return RT_ERR_CUDA;            // ~Owner destroys the streams / events created so far
  }
  owner.c = nullptr;
  *out = reinterpret_cast<RtContext*>(c);
  return RT_OK;
}

int32_t rt_context_destroy(RtContext* cp) {
  if (!cp) return RT_OK;
  DeviceGuard guard;
  Context* c = reinterpret_cast<Context*>(cp);
  context_free_buffers(*c);
  for (CtxDevice& d : c->dev) {
    cudaSetDevice(d.device);
    if (d.done) cudaEventDestroy(d.done);
    if (d.t0) cudaEventDestroy(d.t0);
    if (d.t1) cudaEventDestroy(d.t1);
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  if (c->h_stage) cudaFreeHost(c->h_stage);
  delete c;
  return RT_OK;
}

int32_t rt_context_info(const RtContext* cp, RtContextInfo* info) {
  if (!cp || !info) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  const Context* c = reinterpret_cast<const Context*>(cp);
  std::memset(info, 0, sizeof *info);
  info->n_devices = (int32_t)c->dev.size();
  for (size_t g = 0; g < c->dev.size() && g < 16; g++) info->devices[g] = c->dev[g].device;
  info->reduce_mode = c->mode;
  std::snprintf(info->reduce_note, sizeof info->reduce_note, "%s", c->mode_why.c_str());
  return RT_OK;
}

static int32_t render_multi_impl(RtContext* cp, RtScene* sp, const RtCamera* rc, uint64_t seed, float* rgb_out, float* rgb_sumsq_out,
                                 RtStats* stats);

// no C++ exception crosses the ABI: thread creation (std::system_error) and allocations (std::bad_alloc) are caught here
int32_t rt_render_multi(RtContext* cp, RtScene* sp, const RtCamera* rc, uint64_t seed, float* rgb_out, float* rgb_sumsq_out,
                        RtStats* stats) {
  try {
    return render_multi_impl(cp, sp, rc, seed, rgb_out, rgb_sumsq_out, stats);
  } catch (const std::exception& e) {
    set_error(std::string("rt_render_multi: ") + e.what());
    return RT_ERR_CUDA;
  } catch (...) {
    set_error("rt_render_multi: unknown C++ exception");
    return RT_ERR_CUDA;
  }
}

static int32_t render_multi_impl(RtContext* cp, RtScene* sp, const RtCamera* rc, uint64_t seed, float* rgb_out, float* rgb_sumsq_out,
                                 RtStats* stats) {
  if (!cp || !sp || !rc || !rgb_out) { set_error("null argument"); return RT_ERR_INVALID_ARG; }
  DeviceGuard guard;
  Context& c = *reinterpret_cast<Context*>(cp);
  Scene& s = *reinterpret_cast<Scene*>(sp);
  std::lock_guard<std::mutex> lock(c.mu);
  DCamera cam;
  int32_t st = make_camera(*rc, &cam);
  if (st != RT_OK) return st;
  if (stats) std::memset(stats, 0, sizeof *stats);
  const int G = (int)c.dev.size();
  const size_t elems = (size_t)cam.W * cam.H * 3;
  const size_t unit = 4 * (size_t)G;                         // every slice a whole number of float4
  const size_t n_pad = (elems + unit - 1) / unit * unit;
  st = context_ensure(c, n_pad);
  if (st != RT_OK) return st;
  const size_t cap = c.n_pad;                                // accumulator layout: [sum: cap floats | sumsq: cap floats]
  const bool want_sq = rgb_sumsq_out != nullptr;
  const uint64_t spp = (uint64_t)cam.spp;

  // ---- shares: one host thread per device (scene upload on first use, then the wavefront loop) ------------------------
  std::vector<int32_t> status(G, RT_OK);
  std::vector<std::string> errors(G);
  std::vector<RtStats> part(G);
  auto work = [&](int g) {
    CtxDevice& d = c.dev[g];
    auto run_unchecked = [&]() -> int32_t {
      DeviceScene* ds = nullptr;
      int32_t r = get_device_scene(s, d.device, &ds);       // sets the device
      if (r != RT_OK) return r;
      const uint64_t base = spp / G, rem = spp % G;
      const uint64_t count = base + ((uint64_t)g < rem ? 1 : 0), first = (uint64_t)g * base + std::min<uint64_t>(g, rem);
      RT_CUDA(cudaEventRecord(d.t0, d.stream));
      RT_CUDA(cudaMemsetAsync(d.acc, 0, (want_sq ? 2 * cap : cap) * sizeof(float), d.stream));
      if (count) {
        WorkspaceLease lease;
        r = acquire_workspace(d.device, &lease.w);
        if (r != RT_OK) return r;
        r = ensure_workspace(*lease.w, choose_pool(s, (unsigned long long)cam.W * cam.H * count), 0);
        if (r != RT_OK) return r;
        if (lease.w->pending) { RT_CUDA(cudaStreamWaitEvent(d.stream, lease.w->last_done, 0)); lease.w->pending = false; }
        r = run_wavefront(s, *ds, *lease.w, cam, seed, first, count, d.acc, want_sq ? d.acc + cap : nullptr, d.stream, stats ? &part[g] : nullptr);
        if (r != RT_OK) return r;
      }
      RT_CUDA(cudaEventRecord(d.t1, d.stream));
      RT_CUDA(cudaEventRecord(d.done, d.stream));
      return RT_OK;
    };
    auto run = [&]() -> int32_t {          // an exception must not leave a worker thread (std::terminate)
      try { return run_unchecked(); }
      catch (const std::exception& e) { set_error(std::string("worker: ") + e.what()); return RT_ERR_CUDA; }
      catch (...) { set_error("worker: unknown C++ exception"); return RT_ERR_CUDA; }
    };
    status[g] = run();
    if (status[g] != RT_OK) errors[g] = rt_last_error();     // the message is thread-local: carry it to the caller
  };
  if (G == 1) work(0);
  else {
    std::vector<std::thread> threads;
    threads.reserve(G);
    try {
      for (int g = 0; g < G; g++) threads.emplace_back(work, g);
    } catch (...) {                        // could not start a thread: wait for the ones that run, then report
      for (auto& t : threads) t.join();
      for (CtxDevice& d : c.dev) { cudaSetDevice(d.device); cudaStreamSynchronize(d.stream); }
      throw;
    }
    for (auto& t : threads) t.join();
  }
  for (int g = 0; g < G; g++)
    if (status[g] != RT_OK) {
      for (CtxDevice& d : c.dev) { cudaSetDevice(d.device); cudaStreamSynchronize(d.stream); }
      set_error("device " + std::to_string(c.dev[g].device) + ": " + errors[g]);
      return status[g];
    }

  // ---- reduce-scatter + read-back: device g owns slice g of the image ---------------------------------------------------
  const float scale = 1.0f / (float)cam.spp;                  // Ray.hs:232
  const size_t slice = n_pad / G;
  const size_t img_bytes = elems * sizeof(float);
  const bool direct = is_pinned_host(rgb_out, img_bytes) && (!want_sq || is_pinned_host(rgb_sumsq_out, img_bytes));
  if (!direct && c.h_stage_floats < 2 * elems) {
    if (c.h_stage) cudaFreeHost(c.h_stage);
    c.h_stage = nullptr;
    c.h_stage_floats = 0;
    RT_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c.h_stage), 2 * elems * sizeof(float)));
    c.h_stage_floats = 2 * elems;
  }
  float* h_rgb = direct ? rgb_out : c.h_stage;
  float* h_sq = direct ? rgb_sumsq_out : c.h_stage + elems;
  if (c.mode == kReduceGather) {      // no peer access: gather the shares on the first device, sum there
    CtxDevice& d0 = c.dev[0];
    RT_CUDA(cudaSetDevice(d0.device));
    for (int g = 0; g < G; g++) {
      RT_CUDA(cudaStreamWaitEvent(d0.stream, c.dev[g].done, 0));
      RT_CUDA(cudaMemcpyPeerAsync(d0.gather + (size_t)g * 2 * cap, d0.device, c.dev[g].acc, c.dev[g].device, (want_sq ? 2 * cap : cap) * sizeof(float), d0.stream));
    }
    launch_slice_reduce(c, 0, 0, 0, n_pad, d0.out, scale);
    RT_CUDA(cudaMemcpyAsync(h_rgb, d0.out, img_bytes, cudaMemcpyDeviceToHost, d0.stream));
    if (want_sq) {
      launch_slice_reduce(c, 0, cap, 0, n_pad, d0.out + n_pad, 1.0f);
      RT_CUDA(cudaMemcpyAsync(h_sq, d0.out + n_pad, img_bytes, cudaMemcpyDeviceToHost, d0.stream));
    }
  } else {
    for (int g = 0; g < G; g++) {
      CtxDevice& d = c.dev[g];
      RT_CUDA(cudaSetDevice(d.device));
      for (int h = 0; h < G; h++) if (h != g) RT_CUDA(cudaStreamWaitEvent(d.stream, c.dev[h].done, 0));
      const size_t off = (size_t)g * slice;
      const size_t n_copy = off >= elems ? 0 : std::min(slice, elems - off);   // the padding is not part of the image
      if (G == 1) {
        k_scale<<<d.sm_count * 4, 256, 0, d.stream>>>(d.acc, d.out, (int)slice, scale);
      } else launch_slice_reduce(c, g, 0, off, slice, d.out, scale);
      if (n_copy) RT_CUDA(cudaMemcpyAsync(h_rgb + off, d.out, n_copy * sizeof(float), cudaMemcpyDeviceToHost, d.stream));
      if (want_sq) {
        if (G == 1) RT_CUDA(cudaMemcpyAsync(d.out + slice, d.acc + cap, slice * sizeof(float), cudaMemcpyDeviceToDevice, d.stream));
        else launch_slice_reduce(c, g, cap, off, slice, d.out + slice, 1.0f);
        if (n_copy) RT_CUDA(cudaMemcpyAsync(h_sq + off, d.out + slice, n_copy * sizeof(float), cudaMemcpyDeviceToHost, d.stream));
      }
    }
  }
  cudaError_t e = cudaSuccess;
  for (CtxDevice& d : c.dev) {
    cudaSetDevice(d.device);
    cudaError_t e2 = cudaStreamSynchronize(d.stream);
    if (e == cudaSuccess) e = e2;
  }
  RT_CUDA(e);
  if (!direct) {
    std::memcpy(rgb_out, c.h_stage, img_bytes);
    if (want_sq) std::memcpy(rgb_sumsq_out, c.h_stage + elems, img_bytes);
  }
  if (stats) {
    for (int g = 0; g < G; g++) {
      const RtStats& p = part[g];
      stats->samples += p.samples; stats->rays += p.rays; stats->launches += p.launches + 2;
      stats->iterations = std::max(stats->iterations, p.iterations);
      stats->pool = std::max(stats->pool, p.pool);
      stats->ms_traverse = std::max(stats->ms_traverse, p.ms_traverse); stats->ms_shade = std::max(stats->ms_shade, p.ms_shade);
      stats->ms_raygen = std::max(stats->ms_raygen, p.ms_raygen); stats->ms_finish = std::max(stats->ms_finish, p.ms_finish);
      stats->rays_finish += p.rays_finish; stats->paths_finish += p.paths_finish;
      stats->node_pairs += p.node_pairs; stats->sphere_tests += p.sphere_tests; stats->planar_tests += p.planar_tests;
      stats->box_tests += p.box_tests; stats->rays_counted += p.rays_counted;
      float ms = 0;
      cudaSetDevice(c.dev[g].device);
      if (cudaEventElapsedTime(&ms, c.dev[g].t0, c.dev[g].t1) == cudaSuccess) stats->ms_total = std::max(stats->ms_total, (double)ms);
    }
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

}  // extern "C"
