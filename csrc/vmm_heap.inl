// Symmetric heap on the CUDA virtual-memory API with an NVLS multicast object.
// Included by ddlb_b200.cu. The default heap on NVSwitch systems (2, 4 and 8 GPUs:
// tests/test_vmm_multigpu.py); the cudaIpc heap in runtime.inl is the fallback.
//
// Every rank creates one physical allocation (cuMemCreate, shareable as a POSIX file
// descriptor), maps all ranks' allocations into one reserved VA range (rank r at
// base + r * bytes) and, when the devices support multicast, binds its allocation to a
// multicast object created by rank 0 and maps that too. A store to the multicast address
// lands in every rank's memory (multimem.st), a multimem.ld_reduce returns the sum over
// all ranks computed in the NVSwitch. File descriptors travel between the rank processes
// over a Unix socket (SCM_RIGHTS) - done in Python (ddlb_b200/parallel/vmm_heap.py).
//
// Replaces what the reference gets from NCCL / nvFuser's "cuda" backend with
// multicast_protocol=multimem (ddlb/primitives/TPColumnwise/fuser.py:171-206).

namespace ddlb {

struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
};

static bool load_entry(const char* name, void** out) {
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || *out == nullptr) {
    (void)cudaGetLastError();
    return false;
  }
  return true;
}

static DriverApi& driver_api() {
  static DriverApi api;
  if (!api.ok) {
    bool ok = true;
    ok &= load_entry("cuMemCreate", reinterpret_cast<void**>(&api.MemCreate));
    ok &= load_entry("cuMemRelease", reinterpret_cast<void**>(&api.MemRelease));
    ok &= load_entry("cuMemAddressReserve", reinterpret_cast<void**>(&api.MemAddressReserve));
    ok &= load_entry("cuMemAddressFree", reinterpret_cast<void**>(&api.MemAddressFree));
    ok &= load_entry("cuMemMap", reinterpret_cast<void**>(&api.MemMap));
    ok &= load_entry("cuMemUnmap", reinterpret_cast<void**>(&api.MemUnmap));
    ok &= load_entry("cuMemSetAccess", reinterpret_cast<void**>(&api.MemSetAccess));
    ok &= load_entry("cuMemExportToShareableHandle", reinterpret_cast<void**>(&api.MemExportToShareableHandle));
    ok &= load_entry("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&api.MemImportFromShareableHandle));
    ok &= load_entry("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&api.MemGetAllocationGranularity));
    ok &= load_entry("cuMulticastCreate", reinterpret_cast<void**>(&api.MulticastCreate));
    ok &= load_entry("cuMulticastAddDevice", reinterpret_cast<void**>(&api.MulticastAddDevice));
    ok &= load_entry("cuMulticastBindMem", reinterpret_cast<void**>(&api.MulticastBindMem));
    ok &= load_entry("cuMulticastGetGranularity", reinterpret_cast<void**>(&api.MulticastGetGranularity));
    ok &= load_entry("cuDeviceGetAttribute", reinterpret_cast<void**>(&api.DeviceGetAttribute));
    api.ok = ok;
  }
  return api;
}

#define DDLB_CU(expr)                                                                        \
  do {                                                                                       \
    CUresult _r = (expr);                                                                    \
    if (_r != CUDA_SUCCESS) return ::ddlb::set_error("%s failed with CUresult %d", #expr, (int)_r); \
  } while (0)

// multimem self-test: every rank writes `rank + 1` into its slice of a region through the
// multicast address (so the value appears in EVERY rank's copy), then - after the caller's
// barrier - sums a second region over all ranks with multimem.ld_reduce.
__global__ void multimem_store_kernel(float* mc, int rank, int world, int n_per_rank) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_per_rank) {
    float* dst = mc + static_cast<size_t>(rank) * n_per_rank + i;
    const float v = static_cast<float>(rank + 1);
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
  }
}

__global__ void multimem_reduce_kernel(const float* mc, float* out, int n) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc + i)
                 : "memory");
    *reinterpret_cast<float4*>(out + i) = v;
  }
}

}  // namespace ddlb

struct ddlb_vmm {
  int device = 0, rank = 0, world = 1;
  size_t bytes = 0, gran = 0;
  CUmemGenericAllocationHandle handle[DDLB_MAX_WORLD] = {};
  bool mapped[DDLB_MAX_WORLD] = {};
  CUdeviceptr va = 0;          // world * bytes reserved; rank r at va + r * bytes
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_va = 0;
  bool has_mc = false, mc_bound = false;
};

static int vmm_access(ddlb_vmm* s, CUdeviceptr ptr, size_t size) {
  using namespace ddlb;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = s->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DDLB_CU(driver_api().MemSetAccess(ptr, size, &acc, 1));
  return 0;
}

static CUmemAllocationProp vmm_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

extern "C" int ddlb_vmm_multicast_supported(int device, int* supported) {
  using namespace ddlb;
  if (supported == nullptr) return set_error("null argument");
  *supported = 0;
  DriverApi& api = driver_api();
  if (!api.ok) return set_error("CUDA driver entry points for the VMM API are unavailable");
  DDLB_CUDA(cudaSetDevice(device));
  DDLB_CUDA(cudaFree(nullptr));   // make sure the primary context exists
  DDLB_CU(api.DeviceGetAttribute(supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device));
  return 0;
}

// Step 1 (every rank): allocate, reserve the VA range, map the own allocation.
extern "C" int ddlb_vmm_create(int device, int rank, int world, int64_t bytes, ddlb_vmm** out) {
  using namespace ddlb;
  if (out == nullptr) return set_error("null argument");
  *out = nullptr;
  if (world < 1 || world > DDLB_MAX_WORLD || rank < 0 || rank >= world)
    return set_error("bad rank/world %d/%d", rank, world);
  DriverApi& api = driver_api();
  if (!api.ok) return set_error("CUDA driver entry points for the VMM API are unavailable");
  DDLB_CUDA(cudaSetDevice(device));
  DDLB_CUDA(cudaFree(nullptr));
  auto* s = new ddlb_vmm();
  s->device = device; s->rank = rank; s->world = world;
  CUmemAllocationProp prop = vmm_prop(device);
  size_t gran = 0;
  if (api.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS ||
      gran == 0) { delete s; return set_error("cuMemGetAllocationGranularity failed"); }
  int mc_ok = 0;
  api.DeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device);
  if (mc_ok && world > 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = static_cast<size_t>(bytes);
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mgran = 0;
    if (api.MulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS &&
        mgran > gran) gran = mgran;
    s->has_mc = true;
  }
  s->gran = gran;
  s->bytes = (static_cast<size_t>(bytes) + gran - 1) / gran * gran;
  CUresult r = api.MemCreate(&s->handle[rank], s->bytes, &prop, 0);
  if (r != CUDA_SUCCESS) { delete s; return set_error("cuMemCreate(%zu) failed with CUresult %d", s->bytes, (int)r); }
  r = api.MemAddressReserve(&s->va, s->bytes * world, gran, 0, 0);
  if (r != CUDA_SUCCESS) { api.MemRelease(s->handle[rank]); delete s; return set_error("cuMemAddressReserve failed with CUresult %d", (int)r); }
  const CUdeviceptr mine = s->va + static_cast<CUdeviceptr>(rank) * s->bytes;
  r = api.MemMap(mine, s->bytes, 0, s->handle[rank], 0);
  if (r != CUDA_SUCCESS) { delete s; return set_error("cuMemMap(own) failed with CUresult %d", (int)r); }
  s->mapped[rank] = true;
  if (int rc = vmm_access(s, mine, s->bytes)) { delete s; return rc; }
  DDLB_CUDA(cudaMemset(reinterpret_cast<void*>(mine), 0, s->bytes));
  DDLB_CUDA(cudaDeviceSynchronize());
  *out = s;
  return 0;
}

// Step 2: export the own allocation as a file descriptor (caller sends it to the peers and
// closes it afterwards).
extern "C" int ddlb_vmm_export_fd(ddlb_vmm* s, int* fd) {
  using namespace ddlb;
  if (s == nullptr || fd == nullptr) return set_error("null argument");
  DDLB_CU(driver_api().MemExportToShareableHandle(fd, s->handle[s->rank],
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return 0;
}

// Step 3: map a peer's allocation received as a file descriptor.
extern "C" int ddlb_vmm_import_peer(ddlb_vmm* s, int peer, int fd) {
  using namespace ddlb;
  if (s == nullptr || peer < 0 || peer >= s->world || peer == s->rank) return set_error("bad peer");
  DriverApi& api = driver_api();
  DDLB_CUDA(cudaSetDevice(s->device));
  DDLB_CU(api.MemImportFromShareableHandle(&s->handle[peer], reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  const CUdeviceptr at = s->va + static_cast<CUdeviceptr>(peer) * s->bytes;
  DDLB_CU(api.MemMap(at, s->bytes, 0, s->handle[peer], 0));
  s->mapped[peer] = true;
  return vmm_access(s, at, s->bytes);
}

// Multicast, step A (rank 0 only): create the object and export it.
extern "C" int ddlb_vmm_mc_create(ddlb_vmm* s, int* fd) {
  using namespace ddlb;
  if (s == nullptr || fd == nullptr) return set_error("null argument");
  if (!s->has_mc) return set_error("multicast is not supported on this device");
  DriverApi& api = driver_api();
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = static_cast<unsigned>(s->world);
  mp.size = s->bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DDLB_CU(api.MulticastCreate(&s->mc, &mp));
  DDLB_CU(api.MemExportToShareableHandle(fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return 0;
}

// Multicast, step A' (other ranks): import the object.
extern "C" int ddlb_vmm_mc_import(ddlb_vmm* s, int fd) {
  using namespace ddlb;
  if (s == nullptr) return set_error("null argument");
  DDLB_CU(driver_api().MemImportFromShareableHandle(&s->mc, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                    CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return 0;
}

// Multicast, step B (every rank): add the own device. ALL ranks must have done this
// before ANY rank calls ddlb_vmm_mc_bind (host barrier in between).
extern "C" int ddlb_vmm_mc_add_device(ddlb_vmm* s) {
  using namespace ddlb;
  if (s == nullptr || s->mc == 0) return set_error("no multicast object");
  DDLB_CU(driver_api().MulticastAddDevice(s->mc, s->device));
  return 0;
}

// Multicast, step C (every rank): bind the own allocation and map the multicast range.
extern "C" int ddlb_vmm_mc_bind(ddlb_vmm* s) {
  using namespace ddlb;
  if (s == nullptr || s->mc == 0) return set_error("no multicast object");
  DriverApi& api = driver_api();
  DDLB_CUDA(cudaSetDevice(s->device));
  DDLB_CU(api.MulticastBindMem(s->mc, 0, s->handle[s->rank], 0, s->bytes, 0));
  DDLB_CU(api.MemAddressReserve(&s->mc_va, s->bytes, s->gran, 0, 0));
  DDLB_CU(api.MemMap(s->mc_va, s->bytes, 0, s->mc, 0));
  if (int rc = vmm_access(s, s->mc_va, s->bytes)) return rc;
  s->mc_bound = true;
  return 0;
}

extern "C" void* ddlb_vmm_base(ddlb_vmm* s, int rank) {
  if (s == nullptr || rank < 0 || rank >= s->world || !s->mapped[rank]) return nullptr;
  return reinterpret_cast<void*>(s->va + static_cast<CUdeviceptr>(rank) * s->bytes);
}
extern "C" void* ddlb_vmm_mc_base(ddlb_vmm* s) {
  return (s != nullptr && s->mc_bound) ? reinterpret_cast<void*>(s->mc_va) : nullptr;
}
extern "C" int64_t ddlb_vmm_bytes(ddlb_vmm* s) { return s ? static_cast<int64_t>(s->bytes) : 0; }

// Self-test, phase 1: multimem.st of (rank+1) into slice `rank` of the region at `offset`.
extern "C" int ddlb_vmm_selftest_store(ddlb_vmm* s, int64_t offset, int n_per_rank, void* stream) {
  using namespace ddlb;
  if (s == nullptr || !s->mc_bound) return set_error("multicast range is not mapped");
  if (offset < 0 || offset % 16 != 0 ||
      static_cast<size_t>(offset) + static_cast<size_t>(n_per_rank) * s->world * 4 > s->bytes)
    return set_error("self-test region does not fit");
  multimem_store_kernel<<<(n_per_rank + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<float*>(s->mc_va + offset), s->rank, s->world, n_per_rank);
  DDLB_CUDA(cudaGetLastError());
  return 0;
}
// Self-test, phase 2 (after a barrier): out[i] = sum over ranks of region[i].
extern "C" int ddlb_vmm_selftest_reduce(ddlb_vmm* s, int64_t offset, float* out, int n, void* stream) {
  using namespace ddlb;
  if (s == nullptr || !s->mc_bound) return set_error("multicast range is not mapped");
  if (n % 4 != 0 || offset < 0 || offset % 16 != 0 ||
      static_cast<size_t>(offset) + static_cast<size_t>(n) * 4 > s->bytes)
    return set_error("bad self-test region");
  multimem_reduce_kernel<<<(n / 4 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float*>(s->mc_va + offset), out, n);
  DDLB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddlb_vmm_destroy(ddlb_vmm* s) {
  using namespace ddlb;
  if (s == nullptr) return 0;
  DriverApi& api = driver_api();
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  if (s->mc_bound) { api.MemUnmap(s->mc_va, s->bytes); api.MemAddressFree(s->mc_va, s->bytes); }
  if (s->mc != 0) api.MemRelease(s->mc);
  for (int r = 0; r < s->world; ++r) {
    if (s->mapped[r]) api.MemUnmap(s->va + static_cast<CUdeviceptr>(r) * s->bytes, s->bytes);
    if (s->handle[r] != 0) api.MemRelease(s->handle[r]);
  }
  if (s->va != 0) api.MemAddressFree(s->va, s->bytes * s->world);
  (void)cudaGetLastError();
  delete s;
  return 0;
}

extern "C" int ddlb_vmm_has_multicast(ddlb_vmm* s) { return (s != nullptr && s->has_mc) ? 1 : 0; }
