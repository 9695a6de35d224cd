// Symmetric memory on the CUDA virtual-memory API, with an NVSwitch multicast view (SURVEY 5.8 last row, K1/K8/K9).
//
// The reference has no counterpart: its gradient exchange is NCCL inside torch DDP / Horovod / DeepSpeed
// (harness/determined/pytorch/_pytorch_context.py:296-299,357-368).  NCCL's NVLS algorithm is what those paths get on an
// NVSwitch box; this file gives our own kernels the same hardware feature.
//
// * detb200_vmm_alloc: cuMemCreate (device-local, exportable as a POSIX file descriptor) + reserve + map + RW access.
//   The fd travels to the peer processes over a unix socket (SCM_RIGHTS, done in Python); a peer turns it into its own
//   mapping of OUR memory with detb200_vmm_import.  That replaces cudaIpc* for the flat gradient / parameter buffers.
// * detb200_mc_*: one multicast object per symmetric buffer.  Rank 0 creates it (cuMulticastCreate), everybody adds its
//   device, binds its local allocation at offset 0 and maps the object.  A `multimem.ld_reduce` on the mapped address
//   returns the sum over all ranks' buffers computed inside the switch; a `multimem.st` writes all of them.
//
// Only comm.cu's own kernels ever see a multicast address, and they touch it with multimem.* instructions only; no tensor
// map is ever built over it.
// Driver entry points come from cudaGetDriverEntryPoint, so the library does not link libcuda.
#include <cuda.h>
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <string.h>
#include <unistd.h>

#include <mutex>
#include <unordered_map>

#define DETB200_API extern "C" __attribute__((visibility("default")))
#define DETB200_ERR_BAD_ARG (-1)
#define DETB200_ERR_UNSUPPORTED (-3)
void detb200_set_error(const char* fmt, ...);

namespace {

struct Driver {
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  bool ok = false;
  bool mc_ok = false;
};

Driver g_drv;
std::once_flag g_drv_once;

template <typename F>
bool resolve(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}

void resolve_all() {
  Driver& d = g_drv;
  bool ok = true;
  ok &= resolve("cuGetErrorString", &d.GetErrorString);
  ok &= resolve("cuDeviceGet", &d.DeviceGet);
  ok &= resolve("cuDeviceGetAttribute", &d.DeviceGetAttribute);
  ok &= resolve("cuMemCreate", &d.MemCreate);
  ok &= resolve("cuMemRelease", &d.MemRelease);
  ok &= resolve("cuMemAddressReserve", &d.MemAddressReserve);
  ok &= resolve("cuMemAddressFree", &d.MemAddressFree);
  ok &= resolve("cuMemMap", &d.MemMap);
  ok &= resolve("cuMemUnmap", &d.MemUnmap);
  ok &= resolve("cuMemSetAccess", &d.MemSetAccess);
  ok &= resolve("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
  ok &= resolve("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
  ok &= resolve("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
  d.ok = ok;
  bool mc = true;
  mc &= resolve("cuMulticastCreate", &d.MulticastCreate);
  mc &= resolve("cuMulticastAddDevice", &d.MulticastAddDevice);
  mc &= resolve("cuMulticastBindMem", &d.MulticastBindMem);
  mc &= resolve("cuMulticastUnbind", &d.MulticastUnbind);
  mc &= resolve("cuMulticastGetGranularity", &d.MulticastGetGranularity);
  d.mc_ok = ok && mc;
}

const Driver* drv() {
  cudaFree(nullptr);  // make sure the primary context exists before asking for driver symbols
  std::call_once(g_drv_once, resolve_all);
  return &g_drv;
}

int fail(const char* what, CUresult r) {
  const char* s = nullptr;
  if (g_drv.GetErrorString) g_drv.GetErrorString(r, &s);
  detb200_set_error("%s failed: %s (CUresult %d)", what, s ? s : "?", (int)r);
  return 10000 + (int)r;
}

#define DRV_CHECK(call, what)                 \
  do {                                        \
    CUresult _r = (call);                     \
    if (_r != CUDA_SUCCESS) return fail(what, _r); \
  } while (0)

struct Mapping {
  CUmemGenericAllocationHandle handle;
  size_t bytes;
  bool owns_handle;
};
std::mutex g_mu;
std::unordered_map<uintptr_t, Mapping> g_maps;       // mapped VA -> what to undo
std::unordered_map<int64_t, CUmemGenericAllocationHandle> g_mc;  // multicast id -> handle
int64_t g_next_mc = 1;

int current_device(CUdevice* dev, int* ordinal) {
  int o = 0;
  cudaError_t e = cudaGetDevice(&o);
  if (e != cudaSuccess) { detb200_set_error("cudaGetDevice failed: %s", cudaGetErrorString(e)); return (int)e; }
  DRV_CHECK(drv()->DeviceGet(dev, o), "cuDeviceGet");
  if (ordinal) *ordinal = o;
  return 0;
}

CUmemAllocationProp local_prop(int ordinal) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = ordinal;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_handle(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int ordinal, void** out) {
  const Driver* d = drv();
  CUdeviceptr va = 0;
  DRV_CHECK(d->MemAddressReserve(&va, bytes, align, 0, 0), "cuMemAddressReserve");
  CUresult r = d->MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) { d->MemAddressFree(va, bytes); return fail("cuMemMap", r); }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d->MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) { d->MemUnmap(va, bytes); d->MemAddressFree(va, bytes); return fail("cuMemSetAccess", r); }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace

// caps[0] = VMM allocations exportable as POSIX fds, caps[1] = multicast objects (NVLS) on the current device
DETB200_API int detb200_vmm_query(int* caps) {
  if (!caps) return DETB200_ERR_BAD_ARG;
  caps[0] = caps[1] = 0;
  const Driver* d = drv();
  if (!d->ok) return 0;
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc) return rc;
  int v = 0;
  if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS) caps[0] = v;
  v = 0;
  if (d->mc_ok && d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS) caps[1] = v;
  return 0;
}

// Size granularity that satisfies both cuMemCreate and a `world`-device multicast object.
DETB200_API int detb200_vmm_granularity(int world, int64_t* out) {
  const Driver* d = drv();
  if (!d->ok || !out) { detb200_set_error("VMM driver entry points unavailable"); return DETB200_ERR_UNSUPPORTED; }
  CUdevice dev;
  int ord = 0;
  int rc = current_device(&dev, &ord);
  if (rc) return rc;
  CUmemAllocationProp p = local_prop(ord);
  size_t g = 0;
  DRV_CHECK(d->MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
  if (d->mc_ok && world > 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  *out = (int64_t)g;
  return 0;
}

// `bytes` must be a multiple of detb200_vmm_granularity().  *fd_out is a fresh descriptor the caller sends to its peers
// and then closes.  The memory is zero-filled (synchronises).
DETB200_API int detb200_vmm_alloc(int64_t bytes, void** ptr, int* fd_out) {
  const Driver* d = drv();
  if (!d->ok) { detb200_set_error("VMM driver entry points unavailable"); return DETB200_ERR_UNSUPPORTED; }
  if (!ptr || !fd_out || bytes <= 0) { detb200_set_error("vmm_alloc: bad argument"); return DETB200_ERR_BAD_ARG; }
  CUdevice dev;
  int ord = 0;
  int rc = current_device(&dev, &ord);
  if (rc) return rc;
  CUmemAllocationProp p = local_prop(ord);
  CUmemGenericAllocationHandle h;
  DRV_CHECK(d->MemCreate(&h, (size_t)bytes, &p, 0), "cuMemCreate");
  void* va = nullptr;
  rc = map_handle(h, (size_t)bytes, 0, ord, &va);
  if (rc) { d->MemRelease(h); return rc; }
  int fd = -1;
  CUresult r = d->MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d->MemUnmap((CUdeviceptr)va, (size_t)bytes); d->MemAddressFree((CUdeviceptr)va, (size_t)bytes); d->MemRelease(h);
    return fail("cuMemExportToShareableHandle", r);
  }
  cudaError_t e = cudaMemset(va, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { detb200_set_error("zero-fill of VMM allocation failed: %s", cudaGetErrorString(e)); return (int)e; }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_maps[(uintptr_t)va] = Mapping{h, (size_t)bytes, true};
  }
  *ptr = va;
  *fd_out = fd;
  return 0;
}

// Map a peer's allocation (fd received over the unix socket) into this process, readable and writable by the current
// device.  The fd stays the caller's to close.
DETB200_API int detb200_vmm_import(int fd, int64_t bytes, void** ptr_out) {
  const Driver* d = drv();
  if (!d->ok) { detb200_set_error("VMM driver entry points unavailable"); return DETB200_ERR_UNSUPPORTED; }
  CUdevice dev;
  int ord = 0;
  int rc = current_device(&dev, &ord);
  if (rc) return rc;
  CUmemGenericAllocationHandle h;
  DRV_CHECK(d->MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
  void* va = nullptr;
  rc = map_handle(h, (size_t)bytes, 0, ord, &va);
  if (rc) { d->MemRelease(h); return rc; }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_maps[(uintptr_t)va] = Mapping{h, (size_t)bytes, true};
  }
  *ptr_out = va;
  return 0;
}

DETB200_API int detb200_vmm_free(void* ptr) {
  Mapping m;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_maps.find((uintptr_t)ptr);
    if (it == g_maps.end()) return 0;
    m = it->second;
    g_maps.erase(it);
  }
  const Driver* d = drv();
  d->MemUnmap((CUdeviceptr)ptr, m.bytes);
  d->MemAddressFree((CUdeviceptr)ptr, m.bytes);
  if (m.owns_handle) d->MemRelease(m.handle);
  return 0;
}

// ---- multicast object -------------------------------------------------------------------------------------------------
DETB200_API int detb200_mc_create(int world, int64_t bytes, int64_t* mc_out, int* fd_out) {
  const Driver* d = drv();
  if (!d->mc_ok) { detb200_set_error("multicast driver entry points unavailable"); return DETB200_ERR_UNSUPPORTED; }
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)world;
  mp.size = (size_t)bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DRV_CHECK(d->MulticastCreate(&h, &mp), "cuMulticastCreate");
  int fd = -1;
  CUresult r = d->MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { d->MemRelease(h); return fail("cuMemExportToShareableHandle(multicast)", r); }
  std::lock_guard<std::mutex> lk(g_mu);
  *mc_out = g_next_mc++;
  g_mc[*mc_out] = h;
  *fd_out = fd;
  return 0;
}

DETB200_API int detb200_mc_import(int fd, int64_t* mc_out) {
  const Driver* d = drv();
  if (!d->mc_ok) { detb200_set_error("multicast driver entry points unavailable"); return DETB200_ERR_UNSUPPORTED; }
  CUmemGenericAllocationHandle h;
  DRV_CHECK(d->MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle(multicast)");
  std::lock_guard<std::mutex> lk(g_mu);
  *mc_out = g_next_mc++;
  g_mc[*mc_out] = h;
  return 0;
}

static int mc_lookup(int64_t mc, CUmemGenericAllocationHandle* h) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_mc.find(mc);
  if (it == g_mc.end()) { detb200_set_error("unknown multicast object %lld", (long long)mc); return DETB200_ERR_BAD_ARG; }
  *h = it->second;
  return 0;
}

// Every rank, before any rank binds.
DETB200_API int detb200_mc_add_device(int64_t mc) {
  CUmemGenericAllocationHandle h;
  int rc = mc_lookup(mc, &h);
  if (rc) return rc;
  CUdevice dev;
  rc = current_device(&dev, nullptr);
  if (rc) return rc;
  DRV_CHECK(drv()->MulticastAddDevice(h, dev), "cuMulticastAddDevice");
  return 0;
}

// Bind this rank's allocation (a pointer returned by detb200_vmm_alloc) at offset 0 of the object.
DETB200_API int detb200_mc_bind(int64_t mc, void* local_ptr, int64_t bytes) {
  CUmemGenericAllocationHandle h;
  int rc = mc_lookup(mc, &h);
  if (rc) return rc;
  Mapping m;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_maps.find((uintptr_t)local_ptr);
    if (it == g_maps.end()) { detb200_set_error("mc_bind: pointer was not allocated with detb200_vmm_alloc"); return DETB200_ERR_BAD_ARG; }
    m = it->second;
  }
  if ((size_t)bytes > m.bytes) { detb200_set_error("mc_bind: size exceeds the allocation"); return DETB200_ERR_BAD_ARG; }
  DRV_CHECK(drv()->MulticastBindMem(h, 0, m.handle, 0, (size_t)bytes, 0), "cuMulticastBindMem");
  return 0;
}

// Map the object; the returned address is for multimem.* instructions only.
DETB200_API int detb200_mc_map(int64_t mc, int64_t bytes, void** mc_ptr_out) {
  CUmemGenericAllocationHandle h;
  int rc = mc_lookup(mc, &h);
  if (rc) return rc;
  CUdevice dev;
  int ord = 0;
  rc = current_device(&dev, &ord);
  if (rc) return rc;
  void* va = nullptr;
  rc = map_handle(h, (size_t)bytes, 0, ord, &va);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  g_maps[(uintptr_t)va] = Mapping{h, (size_t)bytes, false};
  *mc_ptr_out = va;
  return 0;
}

DETB200_API int detb200_mc_free(int64_t mc, void* mc_ptr, int64_t bytes) {
  if (mc_ptr) detb200_vmm_free(mc_ptr);
  CUmemGenericAllocationHandle h;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_mc.find(mc);
    if (it == g_mc.end()) return 0;
    h = it->second;
    g_mc.erase(it);
  }
  CUdevice dev;
  if (current_device(&dev, nullptr) == 0 && bytes > 0) drv()->MulticastUnbind(h, dev, 0, (size_t)bytes);
  drv()->MemRelease(h);
  return 0;
}

DETB200_API int detb200_close_fd(int fd) { return fd >= 0 ? close(fd) : 0; }
