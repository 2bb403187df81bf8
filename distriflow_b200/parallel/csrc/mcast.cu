// Virtual-memory-management heap with an NVSwitch multicast (NVLS) mapping.
//
// The plain symmetric heap (comm.cu: cudaMalloc + cudaIpc handles) gives every rank a unicast mapping of every peer's
// buffer.  To let the switch reduce (`multimem.ld_reduce`) and broadcast (`multimem.st`) the same physical memory must
// also be bound to a multicast object, which only works for memory created through the driver's VMM calls.  This file is
// that allocator: cuMemCreate (POSIX-fd exportable) -> local unicast mapping, import of the peers' allocations -> peer
// unicast mappings, cuMulticastCreate / AddDevice / BindMem -> one multicast mapping per rank.  The file descriptors travel
// between the rank processes over an abstract unix socket (Python side, `parallel/symm.py`).
// The reference's transport (socket.io star, /root/reference/src/server/abstract_server.ts:81-89 broadcast of weights,
// /root/reference/src/server/federated_server.ts:98-106 mean of uploads) has no counterpart of in-network reduction.
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace dfb {

struct Drv {
  CUresult (*GetGran)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*Export)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*Import)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*Reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*Unmap)(CUdeviceptr, size_t);
  CUresult (*Release)(CUmemGenericAllocationHandle);
  CUresult (*AddrFree)(CUdeviceptr, size_t);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*McGetGran)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*McUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*DevAttr)(int*, CUdevice_attribute, CUdevice);
  CUresult (*ErrName)(CUresult, const char**);
  bool ok = false;
};
static Drv g_drv;
static std::once_flag g_drv_once;

template <typename F>
static bool load_one(const char* name, F& out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    cudaGetLastError();
    return false;
  }
  out = reinterpret_cast<F>(fn);
  return true;
}

static void load_drv() {
  Drv& d = g_drv;
  d.ok = load_one("cuMemGetAllocationGranularity", d.GetGran) && load_one("cuMemCreate", d.Create) &&
         load_one("cuMemExportToShareableHandle", d.Export) && load_one("cuMemImportFromShareableHandle", d.Import) &&
         load_one("cuMemAddressReserve", d.Reserve) && load_one("cuMemMap", d.Map) && load_one("cuMemSetAccess", d.SetAccess) &&
         load_one("cuMemUnmap", d.Unmap) && load_one("cuMemRelease", d.Release) && load_one("cuMemAddressFree", d.AddrFree) &&
         load_one("cuMulticastCreate", d.McCreate) && load_one("cuMulticastAddDevice", d.McAddDevice) &&
         load_one("cuMulticastBindMem", d.McBindMem) && load_one("cuMulticastGetGranularity", d.McGetGran) &&
         load_one("cuMulticastUnbind", d.McUnbind) && load_one("cuDeviceGetAttribute", d.DevAttr) &&
         load_one("cuGetErrorName", d.ErrName);
}

static const char* drv_err(CUresult r) {
  const char* s = nullptr;
  if (g_drv.ErrName && g_drv.ErrName(r, &s) == CUDA_SUCCESS && s) return s;
  return "?";
}

#define DFB_DRV_CHECK(expr)                                                                        \
  do {                                                                                             \
    CUresult _r = (expr);                                                                          \
    if (_r != CUDA_SUCCESS) {                                                                      \
      dfb::set_error("%s:%d: %s -> CUresult %d (%s)", __FILE__, __LINE__, #expr, (int)_r, dfb::drv_err(_r)); \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

static int ready() {
  std::call_once(g_drv_once, load_drv);
  if (!g_drv.ok) {
    set_error("driver VMM / multicast entry points are not available");
    return 1;
  }
  DFB_CUDA_CHECK(cudaFree(nullptr));          // make sure the calling thread has a current context
  return 0;
}

static CUmemAllocationProp phys_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

static int map_rw(CUmemGenericAllocationHandle h, size_t bytes, int dev, void** va_out) {
  CUdeviceptr va = 0;
  DFB_DRV_CHECK(g_drv.Reserve(&va, bytes, 0, 0, 0));
  DFB_DRV_CHECK(g_drv.Map(va, bytes, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DFB_DRV_CHECK(g_drv.SetAccess(va, bytes, &acc, 1));
  *va_out = reinterpret_cast<void*>(va);
  return 0;
}

}  // namespace dfb

using namespace dfb;

// 1 if the current device can take part in a multicast object and export POSIX-fd handles.
DFB_API int dfb_mc_supported(int* out) {
  *out = 0;
  if (ready()) return 1;
  int dev = 0, mc = 0, fd = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  DFB_DRV_CHECK(g_drv.DevAttr(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  DFB_DRV_CHECK(g_drv.DevAttr(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  *out = (mc && fd) ? 1 : 0;
  return 0;
}

// Allocation size that satisfies both the physical and the multicast granularity for `ndev` devices.
DFB_API int dfb_mc_round_size(size_t bytes, int ndev, size_t* rounded_out) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  CUmemAllocationProp p = phys_prop(dev);
  size_t g0 = 0, g1 = 0;
  DFB_DRV_CHECK(g_drv.GetGran(&g0, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)ndev;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DFB_DRV_CHECK(g_drv.McGetGran(&g1, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  size_t g = g0 > g1 ? g0 : g1;
  if (g == 0) g = 2u << 20;
  *rounded_out = (bytes + g - 1) / g * g;
  return 0;
}

// Physical allocation on the current device + local mapping; zero-filled.  `fd_out` is the exportable descriptor.
DFB_API int dfb_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out, void** va_out) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  CUmemAllocationProp p = phys_prop(dev);
  CUmemGenericAllocationHandle h = 0;
  DFB_DRV_CHECK(g_drv.Create(&h, bytes, &p, 0));
  int fd = -1;
  DFB_DRV_CHECK(g_drv.Export(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  if (map_rw(h, bytes, dev, va_out)) return 1;
  DFB_CUDA_CHECK(cudaMemset(*va_out, 0, bytes));
  DFB_CUDA_CHECK(cudaDeviceSynchronize());
  *handle_out = (unsigned long long)h;
  *fd_out = fd;
  return 0;
}

// Map a peer's allocation (descriptor received from its process) into this process, readable and writable from the
// current device.
DFB_API int dfb_vmm_import(int fd, size_t bytes, unsigned long long* handle_out, void** va_out) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  CUmemGenericAllocationHandle h = 0;
  DFB_DRV_CHECK(g_drv.Import(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  if (map_rw(h, bytes, dev, va_out)) return 1;
  *handle_out = (unsigned long long)h;
  return 0;
}

DFB_API int dfb_vmm_release(unsigned long long handle, void* va, size_t bytes) {
  if (ready()) return 1;
  if (va) {
    DFB_DRV_CHECK(g_drv.Unmap(reinterpret_cast<CUdeviceptr>(va), bytes));
    DFB_DRV_CHECK(g_drv.AddrFree(reinterpret_cast<CUdeviceptr>(va), bytes));
  }
  if (handle) DFB_DRV_CHECK(g_drv.Release((CUmemGenericAllocationHandle)handle));
  return 0;
}

// Rank 0: create the multicast object for `ndev` devices and export its descriptor.
DFB_API int dfb_mc_create(int ndev, size_t bytes, unsigned long long* mc_out, int* fd_out) {
  if (ready()) return 1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)ndev;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mc = 0;
  DFB_DRV_CHECK(g_drv.McCreate(&mc, &mp));
  int fd = -1;
  DFB_DRV_CHECK(g_drv.Export(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc_out = (unsigned long long)mc;
  *fd_out = fd;
  return 0;
}

DFB_API int dfb_mc_import(int fd, unsigned long long* mc_out) {
  if (ready()) return 1;
  CUmemGenericAllocationHandle mc = 0;
  DFB_DRV_CHECK(g_drv.Import(&mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_out = (unsigned long long)mc;
  return 0;
}

// Every rank, before anybody binds memory.
DFB_API int dfb_mc_add_device(unsigned long long mc) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  DFB_DRV_CHECK(g_drv.McAddDevice((CUmemGenericAllocationHandle)mc, dev));
  return 0;
}

// After ALL ranks have added their device: bind this rank's physical allocation at offset 0 and map the multicast
// object into this process.
DFB_API int dfb_mc_bind_map(unsigned long long mc, unsigned long long mem, size_t bytes, void** mc_va_out) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  DFB_DRV_CHECK(g_drv.McBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem, 0, bytes, 0));
  return map_rw((CUmemGenericAllocationHandle)mc, bytes, dev, mc_va_out);
}

DFB_API int dfb_mc_release(unsigned long long mc, void* mc_va, size_t bytes) {
  if (ready()) return 1;
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  if (mc_va) {
    DFB_DRV_CHECK(g_drv.Unmap(reinterpret_cast<CUdeviceptr>(mc_va), bytes));
    DFB_DRV_CHECK(g_drv.AddrFree(reinterpret_cast<CUdeviceptr>(mc_va), bytes));
  }
  if (mc) {
    g_drv.McUnbind((CUmemGenericAllocationHandle)mc, dev, 0, bytes);
    DFB_DRV_CHECK(g_drv.Release((CUmemGenericAllocationHandle)mc));
  }
  return 0;
}
