// S2: symmetric memory and NVLS multicast objects, owned by this library.
//
// The reference has no counterpart (its collectives are NCCL's, behind DDP).  The peer-memory kernels here
// need every rank's buffer mapped in every rank's address space, and the NVLS kernels additionally a
// multicast mapping of all of them.  Round 1 got the former from CUDA IPC and the latter from a private
// torch API; this file provides both from the driver's virtual-memory API:
//
//   rdp_symm_alloc        cuMemCreate (shareable as a POSIX file descriptor) + cuMemAddressReserve + cuMemMap
//   rdp_symm_export_fd    cuMemExportToShareableHandle -> an fd the Python side passes over a unix socket
//   rdp_symm_import_fd    cuMemImportFromShareableHandle + map with read/write access for the local device
//   rdp_mc_create / rdp_mc_export_fd / rdp_mc_import_fd / rdp_mc_add_device / rdp_mc_bind / rdp_mc_map
//                         cuMulticastCreate, handle exchange the same way, cuMulticastAddDevice by every rank,
//                         cuMulticastBindMem of every rank's physical allocation, and a VA mapping of the
//                         multicast object: the address multimem.ld_reduce / multimem.st operate on
//   rdp_symm_free / rdp_mc_free   unmap + release; need no cooperation of peers (a dead peer cannot block them)
//
// The driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the library keeps no
// link-time dependency on libcuda.
#include "common.cuh"
#include "raydp_b200.h"

#include <mutex>
#include <string.h>
#include <unistd.h>

namespace rdp {

struct DriverApi {
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*memRelease)(CUmemGenericAllocationHandle);
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*memAddressFree)(CUdeviceptr, size_t);
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*memUnmap)(CUdeviceptr, size_t);
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*memExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*memImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*memGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*mcCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*mcAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*mcBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*mcGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*mcUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*deviceGet)(CUdevice*, int);
  CUresult (*deviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  bool ok = false;
};

static DriverApi g_drv;
static std::once_flag g_drv_once;

template <typename F>
static bool resolve(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      fn == nullptr)
    return false;
  *out = reinterpret_cast<F>(fn);
  return true;
}

static void load_driver() {
  bool ok = true;
  ok &= resolve("cuMemCreate", &g_drv.memCreate);
  ok &= resolve("cuMemRelease", &g_drv.memRelease);
  ok &= resolve("cuMemAddressReserve", &g_drv.memAddressReserve);
  ok &= resolve("cuMemAddressFree", &g_drv.memAddressFree);
  ok &= resolve("cuMemMap", &g_drv.memMap);
  ok &= resolve("cuMemUnmap", &g_drv.memUnmap);
  ok &= resolve("cuMemSetAccess", &g_drv.memSetAccess);
  ok &= resolve("cuMemExportToShareableHandle", &g_drv.memExport);
  ok &= resolve("cuMemImportFromShareableHandle", &g_drv.memImport);
  ok &= resolve("cuMemGetAllocationGranularity", &g_drv.memGranularity);
  ok &= resolve("cuDeviceGet", &g_drv.deviceGet);
  ok &= resolve("cuDeviceGetAttribute", &g_drv.deviceGetAttribute);
  // multicast entry points are optional (older drivers): checked where used
  resolve("cuMulticastCreate", &g_drv.mcCreate);
  resolve("cuMulticastAddDevice", &g_drv.mcAddDevice);
  resolve("cuMulticastBindMem", &g_drv.mcBindMem);
  resolve("cuMulticastGetGranularity", &g_drv.mcGranularity);
  resolve("cuMulticastUnbind", &g_drv.mcUnbind);
  g_drv.ok = ok;
}

#define RDP_CHECK_CU(expr)                                                                        \
  do {                                                                                            \
    CUresult _r = (expr);                                                                         \
    if (_r != CUDA_SUCCESS) {                                                                     \
      rdp::set_error("%s failed at %s:%d with CUresult %d", #expr, __FILE__, __LINE__, int(_r));  \
      return -2000 - int(_r);                                                                     \
    }                                                                                             \
  } while (0)

struct SymmBlock {            // one physical allocation (own or imported) mapped in this process
  CUmemGenericAllocationHandle handle = 0;
  CUdeviceptr va = 0;
  size_t size = 0;            // rounded to the granularity
  int device = 0;
  bool multicast = false;
  bool mapped = false;
};

static CUmemAllocationProp local_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

static int map_block(SymmBlock* b, int device) {
  RDP_CHECK_CU(g_drv.memAddressReserve(&b->va, b->size, 0, 0, 0));
  CUresult r = g_drv.memMap(b->va, b->size, 0, b->handle, 0);
  if (r != CUDA_SUCCESS) {
    g_drv.memAddressFree(b->va, b->size);
    b->va = 0;
    rdp::set_error("cuMemMap failed with CUresult %d", int(r));
    return -2000 - int(r);
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_drv.memSetAccess(b->va, b->size, &acc, 1);
  if (r != CUDA_SUCCESS) {
    g_drv.memUnmap(b->va, b->size);
    g_drv.memAddressFree(b->va, b->size);
    b->va = 0;
    rdp::set_error("cuMemSetAccess failed with CUresult %d", int(r));
    return -2000 - int(r);
  }
  b->mapped = true;
  b->device = device;
  return 0;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_symm_supported(int device) {
  std::call_once(g_drv_once, load_driver);
  if (!g_drv.ok) return 0;
  CUdevice dev;
  if (g_drv.deviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  int vmm = 0, fd = 0, mc = 0;
  g_drv.deviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  g_drv.deviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  if (g_drv.mcCreate && g_drv.mcAddDevice && g_drv.mcBindMem && g_drv.mcGranularity)
    g_drv.deviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  return (vmm && fd) ? (mc ? 2 : 1) : 0;        // 0: unsupported, 1: symmetric memory, 2: + multicast
}

int rdp_symm_granularity(int device, int multicast, int n_devices, uint64_t* out) {
  std::call_once(g_drv_once, load_driver);
  RDP_REQUIRE(g_drv.ok, "rdp_symm: the CUDA driver's virtual memory API is not available");
  size_t g = 0;
  CUmemAllocationProp prop = local_prop(device);
  RDP_CHECK_CU(g_drv.memGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (multicast) {
    RDP_REQUIRE(g_drv.mcGranularity != nullptr, "rdp_symm: multicast objects are not supported by this driver");
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = unsigned(n_devices);
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    RDP_CHECK_CU(g_drv.mcGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > g) g = mg;
  }
  *out = g;
  return 0;
}

// Allocate `bytes` (a multiple of rdp_symm_granularity) of device memory on `device`, shareable as an fd.
int rdp_symm_alloc(uint64_t bytes, int device, void** handle_out, void** ptr_out) {
  std::call_once(g_drv_once, load_driver);
  RDP_REQUIRE(g_drv.ok, "rdp_symm_alloc: the CUDA driver's virtual memory API is not available");
  SymmBlock* b = new SymmBlock();
  b->size = size_t(bytes);
  CUmemAllocationProp prop = local_prop(device);
  CUresult r = g_drv.memCreate(&b->handle, b->size, &prop, 0);
  if (r != CUDA_SUCCESS) {
    delete b;
    rdp::set_error("cuMemCreate(%llu bytes) failed with CUresult %d", (unsigned long long)bytes, int(r));
    return -2000 - int(r);
  }
  int rc = map_block(b, device);
  if (rc) {
    g_drv.memRelease(b->handle);
    delete b;
    return rc;
  }
  RDP_CHECK_CUDA(cudaMemset(reinterpret_cast<void*>(b->va), 0, b->size));
  RDP_CHECK_CUDA(cudaDeviceSynchronize());
  *handle_out = b;
  *ptr_out = reinterpret_cast<void*>(b->va);
  return 0;
}

int rdp_symm_export_fd(void* handle, int* fd_out) {
  SymmBlock* b = static_cast<SymmBlock*>(handle);
  RDP_REQUIRE(b != nullptr, "rdp_symm_export_fd: null handle");
  int fd = -1;
  RDP_CHECK_CU(g_drv.memExport(&fd, b->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}

// Map a peer's allocation (its fd received over a unix socket) with read/write access for `device`.
int rdp_symm_import_fd(int fd, uint64_t bytes, int device, void** handle_out, void** ptr_out) {
  std::call_once(g_drv_once, load_driver);
  RDP_REQUIRE(g_drv.ok, "rdp_symm_import_fd: the CUDA driver's virtual memory API is not available");
  SymmBlock* b = new SymmBlock();
  b->size = size_t(bytes);
  CUresult r = g_drv.memImport(&b->handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete b;
    rdp::set_error("cuMemImportFromShareableHandle failed with CUresult %d", int(r));
    return -2000 - int(r);
  }
  int rc = map_block(b, device);
  if (rc) {
    g_drv.memRelease(b->handle);
    delete b;
    return rc;
  }
  *handle_out = b;
  *ptr_out = reinterpret_cast<void*>(b->va);
  return 0;
}

int rdp_symm_free(void* handle) {
  SymmBlock* b = static_cast<SymmBlock*>(handle);
  if (b == nullptr) return 0;
  if (b->mapped) {
    g_drv.memUnmap(b->va, b->size);
    g_drv.memAddressFree(b->va, b->size);
  }
  g_drv.memRelease(b->handle);
  delete b;
  return 0;
}

// ---- multicast (NVLS) -------------------------------------------------------------------------------
int rdp_mc_create(uint64_t bytes, int n_devices, void** handle_out) {
  std::call_once(g_drv_once, load_driver);
  RDP_REQUIRE(g_drv.ok && g_drv.mcCreate != nullptr, "rdp_mc_create: multicast objects are not supported by this driver");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = unsigned(n_devices);
  mp.size = size_t(bytes);
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  SymmBlock* b = new SymmBlock();
  b->size = size_t(bytes);
  b->multicast = true;
  CUresult r = g_drv.mcCreate(&b->handle, &mp);
  if (r != CUDA_SUCCESS) {
    delete b;
    rdp::set_error("cuMulticastCreate(%llu bytes, %d devices) failed with CUresult %d", (unsigned long long)bytes,
                   n_devices, int(r));
    return -2000 - int(r);
  }
  *handle_out = b;
  return 0;
}

int rdp_mc_import_fd(int fd, uint64_t bytes, void** handle_out) {
  std::call_once(g_drv_once, load_driver);
  RDP_REQUIRE(g_drv.ok && g_drv.mcAddDevice != nullptr, "rdp_mc_import_fd: multicast objects are not supported");
  SymmBlock* b = new SymmBlock();
  b->size = size_t(bytes);
  b->multicast = true;
  CUresult r = g_drv.memImport(&b->handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete b;
    rdp::set_error("cuMemImportFromShareableHandle(multicast) failed with CUresult %d", int(r));
    return -2000 - int(r);
  }
  *handle_out = b;
  return 0;
}

int rdp_mc_add_device(void* mc, int device) {
  SymmBlock* b = static_cast<SymmBlock*>(mc);
  RDP_REQUIRE(b != nullptr && b->multicast, "rdp_mc_add_device: not a multicast handle");
  CUdevice dev;
  RDP_CHECK_CU(g_drv.deviceGet(&dev, device));
  RDP_CHECK_CU(g_drv.mcAddDevice(b->handle, dev));
  b->device = device;
  return 0;
}

// Bind this rank's physical allocation (all of it) at offset 0 of the multicast object.  Every device of
// the object must have been added (rdp_mc_add_device on every rank) before the first bind.
int rdp_mc_bind(void* mc, void* symm) {
  SymmBlock* m = static_cast<SymmBlock*>(mc);
  SymmBlock* s = static_cast<SymmBlock*>(symm);
  RDP_REQUIRE(m && s && m->multicast && !s->multicast, "rdp_mc_bind: bad handles");
  RDP_REQUIRE(s->size <= m->size, "rdp_mc_bind: allocation larger than the multicast object");
  RDP_CHECK_CU(g_drv.mcBindMem(m->handle, 0, s->handle, 0, s->size, 0));
  return 0;
}

int rdp_mc_map(void* mc, int device, void** ptr_out) {
  SymmBlock* m = static_cast<SymmBlock*>(mc);
  RDP_REQUIRE(m && m->multicast, "rdp_mc_map: not a multicast handle");
  int rc = map_block(m, device);
  if (rc) return rc;
  *ptr_out = reinterpret_cast<void*>(m->va);
  return 0;
}

int rdp_close_fd(int fd) { return close(fd) == 0 ? 0 : -1; }

}  // extern "C"
