// NVLS (NVSwitch multicast / in-switch reduction) plumbing through the CUDA driver API.
//
// SURVEY 5.9 names multimem.st / multimem.ld_reduce as the product path for broadcast / allreduce.  torch's
// symmetric memory reports multicast_ptr = 0 on the boxes this was developed on, so this file asks the driver
// directly: is CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED set, what granularity does a multicast object want, and does
// cuMulticastCreate succeed for `world` devices.  hb_nvls_probe reports every CUresult; the collectives fall back to
// the P2P ld/st kernels of coll.cu when the answer is no (the exact codes are recorded in DESIGN.md).
#include "common.cuh"
#include "../include/harp_b200.h"

#include <stdio.h>
#include <string.h>

namespace {

template <typename Fn> Fn driver_fn(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<Fn>(p);
}

typedef CUresult (*DevGetAttrFn)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*McCreateFn)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*McGranFn)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
typedef CUresult (*MemReleaseFn)(CUmemGenericAllocationHandle);
typedef CUresult (*McAddDevFn)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*CtxGetDevFn)(CUdevice*);

}  // namespace

extern "C" {

// Writes a one-line report into `msg` and returns 1 when a multicast object for `world` devices of `bytes` bytes can
// be created on the current device (the object is released again), 0 when not, negative on a plumbing error.
int hb_nvls_probe(int world, size_t bytes, char* msg, int msg_len) {
  auto get_attr = driver_fn<DevGetAttrFn>("cuDeviceGetAttribute");
  auto mc_create = driver_fn<McCreateFn>("cuMulticastCreate");
  auto mc_gran = driver_fn<McGranFn>("cuMulticastGetGranularity");
  auto mem_release = driver_fn<MemReleaseFn>("cuMemRelease");
  auto mc_add = driver_fn<McAddDevFn>("cuMulticastAddDevice");
  auto ctx_dev = driver_fn<CtxGetDevFn>("cuCtxGetDevice");
  if (!get_attr || !mc_create || !mc_gran || !mem_release || !mc_add || !ctx_dev) {
    snprintf(msg, msg_len, "driver entry points for cuMulticast* not found");
    return -1;
  }
  cudaFree(0);      // make sure the primary context exists
  CUdevice dev;
  CUresult r = ctx_dev(&dev);
  if (r != CUDA_SUCCESS) { snprintf(msg, msg_len, "cuCtxGetDevice -> CUresult %d", (int)r); return -1; }
  int mc = -1, fabric = -1, posix = -1;
  const CUresult ra = get_attr(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  get_attr(&fabric, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev);
  get_attr(&posix, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  const CUresult rg = mc_gran(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (rg == CUDA_SUCCESS && gran > 0) prop.size = (bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h = 0;
  const CUresult rc = mc_create(&h, &prop);
  CUresult radd = (CUresult)-1;
  if (rc == CUDA_SUCCESS) {
    radd = mc_add(h, dev);
    mem_release(h);
  }
  snprintf(msg, msg_len,
           "MULTICAST_SUPPORTED=%d (CUresult %d) FABRIC_HANDLES=%d POSIX_FD_HANDLES=%d | cuMulticastGetGranularity -> %d "
           "(granularity %zu) | cuMulticastCreate(numDevices=%d, size=%zu, POSIX_FD) -> CUresult %d | cuMulticastAddDevice -> %d",
           mc, (int)ra, fabric, posix, (int)rg, gran, world, (size_t)prop.size, (int)rc, (int)radd);
  return (rc == CUDA_SUCCESS && radd == CUDA_SUCCESS) ? 1 : 0;
}

}  // extern "C"

// ----------------------------------------------------------------------------
// VMM slab + multicast binding: the symmetric workspace's own allocator (SURVEY 2.1.3 "cuMem VMM + posix handles
// exchanged once, mapped by all peers, plus an NVLS multicast mapping").  Handles travel between the ranks as
// POSIX file descriptors (the Python side passes them over a unix socket with SCM_RIGHTS).
// ----------------------------------------------------------------------------
namespace {
typedef CUresult (*MemCreateFn)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*MemExportFn)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*MemImportFn)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*MemGranFn)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*MemReserveFn)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*MemMapFn)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*MemSetAccessFn)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*MemUnmapFn)(CUdeviceptr, size_t);
typedef CUresult (*MemFreeVaFn)(CUdeviceptr, size_t);
typedef CUresult (*McBindFn)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);

#define HB_CU_OK(expr)                                                                   \
  do {                                                                                   \
    CUresult _r = (expr);                                                                \
    if (_r != CUDA_SUCCESS) {                                                            \
      hb::set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, (int)_r);     \
      return -1;                                                                         \
    }                                                                                    \
  } while (0)

int current_device(CUdevice* dev) {
  cudaFree(0);
  auto ctx_dev = driver_fn<CtxGetDevFn>("cuCtxGetDevice");
  HB_REQUIRE(ctx_dev != nullptr, "cuCtxGetDevice not found");
  HB_CU_OK(ctx_dev(dev));
  return 0;
}

void fill_prop(CUmemAllocationProp* prop, CUdevice dev) {
  memset(prop, 0, sizeof(*prop));
  prop->type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop->location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop->location.id = (int)dev;
  prop->requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
}
}  // namespace

extern "C" {

// Smallest size step for a slab that is also bound to a multicast object of `world` devices.
int hb_vmm_granularity(int world, size_t* out) {
  CUdevice dev;
  if (current_device(&dev) != 0) return -1;
  auto mem_gran = driver_fn<MemGranFn>("cuMemGetAllocationGranularity");
  auto mc_gran = driver_fn<McGranFn>("cuMulticastGetGranularity");
  HB_REQUIRE(mem_gran && mc_gran, "cuMem* / cuMulticast* entry points not found");
  CUmemAllocationProp prop;
  fill_prop(&prop, dev);
  size_t g = 0, gm = 0;
  HB_CU_OK(mem_gran(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (world > 1) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world; mp.size = g; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (mc_gran(&gm, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS) gm = 0;
  }
  *out = g > gm ? g : gm;
  return 0;
}

// Physical memory on the current device; *fd_out is a POSIX fd other processes can import.
int hb_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out) {
  CUdevice dev;
  if (current_device(&dev) != 0) return -1;
  auto mem_create = driver_fn<MemCreateFn>("cuMemCreate");
  auto mem_export = driver_fn<MemExportFn>("cuMemExportToShareableHandle");
  HB_REQUIRE(mem_create && mem_export, "cuMemCreate / cuMemExportToShareableHandle not found");
  CUmemAllocationProp prop;
  fill_prop(&prop, dev);
  CUmemGenericAllocationHandle h;
  HB_CU_OK(mem_create(&h, bytes, &prop, 0));
  int fd = -1;
  HB_CU_OK(mem_export(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle_out = (unsigned long long)h;
  *fd_out = fd;
  return 0;
}

int hb_vmm_import(int fd, unsigned long long* handle_out) {
  CUdevice dev;
  if (current_device(&dev) != 0) return -1;
  auto mem_import = driver_fn<MemImportFn>("cuMemImportFromShareableHandle");
  HB_REQUIRE(mem_import != nullptr, "cuMemImportFromShareableHandle not found");
  CUmemGenericAllocationHandle h;
  HB_CU_OK(mem_import(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle_out = (unsigned long long)h;
  return 0;
}

// Maps a physical or multicast handle read-write into the current device's address space.
int hb_vmm_map(unsigned long long handle, size_t bytes, size_t align, void** ptr_out) {
  CUdevice dev;
  if (current_device(&dev) != 0) return -1;
  auto reserve = driver_fn<MemReserveFn>("cuMemAddressReserve");
  auto map = driver_fn<MemMapFn>("cuMemMap");
  auto set_access = driver_fn<MemSetAccessFn>("cuMemSetAccess");
  HB_REQUIRE(reserve && map && set_access, "cuMemAddressReserve / cuMemMap / cuMemSetAccess not found");
  CUdeviceptr va = 0;
  HB_CU_OK(reserve(&va, bytes, align, 0, 0));
  HB_CU_OK(map(va, bytes, 0, (CUmemGenericAllocationHandle)handle, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = (int)dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  HB_CU_OK(set_access(va, bytes, &acc, 1));
  *ptr_out = (void*)va;
  return 0;
}

int hb_vmm_unmap(void* ptr, size_t bytes) {
  auto unmap = driver_fn<MemUnmapFn>("cuMemUnmap");
  auto free_va = driver_fn<MemFreeVaFn>("cuMemAddressFree");
  HB_REQUIRE(unmap && free_va, "cuMemUnmap / cuMemAddressFree not found");
  HB_CU_OK(unmap((CUdeviceptr)ptr, bytes));
  HB_CU_OK(free_va((CUdeviceptr)ptr, bytes));
  return 0;
}

int hb_vmm_release(unsigned long long handle) {
  auto mem_release = driver_fn<MemReleaseFn>("cuMemRelease");
  HB_REQUIRE(mem_release != nullptr, "cuMemRelease not found");
  HB_CU_OK(mem_release((CUmemGenericAllocationHandle)handle));
  return 0;
}

// Multicast object for `world` devices (created by ONE rank, imported by the others through its fd).
int hb_mc_create(int world, size_t bytes, unsigned long long* handle_out, int* fd_out) {
  auto mc_create = driver_fn<McCreateFn>("cuMulticastCreate");
  auto mem_export = driver_fn<MemExportFn>("cuMemExportToShareableHandle");
  HB_REQUIRE(mc_create && mem_export, "cuMulticastCreate not found");
  cudaFree(0);
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  HB_CU_OK(mc_create(&h, &prop));
  int fd = -1;
  HB_CU_OK(mem_export(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle_out = (unsigned long long)h;
  *fd_out = fd;
  return 0;
}

int hb_mc_add_device(unsigned long long mc_handle) {
  CUdevice dev;
  if (current_device(&dev) != 0) return -1;
  auto mc_add = driver_fn<McAddDevFn>("cuMulticastAddDevice");
  HB_REQUIRE(mc_add != nullptr, "cuMulticastAddDevice not found");
  HB_CU_OK(mc_add((CUmemGenericAllocationHandle)mc_handle, dev));
  return 0;
}

// Every device must have been added (by its own process) before the first bind.
int hb_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes) {
  auto mc_bind = driver_fn<McBindFn>("cuMulticastBindMem");
  HB_REQUIRE(mc_bind != nullptr, "cuMulticastBindMem not found");
  HB_CU_OK(mc_bind((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle, 0, bytes, 0));
  return 0;
}

}  // extern "C"
