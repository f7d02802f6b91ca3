// Host-side set-up for the NVLS (NVLink SHARP / multicast) allreduce: symmetric regions allocated with the
// CUDA virtual-memory-management API so that (a) every peer can map them unicast through a POSIX file
// descriptor (replacing CUDA IPC for this path) and (b) the same physical memory can be bound to an NVSwitch
// multicast object, giving a second "multicast" address on which multimem.ld_reduce / multimem.st operate
// in the switch.  No kernels here; the kernels are in allreduce.cu.
//
// Driver functions are resolved with cudaGetDriverEntryPoint, so libb200ops.so keeps loading on a machine
// without libcuda (the CPU-only build / test container).
#include "common.cuh"

#include "../include/b200_ops.h"

namespace b200 {
namespace {

template <typename Fn>
Fn driver_fn(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<Fn>(p);
}

using FnMemCreate = CUresult (*)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
using FnMemRelease = CUresult (*)(CUmemGenericAllocationHandle);
using FnMemGran = CUresult (*)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
using FnAddrReserve = CUresult (*)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
using FnAddrFree = CUresult (*)(CUdeviceptr, size_t);
using FnMemMap = CUresult (*)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
using FnMemUnmap = CUresult (*)(CUdeviceptr, size_t);
using FnSetAccess = CUresult (*)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
using FnExport = CUresult (*)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
using FnImport = CUresult (*)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
using FnMcCreate = CUresult (*)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
using FnMcAddDevice = CUresult (*)(CUmemGenericAllocationHandle, CUdevice);
using FnMcBindMem = CUresult (*)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
using FnMcGran = CUresult (*)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
using FnDevGetAttr = CUresult (*)(int*, CUdevice_attribute, CUdevice);

struct Driver {
  FnMemCreate mem_create;
  FnMemRelease mem_release;
  FnMemGran mem_gran;
  FnAddrReserve addr_reserve;
  FnAddrFree addr_free;
  FnMemMap mem_map;
  FnMemUnmap mem_unmap;
  FnSetAccess set_access;
  FnExport export_handle;
  FnImport import_handle;
  FnMcCreate mc_create;
  FnMcAddDevice mc_add_device;
  FnMcBindMem mc_bind_mem;
  FnMcGran mc_gran;
  FnDevGetAttr dev_attr;
  bool ok;
};

const Driver& driver() {
  static Driver d = [] {
    Driver x{};
    x.mem_create = driver_fn<FnMemCreate>("cuMemCreate");
    x.mem_release = driver_fn<FnMemRelease>("cuMemRelease");
    x.mem_gran = driver_fn<FnMemGran>("cuMemGetAllocationGranularity");
    x.addr_reserve = driver_fn<FnAddrReserve>("cuMemAddressReserve");
    x.addr_free = driver_fn<FnAddrFree>("cuMemAddressFree");
    x.mem_map = driver_fn<FnMemMap>("cuMemMap");
    x.mem_unmap = driver_fn<FnMemUnmap>("cuMemUnmap");
    x.set_access = driver_fn<FnSetAccess>("cuMemSetAccess");
    x.export_handle = driver_fn<FnExport>("cuMemExportToShareableHandle");
    x.import_handle = driver_fn<FnImport>("cuMemImportFromShareableHandle");
    x.mc_create = driver_fn<FnMcCreate>("cuMulticastCreate");
    x.mc_add_device = driver_fn<FnMcAddDevice>("cuMulticastAddDevice");
    x.mc_bind_mem = driver_fn<FnMcBindMem>("cuMulticastBindMem");
    x.mc_gran = driver_fn<FnMcGran>("cuMulticastGetGranularity");
    x.dev_attr = driver_fn<FnDevGetAttr>("cuDeviceGetAttribute");
    x.ok = x.mem_create && x.mem_release && x.mem_gran && x.addr_reserve && x.addr_free && x.mem_map && x.mem_unmap &&
           x.set_access && x.export_handle && x.import_handle && x.mc_create && x.mc_add_device && x.mc_bind_mem &&
           x.mc_gran && x.dev_attr;
    return x;
  }();
  return d;
}

#define B200_DRV_OK(expr)                                                                     \
  do {                                                                                        \
    CUresult r_ = (expr);                                                                     \
    if (r_ != CUDA_SUCCESS) return set_error(-6, "%s failed with CUresult %d", #expr, (int)r_); \
  } while (0)

CUmemAllocationProp device_prop(int device) {
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

CUmulticastObjectProp mc_prop(int world, size_t bytes) {
  CUmulticastObjectProp prop{};
  prop.numDevices = static_cast<unsigned>(world);
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200_vmm_supported(int device, int* multicast) {
  const Driver& d = driver();
  *multicast = 0;
  if (!d.ok) return 0;
  int vmm = 0, fd = 0, mc = 0;
  if (d.dev_attr(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  if (d.dev_attr(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  if (d.dev_attr(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS) mc = 0;
  *multicast = mc;
  return vmm && fd;
}

int b200_vmm_granularity(int device, int world, long long* bytes) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  CUmemAllocationProp prop = device_prop(device);
  size_t g = 0, gm = 0;
  B200_DRV_OK(d.mem_gran(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (world > 1) {
    CUmulticastObjectProp mp = mc_prop(world, g);
    if (d.mc_gran(&gm, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) gm = 0;
  }
  *bytes = static_cast<long long>(g > gm ? g : gm);
  return 0;
}

/* Physical allocation on `device` (bytes must be a multiple of the granularity) + a POSIX fd for the peers. */
int b200_vmm_create(int device, long long bytes, unsigned long long* handle, int* fd) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  CUmemAllocationProp prop = device_prop(device);
  CUmemGenericAllocationHandle h;
  B200_DRV_OK(d.mem_create(&h, static_cast<size_t>(bytes), &prop, 0));
  int out_fd = -1;
  B200_DRV_OK(d.export_handle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = h;
  *fd = out_fd;
  return 0;
}

/* Works for memory handles and for multicast handles alike (both are generic allocation handles). */
int b200_vmm_import(int fd, unsigned long long* handle) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  CUmemGenericAllocationHandle h;
  B200_DRV_OK(d.import_handle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = h;
  return 0;
}

/* Reserve a VA range, map `handle` (memory or multicast) into it and give `device` read/write access. */
int b200_vmm_map(unsigned long long handle, long long bytes, int device, long long alignment, void** ptr) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  CUdeviceptr va = 0;
  B200_DRV_OK(d.addr_reserve(&va, static_cast<size_t>(bytes), static_cast<size_t>(alignment), 0, 0));
  CUresult r = d.mem_map(va, static_cast<size_t>(bytes), 0, handle, 0);
  if (r != CUDA_SUCCESS) {
    d.addr_free(va, static_cast<size_t>(bytes));
    return set_error(-6, "cuMemMap failed with CUresult %d", (int)r);
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.set_access(va, static_cast<size_t>(bytes), &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.mem_unmap(va, static_cast<size_t>(bytes));
    d.addr_free(va, static_cast<size_t>(bytes));
    return set_error(-6, "cuMemSetAccess failed with CUresult %d", (int)r);
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

int b200_vmm_unmap(void* ptr, long long bytes) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  B200_DRV_OK(d.mem_unmap(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(bytes)));
  B200_DRV_OK(d.addr_free(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(bytes)));
  return 0;
}

int b200_vmm_release(unsigned long long handle) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  B200_DRV_OK(d.mem_release(handle));
  return 0;
}

/* Multicast object for `world` devices (called by ONE rank; the others import the fd). */
int b200_mc_create(int world, long long bytes, unsigned long long* handle, int* fd) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  CUmulticastObjectProp prop = mc_prop(world, static_cast<size_t>(bytes));
  CUmemGenericAllocationHandle h;
  B200_DRV_OK(d.mc_create(&h, &prop));
  int out_fd = -1;
  B200_DRV_OK(d.export_handle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = h;
  *fd = out_fd;
  return 0;
}

/* Every rank adds its device; ALL devices must have been added before any rank binds memory. */
int b200_mc_add_device(unsigned long long mc, int device) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  B200_DRV_OK(d.mc_add_device(mc, device));
  return 0;
}

/* Bind this rank's physical allocation at offset 0 of the multicast object. */
int b200_mc_bind(unsigned long long mc, unsigned long long mem, long long bytes) {
  const Driver& d = driver();
  if (!d.ok) return set_error(-4, "CUDA VMM driver entry points unavailable");
  B200_DRV_OK(d.mc_bind_mem(mc, 0, mem, 0, static_cast<size_t>(bytes), 0));
  return 0;
}

}  // extern "C"
