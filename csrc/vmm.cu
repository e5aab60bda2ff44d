// Symmetric-memory bootstrap on the CUDA virtual memory management API (cuMem*) and NVSwitch multicast objects
// (cuMulticast*): physical allocations that are exported as POSIX file descriptors, imported and mapped by every
// other rank of the node (peer access over NVLink), and bound into a multicast object whose single virtual address
// reaches ALL ranks' copies at once - the address `multimem.ld_reduce` / `multimem.st` operate on (NVLS: the switch
// adds the operands of a load and replicates the data of a store).
//
// Reference counterpart: Spark's broadcast / collect plumbing (elephas/spark_model.py:206-207,220-227); SURVEY.md 5.8.
// The driver API is reached through cudaGetDriverEntryPoint, so the library does not link against libcuda.
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"

namespace eb {
namespace vmm {

struct Api {
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok;
};

static void* entry(const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return sym;
}

static const Api& api() {
  static Api a;
  static std::once_flag once;
  std::call_once(once, [] {
#define EB_LOAD(field, name) a.field = reinterpret_cast<decltype(a.field)>(entry(name))
    EB_LOAD(MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    EB_LOAD(MemCreate, "cuMemCreate");
    EB_LOAD(MemRelease, "cuMemRelease");
    EB_LOAD(MemAddressReserve, "cuMemAddressReserve");
    EB_LOAD(MemAddressFree, "cuMemAddressFree");
    EB_LOAD(MemMap, "cuMemMap");
    EB_LOAD(MemUnmap, "cuMemUnmap");
    EB_LOAD(MemSetAccess, "cuMemSetAccess");
    EB_LOAD(MemExportToShareableHandle, "cuMemExportToShareableHandle");
    EB_LOAD(MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    EB_LOAD(MulticastCreate, "cuMulticastCreate");
    EB_LOAD(MulticastAddDevice, "cuMulticastAddDevice");
    EB_LOAD(MulticastBindMem, "cuMulticastBindMem");
    EB_LOAD(MulticastGetGranularity, "cuMulticastGetGranularity");
    EB_LOAD(DeviceGetAttribute, "cuDeviceGetAttribute");
    EB_LOAD(GetErrorString, "cuGetErrorString");
#undef EB_LOAD
    a.ok = a.MemCreate && a.MemAddressReserve && a.MemMap && a.MemSetAccess && a.MemExportToShareableHandle &&
           a.MemImportFromShareableHandle && a.MemGetAllocationGranularity && a.MemRelease && a.MemUnmap &&
           a.MemAddressFree && a.DeviceGetAttribute;
  });
  return a;
}

static int drv_fail(const char* what, CUresult r) {
  const char* msg = nullptr;
  if (api().GetErrorString) api().GetErrorString(r, &msg);
  return fail("%s failed: %s (%d)", what, msg ? msg : "unknown driver error", (int)r);
}

#define EB_DRV_OK(call, what)                       \
  do {                                              \
    CUresult _r = (call);                           \
    if (_r != CUDA_SUCCESS) return ::eb::vmm::drv_fail(what, _r); \
  } while (0)

static int current_device(int* dev) {
  EB_CUDA_OK(cudaFree(nullptr));  // make sure the primary context exists and is current
  EB_CUDA_OK(cudaGetDevice(dev));
  return 0;
}

static CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace vmm
}  // namespace eb

using namespace eb;

extern "C" int eb_vmm_supported(int32_t* vmm_out, int32_t* multicast_out) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  const vmm::Api& a = vmm::api();
  int v = 0, m = 0;
  if (a.ok) {
    a.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
    if (a.MulticastCreate && a.MulticastAddDevice && a.MulticastBindMem && a.MulticastGetGranularity)
      a.DeviceGetAttribute(&m, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  }
  if (vmm_out) *vmm_out = v;
  if (multicast_out) *multicast_out = m;
  return 0;
}

// Physical allocation on the current device, exportable as a file descriptor. `multicast_ranks` > 0: the size is also
// rounded to the granularity a multicast object over that many devices needs, so the allocation can be bound whole.
extern "C" int eb_vmm_create(int64_t bytes, int32_t multicast_ranks, uint64_t* handle_out, int64_t* size_out) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  const vmm::Api& a = vmm::api();
  if (!a.ok) return fail("eb_vmm_create: the driver does not expose the cuMem* API");
  CUmemAllocationProp prop = vmm::alloc_prop(dev);
  size_t gran = 0;
  EB_DRV_OK(a.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
  if (multicast_ranks > 0 && a.MulticastGetGranularity) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = multicast_ranks;
    mp.size = static_cast<size_t>(bytes);
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (a.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
  }
  const size_t size = (static_cast<size_t>(bytes) + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  EB_DRV_OK(a.MemCreate(&h, size, &prop, 0), "cuMemCreate");
  *handle_out = static_cast<uint64_t>(h);
  *size_out = static_cast<int64_t>(size);
  return 0;
}

extern "C" int eb_vmm_export_fd(uint64_t handle, int32_t* fd_out) {
  const vmm::Api& a = vmm::api();
  int fd = -1;
  EB_DRV_OK(a.MemExportToShareableHandle(&fd, static_cast<CUmemGenericAllocationHandle>(handle),
                                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
  *fd_out = fd;
  return 0;
}

extern "C" int eb_vmm_import_fd(int32_t fd, uint64_t* handle_out) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  const vmm::Api& a = vmm::api();
  CUmemGenericAllocationHandle h;
  EB_DRV_OK(a.MemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
  *handle_out = static_cast<uint64_t>(h);
  return 0;
}

// Reserve a virtual range, map the allocation (or multicast object) there and give the current device read/write access.
extern "C" int eb_vmm_map(uint64_t handle, int64_t size, void** ptr_out) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  const vmm::Api& a = vmm::api();
  CUdeviceptr va = 0;
  EB_DRV_OK(a.MemAddressReserve(&va, static_cast<size_t>(size), 0, 0, 0), "cuMemAddressReserve");
  CUresult r = a.MemMap(va, static_cast<size_t>(size), 0, static_cast<CUmemGenericAllocationHandle>(handle), 0);
  if (r != CUDA_SUCCESS) {
    a.MemAddressFree(va, static_cast<size_t>(size));
    return vmm::drv_fail("cuMemMap", r);
  }
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = a.MemSetAccess(va, static_cast<size_t>(size), &acc, 1);
  if (r != CUDA_SUCCESS) {
    a.MemUnmap(va, static_cast<size_t>(size));
    a.MemAddressFree(va, static_cast<size_t>(size));
    return vmm::drv_fail("cuMemSetAccess", r);
  }
  *ptr_out = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int eb_vmm_unmap(void* ptr, int64_t size) {
  const vmm::Api& a = vmm::api();
  EB_DRV_OK(a.MemUnmap(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(size)), "cuMemUnmap");
  EB_DRV_OK(a.MemAddressFree(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(size)), "cuMemAddressFree");
  return 0;
}

extern "C" int eb_vmm_release(uint64_t handle) {
  EB_DRV_OK(vmm::api().MemRelease(static_cast<CUmemGenericAllocationHandle>(handle)), "cuMemRelease");
  return 0;
}

// Multicast object over `ranks` devices (created by one rank, exported / imported like an allocation).
extern "C" int eb_mc_create(int32_t ranks, int64_t size, uint64_t* handle_out) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  const vmm::Api& a = vmm::api();
  if (!a.MulticastCreate) return fail("eb_mc_create: the driver does not expose cuMulticastCreate");
  CUmulticastObjectProp mp = {};
  mp.numDevices = ranks;
  mp.size = static_cast<size_t>(size);
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  EB_DRV_OK(a.MulticastCreate(&h, &mp), "cuMulticastCreate");
  *handle_out = static_cast<uint64_t>(h);
  return 0;
}

extern "C" int eb_mc_add_device(uint64_t mc_handle) {
  int dev = 0;
  if (vmm::current_device(&dev)) return 1;
  EB_DRV_OK(vmm::api().MulticastAddDevice(static_cast<CUmemGenericAllocationHandle>(mc_handle), dev), "cuMulticastAddDevice");
  return 0;
}

extern "C" int eb_mc_bind(uint64_t mc_handle, uint64_t mem_handle, int64_t size) {
  EB_DRV_OK(vmm::api().MulticastBindMem(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0,
                                        static_cast<CUmemGenericAllocationHandle>(mem_handle), 0, static_cast<size_t>(size), 0),
            "cuMulticastBindMem");
  return 0;
}
